// host/main_demo.cxx -- what the reference's main() does (src/main.cxx:104-118: size the output, call
// cuda_impl, write a P6 PPM), with the scene, size, spp and depth on the command line.
//   rtiow_demo [--scene simple|huge|synthetic:N] [--width W] [--height H] [--spp S] [--depth D] [--seed X]
//              [--exact] [--book-optics] [--out image.ppm] [--dump-scene file.bin]
//              [--gpus N | --devices a,b,c] [--gather-nccl]        several GPUs behind the same cuda_impl call
//              [--spp-per-pass P [--preview-every K] [--preview-prefix path_]]   progressive passes with previews
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <variant>
#include <vector>

#include "../../include/rt_api.h"
#include "cuda_impl.hxx"
#include "scenes.hxx"

namespace {
// raw dump for tests: n, m, then n x {cx,cy,cz,r,mat}, m x {r,g,b,param,kind}, 13 camera floats
void dump_scene(scenes::scene const &sc, std::string const &path)
{
    std::ofstream f{path, std::ios::binary | std::ios::trunc};
    if (!f) throw std::runtime_error("bad file: " + path);
    std::uint32_t n = static_cast<std::uint32_t>(sc.data.spheres.size()), m = static_cast<std::uint32_t>(sc.data.materials.size());
    f.write(reinterpret_cast<char const *>(&n), 4);
    f.write(reinterpret_cast<char const *>(&m), 4);
    for (auto const &s : sc.data.spheres) {
        float v[4] = {s.center.x, s.center.y, s.center.z, s.radius};
        std::uint32_t mi = static_cast<std::uint32_t>(s.material_index);
        f.write(reinterpret_cast<char const *>(v), 16);
        f.write(reinterpret_cast<char const *>(&mi), 4);
    }
    for (auto const &mt : sc.data.materials) {
        float v[4] = {0, 0, 0, 0};
        std::uint32_t kind = 0;
        std::visit([&](auto const &x) {
            using type = std::decay_t<decltype(x)>;
            v[0] = x.albedo.x; v[1] = x.albedo.y; v[2] = x.albedo.z;
            if constexpr (std::is_same_v<type, material::metal>) { v[3] = x.roughness; kind = 1; }
            else if constexpr (std::is_same_v<type, material::dielectric>) { v[3] = x.refraction_index; kind = 2; }
        }, mt);
        f.write(reinterpret_cast<char const *>(v), 16);
        f.write(reinterpret_cast<char const *>(&kind), 4);
    }
    auto const &c = sc.camera;
    float cam[13] = {c.position().x, c.position().y, c.position().z, c.lookat().x, c.lookat().y, c.lookat().z,
                     c.up().x, c.up().y, c.up().z, c.aspect(), c.vFOV(), c.aperture(), c.focus_distance()};
    f.write(reinterpret_cast<char const *>(cam), sizeof cam);
}
}  // namespace

int main(int argc, char **argv)
{
    std::string scene_name = "simple", out = "image_cuda.ppm", dump, preview_prefix;
    std::uint32_t width = 1920, height = 1080;   // src/main.cxx:29-30
    cuda_impl_options opt;
    try {
        for (int i = 1; i < argc; ++i) {
            std::string a = argv[i];
            auto next = [&]() -> std::string {
                if (i + 1 >= argc) throw std::invalid_argument("missing value after " + a);
                return argv[++i];
            };
            if (a == "--scene") scene_name = next();
            else if (a == "--width") width = static_cast<std::uint32_t>(std::stoul(next()));
            else if (a == "--height") height = static_cast<std::uint32_t>(std::stoul(next()));
            else if (a == "--spp") opt.samples_per_pixel = static_cast<std::uint32_t>(std::stoul(next()));
            else if (a == "--depth") opt.max_depth = static_cast<std::uint32_t>(std::stoul(next()));
            else if (a == "--seed") opt.seed = std::stoull(next());
            else if (a == "--device") opt.device = std::stoi(next());
            else if (a == "--gpus") { int n = std::stoi(next()); opt.devices.clear(); for (int d = 0; d < n; ++d) opt.devices.push_back(d); }
            else if (a == "--devices") {
                std::string list = next(); opt.devices.clear();
                for (std::size_t pos = 0; pos <= list.size();) {
                    auto comma = list.find(',', pos);
                    if (comma == std::string::npos) comma = list.size();
                    opt.devices.push_back(std::stoi(list.substr(pos, comma - pos)));
                    pos = comma + 1;
                }
            }
            else if (a == "--gather-nccl") opt.gather_nccl = true;
            else if (a == "--exact") opt.exact = true;
            else if (a == "--book-optics") opt.book_optics = true;
            else if (a == "--spp-per-pass") opt.samples_per_pass = static_cast<std::uint32_t>(std::stoul(next()));
            else if (a == "--preview-every") opt.preview_every = static_cast<std::uint32_t>(std::stoul(next()));
            else if (a == "--preview-prefix") preview_prefix = next();
            else if (a == "--out") out = next();
            else if (a == "--dump-scene") dump = next();
            else throw std::invalid_argument("unknown option " + a);
        }
        std::cout << "started... \n";                                  // src/main.cxx:106
        auto sc = scenes::by_name(scene_name, width, height);
        if (!dump.empty()) { dump_scene(sc, dump); return 0; }
        std::vector<math::u8vec3> output(static_cast<std::size_t>(width) * height);   // src/main.cxx:112
        if (opt.samples_per_pass) {                                    // progressive: report (and optionally save) the previews
            opt.on_preview = [&](std::uint32_t pass, std::uint32_t done, std::uint8_t const *rgb8) {
                std::cout << "preview: pass " << pass << ", " << done << " spp\n";
                if (!preview_prefix.empty())
                    rt_write_ppm((preview_prefix + std::to_string(done) + ".ppm").c_str(), width, height, rgb8);
                return false;
            };
        }
        cuda_impl(width, height, output, sc.data, sc.camera, opt);
        if (rt_write_ppm(out.c_str(), width, height, reinterpret_cast<std::uint8_t const *>(output.data())) != RT_OK)
            throw std::runtime_error(rt_last_error());
        std::cout << "wrote " << out << " (" << width << "x" << height << ", " << opt.samples_per_pixel << " spp, depth "
                  << opt.max_depth << ")\n";
    } catch (std::exception const &e) {
        std::cerr << "rtiow_demo: " << e.what() << "\n";
        return 1;
    }
    return 0;
}
