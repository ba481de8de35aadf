// host/cuda_impl_shim.cpp -- C++ shim between the reference's interface and the C ABI (include/rt_api.h).
#include "cuda_impl.hxx"

#include <stdexcept>
#include <string>
#include <type_traits>
#include <variant>

#include "../../include/rt_api.h"
#include "scenes.hxx"

namespace {

struct ctx_guard final {
    rt_ctx *ctx{nullptr};
    ~ctx_guard() { rt_destroy(ctx); }
};
struct multi_guard final {
    rt_multi *m{nullptr};
    ~multi_guard() { rt_multi_destroy(m); }
};

void check(int rc, char const *what)
{
    if (rc != RT_OK) throw std::runtime_error(std::string{what} + ": " + rt_last_error());
}

}  // namespace

void cuda_impl(std::uint32_t width, std::uint32_t height, std::vector<math::u8vec3> &image_texels,
               raytracer::data const &data, raytracer::camera const &camera, cuda_impl_options const &options)
{
    if (image_texels.size() != static_cast<std::size_t>(width) * height)
        throw std::runtime_error("cuda_impl: image_texels must be pre-sized to width*height (src/main.cxx:112)");

    // express the scene with the eye at the origin, where the reference's ray formula is exact
    auto const eye = camera.position();
    auto const cam0 = camera.at_origin();

    std::vector<rt_sphere> spheres;
    spheres.reserve(data.spheres.size());
    for (auto const &s : data.spheres) {
        auto c = s.center - eye;
        spheres.push_back(rt_sphere{c.x, c.y, c.z, s.radius, static_cast<std::uint32_t>(s.material_index)});
    }
    std::vector<rt_material> materials;
    materials.reserve(data.materials.size());
    for (auto const &m : data.materials) {
        std::visit([&materials](auto const &mat) {
            using type = std::decay_t<decltype(mat)>;
            if constexpr (std::is_same_v<type, material::lambert>)
                materials.push_back(rt_material{mat.albedo.x, mat.albedo.y, mat.albedo.z, 0.f, RT_LAMBERT});
            else if constexpr (std::is_same_v<type, material::metal>)
                materials.push_back(rt_material{mat.albedo.x, mat.albedo.y, mat.albedo.z, mat.roughness, RT_METAL});
            else
                materials.push_back(rt_material{mat.albedo.x, mat.albedo.y, mat.albedo.z, mat.refraction_index, RT_DIELECTRIC});
        }, m);
    }

    float const cam13[13] = {cam0.position().x, cam0.position().y, cam0.position().z,
                             cam0.lookat().x, cam0.lookat().y, cam0.lookat().z,
                             cam0.up().x, cam0.up().y, cam0.up().z,
                             cam0.aspect(), cam0.vFOV(), cam0.aperture(), cam0.focus_distance()};

    rt_render_params p{};
    p.spp = options.samples_per_pixel;
    p.max_depth = options.max_depth;
    p.seed = options.seed;
    p.sample_base = 0;
    p.flags = (options.exact ? static_cast<std::uint32_t>(RT_FLAG_EXACT) : 0u) |
              (options.book_optics ? static_cast<std::uint32_t>(RT_FLAG_BOOK_OPTICS) : 0u);
    static_assert(sizeof(math::u8vec3) == 3, "texels are tightly packed RGB8");
    auto *texels = reinterpret_cast<std::uint8_t *>(image_texels.data());
    if (options.devices.size() > 1 && options.samples_per_pass == 0) {
        // N GPUs, one call: tiles of 128 x 64 pulled from the device-side queue, texels stored into the first device's frame
        multi_guard mg;
        check(rt_multi_create(&mg.m, options.devices.data(), static_cast<int>(options.devices.size())), "rt_multi_create");
        check(rt_multi_render_frame(mg.m, spheres.data(), static_cast<std::uint32_t>(spheres.size()), materials.data(),
                                    static_cast<std::uint32_t>(materials.size()), cam13, width, height, &p, 128, 64,
                                    options.gather_nccl ? static_cast<std::uint32_t>(RT_MULTI_GATHER_NCCL) : 0u, texels),
              "rt_multi_render_frame");
        return;
    }
    ctx_guard g;
    check(rt_create(&g.ctx, options.devices.size() == 1 ? options.devices[0] : options.device), "rt_create");
    if (options.samples_per_pass == 0) {
        check(rt_render_frame(g.ctx, spheres.data(), static_cast<std::uint32_t>(spheres.size()), materials.data(),
                              static_cast<std::uint32_t>(materials.size()), cam13, width, height, &p, texels),
              "rt_render_frame");
        return;
    }
    check(rt_scene_upload(g.ctx, spheres.data(), static_cast<std::uint32_t>(spheres.size()), materials.data(),
                          static_cast<std::uint32_t>(materials.size())), "rt_scene_upload");
    check(rt_camera_set(g.ctx, cam13, cam13 + 3, cam13 + 6, cam13[9], cam13[10], cam13[11], cam13[12]), "rt_camera_set");
    check(rt_framebuffer(g.ctx, width, height), "rt_framebuffer");
    struct relay { cuda_impl_options const *opt; std::exception_ptr error; } r{&options, nullptr};
    auto trampoline = [](void *user, std::uint32_t pass, std::uint32_t done, std::uint8_t const *rgb8) -> int {
        auto *rl = static_cast<relay *>(user);
        try { return rl->opt->on_preview(pass, done, rgb8) ? 1 : 0; }
        catch (...) { rl->error = std::current_exception(); return 1; }     // no exception crosses the C ABI
    };
    check(rt_render_progressive(g.ctx, &p, options.samples_per_pass, options.preview_every,
                                options.on_preview ? static_cast<rt_preview_fn>(trampoline) : nullptr, &r, texels, nullptr),
          "rt_render_progressive");
    if (r.error) std::rethrow_exception(r.error);
}

void cuda_impl(std::uint32_t width, std::uint32_t height, std::vector<math::u8vec3> &image_texels)
{
    auto const sc = scenes::simple(width, height);
    cuda_impl(width, height, image_texels, sc.data, sc.camera, cuda_impl_options{});
}
