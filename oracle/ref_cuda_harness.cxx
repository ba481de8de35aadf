// oracle/ref_cuda_harness.cxx -- TEST / MEASUREMENT INFRASTRUCTURE, NOT PRODUCT CODE.
//
// "Legacy GPU comparator" (SURVEY.md 8(f) row 2): runs the reference's OWN CUDA path,
// cuda_impl(width, height, image_texels) of src/CUDA/cuda_impl.cu:384-453, on this GPU and times it the way
// the reference's main() uses it (src/main.cxx:104-118: size the vector, call, write the P6 file).
// The reference translation unit is compiled unmodified in arithmetic by oracle/build_ref_cuda.sh (a sed-patched
// temporary copy; only the binary lands in oracle/_ref/).  This file only declares the entry point, calls it and
// prints one JSON line.  The reference hard-codes its scene (5 spheres), 48 spp and depth 32
// (src/CUDA/cuda_impl.cu:59-60,404-417), its own xorshift RNG and a different sky/lambert convention,
// so the number is INFORMATIONAL: "the reference's GPU code on the same box", not the parity oracle.
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include "math.hxx"

extern void cuda_impl(std::uint32_t width, std::uint32_t height, std::vector<math::u8vec3> &image_texels);

int main(int argc, char **argv)
{
    std::uint32_t W = 1280, H = 720;
    int reps = 3;
    const char *out = nullptr;
    for (int i = 1; i < argc; ++i) {
        if (!std::strcmp(argv[i], "--width") && i + 1 < argc) W = (std::uint32_t)std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--height") && i + 1 < argc) H = (std::uint32_t)std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--reps") && i + 1 < argc) reps = std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--out") && i + 1 < argc) out = argv[++i];
    }
    const std::uint32_t spp = 48;          // src/CUDA/cuda_impl.cu:59 (sampling_number)
    std::vector<math::u8vec3> texels((std::size_t)W * H);
    double best = 1e300, first = 0;
    for (int r = 0; r < reps; ++r) {
        auto t0 = std::chrono::steady_clock::now();
        cuda_impl(W, H, texels);           // blocking: the reference synchronises the device itself
        auto t1 = std::chrono::steady_clock::now();
        double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        if (r == 0) first = ms;
        if (ms < best) best = ms;
    }
    unsigned long long sum = 0;
    for (auto const &t : texels) sum += (unsigned)t.x + (unsigned)t.y + (unsigned)t.z;
    if (out) {
        std::ofstream f(out, std::ios::binary);
        f << "P6\n" << W << ' ' << H << "\n255\n";
        f.write(reinterpret_cast<const char *>(texels.data()), (std::streamsize)texels.size() * 3);
    }
    std::printf("{\"impl\": \"reference_cuda_legacy\", \"scene\": \"reference hard-coded (5 spheres)\", \"width\": %u, \"height\": %u, "
                "\"spp\": %u, \"max_depth\": 32, \"reps\": %d, \"first_ms\": %.3f, \"best_ms\": %.3f, "
                "\"Gsamples_per_s\": %.6f, \"mean_texel\": %.4f}\n",
                W, H, spp, reps, first, best, (double)W * H * spp / (best * 1e6),
                (double)sum / (3.0 * (double)texels.size()));
    return 0;
}
