// host/cuda_impl.hxx -- the reference-facing entry points of the B200 backend.
//
//   void cuda_impl(width, height, image_texels)
//       the reference's DECLARATION, symbol for symbol (src/main.cxx:18; defined in src/CUDA/cuda_impl.cu:384-453), so
//       the reference's main.cxx links against this backend unchanged: the caller owns and pre-sizes the vector
//       (src/main.cxx:112), the callee fills width*height RGB8 texels in place, row-major, top row first, and
//       throws std::runtime_error on failure (src/CUDA/cuda_impl.cu:101-114) -- without cudaDeviceReset.
//       Like the reference's, it renders its own hard-coded scene (src/CUDA/cuda_impl.cu:348-352,404-417) at
//       the reference GPU path's 48 spp / 32 bounces (src/CUDA/cuda_impl.cu:59-60).
//       DEVIATION, on purpose: the PICTURE is the reference's CPU path's, not its CUDA path's.  The parity target of this
//       backend is the CPU code: sky blend .5*y + 1.0 (src/main.cxx:71; the CUDA file uses + .5,
//       src/CUDA/cuda_impl.cu:293) and a lens offset with aperture .0625 (src/camera.hxx:46-57; the CUDA build compiles the
//       offset out, src/camera.hxx:48).  The sky is therefore bluer and there is depth of field the replaced function
//       does not have.  cuda_impl_options::book_optics gives the + .5 blend.
//
//   void cuda_impl(width, height, image_texels, data, camera, options)
//       the same call for a caller-supplied scene (what main() would pass if the scene were not hard-coded).
#pragma once
#include <cstdint>
#include <functional>
#include <vector>
#include "camera.hxx"
#include "math.hxx"
#include "raytracer.hxx"

struct cuda_impl_options final {
    std::uint32_t samples_per_pixel{48};   // src/CUDA/cuda_impl.cu:59
    std::uint32_t max_depth{32};           // src/CUDA/cuda_impl.cu:60
    std::uint64_t seed{1};
    bool exact{false};                     // RT_FLAG_EXACT: parity shading (renders bit-identical to the CPU oracle)
    bool book_optics{false};               // RT_FLAG_BOOK_OPTICS: textbook refraction / camera / sky instead of the reference's
    int device{0};
    // several GPUs of this machine behind the same call (rt_multi: device-side tile queue, peer-store gather, NCCL ack);
    // empty = the single `device` above.  Progressive rendering stays on one GPU.
    std::vector<int> devices{};
    bool gather_nccl{false};               // RT_MULTI_GATHER_NCCL: gather by ncclReduce instead of peer stores
    // progressive rendering (rt_render_progressive): passes of this many samples (0 = one shot) and, every
    // `preview_every` passes, the image so far handed to `on_preview` while the next pass renders; return true to stop
    std::uint32_t samples_per_pass{0};
    std::uint32_t preview_every{1};
    std::function<bool(std::uint32_t pass, std::uint32_t samples_done, std::uint8_t const *rgb8)> on_preview{};
};

void cuda_impl(std::uint32_t width, std::uint32_t height, std::vector<math::u8vec3> &image_texels);

void cuda_impl(std::uint32_t width, std::uint32_t height, std::vector<math::u8vec3> &image_texels,
               raytracer::data const &data, raytracer::camera const &camera, cuda_impl_options const &options = {});
