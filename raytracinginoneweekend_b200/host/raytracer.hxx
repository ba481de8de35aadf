// host/raytracer.hxx -- raytracer::data{spheres, materials} as in src/raytracer.hxx:19-30 (without the
// std::mt19937: the backend's RNG is a counter-based Philox keyed by (seed, pixel, sample)).
#pragma once
#include <vector>
#include "material.hxx"
#include "primitives.hxx"

namespace raytracer {
struct data final {
    static auto constexpr bounces_number = 64u;   // the reference's CPU cap (src/raytracer.hxx:20)
    std::vector<primitives::sphere> spheres;
    std::vector<material::types> materials;
};
}  // namespace raytracer
