// host/primitives.hxx -- primitives::sphere with the reference's constructor (src/primitives.hxx:6-17).
#pragma once
#include <cstddef>
#include "math.hxx"

namespace primitives {
struct sphere final {
    math::vec3 center{0};
    float radius{1.f};          // negative: inward-facing normal (hollow glass), src/raytracer.hxx:70
    std::size_t material_index{0};

    sphere() = default;
    sphere(math::vec3 center_, float radius_, std::size_t material_index_) noexcept
        : center{center_}, radius{radius_}, material_index{material_index_} {}
};
}  // namespace primitives
