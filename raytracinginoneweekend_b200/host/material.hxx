// host/material.hxx -- material::{lambert, metal, dielectric, types} with the reference's constructors
// (src/material.hxx:12-51).  On the device a material is float4{albedo, param} + a 2-bit kind.
#pragma once
#include <variant>
#include "math.hxx"

namespace material {
struct lambert final {
    math::vec3 albedo{1};
    lambert() = default;
    explicit lambert(math::vec3 albedo_) noexcept : albedo{albedo_} {}
};
struct metal final {
    math::vec3 albedo{1};
    float roughness{0};
    metal() = default;
    metal(math::vec3 albedo_, float roughness_) noexcept : albedo{albedo_}, roughness{roughness_} {}
};
struct dielectric final {
    math::vec3 albedo{1};
    float refraction_index{1};
    dielectric() = default;
    dielectric(math::vec3 albedo_, float refraction_index_) noexcept : albedo{albedo_}, refraction_index{refraction_index_} {}
};
using types = std::variant<lambert, metal, dielectric>;
}  // namespace material
