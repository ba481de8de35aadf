// host/material.hxx -- material::{lambert, metal, dielectric, types}: the three surface models of the reference
// (src/material.hxx:12-51) with its member names and constructor arguments, so scene code written for the
// reference compiles against this header.  On the device a material is float4{albedo, param} + a 2-bit kind;
// `kind` and `param()` below are what the C-ABI packing (cuda_impl_shim.cpp) reads.
#pragma once
#include <cstdint>
#include <variant>
#include "math.hxx"

namespace material {
namespace detail {
// every surface model carries a tint; the base keeps the member name the reference uses
struct tinted {
    math::vec3 albedo{1};
    tinted() = default;
    explicit tinted(math::vec3 tint) noexcept : albedo{tint} {}
};
}  // namespace detail

// diffuse scatter about the normal; no parameter
struct lambert final : detail::tinted {
    static constexpr std::uint32_t kind = 0;            // RT_LAMBERT
    using detail::tinted::tinted;
    float param() const noexcept { return 0.f; }
};

// mirror reflection blurred by `roughness` times a unit-ball sample
struct metal final : detail::tinted {
    static constexpr std::uint32_t kind = 1;            // RT_METAL
    float roughness{0};
    metal() = default;
    metal(math::vec3 tint, float fuzz) noexcept : detail::tinted{tint}, roughness{fuzz} {}
    float param() const noexcept { return roughness; }
};

// refraction / Schlick reflection with index `refraction_index`
struct dielectric final : detail::tinted {
    static constexpr std::uint32_t kind = 2;            // RT_DIELECTRIC
    float refraction_index{1};
    dielectric() = default;
    dielectric(math::vec3 tint, float index) noexcept : detail::tinted{tint}, refraction_index{index} {}
    float param() const noexcept { return refraction_index; }
};

using types = std::variant<lambert, metal, dielectric>;
}  // namespace material
