// host/math.hxx -- the scene-description vocabulary of the reference (math::vec3, math::u8vec3, math::ray)
// for host code that talks to the B200 backend.  Same names, template signature and memory layout as the
// reference's src/math.hxx:24-45,337-354 (so `std::vector<math::u8vec3>&` mangles identically and the
// reference's main.cxx links against our cuda_impl), written from scratch and limited to what scene
// construction needs.  The render arithmetic itself lives in the CUDA kernels, not here.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <type_traits>

namespace math {

template <std::size_t D, class T>
struct vec;

template <class T>
struct vec<3, T> final {
    T x{}, y{}, z{};

    constexpr vec() noexcept = default;
    template <class X, class Y, class Z>
    constexpr vec(X x_, Y y_, Z z_) noexcept : x{static_cast<T>(x_)}, y{static_cast<T>(y_)}, z{static_cast<T>(z_)} {}
    template <class S, std::enable_if_t<std::is_arithmetic_v<S>, int> = 0>
    constexpr vec(S s) noexcept : x{static_cast<T>(s)}, y{static_cast<T>(s)}, z{static_cast<T>(s)} {}

    constexpr vec operator+(vec const &r) const { return {x + r.x, y + r.y, z + r.z}; }
    constexpr vec operator-(vec const &r) const { return {x - r.x, y - r.y, z - r.z}; }
    constexpr vec operator*(vec const &r) const { return {x * r.x, y * r.y, z * r.z}; }
    constexpr vec operator-() const { return {-x, -y, -z}; }
    template <class S, std::enable_if_t<std::is_arithmetic_v<S>, int> = 0>
    constexpr vec operator*(S s) const { return {x * s, y * s, z * s}; }
    template <class S, std::enable_if_t<std::is_arithmetic_v<S>, int> = 0>
    constexpr vec operator/(S s) const { return {x / s, y / s, z / s}; }
    constexpr vec &operator-=(vec const &r) { x -= r.x; y -= r.y; z -= r.z; return *this; }
    constexpr vec &operator+=(vec const &r) { x += r.x; y += r.y; z += r.z; return *this; }
};

using vec3 = vec<3, float>;
using u8vec3 = vec<3, std::uint8_t>;
static_assert(sizeof(vec3) == 12 && sizeof(u8vec3) == 3, "layout must match the reference (SURVEY App. B)");

inline float dot(vec3 const &a, vec3 const &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float length(vec3 const &a) { return std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
inline float distance(vec3 const &a, vec3 const &b) { return length(a - b); }

struct ray final {
    vec3 origin;
    vec3 direction;
};

}  // namespace math
