// host/scenes.hxx -- the deterministic scene generators of SURVEY App. D in C++ (same bits as
// raytracinginoneweekend_b200/scenes.py; tests/test_host_cpp.py compares them).  Every scene comes back
// translated so that the camera sits at the origin.
#pragma once
#include <algorithm>
#include <cstdint>
#include <random>
#include <string>
#include <utility>
#include "camera.hxx"
#include "raytracer.hxx"

namespace scenes {

struct scene final {
    raytracer::data data;
    raytracer::camera camera;
};

namespace detail {
inline float U(std::mt19937 &g) { return static_cast<float>(g() >> 8) * 5.9604644775390625e-8f; }
inline void base_materials(raytracer::data &d)
{
    d.materials.emplace_back(material::lambert{math::vec3{.1, .2, .5}});
    d.materials.emplace_back(material::metal{math::vec3{.8, .6, .2}, 0.f});
    d.materials.emplace_back(material::dielectric{math::vec3{1}, 1.5f});
    d.materials.emplace_back(material::lambert{math::vec3{.64, .8, .0}});
}
inline scene to_origin(raytracer::data d, raytracer::camera const &c)
{
    for (auto &s : d.spheres) s.center -= c.position();
    return scene{std::move(d), c.at_origin()};
}
}  // namespace detail

// src/CUDA/cuda_impl.cu:404-417 (geometry) and :348-352 (camera): ground + 3 spheres, hollow glass.
inline scene simple(std::uint32_t width, std::uint32_t height)
{
    raytracer::data d;
    detail::base_materials(d);
    d.spheres.emplace_back(math::vec3{0, 0, -1}, .5f, 0);
    d.spheres.emplace_back(math::vec3{0, -100.5f, -1}, 100.f, 3);
    d.spheres.emplace_back(math::vec3{1, 0, -1}, .5f, 1);
    d.spheres.emplace_back(math::vec3{-1, 0, -1}, .5f, 2);
    d.spheres.emplace_back(math::vec3{-1, 0, -1}, -.499f, 2);
    raytracer::camera cam{math::vec3{0}, math::vec3{0, 0, -1}, math::vec3{0, 1, 0},
                          static_cast<float>(width) / static_cast<float>(height), 88.f, .0625f, 1.f};
    return detail::to_origin(std::move(d), cam);
}

// src/main.cxx:120-183 with the `#if 0` block made well-defined.
inline scene huge(std::uint32_t width, std::uint32_t height, std::uint32_t seed = 20260101u)
{
    raytracer::data d;
    detail::base_materials(d);
    d.spheres.emplace_back(math::vec3{0, 1, 0}, 1.f, 0);
    d.spheres.emplace_back(math::vec3{0, -1000.125f, 0}, 1000.f, 3);
    d.spheres.emplace_back(math::vec3{2, 1, 0}, 1.f, 1);
    d.spheres.emplace_back(math::vec3{-2, 1, 0}, 1.f, 2);
    d.spheres.emplace_back(math::vec3{-2, 1, 0}, -.99f, 2);
    std::mt19937 g{seed};
    for (auto a = -11; a < 11; ++a) {
        for (auto b = -11; b < 11; ++b) {
            auto k = std::min(2, static_cast<int>(detail::U(g) * 3.f));
            float cx = .9f * detail::U(g) + static_cast<float>(a);
            float cz = .9f * detail::U(g) + static_cast<float>(b);
            float r0 = detail::U(g), r1 = detail::U(g), r2 = detail::U(g), r3 = detail::U(g);
            math::vec3 center{cx, .2f, cz};
            if (math::distance(center, math::vec3{0, 1, 0}) < 1.f) continue;
            d.spheres.emplace_back(center, .2f, d.materials.size());
            if (k == 0) d.materials.emplace_back(material::lambert{math::vec3{r0, r1, r2}});
            else if (k == 1) d.materials.emplace_back(material::metal{math::vec3{r0, r1, r2}, .5f * r3});
            else d.materials.emplace_back(material::dielectric{math::vec3{r0, r1, r2}, 1.5f});
        }
    }
    math::vec3 eye{-4, 3.2, 5}, lookat{0, 1, 0};
    raytracer::camera cam{eye, lookat, math::vec3{0, 1, 0}, static_cast<float>(width) / static_cast<float>(height), 42.f,
                          0.0625f, math::distance(eye, lookat)};
    return detail::to_origin(std::move(d), cam);
}

// App. D synthetic scene: one sphere per cell of a jittered n_side x n_side grid plus a ground sphere.
inline scene synthetic(std::uint32_t n_side, std::uint32_t width, std::uint32_t height, std::uint32_t seed = 7u)
{
    raytracer::data d;
    d.spheres.reserve(static_cast<std::size_t>(n_side) * n_side + 1);
    d.materials.reserve(static_cast<std::size_t>(n_side) * n_side + 1);
    d.spheres.emplace_back(math::vec3{0, -10000, 0}, 10000.f, 0);
    d.materials.emplace_back(material::lambert{math::vec3{.64, .8, .0}});
    std::mt19937 g{seed};
    const float half = static_cast<float>(n_side / 2);
    for (std::uint32_t i = 0; i < n_side; ++i) {
        for (std::uint32_t j = 0; j < n_side; ++j) {
            float u[8];
            for (auto &x : u) x = detail::U(g);
            float r = .1f + .3f * u[0];
            float cx = (static_cast<float>(i) - half) + .8f * u[1];
            float cz = (static_cast<float>(j) - half) + .8f * u[2];
            auto k = std::min(2, static_cast<int>(u[3] * 3.f));
            d.spheres.emplace_back(math::vec3{cx, r, cz}, r, d.materials.size());
            if (k == 0) d.materials.emplace_back(material::lambert{math::vec3{u[4], u[5], u[6]}});
            else if (k == 1) d.materials.emplace_back(material::metal{math::vec3{u[4], u[5], u[6]}, .5f * u[7]});
            else d.materials.emplace_back(material::dielectric{math::vec3{u[4], u[5], u[6]}, 1.5f});
        }
    }
    const float scale = static_cast<float>(n_side) / 1000.f;
    raytracer::camera cam{math::vec3{0, 30.f * scale, 120.f * scale}, math::vec3{0, 0, 0}, math::vec3{0, 1, 0},
                          static_cast<float>(width) / static_cast<float>(height), 40.f, 0.f, 1.f};
    return detail::to_origin(std::move(d), cam);
}

inline scene by_name(std::string const &name, std::uint32_t width, std::uint32_t height)
{
    if (name == "simple") return simple(width, height);
    if (name == "huge") return huge(width, height);
    if (name.rfind("synthetic", 0) == 0) {
        std::uint32_t n_side = 1000;
        auto colon = name.find(':');
        if (colon != std::string::npos) n_side = static_cast<std::uint32_t>(std::stoul(name.substr(colon + 1)));
        return synthetic(n_side, width, height);
    }
    throw std::invalid_argument("unknown scene '" + name + "'");
}

}  // namespace scenes
