// oracle/ref_harness.cxx -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Thin C-ABI harness around the UNMODIFIED arithmetic of the reference CPU path
// (Alabuta/RaytracingInOneWeekend).  It is compiled by oracle/build_ref.sh from the
// reference sources where they lie under /root/reference (a sed-patched temporary copy
// is needed because the headers use an MSVC-only `* = 0` template default; no arithmetic
// is touched) and only the resulting binary lands in oracle/_ref/.
//
// Everything numerically relevant is the reference's own code, called directly:
//   raytracer::intersect / hit_world / apply_material   (src/raytracer.hxx:52-199)
//   raytracer::camera ctor + camera::ray                (src/camera.hxx:24-57)
//   app::background_color / gamma_correction / normalize_rgb_to_8bit (src/main.cxx:39-85)
//   math::reflect / refract / normalize / mix           (src/math.hxx:219-335)
// What this file restates (because the reference hard-codes it inside main()):
//   * the pixel/sample loop            (src/main.cxx:191-215)
//   * the bounce loop with a RUNTIME depth cap (src/main.cxx:52-75; app::color bakes 64)
//   * deterministic seeding of the two mt19937 generators per image row.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load the library this builds.

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <execution>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <memory>
#include <numeric>
#include <optional>
#include <random>
#include <string>
#include <string_view>
#include <thread>
#include <variant>
#include <vector>

// The camera keeps its derived vectors and its private mt19937 behind `private:`
// (src/camera.hxx:59-76).  The harness must seed that generator to be reproducible.
#define private public
#define main reference_main
#include "main.cxx"
#undef main
#undef private

// main.cxx:18 declares it; the CPU oracle never calls it.
void cuda_impl(std::uint32_t, std::uint32_t, std::vector<math::u8vec3> &) {}

namespace {

struct scene_view {
    const float *spheres;     // n x {cx,cy,cz,r}
    const std::uint32_t *sphere_mat; // n
    std::uint32_t n;
    const float *materials;   // m x {r,g,b,param}
    const std::uint32_t *mat_kind;   // m : 0 lambert 1 metal 2 dielectric
    std::uint32_t m;
};

void fill_data(raytracer::data &d, scene_view const &s)
{
    d.spheres.clear();
    d.materials.clear();
    for (std::uint32_t i = 0; i < s.m; ++i) {
        math::vec3 albedo{s.materials[4 * i + 0], s.materials[4 * i + 1], s.materials[4 * i + 2]};
        float param = s.materials[4 * i + 3];
        switch (s.mat_kind[i]) {
            case 0: d.materials.push_back(material::lambert{albedo}); break;
            case 1: d.materials.push_back(material::metal{albedo, param}); break;
            default: d.materials.push_back(material::dielectric{albedo, param}); break;
        }
    }
    for (std::uint32_t i = 0; i < s.n; ++i)
        d.spheres.emplace_back(math::vec3{s.spheres[4 * i + 0], s.spheres[4 * i + 1], s.spheres[4 * i + 2]},
                               s.spheres[4 * i + 3], static_cast<std::size_t>(s.sphere_mat[i]));
}

// cam = {pos[3], lookat[3], up[3], aspect, vfov, aperture, focus}
std::unique_ptr<raytracer::camera> make_camera(const float *c)
{
    return std::make_unique<raytracer::camera>(
        math::vec3{c[0], c[1], c[2]}, math::vec3{c[3], c[4], c[5]}, math::vec3{c[6], c[7], c[8]},
        c[9], c[10], c[11], c[12]);
}

std::uint32_t mix_seed(std::uint64_t seed, std::uint32_t row, std::uint32_t which)
{
    // splitmix64 finaliser; any fixed rule works, the C port uses the same one.
    std::uint64_t z = seed + 0x9E3779B97F4A7C15ull * (2ull * row + which + 1ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return static_cast<std::uint32_t>(z);
}

// src/main.cxx:52-75 with `bounces_number` replaced by a runtime depth.
// stats: [0] segments (hit_world calls) [1] paths that reached the sky
//        [2] absorbed (scatter failed)  [3] depth exhausted
math::vec3 color_depth(raytracer::data &rd, math::ray ray, std::uint32_t depth, std::uint64_t *stats)
{
    math::vec3 attenuation{1};
    math::vec3 energy_absorption{0};
    auto scattered_ray = ray;

    for (auto bounce = 0u; bounce < depth; ++bounce) {
        ++stats[0];
        if (auto hit = raytracer::hit_world(rd.spheres, scattered_ray); hit) {
            if (auto scattered = raytracer::apply_material(rd, scattered_ray, hit.value()); scattered) {
                std::tie(scattered_ray, energy_absorption) = *scattered;
                attenuation *= energy_absorption;
            }
            else { ++stats[2]; return math::vec3{0}; }
        }
        else {
            ++stats[1];
            return app::background_color(.5f * scattered_ray.unit_direction().y + 1.f) * attenuation;
        }
    }
    ++stats[3];
    return math::vec3{0};
}

} // namespace

extern "C" {

int ref_sizes(std::uint32_t *out)
{
    out[0] = sizeof(math::vec3); out[1] = sizeof(math::u8vec3); out[2] = sizeof(math::ray);
    out[3] = sizeof(primitives::sphere); out[4] = sizeof(primitives::hit);
    out[5] = sizeof(material::lambert); out[6] = sizeof(material::metal);
    out[7] = sizeof(material::dielectric); out[8] = sizeof(material::types);
    out[9] = sizeof(raytracer::camera);
    return 10;
}

// Camera derived state: out = {origin[3], lower_left_corner[3], horizontal[3], vertical[3], lens_radius}
void ref_camera_state(const float *cam, float *out)
{
    auto c = make_camera(cam);
    const math::vec3 *v[4] = {&c->origin, &c->lower_left_corner, &c->horizontal, &c->vertical};
    for (int i = 0; i < 4; ++i) { out[3 * i] = v[i]->x; out[3 * i + 1] = v[i]->y; out[3 * i + 2] = v[i]->z; }
    out[12] = c->lens_radius;
}

// camera::ray(u,v) with the camera's private generator seeded with `seed`.
void ref_camera_ray(const float *cam, std::uint32_t seed, float u, float v, float *origin, float *direction)
{
    auto c = make_camera(cam);
    c->generator = std::mt19937{seed};
    auto r = c->ray(u, v);
    origin[0] = r.origin.x; origin[1] = r.origin.y; origin[2] = r.origin.z;
    direction[0] = r.direction.x; direction[1] = r.direction.y; direction[2] = r.direction.z;
}

// raytracer::intersect for one ray/sphere.  out = {t, p[3], n[3]}; returns 1 on hit.
int ref_intersect(const float *o, const float *d, const float *sph, float tmin, float tmax, float *out)
{
    math::ray ray{math::vec3{o[0], o[1], o[2]}, math::vec3{d[0], d[1], d[2]}};
    primitives::sphere s{math::vec3{sph[0], sph[1], sph[2]}, sph[3], 0};
    auto h = raytracer::intersect(ray, s, tmin, tmax);
    if (!h) return 0;
    out[0] = h->time;
    out[1] = h->position.x; out[2] = h->position.y; out[3] = h->position.z;
    out[4] = h->normal.x; out[5] = h->normal.y; out[6] = h->normal.z;
    return 1;
}

void ref_reflect(const float *I, const float *N, float *out)
{
    auto r = math::reflect(math::vec3{I[0], I[1], I[2]}, math::vec3{N[0], N[1], N[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}

void ref_refract(const float *I, const float *N, float eta, float *out)
{
    auto r = math::refract(math::vec3{I[0], I[1], I[2]}, math::vec3{N[0], N[1], N[2]}, eta);
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}

void ref_normalize(const float *v, float *out)
{
    auto r = math::normalize(math::vec3{v[0], v[1], v[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}

float ref_schlick(float refraction_index, float cosine_theta)
{
    return raytracer::schlick_reflection_probability(refraction_index, cosine_theta);
}

void ref_background(float t, float *out)
{
    auto c = app::background_color(t);
    out[0] = c.x; out[1] = c.y; out[2] = c.z;
}

// gamma_correction + normalize_rgb_to_8bit on n RGB triples.
void ref_gamma8(const float *rgb, std::uint64_t n, float *gamma_out, std::uint8_t *u8_out)
{
    for (std::uint64_t i = 0; i < n; ++i) {
        auto g = app::gamma_correction(math::vec3{rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]});
        if (gamma_out) { gamma_out[3 * i] = g.x; gamma_out[3 * i + 1] = g.y; gamma_out[3 * i + 2] = g.z; }
        auto t = app::normalize_rgb_to_8bit(g);
        u8_out[3 * i] = t.x; u8_out[3 * i + 1] = t.y; u8_out[3 * i + 2] = t.z;
    }
}

// One application of raytracer::apply_material with data.generator seeded by `seed`.
// in: ray o,d; hit p,n,t; material {r,g,b,param,kind}.  out = {o[3], d[3], att[3]}; returns 1 if scattered.
int ref_apply_material(const float *o, const float *d, const float *p, const float *n, float t,
                       const float *mat, std::uint32_t kind, std::uint32_t seed, float *out)
{
    raytracer::data rd;
    scene_view sv{nullptr, nullptr, 0, mat, &kind, 1};
    fill_data(rd, sv);
    rd.generator = std::mt19937{seed};
    math::ray ray{math::vec3{o[0], o[1], o[2]}, math::vec3{d[0], d[1], d[2]}};
    primitives::hit h{math::vec3{p[0], p[1], p[2]}, math::vec3{n[0], n[1], n[2]}, t, 0};
    auto s = raytracer::apply_material(rd, ray, h);
    if (!s) return 0;
    auto &[r, a] = *s;
    out[0] = r.origin.x; out[1] = r.origin.y; out[2] = r.origin.z;
    out[3] = r.direction.x; out[4] = r.direction.y; out[5] = r.direction.z;
    out[6] = a.x; out[7] = a.y; out[8] = a.z;
    return 1;
}

// Closest hit for arbitrary rays through raytracer::hit_world.
// rays: n x {o[3], d[3]}.  idx = sphere index (-1 miss), t, normal[3], position[3].
// The reference hit record carries material_index, not the sphere index; the index is recovered
// as the first sphere whose own intersect() returns exactly the winning time (hit_world's
// stable_partition + first min_element => lowest index on ties, src/raytracer.hxx:107-117).
void ref_closest_hits(const float *spheres, const std::uint32_t *sphere_mat, std::uint32_t n,
                      const float *rays, std::uint64_t nrays,
                      std::int32_t *idx, float *t, float *normal, float *position)
{
    raytracer::data rd;
    std::vector<std::uint32_t> kinds(1, 0);
    float mat[4] = {1, 1, 1, 0};
    // all spheres share a dummy material table sized to the largest index
    std::uint32_t mmax = 0;
    for (std::uint32_t i = 0; i < n; ++i) mmax = std::max(mmax, sphere_mat[i]);
    std::vector<float> mats(4 * (mmax + 1), 1.f);
    std::vector<std::uint32_t> mk(mmax + 1, 0);
    (void)mat; (void)kinds;
    scene_view sv{spheres, sphere_mat, n, mats.data(), mk.data(), mmax + 1};
    fill_data(rd, sv);

    auto constexpr kMAX = std::numeric_limits<float>::max();
    auto constexpr kMIN = .008f;

    for (std::uint64_t i = 0; i < nrays; ++i) {
        const float *r = rays + 6 * i;
        math::ray ray{math::vec3{r[0], r[1], r[2]}, math::vec3{r[3], r[4], r[5]}};
        auto h = raytracer::hit_world(rd.spheres, ray);
        if (!h) {
            idx[i] = -1; t[i] = 0.f;
            for (int k = 0; k < 3; ++k) { normal[3 * i + k] = 0.f; position[3 * i + k] = 0.f; }
            continue;
        }
        std::int32_t found = -1;
        for (std::uint32_t s = 0; s < n && found < 0; ++s) {
            auto hs = raytracer::intersect(ray, rd.spheres[s], kMIN, kMAX);
            if (hs && hs->time == h->time) found = static_cast<std::int32_t>(s);
        }
        idx[i] = found; t[i] = h->time;
        normal[3 * i] = h->normal.x; normal[3 * i + 1] = h->normal.y; normal[3 * i + 2] = h->normal.z;
        position[3 * i] = h->position.x; position[3 * i + 1] = h->position.y; position[3 * i + 2] = h->position.z;
    }
}

// Config-2 pass: primary rays, no jitter (u = x/W, v = y/H), camera aperture as given
// (pass aperture 0 for the deterministic check).  Rows [y0, y0+rows).
void ref_primary_hits(const float *spheres, const std::uint32_t *sphere_mat, std::uint32_t n,
                      const float *cam, std::uint32_t W, std::uint32_t H,
                      std::uint32_t y0, std::uint32_t rows,
                      std::int32_t *idx, float *t, float *normal, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
    std::vector<std::thread> pool;
    std::atomic<std::uint32_t> next{0};
    for (int th = 0; th < nthreads; ++th) pool.emplace_back([&] {
        auto camera = make_camera(cam);
        camera->generator = std::mt19937{1u};
        std::vector<float> rays(6 * static_cast<std::size_t>(W));
        std::vector<float> pos(3 * static_cast<std::size_t>(W));
        for (;;) {
            auto r = next.fetch_add(1);
            if (r >= rows) break;
            auto y = y0 + r;
            auto v = static_cast<float>(y) / static_cast<float>(H);
            for (auto x = 0u; x < W; ++x) {
                auto u = static_cast<float>(x) / static_cast<float>(W);
                auto ray = camera->ray(u, v);
                float *q = rays.data() + 6 * x;
                q[0] = ray.origin.x; q[1] = ray.origin.y; q[2] = ray.origin.z;
                q[3] = ray.direction.x; q[4] = ray.direction.y; q[5] = ray.direction.z;
            }
            auto off = static_cast<std::size_t>(r) * W;
            ref_closest_hits(spheres, sphere_mat, n, rays.data(), W, idx + off, t + off, normal + 3 * off, pos.data());
        }
    });
    for (auto &t_ : pool) t_.join();
}

// Full render of `rows` rows y0, y0 + ystep, y0 + 2 ystep, ... of a W x H image (ystep = 1: the band [y0, y0+rows));
// the restated loop of src/main.cxx:191-215.  A strided row set is how bench.py takes a bounded, unbiased sample of a
// frame in ONE call, so that every worker thread stays busy (rows are dealt to the threads one at a time).
// Per row two generators are seeded: data.generator (jitter + materials) and the camera's private
// one (lens).  Draw order per sample follows main.cxx:199-202: u-jitter, v-jitter, camera.ray, color.
//   sum   : rows*W*3 float, sequential float sum of the spp radiance samples (std::reduce(seq), :205)
//   sumsq : rows*W*3 double or NULL, sum of squares (for the Monte Carlo noise bound)
//   stats : 4 x uint64 (see color_depth) accumulated over all samples
// returns wall seconds.
double ref_render(const float *spheres, const std::uint32_t *sphere_mat, std::uint32_t n,
                  const float *materials, const std::uint32_t *mat_kind, std::uint32_t m,
                  const float *cam, std::uint32_t W, std::uint32_t H,
                  std::uint32_t y0, std::uint32_t rows,
                  std::uint32_t spp, std::uint32_t depth, std::uint64_t seed,
                  float *sum, double *sumsq, std::uint64_t *stats, int nthreads, std::uint32_t ystep)
{
    if (nthreads < 1) nthreads = 1;
    if (ystep < 1) ystep = 1;
    scene_view sv{spheres, sphere_mat, n, materials, mat_kind, m};
    std::atomic<std::uint32_t> next{0};
    std::vector<std::array<std::uint64_t, 4>> tstats(nthreads, {0, 0, 0, 0});
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int th = 0; th < nthreads; ++th) pool.emplace_back([&, th] {
        raytracer::data rd;
        fill_data(rd, sv);
        auto camera = make_camera(cam);
        auto random_distribution = std::uniform_real_distribution{0.f, 1.f};
        for (;;) {
            auto r = next.fetch_add(1);
            if (r >= rows) break;
            auto y = y0 + r * ystep;
            rd.generator = std::mt19937{mix_seed(seed, y, 0)};
            camera->generator = std::mt19937{mix_seed(seed, y, 1)};
            auto v = static_cast<float>(y) / static_cast<float>(H);
            for (auto x = 0u; x < W; ++x) {
                auto u = static_cast<float>(x) / static_cast<float>(W);
                math::vec3 color{0};
                double sq[3] = {0, 0, 0};
                for (auto s = 0u; s < spp; ++s) {
                    auto _u = u + random_distribution(rd.generator) / static_cast<float>(W);
                    auto _v = v + random_distribution(rd.generator) / static_cast<float>(H);
                    auto c = color_depth(rd, camera->ray(_u, _v), depth, tstats[th].data());
                    color = color + c;
                    sq[0] += double(c.x) * c.x; sq[1] += double(c.y) * c.y; sq[2] += double(c.z) * c.z;
                }
                auto off = 3 * (static_cast<std::size_t>(r) * W + x);
                sum[off] = color.x; sum[off + 1] = color.y; sum[off + 2] = color.z;
                if (sumsq) { sumsq[off] = sq[0]; sumsq[off + 1] = sq[1]; sumsq[off + 2] = sq[2]; }
            }
        }
    });
    for (auto &t_ : pool) t_.join();
    auto t1 = std::chrono::steady_clock::now();
    if (stats) for (int k = 0; k < 4; ++k) { stats[k] = 0; for (auto &ts : tstats) stats[k] += ts[k]; }
    return std::chrono::duration<double>(t1 - t0).count();
}

// mean = sum/spp (operator/= , src/main.cxx:207) -> gamma -> 8 bit, exactly the reference tail :207-213.
void ref_resolve(const float *sum, std::uint64_t npix, std::uint32_t spp, std::uint8_t *rgb8)
{
    for (std::uint64_t i = 0; i < npix; ++i) {
        math::vec3 color{sum[3 * i], sum[3 * i + 1], sum[3 * i + 2]};
        color /= static_cast<float>(spp);
        color = app::gamma_correction(color);
        auto t = app::normalize_rgb_to_8bit(std::move(color));
        rgb8[3 * i] = t.x; rgb8[3 * i + 1] = t.y; rgb8[3 * i + 2] = t.z;
    }
}

// app::save_to_file (src/main.cxx:87-101) for the PPM "next" row.
int ref_save_ppm(const char *name, std::uint32_t W, std::uint32_t H, const std::uint8_t *rgb8)
{
    app::data ad;
    ad.width = W; ad.height = H;
    std::vector<math::u8vec3> texels(static_cast<std::size_t>(W) * H);
    std::memcpy(static_cast<void *>(texels.data()), rgb8, texels.size() * 3);
    try { app::save_to_file(name, ad, texels); } catch (...) { return 1; }
    return 0;
}

int ref_hardware_concurrency() { return static_cast<int>(std::thread::hardware_concurrency()); }

} // extern "C"
