// TEST INFRASTRUCTURE ONLY (oracle/). Harness around the UNMODIFIED reference sources
// (/root/reference/{scene.cpp,scene.h,triangles.h,material.h,ray.h,config.h,timer.h,bitmap_image.hpp}),
// built by oracle/Makefile into oracle/_ref/libptref.so.  Nothing in the product links or loads it.
//
// What is restated here (and only this): the sample loop of main.cpp:110-140 (adaptive-skip test,
// jittered primary ray, while(IsValid) TraceRay), with the two racy std engines replaced by the
// Philox contract of oracle/philox.h through oracle/shim/rng_inject.h.  Everything else
// (Scene::LoadModel, Scene::TraceRay, Triangle::Intersect, Material::Process, Factory, Ray) is the
// reference's own code.
#include <array>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>
#include <chrono>
#include <omp.h>

#define private public  // read-only peek at Scene::triangles_ / Triangle::plane_ for hit IDs and dumps
#include "scene.h"
#undef private

#include "philox.h"

namespace ptshim {
PathCtx &ctx() {
    static thread_local PathCtx c = {{42u, 0u}, 0u, 0u, 0u, 0u, {0u, 0u, 0u, 0u}, false};
    return c;
}
float next_uniform() {
    PathCtx &c = ctx();
    if (!c.filled) {
        const uint32_t ctr[4] = {c.pixel, c.sample, c.slot, 0u};
        pt_philox4x32_10(ctr, c.key, c.words);
        c.filled = true;
    }
    const uint32_t w = c.words[c.dim & 3u];
    ++c.dim;
    return pt_u01(w);
}
static inline void set_slot(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t slot) {
    PathCtx &c = ctx();
    c.key[0] = seed; c.key[1] = 0u;
    c.pixel = pixel; c.sample = sample; c.slot = slot;
    c.dim = 0u; c.filled = false;
}
}  // namespace ptshim

using glm::vec3;
using glm::vec4;
using glm::ivec2;

namespace {
std::vector<std::vector<vec3>> g_color, g_color2;
std::vector<std::vector<int>> g_count;
Scene *g_scene = nullptr;

void reset_maps() {
    const int W = Config::get().width, H = Config::get().height;
    g_color.assign(W, std::vector<vec3>(H, vec3(0)));
    g_color2.assign(W, std::vector<vec3>(H, vec3(0)));
    g_count.assign(W, std::vector<int>(H, 0));
}
void new_scene() {
    delete g_scene;
    reset_maps();
    g_scene = new Scene(g_color, g_color2, g_count);
}
}  // namespace

extern "C" {

void ref_set_config(int width, int height, int spp, int max_depth, float eps, float error, const char *skybox) {
    Config &c = Config::get();
    c.width = width; c.height = height; c.rays_per_pixel = spp; c.max_ray_reflections = max_depth;
    c.eps = eps; c.error = error; c.update = 0;
    c.skybox = skybox ? skybox : "";
}

// Scene::LoadModel (scene.cpp:26-109) on <model_path><model_name>; returns the triangle count.
int ref_scene_load_obj(const char *model_path, const char *model_name) {
    Config::get().model_path = model_path;
    Config::get().model_name = model_name;
    new_scene();
    g_scene->LoadModel(std::string(model_path) + model_name);
    return static_cast<int>(g_scene->triangles_.size());
}

// Same construction path as the `f` branch of LoadModel (scene.cpp:98-104) but from arrays:
// Factory() per material, Triangle(vertices, material), optional SetNormal(vn) if has_vn[i].
int ref_scene_from_arrays(int n_tri, const float *verts9, const unsigned char *has_vn, const float *vn3,
                          const int *mat_ids, int n_mat, const float *mats10) {
    new_scene();
    std::vector<Material> mats;
    for (int m = 0; m < n_mat; ++m) {
        MaterialCharacteristics ch;
        const float *p = mats10 + 10 * m;
        ch.Ke = vec3(p[0], p[1], p[2]); ch.Kd = vec3(p[3], p[4], p[5]); ch.Ks = vec3(p[6], p[7], p[8]); ch.Ns = p[9];
        mats.emplace_back(Factory(ch, g_color, g_color2, g_count));
    }
    for (int i = 0; i < n_tri; ++i) {
        const float *v = verts9 + 9 * i;
        std::array<vec4, 3> tri = {vec4(v[0], v[1], v[2], 1.0f), vec4(v[3], v[4], v[5], 1.0f), vec4(v[6], v[7], v[8], 1.0f)};
        Triangle t(tri, mats[mat_ids[i]]);
        if (has_vn && has_vn[i]) t.SetNormal(vec3(vn3[3 * i], vn3[3 * i + 1], vn3[3 * i + 2]));
        g_scene->AddTriangle(t);
    }
    return n_tri;
}

int ref_scene_size() { return g_scene ? static_cast<int>(g_scene->triangles_.size()) : 0; }

// Dump what Triangle's constructor / SetNormal computed (triangles.h:27-44).
void ref_scene_dump(float *planes4, float *verts9, float *squares) {
    const auto &T = g_scene->triangles_;
    for (size_t i = 0; i < T.size(); ++i) {
        planes4[4 * i + 0] = T[i].plane_.x; planes4[4 * i + 1] = T[i].plane_.y;
        planes4[4 * i + 2] = T[i].plane_.z; planes4[4 * i + 3] = T[i].plane_.w;
        for (int k = 0; k < 3; ++k) {
            verts9[9 * i + 3 * k + 0] = T[i].vertices_[k].x;
            verts9[9 * i + 3 * k + 1] = T[i].vertices_[k].y;
            verts9[9 * i + 3 * k + 2] = T[i].vertices_[k].z;
        }
        squares[i] = T[i].square;
    }
}

// Nearest hit exactly as Scene::TraceRay's loop (scene.cpp:114-120) on an explicit ray set.
// rays6 = (ox,oy,oz,dx,dy,dz); the direction goes through Ray's constructor, i.e. is normalised there
// (ray.h:21-25).  ids[i] = -1 and t[i] = +inf on a miss.
void ref_intersect(const float *rays6, long n, int *ids, float *t, int threads) {
    const auto &T = g_scene->triangles_;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 256)
    for (long r = 0; r < n; ++r) {
        const float *q = rays6 + 6 * r;
        Ray ray(vec4(q[0], q[1], q[2], 1.0f), vec4(q[3], q[4], q[5], 0.0f), 0, ivec2(0, 0));
        float distance = INFINITY;
        int cur = -1;
        for (int i = 0; i < static_cast<int>(T.size()); ++i)
            if (T[i].Intersect(ray, distance)) cur = i;
        ids[r] = cur;
        t[r] = distance;
    }
}

// Primary ray of (pixel x,y; sample s) as main.cpp:126-129 builds it, after Ray's normalisation.
void ref_primary_ray(unsigned seed, int x, int y, int s, float *o3, float *d3) {
    const int W = Config::get().width, H = Config::get().height;
    std::default_random_engine generator(seed);
    std::uniform_real_distribution<> distribution(-0.5f, 0.5f);
    ptshim::set_slot(seed, static_cast<uint32_t>(y * W + x), static_cast<uint32_t>(s), 0u);
    const vec4 direction = vec4((x + distribution(generator)) / W - 0.5f,
                                -(y + distribution(generator)) / H + 0.5f, 1, 0);
    Ray ray(vec4(0, 0, -20, 1), direction, 0, ivec2(x, y));
    o3[0] = ray.GetBegin().x; o3[1] = ray.GetBegin().y; o3[2] = ray.GetBegin().z;
    d3[0] = ray.GetDirection().x; d3[1] = ray.GetDirection().y; d3[2] = ray.GetDirection().z;
}

// The sample loop main.cpp:110-140 for samples [s_begin, s_end) stepping by s_step (spp partition of
// the multi-GPU runs; s_begin=0,s_step=1 is the reference).  Outputs are row-major [y][x]:
// sum3/sum2_3 = color_map/color2_map, count = samples_count.  Returns the number of TraceRay calls.
// `accumulate`!=0 keeps the maps from the previous call (needed when the adaptive rule spans calls).
unsigned long long ref_render_accum(unsigned seed, int s_begin, int s_end, int s_step, int threads, int accumulate,
                                    float *sum3, float *sum2_3, int *count, double *seconds) {
    const int W = Config::get().width, H = Config::get().height;
    if (!accumulate) reset_maps();
    std::default_random_engine generator(seed);
    std::uniform_real_distribution<> distribution(-0.5f, 0.5f);
    std::vector<Ray> rays;
    rays.resize(static_cast<size_t>(H) * W);
    // main.cpp:108 leaves default Rays (valid: depth 0, colour 1) in the array; they are always
    // overwritten at s=0 in the reference.  With s_begin>10 a skipped pixel must not trace one.
    for (auto &r : rays) r.MakeInvalid();
    unsigned long long segments = 0;
    const auto t0 = std::chrono::steady_clock::now();
    for (int rays_count = s_begin; rays_count < s_end; rays_count += s_step) {
#pragma omp parallel for num_threads(threads) schedule(dynamic, 4)
        for (int y = 0; y < H; ++y) {
            for (int x = 0; x < W; ++x) {
                const auto sample_count = static_cast<float>(g_count[x][y]);
                if (rays_count > 10 && sample_count > 0) {
                    const vec3 m = g_color[x][y] / sample_count;
                    vec3 D = (g_color2[x][y] / sample_count - m * m);
                    if ((rays_count % 4) && D.r < Config::get().error && D.g < Config::get().error &&
                        D.b < Config::get().error) {
                        continue;
                    }
                }
                ptshim::set_slot(seed, static_cast<uint32_t>(y * W + x), static_cast<uint32_t>(rays_count), 0u);
                const vec4 direction = vec4((x + distribution(generator)) / W - 0.5f,
                                            -(y + distribution(generator)) / H + 0.5f, 1, 0);
                rays[static_cast<size_t>(y) * W + x] = Ray(vec4(0, 0, -20, 1), direction, 0, ivec2(x, y));
            }
        }
#pragma omp parallel for num_threads(threads) schedule(dynamic, 4) reduction(+ : segments)
        for (int y = 0; y < H; ++y) {
            for (int x = 0; x < W; ++x) {
                Ray &ray = rays[static_cast<size_t>(y) * W + x];
                while (ray.IsValid()) {
                    ptshim::set_slot(seed, static_cast<uint32_t>(y * W + x), static_cast<uint32_t>(rays_count),
                                     1u + static_cast<uint32_t>(ray.GetDepth()));
                    g_scene->TraceRay(ray);
                    ++segments;
                }
            }
        }
    }
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const size_t p = static_cast<size_t>(y) * W + x;
            if (sum3) { sum3[3 * p] = g_color[x][y].x; sum3[3 * p + 1] = g_color[x][y].y; sum3[3 * p + 2] = g_color[x][y].z; }
            if (sum2_3) { sum2_3[3 * p] = g_color2[x][y].x; sum2_3[3 * p + 1] = g_color2[x][y].y; sum2_3[3 * p + 2] = g_color2[x][y].z; }
            if (count) count[p] = g_count[x][y];
        }
    return segments;
}

// One path, segment by segment, for debugging parity: writes up to max_seg records of
// (ox,oy,oz,dx,dy,dz,Tr,Tg,Tb) taken BEFORE each TraceRay call; returns the number of segments.
int ref_trace_path(unsigned seed, int x, int y, int s, int max_seg, float *rec9) {
    const int W = Config::get().width;
    float o[3], d[3];
    ref_primary_ray(seed, x, y, s, o, d);
    Ray ray(vec4(o[0], o[1], o[2], 1.0f), vec4(0, 0, 1, 0), 0, ivec2(x, y));
    // re-create through the same constructor path as the loop (direction normalised once)
    {
        std::default_random_engine generator(seed);
        std::uniform_real_distribution<> distribution(-0.5f, 0.5f);
        ptshim::set_slot(seed, static_cast<uint32_t>(y * W + x), static_cast<uint32_t>(s), 0u);
        const int H = Config::get().height;
        const vec4 direction = vec4((x + distribution(generator)) / W - 0.5f,
                                    -(y + distribution(generator)) / H + 0.5f, 1, 0);
        ray = Ray(vec4(0, 0, -20, 1), direction, 0, ivec2(x, y));
    }
    int n = 0;
    while (ray.IsValid()) {
        if (n < max_seg) {
            float *r = rec9 + 9 * n;
            r[0] = ray.GetBegin().x; r[1] = ray.GetBegin().y; r[2] = ray.GetBegin().z;
            r[3] = ray.GetDirection().x; r[4] = ray.GetDirection().y; r[5] = ray.GetDirection().z;
            r[6] = ray.GetColor().x; r[7] = ray.GetColor().y; r[8] = ray.GetColor().z;
        }
        ptshim::set_slot(seed, static_cast<uint32_t>(y * W + x), static_cast<uint32_t>(s),
                         1u + static_cast<uint32_t>(ray.GetDepth()));
        g_scene->TraceRay(ray);
        ++n;
    }
    return n;
}

int ref_max_threads() { return omp_get_max_threads(); }

// ---- main.cpp's free functions (compiled from the reference's own main.cpp, main() renamed) ----
// images cross as row-major [y][x] float rgb; the reference indexes its maps [x][y]
static std::vector<std::vector<vec3>> to_map(const float *rgb, int W, int H) {
    std::vector<std::vector<vec3>> m(W, std::vector<vec3>(H));
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) m[x][y] = vec3(rgb[3 * (y * W + x)], rgb[3 * (y * W + x) + 1], rgb[3 * (y * W + x) + 2]);
    return m;
}
static void from_map(const std::vector<std::vector<vec3>> &m, int W, int H, float *rgb) {
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            rgb[3 * (y * W + x)] = m[x][y].x; rgb[3 * (y * W + x) + 1] = m[x][y].y; rgb[3 * (y * W + x) + 2] = m[x][y].z;
        }
}
}  // extern "C"
std::vector<std::vector<vec3>> GaussBlur(std::vector<std::vector<vec3>> color_map, float r);          // main.cpp:11-33
std::vector<std::vector<vec3>> MedianFilter(std::vector<std::vector<vec3>> color_map, int window_size);  // main.cpp:49-80
vec4 Refract(vec4 incidence_direction, vec4 normal, float refractive);                                 // main.cpp:35-47
extern "C" {
void ref_gauss_blur(int W, int H, const float *rgb_in, float r, float *rgb_out) {
    Config::get().width = W; Config::get().height = H;
    from_map(GaussBlur(to_map(rgb_in, W, H), r), W, H, rgb_out);
}
void ref_median_filter(int W, int H, const float *rgb_in, int window, float *rgb_out) {
    Config::get().width = W; Config::get().height = H;
    from_map(MedianFilter(to_map(rgb_in, W, H), window), W, H, rgb_out);
}
void ref_refract(const float *d3, const float *n3, float refractive, float *out3) {
    const vec4 r = Refract(vec4(d3[0], d3[1], d3[2], 0.0f), vec4(n3[0], n3[1], n3[2], 0.0f), refractive);
    out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}

}  // extern "C"
