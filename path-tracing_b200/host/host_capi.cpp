// host_capi.cpp — a C view of the host-side Scene (loader output) so that tests and bench.py can drive the C++
// loader through ctypes: load an OBJ with Scene::LoadModel and copy out the arrays behind Scene::abi().
#include <cstring>

#include "scene.h"

void pth_check_fast_strtof_impl(const char *text, int64_t *counts);  // scene.cpp

namespace {
Image g_c, g_c2;
std::vector<std::vector<int>> g_n;
Scene *g_scene = nullptr;
PtScene g_view;
}  // namespace

#define PTH_API __attribute__((visibility("default")))
extern "C" {
// returns the triangle count, or -1 if the file cannot be opened (the reference would exit(1))
PTH_API int pth_load_obj(const char *model_path, const char *model_name) {
    Config::get().model_path = model_path;
    Config::get().model_name = model_name;
    const std::string full = std::string(model_path) + model_name;
    FILE *f = std::fopen(full.c_str(), "rb");
    if (!f) return -1;
    std::fclose(f);
    delete g_scene;
    g_scene = new Scene(g_c, g_c2, g_n);
    g_scene->LoadModel(full);
    g_view = g_scene->abi();
    return g_view.n_triangles;
}
// -CACHE <dir> ("" = off); pth_loaded_from_cache: did the last pth_load_obj take the binary image?
PTH_API void pth_set_cache_dir(const char *dir) { Config::get().cache = dir ? dir : ""; }
PTH_API int pth_loaded_from_cache(void) { return g_scene && g_scene->loaded_from_cache; }
PTH_API int pth_parsed_in_pieces(void) { return g_scene ? g_scene->parsed_in_pieces : 0; }
PTH_API int pth_material_count(void) { return g_scene ? g_view.n_materials : 0; }
PTH_API void pth_copy_scene(float *vertices9, float *planes4, float *squares, int32_t *mat_ids, float *mats10) {
    const size_t n = g_view.n_triangles, m = g_view.n_materials;
    std::memcpy(vertices9, g_view.vertices, 9 * n * sizeof(float));
    std::memcpy(planes4, g_view.planes, 4 * n * sizeof(float));
    std::memcpy(squares, g_view.squares, n * sizeof(float));
    std::memcpy(mat_ids, g_view.material_ids, n * sizeof(int32_t));
    std::memcpy(mats10, g_view.materials, 10 * m * sizeof(float));
}
// extension: (Ni, Tr) per material as parsed; pth_has_refraction = some material would get the refractive lobe
PTH_API void pth_copy_materials_ext(float *ext2) {
    const auto &m = g_scene->Materials();
    for (size_t j = 0; j < m.size(); ++j) {
        ext2[2 * j] = m[j].GetCharacteristics().Ni;
        ext2[2 * j + 1] = m[j].GetCharacteristics().Tr;
    }
}
PTH_API int pth_has_refraction(void) { return g_scene && g_view.materials_ext != nullptr; }
// Triangle's constructor (+ optional SetNormal) on raw arrays, for the loader parity test
PTH_API void pth_triangle_setup(const float *v9, int has_vn, const float *vn3, float *plane4, float *square) {
    Triangle t({ptm::vec4(v9[0], v9[1], v9[2], 1), ptm::vec4(v9[3], v9[4], v9[5], 1), ptm::vec4(v9[6], v9[7], v9[8], 1)}, 0);
    if (has_vn) t.SetNormal(ptm::vec3(vn3[0], vn3[1], vn3[2]));
    const ptm::vec4 p = t.GetPlane();
    plane4[0] = p.x; plane4[1] = p.y; plane4[2] = p.z; plane4[3] = p.w;
    *square = t.GetSquare();
}
// the same for n triangles (has_vn / vn3 may be NULL)
PTH_API void pth_triangle_setup_n(const float *v9, const uint8_t *has_vn, const float *vn3, int64_t n, float *planes4,
                                  float *squares) {
    for (int64_t i = 0; i < n; ++i)
        pth_triangle_setup(v9 + 9 * i, has_vn && vn3 && has_vn[i], vn3 ? vn3 + 3 * i : v9, planes4 + 4 * i, squares + i);
}
// Scene::Intersect on the loaded scene: rays6 = n x (origin, direction) -> Ray objects -> pt_intersect
PTH_API int pth_intersect(const float *rays6, int64_t n, const PtConfig *cfg, int32_t *ids, float *t) {
    if (!g_scene) return PT_EINVAL;
    std::vector<Ray> rays;
    rays.reserve(n);
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays6 + 6 * i;
        rays.emplace_back(ptm::vec4(r[0], r[1], r[2], 1), ptm::vec4(r[3], r[4], r[5], 0), 0, 0, 0);
    }
    std::vector<int32_t> vi;
    std::vector<float> vt;
    const int rc = g_scene->Intersect(rays, *cfg, vi, vt);
    if (rc) return rc;
    std::memcpy(ids, vi.data(), n * sizeof(int32_t));
    std::memcpy(t, vt.data(), n * sizeof(float));
    return 0;
}
// Ray::Reflect (ray.h:45-50) on raw arrays: state = (begin4, direction4, color3, depth) in / out
PTH_API void pth_ray_reflect(float *begin4, float *dir4, float *color3, int *depth, const float *nbegin4, const float *ndir4,
                             const float *tint3) {
    Ray r(ptm::vec4(begin4[0], begin4[1], begin4[2], begin4[3]), ptm::vec4(dir4[0], dir4[1], dir4[2], dir4[3]), *depth, 0, 0);
    r.SetColor(ptm::vec3(color3[0], color3[1], color3[2]));
    r.Reflect(ptm::vec4(nbegin4[0], nbegin4[1], nbegin4[2], nbegin4[3]), ptm::vec4(ndir4[0], ndir4[1], ndir4[2], ndir4[3]),
              ptm::vec3(tint3[0], tint3[1], tint3[2]));
    const ptm::vec4 b = r.GetBegin(), d = r.GetDirection();
    const ptm::vec3 c = r.GetColor();
    begin4[0] = b.x; begin4[1] = b.y; begin4[2] = b.z; begin4[3] = b.w;
    dir4[0] = d.x; dir4[1] = d.y; dir4[2] = d.z; dir4[3] = d.w;
    color3[0] = c.x; color3[1] = c.y; color3[2] = c.z;
    *depth = r.GetDepth();
}
// the loader's fast float conversion against strtof on every whitespace-separated token of `text` (NUL-terminated):
// counts[0] tokens, [1] taken by the fast path, [2] mismatches (value bits or end of the number); test hook
PTH_API void pth_check_fast_strtof(const char *text, int64_t *counts) { pth_check_fast_strtof_impl(text, counts); }
// Config::set_config on an argv vector; copies the resulting PtConfig out
PTH_API void pth_parse_config(int argc, char **argv, PtConfig *out) {
    Config::get().set_config(argc, argv);
    *out = Config::get().to_abi();
}
}
