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
// the loader's fast float conversion against strtof on every whitespace-separated token of `text` (NUL-terminated):
// counts[0] tokens, [1] taken by the fast path, [2] mismatches (value bits or end of the number); test hook
PTH_API void pth_check_fast_strtof(const char *text, int64_t *counts) { pth_check_fast_strtof_impl(text, counts); }
// Config::set_config on an argv vector; copies the resulting PtConfig out
PTH_API void pth_parse_config(int argc, char **argv, PtConfig *out) {
    Config::get().set_config(argc, argv);
    *out = Config::get().to_abi();
}
}
