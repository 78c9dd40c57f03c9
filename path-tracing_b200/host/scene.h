// scene.h — Scene with the reference's surface (scene.h:7-24): constructed on the three frame accumulators,
// LoadModel(path), AddTriangle(triangle).  TraceRay(Ray&) — one path segment on the CPU — is replaced by Render(),
// which hands the whole sample loop of main.cpp:107-160 to the GPU through the C-ABI (pt_render_accum) and fills the
// same three accumulators, indexed [x][y] as the reference's.
#ifndef PT_B200_HOST_SCENE_H
#define PT_B200_HOST_SCENE_H
#include <string>
#include <vector>

#include "bitmap_image.h"
#include "triangles.h"

using Image = std::vector<std::vector<ptm::vec3>>;  // indexed [x][y] like the reference's color_map

class Scene {
    std::vector<Triangle> triangles_;
    std::vector<std::vector<ptm::vec3>> &color_map_;
    std::vector<std::vector<ptm::vec3>> &color2_map_;
    std::vector<std::vector<int>> &samples_count_;
    std::vector<Material> materials_;
    bitmap_image skybox_;
    // flat arrays behind abi()
    mutable std::vector<float> vertices_, planes_, squares_, mats_, mats_ext_;
    std::vector<float> sum_, sum2_;    // flat copies of the last Render's accumulators, row-major [y][x]
    std::vector<int32_t> count_;
    bool rr_ = false;                  // the last Render used Russian roulette: the divisor is weight_
    std::vector<float> weight_;  // per-pixel divisor of the last Render, row-major [y][x] (== samples_count without -RR)
    mutable std::vector<int32_t> mat_ids_;

    // binary image of a parsed model (extension): see scene.cpp
    bool LoadCache(const std::string &cache_file, const std::string &obj_path);
    void SaveCache(const std::string &cache_file, const std::string &obj_path) const;
    std::vector<std::string> mtl_files_;  // the MTL files LoadModel read, for the cache's validity record
    // the OBJ text (`size` bytes followed by one ' ') parsed by several threads; false = not attempted or the cuts did
    // not line up with the statements, nothing changed: the caller parses sequentially (scene.cpp)
    bool LoadModelParallel(const char *text, size_t size);

public:
    bool loaded_from_cache = false;  // the last LoadModel took the binary image instead of parsing
    int parsed_in_pieces = 0;        // threads the last LoadModel parsed the OBJ with (0 = the sequential loop)
    Scene(std::vector<std::vector<ptm::vec3>> &color_map, std::vector<std::vector<ptm::vec3>> &color2_map,
          std::vector<std::vector<int>> &samples_count);
    void LoadModel(std::string path);
    void AddTriangle(Triangle triangle);
    int AddMaterial(const MaterialCharacteristics &ch);

    size_t TriangleCount() const { return triangles_.size(); }
    const std::vector<Triangle> &Triangles() const { return triangles_; }
    const std::vector<Material> &Materials() const { return materials_; }

    PtScene abi() const;  // plain-data view for the C-ABI; valid until the scene is modified
    // Nearest hit of every ray, exactly the loop scene.cpp:114-120 (pt_intersect): ids[i] = index into Triangles()
    // (-1 = miss), t[i] = hit distance (+inf = miss).  Uses cfg.eps and cfg.accel.  Returns a PT_* code.
    int Intersect(const std::vector<Ray> &rays, const PtConfig &cfg, std::vector<int32_t> &ids, std::vector<float> &t) const;
    int Render(const PtConfig &cfg);  // the loop main.cpp:107-160 on the GPU; returns a PT_* code
    // per-pixel divisor for the resolve, row-major [y][x]: samples_count, or the weight sum of a Russian-roulette render
    const std::vector<float> &Weights() const { return weight_; }
    // main.cpp:162-201 on the GPU (pt_postprocess) from the accumulators of the last Render: dispersion statistics,
    // gamma resolve, Gauss / median filter, 8-bit conversion.  rgb8 = H*W*3, row-major [y][x]
    int Resolve(const PtPostConfig &post, std::vector<uint8_t> &rgb8, PtDispersion &stats) const;
    // Render + Resolve without the round trip through host accumulators (pt_render_image)
    int RenderImage(const PtConfig &cfg, const PtPostConfig &post, std::vector<uint8_t> &rgb8, PtDispersion &stats);
    // the same with the reference's refresh every `update` samples and its time limit (pt_render_progressive);
    // samples_done = the loop's rays_count at exit
    int RenderProgressive(const PtConfig &cfg, const PtPostConfig &post, int update, long long time_limit_ms,
                          PtPreviewFn on_update, void *user, std::vector<uint8_t> &rgb8, PtDispersion &stats,
                          int &samples_done);
};
#endif
