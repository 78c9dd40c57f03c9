// main.cpp — the reference's driver (main.cpp:87-216) with its sample loop (lines 107-160) replaced by ONE call
// into the B200 library: Scene::Render -> pt_render_accum.  Everything around the loop keeps the reference's
// behaviour: Config flags, Scene::LoadModel, dispersion statistics, gamma resolve, optional Gauss / median filter,
// BMP output under the same two names.  The progressive "../result.bmp" refresh every `update` samples
// (main.cpp:141-158) and the time limit (main.cpp:111-114) have no equivalent: the whole loop is one kernel launch.
#include <chrono>
#include <iostream>

#include "postprocess.h"
#include "scene.h"

int main(int argc, char *argv[]) {
    using clock = std::chrono::steady_clock;
    const auto start = clock::now();
    Config &cfg = Config::get();
    cfg.set_config(argc, argv);

    Image color_map(cfg.width, std::vector<ptm::vec3>(cfg.height));
    Image color2_map(cfg.width, std::vector<ptm::vec3>(cfg.height));
    std::vector<std::vector<int>> samples_count(cfg.width, std::vector<int>(cfg.height, 0));

    Scene scene(color_map, color2_map, samples_count);
    scene.LoadModel(cfg.model_path + cfg.model_name);

    bitmap_image image(cfg.width, cfg.height);
    image.clear();

    const PtConfig abi = cfg.to_abi();
    const int rc = scene.Render(abi);
    if (rc) {
        std::cerr << "pt_render failed (" << rc << "): " << pt_last_error() << std::endl;
        return 2;
    }
    std::cerr << cfg.rays_per_pixel << " rays per pixel were sent" << std::endl;

    const DispersionStats st = ResolveAndStats(color_map, color2_map, samples_count, cfg.width, cfg.height,
                                               cfg.gamma_correction, cfg.rr > 0 ? &scene.Weights() : nullptr);
    if (cfg.gauss) color_map = GaussBlur(color_map, cfg.gauss, cfg.width, cfg.height);
    if (cfg.median) color_map = MedianFilter(color_map, cfg.median, cfg.width, cfg.height);
    for (int y = 0; y < cfg.height; ++y)
        for (int x = 0; x < cfg.width; ++x)
            if (samples_count[x][y]) image.set_pixel(x, y, color_map[x][y].x, color_map[x][y].y, color_map[x][y].z);

    const long long ms = std::chrono::duration_cast<std::chrono::milliseconds>(clock::now() - start).count();
    const time_t t = time(nullptr);
    const tm *now = localtime(&t);
    const std::string name = std::to_string(now->tm_year + 1900) + '-' + std::to_string(now->tm_mon + 1) + '-' +
                             std::to_string(now->tm_mday) + '-' + std::to_string(now->tm_hour) + '-' +
                             std::to_string(now->tm_min) + '-' + std::to_string(now->tm_sec) + "  " + std::to_string(ms) +
                             "   " + std::to_string(cfg.rays_per_pixel) + " of " + std::to_string(cfg.rays_per_pixel) +
                             "  max_disp " + std::to_string(st.max_dispersion) + "  min_disp " +
                             std::to_string(st.min_dispersion) + "  aver_disp " + std::to_string(st.average_dispersion);
    image.save_image(name + ".bmp");
    image.save_image("../result.bmp");
    return 0;
}
