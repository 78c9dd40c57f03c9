// main.cpp — the reference's driver (main.cpp:87-216) with everything between LoadModel and save_image handed to the
// B200 library in ONE call: Scene::RenderImage -> pt_render_image = the sample loop (main.cpp:107-160), the
// dispersion statistics and gamma resolve (:162-185), the optional Gauss / median filter (:187-192) and the 8-bit
// conversion of set_pixel (:193-201).  What stays here is the reference's surface: Config flags, Scene::LoadModel,
// the BMP written under the same two names with the same statistics in the file name.
// The progressive "../result.bmp" refresh every `update` samples (main.cpp:141-158) and the time limit
// (main.cpp:111-114) go through Scene::RenderProgressive -> pt_render_progressive, which cuts the loop into blocks
// of samples on device-resident accumulators (same image bit for bit).  Because a whole frame takes a fraction of a
// second here, that path is taken only when asked for: -UPDATE <n> on the command line (the reference's default of
// 32 is not applied silently) or a non-zero -TL.
#include <algorithm>
#include <chrono>
#include <iostream>

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
    PtPostConfig post{};
    post.width = cfg.width; post.height = cfg.height;
    post.gamma_correction = cfg.gamma_correction;
    post.gauss = cfg.gauss; post.median = cfg.median;
    std::vector<uint8_t> rgb8;
    PtDispersion st{};
    int rays_count = cfg.rays_per_pixel;
    auto to_image = [&](const uint8_t *px) {
        for (int y = 0; y < cfg.height; ++y)
            for (int x = 0; x < cfg.width; ++x) {
                const uint8_t *c = &px[3 * (static_cast<size_t>(y) * cfg.width + x)];
                image.set_pixel(x, y, c[0], c[1], c[2]);  // unsampled pixels are 0 = image.clear()
            }
    };
    int rc;
    if ((cfg.update_given && cfg.update != 0) || cfg.time_limit != 0) {
        struct Ctx { bitmap_image *image; decltype(to_image) *fill; } ctx{&image, &to_image};
        const PtPreviewFn refresh = [](void *user, const uint8_t *px, int32_t samples) -> int {
            Ctx *c = static_cast<Ctx *>(user);
            (*c->fill)(px);
            c->image->save_image("../result.bmp");  // main.cpp:155-157
            std::cerr << "Image update" << std::endl << samples << " rays per pixel were sent" << std::endl;
            return 0;
        };
        const long long spent = std::chrono::duration_cast<std::chrono::milliseconds>(clock::now() - start).count();
        const long long limit = cfg.time_limit != 0 ? std::max(1ll, 1000ll * cfg.time_limit - spent) : 0;
        rc = scene.RenderProgressive(abi, post, cfg.update_given ? cfg.update : 0, limit, refresh, &ctx, rgb8, st, rays_count);
    } else {
        rc = scene.RenderImage(abi, post, rgb8, st);
    }
    if (rc) {
        std::cerr << "render failed (" << rc << "): " << pt_last_error() << std::endl;
        return 2;
    }
    std::cerr << rays_count << " rays per pixel were sent" << std::endl;
    to_image(rgb8.data());

    const long long ms = std::chrono::duration_cast<std::chrono::milliseconds>(clock::now() - start).count();
    const time_t t = time(nullptr);
    const tm *now = localtime(&t);
    const std::string name = std::to_string(now->tm_year + 1900) + '-' + std::to_string(now->tm_mon + 1) + '-' +
                             std::to_string(now->tm_mday) + '-' + std::to_string(now->tm_hour) + '-' +
                             std::to_string(now->tm_min) + '-' + std::to_string(now->tm_sec) + "  " + std::to_string(ms) +
                             "   " + std::to_string(rays_count) + " of " + std::to_string(cfg.rays_per_pixel) +
                             "  max_disp " + std::to_string(st.max_dispersion) + "  min_disp " +
                             std::to_string(st.min_dispersion) + "  aver_disp " + std::to_string(st.average_dispersion);
    image.save_image(name + ".bmp");
    image.save_image("../result.bmp");
    return 0;
}
