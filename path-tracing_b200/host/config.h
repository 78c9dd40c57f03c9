// config.h — host-side run parameters with the reference's surface (config.h:7-105): a Meyers singleton with the
// same public fields, defaults, flag names and the same "pairs of argv" parser, so that the reference's main() reads
// unchanged.  to_abi() produces the plain PtConfig that crosses the C-ABI.
#ifndef PT_B200_HOST_CONFIG_H
#define PT_B200_HOST_CONFIG_H

#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>

#include "../../include/pt_render.h"

class Config {
    Config() = default;
    int seed = 42;  // config.h:10

public:
    Config(const Config &) = delete;
    Config &operator=(Config &) = delete;

    int height = 512, width = 512;  // config.h:16-17
    int rays_per_pixel = 20;        // config.h:18
    int max_ray_reflections = 8;    // config.h:19
    int median = 0, gauss = 0;      // config.h:20-21
    float eps = 1e-4f;              // config.h:22
    float error = 0.001f;           // config.h:23
    int update = 32;                // config.h:24
    bool update_given = false;      // -UPDATE was on the command line (main.cpp decides on it, see there)
    float gamma_correction = 1 / 2.2f;
    std::string model_path = "../models/";
    std::string model_name = "Tor.obj";
    std::string skybox = "";
    int time_limit = 0;
    // extensions of the B200 build (defaults = reference behaviour)
    int gpus = 1;
    int accel = PT_ACCEL_AUTO;
    int isect = PT_ISECT_REFERENCE;  // -ISECT 1: the watertight throughput intersector (not the reference's test)
    int rr = 0;  // -RR <depth>: Russian roulette once a path's depth reaches <depth> (0 = off)
    std::string cache = "";  // -CACHE <dir>: keep a binary image of the parsed model there (scene.cpp, LoadModel)

    static Config &get() {
        static Config instance;
        return instance;
    }

    // Flags come in (name, value) pairs starting at argv[1]; a trailing unpaired flag is ignored, unknown flags are
    // skipped, numeric values go through atoi/atof — all as config.h:35-99.
    void set_config(int argc, char *argv[]) {
        for (int i = 1; i + 1 < argc; i += 2) {
            const char *flag = argv[i], *val = argv[i + 1];
            auto is = [flag](const char *name) { return std::strcmp(flag, name) == 0; };
            if (is("--H")) height = std::atoi(val);
            else if (is("--W")) width = std::atoi(val);
            else if (is("-RPP")) rays_per_pixel = std::atoi(val);
            else if (is("-MRR")) max_ray_reflections = std::atoi(val);
            else if (is("-EPS")) eps = static_cast<float>(std::atof(val));
            else if (is("-ERR")) error = static_cast<float>(std::atof(val));
            else if (is("-MEDIAN")) { median = std::atoi(val); gauss = 0; }
            else if (is("-UPDATE")) { update = std::atoi(val); update_given = true; }
            else if (is("-MODEL_PATH")) model_path = val;
            else if (is("-MODEL_NAME")) model_name = val;
            else if (is("-GAUSS")) { gauss = std::atoi(val); median = 0; }
            else if (is("-GAMMA")) gamma_correction = static_cast<float>(std::atof(val));
            else if (is("-SKYBOX")) skybox = val;
            else if (is("-SEED")) seed = std::atoi(val);
            else if (is("-TL")) time_limit = std::atoi(val);
            else if (is("-GPUS")) gpus = std::atoi(val);
            else if (is("-ACCEL")) accel = std::atoi(val);
            else if (is("-RR")) rr = std::atoi(val);
            else if (is("-ISECT")) isect = std::atoi(val);
            else if (is("-CACHE")) cache = val;
        }
    }

    void setSeed(int s) { seed = s; }
    unsigned getSeed() const {  // config.h:101-104
        return seed < 0 ? static_cast<unsigned>(time(nullptr)) : static_cast<unsigned>(seed);
    }

    PtConfig to_abi() const {
        PtConfig c;
        std::memset(&c, 0, sizeof c);
        c.width = width; c.height = height;
        c.rays_per_pixel = rays_per_pixel;
        c.max_ray_reflections = max_ray_reflections;
        c.eps = eps; c.error = error;
        c.seed = getSeed();
        c.sample_begin = 0; c.sample_step = 1;
        c.intersector = isect;
        c.accel = accel;
        c.n_gpus = gpus;
        c.rr_depth = rr;
        return c;
    }
};

#endif
