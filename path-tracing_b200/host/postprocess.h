// postprocess.h — the host-side steps after the sample loop, with the reference's behaviour (main.cpp:11-33 GaussBlur,
// :49-80 MedianFilter, :162-201 dispersion statistics + gamma resolve).  O(pixels) work that stays on the host in this
// round (SURVEY.md §8f ranks 1-2 move it to the GPU later).
#ifndef PT_B200_HOST_POSTPROCESS_H
#define PT_B200_HOST_POSTPROCESS_H
#include <algorithm>
#include <cmath>
#include <vector>

#include "material.h"
#include "vec.h"

using Image = std::vector<std::vector<ptm::vec3>>;  // indexed [x][y] like the reference's color_map

struct DispersionStats {
    float max_dispersion = 0, min_dispersion = INFINITY, average_dispersion = 0;
};

// main.cpp:162-185: per-pixel D = E[c^2] - E[c]^2 summed over channels; pixels without samples add 1 to the average;
// color_map becomes pow(sum / n, gamma) * 255 in place.
inline DispersionStats ResolveAndStats(Image &color_map, const Image &color2_map,
                                       const std::vector<std::vector<int>> &samples_count, int width, int height,
                                       float gamma, const std::vector<float> *divisor = nullptr) {
    DispersionStats st;
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            if (!samples_count[x][y]) { st.average_dispersion += 1; continue; }
            // Russian-roulette renders divide by the weight sum (row-major [y][x]) instead of the count
            const float n = divisor ? (*divisor)[static_cast<size_t>(y) * width + x] : static_cast<float>(samples_count[x][y]);
            const ptm::vec3 s = color_map[x][y], s2 = color2_map[x][y];
            const float mx = s.x / n, my = s.y / n, mz = s.z / n;
            const float disp = (s2.x / n - mx * mx) + (s2.y / n - my * my) + (s2.z / n - mz * mz);
            st.max_dispersion = std::max(st.max_dispersion, disp);
            st.min_dispersion = std::min(st.min_dispersion, disp);
            st.average_dispersion += disp;
            color_map[x][y] = ptm::vec3(std::pow(mx, gamma) * 255.0f, std::pow(my, gamma) * 255.0f, std::pow(mz, gamma) * 255.0f);
        }
    st.average_dispersion /= width * height;
    return st;
}

// main.cpp:11-33: normalised Gaussian of radius ceil(2.57 r), clamp-to-edge, result rounded per channel
inline Image GaussBlur(const Image &src, int r, int width, int height) {
    const int rs = static_cast<int>(std::ceil(r * 2.57));
    Image out(width, std::vector<ptm::vec3>(height));
    for (int i = 0; i < height; ++i)
        for (int j = 0; j < width; ++j) {
            float ax = 0, ay = 0, az = 0, wsum = 0;
            for (int iy = i - rs; iy <= i + rs; ++iy)
                for (int ix = j - rs; ix <= j + rs; ++ix) {
                    const int x = std::min(width - 1, std::max(0, ix)), y = std::min(height - 1, std::max(0, iy));
                    const float dsq = static_cast<float>((ix - j) * (ix - j) + (iy - i) * (iy - i));
                    const float w = std::exp(-dsq / (2.0f * r * r)) / (pi * 2.0f * r * r);
                    ax += src[x][y].x * w; ay += src[x][y].y * w; az += src[x][y].z * w;
                    wsum += w;
                }
            out[j][i] = ptm::vec3(std::round(ax / wsum), std::round(ay / wsum), std::round(az / wsum));
        }
    return out;
}

// main.cpp:49-80: per channel, the (2w+1)^2 window with clamp-to-edge coordinates is sorted and element w*w/2 is
// taken (the reference's index, which is not the true median — kept)
inline Image MedianFilter(const Image &src, int w, int width, int height) {
    Image out(width, std::vector<ptm::vec3>(height));
    std::vector<float> win[3];
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            for (auto &v : win) v.clear();
            for (int dx = -w; dx <= w; ++dx)
                for (int dy = -w; dy <= w; ++dy) {
                    const int i = std::max(std::min(dx + x, width - 1), 0), j = std::max(std::min(dy + y, height - 1), 0);
                    win[0].push_back(src[i][j].x); win[1].push_back(src[i][j].y); win[2].push_back(src[i][j].z);
                }
            for (auto &v : win) std::sort(v.begin(), v.end());
            const size_t k = static_cast<size_t>(w * w / 2);
            out[x][y] = ptm::vec3(win[0][k], win[1][k], win[2][k]);
        }
    return out;
}
#endif
