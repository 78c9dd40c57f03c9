// bitmap_image.h — the subset of the reference's vendored BMP helper (bitmap_image.hpp, A. Partow) that the render
// path uses: ctor(w,h) / ctor(file), clear, set_pixel(x,y,float r,g,b) with float->uchar truncation
// (bitmap_image.hpp:194-207), get_pixel, width/height, save_image (24-bit, bottom-up rows padded to 4 bytes).
// Written from the BMP format, not from that file.
#ifndef PT_B200_HOST_BITMAP_IMAGE_H
#define PT_B200_HOST_BITMAP_IMAGE_H
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

struct rgb_t {
    unsigned char red, green, blue;
};

class bitmap_image {
    unsigned w_ = 0, h_ = 0;
    std::vector<unsigned char> bgr_;  // top-down rows, 3 bytes per pixel, blue first

    static void put32(unsigned char *p, uint32_t v) { p[0] = v & 255; p[1] = (v >> 8) & 255; p[2] = (v >> 16) & 255; p[3] = v >> 24; }
    static uint32_t get32(const unsigned char *p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }

public:
    bitmap_image() = default;
    bitmap_image(unsigned width, unsigned height) : w_(width), h_(height), bgr_((size_t)width * height * 3, 0) {}
    explicit bitmap_image(const std::string &file) { load(file); }

    bool load(const std::string &file) {
        FILE *f = std::fopen(file.c_str(), "rb");
        if (!f) return false;
        unsigned char hdr[54];
        bool ok = std::fread(hdr, 1, 54, f) == 54 && hdr[0] == 'B' && hdr[1] == 'M' && (hdr[28] | (hdr[29] << 8)) == 24;
        if (ok) {
            const uint32_t off = get32(hdr + 10);
            const int32_t W = (int32_t)get32(hdr + 18), H = (int32_t)get32(hdr + 22);
            const unsigned aw = W < 0 ? -W : W, ah = H < 0 ? -H : H;
            const size_t stride = ((size_t)aw * 3 + 3) & ~(size_t)3;
            std::vector<unsigned char> row(stride);
            w_ = aw; h_ = ah;
            bgr_.assign((size_t)aw * ah * 3, 0);
            std::fseek(f, (long)off, SEEK_SET);
            for (unsigned r = 0; r < ah && ok; ++r) {
                ok = std::fread(row.data(), 1, stride, f) == stride;
                const unsigned y = H > 0 ? ah - 1 - r : r;  // positive height = bottom-up file
                if (ok) std::copy(row.begin(), row.begin() + (size_t)aw * 3, bgr_.begin() + (size_t)y * aw * 3);
            }
        }
        std::fclose(f);
        if (!ok) { w_ = h_ = 0; bgr_.clear(); }
        return ok;
    }

    unsigned width() const { return w_; }
    unsigned height() const { return h_; }
    const unsigned char *data() const { return bgr_.data(); }
    void clear(unsigned char v = 0) { std::fill(bgr_.begin(), bgr_.end(), v); }

    void set_pixel(unsigned x, unsigned y, float r, float g, float b) {
        unsigned char *p = &bgr_[((size_t)y * w_ + x) * 3];
        p[0] = static_cast<unsigned char>(b);
        p[1] = static_cast<unsigned char>(g);
        p[2] = static_cast<unsigned char>(r);
    }
    rgb_t get_pixel(unsigned x, unsigned y) const {
        const unsigned char *p = &bgr_[((size_t)y * w_ + x) * 3];
        return rgb_t{p[2], p[1], p[0]};
    }

    bool save_image(const std::string &file) const {
        FILE *f = std::fopen(file.c_str(), "wb");
        if (!f) return false;
        const size_t stride = ((size_t)w_ * 3 + 3) & ~(size_t)3;
        unsigned char hdr[54] = {0};
        hdr[0] = 'B'; hdr[1] = 'M';
        put32(hdr + 2, (uint32_t)(54 + stride * h_));
        put32(hdr + 10, 54);
        put32(hdr + 14, 40);
        put32(hdr + 18, w_);
        put32(hdr + 22, h_);
        hdr[26] = 1; hdr[28] = 24;
        put32(hdr + 34, (uint32_t)(stride * h_));
        std::fwrite(hdr, 1, 54, f);
        std::vector<unsigned char> row(stride, 0);
        for (unsigned r = 0; r < h_; ++r) {
            const unsigned y = h_ - 1 - r;
            std::copy(bgr_.begin() + (size_t)y * w_ * 3, bgr_.begin() + (size_t)(y + 1) * w_ * 3, row.begin());
            std::fwrite(row.data(), 1, stride, f);
        }
        std::fclose(f);
        return true;
    }
};
#endif
