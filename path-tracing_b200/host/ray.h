// ray.h — Ray with the reference's surface (ray.h:11-57).  On the device the same state lives in registers / 64 B of
// HBM per path (origin, unit direction, depth, pixel, RGB throughput); on the host this type is what
// Scene::Intersect(std::vector<Ray>) takes — explicit ray sets for the nearest-hit query (pt_intersect) — and Reflect
// is the reference's bounce step for callers that drive a path by hand.
#ifndef PT_B200_HOST_RAY_H
#define PT_B200_HOST_RAY_H
#include "config.h"
#include "vec.h"

class Ray {
    ptm::vec4 begin_, direction_;
    ptm::vec4 given_direction_;  // the direction as handed in, before normalisation: what Scene::Intersect passes on
                                 // (pt_intersect normalises as the constructor does; normalising twice moves the last bit)
    int depth_ = 0;
    int coord_x_ = 0, coord_y_ = 0;
    ptm::vec3 color_ = ptm::vec3(1);

    static ptm::vec4 unit(const ptm::vec4 &d) {  // glm::normalize(vec4): (x+y)+(z+w) dot, ray.h:23
        const float xx = d.x * d.x, yy = d.y * d.y, zz = d.z * d.z, ww = d.w * d.w;
        const float inv = 1.0f / std::sqrt((xx + yy) + (zz + ww));
        return ptm::vec4(d.x * inv, d.y * inv, d.z * inv, d.w * inv);
    }

public:
    Ray() = default;
    Ray(ptm::vec4 begin, ptm::vec4 direction, int depth, int px, int py)
        : begin_(begin), direction_(unit(direction)), given_direction_(direction), depth_(depth), coord_x_(px), coord_y_(py) {}
    ptm::vec4 GetBegin() const { return begin_; }
    void SetBegin(ptm::vec4 b) { begin_ = b; }
    ptm::vec4 GetDirection() const { return direction_; }
    void SetDirection(ptm::vec4 d) { direction_ = unit(d); given_direction_ = d; }
    ptm::vec4 GetGivenDirection() const { return given_direction_; }
    int GetDepth() const { return depth_; }
    void SetDepth(int d) { depth_ = d; }
    int GetX() const { return coord_x_; }
    int GetY() const { return coord_y_; }
    ptm::vec3 GetColor() const { return color_; }
    void SetColor(ptm::vec3 c) { color_ = c; }
    // ray.h:45-50: new origin, re-normalised direction, throughput times the closure's tint, one bounce deeper
    void Reflect(ptm::vec4 begin, ptm::vec4 direction, ptm::vec3 color) {
        begin_ = begin;
        direction_ = unit(direction);
        given_direction_ = direction;
        color_ = color_ * color;
        ++depth_;
    }
    bool IsValid() const { return depth_ < Config::get().max_ray_reflections && color_ != ptm::vec3(0); }
    void MakeInvalid() { depth_ = Config::get().max_ray_reflections; }
};
#endif
