// triangles.h — Triangle with the reference's surface (triangles.h:19-74) minus Intersect, which is the device's
// job (csrc/pt_trace.cuh).  The constructor precomputes the unit-normal plane and the parallelogram area exactly as
// triangles.h:27-44; SetNormal replaces the plane by a caller-supplied normal (the `vn` override, scene.cpp:102-104).
// Build with -ffp-contract=off: these values cross the ABI and must equal the reference's bit for bit.
#ifndef PT_B200_HOST_TRIANGLES_H
#define PT_B200_HOST_TRIANGLES_H
#include <array>

#include "material.h"
#include "ray.h"

class Triangle {
    ptm::vec4 plane_;
    int material_id_ = 0;
    std::array<ptm::vec3, 3> vertices_;
    float square = 0;

public:
    Triangle(const std::array<ptm::vec4, 3> &vertices, int material_id) : material_id_(material_id) {
        for (size_t i = 0; i < 3; ++i) vertices_[i] = vertices[i].xyz();
        const ptm::vec3 AB = vertices_[1] - vertices_[0], AC = vertices_[2] - vertices_[0];
        const ptm::vec3 n = ptm::cross(AB, AC);
        SetNormal(n);
        square = ptm::length(n);
    }
    // a triangle exactly as a previous constructor (+ SetNormal) call left it: the binary scene cache (scene.cpp)
    Triangle(const float v9[9], const float plane4[4], float sq, int material_id)
        : plane_(plane4[0], plane4[1], plane4[2], plane4[3]), material_id_(material_id), square(sq) {
        for (size_t i = 0; i < 3; ++i) vertices_[i] = ptm::vec3(v9[3 * i], v9[3 * i + 1], v9[3 * i + 2]);
    }
    ptm::vec4 GetNormal() const { return ptm::vec4(plane_.x, plane_.y, plane_.z, 0); }
    ptm::vec4 GetPlane() const { return plane_; }
    void SetNormal(ptm::vec3 normal) {
        normal = ptm::normalize(normal);
        plane_ = ptm::vec4(normal, -ptm::dot(normal, vertices_[0]));
    }
    int GetMaterialId() const { return material_id_; }
    const std::array<ptm::vec3, 3> &GetVertices() const { return vertices_; }
    float GetSquare() const { return square; }
};
#endif
