// material.h — MaterialCharacteristics / Material with the reference's names (material.h:22-56), as DATA.
// The reference's Factory (material.h:58-106) hides (Ke, Kd, Ks, Ns) inside std::function closures that capture the
// frame buffers; a device cannot read those, so Material here keeps the characteristics and reports which closures
// the Factory would have built.  The closures themselves (emissive / mirror / diffuse) run on the GPU
// (csrc/pt_kernels.cuh, trace_path).
#ifndef PT_B200_HOST_MATERIAL_H
#define PT_B200_HOST_MATERIAL_H
#include "vec.h"

const float pi = 3.141593f;  // material.h:12

struct MaterialCharacteristics {
    ptm::vec3 Ke, Kd, Ks;
    float Ns = 0;
    // extension (not parsed by the reference): index of refraction `Ni` and transmission `Tr` (= 1 - `d`) of the MTL
    // block; Tr > 0 && Ni > 0 adds the refractive lobe built on the reference's unused Refract (main.cpp:35-47)
    float Ni = 0, Tr = 0;
};

class Material {
    MaterialCharacteristics ch_;

public:
    Material() = default;
    explicit Material(const MaterialCharacteristics &ch) : ch_(ch) {}
    const MaterialCharacteristics &GetCharacteristics() const { return ch_; }
    // the closure list of material.h:66-104
    bool IsEmissive() const { return ch_.Ke != ptm::vec3(0); }
    bool HasMirror() const { return !IsEmissive() && ch_.Ns != 0 && ch_.Ks != ptm::vec3(0); }
    bool HasDiffuse() const { return !IsEmissive() && (1 - ch_.Ns / 1000 > 0); }
    float MirrorChance() const { return ch_.Ns / 1000; }
    bool HasRefract() const { return !IsEmissive() && ch_.Tr > 0 && ch_.Ni > 0; }  // extension
};

inline Material Factory(const MaterialCharacteristics &characteristics) { return Material(characteristics); }

#endif
