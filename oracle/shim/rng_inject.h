// TEST INFRASTRUCTURE ONLY (oracle/). Force-included (-include) when the UNMODIFIED reference sources
// are compiled into oracle/_ref/.  It redirects the reference's two random sources
//   material.h:16-20  static std::default_random_engine + std::uniform_real_distribution<float>(0,1)
//   main.cpp:91-92    std::default_random_engine + std::uniform_real_distribution<>(-0.5f,0.5f)
// to the counter-based Philox contract stated in oracle/philox.h, read from a thread-local path
// context that the harness (oracle/ref_harness.cpp) sets before every Scene::TraceRay call.
#ifndef PT_ORACLE_RNG_INJECT_H
#define PT_ORACLE_RNG_INJECT_H
#ifdef __cplusplus
#include <random>
#include <cstdint>

namespace ptshim {
struct PathCtx {
    uint32_t key[2];
    uint32_t pixel, sample, slot;
    uint32_t dim;       // next word of the block to hand out
    uint32_t words[4];
    bool filled;
};
PathCtx &ctx();                 // thread-local
float next_uniform();           // word `dim++` of block (pixel,sample,slot,0) -> (0,1), see oracle/philox.h
}  // namespace ptshim

namespace std {
struct pt_injected_engine {
    pt_injected_engine() {}
    explicit pt_injected_engine(unsigned) {}
};
template <class T = double>
struct pt_injected_uniform {
    T lo, hi;
    pt_injected_uniform(T a, T b) : lo(a), hi(b) {}
    template <class E>
    T operator()(E &) { return lo + (hi - lo) * static_cast<T>(ptshim::next_uniform()); }
};
}  // namespace std

#define default_random_engine pt_injected_engine
#define uniform_real_distribution pt_injected_uniform
#endif
#endif
