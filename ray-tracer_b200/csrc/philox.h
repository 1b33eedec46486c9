// Philox4x32-10 counter-based RNG, usable from host C++ and CUDA device code.
//
// Replaces the reference's `rand` crate call sites (src/util.rs:10,28, src/material.rs:164,
// src/volume.rs:58,80, examples/book-one.rs:70-72) by a stream keyed by (pixel, sample) with the
// counter carrying (bounce, dimension, seed) — see DESIGN.md "RNG contract".
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#else
#define RT_HD inline
#endif

namespace rtb {

struct Philox4 {
    uint32_t x, y, z, w;
};

RT_HD uint32_t rt_mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

RT_HD Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = rt_mulhi32(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = rt_mulhi32(M1, c2), lo1 = M1 * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return Philox4{c0, c1, c2, c3};
}

// 24-bit uniform in [0,1): exact in f32.
RT_HD float u24f(uint32_t v) { return (float)(v >> 8) * (1.0f / 16777216.0f); }

// Dimension tags of the RNG contract (counter word 1).
enum : uint32_t { RT_DIM_CAMERA = 0u, RT_DIM_SHADE = 1u, RT_DIM_MEDIUM = 0x80000000u };

}  // namespace rtb
