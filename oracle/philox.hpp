// ORACLE — TEST INFRASTRUCTURE ONLY. Nothing in the product path may include this file.
//
// Philox4x32-10 counter-based generator (Salmon et al., "Parallel random numbers: as easy as
// 1, 2, 3", SC'11; Random123 v1.09 `philox4x32_R(10, ctr, key)`).  The reference draws from the
// un-vendored crate `rand` (Cargo.toml:10, "*", effectively 0.7.x: OS-seeded ChaCha behind
// thread_rng/random) which is irreproducible by construction, so both the oracle and the
// GPU path replace it by this counter-based stream.  Pinned by Random123's known-answer vectors in
// tests/test_oracle_math.py.
#pragma once
#include <cstdint>

namespace orc {

struct Philox4 {
    uint32_t v[4];
};

inline Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int round = 0; round < 10; ++round) {
        const uint64_t p0 = (uint64_t)M0 * c0;
        const uint64_t p1 = (uint64_t)M1 * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    return Philox4{{c0, c1, c2, c3}};
}

// 24-bit uniform in [0,1): exactly representable in f32, so the f64 oracle (paired mode) and the
// f32 GPU path see the *same* variate.
inline double u24(uint32_t x) { return (double)(x >> 8) * (1.0 / 16777216.0); }
// 53-bit uniform in [0,1) from two words (faithful mode: stands in for rand's f64 uniforms).
inline double u53(uint32_t hi, uint32_t lo) {
    const uint64_t bits = ((uint64_t)hi << 21) ^ (uint64_t)(lo >> 11);  // 32+21 = 53 bits
    return (double)(bits & ((1ull << 53) - 1)) * (1.0 / 9007199254740992.0);
}

}  // namespace orc
