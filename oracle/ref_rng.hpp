// ORACLE — TEST INFRASTRUCTURE ONLY (see ref_math.hpp header).
//
// The reference draws every random number from rand 0.8.4's thread_rng() (utils.rs:24-27;
// Cargo.toml:17) — ChaCha12, OS-seeded, one stream per thread, so no two runs of the reference
// agree and no reference output can be reproduced.  rand is not vendored under /root/reference.
// PARITY UNPINNED at this boundary by construction.  Both the oracle and the CUDA path replace
// it by Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11),
// restated here from the published algorithm, keyed so that a sample's draws depend only on
// (seed, pixel, sample index):
//      key     = (seed_lo, seed_hi)
//      counter = (block, a, b, stream)        block = draw_index / 4
//      stream 0: path draws, a = pixel index, b = sample index
//      stream 1: jitter-table shuffle (a = b = 0)
//      stream 2: scene generation     (a = b = 0)
// f32 mapping = top 24 bits * 2^-24, the same as rand's `Standard` distribution for f32.
#pragma once
#include <cstdint>

namespace ref {

inline void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)M0 * c0, p1 = (uint64_t)M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

struct Rng {
    uint32_t key[2];
    uint32_t a, b, stream;
    uint32_t draw = 0;
    uint32_t buf[4];
    Rng(uint64_t seed, uint32_t a_, uint32_t b_, uint32_t stream_) : a(a_), b(b_), stream(stream_) {
        key[0] = (uint32_t)seed;
        key[1] = (uint32_t)(seed >> 32);
    }
    uint32_t next_u32() {
        if ((draw & 3u) == 0u) {
            uint32_t ctr[4] = {draw >> 2, a, b, stream};
            philox4x32_10(ctr, key, buf);
        }
        return buf[(draw++) & 3u];
    }
    // utils.rs:25 `thread_rng().gen::<f32>()`: uniform [0,1) with 24 bits
    float rand() { return (float)(next_u32() >> 8) * (1.0f / 16777216.0f); }
    // utils.rs:26
    float rand_range(float mn, float mx) { return rand() * (mx - mn) + mn; }
};

enum { STREAM_PATH = 0, STREAM_JITTER = 1, STREAM_SCENE = 2 };

}  // namespace ref
