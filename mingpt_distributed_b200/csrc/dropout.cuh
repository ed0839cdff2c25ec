// Counter-based dropout masks (Philox4x32-7), shared by the elementwise dropout kernel and the attention
// kernels: forward and backward regenerate the mask from (seed, site, element index); only the tcgen05 attention
// kernels read it from packed keep words a generator kernel wrote from the same function (below).
//   seed : 64-bit, read from DEVICE memory (so a captured CUDA graph sees a fresh value on every replay)
//   site : which dropout of the network (layer * 4 + {0 attention probs, 1 attention residual, 2 MLP
//          residual}; 0xffff embeddings): mingpt/model.py:140,150,183,215
//   one Philox call = 16 random bytes = 16 consecutive elements: element e of a 32-bit Philox word (e = 0..3) draws the
//   16-bit value (byte e << 8) | byte (e+1)%4 and is KEPT when that value >= thresh = round(p * 65536). Each value is
//   exactly uniform on [0, 65536), so P(keep) = 1 - thresh / 65536; neighbours share one byte, as the high byte of one
//   and the low byte of the other, so their decisions are independent except when a high byte ties with the
//   threshold's (probability 1/256) -- a correlation of order 1e-4 -- for half the Philox calls per element.
#pragma once

#include <stdint.h>

namespace mgb {

// Philox4x32 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11) with 7 rounds: the smallest round
// count the paper reports as passing BigCrush (10 is its safety-margin default); a dropout mask needs no more, and
// every round is 4 of the ~45 instructions of a call.
constexpr int kPhiloxRounds = 7;
__device__ __forceinline__ uint4 philox4x32(uint2 key, uint4 ctr) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < kPhiloxRounds; ++r) {
        const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += W0;
        key.y += W1;
    }
    return ctr;
}

__host__ __device__ __forceinline__ uint32_t dropout_thresh16(float p) {
    float t = p * 65536.0f + 0.5f;
    return t <= 0.f ? 0u : (t >= 65535.f ? 65535u : (uint32_t)t);
}

__device__ __forceinline__ uint4 dropout_rand16(uint64_t seed, uint32_t site, uint64_t group16) {
    return philox4x32(make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)),
                      make_uint4((uint32_t)group16, (uint32_t)(group16 >> 32), site, 0x6d677062u));
}

// The four elements of one Philox word against the threshold, two per packed 16-bit compare: with H = 0x80008000,
// c = (x | H) - (t & ~H) has bit 15 / 31 set iff the low 15 bits compare >=, and the top bits decide otherwise.
// g01: bit 15 = element 0 kept, bit 31 = element 1 kept (other bits arbitrary); g23: elements 2 and 3.
__device__ __forceinline__ void dropout_ge_word(uint32_t w, uint32_t thresh16, uint32_t& g01, uint32_t& g23) {
    const uint32_t H = 0x80008000u, tt = thresh16 | (thresh16 << 16), tl = tt & ~H;
    uint32_t x, y;
    asm("prmt.b32 %0, %1, %1, 0x1201;" : "=r"(x) : "r"(w));       // lanes (b0:b1), (b1:b2)
    asm("prmt.b32 %0, %1, %1, 0x3023;" : "=r"(y) : "r"(w));       // lanes (b2:b3), (b3:b0)
    const uint32_t cx = (x | H) - tl, cy = (y | H) - tl;
    g01 = (x & ~tt) | (~(x ^ tt) & cx);
    g23 = (y & ~tt) | (~(y ^ tt) & cy);
}

// keep flags (bit i = element 8*group + i is kept) for one group of 8 consecutive elements = half a Philox call
__device__ __forceinline__ uint32_t dropout_keep8(uint64_t seed, uint32_t site, uint64_t group, uint32_t thresh16) {
    const uint4 r = dropout_rand16(seed, site, group >> 1);
    const uint32_t w[2] = {(group & 1) ? r.z : r.x, (group & 1) ? r.w : r.y};
    uint32_t m = 0;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        uint32_t g01, g23;
        dropout_ge_word(w[i], thresh16, g01, g23);
        m |= (((g01 >> 15) & 1u) | ((g01 >> 31) << 1) | (((g23 >> 15) & 1u) << 2) | ((g23 >> 31) << 3)) << (4 * i);
    }
    return m;
}

// ---- pre-generated keep words for the tcgen05 attention kernels ----
// The attention kernels are bound by the instruction stream of their softmax / elementwise warps, and eight Philox
// calls per 64 probabilities doubled it. A well-occupied generator kernel produces the same keep bits (same groups,
// same threshold test as dropout_keep8) at the ALU rate instead, as 32-bit words
//     bits[(bh * ceil(T/32) + key/32) * Tp + query],   Tp = T rounded up to 16,
// one word = 32 consecutive keys of one query. Within a word key k sits at bit 16 * (k & 1) + 15 - (k >> 1): that is
// where the generator's packed 16-bit compares leave it (below), and shifted left by s the word carries keys 2s, 2s+1
// in the sign bits of bytes 1 and 3 and keys 2s+16, 2s+17 in those of bytes 0 and 2, which one PRMT each (sign-
// replicate mode) turns into the AND masks of the packed bf16 pairs s and s + 8.
// The backward kernels own one KEY and 16 consecutive queries per thread, so the generator also writes the transpose,
//     bitsT[(bh * ceil(T/32) + query/32) * Kp + key],   Kp = T rounded up to 32,  bit = query % 32,
// behind the query-major words (a 32 x 32 bit transpose across the warp: five shuffle stages).
__host__ __device__ __forceinline__ int dropout_bit_of_key(int k) { return 16 * (k & 1) + 15 - (k >> 1); }
__host__ __device__ __forceinline__ int dropout_key_of_bit(int b) { return 2 * (15 - (b & 15)) + (b >> 4); }
__host__ __device__ __forceinline__ uint64_t dropout_bits_words_q(int64_t BH, int T) {      // query-major part
    return (uint64_t)BH * (uint64_t)((T + 31) >> 5) * (uint64_t)((T + 15) & ~15);
}
__host__ __device__ __forceinline__ uint64_t dropout_bits_words_t(int64_t BH, int T) {      // transposed part
    return (uint64_t)BH * (uint64_t)((T + 31) >> 5) * (uint64_t)(((T + 31) >> 5) * 32);
}

// lane i holds row i of a 32 x 32 bit matrix; returns row i of its transpose (bit j of lane i <- bit i of lane j)
__device__ __forceinline__ uint32_t warp_transpose32(uint32_t x, int lane) {
#pragma unroll
    for (int j = 16; j >= 1; j >>= 1) {
        const uint32_t m = j == 16 ? 0x0000ffffu : j == 8 ? 0x00ff00ffu : j == 4 ? 0x0f0f0f0fu : j == 2 ? 0x33333333u : 0x55555555u;
        const uint32_t y = __shfl_xor_sync(0xffffffffu, x, j);
        x = (lane & j) ? ((x & ~m) | ((y >> j) & m)) : ((x & m) | ((y & m) << j));
    }
    return x;
}

// group0 = the first of the word's four groups of 8 keys (dropout_keep8 numbering)
__device__ __forceinline__ uint32_t dropout_keep32_word(uint64_t seed, uint32_t site, uint64_t group0, uint32_t thresh16) {
    const uint32_t H = 0x80008000u;
    uint32_t word = 0;
    if ((group0 & 1) == 0) {                       // the usual case (T % 16 == 0): two whole Philox calls
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const uint4 r = dropout_rand16(seed, site, (group0 >> 1) + c);
            const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                uint32_t g01, g23;
                dropout_ge_word(w[i], thresh16, g01, g23);       // keys 16c + 4i + {0,1} and + {2,3}
                word |= (g01 & H) >> (8 * c + 2 * i);
                word |= (g23 & H) >> (8 * c + 2 * i + 1);
            }
        }
    } else {                                       // a row of an odd number of groups started inside a Philox call
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint32_t m = dropout_keep8(seed, site, group0 + g, thresh16);
#pragma unroll
            for (int i = 0; i < 8; ++i) word |= ((m >> i) & 1u) << dropout_bit_of_key(8 * g + i);
        }
    }
    return word;
}

// AND masks for the packed pairs s = (keys 2s, 2s+1) and s + 8 = (keys 2s+16, 2s+17) of a keep word, s = 0 .. 7
__device__ __forceinline__ void dropout_pair_masks(uint32_t word, int s, uint32_t& m_s, uint32_t& m_s8) {
    const uint32_t x = word << s;
    asm("prmt.b32 %0, %1, %1, 0xbb99;" : "=r"(m_s) : "r"(x));
    asm("prmt.b32 %0, %1, %1, 0xaa88;" : "=r"(m_s8) : "r"(x));
}

}  // namespace mgb
