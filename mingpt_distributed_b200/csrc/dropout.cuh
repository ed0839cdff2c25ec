// Counter-based dropout masks (Philox4x32-10), shared by the elementwise dropout kernel and the attention
// kernels. No mask is ever stored: forward and backward regenerate it from (seed, site, element index).
//   seed : 64-bit, read from DEVICE memory (so a captured CUDA graph sees a fresh value on every replay)
//   site : which dropout of the network (layer * 4 + {0 attention probs, 1 attention residual, 2 MLP
//          residual}; 0xffff embeddings): mingpt/model.py:140,150,183,215
//   one Philox call = 128 random bits = 8 elements x 16 bits; an element is KEPT when its 16 bits >= thresh,
//   thresh = round(p * 65536), so P(keep) = 1 - thresh / 65536.
#pragma once

#include <stdint.h>

namespace mgb {

__device__ __forceinline__ uint4 philox4x32_10(uint2 key, uint4 ctr) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
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

// keep flags (bit i = element 8*group + i is kept) for one group of 8 consecutive elements
__device__ __forceinline__ uint32_t dropout_keep8(uint64_t seed, uint32_t site, uint64_t group, uint32_t thresh16) {
    const uint4 r = philox4x32_10(make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)),
                                  make_uint4((uint32_t)group, (uint32_t)(group >> 32), site, 0x6d677062u));
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
    uint32_t m = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        m |= ((w[i] & 0xffffu) >= thresh16 ? 1u : 0u) << (2 * i);
        m |= ((w[i] >> 16) >= thresh16 ? 1u : 0u) << (2 * i + 1);
    }
    return m;
}

}  // namespace mgb
