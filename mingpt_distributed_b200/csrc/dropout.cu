// Elementwise dropout for the three non-attention dropout sites of the reference (embeddings
// mingpt/model.py:215,231; after c_proj :140,167; after the MLP :183), fused with the residual add that
// follows it:   y = (res ? res : 0) + (keep ? x / (1 - p) : 0)
// The mask is a pure function of (seed, site, element index): the backward pass calls the same kernel on the
// incoming gradient (res = NULL) and gets the same mask, nothing is stored.
#include "common.cuh"
#include "dropout.cuh"
#include "mingpt_b200.h"

namespace mgb {

__global__ void __launch_bounds__(256)
dropout_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* res, __nv_bfloat16* y, int64_t n,
               uint32_t thresh, float inv_keep, const uint64_t* __restrict__ seed_dev, uint32_t site) {
    const uint64_t seed = *seed_dev;
    const int64_t groups = (n + 7) >> 3;
    for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t keep = dropout_keep8(seed, site, (uint64_t)g, thresh);
        const int64_t i0 = g * 8;
        if (i0 + 8 <= n) {
            const uint4 vx = ldg128(x + i0);
            uint4 vr = make_uint4(0, 0, 0, 0);
            if (res) vr = ldg128(res + i0);
            const uint32_t wx[4] = {vx.x, vx.y, vx.z, vx.w}, wr[4] = {vr.x, vr.y, vr.z, vr.w};
            uint32_t o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 fx = unpack_bf16x2(wx[e]), fr = unpack_bf16x2(wr[e]);
                const float a = ((keep >> (2 * e)) & 1u) ? fx.x * inv_keep : 0.f;
                const float b = ((keep >> (2 * e + 1)) & 1u) ? fx.y * inv_keep : 0.f;
                o[e] = pack_bf16x2(a + fr.x, b + fr.y);
            }
            stg128(y + i0, make_uint4(o[0], o[1], o[2], o[3]));
        } else {
            for (int e = 0; i0 + e < n; ++e) {
                float v = ((keep >> e) & 1u) ? __bfloat162float(x[i0 + e]) * inv_keep : 0.f;
                if (res) v += __bfloat162float(res[i0 + e]);
                y[i0 + e] = __float2bfloat16(v);
            }
        }
    }
}

// The backward of a residual-branch dropout, y = keep ? dy / (1-p) : 0 over a [rows, cols] gradient, which is also
// the output gradient of the linear layer in front of the dropout: its bias gradient = column sums of y come out of
// the same pass (colsum[c] += sum over rows). The block is (cols/8) x rows_per_block threads; a thread keeps its 8
// columns for the whole grid-stride loop, so the sums stay in registers until one shared-memory and one global atomic
// per column and block (few, fat CTAs: those tails are the fixed cost).
// W = 16 columns per thread when cols % 16 == 0 (a whole Philox call = 16 elements per row and thread), else 8 (half).
template <int W>
__global__ void __launch_bounds__(512)
dropout_colsum_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t rows, int cols,
                      uint32_t thresh, float inv_keep, const uint64_t* __restrict__ seed_dev, uint32_t site,
                      float* __restrict__ colsum) {
    extern __shared__ float dcs_smem[];              // [cols]
    constexpr int NV = W / 8;                        // 16-byte vectors per thread and row
    constexpr int U = 4 / NV;                        // rows in flight per thread: 64 bytes either way
    const int nch = cols / W;
    const int rpb = blockDim.x / nch;
    const int ch = threadIdx.x % nch, r = threadIdx.x / nch;
    for (int c = threadIdx.x; c < cols; c += blockDim.x) dcs_smem[c] = 0.f;
    __syncthreads();
    const uint64_t seed = *seed_dev;
    float acc[W];
#pragma unroll
    for (int e = 0; e < W; ++e) acc[e] = 0.f;
    if (r < rpb) {
        const int64_t step = (int64_t)gridDim.x * rpb;
        // the kernel is a stream (2 x M x C x 2 bytes): 64 bytes in flight per thread (30 KB per SM was not enough)
        for (int64_t row0 = (int64_t)blockIdx.x * rpb + r; row0 < rows; row0 += U * step) {
            uint4 v[U][NV];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t row = row0 + u * step;
                if (row < rows) {
#pragma unroll
                    for (int n = 0; n < NV; ++n) v[u][n] = __ldcs(reinterpret_cast<const uint4*>(x + row * cols + ch * W + n * 8));
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t row = row0 + u * step;
                if (row >= rows) continue;
                // 16 consecutive elements = one Philox call (dropout.cuh); the compare leaves the keep flags in the sign bits
                // of two bytes, which PRMT spreads over the packed bf16 pair
                const uint64_t e0 = (uint64_t)(row * cols + ch * W);              // first element of this thread in the row
                const uint4 rn = dropout_rand16(seed, site, e0 >> 4);
                const bool hi = (W == 8) && (e0 & 8);
                const uint32_t rw[4] = {hi ? rn.z : rn.x, hi ? rn.w : rn.y, rn.z, rn.w};
#pragma unroll
                for (int n = 0; n < NV; ++n) {
                    const uint32_t w[4] = {v[u][n].x, v[u][n].y, v[u][n].z, v[u][n].w};
                    uint32_t o[4];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t gg[2];
                        dropout_ge_word(rw[2 * n + h], thresh, gg[0], gg[1]);
#pragma unroll
                        for (int t = 0; t < 2; ++t) {
                            const int e = 2 * h + t;
                            uint32_t m;
                            asm("prmt.b32 %0, %1, %1, 0xbb99;" : "=r"(m) : "r"(gg[t]));
                            const float2 f = mul2(unpack_bf16x2(w[e] & m), make_float2(inv_keep, inv_keep));
                            o[e] = pack_bf16x2(f.x, f.y);
                            const float2 q = unpack_bf16x2(o[e]);          // the sums see what the GEMMs will read
                            acc[n * 8 + 2 * e] += q.x;
                            acc[n * 8 + 2 * e + 1] += q.y;
                        }
                    }
                    stg128(y + row * cols + ch * W + n * 8, make_uint4(o[0], o[1], o[2], o[3]));
                }
            }
        }
#pragma unroll
        for (int e = 0; e < W; ++e) atomicAdd(&dcs_smem[ch * W + e], acc[e]);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < cols; c += blockDim.x) atomicAdd(colsum + c, dcs_smem[c]);
}

}  // namespace mgb

using namespace mgb;

extern "C" int mgb_dropout_bf16_colsum(const void* x, void* y, int64_t rows, int32_t cols, float p, const uint64_t* seed_dev,
                                       uint32_t site, float* colsum, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (rows < 0 || cols <= 0 || (cols % 8) || cols > 4096 || !(p >= 0.f && p < 1.f) || !seed_dev || !colsum) {
        set_last_error("dropout_colsum: bad arguments (cols must be a multiple of 8, at most 4096)"); return kErrBadShape;
    }
    if (((uintptr_t)x % 16) || ((uintptr_t)y % 16)) { set_last_error("dropout: pointers must be 16-byte aligned"); return kErrBadAlign; }
    if (rows == 0) return kOk;
    const uint32_t thresh = dropout_thresh16(p);
    const float inv_keep = 65536.0f / (65536.0f - (float)thresh);
    const int W = (cols % 16 == 0) ? 16 : 8;
    const int nch = cols / W;
    const int threads = (512 / nch) * nch;
    int64_t g = (rows + (threads / nch) - 1) / (threads / nch);
    if (g > (int64_t)num_sms() * 2) g = (int64_t)num_sms() * 2;
    if (W == 16)
        dropout_colsum_kernel<16><<<(int)g, threads, cols * sizeof(float), s>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, rows, cols,
                                                                               thresh, inv_keep, seed_dev, site, colsum);
    else
        dropout_colsum_kernel<8><<<(int)g, threads, cols * sizeof(float), s>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, rows, cols,
                                                                              thresh, inv_keep, seed_dev, site, colsum);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_dropout_bf16(const void* x, const void* res, void* y, int64_t n, float p, const uint64_t* seed_dev,
                                uint32_t site, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (n < 0 || !(p >= 0.f && p < 1.f) || !seed_dev) { set_last_error("dropout: bad arguments"); return kErrBadShape; }
    if (((uintptr_t)x % 16) || ((uintptr_t)y % 16) || (res && ((uintptr_t)res % 16))) { set_last_error("dropout: pointers must be 16-byte aligned"); return kErrBadAlign; }
    if (n == 0) return kOk;
    const uint32_t thresh = dropout_thresh16(p);
    const float inv_keep = 65536.0f / (65536.0f - (float)thresh);
    int64_t g = ((n + 7) / 8 + 255) / 256;
    if (g > (int64_t)num_sms() * 16) g = (int64_t)num_sms() * 16;
    dropout_kernel<<<(int)g, 256, 0, s>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)res, (__nv_bfloat16*)y, n, thresh,
                                         inv_keep, seed_dev, site);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}
