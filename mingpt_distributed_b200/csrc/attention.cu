// Fused (flash-style) self-attention, forward and backward with recompute.
//
// Replaces the need_weights=True slow path of nn.MultiheadAttention that the reference runs
// (mingpt/model.py:160-165 -> torch/nn/functional.py:6630-6647): q*scale, baddbmm(mask, q, k^T),
// softmax, bmm(P, V) -- which materialises (B*H, T, T) fp32 tensors -- with an online-softmax kernel that
// keeps the T x T scores on chip.  Reads Q/K/V straight from the packed in-proj output [B,T,3,H,d] and
// writes the context token-major [B,T,H,d] so the out-projection GEMM consumes it without a transpose.
//
// This is the warp-level tensor-core path (mma.sync m16n8k16, ldmatrix, cp.async double buffering) used
// for every head size the reference's presets produce (d = 16 gpt-nano, 32 gpt-mini, 64 gpt2 family).
//
// mask modes (see mingpt_b200.h): CAUSAL, ADD_TRIL_ONES (the reference's as-written float mask), NONE.
#include "common.cuh"
#include "dropout.cuh"
#include "mingpt_b200.h"

#include <atomic>
#include <mutex>

namespace mgb {
extern std::atomic<int> g_opt_attn_mma;
namespace attn_tc {   // tcgen05 / TMEM kernels for head size 64 (attention_sm100.cu)
int launch_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int mask_mode, float scale, cudaStream_t stream,
               uint32_t drop_thresh, float inv_keep, const uint64_t* seed_dev, uint32_t site, const uint32_t* drop_bits = nullptr);
int launch_bwd_main(const void* qkv, const void* dout, const float* lse, const float* delta, float* dq_accum,
                    void* dqkv, float* dqkv_colsum, int B, int T, int H, int mask_mode, float scale, cudaStream_t stream,
                    uint32_t drop_thresh, float inv_keep, const uint64_t* seed_dev, uint32_t site,
                    const uint32_t* drop_bits = nullptr);
}
namespace attn {

constexpr int BM = 64;   // queries per CTA (fwd) / per inner step (bwd)
constexpr int BN = 64;   // keys per inner step (fwd) / per CTA (bwd)
constexpr int NTHREADS = 128;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool pred) {
    const uint32_t s = smem_u32(smem);
    const int sz = pred ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(smem_u32(p)));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(smem_u32(p)));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
        "{%0, %1, %2, %3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// 64 rows x D bf16 from global (row stride ld) into smem (row stride D+8); rows >= T are zero-filled.
template <int D>
__device__ __forceinline__ void load_tile(__nv_bfloat16* s, const __nv_bfloat16* gbase, int64_t ld, int t0, int T) {
    constexpr int CH = D / 8;
    constexpr int LDS = D + 8;
    for (int i = threadIdx.x; i < 64 * CH; i += NTHREADS) {
        const int r = i / CH, c = i - r * CH;
        const int t = t0 + r;
        const bool ok = t < T;
        cp_async16(s + r * LDS + c * 8, gbase + (int64_t)(ok ? t : 0) * ld + c * 8, ok);
    }
}

// =====================================================================================
// forward
// =====================================================================================
template <int D>
__global__ void __launch_bounds__(NTHREADS)
attn_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out, float* __restrict__ lse,
                int T, int H, int mask_mode, float scale_log2, uint32_t drop_thresh, float inv_keep,
                const uint64_t* __restrict__ seed_dev, uint32_t site) {
    constexpr int LDS = D + 8;
    constexpr int KD = D / 16;   // k-steps over the head dim
    constexpr int ND = D / 8;    // n-tiles over the head dim
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smem_raw);
    __nv_bfloat16* sK = sQ + 64 * LDS;        // [2][64*LDS]
    __nv_bfloat16* sV = sK + 2 * 64 * LDS;    // [2][64*LDS]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t4 = lane & 3;
    const int b = blockIdx.y / H, h = blockIdx.y % H;
    const int qb = gridDim.x - 1 - blockIdx.x;   // heavy (late) causal blocks first
    const int q0 = qb * BM;
    const int64_t ld = (int64_t)3 * H * D;
    const __nv_bfloat16* qbase = qkv + (int64_t)b * T * ld + (int64_t)h * D;
    const __nv_bfloat16* kbase = qbase + (int64_t)H * D;
    const __nv_bfloat16* vbase = qbase + (int64_t)2 * H * D;
    const bool causal = mask_mode == MGB_MASK_CAUSAL;
    const bool tril = mask_mode == MGB_MASK_ADD_TRIL_ONES;
    const int nkv_all = (T + BN - 1) / BN;
    const int nkv = causal ? min(qb + 1, nkv_all) : nkv_all;

    load_tile<D>(sQ, qbase, ld, q0, T);
    load_tile<D>(sK, kbase, ld, 0, T);
    load_tile<D>(sV, vbase, ld, 0, T);
    cp_async_commit();

    float o[ND][4];
#pragma unroll
    for (int n = 0; n < ND; ++n) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f; }
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    uint32_t qf[KD][4];
    const int r0 = q0 + warp * 16 + g, r1 = r0 + 8;

    for (int j = 0; j < nkv; ++j) {
        cp_async_wait_all();
        __syncthreads();
        const int buf = j & 1;
        if (j == 0) {
#pragma unroll
            for (int kk = 0; kk < KD; ++kk)
                ldsm_x4(qf[kk], sQ + (warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * LDS + kk * 16 + (lane >> 4) * 8);
        }
        if (j + 1 < nkv) {
            load_tile<D>(sK + (buf ^ 1) * 64 * LDS, kbase, ld, (j + 1) * BN, T);
            load_tile<D>(sV + (buf ^ 1) * 64 * LDS, vbase, ld, (j + 1) * BN, T);
            cp_async_commit();
        }
        const __nv_bfloat16* cK = sK + buf * 64 * LDS;
        const __nv_bfloat16* cV = sV + buf * 64 * LDS;

        float s[8][4];
#pragma unroll
        for (int n = 0; n < 8; ++n) { s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f; }
#pragma unroll
        for (int kk = 0; kk < KD; ++kk) {
#pragma unroll
            for (int nb = 0; nb < 4; ++nb) {
                uint32_t bf[4];
                ldsm_x4(bf, cK + (nb * 16 + (lane & 7) + (lane >> 4) * 8) * LDS + kk * 16 + ((lane >> 3) & 1) * 8);
                mma16816(s[2 * nb], qf[kk], bf[0], bf[1]);
                mma16816(s[2 * nb + 1], qf[kk], bf[2], bf[3]);
            }
        }
        // scale (log2 domain), additive tril bias, masking
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int n = 0; n < 8; ++n) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int col = j * BN + n * 8 + 2 * t4 + (e & 1);
                const int row = (e < 2) ? r0 : r1;
                float v = s[n][e] * scale_log2;
                if (tril && col <= row) v += LOG2E;
                if (col >= T || (causal && col > row)) v = -INFINITY;
                s[n][e] = v;
                if (e < 2) mx0 = fmaxf(mx0, v); else mx1 = fmaxf(mx1, v);
            }
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
        // fully-masked rows (only padding rows >= T): keep everything at zero without NaNs
        const float ms0 = (mn0 == -INFINITY) ? 0.f : mn0, ms1 = (mn1 == -INFINITY) ? 0.f : mn1;
        const float c0 = exp2f(m0 - ms0), c1 = exp2f(m1 - ms1);
        m0 = mn0; m1 = mn1;
        float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
        for (int n = 0; n < 8; ++n) {
            s[n][0] = exp2f(s[n][0] - ms0); s[n][1] = exp2f(s[n][1] - ms0);
            s[n][2] = exp2f(s[n][2] - ms1); s[n][3] = exp2f(s[n][3] - ms1);
            rs0 += s[n][0] + s[n][1];
            rs1 += s[n][2] + s[n][3];
        }
        l0 = l0 * c0 + rs0;
        l1 = l1 * c1 + rs1;
        if (drop_thresh) {
            // attention-probability dropout (nn.MultiheadAttention dropout=attn_drop, mingpt/model.py:150 ->
            // torch/nn/functional.py:6644-6645): the softmax normaliser keeps every key, P V only the kept ones
            const uint64_t seed = *seed_dev;
            const uint64_t t8 = (uint64_t)((T + 7) >> 3);
            const uint64_t g0 = ((uint64_t)blockIdx.y * T + (uint64_t)r0) * t8 + (uint64_t)(j * (BN / 8));
            const uint64_t g1 = g0 + 8 * t8;
#pragma unroll
            for (int n = 0; n < 8; ++n) {
                const uint32_t ka = dropout_keep8(seed, site, g0 + n, drop_thresh) >> (2 * t4);
                const uint32_t kb = dropout_keep8(seed, site, g1 + n, drop_thresh) >> (2 * t4);
                s[n][0] = (ka & 1u) ? s[n][0] * inv_keep : 0.f;
                s[n][1] = (ka & 2u) ? s[n][1] * inv_keep : 0.f;
                s[n][2] = (kb & 1u) ? s[n][2] * inv_keep : 0.f;
                s[n][3] = (kb & 2u) ? s[n][3] * inv_keep : 0.f;
            }
        }
#pragma unroll
        for (int n = 0; n < ND; ++n) { o[n][0] *= c0; o[n][1] *= c0; o[n][2] *= c1; o[n][3] *= c1; }
        // O += P V
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            uint32_t a[4];
            a[0] = pack_bf16x2(s[2 * kk][0], s[2 * kk][1]);
            a[1] = pack_bf16x2(s[2 * kk][2], s[2 * kk][3]);
            a[2] = pack_bf16x2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
            a[3] = pack_bf16x2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
            for (int nb = 0; nb < KD; ++nb) {
                uint32_t bf[4];
                ldsm_x4_t(bf, cV + (kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * LDS + nb * 16 + (lane >> 4) * 8);
                mma16816(o[2 * nb], a, bf[0], bf[1]);
                mma16816(o[2 * nb + 1], a, bf[2], bf[3]);
            }
        }
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = l0 > 0.f ? 1.f / l0 : 0.f, i1 = l1 > 0.f ? 1.f / l1 : 0.f;
    const int64_t od = (int64_t)H * D;
    if (r0 < T) {
        __nv_bfloat16* orow = out + ((int64_t)b * T + r0) * od + (int64_t)h * D;
#pragma unroll
        for (int n = 0; n < ND; ++n)
            *reinterpret_cast<uint32_t*>(orow + n * 8 + 2 * t4) = pack_bf16x2(o[n][0] * i0, o[n][1] * i0);
        if (t4 == 0) lse[((int64_t)b * H + h) * T + r0] = (m0 + log2f(l0)) * LN2;
    }
    if (r1 < T) {
        __nv_bfloat16* orow = out + ((int64_t)b * T + r1) * od + (int64_t)h * D;
#pragma unroll
        for (int n = 0; n < ND; ++n)
            *reinterpret_cast<uint32_t*>(orow + n * 8 + 2 * t4) = pack_bf16x2(o[n][2] * i1, o[n][3] * i1);
        if (t4 == 0) lse[((int64_t)b * H + h) * T + r1] = (m1 + log2f(l1)) * LN2;
    }
}

// =====================================================================================
// backward
// =====================================================================================
// delta[b,h,t] = sum_e dO[b,t,h,e] * O[b,t,h,e]
// D/8 consecutive threads share one (b,t,h) row: every load instruction of a warp covers contiguous 16-byte pieces.
template <int D>
__global__ void __launch_bounds__(256)
attn_bwd_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ dout,
                      float* __restrict__ delta, int64_t B, int64_t T, int H) {
    constexpr int TPR = D / 8;                      // threads per row: 2, 4 or 8
    const int64_t total = B * T * H * TPR;          // a multiple of TPR, so the shuffle groups below are always whole
    grid_dep_launch();
    grid_dep_wait();                                // dout is the previous kernel's output (programmatic dependent launch)
    // the loop condition is on the warp's first element, so whole warps stay in it for the shuffles
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31); i0 < total; i0 += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = i0 + (threadIdx.x & 31);
        float s = 0.f;
        if (i < total) {
            const uint4 a = ldg128(o + i * 8), d = ldg128(dout + i * 8);
            const uint32_t wa[4] = {a.x, a.y, a.z, a.w}, wd[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 fa = unpack_bf16x2(wa[e]), fd = unpack_bf16x2(wd[e]);
                s += fa.x * fd.x + fa.y * fd.y;
            }
        }
#pragma unroll
        for (int off = TPR / 2; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
        if (i < total && (i & (TPR - 1)) == 0) {
            const int64_t row = i / TPR;
            const int h = (int)(row % H);
            const int64_t bt = row / H;
            const int64_t b = bt / T, t = bt - b * T;
            delta[(b * H + h) * T + t] = s;
        }
    }
}

template <int D>
__global__ void __launch_bounds__(NTHREADS)
attn_bwd_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ dout,
                const float* __restrict__ lse, const float* __restrict__ delta, float* __restrict__ dq_accum,
                __nv_bfloat16* __restrict__ dqkv, int T, int H, int mask_mode, float scale, float scale_log2,
                uint32_t drop_thresh, float inv_keep, const uint64_t* __restrict__ seed_dev, uint32_t site) {
    constexpr int LDS = D + 8;
    constexpr int LDP = 64 + 8;
    constexpr int KD = D / 16;
    constexpr int ND = D / 8;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __nv_bfloat16* sK = reinterpret_cast<__nv_bfloat16*>(smem_raw);
    __nv_bfloat16* sV = sK + 64 * LDS;
    __nv_bfloat16* sQ = sV + 64 * LDS;
    __nv_bfloat16* sdO = sQ + 64 * LDS;
    __nv_bfloat16* sdS = sdO + 64 * LDS;                        // [64 queries][LDP] keys
    float* sLse = reinterpret_cast<float*>(sdS + 64 * LDP);     // [64], already * log2e
    float* sDelta = sLse + 64;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t4 = lane & 3;
    const int b = blockIdx.y / H, h = blockIdx.y % H;
    const int kb = blockIdx.x;
    const int k0 = kb * BN;
    const int64_t ld = (int64_t)3 * H * D;
    const int64_t od = (int64_t)H * D;
    const __nv_bfloat16* qbase = qkv + (int64_t)b * T * ld + (int64_t)h * D;
    const __nv_bfloat16* kbase = qbase + od;
    const __nv_bfloat16* vbase = qbase + 2 * od;
    const __nv_bfloat16* dobase = dout + (int64_t)b * T * od + (int64_t)h * D;
    const bool causal = mask_mode == MGB_MASK_CAUSAL;
    const bool tril = mask_mode == MGB_MASK_ADD_TRIL_ONES;
    const int nq = (T + BM - 1) / BM;
    const int i_begin = causal ? kb : 0;

    load_tile<D>(sK, kbase, ld, k0, T);
    load_tile<D>(sV, vbase, ld, k0, T);
    cp_async_commit();
    cp_async_wait_all();
    __syncthreads();

    uint32_t kf[KD][4], vf[KD][4];
#pragma unroll
    for (int kk = 0; kk < KD; ++kk) {
        const int off = (warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * LDS + kk * 16 + (lane >> 4) * 8;
        ldsm_x4(kf[kk], sK + off);
        ldsm_x4(vf[kk], sV + off);
    }
    float dk[ND][4], dv[ND][4];
#pragma unroll
    for (int n = 0; n < ND; ++n) {
        dk[n][0] = dk[n][1] = dk[n][2] = dk[n][3] = 0.f;
        dv[n][0] = dv[n][1] = dv[n][2] = dv[n][3] = 0.f;
    }
    const int kr0 = k0 + warp * 16 + g, kr1 = kr0 + 8;   // this thread's two key rows

    for (int i = i_begin; i < nq; ++i) {
        __syncthreads();   // previous step finished with sQ / sdO / sdS / sLse
        load_tile<D>(sQ, qbase, ld, i * BM, T);
        load_tile<D>(sdO, dobase, od, i * BM, T);
        cp_async_commit();
        if (threadIdx.x < 64) {
            const int t = i * BM + threadIdx.x;
            const int64_t idx = ((int64_t)b * H + h) * T + t;
            sLse[threadIdx.x] = (t < T) ? lse[idx] * LOG2E : INFINITY;
            sDelta[threadIdx.x] = (t < T) ? delta[idx] : 0.f;
        }
        cp_async_wait_all();
        __syncthreads();

        // S^T = K Q^T and dP^T = V dO^T : [16 keys x 64 queries] per warp
        float st[8][4], dp[8][4];
#pragma unroll
        for (int n = 0; n < 8; ++n) {
            st[n][0] = st[n][1] = st[n][2] = st[n][3] = 0.f;
            dp[n][0] = dp[n][1] = dp[n][2] = dp[n][3] = 0.f;
        }
#pragma unroll
        for (int kk = 0; kk < KD; ++kk) {
#pragma unroll
            for (int nb = 0; nb < 4; ++nb) {
                const int off = (nb * 16 + (lane & 7) + (lane >> 4) * 8) * LDS + kk * 16 + ((lane >> 3) & 1) * 8;
                uint32_t bq[4], bo[4];
                ldsm_x4(bq, sQ + off);
                ldsm_x4(bo, sdO + off);
                mma16816(st[2 * nb], kf[kk], bq[0], bq[1]);
                mma16816(st[2 * nb + 1], kf[kk], bq[2], bq[3]);
                mma16816(dp[2 * nb], vf[kk], bo[0], bo[1]);
                mma16816(dp[2 * nb + 1], vf[kk], bo[2], bo[3]);
            }
        }
        // P^T (into st) and scaled dS^T (into dp)
#pragma unroll
        for (int n = 0; n < 8; ++n) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int ql = n * 8 + 2 * t4 + (e & 1);
                const int q = i * BM + ql;
                const int key = (e < 2) ? kr0 : kr1;
                float sl = st[n][e] * scale_log2;
                if (tril && key <= q) sl += LOG2E;
                const bool masked = (key >= T) || (q >= T) || (causal && key > q);
                const float p = masked ? 0.f : exp2f(sl - sLse[ql]);
                float pd = p, dpe = dp[n][e];
                if (drop_thresh) {
                    // same mask as the forward: element (query q, key) of this (batch, head)
                    const uint64_t grp = ((uint64_t)blockIdx.y * T + (uint64_t)q) * (uint64_t)((T + 7) >> 3) + (uint64_t)(key >> 3);
                    const bool keep = (q < T) && ((dropout_keep8(*seed_dev, site, grp, drop_thresh) >> (key & 7)) & 1u);
                    pd = keep ? p * inv_keep : 0.f;       // what multiplied V in the forward
                    dpe = keep ? dpe * inv_keep : 0.f;    // d(loss)/d(softmax prob)
                }
                st[n][e] = pd;
                dp[n][e] = p * (dpe - sDelta[ql]) * scale;
            }
        }
        // dV += P^T dO ; dK += dS^T Q   (reduction over the 64 queries)
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            uint32_t ap[4], ad[4];
            ap[0] = pack_bf16x2(st[2 * kk][0], st[2 * kk][1]);
            ap[1] = pack_bf16x2(st[2 * kk][2], st[2 * kk][3]);
            ap[2] = pack_bf16x2(st[2 * kk + 1][0], st[2 * kk + 1][1]);
            ap[3] = pack_bf16x2(st[2 * kk + 1][2], st[2 * kk + 1][3]);
            ad[0] = pack_bf16x2(dp[2 * kk][0], dp[2 * kk][1]);
            ad[1] = pack_bf16x2(dp[2 * kk][2], dp[2 * kk][3]);
            ad[2] = pack_bf16x2(dp[2 * kk + 1][0], dp[2 * kk + 1][1]);
            ad[3] = pack_bf16x2(dp[2 * kk + 1][2], dp[2 * kk + 1][3]);
#pragma unroll
            for (int nb = 0; nb < KD; ++nb) {
                const int off = (kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * LDS + nb * 16 + (lane >> 4) * 8;
                uint32_t bo[4], bq[4];
                ldsm_x4_t(bo, sdO + off);
                ldsm_x4_t(bq, sQ + off);
                mma16816(dv[2 * nb], ap, bo[0], bo[1]);
                mma16816(dv[2 * nb + 1], ap, bo[2], bo[3]);
                mma16816(dk[2 * nb], ad, bq[0], bq[1]);
                mma16816(dk[2 * nb + 1], ad, bq[2], bq[3]);
            }
        }
        // dS (query-major) to shared memory for the dQ product
#pragma unroll
        for (int n = 0; n < 8; ++n) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int ql = n * 8 + 2 * t4 + (e & 1);
                const int kl = warp * 16 + g + ((e < 2) ? 0 : 8);
                sdS[ql * LDP + kl] = __float2bfloat16(dp[n][e]);
            }
        }
        __syncthreads();
        // dQ_i (partial over this key block) = dS K : warp takes 16 query rows
        float dq[ND][4];
#pragma unroll
        for (int n = 0; n < ND; ++n) { dq[n][0] = dq[n][1] = dq[n][2] = dq[n][3] = 0.f; }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            uint32_t a[4];
            ldsm_x4(a, sdS + (warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * LDP + kk * 16 + (lane >> 4) * 8);
#pragma unroll
            for (int nb = 0; nb < KD; ++nb) {
                uint32_t bk[4];
                ldsm_x4_t(bk, sK + (kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * LDS + nb * 16 + (lane >> 4) * 8);
                mma16816(dq[2 * nb], a, bk[0], bk[1]);
                mma16816(dq[2 * nb + 1], a, bk[2], bk[3]);
            }
        }
        const int qr0 = i * BM + warp * 16 + g, qr1 = qr0 + 8;
        if (qr0 < T) {
            float* dst = dq_accum + ((int64_t)b * T + qr0) * od + (int64_t)h * D;
#pragma unroll
            for (int n = 0; n < ND; ++n) {
                atomicAdd(dst + n * 8 + 2 * t4, dq[n][0]);
                atomicAdd(dst + n * 8 + 2 * t4 + 1, dq[n][1]);
            }
        }
        if (qr1 < T) {
            float* dst = dq_accum + ((int64_t)b * T + qr1) * od + (int64_t)h * D;
#pragma unroll
            for (int n = 0; n < ND; ++n) {
                atomicAdd(dst + n * 8 + 2 * t4, dq[n][2]);
                atomicAdd(dst + n * 8 + 2 * t4 + 1, dq[n][3]);
            }
        }
    }
    // dK, dV for this key block
    if (kr0 < T) {
        __nv_bfloat16* row = dqkv + ((int64_t)b * T + kr0) * ld + (int64_t)h * D;
#pragma unroll
        for (int n = 0; n < ND; ++n) {
            *reinterpret_cast<uint32_t*>(row + od + n * 8 + 2 * t4) = pack_bf16x2(dk[n][0], dk[n][1]);
            *reinterpret_cast<uint32_t*>(row + 2 * od + n * 8 + 2 * t4) = pack_bf16x2(dv[n][0], dv[n][1]);
        }
    }
    if (kr1 < T) {
        __nv_bfloat16* row = dqkv + ((int64_t)b * T + kr1) * ld + (int64_t)h * D;
#pragma unroll
        for (int n = 0; n < ND; ++n) {
            *reinterpret_cast<uint32_t*>(row + od + n * 8 + 2 * t4) = pack_bf16x2(dk[n][2], dk[n][3]);
            *reinterpret_cast<uint32_t*>(row + 2 * od + n * 8 + 2 * t4) = pack_bf16x2(dv[n][2], dv[n][3]);
        }
    }
}

// dqkv[b,t,0,h,:] = bf16(dq_accum[b,t,h,:])
__global__ void __launch_bounds__(256)
attn_dq_convert_kernel(const float* __restrict__ dq_accum, __nv_bfloat16* __restrict__ dqkv, int64_t BT, int64_t HD) {
    grid_dep_launch();
    grid_dep_wait();
    const int64_t nch = HD >> 3;
    const int64_t total = BT * nch;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / nch, ch = i - row * nch;
        const float4 a = *reinterpret_cast<const float4*>(dq_accum + row * HD + ch * 8);
        const float4 c = *reinterpret_cast<const float4*>(dq_accum + row * HD + ch * 8 + 4);
        stg128(dqkv + row * 3 * HD + ch * 8,
               make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(c.x, c.y), pack_bf16x2(c.z, c.w)));
    }
}
// Same, and colsum[0 .. HD) += column sums of dQ (the Q third of the qkv bias gradient). The block is
// (HD/8) x rows_per_block threads: a thread keeps its 8 columns for the whole grid-stride loop, so the sums
// stay in registers until one shared-memory and one global atomic per column and block.
__global__ void __launch_bounds__(512)
attn_dq_convert_colsum_kernel(const float* __restrict__ dq_accum, __nv_bfloat16* __restrict__ dqkv, int64_t BT, int HD,
                              float* __restrict__ colsum) {
    extern __shared__ float cvt_smem[];              // [HD]
    const int nch = HD >> 3;
    const int rpb = blockDim.x / nch;                // rows per block iteration
    const int ch = threadIdx.x % nch, r = threadIdx.x / nch;
    for (int c = threadIdx.x; c < HD; c += blockDim.x) cvt_smem[c] = 0.f;
    __syncthreads();
    grid_dep_launch();
    grid_dep_wait();                                 // dq_accum is complete once the main backward kernel has finished
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (r < rpb) {
        // four rows per iteration: eight 16-byte loads in flight per thread (the kernel is pure HBM streaming)
        const int64_t step = (int64_t)gridDim.x * rpb;
        for (int64_t row0 = (int64_t)blockIdx.x * rpb + r; row0 < BT; row0 += 4 * step) {
            float4 a[4], c[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t row = row0 + u * step;
                if (row < BT) {
                    a[u] = __ldcs(reinterpret_cast<const float4*>(dq_accum + row * HD + ch * 8));
                    c[u] = __ldcs(reinterpret_cast<const float4*>(dq_accum + row * HD + ch * 8 + 4));
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t row = row0 + u * step;
                if (row < BT) {
                    acc[0] += a[u].x; acc[1] += a[u].y; acc[2] += a[u].z; acc[3] += a[u].w;
                    acc[4] += c[u].x; acc[5] += c[u].y; acc[6] += c[u].z; acc[7] += c[u].w;
                    stg128(dqkv + row * 3 * HD + ch * 8,
                           make_uint4(pack_bf16x2(a[u].x, a[u].y), pack_bf16x2(a[u].z, a[u].w), pack_bf16x2(c[u].x, c[u].y),
                                      pack_bf16x2(c[u].z, c[u].w)));
                }
            }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) atomicAdd(&cvt_smem[ch * 8 + e], acc[e]);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < HD; c += blockDim.x) atomicAdd(colsum + c, cvt_smem[c]);
}

template <int D>
static int launch_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int mask_mode, float scale,
                      cudaStream_t s, float p_drop = 0.f, const uint64_t* seed_dev = nullptr, uint32_t site = 0) {
    constexpr int smem = (64 + 4 * 64) * (D + 8) * 2;
    dim3 grid((T + BM - 1) / BM, B * H);
    const uint32_t thresh = p_drop > 0.f ? dropout_thresh16(p_drop) : 0u;
    const float inv_keep = thresh ? 65536.0f / (65536.0f - (float)thresh) : 1.0f;
    attn_fwd_kernel<D><<<grid, NTHREADS, smem, s>>>((const __nv_bfloat16*)qkv, (__nv_bfloat16*)out, lse, T, H, mask_mode,
                                                   scale * LOG2E, thresh, inv_keep, seed_dev, site);
    return 0;
}
template <int D>
static int launch_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                      float* dq_accum, void* dqkv, int B, int T, int H, int mask_mode, float scale, cudaStream_t s,
                      float p_drop = 0.f, const uint64_t* seed_dev = nullptr, uint32_t site = 0, float* dqkv_colsum = nullptr,
                      const uint32_t* drop_bits = nullptr) {
    const uint32_t thresh = p_drop > 0.f ? dropout_thresh16(p_drop) : 0u;
    const float inv_keep = thresh ? 65536.0f / (65536.0f - (float)thresh) : 1.0f;
    const bool delta_given = (mask_mode & MGB_ATTN_DELTA_GIVEN) != 0;     // the caller's dgrad GEMM already produced it
    mask_mode &= ~MGB_ATTN_DELTA_GIVEN;
    if (!delta_given) {
        const int64_t rows = (int64_t)B * T * H * (D / 8);
        int g = (int)((rows + 255) / 256);
        if (g > num_sms() * 16) g = num_sms() * 16;
        launch_pdl(attn_bwd_delta_kernel<D>, dim3(g), dim3(256), 0, s, (const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, delta,
                   (int64_t)B, (int64_t)T, H);
    }
    constexpr int smem = 4 * 64 * (D + 8) * 2 + 64 * (64 + 8) * 2 + 2 * 64 * 4;
    if (smem > 48 * 1024) {
        static PerDeviceOnce once;
        once.run([] { cudaFuncSetAttribute(attn_bwd_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); });
    }
    // column sums of dqkv (the qkv bias gradient): fused on the tcgen05 path (K / V thirds in the main kernel's item
    // epilogue, Q third in the convert kernel), a separate pass otherwise
    bool fused_colsum = false;
    if (D == 64 && !g_opt_attn_mma.load(std::memory_order_relaxed)) {
        fused_colsum = dqkv_colsum != nullptr && (H * D / 8) <= 512;
        const int rc = attn_tc::launch_bwd_main(qkv, dout, lse, delta, dq_accum, dqkv, fused_colsum ? dqkv_colsum : nullptr, B, T,
                                                H, mask_mode, scale, s, thresh, inv_keep, seed_dev, site, drop_bits);
        if (rc) return rc;
    } else {
        dim3 grid((T + BN - 1) / BN, B * H);
        attn_bwd_kernel<D><<<grid, NTHREADS, smem, s>>>((const __nv_bfloat16*)qkv, (const __nv_bfloat16*)dout, lse, delta,
                                                       dq_accum, (__nv_bfloat16*)dqkv, T, H, mask_mode, scale, scale * LOG2E,
                                                       thresh, inv_keep, seed_dev, site);
    }
    if (fused_colsum) {
        const int nch = H * D / 8;
        // few, fat CTAs: every CTA ends with one global atomic per column, and with 8 CTAs per SM those tails were a
        // third of the kernel (16384 x 768: 19.3 us with 8 x 256-thread CTAs per SM, 13.2 us with 2 x 512)
        const int threads = (512 / nch) * nch;
        const int64_t rows = (int64_t)B * T;
        int64_t g2 = (rows + (threads / nch) - 1) / (threads / nch);
        if (g2 > num_sms() * 2) g2 = num_sms() * 2;
        launch_pdl(attn_dq_convert_colsum_kernel, dim3((int)g2), dim3(threads), H * D * sizeof(float), s, dq_accum,
                   (__nv_bfloat16*)dqkv, rows, H * D, dqkv_colsum);
        return 0;
    }
    const int64_t chunks = (int64_t)B * T * (H * D / 8);
    int g2 = (int)((chunks + 255) / 256);
    if (g2 > num_sms() * 16) g2 = num_sms() * 16;
    launch_pdl(attn_dq_convert_kernel, dim3(g2), dim3(256), 0, s, dq_accum, (__nv_bfloat16*)dqkv, (int64_t)B * T, (int64_t)H * D);
    return dqkv_colsum ? 1 : 0;      // 1: the caller still owes the column sums
}

}  // namespace attn
}  // namespace mgb

using namespace mgb;

static int check_attn_shape(const char* who, int B, int T, int H, int d, int mask_mode) {
    mask_mode &= ~MGB_ATTN_DELTA_GIVEN;
    if (B <= 0 || T <= 0 || H <= 0) { set_last_error("attention: non-positive B/T/H"); return kErrBadShape; }
    if (d != 16 && d != 32 && d != 64) { set_last_error("attention: head dim must be 16, 32 or 64"); return kErrUnsupported; }
    if (mask_mode < 0 || mask_mode > 2) { set_last_error("attention: bad mask mode"); return kErrBadShape; }
    if ((int64_t)B * H > 65535) { set_last_error("attention: B*H exceeds 65535"); return kErrTooLarge; }
    (void)who;
    return kOk;
}

extern "C" int mgb_attention_fwd(const void* qkv, void* out, float* lse, int32_t B, int32_t T, int32_t H, int32_t d,
                                 int32_t mask_mode, float scale, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    int rc = check_attn_shape("fwd", B, T, H, d, mask_mode);
    if (rc) return rc;
    if (((uintptr_t)qkv % 16) || ((uintptr_t)out % 16)) { set_last_error("attention: pointers must be 16-byte aligned"); return kErrBadAlign; }
    if (d == 64 && !g_opt_attn_mma.load(std::memory_order_relaxed)) {
        // gpt2-family head size: tcgen05 / TMEM kernel (attention_sm100.cu)
        rc = attn_tc::launch_fwd(qkv, out, lse, B, T, H, mask_mode, scale, s, 0u, 1.0f, nullptr, 0u);
        if (rc) return rc;
    } else if (d == 64) attn::launch_fwd<64>(qkv, out, lse, B, T, H, mask_mode, scale, s);
    else if (d == 32) attn::launch_fwd<32>(qkv, out, lse, B, T, H, mask_mode, scale, s);
    else attn::launch_fwd<16>(qkv, out, lse, B, T, H, mask_mode, scale, s);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_colsum_bf16(const void* x, int64_t ld, int64_t rows, int64_t cols, float* out, mgb_stream_t s);

static int attention_bwd_impl(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                              float* dq_accum, void* dqkv, float* dqkv_colsum, int32_t B, int32_t T, int32_t H, int32_t d,
                              int32_t mask_mode, float scale, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    int rc = check_attn_shape("bwd", B, T, H, d, mask_mode);
    if (rc) return rc;
    if (((uintptr_t)qkv % 16) || ((uintptr_t)out % 16) || ((uintptr_t)dout % 16) || ((uintptr_t)dqkv % 16) ||
        ((uintptr_t)dq_accum % 16)) {
        set_last_error("attention: pointers must be 16-byte aligned"); return kErrBadAlign;
    }
    cudaMemsetAsync(dq_accum, 0, (size_t)B * T * H * d * sizeof(float), s);
    int owed;
    if (d == 64) owed = attn::launch_bwd<64>(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, mask_mode, scale, s, 0.f, nullptr, 0, dqkv_colsum);
    else if (d == 32) owed = attn::launch_bwd<32>(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, mask_mode, scale, s, 0.f, nullptr, 0, dqkv_colsum);
    else owed = attn::launch_bwd<16>(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, mask_mode, scale, s, 0.f, nullptr, 0, dqkv_colsum);
    if (owed < 0) return owed;
    count_launch(3);
    MGB_CHECK_LAUNCH();
    if (owed == 1) return mgb_colsum_bf16(dqkv, (int64_t)3 * H * d, (int64_t)B * T, (int64_t)3 * H * d, dqkv_colsum, s_);
    return kOk;
}

extern "C" int mgb_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                                 float* dq_accum, void* dqkv, int32_t B, int32_t T, int32_t H, int32_t d,
                                 int32_t mask_mode, float scale, mgb_stream_t s_) {
    return attention_bwd_impl(qkv, out, dout, lse, delta, dq_accum, dqkv, nullptr, B, T, H, d, mask_mode, scale, s_);
}
extern "C" int mgb_attention_bwd_colsum(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                                        float* dq_accum, void* dqkv, float* dqkv_colsum, int32_t B, int32_t T, int32_t H,
                                        int32_t d, int32_t mask_mode, float scale, mgb_stream_t s_) {
    if (!dqkv_colsum) { set_last_error("attention_bwd_colsum: null colsum"); return kErrBadShape; }
    return attention_bwd_impl(qkv, out, dout, lse, delta, dq_accum, dqkv, dqkv_colsum, B, T, H, d, mask_mode, scale, s_);
}

// =====================================================================================
// attention backward for callers that cannot give us scratch memory
// =====================================================================================
// mgb_attention_bwd needs two scratch arrays from its caller (delta and the fp32 dQ accumulator). Some
// callers have nowhere to get them: a plugin living inside a serving process whose device pool was
// carved up at start-up, or a thin FFI stub (ctypes / cgo style) that merely forwards
// the pointers it was handed and owns no allocator. For them the scratch lives here, in one fixed array
// that every GPU loading this library keeps until the process exits.
namespace mgb {
namespace attn {

constexpr int64_t kNoAllocMaxElems = 4194304;                        // B*T*H*d, e.g. 4 x 1024 x 1024
constexpr int64_t kNoAllocDeltaElems = kNoAllocMaxElems / 16;        // B*H*T at the smallest head size (16)
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocMaxElems + kNoAllocDeltaElems];   // 17 MiB

constexpr int kMaxDevices = 64;
static float* g_noalloc_ptr[kMaxDevices] = {};
static std::mutex g_noalloc_mu;

// In host code `g_noalloc_scratch` is only a handle for the symbol, not a device address: resolve it once
// per device (with that device current) and hand the real pointer to the kernels.
static int noalloc_scratch(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        set_last_error("attention_bwd_noalloc: cannot identify the current device");
        return kErrDriver;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            set_last_error(e != cudaSuccess ? cudaGetErrorString(e) : "cudaGetSymbolAddress returned null");
            return kErrDriver;
        }
        g_noalloc_ptr[dev] = static_cast<float*>(p);
    }
    *out = g_noalloc_ptr[dev];
    return kOk;
}

}  // namespace attn
}  // namespace mgb

extern "C" int mgb_attention_bwd_noalloc(const void* qkv, const void* out, const void* dout, const float* lse,
                                         void* dqkv, int32_t B, int32_t T, int32_t H, int32_t d, int32_t mask_mode,
                                         float scale, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    int rc = check_attn_shape("bwd_noalloc", B, T, H, d, mask_mode);
    if (rc) return rc;
    if (mask_mode & MGB_ATTN_DELTA_GIVEN) { set_last_error("attention_bwd_noalloc: delta lives in the built-in scratch, it cannot be given"); return kErrUnsupported; }
    if (((uintptr_t)qkv % 16) || ((uintptr_t)out % 16) || ((uintptr_t)dout % 16) || ((uintptr_t)dqkv % 16)) {
        set_last_error("attention: pointers must be 16-byte aligned"); return kErrBadAlign;
    }
    const int64_t elems = (int64_t)B * T * H * d;
    if (elems > attn::kNoAllocMaxElems) {
        set_last_error("attention_bwd_noalloc: B*T*H*d exceeds the built-in scratch (4194304 elements); "
                       "use mgb_attention_bwd with caller-provided scratch");
        return kErrTooLarge;
    }
    float* scratch = nullptr;
    rc = attn::noalloc_scratch(&scratch);
    if (rc) return rc;
    float* dq_accum = scratch;                                  // [B,T,H,d]; must start at zero
    float* delta = scratch + attn::kNoAllocMaxElems;            // [B,H,T];   fully overwritten
    // the array still holds the previous call's partial sums: clear the part this call accumulates into
    cudaMemsetAsync(dq_accum, 0, (size_t)elems * sizeof(float), s);
    int owed;
    if (d == 64) owed = attn::launch_bwd<64>(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, mask_mode, scale, s);
    else if (d == 32) owed = attn::launch_bwd<32>(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, mask_mode, scale, s);
    else owed = attn::launch_bwd<16>(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, mask_mode, scale, s);
    if (owed < 0) return owed;       // e.g. a tensor-map encode failed: the main kernel was never launched
    count_launch(3);
    MGB_CHECK_LAUNCH();
    return kOk;
}

// =====================================================================================
// incremental decode: ONE new query per sequence against the cached keys / values
// (GPT.generate, mingpt/model.py:322-356, which recomputes the whole window for every token)
// =====================================================================================
namespace mgb {
namespace attn {
constexpr int DEC_THREADS = 128;
// CTA = one (batch, head). The cache has the packed in-proj layout [B, Tcap, 3, H, d] (what the qkv GEMM writes), the
// query is row `pos`, keys / values are rows 0 .. pos -- causal attention of the last position, so no mask is needed.
// fp32 scores in shared memory; thread = key for the scores, thread = (8 features, key group) for P V.
template <int D>
__global__ void __launch_bounds__(DEC_THREADS)
attn_decode_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out, int Tcap, int pos, int H,
                   float scale_log2) {
    extern __shared__ float dsm[];
    float* sc = dsm;                               // [pos + 1] scores, then probabilities
    float* qs = sc + ((pos + 1 + 3) & ~3);         // [D] query, pre-scaled
    float* red = qs + D;                           // [8] block reductions
    float* part = red + 8;                         // [NG][D] partial outputs
    constexpr int CH = D / 8, NG = DEC_THREADS / CH;
    const int b = blockIdx.x / H, h = blockIdx.x - b * H;
    const int C = H * D;
    const int64_t row_stride = (int64_t)3 * C;
    const __nv_bfloat16* base = qkv + (int64_t)b * Tcap * row_stride + (int64_t)h * D;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = pos + 1;
    if (tid < D) qs[tid] = __bfloat162float(base[(int64_t)pos * row_stride + tid]) * scale_log2;
    __syncthreads();
    float mx = -INFINITY;
    for (int key = tid; key < n; key += DEC_THREADS) {
        const __nv_bfloat16* kp = base + (int64_t)key * row_stride + C;
        float dot = 0.f;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            const uint4 kv = ldg128(kp + c * 8);
            const uint32_t w[4] = {kv.x, kv.y, kv.z, kv.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = unpack_bf16x2(w[e]);
                dot = fmaf(f.x, qs[c * 8 + 2 * e], dot);
                dot = fmaf(f.y, qs[c * 8 + 2 * e + 1], dot);
            }
        }
        sc[key] = dot;
        mx = fmaxf(mx, dot);
    }
    mx = warp_max(mx);
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    mx = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
    float sum = 0.f;
    for (int key = tid; key < n; key += DEC_THREADS) {
        const float p = exp2f(sc[key] - mx);
        sc[key] = p;
        sum += p;
    }
    sum = warp_sum(sum);
    if (lane == 0) red[4 + warp] = sum;
    __syncthreads();
    const float inv = 1.0f / (red[4] + red[5] + red[6] + red[7]);
    const int ch = tid % CH, grp = tid / CH;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int key = grp; key < n; key += NG) {
        const uint4 vv = ldg128(base + (int64_t)key * row_stride + 2 * C + ch * 8);
        const uint32_t w[4] = {vv.x, vv.y, vv.z, vv.w};
        const float p = sc[key];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 f = unpack_bf16x2(w[e]);
            acc[2 * e] = fmaf(p, f.x, acc[2 * e]);
            acc[2 * e + 1] = fmaf(p, f.y, acc[2 * e + 1]);
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) part[grp * D + ch * 8 + e] = acc[e];
    __syncthreads();
    if (tid < D) {
        float o = 0.f;
        for (int g = 0; g < NG; ++g) o += part[g * D + tid];
        out[(int64_t)b * C + (int64_t)h * D + tid] = __float2bfloat16(o * inv);
    }
}
}  // namespace attn
}  // namespace mgb

extern "C" int mgb_attention_decode(const void* qkv_cache, void* out, int32_t B, int32_t Tcap, int32_t pos, int32_t H,
                                    int32_t d, float scale, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    int rc = check_attn_shape("decode", B, Tcap, H, d, MGB_MASK_CAUSAL);
    if (rc) return rc;
    if (pos < 0 || pos >= Tcap) { set_last_error("attention_decode: pos outside the cache"); return kErrBadShape; }
    if ((uintptr_t)qkv_cache % 16) { set_last_error("attention: pointers must be 16-byte aligned"); return kErrBadAlign; }
    const int ng = attn::DEC_THREADS / (d / 8);
    const size_t smem = (size_t)(((pos + 1 + 3) & ~3) + d + 8 + ng * d) * sizeof(float);
    if (smem > 48 * 1024) { set_last_error("attention_decode: cache too long for the score buffer"); return kErrTooLarge; }
    const float sl = scale * attn::LOG2E;
    dim3 grid(B * H);
    if (d == 64) attn::attn_decode_kernel<64><<<grid, attn::DEC_THREADS, smem, s>>>((const __nv_bfloat16*)qkv_cache, (__nv_bfloat16*)out, Tcap, pos, H, sl);
    else if (d == 32) attn::attn_decode_kernel<32><<<grid, attn::DEC_THREADS, smem, s>>>((const __nv_bfloat16*)qkv_cache, (__nv_bfloat16*)out, Tcap, pos, H, sl);
    else attn::attn_decode_kernel<16><<<grid, attn::DEC_THREADS, smem, s>>>((const __nv_bfloat16*)qkv_cache, (__nv_bfloat16*)out, Tcap, pos, H, sl);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

// =====================================================================================
// attention with dropout on the probabilities (training with attn_drop > 0, the reference's default 0.1)
// =====================================================================================
namespace mgb {
namespace attn {
// keep[b,h,q,k] (uint8) exactly as the kernels above generate it: lets tests rebuild the computation in torch
__global__ void __launch_bounds__(256)
attn_dropout_mask_kernel(uint8_t* __restrict__ keep, int64_t BH, int T, uint32_t thresh, const uint64_t* seed_dev,
                         uint32_t site) {
    const int64_t t8 = (T + 7) >> 3;
    const int64_t groups = BH * T * t8;
    const uint64_t seed = *seed_dev;
    for (int64_t gidx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; gidx < groups; gidx += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t m = dropout_keep8(seed, site, (uint64_t)gidx, thresh);
        const int64_t row = gidx / t8;                 // (bh * T + q)
        const int k0 = (int)(gidx - row * t8) * 8;
        for (int i = 0; i < 8 && k0 + i < T; ++i) keep[row * T + k0 + i] = (m >> i) & 1u;
    }
}
}  // namespace attn
}  // namespace mgb

namespace mgb {
namespace attn {
// Keep words for the tcgen05 kernels (layouts and bit order: dropout.cuh). Thread = 32 keys of one query, a warp =
// 32 consecutive queries: it stores its query-major words (coalesced) for the forward, transposes the 32 x 32 bits
// across its lanes and stores the key-major words for the backward. Grid = (query chunks, key words, batch * heads).
// Causal: words right of the diagonal are never read, and their warps leave at once.
__global__ void __launch_bounds__(256)
attn_dropout_bits_kernel(uint32_t* __restrict__ bits, int BH, int T, int causal, uint32_t thresh,
                         const uint64_t* __restrict__ seed_dev, uint32_t site) {
    const int kw = (T + 31) >> 5, tp = (T + 15) & ~15, kp = kw * 32;
    const int w = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int q = blockIdx.x * 256 + threadIdx.x;
    const int q0 = q - lane;                                     // first query of this warp: a multiple of 32
    if (q0 >= T || (causal && w * 32 > q0)) return;              // warp-uniform
    const uint64_t t8 = (uint64_t)((T + 7) >> 3);
    const uint64_t seed = *seed_dev;
    uint32_t* bits_t = bits + dropout_bits_words_q(BH, T);
    const int key_t = w * 32 + dropout_key_of_bit(lane);         // after the transpose lane b holds the word of the key at bit b
    for (int bh = blockIdx.z; bh < BH; bh += gridDim.z) {
        uint32_t word = 0;
        if (q < T) {
            word = dropout_keep32_word(seed, site, ((uint64_t)bh * T + (uint64_t)q) * t8 + (uint64_t)w * 4, thresh);
            bits[((int64_t)bh * kw + w) * tp + q] = word;
        }
        const uint32_t wt = warp_transpose32(word, lane);
        bits_t[((int64_t)bh * kw + (q0 >> 5)) * kp + key_t] = wt;
    }
}
}  // namespace attn
}  // namespace mgb

static int64_t dropout_bits_words(int32_t B, int32_t T, int32_t H) {
    return (int64_t)(dropout_bits_words_q((int64_t)B * H, T) + dropout_bits_words_t((int64_t)B * H, T));
}
extern "C" int mgb_attention_dropout_bits_words(int32_t B, int32_t T, int32_t H, int64_t* words) {
    if (B <= 0 || T <= 0 || H <= 0 || !words) { set_last_error("dropout_bits_words: bad arguments"); return kErrBadShape; }
    *words = dropout_bits_words(B, T, H);
    return kOk;
}

extern "C" int mgb_attention_dropout_bits(uint32_t* bits, int32_t B, int32_t T, int32_t H, int32_t mask_mode, float p_drop,
                                          const uint64_t* seed_dev, uint32_t site, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (B <= 0 || T <= 0 || H <= 0 || !bits || !seed_dev || !(p_drop > 0.f && p_drop < 1.f)) { set_last_error("dropout_bits: bad arguments"); return kErrBadShape; }
    if ((uintptr_t)bits % 16) { set_last_error("dropout_bits: the buffer must be 16-byte aligned"); return kErrBadAlign; }
    const int64_t bh = (int64_t)B * H;
    // four (batch, head) pairs per thread: amortises the CTA prologue (the seed load) without starving the tail
    const int64_t gz = (bh + 3) / 4;
    dim3 grid((T + 255) / 256, (T + 31) / 32, (unsigned)(gz < 65535 ? gz : 65535));
    attn::attn_dropout_bits_kernel<<<grid, 256, 0, s>>>(bits, (int)bh, T, (mask_mode & 0xff) == MGB_MASK_CAUSAL ? 1 : 0,
                                                       dropout_thresh16(p_drop), seed_dev, site);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_attention_fwd_dropout_bits(const void* qkv, void* out, float* lse, int32_t B, int32_t T, int32_t H,
                                              int32_t d, int32_t mask_mode, float scale, float p_drop, const uint32_t* bits,
                                              mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    int rc = check_attn_shape("fwd_dropout_bits", B, T, H, d, mask_mode);
    if (rc) return rc;
    if (d != 64 || g_opt_attn_mma.load(std::memory_order_relaxed)) { set_last_error("attention_fwd_dropout_bits: head size 64 (the tcgen05 kernel) only"); return kErrUnsupported; }
    if (!(p_drop > 0.f && p_drop < 1.f) || !bits) { set_last_error("attention: bad dropout arguments"); return kErrBadShape; }
    if (((uintptr_t)qkv % 16) || ((uintptr_t)out % 16) || ((uintptr_t)bits % 16)) { set_last_error("attention: pointers must be 16-byte aligned"); return kErrBadAlign; }
    const uint32_t thresh = dropout_thresh16(p_drop);
    if (!thresh) return mgb_attention_fwd(qkv, out, lse, B, T, H, d, mask_mode, scale, s_);
    rc = attn_tc::launch_fwd(qkv, out, lse, B, T, H, mask_mode, scale, s, thresh, 65536.0f / (65536.0f - (float)thresh), nullptr, 0u, bits);
    if (rc) return rc;
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_attention_bwd_dropout_bits(const void* qkv, const void* out, const void* dout, const float* lse,
                                              float* delta, float* dq_accum, void* dqkv, int32_t B, int32_t T, int32_t H,
                                              int32_t d, int32_t mask_mode, float scale, float p_drop, const uint32_t* bits,
                                              float* dqkv_colsum, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    int rc = check_attn_shape("bwd_dropout_bits", B, T, H, d, mask_mode);
    if (rc) return rc;
    if (d != 64 || g_opt_attn_mma.load(std::memory_order_relaxed)) { set_last_error("attention_bwd_dropout_bits: head size 64 (the tcgen05 kernel) only"); return kErrUnsupported; }
    if (!(p_drop > 0.f && p_drop < 1.f) || !bits) { set_last_error("attention: bad dropout arguments"); return kErrBadShape; }
    if (((uintptr_t)qkv % 16) || ((uintptr_t)out % 16) || ((uintptr_t)dout % 16) || ((uintptr_t)dqkv % 16) ||
        ((uintptr_t)dq_accum % 16) || ((uintptr_t)bits % 16)) {
        set_last_error("attention: pointers must be 16-byte aligned"); return kErrBadAlign;
    }
    cudaMemsetAsync(dq_accum, 0, (size_t)B * T * H * d * sizeof(float), s);
    const int owed = attn::launch_bwd<64>(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, mask_mode, scale, s, p_drop, nullptr, 0u,
                                          dqkv_colsum, bits);
    if (owed < 0) return owed;
    count_launch(3);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_attention_fwd_dropout(const void* qkv, void* out, float* lse, int32_t B, int32_t T, int32_t H,
                                         int32_t d, int32_t mask_mode, float scale, float p_drop,
                                         const uint64_t* seed_dev, uint32_t site, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    int rc = check_attn_shape("fwd_dropout", B, T, H, d, mask_mode);
    if (rc) return rc;
    if (!(p_drop >= 0.f && p_drop < 1.f) || (p_drop > 0.f && !seed_dev)) { set_last_error("attention: bad dropout arguments"); return kErrBadShape; }
    if (((uintptr_t)qkv % 16) || ((uintptr_t)out % 16)) { set_last_error("attention: pointers must be 16-byte aligned"); return kErrBadAlign; }
    if (p_drop == 0.f) return mgb_attention_fwd(qkv, out, lse, B, T, H, d, mask_mode, scale, s_);
    if (d == 64 && !g_opt_attn_mma.load(std::memory_order_relaxed)) {
        // tcgen05 kernel, same mask definition as the mma.sync ones
        const uint32_t thresh = dropout_thresh16(p_drop);
        rc = attn_tc::launch_fwd(qkv, out, lse, B, T, H, mask_mode, scale, s, thresh, 65536.0f / (65536.0f - (float)thresh), seed_dev, site);
        if (rc) return rc;
    } else if (d == 64) attn::launch_fwd<64>(qkv, out, lse, B, T, H, mask_mode, scale, s, p_drop, seed_dev, site);
    else if (d == 32) attn::launch_fwd<32>(qkv, out, lse, B, T, H, mask_mode, scale, s, p_drop, seed_dev, site);
    else attn::launch_fwd<16>(qkv, out, lse, B, T, H, mask_mode, scale, s, p_drop, seed_dev, site);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_attention_bwd_dropout(const void* qkv, const void* out, const void* dout, const float* lse,
                                         float* delta, float* dq_accum, void* dqkv, int32_t B, int32_t T, int32_t H,
                                         int32_t d, int32_t mask_mode, float scale, float p_drop,
                                         const uint64_t* seed_dev, uint32_t site, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    int rc = check_attn_shape("bwd_dropout", B, T, H, d, mask_mode);
    if (rc) return rc;
    if (!(p_drop >= 0.f && p_drop < 1.f) || (p_drop > 0.f && !seed_dev)) { set_last_error("attention: bad dropout arguments"); return kErrBadShape; }
    if (p_drop == 0.f) return mgb_attention_bwd(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, d, mask_mode, scale, s_);
    if (((uintptr_t)qkv % 16) || ((uintptr_t)out % 16) || ((uintptr_t)dout % 16) || ((uintptr_t)dqkv % 16) ||
        ((uintptr_t)dq_accum % 16)) {
        set_last_error("attention: pointers must be 16-byte aligned"); return kErrBadAlign;
    }
    cudaMemsetAsync(dq_accum, 0, (size_t)B * T * H * d * sizeof(float), s);
    int owed;
    if (d == 64) owed = attn::launch_bwd<64>(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, mask_mode, scale, s, p_drop, seed_dev, site);
    else if (d == 32) owed = attn::launch_bwd<32>(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, mask_mode, scale, s, p_drop, seed_dev, site);
    else owed = attn::launch_bwd<16>(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, mask_mode, scale, s, p_drop, seed_dev, site);
    if (owed < 0) return owed;
    count_launch(3);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_attention_dropout_mask(void* keep, int32_t B, int32_t T, int32_t H, float p_drop,
                                          const uint64_t* seed_dev, uint32_t site, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (B <= 0 || T <= 0 || H <= 0 || !seed_dev || !(p_drop >= 0.f && p_drop < 1.f)) { set_last_error("dropout_mask: bad arguments"); return kErrBadShape; }
    const int64_t groups = (int64_t)B * H * T * ((T + 7) / 8);
    int g = (int)((groups + 255) / 256);
    if (g > num_sms() * 32) g = num_sms() * 32;
    attn::attn_dropout_mask_kernel<<<g, 256, 0, s>>>((uint8_t*)keep, (int64_t)B * H, T, dropout_thresh16(p_drop), seed_dev, site);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}
