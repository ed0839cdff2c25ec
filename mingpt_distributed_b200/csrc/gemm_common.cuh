// Shared between the single-CTA (gemm_sm100.cu) and CTA-pair (gemm2_sm100.cu) tcgen05 GEMM kernels:
// kernel parameters, epilogue math, TMA descriptor construction.
#pragma once

#include "common.cuh"
#include "mingpt_b200.h"

#include <atomic>
#include <mutex>

namespace mgb {
extern std::atomic<int> g_opt_gemm_single;
extern std::atomic<int> g_opt_gemm_direct_store;
extern std::atomic<int> g_opt_gemm_static;        // 1 = static persistent tile walk instead of the hardware work queue
extern std::atomic<int> g_opt_gemm_generic;       // 1 = always the generic (run-time) epilogue instantiation
namespace gemm {

constexpr int BLOCK_M = 128;   // rows per CTA (a CTA pair covers 256)
constexpr int BLOCK_K = 64;    // 64 bf16 = 128 bytes = one SWIZZLE_128B row
constexpr int UMMA_K = 16;
constexpr int MAX_BLOCK_N = 256;
constexpr int kEpiWarps = 8;
constexpr int kThreads = 128 + kEpiWarps * 32;
constexpr int kTmemCols = 512;
constexpr int kAccStride = 256;              // TMEM column offset between the two accumulator stages
constexpr int ATOM_BYTES = 64 * BLOCK_K * 2; // one 64-wide MN-major atom: 8 KiB

struct Params {
    int64_t M, N, K;
    int32_t a_mn, b_mn;
    int32_t block_n;
    int32_t split_k;
    int32_t num_m_tiles, num_n_tiles, k_blocks_total, k_blocks_per_split;
    int32_t num_stages;      // CTA-pair kernel only
    int32_t tma_store;       // CTA-pair kernel only: D (and preact) leave through smem + TMA
    int32_t m_fast;          // CTA-pair kernel only: tile raster order (1 = m index fastest)
    int32_t tma_aux;         // CTA-pair kernel only: 1 residual / 2 gate arrives through TMA, updated in place
    int32_t dynamic;         // CTA-pair kernel only: 1 = one cluster per unit + cluster launch control, 0 = static persistent walk
    // epilogue
    void* d;
    int64_t ldd;
    int32_t d_f32;
    int32_t accumulate;
    const float* bias;
    const __nv_bfloat16* residual;
    int64_t ldr;
    int32_t act;
    int32_t act_grad;        // preact receives act'(v) instead of v (MGB_ACT_STORE_GRAD)
    __nv_bfloat16* preact;
    int64_t ldp;
    const __nv_bfloat16* gate;
    int64_t ldg;
    int32_t gate_kind;
    float alpha;
    float* colsum;           // CTA-pair kernel, COLSUM instantiation: column sums of D (fp32, atomic)
    float* rowdot_out;       // CTA-pair kernel, tma_aux == 3: per (row, 64-column slice) sum of D * X
    int64_t rowdot_T;
    const uint64_t* drop_seed;   // CTA-pair kernel, MODE 4: dropout between bias and residual (dropout.cuh masks)
    uint32_t drop_thresh, drop_site;
    float drop_inv_keep;
};

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == MGB_ACT_GELU_TANH) return gelu_tanh_f(v);
    if (act == MGB_ACT_GELU_ERF) return gelu_erf_f(v);
    return v;
}
__device__ __forceinline__ float apply_gate(float g, int kind) {
    if (kind == MGB_GATE_MUL) return g;
    return kind == MGB_ACT_GELU_TANH ? gelu_tanh_grad_f(g) : gelu_erf_grad_f(g);
}
__device__ __forceinline__ float apply_act_grad(float v, int act) {
    return act == MGB_ACT_GELU_TANH ? gelu_tanh_grad_f(v) : gelu_erf_grad_f(v);
}

// The activation kind is uniform over the launch: branch ONCE around a fully unrolled loop. (Selecting per
// element makes the compiler evaluate both GELU flavours under predication: measured 2x on the epilogue.)
#ifndef MGB_GELU_FFMA2
#define MGB_GELU_FFMA2 1
#endif
__device__ __forceinline__ void act_inplace(float (&v)[32], int act) {
    if (act == MGB_ACT_GELU_TANH) {
#if MGB_GELU_FFMA2
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float2 y = gelu_tanh_2(make_float2(v[2 * j], v[2 * j + 1]));
            v[2 * j] = y.x; v[2 * j + 1] = y.y;
        }
#else
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = gelu_tanh_f(v[j]);
#endif
    } else if (act == MGB_ACT_GELU_ERF) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = gelu_erf_f(v[j]);
    }
}
// The 16 packed bf16x2 words a `preact` output receives for these 32 columns: the pre-activation itself, or
// (act_grad) the activation's derivative -- in which case the activation is applied to v on the way (the two
// share the tanh / erf) and true is returned.
__device__ __forceinline__ bool preact_words(float (&v)[32], uint32_t (&w)[16], int act, int act_grad) {
    if (!act_grad) {
#pragma unroll
        for (int j = 0; j < 16; ++j) w[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
        return false;
    }
    if (act == MGB_ACT_GELU_TANH) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
#if MGB_GELU_FFMA2
            float2 y, g;
            gelu_tanh_both_2(make_float2(v[2 * j], v[2 * j + 1]), y, g);
            v[2 * j] = y.x; v[2 * j + 1] = y.y;
            w[j] = pack_bf16x2(g.x, g.y);
#else
            float g0, g1;
            gelu_tanh_both_f(v[2 * j], v[2 * j], g0);
            gelu_tanh_both_f(v[2 * j + 1], v[2 * j + 1], g1);
            w[j] = pack_bf16x2(g0, g1);
#endif
        }
    } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            float g0, g1;
            gelu_erf_both_f(v[2 * j], v[2 * j], g0);
            gelu_erf_both_f(v[2 * j + 1], v[2 * j + 1], g1);
            w[j] = pack_bf16x2(g0, g1);
        }
    }
    return true;
}
// v[j] *= act'(g[j]) (or v[j] *= g[j] for MGB_GATE_MUL) with g given as 16 packed bf16x2 words
__device__ __forceinline__ void gate_inplace(float (&v)[32], const uint32_t (&g)[16], int kind) {
    if (kind == MGB_GATE_MUL) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float2 r = mul2(make_float2(v[2 * j], v[2 * j + 1]), unpack_bf16x2(g[j]));
            v[2 * j] = r.x; v[2 * j + 1] = r.y;
        }
    } else if (kind == MGB_ACT_GELU_TANH) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float2 f = unpack_bf16x2(g[j]);
            v[2 * j] *= gelu_tanh_grad_f(f.x);
            v[2 * j + 1] *= gelu_tanh_grad_f(f.y);
        }
    } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float2 f = unpack_bf16x2(g[j]);
            v[2 * j] *= gelu_erf_grad_f(f.x);
            v[2 * j + 1] *= gelu_erf_grad_f(f.y);
        }
    }
}
__device__ __forceinline__ void add_bf16_inplace(float (&v)[32], const uint32_t (&r)[16]) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const float2 s = add2(make_float2(v[2 * j], v[2 * j + 1]), unpack_bf16x2(r[j]));
        v[2 * j] = s.x; v[2 * j + 1] = s.y;
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    });
    return fn;
}

// 2-D tensor map (SWIZZLE_128B): `inner` contiguous elements, `outer` rows, `ld` elements between rows.
inline int make_map_2d(CUtensorMap* out, const void* base, int64_t inner, int64_t outer, int64_t ld, int box_inner,
                       int box_outer, bool f32 = false) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) { set_last_error("cuTensorMapEncodeTiled not available"); return kErrDriver; }
    cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
    cuuint64_t strides[1] = {(cuuint64_t)ld * (f32 ? 4 : 2)};
    cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                     const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_last_error("cuTensorMapEncodeTiled failed (check 16-byte alignment of base and leading dimension)");
        return kErrDriver;
    }
    return kOk;
}

// CTA-pair kernel launcher (gemm2_sm100.cu). Fills the remaining Params fields itself.
constexpr int kFusedColsum = 1;   // launch_gemm_pair: launched, and the column sums were part of the kernel
int launch_gemm_pair(const mgb_gemm_args* a, Params p, int num_sms, cudaStream_t stream);

}  // namespace gemm
}  // namespace mgb
