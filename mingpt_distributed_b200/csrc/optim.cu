// Gradient-norm, clipping, multi-tensor AdamW and the small elementwise helpers
// (casts, axpy, bias-gradient column sums, standalone GELU fwd/bwd). All HBM-bound, 16-byte accesses.
#include "common.cuh"
#include "mingpt_b200.h"

#include <atomic>
#include <cmath>
#include <cstring>
#include <mutex>

namespace mgb {

// ---------------------------------------------------------------------------
// library-wide state: last error (thread local), launch counter, SM count
// ---------------------------------------------------------------------------
static thread_local char g_err[512] = "";
void set_last_error(const char* msg) {
    std::strncpy(g_err, msg ? msg : "", sizeof(g_err) - 1);
    g_err[sizeof(g_err) - 1] = 0;
}
static std::atomic<int64_t> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
std::atomic<int> g_opt_reserved_sms{0};
int num_sms_total() {
    static std::atomic<int> cache[128];          // per device ordinal; 0 = not queried yet
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev > 127) dev = 0;
    int n = cache[dev].load(std::memory_order_relaxed);
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cache[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}
int num_sms() {
    const int total = num_sms_total();
    int r = g_opt_reserved_sms.load(std::memory_order_relaxed);
    if (r < 0) r = 0;
    r &= ~1;                                      // CTA pairs: keep the usable count even
    return total - r >= 16 ? total - r : total;
}

// ---------------------------------------------------------------------------
// sum of squares (clip_grad_norm's global L2 norm: torch/nn/utils/clip_grad.py)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(512)
sumsq_kernel(const float* __restrict__ x, int64_t n, float* out) {
    __shared__ float sh[16];
    float s = 0.f;
    const int64_t n4 = n >> 2;
    const float4* x4 = reinterpret_cast<const float4*>(x);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = x4[i];
        s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    if (blockIdx.x == 0)
        for (int64_t i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) s += x[i] * x[i];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        float r = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
        r = warp_sum(r);
        if (threadIdx.x == 0) atomicAdd(out, r);
    }
}

__device__ __forceinline__ float clip_coef(const float* gradnorm_sq, float grad_scale, float max_norm) {
    if (!gradnorm_sq || !(max_norm > 0.f)) return 1.0f;
    const float norm = sqrtf(*gradnorm_sq) * fabsf(grad_scale);
    return fminf(1.0f, max_norm / (norm + 1e-6f));
}

__global__ void __launch_bounds__(256)
clip_scale_kernel(float* g, int64_t n, const float* gradnorm_sq, float max_norm) {
    const float c = clip_coef(gradnorm_sq, 1.0f, max_norm);
    if (c >= 1.0f) return;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        g[i] *= c;
}

// ---------------------------------------------------------------------------
// multi-tensor AdamW (torch/optim/adam.py single-tensor math, decoupled decay)
//   p *= 1 - lr*wd;  m = lerp(m, g, 1-b1);  v = b2*v + (1-b2)*g*g;
//   p -= (lr/bc1) * m / (sqrt(v)/sqrt(bc2) + eps)
// One CTA per table entry (a chunk of one tensor), so the decay is uniform inside a CTA.
// Also writes the bf16 compute copy of the updated parameter when params_bf16 != NULL.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
             __nv_bfloat16* __restrict__ pb, const mgb_adamw_chunk* __restrict__ table, float lr, float beta1,
             float beta2, float eps, float step_size, float inv_bc2_sqrt, float grad_scale,
             const float* gradnorm_sq, float max_norm, const int32_t* __restrict__ step_dev) {
    if (step_dev) {
        // step count lives on the device (CUDA-graph replay: host arguments are frozen at capture time)
        const float t = (float)(*step_dev);
        const float bc1 = 1.0f - powf(beta1, t);
        const float bc2 = 1.0f - powf(beta2, t);
        step_size = lr / bc1;
        inv_bc2_sqrt = rsqrtf(bc2);
    }
    const mgb_adamw_chunk ck = table[blockIdx.x];
    const float gs = grad_scale * clip_coef(gradnorm_sq, grad_scale, max_norm);
    const float decay = 1.0f - lr * ck.weight_decay;
    const int64_t base = ck.offset;
    const int n = ck.numel;
    const bool vec = ((base & 3) == 0);
    const int n4 = vec ? (n >> 2) : 0;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        const int64_t o = base + (int64_t)i * 4;
        float4 pp = *reinterpret_cast<float4*>(p + o);
        const float4 gg = *reinterpret_cast<const float4*>(g + o);
        float4 mm = *reinterpret_cast<float4*>(m + o);
        float4 vv = *reinterpret_cast<float4*>(v + o);
        float* P = &pp.x; const float* G = &gg.x; float* M = &mm.x; float* V = &vv.x;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float gr = G[e] * gs;
            float pe = P[e] * decay;
            const float me = M[e] + (gr - M[e]) * (1.0f - beta1);
            const float ve = V[e] * beta2 + (1.0f - beta2) * gr * gr;
            const float denom = sqrtf(ve) * inv_bc2_sqrt + eps;
            pe -= step_size * (me / denom);
            P[e] = pe; M[e] = me; V[e] = ve;
        }
        *reinterpret_cast<float4*>(p + o) = pp;
        *reinterpret_cast<float4*>(m + o) = mm;
        *reinterpret_cast<float4*>(v + o) = vv;
        if (pb) {
            uint2 w;
            w.x = pack_bf16x2(pp.x, pp.y);
            w.y = pack_bf16x2(pp.z, pp.w);
            *reinterpret_cast<uint2*>(pb + o) = w;
        }
    }
    for (int i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) {
        const int64_t o = base + i;
        const float gr = g[o] * gs;
        float pe = p[o] * decay;
        const float me = m[o] + (gr - m[o]) * (1.0f - beta1);
        const float ve = v[o] * beta2 + (1.0f - beta2) * gr * gr;
        pe -= step_size * (me / (sqrtf(ve) * inv_bc2_sqrt + eps));
        p[o] = pe; m[o] = me; v[o] = ve;
        if (pb) pb[o] = __float2bfloat16(pe);
    }
}

// ---------------------------------------------------------------------------
// casts / axpy
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n) {
    const int64_t n8 = n >> 3;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 a = reinterpret_cast<const float4*>(src)[2 * i];
        const float4 b = reinterpret_cast<const float4*>(src)[2 * i + 1];
        stg128(dst + i * 8, make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y),
                                       pack_bf16x2(b.z, b.w)));
    }
    if (blockIdx.x == 0)
        for (int64_t i = (n8 << 3) + threadIdx.x; i < n; i += blockDim.x) dst[i] = __float2bfloat16(src[i]);
}
__global__ void __launch_bounds__(256)
cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = __bfloat162float(src[i]);
}
__global__ void __launch_bounds__(256)
axpy_kernel(float* __restrict__ dst, const float* __restrict__ src, float alpha, int64_t n) {
    const int64_t n4 = n >> 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 d = reinterpret_cast<float4*>(dst)[i];
        const float4 s = reinterpret_cast<const float4*>(src)[i];
        d.x += alpha * s.x; d.y += alpha * s.y; d.z += alpha * s.z; d.w += alpha * s.w;
        reinterpret_cast<float4*>(dst)[i] = d;
    }
    if (blockIdx.x == 0)
        for (int64_t i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) dst[i] += alpha * src[i];
}

__global__ void __launch_bounds__(256)
axpy_dev_kernel(float* __restrict__ dst, const float* __restrict__ src, const float* __restrict__ alpha_p, int64_t n) {
    const float alpha = *alpha_p;
    const int64_t n4 = n >> 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 d = reinterpret_cast<float4*>(dst)[i];
        const float4 s = reinterpret_cast<const float4*>(src)[i];
        d.x += alpha * s.x; d.y += alpha * s.y; d.z += alpha * s.z; d.w += alpha * s.w;
        reinterpret_cast<float4*>(dst)[i] = d;
    }
    if (blockIdx.x == 0)
        for (int64_t i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) dst[i] += alpha * src[i];
}
__global__ void __launch_bounds__(256)
scale_bf16_dev_kernel(__nv_bfloat16* __restrict__ x, int64_t n, const float* __restrict__ alpha_p) {
    const float alpha = *alpha_p;
    if (alpha == 1.0f) return;
    const int64_t n8 = n >> 3;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
        const uint4 v = ldg128(x + i * 8);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) { const float2 f = unpack_bf16x2(w[e]); o[e] = pack_bf16x2(f.x * alpha, f.y * alpha); }
        stg128(x + i * 8, make_uint4(o[0], o[1], o[2], o[3]));
    }
    if (blockIdx.x == 0)
        for (int64_t i = (n8 << 3) + threadIdx.x; i < n; i += blockDim.x)
            x[i] = __float2bfloat16(__bfloat162float(x[i]) * alpha);
}

// ---------------------------------------------------------------------------
// bias gradient: out[c] += sum_r x[r, c]
// block = 8 warps; a warp covers 256 columns (8 per lane), warps take interleaved rows of a 256-row slab
// ---------------------------------------------------------------------------
constexpr int CS_ROWS = 64;   // small slabs: a [32768 x 768] input still gives 1536 CTAs worth of loads in flight
__global__ void __launch_bounds__(256)
colsum_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld, int64_t rows, int64_t cols, float* out) {
    __shared__ float sh[8][256 + 8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t c0 = (int64_t)blockIdx.x * 256 + lane * 8;
    const int64_t r0 = (int64_t)blockIdx.y * CS_ROWS;
    const int64_t r1 = min(r0 + CS_ROWS, rows);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (c0 < cols) {
#pragma unroll 4
        for (int64_t r = r0 + warp; r < r1; r += 8) {
            const uint4 v = ldg128(x + r * ld + c0);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) { float2 f = unpack_bf16x2(w[e]); acc[2 * e] += f.x; acc[2 * e + 1] += f.y; }
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) sh[warp][lane * 8 + e] = acc[e];
    __syncthreads();
    const int c = threadIdx.x;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += sh[w][c];
    const int64_t col = (int64_t)blockIdx.x * 256 + c;
    if (col < cols) atomicAdd(out + col, s);
}
__global__ void __launch_bounds__(256)
colsum_scalar_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld, int64_t rows, int64_t cols, float* out) {
    const int64_t col = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= cols) return;
    const int64_t r0 = (int64_t)blockIdx.y * CS_ROWS, r1 = min(r0 + CS_ROWS, rows);
    float s = 0.f;
    for (int64_t r = r0; r < r1; ++r) s += __bfloat162float(x[r * ld + col]);
    atomicAdd(out + col, s);
}

// ---------------------------------------------------------------------------
// standalone activation fwd / bwd (nn.GELU at mingpt/model.py:182 when not fused)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
act_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ z,
               __nv_bfloat16* __restrict__ dz, int64_t n, int act) {
    const int64_t n8 = n >> 3;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
        const uint4 a = ldg128(dy + i * 8), b = ldg128(z + i * 8);
        const uint32_t wa[4] = {a.x, a.y, a.z, a.w}, wb[4] = {b.x, b.y, b.z, b.w};
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 fa = unpack_bf16x2(wa[e]), fb = unpack_bf16x2(wb[e]);
            float g0, g1;
            if (act == MGB_GATE_MUL) { g0 = fb.x; g1 = fb.y; }          // z already holds the derivative
            else if (act == MGB_ACT_GELU_TANH) { g0 = gelu_tanh_grad_f(fb.x); g1 = gelu_tanh_grad_f(fb.y); }
            else if (act == MGB_ACT_GELU_ERF) { g0 = gelu_erf_grad_f(fb.x); g1 = gelu_erf_grad_f(fb.y); }
            else { g0 = 1.f; g1 = 1.f; }
            o[e] = pack_bf16x2(fa.x * g0, fa.y * g1);
        }
        stg128(dz + i * 8, make_uint4(o[0], o[1], o[2], o[3]));
    }
    if (blockIdx.x == 0)
        for (int64_t i = (n8 << 3) + threadIdx.x; i < n; i += blockDim.x) {
            const float zz = __bfloat162float(z[i]);
            const float g = act == MGB_GATE_MUL ? zz : act == MGB_ACT_GELU_TANH ? gelu_tanh_grad_f(zz) : (act == MGB_ACT_GELU_ERF ? gelu_erf_grad_f(zz) : 1.f);
            dz[i] = __float2bfloat16(__bfloat162float(dy[i]) * g);
        }
}
__global__ void __launch_bounds__(256)
act_fwd_kernel(const __nv_bfloat16* __restrict__ z, const __nv_bfloat16* res, __nv_bfloat16* y, int64_t n, int act) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float v = __bfloat162float(z[i]);
        v = act == MGB_ACT_GELU_TANH ? gelu_tanh_f(v) : (act == MGB_ACT_GELU_ERF ? gelu_erf_f(v) : v);
        if (res) v += __bfloat162float(res[i]);
        y[i] = __float2bfloat16(v);
    }
}

}  // namespace mgb

using namespace mgb;

static int grid_for(int64_t work_items, int per_block, int max_blocks) {
    int64_t g = (work_items + per_block - 1) / per_block;
    if (g < 1) g = 1;
    if (g > max_blocks) g = max_blocks;
    return (int)g;
}

extern "C" const char* mgb_last_error(void) { return g_err; }
extern "C" int mgb_version(void) { return 100; }
extern "C" int64_t mgb_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
namespace mgb {
std::atomic<int> g_opt_gemm_single{0};
std::atomic<int> g_opt_gemm_direct_store{0};
std::atomic<int> g_opt_gemm_generic{0};
std::atomic<int> g_opt_gemm_static{0};
std::atomic<int> g_opt_pdl{1};   // programmatic dependent launch: -3 % on the launch-bound gpt-nano / gpt-mini steps, 0-1 % on gpt2-124M
bool pdl_enabled() { return g_opt_pdl.load(std::memory_order_relaxed) != 0; }
std::atomic<int> g_opt_attn_mma{0};
std::atomic<int> g_opt_attn_poly{1};
std::atomic<int> g_opt_ln_unstaged{0};
std::atomic<int> g_opt_ce_unstaged{0};
}
extern "C" int mgb_set_option(const char* key, int value) {
    if (!key) return -1;
    if (!std::strcmp(key, "gemm_single")) return mgb::g_opt_gemm_single.exchange(value);
    if (!std::strcmp(key, "gemm_direct_store")) return mgb::g_opt_gemm_direct_store.exchange(value);
    if (!std::strcmp(key, "gemm_generic")) return mgb::g_opt_gemm_generic.exchange(value);
    if (!std::strcmp(key, "gemm_static")) return mgb::g_opt_gemm_static.exchange(value);
    if (!std::strcmp(key, "pdl")) return mgb::g_opt_pdl.exchange(value);
    if (!std::strcmp(key, "attn_mma")) return mgb::g_opt_attn_mma.exchange(value);
    if (!std::strcmp(key, "attn_poly")) return mgb::g_opt_attn_poly.exchange(value);
    if (!std::strcmp(key, "ln_unstaged")) return mgb::g_opt_ln_unstaged.exchange(value);
    if (!std::strcmp(key, "ce_unstaged")) return mgb::g_opt_ce_unstaged.exchange(value);
    if (!std::strcmp(key, "reserved_sms")) return mgb::g_opt_reserved_sms.exchange(value);
    return -1;
}

extern "C" int mgb_sumsq_f32(const float* x, int64_t n, float* out, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (n < 0) { set_last_error("sumsq: bad shape"); return kErrBadShape; }
    if ((uintptr_t)x % 16) { set_last_error("sumsq: x must be 16-byte aligned"); return kErrBadAlign; }
    if (n == 0) return kOk;
    sumsq_kernel<<<grid_for(n, 512 * 16, num_sms() * 4), 512, 0, s>>>(x, n, out);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_clip_scale_f32(float* g, int64_t n, const float* gradnorm_sq, float max_norm, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (n < 0) { set_last_error("clip_scale: bad shape"); return kErrBadShape; }
    if (n == 0) return kOk;
    clip_scale_kernel<<<grid_for(n, 256 * 8, num_sms() * 8), 256, 0, s>>>(g, n, gradnorm_sq, max_norm);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_adamw_multi(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, void* params_bf16,
                               const mgb_adamw_chunk* table, int32_t num_chunks, float lr, float beta1, float beta2,
                               float eps, int32_t step, float grad_scale, const float* gradnorm_sq, float max_norm,
                               const int32_t* step_dev, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (num_chunks < 0 || (!step_dev && step < 1)) { set_last_error("adamw: bad num_chunks/step"); return kErrBadShape; }
    if (num_chunks == 0) return kOk;
    float step_size = 0.f, inv_bc2_sqrt = 0.f;
    if (!step_dev) {
        const double bc1 = 1.0 - std::pow((double)beta1, (double)step);
        const double bc2 = 1.0 - std::pow((double)beta2, (double)step);
        step_size = (float)((double)lr / bc1);
        inv_bc2_sqrt = (float)(1.0 / std::sqrt(bc2));
    }
    adamw_kernel<<<num_chunks, 256, 0, s>>>(params, grads, exp_avg, exp_avg_sq, (__nv_bfloat16*)params_bf16, table, lr,
                                            beta1, beta2, eps, step_size, inv_bc2_sqrt, grad_scale, gradnorm_sq,
                                            max_norm, step_dev);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_cast_f32_to_bf16(const float* src, void* dst, int64_t n, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (n < 0) { set_last_error("cast: bad shape"); return kErrBadShape; }
    if (((uintptr_t)src % 16) || ((uintptr_t)dst % 16)) { set_last_error("cast: pointers must be 16-byte aligned"); return kErrBadAlign; }
    if (n == 0) return kOk;
    cast_f32_bf16_kernel<<<grid_for(n, 256 * 8, num_sms() * 8), 256, 0, s>>>(src, (__nv_bfloat16*)dst, n);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}
extern "C" int mgb_cast_bf16_to_f32(const void* src, float* dst, int64_t n, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (n < 0) { set_last_error("cast: bad shape"); return kErrBadShape; }
    if (n == 0) return kOk;
    cast_bf16_f32_kernel<<<grid_for(n, 256 * 4, num_sms() * 8), 256, 0, s>>>((const __nv_bfloat16*)src, dst, n);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}
extern "C" int mgb_axpy_f32(float* dst, const float* src, float alpha, int64_t n, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (n < 0) { set_last_error("axpy: bad shape"); return kErrBadShape; }
    if (((uintptr_t)src % 16) || ((uintptr_t)dst % 16)) { set_last_error("axpy: pointers must be 16-byte aligned"); return kErrBadAlign; }
    if (n == 0) return kOk;
    axpy_kernel<<<grid_for(n, 256 * 8, num_sms() * 8), 256, 0, s>>>(dst, src, alpha, n);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_axpy_f32_dev(float* dst, const float* src, const float* alpha, int64_t n, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (n < 0 || !alpha) { set_last_error("axpy_dev: bad arguments"); return kErrBadShape; }
    if (((uintptr_t)src % 16) || ((uintptr_t)dst % 16)) { set_last_error("axpy_dev: pointers must be 16-byte aligned"); return kErrBadAlign; }
    if (n == 0) return kOk;
    axpy_dev_kernel<<<grid_for(n, 256 * 8, num_sms() * 8), 256, 0, s>>>(dst, src, alpha, n);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}
extern "C" int mgb_scale_bf16_dev(void* x, int64_t n, const float* alpha, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (n < 0 || !alpha) { set_last_error("scale_dev: bad arguments"); return kErrBadShape; }
    if ((uintptr_t)x % 16) { set_last_error("scale_dev: x must be 16-byte aligned"); return kErrBadAlign; }
    if (n == 0) return kOk;
    scale_bf16_dev_kernel<<<grid_for(n, 256 * 8, num_sms() * 8), 256, 0, s>>>((__nv_bfloat16*)x, n, alpha);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_colsum_bf16(const void* x, int64_t ld, int64_t rows, int64_t cols, float* out, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (rows < 0 || cols <= 0 || ld < cols) { set_last_error("colsum: bad shape"); return kErrBadShape; }
    if (rows == 0) return kOk;
    const int64_t gy = (rows + CS_ROWS - 1) / CS_ROWS;
    if (gy > 65535) { set_last_error("colsum: too many rows"); return kErrTooLarge; }
    const bool vec = (ld % 8 == 0) && (cols % 8 == 0) && ((uintptr_t)x % 16 == 0);
    dim3 grid((unsigned)((cols + 255) / 256), (unsigned)gy);
    if (vec) colsum_kernel<<<grid, 256, 0, s>>>((const __nv_bfloat16*)x, ld, rows, cols, out);
    else colsum_scalar_kernel<<<grid, 256, 0, s>>>((const __nv_bfloat16*)x, ld, rows, cols, out);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_act_bwd_bf16(const void* dy, const void* z, void* dz, int64_t n, int32_t act, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (n < 0) { set_last_error("act_bwd: bad shape"); return kErrBadShape; }
    if (((uintptr_t)dy % 16) || ((uintptr_t)z % 16) || ((uintptr_t)dz % 16)) { set_last_error("act_bwd: pointers must be 16-byte aligned"); return kErrBadAlign; }
    if (n == 0) return kOk;
    act_bwd_kernel<<<grid_for(n, 256 * 8, num_sms() * 8), 256, 0, s>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)z,
                                                                      (__nv_bfloat16*)dz, n, act);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}
extern "C" int mgb_act_fwd_bf16(const void* z, const void* res, void* y, int64_t n, int32_t act, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (n < 0) { set_last_error("act_fwd: bad shape"); return kErrBadShape; }
    if (n == 0) return kOk;
    act_fwd_kernel<<<grid_for(n, 256 * 4, num_sms() * 8), 256, 0, s>>>((const __nv_bfloat16*)z, (const __nv_bfloat16*)res,
                                                                      (__nv_bfloat16*)y, n, act);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}
