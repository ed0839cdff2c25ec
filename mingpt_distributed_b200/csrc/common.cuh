// Common device helpers for the sm_100a kernels: PTX wrappers for mbarrier,
// TMA (cp.async.bulk.tensor), tcgen05 (UMMA + TMEM), and small math utilities.
//
// Everything here is B200-only (compile with -gencode arch=compute_100a,code=sm_100a).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <mutex>

namespace mgb {

// ----------------------------------------------------------------------------
// error codes shared with include/mingpt_b200.h
// ----------------------------------------------------------------------------
enum : int {
    kOk = 0,
    kErrBadShape = -1,
    kErrBadAlign = -2,
    kErrBadDtype = -3,
    kErrUnsupported = -4,
    kErrLaunch = -5,
    kErrDriver = -6,
    kErrTooLarge = -7,
};

#define MGB_CHECK_LAUNCH()                                         \
    do {                                                           \
        cudaError_t e__ = cudaGetLastError();                      \
        if (e__ != cudaSuccess) {                                  \
            mgb::set_last_error(cudaGetErrorString(e__));          \
            return mgb::kErrLaunch;                                \
        }                                                          \
    } while (0)

void set_last_error(const char* msg);
void count_launch(int n);   // bumps mgb_launch_count()
int num_sms();              // SMs of the CURRENT device minus mgb_set_option("reserved_sms"): what persistent grids are sized to
int num_sms_total();        // SMs of the current device

// One-time set-up that is PER DEVICE (cudaFuncSetAttribute opt-ins, attribute queries): run(f) calls f() the first
// time it is reached with each device ordinal current, so a process that drives several GPUs gets the opt-in on
// every one of them (a process-wide std::once_flag would leave all but the first device without it).
struct PerDeviceOnce {
    std::atomic<uint64_t> done[2] = {{0}, {0}};   // ordinals 0..127
    std::mutex mu;
    template <typename F>
    void run(F&& f) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev > 127) dev = 0;
        const uint64_t bit = 1ull << (dev & 63);
        std::atomic<uint64_t>& word = done[dev >> 6];
        if (word.load(std::memory_order_acquire) & bit) return;
        std::lock_guard<std::mutex> lock(mu);
        if (word.load(std::memory_order_relaxed) & bit) return;
        f();
        word.fetch_or(bit, std::memory_order_release);
    }
};
bool pdl_enabled();   // optim.cu: mgb_set_option("pdl", ...)

// Launch with programmatic stream serialization when enabled (the kernel MUST call grid_dep_wait() before it reads
// anything a previous kernel wrote), as an ordinary launch otherwise.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ----------------------------------------------------------------------------
// small utilities
// ----------------------------------------------------------------------------
__host__ __device__ __forceinline__ int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
__host__ __device__ __forceinline__ int ceil_div(int a, int b) { return (a + b - 1) / b; }

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
    __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
    return __bfloat1622float2(v);
}

// Packed fp32 pairs (Blackwell FADD2 / FMUL2 / FFMA2: one issue slot for two fp32 lanes). The epilogues and the softmax
// are bound by instruction issue, not by the fp32 pipes, so halving the instruction count of their elementwise math
// is a direct gain. A float2 lives in an aligned 64-bit register pair.
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
    float2 c;
    asm("add.f32x2 %0, %1, %2;" : "=l"(reinterpret_cast<uint64_t&>(c)) : "l"(reinterpret_cast<const uint64_t&>(a)), "l"(reinterpret_cast<const uint64_t&>(b)));
    return c;
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
    float2 c;
    asm("mul.f32x2 %0, %1, %2;" : "=l"(reinterpret_cast<uint64_t&>(c)) : "l"(reinterpret_cast<const uint64_t&>(a)), "l"(reinterpret_cast<const uint64_t&>(b)));
    return c;
}
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    float2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(reinterpret_cast<uint64_t&>(d))
        : "l"(reinterpret_cast<const uint64_t&>(a)), "l"(reinterpret_cast<const uint64_t&>(b)), "l"(reinterpret_cast<const uint64_t&>(c)));
    return d;
}
__device__ __forceinline__ float2 splat2(float a) { return make_float2(a, a); }

// 2^x for a pair WITHOUT the special-function unit: Cody-Waite range reduction (round-to-nearest through the 1.5 * 2^23
// magic constant) and a cubic minimax polynomial for 2^f on [-0.5, 0.5] (max relative error 7.5e-5, far below the bf16
// rounding of the probabilities it feeds), the integer part added straight into the exponent field. MUFU.EX2 runs at 16
// per clock and SM, which is what bounds the attention softmax; a share of the exponentials moved to the FMA pipe
// (6 packed instructions + 2 integer ones per pair) runs beside it. Inputs are clamped at -126 (result ~1e-38).
__device__ __forceinline__ float2 exp2_poly2(float2 x) {
    x.x = fmaxf(x.x, -126.0f);
    x.y = fmaxf(x.y, -126.0f);
    const float2 fl = add2(x, splat2(12582912.0f));
    const float2 f = add2(x, fma2(fl, splat2(-1.0f), splat2(12582912.0f)));      // x - round(x)
    float2 pl = fma2(splat2(0.0551716648042202f), f, splat2(0.2426111251115799f));
    pl = fma2(pl, f, splat2(0.6932609677314758f));
    pl = fma2(pl, f, splat2(0.9999280571937561f));
    return make_float2(__int_as_float(__float_as_int(pl.x) + (__float_as_int(fl.x) << 23)),
                       __int_as_float(__float_as_int(pl.y) + (__float_as_int(fl.y) << 23)));
}

// GELU, both flavours the reference family uses.
//  tanh ("NewGELU", README.md / north_star MLP) and exact erf (nn.GELU(), mingpt/model.py:182)
__device__ __forceinline__ float gelu_tanh_f(float x) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    float u = k0 * (x + k1 * x * x * x);
    return 0.5f * x * (1.0f + tanhf(u));
}
// with h = (1 + tanh u) / 2:  gelu = x h,  gelu' = h (1 + 2 x (1 - h) u'),  u' = k0 (1 + 3 k1 x^2)
__device__ __forceinline__ void gelu_tanh_both_f(float x, float& y, float& g) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    const float x2 = x * x;
    const float u = x * fmaf(k0 * k1, x2, k0);
    const float h = fmaf(tanhf(u), 0.5f, 0.5f);
    const float q = x * fmaf(6.0f * k0 * k1, x2, 2.0f * k0);      // 2 x u'
    y = x * h;
    g = fmaf(h, fmaf(-h, q, q), h);
}
// two elements at a time: 5 packed ops + 2 tanh for the value, 9 + 2 for value and derivative
__device__ __forceinline__ float2 gelu_tanh_2(float2 x) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    const float2 x2 = mul2(x, x);
    const float2 u = mul2(x, fma2(x2, splat2(k0 * k1), splat2(k0)));
    const float2 t = make_float2(tanhf(u.x), tanhf(u.y));
    return mul2(x, fma2(t, splat2(0.5f), splat2(0.5f)));
}
__device__ __forceinline__ void gelu_tanh_both_2(float2 x, float2& y, float2& g) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    const float2 x2 = mul2(x, x);
    const float2 u = mul2(x, fma2(x2, splat2(k0 * k1), splat2(k0)));
    const float2 t = make_float2(tanhf(u.x), tanhf(u.y));
    const float2 h = fma2(t, splat2(0.5f), splat2(0.5f));
    const float2 omh = fma2(t, splat2(-0.5f), splat2(0.5f));            // 1 - h
    const float2 q = fma2(x2, splat2(6.0f * k0 * k1), splat2(2.0f * k0));  // 2 u'
    y = mul2(x, h);
    g = fma2(mul2(y, omh), q, h);                                        // h + x h (1 - h) 2 u'
}
__device__ __forceinline__ float gelu_tanh_grad_f(float x) {
    float y, g;
    gelu_tanh_both_f(x, y, g);
    return g;
}
__device__ __forceinline__ float gelu_erf_f(float x) {
    return 0.5f * x * (1.0f + erff(x * 0.7071067811865476f));
}
__device__ __forceinline__ void gelu_erf_both_f(float x, float& y, float& g) {
    const float inv_sqrt_2pi = 0.3989422804014327f;
    const float cdf = 0.5f * (1.0f + erff(x * 0.7071067811865476f));
    y = x * cdf;
    g = cdf + x * inv_sqrt_2pi * __expf(-0.5f * x * x);
}
__device__ __forceinline__ float gelu_erf_grad_f(float x) {
    float y, g;
    gelu_erf_both_f(x, y, g);
    return g;
}

// ----------------------------------------------------------------------------
// mbarrier
// ----------------------------------------------------------------------------
// One lane of a CONVERGED warp. The producer / MMA warps run their loops with all 32 lanes (addresses and
// descriptors stay warp-uniform, so they live in uniform registers) and predicate only the TMA / tcgen05
// instructions on this: issuing them from inside an `if (lane == 0)` region makes the compiler wrap every
// single one in an ELECT / R2UR / BRA.U.ANY loop, ~100 cycles per instruction (tools/micro/mma_rate.cu).
__device__ __forceinline__ bool elect_one() {
    uint32_t pred = 0;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return done != 0;
}
// try_wait with a suspend-time hint: the hardware may keep the thread asleep for up to `ns` nanoseconds while the
// phase is incomplete (it is woken by the completing arrive), instead of returning after its short default limit.
__device__ __forceinline__ bool mbar_try_wait_hint(uint64_t* bar, uint32_t parity, uint32_t ns) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity), "r"(ns)
        : "memory");
    return done != 0;
}
// Wait with a watchdog: a protocol bug turns into a trap ("unspecified launch failure")
// after ~4 s instead of a hung GPU box. The plain try_wait already suspends the thread in the barrier unit for a
// hardware time limit, so most waits end inside the first probe and the loop body runs rarely. (Tried and dropped:
// try_wait with a large suspend-time hint -- on this part it returns at once instead of sleeping longer, so the
// loop spun through 28 % of the attention backward's issued instructions: profiles/r2_attn_ncu.md.)
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    uint64_t t0 = 0;
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++spins & 0xfffu) == 0) {
            uint64_t now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > 4000000000ull) { __trap(); }
        }
    }
}
// (kept as a name: the attention and row kernels' waits, where the waiting warps share schedulers with busy ones)
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) { mbar_wait(bar, parity); }
// For warps that wait most of the time and whose wake-up may lag by ~ns nanoseconds (a TMA producer several stages
// ahead, drain warps): a failed probe is followed by a real sleep. try_wait alone comes back after a few tens of
// cycles, and the polling of five such warps was 24 % of the attention backward's issued instructions
// (profiles/r2b_attn_bwd_ncu.md) -- issue slots and, on a power-capped part, energy taken from the busy warps.
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity, uint32_t ns) {
    if (mbar_try_wait(bar, parity)) return;
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        __nanosleep(ns);
        if (++spins > (1u << 25)) __trap();          // seconds: a protocol bug, not a wait
    }
}

__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ----------------------------------------------------------------------------
// TMA
// ----------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tm)) : "memory");
}
// 2-D tiled load: coordinates are (inner, outer) in elements.
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int32_t c0,
                                            int32_t c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
// Programmatic dependent launch: a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start
// while its predecessor in the stream still runs; grid_dep_wait() blocks until that predecessor has completed and
// its writes are visible (a no-op for ordinary launches), grid_dep_launch() lets the successor start early.
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// 1-D bulk copy global -> shared (no tensor map): `bytes` and both addresses multiples of 16.
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// 1-D bulk copy shared -> global (completion through bulk_commit / bulk_wait).
__device__ __forceinline__ void bulk_store_1d(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int32_t c0,
                                            int32_t c1, int32_t c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1),
        "r"(c2)
        : "memory");
}

// ----------------------------------------------------------------------------
// tcgen05 / TMEM
// ----------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// tcgen05.commit: arrive on an mbarrier when all prior tcgen05.mma of this thread retire.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 inputs, fp32 accumulate.
__device__ __forceinline__ void umma_bf16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// A operand from TMEM (bf16 packed), B from smem.
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}

// Instruction descriptor for kind::f16 with BF16 A/B and FP32 D
// (bit layout: cute/arch/mma_sm100_desc.hpp InstrDescriptor).
__host__ __device__ __forceinline__ uint32_t make_idesc_bf16(int m, int n, int a_mn_major, int b_mn_major) {
    uint32_t d = 0;
    d |= 1u << 4;                              // c_format = F32
    d |= 1u << 7;                              // a_format = BF16
    d |= 1u << 10;                             // b_format = BF16
    d |= (uint32_t)(a_mn_major & 1) << 15;     // a_major
    d |= (uint32_t)(b_mn_major & 1) << 16;     // b_major
    d |= (uint32_t)(n >> 3) << 17;             // n_dim
    d |= (uint32_t)(m >> 4) << 24;             // m_dim
    return d;
}

// Shared-memory matrix descriptor, SWIZZLE_128B, version 1 (sm_100).
__device__ __forceinline__ uint64_t make_smem_desc_sw128(uint32_t smem_addr, uint32_t lbo_bytes,
                                                         uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3ffffu) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;
    d |= (uint64_t)1 << 46;   // version = 1
    d |= (uint64_t)2 << 61;   // layout_type = SWIZZLE_128B
    return d;
}

// TMEM -> registers: this warp's 32 lanes x 32 consecutive fp32 columns.
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ----------------------------------------------------------------------------
// CTA pairs (cta_group::2): cluster helpers, 2-SM TMA loads, 2-SM MMA, multicast commit
// ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive (no tx) on the barrier at the same smem offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t cta) {
    asm volatile(
        "{\n\t.reg .b32 remAddr32;\n\t"
        "mapa.shared::cluster.u32 remAddr32, %0, %1;\n\t"
        "mbarrier.arrive.shared::cluster.b64 _, [remAddr32];\n\t}"
        ::"r"(smem_u32(bar)), "r"(cta)
        : "memory");
}
// 2-SM TMA load: data lands in THIS CTA's smem, the transaction bytes are credited to the barrier at the
// same offset in the pair's leader CTA (peer bit of the shared::cluster address cleared).
__device__ __forceinline__ void tma_load_2d_cg2(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int32_t c0,
                                                int32_t c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0),
        "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc_cg2(uint32_t* smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_cg2(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_bf16_ss_cg2(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                 uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on the barrier at this smem offset in every CTA of `mask` once the pair's prior MMAs retire
__device__ __forceinline__ void umma_commit_cg2_mc(uint64_t* bar, uint16_t mask) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
        ::"r"(smem_u32(bar)), "h"(mask)
        : "memory");
}

// ----------------------------------------------------------------------------
// Cluster launch control (Blackwell's hardware work queue): a RUNNING cluster asks to cancel a cluster of the same
// grid that has not started yet and, if it gets one, does that cluster's work itself. The grid is launched with one
// cluster per work unit; whoever holds SMs pulls units until none is pending, so SMs that are busy with something else
// (NCCL's CTAs during the gradient all-reduce) simply pull fewer units instead of holding a static share hostage.
// The 16-byte response lands in shared memory of EVERY CTA of the asking cluster (same offset), each CTA's mbarrier
// at `bar`'s offset receives 16 transaction bytes.
// ----------------------------------------------------------------------------
__device__ __forceinline__ void clc_try_cancel_multicast(void* resp_smem, uint64_t* bar) {
    asm volatile(
        "clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.multicast::cluster::all.b128 [%0], [%1];"
        ::"r"(smem_u32(resp_smem)), "r"(smem_u32(bar))
        : "memory");
}
// true: a pending cluster was cancelled, `ctaid_x` is blockIdx.x of its first CTA; false: nothing left to take
__device__ __forceinline__ bool clc_decode(const void* resp_smem, uint32_t& ctaid_x) {
    uint32_t valid = 0, x = 0, y = 0, z = 0;
    asm volatile(
        "{\n\t.reg .pred p1;\n\t.reg .b128 resp;\n\t"
        "ld.shared.b128 resp, [%4];\n\t"
        "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p1, resp;\n\t"
        "selp.u32 %3, 1, 0, p1;\n\t"
        "@p1 clusterlaunchcontrol.query_cancel.get_first_ctaid.v4.b32.b128 {%0, %1, %2, _}, resp;\n\t}"
        : "=r"(x), "=r"(y), "=r"(z), "=r"(valid)
        : "r"(smem_u32(resp_smem))
        : "memory");
    ctaid_x = x;
    return valid != 0;
}

// ----------------------------------------------------------------------------
// TMA stores (smem -> global), plain and fp32 reduce-add; bulk async-group bookkeeping
// ----------------------------------------------------------------------------
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tm, const void* smem_src, int32_t c0, int32_t c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* tm, const void* smem_src, int32_t c0,
                                                  int32_t c1) {
    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ----------------------------------------------------------------------------
// vector global access helpers
// ----------------------------------------------------------------------------
// Shared-memory vector accesses in the shared state space. Pointers into the dynamic shared-memory carve-up
// have lost their address space (the 1024-byte alignment goes through an integer), so a plain dereference
// compiles to a GENERIC LD.E / ST.E (long-scoreboard latency); these keep them LDS / STS.
__device__ __forceinline__ uint4 lds128(const void* p) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(p)));
    return v;
}
__device__ __forceinline__ float4 lds128f(const void* p) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(smem_u32(p)));
    return v;
}
__device__ __forceinline__ void sts128(void* p, uint4 v) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(p)), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void sts128f(void* p, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(p)), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ldg128(const void* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void stg128(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }
__device__ __forceinline__ void red_add_f32x4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

}  // namespace mgb
