// BF16 GEMM for sm_100a: TMA -> shared memory (SWIZZLE_128B) -> tcgen05.mma -> TMEM -> fused epilogue.
//
// Replaces every nn.Linear / F.linear call on the reference's training step
// (mingpt/model.py:138,180-181,249 and torch/nn/functional.py:6478,6653 inside
// nn.MultiheadAttention) and their autograd backward (dgrad + wgrad).
//
//   D[m, n] = epilogue( sum_k A(m, k) * B(n, k) )
//
// A and B are BF16 and may each be "K-major" (reduction index contiguous in memory) or
// "MN-major" (output index contiguous), so the three layouts of a linear layer map onto one kernel:
//   forward  Y = X W^T      : A = X  [M,K] K-major,  B = W  [N,K] K-major
//   dgrad    dX = dY W      : A = dY [M,N] K-major,  B = W  [N,K] viewed as [Kred=N, Nout=K] MN-major
//   wgrad    dW = dY^T X    : A = dY [M,N] viewed as [Kred=M, Mout=N] MN-major, B = X likewise MN-major
//
// One persistent CTA per SM, warp specialised:
//   warp 0      TMA producer        (one lane)
//   warp 1      tcgen05.mma issuer  (one lane)
//   warp 2      TMEM allocator
//   warps 4-11  epilogue: tcgen05.ld -> bias / GELU / gate / residual -> global stores
// Two TMEM accumulator stages let the epilogue of tile i overlap the MMAs of tile i+1.
#include "gemm_common.cuh"
#include "dropout.cuh"
static_assert(sizeof(mgb_gemm_args) == 232, "mgb_gemm_args is mirrored field by field in _lib.py (GemmArgs) and INTEGRATION.md");

#include <cstdlib>

namespace mgb {
namespace gemm {

constexpr int kStages = 4;
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;        // 16 KiB
constexpr int B_STAGE_BYTES = MAX_BLOCK_N * BLOCK_K * 2;    // 32 KiB
constexpr int SMEM_BYTES = kStages * (A_STAGE_BYTES + B_STAGE_BYTES) + 1024 /*align slack*/ + 256 /*barriers*/;

__global__ void __launch_bounds__(kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
            const Params p) {
    extern __shared__ uint8_t smem_raw[];
    // SWIZZLE_128B tiles need 1024-byte alignment.
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + kStages * A_STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * (A_STAGE_BYTES + B_STAGE_BYTES));
    uint64_t* full_bar = bars;                  // [kStages]
    uint64_t* empty_bar = bars + kStages;       // [kStages]
    uint64_t* tmem_full = bars + 2 * kStages;   // [2]
    uint64_t* tmem_empty = tmem_full + 2;       // [2]
    uint32_t* tmem_base_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // provably warp-uniform
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], kEpiWarps);
        }
        fence_mbar_init();
    }
    if (warp == 2) tmem_alloc(tmem_base_ptr, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_ptr;

    const int tiles = p.num_m_tiles * p.num_n_tiles;
    const int units = tiles * p.split_k;
    const int block_n = p.block_n;
    const uint32_t stage_tx = A_STAGE_BYTES + (uint32_t)block_n * BLOCK_K * 2;

    if (warp == 0) {
        // ===================== TMA producer (whole warp runs the loop, one lane issues) =====================
        {
            const bool issuer = elect_one();
            uint32_t stage = 0, phase = 0;
            for (int u = blockIdx.x; u < units; u += gridDim.x) {
                const int split = u / tiles;
                const int t = u - split * tiles;
                const int m0 = (t / p.num_n_tiles) * BLOCK_M;
                const int n0 = (t % p.num_n_tiles) * block_n;
                const int kb0 = split * p.k_blocks_per_split;
                const int kb1 = min(kb0 + p.k_blocks_per_split, p.k_blocks_total);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem_a + stage * A_STAGE_BYTES;
                    uint8_t* sb = smem_b + stage * B_STAGE_BYTES;
                    const int k0 = kb * BLOCK_K;
                    if (issuer) {
                        mbar_arrive_expect_tx(&full_bar[stage], stage_tx);
                        if (!p.a_mn) {
                            tma_load_2d(sa, &tmap_a, &full_bar[stage], k0, m0);
                        } else {
                            tma_load_2d(sa, &tmap_a, &full_bar[stage], m0, k0);
                            tma_load_2d(sa + ATOM_BYTES, &tmap_a, &full_bar[stage], m0 + 64, k0);
                        }
                        if (!p.b_mn) {
                            tma_load_2d(sb, &tmap_b, &full_bar[stage], k0, n0);
                        } else {
                            for (int a = 0; a < block_n / 64; ++a)
                                tma_load_2d(sb + a * ATOM_BYTES, &tmap_b, &full_bar[stage], n0 + a * 64, k0);
                        }
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (whole warp runs the loop, one lane issues) =====================
        {
            const bool issuer = elect_one();
            const uint32_t idesc = make_idesc_bf16(BLOCK_M, block_n, p.a_mn, p.b_mn);
            // K-major: 8-row groups are 1024 B apart (SBO), LBO unused; advance 32 B per UMMA_K.
            // MN-major: 8 k-rows of 128 B per group (SBO = 1024), 64-wide atoms BLOCK_K*128 B apart (LBO);
            //           advance 16 k-rows = 2048 B per UMMA_K.
            const uint32_t a_lbo = p.a_mn ? ATOM_BYTES : 0, b_lbo = p.b_mn ? ATOM_BYTES : 0;
            const uint32_t a_kstep = p.a_mn ? UMMA_K * 128 : UMMA_K * 2;
            const uint32_t b_kstep = p.b_mn ? UMMA_K * 128 : UMMA_K * 2;
            uint32_t stage = 0, phase = 0;
            uint32_t acc = 0, acc_phase = 0;
            for (int u = blockIdx.x; u < units; u += gridDim.x) {
                const int split = u / tiles;
                const int kb0 = split * p.k_blocks_per_split;
                const int kb1 = min(kb0 + p.k_blocks_per_split, p.k_blocks_total);
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * kAccStride;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem_a + stage * A_STAGE_BYTES);
                    const uint32_t sb = smem_u32(smem_b + stage * B_STAGE_BYTES);
                    if (issuer) {
#pragma unroll
                        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                            const uint64_t da = make_smem_desc_sw128(sa + k * a_kstep, a_lbo, 1024);
                            const uint64_t db = make_smem_desc_sw128(sb + k * b_kstep, b_lbo, 1024);
                            umma_bf16_ss(tmem_d, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                        }
                        umma_commit(&empty_bar[stage]);   // frees the smem slot when these MMAs retire
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                if (issuer) umma_commit(&tmem_full[acc]);   // accumulator ready for the epilogue
                __syncwarp();
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const int ew = warp - 4;
        const int quarter = warp & 3;           // TMEM lane quarter this warp may access
        const int half = ew >> 2;               // which half of the tile's columns
        const int cols_per_warp = block_n >> 1; // multiple of 32
        const bool vec_ok = ((p.ldd & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.d) & 15) == 0) &&
                            (!p.residual || (((p.ldr & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.residual) & 15) == 0))) &&
                            (!p.preact || (((p.ldp & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.preact) & 15) == 0))) &&
                            (!p.gate || (((p.ldg & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.gate) & 15) == 0)));
        uint32_t acc = 0, acc_phase = 0;
        for (int u = blockIdx.x; u < units; u += gridDim.x) {
            const int split = u / tiles;
            const int t = u - split * tiles;
            const int m0 = (t / p.num_n_tiles) * BLOCK_M;
            const int n0 = (t % p.num_n_tiles) * block_n;
            mbar_wait(&tmem_full[acc], acc_phase);
            tc_fence_after();
            const int64_t row = (int64_t)m0 + quarter * 32 + lane;
            const bool row_ok = row < p.M;
            const uint32_t taddr0 = tmem_base + acc * kAccStride + ((uint32_t)(quarter * 32) << 16) +
                                    (uint32_t)(half * cols_per_warp);
            const int nchunks = cols_per_warp >> 5;
            for (int c = 0; c < nchunks; ++c) {
                uint32_t r[32];
                __syncwarp();   // tcgen05.ld is .sync.aligned: reconverge after the predicated body below
                tmem_ld_32x32b_x32(taddr0 + c * 32, r);
                tmem_ld_wait();
                if (c == nchunks - 1) {
                    // all TMEM reads of this accumulator stage are done: hand it back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tmem_empty[acc]);
                }
                const int64_t col0 = (int64_t)n0 + half * cols_per_warp + c * 32;
                if (row_ok && col0 < p.N) {
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;
                const bool full = (col0 + 32 <= p.N) && vec_ok;
                if (p.bias) {
                    if (col0 + 32 <= p.N) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] += __ldg(p.bias + col0 + j);
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (col0 + j < p.N) v[j] += __ldg(p.bias + col0 + j);
                    }
                }
                if (full) {
                    bool acted = false;
                    if (p.preact) {
                        __nv_bfloat16* dst = p.preact + row * p.ldp + col0;
                        uint32_t pw[16];
                        acted = preact_words(v, pw, p.act, p.act_grad);
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            stg128(dst + q * 8, make_uint4(pw[q * 4], pw[q * 4 + 1], pw[q * 4 + 2], pw[q * 4 + 3]));
                    }
                    if (p.act && !acted) act_inplace(v, p.act);
                    if (p.gate) {
                        const __nv_bfloat16* g = p.gate + row * p.ldg + col0;
                        uint32_t gw[16];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const uint4 gi = ldg128(g + q * 8);
                            gw[q * 4 + 0] = gi.x; gw[q * 4 + 1] = gi.y; gw[q * 4 + 2] = gi.z; gw[q * 4 + 3] = gi.w;
                        }
                        gate_inplace(v, gw, p.gate_kind);
                    }
                    if (p.residual) {
                        const __nv_bfloat16* rs = p.residual + row * p.ldr + col0;
                        uint32_t rw[16];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const uint4 ri = ldg128(rs + q * 8);
                            rw[q * 4 + 0] = ri.x; rw[q * 4 + 1] = ri.y; rw[q * 4 + 2] = ri.z; rw[q * 4 + 3] = ri.w;
                        }
                        add_bf16_inplace(v, rw);
                    }
                    if (p.d_f32) {
                        float* dst = reinterpret_cast<float*>(p.d) + row * p.ldd + col0;
                        if (p.accumulate) {
#pragma unroll
                            for (int q = 0; q < 8; ++q)
                                red_add_f32x4(dst + q * 4, v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
                        } else {
#pragma unroll
                            for (int q = 0; q < 8; ++q)
                                *reinterpret_cast<float4*>(dst + q * 4) =
                                    make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
                        }
                    } else {
                        __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.d) + row * p.ldd + col0;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            uint4 o;
                            o.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]);
                            o.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
                            o.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]);
                            o.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
                            stg128(dst + q * 8, o);
                        }
                    }
                } else {
                    // ragged tail / unaligned leading dimension: scalar path
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int64_t col = col0 + j;
                        if (col >= p.N) continue;
                        float x = v[j];
                        if (p.preact) p.preact[row * p.ldp + col] = __float2bfloat16(p.act_grad ? apply_act_grad(x, p.act) : x);
                        if (p.act) x = apply_act(x, p.act);
                        if (p.gate) x *= apply_gate(__bfloat162float(p.gate[row * p.ldg + col]), p.gate_kind);
                        if (p.residual) x += __bfloat162float(p.residual[row * p.ldr + col]);
                        if (p.d_f32) {
                            float* dst = reinterpret_cast<float*>(p.d) + row * p.ldd + col;
                            if (p.accumulate) atomicAdd(dst, x);
                            else *dst = x;
                        } else {
                            reinterpret_cast<__nv_bfloat16*>(p.d)[row * p.ldd + col] = __float2bfloat16(x);
                        }
                    }
                }
                }   // row_ok && col0 < N
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ----------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------
static PerDeviceOnce g_attr_once;

static int pick_block_n(int64_t M, int64_t N, int b_mn, int num_sms) {
    // minimise (waves * block_n): quantisation-aware choice among the legal UMMA N sizes
    const int cands[4] = {256, 192, 128, 64};
    int best = 128;
    double best_cost = 1e30;
    const int64_t mt = ceil_div64(M, BLOCK_M);
    for (int i = 0; i < 4; ++i) {
        const int bn = cands[i];
        if (bn > 64 && bn - 64 >= N) continue;   // do not pad N by more than one 64-wide atom
        const int64_t nt = ceil_div64(N, bn);
        const int64_t waves = ceil_div64(mt * nt, num_sms);
        // per-tile cost ~ bn (MMA time) + fixed overhead; narrower tiles are less smem-efficient
        const double cost = (double)waves * (bn + 24.0);
        if (cost < best_cost) { best_cost = cost; best = bn; }
    }
    (void)b_mn;
    return best;
}

}  // namespace gemm
}  // namespace mgb

using namespace mgb;
using namespace mgb::gemm;

extern "C" int mgb_colsum_bf16(const void* x, int64_t ld, int64_t rows, int64_t cols, float* out, mgb_stream_t s);

extern "C" int mgb_gemm_bf16(const mgb_gemm_args* a, mgb_stream_t stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (!a || !a->a || !a->b || !a->d) { set_last_error("gemm: null pointer"); return kErrBadShape; }
    if (a->m <= 0 || a->n <= 0 || a->k <= 0) { set_last_error("gemm: non-positive dimension"); return kErrBadShape; }
    if (a->m > INT32_MAX - 256 || a->n > INT32_MAX - 256 || a->k > INT32_MAX - 256) {
        set_last_error("gemm: dimension too large"); return kErrTooLarge;
    }
    if ((a->lda & 7) || (a->ldb & 7) || (reinterpret_cast<uintptr_t>(a->a) & 15) ||
        (reinterpret_cast<uintptr_t>(a->b) & 15)) {
        set_last_error("gemm: A/B base must be 16-byte aligned and lda/ldb multiples of 8 elements (TMA)");
        return kErrBadAlign;
    }
    if (a->accumulate && !a->d_f32) { set_last_error("gemm: accumulate requires fp32 D"); return kErrBadDtype; }
    if (a->split_k > 1 && !a->accumulate) { set_last_error("gemm: split_k > 1 requires accumulate"); return kErrUnsupported; }
    {
        const int act_kind = a->act & ~MGB_ACT_STORE_GRAD;
        if (act_kind != MGB_ACT_NONE && act_kind != MGB_ACT_GELU_TANH && act_kind != MGB_ACT_GELU_ERF) {
            set_last_error("gemm: unknown activation"); return kErrUnsupported;
        }
        if ((a->act & MGB_ACT_STORE_GRAD) && (act_kind == MGB_ACT_NONE || !a->preact)) {
            set_last_error("gemm: MGB_ACT_STORE_GRAD needs an activation and a preact output"); return kErrUnsupported;
        }
        if (a->gate && a->gate_kind != MGB_ACT_GELU_TANH && a->gate_kind != MGB_ACT_GELU_ERF && a->gate_kind != MGB_GATE_MUL) {
            set_last_error("gemm: unknown gate_kind"); return kErrUnsupported;
        }
    }

    g_attr_once.run([] { cudaFuncSetAttribute(gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES); });
    const int num_sms = mgb::num_sms();
    // The epilogue (bias, activation, gate, residual, saved pre-activation, column sums) is applied ONCE per output
    // element. Split-K units each run the whole epilogue on their partial sum, so accumulate + an epilogue term
    // never splits: an automatic split is forced to 1, an explicit one is an error.
    const bool has_epilogue = a->bias || a->residual || a->gate || a->preact || a->colsum || a->rowdot_x ||
                              (a->act & ~MGB_ACT_STORE_GRAD) != MGB_ACT_NONE;
    if (has_epilogue && a->split_k > 1) {
        set_last_error("gemm: split_k > 1 cannot be combined with bias / act / gate / residual / preact / colsum");
        return kErrUnsupported;
    }

    Params p{};
    p.M = a->m; p.N = a->n; p.K = a->k;
    p.a_mn = a->a_mn_major ? 1 : 0;
    p.b_mn = a->b_mn_major ? 1 : 0;
    p.d = a->d; p.ldd = a->ldd; p.d_f32 = a->d_f32; p.accumulate = a->accumulate;
    p.bias = a->bias;
    p.residual = reinterpret_cast<const __nv_bfloat16*>(a->residual); p.ldr = a->ldr;
    p.act = a->act & ~MGB_ACT_STORE_GRAD;
    p.preact = reinterpret_cast<__nv_bfloat16*>(a->preact); p.ldp = a->ldp;
    p.act_grad = (a->act & MGB_ACT_STORE_GRAD) ? 1 : 0;
    p.gate = reinterpret_cast<const __nv_bfloat16*>(a->gate); p.ldg = a->ldg; p.gate_kind = a->gate_kind;
    p.alpha = a->alpha == 0.0f ? 1.0f : a->alpha;
    if (a->drop_p != 0.0f && a->drop_seed) {
        if (!(a->drop_p > 0.f && a->drop_p < 1.f) || (a->n & 15)) { set_last_error("gemm: dropout needs 0 <= drop_p < 1 and n % 16 == 0"); return kErrBadShape; }
        p.drop_thresh = dropout_thresh16(a->drop_p);
        p.drop_inv_keep = 65536.0f / (65536.0f - (float)p.drop_thresh);
        p.drop_seed = a->drop_seed;
        p.drop_site = a->drop_site;
    }

    // Two tile engines, chosen by shape: CTA pairs (cta_group::2, 256-row tiles, halves the shared-memory
    // traffic per MMA) whenever there are at least two 128-row tiles; the single-CTA kernel for short M.
    // block_n carries an optional engine request in its high bits (see mingpt_b200.h)
    const int engine = a->block_n & (MGB_GEMM_ENGINE_SINGLE | MGB_GEMM_ENGINE_PAIR);
    mgb_gemm_args a_local = *a;
    a_local.block_n = a->block_n & ~(MGB_GEMM_ENGINE_SINGLE | MGB_GEMM_ENGINE_PAIR);
    if (has_epilogue && a->accumulate) a_local.split_k = 1;
    const bool want_pair = engine == MGB_GEMM_ENGINE_PAIR ||
                           (engine == 0 && a->m > BLOCK_M && !g_opt_gemm_single.load(std::memory_order_relaxed));
    if (a->colsum && a->d_f32) { set_last_error("gemm: colsum needs a bf16 D"); return kErrUnsupported; }
    if (p.drop_thresh && !want_pair) { set_last_error("gemm: the dropout epilogue is a CTA-pair engine epilogue (m > 128)"); return kErrUnsupported; }
    if (a->rowdot_x && !want_pair) { set_last_error("gemm: rowdot is a CTA-pair engine epilogue (m > 128)"); return kErrUnsupported; }
    if (want_pair) {
        const int rc = launch_gemm_pair(&a_local, p, num_sms, stream);
        if (rc == kFusedColsum) return kOk;
        if (rc == kOk && a->colsum) return mgb_colsum_bf16(a->d, a->ldd, a->m, a->n, a->colsum, stream_);
        return rc;
    }

    int bn = a_local.block_n;
    if (bn == 0) bn = (a->accumulate && a->n > 128) ? 256 : pick_block_n(a->m, a->n, p.b_mn, num_sms);
    if (bn != 64 && bn != 128 && bn != 192 && bn != 256) { set_last_error("gemm: block_n must be 64/128/192/256"); return kErrBadShape; }
    p.block_n = bn;
    p.num_m_tiles = (int)ceil_div64(a->m, BLOCK_M);
    p.num_n_tiles = (int)ceil_div64(a->n, bn);
    p.k_blocks_total = (int)ceil_div64(a->k, BLOCK_K);
    int split = a_local.split_k;
    const int tiles = p.num_m_tiles * p.num_n_tiles;
    if (split <= 0) {
        split = 1;
        if (a->accumulate) {
            // fill the machine: choose the split whose unit count wastes the least of the last wave,
            // keeping at least 8 k-blocks per unit
            const int max_split = p.k_blocks_total / 8 > 0 ? p.k_blocks_total / 8 : 1;
            double best = 1e30;
            for (int s = 1; s <= max_split && s <= 64; ++s) {
                const int64_t units = (int64_t)tiles * s;
                const int64_t waves = ceil_div64(units, num_sms);
                const int kbs = (int)ceil_div64(p.k_blocks_total, s);
                const double cost = (double)waves * (kbs + 6.0);
                if (cost < best - 1e-9) { best = cost; split = s; }
            }
        }
    }
    if (split > p.k_blocks_total) split = p.k_blocks_total;
    p.k_blocks_per_split = (int)ceil_div64(p.k_blocks_total, split);
    p.split_k = (int)ceil_div64(p.k_blocks_total, p.k_blocks_per_split);   // no empty splits
    CUtensorMap ta, tb;
    int rc;
    if (!p.a_mn) rc = make_map_2d(&ta, a->a, a->k, a->m, a->lda, BLOCK_K, BLOCK_M);
    else         rc = make_map_2d(&ta, a->a, a->m, a->k, a->lda, 64, BLOCK_K);
    if (rc) return rc;
    if (!p.b_mn) rc = make_map_2d(&tb, a->b, a->k, a->n, a->ldb, BLOCK_K, bn);
    else         rc = make_map_2d(&tb, a->b, a->n, a->k, a->ldb, 64, BLOCK_K);
    if (rc) return rc;

    const int64_t units = (int64_t)tiles * p.split_k;
    const int grid = (int)(units < num_sms ? units : num_sms);
    gemm_kernel<<<grid, kThreads, SMEM_BYTES, stream>>>(ta, tb, p);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    if (a->colsum) return mgb_colsum_bf16(a->d, a->ldd, a->m, a->n, a->colsum, stream_);   // no fused variant on this engine
    return kOk;
}
