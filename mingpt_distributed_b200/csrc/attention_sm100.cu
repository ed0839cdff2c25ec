// Fused attention forward on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), head size 64 --
// the head size of every gpt2 preset (gpt2 / medium / large / xl). Other head sizes (16, 32: gpt-nano,
// gpt-mini) stay on the mma.sync kernels in attention.cu.
//
// One CTA = 128 queries of one (batch, head); keys/values stream through in blocks of 128.
//   S = Q K^T   : tcgen05.mma 128x128x64 (Q, K K-major, SWIZZLE_128B), fp32 S in TMEM columns [0,128)
//   softmax     : warps 0-3, ONE THREAD PER QUERY ROW (TMEM lane == row): row max / sum need no shuffles;
//                 online softmax in the log2 domain; P (bf16 pairs) is written back to TMEM columns [128,192)
//   O += P V    : tcgen05.mma 128x64x128 with the A operand (P) read FROM TMEM and V as the MN-major B
//                 operand straight from the TMA tile; fp32 O in TMEM columns [192,256); O is only rescaled
//                 when the running max grows by more than 2^8
// warp 4 = TMA producer, warp 5 = MMA issuer + TMEM allocator. 48 KiB of shared memory and 256 TMEM columns
// per CTA, so two CTAs share an SM: one does softmax while the other's MMAs run.
#include "gemm_common.cuh"
#include "dropout.cuh"

namespace mgb {
extern std::atomic<int> g_opt_attn_poly;   // optim.cu: share of the forward softmax's exponentials on the FMA pipe (0, 1 = 25 %, 2 = 50 %)
namespace attn_tc {

constexpr int BQ = 128, BKV = 128, D = 64;
constexpr int kThreads = 192;
constexpr int TILE_BYTES = 128 * 128;            // 128 rows x 64 bf16
constexpr int SMEM_BYTES = 3 * TILE_BYTES + 1024 + 256;   // Q, K, V + align slack + barriers
constexpr uint32_t kTmemCols = 256;
constexpr uint32_t kColS = 0, kColP = 128, kColO = 192;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr float kRescaleThreshold = 8.0f;         // log2 units: P stays below 2^8 between rescales

__device__ __forceinline__ void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t (&r)[16]);
__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
          "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
          "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}

// Attention-probability dropout on packed bf16 pairs: words w[0 .. NW) hold 2*NW consecutive keys of one query
// row starting at a multiple of 8; `group` is the mask group of the first 8 (dropout.cuh: group = (bh*T + q) *
// ceil(T/8) + key/8, bit = key % 8 -- the definition the mma.sync kernels and mgb_attention_dropout_mask use).
// Dropped probabilities become exact zeros; the 1/(1-p) factor is applied once, to the output row.
template <int NW>
__device__ __forceinline__ void drop_packed(uint32_t* w, uint64_t seed, uint32_t site, uint64_t group, uint32_t thresh) {
#pragma unroll
    for (int g8 = 0; g8 < NW / 4; ++g8) {
        const uint32_t m = dropout_keep8(seed, site, group + g8, thresh);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t keep = ((m >> (2 * i)) & 1u ? 0x0000ffffu : 0u) | ((m >> (2 * i + 1)) & 1u ? 0xffff0000u : 0u);
            w[g8 * 4 + i] &= keep;
        }
    }
}

// The same on a pre-generated keep word (dropout.cuh): 16 packed pairs = 32 consecutive keys of this query row.
__device__ __forceinline__ void drop_packed_word(uint32_t* w, uint32_t word) {
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        uint32_t m_s, m_s8;
        dropout_pair_masks(word, s, m_s, m_s8);
        w[s] &= m_s;
        w[s + 8] &= m_s8;
    }
}

// DROP: attention-probability dropout (nn.MultiheadAttention dropout=attn_drop: mingpt/model.py:150 ->
// torch/nn/functional.py:6644-6645). The softmax normaliser keeps every key, P V only the kept ones.
template <bool DROP, int POLY>
__global__ void __launch_bounds__(kThreads, 2)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmap_qkv, __nv_bfloat16* __restrict__ out,
                   float* __restrict__ lse, int T, int H, int mask_mode, float scale_log2, uint32_t drop_thresh,
                   float inv_keep, const uint64_t* __restrict__ seed_dev, uint32_t site,
                   const uint32_t* __restrict__ drop_bits) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = smem + TILE_BYTES;
    uint8_t* sV = smem + 2 * TILE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 3 * TILE_BYTES);
    uint64_t* q_full = bars + 0;
    uint64_t* k_full = bars + 1;
    uint64_t* k_empty = bars + 2;
    uint64_t* v_full = bars + 3;
    uint64_t* v_empty = bars + 4;
    uint64_t* s_full = bars + 5;
    uint64_t* p_ready = bars + 6;
    uint64_t* pv_done = bars + 7;
    uint32_t* tmem_base_ptr = reinterpret_cast<uint32_t*>(bars + 8);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // provably warp-uniform
    const int b = blockIdx.y / H, h = blockIdx.y % H;
    const int qb = gridDim.x - 1 - blockIdx.x;       // heavy (late) causal blocks first
    const int q0 = qb * BQ;
    const int C = H * D;
    const bool causal = mask_mode == MGB_MASK_CAUSAL;
    const bool tril = mask_mode == MGB_MASK_ADD_TRIL_ONES;
    const int nkv_all = (T + BKV - 1) / BKV;
    const int nkv = causal ? min(qb + 1, nkv_all) : nkv_all;
    const int row0 = b * T;                          // first row of this sequence in the [B*T, 3C] view

    if (warp == 4 && lane == 0) {
        tma_prefetch_desc(&tmap_qkv);
        mbar_init(q_full, 1); mbar_init(k_full, 1); mbar_init(k_empty, 1); mbar_init(v_full, 1); mbar_init(v_empty, 1);
        mbar_init(s_full, 1); mbar_init(p_ready, 4); mbar_init(pv_done, 1);
        fence_mbar_init();
    }
    if (warp == 5) tmem_alloc(tmem_base_ptr, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_ptr;
    grid_dep_launch();
    grid_dep_wait();                                 // qkv is the previous kernel's output (programmatic dependent launch)

    if (warp == 4) {
        // ===================== TMA producer (whole warp runs the loop, one lane issues) =====================
        const bool issuer = elect_one();
        if (issuer) {
            mbar_arrive_expect_tx(q_full, TILE_BYTES);
            tma_load_2d(sQ, &tmap_qkv, q_full, h * D, row0 + q0);
        }
        __syncwarp();
        for (int j = 0; j < nkv; ++j) {
            const uint32_t ph = (uint32_t)(j & 1);
            mbar_wait_relaxed(k_empty, ph ^ 1);               // S_{j-1} has consumed the previous K
            if (issuer) {
                mbar_arrive_expect_tx(k_full, TILE_BYTES);
                tma_load_2d(sK, &tmap_qkv, k_full, C + h * D, row0 + j * BKV);
            }
            __syncwarp();
            mbar_wait_relaxed(v_empty, ph ^ 1);               // P V_{j-1} has consumed the previous V
            if (issuer) {
                mbar_arrive_expect_tx(v_full, TILE_BYTES);
                tma_load_2d(sV, &tmap_qkv, v_full, 2 * C + h * D, row0 + j * BKV);
            }
            __syncwarp();
        }
    } else if (warp == 5) {
        // ===================== MMA issuer (whole warp runs the loop, one lane issues) =====================
        const bool issuer = elect_one();
        const uint32_t idesc_s = make_idesc_bf16(BQ, BKV, 0, 0);   // S: 128 x 128, both K-major
        const uint32_t idesc_o = make_idesc_bf16(BQ, D, 0, 1);     // O: 128 x 64, V MN-major
        const uint64_t dQ = make_smem_desc_sw128(smem_u32(sQ), 0, 1024);
        const uint64_t dK = make_smem_desc_sw128(smem_u32(sK), 0, 1024);
        const uint64_t dV = make_smem_desc_sw128(smem_u32(sV), TILE_BYTES, 1024);   // MN-major: +2048 B per 16 keys
        auto issue_s = [&]() {
            if (issuer) {
#pragma unroll
                for (int k = 0; k < D / 16; ++k)
                    umma_bf16_ss(tmem_base + kColS, dQ + 2 * k, dK + 2 * k, idesc_s, k > 0 ? 1u : 0u);
                umma_commit(k_empty);
                umma_commit(s_full);
            }
            __syncwarp();
        };
        mbar_wait_relaxed(q_full, 0);
        mbar_wait_relaxed(k_full, 0);
        tc_fence_after();
        issue_s();
        for (int j = 0; j < nkv; ++j) {
            const uint32_t ph = (uint32_t)(j & 1);
            mbar_wait_relaxed(p_ready, ph);               // P_j in TMEM, O rescaled if needed, S consumed
            tc_fence_after();
            if (j + 1 < nkv) {                    // S_{j+1} first: the softmax warps start on it while P V_j runs
                mbar_wait_relaxed(k_full, ph ^ 1);
                tc_fence_after();
                issue_s();
            }
            mbar_wait_relaxed(v_full, ph);
            tc_fence_after();
            if (issuer) {
#pragma unroll
                for (int kk = 0; kk < BKV / 16; ++kk)   // A = P from TMEM: 16 bf16 = 8 packed columns per step
                    umma_bf16_ts(tmem_base + kColO, tmem_base + kColP + kk * 8, dV + 128 * kk, idesc_o,
                                 (j > 0 || kk > 0) ? 1u : 0u);
                umma_commit(v_empty);
                umma_commit(pv_done);
            }
            __syncwarp();
        }
    } else {
        // ===================== softmax: one thread per query row =====================
        const int r = warp * 32 + lane;              // row inside the tile == TMEM lane
        const int q = q0 + r;
        const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16);
        float m_run = -INFINITY, l_run = 0.f;
        const uint64_t drop_seed = (DROP && !drop_bits) ? *seed_dev : 0;
        const uint64_t drop_row = ((uint64_t)blockIdx.y * T + (uint64_t)q) * (uint64_t)((T + 7) >> 3);   // first group of this query
        const int drop_kw = (T + 31) >> 5;                       // keep words per query (pre-generated masks)
        const uint64_t drop_tp = (uint64_t)((T + 15) & ~15);
        // pre-generated masks: this row's keep words for a block's 4 x 32 keys, fetched one block ahead
        uint32_t dwn[4] = {~0u, ~0u, ~0u, ~0u};
        auto fetch_dw = [&](int jj) {
            if (q < T && jj < nkv) {
                const uint32_t* bp = drop_bits + ((uint64_t)blockIdx.y * drop_kw + 4 * jj) * drop_tp + q;
#pragma unroll
                for (int g = 0; g < 4; ++g)
                    if (4 * jj + g < drop_kw) dwn[g] = __ldg(bp + g * drop_tp);
            }
        };
        if (DROP && drop_bits) fetch_dw(0);
        for (int j = 0; j < nkv; ++j) {
            const uint32_t dw[4] = {dwn[0], dwn[1], dwn[2], dwn[3]};
            if (DROP && drop_bits) fetch_dw(j + 1);
            const uint32_t ph = (uint32_t)(j & 1);
            const int kvalid = min(BKV, T - j * BKV);            // keys of this block that exist
            const bool diag = (j == qb);
            const bool need_mask = (kvalid < BKV) || (causal && diag);
            const float bias_all = (tril && j < qb) ? LOG2E : 0.f;
            const bool bias_diag = tril && diag;
            const bool fast = !need_mask && !bias_diag;
            // causal diagonal block with every key present (and no additive bias): handled per 32-column group
            const bool diag_groups = causal && diag && (kvalid == BKV) && !bias_diag;
            mbar_wait_relaxed(s_full, ph);
            tc_fence_after();
            float mx = -INFINITY;
            if (diag_groups) {
                // causal diagonal block, all keys present: this warp's rows are 32w..32w+31, so 32-column groups left
                // of the diagonal group need no mask, the diagonal group a per-element one, and the groups right of
                // it are entirely masked (no TMEM read, no exponentials). All of it warp-uniform.
                float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll 1
                for (int g = 0; g <= warp; ++g) {
                    uint32_t sv[32];
                    tmem_ld_32x32b_x32(taddr + kColS + g * 32, sv);
                    tmem_ld_wait();
                    if (g < warp) {
#pragma unroll
                        for (int e = 0; e < 32; e += 8) {
                            m0 = fmaxf(m0, fmaxf(__uint_as_float(sv[e]), __uint_as_float(sv[e + 1])));
                            m1 = fmaxf(m1, fmaxf(__uint_as_float(sv[e + 2]), __uint_as_float(sv[e + 3])));
                            m2 = fmaxf(m2, fmaxf(__uint_as_float(sv[e + 4]), __uint_as_float(sv[e + 5])));
                            m3 = fmaxf(m3, fmaxf(__uint_as_float(sv[e + 6]), __uint_as_float(sv[e + 7])));
                        }
                    } else {
#pragma unroll
                        for (int e = 0; e < 32; ++e)
                            if (e <= lane) m0 = fmaxf(m0, __uint_as_float(sv[e]));
                    }
                }
                mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)) * scale_log2;
            } else {
                // ---- pass 1: row maximum of this block (log2 domain); 64 columns per TMEM round trip, four
                // independent running maxima (one chain would serialise 64 dependent instructions) ----
                float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll 1
                for (int c = 0; c < 2; ++c) {
                    uint32_t s0[32], s1[32];
                    tmem_ld_32x32b_x32(taddr + kColS + c * 64, s0);
                    tmem_ld_32x32b_x32(taddr + kColS + c * 64 + 32, s1);
                    tmem_ld_wait();
                    if (fast) {
#pragma unroll
                        for (int e = 0; e < 32; e += 4) {
                            m0 = fmaxf(m0, fmaxf(__uint_as_float(s0[e]), __uint_as_float(s1[e])));
                            m1 = fmaxf(m1, fmaxf(__uint_as_float(s0[e + 1]), __uint_as_float(s1[e + 1])));
                            m2 = fmaxf(m2, fmaxf(__uint_as_float(s0[e + 2]), __uint_as_float(s1[e + 2])));
                            m3 = fmaxf(m3, fmaxf(__uint_as_float(s0[e + 3]), __uint_as_float(s1[e + 3])));
                        }
                    } else {
#pragma unroll
                        for (int e = 0; e < 64; ++e) {
                            const int col = c * 64 + e;
                            float x = fmaf(__uint_as_float(e < 32 ? s0[e & 31] : s1[e & 31]), scale_log2, bias_all);
                            if (bias_diag && col <= r) x += LOG2E;
                            if (col >= kvalid || (causal && diag && col > r)) x = -INFINITY;
                            m0 = fmaxf(m0, x);
                        }
                    }
                }
                mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
                if (fast) mx = fmaf(mx, scale_log2, bias_all);
            }
            // ---- running max with lazy rescale ----
            float factor = 1.0f;
            const bool grow = mx > m_run + kRescaleThreshold;    // also true on the first block (m_run = -inf)
            if (grow) {
                factor = exp2f(m_run - mx);                      // 0 on the first block
                m_run = mx;
                l_run *= factor;
            }
            if (j > 0) {
                mbar_wait_relaxed(pv_done, ph ^ 1);                      // P V_{j-1} finished: P columns free, O stable
                tc_fence_after();
                if (__any_sync(0xffffffffu, grow)) {
#pragma unroll 1
                    for (int c = 0; c < 2; ++c) {
                        uint32_t ov[32];
                        tmem_ld_32x32b_x32(taddr + kColO + c * 32, ov);
                        tmem_ld_wait();
#pragma unroll
                        for (int e = 0; e < 32; ++e) ov[e] = __float_as_uint(__uint_as_float(ov[e]) * factor);
                        tmem_st_32x32b_x32(taddr + kColO + c * 32, ov);
                    }
                }
            }
            // ---- pass 2: P = exp2(x - m_run), row sum, bf16 pairs back into TMEM ----
            // Everything on fp32 PAIRS (FFMA2 / FADD2: one issue slot for two columns); POLY of every 4 pairs take
            // their exponential on the FMA pipe (exp2_poly2) instead of the special-function unit.
            const float off = bias_all - m_run;
            const float2 sc2 = splat2(scale_log2), off2 = splat2(off);
            float2 acc_a = make_float2(0.f, 0.f), acc_b = make_float2(0.f, 0.f);
            float rs = 0.f;
            if (diag_groups) {
#pragma unroll 1
                for (int g = 0; g < 4; ++g) {
                    uint32_t pk[16];
                    if (g > warp) {
#pragma unroll
                        for (int e = 0; e < 16; ++e) pk[e] = 0u;
                    } else {
                        uint32_t sv[32];
                        tmem_ld_32x32b_x32(taddr + kColS + g * 32, sv);
                        tmem_ld_wait();
                        if (g < warp) {
#pragma unroll
                            for (int e = 0; e < 16; e += 2) {
                                const float2 xa = fma2(make_float2(__uint_as_float(sv[2 * e]), __uint_as_float(sv[2 * e + 1])), sc2, off2);
                                const float2 xb = fma2(make_float2(__uint_as_float(sv[2 * e + 2]), __uint_as_float(sv[2 * e + 3])), sc2, off2);
                                const float2 pa = make_float2(exp2f(xa.x), exp2f(xa.y));
                                const float2 pb = (POLY == 2 || (POLY == 1 && (e & 2))) ? exp2_poly2(xb) : make_float2(exp2f(xb.x), exp2f(xb.y));
                                acc_a = add2(acc_a, pa);
                                acc_b = add2(acc_b, pb);
                                pk[e] = pack_bf16x2(pa.x, pa.y);
                                pk[e + 1] = pack_bf16x2(pb.x, pb.y);
                            }
                        } else {
#pragma unroll
                            for (int e = 0; e < 16; ++e) {
                                float a0 = exp2f(fmaf(__uint_as_float(sv[2 * e]), scale_log2, off));
                                float a1 = exp2f(fmaf(__uint_as_float(sv[2 * e + 1]), scale_log2, off));
                                if (2 * e > lane) a0 = 0.f;
                                if (2 * e + 1 > lane) a1 = 0.f;
                                rs += a0 + a1;
                                pk[e] = pack_bf16x2(a0, a1);
                            }
                        }
                    }
                    if (DROP && g <= warp) {
                        if (drop_bits) drop_packed_word(pk, g == 0 ? dw[0] : (g == 1 ? dw[1] : (g == 2 ? dw[2] : dw[3])));
                        else drop_packed<16>(pk, drop_seed, site, drop_row + (uint64_t)(j * (BKV / 8) + g * 4), drop_thresh);
                    }
                    tmem_st_32x32b_x16(taddr + kColP + g * 16, pk);
                }
            } else {
#pragma unroll 1
                for (int c = 0; c < 2; ++c) {
                    uint32_t s0[32], s1[32];
                    tmem_ld_32x32b_x32(taddr + kColS + c * 64, s0);
                    tmem_ld_32x32b_x32(taddr + kColS + c * 64 + 32, s1);
                    tmem_ld_wait();
                    uint32_t pk[32];
                    if (fast) {
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            const float2 xa = fma2(make_float2(__uint_as_float(s0[2 * e]), __uint_as_float(s0[2 * e + 1])), sc2, off2);
                            const float2 xb = fma2(make_float2(__uint_as_float(s1[2 * e]), __uint_as_float(s1[2 * e + 1])), sc2, off2);
                            const float2 pa = make_float2(exp2f(xa.x), exp2f(xa.y));
                            const float2 pb = (POLY == 2 || (POLY == 1 && (e & 1))) ? exp2_poly2(xb) : make_float2(exp2f(xb.x), exp2f(xb.y));
                            acc_a = add2(acc_a, pa);
                            acc_b = add2(acc_b, pb);
                            pk[e] = pack_bf16x2(pa.x, pa.y);
                            pk[16 + e] = pack_bf16x2(pb.x, pb.y);
                        }
                    } else {
#pragma unroll
                        for (int e = 0; e < 32; ++e) {
                            float pv[2];
#pragma unroll
                            for (int t2 = 0; t2 < 2; ++t2) {
                                const int idx = 2 * e + t2;                       // 0..63 inside this 64-column group
                                const int col = c * 64 + idx;
                                float x = fmaf(__uint_as_float(idx < 32 ? s0[idx & 31] : s1[idx & 31]), scale_log2, off);
                                if (bias_diag && col <= r) x += LOG2E;
                                const bool dead = col >= kvalid || (causal && diag && col > r);
                                pv[t2] = dead ? 0.f : exp2f(x);
                            }
                            rs += pv[0] + pv[1];
                            pk[e] = pack_bf16x2(pv[0], pv[1]);
                        }
                    }
                    if (DROP) {
                        if (drop_bits) {
                            drop_packed_word(pk, c ? dw[2] : dw[0]);
                            drop_packed_word(pk + 16, c ? dw[3] : dw[1]);
                        } else {
                            drop_packed<32>(pk, drop_seed, site, drop_row + (uint64_t)(j * (BKV / 8) + c * 8), drop_thresh);
                        }
                    }
                    tmem_st_32x32b_x32(taddr + kColP + c * 32, pk);
                }
            }
            rs += (acc_a.x + acc_a.y) + (acc_b.x + acc_b.y);
            l_run += rs;
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_ready);
        }
        // ---- epilogue: O / l, log-sum-exp ----
        mbar_wait_relaxed(pv_done, (uint32_t)((nkv - 1) & 1));
        tc_fence_after();
        const float inv = l_run > 0.f ? (DROP ? inv_keep : 1.0f) / l_run : 0.f;
        __nv_bfloat16* orow = out + ((int64_t)(row0 + q)) * C + (int64_t)h * D;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
            uint32_t ov[32];
            __syncwarp();
            tmem_ld_32x32b_x32(taddr + kColO + c * 32, ov);
            tmem_ld_wait();
            if (q < T) {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    stg128(orow + c * 32 + u * 8,
                           make_uint4(pack_bf16x2(__uint_as_float(ov[u * 8 + 0]) * inv, __uint_as_float(ov[u * 8 + 1]) * inv),
                                      pack_bf16x2(__uint_as_float(ov[u * 8 + 2]) * inv, __uint_as_float(ov[u * 8 + 3]) * inv),
                                      pack_bf16x2(__uint_as_float(ov[u * 8 + 4]) * inv, __uint_as_float(ov[u * 8 + 5]) * inv),
                                      pack_bf16x2(__uint_as_float(ov[u * 8 + 6]) * inv, __uint_as_float(ov[u * 8 + 7]) * inv)));
            }
        }
        if (q < T) lse[((int64_t)b * H + h) * T + q] = (m_run + log2f(l_run)) * LN2;
    }

    __syncwarp();
    tc_fence_before();
    __syncthreads();
    if (warp == 5) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

int launch_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int mask_mode, float scale,
               cudaStream_t stream, uint32_t drop_thresh, float inv_keep, const uint64_t* seed_dev, uint32_t site,
               const uint32_t* drop_bits) {
    static PerDeviceOnce once;
    once.run([] {
        cudaFuncSetAttribute(attn_fwd_tc_kernel<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        cudaFuncSetAttribute(attn_fwd_tc_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        cudaFuncSetAttribute(attn_fwd_tc_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        cudaFuncSetAttribute(attn_fwd_tc_kernel<true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        cudaFuncSetAttribute(attn_fwd_tc_kernel<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        cudaFuncSetAttribute(attn_fwd_tc_kernel<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    });
    CUtensorMap tm;
    const int64_t C3 = (int64_t)3 * H * D;
    int rc = gemm::make_map_2d(&tm, qkv, C3, (int64_t)B * T, C3, 64, 128);
    if (rc) return rc;
    dim3 grid((T + BQ - 1) / BQ, B * H);
    const int poly = g_opt_attn_poly.load(std::memory_order_relaxed);
#define MGB_ATTN_FWD(DROPV, POLYV)                                                                                          \
    launch_pdl(attn_fwd_tc_kernel<DROPV, POLYV>, grid, dim3(kThreads), (size_t)SMEM_BYTES, stream, tm, (__nv_bfloat16*)out, lse, T, H, \
               mask_mode, scale * LOG2E, drop_thresh, inv_keep, seed_dev, site, drop_bits)
    if (drop_thresh) {
        if (poly == 2) MGB_ATTN_FWD(true, 2); else if (poly == 1) MGB_ATTN_FWD(true, 1); else MGB_ATTN_FWD(true, 0);
    } else {
        inv_keep = 1.0f; seed_dev = nullptr; site = 0u; drop_bits = nullptr;
        if (poly == 2) MGB_ATTN_FWD(false, 2); else if (poly == 1) MGB_ATTN_FWD(false, 1); else MGB_ATTN_FWD(false, 0);
    }
#undef MGB_ATTN_FWD
    return kOk;
}

// =====================================================================================================
// backward (recompute), head size 64
// =====================================================================================================
// One CTA = one block of 128 keys of one (batch, head); it walks over the query blocks that see those keys
// in SUB-BLOCKS of 64 queries. Everything is computed TRANSPOSED (rows = keys) so that a thread owns one key
// row in TMEM and the recomputed P^T / dS^T are directly the A operands of the dV / dK products:
//   S^T  = K Q^T          tcgen05.mma 128x64x64    TMEM cols [  0,128): two buffers of 64
//   dP^T = V dO^T         tcgen05.mma 128x64x64    TMEM cols [128,256): two buffers of 64
//                         K and V are this item's CONSTANT A operands: the elementwise warps copy them once per
//                         item into TMEM (bf16, cols [256,288) and [288,320)), so these products read only the
//                         64x64 Q / dO sub-tile from shared memory (an A operand in TMEM costs 42 cycles per
//                         128x64x16 instruction, one in shared memory 75: tools/micro/mma_rate.cu)
//   P^T  = exp2(S^T*c - lse_q), dS^T = P^T (dP^T - delta_q) * scale     (16 warps, thread = key row, 16 queries each)
//                         written back IN PLACE as bf16 pairs over the thread's own fp32 S^T / dP^T columns
//   dV  += P^T dO         A = P^T from TMEM, B = dO sub-tile, MN-major;                                cols [384,448)
//   dK  += dS^T Q         A = dS^T from TMEM, B = Q sub-tile, MN-major;                                cols [320,384)
//   dQ_i = dS K           once per PAIR of sub-blocks (128 queries): A = a shared-memory copy of dS^T read as
//                         an MN-major operand, B = K tile (MN-major); cols [448,512); four drain warps move it
//                         TMEM -> registers -> swizzled box -> TMA fp32 reduce-add into dq_accum
// Pipeline: the MMA thread issues S^T/dP^T of sub-block h+1 BEFORE it waits for the elementwise warps to
// finish sub-block h, so the tensor pipe works on h+1 (and on dV/dK/dQ of h-1) while the warps do the
// exponentials of h. tcgen05.mma instructions of one thread execute in order, which is what makes the
// double buffers safe: S^T / dP^T of h+2 are issued after dV / dK of h, the readers of the bf16 P^T / dS^T that
// live in the same columns, and "S^T of h+4 is complete" implies dQ of the block that read the shared-memory
// dS^T atom written four sub-blocks ago is complete.
constexpr int SUBQ = 64;
constexpr int kEwWarps = 16;                      // elementwise warps: 4 per TMEM lane quarter
constexpr int kEwGroups = kEwWarps / 4;           // warpgroups; each takes SUBQ / kEwGroups query columns
constexpr int CPT = SUBQ / kEwGroups;             // columns per elementwise thread (16)
static_assert(CPT == 16, "the K/V copy and the in-place bf16 write-back assume 16 columns per thread");
constexpr int kDrainWarp = kEwWarps, kTmaWarp = kEwWarps + 4, kMmaWarp = kEwWarps + 5;
constexpr int kBwdThreads = (kEwWarps + 6) * 32;  // elementwise | 4 dQ drain | TMA | MMA
constexpr int SUB_BYTES = SUBQ * 128;             // 64 rows x 64 bf16
constexpr int kSlots = 4;                         // Q/dO sub-tiles in flight
constexpr uint32_t kBwdTmemCols = 512;
constexpr uint32_t cST = 0, cDPT = 128, cKT = 256, cVT = 288, cDK = 320, cDV = 384, cDQ = 448;
// K, V | 4 x (Q, dO) sub-tiles | dS^T: 4 atoms of 64 queries | dQ staging: 2 boxes | lse/delta x2 | barriers
// 2 x (K, V) | 4 x (Q, dO) sub-tiles | dS^T: 4 atoms | dQ staging: 2 boxes | lse/delta x2 | barriers. No alignment slack:
// the kernel has no static shared memory, so the dynamic window starts 1024-aligned (checked at kernel entry).
constexpr int BWD_SMEM = 4 * TILE_BYTES + kSlots * 2 * SUB_BYTES + 4 * TILE_BYTES + 2 * TILE_BYTES + 2 * 128 * 8 + 256;

__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_st_32x32b_x8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
        : "memory");
}
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t (&r)[32]) { tmem_ld_32x32b_x32(taddr, r); }
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t (&r)[16]) { tmem_ld_32x32b_x16(taddr, r); }
__device__ __forceinline__ void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t (&r)[16]);
__device__ __forceinline__ void tmem_st_cols(uint32_t taddr, const uint32_t (&r)[16]) { tmem_st_32x32b_x16(taddr, r); }
__device__ __forceinline__ void tmem_st_cols(uint32_t taddr, const uint32_t (&r)[8]) { tmem_st_32x32b_x8(taddr, r); }
__device__ __forceinline__ void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}


// Work items (one key block of one (batch, head)) are dealt to a PERSISTENT grid, so TMEM allocation, barrier
// set-up and the prologue / epilogue bubbles are paid once per SM instead of once per item, and the TMA
// producer already fetches the next item's Q / dO while the current item drains. The deal is in UNITS of
// equal work whose items share a (batch, head): causal => the pair (key block p, key block nq-1-p), which
// together see nq+1 query blocks; otherwise a single key block. Consecutive units are neighbouring key
// blocks of the same (batch, head), so the CTAs running at any moment share their Q / dO tiles and their
// dQ accumulator lines in L2.
struct BwdItem {
    int b, h, kb, i_begin, nblk;
};
struct BwdSched {
    int H, nq, causal, units_per_bh, nunits, items_per_unit;
    uint32_t m_upb, m_h;       // ceil(2^32 / units_per_bh), ceil(2^32 / H): x / d == umulhi(x, m) for the unit counts here
};
__device__ __forceinline__ BwdSched bwd_sched(int B, int H, int nq, bool causal) {
    BwdSched sc;
    sc.H = H; sc.nq = nq; sc.causal = causal;
    sc.units_per_bh = causal ? (nq + 1) / 2 : nq;
    sc.nunits = sc.units_per_bh * B * H;
    sc.items_per_unit = causal ? 2 : 1;
    // every role decodes every item of its CTA, at the start of the item: three integer divisions there were 8 % of the
    // kernel's warp samples (profiles/r2b_attn_bwd_ncu.md). Exact for x * (m * d - 2^32) < 2^32, i.e. x < 2^32 / d.
    sc.m_upb = (uint32_t)((0x100000000ull + (uint32_t)sc.units_per_bh - 1) / (uint32_t)sc.units_per_bh);
    sc.m_h = (uint32_t)((0x100000000ull + (uint32_t)H - 1) / (uint32_t)H);
    return sc;
}
// q-th item slot of this CTA: -1 past the end, 0 empty slot (the middle key block of an odd count pairs with
// itself), 1 valid. Every role walks the same sequence.
__device__ __forceinline__ int bwd_item(const BwdSched& sc, int q, BwdItem& it) {
    const int u = (int)blockIdx.x + (sc.causal ? (q >> 1) : q) * (int)gridDim.x;
    if (u >= sc.nunits) return -1;
    const int second = sc.causal ? (q & 1) : 0;
    const int bh = sc.units_per_bh == 1 ? u : (int)__umulhi((uint32_t)u, sc.m_upb);
    const int p = u - bh * sc.units_per_bh;
    it.kb = second ? sc.nq - 1 - p : p;
    if (second && it.kb == p) return 0;
    it.b = sc.H == 1 ? bh : (int)__umulhi((uint32_t)bh, sc.m_h);
    it.h = bh - it.b * sc.H;
    it.i_begin = sc.causal ? it.kb : 0;
    it.nblk = sc.nq - it.i_begin;
    return 1;
}

#ifndef MGB_ATTN_BWD_SPLIT_LD
#define MGB_ATTN_BWD_SPLIT_LD 1
#endif
// DROP: attention-probability dropout with the forward's mask. A thread owns ONE key and CPT consecutive queries:
//   drop_bits given -- the TRANSPOSED keep words the generator wrote next to the forward's (dropout.cuh; one word =
//       32 queries of one key): one coalesced 4-byte load per thread and sub-block, 2 instructions per element;
//   else -- the mask is regenerated: a group is 8 consecutive KEYS of one query, so the 32 lanes of a warp (32 keys =
//       4 groups) share the 4 x CPT = 64 groups of a sub-block, two Philox calls per lane, one shuffle per query.
// dS = P (ks dP - delta scale) and dV += (ks P)^T dO / scale with ks = keep ? scale / (1-p) : 0.
template <bool DROP>
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmap_kv, const __grid_constant__ CUtensorMap tmap_q,
                   const __grid_constant__ CUtensorMap tmap_do, const __grid_constant__ CUtensorMap tmap_dq,
                   const float* __restrict__ lse, const float* __restrict__ delta,
                   __nv_bfloat16* __restrict__ dqkv, float* __restrict__ kv_colsum, int B, int T, int H, int mask_mode,
                   float scale, float scale_log2, uint32_t drop_thresh, float inv_keep,
                   const uint64_t* __restrict__ seed_dev, uint32_t site, const uint32_t* __restrict__ drop_bits) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw;
    if (smem_u32(smem) & 1023u) __trap();                      // SWIZZLE_128B tiles need 1024-byte alignment
    uint8_t* sKV = smem;                                       // item w: K at (w&1)*2 tiles, V one tile behind it
    uint8_t* sQ = smem + 4 * TILE_BYTES;                       // slot s: Q at +s*2*SUB_BYTES, dO right behind it
    uint8_t* sdS = sQ + kSlots * 2 * SUB_BYTES;                // atom a: 128 key rows x 128 B (64 queries)
    uint8_t* sDQ = sdS + 4 * TILE_BYTES;                       // two boxes: 128 query rows x 32 fp32
    float2* sLD = reinterpret_cast<float2*>(sDQ + 2 * TILE_BYTES);   // [16 warps][16] (bias - lse*log2e, delta*scale)
    uint64_t* bars = reinterpret_cast<uint64_t*>(sDQ + 2 * TILE_BYTES + 2 * 128 * 8);
    uint64_t* qdo_full = bars + 1;     // [4]
    uint64_t* qdo_empty = bars + 5;    // [4]
    uint64_t* s_full = bars + 9;       // [2]
    uint64_t* p_ready = bars + 11;     // [2]
    uint64_t* dq_full = bars + 13;
    uint64_t* dq_free = bars + 14;
    uint64_t* all_done = bars + 15;    // every product of the item has completed: K / V tiles and dK / dV are final
    uint64_t* acc_free = bars + 16;    // dK / dV accumulators have been read out
    uint64_t* kv_ready = bars + 17;    // K / V copied into TMEM
    uint64_t* kv_full = bars + 18;     // [2] K / V tiles of item w landed in slot w & 1
    uint32_t* tmem_base_ptr = reinterpret_cast<uint32_t*>(bars + 20);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // provably warp-uniform
    const int C = H * D;
    const bool causal = mask_mode == MGB_MASK_CAUSAL;
    const bool tril = mask_mode == MGB_MASK_ADD_TRIL_ONES;
    const int nq = (T + BQ - 1) / BQ;
    const BwdSched sched = bwd_sched(B, H, nq, causal);        // key blocks == query blocks (BKV == BQ)

    if (warp == kTmaWarp && lane == 0) {
        tma_prefetch_desc(&tmap_kv);
        tma_prefetch_desc(&tmap_q);
        tma_prefetch_desc(&tmap_do);
        tma_prefetch_desc(&tmap_dq);
        mbar_init(&kv_full[0], 1); mbar_init(&kv_full[1], 1);
        for (int s = 0; s < kSlots; ++s) { mbar_init(&qdo_full[s], 1); mbar_init(&qdo_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&s_full[s], 1); mbar_init(&p_ready[s], kEwWarps); }
        mbar_init(dq_full, 1); mbar_init(dq_free, 4); mbar_init(all_done, 1); mbar_init(acc_free, kEwWarps); mbar_init(kv_ready, kEwWarps);
        fence_mbar_init();
    }
    if (warp == kMmaWarp) tmem_alloc(tmem_base_ptr, kBwdTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_ptr;
    grid_dep_launch();
    grid_dep_wait();                                 // inputs come from the previous kernels (programmatic dependent launch)

    // running counters (identical in every role): g = sub-blocks so far, w = items so far. Every item has an
    // even number of sub-blocks, so g & 1 is also the sub-block's position inside its 128-query block.
    if (warp == kTmaWarp) {
        // ===================== TMA producer (whole warp runs the loop, one lane issues) =====================
        {
            const bool issuer = elect_one();
            uint32_t g = 0, w = 0;
            auto next_item = [&](int& q, BwdItem& it) {       // advance q to the next non-empty slot; false at the end
                for (;; ++q) {
                    const int st = bwd_item(sched, q, it);
                    if (st < 0) return false;
                    if (st > 0) return true;
                }
            };
            auto load_kv = [&](const BwdItem& it, uint32_t slot) {
                if (issuer) {
                    uint8_t* k = sKV + slot * 2 * TILE_BYTES;
                    mbar_arrive_expect_tx(&kv_full[slot], 2 * TILE_BYTES);
                    tma_load_2d(k, &tmap_kv, &kv_full[slot], C + it.h * D, it.b * T + it.kb * BKV);
                    tma_load_2d(k + TILE_BYTES, &tmap_kv, &kv_full[slot], 2 * C + it.h * D, it.b * T + it.kb * BKV);
                }
                __syncwarp();
            };
            int q = 0;
            BwdItem it;
            bool have = next_item(q, it);
            if (have) load_kv(it, 0);
            while (have) {
                // K / V are double-buffered: the NEXT item's tiles are fetched while this item runs, as soon as the
                // item before this one (the slot's last reader) is finished.
                int q2 = q + 1;
                BwdItem it2;
                const bool have2 = next_item(q2, it2);
                const int row0 = it.b * T;
                const int nsub = 2 * it.nblk;
                const int kv_at = (w > 0) ? 2 : 0;
                for (int n = 0; n <= nsub; ++n) {
                    if (n == kv_at && have2) {
                        if (w > 0) mbar_wait_sleep(all_done, (w - 1) & 1, 256);
                        load_kv(it2, (w + 1) & 1);
                    }
                    if (n == nsub) break;
                    const uint32_t slot = g & (kSlots - 1);
                    mbar_wait_sleep(&qdo_empty[slot], ((g >> 2) & 1) ^ 1, 256);     // four slots ahead: its wake-up may lag
                    const int qrow = row0 + it.i_begin * BQ + n * SUBQ;
                    if (issuer) {
                        mbar_arrive_expect_tx(&qdo_full[slot], 2 * SUB_BYTES);
                        tma_load_2d(sQ + slot * 2 * SUB_BYTES, &tmap_q, &qdo_full[slot], it.h * D, qrow);
                        tma_load_2d(sQ + slot * 2 * SUB_BYTES + SUB_BYTES, &tmap_do, &qdo_full[slot], it.h * D, qrow);
                    }
                    __syncwarp();
                    ++g;
                }
                ++w;
                q = q2; it = it2; have = have2;
            }
        }
    } else if (warp == kMmaWarp) {
        // ===================== MMA issuer (whole warp runs the loop, one lane issues) =====================
        {
            const bool issuer = elect_one();
            const uint32_t idesc_s = make_idesc_bf16(128, SUBQ, 0, 0);   // S^T, dP^T: 128 keys x 64 queries
            const uint32_t idesc_kn = make_idesc_bf16(128, D, 0, 1);     // dV (A in TMEM), dK (A K-major smem)
            const uint32_t idesc_nn = make_idesc_bf16(128, D, 1, 1);     // dQ (A = dS^T read MN-major)
            // descriptor bases; a K step of 16 elements is +32 B (K-major) or +2048 B (MN-major) = +2 / +128 encoded
            const uint64_t dKmn0 = make_smem_desc_sw128(smem_u32(sKV), TILE_BYTES, 1024);
            const uint64_t dQk0 = make_smem_desc_sw128(smem_u32(sQ), 0, 1024);
            const uint64_t dQmn0 = make_smem_desc_sw128(smem_u32(sQ), TILE_BYTES, 1024);
            const uint64_t dSmn0 = make_smem_desc_sw128(smem_u32(sdS), TILE_BYTES, 1024);
            constexpr uint64_t kSlotStep = (2 * SUB_BYTES) >> 4, kSubStep = SUB_BYTES >> 4, kAtomStep = TILE_BYTES >> 4;

            auto issue_scores = [&](uint32_t gg) {    // S^T and dP^T of sub-block gg into buffer gg & 1
                const uint32_t slot = gg & (kSlots - 1);
                const uint32_t bcol = (gg & 1) * SUBQ;
                mbar_wait_relaxed(&qdo_full[slot], (gg >> 2) & 1);
                tc_fence_after();
                const uint64_t dq_ = dQk0 + slot * kSlotStep, do_ = dq_ + kSubStep;
                if (issuer) {
#pragma unroll
                    for (int k = 0; k < D / 16; ++k)
                        umma_bf16_ts(tmem_base + cST + bcol, tmem_base + cKT + k * 8, dq_ + 2 * k, idesc_s, k > 0 ? 1u : 0u);
#pragma unroll
                    for (int k = 0; k < D / 16; ++k)
                        umma_bf16_ts(tmem_base + cDPT + bcol, tmem_base + cVT + k * 8, do_ + 2 * k, idesc_s, k > 0 ? 1u : 0u);
                    umma_commit(&s_full[gg & 1]);
                }
                __syncwarp();
            };

            uint32_t g = 0, w = 0;
            for (int q = 0;; ++q) {
                BwdItem it;
                const int st = bwd_item(sched, q, it);
                if (st < 0) break;
                if (st == 0) continue;
                const int nsub = 2 * it.nblk;
                const uint64_t dKmn = dKmn0 + (uint64_t)(w & 1) * 2 * kAtomStep;   // this item's K tile
                mbar_wait_relaxed(kv_ready, w & 1);              // K / V are in TMEM (and the K tile in shared memory)
                tc_fence_after();
                issue_scores(g);
                for (int n = 0; n < nsub; ++n, ++g) {
                    if (n + 1 < nsub) issue_scores(g + 1);
                    mbar_wait_relaxed(&p_ready[g & 1], (g >> 1) & 1);
                    if (n == 0 && w > 0) mbar_wait_relaxed(acc_free, (w - 1) & 1);   // dK / dV of the previous item are out
                    tc_fence_after();
                    const uint32_t slot = g & (kSlots - 1);
                    const uint64_t qmn = dQmn0 + slot * kSlotStep, omn = qmn + kSubStep;
                    const uint32_t bcol = (g & 1) * SUBQ;    // 16 queries = one warpgroup's 8 packed columns, CPT apart
                    if (issuer) {
#pragma unroll
                        for (int kk = 0; kk < SUBQ / 16; ++kk)   // dV += P^T dO   (reduction over the 64 queries)
                            umma_bf16_ts(tmem_base + cDV, tmem_base + cST + bcol + kk * CPT, omn + 128 * kk,
                                         idesc_kn, (n > 0 || kk > 0) ? 1u : 0u);
#pragma unroll
                        for (int kk = 0; kk < SUBQ / 16; ++kk)   // dK += dS^T Q
                            umma_bf16_ts(tmem_base + cDK, tmem_base + cDPT + bcol + kk * CPT, qmn + 128 * kk,
                                         idesc_kn, (n > 0 || kk > 0) ? 1u : 0u);
                        umma_commit(&qdo_empty[slot]);           // last readers of this Q / dO sub-tile
                    }
                    __syncwarp();
                    if (g & 1) {
                        const uint32_t jg = g >> 1;
                        if (jg > 0) { mbar_wait_relaxed(dq_free, (jg - 1) & 1); tc_fence_after(); }
                        const uint64_t dsmn = dSmn0 + (uint64_t)(jg & 1) * 2 * kAtomStep;
                        if (issuer) {
#pragma unroll
                            for (int kk = 0; kk < BKV / 16; ++kk)   // dQ_j = dS K    (reduction over the 128 keys)
                                umma_bf16_ss(tmem_base + cDQ, dsmn + 128 * kk, dKmn + 128 * kk, idesc_nn, kk > 0 ? 1u : 0u);
                            umma_commit(dq_full);
                        }
                        __syncwarp();
                    }
                }
                if (issuer) umma_commit(all_done);
                __syncwarp();
                ++w;
            }
        }
    } else if (warp >= kDrainWarp) {
        // ===================== dQ drain warps: thread = query row of the block ==========
        const int quarter = warp & 3;
        const int r = quarter * 32 + lane;
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
        const uint32_t swz = (uint32_t)(r & 7);
        const bool elected = (warp == kDrainWarp) && (lane == 0);
        uint32_t jg = 0;
        for (int q = 0;; ++q) {
            BwdItem it;
            const int st = bwd_item(sched, q, it);
            if (st < 0) break;
            if (st == 0) continue;
            for (int j = 0; j < it.nblk; ++j, ++jg) {
                mbar_wait_sleep(dq_full, jg & 1, 64);
                tc_fence_after();
                uint32_t lo[32], hi[32];
                tmem_ld_32x32b_x32(taddr + cDQ, lo);
                tmem_ld_32x32b_x32(taddr + cDQ + 32, hi);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(dq_free);          // the next dQ product may overwrite the accumulator
                if (elected) bulk_wait_read<0>();             // both boxes have been read by the previous reduce
                named_bar_sync(3, 128);
                // rows of queries >= T hold exact zeros (their P is zero), so adding them is harmless
                uint8_t* rowp = sDQ + r * 128;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    sts128(rowp + (((uint32_t)u ^ swz) << 4), make_uint4(lo[u * 4], lo[u * 4 + 1], lo[u * 4 + 2], lo[u * 4 + 3]));
                    sts128(rowp + TILE_BYTES + (((uint32_t)u ^ swz) << 4),
                           make_uint4(hi[u * 4], hi[u * 4 + 1], hi[u * 4 + 2], hi[u * 4 + 3]));
                }
                fence_proxy_async_smem();
                named_bar_sync(3, 128);
                if (elected) {
                    const int qrow = it.b * T + (it.i_begin + j) * BQ;
#ifndef MGB_ATTN_NO_DQ_REDUCE
                    tma_reduce_add_2d(&tmap_dq, sDQ, it.h * D, qrow);
                    tma_reduce_add_2d(&tmap_dq, sDQ + TILE_BYTES, it.h * D + 32, qrow);
#endif
                    bulk_commit();
                }
            }
        }
        if (elected) bulk_wait<0>();                          // every reduce-add has completed before the CTA exits
    } else {
        // ===================== elementwise warps: thread = key row, warpgroup = CPT of the 64 queries ==========
        const int quarter = warp & 3, wg = warp >> 2;
        const int r = quarter * 32 + lane;               // key row inside the block == TMEM lane
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
        const uint32_t swz = (uint32_t)(r & 7);
        const uint64_t drop_seed = (DROP && !drop_bits) ? *seed_dev : 0;
        const uint64_t drop_t_off = dropout_bits_words_q(B * H, T);     // the transposed words follow the query-major ones
        uint32_t g = 0, w = 0;
        for (int q = 0;; ++q) {
            BwdItem it;
            const int st = bwd_item(sched, q, it);
            if (st < 0) break;
            if (st == 0) continue;
            const int kb = it.kb, k0 = kb * BKV, key = k0 + r;
            const int64_t bh_off = ((int64_t)it.b * H + it.h) * T;
            // per query: (bias - lse) in the log2 domain and delta * scale, so the inner loop is
            //   p = exp2(fma(s, c, a));  ds = p * fma(dp, scale, -b).
            // Each warp keeps its OWN 16 pairs of the current sub-block in a private 128-byte shared-memory slot: lanes
            // 0-15 fetch them one sub-block ahead into a register and publish them after a __syncwarp -- no CTA-wide
            // rendezvous, the warps drift apart freely.
            // (the fetch returns the RAW values: touching them here would stall on the load it has just issued)
            // (BQ = 2 SUBQ: sub-block n of the item starts at query i_begin * BQ + n * SUBQ; pointers hoisted per item)
            const int tq0 = it.i_begin * BQ + wg * CPT + (lane & 15);
            const float* lse_p = lse + bh_off + tq0;
            const float* del_p = delta + bh_off + tq0;
            const int nsub_item = 2 * it.nblk;
            auto fetch_ld = [&](int n_local) {
                float2 v = make_float2(INFINITY, 0.f);        // queries past the end (and lanes 16-31): lse = +inf -> p = 0
                if (lane < 16 && n_local < nsub_item && tq0 + n_local * SUBQ < T)
                    v = make_float2(lse_p[n_local * SUBQ], del_p[n_local * SUBQ]);
                return v;
            };
            float2 nxt = fetch_ld(0);
            // DROP with pre-generated masks: the transposed keep words (dropout.cuh) -- one word = 32 queries of this
            // thread's key, consecutive lanes = consecutive keys
            const int kwn = (T + 31) >> 5;
            const uint32_t* kq_p = (DROP && drop_bits && key < kwn * 32)
                ? drop_bits + drop_t_off + (uint64_t)(it.b * H + it.h) * kwn * (uint64_t)(kwn * 32) + key : nullptr;
            auto fetch_kq = [&](int n_local) -> uint32_t {
                const int qs = it.i_begin * BQ + n_local * SUBQ + wg * CPT;      // first query of this thread in that sub-block
                return (kq_p && n_local < nsub_item && qs < T) ? __ldg(kq_p + (uint64_t)(qs >> 5) * (uint64_t)(kwn * 32)) : 0u;
            };
            uint32_t kq_nxt = (DROP && drop_bits) ? fetch_kq(0) : 0u;
            {   // this item's K and V rows -> TMEM (bf16): 16 features of row r per thread. The previous item's products
                // are complete (all_done was waited for in its epilogue).
                mbar_wait_relaxed(&kv_full[w & 1], (w >> 1) & 1);
                const uint8_t* sK = sKV + (w & 1) * 2 * TILE_BYTES;
                const uint8_t* sV = sK + TILE_BYTES;
                uint32_t kreg[8], vreg[8];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const uint32_t off = (uint32_t)r * 128 + ((((uint32_t)(wg * 2 + u)) ^ swz) << 4);
                    const uint4 k4 = lds128(sK + off);
                    const uint4 v4 = lds128(sV + off);
                    kreg[u * 4 + 0] = k4.x; kreg[u * 4 + 1] = k4.y; kreg[u * 4 + 2] = k4.z; kreg[u * 4 + 3] = k4.w;
                    vreg[u * 4 + 0] = v4.x; vreg[u * 4 + 1] = v4.y; vreg[u * 4 + 2] = v4.z; vreg[u * 4 + 3] = v4.w;
                }
                tmem_st_32x32b_x8(taddr + cKT + wg * 8, kreg);
                tmem_st_32x32b_x8(taddr + cVT + wg * 8, vreg);
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(kv_ready);
            }

            for (int n = 0; n < 2 * it.nblk; ++n, ++g) {
                const int j = n >> 1;
                const uint32_t half = g & 1, jg = g >> 1;
                const int i = it.i_begin + j;
                const bool diag = causal && (i == kb);
                const bool need_mask = diag || (k0 + BKV > T) || (i * BQ + BQ > T);
                const bool bias_diag = tril && (i == kb);
                // this warp's slot (its last readers passed the __syncwarp below): a[16] = bias - lse * log2e, then
                // nb[16] = -delta * scale, so that a PAIR of queries is one 8-byte operand of the packed math below
                float* lda = reinterpret_cast<float*>(sLD + warp * CPT);
                float* ldb = lda + CPT;
                if (lane < CPT) {
                    lda[lane] = ((tril && i > kb) ? LOG2E : 0.f) - nxt.x * LOG2E;
                    ldb[lane] = -nxt.y * scale;
                }
                nxt = fetch_ld(n + 1);
                __syncwarp();
                const int qc0 = (int)half * SUBQ + wg * CPT;         // first query column (inside the block) of this thread
                // bit e of kq: the probability of query qc0 + e was kept. The word was fetched a sub-block ahead (a load
                // issued here would sit on this iteration's critical path: the products are usually ready by now).
                uint32_t kq = (DROP && drop_bits) ? (kq_nxt >> ((n * SUBQ + wg * CPT) & 16)) : 0xffffu;
                if (DROP && drop_bits) kq_nxt = fetch_kq(n + 1);
                mbar_wait_relaxed(&s_full[half], jg & 1);
                tc_fence_after();
                uint32_t sv[CPT], dp[CPT];
                tmem_ld_cols(taddr + cST + half * SUBQ + wg * CPT, sv);
#if MGB_ATTN_BWD_SPLIT_LD
                // the scores first; dP^T stays in flight while the exponentials of the unmasked path run (it is not
                // needed before the second half of the arithmetic). Every path waits for it before it touches dp or
                // stores over those TMEM columns.
                tmem_ld_wait();
                tmem_ld_cols(taddr + cDPT + half * SUBQ + wg * CPT, dp);
#else
                tmem_ld_cols(taddr + cDPT + half * SUBQ + wg * CPT, dp);
                tmem_ld_wait();
#endif
                // DROP: ks = keep ? scale / (1-p) : 0 replaces `scale` in dS = P (ks dP - delta scale) and multiplies P
                // where it is packed for dV (the item epilogue takes the factor `scale` out again).
                const float iks = inv_keep * scale;
                if (DROP && !drop_bits) {
                    static_assert(CPT == 16, "two Philox calls per lane cover 4 key groups x 16 queries");
                    const uint64_t t8 = (uint64_t)((T + 7) >> 3);
                    const uint64_t qrow = ((uint64_t)it.b * H + it.h) * T + (uint64_t)(i * BQ + qc0 + (lane >> 2));
                    const uint64_t kgrp = (uint64_t)((k0 + 32 * quarter) >> 3) + (uint64_t)(lane & 3);
                    const uint32_t m0 = dropout_keep8(drop_seed, site, qrow * t8 + kgrp, drop_thresh);         // queries 0-7
                    const uint32_t m1 = dropout_keep8(drop_seed, site, (qrow + 8) * t8 + kgrp, drop_thresh);   // queries 8-15
                    kq = 0u;
#pragma unroll
                    for (int e = 0; e < CPT; ++e) {
                        const uint32_t v = __shfl_sync(0xffffffffu, e < 8 ? m0 : m1, (e & 7) * 4 + (lane >> 3));
                        kq |= ((v >> (lane & 7)) & 1u) << e;
                    }
                }
                uint32_t pk[CPT / 2], dk[CPT / 2];
                // causal diagonal block with every key / query present: key row r sees query column c iff r <= c. This
                // warp's 32 rows against this thread group's 16 columns [qc0, qc0 + 16) are, warp-uniformly, either all
                // visible (rows end before qc0: the unmasked path), all hidden (rows start at or after qc0 + 16: zeros, no
                // exponentials) or -- one warp in four -- mixed (the per-element path below).
                const bool diag_only = diag && !bias_diag && (k0 + BKV <= T) && (i * BQ + BQ <= T);
                const bool all_dead = diag_only && (quarter * 32 >= qc0 + CPT);
                const bool all_live = diag_only && (quarter * 32 + 31 <= qc0);
                if (all_dead) {
#if MGB_ATTN_BWD_SPLIT_LD
                    tmem_ld_wait();
#endif
#pragma unroll
                    for (int e = 0; e < CPT / 2; ++e) { pk[e] = 0u; dk[e] = 0u; }
#if MGB_ATTN_BWD_SPLIT_LD
                } else if ((!need_mask && !bias_diag) || all_live) {
                    // p = exp2(s * c + a) for all 16 queries, then (dP^T has arrived) ds = p * (dp * scale + nb): fp32 pairs
                    const float2 sc2 = splat2(scale_log2), scl2 = splat2(scale);
                    float2 ppv[CPT / 2];
#pragma unroll
                    for (int q4 = 0; q4 < CPT / 4; ++q4) {
                        const float4 a4 = lds128f(lda + 4 * q4);
#pragma unroll
                        for (int h2 = 0; h2 < 2; ++h2) {
                            const int e = 2 * q4 + h2;
                            const float2 av = h2 ? make_float2(a4.z, a4.w) : make_float2(a4.x, a4.y);
                            const float2 x = fma2(make_float2(__uint_as_float(sv[2 * e]), __uint_as_float(sv[2 * e + 1])), sc2, av);
                            ppv[e] = make_float2(exp2f(x.x), exp2f(x.y));
                        }
                    }
                    tmem_ld_wait();
#pragma unroll
                    for (int q4 = 0; q4 < CPT / 4; ++q4) {
                        const float4 b4 = lds128f(ldb + 4 * q4);
#pragma unroll
                        for (int h2 = 0; h2 < 2; ++h2) {
                            const int e = 2 * q4 + h2;
                            const float2 bv = h2 ? make_float2(b4.z, b4.w) : make_float2(b4.x, b4.y);
                            const float2 pp = ppv[e];
                            const float2 ks2 = DROP ? make_float2((kq & (1u << (2 * e))) ? iks : 0.f, (kq & (2u << (2 * e))) ? iks : 0.f) : scl2;
                            const float2 tt = fma2(make_float2(__uint_as_float(dp[2 * e]), __uint_as_float(dp[2 * e + 1])), ks2, bv);
                            const float2 ds = mul2(pp, tt);
                            const float2 pm = DROP ? mul2(pp, ks2) : pp;
                            pk[e] = pack_bf16x2(pm.x, pm.y);
                            dk[e] = pack_bf16x2(ds.x, ds.y);
                        }
                    }
#else
                } else if ((!need_mask && !bias_diag) || all_live) {
                    // p = exp2(s * c + a), ds = p * (dp * scale + nb) on fp32 pairs: 3 packed instructions + 2 exponentials
                    // + 2 packs per pair of queries (the scalar form took 10 issue slots)
                    const float2 sc2 = splat2(scale_log2), scl2 = splat2(scale);
#pragma unroll
                    for (int q4 = 0; q4 < CPT / 4; ++q4) {
                        const float4 a4 = lds128f(lda + 4 * q4), b4 = lds128f(ldb + 4 * q4);
#pragma unroll
                        for (int h2 = 0; h2 < 2; ++h2) {
                            const int e = 2 * q4 + h2;
                            const float2 av = h2 ? make_float2(a4.z, a4.w) : make_float2(a4.x, a4.y);
                            const float2 bv = h2 ? make_float2(b4.z, b4.w) : make_float2(b4.x, b4.y);
                            const float2 x = fma2(make_float2(__uint_as_float(sv[2 * e]), __uint_as_float(sv[2 * e + 1])), sc2, av);
                            const float2 pp = make_float2(exp2f(x.x), exp2f(x.y));
                            const float2 ks2 = DROP ? make_float2((kq & (1u << (2 * e))) ? iks : 0.f, (kq & (2u << (2 * e))) ? iks : 0.f) : scl2;
                            const float2 tt = fma2(make_float2(__uint_as_float(dp[2 * e]), __uint_as_float(dp[2 * e + 1])), ks2, bv);
                            const float2 ds = mul2(pp, tt);
                            const float2 pm = DROP ? mul2(pp, ks2) : pp;
                            pk[e] = pack_bf16x2(pm.x, pm.y);
                            dk[e] = pack_bf16x2(ds.x, ds.y);
                        }
                    }
#endif
                } else {
#if MGB_ATTN_BWD_SPLIT_LD
                    tmem_ld_wait();
#endif
#pragma unroll
                    for (int e = 0; e < CPT / 2; ++e) {
                        float pv[2], dsv[2];
#pragma unroll
                        for (int t2 = 0; t2 < 2; ++t2) {
                            const int col = qc0 + 2 * e + t2;          // query column inside the block
                            const float2 ldv = make_float2(lda[2 * e + t2], ldb[2 * e + t2]);
                            float x = fmaf(__uint_as_float(sv[2 * e + t2]), scale_log2, ldv.x);
                            if (bias_diag && r <= col) x += LOG2E;
                            const bool dead = (key >= T) || (diag && r > col);   // queries >= T carry a = -inf
                            const float p = dead ? 0.f : exp2f(x);
                            const float ks = DROP ? ((kq & (1u << (2 * e + t2))) ? iks : 0.f) : scale;
                            pv[t2] = DROP ? p * ks : p;
                            dsv[t2] = p * fmaf(__uint_as_float(dp[2 * e + t2]), ks, ldv.y);
                        }
                        pk[e] = pack_bf16x2(pv[0], pv[1]);
                        dk[e] = pack_bf16x2(dsv[0], dsv[1]);
                    }
                }
                // bf16 pairs over the first half of this thread's own fp32 columns (both were just read)
                tmem_st_cols(taddr + cST + half * SUBQ + wg * CPT, pk);
                tmem_st_cols(taddr + cDPT + half * SUBQ + wg * CPT, dk);
                // and dS^T once more in shared memory for dQ: row `r` of atom g & 3, queries [wg*CPT, +CPT): CPT/8 units
                uint8_t* rowp = sdS + (g & 3) * TILE_BYTES + r * 128;
#pragma unroll
                for (int u = 0; u < CPT / 8; ++u) {
                    const uint32_t jj = (uint32_t)(wg * (CPT / 8) + u);
                    sts128(rowp + ((jj ^ swz) << 4), make_uint4(dk[u * 4 + 0], dk[u * 4 + 1], dk[u * 4 + 2], dk[u * 4 + 3]));
                }
                tmem_st_wait();
                tc_fence_before();
                // the shared-memory dS^T is read (by dQ) only after the block's second sub-block: this thread's fence
                // there also covers what it wrote in the first
                if (half == 1) fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_ready[half]);
            }
            // ---- item epilogue: dK / dV of this key block ----
            mbar_wait_relaxed(all_done, w & 1);
            tc_fence_after();
            uint32_t vk[CPT], vv[CPT];
            tmem_ld_cols(taddr + cDK + wg * CPT, vk);
            tmem_ld_cols(taddr + cDV + wg * CPT, vv);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_free);            // the next item may start accumulating
            if (DROP) {                                      // P carried scale / (1-p) into dV: take `scale` out again
                const float inv_scale = 1.0f / scale;
#pragma unroll
                for (int e = 0; e < CPT; ++e) vv[e] = __float_as_uint(__uint_as_float(vv[e]) * inv_scale);
            }
            if (kv_colsum) {
                // bias gradient of the qkv projection, K and V thirds: column sums over this warp's 32 key rows by a
                // transpose-reduce (rows of keys >= T hold exact zeros), one atomic per column and warp
                float tk[8], tv[8];
                const bool b16 = lane & 16, b8 = lane & 8, b4 = lane & 4, b2 = lane & 2;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float klo = __uint_as_float(vk[i]), khi = __uint_as_float(vk[i + 8]);
                    const float vlo = __uint_as_float(vv[i]), vhi = __uint_as_float(vv[i + 8]);
                    tk[i] = (b16 ? khi : klo) + __shfl_xor_sync(0xffffffffu, b16 ? klo : khi, 16);
                    tv[i] = (b16 ? vhi : vlo) + __shfl_xor_sync(0xffffffffu, b16 ? vlo : vhi, 16);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    tk[i] = (b8 ? tk[i + 4] : tk[i]) + __shfl_xor_sync(0xffffffffu, b8 ? tk[i] : tk[i + 4], 8);
                    tv[i] = (b8 ? tv[i + 4] : tv[i]) + __shfl_xor_sync(0xffffffffu, b8 ? tv[i] : tv[i + 4], 8);
                }
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    tk[i] = (b4 ? tk[i + 2] : tk[i]) + __shfl_xor_sync(0xffffffffu, b4 ? tk[i] : tk[i + 2], 4);
                    tv[i] = (b4 ? tv[i + 2] : tv[i]) + __shfl_xor_sync(0xffffffffu, b4 ? tv[i] : tv[i + 2], 4);
                }
                float sk = (b2 ? tk[1] : tk[0]) + __shfl_xor_sync(0xffffffffu, b2 ? tk[0] : tk[1], 2);
                float sv = (b2 ? tv[1] : tv[0]) + __shfl_xor_sync(0xffffffffu, b2 ? tv[0] : tv[1], 2);
                sk += __shfl_xor_sync(0xffffffffu, sk, 1);       // lanes l and l^1 hold the two halves of the rows
                sv += __shfl_xor_sync(0xffffffffu, sv, 1);
                if ((lane & 1) == 0) {
                    const int col = it.h * D + wg * CPT + (lane >> 1);      // bits 4..1 of the lane picked the column
                    atomicAdd(kv_colsum + C + col, sk);
                    atomicAdd(kv_colsum + 2 * C + col, sv);
                }
            }
            if (key < T) {
                __nv_bfloat16* krow = dqkv + ((int64_t)(it.b * T + key)) * (3 * C) + C + (int64_t)it.h * D + wg * CPT;
#pragma unroll
                for (int u = 0; u < CPT / 8; ++u) {
                    stg128(krow + u * 8,
                           make_uint4(pack_bf16x2(__uint_as_float(vk[u * 8 + 0]), __uint_as_float(vk[u * 8 + 1])),
                                      pack_bf16x2(__uint_as_float(vk[u * 8 + 2]), __uint_as_float(vk[u * 8 + 3])),
                                      pack_bf16x2(__uint_as_float(vk[u * 8 + 4]), __uint_as_float(vk[u * 8 + 5])),
                                      pack_bf16x2(__uint_as_float(vk[u * 8 + 6]), __uint_as_float(vk[u * 8 + 7]))));
                    stg128(krow + C + u * 8,
                           make_uint4(pack_bf16x2(__uint_as_float(vv[u * 8 + 0]), __uint_as_float(vv[u * 8 + 1])),
                                      pack_bf16x2(__uint_as_float(vv[u * 8 + 2]), __uint_as_float(vv[u * 8 + 3])),
                                      pack_bf16x2(__uint_as_float(vv[u * 8 + 4]), __uint_as_float(vv[u * 8 + 5])),
                                      pack_bf16x2(__uint_as_float(vv[u * 8 + 6]), __uint_as_float(vv[u * 8 + 7]))));
                }
            }
            ++w;
        }
    }

    __syncwarp();
    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kBwdTmemCols);
    }
}

// main backward kernel only: delta must be precomputed and dq_accum zeroed by the caller (attention.cu)
int launch_bwd_main(const void* qkv, const void* dout, const float* lse, const float* delta, float* dq_accum,
                    void* dqkv, float* dqkv_colsum, int B, int T, int H, int mask_mode, float scale, cudaStream_t stream,
                    uint32_t drop_thresh, float inv_keep, const uint64_t* seed_dev, uint32_t site, const uint32_t* drop_bits) {
    static PerDeviceOnce once;
    once.run([] {
        cudaFuncSetAttribute(attn_bwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM);
        cudaFuncSetAttribute(attn_bwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM);
    });
    CUtensorMap tkv, tq, td, tdq;
    const int64_t C = (int64_t)H * D, C3 = 3 * C;
    int rc = gemm::make_map_2d(&tkv, qkv, C3, (int64_t)B * T, C3, 64, 128);
    if (rc) return rc;
    rc = gemm::make_map_2d(&tq, qkv, C3, (int64_t)B * T, C3, 64, SUBQ);
    if (rc) return rc;
    rc = gemm::make_map_2d(&td, dout, C, (int64_t)B * T, C, 64, SUBQ);
    if (rc) return rc;
    rc = gemm::make_map_2d(&tdq, dq_accum, C, (int64_t)B * T, C, 32, 128, /*f32=*/true);
    if (rc) return rc;
    const int nq = (T + BKV - 1) / BKV;
    const int nunits = (mask_mode == MGB_MASK_CAUSAL ? (nq + 1) / 2 : nq) * B * H;
    const int grid = nunits < num_sms() ? nunits : num_sms();
    if (drop_thresh)
        launch_pdl(attn_bwd_tc_kernel<true>, dim3(grid), dim3(kBwdThreads), (size_t)BWD_SMEM, stream, tkv, tq, td, tdq, lse, delta,
                   (__nv_bfloat16*)dqkv, dqkv_colsum, B, T, H, mask_mode, scale, scale * LOG2E, drop_thresh, inv_keep, seed_dev, site,
                   drop_bits);
    else
        launch_pdl(attn_bwd_tc_kernel<false>, dim3(grid), dim3(kBwdThreads), (size_t)BWD_SMEM, stream, tkv, tq, td, tdq, lse, delta,
                   (__nv_bfloat16*)dqkv, dqkv_colsum, B, T, H, mask_mode, scale, scale * LOG2E, 0u, 1.0f,
                   (const uint64_t*)nullptr, 0u, (const uint32_t*)nullptr);
    return kOk;
}

}  // namespace attn_tc
}  // namespace mgb
