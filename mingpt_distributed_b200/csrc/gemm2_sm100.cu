// BF16 GEMM, CTA-pair edition: two SMs of a cluster cooperate on one 256 x BLOCK_N tile
// (tcgen05.mma.cta_group::2). Same contract as gemm_sm100.cu (see mingpt_b200.h, mgb_gemm_bf16).
//
// Why pairs: with both operands in shared memory a single-CTA 128 x 256 MMA needs ~190 B/clk of shared
// memory traffic (operand reads + TMA fills) against the SM's 128 B/clk, which caps the tensor pipe near
// 65 % (measured 62 % in profiles/r1_a). In a pair each SM stages its own 128 rows of A but only HALF of
// B; the other half is read from the peer's shared memory by the hardware, so traffic per SM drops by a
// third while the tile -- and the work per MMA instruction -- doubles.
//
// Per CTA:  warp 0 TMA producer (own A rows, own half of B; bytes are credited to the LEADER's barrier),
//           warp 1 MMA issuer (leader CTA only; commits are multicast to both CTAs' barriers),
//           warp 2 TMEM allocator, warps 4-11 epilogue for this CTA's 128 rows.
// Epilogue: all global traffic of the epilogue goes through 128B-swizzled shared-memory boxes and TMA:
//           the residual / gate tile is TMA-LOADED into the box while the tile's MMAs still run, each
//           thread then updates its own row in place (TMEM -> registers -> bias, GELU, gate, residual) and
//           one thread TMA-STORES the box (or TMA-reduce-adds it in fp32 for split-K / accumulation).
#include "gemm_common.cuh"
#include "dropout.cuh"
#ifndef MGB_GEMM_EPI_SLEEP_NS
#define MGB_GEMM_EPI_SLEEP_NS 100
#endif

namespace mgb {
namespace gemm {

constexpr int kMaxStages = 8;
constexpr int kBoxBytes = 128 * 128;          // 128 rows x 128 bytes
constexpr int kNumBoxes = 4;                  // two per epilogue half
constexpr int kSmemLimit = 232448;            // 227 KiB
constexpr int kBarBytes = 320;                // mbarriers (200 B) + TMEM base + two 16-byte work-queue responses at +256

// Fallback epilogue for outputs TMA cannot address (row pitch not a multiple of 16 bytes): registers -> global.
__device__ __forceinline__ void epilogue_direct(const Params& p, float (&v)[32], int64_t row, int64_t col0,
                                                bool vec_ok) {
    const bool full = (col0 + 32 <= p.N) && vec_ok;
    if (full) {
        bool acted = false;
        if (p.preact) {
            __nv_bfloat16* dst = p.preact + row * p.ldp + col0;
            uint32_t pw[16];
            acted = preact_words(v, pw, p.act, p.act_grad);
#pragma unroll
            for (int q = 0; q < 4; ++q) stg128(dst + q * 8, make_uint4(pw[q * 4], pw[q * 4 + 1], pw[q * 4 + 2], pw[q * 4 + 3]));
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
                for (int q = 0; q < 8; ++q) red_add_f32x4(dst + q * 4, v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
            } else {
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    *reinterpret_cast<float4*>(dst + q * 4) = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
            }
        } else {
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.d) + row * p.ldd + col0;
#pragma unroll
            for (int q = 0; q < 4; ++q)
                stg128(dst + q * 8, make_uint4(pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]), pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]),
                                               pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]), pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7])));
        }
    } else {
        const bool tanh_kind = p.act == MGB_ACT_GELU_TANH;
        const bool gate_tanh = p.gate_kind == MGB_ACT_GELU_TANH;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const int64_t col = col0 + j;
            if (col >= p.N) continue;
            float x = v[j];
            if (p.preact) p.preact[row * p.ldp + col] = __float2bfloat16(p.act_grad ? apply_act_grad(x, p.act) : x);
            if (p.act) x = tanh_kind ? gelu_tanh_f(x) : gelu_erf_f(x);
            if (p.gate) {
                const float g = __bfloat162float(p.gate[row * p.ldg + col]);
                x *= apply_gate(g, p.gate_kind);
            }
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
}

// COLSUM: the epilogue also reduces every output tile over its rows (a 31-shuffle transpose-reduce per 32x32 chunk,
// shared-memory atomics per CTA, one global atomic per column and tile) into p.colsum: the bias gradient that
// goes with an output-gradient GEMM, without another pass over D. A separate instantiation, so the register
// allocation of the plain kernel is untouched.
// MODE: the epilogue specialised at compile time for the combinations the training step uses, so that each hot
// instantiation carries only its own code (the all-in-one epilogue spent ~15 % of its warp time on instruction-cache
// misses: ncu "no_inst"):
//   0  generic: everything decided at run time (erf GELU, stores that TMA cannot address, direct gate / residual loads)
//   1  bf16 D through TMA, one output, no activation; optional bias; optional residual / gate-multiply tile through TMA
//   2  bf16 D + saved GELU' through TMA, NewGELU(tanh), bias                                   (c_fc forward)
//   3  fp32 D through TMA store or reduce-add, no epilogue terms                               (weight gradients)
//   4  mode 1 + dropout between the bias and the residual tile                  (c_proj forward of a block, training
//      with resid_drop > 0): 2 Philox calls per 32 columns of a row, hidden behind the next tile's main loop
template <bool COLSUM, int MODE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_d, const __grid_constant__ CUtensorMap tmap_p,
                 const __grid_constant__ CUtensorMap tmap_x, const Params p) {
    // No static shared memory in this kernel, so the dynamic window starts 1024-aligned (checked): no alignment slack --
    // the 3 KiB left beside five 32 KiB stages and the staging boxes are needed for the barriers and the bias tiles.
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw;
    if (smem_u32(smem) & 1023u) __trap();
    const int block_n = p.block_n;
    const int nb = block_n >> 1;                               // B columns staged by this CTA
    const uint32_t stage_bytes = BLOCK_M * BLOCK_K * 2 + (uint32_t)nb * BLOCK_K * 2;
    uint8_t* boxes = smem + p.num_stages * stage_bytes;        // kNumBoxes x 16 KiB, 1024-aligned
    uint64_t* bars = reinterpret_cast<uint64_t*>(boxes + kNumBoxes * kBoxBytes);
    uint64_t* full_bar = bars;                       // [kMaxStages]  (leader's copy is the live one)
    uint64_t* empty_bar = bars + kMaxStages;         // [kMaxStages]
    uint64_t* tmem_full = bars + 2 * kMaxStages;     // [2]
    uint64_t* tmem_empty = tmem_full + 2;            // [2]           (leader's copy is the live one)
    uint64_t* aux_bar = tmem_empty + 2;              // [kNumBoxes]   residual / gate tile landed in box i
    uint32_t* tmem_base_ptr = reinterpret_cast<uint32_t*>(aux_bar + kNumBoxes);
    uint64_t* clc_full = aux_bar + kNumBoxes + 1;    // [2]  a work-queue response has landed in this CTA
    uint64_t* clc_empty = clc_full + 2;              // [2]  (leader's copy is the live one) every reader of both CTAs is done with it
    uint8_t* clc_resp = reinterpret_cast<uint8_t*>(bars) + 256;                        // [2] x 16 bytes
    float* bias_s = reinterpret_cast<float*>(boxes + kNumBoxes * kBoxBytes + kBarBytes);   // [2][256] bias of the current / next tile
    float* cs_smem = bias_s + (p.bias ? 512 : 0);                                      // [256] (COLSUM only; behind the bias if both)
    if (COLSUM) {
        for (int c = threadIdx.x; c < 256; c += blockDim.x) cs_smem[c] = 0.f;
    }

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // provably warp-uniform
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int cluster_id = blockIdx.x >> 1;
    const int num_clusters = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        if (MODE || p.tma_store) {
            tma_prefetch_desc(&tmap_d);
            if (p.preact) tma_prefetch_desc(&tmap_p);
            if (p.tma_aux) tma_prefetch_desc(&tmap_x);
        }
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < p.num_stages; ++s) {
            mbar_init(&full_bar[s], 2);      // leader's arrive.expect_tx + the peer's remote arrive
            mbar_init(&empty_bar[s], 1);     // one multicast commit
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], 2 * kEpiWarps);   // epilogue warps of BOTH CTAs
        }
        for (int s = 0; s < kNumBoxes; ++s) mbar_init(&aux_bar[s], 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&clc_full[s], 1);                       // this CTA's scheduler lane arms it (arrive + 16 tx bytes)
            mbar_init(&clc_empty[s], 2 * (2 + kEpiWarps) + 1);   // per CTA: producer, scheduler, 8 epilogue warps; + the MMA warp
        }
        fence_mbar_init();
    }
    if (warp == 2) tmem_alloc_cg2(tmem_base_ptr, kTmemCols);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_ptr;
    // Everything above touched only this CTA's shared memory / TMEM: under a programmatic dependent launch it ran
    // while the previous kernel of the stream was still draining. From here on its results are read.
    grid_dep_launch();
    grid_dep_wait();

    const int tiles = p.num_m_tiles * p.num_n_tiles;
    const int units = tiles * p.split_k;

    // Work distribution. Static: cluster c takes units c, c + C, c + 2C, ... of a grid of C persistent clusters.
    // Dynamic (p.dynamic): the grid has one cluster per unit; a cluster does its own unit and then keeps taking the
    // units of clusters that have not started yet through cluster launch control (the hardware's work queue) until
    // none is pending. Warp 3 of both CTAs runs the queue (the leader's asks, the response is multicast into both
    // CTAs); every other role reads the responses in order through next_unit(). With SMs taken away by a concurrent
    // kernel (NCCL during the gradient all-reduce) the remaining clusters just take more units, where a static share
    // would wait for its SM pair.
    auto next_unit = [&](int& u, uint32_t& k) -> bool {
        if (!p.dynamic) { u += num_clusters; return u < units; }
        const uint32_t slot = k & 1u;
        mbar_wait(&clc_full[slot], (k >> 1) & 1u);
        uint32_t x;
        const bool valid = clc_decode(clc_resp + slot * 16, x);
        fence_proxy_async_smem();                  // this read is ordered before the next response overwrites the slot
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&clc_empty[slot], 0);
        ++k;
        u = (int)(x >> 1);
        return valid;
    };
    if (warp == 3 && p.dynamic) {
        const bool issuer = elect_one();
        for (uint32_t k = 0;; ++k) {
            const uint32_t slot = k & 1u, ph = (k >> 1) & 1u;
            if (leader) mbar_wait(&clc_empty[slot], ph ^ 1u);          // all readers of both CTAs released the slot
            if (issuer) {
                mbar_arrive_expect_tx(&clc_full[slot], 16);
                if (leader) clc_try_cancel_multicast(clc_resp + slot * 16, &clc_full[slot]);
            }
            __syncwarp();
            mbar_wait(&clc_full[slot], ph);
            uint32_t x;
            const bool valid = clc_decode(clc_resp + slot * 16, x);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(&clc_empty[slot], 0);
            if (!valid) break;                                          // nothing pending any more: the queue is closed
        }
    }

    if (warp == 0) {
        // ===================== TMA producer (both CTAs; whole warp runs the loop, one lane issues) ==========
        {
            const bool issuer = elect_one();
            uint32_t stage = 0, phase = 0;
            int u = cluster_id;
            uint32_t qk = 0;
            for (bool more = u < units; more; more = next_unit(u, qk)) {
                const int split = u / tiles;
                const int t = u - split * tiles;
                const int mt = p.m_fast ? (t % p.num_m_tiles) : (t / p.num_n_tiles);
                const int nt = p.m_fast ? (t / p.num_m_tiles) : (t % p.num_n_tiles);
                const int m0 = mt * (2 * BLOCK_M) + (int)rank * BLOCK_M;
                const int n0 = nt * block_n + (int)rank * nb;
                const int kb0 = split * p.k_blocks_per_split;
                const int kb1 = min(kb0 + p.k_blocks_per_split, p.k_blocks_total);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * stage_bytes;
                    uint8_t* sb = sa + BLOCK_M * BLOCK_K * 2;
                    const int k0 = kb * BLOCK_K;
                    if (issuer) {
                        if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * stage_bytes);
                        else mbar_arrive_cluster(&full_bar[stage], 0);
                        if (!p.a_mn) {
                            tma_load_2d_cg2(sa, &tmap_a, &full_bar[stage], k0, m0);
                        } else {
                            tma_load_2d_cg2(sa, &tmap_a, &full_bar[stage], m0, k0);
                            tma_load_2d_cg2(sa + ATOM_BYTES, &tmap_a, &full_bar[stage], m0 + 64, k0);
                        }
                        if (!p.b_mn) {
                            tma_load_2d_cg2(sb, &tmap_b, &full_bar[stage], k0, n0);
                        } else {
                            for (int a = 0; a < nb / 64; ++a)
                                tma_load_2d_cg2(sb + a * ATOM_BYTES, &tmap_b, &full_bar[stage], n0 + a * 64, k0);
                        }
                    }
                    __syncwarp();
                    if (++stage == (uint32_t)p.num_stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only; whole warp runs the loop, one lane issues) ======
        if (leader) {
            const bool issuer = elect_one();
            const uint32_t idesc = make_idesc_bf16(2 * BLOCK_M, block_n, p.a_mn, p.b_mn);
            const uint32_t a_lbo = p.a_mn ? ATOM_BYTES : 0, b_lbo = p.b_mn ? ATOM_BYTES : 0;
            const uint32_t a_kstep = p.a_mn ? UMMA_K * 128 : UMMA_K * 2;
            const uint32_t b_kstep = p.b_mn ? UMMA_K * 128 : UMMA_K * 2;
            uint32_t stage = 0, phase = 0;
            uint32_t acc = 0, acc_phase = 0;
            int u = cluster_id;
            uint32_t qk = 0;
            for (bool more = u < units; more; more = next_unit(u, qk)) {
                const int split = u / tiles;
                const int kb0 = split * p.k_blocks_per_split;
                const int kb1 = min(kb0 + p.k_blocks_per_split, p.k_blocks_total);
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * kAccStride;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * stage_bytes);
                    const uint32_t sb = sa + BLOCK_M * BLOCK_K * 2;
                    if (issuer) {
#pragma unroll
                        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                            const uint64_t da = make_smem_desc_sw128(sa + k * a_kstep, a_lbo, 1024);
                            const uint64_t db = make_smem_desc_sw128(sb + k * b_kstep, b_lbo, 1024);
                            umma_bf16_ss_cg2(tmem_d, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                        }
                        umma_commit_cg2_mc(&empty_bar[stage], 0x3);   // frees this stage in both CTAs
                    }
                    __syncwarp();
                    if (++stage == (uint32_t)p.num_stages) { stage = 0; phase ^= 1; }
                }
                if (issuer) umma_commit_cg2_mc(&tmem_full[acc], 0x3);   // accumulators ready in both CTAs
                __syncwarp();
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue (both CTAs, own 128 rows) =====================
        const int ew = warp - 4;
        const int quarter = warp & 3;
        const int half = ew >> 2;
        // one lane of warp quarter 0 owns this half's TMA traffic (elect.sync picks the same lane every time, so the
        // bulk async-groups it commits are the ones it later waits on)
        const bool elected = (quarter == 0) && elect_one();
        const int r_local = quarter * 32 + lane;
        const uint32_t swz = (uint32_t)(r_local & 7);
        const bool vec_ok = ((p.ldd & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.d) & 15) == 0) &&
                            (!p.residual || (((p.ldr & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.residual) & 15) == 0))) &&
                            (!p.preact || (((p.ldp & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.preact) & 15) == 0))) &&
                            (!p.gate || (((p.ldg & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.gate) & 15) == 0)));
        // compile-time facts of the specialised instantiations (MODE 0 reads them from the parameters)
        const bool tma_store = MODE != 0 || p.tma_store;
        const bool d_f32 = MODE == 3 || (MODE == 0 && p.d_f32);
        const bool has_bias = MODE != 3 && p.bias != nullptr;
        const bool two_out = MODE == 2 || (MODE == 0 && p.tma_store && p.preact != nullptr);
        const int aux = (MODE == 2 || MODE == 3) ? 0 : p.tma_aux;   // 0 none, 1 residual, 2 gate: TMA-loaded in place
        const int act = MODE == 2 ? (int)MGB_ACT_GELU_TANH : (MODE == 0 ? p.act : 0);
        const int act_grad = MODE == 2 ? 1 : (MODE == 0 ? p.act_grad : 0);
        const bool direct_gate = MODE == 0 && p.gate != nullptr;       // gate / residual read from global memory by the thread
        const bool direct_res = MODE == 0 && p.residual != nullptr;
        const uint64_t drop_seed = MODE == 4 ? *p.drop_seed : 0;
        const int slice_cols = d_f32 ? 32 : 64;                   // columns per 128-byte output slice
        const int nslices = block_n / slice_cols;
        const int chunks_per_slice = slice_cols >> 5;
        uint8_t* box_base = boxes + half * 2 * kBoxBytes;
        uint64_t* aux_b = aux_bar + half * 2;
        uint32_t buf = 0;
        uint32_t aux_ph = 0;                                      // bit b: parity the next load into box b completes
        uint32_t acc = 0, acc_phase = 0;
        int u = cluster_id;
        uint32_t qk = 0;
        for (bool more = u < units; more; more = next_unit(u, qk)) {
            const int split = u / tiles;
            const int t = u - split * tiles;
            const int mt = p.m_fast ? (t % p.num_m_tiles) : (t / p.num_n_tiles);
            const int nt = p.m_fast ? (t / p.num_m_tiles) : (t % p.num_n_tiles);
            const int m0 = mt * (2 * BLOCK_M) + (int)rank * BLOCK_M;
            const int n0 = nt * block_n;
            if (aux && elected) {
                // bf16 output => at most two slices per half and tile, one per box. Fetch their residual / gate
                // tiles now, while the tile's MMAs still run. Every thread of this half is past the previous
                // tile's last named barrier, so nobody touches the boxes until the loads land.
                bulk_wait_read<0>();
                uint32_t b = buf;
                for (int sl = half; sl < nslices; sl += 2, b ^= 1) {
                    mbar_arrive_expect_tx(&aux_b[b], kBoxBytes);
                    tma_load_2d(box_base + b * kBoxBytes, &tmap_x, &aux_b[b], n0 + sl * slice_cols, m0);
                }
            }
            if (has_bias) {
                // The tile's bias (block_n <= 256 values) goes to shared memory once per tile, while the tile's MMAs
                // still run: with the shared-memory carve-out at its maximum the L1 holds next to nothing, and eight
                // 16-byte bias loads per 32-column chunk from L2 were the epilogue's largest stall (ncu: long_sb).
                // Buffer `acc` is free: its readers finished two tiles ago, before their tmem_empty arrive.
                const int c = (int)threadIdx.x - 128;
                bias_s[acc * 256 + c] = (c < block_n && n0 + c < p.N) ? __ldg(p.bias + n0 + c) : 0.f;
            }
            // the eight epilogue warps wait here for most of a tile's main loop: sleep between probes instead of polling
            // (issue slots next to the MMA / TMA threads and, on a power-capped part, energy)
            mbar_wait_sleep(&tmem_full[acc], acc_phase, MGB_GEMM_EPI_SLEEP_NS);
            tc_fence_after();
            if (has_bias) named_bar_sync(4, kEpiWarps * 32);      // bias visible to all eight epilogue warps
            const int64_t row = (int64_t)m0 + r_local;
            const bool row_ok = row < p.M;
            const uint32_t taddr_row = tmem_base + acc * kAccStride + ((uint32_t)(quarter * 32) << 16);
            bool released = false;
            for (int sl = half; sl < nslices; sl += 2) {
                uint8_t* box = box_base + buf * kBoxBytes;
                uint8_t* box2 = box_base + (buf ^ 1) * kBoxBytes;
                if (tma_store && !aux) {
                    // the box written two slices ago (or both, when two outputs share the ring) must have been read
                    if (elected) { if (two_out) bulk_wait_read<0>(); else bulk_wait_read<1>(); }
                    named_bar_sync(1 + half, 128);
                }
                if (aux) {
                    mbar_wait(&aux_b[buf], (aux_ph >> buf) & 1u);
                    aux_ph ^= (1u << buf);
                }
                uint8_t* rowp = box + r_local * 128;
                float rowdot_acc = 0.f;
                for (int ch = 0; ch < chunks_per_slice; ++ch) {
                    const int ccol = sl * slice_cols + ch * 32;      // column inside the tile
                    uint32_t r[32];
                    __syncwarp();
                    tmem_ld_32x32b_x32(taddr_row + (uint32_t)ccol, r);
                    tmem_ld_wait();
                    if (sl + 2 >= nslices && ch == chunks_per_slice - 1) {
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive_cluster(&tmem_empty[acc], 0);
                        released = true;
                    }
                    const int64_t col0 = (int64_t)n0 + ccol;
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;
                    if (has_bias) {      // this tile's bias lives in shared memory (staged once per tile, below)
                        const float* bs = bias_s + acc * 256 + ccol;
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const float4 b4 = lds128f(bs + q * 4);
                            v[q * 4 + 0] += b4.x; v[q * 4 + 1] += b4.y; v[q * 4 + 2] += b4.z; v[q * 4 + 3] += b4.w;
                        }
                    }
                    if (!tma_store) {
                        if (row_ok && col0 < p.N) epilogue_direct(p, v, row, col0, vec_ok);
                        continue;
                    }
                    // ---- staged path: everything below touches shared memory only ----
                    bool acted = false;
                    if (two_out) {
                        uint8_t* rowp2 = box2 + r_local * 128;
                        uint32_t pw[16];
                        acted = preact_words(v, pw, act, act_grad);
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const uint32_t j = (uint32_t)(ch * 4 + q);
                            sts128(rowp2 + ((j ^ swz) << 4), make_uint4(pw[q * 4], pw[q * 4 + 1], pw[q * 4 + 2], pw[q * 4 + 3]));
                        }
                    }
                    if (act && !acted) act_inplace(v, act);
                    if (MODE == 4) {
                        // the mask of mgb_dropout_bf16 over the contiguous [M, N] output (dropout.cuh): 16 consecutive
                        // elements per Philox call, so this thread's 32 columns are two calls (N % 16 == 0)
                        const uint64_t g0 = ((uint64_t)row * (uint64_t)p.N + (uint64_t)col0) >> 4;
                        const float ik = p.drop_inv_keep;
#pragma unroll
                        for (int c = 0; c < 2; ++c) {
                            const uint4 rn = dropout_rand16(drop_seed, p.drop_site, g0 + c);
                            const uint32_t w[4] = {rn.x, rn.y, rn.z, rn.w};
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                uint32_t g01, g23;
                                dropout_ge_word(w[i], p.drop_thresh, g01, g23);
                                float* vv = v + 16 * c + 4 * i;
                                vv[0] = (g01 & 0x8000u) ? vv[0] * ik : 0.f;
                                vv[1] = (g01 & 0x80000000u) ? vv[1] * ik : 0.f;
                                vv[2] = (g23 & 0x8000u) ? vv[2] * ik : 0.f;
                                vv[3] = (g23 & 0x80000000u) ? vv[3] * ik : 0.f;
                            }
                        }
                    }
                    if (aux) {
                        // this thread's 64 bytes of the TMA-loaded residual / gate tile (swizzled like the output)
                        uint32_t xw[16];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const uint32_t j = (uint32_t)(ch * 4 + q);
                            const uint4 xi = lds128(rowp + ((j ^ swz) << 4));
                            xw[q * 4 + 0] = xi.x; xw[q * 4 + 1] = xi.y; xw[q * 4 + 2] = xi.z; xw[q * 4 + 3] = xi.w;
                        }
                        if (aux == 2) gate_inplace(v, xw, (MODE == 1 || MODE == 4) ? (int)MGB_GATE_MUL : p.gate_kind);
                        else if (aux == 3) {
                            // rowdot: sum over this 64-column slice of D * X, from the fp32 values before D is rounded
                            float2 dacc = make_float2(0.f, 0.f);
#pragma unroll
                            for (int j = 0; j < 16; ++j) dacc = fma2(make_float2(v[2 * j], v[2 * j + 1]), unpack_bf16x2(xw[j]), dacc);
                            rowdot_acc += dacc.x + dacc.y;
                        }
                        else add_bf16_inplace(v, xw);
                    }
                    if (direct_gate && aux != 2 && row_ok) {
                        const __nv_bfloat16* g = p.gate + row * p.ldg + col0;
                        uint32_t gw[16];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            uint4 gi = make_uint4(0, 0, 0, 0);
                            if (col0 + q * 8 < p.N) gi = ldg128(g + q * 8);
                            gw[q * 4 + 0] = gi.x; gw[q * 4 + 1] = gi.y; gw[q * 4 + 2] = gi.z; gw[q * 4 + 3] = gi.w;
                        }
                        gate_inplace(v, gw, p.gate_kind);
                    }
                    if (direct_res && aux != 1 && row_ok) {
                        const __nv_bfloat16* rs = p.residual + row * p.ldr + col0;
                        uint32_t rw[16];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            uint4 ri = make_uint4(0, 0, 0, 0);
                            if (col0 + q * 8 < p.N) ri = ldg128(rs + q * 8);
                            rw[q * 4 + 0] = ri.x; rw[q * 4 + 1] = ri.y; rw[q * 4 + 2] = ri.z; rw[q * 4 + 3] = ri.w;
                        }
                        add_bf16_inplace(v, rw);
                    }
                    if (COLSUM) {
                        // lane l ends up with the sum over this warp's 32 rows of column ccol + l
                        float t16[16], t8[8], t4[4], t2[2];
                        const bool b16 = lane & 16, b8 = lane & 8, b4 = lane & 4, b2 = lane & 2, b1 = lane & 1;
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float lo = row_ok ? v[i] : 0.f, hi = row_ok ? v[i + 16] : 0.f;
                            t16[i] = (b16 ? hi : lo) + __shfl_xor_sync(0xffffffffu, b16 ? lo : hi, 16);
                        }
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            t8[i] = (b8 ? t16[i + 8] : t16[i]) + __shfl_xor_sync(0xffffffffu, b8 ? t16[i] : t16[i + 8], 8);
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            t4[i] = (b4 ? t8[i + 4] : t8[i]) + __shfl_xor_sync(0xffffffffu, b4 ? t8[i] : t8[i + 4], 4);
#pragma unroll
                        for (int i = 0; i < 2; ++i)
                            t2[i] = (b2 ? t4[i + 2] : t4[i]) + __shfl_xor_sync(0xffffffffu, b2 ? t4[i] : t4[i + 2], 2);
                        const float tot = (b1 ? t2[1] : t2[0]) + __shfl_xor_sync(0xffffffffu, b1 ? t2[0] : t2[1], 1);
                        atomicAdd(&cs_smem[ccol + lane], tot);
                    }
                    if (d_f32) {
#pragma unroll
                        for (int q = 0; q < 8; ++q)
                            sts128f(rowp + (((uint32_t)q ^ swz) << 4), make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]));
                    } else {
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const uint32_t j = (uint32_t)(ch * 4 + q);
                            sts128(rowp + ((j ^ swz) << 4),
                                   make_uint4(pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]), pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]),
                                              pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]), pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7])));
                        }
                    }
                }
                if (aux == 3 && row_ok && n0 + sl * slice_cols < p.N) {
                    // out[(b * H + h) * T + t] for row = b * T + t and head h = this 64-column slice
                    const int64_t bb = row / p.rowdot_T, tt = row - bb * p.rowdot_T;
                    const int64_t hh = (n0 + sl * slice_cols) >> 6;
                    p.rowdot_out[(bb * (p.N >> 6) + hh) * p.rowdot_T + tt] = rowdot_acc;
                }
                if (tma_store) {
                    fence_proxy_async_smem();
                    named_bar_sync(1 + half, 128);
                    if (elected) {
                        const int c0 = n0 + sl * slice_cols;
                        if (c0 < p.N && m0 < p.M) {
                            if (p.accumulate) tma_reduce_add_2d(&tmap_d, box, c0, m0);
                            else tma_store_2d(&tmap_d, box, c0, m0);
                            if (two_out) tma_store_2d(&tmap_p, box2, c0, m0);
                        }
                        bulk_commit();
                    }
                    if (!two_out) buf ^= 1;
                }
            }
            if (!released) {   // this half had no slice in the tile (narrow tiles): still hand TMEM back
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(&tmem_empty[acc], 0);
            }
            if (COLSUM) {      // all eight epilogue warps have added their rows: one global atomic per column of the tile
                named_bar_sync(3, kEpiWarps * 32);
                const int c = (int)threadIdx.x - 128;
                const float tot = cs_smem[c];
                cs_smem[c] = 0.f;
                if (c < block_n && n0 + c < p.N) atomicAdd(p.colsum + n0 + c, tot);
                named_bar_sync(3, kEpiWarps * 32);
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (tma_store && elected) bulk_wait<0>();
    }

    __syncwarp();
    tc_fence_before();
    cluster_sync_all();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc_cg2(tmem_base, kTmemCols);
    }
}

// ----------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------
static int pick_block_n_pair(int64_t M, int64_t N, int b_mn, bool accumulate, int num_clusters) {
    // K-major B: 64/128/192/256 (each CTA stages block_n/2 rows); MN-major B: 64-wide atoms per CTA -> 128/256.
    // Cost = waves x time per tile; narrow tiles move more operand bytes per flop (efficiency factors measured
    // with tools/gemm_probe.py). With split-K available (accumulate) the widest tile always wins.
    if (accumulate) return N > 128 ? 256 : 128;
    const int cands[4] = {256, 192, 128, 64};
    const double eff[4] = {1.0, 0.93, 0.70, 0.45};
    int best = 128;
    double best_cost = 1e30;
    const int64_t mt = ceil_div64(M, 2 * BLOCK_M);
    for (int i = 0; i < 4; ++i) {
        const int bn = cands[i];
        if (b_mn && (bn == 192 || bn == 64)) continue;
        if (bn > 64 && bn - 64 >= N && !(b_mn && bn == 128)) continue;
        const int64_t nt = ceil_div64(N, bn);
        const int64_t waves = ceil_div64(mt * nt, num_clusters);
        const double cost = (double)waves * (bn / eff[i] + 8.0);
        if (cost < best_cost) { best_cost = cost; best = bn; }
    }
    return best;
}

static bool tma_ok(const void* base, int64_t ld, int elem_bytes) {
    return base && ((ld * elem_bytes) % 16 == 0) && ((reinterpret_cast<uintptr_t>(base) & 15) == 0);
}

int launch_gemm_pair(const mgb_gemm_args* a, Params p, int num_sms, cudaStream_t stream) {
    static PerDeviceOnce once;
    once.run([] {
        cudaFuncSetAttribute(gemm_pair_kernel<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
        cudaFuncSetAttribute(gemm_pair_kernel<true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
        cudaFuncSetAttribute(gemm_pair_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
        cudaFuncSetAttribute(gemm_pair_kernel<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
        cudaFuncSetAttribute(gemm_pair_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
        cudaFuncSetAttribute(gemm_pair_kernel<false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
        cudaFuncSetAttribute(gemm_pair_kernel<false, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
    });
    const int num_clusters = num_sms / 2;
    int bn = a->block_n;
    if (bn == 0) bn = pick_block_n_pair(a->m, a->n, p.b_mn, a->accumulate != 0, num_clusters);
    if (bn != 64 && bn != 128 && bn != 192 && bn != 256) { set_last_error("gemm: block_n must be 64/128/192/256"); return kErrBadShape; }
    if (p.b_mn && (bn == 64 || bn == 192)) bn = (bn == 64) ? 128 : 256;
    p.block_n = bn;
    p.num_m_tiles = (int)ceil_div64(a->m, 2 * BLOCK_M);
    p.num_n_tiles = (int)ceil_div64(a->n, bn);
    p.k_blocks_total = (int)ceil_div64(a->k, BLOCK_K);
    const int tiles = p.num_m_tiles * p.num_n_tiles;
    int split = a->split_k;
    if (split <= 0) {
        split = 1;
        if (a->accumulate) {
            const int max_split = p.k_blocks_total / 8 > 0 ? p.k_blocks_total / 8 : 1;
            double best = 1e30;
            for (int s = 1; s <= max_split && s <= 64; ++s) {
                const int64_t units = (int64_t)tiles * s;
                const int64_t waves = ceil_div64(units, num_clusters);
                const int kbs = (int)ceil_div64(p.k_blocks_total, s);
                const double cost = (double)waves * (kbs + 6.0);
                if (cost < best - 1e-9) { best = cost; split = s; }
            }
        }
    }
    if (split > p.k_blocks_total) split = p.k_blocks_total;
    p.k_blocks_per_split = (int)ceil_div64(p.k_blocks_total, split);
    p.split_k = (int)ceil_div64(p.k_blocks_total, p.k_blocks_per_split);
    // Raster order: concurrently running tiles share one operand tile and sweep the other operand, which is
    // then re-read once per tile of the shared side. Sweep the SMALLER operand (it stays in the 126 MB L2):
    // activations x weights -> n fastest; the head (16384 x 50257: weights larger than the activation chunk) -> m fastest.
    p.m_fast = a->m < a->n ? 1 : 0;

    // D (and the saved pre-activation) leave through TMA when their row pitch allows it; one bf16 side input
    // (gate, else residual) is TMA-loaded in place when the output is bf16 and there is a single output.
    p.tma_store = tma_ok(a->d, a->ldd, a->d_f32 ? 4 : 2) && (!a->preact || tma_ok(a->preact, a->ldp, 2)) &&
                  !(a->preact && a->d_f32);
    if (g_opt_gemm_direct_store.load(std::memory_order_relaxed)) p.tma_store = 0;
    p.tma_aux = 0;
    if (p.tma_store && !a->d_f32 && !a->preact) {
        if (a->gate && tma_ok(a->gate, a->ldg, 2)) p.tma_aux = 2;
        else if (a->residual && tma_ok(a->residual, a->ldr, 2)) p.tma_aux = 1;
        else if (a->rowdot_x && tma_ok(a->rowdot_x, a->ld_rowdot, 2)) p.tma_aux = 3;
    }

    if (a->rowdot_x) {
        if (p.tma_aux != 3 || (a->n & 63) || !a->rowdot_out || a->rowdot_t <= 0 || (a->m % a->rowdot_t)) {
            set_last_error("gemm: rowdot needs a TMA-addressable bf16 D and X, no gate / residual / preact, n a multiple of 64, "
                           "m a multiple of rowdot_t");
            return kErrUnsupported;
        }
        p.rowdot_out = a->rowdot_out;
        p.rowdot_T = a->rowdot_t;
    }
    const int nb = bn / 2;
    const uint32_t stage_bytes = BLOCK_M * BLOCK_K * 2 + (uint32_t)nb * BLOCK_K * 2;
    // column sums ride along only on the staged bf16 path; otherwise mgb_gemm_bf16 adds them with a separate pass
    const bool fused_colsum = a->colsum && p.tma_store && !a->d_f32;
    p.colsum = fused_colsum ? a->colsum : nullptr;
    const int fixed = kNumBoxes * kBoxBytes + kBarBytes + ((a->bias && fused_colsum) ? 3072 : 2048);   // barriers | bias x 2 / colsum
    int stages = (kSmemLimit - fixed) / (int)stage_bytes;
    if (stages > kMaxStages) stages = kMaxStages;
    p.num_stages = stages;
    const int smem_bytes = stages * (int)stage_bytes + fixed;

    CUtensorMap ta, tb, td, tp, tx;
    int rc;
    if (!p.a_mn) rc = make_map_2d(&ta, a->a, a->k, a->m, a->lda, BLOCK_K, BLOCK_M);
    else         rc = make_map_2d(&ta, a->a, a->m, a->k, a->lda, 64, BLOCK_K);
    if (rc) return rc;
    if (!p.b_mn) rc = make_map_2d(&tb, a->b, a->k, a->n, a->ldb, BLOCK_K, nb);
    else         rc = make_map_2d(&tb, a->b, a->n, a->k, a->ldb, 64, BLOCK_K);
    if (rc) return rc;
    td = ta; tp = ta; tx = ta;   // placeholders when unused
    if (p.tma_store) {
        rc = make_map_2d(&td, a->d, a->n, a->m, a->ldd, a->d_f32 ? 32 : 64, BLOCK_M, a->d_f32 != 0);
        if (rc) return rc;
        if (a->preact) {
            rc = make_map_2d(&tp, a->preact, a->n, a->m, a->ldp, 64, BLOCK_M);
            if (rc) return rc;
        }
        if (p.tma_aux) {
            rc = p.tma_aux == 3 ? make_map_2d(&tx, a->rowdot_x, a->n, a->m, a->ld_rowdot, 64, BLOCK_M)
               : p.tma_aux == 2 ? make_map_2d(&tx, a->gate, a->n, a->m, a->ldg, 64, BLOCK_M)
                                : make_map_2d(&tx, a->residual, a->n, a->m, a->ldr, 64, BLOCK_M);
            if (rc) return rc;
        }
    }
    const int64_t units = (int64_t)tiles * p.split_k;
    // more units than resident clusters: one cluster per unit, handed out by the hardware work queue (see the kernel)
    p.dynamic = (units > num_clusters && !g_opt_gemm_static.load(std::memory_order_relaxed)) ? 1 : 0;
    const int clusters = p.dynamic ? (int)units : (int)(units < num_clusters ? units : num_clusters);
    // epilogue specialisation (see the kernel): pick the leanest instantiation that covers this call
    int mode = 0;
    if (p.tma_store && !g_opt_gemm_generic.load(std::memory_order_relaxed)) {
        const bool side_ok = (!a->rowdot_x || p.tma_aux == 3) && (!a->gate || (p.tma_aux == 2 && a->gate_kind == MGB_GATE_MUL && !a->residual)) &&
                             (!a->residual || (p.tma_aux == 1 && !a->gate));
        if (a->d_f32 && !a->bias && !p.act && !a->gate && !a->residual && !a->preact && !fused_colsum) mode = 3;
        else if (!a->d_f32 && p.preact && p.act == MGB_ACT_GELU_TANH && p.act_grad && a->bias && !a->gate && !a->residual &&
                 !fused_colsum) mode = 2;
        else if (!a->d_f32 && !p.preact && !p.act && side_ok) mode = (p.drop_thresh && !fused_colsum) ? 4 : 1;
    }
    if (p.drop_thresh && mode != 4) {
        set_last_error("gemm: the dropout epilogue needs bf16 D through TMA, no activation / colsum / preact, and a residual that TMA can load");
        return kErrUnsupported;
    }
    {
        cudaError_t e;
#define MGB_PAIR(CS, MD) launch_pdl(gemm_pair_kernel<CS, MD>, dim3(2 * clusters), dim3(kThreads), (size_t)smem_bytes, stream, ta, tb, td, tp, tx, p)
        if (mode == 3) e = MGB_PAIR(false, 3);
        else if (mode == 2) e = MGB_PAIR(false, 2);
        else if (mode == 4) e = MGB_PAIR(false, 4);
        else if (mode == 1) e = fused_colsum ? MGB_PAIR(true, 1) : MGB_PAIR(false, 1);
        else e = fused_colsum ? MGB_PAIR(true, 0) : MGB_PAIR(false, 0);
#undef MGB_PAIR
        if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); return kErrLaunch; }
    }
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return fused_colsum ? kFusedColsum : kOk;
}

}  // namespace gemm
}  // namespace mgb
