// HBM-bound row-wise kernels of the training step: LayerNorm fwd/bwd, embedding fwd/bwd,
// cross-entropy (in place on a logits chunk). Single pass over the data, 16-byte vector accesses,
// fp32 statistics.
#include <atomic>
#include <mutex>

#include "common.cuh"
#include "mingpt_b200.h"

namespace mgb {

extern std::atomic<int> g_opt_ln_unstaged;   // optim.cu: 1 = LayerNorm backward keeps its rows in registers
extern std::atomic<int> g_opt_ce_unstaged;   // optim.cu: 1 = cross-entropy streams its rows from global memory

// =====================================================================================
// LayerNorm  (nn.LayerNorm: mingpt/model.py:176,178,248)
// =====================================================================================
constexpr int LN_WARPS = 8;

// One warp per row; the row lives in registers (MAXCH 16-byte chunks per lane). gamma / beta are loaded ONCE per warp
// (the grid is sized so that a warp walks over many rows) and all elementwise math runs on fp32 pairs (FFMA2): the
// kernel used to be bound by instruction issue (~420 per row and lane), not by HBM.
template <int MAXCH>
__global__ void __launch_bounds__(LN_WARPS * 32, MAXCH <= 4 ? 2 : 1)
ln_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
              __nv_bfloat16* __restrict__ y, float* __restrict__ mean, float* __restrict__ rstd, int64_t rows,
              int cols, float eps) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nch = cols >> 3;
    const float inv_cols = 1.0f / (float)cols;
    float2 gam[MAXCH][4], bet[MAXCH][4];
#pragma unroll
    for (int i = 0; i < MAXCH; ++i) {
        const int ch = lane + 32 * i;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            gam[i][e] = (ch < nch) ? make_float2(gamma[ch * 8 + 2 * e], gamma[ch * 8 + 2 * e + 1]) : make_float2(0.f, 0.f);
            bet[i][e] = (ch < nch) ? make_float2(beta[ch * 8 + 2 * e], beta[ch * 8 + 2 * e + 1]) : make_float2(0.f, 0.f);
        }
    }
    grid_dep_launch();
    grid_dep_wait();                                 // x is the previous kernel's output (programmatic dependent launch)
    const int64_t stride = (int64_t)gridDim.x * LN_WARPS;
    int64_t row = (int64_t)blockIdx.x * LN_WARPS + warp;
    uint4 nxt[MAXCH];
    if (row < rows) {
#pragma unroll
        for (int i = 0; i < MAXCH; ++i)
            if (lane + 32 * i < nch) nxt[i] = ldg128(x + row * cols + (lane + 32 * i) * 8);
    }
    for (; row < rows; row += stride) {
        uint4 buf[MAXCH];
#pragma unroll
        for (int i = 0; i < MAXCH; ++i) buf[i] = nxt[i];
        if (row + stride < rows) {                   // the next row's loads are in flight during this row's math
#pragma unroll
            for (int i = 0; i < MAXCH; ++i)
                if (lane + 32 * i < nch) nxt[i] = ldg128(x + (row + stride) * cols + (lane + 32 * i) * 8);
        }
        // mean first, then the centred second moment (two passes over registers: as accurate as torch's)
        float2 sv = make_float2(0.f, 0.f);
#pragma unroll
        for (int i = 0; i < MAXCH; ++i) {
            if (lane + 32 * i < nch) {
                const uint32_t w[4] = {buf[i].x, buf[i].y, buf[i].z, buf[i].w};
#pragma unroll
                for (int e = 0; e < 4; ++e) sv = add2(sv, unpack_bf16x2(w[e]));
            }
        }
        const float mu = warp_sum(sv.x + sv.y) * inv_cols;
        const float2 nmu = splat2(-mu);
        float2 qv = make_float2(0.f, 0.f);
#pragma unroll
        for (int i = 0; i < MAXCH; ++i) {
            if (lane + 32 * i < nch) {
                const uint32_t w[4] = {buf[i].x, buf[i].y, buf[i].z, buf[i].w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 c = add2(unpack_bf16x2(w[e]), nmu);
                    qv = fma2(c, c, qv);
                }
            }
        }
        const float rs = rsqrtf(warp_sum(qv.x + qv.y) * inv_cols + eps);
        if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
        const float2 rs2 = splat2(rs), nmr2 = splat2(-mu * rs);
        __nv_bfloat16* yr = y + row * cols;
#pragma unroll
        for (int i = 0; i < MAXCH; ++i) {
            const int ch = lane + 32 * i;
            if (ch < nch) {
                const uint32_t w[4] = {buf[i].x, buf[i].y, buf[i].z, buf[i].w};
                uint32_t o[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 r = fma2(fma2(unpack_bf16x2(w[e]), rs2, nmr2), gam[i][e], bet[i][e]);
                    o[e] = pack_bf16x2(r.x, r.y);
                }
                stg128(yr + ch * 8, make_uint4(o[0], o[1], o[2], o[3]));
            }
        }
    }
}

// any cols (scalar accesses): ragged widths such as C = 50
__global__ void __launch_bounds__(LN_WARPS * 32)
ln_fwd_generic_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
                      const float* __restrict__ beta, __nv_bfloat16* __restrict__ y, float* __restrict__ mean,
                      float* __restrict__ rstd, int64_t rows, int cols, float eps) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int64_t row = (int64_t)blockIdx.x * LN_WARPS + warp; row < rows; row += (int64_t)gridDim.x * LN_WARPS) {
        const __nv_bfloat16* xr = x + row * cols;
        float s = 0.f;
        for (int c = lane; c < cols; c += 32) s += __bfloat162float(xr[c]);
        const float mu = warp_sum(s) / (float)cols;
        float q = 0.f;
        for (int c = lane; c < cols; c += 32) { float d = __bfloat162float(xr[c]) - mu; q += d * d; }
        const float rs = rsqrtf(warp_sum(q) / (float)cols + eps);
        if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
        for (int c = lane; c < cols; c += 32)
            y[row * cols + c] = __float2bfloat16((__bfloat162float(xr[c]) - mu) * rs * gamma[c] + beta[c]);
    }
}

template <int MAXCH>
__global__ void __launch_bounds__(LN_WARPS * 32, MAXCH <= 3 ? 2 : 1)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
              const float* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ rstd,
              const __nv_bfloat16* dres, __nv_bfloat16* dx, float* __restrict__ dgamma, float* __restrict__ dbeta,
              int64_t rows, int cols) {
    extern __shared__ float ln_smem[];   // [2][cols] CTA-level dgamma / dbeta accumulators
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nch = cols >> 3;
    const float inv_cols = 1.0f / (float)cols;
    for (int c = threadIdx.x; c < 2 * cols; c += blockDim.x) ln_smem[c] = 0.f;
    __syncthreads();

    // Register diet (occupancy is what hides the HBM latency here): gamma and the dgamma/dbeta partial sums
    // stay in registers, x and dy stay PACKED (bf16) between the two passes and xhat is recomputed.
    float gam[MAXCH][8], dg[MAXCH][8], db[MAXCH][8];
#pragma unroll
    for (int i = 0; i < MAXCH; ++i) {
        const int ch = lane + 32 * i;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            gam[i][e] = (ch < nch) ? gamma[ch * 8 + e] : 0.f;
            dg[i][e] = 0.f;
            db[i][e] = 0.f;
        }
    }
    for (int64_t row = (int64_t)blockIdx.x * LN_WARPS + warp; row < rows; row += (int64_t)gridDim.x * LN_WARPS) {
        const float mu = mean[row], rs = rstd[row];
        uint4 px[MAXCH], pd[MAXCH], pr[MAXCH];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < MAXCH; ++i) {      // all three input streams in flight before the first reduction
            const int ch = lane + 32 * i;
            if (ch < nch) {
                px[i] = ldg128(x + row * cols + ch * 8);
                pd[i] = ldg128(dy + row * cols + ch * 8);
                pr[i] = dres ? ldg128(dres + row * cols + ch * 8) : make_uint4(0, 0, 0, 0);
            }
        }
#pragma unroll
        for (int i = 0; i < MAXCH; ++i) {
            const int ch = lane + 32 * i;
            if (ch < nch) {
                const uint32_t wx[4] = {px[i].x, px[i].y, px[i].z, px[i].w};
                const uint32_t wd[4] = {pd[i].x, pd[i].y, pd[i].z, pd[i].w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 fx = unpack_bf16x2(wx[e]);
                    const float2 fd = unpack_bf16x2(wd[e]);
                    const float h0 = (fx.x - mu) * rs, h1 = (fx.y - mu) * rs;
                    dg[i][2 * e] += fd.x * h0; dg[i][2 * e + 1] += fd.y * h1;
                    db[i][2 * e] += fd.x;      db[i][2 * e + 1] += fd.y;
                    const float g0 = fd.x * gam[i][2 * e], g1 = fd.y * gam[i][2 * e + 1];
                    s1 += g0 + g1;
                    s2 += g0 * h0 + g1 * h1;
                }
            }
        }
        const float m1 = warp_sum(s1) * inv_cols;
        const float m2 = warp_sum(s2) * inv_cols;
#pragma unroll
        for (int i = 0; i < MAXCH; ++i) {
            const int ch = lane + 32 * i;
            if (ch < nch) {
                const uint32_t wx[4] = {px[i].x, px[i].y, px[i].z, px[i].w};
                const uint32_t wd[4] = {pd[i].x, pd[i].y, pd[i].z, pd[i].w};
                float r[8];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 fx = unpack_bf16x2(wx[e]);
                    const float2 fd = unpack_bf16x2(wd[e]);
                    const float h0 = (fx.x - mu) * rs, h1 = (fx.y - mu) * rs;
                    r[2 * e] = rs * (fd.x * gam[i][2 * e] - m1 - h0 * m2);
                    r[2 * e + 1] = rs * (fd.y * gam[i][2 * e + 1] - m1 - h1 * m2);
                }
                {
                    const uint32_t wr[4] = {pr[i].x, pr[i].y, pr[i].z, pr[i].w};   // zeros when there is no dres
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 f = unpack_bf16x2(wr[e]);
                        r[2 * e] += f.x; r[2 * e + 1] += f.y;
                    }
                }
                stg128(dx + row * cols + ch * 8,
                       make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]),
                                  pack_bf16x2(r[6], r[7])));
            }
        }
    }
#pragma unroll
    for (int i = 0; i < MAXCH; ++i) {
        const int ch = lane + 32 * i;
        if (ch < nch) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                atomicAdd(&ln_smem[ch * 8 + e], dg[i][e]);
                atomicAdd(&ln_smem[cols + ch * 8 + e], db[i][e]);
            }
        }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < cols; c += blockDim.x) {
        atomicAdd(dgamma + c, ln_smem[c]);
        atomicAdd(dbeta + c, ln_smem[cols + c]);
    }
}

// Same computation with the three input rows STAGED THROUGH SHARED MEMORY by 1-D bulk copies (the TMA engine's
// non-tensor mode): each warp owns a contiguous range of rows and a ring of LNB_STAGES row slots, lane 0 keeps
// the next rows in flight while the warp works on the current one, so the loads no longer occupy registers and
// the number of bytes in flight per SM is set by shared memory (2 CTAs x 8 warps x 2 rows x 4.5 KiB at C = 768),
// not by occupancy. The row is read back from shared memory in both passes (conflict-free 16-byte accesses).
constexpr int LNB_STAGES = 4;      // x and dy rows (the two that are read twice) per slot; dres comes through registers
// COLSUM: additionally dres_colsum[c] += sum over rows of dres[row, c] -- the bias gradient of the linear layer whose
// output gradient `dres` is (it is being read anyway).
template <int MAXCH, bool COLSUM>
__global__ void __launch_bounds__(LN_WARPS * 32, MAXCH <= 3 ? 2 : 1)
ln_bwd_staged_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const float* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ rstd,
                     const __nv_bfloat16* dres, __nv_bfloat16* dx, float* __restrict__ dgamma, float* __restrict__ dbeta,
                     float* __restrict__ dres_colsum, int64_t rows, int cols, int64_t rows_per_warp) {
    extern __shared__ __align__(128) uint8_t lnb_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nch = cols >> 3;
    const uint32_t rowbytes = (uint32_t)cols * 2;
    const float inv_cols = 1.0f / (float)cols;
    constexpr int NACC = COLSUM ? 4 : 3;
    static_assert(LNB_STAGES * 2 * 2 >= (NACC - 1) * 4, "the per-warp ring must hold the warp's column partials");
    float* gam_s = reinterpret_cast<float*>(lnb_smem);                      // [cols] gamma (read from here: registers are short)
    uint8_t* ring = lnb_smem + (size_t)cols * sizeof(float) + (size_t)warp * LNB_STAGES * 2 * rowbytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(lnb_smem + (size_t)cols * sizeof(float) +
                                                 (size_t)LN_WARPS * LNB_STAGES * 2 * rowbytes) + warp * LNB_STAGES;
    for (int c = threadIdx.x; c < cols; c += blockDim.x) gam_s[c] = gamma[c];
    if (lane == 0) {
        for (int s = 0; s < LNB_STAGES; ++s) mbar_init(&bars[s], 1);
        fence_mbar_init();
    }
    __syncthreads();
    grid_dep_launch();
    grid_dep_wait();                                 // inputs come from the previous kernels (programmatic dependent launch)

    // Rows are dealt to the warps INTERLEAVED (warp w of the grid takes rows w, w + W, w + 2W, ...): the rows in flight
    // across the chip at any moment are then neighbours in memory, which keeps the HBM pages open; contiguous per-warp
    // ranges put 2368 streams 21 KB apart.
    const int64_t gw = (int64_t)blockIdx.x * LN_WARPS + warp;
    const int64_t stride = (int64_t)gridDim.x * LN_WARPS;
    const int64_t row0 = gw;
    const int64_t row1 = rows;
    (void)rows_per_warp;
    const uint32_t tx = 2u * rowbytes;
    auto issue = [&](int64_t row, int s) {       // lane 0 only
        uint8_t* slot = ring + (size_t)s * 2 * rowbytes;
        mbar_arrive_expect_tx(&bars[s], tx);
        bulk_load_1d(slot, x + row * cols, rowbytes, &bars[s]);
        bulk_load_1d(slot + rowbytes, dy + row * cols, rowbytes, &bars[s]);
    };
    // dres is read once per row: it travels through registers, fetched one row ahead (the ring then holds twice as
    // many rows in flight for the same shared memory)
    auto fetch_dres = [&](int64_t row, uint4 (&dst)[MAXCH]) {
#pragma unroll
        for (int i = 0; i < MAXCH; ++i)
            if (lane + 32 * i < nch) dst[i] = __ldcs(reinterpret_cast<const uint4*>(dres + row * cols + (lane + 32 * i) * 8));
    };
    if (lane == 0)
        for (int s = 0; s < LNB_STAGES; ++s)
            if (row0 + s * stride < row1) issue(row0 + s * stride, s);

    // All elementwise math on fp32 PAIRS (FFMA2 / FADD2 / FMUL2): the kernel was bound by instruction issue (~21
    // instructions per element), not by the fp32 pipes or by HBM; packed it needs ~11.
    float2 dg[MAXCH][4], db[MAXCH][4], dr[COLSUM ? MAXCH : 1][4];
#pragma unroll
    for (int i = 0; i < MAXCH; ++i) {
        const int ch = lane + 32 * i;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            dg[i][e] = make_float2(0.f, 0.f);
            db[i][e] = make_float2(0.f, 0.f);
            if (COLSUM) dr[i][e] = make_float2(0.f, 0.f);
        }
    }
    float mu_n = 0.f, rs_n = 0.f;
    uint4 dr_n[MAXCH];
    if (row0 < row1) {
        mu_n = mean[row0]; rs_n = rstd[row0];
        if (dres) fetch_dres(row0, dr_n);
    }
    int s = 0;
    uint32_t ph = 0;
    for (int64_t row = row0; row < row1; row += stride) {
        const float mu = mu_n, rs = rs_n;
        if (row + stride < row1) { mu_n = mean[row + stride]; rs_n = rstd[row + stride]; }
        const float2 rs2 = splat2(rs), nmr2 = splat2(-mu * rs);      // xhat = x * rs - mu * rs
        const uint8_t* xs = ring + (size_t)s * 2 * rowbytes;
        const uint8_t* ds = xs + rowbytes;
        mbar_wait_relaxed(&bars[s], ph);
        float2 s1 = make_float2(0.f, 0.f), s2 = make_float2(0.f, 0.f);
#pragma unroll
        for (int i = 0; i < MAXCH; ++i) {
            const int ch = lane + 32 * i;
            if (ch < nch) {
                const uint4 px = lds128(xs + ch * 16), pd = lds128(ds + ch * 16);
                const uint32_t wx[4] = {px.x, px.y, px.z, px.w};
                const uint32_t wd[4] = {pd.x, pd.y, pd.z, pd.w};
                const float4 ga = lds128f(gam_s + ch * 8), gb = lds128f(gam_s + ch * 8 + 4);
                const float2 gam[4] = {make_float2(ga.x, ga.y), make_float2(ga.z, ga.w), make_float2(gb.x, gb.y), make_float2(gb.z, gb.w)};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 h = fma2(unpack_bf16x2(wx[e]), rs2, nmr2);
                    const float2 fd = unpack_bf16x2(wd[e]);
                    dg[i][e] = fma2(fd, h, dg[i][e]);
                    db[i][e] = add2(db[i][e], fd);
                    const float2 g = mul2(fd, gam[e]);
                    s1 = add2(s1, g);
                    s2 = fma2(g, h, s2);
                }
            }
        }
        const float m1 = warp_sum(s1.x + s1.y) * inv_cols;
        const float m2 = warp_sum(s2.x + s2.y) * inv_cols;
        const float2 nm1rs = splat2(-m1 * rs), nm2rs = splat2(-m2 * rs);   // dx = rs * (g - m1 - xhat * m2)
#pragma unroll
        for (int i = 0; i < MAXCH; ++i) {
            const int ch = lane + 32 * i;
            if (ch < nch) {
                const uint4 px = lds128(xs + ch * 16), pd = lds128(ds + ch * 16);
                const uint32_t wx[4] = {px.x, px.y, px.z, px.w};
                const uint32_t wd[4] = {pd.x, pd.y, pd.z, pd.w};
                const float4 ga = lds128f(gam_s + ch * 8), gb = lds128f(gam_s + ch * 8 + 4);
                const float2 gam[4] = {make_float2(ga.x, ga.y), make_float2(ga.z, ga.w), make_float2(gb.x, gb.y), make_float2(gb.z, gb.w)};
                float2 r[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 h = fma2(unpack_bf16x2(wx[e]), rs2, nmr2);
                    const float2 g = mul2(unpack_bf16x2(wd[e]), gam[e]);
                    r[e] = fma2(h, nm2rs, fma2(g, rs2, nm1rs));
                }
                if (dres) {
                    const uint4 pr = dr_n[i];
                    const uint32_t wr[4] = {pr.x, pr.y, pr.z, pr.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 f = unpack_bf16x2(wr[e]);
                        r[e] = add2(r[e], f);
                        if (COLSUM) dr[i][e] = add2(dr[i][e], f);
                    }
                }
                stg128(dx + row * cols + ch * 8,
                       make_uint4(pack_bf16x2(r[0].x, r[0].y), pack_bf16x2(r[1].x, r[1].y), pack_bf16x2(r[2].x, r[2].y),
                                  pack_bf16x2(r[3].x, r[3].y)));
            }
        }
        // next row's dres into the same registers: a whole row's wait + first pass ahead of its use
        if (dres && row + stride < row1) fetch_dres(row + stride, dr_n);
        __syncwarp();                                            // every lane is done with this slot
        if (lane == 0 && row + LNB_STAGES * stride < row1) issue(row + LNB_STAGES * stride, s);
        if (++s == LNB_STAGES) { s = 0; ph ^= 1; }
    }
    // Column sums across the CTA: each warp has ~rows/(8*grid) rows (a handful), so this tail is a visible share of the
    // kernel. The warp parks its partials in its own (now idle) ring slots with plain 16-byte stores, and after one
    // barrier each thread adds the eight partials of its columns and sends ONE atomic per column to global memory.
    float* part = reinterpret_cast<float*>(ring);                          // [NACC - 1][cols] per warp, <= the ring's 16 * cols bytes
#pragma unroll
    for (int i = 0; i < MAXCH; ++i) {
        const int ch = lane + 32 * i;
        if (ch < nch) {
            float4* pg = reinterpret_cast<float4*>(part + ch * 8);
            pg[0] = make_float4(dg[i][0].x, dg[i][0].y, dg[i][1].x, dg[i][1].y);
            pg[1] = make_float4(dg[i][2].x, dg[i][2].y, dg[i][3].x, dg[i][3].y);
            float4* pb = reinterpret_cast<float4*>(part + cols + ch * 8);
            pb[0] = make_float4(db[i][0].x, db[i][0].y, db[i][1].x, db[i][1].y);
            pb[1] = make_float4(db[i][2].x, db[i][2].y, db[i][3].x, db[i][3].y);
            if (COLSUM) {
                float4* pr = reinterpret_cast<float4*>(part + 2 * cols + ch * 8);
                pr[0] = make_float4(dr[i][0].x, dr[i][0].y, dr[i][1].x, dr[i][1].y);
                pr[1] = make_float4(dr[i][2].x, dr[i][2].y, dr[i][3].x, dr[i][3].y);
            }
        }
    }
    __syncthreads();
    const size_t warp_stride = (size_t)LNB_STAGES * 2 * rowbytes / sizeof(float);
    const float* part0 = reinterpret_cast<const float*>(lnb_smem + (size_t)cols * sizeof(float));
    for (int c = threadIdx.x; c < (NACC - 1) * cols; c += blockDim.x) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < LN_WARPS; ++w) t += part0[w * warp_stride + c];
        float* dst = c < cols ? dgamma + c : (c < 2 * cols ? dbeta + (c - cols) : dres_colsum + (c - 2 * cols));
        atomicAdd(dst, t);
    }
}

__global__ void __launch_bounds__(LN_WARPS * 32)
ln_bwd_generic_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                      const float* __restrict__ gamma, const float* __restrict__ mean,
                      const float* __restrict__ rstd, const __nv_bfloat16* dres, __nv_bfloat16* dx,
                      float* __restrict__ dgamma, float* __restrict__ dbeta, int64_t rows, int cols) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int64_t row = (int64_t)blockIdx.x * LN_WARPS + warp; row < rows; row += (int64_t)gridDim.x * LN_WARPS) {
        const float mu = mean[row], rs = rstd[row];
        float s1 = 0.f, s2 = 0.f;
        for (int c = lane; c < cols; c += 32) {
            const float h = (__bfloat162float(x[row * cols + c]) - mu) * rs;
            const float g = __bfloat162float(dy[row * cols + c]) * gamma[c];
            s1 += g; s2 += g * h;
        }
        const float m1 = warp_sum(s1) / (float)cols, m2 = warp_sum(s2) / (float)cols;
        for (int c = lane; c < cols; c += 32) {
            const float h = (__bfloat162float(x[row * cols + c]) - mu) * rs;
            const float d = __bfloat162float(dy[row * cols + c]);
            float r = rs * (d * gamma[c] - m1 - h * m2);
            if (dres) r += __bfloat162float(dres[row * cols + c]);
            dx[row * cols + c] = __float2bfloat16(r);
            atomicAdd(dgamma + c, d * h);
            atomicAdd(dbeta + c, d);
        }
    }
}

// =====================================================================================
// Embedding  (GPTEmbedding.forward, mingpt/model.py:222-231)
// =====================================================================================
__global__ void __launch_bounds__(256)
embedding_fwd_kernel(const int64_t* __restrict__ idx, const __nv_bfloat16* __restrict__ wte,
                     const __nv_bfloat16* __restrict__ wpe, __nv_bfloat16* __restrict__ out, int64_t BT, int64_t T,
                     int C, int64_t vocab, int32_t* err_flag) {
    const int nch = C >> 3;
    const int64_t total = BT * nch;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / nch;
        const int ch = (int)(i - row * nch);
        const int64_t t = row % T;
        const int64_t tok = idx[row];
        uint4 o = make_uint4(0, 0, 0, 0);
        if (tok >= 0 && tok < vocab) {
            const uint4 a = ldg128(wte + tok * C + ch * 8);
            const uint4 b = ldg128(wpe + t * C + ch * 8);
            const uint32_t wa[4] = {a.x, a.y, a.z, a.w}, wb[4] = {b.x, b.y, b.z, b.w};
            uint32_t r[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 fa = unpack_bf16x2(wa[e]), fb = unpack_bf16x2(wb[e]);
                r[e] = pack_bf16x2(fa.x + fb.x, fa.y + fb.y);
            }
            o = make_uint4(r[0], r[1], r[2], r[3]);
        } else if (err_flag && ch == 0) {
            atomicExch(err_flag, 1);
        }
        stg128(out + row * C + ch * 8, o);
    }
}

__global__ void __launch_bounds__(256)
embedding_fwd_scalar_kernel(const int64_t* __restrict__ idx, const __nv_bfloat16* __restrict__ wte,
                            const __nv_bfloat16* __restrict__ wpe, __nv_bfloat16* __restrict__ out, int64_t BT,
                            int64_t T, int C, int64_t vocab, int32_t* err_flag) {
    const int64_t total = BT * C;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / C;
        const int c = (int)(i - row * C);
        const int64_t tok = idx[row];
        float v = 0.f;
        if (tok >= 0 && tok < vocab) v = __bfloat162float(wte[tok * C + c]) + __bfloat162float(wpe[(row % T) * C + c]);
        else if (err_flag && c == 0) atomicExch(err_flag, 1);
        out[i] = __float2bfloat16(v);
    }
}

// dwte: one thread per (row, 4 columns), fp32 vector atomics into the touched table row.
__global__ void __launch_bounds__(256)
embedding_bwd_wte_kernel(const int64_t* __restrict__ idx, const __nv_bfloat16* __restrict__ dx,
                         float* __restrict__ dwte, int64_t BT, int C, int64_t vocab) {
    const int64_t total = BT * C;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / C;
        const int c = (int)(i - row * C);
        const int64_t tok = idx[row];
        if (tok >= 0 && tok < vocab) atomicAdd(dwte + tok * C + c, __bfloat162float(dx[i]));
    }
}
// dwpe[t, c] += sum_b dx[b, t, c]: no atomics needed inside one launch, but the buffer accumulates
__global__ void __launch_bounds__(256)
embedding_bwd_wpe_kernel(const __nv_bfloat16* __restrict__ dx, float* __restrict__ dwpe, int64_t B, int64_t T,
                         int C) {
    const int64_t total = T * C;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int64_t b = 0; b < B; ++b) s += __bfloat162float(dx[b * total + i]);
        dwpe[i] += s;
    }
}

// =====================================================================================
// Cross entropy, in place  (F.cross_entropy ignore_index=-1: mingpt/model.py:316-318)
// =====================================================================================
constexpr int CE_THREADS = 512;

__device__ __forceinline__ float block_reduce_max(float v, float* sh) {
    v = warp_max(v);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : -INFINITY;
    if (threadIdx.x < 32) { r = warp_max(r); if (threadIdx.x == 0) sh[0] = r; }
    __syncthreads();
    r = sh[0];
    __syncthreads();
    return r;
}
__device__ __forceinline__ float block_reduce_sum(float v, float* sh) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0.f;
    if (threadIdx.x < 32) { r = warp_sum(r); if (threadIdx.x == 0) sh[0] = r; }
    __syncthreads();
    r = sh[0];
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(CE_THREADS)
cross_entropy_kernel(__nv_bfloat16* logits, int64_t ld, const int64_t* __restrict__ targets,
                     float* __restrict__ row_loss, const int32_t* __restrict__ n_valid, int64_t vocab,
                     float grad_scale, int write_grad, int32_t* err_flag) {
    __shared__ float sh[32];
    const int64_t row = blockIdx.x;
    __nv_bfloat16* lr = logits + row * ld;
    int64_t tgt = targets[row];
    bool ignored = (tgt == -1);
    if (!ignored && (tgt < 0 || tgt >= vocab)) {
        if (err_flag && threadIdx.x == 0) atomicExch(err_flag, 2);
        ignored = true;
    }
    const bool vec = ((ld & 7) == 0) && ((reinterpret_cast<uintptr_t>(logits) & 15) == 0);
    const int64_t vmain = vec ? (vocab & ~int64_t(7)) : 0;
    if (ignored) {
        if (threadIdx.x == 0) row_loss[row] = 0.f;
        if (write_grad) {
            for (int64_t c = (int64_t)threadIdx.x * 8; c < vmain; c += (int64_t)blockDim.x * 8)
                stg128(lr + c, make_uint4(0, 0, 0, 0));
            for (int64_t c = vmain + threadIdx.x; c < ld; c += blockDim.x) lr[c] = __float2bfloat16(0.f);
        }
        return;
    }
    // pass 1: online max / sum-exp
    float m = -INFINITY, s = 0.f;
    for (int64_t c = (int64_t)threadIdx.x * 8; c < vmain; c += (int64_t)blockDim.x * 8) {
        const uint4 v = ldg128(lr + c);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        float f[8];
#pragma unroll
        for (int e = 0; e < 4; ++e) { float2 t = unpack_bf16x2(w[e]); f[2 * e] = t.x; f[2 * e + 1] = t.y; }
        float lm = f[0];
#pragma unroll
        for (int e = 1; e < 8; ++e) lm = fmaxf(lm, f[e]);
        if (lm > m) { s *= __expf(m - lm); m = lm; }
#pragma unroll
        for (int e = 0; e < 8; ++e) s += __expf(f[e] - m);
    }
    for (int64_t c = vmain + threadIdx.x; c < vocab; c += blockDim.x) {
        const float f = __bfloat162float(lr[c]);
        if (f > m) { s *= __expf(m - f); m = f; }
        s += __expf(f - m);
    }
    const float gm = block_reduce_max(m, sh);
    s = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
    const float gs = block_reduce_sum(s, sh);
    const float lse = gm + logf(gs);
    if (threadIdx.x == 0) row_loss[row] = lse - __bfloat162float(lr[tgt]);
    if (!write_grad) return;
    __syncthreads();   // the target logit was read above; now the row may be overwritten
    const int nv = max(*n_valid, 1);
    const float sc = grad_scale / (float)nv;
    for (int64_t c = (int64_t)threadIdx.x * 8; c < vmain; c += (int64_t)blockDim.x * 8) {
        const uint4 v = ldg128(lr + c);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        float f[8];
#pragma unroll
        for (int e = 0; e < 4; ++e) { float2 t = unpack_bf16x2(w[e]); f[2 * e] = t.x; f[2 * e + 1] = t.y; }
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = (__expf(f[e] - lse) - ((c + e) == tgt ? 1.f : 0.f)) * sc;
        stg128(lr + c, make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]),
                                  pack_bf16x2(f[6], f[7])));
    }
    for (int64_t c = vmain + threadIdx.x; c < ld; c += blockDim.x) {
        float g = 0.f;
        if (c < vocab) g = (__expf(__bfloat162float(lr[c]) - lse) - (c == tgt ? 1.f : 0.f)) * sc;
        lr[c] = __float2bfloat16(g);
    }
}

// The same loss / gradient with the logits row STAGED IN SHARED MEMORY: a persistent CTA per SM, two row buffers,
// a producer warp that bulk-copies rows in (cp.async.bulk, one 100 KB transfer per row) and the finished
// gradient rows out, 16 compute warps that make three passes over the on-chip row: maximum (packed bf16x2);
// e = exp(x - max) summed in fp32 and written back as bf16; (e / sum - onehot) * scale. One exponential per
// element instead of two, DRAM sees each row exactly once in each direction, and the loads / stores run ahead
// of the arithmetic.
constexpr int CES_WARPS = 16;                       // compute warps; warp CES_WARPS is the producer
__global__ void __launch_bounds__((CES_WARPS + 1) * 32, 1)
cross_entropy_staged_kernel(__nv_bfloat16* logits, int64_t ld, const int64_t* __restrict__ targets,
                            float* __restrict__ row_loss, const int32_t* __restrict__ n_valid, int64_t rows,
                            int64_t vocab, float grad_scale, int32_t* err_flag) {
    extern __shared__ __align__(128) uint8_t ces_smem[];
    const uint32_t rowbytes = (uint32_t)ld * 2;
    __nv_bfloat16* bufs[2] = {reinterpret_cast<__nv_bfloat16*>(ces_smem),
                              reinterpret_cast<__nv_bfloat16*>(ces_smem + rowbytes)};
    uint64_t* full = reinterpret_cast<uint64_t*>(ces_smem + 2 * (size_t)rowbytes);   // [2] row landed
    uint64_t* done = full + 2;                                                       // [2] gradient row complete
    float* sh = reinterpret_cast<float*>(done + 2);                                  // [32] block reductions
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int b = 0; b < 2; ++b) { mbar_init(&full[b], 1); mbar_init(&done[b], 1); }
        fence_mbar_init();
    }
    __syncthreads();
    const int64_t nmine = (rows - blockIdx.x + gridDim.x - 1) / gridDim.x;           // rows of this CTA

    if (warp == CES_WARPS) {
        // ===================== producer: rows in, gradient rows out =====================
        if (lane == 0) {
            for (int64_t k = 0; k < nmine && k < 2; ++k) {
                mbar_arrive_expect_tx(&full[k], rowbytes);
                bulk_load_1d(bufs[k], logits + (blockIdx.x + k * gridDim.x) * ld, rowbytes, &full[k]);
            }
            for (int64_t k = 0; k < nmine; ++k) {
                const int b = (int)(k & 1);
                mbar_wait_relaxed(&done[b], (uint32_t)((k >> 1) & 1));
                bulk_store_1d(logits + (blockIdx.x + k * gridDim.x) * ld, bufs[b], rowbytes);
                bulk_commit();
                if (k + 2 < nmine) {
                    bulk_wait_read<0>();                                  // the store has left the buffer
                    mbar_arrive_expect_tx(&full[b], rowbytes);
                    bulk_load_1d(bufs[b], logits + (blockIdx.x + (k + 2) * gridDim.x) * ld, rowbytes, &full[b]);
                }
            }
            bulk_wait<0>();
        }
        return;
    }

    // ===================== compute warps =====================
    const int tid = threadIdx.x;                       // 0 .. 511
    constexpr int NT = CES_WARPS * 32;
    constexpr float LOG2E_F = 1.4426950408889634f;
    const int64_t vmain = vocab & ~int64_t(7);
    const int nv = max(*n_valid, 1);
    const float sc = grad_scale / (float)nv;
    auto reduce = [&](float v, bool is_max) {          // over the 512 compute threads (named barrier 1)
        v = is_max ? warp_max(v) : warp_sum(v);
        if (lane == 0) sh[warp] = v;
        named_bar_sync(1, NT);
        float r = (lane < CES_WARPS) ? sh[lane] : (is_max ? -INFINITY : 0.f);
        r = is_max ? warp_max(r) : warp_sum(r);
        named_bar_sync(1, NT);
        return r;
    };
    for (int64_t k = 0; k < nmine; ++k) {
        const int b = (int)(k & 1);
        const int64_t row = blockIdx.x + k * gridDim.x;
        __nv_bfloat16* lr = bufs[b];
        int64_t tgt = targets[row];
        bool ignored = (tgt == -1);
        if (!ignored && (tgt < 0 || tgt >= vocab)) {
            if (err_flag && tid == 0) atomicExch(err_flag, 2);
            ignored = true;
        }
        mbar_wait_relaxed(&full[b], (uint32_t)((k >> 1) & 1));
        if (ignored) {
            if (tid == 0) row_loss[row] = 0.f;
            for (int64_t c = (int64_t)tid * 8; c < ld; c += NT * 8) sts128(lr + c, make_uint4(0, 0, 0, 0));
        } else {
            const float tgt_logit = __bfloat162float(lr[tgt]);
            // pass A: maximum (packed bf16x2: the maximum of bf16 values is exact in bf16)
            __nv_bfloat162 m2 = __float2bfloat162_rn(-INFINITY);
            for (int64_t c = (int64_t)tid * 8; c < vmain; c += NT * 8) {
                const uint4 v = lds128(lr + c);
                const __nv_bfloat162 a = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&v.x), *reinterpret_cast<const __nv_bfloat162*>(&v.y));
                const __nv_bfloat162 b2 = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&v.z), *reinterpret_cast<const __nv_bfloat162*>(&v.w));
                m2 = __hmax2(m2, __hmax2(a, b2));
            }
            float m = fmaxf(__low2float(m2), __high2float(m2));
            for (int64_t c = vmain + tid; c < vocab; c += NT) m = fmaxf(m, __bfloat162float(lr[c]));
            const float gm = reduce(m, true);          // (its barriers also order the tgt_logit read before pass B)
            // pass B: e = exp(x - max) = exp2(x * log2e - max * log2e): fp32 sum, bf16 copy in place
            const float gml = -gm * LOG2E_F;
            float ssum = 0.f;
            for (int64_t c = (int64_t)tid * 8; c < vmain; c += NT * 8) {
                const uint4 v = lds128(lr + c);
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
                uint32_t o[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 t = unpack_bf16x2(w[e]);
                    const float e0 = exp2f(fmaf(t.x, LOG2E_F, gml)), e1 = exp2f(fmaf(t.y, LOG2E_F, gml));
                    ssum += e0 + e1;
                    o[e] = pack_bf16x2(e0, e1);
                }
                sts128(lr + c, make_uint4(o[0], o[1], o[2], o[3]));
            }
            for (int64_t c = vmain + tid; c < ld; c += NT) {
                float e0 = 0.f;
                if (c < vocab) { e0 = exp2f(fmaf(__bfloat162float(lr[c]), LOG2E_F, gml)); ssum += e0; }
                lr[c] = __float2bfloat16(e0);
            }
            const float gs = reduce(ssum, false);
            if (tid == 0) row_loss[row] = gm + logf(gs) - tgt_logit;
            // pass C: gradient. Each thread revisits exactly the elements it wrote in pass B.
            const float inv = sc / gs;
            for (int64_t c = (int64_t)tid * 8; c < vmain; c += NT * 8) {
                const uint4 v = lds128(lr + c);
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
                float f[8];
#pragma unroll
                for (int e = 0; e < 4; ++e) { const float2 t = unpack_bf16x2(w[e]); f[2 * e] = t.x * inv; f[2 * e + 1] = t.y * inv; }
                if (tgt >= c && tgt < c + 8) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) if (c + e == tgt) f[e] -= sc;
                }
                sts128(lr + c, make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]),
                                          pack_bf16x2(f[6], f[7])));
            }
            for (int64_t c = vmain + tid; c < ld; c += NT) {
                float g = 0.f;
                if (c < vocab) g = __bfloat162float(lr[c]) * inv - (c == tgt ? sc : 0.f);
                lr[c] = __float2bfloat16(g);
            }
        }
        fence_proxy_async_smem();                      // the row goes out through the async proxy
        named_bar_sync(1, NT);
        if (tid == 0) mbar_arrive(&done[b]);
    }
}

__global__ void __launch_bounds__(256)
count_valid_kernel(const int64_t* __restrict__ targets, int64_t n, int32_t* out) {
    __shared__ int sh[8];
    int c = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        c += (targets[i] != -1) ? 1 : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < 8; ++w) t += sh[w];
        atomicAdd(out, t);
    }
}

__global__ void __launch_bounds__(1024)
loss_finalize_kernel(const float* __restrict__ row_loss, int64_t rows, const int32_t* __restrict__ n_valid,
                     float* loss) {
    __shared__ float sh[32];
    float s = 0.f;
    for (int64_t i = threadIdx.x; i < rows; i += blockDim.x) s += row_loss[i];
    const float t = block_reduce_sum(s, sh);
    if (threadIdx.x == 0) *loss = t / (float)max(*n_valid, 1);
}

}  // namespace mgb

using namespace mgb;

static int grid_for(int64_t work_items, int per_block, int max_blocks) {
    int64_t g = (work_items + per_block - 1) / per_block;
    if (g < 1) g = 1;
    if (g > max_blocks) g = max_blocks;
    return (int)g;
}

extern "C" int mgb_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean,
                                 float* rstd, int64_t rows, int32_t cols, float eps, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (rows < 0 || cols <= 0) { set_last_error("layernorm_fwd: bad shape"); return kErrBadShape; }
    if (rows == 0) return kOk;
    // two resident CTAs per SM, each warp walking over several rows (gamma / beta stay in its registers)
    const int grid = grid_for(rows, LN_WARPS * 4, num_sms() * 2);
    auto X = (const __nv_bfloat16*)x; auto Y = (__nv_bfloat16*)y;
    const bool vec = (cols % 8 == 0) && cols <= 2048 && ((uintptr_t)x % 16 == 0) && ((uintptr_t)y % 16 == 0) &&
                     ((uintptr_t)gamma % 16 == 0) && ((uintptr_t)beta % 16 == 0);
    if (!vec) ln_fwd_generic_kernel<<<grid, LN_WARPS * 32, 0, s>>>(X, gamma, beta, Y, mean, rstd, rows, cols, eps);
    else if (cols <= 256) launch_pdl(ln_fwd_kernel<1>, dim3(grid), dim3(LN_WARPS * 32), 0, s, X, gamma, beta, Y, mean, rstd, rows, cols, eps);
    else if (cols <= 512) launch_pdl(ln_fwd_kernel<2>, dim3(grid), dim3(LN_WARPS * 32), 0, s, X, gamma, beta, Y, mean, rstd, rows, cols, eps);
    else if (cols <= 768) launch_pdl(ln_fwd_kernel<3>, dim3(grid), dim3(LN_WARPS * 32), 0, s, X, gamma, beta, Y, mean, rstd, rows, cols, eps);
    else if (cols <= 1024) launch_pdl(ln_fwd_kernel<4>, dim3(grid), dim3(LN_WARPS * 32), 0, s, X, gamma, beta, Y, mean, rstd, rows, cols, eps);
    else launch_pdl(ln_fwd_kernel<8>, dim3(grid), dim3(LN_WARPS * 32), 0, s, X, gamma, beta, Y, mean, rstd, rows, cols, eps);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_colsum_bf16(const void* x, int64_t ld, int64_t rows, int64_t cols, float* out, mgb_stream_t s);

static int layernorm_bwd_impl(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd,
                              const void* dres, void* dx, float* dgamma, float* dbeta, float* dres_colsum, int64_t rows,
                              int32_t cols, cudaStream_t s) {
    if (rows < 0 || cols <= 0) { set_last_error("layernorm_bwd: bad shape"); return kErrBadShape; }
    if (rows == 0) return kOk;
    auto DY = (const __nv_bfloat16*)dy; auto X = (const __nv_bfloat16*)x;
    auto DR = (const __nv_bfloat16*)dres; auto DX = (__nv_bfloat16*)dx;
    const bool vec = (cols % 8 == 0) && cols <= 2048 && ((uintptr_t)x % 16 == 0) && ((uintptr_t)dy % 16 == 0) &&
                     ((uintptr_t)dx % 16 == 0) && (!dres || (uintptr_t)dres % 16 == 0);
    // persistent: each CTA owns register accumulators for dgamma/dbeta, flushed once
    const size_t stage_smem = (size_t)cols * sizeof(float) + (size_t)LN_WARPS * LNB_STAGES * 2 * cols * 2 +
                              LN_WARPS * LNB_STAGES * sizeof(uint64_t);
    if (dres_colsum && !dres) { set_last_error("layernorm_bwd: dres_colsum needs dres"); return kErrBadShape; }
    if (vec && stage_smem <= 227 * 1024 && !g_opt_ln_unstaged.load(std::memory_order_relaxed)) {
        // rows staged through shared memory by bulk copies; every warp owns a contiguous range of rows
        const bool colsum = dres_colsum != nullptr;
        const size_t smem_used = stage_smem;
        const int per_sm = (cols <= 768 && 2 * smem_used <= 227 * 1024) ? 2 : 1;
        int grid = num_sms() * per_sm;
        const int64_t warps_wanted = (rows + 3) / 4;                 // at least ~4 rows per warp
        if ((int64_t)grid * LN_WARPS > warps_wanted) grid = (int)((warps_wanted + LN_WARPS - 1) / LN_WARPS);
        const int64_t rpw = (rows + (int64_t)grid * LN_WARPS - 1) / ((int64_t)grid * LN_WARPS);
#define MGB_LNB_STAGED(N)                                                                                          \
    do {                                                                                                           \
        static PerDeviceOnce once;                                                                                 \
        once.run([] {                                                                                              \
            cudaFuncSetAttribute(ln_bwd_staged_kernel<N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024); \
            cudaFuncSetAttribute(ln_bwd_staged_kernel<N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);  \
        });                                                                                                        \
        if (colsum)                                                                                                \
            launch_pdl(ln_bwd_staged_kernel<N, true>, dim3(grid), dim3(LN_WARPS * 32), smem_used, s, DY, X, gamma, mean, rstd, DR, \
                       DX, dgamma, dbeta, dres_colsum, rows, cols, rpw);                                           \
        else                                                                                                       \
            launch_pdl(ln_bwd_staged_kernel<N, false>, dim3(grid), dim3(LN_WARPS * 32), smem_used, s, DY, X, gamma, mean, rstd, DR, \
                       DX, dgamma, dbeta, (float*)nullptr, rows, cols, rpw);                                       \
    } while (0)
        if (cols <= 256) MGB_LNB_STAGED(1);
        else if (cols <= 512) MGB_LNB_STAGED(2);
        else if (cols <= 768) MGB_LNB_STAGED(3);
        else if (cols <= 1024) MGB_LNB_STAGED(4);
        else if (cols <= 1536) MGB_LNB_STAGED(6);
        else MGB_LNB_STAGED(8);
#undef MGB_LNB_STAGED
        count_launch(1);
        MGB_CHECK_LAUNCH();
        return kOk;
    }
    const int grid = grid_for(rows, LN_WARPS * 4, num_sms() * (cols <= 768 ? 4 : 2));
    const size_t smem = (size_t)2 * cols * sizeof(float);
    if (!vec) ln_bwd_generic_kernel<<<grid, LN_WARPS * 32, 0, s>>>(DY, X, gamma, mean, rstd, DR, DX, dgamma, dbeta, rows, cols);
    else if (cols <= 256) ln_bwd_kernel<1><<<grid, LN_WARPS * 32, smem, s>>>(DY, X, gamma, mean, rstd, DR, DX, dgamma, dbeta, rows, cols);
    else if (cols <= 512) ln_bwd_kernel<2><<<grid, LN_WARPS * 32, smem, s>>>(DY, X, gamma, mean, rstd, DR, DX, dgamma, dbeta, rows, cols);
    else if (cols <= 768) ln_bwd_kernel<3><<<grid, LN_WARPS * 32, smem, s>>>(DY, X, gamma, mean, rstd, DR, DX, dgamma, dbeta, rows, cols);
    else if (cols <= 1024) ln_bwd_kernel<4><<<grid, LN_WARPS * 32, smem, s>>>(DY, X, gamma, mean, rstd, DR, DX, dgamma, dbeta, rows, cols);
    else if (cols <= 1536) ln_bwd_kernel<6><<<grid, LN_WARPS * 32, smem, s>>>(DY, X, gamma, mean, rstd, DR, DX, dgamma, dbeta, rows, cols);
    else ln_bwd_kernel<8><<<grid, LN_WARPS * 32, smem, s>>>(DY, X, gamma, mean, rstd, DR, DX, dgamma, dbeta, rows, cols);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    if (dres_colsum) return mgb_colsum_bf16(dres, cols, rows, cols, dres_colsum, (mgb_stream_t)s);   // unfused fallback
    return kOk;
}

extern "C" int mgb_layernorm_bwd(const void* dy, const void* x, const float* gamma, const float* mean,
                                 const float* rstd, const void* dres, void* dx, float* dgamma, float* dbeta,
                                 int64_t rows, int32_t cols, mgb_stream_t s_) {
    return layernorm_bwd_impl(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, nullptr, rows, cols, (cudaStream_t)s_);
}
extern "C" int mgb_layernorm_bwd_colsum(const void* dy, const void* x, const float* gamma, const float* mean,
                                        const float* rstd, const void* dres, void* dx, float* dgamma, float* dbeta,
                                        float* dres_colsum, int64_t rows, int32_t cols, mgb_stream_t s_) {
    return layernorm_bwd_impl(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, dres_colsum, rows, cols, (cudaStream_t)s_);
}


// =====================================================================================
// Input pipeline: (x, y) training windows cut out of a device-resident byte corpus
// (CharDataset.__getitem__, mingpt/char_dataset.py:35-47, for a whole batch of window starts at once)
// =====================================================================================
// One thread per 8 consecutive positions of one window: 9 code bytes in, 2 x 8 int64 out (two 64-byte stores, and
// the warp's stores are contiguous). The int64 outputs are what dominates: 16 B written per token.
__global__ void __launch_bounds__(256)
token_windows_kernel(const uint8_t* __restrict__ codes, int64_t n_codes, const int64_t* __restrict__ starts, int64_t B,
                     int T, int64_t* __restrict__ x, int64_t* __restrict__ y, int32_t* err_flag) {
    const int per_row = (T + 7) >> 3;
    const int64_t total = B * per_row;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = i / per_row;
        const int t0 = (int)(i - b * per_row) * 8;
        const int64_t s = starts[b];
        if (s < 0 || s + T + 1 > n_codes) {       // the window (T inputs + 1 shifted target) must lie inside the corpus
            if (err_flag && t0 == 0) atomicExch(err_flag, 3);
            continue;
        }
        const uint8_t* src = codes + s + t0;
        const int n = min(8, T - t0);
        int64_t v[9];
#pragma unroll
        for (int e = 0; e < 9; ++e) v[e] = e <= n ? (int64_t)src[e] : 0;
        int64_t* xo = x + b * T + t0;
        int64_t* yo = y + b * T + t0;
        if (n == 8 && ((T & 1) == 0)) {           // rows stay 16-byte aligned: vector stores
#pragma unroll
            for (int e = 0; e < 8; e += 2) {
                *reinterpret_cast<longlong2*>(xo + e) = make_longlong2(v[e], v[e + 1]);
                *reinterpret_cast<longlong2*>(yo + e) = make_longlong2(v[e + 1], v[e + 2]);
            }
        } else {
            for (int e = 0; e < n; ++e) { xo[e] = v[e]; yo[e] = v[e + 1]; }
        }
    }
}

extern "C" int mgb_token_windows(const void* codes, int64_t n_codes, const int64_t* starts, int64_t B, int32_t T,
                                 int64_t* x, int64_t* y, int32_t* err_flag, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (B < 0 || T <= 0 || n_codes < 0) { set_last_error("token_windows: bad shape"); return kErrBadShape; }
    if (((uintptr_t)x % 16) || ((uintptr_t)y % 16)) { set_last_error("token_windows: x / y must be 16-byte aligned"); return kErrBadAlign; }
    if (B == 0) return kOk;
    const int grid = grid_for(B * ((T + 7) / 8), 256, num_sms() * 8);
    token_windows_kernel<<<grid, 256, 0, s>>>((const uint8_t*)codes, n_codes, starts, B, T, x, y, err_flag);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_embedding_fwd(const int64_t* idx, const void* wte, const void* wpe, void* out, int64_t B,
                                 int64_t T, int32_t C, int64_t vocab, int32_t* err_flag, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (B < 0 || T < 0 || C <= 0 || vocab <= 0) { set_last_error("embedding_fwd: bad shape"); return kErrBadShape; }
    if (B * T == 0) return kOk;
    const bool vec = (C % 8 == 0) && ((uintptr_t)wte % 16 == 0) && ((uintptr_t)wpe % 16 == 0) && ((uintptr_t)out % 16 == 0);
    if (vec) {
        const int grid = grid_for(B * T * (C / 8), 256, num_sms() * 16);
        embedding_fwd_kernel<<<grid, 256, 0, s>>>(idx, (const __nv_bfloat16*)wte, (const __nv_bfloat16*)wpe,
                                                  (__nv_bfloat16*)out, B * T, T, C, vocab, err_flag);
    } else {
        const int grid = grid_for(B * T * C, 256, num_sms() * 16);
        embedding_fwd_scalar_kernel<<<grid, 256, 0, s>>>(idx, (const __nv_bfloat16*)wte, (const __nv_bfloat16*)wpe,
                                                         (__nv_bfloat16*)out, B * T, T, C, vocab, err_flag);
    }
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_embedding_bwd(const int64_t* idx, const void* dx, float* dwte, float* dwpe, int64_t B, int64_t T,
                                 int32_t C, int64_t vocab, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (B < 0 || T < 0 || C <= 0 || vocab <= 0) { set_last_error("embedding_bwd: bad shape"); return kErrBadShape; }
    if (B * T == 0) return kOk;
    embedding_bwd_wte_kernel<<<grid_for(B * T * C, 256, num_sms() * 16), 256, 0, s>>>(
        idx, (const __nv_bfloat16*)dx, dwte, B * T, C, vocab);
    embedding_bwd_wpe_kernel<<<grid_for(T * C, 256, num_sms() * 16), 256, 0, s>>>((const __nv_bfloat16*)dx, dwpe, B, T, C);
    count_launch(2);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_cross_entropy_fwd_bwd(void* logits, int64_t ld, const int64_t* targets, float* row_loss,
                                         const int32_t* n_valid, int64_t rows, int64_t vocab, float grad_scale,
                                         int32_t write_grad, int32_t* err_flag, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (rows < 0 || vocab <= 0 || ld < vocab) { set_last_error("cross_entropy: bad shape"); return kErrBadShape; }
    if (rows > INT32_MAX) { set_last_error("cross_entropy: too many rows"); return kErrTooLarge; }
    if (rows == 0) return kOk;
    const size_t stage_smem = (size_t)4 * ld + 4 * sizeof(uint64_t) + 32 * sizeof(float);
    if (write_grad && (ld % 8 == 0) && ((uintptr_t)logits % 16 == 0) && stage_smem <= 227 * 1024 && rows >= 2 * num_sms() &&
        !g_opt_ce_unstaged.load(std::memory_order_relaxed)) {
        static PerDeviceOnce once;
        once.run([] {
            cudaFuncSetAttribute(cross_entropy_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        });
        cross_entropy_staged_kernel<<<num_sms(), (CES_WARPS + 1) * 32, stage_smem, s>>>(
            (__nv_bfloat16*)logits, ld, targets, row_loss, n_valid, rows, vocab, grad_scale, err_flag);
    } else {
        cross_entropy_kernel<<<(int)rows, CE_THREADS, 0, s>>>((__nv_bfloat16*)logits, ld, targets, row_loss, n_valid, vocab,
                                                             grad_scale, write_grad, err_flag);
    }
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_count_valid_targets(const int64_t* targets, int64_t n, int32_t* n_valid, mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (n < 0) { set_last_error("count_valid: bad shape"); return kErrBadShape; }
    cudaMemsetAsync(n_valid, 0, sizeof(int32_t), s);
    if (n == 0) return kOk;
    count_valid_kernel<<<grid_for(n, 256 * 8, num_sms() * 4), 256, 0, s>>>(targets, n, n_valid);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}

extern "C" int mgb_loss_finalize(const float* row_loss, int64_t rows, const int32_t* n_valid, float* loss,
                                 mgb_stream_t s_) {
    cudaStream_t s = (cudaStream_t)s_;
    if (rows < 0) { set_last_error("loss_finalize: bad shape"); return kErrBadShape; }
    loss_finalize_kernel<<<1, 1024, 0, s>>>(row_loss, rows, n_valid, loss);
    count_launch(1);
    MGB_CHECK_LAUNCH();
    return kOk;
}
