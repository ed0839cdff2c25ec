// Microbenchmark: cycles per tcgen05.mma (cta_group::1, kind::f16, M=128, K=16) for small N, and how the ISSUE
// STYLE bounds it: (a) one thread runs the whole loop inside `if (lane == 0)`; (b) the whole warp runs the loop
// (descriptor arithmetic stays warp-uniform) and only the instruction itself is predicated on one lane.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../mingpt_distributed_b200/csrc -I../../include mma_rate.cu -o mma_rate
#include <cstdio>
#include "common.cuh"
using namespace mgb;

__device__ __forceinline__ bool elect_one() {
    uint32_t pred = 0;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

template <int STYLE, int NACC>
__global__ void __launch_bounds__(128, 1) mma_rate_kernel(long long* out, int n, int ts, int count) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint32_t tmem_ptr;
    __shared__ uint64_t bar;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;
    if (warp == 0) tmem_alloc(&tmem_ptr, 512);
    if (threadIdx.x == 32) { mbar_init(&bar, 1); fence_mbar_init(); }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 0 && (STYLE >= 1 || lane == 0)) {
        const bool leader = (STYLE == 2) ? elect_one() : (lane == 0);
        const uint32_t idesc = make_idesc_bf16(128, n, 0, 0);
        const uint32_t sa = smem_u32(smem), sb = sa + 64 * 1024;
        const uint64_t da = make_smem_desc_sw128(sa, 0, 1024);
        const uint64_t db = make_smem_desc_sw128(sb, 0, 1024);
        const uint32_t tmem_d = tmem_ptr, tmem_a = tmem_ptr + 256;
        const long long t0 = clock64();
        for (int i = 0; i < count; i += 4) {
            const uint64_t a0 = da + (uint64_t)((i >> 2) & 3) * 512, b0 = db + (uint64_t)((i >> 2) & 3) * 512;   // "stage"
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t dcol = tmem_d + (uint32_t)(k % NACC) * 64;
                if (STYLE == 0 || leader) {
                    if (ts) umma_bf16_ts(dcol, tmem_a + k * 8, b0 + 2 * k, idesc, 1u);
                    else umma_bf16_ss(dcol, a0 + 2 * k, b0 + 2 * k, idesc, 1u);
                }
            }
        }
        const long long t1 = clock64();
        if (lane == 0) {
            umma_commit(&bar);
            mbar_wait(&bar, 0);
            const long long t2 = clock64();
            out[0] = t1 - t0; out[1] = t2 - t0;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_ptr, 512); }
}

template <int STYLE, int NACC>
void run(long long* d, int n, int ts) {
    const int count = 4096;
    long long h[2];
    cudaFuncSetAttribute(mma_rate_kernel<STYLE, NACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 170 * 1024);
    mma_rate_kernel<STYLE, NACC><<<1, 128, 170 * 1024>>>(d, n, ts, count);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%s  acc x%d  %s N=%3d: issue %.1f cyc/mma, complete %.1f cyc/mma (floor %d)\n",
           STYLE == 2 ? "warp-uniform loop, elect.sync issues" : STYLE ? "warp-uniform loop, lane 0 issues    " : "single-thread loop                  ", NACC, ts ? "TS" : "SS", n,
           (double)h[0] / count, (double)h[1] / count, 128 * n / 256);
}

int main() {
    long long* d; cudaMalloc(&d, 64);
    for (int ts = 0; ts < 2; ++ts) for (int n : {64, 128}) {
        run<0, 1>(d, n, ts); run<1, 1>(d, n, ts); run<2, 1>(d, n, ts);
        run<2, 2>(d, n, ts);
        if (n == 64) run<2, 4>(d, n, ts);
    }
    return 0;
}
