// Microbenchmark: TMEM read bandwidth / latency of tcgen05.ld.32x32b.x32 on one SM, for 4 and 8 warps.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../mingpt_distributed_b200/csrc -I../../include tmem_bw.cu -o tmem_bw
#include <cstdio>
#include "common.cuh"
using namespace mgb;

__global__ void __launch_bounds__(256, 1) tmem_bw_kernel(long long* out, int iters, int nloads, int dep) {
    __shared__ uint32_t tmem_ptr;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tmem_alloc(&tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t taddr = tmem_ptr + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 256;
    uint32_t acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        uint32_t r[32], q[32];
        if (nloads == 1) {
            tmem_ld_32x32b_x32(taddr + (it & 3) * 32, r);
            tmem_ld_wait();
            acc += r[0] + r[31];
        } else {
            tmem_ld_32x32b_x32(taddr + (it & 3) * 32, r);
            tmem_ld_32x32b_x32(taddr + 128 + (it & 3) * 32, q);
            tmem_ld_wait();
            acc += r[0] + q[31];
        }
        if (dep) {     // some dependent ALU work between loads, like an epilogue
#pragma unroll
            for (int j = 0; j < 32; ++j) acc += __float_as_uint(exp2f(__uint_as_float(r[j])));
        }
    }
    const long long t1 = clock64();
    if ((threadIdx.x & 31) == 0) { out[warp * 2] = t1 - t0; out[warp * 2 + 1] = acc; }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_ptr, 512); }
}

int main() {
    long long* d; cudaMalloc(&d, 64 * 8);
    long long h[64];
    const int iters = 1000;
    for (int threads : {32, 128, 256}) for (int nloads : {1, 2}) for (int dep : {0, 1}) {
        tmem_bw_kernel<<<1, threads>>>(d, iters, nloads, dep);
        tmem_bw_kernel<<<1, threads>>>(d, iters, nloads, dep);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        long long mx = 0; for (int w = 0; w < threads / 32; ++w) mx = h[w * 2] > mx ? h[w * 2] : mx;
        const double bytes = (double)iters * nloads * 4096.0 * (threads / 32);
        printf("warps %d loads/iter %d exp2 %d: %lld cycles, %.1f cycles/iter, %.1f B/clk/SM\n", threads / 32, nloads, dep, mx,
               (double)mx / iters, bytes / mx);
    }
    return 0;
}
