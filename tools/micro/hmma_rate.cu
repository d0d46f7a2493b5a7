// Microbenchmark: issue rate of the legacy tensor path (mma.sync.m16n8k16 bf16 -> fp32, SASS HMMA.16816.F32.BF16)
// on B200, per SM sub-partition, as a function of resident warps.  Decides whether a weight-streaming GEMM at
// M = 32 / 64 rows can reach the HBM roofline with mma.sync (needs one HMMA per 22 / 11 clk per sub-partition)
// or has to move to tcgen05.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hmma_rate hmma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void hmma_loop(float* out, int iters, long long* cycles) {
    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;
    uint32_t a0 = threadIdx.x, a1 = threadIdx.x * 3, a2 = 7, a3 = 11, b0 = 0x3f803f80u, b1 = 0x3f803f80u;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)      // 8 independent accumulators
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(acc[i][0]), "+f"(acc[i][1]), "+f"(acc[i][2]), "+f"(acc[i][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) for (int e = 0; e < 4; ++e) s += acc[i][e];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * sizeof(float));
    cudaMallocManaged(&cyc, sizeof(long long));
    const int iters = 4096;
    for (int warps = 1; warps <= 16; warps *= 2) {          // warps per SM (one CTA per SM), spread over 4 sub-partitions
        for (int rep = 0; rep < 2; ++rep) {
            hmma_loop<<<148, warps * 32>>>(out, iters, cyc);
            cudaDeviceSynchronize();
        }
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        hmma_loop<<<148, warps * 32>>>(out, iters, cyc);
        cudaEventRecord(e1);
        cudaDeviceSynchronize();
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        const double hmma_per_warp = (double)iters * 8;
        const double warps_per_smsp = warps < 4 ? 1.0 : warps / 4.0;
        const double clk_per_hmma_smsp = (double)*cyc / (hmma_per_warp * warps_per_smsp);
        const double tflops = 148.0 * warps * hmma_per_warp * 4096.0 / (ms * 1e-3) / 1e12;
        printf("{\"warps_per_sm\": %d, \"cycles\": %lld, \"clk_per_hmma_per_subpartition\": %.2f, \"ms\": %.4f, \"chip_TFLOPs\": %.1f, \"err\": \"%s\"}\n",
               warps, *cyc, clk_per_hmma_smsp, ms, tflops, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
