// Microbenchmark: issue rate of warp-level mma.sync (tf32 m16n8k8, bf16 m16n8k16) and FFMA on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

template <int KIND>
__global__ void __launch_bounds__(256) k(float* out, int iters) {
    float c[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) c[i][j] = threadIdx.x * 0.001f + i + j;
    unsigned a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    float f0 = a0 * 1e-3f, f1 = 1.0001f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (KIND == 0) {
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                             : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            } else if (KIND == 1) {
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                             : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) c[i][j] = fmaf(c[i][j], f1, f0);
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) s += c[i][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int KIND>
void run(const char* name, double macs_per_inst, int ctas_per_sm) {
    int sms;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* out;
    cudaMalloc(&out, sizeof(float) * sms * ctas_per_sm * 256);
    const int iters = 20000;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    k<KIND><<<sms * ctas_per_sm, 256>>>(out, 100);
    cudaEventRecord(a);
    k<KIND><<<sms * ctas_per_sm, 256>>>(out, iters);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    double warps = (double)sms * ctas_per_sm * 8;
    double inst = warps * iters * 8.0 * (KIND == 2 ? 4 : 1);
    double macs = inst * macs_per_inst;
    printf("%-28s ctas/sm=%d  %.3f ms  %.2f TMAC/s  (%.1f warp-inst/clk/SM at 1.9GHz)\n", name, ctas_per_sm, ms,
           macs / ms * 1e-9, inst / sms / (ms * 1e-3 * 1.9e9));
    cudaFree(out);
}

int main() {
    for (int c = 1; c <= 4; c *= 2) {
        run<0>("mma m16n8k8 tf32", 16 * 8 * 8, c);
        run<1>("mma m16n8k16 bf16", 16 * 8 * 16, c);
        run<2>("ffma", 32, c);
    }
    return 0;
}
