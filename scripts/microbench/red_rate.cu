// Throughput of red.global.add.v4.f32 to pseudo-random 128-B lines (development aid): 8 lanes cover one line, as in the
// backward's scatter.  Compared with plain 128-bit stores to the same addresses.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_rate red_rate.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
template <int MODE>
__global__ void __launch_bounds__(256) k(float* buf, uint32_t lines, int iters) {
    const int lane8 = threadIdx.x & 7;
    uint32_t grp = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    uint32_t st = grp * 2654435761u + 12345u;
    for (int i = 0; i < iters; ++i) {
        st = st * 1664525u + 1013904223u;
        const uint32_t line = (st >> 4) % lines;
        float* p = buf + (size_t)line * 32 + lane8 * 4;
        const float v = (float)i;
        if (MODE == 0) asm volatile("red.global.add.v4.f32 [%0], {%1,%1,%1,%1};" ::"l"(p), "f"(v));
        else if (MODE == 1) asm volatile("st.global.v4.f32 [%0], {%1,%1,%1,%1};" ::"l"(p), "f"(v));
        else if (MODE == 2) asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v));
        else {
            // MODE 3: every lane its own pseudo-random 4-B word (the probability-plane reductions of the backward:
            // one lane per texel, a warp-level instruction touches up to 32 different sectors)
            uint32_t s2 = (st ^ (threadIdx.x * 0x9E3779B9u)) * 1664525u + 1013904223u;
            float* q = buf + (size_t)((s2 >> 3) % (lines * 32u));
            asm volatile("red.global.add.f32 [%0], %1;" ::"l"(q), "f"(v));
        }
    }
}
int main() {
    for (size_t mb : {25, 200, 1600}) {
        const uint32_t lines = mb * 1024 * 1024 / 128;
        float* buf;
        cudaMalloc(&buf, (size_t)lines * 128);
        cudaMemset(buf, 0, (size_t)lines * 128);
        const int iters = 256, grid = 148 * 8;
        for (int mode = 0; mode < 4; ++mode) {
            cudaEvent_t a, b;
            cudaEventCreate(&a); cudaEventCreate(&b);
            for (int rep = 0; rep < 2; ++rep) {
                cudaEventRecord(a);
                if (mode == 0) k<0><<<grid, 256>>>(buf, lines, iters);
                else if (mode == 1) k<1><<<grid, 256>>>(buf, lines, iters);
                else if (mode == 2) k<2><<<grid, 256>>>(buf, lines, iters);
                else k<3><<<grid, 256>>>(buf, lines, iters);
                cudaEventRecord(b);
                cudaEventSynchronize(b);
            }
            float ms;
            cudaEventElapsedTime(&ms, a, b);
            const double bytes = (double)grid * 256 * iters * (mode >= 2 ? 4 : 16);
            printf("%4zu MB buffer  %-18s %8.3f ms  %7.1f GB/s  (%.2f G lane-ops/s)\n", mb,
                   mode == 0 ? "red.add.v4.f32" : mode == 1 ? "st.v4.f32" : mode == 2 ? "red.add.f32 (4 B)" : "red.add.f32 scattered", ms, bytes / ms / 1e6,
                   (double)grid * 256 * iters / ms / 1e6);
        }
        cudaFree(buf);
    }
    return 0;
}
