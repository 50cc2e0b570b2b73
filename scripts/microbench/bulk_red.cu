// Throughput of the TMA engine's bulk reduction (cp.reduce.async.bulk ... .add.f32, shared -> global) to pseudo-random
// feature-plane texels, as an alternative path for the backward's scatter: one elected lane issues a whole 128-B (one
// texel) or 256-B (two x-adjacent texels) reduction from a staged shared-memory row; no LSU data-path traffic.
// Also: the same kernel issuing red.global.add.v4.f32 from 8 lanes per line at the same time (do the two paths add up?).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_red bulk_red.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// MODE 0: bulk reduce only (BYTES per op, one op per lane per iteration: 32 ops per warp-iteration)
// MODE 1: bulk reduce issued by lane 0 only (8 ops per warp-iteration, serial issue)
// MODE 2: red.v4 only (8 lanes per 128-B line), same line count as MODE 0 with BYTES = 128
// MODE 3: both at once: every iteration 4 red.v4 lines per warp (32 lanes) + 4 bulk ops of 128 B (lanes 0..3)
template <int MODE, int BYTES>
__global__ void __launch_bounds__(256) k(float* buf, uint32_t lines, int iters) {
    __shared__ __align__(128) float stage[8][64];   // one 256-B row per warp
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = lane; i < 64; i += 32) stage[warp][i] = 1.0f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    uint32_t st = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    const uint32_t src = smem_u32(stage[warp]);
    const uint32_t span = BYTES / 128;
    for (int i = 0; i < iters; ++i) {
        st = st * 1664525u + 1013904223u;
        const uint32_t line = ((st >> 4) % (lines - span + 1));
        float* dst = buf + (size_t)line * 32;
        if (MODE == 0 || (MODE == 1 && lane == 0)) {
            const int n = MODE == 0 ? 1 : 8;
            for (int r = 0; r < n; ++r) {
                float* d = dst + (size_t)((r * 7919u) % (lines - span + 1 - line)) * 32;
                asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(d), "r"(src),
                             "r"((uint32_t)BYTES)
                             : "memory");
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            if ((i & 7) == 7) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        if (MODE == 2 || MODE == 3) {
            // lines chosen per 8-lane group
            const uint32_t g = __shfl_sync(0xffffffffu, line, lane & 24);
            float* p = buf + (size_t)g * 32 + (lane & 7) * 4;
            asm volatile("red.global.add.v4.f32 [%0], {%1,%1,%1,%1};" ::"l"(p), "f"(1.0f));
        }
        if (MODE == 3 && lane < 4) {
            asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst), "r"(src),
                         "r"(128u)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            if ((i & 7) == 7) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <int MODE, int BYTES>
static void run(const char* name, float* buf, uint32_t lines, double lines_per_thread_iter) {
    const int iters = 256, grid = 148 * 8;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    float ms = 0;
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        k<MODE, BYTES><<<grid, 256>>>(buf, lines, iters);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        cudaEventElapsedTime(&ms, a, b);
    }
    cudaError_t e = cudaGetLastError();
    const double ops = (double)grid * 256 * iters * lines_per_thread_iter;
    printf("  %-46s %8.3f ms  %7.2f G ops/s  %7.1f GB/s  %s\n", name, ms, ops / ms / 1e6, ops * BYTES / ms / 1e6,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main() {
    for (size_t mb : {25, 200}) {
        const uint32_t lines = mb * 1024 * 1024 / 128;
        float* buf;
        cudaMalloc(&buf, (size_t)lines * 128);
        cudaMemset(buf, 0, (size_t)lines * 128);
        printf("%zu MB buffer (ops = 128-B / 256-B reductions)\n", mb);
        run<0, 128>("bulk reduce 128 B, every lane issues", buf, lines, 1.0);
        run<0, 256>("bulk reduce 256 B, every lane issues", buf, lines, 1.0);
        run<1, 128>("bulk reduce 128 B, lane 0 issues 8 per iteration", buf, lines, 8.0 / 32);
        run<1, 256>("bulk reduce 256 B, lane 0 issues 8 per iteration", buf, lines, 8.0 / 32);
        run<2, 128>("red.v4 x 8 lanes per 128-B line", buf, lines, 1.0 / 8);
        run<3, 128>("red.v4 (4 lines) + bulk 128 B (4 ops) per warp-iter", buf, lines, 8.0 / 32);
        cudaFree(buf);
    }
    return 0;
}
