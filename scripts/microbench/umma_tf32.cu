// tcgen05.mma kind::tf32 bring-up (development aid): error-compensated 3xTF32 GEMMs with operands split into
// hi/lo copies in shared memory (canonical no-swizzle core-matrix layout) and the accumulator in TMEM.
//   test 1 (layer form):  D[128 x 64] = A[128 x 32] * B[64 x 32]^T      A, B K-major
//   test 2 (wgrad form):  D[64 x 64]  = G[64 pts x 64]^T * H[64 pts x 64]  both operands MN-major views of the
//                          same physical [point][channel] layout, M padded to 128 by reading past G
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_tf32 umma_tf32.cu
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// shared-memory matrix descriptor, no swizzle: start >> 4 | LBO >> 4 << 16 | SBO >> 4 << 32 | version 1 << 46
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
// instruction descriptor: D fp32, A/B tf32, majors, N >> 3, M >> 4
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, bool a_mn, bool b_mn) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t mbar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mbar) : "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ bool mbar_wait(uint32_t mbar, uint32_t parity) {
    for (int it = 0; it < (1 << 22); ++it) {
        uint32_t ok;
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
            : "=r"(ok)
            : "r"(mbar), "r"(parity)
            : "memory");
        if (ok) return true;
    }
    return false;   // never completed: report instead of hanging the box
}
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
    hi = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
    lo = x - hi;
}
// canonical [row][k] -> byte offset: 16-B chunks of 4 consecutive k, 8 rows per 128-B core matrix, core matrices of
// one k-chunk contiguous over the rows (chunk stride = rows * 16 B)
__device__ __forceinline__ uint32_t canon(int row, int k, int rows) {
    return (uint32_t)((k >> 2) * rows * 16 + (row >> 3) * 128 + (row & 7) * 16 + (k & 3) * 4);
}

#define LD32(r, taddr)                                                                                              \
    asm volatile(                                                                                                   \
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                   \
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,"   \
        "%28,%29,%30,%31}, [%32];"                                                                                  \
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),          \
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),    \
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),  \
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])   \
        : "r"(taddr))

struct __align__(1024) Smem {
    unsigned char a_hi[128 * 32 * 4], a_lo[128 * 32 * 4];   // test 1: A 128 x 32;  test 2: G 64 x 64 (+ slack read as pad rows)
    unsigned char b_hi[64 * 64 * 4], b_lo[64 * 64 * 4];     // test 1: B 64 x 32;   test 2: H 64 x 64
    unsigned char slack[32 * 1024];                         // test 2 reads 16 KB past G for the M padding
    uint64_t mbar;
    uint32_t tmem_base;
    int status;
};

// mode 0: test 1, mode 1: test 2.  128 threads.
__global__ void __launch_bounds__(128) umma_test(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D,
                                                 int mode, int* status_out, uint32_t lbo, uint32_t sbo, uint32_t kstep) {
    extern __shared__ __align__(1024) unsigned char raw[];
    Smem& s = *reinterpret_cast<Smem*>(raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        mbar_init(smem_u32(&s.mbar), 1);
        s.status = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)), "r"(64u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // operands -> hi / lo canonical copies
    if (mode == 0) {
        for (int e = tid; e < 128 * 32; e += 128) {
            float hi, lo;
            split_tf32(A[e], hi, lo);
            const uint32_t o = canon(e / 32, e % 32, 128);
            *reinterpret_cast<float*>(s.a_hi + o) = hi;
            *reinterpret_cast<float*>(s.a_lo + o) = lo;
        }
        for (int e = tid; e < 64 * 32; e += 128) {
            float hi, lo;
            split_tf32(B[e], hi, lo);
            const uint32_t o = canon(e / 32, e % 32, 64);
            *reinterpret_cast<float*>(s.b_hi + o) = hi;
            *reinterpret_cast<float*>(s.b_lo + o) = lo;
        }
    } else {
        for (int e = tid; e < 64 * 64; e += 128) {   // [pt][ch], 64 points
            float hi, lo;
            const uint32_t o = canon(e / 64, e % 64, 64);
            split_tf32(A[e], hi, lo);
            *reinterpret_cast<float*>(s.a_hi + o) = hi;
            *reinterpret_cast<float*>(s.a_lo + o) = lo;
            split_tf32(B[e], hi, lo);
            *reinterpret_cast<float*>(s.b_hi + o) = hi;
            *reinterpret_cast<float*>(s.b_lo + o) = lo;
        }
        for (int e = tid; e < 32 * 1024 / 4; e += 128) reinterpret_cast<float*>(s.slack)[e] = 0.f;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s.tmem_base;
    if (tid == 0) {
        if (mode == 0) {
            const uint32_t idesc = make_idesc(128, 64, false, false);
            const uint32_t CSA = 128 * 16, CSB = 64 * 16;
            for (int ks = 0; ks < 4; ++ks) {   // K = 32 = 4 steps of 8
                const uint64_t ah = make_desc(smem_u32(s.a_hi) + ks * 2 * CSA, CSA, 128);
                const uint64_t al = make_desc(smem_u32(s.a_lo) + ks * 2 * CSA, CSA, 128);
                const uint64_t bh = make_desc(smem_u32(s.b_hi) + ks * 2 * CSB, CSB, 128);
                const uint64_t bl = make_desc(smem_u32(s.b_lo) + ks * 2 * CSB, CSB, 128);
                umma_tf32(tmem, al, bh, idesc, ks > 0);
                umma_tf32(tmem, ah, bl, idesc, 1);
                umma_tf32(tmem, ah, bh, idesc, 1);
            }
        } else {
            const uint32_t idesc = make_idesc(128, 64, true, true);
            for (int ks = 0; ks < 8; ++ks) {   // K = 64 points = 8 steps of 8
                const uint64_t ah = make_desc(smem_u32(s.a_hi) + ks * kstep, lbo, sbo);
                const uint64_t al = make_desc(smem_u32(s.a_lo) + ks * kstep, lbo, sbo);
                const uint64_t bh = make_desc(smem_u32(s.b_hi) + ks * kstep, lbo, sbo);
                const uint64_t bl = make_desc(smem_u32(s.b_lo) + ks * kstep, lbo, sbo);
                umma_tf32(tmem, al, bh, idesc, ks > 0);
                umma_tf32(tmem, ah, bl, idesc, 1);
                umma_tf32(tmem, ah, bh, idesc, 1);
            }
        }
        umma_commit(smem_u32(&s.mbar));
    }
    __syncwarp();
    if (!mbar_wait(smem_u32(&s.mbar), 0)) s.status = 1;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // epilogue: warp w reads TMEM lanes 32w..32w+31 (rows), 64 columns
    uint32_t r[32];
    for (int half = 0; half < 2; ++half) {
        const uint32_t taddr = tmem + ((uint32_t)(32 * warp) << 16) + half * 32;
        LD32(r, taddr);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        const int row = 32 * warp + lane;
        for (int c = 0; c < 32; ++c) D[row * 64 + half * 32 + c] = __uint_as_float(r[c]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64u) : "memory");
    if (tid == 0) *status_out = s.status;
}

static float frand(uint32_t& st) {
    st = st * 1664525u + 1013904223u;
    return ((st >> 8) * (1.0f / 16777216.0f)) * 2.f - 1.f;
}

int main() {
    uint32_t st = 12345;
    int fails = 0;
    const uint32_t cand[][3] = {{0, 0, 0}, {128, 1024, 128}, {1024, 128, 128}, {128, 1024, 256}, {1024, 128, 256},
                                {16, 1024, 128}, {1024, 16, 128}};
    for (int ci = 0; ci < 7; ++ci) {
        const int mode = ci == 0 ? 0 : 1;
        const int na = mode == 0 ? 128 * 32 : 64 * 64, nb = mode == 0 ? 64 * 32 : 64 * 64;
        std::vector<float> A(na), B(nb), D(128 * 64, -1.f);
        for (auto& v : A) v = frand(st) * 3.f;
        for (auto& v : B) v = frand(st);
        float *dA, *dB, *dD;
        int* dS;
        cudaMalloc(&dA, na * 4); cudaMalloc(&dB, nb * 4); cudaMalloc(&dD, 128 * 64 * 4); cudaMalloc(&dS, 4);
        cudaMemcpy(dA, A.data(), na * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(dB, B.data(), nb * 4, cudaMemcpyHostToDevice);
        cudaFuncSetAttribute(umma_test, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
        umma_test<<<1, 128, sizeof(Smem)>>>(dA, dB, dD, mode, dS, cand[ci][0], cand[ci][1], cand[ci][2]);
        cudaError_t e = cudaDeviceSynchronize();
        int status = -1;
        cudaMemcpy(&status, dS, 4, cudaMemcpyDeviceToHost);
        cudaMemcpy(D.data(), dD, 128 * 64 * 4, cudaMemcpyDeviceToHost);
        double maxerr = 0, maxref = 0;
        const int rows = mode == 0 ? 128 : 64;
        for (int m = 0; m < rows; ++m)
            for (int n = 0; n < 64; ++n) {
                double ref = 0;
                if (mode == 0) for (int k = 0; k < 32; ++k) ref += (double)A[m * 32 + k] * B[n * 32 + k];
                else for (int p = 0; p < 64; ++p) ref += (double)A[p * 64 + m] * B[p * 64 + n];
                maxerr = fmax(maxerr, fabs(ref - D[m * 64 + n]));
                maxref = fmax(maxref, fabs(ref));
            }
        printf("lbo %u sbo %u kstep %u: ", cand[ci][0], cand[ci][1], cand[ci][2]);
        printf("mode %d: cuda=%s status=%d max|err|=%.3e max|ref|=%.3e  D[0]=%g D[1]=%g D[64]=%g\n", mode, cudaGetErrorString(e),
               status, maxerr, maxref, D[0], D[1], D[64]);
        if (e != cudaSuccess || status != 0 || !(maxerr < 1e-4 * maxref)) ++fails;
        cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(dS);
        if (e != cudaSuccess) break;
    }
    printf(fails ? "FAIL\n" : "PASS\n");
    return fails;
}
