// 3xTF32 mma.sync GEMM fragments taken from shared-memory operands that live in the backward kernel's
// [channel][point] layout (row stride AS = 132 floats) or in weight matrices [n][k] (row stride = 8 mod 32).
// Fragment conventions of mma.m16n8k8 (g = lane >> 2, t = lane & 3):
//   A: a0 (row g, k t)  a1 (row g+8, k t)  a2 (row g, k t+4)  a3 (row g+8, k t+4)
//   B: b0 (k t, col g)  b1 (k t+4, col g)          C: c0 (g, 2t) c1 (g, 2t+1) c2 (g+8, 2t) c3 (g+8, 2t+1)
// Where the contraction index is a channel, logical k = t / t+4 is mapped to channels k0+2t / k0+2t+1 for both
// operands (any permutation of k is legal as long as A and B agree); this makes every request conflict-free:
// banks 8t+g for the activation rows and one 64-bit load per weight fragment.
#pragma once
#include "mlp_mma.cuh"
#include "tile.cuh"

struct FragA { uint32_t h[4], l[4]; };
struct FragB { uint32_t h0, h1, l0, l1; };

// rows = points m0+g (+8), k = channels k0+2t, k0+2t+1 of M[channel][point]
template <bool FAST = false>
__device__ __forceinline__ void fragA_chpt(const float* __restrict__ M, int k0, int m0, int g, int t, FragA& f) {
    const float* p = M + (k0 + 2 * t) * AS + m0 + g;
    split_tf32<FAST>(p[0], f.h[0], f.l[0]);
    split_tf32<FAST>(p[8], f.h[1], f.l[1]);
    split_tf32<FAST>(p[AS], f.h[2], f.l[2]);
    split_tf32<FAST>(p[AS + 8], f.h[3], f.l[3]);
}
// cols = weight rows n0+g, k = k0+2t, k0+2t+1 of W[n][k] (row stride ws, ws % 32 == 8)
template <bool FAST = false>
__device__ __forceinline__ void fragB_w(const float* __restrict__ Wm, int ws, int k0, int n0, int g, int t, FragB& f) {
    const float2 v = *reinterpret_cast<const float2*>(Wm + (n0 + g) * ws + k0 + 2 * t);
    split_tf32<FAST>(v.x, f.h0, f.l0);
    split_tf32<FAST>(v.y, f.h1, f.l1);
}
// cols = n0+g, k = rows k0+2t, k0+2t+1 of a matrix stored W[k][n] (row stride ws = 8 mod 32: banks 16t+g, a 2-way
// conflict accepted to avoid a transposed copy of the weights in shared memory)
template <bool FAST = false>
__device__ __forceinline__ void fragB_kn(const float* __restrict__ Wm, int ws, int k0, int n0, int g, int t, FragB& f) {
    const float* p = Wm + (k0 + 2 * t) * ws + n0 + g;
    split_tf32<FAST>(p[0], f.h0, f.l0);
    split_tf32<FAST>(p[ws], f.h1, f.l1);
}
// weight gradients: rows = channels m0+g (+8) of M[channel][point], k = points k0+t, k0+t+4
template <bool FAST = false>
__device__ __forceinline__ void fragA_rowk(const float* __restrict__ M, int m0, int k0, int g, int t, FragA& f) {
    const float* p = M + (m0 + g) * AS + k0 + t;
    split_tf32<FAST>(p[0], f.h[0], f.l[0]);
    split_tf32<FAST>(p[8 * AS], f.h[1], f.l[1]);
    split_tf32<FAST>(p[4], f.h[2], f.l[2]);
    split_tf32<FAST>(p[8 * AS + 4], f.h[3], f.l[3]);
}
// cols = channels n0+g of M[channel][point], k = points k0+t, k0+t+4
template <bool FAST = false>
__device__ __forceinline__ void fragB_rowk(const float* __restrict__ M, int n0, int k0, int g, int t, FragB& f) {
    const float* p = M + (n0 + g) * AS + k0 + t;
    split_tf32<FAST>(p[0], f.h0, f.l0);
    split_tf32<FAST>(p[4], f.h1, f.l1);
}

// acc[TM][TN] += A[TM] x B[TN], error-compensated: the two small products first, then the main one; the loop
// order keeps dependent mma TM*TN instructions apart
template <int TM, int TN, bool FAST = false>
__device__ __forceinline__ void mma3_tiles(float (&acc)[TM][TN][4], const FragA (&a)[TM], const FragB (&b)[TN]) {
    if (!FAST) {
#pragma unroll
        for (int mt = 0; mt < TM; ++mt)
#pragma unroll
            for (int nt = 0; nt < TN; ++nt) mma_tf32(acc[mt][nt], a[mt].l, b[nt].h0, b[nt].h1);
#pragma unroll
        for (int mt = 0; mt < TM; ++mt)
#pragma unroll
            for (int nt = 0; nt < TN; ++nt) mma_tf32(acc[mt][nt], a[mt].h, b[nt].l0, b[nt].l1);
    }
#pragma unroll
    for (int mt = 0; mt < TM; ++mt)
#pragma unroll
        for (int nt = 0; nt < TN; ++nt) mma_tf32(acc[mt][nt], a[mt].h, b[nt].h0, b[nt].h1);
}

// out[pt][n] = sum_k Act[k][pt] * W(n,k) for this warp's 32 points (m0, m0+16) and TN n-tiles from n0;
// W(n,k) = Wm[n][k] (KN = false) or Wm[k][n] (KN = true)
template <int K, int TN, bool KN, bool FAST = false>
__device__ __forceinline__ void gemm_pts_mma(const float* __restrict__ Act, const float* __restrict__ Wm, int ws, int m0,
                                             int n0, int g, int t, float (&acc)[2][TN][4]) {
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < TN; ++nt)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[mt][nt][i] = 0.f;
#pragma unroll 2
    for (int k0 = 0; k0 < K; k0 += 8) {
        FragA a[2];
        FragB b[TN];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) fragA_chpt<FAST>(Act, k0, m0 + 16 * mt, g, t, a[mt]);
#pragma unroll
        for (int nt = 0; nt < TN; ++nt) {
            if (KN) fragB_kn<FAST>(Wm, ws, k0, n0 + 8 * nt, g, t, b[nt]);
            else fragB_w<FAST>(Wm, ws, k0, n0 + 8 * nt, g, t, b[nt]);
        }
        mma3_tiles<2, TN, FAST>(acc, a, b);
    }
}

// acc[TN] += sum over points [0, npts) of G[m0+row][pt] * A[n0+8nt+col][pt]   (npts multiple of 8)
template <int TN, bool FAST = false>
__device__ __forceinline__ void wgrad_mma(const float* __restrict__ G, const float* __restrict__ A, int m0, int n0,
                                          int npts, int g, int t, float (&acc)[1][TN][4]) {
#pragma unroll 2
    for (int k0 = 0; k0 < npts; k0 += 8) {
        FragA a[1];
        FragB b[TN];
        fragA_rowk<FAST>(G, m0, k0, g, t, a[0]);
#pragma unroll
        for (int nt = 0; nt < TN; ++nt) fragB_rowk<FAST>(A, n0 + 8 * nt, k0, g, t, b[nt]);
        mma3_tiles<1, TN, FAST>(acc, a, b);
    }
}
