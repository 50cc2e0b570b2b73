// Decoder MLP (32 -> 64 -> 64 -> 4) for 32 points per warp on the tensor pipe.  Replaces libraries/NeRF/net.py:10-27 (StyledMLP) applied point-wise
// (libraries/triplane/triplane_nerf.py:44-47) with W_eff / bias supplied per sample.
//
// Precision: the parity bar is 1e-4 on fp32 outputs, so every product is the error-compensated 3xTF32 form
//   a*b ~= hi(a)*hi(b) + hi(a)*lo(b) + lo(a)*hi(b),   hi(x) = x with the low 13 mantissa bits cleared, lo = x - hi
// (the dropped lo*lo term is ~2^-22 relative), accumulated in fp32 by mma.sync.m16n8k8.tf32.  Measured issue
// rate on B200: 139 TMAC/s (profiles/mma_rate_r1.txt) -> 46 TMAC/s effective, above the 36 TMAC/s FFMA peak,
// and -- the real point -- it takes the decoder off the FMA / shared-memory pipes that the gather needs: the
// FFMA version was bound by LDS wavefronts (2-4 per 128-bit request, profiles/ncu_summary_r1_v3.txt).
//
// Register chaining: the m16n8 accumulator fragment (c0,c1 = row g, cols 2t,2t+1; c2,c3 = row g+8) is used
// directly as the A fragment of the next layer's k-block by choosing the k permutation
//   logical k = t  <->  channel 8j+2t,   logical k = t+4  <->  channel 8j+2t+1
// and loading the B fragment (weights) with the same permutation: one 64-bit load W[n][8j+2t .. 8j+2t+1].
// Layers 1 and 2 each keep all 8 n-tiles x 2 m-tiles in flight (16 independent accumulator chains: one n-tile at
// a time leaves 4 chains of dependent mma and the tensor pipe idle).  Two such accumulator sets do not fit in 128
// registers, so layer 1's activations are parked in the warp's own slab (fragment layout in, same layout out:
// the k permutation above still applies), and layer 2's accumulators feed layer 3 straight from registers.
#pragma once
#include "common.cuh"
#include "gather.cuh"

// Per-warp feature slab X[point][channel]: dense 32-float rows, the four 8-float granules of row r stored at
// granule index j ^ (r & 3).  A-fragment loads (64-bit, rows g / g+8, column 8j+2t) then hit bank
// 8*(j ^ (g&3)) + 2t: conflict-free per half-warp, with no padding.
#define XS 32
__device__ __forceinline__ int xs_col(int row, int col) { return col ^ (8 * (row & 3)); }
#define W1M 40   // weight row strides for the 64-bit B-fragment loads (same bank argument)
#define W2M 72

struct __align__(16) DecoderMma {
    float w1[HID * W1M];   // [o][i]
    float w2[HID * W2M];   // [o][i]
    float w3[8 * W2M];     // [c][o], rows 4..7 zero
    float b1[HID], b2[HID], b3[8];
};

__device__ __forceinline__ void load_decoder_mma(DecoderMma& d, const enarf_scene_t& sc, int b, int tid, int nthr) {
    const float* w1 = sc.w1 + (size_t)b * HID * FEAT;
    const float* w2 = sc.w2 + (size_t)b * HID * HID;
    const float* w3 = sc.w3 + (size_t)b * 4 * HID;
    // Every layer is conv + bias -> LeakyReLU(0.2) * sqrt2.  LeakyReLU is positively homogeneous, so the sqrt2 is folded into
    // the weights and biases here (once per sample and CTA) and the per-element activation is max(v, 0.2 v): two
    // instructions instead of three; the biases are the accumulators' initial values instead of a separate add.
    const float S2 = 1.4142135623730951f;
    for (int e = tid; e < HID * FEAT; e += nthr) d.w1[(e / FEAT) * W1M + (e % FEAT)] = w1[e] * S2;
    for (int e = tid; e < HID * HID; e += nthr) d.w2[(e / HID) * W2M + (e % HID)] = w2[e] * S2;
    for (int e = tid; e < 8 * HID; e += nthr) d.w3[(e / HID) * W2M + (e % HID)] = (e < 4 * HID) ? w3[e] * S2 : 0.f;
    for (int e = tid; e < HID; e += nthr) {
        d.b1[e] = sc.b1[e] * S2;
        d.b2[e] = sc.b2[e] * S2;
    }
    if (tid < 8) d.b3[tid] = tid < 4 ? sc.b3[tid] * S2 : 0.f;
}

// LeakyReLU(0.2) of a pre-activation whose sqrt2 gain is already in the weights (load_decoder_mma)
__device__ __forceinline__ float lrelu_folded(float v) { return fmaxf(v, 0.2f * v); }

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// FAST (bf16-plane modes, parity bar 1e-2): one TF32 product instead of three -- hi is x rounded to nearest TF32
// (unbiased, 2^-11 relative, below the 2^-9 of the bf16 planes themselves) and lo is never formed.
template <bool FAST = false>
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
    if (FAST) {
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(x));
        lo = 0u;
    } else {
        hi = __float_as_uint(x) & 0xffffe000u;
        lo = __float_as_uint(x - __uint_as_float(hi));
    }
}

// One k-block of one 16-row tile against one n-tile: main += hi*hi, small += lo*hi + hi*lo
template <bool FAST = false>
__device__ __forceinline__ void mma3(float (&main)[4], float (&small)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4],
                                     uint32_t bh0, uint32_t bh1, uint32_t bl0, uint32_t bl1) {
    if (!FAST) {
        mma_tf32(small, al, bh0, bh1);
        mma_tf32(small, ah, bl0, bl1);
    }
    mma_tf32(main, ah, bh0, bh1);
}

// X: this warp's slab [32 points][XS] (rows of points it does not own may hold anything finite or not: a row
// only influences its own outputs).  Returns the head's pre-activation outputs as accumulator fragments:
// o3[mt][0] = (point 16mt+g,   channel 2t)   o3[mt][1] = (point 16mt+g,   channel 2t+1)
// o3[mt][2] = (point 16mt+g+8, channel 2t)   o3[mt][3] = (point 16mt+g+8, channel 2t+1);  channels >= 4 are padding.
// A1: at least 32 * A1S floats of the warp's own shared memory, may alias X.
#define A1S 64
template <bool FAST = false>
__device__ __forceinline__ void decode32_mma(const float* X, float* A1, const DecoderMma& d, int lane, float (&o3)[2][4]) {
    const int g = lane >> 2, t = lane & 3;
    float acc1[2][8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
        const float2 b = *reinterpret_cast<const float2*>(d.b1 + 8 * nt + 2 * t);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            acc1[mt][nt][0] = acc1[mt][nt][2] = b.x;
            acc1[mt][nt][1] = acc1[mt][nt][3] = b.y;
        }
    }

    // ---- layer 1: 32 -> 64 -------------------------------------------------------------------------------
#pragma unroll 1
    for (int j = 0; j < FEAT / 8; ++j) {
        uint32_t ah[2][4], al[2][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            const int col = xs_col(g, 8 * j + 2 * t);   // rows g, g+8, g+16, g+24 share (row & 3)
            const float2 v0 = *reinterpret_cast<const float2*>(X + (16 * mt + g) * XS + col);
            const float2 v1 = *reinterpret_cast<const float2*>(X + (16 * mt + g + 8) * XS + col);
            split_tf32<FAST>(v0.x, ah[mt][0], al[mt][0]);
            split_tf32<FAST>(v1.x, ah[mt][1], al[mt][1]);
            split_tf32<FAST>(v0.y, ah[mt][2], al[mt][2]);
            split_tf32<FAST>(v1.y, ah[mt][3], al[mt][3]);
        }
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const float2 w = *reinterpret_cast<const float2*>(d.w1 + (8 * nt + g) * W1M + 8 * j + 2 * t);
            uint32_t bh0, bl0, bh1, bl1;
            split_tf32<FAST>(w.x, bh0, bl0);
            split_tf32<FAST>(w.y, bh1, bl1);
            if (!FAST) {
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) mma_tf32(acc1[mt][nt], al[mt], bh0, bh1);
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) mma_tf32(acc1[mt][nt], ah[mt], bl0, bl1);
            }
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) mma_tf32(acc1[mt][nt], ah[mt], bh0, bh1);
        }
    }
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc1[mt][nt][i] = lrelu_folded(acc1[mt][nt][i]);

    // ---- layer-1 activations -> the warp's slab (X and the record slab behind it are dead by now), as
    // [32 points][64 channels] with the same granule swizzle: frees the 64 accumulator registers, so that layer 2
    // can keep all 8 n-tiles in flight (16 independent mma chains, like layer 1) instead of one n-tile at a time
    __syncwarp();   // every lane has read its X fragments
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
        const int col = xs_col(g, 8 * nt + 2 * t);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            *reinterpret_cast<float2*>(A1 + (16 * mt + g) * A1S + col) = make_float2(acc1[mt][nt][0], acc1[mt][nt][1]);
            *reinterpret_cast<float2*>(A1 + (16 * mt + g + 8) * A1S + col) = make_float2(acc1[mt][nt][2], acc1[mt][nt][3]);
        }
    }
    __syncwarp();

    // ---- layer 2: 64 -> 64 -------------------------------------------------------------------------------
    float acc2[2][8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
        const float2 b = *reinterpret_cast<const float2*>(d.b2 + 8 * nt + 2 * t);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            acc2[mt][nt][0] = acc2[mt][nt][2] = b.x;
            acc2[mt][nt][1] = acc2[mt][nt][3] = b.y;
        }
    }
#pragma unroll 1
    for (int j = 0; j < HID / 8; ++j) {
        uint32_t ah[2][4], al[2][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            const int col = xs_col(g, 8 * j + 2 * t);
            const float2 v0 = *reinterpret_cast<const float2*>(A1 + (16 * mt + g) * A1S + col);
            const float2 v1 = *reinterpret_cast<const float2*>(A1 + (16 * mt + g + 8) * A1S + col);
            split_tf32<FAST>(v0.x, ah[mt][0], al[mt][0]);
            split_tf32<FAST>(v1.x, ah[mt][1], al[mt][1]);
            split_tf32<FAST>(v0.y, ah[mt][2], al[mt][2]);
            split_tf32<FAST>(v1.y, ah[mt][3], al[mt][3]);
        }
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const float2 w = *reinterpret_cast<const float2*>(d.w2 + (8 * nt + g) * W2M + 8 * j + 2 * t);
            uint32_t bh0, bl0, bh1, bl1;
            split_tf32<FAST>(w.x, bh0, bl0);
            split_tf32<FAST>(w.y, bh1, bl1);
            if (!FAST) {
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) mma_tf32(acc2[mt][nt], al[mt], bh0, bh1);
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) mma_tf32(acc2[mt][nt], ah[mt], bl0, bl1);
            }
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) mma_tf32(acc2[mt][nt], ah[mt], bh0, bh1);
        }
    }

    // ---- layer 3 (64 -> 4): the finished n-tiles of layer 2 are its k-blocks, straight from the accumulators ---
    float m3[2][4], s3[2][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int i = 0; i < 4; ++i) m3[mt][i] = s3[mt][i] = 0.f;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
        const float2 w3 = *reinterpret_cast<const float2*>(d.w3 + g * W2M + 8 * nt + 2 * t);
        uint32_t bh0, bl0, bh1, bl1;
        split_tf32<FAST>(w3.x, bh0, bl0);
        split_tf32<FAST>(w3.y, bh1, bl1);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            uint32_t ah[4], al[4];
            split_tf32<FAST>(lrelu_folded(acc2[mt][nt][0]), ah[0], al[0]);
            split_tf32<FAST>(lrelu_folded(acc2[mt][nt][2]), ah[1], al[1]);
            split_tf32<FAST>(lrelu_folded(acc2[mt][nt][1]), ah[2], al[2]);
            split_tf32<FAST>(lrelu_folded(acc2[mt][nt][3]), ah[3], al[3]);
            mma3<FAST>(m3[mt], s3[mt], ah, al, bh0, bh1, bl0, bl1);
        }
    }
    const float2 b3 = *reinterpret_cast<const float2*>(d.b3 + 2 * t);
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
        o3[mt][0] = m3[mt][0] + s3[mt][0] + b3.x;
        o3[mt][1] = m3[mt][1] + s3[mt][1] + b3.y;
        o3[mt][2] = m3[mt][2] + s3[mt][2] + b3.x;
        o3[mt][3] = m3[mt][3] + s3[mt][3] + b3.y;
    }
}
