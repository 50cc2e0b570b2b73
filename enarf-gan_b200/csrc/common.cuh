// Shared device/host helpers for libenarf_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/enarf_b200.h"

// Development build (ENARF_NVCC_EXTRA=-DENARF_DEBUG_BOUNDS): device-side asserts on the indices that address shared
// and global memory (compute-sanitizer is not available on the GPU pool this library is developed on).  A failed
// assert traps the kernel; the next ABI call / synchronisation reports the CUDA error.
#ifdef ENARF_DEBUG_BOUNDS
#include <assert.h>
#define ENARF_DBG_ASSERT(c) assert(c)
#else
#define ENARF_DBG_ASSERT(c) ((void)0)
#endif

#define FEAT ENARF_FEAT_DIM
#define HID ENARF_HIDDEN
#define N_PROBE 32  // depth probes per ray (reference: libraries/NeRF/rendering.py:18)

// ---- host side: thread-local error string, argument / launch checks ---------------------------
void enarf_set_error(const char* fmt, ...);

#define ENARF_REQUIRE(cond, code, ...)   \
    do {                                 \
        if (!(cond)) {                   \
            enarf_set_error(__VA_ARGS__); \
            return (code);               \
        }                                \
    } while (0)

#define ENARF_CHECK_LAUNCH(name)                                                     \
    do {                                                                             \
        cudaError_t e__ = cudaGetLastError();                                        \
        if (e__ != cudaSuccess) {                                                    \
            enarf_set_error("%s: CUDA error: %s", (name), cudaGetErrorString(e__)); \
            return ENARF_ERR_CUDA;                                                   \
        }                                                                            \
    } while (0)

int enarf_check_scene(const enarf_scene_t* sc, const char* who);
int enarf_sm_count();
// Per-launch dynamic-tile-scheduler counter: a slot of a static device array, zeroed on `stream` before use.
// Returns nullptr on failure (error string set).
int* enarf_tile_counter(cudaStream_t stream);

// ---- contract arithmetic: every product / sum rounded separately, never contracted ------------
// (oracle/enarf_oracle.py "ARITHMETIC CONTRACT"; these feed the bit-exact masks and indices)
__device__ __forceinline__ float cmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float cadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float csub(float a, float b) { return __fsub_rn(a, b); }
// a*(1-t) + b*t                      (rendering.py:126-129,198-200)
__device__ __forceinline__ float clerp(float a, float b, float t) {
    return cadd(cmul(a, csub(1.0f, t)), cmul(b, t));
}
// (m0*v0 + m1*v1) + m2*v2
__device__ __forceinline__ float cdot3(float m0, float m1, float m2, float v0, float v1, float v2) {
    return cadd(cadd(cmul(m0, v0), cmul(m1, v1)), cmul(m2, v2));
}

// Part frames staged in shared memory: per part 28 floats.
//   [0..8]  R (row-major 3x3 of the part->camera frame)    [9..11] t
//   [12]    canonical scale                                 [13..21] Rc   [22..24] tc
#define PART_STRIDE 28

__device__ __forceinline__ void load_part_params(float* s_part, const enarf_scene_t& sc, int b, int tid, int nthreads) {
    for (int i = tid; i < sc.P * PART_STRIDE; i += nthreads) {
        int p = i / PART_STRIDE, j = i - p * PART_STRIDE;
        float v = 0.f;
        if (j < 12) {
            int r = (j < 9) ? j / 3 : j - 9, c = (j < 9) ? j % 3 : 3;
            v = sc.pose[((size_t)(b * sc.P + p) * 4 + r) * 4 + c];
        } else if (j == 12) {
            v = sc.cscale[b * sc.P + p];
        } else if (j < 25) {
            int jj = j - 13;
            int r = (jj < 9) ? jj / 3 : jj - 9, c = (jj < 9) ? jj % 3 : 3;
            v = sc.canon_pose[((size_t)p * 4 + r) * 4 + c];
        }
        s_part[i] = v;
    }
}

// local = R^T (x - t); |local| <= 1 on every axis           (models/narf.py:158-162, NeRF/utils.py:39-43)
__device__ __forceinline__ bool part_local(const float* q, float x, float y, float z, float& l0, float& l1, float& l2) {
    float d0 = csub(x, q[9]), d1 = csub(y, q[10]), d2 = csub(z, q[11]);
    l0 = cdot3(q[0], q[3], q[6], d0, d1, d2);
    l1 = cdot3(q[1], q[4], q[7], d0, d1, d2);
    l2 = cdot3(q[2], q[5], q[8], d0, d1, d2);
    return fabsf(l0) <= 1.f && fabsf(l1) <= 1.f && fabsf(l2) <= 1.f;
}

// canon = Rc (local * scale) + tc; valid iff in-cube and |canon| < 1 on every axis (narf.py:165-169,200-201)
__device__ __forceinline__ bool part_canonical(const float* q, float x, float y, float z, float& c0, float& c1, float& c2) {
    float l0, l1, l2;
    bool in = part_local(q, x, y, z, l0, l1, l2);
    float s0 = cmul(l0, q[12]), s1 = cmul(l1, q[12]), s2 = cmul(l2, q[12]);
    c0 = cadd(cdot3(q[13], q[14], q[15], s0, s1, s2), q[22]);
    c1 = cadd(cdot3(q[16], q[17], q[18], s0, s1, s2), q[23]);
    c2 = cadd(cdot3(q[19], q[20], q[21], s0, s1, s2), q[24]);
    return in && fabsf(c0) < 1.f && fabsf(c1) < 1.f && fabsf(c2) < 1.f;
}

// Bilinear footprint, zeros padding, align_corners=False (TriplaneSampler_kernel.cu:40-74 semantics).
// A footprint that hangs over an edge is shifted back inside the image: the texel that moves in takes the weight 0
// the out-of-range texel had, so all four addresses are always valid, loads need no predicate, and the result
// is the same as with zeros padding (0 * finite adds nothing).  Requires W, H >= 2.
struct Bilin {
    int x0, y0;          // top-left texel, 0 <= x0 <= W-2, 0 <= y0 <= H-2
    float w00, w01, w10, w11;  // weights of (y0,x0) (y0,x0+1) (y0+1,x0) (y0+1,x0+1)
};

__device__ __forceinline__ Bilin bilin_setup(float x, float y, int W, int H) {
    float ix = ((x + 1.f) * W - 1.f) * 0.5f;
    float iy = ((y + 1.f) * H - 1.f) * 0.5f;
    float fx = floorf(ix), fy = floorf(iy);
    int x0 = (int)fx, y0 = (int)fy;
    float ax = ix - fx, ay = iy - fy;
    float wx0 = (x0 >= 0 && x0 < W) ? 1.f - ax : 0.f, wx1 = (x0 + 1 >= 0 && x0 + 1 < W) ? ax : 0.f;
    float wy0 = (y0 >= 0 && y0 < H) ? 1.f - ay : 0.f, wy1 = (y0 + 1 >= 0 && y0 + 1 < H) ? ay : 0.f;
    if (x0 < 0) { wx0 = wx1; wx1 = 0.f; }
    if (x0 > W - 2) { wx1 = wx0; wx0 = 0.f; }
    if (y0 < 0) { wy0 = wy1; wy1 = 0.f; }
    if (y0 > H - 2) { wy1 = wy0; wy0 = 0.f; }
    Bilin b;
    b.x0 = min(max(x0, 0), W - 2);
    b.y0 = min(max(y0, 0), H - 2);
    b.w00 = wx0 * wy0;
    b.w01 = wx1 * wy0;
    b.w10 = wx0 * wy1;
    b.w11 = wx1 * wy1;
    return b;
}

// The same footprint from its two 1-D halves.  The three planes of a pair sample (c0,c1), (c1,c2), (c2,c0): every
// coordinate is the x of one plane and the y of another, so with square planes three 1-D set-ups serve all three
// footprints instead of six (identical arithmetic: bilin_setup(x, y, W, H) == bilin_from(lin_setup(x, W), lin_setup(y, H))).
struct Lin1 {
    int i0;
    float w0, w1;
};
__device__ __forceinline__ Lin1 lin_setup(float x, int N) {
    float ix = ((x + 1.f) * N - 1.f) * 0.5f;
    float fx = floorf(ix);
    int x0 = (int)fx;
    float ax = ix - fx;
    float wx0 = (x0 >= 0 && x0 < N) ? 1.f - ax : 0.f, wx1 = (x0 + 1 >= 0 && x0 + 1 < N) ? ax : 0.f;
    if (x0 < 0) { wx0 = wx1; wx1 = 0.f; }
    if (x0 > N - 2) { wx1 = wx0; wx0 = 0.f; }
    Lin1 l;
    l.i0 = min(max(x0, 0), N - 2);
    l.w0 = wx0;
    l.w1 = wx1;
    return l;
}
__device__ __forceinline__ Bilin bilin_from(const Lin1& lx, const Lin1& ly) {
    Bilin b;
    b.x0 = lx.i0;
    b.y0 = ly.i0;
    b.w00 = lx.w0 * ly.w0;
    b.w01 = lx.w1 * ly.w0;
    b.w10 = lx.w0 * ly.w1;
    b.w11 = lx.w1 * ly.w1;
    return b;
}
// footprints of the three planes of a pair at canonical coordinates c[0..2]
__device__ __forceinline__ void bilin_setup3(const float (&c)[3], int W, int H, Bilin (&bl)[3]) {
    Lin1 lx[3], ly[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) lx[i] = lin_setup(c[i], W);
    if (W == H) {
#pragma unroll
        for (int i = 0; i < 3; ++i) ly[i] = lx[i];
    } else {
#pragma unroll
        for (int i = 0; i < 3; ++i) ly[i] = lin_setup(c[i], H);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) bl[k] = bilin_from(lx[k], ly[(k + 1) % 3]);
}

// scalar (1-channel) bilinear fetch from a row-major H x W plane
__device__ __forceinline__ float bilin_fetch1(const float* __restrict__ plane, const Bilin& b, int W) {
    const float* r0 = plane + (b.y0 * W + b.x0);
    const float t00 = __ldg(r0), t01 = __ldg(r0 + 1), t10 = __ldg(r0 + W), t11 = __ldg(r0 + W + 1);
    return fmaf(b.w11, t11, fmaf(b.w10, t10, fmaf(b.w01, t01, b.w00 * t00)));
}

__device__ __forceinline__ float sigmoidf_(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }

// ---- Philox4x32-10 (Salmon et al. 2011); oracle twin: enarf_oracle.philox4x32 ------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t h0 = __umulhi(0xD2511F53u, c.x), l0 = 0xD2511F53u * c.x;
        uint32_t h1 = __umulhi(0xCD9E8D57u, c.z), l1 = 0xCD9E8D57u * c.z;
        c = make_uint4(h1 ^ c.y ^ k.x, l1, h0 ^ c.w ^ k.y, l0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}
__device__ __forceinline__ float u01(uint32_t w) { return (float)(w >> 8) * (1.0f / 16777216.0f); }

// Same predicate as part_canonical() with an early exit after the in-cube test (identical arithmetic, so the
// resulting bit is identical); used where the canonical coordinates themselves are not needed.
__device__ __forceinline__ bool part_valid(const float* q, float x, float y, float z) {
    float l0, l1, l2;
    if (!part_local(q, x, y, z, l0, l1, l2)) return false;
    float s0 = cmul(l0, q[12]), s1 = cmul(l1, q[12]), s2 = cmul(l2, q[12]);
    float c0 = cadd(cdot3(q[13], q[14], q[15], s0, s1, s2), q[22]);
    float c1 = cadd(cdot3(q[16], q[17], q[18], s0, s1, s2), q[23]);
    float c2 = cadd(cdot3(q[19], q[20], q[21], s0, s1, s2), q[24]);
    return fabsf(c0) < 1.f && fabsf(c1) < 1.f && fabsf(c2) < 1.f;
}

// Conservative ray/part pre-filter.  The segment start + u*(end-start), u in [0,1], is clipped against the
// part cube inflated to |local| <= 1 + CULL_MARGIN; a point whose u lies outside [u0,u1] cannot pass the
// exact in-cube test (the margin is ~1000x the worst fp32 rounding of the contract arithmetic at these
// magnitudes), so skipping the exact test there cannot change a mask bit.  Returns false if empty.
#define CULL_MARGIN 2e-3f
__device__ __forceinline__ bool ray_part_interval(const float* q, float sx, float sy, float sz, float ex, float ey,
                                                  float ez, float& u0, float& u1) {
    const float ax = sx - q[9], ay = sy - q[10], az = sz - q[11];
    const float bx = ex - sx, by = ey - sy, bz = ez - sz;
    u0 = -1e-3f;
    u1 = 1.f + 1e-3f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float A = q[i] * ax + q[3 + i] * ay + q[6 + i] * az;   // local coordinate at u = 0
        const float Bd = q[i] * bx + q[3 + i] * by + q[6 + i] * bz;  // d(local)/du
        const float lim = 1.f + CULL_MARGIN + 1e-5f * (fabsf(A) + fabsf(Bd));
        if (fabsf(Bd) < 1e-12f) {
            if (fabsf(A) > lim) return false;
        } else {
            const float inv = 1.f / Bd;
            float t0 = (-lim - A) * inv, t1 = (lim - A) * inv;
            if (t0 > t1) { const float t = t0; t0 = t1; t1 = t; }
            const float pad = 1e-5f * (fabsf(t0) + fabsf(t1)) + 1e-6f;
            u0 = fmaxf(u0, t0 - pad);
            u1 = fminf(u1, t1 + pad);
        }
    }
    return u0 <= u1;
}
