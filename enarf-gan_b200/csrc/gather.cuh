// Device helpers shared by the forward (render.cu) and backward (bwd.cu) fused kernels.
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"

__device__ __forceinline__ float lrelu_s2(float x) { return (x > 0.f ? x : 0.2f * x) * 1.4142135623730951f; }

// linspace(0,1,Nc+1)[j] under the oracle's contract
__device__ __forceinline__ float unit_bin(int j, int Nc, float step) {
    return (j < (Nc + 1) / 2) ? cmul(step, (float)j) : csub(1.0f, cmul(step, (float)(Nc - j)));
}

// four consecutive channels of one texel
__device__ __forceinline__ float4 load_feat4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 load_feat4(const __nv_bfloat16* p) {
    uint2 r = __ldg(reinterpret_cast<const uint2*>(p));
    float4 v;
    v.x = __uint_as_float(r.x << 16);
    v.y = __uint_as_float(r.x & 0xffff0000u);
    v.z = __uint_as_float(r.y << 16);
    v.w = __uint_as_float(r.y & 0xffff0000u);
    return v;
}

__device__ __forceinline__ void fma4(float4& f, float w, const float4& v) {
    f.x = fmaf(w, v.x, f.x);
    f.y = fmaf(w, v.y, f.y);
    f.z = fmaf(w, v.z, f.z);
    f.w = fmaf(w, v.w, f.w);
}

// f += sum over the 12 texels of one (point, part) pair record; feat already points at this lane's 4 channels.
// All 12 loads are issued before the first FMA (the footprints are always in range: common.cuh Bilin).
template <typename T>
__device__ __forceinline__ void gather_pair(float4& f, const T* __restrict__ feat, const float* rec, int row_elems,
                                            float scale) {
    const float4* rp = reinterpret_cast<const float4*>(rec);
    const float4 o4 = rp[0];
    const unsigned off[3] = {__float_as_uint(o4.x), __float_as_uint(o4.y), __float_as_uint(o4.z)};   // never negative
    float4 v[12];
    if (row_elems == 256 * FEAT) {
        // the reference's 256 x 256 planes: the second row of a footprint is a compile-time offset from the first, so the
        // four loads of a plane share one address register (the row-pointer arithmetic was 3.5 % of the forward's instructions)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            ENARF_DBG_ASSERT((off[k] & 31u) == 0u);     // a texel offset is a multiple of the 32 channels
            const T* p0 = feat + off[k];
            v[4 * k + 0] = load_feat4(p0);
            v[4 * k + 1] = load_feat4(p0 + FEAT);
            v[4 * k + 2] = load_feat4(p0 + 256 * FEAT);
            v[4 * k + 3] = load_feat4(p0 + 256 * FEAT + FEAT);
        }
    } else {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            ENARF_DBG_ASSERT((off[k] & 31u) == 0u);
            const T* p0 = feat + off[k];
            const T* p1 = p0 + row_elems;
            v[4 * k + 0] = load_feat4(p0);
            v[4 * k + 1] = load_feat4(p0 + FEAT);
            v[4 * k + 2] = load_feat4(p1);
            v[4 * k + 3] = load_feat4(p1 + FEAT);
        }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float4 w4 = rp[1 + k];
        fma4(f, w4.x * scale, v[4 * k + 0]);
        fma4(f, w4.y * scale, v[4 * k + 1]);
        fma4(f, w4.z * scale, v[4 * k + 2]);
        fma4(f, w4.w * scale, v[4 * k + 3]);
    }
}
