// Device helpers shared by the forward (render.cu) and backward (bwd.cu) fused kernels.
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"

__device__ __forceinline__ float lrelu_s2(float x) { return (x > 0.f ? x : 0.2f * x) * 1.4142135623730951f; }

// linspace(0,1,Nc+1)[j] under the oracle's contract
__device__ __forceinline__ float unit_bin(int j, int Nc, float step) {
    return (j < (Nc + 1) / 2) ? cmul(step, (float)j) : csub(1.0f, cmul(step, (float)(Nc - j)));
}

template <typename T>
__device__ __forceinline__ float4 load_feat4(const T* base, long long elem);
template <>
__device__ __forceinline__ float4 load_feat4<float>(const float* base, long long elem) {
    return __ldg(reinterpret_cast<const float4*>(base + elem));
}
template <>
__device__ __forceinline__ float4 load_feat4<__nv_bfloat16>(const __nv_bfloat16* base, long long elem) {
    uint2 r = __ldg(reinterpret_cast<const uint2*>(base + elem));
    float4 v;
    v.x = __uint_as_float(r.x << 16);
    v.y = __uint_as_float(r.x & 0xffff0000u);
    v.z = __uint_as_float(r.y << 16);
    v.w = __uint_as_float(r.y & 0xffff0000u);
    return v;
}

// Gather for one point by an 8-lane group: lane l owns feature channels 4l..4l+3.
template <typename T>
__device__ __forceinline__ float4 gather_point(const enarf_scene_t& sc, const float* s_part, int b, uint32_t m, float x,
                                               float y, float z, int lane8) {
    float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
    const int H = sc.H, W = sc.W;
    const T* feat = reinterpret_cast<const T*>(sc.feat_planes) + (size_t)b * 3 * H * W * FEAT;
    const float* prob = sc.prob_planes + (size_t)b * sc.prob_batch_stride;
    while (m) {
        const int p = __ffs(m) - 1;
        m &= m - 1;
        float c[3];
        part_canonical(s_part + p * PART_STRIDE, x, y, z, c[0], c[1], c[2]);
        Bilin bl[3];
        float wp = 1.f;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            bl[k] = bilin_setup(c[k], c[(k + 1) % 3], W, H);
            wp *= sigmoidf_(bilin_fetch1(prob + (size_t)(3 * p + k) * H * W, bl[k], W));
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const long long base = ((long long)k * H * W + (long long)bl[k].y0 * W + bl[k].x0) * FEAT + lane8 * 4;
            const float w00 = bl[k].w00 * wp, w01 = bl[k].w01 * wp, w10 = bl[k].w10 * wp, w11 = bl[k].w11 * wp;
            float4 v;
            if (w00 != 0.f) { v = load_feat4<T>(feat, base); f.x = fmaf(w00, v.x, f.x); f.y = fmaf(w00, v.y, f.y); f.z = fmaf(w00, v.z, f.z); f.w = fmaf(w00, v.w, f.w); }
            if (w01 != 0.f) { v = load_feat4<T>(feat, base + FEAT); f.x = fmaf(w01, v.x, f.x); f.y = fmaf(w01, v.y, f.y); f.z = fmaf(w01, v.z, f.z); f.w = fmaf(w01, v.w, f.w); }
            if (w10 != 0.f) { v = load_feat4<T>(feat, base + (long long)W * FEAT); f.x = fmaf(w10, v.x, f.x); f.y = fmaf(w10, v.y, f.y); f.z = fmaf(w10, v.z, f.z); f.w = fmaf(w10, v.w, f.w); }
            if (w11 != 0.f) { v = load_feat4<T>(feat, base + (long long)(W + 1) * FEAT); f.x = fmaf(w11, v.x, f.x); f.y = fmaf(w11, v.y, f.y); f.z = fmaf(w11, v.z, f.z); f.w = fmaf(w11, v.w, f.w); }
        }
    }
    return f;
}
