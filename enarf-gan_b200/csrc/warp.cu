// Deformation-field warp of the canonical feature planes (the DSO / "denarf" tri-plane producer,
// reference models/narf.py:39-58): for sample b and plane k the StyleGAN2 flow head emits a 2-channel
// offset field; the canonical 32-channel plane is resampled at pixel centre + offset (F.grid_sample,
// bilinear, zero padding, align_corners=False).  The reference then concatenates, and its sampler pads,
// permutes and reshapes the result on every call (libraries/triplane/sampling.py:96-99).
//
// Here the canonical planes are kept channels-last (3,H,W,32) and the kernel writes the renderer's own
// input layout (B,3,H,W,32) in fp32 or bf16 directly, so a warped tri-plane never exists as NCHW and is
// never re-packed; the backward takes the renderer's channels-last gradient as it is.
//
// Mapping: 8 lanes per output texel, lane l owns channels 4l..4l+3 -> every texel read is one 128-B line,
// every texel write one 128-B (fp32) / 64-B (bf16) line.  Flows are smooth, so the four source texels of
// neighbouring outputs overlap and come from L1/L2; compulsory HBM traffic per sample is the 8 B flow read
// and the output line: (8 + 128) B x 3*H*W fp32 = 26.7 MB, plus 25 MB of canonical planes once per launch.
#include <cuda_bf16.h>

#include "common.cuh"

namespace {

struct Foot {      // bilinear footprint of one output texel in the canonical plane
    int x0, y0;    // north-west source texel (may be outside the image)
    float wx1, wy1;
};

// (pixel centre + flow) / (size/2) - 1, then grid_sample's unnormalisation ((g + 1) * size - 1) / 2:
// the same fp32 operations as narf.py:49 followed by ATen's grid_sampler_unnormalize, none contracted.
__device__ __forceinline__ float source_coord(int pix, float flow, float half, float size) {
    const float g = __fsub_rn(__fdiv_rn(__fadd_rn((float)pix + 0.5f, flow), half), 1.f);
    return __fmul_rn(__fsub_rn(__fmul_rn(__fadd_rn(g, 1.f), size), 1.f), 0.5f);
}

__device__ __forceinline__ Foot footprint(int x, int y, float fx, float fy, int H, int W) {
    const float ix = source_coord(x, fx, 0.5f * W, (float)W);
    const float iy = source_coord(y, fy, 0.5f * H, (float)H);
    // clamp far-away samples so the int conversion is defined; anything beyond one texel outside is all zeros
    const float cx = fminf(fmaxf(ix, -2.f), (float)W + 1.f), cy = fminf(fmaxf(iy, -2.f), (float)H + 1.f);
    Foot f;
    const float flx = floorf(cx), fly = floorf(cy);
    f.x0 = (int)flx;
    f.y0 = (int)fly;
    f.wx1 = cx - flx;
    f.wy1 = cy - fly;
    return f;
}

__device__ __forceinline__ void store4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void store4(__nv_bfloat16* p, float4 v) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&lo);
    u.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(p) = u;
}
__device__ __forceinline__ float4 load4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 load4(const __nv_bfloat16* p) {
    const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
    const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162*>(&u.x);
    const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
    const float2 a = __bfloat1622float2(lo), b = __bfloat1622float2(hi);
    return make_float4(a.x, a.y, b.x, b.y);
}

__device__ __forceinline__ void red4(float* addr, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// canon (3,H,W,32) fp32; flow (B,6,H,W): channel 2k = x offset, 2k+1 = y offset of plane k (narf.py:46:
// reshape(bs*3, 2, H, W)); blockIdx.z = b*3 + k.
// Work decomposition (forward and backward): a CTA owns a strip of 32 columns x WARP_ROWS rows of one (sample, plane);
// warp w walks columns 4w..4w+3 down the rows.  Flows are smooth, so the source row y0+1 read for output row y is the
// row y0 of output row y+1 and is still in L1: ~1.1 source lines come from L2 per output texel instead of ~2.1 with
// row-major CTAs (which made the kernel L2-bandwidth-bound).  The next row's flow is requested one row ahead so the
// flow -> address -> texel chain costs one memory round trip per row, not two.
#define WARP_ROWS 16
template <typename T>
__global__ void __launch_bounds__(256) warp_fwd_kernel(const float* __restrict__ canon, const float* __restrict__ flow,
                                                       int H, int W, T* __restrict__ out) {
    const int HW = H * W;
    const int sub = threadIdx.x & 7;
    const int x = blockIdx.x * 32 + (threadIdx.x >> 3);
    const int y_begin = blockIdx.y * WARP_ROWS, y_end = min(H, y_begin + WARP_ROWS);
    const int bk = blockIdx.z, k = bk % 3;
    if (x >= W) return;
    const float* fl = flow + (size_t)(2 * bk) * HW + x;
    const float* src = canon + (size_t)k * HW * FEAT + 4 * sub;
    T* dst = out + ((size_t)bk * HW + x) * FEAT + 4 * sub;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    float fx = __ldg(fl + (size_t)y_begin * W), fy = __ldg(fl + HW + (size_t)y_begin * W);
    for (int y = y_begin; y < y_end; ++y) {
        const Foot f = footprint(x, y, fx, fy, H, W);
        if (y + 1 < y_end) {
            fx = __ldg(fl + (size_t)(y + 1) * W);
            fy = __ldg(fl + HW + (size_t)(y + 1) * W);
        }
        const bool xin0 = (unsigned)f.x0 < (unsigned)W, xin1 = (unsigned)(f.x0 + 1) < (unsigned)W;
        const bool yin0 = (unsigned)f.y0 < (unsigned)H, yin1 = (unsigned)(f.y0 + 1) < (unsigned)H;
        const size_t o00 = ((size_t)f.y0 * W + f.x0) * FEAT;
        // all four loads issued before the first use
        const float4 v00 = (yin0 && xin0) ? load4(src + o00) : z;
        const float4 v01 = (yin0 && xin1) ? load4(src + o00 + FEAT) : z;
        const float4 v10 = (yin1 && xin0) ? load4(src + o00 + (size_t)W * FEAT) : z;
        const float4 v11 = (yin1 && xin1) ? load4(src + o00 + (size_t)W * FEAT + FEAT) : z;
        const float wx0 = 1.f - f.wx1, wy0 = 1.f - f.wy1;
        const float w00 = wx0 * wy0, w01 = f.wx1 * wy0, w10 = wx0 * f.wy1, w11 = f.wx1 * f.wy1;
        float4 r;
        r.x = v00.x * w00 + v01.x * w01 + v10.x * w10 + v11.x * w11;
        r.y = v00.y * w00 + v01.y * w01 + v10.y * w10 + v11.y * w11;
        r.z = v00.z * w00 + v01.z * w01 + v10.z * w10 + v11.z * w11;
        r.w = v00.w * w00 + v01.w * w01 + v10.w * w10 + v11.w * w11;
        store4(dst + (size_t)y * W * FEAT, r);
    }
}

__device__ __forceinline__ float dot4(float4 a, float4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
__device__ __forceinline__ float4 scale4(float4 a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }

// d_out (B,3,H,W,32) in G (fp32 or bf16) -> d_canon (3,H,W,32) fp32 accumulated with 128-bit reductions,
// d_flow (B,6,H,W) stored.  d(source index)/d(flow) = (size/2) / (size/2) = 1 on both axes.
template <typename G>
__global__ void __launch_bounds__(256) warp_bwd_kernel(const G* __restrict__ d_out, const float* __restrict__ canon,
                                                       const float* __restrict__ flow, int H, int W,
                                                       float* __restrict__ d_canon, float* __restrict__ d_flow) {
    const int HW = H * W;
    const int sub = threadIdx.x & 7;
    const int x = blockIdx.x * 32 + (threadIdx.x >> 3);
    const int y_begin = blockIdx.y * WARP_ROWS, y_end = min(H, y_begin + WARP_ROWS);
    const int bk = blockIdx.z, k = bk % 3;
    if (x >= W) return;   // W % 4 == 0 (checked by the host): a warp's 4 texels are all inside or all outside
    const float* fl = flow + (size_t)(2 * bk) * HW + x;
    const G* gsrc = d_out + ((size_t)bk * HW + x) * FEAT + 4 * sub;
    const size_t base = (size_t)k * HW * FEAT + 4 * sub;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    float fx = __ldg(fl + (size_t)y_begin * W), fy = __ldg(fl + HW + (size_t)y_begin * W);
    for (int y = y_begin; y < y_end; ++y) {
        const Foot f = footprint(x, y, fx, fy, H, W);
        if (y + 1 < y_end) {
            fx = __ldg(fl + (size_t)(y + 1) * W);
            fy = __ldg(fl + HW + (size_t)(y + 1) * W);
        }
        const float4 g = load4(gsrc + (size_t)y * W * FEAT);
        const bool xin0 = (unsigned)f.x0 < (unsigned)W, xin1 = (unsigned)(f.x0 + 1) < (unsigned)W;
        const bool yin0 = (unsigned)f.y0 < (unsigned)H, yin1 = (unsigned)(f.y0 + 1) < (unsigned)H;
        const size_t o00 = base + ((size_t)f.y0 * W + f.x0) * FEAT;
        const size_t o01 = o00 + FEAT, o10 = o00 + (size_t)W * FEAT, o11 = o10 + FEAT;
        const bool in00 = yin0 && xin0, in01 = yin0 && xin1, in10 = yin1 && xin0, in11 = yin1 && xin1;
        const float wx0 = 1.f - f.wx1, wy0 = 1.f - f.wy1;
        if (d_flow) {
            const float4 v00 = in00 ? load4(canon + o00) : z;
            const float4 v01 = in01 ? load4(canon + o01) : z;
            const float4 v10 = in10 ? load4(canon + o10) : z;
            const float4 v11 = in11 ? load4(canon + o11) : z;
            const float d00 = dot4(g, v00), d01 = dot4(g, v01), d10 = dot4(g, v10), d11 = dot4(g, v11);
            float gx = (d01 - d00) * wy0 + (d11 - d10) * f.wy1;
            float gy = (d10 - d00) * wx0 + (d11 - d01) * f.wx1;
#pragma unroll
            for (int m = 1; m < 8; m <<= 1) {
                gx += __shfl_xor_sync(0xffffffffu, gx, m);
                gy += __shfl_xor_sync(0xffffffffu, gy, m);
            }
            if (sub == 0) {
                float* o = d_flow + (size_t)(2 * bk) * HW + (size_t)y * W + x;
                o[0] = gx;
                o[HW] = gy;
            }
        }
        if (d_canon) {
            if (in00) red4(d_canon + o00, scale4(g, wx0 * wy0));
            if (in01) red4(d_canon + o01, scale4(g, f.wx1 * wy0));
            if (in10) red4(d_canon + o10, scale4(g, wx0 * f.wy1));
            if (in11) red4(d_canon + o11, scale4(g, f.wx1 * f.wy1));
        }
    }
}

}  // namespace

extern "C" int enarf_warp_planes_fwd(const float* canon_feat, const float* flow, int B, int H, int W, void* feat_out,
                                     int out_dtype, void* stream) {
    ENARF_REQUIRE(canon_feat && flow && feat_out && B > 0 && H > 0 && W > 0, ENARF_ERR_INVALID_ARG,
                  "enarf_warp_planes_fwd: bad arguments");
    ENARF_REQUIRE(W % 4 == 0, ENARF_ERR_UNSUPPORTED, "enarf_warp_planes_fwd: W=%d must be a multiple of 4", W);
    ENARF_REQUIRE(3LL * B <= 65535, ENARF_ERR_UNSUPPORTED, "enarf_warp_planes_fwd: batch %d too large", B);
    const dim3 grid((W + 31) / 32, (H + WARP_ROWS - 1) / WARP_ROWS, 3 * B);
    if (out_dtype == ENARF_DTYPE_F32)
        warp_fwd_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(canon_feat, flow, H, W, (float*)feat_out);
    else if (out_dtype == ENARF_DTYPE_BF16)
        warp_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(canon_feat, flow, H, W,
                                                                               (__nv_bfloat16*)feat_out);
    else
        ENARF_REQUIRE(false, ENARF_ERR_INVALID_ARG, "enarf_warp_planes_fwd: bad out_dtype %d", out_dtype);
    ENARF_CHECK_LAUNCH("enarf_warp_planes_fwd");
    return ENARF_OK;
}

extern "C" int enarf_warp_planes_bwd(const void* d_feat, int d_feat_dtype, const float* canon_feat, const float* flow,
                                     int B, int H, int W, float* d_canon_feat, float* d_flow, void* stream) {
    ENARF_REQUIRE(d_feat && canon_feat && flow && (d_canon_feat || d_flow) && B > 0 && H > 0 && W > 0,
                  ENARF_ERR_INVALID_ARG, "enarf_warp_planes_bwd: bad arguments");
    ENARF_REQUIRE(W % 4 == 0, ENARF_ERR_UNSUPPORTED, "enarf_warp_planes_bwd: W=%d must be a multiple of 4", W);
    ENARF_REQUIRE(3LL * B <= 65535, ENARF_ERR_UNSUPPORTED, "enarf_warp_planes_bwd: batch %d too large", B);
    const dim3 grid((W + 31) / 32, (H + WARP_ROWS - 1) / WARP_ROWS, 3 * B);
    if (d_feat_dtype == ENARF_DTYPE_F32)
        warp_bwd_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)d_feat, canon_feat, flow, H, W,
                                                                       d_canon_feat, d_flow);
    else if (d_feat_dtype == ENARF_DTYPE_BF16)
        warp_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)d_feat, canon_feat,
                                                                               flow, H, W, d_canon_feat, d_flow);
    else
        ENARF_REQUIRE(false, ENARF_ERR_INVALID_ARG, "enarf_warp_planes_bwd: bad d_feat_dtype %d", d_feat_dtype);
    ENARF_CHECK_LAUNCH("enarf_warp_planes_bwd");
    return ENARF_OK;
}
