// Plane layout conversion and the stand-alone tri-plane sampler (the literal replacement of the
// reference's pybind module; cuda_extension/TriplaneSampler_kernel.cu:13-229 describes the op).
#include <cuda_bf16.h>

#include "common.cuh"

// ---- NCHW feature channels <-> channels-last (B,3,H,W,32) ---------------------------------------
// One CTA transposes a 32-channel x TPX-texel tile through shared memory: each warp reads 1 KB runs along x of one
// channel row (the rows are H*W floats apart, so short runs would touch a new DRAM page every 128 B), and writes
// 128-B texels along the channel axis; the 257-float row stride keeps both sides bank-conflict free.
#define TPX 256
template <typename T>
__global__ void __launch_bounds__(256) repack_kernel(const float* __restrict__ src, int C_total, int HW,
                                                     T* __restrict__ dst) {
    __shared__ float tile[32][TPX + 1];
    const int bk = blockIdx.y;  // b*3 + k
    const int b = bk / 3, k = bk - b * 3;
    const int x0 = blockIdx.x * TPX;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float* s = src + ((size_t)b * C_total + k * FEAT) * HW + x0;
    for (int c = warp; c < 32; c += 8)
#pragma unroll
        for (int j = 0; j < TPX / 32; ++j) {
            const int x = 32 * j + lane;
            tile[c][x] = (x0 + x < HW) ? __ldg(s + (size_t)c * HW + x) : 0.f;
        }
    __syncthreads();
    T* d = dst + ((size_t)bk * HW + x0) * FEAT;
    for (int p = warp; p < TPX; p += 8)
        if (x0 + p < HW) d[(size_t)p * FEAT + lane] = (T)tile[lane][p];
}

template <bool ACC>
__global__ void __launch_bounds__(256) unpack_grad_kernel(const float* __restrict__ d_feat, int C_total, int HW,
                                                          float* __restrict__ d_nchw) {
    __shared__ float tile[32][TPX + 1];
    const int bk = blockIdx.y;
    const int b = bk / 3, k = bk - b * 3;
    const int x0 = blockIdx.x * TPX;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float* s = d_feat + ((size_t)bk * HW + x0) * FEAT;
    for (int p = warp; p < TPX; p += 8) tile[lane][p] = (x0 + p < HW) ? __ldg(s + (size_t)p * FEAT + lane) : 0.f;
    __syncthreads();
    float* d = d_nchw + ((size_t)b * C_total + k * FEAT) * HW + x0;
    for (int c = warp; c < 32; c += 8)
#pragma unroll
        for (int j = 0; j < TPX / 32; ++j) {
            const int x = 32 * j + lane;
            if (x0 + x < HW) {
                float* o = d + (size_t)c * HW + x;
                *o = ACC ? *o + tile[c][x] : tile[c][x];
            }
        }
}

extern "C" int enarf_repack_planes(const float* tri_plane_nchw, int B, int C_total, int H, int W, void* feat_out,
                                   int out_dtype, void* stream) {
    ENARF_REQUIRE(tri_plane_nchw && feat_out && B > 0 && H > 0 && W > 0 && C_total >= 3 * FEAT, ENARF_ERR_INVALID_ARG,
                  "enarf_repack_planes: bad arguments (C_total=%d)", C_total);
    ENARF_REQUIRE(3 * B <= 65535, ENARF_ERR_UNSUPPORTED, "enarf_repack_planes: batch %d too large", B);
    const int HW = H * W;
    dim3 grid((HW + TPX - 1) / TPX, 3 * B);
    if (out_dtype == ENARF_DTYPE_F32)
        repack_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(tri_plane_nchw, C_total, HW, (float*)feat_out);
    else if (out_dtype == ENARF_DTYPE_BF16)
        repack_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(tri_plane_nchw, C_total, HW,
                                                                             (__nv_bfloat16*)feat_out);
    else
        ENARF_REQUIRE(false, ENARF_ERR_INVALID_ARG, "enarf_repack_planes: bad out_dtype %d", out_dtype);
    ENARF_CHECK_LAUNCH("enarf_repack_planes");
    return ENARF_OK;
}

extern "C" int enarf_unpack_plane_grads(const float* d_feat, int B, int C_total, int H, int W, int accumulate,
                                        float* d_tri_plane_nchw, void* stream) {
    ENARF_REQUIRE(d_feat && d_tri_plane_nchw && B > 0 && H > 0 && W > 0 && C_total >= 3 * FEAT, ENARF_ERR_INVALID_ARG,
                  "enarf_unpack_plane_grads: bad arguments");
    ENARF_REQUIRE(3 * B <= 65535, ENARF_ERR_UNSUPPORTED, "enarf_unpack_plane_grads: batch %d too large", B);
    const int HW = H * W;
    dim3 grid((HW + TPX - 1) / TPX, 3 * B);
    if (accumulate)
        unpack_grad_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(d_feat, C_total, HW, d_tri_plane_nchw);
    else
        unpack_grad_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(d_feat, C_total, HW, d_tri_plane_nchw);
    ENARF_CHECK_LAUNCH("enarf_unpack_plane_grads");
    return ENARF_OK;
}

// ---- stand-alone 3-plane sampler (NCHW in, as the reference's op) --------------------------------
// Source-index computation with its derivative (grid_sample semantics: unnormalise, then border clip or
// reflection).  `mult` receives d(index)/d(coord).
__device__ __forceinline__ float reflect_coord(float in, int twice_low, int twice_high, float& mult) {
    if (twice_low == twice_high) {
        mult = 0.f;
        return 0.f;
    }
    const float mn = 0.5f * twice_low, span = 0.5f * (twice_high - twice_low);
    in -= mn;
    float sgn = 1.f;
    if (in < 0.f) {
        in = -in;
        sgn = -1.f;
    }
    const float extra = fmodf(in, span);
    const int flips = (int)floorf(in / span);
    if (flips & 1) {
        mult *= -sgn;
        return span - extra + mn;
    }
    mult *= sgn;
    return extra + mn;
}
__device__ __forceinline__ float clip_coord(float in, int size, float& mult) {
    if (in <= 0.f) {
        mult = 0.f;
        return 0.f;
    }
    const float mx = (float)(size - 1);
    if (in >= mx) {
        mult = 0.f;
        return mx;
    }
    return in;
}
__device__ __forceinline__ float source_index(float coord, int size, int padding, bool align, float& mult) {
    float v;
    if (align) {
        mult = 0.5f * (size - 1);
        v = (coord + 1.f) * 0.5f * (size - 1);
    } else {
        mult = 0.5f * size;
        v = ((coord + 1.f) * size - 1.f) * 0.5f;
    }
    if (padding == 1) {
        v = clip_coord(v, size, mult);
    } else if (padding == 2) {
        v = align ? reflect_coord(v, 0, 2 * (size - 1), mult) : reflect_coord(v, -1, 2 * size - 1, mult);
        v = clip_coord(v, size, mult);
    }
    return v;
}

struct Tap {  // up to four texels of one plane with their weights and the weight derivatives
    int off[4];
    float w[4], dwx[4], dwy[4];
    int cnt;
};

__device__ __forceinline__ Tap make_tap(float gx, float gy, int H, int W, int interp, int padding, bool align) {
    Tap t;
    t.cnt = 0;
    float mx, my;
    const float ix = source_index(gx, W, padding, align, mx);
    const float iy = source_index(gy, H, padding, align, my);
    if (interp == 0) {
        const float fx = floorf(ix), fy = floorf(iy);
        const int x0 = (int)fx, y0 = (int)fy;
        const float ax = ix - fx, ay = iy - fy;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int dx = c & 1, dy = c >> 1;
            const int x = x0 + dx, y = y0 + dy;
            if (x >= 0 && x < W && y >= 0 && y < H) {
                const float wx = dx ? ax : 1.f - ax, wy = dy ? ay : 1.f - ay;
                t.off[t.cnt] = y * W + x;
                t.w[t.cnt] = wx * wy;
                t.dwx[t.cnt] = (dx ? 1.f : -1.f) * wy * mx;
                t.dwy[t.cnt] = (dy ? 1.f : -1.f) * wx * my;
                ++t.cnt;
            }
        }
    } else {
        const int x = (int)nearbyintf(ix), y = (int)nearbyintf(iy);
        if (x >= 0 && x < W && y >= 0 && y < H) {
            t.off[0] = y * W + x;
            t.w[0] = 1.f;
            t.dwx[0] = t.dwy[0] = 0.f;
            t.cnt = 1;
        }
    }
    return t;
}

#define TS_CH 8  // channels per thread

__global__ void __launch_bounds__(256) triplane_fwd_kernel(const float* __restrict__ input,
                                                           const float* __restrict__ grid, int C, int H, int W,
                                                           long long npts_per_n, int interp, int padding, int align,
                                                           float* __restrict__ output) {
    const long long pt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int nb = blockIdx.z, c0 = blockIdx.y * TS_CH;
    if (pt >= npts_per_n) return;
    const float* g = grid + ((size_t)nb * npts_per_n + pt) * 3;
    const float gv[3] = {g[0], g[1], g[2]};
    float acc[TS_CH];
#pragma unroll
    for (int c = 0; c < TS_CH; ++c) acc[c] = 0.f;
    const size_t HW = (size_t)H * W;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const Tap t = make_tap(gv[k], gv[(k + 1) % 3], H, W, interp, padding, align != 0);
        const float* base = input + ((size_t)nb * 3 * C + (size_t)k * C + c0) * HW;
        for (int i = 0; i < t.cnt; ++i) {
#pragma unroll
            for (int c = 0; c < TS_CH; ++c)
                if (c0 + c < C) acc[c] = fmaf(t.w[i], __ldg(base + c * HW + t.off[i]), acc[c]);
        }
    }
#pragma unroll
    for (int c = 0; c < TS_CH; ++c)
        if (c0 + c < C) output[((size_t)nb * C + c0 + c) * npts_per_n + pt] = acc[c];
}

__global__ void __launch_bounds__(256) triplane_bwd_kernel(const float* __restrict__ grad_output,
                                                           const float* __restrict__ input,
                                                           const float* __restrict__ grid, int C, int H, int W,
                                                           long long npts_per_n, int interp, int padding, int align,
                                                           float* __restrict__ grad_input,
                                                           float* __restrict__ grad_grid) {
    const long long pt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int nb = blockIdx.z, c0 = blockIdx.y * TS_CH;
    if (pt >= npts_per_n) return;
    const float* g = grid + ((size_t)nb * npts_per_n + pt) * 3;
    const float gv[3] = {g[0], g[1], g[2]};
    float go[TS_CH];
#pragma unroll
    for (int c = 0; c < TS_CH; ++c)
        go[c] = (c0 + c < C) ? grad_output[((size_t)nb * C + c0 + c) * npts_per_n + pt] : 0.f;
    const size_t HW = (size_t)H * W;
    float gg[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const Tap t = make_tap(gv[k], gv[(k + 1) % 3], H, W, interp, padding, align != 0);
        const size_t plane = ((size_t)nb * 3 * C + (size_t)k * C + c0) * HW;
        float gx = 0.f, gy = 0.f;
        for (int i = 0; i < t.cnt; ++i) {
#pragma unroll
            for (int c = 0; c < TS_CH; ++c) {
                if (c0 + c >= C) continue;
                if (grad_input) atomicAdd(grad_input + plane + c * HW + t.off[i], t.w[i] * go[c]);
                if (grad_grid) {
                    const float v = __ldg(input + plane + c * HW + t.off[i]) * go[c];
                    gx = fmaf(v, t.dwx[i], gx);
                    gy = fmaf(v, t.dwy[i], gy);
                }
            }
        }
        gg[k] += gx;
        gg[(k + 1) % 3] += gy;
    }
    if (grad_grid) {
        float* o = grad_grid + ((size_t)nb * npts_per_n + pt) * 3;
#pragma unroll
        for (int k = 0; k < 3; ++k)
            if (gg[k] != 0.f) atomicAdd(o + k, gg[k]);
    }
}

// ---- channels-last fast path of the stand-alone sampler (C == 32, bilinear) ------------------------------------
// With NCHW planes every (channel, texel pair) is its own 32-B sector: 192 sectors = 6 KB of L2 traffic per point for
// 1.5 KB of useful data, which is what bounds the reference's kernel (TriplaneSampler_kernel.cu:36-74) and the generic
// kernel above.  On channels-last planes (enarf_repack_planes: 0.04 ms for a 96 x 256 x 256 tri-plane) a texel is one
// 128-B line: 8 lanes per point read it as one float4 each, 12 lines per point, all in flight before the first FMA.
// A warp takes 32 consecutive points; results go through a [channel][point] shared-memory tile so the NCHW output
// (and grad_output in the backward) moves as 128-B rows.
struct Tap4 {  // four texels of one plane, always addressable (clamped); outside texels carry weight 0
    int off[4];
    float w[4], dwx[4], dwy[4];
};

__device__ __forceinline__ Tap4 make_tap4(float gx, float gy, int H, int W, int padding, bool align) {
    Tap4 t;
    float mx, my;
    const float ix = source_index(gx, W, padding, align, mx);
    const float iy = source_index(gy, H, padding, align, my);
    // far-away coordinates: clamp before the int conversion (anything beyond one texel outside has weight 0)
    const float cx = fminf(fmaxf(ix, -2.f), (float)W + 1.f), cy = fminf(fmaxf(iy, -2.f), (float)H + 1.f);
    const float fx = floorf(cx), fy = floorf(cy);
    const int x0 = (int)fx, y0 = (int)fy;
    const float ax = cx - fx, ay = cy - fy;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const int dx = c & 1, dy = c >> 1;
        const int x = x0 + dx, y = y0 + dy;
        const bool in = x >= 0 && x < W && y >= 0 && y < H;
        const float wx = dx ? ax : 1.f - ax, wy = dy ? ay : 1.f - ay;
        t.off[c] = min(max(y, 0), H - 1) * W + min(max(x, 0), W - 1);
        t.w[c] = in ? wx * wy : 0.f;
        t.dwx[c] = in ? (dx ? 1.f : -1.f) * wy * mx : 0.f;
        t.dwy[c] = in ? (dy ? 1.f : -1.f) * wx * my : 0.f;
    }
    return t;
}

__device__ __forceinline__ void ts_red4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(256) triplane_cl_fwd_kernel(const float* __restrict__ feat,
                                                              const float* __restrict__ grid, int H, int W,
                                                              long long npts, int padding, int align,
                                                              float* __restrict__ output) {
    __shared__ float tile[8][FEAT][33];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, sub = lane & 7, q = lane >> 3;
    const int nb = blockIdx.y;
    const size_t HW = (size_t)H * W;
    const float* fb = feat + (size_t)nb * 3 * HW * FEAT + 4 * sub;
    const float* gb = grid + (size_t)nb * npts * 3;
    float* ob = output + (size_t)nb * FEAT * npts;
    for (long long base = ((long long)blockIdx.x * 8 + warp) * 32; base < npts; base += (long long)gridDim.x * 256) {
#pragma unroll 2
        for (int j = 0; j < 8; ++j) {
            const long long p = base + 4 * j + q;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            if (p < npts) {
                const float gv[3] = {__ldg(gb + p * 3), __ldg(gb + p * 3 + 1), __ldg(gb + p * 3 + 2)};
                Tap4 t[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) t[k] = make_tap4(gv[k], gv[(k + 1) % 3], H, W, padding, align != 0);
                float4 v[12];
#pragma unroll
                for (int k = 0; k < 3; ++k)
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        v[4 * k + c] = __ldg(reinterpret_cast<const float4*>(fb + ((size_t)k * HW + t[k].off[c]) * FEAT));
#pragma unroll
                for (int k = 0; k < 3; ++k)
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const float w = t[k].w[c];
                        acc.x = fmaf(w, v[4 * k + c].x, acc.x);
                        acc.y = fmaf(w, v[4 * k + c].y, acc.y);
                        acc.z = fmaf(w, v[4 * k + c].z, acc.z);
                        acc.w = fmaf(w, v[4 * k + c].w, acc.w);
                    }
            }
            tile[warp][4 * sub + 0][4 * j + q] = acc.x;
            tile[warp][4 * sub + 1][4 * j + q] = acc.y;
            tile[warp][4 * sub + 2][4 * j + q] = acc.z;
            tile[warp][4 * sub + 3][4 * j + q] = acc.w;
        }
        __syncwarp();
        if (base + lane < npts) {
#pragma unroll 8
            for (int c = 0; c < FEAT; ++c) __stcs(ob + (size_t)c * npts + base + lane, tile[warp][c][lane]);
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(256) triplane_cl_bwd_kernel(const float* __restrict__ grad_output,
                                                              const float* __restrict__ feat,
                                                              const float* __restrict__ grid, int H, int W,
                                                              long long npts, int padding, int align,
                                                              float* __restrict__ d_feat, float* __restrict__ grad_grid) {
    __shared__ float tile[8][FEAT][33];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, sub = lane & 7, q = lane >> 3;
    const int nb = blockIdx.y;
    const size_t HW = (size_t)H * W;
    const size_t fo = (size_t)nb * 3 * HW * FEAT + 4 * sub;
    const float* gb = grid + (size_t)nb * npts * 3;
    const float* gob = grad_output + (size_t)nb * FEAT * npts;
    for (long long base = ((long long)blockIdx.x * 8 + warp) * 32; base < npts; base += (long long)gridDim.x * 256) {
        {
            const bool live = base + lane < npts;
#pragma unroll 8
            for (int c = 0; c < FEAT; ++c) tile[warp][c][lane] = live ? __ldcs(gob + (size_t)c * npts + base + lane) : 0.f;
        }
        __syncwarp();
        for (int j = 0; j < 8; ++j) {
            const long long p = base + 4 * j + q;
            const bool live = p < npts;
            const long long pc = live ? p : npts - 1;
            const float4 g = make_float4(tile[warp][4 * sub + 0][4 * j + q], tile[warp][4 * sub + 1][4 * j + q],
                                         tile[warp][4 * sub + 2][4 * j + q], tile[warp][4 * sub + 3][4 * j + q]);
            const float gv[3] = {__ldg(gb + pc * 3), __ldg(gb + pc * 3 + 1), __ldg(gb + pc * 3 + 2)};
            float gg[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const Tap4 t = make_tap4(gv[k], gv[(k + 1) % 3], H, W, padding, align != 0);
                float gx = 0.f, gy = 0.f;
                if (grad_grid) {
                    float4 v[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        v[c] = __ldg(reinterpret_cast<const float4*>(feat + fo + ((size_t)k * HW + t.off[c]) * FEAT));
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const float d = g.x * v[c].x + g.y * v[c].y + g.z * v[c].z + g.w * v[c].w;
                        gx = fmaf(d, t.dwx[c], gx);
                        gy = fmaf(d, t.dwy[c], gy);
                    }
                }
                if (d_feat && live) {
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const float w = t.w[c];
                        if (w != 0.f)
                            ts_red4(d_feat + fo + ((size_t)k * HW + t.off[c]) * FEAT, w * g.x, w * g.y, w * g.z, w * g.w);
                    }
                }
                gg[k] += gx;
                gg[(k + 1) % 3] += gy;
            }
            if (grad_grid) {
#pragma unroll
                for (int k = 0; k < 3; ++k) {
#pragma unroll
                    for (int m = 1; m < 8; m <<= 1) gg[k] += __shfl_xor_sync(0xffffffffu, gg[k], m);
                }
                if (live && sub < 3) {   // one point is handled by exactly one 8-lane group: plain read-modify-write
                    float* o = grad_grid + ((size_t)nb * npts + p) * 3 + sub;
                    *o += sub == 0 ? gg[0] : (sub == 1 ? gg[1] : gg[2]);
                }
            }
        }
        __syncwarp();
    }
}

static int check_cl(const char* who, const void* a, const void* b, int N, int H, int W, long long npts, int padding) {
    ENARF_REQUIRE(a && b, ENARF_ERR_INVALID_ARG, "%s: NULL argument", who);
    ENARF_REQUIRE(N > 0 && H > 0 && W > 0 && npts > 0, ENARF_ERR_INVALID_ARG, "%s: bad sizes", who);
    ENARF_REQUIRE(padding >= 0 && padding <= 2, ENARF_ERR_INVALID_ARG, "%s: padding %d", who, padding);
    ENARF_REQUIRE(N <= 65535, ENARF_ERR_UNSUPPORTED, "%s: N too large", who);
    return ENARF_OK;
}

static dim3 cl_grid(long long npts, int N) {
    const long long want = (npts + 255) / 256;
    const long long cap = (long long)enarf_sm_count() * 8 * 4;
    return dim3((unsigned)(want < cap ? want : cap), N);
}

extern "C" int enarf_triplane_sample_cl_fwd(const float* feat_planes, const float* grid, int N, int H, int W,
                                            int64_t points, int padding, int align_corners, float* output,
                                            void* stream) {
    int rc = check_cl("enarf_triplane_sample_cl_fwd", feat_planes, grid, N, H, W, points, padding);
    if (rc) return rc;
    ENARF_REQUIRE(output, ENARF_ERR_INVALID_ARG, "enarf_triplane_sample_cl_fwd: NULL output");
    triplane_cl_fwd_kernel<<<cl_grid(points, N), 256, 0, (cudaStream_t)stream>>>(feat_planes, grid, H, W, points, padding,
                                                                                align_corners, output);
    ENARF_CHECK_LAUNCH("enarf_triplane_sample_cl_fwd");
    return ENARF_OK;
}

extern "C" int enarf_triplane_sample_cl_bwd(const float* grad_output, const float* feat_planes, const float* grid, int N,
                                            int H, int W, int64_t points, int padding, int align_corners,
                                            float* d_feat_planes, float* grad_grid, void* stream) {
    int rc = check_cl("enarf_triplane_sample_cl_bwd", feat_planes, grid, N, H, W, points, padding);
    if (rc) return rc;
    ENARF_REQUIRE(grad_output, ENARF_ERR_INVALID_ARG, "enarf_triplane_sample_cl_bwd: NULL grad_output");
    if (!d_feat_planes && !grad_grid) return ENARF_OK;
    triplane_cl_bwd_kernel<<<cl_grid(points, N), 256, 0, (cudaStream_t)stream>>>(grad_output, feat_planes, grid, H, W, points,
                                                                                padding, align_corners, d_feat_planes,
                                                                                grad_grid);
    ENARF_CHECK_LAUNCH("enarf_triplane_sample_cl_bwd");
    return ENARF_OK;
}

static int check_ts(const char* who, const void* a, const void* b, int N, int C, int H, int W, int Ho, int Wo,
                    int interp, int padding) {
    ENARF_REQUIRE(a && b, ENARF_ERR_INVALID_ARG, "%s: NULL argument", who);
    ENARF_REQUIRE(N > 0 && C > 0 && H > 0 && W > 0 && Ho > 0 && Wo > 0, ENARF_ERR_INVALID_ARG, "%s: bad sizes", who);
    ENARF_REQUIRE(interp == 0 || interp == 1, ENARF_ERR_INVALID_ARG, "%s: interpolation %d", who, interp);
    ENARF_REQUIRE(padding >= 0 && padding <= 2, ENARF_ERR_INVALID_ARG, "%s: padding %d", who, padding);
    ENARF_REQUIRE(N <= 65535 && (C + TS_CH - 1) / TS_CH <= 65535, ENARF_ERR_UNSUPPORTED, "%s: N or C too large", who);
    return ENARF_OK;
}

extern "C" int enarf_triplane_sample_fwd(const float* input, const float* grid, int N, int C, int H, int W, int Ho,
                                         int Wo, int interpolation, int padding, int align_corners, float* output,
                                         void* stream) {
    int rc = check_ts("enarf_triplane_sample_fwd", input, grid, N, C, H, W, Ho, Wo, interpolation, padding);
    if (rc) return rc;
    ENARF_REQUIRE(output, ENARF_ERR_INVALID_ARG, "enarf_triplane_sample_fwd: NULL output");
    const long long npts = (long long)Ho * Wo;
    dim3 g((unsigned)((npts + 255) / 256), (C + TS_CH - 1) / TS_CH, N);
    triplane_fwd_kernel<<<g, 256, 0, (cudaStream_t)stream>>>(input, grid, C, H, W, npts, interpolation, padding,
                                                            align_corners, output);
    ENARF_CHECK_LAUNCH("enarf_triplane_sample_fwd");
    return ENARF_OK;
}

extern "C" int enarf_triplane_sample_bwd(const float* grad_output, const float* input, const float* grid, int N,
                                         int C, int H, int W, int Ho, int Wo, int interpolation, int padding,
                                         int align_corners, float* grad_input, float* grad_grid, void* stream) {
    int rc = check_ts("enarf_triplane_sample_bwd", input, grid, N, C, H, W, Ho, Wo, interpolation, padding);
    if (rc) return rc;
    ENARF_REQUIRE(grad_output, ENARF_ERR_INVALID_ARG, "enarf_triplane_sample_bwd: NULL grad_output");
    if (!grad_input && !grad_grid) return ENARF_OK;
    const long long npts = (long long)Ho * Wo;
    dim3 g((unsigned)((npts + 255) / 256), (C + TS_CH - 1) / TS_CH, N);
    triplane_bwd_kernel<<<g, 256, 0, (cudaStream_t)stream>>>(grad_output, input, grid, C, H, W, npts, interpolation,
                                                            padding, align_corners, grad_input, grad_grid);
    ENARF_CHECK_LAUNCH("enarf_triplane_sample_bwd");
    return ENARF_OK;
}
