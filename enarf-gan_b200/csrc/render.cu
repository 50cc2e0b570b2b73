// Fused forward of the hot path: per-ray point generation (a4/a5) -> 24-part inverse transform and culling
// (a6) -> multi-part tri-plane gather (a7/a8) -> modulated decoder (a10) -> alpha compositing (a11).
// Reference being replaced: libraries/NeRF/rendering.py:119-335, models/narf.py:147-275,
// libraries/triplane/sampling.py:54-127, libraries/triplane/triplane_nerf.py:32-48.
//
// One CTA (256 threads, 2 CTAs/SM) owns whole rays (a tile of rays x samples <= TP_MAX points) so that
// per-point features, activations, densities and colours live only in shared memory / registers:
//   phase 0  per (ray, part): conservative slab clip of the ray segment against the part cube -> candidate
//            part mask + parameter interval per ray (exactness argument: common.cuh ray_part_interval);
//   phase 1  every point: position (contract arithmetic), exact slab tests for the candidate parts only ->
//            uint32 part mask; points that hit at least one part are compacted into a list (the others have
//            density 0 by narf.py:204 and never reach the decoder);
//   phase 2  per chunk of 256 active points: 8 lanes per point gather the channels-last feature texels with
//            128-bit loads (12 full 128-B lines per valid part), weighted by prod sigmoid(prob planes);
//            then one thread per point runs the 32->64->64->4 decoder: layer 1 as an outer product over the
//            inputs (64 register accumulators, transposed weights broadcast from smem), layers 2+3 as row dot
//            products folded straight into the 4 outputs, so no activation ever goes back to shared memory;
//   phase 3  one warp per ray: shuffle scan of sigma*delta -> transmittance -> weights -> colour/mask/disparity.
// The grid is persistent (2 x SM count) and strides over tiles in sample-major order so one sample's planes
// stay L2-resident and decoder weights are reloaded only when the sample changes.
#include <cuda_bf16.h>

#include "common.cuh"
#include "gather.cuh"

#define TP_MAX 1024       // points per tile
#define NT 256            // threads per CTA
#define CHUNK 256         // active points decoded per pass
#define FS 257            // feature column stride (conflict-free for the gather writes and the decoder reads)
#define MAX_RAYS 16       // rays per tile upper bound

enum { MODE_COARSE = 0, MODE_FINE = 1, MODE_QUERY = 2 };

struct RenderArgs {
    enarf_scene_t sc;
    const float* dmin;
    const float* dmax;
    const float* dir;
    const float* bins;
    const float* points;
    int n;              // rays per sample (points per sample in query mode)
    int S;              // samples per ray (1 in query mode)
    int rays_per_tile;  // rays (query: points) per tile
    int tiles_per_sample;
    int* tile_counter;
    float* weights_out;
    float* density_out;
    float* color;
    float* mask;
    float* disp;
    float* fine_weights;
    float* fine_depth;
    float* point_out;
    float* q_density;
    float* q_color;
    uint32_t* q_mask;
};

struct Smem {
    float w1t[FEAT * HID];   // [i][o]  layer 1, transposed
    float w2[HID * HID];     // [o][i]  layer 2, as stored
    float w3t[HID * 4];      // [o][c]  layer 3, transposed
    float b1[HID], b2[HID], b3[4];
    float part[ENARF_MAX_PARTS * PART_STRIDE];
    float feat[FEAT * FS];
    float pos[TP_MAX * 3];
    float out[TP_MAX * 4];   // r,g,b,sigma per point
    float depth[TP_MAX + MAX_RAYS];
    float ray[MAX_RAYS * 5];  // dmin,dmax,dx,dy,dz
    float cull[MAX_RAYS * ENARF_MAX_PARTS * 2];  // u0,u1 per (ray, part)
    uint32_t cand[MAX_RAYS];                     // candidate part mask per ray
    uint32_t pmask[TP_MAX];
    uint16_t list[TP_MAX];
    int n_active;
};

// Decoder for one active point whose 32 features sit in column `col` of s.feat.
__device__ __forceinline__ float4 decode_point(const Smem& s, int col) {
    float h[HID];
#pragma unroll
    for (int o = 0; o < HID; ++o) h[o] = s.b1[o];
#pragma unroll 2
    for (int i = 0; i < FEAT; ++i) {
        const float xi = s.feat[i * FS + col];
        const float4* w = reinterpret_cast<const float4*>(s.w1t + i * HID);
#pragma unroll
        for (int o4 = 0; o4 < HID / 4; ++o4) {
            const float4 v = w[o4];
            h[4 * o4 + 0] = fmaf(v.x, xi, h[4 * o4 + 0]);
            h[4 * o4 + 1] = fmaf(v.y, xi, h[4 * o4 + 1]);
            h[4 * o4 + 2] = fmaf(v.z, xi, h[4 * o4 + 2]);
            h[4 * o4 + 3] = fmaf(v.w, xi, h[4 * o4 + 3]);
        }
    }
#pragma unroll
    for (int o = 0; o < HID; ++o) h[o] = lrelu_s2(h[o]);
    float o3[4] = {s.b3[0], s.b3[1], s.b3[2], s.b3[3]};
#pragma unroll 2
    for (int o = 0; o < HID; ++o) {
        const float4* w = reinterpret_cast<const float4*>(s.w2 + o * HID);
        float a0 = s.b2[o], a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
        for (int i4 = 0; i4 < HID / 4; ++i4) {
            const float4 v = w[i4];
            a0 = fmaf(v.x, h[4 * i4 + 0], a0);
            a1 = fmaf(v.y, h[4 * i4 + 1], a1);
            a2 = fmaf(v.z, h[4 * i4 + 2], a2);
            a3 = fmaf(v.w, h[4 * i4 + 3], a3);
        }
        const float act = lrelu_s2((a0 + a1) + (a2 + a3));
        const float4 w3 = *reinterpret_cast<const float4*>(s.w3t + o * 4);
        o3[0] = fmaf(w3.x, act, o3[0]);
        o3[1] = fmaf(w3.y, act, o3[1]);
        o3[2] = fmaf(w3.z, act, o3[2]);
        o3[3] = fmaf(w3.w, act, o3[3]);
    }
    // rgb = tanh(.) (triplane_nerf.py:46); density = relu(.) * 10 (narf.py:274); every layer incl. the last is
    // conv + bias -> LeakyReLU(0.2) * sqrt2 (custom_stylegan2/net.py:317-318)
    return make_float4(tanhf(lrelu_s2(o3[0])), tanhf(lrelu_s2(o3[1])), tanhf(lrelu_s2(o3[2])),
                       fmaxf(lrelu_s2(o3[3]), 0.f) * 10.f);
}

__device__ __forceinline__ void load_decoder(Smem& s, const enarf_scene_t& sc, int b, int tid) {
    const float* w1 = sc.w1 + (size_t)b * HID * FEAT;
    const float* w2 = sc.w2 + (size_t)b * HID * HID;
    const float* w3 = sc.w3 + (size_t)b * 4 * HID;
    for (int e = tid; e < HID * FEAT; e += NT) {   // e = i*HID + o: conflict-free smem store, strided L2 read
        const int i = e / HID, o = e - i * HID;
        s.w1t[e] = w1[o * FEAT + i];
    }
    for (int e = tid; e < HID * HID / 4; e += NT)
        reinterpret_cast<float4*>(s.w2)[e] = reinterpret_cast<const float4*>(w2)[e];
    for (int e = tid; e < 4 * HID; e += NT) {
        const int o = e >> 2, c = e & 3;
        s.w3t[e] = w3[c * HID + o];
    }
    if (tid < HID) {
        s.b1[tid] = sc.b1[tid];
        s.b2[tid] = sc.b2[tid];
    }
    if (tid < 4) s.b3[tid] = sc.b3[tid];
}

template <int MODE, typename T>
__global__ void __launch_bounds__(NT, 2) render_kernel(const RenderArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem& s = *reinterpret_cast<Smem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const enarf_scene_t& sc = a.sc;
    const int S = a.S, n = a.n, P = sc.P;
    const int total_tiles = sc.B * a.tiles_per_sample;
    const float cstep = (MODE == MODE_COARSE) ? __fdiv_rn(1.0f, (float)S) : 0.f;
    int cur_b = -1;

    __shared__ int s_tile;
    for (;;) {
        __syncthreads();  // everyone is done with the previous tile (and with s_tile)
        if (tid == 0) s_tile = atomicAdd(a.tile_counter, 1);
        __syncthreads();
        const int tile = s_tile;
        if (tile >= total_tiles) break;
        const int b = tile / a.tiles_per_sample;
        const int ray0 = (tile - b * a.tiles_per_sample) * a.rays_per_tile;
        const int rays = min(a.rays_per_tile, n - ray0);
        const int tp = rays * S;
        __syncthreads();  // previous tile fully consumed before smem is reused
        if (b != cur_b) {
            load_part_params(s.part, sc, b, tid, NT);
            load_decoder(s, sc, b, tid);
            cur_b = b;
        }
        if (tid == 0) s.n_active = 0;
        if (MODE != MODE_QUERY && tid < rays) {
            const size_t o = (size_t)b * n + ray0 + tid;
            const float* d = a.dir + (size_t)b * 3 * n + ray0 + tid;
            s.ray[tid * 5 + 0] = a.dmin[o];
            s.ray[tid * 5 + 1] = a.dmax[o];
            s.ray[tid * 5 + 2] = d[0];
            s.ray[tid * 5 + 3] = d[n];
            s.ray[tid * 5 + 4] = d[2 * (size_t)n];
            s.cand[tid] = 0;
        }
        __syncthreads();

        // ---- phase 0: conservative ray/part clip ----------------------------------------------------
        if (MODE != MODE_QUERY) {
            for (int e = tid; e < rays * P; e += NT) {
                const int r = e / P, p = e - r * P;
                const float* ry = s.ray + r * 5;
                float u0, u1;
                if (ray_part_interval(s.part + p * PART_STRIDE, ry[0] * ry[2], ry[0] * ry[3], ry[0] * ry[4],
                                      ry[1] * ry[2], ry[1] * ry[3], ry[1] * ry[4], u0, u1)) {
                    s.cull[(r * ENARF_MAX_PARTS + p) * 2] = u0;
                    s.cull[(r * ENARF_MAX_PARTS + p) * 2 + 1] = u1;
                    atomicOr(&s.cand[r], 1u << p);
                }
            }
            __syncthreads();
        }

        // ---- phase 1: positions, exact part masks, compaction -----------------------------------------
        for (int t = tid; t < ((tp + 31) & ~31); t += NT) {
            uint32_t m = 0;
            if (t < tp) {
                float x, y, z;
                if (MODE == MODE_QUERY) {
                    const float* pt = a.points + (size_t)b * 3 * n + ray0 + t;
                    x = pt[0];
                    y = pt[n];
                    z = pt[2 * (size_t)n];
                    for (int p = 0; p < P; ++p)
                        if (part_valid(s.part + p * PART_STRIDE, x, y, z)) m |= 1u << p;
                } else {
                    const int r = t / S, j = t - r * S;
                    const float* ry = s.ray + r * 5;
                    const float dmin = ry[0], dmax = ry[1];
                    const float sx = cmul(dmin, ry[2]), sy = cmul(dmin, ry[3]), sz = cmul(dmin, ry[4]);
                    const float ex = cmul(dmax, ry[2]), ey = cmul(dmax, ry[3]), ez = cmul(dmax, ry[4]);
                    float u;
                    if (MODE == MODE_COARSE) {
                        // mid-point of consecutive stratified edges (rendering.py:125-131)
                        const float t0 = unit_bin(j, S, cstep), t1 = unit_bin(j + 1, S, cstep);
                        x = cmul(cadd(clerp(sx, ex, t1), clerp(sx, ex, t0)), 0.5f);
                        y = cmul(cadd(clerp(sy, ey, t1), clerp(sy, ey, t0)), 0.5f);
                        z = cmul(cadd(clerp(sz, ez, t1), clerp(sz, ez, t0)), 0.5f);
                        s.depth[r * (S + 1) + j] = clerp(dmin, dmax, t0);
                        if (j == S - 1) s.depth[r * (S + 1) + S] = clerp(dmin, dmax, t1);
                        u = 0.5f * (t0 + t1);
                    } else {
                        const size_t o = ((size_t)b * n + ray0 + r) * S + j;
                        u = a.bins[o];
                        x = clerp(sx, ex, u);
                        y = clerp(sy, ey, u);
                        z = clerp(sz, ez, u);
                        const float dep = clerp(dmin, dmax, u);
                        s.depth[r * (S + 1) + j] = dep;
                        a.fine_depth[o] = dep;
                    }
                    uint32_t cm = s.cand[r];
                    const float* cu = s.cull + r * ENARF_MAX_PARTS * 2;
                    while (cm) {
                        const int p = __ffs(cm) - 1;
                        cm &= cm - 1;
                        if (u >= cu[2 * p] && u <= cu[2 * p + 1] && part_valid(s.part + p * PART_STRIDE, x, y, z))
                            m |= 1u << p;
                    }
                }
                s.pos[t * 3 + 0] = x;
                s.pos[t * 3 + 1] = y;
                s.pos[t * 3 + 2] = z;
                s.pmask[t] = m;
                *reinterpret_cast<float4*>(s.out + t * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, m != 0);
            if (bal) {
                int base = 0;
                if (lane == 0) base = atomicAdd(&s.n_active, __popc(bal));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (m) s.list[base + __popc(bal & ((1u << lane) - 1))] = (uint16_t)t;
            }
        }
        __syncthreads();
        const int n_active = s.n_active;

        // ---- phase 2: gather + decode, CHUNK active points at a time ---------------------------------
        for (int c0 = 0; c0 < n_active; c0 += CHUNK) {
            const int cnt = min(CHUNK, n_active - c0);
            const int lane8 = tid & 7;
            for (int q = tid >> 3; q < cnt; q += NT / 8) {
                const int t = s.list[c0 + q];
                const float4 f = gather_point<T>(sc, s.part, b, s.pmask[t], s.pos[t * 3], s.pos[t * 3 + 1],
                                                 s.pos[t * 3 + 2], lane8);
                float* dst = s.feat + (lane8 * 4) * FS + q;
                dst[0] = f.x;
                dst[FS] = f.y;
                dst[2 * FS] = f.z;
                dst[3 * FS] = f.w;
            }
            __syncthreads();
            if (tid < cnt) *reinterpret_cast<float4*>(s.out + s.list[c0 + tid] * 4) = decode_point(s, tid);
            __syncthreads();
        }

        // ---- phase 3: outputs ---------------------------------------------------------------------------
        if (MODE == MODE_QUERY) {
            for (int t = tid; t < tp; t += NT) {
                const float4 o = *reinterpret_cast<const float4*>(s.out + t * 4);
                const size_t gi = (size_t)b * n + ray0 + t;
                if (a.q_density) a.q_density[gi] = o.w;
                if (a.q_color) {
                    float* c = a.q_color + (size_t)b * 3 * n + ray0 + t;
                    c[0] = o.x;
                    c[n] = o.y;
                    c[2 * (size_t)n] = o.z;
                }
                if (a.q_mask) a.q_mask[gi] = s.pmask[t];
            }
            continue;
        }
        if (MODE == MODE_FINE && a.point_out) {
            float4* dst = reinterpret_cast<float4*>(a.point_out) + ((size_t)b * n + ray0) * S;
            for (int t = tid; t < tp; t += NT) dst[t] = *reinterpret_cast<const float4*>(s.out + t * 4);
        }
        const int M = (MODE == MODE_COARSE) ? S : S - 1;  // intervals that carry weight
        for (int r = warp; r < rays; r += NT / 32) {
            const float* dep = s.depth + r * (S + 1);
            float* po = s.out + (size_t)r * S * 4;
            const size_t gray = (size_t)b * n + ray0 + r;
            float carry = 0.f, cr = 0.f, cg = 0.f, cb = 0.f, cm = 0.f, cd = 0.f;
            for (int base = 0; base < M; base += 32) {
                const int j = base + lane;
                const bool ok = j < M;
                const float sigma = ok ? po[j * 4 + 3] : 0.f;
                const float dd = ok ? sigma * (dep[j + 1] - dep[j]) : 0.f;
                float incl = dd;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const float v = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += v;
                }
                const float excl = carry + (incl - dd);
                const float w = expf(-excl) * (1.f - expf(-dd));
                carry += __shfl_sync(0xffffffffu, incl, 31);
                if (ok) {
                    if (MODE == MODE_COARSE) {
                        if (a.density_out) a.density_out[gray * S + j] = sigma;
                        po[j * 4 + 0] = w;  // colour slot is free in the coarse pass
                    } else {
                        a.fine_weights[gray * M + j] = w;
                        cr = fmaf(w, po[j * 4 + 0], cr);
                        cg = fmaf(w, po[j * 4 + 1], cg);
                        cb = fmaf(w, po[j * 4 + 2], cb);
                        cm += w;
                        cd += w / dep[j];
                    }
                }
            }
            if (MODE == MODE_COARSE) {
                __syncwarp();
                // (max(w[j-1],w[j]) + max(w[j],w[j+1]))/2 + 0.01, zero padded (rendering.py:187-190)
                for (int j = lane; j < S; j += 32) {
                    const float wm = j > 0 ? po[(j - 1) * 4] : 0.f, w0 = po[j * 4], wp = j + 1 < S ? po[(j + 1) * 4] : 0.f;
                    a.weights_out[gray * S + j] = cadd(cmul(cadd(fmaxf(wm, w0), fmaxf(w0, wp)), 0.5f), 0.01f);
                }
            } else {
#pragma unroll
                for (int o = 16; o; o >>= 1) {
                    cr += __shfl_xor_sync(0xffffffffu, cr, o);
                    cg += __shfl_xor_sync(0xffffffffu, cg, o);
                    cb += __shfl_xor_sync(0xffffffffu, cb, o);
                    cm += __shfl_xor_sync(0xffffffffu, cm, o);
                    cd += __shfl_xor_sync(0xffffffffu, cd, o);
                }
                if (lane == 0) {
                    float* c = a.color + (size_t)b * 3 * n + ray0 + r;
                    c[0] = cr;
                    c[n] = cg;
                    c[2 * (size_t)n] = cb;
                    a.mask[gray] = cm;
                    a.disp[gray] = cd;
                }
            }
        }
    }
}

// ---- host side -----------------------------------------------------------------------------------
template <int MODE>
static int launch_render(RenderArgs& a, const char* who, cudaStream_t stream) {
    const int total = a.sc.B * a.tiles_per_sample;
    const int grid = min(total, 2 * enarf_sm_count());
    const size_t smem = sizeof(Smem);
    a.tile_counter = enarf_tile_counter(stream);
    if (!a.tile_counter) return ENARF_ERR_CUDA;
    cudaError_t e;
    if (a.sc.feat_dtype == ENARF_DTYPE_F32) {
        auto k = render_kernel<MODE, float>;
        e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) k<<<grid, NT, smem, stream>>>(a);
    } else {
        auto k = render_kernel<MODE, __nv_bfloat16>;
        e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) k<<<grid, NT, smem, stream>>>(a);
    }
    if (e != cudaSuccess) {
        enarf_set_error("%s: cudaFuncSetAttribute: %s", who, cudaGetErrorString(e));
        return ENARF_ERR_CUDA;
    }
    ENARF_CHECK_LAUNCH(who);
    return ENARF_OK;
}

static int tile_rays(int S) { return max(1, min(MAX_RAYS, TP_MAX / S)); }

extern "C" int enarf_render_coarse(const enarf_scene_t* scene, const float* depth_min, const float* depth_max,
                                   const float* ray_dir, int n, int Nc, float* weights_out, float* density_out,
                                   void* stream) {
    int rc = enarf_check_scene(scene, "enarf_render_coarse");
    if (rc) return rc;
    ENARF_REQUIRE(depth_min && depth_max && ray_dir && weights_out, ENARF_ERR_INVALID_ARG,
                  "enarf_render_coarse: NULL argument");
    ENARF_REQUIRE(n > 0 && Nc >= 2, ENARF_ERR_INVALID_ARG, "enarf_render_coarse: bad sizes n=%d Nc=%d", n, Nc);
    ENARF_REQUIRE(Nc <= TP_MAX, ENARF_ERR_UNSUPPORTED, "enarf_render_coarse: Nc=%d exceeds %d", Nc, TP_MAX);
    RenderArgs a = {};
    a.sc = *scene;
    a.dmin = depth_min;
    a.dmax = depth_max;
    a.dir = ray_dir;
    a.n = n;
    a.S = Nc;
    a.rays_per_tile = tile_rays(Nc);
    a.tiles_per_sample = (n + a.rays_per_tile - 1) / a.rays_per_tile;
    a.weights_out = weights_out;
    a.density_out = density_out;
    return launch_render<MODE_COARSE>(a, "enarf_render_coarse", (cudaStream_t)stream);
}

extern "C" int enarf_render_fine_fwd(const enarf_scene_t* scene, const float* depth_min, const float* depth_max,
                                     const float* ray_dir, const float* bins, int n, int Nf, float* color,
                                     float* mask, float* disparity, float* fine_weights, float* fine_depth,
                                     float* point_out, void* stream) {
    int rc = enarf_check_scene(scene, "enarf_render_fine_fwd");
    if (rc) return rc;
    ENARF_REQUIRE(depth_min && depth_max && ray_dir && bins && color && mask && disparity && fine_weights && fine_depth,
                  ENARF_ERR_INVALID_ARG, "enarf_render_fine_fwd: NULL argument");
    ENARF_REQUIRE(n > 0 && Nf >= 2, ENARF_ERR_INVALID_ARG, "enarf_render_fine_fwd: bad sizes n=%d Nf=%d", n, Nf);
    ENARF_REQUIRE(Nf <= TP_MAX, ENARF_ERR_UNSUPPORTED, "enarf_render_fine_fwd: Nf=%d exceeds %d", Nf, TP_MAX);
    ENARF_REQUIRE(!point_out || ((uintptr_t)point_out & 15) == 0, ENARF_ERR_INVALID_ARG,
                  "enarf_render_fine_fwd: point_out must be 16-byte aligned");
    RenderArgs a = {};
    a.sc = *scene;
    a.dmin = depth_min;
    a.dmax = depth_max;
    a.dir = ray_dir;
    a.bins = bins;
    a.n = n;
    a.S = Nf;
    a.rays_per_tile = tile_rays(Nf);
    a.tiles_per_sample = (n + a.rays_per_tile - 1) / a.rays_per_tile;
    a.color = color;
    a.mask = mask;
    a.disp = disparity;
    a.fine_weights = fine_weights;
    a.fine_depth = fine_depth;
    a.point_out = point_out;
    return launch_render<MODE_FINE>(a, "enarf_render_fine_fwd", (cudaStream_t)stream);
}

extern "C" int enarf_density_query(const enarf_scene_t* scene, const float* points, int N, float* density,
                                   float* color, uint32_t* part_mask, void* stream) {
    int rc = enarf_check_scene(scene, "enarf_density_query");
    if (rc) return rc;
    ENARF_REQUIRE(points && N > 0, ENARF_ERR_INVALID_ARG, "enarf_density_query: bad arguments");
    RenderArgs a = {};
    a.sc = *scene;
    a.points = points;
    a.n = N;
    a.S = 1;
    a.rays_per_tile = TP_MAX;
    a.tiles_per_sample = (N + TP_MAX - 1) / TP_MAX;
    a.q_density = density;
    a.q_color = color;
    a.q_mask = part_mask;
    return launch_render<MODE_QUERY>(a, "enarf_density_query", (cudaStream_t)stream);
}
