// Fused forward of the hot path: per-ray point generation (a4/a5) -> 24-part inverse transform and culling
// (a6) -> multi-part tri-plane gather (a7/a8) -> modulated decoder (a10) -> alpha compositing (a11).
// Reference being replaced: libraries/NeRF/rendering.py:119-335, models/narf.py:147-275,
// libraries/triplane/sampling.py:54-127, libraries/triplane/triplane_nerf.py:32-48.
//
// Two 256-thread CTAs per SM, each owning whole rays (tile.cuh: phase 0/1, compaction, pair prefix).  Inside a tile
// the 8 warps are autonomous: each pulls chunks of 32 active points from a shared-memory counter and, with no
// CTA-wide barrier,
//   sets up the (point, part) pair records of its points (one lane per pair), gathers the features with 8 lanes
//   per point into its own 4 KB slab, and runs the 32->64->64->4 decoder for its 32 points on the tensor pipe
//   (mlp_mma.cuh: error-compensated 3xTF32 mma.sync, activations chained through registers),
// so that the latency-bound gathers of some warps overlap the tensor work of others.  Then one warp per ray
// composites: shuffle scan of sigma*delta -> transmittance -> weights -> colour/mask/disparity.
// Per-point features, activations, densities and colours live only in shared memory / registers.  The grid is
// persistent and pulls tiles in sample-major order from an atomic counter so that one sample's planes stay
// L2-resident and the decoder weights are reloaded only when the sample changes.
#define NT 256
#define MAX_RAYS 16
#define RECW 52
#ifdef ENARF_PHASE_TIMING   // development build: per-warp clock64() sums (lane 0 of every warp), scripts/fwd_phases.py
struct FwdTimer {
    long long t;
    unsigned long long acc[8];
};
#define FWD_TIMER_PARAM , FwdTimer& ftm
#define FWD_TIMER_ARG , ftm
#define FWD_PHASE(i) do { if ((threadIdx.x & 31) == 0) { const long long now_ = clock64(); ftm.acc[i] += (unsigned long long)(now_ - ftm.t); ftm.t = now_; } } while (0)
static __device__ unsigned long long g_fwd_phase[8];
#else
#define FWD_TIMER_PARAM
#define FWD_TIMER_ARG
#define FWD_PHASE(i)
#endif
#include "tile.cuh"
#include "mlp_mma.cuh"

enum { MODE_COARSE = TILE_COARSE, MODE_FINE = TILE_FINE, MODE_QUERY = TILE_QUERY };

struct RenderArgs {
    enarf_scene_t sc;
    TileGeom g;
    int rays_per_tile;  // rays (query: points) per tile
    int tiles_per_sample;
    int* tile_counter;
    float* weights_out;
    float* density_out;
    float* color;
    float* mask;
    float* disp;
    float* fine_weights;
    float* point_out;
    float* feat_out;    // (B,n,S,32) checkpoint of the gathered features for the backward, or NULL
    enarf_pair_ckpt_t ck;   // per-pair checkpoint (rec == NULL: none)
    float* q_density;
    float* q_color;
    uint32_t* q_mask;
    // early ray termination (fine pass): threshold on the prefix transmittance (0 = off), the coarse pass's count of
    // opaque rays (rounds are only worth their barriers when there are some; NULL = always), statistics out
    float ert_eps;
    const int* ert_hint;
    int ert_min_count;
    int* ert_stats;         // [0] += active points evaluated, [1] += active points skipped
    int* opaque_rays;       // coarse pass: += rays whose coarse transmittance ends below OPAQUE_T
    float bg_const;         // fine pass: colour out = rgb + (1 - mask) * bg_const (models/generator.py:105-107; 0 = foreground)
};
#define OPAQUE_T 1e-2f

static_assert((32 * XSW + RECW * REC) >= 32 * 64, "the slab holds the 32 x 64 layer-1 activations");
struct __align__(16) WarpSlab {
    float X[32 * XSW];        // features of the warp's 32 points, [point][channel]
    float rec[RECW * REC];    // pair records of the current sub-batch
};

struct __align__(16) FwdSmem {
    DecoderMma dec;
    float part[ENARF_MAX_PARTS * PART_STRIDE];
    TileSmem t;               // t.pm[i] = (x, y, z, part mask) is overwritten by (r, g, b, sigma) once point i is
                              // decoded; a point outside every part keeps mask bits 0 == sigma 0.0f
    union {
        WarpSlab slab[NWARP];
        float cull[MAX_RAYS * ENARF_MAX_PARTS * 2];   // phase 0/1 only (tile_points sees it as RecSmem::cull)
    } u;
};
static_assert(XSW == XS, "gather slab stride must match the decoder's A-fragment loads");

template <int MODE, typename T>
__global__ void __launch_bounds__(NT, 2) render_kernel(const RenderArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FwdSmem& s = *reinterpret_cast<FwdSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const enarf_scene_t& sc = a.sc;
    const int S = a.g.S, n = a.g.n;
    const int total_tiles = sc.B * a.tiles_per_sample;
    int cur_b = -1;
#ifdef ENARF_PHASE_TIMING
    FwdTimer ftm = {};
    ftm.t = clock64();
#endif

    for (;;) {
        __syncthreads();  // everyone is done with the previous tile (and with s.t.tile)
        FWD_PHASE(5);   // outputs / compositing of the previous tile + wait for the others
        if (tid == 0) s.t.tile = atomicAdd(a.tile_counter, 1);
        __syncthreads();
        const int tile = s.t.tile;
        if (tile >= total_tiles) break;
        const int b = tile / a.tiles_per_sample;
        const int ray0 = (tile - b * a.tiles_per_sample) * a.rays_per_tile;
        const int rays = min(a.rays_per_tile, n - ray0);
        const int tp = rays * S;
        if (b != cur_b) {
            load_part_params(s.part, sc, b, tid, NT);
            load_decoder_mma(s.dec, sc, b, tid, NT);
            cur_b = b;
        }
        const bool ert = MODE != MODE_QUERY && a.ert_eps > 0.f;
        // sample groups are evaluated in separate rounds only when the coarse pass saw opaque rays
        const int groups = ert ? (S + ERT_GROUP - 1) / ERT_GROUP : 1;
        const int rounds = (ert && (!a.ert_hint || *a.ert_hint >= a.ert_min_count)) ? groups : 1;
        tile_masks<MODE>(s.t, *reinterpret_cast<RecSmem*>(&s.u), s.part, a.g, sc.P, b, ray0, rays, tid);
        if (tid == 0) {
            s.t.alive = 0xffffffffu;
            s.t.pad_[0] = 0;   // chunk counter of this round
        }
        if (ert && tid < MAX_RAYS) s.t.carry[tid] = 0.f;
        if (MODE == MODE_QUERY && a.q_mask)
            for (int i = tid; i < tp; i += NT) a.q_mask[(size_t)b * n + ray0 + i] = __float_as_uint(s.t.pm[i].w);
        if (MODE == MODE_FINE && a.ck.rec && a.ck.count)   // valid parts per sample: the backward's shortcut to this tile's list
            for (int i = tid; i < tp; i += NT)
                a.ck.count[((size_t)b * n + ray0) * S + i] = (uint8_t)__popc(__float_as_uint(s.t.pm[i].w));
        float4* out = s.t.pm;

        FWD_PHASE(0);   // tile prologue: rays, phase 0/1, compaction, prefix
        // ---- chunks of 32 active points, one warp each, no CTA barrier; with early termination on an opaque scene the
        // sample groups go front to back in rounds, each followed by a partial composite that retires finished rays
        int n_done = 0;
        bool empty_tile = false;
        for (int rd = 0; rd < rounds; ++rd) {
            // (round 0: every ray is alive -- s.t.alive was only just written by thread 0)
            tile_compact(s.t, rays, S, ert ? ERT_GROUP : S, rounds == 1 ? 0 : rd, rounds == 1 ? groups : rd + 1,
                         rd == 0 ? 0xffffffffu : s.t.alive, nullptr, tid);
            const int n_active = s.t.n_active;
            if (rd == 0 && MODE == MODE_FINE && a.ck.rec) {   // hand this tile its slice of the pair checkpoint
                if (rounds > 1) {   // later rounds append to the list: size the slice for every active point of the tile
                    int mine = 0;
                    for (int i = tid; i < tp; i += NT) {
                        const float w = s.t.pm[i].w;
                        if (w != ERT_SENTINEL) mine += __popc(__float_as_uint(w));
                    }
#pragma unroll
                    for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
                    if (lane == 0) s.t.wsum[warp] = mine;
                    __syncthreads();
                }
                if (tid == 0) {
                    int np_tile = 0;
                    if (rounds > 1) for (int w = 0; w < NWARP; ++w) np_tile += s.t.wsum[w];
                    else np_tile = s.t.pref[n_active];
                    // once the checkpoint is full the cursor is left alone, so it cannot wrap however many tiles follow
                    int base = -1;
                    if (np_tile == 0) {
                        base = -2;      // no pairs at all: the backward skips the tile outright
                    } else if (*reinterpret_cast<volatile int*>(a.ck.cursor) <= a.ck.capacity) {
                        base = atomicAdd(a.ck.cursor, np_tile);
                        if ((long long)base + np_tile > a.ck.capacity) base = -1;
                    }
                    ENARF_DBG_ASSERT(base < 0 || (long long)base + np_tile <= a.ck.capacity);
                    a.ck.tile_base[tile] = base;
                    s.t.pad_[1] = base;
                }
                __syncthreads();
            }
            if (MODE != MODE_QUERY && rounds == 1 && n_active == 0) {
                // Nothing of the tile lies inside any part (45 % of the bench's tiles): write what the compositing below
                // would produce from all-zero densities -- the same values bit for bit -- and take the next tile.
                const int M0 = (MODE == MODE_COARSE) ? S : S - 1;
                const size_t gray0 = (size_t)b * n + ray0;
                if (MODE == MODE_COARSE) {
                    const float w0 = cadd(cmul(cadd(0.f, 0.f), 0.5f), 0.01f);
                    for (int i = tid; i < rays * S; i += NT) {
                        a.weights_out[gray0 * S + i] = w0;
                        if (a.density_out) a.density_out[gray0 * S + i] = 0.f;
                    }
                } else {
                    for (int i = tid; i < rays * M0; i += NT) a.fine_weights[gray0 * M0 + i] = 0.f;
                    if (a.point_out) {
                        float4* dst = reinterpret_cast<float4*>(a.point_out) + gray0 * S;
                        for (int i = tid; i < tp; i += NT) dst[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                    if (tid < rays) {
                        float* c = a.color + (size_t)b * 3 * n + ray0 + tid;
                        c[0] = a.bg_const;
                        c[n] = a.bg_const;
                        c[2 * (size_t)n] = a.bg_const;
                        a.mask[gray0 + tid] = 0.f;
                        a.disp[gray0 + tid] = 0.f;
                    }
                }
                empty_tile = true;      // (n_active is CTA-uniform; the loop top synchronises)
                break;
            }
            WarpSlab& my = s.u.slab[warp];
            const int g = lane >> 2, t4 = lane & 3;
            const int csz = 32;   // (an eighth of the round's active points per chunk, so that no warp idles at the tile-end
                                  // barrier, measured no faster: profiles/micro_opt_r2.txt)
            for (;;) {
                int c = 0;
                if (lane == 0) c = atomicAdd(&s.t.pad_[0], 1);
                c = __shfl_sync(0xffffffffu, c, 0);
                const int a0 = n_done + csz * c;
                if (a0 >= n_active) break;
                const int np = min(csz, n_active - a0);
                ENARF_DBG_ASSERT(a0 >= 0 && n_active <= TP_MAX);
                const int tb = (MODE == MODE_FINE && a.ck.rec) ? s.t.pad_[1] : -1;
                gather32_warp<T>(s.t, my.rec, my.X, s.part, sc, b, a0, np, lane,
                                 tb >= 0 ? a.ck.rec + (size_t)tb * REC : nullptr, tb >= 0 ? a.ck.feat + (size_t)tb * FEAT : nullptr
                                 FWD_TIMER_ARG);
                if (MODE == MODE_FINE && a.feat_out) {   // 8 lanes per point: one 128-B row each
                    float4* fo = reinterpret_cast<float4*>(a.feat_out) + ((size_t)b * n + ray0) * S * (FEAT / 4);
                    for (int q = lane >> 3; q < np; q += 4)
                        __stcs(fo + (size_t)s.t.list[a0 + q] * (FEAT / 4) + (lane & 7),
                               *reinterpret_cast<const float4*>(my.X + q * XSW + xs_col(q, 4 * (lane & 7))));
                }
                float o3[2][4];
                // bf16 planes (parity bar 1e-2): the coarse pass only shapes the sampling distribution and carries no
                // gradient, so its decoder runs single-pass TF32; the fine pass keeps 3xTF32 (with single-pass TF32
                // there the DSO step's plane gradient sat at 1.4e-2 of its max-norm: activation kinks)
                decode32_mma<(sizeof(T) == 2 && MODE == MODE_COARSE)>(my.X, my.X, s.dec, lane, o3);   // layer-1 activations overwrite the slab (X + records)
                __syncwarp();   // every lane is done reading X before the next chunk's gather overwrites it
                // rgb = tanh(.) (triplane_nerf.py:46); density = relu(.) * 10 (narf.py:274); every layer incl. the
                // last is conv + bias -> LeakyReLU(0.2) * sqrt2 (custom_stylegan2/net.py:317-318).
                // Lane (g, t4) holds channels 2*t4, 2*t4+1 of points g and g+8 of each 16-row tile; t4 >= 2 is padding.
                if (t4 < 2) {
#pragma unroll
                    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const int row = 16 * mt + 8 * h + g;
                            if (row < np) {
                                const float v0 = lrelu_folded(o3[mt][2 * h]), v1 = lrelu_folded(o3[mt][2 * h + 1]);
                                const float2 r = t4 == 0 ? make_float2(tanhf(v0), tanhf(v1))
                                                         : make_float2(tanhf(v0), fmaxf(v1, 0.f) * 10.f);
                                reinterpret_cast<float2*>(&out[s.t.list[a0 + row]])[t4] = r;
                            }
                        }
                }
                FWD_PHASE(3);   // checkpoint row store + decoder + head
            }
            __syncthreads();
            if (rd + 1 < rounds) {
                // optical depth of group rd's intervals, added to the ray's running total (exact prefix, sample order)
                for (int r = warp; r < rays; r += NWARP) {
                    const int j = rd * ERT_GROUP + lane;
                    float dd = 0.f;
                    if (lane < ERT_GROUP && j < S - 1) {
                        const float* dep = s.t.depth + r * (S + 1);
                        dd = fmaxf(out[r * S + j].w, 0.f) * (dep[j + 1] - dep[j]);
                    }
#pragma unroll
                    for (int o = 8; o; o >>= 1) dd += __shfl_xor_sync(0xffffffffu, dd, o);
                    if (lane == 0) {
                        const float od = s.t.carry[r] + dd;
                        s.t.carry[r] = od;
                        if (expf(-od) < a.ert_eps) atomicAnd(&s.t.alive, ~(1u << r));
                    }
                }
                if (tid == 0) s.t.pad_[0] = 0;
                __syncthreads();
            }
            n_done = n_active;
        }
        if (empty_tile) continue;
        if (MODE != MODE_QUERY && a.ert_stats && tid == 0) {
            atomicAdd(a.ert_stats, n_done);
            if (s.t.skipped) atomicAdd(a.ert_stats + 1, s.t.skipped);
        }
        FWD_PHASE(4);   // failed chunk claims + idle at the end of the tile's chunk phase(s)

        // ---- outputs ---------------------------------------------------------------------------------
        if (MODE == MODE_QUERY) {
            for (int i = tid; i < tp; i += NT) {
                const bool on = (s.t.act[i >> 5] >> (i & 31)) & 1;
                const float4 o = on ? out[i] : make_float4(0.f, 0.f, 0.f, 0.f);
                const size_t gi = (size_t)b * n + ray0 + i;
                if (a.q_density) a.q_density[gi] = o.w;
                if (a.q_color) {
                    float* c = a.q_color + (size_t)b * 3 * n + ray0 + i;
                    c[0] = o.x;
                    c[n] = o.y;
                    c[2 * (size_t)n] = o.z;
                }
            }
            continue;
        }
        if (MODE == MODE_FINE && a.point_out) {
            float4* dst = reinterpret_cast<float4*>(a.point_out) + ((size_t)b * n + ray0) * S;
            for (int i = tid; i < tp; i += NT)
                dst[i] = ((s.t.act[i >> 5] >> (i & 31)) & 1) ? out[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        const int M = (MODE == MODE_COARSE) ? S : S - 1;  // intervals that carry weight
        float* outf = reinterpret_cast<float*>(out);   // rgb of points outside every part is never weighted (sigma 0)
        for (int r = warp; r < rays; r += NWARP) {
            const float* dep = s.t.depth + r * (S + 1);
            float* po = outf + (size_t)r * S * 4;
            const size_t gray = (size_t)b * n + ray0 + r;
            float carry = 0.f, cr = 0.f, cg = 0.f, cb = 0.f, cm = 0.f, cd = 0.f;
            for (int base = 0; base < M; base += 32) {
                const int j = base + lane;
                const bool ok = j < M;
                const float sigma = ok ? fmaxf(po[j * 4 + 3], 0.f) : 0.f;   // (a dropped sample carries ERT_SENTINEL)
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
                if (a.opaque_rays && lane == 0 && expf(-carry) < OPAQUE_T) atomicAdd(a.opaque_rays, 1);
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
                    const float bgc = (1.f - cm) * a.bg_const;   // constant background composited in the epilogue
                    c[0] = cr + bgc;
                    c[n] = cg + bgc;
                    c[2 * (size_t)n] = cb + bgc;
                    a.mask[gray] = cm;
                    a.disp[gray] = cd;
                }
            }
        }
    }
#ifdef ENARF_PHASE_TIMING
    if ((threadIdx.x & 31) == 0)
        for (int i_ = 0; i_ < 8; ++i_) atomicAdd(&g_fwd_phase[i_], ftm.acc[i_]);
#endif
}

#ifdef ENARF_PHASE_TIMING
extern "C" int enarf_debug_fwd_phases(unsigned long long* host_out, int reset) {
    cudaMemcpyFromSymbol(host_out, g_fwd_phase, sizeof(g_fwd_phase));
    if (reset) {
        unsigned long long z[8] = {};
        cudaMemcpyToSymbol(g_fwd_phase, z, sizeof(z));
    }
    return 0;
}
#endif

// ---- host side -----------------------------------------------------------------------------------
template <int MODE>
static int launch_render(RenderArgs& a, const char* who, cudaStream_t stream) {
    const int total = a.sc.B * a.tiles_per_sample;
    const int grid = min(total, 2 * enarf_sm_count());
    const size_t smem = sizeof(FwdSmem);
    static_assert(sizeof(FwdSmem) <= 113 * 1024, "two CTAs per SM");
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

// bwd.cu uses the same tile geometry (NT, MAX_RAYS): tile t of the backward is tile t of the forward
extern "C" int enarf_fine_tiles_per_sample(int n, int Nf) {
    if (n <= 0 || Nf < 2 || Nf > TP_MAX) return 0;
    const int r = tile_rays(Nf);
    return (n + r - 1) / r;
}

extern "C" int enarf_render_coarse(const enarf_scene_t* scene, const float* depth_min, const float* depth_max,
                                   const float* ray_dir, int n, int Nc, float* weights_out, float* density_out,
                                   int* opaque_rays, float ert_eps, const int* ert_hint, int* ert_stats, void* stream) {
    int rc = enarf_check_scene(scene, "enarf_render_coarse");
    if (rc) return rc;
    ENARF_REQUIRE(depth_min && depth_max && ray_dir && weights_out, ENARF_ERR_INVALID_ARG,
                  "enarf_render_coarse: NULL argument");
    ENARF_REQUIRE(n > 0 && Nc >= 2, ENARF_ERR_INVALID_ARG, "enarf_render_coarse: bad sizes n=%d Nc=%d", n, Nc);
    ENARF_REQUIRE(Nc <= TP_MAX, ENARF_ERR_UNSUPPORTED, "enarf_render_coarse: Nc=%d exceeds %d", Nc, TP_MAX);
    RenderArgs a = {};
    a.sc = *scene;
    a.g.dmin = depth_min;
    a.g.dmax = depth_max;
    a.g.dir = ray_dir;
    a.g.n = n;
    a.g.S = Nc;
    a.rays_per_tile = tile_rays(Nc);
    a.tiles_per_sample = (n + a.rays_per_tile - 1) / a.rays_per_tile;
    a.weights_out = weights_out;
    a.density_out = density_out;
    a.opaque_rays = opaque_rays;
    ENARF_REQUIRE(ert_eps >= 0.f && ert_eps < 1.f, ENARF_ERR_INVALID_ARG, "enarf_render_coarse: ert_eps must be in [0, 1)");
    a.ert_eps = ert_eps;
    a.ert_hint = ert_hint;
    a.ert_min_count = max(1, (int)(((long long)scene->B * n) / 64));
    a.ert_stats = ert_stats;
    return launch_render<MODE_COARSE>(a, "enarf_render_coarse", (cudaStream_t)stream);
}

extern "C" int enarf_render_fine_fwd(const enarf_scene_t* scene, const float* depth_min, const float* depth_max,
                                     const float* ray_dir, const float* bins, int n, int Nf, float* color,
                                     float* mask, float* disparity, float* fine_weights, float* fine_depth,
                                     float* point_out, float* feat_out, const enarf_pair_ckpt_t* pair_ckpt, float ert_eps,
                                     const int* ert_hint, int* ert_stats, float bg_const, void* stream) {
    int rc = enarf_check_scene(scene, "enarf_render_fine_fwd");
    if (rc) return rc;
    ENARF_REQUIRE(depth_min && depth_max && ray_dir && bins && color && mask && disparity && fine_weights && fine_depth,
                  ENARF_ERR_INVALID_ARG, "enarf_render_fine_fwd: NULL argument");
    ENARF_REQUIRE(n > 0 && Nf >= 2, ENARF_ERR_INVALID_ARG, "enarf_render_fine_fwd: bad sizes n=%d Nf=%d", n, Nf);
    ENARF_REQUIRE(Nf <= TP_MAX, ENARF_ERR_UNSUPPORTED, "enarf_render_fine_fwd: Nf=%d exceeds %d", Nf, TP_MAX);
    ENARF_REQUIRE((!point_out || ((uintptr_t)point_out & 15) == 0) && (!feat_out || ((uintptr_t)feat_out & 15) == 0),
                  ENARF_ERR_INVALID_ARG, "enarf_render_fine_fwd: point_out / feat_out must be 16-byte aligned");
    RenderArgs a = {};
    a.sc = *scene;
    a.g.dmin = depth_min;
    a.g.dmax = depth_max;
    a.g.dir = ray_dir;
    a.g.bins = bins;
    a.g.fine_depth = fine_depth;
    a.g.n = n;
    a.g.S = Nf;
    a.rays_per_tile = tile_rays(Nf);
    a.tiles_per_sample = (n + a.rays_per_tile - 1) / a.rays_per_tile;
    a.color = color;
    a.mask = mask;
    a.disp = disparity;
    a.fine_weights = fine_weights;
    a.point_out = point_out;
    a.feat_out = feat_out;
    ENARF_REQUIRE(ert_eps >= 0.f && ert_eps < 1.f, ENARF_ERR_INVALID_ARG, "enarf_render_fine_fwd: ert_eps must be in [0, 1)");
    ENARF_REQUIRE(!(ert_eps > 0.f) || point_out, ENARF_ERR_INVALID_ARG,
                  "enarf_render_fine_fwd: early termination records the dropped samples in point_out");
    a.ert_eps = ert_eps;
    a.ert_hint = ert_hint;
    a.ert_min_count = max(1, (int)(((long long)scene->B * n) / 64));   // rounds pay once ~1.5 % of the rays are opaque
    a.ert_stats = ert_stats;
    a.bg_const = bg_const;
    if (pair_ckpt) {
        ENARF_REQUIRE(pair_ckpt->rec && pair_ckpt->feat && pair_ckpt->tile_base && pair_ckpt->cursor && pair_ckpt->capacity > 0 &&
                          ((uintptr_t)pair_ckpt->rec & 15) == 0 && ((uintptr_t)pair_ckpt->feat & 15) == 0,
                      ENARF_ERR_INVALID_ARG, "enarf_render_fine_fwd: incomplete or misaligned pair checkpoint");
        a.ck = *pair_ckpt;
        a.g.skip_last = pair_ckpt->skip_last ? 1 : 0;
    }
    return launch_render<MODE_FINE>(a, "enarf_render_fine_fwd", (cudaStream_t)stream);
}

extern "C" int enarf_density_query(const enarf_scene_t* scene, const float* points, int N, float* density,
                                   float* color, uint32_t* part_mask, void* stream) {
    int rc = enarf_check_scene(scene, "enarf_density_query");
    if (rc) return rc;
    ENARF_REQUIRE(points && N > 0, ENARF_ERR_INVALID_ARG, "enarf_density_query: bad arguments");
    RenderArgs a = {};
    a.sc = *scene;
    a.g.points = points;
    a.g.n = N;
    a.g.S = 1;
    a.rays_per_tile = TP_MAX;
    a.tiles_per_sample = (N + TP_MAX - 1) / TP_MAX;
    a.q_density = density;
    a.q_color = color;
    a.q_mask = part_mask;
    return launch_render<MODE_QUERY>(a, "enarf_density_query", (cudaStream_t)stream);
}
