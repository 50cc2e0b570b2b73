// Tile framework shared by the fused forward (render.cu) and backward (bwd.cu) kernels.
//
// Persistent CTAs (two 256-thread CTAs per SM, dynamic tile counter) own whole rays: a tile is
// rays x samples <= TP_MAX points.  Per tile
//   phase 0  per (ray, part): conservative clip of the ray segment against the part cube (common.cuh);
//   phase 1  per point: position (contract arithmetic), exact part tests on the candidates -> part mask;
//            points inside at least one part are compacted into `list`; a prefix sum of popc(mask) over
//            the list gives every active point its range of (point, part) pairs;
//   pair records  ONE thread (lane) per (point, part) pair: canonical coordinates, three bilinear footprints,
//            the twelve probability texels, sigmoid product -> a 20-word record in shared memory;
//   gather   8 lanes per point walk the point's records: 12 x 128-bit texel loads per pair, lane l owns
//            channels 4l..4l+3.
// The forward consumes chunks of 32 active points per WARP (gather32_warp below, decoder in mlp_mma.cuh), the
// backward chunks of <= CH points per CTA (setup_pairs / gather_chunk below, GEMMs in mma_smem.cuh) with
// activations as [channel][point] rows of stride AS.  Both run the decoder as error-compensated 3xTF32
// mma.sync (see mlp_mma.cuh for the measured trade-off against FFMA).
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"
#include "gather.cuh"

// Geometry of the CTA; a translation unit pre-defines these before including this header (render.cu and bwd.cu both
// run two 256-thread CTAs per SM; bwd.cu adds 64-point CTA-wide chunks).  The defaults below are the one-CTA-per-SM
// geometry of the earlier versions and still compile.
#ifndef NT
#define NT 512             // threads per CTA
#endif
#define NWARP (NT / 32)
#ifndef CH
#define CH 128             // active points per chunk
#endif
#ifndef AS
#define AS 132             // activation row stride in floats (rows 16-B aligned; AS % 32 == 4: row r starts at bank 4r)
#endif
#define W1S 40             // decoder weight row strides (8 mod 32: 64-bit mma B-fragment loads are conflict-free)
#define W2S 72
#define TP_MAX (2 * NT)    // points per tile
#ifndef MAX_RAYS
#define MAX_RAYS 32        // rays per tile
#endif
#ifndef PB
#define PB 384             // (point, part) pair records per chunk
#endif
#define REC 20             // words per record; 20*r mod 32 hits 8 distinct 4-bank groups for r..r+7
static_assert(TP_MAX == 2 * NT, "the prefix scan handles two list entries per thread");

// record layout (words): 0..2 element offset of texel (y0,x0) of plane k inside the sample's feature planes,
// 3 part id | point slot << 8, 4..15 bilinear weights [k][corner] (forward: already multiplied by the part weight),
// 16 part weight wp = prod sigmoid, 17..19 sigmoid(prob sample k)
struct __align__(16) DecoderSmem {
    float w1[HID * W1S];   // [o][i]
    float w2[HID * W2S];   // [o][i]
    float w3[4 * W2S];     // [c][o]
    float b1[HID], b2[HID], b3[4];
    float pad_[12];
};

struct __align__(16) TileSmem {
    float4 pm[TP_MAX];                 // x, y, z, part mask (bits)
    float depth[TP_MAX + MAX_RAYS];
    float ray[MAX_RAYS * 8];           // dmin, dmax, dx, dy, dz
    uint32_t cand[MAX_RAYS];
    uint16_t list[TP_MAX];
    uint16_t pref[TP_MAX + 8];
    uint32_t act[TP_MAX / 32];         // bit i%32 of word i/32: point i is inside at least one part
    uint32_t sel[TP_MAX / 32];         // compaction scratch: the same bits in the order the list is built in
    float carry[MAX_RAYS];             // early ray termination: optical depth of the sample groups composited so far
    uint32_t alive;                    // early ray termination: bit r set = ray r of the tile is still transparent
    int skipped;                       // active points dropped because their ray had terminated (statistics)
    int wsum[NWARP];
    int n_active;
    int tile;
    int pad_[2];
    int n_next;                        // tile_compact: the list length being built (published as n_active after a barrier)
    int pad2_[3];
};

// Early ray termination (north_star step (e)): the samples of a ray are evaluated front to back in groups
// of ERT_GROUP; once the exact prefix transmittance of a ray has dropped below the caller's threshold its remaining
// samples are not evaluated and composite with density 0.  A dropped sample is marked by a negative density in
// the saved (r,g,b,sigma) record, which is how the backward learns to leave it out too.
#define ERT_GROUP 16
#define ERT_SENTINEL (-1.0f)

union __align__(16) RecSmem {
    float rec[PB * REC];
    float cull[MAX_RAYS * ENARF_MAX_PARTS * 2];   // phase 0/1 only
};

__device__ __forceinline__ void load_decoder_smem(DecoderSmem& d, const enarf_scene_t& sc, int b, int tid) {
    const float* w1 = sc.w1 + (size_t)b * HID * FEAT;
    const float* w2 = sc.w2 + (size_t)b * HID * HID;
    const float* w3 = sc.w3 + (size_t)b * 4 * HID;
    for (int e = tid; e < HID * FEAT; e += NT) d.w1[(e / FEAT) * W1S + (e % FEAT)] = w1[e];
    for (int e = tid; e < HID * HID; e += NT) d.w2[(e / HID) * W2S + (e % HID)] = w2[e];
    for (int e = tid; e < 4 * HID; e += NT) {
        d.w3[(e / HID) * W2S + (e % HID)] = w3[e];
    }
    if (tid < HID) {
        d.b1[tid] = sc.b1[tid];
        d.b2[tid] = sc.b2[tid];
    }
    if (tid < 4) d.b3[tid] = sc.b3[tid];
}

// ---- phase 0 + 1 --------------------------------------------------------------------------------------
enum { TILE_COARSE = 0, TILE_FINE = 1, TILE_QUERY = 2 };

struct TileGeom {
    const float *dmin, *dmax, *dir, *bins, *points;
    float* fine_depth;   // optional global copy of the per-sample depth (fine forward only)
    int n, S;
    int skip_last;       // backward: the last sample of a ray only closes the last interval, no gradient
};

// Loads the tile's rays, clips them against the parts and fills pm / depth / act (the part masks of every point);
// resets the active list.  Ends synchronised.
template <int MODE>
__device__ __forceinline__ void tile_masks(TileSmem& t, RecSmem& r, const float* s_part, const TileGeom& g, int P,
                                            int b, int ray0, int rays, int tid) {
    const int lane = tid & 31, warp = tid >> 5;
    const int S = g.S, n = g.n, tp = rays * S;
    if (tid == 0) {
        t.n_active = 0;
        t.skipped = 0;
    }
    if (MODE != TILE_QUERY && tid < rays) {
        const size_t o = (size_t)b * n + ray0 + tid;
        const float* d = g.dir + (size_t)b * 3 * n + ray0 + tid;
        t.ray[tid * 8 + 0] = g.dmin[o];
        t.ray[tid * 8 + 1] = g.dmax[o];
        t.ray[tid * 8 + 2] = d[0];
        t.ray[tid * 8 + 3] = d[n];
        t.ray[tid * 8 + 4] = d[2 * (size_t)n];
        t.cand[tid] = 0;
    }
    __syncthreads();
    if (MODE != TILE_QUERY) {
        for (int e = tid; e < rays * P; e += NT) {
            const int rr = e / P, p = e - rr * P;
            const float* ry = t.ray + rr * 8;
            float u0, u1;
            if (ray_part_interval(s_part + p * PART_STRIDE, ry[0] * ry[2], ry[0] * ry[3], ry[0] * ry[4], ry[1] * ry[2],
                                  ry[1] * ry[3], ry[1] * ry[4], u0, u1)) {
                r.cull[(rr * ENARF_MAX_PARTS + p) * 2] = u0;
                r.cull[(rr * ENARF_MAX_PARTS + p) * 2 + 1] = u1;
                atomicOr(&t.cand[rr], 1u << p);
            }
        }
        __syncthreads();
    }
    const float cstep = (MODE == TILE_COARSE) ? __fdiv_rn(1.0f, (float)S) : 0.f;
    for (int i = tid; i < ((tp + 31) & ~31); i += NT) {
        uint32_t m = 0;
        if (i < tp) {
            float x, y, z;
            if (MODE == TILE_QUERY) {
                const float* pt = g.points + (size_t)b * 3 * n + ray0 + i;
                x = pt[0];
                y = pt[n];
                z = pt[2 * (size_t)n];
                for (int p = 0; p < P; ++p)
                    if (part_valid(s_part + p * PART_STRIDE, x, y, z)) m |= 1u << p;
            } else {
                const int rr = i / S, j = i - rr * S;
                const float* ry = t.ray + rr * 8;
                const float dmin = ry[0], dmax = ry[1];
                const float sx = cmul(dmin, ry[2]), sy = cmul(dmin, ry[3]), sz = cmul(dmin, ry[4]);
                const float ex = cmul(dmax, ry[2]), ey = cmul(dmax, ry[3]), ez = cmul(dmax, ry[4]);
                float u;
                if (MODE == TILE_COARSE) {
                    // mid-point of consecutive stratified edges (rendering.py:125-131)
                    const float t0 = unit_bin(j, S, cstep), t1 = unit_bin(j + 1, S, cstep);
                    x = cmul(cadd(clerp(sx, ex, t1), clerp(sx, ex, t0)), 0.5f);
                    y = cmul(cadd(clerp(sy, ey, t1), clerp(sy, ey, t0)), 0.5f);
                    z = cmul(cadd(clerp(sz, ez, t1), clerp(sz, ez, t0)), 0.5f);
                    t.depth[rr * (S + 1) + j] = clerp(dmin, dmax, t0);
                    if (j == S - 1) t.depth[rr * (S + 1) + S] = clerp(dmin, dmax, t1);
                    u = 0.5f * (t0 + t1);
                } else {
                    const size_t o = ((size_t)b * n + ray0 + rr) * S + j;
                    u = g.bins[o];
                    x = clerp(sx, ex, u);
                    y = clerp(sy, ey, u);
                    z = clerp(sz, ez, u);
                    const float dep = clerp(dmin, dmax, u);
                    t.depth[rr * (S + 1) + j] = dep;
                    if (g.fine_depth) g.fine_depth[o] = dep;
                }
                if (!(g.skip_last && j == S - 1)) {
                    uint32_t cm = t.cand[rr];
                    const float* cu = r.cull + rr * ENARF_MAX_PARTS * 2;
                    while (cm) {
                        const int p = __ffs(cm) - 1;
                        cm &= cm - 1;
                        if (u >= cu[2 * p] && u <= cu[2 * p + 1] && part_valid(s_part + p * PART_STRIDE, x, y, z))
                            m |= 1u << p;
                    }
                }
            }
            t.pm[i] = make_float4(x, y, z, __uint_as_float(m));
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, m != 0);
        if (lane == 0) t.act[i >> 5] = bal;
    }
    __syncthreads();
}

// ---- compaction -----------------------------------------------------------------------------------------
// Appends to t.list the active points of the sample groups [g0, g1) -- group g = samples [g*K, min(S, (g+1)*K)) of
// every ray -- in (group, ray, sample) order, and extends the pair prefix t.pref over them.  With K >= S there is one
// group and the order is plain point order.  Deterministic: the forward and the backward of a tile build the same
// list, and a chunk's points are consecutive along their rays.  Only rays whose bit is set in `alive` contribute; an
// active point of a dead ray is marked ERT_SENTINEL instead (forward).  `saved` (backward): the forward's
// (r,g,b,sigma) records -- points the forward dropped (sigma < 0) are left out.  Call with t.n_active = 0 for the
// first range of a tile.  Ends synchronised.
__device__ __forceinline__ void tile_compact(TileSmem& t, int rays, int S, int K, int g0, int g1, uint32_t alive,
                                             const float4* saved, int tid) {
    const int lane = tid & 31, warp = tid >> 5;
    const int s0 = g0 * K, s1 = min(S, g1 * K);
    const int cnt = rays * (s1 - s0), cnt32 = (cnt + 31) & ~31;
    const int n0 = t.n_active;
    // position q of the range -> point index i = ray * S + sample
    auto point_of = [&](int q, int& rr) {
        const int g = q / (rays * K);                 // every group but the last holds rays*K positions
        const int within = q - g * rays * K, base = (g0 + g) * K;
        const int kg = min(K, S - base);
        rr = within / kg;
        return rr * S + base + (within - rr * kg);
    };
    const bool all_alive = alive == 0xffffffffu;      // (point queries pass > 32 "rays")
    // plain point order over the whole tile with nothing to filter: the part-mask bitmap already is the selection
    const bool direct = K >= S && all_alive && !saved;
    if (!direct) {
        int dropped = 0;
        for (int q = tid; q < cnt32; q += NT) {
            bool on = false;
            if (q < cnt) {
                int rr;
                const int i = point_of(q, rr);
                on = (t.act[i >> 5] >> (i & 31)) & 1;
                if (on && saved && saved[i].w < 0.f) on = false;
                if (on && !all_alive && !((alive >> rr) & 1)) {
                    on = false;
                    t.pm[i] = make_float4(0.f, 0.f, 0.f, ERT_SENTINEL);
                    ++dropped;
                }
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, on);
            if (lane == 0) t.sel[q >> 5] = bal;
        }
        if (dropped) atomicAdd(&t.skipped, dropped);
        __syncthreads();
    }
    const uint32_t* bits = direct ? t.act : t.sel;
    for (int q = tid; q < cnt32; q += NT) {
        const uint32_t bal = bits[q >> 5];
        int base = n0;
        for (int k = 0; k < (q >> 5); ++k) base += __popc(bits[k]);
        if ((bal >> lane) & 1) {
            int rr;
            ENARF_DBG_ASSERT(base + __popc(bal & ((1u << lane) - 1)) < TP_MAX);
            t.list[base + __popc(bal & ((1u << lane) - 1))] = (uint16_t)(direct ? q : point_of(q, rr));
            ENARF_DBG_ASSERT((direct ? q : point_of(q, rr)) < rays * S);
        }
        // the last block closes the count -- into n_next, NOT n_active: a warp that is late entering this function
        // (the point query stores its part masks to global memory just before) must still read the old n_active as n0
        if (q + 32 >= cnt32 && lane == 0) t.n_next = base + __popc(bal);
    }
    if (cnt32 == 0 && tid == 0) t.n_next = n0;
    __syncthreads();
    // exclusive prefix of pairs-per-point over the new list entries (2 entries per thread)
    const int na = t.n_next - n0;
    if (tid == 0) t.n_active = n0 + na;      // visible to every thread after the two barriers below
    const int p_base = n0 ? t.pref[n0] : 0;
    const int i0 = 2 * tid;
    const int c0 = i0 < na ? __popc(__float_as_uint(t.pm[t.list[n0 + i0]].w)) : 0;
    const int c1 = i0 + 1 < na ? __popc(__float_as_uint(t.pm[t.list[n0 + i0 + 1]].w)) : 0;
    const int s2 = c0 + c1;
    int incl = s2;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) t.wsum[warp] = incl;
    __syncthreads();
    int base = p_base;
    for (int w = 0; w < warp; ++w) base += t.wsum[w];
    const int excl = base + incl - s2;
    if (i0 <= na) t.pref[n0 + i0] = (uint16_t)excl;
    if (i0 + 1 <= na) t.pref[n0 + i0 + 1] = (uint16_t)(excl + c0);
    if (i0 + 2 == na && tid == NT - 1) t.pref[n0 + na] = (uint16_t)(excl + s2);   // na == 2*NT: nobody owns pref[na]
    __syncthreads();
}

// tile_masks + one compaction of every point in point order: the geometry every kernel without early termination uses
template <int MODE>
__device__ __forceinline__ void tile_points(TileSmem& t, RecSmem& r, const float* s_part, const TileGeom& g, int P,
                                            int b, int ray0, int rays, int tid) {
    tile_masks<MODE>(t, r, s_part, g, P, b, ray0, rays, tid);
    tile_compact(t, rays, g.S, g.S, 0, 1, 0xffffffffu, nullptr, tid);
}

// End (exclusive) of the chunk starting at active index a: at most CH points and `cap` pairs (cap >= 32).
__device__ __forceinline__ int chunk_end(const TileSmem& t, int a, int cap = PB) {
    const int na = t.n_active;
    int e = min(a + CH, na);
    const int p0 = t.pref[a];
    if (t.pref[e] - p0 > cap) {
        int lo = a + 1, hi = e;   // invariant: pref[lo]-p0 <= cap (one point has <= 32 pairs), answer in [lo, hi)
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (t.pref[mid] - p0 <= cap) lo = mid; else hi = mid;
        }
        e = lo;
    }
    return e;
}

// ---- pair setup: one thread per (point, part) pair of the chunk [a, e) --------------------------------
// PREMUL: store bilinear weights already multiplied by the part weight (forward); otherwise raw (backward).
template <bool PREMUL>
__device__ __forceinline__ void setup_pairs(const TileSmem& t, float* rec, const float* s_part, const enarf_scene_t& sc,
                                            int b, int a, int e, int tid, int nthr = NT) {
    const int p0 = t.pref[a], npairs = t.pref[e] - p0;
    const int H = sc.H, W = sc.W;
    const float* prob = sc.prob_planes + (size_t)b * sc.prob_batch_stride;
    for (int r = tid; r < npairs; r += nthr) {
        const int target = p0 + r;
        int lo = a, hi = e;   // largest q in [a, e) with pref[q] <= target
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (t.pref[mid] <= target) lo = mid; else hi = mid;
        }
        const float4 pm = t.pm[t.list[lo]];
        uint32_t m = __float_as_uint(pm.w);
        for (int k = target - t.pref[lo]; k > 0; --k) m &= m - 1;
        const int p = __ffs(m) - 1;
        float c[3];
        part_canonical(s_part + p * PART_STRIDE, pm.x, pm.y, pm.z, c[0], c[1], c[2]);
        Bilin bl[3];
        bilin_setup3(c, W, H, bl);
        float sg[3], wp = 1.f;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            sg[k] = sigmoidf_(bilin_fetch1(prob + (size_t)(3 * p + k) * H * W, bl[k], W));
            wp *= sg[k];
        }
        const float ws = PREMUL ? wp : 1.f;
        float4* dst = reinterpret_cast<float4*>(rec + r * REC);
        int off[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) off[k] = (k * H * W + bl[k].y0 * W + bl[k].x0) * FEAT;
        // word 3: part id | (point slot within the chunk << 8)
        dst[0] = make_float4(__int_as_float(off[0]), __int_as_float(off[1]), __int_as_float(off[2]),
                             __int_as_float(p | ((lo - a) << 8)));
#pragma unroll
        for (int k = 0; k < 3; ++k)
            dst[1 + k] = make_float4(bl[k].w00 * ws, bl[k].w01 * ws, bl[k].w10 * ws, bl[k].w11 * ws);
        dst[4] = make_float4(wp, sg[0], sg[1], sg[2]);
    }
}

// ---- gather: 8 lanes per point, lane l owns channels 4l..4l+3; writes X[channel][q] -------------------
// SCALE: multiply the stored weights by the part weight on the fly (records hold raw weights).
template <typename T, bool SCALE>
__device__ __forceinline__ void gather_chunk(const TileSmem& t, const float* rec, float* X, const enarf_scene_t& sc,
                                             int b, int a, int e, int tid, int nthr = NT) {
    const int lane8 = tid & 7, cnt = e - a, p0 = t.pref[a];
    const int W = sc.W;
    const T* feat = reinterpret_cast<const T*>(sc.feat_planes) + (size_t)b * 3 * sc.H * W * FEAT + lane8 * 4;
    for (int q = tid >> 3; q < cnt; q += nthr / 8) {
        float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
        const int r1 = t.pref[a + q + 1] - p0;
        for (int r = t.pref[a + q] - p0; r < r1; ++r)
            gather_pair<T>(f, feat, rec + r * REC, W * FEAT, SCALE ? rec[r * REC + 16] : 1.f);
        float* dst = X + (lane8 * 4) * AS + q;
        dst[0] = f.x;
        dst[AS] = f.y;
        dst[2 * AS] = f.z;
        dst[3 * AS] = f.w;
    }
    // columns [cnt, roundup32(cnt)) feed the decoder tiles of the last partially filled warp: keep them finite
    const int cend = (cnt + 31) & ~31;
    for (int i = tid; i < FEAT * (cend - cnt); i += nthr) {
        const int row = i / (cend - cnt), col = cnt + i - row * (cend - cnt);
        X[row * AS + col] = 0.f;
    }
}

// ---- warp-autonomous gather (forward): one warp owns the 32 active points [a, a+np) of a chunk ----------
// Records go to the warp's own slab `rec` (RECW records), features to its slab X[point][XSW]; only __syncwarp.
#ifndef RECW
#define RECW 48     // pair records per warp sub-batch (one point has at most 32 pairs)
#endif
static_assert(RECW <= 64, "the pair-owner bitmap of gather32_warp is 64 bits");
#define XSW 32      // feature slab row stride; 8-float granules are XOR-swizzled by the row (mlp_mma.cuh xs_col)

#ifndef FWD_PHASE   // render.cu's development-build phase timer
#define FWD_PHASE(i)
#define FWD_TIMER_PARAM
#endif
// Pair checkpoint (include/enarf_b200.h enarf_pair_ckpt_t): ck_rec / ck_feat already point at the tile's first pair,
// or are NULL.  The checkpointed record is the backward's 20-word layout with the weights premultiplied by the
// part weight: 0..2 offsets, 3 part | active index << 8, 4..15 weights, 16 1/part weight (0 if the weight is 0),
// 17..19 sigmoids.
template <typename T>
__device__ __forceinline__ void gather32_warp(const TileSmem& t, float* rec, float* X, const float* s_part,
                                              const enarf_scene_t& sc, int b, int a, int np, int lane,
                                              float* __restrict__ ck_rec, float* __restrict__ ck_feat FWD_TIMER_PARAM) {
    const int H = sc.H, W = sc.W;
    const float* prob = sc.prob_planes + (size_t)b * sc.prob_batch_stride;
    const int lane8 = lane & 7, grp = lane >> 3;
    const T* feat = reinterpret_cast<const T*>(sc.feat_planes) + (size_t)b * 3 * H * W * FEAT + lane8 * 4;
    for (int q0 = 0; q0 < np;) {
        // points [q0, q1) whose pairs fit the record slab
        const int base = t.pref[a + q0];
        const int mine = (lane >= q0 && lane < np) ? t.pref[a + lane + 1] - base : (RECW + 1);
        const int q1 = q0 + __popc(__ballot_sync(0xffffffffu, mine <= RECW));
        const int npairs = t.pref[a + q1] - base;
        // which point owns pair r: bit s of `starts` is set when a point's pairs begin at record s (every point has at
        // least one pair, RECW <= 64) -- the owner of r is the number of set bits at or below r, its first record the
        // highest of them: two warp reductions and a few bit operations instead of a binary search in shared memory
        unsigned long long starts;
        {
            const int st = (lane >= q0 && lane < q1) ? t.pref[a + lane] - base : 64;
            const uint32_t lo32 = __reduce_or_sync(0xffffffffu, st < 32 ? 1u << st : 0u);
            const uint32_t hi32 = __reduce_or_sync(0xffffffffu, (st >= 32 && st < 64) ? 1u << (st - 32) : 0u);
            starts = ((unsigned long long)hi32 << 32) | lo32;
        }
        // ---- setup: one lane per pair ------------------------------------------------------------------
        for (int r = lane; r < npairs; r += 32) {
            const int target = base + r;
            const unsigned long long below = starts & (~0ull >> (63 - r));
            const int lo = a + q0 + __popcll(below) - 1;
            const float4 pm = t.pm[t.list[lo]];
            uint32_t m = __float_as_uint(pm.w);
            for (int k = r - (63 - __clzll(below)); k > 0; --k) m &= m - 1;
            const int p = __ffs(m) - 1;
            ENARF_DBG_ASSERT(p >= 0 && p < sc.P && r < RECW);
            float c[3];
            part_canonical(s_part + p * PART_STRIDE, pm.x, pm.y, pm.z, c[0], c[1], c[2]);
            Bilin bl[3];
            bilin_setup3(c, W, H, bl);
            float sg[3], wp = 1.f;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                ENARF_DBG_ASSERT(bl[k].x0 >= 0 && bl[k].x0 <= W - 2 && bl[k].y0 >= 0 && bl[k].y0 <= H - 2);
                sg[k] = sigmoidf_(bilin_fetch1(prob + (size_t)(3 * p + k) * H * W, bl[k], W));
                wp *= sg[k];
            }
            float4 rw[4];
            rw[0] = make_float4(__int_as_float((bl[0].y0 * W + bl[0].x0) * FEAT),
                                __int_as_float((H * W + bl[1].y0 * W + bl[1].x0) * FEAT),
                                __int_as_float((2 * H * W + bl[2].y0 * W + bl[2].x0) * FEAT), __int_as_float(p | (lo << 8)));
#pragma unroll
            for (int k = 0; k < 3; ++k)
                rw[1 + k] = make_float4(bl[k].w00 * wp, bl[k].w01 * wp, bl[k].w10 * wp, bl[k].w11 * wp);
            float4* dst = reinterpret_cast<float4*>(rec + r * REC);
#pragma unroll
            for (int k = 0; k < 4; ++k) dst[k] = rw[k];
            if (ck_rec) {
                ENARF_DBG_ASSERT(target >= 0);
                // streaming stores: the checkpoint is written once and must not push the planes out of L2
                float4* g = reinterpret_cast<float4*>(ck_rec + (size_t)target * REC);
#pragma unroll
                for (int k = 0; k < 4; ++k) __stcs(g + k, rw[k]);
                __stcs(g + 4, make_float4(wp != 0.f ? __frcp_rn(wp) : 0.f, sg[0], sg[1], sg[2]));
            }
        }
        __syncwarp();
        FWD_PHASE(1);   // pair set-up
        // ---- gather: 8 lanes per point, the four groups take points q0+grp, q0+grp+4, ... ---------------
        for (int q = q0 + grp; q < q1; q += 4) {
            float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
            const int r1 = t.pref[a + q + 1] - base;
            for (int r = t.pref[a + q] - base; r < r1; ++r) {
                // the pair's own contribution, summed over the point's parts in part order (the same arithmetic whether
                // or not it is also kept for the backward)
                float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
                gather_pair<T>(g, feat, rec + r * REC, W * FEAT, 1.f);
                if (ck_feat) __stcs(reinterpret_cast<float4*>(ck_feat + (size_t)(base + r) * FEAT + 4 * lane8), g);
                f.x += g.x; f.y += g.y; f.z += g.z; f.w += g.w;
            }
            *reinterpret_cast<float4*>(X + q * XSW + ((4 * lane8) ^ (8 * (q & 3)))) = f;
        }
        __syncwarp();
        FWD_PHASE(2);   // gather
        q0 = q1;
    }
}
