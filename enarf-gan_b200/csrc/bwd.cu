// Fused backward of the fine pass: compositing backward -> decoder backward (dgrad + wgrad) -> tri-plane
// scatter.  Gradients flow to the feature planes, the part-probability planes and the per-sample decoder
// weights / biases only (pose, rays and bins are non-differentiable: libraries/NeRF/rendering.py:216-217).
// Reference being replaced: autograd through libraries/NeRF/rendering.py:307-335, models/narf.py:213-275,
// libraries/triplane/sampling.py:54-127 (grid_sample backward + scatter_add backward), NeRF/activation.py:12-16.
//
// Same tile framework as the forward (tile.cuh), but with two independent 256-thread CTAs per SM working on
// 64-point chunks: the phases of a chunk are separated by CTA-wide barriers and alternate between latency-bound
// (checkpoint load, pair setup, scatter) and tensor-pipe-bound (the GEMMs) work, so a second resident CTA fills
// the issue slots the first one leaves idle.  Shared memory is kept under 113 KB per CTA for that.  Per tile
//   1. positions + part masks are regenerated with the forward's contract arithmetic (bit-identical masks);
//   2. one warp per ray turns the upstream gradients and the forward's saved (rgb, sigma) into per-sample
//      d/d rgb and d/d sigma (two shuffle scans: transmittance prefix, weighted-gradient suffix), in place;
//   3. per chunk of <= 64 active points / 160 pairs: pair records (the forward's checkpoint, fetched by one TMA bulk
//      copy, or rebuilt) -> features (checkpoint rows, or gathered) -> decoder forward (activations kept as
//      [channel][point] rows) -> head backward -> G2 -> dW2 -> G1 -> dW1 -> dX, all six GEMMs as error-compensated
//      3xTF32 mma.sync from shared-memory fragments; the weight gradients are accumulated in registers over the
//      CTA's whole lifetime and flushed with atomics once per (CTA, sample);
//   4. d(feature) is scattered with 128-bit vector reductions (red.global.add.v4.f32) into the channels-last
//      gradient planes using the same pair records; d(part weight) goes through the sigmoid product into
//      scalar reductions on the probability planes.
#define NT 256
#define CH 64
#define AS 68
#define MAX_RAYS 16
#define PB 160
#include "tile.cuh"
#include "mma_smem.cuh"

#define PBLK (CH / 32)        // 32-point blocks per chunk
// d(features) of a chunk is handed to the scatter as [point][DXS] rows (the dead G2 buffer): the 8 lanes of a pair read
// one 128-bit word each -- 128 contiguous bytes, conflict-free -- and the mma epilogue's 64-bit stores are conflict-free
// with a row stride of 40 words (8 g + 2 t covers the 32 banks once per half warp).  The [channel][point] form it
// replaces cost four 4-way-conflicting 32-bit loads per lane: 34 % of the kernel's excess shared-memory wavefronts.
#define DXS 40
// Pair budget of a chunk.  The record buffer holds PB pairs; with the forward's checkpoint the records beyond PB are fetched
// into the G1 buffer once it is dead (after the dW1 / dX phase), so a chunk is its 64 points (3.6 pairs per point on the
// bench batch) instead of the ~45 points whose pairs fit PB: a quarter fewer chunks, i.e. of everything that costs per chunk
// (six GEMM phases on 64-row tiles, the head, eight barriers).
#define PBX ((HID * AS / REC) & ~7)
static_assert(CH * DXS <= HID * AS, "dX rows fit the G2 buffer");
#define WCOL (NWARP / 4)      // column groups of the weight-gradient tiles (4 row blocks of 16 outs each)
#define TN2 (8 / WCOL)        // 8-column mma tiles per warp: dW2 (64 columns) / dW1 (32 columns)
#define TN1 (4 / WCOL)
static_assert(NWARP / PBLK == 4, "a warp's mma tile is 32 points x 16 outs");

struct BwdArgs {
    enarf_scene_t sc;
    TileGeom g;
    const float* saved;       // (B,n,S,4)
    const float* feat;        // (B,n,S,32) forward's feature checkpoint | NULL (gather again)
    enarf_pair_ckpt_t ck;     // forward's per-pair checkpoint (rec == NULL: none)
    const float* g_color;     // (B,3,n) | NULL
    const float* g_mask;      // (B,n)   | NULL
    const float* g_disp;      // (B,n)   | NULL
    const float* g_fw;        // (B,n,S-1) | NULL
    float* d_feat;
    float* d_prob;
    long long d_prob_stride;
    float *d_w1, *d_w2, *d_w3, *d_b1, *d_b2, *d_b3;
    int rays_per_tile, tiles_per_sample;
    int* tile_counter;
    float bg_const;           // the forward composited colour = rgb + (1 - mask) * bg_const
    float ert_eps;            // > 0: the forward ran with early ray termination (group-major pair order, dropped samples)
};

struct __align__(16) BwdSmem {
    DecoderSmem dec;
    float part[ENARF_MAX_PARTS * PART_STRIDE];
    TileSmem t;
    RecSmem r;
    float X[FEAT * AS];      // features, later d(features)
    float A1[HID * AS];      // layer-1 activations, later their pre-activation gradients
    float A2[HID * AS];
    float G3[4 * AS];
    float4 pg[TP_MAX];       // forward (r,g,b,sigma), overwritten by d/d(r,g,b,sigma)
    float rg[MAX_RAYS * 8];  // upstream d/d(color rgb, mask, disparity) per ray
    float db3[4];
    unsigned long long rec_bar[2];   // mbarriers of the records' bulk copies: first PB pairs, overflow
};

#ifdef ENARF_PHASE_TIMING   // development build: per-phase clock64() sums of thread 0 of every CTA
__device__ unsigned long long g_bwd_phase[16];
#define PHASE_DECL long long ph_t = clock64(); unsigned long long ph_acc[16] = {}
#define PHASE(i) do { if (tid == 0) { const long long now_ = clock64(); ph_acc[i] += (unsigned long long)(now_ - ph_t); ph_t = now_; } } while (0)
#define PHASE_FLUSH do { if (tid == 0) for (int i_ = 0; i_ < 16; ++i_) atomicAdd(&g_bwd_phase[i_], ph_acc[i_]); } while (0)
extern "C" int enarf_debug_bwd_phases(unsigned long long* host_out, int reset) {
    cudaMemcpyFromSymbol(host_out, g_bwd_phase, sizeof(g_bwd_phase));
    if (reset) {
        unsigned long long z[16] = {};
        cudaMemcpyToSymbol(g_bwd_phase, z, sizeof(z));
    }
    return 0;
}
#else
#define PHASE_DECL
#define PHASE(i)
#define PHASE_FLUSH
#endif

// ---- TMA 1-D bulk copy (cp.async.bulk) of a contiguous global range into shared memory, completion on an mbarrier --
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src, uint32_t bytes, unsigned long long* bar) {
    const uint32_t b = (uint32_t)__cvta_generic_to_shared(bar);
    // the destination was last touched through the generic proxy (the previous chunk's scatter read it)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(dst_smem)),
                 "l"(src), "r"(bytes), "r"(b)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    const uint32_t b = (uint32_t)__cvta_generic_to_shared(bar);
    asm volatile(
        "{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(b), "r"(parity) : "memory");
}

// The tile's depths and active set from the forward's per-sample part counts instead of the 24-part geometry (checkpointed
// tiles only: their pair records already hold everything else).  A sample with c valid parts gets the stand-in mask
// (1 << c) - 1: tile_compact only looks at its population count.  Ends synchronised, like tile_masks.
__device__ __forceinline__ void tile_from_counts(TileSmem& t, const TileGeom& g, const uint8_t* __restrict__ count, int b,
                                                 int ray0, int rays, int tid) {
    const int lane = tid & 31;
    const int S = g.S, n = g.n, tp = rays * S;
    if (tid == 0) {
        t.n_active = 0;
        t.skipped = 0;
    }
    if (tid < rays) {
        const size_t o = (size_t)b * n + ray0 + tid;
        t.ray[tid * 8 + 0] = g.dmin[o];
        t.ray[tid * 8 + 1] = g.dmax[o];
    }
    __syncthreads();
    const size_t p0 = ((size_t)b * n + ray0) * S;
    for (int i = tid; i < ((tp + 31) & ~31); i += NT) {
        uint32_t m = 0;
        if (i < tp) {
            const int rr = i / S, j = i - rr * S;
            t.depth[rr * (S + 1) + j] = clerp(t.ray[rr * 8], t.ray[rr * 8 + 1], g.bins[p0 + i]);
            const uint32_t c = count[p0 + i];
            m = c >= 32u ? 0xffffffffu : (1u << c) - 1u;
            t.pm[i] = make_float4(0.f, 0.f, 0.f, __uint_as_float(m));
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, m != 0);
        if (lane == 0) t.act[i >> 5] = bal;
    }
    __syncthreads();
}

__device__ __forceinline__ float act_grad(float a) { return (a > 0.f ? 1.f : 0.2f) * 1.4142135623730951f; }

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    // no "memory" clobber: nothing in this kernel reads the gradient planes, and a clobber would pin every
    // texel load behind the previous reduction (one exposed L2 round trip per corner)
#ifdef ENARF_PROBE_NO_RED     // timing probe (results are wrong): what the kernel costs without its feature-plane reductions
    if (a == 123.456f) asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d));
#else
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d));
#endif
}

// per-thread weight-gradient accumulators, alive across tiles
struct WGrad {
    float w2[1][TN2][4]; // mma tiles: rows (o) 16*(warp&3)+g(+8), cols (i) 8*TN2*(warp>>2) + 8 nt + 2t(+1)
    float w1[1][TN1][4]; // rows (o) 16*(warp&3)+g(+8), cols (c) 8*TN1*(warp>>2) + 8 nt + 2t(+1)
    float w3[4][4];      // [c][j]: row c, col og + 16 j, summed over this thread's point quad
    float b2[4];         // cols og + 16 j, this thread's point quad
    float b1[2][2];      // cols 16*ob + 8 nt + 2t + {0,1}, this lane's rows
};

__device__ __forceinline__ void wgrad_zero(WGrad& g) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) g.w3[i][j] = 0.f;
#pragma unroll
        for (int j = 0; j < TN2; ++j) g.w2[0][j][i] = 0.f;
#pragma unroll
        for (int j = 0; j < TN1; ++j) g.w1[0][j][i] = 0.f;
        g.b2[i] = 0.f;
    }
    g.b1[0][0] = g.b1[0][1] = g.b1[1][0] = g.b1[1][1] = 0.f;
}

// thread -> tile coordinates: (pq, og) for the element-wise FP32 passes (4 points x outs og+16j),
// (pb, ob, g, t) for the point-major mma phases (warp tile: 32 points x 16 outs), (wr, wc) for the
// weight-gradient tiles (rows 16 wr.., column group wc)
struct Maps {
    int pq, og;
    int pb, ob, g, t;
    int wr, wc;
};

__device__ __forceinline__ Maps make_maps(int tid) {
    const int w = tid >> 5, lane = tid & 31;
    Maps m;
    m.pb = w % PBLK;
    m.ob = w / PBLK;
    m.pq = m.pb * 8 + (lane & 7);
    m.og = m.ob * 4 + (lane >> 3);
    m.wr = w & 3;
    m.wc = w >> 2;
    m.g = lane >> 2;
    m.t = lane & 3;
    return m;
}

__device__ __forceinline__ void wgrad_flush(WGrad& wg, const Maps& m, float* db3s, const BwdArgs& a, int b, int tid) {
    float* d1 = a.d_w1 + (size_t)b * HID * FEAT;
    float* d2 = a.d_w2 + (size_t)b * HID * HID;
    float* d3 = a.d_w3 + (size_t)b * 4 * HID;
    const int r0 = 16 * m.wr + m.g;
#pragma unroll
    for (int nt = 0; nt < TN2; ++nt) {
        const int c = 8 * TN2 * m.wc + 8 * nt + 2 * m.t;
        atomicAdd(d2 + r0 * HID + c, wg.w2[0][nt][0]);
        atomicAdd(d2 + r0 * HID + c + 1, wg.w2[0][nt][1]);
        atomicAdd(d2 + (r0 + 8) * HID + c, wg.w2[0][nt][2]);
        atomicAdd(d2 + (r0 + 8) * HID + c + 1, wg.w2[0][nt][3]);
    }
#pragma unroll
    for (int nt = 0; nt < TN1; ++nt) {
        const int c = 8 * TN1 * m.wc + 8 * nt + 2 * m.t;
        atomicAdd(d1 + r0 * FEAT + c, wg.w1[0][nt][0]);
        atomicAdd(d1 + r0 * FEAT + c + 1, wg.w1[0][nt][1]);
        atomicAdd(d1 + (r0 + 8) * FEAT + c, wg.w1[0][nt][2]);
        atomicAdd(d1 + (r0 + 8) * FEAT + c + 1, wg.w1[0][nt][3]);
    }
    // dW3 / db2: the 8 lanes that share an out group hold partial sums over different point quads
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float v = wg.w3[c][j];
            v += __shfl_xor_sync(0xffffffffu, v, 1);
            v += __shfl_xor_sync(0xffffffffu, v, 2);
            v += __shfl_xor_sync(0xffffffffu, v, 4);
            if ((tid & 7) == 0) atomicAdd(d3 + c * HID + m.og + 16 * j, v);
        }
        float v = wg.b2[j];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        if ((tid & 7) == 0) atomicAdd(a.d_b2 + m.og + 16 * j, v);
    }
    // db1: lanes with the same t hold partial sums over different rows
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            float v = wg.b1[nt][i];
            v += __shfl_xor_sync(0xffffffffu, v, 4);
            v += __shfl_xor_sync(0xffffffffu, v, 8);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            if (m.g == 0) atomicAdd(a.d_b1 + 16 * m.ob + 8 * nt + 2 * m.t + i, v);
        }
    __syncthreads();
    if (tid < 4) {
        atomicAdd(a.d_b3 + tid, db3s[tid]);
        db3s[tid] = 0.f;
    }
    wgrad_zero(wg);
}

// forward layer on the tensor pipe: dst[o][pt] = lrelu(sum_k src[k][pt] W[o][k] + bias[o]) * sqrt2
// warp tile: points 32*pb .. +31, outs 16*ob .. +15
template <int K, bool FAST>
__device__ __forceinline__ void decoder_layer(const float* src, const float* Wm, int ws, const float* bias, float* dst,
                                              const Maps& m) {
    float acc[2][2][4];
    gemm_pts_mma<K, 2, false, FAST>(src, Wm, ws, 32 * m.pb, 16 * m.ob, m.g, m.t, acc);
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
        const int col = 16 * m.ob + 8 * nt + 2 * m.t;
        const float b0 = bias[col], b1 = bias[col + 1];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            float* d0 = dst + col * AS + 32 * m.pb + 16 * mt + m.g;
            d0[0] = lrelu_s2(acc[mt][nt][0] + b0);
            d0[AS] = lrelu_s2(acc[mt][nt][1] + b1);
            d0[8] = lrelu_s2(acc[mt][nt][2] + b0);
            d0[AS + 8] = lrelu_s2(acc[mt][nt][3] + b1);
        }
    }
}

// Scatter for the chunk [c0, c1): 8 lanes per (point, part) PAIR -- pairs are independent here, so the work
// is spread evenly over the 64 groups whatever the number of parts per point; lane l owns channels 4l..4l+3.
// CK: the records come from the forward's pair checkpoint (weights already multiplied by the part weight, word 3 =
// part | active index << 8, word 16 = 1 / part weight) and ck_feat points at the chunk's first checkpointed pair
// feature, which replaces the 12 texel loads: <dX, f_p> = <dX, wp f_p> / wp.
// Records [0, PB) of the chunk are in `rec`; with the checkpoint a chunk may hold up to PB + PBX pairs, the rest sit in
// `rec_x` (the dead G1 buffer), fetched by a second bulk copy that completes on `bar_x` (parity `phase_x`).
template <typename T, bool CK>
__device__ __forceinline__ void scatter_chunk(const TileSmem& t, const float* rec, const float* rec_x,
                                              unsigned long long* bar_x, uint32_t phase_x, const float* X, const BwdArgs& a,
                                              int b, int c0, int c1, int tid, const float* __restrict__ ck_feat) {
    const enarf_scene_t& sc = a.sc;
    const int lane8 = tid & 7, npairs = t.pref[c1] - t.pref[c0];
    const int H = sc.H, W = sc.W, HW = H * W;
    const unsigned gmask = 0xffu << (tid & 24);
    const T* feat = reinterpret_cast<const T*>(sc.feat_planes) + (size_t)b * 3 * HW * FEAT + lane8 * 4;
    float* dfeat = a.d_feat + (size_t)b * 3 * HW * FEAT + lane8 * 4;
    float* dprob = a.d_prob + (size_t)b * a.d_prob_stride;
    bool waited = false;
    for (int r = tid >> 3; r < npairs; r += NT / 8) {
        if (CK && r >= PB && !waited) {   // the overflow records have landed (each thread observes the barrier itself)
            mbar_wait(bar_x, phase_x);
            waited = true;
        }
        const float* rw = (CK && r >= PB) ? rec_x + (r - PB) * REC : rec + r * REC;
        const float4* rp = reinterpret_cast<const float4*>(rw);
        const float4 o4 = rp[0];
        const unsigned off[3] = {__float_as_uint(o4.x), __float_as_uint(o4.y), __float_as_uint(o4.z)};
        const int p = __float_as_int(o4.w) & 255, q = (__float_as_int(o4.w) >> 8) - (CK ? c0 : 0);
        const float wp = rp[4].x;   // CK: 1 / part weight
        // all texel loads first (independent, always in range), then the dot products and the reductions
        float4 v[CK ? 1 : 12];
        if (CK) {
            // (fetching the next pair's line one iteration ahead measured 1.4 % slower: profiles/bwd_scatter_r2.txt)
            v[0] = __ldg(reinterpret_cast<const float4*>(ck_feat + (size_t)r * FEAT + 4 * lane8));
        } else {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const T* t0 = feat + off[k];
                const T* t1 = t0 + W * FEAT;
                v[(4 * k + 0) % (CK ? 1 : 12)] = load_feat4(t0);
                v[(4 * k + 1) % (CK ? 1 : 12)] = load_feat4(t0 + FEAT);
                v[(4 * k + 2) % (CK ? 1 : 12)] = load_feat4(t1);
                v[(4 * k + 3) % (CK ? 1 : 12)] = load_feat4(t1 + FEAT);
            }
        }
        const float4 dx = *reinterpret_cast<const float4*>(X + q * DXS + 4 * lane8);
        const float4 dxw = CK ? dx : make_float4(dx.x * wp, dx.y * wp, dx.z * wp, dx.w * wp);
        // the reductions depend only on the weights and dX: they are issued while the texel loads are still in flight
        float w[12];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float4 w4 = rp[1 + k];
            w[4 * k] = w4.x; w[4 * k + 1] = w4.y; w[4 * k + 2] = w4.z; w[4 * k + 3] = w4.w;
            float* d0 = dfeat + off[k];
            float* d1 = d0 + W * FEAT;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                // a zero weight (footprint shifted in at an image edge) adds 0: not worth a branch
                const float wc = w[4 * k + c];
                red_add_v4((c >> 1 ? d1 : d0) + (c & 1) * FEAT, wc * dxw.x, wc * dxw.y, wc * dxw.z, wc * dxw.w);
            }
        }
        float4 fs = make_float4(0.f, 0.f, 0.f, 0.f);   // this pair's raw bilinear feature, this lane's channels
        if (CK) {
            fs = make_float4(v[0].x * wp, v[0].y * wp, v[0].z * wp, v[0].w * wp);
        } else {
#pragma unroll
            for (int e = 0; e < 12; ++e) fma4(fs, w[e], v[e % (CK ? 1 : 12)]);
        }
        float fdot = (fs.x * dx.x + fs.y * dx.y) + (fs.z * dx.z + fs.w * dx.w);   // <dx, f_p> on this lane's channels
        fdot += __shfl_xor_sync(gmask, fdot, 1);
        fdot += __shfl_xor_sync(gmask, fdot, 2);
        fdot += __shfl_xor_sync(gmask, fdot, 4);
        // d w_p = fdot;  w_p = prod sigmoid(v_k)  ->  d v_k = fdot * w_p * (1 - sigmoid(v_k))
        for (int e = lane8; e < 12; e += 8) {
            const int k = e >> 2, c = e & 3;
            const float bw = rw[4 + e];
            if (bw != 0.f) {
                // CK: bw already carries the part weight
                const float dv = CK ? fdot * (1.f - rw[17 + k]) : fdot * wp * (1.f - rw[17 + k]);
                const int pix = (off[k] >> 5) - k * HW + (c >> 1) * W + (c & 1);
                ENARF_DBG_ASSERT(pix >= 0 && pix < HW && p >= 0 && p < sc.P && q >= 0 && q < CH);
                atomicAdd(dprob + (size_t)(3 * p + k) * HW + pix, bw * dv);
            }
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(NT, 2) render_bwd_kernel(const BwdArgs a) {
    // bf16 feature planes (parity bar 1e-2): single-pass TF32 GEMMs instead of the error-compensated 3xTF32
    // single-pass TF32 (mma_smem.cuh FAST) is available but off: in bf16-plane mode it made the DSO step 9 % faster and
    // put the plane gradient at 1.4e-2 of its max-norm (activation kinks move), over the 1e-2 bar of that mode
    constexpr bool FAST = false;
    constexpr bool FASTG = false;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BwdSmem& s = *reinterpret_cast<BwdSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const enarf_scene_t& sc = a.sc;
    const int S = a.g.S, n = a.g.n, M = S - 1;
    const int total_tiles = sc.B * a.tiles_per_sample;
    const Maps m = make_maps(tid);
    int cur_b = -1;
    WGrad wg;
    wgrad_zero(wg);
    if (tid < 4) s.db3[tid] = 0.f;
    if (tid == 0) {
        mbar_init(&s.rec_bar[0], 1);
        mbar_init(&s.rec_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    uint32_t rec_phase = 0, rec_phase1 = 0;   // parities of the next completions of rec_bar[0 / 1] (uniform across the CTA)
    PHASE_DECL;

    for (;;) {
        __syncthreads();  // everyone is done with the previous tile (and with s.t.tile)
        PHASE(15);
        if (tid == 0) s.t.tile = atomicAdd(a.tile_counter, 1);
        __syncthreads();
        const int tile = s.t.tile;
        if (tile >= total_tiles) break;
        const int b = tile / a.tiles_per_sample;
        const int ray0 = (tile - b * a.tiles_per_sample) * a.rays_per_tile;
        const int rays = min(a.rays_per_tile, n - ray0);
        const int tp = rays * S;
        const int tb = (a.ck.rec && a.feat) ? a.ck.tile_base[tile] : -1;   // this tile's slice of the pair checkpoint
        if (tb == -2) continue;   // the forward found no (point, part) pair in this tile: nothing to differentiate
        if (b != cur_b) {
            if (cur_b >= 0) wgrad_flush(wg, m, s.db3, a, cur_b, tid);
            load_part_params(s.part, sc, b, tid, NT);
            load_decoder_smem(s.dec, sc, b, tid);
            cur_b = b;
        }
        if (tid < rays) {
            const size_t o = (size_t)b * n + ray0 + tid;
            const float* gc = a.g_color ? a.g_color + (size_t)b * 3 * n + ray0 + tid : nullptr;
            s.rg[tid * 8 + 0] = gc ? gc[0] : 0.f;
            s.rg[tid * 8 + 1] = gc ? gc[n] : 0.f;
            s.rg[tid * 8 + 2] = gc ? gc[2 * (size_t)n] : 0.f;
            // d mask also collects the composite's share: d/d mask of (1 - mask) * bg summed over the three channels
            s.rg[tid * 8 + 3] = (a.g_mask ? a.g_mask[o] : 0.f) - a.bg_const * (s.rg[tid * 8] + s.rg[tid * 8 + 1] + s.rg[tid * 8 + 2]);
            s.rg[tid * 8 + 4] = a.g_disp ? a.g_disp[o] : 0.f;
        }
        {
            const float4* src = reinterpret_cast<const float4*>(a.saved) + ((size_t)b * n + ray0) * S;
            for (int i = tid; i < tp; i += NT) s.pg[i] = src[i];
        }
        __syncthreads();

        // ---- 1. positions, masks, compaction (identical arithmetic to the forward) -----------------
        PHASE(0);   // tile load
        // With the forward's pair checkpoint the active list must be the forward's own (which evaluates the last sample of
        // a ray too); that sample then flows through with a zero upstream gradient.  Otherwise it is skipped outright.
        TileGeom geom = a.g;
        geom.skip_last = tb < 0 || a.ck.skip_last;      // (the forward's own list when it left the last samples out)
        if (tb >= 0 && a.ck.count) tile_from_counts(s.t, geom, a.ck.count, b, ray0, rays, tid);
        else tile_masks<TILE_FINE>(s.t, s.r, s.part, geom, sc.P, b, ray0, rays, tid);
        {
            // the forward's list: point order, or with early termination group-major without the samples it dropped
            const bool ert = a.ert_eps > 0.f;
            const int K = ert ? ERT_GROUP : S;
            tile_compact(s.t, rays, S, K, 0, (S + K - 1) / K, 0xffffffffu, ert ? s.pg : nullptr, tid);
        }
        const int n_active = s.t.n_active;
        PHASE(1);   // tile_points

        // ---- 3. per chunk: records -> gather -> decoder forward -> decoder backward -> scatter -------
        // Features checkpointed by the forward: 8 lanes per point, one 128-B row each.  A chunk's rows are loaded into
        // registers one chunk ahead (before the previous chunk's scatter) and stored to shared memory at its start.
        static_assert(CH == 2 * (NT / 8), "two checkpoint rows per 8-lane group");
        float4 fx[2];
        const int l8 = tid & 7;
        const float4* fi = reinterpret_cast<const float4*>(a.feat) + ((size_t)b * n + ray0) * S * (FEAT / 4);
        auto load_rows = [&](int first, int count) {
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int q = (tid >> 3) + i * (NT / 8);
                fx[i] = q < count ? __ldg(fi + (size_t)s.t.list[first + q] * (FEAT / 4) + l8) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        };
        auto store_rows = [&](int count) {   // fx -> X[channel][point], columns up to the next multiple of 32 zero-filled
            const int cend_ = (count + 31) & ~31;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int q = (tid >> 3) + i * (NT / 8);
                if (q < cend_) {
                    float* dst = s.X + (l8 * 4) * AS + q;
                    dst[0] = fx[i].x;
                    dst[AS] = fx[i].y;
                    dst[2 * AS] = fx[i].z;
                    dst[3 * AS] = fx[i].w;
                }
            }
        };
        const int pair_cap = tb >= 0 ? PB + PBX : PB;
        int c1 = n_active > 0 ? chunk_end(s.t, 0, pair_cap) : 0;
        // the first chunk's rows are requested now and stored after the compositing backward, which runs under the loads
        if (a.feat && n_active > 0) load_rows(0, c1);

        // ---- 2. compositing backward, one warp per ray, in place on pg ------------------------------
        float* pgf = reinterpret_cast<float*>(s.pg);
        for (int r = warp; r < rays; r += NWARP) {
            const float* dep = s.t.depth + r * (S + 1);
            float* go = pgf + (size_t)r * S * 4;
            const float gr = s.rg[r * 8], gg = s.rg[r * 8 + 1], gb = s.rg[r * 8 + 2], gm = s.rg[r * 8 + 3],
                        gd = s.rg[r * 8 + 4];
            const float* gfw = a.g_fw ? a.g_fw + ((size_t)b * n + ray0 + r) * M : nullptr;
            float carry = 0.f, U = 0.f;
            for (int base = 0; base < M; base += 32) {  // pass A: T, w, G and U = sum G*w
                const int j = base + lane;
                const bool ok = j < M;
                const float4 sv = ok ? *reinterpret_cast<const float4*>(go + j * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
                const float dd = ok ? fmaxf(sv.w, 0.f) * (dep[j + 1] - dep[j]) : 0.f;   // dropped sample: sigma < 0
                float incl = dd;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const float v = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += v;
                }
                const float Tn = expf(-(carry + incl));           // T_{j+1} = T_j * exp(-dd_j)
                const float w = expf(-(carry + (incl - dd))) * (1.f - expf(-dd));
                carry += __shfl_sync(0xffffffffu, incl, 31);
                float G = 0.f;
                if (ok) {
                    G = gr * sv.x + gg * sv.y + gb * sv.z + gm + gd / dep[j] + (gfw ? gfw[j] : 0.f);
                    go[j * 4 + 0] = w;
                    go[j * 4 + 1] = G;
                    go[j * 4 + 2] = Tn;
                }
                float u = ok ? G * w : 0.f;
#pragma unroll
                for (int o = 16; o; o >>= 1) u += __shfl_xor_sync(0xffffffffu, u, o);
                U += u;
            }
            float pre = 0.f;
            for (int base = 0; base < M; base += 32) {  // pass B: suffix sums -> d sigma, d rgb
                const int j = base + lane;
                const bool ok = j < M;
                const float w = ok ? go[j * 4 + 0] : 0.f, G = ok ? go[j * 4 + 1] : 0.f, Tn = ok ? go[j * 4 + 2] : 0.f;
                float incl = G * w;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const float v = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += v;
                }
                const float suffix = U - (pre + incl);          // sum_{k>j} G_k w_k
                pre += __shfl_sync(0xffffffffu, incl, 31);
                if (ok) {
                    const float gsig = (dep[j + 1] - dep[j]) * (G * Tn - suffix);
                    *reinterpret_cast<float4*>(go + j * 4) = make_float4(w * gr, w * gg, w * gb, gsig);
                }
            }
            if (lane == 0) *reinterpret_cast<float4*>(go + M * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        PHASE(2);   // composite backward
        if (a.feat && n_active > 0) store_rows(c1);
        __syncthreads();   // pg (the head's input) and X are complete
        for (int c0 = 0; c0 < n_active;) {
            const int cnt = c1 - c0;
            const int cend = (cnt + 31) & ~31;
            const bool tile_on = 32 * m.pb < cnt;   // this warp's 32 points intersect the chunk
            const int pair0 = s.t.pref[c0];
            if (tb >= 0) {   // records straight from the forward's checkpoint (5 x 128 bit each, coalesced)
                // one TMA bulk copy, issued by one thread and awaited only before the scatter at the end of the chunk (the
                // GEMM phases run under it); it replaced ~3 16-byte cp.async per thread and their shared-memory wavefronts
                if (tid == 0)
                    bulk_load(s.r.rec, a.ck.rec + (size_t)(tb + pair0) * REC,
                              (uint32_t)(min(PB, s.t.pref[c1] - pair0) * REC * sizeof(float)), &s.rec_bar[0]);
            } else {
                setup_pairs<false>(s.t, s.r.rec, s.part, sc, b, c0, c1, tid);
            }
            if (!a.feat) {   // no feature checkpoint: gather again (needs the records)
                __syncthreads();
                gather_chunk<T, true>(s.t, s.r.rec, s.X, sc, b, c0, c1, tid);
                __syncthreads();
            }
            // with the checkpoint, X of this chunk was stored during the previous chunk's scatter (or before the loop)
            PHASE(3);   // records + X load
            if (tile_on) decoder_layer<FEAT, FAST>(s.X, s.dec.w1, W1S, s.dec.b1, s.A1, m);
            __syncthreads();
            PHASE(4);   // L1
            // Head backward from the forward's own outputs (four lanes per point, lane c owns output c): the saved
            // y = tanh(lrelu(o)) gives tanh' = 1 - y^2 and the sign of o (LeakyReLU and tanh keep it); the saved
            // sigma = 10 relu(lrelu(o)) is positive exactly when o is.  The head is therefore not evaluated again (it was
            // 6 % of this kernel: 64 x 4 multiply-adds per point through shared memory and two shuffles).  The value is
            // requested here, ahead of layer 2, and used after it.
            float y_saved = 0.f;
            if (8 * warp < cend && (tid >> 2) < cnt)
                y_saved = __ldg(a.saved + (((size_t)b * n + ray0) * S + s.t.list[c0 + (tid >> 2)]) * 4 + (tid & 3));
            if (tile_on) decoder_layer<HID, FAST>(s.A1, s.dec.w2, W2S, s.dec.b2, s.A2, m);
            __syncthreads();
            PHASE(5);   // L2
            if (8 * warp < cend) {
                const int pt = tid >> 2, c = tid & 3;
                float g3 = 0.f;
                if (pt < cnt) {
                    const float4 gin = s.pg[s.t.list[c0 + pt]];
                    const float ag = (y_saved > 0.f ? 1.f : 0.2f) * 1.4142135623730951f;    // act_grad(o): sign(o) = sign(y)
                    if (c < 3) {   // rgb = tanh(lrelu(o3))
                        const float gc = c == 0 ? gin.x : c == 1 ? gin.y : gin.z;
                        g3 = gc * (1.f - y_saved * y_saved) * ag;
                    } else {       // sigma = 10 * MyReLU(lrelu(o3[3]))
                        const float go = 10.f * gin.w;
                        // MyReLU backward (NeRF/activation.py:12-16): g*[x>=0] + 0.1*g*[x<0]*[g<0]; x > 0 <=> sigma > 0
                        const float graw = (y_saved > 0.f) ? go : (go < 0.f ? 0.1f * go : 0.f);
                        g3 = graw * ag;
                    }
                }
                s.G3[c * AS + pt] = g3;
                float v = g3;   // db3[c]: sum over the warp's 8 points (lanes with the same c)
                v += __shfl_xor_sync(0xffffffffu, v, 4);
                v += __shfl_xor_sync(0xffffffffu, v, 8);
                v += __shfl_xor_sync(0xffffffffu, v, 16);
                if (lane < 4) atomicAdd(&s.db3[c], v);
            }
            __syncthreads();
            PHASE(6);   // head
            if (tile_on) {   // G2 = (W3^T g3) * act'(a2), in place over A2; dW3 partial sums
                float4 g3[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) g3[c] = *reinterpret_cast<const float4*>(s.G3 + c * AS + 4 * m.pq);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int o = m.og + 16 * j;
                    float* ap = s.A2 + o * AS + 4 * m.pq;
                    const float4 a2 = *reinterpret_cast<const float4*>(ap);
                    float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const float w = s.dec.w3[c * W2S + o];
                        d.x = fmaf(w, g3[c].x, d.x);
                        d.y = fmaf(w, g3[c].y, d.y);
                        d.z = fmaf(w, g3[c].z, d.z);
                        d.w = fmaf(w, g3[c].w, d.w);
                        wg.w3[c][j] += (g3[c].x * a2.x + g3[c].y * a2.y) + (g3[c].z * a2.z + g3[c].w * a2.w);
                    }
                    d = make_float4(d.x * act_grad(a2.x), d.y * act_grad(a2.y), d.z * act_grad(a2.z), d.w * act_grad(a2.w));
                    *reinterpret_cast<float4*>(ap) = d;
                    wg.b2[j] += (d.x + d.y) + (d.z + d.w);
                }
            }
            __syncthreads();
            PHASE(7);   // G2 pass
            // dW2 += G2^T A1 over the chunk's points: warp tile rows (o) 16 wr.., cols (i) 8 TN2 wc..
            wgrad_mma<TN2, FASTG>(s.A2, s.A1, 16 * m.wr, 8 * TN2 * m.wc, cend, m.g, m.t, wg.w2);
            __syncthreads();
            PHASE(8);   // dW2
            if (tile_on) {   // G1 = (G2 W2) * act'(a1), in place over A1; db1 partial sums
                float acc[2][2][4];
                gemm_pts_mma<HID, 2, true, FASTG>(s.A2, s.dec.w2, W2S, 32 * m.pb, 16 * m.ob, m.g, m.t, acc);
#pragma unroll
                for (int nt = 0; nt < 2; ++nt) {
                    const int col = 16 * m.ob + 8 * nt + 2 * m.t;
#pragma unroll
                    for (int mt = 0; mt < 2; ++mt) {
                        float* d0 = s.A1 + col * AS + 32 * m.pb + 16 * mt + m.g;
                        const float g0 = acc[mt][nt][0] * act_grad(d0[0]), g1 = acc[mt][nt][1] * act_grad(d0[AS]);
                        const float g2 = acc[mt][nt][2] * act_grad(d0[8]), g3 = acc[mt][nt][3] * act_grad(d0[AS + 8]);
                        d0[0] = g0;
                        d0[AS] = g1;
                        d0[8] = g2;
                        d0[AS + 8] = g3;
                        wg.b1[nt][0] += g0 + g2;
                        wg.b1[nt][1] += g1 + g3;
                    }
                }
            }
            __syncthreads();
            PHASE(9);   // G1
            // dW1 += G1^T X: warp tile rows (o) 16 wr.., cols (c) 8 TN1 wc..
            wgrad_mma<TN1, FASTG>(s.A1, s.X, 16 * m.wr, 8 * TN1 * m.wc, cend, m.g, m.t, wg.w1);
            // no barrier: dX goes to the rows of A2 (G2 is dead after the G1 phase), so X and G1 are only read here
            if (tile_on) {   // dX = G1 W1 -> the G2 buffer as [point][DXS]: warp tile points 32 pb.., channels 8 ob..
                float acc[2][1][4];
                gemm_pts_mma<HID, 1, true, FASTG>(s.A1, s.dec.w1, W1S, 32 * m.pb, 8 * m.ob, m.g, m.t, acc);
                const int col = 8 * m.ob + 2 * m.t;
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) {
                    float* d0 = s.A2 + (32 * m.pb + 16 * mt + m.g) * DXS + col;
                    *reinterpret_cast<float2*>(d0) = make_float2(acc[mt][0][0], acc[mt][0][1]);
                    *reinterpret_cast<float2*>(d0 + 8 * DXS) = make_float2(acc[mt][0][2], acc[mt][0][3]);
                }
            }
            __syncthreads();
            const bool overflow = tb >= 0 && s.t.pref[c1] - pair0 > PB;
            if (overflow && tid == 0)   // G1 is dead: its buffer takes the records beyond PB, awaited inside the scatter
                bulk_load(s.A1, a.ck.rec + (size_t)(tb + pair0 + PB) * REC,
                          (uint32_t)((s.t.pref[c1] - pair0 - PB) * REC * sizeof(float)), &s.rec_bar[1]);
            if (tb >= 0) {   // this chunk's records have landed (every thread observes the barrier phase itself)
                mbar_wait(&s.rec_bar[0], rec_phase);
                rec_phase ^= 1u;
            }
            PHASE(11);  // dX
            const int c1n = c1 < n_active ? chunk_end(s.t, c1, pair_cap) : c1;
            if (a.feat && c1 < n_active) load_rows(c1, c1n - c1);   // next chunk's rows, in flight during the scatter
            if (tb >= 0)
                scatter_chunk<T, true>(s.t, s.r.rec, s.A1, &s.rec_bar[1], rec_phase1, s.A2, a, b, c0, c1, tid,
                                       a.ck.feat + (size_t)(tb + pair0) * FEAT);
            else
                scatter_chunk<T, false>(s.t, s.r.rec, nullptr, nullptr, 0u, s.A2, a, b, c0, c1, tid, nullptr);
            if (overflow) rec_phase1 ^= 1u;
            if (a.feat && c1 < n_active) store_rows(c1n - c1);   // X is free since the dW1 / dX phase
            __syncthreads();
            PHASE(12);  // scatter
            c0 = c1;
            c1 = c1n;
        }
    }
    if (cur_b >= 0) wgrad_flush(wg, m, s.db3, a, cur_b, tid);
    PHASE(13);
    PHASE_FLUSH;
}

extern "C" int enarf_render_fine_bwd(const enarf_scene_t* scene, const float* depth_min, const float* depth_max,
                                     const float* ray_dir, const float* bins, const float* point_saved,
                                     const float* feat_saved, const enarf_pair_ckpt_t* pair_ckpt, int n, int Nf,
                                     const float* g_color, const float* g_mask, const float* g_disparity,
                                     const float* g_fine_weights, float* d_feat, float* d_prob,
                                     int64_t d_prob_batch_stride, float* d_w1, float* d_w2, float* d_w3, float* d_b1,
                                     float* d_b2, float* d_b3, float ert_eps, float bg_const, void* stream) {
    int rc = enarf_check_scene(scene, "enarf_render_fine_bwd");
    if (rc) return rc;
    ENARF_REQUIRE(depth_min && depth_max && ray_dir && bins && point_saved, ENARF_ERR_INVALID_ARG,
                  "enarf_render_fine_bwd: NULL forward argument");
    ENARF_REQUIRE(d_feat && d_prob && d_w1 && d_w2 && d_w3 && d_b1 && d_b2 && d_b3, ENARF_ERR_INVALID_ARG,
                  "enarf_render_fine_bwd: NULL gradient buffer");
    ENARF_REQUIRE(((uintptr_t)d_feat & 15) == 0 && ((uintptr_t)point_saved & 15) == 0 && ((uintptr_t)feat_saved & 15) == 0,
                  ENARF_ERR_INVALID_ARG, "enarf_render_fine_bwd: d_feat / point_saved / feat_saved must be 16-byte aligned");
    ENARF_REQUIRE(n > 0 && Nf >= 2, ENARF_ERR_INVALID_ARG, "enarf_render_fine_bwd: bad sizes n=%d Nf=%d", n, Nf);
    ENARF_REQUIRE(Nf <= TP_MAX, ENARF_ERR_UNSUPPORTED, "enarf_render_fine_bwd: Nf=%d exceeds %d", Nf, TP_MAX);
    BwdArgs a = {};
    a.sc = *scene;
    a.g.dmin = depth_min;
    a.g.dmax = depth_max;
    a.g.dir = ray_dir;
    a.g.bins = bins;
    a.g.n = n;
    a.g.S = Nf;
    a.g.skip_last = 1;
    a.saved = point_saved;
    a.feat = feat_saved;
    if (pair_ckpt) {
        ENARF_REQUIRE(feat_saved && pair_ckpt->rec && pair_ckpt->feat && pair_ckpt->tile_base &&
                          ((uintptr_t)pair_ckpt->rec & 15) == 0 && ((uintptr_t)pair_ckpt->feat & 15) == 0,
                      ENARF_ERR_INVALID_ARG, "enarf_render_fine_bwd: the pair checkpoint needs feat_saved and 16-byte aligned buffers");
        a.ck = *pair_ckpt;
    }
    a.g_color = g_color;
    a.g_mask = g_mask;
    a.g_disp = g_disparity;
    a.g_fw = g_fine_weights;
    a.d_feat = d_feat;
    a.d_prob = d_prob;
    a.d_prob_stride = d_prob_batch_stride;
    a.d_w1 = d_w1;
    a.d_w2 = d_w2;
    a.d_w3 = d_w3;
    a.d_b1 = d_b1;
    a.d_b2 = d_b2;
    a.d_b3 = d_b3;
    a.ert_eps = ert_eps;
    a.bg_const = bg_const;
    a.rays_per_tile = max(1, min(MAX_RAYS, TP_MAX / Nf));
    a.tiles_per_sample = (n + a.rays_per_tile - 1) / a.rays_per_tile;
    const int total = scene->B * a.tiles_per_sample;
    const int grid = min(total, 2 * enarf_sm_count());
    const size_t smem = sizeof(BwdSmem);
    static_assert(sizeof(BwdSmem) <= 113 * 1024, "two CTAs per SM");
    a.tile_counter = enarf_tile_counter((cudaStream_t)stream);
    if (!a.tile_counter) return ENARF_ERR_CUDA;
    cudaError_t e;
    if (scene->feat_dtype == ENARF_DTYPE_F32) {
        auto k = render_bwd_kernel<float>;
        e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) k<<<grid, NT, smem, (cudaStream_t)stream>>>(a);
    } else {
        auto k = render_bwd_kernel<__nv_bfloat16>;
        e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) k<<<grid, NT, smem, (cudaStream_t)stream>>>(a);
    }
    if (e != cudaSuccess) {
        enarf_set_error("enarf_render_fine_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return ENARF_ERR_CUDA;
    }
    ENARF_CHECK_LAUNCH("enarf_render_fine_bwd");
    return ENARF_OK;
}
