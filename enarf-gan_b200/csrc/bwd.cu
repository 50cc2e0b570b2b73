// Fused backward of the fine pass: compositing backward -> decoder backward (dgrad + wgrad) -> tri-plane
// scatter.  Gradients flow to the feature planes, the part-probability planes and the per-sample decoder
// weights / biases only (pose, rays and bins are non-differentiable: libraries/NeRF/rendering.py:216-217).
// Reference being replaced: autograd through libraries/NeRF/rendering.py:307-335, models/narf.py:213-275,
// libraries/triplane/sampling.py:54-127 (grid_sample backward + scatter_add backward), NeRF/activation.py:12-16.
//
// Same tiling as the forward (render.cu): one CTA owns whole rays.  Per tile
//   1. positions + part masks are regenerated with the forward's contract arithmetic (bit-identical masks);
//   2. one warp per ray turns the upstream gradients and the forward's saved (rgb, sigma) into per-sample
//      d/d rgb and d/d sigma (two shuffle scans: transmittance prefix, weighted-gradient suffix);
//   3. per chunk of 128 active points: gather -> decoder forward (activations kept in smem columns) ->
//      decoder backward; the three weight-gradient GEMMs (G^T A over the chunk's points) are accumulated in
//      registers across the CTA's whole lifetime and flushed with one atomicAdd per entry per sample;
//   4. d(feature) is scattered with 128-bit vector reductions (red.global.add.v4.f32) into the channels-last
//      gradient planes, d(part weight) goes through the sigmoid product into scalar reductions on the
//      probability planes.
#include "common.cuh"
#include "gather.cuh"

#define TP_MAX 512
#define NT 128
#define CHUNK 128
#define CS 129          // activation column stride
#define MAX_RAYS 16

struct BwdArgs {
    enarf_scene_t sc;
    const float* dmin;
    const float* dmax;
    const float* dir;
    const float* bins;
    const float* saved;       // (B,n,S,4)
    const float* g_color;     // (B,3,n) | NULL
    const float* g_mask;      // (B,n)   | NULL
    const float* g_disp;      // (B,n)   | NULL
    const float* g_fw;        // (B,n,S-1) | NULL
    float* d_feat;
    float* d_prob;
    long long d_prob_stride;
    float *d_w1, *d_w2, *d_w3, *d_b1, *d_b2, *d_b3;
    int n, S, rays_per_tile, tiles_per_sample;
    int* tile_counter;
};

struct BSmem {
    float w1[HID * FEAT];    // [o][i]
    float w2[HID * HID];
    float w3[4 * HID];
    float b1[HID], b2[HID], b3[4];
    float part[ENARF_MAX_PARTS * PART_STRIDE];
    float X[FEAT * CS];      // features, later d(features)
    float A1[HID * CS];      // layer-1 activations, later their pre-activation gradients
    float A2[HID * CS];
    float G3[4 * CS];
    float pos[TP_MAX * 3];
    float sav[TP_MAX * 4];   // forward (r,g,b,sigma)
    float g[TP_MAX * 4];     // d/d(r,g,b,sigma)
    float depth[TP_MAX + MAX_RAYS];
    float ray[MAX_RAYS * 5];
    float rg[MAX_RAYS * 5];  // upstream d/d(color rgb, mask, disparity) per ray
    uint32_t pmask[TP_MAX];
    uint16_t list[TP_MAX];
    int n_active;
};

__device__ __forceinline__ float act_grad(float a) { return (a > 0.f ? 1.f : 0.2f) * 1.4142135623730951f; }

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// per-thread persistent weight-gradient tiles
struct WGrad {
    float w1[16];  // rows 4*ot..+3, cols 4*it..+3
    float w2[32];  // rows 4*ot..+3, cols 8*it..+7
    float w3[2];   // row tid/32, cols 2*(tid%32)..+1
    float b1, b2, b3;
};

__device__ __forceinline__ void wgrad_zero(WGrad& g) {
#pragma unroll
    for (int i = 0; i < 16; ++i) g.w1[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) g.w2[i] = 0.f;
    g.w3[0] = g.w3[1] = 0.f;
    g.b1 = g.b2 = g.b3 = 0.f;
}

__device__ __forceinline__ void wgrad_flush(WGrad& g, const BwdArgs& a, int b, int tid) {
    const int ot = tid >> 3, it = tid & 7;
    float* d1 = a.d_w1 + (size_t)b * HID * FEAT;
    float* d2 = a.d_w2 + (size_t)b * HID * HID;
    float* d3 = a.d_w3 + (size_t)b * 4 * HID;
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) atomicAdd(d1 + (4 * ot + r) * FEAT + 4 * it + c, g.w1[r * 4 + c]);
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) atomicAdd(d2 + (4 * ot + r) * HID + 8 * it + c, g.w2[r * 8 + c]);
    atomicAdd(d3 + (tid >> 5) * HID + 2 * (tid & 31), g.w3[0]);
    atomicAdd(d3 + (tid >> 5) * HID + 2 * (tid & 31) + 1, g.w3[1]);
    if (tid < HID) {
        atomicAdd(a.d_b1 + tid, g.b1);
        atomicAdd(a.d_b2 + tid, g.b2);
    }
    if (tid < 4) atomicAdd(a.d_b3 + tid, g.b3);
    wgrad_zero(g);
}

// dot(W[row][0..K), v[0..K)) + bias with the weight row broadcast from shared memory
template <int K>
__device__ __forceinline__ float row_dot(const float* __restrict__ wrow, const float (&v)[K], float bias) {
    float a0 = bias, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int i4 = 0; i4 < K / 4; ++i4) {
        const float4 w = *reinterpret_cast<const float4*>(wrow + 4 * i4);
        a0 = fmaf(w.x, v[4 * i4 + 0], a0);
        a1 = fmaf(w.y, v[4 * i4 + 1], a1);
        a2 = fmaf(w.z, v[4 * i4 + 2], a2);
        a3 = fmaf(w.w, v[4 * i4 + 3], a3);
    }
    return (a0 + a1) + (a2 + a3);
}

// Decoder forward for the point in column `col` (activations left in s.A1 / s.A2), then the head's backward:
// writes d/d(o3) into s.G3 and returns d/d(pre-activation of layer 2) in g2[].
__device__ __forceinline__ void decode_fwd_and_head_bwd(BSmem& s, int col, const float4 gin, float (&g2)[HID]) {
    {
        float x[FEAT];
#pragma unroll
        for (int i = 0; i < FEAT; ++i) x[i] = s.X[i * CS + col];
#pragma unroll 2
        for (int o = 0; o < HID; ++o) s.A1[o * CS + col] = lrelu_s2(row_dot<FEAT>(s.w1 + o * FEAT, x, s.b1[o]));
    }
    float o3[4];
    {
        float a1[HID];
#pragma unroll
        for (int i = 0; i < HID; ++i) a1[i] = s.A1[i * CS + col];
#pragma unroll 2
        for (int o = 0; o < HID; ++o) s.A2[o * CS + col] = lrelu_s2(row_dot<HID>(s.w2 + o * HID, a1, s.b2[o]));
    }
    float a2[HID];
#pragma unroll
    for (int i = 0; i < HID; ++i) a2[i] = s.A2[i * CS + col];
#pragma unroll
    for (int c = 0; c < 4; ++c) o3[c] = row_dot<HID>(s.w3 + c * HID, a2, s.b3[c]);
    // head: rgb = tanh(lrelu(o3[0:3])), sigma = 10 * MyReLU(lrelu(o3[3]))
    float g3[4];
    const float gc[3] = {gin.x, gin.y, gin.z};
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float y = tanhf(lrelu_s2(o3[c]));
        g3[c] = gc[c] * (1.f - y * y) * act_grad(o3[c]);
    }
    const float raw = lrelu_s2(o3[3]);
    const float go = 10.f * gin.w;
    // MyReLU backward (NeRF/activation.py:12-16): g*[x>=0] + 0.1*g*[x<0]*[g<0]
    const float graw = (raw >= 0.f) ? go : (go < 0.f ? 0.1f * go : 0.f);
    g3[3] = graw * act_grad(o3[3]);
#pragma unroll
    for (int c = 0; c < 4; ++c) s.G3[c * CS + col] = g3[c];
#pragma unroll
    for (int o = 0; o < HID; ++o) {
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) d = fmaf(s.w3[c * HID + o], g3[c], d);
        g2[o] = d * act_grad(a2[o]);
    }
}

// out[i] = sum_o W[o][i] * G[o][col]   (W row-major [o][K]); accumulators statically indexed
template <int K>
__device__ __forceinline__ void wt_times_g(const float* __restrict__ w, const float* __restrict__ G, int col,
                                           float (&out)[K]) {
#pragma unroll
    for (int i = 0; i < K; ++i) out[i] = 0.f;
#pragma unroll 2
    for (int o = 0; o < HID; ++o) {
        const float g = G[o * CS + col];
        const float4* row = reinterpret_cast<const float4*>(w + o * K);
#pragma unroll
        for (int i4 = 0; i4 < K / 4; ++i4) {
            const float4 v = row[i4];
            out[4 * i4 + 0] = fmaf(v.x, g, out[4 * i4 + 0]);
            out[4 * i4 + 1] = fmaf(v.y, g, out[4 * i4 + 1]);
            out[4 * i4 + 2] = fmaf(v.z, g, out[4 * i4 + 2]);
            out[4 * i4 + 3] = fmaf(v.w, g, out[4 * i4 + 3]);
        }
    }
}

// Scatter for one point by an 8-lane group; lane l owns channels 4l..4l+3 of dx.
template <typename T>
__device__ __forceinline__ void scatter_point(const BwdArgs& a, const float* s_part, int b, uint32_t m, float x,
                                              float y, float z, int lane8, float4 dx, unsigned group_mask) {
    const enarf_scene_t& sc = a.sc;
    const int H = sc.H, W = sc.W;
    const T* feat = reinterpret_cast<const T*>(sc.feat_planes) + (size_t)b * 3 * H * W * FEAT;
    float* dfeat = a.d_feat + (size_t)b * 3 * H * W * FEAT;
    const float* prob = sc.prob_planes + (size_t)b * sc.prob_batch_stride;
    float* dprob = a.d_prob + (size_t)b * a.d_prob_stride;
    while (m) {
        const int p = __ffs(m) - 1;
        m &= m - 1;
        float c[3];
        part_canonical(s_part + p * PART_STRIDE, x, y, z, c[0], c[1], c[2]);
        Bilin bl[3];
        float sg[3], wp = 1.f;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            bl[k] = bilin_setup(c[k], c[(k + 1) % 3], W, H);
            sg[k] = sigmoidf_(bilin_fetch1(prob + (size_t)(3 * p + k) * H * W, bl[k], W));
            wp *= sg[k];
        }
        float fdot = 0.f;  // <dx, f_p> restricted to this lane's channels
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const long long base = ((long long)k * H * W + (long long)bl[k].y0 * W + bl[k].x0) * FEAT + lane8 * 4;
            const float bw[4] = {bl[k].w00, bl[k].w01, bl[k].w10, bl[k].w11};
            const long long off[4] = {0, FEAT, (long long)W * FEAT, (long long)(W + 1) * FEAT};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (bw[q] != 0.f) {
                    const float4 v = load_feat4<T>(feat, base + off[q]);
                    fdot = fmaf(bw[q], v.x * dx.x + v.y * dx.y + v.z * dx.z + v.w * dx.w, fdot);
                    const float sc_ = bw[q] * wp;
                    red_add_v4(dfeat + base + off[q], sc_ * dx.x, sc_ * dx.y, sc_ * dx.z, sc_ * dx.w);
                }
            }
        }
        fdot += __shfl_xor_sync(group_mask, fdot, 1);
        fdot += __shfl_xor_sync(group_mask, fdot, 2);
        fdot += __shfl_xor_sync(group_mask, fdot, 4);
        // d w_p = fdot;  w_p = prod sigmoid(v_k)  ->  d v_k = fdot * w_p * (1 - sigmoid(v_k))
        for (int e = lane8; e < 12; e += 8) {
            const int k = e >> 2, q = e & 3;
            const float bwq = (q == 0) ? bl[k].w00 : (q == 1) ? bl[k].w01 : (q == 2) ? bl[k].w10 : bl[k].w11;
            if (bwq != 0.f) {
                const float dv = fdot * wp * (1.f - sg[k]);
                float* dst = dprob + (size_t)(3 * p + k) * H * W + (ptrdiff_t)(bl[k].y0 + (q >> 1)) * W + bl[k].x0 + (q & 1);
                atomicAdd(dst, bwq * dv);
            }
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(NT, 1) render_bwd_kernel(const BwdArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BSmem& s = *reinterpret_cast<BSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const enarf_scene_t& sc = a.sc;
    const int S = a.S, n = a.n, M = a.S - 1;
    const int total_tiles = sc.B * a.tiles_per_sample;
    int cur_b = -1;
    WGrad wg;
    wgrad_zero(wg);

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
        __syncthreads();
        if (b != cur_b) {
            if (cur_b >= 0) wgrad_flush(wg, a, cur_b, tid);
            load_part_params(s.part, sc, b, tid, NT);
            for (int e = tid; e < HID * FEAT; e += NT) s.w1[e] = sc.w1[(size_t)b * HID * FEAT + e];
            for (int e = tid; e < HID * HID; e += NT) s.w2[e] = sc.w2[(size_t)b * HID * HID + e];
            for (int e = tid; e < 4 * HID; e += NT) s.w3[e] = sc.w3[(size_t)b * 4 * HID + e];
            if (tid < HID) {
                s.b1[tid] = sc.b1[tid];
                s.b2[tid] = sc.b2[tid];
            }
            if (tid < 4) s.b3[tid] = sc.b3[tid];
            cur_b = b;
        }
        if (tid == 0) s.n_active = 0;
        if (tid < rays) {
            const size_t o = (size_t)b * n + ray0 + tid;
            const float* d = a.dir + (size_t)b * 3 * n + ray0 + tid;
            s.ray[tid * 5 + 0] = a.dmin[o];
            s.ray[tid * 5 + 1] = a.dmax[o];
            s.ray[tid * 5 + 2] = d[0];
            s.ray[tid * 5 + 3] = d[n];
            s.ray[tid * 5 + 4] = d[2 * (size_t)n];
            const float* gc = a.g_color ? a.g_color + (size_t)b * 3 * n + ray0 + tid : nullptr;
            s.rg[tid * 5 + 0] = gc ? gc[0] : 0.f;
            s.rg[tid * 5 + 1] = gc ? gc[n] : 0.f;
            s.rg[tid * 5 + 2] = gc ? gc[2 * (size_t)n] : 0.f;
            s.rg[tid * 5 + 3] = a.g_mask ? a.g_mask[o] : 0.f;
            s.rg[tid * 5 + 4] = a.g_disp ? a.g_disp[o] : 0.f;
        }
        {
            const float4* src = reinterpret_cast<const float4*>(a.saved) + ((size_t)b * n + ray0) * S;
            for (int t = tid; t < tp; t += NT) *reinterpret_cast<float4*>(s.sav + t * 4) = src[t];
        }
        __syncthreads();

        // ---- 1. positions, masks, compaction (identical arithmetic to the forward) -----------------
        for (int t = tid; t < ((tp + 31) & ~31); t += NT) {
            uint32_t m = 0;
            if (t < tp) {
                const int r = t / S, j = t - r * S;
                const float* ry = s.ray + r * 5;
                const float dmin = ry[0], dmax = ry[1];
                const float u = a.bins[((size_t)b * n + ray0 + r) * S + j];
                const float x = clerp(cmul(dmin, ry[2]), cmul(dmax, ry[2]), u);
                const float y = clerp(cmul(dmin, ry[3]), cmul(dmax, ry[3]), u);
                const float z = clerp(cmul(dmin, ry[4]), cmul(dmax, ry[4]), u);
                s.depth[r * (S + 1) + j] = clerp(dmin, dmax, u);
                if (j < M) {  // the last sample only closes the last interval: no gradient reaches it
                    for (int p = 0; p < sc.P; ++p) {
                        float c0, c1, c2;
                        if (part_canonical(s.part + p * PART_STRIDE, x, y, z, c0, c1, c2)) m |= 1u << p;
                    }
                }
                s.pos[t * 3 + 0] = x;
                s.pos[t * 3 + 1] = y;
                s.pos[t * 3 + 2] = z;
                s.pmask[t] = m;
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

        // ---- 2. compositing backward, one warp per ray ----------------------------------------------
        for (int r = warp; r < rays; r += NT / 32) {
            const float* dep = s.depth + r * (S + 1);
            const float* sv = s.sav + (size_t)r * S * 4;
            float* go = s.g + (size_t)r * S * 4;
            const float gr = s.rg[r * 5], gg = s.rg[r * 5 + 1], gb = s.rg[r * 5 + 2], gm = s.rg[r * 5 + 3],
                        gd = s.rg[r * 5 + 4];
            const float* gfw = a.g_fw ? a.g_fw + ((size_t)b * n + ray0 + r) * M : nullptr;
            float carry = 0.f, U = 0.f;
            for (int base = 0; base < M; base += 32) {  // pass A: T, w, G and U = sum G*w
                const int j = base + lane;
                const bool ok = j < M;
                const float dd = ok ? sv[j * 4 + 3] * (dep[j + 1] - dep[j]) : 0.f;
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
                    G = gr * sv[j * 4] + gg * sv[j * 4 + 1] + gb * sv[j * 4 + 2] + gm + gd / dep[j] + (gfw ? gfw[j] : 0.f);
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
        __syncthreads();

        // ---- 3. decoder forward/backward + scatter, CHUNK active points at a time -------------------
        const int ot = tid >> 3, it = tid & 7, lane8 = tid & 7;
        const unsigned gmask = 0xffu << (lane & 24);
        for (int c0 = 0; c0 < n_active; c0 += CHUNK) {
            const int cnt = min(CHUNK, n_active - c0);
            for (int q = tid >> 3; q < cnt; q += NT / 8) {
                const int t = s.list[c0 + q];
                const float4 f = gather_point<T>(sc, s.part, b, s.pmask[t], s.pos[t * 3], s.pos[t * 3 + 1],
                                                 s.pos[t * 3 + 2], lane8);
                float* dst = s.X + (lane8 * 4) * CS + q;
                dst[0] = f.x;
                dst[CS] = f.y;
                dst[2 * CS] = f.z;
                dst[3 * CS] = f.w;
            }
            __syncthreads();
            float g2[HID];
            if (tid < cnt) {
                const int t = s.list[c0 + tid];
                decode_fwd_and_head_bwd(s, tid, *reinterpret_cast<const float4*>(s.g + t * 4), g2);
            }
            __syncthreads();
            {   // dW3 += G3^T A2, db3
                const int c = tid >> 5, o0 = 2 * (tid & 31);
                float x0 = 0.f, x1 = 0.f, bb = 0.f;
                for (int pt = 0; pt < cnt; ++pt) {
                    const float g = s.G3[c * CS + pt];
                    x0 = fmaf(g, s.A2[o0 * CS + pt], x0);
                    x1 = fmaf(g, s.A2[(o0 + 1) * CS + pt], x1);
                    if (tid < 4) bb += s.G3[tid * CS + pt];
                }
                wg.w3[0] += x0;
                wg.w3[1] += x1;
                wg.b3 += bb;
            }
            __syncthreads();
            if (tid < cnt) {
#pragma unroll
                for (int o = 0; o < HID; ++o) s.A2[o * CS + tid] = g2[o];
            }
            __syncthreads();
            {   // dW2 += G2^T A1, db2     (G2 now lives in A2)
                float acc[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] = 0.f;
                float bb = 0.f;
                for (int pt = 0; pt < cnt; ++pt) {
                    float g[4], v[8];
#pragma unroll
                    for (int r = 0; r < 4; ++r) g[r] = s.A2[(4 * ot + r) * CS + pt];
#pragma unroll
                    for (int c = 0; c < 8; ++c) v[c] = s.A1[(8 * it + c) * CS + pt];
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int c = 0; c < 8; ++c) acc[r * 8 + c] = fmaf(g[r], v[c], acc[r * 8 + c]);
                    if (tid < HID) bb += s.A2[tid * CS + pt];
                }
#pragma unroll
                for (int i = 0; i < 32; ++i) wg.w2[i] += acc[i];
                wg.b2 += bb;
            }
            __syncthreads();
            if (tid < cnt) {  // g1 = (W2^T g2) * act'(a1)  -> overwrites the own A1 column
                float da1[HID];
                wt_times_g<HID>(s.w2, s.A2, tid, da1);
#pragma unroll
                for (int i = 0; i < HID; ++i) s.A1[i * CS + tid] = da1[i] * act_grad(s.A1[i * CS + tid]);
            }
            __syncthreads();
            {   // dW1 += G1^T X, db1
                float acc[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[i] = 0.f;
                float bb = 0.f;
                for (int pt = 0; pt < cnt; ++pt) {
                    float g[4], v[4];
#pragma unroll
                    for (int r = 0; r < 4; ++r) g[r] = s.A1[(4 * ot + r) * CS + pt];
#pragma unroll
                    for (int c = 0; c < 4; ++c) v[c] = s.X[(4 * it + c) * CS + pt];
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int c = 0; c < 4; ++c) acc[r * 4 + c] = fmaf(g[r], v[c], acc[r * 4 + c]);
                    if (tid < HID) bb += s.A1[tid * CS + pt];
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) wg.w1[i] += acc[i];
                wg.b1 += bb;
            }
            __syncthreads();
            if (tid < cnt) {  // dx = W1^T g1 -> overwrites the own X column
                float dx[FEAT];
                wt_times_g<FEAT>(s.w1, s.A1, tid, dx);
#pragma unroll
                for (int i = 0; i < FEAT; ++i) s.X[i * CS + tid] = dx[i];
            }
            __syncthreads();
            for (int q = tid >> 3; q < cnt; q += NT / 8) {
                const int t = s.list[c0 + q];
                const float* src = s.X + (lane8 * 4) * CS + q;
                const float4 dx = make_float4(src[0], src[CS], src[2 * CS], src[3 * CS]);
                scatter_point<T>(a, s.part, b, s.pmask[t], s.pos[t * 3], s.pos[t * 3 + 1], s.pos[t * 3 + 2], lane8, dx,
                                 gmask);
            }
            __syncthreads();
        }
    }
    if (cur_b >= 0) wgrad_flush(wg, a, cur_b, tid);
}

extern "C" int enarf_render_fine_bwd(const enarf_scene_t* scene, const float* depth_min, const float* depth_max,
                                     const float* ray_dir, const float* bins, const float* point_saved, int n, int Nf,
                                     const float* g_color, const float* g_mask, const float* g_disparity,
                                     const float* g_fine_weights, float* d_feat, float* d_prob,
                                     int64_t d_prob_batch_stride, float* d_w1, float* d_w2, float* d_w3, float* d_b1,
                                     float* d_b2, float* d_b3, void* stream) {
    int rc = enarf_check_scene(scene, "enarf_render_fine_bwd");
    if (rc) return rc;
    ENARF_REQUIRE(depth_min && depth_max && ray_dir && bins && point_saved, ENARF_ERR_INVALID_ARG,
                  "enarf_render_fine_bwd: NULL forward argument");
    ENARF_REQUIRE(d_feat && d_prob && d_w1 && d_w2 && d_w3 && d_b1 && d_b2 && d_b3, ENARF_ERR_INVALID_ARG,
                  "enarf_render_fine_bwd: NULL gradient buffer");
    ENARF_REQUIRE(((uintptr_t)d_feat & 15) == 0 && ((uintptr_t)point_saved & 15) == 0, ENARF_ERR_INVALID_ARG,
                  "enarf_render_fine_bwd: d_feat / point_saved must be 16-byte aligned");
    ENARF_REQUIRE(n > 0 && Nf >= 2, ENARF_ERR_INVALID_ARG, "enarf_render_fine_bwd: bad sizes n=%d Nf=%d", n, Nf);
    ENARF_REQUIRE(Nf <= TP_MAX, ENARF_ERR_UNSUPPORTED, "enarf_render_fine_bwd: Nf=%d exceeds %d", Nf, TP_MAX);
    BwdArgs a = {};
    a.sc = *scene;
    a.dmin = depth_min;
    a.dmax = depth_max;
    a.dir = ray_dir;
    a.bins = bins;
    a.saved = point_saved;
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
    a.n = n;
    a.S = Nf;
    a.rays_per_tile = max(1, min(MAX_RAYS, TP_MAX / Nf));
    a.tiles_per_sample = (n + a.rays_per_tile - 1) / a.rays_per_tile;
    const int total = scene->B * a.tiles_per_sample;
    const int grid = min(total, enarf_sm_count());
    const size_t smem = sizeof(BSmem);
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
