// DSO ("denarf") step around the renderer: ray selection and the sparse photometric loss.
//   enarf_ray_scores        mask_based_sampler's score image (libraries/NeRF/ray_sampler.py:7-28): the foreground mask
//                           dilated by a (2r+1)^2 max-pool (r = 64 in the reference) plus U[0,1) noise;
//   enarf_select_rays       torch.topk(score, k, sorted=False) per image as an exact radix select (three histogram
//                           passes over the float bits, ties broken by pixel order) + the rays' pixel centres
//                           (ray_sampler.py:29-38); the reference sorts nothing either -- only the SET of rays matters;
//   enarf_photometric_loss  libraries/NeRF/loss.py:14-48 fused: gather of the ground-truth pixels at the selected rays,
//                           MSE / truncated-MAE colour loss, MSE mask loss, and their gradients, in one pass.
// These are small streaming kernels (a 512 x 512 image is 1 MB): what they remove is ~25 eager torch launches and
// their temporaries per DSO step, not bandwidth.
#include "common.cuh"

// ---- score image: separable max (rows, then columns) + noise ---------------------------------------------
// pass 1: one CTA per image row, the row staged in shared memory
__global__ void __launch_bounds__(256) dilate_rows_kernel(const float* __restrict__ mask, int h, int w, int r,
                                                          float* __restrict__ out) {
    extern __shared__ float row[];
    const size_t base = (size_t)blockIdx.x * w;     // blockIdx.x = b * h + y
    for (int x = threadIdx.x; x < w; x += blockDim.x) row[x] = mask[base + x];
    __syncthreads();
    for (int x = threadIdx.x; x < w; x += blockDim.x) {
        float m = -3.4e38f;                          // max_pool2d pads with -inf: only in-range taps count
        for (int t = max(0, x - r); t <= min(w - 1, x + r); ++t) m = fmaxf(m, row[t]);
        out[base + x] = m;
    }
}

// pass 2: columns (threads along x: coalesced) + noise
__global__ void __launch_bounds__(256) dilate_cols_kernel(const float* __restrict__ tmp, const float* __restrict__ noise,
                                                          uint64_t seed, int use_philox, int h, int w, int r,
                                                          float* __restrict__ score) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= w) return;
    const float* img = tmp + (size_t)b * h * w;
    float m = -3.4e38f;
    for (int t = max(0, y - r); t <= min(h - 1, y + r); ++t) m = fmaxf(m, img[(size_t)t * w + x]);
    const size_t i = ((size_t)b * h + y) * w + x;
    float u;
    if (use_philox) {
        const uint4 o = philox4x32_10(make_uint4((uint32_t)i, (uint32_t)(i >> 32), 0x6d61736bu, 0u),
                                      make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
        u = u01(o.x);
    } else {
        u = noise[i];
    }
    score[i] = m + u;
}

extern "C" int enarf_ray_scores(const float* mask, int B, int h, int w, int radius, const float* noise, uint64_t seed,
                                float* scratch, float* score, void* stream) {
    ENARF_REQUIRE(mask && scratch && score && B > 0 && h > 0 && w > 0 && radius >= 0, ENARF_ERR_INVALID_ARG,
                  "enarf_ray_scores: bad arguments");
    ENARF_REQUIRE(w <= 8192, ENARF_ERR_UNSUPPORTED, "enarf_ray_scores: width %d exceeds 8192", w);
    cudaStream_t st = (cudaStream_t)stream;
    dilate_rows_kernel<<<B * h, 256, (size_t)w * sizeof(float), st>>>(mask, h, w, radius, scratch);
    ENARF_CHECK_LAUNCH("enarf_ray_scores");
    dilate_cols_kernel<<<dim3((w + 255) / 256, h, B), 256, 0, st>>>(scratch, noise, seed, noise ? 0 : 1, h, w, radius, score);
    ENARF_CHECK_LAUNCH("enarf_ray_scores");
    return ENARF_OK;
}

// ---- the same score image for callers that cannot allocate -------------------------------------------------
// enarf_ray_scores_static: enarf_ray_scores with the row-pass scratch in a file-scope device array instead of a caller
// buffer (no counterpart in the reference; added for callers that cannot call
// cudaMalloc: CUDA-graph capture, tight serving loops).  Every GPU that loads the library holds the 4 MB until exit.
// The array's name is only a host-side stand-in in host code: its device address is resolved once per device with
// cudaGetSymbolAddress (with that device current) and cached.  No cudaMalloc / cudaFree / allocating library call.
// The row pass overwrites every element the column pass reads (B*h*w floats), so nothing has to start at zero; all
// calls share the one array, so two calls that overlap on the GPU (different streams) race.
static __device__ __align__(256) float g_score_scratch[ENARF_STATIC_SCRATCH_FLOATS];

static float* score_scratch_ptr(const char* who) {
    static float* cached[64] = {};      // per device id; benign race: concurrent first calls store the same address
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        enarf_set_error("%s: cudaGetDevice: %s", who, cudaGetErrorString(e));
        return nullptr;
    }
    if (dev >= 0 && dev < 64) {
        float* p = __atomic_load_n(&cached[dev], __ATOMIC_ACQUIRE);
        if (p) return p;
    }
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_score_scratch);
    if (e != cudaSuccess || !p) {
        enarf_set_error("%s: cudaGetSymbolAddress(g_score_scratch) on device %d: %s", who, dev, cudaGetErrorString(e));
        return nullptr;
    }
    if (dev >= 0 && dev < 64) __atomic_store_n(&cached[dev], p, __ATOMIC_RELEASE);
    return p;
}

extern "C" int enarf_ray_scores_static(const float* mask, int B, int h, int w, int radius, const float* noise,
                                       uint64_t seed, float* score, void* stream) {
    ENARF_REQUIRE(mask && score && B > 0 && h > 0 && w > 0 && radius >= 0, ENARF_ERR_INVALID_ARG,
                  "enarf_ray_scores_static: bad arguments");
    ENARF_REQUIRE(w <= 8192, ENARF_ERR_UNSUPPORTED, "enarf_ray_scores_static: width %d exceeds 8192", w);
    const long long need = (long long)B * h * w;
    ENARF_REQUIRE(need <= (long long)ENARF_STATIC_SCRATCH_FLOATS, ENARF_ERR_UNSUPPORTED,
                  "enarf_ray_scores_static: B*h*w = %lld floats exceeds the static scratch of %d floats "
                  "(ENARF_STATIC_SCRATCH_FLOATS); use enarf_ray_scores with a caller buffer", need,
                  ENARF_STATIC_SCRATCH_FLOATS);
    float* scratch = score_scratch_ptr("enarf_ray_scores_static");
    if (!scratch) return ENARF_ERR_CUDA;
    cudaStream_t st = (cudaStream_t)stream;
    dilate_rows_kernel<<<B * h, 256, (size_t)w * sizeof(float), st>>>(mask, h, w, radius, scratch);
    ENARF_CHECK_LAUNCH("enarf_ray_scores_static");
    dilate_cols_kernel<<<dim3((w + 255) / 256, h, B), 256, 0, st>>>(scratch, noise, seed, noise ? 0 : 1, h, w, radius, score);
    ENARF_CHECK_LAUNCH("enarf_ray_scores_static");
    return ENARF_OK;
}

// ---- exact top-k selection ----------------------------------------------------------------------------------
// One 1024-thread CTA per image.  Scores are finite floats; the map below makes unsigned order equal float order.
__device__ __forceinline__ uint32_t order_key(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

#define SEL_THREADS 1024
__global__ void __launch_bounds__(SEL_THREADS) select_rays_kernel(const float* __restrict__ score, int n, int w, int k,
                                                                  int64_t* __restrict__ idx_out, float* __restrict__ homo) {
    __shared__ uint32_t hist[4096];
    __shared__ uint32_t s_prefix, s_want, s_base;
    __shared__ uint32_t wsum[SEL_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, b = blockIdx.x;
    const float* sc = score + (size_t)b * n;
    // radix select of the k-th largest key: 12 + 12 + 8 bits
    uint32_t prefix = 0, want = (uint32_t)k;     // keys with (key >> shift) == prefix are still candidates
    const int bits[3] = {12, 12, 8};
    int shift = 32;
    for (int pass = 0; pass < 3; ++pass) {
        const int nb = bits[pass];
        shift -= nb;
        for (int i = tid; i < (1 << nb); i += SEL_THREADS) hist[i] = 0;
        __syncthreads();
        for (int i = tid; i < n; i += SEL_THREADS) {
            const uint32_t key = order_key(sc[i]);
            if (pass == 0 || (key >> (shift + nb)) == prefix) atomicAdd(&hist[(key >> shift) & ((1u << nb) - 1)], 1u);
        }
        __syncthreads();
        if (tid == 0) {      // walk the bins from the top until `want` keys are covered
            uint32_t acc = 0;
            int bin = (1 << nb) - 1;
            for (; bin > 0; --bin) {
                if (acc + hist[bin] >= want) break;
                acc += hist[bin];
            }
            s_prefix = (prefix << nb) | (uint32_t)bin;
            s_want = want - acc;          // still needed from inside this bin
        }
        __syncthreads();
        prefix = s_prefix;
        want = s_want;
        __syncthreads();
    }
    // prefix is now the exact key T of the k-th largest score; `want` of the keys equal to T are taken (pixel order)
    const uint32_t T = prefix;
    if (tid == 0) s_base = 0;
    __syncthreads();
    uint32_t ties_taken = 0;       // running count of == T keys before this chunk (uniform across the CTA)
    for (int i0 = 0; i0 < n; i0 += SEL_THREADS) {
        const int i = i0 + tid;
        const uint32_t key = i < n ? order_key(sc[i]) : 0u;
        const bool gt = i < n && key > T, eq = i < n && key == T;
        // ordered compaction: rank of this pixel among the chunk's == T keys, then among the selected ones
        const uint32_t eq_bal = __ballot_sync(0xffffffffu, eq);
        if (lane == 0) wsum[warp] = __popc(eq_bal);
        __syncthreads();
        uint32_t eq_before = ties_taken, eq_chunk = 0;
        for (int wv = 0; wv < SEL_THREADS / 32; ++wv) {
            if (wv < warp) eq_before += wsum[wv];
            eq_chunk += wsum[wv];
        }
        const bool take = gt || (eq && eq_before + __popc(eq_bal & ((1u << lane) - 1)) < want);
        __syncthreads();
        const uint32_t tk_bal = __ballot_sync(0xffffffffu, take);
        if (lane == 0) wsum[warp] = __popc(tk_bal);
        __syncthreads();
        uint32_t before = s_base, total = 0;
        for (int wv = 0; wv < SEL_THREADS / 32; ++wv) {
            if (wv < warp) before += wsum[wv];
            total += wsum[wv];
        }
        if (take) {
            const uint32_t slot = before + __popc(tk_bal & ((1u << lane) - 1));
            idx_out[(size_t)b * k + slot] = i;
            if (homo) {      // pixel centre (x + .5, y + .5, 1): ray_sampler.py:30-38
                float* hp = homo + (size_t)b * 3 * k + slot;
                hp[0] = (float)(i % w) + 0.5f;
                hp[k] = (float)(i / w) + 0.5f;
                hp[2 * (size_t)k] = 1.0f;
            }
        }
        ties_taken += eq_chunk;
        __syncthreads();
        if (tid == 0) s_base += total;
        __syncthreads();
    }
}

extern "C" int enarf_select_rays(const float* score, int B, int n, int width, int k, int64_t* idx_out, float* homo_out,
                                 void* stream) {
    ENARF_REQUIRE(score && idx_out && B > 0 && n > 0 && width > 0 && k > 0 && k <= n, ENARF_ERR_INVALID_ARG,
                  "enarf_select_rays: bad arguments (k=%d n=%d)", k, n);
    select_rays_kernel<<<B, SEL_THREADS, 0, (cudaStream_t)stream>>>(score, n, width, k, idx_out, homo_out);
    ENARF_CHECK_LAUNCH("enarf_select_rays");
    return ENARF_OK;
}

// ---- sparse photometric loss ----------------------------------------------------------------------------------
// One thread per selected ray: gather the ground-truth pixel, accumulate the two loss sums, write the gradients with
// respect to the rendered colour / mask (already divided by the element counts and multiplied by the coefficients).
__global__ void __launch_bounds__(256) photometric_kernel(const float* __restrict__ r_color, const float* __restrict__ r_mask,
                                                          const float* __restrict__ gt_color, const float* __restrict__ gt_mask,
                                                          const int64_t* __restrict__ idx, int B, int k, int npix, int mae,
                                                          float color_coef, float mask_coef, float* __restrict__ loss,
                                                          float* __restrict__ d_color, float* __restrict__ d_mask) {
    __shared__ float s_c[8], s_m[8];
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    float lc = 0.f, lm = 0.f;
    if (t < B * k) {
        const int b = t / k, j = t - b * k;
        const int64_t p = idx[t];
        const float inv_c = color_coef / (3.f * (float)B * (float)k), inv_m = mask_coef / ((float)B * (float)k);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const size_t o = ((size_t)b * 3 + c) * k + j;
            const float d = r_color[o] - gt_color[((size_t)b * 3 + c) * npix + p];   // rendered - ground truth
            if (mae) {     // clamp_min(|diff|, 0.01).mean(): no gradient inside the dead zone (loss.py:35-37)
                lc += fmaxf(fabsf(d), 0.01f);
                if (d_color) d_color[o] = fabsf(d) > 0.01f ? copysignf(inv_c, d) : 0.f;
            } else {
                lc += d * d;
                if (d_color) d_color[o] = 2.f * d * inv_c;
            }
        }
        if (gt_mask) {
            const float d = r_mask[t] - gt_mask[(size_t)b * npix + p];
            lm = d * d;
            if (d_mask) d_mask[t] = 2.f * d * inv_m;
        }
        lc *= inv_c;
        lm *= inv_m;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        lc += __shfl_xor_sync(0xffffffffu, lc, o);
        lm += __shfl_xor_sync(0xffffffffu, lm, o);
    }
    if ((threadIdx.x & 31) == 0) {
        s_c[threadIdx.x >> 5] = lc;
        s_m[threadIdx.x >> 5] = lm;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int wv = 1; wv < 8; ++wv) {
            lc += s_c[wv];
            lm += s_m[wv];
        }
        atomicAdd(loss, lc);
        atomicAdd(loss + 1, lm);
    }
}

extern "C" int enarf_photometric_loss(const float* rendered_color, const float* rendered_mask, const float* gt_color,
                                      const float* gt_mask, const int64_t* ray_idx, int B, int k, int npix, int loss_type,
                                      float color_coef, float mask_coef, float* loss, float* d_color, float* d_mask,
                                      void* stream) {
    ENARF_REQUIRE(rendered_color && gt_color && ray_idx && loss && B > 0 && k > 0 && npix > 0, ENARF_ERR_INVALID_ARG,
                  "enarf_photometric_loss: bad arguments");
    ENARF_REQUIRE(loss_type == 0 || loss_type == 1, ENARF_ERR_INVALID_ARG, "enarf_photometric_loss: loss_type 0 (mse) | 1 (mae)");
    ENARF_REQUIRE(!gt_mask || rendered_mask, ENARF_ERR_INVALID_ARG, "enarf_photometric_loss: mask loss needs the rendered mask");
    const int total = B * k;
    photometric_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(rendered_color, rendered_mask, gt_color, gt_mask,
                                                                             ray_idx, B, k, npix, loss_type, color_coef,
                                                                             mask_coef, loss, d_color, d_mask);
    ENARF_CHECK_LAUNCH("enarf_photometric_loss");
    return ENARF_OK;
}
