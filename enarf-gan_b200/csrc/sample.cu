// K_a2: hierarchical (fine) sampling, one warp per ray.
// Reference: libraries/NeRF/rendering.py:192-197 -- torch.multinomial(weights, Nf, replacement=True)/Nc +
// U[0,1)/Nc, sorted.  Two RNG modes (include/enarf_b200.h): GIVEN consumes the caller's draws (bit parity with
// a torch-drawn stream); PHILOX draws in-kernel and inverts the CDF under the oracle's contract
// (sequential fp32 prefix sum, target = u * total, idx = #{j < Nc-1 : cdf[j] <= target}).
#include "common.cuh"

#define WARPS 4

__global__ void __launch_bounds__(WARPS * 32) fine_sample_kernel(
    const float* __restrict__ weights, int R, int Nc, int Nf, int Nf2, int mode, const int32_t* __restrict__ idx_in,
    const float* __restrict__ jitter, uint64_t seed, const uint64_t* __restrict__ seed_dev, int64_t ray_id_offset,
    float* __restrict__ bins_out, int32_t* __restrict__ idx_out) {
    extern __shared__ float smem[];
    if (seed_dev) seed += *seed_dev;   // a seed that lives on the device: a captured CUDA graph draws afresh on every replay
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* cdf = smem + warp * (Nc + Nf2);
    float* val = cdf + Nc;
    const float fNc = (float)Nc;
    for (int ray = blockIdx.x * WARPS + warp; ray < R; ray += gridDim.x * WARPS) {
        if (mode == ENARF_RNG_PHILOX) {
            for (int j = lane; j < Nc; j += 32) cdf[j] = weights[(size_t)ray * Nc + j];
            __syncwarp();
            if (lane == 0) {
                // sequential fp32 prefix sum (the oracle's contract), 16 elements at a time: the chunk is read into
                // registers with independent loads, summed in registers (one dependent add per element) and written
                // back, instead of one shared-memory round trip per element
                float c = 0.f;
                for (int j0 = 0; j0 < Nc; j0 += 16) {
                    float w[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) w[i] = (j0 + i < Nc) ? cdf[j0 + i] : 0.f;
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        c = (j0 + i == 0) ? w[0] : cadd(c, w[i]);
                        w[i] = c;
                    }
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (j0 + i < Nc) cdf[j0 + i] = w[i];
                }
            }
            __syncwarp();
        }
        const float total = (mode == ENARF_RNG_PHILOX) ? cdf[Nc - 1] : 0.f;
        for (int sidx = lane; sidx < Nf2; sidx += 32) {
            float v = __int_as_float(0x7f800000);  // +inf padding sorts to the end
            if (sidx < Nf) {
                int id;
                float jit;
                if (mode == ENARF_RNG_PHILOX) {
                    const uint64_t rid = (uint64_t)(ray_id_offset + ray);
                    const uint4 o = philox4x32_10(make_uint4((uint32_t)rid, (uint32_t)sidx, 0u, 0u),
                                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
                    const float target = cmul(u01(o.x), total);
                    jit = u01(o.y);
                    int lo = 0, hi = Nc - 1;  // first j in [0, Nc-1) with cdf[j] > target
                    while (lo < hi) {
                        const int mid = (lo + hi) >> 1;
                        if (cdf[mid] <= target) lo = mid + 1; else hi = mid;
                    }
                    id = lo;
                } else {
                    id = idx_in[(size_t)ray * Nf + sidx];
                    jit = jitter[(size_t)ray * Nf + sidx];
                }
                if (idx_out) idx_out[(size_t)ray * Nf + sidx] = id;
                v = cadd(__fdiv_rn((float)id, fNc), __fdiv_rn(jit, fNc));
            }
            val[sidx] = v;
        }
        __syncwarp();
        if (Nf2 == 64) {
            // bitonic sort of 64 values in registers: element i lives in lane i & 31, register i >> 5; partners at
            // distance < 32 are exchanged by shuffle, distance 32 is the lane's own second register
            float v0 = val[lane], v1 = val[lane + 32];
#pragma unroll
            for (int k = 2; k <= 64; k <<= 1) {
#pragma unroll
                for (int j = k >> 1; j > 0; j >>= 1) {
                    if (j == 32) {
                        const float lo = fminf(v0, v1), hi = fmaxf(v0, v1);   // k == 64: ascending everywhere
                        v0 = lo;
                        v1 = hi;
                    } else {
                        const float p0 = __shfl_xor_sync(0xffffffffu, v0, j), p1 = __shfl_xor_sync(0xffffffffu, v1, j);
                        const bool lower = (lane & j) == 0;            // this lane holds the lower index of the pair
                        const bool up0 = (lane & k) == 0, up1 = ((lane + 32) & k) == 0;
                        v0 = (lower == up0) ? fminf(v0, p0) : fmaxf(v0, p0);
                        v1 = (lower == up1) ? fminf(v1, p1) : fmaxf(v1, p1);
                    }
                }
            }
            val[lane] = v0;
            val[lane + 32] = v1;
            __syncwarp();
        } else
        for (int k = 2; k <= Nf2; k <<= 1) {          // bitonic sort, ascending
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int i = lane; i < Nf2; i += 32) {
                    const int partner = i ^ j;
                    if (partner > i) {
                        const float x = val[i], y = val[partner];
                        const bool up = (i & k) == 0;
                        if ((x > y) == up) {
                            val[i] = y;
                            val[partner] = x;
                        }
                    }
                }
                __syncwarp();
            }
        }
        for (int sidx = lane; sidx < Nf; sidx += 32) bins_out[(size_t)ray * Nf + sidx] = val[sidx];
        __syncwarp();
    }
}

extern "C" int enarf_fine_sample(const float* weights, int R, int Nc, int Nf, int mode, const int32_t* idx,
                                 const float* jitter, uint64_t seed, const uint64_t* seed_dev, int64_t ray_id_offset,
                                 float* bins_out, int32_t* idx_out, void* stream) {
    ENARF_REQUIRE(bins_out && R > 0 && Nc >= 2 && Nf >= 1, ENARF_ERR_INVALID_ARG, "enarf_fine_sample: bad arguments");
    ENARF_REQUIRE(Nc <= 1024 && Nf <= 1024, ENARF_ERR_UNSUPPORTED, "enarf_fine_sample: Nc=%d Nf=%d too large", Nc, Nf);
    if (mode == ENARF_RNG_GIVEN) {
        ENARF_REQUIRE(idx && jitter, ENARF_ERR_INVALID_ARG, "enarf_fine_sample: GIVEN mode needs idx and jitter");
    } else if (mode == ENARF_RNG_PHILOX) {
        ENARF_REQUIRE(weights, ENARF_ERR_INVALID_ARG, "enarf_fine_sample: PHILOX mode needs weights");
    } else {
        ENARF_REQUIRE(false, ENARF_ERR_INVALID_ARG, "enarf_fine_sample: unknown mode %d", mode);
    }
    int Nf2 = 1;
    while (Nf2 < Nf) Nf2 <<= 1;
    const size_t smem = (size_t)WARPS * (Nc + Nf2) * sizeof(float);
    const int grid = min((R + WARPS - 1) / WARPS, 16 * enarf_sm_count());
    fine_sample_kernel<<<grid, WARPS * 32, smem, (cudaStream_t)stream>>>(weights, R, Nc, Nf, Nf2, mode, idx, jitter, seed,
                                                                         seed_dev, ray_id_offset, bins_out, idx_out);
    ENARF_CHECK_LAUNCH("enarf_fine_sample");
    return ENARF_OK;
}
