// K_b1: batch-global near/far and per-ray depth range (reference: libraries/NeRF/rendering.py:10-79).
// One thread per ray; the 32 probes x P slab tests live in registers, nothing is materialised
// (the reference builds a (B,P,3,n*32) temporary, 1.2 GB at config 2).
#include "common.cuh"

#define SQRT3F 1.7320508075688772f

__global__ void near_far_kernel(const float* __restrict__ pose, int BP, float near0, float far0,
                                float* __restrict__ nearfar) {
    __shared__ float s_min[32], s_max[32];
    float mn = 3.4e38f, mx = -3.4e38f;
    for (int i = threadIdx.x; i < BP; i += blockDim.x) {
        float tz = pose[(size_t)i * 16 + 11];
        mn = fminf(mn, tz);
        mx = fmaxf(mx, tz);
    }
    for (int o = 16; o; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if ((threadIdx.x & 31) == 0) {
        s_min[threadIdx.x >> 5] = mn;
        s_max[threadIdx.x >> 5] = mx;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (blockDim.x >> 5); ++w) {
            mn = fminf(mn, s_min[w]);
            mx = fmaxf(mx, s_max[w]);
        }
        nearfar[0] = fmaxf(csub(mn, SQRT3F), near0);   // rendering.py:16
        nearfar[1] = fmaxf(cadd(mx, SQRT3F), far0);    // rendering.py:17
    }
}

extern "C" int enarf_near_far(const float* pose, int B, int P, float near0, float far0, float* nearfar,
                              void* stream) {
    ENARF_REQUIRE(pose && nearfar && B > 0 && P > 0, ENARF_ERR_INVALID_ARG, "enarf_near_far: bad arguments");
    near_far_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(pose, B * P, near0, far0, nearfar);
    ENARF_CHECK_LAUNCH("enarf_near_far");
    return ENARF_OK;
}

// probe i of linspace(near, far, 32): counts up from near in the first half, down from far in the second
__device__ __forceinline__ float probe_depth(float near, float far, float step, int i) {
    return (i < N_PROBE / 2) ? cadd(near, cmul(step, (float)i)) : csub(far, cmul(step, (float)(N_PROBE - 1 - i)));
}

__global__ void __launch_bounds__(128) depth_range_kernel(
    const float* __restrict__ image_coord, const float* __restrict__ inv_K, const float* __restrict__ pose,
    const float* __restrict__ nearfar, int P, int n, float* __restrict__ depth_min, float* __restrict__ depth_max,
    float* __restrict__ ray_dir, uint8_t* __restrict__ validity) {
    __shared__ float s_pose[ENARF_MAX_PARTS * 12];
    __shared__ float s_K[9];
    const int b = blockIdx.y;
    for (int i = threadIdx.x; i < P * 12; i += blockDim.x) {
        int p = i / 12, j = i - p * 12;
        int r = (j < 9) ? j / 3 : j - 9, c = (j < 9) ? j % 3 : 3;
        s_pose[i] = pose[((size_t)(b * P + p) * 4 + r) * 4 + c];
    }
    if (threadIdx.x < 9) s_K[threadIdx.x] = inv_K[b * 9 + threadIdx.x];
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float near = nearfar[0], far = nearfar[1];
    const float step = __fdiv_rn(csub(far, near), (float)(N_PROBE - 1));
    const float* u = image_coord + (size_t)b * 3 * n;
    const float u0 = u[i], u1 = u[n + i], u2 = u[2 * n + i];
    const float dx = cdot3(s_K[0], s_K[1], s_K[2], u0, u1, u2);
    const float dy = cdot3(s_K[3], s_K[4], s_K[5], u0, u1, u2);
    const float dz = cdot3(s_K[6], s_K[7], s_K[8], u0, u1, u2);
    // Parts whose bounding sphere (radius sqrt 3 of the unit cube, + 2 % for rounding) the ray misses cannot contain a
    // probe: they are left out of the 32 x P cube tests below (conservative, so the result is bit for bit the same; most
    // rays miss most of the 24 parts)
    uint32_t near_parts = 0;
    {
        const float inv_dd = 1.f / fmaxf(dx * dx + dy * dy + dz * dz, 1e-30f);
        for (int p = 0; p < P; ++p) {
            const float* q = s_pose + p * 12;
            const float td = q[9] * dx + q[10] * dy + q[11] * dz;
            const float dist2 = (q[9] * q[9] + q[10] * q[10] + q[11] * q[11]) - td * td * inv_dd;
            if (dist2 <= 3.f * 1.04f + 1e-3f) near_parts |= 1u << p;
        }
    }
    uint32_t hit = 0;  // bit j: probe j lies inside some part cube
    for (int j = 0; j < N_PROBE && near_parts; ++j) {
        const float d = probe_depth(near, far, step, j);
        const float x = cmul(dx, d), y = cmul(dy, d), z = cmul(dz, d);
        bool in = false;
        for (uint32_t pm = near_parts; pm && !in; pm &= pm - 1) {
            const int p = __ffs(pm) - 1;
            const float* q = s_pose + p * 12;
            float e0 = csub(x, q[9]), e1 = csub(y, q[10]), e2 = csub(z, q[11]);
            float l0 = cdot3(q[0], q[3], q[6], e0, e1, e2);
            float l1 = cdot3(q[1], q[4], q[7], e0, e1, e2);
            float l2 = cdot3(q[2], q[5], q[8], e0, e1, e2);
            in = fabsf(l0) <= 1.f && fabsf(l1) <= 1.f && fabsf(l2) <= 1.f;
        }
        hit |= (in ? 1u : 0u) << j;
    }
    float dmin = near, dmax = far;
    if (hit) {
        dmin = probe_depth(near, far, step, __ffs(hit) - 1);
        dmax = probe_depth(near, far, step, 31 - __clz(hit));
    }
    dmin = fmaxf(dmin, near);  // rendering.py:69
    const size_t o = (size_t)b * n + i;
    depth_min[o] = dmin;
    depth_max[o] = dmax;
    validity[o] = hit ? 1 : 0;
    float* rd = ray_dir + (size_t)b * 3 * n;
    rd[i] = dx;
    rd[n + i] = dy;
    rd[2 * n + i] = dz;
}

extern "C" int enarf_depth_range(const float* image_coord, const float* inv_K, const float* pose,
                                 const float* nearfar, int B, int P, int n, float* depth_min, float* depth_max,
                                 float* ray_dir, uint8_t* validity, void* stream) {
    ENARF_REQUIRE(image_coord && inv_K && pose && nearfar && depth_min && depth_max && ray_dir && validity,
                  ENARF_ERR_INVALID_ARG, "enarf_depth_range: NULL argument");
    ENARF_REQUIRE(B > 0 && n > 0 && P > 0 && P <= ENARF_MAX_PARTS && B <= 65535, ENARF_ERR_INVALID_ARG,
                  "enarf_depth_range: bad sizes B=%d P=%d n=%d", B, P, n);
    dim3 grid((n + 127) / 128, B);
    depth_range_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(image_coord, inv_K, pose, nearfar, P, n, depth_min,
                                                              depth_max, ray_dir, validity);
    ENARF_CHECK_LAUNCH("enarf_depth_range");
    return ENARF_OK;
}
