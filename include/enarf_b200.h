/* enarf_b200.h -- C ABI of libenarf_b200.so: the B200 (sm_100a) articulated tri-plane renderer hot path.
 *
 * Drop-in boundary for nogu-atsu/ENARF-GAN (reference paths below are relative to that repo):
 *   - replaces the pybind module `triplane_sampler_cuda` (cuda_extension/TriplaneSampler.cpp:15-59);
 *   - supplies fused kernels for what libraries/NeRF/rendering.py:10-359, libraries/triplane/sampling.py:9-127,
 *     models/narf.py:147-275 and libraries/triplane/triplane_nerf.py:32-48 do with ~60 eager torch ops.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its comment says "host"; the caller owns every buffer
 *     (outputs, zero-initialised gradient buffers, scratch); the library never allocates or retains them;
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream); calls are asynchronous and
 *     never synchronise the device; the current device must be the one owning the pointers (the Python binding
 *     wraps every call in torch.cuda.device(tensor.device)); re-entrant: the only mutable process-wide state is a
 *     per-device SM-count cache and the round-robin cursor of the per-device tile-counter slots (both benign under
 *     concurrent callers), plus the thread-local error string;
 *   - every function returns 0 on success or an ENARF_ERR_* code; enarf_last_error() gives the message of
 *     the last failure on the calling thread; no exceptions cross the boundary;
 *   - all floating-point buffers are fp32 unless stated; tensors are dense row-major with the shape given.
 */
#ifndef ENARF_B200_H
#define ENARF_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ENARF_OK 0
#define ENARF_ERR_INVALID_ARG 1   /* null pointer, bad size, unsupported mode */
#define ENARF_ERR_UNSUPPORTED 2   /* shape outside what the kernels are built for */
#define ENARF_ERR_CUDA 3          /* launch / runtime failure (message has the CUDA error string) */

#define ENARF_FEAT_DIM 32         /* tri-plane feature channels per plane (models/narf.py:21) */
#define ENARF_HIDDEN 64           /* decoder width (models/narf.py:75: StyledMLP(32, 64, 4)) */
#define ENARF_MAX_PARTS 32        /* part mask is one uint32 */

#define ENARF_DTYPE_F32 0
#define ENARF_DTYPE_BF16 1

#define ENARF_RNG_GIVEN 0         /* bin indices + jitter supplied by the caller (torch.multinomial parity) */
#define ENARF_RNG_PHILOX 1        /* in-kernel Philox4x32-10 inverse-CDF sampling */

/* Everything constant across one render call (one generator forward). */
typedef struct enarf_scene {
    int32_t B;                    /* samples in the batch */
    int32_t P;                    /* parts after transform_pose (23 or 24), <= ENARF_MAX_PARTS */
    int32_t H, W;                 /* tri-plane resolution (256 x 256 in the reference) */
    const float* pose;            /* (B,P,4,4) part->camera frames, translation already x coordinate_scale
                                     (libraries/NARF/pose_utils.py:129-148, rendering.py:258-260) */
    const float* cscale;          /* (B,P) canonical_bone_length / bone_length / coordinate_scale (narf.py:165) */
    const float* canon_pose;      /* (P,4,4) canonical part frames (narf.py:84-120) */
    const void* feat_planes;      /* (B,3,H,W,32) channels-last feature planes made by enarf_repack_planes */
    int32_t feat_dtype;           /* ENARF_DTYPE_F32 | ENARF_DTYPE_BF16 */
    int32_t reserved0;
    const float* prob_planes;     /* NCHW part-probability planes: element [b][3*p+k][y][x] at
                                     prob_planes + b*prob_batch_stride + ((3*p+k)*H + y)*W + x (narf.py:239) */
    int64_t prob_batch_stride;    /* elements between consecutive samples of prob_planes */
    const float* w1;              /* (B,64,32) per-sample modulated, normalised decoder weights */
    const float* w2;              /* (B,64,64)   (libraries/custom_stylegan2/net.py:229-247) */
    const float* w3;              /* (B,4,64) */
    const float* b1;              /* (64)  StyledConv.bias (net.py:299) */
    const float* b2;              /* (64) */
    const float* b3;              /* (4) */
} enarf_scene_t;

/* Message of the last error raised on this thread ("" if none). */
const char* enarf_last_error(void);
/* Library / ABI version, e.g. 100 = 0.1.0. */
int enarf_version(void);

/* NCHW tri-plane (B, C_total, H, W) fp32 -> channels-last feature planes (B,3,H,W,32) in out_dtype.
 * Channel 32*k+c of the input becomes [b][k][y][x][c] (channel order of sampling.py:28). */
int enarf_repack_planes(const float* tri_plane_nchw, int B, int C_total, int H, int W,
                        void* feat_out, int out_dtype, void* stream);
/* Gradient of the above: d_feat (B,3,H,W,32) fp32 is added into (accumulate != 0) or stored over (accumulate == 0,
 * the destination channels may be uninitialised) d_tri_plane_nchw channels [0,96). */
int enarf_unpack_plane_grads(const float* d_feat, int B, int C_total, int H, int W, int accumulate,
                             float* d_tri_plane_nchw, void* stream);

/* Deformation-field tri-plane producer of the DSO ("denarf") model (models/narf.py:39-58: flow -> warped grid ->
 * F.grid_sample of the canonical feature planes, bilinear / zero padding / align_corners=False), emitting the
 * renderer's own layout so that the warped planes are never built as NCHW nor re-packed.
 * canon_feat (3,H,W,32) fp32 channels-last (enarf_repack_planes of the (1,C,H,W) parameter); flow (B,6,H,W) fp32 as
 * the flow head emits it (channel 2k = x offset, 2k+1 = y offset of plane k, in texels); feat_out (B,3,H,W,32) in
 * out_dtype, directly usable as enarf_scene_t.feat_planes.  W must be a multiple of 4. */
int enarf_warp_planes_fwd(const float* canon_feat, const float* flow, int B, int H, int W, void* feat_out,
                          int out_dtype, void* stream);
/* Gradient of the above.  d_feat (B,3,H,W,32) in d_feat_dtype (the fp32 buffer enarf_render_fine_bwd fills, or bf16).
 * d_canon_feat (3,H,W,32) fp32 is ADDED into (caller zero-initialises; summed over the batch); d_flow (B,6,H,W) is
 * STORED.  Either output may be NULL. */
int enarf_warp_planes_bwd(const void* d_feat, int d_feat_dtype, const float* canon_feat, const float* flow, int B,
                          int H, int W, float* d_canon_feat, float* d_flow, void* stream);

/* Batch-global near/far (rendering.py:15-17): nearfar[0] = max(min tz - sqrt3, near0),
 * nearfar[1] = max(max tz + sqrt3, far0) over all B*P part translations of `pose`. */
int enarf_near_far(const float* pose, int B, int P, float near0, float far0, float* nearfar, void* stream);

/* decide_frustrum_range (rendering.py:10-79).  image_coord (B,3,n) homogeneous pixels; inv_K (B,3,3);
 * outputs depth_min/depth_max (B,n), ray_dir (B,3,n), validity (B,n) uint8 (1 = some probe in some part). */
int enarf_depth_range(const float* image_coord, const float* inv_K, const float* pose, const float* nearfar,
                      int B, int P, int n, float* depth_min, float* depth_max, float* ray_dir,
                      uint8_t* validity, void* stream);

/* Limits of the fused passes below: 2 <= Nc, Nf <= 512 samples per ray, P <= 32 parts, H, W >= 2
 * (ENARF_ERR_UNSUPPORTED / ENARF_ERR_INVALID_ARG otherwise). */
/* Coarse pass (rendering.py:119-190 + narf.py:176-275, no gradient): Nc stratified mid-points per ray ->
 * density -> alpha weights -> max-blur + 0.01.  weights_out (B,n,Nc); density_out (B,n,Nc) optional (NULL).
 * opaque_rays (one int, optional): += the number of rays whose coarse transmittance ends below 1e-2 -- the hint
 * enarf_render_fine_fwd uses to decide whether early-termination rounds are worth their barriers.
 * ert_eps / ert_hint / ert_stats: early ray termination of the coarse samples themselves, as in
 * enarf_render_fine_fwd below (a dropped coarse sample gets weight 0, i.e. the blurred weight's 0.01 floor; its true
 * weight was below ert_eps); ert_hint is typically the PREVIOUS render call's opaque_rays. */
int enarf_render_coarse(const enarf_scene_t* scene, const float* depth_min, const float* depth_max,
                        const float* ray_dir, int n, int Nc, float* weights_out, float* density_out,
                        int* opaque_rays, float ert_eps, const int* ert_hint, int* ert_stats, void* stream);

/* Hierarchical sampling (rendering.py:192-197): per ray draw Nf bins from the blurred coarse weights,
 * bins = idx/Nc + jitter/Nc, sorted ascending.  weights (R,Nc).  mode GIVEN: idx (R,Nf) int32 and jitter
 * (R,Nf) are inputs.  mode PHILOX: counter = (ray_id_offset + ray, sample), key = seed + *seed_dev (seed_dev: one
 * uint64 in device memory, optional -- a render captured in a CUDA graph keeps its seed there and bumps it with a
 * captured add, so every replay draws afresh); idx_out optional. */
int enarf_fine_sample(const float* weights, int R, int Nc, int Nf, int mode, const int32_t* idx,
                      const float* jitter, uint64_t seed, const uint64_t* seed_dev, int64_t ray_id_offset,
                      float* bins_out, int32_t* idx_out, void* stream);

/* Optional per-(sample point, part) checkpoint written by the fine forward and consumed by the fine backward, so
 * that the backward neither rebuilds its pair records (canonical coordinates, bilinear footprints, 12 probability
 * texels, sigmoids) nor re-reads the 12 feature texels of every pair.  The caller owns all buffers and zeroes
 * `cursor` before the forward; tiles that do not fit `capacity` are marked -1 in tile_base and the backward
 * recomputes them.  tile_base has B * enarf_fine_tiles_per_sample(n, Nf) entries. */
typedef struct {
    float* rec;          /* (capacity, 20): texel offsets, part | active index << 8, 12 weights x part weight,
                            1 / part weight, 3 sigmoids */
    float* feat;         /* (capacity, 32): part-weighted bilinear feature of the pair; 16-byte aligned */
    int32_t* tile_base;  /* first pair of every tile; -1: the tile was not checkpointed (the backward recomputes it);
                            -2: the tile has no pair at all (the backward skips it) */
    int32_t* cursor;     /* one int: pairs handed out so far */
    int64_t capacity;    /* pairs */
    int32_t skip_last;   /* != 0: the forward leaves out the LAST sample of every ray.  That sample only closes the last
                            interval: the reference evaluates it and then drops its density and colour
                            (rendering.py:306-308, [..., :Np-1]), so no rendered output or gradient depends on it; its
                            point_out entry is then 0.  For callers that do not read the per-sample outputs (training). */
    int32_t reserved;
    uint8_t* count;      /* optional (NULL to skip), (B,n,Nf): number of valid parts of every sample, written by the
                            forward; with it the backward rebuilds a checkpointed tile's active list without redoing
                            the 24-part geometry */
} enarf_pair_ckpt_t;
int enarf_fine_tiles_per_sample(int n, int Nf);

/* Fine pass forward, fused gather -> decoder -> compositing (rendering.py:198-200,283-335).
 * bins (B,n,Nf) sorted.  Outputs: color (B,3,n), mask (B,n), disparity (B,n),
 * fine_weights (B,n,Nf-1), fine_depth (B,n,Nf); point_out (B,n,Nf,4) = per-sample (r,g,b,sigma), optional
 * (NULL to skip) -- it is what the backward consumes and what rendering.py:287 exposes as fine_density;
 * feat_out (B,n,Nf,32), optional (NULL to skip, the inference path): the gathered, part-weighted feature of
 * every sample that lies inside at least one part (other rows are left untouched) -- a checkpoint that lets the
 * backward skip its own gather; 16-byte aligned.
 * Early ray termination (north_star step (e); the reference composites every sample, rendering.py:307-335):
 * ert_eps > 0 evaluates a ray's samples front to back in groups of 16 and stops once its exact prefix transmittance
 * T = exp(-sum sigma*delta) is below ert_eps; the remaining samples get density 0 (their weights would sum to less
 * than ert_eps: colour / mask / every fine_weight change by < ert_eps, disparity by < ert_eps / near) and are recorded
 * in point_out with sigma = -1, from which the backward (same ert_eps) leaves them out as well.  point_out is then
 * mandatory.  ert_hint (one int, optional): enarf_render_coarse's opaque_rays -- with fewer than B*n/64 opaque rays
 * everything is evaluated in one round (no extra barriers); NULL = always in rounds.  ert_stats (two ints, optional):
 * [0] += samples evaluated, [1] += samples skipped.  ert_eps = 0: off, the reference's exact semantics.
 * bg_const: constant background composited in the epilogue, color = rgb + (1 - mask) * bg_const (the generator's
 * fg + (1 - mask) * bg with its black background bg = -1, models/generator.py:101-107; 0 = plain foreground). */
int enarf_render_fine_fwd(const enarf_scene_t* scene, const float* depth_min, const float* depth_max,
                          const float* ray_dir, const float* bins, int n, int Nf, float* color, float* mask,
                          float* disparity, float* fine_weights, float* fine_depth, float* point_out,
                          float* feat_out, const enarf_pair_ckpt_t* pair_ckpt /* NULL = none */, float ert_eps,
                          const int* ert_hint, int* ert_stats, float bg_const, void* stream);

/* Fine pass backward.  point_saved (B,n,Nf,4) is the forward's point_out; feat_saved (B,n,Nf,32) is the forward's
 * feat_out or NULL (the features are then gathered again); pair_ckpt requires feat_saved.  Upstream gradients (any may be NULL = zero):
 * g_color (B,3,n), g_mask (B,n), g_disparity (B,n), g_fine_weights (B,n,Nf-1).  Gradient buffers (caller
 * zero-initialises; results are ADDED): d_feat (B,3,H,W,32) fp32 channels-last, d_prob same layout/stride as
 * scene->prob_planes, d_w1 (B,64,32), d_w2 (B,64,64), d_w3 (B,4,64), d_b1 (64), d_b2 (64), d_b3 (4).
 * ert_eps: the value the forward was called with (> 0: the samples it dropped, sigma < 0 in point_saved, take no part
 * and carry no gradient; the tile's pair order is the forward's); bg_const: the forward's (g_color then also feeds
 * d mask through the composite). */
int enarf_render_fine_bwd(const enarf_scene_t* scene, const float* depth_min, const float* depth_max,
                          const float* ray_dir, const float* bins, const float* point_saved, const float* feat_saved,
                          const enarf_pair_ckpt_t* pair_ckpt /* the forward's, or NULL */, int n, int Nf, const float* g_color, const float* g_mask, const float* g_disparity,
                          const float* g_fine_weights, float* d_feat, float* d_prob, int64_t d_prob_batch_stride,
                          float* d_w1, float* d_w2, float* d_w3, float* d_b1, float* d_b2, float* d_b3, float ert_eps,
                          float bg_const, void* stream);

/* ---- DSO step around the renderer ----------------------------------------------------------------
 * mask_based_sampler (libraries/NeRF/ray_sampler.py:7-39) in two calls.
 * enarf_ray_scores: score = max_pool2d(mask, 2*radius+1, stride 1, padding radius) + U[0,1) per pixel.  mask, scratch,
 * score (B,h,w) fp32; noise (B,h,w) caller-drawn uniforms, or NULL for in-kernel Philox keyed by `seed`. */
int enarf_ray_scores(const float* mask, int B, int h, int w, int radius, const float* noise, uint64_t seed,
                     float* scratch, float* score, void* stream);
/* enarf_ray_scores_static: enarf_ray_scores for callers that cannot allocate device memory (code under CUDA-graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same inputs, same `score` output, bit-identical results; the row-pass scratch is a
 * file-scope `static __device__ __align__(256)` array inside the library (ENARF_STATIC_SCRATCH_FLOATS floats = 4 MB,
 * held by every GPU that loads the library until the process exits), its device address resolved once per device with
 * cudaGetSymbolAddress.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 * B*h*w > ENARF_STATIC_SCRATCH_FLOATS returns ENARF_ERR_UNSUPPORTED with a message naming the limit.  The array is
 * shared by every call and keeps the last call's contents (each call overwrites all it reads, nothing needs
 * clearing): calls that overlap on the GPU -- different streams, no ordering between them -- race on it. */
#define ENARF_STATIC_SCRATCH_FLOATS (1 << 20)   /* e.g. four 512 x 512 masks (BASELINE configs[3]) */
int enarf_ray_scores_static(const float* mask, int B, int h, int w, int radius, const float* noise, uint64_t seed,
                            float* score, void* stream);
/* enarf_select_rays: the k largest scores of every image (torch.topk(..., sorted=False): the SET is exact, ties at
 * the threshold go to the lower pixel index; output in ascending pixel order).  score (B,n); idx_out (B,k) int64 flat
 * pixel ids; homo_out (B,3,k), optional: pixel centres (x + .5, y + .5, 1) with x = id % width, y = id / width. */
int enarf_select_rays(const float* score, int B, int n, int width, int k, int64_t* idx_out, float* homo_out,
                      void* stream);
/* PhotometricLoss (libraries/NeRF/loss.py:14-48) fused: gather of the ground truth at ray_idx, colour loss
 * (loss_type 0: MSE, 1: mean of max(|diff|, 0.01)) x color_coef, mask MSE x mask_coef (gt_mask NULL: skipped).
 * rendered_color (B,3,k), rendered_mask (B,k), gt_color (B,3,npix), gt_mask (B,npix), ray_idx (B,k) int64.
 * loss[0] += colour loss, loss[1] += mask loss (caller zeroes); d_color (B,3,k) / d_mask (B,k), optional: the
 * gradients of those two numbers with respect to the rendered colour / mask. */
int enarf_photometric_loss(const float* rendered_color, const float* rendered_mask, const float* gt_color,
                           const float* gt_mask, const int64_t* ray_idx, int B, int k, int npix, int loss_type,
                           float color_coef, float mask_coef, float* loss, float* d_color, float* d_mask, void* stream);

/* Point query = calc_density_and_color_from_camera_coord_v2 (models/narf.py:176-211): points (B,3,N) in
 * the scaled camera frame -> density (B,1,N), color (B,3,N) (NULL to skip), part_mask (B,N) uint32 bit p =
 * part p valid (NULL to skip). */
int enarf_density_query(const enarf_scene_t* scene, const float* points, int N, float* density, float* color,
                        uint32_t* part_mask, void* stream);

/* The literal replacement of triplane_sampler_cuda.triplane_sampler_forward / _backward
 * (cuda_extension/TriplaneSampler.cpp:15-52, TriplaneSampler_kernel.cu:13-229): input (N,3C,H,W) NCHW,
 * grid (N,Ho,Wo,3) -> output (N,C,Ho,Wo) = sum over the three planes of the sample at
 * (x=grid[k], y=grid[(k+1)%3]).  interpolation 0 bilinear | 1 nearest; padding 0 zeros | 1 border | 2 reflection. */
int enarf_triplane_sample_fwd(const float* input, const float* grid, int N, int C, int H, int W, int Ho, int Wo,
                              int interpolation, int padding, int align_corners, float* output, void* stream);
/* grad_input (N,3C,H,W) and grad_grid (N,Ho,Wo,3) are ADDED into (caller zero-initialises); either may be NULL. */
int enarf_triplane_sample_bwd(const float* grad_output, const float* input, const float* grid, int N, int C,
                              int H, int W, int Ho, int Wo, int interpolation, int padding, int align_corners,
                              float* grad_input, float* grad_grid, void* stream);

/* Fast path of the same op for the reference's one real call shape (sampling.py:22-26: C = 32, bilinear): the planes
 * are taken channels-last, feat_planes (N,3,H,W,32) = enarf_repack_planes of `input`; grid (N,points,3);
 * output (N,32,points).  One texel is one 128-B line instead of 32 separate sectors. */
int enarf_triplane_sample_cl_fwd(const float* feat_planes, const float* grid, int N, int H, int W, int64_t points,
                                 int padding, int align_corners, float* output, void* stream);
/* d_feat_planes (N,3,H,W,32) and grad_grid (N,points,3) are ADDED into (caller zero-initialises); either may be NULL.
 * enarf_unpack_plane_grads turns d_feat_planes into the NCHW grad_input. */
int enarf_triplane_sample_cl_bwd(const float* grad_output, const float* feat_planes, const float* grid, int N, int H,
                                 int W, int64_t points, int padding, int align_corners, float* d_feat_planes,
                                 float* grad_grid, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ENARF_B200_H */
