// Fused backward of the fine pass (placeholder until the kernel lands; fails loudly).
#include "common.cuh"

extern "C" int enarf_render_fine_bwd(const enarf_scene_t* scene, const float* depth_min, const float* depth_max,
                                     const float* ray_dir, const float* bins, int n, int Nf, const float* g_color,
                                     const float* g_mask, const float* g_disparity, const float* g_fine_weights,
                                     float* d_feat, float* d_prob, int64_t d_prob_batch_stride, float* d_w1,
                                     float* d_w2, float* d_w3, float* d_b1, float* d_b2, float* d_b3, void* stream) {
    enarf_set_error("enarf_render_fine_bwd: not implemented yet");
    return ENARF_ERR_UNSUPPORTED;
}
