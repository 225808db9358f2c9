/* dmsodom NID keyframing -- C ABI.  SURVEY 8(f) "next" row N4.
 *
 * The normalised-information-distance score ElasticFusion::fuseFrame (ElasticFusion.cpp:639-677) uses to decide
 * whether a frame is fused: joint histograms of the key frame's predicted image / depth against the tracker's live
 * pyramids (nextImg(level) / nextD(level), RGBDOdometry.h:79-90), entropies, score.  Reference kernels:
 * elasticfusion/Core/src/Cuda/cudafuncs.cu:906-917 (bin_idx), :1086-1157 (joint histogram kernels),
 * :1513-1650 (computeNIDImg), :1794-1915 (computeNIDDepth); host: MutualInformation.cpp:154-212, KeyFrame.h:69-160.
 *
 * Differences by construction: histograms are exact u32 counts (the reference adds 1.0f with atomicAdd: identical
 * values below 2^24); entropies are evaluated on the device (no histogram download); samples whose bin falls outside
 * the histogram -- which the reference writes out of bounds (cf.cu:906-917 has no range check) -- are dropped.
 */
#ifndef DMSODOM_NID_C_H
#define DMSODOM_NID_C_H
#include "odom_c.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dms_nid dms_nid;

/* MutualInformation(num_bins_depth, num_bins_img, num_pyr_levels)  MutualInformation.cpp:3-30.  pyr_level in 0..2. */
int dms_nid_create(dms_nid** out, int device, int width, int height, int pyr_level, int num_bins_img, int num_bins_depth);
int dms_nid_destroy(dms_nid* h);
/* KeyFrame(activeImageTex, activeVertexTex, ., oldImageTex, oldVertexTex, ., pose, depthCutoff)  KeyFrame.h:69-160
 * (intensity image + verticesToDepth, cf.cu:597-617) followed by the pyramid part of MutualInformation::nidImg
 * (MutualInformation.cpp:168-178: pyrDownUcharGauss / pyrDownGaussF down to pyr_level).  Dense RGBA32F / RGBA8 device
 * images of the frame size. */
int dms_nid_set_keyframe(dms_nid* h, const float* d_vertex4, const uint8_t* d_rgba, const float* d_old_vertex4,
                         const uint8_t* d_old_rgba, float depthCutoff, void* stream);
/* nidImg + nidDepth + nid_score (EF.cpp:651-662) against the live pyramids of `tracker` (after its initRGB). */
int dms_nid_score(dms_nid* h, dms_odom* tracker, float max_depth_m, float depth_lambda, float* nid_img, float* nid_depth,
                  float* score, void* stream);
/* Joint histograms of the last dms_nid_score call (which: 0 = image, 1 = depth), num_bins^2 u32, row = live bin. */
int dms_nid_get_histogram(dms_nid* h, int which, uint32_t* h_dst, size_t n_words, void* stream);

/* Stateless forms of computeNIDImg (cf.cu:1513-1650) / computeNIDDepth (cf.cu:1794-1915) on pitched device images.
 * d_hist: DMS_NID_SCRATCH_WORDS(num_bins) u32 of 8-byte aligned device scratch (histogram, marginals, partial sums;
 * cleared by the call), d_out: 4 floats. */
#define DMS_NID_SCRATCH_WORDS(nb) ((size_t)(nb) * (nb) + (nb) + 520)
int dms_nid_image(const uint8_t* d_img_kf, const uint8_t* d_img_kf_old, const float* d_dmap_kf, const float* d_dmap_kf_old,
                  const uint8_t* d_img_curr, size_t pitch_u8, size_t pitch_f32, size_t pitch_curr, int rows, int cols,
                  int num_bins, uint32_t* d_hist, float* d_out, float* nid, void* stream);
int dms_nid_depth(const float* d_dmap_kf, const float* d_dmap_kf_old, const float* d_dmap_curr, size_t pitch_f32,
                  size_t pitch_curr, int rows, int cols, int num_bins, float max_depth_mm, uint32_t* d_hist, float* d_out,
                  float* nid, void* stream);

#ifdef __cplusplus
}
#endif
#endif
