/* dmsodom tracking front end -- C ABI.  SURVEY 8(f) "next" row N1.
 *
 * A headless restatement of the TRACKING part of ElasticFusion::processFrame
 * (elasticfusion/Core/src/ElasticFusion.cpp:99-268 and the local loop-closure check :397-442): the call sequence a
 * real caller drives the two RGBDOdometry instances of a Context with (Context.h:319-320), the gates around it and the
 * GLSL passes that produce its inputs:
 *
 *   filterDepth            EF.cpp:759-768, Shaders/depth_bilateral.frag          -> dms_depth_bilateral (odom_c.h)
 *   predict()              EF.cpp:718-746  IndexMap::combinedPredict (surfel splat, OUT OF SCOPE: supplied by the
 *                          caller through dms_predict_fn) + FillIn::vertex/normal/image
 *   FillIn                 Shaders/FillIn.cpp:65-190, fill_vertex.frag:37-52, fill_normal.frag:38-51 +
 *                          geometry.glsl:53-77, fill_rgb.frag:29-36                -> dms_fill_in
 *   denseEnough            EF.cpp:84-97 on Resize::image (Shaders/Resize.cpp:67-98, 1/20 scale, EF.cpp:43-47)
 *                                                                                  -> dms_dense_fraction
 *   frameToModel tracking  EF.cpp:171-201 (initICPModel, initRGBModel, initICP(depth), initRGB, track)
 *   reloc gates            EF.cpp:203-244 (lastICPError < 1e-4, covariance diagonal <= 1e-4, 10-frame counter)
 *   velocity weighting     EF.cpp:251-266 (rodrigues2 EF.cpp:941-986)
 *   modelToModel check     EF.cpp:399-442 (cov <= 8e-5, lastICPCount > 15000, lastICPError < 3e-4)
 *
 * Everything after the gates (fern database, deformation graph, fusion) stays with the caller, which receives the gate
 * decisions in dms_frontend_result.  Texture sampling follows the shaders at texel centres with clamp-to-edge; the GL
 * passes themselves cannot run on this box, so parity of these three helpers is against the CPU oracle only.
 */
#ifndef DMSODOM_FRONTEND_C_H
#define DMSODOM_FRONTEND_C_H
#include "odom_c.h"

#ifdef __cplusplus
extern "C" {
#endif

/* A predicted model view: dense device images of the frame size.  vertex / normal are RGBA32F (w ignored), image RGBA8
 * (IndexMap.cpp:59-92).  A NULL member means "not available". */
typedef struct dms_model_view {
  const float* d_vertex;
  const float* d_normal;
  const uint8_t* d_image;
} dms_model_view;

enum dms_predict_kind {
  DMS_PREDICT_TRACK = 0,   /* predict(context, rf, 0.7): ACTIVE view at the tracking confidence, EF.cpp:165 */
  DMS_PREDICT_GLOBAL = 1,  /* predict(context, rf): ACTIVE view at the map confidence, EF.cpp:273 */
  DMS_PREDICT_INACTIVE = 2 /* IndexMap::INACTIVE view of the old part of the map, EF.cpp:402-406 */
};
/* Called by dms_frontend_process where the reference calls IndexMap::combinedPredict.  Must enqueue its work on
 * `stream` (or finish before returning).  Return 0 on success; non-zero = view not available (the step that needs it is
 * skipped and reported). */
typedef int (*dms_predict_fn)(void* user, const float pose_rowmajor[16], int kind, int tick, void* stream, dms_model_view* out);

typedef struct dms_frontend_options {
  float depthCutoff;       /* maxDepthProcessed / depthCutoff of the reference (EF.h), metres */
  float icpWeight;         /* Options::icp[context.id()], default 10 */
  int rgbOnly;             /* Context::rgbOnly() */
  int pyramid, fastOdom, so3;
  int frameToFrameRGB;
  int reloc;               /* Context::reloc(): enables the tracking-failure gates */
  int closeLoops;          /* enables the model-to-model check */
  int hybridTracking;      /* Options::hybrid_tracking: 0 = trust the prior pose, no dense tracking (EF.cpp:245-247) */
  float weightMultiplier;
} dms_frontend_options;
void dms_frontend_default_options(dms_frontend_options* o);

typedef struct dms_frontend_result {
  int tick;                /* Context::tick() of the processed frame (starts at 1) */
  int tracked;             /* 0 on the first frame (map initialisation, EF.cpp:132-152) */
  float pose[16];          /* Context::currPose() after the frame, row-major */
  int shouldFillIn;        /* !denseEnough(...)  EF.cpp:167 */
  float denseFraction;
  int trackingOk, lost, trackingCount, lastFrameRecovery; /* EF.cpp:203-244 */
  float weighting;         /* fusion weight from the velocity, EF.cpp:251-266 */
  dms_odom_result frameToModel;
  int loopChecked;         /* the model-to-model registration ran (EF.cpp:399) */
  int loopAccepted;        /* covOk && lastICPCount > 15000 && lastICPError < 3e-4 (EF.cpp:441-442) */
  float loopPose[16];      /* estPose, EF.cpp:437-440 */
  double loopCovDiag[6];
  dms_odom_result modelToModel;
} dms_frontend_result;

typedef struct dms_frontend dms_frontend;

int dms_frontend_create(dms_frontend** out, int device, int width, int height, float cx, float cy, float fx, float fy,
                        const dms_frontend_options* opt);
int dms_frontend_destroy(dms_frontend* fe);
int dms_frontend_set_predictor(dms_frontend* fe, dms_predict_fn fn, void* user);
/* One frame: d_rgba = live RGBA8 image, d_depth_raw = raw u16 millimetres (dense rows unless pitch given),
 * inPose = row-major 4x4 pose prior (NULL = identity on the first frame / previous pose afterwards). */
int dms_frontend_process(dms_frontend* fe, const uint8_t* d_rgba, size_t rgba_pitch, const uint16_t* d_depth_raw,
                         size_t depth_pitch, const float* inPose, dms_frontend_result* out, void* stream);
/* The two trackers (0 = frameToModel, 1 = modelToModel, Context.h:319-320), the filtered depth and the fill-in maps of
 * the last frame, for callers that go on to fuse. */
int dms_frontend_get_tracker(dms_frontend* fe, int which, dms_odom** out);
int dms_frontend_get_filtered_depth(dms_frontend* fe, const uint16_t** d_depth, size_t* pitch_bytes);
int dms_frontend_get_fill_in(dms_frontend* fe, dms_model_view* out);
/* Context::lost() / lastFrameRecovery() are also written by the fern relocaliser (EF.cpp:276-296), which the caller owns. */
int dms_frontend_set_state(dms_frontend* fe, int lost, int lastFrameRecovery);

/* ---- stateless helpers (dense device images) */
/* FillIn::vertex / normal / image in one pass.  passthrough_geom = context.lost(), passthrough_img = lost ||
 * frameToFrameRGB (EF.cpp:734-742).  Any of the three outputs may be NULL. */
int dms_fill_in(const float* d_model_vertex4, const float* d_model_normal4, const uint8_t* d_model_rgba,
                const uint16_t* d_depth_filtered, const uint8_t* d_live_rgba, int rows, int cols, float cx, float cy,
                float fx, float fy, int passthrough_geom, int passthrough_img, float* d_out_vertex4,
                float* d_out_normal4, uint8_t* d_out_rgba, void* stream);
/* Fraction of the pixels of the 1/`sample` resized image whose three colour channels are all non-zero (EF.cpp:84-97,
 * sample = 20).  d_scratch = 4 bytes of device memory; synchronises the stream. */
int dms_dense_fraction(const uint8_t* d_rgba, int rows, int cols, int sample, unsigned int* d_scratch, float* fraction,
                       void* stream);
/* max(|t|, |rodrigues2(R)|) -> fusion weight, EF.cpp:251-266; poses row-major 4x4. */
float dms_velocity_weighting(const float currPose[16], const float lastPose[16], float weightMultiplier);

#ifdef __cplusplus
}
#endif
#endif
