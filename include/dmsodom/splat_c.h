/* dmsodom model-view synthesis -- C ABI.  SURVEY 8(f) "next" row N3.
 *
 * IndexMap::combinedPredict (elasticfusion/Core/src/IndexMap.cpp:253-368): the surfel map splatted into the camera at
 * `pose` -- the producer of the image / vertex / normal maps that initICPModel / initRGBModel consume
 * (ElasticFusion.cpp:165-181) and of the "old" maps of the loop-closure check (EF.cpp:402-416).  Shaders restated:
 * Shaders/splat.vert:42-95 (visibility tests, camera-frame position / normal, point size from the four extreme points),
 * Shaders/combo_splat.frag:36-66 (ray / surfel-plane intersection, in-disk test, outputs, gl_FragDepth),
 * Shaders/color.glsl:27-34 (decodeColor).  A CUDA rasteriser: one thread per surfel walks its point-sprite box and
 * z-tests with a 64-bit atomicMin on (24-bit depth, surfel index); a second pass resolves the winners.
 *
 * Rasterisation rules fixed here (the GL pipeline cannot run on this box, parity with it is unpinned): a sprite of
 * size s centred at window position (xw, yw) produces the fragments whose centres (i + 0.5, j + 0.5) lie within s / 2 of
 * the centre in x and y; sprites whose centre is outside the viewport are clipped; depth is a 24-bit fixed-point value,
 * GL_LESS, so of two fragments with equal depth the surfel drawn first (lower index) wins; fragments whose depth is
 * not in [0, 1] are clipped.  Every fp32 operation is evaluated left to right without fusion.
 */
#ifndef DMSODOM_SPLAT_C_H
#define DMSODOM_SPLAT_C_H
#include "odom_c.h"

#ifdef __cplusplus
extern "C" {
#endif

#define DMS_SURFEL_FLOATS 15 /* Vertex::SIZE / 4 (Shaders/Vertex.cpp:49-50): pos.xyz conf | colour(encoded rgb, -, initTime, -) | times[3] | normal.xyz radius */

typedef struct dms_splat_params {
  float t_inv[16];        /* pose.inverse(), row-major (IndexMap.cpp:297) */
  float cx, cy, fx, fy;
  int rows, cols;
  float maxDepth;         /* depthCutoff */
  float confThreshold;
  int time, timeIdx, maxTime, timeDelta;
  int active;             /* IndexMap::ACTIVE = 1, INACTIVE = 0 */
} dms_splat_params;

/* d_surfels: count x DMS_SURFEL_FLOATS floats.  d_zbuf: rows*cols u64 of device scratch.  Outputs are dense images
 * (RGBA8, RGBA32F, RGBA32F, R32UI; IndexMap.cpp:59-122), cleared to zero where nothing is drawn; any may be NULL. */
int dms_splat_predict(const float* d_surfels, int count, const dms_splat_params* p, unsigned long long* d_zbuf,
                      uint8_t* d_image_rgba, float* d_vertex4, float* d_normal4, uint32_t* d_time, void* stream);
/* pose (row-major 4x4) -> t_inv for dms_splat_params */
int dms_splat_invert_pose(const float pose[16], float t_inv[16]);

#ifdef __cplusplus
}
#endif
#endif
