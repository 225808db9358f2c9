/* dmsodom sequence runner -- C ABI.
 *
 * Open-loop tracking of a whole sequence with pose priors (SURVEY 8d configuration 2: every frame is registered
 * against the model view rendered at its prior, ElasticFusion.cpp:158-201 with `currPose = *inPose`), driven natively:
 * `n_trackers` trackers take turns on the frames, each on its own stream and host thread, their tracking kernels
 * sharing the GPU (dms_odom_set_grid / dms_odom_set_block), the SO(3) reference image of frame f handed on the device
 * from the tracker that processed frame f-1 (dms_odom_init_first_rgb_from).  Results are those of one tracker
 * processing the frames one after another.  No reference counterpart: the reference calls its single tracker from
 * the GL thread, one blocking call per frame.
 */
#ifndef DMSODOM_SEQUENCE_C_H
#define DMSODOM_SEQUENCE_C_H
#include "odom_c.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dms_seq_frame {
  const float* vertex4;       /* model view at the prior: RGBA32F vertex / normal maps, RGBA8 image (initICPModel, initRGBModel) */
  const float* normal4;
  const uint8_t* model_rgba;
  const uint16_t* depth;      /* live frame: u16 millimetres, RGBA8 (initICP, initRGB) */
  const uint8_t* live_rgba;
  float prior[16];            /* row-major pose prior = pose of the model view */
} dms_seq_frame;

typedef struct dms_seq_result {
  float pose[16];             /* refined pose, row-major */
  dms_odom_result r;
  float track_ms;             /* device time of the tracking kernel of this frame (options.timing != 0) */
  float host_ms;              /* host-visible latency of this frame: from its first init call to its result in hand (with several
                                 trackers this includes the waits for the image hand-off between them) */
} dms_seq_result;

typedef struct dms_seq_options {
  int n_trackers;             /* trackers taking turns (1 = strictly sequential) */
  int block_threads;          /* dms_odom_set_block, 0 = default */
  int grid;                   /* dms_odom_set_grid per tracker, 0 = all block slots / min(n_trackers, 8) */
  int wait_mode;              /* dms_odom_set_wait */
  int rgbOnly; float icpWeight; int pyramid, fastOdom, so3;
  float depthCutoff;
  int timing;                 /* dms_odom_set_timing of every tracker (0 = off: track_ms of the results stays 0) */
} dms_seq_options;
void dms_sequence_default_options(dms_seq_options* o);

typedef struct dms_sequence dms_sequence;
int dms_sequence_create(dms_sequence** out, int device, int width, int height, float cx, float cy, float fx, float fy,
                        const dms_seq_options* opt);
int dms_sequence_destroy(dms_sequence* s);
/* Tracks frames[0 .. n_frames).  first_prev_rgba = live image of the frame before frames[0] (the SO(3) reference of
 * the first frame).  host_buffers != 0: every pointer is (pinned) host memory and the upload is part of the call.
 * Blocks until all results are in `out`.  Returns 0 or the first error. */
int dms_sequence_run(dms_sequence* s, const dms_seq_frame* frames, int n_frames, const uint8_t* first_prev_rgba,
                     int host_buffers, dms_seq_result* out);
/* Host-side timeline of the last dms_sequence_run, microseconds since its entry: [0] the first tracking kernel was
 * enqueued, [1] the last result was collected, [2] the call returned.  Diagnostic for short runs and many ranks per host. */
int dms_sequence_host_timeline(const dms_sequence* s, double us[3]);
/* resident tracking kernels = min(n_trackers, block slots / grid) */
int dms_sequence_resident(const dms_sequence* s);

#ifdef __cplusplus
}
#endif
#endif
