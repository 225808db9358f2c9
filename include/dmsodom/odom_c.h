/*
 * dmsodom C ABI -- B200-native (sm_100a) dense RGB-D odometry.
 *
 * Drop-in boundary for the tracking hot path of robotvisionmu/DenseMonoSLAM
 * (ElasticFusion's RGBDOdometry).  Each entry point names the reference interface it
 * replaces; paths are relative to elasticfusion/Core/src/ in the reference tree:
 *   RO.h / RO.cpp = Utils/RGBDOdometry.{h,cpp},  cf.cuh = Cuda/cudafuncs.cuh,
 *   reduce.cu = Cuda/reduce.cu,  cf.cu = Cuda/cudafuncs.cu
 *
 * Conventions
 *  - plain C types only; every pointer named d_* is a DEVICE pointer on the handle's
 *    device, every pointer named h_* / unprefixed small array is HOST memory.
 *  - `stream` is a cudaStream_t passed as void* (0 = default stream).  Work is enqueued on
 *    it; entry points that return results to host memory synchronise that stream before
 *    returning (the reference synchronises the whole device at the same places).
 *  - all functions return 0 on success, a negative dms_status or a positive cudaError_t
 *    otherwise, and never call exit() (the reference's cudaSafeCall prints and exit(0)s,
 *    convenience.cuh:64-71).  dms_odom_last_error() gives the message.
 *  - one handle = one device = one host thread at a time.  Distinct handles are
 *    independent (one tracker per sensor / per GPU).
 *  - the GL textures of the reference (GPUTexture*, RO.h:47-66) are replaced by headless
 *    device buffers: RGBA8 colour, R16UI depth in millimetres, RGBA32F vertex / normal
 *    maps, row pitch in bytes.
 */
#ifndef DMSODOM_ODOM_C_H_
#define DMSODOM_ODOM_C_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dms_odom dms_odom;

enum dms_status {
  DMS_OK = 0,
  DMS_ERR_INVALID = -1,   /* bad argument */
  DMS_ERR_NOMEM = -2,     /* device allocation failed */
  DMS_ERR_DEVICE = -3,    /* device lacks a required capability (cooperative launch, sm_100) */
  DMS_ERR_TOO_LARGE = -4  /* size beyond a stated maximum */
};

/* Public state of RGBDOdometry after getIncrementalTransformation (RO.h:70-77) plus the
 * iteration counts actually run. */
typedef struct dms_odom_result {
  float lastICPError, lastICPCount, lastRGBError, lastRGBCount, lastSO3Error, lastSO3Count;
  double lastA[36]; /* row-major 6x6 */
  double lastb[6];
  int iters[3];     /* GN iterations run at pyramid level 0,1,2 */
  int so3_iters;    /* so3Step evaluations */
} dms_odom_result;

/* Stopwatch keys of the reference (RO.cpp:336,444,495,525), filled from CUDA events when
 * timing is enabled with dms_odom_set_timing(); milliseconds of the last track call. */
typedef struct dms_odom_timing {
  float track_ms;   /* whole getIncrementalTransformation on the device */
  float so3Step_ms, computeRgbResidual_ms, icpStep_ms, rgbStep_ms; /* split by the kernel's own phase timers */
  /* diagnostics: derivative prologue, and the part of each key spent in the serial section after a grid barrier
   * (accumulator read-out, solve, pose update: run by one warp of every thread block, the others wait) */
  float prologue_ms, serial_so3_ms, serial_a_ms, serial_b_ms;
  float barrier_wait_ms; /* of the keys above: thread block 0 waiting at grid barriers (includes load imbalance) */
} dms_odom_timing;

/* ---- lifetime: RGBDOdometry::RGBDOdometry / ~RGBDOdometry (RO.cpp:21-116).
 * NOTE the argument order cx, cy, fx, fy of the reference constructor (RO.h:34-36).
 * All device buffers are allocated here once and zero-filled. */
int dms_odom_create(dms_odom** out, int device, int width, int height, float cx, float cy, float fx, float fy,
                    float distThresh, float angleThresh);
int dms_odom_destroy(dms_odom* h);
const char* dms_odom_last_error(const dms_odom* h);
/* Limits the tracking kernel of this handle to `n_ctas` thread blocks (one per SM; 0 = all SMs).  Several handles
 * (sensors, or frames of one open-loop sequence) whose limits add up to the SM count run their tracking kernels
 * concurrently on one GPU, each on its own stream: the barrier / solve latency of one overlaps the pixel phases of
 * the others.  Results do not depend on the limit (the reduction is order-free). */
int dms_odom_set_grid(dms_odom* h, int n_ctas);
/* Threads per thread block of the tracking kernel (multiple of 32 in [64, 1024]; 0 = the default, 768).  With 512 or fewer, two or
 * more blocks (of different handles) are resident per SM, so one tracker's pixel phases fill the issue slots another
 * leaves idle at its barriers.  Resets the grid limit to "all" when it no longer fits.  Results do not depend on it. */
int dms_odom_set_block(dms_odom* h, int threads);
/* How dms_odom_track / dms_odom_track_collect wait for the result: 0 = spin on the stream (lowest latency, default),
 * 1 = sleep on a blocking-sync event, 2 = sleep for 85 % of the running average of the last waits, then poll and yield
 * (many trackers / ranks per host core without the wake-up latency of 1), 3 = as 2 but polling with 20 us sleeps instead
 * of yields (more waiting threads than host cores). */
int dms_odom_set_wait(dms_odom* h, int mode);
/* Arithmetic of the divisions and square roots that feed a truncation, a pixel rounding or a threshold test.
 * DMS_NUMERICS_IEEE (default): correctly rounded IEEE operations, the arithmetic the CPU oracle restates; integer outputs
 *   then differ from the reference BUILD in a bounded handful of places (+-1 on < 1 % of the pyramid pixels, single
 *   correspondences at a threshold), because that build approximates.
 * DMS_NUMERICS_LEGACY: the instructions nvcc emits for the reference under its own flags --ftz=true --prec-div=false
 *   --prec-sqrt=false (Core/src/CMakeLists.txt:48-49): div.full.ftz, rcp/sqrt/rsqrt.approx.ftz, x * 0.001f for x / 1000.f, at
 *   cudafuncs.cu:86-90,116,169,440-442,570-572, operators.cuh:77-86 and reduce.cu:270-271,294-302,569-576,806-815,970-971.
 *   Integer outputs (intensity and depth pyramids, derivative images, correspondence records, inlier counts) are then
 *   identical to the reference's own kernels (tests/test_gpu_vs_legacy.py).  Applies to the init_* calls, the track and
 *   the step operators issued after the switch. */
enum dms_numerics { DMS_NUMERICS_IEEE = 0, DMS_NUMERICS_LEGACY = 1 };
int dms_odom_set_numerics(dms_odom* h, int mode);
/* mode 0 = off (default), 1 = CUDA events around the kernel + per-phase counters of thread block 0 (a few clock reads per
 * phase), 2 = additionally the {tag, timestamp} timeline read by dms_odom_get_timeline (a memset per track), 3 = the CUDA
 * events only (track_ms; the other fields read 0).  Modes 1 and 2 (and dms_odom_set_trace) run the instrumented build of
 * the tracking kernel, which is about 3 % slower than the production build that modes 0 and 3 run. */
int dms_odom_set_timing(dms_odom* h, int mode);
int dms_odom_get_timing(const dms_odom* h, dms_odom_timing* out);

/* ---- pyramid construction.
 * initICP(GPUTexture* filteredDepth, depthCutoff, stream)           RO.cpp:118-142 */
int dms_odom_init_icp_depth(dms_odom* h, const uint16_t* d_depth, size_t pitch_bytes, float depthCutoff, void* stream);
/* initICP(GPUTexture* predictedVertices, predictedNormals, cutoff)  RO.cpp:144-167 */
int dms_odom_init_icp_maps(dms_odom* h, const float* d_vertices4, const float* d_normals4, size_t pitch_bytes,
                           float depthCutoff, void* stream);
/* initICPModel(vertices, normals, cutoff, Eigen::Matrix4f modelPose) RO.cpp:169-207;
 * pose is the row-major 4x4 camera->global transform. */
int dms_odom_init_icp_model(dms_odom* h, const float* d_vertices4, const float* d_normals4, size_t pitch_bytes,
                            float depthCutoff, const float pose_rowmajor[16], void* stream);
/* initRGB / initRGBModel / initFirstRGB(GPUTexture* rgb)             RO.cpp:238-266.
 * As in the reference, the depth pyramid paired with the image comes from the vertex map
 * loaded by the LAST init_icp_maps / init_icp_model call (RO.cpp:213,240,246). */
int dms_odom_init_rgb(dms_odom* h, const uint8_t* d_rgba, size_t pitch_bytes, void* stream);
int dms_odom_init_rgb_model(dms_odom* h, const uint8_t* d_rgba, size_t pitch_bytes, void* stream);
int dms_odom_init_first_rgb(dms_odom* h, const uint8_t* d_rgba, size_t pitch_bytes, void* stream);

/* initFirstRGB with the image `src` received in its most recent initRGB call, taken from src's pyramid on the device
 * (three small device-to-device copies instead of a conversion + two pyrDowns, and no re-upload from the host).  For
 * trackers that take turns on the frames of one sequence: tracker k gets frame f-1 from the tracker that processed it.
 * On the device the copy is ordered after src's initRGB and before src's next write to that memory by events.  On the
 * host the caller orders the calls: src's initRGB has RETURNED before this call, and this call has returned before src's
 * next initRGB / initFirstRGB* call starts (dms_sequence_run does both).  No reference counterpart (the reference has one
 * tracker per sensor). */
int dms_odom_init_first_rgb_from(dms_odom* h, const dms_odom* src, void* stream);

/* Same five calls with HOST buffers (dense rows unless pitch given); the copy to the device
 * is part of the call.  Used by callers without a device-side producer and by bench.py's
 * end-to-end number. */
int dms_odom_init_icp_depth_host(dms_odom* h, const uint16_t* h_depth, size_t pitch_bytes, float depthCutoff, void* stream);
int dms_odom_init_icp_maps_host(dms_odom* h, const float* h_vertices4, const float* h_normals4, size_t pitch_bytes,
                                float depthCutoff, void* stream);
int dms_odom_init_icp_model_host(dms_odom* h, const float* h_vertices4, const float* h_normals4, size_t pitch_bytes,
                                 float depthCutoff, const float pose_rowmajor[16], void* stream);
int dms_odom_init_rgb_host(dms_odom* h, const uint8_t* h_rgba, size_t pitch_bytes, void* stream);
int dms_odom_init_rgb_model_host(dms_odom* h, const uint8_t* h_rgba, size_t pitch_bytes, void* stream);
int dms_odom_init_first_rgb_host(dms_odom* h, const uint8_t* h_rgba, size_t pitch_bytes, void* stream);

/* ---- the tracker: getIncrementalTransformation(trans, rot, rgbOnly, icpWeight, pyramid,
 * fastOdom, so3, interMap, stream)  RO.cpp:268-605.  trans[3] / rot_rowmajor[9] carry the
 * previous pose in and the refined pose out.  The whole SO(3) + 3-level Gauss-Newton loop,
 * including the 6x6 solves and pose updates, runs on the device in one launch; `out`
 * (optional) receives the public fields.  Synchronises `stream`. */
int dms_odom_track(dms_odom* h, float trans[3], float rot_rowmajor[9], int rgbOnly, float icpWeight, int pyramid,
                   int fastOdom, int so3, int interMap, void* stream, dms_odom_result* out);
/* Asynchronous pair for pipelined callers: enqueue, then collect later. */
int dms_odom_track_async(dms_odom* h, const float trans[3], const float rot_rowmajor[9], int rgbOnly, float icpWeight,
                         int pyramid, int fastOdom, int so3, int interMap, void* stream);
int dms_odom_track_collect(dms_odom* h, float trans[3], float rot_rowmajor[9], void* stream, dms_odom_result* out);

/* getICPResidual(trans_curr, rot_curr, trans_prev, rot_prev, icpError, icpCount)  RO.cpp:612-643
 * (deterministic here; the reference's icpRes uses float atomicAdd, reduce.cu:482-540). */
int dms_odom_icp_residual(dms_odom* h, const float trans_curr[3], const float rot_curr[9], const float trans_prev[3],
                          const float rot_prev[9], float* icpError, int* icpCount, void* stream);
/* getCovariance() = lastA.lu().inverse()  RO.cpp:607-610 (host, 6x6 double) */
int dms_odom_get_covariance(const dms_odom* h, double cov_rowmajor[36]);

/* ---- buffer accessors: nextImg / lastImg / nextD / lastVmap (RO.h:79-90) and the rest of
 * the pyramid state for inspection.  Planar maps are 3*rows x cols floats (x-plane rows,
 * then y, then z; RO.cpp:97-101), pitched. */
enum dms_buffer {
  DMS_BUF_VMAP_CURR = 0, DMS_BUF_NMAP_CURR = 1, DMS_BUF_VMAP_G_PREV = 2, DMS_BUF_NMAP_G_PREV = 3,
  DMS_BUF_LAST_DEPTH = 4, DMS_BUF_NEXT_DEPTH = 5, DMS_BUF_LAST_IMAGE = 6, DMS_BUF_NEXT_IMAGE = 7,
  DMS_BUF_LAST_NEXT_IMAGE = 8, DMS_BUF_DIDX = 9, DMS_BUF_DIDY = 10, DMS_BUF_DEPTH_TMP = 11,
  DMS_BUF_CORRES = 12, /* 16-byte records {i16 u0, i16 v0, i16 diff, u8 valid, u8 0, f32 z, u32 0}: DataTerm, types.cuh:76-82, plus the model depth at (u0, v0) */
  DMS_BUF_NEXT_DEPTH_MASKED = 13 /* nextDepth with NaN where the iteration-invariant tests of reduce.cu:777-801 fail */
};
int dms_odom_get_buffer(dms_odom* h, int which, int level, void** d_ptr, size_t* pitch_bytes, int* rows, int* cols,
                        int* elem_bytes);
/* Copies a buffer to dense host memory (rows*cols*elem_bytes, planar maps 3 planes). */
int dms_odom_download(dms_odom* h, int which, int level, void* h_dst, size_t dst_bytes, void* stream);

/* Per-iteration trace of the last track call (for parity tests; enable with dms_odom_set_trace): record i holds
 * {level, iter, icp_inliers, rgb_count, rgb_sigma} and the 28 ICP / 28 RGB reduced values
 * (27 products in JtJJtrSE3 order, types.cuh:117-131, then residual) and the pose after it. */
typedef struct dms_iter_record {
  int32_t level, iter;
  int64_t icp_count, rgb_count, rgb_sigma;
  float icp[28], rgb[28];
  float R[9], t[3];
} dms_iter_record;
int dms_odom_set_trace(dms_odom* h, int enabled); /* off by default: recording costs a little solver time */
int dms_odom_get_trace(dms_odom* h, dms_iter_record* h_records, int max_records, int* n_records);
/* Timing mode 2 only: {tag, nanosecond timestamp} pairs logged by CTA 0 of the tracking kernel at its phase
 * boundaries during the last track call (diagnostic for tools/track_timing.py). */
int dms_odom_get_timeline(dms_odom* h, uint64_t* h_pairs, int max_pairs, int* n_pairs);

/* ---- single-step operators: the free functions of cudafuncs.cuh that RGBDOdometry (and
 * MutualInformation / KeyFrame / FGROdometry) call.  They act on the handle's pyramid
 * level `level` and return the reduced values to the host, like the reference wrappers do.
 * out28: 27 products + residual; counts are exact integers.
 * icpStep            reduce.cu:367-428 */
int dms_odom_icp_step(dms_odom* h, int level, const float Rcurr[9], const float tcurr[3], const float Rprev_inv[9],
                      const float tprev[3], float out28[28], int64_t* inliers, uint8_t* h_mask_or_null, void* stream);
/* computeRgbResidual reduce.cu:865-925 (writes DMS_BUF_CORRES of that level) */
int dms_odom_rgb_residual(dms_odom* h, int level, const float kt[3], const float krkinv[9], int64_t* count,
                          int64_t* sigma, void* stream);
/* rgbStep            reduce.cu:643-685 (reads DMS_BUF_CORRES) */
int dms_odom_rgb_step(dms_odom* h, int level, float sigma, float out28[28], int64_t* count, void* stream);
/* so3Step            reduce.cu:1054-1103 (lastNextImage vs nextImage of `level`) */
int dms_odom_so3_step(dms_odom* h, int level, const float imageBasis[9], const float kinv[9], const float krlr[9],
                      float out10[10], int64_t* count, void* stream);
/* computeDerivativeImages cf.cu:697-725 on nextImage of every level (also refreshes the
 * iteration-invariant photometric candidate mask used by rgb_residual). */
int dms_odom_compute_derivatives(dms_odom* h, void* stream);

/* ---- depth pre-filter: the step immediately before initICP(depth).  Stateless, device pointers.
 * ElasticFusion::filterDepth  (ElasticFusion.cpp:759-768, Shaders/depth_bilateral.frag:31-75): 13x13 bilateral filter on
 *   raw u16 millimetres, values < 300 or > maxDepth*1000 -> 0, uint(round(.)).  exp() is evaluated in explicit fp32 steps so
 *   the result is reproducible (GLSL's exp is implementation defined; parity with the GL shader is therefore unpinned).
 * ElasticFusion::metriciseDepth (ElasticFusion.cpp:748-757, Shaders/depth_metric.frag:29-39): millimetres -> metres with
 *   the same validity test. */
int dms_depth_bilateral(const uint16_t* d_raw, size_t pitch_in, int rows, int cols, float maxDepth, uint16_t* d_filtered,
                        size_t pitch_out, void* stream);
int dms_depth_metric(const uint16_t* d_raw, size_t pitch_in, int rows, int cols, float maxDepth, float* d_metres,
                     size_t pitch_out, void* stream);

/* ---- allocation-free variant of the ICP step.  NO COUNTERPART IN THE REFERENCE.
 *
 * For callers that are not allowed to allocate device memory: a tracker hosted inside another
 * engine's frame loop with a memory budget fixed at start-up, a real-time thread that cannot take
 * cudaMalloc's implicit device-wide synchronisation, or a plug-in that only ever works on buffers it
 * is handed.  It needs no handle: the caller passes the four planar maps (3*rows x cols floats each,
 * x plane rows first, row pitch `pitch_bytes` a multiple of 4), the level intrinsics, the
 * poses as in dms_odom_icp_step, and a DEVICE record `d_out` that receives the same 27 products +
 * residual and the inlier count.  The work is enqueued on `stream`; nothing is copied to the host.
 *
 * The reduction's scratch (the exact-accumulator lines, 24 KB) is a static device array inside
 * the library, present on every GPU that loads it until the process exits.  Consequences:
 *  - the call performs no cudaMalloc / cudaFree and uses no allocating library routine;
 *  - images above DMS_NOALLOC_MAX_PIXELS pixels, or with more than DMS_NOALLOC_MAX_DIM rows or columns (the same
 *    per-dimension limit as dms_odom_create), or whose maps span more than 4 GB (3 * rows * pitch_bytes: the kernels
 *    address a map's planes with 32-bit byte offsets), are refused with DMS_ERR_TOO_LARGE (and a line on stderr);
 *  - all calls on one GPU share the array, which is cleared at the start of every call: two calls
 *    whose GPU work overlaps (different streams, or different host threads) race on it and give
 *    wrong sums -- serialise them;
 *  - the first call on each device looks up the array's device address (cudaGetSymbolAddress,
 *    with that device current); make that first call before starting any stream capture.
 * Results are bit-identical to dms_odom_icp_step on the same inputs. */
#define DMS_NOALLOC_MAX_PIXELS (1 << 24)
#define DMS_NOALLOC_MAX_DIM 16384
typedef struct dms_step_sums {
  float vals[28];     /* 27 products in JtJJtrSE3 order (types.cuh:117-131), then the squared residual */
  int64_t count;      /* inliers */
  int64_t reserved0, reserved1;
  int32_t bad;        /* 1 if a partial sum was NaN / Inf / out of accumulator range: vals are NaN */
  int32_t reserved2;
} dms_step_sums;
int dms_icp_step_noalloc(const float* d_vmap_curr, const float* d_nmap_curr, const float* d_vmap_g_prev,
                         const float* d_nmap_g_prev, size_t pitch_bytes, int rows, int cols, float fx, float fy,
                         float cx, float cy, const float Rcurr[9], const float tcurr[3], const float Rprev_inv[9],
                         const float tprev[3], float distThresh, float angleThresh, dms_step_sums* d_out, void* stream);

/* ---- library info */
const char* dms_version(void);
int dms_device_count(void);
/* Host-side helper, exposed for tests: the squared-domain equivalents of the two ICP acceptance tests of reduce.cu:301
 * (sqrtf(d2) <= distThresh  <=>  d2 <= *dist2Max;  sqrtf(s2) < angleThresh  <=>  s2 < *sine2Lim), exact for every float
 * threshold including infinities.  Needs no GPU. */
void dms_threshold_equiv(float distThresh, float angleThresh, float* dist2Max, float* sine2Lim);
/* bytes copied device->host by one dms_odom_track call (the control block with pose + public fields) */
size_t dms_track_d2h_bytes(void);

#ifdef __cplusplus
}
#endif
#endif /* DMSODOM_ODOM_C_H_ */
