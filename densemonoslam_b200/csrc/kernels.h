// Host-visible launchers of the dmsodom kernels (internal to the library).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "common.cuh"
#include "dmsodom/odom_c.h"

namespace dms {

// ---- pyramid.cu
cudaError_t launch_maps_pyramid(const Level* lv, bool use_prev, const float4* v4, const float4* n4, int pitch4,
                                const float* pose16_or_null, float* zdepth, float zcut, bool legacy, cudaStream_t s);
cudaError_t launch_populate(const Level* lv, int which, const float* zsrc, const uchar4* rgba, int pitch_rgba, bool legacy, cudaStream_t s);
cudaError_t launch_icp_depth(const Level* lv, const unsigned short* d_depth, int pitch_in, float cutoff, bool legacy, cudaStream_t s);

cudaError_t launch_depth_bilateral(const unsigned short* src, int sp, unsigned short* dst, int dp, int rows, int cols, float maxD, cudaStream_t s);
cudaError_t launch_depth_metric(const unsigned short* src, int sp, float* dst, int dp, int rows, int cols, float maxD, cudaStream_t s);

// ---- track.cu
static const int ACC_STRIDE = 16;   // unsigned long long words between accumulator slots (128 B: one L2 line per word)
// Accumulators of the persistent tracking kernel: ACC_SETS rotating sets of SET_SLOTS slots, one set per
// accumulate-and-solve phase (an SO(3) evaluation or a Gauss-Newton iteration).  Every CTA reads the totals of a
// phase itself, so a set is cleared two phases later, when every CTA is known to have read it.
static const int SET_SLOTS = 128;
static const int ACC_SETS = 4;
static const int STEP_SLOTS = 192;  // accumulator slots of the stand-alone step operators (behind the sets)
static const int ACC_SLOTS = ACC_SETS * SET_SLOTS + STEP_SLOTS;
static const int MAX_TRACE = 160;   // >= 3*50 interMap iterations

// Device-resident control block of the persistent tracking kernel.
static const int SYNC_WORDS = 32 * 4;   // barrier words, one 128-byte line each: [0] arrive counter (+ abort flag in word 1), [32],[64] work-list counters
static const int SYNC_LISTN = 32 * 1;   // sync[SYNC_LISTN + 32 * parity]: number of entries of the photometric segment list
// Words of the per-CTA block of constants the solver warp leaves in shared memory for the next phase.
enum { PUB_RCURR = 0, PUB_TCURR = 9, PUB_KRKINV = 12, PUB_KT = 21, PUB_SIGMA = 24, PUB_BRK = 25, PUB_SO3DONE = 26,
       PUB_SO3H = 27, PUB_SO3KINV = 36, PUB_SO3KRLR = 45, PUB_WORDS = 64 };
struct Ctrl {
  unsigned int aborted, pad0;   // 1: a grid barrier timed out (a CTA never arrived); the results are invalid
  int n_trace, n_dbg;
  // timing mode: ns per phase kind (Stopwatch keys) as seen by CTA 0; t_ser_*: of which in the serial section after the
  // barrier (accumulator read-out, solve, pose update), t_wait: of which waiting at grid barriers
  unsigned long long t_so3, t_phaseA, t_phaseB, t_mark;
  unsigned long long t_ser_so3, t_ser_a, t_ser_b, t_start, t_prologue, t_wait;
  dms_odom_result res;
  float out_R[9], out_t[3];
};

struct TrackParams {
  Level lv[3];
  int rgbOnly, pyramid, fastOdom, so3, interMap, icp, rgb;
  float icpWeight;
  float Rprev[9], tprev[3], Rprev_inv[9];
  float dist2Max, sine2Lim;  // squared-domain equivalents of distThres / angleThres (threshold_equiv)
  float distThres, angleThres;
  int iters[3];
  double Kd[3][9], Kinv[3][9];  // level intrinsics matrix and its inverse (RO.cpp:415-421, K.inverse())
  Ctrl* ctrl;
  unsigned int* sync;
  unsigned long long* acc;  // ACC_SETS sets, all zero between launches
  dms_iter_record* trace;
  int timing;
  unsigned int* seglist;    // segments with at least one photometric term in the current iteration (any order), lv[0] segments long
  unsigned long long* dbg;  // timeline mode: CTA 0 logs {tag, globaltimer} pairs here (MAX_DBG pairs)
};
static const int MAX_DBG = 1024;

// accessors of the opaque handle for the other translation units (ring.cu)
const Ctrl* odom_ctrl_device(const dms_odom* h);
int odom_device(const dms_odom* h);

void threshold_equiv(float distThres, float angleThres, float* dist2Max, float* sine2Lim);
int track_block_threads_max();
int track_block_threads_default();
int track_max_grid(int device, int tpb);  // co-resident CTAs of tpb threads for the cooperative launch
cudaError_t launch_track(const TrackParams& p, int grid, int tpb, bool legacy, cudaStream_t s);

// single-step operators (regular launches + finalize)
struct StepOut { float vals[28]; long long count, count2, sigma; int bad; int pad; };
cudaError_t launch_derivatives(const Level* lv, cudaStream_t s);
cudaError_t launch_icp_step(const Level& lv, const float Rcurr[9], const float tcurr[3], const float Rprev_inv[9],
                            const float tprev[3], float distThres, float angleThres, unsigned long long* acc, StepOut* d_out,
                            unsigned char* d_mask_or_null, int mask_pitch, bool legacy, cudaStream_t s);
cudaError_t launch_rgb_residual(const Level& lv, const float kt[3], const float krkinv[9], unsigned long long* acc,
                                StepOut* d_out, bool legacy, cudaStream_t s);
cudaError_t launch_rgb_step(const Level& lv, float sigma, unsigned long long* acc, StepOut* d_out, bool legacy, cudaStream_t s);
cudaError_t launch_so3_step(const Level& lv, const float H[9], const float kinv[9], const float krlr[9],
                            unsigned long long* acc, StepOut* d_out, bool legacy, cudaStream_t s);

}  // namespace dms
