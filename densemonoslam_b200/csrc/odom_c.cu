// C-ABI of dmsodom (include/dmsodom/odom_c.h): handle, device buffers, error codes.
// Host code only; every kernel lives in pyramid.cu / track.cu.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sched.h>
#include <chrono>
#include <thread>
#include <new>
#include <string>
#include "kernels.h"

using namespace dms;

struct dms_odom {
  int device;
  int width, height;
  Intr K;
  float distThres, angleThres;
  Level lv[3];
  char* slab;            // one allocation for every device buffer
  size_t slab_bytes;
  float* zdepth;         // level-0 depth of the last loaded vertex map (stands in for vmaps_tmp, RO.cpp:104,213)
  Ctrl* ctrl;
  unsigned int* sync;
  unsigned long long* acc;       // ACC_SETS sets of the tracking kernel, then the step operators' block
  unsigned long long* step_acc;  // = acc + ACC_SETS * SET_SLOTS * ACC_STRIDE
  dms_iter_record* trace;
  StepOut* step_out;
  unsigned long long* dbg;
  unsigned int* seglist;
  unsigned char* mask;   // level-0 sized scratch for dms_odom_icp_step's inlier mask
  // staging for the *_host entry points (allocated on first use)
  float4 *stage_v4, *stage_n4; uchar4* stage_rgba; unsigned short* stage_depth;
  // pinned host mirror of Ctrl / StepOut
  Ctrl* h_ctrl; StepOut* h_step;
  int grid, grid_max;    // CTAs of the cooperative tracking kernel in use / co-resident maximum
  int tpb;               // threads per CTA of the tracking kernel
  int timing, tracing;
  int legacy;            // dms_odom_set_numerics: 1 = the reference build's approximate division / square roots
  dms_odom_result last;  // public fields of the last track call
  dms_odom_timing last_timing;
  cudaEvent_t ev0, ev1;
  cudaEvent_t ev_took;   // recorded after this handle copied another tracker's pyramid (init_first_rgb_from)
  cudaEvent_t wait_reader;  // ev_took of the tracker that is reading rgb_pyr[]: the next write to an image pyramid waits for it
  cudaEvent_t ev_rgb;    // recorded after every initRGB: the intensity pyramid in rgb_pyr[] is complete (init_first_rgb_from waits on it)
  unsigned char* rgb_pyr[3];  // memory that received the image of the last initRGB (stable across the nextImage / lastNextImage pointer swap)
  cudaEvent_t ev_done;   // blocking-sync event (dms_odom_set_wait): the host thread sleeps instead of spinning
  cudaStream_t track_stream;  // stream of the pending track_async
  int in_flight;         // a track_async has not been collected yet
  int blocking_wait;     // 0 spin, 1 blocking-sync event, 2 sleep for most of the expected wait, then poll and yield, 3 same but poll with 20 us sleeps
  double wait_ema_us;    // running estimate of the wait (mode 2)
  std::string err;
};

namespace {

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int fail(dms_odom* h, int code, const char* what) {
  if (h) {
    char buf[256];
    if (code > 0) snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString((cudaError_t)code));
    else snprintf(buf, sizeof buf, "%s", what);
    h->err = buf;
  }
  return code;
}
#define CK(call)                                                          \
  do {                                                                    \
    cudaError_t e__ = (call);                                             \
    if (e__ != cudaSuccess) return fail(h, (int)e__, #call);              \
  } while (0)

struct Bind {  // makes the handle's device current for the duration of a call
  int prev; bool ok;
  explicit Bind(int dev) : prev(-1), ok(true) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
  }
  ~Bind() { /* keep the handle's device current: callers own one device per thread */ }
};

template <class T>
void mat3_inv_host(const T* m, T* o) {  // same operation order as the oracle's (Eigen .inverse(), RO.cpp:391)
  const T c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
  const T det = m[0] * c00 + m[1] * c01 + m[2] * c02;
  const T id = T(1) / det;
  o[0] = c00 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
  o[3] = c01 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
  o[6] = c02 * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

int populate(dms_odom* h, const uint8_t* d_rgba, size_t pitch, int which, cudaStream_t s) {
  // which: 0 = last (initRGBModel), 1 = next (initRGB), 2 = lastNext image only (initFirstRGB)
  if (which != 0 && h->wait_reader) {  // another tracker may still be copying the image this call overwrites
    CK(cudaStreamWaitEvent(s, h->wait_reader, 0));
    h->wait_reader = nullptr;
  }
  CK(launch_populate(h->lv, which, h->zdepth, reinterpret_cast<const uchar4*>(d_rgba), (int)pitch, h->legacy != 0, s));
  return DMS_OK;
}

int ensure_stage(dms_odom* h) {
  if (h->stage_v4) return DMS_OK;
  const size_t n = (size_t)h->width * h->height;
  CK(cudaMalloc(&h->stage_v4, n * 16));
  CK(cudaMalloc(&h->stage_n4, n * 16));
  CK(cudaMalloc(&h->stage_rgba, n * 4));
  CK(cudaMalloc(&h->stage_depth, n * 2));
  return DMS_OK;
}

}  // namespace

namespace dms {
const Ctrl* odom_ctrl_device(const dms_odom* h) { return h->ctrl; }
int odom_device(const dms_odom* h) { return h->device; }
}  // namespace dms

extern "C" {

const char* dms_version(void) { return "dmsodom 0.1 (sm_100a)"; }

int dms_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

size_t dms_track_d2h_bytes(void) { return sizeof(Ctrl); }

void dms_threshold_equiv(float distThresh, float angleThresh, float* dist2Max, float* sine2Lim) {
  float d = 0.f, a = 0.f;
  threshold_equiv(distThresh, angleThresh, &d, &a);
  if (dist2Max) *dist2Max = d;
  if (sine2Lim) *sine2Lim = a;
}

const char* dms_odom_last_error(const dms_odom* h) { return h ? h->err.c_str() : "null handle"; }

int dms_odom_create(dms_odom** out, int device, int width, int height, float cx, float cy, float fx, float fy, float distThresh,
                    float angleThresh) {
  if (!out || width < 16 || height < 16 || width > 16384 || height > 16384) return DMS_ERR_INVALID;
  *out = nullptr;
  dms_odom* h = new (std::nothrow) dms_odom();
  if (!h) return DMS_ERR_NOMEM;
  h->device = device; h->width = width; h->height = height;
  h->K.fx = fx; h->K.fy = fy; h->K.cx = cx; h->K.cy = cy;
  h->distThres = distThresh; h->angleThres = angleThresh;
  h->slab = nullptr; h->stage_v4 = h->stage_n4 = nullptr; h->stage_rgba = nullptr; h->stage_depth = nullptr;
  h->h_ctrl = nullptr; h->h_step = nullptr; h->timing = 0; h->tracing = 0; h->ev0 = h->ev1 = nullptr;
  h->legacy = 0; h->in_flight = 0; h->track_stream = nullptr; h->blocking_wait = 0; h->wait_ema_us = 0.0; h->wait_reader = nullptr;
  h->ev_rgb = h->ev_took = h->ev_done = nullptr; h->rgb_pyr[0] = h->rgb_pyr[1] = h->rgb_pyr[2] = nullptr;
  memset(&h->last, 0, sizeof h->last); memset(&h->last_timing, 0, sizeof h->last_timing);
  // RO.cpp:26-31 initial public state
  h->last.lastICPCount = h->last.lastRGBCount = h->last.lastSO3Count = (float)(width * height);
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) { delete h; return (int)e; }
  int coop = 0, major = 0;
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device);
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
  if (!coop || major < 10) { delete h; return DMS_ERR_DEVICE; }

  // ---- carve the slab (RO.cpp:44-105 buffer set, minus pointClouds / 16-byte corresImg)
  size_t off = 0;
  struct Req { void** p; size_t bytes; };
  Req reqs[64]; int nreq = 0;
  auto want = [&](void** p, size_t bytes) { reqs[nreq].p = p; reqs[nreq].bytes = bytes; ++nreq; };
  for (int i = 0; i < 3; ++i) {
    Level& L = h->lv[i];
    L.rows = height >> i; L.cols = width >> i;
    const int div = 1 << i;
    L.intr.fx = fx / div; L.intr.fy = fy / div; L.intr.cx = cx / div; L.intr.cy = cy / div;  // types.cuh:110-114
    L.pitch_f32 = (int)align_up((size_t)L.cols * 4, 256);
    L.pitch_u8 = (int)align_up((size_t)L.cols, 256);
    L.pitch_s16 = (int)align_up((size_t)L.cols * 2, 256);
    L.pitch_u16 = L.pitch_s16;
    L.pitch_corr = (int)align_up((size_t)L.cols * 16, 256);
    const size_t map = (size_t)3 * L.rows * L.pitch_f32, img = (size_t)L.rows * L.pitch_u8;
    want((void**)&L.vmap_curr, map); want((void**)&L.nmap_curr, map); want((void**)&L.vmap_prev, map); want((void**)&L.nmap_prev, map);
    want((void**)&L.lastDepth, (size_t)L.rows * L.pitch_f32); want((void**)&L.nextDepth, (size_t)L.rows * L.pitch_f32);
    want((void**)&L.nextDepthMasked, (size_t)L.rows * L.pitch_f32);
    want((void**)&L.lastImage, img); want((void**)&L.nextImage, img); want((void**)&L.lastNextImage, img);
    want((void**)&L.dIdx, (size_t)L.rows * L.pitch_s16); want((void**)&L.dIdy, (size_t)L.rows * L.pitch_s16);
    want((void**)&L.depth_tmp, (size_t)L.rows * L.pitch_u16);
    want((void**)&L.corr, (size_t)L.rows * L.pitch_corr);
  }
  want((void**)&h->zdepth, (size_t)h->lv[0].rows * h->lv[0].pitch_f32);
  want((void**)&h->ctrl, sizeof(Ctrl));
  want((void**)&h->sync, sizeof(unsigned int) * SYNC_WORDS);
  want((void**)&h->acc, (size_t)ACC_SLOTS * ACC_STRIDE * sizeof(unsigned long long));
  want((void**)&h->trace, sizeof(dms_iter_record) * MAX_TRACE);
  want((void**)&h->step_out, sizeof(StepOut));
  want((void**)&h->dbg, sizeof(unsigned long long) * 2 * MAX_DBG);
  want((void**)&h->seglist, sizeof(unsigned int) * (size_t)h->lv[0].rows * ((h->lv[0].cols + 31) / 32));
  want((void**)&h->mask, (size_t)h->lv[0].rows * h->lv[0].pitch_u8);
  for (int r = 0; r < nreq; ++r) off += align_up(reqs[r].bytes, 256);
  h->slab_bytes = off;
  e = cudaMalloc(&h->slab, h->slab_bytes);
  if (e != cudaSuccess) { delete h; return DMS_ERR_NOMEM; }
  e = cudaMemset(h->slab, 0, h->slab_bytes);  // the reference leaves buffers uninitialised
  off = 0;
  for (int r = 0; r < nreq; ++r) { *reqs[r].p = h->slab + off; off += align_up(reqs[r].bytes, 256); }
  h->step_acc = h->acc + (size_t)ACC_SETS * SET_SLOTS * ACC_STRIDE;
  if (e == cudaSuccess) e = cudaMallocHost(&h->h_ctrl, sizeof(Ctrl));
  if (e == cudaSuccess) e = cudaMallocHost(&h->h_step, sizeof(StepOut));
  if (e == cudaSuccess) {
    // the public fields live in the device control block between calls (a call that skips SO(3) leaves
    // lastSO3Error/Count untouched, like the reference's members); seed them as RO.cpp:26-31 does
    memset(h->h_ctrl, 0, sizeof(Ctrl));
    h->h_ctrl->res = h->last;
    e = cudaMemcpy(h->ctrl, h->h_ctrl, sizeof(Ctrl), cudaMemcpyHostToDevice);
  }
  if (e == cudaSuccess) e = cudaEventCreate(&h->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&h->ev1);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_rgb, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_took, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_done, cudaEventBlockingSync | cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { dms_odom_destroy(h); return (int)e; }
  h->tpb = track_block_threads_default();
  if (const char* g = getenv("DMS_TRACK_THREADS")) {  // tuning knob: smaller CTAs (several co-resident per SM)
    const int v = atoi(g);
    if (v >= 64 && v <= track_block_threads_max() && (v & 31) == 0) h->tpb = v;
  }
  h->grid = track_max_grid(device, h->tpb);
  if (h->grid <= 0) { dms_odom_destroy(h); return DMS_ERR_DEVICE; }
  h->grid_max = h->grid;
  if (const char* g = getenv("DMS_TRACK_GRID")) {  // tuning knob: fewer co-resident CTAs than the maximum
    const int v = atoi(g);
    if (v > 0 && v < h->grid) h->grid = v;
  }
  *out = h;
  return DMS_OK;
}

int dms_odom_destroy(dms_odom* h) {
  if (!h) return DMS_OK;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  if (h->slab) cudaFree(h->slab);
  if (h->stage_v4) cudaFree(h->stage_v4);
  if (h->stage_n4) cudaFree(h->stage_n4);
  if (h->stage_rgba) cudaFree(h->stage_rgba);
  if (h->stage_depth) cudaFree(h->stage_depth);
  if (h->h_ctrl) cudaFreeHost(h->h_ctrl);
  if (h->h_step) cudaFreeHost(h->h_step);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->ev_done) cudaEventDestroy(h->ev_done);
  if (h->ev_rgb) cudaEventDestroy(h->ev_rgb);
  if (h->ev_took) cudaEventDestroy(h->ev_took);
  delete h;
  return DMS_OK;
}

int dms_odom_set_grid(dms_odom* h, int n_ctas) {
  if (!h || n_ctas < 0) return DMS_ERR_INVALID;
  h->grid = (n_ctas == 0 || n_ctas > h->grid_max) ? h->grid_max : (n_ctas < 2 ? 2 : n_ctas);
  return DMS_OK;
}
int dms_odom_set_block(dms_odom* h, int threads) {
  if (!h || threads < 0) return DMS_ERR_INVALID;
  const int tpb = threads == 0 ? track_block_threads_default() : threads;
  cudaSetDevice(h->device);
  const int g = track_max_grid(h->device, tpb);
  if (g <= 0) return DMS_ERR_INVALID;
  const bool all = h->grid == h->grid_max;
  h->tpb = tpb; h->grid_max = g;
  if (all || h->grid > g) h->grid = g;
  return DMS_OK;
}
int dms_odom_set_wait(dms_odom* h, int mode) { if (!h || mode < 0 || mode > 3) return DMS_ERR_INVALID; h->blocking_wait = mode; h->wait_ema_us = 0.0; return DMS_OK; }
int dms_odom_set_numerics(dms_odom* h, int mode) { if (!h || mode < 0 || mode > 1) return DMS_ERR_INVALID; h->legacy = mode; return DMS_OK; }
int dms_odom_set_timing(dms_odom* h, int mode) { if (!h || mode < 0 || mode > 3) return DMS_ERR_INVALID; h->timing = mode; return DMS_OK; }
int dms_odom_get_timing(const dms_odom* h, dms_odom_timing* out) { if (!h || !out) return DMS_ERR_INVALID; *out = h->last_timing; return DMS_OK; }

// ------------------------------------------------------------------ pyramid construction
int dms_odom_init_icp_depth(dms_odom* h, const uint16_t* d_depth, size_t pitch, float depthCutoff, void* stream) {
  if (!h || !d_depth) return fail(h, DMS_ERR_INVALID, "init_icp_depth: null argument");
  Bind b(h->device);
  cudaStream_t s = (cudaStream_t)stream;
  if (pitch == 0) pitch = (size_t)h->width * 2;
  CK(launch_icp_depth(h->lv, d_depth, (int)pitch, depthCutoff, h->legacy != 0, s));
  return DMS_OK;
}

int dms_odom_init_icp_maps(dms_odom* h, const float* d_v4, const float* d_n4, size_t pitch, float depthCutoff, void* stream) {
  (void)depthCutoff;  // accepted and unused, as in the reference (RO.cpp:144-167)
  if (!h || !d_v4 || !d_n4) return fail(h, DMS_ERR_INVALID, "init_icp_maps: null argument");
  Bind b(h->device);
  if (pitch == 0) pitch = (size_t)h->width * 16;
  CK(launch_maps_pyramid(h->lv, false, reinterpret_cast<const float4*>(d_v4), reinterpret_cast<const float4*>(d_n4), (int)pitch, nullptr,
                         h->zdepth, 6.0f, h->legacy != 0, (cudaStream_t)stream));
  return DMS_OK;
}

int dms_odom_init_icp_model(dms_odom* h, const float* d_v4, const float* d_n4, size_t pitch, float depthCutoff, const float pose[16],
                            void* stream) {
  (void)depthCutoff;  // RO.cpp:169-207 ignores it too
  if (!h || !d_v4 || !d_n4 || !pose) return fail(h, DMS_ERR_INVALID, "init_icp_model: null argument");
  Bind b(h->device);
  if (pitch == 0) pitch = (size_t)h->width * 16;
  CK(launch_maps_pyramid(h->lv, true, reinterpret_cast<const float4*>(d_v4), reinterpret_cast<const float4*>(d_n4), (int)pitch, pose,
                         h->zdepth, 6.0f, h->legacy != 0, (cudaStream_t)stream));
  return DMS_OK;
}

int dms_odom_init_rgb(dms_odom* h, const uint8_t* d_rgba, size_t pitch, void* stream) {
  if (!h || !d_rgba) return fail(h, DMS_ERR_INVALID, "init_rgb: null argument");
  Bind b(h->device);
  const int rc = populate(h, d_rgba, pitch ? pitch : (size_t)h->width * 4, 1, (cudaStream_t)stream);
  if (rc != DMS_OK) return rc;
  for (int i = 0; i < 3; ++i) h->rgb_pyr[i] = h->lv[i].nextImage;
  CK(cudaEventRecord(h->ev_rgb, (cudaStream_t)stream));
  return DMS_OK;
}
int dms_odom_init_rgb_model(dms_odom* h, const uint8_t* d_rgba, size_t pitch, void* stream) {
  if (!h || !d_rgba) return fail(h, DMS_ERR_INVALID, "init_rgb_model: null argument");
  Bind b(h->device);
  return populate(h, d_rgba, pitch ? pitch : (size_t)h->width * 4, 0, (cudaStream_t)stream);
}
int dms_odom_init_first_rgb(dms_odom* h, const uint8_t* d_rgba, size_t pitch, void* stream) {
  if (!h || !d_rgba) return fail(h, DMS_ERR_INVALID, "init_first_rgb: null argument");
  Bind b(h->device);
  return populate(h, d_rgba, pitch ? pitch : (size_t)h->width * 4, 2, (cudaStream_t)stream);
}

int dms_odom_init_first_rgb_from(dms_odom* h, const dms_odom* src, void* stream) {
  if (!h || !src || src == h) return fail(h, DMS_ERR_INVALID, "init_first_rgb_from: bad argument");
  if (src->device != h->device || src->width != h->width || src->height != h->height || !src->rgb_pyr[0])
    return fail(h, DMS_ERR_INVALID, "init_first_rgb_from: source has another size / device or no image yet");
  Bind b(h->device);
  cudaStream_t s = (cudaStream_t)stream;
  if (h->wait_reader) { CK(cudaStreamWaitEvent(s, h->wait_reader, 0)); h->wait_reader = nullptr; }
  CK(cudaStreamWaitEvent(s, src->ev_rgb, 0));
  for (int i = 0; i < 3; ++i) {
    const Level& L = h->lv[i];
    CK(cudaMemcpy2DAsync(L.lastNextImage, L.pitch_u8, src->rgb_pyr[i], src->lv[i].pitch_u8, (size_t)L.cols, (size_t)L.rows, cudaMemcpyDeviceToDevice, s));
  }
  // src's next write to that memory (its own initFirstRGB / initRGB, possibly on another stream) has to wait for these copies
  CK(cudaEventRecord(h->ev_took, s));
  const_cast<dms_odom*>(src)->wait_reader = h->ev_took;
  return DMS_OK;
}

// ---- host-buffer variants
int dms_odom_init_icp_depth_host(dms_odom* h, const uint16_t* src, size_t pitch, float cutoff, void* stream) {
  if (!h || !src) return fail(h, DMS_ERR_INVALID, "init_icp_depth_host: null argument");
  Bind b(h->device);
  int rc = ensure_stage(h); if (rc) return rc;
  const size_t w = (size_t)h->width * 2;
  CK(cudaMemcpy2DAsync(h->stage_depth, w, src, pitch ? pitch : w, w, h->height, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return dms_odom_init_icp_depth(h, h->stage_depth, w, cutoff, stream);
}
static int stage_maps(dms_odom* h, const float* v, const float* n, size_t pitch, cudaStream_t s) {
  int rc = ensure_stage(h); if (rc) return rc;
  const size_t w = (size_t)h->width * 16;
  CK(cudaMemcpy2DAsync(h->stage_v4, w, v, pitch ? pitch : w, w, h->height, cudaMemcpyHostToDevice, s));
  CK(cudaMemcpy2DAsync(h->stage_n4, w, n, pitch ? pitch : w, w, h->height, cudaMemcpyHostToDevice, s));
  return DMS_OK;
}
int dms_odom_init_icp_maps_host(dms_odom* h, const float* v, const float* n, size_t pitch, float cutoff, void* stream) {
  if (!h || !v || !n) return fail(h, DMS_ERR_INVALID, "init_icp_maps_host: null argument");
  Bind b(h->device);
  int rc = stage_maps(h, v, n, pitch, (cudaStream_t)stream); if (rc) return rc;
  return dms_odom_init_icp_maps(h, (const float*)h->stage_v4, (const float*)h->stage_n4, 0, cutoff, stream);
}
int dms_odom_init_icp_model_host(dms_odom* h, const float* v, const float* n, size_t pitch, float cutoff, const float pose[16], void* stream) {
  if (!h || !v || !n || !pose) return fail(h, DMS_ERR_INVALID, "init_icp_model_host: null argument");
  Bind b(h->device);
  int rc = stage_maps(h, v, n, pitch, (cudaStream_t)stream); if (rc) return rc;
  return dms_odom_init_icp_model(h, (const float*)h->stage_v4, (const float*)h->stage_n4, 0, cutoff, pose, stream);
}
static int stage_rgba(dms_odom* h, const uint8_t* src, size_t pitch, cudaStream_t s) {
  int rc = ensure_stage(h); if (rc) return rc;
  const size_t w = (size_t)h->width * 4;
  CK(cudaMemcpy2DAsync(h->stage_rgba, w, src, pitch ? pitch : w, w, h->height, cudaMemcpyHostToDevice, s));
  return DMS_OK;
}
int dms_odom_init_rgb_host(dms_odom* h, const uint8_t* src, size_t pitch, void* stream) {
  if (!h || !src) return fail(h, DMS_ERR_INVALID, "init_rgb_host: null argument");
  Bind b(h->device);
  int rc = stage_rgba(h, src, pitch, (cudaStream_t)stream); if (rc) return rc;
  return dms_odom_init_rgb(h, (const uint8_t*)h->stage_rgba, 0, stream);
}
int dms_odom_init_rgb_model_host(dms_odom* h, const uint8_t* src, size_t pitch, void* stream) {
  if (!h || !src) return fail(h, DMS_ERR_INVALID, "init_rgb_model_host: null argument");
  Bind b(h->device);
  int rc = stage_rgba(h, src, pitch, (cudaStream_t)stream); if (rc) return rc;
  return dms_odom_init_rgb_model(h, (const uint8_t*)h->stage_rgba, 0, stream);
}
int dms_odom_init_first_rgb_host(dms_odom* h, const uint8_t* src, size_t pitch, void* stream) {
  if (!h || !src) return fail(h, DMS_ERR_INVALID, "init_first_rgb_host: null argument");
  Bind b(h->device);
  int rc = stage_rgba(h, src, pitch, (cudaStream_t)stream); if (rc) return rc;
  return dms_odom_init_first_rgb(h, (const uint8_t*)h->stage_rgba, 0, stream);
}

// ------------------------------------------------------------------ tracking
int dms_odom_track_async(dms_odom* h, const float trans[3], const float rot[9], int rgbOnly, float icpWeight, int pyramid, int fastOdom,
                         int so3, int interMap, void* stream) {
  if (!h || !trans || !rot) return fail(h, DMS_ERR_INVALID, "track: null argument");
  if (h->in_flight) return fail(h, DMS_ERR_INVALID, "track_async: the previous track has not been collected");
  Bind b(h->device);
  cudaStream_t s = (cudaStream_t)stream;
  TrackParams P;
  memset(&P, 0, sizeof P);
  for (int i = 0; i < 3; ++i) P.lv[i] = h->lv[i];
  P.rgbOnly = rgbOnly ? 1 : 0; P.pyramid = pyramid ? 1 : 0; P.fastOdom = fastOdom ? 1 : 0; P.so3 = so3 ? 1 : 0; P.interMap = interMap ? 1 : 0;
  P.icp = (!rgbOnly && icpWeight > 0) ? 1 : 0;   // RO.cpp:278-279
  P.rgb = (rgbOnly || icpWeight < 100) ? 1 : 0;
  P.icpWeight = icpWeight;
  memcpy(P.Rprev, rot, sizeof P.Rprev); memcpy(P.tprev, trans, sizeof P.tprev);
  mat3_inv_host<float>(P.Rprev, P.Rprev_inv);
  threshold_equiv(h->distThres, h->angleThres, &P.dist2Max, &P.sine2Lim);
  P.distThres = h->distThres; P.angleThres = h->angleThres;
  P.iters[0] = interMap ? 50 : fastOdom ? 3 : 10;  // RO.cpp:387-389
  P.iters[1] = interMap ? 50 : pyramid ? 5 : 0;
  P.iters[2] = interMap ? 50 : pyramid ? 4 : 0;
  for (int i = 0; i < 3; ++i) {
    const Intr& k = h->lv[i].intr;
    const double Kd[9] = {k.fx, 0, k.cx, 0, k.fy, k.cy, 0, 0, 1};
    double Ki[9];
    mat3_inv_host<double>(Kd, Ki);
    memcpy(P.Kd[i], Kd, sizeof Kd); memcpy(P.Kinv[i], Ki, sizeof Ki);
  }
  P.ctrl = h->ctrl; P.sync = h->sync; P.acc = h->acc; P.trace = h->tracing ? h->trace : nullptr; P.timing = (h->timing == 1 || h->timing == 2) ? h->timing : 0;  // mode 3: events only, production kernel
  P.dbg = h->timing >= 2 ? h->dbg : nullptr;  // the timeline probes run only on request (set_timing(h, 2))
  P.seglist = h->seglist;
  if (h->timing >= 2) CK(cudaMemsetAsync(h->dbg, 0, sizeof(unsigned long long) * 2 * MAX_DBG, s));
  if (h->timing) CK(cudaEventRecord(h->ev0, s));
  cudaError_t e = launch_track(P, h->grid, h->tpb, h->legacy != 0, s);
  if (e != cudaSuccess) return fail(h, (int)e, "cooperative launch of k_track");
  if (h->timing) CK(cudaEventRecord(h->ev1, s));
  CK(cudaMemcpyAsync(h->h_ctrl, h->ctrl, sizeof(Ctrl), cudaMemcpyDeviceToHost, s));
  CK(cudaEventRecord(h->ev_done, s));  // collect waits on this event, whatever stream it is given
  h->in_flight = 1; h->track_stream = s;
  if (so3)  // RO.cpp:595-601: the kernel holds the old pointers, so the swap can be applied right away
    for (int i = 0; i < 3; ++i) { unsigned char* t = h->lv[i].lastNextImage; h->lv[i].lastNextImage = h->lv[i].nextImage; h->lv[i].nextImage = t; }
  return DMS_OK;
}

int dms_odom_track_collect(dms_odom* h, float trans[3], float rot[9], void* stream, dms_odom_result* out) {
  if (!h || !trans || !rot) return fail(h, DMS_ERR_INVALID, "track_collect: null argument");
  Bind b(h->device);
  cudaError_t e;
  (void)stream;  // the result is ordered by the event recorded on the stream the track was launched on
  if (!h->in_flight) return fail(h, DMS_ERR_INVALID, "track_collect: no track pending");
  h->in_flight = 0;
  if (h->blocking_wait >= 2) {
    const auto t0 = std::chrono::steady_clock::now();
    const bool slept = h->wait_ema_us > 150.0;
    if (slept) std::this_thread::sleep_for(std::chrono::nanoseconds((long long)(h->wait_ema_us * 850.0)));
    e = cudaEventQuery(h->ev_done);
    if (slept && e != cudaErrorNotReady) {
      // the result was already there when the sleep ended: the estimate is too long by an unknown amount.  Measuring
      // this wait would only return the sleep itself, so one long wait (the first frame of a run waits for its uploads)
      // kept the estimate high for dozens of frames and every later result was collected late: shrink it instead.
      h->wait_ema_us *= 0.7;
    } else {
      while (e == cudaErrorNotReady) {
        if (h->blocking_wait == 3) std::this_thread::sleep_for(std::chrono::microseconds(20)); else sched_yield();
        e = cudaEventQuery(h->ev_done);
      }
      const double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
      h->wait_ema_us = h->wait_ema_us == 0.0 ? us : 0.8 * h->wait_ema_us + 0.2 * us;
    }
  } else if (h->blocking_wait == 1) {
    e = cudaEventSynchronize(h->ev_done);
  } else {
    while ((e = cudaEventQuery(h->ev_done)) == cudaErrorNotReady) { }  // spin: lowest latency
  }
  const bool aborted = e == cudaSuccess && h->h_ctrl->aborted != 0u;
  if (e != cudaSuccess || aborted) {
    // a failed launch or an aborted grid barrier leaves the barrier words and accumulators undefined: rearm them
    cudaStreamSynchronize(h->track_stream);
    memset(h->h_ctrl, 0, sizeof(Ctrl));
    h->h_ctrl->res = h->last;
    cudaMemcpy(h->ctrl, h->h_ctrl, sizeof(Ctrl), cudaMemcpyHostToDevice);
    cudaMemset(h->acc, 0, (size_t)ACC_SLOTS * ACC_STRIDE * sizeof(unsigned long long));
    cudaMemset(h->sync, 0, sizeof(unsigned int) * SYNC_WORDS);
    if (aborted) return fail(h, DMS_ERR_DEVICE, "k_track: a grid barrier timed out (a thread block never arrived)");
    return fail(h, (int)e, "k_track");
  }
  const Ctrl* C = h->h_ctrl;
  memcpy(rot, C->out_R, sizeof(float) * 9); memcpy(trans, C->out_t, sizeof(float) * 3);
  h->last = C->res;
  if (out) *out = h->last;
  if (h->timing) {
    float ms = 0.f; cudaEventElapsedTime(&ms, h->ev0, h->ev1);
    h->last_timing.track_ms = ms;
    if (h->timing == 3) {  // events only: the production kernel keeps no phase counters
      const float ms_ = h->last_timing.track_ms;
      memset(&h->last_timing, 0, sizeof h->last_timing);
      h->last_timing.track_ms = ms_;
      return DMS_OK;
    }
    h->last_timing.so3Step_ms = (float)(C->t_so3 * 1e-6);
    h->last_timing.icpStep_ms = (float)(C->t_phaseA * 1e-6);  // ICP rows + photometric residual are one fused pass
    h->last_timing.computeRgbResidual_ms = 0.f;
    h->last_timing.rgbStep_ms = (float)(C->t_phaseB * 1e-6);
    h->last_timing.prologue_ms = (float)(C->t_prologue * 1e-6);
    h->last_timing.serial_so3_ms = (float)(C->t_ser_so3 * 1e-6);
    h->last_timing.serial_a_ms = (float)(C->t_ser_a * 1e-6);
    h->last_timing.serial_b_ms = (float)(C->t_ser_b * 1e-6);
    h->last_timing.barrier_wait_ms = (float)(C->t_wait * 1e-6);
  }
  return DMS_OK;
}

int dms_odom_track(dms_odom* h, float trans[3], float rot[9], int rgbOnly, float icpWeight, int pyramid, int fastOdom, int so3,
                   int interMap, void* stream, dms_odom_result* out) {
  int rc = dms_odom_track_async(h, trans, rot, rgbOnly, icpWeight, pyramid, fastOdom, so3, interMap, stream);
  if (rc) return rc;
  return dms_odom_track_collect(h, trans, rot, stream, out);
}

int dms_odom_set_trace(dms_odom* h, int enabled) { if (!h) return DMS_ERR_INVALID; h->tracing = enabled ? 1 : 0; return DMS_OK; }

int dms_odom_get_trace(dms_odom* h, dms_iter_record* recs, int max_records, int* n_records) {
  if (!h || !n_records) return fail(h, DMS_ERR_INVALID, "get_trace: null argument");
  Bind b(h->device);
  int n = h->tracing ? h->h_ctrl->n_trace : 0;
  if (n > MAX_TRACE) n = MAX_TRACE;
  if (n > max_records) n = max_records;
  if (n > 0 && recs) CK(cudaMemcpy(recs, h->trace, sizeof(dms_iter_record) * n, cudaMemcpyDeviceToHost));
  *n_records = n;
  return DMS_OK;
}

int dms_odom_get_timeline(dms_odom* h, uint64_t* pairs, int max_pairs, int* n_pairs) {
  if (!h || !n_pairs) return fail(h, DMS_ERR_INVALID, "get_timeline: null argument");
  Bind b(h->device);
  int n = MAX_DBG;  // CTA 0 logs from entry 0, the solver from entry 512; unused entries are zero
  if (n > max_pairs) n = max_pairs;
  if (n > 0 && pairs) CK(cudaMemcpy(pairs, h->dbg, sizeof(unsigned long long) * 2 * n, cudaMemcpyDeviceToHost));
  *n_pairs = n;
  return DMS_OK;
}

int dms_odom_get_covariance(const dms_odom* h, double cov[36]) {
  if (!h || !cov) return DMS_ERR_INVALID;
  // lastA.lu().inverse(): Gauss-Jordan with partial pivoting in double (RO.cpp:607-610)
  double m[6][12];
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) { m[i][j] = h->last.lastA[i * 6 + j]; m[i][6 + j] = (i == j) ? 1.0 : 0.0; }
  for (int c = 0; c < 6; ++c) {
    int p = c;
    for (int r = c + 1; r < 6; ++r) if (fabs(m[r][c]) > fabs(m[p][c])) p = r;
    if (p != c) for (int j = 0; j < 12; ++j) { const double t = m[p][j]; m[p][j] = m[c][j]; m[c][j] = t; }
    const double d = m[c][c];
    for (int j = 0; j < 12; ++j) m[c][j] /= d;
    for (int r = 0; r < 6; ++r) if (r != c) { const double f = m[r][c]; for (int j = 0; j < 12; ++j) m[r][j] -= f * m[c][j]; }
  }
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) cov[i * 6 + j] = m[i][6 + j];
  return DMS_OK;
}

// ------------------------------------------------------------------ step operators
static int fetch_step(dms_odom* h, cudaStream_t s) {
  CK(cudaMemcpyAsync(h->h_step, h->step_out, sizeof(StepOut), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  return DMS_OK;
}

int dms_odom_icp_step(dms_odom* h, int level, const float Rcurr[9], const float tcurr[3], const float Rprev_inv[9], const float tprev[3],
                      float out28[28], int64_t* inliers, uint8_t* h_mask, void* stream) {
  if (!h || level < 0 || level > 2 || !Rcurr || !tcurr || !Rprev_inv || !tprev) return fail(h, DMS_ERR_INVALID, "icp_step: bad argument");
  if (h_mask && level != 0) return fail(h, DMS_ERR_INVALID, "icp_step: mask is level-0 only");
  Bind b(h->device);
  cudaStream_t s = (cudaStream_t)stream;
  const Level& L = h->lv[level];
  CK(launch_icp_step(L, Rcurr, tcurr, Rprev_inv, tprev, h->distThres, h->angleThres, h->step_acc, h->step_out, h_mask ? h->mask : nullptr,
                     h->lv[0].pitch_u8, h->legacy != 0, s));
  int rc = fetch_step(h, s); if (rc) return rc;
  if (out28) memcpy(out28, h->h_step->vals, sizeof(float) * 28);
  if (inliers) *inliers = h->h_step->count;
  if (h_mask) CK(cudaMemcpy2D(h_mask, L.cols, h->mask, h->lv[0].pitch_u8, L.cols, L.rows, cudaMemcpyDeviceToHost));
  return DMS_OK;
}

int dms_odom_rgb_residual(dms_odom* h, int level, const float kt[3], const float krkinv[9], int64_t* count, int64_t* sigma, void* stream) {
  if (!h || level < 0 || level > 2 || !kt || !krkinv) return fail(h, DMS_ERR_INVALID, "rgb_residual: bad argument");
  Bind b(h->device);
  cudaStream_t s = (cudaStream_t)stream;
  CK(launch_rgb_residual(h->lv[level], kt, krkinv, h->step_acc, h->step_out, h->legacy != 0, s));
  int rc = fetch_step(h, s); if (rc) return rc;
  if (count) *count = h->h_step->count;
  if (sigma) *sigma = h->h_step->sigma;
  return DMS_OK;
}

int dms_odom_rgb_step(dms_odom* h, int level, float sigma, float out28[28], int64_t* count, void* stream) {
  if (!h || level < 0 || level > 2) return fail(h, DMS_ERR_INVALID, "rgb_step: bad argument");
  Bind b(h->device);
  cudaStream_t s = (cudaStream_t)stream;
  CK(launch_rgb_step(h->lv[level], sigma, h->step_acc, h->step_out, h->legacy != 0, s));
  int rc = fetch_step(h, s); if (rc) return rc;
  if (out28) memcpy(out28, h->h_step->vals, sizeof(float) * 28);
  if (count) *count = h->h_step->count;
  return DMS_OK;
}

int dms_odom_so3_step(dms_odom* h, int level, const float H[9], const float kinv[9], const float krlr[9], float out10[10], int64_t* count,
                      void* stream) {
  if (!h || level < 0 || level > 2 || !H || !kinv || !krlr) return fail(h, DMS_ERR_INVALID, "so3_step: bad argument");
  Bind b(h->device);
  cudaStream_t s = (cudaStream_t)stream;
  CK(launch_so3_step(h->lv[level], H, kinv, krlr, h->step_acc, h->step_out, h->legacy != 0, s));
  int rc = fetch_step(h, s); if (rc) return rc;
  if (out10) memcpy(out10, h->h_step->vals, sizeof(float) * 10);
  if (count) *count = h->h_step->count;
  return DMS_OK;
}

int dms_odom_compute_derivatives(dms_odom* h, void* stream) {
  if (!h) return DMS_ERR_INVALID;
  Bind b(h->device);
  CK(launch_derivatives(h->lv, (cudaStream_t)stream));
  return DMS_OK;
}

// getICPResidual (RO.cpp:612-643): level-0 residual-only ICP.  The reference passes rot_curr.inverse() as
// Rprev_inv and trans_prev unchanged (RO.cpp:623-625); that call shape is kept.
int dms_odom_icp_residual(dms_odom* h, const float trans_curr[3], const float rot_curr[9], const float trans_prev[3],
                          const float rot_prev[9], float* icpError, int* icpCount, void* stream) {
  (void)rot_prev;
  if (!h || !trans_curr || !rot_curr || !trans_prev) return fail(h, DMS_ERR_INVALID, "icp_residual: null argument");
  float Rinv[9]; mat3_inv_host<float>(rot_curr, Rinv);
  float out[28]; int64_t cnt = 0;
  int rc = dms_odom_icp_step(h, 0, rot_curr, trans_curr, Rinv, trans_prev, out, &cnt, nullptr, stream);
  if (rc) return rc;
  if (icpCount) *icpCount = (int)cnt;
  if (icpError) *icpError = sqrtf(out[27]) / (float)(int)cnt;
  return DMS_OK;
}

// ------------------------------------------------------------------ depth pre-filter (stateless)
int dms_depth_bilateral(const uint16_t* d_raw, size_t pitch_in, int rows, int cols, float maxDepth, uint16_t* d_filtered, size_t pitch_out,
                        void* stream) {
  if (!d_raw || !d_filtered || rows <= 0 || cols <= 0) return DMS_ERR_INVALID;
  cudaError_t e = launch_depth_bilateral(d_raw, (int)(pitch_in ? pitch_in : (size_t)cols * 2), d_filtered, (int)(pitch_out ? pitch_out : (size_t)cols * 2),
                                         rows, cols, maxDepth, (cudaStream_t)stream);
  return e == cudaSuccess ? DMS_OK : (int)e;
}
int dms_depth_metric(const uint16_t* d_raw, size_t pitch_in, int rows, int cols, float maxDepth, float* d_metres, size_t pitch_out, void* stream) {
  if (!d_raw || !d_metres || rows <= 0 || cols <= 0) return DMS_ERR_INVALID;
  cudaError_t e = launch_depth_metric(d_raw, (int)(pitch_in ? pitch_in : (size_t)cols * 2), d_metres, (int)(pitch_out ? pitch_out : (size_t)cols * 4), rows,
                                      cols, maxDepth, (cudaStream_t)stream);
  return e == cudaSuccess ? DMS_OK : (int)e;
}

// ------------------------------------------------------------------ buffers
int dms_odom_get_buffer(dms_odom* h, int which, int level, void** d_ptr, size_t* pitch, int* rows, int* cols, int* elem) {
  if (!h || level < 0 || level > 2) return fail(h, DMS_ERR_INVALID, "get_buffer: bad argument");
  const Level& L = h->lv[level];
  void* p = nullptr; size_t pb = 0; int r = L.rows, e = 4;
  switch (which) {
    case DMS_BUF_VMAP_CURR: p = L.vmap_curr; pb = L.pitch_f32; r = 3 * L.rows; break;
    case DMS_BUF_NMAP_CURR: p = L.nmap_curr; pb = L.pitch_f32; r = 3 * L.rows; break;
    case DMS_BUF_VMAP_G_PREV: p = L.vmap_prev; pb = L.pitch_f32; r = 3 * L.rows; break;
    case DMS_BUF_NMAP_G_PREV: p = L.nmap_prev; pb = L.pitch_f32; r = 3 * L.rows; break;
    case DMS_BUF_LAST_DEPTH: p = L.lastDepth; pb = L.pitch_f32; break;
    case DMS_BUF_NEXT_DEPTH: p = L.nextDepth; pb = L.pitch_f32; break;
    case DMS_BUF_LAST_IMAGE: p = L.lastImage; pb = L.pitch_u8; e = 1; break;
    case DMS_BUF_NEXT_IMAGE: p = L.nextImage; pb = L.pitch_u8; e = 1; break;
    case DMS_BUF_LAST_NEXT_IMAGE: p = L.lastNextImage; pb = L.pitch_u8; e = 1; break;
    case DMS_BUF_DIDX: p = L.dIdx; pb = L.pitch_s16; e = 2; break;
    case DMS_BUF_DIDY: p = L.dIdy; pb = L.pitch_s16; e = 2; break;
    case DMS_BUF_DEPTH_TMP: p = L.depth_tmp; pb = L.pitch_u16; e = 2; break;
    case DMS_BUF_CORRES: p = L.corr; pb = L.pitch_corr; e = 16; break;
    case DMS_BUF_NEXT_DEPTH_MASKED: p = L.nextDepthMasked; pb = L.pitch_f32; break;
    default: return fail(h, DMS_ERR_INVALID, "get_buffer: unknown buffer");
  }
  if (d_ptr) *d_ptr = p;
  if (pitch) *pitch = pb;
  if (rows) *rows = r;
  if (cols) *cols = L.cols;
  if (elem) *elem = e;
  return DMS_OK;
}

int dms_odom_download(dms_odom* h, int which, int level, void* dst, size_t dst_bytes, void* stream) {
  void* p; size_t pitch; int rows, cols, elem;
  int rc = dms_odom_get_buffer(h, which, level, &p, &pitch, &rows, &cols, &elem);
  if (rc) return rc;
  const size_t need = (size_t)rows * cols * elem;
  if (!dst || dst_bytes < need) return fail(h, DMS_ERR_INVALID, "download: destination too small");
  Bind b(h->device);
  cudaStream_t s = (cudaStream_t)stream;
  CK(cudaMemcpy2DAsync(dst, (size_t)cols * elem, p, pitch, (size_t)cols * elem, rows, cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  return DMS_OK;
}

}  // extern "C"
