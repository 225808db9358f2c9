// Tracking front end (include/dmsodom/frontend_c.h): the tracking part of ElasticFusion::processFrame
// (ElasticFusion.cpp:99-268, 397-442) over two dms_odom trackers, with the GLSL passes that feed them
// (FillIn, Resize + denseEnough) as CUDA kernels.  Host logic is plain C++; nothing here allocates per frame.
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <new>
#include "common.cuh"
#include "kernels.h"
#include "../../include/dmsodom/frontend_c.h"

namespace dms {

// ------------------------------------------------------------------ FillIn (Shaders/FillIn.cpp:65-190)
// geometry.glsl:53-58: forward-difference vertex from raw u16 millimetres; cam = (cx, cy, 1/fx, 1/fy)
__device__ __forceinline__ float3 fill_vertex_at(const unsigned short* depth, int rows, int cols, int sx, int sy, int x, int y, float4 cam) {
  sx = min(max(sx, 0), cols - 1); sy = min(max(sy, 0), rows - 1);  // clamp-to-edge sampling
  const float z = fdiv((float)depth[(size_t)sy * cols + sx], 1000.0f);
  return make_float3(fmul(fmul(fsub((float)x, cam.x), z), cam.z), fmul(fmul(fsub((float)y, cam.y), z), cam.w), z);
}

__global__ void __launch_bounds__(256) k_fill_in(const float4* mv, const float4* mn, const uchar4* mi, const unsigned short* depth, const uchar4* live,
                                                 int rows, int cols, float4 cam, int pass_geom, int pass_img, float4* ov, float4* on, uchar4* oi) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= cols || y >= rows) return;
  const size_t i = (size_t)y * cols + x;
  if (ov || on) {
    const float3 vp = fill_vertex_at(depth, rows, cols, x, y, x, y, cam);
    if (ov) {  // fill_vertex.frag:37-52
      const float4 s = mv[i];
      ov[i] = (s.z == 0 || pass_geom) ? make_float4(vp.x, vp.y, vp.z, 1.0f) : s;
    }
    if (on) {  // fill_normal.frag:38-51, geometry.glsl:61-77
      const float4 s = mn[i];
      if (s.z == 0 || pass_geom) {
        const float3 vx = fill_vertex_at(depth, rows, cols, x + 1, y, x + 1, y, cam);
        const float3 vy = fill_vertex_at(depth, rows, cols, x, y + 1, x, y + 1, cam);
        const float3 dx = sub3(vx, vp), dy = sub3(vy, vp);
        const float3 c = make_float3(fsub(fmul(dx.y, dy.z), fmul(dx.z, dy.y)), fsub(fmul(dx.z, dy.x), fmul(dx.x, dy.z)),
                                     fsub(fmul(dx.x, dy.y), fmul(dx.y, dy.x)));
        const float inv = fdiv(1.0f, fsqrt(fadd(fadd(fmul(c.x, c.x), fmul(c.y, c.y)), fmul(c.z, c.z))));
        on[i] = make_float4(fmul(c.x, inv), fmul(c.y, inv), fmul(c.z, inv), 1.0f);
      } else {
        on[i] = s;
      }
    }
  }
  if (oi) {  // fill_rgb.frag:29-36 (normalised colours: the sum is zero iff all three bytes are zero)
    const uchar4 s = mi[i];
    oi[i] = ((s.x == 0 && s.y == 0 && s.z == 0) || pass_img) ? live[i] : s;
  }
}

// Resize::image (nearest sample at the destination texel centre) + ElasticFusion::denseEnough (EF.cpp:84-97)
__global__ void __launch_bounds__(256) k_dense_count(const uchar4* img, int rows, int cols, int drows, int dcols, unsigned int* count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  bool ok = false;
  if (i < drows * dcols) {
    const int dy = i / dcols, dx = i - dy * dcols;
    const int sx = min((int)(((long long)(2 * dx + 1) * cols) / (2 * dcols)), cols - 1);
    const int sy = min((int)(((long long)(2 * dy + 1) * rows) / (2 * drows)), rows - 1);
    const uchar4 p = img[(size_t)sy * cols + sx];
    ok = p.x > 0 && p.y > 0 && p.z > 0;
  }
  const unsigned int b = __ballot_sync(0xffffffffu, ok);
  if ((threadIdx.x & 31) == 0 && b) atomicAdd(count, (unsigned int)__popc(b));
}

// ------------------------------------------------------------------ small host algebra
static void mat4_mul(const float* a, const float* b, float* c) {
  float t[16];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      float s = 0.f;
      for (int k = 0; k < 4; ++k) s += a[i * 4 + k] * b[k * 4 + j];
      t[i * 4 + j] = s;
    }
  memcpy(c, t, sizeof(t));
}
// general 4x4 inverse by cofactors (Eigen's Matrix4f::inverse() is also cofactor based)
static bool mat4_inv(const float* m, float* out) {
  double inv[16], a[16];
  for (int i = 0; i < 16; ++i) a[i] = m[i];
  inv[0] = a[5] * a[10] * a[15] - a[5] * a[11] * a[14] - a[9] * a[6] * a[15] + a[9] * a[7] * a[14] + a[13] * a[6] * a[11] - a[13] * a[7] * a[10];
  inv[4] = -a[4] * a[10] * a[15] + a[4] * a[11] * a[14] + a[8] * a[6] * a[15] - a[8] * a[7] * a[14] - a[12] * a[6] * a[11] + a[12] * a[7] * a[10];
  inv[8] = a[4] * a[9] * a[15] - a[4] * a[11] * a[13] - a[8] * a[5] * a[15] + a[8] * a[7] * a[13] + a[12] * a[5] * a[11] - a[12] * a[7] * a[9];
  inv[12] = -a[4] * a[9] * a[14] + a[4] * a[10] * a[13] + a[8] * a[5] * a[14] - a[8] * a[6] * a[13] - a[12] * a[5] * a[10] + a[12] * a[6] * a[9];
  inv[1] = -a[1] * a[10] * a[15] + a[1] * a[11] * a[14] + a[9] * a[2] * a[15] - a[9] * a[3] * a[14] - a[13] * a[2] * a[11] + a[13] * a[3] * a[10];
  inv[5] = a[0] * a[10] * a[15] - a[0] * a[11] * a[14] - a[8] * a[2] * a[15] + a[8] * a[3] * a[14] + a[12] * a[2] * a[11] - a[12] * a[3] * a[10];
  inv[9] = -a[0] * a[9] * a[15] + a[0] * a[11] * a[13] + a[8] * a[1] * a[15] - a[8] * a[3] * a[13] - a[12] * a[1] * a[11] + a[12] * a[3] * a[9];
  inv[13] = a[0] * a[9] * a[14] - a[0] * a[10] * a[13] - a[8] * a[1] * a[14] + a[8] * a[2] * a[13] + a[12] * a[1] * a[10] - a[12] * a[2] * a[9];
  inv[2] = a[1] * a[6] * a[15] - a[1] * a[7] * a[14] - a[5] * a[2] * a[15] + a[5] * a[3] * a[14] + a[13] * a[2] * a[7] - a[13] * a[3] * a[6];
  inv[6] = -a[0] * a[6] * a[15] + a[0] * a[7] * a[14] + a[4] * a[2] * a[15] - a[4] * a[3] * a[14] - a[12] * a[2] * a[7] + a[12] * a[3] * a[6];
  inv[10] = a[0] * a[5] * a[15] - a[0] * a[7] * a[13] - a[4] * a[1] * a[15] + a[4] * a[3] * a[13] + a[12] * a[1] * a[7] - a[12] * a[3] * a[5];
  inv[14] = -a[0] * a[5] * a[14] + a[0] * a[6] * a[13] + a[4] * a[1] * a[14] - a[4] * a[2] * a[13] - a[12] * a[1] * a[6] + a[12] * a[2] * a[5];
  inv[3] = -a[1] * a[6] * a[11] + a[1] * a[7] * a[10] + a[5] * a[2] * a[11] - a[5] * a[3] * a[10] - a[9] * a[2] * a[7] + a[9] * a[3] * a[6];
  inv[7] = a[0] * a[6] * a[11] - a[0] * a[7] * a[10] - a[4] * a[2] * a[11] + a[4] * a[3] * a[10] + a[8] * a[2] * a[7] - a[8] * a[3] * a[6];
  inv[11] = -a[0] * a[5] * a[11] + a[0] * a[7] * a[9] + a[4] * a[1] * a[11] - a[4] * a[3] * a[9] - a[8] * a[1] * a[7] + a[8] * a[3] * a[5];
  inv[15] = a[0] * a[5] * a[10] - a[0] * a[6] * a[9] - a[4] * a[1] * a[10] + a[4] * a[2] * a[9] + a[8] * a[1] * a[6] - a[8] * a[2] * a[5];
  const double det = a[0] * inv[0] + a[1] * inv[4] + a[2] * inv[8] + a[3] * inv[12];
  if (det == 0.0) return false;
  for (int i = 0; i < 16; ++i) out[i] = (float)(inv[i] / det);
  return true;
}
static void mat3_inv_t(const double* m, double* o) {  // o = inverse(m) transposed
  const double c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
  const double det = m[0] * c00 + m[1] * c01 + m[2] * c02, id = 1.0 / det;
  o[0] = c00 * id; o[1] = c01 * id; o[2] = c02 * id;
  o[3] = (m[2] * m[7] - m[1] * m[8]) * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[1] * m[6] - m[0] * m[7]) * id;
  o[6] = (m[1] * m[5] - m[2] * m[4]) * id; o[7] = (m[2] * m[3] - m[0] * m[5]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}
// ElasticFusion::rodrigues2 (elasticfusion/Core/src/ElasticFusion.cpp:941-986), restated without Eigen / OpenCV: the host
// math of the reference's gate, kept in its operation order on purpose (parity of the loop-closure gate needs it).  R = U * V^T of the SVD is the orthogonal polar factor, obtained
// here by Newton's iteration X <- (X + X^-T) / 2 (quadratic convergence from a near-rotation).
static void rodrigues2(const float* Rf, double out[3]) {
  double R[9], T[9];
  for (int i = 0; i < 9; ++i) R[i] = Rf[i];
  for (int it = 0; it < 8; ++it) {
    mat3_inv_t(R, T);
    for (int i = 0; i < 9; ++i) R[i] = 0.5 * (R[i] + T[i]);
  }
  float Rq[9];
  for (int i = 0; i < 9; ++i) Rq[i] = (float)R[i];  // the reference's R is an Eigen::Matrix3f
  double rx = Rq[7] - Rq[5], ry = Rq[2] - Rq[6], rz = Rq[3] - Rq[1];
  const double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
  double c = ((Rq[0] + Rq[4] + Rq[8]) - 1) * 0.5;
  c = c > 1. ? 1. : c < -1. ? -1. : c;
  double theta = acos(c);
  if (s < 1e-5) {
    if (c > 0) {
      rx = ry = rz = 0;
    } else {
      double t = (Rq[0] + 1) * 0.5;
      rx = sqrt(t > 0.0 ? t : 0.0);
      t = (Rq[4] + 1) * 0.5;
      ry = sqrt(t > 0.0 ? t : 0.0) * (Rq[1] < 0 ? -1.0 : 1.0);
      t = (Rq[8] + 1) * 0.5;
      rz = sqrt(t > 0.0 ? t : 0.0) * (Rq[2] < 0 ? -1.0 : 1.0);
      if (fabs(rx) < fabs(ry) && fabs(rx) < fabs(rz) && (Rq[5] > 0) != (ry * rz > 0)) rz = -rz;
      theta /= sqrt(rx * rx + ry * ry + rz * rz);
      rx *= theta; ry *= theta; rz *= theta;
    }
  } else {
    double vth = 1 / (2 * s);
    vth *= theta;
    rx *= vth; ry *= vth; rz *= vth;
  }
  out[0] = (double)(float)rx; out[1] = (double)(float)ry; out[2] = (double)(float)rz;
}

static const float kIdentity[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};

}  // namespace dms

using namespace dms;

struct dms_frontend {
  int device, width, height;
  float cx, cy, fx, fy;
  dms_frontend_options opt;
  dms_odom* frameToModel;
  dms_odom* modelToModel;
  dms_predict_fn predict;
  void* predict_user;
  // Context state (Context.h:280-290, 333)
  int tick, lost, trackingCount, lastFrameRecovery;
  float currPose[16];
  // device scratch: DEPTH_FILTERED and the FillIn textures
  unsigned short* d_depth_filt;
  float *d_fill_v, *d_fill_n;
  unsigned char* d_fill_img;
  unsigned int* d_count;
  unsigned int* h_count;
};

extern "C" {

void dms_frontend_default_options(dms_frontend_options* o) {
  if (!o) return;
  o->depthCutoff = 3.0f; o->icpWeight = 10.0f; o->rgbOnly = 0; o->pyramid = 1; o->fastOdom = 0; o->so3 = 1;
  o->frameToFrameRGB = 0; o->reloc = 0; o->closeLoops = 1; o->hybridTracking = 1; o->weightMultiplier = 1.0f;
}

int dms_frontend_destroy(dms_frontend* fe) {
  if (!fe) return DMS_OK;
  cudaSetDevice(fe->device);
  if (fe->frameToModel) dms_odom_destroy(fe->frameToModel);
  if (fe->modelToModel) dms_odom_destroy(fe->modelToModel);
  if (fe->d_depth_filt) cudaFree(fe->d_depth_filt);
  if (fe->d_fill_v) cudaFree(fe->d_fill_v);
  if (fe->d_fill_n) cudaFree(fe->d_fill_n);
  if (fe->d_fill_img) cudaFree(fe->d_fill_img);
  if (fe->d_count) cudaFree(fe->d_count);
  if (fe->h_count) cudaFreeHost(fe->h_count);
  delete fe;
  return DMS_OK;
}

int dms_frontend_create(dms_frontend** out, int device, int width, int height, float cx, float cy, float fx, float fy,
                        const dms_frontend_options* opt) {
  if (!out || width <= 0 || height <= 0) return DMS_ERR_INVALID;
  *out = nullptr;
  dms_frontend* fe = new (std::nothrow) dms_frontend();
  if (!fe) return DMS_ERR_NOMEM;
  memset(fe, 0, sizeof(*fe));
  fe->device = device; fe->width = width; fe->height = height; fe->cx = cx; fe->cy = cy; fe->fx = fx; fe->fy = fy;
  if (opt) fe->opt = *opt; else dms_frontend_default_options(&fe->opt);
  fe->tick = 1;  // Context.h:43
  memcpy(fe->currPose, kIdentity, sizeof(kIdentity));
  // RGBDOdometry(width, height, cx, cy, fx, fy) with the default thresholds (Context.h:31-40, RO.h:34-36)
  const float angle = sinf(20.f * 3.14159254f / 180.f);
  int rc = dms_odom_create(&fe->frameToModel, device, width, height, cx, cy, fx, fy, 0.10f, angle);
  if (rc == DMS_OK) rc = dms_odom_create(&fe->modelToModel, device, width, height, cx, cy, fx, fy, 0.10f, angle);
  const size_t n = (size_t)width * height;
  cudaError_t e = cudaSuccess;
  if (rc == DMS_OK) {
    cudaSetDevice(device);
    e = cudaMalloc(&fe->d_depth_filt, n * 2);
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_fill_v, n * 16);
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_fill_n, n * 16);
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_fill_img, n * 4);
    if (e == cudaSuccess) e = cudaMalloc(&fe->d_count, 256);
    if (e == cudaSuccess) e = cudaMallocHost(&fe->h_count, 256);
    if (e != cudaSuccess) rc = (int)e;
  }
  if (rc != DMS_OK) { dms_frontend_destroy(fe); return rc; }
  *out = fe;
  return DMS_OK;
}

int dms_frontend_set_predictor(dms_frontend* fe, dms_predict_fn fn, void* user) {
  if (!fe) return DMS_ERR_INVALID;
  fe->predict = fn; fe->predict_user = user;
  return DMS_OK;
}
int dms_frontend_set_state(dms_frontend* fe, int lost, int lastFrameRecovery) {
  if (!fe) return DMS_ERR_INVALID;
  fe->lost = lost ? 1 : 0; fe->lastFrameRecovery = lastFrameRecovery ? 1 : 0;
  return DMS_OK;
}
int dms_frontend_get_tracker(dms_frontend* fe, int which, dms_odom** out) {
  if (!fe || !out || which < 0 || which > 1) return DMS_ERR_INVALID;
  *out = which == 0 ? fe->frameToModel : fe->modelToModel;
  return DMS_OK;
}
int dms_frontend_get_filtered_depth(dms_frontend* fe, const uint16_t** d_depth, size_t* pitch_bytes) {
  if (!fe || !d_depth) return DMS_ERR_INVALID;
  *d_depth = fe->d_depth_filt;
  if (pitch_bytes) *pitch_bytes = (size_t)fe->width * 2;
  return DMS_OK;
}
int dms_frontend_get_fill_in(dms_frontend* fe, dms_model_view* out) {
  if (!fe || !out) return DMS_ERR_INVALID;
  out->d_vertex = fe->d_fill_v; out->d_normal = fe->d_fill_n; out->d_image = fe->d_fill_img;
  return DMS_OK;
}

int dms_fill_in(const float* mv, const float* mn, const uint8_t* mi, const uint16_t* depth, const uint8_t* live, int rows, int cols, float cx, float cy,
                float fx, float fy, int pass_geom, int pass_img, float* ov, float* on, uint8_t* oi, void* stream) {
  if (rows <= 0 || cols <= 0) return DMS_ERR_INVALID;
  if ((ov && (!mv || !depth)) || (on && (!mn || !depth)) || (oi && (!mi || !live))) return DMS_ERR_INVALID;
  const float4 cam = make_float4(cx, cy, 1.0f / fx, 1.0f / fy);  // FillIn.cpp:113-116
  k_fill_in<<<dim3((cols + 31) / 32, (rows + 7) / 8), 256, 0, (cudaStream_t)stream>>>(
      (const float4*)mv, (const float4*)mn, (const uchar4*)mi, depth, (const uchar4*)live, rows, cols, cam, pass_geom, pass_img, (float4*)ov, (float4*)on,
      (uchar4*)oi);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? DMS_OK : (int)e;
}

int dms_dense_fraction(const uint8_t* d_rgba, int rows, int cols, int sample, unsigned int* d_scratch, float* fraction, void* stream) {
  if (!d_rgba || !d_scratch || !fraction || rows <= 0 || cols <= 0 || sample <= 0) return DMS_ERR_INVALID;
  const int drows = rows / sample > 0 ? rows / sample : 1, dcols = cols / sample > 0 ? cols / sample : 1;  // EF.cpp:44-47
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(d_scratch, 0, 4, s);
  if (e != cudaSuccess) return (int)e;
  const int n = drows * dcols;
  k_dense_count<<<(n + 255) / 256, 256, 0, s>>>((const uchar4*)d_rgba, rows, cols, drows, dcols, d_scratch);
  unsigned int h = 0;
  e = cudaMemcpyAsync(&h, d_scratch, 4, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return (int)e;
  *fraction = (float)(int)h / (float)n;  // float(sum) / float(img.rows * img.cols)
  return DMS_OK;
}

float dms_velocity_weighting(const float currPose[16], const float lastPose[16], float weightMultiplier) {
  float inv[16], diff[16];
  if (!mat4_inv(currPose, inv)) return 0.5f * weightMultiplier;
  mat4_mul(inv, lastPose, diff);  // diff = currPose.inverse() * lastPose, EF.cpp:252
  const float tn = sqrtf(diff[3] * diff[3] + diff[7] * diff[7] + diff[11] * diff[11]);
  const float R[9] = {diff[0], diff[1], diff[2], diff[4], diff[5], diff[6], diff[8], diff[9], diff[10]};
  double r[3];
  rodrigues2(R, r);
  const float rn = sqrtf((float)r[0] * (float)r[0] + (float)r[1] * (float)r[1] + (float)r[2] * (float)r[2]);
  float weighting = tn > rn ? tn : rn;
  const float largest = 0.01f, minWeight = 0.5f;
  if (weighting > largest) weighting = largest;
  const float w = 1.0f - (weighting / largest);
  return (w > minWeight ? w : minWeight) * weightMultiplier;
}

static int fe_cov_ok(dms_odom* h, double thresh, double diag[6]) {  // EF.cpp:208-214 / 427-434
  double cov[36];
  const int rc = dms_odom_get_covariance(h, cov);
  int ok = 1;
  for (int i = 0; i < 6; ++i) {
    if (diag) diag[i] = cov[i * 6 + i];
    if (rc != DMS_OK || cov[i * 6 + i] > thresh) ok = 0;  // NaN compares false, as in the reference
  }
  return ok;
}

int dms_frontend_process(dms_frontend* fe, const uint8_t* d_rgba, size_t rgba_pitch, const uint16_t* d_depth_raw, size_t depth_pitch,
                         const float* inPose, dms_frontend_result* out, void* stream) {
  if (!fe || !d_rgba || !d_depth_raw || !out) return DMS_ERR_INVALID;
  cudaSetDevice(fe->device);
  cudaStream_t s = (cudaStream_t)stream;
  const int W = fe->width, H = fe->height;
  const dms_frontend_options& o = fe->opt;
  memset(out, 0, sizeof(*out));
  out->tick = fe->tick;
  if (rgba_pitch && rgba_pitch != (size_t)W * 4) return DMS_ERR_INVALID;  // dense live image (it is also a FillIn input)
  // Preprocess: filterDepth (EF.cpp:116, 759-768)
  int rc = dms_depth_bilateral(d_depth_raw, depth_pitch, H, W, o.depthCutoff, fe->d_depth_filt, (size_t)W * 2, s);
  if (rc != DMS_OK) return rc;

  if (fe->tick == 1) {  // first run of the (single-context) reference frame, EF.cpp:131-152
    memcpy(fe->currPose, inPose ? inPose : kIdentity, sizeof(fe->currPose));
    rc = dms_odom_init_first_rgb(fe->frameToModel, d_rgba, 0, s);
    if (rc != DMS_OK) return rc;
    out->trackingOk = 1; out->weighting = o.weightMultiplier;
  } else {
    float lastPose[16];
    memcpy(lastPose, fe->currPose, sizeof(lastPose));
    int trackingOk = 1;
    if (inPose) memcpy(fe->currPose, inPose, sizeof(fe->currPose));  // EF.cpp:164
    if (!fe->predict) return DMS_ERR_INVALID;
    dms_model_view view = {nullptr, nullptr, nullptr};
    if (fe->predict(fe->predict_user, fe->currPose, DMS_PREDICT_TRACK, fe->tick, stream, &view) != 0 || !view.d_vertex || !view.d_normal || !view.d_image)
      return DMS_ERR_INVALID;  // no model view, no tracking
    // predict() ends with the FillIn passes (EF.cpp:733-743)
    rc = dms_fill_in(view.d_vertex, view.d_normal, view.d_image, fe->d_depth_filt, d_rgba, H, W, fe->cx, fe->cy, fe->fx, fe->fy, fe->lost,
                     fe->lost || o.frameToFrameRGB, fe->d_fill_v, fe->d_fill_n, fe->d_fill_img, s);
    if (rc != DMS_OK) return rc;
    float frac = 0.f;
    rc = dms_dense_fraction(view.d_image, H, W, 20, fe->d_count, &frac, s);  // resize.image + denseEnough, EF.cpp:166-167
    if (rc != DMS_OK) return rc;
    const int shouldFillIn = !(frac > 0.95f);
    out->denseFraction = frac; out->shouldFillIn = shouldFillIn; out->tracked = 1;
    if (o.hybridTracking) {
      dms_odom* f2m = fe->frameToModel;
      // WARNING initICP* must be called before initRGB* (EF.cpp:172)
      rc = dms_odom_init_icp_model(f2m, shouldFillIn ? fe->d_fill_v : view.d_vertex, shouldFillIn ? fe->d_fill_n : view.d_normal, 0, o.depthCutoff,
                                   fe->currPose, s);
      if (rc == DMS_OK) rc = dms_odom_init_rgb_model(f2m, (shouldFillIn || o.frameToFrameRGB) ? fe->d_fill_img : view.d_image, 0, s);
      if (rc == DMS_OK) rc = dms_odom_init_icp_depth(f2m, fe->d_depth_filt, 0, o.depthCutoff, s);
      if (rc == DMS_OK) rc = dms_odom_init_rgb(f2m, d_rgba, 0, s);
      if (rc != DMS_OK) return rc;
      float trans[3] = {fe->currPose[3], fe->currPose[7], fe->currPose[11]};
      float rot[9] = {fe->currPose[0], fe->currPose[1], fe->currPose[2], fe->currPose[4], fe->currPose[5],
                      fe->currPose[6], fe->currPose[8], fe->currPose[9], fe->currPose[10]};
      rc = dms_odom_track(f2m, trans, rot, o.rgbOnly, o.icpWeight, o.pyramid, o.fastOdom, o.so3, 0, s, &out->frameToModel);
      if (rc != DMS_OK) return rc;
      trackingOk = !o.reloc || out->frameToModel.lastICPError < 1e-04f;  // EF.cpp:203-204
      if (o.reloc) {  // EF.cpp:206-244
        if (!fe->lost) {
          if (!fe_cov_ok(f2m, 1e-04, nullptr)) trackingOk = 0;
          if (!trackingOk) {
            fe->trackingCount++;
            if (fe->trackingCount > 10) fe->lost = 1;
          } else {
            fe->trackingCount = 0;
          }
        } else if (fe->lastFrameRecovery) {
          if (!fe_cov_ok(f2m, 1e-04, nullptr)) trackingOk = 0;
          if (trackingOk) { fe->lost = 0; fe->trackingCount = 0; }
          fe->lastFrameRecovery = 0;
        }
      }
      // a non-finite pose (a partial sum overflowed the accumulators, or a NaN came in with the inputs) must not become the
      // state the next frame starts from: keep the pose the frame started with and report the frame as not tracked
      bool finite = true;
      for (int i = 0; i < 3; ++i) finite = finite && isfinite(trans[i]);
      for (int i = 0; i < 9; ++i) finite = finite && isfinite(rot[i]);
      if (finite) {
        fe->currPose[3] = trans[0]; fe->currPose[7] = trans[1]; fe->currPose[11] = trans[2];
        for (int i = 0; i < 3; ++i)
          for (int j = 0; j < 3; ++j) fe->currPose[i * 4 + j] = rot[i * 3 + j];
      } else {
        trackingOk = 0;
      }
    }
    out->trackingOk = trackingOk;
    out->weighting = dms_velocity_weighting(fe->currPose, lastPose, o.weightMultiplier);  // EF.cpp:251-266

    // GlobalPredict (EF.cpp:272-274) and the local loop-closure registration (EF.cpp:399-442; the fern lookup in
    // between is switched off in the reference, so rawGraph is empty here)
    if (!fe->lost && o.closeLoops) {
      dms_model_view cur = {nullptr, nullptr, nullptr}, old = {nullptr, nullptr, nullptr};
      if (fe->predict(fe->predict_user, fe->currPose, DMS_PREDICT_GLOBAL, fe->tick, stream, &cur) != 0 || !cur.d_vertex || !cur.d_normal || !cur.d_image)
        cur = view;
      if (fe->predict(fe->predict_user, fe->currPose, DMS_PREDICT_INACTIVE, fe->tick, stream, &old) == 0 && old.d_vertex && old.d_normal && old.d_image) {
        dms_odom* m2m = fe->modelToModel;
        rc = dms_odom_init_icp_model(m2m, old.d_vertex, old.d_normal, 0, o.depthCutoff, fe->currPose, s);
        if (rc == DMS_OK) rc = dms_odom_init_rgb_model(m2m, old.d_image, 0, s);
        if (rc == DMS_OK) rc = dms_odom_init_icp_maps(m2m, cur.d_vertex, cur.d_normal, 0, o.depthCutoff, s);
        if (rc == DMS_OK) rc = dms_odom_init_rgb(m2m, cur.d_image, 0, s);
        if (rc != DMS_OK) return rc;
        float trans[3] = {fe->currPose[3], fe->currPose[7], fe->currPose[11]};
        float rot[9] = {fe->currPose[0], fe->currPose[1], fe->currPose[2], fe->currPose[4], fe->currPose[5],
                        fe->currPose[6], fe->currPose[8], fe->currPose[9], fe->currPose[10]};
        rc = dms_odom_track(m2m, trans, rot, 0, 10.0f, o.pyramid, o.fastOdom, 0, 0, s, &out->modelToModel);  // EF.cpp:424-425
        if (rc != DMS_OK) return rc;
        const int covOk = fe_cov_ok(m2m, 8e-05, out->loopCovDiag);
        memcpy(out->loopPose, kIdentity, sizeof(kIdentity));
        out->loopPose[3] = trans[0]; out->loopPose[7] = trans[1]; out->loopPose[11] = trans[2];
        for (int i = 0; i < 3; ++i)
          for (int j = 0; j < 3; ++j) out->loopPose[i * 4 + j] = rot[i * 3 + j];
        out->loopChecked = 1;
        out->loopAccepted = covOk && out->modelToModel.lastICPCount > 15000 && out->modelToModel.lastICPError < 0.0003f;
      }
    }
  }
  memcpy(out->pose, fe->currPose, sizeof(fe->currPose));
  out->lost = fe->lost; out->trackingCount = fe->trackingCount; out->lastFrameRecovery = fe->lastFrameRecovery;
  if (!fe->lost) fe->tick++;  // EF.cpp:588-591
  return DMS_OK;
}

}  // extern "C"
