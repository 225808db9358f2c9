// Driver for the reference's own legacy kernels (Cuda/reduce.cu, Cuda/cudafuncs.cu), compiled from
// /root/reference by oracle/ref_build.sh into oracle/_ref/libref_legacy.so.  TEST INFRASTRUCTURE ONLY:
// it pins the CPU oracle against the reference's code and provides the "legacy CUDA" reported baseline.
// Everything here works on dense host arrays; uploads / downloads use the reference's own containers.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstring>
#include <vector>
#include "cudafuncs.cuh"  // from /root/reference/elasticfusion/Core/src/Cuda (via -I)

struct RefImage { const void* ptr; size_t pitch; };

namespace {
template <class T>
void up2d(DeviceArray2D<T>& d, const T* h, int rows, int cols) { d.create(rows, cols); d.upload(h, (size_t)cols * sizeof(T), rows, cols); }
template <class T>
void down2d(const DeviceArray2D<T>& d, T* h, int cols) { d.download(h, (size_t)cols * sizeof(T)); }
mat33 m33(const float* m) { mat33 r; memcpy(r.data, m, sizeof(float) * 9); return r; }
}  // namespace

extern "C" {

int ref_device_ok() { int n = 0; return cudaGetDeviceCount(&n) == cudaSuccess && n > 0; }

// imageBGRToIntensity (cf.cu:657-669)
void ref_bgr_to_intensity(const uint8_t* rgba, int rows, int cols, uint8_t* out) {
  DeviceArray2D<uchar4> src; up2d(src, reinterpret_cast<const uchar4*>(rgba), rows, cols);
  DeviceArray2D<unsigned char> dst; dst.create(rows, cols);
  RefImage img = {src.ptr(0), src.step()};
  imageBGRToIntensity(reinterpret_cast<cudaArray*>(&img), dst);
  cudaDeviceSynchronize();
  down2d(dst, out, cols);
}
void ref_pyrdown_u8_gauss(const uint8_t* s, int rows, int cols, uint8_t* out) {
  DeviceArray2D<unsigned char> a, b; up2d(a, s, rows, cols);
  pyrDownUcharGauss(a, b); cudaDeviceSynchronize(); down2d(b, out, cols / 2);
}
void ref_pyrdown_f32_gauss(const float* s, int rows, int cols, float* out) {
  DeviceArray2D<float> a, b; up2d(a, s, rows, cols);
  pyrDownGaussF(a, b); cudaDeviceSynchronize(); down2d(b, out, cols / 2);
}
void ref_pyrdown_u16(const uint16_t* s, int rows, int cols, uint16_t* out) {
  DeviceArray2D<unsigned short> a, b; up2d(a, s, rows, cols);
  pyrDown(a, b, 0); cudaDeviceSynchronize(); down2d(b, out, cols / 2);
}
// createVMap + createNMap (cf.cu:130-198); maps are pre-filled with zeros
void ref_vmap_nmap(const uint16_t* depth, int rows, int cols, float fx, float fy, float cx, float cy, float cutoff, float* vmap, float* nmap) {
  DeviceArray2D<unsigned short> d; up2d(d, depth, rows, cols);
  std::vector<float> z((size_t)3 * rows * cols, 0.f);
  DeviceArray2D<float> v, n; up2d(v, z.data(), 3 * rows, cols); up2d(n, z.data(), 3 * rows, cols);
  createVMap(CameraModel(fx, fy, cx, cy), d, v, cutoff); createNMap(v, n); cudaDeviceSynchronize();
  down2d(v, vmap, cols); down2d(n, nmap, cols);
}
// copyMaps + resizeVMap/resizeNMap x2 + tranformMaps (RO.cpp:169-207); out arrays hold the three levels back to back
void ref_model_maps(const float* v4, const float* n4, int rows, int cols, const float* R, const float* t, int transform,
                    float* vout, float* nout) {
  DeviceArray<float> vt, nt; vt.create((size_t)rows * cols * 4); nt.create((size_t)rows * cols * 4);
  vt.upload(v4, (size_t)rows * cols * 4); nt.upload(n4, (size_t)rows * cols * 4);
  DeviceArray2D<float> v[3], n[3];
  for (int i = 0; i < 3; ++i) {
    std::vector<float> z((size_t)3 * (rows >> i) * (cols >> i), 0.f);
    up2d(v[i], z.data(), 3 * (rows >> i), cols >> i); up2d(n[i], z.data(), 3 * (rows >> i), cols >> i);
  }
  copyMaps(vt, nt, v[0], n[0]);
  for (int i = 1; i < 3; ++i) { resizeVMap(v[i - 1], v[i]); resizeNMap(n[i - 1], n[i]); }
  if (transform) { mat33 Rm = m33(R); float3 tv = make_float3(t[0], t[1], t[2]); for (int i = 0; i < 3; ++i) tranformMaps(v[i], n[i], Rm, tv, v[i], n[i]); }
  cudaDeviceSynchronize();
  size_t off = 0;
  for (int i = 0; i < 3; ++i) { down2d(v[i], vout + off, cols >> i); down2d(n[i], nout + off, cols >> i); off += (size_t)3 * (rows >> i) * (cols >> i); }
}
void ref_vertices_to_depth(const float* v4, int rows, int cols, float cutoff, float* out) {
  DeviceArray<float> vt; vt.create((size_t)rows * cols * 4); vt.upload(v4, (size_t)rows * cols * 4);
  DeviceArray2D<float> d; d.create(rows, cols);
  verticesToDepth(vt, d, cutoff); cudaDeviceSynchronize(); down2d(d, out, cols);
}
void ref_derivatives(const uint8_t* img, int rows, int cols, int16_t* dx, int16_t* dy) {
  DeviceArray2D<unsigned char> s; up2d(s, img, rows, cols);
  DeviceArray2D<short> a, b; a.create(rows, cols); b.create(rows, cols);
  computeDerivativeImages(s, a, b); cudaDeviceSynchronize(); down2d(a, dx, cols); down2d(b, dy, cols);
}
// icpStep (reduce.cu:367-428): A[36], b[6], residual[2] = {sum r^2, inliers}
void ref_icp_step(const float* Rcurr, const float* tcurr, const float* vmap_curr, const float* nmap_curr, const float* Rprev_inv,
                  const float* tprev, float fx, float fy, float cx, float cy, const float* vmap_g_prev, const float* nmap_g_prev,
                  float distThres, float angleThres, int rows, int cols, int threads, int blocks, float* A, float* b, float* residual) {
  DeviceArray2D<float> vc, nc, vp, np_;
  up2d(vc, vmap_curr, 3 * rows, cols); up2d(nc, nmap_curr, 3 * rows, cols); up2d(vp, vmap_g_prev, 3 * rows, cols); up2d(np_, nmap_g_prev, 3 * rows, cols);
  DeviceArray<JtJJtrSE3> sum, out; sum.create(MAX_THREADS); out.create(1);
  icpStep(m33(Rcurr), make_float3(tcurr[0], tcurr[1], tcurr[2]), vc, nc, m33(Rprev_inv), make_float3(tprev[0], tprev[1], tprev[2]),
          CameraModel(fx, fy, cx, cy), vp, np_, distThres, angleThres, sum, out, A, b, residual, threads, blocks);
}
// computeRgbResidual (reduce.cu:865-925) + rgbStep (reduce.cu:643-685) + projectToPointCloud (cf.cu:743-757).
// corr_out: rows*cols records {i16 u0, i16 v0, i16 x, i16 y, f32 diff, u8 valid, pad[3]} = the reference's DataTerm.
void ref_rgb(float minScale, const int16_t* dIdx, const int16_t* dIdy, const float* lastDepth, const float* nextDepth,
             const uint8_t* lastImage, const uint8_t* nextImage, float maxDepthDelta, const float* kt, const float* krkinv,
             float fx, float fy, float cx, float cy, float sigma_or_nan, int rows, int cols, int threads_res, int blocks_res,
             int threads_step, int blocks_step, void* corr_out, int* count, int* sigmaSum, float* A, float* b) {
  DeviceArray2D<short> dx, dy; up2d(dx, dIdx, rows, cols); up2d(dy, dIdy, rows, cols);
  DeviceArray2D<float> ld, nd; up2d(ld, lastDepth, rows, cols); up2d(nd, nextDepth, rows, cols);
  DeviceArray2D<unsigned char> li, ni; up2d(li, lastImage, rows, cols); up2d(ni, nextImage, rows, cols);
  DeviceArray2D<DataTerm> corr; corr.create(rows, cols);
  cudaMemset2D(corr.ptr(0), corr.step(), 0, (size_t)cols * sizeof(DataTerm), rows);
  DeviceArray<int2> sumRes; sumRes.create(MAX_THREADS);
  computeRgbResidual(minScale, dx, dy, ld, nd, li, ni, corr, sumRes, maxDepthDelta, make_float3(kt[0], kt[1], kt[2]), m33(krkinv),
                     *sigmaSum, *count, threads_res, blocks_res);
  if (corr_out) {
    // the kernels index corresImg linearly (reduce.cu:562,840), ignoring the pitch
    cudaMemcpy(corr_out, corr.ptr(0), (size_t)rows * cols * sizeof(DataTerm), cudaMemcpyDeviceToHost);
  }
  if (A) {
    DeviceArray2D<float3> cloud; cloud.create(rows, cols);
    CameraModel intr(fx, fy, cx, cy);
    projectToPointCloud(ld, cloud, intr, 0);
    DeviceArray<JtJJtrSE3> sum, out; sum.create(MAX_THREADS); out.create(1);
    float sigma = sigma_or_nan;
    if (sigma != sigma) sigma = sqrtf((float)*count);
    rgbStep(corr, sigma, cloud, fx, fy, dx, dy, 0.125f, sum, out, A, b, threads_step, blocks_step);
  }
}
// so3Step (reduce.cu:1054-1103)
void ref_so3_step(const uint8_t* lastImage, const uint8_t* nextImage, const float* H, const float* kinv, const float* krlr, int rows,
                  int cols, int threads, int blocks, float* A, float* b, float* residual) {
  DeviceArray2D<unsigned char> li, ni; up2d(li, lastImage, rows, cols); up2d(ni, nextImage, rows, cols);
  DeviceArray<JtJJtrSO3> sum, out; sum.create(MAX_THREADS); out.create(1);
  so3Step(li, ni, m33(H), m33(kinv), m33(krlr), sum, out, A, b, residual, threads, blocks);
}

// computeNIDImg (cf.cu:1513-1650) / computeNIDDepth (cf.cu:1794-1915): the reference's own histogram kernels and host entropy loops
float ref_nid_img(const uint8_t* ik, const uint8_t* iko, const float* dk, const float* dko, const uint8_t* ic, int rows, int cols, int nb) {
  DeviceArray2D<unsigned char> a, b, c; DeviceArray2D<float> d, e;
  up2d(a, ik, rows, cols); up2d(b, iko, rows, cols); up2d(c, ic, rows, cols); up2d(d, dk, rows, cols); up2d(e, dko, rows, cols);
  float nid = 0.f;
  computeNIDImg(a, b, d, e, c, nid, nb, false);
  return nid;
}
float ref_nid_depth(const float* dk, const float* dko, const float* dc, int rows, int cols, int nb, float max_depth_mm) {
  DeviceArray2D<float> d, e, c;
  up2d(d, dk, rows, cols); up2d(e, dko, rows, cols); up2d(c, dc, rows, cols);
  float nid = 0.f;
  computeNIDDepth(d, e, c, nid, nb, max_depth_mm, false);
  return nid;
}
}  // extern "C"

// =====================================================================================================
// A stateful stand-in for RGBDOdometry.cpp (which needs Eigen + Pangolin + OpenGL and cannot be built here):
// the reference's buffer set (RO.cpp:44-105) and its host loop (RO.cpp:118-605) driving the reference's own
// kernels call for call -- blocking sync + download after every kernel, per-call cudaMalloc/cudaFree inside
// computeRgbResidual / pyrDownGaussF / pyrDownUcharGauss, GPUConfig default launch shapes (GPUConfig.h:53-60).
// Used for the full-track comparison with the oracle and as the "legacy CUDA" reported baseline.
// =====================================================================================================
#include <cmath>
#include <limits>
#include <algorithm>

namespace {

struct RefOdom {
  int W, H; CameraModel intr; float distT, angT;
  DeviceArray2D<unsigned short> depth_tmp[3];
  DeviceArray<float> vmaps_tmp, nmaps_tmp;
  DeviceArray2D<float> vg[3], ng[3], vc[3], nc[3];
  DeviceArray<JtJJtrSE3> sumSE3, outSE3; DeviceArray<int2> sumRGB; DeviceArray<JtJJtrSO3> sumSO3, outSO3;
  DeviceArray2D<float> lastDepth[3], nextDepth[3];
  DeviceArray2D<unsigned char> lastImage[3], nextImage[3], lastNextImage[3];
  DeviceArray2D<short> dIdx[3], dIdy[3];
  DeviceArray2D<DataTerm> corres[3];
  DeviceArray2D<float3> clouds[3];
  // device-resident inputs (stand-ins for the GL textures)
  static const int NSLOT = 16;
  DeviceArray2D<unsigned short> in_depth[NSLOT]; DeviceArray<float> in_v4[NSLOT], in_n4[NSLOT]; DeviceArray2D<uchar4> in_rgba[NSLOT];
  float lastICPError, lastICPCount, lastRGBError, lastRGBCount, lastSO3Error, lastSO3Count;
  double lastA[36], lastb[6];
};

template <class T, int N>
void ldlt(const T* A, const T* b, T* x) {
  T L[N * N], D[N], y[N];
  for (int i = 0; i < N * N; ++i) L[i] = 0;
  for (int j = 0; j < N; ++j) {
    T d = A[j * N + j];
    for (int k = 0; k < j; ++k) d -= L[j * N + k] * L[j * N + k] * D[k];
    D[j] = d;
    for (int i = j + 1; i < N; ++i) { T s = A[i * N + j]; for (int k = 0; k < j; ++k) s -= L[i * N + k] * L[j * N + k] * D[k]; L[i * N + j] = s / d; }
  }
  for (int i = 0; i < N; ++i) { T s = b[i]; for (int k = 0; k < i; ++k) s -= L[i * N + k] * y[k]; y[i] = s; }
  for (int i = 0; i < N; ++i) y[i] /= D[i];
  for (int i = N - 1; i >= 0; --i) { T s = y[i]; for (int k = i + 1; k < N; ++k) s -= L[k * N + i] * x[k]; x[i] = s; }
}
void mul3(const double* a, const double* b, double* c) { for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) c[i * 3 + j] = a[i * 3] * b[j] + a[i * 3 + 1] * b[3 + j] + a[i * 3 + 2] * b[6 + j]; }
template <class T> void inv3(const T* m, T* o) {
  const T c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
  const T id = T(1) / (m[0] * c00 + m[1] * c01 + m[2] * c02);
  o[0] = c00 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
  o[3] = c01 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
  o[6] = c02 * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}
void rodr(const double* r, double* R) {  // OdometryProvider.h:35-71
  double rx = r[0], ry = r[1], rz = r[2]; const double th = std::sqrt(rx * rx + ry * ry + rz * rz);
  const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  if (th >= 2.2204460492503131e-16) {
    const double c = std::cos(th), s = std::sin(th), c1 = 1. - c, it = 1. / th; rx *= it; ry *= it; rz *= it;
    const double rrt[9] = {rx * rx, rx * ry, rx * rz, rx * ry, ry * ry, ry * rz, rx * rz, ry * rz, rz * rz}, sk[9] = {0, -rz, ry, rz, 0, -rx, -ry, rx, 0};
    for (int k = 0; k < 9; ++k) R[k] = c * I[k] + c1 * rrt[k] + s * sk[k];
  } else for (int k = 0; k < 9; ++k) R[k] = I[k];
}
mat33 m33d(const double* m) { float f[9]; for (int i = 0; i < 9; ++i) f[i] = (float)m[i]; mat33 r; memcpy(r.data, f, sizeof f); return r; }

void populate(RefOdom* o, DeviceArray2D<uchar4>& rgb, DeviceArray2D<float>* dD, DeviceArray2D<unsigned char>* dI) {  // RO.cpp:209-236
  verticesToDepth(o->vmaps_tmp, dD[0], 6.0f);
  for (int i = 0; i + 1 < 3; ++i) pyrDownGaussF(dD[i], dD[i + 1]);
  RefImage img = {rgb.ptr(0), rgb.step()};
  imageBGRToIntensity(reinterpret_cast<cudaArray*>(&img), dI[0]);
  for (int i = 0; i + 1 < 3; ++i) pyrDownUcharGauss(dI[i], dI[i + 1]);
  cudaDeviceSynchronize();
}

}  // namespace

extern "C" {

void* ref_odom_create(int w, int h, float cx, float cy, float fx, float fy, float distT, float angT) {  // RO.cpp:21-111
  RefOdom* o = new RefOdom();
  o->W = w; o->H = h; o->intr = CameraModel(fx, fy, cx, cy); o->distT = distT; o->angT = angT;
  o->sumSE3.create(MAX_THREADS); o->outSE3.create(1); o->sumRGB.create(MAX_THREADS); o->sumSO3.create(MAX_THREADS); o->outSO3.create(1);
  for (int i = 0; i < 3; ++i) {
    const int r = h >> i, c = w >> i;
    o->lastDepth[i].create(r, c); o->lastImage[i].create(r, c); o->nextDepth[i].create(r, c); o->nextImage[i].create(r, c);
    o->lastNextImage[i].create(r, c); o->dIdx[i].create(r, c); o->dIdy[i].create(r, c); o->clouds[i].create(r, c); o->corres[i].create(r, c);
    o->depth_tmp[i].create(r, c); o->vg[i].create(r * 3, c); o->ng[i].create(r * 3, c); o->vc[i].create(r * 3, c); o->nc[i].create(r * 3, c);
    cudaMemset2D(o->lastNextImage[i].ptr(0), o->lastNextImage[i].step(), 0, c, r); cudaMemset2D(o->nextImage[i].ptr(0), o->nextImage[i].step(), 0, c, r);
  }
  o->vmaps_tmp.create((size_t)h * 4 * w); o->nmaps_tmp.create((size_t)h * 4 * w);

  o->lastICPError = o->lastRGBError = o->lastSO3Error = 0; o->lastICPCount = o->lastRGBCount = o->lastSO3Count = (float)(w * h);
  return o;
}
void ref_odom_destroy(void* p) { delete (RefOdom*)p; }
// stage inputs on the device (outside any timed region: the reference gets them as GL textures)
void ref_odom_upload(void* p, int slot, const float* v4, const float* n4, const uint8_t* rgba, const uint16_t* depth) {
  RefOdom* o = (RefOdom*)p;
  if (slot < 0 || slot >= RefOdom::NSLOT) return;
  if (v4) { o->in_v4[slot].create((size_t)o->H * 4 * o->W); o->in_v4[slot].upload(v4, (size_t)o->H * 4 * o->W); }
  if (n4) { o->in_n4[slot].create((size_t)o->H * 4 * o->W); o->in_n4[slot].upload(n4, (size_t)o->H * 4 * o->W); }
  if (rgba) { o->in_rgba[slot].create(o->H, o->W); o->in_rgba[slot].upload(reinterpret_cast<const uchar4*>(rgba), (size_t)o->W * 4, o->H, o->W); }
  if (depth) { o->in_depth[slot].create(o->H, o->W); o->in_depth[slot].upload(depth, (size_t)o->W * 2, o->H, o->W); }
  cudaDeviceSynchronize();
}
void ref_odom_init_icp_depth(void* p, int slot, float cutoff) {  // RO.cpp:118-142
  RefOdom* o = (RefOdom*)p;
  cudaMemcpy2D(o->depth_tmp[0].ptr(0), o->depth_tmp[0].step(), o->in_depth[slot].ptr(0), o->in_depth[slot].step(), (size_t)o->W * 2, o->H, cudaMemcpyDeviceToDevice);
  for (int i = 1; i < 3; ++i) pyrDown(o->depth_tmp[i - 1], o->depth_tmp[i], 0);
  for (int i = 0; i < 3; ++i) { createVMap(o->intr(i), o->depth_tmp[i], o->vc[i], cutoff); createNMap(o->vc[i], o->nc[i]); }
  cudaDeviceSynchronize();
}
void ref_odom_init_icp_model(void* p, int slot, const float* pose) {  // RO.cpp:169-207
  RefOdom* o = (RefOdom*)p;
  cudaMemcpy(o->vmaps_tmp.ptr(), o->in_v4[slot].ptr(), o->vmaps_tmp.sizeBytes(), cudaMemcpyDeviceToDevice);
  cudaMemcpy(o->nmaps_tmp.ptr(), o->in_n4[slot].ptr(), o->nmaps_tmp.sizeBytes(), cudaMemcpyDeviceToDevice);
  copyMaps(o->vmaps_tmp, o->nmaps_tmp, o->vg[0], o->ng[0]);
  for (int i = 1; i < 3; ++i) { resizeVMap(o->vg[i - 1], o->vg[i]); resizeNMap(o->ng[i - 1], o->ng[i]); }
  float R[9] = {pose[0], pose[1], pose[2], pose[4], pose[5], pose[6], pose[8], pose[9], pose[10]};
  mat33 Rm = m33(R); float3 t = make_float3(pose[3], pose[7], pose[11]);
  for (int i = 0; i < 3; ++i) tranformMaps(o->vg[i], o->ng[i], Rm, t, o->vg[i], o->ng[i]);
  cudaDeviceSynchronize();
}
void ref_odom_init_icp_maps(void* p, int slot) {  // RO.cpp:144-167
  RefOdom* o = (RefOdom*)p;
  cudaMemcpy(o->vmaps_tmp.ptr(), o->in_v4[slot].ptr(), o->vmaps_tmp.sizeBytes(), cudaMemcpyDeviceToDevice);
  cudaMemcpy(o->nmaps_tmp.ptr(), o->in_n4[slot].ptr(), o->nmaps_tmp.sizeBytes(), cudaMemcpyDeviceToDevice);
  copyMaps(o->vmaps_tmp, o->nmaps_tmp, o->vc[0], o->nc[0]);
  for (int i = 1; i < 3; ++i) { resizeVMap(o->vc[i - 1], o->vc[i]); resizeNMap(o->nc[i - 1], o->nc[i]); }
  cudaDeviceSynchronize();
}
void ref_odom_init_rgb_model(void* p, int slot) { RefOdom* o = (RefOdom*)p; populate(o, o->in_rgba[slot], o->lastDepth, o->lastImage); }
void ref_odom_init_rgb(void* p, int slot) { RefOdom* o = (RefOdom*)p; populate(o, o->in_rgba[slot], o->nextDepth, o->nextImage); }
void ref_odom_init_first_rgb(void* p, int slot) {  // RO.cpp:250-266
  RefOdom* o = (RefOdom*)p;
  RefImage img = {o->in_rgba[slot].ptr(0), o->in_rgba[slot].step()};
  imageBGRToIntensity(reinterpret_cast<cudaArray*>(&img), o->lastNextImage[0]);
  for (int i = 0; i + 1 < 3; ++i) pyrDownUcharGauss(o->lastNextImage[i], o->lastNextImage[i + 1]);
}

// RO.cpp:268-605.  out[0..5] = lastICPError, lastICPCount, lastRGBError, lastRGBCount, lastSO3Error, lastSO3Count
void ref_odom_track(void* p, float* trans, float* rot, int rgbOnly, float icpWeight, int pyramid, int fastOdom, int so3, int interMap, float* out) {
  RefOdom* o = (RefOdom*)p;
  const bool icp = !rgbOnly && icpWeight > 0, rgb = rgbOnly || icpWeight < 100;
  float Rprev[9], tprev[3], Rcurr[9], tcurr[3];
  memcpy(Rprev, rot, sizeof Rprev); memcpy(tprev, trans, sizeof tprev); memcpy(Rcurr, rot, sizeof Rcurr); memcpy(tcurr, trans, sizeof tcurr);
  if (rgb) for (int i = 0; i < 3; ++i) computeDerivativeImages(o->nextImage[i], o->dIdx[i], o->dIdy[i]);
  double resultR[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  if (so3) {
    const int L = 2; CameraModel k = o->intr(L);
    float R_lr[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    const double K[9] = {k.fx, 0, k.cx, 0, k.fy, k.cy, 0, 0, 1}; double Kinv[9]; inv3(K, Kinv);
    float lastError = std::numeric_limits<float>::max() / 2, lastCount = lastError; double lastR[9]; memcpy(lastR, resultR, sizeof lastR);
    for (int it = 0; it < 10; ++it) {
      double KR[9], Hm[9]; mul3(K, resultR, KR); mul3(KR, Kinv, Hm);
      float jtj[9], jtr[3], residual[2];
      so3Step(o->lastNextImage[L], o->nextImage[L], m33d(Hm), m33d(Kinv), m33d(KR), o->sumSO3, o->outSO3, jtj, jtr, residual, 160, 64);
      o->lastSO3Error = std::sqrt(residual[0]) / residual[1]; o->lastSO3Count = residual[1];
      if (o->lastSO3Error < lastError && lastCount == o->lastSO3Count) break;
      else if (o->lastSO3Error > lastError + 0.001) { o->lastSO3Error = lastError; o->lastSO3Count = lastCount; memcpy(resultR, lastR, sizeof lastR); break; }
      lastError = o->lastSO3Error; lastCount = o->lastSO3Count; memcpy(lastR, resultR, sizeof lastR);
      float delta[3]; ldlt<float, 3>(jtj, jtr, delta);
      const double dd[3] = {delta[0], delta[1], delta[2]}; double ru[9]; rodr(dd, ru);
      float nr[9];
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) nr[r * 3 + c] = (float)ru[r * 3] * R_lr[c] + (float)ru[r * 3 + 1] * R_lr[3 + c] + (float)ru[r * 3 + 2] * R_lr[6 + c];
      memcpy(R_lr, nr, sizeof nr); for (int q = 0; q < 9; ++q) resultR[q] = R_lr[q];
    }
  }
  const int iters[3] = {interMap ? 50 : fastOdom ? 3 : 10, interMap ? 50 : pyramid ? 5 : 0, interMap ? 50 : pyramid ? 4 : 0};
  float Rpi[9]; inv3<float>(Rprev, Rpi);
  double Rt4[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  if (so3) for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) Rt4[r * 4 + c] = resultR[r * 3 + c];
  float residual[2] = {0.f, 0.f};
  for (int i = 2; i >= 0; --i) {
    if (rgb) projectToPointCloud(o->lastDepth[i], o->clouds[i], o->intr, i);
    CameraModel k = o->intr(i);
    const double K[9] = {k.fx, 0, k.cx, 0, k.fy, k.cy, 0, 0, 1}; double Kinv[9]; inv3(K, Kinv);
    o->lastRGBError = std::numeric_limits<float>::max();
    for (int j = 0; j < iters[i]; ++j) {
      double R3[9], Ri[9], ti[3];
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) R3[r * 3 + c] = Rt4[r * 4 + c];
      inv3(R3, Ri);
      for (int r = 0; r < 3; ++r) ti[r] = -(Ri[r * 3] * Rt4[3] + Ri[r * 3 + 1] * Rt4[7] + Ri[r * 3 + 2] * Rt4[11]);
      double KR[9], KRKi[9]; mul3(K, Ri, KR); mul3(KR, Kinv, KRKi);
      const float3 kt = make_float3((float)(K[0] * ti[0] + K[1] * ti[1] + K[2] * ti[2]), (float)(K[3] * ti[0] + K[4] * ti[1] + K[5] * ti[2]), (float)(K[6] * ti[0] + K[7] * ti[1] + K[8] * ti[2]));
      int sigma = 0, rgbSize = 0;
      if (rgb) {
        const float g = i == 0 ? 5.f : i == 1 ? 3.f : 1.f;
        computeRgbResidual((float)(std::pow(g, 2.0) / std::pow(0.125, 2.0)), o->dIdx[i], o->dIdy[i], o->lastDepth[i], o->nextDepth[i], o->lastImage[i], o->nextImage[i],
                           o->corres[i], o->sumRGB, 0.07f, kt, m33d(KRKi), sigma, rgbSize, 256, 336);
      }
      float sigmaVal = std::sqrt((float)sigma / rgbSize == 0 ? 1 : rgbSize);
      if (rgbOnly && std::sqrt((double)sigma) / rgbSize > o->lastRGBError) break;
      o->lastRGBError = std::sqrt((double)sigma) / rgbSize; o->lastRGBCount = rgbSize;
      if (rgbOnly) sigmaVal = -1;
      float A_icp[36] = {0}, b_icp[6] = {0}, A_rgb[36] = {0}, b_rgb[6] = {0};
      if (icp) icpStep(m33(Rcurr), make_float3(tcurr[0], tcurr[1], tcurr[2]), o->vc[i], o->nc[i], m33(Rpi), make_float3(tprev[0], tprev[1], tprev[2]), k, o->vg[i], o->ng[i],
                       o->distT, o->angT, o->sumSE3, o->outSE3, A_icp, b_icp, residual, 128, 112);
      o->lastICPError = std::sqrt(residual[0]) / residual[1]; o->lastICPCount = residual[1];
      if (rgb) rgbStep(o->corres[i], sigmaVal, o->clouds[i], k.fx, k.fy, o->dIdx[i], o->dIdy[i], 0.125f, o->sumSE3, o->outSE3, A_rgb, b_rgb, 128, 112);
      double x[6];
      if (icp && rgb) { const double w = icpWeight; for (int q = 0; q < 36; ++q) o->lastA[q] = (double)A_rgb[q] + w * w * (double)A_icp[q]; for (int q = 0; q < 6; ++q) o->lastb[q] = (double)b_rgb[q] + w * (double)b_icp[q]; }
      else if (icp) { for (int q = 0; q < 36; ++q) o->lastA[q] = A_icp[q]; for (int q = 0; q < 6; ++q) o->lastb[q] = b_icp[q]; }
      else { for (int q = 0; q < 36; ++q) o->lastA[q] = A_rgb[q]; for (int q = 0; q < 6; ++q) o->lastb[q] = b_rgb[q]; }
      ldlt<double, 6>(o->lastA, o->lastb, x);
      double Ru[9]; rodr(x + 3, Ru);
      const double U[16] = {Ru[0], Ru[1], Ru[2], x[0], Ru[3], Ru[4], Ru[5], x[1], Ru[6], Ru[7], Ru[8], x[2], 0, 0, 0, 1};
      double nRt[16];
      for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) nRt[r * 4 + c] = U[r * 4] * Rt4[c] + U[r * 4 + 1] * Rt4[4 + c] + U[r * 4 + 2] * Rt4[8 + c] + U[r * 4 + 3] * Rt4[12 + c];
      memcpy(Rt4, nRt, sizeof nRt);
      float oR[9], ot[3], iR[9], it_[3];
      for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) oR[r * 3 + c] = (float)Rt4[r * 4 + c]; ot[r] = (float)Rt4[r * 4 + 3]; }
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) iR[r * 3 + c] = oR[c * 3 + r];
      for (int r = 0; r < 3; ++r) it_[r] = -(iR[r * 3] * ot[0] + iR[r * 3 + 1] * ot[1] + iR[r * 3 + 2] * ot[2]);
      for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) Rcurr[r * 3 + c] = Rprev[r * 3] * iR[c] + Rprev[r * 3 + 1] * iR[3 + c] + Rprev[r * 3 + 2] * iR[6 + c];
        tcurr[r] = Rprev[r * 3] * it_[0] + Rprev[r * 3 + 1] * it_[1] + Rprev[r * 3 + 2] * it_[2] + tprev[r];
      }
    }
  }
  if (rgb) { const float dx = tcurr[0] - tprev[0], dy = tcurr[1] - tprev[1], dz = tcurr[2] - tprev[2]; if (std::sqrt(dx * dx + dy * dy + dz * dz) > 0.3) { memcpy(Rcurr, Rprev, sizeof Rcurr); memcpy(tcurr, tprev, sizeof tcurr); } }
  if (so3) for (int i = 0; i < 3; ++i) std::swap(o->lastNextImage[i], o->nextImage[i]);
  memcpy(trans, tcurr, sizeof tcurr); memcpy(rot, Rcurr, sizeof Rcurr);
  if (out) { out[0] = o->lastICPError; out[1] = o->lastICPCount; out[2] = o->lastRGBError; out[3] = o->lastRGBCount; out[4] = o->lastSO3Error; out[5] = o->lastSO3Count; }
}

}  // extern "C"
