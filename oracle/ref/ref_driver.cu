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

}  // extern "C"
