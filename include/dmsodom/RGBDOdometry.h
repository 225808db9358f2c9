/*
 * Drop-in C++ surface of ElasticFusion's RGBDOdometry (reference: Core/src/Utils/RGBDOdometry.h:32-153) on top of
 * the dmsodom C ABI (dmsodom/odom_c.h).  Same class name, constructor argument order, method names and argument
 * meaning, public fields and accessors, so Context / ReferenceFrame / Ferns / GPUTest compile against it unchanged.
 *
 * Differences that the missing graphics stack forces:
 *  - GPUTexture is a headless description of a device buffer ({dptr, pitch, width, height}); with a GL build the
 *    reference's GPUTexture would be mapped to such a pointer by the caller (RO.cpp:122-128 does the same mapping).
 *  - DeviceArray2D<T> returned by nextImg()/lastImg()/nextD()/lastVmap() is a non-owning view of the tracker's
 *    pitched device buffer (ptr(), step(), rows(), cols(), download()), the subset the reference's callers use.
 *  - Stopwatch: the four keys the reference TICK/TOCKs (RO.cpp:336,444,495,525) are filled from the kernel's own
 *    timers after every getIncrementalTransformation.
 */
#ifndef DMSODOM_RGBDODOMETRY_H_
#define DMSODOM_RGBDODOMETRY_H_

#include <cmath>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>
#include <cuda_runtime_api.h>

#if defined(__has_include)
#if __has_include(<Eigen/Core>)
#include <Eigen/Core>
#include <Eigen/Dense>
#define DMSODOM_HAVE_EIGEN 1
#endif
#endif
#ifndef DMSODOM_HAVE_EIGEN
#include "dmsodom/mini_eigen.h"
#endif
#include "dmsodom/odom_c.h"

struct GPUTexture {  /* headless stand-in for Core/src/GPUTexture.h */
  void* dptr; size_t pitch; int width, height;
  GPUTexture(void* p = 0, size_t pitchBytes = 0, int w = 0, int h = 0) : dptr(p), pitch(pitchBytes), width(w), height(h) {}
};

template <class T>
class DeviceArray2D {  /* view with the accessors of Cuda/containers/device_array.hpp used outside the tracker */
 public:
  DeviceArray2D() : p_(0), step_(0), rows_(0), cols_(0) {}
  DeviceArray2D(void* p, size_t step, int rows, int cols) : p_((char*)p), step_(step), rows_(rows), cols_(cols) {}
  T* ptr(int y = 0) { return (T*)(p_ + (size_t)y * step_); }
  const T* ptr(int y = 0) const { return (const T*)(p_ + (size_t)y * step_); }
  size_t step() const { return step_; }
  int rows() const { return rows_; }
  int cols() const { return cols_; }
  size_t colsBytes() const { return (size_t)cols_ * sizeof(T); }
  void download(void* host, size_t hostStep) const { cudaMemcpy2D(host, hostStep, p_, step_, colsBytes(), rows_, cudaMemcpyDeviceToHost); }
 private:
  char* p_; size_t step_; int rows_, cols_;
};

class Stopwatch {  /* subset of Utils/Stopwatch.h read by GPUTest.cpp:280-283 */
 public:
  static Stopwatch& getInstance() { static Stopwatch s; return s; }
  std::map<std::string, std::vector<float> >& getTimings() { return timings_; }
  void add(const std::string& k, float ms) { timings_[k].push_back(ms); }
 private:
  std::map<std::string, std::vector<float> > timings_;
};

class RGBDOdometry {
 public:
  RGBDOdometry(int width, int height, float cx, float cy, float fx, float fy, float distThresh = 0.10f,
               float angleThresh = sin(20.f * 3.14159254f / 180.f))
      : lastICPError(0), lastICPCount((float)(width * height)), lastRGBError(0), lastRGBCount((float)(width * height)),
        lastSO3Error(0), lastSO3Count((float)(width * height)), h_(0), width(width), height(height) {
    lastA = Eigen::Matrix<double, 6, 6, Eigen::RowMajor>::Zero();
    lastb = Eigen::Matrix<double, 6, 1>::Zero();
    int dev = 0;
    cudaGetDevice(&dev);
    check(dms_odom_create(&h_, dev, width, height, cx, cy, fx, fy, distThresh, angleThresh), "dms_odom_create");
    dms_odom_set_timing(h_, 1);
  }
  virtual ~RGBDOdometry() { dms_odom_destroy(h_); }

  void initICP(GPUTexture* filteredDepth, const float depthCutoff, cudaStream_t stream = 0) {
    check(dms_odom_init_icp_depth(h_, (const uint16_t*)filteredDepth->dptr, filteredDepth->pitch, depthCutoff, stream), "initICP");
  }
  void initICP(GPUTexture* predictedVertices, GPUTexture* predictedNormals, const float depthCutoff, cudaStream_t stream = 0) {
    check(dms_odom_init_icp_maps(h_, (const float*)predictedVertices->dptr, (const float*)predictedNormals->dptr, predictedVertices->pitch,
                                 depthCutoff, stream), "initICP");
  }
  void initICPModel(GPUTexture* predictedVertices, GPUTexture* predictedNormals, const float depthCutoff, const Eigen::Matrix4f& modelPose,
                    cudaStream_t stream = 0) {
    float p[16];
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) p[r * 4 + c] = modelPose(r, c);
    check(dms_odom_init_icp_model(h_, (const float*)predictedVertices->dptr, (const float*)predictedNormals->dptr, predictedVertices->pitch,
                                  depthCutoff, p, stream), "initICPModel");
  }
  void initRGB(GPUTexture* rgb, cudaStream_t stream = 0) { check(dms_odom_init_rgb(h_, (const uint8_t*)rgb->dptr, rgb->pitch, stream), "initRGB"); }
  void initRGBModel(GPUTexture* rgb, cudaStream_t stream = 0) { check(dms_odom_init_rgb_model(h_, (const uint8_t*)rgb->dptr, rgb->pitch, stream), "initRGBModel"); }
  void initFirstRGB(GPUTexture* rgb, cudaStream_t s = 0) { check(dms_odom_init_first_rgb(h_, (const uint8_t*)rgb->dptr, rgb->pitch, s), "initFirstRGB"); }

  void getIncrementalTransformation(Eigen::Vector3f& trans, Eigen::Matrix<float, 3, 3, Eigen::RowMajor>& rot, const bool& rgbOnly,
                                    const float& icpWeight, const bool& pyramid, const bool& fastOdom, const bool& so3,
                                    const bool interMap = false, cudaStream_t stream = 0) {
    dms_odom_result r;
    check(dms_odom_track(h_, trans.data(), rot.data(), rgbOnly, icpWeight, pyramid, fastOdom, so3, interMap, stream, &r),
          "getIncrementalTransformation");
    lastICPError = r.lastICPError; lastICPCount = r.lastICPCount; lastRGBError = r.lastRGBError; lastRGBCount = r.lastRGBCount;
    lastSO3Error = r.lastSO3Error; lastSO3Count = r.lastSO3Count;
    for (int i = 0; i < 6; ++i) { for (int j = 0; j < 6; ++j) lastA(i, j) = r.lastA[i * 6 + j]; lastb(i) = r.lastb[i]; }
    dms_odom_timing t;
    if (dms_odom_get_timing(h_, &t) == 0) {
      Stopwatch& sw = Stopwatch::getInstance();
      sw.add("so3Step", t.so3Step_ms); sw.add("computeRgbResidual", t.computeRgbResidual_ms); sw.add("icpStep", t.icpStep_ms); sw.add("rgbStep", t.rgbStep_ms);
    }
  }

  void getICPResidual(Eigen::Vector3f& trans_curr, Eigen::Matrix<float, 3, 3, Eigen::RowMajor>& rot_curr, Eigen::Vector3f& trans_prev,
                      Eigen::Matrix<float, 3, 3, Eigen::RowMajor>& rot_prev, float& icpError, int& icpCount) {
    check(dms_odom_icp_residual(h_, trans_curr.data(), rot_curr.data(), trans_prev.data(), rot_prev.data(), &icpError, &icpCount, 0), "getICPResidual");
  }

  Eigen::MatrixXd getCovariance() {
    double c[36];
    check(dms_odom_get_covariance(h_, c), "getCovariance");
    Eigen::MatrixXd m(6, 6);
    for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) m(i, j) = c[i * 6 + j];
    return m;
  }

  float lastICPError, lastICPCount, lastRGBError, lastRGBCount, lastSO3Error, lastSO3Count;
  Eigen::Matrix<double, 6, 6, Eigen::RowMajor> lastA;
  Eigen::Matrix<double, 6, 1> lastb;

  DeviceArray2D<unsigned char>& nextImg(int pyr_level = 0) { return view(nextImage_, DMS_BUF_NEXT_IMAGE, pyr_level); }
  DeviceArray2D<unsigned char>& lastImg(int pyr_level = 0) { return view(lastImage_, DMS_BUF_LAST_IMAGE, pyr_level); }
  DeviceArray2D<float>& nextD(int pyr_level = 0) { return view(nextDepth_, DMS_BUF_NEXT_DEPTH, pyr_level); }
  DeviceArray2D<float>& lastVmap(int pyr_level = 0) { return view(lastVmap_, DMS_BUF_VMAP_G_PREV, pyr_level); }

  dms_odom* handle() { return h_; }

 private:
  template <class T>
  DeviceArray2D<T>& view(DeviceArray2D<T> (&cache)[3], int which, int level) {
    void* p; size_t pitch; int rows, cols, elem;
    check(dms_odom_get_buffer(h_, which, level, &p, &pitch, &rows, &cols, &elem), "get_buffer");
    cache[level] = DeviceArray2D<T>(p, pitch, rows, cols);
    return cache[level];
  }
  void check(int rc, const char* what) {
    if (rc != 0) throw std::runtime_error(std::string(what) + ": " + (h_ ? dms_odom_last_error(h_) : "no handle") + " (" + std::to_string(rc) + ")");
  }
  dms_odom* h_;
  DeviceArray2D<unsigned char> nextImage_[3], lastImage_[3];
  DeviceArray2D<float> nextDepth_[3], lastVmap_[3];
  const int width, height;

 public:
  EIGEN_MAKE_ALIGNED_OPERATOR_NEW
};

#endif /* DMSODOM_RGBDODOMETRY_H_ */
