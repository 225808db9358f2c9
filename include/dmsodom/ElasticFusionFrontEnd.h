// C++ mirrors (header only) of the reference classes either side of the tracker, over the C ABI:
//   IndexMap::combinedPredict          (IndexMap.h:66-76, IndexMap.cpp:253-368)        -> dmsodom::IndexMap
//   MutualInformation::nidImg/nidDepth (MutualInformation.h, MutualInformation.cpp:154-212) with the KeyFrame it scores
//                                       against (KeyFrame.h:69-160)                     -> dmsodom::MutualInformation
//   the tracking block of ElasticFusion::processFrame (ElasticFusion.cpp:99-268, 397-442) -> dmsodom::TrackingFrontEnd
// Same headless GPUTexture (device pointer + pitch) and mini Eigen types as RGBDOdometry.h.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>
#include "RGBDOdometry.h"
#include "frontend_c.h"
#include "nid_c.h"
#include "splat_c.h"

namespace dmsodom {

inline void ck(int rc, const char* what) {
  if (rc != DMS_OK) throw std::runtime_error(std::string(what) + " failed (" + std::to_string(rc) + ")");
}
inline void to_rowmajor(const Eigen::Matrix4f& m, float p[16]) {
  for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) p[r * 4 + c] = m(r, c);
}
inline Eigen::Matrix4f from_rowmajor(const float p[16]) {
  Eigen::Matrix4f m;
  for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m(r, c) = p[r * 4 + c];
  return m;
}

// The four output textures of IndexMap (image RGBA8, vertex / normal RGBA32F, time R32UI), owned here.
class IndexMap {
 public:
  enum Prediction { ACTIVE, INACTIVE };
  IndexMap(int width, int height, float cx, float cy, float fx, float fy) : w_(width), h_(height), cx_(cx), cy_(cy), fx_(fx), fy_(fy) {
    const size_t n = (size_t)width * height;
    alloc(&img_, n * 4); alloc(&ver_, n * 16); alloc(&nor_, n * 16); alloc(&tim_, n * 4); alloc(&zbuf_, n * 8);
    imageTexture = GPUTexture(img_, (size_t)width * 4, width, height);
    vertexTexture = GPUTexture(ver_, (size_t)width * 16, width, height);
    normalTexture = GPUTexture(nor_, (size_t)width * 16, width, height);
    timeTexture = GPUTexture(tim_, (size_t)width * 4, width, height);
  }
  ~IndexMap() { cudaFree(img_); cudaFree(ver_); cudaFree(nor_); cudaFree(tim_); cudaFree(zbuf_); }
  // model = device pointer to `count` surfels of Vertex::SIZE bytes (the mapped vertex buffer of GlobalModel::model())
  void combinedPredict(const Eigen::Matrix4f& pose, const float* d_model, int count, const float depthCutoff, const float confThreshold, const int time,
                       const int timeIdx, const int maxTime, const int timeDelta, Prediction predictionType, cudaStream_t stream = 0) {
    dms_splat_params p;
    float pm[16];
    to_rowmajor(pose, pm);
    ck(dms_splat_invert_pose(pm, p.t_inv), "dms_splat_invert_pose");
    p.cx = cx_; p.cy = cy_; p.fx = fx_; p.fy = fy_; p.rows = h_; p.cols = w_; p.maxDepth = depthCutoff; p.confThreshold = confThreshold;
    p.time = time; p.timeIdx = timeIdx; p.maxTime = maxTime; p.timeDelta = timeDelta; p.active = predictionType == ACTIVE;
    ck(dms_splat_predict(d_model, count, &p, (unsigned long long*)zbuf_, (uint8_t*)img_, (float*)ver_, (float*)nor_, (uint32_t*)tim_, stream), "dms_splat_predict");
  }
  GPUTexture* imageTex() { return &imageTexture; }
  GPUTexture* vertexTex() { return &vertexTexture; }
  GPUTexture* normalTex() { return &normalTexture; }
  GPUTexture* timeTex() { return &timeTexture; }

 private:
  IndexMap(const IndexMap&);
  IndexMap& operator=(const IndexMap&);
  static void alloc(void** p, size_t bytes) { if (cudaMalloc(p, bytes) != cudaSuccess) throw std::runtime_error("cudaMalloc failed"); }
  int w_, h_;
  float cx_, cy_, fx_, fy_;
  void *img_, *ver_, *nor_, *tim_, *zbuf_;
  GPUTexture imageTexture, vertexTexture, normalTexture, timeTexture;
};

class MutualInformation {
 public:
  MutualInformation(int width, int height, int num_bins_depth, int num_bins_img, int pyr_level) : h_(0), lastNIDImgScore(0), lastNIDDepthScore(0) {
    int dev = 0;
    cudaGetDevice(&dev);
    ck(dms_nid_create(&h_, dev, width, height, pyr_level, num_bins_img, num_bins_depth), "dms_nid_create");
  }
  ~MutualInformation() { dms_nid_destroy(h_); }
  // KeyFrame(activeImageTex, activeVertexTex, ., oldImageTex, oldVertexTex, ., pose, depthCutoff)  KeyFrame.h:69-160
  void setKeyFrame(GPUTexture& activeImageTex, GPUTexture& activeVertexTex, GPUTexture& oldImageTex, GPUTexture& oldVertexTex, float depthCutoff,
                   cudaStream_t stream = 0) {
    ck(dms_nid_set_keyframe(h_, (const float*)activeVertexTex.dptr, (const uint8_t*)activeImageTex.dptr, (const float*)oldVertexTex.dptr,
                            (const uint8_t*)oldImageTex.dptr, depthCutoff, stream), "dms_nid_set_keyframe");
  }
  // nid_score of ElasticFusion::fuseFrame (EF.cpp:651-662) against tracker.nextImg(level) / nextD(level)
  float nidScore(RGBDOdometry& tracker, float max_depth, float nid_depth_lambda, cudaStream_t stream = 0) {
    return nidScore(tracker.handle(), max_depth, nid_depth_lambda, stream);
  }
  float nidScore(dms_odom* tracker, float max_depth, float nid_depth_lambda, cudaStream_t stream = 0) {
    float s = 0.f;
    ck(dms_nid_score(h_, tracker, max_depth, nid_depth_lambda, &lastNIDImgScore, &lastNIDDepthScore, &s, stream), "dms_nid_score");
    return s;
  }
  float lastNIDImgScore, lastNIDDepthScore;

 private:
  MutualInformation(const MutualInformation&);
  MutualInformation& operator=(const MutualInformation&);
  dms_nid* h_;
};

// The Context-side state and tracking block of processFrame.  Subclass and implement predict() with the map in use
// (e.g. an IndexMap above over the surfel buffer); return false when the view does not exist yet.
class TrackingFrontEnd {
 public:
  struct View { GPUTexture* vertex; GPUTexture* normal; GPUTexture* image; };
  TrackingFrontEnd(int width, int height, float cx, float cy, float fx, float fy, const dms_frontend_options* opt = 0) : h_(0) {
    int dev = 0;
    cudaGetDevice(&dev);
    ck(dms_frontend_create(&h_, dev, width, height, cx, cy, fx, fy, opt), "dms_frontend_create");
    ck(dms_frontend_set_predictor(h_, &TrackingFrontEnd::trampoline, this), "dms_frontend_set_predictor");
  }
  virtual ~TrackingFrontEnd() { dms_frontend_destroy(h_); }
  // kind: dms_predict_kind; tick = Context::tick()
  virtual bool predict(const Eigen::Matrix4f& pose, int kind, int tick, cudaStream_t stream, View* out) = 0;
  dms_frontend_result processFrame(GPUTexture* rgb, GPUTexture* depthRaw, const Eigen::Matrix4f* inPose, cudaStream_t stream = 0) {
    dms_frontend_result r;
    float p[16];
    if (inPose) to_rowmajor(*inPose, p);
    ck(dms_frontend_process(h_, (const uint8_t*)rgb->dptr, rgb->pitch, (const uint16_t*)depthRaw->dptr, depthRaw->pitch, inPose ? p : 0, &r, stream),
       "dms_frontend_process");
    return r;
  }
  dms_frontend* handle() { return h_; }

 private:
  TrackingFrontEnd(const TrackingFrontEnd&);
  TrackingFrontEnd& operator=(const TrackingFrontEnd&);
  static int trampoline(void* user, const float pose[16], int kind, int tick, void* stream, dms_model_view* out) {
    View v = {0, 0, 0};
    if (!static_cast<TrackingFrontEnd*>(user)->predict(from_rowmajor(pose), kind, tick, (cudaStream_t)stream, &v) || !v.vertex || !v.normal || !v.image) return 1;
    out->d_vertex = (const float*)v.vertex->dptr; out->d_normal = (const float*)v.normal->dptr; out->d_image = (const uint8_t*)v.image->dptr;
    return 0;
  }
  dms_frontend* h_;
};

}  // namespace dmsodom
