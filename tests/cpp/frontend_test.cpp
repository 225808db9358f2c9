// Exercises the C++ mirrors of IndexMap::combinedPredict, the processFrame tracking block and MutualInformation the way
// ElasticFusion.cpp:158-201 / 639-662 use them.  Usage: frontend_test <dir> -- reads <dir>/{surfels,live0_rgba,live0_depth,
// live1_rgba,live1_depth,prior1}.bin + meta.txt, writes <dir>/out_fe.txt
#include <cstdio>
#include <fstream>
#include <vector>
#include "dmsodom/ElasticFusionFrontEnd.h"

static std::vector<char> slurp(const std::string& p) {
  std::ifstream f(p.c_str(), std::ios::binary); std::vector<char> b((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>()); return b;
}
static void* to_dev(const std::vector<char>& b) { void* d = 0; cudaMalloc(&d, b.size()); cudaMemcpy(d, b.data(), b.size(), cudaMemcpyHostToDevice); return d; }

struct FrontEnd : dmsodom::TrackingFrontEnd {
  dmsodom::IndexMap& map; const float* d_model; int count; float cutoff;
  FrontEnd(int w, int h, float cx, float cy, float fx, float fy, const dms_frontend_options* o, dmsodom::IndexMap& m, const float* model, int n, float c)
      : dmsodom::TrackingFrontEnd(w, h, cx, cy, fx, fy, o), map(m), d_model(model), count(n), cutoff(c) {}
  bool predict(const Eigen::Matrix4f& pose, int kind, int tick, cudaStream_t stream, View* out) {
    if (kind == DMS_PREDICT_INACTIVE) return false;
    map.combinedPredict(pose, d_model, count, cutoff, 5.0f, tick, 0, tick, 200, dmsodom::IndexMap::ACTIVE, stream);   // EF.cpp:718-731
    out->vertex = map.vertexTex(); out->normal = map.normalTex(); out->image = map.imageTex();
    return true;
  }
};

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  const std::string dir = argv[1];
  int w, h; float cx, cy, fx, fy;
  { std::ifstream m((dir + "/meta.txt").c_str()); m >> w >> h >> cx >> cy >> fx >> fy; }
  const std::vector<char> sb = slurp(dir + "/surfels.bin");
  const int count = (int)(sb.size() / (DMS_SURFEL_FLOATS * sizeof(float)));
  const float* d_model = (const float*)to_dev(sb);
  GPUTexture rgb0(to_dev(slurp(dir + "/live0_rgba.bin")), (size_t)w * 4, w, h), d0(to_dev(slurp(dir + "/live0_depth.bin")), (size_t)w * 2, w, h);
  GPUTexture rgb1(to_dev(slurp(dir + "/live1_rgba.bin")), (size_t)w * 4, w, h), d1(to_dev(slurp(dir + "/live1_depth.bin")), (size_t)w * 2, w, h);
  const std::vector<char> pb = slurp(dir + "/prior1.bin");
  Eigen::Matrix4f prior;
  for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) prior(r, c) = ((const float*)pb.data())[r * 4 + c];
  dms_frontend_options opt;
  dms_frontend_default_options(&opt);
  opt.depthCutoff = 3.5f; opt.closeLoops = 0;
  dmsodom::IndexMap map(w, h, cx, cy, fx, fy);
  FrontEnd fe(w, h, cx, cy, fx, fy, &opt, map, d_model, count, 3.5f);
  Eigen::Matrix4f I = Eigen::Matrix4f::Identity();
  const dms_frontend_result r0 = fe.processFrame(&rgb0, &d0, &I);
  const dms_frontend_result r1 = fe.processFrame(&rgb1, &d1, &prior);
  // fuseFrame's NID score of the tracked frame against the prediction at its pose (EF.cpp:645-662)
  dmsodom::MutualInformation mi(w, h, 500, 64, 1);
  map.combinedPredict(dmsodom::from_rowmajor(r1.pose), d_model, count, 3.5f, 5.0f, 2, 0, 2, 200, dmsodom::IndexMap::ACTIVE);
  mi.setKeyFrame(*map.imageTex(), *map.vertexTex(), *map.imageTex(), *map.vertexTex(), 3.5f);
  dms_odom* f2m = 0;
  dms_frontend_get_tracker(fe.handle(), 0, &f2m);
  const float score = mi.nidScore(f2m, 3.5f, 0.5f);
  FILE* f = fopen((dir + "/out_fe.txt").c_str(), "w");
  fprintf(f, "%d %d %d %d\n", r0.tick, r0.tracked, r1.tick, r1.tracked);
  for (int i = 0; i < 16; ++i) fprintf(f, "%.9g%c", r1.pose[i], i % 4 == 3 ? '\n' : ' ');
  fprintf(f, "%.9g %.9g %.9g %d %d\n", r1.frameToModel.lastICPCount, r1.frameToModel.lastRGBCount, r1.weighting, r1.shouldFillIn, r1.trackingOk);
  fprintf(f, "%.9g %.9g %.9g\n", score, mi.lastNIDImgScore, mi.lastNIDDepthScore);
  fclose(f);
  return 0;
}
