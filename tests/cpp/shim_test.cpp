// Exercises the C++ drop-in surface the way GPUTest.cpp:215,260-283 and ElasticFusion.cpp:173-201 call it.
// Usage: shim_test <dir>  -- reads <dir>/{vmap4,nmap4,model_rgba,live_rgba,depth}.bin + meta.txt, writes <dir>/out.txt
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <vector>
#include "dmsodom/RGBDOdometry.h"

static std::vector<char> slurp(const std::string& p) {
  std::ifstream f(p.c_str(), std::ios::binary); std::vector<char> b((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>()); return b;
}
static void* to_dev(const std::vector<char>& b) { void* d = 0; cudaMalloc(&d, b.size()); cudaMemcpy(d, b.data(), b.size(), cudaMemcpyHostToDevice); return d; }

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  const std::string dir = argv[1];
  int w, h; float cx, cy, fx, fy;
  { std::ifstream m((dir + "/meta.txt").c_str()); m >> w >> h >> cx >> cy >> fx >> fy; }
  GPUTexture V(to_dev(slurp(dir + "/vmap4.bin")), (size_t)w * 16, w, h), N(to_dev(slurp(dir + "/nmap4.bin")), (size_t)w * 16, w, h);
  GPUTexture M(to_dev(slurp(dir + "/model_rgba.bin")), (size_t)w * 4, w, h), L(to_dev(slurp(dir + "/live_rgba.bin")), (size_t)w * 4, w, h);
  GPUTexture D(to_dev(slurp(dir + "/depth.bin")), (size_t)w * 2, w, h);
  RGBDOdometry odom(w, h, cx, cy, fx, fy);  // note the reference's argument order: cx, cy, fx, fy
  Eigen::Matrix4f currPose = Eigen::Matrix4f::Identity();
  odom.initFirstRGB(&M);
  odom.initICPModel(&V, &N, 25.0f, currPose);
  odom.initRGBModel(&M);
  odom.initICP(&D, 25.0f);
  odom.initRGB(&L);
  Eigen::Vector3f trans; Eigen::Matrix<float, 3, 3, Eigen::RowMajor> rot = Eigen::Matrix<float, 3, 3, Eigen::RowMajor>::Identity();
  odom.getIncrementalTransformation(trans, rot, false, 10.0f, true, false, true);
  Eigen::MatrixXd cov = odom.getCovariance();
  std::vector<unsigned char> img((size_t)odom.nextImg(0).rows() * odom.nextImg(0).cols());
  odom.nextImg(0).download(img.data(), odom.nextImg(0).cols());
  unsigned long sum = 0; for (size_t i = 0; i < img.size(); ++i) sum += img[i];
  FILE* f = fopen((dir + "/out.txt").c_str(), "w");
  fprintf(f, "%.9g %.9g %.9g\n", trans(0), trans(1), trans(2));
  for (int i = 0; i < 3; ++i) fprintf(f, "%.9g %.9g %.9g\n", rot(i, 0), rot(i, 1), rot(i, 2));
  fprintf(f, "%.9g %.9g %.9g %.9g %.9g %.9g\n", odom.lastICPError, odom.lastICPCount, odom.lastRGBError, odom.lastRGBCount, odom.lastSO3Error, odom.lastSO3Count);
  fprintf(f, "%.17g %.17g %lu\n", odom.lastA(0, 0), cov(0, 0), sum);
  const std::map<std::string, std::vector<float> >& t = Stopwatch::getInstance().getTimings();
  fprintf(f, "%d %d %d %d\n", (int)t.count("icpStep"), (int)t.count("rgbStep"), (int)t.count("computeRgbResidual"), (int)t.count("so3Step"));
  fclose(f);
  return 0;
}
