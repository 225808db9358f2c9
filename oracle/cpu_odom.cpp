// =====================================================================================
// oracle/cpu_odom.cpp  --  TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement of the dense tracking hot path of robotvisionmu/DenseMonoSLAM
// (ElasticFusion's RGBDOdometry).  It is the checker the CUDA path is compared with.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load this file's shared object; the product library (libdmsodom.so) never
// links, loads or calls it.
//
// Parity status: the reference ships no golden vectors / tests for this path
// (SURVEY.md section 4), so this restatement is pinned two ways:
//   * against the reference's own legacy kernels built from /root/reference into
//     oracle/_ref/ and run on the GPU box (oracle/ref_build.sh, tests/test_legacy_ref.py)
//   * self-consistency (known synthetic motion is recovered, identity -> zero update).
//
// Every function cites the reference file:line it follows.  Abbreviations:
//   RO.cpp    = elasticfusion/Core/src/Utils/RGBDOdometry.cpp
//   reduce.cu = elasticfusion/Core/src/Cuda/reduce.cu
//   cf.cu     = elasticfusion/Core/src/Cuda/cudafuncs.cu
//   OP.h      = elasticfusion/Core/src/Utils/OdometryProvider.h
//
// Arithmetic contract (shared, by construction and not by shared code, with the CUDA
// kernels in densemonoslam_b200/csrc):
//   * fp32 IEEE-754 round-to-nearest everywhere; the fused multiply-adds that nvcc's
//     default -fmad=true produces for the reference's expressions are written out as
//     explicit fmaf() (pattern checked on nvcc 12.9 SASS: a*b + c*d + e*f becomes
//     fma(e,f, fma(a,b, c*d))).  Everything else is a separately rounded operation;
//     this file must be compiled with -ffp-contract=off.
//   * divisions and square roots are IEEE (the reference's --prec-div=false,
//     --prec-sqrt=false, rsqrtf approximations are NOT reproduced; the difference is
//     measured against oracle/_ref, not gated).
//   * float -> int pixel rounding = CUDA __float2int_rn (ties to even, NaN -> 0,
//     saturating).
//   * reductions: "canonical tree" -- every image row is cut into segments of 32
//     consecutive pixels; product (i, j) is accumulated over the 32 pixels of a segment
//     in order (pixel 0 first) as s = fmaf(row[i], row[j], s) in fp32, and segment
//     sums are accumulated EXACTLY in a 2x64-bit fixed-point accumulator, so the result
//     does not depend on how segments are scheduled.  Counts are integers.
// =====================================================================================
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// ------------------------------------------------------------------ small math helpers
struct V3 { float x, y, z; };
struct M33 { V3 r[3]; };  // row-major rows, like mat33 (types.cuh:61-74)

static inline V3 v3(float x, float y, float z) { V3 v = {x, y, z}; return v; }
static inline V3 sub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }  // operators.cuh:57-60
static inline V3 add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }  // operators.cuh:62-65
// operators.cuh:72-75  a.x*b.x + a.y*b.y + a.z*b.z ; nvcc: fma(az,bz, fma(ax,bx, ay*by))
static inline float dot3(V3 a, V3 b) { return fmaf(a.z, b.z, fmaf(a.x, b.x, a.y * b.y)); }
// operators.cuh:67-70 ; nvcc: fma(a.y,b.z, -(a.z*b.y)) ...
static inline V3 cross3(V3 a, V3 b) {
  return v3(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
static inline float norm3(V3 a) { return sqrtf(dot3(a, a)); }  // operators.cuh:77-80
// operators.cuh:82-86 uses rsqrtf; contract: IEEE 1/sqrt
static inline V3 normalized3(V3 a) {
  const float rn = 1.0f / sqrtf(dot3(a, a));
  return v3(a.x * rn, a.y * rn, a.z * rn);
}
static inline V3 mul(const M33& m, V3 a) { return v3(dot3(m.r[0], a), dot3(m.r[1], a), dot3(m.r[2], a)); }  // operators.cuh:88-91

static inline float qnan() { uint32_t u = 0x7fffffffu; float f; memcpy(&f, &u, 4); return f; }

// CUDA __float2int_rn semantics
static inline int f2i_rn(float x) {
  if (std::isnan(x)) return 0;
  if (x >= 2147483648.0f) return INT_MAX;
  if (x <= -2147483648.0f) return INT_MIN;
  return (int)nearbyintf(x);  // FE_TONEAREST: ties to even
}
// CUDA float -> int (truncation, cvt.rzi.s32.f32): NaN -> 0, saturating
static inline int f2i_rz(float x) {
  if (std::isnan(x)) return 0;
  if (x >= 2147483648.0f) return INT_MAX;
  if (x <= -2147483648.0f) return INT_MIN;
  return (int)x;
}

// ------------------------------------------------------------------ exact accumulator
// Segment sums (fp32) are split into integer part and 2^-44 fixed-point fraction and
// accumulated in two int64 words: associative, hence schedule independent.
struct Exact {
  int64_t hi, lo;
  int bad;
  Exact() : hi(0), lo(0), bad(0) {}
  inline void addf(float p) {
    if (!(std::fabs(p) < 4.611686018427387904e18f)) { bad = 1; return; }  // NaN/Inf/|p|>=2^62
    const float hf = truncf(p);
    hi += (int64_t)hf;
    lo += (int64_t)((p - hf) * 17592186044416.0f);  // 2^44, exact scaling, trunc toward 0
  }
  inline void merge(const Exact& o) { hi += o.hi; lo += o.lo; bad |= o.bad; }
  inline float value() const {
    if (bad) return qnan();
    const double v = (double)hi + (double)lo * 5.6843418860808015e-14;  // 2^-44
    return (float)v;
  }
};

static const int SEGW = 32;     // pixels per segment: 32 consecutive pixels of one image row (one GPU warp)

// Generic canonical reduction.  pix(x, y, row[NR]) fills the Jacobian row of one pixel and returns true
// when the pixel contributes.  Within a segment product (i, j), i <= j, is accumulated over the pixels in
// order, pixel 0 first, as s = fma(row[i], row[j], s) in fp32 (the reference's own per-thread accumulation
// `sum.add(getProducts(i))`, reduce.cu:350-356, contracts to the same FFMA under nvcc's default -fmad);
// segment sums are accumulated exactly.  The count of contributing pixels is an integer.
template <int NR, class PixFn>
static void canonical_products(int rows, int cols, PixFn pix, float* out_products, int64_t* out_count) {
  const int NP = NR * (NR + 1) / 2;
  const int spr = (cols + SEGW - 1) / SEGW;
  const long nseg = (long)rows * spr;
  Exact total[NP];
  int64_t count = 0;
#pragma omp parallel
  {
    Exact loc[NP];
    int64_t lcount = 0;
#pragma omp for schedule(static)
    for (long s = 0; s < nseg; ++s) {
      const int y = (int)(s / spr);
      const int x0 = (int)(s - (long)y * spr) * SEGW;
      float segacc[NP];
      for (int k = 0; k < NP; ++k) segacc[k] = 0.0f;
      for (int l = 0; l < SEGW; ++l) {
        const int x = x0 + l;
        float row[NR];
        if (x < cols && pix(x, y, row)) {
          ++lcount;
          int k = 0;
          for (int i = 0; i < NR; ++i)
            for (int j = i; j < NR; ++j, ++k) segacc[k] = fmaf(row[i], row[j], segacc[k]);
        }
      }
      for (int k = 0; k < NP; ++k) loc[k].addf(segacc[k]);
    }
#pragma omp critical
    {
      for (int k = 0; k < NP; ++k) total[k].merge(loc[k]);
      count += lcount;
    }
  }
  for (int k = 0; k < NP; ++k) out_products[k] = total[k].value();
  *out_count = count;
}

// ------------------------------------------------------------------ pyramid kernels (K1..K12)

// cf.cu:643-655 bgr2IntensityKernel.  value = x*0.114f + y*0.299f + z*0.587f, truncated.
static void bgr_to_intensity(const uint8_t* rgba, int rows, int cols, uint8_t* dst) {
#pragma omp parallel for schedule(static)
  for (int y = 0; y < rows; ++y)
    for (int x = 0; x < cols; ++x) {
      const uint8_t* s = rgba + ((size_t)y * cols + x) * 4;
      const float v = fmaf((float)s[2], 0.587f, fmaf((float)s[0], 0.114f, (float)s[1] * 0.299f));
      dst[(size_t)y * cols + x] = (uint8_t)f2i_rz(v);
    }
}

static const float GAUSS5[25] = {1, 4, 6, 4, 1, 4, 16, 24, 16, 4, 6, 24, 36, 24, 6, 4, 16, 24, 16, 4, 1, 4, 6, 4, 1};  // cf.cu:529-530,582-583

// cf.cu:544-573 pyrDownKernelIntensityGauss (window/weight indexing quirks kept)
static void pyrdown_u8_gauss(const uint8_t* src, int srows, int scols, uint8_t* dst) {
  const int drows = srows / 2, dcols = scols / 2;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < drows; ++y)
    for (int x = 0; x < dcols; ++x) {
      const int tx = std::min(2 * x - 2 + 5, scols - 1);
      const int ty = std::min(2 * y - 2 + 5, srows - 1);
      float sum = 0.0f;
      int count = 0;
      for (int cy = std::max(0, 2 * y - 2); cy < ty; ++cy)
        for (int cx = std::max(0, 2 * x - 2); cx < tx; ++cx) {
          const uint8_t v = src[(size_t)cy * scols + cx];
          if (v > 0) {
            const float g = GAUSS5[(ty - cy - 1) * 5 + (tx - cx - 1)];
            sum = fmaf((float)(int)v, g, sum);
            count = f2i_rz((float)count + g);
          }
        }
      const float r = sum / (float)count;  // 0/0 -> NaN -> 0 on store
      dst[(size_t)y * dcols + x] = (uint8_t)f2i_rz(r);
    }
}

// cf.cu:416-443 pyrDownKernelGaussF
static void pyrdown_f32_gauss(const float* src, int srows, int scols, float* dst) {
  const int drows = srows / 2, dcols = scols / 2;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < drows; ++y)
    for (int x = 0; x < dcols; ++x) {
      const int tx = std::min(2 * x - 2 + 5, scols - 1);
      const int ty = std::min(2 * y - 2 + 5, srows - 1);
      float sum = 0.0f;
      int count = 0;
      for (int cy = std::max(0, 2 * y - 2); cy < ty; ++cy)
        for (int cx = std::max(0, 2 * x - 2); cx < tx; ++cx) {
          const float v = src[(size_t)cy * scols + cx];
          if (!std::isnan(v)) {
            const float g = GAUSS5[(ty - cy - 1) * 5 + (tx - cx - 1)];
            sum = fmaf(v, g, sum);
            count = f2i_rz((float)count + g);
          }
        }
      dst[(size_t)y * dcols + x] = sum / (float)count;
    }
}

// cf.cu:57-91 pyrDownGaussKernel (u16 depth, sigma_color = 30 at cf.cu:100)
static void pyrdown_u16(const uint16_t* src, int srows, int scols, uint16_t* dst) {
  const int drows = srows / 2, dcols = scols / 2;
  const float weights[3] = {0.375f, 0.25f, 0.0625f};
  const float thr = 3 * 30.0f;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < drows; ++y)
    for (int x = 0; x < dcols; ++x) {
      const int center = src[(size_t)(2 * y) * scols + 2 * x];
      const int x_mi = std::max(0, 2 * x - 2) - 2 * x;
      const int y_mi = std::max(0, 2 * y - 2) - 2 * y;
      const int x_ma = std::min(scols, 2 * x - 2 + 5) - 2 * x;
      const int y_ma = std::min(srows, 2 * y - 2 + 5) - 2 * y;
      float sum = 0.0f, wall = 0.0f;
      for (int yi = y_mi; yi < y_ma; ++yi)
        for (int xi = x_mi; xi < x_ma; ++xi) {
          const int val = src[(size_t)(2 * y + yi) * scols + 2 * x + xi];
          if ((float)std::abs(val - center) < thr) {
            const float wx = weights[std::abs(xi)], wy = weights[std::abs(yi)];
            sum = fmaf((float)val * wx, wy, sum);
            wall = fmaf(wx, wy, wall);
          }
        }
      dst[(size_t)y * dcols + x] = (uint16_t)f2i_rz(sum / wall);
    }
}

// cf.cu:106-128 computeVmapKernel (+ host wrapper cf.cu:130-147 passes 1.f/fx, 1.f/fy).
// Invalid pixels: only the x plane is written (NaN); y/z planes keep their old content.
static void create_vmap(const uint16_t* depth, int rows, int cols, float fx, float fy, float cx, float cy,
                        float cutoff, float* vmap) {
  const float fx_inv = 1.f / fx, fy_inv = 1.f / fy;
  const size_t plane = (size_t)rows * cols;
#pragma omp parallel for schedule(static)
  for (int v = 0; v < rows; ++v)
    for (int u = 0; u < cols; ++u) {
      const size_t i = (size_t)v * cols + u;
      const float z = (float)(int)depth[i] / 1000.f;
      if (z != 0 && z < cutoff) {
        vmap[i] = z * ((float)u - cx) * fx_inv;
        vmap[plane + i] = z * ((float)v - cy) * fy_inv;
        vmap[2 * plane + i] = z;
      } else {
        vmap[i] = qnan();
      }
    }
}

// cf.cu:149-182 computeNmapKernel
static void create_nmap(const float* vmap, int rows, int cols, float* nmap) {
  const size_t plane = (size_t)rows * cols;
#pragma omp parallel for schedule(static)
  for (int v = 0; v < rows; ++v)
    for (int u = 0; u < cols; ++u) {
      const size_t i = (size_t)v * cols + u;
      if (u == cols - 1 || v == rows - 1) { nmap[i] = qnan(); continue; }
      const float x00 = vmap[i], x01 = vmap[i + 1], x10 = vmap[i + cols];
      if (!std::isnan(x00) && !std::isnan(x01) && !std::isnan(x10)) {
        const V3 v00 = v3(x00, vmap[plane + i], vmap[2 * plane + i]);
        const V3 v01 = v3(x01, vmap[plane + i + 1], vmap[2 * plane + i + 1]);
        const V3 v10 = v3(x10, vmap[plane + i + cols], vmap[2 * plane + i + cols]);
        const V3 r = normalized3(cross3(sub(v01, v00), sub(v10, v00)));
        nmap[i] = r.x; nmap[plane + i] = r.y; nmap[2 * plane + i] = r.z;
      } else {
        nmap[i] = qnan();
      }
    }
}

// cf.cu:313-354 copyMapsKernel (RGBA32F -> planar; both invalidated iff vertex.z == 0)
static void copy_maps(const float* v4, const float* n4, int rows, int cols, float* vmap, float* nmap) {
  const size_t plane = (size_t)rows * cols;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < rows; ++y)
    for (int x = 0; x < cols; ++x) {
      const size_t i = (size_t)y * cols + x;
      const float* vs = v4 + i * 4;
      const float* ns = n4 + i * 4;
      const bool ok = !(vs[2] == 0);
      vmap[i] = ok ? vs[0] : qnan(); vmap[plane + i] = ok ? vs[1] : qnan(); vmap[2 * plane + i] = ok ? vs[2] : qnan();
      nmap[i] = ok ? ns[0] : qnan(); nmap[plane + i] = ok ? ns[1] : qnan(); nmap[2 * plane + i] = ok ? ns[2] : qnan();
    }
}

// cf.cu:445-492 resizeMapKernel<normalize>
static void resize_map(const float* in, int srows, int scols, float* out, bool normalize) {
  const int drows = srows / 2, dcols = scols / 2;
  const size_t sp = (size_t)srows * scols, dp = (size_t)drows * dcols;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < drows; ++y)
    for (int x = 0; x < dcols; ++x) {
      const size_t o = (size_t)y * dcols + x;
      const size_t a = (size_t)(2 * y) * scols + 2 * x, b = a + scols;
      const float x00 = in[a], x01 = in[a + 1], x10 = in[b], x11 = in[b + 1];
      if (std::isnan(x00) || std::isnan(x01) || std::isnan(x10) || std::isnan(x11)) { out[o] = qnan(); continue; }
      V3 n;
      n.x = (x00 + x01 + x10 + x11) / 4;
      n.y = (in[sp + a] + in[sp + a + 1] + in[sp + b] + in[sp + b + 1]) / 4;
      n.z = (in[2 * sp + a] + in[2 * sp + a + 1] + in[2 * sp + b] + in[2 * sp + b + 1]) / 4;
      if (normalize) n = normalized3(n);
      out[o] = n.x; out[dp + o] = n.y; out[2 * dp + o] = n.z;
    }
}

// cf.cu:200-246 tranformMapsKernel (in place: RO.cpp:201-204)
static void transform_maps(float* vmap, float* nmap, int rows, int cols, const M33& R, V3 t) {
  const size_t plane = (size_t)rows * cols;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < rows; ++y)
    for (int x = 0; x < cols; ++x) {
      const size_t i = (size_t)y * cols + x;
      float vx = qnan();
      if (!std::isnan(vmap[i])) {
        const V3 d = add(mul(R, v3(vmap[i], vmap[plane + i], vmap[2 * plane + i])), t);
        vmap[plane + i] = d.y; vmap[2 * plane + i] = d.z; vx = d.x;
      }
      vmap[i] = vx;
      float nx = qnan();
      if (!std::isnan(nmap[i])) {
        const V3 d = mul(R, v3(nmap[i], nmap[plane + i], nmap[2 * plane + i]));
        nmap[plane + i] = d.y; nmap[2 * plane + i] = d.z; nx = d.x;
      }
      nmap[i] = nx;
    }
}

// cf.cu:597-608 verticesToDepthKernel
static void vertices_to_depth(const float* v4, int rows, int cols, float cutoff, float* dst) {
#pragma omp parallel for schedule(static)
  for (int y = 0; y < rows; ++y)
    for (int x = 0; x < cols; ++x) {
      const size_t i = (size_t)y * cols + x;
      const float z = v4[i * 4 + 2];
      dst[i] = (z > cutoff || z <= 0) ? qnan() : z;
    }
}

// cf.cu:674-725 applyKernel + coefficient tables cf.cu:703-707 (border quirk SURVEY A.6)
static void derivatives(const uint8_t* src, int rows, int cols, int16_t* dx, int16_t* dy) {
  const float gsx[9] = {0.52201f, 0.00000f, -0.52201f, 0.79451f, -0.00000f, -0.79451f, 0.52201f, 0.00000f, -0.52201f};
  const float gsy[9] = {0.52201f, 0.79451f, 0.52201f, 0.00000f, 0.00000f, 0.00000f, -0.52201f, -0.79451f, -0.52201f};
#pragma omp parallel for schedule(static)
  for (int y = 0; y < rows; ++y)
    for (int x = 0; x < cols; ++x) {
      float dxv = 0, dyv = 0;
      int k = 8;
      for (int j = std::max(y - 1, 0); j <= std::min(y + 1, rows - 1); ++j)
        for (int i = std::max(x - 1, 0); i <= std::min(x + 1, cols - 1); ++i) {
          const float s = (float)src[(size_t)j * cols + i];
          dxv = fmaf(s, gsx[k], dxv);
          dyv = fmaf(s, gsy[k], dyv);
          --k;
        }
      dx[(size_t)y * cols + x] = (int16_t)f2i_rz(dxv);
      dy[(size_t)y * cols + x] = (int16_t)f2i_rz(dyv);
    }
}

// ------------------------------------------------------------------ depth pre-filter (the step before initICP(depth))

// exp(-t), t >= 0, as plain fp32 arithmetic so that CPU and GPU agree bit for bit (GLSL's exp is implementation
// defined; this one is accurate to ~3e-7 relative): 2^(-a), a = t*log2(e) = n + f with n = rint(a), 2^(-f) by its
// degree-6 Taylor polynomial.
static inline float exp_neg(float t) {
  const float a = t * 1.44269504f;
  const float n = nearbyintf(a);     // ties to even, as rintf on the GPU
  if (!(n < 126.0f)) return 0.0f;    // also catches NaN
  const float f = a - n;             // [-0.5, 0.5]
  float p = 1.54035304e-4f;          // Taylor series of 2^(-f) = exp(-f ln 2), degree 6 (truncation 1.2e-7)
  p = fmaf(p, f, -1.33335581e-3f);
  p = fmaf(p, f, 9.61812911e-3f);
  p = fmaf(p, f, -5.55041087e-2f);
  p = fmaf(p, f, 2.40226507e-1f);
  p = fmaf(p, f, -6.93147181e-1f);
  p = fmaf(p, f, 1.0f);
  uint32_t bits = (uint32_t)(127 - (int)n) << 23;  // 2^(-n)
  float s; memcpy(&s, &bits, 4);
  return p * s;
}

// Shaders/depth_bilateral.frag:31-75 (ElasticFusion::filterDepth, ElasticFusion.cpp:759-768): 13x13 bilateral filter
// on raw u16 millimetres; texel (cx, cy) is read for the tap at (cx, cy).
static void depth_bilateral(const uint16_t* src, int rows, int cols, float maxD, uint16_t* dst) {
  const float ss = 0.024691358f, sc = 0.000555556f;
  const unsigned int hi = (unsigned int)(maxD * 1000.0f);
#pragma omp parallel for schedule(static)
  for (int y = 0; y < rows; ++y)
    for (int x = 0; x < cols; ++x) {
      const unsigned int value = src[(size_t)y * cols + x];
      if (value > hi || value < 300U) { dst[(size_t)y * cols + x] = 0; continue; }
      const int tx = std::min(x - 6 + 13, cols), ty = std::min(y - 6 + 13, rows);
      float sum1 = 0, sum2 = 0;
      for (int cy = std::max(y - 6, 0); cy < ty; ++cy)
        for (int cx = std::max(x - 6, 0); cx < tx; ++cx) {
          const unsigned int tmp = src[(size_t)cy * cols + cx];
          const float dx = (float)x - (float)cx, dy = (float)y - (float)cy;
          const float space2 = fmaf(dx, dx, dy * dy);
          const float dc = (float)value - (float)tmp;
          const float color2 = dc * dc;
          const float weight = exp_neg(fmaf(space2, ss, color2 * sc));
          sum1 = fmaf((float)tmp, weight, sum1);
          sum2 = sum2 + weight;
        }
      dst[(size_t)y * cols + x] = (uint16_t)(unsigned int)roundf(sum1 / sum2);
    }
}

// Shaders/depth_metric.frag:29-39 (ElasticFusion::metriciseDepth, ElasticFusion.cpp:748-757)
static void depth_metric(const uint16_t* src, int rows, int cols, float maxD, float* dst) {
  const unsigned int hi = (unsigned int)(maxD * 1000.0f);
#pragma omp parallel for schedule(static)
  for (int i = 0; i < rows * cols; ++i) {
    const unsigned int value = src[i];
    dst[i] = (value > hi || value < 300U) ? 0.0f : (float)value / 1000.0f;
  }
}

// ------------------------------------------------------------------ GN per-pixel terms

struct Intr { float fx, fy, cx, cy; };
// types.cuh:110-114 CameraModel::operator()(level)
static inline Intr intr_level(Intr k, int level) {
  const int div = 1 << level;
  Intr r = {k.fx / div, k.fy / div, k.cx / div, k.cy / div};
  return r;
}

struct IcpArgs {
  M33 Rcurr; V3 tcurr; M33 Rprev_inv; V3 tprev; Intr intr;
  const float *vmap_curr, *nmap_curr, *vmap_g_prev, *nmap_g_prev;
  float distThres, angleThres; int rows, cols;
};

// reduce.cu:259-344 ICPReduction::search + getProducts (row only)
static inline bool icp_pixel(const IcpArgs& a, int x, int y, float row[7]) {
  const size_t plane = (size_t)a.rows * a.cols, i = (size_t)y * a.cols + x;
  const V3 vcurr = v3(a.vmap_curr[i], a.vmap_curr[plane + i], a.vmap_curr[2 * plane + i]);
  const V3 vcurr_g = add(mul(a.Rcurr, vcurr), a.tcurr);
  const V3 vcurr_cp = mul(a.Rprev_inv, sub(vcurr_g, a.tprev));
  const int ux = f2i_rn(vcurr_cp.x * a.intr.fx / vcurr_cp.z + a.intr.cx);
  const int uy = f2i_rn(vcurr_cp.y * a.intr.fy / vcurr_cp.z + a.intr.cy);
  if (ux < 0 || uy < 0 || ux >= a.cols || uy >= a.rows || vcurr_cp.z < 0) return false;
  const size_t j = (size_t)uy * a.cols + ux;
  const V3 vprev_g = v3(a.vmap_g_prev[j], a.vmap_g_prev[plane + j], a.vmap_g_prev[2 * plane + j]);
  const V3 ncurr = v3(a.nmap_curr[i], a.nmap_curr[plane + i], a.nmap_curr[2 * plane + i]);
  const V3 ncurr_g = mul(a.Rcurr, ncurr);
  const V3 nprev_g = v3(a.nmap_g_prev[j], a.nmap_g_prev[plane + j], a.nmap_g_prev[2 * plane + j]);
  const float dist = norm3(sub(vprev_g, vcurr_g));
  const float sine = norm3(cross3(ncurr_g, nprev_g));
  if (!(sine < a.angleThres && dist <= a.distThres && !std::isnan(ncurr.x) && !std::isnan(nprev_g.x))) return false;
  const V3 s_cp = mul(a.Rprev_inv, sub(vcurr_g, a.tprev));
  const V3 d_cp = mul(a.Rprev_inv, sub(vprev_g, a.tprev));
  const V3 n_cp = mul(a.Rprev_inv, nprev_g);
  const V3 c = cross3(s_cp, n_cp);
  row[0] = n_cp.x; row[1] = n_cp.y; row[2] = n_cp.z;
  row[3] = c.x; row[4] = c.y; row[5] = c.z;
  row[6] = dot3(n_cp, sub(s_cp, d_cp));
  return true;
}

struct RgbCorr { int16_t u0, v0; int16_t diff; uint8_t valid; uint8_t pad; };  // compact DataTerm (types.cuh:76-82)

struct RgbResArgs {
  float minScale; const int16_t *dIdx, *dIdy; const float *lastDepth, *nextDepth;
  const uint8_t *lastImage, *nextImage; float maxDepthDelta; V3 kt; M33 krkinv; int rows, cols;
};

// reduce.cu:767-843 RGBResidual::getProducts
static inline bool rgb_residual_pixel(const RgbResArgs& a, int j0, int i, RgbCorr* out) {
  out->u0 = 0; out->v0 = 0; out->diff = 0; out->valid = 0; out->pad = 0;
  if (!(j0 < a.cols - 5 && i < a.rows - 1)) return false;
  for (int u = std::max(i - 2, 0); u < std::min(i + 2, a.rows); u++)
    for (int v = std::max(j0 - 2, 0); v < std::min(j0 + 2, a.cols); v++)
      if (!(a.nextImage[(size_t)u * a.cols + v] > 0)) return false;
  const size_t k = (size_t)i * a.cols + j0;
  const int valx = a.dIdx[k], valy = a.dIdy[k];
  const float mTwo = (float)((valx * valx) + (valy * valy));
  if (!(mTwo >= a.minScale)) return false;
  const float d1 = a.nextDepth[k];
  if (std::isnan(d1)) return false;
  const float xf = (float)j0, yf = (float)i;
  // k.x*x + k.y*y + k.z  -> fma(kx,x, ky*y) + kz ;  d1*(..) + kt -> fma(d1, s, kt)
  const float s0 = fmaf(a.krkinv.r[0].x, xf, a.krkinv.r[0].y * yf) + a.krkinv.r[0].z;
  const float s1 = fmaf(a.krkinv.r[1].x, xf, a.krkinv.r[1].y * yf) + a.krkinv.r[1].z;
  const float s2 = fmaf(a.krkinv.r[2].x, xf, a.krkinv.r[2].y * yf) + a.krkinv.r[2].z;
  const float transformed_d1 = fmaf(d1, s2, a.kt.z);
  const int u0 = f2i_rn(fmaf(d1, s0, a.kt.x) / transformed_d1);
  const int v0 = f2i_rn(fmaf(d1, s1, a.kt.y) / transformed_d1);
  if (!(u0 >= 0 && v0 >= 0 && u0 < a.cols && v0 < a.rows)) return false;
  const size_t k0 = (size_t)v0 * a.cols + u0;
  const float d0 = a.lastDepth[k0];
  if (!(d0 > 0 && std::fabs(transformed_d1 - d0) <= a.maxDepthDelta && a.lastImage[k0] != 0)) return false;
  out->u0 = (int16_t)u0; out->v0 = (int16_t)v0;
  out->diff = (int16_t)((int)a.nextImage[k] - (int)a.lastImage[k0]);
  out->valid = 1;
  return true;
}

struct RgbStepArgs {
  const RgbCorr* corr; float sigma; const float* lastDepth; Intr intr; const int16_t *dIdx, *dIdy;
  float sobelScale; int rows, cols;
};

// reduce.cu:561-620 RGBReduction::getProducts (row only); point cloud recomputed from
// lastDepth as cf.cu:727-741 projectPointsKernel does (host passes 1.0f/fx: cf.cu:752-754)
static inline bool rgb_step_pixel(const RgbStepArgs& a, int x, int y, float row[7]) {
  const size_t k = (size_t)y * a.cols + x;
  const RgbCorr c = a.corr[k];
  if (!c.valid) return false;
  const float diff = (float)c.diff;
  float w = a.sigma + std::fabs(diff);
  w = w > 1.19209290E-07F ? 1.0f / w : 1.0f;
  if (a.sigma == -1) w = 1;
  row[6] = -w * diff;
  const float z = a.lastDepth[(size_t)c.v0 * a.cols + c.u0];
  const float invFx = 1.0f / a.intr.fx, invFy = 1.0f / a.intr.fy;
  const V3 cp = v3(((float)c.u0 - a.intr.cx) * z * invFx, ((float)c.v0 - a.intr.cy) * z * invFy, z);
  // reduce.cu:584 divides in double and rounds to float; a single IEEE fp32 division gives the same float
  // except on double-rounding ties (probability ~2e-9 per pixel) and is what the kernels use
  const float invz = 1.0f / cp.z;
  const float dI_dx_val = w * a.sobelScale * (float)a.dIdx[k];
  const float dI_dy_val = w * a.sobelScale * (float)a.dIdy[k];
  const float v0 = dI_dx_val * a.intr.fx * invz;
  const float v1 = dI_dy_val * a.intr.fy * invz;
  const float v2 = -(fmaf(v0, cp.x, v1 * cp.y)) * invz;
  row[0] = v0; row[1] = v1; row[2] = v2;
  row[3] = fmaf(-cp.z, v1, cp.y * v2);
  row[4] = fmaf(cp.z, v0, -(cp.x * v2));
  row[5] = fmaf(-cp.y, v0, cp.x * v1);
  return true;
}

struct So3Args { const uint8_t *lastImage, *nextImage; M33 imageBasis, kinv, krlr; int rows, cols; };

// reduce.cu:942-957 SO3Reduction::getGradient
static inline void so3_gradient(const uint8_t* img, int cols, int x, int y, float* gx, float* gy) {
  const float actu = (float)img[(size_t)y * cols + x];
  float back = (float)img[(size_t)y * cols + x - 1], fore = (float)img[(size_t)y * cols + x + 1];
  *gx = ((back + actu) / 2.0f) - ((fore + actu) / 2.0f);
  back = (float)img[(size_t)(y - 1) * cols + x]; fore = (float)img[(size_t)(y + 1) * cols + x];
  *gy = ((back + actu) / 2.0f) - ((fore + actu) / 2.0f);
}

// reduce.cu:959-1032 SO3Reduction::getProducts (row only)
static inline bool so3_pixel(const So3Args& a, int x, int y, float row[4]) {
  const V3 p = v3((float)x, (float)y, 1.0f);
  const V3 wp = mul(a.imageBasis, p);
  const int wx = f2i_rn(wp.x / wp.z), wy = f2i_rn(wp.y / wp.z);
  if (!(wx >= 1 && wx < a.cols - 1 && wy >= 1 && wy < a.rows - 1 && x >= 1 && x < a.cols - 1 && y >= 1 && y < a.rows - 1))
    return false;
  float gnx, gny, glx, gly;
  so3_gradient(a.nextImage, a.cols, wx, wy, &gnx, &gny);
  so3_gradient(a.lastImage, a.cols, x, y, &glx, &gly);
  const float gx = (gnx + glx) / 2.0f, gy = (gny + gly) / 2.0f;
  const V3 point = mul(a.kinv, p);
  const float z2 = point.z * point.z;
  const float A = a.krlr.r[0].x, B = a.krlr.r[0].y, C = a.krlr.r[0].z;
  const float D = a.krlr.r[1].x, E = a.krlr.r[1].y, F = a.krlr.r[1].z;
  const float G = a.krlr.r[2].x, H = a.krlr.r[2].y, I = a.krlr.r[2].z;
  // ((pz*(d*gy + a*gx)) - (gy*g*y) - (gx*g*x)) / z2  with nvcc contraction
  const float yf = p.y, xf = p.x;
  V3 lp;
  lp.x = fmaf(-(gx * G), xf, fmaf(point.z, fmaf(D, gy, A * gx), -((gy * G) * yf))) / z2;
  lp.y = fmaf(-(gx * H), xf, fmaf(point.z, fmaf(E, gy, B * gx), -((gy * H) * yf))) / z2;
  lp.z = fmaf(-(gx * I), xf, fmaf(point.z, fmaf(F, gy, C * gx), -((gy * I) * yf))) / z2;
  const V3 jr = cross3(lp, point);
  row[0] = jr.x; row[1] = jr.y; row[2] = jr.z;
  row[3] = -((float)a.nextImage[(size_t)wy * a.cols + wx] - (float)a.lastImage[(size_t)y * a.cols + x]);
  return true;
}

// ------------------------------------------------------------------ step reductions

// icpStep (reduce.cu:367-428): out[28] = 27 products + residual (row6*row6), inliers
static void icp_step(const IcpArgs& a, float out[28], int64_t* inliers) {
  canonical_products<7>(a.rows, a.cols, [&](int x, int y, float* row) { return icp_pixel(a, x, y, row); }, out, inliers);
}

// computeRgbResidual (reduce.cu:865-925): writes corr[], returns {count, sum diff^2}
static void rgb_residual(const RgbResArgs& a, RgbCorr* corr, int64_t* count, int64_t* sigma) {
  int64_t c = 0, s = 0;
#pragma omp parallel for schedule(static) reduction(+ : c, s)
  for (int y = 0; y < a.rows; ++y)
    for (int x = 0; x < a.cols; ++x) {
      RgbCorr r;
      if (rgb_residual_pixel(a, x, y, &r)) { c += 1; s += (int64_t)r.diff * r.diff; }
      corr[(size_t)y * a.cols + x] = r;
    }
  *count = c; *sigma = s;
}

static void rgb_step(const RgbStepArgs& a, float out[28], int64_t* count) {
  canonical_products<7>(a.rows, a.cols, [&](int x, int y, float* row) { return rgb_step_pixel(a, x, y, row); }, out, count);
}

static void so3_step(const So3Args& a, float out[10], int64_t* count) {
  canonical_products<4>(a.rows, a.cols, [&](int x, int y, float* row) { return so3_pixel(a, x, y, row); }, out, count);
}

// ------------------------------------------------------------------ small dense algebra (double)
// reduce.cu:415-424 unpack: 27 products -> symmetric 6x6 (row-major) + 6-vector
static void unpack_se3(const float p[28], float A[36], float b[6]) {
  int k = 0;
  for (int i = 0; i < 6; ++i)
    for (int j = i; j < 7; ++j) {
      const float v = p[k++];
      if (j == 6) b[i] = v; else A[j * 6 + i] = A[i * 6 + j] = v;
    }
}
static void unpack_so3(const float p[10], float A[9], float b[3]) {  // reduce.cu:1090-1099
  int k = 0;
  for (int i = 0; i < 3; ++i)
    for (int j = i; j < 4; ++j) {
      const float v = p[k++];
      if (j == 3) b[i] = v; else A[j * 3 + i] = A[i * 3 + j] = v;
    }
}

// Reciprocal used by the fp64 solver steps: IEEE fp32 division as the seed, two Newton steps in fp64 with
// explicit fma.  Accurate to ~1 ulp of fp64 and, unlike a hardware reciprocal, identical on CPU and GPU
// (a correctly rounded fp64 division costs ~400 cycles on the one GPU warp every iteration waits for).
static inline double rcp_nr(double d) {
  double r = (double)(1.0f / (float)d);
  r = fma(r, fma(-d, r, 1.0), r);
  r = fma(r, fma(-d, r, 1.0), r);
  return r;
}

// sin / cos for Rodrigues.  Below 0.25 rad (every realistic Gauss-Newton update) an explicit Taylor
// polynomial in plain fp64 mul/add: truncation < 2e-25, and bit-identical on CPU and GPU.  Larger angles
// fall back to the platform's sincos (libm / libdevice agree to ~1 ulp).
static inline void sincos_small(double t, double* sn, double* cs) {
  if (t < 0.25) {
    const double t2 = t * t;
    double s = -7.6471637318198164759e-13;                       // -1/15!
    s = s * t2 + 1.6059043836821614599e-10;                      //  1/13!
    s = s * t2 + -2.5052108385441718775e-08;                     // -1/11!
    s = s * t2 + 2.7557319223985890653e-06;                      //  1/9!
    s = s * t2 + -1.9841269841269841270e-04;                     // -1/7!
    s = s * t2 + 8.3333333333333333333e-03;                      //  1/5!
    s = s * t2 + -1.6666666666666666667e-01;                     // -1/3!
    s = s * t2 + 1.0;
    *sn = t * s;
    double c = 4.7794773323873852974e-14;                        //  1/16!
    c = c * t2 + -1.1470745597729724714e-11;                     // -1/14!
    c = c * t2 + 2.0876756987868098979e-09;                      //  1/12!
    c = c * t2 + -2.7557319223985890653e-07;                     // -1/10!
    c = c * t2 + 2.4801587301587301587e-05;                      //  1/8!
    c = c * t2 + -1.3888888888888888889e-03;                     // -1/6!
    c = c * t2 + 4.1666666666666666667e-02;                      //  1/4!
    c = c * t2 + -0.5;
    c = c * t2 + 1.0;
    *cs = c;
  } else {
    *sn = sin(t); *cs = cos(t);
  }
}

// Reciprocal of an LDL^T pivot.  Eigen's LDLT::solve (RO.cpp:554-566,371) sets the components whose |D| is not above
// its tolerance to zero instead of dividing, so a vanishing pivot (no correspondences: A = 0, b = 0) gives a zero step
// and the prior pose, not NaN.  The threshold is the smallest normal fp32 (the fp32 seed of rcp_nr must be finite).
static inline double pivot_rcp(double d) { const double r = rcp_nr(d); return std::fabs(d) > 1.17549435e-38 ? r : 0.0; }
static inline float pivot_rcp(float d) { return std::fabs(d) > 1.17549435e-38f ? 1.0f / d : 0.0f; }

// LDL^T solve without pivoting (stands in for Eigen ldlt().solve, RO.cpp:371) of the 3x3 fp32 SO(3) system.
// One reciprocal per pivot, then multiplications (the device evaluates the same sequence in every lane).
template <class T, int N>
static void ldlt_solve(const T* A, const T* b, T* x) {
  T L[N * N], D[N], iD[N], y[N];
  for (int i = 0; i < N * N; ++i) L[i] = 0;
  for (int j = 0; j < N; ++j) {
    T d = A[j * N + j];
    for (int k = 0; k < j; ++k) d = d - (L[j * N + k] * L[j * N + k]) * D[k];
    D[j] = d;
    iD[j] = pivot_rcp(d);
    for (int i = j + 1; i < N; ++i) {
      T s = A[i * N + j];
      for (int k = 0; k < j; ++k) s = s - (L[i * N + k] * L[j * N + k]) * D[k];
      L[i * N + j] = s * iD[j];
    }
  }
  for (int i = 0; i < N; ++i) {
    T s = b[i];
    for (int k = 0; k < i; ++k) s = s - L[i * N + k] * y[k];
    y[i] = s;
  }
  for (int i = 0; i < N; ++i) y[i] = y[i] * iD[i];
  for (int i = N - 1; i >= 0; --i) {
    T s = y[i];
    for (int k = N - 1; k > i; --k) s = s - L[k * N + i] * x[k];  // last unknown first (right-looking order)
    x[i] = s;
  }
}

// LDL^T solve of the 6x6 fp64 system (stands in for Eigen ldlt().solve, RO.cpp:554-566) in the form the device's
// solver warp uses: right-looking on the lower triangle of the augmented matrix [A; b^T].  At pivot k every entry to
// the right of column k takes s_rc -= (s_rk * s_ck) / D_k with the unscaled column k, then column k becomes
// L_rk = s_rk / D_k.  Row N carries the right-hand side through the same updates (= forward substitution) and ends as
// y = D^-1 L^-1 b; back substitution takes the last unknown first.
template <int N>
static void ldlt_solve_aug(const double* A, const double* b, double* x) {
  double s[N + 1][N];
  for (int r = 0; r < N; ++r) for (int c = 0; c <= r; ++c) s[r][c] = A[r * N + c];
  for (int c = 0; c < N; ++c) s[N][c] = b[c];
  for (int k = 0; k < N; ++k) {
    const double iDk = pivot_rcp(s[k][k]);
    for (int r = k + 1; r <= N; ++r)
      for (int c = k + 1; c <= (r < N ? r : N - 1); ++c) s[r][c] = s[r][c] - (s[r][k] * s[c][k]) * iDk;
    for (int r = k + 1; r <= N; ++r) s[r][k] = s[r][k] * iDk;
  }
  for (int i = N - 1; i >= 0; --i) {
    double v = s[N][i];
    for (int k = N - 1; k > i; --k) v = v - s[k][i] * x[k];
    x[i] = v;
  }
}

static void mat3_mul(const double* a, const double* b, double* c) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) c[i * 3 + j] = a[i * 3 + 0] * b[0 * 3 + j] + a[i * 3 + 1] * b[1 * 3 + j] + a[i * 3 + 2] * b[2 * 3 + j];
}
// general 3x3 inverse by cofactors (Eigen .inverse() for fixed size 3)
template <class T>
static void mat3_inv(const T* m, T* o) {
  const T c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
  const T det = m[0] * c00 + m[1] * c01 + m[2] * c02;
  const T id = T(1) / det;
  o[0] = c00 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
  o[3] = c01 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
  o[6] = c02 * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

// same cofactor inverse with the Newton reciprocal (device-side use: Rt^-1 of RO.cpp:427)
static void mat3_inv_nr(const double* m, double* o) {
  const double c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
  const double det = m[0] * c00 + m[1] * c01 + m[2] * c02;
  const double id = rcp_nr(det);
  o[0] = c00 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
  o[3] = c01 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
  o[6] = c02 * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

// OP.h:35-71 rodrigues
static void rodrigues(const double r[3], double R[9]) {
  double rx = r[0], ry = r[1], rz = r[2];
  const double theta = sqrt(rx * rx + ry * ry + rz * rz);
  const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  if (theta >= 2.2204460492503131e-16) {
    double s, c;
    sincos_small(theta, &s, &c);
    const double c1 = 1. - c;
    const double itheta = theta ? rcp_nr(theta) : 0.;
    rx *= itheta; ry *= itheta; rz *= itheta;
    const double rrt[9] = {rx * rx, rx * ry, rx * rz, rx * ry, ry * ry, ry * rz, rx * rz, ry * rz, rz * rz};
    const double rx_[9] = {0, -rz, ry, rz, 0, -rx, -ry, rx, 0};
    for (int k = 0; k < 9; ++k) R[k] = c * I[k] + c1 * rrt[k] + s * rx_[k];
  } else {
    for (int k = 0; k < 9; ++k) R[k] = I[k];
  }
}

static M33 m33_from(const double* m) {
  M33 r;
  r.r[0] = v3((float)m[0], (float)m[1], (float)m[2]);
  r.r[1] = v3((float)m[3], (float)m[4], (float)m[5]);
  r.r[2] = v3((float)m[6], (float)m[7], (float)m[8]);
  return r;
}
static M33 m33_fromf(const float* m) {
  M33 r;
  r.r[0] = v3(m[0], m[1], m[2]); r.r[1] = v3(m[3], m[4], m[5]); r.r[2] = v3(m[6], m[7], m[8]);
  return r;
}

// ------------------------------------------------------------------ the tracker object
struct IterLog {  // one record per GN iteration, for per-iteration parity checks
  int level, iter;
  int64_t icp_count, rgb_count, rgb_sigma;
  float icp[28], rgb[28];
  float Rcurr[9], tcurr[3];
};

struct Odom {
  int W, H; Intr K; float distThres, angleThres;
  static const int NP = 3;
  int rows[NP], cols[NP];
  std::vector<uint16_t> depth_tmp[NP];
  std::vector<float> vmap_curr[NP], nmap_curr[NP], vmap_prev[NP], nmap_prev[NP];
  std::vector<float> lastDepth[NP], nextDepth[NP];
  std::vector<uint8_t> lastImage[NP], nextImage[NP], lastNextImage[NP];
  std::vector<int16_t> dIdx[NP], dIdy[NP];
  std::vector<RgbCorr> corr[NP];
  std::vector<float> vmaps_tmp, nmaps_tmp;
  // public state (RO.h:70-77)
  float lastICPError, lastICPCount, lastRGBError, lastRGBCount, lastSO3Error, lastSO3Count;
  double lastA[36], lastb[6];
  int itersDone[NP], so3Iters;
  std::vector<IterLog> log;

  // RO.cpp:21-111 (buffers are zero-filled here; the reference leaves them uninitialised)
  Odom(int w, int h, float cx, float cy, float fx, float fy, float dT, float aT) : W(w), H(h), distThres(dT), angleThres(aT) {
    K.fx = fx; K.fy = fy; K.cx = cx; K.cy = cy;
    for (int i = 0; i < NP; ++i) {
      rows[i] = h >> i; cols[i] = w >> i;
      const size_t n = (size_t)rows[i] * cols[i];
      depth_tmp[i].assign(n, 0);
      vmap_curr[i].assign(3 * n, 0.f); nmap_curr[i].assign(3 * n, 0.f);
      vmap_prev[i].assign(3 * n, 0.f); nmap_prev[i].assign(3 * n, 0.f);
      lastDepth[i].assign(n, 0.f); nextDepth[i].assign(n, 0.f);
      lastImage[i].assign(n, 0); nextImage[i].assign(n, 0); lastNextImage[i].assign(n, 0);
      dIdx[i].assign(n, 0); dIdy[i].assign(n, 0);
      corr[i].resize(n);
    }
    vmaps_tmp.assign((size_t)w * h * 4, 0.f); nmaps_tmp.assign((size_t)w * h * 4, 0.f);
    lastICPError = 0; lastICPCount = (float)(w * h); lastRGBError = 0; lastRGBCount = (float)(w * h);
    lastSO3Error = 0; lastSO3Count = (float)(w * h);
    memset(lastA, 0, sizeof lastA); memset(lastb, 0, sizeof lastb);
    itersDone[0] = itersDone[1] = itersDone[2] = 0; so3Iters = 0;
  }

  // RO.cpp:118-142 initICP(depth)
  void initICPDepth(const uint16_t* depth, float cutoff) {
    memcpy(depth_tmp[0].data(), depth, (size_t)W * H * 2);
    for (int i = 1; i < NP; ++i) pyrdown_u16(depth_tmp[i - 1].data(), rows[i - 1], cols[i - 1], depth_tmp[i].data());
    for (int i = 0; i < NP; ++i) {
      const Intr k = intr_level(K, i);
      create_vmap(depth_tmp[i].data(), rows[i], cols[i], k.fx, k.fy, k.cx, k.cy, cutoff, vmap_curr[i].data());
      create_nmap(vmap_curr[i].data(), rows[i], cols[i], nmap_curr[i].data());
    }
  }
  // RO.cpp:144-167 initICP(V, N)
  void initICPMaps(const float* v4, const float* n4) {
    memcpy(vmaps_tmp.data(), v4, vmaps_tmp.size() * 4); memcpy(nmaps_tmp.data(), n4, nmaps_tmp.size() * 4);
    copy_maps(v4, n4, rows[0], cols[0], vmap_curr[0].data(), nmap_curr[0].data());
    for (int i = 1; i < NP; ++i) {
      resize_map(vmap_curr[i - 1].data(), rows[i - 1], cols[i - 1], vmap_curr[i].data(), false);
      resize_map(nmap_curr[i - 1].data(), rows[i - 1], cols[i - 1], nmap_curr[i].data(), true);
    }
  }
  // RO.cpp:169-207 initICPModel
  void initICPModel(const float* v4, const float* n4, const float pose[16]) {
    memcpy(vmaps_tmp.data(), v4, vmaps_tmp.size() * 4); memcpy(nmaps_tmp.data(), n4, nmaps_tmp.size() * 4);
    copy_maps(v4, n4, rows[0], cols[0], vmap_prev[0].data(), nmap_prev[0].data());
    for (int i = 1; i < NP; ++i) {
      resize_map(vmap_prev[i - 1].data(), rows[i - 1], cols[i - 1], vmap_prev[i].data(), false);
      resize_map(nmap_prev[i - 1].data(), rows[i - 1], cols[i - 1], nmap_prev[i].data(), true);
    }
    M33 R; R.r[0] = v3(pose[0], pose[1], pose[2]); R.r[1] = v3(pose[4], pose[5], pose[6]); R.r[2] = v3(pose[8], pose[9], pose[10]);
    const V3 t = v3(pose[3], pose[7], pose[11]);
    for (int i = 0; i < NP; ++i) transform_maps(vmap_prev[i].data(), nmap_prev[i].data(), rows[i], cols[i], R, t);
  }
  // RO.cpp:209-236 populateRGBDData (maxDepthRGB = 6.0f, RO.cpp:37)
  void populate(const uint8_t* rgba, std::vector<float>* dD, std::vector<uint8_t>* dI) {
    vertices_to_depth(vmaps_tmp.data(), rows[0], cols[0], 6.0f, dD[0].data());
    for (int i = 0; i + 1 < NP; ++i) pyrdown_f32_gauss(dD[i].data(), rows[i], cols[i], dD[i + 1].data());
    bgr_to_intensity(rgba, rows[0], cols[0], dI[0].data());
    for (int i = 0; i + 1 < NP; ++i) pyrdown_u8_gauss(dI[i].data(), rows[i], cols[i], dI[i + 1].data());
  }
  void initRGBModel(const uint8_t* rgba) { populate(rgba, lastDepth, lastImage); }  // RO.cpp:238-242
  void initRGB(const uint8_t* rgba) { populate(rgba, nextDepth, nextImage); }       // RO.cpp:244-248
  void initFirstRGB(const uint8_t* rgba) {                                           // RO.cpp:250-266
    bgr_to_intensity(rgba, rows[0], cols[0], lastNextImage[0].data());
    for (int i = 0; i + 1 < NP; ++i) pyrdown_u8_gauss(lastNextImage[i].data(), rows[i], cols[i], lastNextImage[i + 1].data());
  }

  // RO.cpp:268-605 getIncrementalTransformation
  void track(float trans[3], float rot[9], bool rgbOnly, float icpWeight, bool pyramid, bool fastOdom, bool so3,
             bool interMap, bool keepLog) {
    const bool icp = !rgbOnly && icpWeight > 0;
    const bool rgb = rgbOnly || icpWeight < 100;
    float Rprev[9], tprev[3], Rcurr[9], tcurr[3];
    memcpy(Rprev, rot, sizeof Rprev); memcpy(tprev, trans, sizeof tprev);
    memcpy(Rcurr, rot, sizeof Rcurr); memcpy(tcurr, trans, sizeof tcurr);
    log.clear();
    if (rgb)
      for (int i = 0; i < NP; ++i) derivatives(nextImage[i].data(), rows[i], cols[i], dIdx[i].data(), dIdy[i].data());

    double resultR[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    so3Iters = 0;
    if (so3) {  // RO.cpp:297-385
      const int L = 2;
      const Intr k = intr_level(K, L);
      float R_lr[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
      const double Kd[9] = {k.fx, 0, k.cx, 0, k.fy, k.cy, 0, 0, 1};
      double Kinv[9]; mat3_inv(Kd, Kinv);
      float lastError = std::numeric_limits<float>::max() / 2, lastCount = std::numeric_limits<float>::max() / 2;
      double lastResultR[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
      for (int it = 0; it < 10; ++it) {
        double KR[9], Hm[9];
        mat3_mul(Kd, resultR, KR); mat3_mul(KR, Kinv, Hm);
        So3Args a; a.lastImage = lastNextImage[L].data(); a.nextImage = nextImage[L].data();
        a.imageBasis = m33_from(Hm); a.kinv = m33_from(Kinv); a.krlr = m33_from(KR); a.rows = rows[L]; a.cols = cols[L];
        float prod[10]; int64_t cnt;
        so3_step(a, prod, &cnt);
        ++so3Iters;
        float jtj[9], jtr[3]; unpack_so3(prod, jtj, jtr);
        const float residual0 = prod[9], residual1 = (float)cnt;
        lastSO3Error = sqrtf(residual0) / residual1;
        lastSO3Count = residual1;
        if (lastSO3Error < lastError && lastCount == lastSO3Count) break;
        else if ((double)lastSO3Error > (double)lastError + 0.001) {
          lastSO3Error = lastError; lastSO3Count = lastCount; memcpy(resultR, lastResultR, sizeof resultR); break;
        }
        lastError = lastSO3Error; lastCount = lastSO3Count; memcpy(lastResultR, resultR, sizeof resultR);
        float delta[3]; ldlt_solve<float, 3>(jtj, jtr, delta);
        const double dd[3] = {delta[0], delta[1], delta[2]};
        double ru[9]; rodrigues(dd, ru);
        float ruf[9], nr[9];
        for (int q = 0; q < 9; ++q) ruf[q] = (float)ru[q];
        for (int r = 0; r < 3; ++r)
          for (int c = 0; c < 3; ++c) nr[r * 3 + c] = ruf[r * 3 + 0] * R_lr[0 * 3 + c] + ruf[r * 3 + 1] * R_lr[1 * 3 + c] + ruf[r * 3 + 2] * R_lr[2 * 3 + c];
        memcpy(R_lr, nr, sizeof nr);
        for (int q = 0; q < 9; ++q) resultR[q] = R_lr[q];
      }
    }

    int iterations[NP];
    iterations[0] = interMap ? 50 : fastOdom ? 3 : 10;  // RO.cpp:387-389
    iterations[1] = interMap ? 50 : pyramid ? 5 : 0;
    iterations[2] = interMap ? 50 : pyramid ? 4 : 0;

    float Rprev_inv[9]; mat3_inv<float>(Rprev, Rprev_inv);  // RO.cpp:391
    double resultRt[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    if (so3) for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) resultRt[r * 4 + c] = resultR[r * 3 + c];

    float icpErr = 0.f, icpCnt = 0.f;  // documented deviation: defined (0) when icp is off
    bool anyIter = false;
    for (int i = NP - 1; i >= 0; --i) {
      const Intr k = intr_level(K, i);
      const double Kd[9] = {k.fx, 0, k.cx, 0, k.fy, k.cy, 0, 0, 1};
      double Kinv[9]; mat3_inv(Kd, Kinv);
      lastRGBError = std::numeric_limits<float>::max();
      itersDone[i] = 0;
      for (int j = 0; j < iterations[i]; ++j) {
        // Rt = resultRt.inverse(): affine [R t; 0 1] -> [R^-1, -R^-1 t]
        double R3[9], Ri[9], ti[3];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) R3[r * 3 + c] = resultRt[r * 4 + c];
        mat3_inv_nr(R3, Ri);
        for (int r = 0; r < 3; ++r) ti[r] = -(Ri[r * 3 + 0] * resultRt[3] + Ri[r * 3 + 1] * resultRt[7] + Ri[r * 3 + 2] * resultRt[11]);
        double KR[9], KRKi[9]; mat3_mul(Kd, Ri, KR); mat3_mul(KR, Kinv, KRKi);
        const M33 krkInv = m33_from(KRKi);
        double Kt[3];
        for (int r = 0; r < 3; ++r) Kt[r] = Kd[r * 3 + 0] * ti[0] + Kd[r * 3 + 1] * ti[1] + Kd[r * 3 + 2] * ti[2];
        const V3 kt = v3((float)Kt[0], (float)Kt[1], (float)Kt[2]);

        int sigma = 0, rgbSize = 0;
        int64_t rc = 0, rs = 0;
        if (rgb) {
          RgbResArgs a;
          const double g = (double)(i == 0 ? 5 : i == 1 ? 3 : 1);           // RO.cpp:107-110
          a.minScale = (float)((g * g) / (0.125 * 0.125));                  // RO.cpp:445 pow(g,2)/pow(sobelScale,2)
          a.dIdx = dIdx[i].data(); a.dIdy = dIdy[i].data(); a.lastDepth = lastDepth[i].data(); a.nextDepth = nextDepth[i].data();
          a.lastImage = lastImage[i].data(); a.nextImage = nextImage[i].data(); a.maxDepthDelta = 0.07f;
          a.kt = kt; a.krkinv = krkInv; a.rows = rows[i]; a.cols = cols[i];
          rgb_residual(a, corr[i].data(), &rc, &rs);
          rgbSize = (int)rc; sigma = (int)(int32_t)(uint32_t)(uint64_t)rs;  // reference sums in 32-bit int
        }
        float sigmaVal = (float)std::sqrt((double)(((float)sigma / rgbSize == 0) ? 1 : rgbSize));  // RO.cpp:464 (sic)
        if (rgbOnly && std::sqrt((double)sigma) / rgbSize > (double)lastRGBError) break;          // RO.cpp:466-469
        lastRGBError = (float)(std::sqrt((double)sigma) / rgbSize);
        lastRGBCount = (float)rgbSize;
        if (rgbOnly) sigmaVal = -1;

        IterLog lg; memset(&lg, 0, sizeof lg);
        lg.level = i; lg.iter = j; lg.rgb_count = rc; lg.rgb_sigma = rs;
        float A_icp[36] = {0}, b_icp[6] = {0}, A_rgb[36] = {0}, b_rgb[6] = {0};
        if (icp) {
          IcpArgs a; a.Rcurr = m33_fromf(Rcurr); a.tcurr = v3(tcurr[0], tcurr[1], tcurr[2]);
          a.Rprev_inv = m33_fromf(Rprev_inv); a.tprev = v3(tprev[0], tprev[1], tprev[2]); a.intr = k;
          a.vmap_curr = vmap_curr[i].data(); a.nmap_curr = nmap_curr[i].data();
          a.vmap_g_prev = vmap_prev[i].data(); a.nmap_g_prev = nmap_prev[i].data();
          a.distThres = distThres; a.angleThres = angleThres; a.rows = rows[i]; a.cols = cols[i];
          int64_t inl; icp_step(a, lg.icp, &inl); lg.icp_count = inl;
          unpack_se3(lg.icp, A_icp, b_icp);
          const float residual0 = lg.icp[27], residual1 = (float)inl;
          icpErr = sqrtf(residual0) / residual1; icpCnt = residual1;  // RO.cpp:517-518
        }
        lastICPError = icpErr; lastICPCount = icpCnt;
        if (rgb) {
          RgbStepArgs a; a.corr = corr[i].data(); a.sigma = sigmaVal; a.lastDepth = lastDepth[i].data(); a.intr = k;
          a.dIdx = dIdx[i].data(); a.dIdy = dIdy[i].data(); a.sobelScale = 0.125f; a.rows = rows[i]; a.cols = cols[i];
          int64_t c2; rgb_step(a, lg.rgb, &c2);
          unpack_se3(lg.rgb, A_rgb, b_rgb);
        }
        // RO.cpp:543-567
        double result[6];
        if (icp && rgb) {
          const double w = icpWeight;
          for (int q = 0; q < 36; ++q) lastA[q] = (double)A_rgb[q] + w * w * (double)A_icp[q];
          for (int q = 0; q < 6; ++q) lastb[q] = (double)b_rgb[q] + w * (double)b_icp[q];
        } else if (icp) {
          for (int q = 0; q < 36; ++q) lastA[q] = A_icp[q];
          for (int q = 0; q < 6; ++q) lastb[q] = b_icp[q];
        } else {
          for (int q = 0; q < 36; ++q) lastA[q] = A_rgb[q];
          for (int q = 0; q < 6; ++q) lastb[q] = b_rgb[q];
        }
        ldlt_solve_aug<6>(lastA, lastb, result);
        // OP.h:73-93 computeUpdateSE3
        double Ru[9]; rodrigues(result + 3, Ru);
        double U[16] = {Ru[0], Ru[1], Ru[2], result[0], Ru[3], Ru[4], Ru[5], result[1], Ru[6], Ru[7], Ru[8], result[2], 0, 0, 0, 1};
        double nRt[16];
        for (int r = 0; r < 4; ++r)
          for (int c = 0; c < 4; ++c)
            nRt[r * 4 + c] = U[r * 4 + 0] * resultRt[0 * 4 + c] + U[r * 4 + 1] * resultRt[1 * 4 + c] + U[r * 4 + 2] * resultRt[2 * 4 + c] + U[r * 4 + 3] * resultRt[3 * 4 + c];
        memcpy(resultRt, nRt, sizeof nRt);
        float oR[9], ot[3];
        for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) oR[r * 3 + c] = (float)resultRt[r * 4 + c]; ot[r] = (float)resultRt[r * 4 + 3]; }
        // RO.cpp:577-585: currentT = [Rprev|tprev] * rgbOdom.inverse(), Isometry inverse = [R^T | -R^T t]
        float iR[9], it_[3];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) iR[r * 3 + c] = oR[c * 3 + r];
        for (int r = 0; r < 3; ++r) it_[r] = -(iR[r * 3 + 0] * ot[0] + iR[r * 3 + 1] * ot[1] + iR[r * 3 + 2] * ot[2]);
        for (int r = 0; r < 3; ++r) {
          for (int c = 0; c < 3; ++c) Rcurr[r * 3 + c] = Rprev[r * 3 + 0] * iR[0 * 3 + c] + Rprev[r * 3 + 1] * iR[1 * 3 + c] + Rprev[r * 3 + 2] * iR[2 * 3 + c];
          tcurr[r] = (Rprev[r * 3 + 0] * it_[0] + Rprev[r * 3 + 1] * it_[1] + Rprev[r * 3 + 2] * it_[2]) + tprev[r];
        }
        memcpy(lg.Rcurr, Rcurr, sizeof Rcurr); memcpy(lg.tcurr, tcurr, sizeof tcurr);
        if (keepLog) log.push_back(lg);
        ++itersDone[i];
        anyIter = true;
      }
    }
    (void)anyIter;
    // RO.cpp:589-593
    if (rgb) {
      const float dx = tcurr[0] - tprev[0], dy = tcurr[1] - tprev[1], dz = tcurr[2] - tprev[2];
      const float n = sqrtf(dx * dx + dy * dy + dz * dz);
      if ((double)n > 0.3) { memcpy(Rcurr, Rprev, sizeof Rcurr); memcpy(tcurr, tprev, sizeof tcurr); }
    }
    if (so3) for (int i = 0; i < NP; ++i) std::swap(lastNextImage[i], nextImage[i]);  // RO.cpp:595-601
    memcpy(trans, tcurr, sizeof tcurr); memcpy(rot, Rcurr, sizeof Rcurr);
  }
};

}  // namespace

// ======================================================================================
// C API for ctypes (dense row-major host arrays; planar maps are 3*rows*cols floats)
// ======================================================================================
extern "C" {

int orc_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void orc_set_num_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}

void orc_bgr_to_intensity(const uint8_t* rgba, int rows, int cols, uint8_t* dst) { bgr_to_intensity(rgba, rows, cols, dst); }
void orc_pyrdown_u8_gauss(const uint8_t* s, int r, int c, uint8_t* d) { pyrdown_u8_gauss(s, r, c, d); }
void orc_pyrdown_f32_gauss(const float* s, int r, int c, float* d) { pyrdown_f32_gauss(s, r, c, d); }
void orc_pyrdown_u16(const uint16_t* s, int r, int c, uint16_t* d) { pyrdown_u16(s, r, c, d); }
void orc_create_vmap(const uint16_t* depth, int rows, int cols, float fx, float fy, float cx, float cy, float cutoff, float* vmap) {
  create_vmap(depth, rows, cols, fx, fy, cx, cy, cutoff, vmap);
}
void orc_create_nmap(const float* vmap, int rows, int cols, float* nmap) { create_nmap(vmap, rows, cols, nmap); }
void orc_copy_maps(const float* v4, const float* n4, int rows, int cols, float* vmap, float* nmap) { copy_maps(v4, n4, rows, cols, vmap, nmap); }
void orc_resize_map(const float* in, int srows, int scols, float* out, int normalize) { resize_map(in, srows, scols, out, normalize != 0); }
void orc_transform_maps(float* vmap, float* nmap, int rows, int cols, const float R[9], const float t[3]) {
  transform_maps(vmap, nmap, rows, cols, m33_fromf(R), v3(t[0], t[1], t[2]));
}
void orc_vertices_to_depth(const float* v4, int rows, int cols, float cutoff, float* dst) { vertices_to_depth(v4, rows, cols, cutoff, dst); }
void orc_derivatives(const uint8_t* src, int rows, int cols, int16_t* dx, int16_t* dy) { derivatives(src, rows, cols, dx, dy); }
void orc_depth_bilateral(const uint16_t* src, int rows, int cols, float maxD, uint16_t* dst) { depth_bilateral(src, rows, cols, maxD, dst); }
void orc_depth_metric(const uint16_t* src, int rows, int cols, float maxD, float* dst) { depth_metric(src, rows, cols, maxD, dst); }
float orc_exp_neg(float t) { return exp_neg(t); }

// out[28] products (27 JtJ/Jtr + residual), inliers
void orc_icp_step(const float Rcurr[9], const float tcurr[3], const float* vmap_curr, const float* nmap_curr,
                  const float Rprev_inv[9], const float tprev[3], float fx, float fy, float cx, float cy,
                  const float* vmap_g_prev, const float* nmap_g_prev, float distThres, float angleThres, int rows, int cols,
                  float out[28], int64_t* inliers) {
  IcpArgs a; a.Rcurr = m33_fromf(Rcurr); a.tcurr = v3(tcurr[0], tcurr[1], tcurr[2]); a.Rprev_inv = m33_fromf(Rprev_inv);
  a.tprev = v3(tprev[0], tprev[1], tprev[2]); a.intr.fx = fx; a.intr.fy = fy; a.intr.cx = cx; a.intr.cy = cy;
  a.vmap_curr = vmap_curr; a.nmap_curr = nmap_curr; a.vmap_g_prev = vmap_g_prev; a.nmap_g_prev = nmap_g_prev;
  a.distThres = distThres; a.angleThres = angleThres; a.rows = rows; a.cols = cols;
  icp_step(a, out, inliers);
}

// per-pixel ICP inlier mask (valid-correspondence mask parity)
void orc_icp_mask(const float Rcurr[9], const float tcurr[3], const float* vmap_curr, const float* nmap_curr,
                  const float Rprev_inv[9], const float tprev[3], float fx, float fy, float cx, float cy,
                  const float* vmap_g_prev, const float* nmap_g_prev, float distThres, float angleThres, int rows, int cols,
                  uint8_t* mask) {
  IcpArgs a; a.Rcurr = m33_fromf(Rcurr); a.tcurr = v3(tcurr[0], tcurr[1], tcurr[2]); a.Rprev_inv = m33_fromf(Rprev_inv);
  a.tprev = v3(tprev[0], tprev[1], tprev[2]); a.intr.fx = fx; a.intr.fy = fy; a.intr.cx = cx; a.intr.cy = cy;
  a.vmap_curr = vmap_curr; a.nmap_curr = nmap_curr; a.vmap_g_prev = vmap_g_prev; a.nmap_g_prev = nmap_g_prev;
  a.distThres = distThres; a.angleThres = angleThres; a.rows = rows; a.cols = cols;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < rows; ++y)
    for (int x = 0; x < cols; ++x) { float row[7]; mask[(size_t)y * cols + x] = icp_pixel(a, x, y, row) ? 1 : 0; }
}

// corr: rows*cols records of 8 bytes {i16 u0, i16 v0, i16 diff, u8 valid, u8 pad}
void orc_rgb_residual(float minScale, const int16_t* dIdx, const int16_t* dIdy, const float* lastDepth, const float* nextDepth,
                      const uint8_t* lastImage, const uint8_t* nextImage, float maxDepthDelta, const float kt[3],
                      const float krkinv[9], int rows, int cols, void* corr, int64_t* count, int64_t* sigma) {
  RgbResArgs a; a.minScale = minScale; a.dIdx = dIdx; a.dIdy = dIdy; a.lastDepth = lastDepth; a.nextDepth = nextDepth;
  a.lastImage = lastImage; a.nextImage = nextImage; a.maxDepthDelta = maxDepthDelta; a.kt = v3(kt[0], kt[1], kt[2]);
  a.krkinv = m33_fromf(krkinv); a.rows = rows; a.cols = cols;
  rgb_residual(a, (RgbCorr*)corr, count, sigma);
}

void orc_rgb_step(const void* corr, float sigma, const float* lastDepth, float fx, float fy, float cx, float cy,
                  const int16_t* dIdx, const int16_t* dIdy, float sobelScale, int rows, int cols, float out[28], int64_t* count) {
  RgbStepArgs a; a.corr = (const RgbCorr*)corr; a.sigma = sigma; a.lastDepth = lastDepth; a.intr.fx = fx; a.intr.fy = fy;
  a.intr.cx = cx; a.intr.cy = cy; a.dIdx = dIdx; a.dIdy = dIdy; a.sobelScale = sobelScale; a.rows = rows; a.cols = cols;
  rgb_step(a, out, count);
}

void orc_so3_step(const uint8_t* lastImage, const uint8_t* nextImage, const float imageBasis[9], const float kinv[9],
                  const float krlr[9], int rows, int cols, float out[10], int64_t* count) {
  So3Args a; a.lastImage = lastImage; a.nextImage = nextImage; a.imageBasis = m33_fromf(imageBasis); a.kinv = m33_fromf(kinv);
  a.krlr = m33_fromf(krlr); a.rows = rows; a.cols = cols;
  so3_step(a, out, count);
}

// ---- stateful tracker
void* orc_odom_create(int w, int h, float cx, float cy, float fx, float fy, float distThresh, float angleThresh) {
  return new Odom(w, h, cx, cy, fx, fy, distThresh, angleThresh);
}
void orc_odom_destroy(void* o) { delete (Odom*)o; }
void orc_odom_init_icp_depth(void* o, const uint16_t* depth, float cutoff) { ((Odom*)o)->initICPDepth(depth, cutoff); }
void orc_odom_init_icp_maps(void* o, const float* v4, const float* n4) { ((Odom*)o)->initICPMaps(v4, n4); }
void orc_odom_init_icp_model(void* o, const float* v4, const float* n4, const float pose[16]) { ((Odom*)o)->initICPModel(v4, n4, pose); }
void orc_odom_init_rgb(void* o, const uint8_t* rgba) { ((Odom*)o)->initRGB(rgba); }
void orc_odom_init_rgb_model(void* o, const uint8_t* rgba) { ((Odom*)o)->initRGBModel(rgba); }
void orc_odom_init_first_rgb(void* o, const uint8_t* rgba) { ((Odom*)o)->initFirstRGB(rgba); }

struct orc_result {
  float lastICPError, lastICPCount, lastRGBError, lastRGBCount, lastSO3Error, lastSO3Count;
  double A[36], b[6];
  int iters[3], so3_iters;
};
void orc_odom_track(void* o_, float trans[3], float rot[9], int rgbOnly, float icpWeight, int pyramid, int fastOdom, int so3,
                    int interMap, int keepLog, orc_result* out) {
  Odom* o = (Odom*)o_;
  o->track(trans, rot, rgbOnly != 0, icpWeight, pyramid != 0, fastOdom != 0, so3 != 0, interMap != 0, keepLog != 0);
  if (out) {
    out->lastICPError = o->lastICPError; out->lastICPCount = o->lastICPCount; out->lastRGBError = o->lastRGBError;
    out->lastRGBCount = o->lastRGBCount; out->lastSO3Error = o->lastSO3Error; out->lastSO3Count = o->lastSO3Count;
    memcpy(out->A, o->lastA, sizeof o->lastA); memcpy(out->b, o->lastb, sizeof o->lastb);
    for (int i = 0; i < 3; ++i) out->iters[i] = o->itersDone[i];
    out->so3_iters = o->so3Iters;
  }
}
int orc_odom_log_size(void* o) { return (int)((Odom*)o)->log.size(); }
// flat record: level, iter, icp_count, rgb_count, rgb_sigma (as int64 x5), icp[28], rgb[28], R[9], t[3]
void orc_odom_log_get(void* o_, int idx, int64_t ints[5], float icp[28], float rgb[28], float R[9], float t[3]) {
  const IterLog& l = ((Odom*)o_)->log[idx];
  ints[0] = l.level; ints[1] = l.iter; ints[2] = l.icp_count; ints[3] = l.rgb_count; ints[4] = l.rgb_sigma;
  memcpy(icp, l.icp, sizeof l.icp); memcpy(rgb, l.rgb, sizeof l.rgb); memcpy(R, l.Rcurr, sizeof l.Rcurr); memcpy(t, l.tcurr, sizeof l.tcurr);
}
// which: 0 vmap_curr 1 nmap_curr 2 vmap_g_prev 3 nmap_g_prev 4 lastDepth 5 nextDepth 6 lastImage 7 nextImage
//        8 lastNextImage 9 dIdx 10 dIdy 11 depth_tmp 12 corr
const void* orc_odom_buffer(void* o_, int which, int level, int* rows, int* cols) {
  Odom* o = (Odom*)o_;
  *rows = o->rows[level]; *cols = o->cols[level];
  switch (which) {
    case 0: return o->vmap_curr[level].data();
    case 1: return o->nmap_curr[level].data();
    case 2: return o->vmap_prev[level].data();
    case 3: return o->nmap_prev[level].data();
    case 4: return o->lastDepth[level].data();
    case 5: return o->nextDepth[level].data();
    case 6: return o->lastImage[level].data();
    case 7: return o->nextImage[level].data();
    case 8: return o->lastNextImage[level].data();
    case 9: return o->dIdx[level].data();
    case 10: return o->dIdy[level].data();
    case 11: return o->depth_tmp[level].data();
    case 12: return o->corr[level].data();
  }
  return 0;
}
// RO.cpp:607-610 getCovariance = lastA.lu().inverse(): Gauss-Jordan with partial pivoting in double
void orc_covariance(const double A[36], double cov[36]) {
  double m[6][12];
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) { m[i][j] = A[i * 6 + j]; m[i][6 + j] = (i == j) ? 1.0 : 0.0; }
  for (int c = 0; c < 6; ++c) {
    int p = c;
    for (int r = c + 1; r < 6; ++r) if (std::fabs(m[r][c]) > std::fabs(m[p][c])) p = r;
    if (p != c) for (int j = 0; j < 12; ++j) std::swap(m[p][j], m[c][j]);
    const double d = m[c][c];
    for (int j = 0; j < 12; ++j) m[c][j] /= d;
    for (int r = 0; r < 6; ++r) if (r != c) { const double f = m[r][c]; for (int j = 0; j < 12; ++j) m[r][j] -= f * m[c][j]; }
  }
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) cov[i * 6 + j] = m[i][6 + j];
}

}  // extern "C"
