// Pyramid construction kernels (reference kernels K1..K12), fused:
//   maps_pyramid     : copyMaps + resizeVMap/resizeNMap x2 + tranformMaps      (cf.cu:313-354,445-492,200-246)
//   v2d_intensity    : verticesToDepth + bgr2Intensity                         (cf.cu:597-608,643-655)
//   pyrdown_gauss    : pyrDownGaussF + pyrDownUcharGauss                       (cf.cu:416-443,544-573)
//   pyrdown_u16      : pyrDown (bilateral-ish u16 depth)                       (cf.cu:57-91)
//   vmap_nmap        : createVMap + createNMap for the three levels            (cf.cu:106-128,149-182)
// All outputs are bit-identical to running the reference's separate passes with IEEE arithmetic.
#include "common.cuh"
#include "kernels.h"

namespace dms {

// ------------------------------------------------------------------ maps_pyramid
struct MapsPyrArgs {
  const float4* v4; const float4* n4; int pitch4;  // RGBA32F inputs, pitch in bytes
  float* vmap[3]; float* nmap[3]; int pitch[3]; int rows[3]; int cols[3];
  float* zdepth; float zcut;  // level-0 depth of the vertex map (verticesToDepth, cf.cu:597-608), kept for init_rgb*
  Mat33 R; float3 t;
};

__device__ __forceinline__ void store_plane3(float* base, int pitch, int rows, int y, int x, float3 v) {
  row_ptr(base, pitch, y)[x] = v.x;
  row_ptr(base, pitch, y + rows)[x] = v.y;
  row_ptr(base, pitch, y + 2 * rows)[x] = v.z;
}

template <bool TRANSFORM>
__device__ __forceinline__ float3 xform_v(const MapsPyrArgs& a, float3 v) {
  if (!TRANSFORM) return isnan(v.x) ? make_float3(qnan(), qnan(), qnan()) : v;
  if (isnan(v.x)) return make_float3(qnan(), qnan(), qnan());
  return add3(mul33(a.R, v), a.t);
}
template <bool TRANSFORM>
__device__ __forceinline__ float3 xform_n(const MapsPyrArgs& a, float3 n) {
  if (!TRANSFORM) return isnan(n.x) ? make_float3(qnan(), qnan(), qnan()) : n;
  if (isnan(n.x)) return make_float3(qnan(), qnan(), qnan());
  return mul33(a.R, n);
}

// box average of 4 taps (cf.cu:459-486); NaN in any x tap -> invalid
__device__ __forceinline__ float3 box4(float3 a, float3 b, float3 c, float3 d, bool normalize) {
  if (isnan(a.x) || isnan(b.x) || isnan(c.x) || isnan(d.x)) return make_float3(qnan(), qnan(), qnan());
  float3 n;
  n.x = fmul(fadd(fadd(fadd(a.x, b.x), c.x), d.x), 0.25f);
  n.y = fmul(fadd(fadd(fadd(a.y, b.y), c.y), d.y), 0.25f);
  n.z = fmul(fadd(fadd(fadd(a.z, b.z), c.z), d.z), 0.25f);
  if (normalize) n = normalized3(n);
  return n;
}

// One thread per 4x4 block of level-0 pixels: writes 16 L0, 4 L1 and 1 L2 outputs.
template <bool TRANSFORM>
__global__ void __launch_bounds__(128) k_maps_pyramid(const MapsPyrArgs a) {
  const int bx = blockIdx.x * blockDim.x + threadIdx.x;
  const int by = blockIdx.y * blockDim.y + threadIdx.y;
  const int x0 = bx * 4, y0 = by * 4;
  if (x0 >= a.cols[0] || y0 >= a.rows[0]) return;
  float3 v1[4], n1[4];  // level-1 values (before transform) of the 2x2 sub-blocks
#pragma unroll
  for (int sy = 0; sy < 2; ++sy) {
#pragma unroll
    for (int sx = 0; sx < 2; ++sx) {
      float3 v0[4], n0[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int x = x0 + sx * 2 + (q & 1), y = y0 + sy * 2 + (q >> 1);
        float3 v = make_float3(qnan(), qnan(), qnan()), n = v;
        if (x < a.cols[0] && y < a.rows[0]) {
          const float4 vs = __ldg(row_ptr(a.v4, a.pitch4, y) + x);
          const float4 ns = __ldg(row_ptr(a.n4, a.pitch4, y) + x);
          if (!(vs.z == 0)) { v = make_float3(vs.x, vs.y, vs.z); n = make_float3(ns.x, ns.y, ns.z); }
          row_ptr(a.zdepth, a.pitch[0], y)[x] = (vs.z > a.zcut || vs.z <= 0) ? qnan() : vs.z;
          store_plane3(a.vmap[0], a.pitch[0], a.rows[0], y, x, xform_v<TRANSFORM>(a, v));
          store_plane3(a.nmap[0], a.pitch[0], a.rows[0], y, x, xform_n<TRANSFORM>(a, n));
        }
        v0[q] = v; n0[q] = n;
      }
      const int x1 = bx * 2 + sx, y1 = by * 2 + sy;
      const float3 vv = box4(v0[0], v0[1], v0[2], v0[3], false);
      const float3 nn = box4(n0[0], n0[1], n0[2], n0[3], true);
      v1[sy * 2 + sx] = vv; n1[sy * 2 + sx] = nn;
      if (x1 < a.cols[1] && y1 < a.rows[1]) {
        store_plane3(a.vmap[1], a.pitch[1], a.rows[1], y1, x1, xform_v<TRANSFORM>(a, vv));
        store_plane3(a.nmap[1], a.pitch[1], a.rows[1], y1, x1, xform_n<TRANSFORM>(a, nn));
      }
    }
  }
  if (bx < a.cols[2] && by < a.rows[2]) {
    const float3 vv = box4(v1[0], v1[1], v1[2], v1[3], false);
    const float3 nn = box4(n1[0], n1[1], n1[2], n1[3], true);
    store_plane3(a.vmap[2], a.pitch[2], a.rows[2], by, bx, xform_v<TRANSFORM>(a, vv));
    store_plane3(a.nmap[2], a.pitch[2], a.rows[2], by, bx, xform_n<TRANSFORM>(a, nn));
  }
}

cudaError_t launch_maps_pyramid(const Level* lv, bool use_prev, const float4* v4, const float4* n4, int pitch4,
                                const float* pose16_or_null, float* zdepth, float zcut, cudaStream_t s) {
  MapsPyrArgs a;
  a.v4 = v4; a.n4 = n4; a.pitch4 = pitch4; a.zdepth = zdepth; a.zcut = zcut;
  for (int i = 0; i < 3; ++i) {
    a.vmap[i] = use_prev ? lv[i].vmap_prev : lv[i].vmap_curr;
    a.nmap[i] = use_prev ? lv[i].nmap_prev : lv[i].nmap_curr;
    a.pitch[i] = lv[i].pitch_f32; a.rows[i] = lv[i].rows; a.cols[i] = lv[i].cols;
  }
  const dim3 block(32, 4);
  const dim3 grid(((lv[0].cols + 3) / 4 + 31) / 32, ((lv[0].rows + 3) / 4 + 3) / 4);
  if (pose16_or_null) {
    const float* p = pose16_or_null;
    a.R.r[0] = make_float3(p[0], p[1], p[2]); a.R.r[1] = make_float3(p[4], p[5], p[6]); a.R.r[2] = make_float3(p[8], p[9], p[10]);
    a.t = make_float3(p[3], p[7], p[11]);
    k_maps_pyramid<true><<<grid, block, 0, s>>>(a);
  } else {
    a.R.r[0] = a.R.r[1] = a.R.r[2] = make_float3(0, 0, 0); a.t = make_float3(0, 0, 0);
    k_maps_pyramid<false><<<grid, block, 0, s>>>(a);
  }
  return cudaGetLastError();
}

// ------------------------------------------------------------------ v2d_intensity
// One thread per 4 consecutive pixels.  depth: the level-0 depth of the last loaded vertex map (already cut at
// maxDepthRGB = 6 m by maps_pyramid; cf.cu:604-607, RO.cpp:37,213) is copied into the destination pyramid;
// intensity: int(x*0.114f + y*0.299f + z*0.587f) = trunc(fma(z,.587, fma(x,.114, y*.299)))  (cf.cu:651-654)
__global__ void __launch_bounds__(256) k_v2d_intensity(const float* zsrc, int pitch_z, float* depth, int pitch_d,
                                                       const uchar4* rgba, int pitch_rgba, unsigned char* img, int pitch_i,
                                                       int rows, int cols) {
  const int qx = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  const int x0 = qx * 4;
  if (x0 >= cols || y >= rows) return;
  const int n = min(4, cols - x0);
  if (depth) {
    float d[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      d[p] = qnan();
      if (p < n) d[p] = __ldg(row_ptr(zsrc, pitch_z, y) + x0 + p);
    }
    float* dr = row_ptr(depth, pitch_d, y) + x0;
    if (n == 4) *reinterpret_cast<float4*>(dr) = make_float4(d[0], d[1], d[2], d[3]);
    else for (int p = 0; p < n; ++p) dr[p] = d[p];
  }
  if (img) {
    unsigned char o[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      o[p] = 0;
      if (p < n) {
        const uchar4 s = __ldg(row_ptr(rgba, pitch_rgba, y) + x0 + p);
        const float v = ffma((float)s.z, 0.587f, ffma((float)s.x, 0.114f, fmul((float)s.y, 0.299f)));
        o[p] = (unsigned char)__float2int_rz(v);
      }
    }
    unsigned char* ir = row_ptr(img, pitch_i, y) + x0;
    if (n == 4) *reinterpret_cast<uchar4*>(ir) = make_uchar4(o[0], o[1], o[2], o[3]);
    else for (int p = 0; p < n; ++p) ir[p] = o[p];
  }
}

cudaError_t launch_v2d_intensity(const float* zsrc, int pitch_z, float* depth, int pitch_d, const uchar4* rgba,
                                 int pitch_rgba, unsigned char* img, int pitch_i, int rows, int cols, cudaStream_t s) {
  const dim3 block(256);
  const dim3 grid(((cols + 3) / 4 + 255) / 256, rows);
  k_v2d_intensity<<<grid, block, 0, s>>>(zsrc, pitch_z, depth, pitch_d, rgba, pitch_rgba, img, pitch_i, rows, cols);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ pyrdown_gauss (f32 depth + u8 image)
__constant__ float c_gauss5[25] = {1, 4, 6, 4, 1, 4, 16, 24, 16, 4, 6, 24, 36, 24, 6, 4, 16, 24, 16, 4, 1, 4, 6, 4, 1};

__global__ void __launch_bounds__(256) k_pyrdown_gauss(const float* sd, int spd, float* dd, int dpd, const unsigned char* si,
                                                       int spi, unsigned char* di, int dpi, int srows, int scols) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  const int drows = srows / 2, dcols = scols / 2;
  if (x >= dcols || y >= drows) return;
  const int tx = min(2 * x - 2 + 5, scols - 1);
  const int ty = min(2 * y - 2 + 5, srows - 1);
  const int cx0 = max(0, 2 * x - 2), cy0 = max(0, 2 * y - 2);
  if (sd) {
    float sum = 0.0f; int count = 0;
    for (int cy = cy0; cy < ty; ++cy) {
      const float* r = row_ptr(sd, spd, cy);
      for (int cx = cx0; cx < tx; ++cx) {
        const float v = __ldg(r + cx);
        if (!isnan(v)) {
          const float g = c_gauss5[(ty - cy - 1) * 5 + (tx - cx - 1)];
          sum = ffma(v, g, sum);
          count = __float2int_rz(fadd((float)count, g));
        }
      }
    }
    row_ptr(dd, dpd, y)[x] = fdiv(sum, (float)count);
  }
  if (si) {
    float sum = 0.0f; int count = 0;
    for (int cy = cy0; cy < ty; ++cy) {
      const unsigned char* r = row_ptr(si, spi, cy);
      for (int cx = cx0; cx < tx; ++cx) {
        const unsigned char v = __ldg(r + cx);
        if (v > 0) {
          const float g = c_gauss5[(ty - cy - 1) * 5 + (tx - cx - 1)];
          sum = ffma((float)(int)v, g, sum);
          count = __float2int_rz(fadd((float)count, g));
        }
      }
    }
    const float r = fdiv(sum, (float)count);  // 0/0 -> NaN -> 0
    row_ptr(di, dpi, y)[x] = (unsigned char)__float2int_rz(r);
  }
}

cudaError_t launch_pyrdown_gauss(const float* sd, int spd, float* dd, int dpd, const unsigned char* si, int spi,
                                 unsigned char* di, int dpi, int srows, int scols, cudaStream_t s) {
  const dim3 block(32, 8);
  const dim3 grid((scols / 2 + 31) / 32, (srows / 2 + 7) / 8);
  if (grid.x == 0 || grid.y == 0) return cudaSuccess;
  k_pyrdown_gauss<<<grid, block, 0, s>>>(sd, spd, dd, dpd, si, spi, di, dpi, srows, scols);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ pyrdown_u16 (cf.cu:57-91, sigma_color 30)
__global__ void __launch_bounds__(256) k_pyrdown_u16(const unsigned short* src, int sp, unsigned short* dst, int dp, int srows,
                                                     int scols) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  const int drows = srows / 2, dcols = scols / 2;
  if (x >= dcols || y >= drows) return;
  const int center = __ldg(row_ptr(src, sp, 2 * y) + 2 * x);
  const int x_mi = max(0, 2 * x - 2) - 2 * x, y_mi = max(0, 2 * y - 2) - 2 * y;
  const int x_ma = min(scols, 2 * x - 2 + 5) - 2 * x, y_ma = min(srows, 2 * y - 2 + 5) - 2 * y;
  float sum = 0.0f, wall = 0.0f;
  for (int yi = y_mi; yi < y_ma; ++yi) {
    const unsigned short* r = row_ptr(src, sp, 2 * y + yi);
    const float wy = (yi == 0) ? 0.375f : (abs(yi) == 1 ? 0.25f : 0.0625f);
    for (int xi = x_mi; xi < x_ma; ++xi) {
      const int val = __ldg(r + 2 * x + xi);
      if ((float)abs(val - center) < 90.0f) {
        const float wx = (xi == 0) ? 0.375f : (abs(xi) == 1 ? 0.25f : 0.0625f);
        sum = ffma(fmul((float)val, wx), wy, sum);
        wall = ffma(wx, wy, wall);
      }
    }
  }
  row_ptr(dst, dp, y)[x] = (unsigned short)__float2int_rz(fdiv(sum, wall));
}

cudaError_t launch_pyrdown_u16(const unsigned short* src, int sp, unsigned short* dst, int dp, int srows, int scols,
                               cudaStream_t s) {
  const dim3 block(32, 8);
  const dim3 grid((scols / 2 + 31) / 32, (srows / 2 + 7) / 8);
  if (grid.x == 0 || grid.y == 0) return cudaSuccess;
  k_pyrdown_u16<<<grid, block, 0, s>>>(src, sp, dst, dp, srows, scols);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ vmap_nmap (all three levels, grid.z = level)
struct VNArgs { const unsigned short* depth[3]; int pd[3]; float* vmap[3]; float* nmap[3]; int pf[3]; int rows[3]; int cols[3]; Intr k[3]; float cutoff; };

__device__ __forceinline__ float3 backproject(const unsigned short* depth, int pd, int u, int v, const Intr& k, float fxi, float fyi,
                                              float cutoff) {
  const float z = fdiv((float)(int)__ldg(row_ptr(depth, pd, v) + u), 1000.f);
  if (z != 0 && z < cutoff) return make_float3(fmul(fmul(z, fsub((float)u, k.cx)), fxi), fmul(fmul(z, fsub((float)v, k.cy)), fyi), z);
  return make_float3(qnan(), qnan(), qnan());
}

__global__ void __launch_bounds__(256) k_vmap_nmap(const VNArgs a) {
  const int l = blockIdx.z;
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  const int v = blockIdx.y * blockDim.y + threadIdx.y;
  const int rows = a.rows[l], cols = a.cols[l];
  if (u >= cols || v >= rows) return;
  const Intr k = a.k[l];
  const float fxi = fdiv(1.f, k.fx), fyi = fdiv(1.f, k.fy);
  const float3 v00 = backproject(a.depth[l], a.pd[l], u, v, k, fxi, fyi, a.cutoff);
  store_plane3(a.vmap[l], a.pf[l], rows, v, u, v00);
  float3 n = make_float3(qnan(), qnan(), qnan());
  if (u < cols - 1 && v < rows - 1) {
    const float3 v01 = backproject(a.depth[l], a.pd[l], u + 1, v, k, fxi, fyi, a.cutoff);
    const float3 v10 = backproject(a.depth[l], a.pd[l], u, v + 1, k, fxi, fyi, a.cutoff);
    if (!isnan(v00.x) && !isnan(v01.x) && !isnan(v10.x)) n = normalized3(cross3(sub3(v01, v00), sub3(v10, v00)));
  }
  store_plane3(a.nmap[l], a.pf[l], rows, v, u, n);
}

cudaError_t launch_vmap_nmap(const Level* lv, float cutoff, cudaStream_t s) {
  VNArgs a;
  for (int i = 0; i < 3; ++i) {
    a.depth[i] = lv[i].depth_tmp; a.pd[i] = lv[i].pitch_u16; a.vmap[i] = lv[i].vmap_curr; a.nmap[i] = lv[i].nmap_curr;
    a.pf[i] = lv[i].pitch_f32; a.rows[i] = lv[i].rows; a.cols[i] = lv[i].cols; a.k[i] = lv[i].intr;
  }
  a.cutoff = cutoff;
  const dim3 block(32, 8);
  const dim3 grid((lv[0].cols + 31) / 32, (lv[0].rows + 7) / 8, 3);
  k_vmap_nmap<<<grid, block, 0, s>>>(a);
  return cudaGetLastError();
}

}  // namespace dms
