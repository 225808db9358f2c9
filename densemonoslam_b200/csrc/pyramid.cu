// Pyramid construction kernels (reference kernels K1..K12), fused:
//   maps_pyramid     : copyMaps + resizeVMap/resizeNMap x2 + tranformMaps      (cf.cu:313-354,445-492,200-246)
//   populate         : verticesToDepth + bgr2Intensity + pyrDownGaussF x2 + pyrDownUcharGauss x2   (cf.cu:416-443,544-573,597-655)
//   icp_depth        : pyrDown x2 (bilateral-ish u16) + createVMap + createNMap for the three levels   (cf.cu:57-182)
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
template <bool LEG>
__device__ __forceinline__ float3 box4(float3 a, float3 b, float3 c, float3 d, bool normalize) {
  if (isnan(a.x) || isnan(b.x) || isnan(c.x) || isnan(d.x)) return make_float3(qnan(), qnan(), qnan());
  float3 n;
  n.x = fmul(fadd(fadd(fadd(a.x, b.x), c.x), d.x), 0.25f);
  n.y = fmul(fadd(fadd(fadd(a.y, b.y), c.y), d.y), 0.25f);
  n.z = fmul(fadd(fadd(fadd(a.z, b.z), c.z), d.z), 0.25f);
  if (normalize) n = normalized3_t<LEG>(n);
  return n;
}

// One thread per 2x2 block of level-0 pixels (= one level-1 pixel): writes 4 L0 and 1 L1 outputs; the level-2 pixel of a
// 4x4 block is assembled from its four threads with shuffles.  A warp covers 16 x 2 level-1 pixels (lane = ly * 16 + lx),
// so rows are read as 512 contiguous bytes and the level-0 planes are written as float2 pairs.
__device__ __forceinline__ void store_plane3_pair(float* base, int pitch, int rows, int y, int x, float3 a, float3 b) {
  *reinterpret_cast<float2*>(row_ptr(base, pitch, y) + x) = make_float2(a.x, b.x);
  *reinterpret_cast<float2*>(row_ptr(base, pitch, y + rows) + x) = make_float2(a.y, b.y);
  *reinterpret_cast<float2*>(row_ptr(base, pitch, y + 2 * rows) + x) = make_float2(a.z, b.z);
}
__device__ __forceinline__ float3 shfl3(float3 v, int src) {
  return make_float3(__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src), __shfl_sync(0xffffffffu, v.z, src));
}
template <bool TRANSFORM, bool LEG>
__global__ void __launch_bounds__(128) k_maps_pyramid(const MapsPyrArgs a) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x1 = blockIdx.x * 32 + (warp & 1) * 16 + (lane & 15);   // level-1 pixel of this thread
  const int y1 = blockIdx.y * 4 + (warp >> 1) * 2 + (lane >> 4);
  const int x0 = x1 * 2, y0 = y1 * 2;
  float3 v0[4], n0[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int x = x0 + (q & 1), y = y0 + (q >> 1);
    float3 v = make_float3(qnan(), qnan(), qnan()), n = v;
    if (x < a.cols[0] && y < a.rows[0]) {
      const float4 vs = __ldg(row_ptr(a.v4, a.pitch4, y) + x);
      const float4 ns = __ldg(row_ptr(a.n4, a.pitch4, y) + x);
      if (!(vs.z == 0)) { v = make_float3(vs.x, vs.y, vs.z); n = make_float3(ns.x, ns.y, ns.z); }
      row_ptr(a.zdepth, a.pitch[0], y)[x] = (vs.z > a.zcut || vs.z <= 0) ? qnan() : vs.z;
    }
    v0[q] = v; n0[q] = n;
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int y = y0 + r;
    if (y >= a.rows[0] || x0 >= a.cols[0]) continue;
    const float3 va = xform_v<TRANSFORM>(a, v0[2 * r]), na = xform_n<TRANSFORM>(a, n0[2 * r]);
    if (x0 + 1 < a.cols[0]) {
      store_plane3_pair(a.vmap[0], a.pitch[0], a.rows[0], y, x0, va, xform_v<TRANSFORM>(a, v0[2 * r + 1]));
      store_plane3_pair(a.nmap[0], a.pitch[0], a.rows[0], y, x0, na, xform_n<TRANSFORM>(a, n0[2 * r + 1]));
    } else {
      store_plane3(a.vmap[0], a.pitch[0], a.rows[0], y, x0, va);
      store_plane3(a.nmap[0], a.pitch[0], a.rows[0], y, x0, na);
    }
  }
  const float3 vv = box4<LEG>(v0[0], v0[1], v0[2], v0[3], false);
  const float3 nn = box4<LEG>(n0[0], n0[1], n0[2], n0[3], true);
  if (x1 < a.cols[1] && y1 < a.rows[1]) {
    store_plane3(a.vmap[1], a.pitch[1], a.rows[1], y1, x1, xform_v<TRANSFORM>(a, vv));
    store_plane3(a.nmap[1], a.pitch[1], a.rows[1], y1, x1, xform_n<TRANSFORM>(a, nn));
  }
  // level 2: the 2x2 level-1 values sit in lanes l, l+1, l+16, l+17 (l = even lx of the upper row)
  const float3 vb = shfl3(vv, lane + 1), vc = shfl3(vv, lane + 16), vd = shfl3(vv, lane + 17);
  const float3 nb = shfl3(nn, lane + 1), nc = shfl3(nn, lane + 16), nd = shfl3(nn, lane + 17);
  const int x2 = x1 >> 1, y2 = y1 >> 1;
  if ((lane & 17) == 0 && x2 < a.cols[2] && y2 < a.rows[2]) {
    store_plane3(a.vmap[2], a.pitch[2], a.rows[2], y2, x2, xform_v<TRANSFORM>(a, box4<LEG>(vv, vb, vc, vd, false)));
    store_plane3(a.nmap[2], a.pitch[2], a.rows[2], y2, x2, xform_n<TRANSFORM>(a, box4<LEG>(nn, nb, nc, nd, true)));
  }
}

cudaError_t launch_maps_pyramid(const Level* lv, bool use_prev, const float4* v4, const float4* n4, int pitch4,
                                const float* pose16_or_null, float* zdepth, float zcut, bool legacy, cudaStream_t s) {
  MapsPyrArgs a;
  a.v4 = v4; a.n4 = n4; a.pitch4 = pitch4; a.zdepth = zdepth; a.zcut = zcut;
  for (int i = 0; i < 3; ++i) {
    a.vmap[i] = use_prev ? lv[i].vmap_prev : lv[i].vmap_curr;
    a.nmap[i] = use_prev ? lv[i].nmap_prev : lv[i].nmap_curr;
    a.pitch[i] = lv[i].pitch_f32; a.rows[i] = lv[i].rows; a.cols[i] = lv[i].cols;
  }
  const dim3 block(128);
  const dim3 grid(((lv[0].cols + 1) / 2 + 31) / 32, ((lv[0].rows + 1) / 2 + 3) / 4);
  if (pose16_or_null) {
    const float* p = pose16_or_null;
    a.R.r[0] = make_float3(p[0], p[1], p[2]); a.R.r[1] = make_float3(p[4], p[5], p[6]); a.R.r[2] = make_float3(p[8], p[9], p[10]);
    a.t = make_float3(p[3], p[7], p[11]);
    if (legacy) k_maps_pyramid<true, true><<<grid, block, 0, s>>>(a); else k_maps_pyramid<true, false><<<grid, block, 0, s>>>(a);
  } else {
    a.R.r[0] = a.R.r[1] = a.R.r[2] = make_float3(0, 0, 0); a.t = make_float3(0, 0, 0);
    if (legacy) k_maps_pyramid<false, true><<<grid, block, 0, s>>>(a); else k_maps_pyramid<false, false><<<grid, block, 0, s>>>(a);
  }
  return cudaGetLastError();
}

// ------------------------------------------------------------------ populate: depth + intensity pyramids in ONE launch
// verticesToDepth copy + bgr2Intensity (cf.cu:597-608,643-655) and both levels of pyrDownGaussF / pyrDownUcharGauss
// (cf.cu:416-443,544-573).  A CTA owns a 32x32 block of level 0 (= 16x16 of level 1 = 8x8 of level 2) and stages the
// 41x41 / 19x19 tiles it needs (with halos) in shared memory, so every tap is a shared-memory read and the colour
// image is read once.  Window bounds and weight indexing are the reference's, evaluated on global
// coordinates, so results are bit-identical to the three separate passes.
__constant__ float c_gauss5[25] = {1, 4, 6, 4, 1, 4, 16, 24, 16, 4, 6, 24, 36, 24, 6, 4, 16, 24, 16, 4, 1, 4, 6, 4, 1};

struct PopArgs {
  const float* zsrc; int pitch_z;            // level-0 depth of the last loaded vertex map (already cut at 6 m)
  const uchar4* rgba; int pitch_rgba;
  float* depth[3]; unsigned char* img[3];    // destination pyramids (depth[] all null: image only)
  int pitch_f[3], pitch_i[3], rows[3], cols[3];
};

// One output of pyrDownGaussF / pyrDownUcharGauss (cf.cu:416-443, 544-573) from shared tiles.  The window is
// [cx0, tx) x [cy0, ty) with the weight of tap (cx, cy) taken at (tx - cx - 1, ty - cy - 1).  The counts
// are sums of the integer weights (the reference adds them in float and truncates: exact), so they are kept as ints.
// FULL = the 5x5 window lies inside the image: fixed trip counts, weights as immediates.
template <bool HAS_DEPTH, bool FULL, int TS_D, int TS_I>
__device__ __forceinline__ void gauss_tap_window(const float* sd, const unsigned char* si, int lx0, int ly0, int nx, int ny, float* sumd_o, int* cntd_o,
                                                 float* sumi_o, int* cnti_o) {
  float sumd = 0.f, sumi = 0.f; int cntd = 0, cnti = 0;
  if (FULL) {
#pragma unroll
    for (int r = 0; r < 5; ++r)
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        const int gi = ((r == 0 || r == 4) ? 1 : ((r == 1 || r == 3) ? 4 : 6)) * ((c == 0 || c == 4) ? 1 : ((c == 1 || c == 3) ? 4 : 6));  // symmetric: index order irrelevant
        const float g = (float)gi;
        const unsigned char iv = si[(ly0 + r) * TS_I + lx0 + c];
        if (iv > 0) { sumi = ffma((float)(int)iv, g, sumi); cnti += gi; }
        if (HAS_DEPTH) {
          const float v = sd[(ly0 + r) * TS_D + lx0 + c];
          if (!isnan(v)) { sumd = ffma(v, g, sumd); cntd += gi; }
        }
      }
  } else {
    for (int r = 0; r < ny; ++r)
      for (int c = 0; c < nx; ++c) {
        const float g = c_gauss5[(ny - r - 1) * 5 + (nx - c - 1)];
        const unsigned char iv = si[(ly0 + r) * TS_I + lx0 + c];
        if (iv > 0) { sumi = ffma((float)(int)iv, g, sumi); cnti = __float2int_rz(fadd((float)cnti, g)); }
        if (HAS_DEPTH) {
          const float v = sd[(ly0 + r) * TS_D + lx0 + c];
          if (!isnan(v)) { sumd = ffma(v, g, sumd); cntd = __float2int_rz(fadd((float)cntd, g)); }
        }
      }
  }
  *sumd_o = sumd; *cntd_o = cntd; *sumi_o = sumi; *cnti_o = cnti;
}

template <bool HAS_DEPTH, bool LEG>
__global__ void __launch_bounds__(256) k_populate(const PopArgs a) {
  __shared__ float sd0[41][41];
  __shared__ unsigned char si0[41][44];
  __shared__ float sd1[19][19];
  __shared__ unsigned char si1[19][20];
  const int tid = threadIdx.x, bx = blockIdx.x, by = blockIdx.y;
  const int ox0 = 32 * bx - 6, oy0 = 32 * by - 6, ox1 = 16 * bx - 2, oy1 = 16 * by - 2;
  // ---- level 0 tile (+ own 32x32 block written out)
  for (int idx = tid; idx < 41 * 41; idx += 256) {
    const int ly = idx / 41, lx = idx - ly * 41, gx = ox0 + lx, gy = oy0 + ly;
    if (gx < 0 || gy < 0 || gx >= a.cols[0] || gy >= a.rows[0]) continue;
    const uchar4 c = __ldg(row_ptr(a.rgba, a.pitch_rgba, gy) + gx);
    const unsigned char iv = (unsigned char)__float2int_rz(ffma((float)c.z, 0.587f, ffma((float)c.x, 0.114f, fmul((float)c.y, 0.299f))));
    si0[ly][lx] = iv;
    float d = 0.f;
    if (HAS_DEPTH) { d = __ldg(row_ptr(a.zsrc, a.pitch_z, gy) + gx); sd0[ly][lx] = d; }
    if (lx >= 6 && lx < 38 && ly >= 6 && ly < 38) {
      row_ptr(a.img[0], a.pitch_i[0], gy)[gx] = iv;
      if (HAS_DEPTH) row_ptr(a.depth[0], a.pitch_f[0], gy)[gx] = d;
    }
  }
  __syncthreads();
  // ---- level 1 tile
  for (int idx = tid; idx < 19 * 19; idx += 256) {
    const int ly = idx / 19, lx = idx - ly * 19, x = ox1 + lx, y = oy1 + ly;
    if (x < 0 || y < 0 || x >= a.cols[1] || y >= a.rows[1]) continue;
    const int tx = min(2 * x - 2 + 5, a.cols[0] - 1), ty = min(2 * y - 2 + 5, a.rows[0] - 1);
    const int cx0 = max(0, 2 * x - 2), cy0 = max(0, 2 * y - 2);
    float sumd, sumi; int cntd, cnti;
    if (tx - cx0 == 5 && ty - cy0 == 5) gauss_tap_window<HAS_DEPTH, true, 41, 44>(&sd0[0][0], &si0[0][0], cx0 - ox0, cy0 - oy0, 5, 5, &sumd, &cntd, &sumi, &cnti);
    else gauss_tap_window<HAS_DEPTH, false, 41, 44>(&sd0[0][0], &si0[0][0], cx0 - ox0, cy0 - oy0, tx - cx0, ty - cy0, &sumd, &cntd, &sumi, &cnti);
    const unsigned char iv1 = (unsigned char)__float2int_rz(div_t<LEG>(sumi, (float)cnti));  // 0/0 -> NaN -> 0
    si1[ly][lx] = iv1;
    float d1 = 0.f;
    if (HAS_DEPTH) { d1 = div_t<LEG>(sumd, (float)cntd); sd1[ly][lx] = d1; }
    if (lx >= 2 && lx < 18 && ly >= 2 && ly < 18) {
      row_ptr(a.img[1], a.pitch_i[1], y)[x] = iv1;
      if (HAS_DEPTH) row_ptr(a.depth[1], a.pitch_f[1], y)[x] = d1;
    }
  }
  __syncthreads();
  // ---- level 2 block
  if (tid < 64) {
    const int ly = tid >> 3, lx = tid & 7, x = 8 * bx + lx, y = 8 * by + ly;
    if (x < a.cols[2] && y < a.rows[2]) {
      const int tx = min(2 * x - 2 + 5, a.cols[1] - 1), ty = min(2 * y - 2 + 5, a.rows[1] - 1);
      const int cx0 = max(0, 2 * x - 2), cy0 = max(0, 2 * y - 2);
      float sumd, sumi; int cntd, cnti;
      if (tx - cx0 == 5 && ty - cy0 == 5) gauss_tap_window<HAS_DEPTH, true, 19, 20>(&sd1[0][0], &si1[0][0], cx0 - ox1, cy0 - oy1, 5, 5, &sumd, &cntd, &sumi, &cnti);
      else gauss_tap_window<HAS_DEPTH, false, 19, 20>(&sd1[0][0], &si1[0][0], cx0 - ox1, cy0 - oy1, tx - cx0, ty - cy0, &sumd, &cntd, &sumi, &cnti);
      row_ptr(a.img[2], a.pitch_i[2], y)[x] = (unsigned char)__float2int_rz(div_t<LEG>(sumi, (float)cnti));
      if (HAS_DEPTH) row_ptr(a.depth[2], a.pitch_f[2], y)[x] = div_t<LEG>(sumd, (float)cntd);
    }
  }
}

cudaError_t launch_populate(const Level* lv, int which, const float* zsrc, const uchar4* rgba, int pitch_rgba, bool legacy, cudaStream_t s) {
  // which: 0 = last (initRGBModel), 1 = next (initRGB), 2 = lastNext image only (initFirstRGB)
  PopArgs a;
  a.zsrc = zsrc; a.pitch_z = lv[0].pitch_f32; a.rgba = rgba; a.pitch_rgba = pitch_rgba;
  for (int i = 0; i < 3; ++i) {
    a.depth[i] = which == 0 ? lv[i].lastDepth : which == 1 ? lv[i].nextDepth : nullptr;
    a.img[i] = which == 0 ? lv[i].lastImage : which == 1 ? lv[i].nextImage : lv[i].lastNextImage;
    a.pitch_f[i] = lv[i].pitch_f32; a.pitch_i[i] = lv[i].pitch_u8; a.rows[i] = lv[i].rows; a.cols[i] = lv[i].cols;
  }
  const dim3 grid((lv[0].cols + 31) / 32, (lv[0].rows + 31) / 32);
  if (which == 2) { if (legacy) k_populate<false, true><<<grid, 256, 0, s>>>(a); else k_populate<false, false><<<grid, 256, 0, s>>>(a); }
  else { if (legacy) k_populate<true, true><<<grid, 256, 0, s>>>(a); else k_populate<true, false><<<grid, 256, 0, s>>>(a); }
  return cudaGetLastError();
}

// ------------------------------------------------------------------ icp_depth: depth pyramid + vertex / normal maps in ONE launch
// pyrDown x2 (cf.cu:57-91, sigma_color 30), createVMap and createNMap for the three levels (cf.cu:106-182).
// Same tiling: 45x45 / 21x21 / 9x9 shared-memory tiles of u16 depth (one extra row / column for the forward
// differences of the normals).
struct IcpDepthArgs {
  const unsigned short* depth; int pitch_in;
  unsigned short* dt[3]; float* vmap[3]; float* nmap[3];
  int pitch_u16[3], pitch_f[3], rows[3], cols[3];
  Intr k[3]; float fxi[3], fyi[3];  // 1.f / fx, 1.f / fy as the reference's host code passes them (cf.cu:136-137)
  float cutoff;
};

template <bool LEG>
__device__ __forceinline__ unsigned short pyrdown_u16_tap(const unsigned short* tile, int tstride, int ox, int oy, int x, int y, int srows, int scols) {
  const int center = tile[(2 * y - oy) * tstride + (2 * x - ox)];
  const int x_mi = max(0, 2 * x - 2) - 2 * x, y_mi = max(0, 2 * y - 2) - 2 * y;
  const int x_ma = min(scols, 2 * x - 2 + 5) - 2 * x, y_ma = min(srows, 2 * y - 2 + 5) - 2 * y;
  float sum = 0.0f, wall = 0.0f;
  if (x_mi == -2 && y_mi == -2 && x_ma == 3 && y_ma == 3) {  // whole 5x5 window inside the image: fixed trip counts, immediate weights
#pragma unroll
    for (int yi = -2; yi <= 2; ++yi) {
      const float wy = (yi == 0) ? 0.375f : ((yi == 1 || yi == -1) ? 0.25f : 0.0625f);
#pragma unroll
      for (int xi = -2; xi <= 2; ++xi) {
        const int val = tile[(2 * y + yi - oy) * tstride + (2 * x + xi - ox)];
        if (abs(val - center) < 90) {  // cf.cu:79 compares (float)abs(.) < 90.0f: the same test on integers below 2^24
          const float wx = (xi == 0) ? 0.375f : ((xi == 1 || xi == -1) ? 0.25f : 0.0625f);
          sum = ffma(fmul((float)val, wx), wy, sum);
          wall = ffma(wx, wy, wall);
        }
      }
    }
    return (unsigned short)__float2int_rz(div_t<LEG>(sum, wall));
  }
  for (int yi = y_mi; yi < y_ma; ++yi) {
    const float wy = (yi == 0) ? 0.375f : (abs(yi) == 1 ? 0.25f : 0.0625f);
    for (int xi = x_mi; xi < x_ma; ++xi) {
      const int val = tile[(2 * y + yi - oy) * tstride + (2 * x + xi - ox)];
      if (abs(val - center) < 90) {  // cf.cu:79 compares (float)abs(.) < 90.0f: the same test on integers below 2^24
        const float wx = (xi == 0) ? 0.375f : (abs(xi) == 1 ? 0.25f : 0.0625f);
        sum = ffma(fmul((float)val, wx), wy, sum);
        wall = ffma(wx, wy, wall);
      }
    }
  }
  return (unsigned short)__float2int_rz(div_t<LEG>(sum, wall));
}

// d / 1000.f, correctly rounded, for an integer d in [0, 65535]: product with RN(1/1000), exact remainder, one correction
// (Markstein's scheme; checked against the IEEE quotient for all 65536 inputs) - 3 instructions instead of
// the ~10 of a general IEEE division.
__device__ __forceinline__ float mm_to_m(float d) {
  const float q = fmul(d, 0.001f);
  return ffma(ffma(-q, 1000.f, d), 0.001f, q);
}
template <bool LEG>
__device__ __forceinline__ float3 backproject_d(int d, int u, int v, const Intr& k, float fxi, float fyi, float cutoff) {
  const float z = LEG ? fmul((float)d, 0.001f) : mm_to_m((float)d);  // cf.cu:116; LEG: nvcc folds the division by a constant under --prec-div=false
  if (z != 0 && z < cutoff) return make_float3(fmul(fmul(z, fsub((float)u, k.cx)), fxi), fmul(fmul(z, fsub((float)v, k.cy)), fyi), z);
  return make_float3(qnan(), qnan(), qnan());
}

// vertex + normal of the own block of one level from its shared depth tile
template <bool LEG>
__device__ __forceinline__ void vn_block(const unsigned short* tile, int tstride, int ox, int oy, int bx0, int by0, int bw, int rows, int cols,
                                         const Intr& k, float fxi, float fyi, float cutoff, float* vmap, float* nmap, int pitch) {
  for (int idx = threadIdx.x; idx < bw * bw; idx += blockDim.x) {
    const int ly = idx / bw, lx = idx - ly * bw, u = bx0 + lx, v = by0 + ly;
    if (u >= cols || v >= rows) continue;
    const float3 v00 = backproject_d<LEG>(tile[(v - oy) * tstride + (u - ox)], u, v, k, fxi, fyi, cutoff);
    store_plane3(vmap, pitch, rows, v, u, v00);
    float3 n = make_float3(qnan(), qnan(), qnan());
    if (u < cols - 1 && v < rows - 1) {
      const float3 v01 = backproject_d<LEG>(tile[(v - oy) * tstride + (u + 1 - ox)], u + 1, v, k, fxi, fyi, cutoff);
      const float3 v10 = backproject_d<LEG>(tile[(v + 1 - oy) * tstride + (u - ox)], u, v + 1, k, fxi, fyi, cutoff);
      if (!isnan(v00.x) && !isnan(v01.x) && !isnan(v10.x)) n = normalized3_t<LEG>(cross3(sub3(v01, v00), sub3(v10, v00)));
    }
    store_plane3(nmap, pitch, rows, v, u, n);
  }
}

// Level 0 (three quarters of the pixels) needs no pyramid and no tile: a warp owns a strip of 31 columns x 4 rows, every lane
// back-projects the 5 depths of its column once (1.25 per pixel instead of 3), takes the right neighbour's vertex from the
// next lane (lane 31 only supplies it) and the lower neighbour's from its own next row.
static const int L0_RUN = 4;
template <bool LEG>
__device__ __forceinline__ void icp_depth_level0(const IcpDepthArgs& a, int strip) {
  const int lane = threadIdx.x & 31, rows = a.rows[0], cols = a.cols[0];
  const int wpr = (cols + 30) / 31;
  const int sy = strip / wpr, sx = strip - sy * wpr;
  const int x = sx * 31 + lane, y0 = sy * L0_RUN;
  if (y0 >= rows) return;
  const Intr k = a.k[0];
  const float fxi = a.fxi[0], fyi = a.fyi[0];
  float3 v[L0_RUN + 1];
#pragma unroll
  for (int r = 0; r <= L0_RUN; ++r) {
    const int y = y0 + r;
    const bool in = x < cols && y < rows;
    const unsigned short d = in ? __ldg(row_ptr(a.depth, a.pitch_in, y) + x) : (unsigned short)0;
    if (r < L0_RUN && lane < 31 && in) row_ptr(a.dt[0], a.pitch_u16[0], y)[x] = d;
    v[r] = backproject_d<LEG>(d, x, y, k, fxi, fyi, a.cutoff);
  }
#pragma unroll
  for (int r = 0; r < L0_RUN; ++r) {
    const int y = y0 + r;
    const float3 v00 = v[r], v10 = v[r + 1];
    const float3 v01 = make_float3(__shfl_down_sync(0xffffffffu, v00.x, 1), __shfl_down_sync(0xffffffffu, v00.y, 1), __shfl_down_sync(0xffffffffu, v00.z, 1));
    if (lane == 31 || x >= cols || y >= rows) continue;
    store_plane3(a.vmap[0], a.pitch_f[0], rows, y, x, v00);
    float3 n = make_float3(qnan(), qnan(), qnan());
    if (x < cols - 1 && y < rows - 1 && !isnan(v00.x) && !isnan(v01.x) && !isnan(v10.x)) n = normalized3_t<LEG>(cross3(sub3(v01, v00), sub3(v10, v00)));
    store_plane3(a.nmap[0], a.pitch_f[0], rows, y, x, n);
  }
}

// One launch, two kinds of thread blocks: blocks [0, n_l0) run icp_depth_level0 (8 strips each); the others own a 32x32
// level-0 block each and build the two coarser levels from shared-memory tiles of the u16 depth (45x45 / 21x21 / 9x9, one
// extra row / column for the forward differences of the normals), the tile fetched as 32-bit pairs.
template <bool LEG>
__global__ void __launch_bounds__(256) k_icp_depth(const IcpDepthArgs a, int n_l0, int tiles_x) {
  if ((int)blockIdx.x < n_l0) { icp_depth_level0<LEG>(a, blockIdx.x * 8 + (threadIdx.x >> 5)); return; }
  __shared__ __align__(4) unsigned short t0[45 * 46];
  __shared__ unsigned short t1[21 * 22];
  __shared__ unsigned short t2[9 * 10];
  const int tid = threadIdx.x, tile = blockIdx.x - n_l0, by = tile / tiles_x, bx = tile - by * tiles_x;
  const int ox0 = 32 * bx - 6, oy0 = 32 * by - 6, ox1 = 16 * bx - 2, oy1 = 16 * by - 2, ox2 = 8 * bx, oy2 = 8 * by;
  // rows of 46 = 23 pairs; ox0 is even and rows are 256-byte aligned, so a pair never straddles an alignment boundary
  for (int idx = tid; idx < 45 * 23; idx += 256) {
    const int ly = idx / 23, lp = idx - ly * 23, gx = ox0 + 2 * lp, gy = oy0 + ly;
    unsigned int w = 0u;
    if (gy >= 0 && gy < a.rows[0] && gx >= 0 && gx < a.cols[0]) {
      if (gx + 1 < a.cols[0] && (a.pitch_in & 3) == 0 && ((size_t)a.depth & 3) == 0) w = __ldg(reinterpret_cast<const unsigned int*>(row_ptr(a.depth, a.pitch_in, gy) + gx));
      else w = (unsigned int)__ldg(row_ptr(a.depth, a.pitch_in, gy) + gx) | (gx + 1 < a.cols[0] ? (unsigned int)__ldg(row_ptr(a.depth, a.pitch_in, gy) + gx + 1) << 16 : 0u);
    }
    *reinterpret_cast<unsigned int*>(&t0[ly * 46 + 2 * lp]) = w;
  }
  __syncthreads();
  for (int idx = tid; idx < 21 * 21; idx += 256) {
    const int ly = idx / 21, lx = idx - ly * 21, x = ox1 + lx, y = oy1 + ly;
    if (x < 0 || y < 0 || x >= a.cols[1] || y >= a.rows[1]) continue;
    const unsigned short d = pyrdown_u16_tap<LEG>(t0, 46, ox0, oy0, x, y, a.rows[0], a.cols[0]);
    t1[ly * 22 + lx] = d;
    if (lx >= 2 && lx < 18 && ly >= 2 && ly < 18) row_ptr(a.dt[1], a.pitch_u16[1], y)[x] = d;
  }
  __syncthreads();
  if (tid < 81) {
    const int ly = tid / 9, lx = tid - ly * 9, x = ox2 + lx, y = oy2 + ly;
    if (x < a.cols[2] && y < a.rows[2]) {
      const unsigned short d = pyrdown_u16_tap<LEG>(t1, 22, ox1, oy1, x, y, a.rows[1], a.cols[1]);
      t2[ly * 10 + lx] = d;
      if (lx < 8 && ly < 8) row_ptr(a.dt[2], a.pitch_u16[2], y)[x] = d;
    }
  }
  __syncthreads();
  vn_block<LEG>(t1, 22, ox1, oy1, 16 * bx, 16 * by, 16, a.rows[1], a.cols[1], a.k[1], a.fxi[1], a.fyi[1], a.cutoff, a.vmap[1], a.nmap[1], a.pitch_f[1]);
  vn_block<LEG>(t2, 10, ox2, oy2, 8 * bx, 8 * by, 8, a.rows[2], a.cols[2], a.k[2], a.fxi[2], a.fyi[2], a.cutoff, a.vmap[2], a.nmap[2], a.pitch_f[2]);
}

cudaError_t launch_icp_depth(const Level* lv, const unsigned short* d_depth, int pitch_in, float cutoff, bool legacy, cudaStream_t s) {
  IcpDepthArgs a;
  a.depth = d_depth; a.pitch_in = pitch_in; a.cutoff = cutoff;
  for (int i = 0; i < 3; ++i) {
    a.dt[i] = lv[i].depth_tmp; a.vmap[i] = lv[i].vmap_curr; a.nmap[i] = lv[i].nmap_curr;
    a.pitch_u16[i] = lv[i].pitch_u16; a.pitch_f[i] = lv[i].pitch_f32; a.rows[i] = lv[i].rows; a.cols[i] = lv[i].cols; a.k[i] = lv[i].intr;
    a.fxi[i] = 1.f / lv[i].intr.fx; a.fyi[i] = 1.f / lv[i].intr.fy;
  }
  const int strips = ((lv[0].cols + 30) / 31) * ((lv[0].rows + L0_RUN - 1) / L0_RUN);
  const int n_l0 = (strips + 7) / 8, tiles_x = (lv[0].cols + 31) / 32, tiles_y = (lv[0].rows + 31) / 32;
  const int grid = n_l0 + tiles_x * tiles_y;
  if (legacy) k_icp_depth<true><<<grid, 256, 0, s>>>(a, n_l0, tiles_x); else k_icp_depth<false><<<grid, 256, 0, s>>>(a, n_l0, tiles_x);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ depth pre-filter (the step before initICP(depth))
// exp(-t), t >= 0, in explicit fp32 steps (bit-identical to the oracle's; GLSL's exp is implementation defined)
__device__ __forceinline__ float exp_neg(float t) {
  const float a = fmul(t, 1.44269504f);
  const float n = rintf(a);
  if (!(n < 126.0f)) return 0.0f;
  const float f = fsub(a, n);
  float p = 1.54035304e-4f;
  p = ffma(p, f, -1.33335581e-3f);
  p = ffma(p, f, 9.61812911e-3f);
  p = ffma(p, f, -5.55041087e-2f);
  p = ffma(p, f, 2.40226507e-1f);
  p = ffma(p, f, -6.93147181e-1f);
  p = ffma(p, f, 1.0f);
  return fmul(p, __int_as_float((127 - (int)n) << 23));
}

// Shaders/depth_bilateral.frag:31-75: 13x13 bilateral filter on raw u16 millimetres.  32x8 outputs per CTA from a
// 44x20 shared tile.
__global__ void __launch_bounds__(256) k_depth_bilateral(const unsigned short* src, int sp, unsigned short* dst, int dp, int rows, int cols,
                                                         float maxD) {
  __shared__ unsigned short tile[20][46];
  const int x0 = blockIdx.x * 32, y0 = blockIdx.y * 8;
  for (int idx = threadIdx.x; idx < 20 * 44; idx += 256) {
    const int ly = idx / 44, lx = idx - ly * 44, gx = x0 - 6 + lx, gy = y0 - 6 + ly;
    tile[ly][lx] = (gx >= 0 && gy >= 0 && gx < cols && gy < rows) ? __ldg(row_ptr(src, sp, gy) + gx) : (unsigned short)0;
  }
  __syncthreads();
  const int lx = threadIdx.x & 31, ly = threadIdx.x >> 5, x = x0 + lx, y = y0 + ly;
  if (x >= cols || y >= rows) return;
  const unsigned int value = tile[ly + 6][lx + 6];
  const unsigned int hi = (unsigned int)fmul(maxD, 1000.0f);
  if (value > hi || value < 300U) { row_ptr(dst, dp, y)[x] = 0; return; }
  const int tx = min(x - 6 + 13, cols), ty = min(y - 6 + 13, rows);
  float sum1 = 0.f, sum2 = 0.f;
  for (int cy = max(y - 6, 0); cy < ty; ++cy)
    for (int cx = max(x - 6, 0); cx < tx; ++cx) {
      const unsigned int tmp = tile[cy - y0 + 6][cx - x0 + 6];
      const float dx = fsub((float)x, (float)cx), dy = fsub((float)y, (float)cy);
      const float space2 = ffma(dx, dx, fmul(dy, dy));
      const float dc = fsub((float)value, (float)tmp);
      const float weight = exp_neg(ffma(space2, 0.024691358f, fmul(fmul(dc, dc), 0.000555556f)));
      sum1 = ffma((float)tmp, weight, sum1);
      sum2 = fadd(sum2, weight);
    }
  row_ptr(dst, dp, y)[x] = (unsigned short)(unsigned int)roundf(fdiv(sum1, sum2));
}

// Shaders/depth_metric.frag:29-39
__global__ void __launch_bounds__(256) k_depth_metric(const unsigned short* src, int sp, float* dst, int dp, int rows, int cols, float maxD) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= cols || y >= rows) return;
  const unsigned int value = __ldg(row_ptr(src, sp, y) + x);
  const unsigned int hi = (unsigned int)fmul(maxD, 1000.0f);
  row_ptr(dst, dp, y)[x] = (value > hi || value < 300U) ? 0.0f : fdiv((float)value, 1000.0f);
}

cudaError_t launch_depth_bilateral(const unsigned short* src, int sp, unsigned short* dst, int dp, int rows, int cols, float maxD, cudaStream_t s) {
  k_depth_bilateral<<<dim3((cols + 31) / 32, (rows + 7) / 8), 256, 0, s>>>(src, sp, dst, dp, rows, cols, maxD);
  return cudaGetLastError();
}
cudaError_t launch_depth_metric(const unsigned short* src, int sp, float* dst, int dp, int rows, int cols, float maxD, cudaStream_t s) {
  k_depth_metric<<<dim3((cols + 31) / 32, (rows + 7) / 8), 256, 0, s>>>(src, sp, dst, dp, rows, cols, maxD);
  return cudaGetLastError();
}

}  // namespace dms
