// Model-view synthesis (include/dmsodom/splat_c.h): IndexMap::combinedPredict (IndexMap.cpp:253-368) with
// Shaders/splat.vert:42-95 and Shaders/combo_splat.frag:36-66 restated as a CUDA rasteriser.
#include <math.h>
#include <string.h>
#include "common.cuh"
#include "kernels.h"
#include "../../include/dmsodom/splat_c.h"

namespace dms {

struct SplatK {  // kernel-side copy of dms_splat_params
  float t[16];
  float cx, cy, fx, fy;
  int rows, cols;
  float maxDepth, confThreshold;
  int time, timeIdx, maxTime, timeDelta, active;
};

// unfused, left to right
__device__ __forceinline__ float dot3u(float ax, float ay, float az, float bx, float by, float bz) { return fadd(fadd(fmul(ax, bx), fmul(ay, by)), fmul(az, bz)); }
__device__ __forceinline__ float3 normalize3u(float3 v) {
  const float inv = fdiv(1.0f, fsqrt(dot3u(v.x, v.y, v.z, v.x, v.y, v.z)));
  return make_float3(fmul(v.x, inv), fmul(v.y, inv), fmul(v.z, inv));
}

struct SplatVert { float3 pos; float conf; float3 n; float rad; float xw, yw, size; float colour, initTime; bool visible; };

// splat.vert:42-95
__device__ __forceinline__ SplatVert splat_vertex(const float* s, const SplatK& k) {
  SplatVert v;
  const float px = s[0], py = s[1], pz = s[2];
  v.conf = s[3]; v.colour = s[4]; v.initTime = s[6];
  const float vt = s[8 + k.timeIdx];
  const float* t = k.t;
  v.pos.x = fadd(dot3u(t[0], t[1], t[2], px, py, pz), t[3]);
  v.pos.y = fadd(dot3u(t[4], t[5], t[6], px, py, pz), t[7]);
  v.pos.z = fadd(dot3u(t[8], t[9], t[10], px, py, pz), t[11]);
  const bool t3 = (vt == -3.0f), actv = k.active != 0;
  const bool reject = !(!actv && t3) && (v.pos.z > k.maxDepth || v.pos.z < 0 || v.conf < k.confThreshold || (actv && t3) ||
                                         (!t3 && fsub((float)k.time, vt) > (float)k.timeDelta) || vt > (float)k.maxTime);
  v.visible = !reject;
  v.rad = s[14];
  v.n = normalize3u(make_float3(dot3u(t[0], t[1], t[2], s[11], s[12], s[13]), dot3u(t[4], t[5], t[6], s[11], s[12], s[13]),
                                dot3u(t[8], t[9], t[10], s[11], s[12], s[13])));
  const float3 u = normalize3u(make_float3(fsub(v.n.y, v.n.z), -v.n.x, v.n.x));
  const float3 x1 = make_float3(fmul(fmul(u.x, v.rad), 1.41421356f), fmul(fmul(u.y, v.rad), 1.41421356f), fmul(fmul(u.z, v.rad), 1.41421356f));
  const float3 y1 = make_float3(fsub(fmul(v.n.y, x1.z), fmul(x1.y, v.n.z)), fsub(fmul(v.n.z, x1.x), fmul(x1.z, v.n.x)),
                                fsub(fmul(v.n.x, x1.y), fmul(x1.x, v.n.y)));
  float xmin = 0.f, xmax = 0.f, ymin = 0.f, ymax = 0.f;
#pragma unroll
  for (int c = 0; c < 4; ++c) {  // proj1..proj4 = pos + x1, pos + y1, pos - y1, pos - x1
    const float3 d = c == 0 ? x1 : (c == 1 ? y1 : (c == 2 ? make_float3(-y1.x, -y1.y, -y1.z) : make_float3(-x1.x, -x1.y, -x1.z)));
    const float qx = fadd(v.pos.x, d.x), qy = fadd(v.pos.y, d.y), qz = fadd(v.pos.z, d.z);
    const float ix = fadd(fdiv(fmul(k.fx, qx), qz), k.cx), iy = fadd(fdiv(fmul(k.fy, qy), qz), k.cy);
    if (c == 0) { xmin = xmax = ix; ymin = ymax = iy; }
    else { xmin = fminf(xmin, ix); xmax = fmaxf(xmax, ix); ymin = fminf(ymin, iy); ymax = fmaxf(ymax, iy); }
  }
  v.size = fmaxf(0.0f, fmaxf(fabsf(fsub(xmax, xmin)), fabsf(fsub(ymax, ymin))));
  v.xw = fadd(fdiv(fmul(k.fx, v.pos.x), v.pos.z), k.cx);
  v.yw = fadd(fdiv(fmul(k.fy, v.pos.y), v.pos.z), k.cy);
  // clipped when the centre leaves the viewport (or the size is zero / not a number)
  if (!(v.xw >= 0.0f && v.xw <= (float)k.cols && v.yw >= 0.0f && v.yw <= (float)k.rows && v.size > 0.0f)) v.visible = false;
  return v;
}

// combo_splat.frag:36-66 at pixel (i, j); returns false when the fragment is discarded / clipped
// un-normalised ray components of a pixel column / row: (gl_FragCoord - cam.xy) / cam.zw
__device__ __forceinline__ float splat_ray_x(const SplatK& k, int i) { return fdiv(fsub(fadd((float)i, 0.5f), k.cx), k.fx); }
__device__ __forceinline__ float splat_ray_y(const SplatK& k, int j) { return fdiv(fsub(fadd((float)j, 0.5f), k.cy), k.fy); }
__device__ __forceinline__ bool splat_fragment(const SplatVert& v, const SplatK& k, float rx, float ry, float* z_out, unsigned int* d24) {
  const float3 l = normalize3u(make_float3(rx, ry, 1.0f));
  const float sc = fdiv(dot3u(v.pos.x, v.pos.y, v.pos.z, v.n.x, v.n.y, v.n.z), dot3u(l.x, l.y, l.z, v.n.x, v.n.y, v.n.z));
  const float3 cp = make_float3(fmul(sc, l.x), fmul(sc, l.y), fmul(sc, l.z));
  const float dx = fsub(cp.x, v.pos.x), dy = fsub(cp.y, v.pos.y), dz = fsub(cp.z, v.pos.z);
  if (dot3u(dx, dy, dz, dx, dy, dz) > fmul(v.rad, v.rad)) return false;
  const float depth = fadd(fdiv(cp.z, fmul(2.0f, k.maxDepth)), 0.5f);
  if (!(depth >= 0.0f && depth <= 1.0f)) return false;
  *z_out = cp.z;
  *d24 = (unsigned int)((double)depth * 16777215.0 + 0.5);
  return true;
}

__device__ __forceinline__ void splat_box(const SplatVert& v, const SplatK& k, int* x0, int* x1, int* y0, int* y1) {
  const float h = fmul(v.size, 0.5f);
  *x0 = max(0, (int)ceilf(fsub(fsub(v.xw, h), 0.5f))); *x1 = min(k.cols - 1, (int)floorf(fsub(fadd(v.xw, h), 0.5f)));
  *y0 = max(0, (int)ceilf(fsub(fsub(v.yw, h), 0.5f))); *y1 = min(k.rows - 1, (int)floorf(fsub(fadd(v.yw, h), 0.5f)));
}

static const unsigned long long ZBUF_EMPTY = ~0ull;

__global__ void __launch_bounds__(256) k_splat_raster(const float* surfels, int count, SplatK k, unsigned long long* zbuf) {
  extern __shared__ float s_ray[];  // cols x-components, then rows y-components: the two divisions of a fragment's ray, once per CTA
  for (int i = threadIdx.x; i < k.cols; i += blockDim.x) s_ray[i] = splat_ray_x(k, i);
  for (int j = threadIdx.x; j < k.rows; j += blockDim.x) s_ray[k.cols + j] = splat_ray_y(k, j);
  __syncthreads();
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < count; idx += gridDim.x * blockDim.x) {
    const SplatVert v = splat_vertex(surfels + (size_t)idx * DMS_SURFEL_FLOATS, k);
    if (!v.visible) continue;
    int x0, x1, y0, y1;
    splat_box(v, k, &x0, &x1, &y0, &y1);
    // Conservative pre-cull (never rejects a fragment the shader test accepts): a point of the disk is within rad of its
    // centre in 3-D, so with z > rad its projection stays within f * rad * (1 + max(|x|, |y|) / z) / (z - rad) pixels of
    // the centre's projection; 2 % + half a pixel cover the rounding of the exact test.  Rows are clipped to that circle.
    float R2 = 3.0e38f, Rr = 3.0e38f;
    if (v.pos.z > v.rad * 1.001f) {
      const float f = fmaxf(k.fx, k.fy);
      Rr = f * v.rad * (1.0f + fmaxf(fabsf(v.pos.x), fabsf(v.pos.y)) / v.pos.z) / (v.pos.z - v.rad) * 1.02f + 0.5f;
      R2 = Rr * Rr;
    }
    for (int j = y0; j <= y1; ++j) {
      const float dyp = (float)j + 0.5f - v.yw;
      const float rem = R2 - dyp * dyp;
      if (!(rem >= 0.0f)) continue;
      int xa = x0, xb = x1;
      if (Rr < 1.0e30f) {
        const float hw = sqrtf(rem) + 0.01f;
        xa = max(x0, (int)floorf(v.xw - hw - 0.5f)); xb = min(x1, (int)ceilf(v.xw + hw - 0.5f));
      }
      const float ry = s_ray[k.cols + j];
      for (int i = xa; i <= xb; ++i) {
        float z; unsigned int d;
        if (splat_fragment(v, k, s_ray[i], ry, &z, &d)) atomicMin(&zbuf[(size_t)j * k.cols + i], ((unsigned long long)d << 32) | (unsigned int)idx);
      }
    }
  }
}

__global__ void __launch_bounds__(256) k_splat_resolve(const float* surfels, SplatK k, const unsigned long long* zbuf, uchar4* image, float4* vertex,
                                                       float4* normal, unsigned int* time) {
  const int i = blockIdx.x * 32 + (threadIdx.x & 31), j = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (i >= k.cols || j >= k.rows) return;
  const size_t p = (size_t)j * k.cols + i;
  const unsigned long long key = zbuf[p];
  uchar4 im = make_uchar4(0, 0, 0, 0); float4 ve = make_float4(0.f, 0.f, 0.f, 0.f), no = ve; unsigned int ti = 0u;
  if (key != ZBUF_EMPTY) {
    const SplatVert v = splat_vertex(surfels + (size_t)(unsigned int)key * DMS_SURFEL_FLOATS, k);
    float z = 0.f; unsigned int d;
    splat_fragment(v, k, splat_ray_x(k, i), splat_ray_y(k, j), &z, &d);
    const int c = (int)v.colour;  // color.glsl:27-34; the RGBA8 attachment stores round(c / 255 * 255) = c
    im = make_uchar4((unsigned char)((c >> 16) & 0xFF), (unsigned char)((c >> 8) & 0xFF), (unsigned char)(c & 0xFF), 255);
    const float fxc = fadd((float)i, 0.5f), fyc = fadd((float)j, 0.5f);
    ve = make_float4(fmul(fmul(fsub(fxc, k.cx), z), fdiv(1.0f, k.fx)), fmul(fmul(fsub(fyc, k.cy), z), fdiv(1.0f, k.fy)), z, v.conf);
    no = make_float4(v.n.x, v.n.y, v.n.z, v.rad);
    ti = (unsigned int)v.initTime;
  }
  if (image) image[p] = im;
  if (vertex) vertex[p] = ve;
  if (normal) normal[p] = no;
  if (time) time[p] = ti;
}

}  // namespace dms

using namespace dms;

extern "C" {

int dms_splat_invert_pose(const float pose[16], float t_inv[16]) {
  if (!pose || !t_inv) return DMS_ERR_INVALID;
  // rigid or not, a general inverse as Eigen's Matrix4f::inverse() (IndexMap.cpp:297): Gauss-Jordan in double
  double a[4][8];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) { a[i][j] = pose[i * 4 + j]; a[i][4 + j] = i == j ? 1.0 : 0.0; }
  for (int c = 0; c < 4; ++c) {
    int piv = c;
    for (int r = c + 1; r < 4; ++r) if (fabs(a[r][c]) > fabs(a[piv][c])) piv = r;
    if (a[piv][c] == 0.0) return DMS_ERR_INVALID;
    if (piv != c) for (int j = 0; j < 8; ++j) { const double t = a[c][j]; a[c][j] = a[piv][j]; a[piv][j] = t; }
    const double d = a[c][c];
    for (int j = 0; j < 8; ++j) a[c][j] /= d;
    for (int r = 0; r < 4; ++r) if (r != c) { const double f = a[r][c]; for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j]; }
  }
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) t_inv[i * 4 + j] = (float)a[i][4 + j];
  return DMS_OK;
}

int dms_splat_predict(const float* d_surfels, int count, const dms_splat_params* p, unsigned long long* d_zbuf, uint8_t* d_image, float* d_vertex,
                      float* d_normal, uint32_t* d_time, void* stream) {
  if (!p || !d_zbuf || count < 0 || (count > 0 && !d_surfels) || p->rows <= 0 || p->cols <= 0 || p->timeIdx < 0 || p->timeIdx > 2) return DMS_ERR_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  SplatK k;
  memcpy(k.t, p->t_inv, sizeof(k.t));
  k.cx = p->cx; k.cy = p->cy; k.fx = p->fx; k.fy = p->fy; k.rows = p->rows; k.cols = p->cols; k.maxDepth = p->maxDepth; k.confThreshold = p->confThreshold;
  k.time = p->time; k.timeIdx = p->timeIdx; k.maxTime = p->maxTime; k.timeDelta = p->timeDelta; k.active = p->active;
  cudaError_t e = cudaMemsetAsync(d_zbuf, 0xff, (size_t)p->rows * p->cols * sizeof(unsigned long long), s);  // glClear
  if (e != cudaSuccess) return (int)e;
  if (count > 0) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int grid = (count + 255) / 256;
    if (grid > sms * 8) grid = sms * 8;
    const size_t ray_bytes = (size_t)(p->rows + p->cols) * sizeof(float);
    if (ray_bytes > 48 * 1024) return DMS_ERR_TOO_LARGE;  // rows + cols <= 12288
    k_splat_raster<<<grid, 256, ray_bytes, s>>>(d_surfels, count, k, d_zbuf);
    e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
  }
  k_splat_resolve<<<dim3((p->cols + 31) / 32, (p->rows + 7) / 8), 256, 0, s>>>(d_surfels, k, d_zbuf, (uchar4*)d_image, (float4*)d_vertex, (float4*)d_normal,
                                                                               d_time);
  e = cudaGetLastError();
  return e == cudaSuccess ? DMS_OK : (int)e;
}

}  // extern "C"
