// NID keyframing (include/dmsodom/nid_c.h): joint histograms + entropies on the device.
// Reference: cudafuncs.cu:906-917, 1086-1157, 1513-1650, 1794-1915; MutualInformation.cpp:154-212; KeyFrame.h:69-160.
#include <math.h>
#include <string.h>
#include <new>
#include "common.cuh"
#include "kernels.h"
#include "../../include/dmsodom/nid_c.h"

namespace dms {

// cf.cu:597-608 verticesToDepthKernel on a dense RGBA32F vertex image
__global__ void __launch_bounds__(256) k_vertices_to_depth(const float4* v, int rows, int cols, float cutoff, float* dst, int pitch) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= cols || y >= rows) return;
  const float z = __ldg(&v[(size_t)y * cols + x]).z;
  row_ptr(dst, pitch, y)[x] = (z > cutoff || z <= 0) ? qnan() : z;
}

// The key-frame sample of a pixel (cf.cu:1097-1112 / 1135-1150).  The reference tests validity with
// `d != __int_as_float(0x7fffffff)`, which is true for every value including NaN, so only its first branch is live:
// active if dmap_kf <= dmap_kf_old (false when either is NaN), else old.
__device__ __forceinline__ bool nid_pick_active(float dk, float dko) { return dk <= dko; }

static const int NID_SMEM_BINS = 64;  // bins per side that fit a privatised shared-memory histogram (16 KB)

// joint histogram of image bins: row = live bin, column = key-frame bin (cf.cu:1086-1122)
template <bool SMEM>
__global__ void __launch_bounds__(256) k_nid_hist_img(const unsigned char* ik, const unsigned char* iko, const float* dk, const float* dko,
                                                     const unsigned char* ic, int p8, int pf, int pc, int rows, int cols, int nb,
                                                     unsigned int* hist) {
  __shared__ unsigned int sh[SMEM ? NID_SMEM_BINS * NID_SMEM_BINS : 1];
  if (SMEM) { for (int i = threadIdx.x; i < nb * nb; i += blockDim.x) sh[i] = 0u; __syncthreads(); }
  const int bw = 256 / nb;  // bin_idx(unsigned char&, num_bins), cf.cu:906-911
  const int spr = (cols + 31) >> 5, nseg = rows * spr, lane = threadIdx.x & 31;
  for (int seg = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); seg < nseg; seg += gridDim.x * (blockDim.x >> 5)) {
    const int y = seg / spr, x = (seg - y * spr) * 32 + lane;
    if (x >= cols) continue;
    const bool act = nid_pick_active(__ldg(row_ptr(dk, pf, y) + x), __ldg(row_ptr(dko, pf, y) + x));
    const int a = act ? __ldg(row_ptr(ik, p8, y) + x) : __ldg(row_ptr(iko, p8, y) + x);
    const int b = __ldg(row_ptr(ic, pc, y) + x);
    const int ba = a / bw, bb = b / bw;
    if (ba >= nb || bb >= nb) continue;  // the reference writes out of bounds here
    if (SMEM) atomicAdd(&sh[bb * nb + ba], 1u); else atomicAdd(&hist[bb * nb + ba], 1u);
  }
  if (SMEM) {
    __syncthreads();
    for (int i = threadIdx.x; i < nb * nb; i += blockDim.x) { const unsigned int v = sh[i]; if (v) atomicAdd(&hist[i], v); }
  }
}

// joint histogram of depth bins (cf.cu:1124-1157); bin_idx(float&, max_depth, num_bins) cf.cu:913-917
__device__ __forceinline__ int nid_depth_bin(float d_mm, int bw) { return (int)fdiv(d_mm, (float)bw); }  // NaN -> 0 as on the device

__global__ void __launch_bounds__(256) k_nid_hist_depth(const float* dk, const float* dko, const float* dc, int pf, int pc, int rows, int cols, int nb,
                                                       float max_depth, unsigned int* hist) {
  const int bw = (int)fdiv(max_depth, (float)nb);
  const int spr = (cols + 31) >> 5, nseg = rows * spr, lane = threadIdx.x & 31;
  for (int seg = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); seg < nseg; seg += gridDim.x * (blockDim.x >> 5)) {
    const int y = seg / spr, x = (seg - y * spr) * 32 + lane;
    if (x >= cols) continue;
    const float k = __ldg(row_ptr(dk, pf, y) + x), ko = __ldg(row_ptr(dko, pf, y) + x);
    const float a = fmul(nid_pick_active(k, ko) ? k : ko, 1000.0f);
    const float b = fmul(__ldg(row_ptr(dc, pc, y) + x), 1000.0f);
    const int ba = nid_depth_bin(a, bw), bb = nid_depth_bin(b, bw);
    if (ba < 0 || bb < 0 || ba >= nb || bb >= nb) continue;  // out of bounds in the reference
    // neighbouring pixels mostly fall into the same bin pair: one atomic per distinct pair of the warp
    const int key = bb * nb + ba;
    const unsigned int peers = __match_any_sync(__activemask(), key);
    if ((int)(__ffs((int)peers) - 1) == lane) atomicAdd(&hist[key], (unsigned int)__popc(peers));
  }
}

// Entropies and the score (cf.cu:1556-1615).  p = count / num_points in fp32, p * log2f(p) in fp32, sums in fp64.
// Pass 1 (one CTA per group of histogram rows): joint-entropy and live-marginal partial sums per CTA, key-frame
// marginal counts (columns) by integer atomics.  Pass 2 (one CTA): key-frame marginal entropy, the partial sums in
// CTA order (deterministic), the score.  out[0] = nid, out[1] = joint entropy, out[2] = key-frame, out[3] = live.
static const int NID_ENT_CTAS = 128;
__device__ __forceinline__ double block_sum(double s, double* red) {
  const int t = threadIdx.x;
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  __syncthreads();
  if ((t & 31) == 0) red[t >> 5] = s;
  __syncthreads();
  if (t < 32) {
    s = t < (int)(blockDim.x >> 5) ? red[t] : 0.0;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  }
  return s;  // valid in thread 0
}
__global__ void __launch_bounds__(256) k_nid_entropy_rows(const unsigned int* hist, int nb, int num_points, unsigned int* colsum, double* partial) {
  __shared__ double red[32];
  __shared__ unsigned int rowsum;
  const int t = threadIdx.x;
  const float np = (float)num_points;
  double joint = 0.0, cf = 0.0;
  for (int r = blockIdx.x; r < nb; r += gridDim.x) {
    if (t == 0) rowsum = 0u;
    __syncthreads();
    unsigned int mine = 0u;
    for (int col = t; col < nb; col += blockDim.x) {
      const unsigned int c = hist[(size_t)r * nb + col];
      if (c) {
        mine += c;
        atomicAdd(&colsum[col], c);
        const float p = fdiv((float)c, np);
        joint += (double)fmul(p, log2f(p));
      }
    }
    mine = __reduce_add_sync(0xffffffffu, mine);
    if ((t & 31) == 0 && mine) atomicAdd(&rowsum, mine);
    __syncthreads();
    if (t == 0 && rowsum) { const float p = fdiv((float)rowsum, np); cf += (double)fmul(p, log2f(p)); }
    __syncthreads();
  }
  const double js = block_sum(joint, red);
  if (t == 0) { partial[2 * blockIdx.x] = js; partial[2 * blockIdx.x + 1] = cf; }
}
__global__ void __launch_bounds__(256) k_nid_entropy_final(const unsigned int* colsum, int nb, int num_points, const double* partial, int nparts, float* out) {
  __shared__ double red[32];
  const int t = threadIdx.x;
  const float np = (float)num_points;
  double kf = 0.0;
  for (int i = t; i < nb; i += blockDim.x) {
    const unsigned int a = colsum[i];
    if (a) { const float p = fdiv((float)a, np); kf += (double)fmul(p, log2f(p)); }
  }
  const double kfs = block_sum(kf, red);
  // the per-CTA partial sums: fixed tree (thread i takes partial i), hence deterministic
  const double joint = block_sum(t < nparts ? partial[2 * t] : 0.0, red);
  const double cf = block_sum(t < nparts ? partial[2 * t + 1] : 0.0, red);
  if (t == 0) {
    const float joint_entropy = (float)(-joint), kf_e = (float)(-kfs), cf_e = (float)(-cf);
    const float mutual = fsub(fadd(kf_e, cf_e), joint_entropy);
    out[0] = num_points == 0 ? 1.0f : fdiv(fsub(joint_entropy, mutual), joint_entropy);
    out[1] = joint_entropy; out[2] = kf_e; out[3] = cf_e;
  }
}
// marg: nb u32 column sums followed (8-byte aligned) by 2 * NID_ENT_CTAS doubles
static size_t nid_marg_bytes(int nb) { return ((size_t)nb * 4 + 7) / 8 * 8 + 2 * NID_ENT_CTAS * sizeof(double); }
static cudaError_t launch_nid_entropy(const unsigned int* hist, int nb, int num_points, unsigned int* marg, float* out, cudaStream_t s) {
  double* partial = reinterpret_cast<double*>(reinterpret_cast<char*>(marg) + ((size_t)nb * 4 + 7) / 8 * 8);
  cudaError_t e = cudaMemsetAsync(marg, 0, (size_t)nb * 4, s);
  if (e != cudaSuccess) return e;
  const int ctas = nb < NID_ENT_CTAS ? nb : NID_ENT_CTAS;
  k_nid_entropy_rows<<<ctas, 256, 0, s>>>(hist, nb, num_points, marg, partial);
  k_nid_entropy_final<<<1, 256, 0, s>>>(marg, nb, num_points, partial, ctas, out);
  return cudaGetLastError();
}

static int nid_grid() {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms * 4;
}

}  // namespace dms

using namespace dms;

struct dms_nid {
  int device, width, height, level, nb_img, nb_depth;
  Level lv[3];                 // lastImage/lastDepth = key frame (active), nextImage/nextDepth = key frame (old)
  unsigned char* slab;
  float* zsrc;                 // verticesToDepth output, level-0 pitch
  unsigned int *hist_img, *hist_depth, *marg;
  float* d_out;                // 8 floats
  float* h_out;                // pinned
  int have_keyframe;
};

extern "C" {

int dms_nid_destroy(dms_nid* h) {
  if (!h) return DMS_OK;
  cudaSetDevice(h->device);
  if (h->slab) cudaFree(h->slab);
  if (h->h_out) cudaFreeHost(h->h_out);
  delete h;
  return DMS_OK;
}

int dms_nid_create(dms_nid** out, int device, int width, int height, int pyr_level, int nb_img, int nb_depth) {
  if (!out || width <= 0 || height <= 0 || pyr_level < 0 || pyr_level > 2 || nb_img <= 0 || nb_img > 256 || nb_depth <= 0 || nb_depth > 4096)
    return DMS_ERR_INVALID;
  *out = nullptr;
  dms_nid* h = new (std::nothrow) dms_nid();
  if (!h) return DMS_ERR_NOMEM;
  memset(h, 0, sizeof(*h));
  h->device = device; h->width = width; h->height = height; h->level = pyr_level; h->nb_img = nb_img; h->nb_depth = nb_depth;
  cudaSetDevice(device);
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  size_t total = 0;
  for (int l = 0; l < 3; ++l) {
    Level& L = h->lv[l];
    L.rows = height >> l; L.cols = width >> l;
    L.pitch_f32 = (int)al((size_t)L.cols * 4); L.pitch_u8 = (int)al((size_t)L.cols);
    total += 2 * al((size_t)L.rows * L.pitch_f32) + 2 * al((size_t)L.rows * L.pitch_u8);
  }
  const size_t zbytes = al((size_t)height * h->lv[0].pitch_f32);
  const size_t hi = al((size_t)nb_img * nb_img * 4), hd = al((size_t)nb_depth * nb_depth * 4);
  const size_t mg = al(nid_marg_bytes(nb_img > nb_depth ? nb_img : nb_depth));
  total += zbytes + hi + hd + 2 * mg + 256;
  cudaError_t e = cudaMalloc(&h->slab, total);
  if (e == cudaSuccess) e = cudaMemset(h->slab, 0, total);
  if (e == cudaSuccess) e = cudaMallocHost(&h->h_out, 64);
  if (e != cudaSuccess) { dms_nid_destroy(h); return e == cudaErrorMemoryAllocation ? DMS_ERR_NOMEM : (int)e; }
  unsigned char* p = h->slab;
  for (int l = 0; l < 3; ++l) {
    Level& L = h->lv[l];
    L.lastDepth = (float*)p; p += al((size_t)L.rows * L.pitch_f32);
    L.nextDepth = (float*)p; p += al((size_t)L.rows * L.pitch_f32);
    L.lastImage = p; p += al((size_t)L.rows * L.pitch_u8);
    L.nextImage = p; p += al((size_t)L.rows * L.pitch_u8);
  }
  h->zsrc = (float*)p; p += zbytes;
  h->hist_img = (unsigned int*)p; p += hi;
  h->hist_depth = (unsigned int*)p; p += hd;
  h->marg = (unsigned int*)p; p += 2 * mg;  // one block per histogram
  h->d_out = (float*)p;
  *out = h;
  return DMS_OK;
}

int dms_nid_set_keyframe(dms_nid* h, const float* v4, const uint8_t* rgba, const float* old_v4, const uint8_t* old_rgba, float depthCutoff,
                         void* stream) {
  if (!h || !v4 || !rgba || !old_v4 || !old_rgba) return DMS_ERR_INVALID;
  cudaSetDevice(h->device);
  cudaStream_t s = (cudaStream_t)stream;
  const dim3 grid((h->width + 31) / 32, (h->height + 7) / 8);
  for (int which = 0; which < 2; ++which) {
    k_vertices_to_depth<<<grid, 256, 0, s>>>((const float4*)(which ? old_v4 : v4), h->height, h->width, depthCutoff, h->zsrc, h->lv[0].pitch_f32);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = launch_populate(h->lv, which, h->zsrc, (const uchar4*)(which ? old_rgba : rgba), h->width * 4, false, s);
    if (e != cudaSuccess) return (int)e;
  }
  h->have_keyframe = 1;
  return DMS_OK;
}

static int nid_finish(const unsigned int* hist, int nb, int num_points, unsigned int* marg, float* d_out, float* nid, cudaStream_t s) {
  cudaError_t e = launch_nid_entropy(hist, nb, num_points, marg, d_out, s);
  float r[4];
  if (e == cudaSuccess) e = cudaMemcpyAsync(r, d_out, sizeof(r), cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return (int)e;
  *nid = r[0];
  return DMS_OK;
}

// d_out doubles as the marginal scratch holder in the stateless forms: [0..3] result, marginals right behind the histogram
int dms_nid_image(const uint8_t* ik, const uint8_t* iko, const float* dk, const float* dko, const uint8_t* ic, size_t p8, size_t pf, size_t pc, int rows,
                  int cols, int nb, uint32_t* d_hist, float* d_out, float* nid, void* stream) {
  if (!ik || !iko || !dk || !dko || !ic || !d_hist || !d_out || !nid || rows <= 0 || cols <= 0 || nb <= 0 || nb > 256) return DMS_ERR_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  // histogram + 2*nb marginal words (the caller's scratch holds nb*nb + 2*nb words)
  cudaError_t e = cudaMemsetAsync(d_hist, 0, (size_t)nb * nb * 4, s);
  if (e != cudaSuccess) return (int)e;
  if (nb <= NID_SMEM_BINS)
    k_nid_hist_img<true><<<nid_grid(), 256, 0, s>>>(ik, iko, dk, dko, ic, (int)p8, (int)pf, (int)pc, rows, cols, nb, d_hist);
  else
    k_nid_hist_img<false><<<nid_grid(), 256, 0, s>>>(ik, iko, dk, dko, ic, (int)p8, (int)pf, (int)pc, rows, cols, nb, d_hist);
  e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  return nid_finish(d_hist, nb, rows * cols, d_hist + ((size_t)nb * nb + 1) / 2 * 2, d_out, nid, s);
}

int dms_nid_depth(const float* dk, const float* dko, const float* dc, size_t pf, size_t pc, int rows, int cols, int nb, float max_depth_mm, uint32_t* d_hist,
                  float* d_out, float* nid, void* stream) {
  if (!dk || !dko || !dc || !d_hist || !d_out || !nid || rows <= 0 || cols <= 0 || nb <= 0 || nb > 4096) return DMS_ERR_INVALID;
  if ((int)(max_depth_mm / (float)nb) <= 0) return DMS_ERR_INVALID;  // bin width 0: the reference divides by zero
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(d_hist, 0, (size_t)nb * nb * 4, s);
  if (e != cudaSuccess) return (int)e;
  k_nid_hist_depth<<<nid_grid(), 256, 0, s>>>(dk, dko, dc, (int)pf, (int)pc, rows, cols, nb, max_depth_mm, d_hist);
  e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  return nid_finish(d_hist, nb, rows * cols, d_hist + ((size_t)nb * nb + 1) / 2 * 2, d_out, nid, s);
}

int dms_nid_score(dms_nid* h, dms_odom* tracker, float max_depth_m, float lambda, float* nid_img, float* nid_depth, float* score, void* stream) {
  if (!h || !tracker || !h->have_keyframe) return DMS_ERR_INVALID;
  cudaSetDevice(h->device);
  void *pi = nullptr, *pd = nullptr;
  size_t pitch_i = 0, pitch_d = 0;
  int r = 0, c = 0, eb = 0;
  int rc = dms_odom_get_buffer(tracker, DMS_BUF_NEXT_IMAGE, h->level, &pi, &pitch_i, &r, &c, &eb);   // nextImg(level), EF.cpp:653
  if (rc == DMS_OK) rc = dms_odom_get_buffer(tracker, DMS_BUF_NEXT_DEPTH, h->level, &pd, &pitch_d, &r, &c, &eb);  // nextD(level), EF.cpp:658
  if (rc != DMS_OK) return rc;
  const Level& L = h->lv[h->level];
  if (r != L.rows || c != L.cols) return DMS_ERR_INVALID;
  float ni = 0.f, nd = 0.f;
  // the stateless forms want the marginal scratch right behind the histogram: use the handle's own buffers through the kernels directly
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(h->hist_img, 0, (size_t)h->nb_img * h->nb_img * 4, s);
  if (e == cudaSuccess) e = cudaMemsetAsync(h->hist_depth, 0, (size_t)h->nb_depth * h->nb_depth * 4, s);
  if (e != cudaSuccess) return (int)e;
  if (h->nb_img <= NID_SMEM_BINS)
    k_nid_hist_img<true><<<nid_grid(), 256, 0, s>>>(L.lastImage, L.nextImage, L.lastDepth, L.nextDepth, (const unsigned char*)pi, L.pitch_u8, L.pitch_f32,
                                                   (int)pitch_i, L.rows, L.cols, h->nb_img, h->hist_img);
  else
    k_nid_hist_img<false><<<nid_grid(), 256, 0, s>>>(L.lastImage, L.nextImage, L.lastDepth, L.nextDepth, (const unsigned char*)pi, L.pitch_u8, L.pitch_f32,
                                                    (int)pitch_i, L.rows, L.cols, h->nb_img, h->hist_img);
  const float max_mm = max_depth_m * 1000;  // MutualInformation.cpp:206
  if ((int)(max_mm / (float)h->nb_depth) <= 0) return DMS_ERR_INVALID;
  k_nid_hist_depth<<<nid_grid(), 256, 0, s>>>(L.lastDepth, L.nextDepth, (const float*)pd, L.pitch_f32, (int)pitch_d, L.rows, L.cols, h->nb_depth, max_mm,
                                              h->hist_depth);
  const size_t mgw = (nid_marg_bytes(h->nb_img > h->nb_depth ? h->nb_img : h->nb_depth) + 255) / 256 * 256 / 4;
  e = launch_nid_entropy(h->hist_img, h->nb_img, L.rows * L.cols, h->marg, h->d_out, s);
  if (e == cudaSuccess) e = launch_nid_entropy(h->hist_depth, h->nb_depth, L.rows * L.cols, h->marg + mgw, h->d_out + 4, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(h->h_out, h->d_out, 32, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return (int)e;
  ni = h->h_out[0]; nd = h->h_out[4];
  if (nid_img) *nid_img = ni;
  if (nid_depth) *nid_depth = nd;
  if (score) *score = (lambda * nd) + ((1.0f - lambda) * ni);  // EF.cpp:661-662
  return DMS_OK;
}

int dms_nid_get_histogram(dms_nid* h, int which, uint32_t* dst, size_t n_words, void* stream) {
  if (!h || !dst || which < 0 || which > 1) return DMS_ERR_INVALID;
  const int nb = which ? h->nb_depth : h->nb_img;
  if (n_words < (size_t)nb * nb) return DMS_ERR_INVALID;
  cudaSetDevice(h->device);
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemcpyAsync(dst, which ? h->hist_depth : h->hist_img, (size_t)nb * nb * 4, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  return e == cudaSuccess ? DMS_OK : (int)e;
}

}  // extern "C"
