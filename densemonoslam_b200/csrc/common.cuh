// Shared device helpers for the dmsodom kernels (sm_100a).
//
// Arithmetic contract (DESIGN.md "Arithmetic"): every fp32 operation whose rounding can reach
// an integer output (pixel rounding, threshold compares, u8/s16 stores) is written with an
// explicit-rounding intrinsic so nvcc's -fmad contraction cannot change it; the fused
// multiply-adds are the ones nvcc emits for the reference's expressions.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dms {

// ------------------------------------------------------------------ layout / views
struct Intr { float fx, fy, cx, cy; };

struct Mat33 { float3 r[3]; };

// One pyramid level.  All pitches in BYTES, every row 256-byte aligned.
struct Level {
  int rows, cols;
  Intr intr;
  // planar maps: plane p at base + p*rows*pitch
  float *vmap_curr, *nmap_curr, *vmap_prev, *nmap_prev;
  float *lastDepth, *nextDepth, *nextDepthMasked;
  unsigned char *lastImage, *nextImage, *lastNextImage;
  short *dIdx, *dIdy;
  unsigned short* depth_tmp;
  uint4* corr;  // 16-byte DataTerm {s16 u0, s16 v0, s16 diff, u8 valid, u8 0, f32 z = model depth at (u0, v0), u32 0}
  int pitch_f32, pitch_u8, pitch_s16, pitch_u16, pitch_corr;
};

template <class T>
__device__ __forceinline__ T* row_ptr(T* base, int pitch_bytes, int y) {
  return reinterpret_cast<T*>(reinterpret_cast<char*>(base) + (size_t)y * pitch_bytes);
}
template <class T>
__device__ __forceinline__ const T* row_ptr(const T* base, int pitch_bytes, int y) {
  return reinterpret_cast<const T*>(reinterpret_cast<const char*>(base) + (size_t)y * pitch_bytes);
}

// ------------------------------------------------------------------ explicit-rounding fp32
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float fsqrt(float a) { return __fsqrt_rn(a); }
__device__ __forceinline__ float qnan() { return __int_as_float(0x7fffffff); }

__device__ __forceinline__ float3 sub3(float3 a, float3 b) { return make_float3(fsub(a.x, b.x), fsub(a.y, b.y), fsub(a.z, b.z)); }
__device__ __forceinline__ float3 add3(float3 a, float3 b) { return make_float3(fadd(a.x, b.x), fadd(a.y, b.y), fadd(a.z, b.z)); }
// a.x*b.x + a.y*b.y + a.z*b.z  ->  fma(az,bz, fma(ax,bx, ay*by))   (operators.cuh:72-75 under -fmad)
__device__ __forceinline__ float dot3(float3 a, float3 b) { return ffma(a.z, b.z, ffma(a.x, b.x, fmul(a.y, b.y))); }
__device__ __forceinline__ float3 cross3(float3 a, float3 b) {
  return make_float3(ffma(a.y, b.z, -fmul(a.z, b.y)), ffma(a.z, b.x, -fmul(a.x, b.z)), ffma(a.x, b.y, -fmul(a.y, b.x)));
}
__device__ __forceinline__ float norm3(float3 a) { return fsqrt(dot3(a, a)); }
__device__ __forceinline__ float3 normalized3(float3 a) {
  const float rn = fdiv(1.0f, fsqrt(dot3(a, a)));
  return make_float3(fmul(a.x, rn), fmul(a.y, rn), fmul(a.z, rn));
}
// ------------------------------------------------------------------ legacy numerics (dms_odom_set_numerics)
// The reference is built with --ftz=true --prec-div=false --prec-sqrt=false (Core/src/CMakeLists.txt:48-49): nvcc turns
// its a / b into div.full.ftz.f32, 1.0f / a into rcp.approx.ftz.f32, sqrtf into sqrt.approx.ftz.f32, rsqrtf into
// rsqrt.approx.ftz.f32 and x / 1000.f into a multiplication by 0.001f.  These helpers issue exactly those PTX
// instructions, so a kernel instantiated with LEG = true rounds like the reference's own build at the sites where a
// division or square root feeds a truncation, a pixel rounding or a threshold test (the CPU cannot reproduce the MUFU
// approximations; the GPU can, bit for bit).  LEG = false is the IEEE arithmetic the CPU oracle restates.
__device__ __forceinline__ float div_full_ftz(float a, float b) { float r; asm("div.full.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float rcp_approx_ftz(float a) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ float sqrt_approx_ftz(float a) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ float rsqrt_approx_ftz(float a) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
template <bool LEG> __device__ __forceinline__ float div_t(float a, float b) { return LEG ? div_full_ftz(a, b) : fdiv(a, b); }
template <bool LEG> __device__ __forceinline__ float3 normalized3_t(float3 a) {  // operators.cuh:82-86
  const float rn = LEG ? rsqrt_approx_ftz(dot3(a, a)) : fdiv(1.0f, fsqrt(dot3(a, a)));
  return make_float3(fmul(a.x, rn), fmul(a.y, rn), fmul(a.z, rn));
}

__device__ __forceinline__ float3 mul33(const Mat33& m, float3 a) { return make_float3(dot3(m.r[0], a), dot3(m.r[1], a), dot3(m.r[2], a)); }

// ------------------------------------------------------------------ exact accumulation
// A segment sum p (fp32) is split into trunc(p) and a 2^-44 fixed-point fraction; both are
// accumulated in int64, so the grid-level reduction is associative and may use atomics while
// staying bit-reproducible (and equal to the CPU oracle's).
struct Exact {
  long long hi, lo;
  int bad;
  __device__ __forceinline__ void clear() { hi = 0; lo = 0; bad = 0; }
  // branch-free: a value that sets `bad` also adds a saturated / zero conversion, which no longer matters (the result
  // of a flagged accumulator is NaN)
  __device__ __forceinline__ void addf(float p) {
    bad |= !(fabsf(p) < 4.611686018427387904e18f) ? 1 : 0;
    const float hf = truncf(p);
    hi += __float2ll_rz(hf);
    lo += __float2ll_rz(fmul(fsub(p, hf), 17592186044416.0f));
  }
};
__device__ __forceinline__ float exact_value(long long hi, long long lo, int bad) {
  if (bad) return qnan();
  const double v = __dadd_rn(__ll2double_rn(hi), __dmul_rn(__ll2double_rn(lo), 5.6843418860808015e-14));
  return __double2float_rn(v);
}

// ------------------------------------------------------------------ scoped memory ops
__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// relaxed (no L1 invalidation) load for spin loops: a polling ld.acquire costs a CCTL.IVALL per poll,
// which keeps wiping the L1 of CTAs on the same SM that are still working
__device__ __forceinline__ unsigned int ld_relaxed_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ void st_release_u32(unsigned int* p, unsigned int v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int atom_add_acqrel_u32(unsigned int* p, unsigned int v) {
  unsigned int old;
  asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

}  // namespace dms
