// Force-included (-include) in front of the reference's own Cuda/reduce.cu and Cuda/cudafuncs.cu so they compile
// unmodified, where they lie under /root/reference, with CUDA 12.9 for sm_100a.  Test infrastructure only.
//
//  1. reduce.cu:88-122,174-187,689-690 call the pre-Volta __shfl_down(val, offset), which CUDA 12 no longer declares
//     for sm_70+: provide it on top of __shfl_down_sync with a full mask (the reference always shuffles full warps).
//  2. cudafuncs.cu:641-669 reads the colour image through a legacy texture *reference* (removed in CUDA 12).  A
//     stand-in `texture` type holding a pitched device pointer, with tex2D / cudaBindTextureToArray /
//     cudaUnbindTexture overloads, keeps the reference source unchanged; the "cudaArray*" the driver passes is a
//     pointer to a RefImage descriptor.
#pragma once
#include <cuda_runtime.h>
#include <vector_types.h>
#include <vector_functions.h>
#include <cstddef>

// (in nvcc's host pass __CUDA_ARCH__ is undefined, "< 300" holds and reduce.cu:57-77 declares its own fallback)
#if defined(__CUDA_ARCH__) && __CUDA_ARCH__ >= 300
__device__ __forceinline__ float __shfl_down(float v, int offset) { return __shfl_down_sync(0xffffffffu, v, offset); }
__device__ __forceinline__ int __shfl_down(int v, int offset) { return __shfl_down_sync(0xffffffffu, v, offset); }
#endif

struct RefImage { const void* ptr; size_t pitch; };  // what the driver hands over as "cudaArray*"

template <class T, int D, int M>
struct ref_texture { const T* ptr; size_t pitch; };

template <class T, int D, int M>
__device__ __forceinline__ T tex2D(const ref_texture<T, D, M>& t, int x, int y) {
  return *reinterpret_cast<const T*>(reinterpret_cast<const char*>(t.ptr) + (size_t)y * t.pitch + (size_t)x * sizeof(T));
}
template <class T, int D, int M>
static inline cudaError_t cudaBindTextureToArray(ref_texture<T, D, M>& tex, cudaArray* arr) {
  const RefImage* img = reinterpret_cast<const RefImage*>(arr);
  ref_texture<T, D, M> h;
  h.ptr = static_cast<const T*>(img->ptr); h.pitch = img->pitch;
  return cudaMemcpyToSymbol(tex, &h, sizeof h);
}
template <class T, int D, int M>
static inline cudaError_t cudaUnbindTexture(ref_texture<T, D, M>&) { return cudaSuccess; }

#define texture __device__ ref_texture
