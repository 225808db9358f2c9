// Result ring (include/dmsodom/ring_c.h): per-frame tracker results written device-to-device into one GPU's memory.
#include <cuda_runtime.h>
#include <string.h>
#include <new>
#include "kernels.h"
#include "../../include/dmsodom/ring_c.h"

using namespace dms;

struct dms_ring {
  int device;            // GPU whose memory holds the slots (consumer) or, for an opened ring, the producers' GPU
  int n_sensors, slots;
  dms_ring_record* base; // device pointer valid in this process
  bool owner;            // allocated here (consumer) vs. mapped through IPC (producer process)
  int* d_count;          // consumer: poll result
  int* h_count;          // pinned
  bool peer_enabled[64];
};

namespace {

static_assert(sizeof(dms_ring_record) == 512, "a record is 512 bytes");

// One block.  The tracking kernel has finished (stream order), so its control block holds the final result.
__global__ void k_ring_publish(const Ctrl* __restrict__ c, dms_ring_record* slot, unsigned int sensor, unsigned int frame) {
  const int t = threadIdx.x;
  if (t < 9) slot->R[t] = c->out_R[t];
  else if (t < 12) slot->t[t - 9] = c->out_t[t - 9];
  else if (t < 18) slot->stats[t - 12] = (&c->res.lastICPError)[t - 12];
  else if (t < 21) slot->iters[t - 18] = c->res.iters[t - 18];
  else if (t == 21) slot->so3_iters = c->res.so3_iters;
  else if (t == 22) { slot->sensor = sensor; slot->frame = frame; slot->reserved = 0u; }
  if (t >= 32 && t < 68) slot->A[t - 32] = c->res.lastA[t - 32];
  else if (t >= 68 && t < 74) slot->b[t - 68] = c->res.lastb[t - 68];
  __threadfence_system();   // the record before its sequence number, as seen from any GPU or the host
  __syncthreads();
  if (t == 0) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(&slot->seq), "r"(frame + 1u) : "memory");
}

__global__ void k_ring_count(const dms_ring_record* lane, int slots, unsigned int sensor, int first, int n, int* out) {
  __shared__ int s;
  if (threadIdx.x == 0) s = 0;
  __syncthreads();
  int mine = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const unsigned int f = (unsigned int)(first + i);
    const dms_ring_record* r = lane + (f % (unsigned int)slots);
    unsigned int seq;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seq) : "l"(&r->seq) : "memory");
    if (seq == f + 1u && r->frame == f && r->sensor == sensor) ++mine;
  }
  atomicAdd(&s, mine);
  __syncthreads();
  if (threadIdx.x == 0) *out = s;
}

int make(dms_ring** out, int device, int n_sensors, int slots, bool owner) {
  if (!out || n_sensors < 1 || slots < 1 || (long long)n_sensors * slots > (1 << 22)) return DMS_ERR_INVALID;
  *out = nullptr;
  dms_ring* r = new (std::nothrow) dms_ring();
  if (!r) return DMS_ERR_NOMEM;
  memset(r, 0, sizeof *r);
  r->device = device; r->n_sensors = n_sensors; r->slots = slots; r->owner = owner;
  *out = r;
  return DMS_OK;
}

}  // namespace

extern "C" {

int dms_ring_create(dms_ring** out, int device, int n_sensors, int slots_per_sensor) {
  int rc = make(out, device, n_sensors, slots_per_sensor, true);
  if (rc) return rc;
  dms_ring* r = *out;
  const size_t bytes = sizeof(dms_ring_record) * (size_t)n_sensors * slots_per_sensor;
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaMalloc(&r->base, bytes);
  if (e == cudaSuccess) e = cudaMemset(r->base, 0, bytes);
  if (e == cudaSuccess) e = cudaMalloc(&r->d_count, sizeof(int));
  if (e == cudaSuccess) e = cudaMallocHost(&r->h_count, sizeof(int));
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { dms_ring_destroy(r); *out = nullptr; return (int)e; }
  return DMS_OK;
}

int dms_ring_export(dms_ring* r, uint8_t handle[DMS_RING_HANDLE_BYTES]) {
  if (!r || !r->owner || !handle) return DMS_ERR_INVALID;
  static_assert(sizeof(cudaIpcMemHandle_t) == DMS_RING_HANDLE_BYTES, "handle size");
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaSetDevice(r->device);
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, r->base);
  if (e != cudaSuccess) return (int)e;
  memcpy(handle, &h, sizeof h);
  return DMS_OK;
}

int dms_ring_open(dms_ring** out, int device, const uint8_t handle[DMS_RING_HANDLE_BYTES], int n_sensors, int slots_per_sensor) {
  if (!handle) return DMS_ERR_INVALID;
  int rc = make(out, device, n_sensors, slots_per_sensor, false);
  if (rc) return rc;
  dms_ring* r = *out;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  cudaError_t e = cudaSetDevice(device);
  void* p = nullptr;
  if (e == cudaSuccess) e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) { delete r; *out = nullptr; return (int)e; }
  r->base = static_cast<dms_ring_record*>(p);
  return DMS_OK;
}

int dms_ring_destroy(dms_ring* r) {
  if (!r) return DMS_OK;
  cudaSetDevice(r->device);
  cudaDeviceSynchronize();
  if (r->owner) { if (r->base) cudaFree(r->base); }
  else if (r->base) cudaIpcCloseMemHandle(r->base);
  if (r->d_count) cudaFree(r->d_count);
  if (r->h_count) cudaFreeHost(r->h_count);
  delete r;
  return DMS_OK;
}

int dms_ring_publish(dms_ring* r, dms_odom* h, int sensor, int frame, void* stream) {
  if (!r || !h || sensor < 0 || sensor >= r->n_sensors || frame < 0) return DMS_ERR_INVALID;
  const int dev = odom_device(h);
  cudaError_t e = cudaSetDevice(dev);
  if (e != cudaSuccess) return (int)e;
  if (r->owner && dev != r->device && dev >= 0 && dev < 64 && !r->peer_enabled[dev]) {  // same process, other GPU: direct peer stores
    e = cudaDeviceEnablePeerAccess(r->device, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return (int)e;
    (void)cudaGetLastError();
    r->peer_enabled[dev] = true;
  }
  dms_ring_record* slot = r->base + (size_t)sensor * r->slots + (size_t)(frame % r->slots);
  k_ring_publish<<<1, 96, 0, (cudaStream_t)stream>>>(odom_ctrl_device(h), slot, (unsigned int)sensor, (unsigned int)frame);
  e = cudaGetLastError();
  return e == cudaSuccess ? DMS_OK : (int)e;
}

int dms_ring_clear(dms_ring* r, void* stream) {
  if (!r || !r->owner) return DMS_ERR_INVALID;
  cudaError_t e = cudaSetDevice(r->device);
  if (e == cudaSuccess) e = cudaMemsetAsync(r->base, 0, sizeof(dms_ring_record) * (size_t)r->n_sensors * r->slots, (cudaStream_t)stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)stream);
  return e == cudaSuccess ? DMS_OK : (int)e;
}

int dms_ring_poll(dms_ring* r, int sensor, int first_frame, int n_frames, int* n_complete, void* stream) {
  if (!r || !r->owner || !n_complete || sensor < 0 || sensor >= r->n_sensors || first_frame < 0 || n_frames < 0) return DMS_ERR_INVALID;
  if (n_frames > r->slots) { first_frame += n_frames - r->slots; n_frames = r->slots; }  // older frames have been lapped
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaSetDevice(r->device);
  if (e != cudaSuccess) return (int)e;
  k_ring_count<<<1, 256, 0, s>>>(r->base + (size_t)sensor * r->slots, r->slots, (unsigned int)sensor, first_frame, n_frames, r->d_count);
  e = cudaMemcpyAsync(r->h_count, r->d_count, sizeof(int), cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return (int)e;
  *n_complete = *r->h_count;
  return DMS_OK;
}

int dms_ring_read(dms_ring* r, int sensor, int frame, dms_ring_record* h_out, void* stream) {
  if (!r || !r->owner || !h_out || sensor < 0 || sensor >= r->n_sensors || frame < 0) return DMS_ERR_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaSetDevice(r->device);
  if (e == cudaSuccess) e = cudaMemcpyAsync(h_out, r->base + (size_t)sensor * r->slots + (size_t)(frame % r->slots), sizeof(dms_ring_record), cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return (int)e;
  return (h_out->seq == (uint32_t)frame + 1u && h_out->frame == (uint32_t)frame && h_out->sensor == (uint32_t)sensor) ? DMS_OK : DMS_ERR_INVALID;
}

}  // extern "C"
