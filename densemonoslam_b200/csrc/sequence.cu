// Sequence runner (include/dmsodom/sequence_c.h): K trackers take turns on the frames of one open-loop sequence, one host
// thread and one stream each.  Host-side ordering between the threads (per frame):
//   issued[i]  initRGB of frame i has returned (its pyramid may be taken by frame i + 1)
//   taken[j]   the image tracker j % K holds from frame j - K has been handed on, frame j may load its own
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <new>
#include <thread>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/dmsodom/sequence_c.h"
#include "../../include/dmsodom/ring_c.h"
#include "seq_flags.h"

// Device-side copy of one frame's inputs (host-buffer runs): a tracker owns two of them and a copy stream, so the upload of
// its NEXT frame (frame i + K) runs under the kernels of the current one and the PCIe link always has a copy queued.
struct Stage {
  float* v4 = nullptr; float* n4 = nullptr; uint8_t* mi = nullptr; uint8_t* li = nullptr; uint16_t* d = nullptr;
  cudaEvent_t ready = nullptr;   // the upload of the frame it holds has completed
  int frame = -1;                // frame it holds or is receiving (-1: nothing of the current run)
};

// One run of the sequence, as seen by the worker threads.
struct Job { const dms_seq_frame* frames; int n; const uint8_t* first_prev; int host; dms_seq_result* out; };

struct dms_sequence {
  int device, width, height;
  dms_seq_options opt;
  std::vector<dms_odom*> trk;
  std::vector<cudaStream_t> streams, copy_streams;
  std::vector<Stage> stages;   // two per tracker, allocated by the first host-buffer run
  int grid, resident;
  dms_ring* ring = nullptr; int ring_sensor = 0;   // results of every tracked frame also go to this ring (dms_sequence_attach_ring)
  dms_seq::Flags issued, taken;
  // worker pool: one host thread per tracker, created once and parked between runs (spawning them per run cost a
  // fixed ~2 ms per call, which dominated short runs and multi-rank launches)
  std::vector<std::thread> workers;
  std::mutex m; std::condition_variable cv_go, cv_done;
  unsigned long long run_id = 0; int pending = 0; bool quit = false;
  Job job{}; std::atomic<int> err{0};
  // host timeline of the last run, microseconds since its entry: first tracking kernel enqueued, last result collected, exit
  std::chrono::steady_clock::time_point t_entry; std::atomic<long long> t_first_launch_us{-1}, t_last_collect_us{0}; double timeline[3] = {0, 0, 0};
};

static void seq_work(dms_sequence* s, int k);
static void seq_worker_main(dms_sequence* s, int k) {
  cudaSetDevice(s->device);
  unsigned long long seen = 0;
  for (;;) {
    {
      std::unique_lock<std::mutex> l(s->m);
      s->cv_go.wait(l, [&] { return s->quit || s->run_id != seen; });
      if (s->quit) return;
      seen = s->run_id;
    }
    seq_work(s, k);
    {
      std::lock_guard<std::mutex> l(s->m);
      if (--s->pending == 0) s->cv_done.notify_all();
    }
  }
}

// Frames k, k + K, k + 2K, ... of the running job on tracker k.
static void seq_work(dms_sequence* s, int k) {
  const int K = (int)s->trk.size();
  const dms_seq_options& o = s->opt;
  const dms_seq_frame* frames = s->job.frames; const int n = s->job.n; const uint8_t* first_prev = s->job.first_prev;
  const int host = s->job.host; dms_seq_result* out = s->job.out;
  std::atomic<int>& err = s->err;
  {
    dms_odom* h = s->trk[k];
    void* st = (void*)s->streams[k];
    const size_t px = (size_t)s->width * s->height;
    cudaStream_t cs = s->copy_streams[k];
    Stage* sg = &s->stages[(size_t)2 * k];
    auto upload = [&](int fi, Stage& g) -> int {  // frame fi -> stage g on the copy stream
      cudaError_t e = cudaSuccess;
      if (!g.v4) {
        if (e == cudaSuccess) e = cudaMalloc(&g.v4, px * 16);
        if (e == cudaSuccess) e = cudaMalloc(&g.n4, px * 16);
        if (e == cudaSuccess) e = cudaMalloc(&g.mi, px * 4);
        if (e == cudaSuccess) e = cudaMalloc(&g.li, px * 4);
        if (e == cudaSuccess) e = cudaMalloc(&g.d, px * 2);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g.ready, cudaEventDisableTiming);
      }
      const dms_seq_frame& f = frames[fi];
      if (e == cudaSuccess) e = cudaMemcpyAsync(g.v4, f.vertex4, px * 16, cudaMemcpyHostToDevice, cs);
      if (e == cudaSuccess) e = cudaMemcpyAsync(g.n4, f.normal4, px * 16, cudaMemcpyHostToDevice, cs);
      if (e == cudaSuccess) e = cudaMemcpyAsync(g.mi, f.model_rgba, px * 4, cudaMemcpyHostToDevice, cs);
      if (e == cudaSuccess) e = cudaMemcpyAsync(g.d, f.depth, px * 2, cudaMemcpyHostToDevice, cs);
      if (e == cudaSuccess) e = cudaMemcpyAsync(g.li, f.live_rgba, px * 4, cudaMemcpyHostToDevice, cs);
      if (e == cudaSuccess) e = cudaEventRecord(g.ready, cs);
      g.frame = fi;
      return e == cudaSuccess ? DMS_OK : (int)e;
    };
    if (host) { sg[0].frame = -1; sg[1].frame = -1; }
    for (int i = k; i < n && !err.load(); i += K) {
      const dms_seq_frame& f = frames[i];
      const Stage* cur = nullptr;
      int rc = DMS_OK;
      const std::chrono::steady_clock::time_point t_frame = std::chrono::steady_clock::now();
      if (host) {
        Stage& c = sg[(i / K) & 1];
        if (c.frame != i) rc = upload(i, c);
        if (!rc) { cudaError_t e = cudaStreamWaitEvent((cudaStream_t)st, c.ready, 0); if (e != cudaSuccess) rc = (int)e; }
        cur = &c;
      }
      const float* fv = host ? cur->v4 : f.vertex4; const float* fn = host ? cur->n4 : f.normal4;
      const uint8_t* fmi = host ? cur->mi : f.model_rgba; const uint8_t* fli = host ? cur->li : f.live_rgba;
      const uint16_t* fd = host ? cur->d : f.depth;
      if (!rc) rc = dms_odom_init_icp_model(h, fv, fn, 0, o.depthCutoff, f.prior, st);
      if (!rc) rc = dms_odom_init_rgb_model(h, fmi, 0, st);
      if (!rc) rc = dms_odom_init_icp_depth(h, fd, 0, o.depthCutoff, st);
      if (!rc) {
        if (K > 1) s->taken.wait((size_t)i, (size_t)k, err);
        rc = dms_odom_init_rgb(h, fli, 0, st);
      }
      s->issued.set((size_t)i, (size_t)(i + 1) % K);        // waited for by the thread of frame i + 1
      if (!rc) {
        if (i == 0 || K == 1) {
          // one tracker keeps its own previous image (RO.cpp:595-601); the first frame loads the one it is given
          if (i == 0) rc = host ? dms_odom_init_first_rgb_host(h, first_prev, 0, st) : dms_odom_init_first_rgb(h, first_prev, 0, st);
        } else {
          s->issued.wait((size_t)i - 1, (size_t)k, err);
          if (!err.load()) rc = dms_odom_init_first_rgb_from(h, s->trk[(i - 1) % K], st);
        }
      }
      s->taken.set((size_t)i - 1 + K, (size_t)(i - 1 + K) % K);   // waited for by the thread of frame i - 1 + K
      float trans[3] = {f.prior[3], f.prior[7], f.prior[11]};
      float rot[9] = {f.prior[0], f.prior[1], f.prior[2], f.prior[4], f.prior[5], f.prior[6], f.prior[8], f.prior[9], f.prior[10]};
      if (!rc) rc = dms_odom_track_async(h, trans, rot, o.rgbOnly, o.icpWeight, o.pyramid, o.fastOdom, o.so3, 0, st);
      // this tracker's next frame goes up while the current one is tracked (queued behind the first uploads of the other
      // trackers, so a short run starts all of them before it prefetches).  The other stage held frame i - K, whose kernels
      // finished before its result was collected.
      if (!rc && s->ring) rc = dms_ring_publish(s->ring, h, s->ring_sensor, i, st);   // behind the tracking kernel, device to device
      if (!rc && host && i + K < n) rc = upload(i + K, sg[((i / K) + 1) & 1]);
      if (!rc) {
        long long expect = -1;
        s->t_first_launch_us.compare_exchange_strong(expect, (long long)std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - s->t_entry).count());
        rc = dms_odom_track_collect(h, trans, rot, st, &out[i].r);
        const long long now = (long long)std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - s->t_entry).count();
        long long prev = s->t_last_collect_us.load();
        while (prev < now && !s->t_last_collect_us.compare_exchange_weak(prev, now)) { }
      }
      if (rc) { err.store(rc); s->issued.wake(); s->taken.wake(); break; }
      float* p = out[i].pose;
      for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) p[r * 4 + c] = rot[r * 3 + c]; p[r * 4 + 3] = trans[r]; }
      p[12] = p[13] = p[14] = 0.f; p[15] = 1.f;
      dms_odom_timing tm;
      dms_odom_get_timing(h, &tm);
      out[i].track_ms = tm.track_ms;
      out[i].host_ms = (float)(std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t_frame).count() * 1e-6);
    }
  }
}

extern "C" {

void dms_sequence_default_options(dms_seq_options* o) {
  if (!o) return;
  o->n_trackers = 12; o->block_threads = 512; o->grid = 0; o->wait_mode = 2;
  o->rgbOnly = 0; o->icpWeight = 10.0f; o->pyramid = 1; o->fastOdom = 0; o->so3 = 1; o->depthCutoff = 25.0f;
  o->timing = 0;
}

int dms_sequence_destroy(dms_sequence* s) {
  if (!s) return DMS_OK;
  { std::lock_guard<std::mutex> l(s->m); s->quit = true; }
  s->cv_go.notify_all();
  for (auto& t : s->workers) if (t.joinable()) t.join();
  cudaSetDevice(s->device);
  for (size_t i = 0; i < s->trk.size(); ++i) if (s->trk[i]) dms_odom_destroy(s->trk[i]);
  for (size_t i = 0; i < s->streams.size(); ++i) if (s->streams[i]) cudaStreamDestroy(s->streams[i]);
  for (size_t i = 0; i < s->copy_streams.size(); ++i) if (s->copy_streams[i]) cudaStreamDestroy(s->copy_streams[i]);
  for (Stage& g : s->stages) {
    if (g.v4) cudaFree(g.v4); if (g.n4) cudaFree(g.n4); if (g.mi) cudaFree(g.mi); if (g.li) cudaFree(g.li); if (g.d) cudaFree(g.d);
    if (g.ready) cudaEventDestroy(g.ready);
  }
  delete s;
  return DMS_OK;
}

int dms_sequence_create(dms_sequence** out, int device, int width, int height, float cx, float cy, float fx, float fy, const dms_seq_options* opt) {
  if (!out) return DMS_ERR_INVALID;
  *out = nullptr;
  dms_seq_options o;
  if (opt) o = *opt; else dms_sequence_default_options(&o);
  if (o.n_trackers < 1 || o.n_trackers > 64) return DMS_ERR_INVALID;
  dms_sequence* s = new (std::nothrow) dms_sequence();
  if (!s) return DMS_ERR_NOMEM;
  s->device = device; s->width = width; s->height = height; s->opt = o; s->grid = 0; s->resident = 1;
  cudaSetDevice(device);
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const float angle = sinf(20.f * 3.14159254f / 180.f);
  for (int k = 0; k < o.n_trackers; ++k) {
    dms_odom* h = nullptr;
    int rc = dms_odom_create(&h, device, width, height, cx, cy, fx, fy, 0.10f, angle);
    if (rc != DMS_OK) { dms_sequence_destroy(s); return rc; }
    s->trk.push_back(h);
    cudaStream_t st = nullptr;
    if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) { dms_sequence_destroy(s); return DMS_ERR_DEVICE; }
    s->streams.push_back(st);
    cudaStream_t cs = nullptr;
    if (cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking) != cudaSuccess) { dms_sequence_destroy(s); return DMS_ERR_DEVICE; }
    s->copy_streams.push_back(cs);
    if (o.block_threads > 0) { rc = dms_odom_set_block(h, o.block_threads); if (rc != DMS_OK) { dms_sequence_destroy(s); return rc; } }
    if (o.n_trackers > 1) {
      const int threads = o.block_threads > 0 ? o.block_threads : 1024;
      const int slots = sms * (1024 / threads > 0 ? 1024 / threads : 1);
      // Kernels of different streams run concurrently only while the streams map to different hardware queues: 8 by default,
      // CUDA_DEVICE_MAX_CONNECTIONS of them when the process sets that variable before CUDA starts (measured: 12 trackers on 12
      // queues +10 % over 8 on 8, but -25 % on 8 queues).  So at most that many tracking kernels are sized to be resident.
      int queues = 8;
      if (const char* q = getenv("CUDA_DEVICE_MAX_CONNECTIONS")) { const int v = atoi(q); if (v >= 1 && v <= 32) queues = v; }
      const int resident_max = o.n_trackers < queues ? o.n_trackers : queues;
      s->grid = o.grid > 0 ? o.grid : slots / resident_max;
      dms_odom_set_grid(h, s->grid);
      s->resident = slots / s->grid < o.n_trackers ? slots / s->grid : o.n_trackers;
      dms_odom_set_wait(h, o.wait_mode);
    } else if (o.grid > 0) {
      dms_odom_set_grid(h, o.grid);
    }
    dms_odom_set_timing(h, o.timing);
  }
  if (o.n_trackers > 1)
    for (int k = 0; k < o.n_trackers; ++k) s->workers.emplace_back(seq_worker_main, s, k);
  s->stages.resize((size_t)2 * o.n_trackers);
  *out = s;
  return DMS_OK;
}

int dms_sequence_resident(const dms_sequence* s) { return s ? s->resident : 0; }

int dms_sequence_run(dms_sequence* s, const dms_seq_frame* frames, int n, const uint8_t* first_prev, int host, dms_seq_result* out) {
  if (!s || !frames || !out || n < 0 || (n > 0 && !first_prev)) return DMS_ERR_INVALID;
  const int K = (int)s->trk.size();
  s->issued.reset((size_t)n + 1, (size_t)K);
  s->taken.reset((size_t)n + K + 1, (size_t)K);
  for (int j = 0; j < K; ++j) s->taken.set((size_t)j, (size_t)j);
  s->err.store(0);
  s->t_entry = std::chrono::steady_clock::now(); s->t_first_launch_us.store(-1); s->t_last_collect_us.store(0);
  s->job.frames = frames; s->job.n = n; s->job.first_prev = first_prev; s->job.host = host; s->job.out = out;
  if (K == 1) {
    cudaSetDevice(s->device);
    seq_work(s, 0);
  } else {
    std::unique_lock<std::mutex> l(s->m);
    s->pending = K; ++s->run_id;
    s->cv_go.notify_all();
    s->cv_done.wait(l, [&] { return s->pending == 0; });
  }
  s->timeline[0] = (double)s->t_first_launch_us.load(); s->timeline[1] = (double)s->t_last_collect_us.load();
  s->timeline[2] = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - s->t_entry).count();
  return s->err.load();
}

int dms_sequence_attach_ring(dms_sequence* s, dms_ring* r, int sensor) {
  if (!s || sensor < 0) return DMS_ERR_INVALID;
  s->ring = r; s->ring_sensor = sensor;
  return DMS_OK;
}

int dms_sequence_host_timeline(const dms_sequence* s, double us[3]) {
  if (!s || !us) return DMS_ERR_INVALID;
  us[0] = s->timeline[0]; us[1] = s->timeline[1]; us[2] = s->timeline[2];
  return DMS_OK;
}

}  // extern "C"
