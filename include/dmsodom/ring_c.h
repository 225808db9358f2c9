/* dmsodom result ring -- C ABI.
 *
 * The multi-sensor mode runs one tracker per GPU and exchanges nothing on the data path except the small per-frame result of
 * every sensor, which the reference consumes in its sequential per-sensor loop (GUI/src/MainController.cpp:262-400: pose,
 * lastICPError / lastICPCount gates, covariance from lastA).  This ring is that exchange: a block of fixed-size slots in
 * the memory of ONE GPU (the consumer's); every tracker, in whatever process and on whatever GPU it runs, writes the record
 * of a finished frame straight into its slot with a tiny kernel enqueued behind the tracking kernel (peer stores over
 * NVLink for a remote GPU, through a CUDA IPC mapping for a remote process), sequence number last.  No NCCL, no host hop.
 * No reference counterpart (the reference is one process on one GPU).
 *
 * Slot of (sensor, frame) = sensor * slots_per_sensor + frame % slots_per_sensor; a slot is complete when its `seq` equals
 * frame + 1.  Records are plain stores: the consumer must have read a slot before the producer laps it.
 */
#ifndef DMSODOM_RING_C_H
#define DMSODOM_RING_C_H
#include "odom_c.h"
#include "sequence_c.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dms_ring_record {   /* 512 bytes */
  uint32_t seq;                    /* frame + 1; written last, after a system-scope fence */
  uint32_t sensor, frame, reserved;
  float R[9], t[3];                /* refined pose, row-major rotation */
  float stats[6];                  /* lastICPError, lastICPCount, lastRGBError, lastRGBCount, lastSO3Error, lastSO3Count */
  int32_t iters[3], so3_iters;
  double A[36], b[6];              /* lastA (row-major), lastb: RO.h:76-77 */
  uint8_t pad[512 - 16 - 48 - 24 - 16 - 336];
} dms_ring_record;

#define DMS_RING_HANDLE_BYTES 64
typedef struct dms_ring dms_ring;

/* Consumer side: allocates n_sensors * slots_per_sensor zeroed records on `device`. */
int dms_ring_create(dms_ring** out, int device, int n_sensors, int slots_per_sensor);
/* The CUDA IPC handle of the ring's memory, to be sent to producer processes (any byte transport). */
int dms_ring_export(dms_ring* r, uint8_t handle[DMS_RING_HANDLE_BYTES]);
/* Producer side in ANOTHER process: maps the consumer's ring.  `device` = the GPU the producers of this process run on. */
int dms_ring_open(dms_ring** out, int device, const uint8_t handle[DMS_RING_HANDLE_BYTES], int n_sensors, int slots_per_sensor);
int dms_ring_destroy(dms_ring* r);

/* Enqueues on `stream` (behind the handle's tracking kernel) the write of the handle's last result as record (sensor, frame).
 * Producers in the consumer's own process pass the ring from dms_ring_create (peer access is enabled on first use when the
 * tracker runs on another GPU). */
int dms_ring_publish(dms_ring* r, dms_odom* h, int sensor, int frame, void* stream);
/* Every frame tracked by `s` from now on is published as sensor `sensor` (dms_ring_publish behind each track); r = NULL detaches. */
int dms_sequence_attach_ring(dms_sequence* s, dms_ring* r, int sensor);

/* Consumer: zeroes every slot (synchronises `stream`).  No producer may be publishing. */
int dms_ring_clear(dms_ring* r, void* stream);
/* Consumer: number of complete records among frames [first_frame, first_frame + n_frames) of `sensor` (one small kernel and
 * a 4-byte read-back on `stream`, which is synchronised). */
int dms_ring_poll(dms_ring* r, int sensor, int first_frame, int n_frames, int* n_complete, void* stream);
/* Consumer: copies the record of (sensor, frame) to host memory; returns DMS_ERR_INVALID if it is not complete. */
int dms_ring_read(dms_ring* r, int sensor, int frame, dms_ring_record* h_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif
