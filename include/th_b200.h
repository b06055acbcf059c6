/*
 * th_b200.h -- C ABI of the B200-native reservation hot path (libth_b200.so).
 *
 * The reference (kivicode/TensorHive-Fixed v1.1.0) is pure Python and has NO plugin / FFI seam for
 * this path: the boundary a maintainer binds is the set of Python methods cited on each entry point
 * below (paths relative to the reference root).  INTEGRATION.md shows the ctypes stubs that rebind
 * those methods to these symbols.
 *
 * Conventions
 *   - every function returns 0 on success or a negative TH_E* code; th_last_error() then holds a
 *     human-readable message for the calling thread.  Nothing throws across this boundary.
 *   - all arrays are contiguous.  Columns are int64 unless the name ends in _u8 / _u32.
 *   - `mem` says where EVERY array pointer of that call lives: TH_MEM_HOST (the library stages them
 *     through stream-ordered device scratch and returns only when the outputs are readable) or
 *     TH_MEM_DEVICE (pointers on the snapshot's device; the call only enqueues work on `stream`).
 *     th_conflict_bitmap reads page-locked (cudaHostAlloc / cudaHostRegister) TH_MEM_HOST columns in place:
 *     the kernel streams them over PCIe itself, no staging copy (TH_NO_ZERO_COPY=1 in the environment
 *     forces the staged path).
 *   - `stream` is a cudaStream_t passed as void*.  NULL means: with TH_MEM_DEVICE the legacy default stream; with
 *     TH_MEM_HOST a private non-blocking stream of the CALLING THREAD, so that service threads making host-array
 *     calls neither serialise against each other nor synchronise with the process's other streams.
 *   - a loaded snapshot is immutable; any number of threads may query it concurrently.  Loading is
 *     not concurrent with queries on the same handle (swap handles instead -- see DESIGN.md).
 *   - time is int64 microseconds since the Unix epoch (naive UTC), the grain of the reference's
 *     '%Y-%m-%dT%H:%M:%S.%fZ' format (tensorhive/utils/DateUtils.py:9).  uuid_id is the dense
 *     dictionary code of `resource_id` in [0, n_uuids).  owner is the `user_id` foreign key.
 */
#ifndef TH_B200_H
#define TH_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TH_ABI_VERSION 1

#define TH_MEM_HOST 0
#define TH_MEM_DEVICE 1

/* Interval predicates (tensorhive/models/Reservation.py) */
#define TH_MODE_STRICT 0    /* would_interfere :120-130      _end > qs AND _start < qe, id != self           */
#define TH_MODE_INCLUSIVE 1 /* filter_by_uuids_and_time_range :132-148 / current_events :89-104 (qs = qe = now) */
#define TH_MODE_UPCOMING 2  /* upcoming_events_for_resource :106-118  qs = now, qe = now + period             */

/* th_upcoming_owner_flags tests */
#define TH_OWNER_ANY 0   /* any upcoming reservation             (interferes_with_reservations, allow_own=False) */
#define TH_OWNER_OTHER 1 /* any upcoming reservation of another user (allow_own=True, `r.user is not job.user`)  */
#define TH_OWNER_SAME 2  /* any upcoming reservation of this user (GreedyScheduler, `r.user == job.user`)        */

#define TH_OK 0
#define TH_EINVAL -1   /* bad argument (null pointer, negative size, uuid out of range, bad mode)   */
#define TH_ECUDA -2    /* a CUDA runtime call failed; message carries cudaGetErrorString            */
#define TH_ENOMEM -3   /* device allocation failed                                                   */
#define TH_ESTATE -4   /* snapshot not loaded / wrong device                                         */
#define TH_ECAPACITY -5 /* caller buffer too small; *out_needed holds the required element count     */
#define TH_EINTERNAL -6 /* internal invariant broken (e.g. radix look-back watchdog)                 */

typedef struct th_snapshot th_snapshot;

/* Diagnostics. */
int th_version(void);
const char *th_last_error(void);

/* Diagnostics (no counterpart in the reference): bytes the library's stream-ordered staging pool on `device` has
 * reserved from the driver, bytes of it in use right now, and the high-water mark of the latter since it was last
 * reset (reset_high != 0 resets it after reading).  Any pointer may be NULL. */
int th_pool_stats(int device, int64_t *reserved_bytes, int64_t *used_bytes, int64_t *used_high_bytes, int reset_high);

/* Handle lifecycle.  The handle owns the device copy of the `reservations` table
 * (class Reservation columns, tensorhive/models/Reservation.py:14-37). */
int th_snapshot_create(int device, th_snapshot **out);
int th_snapshot_destroy(th_snapshot *h);

/* Upload the table and index it: rows are radix-sorted by (uuid_id, start_us, id) [K1], per-uuid
 * segment offsets and running maxima of `end` are built.  Replaces what SQLite does on every call in
 * the reference -- a full scan of `reservations` (no index exists; tensorhive/database.py:19-23,
 * migrations ce624ab2c458:35-46) -- and the ORDER BY _start of Reservation.py:117.
 * cancelled_u8[i] = 1 iff `is_cancelled is not None and is_cancelled` (Reservation.py:81-83).
 * ids must be unique and >= 0 (the primary key).  A load that is refused (bad uuid range, negative or duplicate ids,
 * INT64_MIN start, grid too large) leaves a previously loaded snapshot untouched and queryable. */
int th_snapshot_load(th_snapshot *h, int64_t n, int64_t n_uuids, const int64_t *id, const int64_t *uuid_id,
                     const int64_t *start_us, const int64_t *end_us, const int64_t *owner,
                     const uint8_t *cancelled_u8, int mem, void *stream);

/* Snapshot maintenance.  The reference commits every change through CRUDModel.save /
 * destroy (tensorhive/models/CRUDModel.py:21-50); instead of re-sending the whole table after each commit, send the
 * delta: `dst` becomes `src` without the rows whose id is upserted or deleted, plus the upserted rows (insert, update
 * and cancel are all upserts), re-indexed on the device.  Only the delta crosses PCIe.  dst may be src (in place; then
 * no query may run on it meanwhile) or a second handle (readers keep using src until the caller swaps).  n_uuids may
 * grow but not shrink.  Every id may appear at most once in the delta (upserts and deletes together): collapse a
 * batch first (last change wins); TH_EINVAL otherwise, and `dst` is left as it was. */
int th_snapshot_apply_delta(th_snapshot *dst, const th_snapshot *src, int64_t n_uuids, int64_t n_upsert, const int64_t *id,
                            const int64_t *uuid_id, const int64_t *start_us, const int64_t *end_us, const int64_t *owner,
                            const uint8_t *cancelled_u8, int64_t n_delete, const int64_t *delete_ids, int mem,
                            void *stream);

/* Size / bookkeeping of the loaded table. radix_passes = digit passes the last load ran. */
int th_snapshot_info(const th_snapshot *h, int64_t *n, int64_t *n_uuids, int64_t *device_bytes,
                     int64_t *radix_passes);

/* Copy the indexed table back out in snapshot order (tests / debugging).  seg_offsets has
 * n_uuids + 1 entries; any pointer may be NULL to skip that column. */
int th_snapshot_export(const th_snapshot *h, int64_t *id, int64_t *uuid_id, int64_t *start_us, int64_t *end_us,
                       int64_t *owner, uint8_t *cancelled_u8, int64_t *seg_offsets, int mem, void *stream);

/* Batched Reservation.would_interfere (tensorhive/models/Reservation.py:120-130; callers
 * CRUDModel.save -> check_assertions :42-55 and ReservationVerifier.py:103).
 * Query i = (q_uuid[i], q_start[i], q_end[i], q_self_id[i]); q_self_id < 0 means "not saved yet"
 * (SQL `id IS NOT NULL`).  Bit i of out_bits_u32 (word i/32, bit i%32) = the bool the method returns.
 * TH_MODE_INCLUSIVE answers "does filter_by_uuids_and_time_range return anything for this uuid"
 * (cancelled rows count, no self exclusion).  out_bits_u32 has (m + 31) / 32 words, fully written.
 * A q_uuid outside [0, n_uuids) is a resource with no reservations: bit 0. */
int th_conflict_bitmap(const th_snapshot *h, int64_t m, const int64_t *q_uuid, const int64_t *q_start,
                       const int64_t *q_end, const int64_t *q_self_id, int mode, uint32_t *out_bits_u32, int mem,
                       void *stream);

/* th_conflict_bitmap over a narrower wire format (same reference method, Reservation.py:120-130): the host path is bound
 * by the bytes that cross PCIe, so uuid codes and self ids travel as uint32 columns.  q_self_id_u32[i] = 0xFFFFFFFF means
 * "not saved yet"; the column may be NULL when no candidate is a self-update (20 bytes per query, else 24, against 32).
 * Reservation ids must be below 2^32 - 1.  Bitmaps are identical to th_conflict_bitmap's. */
int th_conflict_bitmap_packed(const th_snapshot *h, int64_t m, const uint32_t *q_uuid_u32, const int64_t *q_start,
                              const int64_t *q_end, const uint32_t *q_self_id_u32, int mode, uint32_t *out_bits_u32,
                              int mem, void *stream);

/* The row lists behind the verdicts, CSR: out_offsets[m + 1], then ids and their cancelled flags.
 *   TH_MODE_STRICT     rows of Reservation.py:121-129 `.all()`  (cancelled rows INCLUDED, ascending id)
 *   TH_MODE_INCLUSIVE  rows of filter_by_uuids_and_time_range :148 / current_events :102 before its
 *                      Python filter (cancelled INCLUDED, ascending id)
 *   TH_MODE_UPCOMING   upcoming_events_for_resource :109-118 (cancelled EXCLUDED, ORDER BY start, id)
 * If the total exceeds `capacity`, nothing is written to out_ids/out_cancelled_u8, *out_needed is the
 * required capacity and TH_ECAPACITY is returned (out_offsets is valid).  out_needed is a HOST pointer
 * in both mem modes; the call synchronises `stream` (once for device arrays whose lists are short -- the fill is
 * launched under a device-side guard before the host has seen the total --, twice otherwise). */
int th_conflict_sets(const th_snapshot *h, int64_t m, const int64_t *q_uuid, const int64_t *q_start,
                     const int64_t *q_end, const int64_t *q_self_id, int mode, int64_t *out_offsets,
                     int64_t *out_ids, uint8_t *out_cancelled_u8, int64_t capacity, int64_t *out_needed, int mem,
                     void *stream);

/* Per uuid, `Reservation.current_events(uuid)[0]` as ProtectionService uses it
 * (tensorhive/core/services/ProtectionService.py:90-93): the lowest-id non-cancelled row with
 * start <= now <= end.  out_res_id / out_owner have n_uuids entries, -1 where there is none. */
int th_current_owner(const th_snapshot *h, int64_t now_us, int64_t *out_res_id, int64_t *out_owner, int mem,
                     void *stream);

/* Per requested uuid, the head and length of upcoming_events_for_resource(uuid, period)
 * (Reservation.py:106-118) as JobSchedulingService.check_current_gpu_slots reads it
 * (tensorhive/core/services/JobSchedulingService.py:94-102): first row in (start, id) order and the
 * number of rows; -1/-1/-1/0 when empty or when the uuid is outside [0, n_uuids).  period_us >= 0; a horizon
 * now_us + period_us beyond INT64_MAX is clamped there (no upper bound), here and in th_upcoming_owner_flags. */
int th_upcoming(const th_snapshot *h, int64_t now_us, int64_t period_us, int64_t k, const int64_t *uuids,
                int64_t *out_first_id, int64_t *out_first_start, int64_t *out_first_owner, int64_t *out_count,
                int mem, void *stream);

/* Per job task (uuid, owner): does upcoming_events_for_resource(uuid, period) hold a reservation that
 * passes `test` (TH_OWNER_*)?  JobSchedulingService.interferes_with_reservations :117-132 and
 * GreedyScheduler.schedule_jobs (tensorhive/core/scheduling.py:48-53).  task_uuid < 0 = task without an
 * assigned GPU (get_assigned_gpu_uid returned None): flag 0. */
int th_upcoming_owner_flags(const th_snapshot *h, int64_t now_us, int64_t period_us, int64_t t,
                            const int64_t *task_uuid, const int64_t *task_owner, int test, uint8_t *out_flag_u8,
                            int mem, void *stream);

/* ProtectionService.do_run join (tensorhive/core/services/ProtectionService.py:84-102): per reported GPU
 * process (proc_uuid, proc_owner) -> out_flag_u8 = 1 if it is a violation, out_res_id = id of the
 * violated reservation or -1 (strict mode, no reservation).  A current reservation whose owner is < 0
 * (`reservation.user is None`, :94-95) silences the GPU. */
int th_violation_join(const th_snapshot *h, int64_t now_us, int strict, int64_t p, const int64_t *proc_uuid,
                      const int64_t *proc_owner, uint8_t *out_flag_u8, int64_t *out_res_id, int mem, void *stream);

/* No counterpart in the reference.  Same join as th_violation_join for callers that must not allocate device
 * memory (a protection tick on a GPU whose memory a reserved job has filled; a process with a fixed device
 * footprint): the per-uuid table between the two kernels lives in a fixed 1 MiB array inside the library, so
 * the snapshot may have at most TH_STATIC_JOIN_MAX_UUIDS uuids (TH_ECAPACITY beyond that).  All pointers are
 * DEVICE pointers on the snapshot's device; work is only enqueued on `stream`.  The first call on a device
 * resolves the array's address and should be made outside stream capture.  Every call on a device shares the
 * array: calls whose kernels could overlap on the GPU must be serialised by the caller. */
#define TH_STATIC_JOIN_MAX_UUIDS 65536
int th_violation_join_static(const th_snapshot *h, int64_t now_us, int strict, int64_t p, const int64_t *proc_uuid,
                             const int64_t *proc_owner, uint8_t *out_flag_u8, int64_t *out_res_id, void *stream);

/* Usage averaging (tensorhive/core/services/UsageLoggingService.py:25-30,85-105,196-229): samples are
 * grouped by uuid (uuid_offsets[n_uuids + 1]) and ascending in ts within a uuid; window j sums the
 * samples of win_uuid[j] with win_start[j] <= ts <= win_end[j].  out_avg_* = sum // count, or -1 when
 * count == 0 (the reference's `avg`).  Any of the six outputs may be NULL.  `device` is used for
 * TH_MEM_HOST staging; with TH_MEM_DEVICE the pointers' device must be current. */
int th_usage_reduce(int device, int64_t s, int64_t n_uuids, const int64_t *uuid_offsets, const int64_t *ts_us,
                    const int64_t *gpu_util, const int64_t *mem_util, int64_t w, const int64_t *win_uuid,
                    const int64_t *win_start, const int64_t *win_end, int64_t *out_sum_gpu, int64_t *out_sum_mem,
                    int64_t *out_count, int64_t *out_avg_gpu, int64_t *out_avg_mem, int mem, void *stream);

/* The same reduction over the narrow log format: utilisation percentages stored as one
 * byte each (lossless for 0..255; the reference's samples are ints 0..100, NvidiaSmiParser.py:84-85), replacing the
 * per-reservation JSON logs of UsageLoggingService.py:62-116.  2 bytes instead of 16 are read per sample. */
int th_usage_reduce_u8(int device, int64_t s, int64_t n_uuids, const int64_t *uuid_offsets, const int64_t *ts_us,
                       const uint8_t *gpu_util_u8, const uint8_t *mem_util_u8, int64_t w, const int64_t *win_uuid,
                       const int64_t *win_start, const int64_t *win_end, int64_t *out_sum_gpu, int64_t *out_sum_mem,
                       int64_t *out_count, int64_t *out_avg_gpu, int64_t *out_avg_mem, int mem, void *stream);

/* Device-resident usage log, replacing the per-reservation JSON files that the usage logger
 * re-reads and re-writes on every tick (tensorhive/core/services/UsageLoggingService.py:85-116, 159-171).  One region per
 * GPU uuid inside device columns (ts int64, gpu_util / mem_util uint8).  th_ring_append takes one monitoring tick (or any
 * batch) as HOST arrays -- a sample is appended by the caller only when both metrics are present (:92-99); timestamps of
 * a uuid must not decrease; uuid codes beyond the current count extend the ring -- and costs one small H2D copy plus one
 * scatter kernel whatever the log's length (a full region doubles first).  th_ring_reduce is th_usage_reduce_u8 over the
 * ring in place (handle_expired_logs :196-229 for a list of windows): nothing is re-uploaded but a 32-byte entry per uuid
 * after appends.  While a uuid's log is an exact arithmetic run (ts[k] = t0 + k * cadence) its timestamps are treated as
 * delta-encoded: both window ends follow from divisions and the ts column is not read.  Calls on one ring serialise on
 * an internal lock.  th_ring_info: any pointer may be NULL. */
typedef struct th_ring th_ring;
int th_ring_create(int device, int64_t samples_per_uuid_hint, th_ring **out);
int th_ring_destroy(th_ring *r);
int th_ring_append(th_ring *r, int64_t k, const int64_t *uuid, const int64_t *ts_us, const uint8_t *gpu_util_u8,
                   const uint8_t *mem_util_u8, void *stream);
int th_ring_reduce(th_ring *r, int64_t w, const int64_t *win_uuid, const int64_t *win_start, const int64_t *win_end,
                   int64_t *out_sum_gpu, int64_t *out_sum_mem, int64_t *out_count, int64_t *out_avg_gpu,
                   int64_t *out_avg_mem, int mem, void *stream);
int th_ring_info(th_ring *r, int64_t *n_uuids, int64_t *samples, int64_t *capacity, int64_t *device_bytes,
                 int64_t *regular_uuids, int64_t *relayouts);

/* Multi-GPU exchange (no counterpart in the reference, which is single-process CPU code).
 * One process per GPU; rank r owns a contiguous uuid range.  th_peer_create allocates this rank's slot buffer (two
 * generations x world x words_per_rank 8-byte {word, epoch} slots + barrier flags) and returns its 64-byte CUDA IPC
 * handle; after the handles have been exchanged (rank order, world x 64 bytes) th_peer_connect maps every peer's
 * buffer.  th_conflict_bitmap_exchange is K2 FUSED with the all-gather: the kernel stores each bitmap word, tagged
 * with the step's epoch, into every rank's slots over NVLink, and an unpack kernel on `stream` polls the slots and
 * writes *out_gathered, this rank's compact [world][words_per_rank] words (valid until the step after next).
 * m_per_rank (HOST array, world entries, or NULL = every rank has m queries) says how much each rank contributes.
 * Query arrays are DEVICE pointers, or page-locked host memory (all four columns), which the kernel then streams
 * over PCIe itself.  The gathered words are valid only once th_peer_check has returned TH_OK: it synchronises
 * `stream` and reports a missed arrival (a peer more than TH_PEER_TIMEOUT_MS behind, default 2000; the words of that
 * step then read as 0) as TH_EINTERNAL, once.  With wait = 0 the call only enqueues the kernel; th_peer_wait enqueues
 * the receive half later.
 * th_peer_connect(local_ptrs != NULL) wires buffers of the SAME process together (several ranks emulated on one
 * GPU in tests); th_peer_local_base returns the buffer address for that purpose. */
typedef struct th_peer th_peer;
int th_peer_create(int device, int world, int rank, int64_t words_per_rank, th_peer **out, void *handle_out_64);
int th_peer_connect(th_peer *p, const void *all_handles, void *const *local_ptrs);
int th_peer_local_base(th_peer *p, void **out);
int th_peer_destroy(th_peer *p);
int th_peer_check(th_peer *p, void *stream);
int th_conflict_bitmap_exchange(const th_snapshot *h, th_peer *peer, int64_t m, const int64_t *q_uuid,
                                const int64_t *q_start, const int64_t *q_end, const int64_t *q_self_id, int mode,
                                const int64_t *m_per_rank, int wait, uint32_t **out_gathered, void *stream);
/* th_conflict_bitmap_exchange from the packed wire format of th_conflict_bitmap_packed. */
int th_conflict_bitmap_exchange_packed(const th_snapshot *h, th_peer *peer, int64_t m, const uint32_t *q_uuid_u32,
                                       const int64_t *q_start, const int64_t *q_end, const uint32_t *q_self_id_u32,
                                       int mode, const int64_t *m_per_rank, int wait, uint32_t **out_gathered,
                                       void *stream);
int th_peer_wait(th_peer *p, void *stream);
/* Device-side rendezvous of all ranks on `stream` (no host synchronisation). */
int th_peer_barrier(th_peer *p, void *stream);

/* Stand-alone K1: stable LSD radix sort of (k_hi, k_lo) pairs carrying their original index.
 * Exposed for tests and for sorting windows / queries on the device; all pointers are device memory.
 * out_perm[i] = input position of the i-th smallest key. */
int th_sort_pairs(int device, int64_t n, const int64_t *k_hi, const int64_t *k_lo, int64_t *out_perm, int mem,
                  void *stream);

/* Synthetic workloads generated on the device (bench / tests only). */
int th_synth_samples(int device, int64_t n_uuids, int64_t samples_per_uuid, int64_t t0_us, uint64_t seed,
                     int64_t *uuid_offsets, int64_t *ts_us, int64_t *gpu_util, int64_t *mem_util, void *stream);
int th_synth_windows(int device, int64_t n_uuids, int64_t samples_per_uuid, int64_t t0_us, uint64_t seed, int64_t w,
                     int64_t *win_uuid, int64_t *win_start, int64_t *win_end, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* TH_B200_H */
