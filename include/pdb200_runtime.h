/* C ABI of libpdb200_runtime.so - the host-side native runtime of the paddle_b200 framework:
 * TCP key/value store (rendezvous, handle exchange), communication watchdog, gradient-bucket assignment,
 * flag mirror, symmetric-memory arena (CUDA VMM + NVLS multicast).  Every call returns 0 / a non-negative value
 * on success and a negative value on failure (message via pdb200_rt_last_error()).  Parsed by
 * paddle_b200/lib/__init__.py for the ctypes signatures: one declaration per statement.
 */
#ifndef PDB200_RUNTIME_H_
#define PDB200_RUNTIME_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#define PDB_API

PDB_API const char* pdb200_rt_last_error();

/* TCP store (reference: paddle/phi/core/distributed/store/tcp_store.{h,cc}: MasterDaemon poll() loop with
 * ADD/GET/CHECK/SET/WAIT/STOP, clients with retry connect).  handle = opaque pointer. */
PDB_API void* pdb200_store_create(const char* host, int port, int is_master, int world_size, int timeout_ms);
PDB_API void pdb200_store_destroy(void* store);
PDB_API int pdb200_store_set(void* store, const char* key, const void* value, int64_t nbytes);
PDB_API int64_t pdb200_store_get(void* store, const char* key, void* out, int64_t capacity, int timeout_ms);
PDB_API int64_t pdb200_store_add(void* store, const char* key, int64_t delta);
PDB_API int pdb200_store_check(void* store, const char* key);
PDB_API int pdb200_store_wait(void* store, const char* key, int timeout_ms);
PDB_API int pdb200_store_delete(void* store, const char* key);
PDB_API int pdb200_store_port(void* store);

/* Communication watchdog (reference: paddle/phi/core/distributed/comm_task_manager.cc:65-280,
 * nccl_comm_task.cc:30-268).  A background thread classifies tasks older than the timeout and logs a trace line
 * with group, sequence, op and numel plus the group's last finished task; abort_on_timeout ends the process so
 * the launcher can restart the job. */
PDB_API int pdb200_wd_start(int rank, int timeout_ms, int poll_ms);
PDB_API int pdb200_wd_stop();
PDB_API int pdb200_wd_task_begin(int group, int64_t seq, const char* op, int64_t numel, int sync_op);
PDB_API int pdb200_wd_task_end(int group, int64_t seq);
/* Device-truth variants: begin records a CUDA start event on `stream` (returns 1 when events are in use), end records
 * the end event; the polling thread / pdb200_wd_sweep retire tasks whose end event the GPU has passed, and timeouts are
 * classified "not started" / "started, not finished" from the events (reference: nccl_comm_task.cc:90-140, 238-268). */
PDB_API int pdb200_wd_task_begin_ev(int group, int64_t seq, const char* op, int64_t numel, int sync_op, void* stream);
PDB_API int pdb200_wd_task_end_ev(int group, int64_t seq, void* stream);
PDB_API int pdb200_wd_sweep();
/* 64 bytes of mapped pinned host memory (UVA pointer, 0 without a GPU): word 0 = last barrier epoch passed by the fused
 * NVLink kernels of this process (pass it to pdb200_nvl_set_progress of the kernel library). */
PDB_API uint64_t pdb200_wd_progress_ptr();
PDB_API int pdb200_wd_inflight();
PDB_API int pdb200_wd_timeouts();
PDB_API int pdb200_wd_set_abort(int abort_on_timeout);
PDB_API int64_t pdb200_wd_dump(char* out, int64_t capacity);

/* Native process group over NCCL (reference: fluid/distributed/collective/process_group_nccl.cc:160-1016,
 * phi/core/distributed/nccl_comm_context.cc, comm_context_manager.cc:61-110).  pdb200_pg_create is collective over the
 * group: rank 0 publishes the NCCL unique id in the given pdb200_store under "<prefix>/nccl_id/<gid>".  Every collective
 * returns a task handle (null on failure).  use_calc_stream = 1 runs on `calc_stream` itself; otherwise the group's
 * communication stream runs it after an event recorded on `calc_stream`, and pdb200_pg_task_wait makes a stream wait
 * for the result ON THE DEVICE.  dtype: 0 f32, 1 bf16, 2 f16, 3 f64, 4 i32, 5 i64, 6 u8, 7 i8.  op: 0 sum, 1 max, 2 min,
 * 3 prod, 4 avg.  Tasks are registered with the watchdog (CUDA events); its thread polls ncclCommGetAsyncError and
 * aborts a communicator in error. */
PDB_API int pdb200_pg_available();
PDB_API int pdb200_pg_nccl_version();
PDB_API void* pdb200_pg_create(void* store, const char* prefix, int gid, int rank, int nranks, int timeout_ms);
PDB_API int pdb200_pg_destroy(void* pg);
PDB_API uint64_t pdb200_pg_comm_stream(void* pg);
PDB_API int pdb200_pg_abort(void* pg);
PDB_API int pdb200_pg_check_async_errors();
PDB_API void* pdb200_pg_all_reduce(void* pg, const void* in, void* out, int64_t numel, int dtype, int op, void* calc_stream, int use_calc_stream, int sync_op);
PDB_API void* pdb200_pg_broadcast(void* pg, const void* in, void* out, int64_t numel, int dtype, int root, void* calc_stream, int use_calc_stream, int sync_op);
PDB_API void* pdb200_pg_reduce(void* pg, const void* in, void* out, int64_t numel, int dtype, int op, int root, void* calc_stream, int use_calc_stream, int sync_op);
PDB_API void* pdb200_pg_all_gather(void* pg, const void* in, void* out, int64_t numel_per_rank, int dtype, void* calc_stream, int use_calc_stream, int sync_op);
PDB_API void* pdb200_pg_reduce_scatter(void* pg, const void* in, void* out, int64_t numel_per_rank, int dtype, int op, void* calc_stream, int use_calc_stream, int sync_op);
PDB_API void* pdb200_pg_all_to_all(void* pg, const void* in, void* out, int64_t numel_per_rank, int dtype, void* calc_stream, int use_calc_stream, int sync_op);
PDB_API void* pdb200_pg_send(void* pg, const void* in, int64_t numel, int dtype, int peer, void* calc_stream, int use_calc_stream, int sync_op);
PDB_API void* pdb200_pg_recv(void* pg, void* out, int64_t numel, int dtype, int peer, void* calc_stream, int use_calc_stream, int sync_op);
PDB_API int pdb200_pg_task_wait(void* task, void* calc_stream);
PDB_API int pdb200_pg_task_is_completed(void* task);
PDB_API int pdb200_pg_task_synchronize(void* task);
PDB_API int pdb200_pg_task_free(void* pg, void* task);

/* Gradient bucket assignment (reference: paddle/fluid/distributed/collective/reducer.cc:47-158
 * Eager_AssignGroupBySize).  sizes: bytes per tensor, dtype_ids: small ints, is_sparse: 0/1, limits: bytes;
 * out_group[i] = bucket index of tensor i (buckets ordered by their first tensor).  Returns the bucket count. */
PDB_API int pdb200_assign_group_by_size(const int64_t* sizes, const int* dtype_ids, const int* is_sparse, int n, const int64_t* limits, int n_limits, int* out_group);

/* Flag mirror: Python pushes every FLAGS_* change so native code can read it without the interpreter. */
PDB_API int pdb200_flag_set(const char* name, const char* value);
PDB_API int64_t pdb200_flag_get(const char* name, char* out, int64_t capacity);

/* Symmetric-memory arena (reference has none): every rank of a group allocates one physical segment
 * with cuMemCreate, exports a POSIX file descriptor, passes it to its peers over a Unix-domain socket
 * (SCM_RIGHTS) and maps every peer's segment, so device code holds nranks base pointers; when the fabric
 * supports it an NVLS multicast object is bound to all segments and mapped as one more address.
 * rendezvous_dir: a directory visible to all ranks of the node (socket paths).  Returns an opaque handle. */
PDB_API void* pdb200_symm_create(int rank, int nranks, int device, int64_t bytes, const char* rendezvous_dir, const char* tag, int want_multicast, int timeout_ms);
PDB_API void pdb200_symm_destroy(void* arena);
PDB_API int64_t pdb200_symm_bytes(void* arena);
PDB_API uint64_t pdb200_symm_peer_ptr(void* arena, int peer);
PDB_API uint64_t pdb200_symm_multicast_ptr(void* arena);
PDB_API int pdb200_symm_supported(int device);
/* All-gather on the copy engines: segs = HOST array of nseg int64 triples {src_off, dst_off, shard_bytes}; this rank's
 * shard goes to [dst_off + rank * shard_bytes) of every rank's segment with cudaMemcpyAsync on `stream` (no SM work).
 * Bracket it with the arena's cross-rank barriers like the kernel collectives. */
PDB_API int pdb200_symm_allgather_ce(void* arena, const int64_t* segs, int nseg, void* stream);
/* Broadcast of byte ranges on the copy engines: [src_off, +nbytes) of this rank's segment to [dst_off, +nbytes) of every
 * rank's segment. */
PDB_API int pdb200_symm_bcast_ce(void* arena, const int64_t* segs, int nseg, void* stream);
/* One copy-engine put: [src_off, +nbytes) of this rank's segment -> [dst_off, +nbytes) of rank `peer`'s segment. */
PDB_API int pdb200_symm_put_ce(void* arena, int peer, int64_t src_off, int64_t dst_off, int64_t nbytes, void* stream);
/* Overlapped GEMM + collective plumbing, one native call per collective (flags are 32-bit words in the arena; a flag
 * write is a stream-ordered 32-bit memset into the receiver's mapping, a wait is cuStreamWaitValue32 on a local word).
 * push_signal (K5): nchunks chunks of this rank go to every rank at the same offsets, each followed by its flag.
 * ship_blocks (K6): for every other owner, wait for the local counter of its block, copy the block into slot `rank` of
 * the owner's receive buffer, raise the owner's flag[rank].  wait_flags: the stream waits for all local flags. */
PDB_API int pdb200_symm_streamops_supported();
PDB_API int pdb200_symm_push_signal(void* arena, int64_t src_off, int64_t dst_off, int64_t chunk_bytes, int nchunks, int64_t flag_off, int flag_index0, int64_t value, void* stream);
PDB_API int pdb200_symm_ship_blocks(void* arena, int64_t counters_off, int64_t target, int64_t part_off, int64_t recv_off, int64_t blk_bytes, int64_t flag_off, int64_t value, void* stream);
PDB_API int pdb200_symm_wait_flags(void* arena, int64_t flag_off, int64_t value, int skip, void* stream);
/* All-to-all on the copy engines (transfer half of a reduce-scatter): shard p of this rank's buffer at src_off goes to
 * slot `rank` of rank p's receive buffer at dst_off; a local kernel sums the slots afterwards. */
PDB_API int pdb200_symm_alltoall_ce(void* arena, const int64_t* segs, int nseg, void* stream);

/* Offset heap of a symmetric arena: best fit, coalescing free; offsets depend on the call sequence only, so ranks that
 * allocate and free in the same order agree on every offset.  alloc returns -1 when nothing fits.
 * stat: 0 free bytes, 1 largest free block, 2 live blocks, 3 bytes in use, 4 peak bytes in use, 5 free blocks, 6 capacity. */
PDB_API void* pdb200_heap_create(int64_t capacity, int64_t reserved);
PDB_API void pdb200_heap_destroy(void* heap);
PDB_API int64_t pdb200_heap_alloc(void* heap, int64_t nbytes, int64_t align);
PDB_API int pdb200_heap_free(void* heap, int64_t offset);
PDB_API int64_t pdb200_heap_stat(void* heap, int what);

/* Bookkeeping of the data-parallel reducer: which buckets may be reduced, in bucket order, as gradients arrive from
 * autograd hooks; unused-parameter detection at the end of backward.  Tensor work and collectives stay with the caller. */
PDB_API void* pdb200_reducer_create(const int32_t* bucket_of_param, int n_params, const uint8_t* bucket_sparse, int n_buckets);
PDB_API void pdb200_reducer_destroy(void* reducer);
PDB_API void pdb200_reducer_prepare(void* reducer);
PDB_API int pdb200_reducer_mark_ready(void* reducer, int param, int32_t* out_launch, int cap);
PDB_API int pdb200_reducer_finalize(void* reducer, int tolerate_unused, int32_t* out_missing, int cap_missing, int32_t* out_launch, int cap_launch, int32_t* n_launch);
PDB_API void pdb200_reducer_used_bitmap(void* reducer, int32_t* out);

#ifdef __cplusplus
}
#endif
#endif /* PDB200_RUNTIME_H_ */
