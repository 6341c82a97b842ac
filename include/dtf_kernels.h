/* C ABI of libdtf_kernels.so -- the hand-written sm_100a kernels of distributed_tensorflow_b200.
 *
 * Every function returns 0 on success and a negative value on failure; dtf_last_error() then
 * holds a description (thread local).  Pointers are device pointers unless stated otherwise;
 * `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  All launches are
 * asynchronous with respect to the host.
 *
 * The reference (yaokeepmoving/distributed_tensorflow) has no native code of its own; each entry
 * cites the reference call site whose TF-runtime kernel it stands in for.
 */
#ifndef DTF_KERNELS_H_
#define DTF_KERNELS_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DTF_KERNELS_ABI_VERSION 2

/* ---- bookkeeping ---------------------------------------------------------------------------- */
const char* dtf_last_error(void);
long long dtf_launch_count(void);      /* kernels launched by this library since the last reset */
void dtf_reset_launch_count(void);
int dtf_kernels_abi_version(void);
const char* dtf_kernels_source_hash(void);   /* hash of the sources this library was built from */

/* ---- epilogues of dtf_gemm_bf16 -------------------------------------------------------------- */
enum {
  DTF_EPI_BF16 = 0,           /* D = bf16(acc)                                  */
  DTF_EPI_BIAS_BF16 = 1,      /* D = bf16(acc + bias[n])                        */
  DTF_EPI_BIAS_RELU_BF16 = 2, /* D = bf16(max(acc + bias[n], 0))                */
  DTF_EPI_DRELU_BF16 = 3,     /* D = bf16(mask bit(m,n) ? acc : 0)              */
  DTF_EPI_F32 = 4,            /* D = acc (fp32)                                 */
  DTF_EPI_F32_ATOMIC = 5      /* D += acc (fp32 red.global.add; for split-K)    */
};

/* D[M,N] = epilogue(A * B^T), bf16 operands, fp32 accumulation on tcgen05 tensor cores.
 *   a_mn == 0: A is stored [M,K] (row stride lda elements);  a_mn == 1: A is stored [K,M].
 *   b_mn == 0: B is stored [N,K] (row stride ldb elements);  b_mn == 1: B is stored [K,N].
 * Supported layouts: (0,0) forward, (0,1) data gradient, (1,1) weight gradient.
 * Row strides must be multiples of 8 elements and bases 16-byte aligned (TMA).
 * splits > 1 partitions K across CTAs and requires DTF_EPI_F32_ATOMIC (D pre-zeroed by caller).
 * ReLU masks are bit matrices: uint32 word [m, n/32] bit (n%32); ldmask = words per row, even,
 *   >= ceil(N/32).  DTF_EPI_BIAS_RELU_BF16 writes one to mask_out when it is non-NULL (bit =
 *   output > 0); DTF_EPI_DRELU_BF16 reads mask_in.  D must satisfy the same alignment as A/B
 *   (it is written by TMA stores).
 * bn: N tile (64/128/256) or 0 to let the library choose.
 * Stands in for the Mul/Add/Sum Eigen kernels behind reference trainer/task.py:137-139. */
int dtf_gemm_bf16(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn,
                  void* D, long long ldd, int M, int N, int K, int epilogue, const float* bias,
                  const void* mask_in, void* mask_out, long long ldmask, int splits, int bn,
                  void* stream);

/* CTA-pair tiles: GEMMs with a pair instantiation run as clusters of two CTAs (the two SMs of a TPC)
 * that share one 256 x BN tile through tcgen05.mma.cta_group::2.  Process-wide switch; initial value
 * from the environment variable DTF_PAIR (default 1). */
void dtf_gemm_set_pair_mode(int on);
int dtf_gemm_get_pair_mode(void);

/* Debugging: device buffer (>= 71000 x 8 bytes, zeroed by the caller) that receives %globaltimer stamps
 * of the gather warps and TMA producers of the following fused-pull launches (bench/pull_trace.py);
 * NULL switches the stamps off. */
void dtf_debug_set_pull_trace(void* buf);

/* Fused parameter pull: H[M,N] = act([gather(row_src) | dense_x] * W^T + bias).  row_src[m*slots+f]
 * is the device address (local or peer-mapped) of the 128-wide bf16 embedding row of example m,
 * slot f; gather warps of the same persistent kernel fetch rows over NVLink while its tcgen05
 * warps multiply row-tiles that already arrived.  The gathered operand stays in a_buf
 * [M, slots*128 + dense_cols] (slots <= 63).  scratch: dtf_pull_gemm_scratch_bytes(M) bytes of device memory
 * owned by the caller (cleared by the call).  Replaces the ps->worker variable read (RecvTensor)
 * plus the first Mul of reference trainer/task.py:137. */
size_t dtf_pull_gemm_scratch_bytes(int M);
int dtf_pull_gemm_bf16(const unsigned long long* row_src, const void* dense_x, int slots, int dense_cols,
                       void* a_buf, const void* W, long long ldw, const float* bias, void* H, long long ldh,
                       void* mask_out, long long ldmask, int M, int N, int relu, void* scratch, void* stream);

/* Fused dense-parameter pull (all-gather + GEMM): D = act(A W^T + bias) with W [N,K] row-sharded over
 * `shards` ranks; W_shards is a host array of device pointers (local or peer-mapped), shard s holding
 * rows [s*rows_per_shard, ...).  Each N-tile's TMA loads come straight from the owning rank's HBM
 * through that rank's tensor map; no gathered copy of W exists.  rows_per_shard must be a multiple of
 * 64; maps_scratch = shards*128 bytes of 64-byte aligned device memory.  Replaces the dense variable
 * read ps->worker of reference trainer/task.py:137. */
int dtf_allgather_gemm_bf16(const void* A, long long lda, const void* const* W_shards, int shards,
                            int rows_per_shard, long long ldw, void* D, long long ldd, int M, int N, int K,
                            const float* bias, int relu, void* mask_out, long long ldmask, void* maps_scratch,
                            void* stream);

/* Fused gradient push: dA = dY[M,K] * W[K, :slots*128]; the bf16 row of (m, slot) is stored to the
 * address row_dst[m*slots+slot] (0 = skip), e.g. a mailbox entry in the owning rank's HBM.
 * Replaces the worker->ps gradient Send of reference trainer/task.py:138. */
int dtf_push_gemm_bf16(const void* dY, long long lddy, const void* W, long long ldw,
                       const unsigned long long* row_dst, int slots, int M, int K, void* stream);

/* Same computation as dtf_pull_gemm_bf16 for callers that cannot allocate device memory (a process
 * whose device memory was all handed to its own pool at start-up, a plug-in inside one, or a
 * request path that cannot tolerate cudaMalloc's device-wide synchronisation): the scratch is a
 * static device array inside the library (about 256 KiB per GPU that loads the library, held until
 * exit), so there is no scratch argument and M is limited to dtf_pull_gemm_noalloc_max_rows()
 * (1,048,576); larger M is rejected with an error.  The call allocates nothing.  All calls on a
 * device share that array, which keeps the previous call's contents (the used part is cleared at the
 * start of each call): calls whose GPU work overlaps race -- keep them stream-ordered.  The first
 * call on each device resolves the array's device address; issue it at start-up, before any stream
 * capture.  That address is cached per device for the life of the process: after a cudaDeviceReset
 * it is stale, so a process that resets a device must not call this entry point on it again.
 * This entry point has no counterpart in the reference. */
int dtf_pull_gemm_noalloc_max_rows(void);
int dtf_pull_gemm_bf16_noalloc(const unsigned long long* row_src, const void* dense_x, int slots,
                               int dense_cols, void* a_buf, const void* W, long long ldw, const float* bias,
                               void* H, long long ldh, void* mask_out, long long ldmask, int M, int N,
                               int relu, void* stream);

/* ---- symmetric heap (peer-mapped HBM over NVLink / NVSwitch) --------------------------------- *
 * Data plane that replaces the gRPC Send/Recv traffic of every sess.run in the reference
 * (trainer/task.py:236).  Handles are opaque byte strings exchanged by the host control plane. */
int dtf_symm_alloc(size_t bytes, void** out_ptr);             /* zero-filled; VMM allocation when possible */
int dtf_symm_free(void* ptr);
int dtf_symm_is_vmm(void* ptr);                               /* 1: shareable as a file descriptor */
long long dtf_symm_mapped_size(void* ptr);
int dtf_symm_export_fd(void* ptr, int* fd_out);               /* POSIX fd of a VMM slab */
int dtf_symm_import_fd(int fd, size_t size, void** out_ptr);  /* maps a peer's slab (2 MB pages) */
int dtf_symm_handle_size(void);
int dtf_symm_get_handle(void* ptr, void* handle_out);
int dtf_symm_open_handle(const void* handle, void** out_ptr); /* maps a peer's slab */
int dtf_symm_close_handle(void* ptr);
/* Copy by kernel between device memory and / or pinned host memory (UVA pointers): a transfer that is a
 * kernel node, not a copy-engine node, when captured in a CUDA graph.  blocks x 256 threads. */
int dtf_copy_async(void* dst, const void* src, size_t bytes, int blocks, void* stream);

int dtf_symm_enable_peer(int peer_device);                    /* same-process peers */

/* Multicast objects (NVLS): one object spans the same leading byte range of every rank's slab; on its
 * mapping multimem.ld_reduce returns the sum over all ranks (reduced inside NVSwitch) and multimem.st
 * writes every rank's copy.  Collective life cycle with host barriers between the phases: rank 0
 * dtf_mc_create + passes the fd; others dtf_mc_import; all dtf_mc_add_device; all dtf_mc_bind_map. */
int dtf_mc_supported(void);
long long dtf_mc_usable_bytes(int num_devices, size_t bytes);
int dtf_mc_create(int num_devices, size_t bytes, long long* id_out, int* fd_out);
int dtf_mc_import(int fd, size_t bytes, long long* id_out);
int dtf_mc_add_device(long long id);
int dtf_mc_bind_map(long long id, void* slab, void** mc_ptr_out);
int dtf_mc_destroy(long long id);   /* unmap this process's view and release the object, exactly once */

/* ---- wide-and-deep plan: the training / serving step executor -------------------------------- *
 * Every rank is a worker and a parameter-server shard at once (embedding row r on rank
 * r % world, dense parameters range-partitioned), the B200 mapping of the ps / worker / master
 * roles of reference trainer/task.py:121-127.  One step = route -> fused pull (embedding gather +
 * GEMM-1) -> MLP -> loss -> backward -> fused push (dgrad GEMM scattering rows into the owners'
 * mailboxes) -> [barrier] -> ps-side apply (sparse scatter-add + dense reduce-scatter, each fused
 * with SGD / Adagrad / Adam) -> [barrier].  Stands in for sess.run(train_op) (task.py:236). */
typedef struct dtf_wd_config {
  int world, rank;
  int batch;          /* examples per rank per step */
  int slots;          /* sparse ids per example (F) */
  int emb_dim;        /* must be 128 */
  int dense_in;       /* dense features per example */
  int dense_pad;      /* dense block width inside the GEMM-1 operand: 64 */
  int h1, h2, h3;     /* hidden sizes; h3 multiple of 256 */
  long long emb_rows; /* global rows of the embedding (and wide) table */
  int optimizer;      /* 0 sgd, 1 adadelta, 2 adagrad, 3 adam, 4 ftrl, 5 rmsprop (same ids as dtf_native.h) */
  int sync;           /* 1: sync-replica (sum all replicas, apply once); 0: async / Hogwild */
  int no_device_barrier; /* 1: caller orders ranks itself (several ranks emulated on one GPU) */
  float lr;
  float emb_init_scale;
  int overlap_apply_blocks; /* async overlap: apply blocks per SM (0 = default 2) */
} dtf_wd_config;

typedef struct dtf_wd_info_t {
  int K0;
  long long rows_local, P, P_pad, dense_lo, dense_hi;
  long long off_w1, off_w2, off_w3, off_b1, off_b2, off_b3, off_w4, off_b4, off_wide_b, small_begin;
  long long slab_bytes, step;
} dtf_wd_info_t;

typedef struct dtf_wd_plan dtf_wd_plan;
dtf_wd_plan* dtf_wd_create(const dtf_wd_config* cfg);   /* allocates every buffer incl. the slab */
void dtf_wd_destroy(dtf_wd_plan*);
int dtf_wd_info(dtf_wd_plan*, dtf_wd_info_t* out);
/* named device buffers ("emb_master", "emb_slot1", "wide", "dense_master", "dense_bf16",
 * "dense_grad", "h1", "prob", ...) for checkpointing and tests */
int dtf_wd_buffer(dtf_wd_plan*, const char* name, void** ptr, size_t* bytes);
int dtf_wd_slab_handle(dtf_wd_plan*, void* handle_out);
void* dtf_wd_slab_ptr(dtf_wd_plan*);
int dtf_wd_slab_is_vmm(dtf_wd_plan*);
int dtf_wd_slab_fd(dtf_wd_plan*, int* fd_out);
int dtf_wd_attach_fds(dtf_wd_plan*, const int* fds /* world descriptors, own entry ignored */);
int dtf_wd_attach_multicast(dtf_wd_plan*, void* mc_ptr, long long bytes);   /* NVLS path of the dense apply; NULL = off */
int dtf_wd_attach_ipc(dtf_wd_plan*, const void* handles /* world * handle_size bytes; legacy, slow mappings */);
int dtf_wd_attach_ptrs(dtf_wd_plan*, void* const* slabs /* world pointers */);
int dtf_wd_init_params(dtf_wd_plan*, unsigned long long seed, void* stream);
int dtf_wd_broadcast_dense(dtf_wd_plan*, void* stream);
int dtf_wd_refresh_embedding(dtf_wd_plan*, void* stream);
int dtf_wd_step_compute(dtf_wd_plan*, const int* ids, const float* dense, const float* labels, void* stream);
int dtf_wd_barrier(dtf_wd_plan*, int publish_counts, void* stream);
/* Failure detection on the device (the reference's workers hang in gRPC when a ps dies, SURVEY.md 5.3;
 * here a dead peer must not hang the survivors): a device barrier waits at most this long for a peer
 * (default 30 000 ms, env DTF_BARRIER_TIMEOUT_MS, 0 = forever), then records the missing rank.
 * dtf_wd_check_error: 0 healthy, r + 1 = gave up on rank r (sticky), < 0 CUDA error; does not
 * synchronise the training streams. */
void dtf_wd_set_barrier_timeout_ms(long long ms);
int dtf_wd_check_error(dtf_wd_plan*);
int dtf_wd_step_apply(dtf_wd_plan*, void* stream);
int dtf_wd_step(dtf_wd_plan*, const int* ids, const float* dense, const float* labels, void* stream);
int dtf_wd_predict(dtf_wd_plan*, const int* ids, const float* dense, int n /* <= batch */, void* stream);
int dtf_wd_read_loss(dtf_wd_plan*, float* loss_sum, void* stream);
void* dtf_wd_loss_ptr(dtf_wd_plan*);   /* loss accumulator of the most recent compute phase */
int dtf_wd_last_parity(dtf_wd_plan*);
/* Async mode only: compute of step t on compute_stream, ps-side barrier + apply of step t on
 * apply_stream, overlapping the compute of step t+1 (bounded staleness of one step; mailboxes,
 * gradient buffers and scratch are double-buffered by step parity).  The async counterpart of a TF
 * worker that does not wait for the ps to apply (reference trainer/task.py:138, no sync wrapper). */
int dtf_wd_step_overlapped(dtf_wd_plan*, const int* ids, const float* dense, const float* labels,
                           void* compute_stream, void* apply_stream);
int dtf_wd_join(dtf_wd_plan*, void* stream);      /* stream waits for every apply issued so far */
void dtf_wd_overlap_reset(dtf_wd_plan*);
/* async mode: announce the next step's batch so that its rows are gathered (low-priority stream) during the coming
 * step; 1 = taken, 0 = ignored (sync plan, DTF_PREFETCH=0, profiling) */
int dtf_wd_set_next_batch(dtf_wd_plan*, const int* ids, const float* dense_raw);
long long dtf_wd_prefetched_steps(dtf_wd_plan*);   /* steps whose layer 1 ran on prefetched rows */
/* DTF_STAMPS=1 diagnostics: device-clock stamp at every phase mark, readable while the streams are busy or stuck */
long long dtf_wd_stamps_count(dtf_wd_plan*);       /* marks issued so far; -1 when stamps are off */
int dtf_wd_stamps_read(dtf_wd_plan*, long long from, long long to, int clear, char* buf, size_t cap);
/* per-kernel tracing with CUDA events (the reference only prints wall-clock, trainer/task.py:100-102) */
void dtf_wd_profile_enable(dtf_wd_plan*, int on);
int dtf_wd_profile_read(dtf_wd_plan*, char* buf, size_t cap);
int dtf_wd_profile_read_abs(dtf_wd_plan*, char* buf, size_t cap);  /* offsets from the first mark */

/* The per-element optimizer rule of the ps-side apply kernels (reference trainer/task.py:41-56: sgd,
 * adadelta, adagrad, adam, ftrl, rmsprop with TF-1 defaults), run on the host over n elements: lets
 * CPU tests compare the device math with the NumPy optimizers.  `step` (>= 1) only matters for adam. */
int dtf_opt_update_host(int optimizer, float* w, const float* g, float* slot1, float* slot2, long long n,
                        float lr, long long step);
float dtf_opt_slot1_init(int optimizer);
int dtf_opt_num_slots(int optimizer);

#ifdef __cplusplus
}
#endif
#endif /* DTF_KERNELS_H_ */
