/* C ABI of libdtf_native.so -- the CPU runtime of distributed_tensorflow_b200.
 *
 * The reference (yaokeepmoving/distributed_tensorflow) ships no native code; it leans on the
 * C++ runtime of TensorFlow 1.x.  Each block below names the reference call site whose TF C++
 * component it replaces.
 *
 * Conventions: functions returning int give 0 on success, negative on failure, and
 * dtfn_last_error() (thread local) describes the failure.  Handles are opaque pointers.
 */
#ifndef DTF_NATIVE_H_
#define DTF_NATIVE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* dtfn_last_error(void);

/* ---- crc32c (Castagnoli), as used by TFRecord framing and the tensor bundle ---------------- */
uint32_t dtfn_crc32c(const void* data, size_t n);
uint32_t dtfn_crc32c_extend(uint32_t crc, const void* data, size_t n);
uint32_t dtfn_crc32c_mask(uint32_t crc);   /* rotr15(crc) + 0xa282ead8 */
uint32_t dtfn_crc32c_unmask(uint32_t masked);

/* ---- immutable sorted table (LevelDB / TF table format) ------------------------------------ *
 * Replaces tensorflow/core/lib/io/table_builder + table behind tf.train.Saver
 * (reference trainer/task.py:143).  Keys must be added in strictly increasing bytewise order. */
typedef struct dtfn_table_builder dtfn_table_builder;
dtfn_table_builder* dtfn_table_builder_new(const char* path, int block_size);
int dtfn_table_builder_add(dtfn_table_builder*, const void* key, size_t klen, const void* val, size_t vlen);
int dtfn_table_builder_finish(dtfn_table_builder*);  /* writes footer, closes, frees */
void dtfn_table_builder_abandon(dtfn_table_builder*);

typedef struct dtfn_table dtfn_table;
dtfn_table* dtfn_table_open(const char* path);  /* reads + verifies every block */
long long dtfn_table_count(const dtfn_table*);
int dtfn_table_entry(const dtfn_table*, long long i, const void** key, size_t* klen, const void** val,
                     size_t* vlen);
int dtfn_table_find(const dtfn_table*, const void* key, size_t klen, const void** val, size_t* vlen);
void dtfn_table_close(dtfn_table*);

/* ---- tensor bundle (TF checkpoint V2) ------------------------------------------------------ *
 * Replaces BundleWriter / BundleReader / MergeBundles (SaveV2, RestoreV2, MergeV2Checkpoints)
 * behind tf.train.Saver / Saver(sharded=True) (reference trainer/task.py:143,
 * trainer/task_supervisor.py:145).  dtype values are TF DataType enum numbers. */
typedef struct dtfn_bundle_writer dtfn_bundle_writer;
dtfn_bundle_writer* dtfn_bundle_writer_new(const char* prefix);
int dtfn_bundle_writer_add(dtfn_bundle_writer*, const char* name, int dtype, int ndims, const long long* dims,
                           const void* data, size_t nbytes);
int dtfn_bundle_writer_finish(dtfn_bundle_writer*);  /* closes + frees */
void dtfn_bundle_writer_abandon(dtfn_bundle_writer*);
/* Merge n single-shard bundles into one n-shard bundle at `merged_prefix`; source index files are
 * removed and their data files renamed, like MergeV2Checkpoints(delete_old_dirs). */
int dtfn_bundle_merge(const char* const* prefixes, int n, const char* merged_prefix);

typedef struct dtfn_bundle_reader dtfn_bundle_reader;
dtfn_bundle_reader* dtfn_bundle_reader_open(const char* prefix);
int dtfn_bundle_reader_num_shards(const dtfn_bundle_reader*);
long long dtfn_bundle_reader_count(const dtfn_bundle_reader*);
/* name is returned as pointer+length into reader-owned memory; dims has room for 8 entries */
int dtfn_bundle_reader_entry(const dtfn_bundle_reader*, long long i, const char** name, size_t* name_len,
                             int* dtype, int* ndims, long long* dims, int* shard_id, long long* offset,
                             long long* size, uint32_t* masked_crc);
int dtfn_bundle_reader_lookup(const dtfn_bundle_reader*, const char* name);  /* index or -1 */
int dtfn_bundle_reader_read(const dtfn_bundle_reader*, long long i, void* dst, size_t nbytes);  /* crc checked */
/* > 0: entry i is the full tensor of a partitioned variable (BundleEntryProto.slices); reading it assembles
 * the slices, which live under OrderedCode-encoded keys (start with a zero byte) next to the tensor names */
int dtfn_bundle_reader_num_slices(const dtfn_bundle_reader*, long long i);
void dtfn_bundle_reader_close(dtfn_bundle_reader*);

/* ---- TFRecord framing (event files) -------------------------------------------------------- *
 * Replaces RecordWriter/EventsWriter behind tf.summary.FileWriter (reference task.py:80). */
typedef struct dtfn_record_writer dtfn_record_writer;
dtfn_record_writer* dtfn_record_writer_open(const char* path);
int dtfn_record_writer_write(dtfn_record_writer*, const void* data, size_t n);
int dtfn_record_writer_flush(dtfn_record_writer*);
void dtfn_record_writer_close(dtfn_record_writer*);

/* ---- optimizer apply kernels (CPU parameter server) ---------------------------------------- *
 * Replaces ApplyGradientDescent/ApplyAdagrad/ApplyAdam/ApplyAdadelta/ApplyFtrl/ApplyRMSProp
 * (use_locking=False) behind optimizer.minimize (reference trainer/task.py:41-53,138).
 * All arrays are float32 of length n; hyper-parameters follow TF-1 defaults. */
enum { DTFN_OPT_SGD = 0, DTFN_OPT_ADADELTA = 1, DTFN_OPT_ADAGRAD = 2, DTFN_OPT_ADAM = 3, DTFN_OPT_FTRL = 4,
       DTFN_OPT_RMSPROP = 5 };
void dtfn_apply_sgd(float* var, const float* g, size_t n, float lr);
void dtfn_apply_adagrad(float* var, float* accum, const float* g, size_t n, float lr);
void dtfn_apply_adam(float* var, float* m, float* v, const float* g, size_t n, float lr, float beta1, float beta2,
                     float eps, float beta1_power, float beta2_power);
void dtfn_apply_adadelta(float* var, float* accum, float* accum_update, const float* g, size_t n, float lr,
                         float rho, float eps);
void dtfn_apply_ftrl(float* var, float* accum, float* linear, const float* g, size_t n, float lr, float l1,
                     float l2, float lr_power);
void dtfn_apply_rmsprop(float* var, float* ms, float* mom, const float* g, size_t n, float lr, float decay,
                        float momentum, float eps);

/* ---- serving JSON fast path --------------------------------------------------------------- *
 * The predict endpoint's bodies are objects of rectangular nested numeric arrays (reference
 * README.md:40).  dtfn_json_numeric_array finds `key` at the top level (or inside a top-level "data"
 * envelope) and parses its value into out (row-major doubles): returns the number of values and fills
 * dims[0..*ndims) (ndims <= 4); -1 key absent, -2 not a plain rectangular numeric array (use a generic
 * JSON parser), -3 out too small.  dtfn_json_format_array writes a row-major array as nested JSON
 * arrays (kind 0: float32 in the shortest form that round-trips, 1: int64); returns bytes written or
 * minus the capacity it needs. */
long long dtfn_json_numeric_array(const char* body, size_t n, const char* key, double* out, size_t cap,
                                  long long* dims, int* ndims);
long long dtfn_json_format_array(const void* data, int kind, const long long* dims, int ndims, char* out, size_t cap);

/* ---- parameter-server transport ------------------------------------------------------------ *
 * Replaces the in-process gRPC tf.train.Server (MasterService + WorkerService), the Send/Recv
 * rendezvous and the shared-name FIFOQueue resource (reference trainer/task.py:118-122,
 * auto_stop_ps/task.py:127-150).  One server per task; ps tasks host variables and queues. */
typedef struct dtfn_server dtfn_server;
dtfn_server* dtfn_server_start(const char* host, int port);  /* binds + starts the accept thread */
int dtfn_server_port(const dtfn_server*);
void dtfn_server_join(dtfn_server*);                         /* blocks until the server is stopped */
void dtfn_server_stop(dtfn_server*);                         /* stops threads, closes sockets, frees */
/* test hook: drop the next n requests without answering (fault injection) */
void dtfn_server_inject_drop(dtfn_server*, int n);
/* direct (in-process) access to the variable store of a server, used by the ps itself */
long long dtfn_server_num_variables(const dtfn_server*);

typedef struct dtfn_client dtfn_client;
dtfn_client* dtfn_client_connect(const char* host, int port, int timeout_ms);
void dtfn_client_close(dtfn_client*);
void dtfn_client_set_timeout(dtfn_client*, int timeout_ms);
int dtfn_client_ping(dtfn_client*);
/* create (or overwrite when overwrite != 0) a float32/int32/int64 variable on the server */
int dtfn_client_init_var(dtfn_client*, const char* name, int dtype, int ndims, const long long* dims,
                         const void* data, size_t nbytes, int overwrite);
/* copy a variable's bytes into dst; returns bytes written or negative */
long long dtfn_client_pull(dtfn_client*, const char* name, void* dst, size_t cap);
/* list variable names as a '\n'-joined string into buf; returns needed length */
long long dtfn_client_list(dtfn_client*, char* buf, size_t cap);
/* metadata: dtype, dims (room for 8); returns ndims or negative when absent */
int dtfn_client_var_info(dtfn_client*, const char* name, int* dtype, long long* dims, long long* nbytes);
/* apply one gradient to one variable with the given optimizer (slots are created server-side and
 * named like TF: <var>/Adagrad, <var>/Adam, <var>/Adam_1, ...); lock-free on the server */
int dtfn_client_push_apply(dtfn_client*, const char* name, int optimizer, float lr, const float* grad, size_t n);
/* one fused round trip for a training step: push k gradients, add step_inc to the int64/int32
 * variable `step_name` (may be NULL), and pull the k updated variables back (dst may be NULL).
 * For Adam the server advances beta powers once per call. */
int dtfn_client_push_pull(dtfn_client*, int k, const char* const* names, int optimizer, float lr,
                          const float* const* grads, const size_t* lens, const char* step_name,
                          long long step_inc, float* const* dst, long long* step_out);
int dtfn_client_assign_add_i64(dtfn_client*, const char* name, long long delta, long long* out);
/* named bounded blocking queue of int32 living on the server (shared_name FIFOQueue) */
int dtfn_client_queue_create(dtfn_client*, const char* name, int capacity);
int dtfn_client_enqueue(dtfn_client*, const char* name, int value);
int dtfn_client_dequeue(dtfn_client*, const char* name, int* value, int timeout_ms);  /* -2 on timeout */
/* ask the server to write its variables as a single-shard bundle at prefix (ps-side SaveV2);
 * names listed in the '\n'-joined exclude string (may be NULL) are skipped */
int dtfn_client_save(dtfn_client*, const char* prefix, const char* exclude_newline_joined);
int dtfn_client_shutdown_server(dtfn_client*);

#ifdef __cplusplus
}
#endif
#endif /* DTF_NATIVE_H_ */
