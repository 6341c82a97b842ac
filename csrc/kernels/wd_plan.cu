// Host-side executor of the wide-and-deep training / serving step: owns every device buffer, the
// symmetric slab and the peer table, and launches the whole step (about twenty kernels) from one C
// call so that Python is not on the critical path.  It plays the part of TensorFlow's C++
// master/worker executor behind sess.run(train_op) in the reference (reference
// trainer/task.py:236).
#include <nvtx3/nvToolsExt.h>
#include <math.h>
#include <string.h>

#include <map>
#include <string>
#include <utility>
#include <type_traits>
#include <vector>

#include "../../include/dtf_kernels.h"
#include "gemm_launch.cuh"
#include "wd_kernels.cuh"

using namespace dtf;

namespace {

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Buf {
  void* ptr = nullptr;
  size_t bytes = 0;
};

}  // namespace

struct dtf_wd_plan {
  dtf_wd_config cfg{};
  WdShapes sh{};
  DenseLayout dl{};
  SlabLayout sl{};
  PeerTable pt{};
  long long P_pad = 0, dense_lo = 0, dense_hi = 0;
  uint8_t* slab = nullptr;
  bool owns_slab = true;
  std::vector<void*> opened;  // IPC mappings to close
  std::map<std::string, Buf> bufs;
  std::vector<void*> owned;
  // local buffers
  float *emb_master = nullptr, *emb_s1 = nullptr, *emb_s2 = nullptr, *G = nullptr;
  float *wide_s1 = nullptr, *wide_s2 = nullptr, *wide_G = nullptr;
  unsigned long long* row_state = nullptr;   // sync mode: per-row contribution counters
  float *dense_master = nullptr, *dense_s1 = nullptr, *dense_s2 = nullptr;
  __nv_bfloat16 *a_buf = nullptr, *dense_x = nullptr, *H1 = nullptr, *H2 = nullptr, *H3 = nullptr;
  __nv_bfloat16 *dZ1 = nullptr, *dZ2 = nullptr, *dZ3 = nullptr;
  uint32_t *m1 = nullptr, *m2 = nullptr, *m3 = nullptr;
  int ldm1 = 0, ldm2 = 0, ldm3 = 0;
  unsigned long long *row_src = nullptr, *wide_src = nullptr, *row_dst = nullptr;
  float* wide_vals = nullptr;   // [B*F] wide weights of the batch's ids, filled by the fused pull
  float* dlogit = nullptr;
  float* prob = nullptr;
  // zeroed at the start of every step: cnt[kMaxWorld] | chunk_counter | pad | loss_acc[2] | tile_flags[m_tiles][kFlagWords] (64-bit)
  unsigned int* zero_region = nullptr;
  size_t zero_bytes = 0;
  unsigned int *cnt = nullptr, *chunk_counter = nullptr;
  unsigned long long* tile_flags = nullptr;
  float* loss_acc = nullptr;
  // optional per-kernel tracing: CUDA events recorded between launches
  bool profile = false;
  std::vector<std::pair<std::string, cudaEvent_t>> marks;
  unsigned long long* gemm_dbg = nullptr;    // DTF_GEMM_DBG=1: stall records of the GEMM kernels (GemmParams::dbg)
  int dbg_id = 0;                            // id the next GEMM launch reports under
  unsigned long long* stamps = nullptr;      // DTF_STAMPS=1: ring of device-clock stamps, one per mark()
  std::vector<std::string> stamp_names;
  long long n_stamp = 0;
  StepState* state = nullptr;  // device: step counter, barrier epoch, optimizer constants
  long long step = 0;          // host mirror of the number of applies issued
  // Async mode double-buffers everything a step hands from its compute phase to its apply phase
  // (mailboxes, dense-gradient buffer, per-step scratch) by step parity, so that the ps-side apply of
  // step t can run on a second stream while the compute phase of step t+1 is already running.
  SlabLayout sl_par[2];
  unsigned int* zero_par[2] = {nullptr, nullptr};
  long long n_compute = 0, n_apply = 0;
  cudaEvent_t ev_push[2] = {nullptr, nullptr}, ev_apply[2] = {nullptr, nullptr};
  // side stream: the NVLink-bound fused push runs on `push_ctas` SMs next to wgrad-1 on the others
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  int push_ctas = 0;
  uint8_t* mc = nullptr;   // multicast (NVLS) mapping of all ranks' slabs, nullptr = plain peer loads / stores
  bool ev_apply_valid[2] = {false, false};
  int last_compute_parity = 0;
  int apply_blocks_per_sm = 8;   // lowered while the apply shares the SMs with the next step's GEMMs
  // Async-mode parameter prefetch: the operand buffer and the wide values exist once per step parity; while step t
  // computes, a low-priority stream gathers step t+1's rows into the other parity's buffers (dtf_wd_set_next_batch).
  __nv_bfloat16* a_buf_par[2] = {nullptr, nullptr};
  float* wide_vals_par[2] = {nullptr, nullptr};
  unsigned long long *row_src_pf = nullptr, *wide_src_pf = nullptr;
  cudaStream_t pf_stream = nullptr;
  cudaEvent_t ev_pf[2] = {nullptr, nullptr}, ev_cstart = nullptr;
  bool pf_valid[2] = {false, false};
  const int* pf_ids[2] = {nullptr, nullptr};   // the ids pointer a parity's prefetched rows belong to
  const int* next_ids = nullptr;
  const float* next_dense = nullptr;
  int cur_par = 0;
  long long n_prefetched = 0;                  // steps whose layer 1 ran on prefetched rows

  template <typename T>
  T* alloc(const char* name, size_t count) {
    void* p = nullptr;
    const size_t bytes = count * sizeof(T);
    if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) {
      set_error("wd plan: cudaMalloc(%zu bytes) for '%s' failed", bytes, name);
      (void)cudaGetLastError();
      return nullptr;
    }
    cudaMemset(p, 0, bytes ? bytes : 16);
    owned.push_back(p);
    bufs[name] = Buf{p, bytes};
    return static_cast<T*>(p);
  }
  template <typename T>
  T* in_slab(const char* name, size_t off, size_t bytes) {
    bufs[name] = Buf{slab + off, bytes};
    return reinterpret_cast<T*>(slab + off);
  }
};

// zero fill by kernel instead of cudaMemsetAsync: memset nodes on IPC-exported (peer-mapped) memory
// made CUDA-graph replays of the multi-rank step slow
__global__ void wd_zero_kernel(uint4* __restrict__ p, size_t n16, uint4* __restrict__ q, size_t m16) {
  pdl_launch_dependents();
  pdl_wait();
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t t = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  for (size_t i = t; i < n16; i += stride) p[i] = make_uint4(0, 0, 0, 0);
  for (size_t i = t; i < m16; i += stride) q[i] = make_uint4(0, 0, 0, 0);
}
// zero one or two regions with one launch
static int zero_fill(void* ptr, size_t bytes, cudaStream_t st, void* ptr2 = nullptr, size_t bytes2 = 0) {
  const size_t n16 = (bytes + 15) / 16, m16 = (bytes2 + 15) / 16;
  int grid = static_cast<int>(((n16 > m16 ? n16 : m16) + 255) / 256);
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  if (grid < 1) grid = 1;
  DTF_CUDA_OK(launch_kernel(true, wd_zero_kernel, dim3(grid), dim3(256), 0, st, static_cast<uint4*>(ptr), n16,
                            static_cast<uint4*>(ptr2), m16));
  count_launch();
  return 0;
}

// NVTX (DTF_NVTX=1): one range per executor call and a marker at the end of every kernel group, so that
// a timeline tool shows pull / MLP / loss / backward / push / apply without per-kernel name matching.
static bool nvtx_on() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DTF_NVTX");
    v = (e != nullptr && e[0] == '1') ? 1 : 0;
  }
  return v == 1;
}
struct NvtxRange {
  bool on;
  explicit NvtxRange(const char* name) : on(nvtx_on()) {
    if (on) nvtxRangePushA(name);
  }
  ~NvtxRange() {
    if (on) nvtxRangePop();
  }
};

// DTF_STAMPS=1: every mark also launches a one-thread kernel that writes the device clock into the next slot of
// a ring, under CUDA-graph capture too (a replay overwrites the slots of its capture), so the timeline of the
// overlapped streams -- or, read from a second stream while the job is stuck, the last point every stream
// reached -- can be printed with dtf_wd_stamps_read.  Diagnostics only: the extra kernels cost a few us per step.
constexpr int kStampCap = 8192;
__global__ void wd_stamp_kernel(unsigned long long* slot) { *slot = globaltimer_ns(); }

static void mark(dtf_wd_plan* p, cudaStream_t st, const char* name) {
  if (nvtx_on()) nvtxMarkA(name);
  if (p->stamps != nullptr) {
    const int idx = static_cast<int>(p->n_stamp % kStampCap);
    p->stamp_names[idx] = name;
    p->n_stamp += 1;
    wd_stamp_kernel<<<1, 1, 0, st>>>(p->stamps + idx);
  }
  if (!p->profile) return;
  cudaEvent_t e;
  cudaEventCreate(&e);
  cudaEventRecord(e, st);
  p->marks.emplace_back(name, e);
}

static void select_parity(dtf_wd_plan* p, int par) {
  p->sl = p->sl_par[par];
  p->zero_region = p->zero_par[par];
  p->cnt = p->zero_region;
  p->chunk_counter = p->zero_region + kMaxWorld;
  p->loss_acc = reinterpret_cast<float*>(p->zero_region + kMaxWorld + 4);
  p->tile_flags = reinterpret_cast<unsigned long long*>(p->zero_region + kMaxWorld + 8);
  p->a_buf = p->a_buf_par[par];
  p->wide_vals = p->wide_vals_par[par];
  p->cur_par = par;
}

static int mask_ld(int n) {
  int w = (n + 31) / 32;
  return w + (w & 1);
}


extern "C" {

dtf_wd_plan* dtf_wd_create(const dtf_wd_config* cfg) {
  if (cfg->world < 1 || cfg->world > kMaxWorld || cfg->rank < 0 || cfg->rank >= cfg->world) {
    set_error("wd plan: bad world/rank %d/%d (max world %d)", cfg->world, cfg->rank, kMaxWorld);
    return nullptr;
  }
  if (cfg->emb_dim != kEmbDim) {
    set_error("wd plan: embedding_dim must be %d (one row = two 128-byte swizzle rows)", kEmbDim);
    return nullptr;
  }
  if (cfg->h3 % 256 != 0 || cfg->h3 > 1024 || cfg->h1 % 64 || cfg->h2 % 64 || cfg->dense_pad % 64 ||
      cfg->dense_in > cfg->dense_pad || cfg->batch < 1 || cfg->slots < 1) {
    set_error("wd plan: unsupported shape (h1,h2 multiples of 64; h3 multiple of 256 <= 1024; dense_pad multiple of 64)");
    return nullptr;
  }
  if (cfg->optimizer < WD_SGD || cfg->optimizer > WD_RMSPROP) {
    set_error("wd plan: unknown optimizer id %d (sgd=0, adadelta=1, adagrad=2, adam=3, ftrl=4, rmsprop=5)", cfg->optimizer);
    return nullptr;
  }
  if (cfg->dense_pad != 64) {
    // the gather warps of the fused pull move the dense block as one 64-column chunk per row
    set_error("wd plan: dense_pad must be 64 (got %d)", cfg->dense_pad);
    return nullptr;
  }
  auto* p = new dtf_wd_plan();
  p->cfg = *cfg;
  WdShapes& sh = p->sh;
  sh.B = cfg->batch; sh.F = cfg->slots; sh.dense_in = cfg->dense_in; sh.dense_pad = cfg->dense_pad;
  sh.K0 = cfg->slots * kEmbDim + cfg->dense_pad;
  sh.h1 = cfg->h1; sh.h2 = cfg->h2; sh.h3 = cfg->h3;
  sh.V = cfg->emb_rows;
  sh.rows_local = (cfg->emb_rows - cfg->rank + cfg->world - 1) / cfg->world;
  sh.rows_local_max = (cfg->emb_rows + cfg->world - 1) / cfg->world;
  sh.cap = static_cast<long long>(cfg->batch) * cfg->slots;

  DenseLayout& dl = p->dl;
  long long o = 0;
  dl.w1 = o; o = align_up(o + static_cast<long long>(sh.h1) * sh.K0, 256);
  dl.w2 = o; o = align_up(o + static_cast<long long>(sh.h2) * sh.h1, 256);
  dl.w3 = o; o = align_up(o + static_cast<long long>(sh.h3) * sh.h2, 256);
  dl.small_begin = o;
  dl.b1 = o; o = align_up(o + sh.h1, 256);
  dl.b2 = o; o = align_up(o + sh.h2, 256);
  dl.b3 = o; o = align_up(o + sh.h3, 256);
  dl.w4 = o; o = align_up(o + sh.h3, 256);
  dl.b4 = o; o += 4;
  dl.wide_b = o; o += 4;
  dl.P = align_up(o, 256);
  dl.n_small = static_cast<int>(dl.P - dl.small_begin);
  p->P_pad = align_up(dl.P, 256ll * cfg->world);
  const long long chunk = p->P_pad / cfg->world;
  p->dense_lo = chunk * cfg->rank;
  p->dense_hi = p->dense_lo + chunk;

  SlabLayout& sl = p->sl;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t r = off; off = align_up(off + bytes, 1024); return r; };
  sl.emb_bf16 = take(static_cast<size_t>(sh.rows_local_max) * kEmbDim * 2);
  sl.wide = take(static_cast<size_t>(sh.rows_local_max) * 4);
  sl.wbf16 = take(static_cast<size_t>(p->P_pad) * 2);
  sl.small_f32 = take(static_cast<size_t>(dl.n_small + (p->P_pad - dl.P)) * 4);
  sl.dense_grad = take(static_cast<size_t>(p->P_pad) * 4);
  sl.mbx_payload = take(static_cast<size_t>(cfg->world) * sh.cap * kEmbDim * 2);
  sl.mbx_rows = take(static_cast<size_t>(cfg->world) * sh.cap * 8);
  sl.mbx_dlvec = take(static_cast<size_t>(cfg->world) * sh.B * 4);
  sl.mbx_count = take(kMaxWorld * 4);
  sl.signal = take(kMaxWorld * 64);
  p->sl_par[0] = sl;
  p->sl_par[1] = sl;
  if (!cfg->sync) {
    SlabLayout& s1 = p->sl_par[1];
    s1.dense_grad = take(static_cast<size_t>(p->P_pad) * 4);
    s1.mbx_payload = take(static_cast<size_t>(cfg->world) * sh.cap * kEmbDim * 2);
    s1.mbx_rows = take(static_cast<size_t>(cfg->world) * sh.cap * 8);
    s1.mbx_dlvec = take(static_cast<size_t>(cfg->world) * sh.B * 4);
    s1.mbx_count = take(kMaxWorld * 4);
  }
  sl.total = off;
  p->sl_par[0].total = p->sl_par[1].total = off;

  void* slab = nullptr;
  if (dtf_symm_alloc(sl.total, &slab) != 0) {
    delete p;
    return nullptr;
  }
  p->slab = static_cast<uint8_t*>(slab);
  p->pt.world = cfg->world;
  p->pt.rank = cfg->rank;
  for (int r = 0; r < kMaxWorld; ++r) p->pt.slab[r] = nullptr;
  p->pt.slab[cfg->rank] = p->slab;

  p->in_slab<void>("emb_bf16", sl.emb_bf16, static_cast<size_t>(sh.rows_local) * kEmbDim * 2);
  p->in_slab<void>("wide", sl.wide, static_cast<size_t>(sh.rows_local) * 4);
  p->in_slab<void>("dense_bf16", sl.wbf16, static_cast<size_t>(p->P_pad) * 2);
  p->in_slab<void>("dense_small_f32", sl.small_f32, static_cast<size_t>(dl.n_small) * 4);
  p->in_slab<void>("dense_grad", sl.dense_grad, static_cast<size_t>(p->P_pad) * 4);
  p->in_slab<void>("mbx_payload", sl.mbx_payload, static_cast<size_t>(cfg->world) * sh.cap * kEmbDim * 2);
  p->in_slab<void>("mbx_rows", sl.mbx_rows, static_cast<size_t>(cfg->world) * sh.cap * 8);
  p->in_slab<void>("mbx_dlvec", sl.mbx_dlvec, static_cast<size_t>(cfg->world) * sh.B * 4);
  p->in_slab<void>("mbx_count", sl.mbx_count, kMaxWorld * 4);
  p->in_slab<void>("signal", sl.signal, kMaxWorld * 64);   // barrier pads (diagnostics: bench/stress_overlap.py)

  const size_t ne = static_cast<size_t>(sh.rows_local) * kEmbDim;
  const bool has1 = opt_has_slot1(cfg->optimizer), has2 = opt_has_slot2(cfg->optimizer);
  bool ok = true;
  ok &= (p->emb_master = p->alloc<float>("emb_master", ne)) != nullptr;
  if (has1) ok &= (p->emb_s1 = p->alloc<float>("emb_slot1", ne)) != nullptr;
  if (has2) ok &= (p->emb_s2 = p->alloc<float>("emb_slot2", ne)) != nullptr;
  if (has1) ok &= (p->wide_s1 = p->alloc<float>("wide_slot1", sh.rows_local)) != nullptr;
  if (has2) ok &= (p->wide_s2 = p->alloc<float>("wide_slot2", sh.rows_local)) != nullptr;
  if (cfg->sync) {
    ok &= (p->G = p->alloc<float>("emb_grad_accum", ne)) != nullptr;
    ok &= (p->wide_G = p->alloc<float>("wide_grad_accum", sh.rows_local)) != nullptr;
    ok &= (p->row_state = p->alloc<unsigned long long>("row_state", sh.rows_local)) != nullptr;
  }
  ok &= (p->dense_master = p->alloc<float>("dense_master", chunk)) != nullptr;
  if (has1) ok &= (p->dense_s1 = p->alloc<float>("dense_slot1", chunk)) != nullptr;
  if (has2) ok &= (p->dense_s2 = p->alloc<float>("dense_slot2", chunk)) != nullptr;
  const size_t B = sh.B;
  ok &= (p->a_buf_par[0] = p->alloc<__nv_bfloat16>("a_buf", B * sh.K0)) != nullptr;
  p->a_buf_par[1] = p->a_buf_par[0];
  if (!cfg->sync) {
    ok &= (p->a_buf_par[1] = p->alloc<__nv_bfloat16>("a_buf1", B * sh.K0)) != nullptr;
    ok &= (p->row_src_pf = p->alloc<unsigned long long>("row_src_pf", sh.cap)) != nullptr;
    ok &= (p->wide_src_pf = p->alloc<unsigned long long>("wide_src_pf", sh.cap)) != nullptr;
  }
  ok &= (p->dense_x = p->alloc<__nv_bfloat16>("dense_x", B * sh.dense_pad)) != nullptr;
  ok &= (p->H1 = p->alloc<__nv_bfloat16>("h1", B * sh.h1)) != nullptr;
  ok &= (p->H2 = p->alloc<__nv_bfloat16>("h2", B * sh.h2)) != nullptr;
  ok &= (p->H3 = p->alloc<__nv_bfloat16>("h3", B * sh.h3)) != nullptr;
  ok &= (p->dZ1 = p->alloc<__nv_bfloat16>("dz1", B * sh.h1)) != nullptr;
  ok &= (p->dZ2 = p->alloc<__nv_bfloat16>("dz2", B * sh.h2)) != nullptr;
  ok &= (p->dZ3 = p->alloc<__nv_bfloat16>("dz3", B * sh.h3)) != nullptr;
  p->ldm1 = mask_ld(sh.h1); p->ldm2 = mask_ld(sh.h2); p->ldm3 = mask_ld(sh.h3);
  ok &= (p->m1 = p->alloc<uint32_t>("mask1", B * p->ldm1)) != nullptr;
  ok &= (p->m2 = p->alloc<uint32_t>("mask2", B * p->ldm2)) != nullptr;
  ok &= (p->m3 = p->alloc<uint32_t>("mask3", B * p->ldm3)) != nullptr;
  ok &= (p->row_src = p->alloc<unsigned long long>("row_src", sh.cap)) != nullptr;
  ok &= (p->wide_src = p->alloc<unsigned long long>("wide_src", sh.cap)) != nullptr;
  ok &= (p->wide_vals_par[0] = p->alloc<float>("wide_vals", sh.cap)) != nullptr;
  p->wide_vals_par[1] = p->wide_vals_par[0];
  if (!cfg->sync) ok &= (p->wide_vals_par[1] = p->alloc<float>("wide_vals1", sh.cap)) != nullptr;
  ok &= (p->row_dst = p->alloc<unsigned long long>("row_dst", sh.cap)) != nullptr;
  ok &= (p->dlogit = p->alloc<float>("dlogit", B)) != nullptr;
  ok &= (p->prob = p->alloc<float>("prob", B)) != nullptr;
  const int m_tiles = (sh.B + BM - 1) / BM;
  p->zero_bytes = (kMaxWorld + 4 + 4 + m_tiles * kFlagWords * 2) * 4;   // the allocation below is padded to 16-byte multiples
  ok &= (p->zero_par[0] = p->alloc<unsigned int>("step_scratch", kMaxWorld + 8 + m_tiles * kFlagWords * 2 + 8)) != nullptr;
  if (cfg->sync) p->zero_par[1] = p->zero_par[0];
  else ok &= (p->zero_par[1] = p->alloc<unsigned int>("step_scratch1", kMaxWorld + 8 + m_tiles * kFlagWords * 2 + 8)) != nullptr;
  ok &= (p->state = p->alloc<StepState>("step_state", 1)) != nullptr;
  if (getenv("DTF_GEMM_DBG") != nullptr && getenv("DTF_GEMM_DBG")[0] == '1') {
    ok &= (p->gemm_dbg = p->alloc<unsigned long long>("gemm_dbg", 128 + 96 * 4)) != nullptr;
    if (p->gemm_dbg) cudaMemset(p->gemm_dbg, 0, (128 + 96 * 4) * 8);
  }
  if (getenv("DTF_STAMPS") != nullptr && getenv("DTF_STAMPS")[0] == '1') {
    ok &= (p->stamps = p->alloc<unsigned long long>("stamps", kStampCap)) != nullptr;
    p->stamp_names.assign(kStampCap, std::string());
  }
  if (!ok) {
    dtf_wd_destroy(p);
    return nullptr;
  }
  select_parity(p, 0);
  for (int i = 0; i < 2; ++i) {
    cudaEventCreateWithFlags(&p->ev_push[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&p->ev_apply[i], cudaEventDisableTiming);
    if (i == 0) {
      int lo = 0, hi = 0;
      cudaDeviceGetStreamPriorityRange(&lo, &hi);
      cudaStreamCreateWithPriority(&p->side, cudaStreamNonBlocking, hi);
      // The prefetch stream has the priority of the compute stream: at low priority its blocks were only placed
      // once the compute stream's chain of back-to-back (programmatic) launches had a gap, ~200 us into the
      // step, where the gather then shared NVLink with the push (DTF_STAMPS timeline, 2 GPUs).
      const char* pe = getenv("DTF_PF_PRIO");
      cudaStreamCreateWithPriority(&p->pf_stream, cudaStreamNonBlocking, (pe != nullptr && pe[0] == '0') ? lo : hi);
      cudaEventCreateWithFlags(&p->ev_cstart, cudaEventDisableTiming);
      cudaEventCreateWithFlags(&p->ev_pf[0], cudaEventDisableTiming);
      cudaEventCreateWithFlags(&p->ev_pf[1], cudaEventDisableTiming);
      cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming);
      cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming);
      // measured (bench.py, async): 2 GPUs 0.424 -> 0.409 ms with 48 SMs, 8 GPUs 0.473 -> 0.449 ms with 56;
      // 1 GPU, with dW3 / dW2 moved next to the push: 0.379 -> 0.370 ms with 56, 0.375 with 40
      const char* e = getenv("DTF_SPLIT_PUSH");
      p->push_ctas = e ? atoi(e) : 56;
    }
  }
  return p;
}

void dtf_wd_destroy(dtf_wd_plan* p) {
  if (!p) return;
  cudaDeviceSynchronize();
  for (int i = 0; i < 2; ++i) {
    if (p->ev_push[i]) cudaEventDestroy(p->ev_push[i]);
    if (p->ev_apply[i]) cudaEventDestroy(p->ev_apply[i]);
  }
  if (p->ev_fork) cudaEventDestroy(p->ev_fork);
  if (p->ev_join) cudaEventDestroy(p->ev_join);
  if (p->side) cudaStreamDestroy(p->side);
  if (p->pf_stream) cudaStreamDestroy(p->pf_stream);
  if (p->ev_cstart) cudaEventDestroy(p->ev_cstart);
  for (int i = 0; i < 2; ++i)
    if (p->ev_pf[i]) cudaEventDestroy(p->ev_pf[i]);
  for (void* m : p->opened) dtf_symm_close_handle(m);
  for (void* q : p->owned) cudaFree(q);
  if (p->slab && p->owns_slab) dtf_symm_free(p->slab);
  (void)cudaGetLastError();
  delete p;
}

int dtf_wd_buffer(dtf_wd_plan* p, const char* name, void** ptr, size_t* bytes) {
  auto it = p->bufs.find(name);
  if (it == p->bufs.end()) {
    set_error("wd plan: no buffer named '%s'", name);
    return -2;
  }
  *ptr = it->second.ptr;
  *bytes = it->second.bytes;
  return 0;
}

int dtf_wd_info(dtf_wd_plan* p, dtf_wd_info_t* out) {
  memset(out, 0, sizeof(*out));
  out->K0 = p->sh.K0;
  out->rows_local = p->sh.rows_local;
  out->P = p->dl.P;
  out->P_pad = p->P_pad;
  out->dense_lo = p->dense_lo;
  out->dense_hi = p->dense_hi;
  out->off_w1 = p->dl.w1; out->off_w2 = p->dl.w2; out->off_w3 = p->dl.w3;
  out->off_b1 = p->dl.b1; out->off_b2 = p->dl.b2; out->off_b3 = p->dl.b3;
  out->off_w4 = p->dl.w4; out->off_b4 = p->dl.b4; out->off_wide_b = p->dl.wide_b;
  out->small_begin = p->dl.small_begin;
  out->slab_bytes = static_cast<long long>(p->sl.total);
  out->step = p->step;
  return 0;
}

// Hand the plan a multicast mapping of the first `bytes` of every rank's slab (dtf_mc_bind_map).  The
// dense-parameter apply then reduces the replicas' gradients with multimem.ld_reduce and broadcasts
// the new weights with multimem.st.  Refused when the regions it would address lie beyond `bytes`.
int dtf_wd_attach_multicast(dtf_wd_plan* p, void* mc_ptr, long long bytes) {
  if (mc_ptr == nullptr) {
    p->mc = nullptr;
    return 0;
  }
  for (int par = 0; par < 2; ++par) {
    const SlabLayout& sl = p->sl_par[par];
    const long long need_grad = static_cast<long long>(sl.dense_grad) + p->P_pad * 4;
    const long long need_w = static_cast<long long>(sl.wbf16) + p->P_pad * 2;
    const long long need_small = static_cast<long long>(sl.small_f32) + (p->P_pad - p->dl.small_begin) * 4;
    if (need_grad > bytes || need_w > bytes || need_small > bytes) {
      set_error("dtf_wd_attach_multicast: mapping of %lld bytes does not cover the dense regions", bytes);
      return -2;
    }
  }
  p->mc = static_cast<uint8_t*>(mc_ptr);
  return 0;
}

int dtf_wd_slab_handle(dtf_wd_plan* p, void* handle_out) { return dtf_symm_get_handle(p->slab, handle_out); }
void* dtf_wd_slab_ptr(dtf_wd_plan* p) { return p->slab; }

// handles: world * dtf_symm_handle_size() bytes (own entry ignored)
int dtf_wd_attach_ipc(dtf_wd_plan* p, const void* handles) {
  const int hs = dtf_symm_handle_size();
  for (int r = 0; r < p->cfg.world; ++r) {
    if (r == p->cfg.rank) continue;
    void* m = nullptr;
    if (dtf_symm_open_handle(static_cast<const uint8_t*>(handles) + static_cast<size_t>(r) * hs, &m) != 0) return -1;
    p->opened.push_back(m);
    p->pt.slab[r] = static_cast<uint8_t*>(m);
  }
  return 0;
}

// peers' slabs received as POSIX file descriptors (VMM allocations, 2 MB pages); own entry ignored
int dtf_wd_attach_fds(dtf_wd_plan* p, const int* fds) {
  const long long size = dtf_symm_mapped_size(p->slab);
  for (int r = 0; r < p->cfg.world; ++r) {
    if (r == p->cfg.rank) continue;
    void* m = nullptr;
    if (dtf_symm_import_fd(fds[r], static_cast<size_t>(size), &m) != 0) return -1;
    p->opened.push_back(m);
    p->pt.slab[r] = static_cast<uint8_t*>(m);
  }
  return 0;
}
int dtf_wd_slab_fd(dtf_wd_plan* p, int* fd_out) { return dtf_symm_export_fd(p->slab, fd_out); }
int dtf_wd_slab_is_vmm(dtf_wd_plan* p) { return dtf_symm_is_vmm(p->slab); }

// same-process peers: raw slab pointers of the other plans (tests, single-process multi-GPU)
int dtf_wd_attach_ptrs(dtf_wd_plan* p, void* const* slabs) {
  for (int r = 0; r < p->cfg.world; ++r)
    if (r != p->cfg.rank) p->pt.slab[r] = static_cast<uint8_t*>(slabs[r]);
  return 0;
}

int dtf_wd_init_params(dtf_wd_plan* p, unsigned long long seed, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  const float slot1_init = opt_slot1_init(p->cfg.optimizer);
  wd_init_embedding_kernel<<<num_sms() * 8, 256, 0, st>>>(
      p->emb_master, reinterpret_cast<__nv_bfloat16*>(p->slab + p->sl.emb_bf16), p->emb_s1, p->emb_s2,
      reinterpret_cast<float*>(p->slab + p->sl.wide), p->wide_s1, p->wide_s2, p->sh.rows_local, p->cfg.world,
      p->cfg.rank, seed, p->cfg.emb_init_scale, slot1_init);
  wd_init_dense_kernel<<<num_sms() * 4, 256, 0, st>>>(p->dense_master, p->dense_s1, p->dense_s2, p->dense_lo,
                                                      p->dense_hi, p->dl, p->sh, seed, slot1_init);
  DTF_CUDA_OK(cudaGetLastError());
  count_launch(2);
  return dtf_wd_broadcast_dense(p, stream_);
}

// push this rank's dense master range into every peer's working copy (needs peers attached)
int dtf_wd_broadcast_dense(dtf_wd_plan* p, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  wd_dense_broadcast_kernel<<<num_sms() * 4, 256, 0, st>>>(p->dense_master, p->dense_lo, p->dense_hi, p->pt, p->sl,
                                                           p->dl.small_begin);
  DTF_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// refresh the bf16 working copy of the embedding shard from the fp32 master (after a restore)
__global__ void wd_refresh_emb_kernel(const float* __restrict__ master, __nv_bfloat16* __restrict__ work, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    work[i] = __float2bfloat16(master[i]);
}
int dtf_wd_refresh_embedding(dtf_wd_plan* p, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  wd_refresh_emb_kernel<<<num_sms() * 8, 256, 0, st>>>(
      p->emb_master, reinterpret_cast<__nv_bfloat16*>(p->slab + p->sl.emb_bf16), p->sh.rows_local * kEmbDim);
  DTF_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------ step
// Which GEMMs of the step run on CTA pairs (cta_group::2): bit 0 fused pull, 1 layers 2-3, 2 weight
// gradients, 3 data gradients, 4 fused push.  DTF_PAIR_MASK overrides the default.
static int pair_mask() {
  static int m = -1;
  if (m < 0) {
    const char* e = getenv("DTF_PAIR_MASK");
    m = e ? atoi(e) : (1 | 4 | 16);   // measured: pull 74 -> 70 us, wgrad-1 60 -> 54, push 49 -> 46; the small GEMMs do not gain
  }
  return m;
}
static void set_pair(GemmArgs& a, int bit) { a.pair = ((pair_mask() >> bit) & 1) && pair_eligible(a, a.bn); }

static GemmArgs base_args(cudaStream_t st) {
  GemmArgs a{};
  a.stream = st;
  a.splits = 1;
  a.pdl = true;
  return a;
}

// Parameter prefetch of async mode (dtf_wd_set_next_batch): addresses + row gather of the next batch on the
// low-priority stream, into the operand buffer and wide values of the parity that step will use.  Forked from
// `st` here (everything the previous user of those buffers launched on `st` is ordered before) and joined by the
// event the next step's layer 1 waits for.
static int pf_blocks_per_sm() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DTF_PF_BLOCKS");
    v = e ? atoi(e) : 4;   // 2 GPUs: 0.415 ms/step without prefetch, 0.407 with 2 blocks per SM, 0.398 with 4 or 6
    if (v < 1) v = 1;
  }
  return v;
}

static int launch_prefetch(dtf_wd_plan* p, cudaStream_t st) {
  const WdShapes& sh = p->sh;
  const int n = sh.B * sh.F;
  const int np = p->cur_par ^ 1;
  DTF_CUDA_OK(cudaEventRecord(p->ev_cstart, st));
  DTF_CUDA_OK(cudaStreamWaitEvent(p->pf_stream, p->ev_cstart, 0));
  const int work = n > sh.B * sh.dense_pad ? n : sh.B * sh.dense_pad;
  mark(p, p->pf_stream, "pf:start");
  wd_prefetch_addr_kernel<<<(work + 255) / 256, 256, 0, p->pf_stream>>>(
      p->next_ids, n, sh.V, p->pt, p->sl, p->row_src_pf, p->wide_src_pf, p->next_dense, p->a_buf_par[np],
      static_cast<long long>(sh.K0), sh.F * kEmbDim, sh.B, sh.dense_in, sh.dense_pad);
  DTF_CUDA_OK(cudaGetLastError());
  count_launch();
  wd_gather_rows_kernel<<<num_sms() * pf_blocks_per_sm(), 256, 0, p->pf_stream>>>(
      p->row_src_pf, p->wide_src_pf, p->a_buf_par[np], static_cast<long long>(sh.K0), p->wide_vals_par[np], sh.B, sh.F);
  DTF_CUDA_OK(cudaGetLastError());
  count_launch();
  mark(p, p->pf_stream, "pf:gathered");
  DTF_CUDA_OK(cudaEventRecord(p->ev_pf[np], p->pf_stream));
  p->pf_valid[np] = true;
  p->pf_ids[np] = p->next_ids;
  p->next_ids = nullptr;
  p->next_dense = nullptr;
  return 0;
}

// debug ids of the step's GEMMs: 1 fused pull, 10 layer 1 on prefetched rows, 2-3 layers 2-3, 4 wgrad-3, 5 dgrad-3,
// 6 wgrad-2, 7 dgrad-2, 8 push, 9 wgrad-1
static void tag(dtf_wd_plan* p, GemmArgs& a, int id) {
  a.dbg = p->gemm_dbg;
  a.dbg_id = id;
}

// Where in the step the prefetch of the next batch is forked (DTF_PF_POINT): 0 after the route kernel, 1 after
// layer 1 has been launched, 2 after the loss kernel.
static int pf_point() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DTF_PF_POINT");
    v = e ? atoi(e) : 0;
  }
  return v;
}
static int maybe_prefetch(dtf_wd_plan* p, cudaStream_t st, int point) {
  if (point != pf_point() || p->next_ids == nullptr || p->pf_stream == nullptr || p->cfg.sync) return 0;
  return launch_prefetch(p, st);
}

static int forward(dtf_wd_plan* p, const int* ids, const float* dense_raw, cudaStream_t st, bool train,
                   bool scratch_zeroed = false) {
  const WdShapes& sh = p->sh;
  const DenseLayout& dl = p->dl;
  const int n = sh.B * sh.F;
  mark(p, st, "begin");
  if (!scratch_zeroed && zero_fill(p->zero_region, p->zero_bytes, st) != 0) return -1;
  DTF_CUDA_OK(launch_kernel(true, wd_route_kernel, dim3((n + 255) / 256), dim3(256), 0, st, ids, n, sh.V, p->pt, p->sl,
                            sh.cap, p->cnt, p->row_src, p->wide_src, p->row_dst, dense_raw, p->dense_x, sh.B,
                            sh.dense_in, sh.dense_pad));
  count_launch();
  mark(p, st, "route+pad");
  const __nv_bfloat16* W = reinterpret_cast<const __nv_bfloat16*>(p->slab + p->sl.wbf16);
  const float* small = reinterpret_cast<const float*>(p->slab + p->sl.small_f32);
  // async mode: start fetching the NEXT step's rows into the other parity's operand buffer
  if (train && maybe_prefetch(p, st, 0) != 0) return -1;
  const int par = p->cur_par;
  const bool prefetched = train && p->pf_valid[par] && p->pf_ids[par] == ids;
  if (p->pf_valid[par]) {
    // whether the rows are used or not (a caller that announced one batch and stepped on another), nothing may
    // still be writing this parity's buffers when the step touches them
    DTF_CUDA_OK(cudaStreamWaitEvent(st, p->ev_pf[par], 0));
    p->pf_valid[par] = false;
  }
  if (p->stamps != nullptr) mark(p, st, prefetched ? "gemm1 may start (prefetched)" : "pull may start");
  GemmArgs a = base_args(st);
  a.A = p->a_buf; a.lda = sh.K0; a.B = W + dl.w1; a.ldb = sh.K0; a.D = p->H1; a.ldd = sh.h1;
  a.M = sh.B; a.N = sh.h1; a.K = sh.K0; a.epi = EPI_BIAS_RELU_BF16; a.bias = small + (dl.b1 - dl.small_begin);
  a.mask_out = train ? p->m1 : nullptr; a.ldmask = p->ldm1;
  a.bn = 256;
  set_pair(a, 0);
  if (prefetched) {
    // layer 1 on rows a low-priority stream gathered during the previous step: a plain GEMM
    a.pdl = false;                    // its operand comes from another stream: no early start past the event wait
    tag(p, a, 10);
    if (launch_gemm_kk(a) != 0) return -1;
    p->n_prefetched += 1;
  } else {
    // layer 1: embedding pull fused with the GEMM
    a.gather = true; a.row_src = p->row_src; a.dense_x = p->dense_x; a.slots = sh.F; a.dense_cols = sh.dense_pad;
    a.tile_flags = p->tile_flags; a.chunk_counter = p->chunk_counter;
    // the gather warps also fetch the wide weights of the batch's ids (a separate gather kernel on the side
    // stream next to layers 2-3 measured the same at 2 GPUs and 6 us slower at 1 GPU: run90)
    a.wide_src = p->wide_src; a.wide_out = p->wide_vals;
    tag(p, a, 1);
    if (launch_gemm_gather(a) != 0) return -1;
  }
  mark(p, st, "pull+gemm1");
  if (train && maybe_prefetch(p, st, 1) != 0) return -1;
  // layers 2, 3
  a = base_args(st);
  a.A = p->H1; a.lda = sh.h1; a.B = W + dl.w2; a.ldb = sh.h1; a.D = p->H2; a.ldd = sh.h2;
  a.M = sh.B; a.N = sh.h2; a.K = sh.h1; a.epi = EPI_BIAS_RELU_BF16; a.bias = small + (dl.b2 - dl.small_begin);
  a.mask_out = train ? p->m2 : nullptr; a.ldmask = p->ldm2;
  a.bn = choose_bn(a.M, a.N, 1);
  set_pair(a, 1);
  tag(p, a, 2);
  if (launch_gemm_kk(a) != 0) return -1;
  a = base_args(st);
  a.A = p->H2; a.lda = sh.h2; a.B = W + dl.w3; a.ldb = sh.h2; a.D = p->H3; a.ldd = sh.h3;
  a.M = sh.B; a.N = sh.h3; a.K = sh.h2; a.epi = EPI_BIAS_RELU_BF16; a.bias = small + (dl.b3 - dl.small_begin);
  a.mask_out = train ? p->m3 : nullptr; a.ldmask = p->ldm3;
  a.bn = choose_bn(a.M, a.N, 1);
  set_pair(a, 1);
  tag(p, a, 3);
  if (launch_gemm_kk(a) != 0) return -1;
  mark(p, st, "gemm2+gemm3");
  return 0;
}

static int launch_loss(dtf_wd_plan* p, const float* labels, cudaStream_t st, bool train) {
  const WdShapes& sh = p->sh;
  const DenseLayout& dl = p->dl;
  const float* small = reinterpret_cast<const float*>(p->slab + p->sl.small_f32);
  float* g_small = reinterpret_cast<float*>(p->slab + p->sl.dense_grad) + dl.small_begin;
  const int off_w4 = static_cast<int>(dl.w4 - dl.small_begin), off_b4 = static_cast<int>(dl.b4 - dl.small_begin),
            off_wb = static_cast<int>(dl.wide_b - dl.small_begin);
  const float inv_batch = 1.0f / (static_cast<float>(sh.B) * (p->cfg.sync ? p->cfg.world : 1));
  // 48 registers since the register arrays are sized by H3 (126 before): four blocks per SM (1 GPU: 0.3599 -> 0.3575 ms)
  static const int loss_mult = getenv("DTF_LOSS_GRID") ? atoi(getenv("DTF_LOSS_GRID")) : 4;
  int grid = num_sms() * loss_mult;
  if (grid > (sh.B + 7) / 8) grid = (sh.B + 7) / 8;
  const int maxc = sh.h3 <= 256 ? 1 : (sh.h3 <= 512 ? 2 : 4);
  if (train) {
    const size_t smem = 8 * (sh.h3 + 2) * sizeof(float);
    auto go = [&](auto kern) {
      return launch_kernel(true, kern, dim3(grid), dim3(256), smem, st, p->H3, p->m3, static_cast<long long>(p->ldm3), small,
                           off_w4, off_b4, off_wb, p->wide_vals, p->pt, p->sl, labels, sh.B, sh.F, sh.h3, inv_batch, p->dlogit,
                           p->dZ3, g_small, p->loss_acc, static_cast<float*>(nullptr));
    };
    DTF_CUDA_OK(maxc == 1 ? go(wd_loss_kernel<true, 1>) : (maxc == 2 ? go(wd_loss_kernel<true, 2>) : go(wd_loss_kernel<true, 4>)));
  } else {
    auto go = [&](auto kern) {
      kern<<<grid, 256, 0, st>>>(p->H3, nullptr, 0, small, off_w4, off_b4, off_wb, p->wide_vals, p->pt, p->sl, nullptr, sh.B,
                                 sh.F, sh.h3, 0.f, nullptr, nullptr, nullptr, nullptr, p->prob);
    };
    if (maxc == 1) go(wd_loss_kernel<false, 1>);
    else if (maxc == 2) go(wd_loss_kernel<false, 2>);
    else go(wd_loss_kernel<false, 4>);
  }
  DTF_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

static int wgrad(dtf_wd_plan* p, const __nv_bfloat16* dy, int n_out, const __nv_bfloat16* x, int k_in, long long off,
                 cudaStream_t st, int max_ctas = 0) {
  // dW[n_out, k_in] += dy^T x, reduction over the batch; partial tiles are added by TMA reduce
  GemmArgs a = base_args(st);
  a.A = dy; a.lda = n_out; a.a_mn = true; a.B = x; a.ldb = k_in; a.b_mn = true;
  a.D = reinterpret_cast<float*>(p->slab + p->sl.dense_grad) + off; a.ldd = k_in;
  a.M = n_out; a.N = k_in; a.K = p->sh.B; a.epi = EPI_F32_ATOMIC;
  a.bn = 256;
  set_pair(a, 2);
  const int ctas = a.pair ? 2 : 1;
  const int tiles = ((n_out + BM * ctas - 1) / (BM * ctas)) * ((k_in + 255) / 256);
  a.max_ctas = max_ctas;
  const int sms = (max_ctas > 0 && max_ctas < num_sms()) ? max_ctas : num_sms();
  // split-K factor: fewest k-blocks on the critical path, counting whole waves of the persistent grid
  // plus ~2 k-blocks' worth of epilogue (fp32 TMA reduce-add of the partial tile) per wave
  const int units = sms / ctas > 0 ? sms / ctas : 1;
  const int kb = (p->sh.B + BK - 1) / BK;
  int splits = 1;
  long long best = -1;
  for (int sp = 1; sp <= 16 && sp <= kb; ++sp) {
    const long long waves = (static_cast<long long>(tiles) * sp + units - 1) / units;
    const long long cost = waves * ((kb + sp - 1) / sp + 2);
    if (best < 0 || cost < best) {
      best = cost;
      splits = sp;
    }
  }
  a.splits = splits;
  tag(p, a, p->dbg_id);
  return launch_gemm_mnmn(a);
}

static int dgrad(dtf_wd_plan* p, const __nv_bfloat16* dy, int n, const __nv_bfloat16* w, int k, const uint32_t* mask,
                 int ldmask, __nv_bfloat16* out, cudaStream_t st) {
  GemmArgs a = base_args(st);
  a.A = dy; a.lda = n; a.B = w; a.ldb = k; a.b_mn = true; a.D = out; a.ldd = k;
  a.M = p->sh.B; a.N = k; a.K = n; a.epi = EPI_DRELU_BF16; a.mask_in = mask; a.ldmask = ldmask;
  a.bn = choose_bn(a.M, a.N, 1);
  set_pair(a, 3);
  tag(p, a, p->dbg_id);
  return launch_gemm_kmn(a);
}

// bias gradients of the three hidden layers in one launch
static int colsum3(const __nv_bfloat16* x0, int n0, float* o0, const __nv_bfloat16* x1, int n1, float* o1,
                   const __nv_bfloat16* x2, int n2, float* o2, int M, cudaStream_t st) {
  ColsumJobs j{};
  j.M = M;
  j.rows_per_block = 64;
  const int row_blocks = (M + j.rows_per_block - 1) / j.rows_per_block;
  const __nv_bfloat16* xs[3] = {x0, x1, x2};
  float* os[3] = {o0, o1, o2};
  const int ns[3] = {n0, n1, n2};
  int total = 0;
  for (int i = 0; i < 3; ++i) {
    j.x[i] = xs[i]; j.out[i] = os[i]; j.N[i] = ns[i];
    total += ((ns[i] + 255) / 256) * row_blocks;
    j.block_end[i] = total;
  }
  DTF_CUDA_OK(launch_kernel(true, wd_colsum_kernel, dim3(total), dim3(256), 0, st, j));
  count_launch();
  return 0;
}

// which == 1: tick + sparse (embedding + wide) apply; which == 2: dense apply; 3: both
template <int OPT>
static int apply_t(dtf_wd_plan* p, cudaStream_t st, int which = 3) {
  const SlabLayout& sl = p->sl;
  const StepState* oc = p->state;
  if (which & 1) {
    MailboxView mv{};
    mv.payload = reinterpret_cast<const __nv_bfloat16*>(p->slab + sl.mbx_payload);
    mv.rows = reinterpret_cast<const long long*>(p->slab + sl.mbx_rows);
    mv.dlvec = reinterpret_cast<const float*>(p->slab + sl.mbx_dlvec);
    mv.B = p->sh.B;
    mv.count = reinterpret_cast<const unsigned int*>(p->slab + sl.mbx_count);
    mv.cap = p->sh.cap;
    mv.world = p->cfg.world;
    wd_tick_kernel<<<1, 1, 0, st>>>(p->state, p->cfg.lr, p->cfg.optimizer);
    count_launch();
    __nv_bfloat16* work = reinterpret_cast<__nv_bfloat16*>(p->slab + sl.emb_bf16);
    float* wide = reinterpret_cast<float*>(p->slab + sl.wide);
    int grid = num_sms() * 8;
    if (p->apply_blocks_per_sm != 8) {
      // overlapped with the next step's compute: many short-lived blocks (one pair of entries per warp)
      // instead of a persistent grid, so that SMs are handed to the higher-priority GEMM CTAs as soon
      // as those become runnable and the apply fills whatever room they leave
      const long long want = (p->sh.cap + 15) / 16;
      grid = static_cast<int>(want > grid ? want : grid);
    }
    if (p->cfg.sync) {
      wd_sparse_count_kernel<<<num_sms() * 4, 256, 0, st>>>(mv, p->row_state);
      mark(p, st, "sparse count");
      wd_sparse_apply_sync_kernel<OPT><<<grid, 256, 0, st>>>(mv, p->row_state, p->G, p->wide_G, p->emb_master, p->emb_s1,
                                                             p->emb_s2, work, wide, p->wide_s1, p->wide_s2, oc);
      count_launch(2);
      mark(p, st, "sparse apply (sync)");
    } else {
      // DTF_APPLY_STREAM=1: optimizer state and mailbox entries are read / written with evict-first hints.  Tried
      // against the idea that the overlapped apply pushes the GEMM operands out of L2: 0.3675 -> 0.3657 ms per step
      // at one GPU, i.e. within noise, so L2 pollution is not what the overlap costs; left off.
      static const bool stream_hint = getenv("DTF_APPLY_STREAM") != nullptr && getenv("DTF_APPLY_STREAM")[0] == '1';
      if (stream_hint)
        wd_sparse_apply_entries_kernel<OPT, true><<<grid, 256, 0, st>>>(mv, p->emb_master, p->emb_s1, p->emb_s2, work, wide,
                                                                        p->wide_s1, p->wide_s2, oc);
      else
        wd_sparse_apply_entries_kernel<OPT, false><<<grid, 256, 0, st>>>(mv, p->emb_master, p->emb_s1, p->emb_s2, work, wide,
                                                                         p->wide_s1, p->wide_s2, oc);
      count_launch();
      mark(p, st, "sparse apply entries");
    }
  }
  if (which & 2) {
    const long long dense_n4 = (p->dense_hi - p->dense_lo) / 4;
    // replicas held in registers: none beyond one when the NVLS sum is used; U float4 groups per thread
    const bool mc_sum = p->mc != nullptr && p->cfg.sync;
    const int world = p->cfg.world;
    auto launch = [&](auto kern, int u) {
      const long long per_block = 256LL * u;
      const int dense_grid = p->apply_blocks_per_sm != 8 ? static_cast<int>((dense_n4 + per_block - 1) / per_block)
                                                         : num_sms() * 4;
      kern<<<dense_grid > 0 ? dense_grid : 1, 256, 0, st>>>(p->pt, sl, p->dense_lo, p->dense_hi, p->cfg.sync ? 1 : 0,
                                                             p->dense_master, p->dense_s1, p->dense_s2, oc,
                                                             p->dl.small_begin, p->mc);
    };
    if (mc_sum || world == 1) launch(wd_dense_apply_kernel<OPT, 1, 4>, 4);
    else if (world == 2) launch(wd_dense_apply_kernel<OPT, 2, 4>, 4);
    else if (world <= 4) launch(wd_dense_apply_kernel<OPT, 4, 2>, 2);
    else if (world <= 8) launch(wd_dense_apply_kernel<OPT, 8, 1>, 1);
    else launch(wd_dense_apply_kernel<OPT, kMaxWorld, 1>, 1);
    count_launch();
    mark(p, st, "dense apply");
  }
  DTF_CUDA_OK(cudaGetLastError());
  return 0;
}

static int apply_dispatch(dtf_wd_plan* p, cudaStream_t st, int which) {
  NvtxRange nvtx_range(which == 1 ? "wd_apply_sparse" : (which == 2 ? "wd_apply_dense" : "wd_apply"));
  switch (p->cfg.optimizer) {
    case WD_SGD: return apply_t<WD_SGD>(p, st, which);
    case WD_ADAGRAD: return apply_t<WD_ADAGRAD>(p, st, which);
    case WD_ADAM: return apply_t<WD_ADAM>(p, st, which);
    case WD_ADADELTA: return apply_t<WD_ADADELTA>(p, st, which);
    case WD_FTRL: return apply_t<WD_FTRL>(p, st, which);
    default: return apply_t<WD_RMSPROP>(p, st, which);
  }
}

extern "C" {

// Tracing: with profiling on, every executor call records CUDA events between its launches;
// dtf_wd_profile_read synchronises, writes "name=microseconds" lines (time since the previous mark)
// into buf and clears the marks.  Not for use under CUDA-graph capture.
void dtf_wd_profile_enable(dtf_wd_plan* p, int on) { p->profile = on != 0; }
int dtf_wd_profile_read(dtf_wd_plan* p, char* buf, size_t cap) {
  cudaDeviceSynchronize();
  std::string out;
  for (size_t i = 1; i < p->marks.size(); ++i) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, p->marks[i - 1].second, p->marks[i].second);
    char line[128];
    snprintf(line, sizeof(line), "%s=%.1f\n", p->marks[i].first.c_str(), ms * 1e3f);
    out += line;
  }
  for (auto& m : p->marks) cudaEventDestroy(m.second);
  p->marks.clear();
  if (buf && cap) {
    const size_t n = out.size() < cap - 1 ? out.size() : cap - 1;
    memcpy(buf, out.data(), n);
    buf[n] = 0;
  }
  return static_cast<int>(out.size());
}

// Same marks as absolute offsets (microseconds since the first mark): lets marks recorded on the
// compute stream and on the apply stream be laid on one time line.
int dtf_wd_profile_read_abs(dtf_wd_plan* p, char* buf, size_t cap) {
  cudaDeviceSynchronize();
  std::string out;
  for (size_t i = 0; i < p->marks.size(); ++i) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, p->marks[0].second, p->marks[i].second);
    char line[128];
    snprintf(line, sizeof(line), "%s=%.1f\n", p->marks[i].first.c_str(), ms * 1e3f);
    out += line;
  }
  for (auto& m : p->marks) cudaEventDestroy(m.second);
  p->marks.clear();
  if (buf && cap) {
    const size_t n = out.size() < cap - 1 ? out.size() : cap - 1;
    memcpy(buf, out.data(), n);
    buf[n] = 0;
  }
  return static_cast<int>(out.size());
}

// forward + backward + push of one micro-batch; device pointers: ids int32 [B*F], dense f32
// [B*dense_in], labels f32 [B]
static int weight_grads(dtf_wd_plan* p, cudaStream_t st, int wgrad1_ctas, bool with_colsum = true);
static int weight_grads_deferred(dtf_wd_plan* p, cudaStream_t st, int wgrad1_ctas = 0, bool with_colsum = true);

// dW3 and dW2 are not on the way to the push: by default they run after the push has been forked to the side
// stream, next to it, instead of between the data-gradient GEMMs (DTF_LATE_WGRAD=0 restores the old order).
static bool late_wgrad() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DTF_LATE_WGRAD");
    v = (e == nullptr || e[0] != '0') ? 1 : 0;
  }
  return v == 1;
}

// defer_wgrad: leave out the weight / bias gradient kernels (the caller runs weight_grads() on another
// stream) and keep only the chain the sparse push depends on
static int compute_impl(dtf_wd_plan* p, const int* ids, const float* dense_raw, const float* labels, cudaStream_t st,
                        bool defer_wgrad) {
  NvtxRange nvtx_range("wd_step_compute");
  const WdShapes& sh = p->sh;
  const DenseLayout& dl = p->dl;
  const int par = p->cfg.sync ? 0 : static_cast<int>(p->n_compute & 1);
  select_parity(p, par);
  p->last_compute_parity = par;
  p->n_compute += 1;
  mark(p, st, "step");
  // one launch clears the dense-gradient buffer and the per-step scratch (counters, loss, arrival flags)
  if (zero_fill(p->slab + p->sl.dense_grad, static_cast<size_t>(p->P_pad) * 4, st, p->zero_region, p->zero_bytes) != 0)
    return -1;
  if (forward(p, ids, dense_raw, st, true, true) != 0) return -1;
  if (launch_loss(p, labels, st, true) != 0) return -1;
  mark(p, st, "loss+wide");
  if (maybe_prefetch(p, st, 2) != 0) return -1;
  const __nv_bfloat16* W = reinterpret_cast<const __nv_bfloat16*>(p->slab + p->sl.wbf16);
  float* grad = reinterpret_cast<float*>(p->slab + p->sl.dense_grad);
  const bool split = p->push_ctas > 0 && p->push_ctas < num_sms() && !p->profile;
  const bool late = !defer_wgrad && split && late_wgrad();
  if (!defer_wgrad && !late) {
    p->dbg_id = 4;
    if (wgrad(p, p->dZ3, sh.h3, p->H2, sh.h2, dl.w3, st) != 0) return -1;
    mark(p, st, "wgrad3");
  }
  p->dbg_id = 5;
  if (dgrad(p, p->dZ3, sh.h3, W + dl.w3, sh.h2, p->m2, p->ldm2, p->dZ2, st) != 0) return -1;
  mark(p, st, "dgrad3");
  if (!defer_wgrad && !late) {
    p->dbg_id = 6;
    if (wgrad(p, p->dZ2, sh.h2, p->H1, sh.h1, dl.w2, st) != 0) return -1;
    mark(p, st, "wgrad2");
  }
  p->dbg_id = 7;
  if (dgrad(p, p->dZ2, sh.h2, W + dl.w2, sh.h1, p->m1, p->ldm1, p->dZ1, st) != 0) return -1;
  mark(p, st, "dgrad2");
  // fused push: dA = dZ1 W1[:, :F*128]; every 128-wide row goes to its owner's mailbox
  GemmArgs a = base_args(st);
  a.A = p->dZ1; a.lda = sh.h1; a.B = W + dl.w1; a.ldb = sh.K0; a.b_mn = true;
  a.M = sh.B; a.N = sh.F * kEmbDim; a.K = sh.h1; a.epi = EPI_PUSH_ROWS; a.row_dst = p->row_dst; a.slots = sh.F;
  a.bn = 256;
  set_pair(a, 4);
  tag(p, a, 8);
  if (defer_wgrad) {
    if (launch_gemm_kmn(a) != 0) return -1;
    mark(p, st, "dgrad1+push");
    return 0;
  }
  if (split) {
    // wgrad-1 (tensor-bound) and the push (NVLink-bound once the shards are remote) only depend on
    // dZ1: run them side by side, the push on `push_ctas` SMs of a second stream
    DTF_CUDA_OK(cudaEventRecord(p->ev_fork, st));
    DTF_CUDA_OK(cudaStreamWaitEvent(p->side, p->ev_fork, 0));
    a.stream = p->side;
    a.max_ctas = p->push_ctas;
    a.pdl = false;
    if (launch_gemm_kmn(a) != 0) return -1;
    mark(p, p->side, "side:push done");
    // the three bias gradients (column sums of dZ1..dZ3, 12 us) follow the push here: the other branch, with the
    // three weight-gradient GEMMs on the remaining SMs, is the longer one
    if (colsum3(p->dZ1, sh.h1, grad + dl.b1, p->dZ2, sh.h2, grad + dl.b2, p->dZ3, sh.h3, grad + dl.b3, sh.B, p->side) != 0) return -1;
    mark(p, p->side, "side:colsum x3");
    DTF_CUDA_OK(cudaEventRecord(p->ev_join, p->side));
  }
  (void)grad;
  if (late) {
    if (weight_grads_deferred(p, st, num_sms() - p->push_ctas, !split) != 0) return -1;
  } else if (weight_grads(p, st, split ? num_sms() - p->push_ctas : 0, !split) != 0) {
    return -1;
  }
  if (split) {
    DTF_CUDA_OK(cudaStreamWaitEvent(st, p->ev_join, 0));
  } else {
    if (launch_gemm_kmn(a) != 0) return -1;
  }
  mark(p, st, "dgrad1+push");
  return 0;
}

// dW1 and the three bias gradients (plus dW2, dW3 when they were deferred)
static int weight_grads_deferred(dtf_wd_plan* p, cudaStream_t st, int wgrad1_ctas, bool with_colsum) {
  const WdShapes& sh = p->sh;
  const DenseLayout& dl = p->dl;
  p->dbg_id = 4;
  if (wgrad(p, p->dZ3, sh.h3, p->H2, sh.h2, dl.w3, st, wgrad1_ctas) != 0) return -1;
  p->dbg_id = 6;
  if (wgrad(p, p->dZ2, sh.h2, p->H1, sh.h1, dl.w2, st, wgrad1_ctas) != 0) return -1;
  mark(p, st, "wgrad3+wgrad2");
  return weight_grads(p, st, wgrad1_ctas, with_colsum);
}

static int weight_grads(dtf_wd_plan* p, cudaStream_t st, int wgrad1_ctas, bool with_colsum) {
  const WdShapes& sh = p->sh;
  const DenseLayout& dl = p->dl;
  float* grad = reinterpret_cast<float*>(p->slab + p->sl.dense_grad);
  p->dbg_id = 9;
  if (wgrad(p, p->dZ1, sh.h1, p->a_buf, sh.K0, dl.w1, st, wgrad1_ctas) != 0) return -1;
  mark(p, st, "wgrad1");
  if (!with_colsum) return 0;
  if (colsum3(p->dZ1, sh.h1, grad + dl.b1, p->dZ2, sh.h2, grad + dl.b2, p->dZ3, sh.h3, grad + dl.b3, sh.B, st) != 0) return -1;
  mark(p, st, "colsum x3");
  return 0;
}

// forward + backward + push of one micro-batch; device pointers: ids int32 [B*F], dense f32
// [B*dense_in], labels f32 [B]
int dtf_wd_step_compute(dtf_wd_plan* p, const int* ids, const float* dense_raw, const float* labels, void* stream_) {
  return compute_impl(p, ids, dense_raw, labels, reinterpret_cast<cudaStream_t>(stream_), false);
}

// How long a device barrier waits for a peer before it records an error (DTF_BARRIER_TIMEOUT_MS, default
// 30 s like --ps_timeout_ms; 0 = wait forever).  dtf_wd_set_barrier_timeout_ms overrides the environment.
static long long g_barrier_timeout_ms = -1;
static unsigned long long barrier_timeout_ns() {
  if (g_barrier_timeout_ms < 0) {
    const char* e = getenv("DTF_BARRIER_TIMEOUT_MS");
    g_barrier_timeout_ms = e ? atoll(e) : 30000;
    if (g_barrier_timeout_ms < 0) g_barrier_timeout_ms = 0;
  }
  return static_cast<unsigned long long>(g_barrier_timeout_ms) * 1000000ull;
}
void dtf_wd_set_barrier_timeout_ms(long long ms) { g_barrier_timeout_ms = ms < 0 ? 0 : ms; }

// 0 = healthy; r + 1 = a device barrier of this plan stopped waiting for rank r (dead or stalled peer).
// Reads the flag with a copy on a private stream, so the training streams are not synchronised.
int dtf_wd_check_error(dtf_wd_plan* p) {
  static thread_local cudaStream_t probe = nullptr;
  if (probe == nullptr && cudaStreamCreateWithFlags(&probe, cudaStreamNonBlocking) != cudaSuccess) {
    (void)cudaGetLastError();
    return -1;
  }
  unsigned int v = 0;
  if (cudaMemcpyAsync(&v, &p->state->error, sizeof(v), cudaMemcpyDeviceToHost, probe) != cudaSuccess ||
      cudaStreamSynchronize(probe) != cudaSuccess) {
    set_error("dtf_wd_check_error: %s", cudaGetErrorString(cudaGetLastError()));
    return -1;
  }
  if (v != 0) set_error("device barrier timed out waiting for rank %u (peer dead or stalled)", v - 1);
  return static_cast<int>(v);
}

// cross-rank barrier on the device (no-op publish for a single rank)
int dtf_wd_barrier(dtf_wd_plan* p, int publish_counts, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  select_parity(p, p->cfg.sync ? 0 : static_cast<int>(p->n_apply & 1));
  if (p->cfg.world == 1 || p->cfg.no_device_barrier) {
    if (publish_counts) {
      wd_publish_local_kernel<<<1, kMaxWorld, 0, st>>>(p->pt, p->sl, p->cnt);
      count_launch();
    }
  } else {
    wd_barrier_kernel<<<1, kMaxWorld, 0, st>>>(p->pt, p->sl, p->state, publish_counts ? p->cnt : nullptr,
                                               barrier_timeout_ns());
    count_launch();
  }
  mark(p, st, "barrier");
  DTF_CUDA_OK(cudaGetLastError());
  return 0;
}

// parameter-server side: apply what the mailboxes and gradient buffers hold
int dtf_wd_step_apply(dtf_wd_plan* p, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  select_parity(p, p->cfg.sync ? 0 : static_cast<int>(p->n_apply & 1));
  const int rc = apply_dispatch(p, st, 3);
  if (rc == 0) {
    p->step += 1;
    p->n_apply += 1;
  }
  return rc;
}

// Sync-replica step.  The sparse push only depends on the dgrad chain, so the weight / bias gradient
// kernels (tensor-bound, ~115 us) are taken off the critical path: they run on the side stream while
// the main stream does barrier + sparse apply (HBM-bound, ~160 us); the dense apply waits for both
// (and, across ranks, for one more barrier).  DTF_SYNC_PIPE=0 restores the serial order.
static bool sync_pipe() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DTF_SYNC_PIPE");
    v = (e == nullptr || e[0] != '0') ? 1 : 0;
  }
  return v == 1;
}

int dtf_wd_step(dtf_wd_plan* p, const int* ids, const float* dense_raw, const float* labels, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  if (p->cfg.sync && sync_pipe() && !p->profile) {
    if (compute_impl(p, ids, dense_raw, labels, st, true) != 0) return -1;
    DTF_CUDA_OK(cudaEventRecord(p->ev_fork, st));
    DTF_CUDA_OK(cudaStreamWaitEvent(p->side, p->ev_fork, 0));
    if (weight_grads_deferred(p, p->side) != 0) return -1;
    DTF_CUDA_OK(cudaEventRecord(p->ev_join, p->side));
    if (dtf_wd_barrier(p, 1, stream_) != 0) return -1;          // every sparse push has landed
    select_parity(p, 0);
    // short-lived apply blocks (as in the async overlap): a persistent apply grid would hold every
    // thread slot until it ends and keep the next weight-gradient kernel's CTAs off the SMs
    p->apply_blocks_per_sm = p->cfg.overlap_apply_blocks > 0 ? p->cfg.overlap_apply_blocks : 2;
    const int rc_sparse = apply_dispatch(p, st, 1);
    p->apply_blocks_per_sm = 8;
    if (rc_sparse != 0) return -1;
    DTF_CUDA_OK(cudaStreamWaitEvent(st, p->ev_join, 0));
    if (p->cfg.world > 1 && dtf_wd_barrier(p, 0, stream_) != 0) return -1;   // every rank's dense gradient is complete
    if (apply_dispatch(p, st, 2) != 0) return -1;
    p->step += 1;
    p->n_apply += 1;
    if (p->cfg.world > 1 && dtf_wd_barrier(p, 0, stream_) != 0) return -1;   // updated shards visible to all
    return 0;
  }
  if (dtf_wd_step_compute(p, ids, dense_raw, labels, stream_) != 0) return -1;
  if (dtf_wd_barrier(p, 1, stream_) != 0) return -1;   // every push and gradient has landed
  if (dtf_wd_step_apply(p, stream_) != 0) return -1;
  if (p->cfg.world > 1 && dtf_wd_barrier(p, 0, stream_) != 0) return -1;  // updated shards visible to all
  return 0;
}

}  // extern "C"

// Forward pass for a handful of examples (n <= 4, serving): addresses, then the three layers as swapped-role
// matrix-vector kernels (wd_small_layer_kernel) -- 5 launches with the logit kernel, no tensor-core tile that would
// be 1-12 % full, no scratch to clear.
template <int MAXM>
static int small_layer(const __nv_bfloat16* X, long long ldx, const unsigned long long* row_src,
                       const unsigned long long* wide_src, int F, const __nv_bfloat16* dense_x, int dense_cols,
                       float* wide_out, const __nv_bfloat16* W, long long ldw, const float* bias, __nv_bfloat16* Y,
                       long long ldy, int M, int N, int K, cudaStream_t st) {
  const size_t smem = static_cast<size_t>(M) * K * 2;
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    DTF_CUDA_OK(cudaFuncSetAttribute(wd_small_layer_kernel<MAXM>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    configured = 200 * 1024;
  }
  int grid = (N + 7) / 8;
  if (grid > num_sms()) grid = num_sms();
  wd_small_layer_kernel<MAXM><<<grid, 256, smem, st>>>(X, ldx, row_src, wide_src, F, dense_x, dense_cols, wide_out, W, ldw,
                                                      bias, Y, ldy, M, N, K, 1);
  DTF_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

static int forward_small(dtf_wd_plan* p, const int* ids, const float* dense_raw, int n, cudaStream_t st) {
  const WdShapes& sh = p->sh;
  const DenseLayout& dl = p->dl;
  const int cnt = n * sh.F;
  const int work = cnt > n * sh.dense_pad ? cnt : n * sh.dense_pad;
  // pull addresses of the ids' rows / wide weights, dense features as zero-padded bf16 (no mailbox claims)
  wd_prefetch_addr_kernel<<<(work + 255) / 256, 256, 0, st>>>(ids, cnt, sh.V, p->pt, p->sl, p->row_src, p->wide_src, dense_raw,
                                                              p->dense_x, static_cast<long long>(sh.dense_pad), 0, n,
                                                              sh.dense_in, sh.dense_pad);
  DTF_CUDA_OK(cudaGetLastError());
  count_launch();
  const __nv_bfloat16* W = reinterpret_cast<const __nv_bfloat16*>(p->slab + p->sl.wbf16);
  const float* small = reinterpret_cast<const float*>(p->slab + p->sl.small_f32);
  const float *b1 = small + (dl.b1 - dl.small_begin), *b2 = small + (dl.b2 - dl.small_begin), *b3 = small + (dl.b3 - dl.small_begin);
  auto layers = [&](auto tagm) {
    constexpr int MM = decltype(tagm)::value;
    if (small_layer<MM>(nullptr, 0, p->row_src, p->wide_src, sh.F, p->dense_x, sh.dense_pad, p->wide_vals, W + dl.w1, sh.K0, b1,
                        p->H1, sh.h1, n, sh.h1, sh.K0, st) != 0)
      return -1;
    if (small_layer<MM>(p->H1, sh.h1, nullptr, nullptr, 0, nullptr, 0, nullptr, W + dl.w2, sh.h1, b2, p->H2, sh.h2, n, sh.h2,
                        sh.h1, st) != 0)
      return -1;
    return small_layer<MM>(p->H2, sh.h2, nullptr, nullptr, 0, nullptr, 0, nullptr, W + dl.w3, sh.h2, b3, p->H3, sh.h3, n, sh.h3,
                           sh.h2, st);
  };
  if (n == 1) return layers(std::integral_constant<int, 1>{});
  return layers(std::integral_constant<int, 4>{});
}

extern "C" {

// serving / evaluation forward: probabilities of the first n examples (n <= batch) into the plan's
// "prob" buffer.  Small n takes the small-M path of the same kernels: one 128-row tile, TMA
// zero-fill for the missing rows.
int dtf_wd_predict(dtf_wd_plan* p, const int* ids, const float* dense_raw, int n, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  if (n < 1 || n > p->cfg.batch) {
    set_error("dtf_wd_predict: n=%d outside [1, %d]", n, p->cfg.batch);
    return -2;
  }
  const int full = p->sh.B;
  p->sh.B = n;
  select_parity(p, p->cfg.sync ? 0 : static_cast<int>(p->n_compute & 1));
  static const bool small_off = getenv("DTF_SERVE_SMALL") != nullptr && getenv("DTF_SERVE_SMALL")[0] == '0';
  // measured kernel chain (eager launches, 1 GPU): batch 1 76 -> 44 us, batch 4 70 -> 42 us, batch 16 70 -> 80 us (the
  // 16 activation rows are re-read from shared memory for every row of W), so the path is taken up to 4 examples
  if (n <= 4 && !small_off && p->sh.K0 % 8 == 0 && p->sh.h1 % 8 == 0 && p->sh.h2 % 8 == 0) {
    // rows prefetched for the next training step live in this parity's buffers, which the forward below writes
    // (wide values): wait for the gather and drop it, exactly as forward() does -- that step then pulls again
    if (p->pf_valid[p->cur_par]) {
      if (cudaStreamWaitEvent(st, p->ev_pf[p->cur_par], 0) != cudaSuccess) {
        p->sh.B = full;
        set_error("dtf_wd_predict: %s", cudaGetErrorString(cudaGetLastError()));
        return -1;
      }
      p->pf_valid[p->cur_par] = false;
    }
    const int rc_small = forward_small(p, ids, dense_raw, n, st);
    int rc2 = rc_small;
    if (rc2 == 0) rc2 = launch_loss(p, nullptr, st, false);
    p->sh.B = full;
    return rc2;
  }
  int rc = forward(p, ids, dense_raw, st, false);
  if (rc == 0) rc = launch_loss(p, nullptr, st, false);
  p->sh.B = full;
  return rc;
}

// sum of per-example losses of the last step (device -> host, synchronises the stream)
int dtf_wd_read_loss(dtf_wd_plan* p, float* loss_sum, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  DTF_CUDA_OK(cudaMemcpyAsync(loss_sum, p->loss_acc, sizeof(float), cudaMemcpyDeviceToHost, st));
  DTF_CUDA_OK(cudaStreamSynchronize(st));
  return 0;
}

void* dtf_wd_loss_ptr(dtf_wd_plan* p) {
  return reinterpret_cast<float*>(p->zero_par[p->last_compute_parity] + kMaxWorld + 4);
}
int dtf_wd_last_parity(dtf_wd_plan* p) { return p->last_compute_parity; }

// Async (Hogwild) step with the parameter-server side off the critical path: the compute phase of
// step t runs on `compute_stream`; barrier + apply (+ barrier) of step t run on `apply_stream` and
// overlap the compute phase of step t+1, which pulls whatever the shards hold at that moment (stale
// by at most one step, like a TF worker that does not wait for the ps to apply).  Parity-double-
// buffered mailboxes keep step t+1's pushes away from the entries step t's apply is still reading;
// a parity's buffers are reused only after the apply that consumed them has finished everywhere.
int dtf_wd_step_overlapped(dtf_wd_plan* p, const int* ids, const float* dense_raw, const float* labels,
                           void* compute_stream, void* apply_stream) {
  if (p->cfg.sync) {
    set_error("dtf_wd_step_overlapped: sync-replica mode has no apply/compute overlap; use dtf_wd_step");
    return -2;
  }
  cudaStream_t C = reinterpret_cast<cudaStream_t>(compute_stream), A = reinterpret_cast<cudaStream_t>(apply_stream);
  const int par = static_cast<int>(p->n_compute & 1);
  if (p->ev_apply_valid[par]) DTF_CUDA_OK(cudaStreamWaitEvent(C, p->ev_apply[par], 0));
  if (dtf_wd_step_compute(p, ids, dense_raw, labels, compute_stream) != 0) return -1;
  DTF_CUDA_OK(cudaEventRecord(p->ev_push[par], C));
  DTF_CUDA_OK(cudaStreamWaitEvent(A, p->ev_push[par], 0));
  if (dtf_wd_barrier(p, 1, apply_stream) != 0) return -1;
  // two blocks per SM: the apply must leave thread slots for the next step's persistent GEMM CTAs
  p->apply_blocks_per_sm = p->cfg.overlap_apply_blocks > 0 ? p->cfg.overlap_apply_blocks : 2;
  const int arc = dtf_wd_step_apply(p, apply_stream);
  p->apply_blocks_per_sm = 8;
  if (arc != 0) return -1;
  if (p->cfg.world > 1 && dtf_wd_barrier(p, 0, apply_stream) != 0) return -1;
  DTF_CUDA_OK(cudaEventRecord(p->ev_apply[par], A));
  p->ev_apply_valid[par] = true;
  return 0;
}

// make `stream` wait for every apply issued so far (before reading shards, or to close a capture)
int dtf_wd_join(dtf_wd_plan* p, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  for (int i = 0; i < 2; ++i)
    if (p->ev_apply_valid[i]) DTF_CUDA_OK(cudaStreamWaitEvent(st, p->ev_apply[i], 0));
  for (int i = 0; i < 2; ++i)
    if (p->pf_valid[i]) DTF_CUDA_OK(cudaStreamWaitEvent(st, p->ev_pf[i], 0));   // an announced batch nobody stepped on yet
  return 0;
}

// forget cross-step dependencies (call after a full synchronisation, e.g. before capturing a graph)
void dtf_wd_overlap_reset(dtf_wd_plan* p) {
  p->ev_apply_valid[0] = p->ev_apply_valid[1] = false;
  p->pf_valid[0] = p->pf_valid[1] = false;
  p->next_ids = nullptr;
  p->next_dense = nullptr;
}

// Async mode: announce the batch of the NEXT training step (device pointers that stay valid and unchanged until
// that step has run).  The coming step then starts a low-priority gather of those rows that overlaps its own
// backward pass, and the step after runs layer 1 as a plain GEMM on them instead of the fused pull -- provided it
// is called with the same `ids` pointer; otherwise the prefetched rows are dropped.  The rows are read up to one
// step earlier than the fused pull would read them (one more step of staleness, which Hogwild training allows);
// sync-replica plans ignore the call.  Returns 1 if the announcement was taken, 0 if ignored.
int dtf_wd_set_next_batch(dtf_wd_plan* p, const int* ids, const float* dense_raw) {
  static const bool off = getenv("DTF_PREFETCH") != nullptr && getenv("DTF_PREFETCH")[0] == '0';
  if (p->cfg.sync || p->pf_stream == nullptr || off || p->profile) return 0;
  p->next_ids = ids;
  p->next_dense = dense_raw;
  return 1;
}

long long dtf_wd_prefetched_steps(dtf_wd_plan* p) { return p->n_prefetched; }

// DTF_STAMPS=1 diagnostics.  Count of marks issued so far, and "index name microseconds" lines for the marks
// [from, to) -- times relative to the earliest of them, 'pending' for a slot the device has not written since
// `clear` zeroed it.  Reads on a private stream, so it also works while the training streams are stuck.
long long dtf_wd_stamps_count(dtf_wd_plan* p) { return p->stamps ? p->n_stamp : -1; }
int dtf_wd_stamps_read(dtf_wd_plan* p, long long from, long long to, int clear, char* buf, size_t cap) {
  if (p->stamps == nullptr || cap == 0) return -2;
  static thread_local cudaStream_t probe = nullptr;
  if (probe == nullptr) DTF_CUDA_OK(cudaStreamCreateWithFlags(&probe, cudaStreamNonBlocking));
  std::vector<unsigned long long> h(kStampCap);
  DTF_CUDA_OK(cudaMemcpyAsync(h.data(), p->stamps, kStampCap * 8, cudaMemcpyDeviceToHost, probe));
  if (clear) DTF_CUDA_OK(cudaMemsetAsync(p->stamps, 0, kStampCap * 8, probe));
  DTF_CUDA_OK(cudaStreamSynchronize(probe));
  if (to > p->n_stamp) to = p->n_stamp;
  if (from < to - kStampCap) from = to - kStampCap;
  if (from < 0) from = 0;
  unsigned long long t0 = ~0ull;
  for (long long i = from; i < to; ++i)
    if (h[i % kStampCap] != 0ull && h[i % kStampCap] < t0) t0 = h[i % kStampCap];
  std::string out;
  char line[160];
  for (long long i = from; i < to; ++i) {
    const unsigned long long t = h[i % kStampCap];
    if (t == 0ull) snprintf(line, sizeof(line), "%lld %s pending\n", i, p->stamp_names[i % kStampCap].c_str());
    else snprintf(line, sizeof(line), "%lld %s %.1f\n", i, p->stamp_names[i % kStampCap].c_str(), (t - t0) * 1e-3);
    out += line;
  }
  snprintf(buf, cap, "%s", out.c_str());
  return 0;
}

}  // extern "C"
