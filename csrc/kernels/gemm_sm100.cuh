// Persistent warp-specialised bf16 GEMM for sm_100a: TMA -> 128B-swizzled SMEM ring -> tcgen05.mma
// with fp32 accumulators double-buffered in TMEM -> tcgen05.ld epilogue.
//
//   D[M,N] = epilogue( A[M,K] * B[N,K]^T )
//
// Both operands may be K-major (reduction dim contiguous) or MN-major (the other dim contiguous), so
// forward (x W^T), dgrad (dY W) and wgrad (dY^T X) all run on the same kernel with no transposes.
// Replaces the Eigen Mul/Add/Sum kernels behind /root/reference/trainer/task.py:137-139 for the
// wide-and-deep model.
//
// Warp roles (192 threads, 1 CTA / SM):
//   warp 0      : TMA producer (one elected lane)
//   warp 1      : TMEM allocator + tcgen05.mma issuer (one lane)
//   warps 2..5  : epilogue; warp w owns TMEM lanes [32*(w%4), 32*(w%4)+32)
// The fused pull (GATHER) adds warps 6..17 = gather warps that fetch embedding rows from the owning
// ranks' HBM (local or peer over NVLink) into the A operand buffer, and warp 18 = watcher that turns
// their arrival counters into "k-blocks ready" for the TMA producer, so that the MMAs trail the gather
// k-block by k-block (608 threads).  The fused push is an epilogue that scatters output rows to peer
// mailboxes.  CTAS = 2 runs every variant on CTA pairs (clusters of two, tcgen05.mma.cta_group::2).
#pragma once

#include "sm100_ptx.cuh"

namespace dtf {

constexpr int BM = 128;       // rows per CTA tile == UMMA M
constexpr int BK = 64;        // 64 bf16 = 128 B = one swizzle row
constexpr int UMMA_K = 16;    // fixed for 16-bit inputs
constexpr int kNumMainThreads = 192;
constexpr int kNumGatherWarps = 12;
constexpr int kNumGatherThreads = kNumGatherWarps * 32 + 32;   // gather warps + one watcher warp
constexpr int kSliceRows = 16;          // rows per gather slice (one warp, 8 x LDG.128 per lane)
constexpr int kFlagWords = 4;           // 64-bit words of 16 chunk counters per row-tile: slots + 1 <= 64
constexpr int kSlicesPerTile = BM / kSliceRows;

enum Epilogue : int {
  EPI_BF16 = 0,            // D = bf16(acc)
  EPI_BIAS_BF16 = 1,       // D = bf16(acc + bias[n])
  EPI_BIAS_RELU_BF16 = 2,  // D = bf16(relu(acc + bias[n]))
  EPI_DRELU_BF16 = 3,      // D = bf16(mask bit(m,n) ? acc : 0)         (dgrad through ReLU)
  EPI_F32 = 4,             // D = acc (fp32)
  EPI_F32_ATOMIC = 5,      // D += acc (fp32, red.global.add; split-K wgrad)
  EPI_PUSH_ROWS = 6,       // bf16 row segments scattered to per-(m,slot) destination pointers
};

struct GemmParams {
  int M, N, K;
  int m_tiles, n_tiles, splits, k_blocks_per_split;
  void* D;
  long long ldd;  // elements
  const float* bias;
  const uint32_t* mask_in;   // EPI_DRELU_BF16: bit j of word [m, n/32] = keep column n
  uint32_t* mask_out;        // EPI_BIAS_RELU_BF16 (optional): same format, bit = (output > 0)
  long long ldmask;          // words per row (even)
  // --- fused pull (gather warps): rows of 128 bf16 fetched through row_src pointers ---
  const unsigned long long* row_src;  // [M * slots] global addresses of 256-byte rows (may be peer)
  const __nv_bfloat16* dense_x;       // [M, dense_cols] local bf16 (already zero padded)
  __nv_bfloat16* a_buf;               // [M, lda_buf] gathered operand, consumed through tmA
  long long lda_buf;
  int slots;                          // F
  int dense_cols;                     // 64
  unsigned long long* tile_flags;     // [m_tiles][kFlagWords]: 4-bit arrival counter per (row-tile, chunk), zero before launch
  unsigned int* chunk_counter;        // [1] reserved
  // optional side gather: one fp32 per (row, slot) fetched from wide_src[m*slots+f] into wide_out
  const unsigned long long* wide_src;
  float* wide_out;
  unsigned long long* trace;          // debugging only (bench/pull_trace.py): globaltimer stamps, nullptr = off
  // debugging only (DTF_GEMM_DBG=1, wd plan): word 0 = number of stall records, words 16 + 4 id + {0,1,2} = CTAs of
  // launch `dbg_id` that entered / finished set-up / left, from word 128: 4-word records of waits that lasted > 2 s
  unsigned long long* dbg;
  int dbg_id;
  // --- fused dense pull (all-gather + GEMM): B is row-sharded over ranks; tile n reads its rows
  //     straight from the owning rank through that rank's tensor map (global memory array) ---
  const CUtensorMap* b_maps;          // [shards] or nullptr
  int b_rows_per_shard;               // multiple of the N tile
  // --- fused push epilogue: destination of the 128-wide bf16 row for (m, slot) ---
  const unsigned long long* row_dst;  // [M * slots] global addresses (may be peer mailbox)
};

template <int BN, bool GATHER = false, int CTAS = 1>
struct GemmSmem {
  static constexpr int kABytes = BM * BK * 2;  // 16 KB
  static constexpr int kBBytes = (BN / CTAS) * BK * 2;   // a CTA pair splits the B rows
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (kStageBytes >= 48 * 1024) ? 4 : ((kStageBytes >= 32 * 1024) ? 6 : 8);
  static constexpr int kGatherBytes = 0;
  static constexpr int kStagingPerWarp = 5120;  // 32 rows x 128 B (TMA store) or x 144 B (push)
  static constexpr int kStagingBytes = 4 * kStagingPerWarp;
  static constexpr int kBarBytes = 256;
  static constexpr int kBiasBytes = BN * 4 * 2;
  static constexpr int kTotal =
      kStages * kStageBytes + kStagingBytes + kGatherBytes + kBarBytes + kBiasBytes + 1024 /*align*/;
};

// tile index -> (m, n, split); n fastest so that concurrently running CTAs share A row-tiles.
__device__ __forceinline__ void decode_tile(const GemmParams& p, int tile, int& m_t, int& n_t,
                                            int& s_t, int m_units = -1, int ctas = 1, int rank = 0) {
  if (m_units < 0) m_units = p.m_tiles;
  n_t = tile % p.n_tiles;
  int r = tile / p.n_tiles;
  m_t = (r % m_units) * ctas + rank;   // 128-row tile of this CTA (may be past the end for the odd pair)
  s_t = r / m_units;
}

// CTAS = 2: the grid is launched in clusters of two CTAs (one TPC); the pair computes a 256 x BN tile
// with tcgen05.mma.cta_group::2 -- CTA rank r holds rows [128 r, 128 r + 128) of A and rows
// [BN/2 r, BN/2 r + BN/2) of B, so each SM fills 16 KB + 16 KB per k-block instead of 16 KB + 32 KB.
template <int BN, bool A_MN, bool B_MN, int EPI, bool GATHER, int CTAS = 1>
__global__ void __launch_bounds__(GATHER ? kNumMainThreads + kNumGatherThreads : kNumMainThreads, 1) __maxnreg__(GATHER ? 104 : 255)
gemm_sm100_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmD, const GemmParams p) {
  using S = GemmSmem<BN, GATHER, CTAS>;
  constexpr bool kPair = CTAS == 2;
  constexpr int BNL = BN / CTAS;   // B rows (output columns) staged by one CTA
  constexpr int kStages = S::kStages;
  constexpr uint32_t kTmemCols = (2 * BN <= 32) ? 32 : 2 * BN;  // two accumulator buffers
  constexpr bool kOutF32 = (EPI == EPI_F32 || EPI == EPI_F32_ATOMIC);
  constexpr bool kPush = (EPI == EPI_PUSH_ROWS);
  constexpr int CW = kOutF32 ? 32 : 64;          // output columns per 128-byte staging row
  constexpr int kPitch = kPush ? 144 : 128;      // staging row pitch (push rows are not swizzled)

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t staging_base = smem_base + kStages * S::kStageBytes;  // 1024-aligned
  const uint32_t gather_base = staging_base + S::kStagingBytes;      // 1024-aligned
  const uint32_t bar_base = gather_base + S::kGatherBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
  auto tfull_bar = [&](int b) { return bar_base + 8u * (2 * kStages + b); };
  auto tempty_bar = [&](int b) { return bar_base + 8u * (2 * kStages + 2 + b); };
  uint8_t* misc = smem + kStages * S::kStageBytes + S::kStagingBytes + S::kGatherBytes;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(misc + 8 * (2 * kStages + 4));
  const uint32_t ready_addr = bar_base + 8u * (2 * kStages + 4) + 8u;   // fused pull: chunks published by the watcher
  float* bias_smem = reinterpret_cast<float*>(misc + S::kBarBytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // mbarrier wait that, with the debug buffer attached, leaves a record once it has been spinning for 2 s
  auto wait_bar = [&](uint32_t bar, uint32_t parity, int tag, uint32_t progress) {
    if (p.dbg == nullptr) {
      mbar_wait(bar, parity);
      return;
    }
    unsigned long long t0 = 0;
    bool reported = false;
    for (uint32_t spin = 0; !mbar_try_wait(bar, parity); ++spin) {
      if ((spin & 1023u) == 1023u && !reported) {
        const unsigned long long t = globaltimer_ns();
        if (t0 == 0ull) t0 = t;
        else if (t - t0 > 2000000000ull) {
          const unsigned long long slot = atomicAdd(p.dbg, 1ull) % 96ull;
          unsigned long long* d = p.dbg + 128 + slot * 4;
          d[0] = (static_cast<unsigned long long>(p.dbg_id) << 32) | blockIdx.x;
          d[1] = (static_cast<unsigned long long>(tag) << 32) | static_cast<unsigned int>(warp);
          d[2] = (static_cast<unsigned long long>(parity) << 32) | bar;
          d[3] = progress;
          reported = true;
        }
      }
    }
  };
  if (p.dbg != nullptr && threadIdx.x == 0) atomicAdd(p.dbg + 16 + 4 * p.dbg_id, 1ull);
  // tiles are 128 x BN per CTA, 256 x BN per CTA pair; a pair walks one tile sequence together
  const uint32_t cta_rank = kPair ? cluster_ctarank() : 0u;
  const int m_units = (p.m_tiles + CTAS - 1) / CTAS;
  const int num_tiles = m_units * p.n_tiles * p.splits;
  const int total_kb = (p.K + BK - 1) / BK;
  const int unit_id = blockIdx.x / CTAS;
  // Fused pull: the tiles are dealt over a whole number of row-tile groups per wave, so that wave i of
  // the GEMM consumes exactly the row-tiles of gather group i (CTAs beyond that only gather).
  int tile_step = gridDim.x / CTAS;
  if constexpr (GATHER) {
    if (tile_step >= p.n_tiles) tile_step = (tile_step / p.n_tiles) * p.n_tiles;
  }
  const int tile_begin = unit_id < tile_step ? unit_id : num_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if constexpr (!kPush) tma_prefetch_desc(&tmD);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tfull_bar(b), 1);
      mbar_init(tempty_bar(b), 4 * CTAS);  // one arrival per epilogue warp (of both CTAs on the leader)
    }
    if constexpr (GATHER) st_shared_u32(ready_addr, 0u);
    fence_barrier_init();
  }
  // Pair: both CTAs must be running before either warp issues the cta_group::2 allocation, which acts on the
  // tensor memory of both SMs.  Without this cluster barrier a pair now and then never came back from the
  // allocation (2 GPUs, eager launches: bench/stress_tail.py stalled within ~100 trials of 3 steps, DTF_GEMM_DBG
  // showed two CTAs of one wgrad launch entered and not set up); with it 12 000 trials ran through.
  if constexpr (kPair) cluster_sync_all();
  if (warp == 1) {
    if constexpr (kPair) {
      tmem_alloc_pair(smem_u32(tmem_slot), kTmemCols);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(smem_u32(tmem_slot), kTmemCols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if constexpr (kPair) cluster_sync_all();   // the peer's barriers must exist before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (p.dbg != nullptr && threadIdx.x == 0) atomicAdd(p.dbg + 16 + 4 * p.dbg_id + 1, 1ull);
  // everything above (barrier init, TMEM allocation, descriptor prefetch) may overlap the tail of the
  // previous kernel in the stream; its results are needed only from here on
  pdl_launch_dependents();
  pdl_wait();

  if (warp == 0) {
    // =============================== TMA producer ===============================
    if (lane == 0) {
      uint32_t kiter = 0;
      for (int tile = tile_begin; tile < num_tiles; tile += tile_step) {
        int m_t, n_t, s_t;
        decode_tile(p, tile, m_t, n_t, s_t, m_units, CTAS, cta_rank);
        const int kb0 = s_t * p.k_blocks_per_split;
        const int kb1 = min(total_kb, kb0 + p.k_blocks_per_split);
        // Fused pull: k-blocks become loadable chunk by chunk (one slot = 128 columns = 2 k-blocks, the
        // dense block last) as the gather warps of the grid deliver them, so the MMAs of a tile run
        // right behind its gather instead of after it.
        int ready_kb = kb1;
        if constexpr (GATHER) {
          ready_kb = kb0;
          if (p.trace) p.trace[65536 + (blockIdx.x * 8 + (tile / tile_step)) * 4 + 0] = globaltimer_ns();
        }
        for (int kb = kb0; kb < kb1; ++kb, ++kiter) {
          if constexpr (GATHER) {
            if (kb >= ready_kb) {
              // the watcher warp polls the global arrival counters, performs the gpu-scope acquire
              // and publishes "chunks complete" in shared memory; this thread only pays a shared-memory
              // acquire load and the generic->async proxy fence before TMA reads the rows
              const int chunks = p.slots + 1;
              const int fneed = min(kb / (128 / BK), p.slots);
              const uint32_t base = static_cast<uint32_t>(tile / tile_step) * chunks;
              uint32_t r;
              for (;;) {
                r = ld_acquire_cta_shared_u32(ready_addr);
                if (r > base + fneed) break;
                __nanosleep(20);
              }
              const int c = min(static_cast<int>(r - base), chunks);
              ready_kb = (c >= chunks) ? kb1 : c * (128 / BK);
              asm volatile("fence.proxy.async;" ::: "memory");
              if (p.trace && ready_kb == kb1) {
                p.trace[65536 + (blockIdx.x * 8 + (tile / tile_step)) * 4 + 1] = globaltimer_ns();
                p.trace[65536 + (blockIdx.x * 8 + (tile / tile_step)) * 4 + 3] = static_cast<unsigned long long>(m_t);
              }
            }
          }
          const int s = kiter % kStages;
          const uint32_t ph = (kiter / kStages) & 1;
          wait_bar(empty_bar(s), ph ^ 1, 1, kiter);
          // pair: both CTAs' loads are counted on the leader's barrier, armed by the leader alone
          const uint32_t fbar = kPair ? mapa_shared(full_bar(s), 0) : full_bar(s);
          if (!kPair || cta_rank == 0) mbar_arrive_expect_tx(full_bar(s), CTAS * S::kStageBytes);
          auto tma_load_2d = [&](uint32_t dst, const void* map, uint32_t bar, int c0, int c1) {
            if constexpr (kPair) dtf::tma_load_2d_pair(dst, map, bar, c0, c1);
            else dtf::tma_load_2d(dst, map, bar, c0, c1);
          };
          const uint32_t sa = smem_base + s * S::kStageBytes;
          const uint32_t sb = sa + S::kABytes;
          if constexpr (!A_MN) {
            tma_load_2d(sa, &tmA, fbar, kb * BK, m_t * BM);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_2d(sa + j * (BK * 128), &tmA, fbar, m_t * BM + j * 64, kb * BK);
          }
          if constexpr (!B_MN) {
            const void* mapB = &tmB;
            int nrow = n_t * BN + static_cast<int>(cta_rank) * BNL;
            if (p.b_maps != nullptr) {  // peer-addressed operand: pick the owner's tensor map
              const int shard = nrow / p.b_rows_per_shard;
              mapB = p.b_maps + shard;
              nrow -= shard * p.b_rows_per_shard;
            }
            tma_load_2d(sb, mapB, fbar, kb * BK, nrow);
          } else {
#pragma unroll
            for (int j = 0; j < BNL / 64; ++j)
              tma_load_2d(sb + j * (BK * 128), &tmB, fbar, n_t * BN + static_cast<int>(cta_rank) * BNL + j * 64, kb * BK);
          }
        }
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer ===============================
    if (lane == 0 && cta_rank == 0) {   // of a pair, only the leader issues MMAs
      constexpr uint32_t idesc = make_idesc_bf16(BM * CTAS, BN, A_MN, B_MN);
      auto umma_bf16 = [&](uint32_t d, uint64_t ad, uint64_t bd, uint32_t id, uint32_t acc) {
        if constexpr (kPair) dtf::umma_bf16_pair(d, ad, bd, id, acc);
        else dtf::umma_bf16(d, ad, bd, id, acc);
      };
      auto umma_commit = [&](uint32_t bar) {
        if constexpr (kPair) dtf::umma_commit_pair(bar);
        else dtf::umma_commit(bar);
      };
      uint32_t kiter = 0;
      int it = 0;
      for (int tile = tile_begin; tile < num_tiles; tile += tile_step, ++it) {
        int m_t, n_t, s_t;
        decode_tile(p, tile, m_t, n_t, s_t, m_units, CTAS, cta_rank);
        const int kb0 = s_t * p.k_blocks_per_split;
        const int kb1 = min(total_kb, kb0 + p.k_blocks_per_split);
        const int buf = it & 1;
        wait_bar(tempty_bar(buf), ((it >> 1) & 1) ^ 1, 2, static_cast<uint32_t>(it));
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        for (int kb = kb0; kb < kb1; ++kb, ++kiter) {
          const int s = kiter % kStages;
          const uint32_t ph = (kiter / kStages) & 1;
          wait_bar(full_bar(s), ph, 3, kiter);
          tc_fence_after();
          const uint32_t sa = smem_base + s * S::kStageBytes;
          const uint32_t sb = sa + S::kABytes;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // K-major: 8-row groups 1024 B apart, advance 32 B per UMMA_K inside the swizzle row.
            // MN-major: 64-element column chunks BK*128 B apart (LBO), 8-k groups 1024 B apart
            //           (SBO), advance 2 groups = 2048 B per UMMA_K.
            const uint64_t adesc = A_MN ? make_smem_desc_sw128(sa + k * 2048, BK * 128, 1024)
                                        : make_smem_desc_sw128(sa + k * 32, 16, 1024);
            const uint64_t bdesc = B_MN ? make_smem_desc_sw128(sb + k * 2048, BK * 128, 1024)
                                        : make_smem_desc_sw128(sb + k * 32, 16, 1024);
            umma_bf16(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(empty_bar(s));  // frees the smem stage once these MMAs retire
        }
        umma_commit(tfull_bar(buf));  // accumulator complete -> epilogue
      }
    }
  } else if (warp < 6) {
    // =============================== epilogue ===============================
    // TMEM -> registers -> (bias / ReLU / mask) -> 128B-swizzled staging rows in SMEM -> TMA store
    // (or TMA reduce-add for split-K, or per-row bulk copies to peer mailboxes for the fused push).
    const int q = warp & 3;           // TMEM lane quarter this warp may touch
    const int et = threadIdx.x - 64;  // 0..127 within the epilogue group
    const uint32_t stg = staging_base + q * S::kStagingPerWarp;
    const uint32_t my_row_smem = stg + lane * kPitch;
    int it = 0;
    for (int tile = tile_begin; tile < num_tiles; tile += tile_step, ++it) {
      int m_t, n_t, s_t;
      decode_tile(p, tile, m_t, n_t, s_t, m_units, CTAS, cta_rank);
      const int buf = it & 1;
      const int n0 = n_t * BN;
      const int row0 = m_t * BM + q * 32;
      const int row = row0 + lane;
      float* bsm = bias_smem + buf * BN;
      if constexpr (EPI == EPI_BIAS_BF16 || EPI == EPI_BIAS_RELU_BF16) {
        for (int i = et; i < BN; i += 128) bsm[i] = (n0 + i < p.N) ? __ldg(p.bias + n0 + i) : 0.f;
        named_bar_sync(1, 128);
      }
      // operands of the epilogue that do not depend on the accumulator are fetched before the wait
      uint32_t mw[BN / 32];
      if constexpr (EPI == EPI_DRELU_BF16) {
#pragma unroll
        for (int i = 0; i < BN / 32; ++i)
          mw[i] = (row < p.M && n0 + i * 32 < p.N) ? __ldg(p.mask_in + static_cast<long long>(row) * p.ldmask + (n0 >> 5) + i)
                                                   : 0u;
      }
      unsigned long long pdst[(BN + 127) / 128];
      if constexpr (kPush) {
#pragma unroll
        for (int i = 0; i < (BN + 127) / 128; ++i) {
          const int f = (n0 >> 7) + i;
          pdst[i] = (row < p.M && f < p.slots) ? __ldg(p.row_dst + static_cast<long long>(row) * p.slots + f) : 0ull;
        }
      }
      wait_bar(tfull_bar(buf), (it >> 1) & 1, 4, static_cast<uint32_t>(it));
      tc_fence_after();
      if constexpr (GATHER) {
        if (p.trace && warp == 2 && lane == 0) p.trace[65536 + (blockIdx.x * 8 + it) * 4 + 2] = globaltimer_ns();   // MMAs of the tile done
      }
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN;
#pragma unroll 1
      for (int c = 0; c < BN / CW; ++c) {
        uint32_t w[32];  // 128 bytes of output for this thread's row
        if constexpr (kOutF32) {
          tmem_ld_32x32(t_row + c * 32, w);
          tmem_ld_wait();
        } else {
          uint32_t r0[32], r1[32];
          tmem_ld_32x32(t_row + c * 64, r0);
          tmem_ld_32x32(t_row + c * 64 + 32, r1);
          tmem_ld_wait();
          float v[64];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            v[j] = __uint_as_float(r0[j]);
            v[32 + j] = __uint_as_float(r1[j]);
          }
          if constexpr (EPI == EPI_BIAS_BF16 || EPI == EPI_BIAS_RELU_BF16) {
#pragma unroll
            for (int j = 0; j < 64; ++j) {
              v[j] += bsm[c * 64 + j];
              if (EPI == EPI_BIAS_RELU_BF16) v[j] = fmaxf(v[j], 0.f);
            }
            if constexpr (EPI == EPI_BIAS_RELU_BF16) {
              if (p.mask_out != nullptr) {
                uint32_t m0 = 0, m1 = 0;
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                  m0 |= (v[j] > 0.f ? 1u : 0u) << j;
                  m1 |= (v[32 + j] > 0.f ? 1u : 0u) << j;
                }
                const int col = n0 + c * 64;
                if (row < p.M && col < p.N)
                  *reinterpret_cast<uint2*>(p.mask_out + static_cast<long long>(row) * p.ldmask + (col >> 5)) =
                      make_uint2(m0, m1);
              }
            }
          }
          if constexpr (EPI == EPI_DRELU_BF16) {
            const uint32_t m0 = mw[2 * c], m1 = mw[2 * c + 1];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (!((m0 >> j) & 1u)) v[j] = 0.f;
              if (!((m1 >> j) & 1u)) v[32 + j] = 0.f;
            }
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) w[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
        }
        // the staging rows must have been read out by the previous chunk's store
        if constexpr (kPush) {
          __syncwarp();
        } else {
          if (lane == 0) tma_store_wait_read<0>();
          __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t a = kPush ? (my_row_smem + j * 16) : (my_row_smem + ((j ^ (lane & 7)) << 4));
          st_shared_v4(a, w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
        }
        if constexpr (kPush) {
          // Scatter to the owners' mailboxes with full 128-byte segments per row: eight lanes cover
          // one row's half (8 x 16 B), so every store instruction writes four rows and each row
          // half leaves the SM -- and crosses NVLink -- as one 128-byte transaction.
          __syncwarp();
          const int col = n0 + c * 64;
          const unsigned long long mine = pdst[(c * 64) >> 7];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int r = 4 * u + (lane >> 3);
            const unsigned long long d = __shfl_sync(0xffffffffu, mine, r);
            uint32_t x0, x1, x2, x3;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(x0), "=r"(x1), "=r"(x2), "=r"(x3)
                         : "r"(stg + r * kPitch + (lane & 7) * 16));
            if (d != 0ull)
              st_global_v4(reinterpret_cast<uint8_t*>(d) + (col & 127) * 2 + (lane & 7) * 16, x0, x1, x2, x3);
          }
        } else {
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            if constexpr (EPI == EPI_F32_ATOMIC) tma_reduce_add_2d(&tmD, stg, n0 + c * CW, row0);
            else tma_store_2d(&tmD, stg, n0 + c * CW, row0);
            tma_store_commit();
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (kPair) mbar_arrive_cluster(mapa_shared(tempty_bar(buf), 0));
        else mbar_arrive(tempty_bar(buf));
      }
      if constexpr (GATHER) {
        if (p.trace && warp == 2 && lane == 0) p.trace[70272 + blockIdx.x * 8 + it] = globaltimer_ns();   // epilogue of the tile issued
      }
    }
    tma_store_wait_all();
  } else {
    // =============================== gather warps (fused pull) ===============================
    // Twelve independent warps, no shared memory and no intra-CTA synchronisation.  The work is cut
    // into slices of 16 rows x one chunk (one slot = 256 B per row, 4 KB per slice; the dense block is
    // the last chunk).  A warp loads the 16 row addresses, issues 8 LDG.128 per lane straight from the
    // owning rank's HBM (local, or a peer's over NVLink), stores them into the GEMM operand buffer
    // and bumps the 4-bit arrival counter of (row-tile, chunk).  The loads of the next slice are
    // issued before the previous slice is stored and released (two register sets), so every warp
    // always has 4 KB in flight.
    // Order: row-tiles are taken in groups that match one wave of GEMM tiles; inside a group the
    // slices run chunk-major, so all row-tiles of the wave receive chunk 0, then chunk 1, ... and the
    // MMAs of the whole wave follow the gather k-block by k-block.  (Needs every CTA resident:
    // grid <= SM count, one CTA per SM.)
    if constexpr (GATHER) {
      if (warp == 6 + kNumGatherWarps) {
        // ------------------------------ watcher ------------------------------
        // Follows this CTA's tile sequence, polls the 4-bit arrival counters of the current row-tile
        // (relaxed loads with back-off), and after every advance does the gpu-scope acquire and
        // publishes the running count of complete chunks for the producer.
        if (lane == 0) {
          const int chunks = p.slots + 1;
          uint32_t published = 0;
          for (int tile = tile_begin; tile < num_tiles; tile += tile_step) {
            int m_t, n_t, s_t;
            decode_tile(p, tile, m_t, n_t, s_t, m_units, CTAS, cta_rank);
            int c = 0;
            if (m_t >= p.m_tiles) {   // the odd CTA of the last pair has no rows: nothing to wait for
              published += static_cast<uint32_t>(chunks);
              st_release_cta_shared_u32(ready_addr, published);
              c = chunks;
            }
            while (c < chunks) {
              unsigned long long v;
              asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];"
                           : "=l"(v)
                           : "l"(p.tile_flags + static_cast<long long>(m_t) * kFlagWords + (c >> 4))
                           : "memory");
              int c2 = c;
              while (c2 < chunks && (c2 >> 4) == (c >> 4) && ((v >> ((c2 & 15) * 4)) & 15ull) == kSlicesPerTile) ++c2;
              if (c2 > c) {
                asm volatile("fence.acq_rel.gpu;" ::: "memory");
                published += static_cast<uint32_t>(c2 - c);
                c = c2;
                st_release_cta_shared_u32(ready_addr, published);
              } else {
                __nanosleep(32);
              }
            }
          }
        }
      } else {
      const int gw = warp - 6;
      const int chunks_per_tile = p.slots + 1;
      const long long total_slices = static_cast<long long>(p.m_tiles) * chunks_per_tile * kSlicesPerTile;
      const long long stride = static_cast<long long>(gridDim.x) * kNumGatherWarps;
      const int dense_pieces = p.dense_cols / 8;  // 16-byte pieces per dense row
      const int group_tiles = max(1, tile_step / p.n_tiles) * CTAS;
      const long long group_slices = static_cast<long long>(group_tiles) * chunks_per_tile * kSlicesPerTile;
      const unsigned long long fallback_src = reinterpret_cast<unsigned long long>(p.row_src);
      const long long first = static_cast<long long>(blockIdx.x) * kNumGatherWarps + gw;
      int trace_it = 0;

      // slice index -> (row-tile, chunk, first row)
      auto locate = [&](long long sl, int& m_t, int& f, int& row0) {
        const int g = static_cast<int>(sl / group_slices);
        const int r = static_cast<int>(sl - g * group_slices);
        const int m_base = g * group_tiles;
        const int gsize = min(group_tiles, p.m_tiles - m_base) * kSlicesPerTile;
        f = r / gsize;
        const int rr = r - f * gsize;
        m_t = m_base + rr / kSlicesPerTile;
        row0 = m_t * BM + (rr % kSlicesPerTile) * kSliceRows;
      };
      auto issue = [&](uint4 (&v)[8], float& wv, int f, int row0) {
        if (f < p.slots) {
          // the wide (linear) weight of the same 16 ids rides along: one 4-byte load per row
          if (p.wide_src != nullptr && lane < kSliceRows && row0 + lane < p.M) {
            const unsigned long long wa = __ldg(p.wide_src + static_cast<long long>(row0 + lane) * p.slots + f);
            asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(wv) : "l"(wa));
          }
          // lanes 0..15 fetch the 16 row addresses once; every step broadcasts the one it needs, so
          // the eight 16-byte row loads of a lane are issued back to back.  Rows past M and null
          // addresses read a harmless local location instead of being predicated off: a predicated
          // load has to merge with the old register value, which serialises the loads.
          unsigned long long my_src = fallback_src;
          if (lane < kSliceRows && row0 + lane < p.M) {
            const unsigned long long a = __ldg(p.row_src + static_cast<long long>(row0 + lane) * p.slots + f);
            if (a != 0ull) my_src = a;
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int pc = u * 32 + lane;
            const unsigned long long src = __shfl_sync(0xffffffffu, my_src, pc >> 4);
            v[u] = ld_global_nc_v4(reinterpret_cast<const uint8_t*>(src) + (pc & 15) * 16);
          }
        } else {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int pc = u * 32 + lane;
            const int r = row0 + pc / max(dense_pieces, 1);
            if (pc < kSliceRows * dense_pieces && r < p.M)
              v[u] = ld_global_nc_v4(p.dense_x + static_cast<long long>(r) * p.dense_cols + (pc % dense_pieces) * 8);
          }
        }
      };
      auto retire = [&](const uint4 (&v)[8], float wv, int m_t, int f, int row0) {
        unsigned long long* tr = nullptr;
        if (p.trace && lane == 0 && trace_it < 8) {
          tr = p.trace + ((static_cast<long long>(blockIdx.x) * kNumGatherWarps + gw) * 8 + trace_it) * 4;
          tr[0] = globaltimer_ns();
        }
        if (f < p.slots) {
          if (p.wide_src != nullptr && lane < kSliceRows && row0 + lane < p.M)
            p.wide_out[static_cast<long long>(row0 + lane) * p.slots + f] = wv;
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int pc = u * 32 + lane;
            const int r = row0 + (pc >> 4);
            if (r < p.M)
              st_global_v4(reinterpret_cast<uint8_t*>(p.a_buf + static_cast<long long>(r) * p.lda_buf + f * 128) + (pc & 15) * 16,
                           v[u].x, v[u].y, v[u].z, v[u].w);
          }
        } else {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int pc = u * 32 + lane;
            const int r = row0 + pc / max(dense_pieces, 1);
            if (pc < kSliceRows * dense_pieces && r < p.M)
              st_global_v4(p.a_buf + static_cast<long long>(r) * p.lda_buf + p.slots * 128 + (pc % dense_pieces) * 8,
                           v[u].x, v[u].y, v[u].z, v[u].w);
          }
        }
        if (tr) tr[1] = globaltimer_ns();   // loads consumed, stores issued
        // __syncwarp orders the lanes' stores before lane 0's release, which is cumulative
        __syncwarp();
        if (lane == 0)
          asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(p.tile_flags + static_cast<long long>(m_t) * kFlagWords + (f >> 4)),
                       "l"(1ull << ((f & 15) * 4))
                       : "memory");
        if (tr) tr[2] = globaltimer_ns();
        ++trace_it;
      };

      if (first < total_slices) {
        uint4 va[8], vb[8];
        float wa = 0.f, wb = 0.f;
        int ma, fa, ra, mb, fb, rb;
        long long s0 = first;
        locate(s0, ma, fa, ra);
        issue(va, wa, fa, ra);
        for (;;) {
          const long long s1 = s0 + stride;
          const bool have1 = s1 < total_slices;
          if (have1) {
            locate(s1, mb, fb, rb);
            issue(vb, wb, fb, rb);
          }
          retire(va, wa, ma, fa, ra);
          if (!have1) break;
          s0 = s1 + stride;
          const bool have2 = s0 < total_slices;
          if (have2) {
            locate(s0, ma, fa, ra);
            issue(va, wa, fa, ra);
          }
          retire(vb, wb, mb, fb, rb);
          if (!have2) break;
        }
      }
      }  // gather warp
    }
  }

  tc_fence_before();
  if constexpr (kPair) cluster_sync_all();   // neither CTA may leave while the other still signals it
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if constexpr (kPair) tmem_dealloc_pair(tmem_base, kTmemCols);
    else tmem_dealloc(tmem_base, kTmemCols);
  }
  // Pair: leave together, after both cta_group::2 deallocations have been issued (a precaution that mirrors the
  // barrier before the allocation; the stall above was cured by that one, this one alone did not cure it).
  if constexpr (kPair) cluster_sync_all();
  if (p.dbg != nullptr && threadIdx.x == 0) atomicAdd(p.dbg + 16 + 4 * p.dbg_id + 2, 1ull);
}

}  // namespace dtf
