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
// Fused variants add gather warps (6..9) that pull embedding rows from peer HBM into the A
// buffer while the MMA warps work on earlier tiles, and an epilogue that scatters output rows to
// peer mailboxes (see fused_pull_gemm / fused_push_gemm instantiations).
#pragma once

#include "sm100_ptx.cuh"

namespace dtf {

constexpr int BM = 128;       // rows per CTA tile == UMMA M
constexpr int BK = 64;        // 64 bf16 = 128 B = one swizzle row
constexpr int UMMA_K = 16;    // fixed for 16-bit inputs
constexpr int kNumMainThreads = 192;
constexpr int kNumGatherThreads = 128;

enum Epilogue : int {
  EPI_BF16 = 0,            // D = bf16(acc)
  EPI_BIAS_BF16 = 1,       // D = bf16(acc + bias[n])
  EPI_BIAS_RELU_BF16 = 2,  // D = bf16(relu(acc + bias[n]))
  EPI_DRELU_BF16 = 3,      // D = bf16(aux[m,n] > 0 ? acc : 0)          (dgrad through ReLU)
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
  const __nv_bfloat16* aux;
  long long ldaux;
  // --- fused pull (gather warps): rows of 128 bf16 fetched through row_src pointers ---
  const unsigned long long* row_src;  // [M * slots] global addresses of 256-byte rows (may be peer)
  const __nv_bfloat16* dense_x;       // [M, dense_cols] local bf16 (already zero padded)
  __nv_bfloat16* a_buf;               // [M, lda_buf] gathered operand, consumed through tmA
  long long lda_buf;
  int slots;                          // F
  int dense_cols;                     // 64
  unsigned int* tile_flags;           // [m_tiles] arrival counters, zero before launch
  unsigned int* chunk_counter;        // [1] dynamic chunk claim, zero before launch
  // --- fused push epilogue: destination of the 128-wide bf16 row for (m, slot) ---
  const unsigned long long* row_dst;  // [M * slots] global addresses (may be peer mailbox)
};

template <int BN>
struct GemmSmem {
  static constexpr int kABytes = BM * BK * 2;  // 16 KB
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN == 256) ? 4 : ((BN == 128) ? 6 : 8);
  static constexpr int kBarBytes = 1024;
  static constexpr int kBiasBytes = BN * 4 * 2;
  static constexpr int kTotal = kStages * kStageBytes + kBarBytes + kBiasBytes + 1024 /*align*/;
};

// tile index -> (m, n, split); n fastest so that concurrently running CTAs share A row-tiles.
__device__ __forceinline__ void decode_tile(const GemmParams& p, int tile, int& m_t, int& n_t,
                                            int& s_t) {
  n_t = tile % p.n_tiles;
  int r = tile / p.n_tiles;
  m_t = r % p.m_tiles;
  s_t = r / p.m_tiles;
}

template <int BN, bool A_MN, bool B_MN, int EPI, bool GATHER>
__global__ void __launch_bounds__(GATHER ? kNumMainThreads + kNumGatherThreads : kNumMainThreads, 1)
gemm_sm100_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const GemmParams p) {
  using S = GemmSmem<BN>;
  constexpr int kStages = S::kStages;
  constexpr uint32_t kTmemCols = (2 * BN <= 32) ? 32 : 2 * BN;  // two accumulator buffers

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar_base = smem_base + kStages * S::kStageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
  auto tfull_bar = [&](int b) { return bar_base + 8u * (2 * kStages + b); };
  auto tempty_bar = [&](int b) { return bar_base + 8u * (2 * kStages + 2 + b); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kStages * S::kStageBytes + 8 * (2 * kStages + 4));
  float* bias_smem = reinterpret_cast<float*>(smem + kStages * S::kStageBytes + S::kBarBytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.m_tiles * p.n_tiles * p.splits;
  const int total_kb = (p.K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tfull_bar(b), 1);
      mbar_init(tempty_bar(b), 4);  // one arrival per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // =============================== TMA producer ===============================
    if (lane == 0) {
      uint32_t kiter = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_t, n_t, s_t;
        decode_tile(p, tile, m_t, n_t, s_t);
        const int kb0 = s_t * p.k_blocks_per_split;
        const int kb1 = min(total_kb, kb0 + p.k_blocks_per_split);
        if constexpr (GATHER) {
          // wait until every chunk of this row-tile has been gathered into a_buf
          const unsigned int need = static_cast<unsigned int>(p.slots + 1);
          unsigned int v;
          do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p.tile_flags + m_t) : "memory");
          } while (v < need);
          asm volatile("fence.proxy.async;" ::: "memory");  // generic-proxy writes -> TMA reads
        }
        for (int kb = kb0; kb < kb1; ++kb, ++kiter) {
          const int s = kiter % kStages;
          const uint32_t ph = (kiter / kStages) & 1;
          mbar_wait(empty_bar(s), ph ^ 1);
          mbar_arrive_expect_tx(full_bar(s), S::kStageBytes);
          const uint32_t sa = smem_base + s * S::kStageBytes;
          const uint32_t sb = sa + S::kABytes;
          if constexpr (!A_MN) {
            tma_load_2d(sa, &tmA, full_bar(s), kb * BK, m_t * BM);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_2d(sa + j * (BK * 128), &tmA, full_bar(s), m_t * BM + j * 64, kb * BK);
          }
          if constexpr (!B_MN) {
            tma_load_2d(sb, &tmB, full_bar(s), kb * BK, n_t * BN);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sb + j * (BK * 128), &tmB, full_bar(s), n_t * BN + j * 64, kb * BK);
          }
        }
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer ===============================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN, B_MN);
      uint32_t kiter = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        int m_t, n_t, s_t;
        decode_tile(p, tile, m_t, n_t, s_t);
        const int kb0 = s_t * p.k_blocks_per_split;
        const int kb1 = min(total_kb, kb0 + p.k_blocks_per_split);
        const int buf = it & 1;
        mbar_wait(tempty_bar(buf), ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        for (int kb = kb0; kb < kb1; ++kb, ++kiter) {
          const int s = kiter % kStages;
          const uint32_t ph = (kiter / kStages) & 1;
          mbar_wait(full_bar(s), ph);
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
    const int q = warp & 3;  // TMEM lane quarter this warp may touch
    const int et = threadIdx.x - 64;  // 0..127 within the epilogue group
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      int m_t, n_t, s_t;
      decode_tile(p, tile, m_t, n_t, s_t);
      const int buf = it & 1;
      const int n0 = n_t * BN;
      const int row = m_t * BM + q * 32 + lane;
      float* bsm = bias_smem + buf * BN;
      if constexpr (EPI == EPI_BIAS_BF16 || EPI == EPI_BIAS_RELU_BF16) {
        for (int i = et; i < BN; i += 128) bsm[i] = (n0 + i < p.N) ? __ldg(p.bias + n0 + i) : 0.f;
        named_bar_sync(1, 128);
      }
      mbar_wait(tfull_bar(buf), (it >> 1) & 1);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(t_row + c * 32, r);
        tmem_ld_wait();
        const int col = n0 + c * 32;
        if (row < p.M && col < p.N) {
          const bool full = (col + 32 <= p.N);
          if constexpr (EPI == EPI_F32 || EPI == EPI_F32_ATOMIC) {
            float* drow = reinterpret_cast<float*>(p.D) + static_cast<long long>(row) * p.ldd + col;
            if (full && ((reinterpret_cast<uintptr_t>(drow) & 15) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                if constexpr (EPI == EPI_F32_ATOMIC) {
                  red_add_f32x4(drow + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                                __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
                } else {
                  st_global_v4(drow + j, r[j], r[j + 1], r[j + 2], r[j + 3]);
                }
              }
            } else {
              for (int j = 0; j < 32 && col + j < p.N; ++j) {
                if constexpr (EPI == EPI_F32_ATOMIC) atomicAdd(drow + j, __uint_as_float(r[j]));
                else drow[j] = __uint_as_float(r[j]);
              }
            }
          } else {
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
            if constexpr (EPI == EPI_BIAS_BF16 || EPI == EPI_BIAS_RELU_BF16) {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                v[j] += bsm[c * 32 + j];
                if (EPI == EPI_BIAS_RELU_BF16) v[j] = fmaxf(v[j], 0.f);
              }
            }
            if constexpr (EPI == EPI_DRELU_BF16) {
              const __nv_bfloat16* arow = p.aux + static_cast<long long>(row) * p.ldaux + col;
              if (full && ((reinterpret_cast<uintptr_t>(arow) & 15) == 0)) {
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                  uint4 h = ld_global_nc_v4(arow + j);
                  const uint32_t hw[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    // post-ReLU activations are >= 0, so "> 0" is "low 15 bits or high 15 bits set"
                    if ((hw[e] & 0x7FFFu) == 0u || (hw[e] & 0x8000u)) v[j + 2 * e] = 0.f;
                    if ((hw[e] & 0x7FFF0000u) == 0u || (hw[e] & 0x80000000u)) v[j + 2 * e + 1] = 0.f;
                  }
                }
              } else {
                for (int j = 0; j < 32 && col + j < p.N; ++j)
                  if (!(__bfloat162float(arow[j]) > 0.f)) v[j] = 0.f;
              }
            }
            if constexpr (EPI == EPI_PUSH_ROWS) {
              // 128-wide row of slot f lives behind row_dst[row*slots + f]; this chunk covers
              // 32 of its columns.
              const int f = col >> 7;
              if (f < p.slots) {
                const unsigned long long dst = p.row_dst[static_cast<long long>(row) * p.slots + f];
                if (dst != 0ull) {
                  uint8_t* d8 = reinterpret_cast<uint8_t*>(dst) + (col & 127) * 2;
#pragma unroll
                  for (int j = 0; j < 32; j += 8)
                    st_global_v4(d8 + j * 2, pack_bf16x2(v[j], v[j + 1]), pack_bf16x2(v[j + 2], v[j + 3]),
                                 pack_bf16x2(v[j + 4], v[j + 5]), pack_bf16x2(v[j + 6], v[j + 7]));
                }
              }
            } else {
              __nv_bfloat16* drow =
                  reinterpret_cast<__nv_bfloat16*>(p.D) + static_cast<long long>(row) * p.ldd + col;
              if (full && ((reinterpret_cast<uintptr_t>(drow) & 15) == 0)) {
#pragma unroll
                for (int j = 0; j < 32; j += 8)
                  st_global_v4(drow + j, pack_bf16x2(v[j], v[j + 1]), pack_bf16x2(v[j + 2], v[j + 3]),
                               pack_bf16x2(v[j + 4], v[j + 5]), pack_bf16x2(v[j + 6], v[j + 7]));
              } else {
                for (int j = 0; j < 32 && col + j < p.N; ++j) drow[j] = __float2bfloat16(v[j]);
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(buf));
    }
  } else {
    // =============================== gather warps (fused pull) ===============================
    if constexpr (GATHER) {
      const int gt = threadIdx.x - kNumMainThreads;  // 0..127
      const int chunks_per_tile = p.slots + 1;
      const int total_chunks = p.m_tiles * chunks_per_tile;
      __shared__ int claimed;
      for (;;) {
        if (gt == 0) claimed = static_cast<int>(atomicAdd(p.chunk_counter, 1u));
        named_bar_sync(2, kNumGatherThreads);
        const int chunk = claimed;
        named_bar_sync(3, kNumGatherThreads);
        if (chunk >= total_chunks) break;
        const int m_t = chunk / chunks_per_tile;
        const int f = chunk % chunks_per_tile;
        const int row0 = m_t * BM;
        if (f < p.slots) {
          // 128 rows x 256 B = 2048 16-byte pieces; 16 per thread, 8 loads in flight at a time
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint4 v[8];
            bool ok[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              const int piece = (half * 8 + u) * kNumGatherThreads + gt;
              const int r = row0 + (piece >> 4);
              ok[u] = r < p.M;
              if (ok[u]) {
                const unsigned long long src = __ldg(p.row_src + static_cast<long long>(r) * p.slots + f);
                v[u] = ld_global_nc_v4(reinterpret_cast<const uint8_t*>(src) + (piece & 15) * 16);
              }
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              const int piece = (half * 8 + u) * kNumGatherThreads + gt;
              const int r = row0 + (piece >> 4);
              if (ok[u]) {
                uint8_t* d = reinterpret_cast<uint8_t*>(p.a_buf + static_cast<long long>(r) * p.lda_buf + f * 128) +
                             (piece & 15) * 16;
                st_global_v4(d, v[u].x, v[u].y, v[u].z, v[u].w);
              }
            }
          }
        } else {
          // dense feature block: dense_cols bf16 per row, local copy
          const int pieces_per_row = p.dense_cols / 8;
          for (int piece = gt; piece < BM * pieces_per_row; piece += kNumGatherThreads) {
            const int r = row0 + piece / pieces_per_row;
            const int c = piece % pieces_per_row;
            if (r < p.M) {
              uint4 v = ld_global_nc_v4(p.dense_x + static_cast<long long>(r) * p.dense_cols + c * 8);
              st_global_v4(p.a_buf + static_cast<long long>(r) * p.lda_buf + p.slots * 128 + c * 8, v.x, v.y, v.z, v.w);
            }
          }
        }
        __threadfence();
        named_bar_sync(2, kNumGatherThreads);
        if (gt == 0) {
          asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p.tile_flags + m_t) : "memory");
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

}  // namespace dtf
