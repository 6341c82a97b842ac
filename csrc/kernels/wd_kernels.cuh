// Device kernels of the wide-and-deep step that are not GEMMs: routing, loss / last layer / wide
// part, bias-gradient column sums, the parameter-server side applies (sparse scatter-add and dense
// reduce-scatter fused with SGD / Adagrad / Adam) and the cross-rank barrier.
//
// Reference call sites being replaced: variable reads and Apply* ops on the ps behind
// optimizer.minimize(loss, global_step) (reference trainer/task.py:137-139), use_locking=False
// (async) plus the sync-replica mode required by the north star.
#pragma once

#include "sm100_ptx.cuh"
#include "wd.cuh"

namespace dtf {

// ------------------------------------------------------------------------------------------ utils
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__device__ __forceinline__ float uniform_pm1(uint64_t key) {  // (-1, 1)
  return (static_cast<float>(splitmix64(key) >> 40) + 0.5f) * (2.0f / 16777216.0f) - 1.0f;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Device-resident step state so that a captured CUDA graph of the step stays valid across replays:
// the step counter, the per-step optimizer constants and the barrier epoch all live in HBM.
struct StepState {
  long long step;          // global_step of this shard (K7: folded into the apply)
  unsigned int epoch;      // barrier generation
  unsigned int error;      // != 0: a device barrier gave up waiting for rank (error - 1); sticky, read by the host
  OptConst oc;
};

__global__ void wd_tick_kernel(StepState* st, float base_lr, int opt) {
  const long long t = st->step + 1;
  st->step = t;
  OptConst oc;
  oc.beta1 = 0.9f; oc.beta2 = 0.999f; oc.eps = 1e-8f; oc.lr = base_lr;
  if (opt == WD_ADAM) {
    const double b1t = pow(0.9, static_cast<double>(t)), b2t = pow(0.999, static_cast<double>(t));
    oc.lr = static_cast<float>(base_lr * sqrt(1.0 - b2t) / (1.0 - b1t));
  }
  st->oc = oc;
}

// ------------------------------------------------------------------------------------------ init
__global__ void wd_init_embedding_kernel(float* __restrict__ master, __nv_bfloat16* __restrict__ work,
                                         float* __restrict__ slot1, float* __restrict__ slot2,
                                         float* __restrict__ wide, float* __restrict__ wide_s1,
                                         float* __restrict__ wide_s2, long long rows_local, int world, int rank,
                                         unsigned long long seed, float scale, float slot1_init) {
  const long long n = rows_local * kEmbDim;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long lrow = i / kEmbDim;
    const int col = static_cast<int>(i % kEmbDim);
    const unsigned long long grow = static_cast<unsigned long long>(lrow) * world + rank;
    const float v = scale * uniform_pm1(seed ^ (grow * 131ull + col));
    master[i] = v;
    work[i] = __float2bfloat16(v);
    if (slot1) slot1[i] = slot1_init;
    if (slot2) slot2[i] = 0.f;
    if (col == 0) {
      wide[lrow] = 0.f;  // TF's linear (wide) columns start at zero
      if (wide_s1) wide_s1[lrow] = slot1_init;
      if (wide_s2) wide_s2[lrow] = 0.f;
    }
  }
}

__global__ void wd_init_dense_kernel(float* __restrict__ master, float* __restrict__ s1, float* __restrict__ s2,
                                     long long lo, long long hi, DenseLayout dl, WdShapes sh, unsigned long long seed,
                                     float slot1_init) {
  for (long long i = lo + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < hi;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float fan_in = 0.f;
    if (i < dl.w2) fan_in = static_cast<float>(sh.K0);
    else if (i < dl.w3) fan_in = static_cast<float>(sh.h1);
    else if (i < dl.small_begin) fan_in = static_cast<float>(sh.h2);
    else if (i >= dl.w4 && i < dl.w4 + sh.h3) fan_in = static_cast<float>(sh.h3);
    float v = 0.f;
    if (fan_in > 0.f) v = rsqrtf(fan_in) * uniform_pm1((seed * 0x51ull) ^ static_cast<unsigned long long>(i));
    // padding between segments stays zero
    const bool in_w1 = i < dl.w1 + static_cast<long long>(sh.h1) * sh.K0;
    const bool in_w2 = i >= dl.w2 && i < dl.w2 + static_cast<long long>(sh.h2) * sh.h1;
    const bool in_w3 = i >= dl.w3 && i < dl.w3 + static_cast<long long>(sh.h3) * sh.h2;
    const bool in_w4 = i >= dl.w4 && i < dl.w4 + sh.h3;
    if (!(in_w1 || in_w2 || in_w3 || in_w4)) v = 0.f;
    master[i - lo] = v;
    if (s1) s1[i - lo] = slot1_init;
    if (s2) s2[i - lo] = 0.f;
  }
}

// owner -> every peer: bf16 working copy (+ fp32 copy of the small tail)
__global__ void wd_dense_broadcast_kernel(const float* __restrict__ master, long long lo, long long hi, PeerTable pt,
                                          SlabLayout sl, long long small_begin) {
  for (long long i = lo + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < hi;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float v = master[i - lo];
    const __nv_bfloat16 b = __float2bfloat16(v);
    for (int r = 0; r < pt.world; ++r) {
      reinterpret_cast<__nv_bfloat16*>(pt.slab[r] + sl.wbf16)[i] = b;
      if (i >= small_begin) reinterpret_cast<float*>(pt.slab[r] + sl.small_f32)[i - small_begin] = v;
    }
  }
}

// ------------------------------------------------------------------------------------------ route
// For every (example, slot): where to pull the embedding / wide weight from, and which mailbox entry
// at the owning shard receives this example's gradient.  Mailbox slots are claimed with one global
// atomic per (block, owner).
__global__ void wd_route_kernel(const int* __restrict__ ids, int n, long long V, PeerTable pt, SlabLayout sl,
                                long long cap, unsigned int* __restrict__ cnt,
                                unsigned long long* __restrict__ row_src, unsigned long long* __restrict__ wide_src,
                                unsigned long long* __restrict__ row_dst,
                                const float* __restrict__ dense_raw, __nv_bfloat16* __restrict__ dense_out, int B,
                                int dense_in, int dense_pad) {
  __shared__ unsigned int s_cnt[kMaxWorld];
  __shared__ unsigned int s_base[kMaxWorld];
  pdl_launch_dependents();
  pdl_wait();
  // rides along: the dense features as zero-padded bf16 (the last chunk of the GEMM-1 operand)
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B * dense_pad; i += gridDim.x * blockDim.x) {
    const int m = i / dense_pad, c = i % dense_pad;
    dense_out[i] = __float2bfloat16(c < dense_in ? dense_raw[m * dense_in + c] : 0.f);
  }
  if (threadIdx.x < kMaxWorld) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  int owner = 0;
  long long lrow = 0;
  unsigned int my = 0;
  if (e < n) {
    long long id = ids[e];
    id %= V;
    if (id < 0) id += V;
    owner = static_cast<int>(id % pt.world);
    lrow = id / pt.world;
    my = atomicAdd(&s_cnt[owner], 1u);
  }
  __syncthreads();
  if (threadIdx.x < pt.world && s_cnt[threadIdx.x] > 0)
    s_base[threadIdx.x] = atomicAdd(&cnt[threadIdx.x], s_cnt[threadIdx.x]);
  __syncthreads();
  if (e < n) {
    const unsigned long long entry = static_cast<unsigned long long>(pt.rank) * cap + s_base[owner] + my;
    uint8_t* oslab = pt.slab[owner];
    row_src[e] = reinterpret_cast<unsigned long long>(oslab + sl.emb_bf16) + static_cast<unsigned long long>(lrow) * (kEmbDim * 2);
    wide_src[e] = reinterpret_cast<unsigned long long>(oslab + sl.wide) + static_cast<unsigned long long>(lrow) * 4;
    row_dst[e] = reinterpret_cast<unsigned long long>(oslab + sl.mbx_payload) + entry * (kEmbDim * 2);
    // one 8-byte store per entry: where the gradient row goes at the owner, and which of this rank's examples it
    // belongs to (the owner looks the example's dlogit -- the wide gradient -- up in the vector the loss kernel sends)
    const long long example = e / (n / B);
    reinterpret_cast<long long*>(oslab + sl.mbx_rows)[entry] = lrow | (example << 32);
  }
}

// ------------------------------------------------------------------------------------------ prefetch
// Async mode may fetch the NEXT step's embedding rows while this step's backward pass keeps the tensor cores
// busy and NVLink idle (a parameter prefetch: the rows are one more step stale, which Hogwild semantics allow).
// Two small kernels on a low-priority stream: addresses (no mailbox claims -- those stay with the step itself),
// then a plain gather into the other parity's operand buffer.  The next step then runs GEMM-1 as a plain GEMM.
__global__ void wd_prefetch_addr_kernel(const int* __restrict__ ids, int n, long long V, PeerTable pt, SlabLayout sl,
                                        unsigned long long* __restrict__ row_src, unsigned long long* __restrict__ wide_src,
                                        const float* __restrict__ dense_raw, __nv_bfloat16* __restrict__ a_buf, long long lda,
                                        int dense_col0, int B, int dense_in, int dense_pad) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B * dense_pad; i += gridDim.x * blockDim.x) {
    const int m = i / dense_pad, c = i % dense_pad;
    a_buf[static_cast<long long>(m) * lda + dense_col0 + c] = __float2bfloat16(c < dense_in ? dense_raw[m * dense_in + c] : 0.f);
  }
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < n) {
    long long id = ids[e];
    id %= V;
    if (id < 0) id += V;
    const int owner = static_cast<int>(id % pt.world);
    const long long lrow = id / pt.world;
    uint8_t* oslab = pt.slab[owner];
    row_src[e] = reinterpret_cast<unsigned long long>(oslab + sl.emb_bf16) + static_cast<unsigned long long>(lrow) * (kEmbDim * 2);
    wide_src[e] = reinterpret_cast<unsigned long long>(oslab + sl.wide) + static_cast<unsigned long long>(lrow) * 4;
  }
}

// One warp per slice of 16 rows x one slot (256 B per row): 8 unpredicated 16-byte loads per lane from the owners'
// HBM (local or a peer's over NVLink), stored row-major into the GEMM-1 operand.  256 threads and few registers so
// that a block fits next to a persistent GEMM CTA (192 threads, up to 168 registers) on the same SM.
__global__ void __launch_bounds__(256)
wd_gather_rows_kernel(const unsigned long long* __restrict__ row_src, const unsigned long long* __restrict__ wide_src,
                      __nv_bfloat16* __restrict__ a_buf, long long lda, float* __restrict__ wide_out, int M, int F) {
  const int lane = threadIdx.x & 31;
  const long long warp = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const long long total = static_cast<long long>((M + 15) / 16) * F;
  const unsigned long long fallback = reinterpret_cast<unsigned long long>(row_src);   // harmless local address for rows past M
  for (long long s = warp; s < total; s += nwarps) {
    const int f = static_cast<int>(s % F);
    const int row0 = static_cast<int>(s / F) * 16;
    unsigned long long my_src = fallback;
    float wv = 0.f;
    const bool mine = lane < 16 && row0 + lane < M;
    if (mine) {
      const long long e = static_cast<long long>(row0 + lane) * F + f;
      my_src = __ldg(row_src + e);
      const unsigned long long wa = __ldg(wide_src + e);
      asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(wv) : "l"(wa));
    }
    uint4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int pc = u * 32 + lane;
      const unsigned long long src = __shfl_sync(0xffffffffu, my_src, pc >> 4);
      v[u] = ld_global_nc_v4(reinterpret_cast<const uint8_t*>(src) + (pc & 15) * 16);
    }
    if (mine) wide_out[static_cast<long long>(row0 + lane) * F + f] = wv;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int pc = u * 32 + lane;
      const int r = row0 + (pc >> 4);
      if (r < M)
        st_global_v4(reinterpret_cast<uint8_t*>(a_buf + static_cast<long long>(r) * lda + f * 128) + (pc & 15) * 16, v[u].x, v[u].y,
                     v[u].z, v[u].w);
    }
  }
}

// ------------------------------------------------------------------------------------------ small-batch forward
// Serving requests of a few examples (M <= 4 in practice): a 128-row tensor-core tile would be 1-3 % full and the layer is bound
// by reading W once, so the layer runs as a matrix-vector product on the CUDA cores with the roles swapped -- one
// warp per OUTPUT NEURON streams its row of W (16-byte loads) against the M activation rows held in shared memory,
// fp32 accumulation, warp reduction, bias / ReLU, bf16 out.  For layer 1 (X == nullptr) the block first gathers its
// operand itself: the embedding rows of the ids (addresses from wd_prefetch_addr_kernel, local or on a peer) and
// the padded dense features; block 0 also fetches the wide weights for the logit kernel.  K must be a multiple of 8.
template <int MAXM>
__global__ void __launch_bounds__(256)
wd_small_layer_kernel(const __nv_bfloat16* __restrict__ X, long long ldx, const unsigned long long* __restrict__ row_src,
                      const unsigned long long* __restrict__ wide_src, int F, const __nv_bfloat16* __restrict__ dense_x,
                      int dense_cols, float* __restrict__ wide_out, const __nv_bfloat16* __restrict__ W, long long ldw,
                      const float* __restrict__ bias, __nv_bfloat16* __restrict__ Y, long long ldy, int M, int N, int K,
                      int relu) {
  extern __shared__ uint4 s_x[];   // [M][K / 8]
  const int kc = K >> 3;
  if (X != nullptr) {
    for (int i = threadIdx.x; i < M * kc; i += blockDim.x) {
      const int m = i / kc, c = i - m * kc;
      s_x[i] = *reinterpret_cast<const uint4*>(X + static_cast<long long>(m) * ldx + c * 8);
    }
  } else {
    for (int i = threadIdx.x; i < M * kc; i += blockDim.x) {
      const int m = i / kc, c = i - m * kc;
      if (c < F * 16) {
        const unsigned long long src = __ldg(row_src + static_cast<long long>(m) * F + (c >> 4));
        s_x[i] = ld_global_nc_v4(reinterpret_cast<const uint8_t*>(src) + (c & 15) * 16);
      } else {
        s_x[i] = *reinterpret_cast<const uint4*>(dense_x + static_cast<long long>(m) * dense_cols + (c - F * 16) * 8);
      }
    }
    if (blockIdx.x == 0 && wide_out != nullptr)
      for (int i = threadIdx.x; i < M * F; i += blockDim.x) {
        float v;
        asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(__ldg(wide_src + i)));
        wide_out[i] = v;
      }
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; n < N; n += warps) {
    float acc[MAXM];
#pragma unroll
    for (int m = 0; m < MAXM; ++m) acc[m] = 0.f;
    const __nv_bfloat16* wrow = W + static_cast<long long>(n) * ldw;
#pragma unroll 8
    for (int c = lane; c < kc; c += 32) {
      const uint4 wq = ld_global_nc_v4(wrow + c * 8);
      const __nv_bfloat162* wp = reinterpret_cast<const __nv_bfloat162*>(&wq);
      float2 wf[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) wf[j] = __bfloat1622float2(wp[j]);
#pragma unroll
      for (int m = 0; m < MAXM; ++m) {
        if (m < M) {
          const uint4 xq = s_x[m * kc + c];
          const __nv_bfloat162* xp = reinterpret_cast<const __nv_bfloat162*>(&xq);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 xf = __bfloat1622float2(xp[j]);
            acc[m] = fmaf(wf[j].x, xf.x, acc[m]);
            acc[m] = fmaf(wf[j].y, xf.y, acc[m]);
          }
        }
      }
    }
    const float bn = bias != nullptr ? __ldg(bias + n) : 0.f;
#pragma unroll
    for (int m = 0; m < MAXM; ++m) {
      if (m < M) {
        float v = warp_sum(acc[m]) + bn;
        if (relu) v = fmaxf(v, 0.f);
        if (lane == 0) Y[static_cast<long long>(m) * ldy + n] = __float2bfloat16(v);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------ loss
// One warp per example: sum of the wide weights (gathered from the owning shards by the fused pull
// kernel's gather warps), last layer h3 . w4,
// sigmoid cross-entropy, dlogit; then (training) the wide gradient is written straight into the
// owners' mailboxes, dZ3 = dlogit * w4 masked by ReLU'(h3), and block-reduced gradients of w4 / b4 /
// wide bias.  SURVEY.md 2E rows K5, K6.
// MAXC: register arrays are sized for H3 <= 256 * MAXC (1 for the product shape: 126 -> ~70 registers)
template <bool TRAIN, int MAXC>
__global__ void __launch_bounds__(256)
wd_loss_kernel(const __nv_bfloat16* __restrict__ h3, const uint32_t* __restrict__ mask3, long long ldmask,
               const float* __restrict__ small, int off_w4, int off_b4, int off_wide_b,
               const float* __restrict__ wide_vals, PeerTable pt, SlabLayout sl,
               const float* __restrict__ labels, int B, int F, int H3, float inv_batch,
               float* __restrict__ dlogit, __nv_bfloat16* __restrict__ dz3, float* __restrict__ g_small,
               float* __restrict__ loss_acc, float* __restrict__ prob_out) {
  extern __shared__ float s_red[];  // [8][H3 + 2] (train only)
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int chunks = H3 / 256;  // 8 elements per lane per chunk
  float gw4[MAXC][8];
#pragma unroll
  for (int c = 0; c < MAXC; ++c)
#pragma unroll
    for (int j = 0; j < 8; ++j) gw4[c][j] = 0.f;
  float g_b = 0.f, loss_sum = 0.f;

  for (int m = blockIdx.x * (blockDim.x >> 5) + warp; m < B; m += nwarps) {
    float wide = 0.f;
    // the wide weights of this example's ids were fetched from their owners by the fused pull
    for (int f = lane; f < F; f += 32) wide += wide_vals[static_cast<long long>(m) * F + f];
    float dot = 0.f;
    float hv[MAXC][8];
#pragma unroll
    for (int c = 0; c < MAXC; ++c) {
      if (c < chunks) {
        const int col = c * 256 + lane * 8;
        const uint4 raw = *reinterpret_cast<const uint4*>(h3 + static_cast<long long>(m) * H3 + col);
        const __nv_bfloat162* hp = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 f2 = __bfloat1622float2(hp[j]);
          hv[c][2 * j] = f2.x;
          hv[c][2 * j + 1] = f2.y;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) dot += hv[c][j] * small[off_w4 + col + j];
      }
    }
    dot = warp_sum(dot);
    wide = warp_sum(wide);
    const float logit = dot + small[off_b4] + wide + small[off_wide_b];
    const float p = 1.f / (1.f + __expf(-logit));
    if (!TRAIN) {
      if (lane == 0) prob_out[m] = p;
      continue;
    }
    const float y = labels[m];
    const float dl = (p - y) * inv_batch;
    if (lane == 0) {
      dlogit[m] = dl;
      loss_sum += fmaxf(logit, 0.f) - logit * y + log1pf(__expf(-fabsf(logit)));
      g_b += dl;
    }
    // wide gradient push: dlogit of this example goes ONCE to every shard (lane r -> rank r) instead of once per
    // (example, slot) entry; the owners pick it up by the example index stored with each mailbox entry
    if (lane < pt.world) reinterpret_cast<float*>(pt.slab[lane] + sl.mbx_dlvec)[static_cast<long long>(pt.rank) * B + m] = dl;
#pragma unroll
    for (int c = 0; c < MAXC; ++c) {
      if (c < chunks) {
        const int col = c * 256 + lane * 8;
        const uint32_t mword = mask3[static_cast<long long>(m) * ldmask + (col >> 5)];
        const uint32_t bits = (mword >> (col & 31)) & 0xFFu;
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          gw4[c][j] += dl * hv[c][j];
          o[j] = ((bits >> j) & 1u) ? dl * small[off_w4 + col + j] : 0.f;
        }
        uint4 packed;
        packed.x = pack_bf16x2(o[0], o[1]);
        packed.y = pack_bf16x2(o[2], o[3]);
        packed.z = pack_bf16x2(o[4], o[5]);
        packed.w = pack_bf16x2(o[6], o[7]);
        *reinterpret_cast<uint4*>(dz3 + static_cast<long long>(m) * H3 + col) = packed;
      }
    }
  }
  if (!TRAIN) return;
  // block reduction of the w4 / b4 / loss partials, one atomic per column per block
  float* mine = s_red + warp * (H3 + 2);
#pragma unroll
  for (int c = 0; c < MAXC; ++c)
    if (c < chunks)
#pragma unroll
      for (int j = 0; j < 8; ++j) mine[c * 256 + lane * 8 + j] = gw4[c][j];
  if (lane == 0) {
    mine[H3] = g_b;
    mine[H3 + 1] = loss_sum;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < H3 + 2; i += blockDim.x) {
    float s = 0.f;
    for (int w = 0; w < (blockDim.x >> 5); ++w) s += s_red[w * (H3 + 2) + i];
    if (i < H3) atomicAdd(g_small + off_w4 + i, s);
    else if (i == H3) {
      atomicAdd(g_small + off_b4, s);
      atomicAdd(g_small + off_wide_b, s);
    } else atomicAdd(loss_acc, s);
  }
}

// ------------------------------------------------------------------------------------------ colsum
// out[n] += sum_m x[m, n]  (bias gradients) for up to three matrices in one launch.  A block covers
// 256 columns x `rows_per_block` rows of one matrix; every thread streams 8 columns (one 16-byte load
// per row) for one of 8 row groups.
struct ColsumJobs {
  const __nv_bfloat16* x[3];
  float* out[3];
  int N[3];
  int block_end[3];   // exclusive prefix sums of the blocks per job
  int M;
  int rows_per_block;
};

__global__ void __launch_bounds__(256)
wd_colsum_kernel(ColsumJobs jobs) {
  __shared__ float s[8][256 + 8];
  pdl_launch_dependents();
  pdl_wait();
  int j = 0;
  while (j < 2 && static_cast<int>(blockIdx.x) >= jobs.block_end[j]) ++j;
  const int local = blockIdx.x - (j == 0 ? 0 : jobs.block_end[j - 1]);
  const int N = jobs.N[j], M = jobs.M;
  const long long ld = N;
  const __nv_bfloat16* __restrict__ x = jobs.x[j];
  const int col_blocks = (N + 255) / 256;
  const int bx = local % col_blocks, by = local / col_blocks;
  const int cg = threadIdx.x & 31, rg = threadIdx.x >> 5;
  const int col = bx * 256 + cg * 8;
  const int r0 = by * jobs.rows_per_block, r1 = min(M, r0 + jobs.rows_per_block);
  float acc[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) acc[q] = 0.f;
  if (col + 8 <= N) {
    for (int r = r0 + rg; r < r1; r += 8) {
      const uint4 raw = *reinterpret_cast<const uint4*>(x + static_cast<long long>(r) * ld + col);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 f = __bfloat1622float2(h[q]);
        acc[2 * q] += f.x;
        acc[2 * q + 1] += f.y;
      }
    }
  } else {
    for (int r = r0 + rg; r < r1; r += 8)
      for (int q = 0; q < 8 && col + q < N; ++q) acc[q] += __bfloat162float(x[static_cast<long long>(r) * ld + col + q]);
  }
#pragma unroll
  for (int q = 0; q < 8; ++q) s[rg][cg * 8 + q] = acc[q];
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) t += s[w][threadIdx.x];
  const int c = bx * 256 + threadIdx.x;
  if (c < N) atomicAdd(jobs.out[j] + c, t);
}

// ------------------------------------------------------------------------------------------ sparse
// Mailbox entries of all sources, flattened: entry g belongs to source s with prefix[s] <= g < prefix[s+1].
struct MailboxView {
  const __nv_bfloat16* payload;  // [world, cap, 128]
  const long long* rows;         // [world, cap]  local row (low 32 bits) | example index at the source (high 32)
  const float* dlvec;            // [world, B]    dlogit of every example of each source
  int B;
  const unsigned int* count;     // [world]
  long long cap;
  int world;
};

__device__ __forceinline__ bool mailbox_locate(const MailboxView& mv, const unsigned int* s_prefix, long long g,
                                               long long& entry) {
  int s = 0;
  while (s < mv.world && g >= s_prefix[s + 1]) ++s;
  if (s >= mv.world) return false;
  entry = static_cast<long long>(s) * mv.cap + (g - s_prefix[s]);
  return true;
}

// async (Hogwild): every mailbox entry is applied on its own, straight onto the shard, with no
// duplicate handling and no locks -- concurrent entries for one row race exactly like
// use_locking=False in the reference's Apply* ops.
template <int OPT, bool STREAM>
__global__ void __launch_bounds__(256)
wd_sparse_apply_entries_kernel(MailboxView mv, float* __restrict__ master, float* __restrict__ s1,
                               float* __restrict__ s2, __nv_bfloat16* __restrict__ work, float* __restrict__ wide,
                               float* __restrict__ wide_s1, float* __restrict__ wide_s2, const StepState* __restrict__ stp) {
  const OptConst oc = stp->oc;
  __shared__ unsigned int s_prefix[kMaxWorld + 1];
  if (threadIdx.x == 0) {
    unsigned int acc = 0;
    for (int s = 0; s < mv.world; ++s) {
      s_prefix[s] = acc;
      acc += min(mv.count[s], static_cast<unsigned int>(mv.cap));
    }
    for (int s = mv.world; s <= kMaxWorld; ++s) s_prefix[s] = acc;
  }
  __syncthreads();
  const long long total = s_prefix[mv.world];
  const int lane = threadIdx.x & 31;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const unsigned long long pol = STREAM ? evict_first_policy() : 0ull;
  // two entries per iteration: both rows' loads are in flight before either is updated
  for (long long gi = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; gi < total; gi += 2 * nwarps) {
    long long e[2], row[2] = {0, 0}, o[2] = {0, 0}, ex[2] = {0, 0};
    bool ok[2];
    uint2 raw[2];
    float4 w[2], a[2], v[2];
    float dlv[2], ww[2], wa[2], wv[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long g = gi + u * nwarps;
      ok[u] = g < total && mailbox_locate(mv, s_prefix, g, e[u]);
      if (ok[u]) {
        const long long packed = mv.rows[e[u]];
        row[u] = packed & 0xFFFFFFFFll;
        ex[u] = (e[u] / mv.cap) * mv.B + (packed >> 32);      // [source, example] in the dlogit vectors
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      a[u] = make_float4(0, 0, 0, 0);
      v[u] = make_float4(0, 0, 0, 0);
      wa[u] = wv[u] = 0.f;
      if (!ok[u]) continue;
      o[u] = row[u] * kEmbDim + lane * 4;
      if (STREAM) {
        raw[u] = ld_stream_u2(mv.payload + e[u] * kEmbDim + lane * 4, pol);
        w[u] = ld_stream_f4(master + o[u], pol);
        if (OPT != WD_SGD) a[u] = ld_stream_f4(s1 + o[u], pol);
        if (opt_has_slot2(OPT)) v[u] = ld_stream_f4(s2 + o[u], pol);
      } else {
        raw[u] = *reinterpret_cast<const uint2*>(mv.payload + e[u] * kEmbDim + lane * 4);
        w[u] = *reinterpret_cast<float4*>(master + o[u]);
        if (OPT != WD_SGD) a[u] = *reinterpret_cast<float4*>(s1 + o[u]);
        if (opt_has_slot2(OPT)) v[u] = *reinterpret_cast<float4*>(s2 + o[u]);
      }
      if (lane == 0) {
        dlv[u] = mv.dlvec[ex[u]];
        ww[u] = wide[row[u]];
        if (OPT != WD_SGD) wa[u] = wide_s1[row[u]];
        if (opt_has_slot2(OPT)) wv[u] = wide_s2[row[u]];
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (!ok[u]) continue;
      const float2 ga = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw[u].x));
      const float2 gb = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw[u].y));
      opt_update<OPT>(w[u].x, ga.x, a[u].x, v[u].x, oc);
      opt_update<OPT>(w[u].y, ga.y, a[u].y, v[u].y, oc);
      opt_update<OPT>(w[u].z, gb.x, a[u].z, v[u].z, oc);
      opt_update<OPT>(w[u].w, gb.y, a[u].w, v[u].w, oc);
      uint2 pk;
      pk.x = pack_bf16x2(w[u].x, w[u].y);
      pk.y = pack_bf16x2(w[u].z, w[u].w);
      if (STREAM) {
        st_stream_f4(master + o[u], w[u], pol);
        if (OPT != WD_SGD) st_stream_f4(s1 + o[u], a[u], pol);
        if (opt_has_slot2(OPT)) st_stream_f4(s2 + o[u], v[u], pol);
        st_stream_u2(work + o[u], pk, pol);
      } else {
        *reinterpret_cast<float4*>(master + o[u]) = w[u];
        if (OPT != WD_SGD) *reinterpret_cast<float4*>(s1 + o[u]) = a[u];
        if (opt_has_slot2(OPT)) *reinterpret_cast<float4*>(s2 + o[u]) = v[u];
        *reinterpret_cast<uint2*>(work + o[u]) = pk;
      }
      if (lane == 0) {
        opt_update<OPT>(ww[u], dlv[u], wa[u], wv[u], oc);
        wide[row[u]] = ww[u];
        if (OPT != WD_SGD) wide_s1[row[u]] = wa[u];
        if (opt_has_slot2(OPT)) wide_s2[row[u]] = wv[u];
      }
    }
  }
}

// ---- sync-replica, version 2: count first, then apply in one pass ---------------------------------
// row_state[row] (64 bit): low word = number of mailbox entries of this step that target the row,
// high word = how many of them have been folded in so far.  Rows with a single contribution
// (the common case) are updated straight from the mailbox; contributions to a shared row are summed
// in G with fp32 reductions and whichever warp arrives last performs the one optimizer step.
__global__ void __launch_bounds__(256)
wd_sparse_count_kernel(MailboxView mv, unsigned long long* __restrict__ row_state) {
  __shared__ unsigned int s_prefix[kMaxWorld + 1];
  if (threadIdx.x == 0) {
    unsigned int acc = 0;
    for (int s = 0; s < mv.world; ++s) {
      s_prefix[s] = acc;
      acc += min(mv.count[s], static_cast<unsigned int>(mv.cap));
    }
    for (int s = mv.world; s <= kMaxWorld; ++s) s_prefix[s] = acc;
  }
  __syncthreads();
  const long long total = s_prefix[mv.world];
  for (long long g = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; g < total;
       g += static_cast<long long>(gridDim.x) * blockDim.x) {
    long long e;
    if (mailbox_locate(mv, s_prefix, g, e)) atomicAdd(row_state + (mv.rows[e] & 0xFFFFFFFFll), 1ull);
  }
}

template <int OPT>
__global__ void __launch_bounds__(256)
wd_sparse_apply_sync_kernel(MailboxView mv, unsigned long long* __restrict__ row_state, float* __restrict__ G,
                            float* __restrict__ wide_G, float* __restrict__ master, float* __restrict__ s1,
                            float* __restrict__ s2, __nv_bfloat16* __restrict__ work, float* __restrict__ wide,
                            float* __restrict__ wide_s1, float* __restrict__ wide_s2, const StepState* __restrict__ stp) {
  const OptConst oc = stp->oc;
  __shared__ unsigned int s_prefix[kMaxWorld + 1];
  if (threadIdx.x == 0) {
    unsigned int acc = 0;
    for (int s = 0; s < mv.world; ++s) {
      s_prefix[s] = acc;
      acc += min(mv.count[s], static_cast<unsigned int>(mv.cap));
    }
    for (int s = mv.world; s <= kMaxWorld; ++s) s_prefix[s] = acc;
  }
  __syncthreads();
  const long long total = s_prefix[mv.world];
  const int lane = threadIdx.x & 31;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  // two entries per iteration; every load that only depends on the row id (gradient, weights,
  // slots, row_state) is issued for both before anything is consumed
  for (long long gi = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; gi < total; gi += 2 * nwarps) {
    long long e[2], row[2] = {0, 0}, o[2] = {0, 0}, ex[2] = {0, 0};
    bool ok[2];
    uint2 raw[2];
    float4 w[2], a[2], v[2];
    float dlv[2] = {0.f, 0.f}, ww[2] = {0.f, 0.f}, wa[2] = {0.f, 0.f}, wv[2] = {0.f, 0.f};
    unsigned int n[2] = {1u, 1u};
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long g = gi + u * nwarps;
      ok[u] = g < total && mailbox_locate(mv, s_prefix, g, e[u]);
      if (ok[u]) {
        const long long packed = mv.rows[e[u]];
        row[u] = packed & 0xFFFFFFFFll;
        ex[u] = (e[u] / mv.cap) * mv.B + (packed >> 32);      // [source, example] in the dlogit vectors
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      a[u] = make_float4(0, 0, 0, 0);
      v[u] = make_float4(0, 0, 0, 0);
      if (!ok[u]) continue;
      o[u] = row[u] * kEmbDim + lane * 4;
      n[u] = static_cast<unsigned int>(__ldcg(row_state + row[u]) & 0xFFFFFFFFull);
      raw[u] = *reinterpret_cast<const uint2*>(mv.payload + e[u] * kEmbDim + lane * 4);
      w[u] = *reinterpret_cast<float4*>(master + o[u]);
      if (OPT != WD_SGD) a[u] = *reinterpret_cast<float4*>(s1 + o[u]);
      if (opt_has_slot2(OPT)) v[u] = *reinterpret_cast<float4*>(s2 + o[u]);
      if (lane == 0) {
        dlv[u] = mv.dlvec[ex[u]];
        ww[u] = wide[row[u]];
        if (OPT != WD_SGD) wa[u] = wide_s1[row[u]];
        if (opt_has_slot2(OPT)) wv[u] = wide_s2[row[u]];
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (!ok[u]) continue;
      const float2 ga = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw[u].x));
      const float2 gb = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw[u].y));
      float4 g = make_float4(ga.x, ga.y, gb.x, gb.y);
      float gw = dlv[u];
      if (n[u] > 1) {
        // shared row: fold this contribution into G, the last arriver takes the sum.  Weights and
        // slots of a shared row are only touched by that last arriver, so the copies loaded above
        // are still current when it gets here.
        red_add_f32x4(G + o[u], g.x, g.y, g.z, g.w);
        if (lane == 0) atomicAdd(wide_G + row[u], gw);
        __threadfence();
        __syncwarp();
        unsigned int arrived = 0;
        if (lane == 0) arrived = static_cast<unsigned int>(atomicAdd(row_state + row[u], 1ull << 32) >> 32);
        arrived = __shfl_sync(0xffffffffu, arrived, 0);
        if (arrived + 1 != n[u]) continue;
        __threadfence();
        g = __ldcg(reinterpret_cast<const float4*>(G + o[u]));
        gw = __shfl_sync(0xffffffffu, lane == 0 ? __ldcg(wide_G + row[u]) : 0.f, 0);
        *reinterpret_cast<float4*>(G + o[u]) = make_float4(0, 0, 0, 0);
        if (lane == 0) wide_G[row[u]] = 0.f;
      }
      opt_update<OPT>(w[u].x, g.x, a[u].x, v[u].x, oc);
      opt_update<OPT>(w[u].y, g.y, a[u].y, v[u].y, oc);
      opt_update<OPT>(w[u].z, g.z, a[u].z, v[u].z, oc);
      opt_update<OPT>(w[u].w, g.w, a[u].w, v[u].w, oc);
      *reinterpret_cast<float4*>(master + o[u]) = w[u];
      if (OPT != WD_SGD) *reinterpret_cast<float4*>(s1 + o[u]) = a[u];
      if (opt_has_slot2(OPT)) *reinterpret_cast<float4*>(s2 + o[u]) = v[u];
      uint2 pk;
      pk.x = pack_bf16x2(w[u].x, w[u].y);
      pk.y = pack_bf16x2(w[u].z, w[u].w);
      *reinterpret_cast<uint2*>(work + o[u]) = pk;
      if (lane == 0) {
        opt_update<OPT>(ww[u], gw, wa[u], wv[u], oc);
        wide[row[u]] = ww[u];
        if (OPT != WD_SGD) wide_s1[row[u]] = wa[u];
        if (opt_has_slot2(OPT)) wide_s2[row[u]] = wv[u];
        row_state[row[u]] = 0ull;   // ready for the next step
      }
    }
  }
}

// ------------------------------------------------------------------------------------------ dense
// Owner of [lo, hi): gradient = sum over `nsrc` ranks' gradient buffers read over NVLink
// (reduce-scatter), optimizer step on the fp32 master + slots, bf16 (and fp32 tail) written into
// every peer's working copy (all-gather), all in one pass.  SURVEY.md 2E row K2.
// WMAX: compile-time bound on the number of gradient replicas held in registers (1 when the sum comes
// from multimem.ld_reduce); U: independent float4 groups per thread, all of whose loads are issued
// before anything is consumed (the kernel is a pure stream: ~22 bytes moved per parameter).
template <int OPT, int WMAX, int U>
__global__ void __launch_bounds__(256)
wd_dense_apply_kernel(PeerTable pt, SlabLayout sl, long long lo, long long hi, int sum_first,
                      float* __restrict__ master, float* __restrict__ s1, float* __restrict__ s2,
                      const StepState* __restrict__ stp, long long small_begin, uint8_t* __restrict__ mc) {
  // mc != nullptr: multicast mapping of the slabs (NVLS).  The sum over the replicas' gradients is
  // then one multimem.ld_reduce (added inside NVSwitch) instead of `world` peer loads, and the new
  // values reach every rank's working copy with one multimem.st instead of `world` peer stores.
  // sum_first = 1 (sync-replica): one optimizer step on the sum of all replicas' gradients.
  // sum_first = 0 (async / Hogwild): every replica's gradient is applied on its own, in rank order,
  //   exactly as if each had been sent separately; all of them are read up front so that the
  //   NVLink loads overlap.
  const OptConst oc = stp->oc;
  const long long n4 = (hi - lo) >> 2;
  const bool use_mc_sum = mc != nullptr && sum_first;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long q0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; q0 < n4; q0 += stride * U) {
    float4 g[U][WMAX], w[U], a[U], v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long q = q0 + u * stride;
      a[u] = make_float4(0, 0, 0, 0);
      v[u] = make_float4(0, 0, 0, 0);
      if (q >= n4) continue;
      const long long i = lo + q * 4;
      if (use_mc_sum) {
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(g[u][0].x), "=f"(g[u][0].y), "=f"(g[u][0].z), "=f"(g[u][0].w)
                     : "l"(mc + sl.dense_grad + i * 4)
                     : "memory");
      } else {
#pragma unroll
        for (int r = 0; r < WMAX; ++r)
          if (r < pt.world) g[u][r] = *reinterpret_cast<const float4*>(pt.slab[r] + sl.dense_grad + i * 4);
      }
      w[u] = *reinterpret_cast<float4*>(master + (i - lo));
      if (OPT != WD_SGD) a[u] = *reinterpret_cast<float4*>(s1 + (i - lo));
      if (opt_has_slot2(OPT)) v[u] = *reinterpret_cast<float4*>(s2 + (i - lo));
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long q = q0 + u * stride;
      if (q >= n4) continue;
      const long long i = lo + q * 4;
      if (sum_first) {
        float4 t = g[u][0];
        if (!use_mc_sum) {
#pragma unroll
          for (int r = 1; r < WMAX; ++r)
            if (r < pt.world) { t.x += g[u][r].x; t.y += g[u][r].y; t.z += g[u][r].z; t.w += g[u][r].w; }
        }
        opt_update<OPT>(w[u].x, t.x, a[u].x, v[u].x, oc);
        opt_update<OPT>(w[u].y, t.y, a[u].y, v[u].y, oc);
        opt_update<OPT>(w[u].z, t.z, a[u].z, v[u].z, oc);
        opt_update<OPT>(w[u].w, t.w, a[u].w, v[u].w, oc);
      } else {
#pragma unroll
        for (int r = 0; r < WMAX; ++r)
          if (r < pt.world) {
            opt_update<OPT>(w[u].x, g[u][r].x, a[u].x, v[u].x, oc);
            opt_update<OPT>(w[u].y, g[u][r].y, a[u].y, v[u].y, oc);
            opt_update<OPT>(w[u].z, g[u][r].z, a[u].z, v[u].z, oc);
            opt_update<OPT>(w[u].w, g[u][r].w, a[u].w, v[u].w, oc);
          }
      }
      *reinterpret_cast<float4*>(master + (i - lo)) = w[u];
      if (OPT != WD_SGD) *reinterpret_cast<float4*>(s1 + (i - lo)) = a[u];
      if (opt_has_slot2(OPT)) *reinterpret_cast<float4*>(s2 + (i - lo)) = v[u];
      uint2 pk;
      pk.x = pack_bf16x2(w[u].x, w[u].y);
      pk.y = pack_bf16x2(w[u].z, w[u].w);
      if (mc != nullptr) {
        asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(mc + sl.wbf16 + i * 2),
                     "f"(__uint_as_float(pk.x)), "f"(__uint_as_float(pk.y))
                     : "memory");
        if (i >= small_begin)
          asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + sl.small_f32 +
                                                                                              (i - small_begin) * 4),
                       "f"(w[u].x), "f"(w[u].y), "f"(w[u].z), "f"(w[u].w)
                       : "memory");
      } else {
        for (int r = 0; r < pt.world; ++r) {
          *reinterpret_cast<uint2*>(pt.slab[r] + sl.wbf16 + i * 2) = pk;
          if (i >= small_begin) *reinterpret_cast<float4*>(pt.slab[r] + sl.small_f32 + (i - small_begin) * 4) = w[u];
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------ barrier
// Device-side barrier over the signal pads (one 64-byte line per source on every rank).  Thread r
// optionally publishes how many mailbox entries this rank pushed to rank r, then release-stores
// `epoch` into rank r's pad and acquire-spins on its own pad until rank r has done the same.
__global__ void wd_barrier_kernel(PeerTable pt, SlabLayout sl, StepState* stp, const unsigned int* publish_cnt,
                                  unsigned long long timeout_ns) {
  __shared__ unsigned int s_epoch;
  if (threadIdx.x == 0) {
    s_epoch = stp->epoch + 1;
    stp->epoch = s_epoch;
  }
  __syncthreads();
  const unsigned int epoch = s_epoch;
  const int r = threadIdx.x;
  if (r >= pt.world) return;
  if (publish_cnt != nullptr) reinterpret_cast<unsigned int*>(pt.slab[r] + sl.mbx_count)[pt.rank] = publish_cnt[r];
  __threadfence_system();
  unsigned int* remote = reinterpret_cast<unsigned int*>(pt.slab[r] + sl.signal) + pt.rank * 16;
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(remote), "r"(epoch) : "memory");
  const unsigned int* mine = reinterpret_cast<const unsigned int*>(pt.slab[pt.rank] + sl.signal) + r * 16;
  // Bounded wait: a rank that died or fell out of the job never signals.  After
  // `timeout_ns` the waiter records which rank it missed in StepState::error and gives up; once the flag
  // is set every later barrier returns at once, so the host, which polls the flag (dtf_wd_check_error),
  // gets control back within one timeout instead of a silent hang of all ranks.
  if (*reinterpret_cast<volatile unsigned int*>(&stp->error) != 0u) return;
  unsigned int v;
  unsigned long long t0 = 0;
  for (unsigned int spin = 0;; ++spin) {
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
    if (static_cast<int>(v - epoch) >= 0) break;
    if (timeout_ns != 0ull && (spin & 255u) == 255u) {
      const unsigned long long t = globaltimer_ns();
      if (t0 == 0ull) t0 = t;
      else if (t - t0 > timeout_ns) {
        atomicCAS(&stp->error, 0u, static_cast<unsigned int>(r) + 1u);
        break;
      }
    }
  }
}

// single-rank stand-in for the publish half of the barrier (counts stay local)
__global__ void wd_publish_local_kernel(PeerTable pt, SlabLayout sl, const unsigned int* publish_cnt) {
  const int r = threadIdx.x;
  if (r < pt.world) reinterpret_cast<unsigned int*>(pt.slab[r] + sl.mbx_count)[pt.rank] = publish_cnt[r];
}

}  // namespace dtf
