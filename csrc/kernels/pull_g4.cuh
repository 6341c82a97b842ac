// Fused embedding pull + GEMM-1, second design: the TMA unit gathers the rows.
//
//   H[M, N] = relu(A[M, K0] W[N, K0]^T + bias),   A[m, f*128 : f*128+128] = table row idx[f][m],  A[m, F*128 :] = dense_x[m]
//
// Replaces the variable reads of `X*w + b` behind sess.run(train_op) in the reference (reference
// trainer/task.py:137,236) for the wide-and-deep model.
//
// Every rank's slab is mapped back to back in one virtual-address window and ONE 2-D tensor map covers the
// window as a [rows, 128] bf16 matrix, so an embedding row on any rank is just a row index.
// `cp.async.bulk.tensor.2d ... tile::gather4` fetches four arbitrary rows x 64 columns per instruction
// straight into the 128B-swizzled A stage that tcgen05.mma reads: no gather warps, no registers, no
// staging of the operand in global memory on the way to the tensor core.  Measured (bench/gather4_probe.cu):
// 2.4-3.1 TB/s on local rows, 688 GB/s = link rate on rows of a peer.
//
// One 128-row tile needs N = 1024 accumulator columns, twice what one SM's TMEM holds, so a CLUSTER OF TWO
// CTAs owns a row-tile: CTA r computes columns [r*NH, r*NH + NH) (NH = N/2 = 512 fills its TMEM).  Both need
// the same A tile but every row must cross NVLink once: CTA r gathers the k-blocks with kb % 2 == r into its
// own stage and, when they have landed, forwards the 16 KB stage to its sibling with a DSMEM bulk copy
// (cp.async.bulk.shared::cluster.shared::cta) that completes on the sibling's barrier.  TMA multicast is not
// used for this: TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load
// from another GPU's memory hung the GPU, and these rows can live on a peer.
// The gathering CTA also writes the landed stage to the row-major operand copy that wgrad-1 reads (TMA
// store from the swizzled stage, off the critical path).
//
// Warp roles (320 threads): 0 A producer (all lanes issue gather4) | 1 TMEM alloc + MMA issuer |
// 2 B producer (W1 tiles) | 3 forwarder (DSMEM copy + operand-copy store) | 4..7 epilogue | 8..9 wide gather.
// Barriers per CTA: a_full[s] (1 arrival + 16 KB tx, bytes from the own gather4s or the sibling's copy),
// a_empty[s] (4 arrivals: both CTAs' MMA commits, multicast to both, and both forwarders), b_full / b_empty
// per 32 KB W1 unit (256 rows x 64 columns), tfull / tempty around the accumulator.
#pragma once

#include "sm100_ptx.cuh"

namespace dtf {
namespace g4 {

constexpr int kAStages = 4;            // 16 KB each
constexpr int kBUnits = 4;             // 32 KB each: 256 W1 rows x 64 columns
constexpr int kUnitN = 256;
constexpr int kThreads = 320;
constexpr int kABytes = 128 * 64 * 2;
constexpr int kBBytes = kUnitN * 64 * 2;
constexpr int kStagingPerWarp = 4096;  // 32 rows x 128 B
constexpr int kSmemBytes = kAStages * kABytes + kBUnits * kBBytes + 4 * kStagingPerWarp + 512 /*barriers*/ +
                           512 * 4 /*bias*/ + 1024 /*alignment*/;

struct Params {
  int M, N, NH;                 // N = 2 * NH
  int slots;                    // F; K0 = F * 128 + 64
  int m_tiles;
  const int* idx;               // [slots][idx_ld] window row index of (slot, example); entries past M must be valid rows (0)
  long long idx_ld;
  const float* bias;            // [N]
  uint32_t* mask_out;           // optional ReLU bit mask, bit j of word [m, n/32]
  long long ldmask;
  int relu;
  const unsigned long long* wide_src;   // optional [M * slots] addresses of fp32 scalars (may be peer memory)
  float* wide_out;                      // [M * slots]
  unsigned int* error_flag;             // optional: set to 1 when a wait times out (debug builds of the protocol)
};

__device__ __forceinline__ void tma_gather4(uint32_t dst, const void* tm, uint32_t bar, int c0, int r0, int r1, int r2, int r3) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
      : "memory");
}
// shared memory of this CTA -> shared memory of another CTA of the cluster; bytes are counted on a barrier of the destination
__device__ __forceinline__ void dsmem_copy(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_cluster),
               "r"(src_cta), "r"(bytes), "r"(bar_cluster)
               : "memory");
}
// arrive, once the MMAs issued so far have retired, on the barrier at this offset in BOTH CTAs of the cluster
__device__ __forceinline__ void umma_commit_both(uint32_t bar) {
  asm volatile(
      "{\n"
      " .reg .b16 m;\n"
      " mov.b16 m, 3;\n"
      " tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n"
      "}\n" ::"r"(bar)
      : "memory");
}
// Bounded wait: a protocol error must end the kernel (trap), never hang the GPU.  ~2 s.
__device__ __forceinline__ void mbar_wait_b(uint32_t bar, uint32_t parity) {
  unsigned long long t0 = 0;
  for (uint32_t spin = 0; !mbar_try_wait(bar, parity); ++spin) {
    if ((spin & 4095u) == 4095u) {
      const unsigned long long t = globaltimer_ns();
      if (t0 == 0) t0 = t;
      else if (t - t0 > 2000000000ull) {
        printf("pull_g4: block %d thread %d waited 2 s on barrier 0x%x parity %u\n", blockIdx.x, threadIdx.x, bar, parity);
        __trap();
      }
    }
  }
}

}  // namespace g4
}  // namespace dtf
