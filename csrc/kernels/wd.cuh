// Wide-and-deep training plan: shared definitions between the host executor (wd_plan.cu) and the
// device kernels (wd_kernels.cu).
//
// Every rank is at the same time a worker (runs the model on its own micro-batch) and a
// parameter-server shard (owns embedding rows r with r % world == rank and a contiguous range of
// the flat dense parameter vector) -- the B200 mapping of the reference's ps / worker / master
// roles (reference trainer/task.py:121-127).  Shards live in a symmetric heap that every peer maps.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "wd_optim.cuh"

namespace dtf {

constexpr int kMaxWorld = 16;
constexpr int kEmbDim = 128;  // one embedding row = 128 bf16 = 256 B = two 128-byte swizzle rows


struct PeerTable {
  uint8_t* slab[kMaxWorld];
  int world;
  int rank;
};

// byte offsets of the symmetric buffers inside each rank's slab (identical on every rank)
struct SlabLayout {
  size_t emb_bf16;     // [rows_local_max, 128] bf16   working copy of the embedding shard (pulled by peers)
  size_t wide;         // [rows_local_max] f32         wide weights shard (also the master copy)
  size_t wbf16;        // [P_pad] bf16                 full dense parameters, written by their owners
  size_t small_f32;    // [n_small] f32                biases / last layer in fp32, written by owners
  size_t dense_grad;   // [P_pad] f32                  this rank's dense gradient (read by owners)
  size_t mbx_payload;  // [world, cap, 128] bf16       embedding-gradient rows pushed to this shard
  size_t mbx_rows;     // [world, cap] i64             per pushed entry: local row index (low 32 bits) | example index at the source (high 32)
  size_t mbx_dlvec;    // [world, B] f32               dlogit of every example of each source = wide gradient of its entries
  size_t mbx_count;    // [world] u32                  entries published by each source this step
  size_t signal;       // [world * 16] u32             barrier flags
  size_t total;
};

struct DenseLayout {
  // flat parameter space; big matrices first, then the "small" tail kept in fp32 for epilogues
  long long w1, w2, w3;             // [h1,K0] [h2,h1] [h3,h2]
  long long small_begin;            // == b1
  long long b1, b2, b3, w4, b4, wide_b;
  long long P;                      // total (multiple of 256)
  int n_small;
};

struct WdShapes {
  int B, F, dense_in, dense_pad, K0, h1, h2, h3;
  long long V;            // global embedding rows
  long long rows_local;   // rows owned by this rank
  long long rows_local_max;
  long long cap;          // mailbox capacity per source = B * F
};

}  // namespace dtf
