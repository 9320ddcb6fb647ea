// Planned-op objects behind the C ABI: an op is created once for fixed device
// pointers/shapes (TMA descriptors encoded at creation) and then enqueued any
// number of times on a stream; a sequence replays a list of ops with one call.
#pragma once
#include <vector>

#include "common.cuh"

struct b2seg_op {
  virtual ~b2seg_op() {}
  virtual int run(cudaStream_t st) = 0;
  virtual const char* name() const = 0;
  // number of kernel launches one run() enqueues (for bench accounting)
  virtual int launches() const { return 1; }
};

struct b2seg_seq {
  std::vector<b2seg_op*> ops;
  // lane 0 = the dependent chain; lane 1 = ops whose results nothing in the sequence reads
  // (weight gradients): in the captured graph they fork onto a side branch after the op that
  // precedes them and join at the end, so they fill the chain's latency bubbles.
  std::vector<int> lanes;
  std::vector<size_t> join_at;  // op indices before which the side branch joins the chain
  cudaStream_t side_stream = nullptr;
  cudaEvent_t fork_event = nullptr, join_event = nullptr;
  // CUDA-graph replay: the op list is captured once on a private stream (PyTorch's default stream
  // is the legacy stream, which cannot be captured) and launched into the caller's stream.
  cudaGraphExec_t exec = nullptr;
  cudaGraph_t graph = nullptr;
  cudaStream_t capture_stream = nullptr;
  bool capture_failed = false;
};

// NHWC bf16 view -> rank-4 TMA descriptor with box (64ch, wt, ht, nt), 128B swizzle.
int b2_view_tmap(CUtensorMap* out, const b2seg_tensor4* v, int wt, int ht, int nt);

// pixel-tile shape (wt*ht*nt == 128) for an (n,h,w) iteration grid
static inline void b2_pick_pixel_tile(int n, int h, int w, int* wt, int* ht, int* nt) {
  auto p2 = [](int x) { int p = 1; while (p < x) p <<= 1; return p; };
  int a = p2(w); if (a > 8) a = 8;  // 8-wide bricks: one brick row = one 8-row UMMA group (conv3x3.cu)
  int b = p2(h); if (b > 128 / a) b = 128 / a;
  *wt = a; *ht = b; *nt = 128 / (a * b);
  (void)n;
}

// conv3x3.cu: halo-reuse kernel for 3x3 stride-1 tap tables; *out stays NULL when not applicable.
int b2_halo3x3_try_create(const b2seg_tapgemm_desc* d, b2seg_op** out);
