// Last stage of the loss reduction, done inside the launch that wrote the
// partials (K2, K5, K1a/K1b): the CTA that finishes last sums all partials in
// a fixed order, forms the weighted total (experiment.py:401-407) and, for a
// sharded batch, does the peer exchange.  One launch per call instead of two.
#pragma once
#include "exchange.cuh"
#include "fused_common.cuh"

namespace vtr {

struct FinishArgs {
  // Counter of finished CTAs: zero when the launch starts, set back to zero
  // by the CTA that finishes last.  Null: the stage is off and the host
  // appends K3 (reduce_losses_kernel).
  unsigned int* ticket;
  int num_ctas;      // CTAs that call finish_losses()
  int num_partials;  // double[3] partials the launch writes
  float* losses;
  double* losses_f64;
  // Loss exchange of a sharded batch (exchange.cuh); ex_world == 0 = off.
  PeerBuffers ex_peers;
  int ex_rank, ex_world;
  uint32_t ex_seq;
  uint32_t ex_lag;  // ex_sums receives the sums of step ex_seq - ex_lag
  float* ex_sums;
};

// Launch arguments of the 32-column-tile kernels K1a / K1b.
struct TileArgs {
  FusedArgs f;
  FinishArgs fin;  // ticket == null: the host appends K3
};

// The end of both variants below: one warp holds the three sums of the whole
// batch (every lane the same values) and forms the total, does the peer
// exchange and writes the results.
// `lagged_done`: the sums of step ex_seq - ex_lag (ex_lag > 0) were already
// fetched earlier in this launch (collect_lagged_early).
__device__ __forceinline__ void finish_publish(const FusedArgs& a, const FinishArgs& fin, double pg, double bl,
                                               double ent, bool lagged_done = false) {
  const int lane = threadIdx.x & 31;
  const double total =
      pg + static_cast<double>(a.baseline_cost) * bl + static_cast<double>(a.entropy_cost) * ent;
  if (lane < fin.ex_world)  // lane r stores this rank's four scalars into peer r
    exchange_publish4_to_peer(fin.ex_peers.p[lane], static_cast<float>(pg), static_cast<float>(bl),
                              static_cast<float>(ent), static_cast<float>(total), fin.ex_rank,
                              fin.ex_world, fin.ex_seq);
  if (fin.ex_world > 0 && fin.ex_sums && fin.ex_seq > fin.ex_lag && !(lagged_done && fin.ex_lag > 0)) {
    __syncwarp();
    exchange_collect_warp(fin.ex_peers.p[fin.ex_rank], 4, fin.ex_world, fin.ex_seq - fin.ex_lag,
                          fin.ex_sums, kExchangeTimeoutNs);
  }
  if (lane == 0) {
    if (fin.losses) {
      fin.losses[0] = static_cast<float>(pg), fin.losses[1] = static_cast<float>(bl);
      fin.losses[2] = static_cast<float>(ent), fin.losses[3] = static_cast<float>(total);
    }
    if (fin.losses_f64)
      fin.losses_f64[0] = pg, fin.losses_f64[1] = bl, fin.losses_f64[2] = ent, fin.losses_f64[3] = total;
    if (a.total_out) *a.total_out = static_cast<float>(total);
  }
}

// Sharded batch with exchange_lag > 0: the sums wanted are those of an earlier
// step and do not depend on this launch, so one warp with time to spare can
// fetch them while the launch streams (normally they are already in this
// rank's buffer) instead of the last CTA doing it behind everything else.
// Called by one whole warp of ONE CTA.
__device__ __forceinline__ void collect_lagged_early(const FinishArgs& fin) {
  if (fin.ticket && fin.ex_world > 0 && fin.ex_sums && fin.ex_lag > 0 && fin.ex_seq > fin.ex_lag)
    exchange_collect_warp(fin.ex_peers.p[fin.ex_rank], 4, fin.ex_world, fin.ex_seq - fin.ex_lag, fin.ex_sums,
                          kExchangeTimeoutNs);
}

// Called by ONE whole warp of every CTA, after its lane 0 has written the
// CTA's partial(s).  The summation order depends on the launch shape only:
// partials strided over the 32 lanes, then a fixed tree.
__device__ __forceinline__ void finish_losses(const FusedArgs& a, const FinishArgs& fin) {
  const int lane = threadIdx.x & 31;
  int last = 0;
  if (lane == 0) {
    __threadfence();  // this CTA's partial is visible before its ticket is
    const unsigned int before = atomicAdd(fin.ticket, 1u);
    last = before == static_cast<unsigned int>(fin.num_ctas - 1);
    if (last) *fin.ticket = 0u;  // everybody has drawn: ready for the next launch
  }
  last = __shfl_sync(0xffffffffu, last, 0);
  if (!last) return;
  __threadfence();
  double pg = 0.0, bl = 0.0, ent = 0.0;
  // This is the tail of the launch: four partials per lane in flight at a time
  // (an L2 round trip each).  ld.global.cg: other CTAs wrote them; this SM's
  // L1 has never seen them.  Added in index order.
  for (int i0 = lane; i0 < fin.num_partials; i0 += 128) {
    double v[4][3];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + 32 * u;
      const bool in = i < fin.num_partials;
#pragma unroll
      for (int k = 0; k < 3; ++k) v[u][k] = in ? __ldcg(a.partials + 3 * i + k) : 0.0;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) pg += v[u][0], bl += v[u][1], ent += v[u][2];
  }
  pg = warp_sum_fixed(pg), bl = warp_sum_fixed(bl), ent = warp_sum_fixed(ent);
  finish_publish(a, fin, pg, bl, ent);
}

// The same stage for kernels whose CTAs have several warps: called by EVERY
// thread of every CTA (blockDim.x a multiple of 32), after thread 0 has
// written the CTA's partial(s).  The CTA that finishes last reads the partials
// with all of its threads - one L2 round trip instead of one per 128 partials
// - and sums them in an order that depends on the launch shape only: partials
// strided over the threads, a fixed tree inside each warp, the warps in a
// fixed tree.  `scratch`: 3 doubles of shared memory per warp, free by now.
__device__ __forceinline__ void finish_losses_cta(const FusedArgs& a, const FinishArgs& fin, double* scratch,
                                                  bool lagged_done = false) {
  __shared__ int sh_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int threads = blockDim.x;
  if (threadIdx.x == 0) {
    __threadfence();  // this CTA's partial is visible before its ticket is
    const unsigned int before = atomicAdd(fin.ticket, 1u);
    const int last = before == static_cast<unsigned int>(fin.num_ctas - 1);
    if (last) *fin.ticket = 0u;  // everybody has drawn: ready for the next launch
    sh_last = last;
  }
  __syncthreads();
  if (!sh_last) return;
  __threadfence();
  double pg = 0.0, bl = 0.0, ent = 0.0;
  for (int i0 = threadIdx.x; i0 < fin.num_partials; i0 += 2 * threads) {
    double v[2][3];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int i = i0 + threads * u;
      const bool in = i < fin.num_partials;
#pragma unroll
      for (int k = 0; k < 3; ++k) v[u][k] = in ? __ldcg(a.partials + 3 * i + k) : 0.0;
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) pg += v[u][0], bl += v[u][1], ent += v[u][2];
  }
  pg = warp_sum_fixed(pg), bl = warp_sum_fixed(bl), ent = warp_sum_fixed(ent);
  if (lane == 0) scratch[3 * warp] = pg, scratch[3 * warp + 1] = bl, scratch[3 * warp + 2] = ent;
  __syncthreads();
  if (warp != 0) return;
  const bool has = lane < (threads >> 5);
  pg = has ? scratch[3 * lane] : 0.0, bl = has ? scratch[3 * lane + 1] : 0.0;
  ent = has ? scratch[3 * lane + 2] : 0.0;
  pg = warp_sum_fixed(pg), bl = warp_sum_fixed(bl), ent = warp_sum_fixed(ent);
  finish_publish(a, fin, pg, bl, ent, lagged_done);
}

}  // namespace vtr
