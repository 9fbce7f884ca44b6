// Pieces shared by the fused-pass kernels: launch arguments, row load/store
// helpers and the routine that fills one ring slot.  See vtrace_fused.cu for
// the overview of the kernels.
#pragma once
#include <math.h>
#include <stdint.h>

#include "common.h"
#include "ptx.cuh"
#include "vtrace_math.cuh"

namespace vtr {

struct FusedArgs {
  int T, B, A;
  int depth;        // ring slots
  int slot_floats;  // floats per ring slot: 2*32*A logits + 4*32 per-step operands
  uint32_t flags;
  int reward_clip_mode;
  int tma_in_ok;   // logits rows are 16-byte aligned for every t and tile
  int tma_out_ok;  // same for d_target_logits
  float clip_rho, clip_pg_rho, baseline_cost, entropy_cost, discounting;
  const float* blog;
  const float* tlog;
  const int32_t* actions;
  const float* discounts;
  const float* rewards;
  const float* values;
  const float* bootstrap;
  const uint8_t* done;
  float* vs;
  float* pg;
  float* log_rhos;
  float* b_lp;
  float* t_lp;
  float* d_logits;
  float* d_values;
  double* partials;  // loss partials, or null
  float* losses;       // K1c only: final losses written by the last tile to finish
  double* losses_f64;  // K1c only
  float* total_out;    // the weighted total once more, or null (vtr_fused_params.total_loss)
};

constexpr int kColumnsPerTile = 32;
constexpr int kColumnBarrierBytes = 128;  // K1a: up to 16 "slot full" mbarriers
constexpr int kRelayBarrierBytes = 512;   // K1b: 32 "slot full" + 32 "carry arrived"
constexpr int kMaxDepth = 16;             // K1a, K1c
constexpr int kMaxWarps = 16;
constexpr int kRelayMaxDepth = 32;
constexpr int kRelayMaxWarps = 16;        // K1b: a tile is issue-bound from ~16 warps on (20 and 24 measured the same)

template <int N>
__device__ __forceinline__ void load_row(float (&x)[N], const float* row) {
  if constexpr (N % 4 == 0) {
#pragma unroll
    for (int i = 0; i < N / 4; ++i) {
      float4 v = reinterpret_cast<const float4*>(row)[i];
      x[4 * i] = v.x, x[4 * i + 1] = v.y, x[4 * i + 2] = v.z, x[4 * i + 3] = v.w;
    }
  } else if constexpr (N % 2 == 0) {
#pragma unroll
    for (int i = 0; i < N / 2; ++i) {
      float2 v = reinterpret_cast<const float2*>(row)[i];
      x[2 * i] = v.x, x[2 * i + 1] = v.y;
    }
  } else {
#pragma unroll
    for (int i = 0; i < N; ++i) x[i] = row[i];
  }
}

template <int N>
__device__ __forceinline__ void store_row(float* row, const float (&x)[N]) {
  if constexpr (N % 4 == 0) {
#pragma unroll
    for (int i = 0; i < N / 4; ++i)
      reinterpret_cast<float4*>(row)[i] = make_float4(x[4 * i], x[4 * i + 1], x[4 * i + 2], x[4 * i + 3]);
  } else if constexpr (N % 2 == 0) {
#pragma unroll
    for (int i = 0; i < N / 2; ++i) reinterpret_cast<float2*>(row)[i] = make_float2(x[2 * i], x[2 * i + 1]);
  } else {
#pragma unroll
    for (int i = 0; i < N; ++i) row[i] = x[i];
  }
}

__device__ __forceinline__ double warp_sum_fixed(double v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

// Fills ring slot `s` (shared address `dst`, mbarrier `bar`) with time step `t`
// of the tile starting at column b0.  Called by all 32 lanes of one warp; the
// barrier expects 33 arrivals: every lane's cp.async group + lane 0's arm.
__device__ __forceinline__ void issue_step(const FusedArgs& a, int t, uint32_t dst, uint32_t bar,
                                           int lane, int b0, int nb, int A, bool tma_in,
                                           uint64_t policy) {
  const uint32_t row_bytes = static_cast<uint32_t>(nb) * A * 4u;
  const size_t row0 = (static_cast<size_t>(t) * a.B + b0) * A;
  if (tma_in) {
    if (lane == 0) {
      ptx::mbar_arrive_expect_tx(bar, 2u * row_bytes);
      ptx::bulk_g2s(dst, a.blog + row0, row_bytes, bar, policy);
      ptx::bulk_g2s(dst + 32u * A * 4u, a.tlog + row0, row_bytes, bar, policy);
    }
  } else {
    if (lane == 0) ptx::mbar_arrive(bar);
    for (int i = lane; i < nb * A; i += 32) {
      ptx::cp_async4(dst + 4u * i, a.blog + row0 + i);
      ptx::cp_async4(dst + (32u * A + i) * 4u, a.tlog + row0 + i);
    }
  }
  if (lane < nb) {
    const size_t e = static_cast<size_t>(t) * a.B + b0 + lane;
    const uint32_t sc = dst + (64u * A + lane) * 4u;
    ptx::cp_async4(sc, a.actions + e);
    if (a.discounts) ptx::cp_async4(sc + 128u, a.discounts + e);
    ptx::cp_async4(sc + 256u, a.rewards + e);
    ptx::cp_async4(sc + 384u, a.values + e);
  }
  ptx::cp_async_mbar_arrive_noinc(bar);
}


}  // namespace vtr
