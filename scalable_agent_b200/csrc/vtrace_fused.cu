// Fused V-trace + IMPALA loss, forward and backward, one pass over HBM.
//
// Replaces vtrace.from_logits (vtrace.py:71-161), the three loss helpers
// (experiment.py:324-343), the weighted total (experiment.py:401-407) and the
// autodiff backward of that total (experiment.py:415).
//
// K1 (this file).  A tile is 32 adjacent batch columns, one lane per column,
// walked from t = T-1 down to 0 so that the reverse recurrence
// acc_t = delta_t + gamma_t c_t acc_{t+1} (vtrace.py:250-262) and vs_{t+1}
// (vtrace.py:268-269) are carried, not recomputed.  For one time step the 32
// columns' logits are ONE contiguous run of 32*A floats in the time-major
// [T,B,A] layout; a ring of `depth` slots in shared memory is kept full by TMA
// bulk copies (cp.async.bulk -> mbarrier complete_tx) for both logits tensors
// plus 4-byte cp.async for the four [T,B] operands, all landing on the slot's
// one mbarrier.  d_target_logits rows are staged in shared memory and leave
// with TMA bulk stores.  Two shells run the same arithmetic (vtrace_math.cuh):
//   K1a fused_column_kernel  one warp per tile, deepest ring   (>= 7 tiles/SM)
//   K1b fused_relay_kernel   W warps per tile relaying the carry (< 7 tiles/SM)
// Losses accumulate per lane in double, are reduced with a fixed shuffle tree
// and written as one partial per tile; K3 (reduce_losses_kernel) sums the
// partials in a fixed order.  No atomics.
#include <math.h>
#include <stdlib.h>
#include <string.h>

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
};

constexpr int kColumnsPerTile = 32;
constexpr int kColumnBarrierBytes = 128;  // K1a: up to 16 "slot full" mbarriers
constexpr int kRelayBarrierBytes = 256;   // K1b: 16 "slot full" + 16 "carry arrived"
constexpr int kMaxDepth = 16;
constexpr int kMaxWarps = 16;

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

// ---------------------------------------------------------------------------
// K1a: one warp per tile.  Used when an SM has to hold so many tiles (>= 7)
// that shared memory allows only one warp's worth of ring per tile: the whole
// ring (5 slots at A=18) then runs ahead of the single consumer.
// ---------------------------------------------------------------------------
template <int A_CT, bool GRADS>
__global__ void __launch_bounds__(32) fused_column_kernel(const __grid_constant__ FusedArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int A = A_CT > 0 ? A_CT : a.A;
  const int T = a.T, B = a.B;
  const int lane = threadIdx.x;
  const int b0 = blockIdx.x * kColumnsPerTile;
  const int nb = min(kColumnsPerTile, B - b0);
  const bool active = lane < nb;
  const int b = b0 + (active ? lane : 0);
  const int depth = a.depth;
  const int slot_floats = a.slot_floats;

  float* ring = reinterpret_cast<float*>(smem_raw + kColumnBarrierBytes);
  float* outbuf = ring + static_cast<size_t>(depth) * slot_floats;  // 2 x [32*A], GRADS only
  const uint32_t bars = ptx::smem_u32(smem_raw);
  const uint32_t ring_u32 = ptx::smem_u32(ring);
  const uint32_t row_bytes = static_cast<uint32_t>(nb) * A * 4u;
  const int row_elems = nb * A;
  const bool tma_in = a.tma_in_ok && (row_bytes % 16u == 0);
  const bool tma_out = a.tma_out_ok && (row_bytes % 16u == 0);
  const bool need_softmax = GRADS || a.partials != nullptr;

  if (lane == 0) {
    for (int s = 0; s < depth; ++s) ptx::mbar_init(bars + 8u * s, 33);
    ptx::fence_mbar_init();
    ptx::fence_proxy_async();
  }
  __syncwarp();
  const uint64_t policy = ptx::policy_evict_first();

  auto issue = [&](int t, int s) {
    issue_step(a, t, ring_u32 + static_cast<uint32_t>(s) * slot_floats * 4u, bars + 8u * s, lane, b0,
               nb, A, tma_in, policy);
  };
  const int prologue = min(depth, T);
  for (int k = 0; k < prologue; ++k) issue(T - 1 - k, k);

  math::Carry carry;
  carry.acc = 0.0f;
  carry.v_next = active ? a.bootstrap[b] : 0.0f;
  carry.vs_next = carry.v_next;
  math::LossSums sums = {0.0, 0.0, 0.0};
  uint8_t done_next = 0;
  if (a.done && active) done_next = a.done[static_cast<size_t>(T - 1) * B + b];

  int s = 0;
  uint32_t parity = 0;
  for (int it = 0; it < T; ++it) {
    const int t = T - 1 - it;
    const size_t e = static_cast<size_t>(t) * B + b;
    const uint8_t done_cur = done_next;
    if (a.done && active && t > 0) done_next = a.done[e - B];

    if (GRADS) {
      // The staging buffer written this step was handed to a bulk store two
      // steps ago; wait until that store has finished reading it.
      if (tma_out && lane == 0) ptx::bulk_wait_read<1>();
      __syncwarp();
    }
    ptx::mbar_wait(bars + 8u * s, parity);

    const float* slot = ring + static_cast<size_t>(s) * slot_floats;
    const float* xb_row = slot + lane * A;
    const float* xt_row = slot + 32 * A + lane * A;
    const float* sc = slot + 64 * A;
    float* orow = outbuf + (it & 1) * 32 * A + lane * A;

    // ---- operands and rows out of the slot ----
    int act = 0;
    float disc = 0.0f, rew = 0.0f, val = 0.0f, xa_t = 0.0f, xa_b = 0.0f;
    math::RowStats st = {0.0f, 1.0f, 0.0f}, sb = {0.0f, 1.0f, 0.0f};
    float x[A_CT > 0 ? A_CT : 1], ex[A_CT > 0 ? A_CT : 1];  // target row, its exp terms
    if (active) {
      act = __float_as_int(sc[lane]);
      act = min(max(act, 0), A - 1);
      disc = a.done ? (done_cur ? 0.0f : a.discounting) : sc[32 + lane];
      rew = sc[64 + lane];
      val = sc[96 + lane];
      xa_t = xt_row[act];
      xa_b = xb_row[act];
      if constexpr (A_CT > 0) {
        float xb[A_CT];
        load_row(xb, xb_row);
        sb = math::row_stats(xb);
        load_row(x, xt_row);
      }
    }
    if constexpr (A_CT > 0) {
      // The rows are in registers: the slot can take step it+depth already.
      __syncwarp();
      if (it + depth < T) issue(t - depth, s);
    }

    if (active) {
      rew = math::clip_reward(rew, a.reward_clip_mode);
      // ---- action log-probs of both policies (vtrace.py:143-146) ----
      if constexpr (A_CT > 0) {
        st = math::row_stats(x, ex);
      } else {
        st = math::smem_row_stats(xt_row, A);
        sb = math::smem_row_stats(xb_row, A);
      }
      const float t_lp = math::action_log_prob(xa_t, st);
      const float b_lp = math::action_log_prob(xa_b, sb);
      const float log_rho = __fsub_rn(t_lp, b_lp);  // vtrace.py:147

      // ---- V-trace step (vtrace.py:235-276), walking t downwards ----
      const math::Weights w = math::importance_weights(log_rho, a.flags, a.clip_rho, a.clip_pg_rho);
      float vs, pg;
      math::vtrace_step(w, rew, disc, val, carry, vs, pg);

      if (a.vs) a.vs[e] = vs;
      if (a.pg) a.pg[e] = pg;
      if (a.log_rhos) a.log_rhos[e] = log_rho;
      if (a.b_lp) a.b_lp[e] = b_lp;
      if (a.t_lp) a.t_lp[e] = t_lp;

      // ---- losses (experiment.py:324-343) and their gradients ----
      if (need_softmax) {
        const float inv_sum = __frcp_rn(st.sum);
        float entropy;
        if constexpr (A_CT > 0) entropy = math::row_entropy(x, ex, st, inv_sum);
        else entropy = math::smem_row_entropy(xt_row, A, st, inv_sum);
        if (GRADS) {
          // d/dlogits = (p - onehot) pg_adv + entropy_cost p (log p + H)
          const float ka = __fmul_rn(a.entropy_cost, inv_sum);
          const float kb = __fmul_rn(__fmaf_rn(a.entropy_cost, entropy, pg), inv_sum);
          if constexpr (A_CT > 0) {
            math::row_gradient(ex, x, ka, kb);
            store_row(orow, ex);
          } else {
            math::smem_row_gradient(orow, xt_row, A, st, ka, kb);
          }
          orow[act] = __fsub_rn(orow[act], pg);  // the -onehot(a) * pg_adv term
          if (a.d_values) a.d_values[e] = __fmul_rn(a.baseline_cost, __fsub_rn(val, vs));
        }
        if (a.partials) sums.add(t_lp, pg, vs, val, entropy);
      }
    }

    if (GRADS) {
      float* gdst = a.d_logits + (static_cast<size_t>(t) * B + b0) * A;
      if (tma_out) {
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          ptx::bulk_s2g(gdst, ptx::smem_u32(outbuf + (it & 1) * 32 * A), row_bytes);
          ptx::bulk_commit();
        }
      } else {
        __syncwarp();
        const float* src = outbuf + (it & 1) * 32 * A;
        for (int i = lane; i < row_elems; i += 32) gdst[i] = src[i];
      }
    }
    if constexpr (A_CT == 0) {
      // run-time A re-reads the rows from the slot: refill only now
      __syncwarp();
      if (it + depth < T) issue(t - depth, s);
    }
    if (++s == depth) s = 0, parity ^= 1u;
  }

  if (a.partials) {
    const double pg_sum = warp_sum_fixed(sums.pg);
    const double bl_sum = warp_sum_fixed(sums.baseline);
    const double ent_sum = warp_sum_fixed(sums.entropy);
    if (lane == 0) {
      double* out = a.partials + 3 * static_cast<size_t>(blockIdx.x);
      out[0] = pg_sum, out[1] = bl_sum, out[2] = ent_sum;
    }
  }
  if (GRADS && tma_out && lane == 0) ptx::bulk_wait<0>();
}

// ---------------------------------------------------------------------------
// K1b: several warps per tile, passing the recurrence along ("relay").  Warp w
// owns the time steps it = w, w+W, ... (it counts from the end: t = T-1-it);
// each does its whole step in its own registers and only the 3-float carry
// (acc, V_{t+1}, vs_{t+1}) per lane travels from the owner of step t+1 to the
// owner of step t, through shared memory and an mbarrier.  Everything that
// does not depend on the carry (both softmaxes, log_rho, rho, the entropy:
// most of the work) overlaps between warps, so a tile advances much faster
// than one warp can; used whenever shared memory leaves room for it.
//
// Shared memory of one tile, in bytes from the start:
//   [0,128)    `depth` mbarriers "slot full"      (TMA + cp.async completion)
//   [128,256)  `warps` mbarriers "carry arrived"  (recurrence hand-off)
//   ring       depth slots of slot_floats floats: one time step of the tile each
//   staging    warps x 32*A floats: each warp's d_target_logits row  (GRADS only)
//   carry      warps x 3 x 32 floats: (acc, V_{t+1}, vs_{t+1}) per lane
//   sums       warps x 3 doubles: per-warp loss partials
//
// Launch bound 768 threads (never more than 512 are launched) caps the kernel
// at 80 registers, so that 24 warps fit an SM's register file.
// ---------------------------------------------------------------------------
template <int A_CT, bool GRADS>
__global__ void __launch_bounds__(768) fused_relay_kernel(const __grid_constant__ FusedArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int A = A_CT > 0 ? A_CT : a.A;
  const int T = a.T, B = a.B;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int W = blockDim.x >> 5;
  const int b0 = blockIdx.x * kColumnsPerTile;
  const int nb = min(kColumnsPerTile, B - b0);
  const bool active = lane < nb;
  const int b = b0 + (active ? lane : 0);
  const int depth = a.depth;
  const int slot_floats = a.slot_floats;

  float* ring = reinterpret_cast<float*>(smem_raw + kRelayBarrierBytes);
  float* staging = ring + static_cast<size_t>(depth) * slot_floats;
  float* carry_buf = staging + (GRADS ? W * 32 * A : 0);
  double* warp_sums = reinterpret_cast<double*>(carry_buf + W * 96);
  const uint32_t full_bars = ptx::smem_u32(smem_raw);
  const uint32_t carry_bars = full_bars + 128u;
  const uint32_t ring_u32 = ptx::smem_u32(ring);
  const uint32_t row_bytes = static_cast<uint32_t>(nb) * A * 4u;
  const int row_elems = nb * A;
  const bool tma_in = a.tma_in_ok && (row_bytes % 16u == 0);
  const bool tma_out = a.tma_out_ok && (row_bytes % 16u == 0);
  const bool need_softmax = GRADS || a.partials != nullptr;
  float* orow_base = staging + warp * 32 * A;  // this warp's gradient row
  float* orow = orow_base + lane * A;

  if (threadIdx.x == 0) {
    for (int s = 0; s < depth; ++s) ptx::mbar_init(full_bars + 8u * s, 33);
    for (int w = 0; w < W; ++w) ptx::mbar_init(carry_bars + 8u * w, 32);
    ptx::fence_mbar_init();
    ptx::fence_proxy_async();
  }
  __syncthreads();
  const uint64_t policy = ptx::policy_evict_first();

  auto issue = [&](int t, int s) {
    issue_step(a, t, ring_u32 + static_cast<uint32_t>(s) * slot_floats * 4u, full_bars + 8u * s, lane,
               b0, nb, A, tma_in, policy);
  };
  // Step `it` lives in ring slot it % depth and is loaded by the warp that owns it.
  for (int it = warp; it < min(depth, T); it += W) issue(T - 1 - it, it);

  math::LossSums sums = {0.0, 0.0, 0.0};

  for (int it = warp; it < T; it += W) {
    const int t = T - 1 - it;
    const size_t e = static_cast<size_t>(t) * B + b;
    const int s = it % depth;
    const uint8_t done_cur = (a.done && active) ? a.done[e] : 0;

    ptx::mbar_wait(full_bars + 8u * s, (it / depth) & 1);
    const float* slot = ring + static_cast<size_t>(s) * slot_floats;
    const float* xb_row = slot + lane * A;
    const float* xt_row = slot + 32 * A + lane * A;
    const float* sc = slot + 64 * A;

    // ---- phase A: everything that does not depend on step t+1 ----
    int act = 0;
    float disc = 0.0f, rew = 0.0f, val = 0.0f, xa_t = 0.0f, xa_b = 0.0f;
    math::RowStats st = {0.0f, 1.0f, 0.0f}, sb = {0.0f, 1.0f, 0.0f};
    float x[A_CT > 0 ? A_CT : 1], ex[A_CT > 0 ? A_CT : 1];  // target row, its exp terms
    if (active) {
      act = __float_as_int(sc[lane]);
      act = min(max(act, 0), A - 1);
      disc = a.done ? (done_cur ? 0.0f : a.discounting) : sc[32 + lane];
      rew = sc[64 + lane];
      val = sc[96 + lane];
      xa_t = xt_row[act];
      xa_b = xb_row[act];
      if constexpr (A_CT > 0) {
        // behaviour row first: it reduces to its statistics and frees its registers
        float xb[A_CT];
        load_row(xb, xb_row);
        sb = math::row_stats(xb);
        load_row(x, xt_row);
      }
    }
    if constexpr (A_CT > 0) {
      // The rows are in registers: the slot can take step it+depth already.
      __syncwarp();
      if (it + depth < T) issue(t - depth, s);
    }
    float t_lp = 0.0f, b_lp = 0.0f, log_rho = 0.0f;
    float entropy = 0.0f, inv_sum = 0.0f;
    math::Weights w = {0.0f, 0.0f, 0.0f};
    if (active) {
      rew = math::clip_reward(rew, a.reward_clip_mode);
      // action log-probs of both policies (vtrace.py:143-146)
      if constexpr (A_CT > 0) {
        st = math::row_stats(x, ex);
      } else {
        st = math::smem_row_stats(xt_row, A);
        sb = math::smem_row_stats(xb_row, A);
      }
      t_lp = math::action_log_prob(xa_t, st);
      b_lp = math::action_log_prob(xa_b, sb);
      log_rho = __fsub_rn(t_lp, b_lp);  // vtrace.py:147
      w = math::importance_weights(log_rho, a.flags, a.clip_rho, a.clip_pg_rho);
      if (need_softmax) {
        inv_sum = __frcp_rn(st.sum);
        if constexpr (A_CT > 0) entropy = math::row_entropy(x, ex, st, inv_sum);
        else entropy = math::smem_row_entropy(xt_row, A, st, inv_sum);
      }
    }

    // ---- phase B: the V-trace step (vtrace.py:243-276); needs step t+1 ----
    math::Carry carry;
    if (it == 0) {
      carry.acc = 0.0f;
      carry.v_next = active ? a.bootstrap[b] : 0.0f;
      carry.vs_next = carry.v_next;
    } else {
      // n-th use of this warp's carry slot: steps warp, warp+W, ... (step 0 has none)
      const int use = it / W - (warp == 0 ? 1 : 0);
      ptx::mbar_wait(carry_bars + 8u * warp, use & 1);
      const float* cb = carry_buf + warp * 96;
      carry.acc = cb[lane], carry.v_next = cb[32 + lane], carry.vs_next = cb[64 + lane];
    }
    float vs, pg;
    math::vtrace_step(w, rew, disc, val, carry, vs, pg);
    if (it + 1 < T) {
      const int nw = warp + 1 == W ? 0 : warp + 1;  // owner of step it+1
      float* cb = carry_buf + nw * 96;
      cb[lane] = carry.acc, cb[32 + lane] = carry.v_next, cb[64 + lane] = carry.vs_next;
      ptx::mbar_arrive(carry_bars + 8u * nw);
    }

    // ---- phase C: outputs, gradient row, loss terms ----
    if (GRADS) {
      // this warp's previous bulk store must have read the staging row
      if (tma_out && lane == 0) ptx::bulk_wait_read<0>();
      __syncwarp();
    }
    if (active) {
      if (a.vs) a.vs[e] = vs;
      if (a.pg) a.pg[e] = pg;
      if (a.log_rhos) a.log_rhos[e] = log_rho;
      if (a.b_lp) a.b_lp[e] = b_lp;
      if (a.t_lp) a.t_lp[e] = t_lp;
      if (GRADS) {
        // d/dlogits = (p - onehot) pg_adv + entropy_cost p (log p + H)
        const float ka = __fmul_rn(a.entropy_cost, inv_sum);
        const float kb = __fmul_rn(__fmaf_rn(a.entropy_cost, entropy, pg), inv_sum);
        if constexpr (A_CT > 0) {
          math::row_gradient(ex, x, ka, kb);
          store_row(orow, ex);
        } else {
          math::smem_row_gradient(orow, xt_row, A, st, ka, kb);
        }
        orow[act] = __fsub_rn(orow[act], pg);  // the -onehot(a) * pg_adv term
        if (a.d_values) a.d_values[e] = __fmul_rn(a.baseline_cost, __fsub_rn(val, vs));
      }
      if (a.partials) sums.add(t_lp, pg, vs, val, entropy);
    }
    if (GRADS) {
      float* gdst = a.d_logits + (static_cast<size_t>(t) * B + b0) * A;
      if (tma_out) {
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          ptx::bulk_s2g(gdst, ptx::smem_u32(orow_base), row_bytes);
          ptx::bulk_commit();
        }
      } else {
        __syncwarp();
        for (int i = lane; i < row_elems; i += 32) gdst[i] = orow_base[i];
        __syncwarp();
      }
    }
    if constexpr (A_CT == 0) {
      // run-time A re-reads the rows from the slot: refill only now
      __syncwarp();
      if (it + depth < T) issue(t - depth, s);
    }
  }

  if (a.partials) {
    const double pg_sum = warp_sum_fixed(sums.pg);
    const double bl_sum = warp_sum_fixed(sums.baseline);
    const double ent_sum = warp_sum_fixed(sums.entropy);
    if (lane == 0)
      warp_sums[3 * warp] = pg_sum, warp_sums[3 * warp + 1] = bl_sum, warp_sums[3 * warp + 2] = ent_sum;
    __syncthreads();
    if (threadIdx.x == 0) {
      double p0 = 0.0, p1 = 0.0, p2 = 0.0;
      for (int w = 0; w < W; ++w)
        p0 += warp_sums[3 * w], p1 += warp_sums[3 * w + 1], p2 += warp_sums[3 * w + 2];
      double* out = a.partials + 3 * static_cast<size_t>(blockIdx.x);
      out[0] = p0, out[1] = p1, out[2] = p2;
    }
  }
  if (GRADS && tma_out && lane == 0) ptx::bulk_wait<0>();
}

// ---------------------------------------------------------------------------
// K1c: one persistent tile per SM ("SM tile").  The batch is split into
// G (a multiple of the SM count) contiguous column ranges of equal width, up
// to 256 columns each, so every SM streams the same number of bytes.  All
// column groups of a tile share ONE ring: a producer warp fills slot after slot
// with six TMA bulk copies per time step (two logits rows of ncols*A floats,
// four operand rows of ncols floats: runs of up to 18 KB, which keeps DRAM
// pages open), consumer warps wait on the slot's "full" mbarrier, pull their
// rows into registers and hand the slot back through its "empty" mbarrier
// straight away, long before their arithmetic is done.  Consumer warp
// (g, r) owns columns [32g, 32g+32) of the tile and the time steps
// it = r, r+R, ...; the R warps of a column group relay the recurrence carry
// as in K1b (R = 1: the carry stays in registers).  Needs 16-byte aligned
// rows: B % 4 == 0 and 16-byte aligned base pointers; anything else goes to
// K1a/K1b.
//
// Shared memory of one tile, in bytes from the start:
//   [0,128)    depth mbarriers "slot full"   (count 1 + tx bytes)
//   [128,256)  depth mbarriers "slot empty"  (count = column groups)
//   [256,384)  groups*relay mbarriers "carry arrived" (count 32)
//   ring       depth slots: blog[cols*A] tlog[cols*A] actions[cols] discounts[cols]
//              rewards[cols] values[cols], cols = 32*groups
//   staging    per consumer warp, staging_bufs x 32*A floats   (GRADS only)
//   carry      per consumer warp 3 x 32 floats
//   sums       per consumer warp 3 doubles
// ---------------------------------------------------------------------------
constexpr int kSmBarrierBytes = 512;
constexpr int kSmMaxConsumerWarps = 18;
constexpr int kMaxSms = 256;

// Measured streaming speed of every SM (columns x steps per ns, smoothed over
// calls), written by K1c at the end of each tile and read at the start of the
// next call to size the tiles.  SMs do not get equal shares of the memory
// system (on B200 the spread between the fastest and the slowest SM under full
// load is ~40 %, fixed by where the SM sits), so equal tiles leave the fast SMs
// idle for the last quarter of the kernel.  Only the partition depends on
// this table; every output value is independent of the partition.
static __device__ float g_sm_speed[kMaxSms];

struct SmTileArgs {
  FusedArgs f;
  int groups;        // column groups (warps side by side) per tile = capacity / 32
  int relay;         // warps per column group, owning interleaved time steps
  int staging_bufs;  // 1 or 2 gradient staging rows per consumer warp
  int num_sm;        // SMs of the device
  int waves;         // tiles per SM (grid = num_sm * waves) when balancing
  unsigned int* cursor;  // zeroed column cursor for speed-weighted tiles, or null
};

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned int sm_id() {
  unsigned int s;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(s));
  return s;
}

// Tile sizing with the speed table (warp 0 of every tile computes the same
// numbers from the same table): SM s gets granules in proportion to its
// speed, clamped to [0.7, 1.25] x the even share, plus what is needed for the
// shares to cover the batch.  Returns this tile's share in granules; falls
// back to the even share when a share would exceed `cap` granules.
__device__ __forceinline__ int weighted_share(const SmTileArgs& args, int total_gran, int cap,
                                              int lane, unsigned int my_sm) {
  const int nsm = args.num_sm, grid = nsm * args.waves;
  float v[kMaxSms / 32];
  float sum = 0.0f;
  int cnt = 0;
#pragma unroll
  for (int k = 0; k < kMaxSms / 32; ++k) {
    const int s = lane + 32 * k;
    v[k] = s < nsm ? g_sm_speed[s] : 0.0f;
    if (v[k] > 0.0f) sum += v[k], ++cnt;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    sum += __shfl_xor_sync(0xffffffffu, sum, off);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
  }
  const float mean = cnt > 0 ? sum / cnt : 0.0f;
  const float even = static_cast<float>(total_gran) / grid;
  int total = 0, biggest = 0, mine = 0;
#pragma unroll
  for (int k = 0; k < kMaxSms / 32; ++k) {
    const int s = lane + 32 * k;
    if (s < nsm) {
      float w = (mean > 0.0f && v[k] > 0.0f) ? v[k] / mean : 1.0f;
      w = fminf(fmaxf(w, 0.7f), 1.25f);
      const int share = static_cast<int>(w * even);
      total += share;
      biggest = max(biggest, share);
      if (s == static_cast<int>(my_sm)) mine = share;
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    total += __shfl_xor_sync(0xffffffffu, total, off);
    biggest = max(biggest, __shfl_xor_sync(0xffffffffu, biggest, off));
    mine = max(mine, __shfl_xor_sync(0xffffffffu, mine, off));
  }
  // Shares were rounded down: hand the missing granules out evenly, the
  // remainder to the lowest SM ids, so that the shares add up to the batch.
  const int deficit = max(0, total_gran - total * args.waves);
  const int add = deficit / grid, rem = deficit - add * grid;
  const int bump_below = (rem + args.waves - 1) / args.waves;
  if (biggest + add + 1 > cap || my_sm >= static_cast<unsigned int>(nsm))
    return min(cap, (total_gran + grid - 1) / grid);
  return mine + add + (static_cast<int>(my_sm) < bump_below ? 1 : 0);
}

template <int A_CT, bool GRADS>
__global__ void __launch_bounds__(32 * (kSmMaxConsumerWarps + 1))
    fused_sm_tile_kernel(const __grid_constant__ SmTileArgs args) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const FusedArgs& a = args.f;
  const int A = A_CT > 0 ? A_CT : a.A;
  const int T = a.T, B = a.B;
  const int NG = args.groups, R = args.relay, depth = a.depth;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int consumers = NG * R;
  const int cols = 32 * NG;  // slot stride in columns

  // This tile's column range, in granules of 4 columns (16-byte rows).  With a
  // cursor: speed-weighted shares handed out first come, first served.
  // Without: B/4 granules dealt evenly by block index.
  __shared__ int sh_range[2];
  const int total_gran = B >> 2;
  const unsigned long long t_begin = global_timer_ns();
  if (args.cursor) {
    if (warp == 0) {
      const int share = weighted_share(args, total_gran, cols >> 2, lane, sm_id());
      if (lane == 0) {
        const unsigned int start = atomicAdd(args.cursor, static_cast<unsigned int>(share * 4));
        sh_range[0] = static_cast<int>(min(start, static_cast<unsigned int>(B)));
        sh_range[1] = max(0, min(share * 4, B - sh_range[0]));
      }
    }
  } else if (threadIdx.x == 0) {
    const int G = gridDim.x, c = blockIdx.x;
    const int base = total_gran / G, rem = total_gran % G;
    sh_range[0] = (c * base + min(c, rem)) * 4;
    sh_range[1] = (base + (c < rem ? 1 : 0)) * 4;
  }

  const int slot_floats = cols * (2 * A + 4);
  float* ring = reinterpret_cast<float*>(smem_raw + kSmBarrierBytes);
  float* staging = ring + static_cast<size_t>(depth) * slot_floats;
  float* carry_buf = staging + (GRADS ? consumers * args.staging_bufs * 32 * A : 0);
  double* lane_sums = reinterpret_cast<double*>(carry_buf + consumers * 96);  // [consumer][k][lane]
  const uint32_t full_bars = ptx::smem_u32(smem_raw);
  const uint32_t empty_bars = full_bars + 128u;
  const uint32_t carry_bars = full_bars + 256u;
  const uint32_t ring_u32 = ptx::smem_u32(ring);

  if (threadIdx.x == 0) {
    for (int s = 0; s < depth; ++s) {
      ptx::mbar_init(full_bars + 8u * s, 1);
      ptx::mbar_init(empty_bars + 8u * s, NG);
    }
    for (int w = 0; w < consumers; ++w) ptx::mbar_init(carry_bars + 8u * w, 32);
    ptx::fence_mbar_init();
    ptx::fence_proxy_async();
  }
  __syncthreads();
  const int col0 = sh_range[0], ncols = sh_range[1];

  if (warp == consumers) {
    // ---- producer: one lane keeps the ring full ----
    if (lane == 0 && ncols > 0) {
      const uint64_t policy = ptx::policy_evict_first();
      const uint32_t logit_bytes = static_cast<uint32_t>(ncols) * A * 4u;
      const uint32_t oper_bytes = static_cast<uint32_t>(ncols) * 4u;
      const uint32_t tx = 2u * logit_bytes + (a.discounts ? 4u : 3u) * oper_bytes;
      for (int it = 0; it < T; ++it) {
        const int s = it % depth, n = it / depth;
        if (n > 0) ptx::mbar_wait(empty_bars + 8u * s, (n - 1) & 1);
        const int t = T - 1 - it;
        const uint32_t bar = full_bars + 8u * s;
        const uint32_t dst = ring_u32 + static_cast<uint32_t>(s) * slot_floats * 4u;
        const size_t e0 = static_cast<size_t>(t) * B + col0;
        ptx::mbar_arrive_expect_tx(bar, tx);
        ptx::bulk_g2s(dst, a.blog + e0 * A, logit_bytes, bar, policy);
        ptx::bulk_g2s(dst + cols * A * 4u, a.tlog + e0 * A, logit_bytes, bar, policy);
        const uint32_t op = dst + 2u * cols * A * 4u;
        ptx::bulk_g2s(op, a.actions + e0, oper_bytes, bar, policy);
        if (a.discounts) ptx::bulk_g2s(op + cols * 4u, a.discounts + e0, oper_bytes, bar, policy);
        ptx::bulk_g2s(op + 2u * cols * 4u, a.rewards + e0, oper_bytes, bar, policy);
        ptx::bulk_g2s(op + 3u * cols * 4u, a.values + e0, oper_bytes, bar, policy);
      }
    }
  } else if (ncols > 0) {
    // ---- consumers ----
    const int g = warp % NG, r = warp / NG;
    const int j = 32 * g + lane;  // column within the tile
    const bool active = j < ncols;
    const int b = col0 + (active ? j : 0);
    const int nb = max(0, min(32, ncols - 32 * g));  // columns of this group
    const uint32_t row_bytes = static_cast<uint32_t>(nb) * A * 4u;
    const bool need_softmax = GRADS || a.partials != nullptr;
    float* stage_base = staging + static_cast<size_t>(warp) * args.staging_bufs * 32 * A;
    int stage_sel = 0;

    math::Carry carry;  // R == 1: lives here across steps
    carry.acc = 0.0f;
    carry.v_next = active ? a.bootstrap[b] : 0.0f;
    carry.vs_next = carry.v_next;
    math::LossSums sums = {0.0, 0.0, 0.0};
    // `done` flags (experiment.py:384) come straight from global memory, one
    // owned step ahead.
    uint8_t done_next = 0;
    if (a.done && active && r < T) done_next = a.done[static_cast<size_t>(T - 1 - r) * B + b];

    for (int it = r; it < T; it += R) {
      const int t = T - 1 - it;
      const size_t e = static_cast<size_t>(t) * B + b;
      const int s = it % depth;
      const uint8_t done_cur = done_next;
      if (a.done && active && it + R < T) done_next = a.done[e - static_cast<size_t>(R) * B];

      ptx::mbar_wait(full_bars + 8u * s, (it / depth) & 1);
      const float* slot = ring + static_cast<size_t>(s) * slot_floats;
      const float* xb_row = slot + j * A;
      const float* xt_row = slot + cols * A + j * A;
      const float* op = slot + 2 * cols * A;

      // ---- phase A: everything that does not depend on step t+1 ----
      int act = 0;
      float disc = 0.0f, rew = 0.0f, val = 0.0f, xa_t = 0.0f, xa_b = 0.0f;
      math::RowStats st = {0.0f, 1.0f, 0.0f}, sb = {0.0f, 1.0f, 0.0f};
      float x[A_CT > 0 ? A_CT : 1], ex[A_CT > 0 ? A_CT : 1];  // target row, its exp terms
      if (active) {
        act = __float_as_int(op[j]);
        act = min(max(act, 0), A - 1);
        disc = a.done ? (done_cur ? 0.0f : a.discounting) : op[cols + j];
        rew = op[2 * cols + j];
        val = op[3 * cols + j];
        xa_t = xt_row[act];
        xa_b = xb_row[act];
        if constexpr (A_CT > 0) {
          float xb[A_CT];
          load_row(xb, xb_row);
          sb = math::row_stats(xb);
          load_row(x, xt_row);
        }
      }
      if constexpr (A_CT > 0) {
        // Rows are in registers: give the slot back to the producer.
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(empty_bars + 8u * s);
      }
      float t_lp = 0.0f, b_lp = 0.0f, log_rho = 0.0f, entropy = 0.0f, inv_sum = 0.0f;
      math::Weights w = {0.0f, 0.0f, 0.0f};
      if (active) {
        rew = math::clip_reward(rew, a.reward_clip_mode);
        if constexpr (A_CT > 0) {
          st = math::row_stats(x, ex);
        } else {
          st = math::smem_row_stats(xt_row, A);
          sb = math::smem_row_stats(xb_row, A);
        }
        t_lp = math::action_log_prob(xa_t, st);
        b_lp = math::action_log_prob(xa_b, sb);
        log_rho = __fsub_rn(t_lp, b_lp);  // vtrace.py:147
        w = math::importance_weights(log_rho, a.flags, a.clip_rho, a.clip_pg_rho);
        if (need_softmax) {
          inv_sum = __frcp_rn(st.sum);
          if constexpr (A_CT > 0) entropy = math::row_entropy(x, ex, st, inv_sum);
          else entropy = math::smem_row_entropy(xt_row, A, st, inv_sum);
        }
      }

      // ---- phase B: the V-trace step (vtrace.py:243-276); needs step t+1 ----
      if (R > 1 && it > 0) {
        // n-th use of this warp's carry slot (step 0 of the tile has none)
        const int use = it / R - (r == 0 ? 1 : 0);
        ptx::mbar_wait(carry_bars + 8u * warp, use & 1);
        const float* cb = carry_buf + warp * 96;
        carry.acc = cb[lane], carry.v_next = cb[32 + lane], carry.vs_next = cb[64 + lane];
      }
      float vs, pg;
      math::vtrace_step(w, rew, disc, val, carry, vs, pg);
      if (R > 1 && it + 1 < T) {
        const int nw = (r + 1 == R ? 0 : r + 1) * NG + g;  // owner of step it+1
        float* cb = carry_buf + nw * 96;
        cb[lane] = carry.acc, cb[32 + lane] = carry.v_next, cb[64 + lane] = carry.vs_next;
        ptx::mbar_arrive(carry_bars + 8u * nw);
      }

      // ---- phase C: outputs, gradient row, loss terms ----
      float* orow_base = stage_base + stage_sel * 32 * A;
      float* orow = orow_base + lane * A;
      if (GRADS) {
        // the bulk store that last read this staging row must be done with it
        if (lane == 0) {
          if (args.staging_bufs == 2) ptx::bulk_wait_read<1>();
          else ptx::bulk_wait_read<0>();
        }
        __syncwarp();
      }
      if (active) {
        if (a.vs) a.vs[e] = vs;
        if (a.pg) a.pg[e] = pg;
        if (a.log_rhos) a.log_rhos[e] = log_rho;
        if (a.b_lp) a.b_lp[e] = b_lp;
        if (a.t_lp) a.t_lp[e] = t_lp;
        if (GRADS) {
          // d/dlogits = (p - onehot) pg_adv + entropy_cost p (log p + H)
          const float ka = __fmul_rn(a.entropy_cost, inv_sum);
          const float kb = __fmul_rn(__fmaf_rn(a.entropy_cost, entropy, pg), inv_sum);
          if constexpr (A_CT > 0) {
            math::row_gradient(ex, x, ka, kb);
            store_row(orow, ex);
          } else {
            math::smem_row_gradient(orow, xt_row, A, st, ka, kb);
          }
          orow[act] = __fsub_rn(orow[act], pg);  // the -onehot(a) * pg_adv term
          if (a.d_values) a.d_values[e] = __fmul_rn(a.baseline_cost, __fsub_rn(val, vs));
        }
        if (a.partials) sums.add(t_lp, pg, vs, val, entropy);
      }
      if (GRADS) {
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0 && nb > 0) {
          ptx::bulk_s2g(a.d_logits + (static_cast<size_t>(t) * B + col0 + 32 * g) * A,
                        ptx::smem_u32(orow_base), row_bytes);
          ptx::bulk_commit();
        }
        if (args.staging_bufs == 2) stage_sel ^= 1;
      }
      if constexpr (A_CT == 0) {
        // run-time A re-reads the rows from the slot: release it only now
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(empty_bars + 8u * s);
      }
    }

    if (a.partials) {
      double* mine = lane_sums + static_cast<size_t>(warp) * 96;
      mine[lane] = sums.pg, mine[32 + lane] = sums.baseline, mine[64 + lane] = sums.entropy;
    }
    if (GRADS && lane == 0) ptx::bulk_wait<0>();
  }

  if (a.partials) {
    // One partial per granule of 4 columns, summed over the relay warps and then
    // over the 4 lanes in a fixed order: the same numbers however the batch
    // was cut into tiles.
    __syncthreads();
    for (int q = threadIdx.x; q < (ncols >> 2); q += blockDim.x) {
      const int g = (4 * q) >> 5, l0 = (4 * q) & 31;
      double p[3] = {0.0, 0.0, 0.0};
      for (int l = l0; l < l0 + 4; ++l)
        for (int r = 0; r < R; ++r) {
          const double* src = lane_sums + static_cast<size_t>(r * NG + g) * 96;
          p[0] += src[l], p[1] += src[32 + l], p[2] += src[64 + l];
        }
      double* out = a.partials + 3 * (static_cast<size_t>(col0 >> 2) + q);
      out[0] = p[0], out[1] = p[1], out[2] = p[2];
    }
  }
  if (a.partials && args.cursor && (a.losses || a.losses_f64)) {
    // Second stage inside the same launch: the tile that finishes last (a
    // ticket, the only other atomic besides the cursor; no value is ever
    // accumulated atomically) sums all granule partials in a fixed order.
    __shared__ int sh_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) sh_last = atomicAdd(args.cursor + 1, 1u) == gridDim.x - 1;
    __syncthreads();
    if (sh_last) {
      __threadfence();
      const volatile double* parts = a.partials;
      double p0 = 0.0, p1 = 0.0, p2 = 0.0;
      for (int i = threadIdx.x; i < total_gran; i += blockDim.x)
        p0 += parts[3 * i], p1 += parts[3 * i + 1], p2 += parts[3 * i + 2];
      p0 = warp_sum_fixed(p0), p1 = warp_sum_fixed(p1), p2 = warp_sum_fixed(p2);
      double* warp_tot = lane_sums;  // free again: every tile partial has been written
      if (lane == 0) warp_tot[3 * warp] = p0, warp_tot[3 * warp + 1] = p1, warp_tot[3 * warp + 2] = p2;
      __syncthreads();
      if (threadIdx.x == 0) {
        double pg = 0.0, bl = 0.0, ent = 0.0;
        for (int w = 0; w <= consumers; ++w)
          pg += warp_tot[3 * w], bl += warp_tot[3 * w + 1], ent += warp_tot[3 * w + 2];
        const double total = pg + static_cast<double>(a.baseline_cost) * bl +
                             static_cast<double>(a.entropy_cost) * ent;
        if (a.losses) {
          a.losses[0] = static_cast<float>(pg), a.losses[1] = static_cast<float>(bl);
          a.losses[2] = static_cast<float>(ent), a.losses[3] = static_cast<float>(total);
        }
        if (a.losses_f64)
          a.losses_f64[0] = pg, a.losses_f64[1] = bl, a.losses_f64[2] = ent, a.losses_f64[3] = total;
      }
    }
  }
  if (threadIdx.x == 0 && ncols > 0) {
    const unsigned int sm = sm_id();
    const unsigned long long dt = global_timer_ns() - t_begin;
    // Tiles shorter than 20 us measure launch latency, not the SM: leave the table alone.
    if (sm < kMaxSms && dt > 20000) {
      const float speed = static_cast<float>(ncols) * T / static_cast<float>(dt);
      const float old = g_sm_speed[sm];
      g_sm_speed[sm] = old > 0.0f ? 0.5f * (old + speed) : speed;
    }
  }
}

// ---------------------------------------------------------------------------
// K2: item-parallel kernel for small batches.  One CTA per batch column, one
// thread per time step: every thread loads its own two logits rows straight
// from global memory, does everything that does not depend on step t+1, and
// leaves (gain_t, delta_t) in shared memory; thread 0 then runs the recurrence
// acc_t = gain_t acc_{t+1} + delta_t over the T entries (T dependent FMAs, a
// fraction of a microsecond) and every thread finishes its own step.  The
// arithmetic is the same sequence of operations as in K1, so the results are
// bit-identical to K1's.  With B columns x T steps all in flight at once, a
// small batch costs one memory round trip instead of T sequential steps.
// Needs T <= kItemMaxSteps and a compile-time A.
// ---------------------------------------------------------------------------
constexpr int kItemMaxSteps = 256;
constexpr int kItemMaxColumns = 16384;  // K2 writes one loss partial per column

template <int N>
__device__ __forceinline__ void load_row_global(float (&x)[N], const float* row) {
  if (N % 2 == 0 && (reinterpret_cast<uintptr_t>(row) & 7) == 0) {
#pragma unroll
    for (int i = 0; i < N / 2; ++i) {
      const float2 v = __ldg(reinterpret_cast<const float2*>(row) + i);
      x[2 * i] = v.x, x[2 * i + 1] = v.y;
    }
  } else {
#pragma unroll
    for (int i = 0; i < N; ++i) x[i] = __ldg(row + i);
  }
}

template <int N>
__device__ __forceinline__ void store_row_global(float* row, const float (&x)[N]) {
  if (N % 2 == 0 && (reinterpret_cast<uintptr_t>(row) & 7) == 0) {
#pragma unroll
    for (int i = 0; i < N / 2; ++i) reinterpret_cast<float2*>(row)[i] = make_float2(x[2 * i], x[2 * i + 1]);
  } else {
#pragma unroll
    for (int i = 0; i < N; ++i) row[i] = x[i];
  }
}

template <int A_CT, bool GRADS>
__global__ void __launch_bounds__(kItemMaxSteps) fused_item_kernel(const __grid_constant__ FusedArgs a) {
  static_assert(A_CT > 0, "the item kernel keeps rows in registers");
  __shared__ float sh_gain[kItemMaxSteps], sh_delta[kItemMaxSteps], sh_acc[kItemMaxSteps];
  __shared__ double sh_sums[3][kItemMaxSteps / 32];
  const int T = a.T, B = a.B;
  const int k = threadIdx.x;  // step index from the end: t = T-1-k
  const int b = blockIdx.x;
  const bool active = k < T;
  const int t = T - 1 - (active ? k : 0);
  const size_t e = static_cast<size_t>(t) * B + b;
  const bool need_softmax = GRADS || a.partials != nullptr;

  float disc = 0.0f, rew = 0.0f, val = 0.0f, v_next = 0.0f, t_lp = 0.0f, b_lp = 0.0f, log_rho = 0.0f;
  float entropy = 0.0f, inv_sum = 0.0f;
  int act = 0;
  math::Weights w = {0.0f, 0.0f, 0.0f};
  float x[A_CT], ex[A_CT];
  if (active) {
    act = min(max(__ldg(a.actions + e), 0), A_CT - 1);
    disc = a.done ? (a.done[e] ? 0.0f : a.discounting) : __ldg(a.discounts + e);
    rew = math::clip_reward(__ldg(a.rewards + e), a.reward_clip_mode);
    val = __ldg(a.values + e);
    v_next = k == 0 ? __ldg(a.bootstrap + b) : __ldg(a.values + e + B);  // vtrace.py:243-244
    float xb[A_CT];
    load_row_global(xb, a.blog + e * A_CT);
    load_row_global(x, a.tlog + e * A_CT);
    const float xa_b = __ldg(a.blog + e * A_CT + act), xa_t = __ldg(a.tlog + e * A_CT + act);
    const math::RowStats sb = math::row_stats(xb);
    const math::RowStats st = math::row_stats(x, ex);
    t_lp = math::action_log_prob(xa_t, st);
    b_lp = math::action_log_prob(xa_b, sb);
    log_rho = __fsub_rn(t_lp, b_lp);  // vtrace.py:147
    w = math::importance_weights(log_rho, a.flags, a.clip_rho, a.clip_pg_rho);
    if (need_softmax) {
      inv_sum = __frcp_rn(st.sum);
      entropy = math::row_entropy(x, ex, st, inv_sum);
    }
    sh_gain[k] = math::step_gain(w, disc);
    sh_delta[k] = math::step_delta(w, rew, disc, val, v_next);
  }
  __syncthreads();
  if (k == 0) {
    float acc = 0.0f;  // vtrace.py:256
    for (int i = 0; i < T; ++i) {
      acc = math::step_acc(sh_gain[i], acc, sh_delta[i]);
      sh_acc[i] = acc;
    }
  }
  __syncthreads();

  double s_pg = 0.0, s_bl = 0.0, s_ent = 0.0;
  if (active) {
    const float vs = __fadd_rn(sh_acc[k], val);  // vtrace.py:265
    // vs_{t+1} = acc_{t+1} + V(x_{t+1}); bootstrap_value past the end (:268-269)
    const float vs_next = k == 0 ? v_next : __fadd_rn(sh_acc[k - 1], v_next);
    const float pg = math::step_pg(w, rew, disc, val, vs_next);
    if (a.vs) a.vs[e] = vs;
    if (a.pg) a.pg[e] = pg;
    if (a.log_rhos) a.log_rhos[e] = log_rho;
    if (a.b_lp) a.b_lp[e] = b_lp;
    if (a.t_lp) a.t_lp[e] = t_lp;
    if (GRADS) {
      const float ka = __fmul_rn(a.entropy_cost, inv_sum);
      const float kb = __fmul_rn(__fmaf_rn(a.entropy_cost, entropy, pg), inv_sum);
      math::row_gradient(ex, x, ka, kb);
#pragma unroll
      for (int i = 0; i < A_CT; ++i)
        if (i == act) ex[i] = __fsub_rn(ex[i], pg);  // the -onehot(a) * pg_adv term
      store_row_global(a.d_logits + e * A_CT, ex);
      if (a.d_values) a.d_values[e] = __fmul_rn(a.baseline_cost, __fsub_rn(val, vs));
    }
    if (a.partials) {
      math::LossSums one = {0.0, 0.0, 0.0};
      one.add(t_lp, pg, vs, val, entropy);
      s_pg = one.pg, s_bl = one.baseline, s_ent = one.entropy;
    }
  }
  if (a.partials) {
    // fixed order: xor tree inside each warp, then the warps in index order
    s_pg = warp_sum_fixed(s_pg), s_bl = warp_sum_fixed(s_bl), s_ent = warp_sum_fixed(s_ent);
    if ((k & 31) == 0) sh_sums[0][k >> 5] = s_pg, sh_sums[1][k >> 5] = s_bl, sh_sums[2][k >> 5] = s_ent;
    __syncthreads();
    if (k == 0) {
      double p0 = 0.0, p1 = 0.0, p2 = 0.0;
      for (int wi = 0; wi < (blockDim.x >> 5); ++wi)
        p0 += sh_sums[0][wi], p1 += sh_sums[1][wi], p2 += sh_sums[2][wi];
      double* out = a.partials + 3 * static_cast<size_t>(b);
      out[0] = p0, out[1] = p1, out[2] = p2;
    }
  }
}

// K3: second stage of the loss reduction.  One block, fixed summation order.
__global__ void __launch_bounds__(256) reduce_losses_kernel(const double* __restrict__ partials,
                                                            int n, float baseline_cost,
                                                            float entropy_cost,
                                                            float* __restrict__ losses,
                                                            double* __restrict__ losses_f64) {
  __shared__ double sh[3][256];
  const int tid = threadIdx.x;
  double s0 = 0.0, s1 = 0.0, s2 = 0.0;
  for (int i = tid; i < n; i += 256) {
    s0 += partials[3 * i], s1 += partials[3 * i + 1], s2 += partials[3 * i + 2];
  }
  sh[0][tid] = s0, sh[1][tid] = s1, sh[2][tid] = s2;
  __syncthreads();
  for (int off = 128; off > 0; off >>= 1) {
    if (tid < off) {
      sh[0][tid] += sh[0][tid + off];
      sh[1][tid] += sh[1][tid + off];
      sh[2][tid] += sh[2][tid + off];
    }
    __syncthreads();
  }
  if (tid == 0) {
    const double pg = sh[0][0], bl = sh[1][0], ent = sh[2][0];
    const double total = pg + static_cast<double>(baseline_cost) * bl +
                         static_cast<double>(entropy_cost) * ent;
    if (losses) {
      losses[0] = static_cast<float>(pg), losses[1] = static_cast<float>(bl);
      losses[2] = static_cast<float>(ent), losses[3] = static_cast<float>(total);
    }
    if (losses_f64) losses_f64[0] = pg, losses_f64[1] = bl, losses_f64[2] = ent, losses_f64[3] = total;
  }
}

static inline int column_tiles(int B) { return (B + kColumnsPerTile - 1) / kColumnsPerTile; }

static bool specialised_actions(int A) { return A == 3 || A == 4 || A == 6 || A == 9 || A == 18; }

// vtr_fused_describe(): the launch functions below record what they would
// launch instead of launching it.
struct PlanRecorder {
  bool dry_run = false;
  int kernels = 0;
  char text[256] = "";
};
static thread_local PlanRecorder g_plan;

static void record_plan(const char* fmt, ...) {
  const size_t used = strlen(g_plan.text);
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_plan.text + used, sizeof(g_plan.text) - used, fmt, ap);
  va_end(ap);
  ++g_plan.kernels;
}

template <int A_CT, bool GRADS>
static int launch_tiles(const FusedArgs& args, int tiles, int warps, size_t smem, cudaStream_t stream) {
  if (g_plan.dry_run) {
    record_plan("%s<A=%d,%s> grid=%d warps/tile=%d ring=%d smem=%zu; ",
                warps > 1 ? "fused_relay_kernel" : "fused_column_kernel", A_CT, GRADS ? "grads" : "fwd",
                tiles, warps, args.depth, smem);
    return VTR_OK;
  }
  auto kernel = warps > 1 ? fused_relay_kernel<A_CT, GRADS> : fused_column_kernel<A_CT, GRADS>;
  VTR_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(smem)));
  kernel<<<tiles, 32 * warps, smem, stream>>>(args);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

template <bool GRADS>
static int dispatch_tiles(const FusedArgs& args, int tiles, int warps, size_t smem, cudaStream_t stream) {
  switch (args.A) {
    case 3: return launch_tiles<3, GRADS>(args, tiles, warps, smem, stream);
    case 4: return launch_tiles<4, GRADS>(args, tiles, warps, smem, stream);
    case 6: return launch_tiles<6, GRADS>(args, tiles, warps, smem, stream);
    case 9: return launch_tiles<9, GRADS>(args, tiles, warps, smem, stream);
    case 18: return launch_tiles<18, GRADS>(args, tiles, warps, smem, stream);
    default: return launch_tiles<0, GRADS>(args, tiles, warps, smem, stream);
  }
}

// Tuning overrides for experiments (not part of the ABI):
// VTR_COLUMN_WARPS in [1,16], VTR_COLUMN_DEPTH in [1,16].
static int env_int(const char* name) {
  const char* v = getenv(name);
  return v ? atoi(v) : 0;
}

static size_t slot_bytes_for(int A) {
  return (2u * kColumnsPerTile * A + 4u * kColumnsPerTile) * sizeof(float);
}

// Shared memory of a tile with `warps` warps and `depth` ring slots.
static size_t tile_smem_bytes(int A, int depth, int warps, bool grads) {
  const size_t row = kColumnsPerTile * A * sizeof(float);
  if (warps == 1) return kColumnBarrierBytes + depth * slot_bytes_for(A) + (grads ? 2 * row : 0);
  const size_t per_warp = (grads ? row : 0) + 96 * sizeof(float) + 3 * sizeof(double);
  return kRelayBarrierBytes + depth * slot_bytes_for(A) + warps * per_warp;
}

static void fill_args(FusedArgs& a, const vtr_fused_params* p, double* partials) {
  a.T = p->T, a.B = p->B, a.A = p->A;
  a.flags = p->flags;
  a.reward_clip_mode = p->reward_clip_mode;
  a.clip_rho = p->clip_rho_threshold, a.clip_pg_rho = p->clip_pg_rho_threshold;
  a.baseline_cost = p->baseline_cost, a.entropy_cost = p->entropy_cost;
  a.discounting = p->discounting;
  a.blog = p->behaviour_policy_logits, a.tlog = p->target_policy_logits;
  a.actions = p->actions, a.discounts = p->done ? nullptr : p->discounts;
  a.rewards = p->rewards, a.values = p->values, a.bootstrap = p->bootstrap_value;
  a.done = p->done;
  a.vs = p->vs, a.pg = p->pg_advantages, a.log_rhos = p->log_rhos;
  a.b_lp = p->behaviour_action_log_probs, a.t_lp = p->target_action_log_probs;
  a.d_logits = p->d_target_logits, a.d_values = p->d_values;
  a.partials = partials;
  a.losses = nullptr, a.losses_f64 = nullptr;
  a.depth = 0, a.slot_floats = 0, a.tma_in_ok = 0, a.tma_out_ok = 0;
}

// K1c needs every row it bulk-copies to start and end on 16 bytes.
static bool sm_tile_eligible(const vtr_fused_params* p) {
  if (p->B % 4 != 0) return false;
  const void* ptrs[] = {p->behaviour_policy_logits, p->target_policy_logits, p->actions, p->rewards,
                        p->values, p->done ? nullptr : p->discounts, p->d_target_logits};
  for (const void* q : ptrs)
    if (!aligned(q, 16)) return false;
  return true;
}

// VTR_PATH_AUTO: K1c from two 32-column groups per SM upwards (measured on
// B200 at T=100, A=18: B=16384 81 us vs 90 us for K1b, B=49152 220 us vs
// 278 us); below that its single producer lane is the bottleneck and K1b,
// where every warp loads its own steps, is faster (B=4096: 26 us vs 45 us).
static bool sm_tile_preferred(const vtr_fused_params* p) {
  DeviceInfo dev;
  if (device_info(&dev) != VTR_OK) return false;
  // Short unrolls leave no time to amortise filling the shared ring
  // (T=20 B=16384: 28.8 us against 22.7 us for K1b).
  return sm_tile_eligible(p) && p->B >= 64 * dev.sm_count && p->T >= 32;
}

template <int A_CT, bool GRADS>
static int launch_sm_tile(const SmTileArgs& args, int grid, size_t smem, cudaStream_t stream) {
  if (g_plan.dry_run) {
    record_plan("fused_sm_tile_kernel<A=%d,%s> grid=%d groups=%d relay=%d ring=%d smem=%zu %s; ", A_CT,
                GRADS ? "grads" : "fwd", grid, args.groups, args.relay, args.f.depth, smem,
                args.cursor ? "speed-weighted tiles, in-kernel final reduction" : "even tiles");
    return VTR_OK;
  }
  auto kernel = fused_sm_tile_kernel<A_CT, GRADS>;
  VTR_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(smem)));
  kernel<<<grid, 32 * (args.groups * args.relay + 1), smem, stream>>>(args);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

template <bool GRADS>
static int dispatch_sm_tile(const SmTileArgs& args, int grid, size_t smem, cudaStream_t stream) {
  switch (args.f.A) {
    case 3: return launch_sm_tile<3, GRADS>(args, grid, smem, stream);
    case 4: return launch_sm_tile<4, GRADS>(args, grid, smem, stream);
    case 6: return launch_sm_tile<6, GRADS>(args, grid, smem, stream);
    case 9: return launch_sm_tile<9, GRADS>(args, grid, smem, stream);
    case 18: return launch_sm_tile<18, GRADS>(args, grid, smem, stream);
    default: return launch_sm_tile<0, GRADS>(args, grid, smem, stream);
  }
}

// Workspace layout (vtr_fused_workspace_bytes): a 64-byte header whose first
// word is K1c's column cursor, then the loss partials.
constexpr size_t kWorkspaceHeaderBytes = 64;

// K1c launch: `*num_partials` = partials written (one per granule of 4 columns).
// `header` is the 64-byte workspace header, or null when the caller gave no
// workspace (then tiles are dealt evenly instead of by measured SM speed).
// Returns VTR_ERR_UNSUPPORTED without launching if the shape does not fit this
// kernel's shared memory.
static int run_sm_tile_path(const vtr_fused_params* p, void* header, double* partials,
                            cudaStream_t stream, int* num_partials, bool* losses_done) {
  DeviceInfo dev;
  if (int rc = device_info(&dev)) return rc;
  const bool grads = p->d_target_logits != nullptr;
  const int sm = dev.sm_count;
  const int total_gran = p->B / 4;

  // Grid: a multiple of the SM count once the batch is wide enough to give
  // every SM at least one 32-column group; tiles of at most ~256 columns.
  int waves = 1, grid;
  if (p->B < 32 * sm) {
    grid = (p->B + 31) / 32;
  } else {
    waves = (p->B + sm * 224 - 1) / (sm * 224);
    grid = sm * waves;
  }
  const bool balance = header != nullptr && grid % sm == 0 && sm <= kMaxSms &&
                       !getenv("VTR_SM_NO_BALANCE");
  const int even_cols = ((total_gran + grid - 1) / grid) * 4;
  // Tile capacity: the even share, or with balancing 1.25x + 2 granules of it
  // (weighted_share clamps to 1.25x and adds at most a couple of granules).
  int cap_cols = even_cols;
  if (balance) {
    int headroom = env_int("VTR_SM_HEADROOM");  // percent, experiments only
    if (headroom <= 0) headroom = 25;
    cap_cols = even_cols + even_cols * headroom / 100 + 8;
  }
  SmTileArgs args;
  fill_args(args.f, p, partials);
  args.groups = (cap_cols + 31) / 32;
  args.num_sm = sm;
  args.waves = waves;
  args.cursor = balance ? static_cast<unsigned int*>(header) : nullptr;
  if (balance) args.f.losses = p->losses, args.f.losses_f64 = p->losses_f64;
  *losses_done = balance;

  // Consumer warps: ~16 per SM.  Two tiles share an SM when the batch needs
  // two or more waves, so that the waves overlap instead of queueing.
  const int co_resident = waves >= 2 ? 2 : 1;
  const int max_consumers = kSmMaxConsumerWarps / co_resident;
  if (args.groups > max_consumers) return fail(VTR_ERR_UNSUPPORTED, "sm-tile kernel: tile too wide");
  int relay = max(1, max_consumers / args.groups);
  if (relay * args.groups > 16 / co_resident && relay > 2) relay = max(2, 16 / co_resident / args.groups);
  if (const int forced = env_int("VTR_SM_RELAY")) relay = forced;
  if (relay > p->T) relay = p->T;
  if (relay * args.groups > max_consumers) relay = max_consumers / args.groups;
  if (relay < 1) relay = 1;

  size_t budget = static_cast<size_t>(dev.smem_per_sm) / co_resident;
  budget = budget > 2048 ? budget - 1024 : budget;
  if (budget > static_cast<size_t>(dev.max_smem_optin)) budget = dev.max_smem_optin;
  const size_t row = 32 * static_cast<size_t>(p->A) * sizeof(float);
  const size_t slot_bytes = 32u * args.groups * (2u * p->A + 4u) * sizeof(float);
  size_t fixed = 0;
  long depth = 0;
  for (;; --relay) {
    args.relay = relay;
    args.staging_bufs = relay == 1 ? 2 : 1;
    fixed = kSmBarrierBytes + args.groups * relay *
                                  ((grads ? args.staging_bufs * row : 0) + 96 * sizeof(float) +
                                   96 * sizeof(double));
    depth = budget > fixed ? static_cast<long>((budget - fixed) / slot_bytes) : 0;
    if (const int forced = env_int("VTR_COLUMN_DEPTH")) depth = forced;
    if (depth > kMaxDepth) depth = kMaxDepth;
    if (depth > p->T) depth = p->T;
    // A warp waits for use n of a slot by parity, which only means something
    // once use n-1 has completed.  The relay order guarantees that for steps
    // more than `relay` back, so a reused slot must be further back than that.
    if (relay == 1 || depth >= p->T || relay < depth) break;
  }
  if (depth < 2 && p->T >= 2) return fail(VTR_ERR_UNSUPPORTED, "sm-tile kernel: ring does not fit");
  if (depth < 1) depth = 1;
  args.f.depth = static_cast<int>(depth);
  const size_t smem = fixed + depth * slot_bytes;
  if (balance && !g_plan.dry_run)
    VTR_CUDA_CHECK(cudaMemsetAsync(header, 0, kWorkspaceHeaderBytes, stream));
  *num_partials = total_gran;
  return grads ? dispatch_sm_tile<true>(args, grid, smem, stream)
               : dispatch_sm_tile<false>(args, grid, smem, stream);
}

// VTR_PATH_AUTO takes K2 for the smallest batches only: its threads read their
// rows straight from global memory, 72-byte pieces 2 MB apart, which is fine
// for a few hundred columns and slower than K1b's ring beyond (measured at
// T=100 A=18: B=256 14 us vs 25 us, B=512 20 us vs 25 us, B=1024 29 us vs 26 us,
// B=2048 50 us vs 23 us).
constexpr int kItemAutoColumns = 512;
static bool item_kernel_applies(const vtr_fused_params* p) {
  return p->T <= kItemMaxSteps && specialised_actions(p->A) && p->B <= kItemAutoColumns;
}

template <int A_CT, bool GRADS>
static int launch_item(const FusedArgs& args, cudaStream_t stream) {
  const int threads = ((args.T + 31) / 32) * 32;
  if (g_plan.dry_run) {
    record_plan("fused_item_kernel<A=%d,%s> grid=%d threads=%d; ", A_CT, GRADS ? "grads" : "fwd", args.B,
                threads);
    return VTR_OK;
  }
  fused_item_kernel<A_CT, GRADS><<<args.B, threads, 0, stream>>>(args);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

template <bool GRADS>
static int dispatch_item(const FusedArgs& args, cudaStream_t stream) {
  switch (args.A) {
    case 3: return launch_item<3, GRADS>(args, stream);
    case 4: return launch_item<4, GRADS>(args, stream);
    case 6: return launch_item<6, GRADS>(args, stream);
    case 9: return launch_item<9, GRADS>(args, stream);
    case 18: return launch_item<18, GRADS>(args, stream);
    default: return fail(VTR_ERR_UNSUPPORTED, "item kernel: A=%d is not a compiled-in size", args.A);
  }
}

static int run_item_path(const vtr_fused_params* p, double* partials, cudaStream_t stream,
                         int* num_partials) {
  if (p->T > kItemMaxSteps || p->B > kItemMaxColumns)
    return fail(VTR_ERR_UNSUPPORTED, "item kernel: needs T <= %d and B <= %d", kItemMaxSteps,
                kItemMaxColumns);
  FusedArgs a;
  fill_args(a, p, partials);
  *num_partials = p->B;  // one partial per column
  return p->d_target_logits ? dispatch_item<true>(a, stream) : dispatch_item<false>(a, stream);
}

static int run_column_path(const vtr_fused_params* p, double* partials, cudaStream_t stream,
                           int* num_partials) {
  DeviceInfo dev;
  if (int rc = device_info(&dev)) return rc;
  const int tiles = column_tiles(p->B);
  const bool grads = p->d_target_logits != nullptr;
  *num_partials = tiles;

  FusedArgs a;
  fill_args(a, p, partials);

  // A row of the tile for step t starts at element (t*B + 32*tile)*A; TMA bulk
  // copies need 16-byte alignment of every such address.
  const bool rows_aligned = (static_cast<size_t>(p->B) * p->A) % 4 == 0;
  a.tma_in_ok = rows_aligned && aligned(a.blog, 16) && aligned(a.tlog, 16);
  a.tma_out_ok = rows_aligned && aligned(a.d_logits, 16);

  // Shared memory per tile = what the SM has, divided by the tiles that must be
  // co-resident on it (everything runs in one wave up to 8 tiles per SM).
  const size_t slot_bytes = slot_bytes_for(p->A);
  int resident = (tiles + dev.sm_count - 1) / dev.sm_count;
  resident = resident < 1 ? 1 : (resident > 8 ? 8 : resident);
  size_t budget = static_cast<size_t>(dev.smem_per_sm) / resident;
  budget = budget > 2048 ? budget - 1024 : budget;  // per-block reservation
  if (budget > static_cast<size_t>(dev.max_smem_optin)) budget = dev.max_smem_optin;

  // Warps per tile.  With 7+ tiles per SM the ring is worth more than extra
  // warps (measured: T=100 B=32768 A=18 runs 142 us with one warp and 5 slots,
  // 166 us with 3 warps and 4 slots); below that, aim at ~24 warps per SM while
  // the ring still runs at least 2 slots ahead of the warps.
  int warps = resident >= 7 ? 1 : 24 / resident;
  if (const int forced = env_int("VTR_COLUMN_WARPS")) warps = forced;
  if (warps > kMaxWarps) warps = kMaxWarps;
  if (warps > p->T) warps = p->T;
  if (warps < 1) warps = 1;
  long depth = 0;
  for (;; --warps) {
    const size_t fixed = tile_smem_bytes(p->A, 0, warps, grads);
    depth = budget > fixed ? static_cast<long>((budget - fixed) / slot_bytes) : 0;
    if (warps == 1 || depth >= warps + 2) break;
  }
  if (const int forced = env_int("VTR_COLUMN_DEPTH")) depth = forced;
  if (depth > kMaxDepth) depth = kMaxDepth;
  if (depth > p->T) depth = p->T;
  if (depth < 2) depth = p->T < 2 ? 1 : 2;
  // Waiting on a slot by parity needs the slot's previous use to be complete;
  // the relay order guarantees that only for steps more than `warps` back.
  if (depth < p->T && warps >= depth) warps = static_cast<int>(depth) - 1;
  if (warps < 1) warps = 1;
  const size_t smem = tile_smem_bytes(p->A, static_cast<int>(depth), warps, grads);
  if (smem > static_cast<size_t>(dev.max_smem_optin))
    return fail(VTR_ERR_UNSUPPORTED, "A=%d needs %zu bytes of shared memory, device allows %d",
                p->A, smem, dev.max_smem_optin);
  a.depth = static_cast<int>(depth);
  a.slot_floats = static_cast<int>(slot_bytes / sizeof(float));

  return grads ? dispatch_tiles<true>(a, tiles, warps, smem, stream)
               : dispatch_tiles<false>(a, tiles, warps, smem, stream);
}

}  // namespace vtr

extern "C" size_t vtr_fused_workspace_bytes(int32_t T, int32_t B, int32_t A) {
  if (T <= 0 || B <= 0 || A <= 0) return 0;
  // header + one double[3] per 32-column tile (K1a/K1b), per 4-column granule
  // (K1c) or per column (K2, which only runs below kItemMaxColumns columns)
  size_t partials = static_cast<size_t>(vtr::column_tiles(B));
  if (static_cast<size_t>(B / 4) > partials) partials = B / 4;
  if (B <= vtr::kItemMaxColumns && static_cast<size_t>(B) > partials) partials = B;
  return vtr::kWorkspaceHeaderBytes + partials * 3 * sizeof(double);
}

extern "C" int vtr_reduce_losses(const double* partials, int32_t num_partials,
                                 float baseline_cost, float entropy_cost, float* losses,
                                 double* losses_f64, void* stream) {
  if (!partials || num_partials <= 0) return vtr::fail(VTR_ERR_INVALID_ARGUMENT, "vtr_reduce_losses: no partials");
  if (!losses && !losses_f64) return vtr::fail(VTR_ERR_INVALID_ARGUMENT, "vtr_reduce_losses: no output");
  if (vtr::g_plan.dry_run) {
    vtr::record_plan("reduce_losses_kernel grid=1 partials=%d; ", num_partials);
    return VTR_OK;
  }
  vtr::reduce_losses_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      partials, num_partials, baseline_cost, entropy_cost, losses, losses_f64);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

extern "C" int vtr_from_logits_fused_f32(const vtr_fused_params* p, void* workspace,
                                         size_t workspace_bytes, void* stream) {
  using vtr::fail;
  if (!p) return fail(VTR_ERR_INVALID_ARGUMENT, "params is NULL");
  if (p->T <= 0 || p->B <= 0 || p->A <= 0)
    return fail(VTR_ERR_INVALID_ARGUMENT, "T, B, A must be positive (got %d, %d, %d)", p->T, p->B, p->A);
  if (p->A > VTR_MAX_ACTIONS)
    return fail(VTR_ERR_UNSUPPORTED, "A=%d exceeds VTR_MAX_ACTIONS=%d", p->A, VTR_MAX_ACTIONS);
  if (!p->behaviour_policy_logits || !p->target_policy_logits || !p->actions || !p->rewards ||
      !p->values || !p->bootstrap_value)
    return fail(VTR_ERR_INVALID_ARGUMENT, "a required input pointer is NULL");
  if (!p->discounts && !p->done)
    return fail(VTR_ERR_INVALID_ARGUMENT, "one of discounts / done must be given");
  if (p->reward_clip_mode < VTR_REWARD_CLIP_NONE || p->reward_clip_mode > VTR_REWARD_CLIP_SOFT_ASYMMETRIC)
    return fail(VTR_ERR_UNSUPPORTED, "unknown reward_clip_mode %d", p->reward_clip_mode);
  if (p->path < VTR_PATH_AUTO || p->path > VTR_PATH_ITEM)
    return fail(VTR_ERR_UNSUPPORTED, "unknown path %d", p->path);
  if ((p->d_values != nullptr) && !p->d_target_logits)
    return fail(VTR_ERR_INVALID_ARGUMENT, "d_values requires d_target_logits");
  const void* ptrs[] = {p->behaviour_policy_logits, p->target_policy_logits, p->actions, p->discounts,
                        p->rewards, p->values, p->bootstrap_value, p->vs, p->pg_advantages,
                        p->log_rhos, p->behaviour_action_log_probs, p->target_action_log_probs,
                        p->d_target_logits, p->d_values, p->losses};
  for (const void* q : ptrs)
    if (!vtr::aligned(q, 4)) return fail(VTR_ERR_MISALIGNED, "a float/int32 pointer is not 4-byte aligned");
  if (!vtr::aligned(p->losses_f64, 8)) return fail(VTR_ERR_MISALIGNED, "losses_f64 is not 8-byte aligned");

  const bool want_losses = p->losses || p->losses_f64;
  double* partials = nullptr;
  if (want_losses) {
    // K1a/K1b write one partial per 32-column tile, K1c one per 4-column granule.
    const size_t need = p->path == VTR_PATH_COLUMN
                            ? vtr::kWorkspaceHeaderBytes + vtr::column_tiles(p->B) * 3 * sizeof(double)
                            : vtr_fused_workspace_bytes(p->T, p->B, p->A);
    if (p->path == VTR_PATH_ITEM && p->B > vtr::kItemMaxColumns)
      return fail(VTR_ERR_UNSUPPORTED, "VTR_PATH_ITEM needs B <= %d", vtr::kItemMaxColumns);
    if (!workspace || workspace_bytes < need)
      return fail(VTR_ERR_WORKSPACE, "workspace of %zu bytes needed, %zu given", need, workspace_bytes);
    if (!vtr::aligned(workspace, 8)) return fail(VTR_ERR_MISALIGNED, "workspace is not 8-byte aligned");
    partials = reinterpret_cast<double*>(static_cast<char*>(workspace) + vtr::kWorkspaceHeaderBytes);
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int num_partials = 0;
  int rc = VTR_ERR_UNSUPPORTED;
  const bool want_sm_tile = p->path == VTR_PATH_SM_TILE ||
                            (p->path == VTR_PATH_AUTO && vtr::sm_tile_preferred(p));
  if (p->path == VTR_PATH_SM_TILE && !vtr::sm_tile_eligible(p))
    return fail(VTR_ERR_UNSUPPORTED,
                "VTR_PATH_SM_TILE needs B %% 4 == 0 and 16-byte aligned tensors");
  bool losses_done = false;
  if (want_sm_tile)
    rc = vtr::run_sm_tile_path(p, want_losses ? workspace : nullptr, partials, s, &num_partials,
                               &losses_done);
  if (rc == VTR_ERR_UNSUPPORTED && p->path != VTR_PATH_SM_TILE) {
    losses_done = false;
    if (p->path == VTR_PATH_ITEM || (p->path == VTR_PATH_AUTO && vtr::item_kernel_applies(p)))
      rc = vtr::run_item_path(p, partials, s, &num_partials);
    else
      rc = vtr::run_column_path(p, partials, s, &num_partials);
  }
  if (rc) return rc;
  if (want_losses && !losses_done) {
    return vtr_reduce_losses(partials, num_partials, p->baseline_cost, p->entropy_cost, p->losses,
                             p->losses_f64, stream);
  }
  return VTR_OK;
}

extern "C" int vtr_fused_describe(const vtr_fused_params* p, int have_workspace, int* num_kernels,
                                  char* text, size_t text_bytes) {
  vtr::g_plan = vtr::PlanRecorder();
  vtr::g_plan.dry_run = true;
  // Validation wants a workspace pointer when losses are requested; nothing is
  // written through it in a dry run.
  static double dummy[8];
  const int rc = vtr_from_logits_fused_f32(p, have_workspace ? dummy : nullptr,
                                           have_workspace ? static_cast<size_t>(-1) : 0, nullptr);
  vtr::g_plan.dry_run = false;
  if (rc != VTR_OK) return rc;
  if (text && text_bytes) snprintf(text, text_bytes, "%s", vtr::g_plan.text);
  if (num_kernels) *num_kernels = vtr::g_plan.kernels;
  return VTR_OK;
}

// ---------------------------------------------------------------------------
// Allocation-free variant of the fused pass.
//
// Some callers of this library have no way to hand over a workspace: a
// framework op whose outputs were all declared up front and which gets no
// temporary allocator, or a learner thread inside a process whose device
// memory was carved up at start-up and where cudaMalloc is banned afterwards
// (it can fail once the pool owns the GPU, and it synchronises the device,
// stalling the inference streams that share it).  For them the tile partials
// of the two-stage loss reduction live in the fixed array below instead of a
// caller buffer.  Every GPU that loads the library carries these 192 KiB until
// the process exits, whether or not this entry point is ever called.
// ---------------------------------------------------------------------------
#include <mutex>

namespace vtr {

constexpr int kNoAllocMaxTiles = VTR_NOALLOC_MAX_B / kColumnsPerTile;
constexpr size_t kNoAllocDoubles = kWorkspaceHeaderBytes / sizeof(double) + kNoAllocMaxTiles * 3;
static __device__ __align__(256) double g_noalloc_partials[kNoAllocDoubles];

// In host code `g_noalloc_partials` only names the symbol; the address the
// kernels can use is per device and comes from cudaGetSymbolAddress, asked
// once per device with that device current.
static int noalloc_scratch(double** out) {
  constexpr int kMaxDevices = 64;
  static double* cache[kMaxDevices] = {};
  static std::mutex mu;
  int dev = 0;
  VTR_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(VTR_ERR_UNSUPPORTED, "device index %d", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!cache[dev]) {
    void* ptr = nullptr;
    VTR_CUDA_CHECK(cudaGetSymbolAddress(&ptr, g_noalloc_partials));
    if (!ptr) return fail(VTR_ERR_WORKSPACE, "cudaGetSymbolAddress returned a null scratch address");
    cache[dev] = static_cast<double*>(ptr);
  }
  *out = cache[dev];
  return VTR_OK;
}

}  // namespace vtr

extern "C" int vtr_from_logits_fused_noalloc_f32(const vtr_fused_params* p, void* stream) {
  if (!p) return vtr::fail(VTR_ERR_INVALID_ARGUMENT, "params is NULL");
  if (p->B > VTR_NOALLOC_MAX_B)
    return vtr::fail(VTR_ERR_UNSUPPORTED,
                     "vtr_from_logits_fused_noalloc_f32: B=%d is above VTR_NOALLOC_MAX_B=%d, the most "
                     "the built-in scratch holds; use vtr_from_logits_fused_f32 with your own workspace",
                     p->B, VTR_NOALLOC_MAX_B);
  double* scratch = nullptr;
  if (int rc = vtr::noalloc_scratch(&scratch)) return rc;
  // The scratch holds one partial per 32-column tile of the largest batch.  The
  // other kernels write more partials per column, so they are only eligible
  // while their workspace still fits; beyond that this entry point takes the
  // 32-column-tile kernels.  The partials never have to start at zero: the
  // first stage writes every one the second stage reads, so whatever an
  // earlier call left behind is overwritten or never looked at.  The per-SM
  // kernel keeps a cursor in the scratch header; the library zeroes that
  // itself before every launch that uses it.
  const size_t scratch_bytes = sizeof(double) * vtr::kNoAllocDoubles;
  vtr_fused_params q = *p;
  if (q.path == VTR_PATH_AUTO && vtr_fused_workspace_bytes(q.T, q.B, q.A) > scratch_bytes)
    q.path = VTR_PATH_COLUMN;
  return vtr_from_logits_fused_f32(&q, scratch, scratch_bytes, stream);
}
