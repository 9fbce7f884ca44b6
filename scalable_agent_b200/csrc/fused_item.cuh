// K2 of the fused pass (see vtrace_fused.cu).
#pragma once
#include "fused_common.cuh"
#include "fused_finish.cuh"

namespace vtr {

// ---------------------------------------------------------------------------
// K2: item-parallel kernel for small batches.  One CTA per batch column, one
// thread per time step: every thread loads its own two logits rows straight
// from global memory, does everything that does not depend on step t+1, and
// ends up with its step's affine map acc -> gain_t acc + delta_t; the block
// then runs the recurrence acc_t = gain_t acc_{t+1} + delta_t as a warp-shuffle
// associative scan of those maps, chunked over T by warp (log2(32) shuffle
// rounds plus one fold over at most 8 chunk aggregates instead of T dependent
// steps), and every thread finishes its own step.  With B columns x T steps
// all in flight at once, a small batch costs one memory round trip instead of
// T sequential steps.  The scan re-associates the float32 operations, so acc
// differs from K1's in the last bits (both are within 1e-5 of the float64
// oracle); its tree is fixed, so runs are bitwise reproducible.  Needs
// T <= kItemMaxSteps and a compile-time A.
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

struct ItemArgs {
  FusedArgs f;
  FinishArgs fin;  // ticket == null: the host appends K3
};

template <int A_CT, bool GRADS>
__global__ void __launch_bounds__(kItemMaxSteps) fused_item_kernel(const __grid_constant__ ItemArgs args) {
  static_assert(A_CT > 0, "the item kernel keeps rows in registers");
  __shared__ float sh_gain[kItemMaxSteps / 32], sh_delta[kItemMaxSteps / 32];  // per-warp aggregates
  __shared__ double sh_sums[3 * (kItemMaxSteps / 32)];
  const FusedArgs& a = args.f;
  const int T = a.T, B = a.B;
  const int k = threadIdx.x;  // step index from the end: t = T-1-k
  const int lane = k & 31, warp = k >> 5;
  const int b = blockIdx.x;
  const bool active = k < T;
  const int t = T - 1 - (active ? k : 0);
  const size_t e = static_cast<size_t>(t) * B + b;
  const bool need_softmax = GRADS || a.partials != nullptr;
  // Programmatic dependent launch (launch_config() in vtrace_fused.cu): no
  // tensor is touched before this point.
  ptx::grid_launch_dependents();
  ptx::grid_dependency_wait();

  float disc = 0.0f, rew = 0.0f, val = 0.0f, v_next = 0.0f, t_lp = 0.0f, b_lp = 0.0f, log_rho = 0.0f;
  float entropy = 0.0f, inv_sum = 0.0f;
  int act = 0;
  math::Weights w = {0.0f, 0.0f, 0.0f};
  float x[A_CT], ex[A_CT];
  // acc -> gain acc + delta, the affine map of this step; the identity past the end
  float gain = 1.0f, delta = 0.0f;
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
    gain = math::step_gain(w, disc);
    delta = math::step_delta(w, rew, disc, val, v_next);
  }

  // The reverse recurrence acc_k = gain_k acc_{k-1} + delta_k (vtrace.py:250-262,
  // acc_{-1} = 0) as an associative scan of affine maps, chunked over T: a
  // Kogge-Stone scan by warp shuffles inside each chunk of 32 steps (lane l
  // ends up with the composition of its chunk's steps up to l), the chunks'
  // aggregates folded in order into the value entering each chunk, and one
  // FMA per step to apply it.  The tree is fixed, so the bits are the same on
  // every run.
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const float g_lo = __shfl_up_sync(0xffffffffu, gain, off);
    const float d_lo = __shfl_up_sync(0xffffffffu, delta, off);
    if (lane >= off) {
      delta = __fmaf_rn(gain, d_lo, delta);
      gain = __fmul_rn(gain, g_lo);
    }
  }
  if (lane == 31) sh_gain[warp] = gain, sh_delta[warp] = delta;
  __syncthreads();
  float carry = 0.0f;  // acc entering this warp's chunk (vtrace.py:256 for the first)
  for (int j = 0; j < warp; ++j) carry = __fmaf_rn(sh_gain[j], carry, sh_delta[j]);
  const float acc = __fmaf_rn(gain, carry, delta);
  float acc_next = __shfl_up_sync(0xffffffffu, acc, 1);  // acc of step t+1
  if (lane == 0) acc_next = carry;

  double s_pg = 0.0, s_bl = 0.0, s_ent = 0.0;
  if (active) {
    const float vs = __fadd_rn(acc, val);  // vtrace.py:265
    // vs_{t+1} = acc_{t+1} + V(x_{t+1}); bootstrap_value past the end (:268-269)
    const float vs_next = k == 0 ? v_next : __fadd_rn(acc_next, v_next);
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
    if (lane == 0) sh_sums[3 * warp] = s_pg, sh_sums[3 * warp + 1] = s_bl, sh_sums[3 * warp + 2] = s_ent;
    __syncthreads();
    if (k == 0) {
      double p0 = 0.0, p1 = 0.0, p2 = 0.0;
      for (int wi = 0; wi < (blockDim.x >> 5); ++wi)
        p0 += sh_sums[3 * wi], p1 += sh_sums[3 * wi + 1], p2 += sh_sums[3 * wi + 2];
      double* out = a.partials + 3 * static_cast<size_t>(b);
      out[0] = p0, out[1] = p1, out[2] = p2;
    }
    if (args.fin.ticket) finish_losses_cta(a, args.fin, sh_sums);
  }
}


}  // namespace vtr
