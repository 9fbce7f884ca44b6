// K1a of the fused pass (see vtrace_fused.cu).
#pragma once
#include "fused_common.cuh"
#include "fused_finish.cuh"

namespace vtr {

// ---------------------------------------------------------------------------
// K1a: one warp per tile.  Used when an SM has to hold so many tiles (>= 7)
// that shared memory allows only one warp's worth of ring per tile: the whole
// ring (5 slots at A=18) then runs ahead of the single consumer.
// ---------------------------------------------------------------------------
template <int A_CT, bool GRADS>
__global__ void __launch_bounds__(32) fused_column_kernel(const __grid_constant__ TileArgs args) {
  const FusedArgs& a = args.f;
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
  // Programmatic dependent launch (launch_config() in vtrace_fused.cu): no
  // tensor is touched before this point.
  ptx::grid_launch_dependents();
  ptx::grid_dependency_wait();
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
    if (args.fin.ticket) finish_losses(a, args.fin);
  }
  // the bulk stores must have read their staging rows before the CTA goes away
  if (GRADS && tma_out && lane == 0) ptx::bulk_wait_read<0>();
}


}  // namespace vtr
