// K1b of the fused pass (see vtrace_fused.cu).
#pragma once
#include "fused_common.cuh"
#include "fused_finish.cuh"

namespace vtr {

// ---------------------------------------------------------------------------
// K1b: several warps per tile, passing the recurrence along ("relay").  Warp w
// owns the time steps it = w, w+W, ... (it counts from the end: t = T-1-it);
// each does its whole step in its own registers and only acc_{t+1}, one float
// per lane, travels from the owner of step t+1 to the owner of step t, through
// shared memory and an mbarrier (V_{t+1} is input data and is read ahead of
// time; vs_{t+1} is re-formed as acc_{t+1} + V_{t+1}).  Everything that does
// not depend on the carry (both softmaxes, log_rho, rho, the entropy, delta_t
// and the gain: all but one FMA) overlaps between warps, so a tile advances
// much faster than one warp can; used whenever shared memory leaves room for
// it.
//
// Shared memory of one tile, in bytes from the start:
//   [0,256)    `depth` mbarriers "slot full"      (TMA + cp.async completion)
//   [256,512)  `warps` mbarriers "carry arrived"  (recurrence hand-off)
//   ring       depth slots of slot_floats floats: one time step of the tile each
//   staging    warps x 32*A floats: each warp's d_target_logits row  (GRADS only)
//   carry      warps x 96 floats reserved, the first warps x 32 used: acc per lane
//   sums       warps x 3 doubles: per-warp loss partials
//
// Launch bound 768 threads (never more than 512 are launched) caps the kernel
// at 80 registers, so that 24 warps fit an SM's register file.
// ---------------------------------------------------------------------------
// 768 threads per SM is what the host makes resident (24 warps: two tiles of
// 12, three of 8 ...), so the register budget is that of 768 threads - except
// for a tile that has its SM to itself (SOLO: 16 warps, 128 registers), where
// fewer values have to be recomputed or re-read from the parameter bank in
// every step of a loop that is bound by instruction issue.
// DUO: two tiles of up to 10 warps per SM, 96 registers each.
enum RelayBudget { kRelayShared = 0, kRelaySolo = 1, kRelayDuo = 2 };
template <int A_CT, bool GRADS, int BUDGET = kRelayShared>
__global__ void __launch_bounds__(BUDGET == kRelaySolo ? 512 : (BUDGET == kRelayDuo ? 320 : 768),
                                  BUDGET == kRelayDuo ? 2 : 1)
    fused_relay_kernel(const __grid_constant__ TileArgs args) {
  const FusedArgs& a = args.f;
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
  const uint32_t carry_bars = full_bars + 256u;
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
  // Programmatic dependent launch (launch_config() in vtrace_fused.cu): no
  // tensor is touched before this point.
  ptx::grid_launch_dependents();
  ptx::grid_dependency_wait();
  const uint64_t policy = ptx::policy_evict_first();

  auto issue = [&](int t, int s) {
    issue_step(a, t, ring_u32 + static_cast<uint32_t>(s) * slot_floats * 4u, full_bars + 8u * s, lane,
               b0, nb, A, tma_in, policy);
  };
  // Step `it` lives in ring slot it % depth and is loaded by the warp that owns it.
  for (int it = warp; it < min(depth, T); it += W) issue(T - 1 - it, it);
  // The last warp of the relay has W-1 hops to wait for its first carry.
  if (blockIdx.x == 0 && warp == W - 1) collect_lagged_early(args.fin);

  math::LossSums sums = {0.0, 0.0, 0.0};

  // Running forms of it % depth, (it / depth) & 1 and it / W: no division in the loop.
  // W < depth whenever the ring wraps (the host sees to it), so one
  // conditional subtraction keeps the slot index in range.
  int s = warp % depth;
  uint32_t full_parity = (warp / depth) & 1;
  int my_steps = 0;  // steps this warp has done = it / W
  for (int it = warp; it < T; it += W, ++my_steps) {
    const int t = T - 1 - it;
    const size_t e = static_cast<size_t>(t) * B + b;
    const uint8_t done_cur = (a.done && active) ? a.done[e] : 0;

    ptx::mbar_wait(full_bars + 8u * s, full_parity);
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
    float v_next = 0.0f;  // V(x_{t+1}): input data, not part of the recurrence (vtrace.py:243-244)
    if (active) {
      v_next = it == 0 ? a.bootstrap[b] : __ldg(a.values + e + B);
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

    // ---- phase B: the V-trace step (vtrace.py:243-276).  Only acc_{t+1} comes
    // from the previous warp; everything else of the step is ready. ----
    const float delta = math::step_delta(w, rew, disc, val, v_next);
    const float gain = math::step_gain(w, disc);
    float acc_in = 0.0f;
    if (it > 0) {
      // n-th use of this warp's carry slot: steps warp, warp+W, ... (step 0 has none)
      const int use = my_steps - (warp == 0 ? 1 : 0);
      ptx::mbar_wait(carry_bars + 8u * warp, use & 1);
      acc_in = carry_buf[warp * 32 + lane];
    }
    const float acc = math::step_acc(gain, acc_in, delta);
    if (it + 1 < T) {
      const int nw = warp + 1 == W ? 0 : warp + 1;  // owner of step it+1
      carry_buf[nw * 32 + lane] = acc;
      ptx::mbar_arrive(carry_bars + 8u * nw);
    }
    const float vs = __fadd_rn(acc, val);  // vtrace.py:265
    // vs_{t+1} = acc_{t+1} + V(x_{t+1}), formed as the previous step formed it
    const float pg = math::step_pg(w, rew, disc, val, it == 0 ? v_next : __fadd_rn(acc_in, v_next));

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
    s += W;
    if (s >= depth) s -= depth, full_parity ^= 1u;
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
    if (args.fin.ticket) finish_losses_cta(a, args.fin, warp_sums, true);
  }
  // the bulk stores must have read their staging rows before the CTA goes away
  if (GRADS && tma_out && lane == 0) ptx::bulk_wait_read<0>();
}


}  // namespace vtr
