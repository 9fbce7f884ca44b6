// K5 of the fused pass (see vtrace_fused.cu): time split over a thread-block
// cluster, chunk aggregates of the affine recurrence exchanged through
// distributed shared memory.
#pragma once
#include "fused_common.cuh"
#include "fused_finish.cuh"

namespace vtr {

// ---------------------------------------------------------------------------
// K5: the batch is cut into groups of 32 adjacent columns (lane = column) and
// the T time steps of a group into C chunks of S steps; chunk c of a group is
// CTA c of a cluster of C CTAs (on C different SMs), warp w of that CTA owns
// step it = c*S + w, counted from the end (t = T-1-it).  So a mid-sized batch
// that has fewer column groups than the GPU has SMs still uses all of them,
// and no SM has to issue the instructions of all T steps of a group.
//
// The recurrence acc_t = gain_t acc_{t+1} + delta_t (vtrace.py:250-262) is
// affine in acc_{t+1}, and affine maps compose.  With (P_k, Z_k) = (product of
// the gains, acc given acc = 0 on entry) over the chunk's steps up to k, the
// true value is acc_k = P_k acc_in + Z_k, and the value entering chunk c is the
// composition of the whole-chunk aggregates of chunks 0..c-1 applied to 0.
//   phase A  every warp: its step's rows arrive by TMA bulk copy; both
//            softmaxes, log_rho, rho, entropy - everything but the recurrence -
//            and (gain, delta) to shared memory.  Only a dozen scalars per
//            lane stay in registers; the target rows stay in shared memory.
//   scan     warp 0 runs the S steps from acc = 0, leaves (P_k, Z_k) for every
//            step and stores the chunk's (P, Z) straight into the shared
//            memory of the CTAs of the earlier-time chunks (st.shared::cluster);
//            one cluster barrier.
//   compose  every warp folds the aggregates of chunks 0..c-1, in index order,
//            into the incoming acc and applies it to its own step.
//   phase C  every warp finishes its step: vs, pg_advantages, the gradient row
//            (terms recomputed from the row in shared memory, bulk store),
//            d_values, loss terms; one double[3] partial per CTA; the CTA that
//            finishes last reduces them (fused_finish.cuh).
// The tree of operations is fixed by (T, S), so results are bitwise
// reproducible run to run; acc differs from K1's in the last bits (an FMA with
// the incoming value instead of a chain through it), everything computed
// before the recurrence has the same bits.
//
// Shared memory of one CTA, in bytes from the start:
//   [0,128)   S mbarriers "rows of step w have landed"
//   slots     S x (2*32*A) floats: behaviour and target rows of the warp's step;
//             the behaviour half becomes the staging row of the gradient
//   pz        S x 2 x 32 floats: (gain, delta) of step k, overwritten by the
//             scan with (P_k, Z_k)
//   agg       kChunkMaxChunks x 2 x 32 floats: (P, Z) of the later-time chunks,
//             written by those CTAs
//   sums      S x 3 doubles
// ---------------------------------------------------------------------------
constexpr int kChunkMaxSteps = 13;   // steps per chunk = warps per CTA
constexpr int kChunkMaxChunks = 16;  // CTAs per cluster (8 is the portable limit)
constexpr int kChunkBarrierBytes = 128;

struct ChunkArgs {
  FusedArgs f;
  int steps;   // S
  int chunks;  // C = cluster size = gridDim.x
  FinishArgs fin;
#ifdef VTR_CHUNK_TRACE
  unsigned long long* trace;  // debug builds: 8 time stamps per CTA
#endif
};

#ifdef VTR_CHUNK_TRACE
#define VTR_STAMP(k)                                                                          \
  do {                                                                                        \
    if (args.trace && threadIdx.x == 0) {                                                     \
      unsigned long long now__;                                                               \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now__));                               \
      args.trace[(static_cast<size_t>(blockIdx.y) * gridDim.x + blockIdx.x) * 8 + (k)] = now__; \
    }                                                                                         \
  } while (0)
#else
#define VTR_STAMP(k) \
  do {               \
  } while (0)
#endif

// Registers: with the rows left in shared memory, compile-time A <= 9 fits 32
// registers (four CTAs of 13 warps per SM), A = 18 fits 48 (three).
template <int A_CT, bool GRADS>
__global__ void __launch_bounds__(32 * kChunkMaxSteps, (A_CT > 0 && A_CT <= 9) ? 4 : 3)
    fused_chunk_kernel(const __grid_constant__ ChunkArgs args) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const FusedArgs& a = args.f;
  const int A = A_CT > 0 ? A_CT : a.A;
  const int T = a.T, B = a.B;
  const int S = args.steps, C = args.chunks;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int c = C > 1 ? static_cast<int>(ptx::cluster_cta_rank()) : 0;
  const int b0 = blockIdx.y * kColumnsPerTile;
  const int nb = min(kColumnsPerTile, B - b0);
  const bool active = lane < nb;
  const int b = b0 + (active ? lane : 0);
  const int n = max(0, min(S, T - c * S));  // steps of this chunk
  const bool has_step = warp < n;
  const int it = c * S + warp;
  const int t = has_step ? T - 1 - it : 0;
  const size_t e = static_cast<size_t>(t) * B + b;
  const bool live = has_step && active;

  // Everybody is running (and its shared memory exists) once this phase of the
  // cluster barrier completes; it is waited for just before the first remote
  // store.
  VTR_STAMP(0);
  if (C > 1) ptx::cluster_arrive_relaxed();

  const int slot_floats = 64 * A;
  float* slots = reinterpret_cast<float*>(smem_raw + kChunkBarrierBytes);
  float* sh_pz = slots + static_cast<size_t>(S) * slot_floats;  // [k][0] gain / P_k, [k][1] delta / Z_k
  float* sh_agg = sh_pz + S * 64;
  double* sh_sums = reinterpret_cast<double*>(sh_agg + kChunkMaxChunks * 64);
  float* slot = slots + static_cast<size_t>(warp) * slot_floats;
  const uint32_t slot_u32 = ptx::smem_u32(slot);
  const uint32_t bar = ptx::smem_u32(smem_raw) + 8u * warp;
  const uint32_t row_bytes = static_cast<uint32_t>(nb) * A * 4u;
  const bool tma_in = a.tma_in_ok && (row_bytes % 16u == 0);
  const bool tma_out = a.tma_out_ok && (row_bytes % 16u == 0);
  const bool need_softmax = GRADS || a.partials != nullptr;

  // Programmatic dependent launch (see launch_config() in vtrace_fused.cu):
  // no tensor is touched before this point.
  ptx::grid_launch_dependents();
  ptx::grid_dependency_wait();

  // ---- loads: the two rows through shared memory, the operands directly ----
  // Every warp has its own barrier and waits for nobody else's rows.
  int act = 0;
  float disc = 0.0f, rew = 0.0f, val = 0.0f, v_next = 0.0f;
  if (has_step) {
    if (lane == 0) {
      ptx::mbar_init(bar, 33);
      ptx::fence_mbar_init();
      ptx::fence_proxy_async();
    }
    __syncwarp();
    const size_t row0 = (static_cast<size_t>(t) * B + b0) * A;
    if (tma_in) {
      if (lane == 0) {
        const uint64_t policy = ptx::policy_evict_first();
        ptx::mbar_arrive_expect_tx(bar, 2u * row_bytes);
        ptx::bulk_g2s(slot_u32, a.blog + row0, row_bytes, bar, policy);
        ptx::bulk_g2s(slot_u32 + 32u * A * 4u, a.tlog + row0, row_bytes, bar, policy);
      }
    } else {
      if (lane == 0) ptx::mbar_arrive(bar);
      for (int i = lane; i < nb * A; i += 32) {
        ptx::cp_async4(slot_u32 + 4u * i, a.blog + row0 + i);
        ptx::cp_async4(slot_u32 + (32u * A + i) * 4u, a.tlog + row0 + i);
      }
    }
    ptx::cp_async_mbar_arrive_noinc(bar);
    if (active) {
      act = min(max(__ldg(a.actions + e), 0), A - 1);
      disc = a.done ? (a.done[e] ? 0.0f : a.discounting) : __ldg(a.discounts + e);
      rew = __ldg(a.rewards + e);
      val = __ldg(a.values + e);
      v_next = it == 0 ? __ldg(a.bootstrap + b) : __ldg(a.values + e + B);  // vtrace.py:243-244
    }
  }

  // ---- phase A: everything that does not depend on step t+1 ----
  math::RowStats st = {0.0f, 1.0f, 0.0f};
  const float* xb_row = slot + lane * A;
  const float* xt_row = slot + 32 * A + lane * A;
  VTR_STAMP(1);
  if (has_step) ptx::mbar_wait(bar, 0);
  VTR_STAMP(2);
  float t_lp = 0.0f, b_lp = 0.0f, log_rho = 0.0f, entropy = 0.0f, inv_sum = 0.0f;
  math::Weights w = {0.0f, 0.0f, 0.0f};
  float gain = 0.0f, delta = 0.0f;
  if (live) {
    const float xa_t = xt_row[act], xa_b = xb_row[act];
    rew = math::clip_reward(rew, a.reward_clip_mode);
    math::RowStats sb;
    float dot = 0.0f;
    if constexpr (A_CT > 0) {
      {
        float xb[A_CT];
        load_row(xb, xb_row);
        sb = math::row_stats(xb);
      }
      float x[A_CT];
      load_row(x, xt_row);
      st = math::row_stats_dot(x, dot);
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
      if constexpr (A_CT > 0) entropy = math::entropy_from_dot(st, dot, inv_sum);
      else entropy = math::smem_row_entropy(xt_row, A, st, inv_sum);
    }
    gain = math::step_gain(w, disc);
    delta = math::step_delta(w, rew, disc, val, v_next);
  }
  sh_pz[warp * 64 + lane] = gain, sh_pz[warp * 64 + 32 + lane] = delta;
  VTR_STAMP(3);
  __syncthreads();

  // ---- scan: (P_k, Z_k) for every step, the chunk aggregate to the earlier-time chunks ----
  if (C > 1) ptx::cluster_wait();  // every CTA of the cluster has started
  if (warp == 0) {
    float acc = 0.0f, prod = 1.0f;  // vtrace.py:256
#pragma unroll 4
    for (int k = 0; k < n; ++k) {
      const float g = sh_pz[k * 64 + lane];
      acc = math::step_acc(g, acc, sh_pz[k * 64 + 32 + lane]);
      prod = __fmul_rn(prod, g);
      sh_pz[k * 64 + lane] = prod, sh_pz[k * 64 + 32 + lane] = acc;
    }
    const uint32_t mine = ptx::smem_u32(sh_agg + c * 64 + lane);
    for (int peer = c + 1; peer < C; ++peer) {
      const uint32_t remote = ptx::map_to_cta(mine, static_cast<uint32_t>(peer));
      ptx::st_cluster_f32(remote, prod);
      ptx::st_cluster_f32(remote + 128u, acc);
    }
  }
  if (C > 1) {
    ptx::cluster_arrive_release();
    ptx::cluster_wait();
  } else {
    __syncthreads();
  }
  VTR_STAMP(4);

  // ---- compose: what enters this chunk, applied to this warp's step ----
  math::LossSums sums = {0.0, 0.0, 0.0};
  float* orow_base = slot;  // the behaviour rows are spent
  if (live) {
    float acc_in = 0.0f;
    for (int k = 0; k < c; ++k) acc_in = __fmaf_rn(sh_agg[k * 64 + lane], acc_in, sh_agg[k * 64 + 32 + lane]);
    const float acc = __fmaf_rn(sh_pz[warp * 64 + lane], acc_in, sh_pz[warp * 64 + 32 + lane]);
    const float acc_next =
        warp == 0 ? acc_in : __fmaf_rn(sh_pz[(warp - 1) * 64 + lane], acc_in, sh_pz[(warp - 1) * 64 + 32 + lane]);
    VTR_STAMP(5);

    // ---- phase C: outputs, gradient row, loss terms ----
    const float vs = __fadd_rn(acc, val);  // vtrace.py:265
    // vs_{t+1} = acc_{t+1} + V(x_{t+1}); bootstrap_value past the end (:268-269)
    const float vs_next = it == 0 ? v_next : __fadd_rn(acc_next, v_next);
    const float pg = math::step_pg(w, rew, disc, val, vs_next);
    if (a.vs) a.vs[e] = vs;
    if (a.pg) a.pg[e] = pg;
    if (a.log_rhos) a.log_rhos[e] = log_rho;
    if (a.b_lp) a.b_lp[e] = b_lp;
    if (a.t_lp) a.t_lp[e] = t_lp;
    if (GRADS) {
      // d/dlogits = (p - onehot) pg_adv + entropy_cost p (log p + H)
      float* orow = orow_base + lane * A;
      const float ka = __fmul_rn(a.entropy_cost, inv_sum);
      const float kb = __fmul_rn(__fmaf_rn(a.entropy_cost, entropy, pg), inv_sum);
      if constexpr (A_CT > 0) {
        float x[A_CT];
        load_row(x, xt_row);
        math::row_gradient_from_logits(x, st, ka, kb);
        store_row(orow, x);
      } else {
        math::smem_row_gradient(orow, xt_row, A, st, ka, kb);
      }
      orow[act] = __fsub_rn(orow[act], pg);  // the -onehot(a) * pg_adv term
      if (a.d_values) a.d_values[e] = __fmul_rn(a.baseline_cost, __fsub_rn(val, vs));
    }
    if (a.partials) sums.add(t_lp, pg, vs, val, entropy);
  }
  if (GRADS && has_step) {
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
      for (int i = lane; i < nb * A; i += 32) gdst[i] = orow_base[i];
    }
  }
  VTR_STAMP(6);

  if (a.partials) {
    // fixed order: xor tree inside each warp, then the warps in index order
    const double pg_sum = warp_sum_fixed(sums.pg);
    const double bl_sum = warp_sum_fixed(sums.baseline);
    const double ent_sum = warp_sum_fixed(sums.entropy);
    if (lane == 0) sh_sums[3 * warp] = pg_sum, sh_sums[3 * warp + 1] = bl_sum, sh_sums[3 * warp + 2] = ent_sum;
    __syncthreads();
    if (threadIdx.x == 0) {
      double p0 = 0.0, p1 = 0.0, p2 = 0.0;
      for (int k = 0; k < n; ++k) p0 += sh_sums[3 * k], p1 += sh_sums[3 * k + 1], p2 += sh_sums[3 * k + 2];
      double* out = a.partials + 3 * (static_cast<size_t>(blockIdx.y) * C + c);
      out[0] = p0, out[1] = p1, out[2] = p2;
    }
    if (args.fin.ticket) finish_losses_cta(a, args.fin, sh_sums);
  }
  // the bulk store must have read its staging row before the CTA goes away
  if (GRADS && tma_out && has_step && lane == 0) ptx::bulk_wait_read<0>();
  VTR_STAMP(7);
}

}  // namespace vtr
