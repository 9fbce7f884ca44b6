// K1c of the fused pass (see vtrace_fused.cu).
#pragma once
#include "exchange.cuh"
#include "fused_common.cuh"

namespace vtr {

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
//   sums       per consumer warp 3 x 32 doubles (per-lane loss sums)
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
static __device__ unsigned int g_sm_samples[kMaxSms];  // measurements behind each entry

struct SmTileArgs {
  FusedArgs f;
  int groups;        // column groups (warps side by side) per tile = capacity / 32
  int relay;         // warps per column group, owning interleaved time steps
  int staging_bufs;  // 1 or 2 gradient staging rows per consumer warp
  int num_sm;        // SMs of the device
  int waves;         // tiles per SM (grid = num_sm * waves) when balancing
  unsigned int* cursor;  // 256 counters, zero on entry: [0] column cursor and [1] tiles served
                         // (one 64-bit counter), [2] finished groups, [16 + k] granules
                         // of group k written so far; or null
  int self_clean;        // the tile that finishes the reduction zeroes the counters again
  double* group_partials;  // [groups][3], second level of the loss reduction
  int group_gran;          // granules (of 4 columns) per reduction group
  float smoothing;         // least weight of the newest measurement in the speed table
  int share_pct;           // 100; tests shrink every share to force the leftover passes
  // Loss exchange of a sharded batch, done by the tile that finishes the
  // reduction (exchange.cuh); ex_world == 0 turns it off.
  PeerBuffers ex_peers;
  int ex_rank, ex_world;
  uint32_t ex_seq;
  uint32_t ex_lag;  // ex_sums receives the sums of step ex_seq - ex_lag
  float* ex_sums;   // or null: sum over ranks, collected by the same tile
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
__global__ void __launch_bounds__(32 * (kSmMaxConsumerWarps + 1), 1)
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
  //
  // The weighted shares add up to the batch when every SM runs exactly `waves`
  // tiles.  Nothing guarantees that (another stream may hold some SMs; with
  // three or more waves a fast SM takes an extra tile), so the cursor also
  // counts tiles: the one that draws the last ticket knows that nobody comes
  // after it and owns whatever is left - up to a full tile in this pass, the
  // rest in further passes of the loop below.  Too much handed out is harmless:
  // late tiles find the cursor at B and leave.
  __shared__ int sh_range[4];  // first column, columns, drew the last ticket, drew the first ticket
  const int total_gran = B >> 2;
  if (args.cursor) {
    if (warp == 0) {
      int share = weighted_share(args, total_gran, cols >> 2, lane, sm_id());
      if (args.share_pct != 100) share = max(1, share * args.share_pct / 100);
      if (lane == 0) {
        const unsigned long long drawn =
            atomicAdd(reinterpret_cast<unsigned long long*>(args.cursor),
                      (1ull << 32) | static_cast<unsigned long long>(share * 4));
        const unsigned int start = static_cast<unsigned int>(drawn & 0xffffffffull);
        const bool last = static_cast<unsigned int>(drawn >> 32) == gridDim.x - 1;
        sh_range[0] = static_cast<int>(min(start, static_cast<unsigned int>(B)));
        sh_range[1] = max(0, min(last ? cols : share * 4, B - sh_range[0]));
        sh_range[2] = last;
        sh_range[3] = static_cast<unsigned int>(drawn >> 32) == 0u;
      }
    }
  } else if (threadIdx.x == 0) {
    const int G = gridDim.x, c = blockIdx.x;
    const int base = total_gran / G, rem = total_gran % G;
    sh_range[0] = (c * base + min(c, rem)) * 4;
    sh_range[1] = (base + (c < rem ? 1 : 0)) * 4;
    sh_range[2] = 0, sh_range[3] = 0;
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

  // Everything above ran on set-up data only (kernel arguments, the library's
  // counters and speed table); from here on the tensors are touched.
  ptx::grid_dependency_wait();
  // The SM's speed is measured from here: time spent waiting for the previous
  // kernel says nothing about this SM.
  const unsigned long long t_begin = global_timer_ns();

  // Sharded batch with exchange_lag > 0: the sums wanted are those of an
  // earlier step, which do not depend on this launch.  The first tile to start
  // fetches them now (normally they are already in this rank's buffer), next
  // to the streaming instead of behind it at the end of the launch.
  if (warp == 0 && args.cursor && args.ex_world > 0 && args.ex_sums && args.ex_lag > 0 &&
      args.ex_seq > args.ex_lag) {
    __syncwarp();
    if (sh_range[3])
      exchange_collect_warp(args.ex_peers.p[args.ex_rank], 4, args.ex_world, args.ex_seq - args.ex_lag,
                            args.ex_sums, kExchangeTimeoutNs);
  }

  for (int pass = 0;; ++pass) {  // one pass, except for the last ticket's leftovers
  if (threadIdx.x == 0) {
    for (int s = 0; s < depth; ++s) {
      if (pass > 0) ptx::mbar_inval(full_bars + 8u * s), ptx::mbar_inval(empty_bars + 8u * s);
      ptx::mbar_init(full_bars + 8u * s, 1);
      ptx::mbar_init(empty_bars + 8u * s, NG);
    }
    for (int w = 0; w < consumers; ++w) {
      if (pass > 0) ptx::mbar_inval(carry_bars + 8u * w);
      ptx::mbar_init(carry_bars + 8u * w, 32);
    }
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

  __syncthreads();  // the streaming part of this pass is over
  if (pass == 0 && threadIdx.x == 0) {
    const unsigned int sm = sm_id();
    const unsigned long long dt = global_timer_ns() - t_begin;
    // Tiles shorter than 20 us measure launch latency, not the SM: leave the table alone.
    if (ncols > 0 && sm < kMaxSms && dt > 20000) {
      const float speed = static_cast<float>(ncols) * T / static_cast<float>(dt);
      // Running mean over the first 1 / smoothing measurements, exponential
      // after: one tile's time is noisy (+-2 % from call to call on the same
      // addresses), and noise in the table is imbalance in the next call.
      const unsigned int n = g_sm_samples[sm] + 1u;
      const float weight = fmaxf(args.smoothing, 1.0f / static_cast<float>(n));
      const float old = g_sm_speed[sm];
      g_sm_speed[sm] = old > 0.0f ? old + weight * (speed - old) : speed;
      if (n < 1000000u) g_sm_samples[sm] = n;
      __threadfence();
    }
    // The next launch of the stream may now place its tiles; it starts once
    // every tile of this one has got here, so all of its tiles size
    // themselves from the same, complete speed table.  What is left of this
    // launch - the loss reduction - overlaps that set-up.
    ptx::grid_launch_dependents();
  }

  if (a.partials) {
    // One partial per granule of 4 columns, summed over the relay warps and then
    // over the 4 lanes in a fixed order: the same numbers however the batch
    // was cut into tiles.
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
  if (a.partials && args.cursor && (a.losses || a.losses_f64) && ncols > 0) {
    // Second and third stage inside the same launch, without a pass over all
    // partials by one SM.  Granules are grouped in fixed ranges of
    // `group_gran`; a tile adds the number of granules it wrote to each group
    // it overlaps (a ticket: no value is ever accumulated atomically), and the
    // tile that completes a group sums that group's granule partials in index
    // order.  Whoever completes the last group sums the group partials in
    // index order.  Neither level depends on how the batch was cut into tiles.
    __shared__ int sh_flag;
    __threadfence();
    __syncthreads();
    const int G = args.group_gran;
    const int gran0 = col0 >> 2, gran1 = (col0 + ncols) >> 2;  // this tile's granules [gran0, gran1)
    const int num_groups = (total_gran + G - 1) / G;
    for (int grp = gran0 / G; grp <= (gran1 - 1) / G; ++grp) {
      const int lo = grp * G, hi = min(total_gran, lo + G);
      if (threadIdx.x == 0) {
        const unsigned int mine = static_cast<unsigned int>(min(gran1, hi) - max(gran0, lo));
        const unsigned int before = atomicAdd(args.cursor + 16 + grp, mine);
        sh_flag = before + mine == static_cast<unsigned int>(hi - lo);
      }
      __syncthreads();
      if (sh_flag) {
        __threadfence();
        // ld.global.cg: other tiles wrote these; this SM's L1 has never seen them
        double p0 = 0.0, p1 = 0.0, p2 = 0.0;
        for (int i = lo + threadIdx.x; i < hi; i += blockDim.x)
          p0 += __ldcg(a.partials + 3 * i), p1 += __ldcg(a.partials + 3 * i + 1),
              p2 += __ldcg(a.partials + 3 * i + 2);
        p0 = warp_sum_fixed(p0), p1 = warp_sum_fixed(p1), p2 = warp_sum_fixed(p2);
        double* warp_tot = lane_sums;  // free: this tile's granule partials are written
        if (lane == 0) warp_tot[3 * warp] = p0, warp_tot[3 * warp + 1] = p1, warp_tot[3 * warp + 2] = p2;
        __syncthreads();
        if (warp == 0) {
          double g0 = 0.0, g1 = 0.0, g2 = 0.0;
          if (lane == 0) {
            for (int w = 0; w <= consumers; ++w)
              g0 += warp_tot[3 * w], g1 += warp_tot[3 * w + 1], g2 += warp_tot[3 * w + 2];
            double* out = args.group_partials + 3 * grp;
            out[0] = g0, out[1] = g1, out[2] = g2;
            __threadfence();
            sh_flag = atomicAdd(args.cursor + 2, 1u) == static_cast<unsigned int>(num_groups - 1);
          }
          __syncwarp();
          if (sh_flag) {  // every group is complete: the final sum, in group order
            __threadfence();
            double pg = 0.0, bl = 0.0, ent = 0.0;
            for (int k = lane; k < num_groups; k += 32)
              pg += __ldcg(args.group_partials + 3 * k), bl += __ldcg(args.group_partials + 3 * k + 1),
                  ent += __ldcg(args.group_partials + 3 * k + 2);
            pg = warp_sum_fixed(pg), bl = warp_sum_fixed(bl), ent = warp_sum_fixed(ent);
            const double total = pg + static_cast<double>(a.baseline_cost) * bl +
                                 static_cast<double>(a.entropy_cost) * ent;
            if (lane < args.ex_world) {  // lane r stores this rank's four scalars into peer r
              exchange_publish4_to_peer(args.ex_peers.p[lane], static_cast<float>(pg), static_cast<float>(bl),
                                        static_cast<float>(ent), static_cast<float>(total), args.ex_rank,
                                        args.ex_world, args.ex_seq);
            }
            if (args.ex_world > 0 && args.ex_sums && args.ex_lag == 0) {
              __syncwarp();
              exchange_collect_warp(args.ex_peers.p[args.ex_rank], 4, args.ex_world, args.ex_seq,
                                    args.ex_sums, kExchangeTimeoutNs);
            }
            if (lane == 0) {
              if (a.losses) {
                a.losses[0] = static_cast<float>(pg), a.losses[1] = static_cast<float>(bl);
                a.losses[2] = static_cast<float>(ent), a.losses[3] = static_cast<float>(total);
              }
              if (a.losses_f64)
                a.losses_f64[0] = pg, a.losses_f64[1] = bl, a.losses_f64[2] = ent, a.losses_f64[3] = total;
              if (a.total_out) *a.total_out = static_cast<float>(total);
            }
            if (args.self_clean) {
              // Counters that belong to the library go back to zero for the
              // next launch that is handed them.  Every group is complete, so
              // the only accesses still to come are the cursor draws of tiles
              // that have not started yet (they will find no columns left):
              // wait until all gridDim.x tiles have drawn.
              if (lane == 0) {
                const volatile unsigned int* served = args.cursor + 1;
                while (*served < gridDim.x) __nanosleep(200);
              }
              __syncwarp();
              for (int i = lane; i < 256; i += 32) args.cursor[i] = 0u;
            }
          }
        }
      }
      __syncthreads();
    }
  }
  if (!sh_range[2] || col0 + ncols >= B) break;
  __syncthreads();  // everybody is done with this pass's range and barriers
  if (threadIdx.x == 0) {
    sh_range[0] = col0 + ncols;
    sh_range[1] = min(cols, B - (col0 + ncols));
  }
  }  // pass
}


}  // namespace vtr
