// Fused V-trace + IMPALA loss, forward and backward, one pass over HBM.
//
// Replaces vtrace.from_logits (vtrace.py:71-161), the three loss helpers
// (experiment.py:324-343), the weighted total (experiment.py:401-407) and the
// autodiff backward of that total (experiment.py:415).
//
// K1 "column" kernel (this file): one warp owns 32 adjacent batch columns and
// walks time from T-1 down to 0, one lane per column, so the reverse
// recurrence acc_t = delta_t + gamma_t c_t acc_{t+1} (vtrace.py:250-262) and
// vs_{t+1} (vtrace.py:268-269) are live registers.  For one time step the 32
// columns' logits are ONE contiguous run of 32*A floats in the time-major
// [T,B,A] layout; a ring of `depth` stages in shared memory is kept full by
// TMA bulk copies (cp.async.bulk -> mbarrier complete_tx) for both logits
// tensors plus 4-byte cp.async for the four [T,B] operands, all landing on the
// same mbarrier.  d_target_logits is staged in shared memory and leaves with a
// TMA bulk store.  Losses accumulate per lane in double, are reduced over the
// warp with a fixed shuffle tree and written as one partial per tile; K3
// (reduce_losses_kernel) sums the partials in a fixed order.  No atomics.
#include <math.h>

#include "common.h"
#include "ptx.cuh"

namespace vtr {

struct FusedArgs {
  int T, B, A;
  int depth;         // ring stages
  int stage_floats;  // floats per stage: 2*32*A logits + 4*32 per-step operands
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
  double* partials;  // [tiles][3] or null
};

constexpr int kColumnsPerTile = 32;
constexpr int kBarrierBytes = 128;  // up to 16 mbarriers
constexpr int kMaxDepth = 16;

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

// experiment.py:377-382
__device__ __forceinline__ float clip_reward(float r, int mode) {
  if (mode == VTR_REWARD_CLIP_ABS_ONE) return fminf(fmaxf(r, -1.0f), 1.0f);
  if (mode == VTR_REWARD_CLIP_SOFT_ASYMMETRIC) {
    const float squeezed = tanhf(r / 5.0f);
    return (r < 0.0f ? 0.3f * squeezed : squeezed) * 5.0f;
  }
  return r;
}

__device__ __forceinline__ double warp_sum_fixed(double v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

// Everything one (t, column) step produces besides the gradient row.
struct StepOut {
  float vs, pg, log_rho, b_lp, t_lp;
};

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
  const int stage_floats = a.stage_floats;

  float* ring = reinterpret_cast<float*>(smem_raw + kBarrierBytes);
  float* outbuf = ring + static_cast<size_t>(depth) * stage_floats;  // 2 x [32*A], GRADS only
  const uint32_t bars = ptx::smem_u32(smem_raw);
  const uint32_t ring_u32 = ptx::smem_u32(ring);
  const uint32_t row_bytes = static_cast<uint32_t>(nb) * A * 4u;
  const int row_elems = nb * A;
  const bool tma_in = a.tma_in_ok && (row_bytes % 16u == 0);
  const bool tma_out = a.tma_out_ok && (row_bytes % 16u == 0);

  if (lane == 0) {
    for (int s = 0; s < depth; ++s) ptx::mbar_init(bars + 8u * s, 33);
    ptx::fence_mbar_init();
    ptx::fence_proxy_async();
  }
  __syncwarp();
  const uint64_t policy = ptx::policy_evict_first();

  // Fills stage `s` with time step `t` of this tile.
  auto issue = [&](int t, int s) {
    const uint32_t bar = bars + 8u * s;
    const uint32_t dst = ring_u32 + static_cast<uint32_t>(s) * stage_floats * 4u;
    const size_t row0 = (static_cast<size_t>(t) * B + b0) * A;
    if (tma_in) {
      if (lane == 0) {
        ptx::mbar_arrive_expect_tx(bar, 2u * row_bytes);
        ptx::bulk_g2s(dst, a.blog + row0, row_bytes, bar, policy);
        ptx::bulk_g2s(dst + 32u * A * 4u, a.tlog + row0, row_bytes, bar, policy);
      }
    } else {
      if (lane == 0) ptx::mbar_arrive(bar);
      for (int i = lane; i < row_elems; i += 32) {
        ptx::cp_async4(dst + 4u * i, a.blog + row0 + i);
        ptx::cp_async4(dst + (32u * A + i) * 4u, a.tlog + row0 + i);
      }
    }
    if (active) {
      const size_t e = static_cast<size_t>(t) * B + b;
      const uint32_t sc = dst + (64u * A + lane) * 4u;
      ptx::cp_async4(sc, a.actions + e);
      if (a.discounts) ptx::cp_async4(sc + 128u, a.discounts + e);
      ptx::cp_async4(sc + 256u, a.rewards + e);
      ptx::cp_async4(sc + 384u, a.values + e);
    }
    ptx::cp_async_mbar_arrive_noinc(bar);
  };

  const int prologue = min(depth, T);
  for (int k = 0; k < prologue; ++k) issue(T - 1 - k, k);

  const bool no_clip_rho = a.flags & VTR_FLAG_NO_CLIP_RHO;
  const bool no_clip_pg = a.flags & VTR_FLAG_NO_CLIP_PG_RHO;
  const float ec = a.entropy_cost;

  float acc = 0.0f;  // vs_{t+1} - V(x_{t+1}), initial value vtrace.py:256
  float v_next = active ? a.bootstrap[b] : 0.0f;  // V(x_{t+1}), vtrace.py:243-244
  float vs_next = v_next;                         // vs_{t+1},   vtrace.py:268-269
  double pg_sum = 0.0, bl_sum = 0.0, ent_sum = 0.0;
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

    const float* st = ring + static_cast<size_t>(s) * stage_floats;
    const float* xb_row = st + lane * A;
    const float* xt_row = st + 32 * A + lane * A;
    const float* sc = st + 64 * A;
    float* orow = outbuf + (it & 1) * 32 * A + lane * A;

    if (active) {
      int act = __float_as_int(sc[lane]);
      act = min(max(act, 0), A - 1);
      float disc = a.done ? (done_cur ? 0.0f : a.discounting) : sc[32 + lane];
      const float rew = clip_reward(sc[64 + lane], a.reward_clip_mode);
      const float val = sc[96 + lane];

      // ---- action log-probs of both policies (vtrace.py:143-146) ----
      float m_t, sum_t, m_b, sum_b;
      float d[A_CT > 0 ? A_CT : 1], ex[A_CT > 0 ? A_CT : 1];
      if constexpr (A_CT > 0) {
        load_row(d, xt_row);
        m_t = d[0];
#pragma unroll
        for (int i = 1; i < A_CT; ++i) m_t = fmaxf(m_t, d[i]);
        float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
        for (int i = 0; i < A_CT; ++i) {
          d[i] -= m_t;
          ex[i] = ptx::ex2(d[i] * kLog2e);
          if (i & 1) s1 += ex[i]; else s0 += ex[i];
        }
        sum_t = s0 + s1;
        float xb[A_CT];
        load_row(xb, xb_row);
        m_b = xb[0];
#pragma unroll
        for (int i = 1; i < A_CT; ++i) m_b = fmaxf(m_b, xb[i]);
        s0 = 0.0f, s1 = 0.0f;
#pragma unroll
        for (int i = 0; i < A_CT; ++i) {
          const float v = ptx::ex2((xb[i] - m_b) * kLog2e);
          if (i & 1) s1 += v; else s0 += v;
        }
        sum_b = s0 + s1;
      } else {
        m_t = xt_row[0], m_b = xb_row[0];
        for (int i = 1; i < A; ++i) m_t = fmaxf(m_t, xt_row[i]), m_b = fmaxf(m_b, xb_row[i]);
        sum_t = 0.0f, sum_b = 0.0f;
        for (int i = 0; i < A; ++i) {
          const float v = ptx::ex2((xt_row[i] - m_t) * kLog2e);
          sum_t += v;
          if (GRADS) orow[i] = v;
          sum_b += ptx::ex2((xb_row[i] - m_b) * kLog2e);
        }
      }
      const float ln_sum_t = logf(sum_t);
      const float t_lp = (xt_row[act] - m_t) - ln_sum_t;
      const float b_lp = (xb_row[act] - m_b) - logf(sum_b);
      const float log_rho = t_lp - b_lp;  // vtrace.py:147

      // ---- V-trace step (vtrace.py:235-276), walking t downwards ----
      const float rho = expf(log_rho);
      const float rho_hat = no_clip_rho ? rho : fminf(a.clip_rho, rho);
      const float c = fminf(1.0f, rho);
      const float rho_pg = no_clip_pg ? rho : fminf(a.clip_pg_rho, rho);
      const float delta = rho_hat * (rew + disc * v_next - val);
      acc = delta + disc * c * acc;
      const float vs = acc + val;
      const float pg = rho_pg * (rew + disc * vs_next - val);
      v_next = val;
      vs_next = vs;

      if (a.vs) a.vs[e] = vs;
      if (a.pg) a.pg[e] = pg;
      if (a.log_rhos) a.log_rhos[e] = log_rho;
      if (a.b_lp) a.b_lp[e] = b_lp;
      if (a.t_lp) a.t_lp[e] = t_lp;

      // ---- losses (experiment.py:324-343) and their gradients ----
      float entropy = 0.0f;
      if (GRADS || a.partials) {
        const float inv_sum = 1.0f / sum_t;
        if constexpr (A_CT > 0) {
          float h0 = 0.0f, h1 = 0.0f;
#pragma unroll
          for (int i = 0; i < A_CT; ++i) {
            ex[i] *= inv_sum;      // p_i
            d[i] -= ln_sum_t;      // log p_i
            if (i & 1) h1 = fmaf(-ex[i], d[i], h1); else h0 = fmaf(-ex[i], d[i], h0);
          }
          entropy = h0 + h1;
          if (GRADS) {
            const float k = fmaf(ec, entropy, pg);
#pragma unroll
            for (int i = 0; i < A_CT; ++i) ex[i] *= fmaf(ec, d[i], k);
            store_row(orow, ex);
          }
        } else {
          for (int i = 0; i < A; ++i) {
            const float p = (GRADS ? orow[i] : ptx::ex2((xt_row[i] - m_t) * kLog2e)) * inv_sum;
            const float lp = (xt_row[i] - m_t) - ln_sum_t;
            entropy = fmaf(-p, lp, entropy);
          }
          if (GRADS) {
            const float k = fmaf(ec, entropy, pg);
            for (int i = 0; i < A; ++i) {
              const float p = orow[i] * inv_sum;
              const float lp = (xt_row[i] - m_t) - ln_sum_t;
              orow[i] = p * fmaf(ec, lp, k);
            }
          }
        }
        if (GRADS) {
          orow[act] -= pg;  // the -onehot(a) * pg_adv term
          if (a.d_values) a.d_values[e] = a.baseline_cost * (val - vs);
        }
        if (a.partials) {
          pg_sum = fma(-static_cast<double>(t_lp), static_cast<double>(pg), pg_sum);
          const double diff = static_cast<double>(vs) - static_cast<double>(val);
          bl_sum = fma(0.5 * diff, diff, bl_sum);
          ent_sum -= static_cast<double>(entropy);
        }
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
    __syncwarp();  // every lane is done with stage s
    if (it + depth < T) issue(t - depth, s);
    if (++s == depth) s = 0, parity ^= 1u;
  }

  if (a.partials) {
    pg_sum = warp_sum_fixed(pg_sum);
    bl_sum = warp_sum_fixed(bl_sum);
    ent_sum = warp_sum_fixed(ent_sum);
    if (lane == 0) {
      double* out = a.partials + 3 * static_cast<size_t>(blockIdx.x);
      out[0] = pg_sum, out[1] = bl_sum, out[2] = ent_sum;
    }
  }
  if (GRADS && tma_out && lane == 0) ptx::bulk_wait<0>();
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

template <int A_CT, bool GRADS>
static int launch_column(const FusedArgs& args, int tiles, size_t smem, cudaStream_t stream) {
  auto kernel = fused_column_kernel<A_CT, GRADS>;
  VTR_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(smem)));
  kernel<<<tiles, 32, smem, stream>>>(args);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

template <bool GRADS>
static int dispatch_column(const FusedArgs& args, int tiles, size_t smem, cudaStream_t stream) {
  switch (args.A) {
    case 3: return launch_column<3, GRADS>(args, tiles, smem, stream);
    case 4: return launch_column<4, GRADS>(args, tiles, smem, stream);
    case 6: return launch_column<6, GRADS>(args, tiles, smem, stream);
    case 9: return launch_column<9, GRADS>(args, tiles, smem, stream);
    case 18: return launch_column<18, GRADS>(args, tiles, smem, stream);
    default: return launch_column<0, GRADS>(args, tiles, smem, stream);
  }
}

static int run_column_path(const vtr_fused_params* p, double* partials, cudaStream_t stream) {
  DeviceInfo dev;
  if (int rc = device_info(&dev)) return rc;
  const int tiles = column_tiles(p->B);
  const bool grads = p->d_target_logits != nullptr;

  FusedArgs a;
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

  // A row of the tile for step t starts at element (t*B + 32*tile)*A; TMA bulk
  // copies need 16-byte alignment of every such address.
  const bool rows_aligned = (static_cast<size_t>(p->B) * p->A) % 4 == 0;
  a.tma_in_ok = rows_aligned && aligned(a.blog, 16) && aligned(a.tlog, 16);
  a.tma_out_ok = rows_aligned && aligned(a.d_logits, 16);

  // Ring depth: share the SM's shared memory between the tiles that will be
  // co-resident on it, and keep as many time steps in flight as fit.
  const size_t stage_bytes = (2u * kColumnsPerTile * p->A + 4u * kColumnsPerTile) * sizeof(float);
  const size_t out_bytes = grads ? 2u * kColumnsPerTile * p->A * sizeof(float) : 0;
  int resident = (tiles + dev.sm_count - 1) / dev.sm_count;
  resident = resident < 1 ? 1 : (resident > 8 ? 8 : resident);
  size_t budget = static_cast<size_t>(dev.smem_per_sm) / resident;
  budget = budget > 2048 ? budget - 1024 : budget;  // per-block reservation
  if (budget > static_cast<size_t>(dev.max_smem_optin)) budget = dev.max_smem_optin;
  long depth = (static_cast<long>(budget) - static_cast<long>(out_bytes) - kBarrierBytes) /
               static_cast<long>(stage_bytes);
  if (depth > kMaxDepth) depth = kMaxDepth;
  if (depth > p->T) depth = p->T;
  if (depth < 2) depth = p->T < 2 ? 1 : 2;
  const size_t smem = kBarrierBytes + depth * stage_bytes + out_bytes;
  if (smem > static_cast<size_t>(dev.max_smem_optin))
    return fail(VTR_ERR_UNSUPPORTED, "A=%d needs %zu bytes of shared memory, device allows %d",
                p->A, smem, dev.max_smem_optin);
  a.depth = static_cast<int>(depth);
  a.stage_floats = static_cast<int>(stage_bytes / sizeof(float));

  return grads ? dispatch_column<true>(a, tiles, smem, stream)
               : dispatch_column<false>(a, tiles, smem, stream);
}

}  // namespace vtr

extern "C" size_t vtr_fused_workspace_bytes(int32_t T, int32_t B, int32_t A) {
  if (T <= 0 || B <= 0 || A <= 0) return 0;
  return static_cast<size_t>(vtr::column_tiles(B)) * 3 * sizeof(double);
}

extern "C" int vtr_reduce_losses(const double* partials, int32_t num_partials,
                                 float baseline_cost, float entropy_cost, float* losses,
                                 double* losses_f64, void* stream) {
  if (!partials || num_partials <= 0) return vtr::fail(VTR_ERR_INVALID_ARGUMENT, "vtr_reduce_losses: no partials");
  if (!losses && !losses_f64) return vtr::fail(VTR_ERR_INVALID_ARGUMENT, "vtr_reduce_losses: no output");
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
  if (p->path < VTR_PATH_AUTO || p->path > VTR_PATH_SCAN)
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
    const size_t need = vtr_fused_workspace_bytes(p->T, p->B, p->A);
    if (!workspace || workspace_bytes < need)
      return fail(VTR_ERR_WORKSPACE, "workspace of %zu bytes needed, %zu given", need, workspace_bytes);
    if (!vtr::aligned(workspace, 8)) return fail(VTR_ERR_MISALIGNED, "workspace is not 8-byte aligned");
    partials = static_cast<double*>(workspace);
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (int rc = vtr::run_column_path(p, partials, s)) return rc;
  if (want_losses) {
    return vtr_reduce_losses(partials, vtr::column_tiles(p->B), p->baseline_cost, p->entropy_cost,
                             p->losses, p->losses_f64, stream);
  }
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
static __device__ __align__(256) double g_noalloc_partials[kNoAllocMaxTiles * 3];

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
  // Nothing in the scratch has to start at zero: the first stage writes the
  // partial of every tile the second stage reads, so whatever an earlier call
  // left behind is overwritten or never looked at.
  return vtr_from_logits_fused_f32(p, scratch, sizeof(double) * 3 * vtr::kNoAllocMaxTiles, stream);
}
