// The arithmetic of one (t, column) step, shared by every kernel that performs
// it (fused_column_kernel, fused_relay_kernel).  Written with explicit
// round-to-nearest intrinsics so that no kernel-specific FMA contraction can
// creep in: whichever kernel the launch heuristics pick, a column gets the
// same bits.
//
// Reference: vtrace.py:45-68 (action log-probs), :235-276 (V-trace step),
// experiment.py:324-343 (losses) and the closed-form gradients of the total
// loss.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/vtrace_b200.h"
#include "ptx.cuh"

namespace vtr {
namespace math {

constexpr float kLog2e = 1.4426950408889634f;

// exp(x) for x <= 0 on the MUFU pipe: 2^(x * log2 e).
__device__ __forceinline__ float exp_nonpos(float x) { return ptx::ex2(__fmul_rn(x, kLog2e)); }

// ---- rows held in registers (A known at compile time) ----------------------

// Behaviour row: returns sum_i exp(x_i - max), sets m = max.
template <int N>
__device__ __forceinline__ float row_exp_sum(const float (&x)[N], float& m) {
  m = x[0];
#pragma unroll
  for (int i = 1; i < N; ++i) m = fmaxf(m, x[i]);
  float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const float v = exp_nonpos(__fsub_rn(x[i], m));
    if (i & 1) s1 = __fadd_rn(s1, v); else s0 = __fadd_rn(s0, v);
  }
  return __fadd_rn(s0, s1);
}

// Target row: d <- x - max (d holds x on entry), ex <- exp(d); returns the sum.
template <int N>
__device__ __forceinline__ float row_exp_terms(float (&d)[N], float (&ex)[N], float& m) {
  m = d[0];
#pragma unroll
  for (int i = 1; i < N; ++i) m = fmaxf(m, d[i]);
  float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    d[i] = __fsub_rn(d[i], m);
    ex[i] = exp_nonpos(d[i]);
    if (i & 1) s1 = __fadd_rn(s1, ex[i]); else s0 = __fadd_rn(s0, ex[i]);
  }
  return __fadd_rn(s0, s1);
}

// ex <- softmax p, d <- log-softmax; returns the entropy H = -sum p log p
// (experiment.py:332-334).
template <int N>
__device__ __forceinline__ float row_entropy(float (&d)[N], float (&ex)[N], float inv_sum, float ln_sum) {
  float h0 = 0.0f, h1 = 0.0f;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    ex[i] = __fmul_rn(ex[i], inv_sum);
    d[i] = __fsub_rn(d[i], ln_sum);
    if (i & 1) h1 = __fmaf_rn(-ex[i], d[i], h1); else h0 = __fmaf_rn(-ex[i], d[i], h0);
  }
  return __fadd_rn(h0, h1);
}

// ex <- p * (entropy_cost * log p + k), k = pg_adv + entropy_cost * H: the
// gradient row but for the -onehot(a) * pg_adv term.
template <int N>
__device__ __forceinline__ void row_gradient(float (&ex)[N], const float (&d)[N], float ec, float k) {
#pragma unroll
  for (int i = 0; i < N; ++i) ex[i] = __fmul_rn(ex[i], __fmaf_rn(ec, d[i], k));
}

// ---- rows left in shared memory (A known only at run time) ------------------
// Same operations in the same order as the register versions above.

__device__ __forceinline__ float smem_row_exp_sum(const float* row, int A, float& m) {
  m = row[0];
  for (int i = 1; i < A; ++i) m = fmaxf(m, row[i]);
  float s0 = 0.0f, s1 = 0.0f;
  for (int i = 0; i < A; ++i) {
    const float v = exp_nonpos(__fsub_rn(row[i], m));
    if (i & 1) s1 = __fadd_rn(s1, v); else s0 = __fadd_rn(s0, v);
  }
  return __fadd_rn(s0, s1);
}

__device__ __forceinline__ float smem_row_entropy(const float* row, int A, float m, float inv_sum,
                                                  float ln_sum) {
  float h0 = 0.0f, h1 = 0.0f;
  for (int i = 0; i < A; ++i) {
    const float d = __fsub_rn(row[i], m);
    const float p = __fmul_rn(exp_nonpos(d), inv_sum);
    const float lp = __fsub_rn(d, ln_sum);
    if (i & 1) h1 = __fmaf_rn(-p, lp, h1); else h0 = __fmaf_rn(-p, lp, h0);
  }
  return __fadd_rn(h0, h1);
}

__device__ __forceinline__ void smem_row_gradient(float* out, const float* row, int A, float m,
                                                  float inv_sum, float ln_sum, float ec, float k) {
  for (int i = 0; i < A; ++i) {
    const float d = __fsub_rn(row[i], m);
    const float p = __fmul_rn(exp_nonpos(d), inv_sum);
    out[i] = __fmul_rn(p, __fmaf_rn(ec, __fsub_rn(d, ln_sum), k));
  }
}

// ---- per-step scalars --------------------------------------------------------

// log pi(a) = (x_a - max) - log sum exp(x - max)   (vtrace.py:67)
__device__ __forceinline__ float action_log_prob(float x_a, float m, float ln_sum) {
  return __fsub_rn(__fsub_rn(x_a, m), ln_sum);
}

// experiment.py:377-382
__device__ __forceinline__ float clip_reward(float r, int mode) {
  if (mode == VTR_REWARD_CLIP_ABS_ONE) return fminf(fmaxf(r, -1.0f), 1.0f);
  if (mode == VTR_REWARD_CLIP_SOFT_ASYMMETRIC) {
    const float squeezed = tanhf(__fdiv_rn(r, 5.0f));
    return __fmul_rn(r < 0.0f ? __fmul_rn(0.3f, squeezed) : squeezed, 5.0f);
  }
  return r;
}

struct Weights {
  float rho_hat;  // min(rho_bar, rho) or rho        vtrace.py:236-239
  float c;        // min(1, rho)                      vtrace.py:241
  float rho_pg;   // min(rho_bar_pg, rho) or rho     vtrace.py:270-274
};

__device__ __forceinline__ Weights importance_weights(float log_rho, uint32_t flags, float clip_rho,
                                                      float clip_pg_rho) {
  const float rho = expf(log_rho);  // vtrace.py:235
  Weights w;
  w.rho_hat = (flags & VTR_FLAG_NO_CLIP_RHO) ? rho : fminf(clip_rho, rho);
  w.c = fminf(1.0f, rho);
  w.rho_pg = (flags & VTR_FLAG_NO_CLIP_PG_RHO) ? rho : fminf(clip_pg_rho, rho);
  return w;
}

// What step t needs from step t+1.
struct Carry {
  float acc;      // vs_{t+1} - V(x_{t+1}); 0 past the end (vtrace.py:256)
  float v_next;   // V(x_{t+1}); bootstrap_value past the end (vtrace.py:243-244)
  float vs_next;  // vs_{t+1};   bootstrap_value past the end (vtrace.py:268-269)
};

// One step of the reverse scan (vtrace.py:245-276); updates the carry.
__device__ __forceinline__ void vtrace_step(const Weights& w, float rew, float disc, float val,
                                            Carry& c, float& vs, float& pg) {
  const float delta = __fmul_rn(w.rho_hat, __fsub_rn(__fmaf_rn(disc, c.v_next, rew), val));
  c.acc = __fmaf_rn(__fmul_rn(disc, w.c), c.acc, delta);
  vs = __fadd_rn(c.acc, val);
  pg = __fmul_rn(w.rho_pg, __fsub_rn(__fmaf_rn(disc, c.vs_next, rew), val));
  c.v_next = val;
  c.vs_next = vs;
}

// Per-lane loss sums in double (experiment.py:328,335,343).
struct LossSums {
  double pg, baseline, entropy;
  __device__ __forceinline__ void add(float t_lp, float pg_adv, float vs, float val, float h) {
    pg = fma(-static_cast<double>(t_lp), static_cast<double>(pg_adv), pg);
    const double diff = static_cast<double>(vs) - static_cast<double>(val);
    baseline = fma(0.5 * diff, diff, baseline);
    entropy -= static_cast<double>(h);
  }
};

}  // namespace math
}  // namespace vtr
