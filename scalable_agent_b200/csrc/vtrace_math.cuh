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
constexpr float kLn2 = 0.6931471805599453f;

// Softmax statistics of one logits row.
struct RowStats {
  float m;        // max_i x_i
  float sum;      // sum_i e_i with e_i = 2^(x_i*log2e + c), c = fl(-m*log2e)
  float ln_sum;   // log sum_i exp(x_i - m), i.e. logsumexp(x) - m
};

// The exponent offset c = fl(-m*log2e) lets each term cost one FMA + one MUFU.
// Its rounding error is the same for every i, so it cancels in e_i / sum and
// shows up in log(sum) as the constant (c_exact - c) * ln 2, which is added
// back: ln_sum is as accurate as if x_i - m had been formed first.
struct ExpOffset {
  float c, ln_fix;
  __device__ __forceinline__ explicit ExpOffset(float m) {
    c = __fmul_rn(-m, kLog2e);
    ln_fix = __fmul_rn(__fmaf_rn(-m, kLog2e, -c), kLn2);
  }
  __device__ __forceinline__ float term(float x) const { return ptx::ex2(__fmaf_rn(x, kLog2e, c)); }
  __device__ __forceinline__ float ln_sum(float sum) const { return __fadd_rn(logf(sum), ln_fix); }
};

// ---- rows held in registers (A known at compile time) ----------------------

template <int N>
__device__ __forceinline__ float row_max(const float (&x)[N]) {
  float m = x[0];
#pragma unroll
  for (int i = 1; i < N; ++i) m = fmaxf(m, x[i]);
  return m;
}

// Behaviour row: only the statistics are needed.
template <int N>
__device__ __forceinline__ RowStats row_stats(const float (&x)[N]) {
  RowStats r;
  r.m = row_max(x);
  const ExpOffset off(r.m);
  float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const float v = off.term(x[i]);
    if (i & 1) s1 = __fadd_rn(s1, v); else s0 = __fadd_rn(s0, v);
  }
  r.sum = __fadd_rn(s0, s1);
  r.ln_sum = off.ln_sum(r.sum);
  return r;
}

// Target row: also keeps the terms e_i (softmax p_i = e_i / sum).
template <int N>
__device__ __forceinline__ RowStats row_stats(const float (&x)[N], float (&e)[N]) {
  RowStats r;
  r.m = row_max(x);
  const ExpOffset off(r.m);
  float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    e[i] = off.term(x[i]);
    if (i & 1) s1 = __fadd_rn(s1, e[i]); else s0 = __fadd_rn(s0, e[i]);
  }
  r.sum = __fadd_rn(s0, s1);
  r.ln_sum = off.ln_sum(r.sum);
  return r;
}

// x <- log-softmax (x_i - logsumexp); returns the entropy H = -sum p log p
// = -(sum_i e_i lp_i) / sum   (experiment.py:332-334).
template <int N>
__device__ __forceinline__ float row_entropy(float (&x)[N], const float (&e)[N], const RowStats& r,
                                             float inv_sum) {
  const float lse = __fadd_rn(r.m, r.ln_sum);
  float h0 = 0.0f, h1 = 0.0f;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    x[i] = __fsub_rn(x[i], lse);
    if (i & 1) h1 = __fmaf_rn(e[i], x[i], h1); else h0 = __fmaf_rn(e[i], x[i], h0);
  }
  return __fmul_rn(-inv_sum, __fadd_rn(h0, h1));
}

// Target row for kernels that do not keep the terms e_i (K5): the statistics
// and dot = sum_i e_i (x_i - m), from which H = -sum p_i log p_i
// = ln_sum - dot / sum (log p_i = (x_i - m) - ln_sum).  The terms and their
// sum are formed exactly as in row_stats(), so the action log-probs carry the
// same bits as in the other kernels.
template <int N>
__device__ __forceinline__ RowStats row_stats_dot(const float (&x)[N], float& dot) {
  RowStats r;
  r.m = row_max(x);
  const ExpOffset off(r.m);
  float s0 = 0.0f, s1 = 0.0f, d0 = 0.0f, d1 = 0.0f;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const float e = off.term(x[i]);
    const float c = __fsub_rn(x[i], r.m);
    if (i & 1) s1 = __fadd_rn(s1, e), d1 = __fmaf_rn(e, c, d1);
    else s0 = __fadd_rn(s0, e), d0 = __fmaf_rn(e, c, d0);
  }
  r.sum = __fadd_rn(s0, s1);
  r.ln_sum = off.ln_sum(r.sum);
  dot = __fadd_rn(d0, d1);
  return r;
}
__device__ __forceinline__ float entropy_from_dot(const RowStats& r, float dot, float inv_sum) {
  return __fsub_rn(r.ln_sum, __fmul_rn(dot, inv_sum));
}

// x <- the gradient row but for the -onehot(a) pg_adv term, e_i (ka lp_i + kb),
// with the terms recomputed from the logits (same bits as row_stats()).
template <int N>
__device__ __forceinline__ void row_gradient_from_logits(float (&x)[N], const RowStats& r, float ka,
                                                         float kb) {
  const ExpOffset off(r.m);
  const float lse = __fadd_rn(r.m, r.ln_sum);
#pragma unroll
  for (int i = 0; i < N; ++i)
    x[i] = __fmul_rn(off.term(x[i]), __fmaf_rn(ka, __fsub_rn(x[i], lse), kb));
}

// e <- p_i (entropy_cost lp_i + k) = e_i (ka lp_i + kb) with ka = ec/sum,
// kb = k/sum, k = pg_adv + entropy_cost H: the gradient row but for the
// -onehot(a) pg_adv term.
template <int N>
__device__ __forceinline__ void row_gradient(float (&e)[N], const float (&lp)[N], float ka, float kb) {
#pragma unroll
  for (int i = 0; i < N; ++i) e[i] = __fmul_rn(e[i], __fmaf_rn(ka, lp[i], kb));
}

// ---- rows left in shared memory (A known only at run time) ------------------
// Same operations in the same order as the register versions above.

__device__ __forceinline__ RowStats smem_row_stats(const float* row, int A) {
  RowStats r;
  r.m = row[0];
  for (int i = 1; i < A; ++i) r.m = fmaxf(r.m, row[i]);
  const ExpOffset off(r.m);
  float s0 = 0.0f, s1 = 0.0f;
  for (int i = 0; i < A; ++i) {
    const float v = off.term(row[i]);
    if (i & 1) s1 = __fadd_rn(s1, v); else s0 = __fadd_rn(s0, v);
  }
  r.sum = __fadd_rn(s0, s1);
  r.ln_sum = off.ln_sum(r.sum);
  return r;
}

__device__ __forceinline__ float smem_row_entropy(const float* row, int A, const RowStats& r,
                                                  float inv_sum) {
  const ExpOffset off(r.m);
  const float lse = __fadd_rn(r.m, r.ln_sum);
  float h0 = 0.0f, h1 = 0.0f;
  for (int i = 0; i < A; ++i) {
    const float e = off.term(row[i]);
    const float lp = __fsub_rn(row[i], lse);
    if (i & 1) h1 = __fmaf_rn(e, lp, h1); else h0 = __fmaf_rn(e, lp, h0);
  }
  return __fmul_rn(-inv_sum, __fadd_rn(h0, h1));
}

__device__ __forceinline__ void smem_row_gradient(float* out, const float* row, int A,
                                                  const RowStats& r, float ka, float kb) {
  const ExpOffset off(r.m);
  const float lse = __fadd_rn(r.m, r.ln_sum);
  for (int i = 0; i < A; ++i)
    out[i] = __fmul_rn(off.term(row[i]), __fmaf_rn(ka, __fsub_rn(row[i], lse), kb));
}

// ---- per-step scalars --------------------------------------------------------

// log pi(a) = (x_a - max) - log sum exp(x - max)   (vtrace.py:67)
__device__ __forceinline__ float action_log_prob(float x_a, const RowStats& r) {
  return __fsub_rn(__fsub_rn(x_a, r.m), r.ln_sum);
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

// The pieces of one step of the reverse scan (vtrace.py:245-276).  Kernels that
// walk a column in one thread use vtrace_step(); the item-parallel kernel
// computes delta and gain per (t, b) item, runs only step_acc() sequentially
// and finishes per item -- the same operations in the same order either way.
__device__ __forceinline__ float step_delta(const Weights& w, float rew, float disc, float val,
                                            float v_next) {  // vtrace.py:245
  return __fmul_rn(w.rho_hat, __fsub_rn(__fmaf_rn(disc, v_next, rew), val));
}
__device__ __forceinline__ float step_gain(const Weights& w, float disc) {  // gamma_t c_t, :253
  return __fmul_rn(disc, w.c);
}
__device__ __forceinline__ float step_acc(float gain, float acc_next, float delta) {  // :253
  return __fmaf_rn(gain, acc_next, delta);
}
__device__ __forceinline__ float step_pg(const Weights& w, float rew, float disc, float val,
                                         float vs_next) {  // vtrace.py:275-276
  return __fmul_rn(w.rho_pg, __fsub_rn(__fmaf_rn(disc, vs_next, rew), val));
}

// One whole step; updates the carry.
__device__ __forceinline__ void vtrace_step(const Weights& w, float rew, float disc, float val,
                                            Carry& c, float& vs, float& pg) {
  const float delta = step_delta(w, rew, disc, val, c.v_next);
  c.acc = step_acc(step_gain(w, disc), c.acc, delta);
  vs = __fadd_rn(c.acc, val);  // vtrace.py:265
  pg = step_pg(w, rew, disc, val, c.vs_next);
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
