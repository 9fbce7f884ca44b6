// Stand-alone pieces of the path, for callers that use the reference's
// functions one by one instead of the fused pass:
//   from_importance_weights            vtrace.py:164-280
//   log_probs_from_logits_and_actions  vtrace.py:45-68
//   compute_policy_gradient_loss       experiment.py:338-343  (+ backward)
//   compute_entropy_loss               experiment.py:331-335  (+ backward)
//   compute_baseline_loss              experiment.py:324-328  (+ backward)
// All reductions are two-stage with a fixed order (no atomics).
#include <math.h>
#include <stdlib.h>

#include "common.h"
#include "fused_common.cuh"
#include "ptx.cuh"
#include "vtrace_math.cuh"

namespace vtr {

// ---------------------------------------------------------------------------
// K4: from_importance_weights.  One thread per (b, c) column, walking t down.
// All operands are [T, B*C] (or [T, B] when broadcast), so a warp's accesses
// are coalesced; the loads do not depend on the recurrence, so the unrolled
// loop keeps several time steps of loads in flight per thread.
// ---------------------------------------------------------------------------
struct IwArgs {
  const float* log_rhos;
  const float* discounts;
  const float* rewards;
  const float* values;
  const float* bootstrap;
  float* vs;
  float* pg;
  int T, B, C;
  int rho_bcast, disc_bcast;
  int no_clip_rho, no_clip_pg;
  float clip_rho, clip_pg;
};

__global__ void __launch_bounds__(128) importance_weights_kernel(const __grid_constant__ IwArgs a) {
  const long cols = static_cast<long>(a.B) * a.C;
  const long j = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (j >= cols) return;
  const long jb = j / a.C;  // batch index, for operands with trailing size 1
  const long rho_col = a.rho_bcast ? jb : j, rho_stride = a.rho_bcast ? a.B : cols;
  const long disc_col = a.disc_bcast ? jb : j, disc_stride = a.disc_bcast ? a.B : cols;
  float acc = 0.0f;
  float v_next = a.bootstrap[j];
  float vs_next = v_next;
#pragma unroll 4
  for (int t = a.T - 1; t >= 0; --t) {
    const float log_rho = __ldg(a.log_rhos + t * rho_stride + rho_col);
    const float disc = __ldg(a.discounts + t * disc_stride + disc_col);
    const float rew = __ldg(a.rewards + t * cols + j);
    const float val = __ldg(a.values + t * cols + j);
    const float rho = expf(log_rho);                                   // :235
    const float rho_hat = a.no_clip_rho ? rho : fminf(a.clip_rho, rho);  // :236-239
    const float c = fminf(1.0f, rho);                                  // :241
    const float rho_pg = a.no_clip_pg ? rho : fminf(a.clip_pg, rho);     // :270-274
    const float delta = rho_hat * (rew + disc * v_next - val);         // :245
    acc = delta + disc * c * acc;                                      // :250-253
    const float vs = acc + val;                                        // :265
    const float pg = rho_pg * (rew + disc * vs_next - val);            // :275-276
    a.vs[t * cols + j] = vs;
    a.pg[t * cols + j] = pg;
    v_next = val;
    vs_next = vs;
  }
}

// K4 with the unroll split over the threads of a block: for a batch of a few
// ten thousand columns one thread per column leaves an SM with a couple of
// hundred threads and as many loads in flight - a third of the memory
// bandwidth.  Here 32 adjacent columns share a block of kIwChunks x 32
// threads; thread (c, lane) owns chunk c of its column's time steps.  The
// recurrence acc_t = gain_t acc_{t+1} + delta_t (vtrace.py:250-262) is affine,
// so a chunk first walks its steps from acc = 0 and publishes (prod gain, acc)
// in shared memory, folds the aggregates of the later-time chunks in index
// order into the value that enters it, and then walks its steps again from
// that value - inside a chunk the operations are those of the sequential
// recurrence - writing vs and pg_advantages.  The second walk re-reads what
// the first one brought into L1/L2.
constexpr int kIwChunks = 8;

struct IwStep {
  float gain, delta, rho_pg, rew, disc, val;
};

__device__ __forceinline__ IwStep iw_load(const IwArgs& a, long cols, long j, long rho_col, long rho_stride,
                                          long disc_col, long disc_stride, int t, float v_next) {
  const float log_rho = __ldg(a.log_rhos + t * rho_stride + rho_col);
  IwStep s;
  s.disc = __ldg(a.discounts + t * disc_stride + disc_col);
  s.rew = __ldg(a.rewards + t * cols + j);
  s.val = __ldg(a.values + t * cols + j);
  const float rho = expf(log_rho);                                       // :235
  const float rho_hat = a.no_clip_rho ? rho : fminf(a.clip_rho, rho);    // :236-239
  const float c = fminf(1.0f, rho);                                      // :241
  s.rho_pg = a.no_clip_pg ? rho : fminf(a.clip_pg, rho);                 // :270-274
  s.delta = rho_hat * (s.rew + s.disc * v_next - s.val);                 // :245
  s.gain = s.disc * c;
  return s;
}

__global__ void __launch_bounds__(32 * kIwChunks)
    importance_weights_chunk_kernel(const __grid_constant__ IwArgs a) {
  __shared__ float sh_prod[kIwChunks][32], sh_acc[kIwChunks][32];
  const long cols = static_cast<long>(a.B) * a.C;
  const int lane = threadIdx.x & 31, c = threadIdx.x >> 5;
  const long j = static_cast<long>(blockIdx.x) * 32 + lane;
  const bool live = j < cols;
  const long jj = live ? j : 0;
  const long jb = jj / a.C;
  const long rho_col = a.rho_bcast ? jb : jj, rho_stride = a.rho_bcast ? a.B : cols;
  const long disc_col = a.disc_bcast ? jb : jj, disc_stride = a.disc_bcast ? a.B : cols;
  const int S = (a.T + kIwChunks - 1) / kIwChunks;
  const int it0 = c * S, it1 = min(a.T, it0 + S);  // steps counted from the end: t = T-1-it

  // first walk: the chunk's aggregate
  float acc = 0.0f, prod = 1.0f;
  if (live) {
    float v_next = it0 == 0 ? a.bootstrap[jj] : __ldg(a.values + (a.T - it0) * cols + jj);
#pragma unroll 4
    for (int it = it0; it < it1; ++it) {
      const IwStep s = iw_load(a, cols, jj, rho_col, rho_stride, disc_col, disc_stride, a.T - 1 - it, v_next);
      acc = s.delta + s.gain * acc;  // :250-253
      prod *= s.gain;
      v_next = s.val;
    }
  }
  sh_prod[c][lane] = prod, sh_acc[c][lane] = acc;
  __syncthreads();
  if (!live || it0 >= it1) return;
  float acc_in = 0.0f;
  for (int k = 0; k < c; ++k) acc_in = fmaf(sh_prod[k][lane], acc_in, sh_acc[k][lane]);

  // second walk: the sequential recurrence from the value that enters the chunk
  acc = acc_in;
  float v_next = it0 == 0 ? a.bootstrap[jj] : __ldg(a.values + (a.T - it0) * cols + jj);
  float vs_next = it0 == 0 ? v_next : acc_in + v_next;  // :268-269
#pragma unroll 4
  for (int it = it0; it < it1; ++it) {
    const int t = a.T - 1 - it;
    const IwStep s = iw_load(a, cols, jj, rho_col, rho_stride, disc_col, disc_stride, t, v_next);
    acc = s.delta + s.gain * acc;
    const float vs = acc + s.val;                                    // :265
    a.vs[t * cols + jj] = vs;
    a.pg[t * cols + jj] = s.rho_pg * (s.rew + s.disc * vs_next - s.val);  // :275-276
    v_next = s.val;
    vs_next = vs;
  }
}

// ---------------------------------------------------------------------------
// Row kernels: one block stages R rows of logits ([R, A], contiguous in
// global memory) into shared memory with coalesced loads, then one thread per
// row does the softmax arithmetic out of shared memory (row stride padded to
// an odd word count: conflict-free).  Backward variants overwrite the staged
// rows with the gradient and store them back coalesced.
// ---------------------------------------------------------------------------
enum RowMode { kLogProbs = 0, kPgFwd = 1, kPgBwd = 2, kEntFwd = 3, kEntBwd = 4 };

struct RowArgs {
  const float* logits;
  const int32_t* actions;
  const float* advantages;
  const float* grad_out;  // device scalar
  float* out;             // [rows] (log-probs) or [rows, A] (gradients)
  double* partials;       // one per block (forward losses)
  long rows;
  int A;
  int stride;  // padded shared-memory row stride in floats
};

__device__ __forceinline__ double block_sum_fixed(double v, double* sh) {
  const int tid = threadIdx.x;
  sh[tid] = v;
  __syncthreads();
  for (int off = blockDim.x >> 1; off > 0; off >>= 1) {
    if (tid < off) sh[tid] += sh[tid + off];
    __syncthreads();
  }
  return sh[0];
}

template <int MODE>
__global__ void __launch_bounds__(128) row_kernel(const __grid_constant__ RowArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* tile = reinterpret_cast<float*>(smem_raw);
  __shared__ double sh_red[128];
  const int R = blockDim.x, tid = threadIdx.x, A = a.A, stride = a.stride;
  const long row0 = static_cast<long>(blockIdx.x) * R;
  const int nrows = static_cast<int>(min(static_cast<long>(R), a.rows - row0));
  const int nelem = nrows * A;
  const float* src = a.logits + row0 * A;

  // coalesced load; (r, c) tracks element i = r*A + c without a division
  {
    int r = tid / A, c = tid - r * A;
    const int dr = R / A, dc = R - dr * A;
    for (int i = tid; i < nelem; i += R) {
      tile[r * stride + c] = __ldg(src + i);
      r += dr, c += dc;
      if (c >= A) c -= A, ++r;
    }
  }
  __syncthreads();

  double partial = 0.0;
  if (tid < nrows) {
    float* x = tile + tid * stride;
    float m = x[0];
    for (int i = 1; i < A; ++i) m = fmaxf(m, x[i]);
    float sum = 0.0f;
    for (int i = 0; i < A; ++i) sum += ptx::ex2((x[i] - m) * kLog2e);
    const float ln_sum = logf(sum);
    const long row = row0 + tid;
    if (MODE == kLogProbs || MODE == kPgFwd || MODE == kPgBwd) {
      int act = a.actions[row];
      act = min(max(act, 0), A - 1);
      const float lp = (x[act] - m) - ln_sum;
      if (MODE == kLogProbs) a.out[row] = lp;
      if (MODE == kPgFwd)  // cross_entropy * advantages, experiment.py:339-342
        partial = -static_cast<double>(lp) * static_cast<double>(a.advantages[row]);
      if (MODE == kPgBwd) {
        const float g = __ldg(a.grad_out) * a.advantages[row];
        const float inv = 1.0f / sum;
        for (int i = 0; i < A; ++i) x[i] = g * (ptx::ex2((x[i] - m) * kLog2e) * inv);
        x[act] -= g;
      }
    } else {
      const float inv = 1.0f / sum;
      float h = 0.0f;  // entropy H = -sum p log p
      for (int i = 0; i < A; ++i) {
        const float d = x[i] - m;
        h = fmaf(-(ptx::ex2(d * kLog2e) * inv), d - ln_sum, h);
      }
      if (MODE == kEntFwd) partial = -static_cast<double>(h);  // experiment.py:334-335
      if (MODE == kEntBwd) {
        const float g = __ldg(a.grad_out);
        for (int i = 0; i < A; ++i) {
          const float d = x[i] - m;
          const float lp = d - ln_sum;
          x[i] = g * (ptx::ex2(d * kLog2e) * inv) * (lp + h);
        }
      }
    }
  }

  if (MODE == kPgFwd || MODE == kEntFwd) {
    const double total = block_sum_fixed(partial, sh_red);
    if (tid == 0) a.partials[blockIdx.x] = total;
  }
  if (MODE == kPgBwd || MODE == kEntBwd) {
    __syncthreads();
    float* dst = a.out + row0 * A;
    int r = tid / A, c = tid - r * A;
    const int dr = R / A, dc = R - dr * A;
    for (int i = tid; i < nelem; i += R) {
      dst[i] = tile[r * stride + c];
      r += dr, c += dc;
      if (c >= A) c -= A, ++r;
    }
  }
}

// The same five row operations for the action counts the fused kernels are
// specialised for, with the row arithmetic of vtrace_math.cuh.  A tile of 128
// rows is 128*A contiguous floats: it arrives with ONE TMA bulk copy
// (cp.async.bulk -> mbarrier), every thread pulls its row into registers with
// vector loads (row stride A words: conflict-free for the widths used), and
// the backward variants write the gradient rows back into the tile, which
// leaves with one bulk store.  Per row that is a third of the instructions of
// the generic kernel above, which stays for every other A, for tensors that
// are not 16-byte aligned and for a ragged last tile.
template <int MODE, int A_CT>
__global__ void __launch_bounds__(128) row_tile_kernel(const __grid_constant__ RowArgs a) {
  __shared__ __align__(128) float tile[128 * A_CT];
  __shared__ __align__(8) unsigned long long sh_bar;
  __shared__ double sh_red[128];
  const int tid = threadIdx.x;
  const long row0 = static_cast<long>(blockIdx.x) * 128;
  const int nrows = static_cast<int>(min(128L, a.rows - row0));
  const int nelem = nrows * A_CT;
  const uint32_t bytes = static_cast<uint32_t>(nelem) * 4u;
  const float* src = a.logits + row0 * A_CT;
  const bool bulk = bytes % 16u == 0;  // a ragged last tile may not be
  const uint32_t bar = ptx::smem_u32(&sh_bar);
  if (bulk) {
    if (tid == 0) {
      ptx::mbar_init(bar, 1);
      ptx::fence_mbar_init();
      ptx::fence_proxy_async();
      ptx::mbar_arrive_expect_tx(bar, bytes);
      ptx::bulk_g2s(ptx::smem_u32(tile), src, bytes, bar, ptx::policy_evict_first());
    }
  } else {
    for (int i = tid; i < nelem; i += 128) tile[i] = __ldg(src + i);
  }
  // the per-row operands travel while the tile does
  const long row = row0 + tid;
  const bool live = tid < nrows;
  int act = 0;
  float adv = 0.0f, g = 0.0f;
  if (live) {
    if (MODE == kLogProbs || MODE == kPgFwd || MODE == kPgBwd) act = min(max(__ldg(a.actions + row), 0), A_CT - 1);
    if (MODE == kPgFwd || MODE == kPgBwd) adv = __ldg(a.advantages + row);
    if (MODE == kPgBwd || MODE == kEntBwd) g = __ldg(a.grad_out);
  }
  __syncthreads();  // the barrier is initialised / the plain loads have landed
  if (bulk) ptx::mbar_wait(bar, 0);

  double partial = 0.0;
  if (live) {
    float* mine = tile + tid * A_CT;
    float x[A_CT];
    load_row(x, mine);
    if (MODE == kLogProbs || MODE == kPgFwd) {
      const math::RowStats st = math::row_stats(x);
      const float lp = math::action_log_prob(mine[act], st);  // vtrace.py:67
      if (MODE == kLogProbs) a.out[row] = lp;
      // cross_entropy * advantages, experiment.py:339-342
      if (MODE == kPgFwd) partial = -static_cast<double>(lp) * static_cast<double>(adv);
    } else {
      float e[A_CT];
      const math::RowStats st = math::row_stats(x, e);
      const float inv = __frcp_rn(st.sum);
      if (MODE == kPgBwd) {  // g adv (softmax - onehot)
        const float ga = __fmul_rn(g, adv);
        const float k = __fmul_rn(ga, inv);
#pragma unroll
        for (int i = 0; i < A_CT; ++i) e[i] = __fmul_rn(k, e[i]);
        store_row(mine, e);
        mine[act] = __fsub_rn(mine[act], ga);
      } else {
        const float h = math::row_entropy(x, e, st, inv);          // x <- log-softmax
        if (MODE == kEntFwd) partial = -static_cast<double>(h);  // experiment.py:334-335
        if (MODE == kEntBwd) {                                     // g p (log p + H)
          const float k = __fmul_rn(g, inv);
#pragma unroll
          for (int i = 0; i < A_CT; ++i) e[i] = __fmul_rn(__fmul_rn(k, e[i]), __fadd_rn(x[i], h));
          store_row(mine, e);
        }
      }
    }
  }

  if (MODE == kPgFwd || MODE == kEntFwd) {
    const double total = block_sum_fixed(partial, sh_red);
    if (tid == 0) a.partials[blockIdx.x] = total;
  }
  if (MODE == kPgBwd || MODE == kEntBwd) {
    float* dst = a.out + row0 * A_CT;
    if (bulk) {
      ptx::fence_proxy_async();
      __syncthreads();
      if (tid == 0) {
        ptx::bulk_s2g(dst, ptx::smem_u32(tile), bytes);
        ptx::bulk_commit();
        ptx::bulk_wait_read<0>();  // the tile is read before the block goes away
      }
    } else {
      __syncthreads();
      for (int i = tid; i < nelem; i += 128) dst[i] = tile[i];
    }
  }
}

// Second stage for the single-scalar losses: out = scale * sum(partials).
__global__ void __launch_bounds__(256) reduce_scalar_kernel(const double* __restrict__ partials,
                                                            long n, double scale,
                                                            float* __restrict__ out) {
  __shared__ double sh[256];
  double s = 0.0;
  for (long i = threadIdx.x; i < n; i += 256) s += partials[i];
  const double total = block_sum_fixed(s, sh);
  if (threadIdx.x == 0) out[0] = static_cast<float>(scale * total);
}

constexpr int kSquareBlock = 256;
constexpr int kSquarePerThread = 8;

// Baseline loss first stage: each block sums a fixed contiguous chunk.
__global__ void __launch_bounds__(kSquareBlock) sum_squares_kernel(const float* __restrict__ x,
                                                                  long n,
                                                                  double* __restrict__ partials) {
  __shared__ double sh[kSquareBlock];
  const long base = static_cast<long>(blockIdx.x) * kSquareBlock * kSquarePerThread;
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < kSquarePerThread; ++k) {
    const long i = base + k * kSquareBlock + threadIdx.x;
    if (i < n) {
      const double v = static_cast<double>(__ldg(x + i));
      s = fma(v, v, s);
    }
  }
  const double total = block_sum_fixed(s, sh);
  if (threadIdx.x == 0) partials[blockIdx.x] = total;
}

__global__ void __launch_bounds__(256) scale_kernel(const float* __restrict__ x,
                                                    const float* __restrict__ grad_out, long n,
                                                    float* __restrict__ y) {
  const long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) y[i] = __ldg(grad_out) * x[i];
}

// In-place x *= *scale on a fixed-size grid; a whole-grid no-op when *scale == 1.
__global__ void __launch_bounds__(256) scale_inplace_kernel(float* __restrict__ x,
                                                            const float* __restrict__ scale,
                                                            long n) {
  const float g = __ldg(scale);
  if (g == 1.0f) return;
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  for (long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    x[i] *= g;
}

static int rows_per_block(int A) { return A <= 96 ? 128 : 64; }

static long row_blocks(long rows, int A) {
  const int R = rows_per_block(A);
  return (rows + R - 1) / R;
}

template <int MODE, int A_CT>
static int launch_row_tiles(const RowArgs& a, cudaStream_t stream) {
  const long blocks = (a.rows + 127) / 128;
  if (blocks > 0x7fffffffL) return fail(VTR_ERR_UNSUPPORTED, "too many rows (%ld)", a.rows);
  row_tile_kernel<MODE, A_CT><<<static_cast<unsigned>(blocks), 128, 0, stream>>>(a);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

template <int MODE>
static int launch_rows(RowArgs a, cudaStream_t stream) {
  // compiled-in action counts, 16-byte aligned tensors: the bulk-copy kernel
  const bool writes_rows = MODE == kPgBwd || MODE == kEntBwd;
  if (aligned(a.logits, 16) && (!writes_rows || aligned(a.out, 16)) && !getenv("VTR_ROWS_GENERIC")) {
    switch (a.A) {
      case 3: return launch_row_tiles<MODE, 3>(a, stream);
      case 4: return launch_row_tiles<MODE, 4>(a, stream);
      case 6: return launch_row_tiles<MODE, 6>(a, stream);
      case 9: return launch_row_tiles<MODE, 9>(a, stream);
      case 18: return launch_row_tiles<MODE, 18>(a, stream);
      default: break;
    }
  }
  const int R = rows_per_block(a.A);
  a.stride = a.A | 1;
  const size_t smem = static_cast<size_t>(R) * a.stride * sizeof(float);
  auto kernel = row_kernel<MODE>;
  if (int rc = allow_dynamic_smem(reinterpret_cast<const void*>(kernel), smem)) return rc;
  const long blocks = row_blocks(a.rows, a.A);
  if (blocks > 0x7fffffffL) return fail(VTR_ERR_UNSUPPORTED, "too many rows (%ld)", a.rows);
  kernel<<<static_cast<unsigned>(blocks), R, smem, stream>>>(a);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

static int check_rows(const char* who, const void* logits, long rows, int A) {
  if (!logits) return fail(VTR_ERR_INVALID_ARGUMENT, "%s: logits is NULL", who);
  if (rows <= 0 || A <= 0) return fail(VTR_ERR_INVALID_ARGUMENT, "%s: rows and A must be positive", who);
  if (A > VTR_MAX_ACTIONS) return fail(VTR_ERR_UNSUPPORTED, "%s: A=%d exceeds VTR_MAX_ACTIONS=%d", who, A, VTR_MAX_ACTIONS);
  if (!aligned(logits, 4)) return fail(VTR_ERR_MISALIGNED, "%s: logits not 4-byte aligned", who);
  return VTR_OK;
}

static int check_workspace(const char* who, long units, void* workspace, size_t bytes) {
  const size_t need = vtr_loss_workspace_bytes(units);
  if (!workspace || bytes < need)
    return fail(VTR_ERR_WORKSPACE, "%s: workspace of %zu bytes needed, %zu given", who, need, bytes);
  if (!aligned(workspace, 8)) return fail(VTR_ERR_MISALIGNED, "%s: workspace not 8-byte aligned", who);
  return VTR_OK;
}

}  // namespace vtr

using vtr::fail;

extern "C" int vtr_from_importance_weights_f32(const float* log_rhos, const float* discounts,
                                               const float* rewards, const float* values,
                                               const float* bootstrap_value, int32_t T, int32_t B,
                                               int32_t C, uint32_t bcast_mask,
                                               float clip_rho_threshold,
                                               float clip_pg_rho_threshold, uint32_t flags,
                                               float* vs, float* pg_advantages, void* stream) {
  if (!log_rhos || !discounts || !rewards || !values || !bootstrap_value || !vs || !pg_advantages)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_from_importance_weights_f32: NULL pointer");
  if (T <= 0 || B <= 0 || C <= 0)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_from_importance_weights_f32: T, B, C must be positive");
  vtr::IwArgs a;
  a.log_rhos = log_rhos, a.discounts = discounts, a.rewards = rewards, a.values = values;
  a.bootstrap = bootstrap_value, a.vs = vs, a.pg = pg_advantages;
  a.T = T, a.B = B, a.C = C;
  a.rho_bcast = (bcast_mask & 1u) != 0 && C > 1;
  a.disc_bcast = (bcast_mask & 2u) != 0 && C > 1;
  a.no_clip_rho = (flags & VTR_FLAG_NO_CLIP_RHO) != 0;
  a.no_clip_pg = (flags & VTR_FLAG_NO_CLIP_PG_RHO) != 0;
  a.clip_rho = clip_rho_threshold, a.clip_pg = clip_pg_rho_threshold;
  const long cols = static_cast<long>(B) * C;
  // Up to a few hundred thousand columns the unroll is split over the threads
  // of a block (one thread per column would leave the SMs short of loads in
  // flight); beyond, one thread per column fills the GPU by itself.
  vtr::DeviceInfo dev;
  if (int rc = vtr::device_info(&dev)) return rc;
  if (T >= 2 * vtr::kIwChunks && cols <= 2048L * dev.sm_count && !getenv("VTR_IW_PLAIN")) {
    const long blocks = (cols + 31) / 32;
    vtr::importance_weights_chunk_kernel<<<static_cast<unsigned>(blocks), 32 * vtr::kIwChunks, 0,
                                           static_cast<cudaStream_t>(stream)>>>(a);
    VTR_CUDA_CHECK(cudaGetLastError());
    return VTR_OK;
  }
  const long blocks = (cols + 127) / 128;
  if (blocks > 0x7fffffffL) return fail(VTR_ERR_UNSUPPORTED, "too many columns");
  vtr::importance_weights_kernel<<<static_cast<unsigned>(blocks), 128, 0,
                                   static_cast<cudaStream_t>(stream)>>>(a);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

extern "C" int vtr_log_probs_from_logits_and_actions_f32(const float* policy_logits,
                                                         const int32_t* actions, int64_t rows,
                                                         int32_t A, float* log_probs,
                                                         void* stream) {
  if (int rc = vtr::check_rows("vtr_log_probs", policy_logits, rows, A)) return rc;
  if (!actions || !log_probs) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_log_probs: NULL pointer");
  vtr::RowArgs a = {};
  a.logits = policy_logits, a.actions = actions, a.out = log_probs, a.rows = rows, a.A = A;
  return vtr::launch_rows<vtr::kLogProbs>(a, static_cast<cudaStream_t>(stream));
}

extern "C" size_t vtr_loss_workspace_bytes(int64_t rows) {
  if (rows <= 0) return 0;
  return static_cast<size_t>((rows + 63) / 64) * sizeof(double);
}

extern "C" int vtr_policy_gradient_loss_f32(const float* logits, const int32_t* actions,
                                            const float* advantages, int64_t rows, int32_t A,
                                            float* loss, void* workspace, size_t workspace_bytes,
                                            void* stream) {
  if (int rc = vtr::check_rows("vtr_policy_gradient_loss", logits, rows, A)) return rc;
  if (!actions || !advantages || !loss)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_policy_gradient_loss: NULL pointer");
  if (int rc = vtr::check_workspace("vtr_policy_gradient_loss", rows, workspace, workspace_bytes)) return rc;
  vtr::RowArgs a = {};
  a.logits = logits, a.actions = actions, a.advantages = advantages, a.rows = rows, a.A = A;
  a.partials = static_cast<double*>(workspace);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (int rc = vtr::launch_rows<vtr::kPgFwd>(a, s)) return rc;
  vtr::reduce_scalar_kernel<<<1, 256, 0, s>>>(a.partials, vtr::row_blocks(rows, A), 1.0, loss);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

extern "C" int vtr_policy_gradient_loss_bwd_f32(const float* logits, const int32_t* actions,
                                                const float* advantages, const float* grad_out,
                                                int64_t rows, int32_t A, float* d_logits,
                                                void* stream) {
  if (int rc = vtr::check_rows("vtr_policy_gradient_loss_bwd", logits, rows, A)) return rc;
  if (!actions || !advantages || !grad_out || !d_logits)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_policy_gradient_loss_bwd: NULL pointer");
  vtr::RowArgs a = {};
  a.logits = logits, a.actions = actions, a.advantages = advantages, a.grad_out = grad_out;
  a.out = d_logits, a.rows = rows, a.A = A;
  return vtr::launch_rows<vtr::kPgBwd>(a, static_cast<cudaStream_t>(stream));
}

extern "C" int vtr_entropy_loss_f32(const float* logits, int64_t rows, int32_t A, float* loss,
                                    void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = vtr::check_rows("vtr_entropy_loss", logits, rows, A)) return rc;
  if (!loss) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_entropy_loss: NULL pointer");
  if (int rc = vtr::check_workspace("vtr_entropy_loss", rows, workspace, workspace_bytes)) return rc;
  vtr::RowArgs a = {};
  a.logits = logits, a.rows = rows, a.A = A;
  a.partials = static_cast<double*>(workspace);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (int rc = vtr::launch_rows<vtr::kEntFwd>(a, s)) return rc;
  vtr::reduce_scalar_kernel<<<1, 256, 0, s>>>(a.partials, vtr::row_blocks(rows, A), 1.0, loss);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

extern "C" int vtr_entropy_loss_bwd_f32(const float* logits, const float* grad_out, int64_t rows,
                                        int32_t A, float* d_logits, void* stream) {
  if (int rc = vtr::check_rows("vtr_entropy_loss_bwd", logits, rows, A)) return rc;
  if (!grad_out || !d_logits) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_entropy_loss_bwd: NULL pointer");
  vtr::RowArgs a = {};
  a.logits = logits, a.grad_out = grad_out, a.out = d_logits, a.rows = rows, a.A = A;
  return vtr::launch_rows<vtr::kEntBwd>(a, static_cast<cudaStream_t>(stream));
}

extern "C" int vtr_baseline_loss_f32(const float* advantages, int64_t n, float* loss,
                                     void* workspace, size_t workspace_bytes, void* stream) {
  if (!advantages || !loss || n <= 0)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_baseline_loss: NULL pointer or n <= 0");
  if (int rc = vtr::check_workspace("vtr_baseline_loss", n, workspace, workspace_bytes)) return rc;
  const long per_block = vtr::kSquareBlock * vtr::kSquarePerThread;
  const long blocks = (n + per_block - 1) / per_block;
  if (blocks > 0x7fffffffL) return fail(VTR_ERR_UNSUPPORTED, "vtr_baseline_loss: n too large");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  double* partials = static_cast<double*>(workspace);
  vtr::sum_squares_kernel<<<static_cast<unsigned>(blocks), vtr::kSquareBlock, 0, s>>>(advantages, n, partials);
  VTR_CUDA_CHECK(cudaGetLastError());
  vtr::reduce_scalar_kernel<<<1, 256, 0, s>>>(partials, blocks, 0.5, loss);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

extern "C" int vtr_baseline_loss_bwd_f32(const float* advantages, const float* grad_out, int64_t n,
                                         float* d_advantages, void* stream) {
  if (!advantages || !grad_out || !d_advantages || n <= 0)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_baseline_loss_bwd: NULL pointer or n <= 0");
  const long blocks = (n + 255) / 256;
  if (blocks > 0x7fffffffL) return fail(VTR_ERR_UNSUPPORTED, "vtr_baseline_loss_bwd: n too large");
  vtr::scale_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      advantages, grad_out, n, d_advantages);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

extern "C" int vtr_scale_inplace_f32(float* x, const float* scale, int64_t n, void* stream) {
  if (n < 0) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_scale_inplace: n < 0");
  if (n == 0) return VTR_OK;  // nothing to scale (e.g. a gradient nobody asked for)
  if (!x || !scale) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_scale_inplace: NULL pointer");
  vtr::DeviceInfo dev;
  if (int rc = vtr::device_info(&dev)) return rc;
  long blocks = (n + 255) / 256;
  const long cap = static_cast<long>(dev.sm_count) * 8;
  if (blocks > cap) blocks = cap;
  vtr::scale_inplace_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, scale, n);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}
