// The optimiser step that follows the path:
//   experiment.py:410-412  tf.train.polynomial_decay(lr, num_env_frames, total_frames, 0)
//   experiment.py:413-415  tf.train.RMSPropOptimizer(lr, decay, momentum, epsilon).minimize(loss)
//
// ApplyRMSProp, as TensorFlow 1.9.0 defines it:
//   ms  <- ms + (1 - rho) (g^2 - ms)
//   mom <- momentum mom + lr g / sqrt(ms + epsilon)
//   var <- var - mom
//
// Two forms.  vtr_rmsprop_f32 is the update on one GPU: 16 bytes read and 12
// written per parameter (12 and 8 without the momentum slot), HBM-bound.
// vtr_rmsprop_allreduce_f32 is the data-parallel learner's step in one launch:
// with the batch sharded over the GPUs of a box the gradient the optimiser
// needs is the sum of the ranks' gradients, so each rank
//   0. tells its peers that its gradient buffer is final and waits for theirs,
//   1. sums, for ITS slice of the parameters, every rank's gradient straight
//      out of the peers' memory (NVLink loads, rank order: the same bits on
//      every run), applies the update with its slice of the slots, and stores
//      the new values into every rank's parameter buffer (NVLink stores),
//   2. tells its peers that its slice is in place and waits for theirs.
// That is a reduce-scatter, the update and an all-gather without the gradient
// sum or the updated slice ever going through HBM a second time, and the slots
// exist once per box instead of once per GPU.
#include <math.h>

#include <atomic>

#include "common.h"
#include "exchange.cuh"

namespace vtr {

struct RmsHyper {
  float lr, one_minus_decay, momentum, epsilon;
};

__device__ __forceinline__ void rms_update(float& var, float& ms, float& mom, float g, const RmsHyper& h) {
  ms = __fmaf_rn(h.one_minus_decay, __fsub_rn(__fmul_rn(g, g), ms), ms);
  const float step = __fdiv_rn(__fmul_rn(h.lr, g), __fsqrt_rn(__fadd_rn(ms, h.epsilon)));
  mom = __fmaf_rn(h.momentum, mom, step);
  var = __fsub_rn(var, mom);
}

__device__ __forceinline__ void rms_update4(float4& v, float4& m, float4& mo, const float4& g, const RmsHyper& h) {
  rms_update(v.x, m.x, mo.x, g.x, h);
  rms_update(v.y, m.y, mo.y, g.y, h);
  rms_update(v.z, m.z, mo.z, g.z, h);
  rms_update(v.w, m.w, mo.w, g.w, h);
}

// ---------------------------------------------------------------------------
// One GPU.  Whole float4 groups grid-stride; the last n % 4 elements by the
// first threads of block 0.
// ---------------------------------------------------------------------------
template <bool HAS_MOM>
__global__ void __launch_bounds__(256) rmsprop_kernel(float* __restrict__ var, float* __restrict__ ms,
                                                      float* __restrict__ mom,
                                                      const float* __restrict__ grad, long n, RmsHyper h) {
  const long groups = n >> 2;
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  float4* var4 = reinterpret_cast<float4*>(var);
  float4* ms4 = reinterpret_cast<float4*>(ms);
  float4* mom4 = reinterpret_cast<float4*>(mom);
  const float4* grad4 = reinterpret_cast<const float4*>(grad);
  for (long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x; i < groups; i += stride) {
    float4 v = var4[i], m = ms4[i];
    const float4 g = __ldcs(grad4 + i);  // read once, never again
    float4 mo = HAS_MOM ? mom4[i] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    rms_update4(v, m, mo, g, h);
    var4[i] = v, ms4[i] = m;
    if (HAS_MOM) mom4[i] = mo;
  }
  const long i = (groups << 2) + threadIdx.x;
  if (blockIdx.x == 0 && i < n) {
    float v = var[i], m = ms[i], mo = HAS_MOM ? mom[i] : 0.0f;
    rms_update(v, m, mo, grad[i], h);
    var[i] = v, ms[i] = m;
    if (HAS_MOM) mom[i] = mo;
  }
}

// ---------------------------------------------------------------------------
// All GPUs of a box, one launch per rank.  Sync words of one rank (uint32,
// zeroed once): [r] = rank r's gradient is final for step seq; [16 + r] =
// rank r's slice of the parameters is in place for step seq; [32] = blocks of
// this launch that are done (local ticket); [33] = sticky status, non-zero
// once a step failed; [34] = this launch's go / no-go, decided once by block 0
// and read by every block.  A flag word holds seq, or seq | kSyncFailed when
// its writer gave the step up (a peer did not show up in time): whoever waits
// on it then gives up too, at once, instead of using half-finished data.
// ---------------------------------------------------------------------------
constexpr int kSyncGradFlags = 0;
constexpr int kSyncDoneFlags = 16;
constexpr int kSyncTicket = 32;
constexpr int kSyncStatus = 33;
constexpr int kSyncGo = 34;
constexpr uint32_t kSyncFailed = 0x80000000u;  // seq stays below 2^31

struct RmsShardArgs {
  PeerBuffers var;   // every rank's full parameter buffer, as mapped here
  PeerBuffers grad;  // every rank's full gradient buffer
  PeerBuffers sync;  // every rank's sync words
  float* ms;         // this rank's slice of the slots, [hi - lo]
  float* mom;        // or null when momentum == 0
  long n, lo, hi;    // parameters; this rank's slice [lo, hi), lo % 4 == 0
  int rank, world;
  uint32_t seq;
  RmsHyper h;
  unsigned long long timeout_ns;
  uint32_t* host_status;  // pinned host word the kernel sets to seq on failure, or null
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Spins until *flag carries step `want`: true if its writer went ahead, false
// if it gave the step up or nothing came within timeout_ns (a peer is missing).
__device__ __forceinline__ bool wait_flag(const uint32_t* flag, uint32_t want, unsigned long long timeout_ns) {
  unsigned long long t0, now;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
    if (v == want) return true;
    if (v == (want | kSyncFailed)) return false;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
    if (now - t0 > timeout_ns) return false;
    __nanosleep(64);
  }
}

// This rank gives step `seq` up: sticky status here, the host word if there is
// one, and the failure marker in both flags of every peer, so that a peer
// that is waiting for this rank - or turns up later - gives up as well.
__device__ __forceinline__ void give_up(const RmsShardArgs& a, uint32_t* my_sync, int tid) {
  if (tid < a.world) {
    uint32_t* peer = reinterpret_cast<uint32_t*>(a.sync.p[tid]);
    st_release_sys(peer + kSyncGradFlags + a.rank, a.seq | kSyncFailed);
    st_release_sys(peer + kSyncDoneFlags + a.rank, a.seq | kSyncFailed);
  }
  if (tid == 0) {
    my_sync[kSyncStatus] = 1u;
    if (a.host_status) {
      *reinterpret_cast<volatile uint32_t*>(a.host_status) = a.seq;
      __threadfence_system();
    }
  }
}

__device__ __forceinline__ float4 add4(const float4& a, const float4& b) {
  return make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), __fadd_rn(a.w, b.w));
}

// U float4 groups per thread and iteration, their loads from 8 / U ranks at a
// time: always eight 16-byte loads in flight per thread.  NVLink loads take a
// few microseconds to come back, so with one or two peers a thread has to
// cover several groups at once to keep the links busy.
template <bool HAS_MOM, int U>
__global__ void __launch_bounds__(256) rmsprop_allreduce_kernel(const __grid_constant__ RmsShardArgs a) {
  constexpr int RC = 8 / U;
  __shared__ int sh_ok, sh_last;
  uint32_t* my_sync = reinterpret_cast<uint32_t*>(a.sync.p[a.rank]);
  const int tid = threadIdx.x;

  // ---- 0: gradients are final everywhere ----
  // This rank's gradient was written by earlier work on this stream, so it is
  // complete when this kernel starts; block 0 says so to every peer, waits for
  // theirs and decides for the whole grid: every block takes the same branch,
  // so a slice is updated completely or not at all.
  if (blockIdx.x == 0) {
    if (tid < a.world) {
      __threadfence_system();
      st_release_sys(reinterpret_cast<uint32_t*>(a.sync.p[tid]) + kSyncGradFlags + a.rank, a.seq);
    }
    if (tid == 0) sh_ok = 1;
    __syncthreads();
    if (tid < a.world && !wait_flag(my_sync + kSyncGradFlags + tid, a.seq, a.timeout_ns)) sh_ok = 0;
    __syncthreads();
    if (!sh_ok) give_up(a, my_sync, tid);
    if (tid == 0) st_release_sys(my_sync + kSyncGo, sh_ok ? a.seq : (a.seq | kSyncFailed));
  } else {
    // block 0 is resident before any other block starts; its wait is bounded
    if (tid == 0) sh_ok = wait_flag(my_sync + kSyncGo, a.seq, 4 * a.timeout_ns) ? 1 : 0;
    __syncthreads();
  }
  const bool ok = sh_ok != 0;

  // ---- 1: reduce-scatter + update + all-gather of this rank's slice ----
  if (ok) {
    const long groups = (a.hi - a.lo) >> 2;  // whole float4 groups of the slice
    const long stride = static_cast<long>(gridDim.x) * blockDim.x;
    float4* ms4 = reinterpret_cast<float4*>(a.ms);
    float4* mom4 = reinterpret_cast<float4*>(a.mom);
    for (long i0 = static_cast<long>(blockIdx.x) * blockDim.x + tid; i0 < groups; i0 += stride * U) {
      float4 g[U];
      for (int r0 = 0; r0 < a.world; r0 += RC) {
        float4 part[U][RC];
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int k = 0; k < RC; ++k)
            if (r0 + k < a.world && i0 + u * stride < groups)
              part[u][k] = __ldcg(reinterpret_cast<const float4*>(a.grad.p[r0 + k] + a.lo +
                                                                  ((i0 + u * stride) << 2)));
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int k = 0; k < RC; ++k)  // rank order
            if (r0 + k < a.world) g[u] = (r0 + k == 0) ? part[u][0] : add4(g[u], part[u][k]);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + u * stride;
        if (i >= groups) break;
        const long e = a.lo + (i << 2);
        float4 v = __ldcg(reinterpret_cast<const float4*>(a.var.p[a.rank] + e));
        float4 m = ms4[i];
        float4 mo = HAS_MOM ? mom4[i] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        rms_update4(v, m, mo, g[u], a.h);
        ms4[i] = m;
        if (HAS_MOM) mom4[i] = mo;
        for (int r = 0; r < a.world; ++r) *reinterpret_cast<float4*>(a.var.p[r] + e) = v;
      }
    }
    // the last, partial group of the whole buffer belongs to whoever owns it
    const long e = a.lo + (groups << 2) + tid;
    if (blockIdx.x == 0 && e < a.hi) {
      float g = 0.0f;
      for (int r = 0; r < a.world; ++r) g = r == 0 ? __ldcg(a.grad.p[0] + e) : __fadd_rn(g, __ldcg(a.grad.p[r] + e));
      const long i = e - a.lo;
      float v = __ldcg(a.var.p[a.rank] + e), m = a.ms[i], mo = HAS_MOM ? a.mom[i] : 0.0f;
      rms_update(v, m, mo, g, a.h);
      a.ms[i] = m;
      if (HAS_MOM) a.mom[i] = mo;
      for (int r = 0; r < a.world; ++r) a.var.p[r][e] = v;
    }
  }

  // ---- 2: every slice is in place everywhere ----
  __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    const unsigned int before = atomicAdd(my_sync + kSyncTicket, 1u);
    sh_last = before == gridDim.x - 1;
  }
  __syncthreads();
  if (!sh_last) return;
  if (tid == 0) my_sync[kSyncTicket] = 0u;  // ready for the next launch
  if (!ok) return;                          // given up in phase 0: the peers were told there
  __threadfence_system();  // the other blocks' stores are ordered before the flag below
  if (tid == 0) sh_ok = 1;
  __syncthreads();
  if (tid < a.world) {
    st_release_sys(reinterpret_cast<uint32_t*>(a.sync.p[tid]) + kSyncDoneFlags + a.rank, a.seq);
    // Until every peer has said the same, some of them may still be reading
    // this rank's gradient or writing into its parameters.  A peer that gave
    // the step up (or never came) leaves the parameters out of step: flagged.
    if (!wait_flag(my_sync + kSyncDoneFlags + tid, a.seq, a.timeout_ns)) sh_ok = 0;
  }
  __syncthreads();
  if (!sh_ok) give_up(a, my_sync, tid);
}

}  // namespace vtr

using vtr::fail;

extern "C" double vtr_polynomial_decay(double learning_rate, int64_t global_step, int64_t decay_steps,
                                       double end_learning_rate, double power) {
  if (decay_steps <= 0) return end_learning_rate;
  double step = static_cast<double>(global_step < 0 ? 0 : global_step);
  if (step > static_cast<double>(decay_steps)) step = static_cast<double>(decay_steps);
  const double frac = 1.0 - step / static_cast<double>(decay_steps);
  return (learning_rate - end_learning_rate) * pow(frac, power) + end_learning_rate;
}

static int check_hyper(const char* who, float decay, float momentum, float epsilon) {
  if (!(decay >= 0.0f && decay <= 1.0f)) return fail(VTR_ERR_INVALID_ARGUMENT, "%s: decay must be in [0, 1]", who);
  if (!(momentum >= 0.0f)) return fail(VTR_ERR_INVALID_ARGUMENT, "%s: momentum must be >= 0", who);
  if (!(epsilon >= 0.0f)) return fail(VTR_ERR_INVALID_ARGUMENT, "%s: epsilon must be >= 0", who);
  return VTR_OK;
}

static int blocks_for(long groups) {
  vtr::DeviceInfo dev;
  int sm = 148;
  if (vtr::device_info(&dev) == VTR_OK) sm = dev.sm_count;
  long blocks = (groups + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 8L * sm) blocks = 8L * sm;
  return static_cast<int>(blocks);
}

extern "C" int vtr_rmsprop_f32(float* var, float* ms, float* mom, const float* grad, int64_t n,
                               float learning_rate, float decay, float momentum, float epsilon,
                               void* stream) {
  if (n < 0) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_rmsprop: n < 0");
  if (n == 0) return VTR_OK;
  if (!var || !ms || !grad) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_rmsprop: NULL pointer");
  if (!mom && momentum != 0.0f)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_rmsprop: the momentum slot may only be NULL when momentum == 0");
  if (int rc = check_hyper("vtr_rmsprop", decay, momentum, epsilon)) return rc;
  if (!vtr::aligned(var, 16) || !vtr::aligned(ms, 16) || !vtr::aligned(mom, 16) || !vtr::aligned(grad, 16))
    return fail(VTR_ERR_MISALIGNED, "vtr_rmsprop: buffers must be 16-byte aligned");
  const vtr::RmsHyper h = {learning_rate, 1.0f - decay, momentum, epsilon};
  const int blocks = blocks_for(n >> 2);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (mom) vtr::rmsprop_kernel<true><<<blocks, 256, 0, s>>>(var, ms, mom, grad, n, h);
  else vtr::rmsprop_kernel<false><<<blocks, 256, 0, s>>>(var, ms, nullptr, grad, n, h);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

extern "C" int vtr_rmsprop_slice(int64_t n, int32_t world, int32_t rank, int64_t* lo, int64_t* hi) {
  if (n < 0 || world <= 0 || world > VTR_EXCHANGE_MAX_RANKS || rank < 0 || rank >= world || !lo || !hi)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_rmsprop_slice: bad arguments");
  const int64_t groups = (n + 3) / 4;
  const int64_t per = ((groups + world - 1) / world) * 4;
  *lo = rank * per < n ? rank * per : n;
  *hi = *lo + per < n ? *lo + per : n;
  return VTR_OK;
}

extern "C" size_t vtr_rmsprop_sync_bytes(void) { return 64 * sizeof(uint32_t); }

// How long a rank waits for its peers before it gives a step up.  Ranks of a
// learner drift apart by far more than a kernel's length now and then (an
// actor batch arriving late, a checkpoint being written), so the default is
// generous; it only has to keep a dead peer from hanging the GPU for good.
static std::atomic<unsigned long long> g_peer_timeout_ns{20000000000ull};  // 20 s

extern "C" void vtr_set_peer_timeout_ms(uint64_t milliseconds) {
  if (milliseconds == 0) milliseconds = 1;
  g_peer_timeout_ns.store(static_cast<unsigned long long>(milliseconds) * 1000000ull);
}
extern "C" uint64_t vtr_get_peer_timeout_ms(void) { return g_peer_timeout_ns.load() / 1000000ull; }

extern "C" int vtr_rmsprop_allreduce_f32(void* const* var_peers, void* const* grad_peers,
                                         void* const* sync_peers, float* ms_slice, float* mom_slice,
                                         int64_t n, int32_t rank, int32_t world, uint32_t seq,
                                         float learning_rate, float decay, float momentum,
                                         float epsilon, uint32_t* host_status, void* stream) {
  if (n <= 0) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_rmsprop_allreduce: n must be positive");
  if (world <= 0 || world > VTR_EXCHANGE_MAX_RANKS || rank < 0 || rank >= world)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_rmsprop_allreduce: bad rank/world %d/%d", rank, world);
  if (seq == 0 || seq >= vtr::kSyncFailed)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_rmsprop_allreduce: seq must be in [1, 2^31)");
  if (!vtr::aligned(host_status, 4)) return fail(VTR_ERR_MISALIGNED, "vtr_rmsprop_allreduce: host_status");
  if (int rc = check_hyper("vtr_rmsprop_allreduce", decay, momentum, epsilon)) return rc;
  vtr::RmsShardArgs a;
  if (int rc = vtr::fill_peer_buffers(&a.var, var_peers, world)) return rc;
  if (int rc = vtr::fill_peer_buffers(&a.grad, grad_peers, world)) return rc;
  if (int rc = vtr::fill_peer_buffers(&a.sync, sync_peers, world)) return rc;
  for (int r = 0; r < world; ++r)
    if (!vtr::aligned(a.var.p[r], 16) || !vtr::aligned(a.grad.p[r], 16) || !vtr::aligned(a.sync.p[r], 4))
      return fail(VTR_ERR_MISALIGNED, "vtr_rmsprop_allreduce: buffers must be 16-byte aligned");
  int64_t lo = 0, hi = 0;
  if (int rc = vtr_rmsprop_slice(n, world, rank, &lo, &hi)) return rc;
  if (hi > lo) {
    if (!ms_slice) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_rmsprop_allreduce: ms_slice is NULL");
    if (!mom_slice && momentum != 0.0f)
      return fail(VTR_ERR_INVALID_ARGUMENT,
                  "vtr_rmsprop_allreduce: the momentum slot may only be NULL when momentum == 0");
    if (!vtr::aligned(ms_slice, 16) || !vtr::aligned(mom_slice, 16))
      return fail(VTR_ERR_MISALIGNED, "vtr_rmsprop_allreduce: slots must be 16-byte aligned");
  }
  a.ms = ms_slice, a.mom = mom_slice;
  a.n = n, a.lo = lo, a.hi = hi;
  a.rank = rank, a.world = world, a.seq = seq;
  a.h = {learning_rate, 1.0f - decay, momentum, epsilon};
  a.timeout_ns = g_peer_timeout_ns.load();
  a.host_status = host_status;
  // Every rank launches, even one whose slice is empty: its peers wait for its flags.
  const int unroll = world <= 2 ? 4 : world <= 4 ? 2 : 1;
  const int blocks = blocks_for((((hi - lo) >> 2) + unroll - 1) / unroll);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool m = mom_slice != nullptr;
  if (unroll == 4) {
    if (m) vtr::rmsprop_allreduce_kernel<true, 4><<<blocks, 256, 0, s>>>(a);
    else vtr::rmsprop_allreduce_kernel<false, 4><<<blocks, 256, 0, s>>>(a);
  } else if (unroll == 2) {
    if (m) vtr::rmsprop_allreduce_kernel<true, 2><<<blocks, 256, 0, s>>>(a);
    else vtr::rmsprop_allreduce_kernel<false, 2><<<blocks, 256, 0, s>>>(a);
  } else {
    if (m) vtr::rmsprop_allreduce_kernel<true, 1><<<blocks, 256, 0, s>>>(a);
    else vtr::rmsprop_allreduce_kernel<false, 1><<<blocks, 256, 0, s>>>(a);
  }
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}
