// Fused V-trace + IMPALA loss, forward and backward, one pass over HBM:
// kernel selection, launch and the C entry points.
//
// Replaces vtrace.from_logits (vtrace.py:71-161), the three loss helpers
// (experiment.py:324-343), the weighted total (experiment.py:401-407) and the
// autodiff backward of that total (experiment.py:415).
//
// A lane owns a batch column and time is walked from t = T-1 down to 0, so the
// reverse recurrence acc_t = delta_t + gamma_t c_t acc_{t+1} (vtrace.py:250-262)
// and vs_{t+1} (vtrace.py:268-269) are carried, not recomputed.  For one time
// step the logits of adjacent columns are ONE contiguous run in the time-major
// [T,B,A] layout; rings of shared-memory slots are kept full by TMA bulk copies
// (cp.async.bulk -> mbarrier complete_tx), gradient rows leave with TMA bulk
// stores.  All kernels take their arithmetic from vtrace_math.cuh:
//   K1a fused_column_kernel   fused_column.cuh   32-column tile, one warp
//   K1b fused_relay_kernel    fused_relay.cuh    32-column tile, W warps relaying the carry
//   K1c fused_sm_tile_kernel  fused_sm_tile.cuh  one wide tile per SM, producer/consumer ring,
//                                                speed-weighted tiles, in-kernel final reduction
//   K2  fused_item_kernel     fused_item.cuh     one thread per (t, b) item, tiny batches;
//                                                warp-shuffle affine scan over T
//   K5  fused_chunk_kernel    fused_chunk.cuh    T split over a thread-block cluster, chunk
//                                                aggregates through distributed shared memory
//   K3  reduce_losses_kernel  (below)            second stage of the loss reduction, when it
//                                                is not done by the last CTA of K2/K5
// Losses accumulate per lane in double and are reduced in fixed orders; the
// only atomics are tickets (which tile next in K1c, who finished last).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "common.h"
#include "fused_chunk.cuh"
#include "fused_column.cuh"
#include "fused_common.cuh"
#include "fused_finish.cuh"
#include "fused_item.cuh"
#include "fused_relay.cuh"
#include "fused_sm_tile.cuh"
#include "ptx.cuh"
#include "vtrace_math.cuh"

namespace vtr {

// K3: second stage of the loss reduction.  One block, fixed summation order.
__global__ void __launch_bounds__(256) reduce_losses_kernel(const double* __restrict__ partials,
                                                            int n, float baseline_cost,
                                                            float entropy_cost,
                                                            float* __restrict__ losses,
                                                            double* __restrict__ losses_f64,
                                                            float* __restrict__ total_out) {
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
    if (total_out) *total_out = static_cast<float>(total);
  }
}

static inline int column_tiles(int B) { return (B + kColumnsPerTile - 1) / kColumnsPerTile; }

static bool specialised_actions(int A) { return A == 3 || A == 4 || A == 6 || A == 9 || A == 18; }

// vtr_fused_describe(): the launch functions below record what they would
// launch instead of launching it.
struct PlanRecorder {
  bool dry_run = false;
  int kernels = 0;
  char text[256] = "";
};
static thread_local PlanRecorder g_plan;

static void record_plan(const char* fmt, ...) {
  const size_t used = strlen(g_plan.text);
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_plan.text + used, sizeof(g_plan.text) - used, fmt, ap);
  va_end(ap);
  ++g_plan.kernels;
}

// Launch configuration of the fused kernels.  Every one of them is launched
// with programmatic stream serialization: its CTAs may be placed on SMs while
// the previous kernel of the stream is still draining and do their set-up
// there; they touch no tensor before griddepcontrol.wait, which returns when
// that kernel has completed and its writes are visible.  Stream order as the
// caller sees it is unchanged; the launch latency between two kernels is
// hidden.  VTR_NO_PDL=1 launches them the plain way.
struct LaunchConfig {
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[2];
};
static void launch_config(LaunchConfig& l, dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                          int cluster_x = 0) {
  static const bool no_pdl = getenv("VTR_NO_PDL") != nullptr;
  l.cfg = cudaLaunchConfig_t{};
  l.cfg.gridDim = grid, l.cfg.blockDim = block;
  l.cfg.dynamicSmemBytes = smem;
  l.cfg.stream = stream;
  l.cfg.attrs = l.attr;
  unsigned int n = 0;
  if (cluster_x > 0) {
    l.attr[n].id = cudaLaunchAttributeClusterDimension;
    l.attr[n].val.clusterDim.x = cluster_x;
    l.attr[n].val.clusterDim.y = 1;
    l.attr[n].val.clusterDim.z = 1;
    ++n;
  }
  if (!no_pdl) {
    l.attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    l.attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  l.cfg.numAttrs = n;
}

template <int A_CT, bool GRADS>
static int launch_tiles(const TileArgs& args, int tiles, int warps, size_t smem, cudaStream_t stream) {
  if (g_plan.dry_run) {
    record_plan("%s<A=%d,%s> grid=%d warps/tile=%d ring=%d smem=%zu%s%s; ",
                warps > 1 ? "fused_relay_kernel" : "fused_column_kernel", A_CT, GRADS ? "grads" : "fwd",
                tiles, warps, args.f.depth, smem, args.fin.ticket ? " + in-kernel final reduction" : "",
                args.fin.ex_world > 0 ? (args.fin.ex_sums ? " + peer publish + collect" : " + peer publish")
                                      : "");
    return VTR_OK;
  }
  DeviceInfo dev;
  const bool known = device_info(&dev) == VTR_OK && !getenv("VTR_RELAY_NO_SOLO");
  const bool solo = known && warps > 1 && warps <= 16 && tiles <= dev.sm_count;
  const bool duo = known && !solo && warps > 1 && warps <= 10 && tiles <= 2 * dev.sm_count;
  auto kernel = warps > 1 ? (solo ? fused_relay_kernel<A_CT, GRADS, kRelaySolo>
                                  : (duo ? fused_relay_kernel<A_CT, GRADS, kRelayDuo>
                                         : fused_relay_kernel<A_CT, GRADS, kRelayShared>))
                          : fused_column_kernel<A_CT, GRADS>;
  if (int rc = allow_dynamic_smem(reinterpret_cast<const void*>(kernel), smem)) return rc;
  LaunchConfig l;
  launch_config(l, dim3(tiles), dim3(32 * warps), smem, stream);
  VTR_CUDA_CHECK(cudaLaunchKernelEx(&l.cfg, kernel, args));
  return VTR_OK;
}

template <bool GRADS>
static int dispatch_tiles(const TileArgs& args, int tiles, int warps, size_t smem, cudaStream_t stream) {
  switch (args.f.A) {
    case 3: return launch_tiles<3, GRADS>(args, tiles, warps, smem, stream);
    case 4: return launch_tiles<4, GRADS>(args, tiles, warps, smem, stream);
    case 6: return launch_tiles<6, GRADS>(args, tiles, warps, smem, stream);
    case 9: return launch_tiles<9, GRADS>(args, tiles, warps, smem, stream);
    case 18: return launch_tiles<18, GRADS>(args, tiles, warps, smem, stream);
    default: return launch_tiles<0, GRADS>(args, tiles, warps, smem, stream);
  }
}

// Tuning overrides for experiments (not part of the ABI):
// VTR_COLUMN_WARPS in [1,16], VTR_COLUMN_DEPTH in [1,16].
static int env_int(const char* name) {
  const char* v = getenv(name);
  return v ? atoi(v) : 0;
}

static size_t slot_bytes_for(int A) {
  return (2u * kColumnsPerTile * A + 4u * kColumnsPerTile) * sizeof(float);
}

// Shared memory of a tile with `warps` warps and `depth` ring slots.
static size_t tile_smem_bytes(int A, int depth, int warps, bool grads) {
  const size_t row = kColumnsPerTile * A * sizeof(float);
  if (warps == 1) return kColumnBarrierBytes + depth * slot_bytes_for(A) + (grads ? 2 * row : 0);
  const size_t per_warp = (grads ? row : 0) + 96 * sizeof(float) + 3 * sizeof(double);
  return kRelayBarrierBytes + depth * slot_bytes_for(A) + warps * per_warp;
}

static void fill_args(FusedArgs& a, const vtr_fused_params* p, double* partials) {
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
  a.losses = nullptr, a.losses_f64 = nullptr;
  a.total_out = p->total_loss;
  a.depth = 0, a.slot_floats = 0, a.tma_in_ok = 0, a.tma_out_ok = 0;
}

// K1c needs every row it bulk-copies to start and end on 16 bytes.
static bool sm_tile_eligible(const vtr_fused_params* p) {
  if (p->B % 4 != 0) return false;
  const void* ptrs[] = {p->behaviour_policy_logits, p->target_policy_logits, p->actions, p->rewards,
                        p->values, p->done ? nullptr : p->discounts, p->d_target_logits};
  for (const void* q : ptrs)
    if (!aligned(q, 16)) return false;
  return true;
}

// VTR_PATH_AUTO: K1c from two 32-column groups per SM upwards (measured on
// B200 at T=100, A=18: B=16384 81 us vs 90 us for K1b, B=49152 220 us vs
// 278 us); below that its single producer lane is the bottleneck and K1b,
// where every warp loads its own steps, is faster (B=4096: 26 us vs 45 us).
static bool sm_tile_preferred(const vtr_fused_params* p) {
  DeviceInfo dev;
  if (device_info(&dev) != VTR_OK) return false;
  // Short unrolls leave no time to amortise filling the shared ring
  // (T=20 B=16384: 28.8 us against 22.7 us for K1b).
  return sm_tile_eligible(p) && p->B >= 64 * dev.sm_count && p->T >= 32;
}

template <int A_CT, bool GRADS>
static int launch_sm_tile(const SmTileArgs& args, int grid, size_t smem, cudaStream_t stream) {
  if (g_plan.dry_run) {
    record_plan("fused_sm_tile_kernel<A=%d,%s> grid=%d groups=%d relay=%d ring=%d smem=%zu %s%s; ", A_CT,
                GRADS ? "grads" : "fwd", grid, args.groups, args.relay, args.f.depth, smem,
                args.cursor ? "speed-weighted tiles, in-kernel final reduction" : "even tiles",
                args.ex_world > 0 ? (args.ex_sums ? " + peer publish + collect" : " + peer publish") : "");
    return VTR_OK;
  }
  auto kernel = fused_sm_tile_kernel<A_CT, GRADS>;
  if (int rc = allow_dynamic_smem(reinterpret_cast<const void*>(kernel), smem)) return rc;
  LaunchConfig l;
  launch_config(l, dim3(grid), dim3(32 * (args.groups * args.relay + 1)), smem, stream);
  VTR_CUDA_CHECK(cudaLaunchKernelEx(&l.cfg, kernel, args));
  return VTR_OK;
}

template <bool GRADS>
static int dispatch_sm_tile(const SmTileArgs& args, int grid, size_t smem, cudaStream_t stream) {
  switch (args.f.A) {
    case 3: return launch_sm_tile<3, GRADS>(args, grid, smem, stream);
    case 4: return launch_sm_tile<4, GRADS>(args, grid, smem, stream);
    case 6: return launch_sm_tile<6, GRADS>(args, grid, smem, stream);
    case 9: return launch_sm_tile<9, GRADS>(args, grid, smem, stream);
    case 18: return launch_sm_tile<18, GRADS>(args, grid, smem, stream);
    default: return launch_sm_tile<0, GRADS>(args, grid, smem, stream);
  }
}

// ---------------------------------------------------------------------------
// Tickets for the in-kernel final reduction (fused_finish.cuh).  The "who
// finished last" counter must be zero when a launch starts.  Zeroing it in the
// caller's workspace would cost a second stream operation per call - as much
// as the kernel itself on the small shapes - so the counters live in a ring
// inside the library: zero at load, each launch takes the next one, and the
// CTA that finishes last puts it back to zero.  A counter is handed out again
// kTicketRing launches later; a GPU runs at most 128 kernels at a time, so two
// launches never hold the same counter.  A launch that is being captured into
// a graph will be replayed with the counter it was captured with, any number
// of times and at any later moment, so it gets one of kTicketCaptured counters
// that are never handed out again; once those are used up captured launches
// leave the stage to K3.
// ---------------------------------------------------------------------------
constexpr int kTicketRing = 256;
constexpr int kTicketCaptured = 256;
constexpr int kTicketWords = 256;  // counters of one launch: K1c uses all, the others one
static __device__ unsigned int g_finish_tickets[(kTicketRing + kTicketCaptured) * kTicketWords];

// *out: kTicketWords zeroed counters for one launch, or null (then the caller
// zeroes counters of its own, or leaves the last stage to K3).
static int next_ticket(cudaStream_t stream, unsigned int** out) {
  constexpr int kMaxDevices = 64;
  static unsigned int* base[kMaxDevices] = {};
  static std::mutex mu;
  static std::atomic<unsigned int> turn{0}, captured{0};
  *out = nullptr;
  cudaStreamCaptureStatus capturing = cudaStreamCaptureStatusNone;
  VTR_CUDA_CHECK(cudaStreamIsCapturing(stream, &capturing));
  int dev = 0;
  VTR_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(VTR_ERR_UNSUPPORTED, "device index %d", dev);
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!base[dev]) {
      if (capturing != cudaStreamCaptureStatusNone) return VTR_OK;  // no symbol lookup while capturing
      void* ptr = nullptr;
      VTR_CUDA_CHECK(cudaGetSymbolAddress(&ptr, g_finish_tickets));
      base[dev] = static_cast<unsigned int*>(ptr);
    }
  }
  if (capturing != cudaStreamCaptureStatusNone) {
    const unsigned int k = captured.fetch_add(1u, std::memory_order_relaxed);
    if (k < static_cast<unsigned int>(kTicketCaptured))
      *out = base[dev] + static_cast<size_t>(kTicketRing + k) * kTicketWords;
    return VTR_OK;
  }
  *out = base[dev] + static_cast<size_t>(turn.fetch_add(1u, std::memory_order_relaxed) % kTicketRing) * kTicketWords;
  return VTR_OK;
}

// Workspace layout (vtr_fused_workspace_bytes): an 8 KB header -- K1c's column
// cursor, its finished-group ticket and per-group granule counters in the
// first KB (zeroed by the library before every K1c launch), its group
// partials after that -- then the loss partials.
constexpr size_t kWorkspaceHeaderBytes = 8192;  // counters in [0,1024), group partials after
constexpr size_t kWorkspaceCounterBytes = 1024;
constexpr int kMaxReductionGroups = 240;
constexpr int kReductionGroupGranules = 512;

// K1c launch: `*num_partials` = partials written (one per granule of 4 columns).
// `header` is the 64-byte workspace header, or null when the caller gave no
// workspace (then tiles are dealt evenly instead of by measured SM speed).
// Returns VTR_ERR_UNSUPPORTED without launching if the shape does not fit this
// kernel's shared memory.
static int run_sm_tile_path(const vtr_fused_params* p, void* header, double* partials,
                            cudaStream_t stream, int* num_partials, bool* losses_done,
                            bool* exchange_done) {
  DeviceInfo dev;
  if (int rc = device_info(&dev)) return rc;
  const bool grads = p->d_target_logits != nullptr;
  const int sm = dev.sm_count;
  const int total_gran = p->B / 4;

  // Grid: a multiple of the SM count once the batch is wide enough to give
  // every SM at least one 32-column group; tiles of at most ~256 columns.
  int waves = 1, grid;
  if (p->B < 32 * sm) {
    grid = (p->B + 31) / 32;
  } else {
    waves = (p->B + sm * 224 - 1) / (sm * 224);
    grid = sm * waves;
  }
  const bool balance = header != nullptr && grid % sm == 0 && sm <= kMaxSms &&
                       !getenv("VTR_SM_NO_BALANCE");
  const int even_cols = ((total_gran + grid - 1) / grid) * 4;
  // Tile capacity: the even share, or with balancing 1.25x + 2 granules of it
  // (weighted_share clamps to 1.25x and adds at most a couple of granules).
  int cap_cols = even_cols;
  if (balance) {
    int headroom = env_int("VTR_SM_HEADROOM");  // percent, experiments only
    if (headroom <= 0) headroom = 25;
    cap_cols = even_cols + even_cols * headroom / 100 + 8;
  }
  SmTileArgs args;
  fill_args(args.f, p, partials);
  args.groups = (cap_cols + 31) / 32;
  args.num_sm = sm;
  args.waves = waves;
  // The counters: from the library's ring, which needs no zeroing (the tile
  // that finishes the reduction leaves them zero), else the first KB of the
  // workspace, zeroed with a memset before the launch.
  args.cursor = nullptr, args.self_clean = 0;
  if (balance) {
    unsigned int* ring = nullptr;
    const bool in_kernel_finish = p->losses || p->losses_f64;
    if (in_kernel_finish && !g_plan.dry_run && !getenv("VTR_SM_MEMSET"))
      if (int rc = next_ticket(stream, &ring)) return rc;
    args.cursor = ring ? ring : static_cast<unsigned int*>(header);
    args.self_clean = ring != nullptr;
  }
  args.group_partials = balance ? reinterpret_cast<double*>(static_cast<char*>(header) +
                                                            kWorkspaceCounterBytes)
                                : nullptr;
  args.group_gran = kReductionGroupGranules;
  if ((total_gran + args.group_gran - 1) / args.group_gran > kMaxReductionGroups)
    args.group_gran = (total_gran + kMaxReductionGroups - 1) / kMaxReductionGroups;
  args.smoothing = 0.05f;
  if (const int pct = env_int("VTR_SM_SMOOTHING")) args.smoothing = 0.01f * pct;  // experiments only
  args.share_pct = 100;
  if (const int pct = env_int("VTR_SM_SHARE_PCT")) args.share_pct = min(100, max(1, pct));  // tests only

  if (balance) args.f.losses = p->losses, args.f.losses_f64 = p->losses_f64;
  *losses_done = balance;
  // The tile that finishes the reduction also publishes to the peers.
  args.ex_world = 0, args.ex_rank = 0, args.ex_seq = 0, args.ex_lag = 0, args.ex_sums = nullptr;
  for (float*& q : args.ex_peers.p) q = nullptr;
  if (balance && p->exchange_peer_buffers) {
    if (int rc = fill_peer_buffers(&args.ex_peers, p->exchange_peer_buffers, p->exchange_world)) return rc;
    args.ex_world = p->exchange_world, args.ex_rank = p->exchange_rank, args.ex_seq = p->exchange_seq;
    args.ex_lag = p->exchange_lag, args.ex_sums = p->exchange_sums;
    *exchange_done = true;
  }

  // Consumer warps: ~16 per SM.  Two tiles share an SM when the batch needs
  // two or more waves, so that the waves overlap instead of queueing.
  const int co_resident = waves >= 2 ? 2 : 1;
  const int max_consumers = kSmMaxConsumerWarps / co_resident;
  if (args.groups > max_consumers) return fail(VTR_ERR_UNSUPPORTED, "sm-tile kernel: tile too wide");
  int relay = max(1, max_consumers / args.groups);
  if (relay * args.groups > 16 / co_resident && relay > 2) relay = max(2, 16 / co_resident / args.groups);
  if (const int forced = env_int("VTR_SM_RELAY")) relay = forced;
  if (relay > p->T) relay = p->T;
  if (relay * args.groups > max_consumers) relay = max_consumers / args.groups;
  if (relay < 1) relay = 1;

  size_t budget = static_cast<size_t>(dev.smem_per_sm) / co_resident;
  budget = budget > 2048 ? budget - 1024 : budget;
  if (budget > static_cast<size_t>(dev.max_smem_optin)) budget = dev.max_smem_optin;
  const size_t row = 32 * static_cast<size_t>(p->A) * sizeof(float);
  const size_t slot_bytes = 32u * args.groups * (2u * p->A + 4u) * sizeof(float);
  size_t fixed = 0;
  long depth = 0;
  for (;; --relay) {
    args.relay = relay;
    args.staging_bufs = relay == 1 ? 2 : 1;
    fixed = kSmBarrierBytes + args.groups * relay *
                                  ((grads ? args.staging_bufs * row : 0) + 96 * sizeof(float) +
                                   96 * sizeof(double));
    depth = budget > fixed ? static_cast<long>((budget - fixed) / slot_bytes) : 0;
    if (const int forced = env_int("VTR_COLUMN_DEPTH")) depth = forced;
    if (depth > kMaxDepth) depth = kMaxDepth;
    if (depth > p->T) depth = p->T;
    // A warp waits for use n of a slot by parity, which only means something
    // once use n-1 has completed.  The relay order guarantees that for steps
    // more than `relay` back, so a reused slot must be further back than that.
    if (relay == 1 || depth >= p->T || relay < depth) break;
  }
  if (depth < 2 && p->T >= 2) return fail(VTR_ERR_UNSUPPORTED, "sm-tile kernel: ring does not fit");
  if (depth < 1) depth = 1;
  args.f.depth = static_cast<int>(depth);
  const size_t smem = fixed + depth * slot_bytes;
  if (balance && !args.self_clean && !g_plan.dry_run)
    VTR_CUDA_CHECK(cudaMemsetAsync(header, 0, kWorkspaceCounterBytes, stream));
  *num_partials = total_gran;
  return grads ? dispatch_sm_tile<true>(args, grid, smem, stream)
               : dispatch_sm_tile<false>(args, grid, smem, stream);
}

// Fills the in-kernel final stage of K2/K5.  `wanted`: losses were asked for
// and the last CTA can afford the sum.  Sets *losses_done / *exchange_done to
// what the launch will do itself.
static int fill_finish(FinishArgs* fin, const vtr_fused_params* p, bool wanted, int num_ctas,
                       int num_partials, cudaStream_t stream, bool* losses_done, bool* exchange_done) {
  fin->ticket = nullptr;
  fin->num_ctas = num_ctas, fin->num_partials = num_partials;
  fin->losses = p->losses, fin->losses_f64 = p->losses_f64;
  fin->ex_world = 0, fin->ex_rank = 0, fin->ex_seq = 0, fin->ex_lag = 0, fin->ex_sums = nullptr;
  for (float*& q : fin->ex_peers.p) q = nullptr;
  *losses_done = *exchange_done = false;
  if (!wanted || getenv("VTR_NO_INKERNEL_FINISH")) return VTR_OK;
  if (g_plan.dry_run) {
    static unsigned int dummy;
    fin->ticket = &dummy;  // never dereferenced
  } else if (int rc = next_ticket(stream, &fin->ticket)) {
    return rc;
  }
  if (!fin->ticket) return VTR_OK;
  *losses_done = true;
  if (p->exchange_peer_buffers) {
    if (int rc = fill_peer_buffers(&fin->ex_peers, p->exchange_peer_buffers, p->exchange_world)) return rc;
    fin->ex_world = p->exchange_world, fin->ex_rank = p->exchange_rank, fin->ex_seq = p->exchange_seq;
    fin->ex_lag = p->exchange_lag, fin->ex_sums = p->exchange_sums;
    *exchange_done = true;
  }
  return VTR_OK;
}

// K5.  Steps per chunk S (= warps per CTA) and chunks C (= CTAs per cluster):
// about 12 steps per chunk, at most 8 chunks (the portable cluster size); a
// longer unroll takes up to kChunkMaxChunks chunks if the device schedules
// such clusters, and falls back to the 32-column-tile kernels otherwise.
constexpr int kChunkMaxColumns = 16384;  // C partials per group stay within one per column

static size_t chunk_smem_bytes(int A, int S) {
  return kChunkBarrierBytes + static_cast<size_t>(S) * 64 * A * sizeof(float) +
         (static_cast<size_t>(S) * 64 + kChunkMaxChunks * 64) * sizeof(float) +
         static_cast<size_t>(S) * 3 * sizeof(double);
}

static bool chunk_shape(const vtr_fused_params* p, int* steps, int* chunks) {
  if (p->B > kChunkMaxColumns) return false;
  DeviceInfo dev;
  if (device_info(&dev) != VTR_OK) return false;
  int S = env_int("VTR_CHUNK_STEPS");
  if (S <= 0) {
    const int c0 = min(8, max(1, (p->T + 11) / 12));
    S = (p->T + c0 - 1) / c0;
  }
  S = max(1, min(S, min(p->T, kChunkMaxSteps)));
  while (S > 1 && chunk_smem_bytes(p->A, S) > static_cast<size_t>(dev.max_smem_optin)) --S;
  if (chunk_smem_bytes(p->A, S) > static_cast<size_t>(dev.max_smem_optin)) return false;
  const int C = (p->T + S - 1) / S;
  if (C > kChunkMaxChunks) return false;
  *steps = S, *chunks = C;
  return true;
}

template <int A_CT, bool GRADS>
static int launch_chunk(const ChunkArgs& args, int groups, size_t smem, cudaStream_t stream) {
  if (g_plan.dry_run) {
    record_plan("fused_chunk_kernel<A=%d,%s> grid=%dx%d cluster=%d warps/cta=%d smem=%zu%s%s; ", A_CT,
                GRADS ? "grads" : "fwd", args.chunks, groups, args.chunks, args.steps, smem,
                args.fin.ticket ? " + in-kernel final reduction" : "",
                args.fin.ex_world > 0 ? (args.fin.ex_sums ? " + peer publish + collect" : " + peer publish")
                                      : "");
    return VTR_OK;
  }
  auto kernel = fused_chunk_kernel<A_CT, GRADS>;
  if (int rc = allow_dynamic_smem(reinterpret_cast<const void*>(kernel), smem)) return rc;
  if (args.chunks > 8) {
    // Clusters of more than 8 CTAs are opt-in and may not be schedulable.
    static std::mutex mu;
    static int allowed[64] = {};  // per device: 0 unknown, 1 yes, -1 no (per template instance)
    int dev = 0;
    VTR_CUDA_CHECK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && allowed[dev] == 0)
      allowed[dev] = cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) ==
                             cudaSuccess
                         ? 1
                         : -1;
    cudaGetLastError();
    if (dev < 0 || dev >= 64 || allowed[dev] < 0)
      return fail(VTR_ERR_UNSUPPORTED, "chunk kernel: clusters of %d CTAs are not available", args.chunks);
  }
  LaunchConfig l;
  launch_config(l, dim3(args.chunks, groups, 1), dim3(32 * args.steps), smem, stream, args.chunks);
  cudaLaunchConfig_t& cfg = l.cfg;
  if (args.chunks > 8) {
    int clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&clusters, kernel, &cfg) != cudaSuccess || clusters < 1) {
      cudaGetLastError();
      return fail(VTR_ERR_UNSUPPORTED, "chunk kernel: a cluster of %d CTAs does not fit this device",
                  args.chunks);
    }
  }
  VTR_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kernel, args));
  return VTR_OK;
}

template <bool GRADS>
static int dispatch_chunk(const ChunkArgs& args, int groups, size_t smem, cudaStream_t stream) {
  switch (args.f.A) {
    case 3: return launch_chunk<3, GRADS>(args, groups, smem, stream);
    case 4: return launch_chunk<4, GRADS>(args, groups, smem, stream);
    case 6: return launch_chunk<6, GRADS>(args, groups, smem, stream);
    case 9: return launch_chunk<9, GRADS>(args, groups, smem, stream);
    case 18: return launch_chunk<18, GRADS>(args, groups, smem, stream);
    default: return launch_chunk<0, GRADS>(args, groups, smem, stream);
  }
}

static int run_chunk_path(const vtr_fused_params* p, double* partials, cudaStream_t stream,
                          int* num_partials, bool* losses_done, bool* exchange_done) {
  int S = 0, C = 0;
  if (!chunk_shape(p, &S, &C))
    return fail(VTR_ERR_UNSUPPORTED, "chunk kernel: needs B <= %d and T <= %d steps per chunk x %d chunks",
                kChunkMaxColumns, kChunkMaxSteps, kChunkMaxChunks);
  const int groups = column_tiles(p->B);
  ChunkArgs args;
  fill_args(args.f, p, partials);
  const bool rows_aligned = (static_cast<size_t>(p->B) * p->A) % 4 == 0;
  args.f.tma_in_ok = rows_aligned && aligned(args.f.blog, 16) && aligned(args.f.tlog, 16);
  args.f.tma_out_ok = rows_aligned && aligned(args.f.d_logits, 16);
  args.steps = S, args.chunks = C;
  *num_partials = groups * C;  // one partial per CTA
  if (int rc = fill_finish(&args.fin, p, partials != nullptr, groups * C, groups * C, stream, losses_done,
                           exchange_done))
    return rc;
#ifdef VTR_CHUNK_TRACE
  args.trace = nullptr;
  if (const char* v = getenv("VTR_CHUNK_TRACE_PTR")) args.trace = reinterpret_cast<unsigned long long*>(strtoull(v, nullptr, 16));
#endif
  const size_t smem = chunk_smem_bytes(p->A, S);
  return p->d_target_logits ? dispatch_chunk<true>(args, groups, smem, stream)
                            : dispatch_chunk<false>(args, groups, smem, stream);
}

// VTR_PATH_AUTO takes K2 (a CTA per column, a lane per time step) where
// lane-per-column kernels would leave most lanes idle, and for long unrolls of
// small batches that K5 cannot split (T beyond its steps x chunks): its
// threads read their rows straight from global memory, 72-byte pieces 2 MB
// apart, which is fine for a few hundred columns and slower than K1b's ring
// beyond (measured at T=100 A=18: B=256 14 us vs 25 us, B=512 20 us vs 25 us,
// B=1024 29 us vs 26 us, B=2048 50 us vs 23 us).
constexpr int kItemTinyColumns = 16;
constexpr int kItemAutoColumns = 512;
static bool item_kernel_applies(const vtr_fused_params* p, int max_columns) {
  return p->T <= kItemMaxSteps && specialised_actions(p->A) && p->B <= max_columns;
}

// VTR_PATH_AUTO takes K5 while its CTAs (column groups x chunks) fit the GPU in
// little more than one wave: there every SM gets work however few column
// groups there are, and the launch costs one CTA's latency.  A CTA does load,
// compute, store one after the other, so with several waves the 32-column
// tiles, whose ring keeps loads ahead of the arithmetic, are faster (measured
// on B200, T=100 A=18: B=1024 12.3 vs 22.6 us, 2048 19.6 vs 23.9, 3072 25.1 vs
// 24.6, 4096 31.5 vs 24.6; T=20: 4096 10.3 vs 11.6, 8192 14.1 vs 13.9, 16384
// 24.7 vs 20.6).
static bool chunk_kernel_applies(const vtr_fused_params* p) {
  int S = 0, C = 0;
  if (!chunk_shape(p, &S, &C) || C > 8 || getenv("VTR_NO_CHUNK")) return false;
  DeviceInfo dev;
  if (device_info(&dev) != VTR_OK) return false;
  // CTAs an SM holds: registers (32 for a compiled-in A <= 9, else 48), shared
  // memory, threads.
  const int regs = (specialised_actions(p->A) && p->A <= 9) ? 32 : 48;
  int per_sm = 65536 / (regs * 32 * S);
  per_sm = min(per_sm, static_cast<int>(dev.smem_per_sm / (chunk_smem_bytes(p->A, S) + 1024)));
  per_sm = min(per_sm, 2048 / (32 * S));
  if (per_sm < 1) return false;
  const long ctas = static_cast<long>(column_tiles(p->B)) * C;
  return 4 * ctas <= 5L * per_sm * dev.sm_count;
}

template <int A_CT, bool GRADS>
static int launch_item(const ItemArgs& args, cudaStream_t stream) {
  const int threads = ((args.f.T + 31) / 32) * 32;
  if (g_plan.dry_run) {
    record_plan("fused_item_kernel<A=%d,%s> grid=%d threads=%d%s%s; ", A_CT, GRADS ? "grads" : "fwd", args.f.B,
                threads, args.fin.ticket ? " + in-kernel final reduction" : "",
                args.fin.ex_world > 0 ? (args.fin.ex_sums ? " + peer publish + collect" : " + peer publish")
                                      : "");
    return VTR_OK;
  }
  LaunchConfig l;
  launch_config(l, dim3(args.f.B), dim3(threads), 0, stream);
  VTR_CUDA_CHECK(cudaLaunchKernelEx(&l.cfg, fused_item_kernel<A_CT, GRADS>, args));
  return VTR_OK;
}

template <bool GRADS>
static int dispatch_item(const ItemArgs& args, cudaStream_t stream) {
  switch (args.f.A) {
    case 3: return launch_item<3, GRADS>(args, stream);
    case 4: return launch_item<4, GRADS>(args, stream);
    case 6: return launch_item<6, GRADS>(args, stream);
    case 9: return launch_item<9, GRADS>(args, stream);
    case 18: return launch_item<18, GRADS>(args, stream);
    default: return fail(VTR_ERR_UNSUPPORTED, "item kernel: A=%d is not a compiled-in size", args.f.A);
  }
}

// The last CTA of K2 sums one partial per column with at most 256 threads:
// fine for the small batches this kernel is meant for, K3 beyond.
constexpr int kItemFinishColumns = 1024;

static int run_item_path(const vtr_fused_params* p, double* partials, cudaStream_t stream,
                         int* num_partials, bool* losses_done, bool* exchange_done) {
  if (p->T > kItemMaxSteps || p->B > kItemMaxColumns)
    return fail(VTR_ERR_UNSUPPORTED, "item kernel: needs T <= %d and B <= %d", kItemMaxSteps,
                kItemMaxColumns);
  ItemArgs a;
  fill_args(a.f, p, partials);
  *num_partials = p->B;  // one partial per column
  if (int rc = fill_finish(&a.fin, p, partials != nullptr && p->B <= kItemFinishColumns, p->B, p->B, stream,
                           losses_done, exchange_done))
    return rc;
  return p->d_target_logits ? dispatch_item<true>(a, stream) : dispatch_item<false>(a, stream);
}

// The last CTA of K1a/K1b sums one partial per tile with one warp: fine up to
// a thousand tiles, K3 beyond.
constexpr int kTileFinishTiles = 1024;

static int run_column_path(const vtr_fused_params* p, double* partials, cudaStream_t stream,
                           int* num_partials, bool* losses_done, bool* exchange_done) {
  DeviceInfo dev;
  if (int rc = device_info(&dev)) return rc;
  const int tiles = column_tiles(p->B);
  const bool grads = p->d_target_logits != nullptr;
  *num_partials = tiles;

  TileArgs targs;
  FusedArgs& a = targs.f;
  fill_args(a, p, partials);
  if (int rc = fill_finish(&targs.fin, p, partials != nullptr && tiles <= kTileFinishTiles, tiles, tiles,
                           stream, losses_done, exchange_done))
    return rc;

  // A row of the tile for step t starts at element (t*B + 32*tile)*A; TMA bulk
  // copies need 16-byte alignment of every such address.
  const bool rows_aligned = (static_cast<size_t>(p->B) * p->A) % 4 == 0;
  a.tma_in_ok = rows_aligned && aligned(a.blog, 16) && aligned(a.tlog, 16);
  a.tma_out_ok = rows_aligned && aligned(a.d_logits, 16);

  // Shared memory per tile = what the SM has, divided by the tiles that must be
  // co-resident on it (everything runs in one wave up to 8 tiles per SM).
  const size_t slot_bytes = slot_bytes_for(p->A);
  int resident = (tiles + dev.sm_count - 1) / dev.sm_count;
  resident = resident < 1 ? 1 : (resident > 8 ? 8 : resident);
  size_t budget = static_cast<size_t>(dev.smem_per_sm) / resident;
  budget = budget > 2048 ? budget - 1024 : budget;  // per-block reservation
  if (budget > static_cast<size_t>(dev.max_smem_optin)) budget = dev.max_smem_optin;

  // Warps per tile.  With 7+ tiles per SM the ring is worth more than extra
  // warps (measured: T=100 B=32768 A=18 runs 142 us with one warp and 5 slots,
  // 166 us with 3 warps and 4 slots); below that, aim at ~24 warps per SM while
  // the ring still runs at least 2 slots ahead of the warps.
  int warps = resident >= 7 ? 1 : 24 / resident;
  // Two tiles per SM: 10 warps each instead of 12 buys the kernel 96 registers
  // instead of 80 (the DUO variant of K1b).
  if (resident == 2 && !getenv("VTR_RELAY_NO_SOLO")) warps = 10;
  if (const int forced = env_int("VTR_COLUMN_WARPS")) warps = forced;
  if (warps > kRelayMaxWarps) warps = kRelayMaxWarps;
  if (warps > p->T) warps = p->T;
  if (warps < 1) warps = 1;
  long depth = 0;
  for (;; --warps) {
    const size_t fixed = tile_smem_bytes(p->A, 0, warps, grads);
    depth = budget > fixed ? static_cast<long>((budget - fixed) / slot_bytes) : 0;
    if (warps == 1 || depth >= warps + 2) break;
  }
  if (const int forced = env_int("VTR_COLUMN_DEPTH")) depth = forced;
  if (depth > (warps > 1 ? kRelayMaxDepth : kMaxDepth)) depth = warps > 1 ? kRelayMaxDepth : kMaxDepth;
  if (depth > p->T) depth = p->T;
  if (depth < 2) depth = p->T < 2 ? 1 : 2;
  // Waiting on a slot by parity needs the slot's previous use to be complete;
  // the relay order guarantees that only for steps more than `warps` back.
  if (depth < p->T && warps >= depth) warps = static_cast<int>(depth) - 1;
  if (warps < 1) warps = 1;
  const size_t smem = tile_smem_bytes(p->A, static_cast<int>(depth), warps, grads);
  if (smem > static_cast<size_t>(dev.max_smem_optin))
    return fail(VTR_ERR_UNSUPPORTED, "A=%d needs %zu bytes of shared memory, device allows %d",
                p->A, smem, dev.max_smem_optin);
  a.depth = static_cast<int>(depth);
  a.slot_floats = static_cast<int>(slot_bytes / sizeof(float));

  return grads ? dispatch_tiles<true>(targs, tiles, warps, smem, stream)
               : dispatch_tiles<false>(targs, tiles, warps, smem, stream);
}

// Which kernels a call may run, in the order they are tried: K1c if wanted,
// then `second`; under VTR_PATH_AUTO a kernel that turns the shape down
// (VTR_ERR_UNSUPPORTED) hands over to the 32-column-tile kernels, which take
// everything.
enum SecondChoice { kSecondColumn, kSecondItem, kSecondChunk };
struct KernelChoice {
  bool sm_tile;
  SecondChoice second;
};

static KernelChoice choose_kernels(const vtr_fused_params* p) {
  KernelChoice c;
  c.sm_tile = p->path == VTR_PATH_SM_TILE || (p->path == VTR_PATH_AUTO && sm_tile_preferred(p));
  c.second = kSecondColumn;
  if (p->path == VTR_PATH_ITEM) c.second = kSecondItem;
  else if (p->path == VTR_PATH_CHUNK) c.second = kSecondChunk;
  else if (p->path == VTR_PATH_AUTO) {
    if (item_kernel_applies(p, kItemTinyColumns)) c.second = kSecondItem;
    else if (chunk_kernel_applies(p)) c.second = kSecondChunk;
    else if (item_kernel_applies(p, kItemAutoColumns)) c.second = kSecondItem;
  }
  return c;
}

// Loss partials (double[3] each) the kernels of `c` may write: one per
// 4-column granule (K1c), per column (K2), per CTA (K5) or per 32-column tile
// (K1a/K1b); the largest of those the call can end up running.
static size_t partials_needed(const vtr_fused_params* p, const KernelChoice& c) {
  const size_t tiles = static_cast<size_t>(column_tiles(p->B));
  size_t need = 0;
  if (c.sm_tile) need = static_cast<size_t>(p->B / 4);
  if (p->path != VTR_PATH_SM_TILE) {
    size_t second = tiles;
    if (c.second == kSecondItem) second = static_cast<size_t>(p->B);
    if (c.second == kSecondChunk) {
      int S = 0, C = 1;
      if (chunk_shape(p, &S, &C)) second = tiles * C;  // else the call fails or falls back
    }
    if (second > need) need = second;
    if (p->path == VTR_PATH_AUTO && tiles > need) need = tiles;
  }
  return need;
}

}  // namespace vtr

extern "C" size_t vtr_fused_workspace_bytes(int32_t T, int32_t B, int32_t A) {
  if (T <= 0 || B <= 0 || A <= 0) return 0;
  // header + one double[3] per 32-column tile (K1a/K1b), per 4-column granule
  // (K1c), per column (K2, which only runs below kItemMaxColumns columns) or
  // per CTA (K5: at most kChunkMaxChunks per tile, below kChunkMaxColumns)
  const size_t tiles = static_cast<size_t>(vtr::column_tiles(B));
  size_t partials = tiles;
  if (static_cast<size_t>(B / 4) > partials) partials = B / 4;
  if (B <= vtr::kItemMaxColumns && static_cast<size_t>(B) > partials) partials = B;
  if (B <= vtr::kChunkMaxColumns && tiles * vtr::kChunkMaxChunks > partials)
    partials = tiles * vtr::kChunkMaxChunks;
  return vtr::kWorkspaceHeaderBytes + partials * 3 * sizeof(double);
}

static int reduce_losses(const double* partials, int32_t num_partials, float baseline_cost,
                         float entropy_cost, float* losses, double* losses_f64, float* total_out,
                         void* stream) {
  if (!partials || num_partials <= 0) return vtr::fail(VTR_ERR_INVALID_ARGUMENT, "vtr_reduce_losses: no partials");
  if (!losses && !losses_f64) return vtr::fail(VTR_ERR_INVALID_ARGUMENT, "vtr_reduce_losses: no output");
  if (vtr::g_plan.dry_run) {
    vtr::record_plan("reduce_losses_kernel grid=1 partials=%d; ", num_partials);
    return VTR_OK;
  }
  vtr::reduce_losses_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      partials, num_partials, baseline_cost, entropy_cost, losses, losses_f64, total_out);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

extern "C" int vtr_reduce_losses(const double* partials, int32_t num_partials,
                                 float baseline_cost, float entropy_cost, float* losses,
                                 double* losses_f64, void* stream) {
  return reduce_losses(partials, num_partials, baseline_cost, entropy_cost, losses, losses_f64, nullptr,
                       stream);
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
  if (p->path < VTR_PATH_AUTO || p->path > VTR_PATH_CHUNK)
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

  if (p->total_loss && !p->losses) return fail(VTR_ERR_INVALID_ARGUMENT, "total_loss needs `losses`");
  if (!vtr::aligned(p->total_loss, 4)) return fail(VTR_ERR_MISALIGNED, "total_loss is not 4-byte aligned");
  const bool want_exchange = p->exchange_peer_buffers != nullptr;
  if (want_exchange) {
    if (!p->losses) return fail(VTR_ERR_INVALID_ARGUMENT, "exchange_peer_buffers needs `losses`");
    if (p->exchange_world <= 0 || p->exchange_world > VTR_EXCHANGE_MAX_RANKS || p->exchange_rank < 0 ||
        p->exchange_rank >= p->exchange_world)
      return fail(VTR_ERR_INVALID_ARGUMENT, "bad exchange rank/world %d/%d", p->exchange_rank,
                  p->exchange_world);
    if (p->exchange_seq == 0) return fail(VTR_ERR_INVALID_ARGUMENT, "exchange_seq must start at 1");
    if (p->exchange_lag > 1) return fail(VTR_ERR_INVALID_ARGUMENT, "exchange_lag must be 0 or 1");
    if (!vtr::aligned(p->exchange_sums, 4)) return fail(VTR_ERR_MISALIGNED, "exchange_sums is not 4-byte aligned");
    for (int r = 0; r < p->exchange_world; ++r)
      if (!p->exchange_peer_buffers[r])
        return fail(VTR_ERR_INVALID_ARGUMENT, "exchange_peer_buffers[%d] is NULL", r);
  }
  if (p->path == VTR_PATH_SM_TILE && !vtr::sm_tile_eligible(p))
    return fail(VTR_ERR_UNSUPPORTED,
                "VTR_PATH_SM_TILE needs B %% 4 == 0 and 16-byte aligned tensors");
  if (p->path == VTR_PATH_ITEM && p->B > vtr::kItemMaxColumns)
    return fail(VTR_ERR_UNSUPPORTED, "VTR_PATH_ITEM needs B <= %d", vtr::kItemMaxColumns);
  const vtr::KernelChoice choice = vtr::choose_kernels(p);
  const bool want_losses = p->losses || p->losses_f64;
  double* partials = nullptr;
  if (want_losses) {
    // What the kernels this call may run really write; never more than
    // vtr_fused_workspace_bytes() says.
    const size_t need = vtr::kWorkspaceHeaderBytes + vtr::partials_needed(p, choice) * 3 * sizeof(double);
    if (!workspace || workspace_bytes < need)
      return fail(VTR_ERR_WORKSPACE, "workspace of %zu bytes needed, %zu given", need, workspace_bytes);
    if (!vtr::aligned(workspace, 8)) return fail(VTR_ERR_MISALIGNED, "workspace is not 8-byte aligned");
    partials = reinterpret_cast<double*>(static_cast<char*>(workspace) + vtr::kWorkspaceHeaderBytes);
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int num_partials = 0;
  int rc = VTR_ERR_UNSUPPORTED;
  const bool want_sm_tile = choice.sm_tile;
  bool losses_done = false, exchange_done = false;
  const vtr::SecondChoice second = choice.second;
  if (want_sm_tile)
    rc = vtr::run_sm_tile_path(p, want_losses ? workspace : nullptr, partials, s, &num_partials,
                               &losses_done, &exchange_done);
  if (rc == VTR_ERR_UNSUPPORTED && p->path != VTR_PATH_SM_TILE) {
    losses_done = exchange_done = false;
    if (second == vtr::kSecondItem)
      rc = vtr::run_item_path(p, partials, s, &num_partials, &losses_done, &exchange_done);
    else if (second == vtr::kSecondChunk)
      rc = vtr::run_chunk_path(p, partials, s, &num_partials, &losses_done, &exchange_done);
    else
      rc = vtr::run_column_path(p, partials, s, &num_partials, &losses_done, &exchange_done);
    // a cluster the device cannot schedule: the 32-column-tile kernels always run
    if (rc == VTR_ERR_UNSUPPORTED && second != vtr::kSecondColumn && p->path == VTR_PATH_AUTO) {
      rc = vtr::run_column_path(p, partials, s, &num_partials, &losses_done, &exchange_done);
    }
  }
  if (rc) return rc;
  if (want_losses && !losses_done) {
    rc = reduce_losses(partials, num_partials, p->baseline_cost, p->entropy_cost, p->losses,
                       p->losses_f64, p->total_loss, stream);
    if (rc) return rc;
  }
  if (want_exchange && !exchange_done) {
    if (vtr::g_plan.dry_run) {
      vtr::record_plan("exchange_publish_kernel grid=1; ");
      if (p->exchange_sums && p->exchange_seq > p->exchange_lag)
        vtr::record_plan("exchange_collect_kernel grid=1; ");
      return VTR_OK;
    }
    rc = vtr_exchange_publish_f32(p->losses, 4, p->exchange_peer_buffers, p->exchange_rank,
                                  p->exchange_world, p->exchange_seq, stream);
    if (rc || !p->exchange_sums || p->exchange_seq <= p->exchange_lag) return rc;
    return vtr_exchange_collect_f32(p->exchange_peer_buffers[p->exchange_rank], 4, p->exchange_world,
                                    p->exchange_seq - p->exchange_lag, p->exchange_sums, stream);
  }
  return VTR_OK;
}

extern "C" int vtr_fused_describe(const vtr_fused_params* p, int have_workspace, int* num_kernels,
                                  char* text, size_t text_bytes) {
  vtr::g_plan = vtr::PlanRecorder();
  vtr::g_plan.dry_run = true;
  // Validation wants a workspace pointer when losses are requested; nothing is
  // written through it in a dry run.
  static double dummy[8];
  const int rc = vtr_from_logits_fused_f32(p, have_workspace ? dummy : nullptr,
                                           have_workspace ? static_cast<size_t>(-1) : 0, nullptr);
  vtr::g_plan.dry_run = false;
  if (rc != VTR_OK) return rc;
  if (text && text_bytes) snprintf(text, text_bytes, "%s", vtr::g_plan.text);
  if (num_kernels) *num_kernels = vtr::g_plan.kernels;
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
// caller buffer.  Every GPU that loads the library carries these 200 KiB until
// the process exits, whether or not this entry point is ever called.
// ---------------------------------------------------------------------------
#include <mutex>

namespace vtr {

constexpr int kNoAllocMaxTiles = VTR_NOALLOC_MAX_B / kColumnsPerTile;
constexpr size_t kNoAllocDoubles = kWorkspaceHeaderBytes / sizeof(double) + kNoAllocMaxTiles * 3;
static __device__ __align__(256) double g_noalloc_partials[kNoAllocDoubles];

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
  if (!p->losses && !p->losses_f64)  // no losses, no partials: the scratch is not touched
    return vtr_from_logits_fused_f32(p, nullptr, 0, stream);
  if (p->T <= 0 || p->B <= 0 || p->A <= 0 || p->path < VTR_PATH_AUTO || p->path > VTR_PATH_CHUNK)
    return vtr_from_logits_fused_f32(p, nullptr, 0, stream);  // reports the bad argument
  double* scratch = nullptr;
  if (int rc = vtr::noalloc_scratch(&scratch)) return rc;
  // The scratch holds one partial per 32-column tile of the largest batch.
  // The other kernels write more partials per column, so the automatic choice
  // keeps them only while what they write still fits, and takes the 32-column
  // tiles beyond that.  The partials never have to start at zero: the first
  // stage writes every one the second stage reads, so whatever an earlier
  // call left behind is overwritten or never looked at.  The per-SM kernel
  // keeps a cursor in the scratch header; the library zeroes that itself
  // before every launch that uses it.
  const size_t scratch_bytes = sizeof(double) * vtr::kNoAllocDoubles;
  vtr_fused_params q = *p;
  const size_t need = vtr::kWorkspaceHeaderBytes +
                      vtr::partials_needed(&q, vtr::choose_kernels(&q)) * 3 * sizeof(double);
  if (need > scratch_bytes) {
    if (q.path != VTR_PATH_AUTO)
      return vtr::fail(VTR_ERR_UNSUPPORTED,
                       "vtr_from_logits_fused_noalloc_f32: the kernel asked for with path=%d writes %zu "
                       "bytes of loss partials at B=%d, the built-in scratch holds %zu; leave the "
                       "choice to the library (VTR_PATH_AUTO) or pass a workspace to "
                       "vtr_from_logits_fused_f32",
                       q.path, need, q.B, scratch_bytes);
    q.path = VTR_PATH_COLUMN;
  }
  return vtr_from_logits_fused_f32(&q, scratch, scratch_bytes, stream);
}
