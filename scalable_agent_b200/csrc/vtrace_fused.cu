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
//   K2  fused_item_kernel     fused_item.cuh     one thread per (t, b) item, tiny batches
//   K3  reduce_losses_kernel  (below)            second stage of the loss reduction
// Losses accumulate per lane in double and are reduced in fixed orders; the
// only atomics are two tickets in K1c (which tile next, who finished last).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "common.h"
#include "fused_column.cuh"
#include "fused_common.cuh"
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

template <int A_CT, bool GRADS>
static int launch_tiles(const FusedArgs& args, int tiles, int warps, size_t smem, cudaStream_t stream) {
  if (g_plan.dry_run) {
    record_plan("%s<A=%d,%s> grid=%d warps/tile=%d ring=%d smem=%zu; ",
                warps > 1 ? "fused_relay_kernel" : "fused_column_kernel", A_CT, GRADS ? "grads" : "fwd",
                tiles, warps, args.depth, smem);
    return VTR_OK;
  }
  auto kernel = warps > 1 ? fused_relay_kernel<A_CT, GRADS> : fused_column_kernel<A_CT, GRADS>;
  if (int rc = allow_dynamic_smem(reinterpret_cast<const void*>(kernel), smem)) return rc;
  kernel<<<tiles, 32 * warps, smem, stream>>>(args);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

template <bool GRADS>
static int dispatch_tiles(const FusedArgs& args, int tiles, int warps, size_t smem, cudaStream_t stream) {
  switch (args.A) {
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
  kernel<<<grid, 32 * (args.groups * args.relay + 1), smem, stream>>>(args);
  VTR_CUDA_CHECK(cudaGetLastError());
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
  args.cursor = balance ? static_cast<unsigned int*>(header) : nullptr;
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
  args.ex_world = 0, args.ex_rank = 0, args.ex_seq = 0, args.ex_sums = nullptr;
  for (float*& q : args.ex_peers.p) q = nullptr;
  if (balance && p->exchange_peer_buffers) {
    if (int rc = fill_peer_buffers(&args.ex_peers, p->exchange_peer_buffers, p->exchange_world)) return rc;
    args.ex_world = p->exchange_world, args.ex_rank = p->exchange_rank, args.ex_seq = p->exchange_seq;
    args.ex_sums = p->exchange_sums;
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
  if (balance && !g_plan.dry_run)
    VTR_CUDA_CHECK(cudaMemsetAsync(header, 0, kWorkspaceCounterBytes, stream));
  *num_partials = total_gran;
  return grads ? dispatch_sm_tile<true>(args, grid, smem, stream)
               : dispatch_sm_tile<false>(args, grid, smem, stream);
}

// VTR_PATH_AUTO takes K2 for the smallest batches only: its threads read their
// rows straight from global memory, 72-byte pieces 2 MB apart, which is fine
// for a few hundred columns and slower than K1b's ring beyond (measured at
// T=100 A=18: B=256 14 us vs 25 us, B=512 20 us vs 25 us, B=1024 29 us vs 26 us,
// B=2048 50 us vs 23 us).
constexpr int kItemAutoColumns = 512;
static bool item_kernel_applies(const vtr_fused_params* p) {
  return p->T <= kItemMaxSteps && specialised_actions(p->A) && p->B <= kItemAutoColumns;
}

template <int A_CT, bool GRADS>
static int launch_item(const FusedArgs& args, cudaStream_t stream) {
  const int threads = ((args.T + 31) / 32) * 32;
  if (g_plan.dry_run) {
    record_plan("fused_item_kernel<A=%d,%s> grid=%d threads=%d; ", A_CT, GRADS ? "grads" : "fwd", args.B,
                threads);
    return VTR_OK;
  }
  fused_item_kernel<A_CT, GRADS><<<args.B, threads, 0, stream>>>(args);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

template <bool GRADS>
static int dispatch_item(const FusedArgs& args, cudaStream_t stream) {
  switch (args.A) {
    case 3: return launch_item<3, GRADS>(args, stream);
    case 4: return launch_item<4, GRADS>(args, stream);
    case 6: return launch_item<6, GRADS>(args, stream);
    case 9: return launch_item<9, GRADS>(args, stream);
    case 18: return launch_item<18, GRADS>(args, stream);
    default: return fail(VTR_ERR_UNSUPPORTED, "item kernel: A=%d is not a compiled-in size", args.A);
  }
}

static int run_item_path(const vtr_fused_params* p, double* partials, cudaStream_t stream,
                         int* num_partials) {
  if (p->T > kItemMaxSteps || p->B > kItemMaxColumns)
    return fail(VTR_ERR_UNSUPPORTED, "item kernel: needs T <= %d and B <= %d", kItemMaxSteps,
                kItemMaxColumns);
  FusedArgs a;
  fill_args(a, p, partials);
  *num_partials = p->B;  // one partial per column
  return p->d_target_logits ? dispatch_item<true>(a, stream) : dispatch_item<false>(a, stream);
}

static int run_column_path(const vtr_fused_params* p, double* partials, cudaStream_t stream,
                           int* num_partials) {
  DeviceInfo dev;
  if (int rc = device_info(&dev)) return rc;
  const int tiles = column_tiles(p->B);
  const bool grads = p->d_target_logits != nullptr;
  *num_partials = tiles;

  FusedArgs a;
  fill_args(a, p, partials);

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
  if (const int forced = env_int("VTR_COLUMN_WARPS")) warps = forced;
  if (warps > kMaxWarps) warps = kMaxWarps;
  if (warps > p->T) warps = p->T;
  if (warps < 1) warps = 1;
  long depth = 0;
  for (;; --warps) {
    const size_t fixed = tile_smem_bytes(p->A, 0, warps, grads);
    depth = budget > fixed ? static_cast<long>((budget - fixed) / slot_bytes) : 0;
    if (warps == 1 || depth >= warps + 2) break;
  }
  if (const int forced = env_int("VTR_COLUMN_DEPTH")) depth = forced;
  if (depth > kMaxDepth) depth = kMaxDepth;
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

  return grads ? dispatch_tiles<true>(a, tiles, warps, smem, stream)
               : dispatch_tiles<false>(a, tiles, warps, smem, stream);
}

}  // namespace vtr

extern "C" size_t vtr_fused_workspace_bytes(int32_t T, int32_t B, int32_t A) {
  if (T <= 0 || B <= 0 || A <= 0) return 0;
  // header + one double[3] per 32-column tile (K1a/K1b), per 4-column granule
  // (K1c) or per column (K2, which only runs below kItemMaxColumns columns)
  size_t partials = static_cast<size_t>(vtr::column_tiles(B));
  if (static_cast<size_t>(B / 4) > partials) partials = B / 4;
  if (B <= vtr::kItemMaxColumns && static_cast<size_t>(B) > partials) partials = B;
  return vtr::kWorkspaceHeaderBytes + partials * 3 * sizeof(double);
}

extern "C" int vtr_reduce_losses(const double* partials, int32_t num_partials,
                                 float baseline_cost, float entropy_cost, float* losses,
                                 double* losses_f64, void* stream) {
  if (!partials || num_partials <= 0) return vtr::fail(VTR_ERR_INVALID_ARGUMENT, "vtr_reduce_losses: no partials");
  if (!losses && !losses_f64) return vtr::fail(VTR_ERR_INVALID_ARGUMENT, "vtr_reduce_losses: no output");
  if (vtr::g_plan.dry_run) {
    vtr::record_plan("reduce_losses_kernel grid=1 partials=%d; ", num_partials);
    return VTR_OK;
  }
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
  if (p->path < VTR_PATH_AUTO || p->path > VTR_PATH_ITEM)
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

  const bool want_exchange = p->exchange_peer_buffers != nullptr;
  if (want_exchange) {
    if (!p->losses) return fail(VTR_ERR_INVALID_ARGUMENT, "exchange_peer_buffers needs `losses`");
    if (p->exchange_world <= 0 || p->exchange_world > VTR_EXCHANGE_MAX_RANKS || p->exchange_rank < 0 ||
        p->exchange_rank >= p->exchange_world)
      return fail(VTR_ERR_INVALID_ARGUMENT, "bad exchange rank/world %d/%d", p->exchange_rank,
                  p->exchange_world);
    if (p->exchange_seq == 0) return fail(VTR_ERR_INVALID_ARGUMENT, "exchange_seq must start at 1");
    if (!vtr::aligned(p->exchange_sums, 4)) return fail(VTR_ERR_MISALIGNED, "exchange_sums is not 4-byte aligned");
    for (int r = 0; r < p->exchange_world; ++r)
      if (!p->exchange_peer_buffers[r])
        return fail(VTR_ERR_INVALID_ARGUMENT, "exchange_peer_buffers[%d] is NULL", r);
  }
  const bool want_losses = p->losses || p->losses_f64;
  double* partials = nullptr;
  if (want_losses) {
    // K1a/K1b write one partial per 32-column tile, K1c one per 4-column granule.
    const size_t need = p->path == VTR_PATH_COLUMN
                            ? vtr::kWorkspaceHeaderBytes + vtr::column_tiles(p->B) * 3 * sizeof(double)
                            : vtr_fused_workspace_bytes(p->T, p->B, p->A);
    if (p->path == VTR_PATH_ITEM && p->B > vtr::kItemMaxColumns)
      return fail(VTR_ERR_UNSUPPORTED, "VTR_PATH_ITEM needs B <= %d", vtr::kItemMaxColumns);
    if (!workspace || workspace_bytes < need)
      return fail(VTR_ERR_WORKSPACE, "workspace of %zu bytes needed, %zu given", need, workspace_bytes);
    if (!vtr::aligned(workspace, 8)) return fail(VTR_ERR_MISALIGNED, "workspace is not 8-byte aligned");
    partials = reinterpret_cast<double*>(static_cast<char*>(workspace) + vtr::kWorkspaceHeaderBytes);
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int num_partials = 0;
  int rc = VTR_ERR_UNSUPPORTED;
  const bool want_sm_tile = p->path == VTR_PATH_SM_TILE ||
                            (p->path == VTR_PATH_AUTO && vtr::sm_tile_preferred(p));
  if (p->path == VTR_PATH_SM_TILE && !vtr::sm_tile_eligible(p))
    return fail(VTR_ERR_UNSUPPORTED,
                "VTR_PATH_SM_TILE needs B %% 4 == 0 and 16-byte aligned tensors");
  bool losses_done = false, exchange_done = false;
  if (want_sm_tile)
    rc = vtr::run_sm_tile_path(p, want_losses ? workspace : nullptr, partials, s, &num_partials,
                               &losses_done, &exchange_done);
  if (rc == VTR_ERR_UNSUPPORTED && p->path != VTR_PATH_SM_TILE) {
    losses_done = exchange_done = false;
    if (p->path == VTR_PATH_ITEM || (p->path == VTR_PATH_AUTO && vtr::item_kernel_applies(p)))
      rc = vtr::run_item_path(p, partials, s, &num_partials);
    else
      rc = vtr::run_column_path(p, partials, s, &num_partials);
  }
  if (rc) return rc;
  if (want_losses && !losses_done) {
    rc = vtr_reduce_losses(partials, num_partials, p->baseline_cost, p->entropy_cost, p->losses,
                           p->losses_f64, stream);
    if (rc) return rc;
  }
  if (want_exchange && !exchange_done) {
    if (vtr::g_plan.dry_run) {
      vtr::record_plan("exchange_publish_kernel grid=1; ");
      if (p->exchange_sums) vtr::record_plan("exchange_collect_kernel grid=1; ");
      return VTR_OK;
    }
    rc = vtr_exchange_publish_f32(p->losses, 4, p->exchange_peer_buffers, p->exchange_rank,
                                  p->exchange_world, p->exchange_seq, stream);
    if (rc || !p->exchange_sums) return rc;
    return vtr_exchange_collect_f32(p->exchange_peer_buffers[p->exchange_rank], 4, p->exchange_world,
                                    p->exchange_seq, p->exchange_sums, stream);
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
  double* scratch = nullptr;
  if (int rc = vtr::noalloc_scratch(&scratch)) return rc;
  // The scratch holds one partial per 32-column tile of the largest batch.  The
  // other kernels write more partials per column, so they are only eligible
  // while their workspace still fits; beyond that this entry point takes the
  // 32-column-tile kernels.  The partials never have to start at zero: the
  // first stage writes every one the second stage reads, so whatever an
  // earlier call left behind is overwritten or never looked at.  The per-SM
  // kernel keeps a cursor in the scratch header; the library zeroes that
  // itself before every launch that uses it.
  const size_t scratch_bytes = sizeof(double) * vtr::kNoAllocDoubles;
  vtr_fused_params q = *p;
  if (q.path == VTR_PATH_AUTO && vtr_fused_workspace_bytes(q.T, q.B, q.A) > scratch_bytes)
    q.path = VTR_PATH_COLUMN;
  return vtr_from_logits_fused_f32(&q, scratch, scratch_bytes, stream);
}

