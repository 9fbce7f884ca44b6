// The policy and baseline heads that feed the path:
//   experiment.py:199-203  policy_logits = snt.Linear(num_actions)(core_output)
//                          baseline      = snt.Linear(1)(core_output)
//   experiment.py:237      applied over [T*B, 256] by snt.BatchApply
// i.e. one [rows, K] x [K, A+1] contraction, K = 256 - the one dense
// contraction next to the V-trace path, and the one place tensor cores are
// owed.  It reads 4K bytes per row and writes 4(A+1): HBM-bound on the core
// output, so the job of the kernel is to stream [rows, K] at copy speed with
// the arithmetic hidden behind it.
//
// tcgen05 kernel, one persistent CTA per SM, roles by warp:
//   warps 0-3  epilogue   TMEM -> registers (tcgen05.ld), + bias, -> HBM
//   warps 4-7  splitters  a thread pulls its row of the landed fp32 tile out of
//                         shared memory, splits it into (hi, lo) TF32 halves and
//                         stores them into TENSOR MEMORY (tcgen05.st)
//   warp  8    producer   TMA tensor-map loads (cp.async.bulk.tensor) of the
//                         core-output tiles into a ring, 128B-swizzled
//   warp  9    MMA        one elected lane issues tcgen05.mma.kind::tf32 with
//                         the A operand in TMEM and the weights in shared
//                         memory, accumulators in TMEM, completion by
//                         tcgen05.commit
// The A operand goes through TMEM, not shared memory, because shared-memory
// bandwidth is what bounded the first version (A halves written to and read
// back from shared memory by three MMAs per k-step: 124 KB of shared-memory
// traffic per 16 KB landed, 3.9 us per tile against 2.9 us of HBM time; 902 us
// at T*B = 3.3 M rows).  Now a landed tile is read from shared memory once.
// Precision: the V-trace outputs are held to 1e-5 of a float64 oracle and the
// logits feed them, so one TF32 pass (10-bit mantissa) is not enough.  Each
// operand is split x = hi + lo with hi = x with the low 13 mantissa bits
// cleared (exactly what a TF32 multiplier keeps) and lo = x - hi (exact), and
// the product is accumulated as hi*hi + hi*lo + lo*hi + lo*lo in fp32 in TMEM
// (the "3xTF32" scheme plus the fourth term, which comes for free here; what is
// left is the truncation of the lo parts, 2^-22 relative).
// The tensor core's fp32 accumulation does not round to nearest - a chain of 96
// accumulations into one accumulator came out 1.5e-6 low on sums of magnitude 2,
// with the same sign everywhere - so a tile has four accumulators, one per
// k-step of a k-block (8 to 16 accumulations each), and the epilogue adds them
// in registers with round-to-nearest.  Spreading consecutive MMAs over
// different accumulators is also what lets the tensor pipe overlap them.
// The weights are split once (vtr_policy_head_prepare_f32); the activations are
// split on their way from shared memory to tensor memory.
#include <cuda.h>
#include <stdlib.h>

#include "common.h"
#include "ptx.cuh"

namespace vtr {
namespace head {

constexpr int kTileM = 128;          // rows per tile = TMEM lanes
constexpr int kTileN = 32;           // accumulator columns: A + 1 <= 32
constexpr int kBlockK = 32;          // fp32 elements per 128-byte swizzled row
constexpr int kUmmaK = 8;            // K of one tcgen05.mma.kind::tf32
constexpr int kMaxK = 256;
constexpr int kStages = 8;           // ring of landed core-output tiles (16 KB each)
constexpr int kAccStages = 1;        // tiles' accumulators in flight (the epilogue drains them in ~0.3 us)
constexpr int kAccSets = kBlockK / kUmmaK;  // 4 accumulators per tile: k-step kk of every k-block goes to set kk
constexpr int kAccN = 2 * kTileN;    // an accumulator: 32 columns of hi*hi, 32 of the two small terms
constexpr int kAccCols = kAccSets * kAccN;  // 4 x 64 columns per tile
// (hi, lo) A operands in tensor memory, 64 columns each.  Their number is what
// sets the pace: a buffer is busy from the splitters' tcgen05.st to the commit
// of the MMAs that read it, a round trip of ~3.5k cycles; with 3 buffers the
// kernel ran at one k-block per 1150 cycles (0.61 of the copy bandwidth), the
// splitters waiting for a free buffer a quarter of the time.  (More buffers
// did not help: see the note on the MMA count at the issue loop.)
constexpr int kABufs = 4;
constexpr int kABase = kAccStages * kAccCols;                     // columns [256, 512)
constexpr int kTmemCols = 512;       // all of an SM's tensor memory (power of two)
constexpr int kThreads = 320;
constexpr int kStageBytes = kTileM * kBlockK * 4;   // 16 KB
constexpr int kWBlockBytes = kTileN * kBlockK * 4;  // 4 KB: [32 n][32 k] of W_hi, then the same of W_lo
constexpr unsigned long long kWatchdogNs = 4000000000ull;  // a wait this long is a bug: trap, do not hang

struct HeadArgs {
  float* logits;    // [rows, A]
  float* baseline;  // [rows]
  const float* bias;  // [32]: b_policy, b_baseline, zeros
  long rows;
  int K, A;
  int num_tiles;
  int terms;  // 3; experiments: 1 = hi*hi only (TF32 accuracy), 0 = no MMA at all
};

// Shared memory, 1024-byte aligned (the 128B swizzle atom is 8 rows x 128 B):
//   raw   kStages x 16 KB   core tiles as they land ([128 rows][32 k], 128B-swizzled)
//   w_hi  (K/32) x 4 KB     weights^T [32 n][32 k] per k-block, hi / lo
//   w_lo  (K/32) x 4 KB
//   barriers, TMEM base address, bias
struct Barriers {
  unsigned long long full[kStages], empty[kStages];      // landed tile: TMA -> splitters -> producer
  unsigned long long a_full[kABufs], a_empty[kABufs];    // A operand in TMEM: splitters -> MMA -> splitters
  unsigned long long acc_full[kAccStages], acc_empty[kAccStages];
  unsigned long long weights;
  uint32_t tmem_base;
  float bias[kTileN];
};

static size_t smem_bytes(int K) {
  return 1024 + kStages * kStageBytes + 2u * (K / kBlockK) * kWBlockBytes + sizeof(Barriers);
}

// ---- PTX ---------------------------------------------------------------------
__device__ __forceinline__ void wait_or_trap(uint32_t bar, uint32_t parity) {
  // A failed try has been suspended for up to ptx::kMbarSuspendNs already:
  // failures are rare and looking at the clock on each of them costs nothing.
  unsigned long long t0 = 0;
  while (!ptx::mbar_try_wait(bar, parity)) {
    unsigned long long now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
    if (t0 == 0) t0 = now;
    else if (now - t0 > kWatchdogNs) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int x, int y, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void prefetch_tensormap(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem] * B[smem], M = 128, N = 32, K = 8, TF32 inputs, fp32 accumulate.
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// The same with the A operand in tensor memory (lane = row, one column per k).
__device__ __forceinline__ void mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 32 lanes x 32 columns from the warp's registers into tensor memory.
__device__ __forceinline__ void tmem_store_32x32(uint32_t addr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(addr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]),
      "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]),
      "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_store_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
// Arrives on the mbarrier once every tcgen05.mma issued so far by this thread has completed.
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_load_32x32(uint32_t addr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(addr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor of a K-major tile whose rows are 128 bytes,
// 128B-swizzled (what the tensor-map loads below write): start address and
// offsets in 16-byte units, 8-row groups 1024 bytes apart, version 1 (sm_100),
// layout type 2 = SWIZZLE_128B.  `byte_offset` steps along K inside the row.
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t byte_offset) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>(((addr + byte_offset) >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>(1) << 16;            // leading byte offset: unused with swizzle, 1 by convention
  d |= static_cast<uint64_t>(1024 >> 4) << 32;    // stride byte offset between 8-row groups
  d |= static_cast<uint64_t>(1) << 46;            // descriptor version
  d |= static_cast<uint64_t>(2) << 61;            // SWIZZLE_128B
  return d;
}
// Instruction descriptor: fp32 accumulator, TF32 A and B, both K-major, M = 128, N = 32 or 64.
constexpr uint32_t instr_desc(int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(n >> 3) << 17) | ((kTileM >> 4) << 24);
}
constexpr uint32_t kInstrDesc64 = instr_desc(2 * kTileN);

// ---- the kernel --------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1)
    policy_head_kernel(const __grid_constant__ CUtensorMap map_core, const __grid_constant__ CUtensorMap map_w_hi,
                       const __grid_constant__ CUtensorMap map_w_lo, const __grid_constant__ HeadArgs a) {
  extern __shared__ unsigned char smem_unaligned[];
  // the 128B swizzle pattern is a function of the address: tiles start on 1024 bytes
  unsigned char* smem = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_unaligned) + 1023) & ~static_cast<uintptr_t>(1023));
  const int num_kb = a.K / kBlockK;
  unsigned char* raw = smem;
  unsigned char* w_cat = raw + kStages * kStageBytes;  // per k-block 8 KB: [W_hi rows 0-31 | W_lo rows 32-63] x 32 k
  Barriers* bars = reinterpret_cast<Barriers*>(w_cat + 2 * num_kb * kWBlockBytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(ptx::smem_u32(&bars->full[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&bars->empty[s]), 4);  // one arrival per splitter warp
    }
    for (int j = 0; j < kABufs; ++j) {
      ptx::mbar_init(ptx::smem_u32(&bars->a_full[j]), 4);   // one arrival per splitter warp
      ptx::mbar_init(ptx::smem_u32(&bars->a_empty[j]), 1);  // tcgen05.commit
    }
    for (int s = 0; s < kAccStages; ++s) {
      ptx::mbar_init(ptx::smem_u32(&bars->acc_full[s]), 1);   // tcgen05.commit
      ptx::mbar_init(ptx::smem_u32(&bars->acc_empty[s]), 4);  // one arrival per epilogue warp
    }
    ptx::mbar_init(ptx::smem_u32(&bars->weights), 1);
    ptx::fence_mbar_init();
    ptx::fence_proxy_async();
  }
  if (threadIdx.x < kTileN) bars->bias[threadIdx.x] = a.bias[threadIdx.x];
  if (warp == 9) tmem_alloc(ptx::smem_u32(&bars->tmem_base), kTmemCols);  // whole warp; it also frees
  if (warp == 8 && lane == 0) {
    prefetch_tensormap(&map_core);
    prefetch_tensormap(&map_w_hi);
    prefetch_tensormap(&map_w_lo);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 8) {
    // ===== producer: weights once, then the core-output tiles through the ring =====
    if (lane == 0) {
      const uint32_t wbar = ptx::smem_u32(&bars->weights);
      ptx::mbar_arrive_expect_tx(wbar, 2u * num_kb * kWBlockBytes);
      for (int kb = 0; kb < num_kb; ++kb) {
        tma_load_2d(ptx::smem_u32(w_cat + 2 * kb * kWBlockBytes), &map_w_hi, kb * kBlockK, 0, wbar);
        tma_load_2d(ptx::smem_u32(w_cat + (2 * kb + 1) * kWBlockBytes), &map_w_lo, kb * kBlockK, 0, wbar);
      }
      int it = 0;
      for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kStages, use = it / kStages;
          if (use > 0) wait_or_trap(ptx::smem_u32(&bars->empty[s]), (use - 1) & 1);
          const uint32_t bar = ptx::smem_u32(&bars->full[s]);
          ptx::mbar_arrive_expect_tx(bar, kStageBytes);
          tma_load_2d(ptx::smem_u32(raw + s * kStageBytes), &map_core, kb * kBlockK, tile * kTileM, bar);
        }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ===== splitters: thread = row.  x -> hi (low 13 mantissa bits cleared) and lo = x - hi,
    // from the landed tile in shared memory into the A operand in tensor memory =====
    const int q = warp - 4;              // the warp's quarter of the TMEM lanes
    const int row = 32 * q + lane;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(32 * q) << 16);
    int it = 0;
    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int s = it % kStages, j = it % kABufs;
        wait_or_trap(ptx::smem_u32(&bars->full[s]), (it / kStages) & 1);
        // A row is 128 bytes = 8 chunks of 16; chunk c sits at position c ^ (row % 8) (128B swizzle),
        // which is also what makes 8 neighbouring rows hit different banks.
        const float4* src = reinterpret_cast<const float4*>(raw + s * kStageBytes + row * 128);
        uint32_t hi[32], lo[32];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 x = src[c ^ (row & 7)];
          const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint32_t h = __float_as_uint(xs[e]) & 0xffffe000u;
            hi[4 * c + e] = h;
            lo[4 * c + e] = __float_as_uint(__fsub_rn(xs[e], __uint_as_float(h)));
          }
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&bars->empty[s]));  // the landed tile is in registers
        if (it >= kABufs) wait_or_trap(ptx::smem_u32(&bars->a_empty[j]), ((it / kABufs) - 1) & 1);
        tc_fence_after();
        const uint32_t dst = lane_base + static_cast<uint32_t>(kABase + 64 * j);
        tmem_store_32x32(dst, hi);
        tmem_store_32x32(dst + 32, lo);
        tmem_store_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&bars->a_full[j]));
      }
    }
  } else if (warp == 9) {
    // ===== MMA issuer: A from tensor memory, weights from shared memory =====
    // The whole warp walks the loop (waits, addresses: warp-uniform values, cheap) and one elected
    // lane issues.  The instruction stream of this one warp was the bottleneck of earlier versions:
    // ~180 instructions per k-block around 12 MMAs, with every operand re-derived and moved to the
    // uniform registers behind an election, is ~1150 cycles - more than the 830 the HBM stream
    // leaves per k-block - while the tensor pipe itself was mostly idle (0.61 of the copy bandwidth
    // then, 0.90 with the MMAs switched off).
    wait_or_trap(ptx::smem_u32(&bars->weights), 0);
    const uint64_t wdesc0 = smem_desc(ptx::smem_u32(w_cat), 0);
    int it = 0, local_tile = 0;
    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x, ++local_tile) {
      const int acc = local_tile % kAccStages, acc_use = local_tile / kAccStages;
      if (acc_use > 0) wait_or_trap(ptx::smem_u32(&bars->acc_empty[acc]), (acc_use - 1) & 1);
      const uint32_t d0 = tmem_base + static_cast<uint32_t>(acc * kAccCols);
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int j = it % kABufs;
        wait_or_trap(ptx::smem_u32(&bars->a_full[j]), (it / kABufs) & 1);
        tc_fence_after();
        const uint32_t ah = tmem_base + static_cast<uint32_t>(kABase + 64 * j);
        // descriptor of k-block kb: the start address field moves by 8 KB (in 16-byte units)
        const uint64_t wd = wdesc0 + static_cast<uint64_t>(kb * (2 * kWBlockBytes >> 4));
        // Two MMAs per k-step of 8, both M128 x N64 x K8 against B = [W_hi | W_lo]: A_hi gives hi*hi in
        // columns 0-31 of the accumulator and hi*lo in 32-63, A_lo adds lo*hi and lo*lo on top - all
        // four partial products in two instructions.  Never two in a row into the same accumulator
        // (back-to-back MMAs on the same TMEM columns wait for each other): k-step kk accumulates
        // into set kk, and the four A_lo MMAs follow the four A_hi ones.
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kAccSets; ++kk)
            if (a.terms >= 1)
              mma_tf32_ts(d0 + kk * kAccN, ah + kk * kUmmaK, wd + 2 * kk, kInstrDesc64, kb != 0);
#pragma unroll
          for (int kk = 0; kk < kAccSets; ++kk)
            if (a.terms >= 3) mma_tf32_ts(d0 + kk * kAccN, ah + 32 + kk * kUmmaK, wd + 2 * kk, kInstrDesc64, 1);
          mma_commit(ptx::smem_u32(&bars->a_empty[j]));  // the A operand is free once these have read it
          if (kb == num_kb - 1) mma_commit(ptx::smem_u32(&bars->acc_full[acc]));  // the tile is complete
        }
        __syncwarp();
      }
    }
  } else {
    // ===== epilogue: warp w owns TMEM lanes 32w..32w+31 = rows of the tile =====
    const int num_acc = kAccSets;
    int local_tile = 0;
    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x, ++local_tile) {
      const int acc = local_tile % kAccStages, acc_use = local_tile / kAccStages;
      wait_or_trap(ptx::smem_u32(&bars->acc_full[acc]), acc_use & 1);
      tc_fence_after();
      float y[kTileN];
#pragma unroll
      for (int n = 0; n < kTileN; ++n) y[n] = 0.0f;
      for (int g = 0; g < 2 * num_acc; ++g) {  // small terms first (odd halves), then hi*hi; round-to-nearest
        const int half = g < num_acc ? 1 : 0, which = g < num_acc ? g : g - num_acc;
        uint32_t v[32];
        tmem_load_32x32(tmem_base + (static_cast<uint32_t>(warp * 32) << 16) +
                            static_cast<uint32_t>(acc * kAccCols + which * kAccN + half * kTileN), v);
#pragma unroll
        for (int n = 0; n < kTileN; ++n) y[n] = __fadd_rn(y[n], __uint_as_float(v[n]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&bars->acc_empty[acc]));  // registers hold the tile
      const long row = static_cast<long>(tile) * kTileM + warp * 32 + lane;
      if (row < a.rows) {
        float* out = a.logits + row * a.A;
#pragma unroll
        for (int n = 0; n < kTileN; ++n) {  // unrolled: y[] stays in registers
          const float out_n = __fadd_rn(y[n], bars->bias[n]);
          if (n < a.A) out[n] = out_n;
          else if (n == a.A) a.baseline[row] = out_n;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// Packs the two heads' weights for the kernel above: W^T padded to 32 output
// rows ([32][K], row n < A = policy column n, row A = baseline, rest zero),
// split into hi / lo TF32 halves, then the 32 biases.
__global__ void __launch_bounds__(256) pack_weights_kernel(const float* __restrict__ w_policy,
                                                           const float* __restrict__ b_policy,
                                                           const float* __restrict__ w_baseline,
                                                           const float* __restrict__ b_baseline, int K, int A,
                                                           float* __restrict__ packed) {
  float* w_hi = packed;
  float* w_lo = packed + kTileN * K;
  float* bias = packed + 2 * kTileN * K;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < kTileN * K; i += gridDim.x * blockDim.x) {
    const int n = i / K, k = i - n * K;
    float w = 0.0f;
    if (n < A) w = w_policy[static_cast<size_t>(k) * A + n];  // snt.Linear weights are [in, out]
    else if (n == A) w = w_baseline[k];
    const float hi = __uint_as_float(__float_as_uint(w) & 0xffffe000u);
    w_hi[i] = hi;
    w_lo[i] = __fsub_rn(w, hi);
  }
  if (blockIdx.x == 0 && threadIdx.x < kTileN) {
    const int n = threadIdx.x;
    bias[n] = n < A ? b_policy[n] : (n == A ? b_baseline[0] : 0.0f);
  }
}

// cuTensorMapEncodeTiled through the runtime's driver entry point: the library
// does not link libcuda.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int encode_fn(EncodeTiledFn* out) {
  static EncodeTiledFn cached = nullptr;
  if (!cached) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult status;
    VTR_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &status));
    if (!fn || status != cudaDriverEntryPointSuccess)
      return fail(VTR_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled is not available from this driver");
    cached = reinterpret_cast<EncodeTiledFn>(fn);
  }
  *out = cached;
  return VTR_OK;
}

// [rows, K] fp32 row-major, boxes of box_rows x 32 elements, 128B swizzle.
static int make_map(CUtensorMap* map, const void* base, uint64_t rows, uint64_t K, uint32_t box_rows) {
  EncodeTiledFn encode = nullptr;
  if (int rc = encode_fn(&encode)) return rc;
  const cuuint64_t dims[2] = {K, rows};
  const cuuint64_t strides[1] = {K * sizeof(float)};
  const cuuint32_t box[2] = {static_cast<cuuint32_t>(kBlockK), box_rows};
  const cuuint32_t elem[2] = {1, 1};
  const CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box,
                            elem, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(VTR_ERR_INVALID_ARGUMENT, "cuTensorMapEncodeTiled failed (%d)", static_cast<int>(r));
  return VTR_OK;
}

}  // namespace head
}  // namespace vtr

using vtr::fail;

static int check_head_shape(const char* who, int32_t K, int32_t A) {
  if (K <= 0 || K > vtr::head::kMaxK || K % vtr::head::kBlockK != 0)
    return fail(VTR_ERR_UNSUPPORTED, "%s: K must be a multiple of %d up to %d (got %d)", who, vtr::head::kBlockK,
                vtr::head::kMaxK, K);
  if (A <= 0 || A + 1 > vtr::head::kTileN)
    return fail(VTR_ERR_UNSUPPORTED, "%s: A must be in 1..%d (got %d)", who, vtr::head::kTileN - 1, A);
  return VTR_OK;
}

extern "C" size_t vtr_policy_head_packed_bytes(int32_t K, int32_t A) {
  if (K <= 0 || A <= 0) return 0;
  return (2u * vtr::head::kTileN * static_cast<size_t>(K) + vtr::head::kTileN) * sizeof(float);
}

extern "C" int vtr_policy_head_prepare_f32(const float* w_policy, const float* b_policy, const float* w_baseline,
                                           const float* b_baseline, int32_t K, int32_t A, void* packed,
                                           void* stream) {
  if (int rc = check_head_shape("vtr_policy_head_prepare", K, A)) return rc;
  if (!w_policy || !b_policy || !w_baseline || !b_baseline || !packed)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_policy_head_prepare: NULL pointer");
  if (!vtr::aligned(packed, 128)) return fail(VTR_ERR_MISALIGNED, "vtr_policy_head_prepare: packed must be 128-byte aligned");
  vtr::head::pack_weights_kernel<<<32, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w_policy, b_policy, w_baseline, b_baseline, K, A, static_cast<float*>(packed));
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

extern "C" int vtr_policy_head_f32(const float* core_output, const void* packed, int64_t rows, int32_t K, int32_t A,
                                   float* policy_logits, float* baseline, void* stream) {
  using namespace vtr::head;
  if (int rc = check_head_shape("vtr_policy_head", K, A)) return rc;
  if (!core_output || !packed || !policy_logits || !baseline)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_policy_head: NULL pointer");
  if (rows <= 0 || rows > (1LL << 31) - kTileM) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_policy_head: bad row count");
  if (!vtr::aligned(core_output, 16) || !vtr::aligned(packed, 128))
    return fail(VTR_ERR_MISALIGNED, "vtr_policy_head: core_output must be 16-byte, packed 128-byte aligned");
  if (!vtr::aligned(policy_logits, 4) || !vtr::aligned(baseline, 4))
    return fail(VTR_ERR_MISALIGNED, "vtr_policy_head: outputs must be 4-byte aligned");
  vtr::DeviceInfo dev;
  if (int rc = vtr::device_info(&dev)) return rc;
  const float* p = static_cast<const float*>(packed);
  CUtensorMap map_core, map_w_hi, map_w_lo;
  if (int rc = make_map(&map_core, core_output, static_cast<uint64_t>(rows), K, kTileM)) return rc;
  if (int rc = make_map(&map_w_hi, p, kTileN, K, kTileN)) return rc;
  if (int rc = make_map(&map_w_lo, p + kTileN * K, kTileN, K, kTileN)) return rc;
  HeadArgs a;
  a.logits = policy_logits, a.baseline = baseline, a.bias = p + 2 * kTileN * K;
  a.rows = rows, a.K = K, a.A = A;
  a.num_tiles = static_cast<int>((rows + kTileM - 1) / kTileM);
  a.terms = 3;
  if (const char* v = getenv("VTR_HEAD_TERMS")) a.terms = atoi(v);  // experiments only
  const size_t smem = smem_bytes(K);
  if (int rc = vtr::allow_dynamic_smem(reinterpret_cast<const void*>(policy_head_kernel), smem)) return rc;
  const int grid = a.num_tiles < dev.sm_count ? a.num_tiles : dev.sm_count;
  policy_head_kernel<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(map_core, map_w_hi, map_w_lo, a);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}
