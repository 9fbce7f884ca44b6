// Device side of the peer-memory loss exchange, shared by the stand-alone
// publish kernel (vtrace_exchange.cu) and the per-SM fused kernel, which
// publishes from the tile that finishes the reduction (fused_sm_tile.cuh).
#pragma once
#include <stdint.h>

#include "../../include/vtrace_b200.h"

namespace vtr {

constexpr int kSlotFloats = VTR_EXCHANGE_SLOT_FLOATS;
// Buffer of one rank: float[kExchangeSlots][world][8]; slot [seq % 4][r] holds
// rank r's values of step seq in [0..5] and seq itself in [7].  Four slots let
// a rank run ahead: it publishes step s+4 (which reuses the slot of step s)
// only after it has collected step s+2, which needs every peer's publish of
// s+2, which that peer's stream orders after its own collect of step s.
constexpr uint32_t kExchangeSlots = VTR_EXCHANGE_SLOTS;

struct PeerBuffers {
  float* p[VTR_EXCHANGE_MAX_RANKS];
};

// Stores values[0..n) and then the sequence number into slot [seq % 4][rank]
// of ONE peer's buffer: values first, then the flag with system-scope release
// semantics (which orders this thread's earlier stores before it), so a peer
// that sees the flag sees the values.
__device__ __forceinline__ void exchange_publish_to_peer(float* peer, const float* values, int n,
                                                         int rank, int world, uint32_t seq) {
  float* slot = peer + (static_cast<size_t>(seq % kExchangeSlots) * world + rank) * kSlotFloats;
  for (int i = 0; i < n; ++i) slot[i] = values[i];
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(reinterpret_cast<uint32_t*>(slot + 7)),
               "r"(seq)
               : "memory");
}

// The same for exactly four values held in registers.
__device__ __forceinline__ void exchange_publish4_to_peer(float* peer, float v0, float v1, float v2,
                                                          float v3, int rank, int world, uint32_t seq) {
  float* slot = peer + (static_cast<size_t>(seq % kExchangeSlots) * world + rank) * kSlotFloats;
  *reinterpret_cast<float4*>(slot) = make_float4(v0, v1, v2, v3);  // slots are 32-byte aligned
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(reinterpret_cast<uint32_t*>(slot + 7)),
               "r"(seq)
               : "memory");
}

constexpr unsigned long long kExchangeTimeoutNs = 2000000000ull;  // 2 s

// Called by one whole warp.  Lane r waits until rank r's slot of step `seq`
// in this rank's own buffer carries `seq`, then lane 0 sums the slots
// in rank order in double.  A peer that never publishes must not hang the GPU:
// after the timeout the sums are NaN.
__device__ __forceinline__ void exchange_collect_warp(const float* mine, int n, int world, uint32_t seq,
                                                      float* sums, unsigned long long timeout_ns) {
  const int r = threadIdx.x & 31;
  const float* base = mine + static_cast<size_t>(seq % kExchangeSlots) * world * kSlotFloats;
  int ok = 1;
  if (r < world) {
    const uint32_t* flag = reinterpret_cast<const uint32_t*>(base + r * kSlotFloats + 7);
    unsigned long long t0, now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      uint32_t v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
      if (v == seq) break;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (now - t0 > timeout_ns) {
        ok = 0;
        break;
      }
      __nanosleep(100);
    }
  }
  const bool all_ok = __all_sync(0xffffffffu, ok);
  if (r == 0) {
    for (int i = 0; i < n; ++i) {
      double s = 0.0;  // rank order: the same bits on every rank
      for (int q = 0; q < world; ++q)
        s += static_cast<double>(*reinterpret_cast<const volatile float*>(base + q * kSlotFloats + i));
      sums[i] = all_ok ? static_cast<float>(s) : __int_as_float(0x7fc00000);
    }
  }
}

// Host: validates and copies the caller's pointer array.
int fill_peer_buffers(PeerBuffers* out, void* const* peer_buffers, int world);

}  // namespace vtr
