// Device side of the peer-memory loss exchange, shared by the stand-alone
// publish kernel (vtrace_exchange.cu) and the per-SM fused kernel, which
// publishes from the tile that finishes the reduction (fused_sm_tile.cuh).
#pragma once
#include <stdint.h>

#include "../../include/vtrace_b200.h"

namespace vtr {

constexpr int kSlotFloats = VTR_EXCHANGE_SLOT_FLOATS;
// Buffer of one rank: float[kExchangeSlots][world][16]; slot [seq % 4][r] is
// eight 8-byte words, word i = {value i of rank r, seq}.  A word is written
// with one 64-bit store, which is single-copy atomic, so every word says for
// itself which step it belongs to: no flag behind a release fence, no NVLink
// round trip between the values and a flag.  Four slots let a rank run ahead:
// it publishes step s+4 (which reuses the slot of step s) only after it has
// collected step s+2, which needs every peer's publish of s+2, which that
// peer's stream orders after its own collect of step s.
constexpr uint32_t kExchangeSlots = VTR_EXCHANGE_SLOTS;
constexpr int kExchangeMaxValues = 6;

struct PeerBuffers {
  float* p[VTR_EXCHANGE_MAX_RANKS];
};

__device__ __forceinline__ void exchange_store_word(float* slot, int i, float value, uint32_t seq) {
  const unsigned long long word =
      (static_cast<unsigned long long>(seq) << 32) | static_cast<unsigned long long>(__float_as_uint(value));
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(reinterpret_cast<unsigned long long*>(slot) + i),
               "l"(word)
               : "memory");
}

// Stores values[0..n) of step `seq` into slot [seq % 4][rank] of ONE peer's buffer.
__device__ __forceinline__ void exchange_publish_to_peer(float* peer, const float* values, int n,
                                                         int rank, int world, uint32_t seq) {
  float* slot = peer + (static_cast<size_t>(seq % kExchangeSlots) * world + rank) * kSlotFloats;
  for (int i = 0; i < n; ++i) exchange_store_word(slot, i, values[i], seq);
}

// The same for exactly four values held in registers.
__device__ __forceinline__ void exchange_publish4_to_peer(float* peer, float v0, float v1, float v2,
                                                          float v3, int rank, int world, uint32_t seq) {
  float* slot = peer + (static_cast<size_t>(seq % kExchangeSlots) * world + rank) * kSlotFloats;
  exchange_store_word(slot, 0, v0, seq);
  exchange_store_word(slot, 1, v1, seq);
  exchange_store_word(slot, 2, v2, seq);
  exchange_store_word(slot, 3, v3, seq);
}

constexpr unsigned long long kExchangeTimeoutNs = 2000000000ull;  // 2 s

// Called by one whole warp.  Lane r reads the n words of rank r's slot of
// step `seq` in this rank's own buffer until all of them carry `seq`; the
// values are then summed in rank order in double (shuffles, the same order on
// every rank, so the same bits).  A peer that never publishes must not hang
// the GPU: after the timeout the sums are NaN.
__device__ __forceinline__ void exchange_collect_warp(const float* mine, int n, int world, uint32_t seq,
                                                      float* sums, unsigned long long timeout_ns) {
  const int r = threadIdx.x & 31;
  const float* base = mine + static_cast<size_t>(seq % kExchangeSlots) * world * kSlotFloats;
  int ok = 1;
  float v[kExchangeMaxValues];
#pragma unroll
  for (int i = 0; i < kExchangeMaxValues; ++i) v[i] = 0.0f;
  if (r < world) {
    const unsigned long long* words = reinterpret_cast<const unsigned long long*>(base + r * kSlotFloats);
    unsigned long long t0, now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      bool all = true;
#pragma unroll
      for (int i = 0; i < kExchangeMaxValues; ++i) {
        if (i < n) {
          unsigned long long w;
          asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w) : "l"(words + i) : "memory");
          all = all && static_cast<uint32_t>(w >> 32) == seq;
          v[i] = __uint_as_float(static_cast<uint32_t>(w));
        }
      }
      if (all) break;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (now - t0 > timeout_ns) {
        ok = 0;
        break;
      }
      __nanosleep(100);
    }
  }
  const bool all_ok = __all_sync(0xffffffffu, ok);
#pragma unroll
  for (int i = 0; i < kExchangeMaxValues; ++i) {
    if (i < n) {  // n is the same in all lanes
      double s = 0.0;
      for (int q = 0; q < world; ++q) s += static_cast<double>(__shfl_sync(0xffffffffu, v[i], q));
      if (r == 0) sums[i] = all_ok ? static_cast<float>(s) : __int_as_float(0x7fc00000);
    }
  }
}

// Host: validates and copies the caller's pointer array.
int fill_peer_buffers(PeerBuffers* out, void* const* peer_buffers, int world);

}  // namespace vtr
