// Cross-GPU sum of the loss scalars over NVLink peer memory.
//
// Sharded over the GPUs of one box the path needs one
// exchange per call: every rank's 4 loss scalars.  A collective library call
// costs a kernel launch per rank plus its own rendezvous for 16 bytes; here
// every rank stores its values straight into every peer's buffer (P2P stores
// through NVLink / NVSwitch, buffers from torch symmetric memory) followed by a
// sequence number, and sums what has arrived in rank order.  The sum is
// bitwise identical on all ranks.
//
// Buffer of one rank: float[2][world][8]; slot [seq & 1][r] holds rank r's
// values in [0..5] and the sequence number in [7].  Two parities suffice: rank
// Y can publish step s+2 only after collecting step s+1, which needs X's
// publish of s+1, which X's stream orders after its collect of step s.
#include "common.h"

namespace vtr {

constexpr int kSlotFloats = VTR_EXCHANGE_SLOT_FLOATS;

struct PeerBuffers {
  float* p[VTR_EXCHANGE_MAX_RANKS];
};

__global__ void __launch_bounds__(32) exchange_publish_kernel(const float* __restrict__ values, int n,
                                                              PeerBuffers peers, int rank, int world,
                                                              uint32_t seq) {
  const int r = threadIdx.x;
  if (r >= world) return;
  float* slot = peers.p[r] + (static_cast<size_t>(seq & 1u) * world + rank) * kSlotFloats;
  for (int i = 0; i < n; ++i) slot[i] = values[i];
  __threadfence_system();  // values before the sequence number, system-wide
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(reinterpret_cast<uint32_t*>(slot + 7)),
               "r"(seq)
               : "memory");
}

__global__ void __launch_bounds__(32) exchange_collect_kernel(const float* __restrict__ mine, int n,
                                                              int world, uint32_t seq,
                                                              float* __restrict__ sums,
                                                              unsigned long long timeout_ns) {
  __shared__ int sh_ok[32];
  const int r = threadIdx.x;
  const float* base = mine + static_cast<size_t>(seq & 1u) * world * kSlotFloats;
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
      if (now - t0 > timeout_ns) {  // a peer never published: do not hang the GPU
        ok = 0;
        break;
      }
      __nanosleep(100);
    }
  }
  sh_ok[r] = ok;
  __syncwarp();
  if (r == 0) {
    bool all_ok = true;
    for (int i = 0; i < world; ++i) all_ok = all_ok && sh_ok[i];
    for (int i = 0; i < n; ++i) {
      double s = 0.0;  // rank order: the same bits on every rank
      for (int q = 0; q < world; ++q)
        s += static_cast<double>(*reinterpret_cast<const volatile float*>(base + q * kSlotFloats + i));
      sums[i] = all_ok ? static_cast<float>(s) : __int_as_float(0x7fc00000);
    }
  }
}

}  // namespace vtr

using vtr::fail;

extern "C" size_t vtr_exchange_buffer_bytes(int32_t world) {
  if (world <= 0 || world > VTR_EXCHANGE_MAX_RANKS) return 0;
  return sizeof(float) * 2 * world * VTR_EXCHANGE_SLOT_FLOATS;
}

extern "C" int vtr_exchange_publish_f32(const float* values, int32_t n, void* const* peer_buffers,
                                        int32_t rank, int32_t world, uint32_t seq, void* stream) {
  if (!values || !peer_buffers) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_exchange_publish: NULL pointer");
  if (world <= 0 || world > VTR_EXCHANGE_MAX_RANKS || rank < 0 || rank >= world)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_exchange_publish: bad rank/world %d/%d", rank, world);
  if (n <= 0 || n > 6) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_exchange_publish: n must be 1..6");
  if (seq == 0) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_exchange_publish: seq must start at 1");
  vtr::PeerBuffers peers;
  for (int r = 0; r < world; ++r) {
    if (!peer_buffers[r]) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_exchange_publish: peer %d is NULL", r);
    peers.p[r] = static_cast<float*>(peer_buffers[r]);
  }
  vtr::exchange_publish_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(values, n, peers, rank,
                                                                                 world, seq);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}

extern "C" int vtr_exchange_collect_f32(const void* my_buffer, int32_t n, int32_t world, uint32_t seq,
                                        float* sums, void* stream) {
  if (!my_buffer || !sums) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_exchange_collect: NULL pointer");
  if (world <= 0 || world > VTR_EXCHANGE_MAX_RANKS)
    return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_exchange_collect: bad world %d", world);
  if (n <= 0 || n > 6) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_exchange_collect: n must be 1..6");
  if (seq == 0) return fail(VTR_ERR_INVALID_ARGUMENT, "vtr_exchange_collect: seq must start at 1");
  const unsigned long long timeout_ns = 2000000000ull;  // 2 s
  vtr::exchange_collect_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const float*>(my_buffer), n, world, seq, sums, timeout_ns);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}
