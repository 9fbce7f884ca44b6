// Cross-GPU sum of the loss scalars over NVLink peer memory.
//
// Sharded over the GPUs of one box the path needs one
// exchange per call: every rank's 4 loss scalars.  A collective library call
// costs a kernel launch per rank plus its own rendezvous for 16 bytes; here
// every rank stores its values straight into every peer's buffer (P2P stores
// through NVLink / NVSwitch, buffers from torch symmetric memory), each value
// in one 8-byte word together with the step's sequence number, and sums what
// has arrived in rank order.  The sum is bitwise identical on all ranks.
//
// Buffer layout and the slot-reuse argument: exchange.cuh.
#include "common.h"
#include "exchange.cuh"

namespace vtr {

__global__ void __launch_bounds__(32) exchange_publish_kernel(const float* __restrict__ values, int n,
                                                              PeerBuffers peers, int rank, int world,
                                                              uint32_t seq) {
  const int r = threadIdx.x;
  if (r >= world) return;
  float v[6];
  for (int i = 0; i < n; ++i) v[i] = values[i];
  exchange_publish_to_peer(peers.p[r], v, n, rank, world, seq);
}

__global__ void __launch_bounds__(32) exchange_collect_kernel(const float* __restrict__ mine, int n,
                                                              int world, uint32_t seq,
                                                              float* __restrict__ sums,
                                                              unsigned long long timeout_ns) {
  exchange_collect_warp(mine, n, world, seq, sums, timeout_ns);
}

}  // namespace vtr

using vtr::fail;

extern "C" size_t vtr_exchange_buffer_bytes(int32_t world) {
  if (world <= 0 || world > VTR_EXCHANGE_MAX_RANKS) return 0;
  return sizeof(float) * VTR_EXCHANGE_SLOTS * world * VTR_EXCHANGE_SLOT_FLOATS;
}

namespace vtr {
int fill_peer_buffers(PeerBuffers* out, void* const* peer_buffers, int world) {
  if (!peer_buffers) return fail(VTR_ERR_INVALID_ARGUMENT, "exchange: peer_buffers is NULL");
  if (world <= 0 || world > VTR_EXCHANGE_MAX_RANKS)
    return fail(VTR_ERR_INVALID_ARGUMENT, "exchange: world must be 1..%d", VTR_EXCHANGE_MAX_RANKS);
  for (int r = 0; r < VTR_EXCHANGE_MAX_RANKS; ++r) out->p[r] = nullptr;
  for (int r = 0; r < world; ++r) {
    if (!peer_buffers[r]) return fail(VTR_ERR_INVALID_ARGUMENT, "exchange: peer %d is NULL", r);
    out->p[r] = static_cast<float*>(peer_buffers[r]);
  }
  return VTR_OK;
}
}  // namespace vtr

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
  const unsigned long long timeout_ns = vtr::kExchangeTimeoutNs;
  vtr::exchange_collect_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const float*>(my_buffer), n, world, seq, sums, timeout_ns);
  VTR_CUDA_CHECK(cudaGetLastError());
  return VTR_OK;
}
