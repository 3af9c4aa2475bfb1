// Launchers of the peer-memory exchange kernels (peer.cu).
#pragma once
#include "kernels.cuh"

namespace nmfb {

constexpr int kMaxPeers = 16;

cudaError_t peer_push_launch(const void* src, long src_pitch, void* const* peers_host, int world, long dst_off,
                             long dst_pitch, int rows, long row_bytes, cudaStream_t stream);
cudaError_t peer_scatter_launch(const void* src, long src_pitch, long src_peer_stride, void* const* peers_host,
                                int world, long dst_off, long dst_pitch, int rows, const long* row_bytes_host,
                                cudaStream_t stream);
cudaError_t peer_push_mc_launch(const void* src, long src_pitch, void* mc_base, long dst_off, long dst_pitch,
                                int rows, long row_bytes, cudaStream_t stream);
cudaError_t peer_barrier_launch(void* const* flags_host, int rank, int world, unsigned int seq, cudaStream_t stream);
cudaError_t sum_rank_slots_launch(const double* slots, int world, long stride, int n, double* out64, float* out32,
                                  cudaStream_t stream);

}  // namespace nmfb
