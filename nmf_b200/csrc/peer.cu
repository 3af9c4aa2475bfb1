// Exchange step of the sharded runs (MPI-FAUN 1-D layout) over NVLink / NVSwitch peer memory.
//
// Every rank owns a window of each factor.  After its update kernel it STORES that window (fp32 master rows)
// straight into an exchange buffer of every peer -- plain st.global to the peers' device pointers, which the
// caller obtained from symmetric memory (torch.distributed._symmetric_memory) or CUDA IPC --, together with
// its partial of the k x k Gram matrix and its partial objective; one cross-GPU barrier later every rank owns
// all windows and reduces the Gram partials in rank order (so that all ranks hold bit-identical results).
// No staging copies, no collective library call on the data path.
//
// Ordering.  push(t) -> barrier(t) -> consume(t).  The barrier is a kernel on the same stream: it starts
// after the local pushes have completed (kernel boundary), publishes `seq` to every peer with a system-scope
// release store and spins until every peer has published >= seq.  Buffers are double-buffered by the caller
// (parity of an exchange counter): a peer's push(t+1) into buffer (t+1)&1 can then never overtake this rank's
// consume(t) of buffer t&1, and its push(t+2) is ordered behind barrier(t+1), which this rank enters only after
// consume(t) in stream order.
//
// nmfb_exchange_factor (capi.cu) sequences one exchange: window store on the caller's stream, the rank's Gram term on
// a side stream under it (joined by events before the barrier), barrier, adoption, sum of the Gram terms.
#include "elementwise.cuh"
#include "kernels.cuh"
#include "peer.cuh"

namespace nmfb {

namespace {

struct PeerPtrs {
    void* p[kMaxPeers];
};

// rows x row_bytes (a multiple of 16) from src (pitch src_pitch bytes) to dst_off bytes into every peer's
// buffer (pitch dst_pitch).  grid = (blocks, world): blockIdx.y is the destination, so that the stores of one
// CTA stream over one NVLink route.
__global__ void __launch_bounds__(256)
peer_push_kernel(const char* __restrict__ src, long src_pitch, PeerPtrs dst, long dst_off, long dst_pitch,
                 int rows, long row_bytes) {
    char* out = static_cast<char*>(dst.p[blockIdx.y]) + dst_off;
    const long vec_per_row = row_bytes >> 4;
    const long total = vec_per_row * rows;
    for (long i = static_cast<long>(blockIdx.x) * 256 + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * 256) {
        const long r = i / vec_per_row, c = i - r * vec_per_row;
        const uint4 v = *reinterpret_cast<const uint4*>(src + r * src_pitch + (c << 4));
        *reinterpret_cast<uint4*>(out + r * dst_pitch + (c << 4)) = v;
    }
}

// All-to-all of row blocks: destination q receives rows x row_bytes[q] bytes taken from src + q * src_peer_stride
// (the columns of this rank's rows that belong to rank q), at dst_off of its buffer.  VEC = bytes per access.
struct ScatterWidths {
    long row_bytes[kMaxPeers];
};
template <int VEC>
__global__ void __launch_bounds__(256)
peer_scatter_kernel(const char* __restrict__ src, long src_pitch, long src_peer_stride, PeerPtrs dst, long dst_off,
                    long dst_pitch, int rows, ScatterWidths w) {
    const int q = blockIdx.y;
    const char* in = src + q * src_peer_stride;
    char* out = static_cast<char*>(dst.p[q]) + dst_off;
    const long per_row = w.row_bytes[q] / VEC;
    const long total = per_row * rows;
    for (long i = static_cast<long>(blockIdx.x) * 256 + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * 256) {
        const long r = i / per_row, c = i - r * per_row;
        if (VEC == 16)
            *reinterpret_cast<uint4*>(out + r * dst_pitch + c * 16) =
                *reinterpret_cast<const uint4*>(in + r * src_pitch + c * 16);
        else
            *reinterpret_cast<unsigned int*>(out + r * dst_pitch + c * 4) =
                *reinterpret_cast<const unsigned int*>(in + r * src_pitch + c * 4);
    }
}

// The same copy through the NVSwitch multicast address of the arena (multimem.st): ONE store per 16 bytes leaves
// this GPU and the switch replicates it into every rank's arena (this rank's included), instead of `world` unicast
// stores -- 1/world of the NVLink egress.
__global__ void __launch_bounds__(256)
peer_push_mc_kernel(const char* __restrict__ src, long src_pitch, char* __restrict__ mc, long dst_pitch, int rows,
                    long row_bytes) {
    const long vec_per_row = row_bytes >> 4;
    const long total = vec_per_row * rows;
    for (long i = static_cast<long>(blockIdx.x) * 256 + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * 256) {
        const long r = i / vec_per_row, c = i - r * vec_per_row;
        const float4 v = *reinterpret_cast<const float4*>(src + r * src_pitch + (c << 4));
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                     ::"l"(mc + r * dst_pitch + (c << 4)), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    }
}

// flags: [world] uint32 on every rank (symmetric).  Thread p publishes seq into peer p's flags[rank] and
// waits for flags[p] >= seq on this rank.
__global__ void __launch_bounds__(kMaxPeers)
peer_barrier_kernel(PeerPtrs flags, int rank, int world, unsigned int seq) {
    const int p = threadIdx.x;
    if (p >= world) return;
    asm volatile("fence.proxy.alias;" ::: "memory");       // earlier multimem.st went through the multicast alias
    __threadfence_system();
    unsigned int* theirs = static_cast<unsigned int*>(flags.p[p]) + rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(seq) : "memory");
    const unsigned int* mine = static_cast<const unsigned int*>(flags.p[rank]) + p;
    unsigned int seen;
    do {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(mine) : "memory");
    } while (static_cast<int>(seen - seq) < 0);
    __threadfence_system();
    asm volatile("fence.proxy.alias;" ::: "memory");
}

// out[i] = sum over ranks (fixed order) of slots[p * stride + i]; optional fp32 copy.
__global__ void __launch_bounds__(256)
sum_rank_slots_kernel(const double* __restrict__ slots, int world, long stride, int n, double* __restrict__ out64,
                      float* __restrict__ out32) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    double acc = 0.0;
    for (int p = 0; p < world; ++p) acc += slots[p * stride + i];
    if (out64) out64[i] = acc;
    if (out32) out32[i] = static_cast<float>(acc);
}

}  // namespace

cudaError_t peer_push_launch(const void* src, long src_pitch, void* const* peers_host, int world, long dst_off,
                             long dst_pitch, int rows, long row_bytes, cudaStream_t stream) {
    if (world < 1 || world > kMaxPeers || (row_bytes & 15) || rows < 1) return cudaErrorInvalidValue;
    PeerPtrs d;
    for (int p = 0; p < kMaxPeers; ++p) d.p[p] = p < world ? peers_host[p] : nullptr;
    const long total = (row_bytes >> 4) * rows;
    long blocks = (total + 255) / 256;
    if (blocks > 592 / world) blocks = 592 / world;   // ~4 CTAs per SM over all destinations
    if (blocks < 1) blocks = 1;
    peer_push_kernel<<<dim3(static_cast<unsigned>(blocks), world), 256, 0, stream>>>(
        static_cast<const char*>(src), src_pitch, d, dst_off, dst_pitch, rows, row_bytes);
    count_launch();
    return cudaGetLastError();
}

cudaError_t peer_scatter_launch(const void* src, long src_pitch, long src_peer_stride, void* const* peers_host,
                                int world, long dst_off, long dst_pitch, int rows, const long* row_bytes_host,
                                cudaStream_t stream) {
    if (world < 1 || world > kMaxPeers || rows < 1) return cudaErrorInvalidValue;
    PeerPtrs d;
    ScatterWidths w;
    bool vec16 = ((src_pitch | src_peer_stride | dst_off | dst_pitch) & 15) == 0 &&
                 (reinterpret_cast<uintptr_t>(src) & 15) == 0;
    for (int p = 0; p < kMaxPeers; ++p) {
        d.p[p] = p < world ? peers_host[p] : nullptr;
        w.row_bytes[p] = p < world ? row_bytes_host[p] : 0;
        if (w.row_bytes[p] & 3) return cudaErrorInvalidValue;
        if (w.row_bytes[p] & 15) vec16 = false;
    }
    const dim3 grid(148, world);
    if (vec16)
        peer_scatter_kernel<16><<<grid, 256, 0, stream>>>(static_cast<const char*>(src), src_pitch, src_peer_stride, d,
                                                         dst_off, dst_pitch, rows, w);
    else
        peer_scatter_kernel<4><<<grid, 256, 0, stream>>>(static_cast<const char*>(src), src_pitch, src_peer_stride, d,
                                                        dst_off, dst_pitch, rows, w);
    count_launch();
    return cudaGetLastError();
}

cudaError_t peer_push_mc_launch(const void* src, long src_pitch, void* mc_base, long dst_off, long dst_pitch,
                                int rows, long row_bytes, cudaStream_t stream) {
    if (!mc_base || (row_bytes & 15) || rows < 1) return cudaErrorInvalidValue;
    const long total = (row_bytes >> 4) * rows;
    long blocks = (total + 255) / 256;
    if (blocks > 592) blocks = 592;               // 4 CTAs per SM: the copy is a load-latency chain per thread
    if (blocks < 1) blocks = 1;
    peer_push_mc_kernel<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(
        static_cast<const char*>(src), src_pitch, static_cast<char*>(mc_base) + dst_off, dst_pitch, rows, row_bytes);
    count_launch();
    return cudaGetLastError();
}

cudaError_t peer_barrier_launch(void* const* flags_host, int rank, int world, unsigned int seq, cudaStream_t stream) {
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world) return cudaErrorInvalidValue;
    PeerPtrs f;
    for (int p = 0; p < kMaxPeers; ++p) f.p[p] = p < world ? flags_host[p] : nullptr;
    peer_barrier_kernel<<<1, kMaxPeers, 0, stream>>>(f, rank, world, seq);
    count_launch();
    return cudaGetLastError();
}

cudaError_t sum_rank_slots_launch(const double* slots, int world, long stride, int n, double* out64, float* out32,
                                  cudaStream_t stream) {
    sum_rank_slots_kernel<<<(n + 255) / 256, 256, 0, stream>>>(slots, world, stride, n, out64, out32);
    count_launch();
    return cudaGetLastError();
}

}  // namespace nmfb
