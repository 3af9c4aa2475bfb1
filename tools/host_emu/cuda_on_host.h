// Runs CUDA-core kernels of libnmf_b200 on host threads, one std::thread per CUDA thread, one block at a
// time: __syncthreads is a std::barrier over the block, __shfl_xor_sync an exchange inside the 32-lane warp,
// __shared__ variables are function-level statics (blocks run sequentially), threadIdx / blockIdx are
// thread_local.  Test infrastructure only (tests/test_host_emulation.py): it checks the indexing, the
// synchronisation structure and the arithmetic of kernels that have no tcgen05 / TMA / mbarrier code, on the
// CPU, against the oracle.  It says nothing about performance and is never part of the product.
#pragma once
#include <cuda_bf16.h>      // __nv_bfloat16 with host conversions; pulls in the vector types (float4, double2, ...)

#include <atomic>
#include <barrier>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <memory>
#include <thread>
#include <vector>

#undef __global__
#undef __device__
#undef __host__
#undef __shared__
#undef __launch_bounds__
#define __global__
#define __device__
#define __host__
#define __shared__ static
#define __launch_bounds__(...)

namespace emu {

struct Idx { unsigned x, y, z; };

struct Block {
    std::barrier<> all;
    std::vector<std::unique_ptr<std::barrier<>>> warp;
    std::vector<std::uint64_t> xchg;                 // one 8-byte slot per thread
    std::vector<unsigned char> dyn_smem;
    explicit Block(int nthreads, size_t smem) : all(nthreads), xchg(nthreads), dyn_smem(smem + 16) {
        for (int w = 0; w < (nthreads + 31) / 32; ++w) {
            const int lanes = std::min(32, nthreads - w * 32);
            warp.emplace_back(std::make_unique<std::barrier<>>(lanes));
        }
    }
};

inline thread_local Block* tl_block = nullptr;
inline thread_local int tl_tid = 0;

template <typename T>
inline T* dyn_smem() {
    auto p = reinterpret_cast<std::uintptr_t>(tl_block->dyn_smem.data());
    return reinterpret_cast<T*>((p + 15) & ~static_cast<std::uintptr_t>(15));
}

}  // namespace emu

inline thread_local emu::Idx threadIdx{0, 0, 0}, blockIdx{0, 0, 0};
inline emu::Idx blockDim{1, 1, 1}, gridDim{1, 1, 1};

inline void __syncthreads() { emu::tl_block->all.arrive_and_wait(); }

template <typename T>
inline T __shfl_xor_sync(unsigned, T v, int lane_mask) {
    static_assert(sizeof(T) <= 8, "shuffle of at most 8 bytes");
    emu::Block& b = *emu::tl_block;
    const int tid = emu::tl_tid, w = tid >> 5;
    std::uint64_t bits = 0;
    std::memcpy(&bits, &v, sizeof(T));
    b.xchg[tid] = bits;
    b.warp[w]->arrive_and_wait();
    const std::uint64_t got = b.xchg[(tid & ~31) | ((tid ^ lane_mask) & 31)];
    b.warp[w]->arrive_and_wait();                    // nobody overwrites a slot before it has been read
    T out;
    std::memcpy(&out, &got, sizeof(T));
    return out;
}

template <typename T>
inline T __shfl_sync(unsigned, T v, int src_lane) {
    static_assert(sizeof(T) <= 8, "shuffle of at most 8 bytes");
    emu::Block& b = *emu::tl_block;
    const int tid = emu::tl_tid, w = tid >> 5;
    std::uint64_t bits = 0;
    std::memcpy(&bits, &v, sizeof(T));
    b.xchg[tid] = bits;
    b.warp[w]->arrive_and_wait();
    const std::uint64_t got = b.xchg[(tid & ~31) | (src_lane & 31)];
    b.warp[w]->arrive_and_wait();
    T out;
    std::memcpy(&out, &got, sizeof(T));
    return out;
}

// mma.sync.m8n8k4.f64 (ptx.cuh: dmma_884) on the host: every lane publishes its A and B element, then computes its
// two D elements from the fragments' documented lane layout.
namespace nmfb {
inline void dmma_884(double (&d)[2], double a, double b) {
    emu::Block& blk = *emu::tl_block;
    const int tid = emu::tl_tid, w = tid >> 5, lane = tid & 31, base = tid & ~31;
    double av[32], bv[32];
    std::uint64_t bits;
    std::memcpy(&bits, &a, 8);
    blk.xchg[tid] = bits;
    blk.warp[w]->arrive_and_wait();
    for (int l = 0; l < 32; ++l) std::memcpy(&av[l], &blk.xchg[base + l], 8);
    blk.warp[w]->arrive_and_wait();
    std::memcpy(&bits, &b, 8);
    blk.xchg[tid] = bits;
    blk.warp[w]->arrive_and_wait();
    for (int l = 0; l < 32; ++l) std::memcpy(&bv[l], &blk.xchg[base + l], 8);
    blk.warp[w]->arrive_and_wait();
    const int m = lane >> 2;
    for (int j = 0; j < 2; ++j) {
        const int n = 2 * (lane & 3) + j;
        for (int k = 0; k < 4; ++k) d[j] = std::fma(av[m * 4 + k], bv[n * 4 + k], d[j]);   // A[m][k], B[k][n]
    }
}
}  // namespace nmfb

inline unsigned atomicMax(unsigned* a, unsigned v) {
    unsigned old = __atomic_load_n(a, __ATOMIC_RELAXED);
    while (old < v && !__atomic_compare_exchange_n(a, &old, v, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
    return old;
}
inline unsigned atomicMin(unsigned* a, unsigned v) {
    unsigned old = __atomic_load_n(a, __ATOMIC_RELAXED);
    while (old > v && !__atomic_compare_exchange_n(a, &old, v, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
    return old;
}
inline unsigned atomicOr(unsigned* a, unsigned v) { return __atomic_fetch_or(a, v, __ATOMIC_RELAXED); }

inline unsigned __float_as_uint(float f) { unsigned u; std::memcpy(&u, &f, 4); return u; }
inline float __uint_as_float(unsigned u) { float f; std::memcpy(&f, &u, 4); return f; }
inline int lane_id() { return emu::tl_tid & 31; }        // ptx.cuh reads %laneid
using std::isfinite;
using std::max;
using std::min;

namespace emu {

// grid / block as (x, y, z); `body` is the kernel call.  A thread that returns from the kernel drops out of
// the barriers, like an exited CUDA thread.
inline void launch(Idx grid, Idx block, size_t smem_bytes, const std::function<void()>& body) {
    const int nthreads = static_cast<int>(block.x * block.y * block.z);
    ::blockDim = block;
    ::gridDim = grid;
    for (unsigned bz = 0; bz < grid.z; ++bz)
        for (unsigned by = 0; by < grid.y; ++by)
            for (unsigned bx = 0; bx < grid.x; ++bx) {
                Block blk(nthreads, smem_bytes);
                std::vector<std::thread> pool;
                pool.reserve(nthreads);
                for (int t = 0; t < nthreads; ++t)
                    pool.emplace_back([&, t] {
                        tl_block = &blk;
                        tl_tid = t;
                        ::threadIdx = Idx{t % block.x, (t / block.x) % block.y, t / (block.x * block.y)};
                        ::blockIdx = Idx{bx, by, bz};
                        body();
                        blk.warp[t >> 5]->arrive_and_drop();
                        blk.all.arrive_and_drop();
                    });
                for (auto& th : pool) th.join();
            }
}

}  // namespace emu
