// Host emulation shim: lets the kernel source `mantrap_b200/csrc/mtb_kernels.cu` compile with g++ and run every
// CUDA thread of a block as an OS thread (barriers for __syncthreads / warp shuffles, CAS loops for atomics).
//
// TEST INFRASTRUCTURE ONLY (lives under tests/). It exists because the build container has no GPU: the CPU test
// suite runs the *actual kernel code* against the golden vectors before any GPU time is spent. It is never built
// into, loaded by, or reachable from the product package; the product path fails loudly without its CUDA library.
#pragma once
#include <algorithm>
#include <atomic>
#include <barrier>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#define __global__
#define __device__
#define __align__(n) __attribute__((aligned(n)))
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __restrict__

struct float2 { float x, y; };
struct alignas(8) uint2 { unsigned x, y; };
static inline uint2 make_uint2(unsigned x, unsigned y) { uint2 r; r.x = x; r.y = y; return r; }
struct alignas(16) float4 { float x, y, z, w; };
static inline float2 make_float2(float x, float y) { float2 r; r.x = x; r.y = y; return r; }
static inline float4 make_float4(float x, float y, float z, float w) { float4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
struct dim3 { unsigned x = 1, y = 1, z = 1; };

typedef void* cudaStream_t;
typedef int cudaError_t;
#define cudaSuccess 0

namespace emul {
struct Warp {
    std::barrier<> bar{32};
    uint32_t buf[32];
};
struct Block {
    explicit Block(int threads, size_t smem_bytes)
        : bar(threads), warps((threads + 31) / 32), smem(static_cast<char*>(std::aligned_alloc(128, ((smem_bytes + 127) / 128 + 1) * 128))) {}
    ~Block() { std::free(smem); }
    std::barrier<> bar;
    std::vector<Warp> warps;
    char* smem;
    // named barriers (bar.sync id, count), created on first use
    std::mutex named_mutex;
    std::unique_ptr<std::barrier<>> named[16];
    std::barrier<>& named_barrier(int id, int count) {
        std::lock_guard<std::mutex> lock(named_mutex);
        if (!named[id]) named[id] = std::make_unique<std::barrier<>>(count);
        return *named[id];
    }
};
inline thread_local Block* tl_block = nullptr;
}  // namespace emul

inline thread_local dim3 threadIdx, blockIdx, blockDim, gridDim;

static inline void __syncthreads() { emul::tl_block->bar.arrive_and_wait(); }
static inline emul::Warp& emul_warp() { return emul::tl_block->warps[threadIdx.x >> 5]; }
static inline void __syncwarp(unsigned = 0xffffffffu) { emul_warp().bar.arrive_and_wait(); }
#ifdef MTB_EMUL_BREAK_GROUP_BARRIER   // self-test of tools/tsan_emulation.py: without the group barrier ThreadSanitizer must report races
static inline void mtb_bar_sync(int, int) {}
#else
static inline void mtb_bar_sync(int id, int count) { emul::tl_block->named_barrier(id, count).arrive_and_wait(); }
#endif
static inline void mtb_bar_arrive(int id, int count) { (void)emul::tl_block->named_barrier(id, count).arrive(); }
template <int ID> static inline void mtb_bar_sync_id(int count) { mtb_bar_sync(ID, count); }
template <int ID> static inline void mtb_bar_arrive_id(int count) { mtb_bar_arrive(ID, count); }

template <class T>
static inline T emul_exchange(T v, int src_lane) {
    static_assert(sizeof(T) == 4, "32-bit shuffles only");
    emul::Warp& w = emul_warp();
    const int lane = threadIdx.x & 31;
    std::memcpy(&w.buf[lane], &v, 4);
    w.bar.arrive_and_wait();
    T r;
    std::memcpy(&r, &w.buf[src_lane & 31], 4);
    w.bar.arrive_and_wait();
    return r;
}
template <class T>
static inline T __shfl_xor_sync(unsigned, T v, int lane_mask) { return emul_exchange(v, (threadIdx.x & 31) ^ lane_mask); }
template <class T>
static inline T __shfl_sync(unsigned, T v, int src_lane) { return emul_exchange(v, src_lane); }
template <class T>
static inline T __shfl_down_sync(unsigned, T v, int delta) {
    const int lane = threadIdx.x & 31;
    return emul_exchange(v, lane + delta < 32 ? lane + delta : lane);
}
static inline unsigned __ballot_sync(unsigned, int pred) {
    emul::Warp& w = emul_warp();
    const int lane = threadIdx.x & 31;
    w.buf[lane] = pred ? 1u : 0u;
    w.bar.arrive_and_wait();
    unsigned r = 0;
    for (int l = 0; l < 32; ++l) r |= (w.buf[l] << l);
    w.bar.arrive_and_wait();
    return r;
}
static inline int __any_sync(unsigned mask, int pred) { return __ballot_sync(mask, pred) != 0u; }

static inline float atomicAdd(float* addr, float val) {
    uint32_t* a = reinterpret_cast<uint32_t*>(addr);
    uint32_t old_bits = __atomic_load_n(a, __ATOMIC_RELAXED);
    for (;;) {
        float old_f;
        std::memcpy(&old_f, &old_bits, 4);
        const float new_f = old_f + val;
        uint32_t new_bits;
        std::memcpy(&new_bits, &new_f, 4);
        if (__atomic_compare_exchange_n(a, &old_bits, new_bits, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) return old_f;
    }
}
static inline unsigned atomicOr(unsigned* addr, unsigned val) { return __atomic_fetch_or(addr, val, __ATOMIC_RELAXED); }
static inline unsigned atomicAdd(unsigned* addr, unsigned val) { return __atomic_fetch_add(addr, val, __ATOMIC_RELAXED); }
static inline unsigned long long atomicAdd(unsigned long long* addr, unsigned long long val) {
    return __atomic_fetch_add(addr, val, __ATOMIC_RELAXED);
}
static inline unsigned long long atomicCAS(unsigned long long* addr, unsigned long long compare, unsigned long long val) {
    __atomic_compare_exchange_n(addr, &compare, val, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED);
    return compare;
}
static inline unsigned long long atomicMin(unsigned long long* addr, unsigned long long val) {
    unsigned long long old = __atomic_load_n(addr, __ATOMIC_RELAXED);
    while (val < old && !__atomic_compare_exchange_n(addr, &old, val, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
    return old;
}
template <class T>
static inline T __ldg(const T* p) { return *p; }
static inline void mtb_red_add_f4(float4* addr, float a, float b, float c, float d) {
    float* f = reinterpret_cast<float*>(addr);
    atomicAdd(f, a); atomicAdd(f + 1, b); atomicAdd(f + 2, c); atomicAdd(f + 3, d);
}
static inline float4 mtb_ldcg_f4(const float4* addr) { return *addr; }
static inline void mtb_stcg_f4(float4* addr, float4 v) { *addr = v; }
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline float rsqrtf(float x) { return 1.0f / sqrtf(x); }
static inline float __fdividef(float a, float b) { return a / b; }
static inline unsigned __float_as_uint(float f) { unsigned u; std::memcpy(&u, &f, 4); return u; }
static inline float __uint_as_float(unsigned u) { float f; std::memcpy(&f, &u, 4); return f; }

namespace emul {
template <class Kernel, class... Args>
void launch(Kernel kernel, int grid, int block, size_t smem_bytes, Args... args) {
    if (block % 32 != 0) { std::fprintf(stderr, "emul: block size must be a multiple of 32\n"); std::abort(); }
    for (int b = 0; b < grid; ++b) {
        Block blk(block, smem_bytes);
        std::vector<std::thread> threads;
        threads.reserve(block);
        for (int t = 0; t < block; ++t) {
            threads.emplace_back([&, t]() {
                tl_block = &blk;
                threadIdx.x = t;
                blockIdx.x = b;
                blockDim.x = block;
                gridDim.x = grid;
                kernel(args...);
            });
        }
        for (auto& th : threads) th.join();
    }
}
}  // namespace emul

#define MTB_SHARED_FLOAT4(name) float4* name = reinterpret_cast<float4*>(emul::tl_block->smem)
#define MTB_LAUNCH(kernel, grid, block, smem, stream, ...) emul::launch(kernel, grid, block, smem, __VA_ARGS__)
