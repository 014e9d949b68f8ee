// Host SIMT emulator — TEST INFRASTRUCTURE ONLY (never part of libgraspqp_b200.so).
//
// When a kernel source is compiled by g++ with -DGQP_EMU, this header supplies stand-ins for the CUDA built-ins the
// kernels use, so the *same kernel text* runs on the CPU: one std::thread per CUDA thread, one block at a time,
// __syncthreads() = a std::barrier over the block, warp shuffles/ballots = an exchange buffer + a 32-thread barrier.
// It is slow and only meant for tiny parity cases in `pytest -m "not gpu"` (tests/_emu/libgraspqp_b200_emu.so).
// Rules the kernels obey so that the emulation is faithful:
//   * blockDim.x is a multiple of 32, 1-D blocks and 1-D grids only;
//   * every lane of a warp executes every warp-collective (no early return before a shuffle / __syncwarp);
//   * every thread of a block executes every __syncthreads().
#pragma once
#ifdef GQP_EMU
#include <vector_types.h>
#include <vector_functions.h>

#include <atomic>
#include <barrier>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <memory>
#include <thread>
#include <vector>

namespace gqp_emu {
// Sense-reversing barrier that spins briefly and then yields: the emulated threads of a block outnumber the cores many
// times over, and a futex-based std::barrier pays a sleep / wake-up round trip per thread per warp collective (measured:
// 370-420 us per 32-thread barrier on a loaded 8-core host against 140 us for this one).
struct SpinBarrier {
    std::atomic<int> count{0}, gen{0};
    int n;
    explicit SpinBarrier(int n_) : n(n_) {}
    void arrive_and_wait() {
        const int g = gen.load(std::memory_order_acquire);
        if (count.fetch_add(1, std::memory_order_acq_rel) == n - 1) {
            count.store(0, std::memory_order_relaxed);
            gen.store(g + 1, std::memory_order_release);
        } else {
            int spins = 0;
            while (gen.load(std::memory_order_acquire) == g)
                if (++spins > 64) std::this_thread::yield();
        }
    }
};
struct WarpCtx {
    SpinBarrier bar;
    uint64_t xchg[32];
    explicit WarpCtx(int n) : bar(n) {}
};
inline thread_local uint3 t_idx, b_idx;
inline dim3 b_dim, g_dim;
inline SpinBarrier *block_bar = nullptr;
inline thread_local WarpCtx *warp = nullptr;
inline unsigned char *dyn_smem = nullptr;

template <class F>
void launch(dim3 grid, dim3 block, size_t smem, F body) {
    b_dim = block;
    g_dim = grid;
    std::vector<unsigned char> sm(smem + 64);
    dyn_smem = (unsigned char *)(((uintptr_t)sm.data() + 63) & ~(uintptr_t)63);
    int nt = (int)block.x, nw = (nt + 31) / 32;
    for (unsigned b = 0; b < grid.x; ++b) {
        SpinBarrier bb(nt);
        block_bar = &bb;
        std::vector<std::unique_ptr<WarpCtx>> warps;
        for (int w = 0; w < nw; ++w) warps.emplace_back(new WarpCtx(std::min(32, nt - 32 * w)));
        std::vector<std::thread> th;
        for (int t = 0; t < nt; ++t)
            th.emplace_back([&, t, b]() {
                t_idx = uint3{(unsigned)t, 0, 0};
                b_idx = uint3{b, 0, 0};
                warp = warps[t / 32].get();
                body();
            });
        for (auto &x : th) x.join();
    }
}
template <class T>
inline T shfl_any(T v, int src) {
    static_assert(sizeof(T) <= 8, "");
    int lane = t_idx.x & 31;
    uint64_t u = 0;
    std::memcpy(&u, &v, sizeof(T));
    warp->xchg[lane] = u;
    warp->bar.arrive_and_wait();
    uint64_t r = warp->xchg[src & 31];
    warp->bar.arrive_and_wait();
    T o;
    std::memcpy(&o, &r, sizeof(T));
    return o;
}
// ballot in ONE exchange (two barrier waits) instead of 32 shuffles: every lane posts its predicate, then reads all of them
inline unsigned ballot_any(int pred) {
    int lane = t_idx.x & 31;
    warp->xchg[lane] = pred ? 1u : 0u;
    warp->bar.arrive_and_wait();
    unsigned r = 0;
    const int n = warp->bar.n;  // lanes of a partial last warp beyond n do not exist: their bits stay 0
    for (int i = 0; i < n; ++i) r |= (warp->xchg[i] ? 1u : 0u) << i;
    warp->bar.arrive_and_wait();
    return r;
}
}  // namespace gqp_emu

#define threadIdx gqp_emu::t_idx
#define blockIdx gqp_emu::b_idx
#define blockDim gqp_emu::b_dim
#define gridDim gqp_emu::g_dim
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __shared__ static
#define __restrict__
#define __launch_bounds__(...)
#define GQP_DYN_SMEM(type, name) type *name = reinterpret_cast<type *>(gqp_emu::dyn_smem)
#define GQP_LAUNCH(kern, grid, block, smem, stream, ...) \
    gqp_emu::launch(dim3(grid), dim3(block), (smem), [&]() { kern(__VA_ARGS__); })

inline void __syncthreads() { gqp_emu::block_bar->arrive_and_wait(); }
inline void __syncwarp(unsigned = 0xffffffffu) { gqp_emu::warp->bar.arrive_and_wait(); }
template <class T>
inline T __shfl_sync(unsigned, T v, int src) { return gqp_emu::shfl_any(v, src); }
template <class T>
inline T __shfl_xor_sync(unsigned, T v, int m) { return gqp_emu::shfl_any(v, (int)(threadIdx.x & 31) ^ m); }
template <class T>
inline T __shfl_down_sync(unsigned, T v, int d) {
    int l = (threadIdx.x & 31) + d;
    return gqp_emu::shfl_any(v, l > 31 ? (int)(threadIdx.x & 31) : l);
}
template <class T>
inline T __shfl_up_sync(unsigned, T v, int d) {
    int l = (int)(threadIdx.x & 31) - d;
    return gqp_emu::shfl_any(v, l < 0 ? (int)(threadIdx.x & 31) : l);
}
inline unsigned __ballot_sync(unsigned, int pred) { return gqp_emu::ballot_any(pred); }
inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0; }
inline int __all_sync(unsigned m, int pred) { return __ballot_sync(m, pred) == 0xffffffffu; }
inline float atomicAdd(float *p, float v) {
    std::atomic_ref<float> a(*p);
    return a.fetch_add(v, std::memory_order_relaxed);
}
inline double atomicAdd(double *p, double v) {
    std::atomic_ref<double> a(*p);
    return a.fetch_add(v, std::memory_order_relaxed);
}
inline int atomicAdd(int *p, int v) {
    std::atomic_ref<int> a(*p);
    return a.fetch_add(v, std::memory_order_relaxed);
}
inline unsigned atomicAdd(unsigned *p, unsigned v) {
    std::atomic_ref<unsigned> a(*p);
    return a.fetch_add(v, std::memory_order_relaxed);
}
inline unsigned long long atomicAdd(unsigned long long *p, unsigned long long v) {
    std::atomic_ref<unsigned long long> a(*p);
    return a.fetch_add(v, std::memory_order_relaxed);
}
inline long long __double2ll_rn(double a) { return std::llrint(a); }
inline unsigned atomicMax(unsigned *p, unsigned v) {
    std::atomic_ref<unsigned> a(*p);
    unsigned old = a.load(std::memory_order_relaxed);
    while (old < v && !a.compare_exchange_weak(old, v, std::memory_order_relaxed)) {}
    return old;
}
inline unsigned atomicOr(unsigned *p, unsigned v) {
    std::atomic_ref<unsigned> a(*p);
    return a.fetch_or(v, std::memory_order_relaxed);
}
template <class T>
inline T __ldg(const T *p) { return *p; }
inline float __fmaf_rn(float a, float b, float c) { return std::fmaf(a, b, c); }
inline float __fmul_rn(float a, float b) { volatile float r = a * b; return r; }
inline float __fadd_rn(float a, float b) { volatile float r = a + b; return r; }
inline float __fsub_rn(float a, float b) { volatile float r = a - b; return r; }
inline float __fdiv_rn(float a, float b) { volatile float r = a / b; return r; }
inline float __fsqrt_rn(float a) { return std::sqrt(a); }
inline float rsqrtf(float a) { return 1.0f / std::sqrt(a); }
inline double rsqrt(double a) { return 1.0 / std::sqrt(a); }
inline float __expf(float a) { return std::exp(a); }
inline float __int_as_float(int i) { float f; std::memcpy(&f, &i, 4); return f; }
inline int __float_as_int(float f) { int i; std::memcpy(&i, &f, 4); return i; }
inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((uint64_t)a * b) >> 32); }
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __ffs(int v) { return __builtin_ffs(v); }
inline float fminf_(float a, float b) { return a < b ? a : b; }
using std::fmaxf;
using std::fminf;
using std::max;
using std::min;
#else  // real CUDA
#include <cuda_runtime.h>
#define GQP_DYN_SMEM(type, name)                              \
    extern __shared__ __align__(128) unsigned char name##_raw[]; \
    type *name = reinterpret_cast<type *>(name##_raw)
#define GQP_LAUNCH(kern, grid, block, smem, stream, ...) kern<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#endif
