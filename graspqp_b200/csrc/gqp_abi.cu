// Error plumbing, version and launch counter of the C ABI (include/graspqp_b200.h).
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "gqp_common.cuh"
#include "graspqp_b200.h"

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

int gqp_set_error(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
void gqp_count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

extern "C" int gqp_abi_version(void) { return 2; }  // 2: after_reset argument of gqp_mala_accept, (f3) entry points
extern "C" const char *gqp_last_error(void) { return g_err; }
extern "C" uint64_t gqp_launch_count(void) { return g_launches.load(); }

#ifndef GQP_EMU
// FMA probes: 8 independent dependent-chains per lane, 16 rounds per loop trip (128 FMAs per branch: the loop overhead
// of the round-1 probe cost 15 % of the issue slots), 2 flop per FMA.
template <typename T>
__global__ void __launch_bounds__(256) fma_probe_kernel(int trips, T *sink) {
    T a0 = (T)threadIdx.x * (T)1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6,
      a7 = a0 + 7;
    const T m = (T)0.999, c = (T)1e-3;
    for (int i = 0; i < trips; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
            a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
        }
    }
    T s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (s == (T)123.456) sink[0] = s;
}
template <typename T>
static int fma_probe(int iters, T *sink, uint64_t *n_fma_host, void *stream) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = sms * 8, trips = (iters + 15) / 16;
    fma_probe_kernel<T><<<grid, 256, 0, (cudaStream_t)stream>>>(trips, sink);
    GQP_CHECK_LAUNCH("gqp_fma_peak_probe");
    if (n_fma_host) *n_fma_host = (uint64_t)grid * 256ull * 8ull * 16ull * (uint64_t)trips;
    return GQP_OK;
}
extern "C" int gqp_fma_peak_probe(int iters, float *sink, uint64_t *n_fma_host, void *stream) {
    return fma_probe<float>(iters, sink, n_fma_host, stream);
}
extern "C" int gqp_dfma_peak_probe(int iters, double *sink, uint64_t *n_fma_host, void *stream) {
    return fma_probe<double>(iters, sink, n_fma_host, stream);
}
#else
extern "C" int gqp_fma_peak_probe(int, float *, uint64_t *, void *) { return gqp_set_error(GQP_ERR_ARG, "emu"); }
extern "C" int gqp_dfma_peak_probe(int, double *, uint64_t *, void *) { return gqp_set_error(GQP_ERR_ARG, "emu"); }
#endif

// zero a device buffer on the stream (the per-step reset of the link-transform gradient)
extern "C" int gqp_memset_zero(void *ptr, uint64_t bytes, void *stream) {
    if (!ptr && bytes) return gqp_set_error(GQP_ERR_ARG, "gqp_memset_zero: null pointer");
#ifndef GQP_EMU
    cudaError_t e = cudaMemsetAsync(ptr, 0, bytes, (cudaStream_t)stream);
    if (e != cudaSuccess) return gqp_set_error(GQP_ERR_CUDA, "gqp_memset_zero: %s", cudaGetErrorString(e));
#else
    std::memset(ptr, 0, bytes);
#endif
    return GQP_OK;
}
