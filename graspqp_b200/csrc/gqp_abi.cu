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

extern "C" int gqp_abi_version(void) { return 1; }
extern "C" const char *gqp_last_error(void) { return g_err; }
extern "C" uint64_t gqp_launch_count(void) { return g_launches.load(); }

#ifndef GQP_EMU
// FP32 FMA probe: 8 independent dependent-chains per lane, 2 flop per FMA.
__global__ void __launch_bounds__(256) fma_probe_kernel(int iters, float *sink) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6,
          a7 = a0 + 7;
    const float m = 0.999f, c = 1e-3f;
    for (int i = 0; i < iters; ++i) {
        a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
        a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
    }
    float s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (s == 123.456f) sink[0] = s;
}
extern "C" int gqp_fma_peak_probe(int iters, float *sink, uint64_t *n_fma_host, void *stream) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int grid = sms * 8;
    fma_probe_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(iters, sink);
    GQP_CHECK_LAUNCH("gqp_fma_peak_probe");
    if (n_fma_host) *n_fma_host = (uint64_t)grid * 256ull * 8ull * (uint64_t)iters;
    return GQP_OK;
}
#else
extern "C" int gqp_fma_peak_probe(int, float *, uint64_t *, void *) { return gqp_set_error(GQP_ERR_ARG, "emu"); }
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
