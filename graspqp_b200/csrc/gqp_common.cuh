// Shared device helpers: small vector math, reductions, error plumbing.  Compiled by nvcc for sm_100a, or by g++
// with -DGQP_EMU for the host SIMT emulator (tests only, see gqp_emu.h).
#pragma once
#include "gqp_emu.h"

#include <stdint.h>

#define GQP_WARP 32
#define GQP_FULL 0xffffffffu

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 v3(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross3(V3 a, V3 b) {
    return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ V3 ld3(const float *p) { return V3{p[0], p[1], p[2]}; }
__device__ __forceinline__ void st3(float *p, V3 a) { p[0] = a.x; p[1] = a.y; p[2] = a.z; }

// Row-major 3x4 rigid transform [R|t].
struct T34 {
    float m[12];
};
__device__ __forceinline__ T34 t34_load(const float *p) {
    T34 r;
#pragma unroll
    for (int i = 0; i < 12; ++i) r.m[i] = p[i];
    return r;
}
// the same from a 16-byte aligned address: three 128-bit loads
__device__ __forceinline__ T34 t34_load4(const float *p) {
    const float4 *q = reinterpret_cast<const float4 *>(p);
    const float4 a = q[0], b = q[1], c = q[2];
    T34 r;
    r.m[0] = a.x; r.m[1] = a.y; r.m[2] = a.z; r.m[3] = a.w;
    r.m[4] = b.x; r.m[5] = b.y; r.m[6] = b.z; r.m[7] = b.w;
    r.m[8] = c.x; r.m[9] = c.y; r.m[10] = c.z; r.m[11] = c.w;
    return r;
}
__device__ __forceinline__ T34 t34_identity() {
    T34 r;
#pragma unroll
    for (int i = 0; i < 12; ++i) r.m[i] = 0.f;
    r.m[0] = r.m[5] = r.m[10] = 1.f;
    return r;
}
// c = a * b (both rigid 3x4, implicit last row 0 0 0 1)
__device__ __forceinline__ T34 t34_mul(const T34 &a, const T34 &b) {
    T34 c;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float s = a.m[4 * i + 0] * b.m[j] + a.m[4 * i + 1] * b.m[4 + j] + a.m[4 * i + 2] * b.m[8 + j];
            if (j == 3) s += a.m[4 * i + 3];
            c.m[4 * i + j] = s;
        }
    }
    return c;
}
__device__ __forceinline__ V3 t34_point(const T34 &a, V3 p) {
    return V3{a.m[0] * p.x + a.m[1] * p.y + a.m[2] * p.z + a.m[3], a.m[4] * p.x + a.m[5] * p.y + a.m[6] * p.z + a.m[7],
              a.m[8] * p.x + a.m[9] * p.y + a.m[10] * p.z + a.m[11]};
}
__device__ __forceinline__ V3 t34_vec(const T34 &a, V3 p) {
    return V3{a.m[0] * p.x + a.m[1] * p.y + a.m[2] * p.z, a.m[4] * p.x + a.m[5] * p.y + a.m[6] * p.z,
              a.m[8] * p.x + a.m[9] * p.y + a.m[10] * p.z};
}
// R^T (p - t): world -> local
__device__ __forceinline__ V3 t34_inv_point(const T34 &a, V3 p) {
    V3 q = V3{p.x - a.m[3], p.y - a.m[7], p.z - a.m[11]};
    return V3{a.m[0] * q.x + a.m[4] * q.y + a.m[8] * q.z, a.m[1] * q.x + a.m[5] * q.y + a.m[9] * q.z,
              a.m[2] * q.x + a.m[6] * q.y + a.m[10] * q.z};
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(GQP_FULL, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(GQP_FULL, v, o);
    return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(GQP_FULL, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(GQP_FULL, v, o));
    return v;
}

// ---- host-side error plumbing (defined in gqp_abi.cu) ----
int gqp_set_error(int code, const char *fmt, ...);
void gqp_count_launch(int n = 1);
#ifndef GQP_EMU
#define GQP_CHECK_LAUNCH(name)                                                                   \
    do {                                                                                         \
        cudaError_t e_ = cudaGetLastError();                                                     \
        if (e_ != cudaSuccess) return gqp_set_error(-2, "%s: %s", name, cudaGetErrorString(e_)); \
        gqp_count_launch();                                                                      \
    } while (0)
#else
#define GQP_CHECK_LAUNCH(name) gqp_count_launch()
#endif

// ---- "device" memory for the immutable tables behind the handles (host heap under the emulator) ----
#include <cstdlib>
#include <cstring>
static inline void *gqp_dev_alloc_upload(const void *src, size_t bytes) {
    if (bytes == 0) {  // empty table (e.g. a hand without penetration spheres): a valid pointer, nothing to read
        bytes = 16;
        src = nullptr;
    }
    void *d = nullptr;
#ifndef GQP_EMU
    if (cudaMalloc(&d, bytes) != cudaSuccess) return nullptr;
    if (src && cudaMemcpy(d, src, bytes, cudaMemcpyHostToDevice) != cudaSuccess) {
        cudaFree(d);
        return nullptr;
    }
#else
    d = std::aligned_alloc(256, (bytes + 255) / 256 * 256);
    if (d && src) std::memcpy(d, src, bytes);
#endif
    return d;
}
static inline void gqp_dev_free(void *d) {
    if (!d) return;
#ifndef GQP_EMU
    cudaFree(d);
#else
    std::free(d);
#endif
}
