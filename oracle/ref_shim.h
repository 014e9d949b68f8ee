/* TEST INFRASTRUCTURE. Host stand-ins for the few CUDA built-ins the reference TorchSDF kernel body uses, so that
 * the kernel text of /root/reference/thirdparty/TorchSDF/torchsdf/csrc/unbatched_triangle_distance_cuda.cu can be
 * compiled by g++ unchanged (see oracle/build_ref.sh).  One emulated thread: blockDim = gridDim = 1, so the
 * kernel's own block-stride / grid-stride loops visit every face and every point. */
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
struct float3 { float x, y, z; };
struct double3 { double x, y, z; };
static inline float3 make_float3(float x, float y, float z) { return float3{x, y, z}; }
static inline double3 make_double3(double x, double y, double z) { return double3{x, y, z}; }
struct dim3_ { int x, y, z; };
static const dim3_ threadIdx{0, 0, 0}, blockIdx{0, 0, 0}, blockDim{1, 1, 1}, gridDim{1, 1, 1};
#define __device__
#define __global__
#define __forceinline__ inline
#define __shared__ static thread_local
static inline void __syncthreads() {}
static inline float rsqrt(float x) { return 1.0f / std::sqrt(x); }
static inline double rsqrt(double x) { return 1.0 / std::sqrt(x); }
using std::max;
using std::min;
