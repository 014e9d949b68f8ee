/* TEST INFRASTRUCTURE. C entry points around the reference's own kernel body (extracted at build time into
 * oracle/_ref/torchsdf_kernel_body.inc by oracle/build_ref.sh; never committed). */
#include "ref_shim.h"
#include "torchsdf_kernel_body.inc"

extern "C" void torchsdf_ref_forward(const float *points, const float *faces, int P, int F, float *dist,
                                     int32_t *sign, float *normals, float *clst) {
    const int chunk = 256;
#pragma omp parallel for schedule(dynamic)
    for (int s = 0; s < P; s += chunk) {
        int n = std::min(chunk, P - s);
        kaolin::unbatched_triangle_distance_forward_cuda_kernel<float, float3, 512>(
            (const float3 *)points + s, (const float3 *)faces, n, F, dist + s, sign + s, (float3 *)normals + s,
            (float3 *)clst + s);
    }
}
extern "C" void torchsdf_ref_backward(const float *grad_dist, const float *points, const float *clst, int P,
                                      float *grad_points) {
    kaolin::unbatched_triangle_distance_backward_cuda_kernel<float, float3>(
        grad_dist, (const float3 *)points, (const float3 *)clst, P, (float3 *)grad_points);
}
