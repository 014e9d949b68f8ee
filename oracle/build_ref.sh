#!/bin/bash
# TEST INFRASTRUCTURE. Builds oracle/_ref/libtorchsdf_ref.so from the reference's own kernel text where it lies
# (/root/reference, build container only) and oracle/_build/libsdf_oracle.so (our C restatement).  Outputs only
# under oracle/_ref and oracle/_build (both git-ignored; both travel to the GPU box with the snapshot).
set -e
cd "$(dirname "$0")"
mkdir -p _build _ref
gcc -O2 -ffp-contract=off -fopenmp -fPIC -shared sdf_oracle.c -o _build/libsdf_oracle.so -lm
REF=${GRASPQP_REFERENCE:-/root/reference}
SRC=$REF/thirdparty/TorchSDF/torchsdf/csrc/unbatched_triangle_distance_cuda.cu
if [ -f "$SRC" ]; then
  # kernel body = from "namespace kaolin {" up to (not including) the first host launcher
  awk '/^namespace kaolin/{on=1} /^void unbatched_triangle_distance_forward_cuda_impl/{on=0} on{print}' "$SRC" \
      > _ref/torchsdf_kernel_body.inc
  echo "}  // namespace kaolin" >> _ref/torchsdf_kernel_body.inc
  g++ -O2 -ffp-contract=off -fopenmp -fPIC -shared -I. -I_ref ref_wrap.cpp -o _ref/libtorchsdf_ref.so
  rm -f _ref/torchsdf_kernel_body.inc  # the extracted kernel text is a build intermediate, never kept
  echo "built oracle/_ref/libtorchsdf_ref.so"
else
  echo "reference not present; keeping prebuilt oracle/_ref if any"
fi
