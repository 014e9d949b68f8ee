#!/usr/bin/env bash
# Baseline G: the reference's ONE native component, TorchSDF, compiled unmodified for sm_100 from
# where it lies under /root/reference, plus the reference's pure-Python packages (graspqp, pytorch_kinematics), into
# baseline/_ref/ (git-ignored; travels to the GPU box with gpurun).
# Runs in the build container only (the GPU box has no /root/reference); no-op when already built.
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
REF="${GRASPQP_REFERENCE:-/root/reference}"
OUT="$ROOT/baseline/_ref"
if [ -f "$OUT/torchsdf/_C.so" ] && [ -d "$OUT/graspqp" ]; then echo "baseline/_ref present"; exit 0; fi
if [ ! -d "$REF/thirdparty/TorchSDF" ]; then echo "reference not present: cannot build baseline/_ref"; exit 0; fi
mkdir -p "$OUT"
if [ ! -f "$OUT/torchsdf/_C.so" ]; then
  TMP="$(mktemp -d /tmp/torchsdf_build.XXXXXX)"
  cp -r "$REF/thirdparty/TorchSDF/." "$TMP/"        # /root/reference is read-only: setup.py writes version.py
  ( cd "$TMP" && FORCE_CUDA=1 TORCH_CUDA_ARCH_LIST="10.0" MAX_JOBS=8 python setup.py build_ext --inplace >"$TMP/build.log" 2>&1 ) \
    || { tail -30 "$TMP/build.log"; exit 1; }
  mkdir -p "$OUT/torchsdf"
  cp "$TMP"/torchsdf/__init__.py "$TMP"/torchsdf/sdf.py "$TMP"/torchsdf/version.py "$OUT/torchsdf/"
  cp "$TMP"/torchsdf/_C*.so "$OUT/torchsdf/_C.so"
  rm -rf "$TMP"
  echo "built baseline/_ref/torchsdf/_C.so (TorchSDF, sm_100, unmodified sources)"
fi
# the reference's own Python for the path (optimizer, energy, metrics, pk): the installed tree, loaded module by module
# by baseline/gpu_reference.py (the package __init__ files import trimesh / pytorch3d / lxml, which are absent)
if [ ! -d "$OUT/graspqp" ]; then
  cp -r "$REF/graspqp/src/graspqp" "$OUT/graspqp"
  cp -r "$REF/thirdparty/pytorch_kinematics/pytorch_kinematics" "$OUT/pytorch_kinematics"
  find "$OUT" -name "__pycache__" -prune -exec rm -rf {} + 2>/dev/null || true
  chmod -R u+w "$OUT"
  echo "installed baseline/_ref/graspqp, baseline/_ref/pytorch_kinematics (unmodified)"
fi
