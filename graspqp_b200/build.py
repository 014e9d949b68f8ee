"""Build the native libraries in-tree.

* build_cuda(): nvcc -> graspqp_b200/libgraspqp_b200.so (sm_100a only; the product; cross-compiles without a GPU).
* build_emu():  g++ -DGQP_EMU -> tests/_emu/libgraspqp_b200_emu.so — the SAME kernel sources run by a host SIMT
  emulator (csrc/gqp_emu.h).  Test infrastructure only: the Python package never loads it.
Both are incremental per source file (mtime of the source and of every header).
"""
import glob
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "graspqp_b200", "csrc")
INC = os.path.join(ROOT, "include")
CUDA_LIB = os.path.join(ROOT, "graspqp_b200", "libgraspqp_b200.so")
EMU_DIR = os.path.join(ROOT, "tests", "_emu")
EMU_LIB = os.path.join(EMU_DIR, "libgraspqp_b200_emu.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")


def _sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cpp")))


def _headers_mtime():
    hs = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(INC, "*.h"))
    return max(os.path.getmtime(h) for h in hs)


def _stale(out, src, hm):
    return (not os.path.exists(out)) or os.path.getmtime(out) < max(os.path.getmtime(src), hm)


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError("build failed: " + os.path.basename(cmd[-1]))
    return r.stdout + r.stderr


def build_cuda(verbose=False):
    objdir = os.path.join(ROOT, "build", "cuda")
    os.makedirs(objdir, exist_ok=True)
    hm = _headers_mtime()
    flags = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
             "-I", INC, "-I", CSRC, "--expt-relaxed-constexpr", "--expt-extended-lambda"]
    if verbose:
        flags += ["-Xptxas", "-v"]
    jobs, objs = [], []
    for s in _sources():
        o = os.path.join(objdir, os.path.basename(s) + ".o")
        objs.append(o)
        if _stale(o, s, hm) or verbose:
            jobs.append([NVCC] + flags + ["-x", "cu", "-c", s, "-o", o])
    with ThreadPoolExecutor(8) as ex:
        outs = list(ex.map(_run, jobs))
    if verbose:
        print("\n".join(outs))
    if jobs or not os.path.exists(CUDA_LIB):
        _run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", CUDA_LIB] + objs)
    return CUDA_LIB


def build_emu():
    objdir = os.path.join(ROOT, "build", "emu")
    os.makedirs(objdir, exist_ok=True)
    os.makedirs(EMU_DIR, exist_ok=True)
    hm = _headers_mtime()
    flags = ["-DGQP_EMU", "-O1", "-g", "-std=c++20", "-fPIC", "-ffp-contract=off", "-pthread", "-I", INC, "-I", CSRC,
             "-I", "/usr/local/cuda/include", "-Wno-attributes", "-Wno-unused-value"]
    jobs, objs = [], []
    for s in _sources():
        o = os.path.join(objdir, os.path.basename(s) + ".o")
        objs.append(o)
        if _stale(o, s, hm):
            jobs.append(["g++"] + flags + ["-x", "c++", "-c", s, "-o", o])
    with ThreadPoolExecutor(8) as ex:
        list(ex.map(_run, jobs))
    if jobs or not os.path.exists(EMU_LIB):
        _run(["g++", "-shared", "-pthread", "-o", EMU_LIB] + objs)
    return EMU_LIB


if __name__ == "__main__":
    what = sys.argv[1:] or ["cuda", "emu"]
    if "cuda" in what:
        print(build_cuda(verbose="-v" in what))
    if "emu" in what:
        print(build_emu())
