"""Builds libmsda_b200.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python -m kinet_b200.build [--force] [-v]

nvcc cross-compiles without a GPU.  The dtype families live in separate translation units that
compile in parallel; objects go to kinet_b200/build/ (git-ignored).  The .so is git-ignored but
travels to the GPU box with the gpurun snapshot.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libmsda_b200.so")
SOURCES = ["msda_abi.cu", "msda_inst_f32.cu", "msda_inst_bf16.cu", "msda_inst_f16.cu", "msda_static.cu", "layer_ops.cu", "small_mha.cu"]
HEADERS = ["msda_common.cuh", "msda_fwd.cuh", "msda_bwd.cuh", "msda_dense.cuh", "msda_launch.cuh",
           os.path.join("..", "..", "include", "msda_b200.h"),
           os.path.join("..", "..", "include", "kinet_layers_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def _nvcc():
    cuda_home = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cand = os.path.join(cuda_home, "bin", "nvcc")
    return cand if os.path.exists(cand) else "nvcc"


def _existing(names):
    return [os.path.join(CSRC, n) for n in names if os.path.exists(os.path.join(CSRC, n))]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = _existing(SOURCES + HEADERS) + [os.path.abspath(__file__)]
    return any(os.path.getmtime(p) > t for p in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    env = dict(os.environ)
    # the image's CC points at a gcc without libgomp; nvcc only needs a host g++
    env.pop("CC", None)
    env.pop("CXX", None)
    hdr_t = max(os.path.getmtime(p) for p in _existing(HEADERS) + [os.path.abspath(__file__)])
    procs, objs = [], []
    for src in _existing(SOURCES):
        obj = os.path.join(OBJ, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        if (not force and os.path.exists(obj) and os.path.getmtime(obj) > max(hdr_t, os.path.getmtime(src))):
            continue
        extra = os.environ.get("MSDA_NVCC_EXTRA", "").split()
        cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        procs.append((src, subprocess.Popen(cmd, cwd=CSRC, env=env, stdout=subprocess.PIPE,
                                            stderr=subprocess.STDOUT, text=True)))
    failed = []
    for src, p in procs:
        out, _ = p.communicate()
        if out and (verbose or p.returncode != 0):
            sys.stderr.write(out)
        if p.returncode != 0:
            failed.append(src)
    if failed:
        raise RuntimeError("nvcc failed for: " + ", ".join(failed))
    subprocess.check_call([_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB] + objs, env=env)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
