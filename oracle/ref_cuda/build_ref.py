#!/usr/bin/env python
"""Compile the REFERENCE's CUDA kernels for sm_100a from /root/reference (build container only;
the GPU box uses the prebuilt oracle/_ref/libmsda_ref_cuda.so that travels with the snapshot).
The reference's own build (ops/setup.py) needs a visible GPU and torch headers and fails against
torch 2.11 (ms_deform_attn_cuda.cu:69,139); its kernel header, however, is plain CUDA, so it is
compiled directly with three empty stand-in headers for its unused ATen/THC includes."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/src/trackformer/models/ops/src"
OUT_DIR = os.path.join(os.path.dirname(HERE), "_ref")
OUT = os.path.join(OUT_DIR, "libmsda_ref_cuda.so")


def main():
    if not os.path.isdir(REF_SRC):
        print("reference sources not present; keeping prebuilt", OUT)
        return 0
    os.makedirs(OUT_DIR, exist_ok=True)
    src = os.path.join(HERE, "ref_wrap.cu")
    hdr = os.path.join(REF_SRC, "cuda", "ms_deform_im2col_cuda.cuh")
    if os.path.exists(OUT) and os.path.getmtime(OUT) > max(os.path.getmtime(src), os.path.getmtime(hdr)):
        return 0
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
           "-Xcompiler", "-fPIC", "-shared", "-w", "-I", os.path.join(HERE, "shim"), "-I", REF_SRC,
           "-o", OUT, src]
    subprocess.check_call(cmd, env=env)
    print("built", OUT)
    return 0


if __name__ == "__main__":
    sys.exit(main())
