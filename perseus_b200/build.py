"""Builds libperseus_b200.so (hand-written sm_100a CUDA behind the C ABI of include/perseus_b200.h).

Plain nvcc, in-tree output (perseus_b200/libperseus_b200.so) so the built library travels to
the GPU box with the repo snapshot.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libperseus_b200.so")
STAMP = os.path.join(HERE, ".build_stamp")

SOURCES = ["runtime.cu", "frames.cu", "pool_head.cu", "conv_simt.cu", "conv_umma.cu", "conv_umma2.cu", "conv_halo.cu", "stem_pool.cu", "conv_p8.cu", "factors.cu", "plan.cu", "wgrad_umma.cu", "train.cu", "augment.cu", "smoother.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "--shared", "-Xcompiler", "-fPIC",
    "--fmad=true",
]


def _source_hash() -> str:
    h = hashlib.sha256()
    files = sorted(os.listdir(CSRC)) + [os.path.join("..", "..", "include", "perseus_b200.h")]
    for f in files:
        path = os.path.join(CSRC, f)
        if os.path.isfile(path):
            h.update(f.encode())
            with open(path, "rb") as fh:
                h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile if sources changed; returns the library path."""
    want = _source_hash()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP):
        with open(STAMP) as fh:
            if fh.read().strip() == want:
                return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    for src in SOURCES:
        obj = os.path.join(HERE, "build", src.replace(".cu", ".o"))
        cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
               "-Xcompiler", "-fPIC", "-c",
               os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas")
            cmd.insert(2, "-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            failed = True
            sys.stderr.write(f"--- nvcc failed on {src} ---\n{out}\n")
        elif verbose or out.strip():
            sys.stderr.write(f"--- {src} ---\n{out}\n")
    if failed:
        raise RuntimeError("nvcc failed building libperseus_b200.so")
    # shared cudart: ONE CUDA runtime per process (the one torch already loaded), so stream capture into CUDA graphs
    # and the current-device state are shared with PyTorch; rpath = the toolkit copy for torch-free loads
    link = [nvcc, "--shared", "--cudart", "shared", "-Xlinker", "-rpath=/usr/local/cuda/lib64", "-o", LIB] + objs + [
        "-gencode", "arch=compute_100a,code=sm_100a"]
    subprocess.check_call(link)
    with open(STAMP, "w") as fh:
        fh.write(want)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
