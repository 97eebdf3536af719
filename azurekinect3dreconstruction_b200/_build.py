"""In-tree build of libb2recon.so (hand-written sm_100a CUDA behind a C-ABI) and of the CPU oracle.

``python -m azurekinect3dreconstruction_b200._build`` or ``__graft_entry__.build()``.
nvcc cross-compiles without a GPU; the .so stays in-tree so it travels with the repo snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libb2recon.so")
SOURCES = ["tsdf.cu", "unproject.cu", "extract.cu", "odometry.cu", "icp.cu", "pointcloud.cu", "posegraph.cu", "features.cu", "peer.cu"]
HEADERS = ["common.cuh", "host_math.h", "tsdf_internal.h", "mc_tables.h", "rigid_fit.cuh"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false",                       # parity: never contract a*b+c (see DESIGN.md "Arithmetic parity")
    "-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=hidden",
    "-Xptxas", "-v",
    "--shared", "-cudart", "static",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build_cuda(force: bool = False, verbose: bool = False) -> str:
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    deps = srcs + [os.path.join(CSRC, h) for h in HEADERS] + [os.path.abspath(__file__)]
    if not force and not _stale(LIB, deps) and not os.environ.get("B2R_LIB_OUT"):
        return LIB
    env = dict(os.environ)
    # the image exports CC/CXX pointing at a wrapper without libgomp specs; use the system compiler
    ccbin = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    # tuning experiments: B2R_EXTRA_DEFS="-DFUSE_MIN_CTAS=2" B2R_LIB_OUT=/path/variant.so (load it with B2R_LIB=...)
    extra = os.environ.get("B2R_EXTRA_DEFS", "").split()
    out = os.environ.get("B2R_LIB_OUT", LIB)
    cmd = [_nvcc(), "-ccbin", ccbin] + NVCC_FLAGS + extra + ["-I", CSRC, "-o", out] + srcs
    res = subprocess.run(cmd, env=env, capture_output=True, text=True)
    log = os.path.join(PKG, "csrc", "build.log")
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed (see %s)" % log)
    return LIB


def build_oracle(force: bool = False) -> str:
    odir = os.path.join(ROOT, "oracle")
    lib = os.path.join(odir, "liboracle.so")
    deps = [os.path.join(odir, "oracle.cpp"), os.path.join(odir, "oracle_pcd.cpp"), os.path.join(CSRC, "mc_tables.h"), os.path.join(odir, "Makefile")]
    if force or _stale(lib, deps):
        res = subprocess.run(["make", "-C", odir] + (["-B"] if force else []), capture_output=True, text=True)
        if res.returncode != 0:
            sys.stderr.write(res.stdout + res.stderr)
            raise RuntimeError("oracle build failed")
    return lib


if __name__ == "__main__":
    force = "--force" in sys.argv
    print(build_cuda(force=force, verbose=True))
    print(build_oracle(force=force))
