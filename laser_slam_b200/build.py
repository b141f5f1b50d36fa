"""In-tree build of libzhsm.so (nvcc, sm_100a only).  python -m laser_slam_b200.build"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libzhsm.so")
SOURCES = ["zhsm_api.cu", "plicp_sm100.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
# -fmad=false: the reference links an x86-64 (no FMA contraction) CSM; every a*b+c must round twice like it does
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-fmad=false",
         "-Xcompiler", "-fPIC,-ffp-contract=off,-O2", "-shared", "-Xptxas", "-v"]


def newest(paths):
    return max(os.path.getmtime(p) for p in paths)


def build(force=False, verbose=False):
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "zhsm.h")]
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= newest(deps):
        return OUT
    ccbin = [] if not os.path.exists("/usr/bin/g++") else ["-ccbin", "/usr/bin/g++"]
    cmd = [NVCC] + ccbin + FLAGS + ["-o", OUT] + srcs
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    log = os.path.join(HERE, "build.log")
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout)
    if verbose or r.returncode != 0:
        print(r.stdout)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed, see " + log)
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
