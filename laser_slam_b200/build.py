"""In-tree build of libzhsm.so (nvcc, sm_100a only).  python -m laser_slam_b200.build"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libzhsm.so")
SOURCES = ["zhsm_api.cu", "plicp2_dispatch.cu", "plicp2_t64.cu", "plicp2_t128.cu", "plicp2_t160.cu", "plicp2_t192.cu", "plicp2_t256.cu", "plicp2_t256r.cu", "plicp2_t512.cu", "plicp2_final.cu", "plicp2_jump.cu", "plicp2_debug.cu", "psm_sm100.cu", "zhsm_psm_api.cu", "icp_sm100.cu", "zhsm_icp_api.cu", "graph_sm100.cu", "submap_sm100.cu", "synth_sm100.cu", "csm_compat.cu", "zhsm_noalloc.cu", "zhsm_handoff.cu", "zhsm_group.cu", "plicp2_ms128.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
# -fmad=false: the reference links an x86-64 (no FMA contraction) CSM; every a*b+c must round twice like it does
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-fmad=false",
         "-Xcompiler", "-fPIC,-ffp-contract=off,-O2", "-shared", "-Xptxas", "-v"]


def newest(paths):
    return max(os.path.getmtime(p) for p in paths)


def build(force=False, verbose=False, variant=None, defs=()):
    """Per-source objects (csrc/_obj/*.o, rebuilt only when a dependency is newer) linked into libzhsm.so.
    variant/defs: a tuning build with extra -D flags into libzhsm_<variant>.so (loaded with ZHSM_LIB_VARIANT=<variant>;
    e.g. variant="timing", defs=("PLICP_PHASE_TIMING",) for tools/gpu_probe_phases.py); never the product."""
    objdir = os.path.join(CSRC, "_obj" + ("_" + variant if variant else ""))
    OUT = os.path.join(HERE, "libzhsm%s.so" % ("_" + variant if variant else ""))
    extra = ["-D" + d for d in defs]
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh", ".inc"))]
    headers.append(os.path.join(HERE, "..", "include", "zhsm.h"))
    headers.append(os.path.join(HERE, "..", "include", "csm_compat.h"))
    ccbin = [] if not os.path.exists("/usr/bin/g++") else ["-ccbin", "/usr/bin/g++"]
    log = os.path.join(HERE, "build%s.log" % ("_" + variant if variant else ""))
    objs, rebuilt = [], False
    procs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(objdir, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < newest([src] + headers):
            cmd = [NVCC] + ccbin + [f for f in FLAGS if f != "-shared"] + extra + ["-c", "-o", obj, src]
            procs.append((s, cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
            rebuilt = True
    out_all = ""
    for s, cmd, pr in procs:
        out, _ = pr.communicate()
        out_all += " ".join(cmd) + "\n" + out
        if pr.returncode != 0:
            print(out)
            raise RuntimeError("nvcc failed on " + s)
    if rebuilt or not os.path.exists(OUT):
        cmd = [NVCC] + ccbin + ["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", OUT] + objs
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        out_all += " ".join(cmd) + "\n" + r.stdout
        if r.returncode != 0:
            print(r.stdout)
            raise RuntimeError("link failed")
    if out_all:
        with open(log, "w") as f:
            f.write(out_all)
        if verbose:
            print(out_all)
    return OUT


def build_host(verbose=False):
    """The C++ host shim (laser_slam_b200/host: CCanonicalMatcher with the reference's interface) as
    libzhsm_host.so, plus the C++ test driver tests/cpp/_build/shim_runlog that the GPU tests execute."""
    root = os.path.dirname(HERE)
    host = os.path.join(HERE, "host")
    out = os.path.join(HERE, "libzhsm_host.so")
    srcs = [os.path.join(host, f) for f in ("ZHCanonical_Matcher.cpp", "ZHPolar_Match.cpp", "ZHIcp_Warpper.cpp", "zh_carmen.cpp")]
    deps = srcs + [os.path.join(host, f) for f in os.listdir(host) if f.endswith(".h")] + [os.path.join(root, "include", "zhsm.h")]
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    if not os.path.exists(out) or os.path.getmtime(out) < newest(deps):
        cmd = [cxx, "-O2", "-std=c++17", "-Wall", "-fPIC", "-shared", "-I" + os.path.join(root, "include"), "-o", out] + srcs + \
              ["-L" + HERE, "-lzhsm", "-Wl,-rpath,$ORIGIN"]
        subprocess.run(cmd, check=True)
    tdir = os.path.join(root, "tests", "cpp")
    bdir = os.path.join(tdir, "_build")
    os.makedirs(bdir, exist_ok=True)
    exe = os.path.join(bdir, "shim_runlog")
    for name in ("shim_runlog", "shim_icp"):
        exe_k = os.path.join(bdir, name)
        src = os.path.join(tdir, name + ".cpp")
        if not os.path.exists(exe_k) or os.path.getmtime(exe_k) < newest(deps + [src, out]):
            cmd = [cxx, "-O2", "-std=c++17", "-Wall", "-I" + os.path.join(root, "include"), "-I" + host, "-o", exe_k, src,
                   "-L" + HERE, "-lzhsm_host", "-lzhsm", "-Wl,-rpath," + HERE]
            subprocess.run(cmd, check=True)
    # plain C against the CSM-shaped entry: includes <csm/csm_all.h>, links libzhsm.so only
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
    inc = os.path.join(root, "include")
    exe_c = os.path.join(bdir, "csm_plain")
    src_c = os.path.join(tdir, "csm_plain.c")
    lib = os.path.join(HERE, "libzhsm.so")
    if not os.path.exists(exe_c) or os.path.getmtime(exe_c) < newest([src_c, lib, os.path.join(inc, "csm_compat.h")]):
        subprocess.run([cc, "-O2", "-std=gnu99", "-Wall", "-I" + inc, "-o", exe_c, src_c, "-L" + HERE, "-lzhsm",
                        "-Wl,-rpath," + HERE], check=True)
    # the REFERENCE'S OWN CCanonicalMatcher, compiled unmodified where it lies (never copied) against include/csm/*.h
    # and linked against libzhsm.so instead of libcsm (only where /root/reference exists; the binary travels)
    ref = os.environ.get("ZHSM_REFERENCE", "/root/reference")
    ref_src = os.path.join(ref, "src", "zhcsm", "ZHCanonical_Matcher.cpp")
    if os.path.exists(ref_src):
        exe_r = os.path.join(bdir, "ref_zhcsm_on_zhsm")
        drv = os.path.join(tdir, "ref_zhcsm_driver.cpp")
        zhpsm = os.path.join(ref, "src", "zhpsm")
        if not os.path.exists(exe_r) or os.path.getmtime(exe_r) < newest([drv, lib, ref_src, os.path.join(inc, "csm_compat.h")]):
            subprocess.run([cxx, "-O2", "-w", "-I" + inc, "-I" + os.path.join(tdir, "stubs"), "-I" + zhpsm,
                            "-I" + os.path.join(ref, "src", "zhcsm"), "-o", exe_r, drv, ref_src,
                            os.path.join(zhpsm, "ZHPolar_Match.cpp"), os.path.join(zhpsm, "point.cpp"),
                            "-L" + HERE, "-lzhsm", "-Wl,-rpath," + HERE], check=True)
    return out, exe


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    build_host()
