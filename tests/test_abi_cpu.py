"""CPU tests of the C-ABI boundary (no GPU, no compute calls): libzhsm.so loads, exports every symbol that
include/zhsm.h declares, has no CPU fallback (zhsm_create fails without a device), struct layouts and the
reference's parameter defaults (CCanonicalMatcher::sm_init, src/zhcsm/ZHCanonical_Matcher.cpp:83-157)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import laser_slam_b200 as zh

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(zh.lib_path()):
        from laser_slam_b200 import build
        build.build()
    return zh.load_library()


def _declared_symbols():
    syms = set()
    for hdr in ("zhsm.h", "csm_compat.h"):
        txt = open(os.path.join(ROOT, "include", hdr)).read()
        txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
        txt = re.sub(r"static inline[^{]*\{[^}]*\}", "", txt)
        syms |= set(re.findall(r"\b(zhsm_[a-z0-9_]+|sm_icp|sm_debug_write|ld_alloc_new|ld_free)\s*\(", txt))
    return sorted(syms)


def test_header_symbols_exported(lib):
    syms = _declared_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(lib, s), f"libzhsm.so does not export {s}"


def test_header_is_plain_c():
    """The boundary must compile as C (no C++/torch types in the signatures)."""
    src = '#include "zhsm.h"\nint main(void){ zhsm_result r; (void)r; return sizeof(zhsm_result) == 128 ? 0 : 1; }\n'
    p = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-x", "c", "-",
                        "-o", "/tmp/zhsm_hdr_test"], input=src.encode(), capture_output=True)
    assert p.returncode == 0, p.stderr.decode()
    assert subprocess.run(["/tmp/zhsm_hdr_test"]).returncode == 0


def test_no_cpu_fallback(lib):
    """On a box without a GPU every compute path must fail loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    st = lib.zhsm_create(0, C.byref(h))
    assert st != 0 and not h
    with pytest.raises(zh.ZhsmError):
        zh.Engine(0)


def test_result_record_layout():
    d = zh.RESULT_DTYPE
    assert d.itemsize == 128
    assert d.fields["x"][1] == 0 and d.fields["cov"][1] == 24 and d.fields["error"][1] == 72
    assert d.fields["valid"][1] == 80 and d.fields["iterations"][1] == 84 and d.fields["nvalid"][1] == 88
    assert d.fields["status"][1] == 92 and d.fields["corr_hash"][1] == 104 and d.fields["n_loops"][1] == 108


def test_default_params_are_sm_init(lib):
    p = zh.PlicpParams()
    lib.zhsm_plicp_default_params(C.byref(p), 10, 5000)
    want = dict(max_angular_correction_deg=45, max_linear_correction=0.5, max_iterations=10, epsilon_xy=0.001,
                epsilon_theta=0.001, max_correspondence_dist=2, sigma=0.04, use_corr_tricks=1, restart=1,
                restart_threshold_mean_error=0.01, restart_dt=0.01, restart_dtheta=0.0261799,
                use_point_to_line_distance=1, outliers_maxPerc=0.95, outliers_adaptive_order=0.7,
                outliers_adaptive_mult=2.0, outliers_remove_doubles=1, do_compute_covariance=0,
                min_reading=0.10, max_reading=50.0)
    for k, v in want.items():
        assert getattr(p, k) == pytest.approx(v, abs=0, rel=1e-15), k
    # and the oracle's copy of sm_init agrees field by field
    from oracle import pyoracle as po
    o = po.default_params(10, 5000)
    for k in want:
        assert getattr(o, k) == getattr(p, k), k


def test_laser_tables_match_reference_headers():
    """Base_PARAM arithmetic (src/zhpsm/PolarParameter.h:57-83): pm_dfi = fov*D2R/(points-1) in float, with the
    reference's truncated pi; cross-checked against the compiled reference when it is available."""
    from oracle import pyoracle as po
    for name in zh.LASER_TABLE:
        lz = zh.laser_from_table(name)
        assert lz.n == zh.LASER_TABLE[name][0]
        if po.have_ref():
            m = po.RefPolarMatch(name)
            assert (lz.n, lz.min_range_cm, lz.max_range_cm, lz.min_valid_points, lz.scan_window) == \
                   (m.n, m.min_range, m.max_range, m.min_valid, m.window)
            assert np.float32(lz.dfi) == m.dfi and np.float32(lz.fi_min) == m.fi_min


def test_corr_hash_chunk_algebra(lib):
    """ld_corr_hash is affine over GF(2); the kernel evaluates it as XOR over chunks of matrix*chunk-hash.  The host
    emulation of that decomposition (same matrices the device uses) must equal the sequential hash."""
    rng = np.random.default_rng(3)
    lib.zhsm_debug_corr_hash.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.zhsm_debug_corr_hash.restype = C.c_int
    for n in (10, 11, 181, 361, 541, 1080, 1081, 2048, 4096):
        for _ in range(3):
            j1 = rng.integers(-1, n, n)
            j2 = rng.integers(0, n, n)
            v = np.where(j1 < 0, -1, j1 + 1000 * j2).astype(np.int32)
            seq, par = C.c_uint32(), C.c_uint32()
            assert lib.zhsm_debug_corr_hash(v.ctypes.data, n, C.byref(seq), C.byref(par)) == 0
            # independent python restatement of ld_corr_hash
            h = 0
            for i, x in enumerate(v.tolist()):
                x &= 0xFFFFFFFF
                if i % 2 == 0:
                    h ^= ((h << 7) & 0xFFFFFFFF) ^ x ^ (h >> 3)
                else:
                    h ^= (~(((h << 11) & 0xFFFFFFFF) ^ x ^ (h >> 5))) & 0xFFFFFFFF
            assert seq.value == (h & 0x7FFFFFFF)
            assert par.value == seq.value, n


# ---------------------------------------------------------------------------------------------------------------
# the CSM-shaped entry (include/csm_compat.h): what can be checked without a GPU
# ---------------------------------------------------------------------------------------------------------------
CSM_PROBE = r"""
#include <csm/csm_all.h>
#include <csm/algos.h>
#include <csm/logging.h>
#include <math.h>
#include <stdio.h>
#include <string.h>
static void fill(LDP ld, float fi_min, float dfi, double reading) {
    int i;
    for (i = 0; i < ld->nrays; i++) {
        ld->valid[i] = 1; ld->readings[i] = reading; ld->theta[i] = fi_min + dfi * (unsigned)i; ld->cluster[i] = -1;
    }
    ld->min_theta = ld->theta[0]; ld->max_theta = ld->theta[ld->nrays - 1];
}
int main(void) {
    struct sm_params in; struct sm_result out;
    LDP a = ld_alloc_new(181), b = ld_alloc_new(181);
    int i, bad = 0;
    if (!a || !b || a->nrays != 181) return 1;
    for (i = 0; i < 181; i++) bad |= a->valid[i] != 0 || a->cluster[i] != -1 || a->corr[i].valid != 0 || a->corr[i].j1 != -1 || !isnan(a->readings[i]);
    if (bad) return 2;
    memset(&in, 0, sizeof(in));
    in.laser_ref = a; in.laser_sens = b; in.use_corr_tricks = 1; in.use_point_to_line_distance = 1; in.max_iterations = 10;
    in.min_reading = 0.1; in.max_reading = 50;
    sm_debug_write(0);
    /* 1: unfilled scans (NaN theta) are rejected like ld_valid_fields rejects them */
    sm_icp(&in, &out);
    if (out.valid != 0 || out.status != -1) return 3;
    /* 2: a theta[] that is not a float progression is refused, loudly */
    fill(a, -1.5707964f, 0.017453292f, (double)123.5f / 100.0); fill(b, -1.5707964f, 0.017453292f, (double)123.5f / 100.0);
    a->theta[7] += 1e-9; b->theta[7] += 1e-9;
    sm_icp(&in, &out);
    if (out.valid != 0 || out.status != -1) return 4;
    /* 3: readings that are not float-centimetre values are refused */
    fill(a, -1.5707964f, 0.017453292f, 1.2345678901234); fill(b, -1.5707964f, 0.017453292f, 1.2345678901234);
    sm_icp(&in, &out);
    if (out.valid != 0 || out.status != -1) return 5;
    /* 4: unsupported settings are refused */
    fill(a, -1.5707964f, 0.017453292f, (double)123.5f / 100.0); fill(b, -1.5707964f, 0.017453292f, (double)123.5f / 100.0);
    in.use_corr_tricks = 0;
    sm_icp(&in, &out);
    if (out.valid != 0 || out.status != -1) return 6;
    in.use_corr_tricks = 1;
    /* 5: a well-formed call reaches the engine: valid only with a GPU; without one valid = 0 (no CPU matcher) */
    sm_icp(&in, &out);
    printf("engine_valid=%d status=%d\n", out.valid, out.status);
    ld_free(a); ld_free(b);
    { gsl_matrix *m = zhsm_gsl_matrix_alloc(3, 3); gsl_matrix_set(m, 1, 2, 4.5); if (gsl_matrix_get(m, 1, 2) != 4.5 || gsl_matrix_get(m, 2, 1) != 0) return 7; zhsm_gsl_matrix_free(m); }
    return 0;
}
"""


def test_csm_entry_plain_c_links_against_libzhsm_only(lib, tmp_path):
    """include/csm/*.h forward to csm_compat.h; a plain-C program using sm_icp / ld_alloc_new / ld_free /
    gsl_matrix_get links against libzhsm.so alone; malformed inputs are refused with valid = 0 (never a different
    algorithm); without a GPU a well-formed call reports valid = 0 (there is no CPU matcher)."""
    src = tmp_path / "probe.c"
    src.write_text(CSM_PROBE)
    exe = tmp_path / "probe"
    libdir = os.path.dirname(zh.lib_path())
    p = subprocess.run(["gcc", "-std=gnu99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                        "-L" + libdir, "-lzhsm", "-lm", "-Wl,-rpath," + libdir], capture_output=True)
    assert p.returncode == 0, p.stderr.decode()
    r = subprocess.run([str(exe)], capture_output=True)
    assert r.returncode == 0, (r.returncode, r.stderr.decode())
    import torch
    if not torch.cuda.is_available():
        assert b"engine_valid=0 status=-1" in r.stdout
        assert b"no usable sm_100 GPU" in r.stderr
    needed = subprocess.run(["readelf", "-d", str(exe)], capture_output=True).stdout.decode()
    assert "libzhsm.so" in needed and "csm" not in needed.replace("libzhsm", "").lower()


def test_frame_record_wire_format(lib):
    """the client -> server frame record as clientLocal.cpp:189-199 writes it and serverSocket.cpp:60-115
    parses it (QDataStream: big-endian, quint16 block length, reals as 8-byte doubles since Qt 4.6).  A hand-built
    byte vector pins the layout; round trips cover both precisions, with and without covariance, and the server's
    "wait for more bytes" rule."""
    import struct
    from laser_slam_b200 import engine as E
    r = np.array([123.0, 4567.5, 5001.0], np.float32)
    cov = [1e-4, 2e-5, -3e-6, 4e-4, 5e-6, 6e-3]
    blk = E.frame_pack(1.5, -2.25, 0.125, 42, r, cov=cov)
    want = struct.pack(">dddi", 1.5, -2.25, 0.125, 42) + struct.pack(">6d", *cov) + struct.pack(">H", 3) + \
        struct.pack(">3d", 123.0, 4567.5, 5001.0)
    assert blk == struct.pack(">H", len(want)) + want
    blk_nc = E.frame_pack(1.5, -2.25, 0.125, 42, r)
    want_nc = struct.pack(">dddi", 1.5, -2.25, 0.125, 42) + struct.pack(">H", 3) + struct.pack(">3d", 123.0, 4567.5, 5001.0)
    assert blk_nc == struct.pack(">H", len(want_nc)) + want_nc
    blk_s = E.frame_pack(1.5, -2.25, 0.125, 42, r, single=True)
    want_s = struct.pack(">fffi", 1.5, -2.25, 0.125, 42) + struct.pack(">H", 3) + struct.pack(">3f", 123.0, 4567.5, 5001.0)
    assert blk_s == struct.pack(">H", len(want_s)) + want_s
    # round trips on real scans (LMS151-sized, 541 beams: the sensor the TCP client was written for)
    rng = np.random.default_rng(1)
    for single in (False, True):
        for with_cov in (False, True):
            scan = rng.uniform(10, 5001, 541).astype(np.float32)
            c = rng.normal(0, 1e-3, 6) if with_cov else None
            b = E.frame_pack(3.25, 7.5, -1.0, 7, scan, cov=c, single=single)
            n, f = E.frame_unpack(b + b"tail", with_cov=with_cov, single=single)
            assert n == len(b) and f["syn"] == 7 and (f["x"], f["y"], f["th"]) == (3.25, 7.5, -1.0)
            assert np.array_equal(f["r"], scan)
            if with_cov:
                assert np.array_equal(f["cov"], c)
            # a block that has not arrived completely: the server waits (consumed = 0), whatever the cut
            for cut in (0, 1, 2, len(b) // 2, len(b) - 1):
                assert E.frame_unpack(b[:cut], with_cov=with_cov, single=single)[0] == 0
    # a frame whose block would not fit the quint16 length field is refused, like the reference's format cannot say it
    with pytest.raises(zh.ZhsmError):
        E.frame_pack(0, 0, 0, 0, np.zeros(9000, np.float32))
    # a malformed block (np says more ranges than the block holds)
    bad = bytearray(E.frame_pack(0, 0, 0, 0, np.ones(4, np.float32)))
    bad[2 + 28 + 1] = 200
    with pytest.raises(zh.ZhsmError):
        E.frame_unpack(bytes(bad))
