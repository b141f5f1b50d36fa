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
    txt = open(os.path.join(ROOT, "include", "zhsm.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(zhsm_[a-z0-9_]+)\s*\(", txt)))


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
