#!/usr/bin/env python
"""Freeze tests/golden/zhicp_testcpp.npz (run HERE, where /root/reference exists): the two 361-beam scans the
reference embeds in src/zhicp/test.cpp:54-60 (the active `#else` branch: SCAN_RANGES_1/2 + SCAN_VALID_1/2, metres,
aperture pi, right to left) and the initial pose of its run, (0.8, 0, 0) (:112).  The INPUTS are reference-held; the
file records no outputs ("parity unpinned")."""
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/src/zhicp/test.cpp"


def main():
    txt = open(SRC).read()
    active = txt.split("#else", 1)[1].split("#endif", 1)[0]   # "A more usual case"
    out = {}
    for name, dt in (("SCAN_RANGES_1", np.float32), ("SCAN_VALID_1", np.int8), ("SCAN_RANGES_2", np.float32), ("SCAN_VALID_2", np.int8)):
        body = re.search(name + r"\[\]\s*=\s*\{([^}]*)\}", active).group(1)
        vals = [v.strip().rstrip("f") for v in body.split(",") if v.strip()]
        a = np.array([float(v) for v in vals]).astype(dt)
        assert a.size == 361, (name, a.size)
        out[name.lower()] = a
    out["initial_pose"] = np.array([0.8, 0.0, 0.0])
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "zhicp_testcpp.npz"), **out)
    print({k: (v.shape, v.dtype) for k, v in out.items()})


if __name__ == "__main__":
    main()
