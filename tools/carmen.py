"""CARMEN ROBOTLASER1 log reader used to FREEZE test fixtures (tools/make_goldens.py).

Restates what CPolarMatch::readCarmon (src/zhpsm/ZHPolar_Match.cpp:172-284) extracts from a log:
per scan the range readings (m -> cm), the laser pose, and from the first record the start angle
and max range that overwrite the Base_PARAM table.  Used only here, never at GPU-test run time.
"""
import numpy as np

PM_PI = 3.141592654          # src/zhpsm/PolarParameter.h:38 (truncated pi, reproduced on purpose)
PM_D2R = PM_PI / 180.0

# Base_PARAM tables, src/zhpsm/PolarParameter.h:70-83: name -> (points, fov_deg, max_range_cm,
# min_valid_points, scan_window, fi_min_deg, fi_max_deg)
LASERS = {
    "LMS511": (361, 180, 5000, 80, 40, 0, 180),
    "LMS211": (181, 180, 5000, 40, 20, -90, 90),
    "LMS151": (541, 270, 5000, 100, 50, -45, 225),
}


def read_carmen(path, tag="ROBOTLASER1"):
    ranges, poses = [], []
    start = fov = max_range = None
    with open(path) as f:
        for line in f:
            tok = line.split()  # strtok semantics: runs of blanks collapse
            if not tok or tok[0] != tag:
                continue
            num = int(tok[8])
            if num == 0:
                continue
            if start is None:
                start, fov, max_range = float(tok[2]), float(tok[3]), float(tok[5])
            vals = [float(t) for t in tok[9:9 + num]]
            # remission count token, then the laser pose (ZHPolar_Match.cpp:262-270)
            nrem = 9 + num
            pose = [float(tok[nrem + 1]), float(tok[nrem + 2]), float(tok[nrem + 3])]
            ranges.append(vals)
            poses.append(pose)
    r_m = np.asarray(ranges, dtype=np.float64)
    # ls.r[i] = atof(tok)*100.0 narrowed to float (:252)
    r_cm = (r_m * 100.0).astype(np.float32)
    return dict(r_cm=r_cm, pose=np.asarray(poses, np.float64), start=start, fov=fov, max_range_m=max_range)
