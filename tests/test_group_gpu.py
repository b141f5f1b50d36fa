"""GPU test of zhsm_group_* (include/zhsm.h): several GPUs of one box from ONE process -- a handle per
device, jobs in contiguous blocks, one ncclAllGather of the 128-byte records.  Runs on as many GPUs as the box shows
(1 under the default gpurun, 2 under `gpurun --gpus 2`): the sharded result must be the single-GPU result, record for
record, and every GPU must hold all records after the gather."""
import numpy as np
import pytest

from tests.support import load_log, make_laser

pytestmark = pytest.mark.gpu


def _cudart():
    import ctypes
    import glob
    import os
    import torch
    cands = ["libcudart.so", "libcudart.so.12", "/usr/local/cuda/lib64/libcudart.so"]
    cands += glob.glob(os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*"))
    for c in cands:
        try:
            return ctypes.CDLL(c)
        except OSError:
            pass
    raise RuntimeError("libcudart not found")


def test_group_sharded_batch_equals_single_gpu():
    import torch
    import laser_slam_b200 as zh
    from laser_slam_b200 import engine as E
    ndev = min(torch.cuda.device_count(), 8)
    g = load_log("fr079")
    lz = make_laser(g)
    raw = np.ascontiguousarray(g["scans"][:1001])
    S = raw.shape[0]
    eng = zh.Engine(0)
    eng.set_laser(lz)
    grp = E.Group(list(range(ndev)))
    try:
        grp.set_laser(lz)
        rng = np.random.default_rng(3)
        cases = [
            (np.arange(0, S - 1), np.arange(1, S), None),                                    # consecutive-scan odometry
            (np.arange(0, 997), np.arange(3, 1000), rng.normal(0, 1, (997, 3)) * [0.03, 0.03, 0.01]),  # ragged: 997 jobs
            (np.array([5, 5, 5]), np.array([6, 7, 8]), None),                                 # fewer jobs than some blocks
        ]
        for ri, si, fg in cases:
            want = eng.plicp(raw, ri, si, first_guess=fg, filter_scans=True)
            got = grp.plicp(raw, ri, si, first_guess=fg, filter_scans=True)
            assert np.array_equal(got.tobytes(), want.tobytes()), (ri.size, np.flatnonzero(got["corr_hash"] != want["corr_hash"])[:10])
            # after the all-gather every GPU holds every record, in job order
            for k in range(ndev):
                ptr, block, B = grp.gathered(k)
                assert B == ri.size and block == -(-ri.size // ndev)
                with torch.cuda.device(k):
                    buf = torch.empty((B, 128), dtype=torch.uint8, device=f"cuda:{k}")
                    cudart = _cudart()
                    import ctypes
                    assert cudart.cudaMemcpy(ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(ptr), ctypes.c_size_t(B * 128), 3) == 0
                    rec = buf.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
                assert np.array_equal(rec.tobytes(), want.tobytes()), k
        assert (want["valid"] == 1).all()
    finally:
        grp.close()
        eng.close()
