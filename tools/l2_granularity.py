"""Does cudaLimitMaxL2FetchGranularity change what the window-gather kernels pay per 16-byte window row?
The fused f16 lookup reads 1751 B of DRAM per lookup for 512 B of window data (every 16-byte row
of a window costs a whole DRAM burst).  The limit is a per-device hint (32 / 64 / 128 bytes); this tool times the
frontend lookup, the CorrBlock build, one altcorr chunk and the c3 `ba` call under each setting, in one process.

  python tools/l2_granularity.py            # on the GPU box
"""
import ctypes as C
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import droid_backends  # noqa: E402
from droid_b200 import synth  # noqa: E402

cudart = C.CDLL("libcudart.so.12")
LIMIT = 0x05  # cudaLimitMaxL2FetchGranularity


def get_limit():
    v = C.c_size_t(0)
    rc = cudart.cudaDeviceGetLimit(C.byref(v), LIMIT)
    return int(v.value) if rc == 0 else f"rc={rc}"


def set_limit(n):
    return cudart.cudaDeviceSetLimit(LIMIT, C.c_size_t(n))


def timed(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    torch.zeros(1, device=dev)
    p = synth.config_c3(stereo_frac=0.2, sensor_frac=0.2)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    poses0, disps0 = t(p.poses), t(p.disps)
    intr, sens, tgt, wgt, eta, ii, jj = t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)

    def ba():
        poses, disps = poses0.clone(), disps0.clone()
        droid_backends.ba(poses, disps, intr, sens, tgt, wgt, eta, ii, jj, p.t0, p.t1, 2, 1e-5, 1e-2, False)

    rows = []
    for want in (None, 32, 64, 128, 32):
        rc = set_limit(want) if want is not None else 0
        got = get_limit()
        lk = bench.lookup_bench(dev, 20, 3, 6527.1)
        vb = bench.volume_bench(dev, 10, 3, 6527.1)
        row = dict(requested=want, rc=rc, limit=got, lookup_ms=lk["ms_per_launch"], lookup_frac=lk["roofline"]["frac"],
                   volume_ms=vb["ms_per_call"], ba_ms=timed(ba, 10, 3))
        rows.append(row)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
