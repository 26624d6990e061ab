"""droid_b200 — host-side mirror of the reference interfaces around the B200 hot path:
`CorrBlock` / `AltCorrBlock` (droid_slam/modules/corr.py), the BA / distance entry points of
`DepthVideo` (droid_slam/depth_video.py:160-204), edge sharding + allreduce for multi-GPU BA, and
the synthetic EuRoC-shaped inputs used by tests and bench.py.  The kernels live in csrc/ behind the
C-ABI of include/droid_b200.h; the drop-in operator module is `droid_backends`."""
from . import synth  # noqa: F401
