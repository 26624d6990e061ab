"""Double-buffered host -> device staging for callers whose BA inputs live in (pinned) host memory.

Inside DROID-SLAM the `target`/`weight` maps come from the update operator and are already on the GPU
(`droid_slam/factor_graph.py:214-241`); offline tools and the multisession backend
(`Euroc_Multisession_Stereo/BackendAllMaps.py:101-138`, which loads reconstructions from `.npy`) start from
host arrays.  One `droid_backends.ba` call at 512 keyframes / 6k edges moves 295 MB over PCIe — longer than the
BA itself — so the upload of call k+1 is put on a copy stream and overlaps the kernels of call k:

    st = HostStagedCalls(device, host_inputs_template)
    st.upload(inputs_0)
    for k in range(n):
        if k + 1 < n: st.upload(inputs_{k+1})            # async, copy stream
        dev = st.acquire()                                # compute stream waits for upload k only
        dx, dz = droid_backends.ba(*dev, ...)             # in place on the staged poses/disps
        host_out.copy_(..., non_blocking=True)
        st.release()                                      # slot may be overwritten once the above has run
        torch.cuda.current_stream().synchronize()         # the caller reads call k's result

Plumbing only (streams and events); no kernels here.
"""
from __future__ import annotations

import torch


class HostStagedCalls:
    def __init__(self, device, template, depth: int = 2):
        self.device, self.depth = device, depth
        self.copy_stream = torch.cuda.Stream(device)
        self.slots = [[torch.empty(a.shape, dtype=a.dtype, device=device) for a in template] for _ in range(depth)]
        self.ready = [torch.cuda.Event() for _ in range(depth)]
        self.free = [torch.cuda.Event() for _ in range(depth)]
        self._used = [False] * depth
        self.n_up = self.n_acq = 0
        self.bytes_per_upload = sum(a.numel() * a.element_size() for a in template)

    def upload(self, host_inputs):
        """Enqueue the H2D copies of one call's inputs (pinned host tensors) on the copy stream."""
        if self.n_up - self.n_acq >= self.depth:
            raise RuntimeError("HostStagedCalls: all slots hold uploads that were not consumed yet")
        k = self.n_up % self.depth
        with torch.cuda.stream(self.copy_stream):
            if self._used[k]:
                self.copy_stream.wait_event(self.free[k])
            for d, h in zip(self.slots[k], host_inputs):
                if not h.is_pinned():
                    raise RuntimeError("HostStagedCalls.upload: host tensors must be pinned (page-locked)")
                d.copy_(h, non_blocking=True)
            self.ready[k].record(self.copy_stream)
        self.n_up += 1

    def acquire(self):
        """Device tensors of the oldest pending upload; the current stream waits for that upload only."""
        if self.n_acq >= self.n_up:
            raise RuntimeError("HostStagedCalls.acquire: nothing uploaded")
        k = self.n_acq % self.depth
        torch.cuda.current_stream(self.device).wait_event(self.ready[k])
        return self.slots[k]

    def release(self):
        """Call after the last kernel/copy that reads the acquired slot has been enqueued on the current stream."""
        k = self.n_acq % self.depth
        self.free[k].record(torch.cuda.current_stream(self.device))
        self._used[k] = True
        self.n_acq += 1
