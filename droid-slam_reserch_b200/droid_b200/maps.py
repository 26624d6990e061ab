"""Reconstruction `.npy` folders -> keyframe buffers, sharded by owner GPU.

Format: `Droid.save_reconstruction` (droid_slam/droid.py:92-106) writes one folder per map with
`tstamps, images, disps, poses, intrinsics, fmaps, inps, nets` `.npy`, first axis = keyframe.  The multisession
backend (`Euroc_Multisession_Stereo/BackendAllMaps.py:21-61,101-138`) sorts the folders by the first number in their
name, keeps every N-th keyframe of every array (`[::N]`), and lays the maps end to end in the video buffer;
`counter` becomes the total.

Here the same concatenation is produced, but each rank only materialises the HEAVY per-keyframe arrays (fmaps, nets,
inps, images, disps) of the keyframes it owns under the edge partition of `dist.partition_frames` (contiguous frame
ranges): the files are memory-mapped, the strided rows are gathered into pinned host memory and copied to the
rank's GPU.  poses / intrinsics / tstamps are small and replicated (the reduced pose system is replicated too).
Host-side logic only; no kernels.
"""
from __future__ import annotations

import os
import re

import numpy as np
import torch

ARRAYS = ("tstamps", "images", "disps", "poses", "intrinsics", "fmaps", "inps", "nets")
REPLICATED = ("tstamps", "poses", "intrinsics")
SAVE_NAMES = {"tstamps": "tstamp"}        # BackendAllMaps.py:143 writes the fused map's timestamps as tstamp.npy


def _folder_number(name: str) -> int:
    """BackendAllMaps.py:21-24: the first run of digits in the folder name, 0 if none."""
    m = re.findall(r"\d+", name)
    return int(m[0]) if m else 0


def sorted_map_folders(path: str):
    """BackendAllMaps.py:26-41 (stable sort on the number, ties keep os.listdir order)."""
    if not os.path.exists(path):
        return []
    folders = [d for d in os.listdir(path) if os.path.isdir(os.path.join(path, d))]
    folders.sort(key=_folder_number)
    return [os.path.join(path, d) for d in folders]


def save_reconstruction(path: str, **arrays):
    """droid.py:92-106 (arrays already cut to `counter` keyframes)."""
    os.makedirs(path, exist_ok=True)
    for k in ARRAYS:
        if k in arrays:
            np.save(os.path.join(path, k + ".npy"), np.asarray(arrays[k]))


def map_index(path: str, every: int = 2):
    """Per-map keyframe counts after `[::every]` sampling and their offsets in the fused buffer.
    Returns (folders, counts int64[n_maps], offsets int64[n_maps+1])."""
    folders = sorted_map_folders(path)
    counts = []
    for f in folders:
        n = int(np.load(os.path.join(f, "poses.npy"), mmap_mode="r").shape[0])
        counts.append(len(range(0, n, every)))
    counts = np.asarray(counts, np.int64)
    return folders, counts, np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)


def load_maps(path: str, every: int = 2, own_lo: int = 0, own_hi=None, device=None, names=ARRAYS):
    """Fused keyframe arrays of all maps under `path`.

    Replicated arrays come back whole ([T, ...]); heavy arrays only for keyframes [own_lo, own_hi) of the fused
    numbering (`own_hi=None` -> all).  Returns dict(name -> tensor) plus 'counter', 'offsets', 'own' = (lo, hi)."""
    folders, counts, offsets = map_index(path, every)
    T = int(offsets[-1])
    own_hi = T if own_hi is None else min(int(own_hi), T)
    own_lo = max(0, min(int(own_lo), own_hi))
    out = {"counter": T, "offsets": offsets, "own": (own_lo, own_hi), "folders": folders}
    for name in names:
        lo, hi = (0, T) if name in REPLICATED else (own_lo, own_hi)
        parts = []
        for m, f in enumerate(folders):
            fp = os.path.join(f, name + ".npy")
            if not os.path.exists(fp):
                parts = None
                break
            a = np.load(fp, mmap_mode="r")
            # rows of this map inside [lo, hi) of the fused numbering -> rows of the sampled map -> rows of the file
            s0, s1 = max(lo, int(offsets[m])) - int(offsets[m]), min(hi, int(offsets[m + 1])) - int(offsets[m])
            if s1 > s0:
                parts.append(a[s0 * every:(s1 - 1) * every + 1:every])
        if parts is None:
            continue
        if parts:
            n_rows = sum(p.shape[0] for p in parts)
            host = torch.empty((n_rows,) + parts[0].shape[1:], dtype=torch.from_numpy(np.empty(0, parts[0].dtype)).dtype)
            if device is not None and torch.cuda.is_available():
                host = host.pin_memory()
            r = 0
            for p in parts:
                host[r:r + p.shape[0]] = torch.from_numpy(np.array(p))         # strided gather out of the memory map
                r += p.shape[0]
        else:
            host = torch.empty((0,))
        out[name] = host.to(device, non_blocking=True) if device is not None else host
    return out


def owned_range_for_rank(ii, jj, n_frames: int, rank: int, world: int):
    """The keyframes whose heavy arrays rank `rank` needs: the source frames it owns under dist.partition_frames."""
    from .dist import partition_frames
    b = partition_frames(ii, jj, n_frames, world)
    return int(b[rank]), int(b[rank + 1])
