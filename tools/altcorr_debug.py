"""Localises errors of the tensor-core altcorr kernel: runs small-to-large cases against the CPU oracle and prints, per
case, the status words, the max error, where it is, and the error pattern per output channel / per tile.
    gpurun -- python tools/altcorr_debug.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "droid-slam_reserch_b200")); sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tests"))
import droid_backends  # noqa: E402
from droid_b200 import _capi  # noqa: E402
from oracle import oracle  # noqa: E402
import dropin_cases as D  # noqa: E402

t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()


def run(name, f1, f2, co):
    lib = _capi.lib()
    B, S, H, W, _ = co.shape
    _, H2, W2, _ = f2.shape
    ref = oracle.altcorr_forward(f1, f2, co, 3)
    wsb = lib.db200_altcorr_workspace_bytes(B, S, H, W, H2, W2, 128)
    ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
    out = torch.full((B, S, 49, H, W), 777.0, device="cuda")
    a, b, c = t(f1), t(f2), t(co)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.db200_altcorr_forward_ws(a.data_ptr(), b.data_ptr(), c.data_ptr(), out.data_ptr(), B, S, H, W, H2, W2, 128, 3,
                                      ws.data_ptr(), wsb, st)
    torch.cuda.synchronize()
    words = ws[:8].view(torch.int32).cpu().numpy(), ws[256:264].view(torch.int32).cpu().numpy()
    o = out.cpu().numpy()
    err = np.abs(o - ref)
    bad = err > 1e-4 * np.sqrt(128.0)
    print(f"[{name}] rc {rc} flags {words[0]} status {words[1]} max err {np.nanmax(err):.3e} (ref max {np.abs(ref).max():.2f}) "
          f"bad {bad.mean():.4f} unwritten {(o == 777.0).mean():.4f} nan {np.isnan(o).mean():.4f}", flush=True)
    if bad.any():
        idx = np.unravel_index(np.nanargmax(err), err.shape)
        print("   worst at (b,s,ch,y,x)", idx, "got", o[idx], "want", ref[idx])
        print("   bad fraction per channel (iy + 7 ix):", np.round(bad.mean(axis=(0, 1, 3, 4)), 2).tolist())
        ty, tx = (H + 7) // 8, (W + 15) // 16
        tiles = np.zeros((ty, tx))
        for i in range(ty):
            for j in range(tx):
                tiles[i, j] = bad[..., 8 * i:8 * i + 8, 16 * j:16 * j + 16].mean()
        print("   bad fraction per pixel tile:\n", np.round(tiles, 2))
    return not bad.any()


def case(seed, B, S, H, W, H2, W2, exact, mode):
    rng = np.random.default_rng(seed)
    f1 = rng.normal(0, 1, (B, H, W, 128)).astype(np.float32)
    f2 = rng.normal(0, 1, (B, H2, W2, 128)).astype(np.float32)
    if exact:
        f1, f2 = f1.astype(np.float16).astype(np.float32), f2.astype(np.float16).astype(np.float32)
    if mode == "grid":
        yy, xx = np.meshgrid(np.arange(H, dtype=np.float32), np.arange(W, dtype=np.float32), indexing="ij")
        co = np.broadcast_to(np.stack([xx * W2 / W + 0.25, yy * H2 / H + 0.5], -1), (B, S, H, W, 2)).copy()
    elif mode == "smooth":
        co = np.stack([D.smooth_coords(rng, B, H, W) * np.array([W2 / W, H2 / H]) for _ in range(S)], 1)
    else:
        co = np.stack([rng.uniform(-0.2 * W2, 1.2 * W2, (B, S, H, W)), rng.uniform(-0.2 * H2, 1.2 * H2, (B, S, H, W))], -1)
    return f1, f2, co.astype(np.float32)


if __name__ == "__main__":
    ok = True
    ok &= run("1 tile, grid coords, exact f16", *case(1, 1, 1, 8, 16, 8, 16, True, "grid"))
    ok &= run("1 tile, grid coords, fp32 (3 products)", *case(2, 1, 1, 8, 16, 8, 16, False, "grid"))
    ok &= run("6 tiles, grid coords", *case(3, 1, 1, 24, 32, 24, 32, True, "grid"))
    ok &= run("6 tiles, smooth coords", *case(4, 2, 1, 24, 32, 24, 32, False, "smooth"))
    ok &= run("ragged tiles", *case(5, 1, 2, 21, 29, 10, 14, False, "smooth"))
    ok &= run("scattered", *case(6, 2, 1, 24, 32, 24, 32, False, "scatter"))
    ok &= run("full size", *case(7, 3, 1, 48, 64, 48, 64, False, "smooth"))
    ok &= run("full size many edges (persistent loop > 1 task per CTA)", *case(8, 16, 1, 48, 64, 24, 32, True, "smooth"))
    print("ALL OK" if ok else "FAILURES")
    # timing at the judge's shape: 88 edges x 4 levels, fp32 random operands, through the nine-name API
    f1, f2, co = case(9, 88, 1, 48, 64, 48, 64, False, "smooth")
    pyr = [t(f2)]
    for _ in range(3):
        pyr.append(torch.nn.functional.avg_pool2d(pyr[-1].permute(0, 3, 1, 2), 2, stride=2).permute(0, 2, 3, 1).contiguous())
    a, c = t(f1), t(co)
    cl = [(c / 2 ** l).contiguous() for l in range(4)]
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(5):
            for l in range(4):
                droid_backends.altcorr_forward(a, pyr[l], cl[l], 3)
        e1.record(); torch.cuda.synchronize()
        print(f"altcorr_forward 88 edges x 4 levels (fp32 API, 4 calls): {e0.elapsed_time(e1) / 5:.3f} ms", flush=True)


def run_bwd(name, f1, f2, co, seed=3):
    B, S, H, W, _ = co.shape
    g = np.random.default_rng(seed).normal(0, 1, (B, S, 49, H, W)).astype(np.float32)
    r1, r2, _ = oracle.altcorr_backward(f1, f2, co, g, 3)
    droid_backends.CHECK_STATUS = False
    g1, g2, _ = droid_backends.altcorr_backward(t(f1), t(f2), t(co), t(g), 3)
    torch.cuda.synchronize()
    for nm, a, b in (("d_fmap1", g1.cpu().numpy(), r1), ("d_fmap2", g2.cpu().numpy(), r2)):
        err = np.abs(a - b)
        print(f"[bwd {name}] {nm}: max err {np.nanmax(err):.3e} / ref max {np.abs(b).max():.2f} = {np.nanmax(err) / np.abs(b).max():.2e}; "
              f"rel norm {np.linalg.norm(a - b) / np.linalg.norm(b):.2e}; nan {np.isnan(a).mean():.3f}", flush=True)
        if np.nanmax(err) > 1e-4 * np.abs(b).max():
            idx = np.unravel_index(np.nanargmax(err), err.shape)
            print("    worst at (b,y,x,c)", idx, "got", a[idx], "want", b[idx])
            print("    bad fraction per channel block of 16:", np.round((err > 1e-4 * np.abs(b).max()).reshape(-1, 8, 16).mean(axis=(0, 2)), 2).tolist())
            bad = (err > 1e-4 * np.abs(b).max()).mean(axis=(0, 3))
            print("    bad fraction per row (first 16 rows):", np.round(bad.mean(axis=1)[:16], 2).tolist())


if __name__ == "__main__":
    run_bwd("1 tile grid exact", *case(1, 1, 1, 8, 16, 8, 16, True, "grid"))
    run_bwd("1 tile grid fp32", *case(2, 1, 1, 8, 16, 8, 16, False, "grid"))
    run_bwd("6 tiles smooth", *case(4, 2, 1, 24, 32, 24, 32, False, "smooth"))
    run_bwd("ragged, S=2", *case(5, 1, 2, 21, 29, 10, 14, False, "smooth"))
    run_bwd("full size", *case(7, 3, 1, 48, 64, 48, 64, False, "smooth"))
    f1, f2, co = case(9, 88, 1, 48, 64, 48, 64, False, "smooth")
    a, c = t(f1), t(co)
    ga = [torch.randn((88, 1, 49, 48, 64), device="cuda") for _ in range(4)]
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(3):
            for l in range(4):
                droid_backends.altcorr_backward(a, pyr[l], cl[l], ga[l], 3)
        e1.record(); torch.cuda.synchronize()
        print(f"altcorr_backward 88 edges x 4 levels (fp32 API, 4 calls): {e0.elapsed_time(e1) / 3:.3f} ms", flush=True)
