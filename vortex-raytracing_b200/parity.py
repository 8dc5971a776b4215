"""Parity metrics between two result sets of the hot path (numpy only; no oracle is imported here — callers bring the
reference arrays): the north_star gates are ID equality on >= 99.99 % of rays with the residue attributed to equal-t ties,
t within 1e-5 relative, pixel colour within 1/255 per channel."""
import numpy as np


def channel_diff(a, b):
    """max per-channel |a-b| of two 0x00RRGGBB images"""
    d = 0
    for sh in (16, 8, 0):
        ca = ((a >> sh) & 255).astype(np.int32)
        cb = ((b >> sh) & 255).astype(np.int32)
        if ca.size:
            d = max(d, int(np.abs(ca - cb).max()))
    return d


def hit_parity(gpu, ref):
    """ID mismatch statistics between two ray_hit_t arrays (tests/regression/raycast/common.h:24-29)."""
    n = len(ref)
    miss_g, miss_r = gpu["dist"] == np.float32(1e30), ref["dist"] == np.float32(1e30)
    same = (miss_g & miss_r) | ((~miss_g) & (~miss_r) & (gpu["triIdx"] == ref["triIdx"]) & (gpu["blasIdx"] == ref["blasIdx"]))
    bad = np.flatnonzero(~same)
    both = (~miss_g) & (~miss_r)
    rel = np.zeros(n, dtype=np.float64)
    rel[both] = np.abs(gpu["dist"][both].astype(np.float64) - ref["dist"][both]) / np.maximum(np.abs(ref["dist"][both]), 1e-30)
    # a mismatch is a "tie" when both sides hit and report (nearly) the same distance
    tie = np.array([both[i] and rel[i] <= 1e-5 for i in bad], dtype=bool) if len(bad) else np.zeros(0, bool)
    return {"n": n, "mismatch": len(bad), "mismatch_frac": len(bad) / max(n, 1), "ties": int(tie.sum()),
            "non_tie": int(len(bad) - tie.sum()), "max_rel_t": float(rel[same & both].max()) if (same & both).any() else 0.0,
            "bit_exact_t": int((gpu["dist"][same] == ref["dist"][same]).sum()), "same": int(same.sum())}


def primary_rays(karg, rows):
    """GenerateRay (tests/regression/raycast/render.h:190-208) in numpy float32 for the pixels of `rows` (iterable of y):
    (n, 6) origin + direction.  Used only to feed the SAME rays to the GPU and to the CPU checker, so a last-bit difference
    from the reference's own generator would not matter."""
    f = np.float32
    W, H = int(karg["dst_width"][0]), int(karg["dst_height"][0])
    ys = np.asarray(list(rows), dtype=np.float32)
    xs = np.arange(W, dtype=np.float32)
    xn = ((xs + f(0.5)) / f(W) - f(0.5)) * f(karg["viewplane"][0][0])
    yn = ((ys + f(0.5)) / f(H) - f(0.5)) * f(karg["viewplane"][0][1])
    right, up, fwd, pos = (np.asarray(karg[k][0], dtype=np.float32) for k in ("camera_right", "camera_up", "camera_forward", "camera_pos"))
    pt = xn[None, :, None] * right[None, None, :] + yn[:, None, None] * up[None, None, :] + fwd[None, None, :]
    d = ((pt + pos) - pos).astype(np.float32)
    d = d * (f(1.0) / np.sqrt((d * d).sum(axis=2, keepdims=True, dtype=np.float32))).astype(np.float32)
    rays = np.empty((len(ys), W, 6), dtype=np.float32)
    rays[..., :3] = pos
    rays[..., 3:] = d
    return rays.reshape(-1, 6)
