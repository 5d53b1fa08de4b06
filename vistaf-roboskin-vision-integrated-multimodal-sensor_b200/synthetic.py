"""Synthetic fringe + TLC frames of the camera's resolution.

Host-side numpy generator: the single definition of the synthetic workload, so the oracle, the
reference (when golden vectors are made) and the CUDA path all see the same bytes.  A frame is a
2160x3840 BGR uint8 image of the skin disc: vertical two-tone stripes (black TLC / coloured TLC)
of the grating period, phase-modulated by a Gaussian indentation, coloured from a palette indexed
by a smooth temperature field, with illumination fall-off and sensor noise.

Frames are pre-registered (no shift / rotation between reference and deformed), so throughput
configs run with APPLY_GLOBAL_SHIFT=False and USE_ECC_CROP_ALIGNMENT=False on both sides.
"""
from __future__ import annotations

import json
import os

import numpy as np

H, W = 2160, 3840
DISC_C = (2012.0, 1129.0)
DISC_R = 720.0
PERIOD_PX = 65.84
_PAL = None


def _palette():
    global _PAL
    if _PAL is None:
        p = os.path.join(os.path.dirname(os.path.abspath(__file__)), "calibration", "synthetic_palette.json")
        rows = json.load(open(p))["rows"]
        _PAL = (np.array([r["T"] for r in rows], np.float32),
                np.array([r["light"] for r in rows], np.float32),
                np.array([r["dark"] for r in rows], np.float32))
    return _PAL


def frame_params(seed: int, reference: bool = False, amp_max: float = 1.3, sigma_min: float = 60.0,
                 dT_max: float = 10.0) -> dict:
    """Scene parameters drawn from default_rng(seed); the reference frame has no indentation.  The keyword ranges
    widen single draws without changing the draw sequence, so the default frames keep their bytes: amp_max up to
    3 pi brings wrap counts k != 0 into the unwrap, sigma_min keeps the carrier deviation inside
    the +-10-bin patch for such amplitudes, dT_max = 35 reaches the 32-34 degC blend band of the fusion."""
    rng = np.random.default_rng(int(seed) + (1 << 20 if reference else 0))
    ang = rng.uniform(0, 2 * np.pi)
    rad = rng.uniform(0, 250.0)
    p = dict(
        A=0.0 if reference else float(rng.uniform(0.1, amp_max)),
        sigma=float(rng.uniform(sigma_min, 160.0)),
        x0=float(DISC_C[0] + rad * np.cos(ang)), y0=float(DISC_C[1] + rad * np.sin(ang)),
        T0=float(rng.uniform(20.5, 28.0)), dT=float(rng.uniform(0.0, dT_max)),
        sT=float(rng.uniform(150.0, 400.0)),
        tx=float(DISC_C[0] + rng.uniform(-300, 300)), ty=float(DISC_C[1] + rng.uniform(-300, 300)),
        glare=(not reference) and (int(seed) % 16 == 15),
        noise_seed=int(rng.integers(0, 2 ** 31)),
    )
    if reference:
        p.update(T0=22.0, dT=0.0)
    return p


def make_frame(seed: int, reference: bool = False, amp_max: float = 1.3, h: int = H, w: int = W,
               sigma_min: float = 60.0, dT_max: float = 10.0, tilt_deg: float = 0.0) -> np.ndarray:
    """BGR uint8 (h, w, 3).  tilt_deg rotates the stripes (carrier angle != 0: the oriented blur of T12)."""
    p = frame_params(seed, reference, amp_max, sigma_min, dT_max)
    rng = np.random.default_rng(p["noise_seed"])
    img = np.full((h, w, 3), 25.0, np.float32)
    # work inside the disc's bounding box only
    x_lo, x_hi = int(max(0, DISC_C[0] - DISC_R - 2)), int(min(w, DISC_C[0] + DISC_R + 3))
    y_lo, y_hi = int(max(0, DISC_C[1] - DISC_R - 2)), int(min(h, DISC_C[1] + DISC_R + 3))
    yy, xx = np.mgrid[y_lo:y_hi, x_lo:x_hi].astype(np.float32)
    r2 = (xx - DISC_C[0]) ** 2 + (yy - DISC_C[1]) ** 2
    inside = r2 <= DISC_R ** 2
    phi = -p["A"] * np.exp(-((xx - p["x0"]) ** 2 + (yy - p["y0"]) ** 2) / (2.0 * p["sigma"] ** 2))
    if tilt_deg:
        ct, st = np.float32(np.cos(np.deg2rad(tilt_deg))), np.float32(np.sin(np.deg2rad(tilt_deg)))
        theta = (2.0 * np.pi / PERIOD_PX) * (xx * ct + (yy - DISC_C[1]) * st) + phi
    else:
        theta = (2.0 * np.pi / PERIOD_PX) * xx + phi
    wl = 0.5 * (1.0 + np.tanh(np.sin(theta) / (2.0 * np.pi * 3.0 / PERIOD_PX)))      # light-stripe weight
    T = p["T0"] + p["dT"] * np.exp(-((xx - p["tx"]) ** 2 + (yy - p["ty"]) ** 2) / (2.0 * p["sT"] ** 2))
    pt, pl, pd = _palette()
    illum = 1.0 + 0.15 * np.cos(np.sqrt(r2) / DISC_R * (np.pi / 2.0))
    box = img[y_lo:y_hi, x_lo:x_hi]
    for c in range(3):
        col = wl * np.interp(T, pt, pl[:, c]).astype(np.float32) + (1.0 - wl) * np.interp(T, pt, pd[:, c]).astype(np.float32)
        box[..., c] = np.where(inside, col * illum, box[..., c])
    if p["glare"]:
        for gx, gy in ((DISC_C[0] - 210.0, DISC_C[1] - 160.0), (DISC_C[0] + 260.0, DISC_C[1] + 120.0)):
            box[((xx - gx) ** 2 + (yy - gy) ** 2) <= 12.0 ** 2] = 255.0
    img += 2.0 * rng.standard_normal((h, w, 3), dtype=np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def make_batch(seeds, **kw) -> np.ndarray:
    return np.stack([make_frame(s, **kw) for s in seeds], axis=0)


# ---------------------------------------------------------------- configs[4]: 4K-upscaled fringe crops (FTP only)
CROP4K = dict(h=3904, w=3904, cx=1952, cy=1952, r=1951)        # 1182-px crop x 3.303; + 2 x 96 px of padding = 4096^2 FFT
SCALE4K = 3904.0 / 1182.0


def make_crop4k(seed: int, reference: bool = False, amp_max: float = float(3 * np.pi), size: int = 3904) -> np.ndarray:
    """uint8 gray crop (size, size) of the skin disc upscaled by 3.303: stripes of period 65.84 x 3.303 px
    phase-modulated by a Gaussian indentation of up to 3 pi rad (wrap counts k != 0), two grey levels with the same
    illumination fall-off and noise model as make_frame.  Generated directly as the gray crop the FTP chain consumes."""
    p = frame_params(seed, reference, amp_max, sigma_min=130.0)
    rng = np.random.default_rng(p["noise_seed"])
    sc = size / 1182.0
    c = (size - 1) / 2.0
    yy, xx = np.mgrid[0:size, 0:size].astype(np.float32)
    r2 = (xx - c) ** 2 + (yy - c) ** 2
    x0, y0 = c + (p["x0"] - DISC_C[0]) * sc, c + (p["y0"] - DISC_C[1]) * sc
    phi = -p["A"] * np.exp(-((xx - x0) ** 2 + (yy - y0) ** 2) / (2.0 * (p["sigma"] * sc) ** 2))
    theta = (2.0 * np.pi / (PERIOD_PX * sc)) * xx + phi
    wl = 0.5 * (1.0 + np.tanh(np.sin(theta) / (2.0 * np.pi * 3.0 / PERIOD_PX)))
    illum = 1.0 + 0.15 * np.cos(np.sqrt(r2) / (DISC_R * sc) * (np.pi / 2.0))
    img = (60.0 + 95.0 * wl) * illum
    img = np.where(r2 <= (DISC_R * sc) ** 2, img, 25.0).astype(np.float32)
    img += 2.0 * rng.standard_normal((size, size), dtype=np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)
