"""ORACLE — test infrastructure, never shipped, never on the product path.

CPU restatement (numpy + OpenCV, the same third-party calls the reference makes) of the
Fourier-Transform-Profilometry shape path `Code/shape_ftp.py` and the force model
`Code/force_sensor.py`.  Each function cites the reference lines it follows.  The quality-guided
flood unwrap (a pure-Python heap loop in the reference, ~35 s) is restated in C
(oracle/native/unwrap_heap.c) with the same total pop order and float32 arithmetic.

Parity status: PINNED — tools/make_golden.py runs the real reference in the build container on the
demo pairs and synthetic pairs; tests/test_oracle_golden.py checks this file against those digests
and against the reference's stored height_map_bundle.npz / multimodal_summary.json anchors.
Telea inpainting, ECC and phaseCorrelate are OpenCV calls here exactly as in the reference.
"""
from __future__ import annotations

import ctypes
import json
import os
import subprocess
from dataclasses import dataclass

import cv2
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
CAL_DIR = os.path.join(_HERE, "..", "vistaf-roboskin-vision-integrated-multimodal-sensor_b200", "calibration")


@dataclass
class FtpParams:
    """shape_ftp.py:16-218 (module constants that the numeric path reads)."""
    circle_pts: tuple = ((1873, 1703), (1599, 707), (2575, 950))
    patch_half: int = 10
    dc_exclusion: int = 10
    n_peaks: int = 12
    fft_pad: int = 96
    pre_blur_sigma: float = 1.5
    roi_erode_px: int = 0
    apod_taper: int = 120
    amp_percentile: float = 25.0
    quality_sigma: float = 6.0
    edge_margin: int = 6
    poly_order: int = 2
    smooth_sigma: float = 2.5
    frontier_band: float = 200.0
    illum_sigma: float = 45.0
    close_kernel: int = 7
    bad_int_pct: float = 99.9
    bad_grad_pct: float = 99.7
    bad_dilate: int = 5
    bad_inpaint_radius: int = 3
    core_pct: float = 8.0
    contact_pct: float = 92.0
    dilate_kernel: int = 15
    dilate_iters: int = 2
    min_contact_frac: float = 0.002
    max_contact_frac: float = 0.40
    hole_nb: int = 11
    hole_frac: float = 0.70
    hole_min_dist: float = 4.0
    inpaint_radius: int = 5
    unrel_sigma: float = 9.0
    max_dy_frac: float = 0.12
    blob_min_peak_mm: float = 0.1
    blob_rel_frac: float = 1.0 / 3.0
    apply_global_shift: bool = True
    use_ecc: bool = True
    ecc_iters: int = 300
    ecc_eps: float = 1e-7
    ecc_gauss: float = 5.0


# ------------------------------------------------------------------ native heap unwrap
_lib = None


def _native():
    global _lib
    if _lib is None:
        so = os.path.join(_HERE, "native", "libvistaf_oracle.so")
        src = os.path.join(_HERE, "native", "unwrap_heap.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", so, src, "-lm"])
        _lib = ctypes.CDLL(so)
        _lib.vo_unwrap_quality_guided.restype = ctypes.c_long
    return _lib


def unwrap_quality_guided(wrapped, mask, quality):
    """shape_ftp.py:1043-1080."""
    h, w = wrapped.shape
    wr = np.ascontiguousarray(wrapped, np.float32)
    q = np.ascontiguousarray(quality, np.float32)
    m = np.ascontiguousarray(mask, np.uint8)
    out = np.empty((h, w), np.float32)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    _native().vo_unwrap_quality_guided(p(wr), p(m), p(q), ctypes.c_int(h), ctypes.c_int(w), p(out), None)
    return out


# ------------------------------------------------------------------ small helpers
def circle3_int(p1, p2, p3):
    """shape_ftp.py:406-414."""
    (x1, y1), (x2, y2), (x3, y3) = p1, p2, p3
    A = np.array([[2 * (x2 - x1), 2 * (y2 - y1)], [2 * (x3 - x1), 2 * (y3 - y1)]], float)
    b = np.array([x2 * x2 + y2 * y2 - x1 * x1 - y1 * y1, x3 * x3 + y3 * y3 - x1 * x1 - y1 * y1], float)
    cx, cy = np.linalg.solve(A, b)
    return int(round(cx)), int(round(cy)), int(round(float(np.hypot(cx - x1, cy - y1))))


def disc(h, w, cx, cy, r):
    """shape_ftp.py:383-386."""
    Y, X = np.ogrid[:h, :w]
    return ((X - cx) ** 2 + (Y - cy) ** 2) <= r ** 2


def apodization(h, w, cx, cy, r, taper):
    """shape_ftp.py:389-403 — flat inside r-taper, raised cosine to r, zero outside."""
    yy, xx = np.mgrid[0:h, 0:w]
    d = np.sqrt((xx - cx) ** 2 + (yy - cy) ** 2)
    apo = np.zeros((h, w), np.float32)
    r_in = max(0.0, float(r - taper))
    apo[d <= r_in] = 1.0
    ring = (d > r_in) & (d <= r)
    if taper > 0:
        apo[ring] = 0.5 * (1.0 + np.cos(np.pi * (d[ring] - r_in) / max(1e-6, float(taper))))
    return apo


def _fin(a, mask=None):
    v = np.asarray(a).ravel() if mask is None else np.asarray(a)[mask]
    return v[np.isfinite(v)]


def pct(a, q, mask=None, fallback=None):
    """shape_ftp.py:342-346."""
    v = _fin(a, mask)
    return fallback if v.size == 0 else float(np.nanpercentile(v, q))


def med(a, mask=None, fallback=None):
    """shape_ftp.py:349-353."""
    v = _fin(a, mask)
    return fallback if v.size == 0 else float(np.nanmedian(v))


def ellipse(k):
    return cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k))


def top_peaks(mag, dc, n):
    """shape_ftp.py:420-441."""
    h, w = mag.shape
    cy, cx = h // 2, w // 2
    ms = np.asarray(mag).copy()
    ms[max(0, cy - dc):min(h, cy + dc), max(0, cx - dc):min(w, cx + dc)] = 0
    flat = ms.ravel()
    n = min(n, flat.size)
    idx = np.argpartition(flat, -n)[-n:]
    idx = idx[np.argsort(flat[idx])[::-1]]
    return [(int(i % w), int(i // w), float(flat[i])) for i in idx]


def choose_peak(peaks, h, w, max_dy_frac):
    """shape_ftp.py:444-461."""
    cy, cx = h // 2, w // 2
    cand = [p for p in peaks if p[0] > cx] or peaks[:]
    md = int(max_dy_frac * h)
    cand = [p for p in cand if abs(p[1] - cy) <= md] or cand
    best = max(cand, key=lambda t: t[2])
    return best[0], best[1]


def _parab(fm, f0, fp):
    den = fm - 2.0 * f0 + fp
    return 0.0 if abs(den) < 1e-12 else 0.5 * (fm - fp) / den


def refine_peak(mag, px, py):
    """shape_ftp.py:473-483 — parabolic fit on log magnitude."""
    h, w = mag.shape
    x, y = int(px), int(py)
    if x <= 0 or x >= w - 1 or y <= 0 or y >= h - 1:
        return float(x), float(y)
    lm = np.log(mag + 1e-12)
    return (float(x + _parab(lm[y, x - 1], lm[y, x], lm[y, x + 1])),
            float(y + _parab(lm[y - 1, x], lm[y, x], lm[y + 1, x])))


# ------------------------------------------------------------------ pre-processing
def detect_bad_pixels(img, valid, P):
    """shape_ftp.py:625-649."""
    v = img[valid]
    v = v[np.isfinite(v)]
    hi = float(np.percentile(v, P.bad_int_pct)) if v.size else 255.0
    gx = cv2.Sobel(img, cv2.CV_32F, 1, 0, ksize=3)
    gy = cv2.Sobel(img, cv2.CV_32F, 0, 1, ksize=3)
    grad = np.sqrt(gx * gx + gy * gy)
    gv = grad[valid]
    gv = gv[np.isfinite(gv)]
    gthr = float(np.percentile(gv, P.bad_grad_pct)) if gv.size else 0.0
    bad = ((img >= hi) | (grad >= gthr)) & valid
    if P.bad_dilate and P.bad_dilate > 1:
        k = max(3, int(P.bad_dilate) | 1)
        bad = cv2.dilate(bad.astype(np.uint8) * 255, ellipse(k), iterations=1) > 0
    return bad


def inpaint_f32(img, mask, radius):
    """shape_ftp.py:652-666 — cv2 Telea on float32."""
    if not mask.any():
        return img
    m = np.zeros(img.shape, np.uint8)
    m[mask] = 255
    return cv2.inpaint(img.astype(np.float32).copy(), m, float(radius), cv2.INPAINT_TELEA).astype(np.float32)


def demod(gray_u8, apo, P, carrier_refined=None):
    """shape_ftp.py:810-1040 (`patch_shift` branch).  carrier_refined=None: reference-frame mode
    (search + refine the carrier peak); otherwise the carrier is locked to the given float peak."""
    img0 = gray_u8.astype(np.float32)
    valid = apo > 1e-6
    bad = detect_bad_pixels(img0, valid, P)
    if bad.any():
        img0 = inpaint_f32(img0, bad, P.bad_inpaint_radius)
    blur = cv2.GaussianBlur(img0, (0, 0), sigmaX=P.illum_sigma, sigmaY=P.illum_sigma)
    I = img0 / (blur + 1e-6) - 1.0
    I = cv2.GaussianBlur(I, (0, 0), P.pre_blur_sigma).astype(np.float32)
    Iw = I * apo
    Iw = Iw - med(Iw, valid, 0.0)
    pad = int(P.fft_pad)
    Ip = cv2.copyMakeBorder(Iw, pad, pad, pad, pad, borderType=cv2.BORDER_REFLECT)
    hf, wf = Ip.shape
    cy, cx = hf // 2, wf // 2
    F = np.fft.fftshift(np.fft.fft2(Ip))
    mag = np.abs(F)
    if carrier_refined is None:
        peaks = top_peaks(mag, P.dc_exclusion, P.n_peaks)
        px, py = choose_peak(peaks, hf, wf, P.max_dy_frac)
        pxf, pyf = refine_peak(mag, px, py)
    else:
        pxf, pyf = float(carrier_refined[0]), float(carrier_refined[1])
    kx, ky = pxf - cx, pyf - cy
    pxi, pyi = int(np.round(pxf)), int(np.round(pyf))
    bw = int(max(3, P.patch_half))
    x0, x1 = max(0, pxi - bw), min(wf, pxi + bw + 1)
    y0, y1 = max(0, pyi - bw), min(hf, pyi + bw + 1)
    patch = F[y0:y1, x0:x1].copy()
    ph, pw = patch.shape
    win = (np.hanning(ph).astype(np.float32)[:, None] * np.hanning(pw).astype(np.float32)[None, :]).astype(np.float32)
    patch *= win
    Fd = np.zeros_like(F)
    Fd[cy - ph // 2:cy - ph // 2 + ph, cx - pw // 2:cx - pw // 2 + pw] = patch
    field = np.fft.ifft2(np.fft.ifftshift(Fd))
    dpx, dpy = float(pxf - pxi), float(pyf - pyi)
    if abs(dpx) > 1e-6 or abs(dpy) > 1e-6:
        yy, xx = np.mgrid[0:hf, 0:wf]
        field = field * np.exp(-1j * 2.0 * np.pi * (dpx * (xx / wf) + dpy * (yy / hf)))
    h, w = Iw.shape
    c = field[pad:pad + h, pad:pad + w]
    return dict(c=c, amp=np.abs(c).astype(np.float32), peak_f=(pxf, pyf), k=(float(kx), float(ky)),
                hf_wf=(hf, wf), bad=bad, img_inpainted=img0, Iw=Iw, mag=mag)


# ------------------------------------------------------------------ masks
def largest_cc(m):
    """shape_ftp.py:707-718."""
    u = m.astype(np.uint8)
    if np.count_nonzero(u) == 0:
        return m
    n, lab, stats, _ = cv2.connectedComponentsWithStats(u, connectivity=8)
    if n <= 1:
        return m
    return lab == 1 + int(np.argmax(stats[1:, cv2.CC_STAT_AREA]))


def reliable_mask(amp_ref, amp_def, roi, circ, P):
    """shape_ftp.py:739-776."""
    q = cv2.GaussianBlur((amp_ref * amp_def).astype(np.float32), (0, 0), P.quality_sigma).astype(np.float32)
    thr = pct(q, P.amp_percentile, roi)
    if thr is None:
        thr = pct(q, P.amp_percentile, circ, 0.0)
    rel = roi & (q >= float(thr)) & np.isfinite(q)
    if rel.any():
        k = max(3, int(P.close_kernel) | 1)
        rel = (cv2.morphologyEx(rel.astype(np.uint8) * 255, cv2.MORPH_CLOSE, ellipse(k), iterations=1) > 0) & roi
    if rel.any():
        rel = largest_cc(rel) & roi
    if P.edge_margin > 0 and rel.any():
        dist = cv2.distanceTransform(rel.astype(np.uint8) * 255, cv2.DIST_L2, 3)
        rel = (dist > float(P.edge_margin)) & rel
    return rel, q, float(thr)


# ------------------------------------------------------------------ IRLS polynomial surface
def robust_polyfit(z, mask, order, iters=6, c=4.685):
    """shape_ftp.py:1100-1136 — Cauchy-weighted IRLS, float32 lstsq as in the reference."""
    h, w = z.shape
    m = mask & np.isfinite(z)
    nco = 6 if order >= 2 else 3
    if np.count_nonzero(m) < 200:
        return np.zeros(nco, np.float32), np.zeros_like(z, np.float32)
    yy, xx = np.indices((h, w))
    xn = (xx[m].astype(np.float32) - (w - 1) / 2.0) / ((w - 1) / 2.0)
    yn = (yy[m].astype(np.float32) - (h - 1) / 2.0) / ((h - 1) / 2.0)
    zz = z[m].astype(np.float32)
    cols = [xn, yn, np.ones_like(xn)] + ([xn * xn, xn * yn, yn * yn] if order >= 2 else [])
    A = np.stack(cols, axis=1)
    wts = np.ones_like(zz, np.float32)
    for _ in range(iters):
        coef, *_ = np.linalg.lstsq(A * wts[:, None], zz * wts, rcond=None)
        r = zz - A @ coef
        m0 = np.median(r)
        sigma = 1.4826 * (np.median(np.abs(r - m0)) + 1e-6)
        u = r / (c * sigma)
        wts = 1.0 / (1.0 + u * u)
    coef = coef.astype(np.float32)
    xf = (xx.astype(np.float32) - (w - 1) / 2.0) / ((w - 1) / 2.0)
    yf = (yy.astype(np.float32) - (h - 1) / 2.0) / ((h - 1) / 2.0)
    fit = coef[0] * xf + coef[1] * yf + coef[2]
    if order >= 2:
        fit = fit + coef[3] * xf * xf + coef[4] * xf * yf + coef[5] * yf * yf
    return coef, fit.astype(np.float32)


def masked_smooth(z, mask, sigma):
    """shape_ftp.py:1139-1147 — normalised convolution (sigma <= 0: the input, :1140)."""
    if sigma <= 0:
        return z
    z0 = z.copy().astype(np.float32)
    z0[~mask] = 0.0
    num = cv2.GaussianBlur(z0, (0, 0), sigma)
    den = cv2.GaussianBlur(mask.astype(np.float32), (0, 0), sigma) + 1e-6
    return (num / den).astype(np.float32)


def internal_holes(container, known, P):
    """shape_ftp.py:1153-1175."""
    known = known & container
    holes = container & ~known
    if not holes.any():
        return np.zeros_like(container, bool)
    k = max(3, int(P.hole_nb) | 1)
    ck = cv2.boxFilter(known.astype(np.float32), ddepth=-1, ksize=(k, k), normalize=False)
    cc = cv2.boxFilter(container.astype(np.float32), ddepth=-1, ksize=(k, k), normalize=False)
    dist = cv2.distanceTransform(container.astype(np.uint8) * 255, cv2.DIST_L2, 3)
    return holes & (ck / (cc + 1e-6) >= P.hole_frac) & (dist >= P.hole_min_dist)


def fill_internal_holes(hm, rel, P):
    """shape_ftp.py:1770-1801 — hole candidates inside the reliable mask, Telea fill (inpaint_only_mask :1178-1203).
    Returns (height_rel_filled, hole_candidates, output_reliable)."""
    h, w = hm.shape
    known = rel & np.isfinite(hm)
    filled = np.full((h, w), np.nan, np.float32)
    filled[known] = hm[known]
    holes = internal_holes(rel, known, P)
    if holes.any():
        tmp = filled.copy()
        tmp[rel & ~known] = float(np.nanmedian(tmp[known])) if known.any() else 0.0
        kn = rel & ~holes & np.isfinite(tmp)
        zin = np.full_like(tmp, float(np.nanmedian(tmp[kn])) if kn.any() else 0.0)
        zin[kn] = tmp[kn]
        zo = cv2.inpaint(zin, holes.astype(np.uint8) * 255, float(P.inpaint_radius), cv2.INPAINT_TELEA)
        filled[holes] = zo[holes]
    return filled, holes, rel & np.isfinite(filled)


def frontier(height, roi, rel, band, inside, outside):
    """shape_ftp.py:1287-1330 — smoothstep taper towards the reliable frontier."""
    out = height.astype(np.float32).copy()
    rel = rel & roi
    if not rel.any() or band <= 0:
        return out
    if inside:
        d = cv2.distanceTransform(rel.astype(np.uint8) * 255, cv2.DIST_L2, 3).astype(np.float32)
        t = np.clip(np.maximum(d - 1.0, 0.0) / max(1e-6, band), 0.0, 1.0).astype(np.float32)
        wgt = (t * t * (3.0 - 2.0 * t)).astype(np.float32)
        sel = rel & np.isfinite(out)
        out[sel] = 0.0 + (out[sel] - 0.0) * wgt[sel]
    if outside:
        d = cv2.distanceTransform((~rel).astype(np.uint8) * 255, cv2.DIST_L2, 3).astype(np.float32)
        out[roi & ~rel & (np.maximum(d - 1.0, 0.0) <= band)] = 0.0
    return out


# ------------------------------------------------------------------ calibrated models
def load_phase_to_height():
    """shape_ftp.py:672-680."""
    cal = json.load(open(os.path.join(CAL_DIR, "phase_to_height_calibration_model.json")))
    return cal["best_model"], bool(cal.get("use_negated_height_for_fit", True))


def model_predict(model, xs):
    """shape_ftp.py:682-700 / force_sensor.py:129-167 — the six calibrated curve families (float64)."""
    t, p = model["type"], model["params"]
    xs = np.asarray(xs, float)
    if t == "linear0":
        return p["a"] * xs
    if t == "linear":
        return p["a"] * xs + p["b"]
    if t == "poly2":
        return p["c2"] * xs * xs + p["c1"] * xs + p["c0"]
    if t == "sat_exp":
        return p["a"] * (1.0 - np.exp(-p["b"] * np.maximum(xs, 0.0)))
    if t == "growth":
        return p["a"] * (np.exp(p["b"] * np.maximum(xs, 0.0)) - 1.0)
    if t == "hinge_saturating":
        a, b, c = float(p["a"]), float(p["b"]), float(p["c"])
        return a * ((1.0 - np.exp(-b * np.maximum(xs - c, 0.0))) - (1.0 - np.exp(-b * np.maximum(0.0 - c, 0.0))))
    raise ValueError(f"Unknown model type in calibration: {t}")


def unitless_to_mm(hmap, model, use_neg):
    """shape_ftp.py:702-705 (model_predict clamps its argument at 0, :687)."""
    h = np.asarray(hmap, np.float32)
    x = np.maximum(np.asarray(-h if use_neg else h, float), 0.0)
    return model_predict(model, x).astype(np.float32)


def filter_blobs(height_mm, roi, P):
    """shape_ftp.py:1215-1281."""
    out = height_mm.astype(np.float32).copy()
    r = roi & np.isfinite(out)
    cand = r & (out > 0.0)
    if not cand.any():
        return out, np.zeros_like(roi, bool)
    gmax = float(np.max(out[cand]))
    thr = max(float(P.blob_min_peak_mm), float(P.blob_rel_frac) * gmax)
    n, lab, stats, _ = cv2.connectedComponentsWithStats(cand.astype(np.uint8), connectivity=8)
    peak = np.zeros(n, np.float64)
    np.maximum.at(peak, lab[cand], out[cand].astype(np.float64))
    keep = peak >= thr
    keep[0] = False
    kept = keep[lab] & cand
    out[cand & ~kept] = 0.0
    return out, kept


def volume_area_depth(height_mm, roi, mm_per_px, eps=0.01):
    """force_sensor.py:93-123."""
    Z = np.asarray(height_mm, np.float32).copy()
    pos = np.clip(Z, 0.0, np.inf)
    neg = np.clip(-Z, 0.0, np.inf)
    depth = neg if float(np.nansum(neg)) > float(np.nansum(pos)) else pos
    depth[~roi] = 0.0
    depth = np.where(np.isfinite(depth), depth, 0.0).astype(np.float32)
    contact = depth > float(eps)
    if not contact.any():
        return 0.0, 0.0, 0.0
    a = float(mm_per_px) ** 2
    return (float(np.sum(depth[contact]) * a) / 1000.0, float(np.count_nonzero(contact) * a),
            float(np.max(depth[contact])))


def load_force_model():
    return json.load(open(os.path.join(CAL_DIR, "height_to_force_calibration_model.json")))["best_model"]


def force_from_volume(best_model, v):
    """force_sensor.py:149-167."""
    return float(model_predict(best_model, float(v)))


# ------------------------------------------------------------------ alignment front-end (F-next row)
def align(ref_bgr, def_bgr, P, geom):
    """shape_ftp.py:1483-1537: global phase-correlation shift, crop, ECC euclidean alignment."""
    H, W = ref_bgr.shape[:2]
    x1, x2, y1, y2, circ = geom["x1"], geom["x2"], geom["y1"], geom["y2"], geom["circ"]
    info = {}
    if P.apply_global_shift:
        rg = cv2.cvtColor(ref_bgr, cv2.COLOR_BGR2GRAY).astype(np.float32)
        dg = cv2.cvtColor(def_bgr, cv2.COLOR_BGR2GRAY).astype(np.float32)
        rb, db = cv2.GaussianBlur(rg, (0, 0), 7), cv2.GaussianBlur(dg, (0, 0), 7)
        hann = cv2.createHanningWindow((W, H), cv2.CV_32F)
        shift, _ = cv2.phaseCorrelate(np.float32(rb), np.float32(db), hann)
        M = np.array([[1, 0, shift[0]], [0, 1, shift[1]]], np.float32)
        def_bgr = cv2.warpAffine(def_bgr, M, (W, H), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT)
        info["shift"] = shift
    ref_gray = cv2.cvtColor(ref_bgr[y1:y2, x1:x2], cv2.COLOR_BGR2GRAY)
    def_gray = cv2.cvtColor(def_bgr[y1:y2, x1:x2], cv2.COLOR_BGR2GRAY)
    if P.use_ecc:
        h, w = ref_gray.shape
        r = cv2.GaussianBlur(ref_gray.astype(np.float32) / 255.0, (0, 0), P.ecc_gauss)
        m = cv2.GaussianBlur(def_gray.astype(np.float32) / 255.0, (0, 0), P.ecc_gauss)
        warp = np.eye(2, 3, dtype=np.float32)
        crit = (cv2.TERM_CRITERIA_EPS | cv2.TERM_CRITERIA_COUNT, int(P.ecc_iters), float(P.ecc_eps))
        try:
            cc, warp = cv2.findTransformECC(r, m, warp, cv2.MOTION_EUCLIDEAN, crit,
                                            inputMask=circ.astype(np.uint8) * 255, gaussFiltSize=1)
            def_gray = cv2.warpAffine(def_gray, warp, (w, h), flags=cv2.INTER_LINEAR | cv2.WARP_INVERSE_MAP,
                                      borderMode=cv2.BORDER_REFLECT)
            info["ecc"] = float(cc)
        except cv2.error:
            info["ecc"] = float("nan")
    return ref_gray, def_gray, info


def geometry(H, W, P):
    """shape_ftp.py:1494-1528 — crop box, ROI discs and apodization from the fixed 3-point circle."""
    cx, cy, r = circle3_int(*P.circle_pts)
    x1, x2, y1, y2 = max(0, cx - r), min(W, cx + r), max(0, cy - r), min(H, cy + r)
    h, w = y2 - y1, x2 - x1
    cxl, cyl = cx - x1, cy - y1
    rl = int(min(r, cxl, cyl, w - 1 - cxl, h - 1 - cyl))
    circ = disc(h, w, cxl, cyl, rl)
    roi = disc(h, w, cxl, cyl, max(0, rl - int(P.roi_erode_px)))
    apo = apodization(h, w, cxl, cyl, rl, P.apod_taper)
    return dict(x1=x1, x2=x2, y1=y1, y2=y2, h=h, w=w, cx=cxl, cy=cyl, r=rl, circ=circ, roi=roi, apo=apo,
                cx_full=cx, cy_full=cy, r_full=r)


# ------------------------------------------------------------------ per-pair pipeline
def reference_products(ref_gray, geom, P):
    """The reference-frame half of shape_ftp.py:1632-1639 (cached once per batch by callers)."""
    return demod(ref_gray, geom["apo"], P, carrier_refined=None)


def ftp_pair(ref_gray, def_gray, geom, P=None, ref=None, keep=False):
    """shape_ftp.py:1632-1865 on already cropped / aligned uint8 gray crops.
    Returns None when the reliable mask is empty (shape_ftp.py:1677-1679)."""
    P = P or FtpParams()
    roi, circ, apo = geom["roi"], geom["circ"], geom["apo"]
    h, w = ref_gray.shape
    ref = ref or reference_products(ref_gray, geom, P)
    de = demod(def_gray, apo, P, carrier_refined=ref["peak_f"])
    dkx, dky = de["k"][0] - ref["k"][0], de["k"][1] - ref["k"][1]
    hf, wf = de["hf_wf"]
    rel, quality, thr = reliable_mask(ref["amp"], de["amp"], roi, circ, P)
    if not rel.any():
        return None
    ratio = de["c"] * np.conj(ref["c"])
    if abs(dkx) > 1e-6 or abs(dky) > 1e-6:
        yy, xx = np.mgrid[0:h, 0:w]
        ratio = ratio * np.exp(1j * 2.0 * np.pi * (dkx * (xx / max(1, wf)) + dky * (yy / max(1, hf))))
    wrapped = np.angle(ratio).astype(np.float32)
    unw = unwrap_quality_guided(wrapped, rel, quality)
    unw[~rel] = np.nan
    # debug_ramp (shape_ftp.py:1357-1420): always-on order-1 robust plane removal
    if unw[rel].size >= 500:
        _, fit1 = robust_polyfit(unw, rel, 1)
        unw = (unw - fit1).astype(np.float32)
    # two-pass detrend (shape_ftp.py:1716-1751)
    _, fit0 = robust_polyfit(unw, rel, P.poly_order)
    abs_res = np.abs((unw - fit0).astype(np.float32)).astype(np.float32)
    t = pct(abs_res, P.contact_pct, rel)
    if t is None or not np.isfinite(t):
        t = pct(abs_res, 95, rel, 0.0)
    with np.errstate(invalid="ignore"):
        contact = (abs_res >= float(t)) & rel & np.isfinite(abs_res)
        frac = contact.sum() / max(1, rel.sum())
        if frac < P.min_contact_frac:
            t = pct(abs_res, 95, rel, t)
            contact = (abs_res >= float(t)) & rel & np.isfinite(abs_res)
        elif frac > P.max_contact_frac:
            t = pct(abs_res, 98, rel, t)
            contact = (abs_res >= float(t)) & rel & np.isfinite(abs_res)
    contact_d = (cv2.dilate(contact.astype(np.uint8) * 255, ellipse(P.dilate_kernel), iterations=P.dilate_iters) > 0) & rel
    bg = rel & ~contact_d
    if bg.sum() < int(0.15 * rel.sum()):
        bg = rel.copy()
    _, fit = robust_polyfit(unw, bg, P.poly_order)
    det = (unw - fit).astype(np.float32)
    bm = med(det, bg)
    if bm is None or not np.isfinite(bm):
        bm = med(det, rel, 0.0)
    hm = det - float(bm)
    hm = masked_smooth(hm, rel & np.isfinite(hm), P.smooth_sigma)
    core_thr = pct(hm, P.core_pct, rel)
    flipped = False
    if core_thr is not None and np.isfinite(core_thr):
        core = rel & np.isfinite(hm) & (hm <= float(core_thr))
        if core.any() and float(np.median(hm[core])) > 0:
            hm = hm * -1.0
            flipped = True
    filled, holes, out_rel = fill_internal_holes(hm, rel, P)   # holes: empty whenever smooth_sigma > 0
    filled = frontier(filled, roi, out_rel, P.frontier_band, True, False)
    hf_ = np.full((h, w), np.nan, np.float32)
    hf_[roi] = 0.0
    hf_[out_rel] = filled[out_rel]
    sm = masked_smooth(hf_, roi, P.unrel_sigma)
    upd = roi & ~out_rel
    hf_[upd] = sm[upd]
    hf_ = frontier(hf_, roi, out_rel, P.frontier_band, False, True)
    sel = roi & np.isfinite(hf_)
    hf_[sel] = np.minimum(hf_[sel], 0.0)
    model, use_neg = load_phase_to_height()
    mm = unitless_to_mm(hf_, model, use_neg)
    mm, kept = filter_blobs(mm, roi, P)
    vals = [float(hw[1]) / abs(float(k[0])) for k, hw in ((ref["k"], ref["hf_wf"]), (de["k"], de["hf_wf"]))
            if abs(float(k[0])) > 1e-9]
    period = float(np.mean(vals)) if vals else None
    res = dict(height_mm=mm, roi=roi, reliable=rel, output_reliable=out_rel, contact_dilated=contact_d,
               contact_kept=kept, hole_candidates=holes, period_px=period, flipped=flipped, amp_thr=thr,
               peak_f=ref["peak_f"])
    if keep:
        res.update(wrapped=wrapped, quality=quality, unwrapped=unw, height_unitless=hf_, cdef=de["c"],
                   cref=ref["c"], bad_def=de["bad"], Iw_def=de["Iw"], img_inp_def=de["img_inpainted"], hm=hm,
                   abs_res=abs_res, contact_thr=float(t))
    return res


def force_readings(res, pitch_mm=2.0, eps=0.01):
    """multimodal_sensor.py:388-419."""
    mm_per_px = float(pitch_mm) / float(res["period_px"])
    roi = np.isfinite(res["height_mm"])
    V, A, D = volume_area_depth(res["height_mm"], roi, mm_per_px, eps)
    return dict(force_N=force_from_volume(load_force_model(), V), volume_cm3=V, contact_area_mm2=A,
                max_depth_mm=D, scale_mm_per_px=mm_per_px)


def shape_pair_from_frames(ref_bgr, def_bgr, P=None, keep=False):
    """shape_ftp.main numeric path from two full BGR frames (shape_ftp.py:1471-1865)."""
    P = P or FtpParams()
    if ref_bgr.shape != def_bgr.shape:
        raise RuntimeError("Reference and deformed images have different sizes.")
    H, W = ref_bgr.shape[:2]
    geom = geometry(H, W, P)
    rg, dg, info = align(ref_bgr, def_bgr, P, geom)
    res = ftp_pair(rg, dg, geom, P, keep=keep)
    if res is not None:
        res["geom"], res["align"] = geom, info
    return res
