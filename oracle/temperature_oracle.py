"""ORACLE — test infrastructure, never shipped, never on the product path.

CPU restatement (numpy + OpenCV, the same third-party calls the reference makes) of the
temperature path of the reference, `Code/temperature_sensor.py`.  Each function cites the
reference lines it follows.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.

Parity status: PINNED.  tools/make_golden.py imports the real reference in the build
container, runs it on the demo frames and on synthetic frames, and commits digests of its
outputs under tests/golden/; tests/test_oracle_golden.py checks this restatement against them
(and against the reference's own stored mask PNGs / summary JSONs).

Model constants come from vistaf.../calibration/*.json (numbers exported from the reference's
joblib pickles by tools/extract_calibration.py); the polynomial evaluation below restates
sklearn's StandardScaler -> PolynomialFeatures -> linear predict (temperature_sensor.py:236-240).
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field

import cv2
import numpy as np

_PKG = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..",
                    "vistaf-roboskin-vision-integrated-multimodal-sensor_b200")
CAL_DIR = os.path.join(_PKG, "calibration")


@dataclass
class TempParams:
    """temperature_sensor.py:34-98 (module constants)."""
    outer: tuple = ((1845, 1818), (1517, 623), (2687, 914))
    blur_ksize: int = 5
    color_t_min: float = 20.0
    color_t_max: float = 33.0
    guard: float = 0.5
    switch_margin: float = 1.0
    final_t_min: float = 20.0
    final_t_max: float = 75.0
    band_radius: int = 22
    dc_exclusion: int = 28
    max_dy_frac: float = 0.14
    illum_sigma: int = 20
    sat_thresh: int = 245
    sat_dilate: int = 13
    close_k: tuple = (3, 31)
    open_k: tuple = (3, 7)
    chroma_min: float = 10.0
    support_dilate: int = 3
    sigma_across: float = 6.0
    sigma_along: float = 1.0
    crop_pad: int = 10


def circle3(p1, p2, p3):
    """temperature_sensor.py:157-177 — circumcircle in float64."""
    (x1, y1), (x2, y2), (x3, y3) = [tuple(map(float, p)) for p in (p1, p2, p3)]
    a, b, c, d = x1 - x2, y1 - y2, x1 - x3, y1 - y3
    e = (x1 ** 2 - x2 ** 2 + y1 ** 2 - y2 ** 2) / 2.0
    f = (x1 ** 2 - x3 ** 2 + y1 ** 2 - y3 ** 2) / 2.0
    det = a * d - b * c
    if abs(det) < 1e-12:
        raise RuntimeError("Cannot define circle: points are collinear (or nearly collinear).")
    cx = (d * e - b * f) / det
    cy = (-c * e + a * f) / det
    return float(cx), float(cy), float(np.hypot(x1 - cx, y1 - cy))


def roi_mask(h, w, pts):
    """temperature_sensor.py:180-184."""
    cx, cy, r = circle3(*pts)
    Y, X = np.ogrid[:h, :w]
    return ((X - cx) ** 2 + (Y - cy) ** 2) <= r ** 2


def bbox(mask, pad):
    """temperature_sensor.py:195-208."""
    h, w = mask.shape
    ys, xs = np.where(mask)
    if ys.size == 0:
        return 0, h, 0, w
    return (int(max(0, ys.min() - pad)), int(min(h, ys.max() + pad + 1)),
            int(max(0, xs.min() - pad)), int(min(w, xs.max() + pad + 1)))


# --------------------------------------------------------------------------- models
@dataclass
class PolyModel:
    """StandardScaler -> PolynomialFeatures -> HuberRegressor(.predict) [-> Isotonic]."""
    features: tuple
    mean: np.ndarray
    scale: np.ndarray
    powers: np.ndarray
    coef: np.ndarray
    intercept: float
    iso_x: np.ndarray | None = None
    iso_y: np.ndarray | None = None

    @staticmethod
    def load(name):
        d = json.load(open(os.path.join(CAL_DIR, name)))
        iso = d.get("isotonic")
        return PolyModel(tuple(d["use_features"]), np.array(d["scaler_mean"]), np.array(d["scaler_scale"]),
                         np.array(d["powers"], int), np.array(d["coef"]), float(d["intercept"]),
                         np.array(iso["x"]) if iso else None, np.array(iso["y"]) if iso else None)

    def predict(self, X):
        """temperature_sensor.py:236-240.  X float32 (n, nfeat).  sklearn keeps float32 through the
        scaler and the polynomial terms; the final dot with float64 coefficients is float64."""
        X = np.asarray(X, np.float32)
        # StandardScaler.transform: `X -= astype(mean_, X.dtype); X /= astype(scale_, X.dtype)` (float32)
        z = (X - self.mean.astype(np.float32)) / self.scale.astype(np.float32)
        acc = np.full(X.shape[0], self.intercept, np.float64)
        for pw, c in zip(self.powers, self.coef):
            # PolynomialFeatures builds a degree-d column as (degree-(d-1) column of the LATER
            # features) * x_i, so the float32 product associates from the last feature backwards
            idx = [j for j, p in enumerate(pw) for _ in range(int(p))]
            t = np.ones(X.shape[0], np.float32) if not idx else z[:, idx[-1]].copy()
            for j in reversed(idx[:-1]):
                t = t * z[:, j]
            acc += c * t.astype(np.float64)
        pred = acc.astype(np.float32)
        if self.iso_x is not None:
            # IsotonicRegression(out_of_bounds='clip').predict: clip then linear interp (float64)
            v = np.clip(pred.astype(np.float64), self.iso_x[0], self.iso_x[-1])
            pred = np.interp(v, self.iso_x, self.iso_y).astype(np.float32)
        return pred


def _odd(k):
    k = int(k)
    return 1 if k <= 1 else (k + 1 if k % 2 == 0 else k)


def saturation_mask(img, roi, P):
    """temperature_sensor.py:378-387."""
    gray = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
    sat = (gray >= P.sat_thresh) & roi
    k = _odd(P.sat_dilate)
    if k > 1 and sat.any():
        ker = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k))
        sat = (cv2.dilate(sat.astype(np.uint8) * 255, ker, iterations=1) > 127) & roi
    return sat


def illum_normalize(g, roi, sigma):
    """temperature_sensor.py:363-375 (sigma > 0 branch)."""
    g = g.astype(np.float32)
    blur = cv2.GaussianBlur(g, (0, 0), sigmaX=float(sigma), sigmaY=float(sigma))
    blur[blur < 1e-6] = 1.0
    norm = g / blur
    mu = float(np.mean(norm[roi])) if roi.any() else float(np.mean(norm))
    mu = mu if abs(mu) > 1e-9 else 1.0
    return (norm / mu).astype(np.float32)


def top_peaks(mag, dc, n):
    """temperature_sensor.py:317-338 — the n largest bins outside the DC box, sorted descending."""
    h, w = mag.shape
    cy, cx = h // 2, w // 2
    ms = np.asarray(mag, float).copy()
    ms[max(0, cy - dc):min(h, cy + dc), max(0, cx - dc):min(w, cx + dc)] = 0.0
    flat = ms.ravel()
    idx = np.argpartition(flat, -n)[-n:]
    idx = idx[np.argsort(flat[idx])[::-1]]
    return [(int(i % w), int(i // w), float(flat[i])) for i in idx]


def choose_peak(peaks, h, w, max_dy_frac):
    """temperature_sensor.py:341-360."""
    cy, cx = h // 2, w // 2
    cand = [p for p in peaks if p[0] > cx] or peaks[:]
    md = int(max_dy_frac * h)
    cand = [p for p in cand if abs(p[1] - cy) <= md] or cand
    px, py, _ = max(cand, key=lambda t: t[2])
    return px, py


def postprocess(m, roi, P):
    """temperature_sensor.py:390-406 — close RECT(3x31) then open RECT(3x7)."""
    if not m.any():
        return m
    u = m.astype(np.uint8) * 255
    kc = cv2.getStructuringElement(cv2.MORPH_RECT, (_odd(P.close_k[0]), _odd(P.close_k[1])))
    ko = cv2.getStructuringElement(cv2.MORPH_RECT, (_odd(P.open_k[0]), _odd(P.open_k[1])))
    u = cv2.morphologyEx(u, cv2.MORPH_CLOSE, kc)
    u = cv2.morphologyEx(u, cv2.MORPH_OPEN, ko)
    return (u > 127) & roi


def segment(img, roi_full, P):
    """temperature_sensor.py:437-531 — FFT-sideband stripe segmentation."""
    h, w = img.shape[:2]
    gray = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY).astype(np.float32)
    sat = saturation_mask(img, roi_full, P)
    roi_eff = roi_full & ~sat
    if not roi_eff.any():
        raise RuntimeError("ROI became empty after saturation exclusion. Lower SAT_THRESH_GRAY / dilation.")
    g = gray.copy()
    med = float(np.median(g[roi_eff]))
    g[~roi_full] = med
    I = illum_normalize(g, roi_eff, P.illum_sigma)
    Fs = np.fft.fftshift(np.fft.fft2(I))
    mag = np.abs(Fs)
    peaks = top_peaks(mag, P.dc_exclusion, 16)
    px, py = choose_peak(peaks, h, w, P.max_dy_frac)
    Y, X = np.ogrid[:h, :w]
    bp = ((X - px) ** 2 + (Y - py) ** 2) <= float(P.band_radius) ** 2
    z = np.fft.ifft2(np.fft.ifftshift(Fs * bp))
    m = (I - 1.0).astype(np.float32)
    c = np.sum(z[roi_eff] * m[roi_eff])
    phi0 = float(np.angle(c)) if np.isfinite(c) else 0.0
    s = np.real(z * np.exp(-1j * phi0)).astype(np.float32)
    A = (s >= 0) & roi_eff
    B = (s < 0) & roi_eff
    mean_a = float(np.mean(gray[A])) if A.any() else 1e9
    mean_b = float(np.mean(gray[B])) if B.any() else 1e9
    dark = A if mean_a <= mean_b else B
    dark = postprocess(dark, roi_eff, P) & roi_eff
    light = roi_eff & ~dark
    cy, cx = h // 2, w // 2
    angle = float(np.arctan2(float(py - cy), float(px - cx)))
    return dict(dark=dark, light=light, sat=sat, roi_eff=roi_eff, peak=(px, py), phi0=phi0, s=s,
                angle=angle, I_norm=I, median=med, mean_a=mean_a, mean_b=mean_b, c=c)


def feature_planes(img, k):
    """temperature_sensor.py:279-292."""
    k = _odd(k)
    if k > 1:
        img = cv2.GaussianBlur(img, (k, k), 0)
    lab = cv2.cvtColor(img, cv2.COLOR_BGR2LAB).astype(np.float32)
    gray = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY).astype(np.float32)
    return {"L": lab[:, :, 0], "a": lab[:, :, 1], "b": lab[:, :, 2], "gray": gray}


def predict_map(model, planes, mask):
    """temperature_sensor.py:295-311."""
    out = np.full(mask.shape, np.nan, np.float32)
    if mask.any():
        X = np.stack([planes[f][mask] for f in model.features], axis=1).astype(np.float32)
        out[mask] = model.predict(X)
    return out


def clamp_map(m, roi, lo, hi):
    """temperature_sensor.py:538-543."""
    out = m.copy()
    sel = roi & np.isfinite(out)
    out[sel] = np.clip(out[sel], float(lo), float(hi))
    out[~roi] = np.nan
    return out


def inpaint_map(t, roi, radius):
    """temperature_sensor.py:546-577 — Telea fill of non-finite pixels inside roi (u8-quantised)."""
    out = t.copy()
    known = roi & np.isfinite(out)
    missing = roi & ~np.isfinite(out)
    if not missing.any() or not known.any():
        out[~roi] = np.nan
        return out
    vals = out[known]
    vmin, vmax = float(np.nanmin(vals)), float(np.nanmax(vals))
    if vmax - vmin < 1e-6:
        out[missing] = vmin
        out[~roi] = np.nan
        return out
    sc = np.zeros(out.shape, np.uint8)
    sc[known] = ((out[known] - vmin) / (vmax - vmin) * 255.0).clip(0, 255).astype(np.uint8)
    filled = cv2.inpaint(sc, missing.astype(np.uint8) * 255, inpaintRadius=int(radius), flags=cv2.INPAINT_TELEA)
    res = out.copy()
    res[roi] = (filled[roi].astype(np.float32) / 255.0) * (vmax - vmin) + vmin
    res[~roi] = np.nan
    return res


def fuse(roi, wide, color, P):
    """temperature_sensor.py:594-641."""
    final = wide.copy()
    src = np.zeros(final.shape, np.uint8)
    with np.errstate(invalid="ignore"):
        wide_ok = roi & np.isfinite(wide)
        color_ok = (roi & np.isfinite(color) & (color >= P.color_t_min - P.guard)
                    & (color <= P.color_t_max + P.guard))
        final[color_ok] = color[color_ok]
        src[color_ok] = 255
        lo, hi = P.color_t_max - P.switch_margin, P.color_t_max + P.switch_margin
        bz = wide_ok & color_ok & (wide > lo) & (wide < hi)
    if bz.any():
        wgt = np.clip((hi - wide[bz]) / (hi - lo), 0.0, 1.0).astype(np.float32)
        final[bz] = wgt * color[bz] + (1.0 - wgt) * wide[bz]
        src[bz] = 128
    return clamp_map(final, roi, P.final_t_min, P.final_t_max).astype(np.float32), src, color_ok


def oriented_blur(m, roi, angle, sx, sy):
    """temperature_sensor.py:705-742."""
    h, w = m.shape
    ctr = (w / 2.0, h / 2.0)
    deg = -float(angle) * 180.0 / float(np.pi)
    m0 = m.copy()
    m0[~np.isfinite(m0)] = 0.0
    M = cv2.getRotationMatrix2D(ctr, deg, 1.0)
    rot = cv2.warpAffine(m0, M, (w, h), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT)
    rroi = cv2.warpAffine(roi.astype(np.uint8) * 255, M, (w, h), flags=cv2.INTER_NEAREST,
                          borderMode=cv2.BORDER_CONSTANT, borderValue=0) > 127
    bl = cv2.GaussianBlur(rot.astype(np.float32), (0, 0), sigmaX=max(0.0, sx), sigmaY=max(0.0, sy))
    Mi = cv2.getRotationMatrix2D(ctr, -deg, 1.0)
    back = cv2.warpAffine(bl, Mi, (w, h), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT)
    broi = cv2.warpAffine(rroi.astype(np.uint8) * 255, Mi, (w, h), flags=cv2.INTER_NEAREST,
                          borderMode=cv2.BORDER_CONSTANT, borderValue=0) > 127
    out = back.astype(np.float32)
    out[~broi] = np.nan
    return out


_MODELS = {}


def models():
    if not _MODELS:
        _MODELS["color"] = PolyModel.load("color_model_global_huber_deg2.json")
        _MODELS["wide"] = PolyModel.load("black_model_global_huber_deg3.json")
    return _MODELS["color"], _MODELS["wide"]


def temperature_frame(img, P=None):
    """temperature_sensor.py:749-870 (numeric part of main) + multimodal_sensor.py:555-567 (stats)."""
    P = P or TempParams()
    h, w = img.shape[:2]
    roi = roi_mask(h, w, P.outer)
    color_model, wide_model = models()
    seg = segment(img, roi, P)
    roi_eff, sat = seg["roi_eff"], seg["sat"]
    planes = feature_planes(img, P.blur_ksize)
    chroma = np.sqrt((planes["a"] - 128.0) ** 2 + (planes["b"] - 128.0) ** 2).astype(np.float32)
    light_d = seg["light"]
    k = _odd(P.support_dilate)
    if k > 1 and light_d.any():
        ker = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k))
        light_d = cv2.dilate(light_d.astype(np.uint8) * 255, ker, iterations=1) > 127
    support = light_d & roi_eff & ~sat & (chroma >= P.chroma_min)
    wide_raw = predict_map(wide_model, planes, roi_eff)
    color_raw = predict_map(color_model, planes, support)
    wide = clamp_map(inpaint_map(wide_raw, roi, 7), roi, P.final_t_min, P.final_t_max)
    color = clamp_map(inpaint_map(color_raw, support, 5), support, P.color_t_min - 5.0, P.color_t_max + 5.0)
    fused, src, color_ok = fuse(roi, wide, color, P)
    final = clamp_map(oriented_blur(fused, roi, seg["angle"], P.sigma_across, P.sigma_along),
                      roi, P.final_t_min, P.final_t_max)
    valid = roi & np.isfinite(final)
    v = final[valid]
    stats = dict(mean_C=float(np.mean(v)), median_C=float(np.median(v)), std_C=float(np.std(v)),
                 min_C=float(np.min(v)), max_C=float(np.max(v)), valid_pixels=int(valid.sum())) if v.size else {}
    return dict(roi=roi, roi_eff=roi_eff, sat=sat, dark=seg["dark"], light=seg["light"], color_support=support,
                color_ok=color_ok, source=src, fused=fused, final=final, stats=stats, seg=seg,
                wide_raw=wide_raw, color_raw=color_raw, planes=planes, bbox=bbox(roi, P.crop_pad))
