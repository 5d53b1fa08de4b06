"""temperature_sensor — drop-in for the reference's Code/temperature_sensor.py, computed on B200.

Same module-level configuration names, same `main()` convention (argument-less; the caller mutates
INPUT_IMAGE_PATH / OUTPUT_DIR, multimodal_sensor.py:492-503), same outputs in OUTPUT_DIR
(temperature_map_final.npy, temperature_map_fused.npy, mask_*.png, debug_fused_source.png), same
RuntimeErrors.  The numeric body (temperature_sensor.py:749-870) runs as hand-written sm_100a kernels
behind libvistaf (`vistaf_temp_batch`); matplotlib figures are not produced (headless).
`run_batch` is the batched entry point beside the reference-compatible single-frame `main`.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Tuple

import numpy as np

import _native as N

# ---------------------------------------------------------------- USER CONFIG (temperature_sensor.py:34-98)
INPUT_IMAGE_PATH = "./Final_demos_images/FINAL_TEMP_DEMO.jpg"
OUTPUT_DIR = "./Temperature/temp_inference"
OUTER_CIRCLE_P1 = (1845, 1818)
OUTER_CIRCLE_P2 = (1517, 623)
OUTER_CIRCLE_P3 = (2687, 914)
USE_INNER_CIRCLE = False
INNER_CIRCLE_P1 = (1881, 1749)
INNER_CIRCLE_P2 = (1579, 665)
INNER_CIRCLE_P3 = (2616, 936)
CROP_OUTPUT_TO_OUTER_ROI = True
CROP_PAD_PX = 10
BLUR_KSIZE = 5
COLOR_T_MIN = 20.0
COLOR_T_MAX = 33.0
COLOR_GUARD_BAND = 0.5
SWITCH_MARGIN_C = 1.0
FINAL_T_MIN = 20.0
FINAL_T_MAX = 75.0
SEG_BAND_RADIUS = 22
SEG_DC_EXCLUSION = 28
SEG_FORCE_RIGHT_HALF_PLANE = True
SEG_PREFER_PEAK_NEAR_CENTER_ROW = True
SEG_PEAK_MAX_DY_FROM_CENTER = 0.14
SEG_ILLUM_SIGMA = 20
SAT_THRESH_GRAY = 245
SAT_DILATE_KSIZE = 13
POST_CLOSE_KX = 3
POST_CLOSE_KY = 31
POST_OPEN_KX = 3
POST_OPEN_KY = 7
COLOR_CHROMA_MIN = 10.0
COLOR_SUPPORT_DILATE = 3
FINAL_SMOOTH_ENABLE = True
FINAL_SMOOTH_SIGMA_ACROSS = 6.0
FINAL_SMOOTH_SIGMA_ALONG = 1.0
DEBUG_SAVE = True

# model constants (numbers exported from the reference's joblib pickles, tools/extract_calibration.py)
COLOR_MODEL_JSON = os.path.join(N.CAL_DIR, "color_model_global_huber_deg2.json")
WIDE_MODEL_JSON = os.path.join(N.CAL_DIR, "black_model_global_huber_deg3.json")
for _p in (COLOR_MODEL_JSON, WIDE_MODEL_JSON):      # the reference raises at import when a model is absent (:105-150)
    if not os.path.exists(_p):
        raise RuntimeError(f"No model matches pattern:\n  {_p}")


# ---------------------------------------------------------------- geometry (temperature_sensor.py:157-208)
def circle_from_three_points(p1, p2, p3, eps: float = 1e-12) -> Tuple[float, float, float]:
    x1, y1 = map(float, p1)
    x2, y2 = map(float, p2)
    x3, y3 = map(float, p3)
    a, b, c, d = x1 - x2, y1 - y2, x1 - x3, y1 - y3
    e = (x1 ** 2 - x2 ** 2 + y1 ** 2 - y2 ** 2) / 2.0
    f = (x1 ** 2 - x3 ** 2 + y1 ** 2 - y3 ** 2) / 2.0
    det = a * d - b * c
    if abs(det) < eps:
        raise RuntimeError("Cannot define circle: points are collinear (or nearly collinear).")
    cx = (d * e - b * f) / det
    cy = (-c * e + a * f) / det
    return float(cx), float(cy), float(np.hypot(x1 - cx, y1 - cy))


def roi_mask_from_circle(h: int, w: int, p1, p2, p3) -> np.ndarray:
    cx, cy, r = circle_from_three_points(p1, p2, p3)
    Y, X = np.ogrid[:h, :w]
    return ((X - cx) ** 2 + (Y - cy) ** 2) <= (r ** 2)


def annulus_mask(h, w, inner_p1, inner_p2, inner_p3, outer_p1, outer_p2, outer_p3) -> np.ndarray:
    return roi_mask_from_circle(h, w, outer_p1, outer_p2, outer_p3) & ~roi_mask_from_circle(h, w, inner_p1, inner_p2, inner_p3)


def bbox_from_mask(mask: np.ndarray, pad: int = 0) -> Tuple[int, int, int, int]:
    h, w = mask.shape[:2]
    ys, xs = np.where(mask)
    if ys.size == 0:
        return 0, h, 0, w
    return (int(max(0, ys.min() - int(pad))), int(min(h, ys.max() + int(pad) + 1)),
            int(max(0, xs.min() - int(pad))), int(min(w, xs.max() + int(pad) + 1)))


# ---------------------------------------------------------------- plan cache
_PLANS = {}
LAST_RECORD = None       # VistafTempRecord of the most recent main() (statistics reduced on the device, M1)
MASK_NAMES = ("sat", "dark", "light", "color_support", "color_ok", "source")


def _params(h: int, w: int) -> N.TempParams:
    if USE_INNER_CIRCLE:
        raise RuntimeError("USE_INNER_CIRCLE=True (annulus ROI) is not supported by the B200 path")
    if int(BLUR_KSIZE) != 5 or not FINAL_SMOOTH_ENABLE:
        raise RuntimeError("B200 path implements the reference configuration BLUR_KSIZE=5, FINAL_SMOOTH_ENABLE=True")
    P = N.TempParams()
    P.height, P.width = h, w
    P.roi_cx, P.roi_cy, P.roi_r = circle_from_three_points(OUTER_CIRCLE_P1, OUTER_CIRCLE_P2, OUTER_CIRCLE_P3)
    P.color_t_min, P.color_t_max, P.guard_band, P.switch_margin = COLOR_T_MIN, COLOR_T_MAX, COLOR_GUARD_BAND, SWITCH_MARGIN_C
    P.final_t_min, P.final_t_max = FINAL_T_MIN, FINAL_T_MAX
    P.band_radius, P.dc_exclusion = int(SEG_BAND_RADIUS), int(SEG_DC_EXCLUSION)
    P.peak_max_dy_frac, P.illum_sigma = SEG_PEAK_MAX_DY_FROM_CENTER, float(SEG_ILLUM_SIGMA)
    P.sat_thresh, P.sat_dilate = int(SAT_THRESH_GRAY), int(SAT_DILATE_KSIZE)
    P.close_kx, P.close_ky, P.open_kx, P.open_ky = POST_CLOSE_KX, POST_CLOSE_KY, POST_OPEN_KX, POST_OPEN_KY
    P.chroma_min, P.support_dilate = COLOR_CHROMA_MIN, int(COLOR_SUPPORT_DILATE)
    P.sigma_across, P.sigma_along = FINAL_SMOOTH_SIGMA_ACROSS, FINAL_SMOOTH_SIGMA_ALONG
    P.color_model = N.load_poly_model(os.path.basename(COLOR_MODEL_JSON))
    P.wide_model = N.load_poly_model(os.path.basename(WIDE_MODEL_JSON))
    if P.color_model.n_features != 3:
        raise RuntimeError(f"Color model must use ('L','a','b'), got {P.color_model.n_features} features")
    if P.wide_model.n_features != 4:
        raise RuntimeError(f"Wide model must use ('L','a','b','gray'), got {P.wide_model.n_features} features")
    g, c = N.lab_tables()
    P.lab_gamma[:] = g
    P.lab_cbrt[:] = c
    return P


class TempPlan:
    """Owns the opaque library plan + a torch workspace for `chunk` frames."""

    def __init__(self, h: int, w: int, chunk: int = 8, lanes: int = 1):
        self.torch = N.require_cuda()
        self.h, self.w, self.chunk = h, w, int(chunk)
        self.params = _params(h, w)
        self.handle = C.c_void_p()
        N.check(N.lib().vistaf_temp_plan_create(C.byref(self.params), C.byref(self.handle)), "temp_plan_create")
        self.lanes = max(1, int(lanes))               # see shape_ftp.FtpPlan: slices of a batch on concurrent streams
        self.lane_frames = -(-self.chunk // self.lanes)
        self.ws_bytes = int(N.lib().vistaf_temp_workspace_bytes(self.handle, self.lane_frames))
        self.lane_ws = [self.torch.empty(self.ws_bytes, dtype=self.torch.uint8, device="cuda") for _ in range(self.lanes)]
        self.lane_streams = [self.torch.cuda.Stream() for _ in range(self.lanes)] if self.lanes > 1 else []
        self.ws = self.lane_ws[0]
        self.active_lanes = self.lanes        # set to 1 to serialise a batch on the current stream (stage timing)

    def __del__(self):
        try:
            if self.handle:
                N.lib().vistaf_temp_plan_destroy(self.handle)
        except Exception:
            pass

    def workspace_plane(self, name: str):
        off, nb = C.c_size_t(), C.c_size_t()
        N.check(N.lib().vistaf_temp_workspace_plane(self.handle, self.lane_frames, name.encode(), C.byref(off), C.byref(nb)))
        return self.ws[off.value:off.value + nb.value]

    def run(self, frames_dev, masks: bool = True, out: Optional[dict] = None, ready=None) -> dict:
        """frames_dev: torch uint8 CUDA tensor [B,H,W,3] (BGR).  Returns dict of CUDA tensors."""
        t = self.torch
        assert frames_dev.is_cuda and frames_dev.dtype == t.uint8 and frames_dev.is_contiguous()
        B = frames_dev.shape[0]
        assert tuple(frames_dev.shape[1:]) == (self.h, self.w, 3)
        if out is None:
            out = {"fused": t.empty((B, self.h, self.w), dtype=t.float32, device="cuda"),
                   "final": t.empty((B, self.h, self.w), dtype=t.float32, device="cuda"),
                   "records": t.empty((B, 80), dtype=t.uint8, device="cuda")}
            if masks:
                for k in MASK_NAMES:
                    out[k] = t.empty((B, self.h, self.w), dtype=t.uint8, device="cuda")
        # output planes filled once by this plan keep their background (NaN / 0 outside the ROI bounding box): later calls on
        # the same buffers tell the library not to write those three quarters of every plane again
        bg_valid = int(bool(out.get("_background_written")))

        def call(lo, hi, ws, stream):
            O = N.TempOutputs()
            O.background_valid = bg_valid
            O.fused, O.final_map, O.records = out["fused"][lo:].data_ptr(), out["final"][lo:].data_ptr(), out["records"][lo:].data_ptr()
            for k in MASK_NAMES:
                setattr(O, k, out[k][lo:].data_ptr() if k in out else None)
            N.check(N.lib().vistaf_temp_batch(self.handle, frames_dev[lo:].data_ptr(), hi - lo, C.byref(O), ws.data_ptr(),
                                              self.ws_bytes, C.c_void_p(stream.cuda_stream)), "temp_batch")

        cur = t.cuda.current_stream()
        L = min(self.lanes, max(1, int(self.active_lanes)))
        if L == 1 or B < 2 * L:
            N.wait_ready(cur, ready, 0, B)
            call(0, B, self.ws, cur)
            out["_background_written"] = True
            return out
        per = -(-B // L)
        used = []
        for k in range(L):
            lo, hi = k * per, min(B, (k + 1) * per)
            if lo >= hi:
                break
            s = self.lane_streams[k]
            s.wait_stream(cur)
            # one library call per workspace-sized piece, each waiting only for the staged slices it covers
            for a in range(lo, hi, self.lane_frames):
                b = min(hi, a + self.lane_frames)
                N.wait_ready(s, ready, a, b)
                call(a, b, self.lane_ws[k], s)
            used.append(s)
        for s in used:
            cur.wait_stream(s)
        out["_background_written"] = True
        return out

    def roi_bbox(self):
        """(y0, y1, x0, x1) of the outer ROI in frame coordinates (no padding)."""
        return bbox_from_mask(roi_mask_from_circle(self.h, self.w, OUTER_CIRCLE_P1, OUTER_CIRCLE_P2, OUTER_CIRCLE_P3), pad=0)

    def redo_flagged(self, frames_dev, out) -> np.ndarray:
        """T10 (saturated frames) and the warpAffine path of T12 (tilted carrier) run on at most an eighth of a chunk's
        frames; a frame beyond that comes back with ST_RERUN_ALONE.  Here such frames are run again alone (a one-frame
        chunk always has a slot) and patched into `out`.  Reads the records, i.e. synchronises; returns them as numpy."""
        rec = records_to_numpy(out["records"])
        redo = [i for i, r in enumerate(rec) if int(r["status"]) & N.ST_RERUN_ALONE]
        if redo:
            one = get_plan(self.h, self.w, 1)
            masks = all(k in out for k in MASK_NAMES)
            for i in redo:
                o1 = one.run(frames_dev[i:i + 1].contiguous(), masks=masks)
                for k, v in o1.items():
                    if hasattr(v, "shape"):
                        out[k][i:i + 1].copy_(v)
            rec = records_to_numpy(out["records"])
        return rec


def get_plan(h: int, w: int, chunk: int = 8) -> TempPlan:
    key = (h, w, chunk, OUTER_CIRCLE_P1, OUTER_CIRCLE_P2, OUTER_CIRCLE_P3, SEG_BAND_RADIUS, SEG_DC_EXCLUSION,
           SEG_ILLUM_SIGMA, SAT_THRESH_GRAY, SAT_DILATE_KSIZE, COLOR_CHROMA_MIN, FINAL_SMOOTH_SIGMA_ACROSS,
           FINAL_SMOOTH_SIGMA_ALONG, FINAL_T_MIN, FINAL_T_MAX, COLOR_T_MIN, COLOR_T_MAX)
    if key not in _PLANS:
        _PLANS[key] = TempPlan(h, w, chunk)
    return _PLANS[key]


def records_to_numpy(rec_dev) -> np.ndarray:
    return rec_dev.cpu().numpy().view(N.TEMP_RECORD_DTYPE).reshape(-1)


def run_batch(frames_bgr, chunk: int = 8, masks: bool = True) -> dict:
    """Batched temperature sensing.  frames_bgr: numpy uint8 [B,H,W,3] (host) or torch CUDA tensor.
    Host input is staged through pinned memory; outputs stay on the device (torch tensors) except
    `records`, returned as a numpy structured array (one VistafTempRecord per frame)."""
    t = N.require_cuda()
    if isinstance(frames_bgr, np.ndarray):
        frames_bgr = t.from_numpy(np.ascontiguousarray(frames_bgr)).pin_memory().cuda(non_blocking=True)
    plan = get_plan(int(frames_bgr.shape[1]), int(frames_bgr.shape[2]), chunk)
    out = plan.run(frames_bgr, masks=masks)
    out["records_np"] = plan.redo_flagged(frames_bgr, out)     # frames that found no T10 / T12 slot are redone alone
    return out


# ---------------------------------------------------------------- reference-compatible entry point
def main() -> None:
    global LAST_RECORD
    import cv2
    os.makedirs(OUTPUT_DIR, exist_ok=True)
    if not os.path.exists(INPUT_IMAGE_PATH):
        raise RuntimeError(f"INPUT_IMAGE_PATH not found:\n  {os.path.abspath(INPUT_IMAGE_PATH)}")
    img = cv2.imread(INPUT_IMAGE_PATH, cv2.IMREAD_COLOR)
    if img is None:
        raise RuntimeError(f"Could not read image: {INPUT_IMAGE_PATH}")
    h, w = img.shape[:2]
    roi_outer = roi_mask_from_circle(h, w, OUTER_CIRCLE_P1, OUTER_CIRCLE_P2, OUTER_CIRCLE_P3)
    crop_bbox = bbox_from_mask(roi_outer, pad=int(CROP_PAD_PX)) if bool(CROP_OUTPUT_TO_OUTER_ROI) else (0, h, 0, w)
    out = run_batch(img[None], chunk=1, masks=True)
    rec = out["records_np"][0]
    LAST_RECORD = rec
    if rec["status"] & N.ST_EMPTY_ROI:
        raise RuntimeError("ROI became empty after saturation exclusion. Lower SAT_THRESH_GRAY / dilation.")
    if rec["status"] & N.ST_NO_PEAK:
        raise RuntimeError("Could not find FFT peaks for stripe carrier.")
    y0, y1, x0, x1 = crop_bbox
    host = {k: out[k][0].cpu().numpy() for k in ("fused", "final") + MASK_NAMES}
    roi_eff = roi_outer & ~(host["sat"] > 0)
    cv2.imwrite(os.path.join(OUTPUT_DIR, "mask_roi.png"), roi_outer[y0:y1, x0:x1].astype(np.uint8) * 255)
    cv2.imwrite(os.path.join(OUTPUT_DIR, "mask_roi_eff.png"), roi_eff[y0:y1, x0:x1].astype(np.uint8) * 255)
    for k in ("sat", "dark", "light", "color_support", "color_ok"):
        cv2.imwrite(os.path.join(OUTPUT_DIR, f"mask_{k}.png"), (host[k][y0:y1, x0:x1] > 0).astype(np.uint8) * 255)
    cv2.imwrite(os.path.join(OUTPUT_DIR, "debug_fused_source.png"), host["source"][y0:y1, x0:x1])
    np.save(os.path.join(OUTPUT_DIR, "temperature_map_fused.npy"), host["fused"].astype(np.float32))
    np.save(os.path.join(OUTPUT_DIR, "temperature_map_final.npy"), host["final"].astype(np.float32))
    roi_name = "OUTER ROI" if bool(CROP_OUTPUT_TO_OUTER_ROI) else "ROI"
    print(f"Final temperature min/max in {roi_name}: {rec['min_C']:.3f} / {rec['max_C']:.3f} °C")
    print(f"Segmentation: peak=({rec['peak_x']},{rec['peak_y']}) phi0={rec['phi0']:.6f} "
          f"roi_eff_pixels={rec['roi_eff_pixels']} sat_pixels={rec['sat_pixels']} dark_pixels={rec['dark_pixels']}")
    print(f"Saved outputs to: {os.path.abspath(OUTPUT_DIR)}")


if __name__ == "__main__":
    main()
