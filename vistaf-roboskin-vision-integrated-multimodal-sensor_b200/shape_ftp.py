"""shape_ftp — drop-in for the reference's Code/shape_ftp.py, computed on B200.

Same module-level configuration names (shape_ftp.py:16-218), same `main(...)` signature and return
dict (:1428-1438, :2029-2037), same exported files (height_map_crop.npy / _full.npy / _crop.csv /
_bundle.npz, :260-310, :1875-1932), same RuntimeErrors / `None` result.  The numeric body
(:1632-1865: demodulation, reliable mask, unwrap, detrend, smoothing, frontier, phase->mm, blob filter)
runs as hand-written sm_100a kernels behind libvistaf (`vistaf_ftp_set_reference`, `vistaf_ftp_batch`).
`run_batch` / `FtpPlan` are the batched entry points beside the reference-compatible single-pair `main`:
one reference frame, many deformed frames.

Alignment front-end (phase correlation + ECC, :529-587): SURVEY.md section 8(f) row 1 — performed with
the same cv2 host calls as the reference for real pairs (APPLY_GLOBAL_SHIFT / USE_ECC_CROP_ALIGNMENT),
disabled for pre-registered synthetic batches.  matplotlib figures are not produced (headless).
"""
from __future__ import annotations

import ctypes as C
import json
import os
from typing import Optional

import numpy as np

import _native as N

# ---------------------------------------------------------------- USER CONFIG (shape_ftp.py:16-218)
REFERENCE_PATH = "./Final_demos_images/FINAL_reference.jpg"
DEFORMED_PATH = "./Final_demos_images/FINAL_E_deformed.jpg"
OUTPUT_DIR = "./Force/FTP/output"
CALIBRATION_JSON = os.path.join(N.CAL_DIR, "phase_to_height_calibration_model.json")

FFT_SIDEBAND_METHOD = "patch_shift"
PATCH_HALF_WIDTH_BINS = 10
PATCH_WINDOW = "hann"
DC_EXCLUSION = 10
FFT_PAD_PX = 96
PRE_BLUR_SIGMA_PX = 1.5
USE_FIXED_ROI = True
OUTER_CIRCLE_P1 = (1873, 1703)
OUTER_CIRCLE_P2 = (1599, 707)
OUTER_CIRCLE_P3 = (2575, 950)
OUTPUT_HEIGHT_IN_MM = True
MM_KEEP_INDENTATION_NEGATIVE = False
FILTER_SMALL_CONTACT_BLOBS = True
CONTACT_BLOB_MIN_PEAK_MM = 0.1
CONTACT_BLOB_MIN_PEAK_REL_FRAC = 1.0 / 3.0
ROI_ERODE_PX = 0
USE_CIRCULAR_APODIZATION = True
APOD_TAPER_PX = 120
AMP_VALID_PERCENTILE = 25.0
QUALITY_SMOOTH_SIGMA_PX = 6.0
RELIABLE_KEEP_LARGEST_CC = True
RELIABLE_EDGE_MARGIN_PX = 6
POLY_ORDER = 2
RELIABLE_SMOOTH_SIGMA_PX = 2.5
FRONTIER_ZERO_ENABLE = True
FRONTIER_ZERO_BAND_PX = 200
ILLUM_SIGMA_PX = 45
VALID_CLOSE_KERNEL = 7
BAD_PIXEL_ENABLE = True
BAD_INTENSITY_PERCENTILE = 99.9
BAD_GRADIENT_PERCENTILE = 99.7
BAD_DILATE_KSIZE = 5
BAD_INPAINT_RADIUS = 3
FILL_INTERNAL_HOLES_IN_RELIABLE = True
HOLE_NEIGHBORHOOD_PX = 11
HOLE_KNOWN_FRACTION = 0.70
HOLE_MIN_DIST_FROM_RELIABLE_EDGE_PX = 4
INPAINT_RADIUS = 5
INPAINT_METHOD = "telea"
CONTACT_CORE_PERCENTILE = 8.0
CONTACT_PERCENTILE = 92
DILATE_KERNEL_SIZE = 15
DILATE_ITERS = 2
MIN_CONTACT_FRAC = 0.002
MAX_CONTACT_FRAC = 0.40
UNRELIABLE_SMOOTH_SIGMA_PX = 9.0
EXPORT_HEIGHTMAP_FILES = True
HEIGHTMAP_EXPORT_BASENAME = "height_map"
HEIGHTMAP_EXPORT_SAVE_CROP_CSV = True
HEIGHTMAP_EXPORT_SAVE_FULL_CSV = False
DEBUG = True
DEBUG_LOG_TO_FILE = True
DEBUG_N_FFT_PEAKS = 12
BATCH_MODE = False
SAVE_SUMMARY_FIGURES = True
APPLY_GLOBAL_SHIFT = True
USE_ECC_CROP_ALIGNMENT = True
ECC_WARP_MODE = "euclidean"
ECC_ITERS = 300
ECC_EPS = 1e-7
ECC_GAUSS_FILT = 5
PEAK_MAX_DY_FROM_CENTER = 0.12
LOCK_CARRIER_TO_REFERENCE = True

# force constants that travel in the same params struct (force_sensor.py:27-36)
FORCE_CALIBRATION_JSON = os.path.join(N.CAL_DIR, "height_to_force_calibration_model.json")
GRATING_PITCH_MM = 2.0
DEPTH_EPS_MM = 0.01


def log(msg: str) -> None:
    if DEBUG:
        print(msg)


# ---------------------------------------------------------------- geometry (shape_ftp.py:383-414, 1494-1528)
def circle_from_3_points(p1, p2, p3):
    (x1, y1), (x2, y2), (x3, y3) = p1, p2, p3
    A = np.array([[2 * (x2 - x1), 2 * (y2 - y1)], [2 * (x3 - x1), 2 * (y3 - y1)]], float)
    b = np.array([x2 * x2 + y2 * y2 - x1 * x1 - y1 * y1, x3 * x3 + y3 * y3 - x1 * x1 - y1 * y1], float)
    cx, cy = np.linalg.solve(A, b)
    return int(round(cx)), int(round(cy)), int(round(float(np.hypot(cx - x1, cy - y1))))


def crop_geometry(H: int, W: int) -> dict:
    """Crop box and local ROI circle from the fixed 3-point circle (shape_ftp.py:1494-1516)."""
    cx, cy, r = circle_from_3_points(OUTER_CIRCLE_P1, OUTER_CIRCLE_P2, OUTER_CIRCLE_P3)
    x1, x2, y1, y2 = max(0, cx - r), min(W, cx + r), max(0, cy - r), min(H, cy + r)
    h, w = y2 - y1, x2 - x1
    cxl, cyl = cx - x1, cy - y1
    rl = int(min(r, cxl, cyl, w - 1 - cxl, h - 1 - cyl))
    return dict(x1=x1, x2=x2, y1=y1, y2=y2, h=h, w=w, cx=cxl, cy=cyl, r=rl, cx_full=cx, cy_full=cy, r_full=r)


def load_calibration(path: str):
    """shape_ftp.py:672-680."""
    cal = json.load(open(path))
    if "best_model" not in cal:
        raise ValueError("Calibration JSON missing 'best_model'")
    return cal["best_model"], bool(cal.get("use_negated_height_for_fit", True))


# ---------------------------------------------------------------- plan
def _params(geom: dict, calibration_json: str, force_json: str) -> N.FtpParams:
    if FFT_SIDEBAND_METHOD != "patch_shift" or not LOCK_CARRIER_TO_REFERENCE or int(POLY_ORDER) != 2:
        raise RuntimeError("B200 path implements the reference configuration: patch_shift sideband, "
                           "carrier locked to the reference, POLY_ORDER=2")
    if not (BAD_PIXEL_ENABLE and FRONTIER_ZERO_ENABLE and FILTER_SMALL_CONTACT_BLOBS and RELIABLE_KEEP_LARGEST_CC):
        raise RuntimeError("B200 path implements the reference configuration (bad-pixel repair, frontier taper, "
                           "blob filter and largest-component selection enabled)")
    model, use_neg = load_calibration(calibration_json)
    force_model = json.load(open(force_json))
    if "best_model" not in force_model:
        raise ValueError("Force calibration JSON missing 'best_model'")
    P = N.FtpParams()
    P.crop_h, P.crop_w, P.cx, P.cy, P.r = geom["h"], geom["w"], geom["cx"], geom["cy"], geom["r"]
    P.roi_erode_px, P.apod_taper = int(ROI_ERODE_PX), int(APOD_TAPER_PX) if USE_CIRCULAR_APODIZATION else 0
    P.patch_half, P.dc_exclusion, P.n_peaks, P.fft_pad = int(PATCH_HALF_WIDTH_BINS), int(DC_EXCLUSION), int(DEBUG_N_FFT_PEAKS), int(FFT_PAD_PX)
    P.peak_max_dy_frac = PEAK_MAX_DY_FROM_CENTER
    P.illum_sigma, P.pre_blur_sigma, P.quality_sigma = ILLUM_SIGMA_PX, PRE_BLUR_SIGMA_PX, QUALITY_SMOOTH_SIGMA_PX
    P.smooth_sigma, P.unrel_sigma, P.amp_percentile = RELIABLE_SMOOTH_SIGMA_PX, UNRELIABLE_SMOOTH_SIGMA_PX, AMP_VALID_PERCENTILE
    P.close_kernel, P.edge_margin = int(VALID_CLOSE_KERNEL), int(RELIABLE_EDGE_MARGIN_PX)
    P.bad_int_pct, P.bad_grad_pct = BAD_INTENSITY_PERCENTILE, BAD_GRADIENT_PERCENTILE
    P.bad_dilate, P.bad_inpaint_radius = int(BAD_DILATE_KSIZE), int(BAD_INPAINT_RADIUS)
    P.core_pct, P.contact_pct = CONTACT_CORE_PERCENTILE, float(CONTACT_PERCENTILE)
    P.dilate_kernel, P.dilate_iters = int(DILATE_KERNEL_SIZE), int(DILATE_ITERS)
    P.min_contact_frac, P.max_contact_frac, P.frontier_band = MIN_CONTACT_FRAC, MAX_CONTACT_FRAC, float(FRONTIER_ZERO_BAND_PX)
    P.use_negated_height = int(use_neg)
    P.blob_min_peak_mm, P.blob_rel_frac = float(CONTACT_BLOB_MIN_PEAK_MM), float(CONTACT_BLOB_MIN_PEAK_REL_FRAC)
    P.height_model = N.curve_model(model)
    P.force_model = N.curve_model(force_model["best_model"])
    P.grating_pitch_mm, P.depth_eps_mm = float(GRATING_PITCH_MM), float(DEPTH_EPS_MM)
    if str(INPAINT_METHOD).lower() != "telea":
        raise RuntimeError("B200 path implements INPAINT_METHOD='telea'")
    P.fill_internal_holes, P.hole_neighborhood = int(bool(FILL_INTERNAL_HOLES_IN_RELIABLE)), int(HOLE_NEIGHBORHOOD_PX)
    P.hole_known_frac, P.hole_min_dist = float(HOLE_KNOWN_FRACTION), float(HOLE_MIN_DIST_FROM_RELIABLE_EDGE_PX)
    P.hole_inpaint_radius = int(INPAINT_RADIUS)
    return P


MASK_NAMES = ("reliable", "output_reliable", "contact_dilated", "contact_kept", "hole_candidates")


class FtpPlan:
    """Owns the opaque library plan, its torch workspace for `chunk` frames and the cached reference frame."""

    def __init__(self, H: int, W: int, chunk: int = 8, calibration_json: Optional[str] = None,
                 force_json: Optional[str] = None, lanes: int = 1, geom: Optional[dict] = None):
        """geom: explicit crop geometry {h, w, cx, cy, r[, x1, y1]} for callers that bring their own gray crops (the
        3904^2 / 4096^2-FFT streaming geometry of BASELINE.json configs[4]); default: the fixed 3-point circle in an
        H x W camera frame (shape_ftp.py:1494-1516)."""
        self.torch = N.require_cuda()
        self.H, self.W, self.chunk = int(H), int(W), int(chunk)
        self.geom = dict(geom) if geom is not None else crop_geometry(H, W)
        self.h, self.w = self.geom["h"], self.geom["w"]
        self.params = _params(self.geom, calibration_json or CALIBRATION_JSON, force_json or FORCE_CALIBRATION_JSON)
        self.handle = C.c_void_p()
        N.check(N.lib().vistaf_ftp_plan_create(C.byref(self.params), C.byref(self.handle)), "ftp_plan_create")
        # `lanes` > 1: a batch is cut into that many contiguous slices, each marched through the library on its own
        # stream with its own workspace, so one slice's latency-bound stages (Telea march, selection tails, small
        # per-frame kernels) overlap the other slices' bandwidth-bound sweeps.  The plan itself is read-only there.
        self.lanes = max(1, int(lanes))
        self.lane_frames = -(-self.chunk // self.lanes)
        self.ws_bytes = int(N.lib().vistaf_ftp_workspace_bytes(self.handle, self.lane_frames))
        self.lane_ws = [self.torch.empty(self.ws_bytes, dtype=self.torch.uint8, device="cuda") for _ in range(self.lanes)]
        self.lane_streams = [self.torch.cuda.Stream() for _ in range(self.lanes)] if self.lanes > 1 else []
        self.ws = self.lane_ws[0]
        # the reference frame is demodulated in its own one-frame workspace, so it can run beside a batch's begin()
        self.ref_ws_bytes = int(N.lib().vistaf_ftp_workspace_bytes(self.handle, 1))
        self.ref_ws = self.torch.empty(self.ref_ws_bytes, dtype=self.torch.uint8, device="cuda")
        self._begun = None
        self.active_lanes = self.lanes        # set to 1 to serialise a batch on the current stream (stage timing)
        self.ref_info = None

    def __del__(self):
        try:
            if self.handle:
                N.lib().vistaf_ftp_plan_destroy(self.handle)
        except Exception:
            pass

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    def crop_gray(self, frames_dev, out=None):
        """cv2 BGR2GRAY of the ROI crop (shape_ftp.py:1508-1511): uint8 [B,H,W,3] -> uint8 [B,h,w] (into `out` if given)."""
        t, g = self.torch, self.geom
        assert frames_dev.is_cuda and frames_dev.dtype == t.uint8 and frames_dev.is_contiguous()
        B = frames_dev.shape[0]
        gray = out if out is not None else t.empty((B, self.h, self.w), dtype=t.uint8, device="cuda")
        N.check(N.lib().vistaf_crop_gray(frames_dev.data_ptr(), B, self.H, self.W, g["x1"], g["y1"], self.h, self.w,
                                         gray.data_ptr(), self._stream()), "crop_gray")
        return gray

    def set_reference(self, ref_gray_dev, sync_info: bool = True):
        """ref_gray_dev: uint8 CUDA [h,w] crop of the reference frame.  sync_info=True waits for the carrier and returns it
        (VistafFtpReferenceInfo; raises when no carrier peak is found); False leaves the carrier on the device -- the
        call is then asynchronous, and frames of a reference without a carrier come back with ST_NO_PEAK."""
        assert tuple(ref_gray_dev.shape) == (self.h, self.w) and ref_gray_dev.is_contiguous()
        info = N.FtpReferenceInfo() if sync_info else None
        N.check(N.lib().vistaf_ftp_set_reference(self.handle, ref_gray_dev.data_ptr(), self.ref_ws.data_ptr(), self.ref_ws_bytes,
                                                 C.byref(info) if sync_info else None, self._stream()), "ftp_set_reference")
        self.ref_info = info if sync_info else True
        return info

    def masks(self):
        t = self.torch
        roi = t.empty((self.h, self.w), dtype=t.uint8, device="cuda")
        circ = t.empty_like(roi)
        apo = t.empty((self.h, self.w), dtype=t.float32, device="cuda")
        N.check(N.lib().vistaf_ftp_masks(self.handle, roi.data_ptr(), circ.data_ptr(), apo.data_ptr(), self._stream()))
        return roi, circ, apo

    def workspace_plane(self, name: str, chunk: Optional[int] = None):
        off, nb = C.c_size_t(), C.c_size_t()
        N.check(N.lib().vistaf_ftp_workspace_plane(self.handle, chunk or self.lane_frames, name.encode(), C.byref(off), C.byref(nb)))
        return self.ws[off.value:off.value + nb.value]

    def _slices(self, B):
        L = min(self.lanes, max(1, int(self.active_lanes)))
        if L == 1 or B < 2 * L:
            return [(0, B)]
        per = -(-B // L)
        return [(k * per, min(B, (k + 1) * per)) for k in range(L) if k * per < B]

    def begin(self, def_gray_dev=None, frames_dev=None, ready=None):
        """Issue the reference-independent front of the chain (bad pixels, Telea fill, normalisation) for a batch, one
        slice per lane; may be called BEFORE set_reference.  Input: cropped gray frames, or full BGR frames (cropped here,
        slice by slice on the lane streams; `ready` = [(lo, hi, event)] marks when slices of frames_dev become valid).
        Returns the gray tensor to hand to run().  When a slice does not fit its lane workspace in one chunk nothing is
        issued ahead and run() does the whole chain."""
        t = self.torch
        cur = t.cuda.current_stream()
        B = int((def_gray_dev if def_gray_dev is not None else frames_dev).shape[0])
        sl = self._slices(B)
        fits = all(hi - lo <= self.lane_frames for lo, hi in sl)
        if def_gray_dev is None:
            if getattr(self, "_gray", None) is None or self._gray.shape[0] != B:
                self._gray = t.empty((B, self.h, self.w), dtype=t.uint8, device="cuda")
            gray = self._gray
            if not fits or len(sl) == 1:
                N.wait_ready(cur, ready, 0, B)
                gray = self.crop_gray(frames_dev)
        else:
            gray = def_gray_dev
        if not fits:
            return gray
        for k, (lo, hi) in enumerate(sl):
            s = self.lane_streams[k] if len(sl) > 1 else cur
            if s is not cur:
                s.wait_stream(cur)
                if def_gray_dev is None:
                    N.wait_ready(s, ready, lo, hi)
                    with t.cuda.stream(s):
                        g = self.geom
                        N.check(N.lib().vistaf_crop_gray(frames_dev[lo:].data_ptr(), hi - lo, self.H, self.W, g["x1"], g["y1"], self.h,
                                                         self.w, gray[lo:].data_ptr(), C.c_void_p(s.cuda_stream)), "crop_gray")
            N.check(N.lib().vistaf_ftp_batch_begin(self.handle, gray[lo:].data_ptr(), hi - lo, self.lane_ws[k].data_ptr(),
                                                   self.ws_bytes, C.c_void_p(s.cuda_stream)), "ftp_batch_begin")
        self._begun = (gray.data_ptr(), B, sl)
        return gray

    def run(self, def_gray_dev, masks: bool = True, out: Optional[dict] = None) -> dict:
        """def_gray_dev: uint8 CUDA [B,h,w] (aligned crops).  Returns dict of CUDA tensors.  After begin() on the same
        tensor only the reference-dependent rest of the chain is issued."""
        t = self.torch
        if self.ref_info is None:
            raise RuntimeError("FtpPlan.run: call set_reference first")
        assert def_gray_dev.is_cuda and def_gray_dev.dtype == t.uint8 and def_gray_dev.is_contiguous()
        B = int(def_gray_dev.shape[0])
        assert tuple(def_gray_dev.shape[1:]) == (self.h, self.w)
        if out is None:
            out = {"height_mm": t.empty((B, self.h, self.w), dtype=t.float32, device="cuda"),
                   "records": t.empty((B, 104), dtype=t.uint8, device="cuda")}
            if masks:
                for k in MASK_NAMES:
                    out[k] = t.empty((B, self.h, self.w), dtype=t.uint8, device="cuda")
        begun = self._begun is not None and self._begun[:2] == (def_gray_dev.data_ptr(), B)
        sl = self._begun[2] if begun else self._slices(B)
        self._begun = None

        def call(lo, hi, ws, stream):
            O = N.FtpOutputs()
            O.height_mm, O.records = out["height_mm"][lo:].data_ptr(), out["records"][lo:].data_ptr()
            for k in MASK_NAMES:
                setattr(O, k, out[k][lo:].data_ptr() if k in out else None)
            fn = N.lib().vistaf_ftp_batch_finish if begun else N.lib().vistaf_ftp_batch
            N.check(fn(self.handle, def_gray_dev[lo:].data_ptr(), hi - lo, C.byref(O), ws.data_ptr(),
                       self.ws_bytes, C.c_void_p(stream.cuda_stream)), "ftp_batch")

        cur = t.cuda.current_stream()
        if len(sl) == 1:
            call(0, B, self.lane_ws[0], cur)
            return out
        used = []
        for k, (lo, hi) in enumerate(sl):
            s = self.lane_streams[k]
            if not begun:
                s.wait_stream(cur)             # (a begun slice is already ordered after the inputs on its lane stream)
            call(lo, hi, self.lane_ws[k], s)
            used.append(s)
        for s in used:
            cur.wait_stream(s)
        return out


_PLANS = {}


def _config_key():
    return (OUTER_CIRCLE_P1, OUTER_CIRCLE_P2, OUTER_CIRCLE_P3, PATCH_HALF_WIDTH_BINS, DC_EXCLUSION, FFT_PAD_PX,
            PRE_BLUR_SIGMA_PX, ROI_ERODE_PX, APOD_TAPER_PX, AMP_VALID_PERCENTILE, QUALITY_SMOOTH_SIGMA_PX,
            RELIABLE_EDGE_MARGIN_PX, RELIABLE_SMOOTH_SIGMA_PX, FRONTIER_ZERO_BAND_PX, ILLUM_SIGMA_PX, VALID_CLOSE_KERNEL,
            BAD_INTENSITY_PERCENTILE, BAD_GRADIENT_PERCENTILE, BAD_DILATE_KSIZE, BAD_INPAINT_RADIUS,
            CONTACT_CORE_PERCENTILE, CONTACT_PERCENTILE, DILATE_KERNEL_SIZE, DILATE_ITERS, UNRELIABLE_SMOOTH_SIGMA_PX,
            CONTACT_BLOB_MIN_PEAK_MM, CONTACT_BLOB_MIN_PEAK_REL_FRAC, CALIBRATION_JSON, FORCE_CALIBRATION_JSON,
            FILL_INTERNAL_HOLES_IN_RELIABLE, HOLE_NEIGHBORHOOD_PX, HOLE_KNOWN_FRACTION, HOLE_MIN_DIST_FROM_RELIABLE_EDGE_PX,
            INPAINT_RADIUS)


def get_plan(H: int, W: int, chunk: int = 8) -> FtpPlan:
    key = (H, W, chunk) + _config_key()
    if key not in _PLANS:
        _PLANS[key] = FtpPlan(H, W, chunk)
    return _PLANS[key]


def records_to_numpy(rec_dev) -> np.ndarray:
    return rec_dev.cpu().numpy().view(N.FTP_RECORD_DTYPE).reshape(-1)


def run_batch(reference_bgr, deformed_bgr, chunk: int = 8, masks: bool = True) -> dict:
    """Batched FTP shape + force: one reference frame [H,W,3], deformed frames [B,H,W,3] (numpy uint8 on the
    host or torch CUDA tensors), already registered to the reference.  Outputs stay on the device except
    `records_np` (one VistafFtpRecord per frame: force, volume, area, max depth, scale, status ...)."""
    t = N.require_cuda()

    def dev(a):
        if isinstance(a, np.ndarray):
            return t.from_numpy(np.ascontiguousarray(a)).pin_memory().cuda(non_blocking=True)
        return a
    ref, de = dev(reference_bgr), dev(deformed_bgr)
    plan = get_plan(int(de.shape[1]), int(de.shape[2]), chunk)
    plan.set_reference(plan.crop_gray(ref[None] if ref.dim() == 3 else ref)[0])
    out = plan.run(plan.crop_gray(de), masks=masks)
    out["records_np"] = records_to_numpy(out["records"])
    out["geom"] = plan.geom
    return out


# ---------------------------------------------------------------- alignment front-end on the host (shape_ftp.py:529-587)
def _align_host(ref_bgr, def_bgr, geom):
    """Same cv2 calls as the reference (phaseCorrelate on sigma-7 blurred frames, translation warp, crop,
    euclidean ECC on sigma-5 blurred crops, inverse warp); cv2.error from ECC is swallowed as in :576-578."""
    import cv2
    H, W = ref_bgr.shape[:2]
    x1, x2, y1, y2 = geom["x1"], geom["x2"], geom["y1"], geom["y2"]
    if APPLY_GLOBAL_SHIFT:
        rg = cv2.cvtColor(ref_bgr, cv2.COLOR_BGR2GRAY).astype(np.float32)
        dg = cv2.cvtColor(def_bgr, cv2.COLOR_BGR2GRAY).astype(np.float32)
        hann = cv2.createHanningWindow((W, H), cv2.CV_32F)
        shift, _ = cv2.phaseCorrelate(np.float32(cv2.GaussianBlur(rg, (0, 0), 7)), np.float32(cv2.GaussianBlur(dg, (0, 0), 7)), hann)
        M = np.array([[1, 0, shift[0]], [0, 1, shift[1]]], np.float32)
        def_bgr = cv2.warpAffine(def_bgr, M, (W, H), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT)
    ref_gray = cv2.cvtColor(ref_bgr[y1:y2, x1:x2], cv2.COLOR_BGR2GRAY)
    def_gray = cv2.cvtColor(def_bgr[y1:y2, x1:x2], cv2.COLOR_BGR2GRAY)
    if USE_ECC_CROP_ALIGNMENT:
        h, w = ref_gray.shape
        Y, X = np.ogrid[:h, :w]
        circ = ((X - geom["cx"]) ** 2 + (Y - geom["cy"]) ** 2) <= geom["r"] ** 2
        r = cv2.GaussianBlur(ref_gray.astype(np.float32) / 255.0, (0, 0), ECC_GAUSS_FILT)
        m = cv2.GaussianBlur(def_gray.astype(np.float32) / 255.0, (0, 0), ECC_GAUSS_FILT)
        warp = np.eye(2, 3, dtype=np.float32)
        crit = (cv2.TERM_CRITERIA_EPS | cv2.TERM_CRITERIA_COUNT, int(ECC_ITERS), float(ECC_EPS))
        try:
            _, warp = cv2.findTransformECC(r, m, warp, cv2.MOTION_EUCLIDEAN, crit, inputMask=circ.astype(np.uint8) * 255,
                                           gaussFiltSize=1)
            def_gray = cv2.warpAffine(def_gray, warp, (w, h), flags=cv2.INTER_LINEAR | cv2.WARP_INVERSE_MAP,
                                      borderMode=cv2.BORDER_REFLECT)
        except cv2.error as e:
            log(f"[ALIGN][ECC] failed: {e}")
    return ref_gray, def_gray


def export_heightmap_files(output_dir, basename, height_crop, height_full=None, crop_masks=None, full_masks=None,
                           meta=None, save_crop_csv=True, save_full_csv=False):
    """Same files, names and array keys as shape_ftp.py:260-310."""
    os.makedirs(output_dir, exist_ok=True)
    np.save(os.path.join(output_dir, f"{basename}_crop.npy"), height_crop.astype(np.float32))
    if height_full is not None:
        np.save(os.path.join(output_dir, f"{basename}_full.npy"), height_full.astype(np.float32))
    if save_crop_csv:
        np.savetxt(os.path.join(output_dir, f"{basename}_crop.csv"), height_crop.astype(np.float32), delimiter=",", fmt="%.9g")
    if save_full_csv and height_full is not None:
        np.savetxt(os.path.join(output_dir, f"{basename}_full.csv"), height_full.astype(np.float32), delimiter=",", fmt="%.9g")
    bundle = {"height_crop": height_crop.astype(np.float32)}
    if height_full is not None:
        bundle["height_full"] = height_full.astype(np.float32)
    for prefix, d in (("crop_", crop_masks), ("full_", full_masks), ("meta_", meta)):
        for k, v in (d or {}).items():
            bundle[prefix + k] = np.asarray(v)
    np.savez_compressed(os.path.join(output_dir, f"{basename}_bundle.npz"), **bundle)


# ---------------------------------------------------------------- reference-compatible entry point
def main(reference_path=None, deformed_path=None, output_dir=None, calibration_json=None, batch_mode=False,
         save_summary_figures=True, export_heightmaps=None, debug=None, return_results=False):
    global REFERENCE_PATH, DEFORMED_PATH, OUTPUT_DIR, CALIBRATION_JSON
    global BATCH_MODE, SAVE_SUMMARY_FIGURES, EXPORT_HEIGHTMAP_FILES, DEBUG, DEBUG_LOG_TO_FILE
    import cv2
    if reference_path is not None:
        REFERENCE_PATH = reference_path
    if deformed_path is not None:
        DEFORMED_PATH = deformed_path
    if output_dir is not None:
        OUTPUT_DIR = output_dir
    if calibration_json is not None:
        CALIBRATION_JSON = calibration_json
    BATCH_MODE, SAVE_SUMMARY_FIGURES = bool(batch_mode), bool(save_summary_figures)
    if export_heightmaps is not None:
        EXPORT_HEIGHTMAP_FILES = bool(export_heightmaps)
    if debug is not None:
        DEBUG = bool(debug)
        if not DEBUG:
            DEBUG_LOG_TO_FILE = False
    os.makedirs(OUTPUT_DIR, exist_ok=True)
    ref_bgr, def_bgr = cv2.imread(REFERENCE_PATH), cv2.imread(DEFORMED_PATH)
    if ref_bgr is None or def_bgr is None:
        raise RuntimeError("Could not read one or both images. Check file paths.")
    if ref_bgr.shape != def_bgr.shape:
        raise RuntimeError("Reference and deformed images have different sizes.")
    H, W = ref_bgr.shape[:2]
    t = N.require_cuda()
    plan = get_plan(H, W, chunk=1)
    g = plan.geom
    ref_gray, def_gray = _align_host(ref_bgr, def_bgr, g)
    plan.set_reference(t.from_numpy(np.ascontiguousarray(ref_gray)).cuda())
    out = plan.run(t.from_numpy(np.ascontiguousarray(def_gray))[None].cuda(), masks=True)
    rec = records_to_numpy(out["records"])[0]
    if rec["status"] & N.ST_EMPTY_ROI:
        log("[ERROR] reliable mask empty.")
        return None
    height = out["height_mm"][0].cpu().numpy()
    m = {k: out[k][0].cpu().numpy() > 0 for k in MASK_NAMES}
    roi_dev, circ_dev, _ = plan.masks()
    roi_eroded, circ_mask = roi_dev.cpu().numpy() > 0, circ_dev.cpu().numpy() > 0
    if EXPORT_HEIGHTMAP_FILES:
        x1, x2, y1, y2 = g["x1"], g["x2"], g["y1"], g["y2"]

        def full(a, fill=False, dtype=bool):
            f = np.full((H, W), fill, dtype)
            f[y1:y2, x1:x2] = a
            return f
        crop_masks = {"roi_eroded": roi_eroded, "reliable": m["reliable"], "output_reliable": m["output_reliable"],
                      "circ_mask": circ_mask, "contact_kept_by_depth": m["contact_kept"], "hole_candidates": m["hole_candidates"],
                      "contact_dilated": m["contact_dilated"]}
        export_heightmap_files(OUTPUT_DIR, HEIGHTMAP_EXPORT_BASENAME, height, full(height, np.nan, np.float32), crop_masks,
                               {k: full(v) for k, v in crop_masks.items()},
                               {"crop_x1": np.int32(x1), "crop_y1": np.int32(y1), "crop_x2": np.int32(x2), "crop_y2": np.int32(y2),
                                "roi_center_x_full": np.int32(g["cx_full"]), "roi_center_y_full": np.int32(g["cy_full"]),
                                "roi_radius_full": np.int32(g["r_full"]), "roi_center_x_crop": np.int32(g["cx"]),
                                "roi_center_y_crop": np.int32(g["cy"]), "roi_radius_crop": np.int32(g["r"])},
                               HEIGHTMAP_EXPORT_SAVE_CROP_CSV, HEIGHTMAP_EXPORT_SAVE_FULL_CSV)
    period = float(rec["period_px"]) if rec["period_px"] > 0 else None
    log(f"[FTP] force={rec['force_N']:.4f} N volume={rec['volume_cm3']:.6f} cm^3 period={period} px status=0x{int(rec['status']):x}")
    if return_results:
        return {"height_map_mm_crop": height.astype(np.float32), "roi_eroded_crop": roi_eroded.astype(bool),
                "output_reliable_crop": m["output_reliable"].astype(bool), "estimated_grating_period_px": period}


if __name__ == "__main__":
    main()
