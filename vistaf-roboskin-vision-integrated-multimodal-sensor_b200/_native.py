"""ctypes binding of libvistaf.so (include/vistaf.h).  PyTorch is used only for device memory and
streams.  There is NO CPU fallback: importing this module without the built library, or calling
into it without a CUDA device, raises."""
from __future__ import annotations

import ctypes as C
import json
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libvistaf.so")
CAL_DIR = os.path.join(_HERE, "calibration")

MAX_TERMS, MAX_ISO = 35, 64

ST_EMPTY_ROI, ST_NO_PEAK, ST_SATURATED, ST_ROTATED_CARRIER = 0x1, 0x2, 0x4, 0x8
ST_RESIDUES, ST_FEW_RELIABLE, ST_SIGN_FLIPPED, ST_CONTACT_P95, ST_BG_FALLBACK = 0x10, 0x20, 0x40, 0x80, 0x100
ST_SAT_UNFILLED, ST_SPLIT_RELIABLE, ST_HOLES_FILLED, ST_TELEA_OVERFLOW = 0x200, 0x400, 0x800, 0x8000
ST_RERUN_ALONE = ST_SAT_UNFILLED


class PolyModel(C.Structure):
    _fields_ = [("n_features", C.c_int32), ("n_terms", C.c_int32), ("n_iso", C.c_int32), ("pad_", C.c_int32),
                ("mean", C.c_double * 4), ("scale", C.c_double * 4),
                ("powers", (C.c_int32 * 4) * MAX_TERMS), ("coef", C.c_double * MAX_TERMS),
                ("intercept", C.c_double), ("iso_x", C.c_double * MAX_ISO), ("iso_y", C.c_double * MAX_ISO)]


class TempParams(C.Structure):
    _fields_ = [("height", C.c_int32), ("width", C.c_int32),
                ("roi_cx", C.c_double), ("roi_cy", C.c_double), ("roi_r", C.c_double),
                ("color_t_min", C.c_float), ("color_t_max", C.c_float), ("guard_band", C.c_float),
                ("switch_margin", C.c_float), ("final_t_min", C.c_float), ("final_t_max", C.c_float),
                ("band_radius", C.c_int32), ("dc_exclusion", C.c_int32),
                ("peak_max_dy_frac", C.c_float), ("illum_sigma", C.c_float),
                ("sat_thresh", C.c_int32), ("sat_dilate", C.c_int32),
                ("close_kx", C.c_int32), ("close_ky", C.c_int32), ("open_kx", C.c_int32), ("open_ky", C.c_int32),
                ("chroma_min", C.c_float), ("support_dilate", C.c_int32),
                ("sigma_across", C.c_float), ("sigma_along", C.c_float),
                ("color_model", PolyModel), ("wide_model", PolyModel),
                ("lab_gamma", C.c_int32 * 256), ("lab_cbrt", C.c_int32 * 3072)]


class TempRecord(C.Structure):
    _fields_ = [("mean_C", C.c_float), ("median_C", C.c_float), ("std_C", C.c_float), ("min_C", C.c_float),
                ("max_C", C.c_float), ("valid_pixels", C.c_int32), ("peak_x", C.c_int32), ("peak_y", C.c_int32),
                ("phi0", C.c_float), ("mean_gray_a", C.c_float), ("mean_gray_b", C.c_float),
                ("median_gray", C.c_float), ("sat_pixels", C.c_int32), ("roi_eff_pixels", C.c_int32),
                ("dark_pixels", C.c_int32), ("color_ok_pixels", C.c_int32), ("status", C.c_int32),
                ("pad_", C.c_int32 * 3)]


TEMP_RECORD_DTYPE = np.dtype([(n, np.float32 if t is C.c_float else np.int32) for n, t in TempRecord._fields_[:-1]]
                             + [("pad_", np.int32, (3,))])
assert TEMP_RECORD_DTYPE.itemsize == C.sizeof(TempRecord) == 80


class TempOutputs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("fused", "final_map", "sat", "dark", "light", "color_support",
                                          "color_ok", "source", "records")] + [("background_valid", C.c_int32), ("pad_", C.c_int32)]


# ------------------------------------------------------------------ FTP shape + force path (include/vistaf.h)
MODEL_TYPES = {"linear0": 0, "linear": 1, "poly2": 2, "sat_exp": 3, "growth": 4, "hinge_saturating": 5}
MODEL_PARAM_NAMES = {"linear0": ("a",), "linear": ("a", "b"), "poly2": ("c2", "c1", "c0"), "sat_exp": ("a", "b"),
                     "growth": ("a", "b"), "hinge_saturating": ("a", "b", "c")}


class CurveModel(C.Structure):
    _fields_ = [("type", C.c_int32), ("pad_", C.c_int32), ("p", C.c_double * 3)]


def curve_model(best_model: dict) -> CurveModel:
    """best_model = {"type": ..., "params": {...}} as stored in the calibration JSONs (shape_ftp.py:672, force_sensor.py:142)."""
    t = best_model["type"]
    if t not in MODEL_TYPES:
        raise ValueError(f"Unknown model type in calibration: {t}")
    m = CurveModel()
    m.type = MODEL_TYPES[t]
    for i, k in enumerate(MODEL_PARAM_NAMES[t]):
        m.p[i] = float(best_model["params"][k])
    return m


class FtpParams(C.Structure):
    _fields_ = [("crop_h", C.c_int32), ("crop_w", C.c_int32), ("cx", C.c_int32), ("cy", C.c_int32), ("r", C.c_int32),
                ("roi_erode_px", C.c_int32), ("apod_taper", C.c_int32),
                ("patch_half", C.c_int32), ("dc_exclusion", C.c_int32), ("n_peaks", C.c_int32), ("fft_pad", C.c_int32),
                ("peak_max_dy_frac", C.c_float),
                ("illum_sigma", C.c_float), ("pre_blur_sigma", C.c_float), ("quality_sigma", C.c_float),
                ("smooth_sigma", C.c_float), ("unrel_sigma", C.c_float), ("amp_percentile", C.c_float),
                ("close_kernel", C.c_int32), ("edge_margin", C.c_int32),
                ("bad_int_pct", C.c_float), ("bad_grad_pct", C.c_float),
                ("bad_dilate", C.c_int32), ("bad_inpaint_radius", C.c_int32),
                ("core_pct", C.c_float), ("contact_pct", C.c_float),
                ("dilate_kernel", C.c_int32), ("dilate_iters", C.c_int32),
                ("min_contact_frac", C.c_float), ("max_contact_frac", C.c_float), ("frontier_band", C.c_float),
                ("use_negated_height", C.c_int32),
                ("blob_min_peak_mm", C.c_double), ("blob_rel_frac", C.c_double),
                ("height_model", CurveModel), ("force_model", CurveModel),
                ("grating_pitch_mm", C.c_double), ("depth_eps_mm", C.c_double),
                ("fill_internal_holes", C.c_int32), ("hole_neighborhood", C.c_int32),
                ("hole_known_frac", C.c_float), ("hole_min_dist", C.c_float),
                ("hole_inpaint_radius", C.c_int32), ("pad2_", C.c_int32)]


class FtpRecord(C.Structure):
    _fields_ = [("force_N", C.c_double), ("volume_cm3", C.c_double), ("contact_area_mm2", C.c_double),
                ("max_depth_mm", C.c_double), ("mm_per_px", C.c_double), ("period_px", C.c_double),
                ("amp_thr", C.c_float), ("contact_thr", C.c_float), ("bg_median", C.c_float), ("core_median", C.c_float),
                ("reliable_px", C.c_int32), ("output_reliable_px", C.c_int32), ("contact_dilated_px", C.c_int32),
                ("contact_kept_px", C.c_int32), ("bad_px", C.c_int32), ("residues", C.c_int32),
                ("seed_x", C.c_int32), ("seed_y", C.c_int32), ("status", C.c_int32), ("pad_", C.c_int32)]


_T = {C.c_double: np.float64, C.c_float: np.float32, C.c_int32: np.int32}
FTP_RECORD_DTYPE = np.dtype([(n, _T[t]) for n, t in FtpRecord._fields_])
assert FTP_RECORD_DTYPE.itemsize == C.sizeof(FtpRecord) == 104


class FtpOutputs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("height_mm", "reliable", "output_reliable", "contact_dilated",
                                          "contact_kept", "records", "hole_candidates")]


class FtpReferenceInfo(C.Structure):
    _fields_ = [("peak_x", C.c_float), ("peak_y", C.c_float), ("fft_h", C.c_int32), ("fft_w", C.c_int32),
                ("kx", C.c_float), ("ky", C.c_float), ("period_px", C.c_float), ("bad_px", C.c_int32)]


class StageTime(C.Structure):
    _fields_ = [("name", C.c_char * 48), ("ms", C.c_double), ("calls", C.c_int32), ("launches", C.c_int32)]


def profile_enable(on: bool) -> None:
    check(lib().vistaf_profile_enable(int(bool(on))), "profile_enable")


def profile_read() -> list:
    """[(stage, ms, calls, launches)] accumulated since the last read (synchronises the last recorded event)."""
    buf = (StageTime * 64)()
    n = C.c_int()
    check(lib().vistaf_profile_read(buf, 64, C.byref(n)), "profile_read")
    return [(buf[i].name.decode(), float(buf[i].ms), int(buf[i].calls), int(buf[i].launches)) for i in range(n.value)]


def launch_count() -> int:
    return int(lib().vistaf_launch_count())


_lib = None


def lib():
    """Load libvistaf.so; fail loudly if it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"libvistaf.so not built: {LIB_PATH} (run `python __graft_entry__.py`); "
                               "there is no CPU fallback for the sensing path")
        L = C.CDLL(LIB_PATH)
        L.vistaf_last_error.restype = C.c_char_p
        L.vistaf_temp_workspace_bytes.restype = C.c_size_t
        L.vistaf_temp_workspace_bytes.argtypes = [C.c_void_p, C.c_int]
        L.vistaf_temp_plan_create.argtypes = [C.POINTER(TempParams), C.POINTER(C.c_void_p)]
        L.vistaf_temp_plan_destroy.argtypes = [C.c_void_p]
        L.vistaf_temp_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(TempOutputs), C.c_void_p,
                                        C.c_size_t, C.c_void_p]
        L.vistaf_temp_roi_mask.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.vistaf_temp_workspace_plane.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.POINTER(C.c_size_t),
                                                  C.POINTER(C.c_size_t)]
        L.vistaf_ftp_plan_create.argtypes = [C.POINTER(FtpParams), C.POINTER(C.c_void_p)]
        L.vistaf_ftp_plan_destroy.argtypes = [C.c_void_p]
        L.vistaf_ftp_workspace_bytes.restype = C.c_size_t
        L.vistaf_ftp_workspace_bytes.argtypes = [C.c_void_p, C.c_int]
        L.vistaf_crop_gray.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_void_p, C.c_void_p]
        L.vistaf_ftp_set_reference.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                               C.POINTER(FtpReferenceInfo), C.c_void_p]
        L.vistaf_ftp_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(FtpOutputs), C.c_void_p, C.c_size_t,
                                       C.c_void_p]
        L.vistaf_ftp_batch_begin.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
        L.vistaf_ftp_batch_finish.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(FtpOutputs), C.c_void_p, C.c_size_t,
                                              C.c_void_p]
        L.vistaf_ftp_fill_holes.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_size_t, C.c_void_p]
        L.vistaf_ftp_unwrap.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                        C.c_void_p, C.c_size_t, C.c_void_p]
        L.vistaf_ftp_set_debug.argtypes = [C.c_void_p, C.c_int]
        L.vistaf_graph_replays.restype = C.c_longlong
        L.vistaf_ftp_masks.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.vistaf_ftp_workspace_plane.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.POINTER(C.c_size_t),
                                                 C.POINTER(C.c_size_t)]
        L.vistaf_depth_to_volume.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_longlong, C.c_double, C.c_double,
                                             C.c_void_p, C.c_void_p]
        L.vistaf_depth_to_volume_static.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_longlong, C.c_double, C.c_double,
                                                    C.c_void_p, C.c_void_p]
        L.vistaf_profile_read.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.vistaf_launch_count.restype = C.c_longlong
        L.vistaf_curve_predict.restype = C.c_double
        L.vistaf_curve_predict.argtypes = [C.POINTER(CurveModel), C.c_double]
        _lib = L
    return _lib


def check(rc, what=""):
    if rc != 0:
        raise RuntimeError(f"libvistaf {what} failed: {lib().vistaf_last_error().decode()}")


def load_poly_model(json_name: str) -> PolyModel:
    d = json.load(open(os.path.join(CAL_DIR, json_name)))
    m = PolyModel()
    m.n_features, m.n_terms = len(d["use_features"]), len(d["coef"])
    for i, (a, b) in enumerate(zip(d["scaler_mean"], d["scaler_scale"])):
        m.mean[i], m.scale[i] = a, b
    for t, (pw, c) in enumerate(zip(d["powers"], d["coef"])):
        m.coef[t] = c
        for f, p in enumerate(pw):
            m.powers[t][f] = p
    m.intercept = d["intercept"]
    iso = d.get("isotonic")
    m.n_iso = len(iso["x"]) if iso else 0
    if iso:
        for i, (x, y) in enumerate(zip(iso["x"], iso["y"])):
            m.iso_x[i], m.iso_y[i] = x, y
    return m


def lab_tables():
    d = json.load(open(os.path.join(CAL_DIR, "lab_lut.json")))
    return d["gamma"], d["cbrt"]


def require_cuda():
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("VISTAF B200 path needs a CUDA device; there is no CPU fallback")
    return torch


def wait_ready(stream, ready, lo, hi):
    """`ready`: optional list of (lo, hi, torch.cuda.Event) -- frames [lo, hi) of the batch are valid once the event has
    completed (sliced host->device staging).  Makes `stream` wait for every slice overlapping frames [lo, hi)."""
    if ready:
        for a, b, ev in ready:
            if a < hi and b > lo:
                stream.wait_event(ev)
