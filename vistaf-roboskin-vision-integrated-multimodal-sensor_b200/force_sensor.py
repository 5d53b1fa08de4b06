"""force_sensor — drop-in for the reference's Code/force_sensor.py, computed on B200.

Same constants (force_sensor.py:22-44), same functions and error behaviour:
  depth_map_to_volume_cm3 (:93-123)   -> vistaf_depth_to_volume (sign choice + masked sum / count / max as
                                         one warp-shuffle block reduction on the device, float64 accumulators)
  predict_force_from_volume (:149-167) -> vistaf_curve_predict (the six calibrated curve families)
  load_force_calibration (:142), estimate_mm_per_px (:173), main() (:193; writes result.json / result.csv
  with the same keys / columns).
`volumes_batch` is the batched form used by multimodal_sensor.run_batch: it reduces a [B,h,w] device
tensor of height maps without leaving the GPU.  The interactive matplotlib 3-D window is not produced
(headless).  There is no CPU fallback.
"""
from __future__ import annotations

import csv
import ctypes as C
import json
import os
from typing import Any, Dict, Optional, Tuple

import numpy as np

import _native as N
import shape_ftp

# ---------------------------------------------------------------- PATHS / CONFIG (force_sensor.py:22-44)
REFERENCE_PATH = "./Final_demos_images/FINAL_reference.jpg"
DEFORMED_PATH = "./Final_demos_images/FINAL_E_deformed.jpg"
OUTPUT_DIR = "./Force/force_sensor_out"
FORCE_CALIBRATION_JSON = os.path.join(N.CAL_DIR, "height_to_force_calibration_model.json")
GRATING_PITCH_MM = 2.0
DEPTH_EPS_MM = 0.01
OVERRIDE_MM_PER_PX = None
SHAPE_CALIBRATION_JSON = getattr(shape_ftp, "CALIBRATION_JSON", None)
SAVE_SUMMARY_FIGURES = True
EXPORT_HEIGHTMAPS = False
DEBUG = False
SHOW_3D_HEIGHTMAP_POPUP = False      # no window on a headless box


def ensure_dir(p: str):
    os.makedirs(p, exist_ok=True)


def safe_float(x, fallback=np.nan) -> float:
    """float(x) when that is a finite number, else the fallback."""
    try:
        v = float(x)
    except (TypeError, ValueError):
        return float(fallback)
    return v if np.isfinite(v) else float(fallback)


# ---------------------------------------------------------------- volume integration on the device
def volumes_batch(height_mm_dev, roi_dev, mm_per_px: float, depth_eps_mm: float = 0.01):
    """height_mm_dev: torch CUDA float32 [B,h,w]; roi_dev: CUDA uint8 [h,w] or None (ROI = finite pixels).
    Returns a CUDA float64 tensor [B,3] = (volume_cm3, contact_area_mm2, max_depth_mm) per frame."""
    t = N.require_cuda()
    assert height_mm_dev.is_cuda and height_mm_dev.dtype == t.float32 and height_mm_dev.is_contiguous()
    B = int(height_mm_dev.shape[0])
    npix = int(height_mm_dev[0].numel())
    out3 = t.empty((B, 3), dtype=t.float64, device="cuda")
    if roi_dev is not None:
        assert roi_dev.is_cuda and roi_dev.dtype == t.uint8 and roi_dev.numel() == npix and roi_dev.is_contiguous()
    N.check(N.lib().vistaf_depth_to_volume(height_mm_dev.data_ptr(), roi_dev.data_ptr() if roi_dev is not None else None,
                                           B, npix, float(mm_per_px), float(depth_eps_mm), out3.data_ptr(),
                                           C.c_void_p(t.cuda.current_stream().cuda_stream)), "depth_to_volume")
    return out3


def depth_map_to_volume_cm3(height_map_mm: np.ndarray, roi_mask: np.ndarray, mm_per_px: float,
                            depth_eps_mm: float = 0.01) -> Tuple[float, float, float]:
    """(volume_cm3, contact_area_mm2, max_depth_mm) of depth > eps inside the ROI (force_sensor.py:93-123)."""
    t = N.require_cuda()
    Z = np.ascontiguousarray(np.asarray(height_map_mm, dtype=np.float32))
    roi = np.ascontiguousarray(np.asarray(roi_mask, dtype=bool).astype(np.uint8))
    if roi.shape != Z.shape:
        raise IndexError(f"boolean index did not match indexed array: {roi.shape} vs {Z.shape}")
    out3 = volumes_batch(t.from_numpy(Z)[None].cuda(), t.from_numpy(roi).cuda(), mm_per_px, depth_eps_mm)
    v, a, d = out3[0].cpu().tolist()
    return float(v), float(a), float(d)


# ---------------------------------------------------------------- force model
def load_force_calibration(path: str) -> Dict[str, Any]:
    with open(path, "r", encoding="utf-8") as f:
        data = json.load(f)
    if "best_model" not in data:
        raise ValueError("Invalid force calibration JSON: missing 'best_model'")
    return data


def predict_force_from_volume(best_model: Dict[str, Any], volume_cm3: float) -> float:
    t = best_model["type"]
    if t not in N.MODEL_TYPES:
        raise ValueError(f"Unknown model type in force calibration JSON: {t}")
    m = N.curve_model(best_model)
    return float(N.lib().vistaf_curve_predict(C.byref(m), float(volume_cm3)))


def estimate_mm_per_px(estimated_grating_period_px: Optional[float]) -> float:
    if OVERRIDE_MM_PER_PX is not None:
        return float(OVERRIDE_MM_PER_PX)
    if estimated_grating_period_px is None:
        raise RuntimeError("shape_ftp did not return estimated_grating_period_px and OVERRIDE_MM_PER_PX is not set.")
    est = float(estimated_grating_period_px)
    if (not np.isfinite(est)) or est <= 1e-12:
        raise RuntimeError(f"Invalid estimated_grating_period_px={estimated_grating_period_px}. "
                           "Fix shape_ftp return or set OVERRIDE_MM_PER_PX.")
    return float(GRATING_PITCH_MM) / est


RESULT_CSV_FIELDS = ["reference_path", "deformed_path", "volume_cm3", "force_N", "contact_area_mm2", "max_depth_mm",
                     "mm_per_px", "estimated_grating_period_px", "ftp_output_dir", "force_model_type"]


def write_result_files(out_dir: str, out: Dict[str, Any]) -> Tuple[str, str]:
    """result.json + result.csv with the reference's keys / columns (force_sensor.py:240-291)."""
    json_path = os.path.join(out_dir, "result.json")
    with open(json_path, "w", encoding="utf-8") as f:
        json.dump(out, f, indent=2)
    csv_path = os.path.join(out_dir, "result.csv")
    with open(csv_path, "w", newline="", encoding="utf-8") as f:
        w = csv.DictWriter(f, fieldnames=RESULT_CSV_FIELDS)
        w.writeheader()
        row = {k: out.get(k) for k in RESULT_CSV_FIELDS}
        row["force_model_type"] = out["force_model"].get("type", "")
        w.writerow(row)
    return json_path, csv_path


def main():
    ensure_dir(OUTPUT_DIR)
    for name, p in (("REFERENCE_PATH", REFERENCE_PATH), ("DEFORMED_PATH", DEFORMED_PATH),
                    ("FORCE_CALIBRATION_JSON", FORCE_CALIBRATION_JSON)):
        if not os.path.isfile(p):
            raise FileNotFoundError(f"{name} not found: {p}")
    best_model = load_force_calibration(FORCE_CALIBRATION_JSON)["best_model"]
    ftp_out_dir = os.path.join(OUTPUT_DIR, "ftp_run")
    ensure_dir(ftp_out_dir)
    res = shape_ftp.main(reference_path=REFERENCE_PATH, deformed_path=DEFORMED_PATH, output_dir=ftp_out_dir,
                         calibration_json=SHAPE_CALIBRATION_JSON, batch_mode=True, save_summary_figures=SAVE_SUMMARY_FIGURES,
                         export_heightmaps=EXPORT_HEIGHTMAPS, debug=DEBUG, return_results=True)
    height_mm, roi = res["height_map_mm_crop"], res["roi_eroded_crop"]
    est_period_px = res.get("estimated_grating_period_px", None)
    mm_per_px = estimate_mm_per_px(est_period_px)
    volume_cm3, area_mm2, max_depth_mm = depth_map_to_volume_cm3(height_mm, roi, mm_per_px, DEPTH_EPS_MM)
    force_n = predict_force_from_volume(best_model, volume_cm3)
    out = {"reference_path": REFERENCE_PATH, "deformed_path": DEFORMED_PATH, "output_dir": OUTPUT_DIR,
           "ftp_output_dir": ftp_out_dir, "grating_pitch_mm": float(GRATING_PITCH_MM), "depth_eps_mm": float(DEPTH_EPS_MM),
           "estimated_grating_period_px": None if est_period_px is None else safe_float(est_period_px, np.nan),
           "mm_per_px": float(mm_per_px), "volume_cm3": float(volume_cm3), "contact_area_mm2": float(area_mm2),
           "max_depth_mm": float(max_depth_mm), "force_N": float(force_n),
           "force_model": {"type": best_model.get("type", ""), "params": best_model.get("params", {}),
                           "equation": best_model.get("equation", ""), "rmse": best_model.get("rmse", None),
                           "r2": best_model.get("r2", None)}}
    json_path, csv_path = write_result_files(OUTPUT_DIR, out)
    print(f"volume_cm3    = {volume_cm3:.6g}")
    print(f"force_N       = {force_n:.6g}")
    print(f"max_depth_mm  = {max_depth_mm:.6g}")
    print(f"saved         = {json_path}")
    print(f"saved         = {csv_path}")


if __name__ == "__main__":
    main()
