"""multimodal_sensor — drop-in for the reference's Code/multimodal_sensor.py, computed on B200.

`main()` keeps the reference's flow (multimodal_sensor.py:349-700): session directory with
force_sensing/ temperature_sensing/ combined_outputs/, shape_ftp.main -> force_sensor -> result.json /
result.csv, temperature_sensor.main (configured by mutating its module attributes), temperature
statistics over the outer ROI (:555-567; reduced on the device, M1), and the same
combined_outputs/multimodal_summary.json key for key (:592-644).  matplotlib windows / figure crops are
not produced (headless box).

`MultimodalPlan` / `run_batch` are the batched form: B full camera frames already resident on the GPU ->
FTP shape + force record and temperature record per frame, nothing leaves the device except the
per-frame scalar records.  `run_sharded` splits a batch of frame indices over the ranks of a
torch.distributed job (contiguous shards, no data-path collective) and gathers the fixed-size records on
rank 0.
"""
from __future__ import annotations

import json
import os
import shutil
from datetime import datetime
from typing import Any, Dict, Optional

import numpy as np

import _native as N
import force_sensor
import shape_ftp
import temperature_sensor

# ---------------------------------------------------------------- USER CONFIG (multimodal_sensor.py:34-39)
REFERENCE_IMAGE = "./Final_demos_images/FINAL_reference.jpg"
DEFORMED_IMAGE = "./Final_demos_images/FINAL_E_deformed.jpg"
OUTPUT_ROOT = "./Multimodal_Sensor/run_output"
SHOW_3D_HEIGHTMAP_INTERACTIVE = False        # no window on a headless box

# calibration metric JSONs (multimodal_sensor.py:60-69); the numbers travel with the package
PHASE_TO_HEIGHT_JSON = os.path.join(N.CAL_DIR, "phase_to_height_calibration_model.json")
HEIGHT_TO_FORCE_JSON = os.path.join(N.CAL_DIR, "height_to_force_calibration_model.json")
TEMP_COLOR_METRICS_JSON = os.path.join(N.CAL_DIR, "color_models_final_summary_metrics.json")
TEMP_BLACK_METRICS_JSON = os.path.join(N.CAL_DIR, "black_models_final_summary_metrics.json")


def get_timestamp() -> str:
    return datetime.now().strftime("%Y%m%d_%H%M%S")


TIMESTAMP = get_timestamp()
SESSION_DIR = os.path.join(OUTPUT_ROOT, f"session_{TIMESTAMP}")
FORCE_SUBDIR = os.path.join(SESSION_DIR, "force_sensing")
TEMP_SUBDIR = os.path.join(SESSION_DIR, "temperature_sensing")
COMBINED_SUBDIR = os.path.join(SESSION_DIR, "combined_outputs")


ensure_dir = force_sensor.ensure_dir


safe_float = force_sensor.safe_float          # finite float or the fallback (one definition, force_sensor.py)


def load_json_safe(path: str) -> Optional[Dict[str, Any]]:
    """A calibration-metrics JSON, or None (with a console note) when it is missing or unreadable."""
    try:
        with open(path, "r", encoding="utf-8") as fh:
            return json.load(fh)
    except FileNotFoundError:
        print(f"[WARN] Calibration JSON not found: {path}")
    except (OSError, ValueError) as e:
        print(f"[ERROR] Failed to load {path}: {e}")
    return None


# ---------------------------------------------------------------- metric extraction (multimodal_sensor.py:214-281)
def extract_phase_to_height_metrics(calib: Optional[Dict]) -> Dict[str, Any]:
    if calib is None:
        return {}
    best = calib.get("best_model", {})
    return {"calibration_type": "phase_to_height", "model_type": best.get("type", "unknown"),
            "equation": best.get("equation", ""), "r2": safe_float(best.get("r2", np.nan)),
            "rmse": safe_float(best.get("rmse", np.nan)), "n_samples": int(best.get("n", 0)),
            "x_definition": calib.get("x_definition", "")}


def extract_height_to_force_metrics(calib: Optional[Dict]) -> Dict[str, Any]:
    if calib is None:
        return {}
    best = calib.get("best_model", {})
    return {"calibration_type": "height_to_force", "model_type": best.get("type", "unknown"),
            "equation": best.get("equation", ""), "r2": safe_float(best.get("r2", np.nan)),
            "rmse": safe_float(best.get("rmse", np.nan)), "n_fit": int(best.get("n_fit", 0)),
            "n_samples": int(best.get("n_samples", 0)), "volume_definition": calib.get("volume_definition", "")}


_ERR_KEYS = ("rmse_C", "mae_C", "r2", "max_abs_err_C", "p95_abs_err_C")


def extract_temp_model_metrics(calib: Optional[Dict], model_name: str) -> Dict[str, Any]:
    if calib is None:
        return {}
    models = calib.get("models_final", {})
    if model_name not in models:
        return {}
    m = models[model_name]

    def grp(d):
        out = {k: safe_float(d.get(k, np.nan)) for k in _ERR_KEYS}
        out["n"] = int(d.get("n", 0))
        return out
    return {"model": model_name, "degree": int(m.get("degree", 0)), "equation": m.get("equation", ""),
            "frames": grp(m.get("metrics_frames", {})), "means": grp(m.get("metrics_means", {}))}


def print_section(title: str):
    print("\n" + "=" * 70)
    print(title)
    print("=" * 70)


def print_calibration_metrics():
    print_section("CALIBRATION MODEL PERFORMANCE METRICS (GLOBAL TLC ONLY)")
    for path, title in ((TEMP_COLOR_METRICS_JSON, "TLC COLORED (LAB) GLOBAL"), (TEMP_BLACK_METRICS_JSON, "TLC BLACK (LAB+GRAY) GLOBAL")):
        metrics = extract_temp_model_metrics(load_json_safe(path), "global")
        if not metrics:
            continue
        print(f"\n--- {title} ---")
        print(f"Degree: {metrics['degree']}")
        for name, g in (("Frame-level:", metrics["frames"]), ("Per-temp means:", metrics["means"])):
            print(name)
            print(f"  RMSE = {g['rmse_C']:.3f} °C")
            print(f"  MAE  = {g['mae_C']:.3f} °C")
            print(f"  R²   = {g['r2']:.4f}")
            print(f"  Max |err| = {g['max_abs_err_C']:.3f} °C")
            print(f"  95th pct |err| = {g['p95_abs_err_C']:.3f} °C")


def calibration_performance() -> Dict[str, Any]:
    color, black = load_json_safe(TEMP_COLOR_METRICS_JSON), load_json_safe(TEMP_BLACK_METRICS_JSON)
    three = lambda c: {k: extract_temp_model_metrics(c, k) for k in ("heating", "cooling", "global")} if c else {}
    return {"phase_to_height": extract_phase_to_height_metrics(load_json_safe(PHASE_TO_HEIGHT_JSON)),
            "height_to_force": extract_height_to_force_metrics(load_json_safe(HEIGHT_TO_FORCE_JSON)),
            "temperature_color_model": three(color), "temperature_black_model": three(black)}


def build_summary(force: Dict[str, float], temp: Dict[str, Any], reference: str, deformed: str, session_dir: str,
                  force_subdir: str, temp_subdir: str, combined_subdir: str, session_id: str) -> Dict[str, Any]:
    """The multimodal_summary.json document, key for key (multimodal_sensor.py:592-639)."""
    return {
        "session_id": session_id,
        "timestamp": datetime.now().isoformat(),
        "input_images": {"reference": reference, "deformed": deformed},
        "output_directory": session_dir,
        "sensor_readings": {
            "force": {k: force[k] for k in ("force_N", "volume_cm3", "contact_area_mm2", "max_depth_mm", "scale_mm_per_px")},
            "temperature": {k: temp[k] for k in ("mean_C", "median_C", "std_C", "min_C", "max_C", "valid_pixels")},
        },
        "calibration_performance": calibration_performance(),
        "file_paths": {"force_subdir": force_subdir, "temperature_subdir": temp_subdir, "combined_subdir": combined_subdir},
    }


# ---------------------------------------------------------------- reference-compatible entry point
def main():
    for d in (SESSION_DIR, FORCE_SUBDIR, TEMP_SUBDIR, COMBINED_SUBDIR):
        ensure_dir(d)
    print_section(f"MULTIMODAL TACTILE SENSOR - Session {TIMESTAMP}")
    print(f"Reference: {REFERENCE_IMAGE}")
    print(f"Deformed:  {DEFORMED_IMAGE}")
    print(f"Output:    {SESSION_DIR}")

    # ---- force & shape (multimodal_sensor.py:360-474)
    print_section("FORCE & SHAPE SENSING")
    ftp_out_dir = os.path.join(FORCE_SUBDIR, "ftp_run")
    ensure_dir(ftp_out_dir)
    shape_results = shape_ftp.main(reference_path=REFERENCE_IMAGE, deformed_path=DEFORMED_IMAGE, output_dir=ftp_out_dir,
                                   calibration_json=getattr(shape_ftp, "CALIBRATION_JSON", None), batch_mode=True,
                                   save_summary_figures=True, export_heightmaps=True, debug=False, return_results=True)
    height_mm = shape_results["height_map_mm_crop"]
    force_roi_mask = np.isfinite(height_mm)
    est_period_px = shape_results.get("estimated_grating_period_px", None)
    best_model = force_sensor.load_force_calibration(force_sensor.FORCE_CALIBRATION_JSON)["best_model"]
    if est_period_px is not None and np.isfinite(est_period_px) and est_period_px > 1e-12:
        mm_per_px = float(force_sensor.GRATING_PITCH_MM) / float(est_period_px)
    else:
        raise RuntimeError("Invalid estimated_grating_period_px from shape_ftp")
    volume_cm3, contact_area_mm2, max_depth_mm = force_sensor.depth_map_to_volume_cm3(
        height_map_mm=height_mm, roi_mask=force_roi_mask, mm_per_px=mm_per_px, depth_eps_mm=force_sensor.DEPTH_EPS_MM)
    force_N = force_sensor.predict_force_from_volume(best_model, volume_cm3)
    force_results = {
        "reference_path": REFERENCE_IMAGE, "deformed_path": DEFORMED_IMAGE, "output_dir": FORCE_SUBDIR,
        "ftp_output_dir": ftp_out_dir, "grating_pitch_mm": float(force_sensor.GRATING_PITCH_MM),
        "depth_eps_mm": float(force_sensor.DEPTH_EPS_MM),
        "estimated_grating_period_px": float(est_period_px) if est_period_px else None, "mm_per_px": float(mm_per_px),
        "volume_cm3": float(volume_cm3), "contact_area_mm2": float(contact_area_mm2), "max_depth_mm": float(max_depth_mm),
        "force_N": float(force_N),
        "force_model": {"type": best_model.get("type", ""), "params": best_model.get("params", {}),
                        "equation": best_model.get("equation", ""), "rmse": best_model.get("rmse", None),
                        "r2": best_model.get("r2", None)}}
    force_sensor.write_result_files(FORCE_SUBDIR, force_results)
    print(f"\n[FORCE] {force_N:.3f} N")
    print(f"[VOLUME] {volume_cm3:.6g} cm³")
    print(f"[CONTACT AREA] {contact_area_mm2:.3f} mm²")
    print(f"[MAX DEPTH] {max_depth_mm:.3f} mm")
    print(f"[SCALE] {mm_per_px:.6g} mm/px")

    # ---- temperature (multimodal_sensor.py:489-573): configured through module attributes, as the reference does
    print_section("TEMPERATURE SENSING")
    original_in, original_out = temperature_sensor.INPUT_IMAGE_PATH, temperature_sensor.OUTPUT_DIR
    temperature_sensor.INPUT_IMAGE_PATH, temperature_sensor.OUTPUT_DIR = DEFORMED_IMAGE, TEMP_SUBDIR
    try:
        temperature_sensor.main()
    finally:
        temperature_sensor.INPUT_IMAGE_PATH, temperature_sensor.OUTPUT_DIR = original_in, original_out
    rec = temperature_sensor.LAST_RECORD          # statistics over roi_outer & isfinite(final), reduced on the device
    if rec is not None and int(rec["valid_pixels"]) > 0:
        temp = {k: float(rec[k]) for k in ("mean_C", "median_C", "std_C", "min_C", "max_C")}
    else:
        temp = {k: float("nan") for k in ("mean_C", "median_C", "std_C", "min_C", "max_C")}
    temp["valid_pixels"] = int(rec["valid_pixels"]) if rec is not None else 0
    print(f"\n[TEMPERATURE] Mean: {temp['mean_C']:.2f} °C")
    print(f"[TEMPERATURE] Median: {temp['median_C']:.2f} °C")
    print(f"[TEMPERATURE] Std: {temp['std_C']:.2f} °C")
    print(f"[TEMPERATURE] Range: [{temp['min_C']:.2f}, {temp['max_C']:.2f}] °C")

    print_calibration_metrics()

    # ---- combined output (multimodal_sensor.py:586-665)
    print_section("GENERATING COMBINED OUTPUTS")
    force = {"force_N": force_N, "volume_cm3": volume_cm3, "contact_area_mm2": contact_area_mm2,
             "max_depth_mm": max_depth_mm, "scale_mm_per_px": mm_per_px}
    summary = build_summary(force, temp, REFERENCE_IMAGE, DEFORMED_IMAGE, SESSION_DIR, FORCE_SUBDIR, TEMP_SUBDIR,
                            COMBINED_SUBDIR, TIMESTAMP)
    summary_json = os.path.join(COMBINED_SUBDIR, "multimodal_summary.json")
    with open(summary_json, "w", encoding="utf-8") as f:
        json.dump(summary, f, indent=2)
    print(f"\nSaved comprehensive summary: {summary_json}")
    for fname in ("result.json", "result.csv"):
        src = os.path.join(FORCE_SUBDIR, fname)
        if os.path.exists(src):
            shutil.copy2(src, os.path.join(COMBINED_SUBDIR, f"force_{fname}"))
    print_section("MULTIMODAL SENSING COMPLETE")
    print(f"\nSession directory: {SESSION_DIR}")
    print("\n--- SENSOR READINGS ---")
    print(f"Force:       {force_N:.3f} N")
    print(f"Temperature: {temp['mean_C']:.2f} ± {temp['std_C']:.2f} °C")
    print(f"Temperature min/max range: [{temp['min_C']:.2f}, {temp['max_C']:.2f}] °C")
    print(f"Max Depth:   {max_depth_mm:.3f} mm")
    print(f"Contact Area: {contact_area_mm2:.1f} mm²")
    print("\nAll outputs saved successfully!")
    return summary


# ---------------------------------------------------------------- batched multimodal path
# One fixed-size scalar record per frame = VistafFtpRecord (104 B) followed by VistafTempRecord (80 B).
RECORD_DTYPE = np.dtype([("ftp", N.FTP_RECORD_DTYPE), ("temp", N.TEMP_RECORD_DTYPE)])
RECORD_BYTES = RECORD_DTYPE.itemsize


class MultimodalPlan:
    """FTP + force + temperature over batches of full camera frames resident on the GPU.
    mode: "multimodal" | "ftp" | "temperature" selects which halves run (BASELINE.json configs[1..3])."""

    def __init__(self, H: int, W: int, chunk_ftp: int = 64, chunk_temp: int = 32, mode: str = "multimodal",
                 lanes_ftp: int = 1, lanes_temp: int = 1):
        if mode not in ("multimodal", "ftp", "temperature"):
            raise ValueError(f"unknown mode {mode}")
        self.torch = N.require_cuda()
        self.H, self.W, self.mode = int(H), int(W), mode
        self.ftp = shape_ftp.FtpPlan(H, W, chunk=chunk_ftp, lanes=lanes_ftp) if mode != "temperature" else None
        self.temp = temperature_sensor.TempPlan(H, W, chunk=chunk_temp, lanes=lanes_temp) if mode != "ftp" else None
        self._out = {}
        self._side = self.torch.cuda.Stream() if mode == "multimodal" else None
        self._tstream = self.torch.cuda.Stream() if mode == "multimodal" else None
        self.concurrent = True                # False: temperature half, then FTP half, on the current stream (stage timing)

    def step(self, ref_frame_dev, frames_dev, masks: bool = False, reuse_outputs: bool = True, ready=None, ref_ready=None) -> dict:
        """One batch against a (new) reference frame.  The reference-frame half (demodulation of one frame: a short,
        latency-bound chain of kernels; the carrier peak stays on the device) runs on a side stream
        while the temperature half of the batch runs on the current stream; the FTP batch follows both."""
        t = self.torch
        if self._side is None:
            if ref_ready is not None:
                t.cuda.current_stream().wait_event(ref_ready)
            if self.ftp is not None:
                gray = self.ftp.begin(frames_dev=frames_dev, ready=ready)
                self.set_reference(ref_frame_dev)
                for s in self.ftp.lane_streams:       # the reference is set asynchronously on the current stream
                    s.wait_stream(t.cuda.current_stream())
                B = int(frames_dev.shape[0])
                key = (B, bool(masks))
                cache = self._out.get(key) if reuse_outputs else None
                res = {"ftp": self.ftp.run(gray, masks=masks, out=cache["ftp"] if cache else None)}
                return self._finish(res, B, cache, key, reuse_outputs)
            return self.run(frames_dev, masks, reuse_outputs, ready=ready)
        main = t.cuda.current_stream()
        B = int(frames_dev.shape[0])
        key = (B, bool(masks))
        cache = self._out.get(key) if reuse_outputs else None
        # the temperature half runs on its own stream (and lanes) beside the FTP half: neither reads the other's data
        self._side.wait_stream(main)              # inputs are ready; the FTP workspace is free of the previous batch
        tstream = self._tstream if self.concurrent else main
        tstream.wait_stream(main)
        with t.cuda.stream(tstream):
            res = {"temp": self.temp.run(frames_dev, masks=masks, out=cache["temp"] if cache else None, ready=ready)}    # async
        # the front of the FTP chain (bad pixels, Telea fill, normalisation) does not read the reference frame: it is
        # issued first, so the reference's own serial chain (one frame, one host read of the carrier) runs beside it
        gray = self.ftp.begin(frames_dev=frames_dev, ready=ready)
        with t.cuda.stream(self._side):
            if ref_ready is not None:
                self._side.wait_event(ref_ready)
            self.set_reference(ref_frame_dev)     # asynchronous: the carrier stays on the device
        main.wait_stream(self._side)
        for s in self.ftp.lane_streams:
            s.wait_stream(self._side)
        res["ftp"] = self.ftp.run(gray, masks=masks, out=cache["ftp"] if cache else None)
        main.wait_stream(tstream)
        return self._finish(res, B, cache, key, reuse_outputs)

    def set_reference(self, ref_frame_dev, sync_info: bool = False):
        """ref_frame_dev: uint8 CUDA [H,W,3] reference (undeformed) frame.  Asynchronous by default: the carrier stays on the
        device (a reference without a carrier peak shows as ST_NO_PEAK in the frames' records)."""
        if self.ftp is not None:
            if getattr(self, "_ref_gray", None) is None:       # fixed buffer: the same call is replayed from its CUDA graph
                self._ref_gray = self.torch.empty((1, self.ftp.h, self.ftp.w), dtype=self.torch.uint8, device="cuda")
            self.ftp.crop_gray(ref_frame_dev[None], out=self._ref_gray)
            return self.ftp.set_reference(self._ref_gray[0], sync_info=sync_info)

    def run(self, frames_dev, masks: bool = False, reuse_outputs: bool = True, ready=None) -> dict:
        """frames_dev: uint8 CUDA [B,H,W,3].  Returns {"ftp": {...}, "temp": {...}, "records": uint8 [B,184]}."""
        t = self.torch
        B = int(frames_dev.shape[0])
        key = (B, bool(masks))
        cache = self._out.get(key) if reuse_outputs else None
        res = {}
        cur = t.cuda.current_stream()
        forked = self.temp is not None and self.ftp is not None and self._tstream is not None and self.concurrent
        if self.temp is not None:
            if forked:                            # temperature half beside the FTP half
                self._tstream.wait_stream(cur)
                with t.cuda.stream(self._tstream):
                    res["temp"] = self.temp.run(frames_dev, masks=masks, out=cache["temp"] if cache else None, ready=ready)
            else:
                res["temp"] = self.temp.run(frames_dev, masks=masks, out=cache["temp"] if cache else None, ready=ready)
        if self.ftp is not None:
            N.wait_ready(cur, ready, 0, B)
            gray = self.ftp.crop_gray(frames_dev)
            res["ftp"] = self.ftp.run(gray, masks=masks, out=cache["ftp"] if cache else None)
        if forked:
            cur.wait_stream(self._tstream)
        return self._finish(res, B, cache, key, reuse_outputs)

    def _finish(self, res, B, cache, key, reuse_outputs):
        t = self.torch
        rec = cache["records"] if cache else t.zeros((B, RECORD_BYTES), dtype=t.uint8, device="cuda")
        if "ftp" in res:
            rec[:, :104].copy_(res["ftp"]["records"])
        if "temp" in res:
            rec[:, 104:].copy_(res["temp"]["records"])
        res["records"] = rec
        if reuse_outputs:
            self._out[key] = res
        return res


def _records_host(self, res, frames_dev) -> np.ndarray:
    """Records of a finished step on the host (synchronises).  Temperature frames that found no T10 / T12 slot in their
    chunk (ST_RERUN_ALONE: more than an eighth of a chunk saturated or tilted) are re-run alone first and patched into
    `res`, so a caller of the batched API never consumes a frame that skipped a step."""
    rec = records_to_numpy(res["records"])
    if self.temp is not None and any(int(r["temp"]["status"]) & N.ST_RERUN_ALONE for r in rec):
        self.temp.redo_flagged(frames_dev, res["temp"])
        res["records"][:, 104:].copy_(res["temp"]["records"])
        rec = records_to_numpy(res["records"])
    return rec


MultimodalPlan.records_host = _records_host


def records_to_numpy(rec) -> np.ndarray:
    a = rec.cpu().numpy() if hasattr(rec, "cpu") else np.asarray(rec)
    return np.ascontiguousarray(a).view(RECORD_DTYPE).reshape(-1)


def run_batch(reference_bgr, frames_bgr, chunk: int = 8, masks: bool = False, mode: str = "multimodal") -> dict:
    """Batched multimodal sensing of pre-registered frames against one reference frame.  Inputs: numpy uint8
    on the host (staged through pinned memory) or torch CUDA tensors.  `records_np` has one RECORD_DTYPE row per frame."""
    t = N.require_cuda()

    def dev(a):
        if isinstance(a, np.ndarray):
            return t.from_numpy(np.ascontiguousarray(a)).pin_memory().cuda(non_blocking=True)
        return a
    ref, fr = dev(reference_bgr), dev(frames_bgr)
    plan = MultimodalPlan(int(fr.shape[1]), int(fr.shape[2]), chunk, chunk, mode)
    plan.set_reference(ref)
    out = plan.run(fr, masks=masks, reuse_outputs=False)
    out["records_np"] = plan.records_host(out, fr)
    return out


def summaries_from_records(records_np, names=None) -> list:
    """One `sensor_readings` block (multimodal_summary.json schema) per gathered record."""
    out = []
    for i, r in enumerate(records_np):
        f, tr = r["ftp"], r["temp"]
        out.append({"frame": names[i] if names else i,
                    "sensor_readings": {
                        "force": {"force_N": float(f["force_N"]), "volume_cm3": float(f["volume_cm3"]),
                                  "contact_area_mm2": float(f["contact_area_mm2"]), "max_depth_mm": float(f["max_depth_mm"]),
                                  "scale_mm_per_px": float(f["mm_per_px"])},
                        "temperature": {"mean_C": float(tr["mean_C"]), "median_C": float(tr["median_C"]),
                                        "std_C": float(tr["std_C"]), "min_C": float(tr["min_C"]), "max_C": float(tr["max_C"]),
                                        "valid_pixels": int(tr["valid_pixels"])}},
                    "status": {"ftp": int(f["status"]), "temperature": int(tr["status"])}})
    return out


if __name__ == "__main__":
    main()
