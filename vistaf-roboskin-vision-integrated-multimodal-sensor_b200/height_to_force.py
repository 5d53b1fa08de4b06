"""Batch caller of the FTP path: volume -> force calibration data set (Code/height_to_force.py:316-446 of the reference).

The reference walks sphere-1.jpg ... sphere-75.jpg, calls shape_ftp.main once per image (one reference demodulation
and ~40 s of CPU each), and appends a row per image to per_image_results.csv / .jsonl so that an interrupted run
resumes.  Here the same loop, file layout, columns and resume rule sit on the batched device path: every pending
image is decoded and registered on the host exactly as shape_ftp.main does (alignment front-end),
the gray crops go through ONE FtpPlan (reference demodulated once, frames in chunks), and the per-image readings come
back as records.  The model fit that follows in the reference (:448-520) is offline calibration and stays out of the
hot path; `rows` are returned so a caller can fit them.
"""
from __future__ import annotations

import csv
import json
import os
from typing import List, Optional

import numpy as np

import _native as N
import force_sensor
import shape_ftp

# ---------------------------------------------------------------- PATHS / DATASET (height_to_force.py:38-62)
REFERENCE_PATH = "./Force/FINAL_reference.jpg"
DEFORMED_DIR = "./Force/Height_to_force/Loading"
OUTPUT_DIR = "./Force/Height_to_force/calibration_out"
FORCE_LEVELS_N = [0.5, 1.0, 2.0, 3.0, 4.0, 6.0, 8.0, 10.0, 15.0, 20.0, 25.0, 30.0, 35.0, 40.0, 45.0]
IMAGES_PER_LEVEL = 5
IMAGE_PATTERN = "sphere-{}.jpg"
IMAGE_START_INDEX = 1
GRATING_PITCH_MM = 2.0
DEPTH_EPS_MM = 0.01
OVERRIDE_MM_PER_PX = None
CHUNK = 16                      # frames per device batch

FIELDNAMES = ["file", "force_N", "volume_cm3", "contact_area_mm2", "max_depth_mm", "mm_per_px",
              "estimated_grating_period_px", "ftp_output_dir"]


def _pending(csv_path: str):
    """Rows already on disk and the file names they cover (the resume rule of height_to_force.py:341-350)."""
    rows, done = [], set()
    if os.path.exists(csv_path):
        with open(csv_path, "r", newline="", encoding="utf-8") as f:
            for row in csv.DictReader(f):
                rows.append(row)
                if "file" in row:
                    done.add(row["file"])
    return rows, done


def main(chunk: Optional[int] = None) -> List[dict]:
    import cv2
    t = N.require_cuda()
    os.makedirs(OUTPUT_DIR, exist_ok=True)
    runs_dir = os.path.join(OUTPUT_DIR, "ftp_runs")
    os.makedirs(runs_dir, exist_ok=True)
    csv_path = os.path.join(OUTPUT_DIR, "per_image_results.csv")
    jsonl_path = os.path.join(OUTPUT_DIR, "per_image_results.jsonl")
    rows, done = _pending(csv_path)
    todo = []
    for i in range(len(FORCE_LEVELS_N) * IMAGES_PER_LEVEL):
        idx = IMAGE_START_INDEX + i
        fname = IMAGE_PATTERN.format(idx)
        path = os.path.join(DEFORMED_DIR, fname)
        if fname in done:
            print(f"[SKIP] Already processed: {fname}")
        elif not os.path.isfile(path):
            print(f"[SKIP] Missing: {path}")
        else:
            todo.append((idx, fname, path, float(FORCE_LEVELS_N[i // IMAGES_PER_LEVEL])))
    if todo:
        ref_bgr = cv2.imread(REFERENCE_PATH)
        if ref_bgr is None:
            raise RuntimeError("Could not read one or both images. Check file paths.")
        H, W = ref_bgr.shape[:2]
        plan = shape_ftp.get_plan(H, W, int(chunk or CHUNK))
        ref_gray = None
        mode = "a" if os.path.exists(csv_path) else "w"
        with open(csv_path, mode, newline="", encoding="utf-8") as fcsv, open(jsonl_path, "a", encoding="utf-8") as fjsonl:
            w = csv.DictWriter(fcsv, fieldnames=FIELDNAMES)
            if mode == "w":
                w.writeheader()
            step = int(chunk or CHUNK)
            for c0 in range(0, len(todo), step):
                part = todo[c0:c0 + step]
                grays = []
                for idx, fname, path, force in part:
                    img = cv2.imread(path)
                    if img is None or img.shape != ref_bgr.shape:
                        raise RuntimeError(f"{fname}: unreadable or of a different size than the reference")
                    rg, dg = shape_ftp._align_host(ref_bgr, img, plan.geom)        # registration as shape_ftp.main does
                    ref_gray = rg if ref_gray is None else ref_gray
                    grays.append(dg)
                if c0 == 0:
                    plan.set_reference(t.from_numpy(np.ascontiguousarray(ref_gray)).cuda())
                out = plan.run(t.from_numpy(np.stack(grays)).cuda(), masks=False)
                recs = shape_ftp.records_to_numpy(out["records"])
                for (idx, fname, path, force), rec in zip(part, recs):
                    if int(rec["status"]) & N.ST_EMPTY_ROI:
                        raise RuntimeError(f"{fname}: reliable mask empty")
                    period = float(rec["period_px"])
                    if OVERRIDE_MM_PER_PX is not None:
                        mm_per_px = float(OVERRIDE_MM_PER_PX)
                    elif not np.isfinite(period) or period <= 1e-9:
                        raise RuntimeError(f"{fname}: invalid estimated_grating_period_px. "
                                           "Fix shape_ftp return or set OVERRIDE_MM_PER_PX.")
                    else:
                        mm_per_px = float(GRATING_PITCH_MM) / period
                    base = os.path.splitext(fname)[0]
                    out_dir = os.path.join(runs_dir, f"{idx:03d}_{base}_F{force:g}N")
                    os.makedirs(out_dir, exist_ok=True)
                    scale = (mm_per_px / float(rec["mm_per_px"])) ** 2 if OVERRIDE_MM_PER_PX is not None else 1.0
                    row = {"file": fname, "force_N": force, "volume_cm3": float(rec["volume_cm3"]) * scale,
                           "contact_area_mm2": float(rec["contact_area_mm2"]) * scale, "max_depth_mm": float(rec["max_depth_mm"]),
                           "mm_per_px": mm_per_px, "estimated_grating_period_px": force_sensor.safe_float(period, np.nan),
                           "ftp_output_dir": out_dir}
                    w.writerow(row)
                    fjsonl.write(json.dumps(row) + "\n")
                    rows.append(row)
                    print(f"[OK] {fname} | F={force:.3g} N | V={row['volume_cm3']:.6g} cm^3 | area={row['contact_area_mm2']:.3g} mm^2 | "
                          f"maxDepth={row['max_depth_mm']:.3g} mm | mm/px={mm_per_px:.6g} | out={out_dir}")
                fcsv.flush()
                fjsonl.flush()
    return rows


if __name__ == "__main__":
    main()
