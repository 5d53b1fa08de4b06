"""Pins the oracle restatement (oracle/*.py, oracle/native/unwrap_heap.c) against outputs of the
UNMODIFIED reference recorded by tools/make_golden.py (tests/golden/reference_outputs.*) and
against the reference's own stored summary JSONs.  CPU only."""
import hashlib
import json
import os

import cv2
import numpy as np
import pytest

import ftp_oracle as F
import temperature_oracle as T
import synthetic as S

G = os.path.join(os.path.dirname(__file__), "golden")


def sha(m):
    return hashlib.sha1(np.packbits(np.asarray(m, bool)).tobytes()).hexdigest()


def frames(case):
    if case.startswith("demo_"):
        return (cv2.imread(os.path.join(G, "FINAL_reference.jpg")),
                cv2.imread(os.path.join(G, "FINAL_" + case[5:] + ".jpg")))
    # the golden run stored synthetic frames as PNG (lossless) -> identical bytes
    return S.make_frame(0, reference=True), S.make_frame(int(case.split("_")[1]))


CASES = ["demo_TEMP_DEMO", "demo_ROUND_METAL", "syn_3", "syn_15"]


@pytest.mark.parametrize("case", CASES)
def test_ftp_oracle_matches_reference(case, golden):
    sc, arr = golden
    g = sc[case]["ftp"]
    ref, de = frames(case)
    P = F.FtpParams(apply_global_shift=sc[case]["aligned"], use_ecc=sc[case]["aligned"])
    r = F.shape_pair_from_frames(ref, de, P)
    fr = F.force_readings(r)
    # masks: bit-exact
    for k, k2 in [("roi_eroded", "roi"), ("reliable", "reliable"), ("output_reliable", "output_reliable"),
                  ("contact_kept_by_depth", "contact_kept"), ("hole_candidates", "hole_candidates"),
                  ("contact_dilated", "contact_dilated")]:
        assert int(r[k2].sum()) == g["mask_" + k]["count"], k
        assert sha(r[k2]) == g["mask_" + k]["sha1"], k
    # height map: the heap unwrap is restated in C (libm atan2f/cexpf instead of numpy's) -> float rounding only
    sub = arr[case + "__height_sub6"]
    np.testing.assert_array_equal(np.isnan(sub), np.isnan(r["height_mm"][::6, ::6]))
    assert np.nanmax(np.abs(sub - r["height_mm"][::6, ::6])) <= 2e-6          # mm
    assert abs(g["height_nansum"] - float(np.nansum(r["height_mm"].astype(np.float64)))) <= 1e-6 * abs(g["height_nansum"]) + 1e-3
    assert r["period_px"] == pytest.approx(g["period_px"], rel=1e-12)
    for k in ("volume_cm3", "contact_area_mm2", "max_depth_mm", "force_N"):
        assert fr[k] == pytest.approx(g[k], rel=2e-6), k


@pytest.mark.parametrize("case", CASES)
def test_temperature_oracle_matches_reference(case, golden):
    sc, arr = golden
    g = sc[case]["temperature"]
    _, de = frames(case)
    r = T.temperature_frame(de)
    y0, y1, x0, x1 = r["bbox"]
    for k in ["roi", "roi_eff", "sat", "dark", "light", "color_support", "color_ok"]:
        m = r[k][y0:y1, x0:x1]
        assert list(m.shape) == g["mask_" + k]["shape"]
        assert int(m.sum()) == g["mask_" + k]["count"], k
        assert sha(m) == g["mask_" + k]["sha1"], k
    assert hashlib.sha1(np.ascontiguousarray(r["source"][y0:y1, x0:x1]).tobytes()).hexdigest() == g["source_sha1"]
    for name in ("final", "fused"):
        np.testing.assert_array_equal(arr[f"{case}__{name}_sub8"], r[name][::8, ::8])      # bit-exact
        assert float(np.nansum(r[name].astype(np.float64))) == pytest.approx(g[name + "_nansum"], rel=1e-12)
    for k in ("mean_C", "median_C", "std_C", "min_C", "max_C", "valid_pixels"):
        assert r["stats"][k] == g[k], k


@pytest.mark.parametrize("name", ["TEMP_DEMO", "ROUND_METAL"])
def test_stored_summary_anchor(name, golden):
    """The reference's own committed multimodal_summary.json (author's machine): temperature stats
    agree to float rounding; force agrees loosely (ECC / OpenCV version drift)."""
    sc, _ = golden
    st = json.load(open(os.path.join(G, f"FINAL_{name}_multimodal_summary.json")))["sensor_readings"]
    g = sc["demo_" + name]
    for k in ("mean_C", "median_C", "std_C", "min_C", "max_C"):
        assert g["temperature"][k] == pytest.approx(st["temperature"][k], abs=2e-5)
    assert g["temperature"]["valid_pixels"] == st["temperature"]["valid_pixels"]
    assert g["ftp"]["force_N"] == pytest.approx(st["force"]["force_N"], rel=2e-3)
    assert g["ftp"]["mm_per_px"] == pytest.approx(st["force"]["scale_mm_per_px"], rel=1e-6)
