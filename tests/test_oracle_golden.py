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


def frames(case, sc=None):
    if case.startswith("demo_"):
        name = case[5:] + ("_deformed" if len(case) == 6 else "")
        return (cv2.imread(os.path.join(G, "FINAL_reference.jpg")), cv2.imread(os.path.join(G, "FINAL_" + name + ".jpg")))
    # the golden run stored synthetic frames as PNG (lossless) -> identical bytes
    if sc is not None and "gen" in sc[case]:
        return S.make_frame(0, reference=True), S.make_frame(**sc[case]["gen"])
    return S.make_frame(0, reference=True), S.make_frame(int(case.split("_")[1]))


CASES = ["demo_TEMP_DEMO", "demo_ROUND_METAL", "syn_3", "syn_15"]
# round 2 (tools/make_golden_r2.py): phase amplitude up to 3 pi (wrap counts != 0), the three remaining demo pairs
FTP_CASES = CASES + ["syn3pi_21", "syn3pi_56", "demo_E", "demo_F", "demo_P", "synflat_33"]      # synflat: no indentation, force 0
# ... a hot temperature field (blend band, isotonic clip), tilted stripes (warpAffine path of the oriented blur)
TEMP_CASES = CASES + ["synhot_9", "synhot_12", "syntilt_5", "syntilt_6", "demo_E", "demo_F", "demo_P", "synflat_33"]


def test_golden_environment(golden):
    """The goldens and the oracle's bit-exact claims rest on NumPy >= 2 scalar semantics (float32 percentile virtual
    index, float32 * python-float products): pin that here."""
    sc, _ = golden
    assert int(np.__version__.split(".")[0]) >= 2
    assert int(sc["_versions"]["numpy"].split(".")[0]) >= 2


@pytest.mark.parametrize("case", FTP_CASES)
def test_ftp_oracle_matches_reference(case, golden):
    sc, arr = golden
    g = sc[case]["ftp"]
    ref, de = frames(case, sc)
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


@pytest.mark.parametrize("case", TEMP_CASES)
def test_temperature_oracle_matches_reference(case, golden):
    sc, arr = golden
    g = sc[case]["temperature"]
    _, de = frames(case, sc)
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


def test_lab8_integer_recipe_matches_cv2_on_all_colours():
    """T7: the integer BGR->Lab recipe and the two tables shipped in calibration/lab_lut.json (uploaded to the GPU at plan
    creation) equal cv2.cvtColor(BGR2LAB) on all 2^24 colours (temperature_sensor.py:284)."""
    import _native as N
    gam, cb = (np.asarray(t, np.int64) for t in N.lab_tables())
    C = np.array([[1777, 1541, 778], [871, 2929, 296], [73, 448, 3575]], np.int64)
    g_, r_ = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    bad = 0
    for b0 in range(256):
        img = np.stack([np.full_like(g_, b0), g_, r_], -1)
        B, Gc, R = [gam[img[..., c]] for c in range(3)]
        f = [cb[(R * C[r, 0] + Gc * C[r, 1] + B * C[r, 2] + 2048) >> 12] for r in range(3)]
        L = (296 * f[1] - 1336934 + 16384) >> 15
        a = (500 * (f[0] - f[1]) + 128 * 32768 + 16384) >> 15
        b = (200 * (f[1] - f[2]) + 128 * 32768 + 16384) >> 15
        lab = np.stack([np.clip(v, 0, 255) for v in (L, a, b)], -1).astype(np.uint8)
        bad += int((lab != cv2.cvtColor(img, cv2.COLOR_BGR2LAB)).sum())
    assert bad == 0


def test_unwrap_oracle_wraps_and_is_path_independent_without_residues(golden):
    """F10 where it does something: on the 3 pi case the heap flood's wrap counts reach |k| >= 1, and (no residues)
    the field equals wrapped + 2 pi k for the k of ANY spanning order -- here a plain raster BFS."""
    sc, _ = golden
    ref, de = frames("syn3pi_21", sc)
    r = F.shape_pair_from_frames(ref, de, F.FtpParams(apply_global_shift=False, use_ecc=False), keep=True)
    rel = r["reliable"]
    u = F.unwrap_quality_guided(r["wrapped"], rel, r["quality"])
    k = np.rint((u - r["wrapped"]) / (2 * np.pi))
    assert np.nanmax(np.abs(k[rel])) >= 1
    assert np.nanmax(np.abs(u[rel] - (r["wrapped"][rel] + 2 * np.pi * k[rel]))) < 1e-5


def test_internal_hole_fill_equals_the_reference_functions():
    """F14 (shape_ftp.py:1153-1203 composed as :1770-1801): the oracle's fill_internal_holes against a vector recorded
    from the UNMODIFIED reference's own compute_internal_holes_within_mask + inpaint_only_mask on injected NaN islands
    (tools/make_golden_f14.py).  Candidates and output_reliable bit-exact, filled heights bit-exact (same cv2.inpaint
    input).  The CUDA path is compared with this oracle on the same input in tests/test_round2_gpu.py."""
    import ftp_oracle as FO
    from parity_util import f14_case
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "f14_reference.npz"))
    hm, rel, islands = f14_case(int(g["shape"][0]), int(g["shape"][1]))
    P = FO.FtpParams()
    assert [P.hole_nb, P.hole_frac, P.hole_min_dist, P.inpaint_radius] == [float(x) if i in (1, 2) else int(x) for i, x in enumerate(g["constants"])]
    filled, cand, out_rel = FO.fill_internal_holes(hm, rel, P)
    n = hm.size
    want_c = np.unpackbits(g["candidates_packed"])[:n].reshape(hm.shape).astype(bool)
    want_or = np.unpackbits(g["output_reliable_packed"])[:n].reshape(hm.shape).astype(bool)
    assert 20 < int(want_c.sum()) < int((islands & rel).sum())                 # some islands are candidates, some are not
    np.testing.assert_array_equal(cand, want_c)
    np.testing.assert_array_equal(out_rel, want_or)
    np.testing.assert_array_equal(filled[g["cand_y"], g["cand_x"]].view(np.uint32), g["cand_value"].view(np.uint32))
    assert np.isfinite(g["cand_value"]).all()


@pytest.mark.parametrize("name,kw", [("rugged", dict(n=320, levels=0, seed=3)), ("ties64", dict(n=320, levels=64, seed=4)),
                                      ("ties8", dict(n=256, levels=8, seed=5))])
def test_unwrap_oracle_equals_the_reference_flood_where_the_path_matters(name, kw):
    """F10 with phase residues, a hole in the mask and quality ties (13 ... 70 030 distinct quality values over
    45 - 70 k pixels): the result depends on the exact pop order of the reference's heapq flood, i.e. on the tuple ordering
    (-q, y, x, py, px) (shape_ftp.py:1043-1080).  The oracle's C heap must give the same wrap count on every pixel as
    the field recorded from the UNMODIFIED reference (tools/make_golden_unwrap.py)."""
    from parity_util import residue_case
    g = np.load(os.path.join(G, "unwrap_reference.npz"))
    wrapped, mask, q = residue_case(**kw)
    u = F.unwrap_quality_guided(wrapped, mask, q)
    np.testing.assert_array_equal(np.isfinite(u), mask)
    k = np.zeros(wrapped.shape, np.int8)
    k[mask] = np.rint((u[mask].astype(np.float64) - wrapped[mask]) / (2 * np.pi)).astype(np.int8)
    want = g[name + "__k"]
    assert int(np.abs(want).max()) >= 1 and int((k != want).sum()) == 0
    sub = g[name + "__u_sub4"]
    with np.errstate(invalid="ignore"):
        assert float(np.nanmax(np.abs(u[::4, ::4] - sub))) <= 5e-5          # rad: float32 path sums vs libm, as in the chain goldens
