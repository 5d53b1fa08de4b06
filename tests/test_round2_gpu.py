"""GPU parity, round 2: the stages the first round never exercised where they do something.

  F10  unwrap with wrap counts k != 0 (phase amplitude up to 3 pi), through the chain and on its own; frames WITH
       phase residues through the sequential-flood kernel, against the oracle's heap flood (oracle/native/unwrap_heap.c)
  F14  internal-hole detection + Telea r = 5 on a height plane with NaN islands inside the reliable mask
  T11  hot frames (32-34 degC blend band, isotonic clip), T12 tilted stripes (warpAffine path of the oriented blur)
  T10  more glare frames in a chunk than the old slot count
  the three remaining demo pairs, and the reference's own entry points multimodal_sensor.main() /
  temperature_sensor.main() / force_sensor.main() with the files they write (configs[0]).

Every comparison is CUDA path (through the C ABI) vs oracle or vs vectors recorded from the UNMODIFIED reference."""
import csv
import hashlib
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

G = os.path.join(os.path.dirname(__file__), "golden")
TOL_MM, TOL_FORCE, TOL_C = 1e-3, 1e-3, 1e-2


def sha(m):
    return hashlib.sha1(np.packbits(np.asarray(m, bool)).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def mods():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import synthetic as S
    import shape_ftp as SF
    import temperature_sensor as TS
    import ftp_oracle as FO
    import temperature_oracle as TO
    return S, SF, TS, FO, TO


@pytest.fixture(scope="module")
def syn_ref(mods):
    import cv2
    S, SF, TS, FO, TO = mods
    ref = S.make_frame(0, reference=True)
    P = FO.FtpParams(apply_global_shift=False, use_ecc=False)
    geom = FO.geometry(ref.shape[0], ref.shape[1], P)
    rg = cv2.cvtColor(ref[geom["y1"]:geom["y2"], geom["x1"]:geom["x2"]], cv2.COLOR_BGR2GRAY)
    return ref, P, geom, rg, FO.reference_products(rg, geom, P)


def _gray(frame, geom):
    import cv2
    return cv2.cvtColor(frame[geom["y1"]:geom["y2"], geom["x1"]:geom["x2"]], cv2.COLOR_BGR2GRAY)


def _unwrap_gpu(SF, plan, wrapped, quality, rel, mode):
    import ctypes as C
    import torch
    import _native as N
    h, w = wrapped.shape
    wr, q = torch.from_numpy(np.ascontiguousarray(wrapped))[None].cuda(), torch.from_numpy(np.ascontiguousarray(quality))[None].cuda()
    m = torch.from_numpy(np.ascontiguousarray(rel.astype(np.uint8)))[None].cuda()
    out = torch.empty((1, h, w), dtype=torch.float32, device="cuda")
    res = torch.zeros(1, dtype=torch.int32, device="cuda")
    N.check(N.lib().vistaf_ftp_unwrap(plan.handle, wr.data_ptr(), q.data_ptr(), m.data_ptr(), 1, out.data_ptr(), res.data_ptr(),
                                      int(mode), plan.ws.data_ptr(), plan.ws_bytes,
                                      C.c_void_p(torch.cuda.current_stream().cuda_stream)), "ftp_unwrap")
    torch.cuda.synchronize()
    return out[0].cpu().numpy(), int(res.item())


# ------------------------------------------------------------------ F10
@pytest.mark.parametrize("case", ["syn3pi_21", "syn3pi_56"])
def test_wrapping_amplitude_chain_matches_reference_and_oracle(mods, syn_ref, golden, case):
    """Phase amplitude 6.7 / 9.3 rad: the unwrap adds +-2 pi on part of the contact.  Chain vs the UNMODIFIED reference
    (golden) and vs the oracle, masks bit-exact."""
    S, SF, TS, FO, TO = mods
    sc, arr = golden
    ref, P, geom, rg, refp = syn_ref
    fr = S.make_frame(**sc[case]["gen"])
    out = SF.run_batch(ref, fr[None], chunk=1)
    rec = out["records_np"][0]
    g = sc[case]["ftp"]
    for k, k2 in (("reliable", "reliable"), ("output_reliable", "output_reliable"), ("contact_dilated", "contact_dilated"),
                  ("hole_candidates", "hole_candidates")):
        m = out[k2][0].cpu().numpy().astype(bool)
        assert int(m.sum()) == g["mask_" + k]["count"] and sha(m) == g["mask_" + k]["sha1"], k
    hm = out["height_mm"][0].cpu().numpy()
    sub = arr[case + "__height_sub6"]
    np.testing.assert_array_equal(np.isnan(sub), np.isnan(hm[::6, ::6]))
    assert float(np.nanmax(np.abs(sub - hm[::6, ::6]))) <= TOL_MM
    assert float(rec["force_N"]) == pytest.approx(g["force_N"], rel=TOL_FORCE)
    assert float(rec["volume_cm3"]) == pytest.approx(g["volume_cm3"], rel=TOL_FORCE)
    assert float(rec["max_depth_mm"]) == pytest.approx(g["max_depth_mm"], abs=TOL_MM)
    o = FO.ftp_pair(rg, _gray(fr, geom), geom, P, ref=refp, keep=True)
    assert float(np.nanmax(np.abs(hm - o["height_mm"]))) <= TOL_MM
    for k in ("reliable", "output_reliable", "contact_dilated"):
        assert int((out[k][0].cpu().numpy().astype(bool) != o[k]).sum()) == 0, k
    # the wrap-count field itself, on the chain's own inputs: parallel union-find AND sequential flood vs the oracle's heap
    plan = SF.get_plan(ref.shape[0], ref.shape[1], 1)
    want = FO.unwrap_quality_guided(o["wrapped"], o["reliable"], o["quality"])
    k_want = np.rint((want - o["wrapped"]) / (2 * np.pi))
    assert np.nanmax(np.abs(k_want[o["reliable"]])) >= 1, "test input must wrap"
    for mode in (0, 1):
        got, nres = _unwrap_gpu(SF, plan, o["wrapped"], o["quality"], o["reliable"], mode)
        np.testing.assert_array_equal(np.isnan(got), np.isnan(want))
        k_got = np.rint((got - o["wrapped"]) / (2 * np.pi))
        assert int((k_got[o["reliable"]] != k_want[o["reliable"]]).sum()) == 0, f"wrap counts differ (mode {mode})"
        assert float(np.nanmax(np.abs(got - want))) <= 5e-5          # rad: the reference sums float32 steps along its flood path, here wrapped + 2 pi k
        assert nres == 0


def test_unwrap_with_residues_matches_the_sequential_flood(mods, syn_ref, golden):
    """Two opposite phase singularities injected inside the reliable mask make the result path dependent: the frame is
    flagged (residues > 0) and unwrapped by the device restatement of the reference's heap flood -- identical wrap
    counts to oracle/native/unwrap_heap.c."""
    S, SF, TS, FO, TO = mods
    sc, _ = golden
    ref, P, geom, rg, refp = syn_ref
    fr = S.make_frame(**sc["syn3pi_21"]["gen"])
    o = FO.ftp_pair(rg, _gray(fr, geom), geom, P, ref=refp, keep=True)
    h, w = o["wrapped"].shape
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    vort = np.arctan2(yy - 500.5, xx - 480.5) - np.arctan2(yy - 560.5, xx - 700.5)
    wrapped = np.angle(np.exp(1j * (o["wrapped"].astype(np.float64) + vort))).astype(np.float32)
    rel = o["reliable"].copy()
    assert rel[500, 480] and rel[560, 700]
    rng = np.random.default_rng(5)
    quality = (o["quality"] * (1.0 + 0.05 * rng.standard_normal(o["quality"].shape))).astype(np.float32)   # a rugged flood order
    want = FO.unwrap_quality_guided(wrapped, rel, quality)
    plan = SF.get_plan(ref.shape[0], ref.shape[1], 1)
    got, nres = _unwrap_gpu(SF, plan, wrapped, quality, rel, 0)
    assert nres > 0
    np.testing.assert_array_equal(np.isnan(got), np.isnan(want))
    k_got, k_want = np.rint((got - wrapped) / (2 * np.pi)), np.rint((want - wrapped) / (2 * np.pi))
    assert int((k_got[rel] != k_want[rel]).sum()) == 0
    assert np.abs(k_want[rel]).max() >= 1
    assert float(np.nanmax(np.abs(got - want))) <= 5e-5


# ------------------------------------------------------------------ F14
def test_internal_holes_are_found_and_filled(mods, syn_ref):
    """A smooth height plane with NaN islands inside a reliable disc: small interior islands are candidates and get the
    Telea r = 5 fill, an island cut by the mask edge and a large one are not (shape_ftp.py:1153-1203, 1776-1801)."""
    import ctypes as C
    import torch
    import _native as N
    S, SF, TS, FO, TO = mods
    ref, P, geom, rg, refp = syn_ref
    h = w = geom["h"]
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    rel = ((xx - 590) ** 2 + (yy - 600) ** 2) <= 400 ** 2
    hm = (0.3 * np.sin(xx / 37.0) * np.cos(yy / 51.0) - 0.2).astype(np.float32)
    hm[~rel] = np.nan
    holes = np.zeros_like(rel)
    for cy, cx, r in ((500, 500, 2), (620, 700, 3), (601, 595, 1), (700, 450, 3)):       # small interior islands
        holes |= ((xx - cx) ** 2 + (yy - cy) ** 2) <= r * r
    holes |= ((xx - 800) ** 2 + (yy - 600) ** 2) <= 30 ** 2                               # large: known fraction too low
    holes |= ((xx - 590) ** 2 + (yy - 203) ** 2) <= 3 ** 2                                # at the mask edge: too close
    hm[holes & rel] = np.nan
    want, want_c, want_or = FO.fill_internal_holes(hm, rel, P)
    assert 20 < int(want_c.sum()) < int((holes & rel).sum())
    plan = SF.get_plan(ref.shape[0], ref.shape[1], 1)
    z = torch.from_numpy(hm.copy())[None].cuda()
    m = torch.from_numpy(rel.astype(np.uint8))[None].cuda()
    o_rel, o_c = torch.empty_like(m), torch.empty_like(m)
    N.check(N.lib().vistaf_ftp_fill_holes(plan.handle, z.data_ptr(), m.data_ptr(), 1, o_rel.data_ptr(), o_c.data_ptr(),
                                          plan.ws.data_ptr(), plan.ws_bytes, C.c_void_p(torch.cuda.current_stream().cuda_stream)),
            "ftp_fill_holes")
    torch.cuda.synchronize()
    got = z[0].cpu().numpy()
    assert int((o_c[0].cpu().numpy().astype(bool) != want_c).sum()) == 0
    assert int((o_rel[0].cpu().numpy().astype(bool) != want_or).sum()) == 0
    np.testing.assert_array_equal(np.isnan(got[rel]), np.isnan(want[rel]))
    np.testing.assert_array_equal(got[want_c], want[want_c])          # Telea on float32: bit-exact vs cv2.inpaint


def test_chain_without_reliable_smoothing(mods, syn_ref):
    """RELIABLE_SMOOTH_SIGMA_PX = 0 (masked_gaussian_smooth returns its input, shape_ftp.py:1140): the only
    configuration in which F14 can see a NaN inside the reliable mask."""
    S, SF, TS, FO, TO = mods
    ref, P, geom, rg, refp = syn_ref
    import dataclasses
    P0 = dataclasses.replace(P, smooth_sigma=0.0)
    fr = S.make_frame(7)
    old = SF.RELIABLE_SMOOTH_SIGMA_PX
    SF.RELIABLE_SMOOTH_SIGMA_PX = 0.0
    try:
        out = SF.run_batch(ref, fr[None], chunk=1)
    finally:
        SF.RELIABLE_SMOOTH_SIGMA_PX = old
    o = FO.ftp_pair(rg, _gray(fr, geom), geom, P0, ref=refp)
    hm = out["height_mm"][0].cpu().numpy()
    np.testing.assert_array_equal(np.isnan(hm), np.isnan(o["height_mm"]))
    assert float(np.nanmax(np.abs(hm - o["height_mm"]))) <= TOL_MM
    for k in ("reliable", "output_reliable", "contact_dilated", "hole_candidates"):
        assert int((out[k][0].cpu().numpy().astype(bool) != o[k]).sum()) == 0, k


# ------------------------------------------------------------------ demo pairs E / F / P (alignment on, as shipped)
@pytest.mark.parametrize("name", ["E", "F", "P"])
def test_demo_pairs_match_reference_golden(mods, golden, tmp_path, name):
    S, SF, TS, FO, TO = mods
    sc, arr = golden
    g = sc["demo_" + name]["ftp"]
    r = SF.main(reference_path=os.path.join(G, "FINAL_reference.jpg"), deformed_path=os.path.join(G, f"FINAL_{name}_deformed.jpg"),
                output_dir=str(tmp_path), batch_mode=True, save_summary_figures=False, export_heightmaps=True, debug=False,
                return_results=True)
    hm = r["height_map_mm_crop"]
    sub = arr[f"demo_{name}__height_sub6"]
    np.testing.assert_array_equal(np.isnan(sub), np.isnan(hm[::6, ::6]))
    assert float(np.nanmax(np.abs(sub - hm[::6, ::6]))) <= TOL_MM
    b = np.load(os.path.join(str(tmp_path), "height_map_bundle.npz"))
    for k in ("roi_eroded", "reliable", "output_reliable", "hole_candidates", "contact_dilated"):
        m = b["crop_" + k]
        assert int(m.sum()) == g["mask_" + k]["count"] and sha(m) == g["mask_" + k]["sha1"], k
    import force_sensor as FS
    mmpp = FS.GRATING_PITCH_MM / r["estimated_grating_period_px"]
    V, A, D = FS.depth_map_to_volume_cm3(hm, np.isfinite(hm), mmpp, FS.DEPTH_EPS_MM)
    F = FS.predict_force_from_volume(FS.load_force_calibration(FS.FORCE_CALIBRATION_JSON)["best_model"], V)
    assert F == pytest.approx(g["force_N"], rel=TOL_FORCE)


# ------------------------------------------------------------------ temperature: hot field, tilted stripes, many glare frames
def _compare_temp(out, i, o, rec):
    for k in ("sat", "dark", "light", "color_support", "color_ok"):
        assert int((out[k][i].cpu().numpy().astype(bool) != o[k]).sum()) == 0, f"mask {k} differs"
    assert int((out["source"][i].cpu().numpy() != o["source"]).sum()) == 0
    for k in ("fused", "final"):
        g = out[k][i].cpu().numpy()
        np.testing.assert_array_equal(np.isnan(g), np.isnan(o[k]))
        assert float(np.nanmax(np.abs(g - o[k]))) <= TOL_C, k
    st = o["stats"]
    assert int(rec["valid_pixels"]) == st["valid_pixels"]
    for k in ("mean_C", "median_C", "std_C", "min_C", "max_C"):
        assert float(rec[k]) == pytest.approx(st[k], abs=TOL_C), k
    assert (int(rec["peak_x"]), int(rec["peak_y"])) == tuple(o["seg"]["peak"])


@pytest.mark.parametrize("case", ["synhot_9", "synhot_12", "syntilt_5", "syntilt_6"])
def test_hot_and_tilted_frames_match_reference_and_oracle(mods, golden, case):
    S, SF, TS, FO, TO = mods
    sc, arr = golden
    g = sc[case]["temperature"]
    fr = S.make_frame(**sc[case]["gen"])
    out = TS.run_batch(fr[None], chunk=1)
    rec = out["records_np"][0]
    o = TO.temperature_frame(fr)
    if case.startswith("syntilt"):
        assert o["seg"]["angle"] != 0.0 and int(rec["status"]) & 0x8, "test input must have a tilted carrier"
        assert g["valid_pixels"] < 1576625            # the double nearest warp of the ROI blanks a few rim pixels
    else:
        assert int((out["source"][0].cpu().numpy() == 128).sum()) > 1000, "test input must reach the blend band"
    _compare_temp(out, 0, o, rec)
    y0, y1, x0, x1 = o["bbox"]
    for k in ("sat", "dark", "light", "color_support", "color_ok"):
        m = out[k][0].cpu().numpy()[y0:y1, x0:x1] > 0
        assert int(m.sum()) == g["mask_" + k]["count"] and sha(m) == g["mask_" + k]["sha1"], k
    for name in ("final", "fused"):
        sub = arr[f"{case}__{name}_sub8"]
        got = out[name][0].cpu().numpy()[::8, ::8]
        np.testing.assert_array_equal(np.isnan(sub), np.isnan(got))
        assert float(np.nanmax(np.abs(sub - got))) <= TOL_C
    assert int(rec["valid_pixels"]) == g["valid_pixels"]
    for k in ("mean_C", "median_C", "std_C", "min_C", "max_C"):
        assert float(rec[k]) == pytest.approx(g[k], abs=TOL_C), k


def test_every_frame_of_a_chunk_saturated(mods):
    """T10 with glare in EVERY frame of a chunk: the library fills one frame in eight per chunk and flags the rest
    (ST_RERUN_ALONE); temperature_sensor.run_batch AND the multimodal plan (records_host) re-run those alone, so no
    caller of the module API consumes an unfilled frame."""
    import torch
    import multimodal_sensor as MS
    S, SF, TS, FO, TO = mods
    seeds = [15, 31, 47]
    fr = np.stack([S.make_frame(s) for s in seeds])
    out = TS.run_batch(fr, chunk=3)
    for i in range(3):
        assert int(out["records_np"][i]["status"]) & 0x4 and not int(out["records_np"][i]["status"]) & 0x200
        _compare_temp(out, i, TO.temperature_frame(fr[i]), out["records_np"][i])
    plan = MS.MultimodalPlan(fr.shape[1], fr.shape[2], 3, 3, "temperature")
    dev = torch.from_numpy(fr).cuda()
    res = plan.run(dev, masks=True)
    assert any(int(r["temp"]["status"]) & 0x200 for r in MS.records_to_numpy(res["records"])), "test must overflow the slots"
    recs = plan.records_host(res, dev)
    for i in range(3):
        assert not int(recs[i]["temp"]["status"]) & 0x200
        assert torch.equal(torch.nan_to_num(res["temp"]["fused"][i]), torch.nan_to_num(out["fused"][i]))


# ------------------------------------------------------------------ the reference's entry points (configs[0])
def test_multimodal_main_writes_the_reference_outputs(mods, golden, tmp_path):
    """multimodal_sensor.main() on FINAL_reference + FINAL_ROUND_METAL with the shipped defaults: multimodal_summary.json
    key for key against the reference's stored summary, readings against the in-container reference run, the mask PNGs
    against the reference's digests, result.json / result.csv columns (multimodal_sensor.py:445-467)."""
    import cv2
    import multimodal_sensor as MS
    sc, arr = golden
    g = sc["demo_ROUND_METAL"]
    root = str(tmp_path)
    MS.REFERENCE_IMAGE, MS.DEFORMED_IMAGE = os.path.join(G, "FINAL_reference.jpg"), os.path.join(G, "FINAL_ROUND_METAL.jpg")
    MS.SESSION_DIR = os.path.join(root, "session")
    MS.FORCE_SUBDIR, MS.TEMP_SUBDIR, MS.COMBINED_SUBDIR = (os.path.join(MS.SESSION_DIR, d) for d in
                                                           ("force_sensing", "temperature_sensing", "combined"))
    MS.main()
    got = json.load(open(os.path.join(MS.COMBINED_SUBDIR, "multimodal_summary.json")))
    stored = json.load(open(os.path.join(G, "FINAL_ROUND_METAL_multimodal_summary.json")))

    def keys(d, pre=""):
        return sorted(k2 for k, v in d.items() for k2 in ([pre + k] + (keys(v, pre + k + ".") if isinstance(v, dict) else [])))
    assert keys(got) == keys(stored)
    f, t = got["sensor_readings"]["force"], got["sensor_readings"]["temperature"]
    assert f["force_N"] == pytest.approx(g["ftp"]["force_N"], rel=TOL_FORCE)
    assert f["volume_cm3"] == pytest.approx(g["ftp"]["volume_cm3"], rel=TOL_FORCE)
    assert f["max_depth_mm"] == pytest.approx(g["ftp"]["max_depth_mm"], abs=TOL_MM)
    assert f["scale_mm_per_px"] == pytest.approx(g["ftp"]["mm_per_px"], rel=1e-6)
    for k in ("mean_C", "median_C", "std_C", "min_C", "max_C"):
        assert t[k] == pytest.approx(g["temperature"][k], abs=TOL_C)
        assert t[k] == pytest.approx(stored["sensor_readings"]["temperature"][k], abs=TOL_C)
    assert t["valid_pixels"] == g["temperature"]["valid_pixels"] == stored["sensor_readings"]["temperature"]["valid_pixels"]
    assert got["calibration_performance"] == stored["calibration_performance"]
    for k in ("roi", "roi_eff", "sat", "dark", "light", "color_support", "color_ok"):
        m = cv2.imread(os.path.join(MS.TEMP_SUBDIR, f"mask_{k}.png"), 0) > 127
        assert list(m.shape) == g["temperature"]["mask_" + k]["shape"]
        assert sha(m) == g["temperature"]["mask_" + k]["sha1"], k
    src = cv2.imread(os.path.join(MS.TEMP_SUBDIR, "debug_fused_source.png"), 0)
    assert hashlib.sha1(src.tobytes()).hexdigest() == g["temperature"]["source_sha1"]
    for name in ("final", "fused"):
        a = np.load(os.path.join(MS.TEMP_SUBDIR, f"temperature_map_{name}.npy"))
        assert a.dtype == np.float32 and a.shape == (2160, 3840)
        sub = arr[f"demo_ROUND_METAL__{name}_sub8"]
        np.testing.assert_array_equal(np.isnan(sub), np.isnan(a[::8, ::8]))
        assert float(np.nanmax(np.abs(sub - a[::8, ::8]))) <= TOL_C
    for fn in ("height_map_crop.npy", "height_map_full.npy", "height_map_crop.csv", "height_map_bundle.npz"):
        assert os.path.exists(os.path.join(MS.FORCE_SUBDIR, "ftp_run", fn)), fn
    res = json.load(open(os.path.join(MS.FORCE_SUBDIR, "result.json")))
    assert res["force_N"] == pytest.approx(g["ftp"]["force_N"], rel=TOL_FORCE) and res["force_model"]["type"] == "growth"
    rows = list(csv.reader(open(os.path.join(MS.FORCE_SUBDIR, "result.csv"))))
    assert rows[0] == ["reference_path", "deformed_path", "volume_cm3", "force_N", "contact_area_mm2", "max_depth_mm", "mm_per_px",
                       "estimated_grating_period_px", "ftp_output_dir", "force_model_type"]       # multimodal_sensor.py:450-452
    assert float(rows[1][3]) == pytest.approx(g["ftp"]["force_N"], rel=TOL_FORCE) and rows[1][9] == "growth"
    for fn in ("force_result.json", "force_result.csv"):
        assert os.path.exists(os.path.join(MS.COMBINED_SUBDIR, fn))


def test_force_sensor_main(mods, golden, tmp_path):
    """force_sensor.main() (force_sensor.py:190-300): FTP run + volume + force, result.json / result.csv."""
    import force_sensor as FS
    sc, _ = golden
    FS.REFERENCE_PATH, FS.DEFORMED_PATH = os.path.join(G, "FINAL_reference.jpg"), os.path.join(G, "FINAL_F_deformed.jpg")
    FS.OUTPUT_DIR = str(tmp_path)
    FS.main()
    res = json.load(open(os.path.join(str(tmp_path), "result.json")))
    assert res["force_N"] == pytest.approx(sc["demo_F"]["ftp"]["force_N"], rel=TOL_FORCE)
    assert res["contact_area_mm2"] == pytest.approx(sc["demo_F"]["ftp"]["contact_area_mm2"], rel=1e-4)
    assert os.path.exists(os.path.join(str(tmp_path), "result.csv"))
