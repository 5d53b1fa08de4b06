"""GPU parity of the FTP shape + force path (vistaf_ftp_set_reference / vistaf_ftp_batch /
vistaf_depth_to_volume through the ctypes C-ABI binding) against the oracle restatement of
Code/shape_ftp.py + Code/force_sensor.py, on seeded synthetic pairs and the demo pair, and against the
digests recorded from the UNMODIFIED reference (tests/golden/reference_outputs.*).

Bars (BASELINE.json north_star): masks bit-exact, height <= 1e-3 mm, force <= 0.1 %."""
import hashlib
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

G = os.path.join(os.path.dirname(__file__), "golden")
TOL_MM = 1e-3          # stated tolerance of the height map
TOL_FORCE = 1e-3       # relative (0.1 %)
MASKS = ("reliable", "output_reliable", "contact_dilated", "contact_kept")


def sha(m):
    return hashlib.sha1(np.packbits(np.asarray(m, bool)).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def mods():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import synthetic as S
    import shape_ftp as SF
    import ftp_oracle as FO
    return S, SF, FO


@pytest.fixture(scope="module")
def syn_ref(mods):
    import cv2
    S, SF, FO = mods
    ref = S.make_frame(0, reference=True)
    P = FO.FtpParams(apply_global_shift=False, use_ecc=False)
    geom = FO.geometry(ref.shape[0], ref.shape[1], P)
    rg = cv2.cvtColor(ref[geom["y1"]:geom["y2"], geom["x1"]:geom["x2"]], cv2.COLOR_BGR2GRAY)
    return ref, P, geom, rg, FO.reference_products(rg, geom, P)


def _oracle(mods, syn_ref, frame):
    import cv2
    S, SF, FO = mods
    ref, P, geom, rg, refp = syn_ref
    dg = cv2.cvtColor(frame[geom["y1"]:geom["y2"], geom["x1"]:geom["x2"]], cv2.COLOR_BGR2GRAY)
    o = FO.ftp_pair(rg, dg, geom, P, ref=refp)
    return o, FO.force_readings(o)


def _compare(out, i, o, fr):
    rec = out["records_np"][i]
    hm = out["height_mm"][i].cpu().numpy()
    np.testing.assert_array_equal(np.isnan(hm), np.isnan(o["height_mm"]))
    assert float(np.nanmax(np.abs(hm - o["height_mm"]))) <= TOL_MM
    for k in MASKS:
        g = out[k][i].cpu().numpy().astype(bool)
        diff = g != o[k]
        if k == "contact_kept":
            # contact_kept is the SIGN of a float field (depth_mm > 0 inside kept blobs, shape_ftp.py:1244): along its
            # zero contour the two sides may differ by float rounding.  Every differing pixel must be such a pixel
            # (|depth| below the stated height tolerance on both sides), and there must be next to none of them.
            z = np.maximum(np.abs(np.nan_to_num(hm)), np.abs(np.nan_to_num(o["height_mm"])))
            assert not (diff & (z > TOL_MM)).any(), "contact_kept differs away from the zero contour"
            assert int(diff.sum()) <= 1e-4 * max(1, int(o[k].sum())), f"{int(diff.sum())} sign flips on the zero contour"
        else:
            assert int(diff.sum()) == 0, f"mask {k} differs"
    assert float(rec["force_N"]) == pytest.approx(fr["force_N"], rel=TOL_FORCE)
    assert float(rec["volume_cm3"]) == pytest.approx(fr["volume_cm3"], rel=TOL_FORCE)
    # area = (number of pixels with depth > 0.01 mm) x pixel area: pixels whose depth is within float rounding of the
    # 0.01 mm threshold may fall either side, so the count is held to 1e-4 relative (a handful of pixels), not to 0
    assert float(rec["contact_area_mm2"]) == pytest.approx(fr["contact_area_mm2"], rel=1e-4)
    assert float(rec["max_depth_mm"]) == pytest.approx(fr["max_depth_mm"], abs=TOL_MM)
    assert float(rec["mm_per_px"]) == pytest.approx(fr["scale_mm_per_px"], rel=1e-6)
    assert int(rec["reliable_px"]) == int(o["reliable"].sum())
    assert bool(int(rec["status"]) & 0x40) == bool(o["flipped"])


def test_synthetic_batch_matches_oracle(mods, syn_ref):
    """Three seeded pairs (one with glare blobs: seed 15) in one batch with chunk < batch."""
    S, SF, FO = mods
    seeds = [3, 15, 7]
    fr = np.stack([S.make_frame(s) for s in seeds])
    out = SF.run_batch(syn_ref[0], fr, chunk=2)
    for i in range(len(seeds)):
        o, f = _oracle(mods, syn_ref, fr[i])
        assert not (int(out["records_np"][i]["status"]) & 0x11), "empty ROI / residues on a smooth synthetic frame"
        _compare(out, i, o, f)


def test_synthetic_matches_reference_golden(mods, syn_ref, golden):
    """Same pair the UNMODIFIED reference was run on (tools/make_golden.py, case syn_3)."""
    S, SF, FO = mods
    sc, arr = golden
    g = sc["syn_3"]["ftp"]
    out = SF.run_batch(syn_ref[0], S.make_frame(3)[None], chunk=1)
    rec = out["records_np"][0]
    for k, k2 in (("reliable", "reliable"), ("output_reliable", "output_reliable"),
                  ("contact_kept_by_depth", "contact_kept"), ("contact_dilated", "contact_dilated")):
        m = out[k2][0].cpu().numpy().astype(bool)
        assert int(m.sum()) == g["mask_" + k]["count"] and sha(m) == g["mask_" + k]["sha1"], k
    hm = out["height_mm"][0].cpu().numpy()
    sub = arr["syn_3__height_sub6"]
    np.testing.assert_array_equal(np.isnan(sub), np.isnan(hm[::6, ::6]))
    assert float(np.nanmax(np.abs(sub - hm[::6, ::6]))) <= TOL_MM
    assert float(rec["force_N"]) == pytest.approx(g["force_N"], rel=TOL_FORCE)
    assert float(rec["period_px"]) == pytest.approx(g["period_px"], rel=1e-6)
    assert float(rec["contact_area_mm2"]) == pytest.approx(g["contact_area_mm2"], rel=1e-6)


def test_demo_pair_main_matches_reference_golden(mods, golden, tmp_path):
    """configs[0]: shape_ftp.main on FINAL_reference + FINAL_ROUND_METAL with the reference defaults
    (global shift + ECC on the host as in the reference), exported files and returned dict."""
    S, SF, FO = mods
    sc, arr = golden
    g = sc["demo_ROUND_METAL"]["ftp"]
    r = SF.main(reference_path=os.path.join(G, "FINAL_reference.jpg"), deformed_path=os.path.join(G, "FINAL_ROUND_METAL.jpg"),
                output_dir=str(tmp_path), batch_mode=True, save_summary_figures=False, export_heightmaps=True,
                debug=False, return_results=True)
    assert set(r) == {"height_map_mm_crop", "roi_eroded_crop", "output_reliable_crop", "estimated_grating_period_px"}
    hm = r["height_map_mm_crop"]
    assert hm.dtype == np.float32 and hm.shape == (1182, 1182)
    sub = arr["demo_ROUND_METAL__height_sub6"]
    np.testing.assert_array_equal(np.isnan(sub), np.isnan(hm[::6, ::6]))
    assert float(np.nanmax(np.abs(sub - hm[::6, ::6]))) <= TOL_MM
    assert r["estimated_grating_period_px"] == pytest.approx(g["period_px"], rel=1e-6)
    b = np.load(os.path.join(str(tmp_path), "height_map_bundle.npz"))
    for k in ("roi_eroded", "reliable", "output_reliable", "contact_kept_by_depth", "hole_candidates", "contact_dilated"):
        m = b["crop_" + k]
        assert int(m.sum()) == g["mask_" + k]["count"] and sha(m) == g["mask_" + k]["sha1"], k
    for f in ("height_map_crop.npy", "height_map_full.npy", "height_map_crop.csv"):
        assert os.path.exists(os.path.join(str(tmp_path), f))
    # force through the drop-in force_sensor module, as multimodal_sensor.py:388-419 does
    import force_sensor as FS
    mmpp = FS.GRATING_PITCH_MM / r["estimated_grating_period_px"]
    V, A, D = FS.depth_map_to_volume_cm3(hm, np.isfinite(hm), mmpp, FS.DEPTH_EPS_MM)
    F = FS.predict_force_from_volume(FS.load_force_calibration(FS.FORCE_CALIBRATION_JSON)["best_model"], V)
    assert F == pytest.approx(g["force_N"], rel=TOL_FORCE)
    assert A == pytest.approx(g["contact_area_mm2"], rel=1e-6)
    assert D == pytest.approx(g["max_depth_mm"], abs=TOL_MM)


def test_batch_independent_and_repeatable(mods, syn_ref):
    """Property at any size: a frame's outputs do not depend on batch position or chunking."""
    import torch
    S, SF, FO = mods
    a, b = S.make_frame(5), S.make_frame(6)
    o1 = SF.run_batch(syn_ref[0], np.stack([a, b, a]), chunk=3)
    h1, r1 = o1["height_mm"].clone(), o1["reliable"].clone()
    rec1 = o1["records_np"].copy()
    o2 = SF.run_batch(syn_ref[0], np.stack([b, a]), chunk=1)
    assert torch.equal(torch.nan_to_num(h1[0]), torch.nan_to_num(h1[2]))
    assert torch.equal(torch.nan_to_num(h1[0]), torch.nan_to_num(o2["height_mm"][1]))
    assert torch.equal(r1[1], o2["reliable"][0])
    assert rec1[0]["force_N"] == rec1[2]["force_N"] == o2["records_np"][1]["force_N"]


def test_flat_frame_gives_zero_force(mods, syn_ref):
    """Deformed == reference: no contact survives the blob filter's 0.1 mm floor."""
    S, SF, FO = mods
    out = SF.run_batch(syn_ref[0], syn_ref[0][None], chunk=1)
    rec = out["records_np"][0]
    o, f = _oracle(mods, syn_ref, syn_ref[0])
    assert float(rec["force_N"]) == pytest.approx(f["force_N"], abs=1e-6)
    hm = out["height_mm"][0].cpu().numpy()
    assert float(np.nanmax(np.abs(hm - o["height_mm"]))) <= TOL_MM


def test_depth_to_volume_matches_oracle(mods):
    """force_sensor.depth_map_to_volume_cm3 (:93-123) incl. the sign choice and the ROI / NaN handling."""
    S, SF, FO = mods
    import force_sensor as FS
    rng = np.random.default_rng(0)
    for sign in (1.0, -1.0):
        z = (sign * np.abs(rng.normal(0, 0.5, (300, 257)))).astype(np.float32)
        z[rng.random(z.shape) < 0.5] *= -0.01
        z[:10] = np.nan
        roi = rng.random(z.shape) < 0.8
        got = FS.depth_map_to_volume_cm3(z, roi, 0.03, 0.01)
        want = FO.volume_area_depth(z, roi, 0.03, 0.01)
        assert got[0] == pytest.approx(want[0], rel=1e-6)
        assert got[1] == pytest.approx(want[1], rel=1e-12)
        assert got[2] == pytest.approx(want[2], rel=1e-7)
    assert FS.depth_map_to_volume_cm3(np.zeros((4, 4), np.float32), np.ones((4, 4), bool), 0.03) == (0.0, 0.0, 0.0)


def test_errors_are_loud(mods, tmp_path):
    S, SF, FO = mods
    import torch
    with pytest.raises(RuntimeError):
        SF.main(reference_path=str(tmp_path / "missing.jpg"), deformed_path=str(tmp_path / "missing2.jpg"),
                output_dir=str(tmp_path), debug=False)
    plan = SF.FtpPlan(2160, 3840, chunk=1)
    with pytest.raises(RuntimeError):
        plan.run(torch.zeros((1, plan.h, plan.w), dtype=torch.uint8, device="cuda"))      # no reference set


@pytest.mark.parametrize("seed", [15, 7])
def test_telea_inpaint_bit_exact(mods, syn_ref, seed):
    """F2/F3: the bad-pixel mask and the Telea fill (cluster march with the warp heap) equal cv2.inpaint's result bit
    for bit — OpenCV's global (T, insertion order) processing order is reproduced inside every independent cluster."""
    import torch
    S, SF, FO = mods
    import cv2
    ref, P, geom, rg, refp = syn_ref
    fr = S.make_frame(seed)
    dg = cv2.cvtColor(fr[geom["y1"]:geom["y2"], geom["x1"]:geom["x2"]], cv2.COLOR_BGR2GRAY)
    o = FO.ftp_pair(rg, dg, geom, P, ref=refp, keep=True)
    plan = SF.FtpPlan(ref.shape[0], ref.shape[1], chunk=1)
    plan.set_reference(torch.from_numpy(rg).cuda())
    plan.run(torch.from_numpy(dg)[None].cuda(), masks=False)
    torch.cuda.synchronize()
    h, w = plan.h, plan.w
    bad = plan.workspace_plane("bad", 1).cpu().numpy().reshape(h, w) > 0
    img = plan.workspace_plane("img", 1).cpu().numpy().view(np.float32).reshape(h, w)
    assert int((bad != o["bad_def"]).sum()) == 0
    assert int(bad.sum()) > 1000
    np.testing.assert_array_equal(img, o["img_inp_def"])
