"""GPU parity of the rows either side of the hot path and of the streaming geometry of configs[4].

  configs[4]  3904^2 crop / 4096^2 FFT, phase amplitude 6.7 rad (wrap counts != 0) -- a non-default geometry end to end
  8f rank 2   the batch callers: phase_to_height's variant constants (ROI_ERODE_PX = 80, frontier band 300) and the
              height_to_force loop with its CSV / JSONL resume rule
"""
import csv
import dataclasses
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL_MM, TOL_FORCE = 1e-3, 1e-3


@pytest.fixture(scope="module")
def mods():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import synthetic as S
    import shape_ftp as SF
    import ftp_oracle as FO
    return S, SF, FO


def test_streaming_geometry_4k_matches_oracle(mods):
    """configs[4]: one 3904^2 gray crop against its reference crop through an FtpPlan built on the explicit geometry,
    against the oracle on the same bytes; masks bit-exact, the unwrap really wraps."""
    import torch
    S, SF, FO = mods
    g = S.CROP4K
    ref, de = S.make_crop4k(0, reference=True), S.make_crop4k(21)
    P = FO.FtpParams(apply_global_shift=False, use_ecc=False)
    disc = FO.disc(g["h"], g["w"], g["cx"], g["cy"], g["r"])
    geom = dict(g, circ=disc, roi=disc, apo=FO.apodization(g["h"], g["w"], g["cx"], g["cy"], g["r"], P.apod_taper))
    o = FO.ftp_pair(ref, de, geom, P, keep=True)
    fr = FO.force_readings(o)
    u = FO.unwrap_quality_guided(o["wrapped"], o["reliable"], o["quality"])
    assert np.nanmax(np.abs(np.rint((u - o["wrapped"]) / (2 * np.pi))[o["reliable"]])) >= 1
    plan = SF.FtpPlan(g["h"], g["w"], chunk=1, geom=g)
    info = plan.set_reference(torch.from_numpy(ref).cuda())
    assert (info.fft_h, info.fft_w) == (4096, 4096)
    out = plan.run(torch.from_numpy(de)[None].cuda(), masks=True)
    rec = SF.records_to_numpy(out["records"])[0]
    hm = out["height_mm"][0].cpu().numpy()
    # 11.9 M ROI pixels: a few quality values sit within float32 rounding of the p25 threshold, so masks are held to
    # "identical except proven threshold ties" (tests/parity_util.py), the height to 1e-3 mm away from those pixels
    from parity_util import check_ftp_masks
    gm = {k: out[k][0].cpu().numpy().astype(bool) for k in ("reliable", "output_reliable", "contact_dilated", "contact_kept")}
    st = check_ftp_masks(gm, hm, o, P, max_ties=64)
    print("4K tie statistics:", st)
    assert st["reliable_px"] <= 64 and st["contact_dilated_px"] <= 2000
    assert float(rec["period_px"]) == pytest.approx(o["period_px"], rel=1e-6)
    assert float(rec["force_N"]) == pytest.approx(fr["force_N"], rel=TOL_FORCE)
    assert float(rec["volume_cm3"]) == pytest.approx(fr["volume_cm3"], rel=TOL_FORCE)
    assert not int(rec["status"]) & 0x11


def test_phase_to_height_variant_constants(mods):
    """phase_to_height.compute_height_map_pair runs the same chain with ROI_ERODE_PX = 80 and a 300-px frontier band
    (phase_to_height.py:774-1006): both are plan parameters here."""
    import cv2
    S, SF, FO = mods
    ref, fr = S.make_frame(0, reference=True), S.make_frame(5)
    P = FO.FtpParams(apply_global_shift=False, use_ecc=False, roi_erode_px=80, frontier_band=300.0)
    geom = FO.geometry(ref.shape[0], ref.shape[1], P)
    crop = lambda f: cv2.cvtColor(f[geom["y1"]:geom["y2"], geom["x1"]:geom["x2"]], cv2.COLOR_BGR2GRAY)
    o = FO.ftp_pair(crop(ref), crop(fr), geom, P, keep=True)
    old = SF.ROI_ERODE_PX, SF.FRONTIER_ZERO_BAND_PX
    SF.ROI_ERODE_PX, SF.FRONTIER_ZERO_BAND_PX = 80, 300
    try:
        out = SF.run_batch(ref, fr[None], chunk=1)
    finally:
        SF.ROI_ERODE_PX, SF.FRONTIER_ZERO_BAND_PX = old
    hm = out["height_mm"][0].cpu().numpy()
    np.testing.assert_array_equal(np.isnan(hm), np.isnan(o["height_mm"]))
    assert int(np.isfinite(hm).sum()) == int(geom["roi"].sum()) < int(geom["circ"].sum())
    from parity_util import check_ftp_masks
    gm = {k: out[k][0].cpu().numpy().astype(bool) for k in ("reliable", "output_reliable", "contact_dilated", "contact_kept")}
    st = check_ftp_masks(gm, hm, o, P)
    print("variant tie statistics:", st)
    if st["reliable_px"] == 0:
        assert float(out["records_np"][0]["force_N"]) == pytest.approx(FO.force_readings(o)["force_N"], rel=TOL_FORCE)


def test_height_to_force_batch_loop_and_resume(mods, tmp_path):
    """The calibration loop of height_to_force.py:316-446 on the batched path: per-image rows with the reference's columns,
    values equal to shape_ftp.main + force_sensor on the same pair, and a second run skips what is on disk."""
    import cv2
    import force_sensor as FS
    import height_to_force as HF
    S, SF, FO = mods
    d = tmp_path / "loading"
    d.mkdir()
    refp = str(tmp_path / "ref.png")
    cv2.imwrite(refp, S.make_frame(0, reference=True))
    for i, seed in enumerate((3, 5, 7)):
        cv2.imwrite(str(d / f"sphere-{i + 1}.png"), S.make_frame(seed))
    HF.REFERENCE_PATH, HF.DEFORMED_DIR, HF.OUTPUT_DIR = refp, str(d), str(tmp_path / "out")
    HF.IMAGE_PATTERN, HF.FORCE_LEVELS_N, HF.IMAGES_PER_LEVEL = "sphere-{}.png", [1.0, 2.0], 2     # 4 expected, 3 present
    old = SF.APPLY_GLOBAL_SHIFT, SF.USE_ECC_CROP_ALIGNMENT
    SF.APPLY_GLOBAL_SHIFT = SF.USE_ECC_CROP_ALIGNMENT = False
    try:
        rows = HF.main(chunk=2)
        assert [r["file"] for r in rows] == ["sphere-1.png", "sphere-2.png", "sphere-3.png"]
        assert [r["force_N"] for r in rows] == [1.0, 1.0, 2.0]
        got = list(csv.DictReader(open(os.path.join(HF.OUTPUT_DIR, "per_image_results.csv"))))
        assert list(got[0].keys()) == HF.FIELDNAMES and len(got) == 3
        assert len(open(os.path.join(HF.OUTPUT_DIR, "per_image_results.jsonl")).read().splitlines()) == 3
        r = SF.main(reference_path=refp, deformed_path=str(d / "sphere-2.png"), output_dir=str(tmp_path / "single"),
                    batch_mode=True, save_summary_figures=False, export_heightmaps=False, debug=False, return_results=True)
        mmpp = FS.GRATING_PITCH_MM / r["estimated_grating_period_px"]
        V, A, D = FS.depth_map_to_volume_cm3(r["height_map_mm_crop"], r["roi_eroded_crop"], mmpp, FS.DEPTH_EPS_MM)
        assert float(got[1]["volume_cm3"]) == pytest.approx(V, rel=1e-6) and float(got[1]["max_depth_mm"]) == pytest.approx(D, rel=1e-6)
        again = HF.main(chunk=2)                       # nothing left to do: rows come from the CSV
        assert len(again) == 3 and len(list(csv.DictReader(open(os.path.join(HF.OUTPUT_DIR, "per_image_results.csv"))))) == 3
    finally:
        SF.APPLY_GLOBAL_SHIFT, SF.USE_ECC_CROP_ALIGNMENT = old


# ------------------------------------------------------------------ parity at the size of the configs
def _oracle_frame(args):
    """Worker: oracle FTP (keep=True fields needed by the tie proof) and / or temperature readings of one seed."""
    import sys
    seed, want_ftp, want_temp = args
    import cv2
    cv2.setNumThreads(1)
    import synthetic as S
    import ftp_oracle as FO
    import temperature_oracle as TO
    fr = S.make_frame(seed)
    res = {"seed": seed}
    if want_ftp:
        ref = S.make_frame(0, reference=True)
        P = FO.FtpParams(apply_global_shift=False, use_ecc=False)
        g = FO.geometry(ref.shape[0], ref.shape[1], P)
        crop = lambda f: cv2.cvtColor(f[g["y1"]:g["y2"], g["x1"]:g["x2"]], cv2.COLOR_BGR2GRAY)
        o = FO.ftp_pair(crop(ref), crop(fr), g, P, keep=True)
        res["ftp"] = {k: o[k] for k in ("roi", "reliable", "output_reliable", "contact_dilated", "contact_kept", "quality",
                                        "amp_thr", "abs_res", "contact_thr", "height_mm")}
        res["force"] = FO.force_readings(o)
    if want_temp:
        t = TO.temperature_frame(fr)
        sig = t["seg"]["s"]
        tie = t["roi_eff"] & (np.abs(sig) <= 2e-6 * float(np.abs(sig[t["roi_eff"]]).max()))      # sign of s undecided in float32
        res["temp"] = {"stats": t["stats"], "dark": np.packbits(t["dark"]), "color_ok": np.packbits(t["color_ok"]),
                       "final": t["final"], "tie": np.packbits(tie), "shape": t["dark"].shape}
    return res


def test_seed_sweep_of_the_configs(mods):
    """configs[1] / configs[3] seeds 0..31 (FTP + temperature) and configs[2] seeds 1000..1031 (temperature): every frame
    against the oracle (computed on a process pool), with the per-mask tie statistics of tests/parity_util.py."""
    import multiprocessing as mp
    import cv2
    import torch
    import multimodal_sensor as MS
    from parity_util import check_ftp_masks
    S, SF, FO = mods
    jobs = [(s, True, True) for s in range(32)] + [(s, False, True) for s in range(1000, 1032)]
    avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else 8
    with mp.get_context("spawn").Pool(max(1, min(16, avail))) as pool:
        orc = pool.map(_oracle_frame, jobs, chunksize=1)
    ref = S.make_frame(0, reference=True)
    P = FO.FtpParams(apply_global_shift=False, use_ecc=False)
    summary = {"frames_ftp": 0, "frames_temp": 0, "reliable_px": 0, "contact_dilated_px": 0, "contact_kept_px": 0,
               "frames_with_any_mask_diff": 0, "max_rel_force": 0.0, "max_height_mm_far": 0.0, "max_abs_T": 0.0,
               "temp_mask_px": 0, "temp_sign_ties": 0}
    plan = MS.MultimodalPlan(ref.shape[0], ref.shape[1], 8, 8, "multimodal")
    ref_dev = torch.from_numpy(ref).cuda()
    for c0 in range(0, len(jobs), 8):
        part = orc[c0:c0 + 8]
        frames = torch.from_numpy(np.stack([S.make_frame(o["seed"]) for o in part])).cuda()
        res = plan.step(ref_dev, frames, masks=True, reuse_outputs=False)
        recs = plan.records_host(res, frames)
        for i, o in enumerate(part):
            if "ftp" in o:
                gm = {k: res["ftp"][k][i].cpu().numpy().astype(bool) for k in ("reliable", "output_reliable", "contact_dilated", "contact_kept")}
                st = check_ftp_masks(gm, res["ftp"]["height_mm"][i].cpu().numpy(), o["ftp"], P)
                summary["frames_ftp"] += 1
                for k in ("reliable_px", "contact_dilated_px", "contact_kept_px"):
                    summary[k] += st[k]
                summary["frames_with_any_mask_diff"] += int(st["reliable_px"] + st["contact_dilated_px"] + st["contact_kept_px"] > 0)
                summary["max_height_mm_far"] = max(summary["max_height_mm_far"], st["height_max_far"])
                if st["reliable_px"] == 0:
                    rf = abs(float(recs[i]["ftp"]["force_N"]) - o["force"]["force_N"]) / max(1e-9, abs(o["force"]["force_N"]))
                    summary["max_rel_force"] = max(summary["max_rel_force"], rf)
                    assert rf <= TOL_FORCE, (o["seed"], rf)
            t = o["temp"]
            summary["frames_temp"] += 1
            H, W = t["shape"]
            unp = lambda a: np.unpackbits(a)[:H * W].reshape(H, W).astype(bool)
            tie = unp(t["tie"])
            # dark / light = sign of the band-passed stripe signal s, then close 3x31 / open 3x7: a pixel whose |s| is below
            # float32 rounding of the FFT sums may fall either way and moves the mask within the reach of that morphology;
            # the fused / final maps then differ inside the final blur's radius around it (the colour and the wide model
            # disagree by degrees).  Everything else must be identical / within 1e-2 degC.
            near_tie = cv2.dilate(tie.astype(np.uint8), np.ones((41, 9), np.uint8)) > 0 if tie.any() else tie
            bad = np.zeros((H, W), bool)
            for k in ("dark", "color_ok"):
                d = res["temp"][k][i].cpu().numpy().astype(bool) != unp(t[k])
                summary["temp_mask_px"] += int(d.sum())
                assert not (d & ~near_tie).any(), f"seed {o['seed']}: {k} differs away from any sign tie of the stripe signal"
                bad |= d
            summary["temp_sign_ties"] += int(tie.sum())
            far = ~(cv2.dilate(bad.astype(np.uint8), np.ones((61, 61), np.uint8)) > 0) if bad.any() else np.ones((H, W), bool)
            fin = res["temp"]["final"][i].cpu().numpy()
            assert np.array_equal(np.isnan(fin), np.isnan(t["final"]))
            summary["max_abs_T"] = max(summary["max_abs_T"], float(np.nanmax(np.abs(np.where(far, fin - t["final"], 0.0)))))
            if not bad.any():
                for k in ("mean_C", "median_C", "max_C"):
                    assert float(recs[i]["temp"][k]) == pytest.approx(t["stats"][k], abs=1e-2)
            assert int(recs[i]["temp"]["valid_pixels"]) == t["stats"]["valid_pixels"]
    print("seed sweep:", json.dumps(summary))
    assert summary["max_abs_T"] <= 1e-2 and summary["temp_mask_px"] <= 2 * summary["frames_temp"]
    if os.path.isdir(os.path.join(os.path.dirname(__file__), "..", "gpurun_out")):
        json.dump(summary, open(os.path.join(os.path.dirname(__file__), "..", "gpurun_out", "r02_parity_sweep.json"), "w"))
