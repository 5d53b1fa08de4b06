"""Build-container-only: run the UNMODIFIED reference (imported from /root/reference via
tools/ref_harness.py) on the demo pairs and on synthetic pairs, and commit compact golden vectors
to tests/golden/reference_outputs.npz (+ .json with scalars).  tests/test_oracle_golden.py pins
the oracle restatement against these; they are also the end-to-end anchors for the CUDA path.

Stored per case: force/temperature scalars, mask pixel counts + SHA1 of np.packbits(mask),
every-6th-pixel subsample of the height crop, every-8th-pixel subsample of the temperature maps.
"""
import hashlib, json, os, sys, tempfile, time
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.abspath(os.path.join(HERE, ".."))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(REPO, "vistaf-roboskin-vision-integrated-multimodal-sensor_b200"))
import cv2
import synthetic as S
from ref_harness import import_reference

G = os.path.join(REPO, "tests", "golden")
sf, fs, ts = import_reference()


def sha(m):
    return hashlib.sha1(np.packbits(np.asarray(m, bool)).tobytes()).hexdigest()


def run_ftp(ref_path, def_path, aligned):
    sf.APPLY_GLOBAL_SHIFT = bool(aligned)
    sf.USE_ECC_CROP_ALIGNMENT = bool(aligned)
    out = tempfile.mkdtemp()
    t = time.time()
    r = sf.main(reference_path=ref_path, deformed_path=def_path, output_dir=out, batch_mode=True,
                save_summary_figures=False, export_heightmaps=True, debug=False, return_results=True)
    dt = time.time() - t
    b = np.load(os.path.join(out, "height_map_bundle.npz"))
    hm = r["height_map_mm_crop"]
    mmpp = 2.0 / r["estimated_grating_period_px"]
    V, A, D = fs.depth_map_to_volume_cm3(hm, np.isfinite(hm), mmpp, 0.01)
    F = fs.predict_force_from_volume(fs.load_force_calibration(fs.FORCE_CALIBRATION_JSON)["best_model"], V)
    sc = dict(period_px=r["estimated_grating_period_px"], mm_per_px=mmpp, volume_cm3=V, contact_area_mm2=A,
              max_depth_mm=D, force_N=F, seconds=dt, height_nansum=float(np.nansum(hm.astype(np.float64))))
    arrs = {"height_sub6": hm[::6, ::6].copy()}
    for k in ["roi_eroded", "reliable", "output_reliable", "contact_kept_by_depth", "hole_candidates", "contact_dilated"]:
        m = b["crop_" + k]
        sc["mask_" + k] = dict(count=int(m.sum()), sha1=sha(m))
    return sc, arrs, hm


def run_temp(path):
    out = tempfile.mkdtemp()
    ts.INPUT_IMAGE_PATH, ts.OUTPUT_DIR = path, out
    t = time.time()
    ts.main()
    dt = time.time() - t
    final = np.load(os.path.join(out, "temperature_map_final.npy"))
    fused = np.load(os.path.join(out, "temperature_map_fused.npy"))
    h, w = final.shape
    roi = ts.roi_mask_from_circle(h, w, ts.OUTER_CIRCLE_P1, ts.OUTER_CIRCLE_P2, ts.OUTER_CIRCLE_P3)
    v = final[roi & np.isfinite(final)]
    sc = dict(mean_C=float(np.mean(v)), median_C=float(np.median(v)), std_C=float(np.std(v)), min_C=float(np.min(v)),
              max_C=float(np.max(v)), valid_pixels=int(v.size), seconds=dt,
              final_nansum=float(np.nansum(final.astype(np.float64))), fused_nansum=float(np.nansum(fused.astype(np.float64))))
    for k in ["roi", "roi_eff", "sat", "dark", "light", "color_support", "color_ok"]:
        m = cv2.imread(os.path.join(out, f"mask_{k}.png"), 0) > 127          # cropped to ROI bbox + 10
        sc["mask_" + k] = dict(count=int(m.sum()), sha1=sha(m), shape=list(m.shape))
    src = cv2.imread(os.path.join(out, "debug_fused_source.png"), 0)
    sc["source_sha1"] = hashlib.sha1(src.tobytes()).hexdigest()
    return sc, {"final_sub8": final[::8, ::8].copy(), "fused_sub8": fused[::8, ::8].copy()}


if __name__ == "__main__":
    scal, arrs = {}, {}
    tmp = tempfile.mkdtemp()
    cases = [("demo_TEMP_DEMO", os.path.join(G, "FINAL_reference.jpg"), os.path.join(G, "FINAL_TEMP_DEMO.jpg"), True),
             ("demo_ROUND_METAL", os.path.join(G, "FINAL_reference.jpg"), os.path.join(G, "FINAL_ROUND_METAL.jpg"), True)]
    refp = os.path.join(tmp, "syn_ref.png")
    cv2.imwrite(refp, S.make_frame(0, reference=True))
    for seed in (3, 15):
        p = os.path.join(tmp, f"syn_{seed}.png")
        cv2.imwrite(p, S.make_frame(seed))
        cases.append((f"syn_{seed}", refp, p, False))
    for name, rp, dp, aligned in cases:
        s1, a1, _ = run_ftp(rp, dp, aligned)
        s2, a2 = run_temp(dp)
        scal[name] = dict(ftp=s1, temperature=s2, aligned=aligned)
        for k, v in {**a1, **a2}.items():
            arrs[f"{name}__{k}"] = v
        print(name, json.dumps(scal[name])[:400], flush=True)
    np.savez_compressed(os.path.join(G, "reference_outputs.npz"), **arrs)
    json.dump(scal, open(os.path.join(G, "reference_outputs.json"), "w"), indent=1)
