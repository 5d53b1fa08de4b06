"""GPU-box debugging aid: run one synthetic pair through the CUDA FTP path (chunk 1) and through the
oracle, and print per-stage differences of the workspace planes.  Not part of the product path."""
import os, sys, time
REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
for p in (os.path.join(REPO, "vistaf-roboskin-vision-integrated-multimodal-sensor_b200"), os.path.join(REPO, "oracle")):
    sys.path.insert(0, p)
import numpy as np, cv2, torch
import synthetic as S, ftp_oracle as FO, shape_ftp as SF, _native as N

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 3
amp = float(sys.argv[2]) if len(sys.argv) > 2 else 1.3
ref, de = S.make_frame(0, reference=True), S.make_frame(seed, amp_max=amp)
P = FO.FtpParams(apply_global_shift=False, use_ecc=False)
geom = FO.geometry(ref.shape[0], ref.shape[1], P)
rg = cv2.cvtColor(ref[geom["y1"]:geom["y2"], geom["x1"]:geom["x2"]], cv2.COLOR_BGR2GRAY)
dg = cv2.cvtColor(de[geom["y1"]:geom["y2"], geom["x1"]:geom["x2"]], cv2.COLOR_BGR2GRAY)
t0 = time.time()
refp = FO.reference_products(rg, geom, P)
o = FO.ftp_pair(rg, dg, geom, P, ref=refp, keep=True)
fr = FO.force_readings(o)
print(f"oracle: {time.time() - t0:.2f} s  peak={refp['peak_f']}  force={fr}")

plan = SF.FtpPlan(ref.shape[0], ref.shape[1], chunk=1)
g_ref = plan.crop_gray(torch.from_numpy(ref)[None].cuda())
g_def = plan.crop_gray(torch.from_numpy(de)[None].cuda())
print("crop_gray equal:", np.array_equal(g_ref[0].cpu().numpy(), rg), np.array_equal(g_def[0].cpu().numpy(), dg))
info = plan.set_reference(g_ref[0].contiguous())
print(f"gpu ref: peak=({info.peak_x}, {info.peak_y}) kx={info.kx} ky={info.ky} period={info.period_px} bad={info.bad_px}")
roi, circ, apo = plan.masks()
print("roi equal:", np.array_equal(roi.cpu().numpy() > 0, geom["roi"]), "circ equal:", np.array_equal(circ.cpu().numpy() > 0, geom["circ"]),
      "apo max diff:", float(np.abs(apo.cpu().numpy() - geom["apo"]).max()))
torch.cuda.synchronize()
t0 = time.time()
out = plan.run(g_def, masks=True)
torch.cuda.synchronize()
print(f"gpu batch(1): {1e3 * (time.time() - t0):.1f} ms")
rec = SF.records_to_numpy(out["records"])[0]
print("record:", {k: rec[k].item() for k in rec.dtype.names})
h, w = plan.h, plan.w


def plane(name, dt):
    return plan.workspace_plane(name, 1).cpu().numpy().view(dt).reshape(h, w)


def cmp(name, a, b, mask=None):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    if mask is not None:
        a, b = a[mask], b[mask]
    nan_mis = int((np.isnan(a) != np.isnan(b)).sum())
    d = np.abs(a - b)
    d = d[np.isfinite(d)]
    print(f"  {name:14s} max|d|={d.max() if d.size else 0:.3e} mean|d|={d.mean() if d.size else 0:.3e} nan-mismatch={nan_mis} n={a.size}")


def cmpm(name, a, b):
    a, b = np.asarray(a) > 0, np.asarray(b) > 0
    print(f"  {name:14s} count gpu={int(a.sum())} oracle={int(b.sum())} mismatches={int((a != b).sum())}")


valid = geom["apo"] > 1e-6
cmpm("bad", plane("bad", np.uint8), o["bad_def"])
cmp("img(inpaint)", plane("img", np.float32), o["img_inp_def"])
badm = o["bad_def"]
cmp("img on bad", plane("img", np.float32), o["img_inp_def"], badm)
Iw_g = plane("Iw", np.float32)
cmp("Iw - median", Iw_g - np.float32(np.median(Iw_g[valid])), o["Iw_def"])
cmp("wrapped", plane("wrapped", np.float32), o["wrapped"], geom["roi"])
cmp("quality", plane("quality", np.float32), o["quality"])
cmpm("reliable", out["reliable"][0].cpu().numpy(), o["reliable"])
rel = o["reliable"] & (out["reliable"][0].cpu().numpy() > 0)
cmp("unw(detr1)", plane("unw", np.float32), o["unwrapped"], rel)
hm_g = plane("hm", np.float32) * (-1.0 if (rec["status"] & N.ST_SIGN_FLIPPED) else 1.0)
cmp("hm", hm_g, o["hm"], rel)
cmpm("contact_d", out["contact_dilated"][0].cpu().numpy(), o["contact_dilated"])
cmpm("contact_kept", out["contact_kept"][0].cpu().numpy(), o["contact_kept"])
cmp("height_mm", out["height_mm"][0].cpu().numpy(), o["height_mm"])
for k, ok in (("force_N", "force_N"), ("volume_cm3", "volume_cm3"), ("contact_area_mm2", "contact_area_mm2"),
              ("max_depth_mm", "max_depth_mm"), ("mm_per_px", "scale_mm_per_px")):
    print(f"  {k:18s} gpu={rec[k]:.9g} oracle={fr[ok]:.9g} rel={abs(rec[k] - fr[ok]) / max(1e-12, abs(fr[ok])):.3e}")
print("  amp_thr", rec["amp_thr"], o["amp_thr"], " flipped", bool(rec["status"] & N.ST_SIGN_FLIPPED), o["flipped"])
# distance transforms vs OpenCV
relg = (out["reliable"][0].cpu().numpy() > 0).astype(np.uint8)
cmp("dist_in", plane("dist_in", np.float32), cv2.distanceTransform(relg * 255, cv2.DIST_L2, 3), relg > 0)
cmp("dist_out", plane("dist_out", np.float32), cv2.distanceTransform((1 - relg) * 255, cv2.DIST_L2, 3), relg == 0)
