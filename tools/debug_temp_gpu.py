"""GPU-box diagnostic: run the CUDA temperature path on a few frames and print per-stage differences
against the oracle (test infrastructure).  Usage: python tools/debug_temp_gpu.py [seed ...|demo]"""
import os, sys, time
import numpy as np
REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(REPO, "vistaf-roboskin-vision-integrated-multimodal-sensor_b200"))
sys.path.insert(0, os.path.join(REPO, "oracle"))
import cv2, torch
import synthetic as S, temperature_sensor as TS, temperature_oracle as TO

args = sys.argv[1:] or ["3"]
frames, names = [], []
for a in args:
    if a == "demo":
        frames.append(cv2.imread(os.path.join(REPO, "tests/golden/FINAL_TEMP_DEMO.jpg"))); names.append("TEMP_DEMO")
    else:
        frames.append(S.make_frame(int(a))); names.append(f"syn{a}")
fr = np.stack(frames)
plan = TS.get_plan(fr.shape[1], fr.shape[2], chunk=len(frames))
dev = torch.from_numpy(fr).cuda()
torch.cuda.synchronize(); t = time.time()
out = plan.run(dev); torch.cuda.synchronize(); print("gpu first call s", time.time() - t)
t = time.time(); out = plan.run(dev); torch.cuda.synchronize(); print("gpu second call s", time.time() - t)
rec = TS.records_to_numpy(out["records"])
roi_gpu = torch.empty((fr.shape[1], fr.shape[2]), dtype=torch.uint8, device="cuda")
import _native as N, ctypes as C
N.check(N.lib().vistaf_temp_roi_mask(plan.handle, roi_gpu.data_ptr(), None)); torch.cuda.synchronize()
for i, name in enumerate(names):
    o = TO.temperature_frame(fr[i])
    print(f"== {name}: record", {k: rec[i][k] for k in rec.dtype.names if k != 'pad_'})
    print("   oracle stats", o["stats"], "peak", o["seg"]["peak"], "phi0", o["seg"]["phi0"], "median", o["seg"]["median"],
          "meanA/B", o["seg"]["mean_a"], o["seg"]["mean_b"])
    if i == 0:
        print("   roi mismatch", int((roi_gpu.cpu().numpy().astype(bool) != o["roi"]).sum()))
        H, W = fr.shape[1:3]
        g8 = plan.workspace_plane("gray8").cpu().numpy().reshape(H, W)
        print("   gray8 mismatch", int((g8 != cv2.cvtColor(fr[0], cv2.COLOR_BGR2GRAY)).sum()))
        nrm = plan.workspace_plane("norm").view(torch.float32).cpu().numpy().reshape(H, W)
        I = nrm / np.float32(o["seg"]["I_norm"].dtype.type(1))  # norm (before /mu)
        mu = float(np.mean((o["seg"]["I_norm"])[o["roi_eff"]]))
        # oracle I_norm = norm/mu_o ; compare shapes up to scale
        ratio = nrm[o["roi_eff"]] / o["seg"]["I_norm"][o["roi_eff"]]
        print("   norm/I_norm ratio min/max (should be constant = mu)", ratio.min(), ratio.max())
        mA = plan.workspace_plane("maskA").cpu().numpy().reshape(H, W).astype(bool)
        A_or = (o["seg"]["s"] >= 0) & o["roi_eff"]
        d1 = int((mA != A_or).sum()); d2 = int((mA != ((o["seg"]["s"] < 0) & o["roi_eff"])).sum())
        print("   maskA mismatch vs oracle A / B", d1, d2)
    for k in ("sat", "dark", "light", "color_support", "color_ok"):
        g = out[k][i].cpu().numpy().astype(bool)
        print(f"   mask {k}: mismatching px {int((g != o[k]).sum())} of {int(o[k].sum())}")
    print("   source mismatch", int((out["source"][i].cpu().numpy() != o["source"]).sum()))
    for k in ("fused", "final"):
        g = out[k][i].cpu().numpy()
        nanmis = int((np.isnan(g) != np.isnan(o[k])).sum())
        d = np.abs(g - o[k]); d = d[np.isfinite(d)]
        print(f"   {k}: nan-mismatch {nanmis}  max|diff| {d.max() if d.size else -1:.3e}  mean|diff| {d.mean() if d.size else -1:.3e}  n>1e-2 {(d>1e-2).sum()}")
