"""Debug aid: saturated synthetic frames (every 16th seed carries glare blobs) through the CUDA temperature path and the
oracle; prints the mismatch statistics of the T10 (quantise + Telea fill) step."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vistaf-roboskin-vision-integrated-multimodal-sensor_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import synthetic as S
import temperature_sensor as TS
import temperature_oracle as TO

for seed in [int(x) for x in (sys.argv[1:] or ["15", "31"])]:
    fr = S.make_frame(seed)[None]
    out = TS.run_batch(fr, chunk=1)
    o = TO.temperature_frame(fr[0])
    rec = out["records_np"][0]
    print("seed", seed, "status", hex(int(rec["status"])), "sat px", int(rec["sat_pixels"]), int(o["sat"].sum()))
    for k in ("fused", "final"):
        g = out[k][0].cpu().numpy()
        nn = np.isnan(g) != np.isnan(o[k])
        d = np.abs(np.nan_to_num(g) - np.nan_to_num(o[k]))
        print(" ", k, "nan mismatch", int(nn.sum()), "max", float(d.max()), "n>1e-2", int((d > 1e-2).sum()), "n>1e-4", int((d > 1e-4).sum()))
        if k == "fused":
            sat = o["sat"]
            print("    inside sat: max", float(d[sat].max()), "n>1e-2", int((d[sat] > 1e-2).sum()), "of", int(sat.sum()))
    for k in ("color_ok", "source"):
        g = out[k][0].cpu().numpy()
        print(" ", k, "mismatch", int((g.astype(np.int32) != o[k].astype(np.int32)).sum()))
    st = o["stats"]
    print("  stats", {k: (float(rec[k]), st[k]) for k in ("mean_C", "median_C", "std_C", "min_C", "max_C")})
