"""One-off (build container only): sample a stripe-colour palette for the synthetic frame
generator from the reference's demo photograph FINAL_TEMP_DEMO.jpg (DATA, not code).

The oracle's temperature map of that frame is binned in 1 degC bins; per bin the mean BGR of the
light (coloured TLC) stripes and of the dark (black TLC) stripes goes into
calibration/synthetic_palette.json.  The synthetic generator interpolates this table, so its frames
exercise both temperature models over 20..58 degC with realistic chroma.
"""
import json, os, sys
import cv2, numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "oracle"))
import temperature_oracle as T

img = cv2.imread(os.path.join(HERE, "..", "tests", "golden", "FINAL_TEMP_DEMO.jpg"))
r = T.temperature_frame(img)
er = cv2.erode(r["roi"].astype(np.uint8), np.ones((41, 41), np.uint8)) > 0       # stay off the rim
core_l = cv2.erode(r["light"].astype(np.uint8), np.ones((9, 9), np.uint8)) > 0   # stripe cores only
core_d = cv2.erode(r["dark"].astype(np.uint8), np.ones((9, 9), np.uint8)) > 0
Tm = r["final"]
rows = []
for t in range(20, 58):
    sel = er & (Tm >= t) & (Tm < t + 1)
    l, d = sel & core_l, sel & core_d
    if l.sum() < 500 or d.sum() < 500:
        continue
    rows.append(dict(T=t + 0.5, light=[float(img[..., c][l].mean()) for c in range(3)],
                     dark=[float(img[..., c][d].mean()) for c in range(3)], n=int(l.sum())))
    print(rows[-1])
out = os.path.join(HERE, "..", "vistaf-roboskin-vision-integrated-multimodal-sensor_b200", "calibration",
                   "synthetic_palette.json")
json.dump(dict(source="Final_demos_images/FINAL_TEMP_DEMO.jpg binned by oracle temperature", rows=rows),
          open(out, "w"), indent=1)
