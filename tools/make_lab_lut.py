"""Build-container one-off: derive the two lookup tables of OpenCV's integer BGR->Lab (8-bit) path and
verify the integer recipe bit-exactly against cv2.cvtColor on ALL 2^24 colours.  The tables are
written to calibration/lab_lut.json and uploaded to the GPU at plan creation.
(cv2.cvtColor(BGR2LAB) is the third-party call at temperature_sensor.py:284.)"""
import json, os
import numpy as np, cv2

i = np.arange(256, dtype=np.float32) / np.float32(255)
lin = np.where(i <= np.float32(0.04045), i / np.float32(12.92),
               np.power((i + np.float32(0.055)) / np.float32(1.055), np.float32(2.4))).astype(np.float32)
gam = np.rint(np.float32(2040) * lin).astype(np.int32)
x = np.arange(3072, dtype=np.float32) / np.float32(2040)
cb = np.rint(np.float32(32768) * np.where(x < np.float32(216 / 24389), x * np.float32(841 / 108) + np.float32(16 / 116),
                                          np.cbrt(x))).astype(np.int32)
C = np.array([[1777, 1541, 778], [871, 2929, 296], [73, 448, 3575]], np.int64)


def lab8(bgr):
    B, G, R = [gam[bgr[..., c]].astype(np.int64) for c in range(3)]
    f = [cb[(R * C[r, 0] + G * C[r, 1] + B * C[r, 2] + 2048) >> 12].astype(np.int64) for r in range(3)]
    L = (296 * f[1] - 1336934 + 16384) >> 15
    a = (500 * (f[0] - f[1]) + 128 * 32768 + 16384) >> 15
    b = (200 * (f[1] - f[2]) + 128 * 32768 + 16384) >> 15
    return np.stack([np.clip(v, 0, 255) for v in (L, a, b)], -1).astype(np.uint8)


bad = 0
for b0 in range(256):
    g_, r_ = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    img = np.stack([np.full_like(g_, b0), g_, r_], -1)
    bad += int((lab8(img) != cv2.cvtColor(img, cv2.COLOR_BGR2LAB)).sum())
print("mismatching channel values over 2^24 colours:", bad)
assert bad == 0
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..",
                   "vistaf-roboskin-vision-integrated-multimodal-sensor_b200", "calibration", "lab_lut.json")
json.dump(dict(gamma=gam.tolist(), cbrt=cb.tolist(), coeffs=C.tolist(), verified_against="cv2 %s, all 2^24 colours" % cv2.__version__),
          open(out, "w"))
