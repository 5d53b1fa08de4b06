"""One-off (run in the build container, where /root/reference exists): copy the reference's
calibration CONSTANTS into the package as plain JSON so they travel to the GPU box.

Sources (data, not code):
  Force/Phase_to_height/calibration_out/calibration_model.json      (shape_ftp.py:20, :672)
  Force/Height_to_force/calibration_out/calibration_model.json      (force_sensor.py:27, :142)
  Temperature/Colored_Model/calibration_out/color_model_global_huber_deg2.joblib       (temperature_sensor.py:142)
  Temperature/MixedColorBlack_Model/calibration_out/black_model_global_huber_deg3.joblib (temperature_sensor.py:147)
  Temperature/*/calibration_out/models_final_summary_metrics.json   (multimodal_sensor.py:67-70)
The joblib pickles hold an sklearn Pipeline(StandardScaler, PolynomialFeatures, HuberRegressor)
(+ IsotonicRegression for the colour model); only their fitted numbers are exported.
"""
import json, os, shutil, sys, warnings
import numpy as np
warnings.filterwarnings("ignore")
import joblib

REF = "/root/reference"
PKG = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..",
                   "vistaf-roboskin-vision-integrated-multimodal-sensor_b200", "calibration")
os.makedirs(PKG, exist_ok=True)

def dump_model(path, out_name):
    o = joblib.load(path)
    pipe = o["model"]
    sc, pf, reg = (pipe.named_steps[k] for k in ("standardscaler", "polynomialfeatures", "huberregressor"))
    d = {
        "source": os.path.relpath(path, REF),
        "use_features": list(o["use_features"]),
        "degree": int(pf.degree),
        "scaler_mean": [float(x) for x in sc.mean_],
        "scaler_scale": [float(x) for x in sc.scale_],
        "powers": np.asarray(pf.powers_).astype(int).tolist(),
        "coef": [float(x) for x in reg.coef_],
        "intercept": float(reg.intercept_),
    }
    iso = o.get("isotonic_calibrator", None)
    if iso is not None:
        d["isotonic"] = {
            "x": [float(x) for x in iso.X_thresholds_],
            "y": [float(x) for x in iso.y_thresholds_],
            "x_min": float(iso.X_min_), "x_max": float(iso.X_max_),
            "out_of_bounds": str(iso.out_of_bounds),
        }
    with open(os.path.join(PKG, out_name), "w") as f:
        json.dump(d, f, indent=1)
    print("wrote", out_name)

dump_model(f"{REF}/Temperature/Colored_Model/calibration_out/color_model_global_huber_deg2.joblib", "color_model_global_huber_deg2.json")
dump_model(f"{REF}/Temperature/MixedColorBlack_Model/calibration_out/black_model_global_huber_deg3.joblib", "black_model_global_huber_deg3.json")
for src, dst in [
    ("Force/Phase_to_height/calibration_out/calibration_model.json", "phase_to_height_calibration_model.json"),
    ("Force/Height_to_force/calibration_out/calibration_model.json", "height_to_force_calibration_model.json"),
    ("Temperature/Colored_Model/calibration_out/models_final_summary_metrics.json", "color_models_final_summary_metrics.json"),
    ("Temperature/MixedColorBlack_Model/calibration_out/models_final_summary_metrics.json", "black_models_final_summary_metrics.json"),
]:
    shutil.copyfile(os.path.join(REF, src), os.path.join(PKG, dst)); os.chmod(os.path.join(PKG, dst), 0o644)
    print("copied", dst)
