"""ORACLE / baseline infrastructure (never imported by the product path): vendor the UNMODIFIED reference scripts of the
hot path and the calibration files they read, from /root/reference into oracle/_ref/ (git-ignored, NOT gpurun-ignored,
so it travels to the GPU box like a built .so).  `bench.py --impl reference` then times the reference itself on the
box's host cores.  Run by __graft_entry__.build() whenever /root/reference is present; nothing is edited, nothing from
here is ever committed."""
import os
import shutil
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
FILES = [
    "Code/shape_ftp.py", "Code/force_sensor.py", "Code/temperature_sensor.py", "Code/multimodal_sensor.py",
    "Force/Phase_to_height/calibration_out/calibration_model.json",
    "Force/Height_to_force/calibration_out/calibration_model.json",
    "Temperature/Colored_Model/calibration_out/color_model_global_huber_deg2.joblib",
    "Temperature/MixedColorBlack_Model/calibration_out/black_model_global_huber_deg3.joblib",
    "Temperature/Colored_Model/calibration_out/models_final_summary_metrics.json",
    "Temperature/MixedColorBlack_Model/calibration_out/models_final_summary_metrics.json",
]


def main() -> int:
    if not os.path.isdir(REF):
        print("oracle/make_ref.py: /root/reference absent (GPU box): using the prebuilt oracle/_ref if any")
        return 0
    for f in FILES:
        src, dst = os.path.join(REF, f), os.path.join(DST, f)
        if not os.path.exists(src):
            print(f"oracle/make_ref.py: missing {src}", file=sys.stderr)
            return 1
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not os.path.exists(dst) or os.path.getmtime(src) > os.path.getmtime(dst) or os.path.getsize(src) != os.path.getsize(dst):
            shutil.copyfile(src, dst)
    return 0


if __name__ == "__main__":
    sys.exit(main())
