"""CPU-only checks of the drop-in boundary: libvistaf.so loads, exports every symbol include/vistaf.h declares,
the ctypes mirrors of the structs have the C sizes, and the host-side scalar entry point agrees with the
oracle's restatement of the calibrated curves.  No GPU compute is called here."""
import ctypes as C
import json
import os
import re
import subprocess

import numpy as np
import pytest

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
HDR = os.path.join(REPO, "include", "vistaf.h")


@pytest.fixture(scope="module")
def native():
    import __graft_entry__ as G
    G.build()
    import _native as N
    return N


def declared_functions():
    src = re.sub(r"/\*.*?\*/", "", open(HDR).read(), flags=re.S)
    return sorted(set(re.findall(r"\b(vistaf_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    fns = declared_functions()
    for f in ("vistaf_temp_batch", "vistaf_ftp_batch", "vistaf_ftp_set_reference", "vistaf_depth_to_volume",
              "vistaf_curve_predict", "vistaf_crop_gray", "vistaf_last_error", "vistaf_profile_read"):
        assert f in fns


def test_library_exports_every_declared_symbol(native):
    lib = native.lib()
    for f in declared_functions():
        assert hasattr(lib, f), f"libvistaf.so does not export {f}"
    # and nothing in the ABI leaks C++ types: all exported vistaf_* symbols are unmangled
    out = subprocess.run(["nm", "-D", "--defined-only", native.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert set(declared_functions()) <= exported


def test_struct_sizes_match_the_header(native, tmp_path):
    """Compile a tiny C program against include/vistaf.h and compare sizeof() with the ctypes mirrors."""
    names = {"VistafPolyModel": native.PolyModel, "VistafTempParams": native.TempParams, "VistafTempRecord": native.TempRecord,
             "VistafTempOutputs": native.TempOutputs, "VistafCurveModel": native.CurveModel, "VistafFtpParams": native.FtpParams,
             "VistafFtpRecord": native.FtpRecord, "VistafFtpOutputs": native.FtpOutputs,
             "VistafFtpReferenceInfo": native.FtpReferenceInfo, "VistafStageTime": native.StageTime}
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "vistaf.h"\nint main(void){' +
                   "".join(f'printf("{n} %zu\\n", sizeof({n}));' for n in names) + "return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(REPO, "include"), "-o", str(exe), str(src)])
    got = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True).stdout.splitlines())
    for n, ct in names.items():
        assert int(got[n]) == C.sizeof(ct), n
    assert C.sizeof(native.FtpRecord) == 104 and C.sizeof(native.TempRecord) == 80


def test_curve_predict_matches_oracle(native):
    import ftp_oracle as FO
    lib = native.lib()
    force = json.load(open(os.path.join(native.CAL_DIR, "height_to_force_calibration_model.json")))["best_model"]
    m = native.curve_model(force)
    for v in (0.0, 0.01, 0.2254, 0.5, -0.1):
        assert lib.vistaf_curve_predict(C.byref(m), v) == pytest.approx(FO.force_from_volume(force, v), rel=1e-14, abs=1e-300)
    model, _ = FO.load_phase_to_height()
    mh = native.curve_model(model)
    xs = np.array([0.0, 1e-3, 0.3, 1.2, 5.0])
    want = FO.model_predict(model, xs)
    for x, w in zip(xs, want):
        assert lib.vistaf_curve_predict(C.byref(mh), float(x)) == pytest.approx(float(w), rel=1e-13, abs=1e-15)
    generic = {"linear0": {"a": 2.0}, "linear": {"a": 2.0, "b": 1.0}, "poly2": {"c2": 1.0, "c1": 2.0, "c0": 3.0},
               "sat_exp": {"a": 2.0, "b": 3.0}}
    expect = {"linear0": 1.0, "linear": 2.0, "poly2": 4.25, "sat_exp": 2.0 * (1 - np.exp(-1.5))}
    for t, p in generic.items():
        mm = native.curve_model({"type": t, "params": p})
        assert lib.vistaf_curve_predict(C.byref(mm), 0.5) == pytest.approx(expect[t], rel=1e-14)
    with pytest.raises(ValueError):
        native.curve_model({"type": "spline", "params": {}})


def test_no_cpu_fallback_without_cuda(native):
    """The product modules refuse to compute without a CUDA device instead of silently using the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    import force_sensor as FS
    import shape_ftp as SF
    import temperature_sensor as TS
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        TS.run_batch(np.zeros((1, 8, 8, 3), np.uint8))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        SF.run_batch(np.zeros((8, 8, 3), np.uint8), np.zeros((1, 8, 8, 3), np.uint8))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        FS.depth_map_to_volume_cm3(np.zeros((4, 4), np.float32), np.ones((4, 4), bool), 0.03)
    # pure host pieces keep working: force model, calibration loading, error behaviour of the reference API
    best = FS.load_force_calibration(FS.FORCE_CALIBRATION_JSON)["best_model"]
    assert FS.predict_force_from_volume(best, 0.0) == 0.0
    with pytest.raises(ValueError):
        FS.predict_force_from_volume({"type": "nope", "params": {}}, 1.0)
    with pytest.raises(RuntimeError):
        FS.estimate_mm_per_px(None)
    assert FS.estimate_mm_per_px(65.0) == pytest.approx(2.0 / 65.0)
