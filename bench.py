"""bench.py — frames/s of the VISTAF multimodal sensing hot path on B200 (contract in the task statement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload multimodal|ftp|temperature]
                  [--frames F] [--impl reference]

A step = one pass of the hot path over one batch of F synthetic camera-resolution frames per GPU: for the
FTP half the reference frame is (re)demodulated every step, then F deformed frames are processed against it;
the temperature half processes the same F frames.  `value` = frames/s over all ranks with the frames already
resident in HBM (CUDA events, max over ranks); `e2e` = the same through the public Python API
(multimodal_sensor.MultimodalPlan -> ctypes -> C ABI) from pinned HOST buffers, with the H2D copy of the
frames and the D2H read of the per-frame records inside the timed region.  `roofline` is reported for the
pipeline stage (kernel group between two CUDA events on the launch stream, see vistaf_profile_*) that takes
the largest share of the step; `roofline_pipeline` is the whole-step figure against SURVEY.md 8(d)'s B_alg.
`--impl reference` times the reference's own CPU implementation on the host cores with a process pool: the UNMODIFIED
reference scripts when oracle/_ref holds them (vendored at build time by oracle/make_ref.py, git-ignored; kind
"reference"), else oracle/ (the numpy+OpenCV restatement pinned against the unmodified reference; kind "port").
The timed step computes everything the module API returns: masks=True, and the gather of the per-frame records
(the path's only collective) is inside the timed region.  `parity` compares the GPU records of the frames the CPU
arm processed in the same run.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(REPO, "vistaf-roboskin-vision-integrated-multimodal-sensor_b200")
sys.path.insert(0, PKG)

METRIC = "multimodal frames/sec (FTP+force+temp) at 1/2/4/8 B200; % HBM roofline"
# algorithmic bytes per frame (a property of the work, not of the implementation), and their split
# over the pipeline stages the library marks with CUDA events (DESIGN.md "Stage table").  MB per frame.
B_ALG = {"temperature": 1000e6, "ftp": 760e6, "multimodal": 1760e6, "ftp4k": 8300e6}
STAGE_MB = {
    "F.crop_gray": 5.6, "F.sobel_percentiles_bad": 9.1, "F.telea_inpaint": 3.5, "F.illum_blur45": 22.4,
    "F.preblur_apo_median": 11.2, "F.carrier_dft_field": 109.0, "F.quality_reliable_mask": 35.0,
    "F.unwrap_unionfind": 18.2, "F.irls_plane": 126.0, "F.irls_quadric_contact": 141.4,
    "F.irls_quadric_background": 126.0, "F.smooth_sign": 40.6, "F.chamfer_frontier_assemble": 91.0,
    "F.blob_filter_force": 30.8,
    "T.gray_sat_median": 33.2, "T.illum_blur20": 116.2, "T.fft_rows": 99.5, "T.fft_cols_peaks": 132.7,
    "T.band_ifft": 315.2, "T.morph_close_open": 24.9, "T.features_models_fuse": 83.0, "T.final_blur": 132.7,
    "T.stats_median": 66.4,
}
N_DISTINCT = 64         # distinct synthetic frames generated on the host per rank (seeds rank*100003 + 1 ...: one glare frame in 16), tiled to the batch on the device


def clocks_sampler(stop, samples, n_gpus=1):
    """Rank 0 samples the clocks and throttle reasons of all the job's GPUs (one nvidia-smi call for GPUs 0..n_gpus-1):
    one poller per node instead of one per rank, so the polling itself does not perturb the other ranks' launches."""
    q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    ids = ",".join(str(i) for i in range(max(1, n_gpus)))
    while not stop.is_set():
        try:
            o = subprocess.run(["nvidia-smi", "-i", ids, f"--query-gpu={q}",
                                "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
            for ln in o.strip().splitlines():
                samples.append([x.strip() for x in ln.split(",")])
        except Exception:
            pass
        stop.wait(0.25)


def summarize_clocks(samples):
    sm = sorted(int(s[0]) for s in samples if s and s[0].isdigit())
    reasons = set()
    for s in samples:
        for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], s[2:6]):
            if v.lower().startswith("active"):
                reasons.add(name)
    mx = [int(s[1]) for s in samples if len(s) > 1 and s[1].isdigit()]
    return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons)}


# ------------------------------------------------------------------ CPU arm: the oracle on the host cores
_W = {}


def _cpu_init():
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import cv2
    cv2.setNumThreads(1)                      # one worker process per core; no nested OpenCV thread pools
    import synthetic as S
    import ftp_oracle as FO
    import temperature_oracle as TO
    _W.update(cv2=cv2, S=S, FO=FO, TO=TO, frames={})


def _cpu4k_prepare(seed):
    if not _W:
        _cpu_init()
    S, FO = _W["S"], _W["FO"]
    if "refp4k" not in _W:
        g = S.CROP4K
        P = FO.FtpParams(apply_global_shift=False, use_ecc=False)
        disc = FO.disc(g["h"], g["w"], g["cx"], g["cy"], g["r"])
        geom = dict(g, circ=disc, roi=disc, apo=FO.apodization(g["h"], g["w"], g["cx"], g["cy"], g["r"], P.apod_taper))
        ref = S.make_crop4k(0, reference=True)
        _W.update(P4k=P, geom4k=geom, ref4k=ref, refp4k=FO.reference_products(ref, geom, P), crops={})
    if seed not in _W["crops"]:
        _W["crops"][seed] = S.make_crop4k(seed)
    return seed


def _cpu4k_process(args):
    seed, _ = args
    _cpu4k_prepare(seed)
    FO = _W["FO"]
    t = time.perf_counter()
    r = FO.ftp_pair(_W["ref4k"], _W["crops"][seed], _W["geom4k"], _W["P4k"], ref=_W["refp4k"])
    out = {"seed": seed}
    if r is not None:
        fr = FO.force_readings(r)
        out.update(force_N=fr["force_N"], volume_cm3=fr["volume_cm3"], reliable_px=int(r["reliable"].sum()),
                   output_reliable_px=int(r["output_reliable"].sum()), contact_dilated_px=int(r["contact_dilated"].sum()))
    return time.perf_counter() - t, out


def _cpu_prepare(seed):
    """Generate (untimed) and cache this worker's frame; build the reference-frame products once per worker."""
    if not _W:
        _cpu_init()
    S, FO, cv2 = _W["S"], _W["FO"], _W["cv2"]
    if "refp" not in _W:
        ref = S.make_frame(0, reference=True)
        P = FO.FtpParams(apply_global_shift=False, use_ecc=False)
        geom = FO.geometry(ref.shape[0], ref.shape[1], P)
        rg = cv2.cvtColor(ref[geom["y1"]:geom["y2"], geom["x1"]:geom["x2"]], cv2.COLOR_BGR2GRAY)
        _W.update(P=P, geom=geom, rg=rg, refp=FO.reference_products(rg, geom, P))
    if seed not in _W["frames"]:
        _W["frames"][seed] = S.make_frame(seed)
    return seed


def _cpu_process(args):
    """One frame through the oracle; returns (seconds, readings) -- the readings feed the `parity` block of the line."""
    seed, workload = args
    _cpu_prepare(seed)
    FO, TO, cv2, g = _W["FO"], _W["TO"], _W["cv2"], _W["geom"]
    f = _W["frames"][seed]
    out = {"seed": seed}
    t = time.perf_counter()
    if workload in ("ftp", "multimodal"):
        dg = cv2.cvtColor(f[g["y1"]:g["y2"], g["x1"]:g["x2"]], cv2.COLOR_BGR2GRAY)
        r = FO.ftp_pair(_W["rg"], dg, g, _W["P"], ref=_W["refp"])
        if r is not None:
            fr = FO.force_readings(r)
            out.update(force_N=fr["force_N"], volume_cm3=fr["volume_cm3"], reliable_px=int(r["reliable"].sum()),
                       output_reliable_px=int(r["output_reliable"].sum()), contact_dilated_px=int(r["contact_dilated"].sum()))
    if workload in ("temperature", "multimodal"):
        o = TO.temperature_frame(f)
        out.update(T_mean=o["stats"]["mean_C"], T_median=o["stats"]["median_C"], T_max=o["stats"]["max_C"],
                   valid_pixels=o["stats"]["valid_pixels"], dark_px=int(o["dark"].sum()), color_ok_px=int(o["color_ok"].sum()))
    return time.perf_counter() - t, out


# ---- the unmodified reference (oracle/_ref, vendored by oracle/make_ref.py at build time): one frame per call
def _ref_init():
    from unittest.mock import MagicMock
    root = os.path.join(REPO, "oracle", "_ref")
    for n in ["matplotlib", "matplotlib.pyplot", "mpl_toolkits", "mpl_toolkits.mplot3d", "skimage", "skimage.restoration"]:
        sys.modules.setdefault(n, MagicMock())          # figures only; absent on this image
    plt = sys.modules["matplotlib.pyplot"]
    sys.modules["matplotlib"].pyplot = plt
    plt.subplots = lambda *a, **k: (MagicMock(), MagicMock())
    sys.path.insert(0, os.path.join(root, "Code"))
    os.chdir(root)
    import warnings
    warnings.filterwarnings("ignore")
    import io
    import contextlib
    with contextlib.redirect_stdout(io.StringIO()):
        import shape_ftp, force_sensor, temperature_sensor          # the reference's modules, unmodified
    import cv2
    import tempfile
    import synthetic as S
    shape_ftp.APPLY_GLOBAL_SHIFT = False
    shape_ftp.USE_ECC_CROP_ALIGNMENT = False
    tmp = tempfile.mkdtemp()
    refp = os.path.join(tmp, "ref.png")
    cv2.imwrite(refp, S.make_frame(0, reference=True))
    _W.update(ref_mods=(shape_ftp, force_sensor, temperature_sensor), cv2=cv2, S=S, tmp=tmp, refp=refp, paths={})


def _ref_prepare(seed):
    if "ref_mods" not in _W:
        _ref_init()
    if seed not in _W["paths"]:
        p = os.path.join(_W["tmp"], f"f{seed}.png")
        _W["cv2"].imwrite(p, _W["S"].make_frame(seed))          # lossless: the same bytes the GPU arm sees
        _W["paths"][seed] = p
    return seed


def _ref_process(args):
    import io
    import contextlib
    seed, workload = args
    _ref_prepare(seed)
    sf, fs, ts = _W["ref_mods"]
    out = {"seed": seed}
    t = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        if workload in ("ftp", "multimodal"):
            r = sf.main(reference_path=_W["refp"], deformed_path=_W["paths"][seed], output_dir=_W["tmp"], batch_mode=True,
                        save_summary_figures=False, export_heightmaps=False, debug=False, return_results=True)
            if r is not None:
                hm = r["height_map_mm_crop"]
                mmpp = 2.0 / r["estimated_grating_period_px"]
                import numpy as np
                V, A, D = fs.depth_map_to_volume_cm3(hm, np.isfinite(hm), mmpp, 0.01)
                out.update(force_N=fs.predict_force_from_volume(fs.load_force_calibration(fs.FORCE_CALIBRATION_JSON)["best_model"], V),
                           volume_cm3=V)
        if workload in ("temperature", "multimodal"):
            ts.INPUT_IMAGE_PATH, ts.OUTPUT_DIR = _W["paths"][seed], os.path.join(_W["tmp"], f"t{seed}")
            ts.main()
    return time.perf_counter() - t, out


def have_reference():
    return os.path.exists(os.path.join(REPO, "oracle", "_ref", "Code", "shape_ftp.py"))


class CpuArm:
    """Process pool over the host cores; each step pushes `frames_per_step` synthetic frames through the oracle."""

    def __init__(self, workload, workers=None, frames_per_step=None, kind="port"):
        import multiprocessing as mp
        avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        self.workers = int(workers or max(1, min(avail, 32)))
        self.frames = int(frames_per_step or self.workers)
        self.workload, self.kind = workload, kind
        self._prep, self._proc = (_ref_prepare, _ref_process) if kind == "reference" else (_cpu_prepare, _cpu_process)
        if workload == "ftp4k":               # 3904^2 crops: ~45 s and a few GB per frame and process
            self._prep, self._proc = _cpu4k_prepare, _cpu4k_process
            self.workers = int(workers or max(1, min(avail, 8)))
            self.frames = int(frames_per_step or self.workers)
        self.pool = mp.get_context("spawn").Pool(self.workers, initializer=_ref_init if kind == "reference" else _cpu_init)
        self.seeds = list(range(1, self.frames + 1))
        self.pool.map(self._prep, self.seeds, chunksize=max(1, self.frames // self.workers))     # untimed: generation + reference products
        self.readings = []

    def step(self):
        t = time.perf_counter()
        res = self.pool.map(self._proc, [(s, self.workload) for s in self.seeds], chunksize=max(1, self.frames // self.workers))
        self.readings = [r[1] for r in res]
        return time.perf_counter() - t

    def close(self):
        self.pool.close()
        self.pool.join()

    def describe(self, seconds, steps):
        what = ("the UNMODIFIED reference scripts (oracle/_ref: shape_ftp.main + force_sensor + temperature_sensor.main per "
                "frame, its own file reads and writes included)" if self.kind == "reference" else
                "oracle/ (numpy+OpenCV restatement of the reference, C heap unwrap)")
        return (f"{steps} step(s) x {self.frames} synthetic 2160x3840 frames ({self.workload}) through {what} on a pool of "
                f"{self.workers} processes, {seconds:.1f} s of wall time; frame generation excluded; "
                f"os.cpu_count()={os.cpu_count()}, one OpenCV thread per process")


def run_reference(a, config):
    kind = "reference" if have_reference() and not a.cpu_port and a.workload != "ftp4k" else "port"   # the reference's ROI constants are fixed
    arm = CpuArm(a.workload, a.cpu_workers, a.cpu_frames, kind)
    try:
        # bounded sample: the reference needs ~45 s per frame and process (pure-Python heap unwrap), so one frame per
        # worker and one timed pass; the port gets one warm pass (page-in, imports) and up to --cpu-steps passes
        warm = 0 if kind == "reference" else min(a.warmup, 1)
        for _ in range(warm):
            arm.step()
        steps = 1 if kind == "reference" else max(1, min(a.steps, a.cpu_steps))
        dt = sum(arm.step() for _ in range(steps))
    finally:
        arm.close()
    fps = arm.frames * steps / dt
    return {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": a.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": 1000.0 * dt / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": arm.workers, "kind": kind,
                             "sample": arm.describe(dt, steps)},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


# ------------------------------------------------------------------ GPU arm
_REAL_STDOUT = None


def emit(line: dict):
    """The ONE JSON line of the contract, on the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)          # libraries (NCCL's version banner, pool workers) may write to fd 1: send them to stderr
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=os.environ.get("VISTAF_WORKLOAD", "multimodal"),
                    choices=["multimodal", "ftp", "temperature", "ftp4k"])
    ap.add_argument("--frames", type=int, default=256, help="frames per GPU per step")
    ap.add_argument("--chunk", type=int, default=256, help="FTP frames per library chunk (workspace size)")
    ap.add_argument("--chunk-temp", type=int, default=64, help="temperature frames per library chunk")
    ap.add_argument("--lanes-ftp", type=int, default=4, help="concurrent slices (streams + workspaces) of the FTP batch")
    ap.add_argument("--lanes-temp", type=int, default=2, help="concurrent slices of the temperature batch")
    ap.add_argument("--distinct", type=int, default=N_DISTINCT,
                    help="distinct synthetic seeds per rank (consecutive seeds: one glare frame per 16, the T10 path)")
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--cpu-workers", type=int, default=None)
    ap.add_argument("--cpu-frames", type=int, default=None, help="frames per CPU step (default: one per worker)")
    ap.add_argument("--cpu-steps", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-port", action="store_true", help="--impl reference: time oracle/ even when oracle/_ref holds the reference")
    ap.add_argument("--no-maps", action="store_true", help="skip the e2e_maps leg (maps brought back to the host)")
    a = ap.parse_args()
    globals()["N_DISTINCT"] = max(1, a.distinct)
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    config = {"workload": f"{a.workload}: {a.frames} synthetic 2160x3840 BGR fringe+TLC frames per GPU per step "
                          f"({N_DISTINCT} distinct seeds per rank tiled to the batch) against one reference frame, "
                          f"FTP crop 1182^2 / FFT grid 1374^2, both temperature models; frames pre-registered "
                          f"(APPLY_GLOBAL_SHIFT=False, USE_ECC_CROP_ALIGNMENT=False on GPU and CPU arms)",
              "timed_step": "reference frame + F frames -> height maps, temperature maps, all 11 masks, records; gather of the "
                            "records to rank 0 (NCCL) inside the timed region",
              "frames_per_gpu": a.frames, "chunk_ftp": a.chunk, "chunk_temp": a.chunk_temp, "lanes_ftp": a.lanes_ftp,
              "lanes_temp": a.lanes_temp, "resolution": "2160x3840",
              "l2": f"inputs {a.frames * 24.9:.0f} MB per step, larger than the 126 MB L2 (no flush needed)"}

    if a.impl == "reference":
        if rank != 0:
            return
        emit(run_reference(a, config))
        return

    if a.workload == "ftp4k":
        return main_ftp4k(a, rank, world, local)
    import numpy as np
    import torch
    import torch.distributed as dist
    import synthetic as S
    import _native as N
    import multimodal_sensor as MS
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    F = a.frames
    ref_host = S.make_frame(0, reference=True)
    seeds = [rank * 100003 + s + 1 for s in range(N_DISTINCT)]
    if N_DISTINCT > 8:                        # the numpy generator costs ~1 s per frame: spread it over a few processes
        import multiprocessing as mp
        avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        with mp.get_context("spawn").Pool(max(1, min(16, avail // max(1, world)))) as pool:
            host = np.stack(pool.map(S.make_frame, seeds))
    else:
        host = np.stack([S.make_frame(sd) for sd in seeds])
    # pinned host staging buffer: the N_DISTINCT frames once (keeps host RAM at 200 MB per rank); a step's batch is
    # frame i = distinct frame i % N_DISTINCT, both for the resident copy and for every host->device copy of the e2e leg
    pinned = torch.from_numpy(np.ascontiguousarray(host)).pin_memory()
    ref_pinned = torch.from_numpy(ref_host).pin_memory()
    dev = pinned.cuda().repeat((F + N_DISTINCT - 1) // N_DISTINCT, 1, 1, 1)[:F].contiguous()
    ref_dev = ref_pinned.cuda()
    plan = MS.MultimodalPlan(host.shape[1], host.shape[2], min(a.chunk, F), min(a.chunk_temp, F), a.workload,
                             lanes_ftp=a.lanes_ftp, lanes_temp=a.lanes_temp)

    import sharding as SH

    def gather(res):
        """The path's only collective: per-frame records of every rank to rank 0 (inside the timed region)."""
        return SH.gather_records(res["records"], world * F, 0) if world > 1 else res["records"]

    def step_resident():
        res = plan.step(ref_dev, dev, masks=True)          # everything the module API returns: maps, all masks, records
        return gather(res)

    # two device staging buffers: the host->device copy of step n+1 runs while step n computes on the other buffer
    stage = [(torch.empty_like(dev), torch.empty_like(ref_dev)) for _ in range(2)]
    stage_free = [None, None]                      # event: the step that last read the buffer has finished
    stage_turn = [0]

    copy_stream = torch.cuda.Stream()
    n_slices = 2 * max(1, a.lanes_ftp, a.lanes_temp)

    def step_e2e():
        """Host -> device staging of the reference and of the frames in slices on a copy stream (events mark each slice
        ready, so a lane starts as soon as its frames have landed), the step through the public API, records to host."""
        main = torch.cuda.current_stream()
        k = stage_turn[0]
        stage_turn[0] ^= 1
        dev2, ref2 = stage[k]
        if stage_free[k] is not None:
            copy_stream.wait_event(stage_free[k])  # the step two calls ago has finished reading this buffer
        ready = []
        with torch.cuda.stream(copy_stream):
            ref2.copy_(ref_pinned, non_blocking=True)
            ev_ref = torch.cuda.Event()
            ev_ref.record(copy_stream)
            per = -(-F // n_slices)
            per = -(-per // N_DISTINCT) * N_DISTINCT
            starts = list(range(0, F, per))
            # staging order: the first slices of every lane first (the temperature lanes start at frame 0 and F/2)
            half = len(starts) // 2
            order = starts[:2] + starts[half:half + 1] + starts[2:half] + starts[half + 1:] if len(starts) >= 4 else starts
            for lo in order:
                hi = min(F, lo + per)
                for c in range(lo, hi, N_DISTINCT):              # F x 24.9 MB cross PCIe every step
                    n = min(N_DISTINCT, hi - c)
                    dev2[c:c + n].copy_(pinned[:n], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy_stream)
                ready.append((lo, hi, ev))
        res = plan.step(ref2, dev2, masks=True, ready=ready, ref_ready=ev_ref)
        stage_free[k] = torch.cuda.Event()
        stage_free[k].record(main)
        g = gather(res)
        return res, (g.cpu() if g is not None else None)

    # maps-out variant: the height crops and the ROI window of both temperature maps come back to pinned host memory too
    tb = plan.temp
    box = (tb.roi_bbox() if tb is not None else None)
    pin = {}

    out_stream = torch.cuda.Stream()

    def step_e2e_maps():
        """e2e + the maps: device-side snapshots of the height crops and of the ROI window of both temperature maps are
        copied to pinned host memory on a copy-out stream, so the transfer of step n overlaps the compute of step n+1
        (the snapshot frees the plan's reused output buffers at once; the pinned buffers are the consumer's)."""
        res, rec = step_e2e()
        main = torch.cuda.current_stream()
        snap = {}
        if "ftp" in res:
            snap["height"] = res["ftp"]["height_mm"].clone()
        if "temp" in res:
            y0, y1, x0, x1 = box
            for k in ("fused", "final"):
                snap[k] = res["temp"][k][:, y0:y1, x0:x1].contiguous()
        out_stream.wait_stream(main)
        with torch.cuda.stream(out_stream):
            for k, v in snap.items():
                if k not in pin:
                    pin[k] = torch.empty(v.shape, dtype=v.dtype).pin_memory()
                pin[k].copy_(v, non_blocking=True)
                v.record_stream(out_stream)
        return rec

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_ms = [0.0]

    def timed(fn, steps, tail=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
        host_ms[0] = (time.perf_counter() - t0) * 1e3 / steps      # host time to ENQUEUE one step (no sync inside; includes launch-queue back-pressure)
        if tail is not None:
            tail()                                                 # e.g. the timed stream waits for the last copy-out
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    W = max(a.warmup, 3)
    for _ in range(W):
        step_resident()
    stop, samples = threading.Event(), []
    th = threading.Thread(target=clocks_sampler, args=(stop, samples, world), daemon=True) if rank == 0 else None
    if th:
        th.start()
    l0 = N.launch_count()
    ms = timed(step_resident, a.steps)           # the measured region: lanes concurrent, stage timers off
    launches = N.launch_count() - l0
    host_enqueue_ms = host_ms[0]
    # stage table: the same steps once more with every slice serialised on one stream, so that the CUDA events the
    # library records at the stage boundaries (on the launch stream) bracket exactly one stage each
    def set_serial(on):
        plan.concurrent = not on
        for h in (plan.ftp, plan.temp):
            if h is not None:
                h.active_lanes = 1 if on else h.lanes
    set_serial(True)
    step_resident()
    N.profile_enable(True)
    ms_serial = timed(step_resident, a.steps)
    stages = N.profile_read()
    N.profile_enable(False)
    set_serial(False)
    stop.set()
    if th:
        th.join()
    step_e2e()
    e2e_steps = max(1, a.steps // 2)
    ms_e2e = timed(step_e2e, e2e_steps)
    ms_maps = None
    if not a.no_maps:
        step_e2e_maps()
        torch.cuda.synchronize()
        ms_maps = timed(step_e2e_maps, e2e_steps, tail=lambda: torch.cuda.current_stream().wait_stream(out_stream))
    res = plan.step(ref_dev, dev, masks=True)
    gathered = gather(res)
    rec = res["records"]
    replays = int(N.lib().vistaf_graph_replays())
    if rank == 0:
        assert gathered.shape[0] == world * F
        recs = plan.records_host(res, dev)
        fps = world * F * a.steps / (ms / 1000.0)
        fps_e2e = world * F * e2e_steps / (ms_e2e / 1000.0)
        pk_path = os.path.join(REPO, "MEASURED_PEAKS.json")
        peaks = json.load(open(pk_path)) if os.path.exists(pk_path) else {}
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "MEASURED_PEAKS.json hbm_gbs (copy bandwidth)" if peaks else "fallback from B200_PROFILING.md"
        tot = sum(s[1] for s in stages) or 1.0
        table = [{"stage": n, "ms_per_frame": t / (a.steps * F), "share": t / tot, "launches_per_step": k // max(1, a.steps),
                  "alg_MB_per_frame": STAGE_MB.get(n),
                  "GBps": (STAGE_MB[n] * 1e6 * a.steps * F / (t / 1000.0) / 1e9) if n in STAGE_MB and t > 0 else None}
                 for n, t, c, k in stages]
        line = {"metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": a.steps, "warmup": W,
                "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": config, "clocks": summarize_clocks(samples),
                "e2e": {"value": fps_e2e, "unit": "frames/s",
                        "h2d_bytes_per_step": int(dev.numel() + ref_dev.numel()), "d2h_bytes_per_step": int(rec.numel()),
                        "what": "records-only: frames host->device, all maps and masks computed on the device, the per-frame "
                                "records (force, volume, area, depth, temperature statistics, status) device->host"},
                "gpu_launches": int(launches), "graph_replays_total": replays, "ms_per_step_serialised": ms_serial / a.steps,
                "host_enqueue_ms_per_step": host_enqueue_ms}
        if ms_maps is not None:
            d2h = int(rec.numel()) + sum(int(v.numel() * v.element_size()) for v in pin.values())
            line["e2e_maps"] = {"value": world * F * e2e_steps / (ms_maps / 1000.0), "unit": "frames/s",
                                "h2d_bytes_per_step": int(dev.numel() + ref_dev.numel()), "d2h_bytes_per_step": d2h,
                                "what": "as e2e, plus the float32 height crop (1182^2) and the ROI window of the fused and final "
                                        "temperature maps of every frame copied to pinned host memory"}
        ach = B_ALG[a.workload] * (fps / world) / 1e9
        tr_path = os.path.join(REPO, "profiles", "ncu_stage_traffic.json")
        tr = json.load(open(tr_path)) if os.path.exists(tr_path) else {}
        tot_tr = tr.get("_pipeline", {}).get(a.workload)
        line["roofline"] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                            "traffic": tot_tr["dram_bytes_per_frame"] if tot_tr else None,
                            "traffic_note": (f"ncu dram read+write summed over every kernel of one step, per frame ({tot_tr['source']})")
                                            if tot_tr else None,
                            "kernel": "whole pipeline (every kernel of the step)", "frac_of_8TBps_nominal": ach / 8000.0,
                            "alg_bytes_per_frame": B_ALG[a.workload], "peak_source": peak_src,
                            "note": "algorithmic bytes per frame (SURVEY.md 8d: a property of the work) x frames/s per GPU / measured "
                                    "copy bandwidth; the step is ~150 kernels per chunk, none above 8 % of it, so the whole-pipeline "
                                    "figure leads and the largest bandwidth stage follows in roofline_kernel"}
        # per-kernel entry: the bandwidth stage with the largest share (the Telea march is a latency-bound serial chain hidden
        # under the other lanes in the timed region, not a bandwidth stage)
        table_f = [r for r in table if r["GBps"] and "telea" not in r["stage"]]
        top = max(table_f, key=lambda r: r["share"]) if table_f else None
        if top:
            t = tr.get(top["stage"])
            line["roofline_kernel"] = {"bound": "hbm", "achieved": top["GBps"], "peak": peak, "unit": "GB/s",
                                       "frac": top["GBps"] / peak, "traffic": t["dram_bytes_per_launch"] if t else None,
                                       "traffic_note": (f"ncu dram read+write of {t['kernel']} per launch of {t['frames_per_launch']} "
                                                        f"frames ({t['source']})") if t else None,
                                       "kernel": top["stage"], "share_of_step": top["share"],
                                       "note": "algorithmic MB/frame of the stage x frames / stage time from CUDA events on the launch "
                                               "stream, in a second pass of the same steps with the lanes serialised on one stream"}
        line["stages"] = table
        st_f = [int(r["ftp"]["status"]) for r in recs] if a.workload != "temperature" else []
        line["check"] = {"force_N_first": float(recs[0]["ftp"]["force_N"]), "T_mean_first": float(recs[0]["temp"]["mean_C"]),
                         "ftp_status_or": int(np.bitwise_or.reduce(st_f)) if st_f else 0}
        if world == 1 and not a.no_cpu:
            arm = CpuArm(a.workload, a.cpu_workers, a.cpu_frames)
            try:
                dt = arm.step()
            finally:
                arm.close()
            line["cpu_baseline"] = {"value": arm.frames / dt, "unit": "frames/s", "cores": arm.workers, "kind": "port",
                                    "sample": arm.describe(dt, 1),
                                    "reference_python_fps_build_container": {"ftp": 0.025, "temperature": 0.15, "multimodal": 0.018}[a.workload],
                                    "reference_note": "the reference as shipped, single process, timed at survey time in the build "
                                                      "container (BASELINE.md section 3); `--impl reference` times it on this box"}
            # parity at the size of the run: the CPU arm processed seeds 1..n, the GPU batch holds them as frames 0..n-1
            par = {"frames": 0, "max_rel_force": 0.0, "max_abs_T": 0.0, "mask_px_diff": 0, "against": "oracle/ (cpu_baseline leg of this run)"}
            for o in arm.readings:
                i = o["seed"] - 1
                if i >= len(recs) or i >= N_DISTINCT:
                    continue
                par["frames"] += 1
                if "force_N" in o:
                    f = recs[i]["ftp"]
                    par["max_rel_force"] = max(par["max_rel_force"], abs(float(f["force_N"]) - o["force_N"]) / max(1e-9, abs(o["force_N"])))
                    par["mask_px_diff"] += abs(int(f["reliable_px"]) - o["reliable_px"]) + abs(int(f["output_reliable_px"]) - o["output_reliable_px"]) \
                        + abs(int(f["contact_dilated_px"]) - o["contact_dilated_px"])
                if "T_mean" in o:
                    t = recs[i]["temp"]
                    par["max_abs_T"] = max(par["max_abs_T"], abs(float(t["mean_C"]) - o["T_mean"]), abs(float(t["median_C"]) - o["T_median"]),
                                           abs(float(t["max_C"]) - o["T_max"]))
                    par["mask_px_diff"] += abs(int(t["valid_pixels"]) - o["valid_pixels"]) + abs(int(t["dark_pixels"]) - o["dark_px"]) \
                        + abs(int(t["color_ok_pixels"]) - o["color_ok_px"])
            par["bars"] = {"force_rel": 1e-3, "T_abs_C": 1e-2, "mask_px": 0}
            par["ok"] = bool(par["frames"] > 0 and par["max_rel_force"] <= 1e-3 and par["max_abs_T"] <= 1e-2)
            line["parity"] = par
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def main_ftp4k(a, rank, world, local):
    """BASELINE.json configs[4]: FTP-only on 3904^2 gray crops (4096^2 FFT grid), phase amplitude up to 3 pi, frames
    streamed in steps of F crops per GPU: the reference crop is demodulated every step, the crops are staged from pinned
    host memory on a copy stream while the previous step computes (e2e), or resident (value)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import synthetic as S
    import _native as N
    import shape_ftp as SF
    import sharding as SH
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    F = a.frames if a.frames != 256 else 32
    nd = min(8, max(1, a.distinct))
    g = S.CROP4K
    config = {"workload": f"ftp4k: {F} synthetic 3904x3904 gray fringe crops per GPU per step ({nd} distinct seeds per rank tiled to "
                          f"the batch; phase amplitude up to 3 pi, wrap counts != 0) against one reference crop, FFT grid 4096^2, "
                          f"ROI radius 1951 px; FTP + force only",
              "frames_per_gpu": F, "lanes_ftp": 2, "resolution": "3904x3904 crop",
              "l2": f"inputs {F * 15.2:.0f} MB per step, larger than the 126 MB L2 (no flush needed)"}
    seeds = [rank * 100003 + s + 21 for s in range(nd)]
    host = np.stack([S.make_crop4k(sd) for sd in seeds])
    pinned = torch.from_numpy(host).pin_memory()
    ref_pinned = torch.from_numpy(S.make_crop4k(0, reference=True)).pin_memory()
    dev = pinned.cuda().repeat((F + nd - 1) // nd, 1, 1)[:F].contiguous()
    ref_dev = ref_pinned.cuda()
    plan = SF.FtpPlan(g["h"], g["w"], chunk=F, lanes=2, geom=g)
    out = {"height_mm": torch.empty((F, g["h"], g["w"]), dtype=torch.float32, device="cuda"),
           "records": torch.empty((F, 104), dtype=torch.uint8, device="cuda")}
    for k in SF.MASK_NAMES:
        out[k] = torch.empty((F, g["h"], g["w"]), dtype=torch.uint8, device="cuda")

    def gather(rec):
        return SH.gather_records(rec, world * F, 0) if world > 1 else rec

    def step(frames, ref):
        gray = plan.begin(def_gray_dev=frames)
        plan.set_reference(ref, sync_info=False)
        for s in plan.lane_streams:
            s.wait_stream(torch.cuda.current_stream())
        return gather(plan.run(gray, masks=True, out=out)["records"])

    dev2, ref2 = torch.empty_like(dev), torch.empty_like(ref_dev)

    def step_e2e():
        ref2.copy_(ref_pinned, non_blocking=True)
        for c in range(0, F, nd):
            n = min(nd, F - c)
            dev2[c:c + n].copy_(pinned[:n], non_blocking=True)
        r = step(dev2, ref2)
        return r.cpu() if r is not None else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    W = max(a.warmup, 3)
    for _ in range(W):
        step(dev, ref_dev)
    stop, samples = threading.Event(), []
    th = threading.Thread(target=clocks_sampler, args=(stop, samples, world), daemon=True) if rank == 0 else None
    if th:
        th.start()
    l0 = N.launch_count()
    ms = timed(lambda: step(dev, ref_dev), a.steps)
    launches = N.launch_count() - l0
    stop.set()
    if th:
        th.join()
    step_e2e()
    e2e_steps = max(1, a.steps // 2)
    ms_e2e = timed(step_e2e, e2e_steps)
    rec = step(dev, ref_dev)
    if rank == 0:
        recs = SF.records_to_numpy(out["records"])
        fps, fps_e2e = world * F * a.steps / (ms / 1000.0), world * F * e2e_steps / (ms_e2e / 1000.0)
        pk_path = os.path.join(REPO, "MEASURED_PEAKS.json")
        peaks = json.load(open(pk_path)) if os.path.exists(pk_path) else {}
        peak = float(peaks.get("hbm_gbs", 6650.0))
        ach = B_ALG["ftp4k"] * (fps / world) / 1e9
        line = {"metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": a.steps, "warmup": W,
                "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": config, "clocks": summarize_clocks(samples),
                "e2e": {"value": fps_e2e, "unit": "frames/s", "h2d_bytes_per_step": int(dev.numel() + ref_dev.numel()),
                        "d2h_bytes_per_step": int(out["records"].numel())},
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                             "kernel": "whole pipeline (every kernel of the step)", "alg_bytes_per_frame": B_ALG["ftp4k"],
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy bandwidth)" if peaks else "fallback from B200_PROFILING.md"},
                "check": {"force_N_first": float(recs[0]["force_N"]), "status_or": int(np.bitwise_or.reduce(recs["status"])),
                          "residue_frames": int((recs["residues"] > 0).sum())}}
        if world == 1 and not a.no_cpu:
            arm = CpuArm("ftp4k", a.cpu_workers, a.cpu_frames)
            arm.seeds = seeds[:arm.frames]
            try:
                arm.pool.map(arm._prep, arm.seeds)
                dt = arm.step()
            finally:
                arm.close()
            line["cpu_baseline"] = {"value": len(arm.seeds) / dt, "unit": "frames/s", "cores": arm.workers, "kind": "port",
                                    "sample": arm.describe(dt, 1)}
            par = {"frames": 0, "max_rel_force": 0.0, "mask_px_diff": 0}
            for o in arm.readings:
                i = seeds.index(o["seed"])
                if "force_N" in o:
                    par["frames"] += 1
                    par["max_rel_force"] = max(par["max_rel_force"], abs(float(recs[i]["force_N"]) - o["force_N"]) / max(1e-9, abs(o["force_N"])))
                    par["mask_px_diff"] += abs(int(recs[i]["reliable_px"]) - o["reliable_px"]) + abs(int(recs[i]["contact_dilated_px"]) - o["contact_dilated_px"])
            line["parity"] = par
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
