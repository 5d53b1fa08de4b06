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
`--impl reference` times the reference's CPU algorithm (oracle/, the numpy+OpenCV restatement pinned against
the unmodified reference; the reference itself is a Python script under /root/reference that cannot travel
to the GPU box) on the host cores with a process pool.
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
B_ALG = {"temperature": 1000e6, "ftp": 760e6, "multimodal": 1760e6}
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
N_DISTINCT = 8          # distinct synthetic frames generated on the host per rank, tiled to the batch on the device


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
    seed, workload = args
    _cpu_prepare(seed)
    FO, TO, cv2, g = _W["FO"], _W["TO"], _W["cv2"], _W["geom"]
    f = _W["frames"][seed]
    t = time.perf_counter()
    if workload in ("ftp", "multimodal"):
        dg = cv2.cvtColor(f[g["y1"]:g["y2"], g["x1"]:g["x2"]], cv2.COLOR_BGR2GRAY)
        r = FO.ftp_pair(_W["rg"], dg, g, _W["P"], ref=_W["refp"])
        if r is not None:
            FO.force_readings(r)
    if workload in ("temperature", "multimodal"):
        TO.temperature_frame(f)
    return time.perf_counter() - t


class CpuArm:
    """Process pool over the host cores; each step pushes `frames_per_step` synthetic frames through the oracle."""

    def __init__(self, workload, workers=None, frames_per_step=None):
        import multiprocessing as mp
        avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        self.workers = int(workers or max(1, min(avail, 32)))
        self.frames = int(frames_per_step or self.workers)
        self.workload = workload
        self.pool = mp.get_context("spawn").Pool(self.workers, initializer=_cpu_init)
        self.seeds = list(range(1, self.frames + 1))
        self.pool.map(_cpu_prepare, self.seeds, chunksize=max(1, self.frames // self.workers))     # untimed: generation + reference products

    def step(self):
        t = time.perf_counter()
        self.pool.map(_cpu_process, [(s, self.workload) for s in self.seeds], chunksize=max(1, self.frames // self.workers))
        return time.perf_counter() - t

    def close(self):
        self.pool.close()
        self.pool.join()

    def describe(self, seconds, steps):
        return (f"{steps} step(s) x {self.frames} synthetic 2160x3840 frames ({self.workload}) through oracle/ "
                f"(numpy+OpenCV restatement of the reference, C heap unwrap) on a pool of {self.workers} processes, "
                f"{seconds:.1f} s of wall time; frame generation and reference-frame products excluded")


def run_reference(a, config):
    arm = CpuArm(a.workload, a.cpu_workers, a.cpu_frames)
    try:
        for _ in range(min(a.warmup, 1)):                  # one warm pass is enough for a CPU pool (page-in, imports)
            arm.step()
        steps = max(1, min(a.steps, a.cpu_steps))
        dt = sum(arm.step() for _ in range(steps))
    finally:
        arm.close()
    fps = arm.frames * steps / dt
    return {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": a.gpus, "steps": steps,
            "warmup": min(a.warmup, 1), "ms_per_step": 1000.0 * dt / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": arm.workers, "kind": "port",
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
                    choices=["multimodal", "ftp", "temperature"])
    ap.add_argument("--frames", type=int, default=256, help="frames per GPU per step")
    ap.add_argument("--chunk", type=int, default=256, help="FTP frames per library chunk (workspace size)")
    ap.add_argument("--chunk-temp", type=int, default=64, help="temperature frames per library chunk")
    ap.add_argument("--lanes-ftp", type=int, default=4, help="concurrent slices (streams + workspaces) of the FTP batch")
    ap.add_argument("--lanes-temp", type=int, default=2, help="concurrent slices of the temperature batch")
    ap.add_argument("--distinct", type=int, default=8,
                    help="distinct synthetic seeds per rank (16 brings in one glare frame per 16: the T10 path)")
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--cpu-workers", type=int, default=None)
    ap.add_argument("--cpu-frames", type=int, default=None, help="frames per CPU step (default: one per worker)")
    ap.add_argument("--cpu-steps", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    a = ap.parse_args()
    globals()["N_DISTINCT"] = max(1, a.distinct)
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    config = {"workload": f"{a.workload}: {a.frames} synthetic 2160x3840 BGR fringe+TLC frames per GPU per step "
                          f"({N_DISTINCT} distinct seeds per rank tiled to the batch) against one reference frame, "
                          f"FTP crop 1182^2 / FFT grid 1374^2, both temperature models; frames pre-registered "
                          f"(APPLY_GLOBAL_SHIFT=False, USE_ECC_CROP_ALIGNMENT=False on GPU and CPU arms)",
              "frames_per_gpu": a.frames, "chunk_ftp": a.chunk, "chunk_temp": a.chunk_temp, "lanes_ftp": a.lanes_ftp,
              "lanes_temp": a.lanes_temp, "resolution": "2160x3840",
              "l2": f"inputs {a.frames * 24.9:.0f} MB per step, larger than the 126 MB L2 (no flush needed)"}

    if a.impl == "reference":
        if rank != 0:
            return
        emit(run_reference(a, config))
        return

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
    host = np.stack([S.make_frame(rank * 100003 + s + 1) for s in range(N_DISTINCT)])
    # pinned host staging buffer: the N_DISTINCT frames once (keeps host RAM at 200 MB per rank); a step's batch is
    # frame i = distinct frame i % N_DISTINCT, both for the resident copy and for every host->device copy of the e2e leg
    pinned = torch.from_numpy(np.ascontiguousarray(host)).pin_memory()
    ref_pinned = torch.from_numpy(ref_host).pin_memory()
    dev = pinned.cuda().repeat((F + N_DISTINCT - 1) // N_DISTINCT, 1, 1, 1)[:F].contiguous()
    ref_dev = ref_pinned.cuda()
    plan = MS.MultimodalPlan(host.shape[1], host.shape[2], min(a.chunk, F), min(a.chunk_temp, F), a.workload,
                             lanes_ftp=a.lanes_ftp, lanes_temp=a.lanes_temp)

    def step_resident():
        return plan.step(ref_dev, dev, masks=False)

    dev2, ref2 = torch.empty_like(dev), torch.empty_like(ref_dev)

    copy_stream = torch.cuda.Stream()
    n_slices = 2 * max(1, a.lanes_ftp, a.lanes_temp)

    def step_e2e():
        """Host -> device staging of the reference and of the frames in slices on a copy stream (events mark each slice
        ready, so a lane starts as soon as its frames have landed), the step through the public API, records to host."""
        main = torch.cuda.current_stream()
        copy_stream.wait_stream(main)             # the previous step has finished reading dev2 / ref2
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
        return plan.step(ref2, dev2, masks=False, ready=ready, ref_ready=ev_ref)["records"].cpu()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_ms = [0.0]

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
        host_ms[0] = (time.perf_counter() - t0) * 1e3 / steps      # host time to ENQUEUE one step (no sync inside; includes launch-queue back-pressure)
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
    # the only collective of the path: gather of the per-frame scalar records to rank 0
    rec = plan.run(dev, masks=False)["records"]
    import sharding as SH
    gathered = SH.gather_records(rec, world * F, 0) if world > 1 else rec
    if rank == 0:
        assert gathered.shape[0] == world * F
        recs = MS.records_to_numpy(rec)
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
                        "h2d_bytes_per_step": int(dev.numel() + ref_dev.numel()), "d2h_bytes_per_step": int(rec.numel())},
                "gpu_launches": int(launches), "ms_per_step_serialised": ms_serial / a.steps,
                "host_enqueue_ms_per_step": host_enqueue_ms}
        table_f = [r for r in table if r["GBps"]]
        top = max(table_f, key=lambda r: r["share"]) if table_f else None
        tr_path = os.path.join(REPO, "profiles", "ncu_stage_traffic.json")
        tr = json.load(open(tr_path)) if os.path.exists(tr_path) else {}
        if top and top["GBps"]:
            t = tr.get(top["stage"])
            line["roofline"] = {"bound": "hbm", "achieved": top["GBps"], "peak": peak, "unit": "GB/s",
                                "frac": top["GBps"] / peak, "traffic": t["dram_bytes_per_launch"] if t else None,
                                "traffic_note": (f"ncu dram read+write of {t['kernel']} per launch of {t['frames_per_launch']} "
                                                 f"frames ({t['source']})") if t else None,
                                "kernel": top["stage"],
                                "share_of_step": top["share"], "peak_source": peak_src,
                                "note": "dominant stage of the step: algorithmic MB/frame of that stage (DESIGN.md stage table) "
                                        "x frames / stage time from CUDA events on the launch stream, taken in a second pass of "
                                        "the same steps with the concurrent lanes serialised on one stream (in the timed region "
                                        "the lanes overlap, so per-stage times are not separable there)"}
        ach = B_ALG[a.workload] * (fps / world) / 1e9
        line["roofline_pipeline"] = {"achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                                     "frac_of_8TBps_nominal": ach / 8000.0, "alg_bytes_per_frame": B_ALG[a.workload]}
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
                                    "sample": arm.describe(dt, 1)}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
