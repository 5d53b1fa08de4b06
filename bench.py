"""bench.py — frames/s of the VISTAF sensing hot path on B200 (contract in the task statement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload temperature|ftp|multimodal]
                  [--frames F] [--impl reference]

A step = one pass of the hot path over one batch of F synthetic camera-resolution frames per GPU
(resident in HBM).  `value` = frames/s over all ranks with device-resident inputs (CUDA events, max
over ranks); `e2e` = the same through the public Python API from pinned HOST buffers with the H2D copy
of the frames and the D2H read of the per-frame records inside the timed region.
`--impl reference` times the reference's CPU algorithm (the oracle restatement — the reference itself is
a Python script that cannot travel to the GPU box) on the host cores.
"""
import argparse, json, os, subprocess, sys, threading, time

REPO = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(REPO, "vistaf-roboskin-vision-integrated-multimodal-sensor_b200")
sys.path.insert(0, PKG)

# algorithmic bytes per frame (fixed property of the work, not of the implementation)
B_ALG = {"temperature": 1000e6, "ftp": 760e6, "multimodal": 1760e6}
N_DISTINCT = 8          # distinct synthetic frames generated on the host, tiled to the batch on the device


def clocks_sampler(stop, samples):
    q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    while not stop.is_set():
        try:
            o = subprocess.run(["nvidia-smi", "-i", os.environ.get("LOCAL_RANK", "0"), f"--query-gpu={q}",
                                "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
            samples.append([x.strip() for x in o.strip().split(",")])
        except Exception:
            pass
        stop.wait(0.2)


def summarize_clocks(samples):
    sm = sorted(int(s[0]) for s in samples if s and s[0].isdigit())
    reasons = set()
    for s in samples:
        for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], s[2:6]):
            if v.lower().startswith("active"):
                reasons.add(name)
    mx = [int(s[1]) for s in samples if len(s) > 1 and s[1].isdigit()]
    return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons)}


def cpu_reference_run(workload, nframes, seeds):
    """Times the oracle (CPU restatement of the reference) on `nframes` synthetic frames."""
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import numpy as np
    import synthetic as S
    import ftp_oracle as FO, temperature_oracle as TO
    ref = S.make_frame(0, reference=True)
    frames = [S.make_frame(s) for s in seeds[:nframes]]
    P = FO.FtpParams(apply_global_shift=False, use_ecc=False)
    geom = FO.geometry(ref.shape[0], ref.shape[1], P)
    import cv2
    rg = cv2.cvtColor(ref[geom["y1"]:geom["y2"], geom["x1"]:geom["x2"]], cv2.COLOR_BGR2GRAY)
    refp = FO.reference_products(rg, geom, P)
    t = time.time()
    for f in frames:
        if workload in ("ftp", "multimodal"):
            dg = cv2.cvtColor(f[geom["y1"]:geom["y2"], geom["x1"]:geom["x2"]], cv2.COLOR_BGR2GRAY)
            r = FO.ftp_pair(rg, dg, geom, P, ref=refp)
            if r is not None:
                FO.force_readings(r)
        if workload in ("temperature", "multimodal"):
            TO.temperature_frame(f)
    dt = time.time() - t
    return nframes / dt, dt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=os.environ.get("VISTAF_WORKLOAD", "temperature"))
    ap.add_argument("--frames", type=int, default=64, help="frames per GPU per step")
    ap.add_argument("--chunk", type=int, default=16)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--cpu-frames", type=int, default=4)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    metric = "multimodal frames/sec (FTP+force+temp) at 1/2/4/8 B200; % HBM roofline"
    config = {"workload": f"{a.workload}: {a.frames} synthetic 2160x3840 BGR frames per GPU per step "
                          f"({N_DISTINCT} distinct seeds tiled), chunk {a.chunk}",
              "frames_per_gpu": a.frames, "resolution": "2160x3840", "l2": "inputs larger than L2 (no flush needed)"}

    if a.impl == "reference":
        if rank != 0:
            return
        fps, dt = cpu_reference_run(a.workload, a.cpu_frames, list(range(a.cpu_frames)))
        line = {"impl": "reference", "metric": metric, "value": fps, "unit": "frames/s", "n_gpus": a.gpus,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1000.0 * dt, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": 1, "kind": "port",
                                 "sample": f"{a.cpu_frames} frames of the same workload through oracle/ (numpy+OpenCV restatement; OpenCV/numpy internal threads as shipped)"},
                "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    import synthetic as S
    import temperature_sensor as TS
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    F = a.frames
    host = np.stack([S.make_frame(rank * 100003 + s) for s in range(N_DISTINCT)])
    pinned = torch.from_numpy(np.ascontiguousarray(np.tile(host, (F // N_DISTINCT + 1, 1, 1, 1))[:F])).pin_memory()
    dev = pinned.cuda()
    plan = TS.get_plan(host.shape[1], host.shape[2], chunk=a.chunk)
    out = plan.run(dev, masks=False)

    def step_resident():
        plan.run(dev, masks=False, out=out)

    dev2 = torch.empty_like(dev)

    def step_e2e():
        dev2.copy_(pinned, non_blocking=True)
        plan.run(dev2, masks=False, out=out)
        return out["records"].cpu()

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

    for _ in range(max(a.warmup, 3)):
        step_resident()
    stop, samples = threading.Event(), []
    th = threading.Thread(target=clocks_sampler, args=(stop, samples), daemon=True)
    th.start()
    ms = timed(step_resident, a.steps)
    stop.set(); th.join()
    step_e2e()
    ms_e2e = timed(step_e2e, max(1, a.steps // 2))
    e2e_steps = max(1, a.steps // 2)
    # per-frame scalar gather (the only collective of the path)
    if world > 1:
        rec = out["records"]
        gathered = [torch.empty_like(rec) for _ in range(world)] if rank == 0 else None
        dist.gather(rec, gathered, dst=0)
    if rank == 0:
        fps = world * F * a.steps / (ms / 1000.0)
        fps_e2e = world * F * e2e_steps / (ms_e2e / 1000.0)
        peaks = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(REPO, "MEASURED_PEAKS.json")) else {}
        peak = float(peaks.get("hbm_gbs", 6650.0))
        achieved = B_ALG[a.workload] * (fps / world) / 1e9
        cpu_fps, cpu_dt = cpu_reference_run(a.workload, a.cpu_frames, list(range(a.cpu_frames))) if world == 1 else (None, None)
        line = {"metric": metric, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
                "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": config, "clocks": summarize_clocks(samples),
                "e2e": {"value": fps_e2e, "unit": "frames/s", "h2d_bytes_per_step": int(dev.numel()),
                        "d2h_bytes_per_step": int(out["records"].numel())},
                "gpu_launches": None,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": None, "note": "whole-pipeline algorithmic bytes (SURVEY 8d) / step time, of measured" if peaks else "of fallback"}}
        if cpu_fps is not None:
            line["cpu_baseline"] = {"value": cpu_fps, "unit": "frames/s", "cores": 1, "kind": "port",
                                    "sample": f"{a.cpu_frames} frames through oracle/ in {cpu_dt:.1f} s"}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
