"""Host-side logic of the multi-GPU path on CPU: contiguous frame shards (ragged included) and the gather of the
fixed-size per-frame records to rank 0 under torch.distributed (gloo, world_size 2 and 3); plus the
multimodal_summary.json schema against the reference's own stored summary."""
import json
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

G = os.path.join(os.path.dirname(__file__), "golden")


def test_shard_ranges_cover_and_partition():
    import sharding as SH
    for n in (0, 1, 7, 8, 4096, 4097):
        for world in (1, 2, 3, 4, 8):
            r = [SH.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1 and sizes == SH.shard_sizes(n, world)
    with pytest.raises(ValueError):
        SH.shard_range(4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_frames, q):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.join(here, "..", "vistaf-roboskin-vision-integrated-multimodal-sensor_b200"))
    import sharding as SH
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = SH.shard_range(n_frames, rank, world)
        # a record whose bytes encode its global frame index
        local = torch.zeros((hi - lo, 184), dtype=torch.uint8)
        for i in range(lo, hi):
            local[i - lo] = torch.from_numpy(np.frombuffer(np.full(23, i, np.int64).tobytes(), np.uint8).copy())
        out = SH.gather_records(local, n_frames, 0)
        if rank == 0:
            idx = out.numpy().view(np.int64).reshape(n_frames, 23)
            q.put(bool((idx == np.arange(n_frames)[:, None]).all()))
        else:
            assert out is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_frames", [(2, 8), (2, 7), (3, 5)])
def test_gather_records_gloo(world, n_frames):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=10) is True


def test_gather_single_process_passthrough():
    import sharding as SH
    a = torch.arange(3 * 184, dtype=torch.int64).to(torch.uint8).reshape(3, 184)
    assert SH.gather_records(a, 3) is a
    with pytest.raises(ValueError):
        SH.gather_records(a, 4)


def test_summary_schema_matches_reference_json():
    """Key-for-key equality of the summary document with the reference's stored multimodal_summary.json, and
    equality of the calibration_performance block (same numbers, shipped with the package)."""
    import multimodal_sensor as MS
    ref = json.load(open(os.path.join(G, "FINAL_TEMP_DEMO_multimodal_summary.json")))
    force = dict(ref["sensor_readings"]["force"])
    temp = dict(ref["sensor_readings"]["temperature"])
    s = MS.build_summary(force, temp, "a.jpg", "b.jpg", "sess", "f", "t", "c", "id")

    def keys(d, prefix=""):
        out = set()
        for k, v in d.items():
            out.add(prefix + k)
            if isinstance(v, dict):
                out |= keys(v, prefix + k + ".")
        return out
    assert keys(s) == keys(ref)
    assert s["calibration_performance"] == ref["calibration_performance"]
    assert s["sensor_readings"] == ref["sensor_readings"]
    assert MS.RECORD_BYTES == 184


def test_records_roundtrip_to_summaries():
    import multimodal_sensor as MS
    rec = np.zeros(2, MS.RECORD_DTYPE)
    rec["ftp"]["force_N"] = [1.5, 2.5]
    rec["temp"]["mean_C"] = [21.0, 30.0]
    rec["temp"]["valid_pixels"] = [10, 20]
    back = MS.records_to_numpy(torch.from_numpy(rec.view(np.uint8).reshape(2, -1)))
    out = MS.summaries_from_records(back, ["a", "b"])
    assert out[1]["frame"] == "b" and out[1]["sensor_readings"]["force"]["force_N"] == 2.5
    assert out[0]["sensor_readings"]["temperature"]["valid_pixels"] == 10
