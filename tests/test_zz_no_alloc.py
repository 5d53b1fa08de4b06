"""vistaf_depth_to_volume_static: the allocation-free variant of vistaf_depth_to_volume (static __device__ scratch).

No counterpart in the reference.  Called through the C ABI and compared
with the existing entry point on the same device planes, twice in a row (the scratch array keeps the previous call's
sums and must be cleared by the call itself), with the tolerances tests/test_ftp_gpu.py uses for this op: volume 1e-6
relative (double atomics in launch order), area and max depth exact.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import _native as N
    return torch, N


def _planes(torch, seed, batch, h, w):
    rng = np.random.default_rng(seed)
    z = np.abs(rng.normal(0, 0.5, (batch, h, w))).astype(np.float32)
    z[1::2] *= -1.0                                   # odd frames: the indentation is the NEGATIVE side (sign choice)
    z[rng.random(z.shape) < 0.5] *= -0.01
    z[:, :7] = np.nan
    if batch > 2:
        z[2] = 0.0                                    # a frame with nothing above depth_eps -> (0, 0, 0)
    roi = (rng.random((h, w)) < 0.8).astype(np.uint8)
    return torch.from_numpy(z).cuda(), torch.from_numpy(roi).cuda()


def _call(torch, N, fn, z, roi, out):
    b, h, w = z.shape
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    N.check(fn(z.data_ptr(), roi.data_ptr() if roi is not None else None, b, h * w, 0.03, 0.01, out.data_ptr(), st), fn.__name__)


def test_static_scratch_variant_equals_the_allocating_entry_point_twice_in_a_row(env):
    torch, N = env
    L = N.lib()
    assert L.vistaf_depth_to_volume_static_max_batch() == 4096
    # outputs exist before the no-alloc calls; the variant itself must not change the device's free memory
    cases = [(_planes(torch, 11, 5, 300, 257), True), (_planes(torch, 12, 3, 301, 129), False)]
    outs = [(torch.full((z.shape[0], 3), -1.0, dtype=torch.float64, device="cuda"),
             torch.full((z.shape[0], 3), -2.0, dtype=torch.float64, device="cuda")) for (z, _), _ in cases]
    for ((z, roi), use_roi), (want, _) in zip(cases, outs):
        _call(torch, N, L.vistaf_depth_to_volume, z, roi if use_roi else None, want)
    # once-per-device set-up (symbol address; lazy module loading may take device memory on a first launch)
    _call(torch, N, L.vistaf_depth_to_volume_static, cases[1][0][0], None, outs[1][1])
    outs[1][1].fill_(-2.0)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    # second call: smaller batch, different data -- stale sums of the first call would show in frames 0..2
    for ((z, roi), use_roi), (_, got) in zip(cases, outs):
        _call(torch, N, L.vistaf_depth_to_volume_static, z, roi if use_roi else None, got)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the static variant changed the device's free memory"
    for (want, got) in outs:
        w_, g_ = want.cpu().numpy(), got.cpu().numpy()
        assert np.isfinite(g_).all()
        np.testing.assert_allclose(g_[:, 0], w_[:, 0], rtol=1e-6, atol=0)
        np.testing.assert_array_equal(g_[:, 1], w_[:, 1])
        np.testing.assert_array_equal(g_[:, 2], w_[:, 2])
    assert outs[0][1][2].tolist() == [0.0, 0.0, 0.0] and float(outs[0][1][0, 0]) > 0
    # a third call on the first case again: same answer as before
    again = torch.empty_like(outs[0][1])
    (z, roi), _ = cases[0]
    _call(torch, N, L.vistaf_depth_to_volume_static, z, roi, again)
    torch.cuda.synchronize()
    np.testing.assert_allclose(again.cpu().numpy(), outs[0][0].cpu().numpy(), rtol=1e-6, atol=0)


def test_static_scratch_variant_rejects_batches_beyond_its_maximum(env):
    torch, N = env
    L = N.lib()
    z = torch.zeros((1, 4, 4), dtype=torch.float32, device="cuda")
    out = torch.zeros((1, 3), dtype=torch.float64, device="cuda")
    rc = L.vistaf_depth_to_volume_static(z.data_ptr(), None, 4097, 16, 0.03, 0.01, out.data_ptr(), None)
    assert rc != 0
    msg = L.vistaf_last_error().decode()
    assert "4097" in msg and "4096" in msg and "static scratch" in msg
    with pytest.raises(RuntimeError):
        N.check(L.vistaf_depth_to_volume_static(None, None, 1, 16, 0.03, 0.01, out.data_ptr(), None), "null input")
