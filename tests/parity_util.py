"""Mask parity of the FTP chain off the pinned seeds (shared by the GPU tests).

Every FTP mask is a threshold on a float field followed by integer morphology: reliable = erode(largest component(close(
quality >= p25))), contact = |residual| >= p92, contact_kept = depth > 0.  The CUDA path and the reference (NumPy FFT,
OpenCV blurs, LAPACK lstsq) sum the same terms in different orders, so a pixel whose value sits within a few dozen float32
ulps of its threshold may fall on the other side -- and one such pixel moves the mask inside the reach of the morphology
that follows, and the height map inside the 200-px frontier taper around it (the reference's own algorithm is
discontinuous there).  The bar: masks identical EXCEPT pixels provably explained by such a tie, with the ties counted."""
import numpy as np
import cv2

ULPS = 32           # a tie = |value - threshold| <= ULPS float32 ulps of the threshold (observed field differences: <= 22 ulp)


def _near(mask, radius):
    k = 2 * int(radius) + 1
    return cv2.dilate(mask.astype(np.uint8), np.ones((k, k), np.uint8)) > 0


def check_ftp_masks(gpu_masks, hm_gpu, o, P, tol_mm=1e-3, max_ties=8):
    """gpu_masks: dict name -> bool array; o: oracle result with keep=True.  Returns a dict of tie statistics."""
    roi, rel_o = o["roi"], o["reliable"]
    thr = np.float32(o["amp_thr"])
    tie_q = roi & (np.abs(o["quality"] - thr) <= ULPS * np.spacing(thr))
    stats = {"quality_ties": int(tie_q.sum())}
    reach_rel = P.close_kernel // 2 + 1 + int(P.edge_margin) + 2
    d_rel = gpu_masks["reliable"] != rel_o
    stats["reliable_px"] = int(d_rel.sum())
    if d_rel.any():
        assert tie_q.any() and not (d_rel & ~_near(tie_q, reach_rel)).any(), "reliable differs away from any threshold tie"
        assert stats["quality_ties"] <= max_ties, f"{stats['quality_ties']} quality pixels within {ULPS} ulp of p25"
    d_or = gpu_masks["output_reliable"] != o["output_reliable"]
    assert not (d_or & ~d_rel).any(), "output_reliable differs where reliable does not"
    cthr = np.float32(o["contact_thr"])
    with np.errstate(invalid="ignore"):
        tie_c = rel_o & (np.abs(o["abs_res"] - cthr) <= ULPS * np.spacing(cthr))
    d_cd = gpu_masks["contact_dilated"] != o["contact_dilated"]
    stats["contact_dilated_px"], stats["contact_ties"] = int(d_cd.sum()), int(tie_c.sum())
    if d_cd.any():
        reach_c = (P.dilate_kernel // 2) * P.dilate_iters + 1
        ok = _near(tie_c, reach_c) | _near(d_rel, reach_c + 1)
        # a reliable tie also perturbs the fits slightly: then residual ties count with the mask's own reach
        assert not (d_cd & ~ok).any(), "contact_dilated differs away from any threshold tie"
        assert stats["contact_ties"] <= 4 * max_ties
    # contact_kept = sign of the depth inside kept blobs: differences only along the zero contour, or inside the frontier
    # neighbourhood of a reliable tie
    far = ~_near(d_rel, int(P.frontier_band) + 16) if d_rel.any() else np.ones_like(rel_o)
    z = np.maximum(np.abs(np.nan_to_num(hm_gpu)), np.abs(np.nan_to_num(o["height_mm"])))
    d_ck = gpu_masks["contact_kept"] != o["contact_kept"]
    stats["contact_kept_px"] = int(d_ck.sum())
    assert not (d_ck & far & (z > tol_mm)).any(), "contact_kept differs away from the zero contour"
    # height: within tolerance everywhere outside the frontier neighbourhood of reliable ties
    assert np.array_equal(np.isnan(hm_gpu), np.isnan(o["height_mm"]))
    d = np.abs(np.nan_to_num(hm_gpu) - np.nan_to_num(o["height_mm"]))
    stats["height_max_far"] = float(d[far].max()) if far.any() else 0.0
    stats["height_px_near_ties_over_tol"] = int((d[~far] > tol_mm).sum()) if (~far).any() else 0
    assert stats["height_max_far"] <= tol_mm, f"height differs by {stats['height_max_far']} mm away from any mask tie"
    return stats


def f14_case(h=1182, w=1182):
    """The F14 test input (tests/test_round2_gpu.py::test_internal_holes_are_found_and_filled, tools/make_golden_f14.py):
    a smooth height plane over a reliable disc with NaN islands -- four small interior ones (candidates), one large
    (known fraction too low) and one at the mask edge (too close).  Returns (height, reliable, islands)."""
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    rel = ((xx - 590) ** 2 + (yy - 600) ** 2) <= 400 ** 2
    hm = (0.3 * np.sin(xx / 37.0) * np.cos(yy / 51.0) - 0.2).astype(np.float32)
    hm[~rel] = np.nan
    holes = np.zeros_like(rel)
    for cy, cx, r in ((500, 500, 2), (620, 700, 3), (601, 595, 1), (700, 450, 3)):
        holes |= ((xx - cx) ** 2 + (yy - cy) ** 2) <= r * r
    holes |= ((xx - 800) ** 2 + (yy - 600) ** 2) <= 30 ** 2
    holes |= ((xx - 590) ** 2 + (yy - 203) ** 2) <= 3 ** 2
    hm[holes & rel] = np.nan
    return hm, rel, holes


def residue_case(n=320, levels=0, seed=3):
    """A path-dependent unwrap input (tools/make_golden_unwrap.py, tests/test_oracle_golden.py): a smooth phase of
    amplitude ~3 pi wrapped into (-pi, pi], two opposite phase singularities inside a disc mask with a hole, and a rugged
    quality map -- quantised to `levels` values when levels > 0, so that the flood order is decided by the heap's
    tie-breaking (-q, y, x, py, px) almost everywhere.  Returns (wrapped float32, mask bool, quality float32)."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:n, 0:n].astype(np.float32)
    c = (n - 1) / 2.0
    phase = 9.0 * np.exp(-(((xx - 0.45 * n) ** 2 + (yy - 0.55 * n) ** 2) / (2 * (0.18 * n) ** 2))) + 0.01 * xx
    vort = np.arctan2(yy - (0.40 * n + 0.5), xx - (0.35 * n + 0.5)) - np.arctan2(yy - (0.62 * n + 0.5), xx - (0.70 * n + 0.5))
    wrapped = np.angle(np.exp(1j * (phase.astype(np.float64) + vort))).astype(np.float32)
    mask = ((xx - c) ** 2 + (yy - c) ** 2) <= (0.47 * n) ** 2
    mask &= ~(((xx - 0.6 * n) ** 2 + (yy - 0.3 * n) ** 2) <= (0.05 * n) ** 2)                   # a hole: loops around it
    q = np.exp(-(((xx - c) ** 2 + (yy - c) ** 2) / (2 * (0.4 * n) ** 2))).astype(np.float32)
    q = (q * (1.0 + 0.2 * rng.standard_normal(q.shape))).astype(np.float32)
    if levels > 0:
        q = (np.round(q * levels) / levels).astype(np.float32)
    return wrapped, mask, q
