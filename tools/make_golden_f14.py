"""Build-container-only: golden vector of F14 (internal holes inside the reliable mask + Telea r = 5) from the UNMODIFIED
reference's own functions (Code/shape_ftp.py: compute_internal_holes_within_mask :1153-1175, inpaint_only_mask :1178-1203,
composed exactly as main() does at :1770-1801 with the module's constants) on the injected-island input of
tests/parity_util.f14_case.  F14 never fires on the reference's real frames, so this is the only way to
pin the oracle's restatement of it to the reference.  Writes tests/golden/f14_reference.npz."""
import hashlib, os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, "..", "tests"))
G = os.path.join(HERE, "..", "tests", "golden")
from parity_util import f14_case
from ref_harness import import_reference

if __name__ == "__main__":
    R, _, _ = import_reference()
    height_map, reliable, _ = f14_case()
    h, w = height_map.shape
    # ---- shape_ftp.py:1770-1801, verbatim composition of the reference's functions
    known_height = reliable & np.isfinite(height_map)
    height_rel_filled = np.full((h, w), np.nan, np.float32)
    height_rel_filled[known_height] = height_map[known_height]
    assert R.FILL_INTERNAL_HOLES_IN_RELIABLE
    hole_candidates = R.compute_internal_holes_within_mask(container_mask=reliable, known_mask=known_height,
                                                           ksize=R.HOLE_NEIGHBORHOOD_PX, frac_thr=R.HOLE_KNOWN_FRACTION,
                                                           min_dist_edge_px=R.HOLE_MIN_DIST_FROM_RELIABLE_EDGE_PX)
    assert np.any(hole_candidates)
    tmp = height_rel_filled.copy()
    med = float(np.nanmedian(tmp[known_height])) if np.any(known_height) else 0.0
    tmp[reliable & ~known_height] = med
    filled = R.inpaint_only_mask(z_known=tmp, roi_mask=reliable, inpaint_mask=hole_candidates, radius=R.INPAINT_RADIUS,
                                 method=R.INPAINT_METHOD)
    height_rel_filled[hole_candidates] = filled[hole_candidates]
    output_reliable = reliable & np.isfinite(height_rel_filled)
    # ----
    ys, xs = np.nonzero(hole_candidates)
    np.savez_compressed(os.path.join(G, "f14_reference.npz"),
                        candidates_packed=np.packbits(hole_candidates), output_reliable_packed=np.packbits(output_reliable),
                        cand_y=ys.astype(np.int32), cand_x=xs.astype(np.int32), cand_value=height_rel_filled[ys, xs].astype(np.float32),
                        shape=np.array([h, w], np.int32),
                        constants=np.array([R.HOLE_NEIGHBORHOOD_PX, R.HOLE_KNOWN_FRACTION, R.HOLE_MIN_DIST_FROM_RELIABLE_EDGE_PX,
                                            R.INPAINT_RADIUS], np.float64))
    print("candidates", int(hole_candidates.sum()), "output_reliable", int(output_reliable.sum()), "reliable", int(reliable.sum()),
          "sha", hashlib.sha1(np.packbits(hole_candidates).tobytes()).hexdigest()[:12])
