"""Build-container-only: golden wrap-count fields of the UNMODIFIED reference's quality-guided flood
(Code/shape_ftp.py: unwrap_quality_guided :1043-1080, Python heapq) on path-dependent inputs (phase residues, a hole in the
mask, quality ties): tests/parity_util.residue_case.  With residues the result depends on the exact pop order of the
heap, i.e. on the tuple ordering (-q, y, x, py, px); the oracle's C restatement (oracle/native/unwrap_heap.c) and the
device flood (k_ftp_unwrap_heap) claim to reproduce it.  Writes tests/golden/unwrap_reference.npz."""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, "..", "tests"))
G = os.path.join(HERE, "..", "tests", "golden")
from parity_util import residue_case
from ref_harness import import_reference

CASES = {"rugged": dict(n=320, levels=0, seed=3), "ties64": dict(n=320, levels=64, seed=4), "ties8": dict(n=256, levels=8, seed=5)}

if __name__ == "__main__":
    R, _, _ = import_reference()
    out = {}
    for name, kw in CASES.items():
        wrapped, mask, q = residue_case(**kw)
        u = R.unwrap_quality_guided(wrapped, mask, q)
        assert np.array_equal(np.isfinite(u), mask), "the flood reaches every masked pixel of this input"
        k = np.zeros(wrapped.shape, np.int8)
        k[mask] = np.rint((u[mask].astype(np.float64) - wrapped[mask]) / (2 * np.pi)).astype(np.int8)
        out[name + "__k"] = k
        out[name + "__u_sub4"] = u[::4, ::4].astype(np.float32)
        print(name, kw, "k range", int(k.min()), int(k.max()), "nonzero", int((k != 0).sum()), "of", int(mask.sum()),
              "distinct q", int(np.unique(q[mask]).size), flush=True)
    np.savez_compressed(os.path.join(G, "unwrap_reference.npz"), **out)
