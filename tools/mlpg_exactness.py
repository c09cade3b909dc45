"""Diagnostic (GPU): how far the MLPG output is from the oracle, per stream; exact equality expected
without GV (same operations in the same order), <= 1e-9 relative with GV (block reductions)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import jbonsai_b200 as jb
from jbonsai_b200 import workload as wl
from oracle_lib import Oracle
from labels import SAMPLE_SENTENCE_1, SAMPLE_SENTENCE_2

for name in ("nitech", "tohoku-f01-neutral"):
    e, o = jb.Engine.load([wl.VOICES[name]]), Oracle([wl.VOICES[name]])
    for gvw in (1.0, 0.0):
        for s in range(2):
            e.set("gv_weight", gvw, s); o.set("gv_weight", gvw, s)
        for lab in (SAMPLE_SENTENCE_1, SAMPLE_SENTENCE_2):
            dbg = e.debug_batch([lab])
            mcp, lf0, lpf = o.parameters(lab)
            out = []
            for k, ref in (("mcp", mcp), ("lf0", lf0), ("lpf", lpf)):
                got = dbg[k].reshape(ref.shape)
                d = np.abs(got - ref)
                out.append(f"{k}: equal={np.array_equal(got, ref)} maxrel={float((d / np.maximum(np.abs(ref), 1e-300)).max()):.2e}")
            print(name, "gv_weight", gvw, len(lab), " | ".join(out))
