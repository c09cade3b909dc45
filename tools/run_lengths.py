"""Voiced-run statistics of the bench batch (what bounds the pulse walks): python tools/run_lengths.py [n_utt]"""
import sys
import numpy as np
sys.path.insert(0, "tests")
sys.path.insert(0, ".")
import jbonsai_b200 as jb
from jbonsai_b200 import workload as wl
from conftest import NITECH

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
e = jb.Engine.load([NITECH])
utts = wl.timed_batch(e, n, wl.SEED_NITECH)
dbg = e.debug_batch(utts)
fo, lf0 = dbg["frame_off"], dbg["lf0"]
runs = []
for u in range(n):
    v = lf0[fo[u]:fo[u + 1]] != -1e10
    k = 0
    while k < len(v):
        if v[k]:
            j = k
            while j < len(v) and v[j]:
                j += 1
            runs.append(j - k)
            k = j
        else:
            k += 1
runs = np.array(runs)
print("utterances", n, "frames", int(fo[n]), "voiced", int(runs.sum()), "runs", len(runs), "mean", runs.mean(),
      "p50/p90/p99/max", np.percentile(runs, [50, 90, 99]).tolist(), int(runs.max()))
per = np.exp(-lf0[lf0 != -1e10]) * 48000
print("pitch period samples p5/p50/p95", np.percentile(per, [5, 50, 95]).tolist())
