"""Is the pitch period the device computes, rate / exp(lf0) with CUDA's exp, the one the host libm gives?

The north star lists pulse positions under "bit-exact". A pulse falls where pitch_counter >= pitch_of_curr_point
(excitation.rs:72-86); the pitch period comes from exp() (vocoder/mod.rs:73-77), which CUDA rounds to <= 1 ulp and
glibc to < 1 ulp: where the two differ the period differs by one ulp. This tool counts it (GPU box):
  1. a dense grid over the whole clamp range [MIN_LF0, MAX_LF0] (every lf0 value a voice can emit is in it);
  2. the lf0 track of 256 utterances of the bench batch (MLPG + GV output, the values that actually occur);
  3. for every frame of (2) whose period differs: does the pulse train of its utterance differ? (oracle vs device
     pulse positions are compared directly by bench.py's parity block: pulse_position_mismatches.)
Usage: python tools/exp_sweep.py > profiles/r2_exp_sweep.json"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import jbonsai_b200 as jb  # noqa: E402
from jbonsai_b200 import workload as wl  # noqa: E402
import oracle_lib  # noqa: E402

MIN_LF0, MAX_LF0 = 2.995732273553991, 9.903487552536127


def ulps(a, b):
    return np.abs(a.view(np.int64) - b.view(np.int64))


def compare(x, name):
    dev = jb.device_pitch_period(x)
    host = oracle_lib.pitch_period(x)
    u = ulps(dev, host)
    return {"set": name, "values": int(x.size), "period_mismatches": int(np.count_nonzero(u)),
            "max_ulp": int(u.max(initial=0)), "mismatch_fraction": float(np.count_nonzero(u) / max(x.size, 1))}, u


out = {"what": "rate / exp(lf0): device (CUDA exp, k_frame_prep) vs host libm (the oracle's std::exp)", "sets": []}
grid = np.linspace(MIN_LF0, MAX_LF0, 1 << 24)
r, _ = compare(grid, "dense grid, 2^24 points over [MIN_LF0, MAX_LF0]")
out["sets"].append(r)
rng = np.random.default_rng(1)
r, _ = compare(rng.uniform(MIN_LF0, MAX_LF0, 1 << 24), "2^24 uniform random points")
out["sets"].append(r)

e = jb.Engine.load([wl.VOICES["nitech"]])
utts = wl.timed_batch(e, 256, wl.SEED_NITECH)
dbg = e.debug_batch(utts)
lf0 = dbg["lf0"]
voiced = lf0 != -1e10
r, u = compare(lf0[voiced], "voiced frames of 256 bench utterances (MLPG + GV output)")
out["sets"].append(r)
# frames whose period differs: do the device pulse positions of their utterances differ from the oracle's?
bad_frames = np.flatnonzero(voiced)[np.flatnonzero(u)]
fo = dbg["frame_off"]
bad_utts = sorted(set(int(np.searchsorted(fo, f, side="right") - 1) for f in bad_frames))
fp = int(e.get("fperiod"))
o = oracle_lib.Oracle([wl.VOICES["nitech"]])
pos_mismatch = 0
for uu in bad_utts:
    tap = o.synthesize_tap(utts[uu], fma=True)
    g = dbg["pulse"][fo[uu] * fp: fo[uu + 1] * fp]
    pos_mismatch += int(np.count_nonzero((g != 0.0) != (tap["pulse"] != 0.0)))
out["utterances_with_a_differing_period"] = len(bad_utts)
out["pulse_position_mismatches_in_those_utterances"] = pos_mismatch
print(json.dumps(out, indent=1))
