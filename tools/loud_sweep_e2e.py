"""Cuts to repair per synthesize_batch call in the default delivery mode (three sub-batches, ~21 k segments of ~49
own frames), warm-up by loudness on and off, eight batches x two voices. Usage: python tools/loud_sweep_e2e.py"""
import os, sys, time, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import jbonsai_b200 as jb
from jbonsai_b200 import workload as wl, _flatten

CASES = [(40, 0, 2.0), (40, 24, 2.0), (40, 24, 3.0), (40, 28, 2.0), (48, 24, 2.0), (48, 28, 2.0)]
if len(sys.argv) > 1:
    CASES = [eval(a) for a in sys.argv[1:]]
tot = {c: [0, 0, []] for c in CASES}
for voice, seed in (("nitech", wl.SEED_NITECH), ("tohoku-f01-neutral", wl.SEED_TOHOKU)):
    for k in range(8):
        e = jb.Engine.load([wl.VOICES[voice]])
        flat = _flatten(wl.timed_batch(e, 1024, seed, first=k * 1024))
        row = []
        for c in CASES:
            warm, loud, thr = c
            e.set_mlsa_segmenting(warm, 5e-7)
            e.set_mlsa_loud_warm(loud, thr)
            ts = []
            for it in range(4):
                t0 = time.perf_counter()
                out, _, tm = e.synthesize_batch_raw(flat); jb.lib.jb_free(out)
                ts.append((time.perf_counter() - t0) * 1e3)
            tot[c][0] += tm["mlsa_repaired"]; tot[c][1] += tm["mlsa_fallback"]; tot[c][2].append(min(ts[1:]))
            row.append(f"{c}: rep {tm['mlsa_repaired']:.0f} fb {tm['mlsa_fallback']:.0f} worst {tm['mlsa_cut_worst']:.1e} e2e {min(ts[1:]):.1f}")
        print(voice, k, " | ".join(row), flush=True)
for c in CASES:
    print(c, "repaired", tot[c][0], "fallbacks", tot[c][1], "e2e mean of min", round(statistics.mean(tot[c][2]), 2))
