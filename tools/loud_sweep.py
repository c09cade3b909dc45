"""Loudness-sized warm-up (Engine.set_mlsa_loud_warm) on the eight 1024-utterance batches of the scaling run, both
voices: cuts to repair, worst cut bound, filter time. Usage: python tools/loud_sweep.py [batches]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import jbonsai_b200 as jb
from jbonsai_b200 import workload as wl, _flatten

nb = int(sys.argv[1]) if len(sys.argv) > 1 else 8
CASES = [(40, 0, 1.0), (40, 24, 1.5), (40, 24, 2.0), (40, 24, 2.5), (40, 24, 3.0), (40, 20, 2.0), (40, 20, 3.0)]
if len(sys.argv) > 2:
    CASES = [eval(a) for a in sys.argv[2:]]
for voice, seed in (("nitech", wl.SEED_NITECH), ("tohoku-f01-neutral", wl.SEED_TOHOKU)):
    for k in range(nb):
        e = jb.Engine.load([wl.VOICES[voice]])
        e.set_pipeline_chunks(1)
        e.set_output_mode(0)
        flat = _flatten(wl.timed_batch(e, 1024, seed, first=k * 1024))
        row = []
        for warm, loud, thr in CASES:
            e.set_mlsa_segmenting(warm, 5e-7)
            e.set_mlsa_loud_warm(loud, thr)
            for it in range(2):
                out, _, tm = e.synthesize_batch_raw(flat); jb.lib.jb_free(out)
            row.append(f"({warm},{loud},{thr:g}): rep {tm['mlsa_repaired']:.0f} fb {tm['mlsa_fallback']:.0f} worst {tm['mlsa_cut_worst']:.1e} "
                       f"filt {tm['mlsa_seg_frames'] / 1e6:.3f}M mlsa {tm['mlsa_ms']:.2f}")
        print(voice, "batch", k, " | ".join(row), flush=True)
