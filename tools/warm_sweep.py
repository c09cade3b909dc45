"""Warm-up length of the MLSA time segments against repairs and time, on the bench batch.
Usage: python tools/warm_sweep.py [n_utt]"""
import os, sys, time, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import jbonsai_b200 as jb
from jbonsai_b200 import workload as wl, _flatten

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
for voice, seed in (("nitech", wl.SEED_NITECH), ("tohoku-f01-neutral", wl.SEED_TOHOKU)):
    e = jb.Engine.load([wl.VOICES[voice]])
    utts = wl.timed_batch(e, n, seed)
    flat = _flatten(utts)
    for warm, loud in ((40, 0), (40, 28), (40, 24), (40, 20), (40, 16), (44, 24), (36, 0), (32, 0), (24, 0)):
        e.set_mlsa_segmenting(warm, 5e-7)
        e.set_mlsa_loud_warm(loud, 1.0)
        for chunks in (1, 3):
            e.set_pipeline_chunks(chunks)
            e.set_output_mode(0 if chunks == 1 else 1)
            ts, rep, worst, mlsa = [], 0, 0.0, 0.0
            for it in range(6):
                t0 = time.perf_counter()
                out, _, tm = e.synthesize_batch_raw(flat); jb.lib.jb_free(out)
                if it >= 2:
                    ts.append((time.perf_counter() - t0) * 1e3)
                rep, worst, mlsa, fb = tm["mlsa_repaired"], tm["mlsa_cut_worst"], tm["mlsa_ms"], tm["mlsa_fallback"]
            print(f"{voice} warm={warm} loud={loud} chunks={chunks}: e2e p50 {statistics.median(ts):.1f} ms mlsa {mlsa:.2f} ms "
                  f"repaired {rep:.0f} fallback {fb:.0f} worst bound {worst:.2e} segments {tm['mlsa_segments']:.0f} filtered frames {tm['mlsa_seg_frames']:.0f}", flush=True)
