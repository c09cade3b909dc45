"""e2e timing breakdown of jb_engine_synthesize_batch for the bench workload under different
delivery settings (A/B tool). Usage: python tools/e2e_probe.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import jbonsai_b200 as jb
from jbonsai_b200 import workload as wl, _flatten

e = jb.Engine.load([wl.VOICES["nitech"]])
utts = wl.timed_batch(e, 1024, wl.SEED_NITECH)
flat = _flatten(utts)
import statistics
CASES = [(2, [0.35, 0.40, 0.25]), (2, [0.30, 0.45, 0.25]), (2, [0.25, 0.45, 0.30]), (2, [0.2, 0.3, 0.3, 0.2]),
         (2, [0.4, 0.6]), (2, [0.5, 0.5]), (2, [0.3, 0.7]), (3, [0.35, 0.40, 0.25]), (3, [0.5, 0.5]), (3, [0.4, 0.6]),
         (3, [1.0]), (4, [0.5, 0.5]), (4, [1.0]), (4, [0.4, 0.6]), (6, [1.0]), (8, [1.0]), (1, [0.35, 0.40, 0.25])]
if len(sys.argv) > 1:
    CASES = [eval(a) for a in sys.argv[1:]]
for chunks, fr in CASES:
    e.set_pipeline_chunks(chunks)
    e.set_output_fractions(fr)
    for _ in range(4):
        out, _, tm = e.synthesize_batch_raw(flat); jb.lib.jb_free(out)
    ts = []
    for _ in range(8):
        t0 = time.perf_counter()
        out, _, tm = e.synthesize_batch_raw(flat); jb.lib.jb_free(out)
        ts.append((time.perf_counter() - t0) * 1e3)
    print(f"chunks={chunks} fractions={fr}: p50 {statistics.median(ts):.1f} min {min(ts):.1f} ms", {k: round(tm[k], 2) for k in ("host_frontend_ms", "kernels_ms", "mlpg_ms", "mlsa_ms", "d2h_ms", "mlsa_segments")}, flush=True)
