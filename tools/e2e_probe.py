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
for chunks, fr in ((2, [0.30, 0.45, 0.25]), (1, [0.30, 0.45, 0.25]), (2, [0.25] * 4), (2, [0.25, 0.45, 0.30]),
                   (2, [0.35, 0.45, 0.20]), (2, [0.2, 0.3, 0.3, 0.2]), (2, [0.35, 0.40, 0.25]), (2, [0.4, 0.6]), (2, [0.3, 0.5, 0.2])):
    e.set_pipeline_chunks(chunks)
    e.set_output_fractions(fr)
    for _ in range(2):
        out, _, tm = e.synthesize_batch_raw(flat); jb.lib.jb_free(out)
    t0 = time.perf_counter()
    for _ in range(4):
        out, _, tm = e.synthesize_batch_raw(flat); jb.lib.jb_free(out)
    ms = (time.perf_counter() - t0) / 4 * 1e3
    print(f"chunks={chunks} fractions={fr}: {ms:.1f} ms", {k: round(tm[k], 2) for k in ("host_frontend_ms", "pack_ms", "kernels_ms", "mlpg_ms", "excitation_ms", "mlsa_ms", "d2h_ms", "mlsa_segments", "launches")})
