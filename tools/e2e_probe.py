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
for chunks, mode in ((1, 1), (2, 1), (2, 12), (2, 16), (3, 1), (4, 1), (1, 0), (2, 0)):
    e.set_pipeline_chunks(chunks)
    e.set_output_mode(mode)
    for _ in range(2):
        out, _, tm = e.synthesize_batch_raw(flat); jb.lib.jb_free(out)
    t0 = time.perf_counter()
    for _ in range(4):
        out, _, tm = e.synthesize_batch_raw(flat); jb.lib.jb_free(out)
    ms = (time.perf_counter() - t0) / 4 * 1e3
    print(f"chunks={chunks} mode={mode}: {ms:.1f} ms", {k: round(tm[k], 2) for k in ("host_frontend_ms", "pack_ms", "kernels_ms", "mlpg_ms", "excitation_ms", "mlsa_ms", "d2h_ms", "mlsa_segments", "launches")})
