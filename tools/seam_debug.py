"""GPU-box helper: prints where the MGLSA seam leaves the oracle (development aid)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import jbonsai_b200 as jb
import oracle_lib as O
from test_gpu_vocoder_seam import lsp_track, lf0_track, lpf_track
from conftest import NITECH
e = jb.Engine.load([NITECH])
mode = sys.argv[1] if len(sys.argv) > 1 else "rows"
if mode == "rows":
    for stage, M, T in [(1, 25, 40), (1, 35, 40), (2, 35, 40), (3, 36, 97), (5, 35, 40), (8, 35, 40), (2, 35, 700)]:
        s, a, b = lsp_track(T, M, 7), lf0_track(T, 8), lpf_track(T)
        ref = O.vocoder_ex(s, a, b, stage=stage, use_log_gain=True, alpha=0.42)
        rows = np.stack([O.lsp_coefficients(f, stage, 0.42, True) for f in s])
        first = O.lsp_coefficients(s[0], stage, 0.42, True, filtered=False)
        r = e.prepare_tracks(rows, a, b, stage=stage, use_log_gain=True, alpha=0.42, first_rows=first[None, :])
        tm = r.run(); w, _ = r.fetch(); r.close()
        r = e.prepare_tracks(s, a, b, stage=stage, use_log_gain=True, alpha=0.42)
        r.run(); w2, _ = r.fetch(); grow, gfirst = r.rows(T, M); r.close()
        print(stage, M, T, "rows given: maxdiff", np.abs(w - ref).max(), "equal", np.array_equal(w, ref), "segs", tm["mlsa_segments"],
              "| lsp on device: wave diff", np.abs(w2 - ref).max(), "rows rel", np.abs(grow - rows).max() / np.abs(rows).max(),
              "first rel", np.abs(gfirst[0] - first).max() / np.abs(first).max(), flush=True)
else:
    stage, M = 8, 13
    lens = [40, 0, 1, 97, 7]
    spec = [lsp_track(T, M, 10 * stage + i) for i, T in enumerate(lens)]
    lf0 = [lf0_track(T, 50 + i) for i, T in enumerate(lens)]
    lpf = [lpf_track(T) for T in lens]
    waves = e.vocoder_synthesize(spec, lf0, lpf, stage=stage, use_log_gain=True, alpha=0.42)
    print([len(w) for w in waves], [float(np.abs(w).max()) if len(w) else 0 for w in waves])
if mode == "warm":
    M = 35
    lens = [1500, 700]
    for stage in (2, 3):
        spec = [lsp_track(T, M, 20 + i) for i, T in enumerate(lens)]
        lf0 = [lf0_track(T, 30 + i) for i, T in enumerate(lens)]
        lpf = [lpf_track(T) for T in lens]
        for warm in (1, 5, 10, 20, 40, 80, 160):
            e.set_mlsa_segmenting(warm, 1e-7)
            r = e.prepare_tracks(spec, lf0, lpf, stage=stage, use_log_gain=True, alpha=0.42)
            tm = r.run(); r.close()
            print(stage, "warm", warm, "segs", tm["mlsa_segments"], "cut worst", tm["mlsa_cut_worst"], "fallback", tm["mlsa_fallback"], "mlsa ms", tm["mlsa_ms"], flush=True)
if mode == "first":
    for stage, M, T in [(3, 36, 97), (2, 35, 70), (1, 25, 70), (1, 13, 200)]:
        s, a, b = lsp_track(T, M, 7), lf0_track(T, 8), lpf_track(T)
        ref, exc = O.vocoder_ex(s, a, b, stage=stage, use_log_gain=True, alpha=0.42, tap=True)
        rows = np.stack([O.lsp_coefficients(f, stage, 0.42, True) for f in s])
        first = O.lsp_coefficients(s[0], stage, 0.42, True, filtered=False)
        r = e.prepare_tracks(rows, a, b, stage=stage, use_log_gain=True, alpha=0.42, first_rows=first[None, :])
        tm = r.run(); w, _ = r.fetch(); r.close()
        bad = np.nonzero(w != ref)[0]
        print(stage, M, T, "segs", tm["mlsa_segments"], "n bad", len(bad), "first", bad[:3], "frame", bad[:1] // 240, "in-frame", bad[:1] % 240,
              "voiced frames", np.nonzero(a > -1e9)[0][:5], flush=True)
        if len(bad):
            i = bad[0]; print("   ", w[i], ref[i], "lf0 around", a[max(i // 240 - 1, 0):i // 240 + 2])
