"""Device-time measurement of the stage != 0 arm (LSP transform + MGLSA filter) on a synthetic batch.

No shipped voice uses this arm (none sets GAMMA), so the workload is synthetic line-spectral-pair tracks of
config-3 shape: N utterances of 700-1300 frames, 35 coefficients, 240 samples per frame. Prints one JSON line
per stage: device ms of the whole vocoder pass, the filter's share, samples/s and the FP64 instruction rate
against the measured FMA peak (the default kernel spends 3 fused multiply-adds and the coefficient add on a
tap: 4 FP64 instructions).

Usage (GPU box): python tools/bench_mglsa.py [--utts 512] [--stages 1,2,3]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import jbonsai_b200 as jb  # noqa: E402
from jbonsai_b200 import workload as wl  # noqa: E402


def tracks(n, M, seed):
    rng = np.random.default_rng(seed)
    spec, lf0, lpf = [], [], []
    base = np.linspace(0.12, 3.0, M)
    for u in range(n):
        T = int(rng.integers(700, 1301))
        t = np.arange(T)[:, None]
        x = base[None, :] + 0.03 * np.sin(0.07 * t + rng.uniform(0, 6.28, M)[None, :])
        x[:, 0] = 0.05 + 0.03 * np.sin(0.05 * t[:, 0])
        spec.append(np.sort(x, axis=1))
        v = np.full(T, -1e10)
        k = 0
        while k < T:
            m = int(rng.integers(5, 60))
            if rng.random() < 0.7:
                v[k:k + m] = np.log(rng.uniform(90, 260))
            k += m
        lf0.append(v)
        h = np.zeros((T, 31))
        h[:, 15] = 0.6
        h[:, 14] = h[:, 16] = 0.2
        lpf.append(h)
    return spec, lf0, lpf


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utts", type=int, default=512)
    ap.add_argument("--stages", default="1,2,3")
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    e = jb.Engine.load([wl.VOICES["nitech"]])
    fma_tflops, _ = jb.measure_peaks(0) if hasattr(jb, "measure_peaks") else (None, None)
    M = 35
    spec, lf0, lpf = tracks(a.utts, M, 1)
    for stage in [int(x) for x in a.stages.split(",")]:
        r = e.prepare_tracks(spec, lf0, lpf, stage=stage, use_log_gain=True, alpha=0.42)
        r.run()
        runs = [r.run() for _ in range(a.reps)]
        r.close()
        ms = float(np.median([d["device_ms"] for d in runs]))
        fms = float(np.median([d["mlsa_ms"] for d in runs]))
        n = int(runs[0]["samples"])
        segf = runs[0]["mlsa_seg_frames"]
        instr = (4 * (M - 1) + 6) * stage * segf * 240  # FP64 instructions issued (fused flavour), warm-up frames included
        line = {"workload": f"{a.utts} synthetic LSP utterances (700-1300 frames), stage {stage}, 35 coefficients",
                "device_ms": ms, "filter_ms": fms, "samples": n, "audio_s_per_s": n / 48000.0 / (ms * 1e-3),
                "filter_msamples_per_s": n / (fms * 1e-3) / 1e6, "segments": runs[0]["mlsa_segments"],
                "filtered_frames_incl_warmup": segf, "fallback": runs[0]["mlsa_fallback"],
                "cut_check_worst": runs[0]["mlsa_cut_worst"],
                "fp64_tinstr_per_s": instr / (fms * 1e-3) / 1e12}
        if fma_tflops:
            line["fp64_peak_tinstr_per_s"] = fma_tflops / 2
            line["frac_of_fp64_issue_peak"] = line["fp64_tinstr_per_s"] / (fma_tflops / 2)
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
