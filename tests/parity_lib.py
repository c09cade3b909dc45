"""Parity of one batch against the oracle — TEST INFRASTRUCTURE (used by tests/ and by bench.py's
cpu_baseline leg only; the product never imports this or anything under oracle/).

north star (BASELINE.json): excitation RNG sequence, pulse positions, frame/sample indexing bit-exact;
generated parameters <= 1e-9 relative; waveform <= 1e-6 absolute, identical after 16-bit quantisation."""
from __future__ import annotations

import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from oracle_lib import Oracle

NODATA = -1e10
PARAM_RTOL = 1e-9
WAVE_ATOL = 1e-6


def _pcm16(x: np.ndarray) -> np.ndarray:
    # the clamp of the reference's examples (examples/genji/main.rs:34-35): clamp, then truncate toward zero
    return np.trunc(np.clip(x, -32768.0, 32767.0)).astype(np.int16)


def _rel(a: np.ndarray, b: np.ndarray) -> float:
    """Element-wise relative error: max |a - b| / |b| (what tests/test_gpu_parity.py uses on short utterances)."""
    d = np.abs(a - b)
    return float((d / np.maximum(np.abs(b), 1e-300)).max()) if d.size else 0.0


def _rel_scaled(a: np.ndarray, b: np.ndarray) -> float:
    """Relative error with the denominator floored at 1e-3 of the RMS of the coefficient's own track: over
    10^5 frames some coefficient passes within 1e-4 of zero, where an absolute difference of 1e-13 (the GV sums
    are block reductions, not left-to-right) reads as 1e-9 "relative" although nothing is wrong with it."""
    if not a.size:
        return 0.0
    a2, b2 = a.reshape(len(a), -1), b.reshape(len(b), -1)
    floor = 1e-3 * np.sqrt(np.mean(b2 * b2, axis=0, keepdims=True))
    return float((np.abs(a2 - b2) / np.maximum(np.abs(b2), np.maximum(floor, 1e-300))).max())


def wave_parity(gpu_wave, gpu_off, orc_wave, orc_off) -> dict:
    """Waveform of the same utterances from the CUDA path and from the oracle."""
    gpu_off, orc_off = np.asarray(gpu_off, dtype=np.int64), np.asarray(orc_off, dtype=np.int64)
    n = len(orc_off) - 1
    same_len = bool(np.array_equal(gpu_off[: n + 1] - gpu_off[0], orc_off - orc_off[0]))
    out = {"utterances": n, "samples": int(orc_off[-1] - orc_off[0]), "length_mismatches": 0 if same_len else 1}
    if not same_len:
        out.update({"max_abs": float("inf"), "pcm16_mismatches": -1})
        return out
    g = np.asarray(gpu_wave)[int(gpu_off[0]): int(gpu_off[n])]
    o = np.asarray(orc_wave)[int(orc_off[0]): int(orc_off[n])]
    out["max_abs"] = float(np.abs(g - o).max()) if g.size else 0.0
    # "identical after 16-bit quantisation": a sample within 1e-6 of an integer can still truncate to the
    # neighbouring code; count those separately from real mismatches
    qg, qo = _pcm16(g), _pcm16(o)
    bad = np.flatnonzero(qg != qo)
    on_edge = int(np.count_nonzero(np.abs(o[bad] - np.rint(o[bad])) <= WAVE_ATOL)) if bad.size else 0
    out["pcm16_mismatches"] = int(bad.size) - on_edge
    out["pcm16_rounding_edge_cases"] = on_edge
    out["non_finite"] = int(np.count_nonzero(~np.isfinite(g)))
    return out


def stage_parity(engine, voice_paths, utts, nthreads: int | None = None, fma: bool = True) -> dict:
    """MLPG parameters, voiced/unvoiced mask and pulse train of `utts`: engine.debug_batch (CUDA) against
    Oracle.parameters / Oracle.synthesize_tap (one utterance per host thread)."""
    nthreads = nthreads or os.cpu_count() or 1
    dbg = engine.debug_batch(utts)
    fo = dbg["frame_off"]
    fp = int(engine.get("fperiod"))
    orc = Oracle(voice_paths)

    def one(u):
        labels = utts[u]
        f0, f1 = int(fo[u]), int(fo[u + 1])
        if not labels or f1 == f0:
            return (0, 0, 0.0, 0, 0.0, f1 - f0, 0.0, 0.0)
        mcp, lf0, lpf = orc.parameters(labels)
        tap = orc.synthesize_tap(labels, fma=fma)
        g_lf0 = dbg["lf0"][f0:f1]
        mask_bad = int(np.count_nonzero((g_lf0 == NODATA) != (lf0 == NODATA)))
        v = lf0 != NODATA
        prel = max(_rel_scaled(dbg["mcp"][f0:f1], mcp), _rel_scaled(dbg["lpf"][f0:f1], lpf),
                   _rel_scaled(g_lf0[v], lf0[v]) if not mask_bad else float("inf"))
        prel_el = max(_rel(dbg["mcp"][f0:f1], mcp), _rel(dbg["lpf"][f0:f1], lpf),
                      _rel(g_lf0[v], lf0[v]) if not mask_bad else float("inf"))
        pabs = max(float(np.abs(dbg["mcp"][f0:f1] - mcp).max()), float(np.abs(dbg["lpf"][f0:f1] - lpf).max()),
                   float(np.abs(g_lf0[v] - lf0[v]).max(initial=0.0)) if not mask_bad else float("inf"))
        gp = dbg["pulse"][f0 * fp: f1 * fp]
        pulse_bad = int(np.count_nonzero((gp != 0.0) != (tap["pulse"] != 0.0)))
        exc_abs = float(np.abs(dbg["exc"][f0 * fp: f1 * fp] - tap["gained"]).max())
        return (mask_bad, pulse_bad, prel, int(np.count_nonzero(tap["pulse"] != 0.0)), exc_abs, f1 - f0, prel_el, pabs)

    with ThreadPoolExecutor(max_workers=nthreads) as ex:
        res = list(ex.map(one, range(len(utts))))
    orc.close()
    return {
        "stage_utterances": len(utts), "stage_frames": int(sum(r[5] for r in res)),
        "lf0_mask_mismatches": int(sum(r[0] for r in res)),
        "pulse_position_mismatches": int(sum(r[1] for r in res)),
        "pulses": int(sum(r[3] for r in res)),
        "param_max_rel": float(max([r[2] for r in res] + [0.0])),
        "param_max_rel_definition": "max |gpu - oracle| / max(|oracle|, 1e-3 RMS of that coefficient's track)",
        "param_max_rel_elementwise": float(max([r[6] for r in res] + [0.0])),
        "param_max_abs": float(max([r[7] for r in res] + [0.0])),
        "excitation_max_abs": float(max([r[4] for r in res] + [0.0])),
    }


def verdict(p: dict) -> bool:
    return (p.get("length_mismatches", 0) == 0 and p.get("max_abs", 0.0) <= WAVE_ATOL and
            p.get("pcm16_mismatches", 0) == 0 and p.get("non_finite", 0) == 0 and
            p.get("lf0_mask_mismatches", 0) == 0 and p.get("pulse_position_mismatches", 0) == 0 and
            p.get("param_max_rel", 0.0) <= PARAM_RTOL)
