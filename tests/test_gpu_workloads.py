"""Parity ON THE TIMED WORKLOADS (BASELINE.json configs 2-5), not only on short utterances: the default filter
is time-segmented with device-checked cuts, and whether a cut holds is a property of the data, so the batches
bench.py and tools/bench_configs.py time are compared with the oracle here at sizes the oracle finishes in
seconds. Tolerances: parity_lib (north star)."""
import numpy as np
import pytest

import jbonsai_b200 as jb
import parity_lib
from conftest import NITECH, TOHOKU_NEUTRAL
from jbonsai_b200 import workload as wl
from oracle_lib import Oracle

pytestmark = pytest.mark.gpu

NTHREADS = 16


def _batch_parity(voice, seed, n_utt, n_stage):
    e = jb.Engine.load([voice])
    utts = wl.timed_batch(e, n_utt, seed)  # the generator of bench.py (config 3: nitech, config 4: tohoku-f01)
    tm = {}
    w, off = e.synthesize_batch(utts, tm)
    o = Oracle([voice])
    ow, ooff = o.synthesize_batch(utts, fma=True, nthreads=NTHREADS)
    p = parity_lib.wave_parity(w, off, ow, ooff)
    p.update(parity_lib.stage_parity(e, [voice], utts[:n_stage], NTHREADS))
    return p, tm


def test_config3_generator_256_utterances_nitech():
    p, tm = _batch_parity(NITECH, wl.SEED_NITECH, 256, 48)
    assert tm["mlsa_segments"] > 1000 and tm["mlsa_fallback"] == 0
    assert parity_lib.verdict(p), p
    assert p["stage_frames"] > 40000 and p["pulses"] > 10000


def test_config4_generator_256_utterances_tohoku():
    p, tm = _batch_parity(TOHOKU_NEUTRAL, wl.SEED_TOHOKU, 256, 48)
    assert tm["mlsa_segments"] > 1000 and tm["mlsa_fallback"] == 0
    assert parity_lib.verdict(p), p


@pytest.mark.parametrize("voice", [NITECH, TOHOKU_NEUTRAL])
def test_config2_full_genji_lab_as_one_utterance(voice, genji_labels):
    # all 1456 labels: 164 s (nitech) / 129 s (tohoku-f01) of audio in one serial filter, cut into time segments
    e, o = jb.Engine.load([voice]), Oracle([voice])
    tm = {}
    w, off = e.synthesize_batch([genji_labels], tm)
    ow = o.synthesize(genji_labels, fma=True)
    assert len(w) == len(ow) and tm["mlsa_segments"] > 100 and tm["mlsa_fallback"] == 0
    p = parity_lib.wave_parity(w, off, ow, [0, len(ow)])
    assert p["max_abs"] <= parity_lib.WAVE_ATOL and p["pcm16_mismatches"] == 0, p
    # the sequential kernel (no cuts) agrees to rounding
    e.set_mlsa_variant(2)
    w2, _ = e.synthesize_batch([genji_labels])
    assert np.abs(w - w2).max() <= 1e-8


def test_config5_4096_streams_chunked_subset_vs_oracle():
    # 4096 concurrent streams stepped 20 frames at a time (SpeechGenerator::generate_step for all of them per
    # call); 64 of them, spread over the batch, are compared with the oracle sample for sample
    e = jb.Engine.load([NITECH])
    utts = wl.synthetic_batch(4096, wl.SEED_NITECH, lmin=6, lmax=12)  # ~1 s each: the test stays in seconds
    pick = list(range(0, 4096, 64))
    g = e.generator(utts)
    fp = g.fperiod()
    parts = {u: [] for u in pick}
    buf = None
    while True:
        buf, written, _ = g.generate_step(20, buf)
        if not written.any():
            break
        for u in pick:
            if written[u]:
                parts[u].append(buf[u, : written[u]].copy())
    g.close()
    o = Oracle([NITECH])
    ow, ooff = o.synthesize_batch([utts[u] for u in pick], fma=True, nthreads=NTHREADS)
    worst = 0.0
    for k, u in enumerate(pick):
        got = np.concatenate(parts[u]) if parts[u] else np.zeros(0)
        want = ow[ooff[k]: ooff[k + 1]]
        assert len(got) == len(want) and len(got) % fp == 0
        worst = max(worst, float(np.abs(got - want).max(initial=0.0)))
    assert worst <= parity_lib.WAVE_ATOL


def test_config3_full_size_properties():
    # BASELINE config 3 at its FULL size (1024 utterances, 248.7 M samples), where the oracle would take minutes:
    # properties that do not need it. (a) the time-segmented filter (default: ~7000 cuts, warm-up by loudness) against
    # the sequential kernel, which has no cuts at all; (b) sample counts = frames x fperiod per utterance; (c) the
    # same call again returns the same bits; (d) 16-bit delivery = the reference examples' clamp of the f64 samples.
    e = jb.Engine.load([NITECH])
    utts = wl.timed_batch(e, 1024, wl.SEED_NITECH)
    tm = {}
    w, off = e.synthesize_batch(utts, tm)
    assert tm["mlsa_fallback"] == 0 and tm["mlsa_segments"] > 7000
    assert len(w) == off[-1] == int(tm["samples"]) == 240 * int(tm["frames"]) and np.all(np.diff(off) % 240 == 0) and np.isfinite(w).all()
    w_again, _ = e.synthesize_batch(utts)
    assert np.array_equal(w, w_again)
    q, _ = e.synthesize_batch_pcm16(utts)
    assert np.array_equal(q, np.trunc(np.clip(w, -32768.0, 32767.0)).astype(np.int16))
    del w_again, q
    e.set_mlsa_variant(2)
    w_seq, off_seq = e.synthesize_batch(utts)
    assert list(off_seq) == list(off)
    assert np.abs(w - w_seq).max() <= 1e-8  # the cuts cost rounding only; against the oracle: the tests above
