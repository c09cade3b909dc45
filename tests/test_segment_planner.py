"""The host-side planner of the time-segmented filter (Engine::cut_segments through jb_debug_plan_segments; no GPU):
the invariants the kernels rely on, with one warm-up for all cuts and with the warm-up by loudness.

Reference background: Vocoder::synthesize filters an utterance as one serial chain (src/vocoder/mod.rs:112-141);
the segments are this implementation's way of filling the chip, so their only contract is with mlsa_seg.cu:
segments tile every utterance, every segment but the first of an utterance starts `warm` frames early and knows
the segment it continues (`prev`)."""
import numpy as np
import pytest

import jbonsai_b200 as jb


def by_utterance(seg):
    out = {}
    for sid, (u, b, e, w, p) in enumerate(seg):
        out.setdefault(int(u), []).append((int(b), int(e), int(w), int(p), sid))
    return {u: sorted(v) for u, v in out.items()}


def check_invariants(seg, frames, warm_max):
    per = by_utterance(seg)
    for u, T in enumerate(frames):
        if T == 0:
            assert u not in per
            continue
        chain = per[u]
        assert chain[0][0] == 0 and chain[-1][1] == T  # the segments tile the utterance ...
        for k, (b, e, w, p, sid) in enumerate(chain):
            assert b < e and 0 <= w <= b and b - w <= warm_max
            if k == 0:
                assert p == -1 and w == 0  # the zero state at the start of an utterance is the true state
            else:
                pb, pe, _, _, psid = chain[k - 1]
                assert pe == b and p == psid  # ... without gaps, and `prev` is the segment that ends where this begins
                assert w < b or b == 0


@pytest.mark.parametrize("target", [24, 600, 7104])
def test_segments_tile_the_utterances_and_link_back(target):
    rng = np.random.default_rng(7)
    frames = list(rng.integers(800, 1201, 64)) + [0, 1, 47, 48, 49, 95, 97, 5000]
    seg = jb.plan_segments(frames, warm=40, target=target)
    check_invariants(seg, frames, 40)
    filt = seg[:, 2] - seg[:, 3]
    per = by_utterance(seg)
    for u, chain in per.items():  # equal filtered length within an utterance (the last one takes the remainder)
        lens = [e - w for b, e, w, p, _ in chain]
        assert max(lens[:-1], default=lens[-1]) - min(lens[:-1], default=lens[-1]) <= 1
        assert lens[-1] <= max(lens) + 1
    # the table is sorted for the warps: six consecutive segments run as long as the longest of them
    first_round = min(len(seg), target // 2)
    assert np.all(np.diff(filt[:first_round]) <= 0)


def test_short_utterances_are_not_cut_and_many_segments_are_not_forced():
    seg = jb.plan_segments([30, 47, 60], warm=40, target=7104)
    assert len(seg) == 3 and np.all(seg[:, 3] == 0) and np.all(seg[:, 4] == -1)  # whole utterances, no warm-up
    seg = jb.plan_segments([90], warm=40, target=7104)  # a segment is worth its warm-up from 24 own frames on
    assert len(seg) == 2 and min(seg[:, 2] - seg[:, 1]) >= 24
    check_invariants(seg, [90], 40)
    seg = jb.plan_segments([100000], warm=40, target=7104)  # never less than 48 own frames per segment
    assert len(seg) == 100000 // 48 and min(seg[:, 2] - seg[:, 1]) >= 24
    check_invariants(seg, [100000], 40)
    seg = jb.plan_segments([600000], warm=40, target=7104)  # one long utterance fills the chip on its own, once
    assert 0.9 * 7104 <= len(seg) <= 7104
    check_invariants(seg, [600000], 40)


def test_warm_up_by_loudness_gives_loud_cuts_the_short_warm_up():
    rng = np.random.default_rng(3)
    frames = [int(x) for x in rng.integers(800, 1201, 40)]
    # log gain: speech at 3.5, pauses at 0.3: a pause of 60-120 frames every ~300 frames
    gain = []
    for T in frames:
        g = np.full(T, 3.5)
        t = int(rng.integers(100, 300))
        while t < T:
            n = int(rng.integers(60, 121))
            g[t:t + n] = 0.3
            t += n + int(rng.integers(200, 400))
        gain.append(g)
    flat = np.concatenate(gain)
    base = jb.plan_segments(frames, warm=40, target=280)
    seg = jb.plan_segments(frames, warm=40, warm_loud=24, threshold=2.0, gain=flat, target=280)
    assert len(seg) == len(base)  # same number of segments per utterance: only the cuts move
    check_invariants(seg, frames, 40)
    off = np.concatenate([[0], np.cumsum(frames)])
    n_loud = n_quiet = 0
    for u, b, e, w, p in seg:
        if p < 0:
            continue
        W = b - w
        if b <= 24:
            assert W == b  # reaches the start of the utterance
            continue
        mean = flat[off[u] + b - 24: off[u] + b].mean()
        assert W == (24 if mean > 2.0 else 40)  # the rule, at the position the cut ended up at
        n_loud += W == 24
        n_quiet += W == 40
    assert n_loud > 2 * n_quiet > 0
    per = by_utterance(seg)
    for u, chain in per.items():  # still equal filtered length within an utterance
        lens = [e - w for b, e, w, p, _ in chain]
        assert max(lens[:-1], default=lens[-1]) - min(lens[:-1], default=lens[-1]) <= 2
    assert (seg[:, 2] - seg[:, 3]).sum() < 0.97 * (base[:, 2] - base[:, 3]).sum()  # fewer frames filtered
    # nothing loud enough: the plan of the rule switched off
    same = jb.plan_segments(frames, warm=40, warm_loud=24, threshold=1e9, gain=flat, target=280)
    assert np.array_equal(same, base)
