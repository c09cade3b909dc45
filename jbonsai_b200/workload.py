"""Synthetic label batches for BASELINE.json's configs.

Utterance u (seed base + u, splitmix64): the first line of SAMPLE_SENTENCE_1 (sil), a contiguous
slice of genji.lab of L in U[30,50] labels starting at a uniform random offset, then the last
line of SAMPLE_SENTENCE_1 (sil). Every label is a real full-context label, so every tree path
is valid. Pure host data generation: no dependency on the oracle or the CUDA library.
"""
from __future__ import annotations

import os
from typing import List

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GENJI = os.path.join(_ROOT, "tests", "fixtures", "genji.lab")
MODELS = os.path.join(_ROOT, "tests", "fixtures", "models")
VOICES = {
    "nitech": os.path.join(MODELS, "nitech_jp_atr503_m001.htsvoice"),
    "tohoku-f01-neutral": os.path.join(MODELS, "tohoku-f01-neutral.htsvoice"),
    "tohoku-f01-happy": os.path.join(MODELS, "tohoku-f01-happy.htsvoice"),
}
SEED_NITECH = 0xB0054100   # config 3 / 5
SEED_TOHOKU = 0x70F01000   # config 4

SIL_HEAD = ("xx^xx-sil+b=o/A:xx+xx+xx/B:xx-xx_xx/C:xx_xx+xx/D:xx+xx_xx/E:xx_xx!xx_xx-xx/"
            "F:xx_xx#xx_xx@xx_xx|xx_xx/G:4_4%0_xx_xx/H:xx_xx/I:xx-xx@xx+xx&xx-xx|xx+xx/J:1_4/K:1+1-4")
SIL_TAIL = ("a^i-sil+xx=xx/A:xx+xx+xx/B:xx-xx_xx/C:xx_xx+xx/D:xx+xx_xx/E:4_4!0_xx-xx/"
            "F:xx_xx#xx_xx@xx_xx|xx_xx/G:xx_xx%xx_xx_xx/H:1_4/I:xx-xx@xx+xx&xx-xx|xx+xx/J:xx_xx/K:1+1-4")

_M = (1 << 64) - 1


def splitmix64(state: int):
    state = (state + 0x9E3779B97F4A7C15) & _M
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M
    return state, z ^ (z >> 31)


def genji_labels() -> List[str]:
    with open(GENJI, encoding="utf-8") as f:
        return [l.strip() for l in f if l.strip()]


def _is_pause(label: str) -> bool:
    c = label.split("-", 1)[1].split("+", 1)[0]
    return c in ("sil", "pau")


def synthetic_utterance(genji: List[str], seed: int, lmin: int = 30, lmax: int = 50) -> List[str]:
    st, r1 = splitmix64(seed)
    st, r2 = splitmix64(st)
    L = lmin + r1 % (lmax - lmin + 1)
    start = r2 % (len(genji) - L)
    body = genji[start:start + L]
    while body and _is_pause(body[0]):
        body = body[1:]
    while body and _is_pause(body[-1]):
        body = body[:-1]
    return [SIL_HEAD] + body + [SIL_TAIL]


def synthetic_batch(n_utt: int, seed_base: int, first: int = 0, lmin: int = 30, lmax: int = 50):
    g = genji_labels()
    return [synthetic_utterance(g, seed_base + first + u, lmin, lmax) for u in range(n_utt)]


def label_frames(engine, labels: List[str]) -> List[int]:
    """Frames each label contributes at speed 1.0 (state durations are per label then:
    src/duration.rs:68-73 with rho = 0), from the host front-end of `engine`."""
    return [int(engine.durations([l]).sum()) for l in labels]


def timed_batch(engine, n_utt: int, seed_base: int, first: int = 0, lo: int = 800, hi: int = 1200):
    """BASELINE "~5 s each": utterance u (seed base + u) = sil, a contiguous slice of genji.lab
    starting at a uniform random offset and extended label by label until the utterance reaches
    a target drawn uniformly from [lo, hi] frames (4..6 s at 5 ms), then sil."""
    g = genji_labels()
    fr = label_frames(engine, g)
    edge = sum(label_frames(engine, [SIL_HEAD, SIL_TAIL]))
    out = []
    for u in range(n_utt):
        st, r1 = splitmix64(seed_base + first + u)
        st, r2 = splitmix64(st)
        target = lo + r1 % (hi - lo + 1)
        i = r2 % len(g)
        while _is_pause(g[i]):
            i = (i + 1) % len(g)
        body, total = [], edge
        while total < target:
            body.append(g[i])
            total += fr[i]
            i = (i + 1) % len(g)
        while len(body) > 1 and _is_pause(body[-1]):
            body.pop()
        out.append([SIL_HEAD] + body + [SIL_TAIL])
    return out
