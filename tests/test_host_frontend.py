"""Host front-end of the product library (labels, tree lookup, interpolation, durations,
Condition clamps, errors) against the reference's own unit-test goldens:
src/duration.rs:141-180, src/model/mod.rs:184-431, src/label.rs:166-197, src/engine.rs:128-242."""
import numpy as np
import pytest

import jbonsai_b200 as jb
from conftest import NITECH, TOHOKU_HAPPY, TOHOKU_NEUTRAL
from labels import SAMPLE_SENTENCE_1

ALIGNED = [
    (0, 14925000), (14925000, 16725000), (16725000, 17525000), (17525000, 18125000),
    (18125000, 19725000), (19725000, 20825000), (20825000, 22725000), (22725000, 30325000),
]


@pytest.fixture(scope="module")
def engine():
    return jb.Engine.load([NITECH])


def test_metadata(engine):  # src/model/mod.rs:184-189
    assert engine.get("sampling_frequency") == 48000
    assert engine.get("fperiod") == 240
    assert len(engine.durations(SAMPLE_SENTENCE_1)) == 8 * 5


def test_tree_index(engine):  # src/model/mod.rs:192-212
    label = SAMPLE_SENTENCE_1[2]
    assert engine.tree_index(0, "duration", 0, 2, label) == (2, 144)
    assert engine.tree_index(0, "stream", 1, 2, label) == (2, 234)
    assert engine.tree_index(0, "gv", 1, 2, label) == (2, 3)


def test_durations_without_alignment():  # src/duration.rs:141-158
    e = jb.Engine.load([NITECH])
    assert list(e.durations(SAMPLE_SENTENCE_1)) == [
        8, 17, 14, 25, 15, 3, 4, 2, 2, 2, 2, 3, 3, 3, 3, 4, 3, 2, 2, 2, 3, 3, 6, 3, 2, 3,
        3, 3, 3, 2, 2, 1, 3, 2, 14, 22, 14, 26, 38, 5]
    e.set("speed", 1.2)
    assert list(e.durations(SAMPLE_SENTENCE_1)) == [
        6, 12, 11, 19, 14, 3, 4, 2, 2, 2, 2, 3, 3, 3, 3, 4, 3, 2, 2, 2, 3, 3, 6, 3, 2, 3,
        3, 3, 3, 2, 2, 1, 3, 2, 14, 18, 11, 16, 27, 4]


def test_durations_with_alignment():  # src/duration.rs:161-180 + src/label.rs:166-197
    e = jb.Engine.load([NITECH])
    e.set("phoneme_alignment", 1)
    lines = [f"{a} {b} {l}" for (a, b), l in zip(ALIGNED, SAMPLE_SENTENCE_1)]
    assert list(e.durations(lines)) == [
        36, 86, 48, 102, 27, 7, 11, 6, 6, 6, 2, 4, 3, 4, 3, 3, 3, 2, 2, 2, 3, 6, 14, 6, 3,
        4, 5, 6, 4, 3, 3, 1, 4, 4, 26, 28, 19, 42, 55, 8]


def test_get_parameters(engine):  # src/model/mod.rs:231-393
    mean, vari = engine.duration_parameters(SAMPLE_SENTENCE_1)
    assert len(mean) == 40
    assert mean[0] == 7.939206123352051 and vari[0] == 145.76211547851563
    assert mean[4] == 15.016390800476074 and vari[4] == 62.81060791015625
    assert mean[14] == 2.7264480590820313 and vari[14] == 3.725647211074829
    m, v, msd = engine.stream_parameters(SAMPLE_SENTENCE_1, 1)
    assert m.shape == (40, 3)
    assert list(m[0]) == [4.708907127380371, 0.010573429986834526, -0.019542237743735313]
    assert list(v[0]) == [0.027746843174099922, 0.0006717125070281327, 0.002855533268302679]
    assert msd[0] == 0.05000000074505806
    assert list(m[3]) == [0.0, 0.0, 0.0] and list(v[3]) == [1.0, 1.0, 1.0]
    assert list(m[8]) == [4.785215377807617, -0.0017961699049919844, -0.03521687909960747]
    assert msd[8] == 0.47957301139831543 and msd[9] == 0.9500000476837158
    assert list(v[14]) == [0.008521423675119877, 5.143996168044396e-5, 7.035945600364357e-5]
    gm, gv, sw = engine.gv_parameters(SAMPLE_SENTENCE_1, 1)
    assert list(gm) == [0.03621548041701317] and list(gv) == [0.00010934889724012464]
    assert list(sw) == [False] * 5 + [True] * 30 + [False] * 5


def test_multiple_models():  # src/model/mod.rs:396-431
    e = jb.Engine.load([TOHOKU_NEUTRAL, TOHOKU_HAPPY])
    e.set_interpolation_weight("duration", 0, [0.7, 0.3])
    e.set_interpolation_weight("parameter", 1, [0.7, 0.3])
    mean, vari = e.duration_parameters([SAMPLE_SENTENCE_1[2]])
    assert list(mean) == [3.345043873786926, 9.866290760040282, 5.616884994506836,
                          1.7678393721580503, 1.3566675186157227]
    assert list(vari) == [6.943870377540589, 59.23959312438964, 16.154539680480955,
                          0.9487730085849762, 1.2509666562080382]
    m, v, msd = e.stream_parameters([SAMPLE_SENTENCE_1[2]], 1)
    assert list(m[0]) == [5.354794883728027, -0.004957371624186635, 0.010301648452877997]
    assert list(v[0]) == [0.00590993594378233, 0.00017984867736231536, 0.00044686400215141473]
    assert msd[0] == 0.9955164790153503
    assert e.gv_parameters([SAMPLE_SENTENCE_1[2]], 1) is None  # tohoku: USE_GV=0


def test_condition_clamps():  # src/engine.rs:128-242
    e = jb.Engine.load([NITECH])
    e.set("speed", -3.0)
    assert e.get("speed") == 1.0e-06
    e.set("alpha", 1.5)
    assert e.get("alpha") == 1.0
    e.set("beta", -1.0)
    assert e.get("beta") == 0.0
    e.set("msd_threshold", 2.0, 1)
    assert e.get("msd_threshold", 1) == 1.0
    e.set("gv_weight", -2.0, 0)
    assert e.get("gv_weight", 0) == 0.0
    e.set("volume", 6.0)
    assert abs(e.get("volume") - 6.0) < 1e-12
    with pytest.raises(jb.JbError):
        e.set("no_such_key", 1.0)


def test_errors():
    with pytest.raises(jb.JbError) as ei:
        jb.Engine.load([])
    assert ei.value.kind == "EmptyVoice"
    with pytest.raises(jb.JbError) as ei:
        jb.Engine.load(["/nonexistent/voice.htsvoice"])
    assert ei.value.kind == "Io"
    with pytest.raises(jb.JbError) as ei:
        jb.Engine.load_from_bytes([b"this is not a voice"])
    assert ei.value.kind == "ParserError"
    e = jb.Engine.load([TOHOKU_NEUTRAL, TOHOKU_HAPPY])
    with pytest.raises(jb.JbError) as ei:
        e.set_interpolation_weight("duration", 0, [0.7, 0.2])  # must sum to 1
    assert ei.value.kind == "WeightError"
    with pytest.raises(jb.JbError):
        jb.Engine.load([NITECH]).durations(["garbage label"])


def test_load_from_bytes_equals_load():
    a = jb.Engine.load([NITECH])
    b = jb.Engine.load_from_bytes([open(NITECH, "rb").read()])
    assert list(a.durations(SAMPLE_SENTENCE_1)) == list(b.durations(SAMPLE_SENTENCE_1))


def test_mixed_voice_metadata_rejected():
    # nitech and tohoku share the shape (48k/240/5 states, 3 streams) -> accepted as a set,
    # like the reference's VoiceSet::new which only compares metadata (voice_set.rs:22-43).
    # gv differs (nitech has GV) -> MetadataError.
    with pytest.raises(jb.JbError) as ei:
        jb.Engine.load([NITECH, TOHOKU_NEUTRAL])
    assert ei.value.kind == "MetadataError"


def test_noise_matches_rng_definition(engine):
    # independent numpy restatement of excitation.rs:197-237 for the first values
    nxt, vals = 1, []
    def rnd():
        nonlocal nxt
        nxt = (nxt * 1103515245 + 12345) & 0xFFFFFFFFFFFFFFFF
        return float((nxt // 65536) % 32768) / 32767.0
    while len(vals) < 128:
        q, ss = [], []
        while len(ss) < 32:
            r1 = 2.0 * rnd() - 1.0
            r2 = 2.0 * rnd() - 1.0
            s = r1 * r1 + r2 * r2
            if 0.0 < s < 1.0:
                q += [r1, r2]
                ss.append(s)
        for k in range(32):
            m = float(np.sqrt(-2.0 * np.log(ss[k]) / ss[k]))
            vals += [q[2 * k] * m, q[2 * k + 1] * m]
    got = engine.noise(128)
    assert np.allclose(got, np.array(vals), rtol=0, atol=1e-15)
