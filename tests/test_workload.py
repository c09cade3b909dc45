"""The synthetic label generator (bench workload) is deterministic and yields
labels the host front-end accepts."""
import jbonsai_b200 as jb
from conftest import NITECH
from jbonsai_b200 import workload as wl


def test_deterministic_and_valid():
    a = wl.synthetic_batch(8, wl.SEED_NITECH)
    b = wl.synthetic_batch(8, wl.SEED_NITECH)
    assert a == b
    assert wl.synthetic_batch(2, wl.SEED_NITECH, first=3) == a[3:5]
    e = jb.Engine.load([NITECH])
    frames = [int(e.durations(u).sum()) for u in a]
    assert all(300 <= f <= 2500 for f in frames), frames
    assert all(u[0] == wl.SIL_HEAD and u[-1] == wl.SIL_TAIL for u in a)


def test_timed_batch_is_about_five_seconds():
    e = jb.Engine.load([NITECH])
    a = wl.timed_batch(e, 16, wl.SEED_NITECH)
    assert a == wl.timed_batch(e, 16, wl.SEED_NITECH)
    assert wl.timed_batch(e, 4, wl.SEED_NITECH, first=5) == a[5:9]
    frames = [int(e.durations(u).sum()) for u in a]
    assert all(650 <= f <= 1320 for f in frames), frames
