import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import jbonsai_b200 as jb
from jbonsai_b200 import workload as wl
from conftest import NITECH
from labels import SAMPLE_SENTENCE_1
labels = wl.genji_labels()[:22]
e = jb.Engine.load([NITECH])
for warm in (4, 8, 12, 16, 20, 24, 32):
    e.set_mlsa_segmenting(warm, 1e-7)
    tm = {}
    e.synthesize_batch([SAMPLE_SENTENCE_1, labels], timings=tm)
    print(warm, {k: tm[k] for k in ("mlsa_segments", "mlsa_cut_worst", "mlsa_fallback", "mlsa_repaired")}, flush=True)
