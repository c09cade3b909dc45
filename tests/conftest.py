"""Test configuration. `-m "not gpu"` = oracle vs the reference's goldens, host logic, C-ABI
symbols; `-m gpu` = CUDA path vs oracle, through the C ABI."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

MODELS = os.path.join(ROOT, "tests", "fixtures", "models")
NITECH = os.path.join(MODELS, "nitech_jp_atr503_m001.htsvoice")
TOHOKU_NEUTRAL = os.path.join(MODELS, "tohoku-f01-neutral.htsvoice")
TOHOKU_HAPPY = os.path.join(MODELS, "tohoku-f01-happy.htsvoice")
GENJI = os.path.join(ROOT, "tests", "fixtures", "genji.lab")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # build what is missing (make is incremental; nvcc cross-compiles without a GPU)
    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    if not os.path.exists(os.path.join(ROOT, "jbonsai_b200", "libjbonsai_b200.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "jbonsai_b200", "csrc")])


def _has_gpu():
    import jbonsai_b200 as jb
    return jb.device_count() > 0


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def genji_labels():
    with open(GENJI, encoding="utf-8") as f:
        return [l.strip() for l in f if l.strip()]
