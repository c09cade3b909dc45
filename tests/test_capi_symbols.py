"""The C-ABI library loads without a GPU and exports every symbol include/jbonsai_b200.h
declares; synthesis fails loudly (no CPU fallback) when no device is usable."""
import ctypes
import os
import re

import pytest

import jbonsai_b200 as jb
from conftest import NITECH, ROOT
from labels import SAMPLE_SENTENCE_1


def _declared():
    src = open(os.path.join(ROOT, "include", "jbonsai_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(jb_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_exported():
    names = _declared()
    assert len(names) >= 25
    lib = ctypes.CDLL(jb.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/jbonsai_b200.h but not exported"
    assert sorted(jb.C_ABI_SYMBOLS) == names


def test_version_and_load():
    assert "jbonsai_b200" in jb.version()
    e = jb.Engine.load([NITECH])
    assert e.get("sampling_frequency") == 48000
    assert e.get("fperiod") == 240
    assert e.get("alpha") == 0.55


def test_no_cpu_fallback():
    if jb.device_count() > 0:
        pytest.skip("a GPU is present")
    e = jb.Engine.load([NITECH])
    with pytest.raises(jb.JbError) as ei:
        e.synthesize(SAMPLE_SENTENCE_1)
    assert ei.value.kind == "Cuda"


def test_product_does_not_link_oracle():
    out = os.popen(f"ldd {jb.LIB_PATH}").read()
    assert "oracle" not in out
    for root, _, files in os.walk(os.path.join(ROOT, "jbonsai_b200")):
        for f in files:
            if f.endswith((".cc", ".cu", ".h", ".py")) and "build" not in root:
                txt = open(os.path.join(root, f), errors="replace").read()
                assert "oracle/" not in txt and "liboracle" not in txt, os.path.join(root, f)
