"""Host front-end in index mode (the default, SURVEY.md §8f.1: the host sends one pdf row per state and voice, the
device gathers): build_utterance resolves a label with one cached lookup per model (Model::get_all, the label hashed
once for the three streams) - checked here against Model::get_index state by state, the reference's
Model::get_index (src/model/voice/model.rs:50-67), on a separately parsed copy of the voice whose caches are cold.
No GPU: tests/native/frontend_rows_check.cc links the host sources only."""
import os
import shutil
import subprocess
import sys

import pytest

import jbonsai_b200 as jb
from jbonsai_b200 import workload as wl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "jbonsai_b200", "csrc", "host")


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("needs g++")
    exe = str(tmp_path_factory.mktemp("fe") / "frontend_rows_check")
    subprocess.check_call([cxx, "-O1", "-std=c++17", "-ffp-contract=off", "-pthread", "-I", HOST, "-o", exe,
                           os.path.join(ROOT, "tests", "native", "frontend_rows_check.cc"),
                           os.path.join(HOST, "frontend.cc"), os.path.join(HOST, "voice.cc")])
    return exe


@pytest.mark.parametrize("voice,seed", [("nitech", wl.SEED_NITECH), ("tohoku-f01-neutral", wl.SEED_TOHOKU)])
def test_index_mode_rows_equal_get_index_per_state(checker, tmp_path, voice, seed):
    utts = wl.synthetic_batch(24, seed)
    path = tmp_path / "utts.txt"
    path.write_text("".join("\n".join(u) + "\n\n" for u in utts), encoding="utf-8")
    out = subprocess.run([checker, wl.VOICES[voice], str(path)], capture_output=True, text=True)
    sys.stdout.write(out.stdout + out.stderr)
    assert out.returncode == 0 and ", 0 differ" in out.stdout
    n_labels = sum(len(u) for u in utts)
    assert int(out.stdout.split()[0]) >= 2 * 3 * 5 * n_labels  # two passes x streams x states (LPF of tohoku included)
