"""N > 1 host path on CPU: two ranks under torchrun with the gloo backend (the GPU box uses nccl).
Checks the weak-scaling shard (rank r owns utterances [r n, (r + 1) n), disjoint and contiguous) and
the reductions bench.py relies on (units summed, time = max over ranks)."""
import json
import os
import socket
import subprocess
import sys

import jbonsai_b200 as jb
from jbonsai_b200 import workload as wl
from jbonsai_b200.dist import Reducer, rank_utterance_range

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "dist_worker.py")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(world, n):
    env = dict(os.environ, PYTHONHASHSEED="0", OMP_NUM_THREADS="1")
    if world == 1:
        cmd = [sys.executable, WORKER, str(n)]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
               "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), WORKER, str(n)]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    return json.loads(line)


def test_rank_ranges_are_disjoint_and_contiguous():
    r = [rank_utterance_range(7, k) for k in range(4)]
    assert r == [(0, 7), (7, 14), (14, 21), (21, 28)]


def test_single_process_reducer_is_identity():
    red = Reducer(1, None)
    red.barrier()
    assert red.max(3.5) == 3.5 and red.sum(2.0) == 2.0


def test_two_ranks_gloo_match_one_process():
    n = 3
    two = _run(2, n)
    assert two["world"] == 2 and (two["first"], two["last"]) == (0, n)
    # the same 2n utterances built by one process
    eng = jb.Engine.load([wl.VOICES["nitech"]])
    utts = wl.timed_batch(eng, 2 * n, wl.SEED_NITECH, first=0)
    frames = [int(eng.durations(u).sum()) for u in utts]
    assert two["frames_rank0"] == sum(frames[:n])
    assert two["total_frames"] == sum(frames)          # units: sum over ranks
    assert two["max_ms"] == 11.0                        # time: max over ranks
    one = _run(1, n)
    assert one["total_frames"] == sum(frames[:n]) and one["max_ms"] == 10.0
