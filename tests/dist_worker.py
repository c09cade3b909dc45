"""Worker of tests/test_dist_gloo.py: one rank of the bench's N > 1 host logic on CPU (gloo).

Each rank builds its own utterance range of the synthetic workload (weak scaling, no data-path
collective), counts frames with the host front-end, and reduces counters exactly like bench.py
(sum of units, max of time). Rank 0 prints one JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import jbonsai_b200 as jb  # noqa: E402
from jbonsai_b200 import workload as wl  # noqa: E402
from jbonsai_b200.dist import Reducer, dist_env, rank_utterance_range  # noqa: E402


def main():
    n = int(sys.argv[1])
    rank, world, local = dist_env()
    red = Reducer(world, "gloo" if world > 1 else None, local)
    eng = jb.Engine.load([wl.VOICES["nitech"]])  # host front-end only: no GPU call below
    first, last = rank_utterance_range(n, rank)
    utts = wl.timed_batch(eng, n, wl.SEED_NITECH, first=first)
    frames = sum(int(eng.durations(u).sum()) for u in utts)
    digest = sum(hash(tuple(u)) % 1000003 for u in utts) % 1000003
    red.barrier()
    total_frames = red.sum(frames)
    slowest = red.max(10.0 + rank)          # stands for the per-rank step time
    digest_sum = red.sum(digest)
    red.barrier()
    if rank == 0:
        print(json.dumps({"world": world, "first": first, "last": last, "frames_rank0": frames,
                          "total_frames": total_frames, "max_ms": slowest, "digest_sum": digest_sum}))
    red.close()


if __name__ == "__main__":
    main()
