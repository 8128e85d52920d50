"""The CTA-pair GEMM's work distribution (csrc/gemm.cu TileSched), evaluated on the host through
b200q_debug_gemm_schedule: no GPU needed.  Invariants of the stream-K schedule for every shape class:
  * every 16-k step of every tile belongs to exactly one segment of exactly one pair;
  * a pair writes at most ONE partial tile (its slot in the workspace) and it is the first segment it processes;
  * a tile's finisher (kind 2) owns the tile's first steps, and the pairs it waits for - (pair, contributors_end) -
    are exactly the writers of the rest of that tile, each with exactly one partial of it;
  * a writer never waits (so no deadlock whatever the residency), and the finisher's segment comes right after the
    pair's own writer segment, before its whole tiles;
  * equal shares: the pairs' step counts differ by at most one chunk;
  * streamk = 0 is the classic round-robin of whole tiles.
"""
import ctypes

import numpy as np
import pytest

from mlc_llm_b200 import _lib


def schedule(streamk, n_tiles, n_ksteps, chunk, n_pairs):
    L = _lib.lib()
    out = (ctypes.c_int * (5 * 4096))()
    per_pair = []
    for p in range(n_pairs):
        n = L.b200q_debug_gemm_schedule(streamk, n_tiles, n_ksteps, chunk, n_pairs, p, out, 4096)
        assert 0 <= n <= 4096
        per_pair.append([tuple(out[5 * i:5 * i + 5]) for i in range(n)])
    return per_pair


CASES = [(80, 160, 4, 74), (80, 640, 4, 74), (320, 160, 4, 74), (8, 40, 4, 74), (20, 128, 20, 74), (240, 160, 4, 74),
         (1, 8, 4, 74), (74, 64, 4, 74), (75, 64, 4, 74), (7, 13, 4, 3), (100, 17, 20, 74), (3, 1000, 4, 74)]


@pytest.mark.parametrize("n_tiles,n_ksteps,chunk,n_pairs", CASES)
def test_streamk_schedule_invariants(n_tiles, n_ksteps, chunk, n_pairs):
    # (the launcher never starts more pairs than there are chunks: launch_gemm_pair, pairs_sk = min(W, n_pairs))
    n_pairs = min(n_pairs, n_tiles * -(-n_ksteps // chunk))
    sched = schedule(1, n_tiles, n_ksteps, chunk, n_pairs)
    assert all(len(segs) > 0 for segs in sched), "no idle pair"
    cover = np.zeros((n_tiles, n_ksteps), np.int32)
    writers = {}  # tile -> [pairs that hold a partial of it]
    steps = []
    for p, segs in enumerate(sched):
        kinds = [s[3] for s in segs]
        assert kinds.count(1) <= 1 and kinds.count(2) <= 1
        if 1 in kinds:
            assert kinds[0] == 1, "the partial a pair writes is the first thing it processes"
        if 2 in kinds:
            assert kinds.index(2) == (1 if 1 in kinds else 0), "the finisher's head comes before the pair's whole tiles"
        tot = 0
        for tile, ks0, ks1, kind, q_end in segs:
            assert 0 <= tile < n_tiles and 0 <= ks0 < ks1 <= n_ksteps
            assert ks0 % chunk == 0 and (ks1 % chunk == 0 or ks1 == n_ksteps)
            cover[tile, ks0:ks1] += 1
            tot += ks1 - ks0
            if kind == 0:
                assert (ks0, ks1) == (0, n_ksteps)
            elif kind == 1:
                assert ks0 > 0
                writers.setdefault(tile, []).append(p)
            else:
                assert ks0 == 0 and ks1 < n_ksteps
        steps.append(tot)
    assert (cover == 1).all(), "every step of every tile exactly once"
    busy = [s for s in steps if s]
    assert max(busy) - min(busy) <= chunk or n_ksteps % chunk, (min(busy), max(busy))
    # finishers wait for exactly the writers of their tile, and those are the pairs right behind them
    for p, segs in enumerate(sched):
        for tile, ks0, ks1, kind, q_end in segs:
            if kind == 2:
                assert list(range(p + 1, q_end)) == sorted(writers.get(tile, [])), (p, tile)
    for tile, ws in writers.items():
        fins = [p for p, segs in enumerate(sched) for s in segs if s[0] == tile and s[3] == 2]
        assert len(fins) == 1 and fins[0] == min(ws) - 1


@pytest.mark.parametrize("n_tiles,n_ksteps,chunk,n_pairs", CASES)
def test_whole_tile_schedule(n_tiles, n_ksteps, chunk, n_pairs):
    sched = schedule(0, n_tiles, n_ksteps, chunk, n_pairs)
    seen = []
    for p, segs in enumerate(sched):
        assert [s[0] for s in segs] == list(range(p, n_tiles, n_pairs))
        assert all(s[1:4] == (0, n_ksteps, 0) for s in segs)
        seen += [s[0] for s in segs]
    assert sorted(seen) == list(range(n_tiles))


def test_schedule_bad_arguments():
    L = _lib.lib()
    out = (ctypes.c_int * 16)()
    assert L.b200q_debug_gemm_schedule(1, 0, 8, 4, 74, 0, out, 3) == -1
    assert L.b200q_debug_gemm_schedule(1, 8, 8, 4, 74, 74, out, 3) == -1
    assert L.b200q_debug_gemm_schedule(1, 8, 8, 4, 74, 0, None, 3) == -1
