"""Work distribution of one decode-program step, evaluated on the host through b200q_debug_gemv_schedule (no GPU
needed): a step of n_panels x row_blocks blocks is cut into equal shares of consecutive blocks, one per CTA; a panel
whose blocks fall into several shares is summed through partial-sum slots in the workspace.  What the kernels rely on
(csrc/gemv_prog.cu cta_begin / block_owner / partial_slots_per_panel, csrc/gemv_tc.cuh epilogue warps):

  * every block belongs to exactly one CTA, shares differ by at most one block;
  * a panel is either one whole segment, or has exactly ONE finisher whose contributor count is the number of CTAs
    that hold a piece of it; the writers' slots are distinct, inside [1, contributors) for the tcgen05 kernel's rule
    (the finisher keeps its own sum) and below the slot count the builder reserves per panel in both rules.

Shapes: the Llama-7B / 13B / 65B and RedPajama / dolly steps on 148 CTAs (the reference's decode functions,
mlc_llm/relax_model/llama.py:340-401), tensor-parallel shards of them, tiny and ragged ones, steps with fewer blocks
than CTAs.
"""
import ctypes

import pytest

from mlc_llm_b200 import _lib

SEG = 6


def _segments(L, n_panels, rb, G, cta, finisher_first):
    cap = n_panels // G + 3  # a share of T / G (+1) consecutive blocks touches at most this many panels
    out = (ctypes.c_int * (SEG * cap))()
    slots = ctypes.c_int(0)
    n = L.b200q_debug_gemv_schedule(n_panels, rb, G, cta, finisher_first, out, cap, ctypes.byref(slots))
    assert 0 <= n <= cap, (n_panels, rb, G, cta, n)
    return [tuple(out[SEG * i:SEG * i + SEG]) for i in range(n)], slots.value


def _panels(N):
    return (N + 127) // 128


# (panels, row blocks): K / 128 row blocks for q4 (a block = 16 packed word rows = 128 k), K / 64 for q8
STEPS = [
    (_panels(4096) * 3, 32), (_panels(4096), 32), (_panels(11008) * 2, 32), (_panels(4096), 86), (_panels(32000), 32),  # 7B q4
    (_panels(5120) * 3, 80), (_panels(13824) * 2, 80), (_panels(5120), 216),  # 13B q8
    (_panels(8192 // 8) * 3, 64), (_panels(22016 // 8) * 2, 64), (_panels(8192), 22),  # 65B, 8 shards
    (_panels(2560) * 3, 20), (_panels(10240), 20), (_panels(2560), 80), (_panels(50432), 20),  # RedPajama / dolly
    (1, 1), (1, 3), (1, 200), (2, 1), (3, 5), (7, 9), (147, 1), (148, 1), (149, 1), (5, 29), (1000, 2), (33, 43),
]


@pytest.mark.parametrize("finisher_first", [1, 0])
@pytest.mark.parametrize("G", [148, 132, 8, 1])
def test_every_block_once_one_finisher_per_panel_slots_in_range(G, finisher_first):
    L = _lib.lib()
    for n_panels, rb in STEPS:
        T = n_panels * rb
        cover = {}
        per_panel = {}
        shares = []
        slots_reserved = None
        for cta in range(G):
            segs, slots = _segments(L, n_panels, rb, G, cta, finisher_first)
            assert slots_reserved in (None, slots)
            slots_reserved = slots
            blocks = 0
            for panel, r0, r1, kind, slot, contrib in segs:
                assert 0 <= panel < n_panels and 0 <= r0 < r1 <= rb, (n_panels, rb, G, cta, segs)
                for r in range(r0, r1):
                    assert (panel, r) not in cover, ("block twice", n_panels, rb, G, cta, panel, r)
                    cover[(panel, r)] = cta
                blocks += r1 - r0
                per_panel.setdefault(panel, []).append((cta, r0, r1, kind, slot, contrib))
            shares.append(blocks)
        assert len(cover) == T, (n_panels, rb, G, len(cover))
        assert max(shares) - min(shares) <= 1 and sum(shares) == T
        for panel, pieces in per_panel.items():
            kinds = [p[3] for p in pieces]
            if kinds == [0]:
                assert pieces[0][1:3] == (0, rb) and pieces[0][5] == 1
                continue
            assert 0 not in kinds and kinds.count(2) == 1, (n_panels, rb, G, panel, pieces)
            fin = next(p for p in pieces if p[3] == 2)
            writers = [p for p in pieces if p[3] == 1]
            assert fin[5] == len(pieces) or T < G, (n_panels, rb, G, panel, pieces)
            # T < G: one CTA per block, the slot is the row block and the finisher adds up all row_blocks slots
            if T < G:
                assert fin[5] == rb and all(w[4] == w[1] for w in writers)
            # the finisher holds the panel's first (tcgen05 rule) or last (register-level rule) block
            assert (fin[1] == 0) if finisher_first else (fin[2] == rb)
            wslots = [w[4] for w in writers]
            assert len(set(wslots)) == len(wslots), ("slot collision", n_panels, rb, G, panel, pieces)
            assert all(0 <= s < slots_reserved for s in wslots + [fin[4]]), (n_panels, rb, G, panel, pieces, slots_reserved)
            if finisher_first:  # the finisher polls slots 1 .. contributors - 1
                assert sorted(wslots) == list(range(1, fin[5])) or T < G, (n_panels, rb, G, panel, pieces)
            else:  # the finisher adds the slots of the contributors AHEAD of it: 0 .. its own index - 1
                assert sorted(wslots) == list(range(0, fin[4])), (n_panels, rb, G, panel, pieces)


def test_rejects_bad_arguments():
    L = _lib.lib()
    out = (ctypes.c_int * 12)()
    assert L.b200q_debug_gemv_schedule(0, 4, 148, 0, 1, out, 2, None) == -1
    assert L.b200q_debug_gemv_schedule(4, 0, 148, 0, 1, out, 2, None) == -1
    assert L.b200q_debug_gemv_schedule(4, 4, 148, 148, 1, out, 2, None) == -1
    assert L.b200q_debug_gemv_schedule(4, 4, 148, 0, 1, None, 2, None) == -1
    assert L.b200q_debug_gemv_schedule(1 << 20, 1 << 12, 148, 0, 1, out, 2, None) == -1
    # more segments than the caller has room for: the count is still returned, nothing is written past max_seg
    n = L.b200q_debug_gemv_schedule(1000, 2, 8, 3, 1, out, 2, None)
    assert n > 2
