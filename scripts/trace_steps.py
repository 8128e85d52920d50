#!/usr/bin/env python
"""Reads the TR lines of a -DB200Q_TC_TRACE build (one launch, all CTAs) and prints, per traced step, when
(relative to the step's earliest event, in us) the CTAs reach each event:
  5 activation warp starts waiting for the first block's rows   0 its B operand is written
  2 epilogue sees the first accumulator   3 last block of the CTA's share read   4 share finalized/published"""
import sys, collections, warnings
import numpy as np
warnings.simplefilter('ignore')
rows = collections.defaultdict(dict)
for ln in open(sys.argv[1]):
    f = ln.split()
    if f[0] != "TR" or len(f) != 9: continue
    b, j = int(f[1]), int(f[2])
    rows[j][b] = [int(v) for v in f[3:]]
t_prev_end = None
for j in sorted(rows):
    a = np.array([rows[j][b] for b in sorted(rows[j])], dtype=np.float64)  # [cta][event]
    a[a == 0] = np.nan
    base = np.nanmin(a) if t_prev_end is None else t_prev_end
    r = (a - base) / 1e3
    def st(e): return f"{np.nanmin(r[:, e]):6.2f}/{np.nanmedian(r[:, e]):6.2f}/{np.nanmax(r[:, e]):6.2f}"
    print(f"step {j}: wait {st(5)} | B {st(0)} | acc0 {st(2)} | last {st(3)} | fin {st(4)}   (min/med/max us after previous step's last fin)")
    t_prev_end = np.nanmax(a[:, 4])
