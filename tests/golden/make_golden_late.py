"""Two more recordings of the reference itself (build container only; see make_golden.py): prefixes of the
SuperUser- and Wiki-Talk-shaped streams (BASELINE.json configs[3], configs[4]) moved to the END of their
2,773- / 2,320-day spans, where timestamps are ~2e8 s and float32(ts) has a 16 s grid — the reference's
`dt = t_now - float32(ts)` (utils/util.py:450-465) then differs visibly from the exact difference and can
even be negative.  Million-node id space for the second case.

Run:  PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_late.py
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)
sys.dont_write_bytecode = True

from zebra_b200 import synth  # noqa: E402
import make_golden as G  # noqa: E402


def main():
    ref = G._ref()
    for name, case, n, shift in (("superuser", "stream_su_late", 12000, 2773.0 * 86400 - 3.4e6),
                                 ("wikitalk", "stream_wt_late", 12000, 2320.0 * 86400 - 5.0e5)):
        st = synth.make_stream(name, n_edges=n)
        neg = synth.make_negatives(st, 200)
        ts = st.ts + shift
        assert ts.max() > 1.9e8 and np.any(ts.astype(np.float32).astype(np.float64) != ts)
        G.run_stream_case(ref, case, st.src, st.dst, neg, ts, st.eidx, st.shape.table_nodes, 20, [0.1, 0.1], [0.5, 0.95], 200)


if __name__ == "__main__":
    main()
