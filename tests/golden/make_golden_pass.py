"""Recording of the reference's own evaluation loop head (build container only; see make_golden.py): the slices
and the seeded negatives `evaluation/evaluation.py:17-30` produces batch by batch with the reference's
`RandEdgeSampler` (utils/util.py:69-86), for a validation-like split of a Wikipedia-shaped stream — two passes in
a row with the same sampler (the reference resets it at the top of every pass), batch sizes 200 and 37 (ragged).

Run:  PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_pass.py
"""
from __future__ import annotations

import math
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
    ref = G._ref()                                   # the reference's utils.util, imported from /root/reference
    st = synth.make_stream("wikipedia", n_edges=6000)
    lo, hi = 4200, 5113                              # a "validation" slice; the sampler is built on the full data
    out = {"src": st.src, "dst": st.dst, "ts": st.ts, "eidx": st.eidx, "lo": lo, "hi": hi}
    for seed in (0, 3):
        sampler = ref.RandEdgeSampler(st.src, st.dst, seed=seed)
        for bs in (200, 37):
            sampler.reset_random_state()             # evaluation.py:10
            n = hi - lo
            negs = []
            for b in range(math.ceil(n / bs)):       # evaluation.py:19-27
                s, e = b * bs, min(n, (b + 1) * bs)
                _, neg = sampler.sample(e - s)
                negs.append(neg)
            out[f"neg_seed{seed}_bs{bs}"] = np.concatenate(negs)
    # the training loop's sampler: unseeded, numpy's global stream (train.py:88,140,203), two epochs over the
    # "training" slice [0, lo) with a ragged last batch, then one more draw that pins the stream position
    train = ref.RandEdgeSampler(st.src[:lo], st.dst[:lo])
    np.random.seed(0)
    for ep in range(2):
        negs = []
        for b in range(math.ceil(lo / 200)):
            s, e = b * 200, min(lo, (b + 1) * 200)
            _, neg = train.sample(e - s)
            negs.append(neg)
        out[f"train_neg_epoch{ep}"] = np.concatenate(negs)
    out["train_after"] = np.random.randint(0, 1 << 30, 4)
    np.savez_compressed(os.path.join(HERE, "pass_schedule.npz"), **out)
    print("wrote pass_schedule.npz", {k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    main()
