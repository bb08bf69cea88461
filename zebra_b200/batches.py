"""Batch assembly of a whole pass before its loop (reference: train.py:196-203,
evaluation/evaluation.py:19-33, utils/util.py:69-86, utils/data_processing.py:80-93).

The reference slices every batch out of the split on the host, draws its negatives, and hands the five arrays to
the model, one batch at a time; the T-PPR query of the batch then runs synchronously inside the model call.  For
the validation / test passes everything about those batches is fixed before the loop starts: the split is static
and the sampler is seeded and reset at the top of the pass (evaluation.py:9-10).  `PassSchedule` therefore

* draws the negatives of every batch up front — with the sampler's own `sample(size)` calls in the loop's order,
  so the numbers ARE the reference's (numpy's MT19937 stream; nothing is re-implemented);
* lays the pass out as the five flat arrays the loop slices (`batch(b)` returns the same views the reference's
  loop builds), and keeps a device-resident copy of them (`device_arrays()`): one H2D per pass;
* announces the pass to the streaming finder (`tppr_finder.schedule_pass`): the span is uploaded once, streamed
  `window` batches per launch ahead of the model, and each of the model's per-batch queries is answered from rows
  that are already in HBM.  A query that does not match the announcement falls back to the per-call path after a
  roll-back (finder.py), so announcing never changes a result.

`RandEdgeSampler` mirrors utils/util.py:69-86 (same constructor, same draws) for callers that do not import the
reference; `read_edge_csv` reads the `ml_{name}.csv` edge list (utils/data_processing.py:81-91).
The training loop draws its negatives from numpy's GLOBAL stream (train.py:88 seeds it, train.py:203 /
utils/util.py:79-80 are its only consumers in the reference), so its epoch is predictable as well: `TrainLookahead`
(installed on the reference's RandEdgeSampler by `install()`, launcher level 3) draws the whole epoch at the epoch's
first `sample()` call — same calls, same order, hence the same numbers and the same stream position at the end of
the epoch — announces the epoch to the finder and hands the draws out call by call.  A call that asks for another
size than predicted puts the global stream back to where the reference's loop would have it and ends the lookahead.
"""
from __future__ import annotations

import math

import numpy as np

PASS_LOG = []          # one stats dict per announced pass (diagnostics; the launcher prints a summary at exit)


class RandEdgeSampler:
    """Same constructor, attributes and draws as utils/util.py:69-86: negatives are uniform over the DISTINCT
    sources / destinations; `size` source indices are drawn first, then `size` destination indices, from the
    sampler's own RandomState when seeded and from numpy's global stream otherwise."""

    def __init__(self, src_list, dst_list, seed=None):
        self.src_list, self.dst_list = np.unique(src_list), np.unique(dst_list)
        self.seed = seed
        if seed is not None:
            self.reset_random_state()

    def sample(self, size):
        rng = np.random if self.seed is None else self.random_state
        picks = [rng.randint(0, len(pool), size) for pool in (self.src_list, self.dst_list)]     # sources first
        return self.src_list[picks[0]], self.dst_list[picks[1]]

    def reset_random_state(self):
        self.random_state = np.random.RandomState(self.seed)


def read_edge_csv(path):
    """The edge list of `ml_{name}.csv` as the reference reads it (utils/data_processing.py:81-91): columns u, i,
    ts, label, idx -> (sources, destinations, timestamps, edge_idxs, labels), in file (chronological) order."""
    import pandas as pd
    df = pd.read_csv(path)
    return df.u.values, df.i.values, df.ts.values, df.idx.values, df.label.values


class PassSchedule:
    """All batches of one pass over `data` (an object with sources, destinations, timestamps, edge_idxs — the
    reference's Data), assembled before the loop.  `sampler` is the pass's negative sampler (already reset); its
    `sample(size)` is called once per batch, in batch order, exactly as the reference's loop does."""

    def __init__(self, data, sampler, batch_size):
        self.batch_size = int(batch_size)
        self.sources = np.asarray(data.sources)
        self.destinations = np.asarray(data.destinations)
        self.timestamps = np.asarray(data.timestamps)
        self.edge_idxs = np.asarray(data.edge_idxs)
        self.n = len(self.sources)
        self.n_batches = math.ceil(self.n / self.batch_size) if self.n else 0
        negs = []
        for b in range(self.n_batches):
            lo, hi = self.bounds(b)
            _, neg = sampler.sample(hi - lo)            # evaluation.py:27, train.py:203
            negs.append(np.asarray(neg))
        self.negatives = np.concatenate(negs) if negs else np.zeros(0, self.destinations.dtype)
        self._finder = None
        self._announced = self._last = None
        self._dev = None

    def bounds(self, b):
        lo = b * self.batch_size
        return lo, min(self.n, lo + self.batch_size)

    def batch(self, b):
        """(sources, destinations, negatives, timestamps, edge_idxs) of batch b: what the reference's loop hands to
        `compute_edge_probabilities` (train.py:199-203, evaluation.py:22-30)."""
        lo, hi = self.bounds(b)
        return (self.sources[lo:hi], self.destinations[lo:hi], self.negatives[lo:hi], self.timestamps[lo:hi],
                self.edge_idxs[lo:hi])

    def device_arrays(self, device=None):
        """The pass resident on the device: dict of CUDA tensors (sources, destinations, negatives int64;
        timestamps float64; edge_idxs int64), batch b = rows [b*bs, (b+1)*bs).  One H2D per array per pass."""
        import torch
        if self._dev is None:
            dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
            self._dev = {
                "sources": torch.from_numpy(np.ascontiguousarray(self.sources, dtype=np.int64)).to(dev),
                "destinations": torch.from_numpy(np.ascontiguousarray(self.destinations, dtype=np.int64)).to(dev),
                "negatives": torch.from_numpy(np.ascontiguousarray(self.negatives, dtype=np.int64)).to(dev),
                "timestamps": torch.from_numpy(np.ascontiguousarray(self.timestamps, dtype=np.float64)).to(dev),
                "edge_idxs": torch.from_numpy(np.ascontiguousarray(self.edge_idxs, dtype=np.int64)).to(dev),
            }
        return self._dev

    # ---- the streaming finder runs ahead of the loop ----
    def announce(self, model_or_finder, window=64, kind="eval"):
        """Announce the pass to the streaming finder of `model_or_finder` (a zebra_b200 tppr_finder, or a model whose
        embedding module holds one with tppr_strategy 'streaming').  Returns True if it was announced."""
        from .finder import tppr_finder
        f = model_or_finder
        if not isinstance(f, tppr_finder):
            em = getattr(model_or_finder, "embedding_module", None)
            if getattr(em, "tppr_strategy", None) != "streaming":
                return False
            f = getattr(em, "tppr_finder", None)
        if not isinstance(f, tppr_finder) or not self.n_batches:
            return False
        self._finder = f
        self._announced = f.schedule_pass(self.sources, self.destinations, self.negatives, self.timestamps,
                                          self.edge_idxs, self.batch_size, window)
        PASS_LOG.append({"kind": kind, "pass": self._announced})
        return True

    def stats(self):
        a = self._announced or self._last
        return None if a is None else {"batches": a.n_batches, "served": a.hits, "launches": a.launches,
                                       "rolled_back": a.rolled_back}

    def close(self):
        """End of the pass (normal or not): anything still announced is settled."""
        a, self._announced = self._announced, None
        self._last = a or self._last
        if a is not None and self._finder is not None and self._finder._sched is a:
            a.abandon()


class TrainLookahead:
    """The training epoch of the reference's loop (train.py:195-203), assembled at its first batch.

    `train_data`: the split the loop slices (utils/data_processing.py:107, element 1 of get_data's result).
    `take(sampler, size, draw)` stands in for `sampler.sample(size)` of the UNSEEDED sampler; `draw(size)` is the
    sampler's original method.  Nothing else in the reference reads numpy's global stream, so drawing batch b+1..B
    before batch b is consumed changes no number; the stream position after the epoch's last call is the
    reference's.  Between epochs nothing is drawn ahead."""

    def __init__(self, train_data, window=64, finder=None):
        self.data = train_data
        self.window = int(window)
        self.finder = finder            # callable returning the streaming finder to announce to (or None)
        self.sizes, self.draws, self.states, self.k = [], [], [], 0
        self.enabled = True
        self.epochs = 0
        self._announced = None

    def _begin_epoch(self, size, draw):
        n = len(self.data.sources)
        bs = int(size)
        self.sizes = [bs] * (n // bs) + ([n % bs] if n % bs else [])
        self.state0 = np.random.get_state()
        self.draws, self.states = [], []
        for s in self.sizes:
            self.draws.append(draw(s))
            self.states.append(np.random.get_state())
        self.k = 0
        self.epochs += 1
        f = self.finder() if self.finder is not None else None
        if f is not None and self.sizes:
            neg = np.concatenate([d[1] for d in self.draws])
            self._announced = f.schedule_pass(self.data.sources, self.data.destinations, neg, self.data.timestamps,
                                              self.data.edge_idxs, bs, self.window)
            PASS_LOG.append({"kind": "train", "pass": self._announced})

    def take(self, sampler, size, draw):
        n = len(self.data.sources)
        if not self.enabled or n == 0 or size <= 0:
            return draw(size)
        if self.k >= len(self.sizes):
            if size > n:
                return draw(size)
            self._begin_epoch(size, draw)
        if size != self.sizes[self.k]:
            # not the loop that was predicted: the global stream goes back to where that loop has it
            np.random.set_state(self.states[self.k - 1] if self.k else self.state0)
            a, self._announced = self._announced, None
            if a is not None:
                a.abandon()
            self.enabled = False
            self.sizes, self.draws, self.states, self.k = [], [], [], 0
            return draw(size)
        out = self.draws[self.k]
        self.draws[self.k] = None
        self.k += 1
        return out


def pass_summary():
    """Counts over PASS_LOG: {kind: {passes, batches, served, launches, rolled_back}}."""
    out = {}
    for e in PASS_LOG:
        a = e["pass"]
        d = out.setdefault(e["kind"], {"passes": 0, "batches": 0, "served": 0, "launches": 0, "rolled_back": 0})
        d["passes"] += 1
        d["batches"] += a.n_batches
        d["served"] += a.hits
        d["launches"] += a.launches
        d["rolled_back"] += int(a.rolled_back)
    return out


def _live_streaming_finder():
    from .finder import tppr_finder
    for r in reversed(tppr_finder._live):      # train.py builds one model per run: the newest finder is the one in use
        f = r()
        if f is not None:
            return f
    return None


def install(window=64):
    """Launcher level 3: the reference's unseeded RandEdgeSampler (the training sampler, train.py:140) hands out an
    epoch drawn at its first batch, and that epoch is announced to the streaming finder.  The seeded samplers
    (validation / test) are left alone — their passes are announced by metrics.eval_edge_prediction."""
    import utils.data_processing as dp          # the reference's modules (reference root is on sys.path)
    import utils.util as U
    state = {}
    get_data = dp.get_data

    def get_data_capture(*a, **kw):
        res = get_data(*a, **kw)
        state["look"] = TrainLookahead(res[1], window, _live_streaming_finder)
        return res
    dp.get_data = get_data_capture
    sample = U.RandEdgeSampler.sample

    def sample_ahead(self, size):
        look = state.get("look")
        if self.seed is not None or look is None or len(self.src_list) == 0:
            return sample(self, size)
        if state.setdefault("sampler", self) is not self:      # a second unseeded sampler: not the loop we know
            return sample(self, size)
        return look.take(self, size, lambda s: sample(self, s))
    U.RandEdgeSampler.sample = sample_ahead
    return state
