"""Run the reference's own `train.py`, untouched, with the B200 hot path dropped in.

    python -m zebra_b200.launch --ref /path/to/Zebra [--level 0|1|2|3|ref] [--workdir DIR] -- <train.py flags>

`--ref`     checkout of LuckyLYM/Zebra (its train.py, model/, modules/, utils/ are imported as they are).
`--level`   1 (default): `utils.util.tppr_finder`, `utils.util.NeighborFinder`/`get_neighbor_finder` are
               replaced by the CUDA drop-ins and `GraphDiffusionEmbedding.compute_embedding_tppr_ensemble`
               by the device-resident version (fused gather);
            2: level 1 plus the lazy memory update on the device (zebra_pending_select + zebra_gather_*_lazy:
               no per-batch D2H of the neighbour ids, no host-side unique, no clone of the whole memory);
            3: level 2 plus the raw messages of a batch built by one kernel (zebra_b200/messages.py) and the
               link-prediction metrics off scikit-learn (zebra_b200/metrics.py): validation and
               test passes score their batches on the device with one D2H per pass, the training loop's per-batch
               metrics use cheap host restatements of the same definitions; every pass (training epochs included) is
               assembled before its loop and announced to the streaming finder, which streams it a window of
               batches ahead of the model's per-batch queries (zebra_b200/batches.py; ZEBRA_PASS_SCHEDULE=0 turns
               it off);
            0: only the finder classes are replaced (numpy in / numpy out, the strict drop-in);
            ref: nothing is replaced (the reference's numba CPU path; used for the side-by-side runs).
`--workdir` directory to run in; the reference reads `../data/{name}/ml_{name}.csv|.npy` relative to it
            and writes `./log`, `./saved_checkpoints` (train.py:74,110; utils/data_processing.py:70-81).

Two compatibility shims the reference needs on current Python/torch, independent of this project
(SURVEY.md H2/H3): `random.sample` on a set (utils/data_processing.py:102) and
`torch.load(weights_only=False)` for its pickled Memory module (train.py:285).
"""
from __future__ import annotations

import argparse
import os
import random
import runpy
import sys


def _shims():
    import torch
    _sample = random.sample

    def sample(population, k, **kw):
        if isinstance(population, (set, frozenset)):
            population = sorted(population)
        return _sample(population, k, **kw)
    random.sample = sample
    _load = torch.load

    def load(*a, **kw):
        kw.setdefault("weights_only", False)
        return _load(*a, **kw)
    torch.load = load


def install(level: str):
    """Patch the reference modules that are already importable (reference root first on sys.path)."""
    if level == "ref":
        return
    import utils.util as ref_util
    from . import finder, neighbor_finder
    ref_util.tppr_finder = finder.tppr_finder
    ref_util.NeighborFinder = neighbor_finder.NeighborFinder
    ref_util.get_neighbor_finder = neighbor_finder.get_neighbor_finder
    import modules.embedding_module as em          # binds `from utils.util import tppr_finder` -> ours
    em.tppr_finder = finder.tppr_finder
    if level in ("1", "2", "3"):
        from . import embedding
        embedding.LAZY_MEMORY = level in ("2", "3")
        embedding.patch(em.GraphDiffusionEmbedding)
    if level == "3":
        from . import batches, messages, metrics
        metrics.install()
        messages.install()
        if os.environ.get("ZEBRA_PASS_SCHEDULE", "1") != "0":
            import atexit
            import json
            batches.install(int(os.environ.get("ZEBRA_PASS_WINDOW", "64")))
            atexit.register(lambda: print("zebra_b200 announced passes: " + json.dumps(batches.pass_summary()), file=sys.stderr))


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--ref", required=True)
    ap.add_argument("--level", default="1", choices=["0", "1", "2", "3", "ref"])
    ap.add_argument("--workdir", default=None)
    ap.add_argument("rest", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    ref = os.path.abspath(a.ref)
    rest = a.rest[1:] if a.rest[:1] == ["--"] else a.rest
    sys.dont_write_bytecode = True
    sys.path.insert(0, ref)
    if a.workdir:
        os.makedirs(a.workdir, exist_ok=True)
        os.chdir(a.workdir)
    _shims()
    install(a.level)
    sys.argv = [os.path.join(ref, "train.py")] + rest
    runpy.run_path(sys.argv[0], run_name="__main__")


if __name__ == "__main__":
    main()
