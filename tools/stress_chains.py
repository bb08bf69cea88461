"""Developer stress run for the chain-warp protocol: many short adversarial spans (few nodes, most edges
between chain nodes, negatives on chain nodes) through plans, epochs and long calls with a SHORT wait
limit, so that a lost hand-off shows up as a reported time-out (with the wait kind) instead of a hang.
    ZEBRA_WAIT_TIMEOUT_MS=2000 python tools/stress_chains.py [rounds]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("ZEBRA_WAIT_TIMEOUT_MS", "2000")
import numpy as np
import oracle
from tests.test_gpu_chains import _stream, AL, BE
from zebra_b200.finder import tppr_finder

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 40
bad = 0
t0 = time.time()
for r in range(rounds):
    rng = np.random.default_rng(1000 + r)
    n_nodes = int(rng.choice([12, 40, 400, 5000]))
    hot = int(rng.integers(1, min(n_nodes - 2, 40)))
    E = int(rng.choice([4096, 9000, 24000]))
    src, dst, neg, ts, eidx = _stream(r, n_nodes, E, hot, p_hot=float(rng.choice([0.3, 0.7, 0.95])))
    N = n_nodes + 1
    o = oracle.OracleFinder(N, 20, 2, AL, BE)
    o.compute_val_tppr(src.astype(np.int64), dst.astype(np.int64), ts, eidx)
    nodes = np.arange(1, N)
    for mode in ("plan", "epoch", "call", "val"):
        f = tppr_finder(N, 20, 2, AL, BE)
        try:
            if mode == "plan":
                plan = f.make_plan(src, dst, neg, ts, eidx, 200)
                w = int(rng.choice([1, 5, 40, 200]))
                plan.run(0, plan.n_batches, w, out_slots=w)
                f.sync_check()
            elif mode == "epoch":
                for _ in f.stream_epoch(src, dst, neg, ts, eidx, 200, window=int(rng.choice([8, 64]))):
                    pass
            elif mode == "call":
                f.streaming_topk(np.concatenate([src, dst, neg]).astype(np.int32), np.concatenate([ts] * 3), eidx)
            else:
                f.compute_val_tppr(src.astype(np.int64), dst.astype(np.int64), ts, eidx)
            for j in range(2):
                for v, (e, n, t, s, nm) in zip(nodes, f.state(j, nodes)):
                    oe, on, ot, os_, onm = o.state(j, v)
                    if not (np.array_equal(e, oe) and np.array_equal(s, os_) and nm == onm):
                        raise AssertionError(f"state differs: inst {j} node {v}")
        except Exception as ex:  # noqa: BLE001
            bad += 1
            print(f"round {r} mode {mode} nodes {n_nodes} hot {hot} E {E}: {type(ex).__name__}: {ex}", flush=True)
print(f"{rounds} rounds, {bad} failures, {time.time() - t0:.0f} s")
