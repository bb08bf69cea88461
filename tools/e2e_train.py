"""End-to-end check: the reference's own train.py on a synthetic stream, (a) untouched (numba T-PPR on the
host + torch on the GPU) and (b) with the zebra_b200 drop-ins, same seed, same flags; reports per-epoch
tppr / train seconds (train.py:275) and validation / test AP (train.py:278,316-317).

    python tools/e2e_train.py --ref baseline/_ref --name wikipedia --epochs 3 [--edges N] [--levels ref 1]

Developer experiment (needs a checkout of the reference, which the tests and bench.py never read).
"""
import argparse
import json
import os
import re
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", required=True)
    ap.add_argument("--name", default="wikipedia")
    ap.add_argument("--edges", type=int, default=None)
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--strategy", default="streaming")
    ap.add_argument("--levels", nargs="+", default=["ref", "1"])
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "e2e"))
    a = ap.parse_args()
    from zebra_b200 import synth
    st = synth.make_stream(a.name, n_edges=a.edges)
    scratch = os.environ.get("ZEBRA_E2E_SCRATCH", "/tmp/zebra_e2e")       # data, checkpoints: not worth bringing back
    data_root = os.path.join(scratch, "data")
    synth.write_ml_format(st, data_root, a.name)
    results = {}
    os.makedirs(a.out, exist_ok=True)
    for level in a.levels:
        work = os.path.join(scratch, f"run_{level}")        # ../data relative to it
        os.makedirs(work, exist_ok=True)
        cmd = [sys.executable, "-m", "zebra_b200.launch", "--ref", a.ref, "--level", level, "--workdir", work, "--",
               "-d", a.name, "--bs", "200", "--n_epoch", str(a.epochs), "--tppr_strategy", a.strategy, "--topk", "20",
               "--alpha_list", "0.1", "0.1", "--beta_list", "0.5", "0.95", "--gpu", "0", "--n_degree", "10", "--n_layer", "2"]
        env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), PYTHONDONTWRITEBYTECODE="1")
        t = time.time()
        r = subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True)
        wall = time.time() - t
        open(os.path.join(a.out, f"train_{level}.log"), "w").write(r.stdout[-20000:] + "\n---stderr---\n" + r.stderr[-20000:])
        logdir = os.path.join(work, "log", a.name)
        text = ""
        if os.path.isdir(logdir):
            for f in os.listdir(logdir):
                text += open(os.path.join(logdir, f)).read()
        ep = re.findall(r"epoch: (\d+), tppr: ([\d.e+-]+), train: ([\d.e+-]+), val: ([\d.e+-]+)", text)
        val = re.findall(r"val ap: ([\d.]+), new node val ap: ([\d.]+)", text)
        test = re.findall(r"Test statistics: (Old|New) nodes -- auc: ([\d.]+), ap: ([\d.]+)", text)
        results[level] = {"rc": r.returncode, "wall_s": wall,
                          "epochs": [{"epoch": int(e), "tppr_s": float(t1), "train_s": float(t2), "val_s": float(t3)} for e, t1, t2, t3 in ep],
                          "val_ap": [[float(x), float(y)] for x, y in val],
                          "test_ap": {k: float(ap_) for k, _, ap_ in test}}
        print(level, json.dumps(results[level]), flush=True)
    json.dump({"name": a.name, "edges": st.n_edges, "strategy": a.strategy, "results": results},
              open(os.path.join(a.out, f"e2e_{a.name}_{a.strategy}.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
