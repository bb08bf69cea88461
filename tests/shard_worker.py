"""torchrun worker of tests/test_gpu_sharded_mp.py: one process per GPU, node-sharded T-PPR over
CUDA IPC + NVLink, every rank's result compared with the C oracle byte for byte."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import oracle
    from zebra_b200 import synth
    from zebra_b200.sharded import create_sharded, exchange_stats

    name, n_edges, bs, window = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    k, al, be = 20, [0.1, 0.1], [0.5, 0.95]
    st = synth.make_stream(name, n_edges=n_edges, seed=77)
    neg = synth.make_negatives(st, bs)
    N = st.shape.table_nodes
    n_batches = (n_edges + bs - 1) // bs
    wep = max(world, (min(window, n_batches) // world) * world)      # window of the pipelined epoch: a multiple of world
    f = create_sharded(N, k, 2, al, be, window * bs, max(n_batches, 2 * wep) * 2 * 16 * 3 * bs * k)
    plan = f.make_plan(st.src, st.dst, neg, st.ts, st.eidx, bs)
    for rep in range(2):                       # second pass: reset + rerun must give the same bytes
        f.reset_tppr()
        torch.cuda.synchronize()
        dist.barrier()
        plan.run(0, n_batches, window, out_slots=n_batches)
        f.sync_check()
        dist.barrier()
        torch.cuda.synchronize()
        cpu = oracle.OracleFinder(N, k, 2, al, be)
        for b in range(n_batches):
            lo, hi = b * bs, min(n_edges, (b + 1) * bs)
            nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
            ref = cpu.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
            got = plan.views(f.out, b, 0, n_batches)
            for la, lb in zip(got, ref):
                for x, y in zip(la, lb):
                    assert np.array_equal(x.cpu().numpy(), y), f"rank {rank} batch {b} pass {rep}"
        ids = np.unique(np.concatenate([st.src, st.dst]))
        mine = ids[(ids & (world - 1)) == rank][:300]
        for j in range(2):
            for a, (e, n, t, s, nm) in zip(mine, f.state(j, mine)):
                oe, on, ot, os_, onm = cpu.state(j, int(a))
                assert np.array_equal(e, oe) and np.array_equal(s, os_) and nm == onm
        dist.barrier()
    # pipelined sharded epoch (host arrays in, this rank's batches out in pinned host memory), twice
    f.set_row_mode(1)
    for rep in range(2):
        f.reset_tppr()
        cpu = oracle.OracleFinder(N, k, 2, al, be)
        seen = []
        ep = f.epoch(st.src, st.dst, neg, st.ts, st.eidx, bs, window=wep)
        nxt = 0
        for b, nd, ei, dt, w in ep:
            while nxt <= b:                      # the oracle walks every batch, this rank checks its own
                lo, hi = nxt * bs, min(n_edges, (nxt + 1) * bs)
                nodes = np.concatenate([st.src[lo:hi], st.dst[lo:hi], neg[lo:hi]]).astype(np.int32)
                ref = cpu.streaming_topk(nodes, np.concatenate([st.ts[lo:hi]] * 3), st.eidx[lo:hi])
                nxt += 1
            for got, want in zip((nd, ei, dt, w), ref):
                for x, y in zip(got, want):
                    assert np.array_equal(x, y), f"epoch: rank {rank} batch {b} pass {rep}"
            seen.append(b)
        assert seen == list(range(rank, n_batches, world)), (rank, seen[:5], len(seen))
        dist.barrier()
    f.set_row_mode(0)
    if rank == 0:
        print("exchange:", exchange_stats(st.src, st.dst, neg, world, k, 2), flush=True)
    print(f"SHARD_OK rank {rank}/{world} kernel_ms {f.last_kernel_ms():.2f} {f.counters()}", flush=True)
    f.disconnect()
    dist.barrier()
    f.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
