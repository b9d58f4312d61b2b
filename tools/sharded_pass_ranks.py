"""N processes, one GPU each: nosplan_score_pass_dev over a cluster whose nodes are sharded across the ranks, with
the exchange push fused into the scorers' epilogue, against the UNSHARDED answer of the CPU oracle (one oracle row
per distinct request vector) and against an unsharded single-GPU pass.  Handles travel over a gloo group; no NCCL.
Run: python tools/sharded_pass_ranks.py [world] [config] [nodes] [pods]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(rank, world, port, cfgno, nodes, pods):
    import torch
    import torch.distributed as dist

    import walkai_nos_b200 as pkg
    from oracle import oracle as orc
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(rank)
    cfg = pkg.synth.make_config(cfgno, nodes, pods)
    p = pkg.Planner(rank)
    p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
    mig, mig_lo = pkg.dist.shard_snapshot(cfg["mig"], rank, world)
    mps, mps_lo = pkg.dist.shard_snapshot(cfg["mps"], rank, world)
    if mig is not None and mig["n_nodes"] > 0:
        p.set_mig_snapshot(mig)
    if mps is not None and mps["n_nodes"] > 0:
        p.set_mps_snapshot(cfg["sizes"], mps)
    rows = [pkg.synth.pods_of_kind(cfg["pods"], k)[:3] for k in (0, 1)]
    D = [len(r[0]) - 1 if cfg[name] is not None else 0 for r, name in zip(rows, ("mig", "mps"))]
    # expected: the oracle on the whole cluster, one row per distinct vector
    o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
    exp = []
    for kind, name, width in ((0, "mig", len(cfg["names"])), (1, "mps", len(cfg["sizes"]))):
        if not D[kind]:
            continue
        off, key, qty = rows[kind]
        sig = key[:D[kind]].astype(np.int64) * 1000003 + qty[:D[kind]]
        uniq, fi = np.unique(sig, return_index=True)
        req, mask = pkg.synth.csr_to_dense(np.arange(len(uniq) + 1, dtype=np.int32), key[fi], qty[fi], width)
        if kind == 0:
            _, _, ef = o.score_mig(cfg[name], req, mask, threads=4, want_choice=False)
        else:
            _, _, _, ef = o.score_mps(cfg[name], req, mask, threads=4, want_free=False)
        exp.append(ef[np.searchsorted(uniq, sig)])
    exp = np.concatenate(exp).astype(np.int32)
    pkg.dist.open_exchange(p, D[0] + D[1])
    p.set_node_base(mig_lo, mps_lo)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    t = [[dev(x) for x in r] for r in rows]
    stream = torch.cuda.current_stream().cuda_stream
    ptr = lambda k, i: t[k][i].data_ptr() if D[k] else 0

    def run(flags, out):
        p.score_pass_dev(D[0], ptr(0, 0), ptr(0, 1), ptr(0, 2), D[1], ptr(1, 0), ptr(1, 1), ptr(1, 2), out.data_ptr(), flags, stream)
    for flags in (1, 3, 0, 1):
        out = torch.full((D[0] + D[1],), -7, dtype=torch.int32, device="cuda")
        run(flags, out)
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        assert np.array_equal(got, exp), (rank, flags, int((got != exp).sum()), got[:8], exp[:8])
    p.pass_status()
    # back to back without host synchronisation (the two key halves and the flags must hold up), timed
    outs = [torch.full((D[0] + D[1],), -7, dtype=torch.int32, device="cuda") for _ in range(30)]
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for o_ in outs:
        run(1, o_)
    e1.record()
    torch.cuda.synchronize()
    for i, o_ in enumerate(outs):
        assert np.array_equal(o_.cpu().numpy(), exp), (rank, "burst", i)
    p.pass_status()
    dist.barrier()
    if rank == 0:
        print(f"sharded pass ok: world {world}, cfg{cfgno} {nodes or 'full'} nodes, {D[0]}+{D[1]} rows, "
              f"{e0.elapsed_time(e1) / 30 * 1e3:.1f} us per pass incl. exchange; winners found for {(exp >= 0).mean():.2f} of the rows")
    pkg.dist.close_exchange(p)
    p.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    import torch.multiprocessing as mp
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    cfgno = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    nodes = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    pods = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    mp.spawn(worker, args=(world, 29000 + os.getpid() % 2000, cfgno, nodes, pods), nprocs=world, join=True)
