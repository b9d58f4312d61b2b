"""world_size-2 gloo test (CPU) of the multi-GPU path's host logic: contiguous node sharding and the
first-node-wins merge (one all-reduce MIN).  Each rank scores its shard with the CPU oracle (this is a
test; the product scores on the GPU), merges, and the result must equal scoring the whole cluster."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import walkai_nos_b200 as pkg
    from oracle import oracle as orc
    cfg = pkg.synth.make_config(4, 48, 300)  # 24 MIG + 24 MPS nodes
    o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
    res = {}
    for kind, name, width in ((0, "mig", len(cfg["names"])), (1, "mps", len(cfg["sizes"]))):
        off, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], kind)
        req, mask = pkg.synth.csr_to_dense(off, key, qty, width)
        shard, lo = pkg.dist.shard_snapshot(cfg[name], rank, world)
        if kind == 0:
            _, _, first = o.score_mig(shard, req, mask, want_choice=False)
            _, _, full = o.score_mig(cfg[name], req, mask, want_choice=False)
        else:
            _, _, _, first = o.score_mps(shard, req, mask, want_free=False)
            _, _, _, full = o.score_mps(cfg[name], req, mask, want_free=False)
        merged = pkg.dist.merge_first_node(torch.from_numpy(first.copy()), lo).numpy()
        both = pkg.dist.merge_first_node_keys(torch.from_numpy(first.copy()), torch.full((len(first),), lo, dtype=torch.int32)).numpy()
        assert np.array_equal(merged, both)
        res[name] = (merged, full)
    # the gate's ingredient: per-shard sums of used + free per key, one all-reduce(SUM)
    pres_ok = True
    for name in ("mig", "mps"):
        shard, _ = pkg.dist.shard_snapshot(cfg[name], rank, world)
        v = shard["valid"].astype(bool)
        per_gpu_valid = np.repeat(v, np.diff(shard["gpu_off"]))
        local = (shard["used"] + shard["free"])[per_gpu_valid].sum(axis=0).astype(np.int64)
        whole = cfg[name]
        wv = np.repeat(whole["valid"].astype(bool), np.diff(whole["gpu_off"]))
        full_sum = (whole["used"] + whole["free"])[wv].sum(axis=0).astype(np.int64)
        merged_p = pkg.dist.merge_presence(torch.from_numpy(local)).numpy()
        pres_ok &= bool(np.array_equal(merged_p, full_sum))
    ok = pres_ok and all(np.array_equal(m, f) for m, f in res.values())
    some = any((f >= 0).any() for _, f in res.values())
    np.save(os.path.join(out_dir, f"r{rank}.npy"), np.asarray([int(ok), int(some)]))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_and_merge_first_node_world2(tmp_path):
    world = 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        ok, some = np.load(os.path.join(str(tmp_path), f"r{r}.npy"))
        assert ok == 1 and some == 1


def test_shard_bounds_cover_all_nodes():
    sys.path.insert(0, ROOT)
    import walkai_nos_b200 as pkg
    for n in (0, 1, 7, 4096):
        for w in (1, 2, 3, 8):
            covered = []
            for r in range(w):
                lo, hi = pkg.dist.shard_bounds(n, r, w)
                covered += list(range(lo, hi))
            assert covered == list(range(n))
