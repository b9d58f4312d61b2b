"""Two (or N) processes, one GPU each: the library's peer-memory first-node exchange against numpy.
Handles travel over a gloo group; no NCCL involved.  Run: python tools/xchg_two_ranks.py [world]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def worker(rank, world, port):
    import torch
    import torch.distributed as dist
    import walkai_nos_b200 as pkg
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(rank)
    p = pkg.Planner(rank)
    pkg.dist.open_exchange(p, 60000)
    stream = torch.cuda.current_stream().cuda_stream
    for step, n in enumerate((60000, 33, 50000, 1, 60000, 777, 60000, 60000)):
        # every rank can compute every rank's keys: seeded by (step, rank)
        allk = []
        for r in range(world):
            rng = np.random.default_rng(1000 * step + r)
            k = rng.integers(0, 1 << 22, size=n).astype(np.int32)
            k[rng.random(n) < 0.4] = -1
            allk.append(k)
        big = np.where(np.stack(allk) >= 0, np.stack(allk), np.iinfo(np.int32).max).min(axis=0)
        exp = np.where(big == np.iinfo(np.int32).max, -1, big).astype(np.int32)
        d_in = torch.from_numpy(allk[rank]).cuda()
        d_out = torch.full((n,), -7, dtype=torch.int32, device="cuda")
        p.exchange_min_dev(n, d_in.data_ptr(), d_out.data_ptr(), stream)
        torch.cuda.synchronize()
        got = d_out.cpu().numpy()
        assert np.array_equal(got, exp), (rank, step, int((got != exp).sum()))
    p.exchange_status()
    # back-to-back steps without host synchronisation in between (flags and the two key halves must hold up)
    n = 60000
    ins, outs, exps = [], [], []
    for step in range(40):
        allk = []
        for r in range(world):
            rng = np.random.default_rng(77000 + 100 * step + r)
            k = rng.integers(0, 1 << 22, size=n).astype(np.int32)
            k[rng.random(n) < 0.4] = -1
            allk.append(k)
        big = np.where(np.stack(allk) >= 0, np.stack(allk), np.iinfo(np.int32).max).min(axis=0)
        exps.append(np.where(big == np.iinfo(np.int32).max, -1, big).astype(np.int32))
        ins.append(torch.from_numpy(allk[rank]).cuda())
        outs.append(torch.full((n,), -7, dtype=torch.int32, device="cuda"))
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(40):
        p.exchange_min_dev(n, ins[i].data_ptr(), outs[i].data_ptr(), stream)
    e1.record()
    torch.cuda.synchronize()
    for i in range(40):
        assert np.array_equal(outs[i].cpu().numpy(), exps[i]), (rank, "burst", i)
    p.exchange_status()
    dist.barrier()
    if rank == 0:
        print(f"exchange ok: world {world}, {e0.elapsed_time(e1) / 40 * 1e3:.1f} us per exchange of {n} keys")
    pkg.dist.close_exchange(p)
    p.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    import torch.multiprocessing as mp
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    mp.spawn(worker, args=(world, 29000 + os.getpid() % 2000), nprocs=world, join=True)
