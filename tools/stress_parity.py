"""Randomised parity stress outside the test suite: many seeds and sizes of the ragged random clusters of
tests/test_gpu_parity.py (mixed models, out-of-table profiles, invalid nodes, present-but-zero MPS keys,
multi-key rows, quantities 0..9), full per-(row, node, GPU) tables of both scorers against the oracle,
through every output mode, plus the planner in the flag combinations.
usage: python tools/stress_parity.py [first_seed] [n_seeds]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import walkai_nos_b200 as pkg
import test_gpu_parity as T


def run_seed(planner, seed, max_mig=400, max_mps=300, max_rows=700, max_pods=900):
    """One seed: both scorers in every output mode and the planner in every flag combination, all compared
    bit-exactly with the oracle.  Returns (n_mig_nodes, n_mps_nodes, rows, pods)."""
    rng = np.random.default_rng(seed * 7 + 1)
    n_mig, n_mps = int(rng.integers(8, max_mig)), int(rng.integers(8, max_mps))
    cfg, rand_demands, _ = T._random_cluster(pkg, seed, n_mig=n_mig, n_mps=n_mps)
    o = T._load(planner, cfg)
    D = int(rng.integers(1, max_rows))
    off, key, qty = rand_demands(D, len(cfg["names"]))
    T._check_score_mig(pkg, planner, o, cfg, off, key, qty, threads=8)
    off, key, qty = rand_demands(D, len(cfg["sizes"]))
    T._check_score_mps(pkg, planner, o, cfg, off, key, qty, threads=8)
    # single-key rows only (the ordered kernels alone), larger quantities on a few
    for width, check in ((len(cfg["names"]), T._check_score_mig), (len(cfg["sizes"]), T._check_score_mps)):
        k = rng.integers(0, width, size=D).astype(np.int32)
        q = rng.choice([1, 1, 1, 2, 2, 3, 7, 40], size=D).astype(np.int32)
        check(pkg, planner, o, cfg, np.arange(D + 1, dtype=np.int32), k, q, threads=8)
    # the planner: pods drawn from a pool of request vectors; chain-eligible pods (one key, small quantity) and
    # general ones in turn, every flag combination
    M = int(rng.integers(1, max_pods))
    for chainish in (False, True):
        if chainish:
            pool = []
            for width in (len(cfg["names"]), len(cfg["sizes"])):
                k = rng.integers(0, width, size=12).astype(np.int32)
                q = rng.choice([1, 1, 1, 2, 2, 3], size=12).astype(np.int32)
                pool.append((np.arange(13, dtype=np.int32), k, q))
        else:
            pool = [rand_demands(12, len(cfg["names"])), rand_demands(12, len(cfg["sizes"]))]
        kind = rng.integers(0, 2, size=M).astype(np.uint8)
        poff, pkey, pqty = [0], [], []
        for i in range(M):
            po, pk, pq = pool[kind[i]]
            d = int(rng.integers(0, len(po) - 1))
            for e in range(po[d], po[d + 1]):
                pkey.append(int(pk[e]))
                pqty.append(int(pq[e]))
            poff.append(len(pkey))
        pods = dict(kind=kind, off=np.asarray(poff, np.int32), key=np.asarray(pkey if pkey else [0], np.int32),
                    qty=np.asarray(pqty if pqty else [0], np.int32))
        for gate, sched in ((True, True), (True, False), (False, True), (False, False)):
            T._check_plan(pkg, planner, o, cfg, pods, gate, sched)
    return n_mig, n_mps, 4 * D, M


def main():
    s0 = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    ns = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    planner = pkg.Planner(0)
    t0 = time.time()
    rows = 0
    for seed in range(s0, s0 + ns):
        n_mig, n_mps, r, M = run_seed(planner, seed)
        rows += r
        print(f"seed {seed}: {n_mig} MIG + {n_mps} MPS nodes, {r} rows, {M} pods ok", flush=True)
    planner.close()
    print(f"stress ok: {ns} seeds, {rows} rows, {time.time() - t0:.1f} s")


if __name__ == "__main__":
    main()
