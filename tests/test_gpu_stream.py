"""-m gpu: the streaming MIG scorer (mig_stream_kernel: few rows, one thread per node in storage order, no CTA
barrier) against the oracle.  It is what nosplan_score_mig / _dev run when a call has at most 96 rows, wants the
winners only and no node has more than 8 GPUs; the tests below stay inside that rule and check it really was taken
by comparing with a child process that has the path switched off (NOSPLAN_STREAM_ROWS=0)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _first(planner, off, key, qty):
    return planner.score_mig(off, key, qty, want_updated=False, want_choice=False, want_first_choice=True)


@pytest.mark.parametrize("seed", [41, 42, 43, 44])
def test_stream_random_clusters(seed):
    import test_gpu_parity as T
    import walkai_nos_b200 as pkg
    cfg, rand_demands, rng = T._random_cluster(pkg, seed, n_mig=int(50 + 137 * (seed - 40)), n_mps=8)
    p = pkg.Planner(0)
    try:
        o = T._load(p, cfg)
        for D in (1, 2, 5, 13, 24, 77):
            # single-key rows (the streaming kernel) and rows with 0..3 keys (the general kernel takes those of the call)
            k = rng.integers(0, len(cfg["names"]), size=D).astype(np.int32)
            q = rng.choice([1, 1, 2, 3, 7, 40, 300], size=D).astype(np.int32)
            for off, key, qty in ((np.arange(D + 1, dtype=np.int32), k, q), rand_demands(D, len(cfg["names"]))):
                req, mask = pkg.synth.csr_to_dense(off, key, qty, len(cfg["names"]))
                eu, ec, ef = o.score_mig(cfg["mig"], req, mask, threads=4)
                r = _first(p, off, key, qty)
                assert np.array_equal(r["first_node"], ef), (D, r["first_node"], ef)
                offs = cfg["mig"]["gpu_off"]
                for d in range(D):
                    if ef[d] >= 0:
                        g0, g1 = offs[ef[d]], offs[ef[d] + 1]
                        assert np.array_equal(r["first_choice"][d, :g1 - g0], ec[d, g0:g1])
    finally:
        p.close()


@pytest.mark.parametrize("cfgno,nodes", [(1, 0), (2, 0), (5, 4096), (5, 60000)])
def test_stream_baseline_tables(cfgno, nodes):
    import walkai_nos_b200 as pkg
    cfg = pkg.synth.make_config(cfgno, nodes, 200)
    p = pkg.Planner(0)
    try:
        p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        p.set_mig_snapshot(cfg["mig"])
        o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
        h100 = [cfg["names"].index(x) for x in ["1g.10gb", "2g.20gb", "3g.20gb", "3g.40gb", "4g.40gb", "7g.79gb"]]
        key = np.asarray(h100 * 4, np.int32)
        qty = np.asarray([1] * 6 + [2] * 6 + [5] * 6 + [60] * 6, np.int32)
        off = np.arange(25, dtype=np.int32)
        req, mask = pkg.synth.csr_to_dense(off, key, qty, len(cfg["names"]))
        _, _, ef = o.score_mig(cfg["mig"], req, mask, threads=8, want_choice=False)
        assert np.array_equal(_first(p, off, key, qty)["first_node"], ef)
        for D in (1, 7):
            assert np.array_equal(_first(p, off[:D + 1], key[:D], qty[:D])["first_node"], ef[:D])
        # device-resident form with a node base
        import torch
        t = [torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (off, key, qty)]
        first = torch.empty(24, dtype=torch.int32, device="cuda")
        p.set_node_base(1000, 0)
        p.score_mig_dev(24, t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), first.data_ptr(), 0, 0, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert np.array_equal(first.cpu().numpy(), np.where(ef >= 0, ef + 1000, -1))
    finally:
        p.close()


def test_stream_path_is_the_one_that_ran():
    """Same call with the path switched off in a child process: same winners, and the launch really differs
    (the child reports which kernels its library would pick through the timing of one large call is not needed:
    the knob is read by mig_streaming() only)."""
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r); import walkai_nos_b200 as pkg\n"
        "cfg = pkg.synth.make_config(5, 2048, 50); p = pkg.Planner(0)\n"
        "p.set_mig_tables(cfg['names'], cfg['geom_off'], cfg['geom_counts']); p.set_mig_snapshot(cfg['mig'])\n"
        "off, key, qty, _ = pkg.synth.pods_of_kind(cfg['pods'], 0); D = min(20, len(off) - 1)\n"
        "r = p.score_mig(off[:D + 1], key[:D], qty[:D], want_updated=False, want_choice=False)\n"
        "print('WIN', r['first_node'].tolist())\n" % ROOT)
    outs = []
    for knob in ("96", "0"):
        env = dict(os.environ, NOSPLAN_STREAM_ROWS=knob)
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append([l for l in r.stdout.splitlines() if l.startswith("WIN")][0])
    assert outs[0] == outs[1]
