"""-m gpu: the library's own first-node exchange (peer-memory atomicMin + step flags, nosplan_xchg.cuh)
and the node base of the sharded scorer.  One rank runs everywhere; the two-rank case needs two GPUs
and is skipped on a one-GPU box (run it with `gpurun --gpus 2`)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_one_rank_exchange_is_identity_and_rearms():
    import torch

    import walkai_nos_b200 as pkg
    p = pkg.Planner(0)
    try:
        with pytest.raises(pkg.NosplanError):
            p.exchange_min_dev(1, 0, 0)  # not open
        h = p.exchange_create(0, 1, 5000)
        assert len(h) == 64
        p.exchange_open([h])
        rng = np.random.default_rng(5)
        stream = torch.cuda.current_stream().cuda_stream
        for step, n in enumerate((5000, 17, 4096, 1, 5000, 0, 3000)):
            keys = rng.integers(0, 1 << 20, size=max(n, 1)).astype(np.int32)
            keys[rng.random(max(n, 1)) < 0.3] = -1
            d_in = torch.from_numpy(keys).cuda()
            d_out = torch.full((max(n, 1),), -7, dtype=torch.int32, device="cuda")
            p.exchange_min_dev(n, d_in.data_ptr(), d_out.data_ptr(), stream)
            torch.cuda.synchronize()
            assert np.array_equal(d_out.cpu().numpy()[:n], keys[:n]), step
        p.exchange_status()
        with pytest.raises(pkg.NosplanError):
            p.exchange_min_dev(5001, d_in.data_ptr(), d_out.data_ptr(), stream)
        p.exchange_destroy()
        with pytest.raises(pkg.NosplanError):
            p.exchange_min_dev(1, d_in.data_ptr(), d_out.data_ptr(), stream)
    finally:
        p.close()


def test_node_base_makes_dev_scorers_report_global_indices():
    import torch

    import walkai_nos_b200 as pkg
    cfg = pkg.synth.make_config(4, scale_nodes=64, scale_pods=400)
    p = pkg.Planner(0)
    try:
        p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        p.set_mig_snapshot(cfg["mig"])
        p.set_mps_snapshot(cfg["sizes"], cfg["mps"])
        stream = torch.cuda.current_stream().cuda_stream
        res = {}
        for base in ((0, 0), (1000, 70000)):
            p.set_node_base(*base)
            for kind, fn in ((0, p.score_mig_dev), (1, p.score_mps_dev)):
                off, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], kind)
                D = len(off) - 1
                t = [torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (off, key, qty)]
                first = torch.empty(D, dtype=torch.int32, device="cuda")
                if kind == 0:
                    fn(D, t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), first.data_ptr(), 0, 0, stream)
                else:
                    fn(D, t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), first.data_ptr(), 0, 0, 0, stream)
                torch.cuda.synchronize()
                res[(base, kind)] = first.cpu().numpy()
        for kind in (0, 1):
            a, b = res[((0, 0), kind)], res[((1000, 70000), kind)]
            assert (a >= 0).any()
            shift = (1000, 70000)[kind]
            assert np.array_equal(np.where(a >= 0, a + shift, -1), b)
        # the host-buffer scorers are not sharded: local indices whatever the base
        off, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], 0)
        assert np.array_equal(p.score_mig(off, key, qty, want_updated=False, want_choice=False)["first_node"], res[((0, 0), 0)])
        off, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], 1)
        assert np.array_equal(p.score_mps(off, key, qty, want_updated=False, want_free=False)["first_node"], res[((0, 0), 1)])
    finally:
        p.close()


def test_two_ranks_min_over_peer_memory():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "xchg_two_ranks.py")], capture_output=True, text=True,
                       timeout=240)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "exchange ok" in r.stdout


@pytest.mark.parametrize("args", [["5", "512", "6000"], ["4", "96", "900"], ["3", "40", "700"]])
def test_two_ranks_sharded_pass_equals_unsharded_oracle(args):
    """The sharded answer (fused push from the scorers + finish) == the oracle on the whole node list."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sharded_pass_ranks.py"), "2"] + args, capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "sharded pass ok" in r.stdout


def test_presence_counts_follow_the_working_state():
    """nosplan_presence = the sums Controller.profileAlreadyPresent looks at (mig_controller.go:121-144): used + free
    per key over valid nodes, of the snapshot and of what a plan left behind."""
    import walkai_nos_b200 as pkg
    cfg = pkg.synth.make_config(4, scale_nodes=96, scale_pods=600)
    p = pkg.Planner(0)
    try:
        with pytest.raises(pkg.NosplanError):
            p.presence(0)
        p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        mig = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in cfg["mig"].items()}
        mig["valid"][3] = 0
        p.set_mig_snapshot(mig)
        p.set_mps_snapshot(cfg["sizes"], cfg["mps"])

        def expect(snap, used, free):
            v = np.repeat(snap["valid"].astype(bool), np.diff(snap["gpu_off"]))
            return (used + free)[v].sum(axis=0).astype(np.int64)
        assert np.array_equal(p.presence(0), expect(mig, mig["used"], mig["free"]))
        assert np.array_equal(p.presence(1), expect(cfg["mps"], cfg["mps"]["used"], cfg["mps"]["free"]))
        pods = cfg["pods"]
        p.plan(pods["kind"], pods["off"], pods["key"], pods["qty"], gate=False, sched=True)
        u, f = p.get_mig_state()
        assert np.array_equal(p.presence(0), expect(mig, u, f))
        u, f, _ = p.get_mps_state()
        assert np.array_equal(p.presence(1), expect(cfg["mps"], u, f))
    finally:
        p.close()
