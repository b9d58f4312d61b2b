"""-m gpu: nosplan_score_pass_dev (both kinds in one call: one grouping launch, the two scorers side by side, the
exchange push fused into the scorers' epilogue) against the separate device calls and the oracle, in every flag
combination, with and without an open (one-rank) exchange; the measurement hooks that ride on it."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def _dev(torch, *arrays):
    return [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in arrays]


def _expected(pkg, cfg, o, rows_m, rows_p):
    exp = []
    for kind, rows, width in ((0, rows_m, len(cfg["names"])), (1, rows_p, len(cfg["sizes"]))):
        req, mask = pkg.synth.csr_to_dense(*rows, width)
        if kind == 0:
            _, _, ef = o.score_mig(cfg["mig"], req, mask, threads=8, want_choice=False)
        else:
            _, _, _, ef = o.score_mps(cfg["mps"], req, mask, threads=8, want_free=False)
        exp.append(ef)
    return np.concatenate(exp)


def _run_pass(torch, pkg, p, rows_m, rows_p, flags):
    Dm, Dp = len(rows_m[0]) - 1, len(rows_p[0]) - 1
    tm, tp = _dev(torch, *rows_m), _dev(torch, *rows_p)
    out = torch.full((Dm + Dp,), -7, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    p.score_pass_dev(Dm, tm[0].data_ptr(), tm[1].data_ptr(), tm[2].data_ptr(), Dp, tp[0].data_ptr(), tp[1].data_ptr(),
                     tp[2].data_ptr(), out.data_ptr(), flags, stream)
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize("seed", [21, 22])
def test_pass_equals_oracle_on_random_clusters(seed):
    import torch

    import test_gpu_parity as T
    import walkai_nos_b200 as pkg
    cfg, rand_demands, rng = T._random_cluster(pkg, seed, n_mig=150, n_mps=140)
    p = pkg.Planner(0)
    try:
        o = T._load(p, cfg)
        rows_m = rand_demands(300, len(cfg["names"]))    # rows with 0..3 keys
        rows_p = rand_demands(260, len(cfg["sizes"]))
        exp = _expected(pkg, cfg, o, rows_m, rows_p)
        for flags in (0, pkg.nosplan.PASS_SERIAL):
            assert np.array_equal(_run_pass(torch, pkg, p, rows_m, rows_p, flags), exp), flags
        p.pass_status()
        # single-key rows: every flag combination, then the same through a one-rank exchange (fused push) with a node base
        D = 400
        one = lambda width: (np.arange(D + 1, dtype=np.int32), rng.integers(0, width, size=D).astype(np.int32),
                             rng.choice([1, 1, 2, 3, 7], size=D).astype(np.int32))
        rows_m, rows_p = one(len(cfg["names"])), one(len(cfg["sizes"]))
        exp = _expected(pkg, cfg, o, rows_m, rows_p)
        for flags in (0, 1, 2, 3):
            assert np.array_equal(_run_pass(torch, pkg, p, rows_m, rows_p, flags), exp), flags
        # one kind only
        assert np.array_equal(_run_pass(torch, pkg, p, rows_m, (np.zeros(1, np.int32),) * 3, 1), exp[:D])
        assert np.array_equal(_run_pass(torch, pkg, p, (np.zeros(1, np.int32),) * 3, rows_p, 1), exp[D:])
        p.exchange_open([p.exchange_create(0, 1, 2 * D)])
        p.set_node_base(1000, 5000)
        shift = np.concatenate([np.full(D, 1000), np.full(D, 5000)])
        want = np.where(exp >= 0, exp + shift, -1)
        for rep in range(3):  # the two key halves and the step flags re-arm
            for flags in (1, 3, 0):  # fused push (side by side / serial), then the two-kernel exchange
                assert np.array_equal(_run_pass(torch, pkg, p, rows_m, rows_p, flags), want), (rep, flags)
        p.pass_status()
        p.exchange_destroy()
        p.set_node_base(0, 0)
        # the single-key promise is checked, not trusted
        rows_bad = (np.asarray([0, 1, 3], np.int32), np.asarray([0, 0, 1], np.int32), np.asarray([1, 1, 1], np.int32))
        _run_pass(torch, pkg, p, rows_bad, rows_p, pkg.nosplan.PASS_SINGLE_KEY)
        with pytest.raises(pkg.NosplanError) as e:
            p.pass_status()
        assert "MIG demand row 1" in str(e.value)
        _run_pass(torch, pkg, p, rows_m, rows_p, 1)
        p.pass_status()  # the next pass clears it
    finally:
        p.close()


@pytest.mark.parametrize("cfgno,nodes,pods", [(3, 40, 700), (2, 16, 300)])
def test_status_after_a_pass_over_one_kind_only(cfgno, nodes, pods):
    """A handle that only ever saw one kind of snapshot: the status word of the other kind was never written and must
    read as clean (BASELINE configs[2] is MPS only, configs[1] MIG only)."""
    import torch

    import walkai_nos_b200 as pkg
    cfg = pkg.synth.make_config(cfgno, nodes, pods)
    p = pkg.Planner(0)
    try:
        p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
        none = (np.zeros(1, np.int32),) * 3
        if cfg["mig"] is not None:
            p.set_mig_snapshot(cfg["mig"])
            rows_m, rows_p = pkg.synth.pods_of_kind(cfg["pods"], 0)[:3], none
            req, mask = pkg.synth.csr_to_dense(*rows_m, len(cfg["names"]))
            exp = o.score_mig(cfg["mig"], req, mask, threads=8, want_choice=False)[2]
        else:
            p.set_mps_snapshot(cfg["sizes"], cfg["mps"])
            rows_m, rows_p = none, pkg.synth.pods_of_kind(cfg["pods"], 1)[:3]
            req, mask = pkg.synth.csr_to_dense(*rows_p, len(cfg["sizes"]))
            exp = o.score_mps(cfg["mps"], req, mask, threads=8, want_free=False)[3]
        for with_exchange in (False, True):
            if with_exchange:
                p.exchange_open([p.exchange_create(0, 1, len(exp))])
            for flags in (1, 3, 0):
                assert np.array_equal(_run_pass(torch, pkg, p, rows_m, rows_p, flags), exp), (with_exchange, flags)
                p.pass_status()
        p.exchange_destroy()
    finally:
        p.close()


def test_pass_measurement_hooks_cfg5_shard():
    import torch

    import walkai_nos_b200 as pkg
    cfg = pkg.synth.make_config(5, 256, 3000)
    p = pkg.Planner(0)
    try:
        p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        p.set_mig_snapshot(cfg["mig"])
        p.set_mps_snapshot(cfg["sizes"], cfg["mps"])
        o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
        rows_m = pkg.synth.pods_of_kind(cfg["pods"], 0)[:3]
        rows_p = pkg.synth.pods_of_kind(cfg["pods"], 1)[:3]
        exp = _expected(pkg, cfg, o, rows_m, rows_p)
        p.kernel_timing(True)
        for _ in range(3):  # both kinds share one launch: its events are filed under kind 0
            assert np.array_equal(_run_pass(torch, pkg, p, rows_m, rows_p, pkg.nosplan.PASS_SINGLE_KEY), exp)
        ms, n = p.kernel_time_ms(0)
        assert n == 3 and ms > 0
        assert p.kernel_time_ms(1) == (0.0, 0)
        for _ in range(2):  # the kinds one after the other: a pair of events around each scorer kernel
            assert np.array_equal(_run_pass(torch, pkg, p, rows_m, rows_p, pkg.nosplan.PASS_SINGLE_KEY | pkg.nosplan.PASS_SERIAL), exp)
        for kind in (0, 1):
            ms, n = p.kernel_time_ms(kind)
            assert n == 2 and ms > 0
        assert p.kernel_time_ms(0) == (0.0, 0)
        p.kernel_timing(False)
        # executed evaluations: counted by the measurement instantiation, same winners, never more than the nominal count
        got = _run_pass(torch, pkg, p, rows_m, rows_p, pkg.nosplan.PASS_SINGLE_KEY | pkg.nosplan.PASS_COUNT)
        assert np.array_equal(got, exp)
        em, ep = p.executed_evals()
        Dm, Dp = len(rows_m[0]) - 1, len(rows_p[0]) - 1
        assert 0 < em <= Dm * p.mig_evals_per_demand() and 0 < ep <= Dp * p.mps_evals_per_demand()
        # MPS is easy to count by hand: GPUs of the walk for which the size is still missing (slicing/gpu.go:222-232)
        sizes, mps = cfg["sizes"], cfg["mps"]
        hand = 0
        for d in range(Dp):
            sid, q = int(rows_p[1][d]), int(rows_p[2][d])
            for n_ in range(mps["n_nodes"]):
                res = q
                for g in range(mps["gpu_off"][n_], mps["gpu_off"][n_ + 1]):
                    if res <= 0:
                        break
                    fj = int(mps["free"][g, sid])
                    if res - fj <= 0:
                        break
                    hand += 1
                    # what the GPU holds afterwards comes from the oracle-checked kernel; quantity-1 rows end at the
                    # first GPU that can create or holds a slice, otherwise the walk goes on with the same residual
                    cap = int(mps["mem_gb"][g] - (mps["used"][g] * np.asarray(sizes)).sum())
                    if cap >= sizes[sid]:
                        break
                    res -= fj
        assert ep == hand
        assert p.executed_evals() == (0, 0)
    finally:
        p.close()
