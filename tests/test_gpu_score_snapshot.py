"""-m gpu: nosplan_score_snapshot (one call for a whole scoring pass, MIG and MPS halves side by side on
private lanes) must return exactly what the four separate calls return, leave the handle's installed
snapshots alone, and hand back a half's error with its code."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rows(pkg, pods, kind):
    off, key, qty, _ = pkg.synth.pods_of_kind(pods, kind)
    return tuple(np.ascontiguousarray(x, np.int32) for x in (off, key, qty))


def test_one_call_equals_four_calls():
    import walkai_nos_b200 as pkg
    p = pkg.Planner(0)
    q = pkg.Planner(0)
    try:
        for cfgno, nn, mm in ((4, 300, 3000), (2, 0, 0), (3, 0, 0)):
            cfg = pkg.synth.make_config(cfgno, scale_nodes=nn, scale_pods=mm)
            for pl in (p, q):
                pl.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
            exp = {}
            mr = pr = None
            if cfg["mig"] is not None:
                mr = _rows(pkg, cfg["pods"], 0)
                q.set_mig_snapshot(cfg["mig"])
                r = q.score_mig(*mr, want_updated=False, want_choice=False)
                exp.update(mig_first_node=r["first_node"], mig_first_choice=r["first_choice"], ne=r["n_evals"])
            if cfg["mps"] is not None:
                pr = _rows(pkg, cfg["pods"], 1)
                q.set_mps_snapshot(cfg["sizes"], cfg["mps"])
                r = q.score_mps(*pr, want_updated=False, want_free=False)
                exp.update(mps_first_node=r["first_node"])
                exp["ne"] = exp.get("ne", 0) + r["n_evals"]
            for rep in range(3):  # the lanes are re-used
                got = p.score_snapshot(cfg["mig"], mr, cfg["sizes"], cfg["mps"], pr)
                for k in ("mig_first_node", "mig_first_choice", "mps_first_node"):
                    if k in exp:
                        assert np.array_equal(got[k], exp[k]), (cfgno, k, rep)
                assert got["n_evals"] == exp["ne"]
        # nothing was installed on p: it still has no snapshot of its own
        with pytest.raises(pkg.NosplanError):
            p.score_mig(np.asarray([0, 1], np.int32), np.asarray([0], np.int32), np.asarray([1], np.int32))
    finally:
        p.close()
        q.close()


def test_errors_of_a_half_come_back():
    import walkai_nos_b200 as pkg
    cfg = pkg.synth.make_config(4, scale_nodes=64, scale_pods=200)
    p = pkg.Planner(0)
    try:
        mr, pr = _rows(pkg, cfg["pods"], 0), _rows(pkg, cfg["pods"], 1)
        with pytest.raises(pkg.NosplanError):  # tables first
            p.score_snapshot(cfg["mig"], mr, cfg["sizes"], cfg["mps"], pr)
        p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        bad = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in cfg["mps"].items()}
        bad["mem_gb"][5] = -1
        with pytest.raises(pkg.NosplanError) as e:
            p.score_snapshot(cfg["mig"], mr, cfg["sizes"], bad, pr)
        assert e.value.code == -2 and "MPS: GPU 5: memory -1 GB out of range" in str(e.value)
        badrows = (mr[0], mr[1].copy(), mr[2])
        badrows[1][3] = 77
        with pytest.raises(pkg.NosplanError) as e:
            p.score_snapshot(cfg["mig"], badrows, cfg["sizes"], cfg["mps"], pr)
        assert e.value.code == -2 and "MIG: demand row 3: key 77 out of range" in str(e.value)
        got = p.score_snapshot(cfg["mig"], mr, cfg["sizes"], cfg["mps"], pr)  # and it still works
        assert (got["mig_first_node"] >= -1).all() and (got["mps_first_node"] >= -1).all()
        only = p.score_snapshot(None, None, cfg["sizes"], cfg["mps"], pr)
        assert np.array_equal(only["mps_first_node"], got["mps_first_node"]) and "mig_first_node" not in only
    finally:
        p.close()
