"""-m gpu: parity at the documented limits of the ABI — 16 profiles, 32 geometries per model (the generic
candidate-word path), duplicate geometries in a table (index tie-break), 16 GPUs per node, counts up to
the byte range, 32 MPS sizes, 8 keys per request, large quantities."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def _mig_case(pkg, rng, n_models, geoms_per_model, n_nodes, max_gpus, big_counts):
    names = [f"{i + 1}g.{(i + 1) * 5}gb" for i in range(16)]
    names = sorted(names)
    tables = {}
    for m in range(n_models):
        geoms = []
        for c in range(geoms_per_model):
            g = {}
            for p in rng.choice(16, size=int(rng.integers(1, 6)), replace=False):
                g[names[int(p)]] = int(rng.integers(1, 60 if big_counts else 8))
            geoms.append(g)
        if geoms_per_model > 3:
            geoms[3] = dict(geoms[1])  # a duplicate geometry: equal Id, the lower index must win
        tables[f"model-{m}"] = geoms
    model_names, geom_off, geom_counts = pkg.synth.flatten_tables(tables, names)
    off, model, used, free, valid = [0], [], [], [], []
    for n in range(n_nodes):
        cnt = int(rng.integers(0, max_gpus + 1))
        valid.append(1)
        for _ in range(cnt):
            m = int(rng.integers(0, n_models))
            u = np.zeros(16, np.int32)
            f = np.zeros(16, np.int32)
            if rng.random() < 0.7:
                g = tables[model_names[m]][int(rng.integers(0, geoms_per_model))]
                for k, q in g.items():
                    nu = int(rng.integers(0, q + 1))
                    u[names.index(k)] = nu
                    f[names.index(k)] = q - nu
            else:
                for _ in range(3):
                    p = int(rng.integers(0, 16))
                    u[p] = int(rng.integers(0, 100 if big_counts else 3))
                    f[p] = int(rng.integers(0, 150 if big_counts else 3))
            model.append(m)
            used.append(u)
            free.append(f)
        off.append(len(model))
    mig = dict(n_nodes=n_nodes, gpu_off=np.asarray(off, np.int32), valid=np.asarray(valid, np.uint8),
               model=np.asarray(model, np.int32), used=np.stack(used), free=np.stack(free))
    mig["used_mask"] = pkg.synth._mask_of(mig["used"])
    mig["free_mask"] = pkg.synth._mask_of(mig["free"])
    return dict(tables=tables, names=names, geom_off=geom_off, geom_counts=geom_counts, mig=mig)


def _demands(rng, D, width, kmax, qmax):
    off, key, qty = [0], [], []
    for _ in range(D):
        k = int(rng.integers(1, kmax + 1))
        for kk in sorted(rng.choice(width, size=k, replace=False).tolist()):
            key.append(int(kk))
            qty.append(int(rng.integers(0, qmax + 1)))
        off.append(len(key))
    return np.asarray(off, np.int32), np.asarray(key, np.int32), np.asarray(qty, np.int32)


@pytest.mark.parametrize("geoms,big", [(32, False), (32, True), (20, False), (7, True)])
def test_mig_limits(geoms, big):
    import walkai_nos_b200 as pkg
    rng = np.random.default_rng(100 + geoms + int(big))
    cfg = _mig_case(pkg, rng, n_models=3, geoms_per_model=geoms, n_nodes=70, max_gpus=16, big_counts=big)
    p = pkg.Planner(0)
    try:
        p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        p.set_mig_snapshot(cfg["mig"])
        o = orc.Oracle(cfg["tables"], extra_mig_profiles=cfg["names"], builtin_profiles=False)
        assert o.mig_names == cfg["names"]
        for kmax, qmax in ((1, 3), (8, 300 if big else 6)):
            off, key, qty = _demands(rng, 120, 16, kmax, qmax)
            req, mask = pkg.synth.csr_to_dense(off, key, qty, 16)
            eu, ec, ef = o.score_mig(cfg["mig"], req, mask, threads=4)
            r = p.score_mig(off, key, qty)
            assert np.array_equal(r["updated"], eu)
            assert np.array_equal(r["choice"], ec)
            assert np.array_equal(r["first_node"], ef)
            kind = np.zeros(len(off) - 1, np.uint8)
            for gate in (True, False):
                exp, ms, _ = o.plan(cfg["mig"], None, kind, req, mask, gate=gate, sched=True)
                got = p.plan(kind, off, key, qty, gate=gate, sched=True)
                for kk in exp:
                    assert np.array_equal(got[kk], exp[kk]), (kk, gate)
                used, free = p.get_mig_state()
                assert np.array_equal(used, ms["used"]) and np.array_equal(free, ms["free"])
    finally:
        p.close()


def test_mps_limits():
    import walkai_nos_b200 as pkg
    rng = np.random.default_rng(77)
    sizes = sorted(set(int(x) for x in rng.choice(np.arange(1, 200), size=32, replace=False)))
    S = len(sizes)
    off, mem, used, free, fmask = [0], [], [], [], []
    for n in range(60):
        for _ in range(int(rng.integers(0, 17))):
            m = int(rng.integers(50, 2000))
            u = np.zeros(S, np.int32)
            f = np.zeros(S, np.int32)
            fm, tot = 0, 0
            for _ in range(int(rng.integers(0, 12))):
                s = int(rng.integers(0, S))
                if tot + sizes[s] > m or u[s] + f[s] >= 200:
                    continue
                tot += sizes[s]
                if rng.random() < 0.5:
                    u[s] += 1
                else:
                    f[s] += 1
                    fm |= 1 << s
            for s in range(S):
                if rng.random() < 0.1:
                    fm |= 1 << s
            mem.append(m)
            used.append(u)
            free.append(f)
            fmask.append(fm)
        off.append(len(mem))
    mps = dict(n_nodes=60, gpu_off=np.asarray(off, np.int32), valid=np.ones(60, np.uint8), mem_gb=np.asarray(mem, np.int32),
               used=np.stack(used), free=np.stack(free), free_mask=np.asarray(fmask, np.uint32))
    mps["used_mask"] = pkg.synth._mask_of(mps["used"])
    p = pkg.Planner(0)
    try:
        p.set_mps_snapshot(sizes, mps)
        o = orc.Oracle(pkg.synth.yaml_tables(), mps_sizes=sizes)
        for kmax, qmax in ((1, 4), (8, 40)):
            doff, key, qty = _demands(rng, 150, S, kmax, qmax)
            req, mask = pkg.synth.csr_to_dense(doff, key, qty, S)
            eu, efree, emask, ef = o.score_mps(mps, req, mask, threads=4)
            r = p.score_mps(doff, key, qty)
            assert np.array_equal(r["updated"], eu) and np.array_equal(r["first_node"], ef)
            assert np.array_equal(r["out_free_mask"], emask) and np.array_equal(r["out_free"].astype(np.int32), efree)
            r2 = p.score_mps(doff, key, qty, want_free=False)
            assert np.array_equal(r2["updated"], eu) and np.array_equal(r2["first_node"], ef)
            kind = np.ones(len(doff) - 1, np.uint8)
            W = max(S, 13)
            reqw, maskw = pkg.synth.csr_to_dense(doff, key, qty, W)
            for gate in (True, False):
                exp, _, ps = o.plan(None, mps, kind, reqw, maskw, gate=gate, sched=True)
                got = p.plan(kind, doff, key, qty, gate=gate, sched=True)
                for kk in exp:
                    assert np.array_equal(got[kk], exp[kk]), (kk, gate)
                u2, f2, fm2 = p.get_mps_state()
                assert np.array_equal(u2, ps["used"]) and np.array_equal(f2, ps["free"]) and np.array_equal(fm2, ps["free_mask"])
    finally:
        p.close()


def test_huge_quantities_behave_like_uncoverable_ones():
    """The scorers clamp a requested quantity (MIG: 32767) because nothing beyond 16 GPUs x 255 devices can be
    covered; the oracle does no such thing, so the outputs must still agree up to NOSPLAN_MAX_DEMAND_QTY."""
    import walkai_nos_b200 as pkg
    cfg = pkg.synth.make_config(4, scale_nodes=96, scale_pods=10)
    p = pkg.Planner(0)
    try:
        p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        p.set_mig_snapshot(cfg["mig"])
        p.set_mps_snapshot(cfg["sizes"], cfg["mps"])
        o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
        qs = [1, 2, 7, 255, 256, 4079, 4080, 4081, 32767, 32768, 65535, 65536, 999999, 1000000]
        h100 = [cfg["names"].index(x) for x in ["1g.10gb", "2g.20gb", "3g.40gb", "7g.79gb"]]
        key = np.asarray([k for k in h100 for _ in qs], np.int32)
        qty = np.asarray(qs * len(h100), np.int32)
        off = np.arange(len(key) + 1, dtype=np.int32)
        req, mask = pkg.synth.csr_to_dense(off, key, qty, len(cfg["names"]))
        eu, ec, ef = o.score_mig(cfg["mig"], req, mask, threads=4)
        r = p.score_mig(off, key, qty)
        assert np.array_equal(r["updated"], eu) and np.array_equal(r["choice"], ec) and np.array_equal(r["first_node"], ef)
        r = p.score_mig(off, key, qty, want_updated=False, want_choice=False)
        assert np.array_equal(r["first_node"], ef)
        S = len(cfg["sizes"])
        key = np.asarray([s for s in range(S) for _ in qs], np.int32)
        qty = np.asarray(qs * S, np.int32)
        off = np.arange(len(key) + 1, dtype=np.int32)
        req, mask = pkg.synth.csr_to_dense(off, key, qty, S)
        eu, efree, emask, ef = o.score_mps(cfg["mps"], req, mask, threads=4)
        r = p.score_mps(off, key, qty)
        assert np.array_equal(r["updated"], eu) and np.array_equal(r["first_node"], ef)
        assert np.array_equal(r["out_free_mask"], emask) and np.array_equal(r["out_free"].astype(np.int32), efree)
        r = p.score_mps(off, key, qty, want_free=False)
        assert np.array_equal(r["updated"], eu) and np.array_equal(r["first_node"], ef)
    finally:
        p.close()
