"""-m gpu: bit-exact parity of the CUDA path (through the C ABI) against the CPU oracle on the same
seeded inputs: the five BASELINE configs (cfg4/cfg5 also at reduced node counts so the full
per-(demand,node,GPU) tables can be compared), randomised clusters covering the edge cases
(empty/ragged nodes, invalid nodes, mixed models, out-of-table profiles, zero-quantity keys,
present-but-zero MPS keys, multi-profile pods), and size-independent properties at full size."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    import walkai_nos_b200 as p
    return p


@pytest.fixture(scope="module")
def planner(pkg):
    p = pkg.Planner(0)
    yield p
    p.close()


def _load(planner, cfg):
    planner.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
    if cfg["mig"] is not None:
        planner.set_mig_snapshot(cfg["mig"])
    if cfg["mps"] is not None:
        planner.set_mps_snapshot(cfg["sizes"], cfg["mps"])
    return orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])


def _check_score_mig(pkg, planner, o, cfg, off, key, qty, threads=8):
    req, mask = pkg.synth.csr_to_dense(off, key, qty, len(cfg["names"]))
    eu, ec, ef = o.score_mig(cfg["mig"], req, mask, threads=threads)
    r = planner.score_mig(off, key, qty)
    assert np.array_equal(r["updated"], eu)
    assert np.array_equal(r["choice"], ec)
    assert np.array_equal(r["first_node"], ef)
    # the compact output agrees with the full table
    offs = cfg["mig"]["gpu_off"]
    for d in range(len(ef)):
        if ef[d] >= 0:
            g0, g1 = offs[ef[d]], offs[ef[d] + 1]
            assert np.array_equal(r["first_choice"][d, :g1 - g0], ec[d, g0:g1])
    assert r["n_evals"] == (len(off) - 1) * planner.mig_evals_per_demand()


def _check_score_mps(pkg, planner, o, cfg, off, key, qty, threads=8):
    req, mask = pkg.synth.csr_to_dense(off, key, qty, len(cfg["sizes"]))
    eu, efree, emask, ef = o.score_mps(cfg["mps"], req, mask, threads=threads)
    r = planner.score_mps(off, key, qty)
    assert np.array_equal(r["updated"], eu)
    assert np.array_equal(r["first_node"], ef)
    assert np.array_equal(r["out_free_mask"], emask)
    assert np.array_equal(r["out_free"].astype(np.int32), efree)
    # without the full layout table the kernel takes its closed-form single-size path: same answers
    r2 = planner.score_mps(off, key, qty, want_updated=True, want_free=False)
    assert np.array_equal(r2["updated"], eu)
    assert np.array_equal(r2["first_node"], ef)


def _check_plan(pkg, planner, o, cfg, pods, gate, sched):
    W = max(len(cfg["names"]), len(cfg["sizes"]))
    req, mask = pkg.synth.csr_to_dense(pods["off"], pods["key"], pods["qty"], W)
    exp, ms, ps = o.plan(cfg["mig"], cfg["mps"], pods["kind"], req, mask, gate=gate, sched=sched)
    got = planner.plan(pods["kind"], pods["off"], pods["key"], pods["qty"], gate=gate, sched=sched)
    for k in ("gate_hit", "repart_node", "sched_node", "sched_gpu"):
        bad = np.nonzero(got[k] != exp[k])[0]
        assert bad.size == 0, f"{k}: first mismatch at pod {bad[:5]} got {got[k][bad[:5]]} exp {exp[k][bad[:5]]}"
    if cfg["mig"] is not None:
        used, free = planner.get_mig_state()
        assert np.array_equal(used, ms["used"]) and np.array_equal(free, ms["free"])
    if cfg["mps"] is not None:
        used, free, fm = planner.get_mps_state()
        assert np.array_equal(used, ps["used"]) and np.array_equal(free, ps["free"])
        assert np.array_equal(fm, ps["free_mask"])
    return exp


# ---- BASELINE configs ---------------------------------------------------------------------------------
@pytest.mark.parametrize("cfgno,nodes,pods", [(1, 0, 0), (2, 0, 0), (4, 64, 600), (5, 128, 800)])
def test_score_mig_configs(pkg, planner, cfgno, nodes, pods):
    cfg = pkg.synth.make_config(cfgno, nodes, pods)
    o = _load(planner, cfg)
    off, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], 0)
    _check_score_mig(pkg, planner, o, cfg, off, key, qty)


@pytest.mark.parametrize("cfgno,nodes,pods", [(3, 64, 1500), (4, 64, 600), (5, 128, 800)])
def test_score_mps_configs(pkg, planner, cfgno, nodes, pods):
    cfg = pkg.synth.make_config(cfgno, nodes, pods)
    o = _load(planner, cfg)
    off, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], 1)
    _check_score_mps(pkg, planner, o, cfg, off, key, qty)


@pytest.mark.parametrize("gate,sched", [(True, True), (False, True), (False, False), (True, False)])
@pytest.mark.parametrize("cfgno,nodes,pods", [(1, 0, 0), (2, 0, 0), (3, 0, 3000), (4, 256, 4000), (5, 512, 6000)])
def test_plan_configs(pkg, planner, cfgno, nodes, pods, gate, sched):
    cfg = pkg.synth.make_config(cfgno, nodes, pods)
    o = _load(planner, cfg)
    _check_plan(pkg, planner, o, cfg, cfg["pods"], gate, sched)


# ---- randomised clusters ---------------------------------------------------------------------------------
def _random_cluster(pkg, seed, n_mig=37, n_mps=29):
    rng = np.random.default_rng(seed)
    names = pkg.synth.MIG_PROFILES
    tables = pkg.synth.yaml_tables()
    model_names, geom_off, geom_counts = pkg.synth.flatten_tables(tables, names)
    P = len(names)
    # MIG: ragged nodes, mixed models inside a node, out-of-table profiles, invalid nodes
    off = [0]
    model, used, free = [], [], []
    valid = []
    for n in range(n_mig):
        cnt = int(rng.integers(0, 9))
        valid.append(0 if rng.random() < 0.1 else 1)
        for _ in range(cnt):
            m = int(rng.integers(0, len(model_names)))
            u = np.zeros(P, np.int32)
            f = np.zeros(P, np.int32)
            mode = rng.random()
            if mode < 0.6:  # consistent with some allowed geometry
                geoms = tables[model_names[m]]
                g = geoms[int(rng.integers(0, len(geoms)))]
                for k, q in g.items():
                    nu = int(rng.integers(0, q + 1))
                    u[names.index(k)] = nu
                    f[names.index(k)] = q - nu
            elif mode < 0.9:  # arbitrary small counts, possibly profiles the model does not have
                for _ in range(int(rng.integers(0, 4))):
                    p = int(rng.integers(0, P))
                    u[p] = int(rng.integers(0, 3))
                    f[p] = int(rng.integers(0, 4))
            model.append(m)
            used.append(u)
            free.append(f)
        off.append(len(model))
    tg = len(model)
    mig = dict(n_nodes=n_mig, gpu_off=np.asarray(off, np.int32), valid=np.asarray(valid, np.uint8),
               model=np.asarray(model, np.int32), used=np.stack(used), free=np.stack(free))
    mig["used_mask"] = pkg.synth._mask_of(mig["used"])
    mig["free_mask"] = pkg.synth._mask_of(mig["free"])
    # invalid nodes are not interpreted by the library: zero them for the oracle too
    for n in range(n_mig):
        if not valid[n]:
            sl = slice(off[n], off[n + 1])
            mig["used"][sl] = 0
            mig["free"][sl] = 0
            mig["used_mask"][sl] = 0
            mig["free_mask"][sl] = 0
    # MPS
    sizes = sorted(set(int(x) for x in rng.choice(np.arange(1, 41), size=7, replace=False)))
    S = len(sizes)
    off = [0]
    mem, used, free, fmask = [], [], [], []
    valid = []
    for n in range(n_mps):
        cnt = int(rng.integers(0, 9))
        valid.append(0 if rng.random() < 0.1 else 1)
        for _ in range(cnt):
            m = int(rng.integers(8, 100))
            u = np.zeros(S, np.int32)
            f = np.zeros(S, np.int32)
            fm = 0
            tot = 0
            for _ in range(int(rng.integers(0, 6))):
                s = int(rng.integers(0, S))
                if tot + sizes[s] > m:
                    continue
                tot += sizes[s]
                if rng.random() < 0.5:
                    u[s] += 1
                else:
                    f[s] += 1
                    fm |= 1 << s
            for s in range(S):  # present-but-zero keys (left behind by AddPod, slicing/gpu.go:153)
                if rng.random() < 0.15:
                    fm |= 1 << s
            mem.append(m)
            used.append(u)
            free.append(f)
            fmask.append(fm)
        off.append(len(mem))
    tgp = len(mem)
    mps = dict(n_nodes=n_mps, gpu_off=np.asarray(off, np.int32), valid=np.asarray(valid, np.uint8),
               mem_gb=np.asarray(mem, np.int32), used=np.stack(used), free=np.stack(free),
               free_mask=np.asarray(fmask, np.uint32))
    mps["used_mask"] = pkg.synth._mask_of(mps["used"])
    for n in range(n_mps):
        if not valid[n]:
            sl = slice(off[n], off[n + 1])
            for k in ("used", "free", "used_mask", "free_mask", "mem_gb"):
                mps[k][sl] = 0
    cfg = dict(tables=tables, names=names, model_names=model_names, geom_off=geom_off, geom_counts=geom_counts, mig=mig,
               mps=mps, sizes=sizes)

    def rand_demands(D, width):
        off, key, qty = [0], [], []
        for _ in range(D):
            k = int(rng.choice([0, 1, 1, 1, 1, 2, 2, 3]))
            ks = sorted(rng.choice(np.arange(width), size=k, replace=False).tolist())
            for kk in ks:
                key.append(int(kk))
                qty.append(int(rng.choice([0, 1, 1, 1, 2, 2, 3, 5, 9])))
            off.append(len(key))
        return (np.asarray(off, np.int32), np.asarray(key if key else [0], np.int32), np.asarray(qty if qty else [0], np.int32))
    return cfg, rand_demands, rng


@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5, 6])
def test_score_random(pkg, planner, seed):
    cfg, rand_demands, _ = _random_cluster(pkg, seed)
    o = _load(planner, cfg)
    off, key, qty = rand_demands(150, len(cfg["names"]))
    _check_score_mig(pkg, planner, o, cfg, off, key, qty, threads=4)
    off, key, qty = rand_demands(150, len(cfg["sizes"]))
    _check_score_mps(pkg, planner, o, cfg, off, key, qty, threads=4)


@pytest.mark.parametrize("gate,sched", [(True, True), (False, True), (False, False)])
@pytest.mark.parametrize("seed", [11, 12, 13, 14])
def test_plan_random(pkg, planner, seed, gate, sched):
    cfg, rand_demands, rng = _random_cluster(pkg, seed)
    o = _load(planner, cfg)
    M = 400
    # a pool of request vectors per kind, then pods drawn from the pool (so vectors repeat)
    pool = [rand_demands(12, len(cfg["names"])), rand_demands(12, len(cfg["sizes"]))]
    kind = rng.integers(0, 2, size=M).astype(np.uint8)
    off, key, qty = [0], [], []
    for i in range(M):
        po, pk, pq = pool[kind[i]]
        d = int(rng.integers(0, len(po) - 1))
        for e in range(po[d], po[d + 1]):
            key.append(int(pk[e]))
            qty.append(int(pq[e]))
        off.append(len(key))
    pods = dict(kind=kind, off=np.asarray(off, np.int32), key=np.asarray(key if key else [0], np.int32),
                qty=np.asarray(qty if qty else [0], np.int32))
    _check_plan(pkg, planner, o, cfg, pods, gate, sched)


def test_empty_inputs(pkg, planner):
    cfg = pkg.synth.make_config(1)
    _load(planner, cfg)
    z = np.zeros(1, np.int32)
    r = planner.score_mig(z, z, z)
    assert r["first_node"].shape == (0,)
    pr = planner.plan(np.zeros(0, np.uint8), z, z, z)
    assert pr["sched_node"].shape == (0,)
    # a cluster with no nodes at all
    empty = dict(n_nodes=0, gpu_off=np.zeros(1, np.int32), valid=np.zeros(0, np.uint8), model=np.zeros(0, np.int32),
                 used=np.zeros((0, len(cfg["names"])), np.int32), free=np.zeros((0, len(cfg["names"])), np.int32))
    planner.set_mig_snapshot(empty)
    off, key, qty = np.asarray([0, 1], np.int32), np.asarray([0], np.int32), np.asarray([1], np.int32)
    r = planner.score_mig(off, key, qty)
    assert r["first_node"][0] == -1
    pr = planner.plan(np.zeros(1, np.uint8), off, key, qty)
    assert pr["repart_node"][0] == -1 and pr["sched_node"][0] == -1 and pr["gate_hit"][0] == 0


def test_limits_are_rejected(pkg, planner):
    cfg = pkg.synth.make_config(1)
    _load(planner, cfg)
    bad = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in cfg["mig"].items()}
    bad["used"][0, 0] = 300
    with pytest.raises(pkg.NosplanError) as e:
        planner.set_mig_snapshot(bad)
    assert e.value.code == -2
    with pytest.raises(pkg.NosplanError):
        planner.score_mig(np.asarray([0, 1], np.int32), np.asarray([99], np.int32), np.asarray([1], np.int32))


# ---- full size: properties that do not need the oracle's full table -----------------------------------------
def test_full_size_cfg5_properties(pkg, planner):
    cfg = pkg.synth.make_config(5)
    o = _load(planner, cfg)
    pods = cfg["pods"]
    # (1) the scorer on ALL pods: first_node must equal the oracle's on a deduplicated demand set
    off, key, qty, sel = pkg.synth.pods_of_kind(pods, 0)
    r = planner.score_mig(off, key, qty, want_updated=False, want_choice=False, want_first_choice=True)
    sig = key.astype(np.int64) * 100 + qty
    uniq, first_idx = np.unique(sig, return_index=True)
    uo = np.arange(len(uniq) + 1, dtype=np.int32)
    req, mask = pkg.synth.csr_to_dense(uo, key[first_idx], qty[first_idx], len(cfg["names"]))
    eu, ec, ef = o.score_mig(cfg["mig"], req, mask, threads=8)
    for j, s in enumerate(uniq):
        rows = np.nonzero(sig == s)[0]
        assert np.all(r["first_node"][rows] == ef[j])  # identical demand rows give identical answers
        if ef[j] >= 0:
            g0 = cfg["mig"]["gpu_off"][ef[j]]
            assert np.all(r["first_choice"][rows, :8] == ec[j, g0:g0 + 8])
    assert r["n_evals"] == len(sel) * 2048 * 8 * 12
    # (2) plan: idempotent (state is reset from the snapshot), conservation of devices, and the
    # schedulable set equals the oracle's (the oracle finishes this size in seconds with the gate on)
    exp = _check_plan(pkg, planner, o, cfg, pods, True, True)
    again = planner.plan(pods["kind"], pods["off"], pods["key"], pods["qty"])
    assert np.array_equal(again["sched_node"], exp["sched_node"])
    used, free = planner.get_mig_state()
    assert np.array_equal((used + free).sum(axis=0), (cfg["mig"]["used"] + cfg["mig"]["free"]).sum(axis=0))
    # (3) MPS at full size, and the full per-node tables of one row per distinct request (both kinds)
    poff, pkey, pqty, psel = pkg.synth.pods_of_kind(pods, 1)
    rp = planner.score_mps(poff, pkey, pqty, want_free=False)
    psig = pkey.astype(np.int64) * 100 + pqty
    puniq, pfirst = np.unique(psig, return_index=True)
    S = len(cfg["sizes"])
    preq, pmask = pkg.synth.csr_to_dense(np.arange(len(puniq) + 1, dtype=np.int32), pkey[pfirst], pqty[pfirst], S)
    peu, _, _, pef = o.score_mps(cfg["mps"], preq, pmask, threads=8, want_free=False)
    for j, s in enumerate(puniq):
        rows = np.nonzero(psig == s)[0]
        assert np.all(rp["first_node"][rows] == pef[j])
        assert np.array_equal(rp["updated"][rows[0]], peu[j]) and np.array_equal(rp["updated"][rows[-1]], peu[j])
    assert rp["n_evals"] == len(psel) * 2048 * 8
    rm = planner.score_mig(off, key, qty, want_updated=True, want_choice=False, want_first_choice=False)
    for j, s in enumerate(uniq):
        rows = np.nonzero(sig == s)[0]
        assert np.array_equal(rm["updated"][rows[0]], eu[j]) and np.array_equal(rm["updated"][rows[-1]], eu[j])
    # (4) the one-call form agrees with the separate calls at full size
    c32 = lambda *xs: tuple(np.ascontiguousarray(x, np.int32) for x in xs)
    both = planner.score_snapshot(cfg["mig"], c32(off, key, qty), cfg["sizes"], cfg["mps"], c32(poff, pkey, pqty))
    assert np.array_equal(both["mig_first_node"], r["first_node"]) and np.array_equal(both["mig_first_choice"], r["first_choice"])
    assert np.array_equal(both["mps_first_node"], rp["first_node"])


# ---- reference behaviour (gate on) when the gate does not stop everything ------------------------------------------
def _without_profiles(cfg, cols):
    snap = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in cfg["mig"].items()}
    for c in cols:
        snap["used"][:, c] = 0
        snap["free"][:, c] = 0
    snap["used_mask"] = np.zeros_like(snap["used_mask"])
    snap["free_mask"] = np.zeros_like(snap["free_mask"])
    for k in range(snap["used"].shape[1]):
        snap["used_mask"] |= (snap["used"][:, k] != 0).astype(np.uint32) << np.uint32(k)
        snap["free_mask"] |= (snap["free"][:, k] != 0).astype(np.uint32) << np.uint32(k)
    out = dict(cfg)
    out["mig"] = snap
    return out


@pytest.mark.parametrize("sched", [True, False])
@pytest.mark.parametrize("case", ["one_absent", "three_absent", "unreachable", "mps_absent"])
def test_plan_gate_on_with_absent_profiles(pkg, planner, case, sched):
    """Pods whose profile is absent from the whole cluster fall through the gate to tryRepartition
    (mig_controller.go:83-98); everything else stays with the per-profile chains.  One (profile, quantity) per pod."""
    cfg = pkg.synth.make_config(5, 512, 6000)
    names = cfg["names"]
    pods = {k: v.copy() for k, v in cfg["pods"].items()}
    if case == "one_absent":
        cfg = _without_profiles(cfg, [names.index("2g.20gb")])
    elif case == "three_absent":
        cfg = _without_profiles(cfg, [names.index(x) for x in ("1g.10gb", "3g.40gb", "7g.79gb")])
    elif case == "unreachable":
        # a profile no H100 geometry holds: absent, and no node can ever provide it (every such pod: no gate, no node)
        cfg = _without_profiles(cfg, [names.index("4g.40gb")])
        mig_pods = np.nonzero(pods["kind"] == 0)[0]
        pods["key"][pods["off"][mig_pods[::7]]] = names.index("1g.5gb")
    else:
        mps = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in cfg["mps"].items()}
        mps["used"][:, 2] = 0  # no 20 GB slice anywhere (capacity grows accordingly)
        mps["free"][:, 2] = 0
        mps["used_mask"] = pkg.synth._mask_of(mps["used"])
        mps["free_mask"] = pkg.synth._mask_of(mps["free"])
        cfg = dict(cfg)
        cfg["mps"] = mps
    o = _load(planner, cfg)
    exp = _check_plan(pkg, planner, o, cfg, pods, True, sched)
    assert (exp["gate_hit"] == 0).any() and (exp["gate_hit"] == 1).any()
    if case != "unreachable":
        assert (exp["repart_node"] >= 0).any()


def test_plan_gate_on_absent_profile_full_size(pkg, planner):
    cfg = pkg.synth.make_config(5)
    cfg = _without_profiles(cfg, [cfg["names"].index("2g.20gb"), cfg["names"].index("3g.40gb")])
    o = _load(planner, cfg)
    exp = _check_plan(pkg, planner, o, cfg, cfg["pods"], True, True)
    assert (exp["repart_node"] >= 0).sum() >= 1 and (exp["gate_hit"] == 0).any()
