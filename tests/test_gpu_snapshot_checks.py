"""-m gpu: the snapshot range checks run inside the pack kernels (the host only walks gpu_off).  Each
error class must come back through the C ABI with the code and the message of the FIRST bad entry in
GPU order, rows of invalid nodes must not be interpreted, and a rejected snapshot leaves none behind."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture()
def planner():
    import walkai_nos_b200 as pkg
    p = pkg.Planner(0)
    yield p
    p.close()


def _mig(pkg, planner):
    cfg = pkg.synth.make_config(2)
    planner.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
    snap = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in cfg["mig"].items()}
    return cfg, snap


def _mps(pkg):
    cfg = pkg.synth.make_config(3)
    snap = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in cfg["mps"].items()}
    return cfg, snap


def test_mig_first_bad_entry_in_gpu_order(planner):
    import walkai_nos_b200 as pkg
    cfg, snap = _mig(pkg, planner)
    P = len(cfg["names"])
    snap["used"][400, P - 1] = 999          # later GPU
    snap["free"][37, 5] = -1                # first bad GPU, later profile
    snap["used"][37, 2] = 200
    snap["free"][37, 2] = 100               # used + free > 255, earliest profile of GPU 37
    with pytest.raises(pkg.NosplanError) as e:
        planner.set_mig_snapshot(snap)
    assert e.value.code == -2
    assert f"GPU 37 profile {cfg['names'][2]}: used 200 free 100" in str(e.value)
    # the model check of a GPU comes before its counts
    snap["model"][37] = 99
    with pytest.raises(pkg.NosplanError) as e:
        planner.set_mig_snapshot(snap)
    assert "GPU 37: model 99 unknown" in str(e.value)
    snap["model"][12] = -1
    with pytest.raises(pkg.NosplanError) as e:
        planner.set_mig_snapshot(snap)
    assert "GPU 12: model -1 unknown" in str(e.value)
    # a rejected snapshot leaves none behind
    with pytest.raises(pkg.NosplanError):
        planner.score_mig(np.asarray([0, 1], np.int32), np.asarray([0], np.int32), np.asarray([1], np.int32))


def test_mig_invalid_nodes_are_not_interpreted(planner):
    import walkai_nos_b200 as pkg
    cfg, snap = _mig(pkg, planner)
    good = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in snap.items()}
    planner.set_mig_snapshot(good)
    off, key, qty = np.asarray([0, 1, 3], np.int32), np.asarray([1, 0, 3], np.int32), np.asarray([1, 2, 1], np.int32)
    base = planner.score_mig(off, key, qty)
    n_bad = 5
    g0, g1 = snap["gpu_off"][n_bad], snap["gpu_off"][n_bad + 1]
    snap["valid"][n_bad] = 0
    snap["model"][g0:g1] = 1234
    snap["used"][g0:g1] = -7
    snap["free"][g0:g1] = 100000
    planner.set_mig_snapshot(snap)
    r = planner.score_mig(off, key, qty)
    keep = np.arange(snap["n_nodes"]) != n_bad
    assert np.array_equal(r["updated"][:, keep], base["updated"][:, keep])
    assert not r["updated"][:, n_bad].any()
    geoms = np.diff(cfg["geom_off"])[good["model"][g0:g1]].sum()
    assert r["n_evals"] == base["n_evals"] - (len(off) - 1) * int(geoms)
    used, free = planner.get_mig_state()
    assert not used[g0:g1].any() and not free[g0:g1].any()
    assert np.array_equal(used[:g0], good["used"][:g0]) and np.array_equal(free[g1:], good["free"][g1:])


def test_mps_error_classes(planner):
    import walkai_nos_b200 as pkg
    cfg, snap = _mps(pkg)
    sizes = cfg["sizes"]
    planner.set_mps_snapshot(sizes, snap)
    evals = planner.mps_evals_per_demand()
    assert evals == int(snap["gpu_off"][-1])

    def rejected(mut, code, text):
        s = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in snap.items()}
        mut(s)
        with pytest.raises(pkg.NosplanError) as e:
            planner.set_mps_snapshot(sizes, s)
        assert e.value.code == code, str(e.value)
        assert text in str(e.value), str(e.value)
        with pytest.raises(pkg.NosplanError):  # no snapshot left behind
            planner.score_mps(np.asarray([0, 1], np.int32), np.asarray([0], np.int32), np.asarray([1], np.int32))

    def mem_range(s):
        s["mem_gb"][9] = -3
        s["mem_gb"][700] = 1 << 20
    rejected(mem_range, -2, "GPU 9: memory -3 GB out of range")

    def count_range(s):
        s["used"][20, 1] = 256
        s["free"][20, 0] = -1
        s["used"][33, 0] = 300
    rejected(count_range, -2, f"GPU 20 size {sizes[0]}gb: used {snap['used'][20, 0]} free -1 out of range")

    def mask_missing(s):
        g = int(np.nonzero(s["free"].sum(axis=1) > 0)[0][3])
        k = int(np.nonzero(s["free"][g])[0][0])
        s["free_mask"][g] &= ~np.uint32(1 << k)
        mask_missing.g = g
    s_probe = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in snap.items()}
    mask_missing(s_probe)
    rejected(mask_missing, -1, f"GPU {mask_missing.g}: free_mask misses a key")

    def too_much(s):
        s["mem_gb"][50] = 3
        s["used"][50] = 0
        s["free"][50] = 0
        s["used"][50, len(sizes) - 1] = 1
    rejected(too_much, -2, f"GPU 50: slices total {sizes[-1]} GB exceed memory 3 GB")

    # an invalid node hides all of it
    s = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in snap.items()}
    n_bad = 2
    g0, g1 = s["gpu_off"][n_bad], s["gpu_off"][n_bad + 1]
    s["valid"][n_bad] = 0
    s["mem_gb"][g0:g1] = -1
    s["used"][g0:g1] = 1000
    planner.set_mps_snapshot(sizes, s)
    assert planner.mps_evals_per_demand() == evals - (g1 - g0)
    used, free, _ = planner.get_mps_state()
    assert not used[g0:g1].any() and not free[g0:g1].any()


def test_demand_rows_are_checked_on_the_device(planner):
    """The scorers validate the demand rows while grouping them (rows_group_kernel): same codes and messages as
    the host walk, the first bad row in row order wins, and a bad row never reaches a scorer."""
    import walkai_nos_b200 as pkg
    cfg, snap = _mig(pkg, planner)
    planner.set_mig_snapshot(snap)
    mcfg, msnap = _mps(pkg)
    planner.set_mps_snapshot(mcfg["sizes"], msnap)
    P, S = len(cfg["names"]), len(mcfg["sizes"])
    good_off = np.arange(0, 41, dtype=np.int32)
    good_key = (np.arange(40) % 6).astype(np.int32)
    good_qty = np.ones(40, np.int32)

    def both(off, key, qty, code, text, kinds=(0, 1)):
        for kind in kinds:
            k = key
            fn = planner.score_mig if kind == 0 else planner.score_mps
            kw = dict(want_updated=False, want_choice=False) if kind == 0 else dict(want_updated=False, want_free=False)
            with pytest.raises(pkg.NosplanError) as e:
                fn(off, k, qty, **kw)
            assert e.value.code == code, (kind, str(e.value))
            assert text in str(e.value), (kind, str(e.value))

    # key out of range (RANGE), later rows are bad too: the first one is reported
    key = good_key.copy(); key[7] = 99; key[30] = -4
    both(good_off, key, good_qty, -2, "demand row 7: key 99 out of range", kinds=(0,))
    key = good_key.copy() % S; key[7] = 99
    both(good_off, key, good_qty, -2, "demand row 7: key 99 out of range", kinds=(1,))
    # quantity out of range
    qty = good_qty.copy(); qty[3] = -1; qty[9] = 2_000_000
    both(good_off, good_key % S, qty, -2, "demand row 3: quantity -1 out of range")
    qty = good_qty.copy(); qty[9] = 2_000_000
    both(good_off, good_key % S, qty, -2, "demand row 9: quantity 2000000 out of range")
    # keys of a row must be strictly ascending (ARG)
    off = np.asarray([0, 1, 4, 5], np.int32)
    key = np.asarray([0, 1, 3, 3, 2], np.int32)
    qty = np.ones(5, np.int32)
    both(off, key, qty, -1, "demand row 1: keys must be strictly ascending")
    # too many keys in a row (RANGE)
    off = np.asarray([0, 1, 10], np.int32)
    both(off, np.arange(10, dtype=np.int32) % 3, np.ones(10, np.int32), -2, "demand row 1 has 9 keys")
    # dem_off going backwards (ARG)
    off = np.asarray([0, 3, 2, 4], np.int32)
    both(off, np.asarray([0, 1, 2, 3], np.int32), np.ones(4, np.int32), -1, "dem_off not ascending at row 1")
    # and the handle still scores good rows afterwards, identically to before
    r1 = planner.score_mig(good_off, good_key, good_qty, want_updated=False, want_choice=False)
    r2 = planner.score_mig(good_off, good_key, good_qty, want_updated=False, want_choice=False)
    assert np.array_equal(r1["first_node"], r2["first_node"]) and (r1["first_node"] >= -1).all()
