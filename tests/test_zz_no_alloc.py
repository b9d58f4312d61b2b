"""-m gpu, collected last on purpose: the allocation-free scorer entry point.

nosplan_score_mig_static exists for callers that are not allowed to allocate device memory when
they need a score (they work inside a fixed set of buffers, or cannot take the device-wide
synchronisation an allocation implies).  They pass every input and output as a device pointer they
already own; the library's only scratch is a fixed array inside the library image.  This test
drives it through the C ABI and requires the same answers, bit for bit, as the ordinary entry
point nosplan_score_mig — on two calls in a row that share that fixed array, the second on a
smaller cluster so that anything left over from the first call would show up as a wrong answer.
The device buffers are made here by torch; the entry point itself allocates nothing.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run_pair(pkg, planner, torch, cfg_no, nodes, pods):
    cfg = pkg.synth.make_config(cfg_no, nodes, pods)
    planner.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
    mig = cfg["mig"]
    off, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], 0)
    D = len(off) - 1
    # the existing entry point: snapshot uploaded by the library, host buffers
    planner.set_mig_snapshot(mig)
    want = planner.score_mig(off, key, qty, want_updated=False, want_choice=False, want_first_choice=False)["first_node"]
    # the allocation-free entry point: caller-owned device buffers only
    dev = torch.device("cuda", 0)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    d_model, d_used, d_free, d_valid = t(mig["model"]), t(mig["used"]), t(mig["free"]), t(mig["valid"])
    d_off, d_key, d_qty = t(off), t(key), t(qty)
    d_first = torch.full((D,), 12345, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    planner.score_mig_static(int(mig["n_nodes"]), 8, d_valid.data_ptr(), d_model.data_ptr(), d_used.data_ptr(),
                             d_free.data_ptr(), D, d_off.data_ptr(), d_key.data_ptr(), d_qty.data_ptr(), d_first.data_ptr(),
                             torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = d_first.cpu().numpy()
    assert np.array_equal(got, want)  # integer result: exact
    return want


def test_static_scorer_matches_existing_entry_point_twice_in_a_row():
    import torch

    import walkai_nos_b200 as pkg
    planner = pkg.Planner(0)
    try:
        a = _run_pair(pkg, planner, torch, 5, 1024, 3000)  # 512 MIG nodes x 8 GPUs
        b = _run_pair(pkg, planner, torch, 2, 48, 500)     # smaller cluster right after, same fixed array
        assert (a >= 0).any() or (b >= 0).any()
    finally:
        planner.close()


def test_static_scorer_rejects_sizes_beyond_its_fixed_array():
    import torch

    import walkai_nos_b200 as pkg
    planner = pkg.Planner(0)
    try:
        cfg = pkg.synth.make_config(1)
        planner.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        x = torch.zeros(16, dtype=torch.int32, device="cuda:0")
        with pytest.raises(pkg.NosplanError) as e:
            planner.score_mig_static(4097, 8, 0, x.data_ptr(), x.data_ptr(), x.data_ptr(), 1, x.data_ptr(), x.data_ptr(),
                                     x.data_ptr(), x.data_ptr())
        assert e.value.code == -2 and "4096" in str(e.value)
        with pytest.raises(pkg.NosplanError) as e:
            planner.score_mig_static(8, 9, 0, x.data_ptr(), x.data_ptr(), x.data_ptr(), 1, x.data_ptr(), x.data_ptr(),
                                     x.data_ptr(), x.data_ptr())
        assert e.value.code == -2
    finally:
        planner.close()


def test_static_scorer_leaves_the_device_memory_pool_alone():
    """The one promise of the entry point: no device memory is taken or given back by a call.  After a first call
    (which loads the kernels' code and looks the fixed array's address up), the free-memory figure the driver reports
    must be the same before and after further calls."""
    import torch

    import walkai_nos_b200 as pkg
    planner = pkg.Planner(0)
    try:
        cfg = pkg.synth.make_config(5, 512, 800)
        planner.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        mig = cfg["mig"]
        off, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], 0)
        D = len(off) - 1
        dev = torch.device("cuda", 0)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        bufs = [t(mig[k]) for k in ("valid", "model", "used", "free")] + [t(off), t(key), t(qty)]
        first = torch.empty(D, dtype=torch.int32, device=dev)
        stream = torch.cuda.current_stream().cuda_stream
        call = lambda: planner.score_mig_static(int(mig["n_nodes"]), 8, *[b.data_ptr() for b in bufs[:4]], D,
                                                *[b.data_ptr() for b in bufs[4:]], first.data_ptr(), stream)
        call()
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info(0)
        for _ in range(5):
            call()
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info(0)
        assert free1 == free0
    finally:
        planner.close()
