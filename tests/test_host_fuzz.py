"""No GPU: adversarial and random inputs for libnoshost's string parsers (the wire format of pkg/gpu/annotation.go:29-224,
the resource-name scanners of pkg/gpu/mig/util.go and slicing/profile.go, the memory label of pkg/gpu/util.go:59-73, the
geometry-table loader).  The library must answer every input with a value or an error code — never crash, never write
past the caller's buffer — and well-formed keys must survive a format -> parse round trip.  tools/run_host_sanitizers.sh
runs this file against the AddressSanitizer + UBSan build as well."""
import ctypes as C
import random
import string

import pytest

import walkai_nos_b200 as pkg

host = pkg.host
PREFIX_STATUS = "nos.nebuly.com/status-gpu-"
PREFIX_SPEC = "nos.nebuly.com/spec-gpu-"
ALPHABET = string.ascii_letters + string.digits + "-._/: \t\"'[]{}#,\\%+" + "é中"


def _rand(rng, n):
    return "".join(rng.choice(ALPHABET) for _ in range(n))


def _mutations(rng, s):
    yield s
    for _ in range(6):
        i = rng.randrange(len(s) + 1)
        yield s[:i] + _rand(rng, rng.randrange(1, 4)) + s[i:]       # insertion
        if s:
            j = rng.randrange(len(s))
            yield s[:j] + s[j + 1:]                                  # deletion
            yield s[:j]                                              # truncation
    yield s * 3
    yield s + "9" * 40                                               # digits that overflow an int


def test_annotation_parsers_survive_garbage():
    rng = random.Random(20261107)
    seeds = [PREFIX_STATUS + "0-1g.10gb-used", PREFIX_STATUS + "12-7g.79gb-free", PREFIX_SPEC + "3-2g.20gb", PREFIX_STATUS + "1-10gb-used",
             PREFIX_SPEC + "0-40gb", PREFIX_STATUS, PREFIX_SPEC, "", "-", "nos.nebuly.com/", PREFIX_STATUS + "-1-1g.5gb-used",
             PREFIX_STATUS + "0--used", PREFIX_STATUS + "99999999999999999999-1g.5gb-free", PREFIX_SPEC + "0-" + "x" * 400]
    values = ["1", "0", "-1", "", " 1", "1 ", "+3", "007", "2147483647", "2147483648", "99999999999999999999", "1e3", "0x10", "one", "١"]
    n = 0
    for seed in seeds:
        for key in _mutations(rng, seed):
            for value in values + [_rand(rng, rng.randrange(0, 12))]:
                for parse in (host.parse_status_annotation, host.parse_spec_annotation):
                    r = parse(key, value)
                    n += 1
                    if r is not None:  # accepted (strconv.Atoi takes signs, so negative numbers parse, annotation.go:43-50):
                        # the numbers are what the text says and the profile fits the caller's buffer
                        assert int(value) == r["quantity"] and -2**31 <= r["index"] < 2**31 and len(r["profile"]) < 256
    assert n > 5000


def test_small_output_buffers_are_respected():
    L = host.lib()
    idx, st, q = C.c_int32(), C.c_int32(), C.c_int32()
    key = (PREFIX_STATUS + "0-" + "p" * 100 + "-used").encode()
    for cap in (0, 1, 2, 8, 100, 101, 102):
        buf = C.create_string_buffer(b"\xaa" * 128, 128)
        rc = L.noshost_parse_status_annotation(key, b"1", C.byref(idx), buf, cap, C.byref(st), C.byref(q))
        assert buf.raw[cap:] == b"\xaa" * (128 - cap), cap          # nothing written past the capacity given
        if cap <= 100:
            assert rc == 2 and buf.raw == b"\xaa" * 128, cap          # a name that does not fit is refused, not truncated
        else:
            assert rc == 0 and buf.raw[:cap].find(b"\0") == 100, cap  # the 100-byte name and its terminator
    for cap in (0, 1, 5, 30):
        buf = C.create_string_buffer(b"\xaa" * 128, 128)
        L.noshost_format_status_annotation(7, b"1g.10gb", 1, buf, cap)
        assert buf.raw[cap:] == b"\xaa" * (128 - cap), cap
        buf = C.create_string_buffer(b"\xaa" * 128, 128)
        L.noshost_format_spec_annotation(7, b"1g.10gb", buf, cap)
        assert buf.raw[cap:] == b"\xaa" * (128 - cap), cap
        buf = C.create_string_buffer(b"\xaa" * 128, 128)
        rc = L.noshost_mig_profile_of_resource(b"nvidia.com/mig-1g.10gb", buf, cap)
        assert buf.raw[cap:] == b"\xaa" * (128 - cap) and rc == (0 if cap > 7 else 2), cap
        buf = C.create_string_buffer(b"\xaa" * 128, 128)
        rc = L.noshost_mps_profile_of_resource(b"nvidia.com/gpu-10gb", buf, cap)
        assert buf.raw[cap:] == b"\xaa" * (128 - cap) and rc == (0 if cap > 4 else 2), cap


def test_format_parse_round_trip():
    rng = random.Random(7)
    for _ in range(500):
        index = rng.randrange(0, 64)
        profile = rng.choice(pkg.synth.MIG_PROFILES + ["5gb", "10gb", "40gb"])
        status = rng.choice(["used", "free"])
        q = rng.randrange(0, 50)
        k = host.format_status_annotation(index, profile, status)
        assert host.parse_status_annotation(k, str(q)) == dict(index=index, profile=profile, status=status, quantity=q)
        k = host.format_spec_annotation(index, profile)
        assert host.parse_spec_annotation(k, str(q)) == dict(index=index, profile=profile, quantity=q)


def test_scalar_scanners_survive_garbage():
    rng = random.Random(11)
    seeds = ["nvidia.com/mig-1g.10gb", "nvidia.com/gpu-10gb", "nvidia.com/gpu", "nvidia.com/mig-", "nvidia.com/gpu-gb", "40960", "81920", "1gb",
             "999999999999gb", "-5gb", "0gb", "", "gb"]
    n = 0
    for seed in seeds:
        for s in _mutations(rng, seed):
            r = host.mig_profile_of_resource(s)
            assert r is None or (len(r) < 256 and s.endswith(r))
            r = host.mps_profile_of_resource(s)
            assert r is None or (len(r) < 256 and s.endswith(r))
            assert -2**31 <= host.mps_profile_size_gb(s) < 2**31   # Atoi of what is between the prefix and "gb", else 0
                                                                   # (slicing/profile.go:52-59; "-5gb" is -5 there too)
            g = host.memory_gb(s)
            assert g is None or g >= 0
            n += 1
    assert n > 200


def test_snapshot_builder_survives_garbage_nodes():
    rng = random.Random(3)
    for kind in ("mig", "mps"):
        nodes = []
        for i in range(60):
            ann = {}
            for _ in range(rng.randrange(0, 8)):
                seed = rng.choice([PREFIX_STATUS + "0-1g.10gb-used", PREFIX_STATUS + "1-10gb-free", PREFIX_SPEC + "0-2g.20gb", "other/annotation"])
                ann[rng.choice(list(_mutations(rng, seed)))] = rng.choice(["1", "2", "x", "", "-3", "99999999999"])
            labels = {}
            if rng.random() < 0.8:
                labels[host.LABEL_PRODUCT] = rng.choice(["NVIDIA-A100-80GB-PCIe", "NVIDIA-H100-80GB", "A30", _rand(rng, 6), ""])
            if rng.random() < 0.8:
                labels[host.LABEL_COUNT] = rng.choice(["8", "1", "0", "-2", "x", "16", "17", "4000000000", ""])
            if rng.random() < 0.8:
                labels[host.LABEL_MEMORY] = rng.choice(["81920", "40960", "0", "-1", "abc", "1e9", ""])
            nodes.append(dict(name=f"n{i}", labels=labels, annotations=ann))
        s = host.Snapshot(kind, nodes, model_names=["A30", "NVIDIA-A100-80GB-PCIe", "NVIDIA-H100-80GB"])
        a = s.arrays()
        assert a["n_nodes"] == len(nodes) and len(a["gpu_off"]) == len(nodes) + 1
        assert (a["gpu_off"][1:] >= a["gpu_off"][:-1]).all() and a["gpu_off"][-1] == s.tg
        # a status annotation whose value is "-3" gives a count of -3 here as it does in extractGPUs (mig/node.go:65-100 does
        # not look at the sign); nosplan_set_*_snapshot is where such a count is refused (NOSPLAN_ERR_RANGE)
        assert a["used"].min() >= -3 and a["free"].min() >= -3 and a["used"].max() <= 2 and a["free"].max() <= 2
        s.close()


def test_table_loader_survives_garbage():
    rng = random.Random(5)
    good = pkg.synth.table_yaml_text()
    n_ok = 0
    for text in list(_mutations(rng, good)) * 8 + ["", "\n", "- ", "-", "- models:", "- models: [", "- models: [ a ]\n  allowedGeometries:",
                                                   "- models: [ a ]\n  allowedGeometries:\n    - : 1\n", "\t- models: [ a ]", "- models: [ a ]\n allowedGeometries:\n  - p: 99999999999999999999\n",
                                                   "- models: [ " + "m, " * 500 + "z ]\n  allowedGeometries:\n    - p: 1\n", "x" * 5000, "- " * 3000]:
        try:
            entries, g = host.load_tables(text)
        except ValueError as e:
            assert str(e)  # an error always carries its reason
            continue
        n_ok += 1
        assert len(g["geom_off"]) == len(g["models"]) + 1
    assert n_ok >= 1


def _vocab():
    nodes = [dict(name="n0", labels={host.LABEL_PRODUCT: "NVIDIA-H100-80GB", host.LABEL_COUNT: "2", host.LABEL_MEMORY: "81920"},
                  annotations={PREFIX_STATUS + "0-1g.10gb-free": "2", PREFIX_STATUS + "1-10gb-free": "1"})]
    mig = host.Snapshot("mig", nodes, model_names=["NVIDIA-H100-80GB"], extra_names=pkg.synth.MIG_PROFILES)
    mps = host.Snapshot("mps", nodes, extra_sizes=[5, 10, 20, 40])
    return mig, mps


def test_pod_rows_survive_garbage_specs():
    rng = random.Random(13)
    mig, mps = _vocab()
    names = ["nvidia.com/mig-1g.10gb", "nvidia.com/mig-2g.20gb", "nvidia.com/gpu-10gb", "nvidia.com/gpu-40gb", "cpu", "memory", "nvidia.com/gpu",
             "nvidia.com/mig-9g.99gb", "nvidia.com/gpu-7gb", "nvidia.com/mig-", ""]
    qtys = [0, 1, 2, 7, -1, 2**31 - 1, 2**31, 2**62, 2**63 - 1, -2**63]
    accepted = refused = 0
    for _ in range(400):
        pods = []
        for i in range(rng.randrange(0, 6)):
            mk = lambda: {rng.choice(list(_mutations(rng, rng.choice(names)))) if rng.random() < 0.2 else rng.choice(names): rng.choice(qtys)
                          for _ in range(rng.randrange(0, 4))}
            pods.append(dict(name=rng.choice(["ns/a", "ns/b", "", "x" * 300]), priority=rng.choice([0, 5, -5, 2**31 - 1]),
                             created=rng.choice([0, 1, -1, 2**62]), phase=rng.choice(["Pending", "Running"]),
                             node_name=rng.choice(["", "n0"]), unschedulable=rng.random() < 0.8,
                             containers=[mk() for _ in range(rng.randrange(0, 3))], init_containers=[mk() for _ in range(rng.randrange(0, 2))]))
        try:
            r = host.build_pods(pods, mig, mps)
        except ValueError as e:
            assert str(e)
            refused += 1
            continue
        accepted += 1
        n = len(r["source"])
        assert len(r["off"]) == n + 1 and (r["off"][1:] >= r["off"][:-1]).all()
        assert ((r["source"] >= 0) & (r["source"] < max(len(pods), 1))).all()
    assert accepted > 50 and refused > 50
    mig.close()
    mps.close()


def test_pod_builder_refuses_broken_csr_arrays():
    L = host.lib()
    mig, mps = _vocab()
    u8 = lambda *v: (C.c_uint8 * len(v))(*v)
    i32 = lambda *v: (C.c_int32 * len(v))(*v)
    i64 = lambda *v: (C.c_int64 * len(v))(*v)
    strs = lambda *v: (C.c_char_p * len(v))(*v)
    good = dict(n=1, pending=u8(1), sched=u8(0), unsched=u8(1), prio=i32(0), created=i64(0), name=strs(b"ns/p"), cont_off=i32(0, 1),
                cont_init=u8(0), res_off=i32(0, 1), res_name=strs(b"nvidia.com/mig-1g.10gb"), res_qty=i64(1))

    def call(**kw):
        a = dict(good, **kw)
        h = C.c_void_p(L.noshost_pods_build(a["n"], a["pending"], a["sched"], a["unsched"], a["prio"], a["created"], a["name"], a["cont_off"],
                                            a["cont_init"], a["res_off"], a["res_name"], a["res_qty"], mig.h, mps.h))
        err, n = L.noshost_pods_error(h).decode(), L.noshost_pods_count(h)
        L.noshost_pods_free(h)
        return err, n
    assert call() == ("", 1)
    for kw in (dict(n=-1), dict(cont_off=i32(1, 0)), dict(cont_off=i32(-1, 1)), dict(res_off=i32(1, 0)), dict(res_off=i32(-4, 1)),
               dict(res_name=strs(None)), dict(name=None), dict(cont_off=None), dict(res_off=None), dict(res_qty=None)):
        err, n = call(**kw)
        assert err and n == 0, kw
    assert call(name=strs(None)) == ("", 1)          # a pod without a name sorts as ""
    assert call(res_qty=i64(2**40))[0]               # does not fit the planner's int32
    mig.close()
    mps.close()


def test_snapshot_builder_refuses_broken_arguments():
    L = host.lib()
    strs = lambda *v: (C.c_char_p * max(len(v), 1))(*v)
    i32 = lambda *v: (C.c_int32 * max(len(v), 1))(*v)
    good = dict(kind=0, n=1, product=strs(b"A30"), count=strs(b"1"), memory=strs(b"24576"), ann_off=i32(0, 1),
                ann_key=strs((PREFIX_STATUS + "0-1g.6gb-free").encode()), ann_val=strs(b"4"), n_models=1, models=strs(b"A30"),
                n_names=0, names=None, n_sizes=0, sizes=None)

    def call(**kw):
        a = dict(good, **kw)
        h = L.noshost_snapshot_build(a["kind"], a["n"], a["product"], a["count"], a["memory"], a["ann_off"], a["ann_key"], a["ann_val"],
                                     a["n_models"], a["models"], a["n_names"], a["names"], a["n_sizes"], a["sizes"])
        h = C.c_void_p(h)
        out = (bool(h.value), L.noshost_snapshot_n_nodes(h), L.noshost_snapshot_total_gpus(h))
        if h.value:
            L.noshost_snapshot_free(h)
        return out
    assert call() == (True, 1, 1)
    for kw in (dict(kind=2), dict(n=-1), dict(product=None), dict(count=None), dict(ann_off=None), dict(ann_off=i32(1, 0)), dict(ann_off=i32(-1, 0)),
               dict(ann_key=None), dict(ann_val=None), dict(n_models=-1), dict(n_models=1, models=None), dict(n_names=2, names=None),
               dict(n_sizes=1, sizes=None), dict(kind=1, memory=None)):
        assert call(**kw) == (False, 0, 0), kw           # a null handle answers like an empty snapshot
    assert call(ann_key=strs(None)) == (True, 1, 1)      # a null key is not an annotation: the GPU comes from the count label
    assert call(models=strs(None)) == (True, 1, 0)       # no model table entry: the node is invalid (NewGPU: unknown model)
    # the accessors of a null handle and out-of-range indexes
    assert L.noshost_snapshot_key_name(None, 0) == b"" and L.noshost_snapshot_key_size(None, 3) == 0
    s = host.Snapshot("mig", [dict(labels={host.LABEL_PRODUCT: "A30", host.LABEL_COUNT: "1"}, annotations={})], model_names=["A30"],
                      extra_names=["1g.6gb"])
    assert L.noshost_snapshot_key_name(s.h, 99) == b"" and L.noshost_snapshot_key_name(s.h, -1) == b""
    z = i32(0)
    assert L.noshost_spec_annotations(s.h, 5, z, z, None, 0) == -1 and L.noshost_spec_annotations(s.h, -1, z, z, None, 0) == -1
    assert L.noshost_spec_annotations(None, 0, z, z, None, 0) == -1 and L.noshost_spec_annotations(s.h, 0, None, z, None, 0) == -1
    s.close()
    # NodePartitioning.Equal with empty and null sides
    assert L.noshost_node_partitioning_equal(3, 0, None, None, 0, None, None) == 1
    assert L.noshost_node_partitioning_equal(3, 1, None, None, 1, None, None) == 0
    assert L.noshost_node_partitioning_equal(-1, 1, z, z, 1, z, z) == 0


def test_table_accessors_are_bounds_checked():
    L = host.lib()
    err = C.create_string_buffer(64)
    L.noshost_tables_load.restype = C.c_void_p
    t = C.c_void_p(L.noshost_tables_load(pkg.synth.table_yaml_text().encode(), err, 64))
    assert t.value and L.noshost_tables_n_entries(t) == 3
    L.noshost_tables_entry_model.restype = L.noshost_tables_entry_geometry_profile.restype = C.c_char_p
    L.noshost_tables_model.restype = L.noshost_tables_profile.restype = C.c_char_p
    L.noshost_tables_entry_geometry_quantity.restype = C.c_int64
    for e, g, k in ((-1, 0, 0), (3, 0, 0), (0, -1, 0), (0, 99, 0), (0, 0, -1), (0, 0, 99)):
        assert L.noshost_tables_entry_geometry_profile(t, e, g, k) == b"" and L.noshost_tables_entry_geometry_quantity(t, e, g, k) == 0
    assert L.noshost_tables_entry_n_models(t, 7) == 0 and L.noshost_tables_entry_model(t, 0, 9) == b"" and L.noshost_tables_entry_n_geometries(t, -2) == 0
    assert L.noshost_tables_entry_geometry_size(t, 0, 50) == 0 and L.noshost_tables_model(t, 100) == b"" and L.noshost_tables_profile(t, -1) == b""
    assert L.noshost_tables_entry_geometry_profile(t, 0, 0, 0) == b"1g.6gb" and L.noshost_tables_entry_geometry_quantity(t, 0, 0, 0) == 4
    L.noshost_tables_free(t)
    assert L.noshost_tables_n_entries(None) == 0 and L.noshost_tables_n_models(None) == 0 and L.noshost_tables_model(None, 0) == b""
    # an error text longer than the caller's buffer is cut, not overrun
    small = C.create_string_buffer(b"\xaa" * 32, 32)
    assert not L.noshost_tables_load(b"- models: 3\n", small, 8)
    assert small.raw[8:] == b"\xaa" * 24 and small.raw[:8].find(b"\0") >= 0


def test_null_arguments_of_the_small_entry_points():
    L = host.lib()
    buf = C.create_string_buffer(64)
    assert L.noshost_format_status_annotation(3, None, 1, buf, 64) > 0 and buf.value == (PREFIX_STATUS + "3--free").encode()
    assert L.noshost_format_spec_annotation(3, None, buf, 64) > 0 and buf.value == (PREFIX_SPEC + "3-").encode()
    assert L.noshost_pods_count(None) == 0
    L.noshost_pods_error.restype = C.c_char_p
    assert L.noshost_pods_error(None) == b"null handle"
    out = C.c_int32(7)
    assert L.noshost_memory_gb(None, C.byref(out)) == 1 and out.value == 7
    assert L.noshost_mps_profile_size_gb(None) == 0
    assert L.noshost_mig_profile_of_resource(None, buf, 64) == 1 and L.noshost_mps_profile_of_resource(None, buf, 64) == 1
    idx, st, q = C.c_int32(), C.c_int32(), C.c_int32()
    assert L.noshost_parse_status_annotation(None, b"1", C.byref(idx), buf, 64, C.byref(st), C.byref(q)) == 1
    assert L.noshost_parse_spec_annotation(b"x", None, C.byref(idx), buf, 64, C.byref(q)) == 1
    L.noshost_snapshot_free(None)
    L.noshost_pods_free(None)
    L.noshost_tables_free(None)
