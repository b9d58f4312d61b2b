"""CPU restatement (test infrastructure) of the agent-side diff planner, device level:
plan.NewMigConfigPlan — internal/controllers/migagent/plan/plan.go:31-92 with
extractResourcesToRecreate :94-110, extractCandidatesForDeletion :112-137,
getResourcesNotIncludedInSpec :168-182 and MigState.WithoutMigProfiles (mig_state.go:76-87).
Pure Python: used for the reference's table tests and small random cases only.

state: {gpu_index: [(device_id: str, profile: str, status: "used"|"free"|"unknown"), ...]}
spec:  [(gpu_index, profile, quantity), ...]            (parsed spec annotations)
Returns (create_ops, delete_ops): create op = (gpu, profile, qty); delete op = list of
(gpu, device_id, profile, status).  Map iteration is canonicalised (GPUs ascending, profiles by name);
the reference's tests compare with ElementsMatch, so order inside the lists is free."""
from collections import OrderedDict


def new_mig_config_plan(state, spec):
    create_ops, delete_ops = [], []
    flat = lambda st: [(g, d, p, s) for g in sorted(st) for (d, p, s) in st[g]]

    def group_by_profile(devs):
        out = OrderedDict()
        for r in devs:
            out.setdefault((r[0], r[2]), []).append(r)
        return out

    # getResourcesNotIncludedInSpec: per GPU that has spec annotations, drop the resources of the
    # profiles named there; everything left (incl. GPUs without any spec annotation) is deleted
    spec_by_gpu = OrderedDict()
    for (g, p, q) in spec:
        spec_by_gpu.setdefault(g, []).append((g, p, q))
    updated = {g: list(v) for g, v in state.items()}
    for g, anns in spec_by_gpu.items():
        profs = [a[1] for a in anns]
        updated[g] = [r for r in state.get(g, []) if r[1] not in profs]
    for _, lst in sorted(group_by_profile(flat(updated)).items()):
        delete_ops.append(list(lst))

    # resources of the GPUs named in spec, sorted by device id (string compare), grouped by GPU
    by_gpu = {}
    for r in sorted(flat(state), key=lambda r: r[1]):
        by_gpu.setdefault(r[0], []).append(r)
    for g in sorted(spec_by_gpu):
        anns = spec_by_gpu[g]
        gpu_res = group_by_profile(by_gpu.get(g, []))
        n_create = 0
        by_prof = OrderedDict()
        for a in anns:
            by_prof.setdefault(a[1], []).append(a)
        for p in sorted(by_prof):
            actual = gpu_res.get((g, p), [])
            desired = sum(a[2] for a in by_prof[p])
            diff = desired - len(actual)
            if diff > 0:
                create_ops.append((g, p, diff))
                n_create += 1
            if diff < 0:
                n = -diff
                cand = []
                for r in actual:  # free devices first
                    if r[3] == "free":
                        cand.append(r)
                    if len(cand) == n:
                        break
                if len(cand) < n:
                    for r in actual:
                        if r[3] != "free":
                            cand.append(r)
                        if len(cand) == n:
                            break
                delete_ops.append(cand)
        if n_create == 0:
            continue
        # re-create the free resources not already scheduled for deletion
        doomed = {(r[0], r[1]) for op in delete_ops for r in op}
        again = [r for r in by_gpu.get(g, []) if r[3] == "free" and (r[0], r[1]) not in doomed]
        if again:
            delete_ops.append(again)
            for (gg, p), lst in group_by_profile(again).items():
                create_ops.append((gg, p, len(lst)))
    return create_ops, delete_ops


def counts(create_ops, delete_ops):
    """Per (gpu, profile): devices created, free devices deleted, used (non-free) devices deleted."""
    c, df, du = {}, {}, {}
    for (g, p, q) in create_ops:
        c[(g, p)] = c.get((g, p), 0) + q
    for op in delete_ops:
        for (g, _, p, s) in op:
            d = df if s == "free" else du
            d[(g, p)] = d.get((g, p), 0) + 1
    return c, df, du
