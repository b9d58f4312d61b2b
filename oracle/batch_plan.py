"""CPU definition (test infrastructure) of the batch-demand planning mode — NEW BEHAVIOUR, not in the reference.

The reference's docs still describe it (docs/en/docs/dynamic-gpu-partitioning/configuration.md:6-11, 19, 38: pending pods
are processed in batches; the lacking resources of the batch decide the new partitioning; scheduling is simulated and the
partitioning that makes the most pods schedulable is kept) and its knobs survive as dead configuration
(pkg/api/nos.nebuly.com/config/v1alpha1/gpu_partitioner_config.go:28-50, pkg/util/batcher.go:25-130), but the code is gone
from this fork.  What follows is this project's definition, built only from functions the reference still has —
Node.AddPod first-fit (mig/node.go:201-210, slicing/node.go:196-205) and Node.UpdateGeometryFor (mig/node.go:145-177,
slicing/node.go:120-152) — and shaped so that the expensive step is one call of the batched scorer per round:

  A. first-fit every pod of the batch, in planner order, on the devices that are free now;
  B. up to `max_rounds` rounds: lacking[k] = quantity of profile k the unplaced pods still ask for; every row
     {k: lacking[k]} is scored against every node ON THE STATE AT THE START OF THE ROUND (Node.UpdateGeometryFor on a copy);
     then, profile by profile (ascending or descending profile index = the candidate order), nodes in list order whose
     trial update succeeded and that no other profile took in this round are committed until the free devices of the
     profile on the committed nodes cover lacking[k]; the unplaced pods are then first-fitted, in planner order, on the
     new state (what they take becomes used and cannot be deleted by a later round); a round that places nobody ends the loop.
  Both profile orders are simulated from the same starting state; the one that schedules more pods is kept (ties: ascending).

Pods ask for one profile each (1 <= quantity <= 8)."""
import numpy as np


def _copy(snap):
    return {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in snap.items()}


def _first_fit(snap, kind, pods, todo, sched_node, sched_gpu):
    """Node.AddPod first-fit for pods[todo] in order; mutates snap."""
    off = snap["gpu_off"]
    tg = int(off[-1])
    gnode = np.repeat(np.arange(int(snap["n_nodes"])), np.diff(off))
    ok = np.repeat(snap["valid"].astype(bool), np.diff(off)) if tg else np.zeros(0, bool)
    free, used = snap["free"], snap["used"]
    for i in todo:
        k, q = pods[i]
        hit = np.nonzero(ok & (free[:, k] >= q))[0]
        if hit.size == 0:
            continue
        g = int(hit[0])
        free[g, k] -= q
        used[g, k] += q
        if kind == "mps":
            snap["free_mask"][g] |= np.uint32(1 << k)  # FreeProfiles[k] -= q keeps the key (slicing/gpu.go:153)
            snap["used_mask"][g] |= np.uint32(1 << k)
        else:
            snap["used_mask"][g] |= np.uint32(1 << k)
        sched_node[i] = gnode[g]
        sched_gpu[i] = g - off[gnode[g]]


def _one_order(o, kind, snap1, pods, placed_a, descending, max_rounds):
    snap = _copy(snap1)
    N = int(snap["n_nodes"])
    off = snap["gpu_off"]
    K = snap["free"].shape[1]
    unplaced = [i for i in range(len(pods)) if not placed_a[i]]
    repart = np.zeros(N, np.uint8)
    sched_node = np.full(len(pods), -1, np.int32)
    sched_gpu = np.full(len(pods), -1, np.int32)
    rounds = 0
    for _ in range(max_rounds):
        lacking = np.zeros(K, np.int64)
        for i in unplaced:
            lacking[pods[i][0]] += pods[i][1]
        keys = [k for k in range(K) if lacking[k] > 0]
        if not keys:
            break
        req = np.zeros((len(keys), K), np.int32)
        mask = np.zeros(len(keys), np.uint32)
        for r, k in enumerate(keys):
            req[r, k] = min(int(lacking[k]), 1000000)
            mask[r] = 1 << k
        if kind == "mig":
            upd, choice, _ = o.score_mig(snap, req, mask, threads=4)
        else:
            upd, out_free, out_mask, _ = o.score_mps(snap, req, mask, threads=4)
        taken = np.zeros(N, bool)
        order = list(range(len(keys)))
        if descending:
            order.reverse()
        for r in order:
            k = keys[r]
            rem = int(lacking[k])
            for n in range(N):
                if rem <= 0:
                    break
                if not upd[r, n] or taken[n]:
                    continue
                g0, g1 = int(off[n]), int(off[n + 1])
                if kind == "mig":
                    for g in range(g0, g1):
                        c = int(choice[r, g])
                        if c >= 0:  # ApplyGeometry (mig/gpu.go:134-154): free = cand - used, keys outside cand vanish
                            cand = o.geom_counts[o.geom_off[int(snap["model"][g])] + c]
                            snap["free"][g] = cand - snap["used"][g]
                            snap["free_mask"][g] = sum(1 << p for p in range(K) if cand[p] > 0)
                else:
                    snap["free"][g0:g1] = out_free[r, g0:g1]
                    snap["free_mask"][g0:g1] = out_mask[r, g0:g1]
                taken[n] = True
                repart[n] = 1
                rem -= int(snap["free"][g0:g1, k].sum())
        rounds += 1
        # the round's gains are locked in at once: pods that fit now are placed (their devices become used, which no
        # later ApplyGeometry may delete, mig/gpu.go:105-109), the others carry their demand into the next round
        _first_fit(snap, kind, pods, unplaced, sched_node, sched_gpu)
        still = [i for i in unplaced if sched_node[i] < 0]
        if len(still) == len(unplaced):
            break  # a round that places nobody ends the search
        unplaced = still
    return dict(snap=snap, repart=repart, sched_node=sched_node, sched_gpu=sched_gpu, rounds=rounds,
                scheduled=int((sched_node >= 0).sum()))


def batch_plan(o, kind, snap0, pod_key, pod_qty, max_rounds=8):
    """Returns dict(sched_node, sched_gpu [M] (-1: stays pending), phase [M] (1: placed on what was free, 2: placed after
    the re-partitioning, 0: pending), repartitioned [N], order (0 ascending / 1 descending), final snapshot)."""
    pods = [(int(k), int(q)) for k, q in zip(pod_key, pod_qty)]
    for k, q in pods:
        assert 1 <= q <= 8
    snap1 = _copy(snap0)
    M = len(pods)
    node_a = np.full(M, -1, np.int32)
    gpu_a = np.full(M, -1, np.int32)
    _first_fit(snap1, kind, pods, range(M), node_a, gpu_a)
    placed_a = node_a >= 0
    best = None
    for descending in (False, True):
        r = _one_order(o, kind, snap1, pods, placed_a, descending, max_rounds)
        if best is None or r["scheduled"] > best["scheduled"]:
            best = r
            best["order"] = 1 if descending else 0
    sched_node = np.where(placed_a, node_a, best["sched_node"]).astype(np.int32)
    sched_gpu = np.where(placed_a, gpu_a, best["sched_gpu"]).astype(np.int32)
    phase = np.where(placed_a, 1, np.where(best["sched_node"] >= 0, 2, 0)).astype(np.uint8)
    return dict(sched_node=sched_node, sched_gpu=sched_gpu, phase=phase, repartitioned=best["repart"], order=best["order"],
                rounds=best["rounds"], snap=best["snap"])
