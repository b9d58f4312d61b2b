/* nosplan.h — C ABI of libnosplan.so, the B200-native GPU-partitioning planner hot path.
 *
 * Drop-in boundary for the walkai-nos (Nebuly nos fork) gpupartitioner.  The reference has NO FFI for
 * this path today (build/gpupartitioner/Dockerfile:32 builds with CGO_ENABLED=0); its only cgo
 * precedent is the `nvml` build tag (pkg/gpu/nvml/client.go:1, client_stub.go:1).  These entry points
 * are what a `//go:build cuda` cgo shim binds (INTEGRATION.md shows the shim).  Every entry point
 * cites the reference function(s) whose work it replaces; paths are relative to the reference root.
 *
 * Conventions
 *  - Plain C: pointers + sizes, caller-owned host buffers unless the name ends in `_dev`.
 *    The library copies what it needs during the call and keeps no caller pointer afterwards.
 *  - Return value: 0 = ok, negative = error class (below); nosplan_last_error(h) gives the text.
 *    A CUDA failure is always an error.  There is no CPU fallback anywhere in this library.
 *  - A handle is bound to one CUDA device and is NOT re-entrant: one call at a time per handle
 *    (the reference's planner is single-threaded too: MaxConcurrentReconciles 1,
 *    internal/controllers/gpupartitioner/mig_controller.go:204).
 *  - "Maps" of the reference (map[ProfileName]int) cross the boundary as dense rows over a
 *    vocabulary: MIG rows have one int32 column per profile name, names strictly ascending bytewise
 *    (the canonical order of SURVEY §D); MPS rows have one column per slice size in GB, strictly
 *    ascending.  Demand vectors (pod requests) are CSR: off[D+1], key[nnz], qty[nnz], keys ascending
 *    within a row; a key with qty 0 is a present map key (it matters for the gate).
 *  - Nodes appear in the reference's list order (snapshot order); "first node wins" = lowest index.
 *    GPUs of a node appear in the canonical order of SURVEY §D; gpu_off[N+1] is the CSR over GPUs.
 *    valid[n] = 0 marks a node whose NewNode failed in the reference (mig/node.go:40-64): it is
 *    skipped by the gate, by repartitioning and by scheduling, exactly as the controller does
 *    (mig_controller.go:135-137, 153-157).
 */
#ifndef NOSPLAN_H_
#define NOSPLAN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nosplan nosplan_t;

/* limits (checked; exceeding one returns NOSPLAN_ERR_RANGE) */
#define NOSPLAN_MAX_MIG_PROFILES 16   /* MIG vocabulary columns                                   */
#define NOSPLAN_MAX_MODELS 32         /* GPU models in the geometry table                          */
#define NOSPLAN_MAX_GEOMS 32          /* allowed geometries per model                              */
#define NOSPLAN_MAX_MPS_SIZES 32      /* MPS vocabulary columns                                    */
#define NOSPLAN_MAX_GPUS_PER_NODE 16
#define NOSPLAN_MAX_COUNT 255         /* used[p] + free[p] per GPU and profile — and per GPU and MPS size: counts are
                                         bytes on the device, a GPU cut into more than 255 slices of one size is
                                         refused at upload, never wrapped                                        */
#define NOSPLAN_MAX_MEM_GB 65535
#define NOSPLAN_MAX_DEMAND_KEYS 8     /* distinct profiles one pod may request                     */
#define NOSPLAN_MAX_DEMAND_QTY 1000000

/* error classes */
#define NOSPLAN_OK 0
#define NOSPLAN_ERR_ARG (-1)    /* null pointer, bad size, call order                               */
#define NOSPLAN_ERR_RANGE (-2)  /* a documented limit exceeded / value out of range                */
#define NOSPLAN_ERR_CUDA (-3)   /* CUDA runtime or kernel failure                                   */
#define NOSPLAN_ERR_CONFIG (-4) /* geometry table rejected (mirrors ValidateConfigs errors)        */

/* plan flags */
#define NOSPLAN_PLAN_GATE 1  /* apply profileAlreadyPresent before repartitioning (reference: on)  */
#define NOSPLAN_PLAN_SCHED 2 /* simulate scheduling with Node.AddPod first-fit                    */

/* ---- lifetime ---------------------------------------------------------------------------------- */
/* Binds a planner to CUDA device `device` (ordinal).  Replaces nothing in the reference: it is where
 * cmd/gpupartitioner/gpupartitioner.go:49-120 would construct the planner handle. */
int nosplan_create(int device, nosplan_t** out);
void nosplan_destroy(nosplan_t* h);
/* Text of the last error on this handle ("" if none).  Valid until the next call on the handle. */
const char* nosplan_last_error(const nosplan_t* h);
const char* nosplan_version(void);

/* ---- geometry tables ---------------------------------------------------------------------------- */
/* Replaces mig.SetKnownGeometries + ValidateConfigs (pkg/gpu/mig/known_configs.go:144-185).
 *   profile_names[n_profiles]  MIG vocabulary, strictly ascending bytewise; each must match
 *                              \d+g\.\d+gb (pkg/constant/constants.go:40) if it occurs in a geometry
 *   geom_off[n_models+1]       CSR over models; list order inside a model is kept (it decides
 *                              InitGeometry, pkg/gpu/partitioning.go:67-79)
 *   geom_counts[total][n_profiles]  quantity of each profile, 0 = key absent, negative rejected
 * The library renders Geometry.Id() (pkg/gpu/partitioning.go:37-56) for every geometry and ranks
 * the strings bytewise per model: that rank is the 4th key of the selection order (gpu.go:242). */
int nosplan_set_mig_tables(nosplan_t* h, int n_profiles, const char* const* profile_names, int n_models,
                           const int32_t* geom_off, const int32_t* geom_counts);

/* ---- cluster snapshot ---------------------------------------------------------------------------- */
/* Replaces, per node, mig.NewNode/extractGPUs (pkg/gpu/mig/node.go:40-100) once the Go side has
 * parsed labels and annotations: the snapshot is the flattened []mig.Node.
 *   model[total_gpus]               index into the model table (per GPU: tests build mixed nodes)
 *   used/free_[total_gpus][n_profiles]  usedMigDevices / freeMigDevices (gpu.go:29-35), >= 0
 * The arrays are uploaded as they are and range-checked on the device while they are packed; the
 * error reported is the first bad entry in GPU order (model before counts).  Rows of invalid nodes
 * are not interpreted.  A rejected snapshot leaves the handle without one (the same holds for
 * nosplan_set_mps_snapshot). */
int nosplan_set_mig_snapshot(nosplan_t* h, int n_nodes, const int32_t* gpu_off, const uint8_t* valid,
                             const int32_t* model, const int32_t* used, const int32_t* free_);

/* Replaces slicing.NewNode/extractGPUs (pkg/gpu/slicing/node.go:32-105).
 *   sizes_gb[n_sizes]               MPS vocabulary ("<k>gb", slicing/profile.go:48-59), ascending, >= 1
 *   mem_gb[total_gpus]              gpu.GetMemoryGB of the node label (pkg/gpu/util.go:59-73)
 *   used/free_[total_gpus][n_sizes] UsedProfiles / FreeProfiles (slicing/gpu.go:27-33)
 *   free_mask[total_gpus]           bit s set = key s present in FreeProfiles (a present key with
 *                                   value 0 behaves differently from an absent one in
 *                                   slicing/gpu.go:199-216); NULL = present iff value > 0 */
int nosplan_set_mps_snapshot(nosplan_t* h, int n_sizes, const int32_t* sizes_gb, int n_nodes,
                             const int32_t* gpu_off, const uint8_t* valid, const int32_t* mem_gb,
                             const int32_t* used, const int32_t* free_, const uint32_t* free_mask);

/* ---- batched scorer (the throughput path) ----------------------------------------------------------- */
/* For every demand row d and every valid node n, runs mig.Node.UpdateGeometryFor
 * (pkg/gpu/mig/node.go:145-177 -> mig.GPU.UpdateGeometryFor pkg/gpu/mig/gpu.go:160-196: CanApplyGeometry :99,
 * countProvidedProfiles :198, geometryDistance :245, totalSliceCount :262, isBetterGeometryScore
 * :232, ApplyGeometry :134) on a fresh copy of the snapshot node, as updateNodeGeometry does
 * (mig_controller.go:168-198).  One candidate evaluation = one iteration of gpu.go:168-188.
 * Outputs (each may be NULL):
 *   first_node[D]              lowest n that would update, -1 if none (tryRepartition, :146-166)
 *   first_choice[D][max_gpus]  for that node, index of the geometry applied to each GPU, -1 = kept
 *   updated[D][N]              1 if node n would update
 *   choice[D][total_gpus]      applied geometry index per GPU for every (d, n), -1 = kept
 * n_evals (may be NULL) receives the number of reference-equivalent candidate evaluations: every (row, node, GPU,
 * allowed geometry) tuple the reference's loop would visit, n_demands x nosplan_mig_evals_per_demand.  The kernels
 * execute fewer (a node's walk ends once the request is covered, a GPU that already holds enough skips its
 * candidates, the demand-independent half of an evaluation is done once per snapshot): nosplan_executed_evals counts those. */
int nosplan_score_mig(nosplan_t* h, int n_demands, const int32_t* dem_off, const int32_t* dem_key,
                      const int32_t* dem_qty, int32_t* first_node, int8_t* first_choice, uint8_t* updated,
                      int8_t* choice, int64_t* n_evals);

/* MPS twin: slicing.Node.UpdateGeometryFor (pkg/gpu/slicing/node.go:120-152 ->
 * slicing.GPU.UpdateGeometryFor pkg/gpu/slicing/gpu.go:162-220 with createSlices :238, canCreateMoreSlices :250).
 * One evaluation = one (demand, node, GPU) greedy run.
 *   out_free[D][total_gpus][n_sizes] (uint8) / out_free_mask[D][total_gpus]: FreeProfiles after the
 *   trial update, for every (d, n). */
int nosplan_score_mps(nosplan_t* h, int n_demands, const int32_t* dem_off, const int32_t* dem_key,
                      const int32_t* dem_qty, int32_t* first_node, uint8_t* updated, uint8_t* out_free,
                      uint32_t* out_free_mask, int64_t* n_evals);

/* ---- planner ------------------------------------------------------------------------------------------ */
/* Plan(snapshot, pendingPods): replays Controller.Reconcile (mig_controller.go:56-98) pod by pod on
 * the evolving state, starting from the uploaded snapshot (the snapshot itself is not modified):
 * gate profileAlreadyPresent (:121-144), tryRepartition first-node-wins (:146-166), then — with
 * NOSPLAN_PLAN_SCHED — Node.AddPod first-fit (mig/node.go:201-210 -> gpu.go:289-303; slicing twins
 * node.go:196-205, gpu.go:142-156).  Pods must already be filtered by shouldConsiderPod (:100-111)
 * and ordered (priority desc — pkg/util/pod/pod.go:84-88 — then creation time, then ns/name).
 *   kind[M]   0 = MIG pod (keys index the MIG vocabulary), 1 = MPS pod (keys index sizes)
 * Outputs per pod (each may be NULL): gate_hit, repart_node (-1 none), sched_node, sched_gpu (slot in
 * the node, -1 none).  The final state is read with nosplan_get_*_state. */
int nosplan_plan(nosplan_t* h, int n_pods, const uint8_t* kind, const int32_t* dem_off,
                 const int32_t* dem_key, const int32_t* dem_qty, int flags, uint8_t* gate_hit,
                 int32_t* repart_node, int32_t* sched_node, int32_t* sched_gpu);

/* State after the last nosplan_plan (or the snapshot if none ran), same layout as the setters.
 * mig: the desired geometry of GPU g is used+free where > 0 — BuildNodePartitioning
 * (internal/partitioning/mig/state.go:25-45) over these rows gives state.NodePartitioning. */
int nosplan_get_mig_state(nosplan_t* h, int32_t* used, int32_t* free_);
int nosplan_get_mps_state(nosplan_t* h, int32_t* used, int32_t* free_, uint32_t* free_mask);

/* Replaces NodeInitializer.InitNodePartitioning's per-GPU step (internal/partitioning/mig/
 * initializer.go:40-79 -> GPU.InitGeometry gpu.go:120-129 -> GetFewestSlicesGeometry
 * pkg/gpu/partitioning.go:67-79) over every GPU of the snapshot:
 *   err[total_gpus]  1 where the reference returns an error (used devices would be deleted)
 *   out_free[total_gpus][n_profiles]  free devices after the call (unchanged where err) */
int nosplan_init_geometry_mig(nosplan_t* h, uint8_t* err, int32_t* out_free);

/* Agent-side diff planner, count level: plan.NewMigConfigPlan (internal/controllers/migagent/plan/plan.go:31-92,
 * extractCandidatesForDeletion :112-137, extractResourcesToRecreate :94-110, getResourcesNotIncludedInSpec
 * :168-182) for every GPU of a batch.  Rows are [n_gpus][n_profiles]: used/free_ = devices the GPU has
 * ("unknown" status counts as used), spec = summed spec-annotation quantities, spec_mask[g] bit p = a spec
 * annotation names (g, p).  Outputs: create (new + re-created free devices), del_free, del_used — which
 * concrete device ids go is the host's choice (free first, ascending id, as the reference does). */
int nosplan_diff_plan_mig(nosplan_t* h, int n_gpus, int n_profiles, const int32_t* used, const int32_t* free_,
                          const int32_t* spec, const uint32_t* spec_mask, int32_t* create, int32_t* del_free,
                          int32_t* del_used);

/* The same planner down to the devices: which device ids each DeleteOperation names and which CreateOperations
 * exist (plan.go:31-92; MigConfigPlan{DeleteOperations, CreateOperations}, operation.go:25-38).
 *   dev_off[n_gpus + 1]     CSR over the GPUs' devices; the devices of a GPU in the order of
 *                           state.Flatten().SortByDeviceId() (plan.go:45: ascending device id) — the ids
 *                           themselves stay with the caller, index i of these arrays stands for them
 *   dev_profile[n_devices]  column of the device's profile, dev_status[n_devices] 0 used / 1 free / 2 unknown
 *   spec / spec_mask        as for nosplan_diff_plan_mig
 * Outputs:
 *   dev_op[n_devices]  0 keep | 1 deleted: its profile is not in the GPU's spec annotations (plan.go:37-41; one
 *                      DeleteOperation per (GPU, profile), mig.GroupDevicesByMigProfile) | 2 deleted to shrink the
 *                      profile (:60-70, extractCandidatesForDeletion :112-137: free devices first, ascending id; one
 *                      operation per (GPU, profile)) | 3 deleted to be re-created because the GPU gets a create
 *                      (:73-88; one DeleteOperation per GPU)
 *   create_new[n_gpus][n_profiles]    CreateOperation{profile, quantity} for diff > 0 (:61-64), 0 = none
 *   create_again[n_gpus][n_profiles]  CreateOperation for the re-created free devices (:85-87), 0 = none
 *   matches[n_gpus]    1 where MigState.Matches (mig_state.go:40-58) holds for the GPU's part of both maps; the
 *                      reference's answer for a node is the AND over its GPUs */
int nosplan_diff_plan_mig_devices(nosplan_t* h, int n_gpus, int n_profiles, const int32_t* dev_off, const int32_t* dev_profile,
                                  const uint8_t* dev_status, const int32_t* spec, const uint32_t* spec_mask, uint8_t* dev_op,
                                  int32_t* create_new, int32_t* create_again, uint8_t* matches);

/* ---- batch-demand planning mode: NEW BEHAVIOUR, not in the reference ----------------------------------------------
 * The reference's documentation still describes a planner that works on a batch of pending pods — aggregate what the
 * batch lacks, re-partition, simulate the scheduling, keep the partitioning that makes the most pods schedulable
 * (docs/en/docs/dynamic-gpu-partitioning/configuration.md:6-11, 19, 38; the knobs BatchWindowTimeoutSeconds /
 * BatchWindowIdleSeconds, pkg/api/nos.nebuly.com/config/v1alpha1/gpu_partitioner_config.go:28-50, and pkg/util/batcher.go
 * are what is left of it) — but this fork's code reconciles pod by pod.  This entry point is this project's definition of
 * that mode, written down as executable code in oracle/batch_plan.py and built only from functions the reference still
 * has (Node.AddPod first-fit, Node.UpdateGeometryFor); the batched scorer does the re-partitioning step of every round:
 *   A. first-fit every pod, in order, on the devices that are free now;
 *   B. rounds (at most max_rounds, 0 = 8): lacking[k] = what the unplaced pods still ask of profile / size k; the rows
 *      {k: lacking[k]} are scored against every node of the round's starting state (nosplan_score_mig / _mps with full
 *      tables); profile by profile, nodes in list order whose trial update succeeded and that no other profile took in
 *      this round are committed until the profile's free devices on the committed nodes cover lacking[k]; the unplaced
 *      pods are first-fitted on the new state; a round that places nobody ends the loop;
 *   both profile orders (ascending, descending) are simulated from the state after A; the one that places more pods is
 *   kept (ties: ascending).
 * Pods of ONE kind per call (kind 0: keys index the MIG vocabulary, 1: the sizes), one key each, 1 <= quantity <= 8.
 * Starts from the uploaded snapshot, which is left as it was; the planned state is read with nosplan_get_*_state.
 * Outputs: sched_node / sched_gpu [n_pods] (-1: still pending), phase [n_pods] (1 placed in A, 2 placed after a
 * re-partitioning, 0 pending), repartitioned [n_nodes] (1: the node's geometry was changed), order_out (0 ascending, 1
 * descending), rounds_out. */
int nosplan_plan_batch(nosplan_t* h, int kind, int n_pods, const int32_t* pod_key, const int32_t* pod_qty, int max_rounds,
                       int32_t* sched_node, int32_t* sched_gpu, uint8_t* phase, uint8_t* repartitioned, int32_t* order_out,
                       int32_t* rounds_out);

/* ---- device-resident variants (inputs/outputs already in HBM; used for kernel-only timing) ------------ */
/* Same as nosplan_score_mig/_mps but every pointer is a device pointer on the handle's device and the
 * work is enqueued on `stream` (a cudaStream_t) without synchronising.  first_node must be
 * pre-filled by the callee (it is); updated/choice may be NULL. */
int nosplan_score_mig_dev(nosplan_t* h, int n_demands, const int32_t* dem_off, const int32_t* dem_key,
                          const int32_t* dem_qty, int32_t* first_node, uint8_t* updated, int8_t* choice,
                          void* stream);
int nosplan_score_mps_dev(nosplan_t* h, int n_demands, const int32_t* dem_off, const int32_t* dem_key,
                          const int32_t* dem_qty, int32_t* first_node, uint8_t* updated, uint8_t* out_free,
                          uint32_t* out_free_mask, void* stream);

/* A whole device-resident scoring pass in one call: the MIG rows against the MIG snapshot and the MPS rows against
 * the MPS snapshot (same per-row semantics as nosplan_score_mig_dev / nosplan_score_mps_dev: mig/node.go:145-177,
 * slicing/node.go:120-152), first_node[0 .. n_mig_rows) for the MIG rows and [n_mig_rows .. + n_mps_rows) for the
 * MPS rows.  Both kinds are grouped by one launch and scored side by side (two streams, forked from and joined
 * back into `stream`); nothing is synchronised.
 * Without an exchange on the handle, first_node holds local indices + the node base, -1 = none.  With an open
 * exchange (nosplan_exchange_open) the call is collective: every rank calls it with the same row counts, and
 * first_node receives tryRepartition's answer over the WHOLE node list (mig_controller.go:146-166): the lowest
 * global index over ranks.  With NOSPLAN_PASS_SINGLE_KEY the scorers' epilogues fold their winners straight into
 * every rank's exchange buffer over NVLink while the other tiles are still being scored (no separate push, no local
 * round trip), and the last scorer CTA signals the step.
 * flags: NOSPLAN_PASS_SINGLE_KEY  every row has exactly one key (checked on the device: a row that does not is
 *                                 reported by nosplan_pass_status, never scored wrongly); the launches for
 *                                 rows with several keys are then left out
 *        NOSPLAN_PASS_SERIAL      score the kinds one after the other on `stream` (for timing a kernel alone)
 *        NOSPLAN_PASS_COUNT       measurement: count the evaluations that really execute (nosplan_executed_evals)
 * The first call (and any call with more rows than before) sizes the grouping scratch: it may allocate and thereby
 * synchronise the device; make it before a latency-critical loop or a stream capture.  Demand rows are validated on
 * the device; nosplan_pass_status (synchronises the device) reports the first bad row of the last pass and
 * exchange time-outs. */
#define NOSPLAN_PASS_SINGLE_KEY 1
#define NOSPLAN_PASS_SERIAL 2
#define NOSPLAN_PASS_COUNT 4
int nosplan_score_pass_dev(nosplan_t* h, int n_mig_rows, const int32_t* mig_off, const int32_t* mig_key, const int32_t* mig_qty,
                           int n_mps_rows, const int32_t* mps_off, const int32_t* mps_key, const int32_t* mps_qty,
                           int32_t* first_node, int flags, void* stream);
int nosplan_pass_status(nosplan_t* h);

/* Measurement hooks (no counterpart in the reference).  nosplan_kernel_timing(h, 1): nosplan_score_pass_dev records
 * CUDA events on the launching stream around its two dominant kernels (mig_score16_kernel, mps_score1_kernel);
 * nosplan_kernel_time_ms (synchronises) returns the summed duration and the number of launches of kind 0 (MIG) /
 * 1 (MPS) since the last read, at most the last 64.  nosplan_executed_evals (synchronises) returns and clears the
 * counts gathered by passes run with NOSPLAN_PASS_COUNT: MIG = (row, node, GPU, candidate) evaluations whose
 * provided-count was really computed (a (row, GPU) pair that enters the candidate core counts the model's
 * candidates), MPS = (row, node, GPU) runs of the closed form. */
int nosplan_kernel_timing(nosplan_t* h, int enable);
int nosplan_kernel_time_ms(nosplan_t* h, int kind, double* total_ms, int* launches);
int nosplan_executed_evals(nosplan_t* h, int64_t* mig, int64_t* mps);

/* Allocation-free MIG scorer.  It has no counterpart in the reference.
 * For callers that may not allocate device memory when they need a score (a sidecar working inside
 * a fixed buffer budget handed to it by another process's CUDA context, a control loop that cannot
 * afford cudaMalloc's implicit device synchronisation): same result as nosplan_score_mig_dev's
 * first_node, but the snapshot arrives as caller-owned DEVICE arrays (same row layout as
 * nosplan_set_mig_snapshot; every node has gpus_per_node GPUs, node n owns rows
 * [n*gpus_per_node, (n+1)*gpus_per_node); valid may be NULL = all valid) and the packed copy the
 * kernel needs lives in a fixed array inside the library (about 1.1 MB per GPU that loads the
 * library, held until exit).  No cudaMalloc/cudaFree or other allocating call is made.
 * Needs nosplan_set_mig_tables first (set-up time, may allocate).  Sizes beyond the two limits below
 * are rejected with NOSPLAN_ERR_RANGE.  The fixed array is shared by every call on a device: calls
 * whose kernels overlap on the GPU race, so serialise them.  The first call on a device performs a
 * one-time address lookup; make it before beginning any stream capture.  Values are not
 * range-checked (they are on the device): rows must satisfy the limits of nosplan_set_mig_snapshot. */
#define NOSPLAN_STATIC_MAX_NODES 4096
#define NOSPLAN_STATIC_MAX_GPUS_PER_NODE 8
int nosplan_score_mig_static(nosplan_t* h, int n_nodes, int gpus_per_node, const uint8_t* valid, const int32_t* model,
                             const int32_t* used, const int32_t* free_, int n_demands, const int32_t* dem_off,
                             const int32_t* dem_key, const int32_t* dem_qty, int32_t* first_node, void* stream);

/* Candidate evaluations one scorer pass performs for D demand rows on the current snapshot
 * (sum over valid nodes and their GPUs of the model's geometry count; for MPS, GPUs of valid nodes). */
int64_t nosplan_mig_evals_per_demand(const nosplan_t* h);
int64_t nosplan_mps_evals_per_demand(const nosplan_t* h);

/* ---- a whole scoring pass in one call ----------------------------------------------------------------- */
/* nosplan_set_mig_snapshot + nosplan_score_mig and nosplan_set_mps_snapshot + nosplan_score_mps for one
 * cluster snapshot (the structs carry exactly those calls' arguments), with the MIG half and the MPS half
 * running side by side inside the library (two private lanes with their own streams and buffers, the
 * second one on a worker thread owned by the handle), so that the copies and kernels of one kind overlap
 * those of the other.  This is the call the reference's Reconcile would make once per batch of pending
 * pods (mig_controller.go:56-98 lists nodes and pods, then decides): snapshot in, per request vector the
 * first node that would be re-partitioned out.  Either kind may be NULL.  Needs nosplan_set_mig_tables for
 * the MIG half.  The snapshots are NOT installed on the handle (later nosplan_plan / nosplan_score_* calls
 * still see what nosplan_set_*_snapshot installed).  Errors of a half come back with its code and its
 * message prefixed by "MIG: " / "MPS: ".  n_evals (may be NULL): evaluations of both halves. */
typedef struct nosplan_mig_snapshot {
  int32_t n_nodes;
  const int32_t* gpu_off;
  const uint8_t* valid;
  const int32_t* model;
  const int32_t* used;
  const int32_t* free_;
} nosplan_mig_snapshot_t;
typedef struct nosplan_mps_snapshot {
  int32_t n_sizes;
  const int32_t* sizes_gb;
  int32_t n_nodes;
  const int32_t* gpu_off;
  const uint8_t* valid;
  const int32_t* mem_gb;
  const int32_t* used;
  const int32_t* free_;
  const uint32_t* free_mask;
} nosplan_mps_snapshot_t;
typedef struct nosplan_demands {
  int32_t n_rows;
  const int32_t* off;
  const int32_t* key;
  const int32_t* qty;
} nosplan_demands_t;
int nosplan_score_snapshot(nosplan_t* h, const nosplan_mig_snapshot_t* mig, const nosplan_demands_t* mig_rows,
                           int32_t* mig_first_node, int8_t* mig_first_choice, const nosplan_mps_snapshot_t* mps,
                           const nosplan_demands_t* mps_rows, int32_t* mps_first_node, int64_t* n_evals);

/* ---- sharded scoring: one process per GPU, nodes split into contiguous ranges ------------- */
/* tryRepartition walks the whole node list and commits the first node whose UpdateGeometryFor
 * returns true (internal/controllers/gpupartitioner/mig_controller.go:146-166).  With the list split
 * across ranks that is the minimum global node index over ranks, one exchange per scoring pass.
 *
 * nosplan_set_node_base: the global index of this handle's first MIG / MPS node; the _dev scorers
 * then report first_node as a global index (-1 stays -1). */
int nosplan_set_node_base(nosplan_t* h, int mig_first_node, int mps_first_node);

/* The gate's ingredient, per shard: Controller.profileAlreadyPresent (mig_controller.go:121-144) asks, for
 * every requested profile, whether some node's geometry holds it: the sum over valid nodes and their GPUs
 * of used + free.  count[n_keys] (host array, n_keys = profiles for kind 0, sizes for kind 1) receives those
 * sums for the handle's WORKING state of the kind (the snapshot, or what the last nosplan_plan left); with
 * the node list sharded the gate is "sum over ranks > 0" (one all-reduce SUM over n_keys integers). */
int nosplan_presence(nosplan_t* h, int kind, int64_t* count);

/* The exchange itself, done by the library's own kernels over peer memory (NVLink), no library collective: every rank
 * keeps its keys of a step in its OWN buffer and signals a per-rank step flag on every peer; a finishing kernel on each
 * rank then reads all ranks' buffers with coalesced loads and keeps the minimum (csrc/nosplan_xchg.cuh).  Set-up
 * (allocates; once per process group):
 *   1. nosplan_exchange_create on every rank -> a 64-byte CUDA IPC handle of the rank's buffer
 *   2. gather the handles of all ranks (any transport) into handles[world][64]
 *   3. nosplan_exchange_open on every rank, then a barrier: nobody may run a pass before everybody has mapped
 * Per pass, on every rank, in the same order: nosplan_exchange_min_dev(n_keys, keys_in, keys_out) —
 * keys_in[d] = this rank's global node index for row d or -1; keys_out[d] = the minimum over ranks of the
 * non-negative entries, or -1 — or nosplan_score_pass_dev / nosplan_score_snapshot, which do the same from inside the
 * scorers.  Device pointers; enqueued on `stream`, not synchronised; n_keys must be the same on all ranks and at most
 * max_keys.  world <= 16.
 * Rules: all passes of a handle go down ONE stream (the step protocol is ordered by it).  A rank that waits more than
 * 2 s for a peer gives up: that pass's results read -2, nosplan_exchange_status / nosplan_pass_status (synchronise)
 * report NOSPLAN_ERR_CUDA, and the handle refuses further passes until the exchange is destroyed and created again on
 * every rank — so the 2 s must cover the ranks' start-up skew (barrier before the first pass).  Before any rank calls
 * nosplan_exchange_destroy (or nosplan_destroy), every rank must have finished its last pass (status call) and the
 * ranks must have met at a barrier: a peer still in flight would write flags into freed memory. */
#define NOSPLAN_IPC_HANDLE_BYTES 64
int nosplan_exchange_create(nosplan_t* h, int rank, int world, int max_keys, unsigned char* handle_out);
int nosplan_exchange_open(nosplan_t* h, const unsigned char* handles);
int nosplan_exchange_min_dev(nosplan_t* h, int n_keys, const int32_t* keys_in, int32_t* keys_out, void* stream);
int nosplan_exchange_status(nosplan_t* h);
int nosplan_exchange_destroy(nosplan_t* h);

#ifdef __cplusplus
}
#endif
#endif /* NOSPLAN_H_ */
