/* noshost.h — C ABI of libnoshost.so: the host side of the planner drop-in (no CUDA in here).
 *
 * It is the native form of what the reference does in Go on either side of the hot path — turning
 * node labels/annotations and pod specs into the flat arrays of nosplan.h, and turning the planned
 * state back into the annotation wire format — so a snapshot taken from a cluster dump can be fed to
 * libnosplan without going through per-pod, per-node string parsing (SURVEY §8 rows a10-a12, a15, a21,
 * f1).  Every function cites the reference code it mirrors (paths relative to the reference root).
 * Canonical orders are those of SURVEY §D.  Return: 0 ok, 1 where the reference returns an error.
 */
#ifndef NOSHOST_H_
#define NOSHOST_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- wire format: pkg/gpu/annotation.go, pkg/api/nos.nebuly.com/v1alpha1/annotations.go:21-58 ------------ */
#define NOSHOST_STATUS_USED 0
#define NOSHOST_STATUS_FREE 1
#define NOSHOST_STATUS_UNKNOWN 2
/* Return codes of the parsers and name scanners below: 0 = parsed, 1 = the text is not of that form (what the reference
 * reports as an error), 2 = the name does not fit the caller's buffer with its terminator (nothing is written: a parsed
 * name is never truncated).  Indexes and quantities are int32 in this ABI: strconv.Atoi would accept up to 64 bits, a
 * number beyond 32 bits is answered with 1 here (the annotation is ignored) instead of being wrapped.
 * ParseStatusAnnotation (annotation.go:56-85): "nos.nebuly.com/status-gpu-<idx>-<profile>-<status>" = qty */
int noshost_parse_status_annotation(const char* key, const char* value, int32_t* index, char* profile, int profile_cap,
                                    int32_t* status, int32_t* quantity);
/* ParseSpecAnnotation (annotation.go:29-54): "nos.nebuly.com/spec-gpu-<idx>-<profile>" = qty */
int noshost_parse_spec_annotation(const char* key, const char* value, int32_t* index, char* profile, int profile_cap,
                                  int32_t* quantity);
/* StatusAnnotation.String / SpecAnnotation.String (annotation.go:125-127, 161-163); returns the length */
int noshost_format_status_annotation(int32_t index, const char* profile, int32_t status, char* out, int cap);
int noshost_format_spec_annotation(int32_t index, const char* profile, char* out, int cap);
/* gpu.GetMemoryGB (pkg/gpu/util.go:59-73): Atoi the nvidia.com/gpu.memory label, ceil(MB / 1000); 1 when the label does
 * not parse or the result does not fit int32 */
int noshost_memory_gb(const char* label, int32_t* out);
/* mig.ExtractProfileName (pkg/gpu/mig/util.go:46-52): resource must match nvidia\.com\/mig-\d+g\.\d+gb */
int noshost_mig_profile_of_resource(const char* resource, char* profile, int cap);
/* slicing.ExtractProfileName (pkg/gpu/slicing/util.go:34-40): resource must match nvidia\.com/gpu-\d+gb */
int noshost_mps_profile_of_resource(const char* resource, char* profile, int cap);
/* slicing.ProfileName.GetMemorySizeGB (pkg/gpu/slicing/profile.go:52-59): 0 when it does not parse */
int noshost_mps_profile_size_gb(const char* profile);

/* ---- snapshot builder: mig.NewNode/extractGPUs (pkg/gpu/mig/node.go:40-100) and slicing.NewNode/extractGPUs
 * (pkg/gpu/slicing/node.go:32-105) over a whole node list -------------------------------------------------- */
typedef struct noshost_snapshot noshost_snapshot;
#define NOSHOST_KIND_MIG 0
#define NOSHOST_KIND_MPS 1
/* Nodes in list order.  product/count/memory[n] = the nvidia.com/gpu.product / gpu.count / gpu.memory
 * label values, NULL where the label is missing.  Annotations of node n are ann_key/ann_val
 * [ann_off[n], ann_off[n+1]).  MIG: model_names[n_models] are the models of the geometry table (a node
 * whose model is not among them fails like NewGPU does, mig/gpu.go:46-49); extra_names (may be NULL)
 * are profile names that must be in the vocabulary besides those found in annotations (table and pod
 * profiles).  MPS: extra_sizes are slice sizes requested by pods.  A node for which the reference's
 * NewNode returns an error is kept with valid = 0 and no GPUs.  GPUs of a node: annotated GPUs by
 * ascending index, then the "missing" ones with index len..count-1 (node.go:89-96, quirk kept); a gpu.count label above
 * 4096 leaves the node invalid instead of materialising that many GPUs.
 * Returns NULL when the arguments cannot be used (unknown kind, negative count, a required array missing, ann_off not
 * ascending from >= 0) or memory runs out; a NULL key, value, model or extra name is skipped.  Every accessor below
 * answers a NULL handle or an index outside the snapshot with 0 / "" / NULL. */
noshost_snapshot* noshost_snapshot_build(int kind, int n_nodes, const char* const* product, const char* const* count,
                                         const char* const* memory, const int32_t* ann_off, const char* const* ann_key,
                                         const char* const* ann_val, int n_models, const char* const* model_names,
                                         int n_extra_names, const char* const* extra_names, int n_extra_sizes,
                                         const int32_t* extra_sizes);
void noshost_snapshot_free(noshost_snapshot* s);
int noshost_snapshot_n_nodes(const noshost_snapshot* s);
int noshost_snapshot_total_gpus(const noshost_snapshot* s);
int noshost_snapshot_n_keys(const noshost_snapshot* s);              /* vocabulary size P (MIG) or S (MPS) */
const char* noshost_snapshot_key_name(const noshost_snapshot* s, int k); /* MIG profile name / "<k>gb" */
int noshost_snapshot_key_size(const noshost_snapshot* s, int k);      /* MPS size in GB (0 for MIG) */
const int32_t* noshost_snapshot_gpu_off(const noshost_snapshot* s);   /* [n_nodes+1] */
const uint8_t* noshost_snapshot_valid(const noshost_snapshot* s);     /* [n_nodes] */
const int32_t* noshost_snapshot_model(const noshost_snapshot* s);     /* [total_gpus] (MIG) */
const int32_t* noshost_snapshot_mem_gb(const noshost_snapshot* s);    /* [total_gpus] (MPS) */
const int32_t* noshost_snapshot_gpu_index(const noshost_snapshot* s); /* [total_gpus] the reference's GPU index */
const int32_t* noshost_snapshot_used(const noshost_snapshot* s);      /* [total_gpus][n_keys] */
const int32_t* noshost_snapshot_free_(const noshost_snapshot* s);     /* [total_gpus][n_keys] */
const uint32_t* noshost_snapshot_used_mask(const noshost_snapshot* s);/* key-present masks */
const uint32_t* noshost_snapshot_free_mask(const noshost_snapshot* s);

/* ---- pods: shouldConsiderPod (mig_controller.go:100-111, pkg/util/pod/pod.go:37-56), ComputePodRequest
 * (pkg/resource/resource.go:127-146: sum over containers, element-wise max with the init containers; pod
 * overhead is computed and dropped there, so it is not an input), GetRequestedProfiles (mig/util.go:87-95,
 * slicing/util.go:59-67), IsMoreImportant (pod.go:84-88) + creation time + name ordering ----- */
typedef struct noshost_pods noshost_pods;
/* Per pod: pending/scheduled/unschedulable flags, priority (0 when unset), creation time, "ns/name".
 * Containers of pod i: [cont_off[i], cont_off[i+1]); cont_init[c] != 0 marks an init container; the
 * resource requests of container c: res_name/res_qty [res_off[c], res_off[c+1]).
 * vocab: the snapshot(s) the demand keys must index (either may be NULL).
 * The arrays are checked before they are used (offsets ascending from >= 0, no null resource name; a pod without a
 * name sorts as ""); request quantities are summed with saturation and must fit int32 in the end.  A refusal leaves
 * count 0 and its reason in noshost_pods_error. */
noshost_pods* noshost_pods_build(int n_pods, const uint8_t* pending, const uint8_t* scheduled, const uint8_t* unschedulable,
                                 const int32_t* priority, const int64_t* created, const char* const* name,
                                 const int32_t* cont_off, const uint8_t* cont_init, const int32_t* res_off,
                                 const char* const* res_name, const int64_t* res_qty, const noshost_snapshot* mig,
                                 const noshost_snapshot* mps);
void noshost_pods_free(noshost_pods* p);
int noshost_pods_count(const noshost_pods* p);            /* rows (a pod asking for both kinds gives two rows) */
const int32_t* noshost_pods_source(const noshost_pods* p); /* [count] index of the input pod, planner order */
const uint8_t* noshost_pods_kind(const noshost_pods* p);   /* [count] 0 MIG, 1 MPS */
const int32_t* noshost_pods_off(const noshost_pods* p);    /* CSR of nosplan_plan */
const int32_t* noshost_pods_key(const noshost_pods* p);
const int32_t* noshost_pods_qty(const noshost_pods* p);
const char* noshost_pods_error(const noshost_pods* p);     /* "" or why a request could not be mapped */

/* ---- plan output: BuildNodePartitioning (internal/partitioning/mig/state.go:25-45) + getGPUSpecAnnotationList
 * (internal/partitioning/mig/partitioner.go:74-91): spec annotations of node n from the planned state.
 * Writes "key=value\n" lines ("nos.nebuly.com/spec-gpu-<idx>-<profile>=<qty>", quantity > 0 only, GPUs in
 * snapshot order, profiles ascending) into out; returns the length needed (call again if > cap), or -1 for a NULL
 * argument or a node index outside the snapshot. */
int noshost_spec_annotations(const noshost_snapshot* s, int node, const int32_t* used, const int32_t* free_, char* out,
                             int cap);
/* state.NodePartitioning.Equal (internal/partitioning/state/partitioning.go:33-38): same multiset of
 * (GPUIndex, Resources) — order-insensitive.  GPUs given as index + resource rows over n_keys columns
 * (0 = key absent).  Returns 1 if equal. */
int noshost_node_partitioning_equal(int n_keys, int n_a, const int32_t* idx_a, const int32_t* res_a, int n_b,
                                    const int32_t* idx_b, const int32_t* res_b);

/* ---- geometry tables: the loader of config/gpupartitioner/manager/known_mig_geometries.yaml
 * (cmd/gpupartitioner/gpupartitioner.go:122-132 -> sigs.k8s.io/yaml -> AllowedMigGeometries.UnmarshalJSON,
 * pkg/gpu/mig/allowed_geometries.go:30-59: `models` and `allowedGeometries` are both required, a list of strings and
 * a list of {profile: quantity} maps) and AllowedMigGeometriesList.GroupByModel (:74-82: a model named by several
 * entries keeps the last one's list).  text = the YAML document.  Returns NULL and writes the reason into err where
 * the reference's Unmarshal returns an error.  The entry accessors give the list as parsed; the flat accessors give
 * GroupByModel in the layout of nosplan_set_mig_tables: models in order of first appearance, profile vocabulary
 * ascending bytewise, geom_off[n_models + 1], geom_counts[total][n_profiles].  Names and quantities are not
 * validated here (the reference validates in SetKnownGeometries; nosplan_set_mig_tables does).  The reason is cut to
 * err_cap bytes including its terminator; accessors answer a NULL handle or an index outside the table with 0 / "" / NULL. */
typedef struct noshost_tables noshost_tables;
noshost_tables* noshost_tables_load(const char* text, char* err, int err_cap);
void noshost_tables_free(noshost_tables* t);
int noshost_tables_n_entries(const noshost_tables* t);
int noshost_tables_entry_n_models(const noshost_tables* t, int e);
const char* noshost_tables_entry_model(const noshost_tables* t, int e, int i);
int noshost_tables_entry_n_geometries(const noshost_tables* t, int e);
int noshost_tables_entry_geometry_size(const noshost_tables* t, int e, int g);
const char* noshost_tables_entry_geometry_profile(const noshost_tables* t, int e, int g, int k);
int64_t noshost_tables_entry_geometry_quantity(const noshost_tables* t, int e, int g, int k);
int noshost_tables_n_models(const noshost_tables* t);
const char* noshost_tables_model(const noshost_tables* t, int m);
int noshost_tables_n_profiles(const noshost_tables* t);
const char* noshost_tables_profile(const noshost_tables* t, int p);
const int32_t* noshost_tables_geom_off(const noshost_tables* t);
const int32_t* noshost_tables_geom_counts(const noshost_tables* t);

#ifdef __cplusplus
}
#endif
#endif /* NOSHOST_H_ */
