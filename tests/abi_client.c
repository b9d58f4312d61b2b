/* A plain-C client of both ABIs: proves the headers are valid C and that the libraries link and answer
 * without Python in the way.  Exit code 0 = ok.  With no CUDA device nosplan_create must FAIL (no fallback). */
#include <stdio.h>
#include <string.h>

#include "nosplan.h"
#include "noshost.h"

int main(int argc, char** argv) {
  int expect_gpu = argc > 1 && strcmp(argv[1], "gpu") == 0;
  if (strstr(nosplan_version(), "nosplan") == NULL) return 1;
  int32_t idx = -1, st = -1, q = -1;
  char prof[64];
  if (noshost_parse_status_annotation("nos.nebuly.com/status-gpu-1-1g.10gb-used", "1", &idx, prof, sizeof prof, &st, &q)) return 2;
  if (idx != 1 || q != 1 || st != NOSHOST_STATUS_USED || strcmp(prof, "1g.10gb") != 0) return 3;
  int32_t gb = 0;
  if (noshost_memory_gb("1100", &gb) || gb != 2) return 4;
  nosplan_t* h = NULL;
  int rc = nosplan_create(0, &h);
  if (!expect_gpu) {
    if (rc == NOSPLAN_OK) { nosplan_destroy(h); return 0; } /* a GPU happens to be there: fine */
    return (h == NULL && rc == NOSPLAN_ERR_CUDA) ? 0 : 5;
  }
  if (rc != NOSPLAN_OK) return 6;
  /* one A30 GPU, empty; request 1g.6gb:2 -> the planner must repartition to {1g.6gb:4} (node_test.go:336) */
  const char* names[3] = {"1g.6gb", "2g.12gb", "4g.24gb"};
  int32_t geom_off[2] = {0, 4};
  int32_t geoms[4 * 3] = {0, 0, 1, 0, 2, 0, 2, 1, 0, 4, 0, 0};
  if (nosplan_set_mig_tables(h, 3, names, 1, geom_off, geoms)) return 7;
  int32_t gpu_off[2] = {0, 1}, model[1] = {0}, used[3] = {1, 0, 0}, fr[3] = {0, 0, 0};
  uint8_t valid[1] = {1};
  if (nosplan_set_mig_snapshot(h, 1, gpu_off, valid, model, used, fr)) return 8;
  uint8_t kind[1] = {0}, gate[1];
  int32_t doff[2] = {0, 1}, dkey[1] = {0}, dqty[1] = {2}, rep[1], node[1], slot[1];
  if (nosplan_plan(h, 1, kind, doff, dkey, dqty, NOSPLAN_PLAN_SCHED, gate, rep, node, slot)) { puts(nosplan_last_error(h)); return 9; }
  if (nosplan_get_mig_state(h, used, fr)) return 10;
  nosplan_destroy(h);
  /* geometry 1g.6gb:4 applied (used 1 + free 3), then the pod (2 devices) placed on it */
  return (rep[0] == 0 && node[0] == 0 && slot[0] == 0 && used[0] == 3 && fr[0] == 1) ? 0 : 11;
}
