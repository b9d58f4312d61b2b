/* A plain-C client of both ABIs: proves the headers are valid C and that the libraries link and answer
 * without Python in the way.  Exit code 0 = ok.  With no CUDA device nosplan_create must FAIL (no fallback). */
#include <stddef.h>
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
  /* geometry 1g.6gb:4 applied (used 1 + free 3), then the pod (2 devices) placed on it */
  if (!(rep[0] == 0 && node[0] == 0 && slot[0] == 0 && used[0] == 3 && fr[0] == 1)) return 11;

  /* MPS: one 40 GB GPU with two used 10gb slices, request 20gb:1 -> {10gb:2, 20gb:1} (slicing/gpu_test.go:265) */
  int32_t sizes[2] = {10, 20}, mem[1] = {40}, pused[2] = {2, 0}, pfree[2] = {0, 0};
  uint32_t pmask[1] = {0};
  if (nosplan_set_mps_snapshot(h, 2, sizes, 1, gpu_off, valid, mem, pused, pfree, pmask)) { puts(nosplan_last_error(h)); return 12; }
  int32_t poff[2] = {0, 1}, pkey[1] = {1}, pqty[1] = {1}, pfirst[1] = {-7};
  uint8_t pupd[1] = {9}, out_free[2] = {9, 9};
  uint32_t out_mask[1] = {0};
  int64_t evals = 0;
  if (nosplan_score_mps(h, 1, poff, pkey, pqty, pfirst, pupd, out_free, out_mask, &evals)) { puts(nosplan_last_error(h)); return 13; }
  if (!(pfirst[0] == 0 && pupd[0] == 1 && out_free[0] == 0 && out_free[1] == 1 && out_mask[0] == 2u && evals == 1)) return 14;
  uint8_t pkind[1] = {1};
  if (nosplan_plan(h, 1, pkind, poff, pkey, pqty, NOSPLAN_PLAN_GATE | NOSPLAN_PLAN_SCHED, gate, rep, node, slot)) return 15;
  if (nosplan_get_mps_state(h, pused, pfree, pmask)) return 16;
  /* 20gb absent -> no gate; the slice is created and the pod takes it: used {10gb:2, 20gb:1}, free key 20gb present with 0 */
  if (!(gate[0] == 0 && rep[0] == 0 && node[0] == 0 && pused[1] == 1 && pfree[1] == 0 && (pmask[0] & 2u))) return 17;

  /* the one-call scoring pass: both snapshots in from host arrays, winners out; nothing is installed on the handle */
  int32_t mused[3] = {1, 0, 0}, mfree[3] = {0, 0, 0}, m2used[2] = {2, 0}, m2free[2] = {0, 0}, mfirst[1] = {-7};
  int8_t mchoice[1] = {-7};
  nosplan_mig_snapshot_t ms = {1, gpu_off, valid, model, mused, mfree};
  nosplan_mps_snapshot_t ps = {2, sizes, 1, gpu_off, valid, mem, m2used, m2free, NULL};
  nosplan_demands_t mrows = {1, doff, dkey, dqty}, prows = {1, poff, pkey, pqty};
  pfirst[0] = -7;
  if (nosplan_score_snapshot(h, &ms, &mrows, mfirst, mchoice, &ps, &prows, pfirst, &evals)) { puts(nosplan_last_error(h)); return 18; }
  if (!(mfirst[0] == 0 && mchoice[0] == 3 && pfirst[0] == 0 && evals == 4 + 1)) return 19;  /* geometry index 3 = {1g.6gb:4} */

  /* a one-rank exchange: the winners come back unchanged (plus the node base); then the same pass through it */
  unsigned char ipc[NOSPLAN_IPC_HANDLE_BYTES];
  if (nosplan_exchange_create(h, 0, 1, 16, ipc) || nosplan_exchange_open(h, ipc)) { puts(nosplan_last_error(h)); return 20; }
  if (nosplan_set_node_base(h, 100, 200)) return 21;
  mfirst[0] = pfirst[0] = -7;
  if (nosplan_score_snapshot(h, &ms, &mrows, mfirst, NULL, &ps, &prows, pfirst, NULL)) { puts(nosplan_last_error(h)); return 22; }
  if (!(mfirst[0] == 100 && pfirst[0] == 200)) return 23;
  if (nosplan_exchange_status(h) || nosplan_exchange_destroy(h)) return 24;
  nosplan_destroy(h);
  return 0;
}
