#!/bin/bash
# Work-item size against shard size: pass time of bench.py's step on rank 0's shard of a W-rank run, for the rows-per-item
# floor (NOSPLAN_MIN_CHUNK, per 128 nodes) and the CTA waves the split aims at (NOSPLAN_SCORE_WAVES).  One GPU.
for so in ${SHARDS:-8 4 1}; do for c in ${CHUNKS:-64 128 256}; do for w in ${WAVES:-4 2}; do
  NOSPLAN_MIN_CHUNK=$c NOSPLAN_SCORE_WAVES=$w NOSPLAN_MPS_MIN_CHUNK=${MPSCHUNK:-128} python bench.py --steps 10 --warmup 3 --no-plan --no-cpu-baseline --shard-of $so 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); r=d['roofline']
print('shard-of $so minchunk $c waves $w pass_ms %.4f mig_alone %.4f mps_alone %.4f both_in %.4f' % (d['ms_per_step'], r['kernel_ms'], r['mps']['kernel_ms'], r['scoring_launch_ms_in_step']))"
done; done; done
