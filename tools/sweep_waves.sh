#!/bin/bash
# kernel time of the two scorers (alone, CUDA events) for shard sizes x waves of CTAs per launch.  One GPU.
for so in 1 2 4 8; do for w in 1 2 3 4 6; do
  NOSPLAN_SCORE_WAVES=$w python bench.py --steps 10 --warmup 3 --no-plan --no-cpu-baseline --shard-of $so 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); r=d['roofline']
print('shard-of $so waves $w pass_ms %.4f mig_alone %.4f mps_alone %.4f both_in %.4f e2e %.4f' % (d['ms_per_step'], r['kernel_ms'], r['mps']['kernel_ms'], r['scoring_launch_ms_in_step'], d['e2e']['ms_per_step']))"
done; done
