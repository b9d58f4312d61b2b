#!/bin/bash
# kernel time of the MIG scorer (alone, CUDA events) for shard sizes x nodes per CTA x waves.  One GPU.
for so in 1 8; do for t in 128 64 32; do for w in 2 4; do
  NOSPLAN_MIG_TILE=$t NOSPLAN_SCORE_WAVES=$w python bench.py --steps 10 --warmup 3 --no-plan --no-cpu-baseline --shard-of $so 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); r=d['roofline']
print('shard-of $so tile $t waves $w pass_ms %.4f mig_alone %.4f mps_alone %.4f both_in %.4f' % (d['ms_per_step'], r['kernel_ms'], r['mps']['kernel_ms'], r['scoring_launch_ms_in_step']))"
done; done; done
