#!/bin/bash
for so in 8 4; do for t in 128 64; do for c in 16 32 64; do
  NOSPLAN_MIG_TILE=$t NOSPLAN_MIN_CHUNK=$c python bench.py --steps 10 --warmup 3 --no-plan --no-cpu-baseline --shard-of $so 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); r=d['roofline']
print('shard-of $so tile $t minchunk $c pass_ms %.4f mig_alone %.4f mps_alone %.4f both_in %.4f' % (d['ms_per_step'], r['kernel_ms'], r['mps']['kernel_ms'], r['scoring_launch_ms_in_step']))"
done; done; done
