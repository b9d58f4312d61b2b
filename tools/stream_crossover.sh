#!/bin/bash
# per-call time of the MIG scorer for a few rows: streaming kernel (forced) against the ordered kernel (forced), at
# a node list far larger than L2 and at the planner's usual size.  One GPU.
for n in 1000000 2048; do for d in 1 6 12 24 48 96 192; do for knob in 100000 0; do
  NOSPLAN_STREAM_ROWS=$knob python tools/stream_sweep.py $n $d 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('nodes %8d rows %4d %s ms %.4f  evals/s %.3e' % (d['nodes'], d['demands'], 'stream ' if '$knob'!='0' else 'ordered', d['ms'], d['evals_per_s']))"
done; done; done
