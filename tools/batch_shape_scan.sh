#!/bin/bash
# GPU-box tool: headline value / e2e for several (frames per batch, batches in flight) shapes.
for fd in "32 4" "37 4" "24 6" "16 8" "74 2" "48 3" "64 2"; do
  set -- $fd
  timeout 120 python bench.py --only euroc_sequence --no-sweep --no-cpu-baseline --steps 100 --warmup 5 --frames-per-batch $1 --in-flight $2 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('F %s D %s value %.0f e2e %.0f latency %s' % ('$1', '$2', d['value'], d['e2e']['value'], d['extra']['single_frame_latency']['ms']))"
done
