#!/bin/bash
# sweep the number of frame ranges of the host pipeline (development probe)
mkdir -p gpurun_out/e2e
for c in "$@"; do
  python bench.py --steps 3 --warmup 2 --no-cpu-baseline --e2e-chunks $c > gpurun_out/e2e/c$c.json 2>/dev/null
  python -c "import json; d=json.load(open('gpurun_out/e2e/c$c.json')); print($c, round(d['value']), round(d['e2e']['value']), round(d['e2e']['ms_per_step'],1))"
done
