#!/bin/bash
# e2e (pageable host planes) rate for a few staging-pool sizes and caller-thread counts; one JSON summary line per setting
cd "$(dirname "$0")/.." || exit 1
for st in 0 4 8 12 16; do
  for th in 4 6 8; do
    UHDRB200_STAGE_THREADS=$st python bench.py --steps 3 --warmup 3 --distinct 4 --batch 16 --no-cpu-baseline --no-other-stages --e2e-threads $th 2>/dev/null |
      python -c "import sys,json; d=json.loads(sys.stdin.read()); e=d['e2e']; print(json.dumps({'stage_threads': $st, 'callers': $th, 'pageable': round(e['value']), 'pinned': round(e['pinned_planes_mpix_s']), 'ceiling': round(e['link_ceiling_mpix_s'])}))"
  done
done
