#!/bin/bash
# A/B of a compile-time knob through bench.py on one box: MACRO=UHDR_X VALUES="3 4" FILTER=scale bash scripts/bench_macro_ab.sh
cd "$(dirname "$0")/.." || exit 1
for v in $VALUES; do
  UHDR_NVCC_EXTRA="-D$MACRO=$v" python -m libultrahdr_dev_b200.build --force > /dev/null 2>&1 || { echo build failed; continue; }
  python bench.py --no-cpu-baseline --no-e2e --steps 5 2>/dev/null | F="$FILTER" python -c "
import os, sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
o = d.get('other_stages') or d['config'].get('other_stages')
print('$MACRO=$v', 'step', round(d['roofline']['frac'], 3), ' '.join('%s %.3f' % (k.replace('apply_1010102_hlg_', ''), o[k].get('frac_of_hbm_peak', 0)) for k in o if any(f in k for f in os.environ['F'].split(','))))
"
done
python -m libultrahdr_dev_b200.build --force > /dev/null 2>&1
