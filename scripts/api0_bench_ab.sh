cd /root/repo
for cfg in "1 1" "5 1" "5 4" "5 5"; do
  set -- $cfg
  UHDR_NVCC_EXTRA="-DUHDR_API0_CONVERT_MIN_BLOCKS=$1 -DUHDR_API0_MIN_BLOCKS=$2" python -m libultrahdr_dev_b200.build --force > /dev/null 2>&1 || { echo build failed; continue; }
  python bench.py --no-cpu-baseline --no-e2e --steps 5 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
o = d.get('other_stages') or d['config'].get('other_stages')
print('convert_blocks=$1 fused_blocks=$2', 'fused', round(o['api0_fused_tonemap_generate']['frac_of_hbm_peak'], 3), 'fused+convert', round(o['api0_fused_tonemap_generate_convert601']['frac_of_hbm_peak'], 3), 'tonemap', round(o['tonemap']['frac_of_hbm_peak'], 3))
"
done
