cd /root/repo
for S in 256 512 1024 2048; do
  UHDR_NVCC_EXTRA="-DUHDR_JPEG_SUBSEQ_BITS=$S" python -m libultrahdr_dev_b200.build --force > /dev/null 2>&1 || { echo build failed $S; continue; }
  echo "S=$S"; python scripts/jpeg_decode_bench.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('  q%d: %.3f ms (%d Mpix/s), passes %.0f' % (d['quality'], d['decode_dev_ms'], d['mpix_s'], d['sync_passes']))"
done
