for cap in 0 17 14 12 10 8; do for b in 1000 500 250 125; do
  if [ $cap = 0 ]; then unset VO_STENCIL_CAP; else export VO_STENCIL_CAP=$cap; fi
  python tools/variant_bench.py --child --batch $b --passes 5 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('cap $cap batch $b step', d['step_ms'], 'stencil_fused_alone', d['stencil_fused_ms'], d['sha_sequence'])"
done; done
