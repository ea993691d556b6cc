for i in 1 2 3; do
  for o in forward snake; do
    RNB_B200_CLIP_ORDER=$o python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$o', round(d['ms_per_step']*1000,1), round(d['value']), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
  done
done
