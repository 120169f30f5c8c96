run() { python bench.py --steps 20 --warmup 3 --no-cpu-baseline | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['stages']; print('$1', d['ms_per_step'], 'fwd', s['gsb_rasterize_fwd_cap']['ms_per_step'], 'bwd', s['gsb_rasterize_bwd']['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])"; }
