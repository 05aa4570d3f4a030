# A/B helper used for the tuning sweeps under profiles/: usage (on the GPU box)  bash profiles/tools/ab_env.sh "ENV=.. ENV2=.." ...
for spec in "$@"; do for r in 1 2; do env $spec GB_BENCH_SKIP_CPU=1 python bench.py --steps 10 --warmup 3 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms']
print('$spec', '%.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'row %.3f col %.3f'%(k['acq_inv_row']['total_ms'],k['acq_inv_col']['total_ms']))"; done; done
