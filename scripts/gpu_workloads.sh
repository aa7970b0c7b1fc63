#!/bin/bash
# On the GPU box: the secondary workloads of bench.py on one GPU (BASELINE.json configs[2..4] + the AdaGrad target).
mkdir -p gpurun_out
for w in ${@:-adagrad fm16 adam26 fm64 pa1 arow}; do
  timeout 600 python bench.py --workload $w --steps 8 --warmup 3 --no-cpu --no-sweep > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
  echo -n "$w rc=$? "; python - $w <<'PY'
import json,sys
w=sys.argv[1]
try:
    j=json.loads([l for l in open(f'gpurun_out/bench_{w}.json') if l.startswith('{')][0])
    print(round(j['value']/1e6,1), round(j['ms_per_step'],3), 'e2e', round(j['e2e']['value']/1e6,1), 'frac', round(j['roofline']['pipeline_frac'],3), j['roofline']['kernel'])
except Exception as e:
    print('ERR', e)
PY
done
