#!/bin/bash
# profiles/tools/r02_final.sh -- the last one-GPU run of round 2: the whole GPU suite, C1 at 32 and 320 walkers, the default bench line and the
# reference arm, all plain runs.
set -u
mkdir -p gpurun_out
O=gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -6) > $O/final_tests.log
: > $O/final_c1.log
for w in 32 320; do
  MCMCPP_BENCH_W=$w timeout 120 python bench.py --workload c1 --no-cpu --steps 20000 --warmup 2000 2>>$O/final.err | grep "^{" |
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('C1 W=$w value %.4g us/step %.3f' % (d['value'], 1000*d['ms_per_step']))" >> $O/final_c1.log
done
timeout 600 python bench.py --steps 20 --warmup 5 > $O/r02_bench_default_n1.json 2>> $O/final.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $O/r02_bench_reference_arm.json 2>> $O/final.err
cat $O/final_tests.log $O/final_c1.log
python - <<'PY'
import json
d = json.loads([l for l in open('gpurun_out/r02_bench_default_n1.json') if l.startswith('{')][-1])
print('C4 value %.4g ms/step %.4f frac_actual %.3f e2e %.4g clocks %s' % (d['value'], d['ms_per_step'], d['roofline']['frac_actual'], d['e2e']['value'], d['clocks']))
for k, v in d['other_workloads'].items():
    print(k, v.get('error') or ('%.4g' % v['value'], v['roofline'].get('frac'), v['roofline'].get('issue', {}).get('frac'), '%.4g' % (v.get('cpu_baseline') or {'value': float('nan')})['value']))
r = json.loads([l for l in open('gpurun_out/r02_bench_reference_arm.json') if l.startswith('{')][-1]); print('reference arm %.4g' % r['value'])
PY
tail -3 $O/final.err
