#!/bin/bash
# profiles/tools/r02_c2_late.sh -- late round 2: the GPU suite, C2 with both engines (plain runs), then the instruction count of one
# stretchHalfStepThread launch per engine (ncu, metrics only).  Numbers printed under ncu are never bench values.
set -u
mkdir -p gpurun_out
O=gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -6) > $O/late_tests.log
: > $O/late_c2.log
for rep in 1 2; do for rng in sequential per_walker32; do
  timeout 120 python bench.py --workload c2 --rng $rng --no-cpu --steps 2000 --warmup 50 2>>$O/late.err | grep "^{" |
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('C2 rng $rng value %.4g us/colour %.3f sm_mhz %s' % (d['value'], 500*d['ms_per_step'], d['clocks'].get('sm_mhz')))" >> $O/late_c2.log
done; done
for rng in sequential per_walker32; do
  B="python bench.py --workload c2 --rng $rng --steps 4 --warmup 3 --no-cpu"
  $B > $O/plain_c2.log 2>&1 && ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum --clock-control none -k regex:stretchHalfStepThread -s 6 -c 2 --csv \
      --log-file $O/r02_c2_late_inst_$rng.csv $B > $O/ncu_c2_$rng.log 2>&1
done
cat $O/late_tests.log $O/late_c2.log; grep -h "inst_executed" $O/r02_c2_late_inst_*.csv
