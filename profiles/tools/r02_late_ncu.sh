#!/bin/bash
# profiles/tools/r02_late_ncu.sh -- late round 2: `ncu --set full` of the two kernels that changed late (each only after the same command
# exited 0 plainly): stretchFusedAhead (C1) and stretchHalfStepThread with the pcg32 engine (C2).  Summaries are made on the box.
set -u
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/*.ncu-rep
B="python bench.py --workload c1 --steps 2000 --warmup 200 --no-cpu"
$B > $O/plain_c1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:stretchFusedAhead -s 1 -c 1 -o $O/r02_c1_fused_ahead $B > $O/ncu_c1.log 2>&1
B="python bench.py --workload c2 --rng per_walker32 --steps 4 --warmup 3 --no-cpu"
$B > $O/plain_c2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:stretchHalfStepThread -s 6 -c 1 -o $O/r02_c2_thread_per_walker32_late $B > $O/ncu_c2.log 2>&1
ls -la $O/*.ncu-rep
for rep in $O/*.ncu-rep; do
  pre=${rep%.ncu-rep}
  python profiles/tools/summarise_ncu.py $rep $pre > ${pre}_summary.txt 2>&1
  cat ${pre}_summary.txt | head -20
done
rm -f $O/*.ncu-rep
echo done
