#!/bin/bash
# profiles/tools/r02_capture.sh -- what the builder ran on the GPU box in round 2 (gpurun -- 'bash profiles/tools/r02_capture.sh <stage>').
# Numbers printed by a command under ncu are never bench values; the bench lines come from the plain runs.
set -u
mkdir -p gpurun_out
O=gpurun_out
stage=${1:-all}

if [ "$stage" = "tests" ] || [ "$stage" = "all" ]; then
  timeout 1500 python -m pytest tests -m gpu -q -rf --deselect tests/test_gpu_multi.py > $O/pytest_gpu.log 2>&1; tail -15 $O/pytest_gpu.log
fi

if [ "$stage" = "sweep" ] || [ "$stage" = "all" ]; then
  : > $O/sweep.log
  for cfg in "8 2 2" "16 2 2" "8 3 2" "16 3 1" "8 4 1" "16 4 2" "8 6 1" "32 2 1"; do set -- $cfg
    MCMCPP_B200_DIM_STAGE_KB=$1 MCMCPP_B200_DIM_PREFETCH=$2 MCMCPP_B200_DIM_LAG=$3 MCMCPP_BENCH_D=126 timeout 120 python bench.py --shard dim --no-cpu --steps 20 --warmup 5 2>>$O/sweep.err |
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print('D126 stage/P/L $cfg ms/step %.4f' % d['ms_per_step'])" >> $O/sweep.log
  done
  for d in 250 500; do for cfg in "16 2 2" "16 3 1" "32 2 1" "32 3 1"; do set -- $cfg
    MCMCPP_B200_DIM_STAGE_KB=$1 MCMCPP_B200_DIM_PREFETCH=$2 MCMCPP_B200_DIM_LAG=$3 MCMCPP_BENCH_D=$d timeout 120 python bench.py --shard dim --no-cpu --steps 20 --warmup 5 2>>$O/sweep.err |
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print('D$d stage/P/L $cfg ms/step %.4f' % d['ms_per_step'])" >> $O/sweep.log
  done; done
  for shape in 2x8 4x4; do for pass in 0 1; do
    MCMCPP_B200_DENSE_SHAPE=$shape MCMCPP_B200_DENSE_PASS_EARLY=$pass timeout 120 python bench.py --workload c3 --no-cpu --steps 400 --warmup 20 2>>$O/sweep.err |
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print('C3 shape $shape passEarly $pass us/colour %.3f frac %.3f' % (500*d['ms_per_step'], d['roofline']['frac']))" >> $O/sweep.log
  done; done
  for rng in sequential per_walker32; do
    timeout 120 python bench.py --workload c2 --rng $rng --no-cpu --steps 2000 --warmup 50 2>>$O/sweep.err |
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print('C2 rng $rng value %.4g us/colour %.3f' % (d['value'], 500*d['ms_per_step']))" >> $O/sweep.log
  done
  cat $O/sweep.log
fi

if [ "$stage" = "bench" ] || [ "$stage" = "all" ]; then
  timeout 500 python bench.py --steps 20 --warmup 5 > $O/bench_c4.json 2> $O/bench_c4.err; tail -c 300 $O/bench_c4.err
  timeout 200 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2>> $O/bench_c4.err
fi

if [ "$stage" = "ncu" ] || [ "$stage" = "all" ]; then
  rm -f $O/*.ncu-rep
  B="python bench.py --steps 2 --warmup 3 --no-cpu --no-others"
  $B > $O/plain_c4.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/r02_c4_launches.csv $B > $O/ncu_c4_list.log 2>&1
  $B > $O/plain_c4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:stretchHalfStepBulk -s 4 -c 1 -o $O/r02_c4_bulk $B > $O/ncu_c4.log 2>&1
  export MCMCPP_BENCH_D=126
  B="python bench.py --shard dim --steps 2 --warmup 3 --no-cpu"
  $B > $O/plain_dim.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:dimShardFused -s 4 -c 1 -o $O/r02_dim_fused_d126 $B > $O/ncu_dim.log 2>&1
  unset MCMCPP_BENCH_D
  for shape in 2x8 4x4; do
    export MCMCPP_B200_DENSE_SHAPE=$shape
    B="python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu"
    $B > $O/plain_c3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:denseGaussHalfStep -s 4 -c 1 -o $O/r02_c3_dense_$shape $B > $O/ncu_c3_$shape.log 2>&1
  done
  unset MCMCPP_B200_DENSE_SHAPE
  for rng in sequential per_walker32; do
    B="python bench.py --workload c2 --rng $rng --steps 4 --warmup 3 --no-cpu"
    $B > $O/plain_c2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:stretchHalfStepThread -s 6 -c 1 -o $O/r02_c2_thread_$rng $B > $O/ncu_c2_$rng.log 2>&1
  done
  ls -la $O/*.ncu-rep
  # summaries are made on the box (the reports carry the source and are ~30 MB each; gpurun_out/ comes back only below 64 MiB)
  for rep in $O/*.ncu-rep; do
    pre=${rep%.ncu-rep}
    python profiles/tools/summarise_ncu.py $rep $pre > ${pre}_summary.txt 2>&1
  done
  if [ "${KEEP_REPS:-}" != "" ]; then for rep in $O/*.ncu-rep; do case "$rep" in *${KEEP_REPS}*) ;; *) rm -f $rep;; esac; done; else rm -f $O/*.ncu-rep; fi
fi
echo done
