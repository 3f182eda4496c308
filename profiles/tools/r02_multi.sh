#!/bin/bash
# profiles/tools/r02_multi.sh N -- multi-GPU checks and bench lines of round 2 (gpurun --gpus N -- 'bash profiles/tools/r02_multi.sh N')
set -u
N=${1:-2}
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi -L | head -8
if [ "${ONLY:-}" = "walker" ]; then      # the walker-sharded modes north_star words (all-gather) and its row-gather variant, one launch
  timeout 600 $TR --master-port 29617 bench.py --gpus $N --shard-list peer,allgather --steps 6 --warmup 3 > $O/r02_bench_c4_n${N}_walker_shards.jsonl 2> $O/bench_walker_n$N.err
  tail -c 600 $O/bench_walker_n$N.err
  python - $O/r02_bench_c4_n${N}_walker_shards.jsonl <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l); print('shard %s value %.4g ms/step %.4f accepted %s' % (d['config']['shard'], d['value'], d['ms_per_step'], d['acceptance']['accepted']))
PY
  echo done; exit 0
fi
if [ "${ONLY:-}" = "check" ]; then       # regression check of a later tree: the four shard modes against the oracle + one DIM line
  timeout 900 $TR --master-port 29611 tests/multigpu_check.py > $O/r02_multigpu_check_n${N}_final.log 2>&1; tail -4 $O/r02_multigpu_check_n${N}_final.log
  timeout 600 $TR --master-port 29613 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu > $O/r02_bench_c4_n${N}_dim_final.json 2> $O/bench_n$N.err
  python -c "
import json
d=json.loads([l for l in open('$O/r02_bench_c4_n${N}_dim_final.json') if l.startswith('{')][-1]); print('value %.4g ms/step %.4f e2e %.4g accepted %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['acceptance']['accepted']))"
  echo done; exit 0
fi
timeout 900 $TR --master-port 29611 tests/multigpu_check.py > $O/r02_multigpu_check_n$N.log 2>&1; tail -4 $O/r02_multigpu_check_n$N.log
timeout 600 $TR --master-port 29612 bench.py --gpus $N --steps 20 --warmup 5 > $O/r02_bench_c4_n${N}_dim.json 2> $O/bench_n$N.err; tail -c 400 $O/bench_n$N.err
timeout 600 $TR --master-port 29613 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu > $O/r02_bench_c4_n${N}_dim_b.json 2>> $O/bench_n$N.err
if [ "${QUICK:-}" = "1" ]; then
  for f in $O/r02_bench_c4_n${N}_dim.json $O/r02_bench_c4_n${N}_dim_b.json; do python -c "
import json,sys
d=json.loads([l for l in open('$f') if l.startswith('{')][-1]); print('$f value %.4g ms/step %.4f e2e %.4g accepted %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['acceptance']['accepted']))"; done
  echo done; exit 0
fi
timeout 600 $TR --master-port 29614 bench.py --gpus $N --shard peer --steps 10 --warmup 3 --no-cpu > $O/r02_bench_c4_n${N}_peer.json 2>> $O/bench_n$N.err
timeout 600 $TR --master-port 29615 bench.py --gpus $N --shard allgather --steps 6 --warmup 3 --no-cpu > $O/r02_bench_c4_n${N}_allgather.json 2>> $O/bench_n$N.err
timeout 600 $TR --master-port 29616 bench.py --gpus $N --workload c5 --steps 5 --warmup 3 --no-cpu > $O/r02_bench_c5_n${N}_data.json 2>> $O/bench_n$N.err
for f in $O/r02_bench_c*_n${N}_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
    print(sys.argv[1].split('/')[-1], 'value %.4g ms/step %.4f e2e %.4g shard %s accepted %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['shard'], d['acceptance']['accepted']))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
done
tail -5 $O/bench_n$N.err
echo done
