# usage (under gpurun --gpus N): bash tools/scale_run.sh N "<configs>"   -> gpurun_out/bench_<config>_n<N>.json
N=$1; shift
for c in $@; do
  S=2; W=3; if [ $c = 4k16 ] || [ $c = stress64 ]; then S=1; W=1; fi
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --config $c --steps $S --warmup $W --no-e2e > gpurun_out/bench_${c}_n$N.json 2> gpurun_out/bench_${c}_n$N.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_${c}_n$N.json')); print('$c', 'N=$N', round(d['value'],3), 'Msamples/s', round(d['ms_per_step'],1), 'ms')" || tail -5 gpurun_out/bench_${c}_n$N.err
done
