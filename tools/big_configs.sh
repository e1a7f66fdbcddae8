# usage: tools/big_configs.sh <tag>   (under gpurun, ONE GPU) -- BASELINE configs[3] and [4] on one GPU: one timed step each
T=$1
for c in 4k16 stress64; do
  python bench.py --config $c --steps 1 --warmup 0 --no-cpu-baseline --no-parity --no-e2e > gpurun_out/bench_${c}_n1_$T.json 2> gpurun_out/bench_${c}_$T.err || tail -2 gpurun_out/bench_${c}_$T.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_${c}_n1_$T.json')); print('$c', round(d['ms_per_step']), 'ms', round(d['value'],2), 'Msamples/s meanN', round(d['mean_neighbourhood']), 'nan', d['nan_channels'], {k: round(v['ms_per_step']['max']) for k, v in d['kernels'].items()})"
done
