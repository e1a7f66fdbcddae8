# usage (under gpurun, one GPU): bash tools/measure_n1.sh <tag>
# 1. default bench (BASELINE.json configs[1]) -> gpurun_out/bench_4box_<tag>.json
# 2. ncu launch list of the same command -> gpurun_out/launches_4box_<tag>.csv
# 3. ncu --set full of the four dependency_kernel launches of one step at full size -> gpurun_out/dep_full_<tag>.ncu-rep
T=$1
python bench.py > gpurun_out/bench_4box_$T.json 2> gpurun_out/bench_4box_$T.err || exit 1
tail -c 600 gpurun_out/bench_4box_$T.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_4box_$T.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launches_$T.log 2>&1
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e"
ncu --set full --clock-control none --import-source on -k regex:dependency_kernel -c 4 -o gpurun_out/dep_full_$T $CMD > gpurun_out/ncu_dep_full_$T.log 2>&1
tail -2 gpurun_out/ncu_dep_full_$T.log
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref_$T.json 2>/dev/null; tail -c 400 gpurun_out/bench_ref_$T.json
