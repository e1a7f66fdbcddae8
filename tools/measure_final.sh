# usage (under gpurun, one GPU): bash tools/measure_final.sh <tag>   default bench + ncu launch list of the same command
T=$1
python bench.py > gpurun_out/bench_4box_$T.json 2> gpurun_out/bench_4box_$T.err || exit 1
tail -c 400 gpurun_out/bench_4box_$T.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_4box_$T.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launches_$T.log 2>&1
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref_$T.json 2>/dev/null
python bench.py --boxes 7 > gpurun_out/bench_box7_$T.json 2>/dev/null; tail -c 300 gpurun_out/bench_box7_$T.json
