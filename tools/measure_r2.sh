# usage: tools/measure_r2.sh <tag>   (under gpurun, ONE GPU) -- the round's committed measurements of the default workload
# (BASELINE.json configs[2]: 1920x1080x8, boxes 55/35/17/7). Numbers printed under ncu are never bench values: the bench
# line comes from the plain run in step 1.
T=$1; O=gpurun_out
Q="--no-cpu-baseline --no-e2e --no-parity"
# 1. the bench line (plain run)
python bench.py > $O/bench_fullhd_n1_$T.json 2> $O/bench_fullhd_n1_$T.err || { tail -5 $O/bench_fullhd_n1_$T.err; exit 1; }
# 2. launch list of the same workload: every kernel's duration, cold-cache and serialised -- the SHARES must agree with step 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_fullhd_$T.csv python bench.py --steps 2 --warmup 1 $Q > $O/ncu_launches_$T.log 2>&1
# 3. ncu --set full of every dependency_kernel launch of one step at full size (traffic + where the LSU pipe goes)
ncu --set full --clock-control none --import-source on -k regex:dependency_kernel -c 8 -o $O/dep_fullhd_$T python bench.py --steps 1 --warmup 0 $Q > $O/ncu_dep_full_$T.log 2>&1
tail -2 $O/ncu_dep_full_$T.log
# 4. DRAM traffic of the dependency launches of ONE BAND of the sharded frame (the middle band, alone on this GPU)
for B in 1/2 1/4 3/8; do
  N=${B#*/}
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:dependency_kernel -c 12 --csv \
      --log-file $O/traffic_band_n${N}_$T.csv python bench.py --band $B --steps 1 --warmup 0 $Q > $O/ncu_band_n${N}_$T.log 2>&1
done
ls -la $O | tail -15
