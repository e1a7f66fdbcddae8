# usage: tools/measure_final_r2.sh <tag>   (under gpurun, ONE GPU) -- the committed single-GPU measurements of a build
T=$1; O=gpurun_out
Q="--no-cpu-baseline --no-e2e --no-parity"
python bench.py > $O/bench_fullhd_n1_$T.json 2> $O/bench_fullhd_n1_$T.err || { tail -5 $O/bench_fullhd_n1_$T.err; exit 1; }
python bench.py --impl reference --steps 1 --warmup 0 > $O/bench_ref_fullhd_$T.json 2> $O/bench_ref_$T.err
python bench.py --config sample720 > $O/bench_sample720_n1_$T.json 2>> $O/bench_fullhd_n1_$T.err
python bench.py --boxes 7 > $O/bench_fullhd_box7_$T.json 2>> $O/bench_fullhd_n1_$T.err
# a RENDERED feature buffer: pbrt (the plugin binary) renders tests/scenes/spheres.pbrt.in at 640x480x8 and dumps it; bench.py filters the file
sed "s|@INTEGRATOR@|Integrator \"rpf\" \"integer boxsizes\" [17 7] \"string rpfdump\" \"$PWD/$O/spheres_640x480x8.rpfb\"|; s|@W@|640|; s|@H@|480|; s|@SPP@|8|; s|@OUT@|$PWD/$O/spheres.pfm|" tests/scenes/spheres.pbrt.in > $O/spheres.pbrt
raytracer-rpf_b200/host/_build/pbrt_rpf_b200 --quiet $O/spheres.pbrt > $O/pbrt_spheres_$T.log 2>&1; grep "RPF on" $O/pbrt_spheres_$T.log
python bench.py --buffer $O/spheres_640x480x8.rpfb --boxes 17,7 > $O/bench_buffer_spheres_$T.json 2>> $O/bench_fullhd_n1_$T.err
rm -f $O/spheres_640x480x8.rpfb
# launch list of the default workload
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_fullhd_$T.csv python bench.py --steps 2 --warmup 1 $Q > $O/ncu_launches_$T.log 2>&1
# DRAM traffic of every general-path kernel: one full-size step, and the middle band of 2 / 4 / 8 alone on this GPU
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"dependency_kernel|blend_kernel|ordered_stats|membership" -c 32 --csv \
    --log-file $O/traffic_all_fullhd_$T.csv python bench.py --steps 1 --warmup 0 $Q > $O/ncu_traffic_$T.log 2>&1
for B in 1/2 1/4 3/8; do
  N=${B#*/}
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:dependency_kernel -c 12 --csv \
      --log-file $O/traffic_band_n${N}_$T.csv python bench.py --band $B --steps 1 --warmup 0 $Q > $O/ncu_band_n${N}_$T.log 2>&1
done
python - <<PY
import json
for f in ("bench_fullhd_n1", "bench_sample720_n1", "bench_fullhd_box7", "bench_buffer_spheres"):
    try:
        d = json.load(open("$O/%s_$T.json" % f))
        print(f, round(d["ms_per_step"], 1), "ms", round(d["value"], 2), "Msamples/s e2e", round(d["e2e"]["value"], 2), "parity", d["parity"] and d["parity"]["pass"], d["parity_sigma2"] and d["parity_sigma2"]["pass"], "nan", d["nan_channels"], "meanN", round(d["mean_neighbourhood"], 1))
    except Exception as e:
        print(f, "FAILED", e)
d = json.load(open("$O/bench_ref_fullhd_$T.json")); print("reference arm", d["value"], d["cpu_baseline"]["cores"], d["ms_per_step"])
PY
