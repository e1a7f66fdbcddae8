# usage: tools/ncu_dep.sh <box> <tag>   (run under gpurun; captures the dependency kernel at 320x180)
B=$1; T=$2
CMD="python bench.py --steps 1 --warmup 0 --width 320 --height 180 --boxes $B --no-cpu-baseline --no-e2e --no-parity"
$CMD > gpurun_out/plain_$T.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:dependency -c 1 -o gpurun_out/$T $CMD > gpurun_out/ncu_$T.log 2>&1
tail -1 gpurun_out/ncu_$T.log
