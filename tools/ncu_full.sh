# usage: tools/ncu_full.sh <kernel regex> <tag> [box]   (under gpurun, ONE GPU) -- one `--set full` capture with source of the first launch of a
# kernel at 1280x720x8, one box size; the report comes back in gpurun_out/ for `ncu -i ... --page details|source --csv` here
K=$1; T=$2; B=${3:-55}
ncu --set full --import-source on --clock-control none -k regex:$K -c 1 -f -o gpurun_out/full_$T \
    python bench.py --config sample720 --boxes $B --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_full_$T.log 2>&1
ls -la gpurun_out/full_$T.ncu-rep
