# usage: tools/ncu_sections.sh <kernel regex> <tag> [box] [lib variant]   (under gpurun, ONE GPU) -- speed-of-light, memory, warp-state,
# scheduler and occupancy sections (a dozen replay passes, not the ~45 of --set full) of the FIRST launch of a kernel at 1280x720x8
K=$1; T=$2; B=${3:-55}
if [ -n "$4" ]; then export RPF_B200_LIB=$PWD/raytracer-rpf_b200/lib/librpf_b200_$4.so; fi
ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section Occupancy --section SchedulerStats --section LaunchStats \
    --clock-control none -k regex:$K -c 1 --log-file gpurun_out/sections_${T}.txt \
    python bench.py --config sample720 --boxes $B --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_sections_$T.log 2>&1
grep -E "Duration|Throughput|Hit Rate|Issue|Eligible|Achieved Occ|Theoretical Occ|Registers Per|Cycles Per Issued" gpurun_out/sections_${T}.txt | head -40
