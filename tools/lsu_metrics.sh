# usage: tools/lsu_metrics.sh <tag>   (under gpurun, ONE GPU) -- where the LSU pipe of dependency_kernel goes, every launch of one
# FULL-SIZE step (boxes 55 / 35 / 17), metrics only (a few replay passes per launch instead of the ~45 of --set full)
T=$1
M=l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum.pct_of_peak_sustained_elapsed,l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum,smsp__inst_executed_op_shared_atom.sum,sm__issue_active.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
ncu --metrics $M --clock-control none -k regex:dependency_kernel -c 8 --csv --log-file gpurun_out/dep_lsu_fullhd_$T.csv \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_lsu_$T.log 2>&1
tail -1 gpurun_out/ncu_lsu_$T.log | cut -c1-200
