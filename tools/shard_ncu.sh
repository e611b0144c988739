#!/bin/bash
# whole frame against one shard of 8 on one GPU: the same kernel's counters (see tools/shard_probe.py)
ncu --clock-control none -k regex:render_kernel --launch-skip 2 --launch-count 2 \
  --metrics gpu__time_duration.sum,sm__inst_executed.sum,smsp__thread_inst_executed.sum,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum,l1tex__data_pipe_lsu_wavefronts.sum,lts__t_sectors.sum,smsp__warps_active.avg.per_cycle_active,smsp__issue_active.avg.pct,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum \
  --csv --log-file gpurun_out/r2e_shard_ncu.csv python tools/shard_probe.py scene6 3840 2160 8 -g path -n 64 > gpurun_out/r2e_shard_ncu.log 2>&1
tail -2 gpurun_out/r2e_shard_ncu.log
