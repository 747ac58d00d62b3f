#!/bin/bash
# ncu evidence of the bench command (run under gpurun, one GPU): (1) the launch list of one bench run with device time,
# DRAM bytes, tensor-pipe activity, issue utilisation and shared-memory wavefronts per launch; (2) a `--set full` capture
# with source correlation of the halo and fused up-convolution kernels (one step's worth of launches).
# Usage: bash tools/gpu_profile.sh <tag>
tag=${1:-r02}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-extras"
METRICS=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_elapsed,sm__inst_executed.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics $METRICS --clock-control none -c 700 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
echo "launch list rc $?"
$CMD > gpurun_out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"halo_kernel|halo2_kernel|halo_pair_kernel|upconv_fused" -s 18 -c 10 \
    -o gpurun_out/${tag}_convs $CMD > gpurun_out/${tag}_ncu2.log 2>&1
echo "full capture rc $?"
ls -la gpurun_out | head -20
