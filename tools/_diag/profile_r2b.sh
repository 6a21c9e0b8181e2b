#!/bin/bash
# Launch list (+ DRAM bytes) of one eager step and --set full captures of the GEMM / attention kernels (round 2b).
# The .ncu-rep files are exported to CSV on the box and deleted (gpurun_out/ is capped at 64 MiB).
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
CMD="python tools/profile_step.py --steps 2"
$CMD > gpurun_out/r2b_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r2b_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file gpurun_out/r2b_launches.csv $CMD > gpurun_out/r2b_ncu1.log 2>&1
cap() {  # name, kernel regex, skip, count
  ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c $4 -o /tmp/$1 $CMD > gpurun_out/$1.log 2>&1
  ncu -i /tmp/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null
  ncu -i /tmp/$1.ncu-rep --page details --csv > gpurun_out/$1_details.csv 2>/dev/null
}
cap r2b_gemm_fwd gemm_tc 224 8
cap r2b_gemm_bwd gemm_tc 330 12
cap r2b_attn attn_ 57 9
ncu -i /tmp/r2b_gemm_fwd.ncu-rep --page source --csv --kernel-name regex:gemm_tc --launch-skip 2 --launch-count 1 > gpurun_out/r2b_gemm_fwd_up_source.csv 2>/dev/null
du -sh gpurun_out
