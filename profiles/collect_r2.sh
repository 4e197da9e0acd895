#!/bin/bash
# Collects the round-2 evidence on one B200 (run under gpurun from the repo root).  Every ncu pass profiles a command
# that has already exited 0 without ncu in the same call.  Outputs land in gpurun_out/; the summaries are copied to profiles/.
set -u
O=gpurun_out
mkdir -p $O
python profiles/per_launch.py bf16 512 > $O/r2_per_launch_bf16.txt 2>&1 || exit 1
python profiles/per_launch.py fp32 512 > $O/r2_per_launch_fp32.txt 2>&1 || exit 1
python bench.py --steps 2 --warmup 3 --no-dream --no-cpu-baseline --no-extras > $O/r2_bench_plain.json 2>/dev/null || exit 1
# launch list of the bench command (per-launch times are cold-cache and serialised: shares, not absolutes)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_ncu_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-dream --no-cpu-baseline --no-extras > $O/r2_ncu_bench.log 2>&1
# one 512-projection step: time, DRAM bytes, tensor-pipe activity, L2 hit rate per launch
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
    --clock-control none --csv --log-file $O/r2_ncu_step512.csv python profiles/per_launch.py bf16 512 > $O/r2_ncu_step.log 2>&1
# full sections of the fused tail kernel (one 64-projection step)
ncu --set full --clock-control none --import-source on -k regex:"deconv_strip" -c 1 -o $O/r2_ncu_full_strip -f \
    python profiles/per_launch.py bf16 64 > $O/r2_ncu_full.log 2>&1
ncu -i $O/r2_ncu_full_strip.ncu-rep --page raw --csv > $O/r2_ncu_full_strip_raw.csv 2>/dev/null
ls -la $O | tail -20
