#!/bin/bash
# Collects the round-1 evidence on one B200 (run under gpurun from the repo root); every ncu pass runs a command
# that has already exited 0 without ncu in the same call.  Outputs land in gpurun_out/ and are copied to profiles/.
set -u
O=gpurun_out
mkdir -p $O
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err || exit 1
python bench.py --precision fp32 --no-dream --no-cpu-baseline --steps 5 --warmup 3 > $O/bench_fp32.json 2>/dev/null || exit 1
python profiles/per_launch.py bf16 512 > $O/per_launch_bf16.txt 2>&1 || exit 1
python profiles/api_latency.py > $O/api_latency.txt 2>&1
python profiles/stream_kernels.py $O/stream.json > $O/stream.txt 2>&1
python profiles/dream_classes.py bf16 512 > $O/dream_bf16_512.txt 2>&1
python profiles/dream_classes.py fp32 512 | head -1 > $O/dream_fp32_512.txt 2>&1
# launch list of the bench command (per-launch times are cold-cache and serialised: shares, not absolutes)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/ncu_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-dream --no-cpu-baseline > $O/ncu_bench.log 2>&1
# one 512-projection step: time, DRAM bytes, tensor-pipe activity, L2 hit rate per launch
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
    --clock-control none --csv --log-file $O/ncu_step512.csv python profiles/per_launch.py bf16 512 > $O/ncu_step.log 2>&1
# full sections of the tcgen05 kernels of one (64-projection) step
ncu --set full --clock-control none --import-source on -k regex:"deconv_" --launch-skip 27 -c 9 -o $O/ncu_full_deconv -f \
    python profiles/per_launch.py bf16 64 > $O/ncu_full.log 2>&1
ncu -i $O/ncu_full_deconv.ncu-rep --page raw --csv > $O/ncu_full_deconv_raw.csv 2>/dev/null
ls -la $O
