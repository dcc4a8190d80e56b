#!/bin/bash
# tools/gpu_profile.sh TAG -- the evidence run of a round, one gpurun call (1 GPU):
#   bench line, ncu launch list of the SAME bench command, one full capture of a timed lutfir launch of it (DRAM traffic,
#   shared-memory wavefronts, stalls), one full capture of lm_kernel (config 3b).  Everything lands in gpurun_out/.
TAG=${1:-dev}
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err || exit 1
tail -c 600 gpurun_out/bench_$TAG.json
# launch list: every launch of the bench command with its device time (shares, not absolutes)
python bench.py --steps 2 --warmup 3 --no-cpu --no-paths > gpurun_out/bench_plain_$TAG.json 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/bench_launches_$TAG.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-paths > gpurun_out/bench_ncu1_$TAG.log 2>&1
# full capture of one wave launch of the same command (skip the warm-up launches)
python bench.py --steps 1 --warmup 3 --no-cpu --no-paths --total-files 64 > gpurun_out/bench_plain2_$TAG.json 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:lutfir -s 12 -c 1 -f -o gpurun_out/bench_${TAG}_full \
    python bench.py --steps 1 --warmup 3 --no-cpu --no-paths --total-files 64 > gpurun_out/bench_ncu2_$TAG.log 2>&1
tail -2 gpurun_out/bench_ncu2_$TAG.log
python tools/run_path.py --dsd-rate 4 --out-rate 192000 --bits 32 --seconds 120 --reps 1 --warmup 0 > gpurun_out/plain_l.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lm_kernel -c 1 -f -o gpurun_out/lm_${TAG}_full \
    python tools/run_path.py --dsd-rate 4 --out-rate 192000 --bits 32 --seconds 120 --reps 1 --warmup 0 > gpurun_out/ncu_l_$TAG.log 2>&1
tail -2 gpurun_out/ncu_l_$TAG.log
ls -la gpurun_out | tail -12
