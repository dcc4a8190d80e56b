#!/bin/bash
# tools/gpu_run.sh TAG [sections]  -- one gpurun call: tests, microbench, path timings, bench, ncu.
# sections: any of t(ests) m(icrobench) p(aths) b(ench) n(cu lutfir) l(cu lm) ; default "tmpb"
TAG=${1:-dev}; SEC=${2:-tmpb}
mkdir -p gpurun_out
if [[ $SEC == *t* ]]; then
  python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
  tail -8 gpurun_out/pytest_gpu_$TAG.log
fi
if [[ $SEC == *m* ]]; then ./tools/microbench > gpurun_out/microbench_$TAG.json 2>&1; cat gpurun_out/microbench_$TAG.json; fi
if [[ $SEC == *p* ]]; then
  for cfg in "--dsd-rate 4 --out-rate 352800" "--dsd-rate 4 --out-rate 352800 --dither T" "--dsd-rate 4 --out-rate 176400 --bits 20" \
     "--dsd-rate 8 --out-rate 352800 --channels 6 --layout I --msb-first" "--dsd-rate 2 --out-rate 88200 --filter C" \
     "--dsd-rate 1 --out-rate 352800" "--dsd-rate 4 --out-rate 88200" "--dsd-rate 1 --out-rate 88200" \
     "--dsd-rate 4 --out-rate 192000 --bits 32" "--dsd-rate 4 --out-rate 96000 --bits 32" "--dsd-rate 1 --out-rate 96000" \
     "--dsd-rate 8 --out-rate 384000 --channels 6 --layout I --msb-first"; do
    echo "== $cfg"; timeout 300 python tools/run_path.py $cfg --seconds 120 2>&1 | tail -1
  done > gpurun_out/paths_$TAG.txt 2>&1
  grep -o '== .*\|"kernel_ms": [0-9.]*\|"dsd_samples_per_s": [0-9.e+]*' gpurun_out/paths_$TAG.txt | paste - - -
fi
if [[ $SEC == *b* ]]; then
  python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; tail -c 3000 gpurun_out/bench_$TAG.json; tail -3 gpurun_out/bench_$TAG.err
fi
if [[ $SEC == *n* ]]; then
  python tools/run_path.py --dsd-rate 4 --out-rate 352800 --seconds 120 --reps 1 --warmup 0 > gpurun_out/plain_n.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:lutfir -c 1 -f -o gpurun_out/lutfir_${TAG}_full \
     python tools/run_path.py --dsd-rate 4 --out-rate 352800 --seconds 120 --reps 1 --warmup 0 > gpurun_out/ncu_n_$TAG.log 2>&1
  tail -2 gpurun_out/ncu_n_$TAG.log
fi
if [[ $SEC == *l* ]]; then
  python tools/run_path.py --dsd-rate 4 --out-rate 192000 --bits 32 --seconds 120 --reps 1 --warmup 0 > gpurun_out/plain_l.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:lm_kernel -c 1 -f -o gpurun_out/lm_${TAG}_full \
     python tools/run_path.py --dsd-rate 4 --out-rate 192000 --bits 32 --seconds 120 --reps 1 --warmup 0 > gpurun_out/ncu_l_$TAG.log 2>&1
  tail -2 gpurun_out/ncu_l_$TAG.log
fi
ls -la gpurun_out | tail -20
if [[ $SEC == *B* ]]; then   # ncu of the bench command itself: launch list + full capture of one timed launch
  python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_plain_$TAG.json 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/bench_launches_$TAG.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_ncu1_$TAG.log 2>&1
  python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_plain2_$TAG.json 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:lutfir -s 3 -c 1 -f -o gpurun_out/bench_${TAG}_full \
      python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_ncu2_$TAG.log 2>&1
  tail -2 gpurun_out/bench_ncu2_$TAG.log
fi
