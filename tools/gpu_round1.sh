set -x
mkdir -p gpurun_out
TAG=${1:-v2}
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -15 gpurun_out/pytest_gpu_$TAG.log
for cfg in "--dsd-rate 4 --out-rate 352800" "--dsd-rate 4 --out-rate 352800 --dither T" "--dsd-rate 4 --out-rate 176400 --bits 20" "--dsd-rate 8 --out-rate 352800 --channels 6 --layout I --msb-first" "--dsd-rate 2 --out-rate 88200 --filter C" "--dsd-rate 1 --out-rate 352800" "--dsd-rate 4 --out-rate 88200" "--dsd-rate 1 --out-rate 88200"; do
  echo "== $cfg"; timeout 300 python tools/run_path.py $cfg --seconds 120 2>&1 | tail -1
done > gpurun_out/paths_$TAG.txt 2>&1
cat gpurun_out/paths_$TAG.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lutfir -c 1 -o gpurun_out/lutfir_${TAG}_full python tools/run_path.py --dsd-rate 4 --out-rate 352800 --seconds 120 --reps 1 --warmup 0 > gpurun_out/ncu_$TAG.log 2>&1
ls -la gpurun_out
