#!/bin/bash
# File-to-file rate of d2p_convert_file_ex through the native CLI on a RAM-backed file system: a 2 GiB raw DSD256 stereo
# payload (about 32 slices of 64 MiB) -> 24-bit / 352.8 kHz WAV, with the slice overlap (reader and writer helpers) on
# and off (D2P_FILE_SERIAL=1).  Prints the reference-style summary line of each run.
make -C tools -s d2p_cli
D=/dev/shm/d2p_file_rate; mkdir -p $D
python - <<PY
import numpy as np
np.random.default_rng(1).integers(0, 256, size=2 * 4096 * 262144, dtype=np.uint8).tofile("$D/in.dsd")   # 761 s of DSD256 stereo
PY
run() {
  local t0=$(date +%s.%N)
  tools/d2p_cli -c 2 -f P -e L -i 4 -r 352800 -b 24 -d T --seed 5 -o W -p $D/$1 $D/in.dsd 2>&1 | tail -1
  python -c "import sys; dt = $(date +%s.%N) - $t0; print('  %.3f s wall, %.2f GB/s in, %.3e DSD samples/s' % (dt, 2147483648 / dt / 1e9, 2147483648 * 8 / dt))"
}
mkdir -p $D/a $D/b
tools/d2p_cli -c 2 -f P -e L -i 4 -r 352800 -b 24 -d T --seed 5 -o W -p $D/a $D/in.dsd > /dev/null 2>&1     # warm: CUDA context, page cache
echo "overlapped (default):"; run a; run a
echo "serial (D2P_FILE_SERIAL=1):"; D2P_FILE_SERIAL=1 run b; D2P_FILE_SERIAL=1 run b
cmp $D/a/in.wav $D/b/in.wav && echo "outputs identical ($(stat -c %s $D/a/in.wav) bytes)"
rm -rf $D
