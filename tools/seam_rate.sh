#!/bin/bash
# What the drop-in seam sustains from a native caller: tools/d2p_cli reads raw DSD256 stereo from a pipe and drives
# d2p_push / d2p_pull one reader block group (2 x 4096 bytes) per call, exactly like process_blocks
# (rd/conversion_context.rs:295-328).  Default push threshold (8 MiB, pushes gathered in pinned memory) and threshold 0
# (every 8 KiB push shipped on its own: one copy + one launch per call).
make -C tools -s d2p_cli
python - <<'PY'
import numpy as np
np.random.default_rng(1).integers(0, 256, size=2 * 4096 * 20000, dtype=np.uint8).tofile("/tmp/seam_in.dsd")   # 58 s of DSD256 stereo
PY
echo "threshold 8 MiB (default):"; tools/d2p_cli -c 2 -f P -e L -i 4 -r 352800 -b 24 -d T - < /tmp/seam_in.dsd 2>&1 >/dev/null | tail -2
echo "threshold 0 (unbatched), first 8 s:"; head -c 22000000 /tmp/seam_in.dsd | D2P_CLI_PUSH_THRESHOLD=0 tools/d2p_cli -c 2 -f P -e L -i 4 -r 352800 -b 24 -d T - 2>&1 >/dev/null | tail -2
echo "file to file (d2p_convert_file_ex, WAV):"; tools/d2p_cli -c 2 -f P -e L -i 4 -r 352800 -b 24 -d T -o W /tmp/seam_in.dsd 2>&1 | tail -1
