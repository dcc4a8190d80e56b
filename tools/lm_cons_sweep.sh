#!/bin/bash
# every L/M topology with 448 and with 896 consumer threads: which lm_kernel build should the plan pick?
for r in 1 2 4 8; do for o in 96000 192000 384000; do
  a=$(D2P_LM_CONS=448 python tools/run_path.py --dsd-rate $r --out-rate $o --bits 32 --seconds 60 | grep -o '"kernel_ms": [0-9.]*' | cut -d' ' -f2)
  b=$(D2P_LM_CONS=896 python tools/run_path.py --dsd-rate $r --out-rate $o --bits 32 --seconds 60 | grep -o '"kernel_ms": [0-9.]*' | cut -d' ' -f2)
  echo "dsd_rate $r -> $o : 448 consumers $a ms, 896 consumers $b ms"
done; done
