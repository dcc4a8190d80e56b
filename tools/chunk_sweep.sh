#!/bin/bash
# e2e throughput of bench.py for several host-pipeline chunk sizes (D2P_HOST_CHUNK_MB)
for mb in 8 16 32 64 128; do
  D2P_HOST_CHUNK_MB=$mb python bench.py --no-cpu --steps 3 2>/dev/null > /tmp/cs_$mb.json
  python - "$mb" <<'PY'
import sys, json
mb = sys.argv[1]
d = json.loads(open(f"/tmp/cs_{mb}.json").read().strip().splitlines()[-1])
print(mb, "MiB e2e", d["e2e"]["value"], "x copy time", d["e2e"]["copies_alone"]["e2e_over_copy_time"])
PY
done
