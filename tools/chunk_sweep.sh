#!/bin/bash
# e2e throughput of bench.py for several host-pipeline chunk sizes (D2P_HOST_CHUNK_MB), beside the all-rank copy ceiling
for mb in 8 16 32 64 128 256; do
  D2P_HOST_CHUNK_MB=$mb python bench.py --no-cpu --no-paths --steps 2 --warmup 1 --total-files 16 2>/dev/null > /tmp/cs_$mb.json
  python - "$mb" <<'PY'
import sys, json
mb = sys.argv[1]
d = json.loads(open(f"/tmp/cs_{mb}.json").read().strip().splitlines()[-1])
print(mb, "MiB chunks: e2e %.3g DSD samples/s, %.3f x the bare copies" % (d["e2e"]["value"], d["e2e"]["platform_ceiling"]["e2e_over_ceiling"]))
PY
done
