#!/usr/bin/env python3
"""First 256 bytes + length of every DSF / DSDIFF fixture of the reference -> tests/golden/container_heads.json,
with what oracle/containers.py parses out of the WHOLE file.  Run in the build container (needs /root/reference)."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
REF = os.environ.get("D2P_REFERENCE", "/root/reference")
from oracle import containers as orc  # noqa: E402

out = {}
for sub in ("test", "id3_test", "id3_test/dff"):
    for fn in sorted(os.listdir(os.path.join(REF, sub))):
        if not fn.endswith((".dsf", ".dff")):
            continue
        data = open(os.path.join(REF, sub, fn), "rb").read()
        exp = (orc.parse_dsf if fn.endswith(".dsf") else orc.parse_dff)(data)
        out[f"{sub}/{fn}"] = {"file_bytes": len(data), "head_hex": data[:256].hex(), "expect": exp}
json.dump(out, open(os.path.join(HERE, "container_heads.json"), "w"), indent=1)
print(len(out), "fixtures")
