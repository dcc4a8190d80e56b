#!/bin/bash
# Packs the reference's own DSD fixtures (test/ and id3_test/, 11 files, 19 MB) WHOLE into tests/golden/ref_fixtures.tar.xz
# (6 MB: most of them share one payload), so that the GPU box -- which has no /root/reference -- can run every one of
# them end to end.  Inputs only: no reference source code, no reference output.  Run in the build container.
REF=${D2P_REFERENCE:-/root/reference}
tar -C "$REF" -cf - test id3_test | xz -9e -c > "$(dirname "$0")/ref_fixtures.tar.xz"
ls -la "$(dirname "$0")/ref_fixtures.tar.xz"
