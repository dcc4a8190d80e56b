#!/bin/bash
# tools/sass_hist.sh TAG -- static opcode histograms of the benched kernel instantiations taken from the .so in the tree,
# stamped with the commit and the library's sha256 (needs no GPU): profiles/<TAG>_lutfir_sass.txt, profiles/<TAG>_lm_sass.txt.
TAG=${1:-dev}
SO=dsd2dxd_b200/libdsd2pcm_b200.so
STAMP="head $(git rev-parse --short HEAD)$(git diff --quiet -- dsd2dxd_b200/csrc || echo +dirty), $SO sha256 $(sha256sum $SO | cut -c1-16)"
hist() {   # mangled kernel name
  echo "# cuobjdump -sass -fun $1 $SO : static opcode histogram (count opcode); $STAMP"
  cuobjdump -sass -fun "$1" $SO 2>/dev/null | grep -E '^\s+/\*[0-9a-f]{4}\*/' | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//; s/^@!?U?P[0-9T]+\s+//' | awk '{print $1}' | sed 's/;$//' | sort | uniq -c | sort -k1,1nr -k2,2
  echo
}
names=$(cuobjdump -elf $SO 2>/dev/null | grep -o '_ZN3d2p[0-9]*\(lutfir\|lm\)_kernel[A-Za-z0-9_]*' | grep -v _param_ | sort -u)
{ for k in $(echo "$names" | grep 'lutfir_kernelILi1ELi84ELi3ELi1E\|lutfir_kernelILi0ELi0ELi3ELi1E\|lutfir_kernelILi1ELi0ELi3ELi1E'); do hist $k; done
  echo "# TMA / mbarrier mnemonics per lutfir_kernel instantiation (UBLKCP = cp.async.bulk, SYNCS = mbarrier)"
  for k in $(echo "$names" | grep lutfir); do printf "%s " $k; cuobjdump -sass -fun $k $SO 2>/dev/null | grep -c 'UBLKCP\|SYNCS\|UTMACMDFLUSH'; done
} > profiles/${TAG}_lutfir_sass.txt
{ for k in $(echo "$names" | grep 'lm_kernelILi5ELi448E\|lm_kernelILi5ELi896E\|lm_kernelILi2ELi896E'); do hist $k; done
  echo "# DFMA count per kernel (must be 0: the reference's f64 order has no fused multiply-add)"
  for k in $names; do printf "%s " $k; cuobjdump -sass -fun $k $SO 2>/dev/null | grep -c 'DFMA'; done
} > profiles/${TAG}_lm_sass.txt
grep -c . profiles/${TAG}_lutfir_sass.txt profiles/${TAG}_lm_sass.txt
