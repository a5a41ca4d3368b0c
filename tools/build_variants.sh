#!/bin/bash
# Build tuning variants of libb200recon.so into lib/variants/ (selected at run time with B2R_LIB=...).
# usage: tools/build_variants.sh name "-DB2R_IW=4 -DB2R_MINB=6" [name2 "flags2" ...]
set -e
cd "$(dirname "$0")/../3d-reconstruction-planning_b200/csrc"
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  make -j8 EXTRA="$flags" BUILD=build_$name OUT=../lib/variants/lib_$name.so > /dev/null
  echo "== $name: $flags"; grep -E "registers|spill" build_$name/tsdf_integrate.ptxas.log | sort | uniq -c | sort -rn | head -4
done
