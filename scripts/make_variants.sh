#!/bin/bash
# Build side-by-side trees for scripts/ab_bench.sh: _base = HEAD as committed, _vN = the working tree compiled with
# extra nvcc flags.  usage: scripts/make_variants.sh ["-DFLAG_A" ["-DFLAG_B" ...]]   (one _vN per argument)
set -e
cd "$(dirname "$0")/.."
rm -rf _base _v[0-9]*
mkdir _base && git archive HEAD | tar -x -C _base
(cd _base && python -m enarf_gan_b200.build > /dev/null)
i=1
for flags in "$@"; do
  d=_v$i; mkdir $d
  tar -c --exclude=.git --exclude='./_base' --exclude='./_v[0-9]*' --exclude=./gpurun_out --exclude=./oracle/_ref --exclude=./profiles \
      --exclude='*.so' --exclude='*.o' --exclude=__pycache__ . | tar -x -C $d
  (cd $d && ENARF_NVCC_EXTRA="$flags" python -m enarf_gan_b200.build > /dev/null)
  echo "$d: $flags" | tee $d/VARIANT
  i=$((i+1))
done
