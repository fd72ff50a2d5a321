#!/bin/bash
# Builds libsfpe.so of a git revision (default HEAD) into gpurun_variants/<name>.so for tools/run_ab.sh.
# usage: tools/build_head_variant.sh [name] [rev]
set -e
name=${1:-head}; rev=${2:-HEAD}
root=$(cd "$(dirname "$0")/.." && pwd)
tmp=$(mktemp -d)
mkdir -p $tmp/pkg/csrc $tmp/include $root/gpurun_variants
git -C $root show $rev:include/sfpe.h > $tmp/include/sfpe.h
for f in $(git -C $root ls-tree --name-only $rev sensor-fusion-pose-estimation-challenge_b200/csrc/); do
  git -C $root show $rev:$f > $tmp/pkg/csrc/$(basename $f)
done
make -C $tmp/pkg/csrc -j8 > /dev/null 2>&1
cp $tmp/pkg/libsfpe.so $root/gpurun_variants/$name.so
rm -rf $tmp
echo "built $name from $rev"
