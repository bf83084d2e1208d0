#!/bin/bash
# build a tagged variant of the library next to the default one: tools/build_variant.sh <tag> "<-DFLAGS>"  ->  rustraytracer_b200/librtcore_<tag>.so
# (compiled in a scratch copy of csrc/ so the default build's objects stay valid)
set -e
tag=$1; flags=$2
root=$(cd "$(dirname "$0")/.." && pwd)
tmp=$(mktemp -d)
mkdir -p $tmp/pkg/csrc $tmp/include
cp $root/rustraytracer_b200/csrc/*.cu $root/rustraytracer_b200/csrc/*.cuh $root/rustraytracer_b200/csrc/*.cpp $root/rustraytracer_b200/csrc/*.h $root/rustraytracer_b200/csrc/Makefile $tmp/pkg/csrc/
cp $root/include/*.h $tmp/include/
make -s -j6 -C $tmp/pkg/csrc EXTRA="$flags" OUT=$root/rustraytracer_b200/librtcore_$tag.so > $tmp/build.log 2>&1 || { tail -30 $tmp/build.log; exit 1; }
rm -rf $tmp
echo built librtcore_$tag.so
