#!/bin/bash
# tools/build_variant.sh NAME "EXTRA nvcc flags" -- a copy of the current tree's library built with other -D switches under ab/NAME
# (git-ignored, travels to the GPU box), for same-box A/B runs: python ab/NAME/tools/quick_bench.py scenes/teapot.yml 64
set -e
cd "$(dirname "$0")/.."
name=$1; shift
rm -rf ab/$name; mkdir -p ab/$name
cp -r paths_b200 include tools ab/$name/
rm -rf ab/$name/paths_b200/build ab/$name/paths_b200/lib ab/$name/paths_b200/bin ab/$name/paths_b200/__pycache__
ln -sfn ../../scenes ab/$name/scenes; ln -sfn ../../assets ab/$name/assets; ln -sfn ../../tests ab/$name/tests
make -C ab/$name/paths_b200 -s -j8 EXTRA="$*" 2>&1 | grep -v "^$" | head -20
rm -rf ab/$name/paths_b200/build # objects are not needed on the GPU box (the snapshot is capped at 512 MiB)
ls -la ab/$name/paths_b200/lib/libpaths_b200.so
