#!/bin/bash
# Build an A/B variant of libcuorb with extra nvcc flags into csrc/_obj/variants/<name>/libcuorb.so (git-ignored;
# travels to the GPU box).  usage: scripts/build_variant.sh <name> [-DCUORB_...=... ...]
set -e
name=$1; shift
cd "$(dirname "$0")/.."
out=orbslam3_cpp_fork_b200/csrc/_obj/variants/$name
mkdir -p $out
for f in pyramid blur fast octree describe stereo project preproc frameglue bfmatch bow search synth cuorb_api; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-ffp-contract=off "$@" \
    -c orbslam3_cpp_fork_b200/csrc/$f.cu -o $out/$f.o &
done
wait
/usr/local/cuda/bin/nvcc -shared -o $out/libcuorb.so $out/*.o
echo $out/libcuorb.so
