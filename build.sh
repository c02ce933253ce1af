#!/bin/bash
# Builds libytkgbdt.so (CUDA kernels + C ABI) for sm_100a, in-tree.
set -e
cd "$(dirname "$0")"
OUT=ytk-learn_b200/libytkgbdt.so
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
     -Xcompiler -fPIC,-ffp-contract=off,-Wall -shared -I include -I ytk-learn_b200/csrc \
     ${YTK_NVCC_EXTRA} -o "$OUT" ytk-learn_b200/csrc/rows.cu ytk-learn_b200/csrc/hist.cu ytk-learn_b200/csrc/split.cu ytk-learn_b200/csrc/partition.cu ytk-learn_b200/csrc/boost_kernels.cu ytk-learn_b200/csrc/peer.cu ytk-learn_b200/csrc/tree_maker.cu ytk-learn_b200/csrc/grow.cu ytk-learn_b200/csrc/boost.cu ytk-learn_b200/csrc/hooks.cu ytk-learn_b200/csrc/no_alloc.cu -ldl
echo "built $OUT"
