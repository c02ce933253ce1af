#!/bin/bash
# Builds libytkgbdt.so (CUDA kernels + C ABI) for sm_100a, in-tree.  Objects are compiled in parallel into
# build/ and only when a source or header is newer (FORCE=1 rebuilds everything).
set -e
cd "$(dirname "$0")"
OUT=ytk-learn_b200/libytkgbdt.so
SRC=ytk-learn_b200/csrc
FILES="rows hist split partition level boost_kernels peer tree_maker grow engine boost refine hooks no_alloc formats"
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-ffp-contract=off,-Wall -I include -I $SRC ${YTK_NVCC_EXTRA}"
mkdir -p build
newest_hdr=$(ls -t $SRC/*.cuh $SRC/*.h include/*.h build.sh | head -1)
pids=()
for f in $FILES; do
  o=build/$f.o
  if [ -n "$FORCE" ] || [ ! -f $o ] || [ $SRC/$f.cu -nt $o ] || [ $newest_hdr -nt $o ]; then
    nvcc $FLAGS -c $SRC/$f.cu -o $o &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
objs=""
for f in $FILES; do objs="$objs build/$f.o"; done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT" $objs -ldl
echo "built $OUT"
