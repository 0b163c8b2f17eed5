#!/bin/bash
# Builds absolutetrack_b200/libabsolutetrack_b200.so (sm_100a) in-tree.  Used by __graft_entry__.build().
set -e
cd "$(dirname "$0")"
SRC=absolutetrack_b200/csrc
OUT=absolutetrack_b200/libabsolutetrack_b200.so
mkdir -p build
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Iinclude -I$SRC"
pids=()
for f in at_error fk_skin crop_resample crop_cameras conv_f32 conv_tc model tail; do
  if [ ! -f build/$f.o ] || [ $SRC/$f.cu -nt build/$f.o ] || [ -n "$(find $SRC include -name '*.cuh' -newer build/$f.o -o -name '*.h' -newer build/$f.o)" ]; then
    nvcc $FLAGS -c $SRC/$f.cu -o build/$f.o &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
# only the C-ABI (at_*) is exported: internal C++ symbols (namespace abt) stay local to the library
printf '{ global: at_*; local: *; };\n' > build/exports.map
nvcc -shared -o $OUT build/*.o -lcudart -Xlinker --version-script=build/exports.map
echo "built $OUT"
