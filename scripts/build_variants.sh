#!/bin/bash
# Builds A/B variants of libraht_b200.so into scratch/variants/<name>/libraht_b200.so (git-ignored, travels with gpurun).
# usage: scripts/build_variants.sh name "-DRAHT_TR_PREFETCH=1 ..." [name2 "flags2" ...]
set -e
cd "$(dirname "$0")/../raht-3dgs-codec_b200"
ARCH="-gencode arch=compute_100a,code=sm_100a"
FLAGS="$ARCH -O3 -lineinfo -std=c++17 --fmad=false -Xcompiler -fPIC,-fvisibility=hidden"
while [ $# -ge 2 ]; do
  name="$1"; extra="$2"; shift 2
  out="../scratch/variants/$name"; mkdir -p "$out/obj"
  for f in csrc/*.cu; do
    b=$(basename "$f" .cu)
    # only the translation units that see the macros need a variant build; the rest are linked from build/
    case "$b" in transform|tree|scene|shard) (nvcc $FLAGS $extra -c "$f" -o "$out/obj/$b.o" 2> "$out/obj/$b.log" || (cat "$out/obj/$b.log"; exit 1)) & ;; *) cp "build/$b.o" "$out/obj/$b.o" ;; esac
  done
  wait
  nvcc $ARCH -shared -o "$out/libraht_b200.so" "$out"/obj/*.o
  echo "built $out/libraht_b200.so ($extra)"
done
