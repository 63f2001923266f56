#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Builds the UNMODIFIED RabbitKSSD reference binary from the sources
# where they lie (default /root/reference) straight with g++ -- the reference's own cmake build is
# not used (cmake 4.x rejects its cmake_minimum_required and its FindOpenMP probe).
# Outputs go only into oracle/_ref/ (git-ignored, but shipped to the GPU box by gpurun).
# Nothing is copied from the reference into this repository.
set -euo pipefail
REF="${KSSD_REFERENCE:-/root/reference}"
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then
  echo "build_ref.sh: reference tree $REF not present; keeping any prebuilt $OUT/rabbit_kssd" >&2
  exit 0
fi
mkdir -p "$OUT/obj"
if [ -x "$OUT/rabbit_kssd" ] && [ "$OUT/rabbit_kssd" -nt "$REF/src/sketch.cpp" ] && [ "$OUT/rabbit_kssd" -nt "$REF/src/dist.cpp" ]; then
  exit 0
fi
# flags mirror RabbitFX/CMakeLists.txt:11,25 and CMakeLists.txt:21-37; -include cstdint is the
# GCC-13 fix for RabbitFX/io/Reference.h:16.  -march=x86-64-v3 instead of -march=native so that a
# binary built here also runs on the GPU box's host CPU (AVX2 is the widest ISA the reference uses).
pids=()
for f in "$REF"/RabbitFX/io/*.cpp; do
  g++ -std=c++11 -O3 -g -ffast-math -fopenmp -DVERB -include cstdint -I"$REF/RabbitFX/io" \
      -c "$f" -o "$OUT/obj/fx_$(basename "$f" .cpp).o" 2>/dev/null &
  pids+=($!)
done
for f in "$REF"/src/*.cpp; do
  g++ -std=c++14 -g -O3 -march=x86-64-v3 -fopenmp -D DEBUG -D Timer -D Timer_inner -include cstdint \
      -I"$REF/src" -I"$REF/RabbitFX/io" -w -c "$f" -o "$OUT/obj/kssd_$(basename "$f" .cpp).o" &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
# static archive, as the reference's add_library(rabbitfx) does: Formater.o and verbos_Formater.o define
# the same symbols and the archive lets the linker take the first (FastxStream, Formater) only.
rm -f "$OUT/obj/librabbitfx.a"
ar rcs "$OUT/obj/librabbitfx.a" "$OUT"/obj/fx_*.o
g++ -fopenmp "$OUT"/obj/kssd_*.o -L"$OUT/obj" -lrabbitfx -lz -lpthread -o "$OUT/rabbit_kssd"
echo "built $OUT/rabbit_kssd"
