#!/usr/bin/env bash
# Builds the drop-in CLI: the UNMODIFIED reference objects (main, subCommand, common, shuffle, sketch I/O, ...,
# compiled by oracle/build_ref.sh from the sources where they lie) with the five hot functions replaced by the
# shims in this directory, linked against libkssd_b200.so.  This is what INTEGRATION.md asks a maintainer to do
# in CMake; here it is done with the linker alone so that no reference source is touched or copied:
# our objects come first and --allow-multiple-definition makes the linker keep the first definition of
# sketchFastaFile / sketchFastqFile / transSketches / index_tridist / index_dist.
# Output: rabbitkssd_b200/_dropin/rabbit_kssd_b200 (git-ignored; travels to the GPU box).
set -euo pipefail
REF="${KSSD_REFERENCE:-/root/reference}"
HERE="$(cd "$(dirname "$0")" && pwd)"
PKG="$(cd "$HERE/../.." && pwd)"
ROOT="$(cd "$PKG/.." && pwd)"
OBJ="$ROOT/oracle/_ref/obj"
OUT="$PKG/_dropin"
if [ ! -d "$REF/src" ]; then
  echo "build_dropin.sh: reference tree $REF not present; keeping any prebuilt $OUT/rabbit_kssd_b200" >&2
  exit 0
fi
[ -f "$OBJ/kssd_main.o" ] || bash "$ROOT/oracle/build_ref.sh"
mkdir -p "$OUT"
CXX="${ORACLE_CXX:-/usr/bin/g++}"
for f in sketch_gpu dist_gpu setops_gpu; do
  "$CXX" -std=c++14 -O2 -g -fopenmp -include cstdint -I"$REF/src" -I"$REF/RabbitFX/io" -I"$ROOT/include" -w \
      -c "$HERE/$f.cpp" -o "$OUT/$f.o"
done
"$CXX" -fopenmp -Wl,--allow-multiple-definition "$OUT/sketch_gpu.o" "$OUT/dist_gpu.o" "$OUT/setops_gpu.o" "$OBJ"/kssd_*.o \
    -L"$OBJ" -lrabbitfx -L"$PKG" -lkssd_b200 -Wl,-rpath,'$ORIGIN/..' -Wl,-rpath,/usr/local/cuda/lib64 \
    -lz -lpthread -o "$OUT/rabbit_kssd_b200"
echo "built $OUT/rabbit_kssd_b200"
