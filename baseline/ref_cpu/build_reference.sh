#!/bin/sh
# Build yusupov1alik/MPI-Jordan-crazy-acceleration EXACTLY as shipped (main.cpp untouched, the flags
# of its Makefile: -Ofast) against the in-tree MPI shim, because no MPI exists on this image.
#   baseline/_ref/main.cpp   <- byte-identical copy of $REF/main.cpp
#   baseline/_ref/a.out      <- `make` product name of the reference's Makefile
#   baseline/_ref/a_hilbert.out  <- CPPFLAGS=-DHILBERT variant (Makefile:4)
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(cd "$HERE/../.." && pwd)
REF=${REF:-/root/reference}
OUT="$ROOT/baseline/_ref"
mkdir -p "$OUT"
if [ -f "$REF/main.cpp" ]; then
  cp "$REF/main.cpp" "$REF/Makefile" "$OUT/"
fi
[ -f "$OUT/main.cpp" ] || { echo "reference source not found (REF=$REF)"; exit 3; }
python -m mpi_jordan_crazy_acceleration_b200._build --only shim >/dev/null
LIB="$ROOT/mpi_jordan_crazy_acceleration_b200/lib/libjbmpi.a"
cd "$OUT"
g++ $CPPFLAGS -Ofast -I"$ROOT/mpi_shim" main.cpp "$LIB" -o a.out 2> build.log || { cat build.log; exit 4; }
g++ -DHILBERT -Ofast -I"$ROOT/mpi_shim" main.cpp "$LIB" -o a_hilbert.out 2>> build.log || { cat build.log; exit 4; }
echo "built $OUT/a.out"
