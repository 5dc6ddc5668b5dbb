#!/usr/bin/env bash
# Builds the UNMODIFIED reference into baseline/_ref (git-ignored, travels with gpurun):
#   _ref/src/*        verbatim copy of the reference sources (taken from $KD_REFERENCE_DIR,
#                     default /root/reference; an existing copy is reused when that is absent)
#   _ref/sequential   `make sequential` with the reference Makefile
#   _ref/mpi          kdtree_mpi.cpp with the reference MPICXX flags; mpi.h = baseline/mpi_shim
#   _ref/oracle_seq, _ref/oracle_mpi   reference functions behind a command dispatcher (oracle B)
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ref="${KD_REFERENCE_DIR:-/root/reference}"
dst="$here/_ref"
mkdir -p "$dst/src"
if [ -d "$ref" ] && [ -f "$ref/kdtree_sequential.cpp" ]; then
    for f in Node.cpp Node.hpp Utility.cpp Utility.hpp kdtree_sequential.cpp kdtree_mpi.cpp Makefile; do
        cp -f "$ref/$f" "$dst/src/$f"
    done
fi
[ -f "$dst/src/kdtree_sequential.cpp" ] || { echo "no reference sources available" >&2; exit 1; }
( cd "$dst/src" && make -s sequential 2>/dev/null && cp -f sequential "$dst/sequential" )
REF_FLAGS="--std=c++17 -mavx -O3 -g -DOMPI_SKIP_MPICXX"          # reference Makefile:6 (minus warnings)
g++ $REF_FLAGS -I"$here/mpi_shim" -o "$dst/mpi" "$dst/src/kdtree_mpi.cpp" "$dst/src/Node.cpp" "$dst/src/Utility.cpp" -lpthread
SEQ_FLAGS="-O3 -std=c++17 -mavx"                                   # reference Makefile:2
g++ $SEQ_FLAGS -I"$here" -o "$dst/oracle_seq" "$here/oracle_seq.cpp" "$dst/src/Node.cpp" "$dst/src/Utility.cpp" -lm
g++ $SEQ_FLAGS -I"$here" -I"$here/mpi_shim" -o "$dst/oracle_mpi" "$here/oracle_mpi.cpp" "$dst/src/Node.cpp" "$dst/src/Utility.cpp" -lm -lpthread
echo "reference built into $dst"
