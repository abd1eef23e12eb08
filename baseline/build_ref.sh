#!/usr/bin/env bash
# Build the UNMODIFIED reference (Tomaszbrauntsch/Bidirectional-BFS) out of tree into baseline/_ref.
#   baseline/_ref/src   verbatim copy of the reference tree (from /root/reference, or already present)
#   baseline/_ref/bin   v1 as shipped (no -O: Makefile typo) and -O3; v2; v3 and v4 for sm_100
# v2/v4 need MPI, which this image lacks: they are compiled against baseline/mpi_shim (mpi.h +
# mpicxx + mpirun over POSIX shared memory).  No reference source file is edited.
set -uo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF_SRC="${REF_SRC:-/root/reference}"
OUT="$HERE/_ref"
mkdir -p "$OUT/bin"
if [ -d "$REF_SRC" ] && [ ! -d "$OUT/src" ]; then
  mkdir -p "$OUT/src" && cp -r "$REF_SRC"/. "$OUT/src/"
fi
[ -d "$OUT/src" ] || { echo "no reference sources (looked in $REF_SRC and $OUT/src)"; exit 1; }
SHIM="$HERE/mpi_shim"
log="$OUT/build.log"; : > "$log"
ok=()
( cd "$OUT/src/v1" && make clean >/dev/null 2>&1; make >>"$log" 2>&1 && cp bibfs_serial "$OUT/bin/v1_as_shipped" ) && ok+=(v1_as_shipped)
g++ -std=c++11 -O3 "$OUT/src/v1/main-v1.cpp" -o "$OUT/bin/v1_O3" >>"$log" 2>&1 && ok+=(v1_O3)
( cd "$OUT/src/v2" && make clean >/dev/null 2>&1; make MPICXX="$SHIM/mpicxx" >>"$log" 2>&1 && cp second_try "$OUT/bin/v2" ) && ok+=(v2)
"$SHIM/mpicxx" -std=c++11 -O3 "$OUT/src/v2/main-v2.cpp" -o "$OUT/bin/v2_prototype" >>"$log" 2>&1 && ok+=(v2_prototype)
if command -v nvcc >/dev/null 2>&1 || [ -x /usr/local/cuda/bin/nvcc ]; then
  ( cd "$OUT/src/v3" && rm -f bibfs_cuda_only bibfs_cuda_only.o; make SMS=100 >>"$log" 2>&1 && cp bibfs_cuda_only "$OUT/bin/v3" ) && ok+=(v3)
  ( cd "$OUT/src/v4" && make clean >/dev/null 2>&1; make SMS=100 MPICXX="$SHIM/mpicxx" >>"$log" 2>&1 && cp mpi_bibfs "$OUT/bin/v4" ) && ok+=(v4)
fi
echo "built: ${ok[*]}" | tee -a "$log"
