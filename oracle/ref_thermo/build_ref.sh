#!/bin/sh
# TEST INFRASTRUCTURE (oracle/_ref): builds the REFERENCE's own device thermo kernels for sm_100a.
# The kernels, their __constant__ tables and init_const_coeff_ptr are lines 8-357 of
# /root/reference/src_gpu/dfThermo.cu (self-contained CUDA; the rest of that file needs dfMatrixDataBase.H ->
# AmgX / NCCL / MPI and is unbuildable here).  The recipe extracts them, unmodified, from where they lie into
# the git-ignored oracle/_ref/ (only the `extern __global__ correct_internal_boundary_field_scalar`
# declaration at :322-323 is dropped: it belongs to another translation unit) and compiles them together with
# our own driver (ref_thermo_driver.cu), which replays the launch sequence of dfThermo::correctThermo
# (dfThermo.cu:588-601) with the reference's block sizes and its compile flag -fmad=false
# (src_gpu/CMakeLists.txt:17).  No reference source is committed.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../_ref"
SRC=/root/reference/src_gpu/dfThermo.cu
[ -f "$SRC" ] || { echo "reference not present: keeping the prebuilt $OUT/libdfthermo_ref.so"; exit 0; }
mkdir -p "$OUT"
sed -n '8,357p' "$SRC" | sed '/^extern void __global__ correct_internal_boundary_field_scalar/,/);/d' > "$OUT/dfthermo_lifted.cuh"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -std=c++17 -Xcompiler -fPIC -shared \
     -I "$OUT" -o "$OUT/libdfthermo_ref.so" "$HERE/ref_thermo_driver.cu"
echo "built $OUT/libdfthermo_ref.so"
