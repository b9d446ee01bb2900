#!/bin/bash
# Builds, from /root/reference (read-only, never copied into the repository), into baseline/_ref/ (git-ignored, ships
# to the GPU box with gpurun):
#   stock/    libparsec.so of the UNMODIFIED reference with its own CUDA module          (comparator + CPU arm)
#   overlay/  libparsec.so with overlay/device_cuda_b200_shim.c in place of
#             parsec/mca/device/cuda/device_cuda_module.c, device_gpu.c, transfer_gpu.c   (the drop-in)
#   bin/      parsec-ptgpp, dgemm_NN_app = overlay/apps/dgemm_NN.jdf through parsec-ptgpp + dgemm_main.c, and
#             dpotrf_L_app = overlay/apps/dpotrf_L.jdf + dpotrf_main.c (tiled Cholesky, TRSM / SYRK / GEMM bodies)
# The work tree is a scratch copy under $WORK (default /tmp/parsec_b200_overlay); only binaries are installed.
# Recipe for a build without MPI: SURVEY.md Appendix D.
set -euo pipefail
REPO="$(cd "$(dirname "$0")/.." && pwd)"
REF="${PARSEC_REFERENCE:-/root/reference}"
WORK="${WORK:-/tmp/parsec_b200_overlay}"
OUT="$REPO/baseline/_ref"
JOBS="${JOBS:-$(nproc)}"
[ -d "$REF/parsec" ] || { echo "no reference tree at $REF" >&2; exit 3; }
[ -f "$REPO/parsec-1_b200/libparsec_b200.so" ] || make -C "$REPO/parsec-1_b200" -j"$JOBS"
mkdir -p "$WORK" "$OUT/stock" "$OUT/overlay" "$OUT/bin"
gcc -O2 -fPIC -c "$REPO/overlay/nompi_stub.c" -o "$WORK/nompi_stub.o"

prepare_tree() {   # $1 = destination
    rm -rf "$1"; cp -r "$REF" "$1"; chmod -R u+w "$1"
    sed -i '/^  parsec_mpi_funnelled.c$/d; /^  remote_dep_mpi.c$/d' "$1/parsec/CMakeLists.txt"
}
configure_and_build() {   # $1 = source tree, $2 = build dir, $3 = extra C flags, $4 = extra shared-linker flags, $5 = extra libraries
    mkdir -p "$2"
    ( cd "$2" && cmake "$1" -G Ninja -DCMAKE_BUILD_TYPE=Release -DPARSEC_DIST_WITH_MPI=OFF -DSUPPORT_FORTRAN=OFF \
        -DBUILD_TOOLS=OFF -DBUILD_TESTING=${TESTING:-OFF} -DPARSEC_PROF_TRACE=OFF -DPARSEC_GPU_WITH_CUDA=ON \
        -DCMAKE_CUDA_ARCHITECTURES=100 -DCMAKE_C_FLAGS="-include $REPO/overlay/nompi_shim.h $3" \
        -DCMAKE_SHARED_LINKER_FLAGS="$WORK/nompi_stub.o $4" -DCMAKE_C_STANDARD_LIBRARIES="-lm ${5:-}" > cmake.log 2>&1 \
      && ninja -j"$JOBS" parsec parsec-ptgpp > ninja.log 2>&1 ) || { tail -30 "$2/ninja.log" "$2/cmake.log" >&2; exit 4; }
}

MODE="${1:-all}"   # all | overlay-only (skip the stock library) | app-only (only the application, trees must exist)
if [ "$MODE" = "all" ]; then
    prepare_tree "$WORK/src_stock"
    configure_and_build "$WORK/src_stock" "$WORK/build_stock" "" "" ""
    cp -L "$WORK/build_stock/parsec/libparsec.so.4" "$OUT/stock/libparsec.so.4"
    cp "$WORK/build_stock/parsec/interfaces/ptg/ptg-compiler/parsec-ptgpp" "$OUT/bin/parsec-ptgpp"
fi

# the overlay: same tree, three files out, one file in
if [ "$MODE" != "app-only" ]; then
prepare_tree "$WORK/src_overlay"
rm "$WORK/src_overlay/parsec/mca/device/cuda/device_cuda_module.c"
cp "$REPO/overlay/device_cuda_b200_shim.c" "$WORK/src_overlay/parsec/mca/device/cuda/device_cuda_b200_shim.c"
sed -i 's|list(APPEND MCA_${COMPONENT}_SOURCES mca/device/device_gpu.c mca/device/transfer_gpu.c)|# device_gpu.c and transfer_gpu.c are provided by device_cuda_b200_shim.c|' \
    "$WORK/src_overlay/parsec/mca/device/CMakeLists.txt"
# libparsec.so finds libparsec_b200.so next to the repository it ships with: baseline/_ref/overlay -> parsec-1_b200
# the reference's own CUDA tests are built in this tree too (BUILD_TESTING); some of them name MPI without a guard:
# overlay/teststub/mpi.h gives those names single-rank meanings
TESTING=ON
configure_and_build "$WORK/src_overlay" "$WORK/build_overlay" "-I$REPO/include -include $REPO/overlay/teststub/mpi.h -I$REPO/overlay/teststub" \
    "-Wl,-rpath,\\\$ORIGIN/../../../parsec-1_b200" "-L$REPO/parsec-1_b200 -lparsec_b200"
cp -L "$WORK/build_overlay/parsec/libparsec.so.4" "$OUT/overlay/libparsec.so.4"
# the reference's own CUDA tests, unmodified sources, linked against this tree (they run over either library)
mkdir -p "$OUT/reftests"
for t in nvlink stress stage testing_get_best_device; do
    ( cd "$WORK/build_overlay" && ninja -j"$JOBS" $t > ninja_$t.log 2>&1 ) && cp "$WORK/build_overlay/tests/dsl/ptg/cuda/$t" "$OUT/reftests/$t" \
        || echo "reference test $t did not build" >&2
done
for t in dtd_test_cuda_task_insert dtd_test_new_tile dtd_test_simple_gemm; do
    ( cd "$WORK/build_overlay" && ninja -j"$JOBS" $t > ninja_$t.log 2>&1 ) && cp "$WORK/build_overlay/tests/dsl/dtd/$t" "$OUT/reftests/$t" \
        || echo "reference test $t did not build" >&2
done
gcc -O1 -g -shared -fPIC "$REPO/overlay/teststub/segv_trace.c" -o "$OUT/reftests/libsegv_trace.so" -ldl -lpthread   # backtrace on a crash
fi

# the application: JDF -> C with the reference's own compiler, then gcc against the reference's headers
B="$WORK/build_overlay"; S="$WORK/src_overlay"; A="$WORK/app"
mkdir -p "$A" && cp "$REPO/overlay/apps/dgemm_NN.jdf" "$REPO/overlay/apps/dgemm_main.c" "$A/"
( cd "$A" && "$B/parsec/interfaces/ptg/ptg-compiler/parsec-ptgpp" -E -i dgemm_NN.jdf -o dgemm_NN -f dgemm_NN > ptgpp.log 2>&1 ) \
    || { cat "$A/ptgpp.log" >&2; exit 5; }
INC="-include $REPO/overlay/nompi_shim.h -I$S -I$S/parsec/include -I$B -I$B/parsec/include -I/usr/local/cuda/include"
gcc -O3 -fopenmp -ffp-contract=off $INC -c "$A/dgemm_NN.c" -o "$A/dgemm_NN.o"
gcc -O3 -fopenmp -ffp-contract=off $INC -I"$A" -c "$A/dgemm_main.c" -o "$A/dgemm_main.o"
gcc -fopenmp "$A/dgemm_main.o" "$A/dgemm_NN.o" -o "$OUT/bin/dgemm_NN_app" -L"$OUT/overlay" -l:libparsec.so.4 \
    -L/usr/local/cuda/lib64 -lcublas -lcudart -lm -lpthread -ldl -Wl,-rpath-link,"$REPO/parsec-1_b200"
# the second application: tiled Cholesky over the same module, CUDA bodies calling the kernel library directly
cp "$REPO/overlay/apps/dpotrf_L.jdf" "$REPO/overlay/apps/dpotrf_main.c" "$A/"
( cd "$A" && "$B/parsec/interfaces/ptg/ptg-compiler/parsec-ptgpp" -E -i dpotrf_L.jdf -o dpotrf_L -f dpotrf_L > ptgpp_potrf.log 2>&1 ) \
    || { cat "$A/ptgpp_potrf.log" >&2; exit 5; }
gcc -O3 -fopenmp $INC -I"$REPO/include" -c "$A/dpotrf_L.c" -o "$A/dpotrf_L.o"
gcc -O3 -fopenmp $INC -I"$REPO/include" -I"$A" -c "$A/dpotrf_main.c" -o "$A/dpotrf_main.o"
gcc -fopenmp "$A/dpotrf_main.o" "$A/dpotrf_L.o" -o "$OUT/bin/dpotrf_L_app" -L"$OUT/overlay" -l:libparsec.so.4 \
    -L"$REPO/parsec-1_b200" -lparsec_b200 -L/usr/local/cuda/lib64 -lcudart -lm -lpthread -ldl \
    -Wl,-rpath,'$ORIGIN/../../../parsec-1_b200'
[ -f "$OUT/bin/parsec-ptgpp" ] || cp "$B/parsec/interfaces/ptg/ptg-compiler/parsec-ptgpp" "$OUT/bin/parsec-ptgpp"
nm -D --defined-only "$OUT/overlay/libparsec.so.4" | grep -E ' (parsec_cuda_kernel_scheduler|parsec_default_cuda_stage_(in|out)|parsec_cuda_module_(init|fini)|device_cuda_static_component|parsec_gpu_(pop|push|free)_workspace)$' > "$OUT/overlay/exported_symbols.txt"
echo "installed:"; ls -la "$OUT/stock" "$OUT/overlay" "$OUT/bin"
