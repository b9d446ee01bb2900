#!/bin/bash
# Builds the DISTRIBUTED flavour of the overlay from /root/reference into baseline/_ref/dist/ (git-ignored, ships with
# gpurun): the reference runtime configured with PARSEC_DIST_WITH_MPI=ON, so that remote_dep.c / remote_dep_mpi.c and
# the generated code's remote-dependency paths are compiled in, with
#   overlay/comm_engine_b200.c   in place of parsec/parsec_comm_engine.c + parsec/parsec_mpi_funnelled.c   (the seam)
#   overlay/device_cuda_b200_shim.c in place of the CUDA module, as in the single-rank overlay
#   libpb2mpi.so (overlay/pb2mpi) as "the MPI" CMake finds: datatypes + the few communicator calls, no point-to-point
#   dist/libparsec.so.4, dist/libpb2mpi.so, bin/dgemm_NN_dist (the same JDF and driver, PARSEC_HAVE_MPI defined),
#   reftests/dtd_test_ce (the reference's own comm-engine test) and reftests/dist/* (its multi-rank tests), unmodified
# Ranks are started by overlay/pb2run.  The single-rank trees (build_overlay.sh) are not touched.
set -euo pipefail
REPO="$(cd "$(dirname "$0")/.." && pwd)"
REF="${PARSEC_REFERENCE:-/root/reference}"
WORK="${WORK:-/tmp/parsec_b200_overlay}"
OUT="$REPO/baseline/_ref"
JOBS="${JOBS:-$(nproc)}"
[ -d "$REF/parsec" ] || { echo "no reference tree at $REF" >&2; exit 3; }
[ -f "$REPO/parsec-1_b200/libparsec_b200.so" ] || make -C "$REPO/parsec-1_b200" -j"$JOBS"
mkdir -p "$WORK" "$OUT/dist" "$OUT/bin" "$OUT/reftests"

# 1. the MPI stand-in
gcc -O2 -g -Wall -shared -fPIC -I"$REPO/include" -I"$REPO/overlay/pb2mpi" "$REPO/overlay/pb2mpi/pb2mpi.c" \
    -o "$OUT/dist/libpb2mpi.so" -L"$REPO/parsec-1_b200" -lparsec_b200 -lpthread '-Wl,-rpath,$ORIGIN/../../../parsec-1_b200'

# 2. the tree: reference + three substitutions
S="$WORK/src_dist"; B="$WORK/build_dist"
rm -rf "$S"; cp -r "$REF" "$S"; chmod -R u+w "$S"
rm "$S/parsec/mca/device/cuda/device_cuda_module.c" "$S/parsec/parsec_mpi_funnelled.c" "$S/parsec/parsec_comm_engine.c"
cp "$REPO/overlay/device_cuda_b200_shim.c" "$S/parsec/mca/device/cuda/device_cuda_b200_shim.c"
cp "$REPO/overlay/comm_engine_b200.c" "$S/parsec/comm_engine_b200.c"
sed -i 's|list(APPEND MCA_${COMPONENT}_SOURCES mca/device/device_gpu.c mca/device/transfer_gpu.c)|# device_gpu.c and transfer_gpu.c are provided by device_cuda_b200_shim.c|' \
    "$S/parsec/mca/device/CMakeLists.txt"
sed -i '/^  parsec_mpi_funnelled.c$/d; s/^  parsec_comm_engine.c$/  comm_engine_b200.c/' "$S/parsec/CMakeLists.txt"

# 3. configure with "an MPI" that is libpb2mpi (FindMPI's explicit-library mode), build the library and the compiler
mkdir -p "$B"
( cd "$B" && cmake "$S" -G Ninja -DCMAKE_BUILD_TYPE=Release -DPARSEC_DIST_WITH_MPI=ON -DSUPPORT_FORTRAN=OFF \
    -DBUILD_TOOLS=OFF -DBUILD_TESTING=ON -DPARSEC_PROF_TRACE=OFF -DPARSEC_GPU_WITH_CUDA=ON -DCMAKE_CUDA_ARCHITECTURES=100 \
    -DMPI_C_LIB_NAMES=pb2mpi -DMPI_pb2mpi_LIBRARY="$OUT/dist/libpb2mpi.so" -DMPI_C_HEADER_DIR="$REPO/overlay/pb2mpi" \
    -DMPI_C_ADDITIONAL_INCLUDE_DIRS="$REPO/include" -DMPI_CXX_SKIP_MPICXX=ON -DMPIEXEC_EXECUTABLE="$REPO/overlay/pb2run" \
    -DCMAKE_C_FLAGS="-I$REPO/include" \
    -DCMAKE_SHARED_LINKER_FLAGS="-Wl,-rpath,\\\$ORIGIN/../../../parsec-1_b200 -Wl,-rpath,\\\$ORIGIN" \
    -DCMAKE_C_STANDARD_LIBRARIES="-lm -L$REPO/parsec-1_b200 -lparsec_b200" > cmake.log 2>&1 \
  && ninja -j"$JOBS" parsec parsec-ptgpp > ninja.log 2>&1 ) || { tail -40 "$B/ninja.log" "$B/cmake.log" >&2; exit 4; }
cp -L "$B/parsec/libparsec.so.4" "$OUT/dist/libparsec.so.4"
grep -q "PARSEC_HAVE_MPI" "$B/parsec/include/parsec/parsec_options.h" || { echo "the dist tree was configured without MPI" >&2; exit 4; }

# 4. the reference's own comm-engine test and its multi-rank tests, unmodified sources (run by overlay/run_dist_reftests.sh)
( cd "$B" && ninja -j"$JOBS" dtd_test_ce > ninja_dtd_test_ce.log 2>&1 ) && cp "$B/tests/dsl/dtd/dtd_test_ce" "$OUT/reftests/dtd_test_ce" \
    || echo "reference test dtd_test_ce did not build (see $B/ninja_dtd_test_ce.log)" >&2
mkdir -p "$OUT/reftests/dist"
for t in dtd_test_pingpong dtd_test_task_inserting_task dtd_test_task_insertion dtd_test_war dtd_test_interleave_actions \
         dtd_test_allreduce dtd_test_new_tile jdf_forward_RW_NULL jdf_forward_READ_NULL write_check compose project merge_sort \
         reshape input_dep_reshape_single_copy remote_multiple_outs_same_pred_flow testing_redistribute \
         testing_redistribute_random testing_band testing_stencil_1D bw_test schedmicro rtt; do
    if ( cd "$B" && ninja -j"$JOBS" $t > ninja_$t.log 2>&1 ); then
        f=$(cd "$B" && find tests -type f -name $t -perm -u+x | head -1)
        [ -n "$f" ] && cp "$B/$f" "$OUT/reftests/dist/$t"
    else
        echo "reference test $t did not build (see $B/ninja_$t.log)" >&2
    fi
done

# 5. the application, against this tree's headers (PARSEC_HAVE_MPI and DISTRIBUTED come from them)
A="$WORK/app_dist"
mkdir -p "$A" && cp "$REPO/overlay/apps/dgemm_NN.jdf" "$REPO/overlay/apps/dgemm_main.c" "$A/"
( cd "$A" && "$B/parsec/interfaces/ptg/ptg-compiler/parsec-ptgpp" -E -i dgemm_NN.jdf -o dgemm_NN -f dgemm_NN > ptgpp.log 2>&1 ) \
    || { cat "$A/ptgpp.log" >&2; exit 5; }
INC="-I$REPO/overlay/pb2mpi -I$REPO/include -I$S -I$S/parsec/include -I$B -I$B/parsec/include -I/usr/local/cuda/include"
gcc -O3 -fopenmp -ffp-contract=off $INC -c "$A/dgemm_NN.c" -o "$A/dgemm_NN.o"
gcc -O3 -fopenmp -ffp-contract=off $INC -I"$A" -c "$A/dgemm_main.c" -o "$A/dgemm_main.o"
gcc -fopenmp "$A/dgemm_main.o" "$A/dgemm_NN.o" -o "$OUT/bin/dgemm_NN_dist" -L"$OUT/dist" -l:libparsec.so.4 -lpb2mpi \
    -L/usr/local/cuda/lib64 -lcublas -lcudart -lm -lpthread -ldl -Wl,-rpath-link,"$REPO/parsec-1_b200"
# the Cholesky application on several ranks
cp "$REPO/overlay/apps/dpotrf_L.jdf" "$REPO/overlay/apps/dpotrf_main.c" "$A/"
( cd "$A" && "$B/parsec/interfaces/ptg/ptg-compiler/parsec-ptgpp" -E -i dpotrf_L.jdf -o dpotrf_L -f dpotrf_L > ptgpp_potrf.log 2>&1 ) \
    || { cat "$A/ptgpp_potrf.log" >&2; exit 5; }
gcc -O3 -fopenmp $INC -c "$A/dpotrf_L.c" -o "$A/dpotrf_L.o"
gcc -O3 -fopenmp $INC -I"$A" -c "$A/dpotrf_main.c" -o "$A/dpotrf_main.o"
gcc -fopenmp "$A/dpotrf_main.o" "$A/dpotrf_L.o" -o "$OUT/bin/dpotrf_L_dist" -L"$OUT/dist" -l:libparsec.so.4 -lpb2mpi \
    -L"$REPO/parsec-1_b200" -lparsec_b200 -L/usr/local/cuda/lib64 -lcudart -lm -lpthread -ldl \
    -Wl,-rpath,'$ORIGIN/../../../parsec-1_b200'
echo "installed:"; ls -la "$OUT/dist" "$OUT/bin/dgemm_NN_dist" "$OUT/reftests/dtd_test_ce" 2>/dev/null || true
