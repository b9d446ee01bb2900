/*
 * dpotrf_main.c -- driver of the dpotrf_L taskpool (dpotrf_L.jdf) on the reference runtime; one rank, or, built against
 * the distributed tree (overlay/build_dist.sh), one rank of a P x Q grid started by overlay/pb2run.
 *
 * Same shape as dgemm_main.c: parsec_init, one parsec_matrix_block_cyclic_t in tile storage, parsec_dpotrf_L_new,
 * add_taskpool / start / wait, then what the parity test reads: per-class task counts, device->executed_tasks, the
 * bytes the device module moved, and the factor itself (--dump).  The matrix: the lower triangle from the LCG of
 * dtd_test_simple_gemm.c:43-91 (seed 3872), N added to the diagonal (strictly diagonally dominant, so positive
 * definite); the upper triangle is never referenced.
 *
 * usage: dpotrf_L_app [-N n] [-t nb] [-c cores] [--dump file] [--cpu-trsm] [--cpu-potrf] [-P p] [-Q q] [-- parsec/MCA arguments]
 * Several ranks: every rank prints its own line and writes its own tiles into the one --dump file.
 */
#include "parsec.h"
#include "parsec/data_dist/matrix/matrix.h"
#include "parsec/data_dist/matrix/two_dim_rectangle_cyclic.h"
#include "parsec/mca/device/device.h"
#include "parsec/data_internal.h"
#include "parsec/arena.h"
#include "parsec/execution_stream.h"

#include <fcntl.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>
#include <unistd.h>

#if defined(PARSEC_HAVE_MPI)
#include <mpi.h>
#endif

#include "dpotrf_L.h"

int64_t dpotrf_L_count[9];

/* ---- inputs: the generator of dtd_test_simple_gemm.c:43-91 (64-bit LCG with a jump-ahead per tile column; the value
 * is formed in single precision by a multiply and a subtract, each rounded) ---- */
#define RND64_A 6364136223846793005ULL
#define RND64_C 1ULL
static uint64_t rnd64_jump(uint64_t n, uint64_t seed)
{
    uint64_t a_k = RND64_A, c_k = RND64_C, ran = seed;
    for( ; n; n >>= 1 ) {
        if( n & 1 ) ran = a_k * ran + c_k;
        c_k *= (a_k + 1);
        a_k *= a_k;
    }
    return ran;
}

static void fill_tile(double *tile, int m, int n, int mb, int nb, int M, unsigned seed)
{
    uint64_t jump = (uint64_t)m + (uint64_t)n * (uint64_t)M;
    for( int j = 0; j < nb; j++, jump += (uint64_t)M ) {
        uint64_t ran = rnd64_jump(jump, seed);
        for( int i = 0; i < mb; i++ ) {
            volatile float scaled = (float)ran * 5.4210108624275222e-20f;
            float v = 0.5f - scaled;
            tile[(size_t)j * mb + i] = (double)v;
            ran = RND64_A * ran + RND64_C;
        }
    }
}

static double now(void)
{
    struct timeval tv;
    gettimeofday(&tv, NULL);
    return (double)tv.tv_sec + 1e-6 * (double)tv.tv_usec;
}

int main(int argc, char **argv)
{
    int N = 2048, NB = 256, cores = -1, gpu_trsm = 1, gpu_potrf = 1, rank = 0, world = 1, P = 0, Q = 0;
    const char *dump = NULL;
    int pargc = 1;
    char **pargv = (char**)calloc((size_t)argc + 1, sizeof(char*));
    pargv[0] = argv[0];
    for( int i = 1; i < argc; i++ ) {
        if( !strcmp(argv[i], "--") ) { for( i++; i < argc; i++ ) pargv[pargc++] = argv[i]; break; }
        else if( !strcmp(argv[i], "-N") && i + 1 < argc ) N = atoi(argv[++i]);
        else if( !strcmp(argv[i], "-t") && i + 1 < argc ) NB = atoi(argv[++i]);
        else if( !strcmp(argv[i], "-c") && i + 1 < argc ) cores = atoi(argv[++i]);
        else if( !strcmp(argv[i], "--dump") && i + 1 < argc ) dump = argv[++i];
        else if( !strcmp(argv[i], "--cpu-trsm") ) gpu_trsm = 0;
        else if( !strcmp(argv[i], "--cpu-potrf") ) gpu_potrf = 0;
        else if( !strcmp(argv[i], "-P") && i + 1 < argc ) P = atoi(argv[++i]);
        else if( !strcmp(argv[i], "-Q") && i + 1 < argc ) Q = atoi(argv[++i]);
        else { fprintf(stderr, "unknown argument %s\n", argv[i]); return 2; }
    }
#if defined(PARSEC_HAVE_MPI)
    {
        int provided;
        if( MPI_SUCCESS != MPI_Init_thread(&argc, &argv, MPI_THREAD_SERIALIZED, &provided) ) return 1;
        MPI_Comm_size(MPI_COMM_WORLD, &world);
        MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    }
#endif
    if( P <= 0 ) P = Q > 0 ? world / Q : 1;
    if( Q <= 0 ) Q = world / P;
    if( P * Q != world ) { fprintf(stderr, "a %d x %d grid does not hold %d ranks\n", P, Q, world); return 2; }
    parsec_context_t *parsec = parsec_init(cores, &pargc, &pargv);
    if( NULL == parsec ) { fprintf(stderr, "parsec_init failed\n"); return 1; }
    int nb_gpus = 0;
    for( int d = 0; d < (int)parsec_nb_devices; d++ ) {
        parsec_device_module_t *dev = parsec_mca_device_get(d);
        if( NULL != dev && PARSEC_DEV_CUDA == dev->type ) nb_gpus++;
    }

    parsec_matrix_block_cyclic_t *dcA = (parsec_matrix_block_cyclic_t*)calloc(1, sizeof(*dcA));
    parsec_matrix_block_cyclic_init(dcA, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_TILE, rank, NB, NB, N, N, 0, 0, N, N, P, Q, 1, 1, 0, 0);
    size_t bytes = (size_t)dcA->super.nb_local_tiles * (size_t)dcA->super.bsiz * sizeof(double);
    dcA->mat = parsec_data_allocate(bytes);
    if( NULL == dcA->mat ) { fprintf(stderr, "cannot allocate %zu bytes\n", bytes); return 2; }
    parsec_data_collection_set_key((parsec_data_collection_t*)dcA, "A");
    #pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for( int tn = 0; tn < dcA->super.nt; tn++ )
        for( int tm = 0; tm < dcA->super.mt; tm++ ) {
            if( (int)dcA->super.super.rank_of(&dcA->super.super, tm, tn) != rank ) continue;
            double *tile = (double*)parsec_data_get_ptr(dcA->super.super.data_of(&dcA->super.super, tm, tn), 0);
            fill_tile(tile, tm, tn, NB, NB, N, 3872);
            if( tm == tn ) for( int i = 0; i < NB; i++ ) tile[i + (size_t)i * NB] += (double)N;
        }
    for( int d = 0; d < (int)parsec_nb_devices; d++ ) {
        parsec_device_module_t *dev = parsec_mca_device_get(d);
        if( NULL != dev && PARSEC_DEV_CUDA == dev->type ) dcA->super.super.register_memory(&dcA->super.super, dev);
    }

    parsec_dpotrf_L_taskpool_t *tp = parsec_dpotrf_L_new(&dcA->super, gpu_trsm, gpu_potrf);
    parsec_add2arena(&tp->arenas_datatypes[PARSEC_dpotrf_L_DEFAULT_ADT_IDX], parsec_datatype_double_t,
                     PARSEC_MATRIX_FULL, 1, NB, NB, NB, PARSEC_ARENA_ALIGNMENT_SSE, -1);
    double t0 = now();
    parsec_context_add_taskpool(parsec, &tp->super);
    parsec_context_start(parsec);
    parsec_context_wait(parsec);
    double t = now() - t0;
    parsec_del2arena(&tp->arenas_datatypes[PARSEC_dpotrf_L_DEFAULT_ADT_IDX]);
    parsec_taskpool_free(&tp->super);

    uint64_t executed_on_gpu = 0, d2d_bytes = 0, h2d_bytes = 0, d2h_bytes = 0;
    for( int d = 0; d < (int)parsec_nb_devices; d++ ) {
        parsec_device_module_t *dev = parsec_mca_device_get(d);
        if( NULL == dev || PARSEC_DEV_CUDA != dev->type ) continue;
        executed_on_gpu += dev->executed_tasks;
        d2h_bytes += dev->data_out_to_host;
        for( int j = 0; j < (int)dev->data_in_array_size; j++ )
            if( j >= 2 ) d2d_bytes += dev->data_in_from_device[j]; else h2d_bytes += dev->data_in_from_device[j];
    }
    if( NULL != dump ) {   /* tile storage, tiles column-major over tiles; the lower triangle holds L */
        int fd = open(dump, O_CREAT | O_WRONLY, 0644);
        if( fd < 0 ) { perror(dump); return 1; }
        const size_t tile_bytes = sizeof(double) * (size_t)NB * NB;
        for( int tn = 0; tn < dcA->super.nt; tn++ )
            for( int tm = 0; tm < dcA->super.mt; tm++ ) {
                if( (int)dcA->super.super.rank_of(&dcA->super.super, tm, tn) != rank ) continue;
                const char *src = (const char*)parsec_data_get_ptr(dcA->super.super.data_of(&dcA->super.super, tm, tn), 0);
                off_t at = (off_t)((size_t)tn * dcA->super.mt + tm) * (off_t)tile_bytes;
                for( size_t done = 0; done < tile_bytes; ) {
                    ssize_t w = pwrite(fd, src + done, tile_bytes - done, at + (off_t)done);
                    if( w <= 0 ) { perror("pwrite"); return 1; }
                    done += (size_t)w;
                }
            }
        close(fd);
    }
    printf("{\"rank\": %d, \"world\": %d, \"N\": %d, \"NB\": %d, \"cores\": %d, \"gpus\": %d, \"seconds\": %.6f, \"gflops\": %.3f, "
           "\"potrf\": %lld, \"potrf_gpu\": %lld, \"trsm_cpu\": %lld, \"trsm_gpu\": %lld, \"syrk_cpu\": %lld, \"syrk_gpu\": %lld, "
           "\"gemm_cpu\": %lld, \"gemm_gpu\": %lld, \"potrf_failures\": %lld, \"device_executed_tasks\": %llu, "
           "\"h2d_bytes\": %llu, \"d2d_bytes\": %llu, \"d2h_bytes\": %llu}\n",
           rank, world, N, NB, parsec->virtual_processes[0]->nb_cores, nb_gpus, t, (double)N * N * N / 3.0 / t * 1e-9,
           (long long)(dpotrf_L_count[0] + dpotrf_L_count[8]), (long long)dpotrf_L_count[8], (long long)dpotrf_L_count[1], (long long)dpotrf_L_count[2], (long long)dpotrf_L_count[3],
           (long long)dpotrf_L_count[4], (long long)dpotrf_L_count[5], (long long)dpotrf_L_count[6], (long long)dpotrf_L_count[7],
           (unsigned long long)executed_on_gpu, (unsigned long long)h2d_bytes, (unsigned long long)d2d_bytes,
           (unsigned long long)d2h_bytes);
    fflush(stdout);
    for( int d = 0; d < (int)parsec_nb_devices; d++ ) {
        parsec_device_module_t *dev = parsec_mca_device_get(d);
        if( NULL != dev && PARSEC_DEV_CUDA == dev->type ) dcA->super.super.unregister_memory(&dcA->super.super, dev);
    }
    parsec_data_free(dcA->mat);
    parsec_tiled_matrix_destroy(&dcA->super);
    free(dcA);
    parsec_fini(&parsec);
#if defined(PARSEC_HAVE_MPI)
    MPI_Finalize();
#endif
    return 0;
}
