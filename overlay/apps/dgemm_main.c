/*
 * dgemm_main.c -- driver of the dgemm_NN taskpool (dgemm_NN.jdf) on the reference runtime.
 *
 * What a user of the reference writes around a PTG taskpool, in the shape of tests/dsl/ptg/cuda/stress_main.c and
 * tests/dsl/dtd/dtd_test_simple_gemm.c: parsec_init, three parsec_matrix_block_cyclic_t in tile storage filled by
 * the LCG of dtd_test_simple_gemm.c:43-91 (seeds 1789 / 1805 / 1901, :643-648), parsec_dgemm_NN_new,
 * parsec_add2arena, add_taskpool / start / wait timed with gettimeofday (:664-666), then the figures the parity
 * tests read: per-class task counts, device->executed_tasks (device.h:136), the version of every C tile's host
 * parsec_data_copy_t, and C itself (--dump).
 *
 * The same binary runs over the stock libparsec (reference CUDA module, --cublas = its own cuBLAS body: the
 * like-for-like comparator; --mca device_cuda_enabled 0 = CPU bodies: the DAG golden and the CPU baseline) and over
 * the overlay libparsec (overlay/device_cuda_b200_shim.c + libparsec_b200.so); the library is chosen at run time
 * (LD_LIBRARY_PATH), nothing here knows which one it got.
 *
 * Built against the distributed tree (overlay/build_dist.sh: PARSEC_HAVE_MPI, the comm engine of
 * overlay/comm_engine_b200.c, ranks started by overlay/pb2run) the same file is one rank of a P x Q grid: it creates and
 * fills the tiles it owns, and every rank prints its own line; --dump / --versions then name ONE file that the ranks
 * fill together, each at the offsets of its own tiles.
 *
 * usage: dgemm_NN_app [-N n | -M m -n n -K k] [-t nb] [-c cores] [-r runs] [--cublas] [--dump file] [--release] [--decline k]
 *                     [--versions file.json] [--cpu-tile lib.so:symbol] [-P p] [-Q q] [-- parsec/MCA arguments]
 */
#include "parsec.h"
#include "parsec/data_dist/matrix/matrix.h"
#include "parsec/data_dist/matrix/two_dim_rectangle_cyclic.h"
#include "parsec/mca/device/device.h"
#include "parsec/data_internal.h"
#include "parsec/arena.h"
#include "parsec/class/info.h"
#include "parsec/execution_stream.h"

#include <dlfcn.h>
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

#if defined(PARSEC_HAVE_CUDA)
#include <cublas_v2.h>
#include "parsec/mca/device/cuda/device_cuda.h"
#endif

typedef void (*dgemm_cpu_tile_t)(int m, int n, int k, double alpha, const double *A, int lda, const double *B, int ldb,
                                 double beta, double *C, int ldc);
#include "dgemm_NN.h"

int64_t dgemm_NN_count[4];

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

static int grid_rank = 0, grid_P = 1, grid_Q = 1;

static int is_mine(parsec_matrix_block_cyclic_t *dc, int tm, int tn)
{
    return (int)dc->super.super.rank_of(&dc->super.super, tm, tn) == grid_rank;
}

static parsec_matrix_block_cyclic_t *new_matrix(const char *name, int m, int n, int mb, int nb, unsigned seed)
{
    parsec_matrix_block_cyclic_t *dc = (parsec_matrix_block_cyclic_t*)calloc(1, sizeof(*dc));
    parsec_matrix_block_cyclic_init(dc, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_TILE, grid_rank, mb, nb, m, n, 0, 0, m, n,
                                    grid_P, grid_Q, 1, 1, 0, 0);
    size_t bytes = (size_t)dc->super.nb_local_tiles * (size_t)dc->super.bsiz * sizeof(double);
    dc->mat = parsec_data_allocate(bytes);
    if( NULL == dc->mat ) { fprintf(stderr, "cannot allocate %zu bytes for %s\n", bytes, name); exit(2); }
    parsec_data_collection_set_key((parsec_data_collection_t*)dc, (char*)name);
    #pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for( int tn = 0; tn < dc->super.nt; tn++ )
        for( int tm = 0; tm < dc->super.mt; tm++ ) {
            if( !is_mine(dc, tm, tn) ) continue;
            parsec_data_t *d = dc->super.super.data_of(&dc->super.super, tm, tn);
            fill_tile((double*)parsec_data_get_ptr(d, 0), tm, tn, mb, nb, m, seed);
        }
    return dc;
}

#if defined(PARSEC_HAVE_CUDA)
/* per-stream cuBLAS handle, the pattern of tests/dsl/ptg/cuda/nvlink_wrapper.c:23-58 */
static void *create_cublas_handle(void *obj, void *p)
{
    parsec_cuda_exec_stream_t *stream = (parsec_cuda_exec_stream_t*)obj;
    cublasHandle_t handle;
    (void)p;
    if( CUBLAS_STATUS_SUCCESS != cublasCreate(&handle) ) return NULL;
    cublasSetStream(handle, stream->cuda_stream);
    return (void*)handle;
}
static void destroy_cublas_handle(void *h, void *p)
{
    (void)p;
    if( NULL != h ) cublasDestroy((cublasHandle_t)h);
}
#endif

static double now(void)
{
    struct timeval tv;
    gettimeofday(&tv, NULL);
    return (double)tv.tv_sec + 1e-6 * (double)tv.tv_usec;
}

int main(int argc, char **argv)
{
    int M = 2048, N = 2048, K = 2048, NB = 256, cores = -1, runs = 1, use_cublas = 0, release = 0, decline_k = -1;
    const char *dump = NULL, *cpu_tile_spec = NULL, *versions_path = NULL;
    dgemm_cpu_tile_t cpu_tile = NULL;
    int pargc = 1, world = 1, want_P = 0, want_Q = 0;
    char **pargv = (char**)calloc((size_t)argc + 1, sizeof(char*));
    pargv[0] = argv[0];
    for( int i = 1; i < argc; i++ ) {
        if( !strcmp(argv[i], "--") ) { for( i++; i < argc; i++ ) pargv[pargc++] = argv[i]; break; }
        else if( !strcmp(argv[i], "-N") && i + 1 < argc ) M = N = K = atoi(argv[++i]);
        else if( !strcmp(argv[i], "-M") && i + 1 < argc ) M = atoi(argv[++i]);
        else if( !strcmp(argv[i], "-n") && i + 1 < argc ) N = atoi(argv[++i]);
        else if( !strcmp(argv[i], "-K") && i + 1 < argc ) K = atoi(argv[++i]);
        else if( !strcmp(argv[i], "-t") && i + 1 < argc ) NB = atoi(argv[++i]);
        else if( !strcmp(argv[i], "-c") && i + 1 < argc ) cores = atoi(argv[++i]);
        else if( !strcmp(argv[i], "-r") && i + 1 < argc ) runs = atoi(argv[++i]);
        else if( !strcmp(argv[i], "--cublas") ) use_cublas = 1;
        else if( !strcmp(argv[i], "--release") ) release = 1;
        else if( !strcmp(argv[i], "--decline") && i + 1 < argc ) decline_k = atoi(argv[++i]);
        else if( !strcmp(argv[i], "--dump") && i + 1 < argc ) dump = argv[++i];
        else if( !strcmp(argv[i], "--cpu-tile") && i + 1 < argc ) cpu_tile_spec = argv[++i];
        else if( !strcmp(argv[i], "--versions") && i + 1 < argc ) versions_path = argv[++i];
        else if( !strcmp(argv[i], "-P") && i + 1 < argc ) want_P = atoi(argv[++i]);
        else if( !strcmp(argv[i], "-Q") && i + 1 < argc ) want_Q = atoi(argv[++i]);
        else { fprintf(stderr, "unknown argument %s\n", argv[i]); return 2; }
    }
    if( NULL != cpu_tile_spec ) {   /* lib.so:symbol -- a faster tile routine for the CPU incarnation */
        char *spec = strdup(cpu_tile_spec), *colon = strrchr(spec, ':');
        if( NULL == colon ) { fprintf(stderr, "--cpu-tile wants lib.so:symbol\n"); return 2; }
        *colon = '\0';
        void *h = dlopen(spec, RTLD_NOW);
        cpu_tile = h ? (dgemm_cpu_tile_t)dlsym(h, colon + 1) : NULL;
        if( NULL == cpu_tile ) { fprintf(stderr, "cannot resolve %s: %s\n", cpu_tile_spec, dlerror()); return 2; }
        free(spec);
    }

#if defined(PARSEC_HAVE_MPI)
    {
        int provided;
        if( MPI_SUCCESS != MPI_Init_thread(&argc, &argv, MPI_THREAD_SERIALIZED, &provided) ) return 1;
        MPI_Comm_size(MPI_COMM_WORLD, &world);
        MPI_Comm_rank(MPI_COMM_WORLD, &grid_rank);
    }
#endif
    grid_P = want_P > 0 ? want_P : (want_Q > 0 ? world / want_Q : 1);
    grid_Q = want_Q > 0 ? want_Q : world / grid_P;
    if( grid_P * grid_Q != world ) { fprintf(stderr, "a %d x %d grid does not hold %d ranks\n", grid_P, grid_Q, world); return 2; }

    parsec_context_t *parsec = parsec_init(cores, &pargc, &pargv);
    if( NULL == parsec ) { fprintf(stderr, "parsec_init failed\n"); return 1; }

    int nb_gpus = 0;
    for( int d = 0; d < (int)parsec_nb_devices; d++ ) {
        parsec_device_module_t *dev = parsec_mca_device_get(d);
        if( NULL != dev && PARSEC_DEV_CUDA == dev->type ) nb_gpus++;
    }

    parsec_matrix_block_cyclic_t *dcA = new_matrix("A", M, K, NB, NB, 1789);
    parsec_matrix_block_cyclic_t *dcB = new_matrix("B", K, N, NB, NB, 1805);
    parsec_matrix_block_cyclic_t *dcC = new_matrix("C", M, N, NB, NB, 1901);

    /* page-lock the matrices before the clock starts; the taskpool's startup would do it otherwise
     * (twoDBC_memory_register, two_dim_rectangle_cyclic.c:39-49) */
    for( int d = 0; d < (int)parsec_nb_devices; d++ ) {
        parsec_device_module_t *dev = parsec_mca_device_get(d);
        if( NULL == dev || PARSEC_DEV_CUDA != dev->type ) continue;
        dcA->super.super.register_memory(&dcA->super.super, dev);
        dcB->super.super.register_memory(&dcB->super.super, dev);
        dcC->super.super.register_memory(&dcC->super.super, dev);
    }

    parsec_info_id_t CuHI = -1;
#if defined(PARSEC_HAVE_CUDA)
    CuHI = parsec_info_register(&parsec_per_stream_infos, "CUBLAS::HANDLE", destroy_cublas_handle, NULL,
                                create_cublas_handle, NULL, NULL);
#endif

    double best = 1e30, first = 0;
    uint64_t executed_on_gpu = 0;
    for( int r = 0; r < runs; r++ ) {
        if( release ) parsec_devices_release_memory();   /* start every run with an empty device cache */
        parsec_dgemm_NN_taskpool_t *tp = parsec_dgemm_NN_new(&dcA->super, &dcB->super, &dcC->super, CuHI, use_cublas,
                                                             cpu_tile, decline_k);
        parsec_add2arena(&tp->arenas_datatypes[PARSEC_dgemm_NN_DEFAULT_ADT_IDX], parsec_datatype_double_t,
                         PARSEC_MATRIX_FULL, 1, NB, NB, NB, PARSEC_ARENA_ALIGNMENT_SSE, -1);
        double t0 = now();
        parsec_context_add_taskpool(parsec, &tp->super);
        parsec_context_start(parsec);
        parsec_context_wait(parsec);
        double t = now() - t0;
        if( 0 == r ) first = t;
        if( t < best ) best = t;
        parsec_del2arena(&tp->arenas_datatypes[PARSEC_dgemm_NN_DEFAULT_ADT_IDX]);
        parsec_taskpool_free(&tp->super);
    }
    uint64_t d2d_bytes = 0, h2d_bytes = 0, d2h_bytes = 0;
    for( int d = 0; d < (int)parsec_nb_devices; d++ ) {
        parsec_device_module_t *dev = parsec_mca_device_get(d);
        if( NULL == dev || PARSEC_DEV_CUDA != dev->type ) continue;
        executed_on_gpu += dev->executed_tasks;
        d2h_bytes += dev->data_out_to_host;
        for( int j = 0; j < (int)dev->data_in_array_size; j++ )   /* bytes staged in, by source device (device.h:130) */
            if( j >= 2 ) d2d_bytes += dev->data_in_from_device[j]; else h2d_bytes += dev->data_in_from_device[j];
    }

    /* versions of the host copies of C (KT per run) and of A, B (0) */
    uint32_t vmin = ~0u, vmax = 0, vab = 0;
    for( int tn = 0; tn < dcC->super.nt; tn++ )
        for( int tm = 0; tm < dcC->super.mt; tm++ ) {
            if( !is_mine(dcC, tm, tn) ) continue;
            parsec_data_t *d = dcC->super.super.data_of(&dcC->super.super, tm, tn);
            uint32_t v = parsec_data_get_copy(d, 0)->version;
            if( v < vmin ) vmin = v;
            if( v > vmax ) vmax = v;
        }
    for( int tn = 0; tn < dcA->super.nt; tn++ )
        for( int tm = 0; tm < dcA->super.mt; tm++ ) {
            if( !is_mine(dcA, tm, tn) ) continue;
            uint32_t v = parsec_data_get_copy(dcA->super.super.data_of(&dcA->super.super, tm, tn), 0)->version;
            if( v > vab ) vab = v;
        }
    for( int tn = 0; tn < dcB->super.nt; tn++ )
        for( int tm = 0; tm < dcB->super.mt; tm++ ) {
            if( !is_mine(dcB, tm, tn) ) continue;
            uint32_t v = parsec_data_get_copy(dcB->super.super.data_of(&dcB->super.super, tm, tn), 0)->version;
            if( v > vab ) vab = v;
        }

    if( NULL != versions_path ) {   /* version of the host copy of every tile: A, B, C, tiles column-major */
        char path[4096];   /* several ranks: one file per rank, -1 for the tiles of the others */
        if( world > 1 ) snprintf(path, sizeof(path), "%s.%d", versions_path, grid_rank);
        else snprintf(path, sizeof(path), "%s", versions_path);
        FILE *f = fopen(path, "w");
        if( NULL == f ) { perror(path); return 1; }
        parsec_matrix_block_cyclic_t *dcs[3] = { dcA, dcB, dcC };
        fprintf(f, "{");
        for( int which = 0; which < 3; which++ ) {
            fprintf(f, "%s\"%c\": [", which ? ", " : "", "ABC"[which]);
            for( int tn = 0; tn < dcs[which]->super.nt; tn++ )
                for( int tm = 0; tm < dcs[which]->super.mt; tm++ ) {
                    if( !is_mine(dcs[which], tm, tn) ) { fprintf(f, "%s-1", (tn || tm) ? ", " : ""); continue; }
                    parsec_data_t *d = dcs[which]->super.super.data_of(&dcs[which]->super.super, tm, tn);
                    fprintf(f, "%s%u", (tn || tm) ? ", " : "", parsec_data_get_copy(d, 0)->version);
                }
            fprintf(f, "]");
        }
        fprintf(f, "}\n");
        fclose(f);
    }

    if( NULL != dump ) {   /* C in tile storage, tiles in data_of order (column-major over tiles); every rank writes its own */
        int fd = open(dump, O_CREAT | O_WRONLY, 0644);
        if( fd < 0 ) { perror(dump); return 1; }
        const size_t tile_bytes = sizeof(double) * (size_t)NB * NB;
        for( int tn = 0; tn < dcC->super.nt; tn++ )
            for( int tm = 0; tm < dcC->super.mt; tm++ ) {
                if( !is_mine(dcC, tm, tn) ) continue;
                parsec_data_t *d = dcC->super.super.data_of(&dcC->super.super, tm, tn);
                const char *src = (const char*)parsec_data_get_ptr(d, 0);
                off_t at = (off_t)((size_t)tn * dcC->super.mt + tm) * (off_t)tile_bytes;
                for( size_t done = 0; done < tile_bytes; ) {
                    ssize_t w = pwrite(fd, src + done, tile_bytes - done, at + (off_t)done);
                    if( w <= 0 ) { perror("pwrite"); return 1; }
                    done += (size_t)w;
                }
            }
        close(fd);
    }

    double flops = 2.0 * (double)M * (double)N * (double)K;
    printf("{\"rank\": %d, \"world\": %d, \"P\": %d, \"Q\": %d, "
           "\"M\": %d, \"N\": %d, \"K\": %d, \"NB\": %d, \"runs\": %d, \"cores\": %d, \"gpus\": %d, \"cublas\": %d, "
           "\"seconds_best\": %.6f, \"seconds_first\": %.6f, \"gflops_best\": %.3f, \"gflops_first\": %.3f, "
           "\"read_a\": %lld, \"read_b\": %lld, \"gemm_cpu\": %lld, \"gemm_gpu\": %lld, \"device_executed_tasks\": %llu, "
           "\"c_version_min\": %u, \"c_version_max\": %u, \"ab_version_max\": %u, "
           "\"h2d_bytes\": %llu, \"d2d_bytes\": %llu, \"d2h_bytes\": %llu}\n",
           grid_rank, world, grid_P, grid_Q,
           M, N, K, NB, runs, parsec->virtual_processes[0]->nb_cores, nb_gpus, use_cublas, best, first,
           flops / best * 1e-9, flops / first * 1e-9, (long long)dgemm_NN_count[0], (long long)dgemm_NN_count[1],
           (long long)dgemm_NN_count[2], (long long)dgemm_NN_count[3], (unsigned long long)executed_on_gpu,
           vmin, vmax, vab, (unsigned long long)h2d_bytes, (unsigned long long)d2d_bytes, (unsigned long long)d2h_bytes);
    fflush(stdout);

    for( int d = 0; d < (int)parsec_nb_devices; d++ ) {
        parsec_device_module_t *dev = parsec_mca_device_get(d);
        if( NULL == dev || PARSEC_DEV_CUDA != dev->type ) continue;
        dcA->super.super.unregister_memory(&dcA->super.super, dev);
        dcB->super.super.unregister_memory(&dcB->super.super, dev);
        dcC->super.super.unregister_memory(&dcC->super.super, dev);
    }
#if defined(PARSEC_HAVE_CUDA)
    parsec_info_unregister(&parsec_per_stream_infos, CuHI, NULL);
#endif
    parsec_data_free(dcA->mat); parsec_data_free(dcB->mat); parsec_data_free(dcC->mat);
    parsec_tiled_matrix_destroy(&dcA->super); parsec_tiled_matrix_destroy(&dcB->super); parsec_tiled_matrix_destroy(&dcC->super);
    free(dcA); free(dcB); free(dcC);
    parsec_fini(&parsec);
#if defined(PARSEC_HAVE_MPI)
    MPI_Finalize();
#endif
    return 0;
}
