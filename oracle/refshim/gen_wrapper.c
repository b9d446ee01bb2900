/* Compiles the reference's tests/dsl/dtd/dtd_test_simple_gemm.c where it lies (its main renamed) so that its own
 * initialize_tile() -- the synthetic-input generator of the path -- can be executed.  initialize_tile fetches its
 * arguments with parsec_dtd_unpack_args(); the definition below hands it the arguments set by ref_init_tile(). */
#include <stdarg.h>
#include <stddef.h>
/* the test only compiles with CUDA and MPI present: give it the macro / symbols it names (never executed here) */
#define PARSEC_HAVE_CUDA 1
#define MPI_COMM_WORLD 0
int MPI_Abort(int comm, int code);
#define main ref_simple_gemm_main
#include "tests/dsl/dtd/dtd_test_simple_gemm.c"
#undef main

static struct { double* data; int m, n, mb, nb, M, ld; unsigned seed; } g_args;

void parsec_dtd_unpack_args(parsec_task_t* this_task, ...) {
    (void)this_task;
    va_list ap;
    va_start(ap, this_task);
    *va_arg(ap, double**) = g_args.data;
    *va_arg(ap, int*) = g_args.m;
    *va_arg(ap, int*) = g_args.n;
    *va_arg(ap, int*) = g_args.mb;
    *va_arg(ap, int*) = g_args.nb;
    *va_arg(ap, int*) = g_args.M;
    *va_arg(ap, int*) = g_args.ld;
    *va_arg(ap, unsigned int*) = g_args.seed;
    va_end(ap);
}

int ref_init_tile(double* data, int m, int n, int mb, int nb, int M, int ld, unsigned seed) {
    g_args.data = data; g_args.m = m; g_args.n = n; g_args.mb = mb; g_args.nb = nb; g_args.M = M; g_args.ld = ld;
    g_args.seed = seed;
    return initialize_tile(NULL, NULL);
}
