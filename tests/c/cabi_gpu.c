/* The reference's own test shape in plain C against the C ABI: create the device module, three block-cyclic
 * descriptors over user memory (registered, like parsec_matrix_block_cyclic_init + dc->mat + register_memory),
 * C += A * B through the taskpool, then compare with a triple loop.  Built and run by tests/test_cabi_from_c.py on a
 * GPU box; exits 0 when the result matches.
 * Usage: cabi_gpu N NB */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "parsec_b200.h"

#define CHECK(x)                                                    \
    do {                                                            \
        if (!(x)) {                                                 \
            fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, #x); \
            return 1;                                               \
        }                                                           \
    } while (0)

/* element (i, j) of a tile-stored square matrix: tile (i / nb, j / nb) at position n * MT + m, column-major inside */
static double* at(double* mat, int MT, int nb, int i, int j) {
    return mat + ((size_t)(j / nb) * MT + (size_t)(i / nb)) * nb * nb + (size_t)(j % nb) * nb + (i % nb);
}

int main(int argc, char** argv) {
    const int N = argc > 1 ? atoi(argv[1]) : 256, NB = argc > 2 ? atoi(argv[2]) : 64, MT = N / NB;
    CHECK(N > 0 && NB > 0 && N % NB == 0);
    pb2_device_params_t params;
    pb2_device_params_default(&params);
    params.memory_number_of_blocks = 256; /* 128 MiB of HBM is plenty here */
    pb2_device_t* dev = NULL;
    CHECK(pb2_device_create(0, &params, &dev) == PB2_SUCCESS);

    pb2_matrix_block_cyclic_t* dc[3];
    double* mat[3];
    const size_t bytes = (size_t)N * N * sizeof(double);
    for (int i = 0; i < 3; ++i) {
        dc[i] = pb2_matrix_block_cyclic_new(PB2_MATRIX_DOUBLE, 0, NB, NB, N, N, 0, 0, N, N, 1, 1, 1, 1, 0, 0);
        CHECK(dc[i] != NULL && pb2_matrix_nb_local_tiles(dc[i]) == (int64_t)MT * MT);
        CHECK(posix_memalign((void**)&mat[i], 4096, bytes) == 0);
        CHECK(pb2_matrix_set_mat(dc[i], mat[i]) == PB2_SUCCESS);
        CHECK(pb2_matrix_register_memory(dc[i], dev) == PB2_SUCCESS); /* pins it for the copy engines */
        CHECK(pb2_matrix_fill_lcg(dc[i], 1789u + 16u * (unsigned)i, dev) == PB2_SUCCESS);
    }
    double* ref = (double*)malloc(bytes);
    CHECK(ref != NULL);
    for (size_t i = 0; i < (size_t)N * N; ++i) ref[i] = mat[2][i];

    pb2_taskpool_t* tp = pb2_gemm_taskpool_new(dc[0], dc[1], dc[2], dev);
    CHECK(tp != NULL);
    pb2_gemm_stats_t st;
    CHECK(pb2_gemm_taskpool_run(tp, &st) == PB2_SUCCESS);
    CHECK(st.nb_gemm == (int64_t)MT * MT * MT && st.nb_read_a == (int64_t)MT * MT && st.nb_read_b == (int64_t)MT * MT);
    pb2_gemm_taskpool_free(tp);

    double worst = 0.0, amax = 0.0, bmax = 0.0;
    for (int i = 0; i < N; ++i) {
        double ra = 0.0, rb = 0.0;
        for (int k = 0; k < N; ++k) {
            ra += fabs(*at(mat[0], MT, NB, i, k));
            rb += fabs(*at(mat[1], MT, NB, i, k));
        }
        amax = ra > amax ? ra : amax;
        bmax = rb > bmax ? rb : bmax;
    }
    for (int i = 0; i < N; ++i) {
        double rowerr = 0.0;
        for (int j = 0; j < N; ++j) {
            double acc = *at(ref, MT, NB, i, j);
            for (int k = 0; k < N; ++k) acc += *at(mat[0], MT, NB, i, k) * *at(mat[1], MT, NB, k, j);
            rowerr += fabs(acc - *at(mat[2], MT, NB, i, j));
        }
        worst = rowerr > worst ? rowerr : worst;
    }
    const double norm = worst / (amax * bmax * N * 1.1102230246251565e-16); /* the north-star norm, eps = 2^-53 */
    printf("C ABI gpu ok: N=%d NB=%d tasks=%lld norm=%.3g\n", N, NB, (long long)st.nb_gemm, norm);
    CHECK(norm < 10.0);

    for (int i = 0; i < 3; ++i) {
        uint32_t v = 99;
        if (i == 2) { /* every C tile ends at version KT */
            for (int n = 0; n < MT; ++n)
                for (int m = 0; m < MT; ++m)
                    CHECK(pb2_data_copy_version(pb2_matrix_data_of(dc[i], m, n), 0, &v) == PB2_SUCCESS && v == (uint32_t)MT);
        }
        CHECK(pb2_matrix_unregister_memory(dc[i], dev) == PB2_SUCCESS);
        pb2_matrix_block_cyclic_free(dc[i]);
        free(mat[i]);
    }
    free(ref);
    CHECK(pb2_device_destroy(dev) == PB2_SUCCESS);
    return 0;
}
