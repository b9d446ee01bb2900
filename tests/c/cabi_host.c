/* A plain C11 client of the C ABI (host-only entry points: no GPU needed).  Built and run by
 * tests/test_cabi_from_c.py; exits 0 when every check holds. */
#include <stdio.h>
#include <stdlib.h>

#include "parsec_b200.h"

#define CHECK(x)                                                  \
    do {                                                          \
        if (!(x)) {                                               \
            fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, #x); \
            return 1;                                             \
        }                                                         \
    } while (0)

int main(void) {
    /* HBM heap hooks (the job of zone_malloc_init / zone_malloc / zone_free, parsec/utils/zone_malloc.h:41-63) */
    void* z = pb2_zone_create(8, 4096);
    CHECK(z != NULL);
    int64_t a = pb2_zone_malloc(z, 4096), b = pb2_zone_malloc(z, 3 * 4096);
    CHECK(a >= 0 && b >= 0 && a % 4096 == 0 && b % 4096 == 0 && (a + 4096 <= b || b + 3 * 4096 <= a));
    CHECK(pb2_zone_in_use(z) == 4 * 4096);
    CHECK(pb2_zone_free(z, a) == PB2_SUCCESS && pb2_zone_malloc(z, 5 * 4096) == -1);
    pb2_zone_destroy(z);

    /* parsec_matrix_block_cyclic_init + rank_of / data_of on a 2 x 3 grid with 2 x 1 super-tiles */
    pb2_matrix_block_cyclic_t* dc = pb2_matrix_block_cyclic_new(PB2_MATRIX_DOUBLE, 4, 8, 8, 100, 60, 0, 0, 100, 60, 2, 3, 2, 1, 0, 0);
    CHECK(dc != NULL && pb2_matrix_mt(dc) == 13 && pb2_matrix_nt(dc) == 8);
    int64_t mine = 0;
    for (int n = 0; n < pb2_matrix_nt(dc); ++n)
        for (int m = 0; m < pb2_matrix_mt(dc); ++m) {
            int owner = pb2_matrix_rank_of(dc, m, n);
            CHECK(owner == ((m / 2) % 2) * 3 + n % 3);
            if (owner == 4) {
                ++mine;
                CHECK(pb2_matrix_data_of(dc, m, n) != NULL);
            } else {
                CHECK(pb2_matrix_data_of(dc, m, n) == NULL);
            }
        }
    CHECK(mine == pb2_matrix_nb_local_tiles(dc));
    pb2_matrix_block_cyclic_free(dc);

    /* the broadcast plan of one rank of a 2 x 2 grid */
    int64_t nops = pb2_comm_plan(2, 2, 3, 4, 4, 4, NULL, 0);
    CHECK(nops == 4 * (2 + 2));
    pb2_bcast_op_t* ops = (pb2_bcast_op_t*)malloc((size_t)nops * sizeof(*ops));
    CHECK(pb2_comm_plan(2, 2, 3, 4, 4, 4, ops, nops) == nops);
    for (int64_t i = 1; i < nops; ++i) CHECK(ops[i].k >= ops[i - 1].k);
    free(ops);

    /* the device module needs a CUDA device and says so instead of falling back */
    pb2_device_params_t p;
    pb2_device_params_default(&p);
    CHECK(p.max_streams == 18 && p.max_inflight_tasks == 256);
    printf("C ABI ok\n");
    return 0;
}
