/* Force-included into every file of the reference when it is built without MPI (overlay/build_overlay.sh).
 * This fork names MPI_COMM_WORLD and MPI_Pack_size outside #ifdef PARSEC_HAVE_MPI (parsec/parsec.c:2156,
 * parsec/parsec_reshape.c:449); there is no MPI in the image, so the two names are given single-rank meanings. */
#include <stdint.h>
#define MPI_COMM_WORLD 0
extern int parsec_type_size(intptr_t type, int *size);
static inline int MPI_Pack_size(int count, intptr_t dtt, int comm, int *sz)
{
    int s = 0;
    (void)comm;
    parsec_type_size(dtt, &s);
    *sz = s * count;
    return 0;
}
