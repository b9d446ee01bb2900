/* Link-time stand-ins for the three symbols of the two MPI-only sources (parsec_mpi_funnelled.c, remote_dep_mpi.c)
 * that a build without MPI drops; none is reached with one rank (parsec_comm_engine.c:17,27, parsec_reshape.c:135). */
#include <stdio.h>
#include <stdlib.h>
void *mpi_funnelled_init(void *ctx) { (void)ctx; return NULL; }
int mpi_funnelled_fini(void *ce) { (void)ce; return 0; }
void parsec_local_reshape(void *future, void **in_data)
{
    (void)future; (void)in_data;
    fprintf(stderr, "parsec_local_reshape: not available in a build without MPI\n");
    abort();
}
