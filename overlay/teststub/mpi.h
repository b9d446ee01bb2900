/*
 * mpi.h -- single-rank stand-ins for the handful of MPI names the reference's own CUDA tests use
 * (tests/dsl/ptg/cuda/stage_main.c, get_best_device_check.jdf, testing_get_best_device.c,
 * tests/dsl/dtd/dtd_test_simple_gemm.c).  The image has no MPI; these tests only ask for their rank and size and sum
 * one int.  Only on the include path of those test programs (overlay/build_overlay.sh), never of the runtime.
 */
#ifndef PB2_TESTSTUB_MPI_H
#define PB2_TESTSTUB_MPI_H
#include <stdlib.h>
#include <string.h>
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
#ifndef MPI_COMM_WORLD
#define MPI_COMM_WORLD 0
#endif
#define MPI_SUCCESS 0
#define MPI_INT 4
#define MPI_SUM 1
#define MPI_THREAD_SERIALIZED 2
#define MPI_THREAD_MULTIPLE 3
static inline int MPI_Init_thread(int *argc, char ***argv, int required, int *provided)
{ (void)argc; (void)argv; if( provided ) *provided = required; return MPI_SUCCESS; }
static inline int MPI_Comm_size(MPI_Comm comm, int *size) { (void)comm; *size = 1; return MPI_SUCCESS; }
static inline int MPI_Comm_rank(MPI_Comm comm, int *rank) { (void)comm; *rank = 0; return MPI_SUCCESS; }
static inline int MPI_Finalize(void) { return MPI_SUCCESS; }
static inline int MPI_Abort(MPI_Comm comm, int code) { (void)comm; exit(code); return MPI_SUCCESS; }
static inline int MPI_Allreduce(const void *send, void *recv, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm)
{ (void)op; (void)comm; memcpy(recv, send, (size_t)count * (size_t)dt); return MPI_SUCCESS; }
#endif
