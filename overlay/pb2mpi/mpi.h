/*
 * mpi.h -- the subset of MPI the reference runtime names outside its MPI comm engine, for the ranks of ONE box.
 *
 * There is no MPI library in this image, and the reference cannot be built distributed without one: its datatype
 * layer is MPI's (parsec/datatype/datatype_mpi.c, parsec/datatype.h:12-41), and remote_dep_mpi.c / parsec.c /
 * vpmap.c name a handful of communicator calls.  libpb2mpi.so (pb2mpi.c) provides exactly those: derived datatypes
 * (contiguous, vector, hvector, indexed, indexed_block, struct, resized) with Pack / Unpack, communicators that are all
 * the one world of the job, and the three collectives the runtime uses, over libparsec_b200's pb2_ce_* transport.
 * Point-to-point (MPI_Isend / Irecv / Testsome ...) is NOT here: the file that used it, parsec_mpi_funnelled.c, is
 * replaced by overlay/comm_engine_b200.c.  A rank learns who it is from PB2_CE_RANK / PB2_CE_WORLD / PB2_CE_JOB
 * (overlay/pb2run sets them), or from torchrun's RANK / WORLD_SIZE / MASTER_PORT; neither: a world of one.
 */
#ifndef PB2MPI_MPI_H
#define PB2MPI_MPI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPI_VERSION 3
#define MPI_SUBVERSION 1
#define PB2MPI 1

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Info;
typedef int MPI_Request;
typedef int MPI_Errhandler;
typedef long MPI_Aint;
typedef long long MPI_Offset;
typedef long long MPI_Count;
typedef struct {
    int MPI_SOURCE, MPI_TAG, MPI_ERROR, pb2_count;
} MPI_Status;

#define MPI_SUCCESS 0
#define MPI_ERR_ARG 12
#define MPI_ERR_OTHER 15

#define MPI_COMM_NULL 0
#define MPI_COMM_WORLD 1
#define MPI_COMM_SELF 2
#define MPI_INFO_NULL 0
#define MPI_REQUEST_NULL 0
#define MPI_PROC_NULL (-2)
#define MPI_ANY_SOURCE (-1)
#define MPI_ANY_TAG (-1)
#define MPI_UNDEFINED (-32766)
#define MPI_IN_PLACE ((void*)-1)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_MAX_OBJECT_NAME 64
#define MPI_MAX_PROCESSOR_NAME 128
#define MPI_MAX_ERROR_STRING 128
#define MPI_COMM_TYPE_SHARED 1

#define MPI_IDENT 0
#define MPI_CONGRUENT 1
#define MPI_SIMILAR 2
#define MPI_UNEQUAL 3

#define MPI_THREAD_SINGLE 0
#define MPI_THREAD_FUNNELED 1
#define MPI_THREAD_SERIALIZED 2
#define MPI_THREAD_MULTIPLE 3

/* predefined datatypes: handles below PB2MPI_FIRST_DERIVED index a table of sizes (pb2mpi.c) */
#define MPI_DATATYPE_NULL 0
#define MPI_BYTE 1
#define MPI_PACKED 2
#define MPI_CHAR 3
#define MPI_SIGNED_CHAR 4
#define MPI_UNSIGNED_CHAR 5
#define MPI_SHORT 6
#define MPI_UNSIGNED_SHORT 7
#define MPI_INT 8
#define MPI_UNSIGNED 9
#define MPI_UINT MPI_UNSIGNED
#define MPI_LONG 10
#define MPI_UNSIGNED_LONG 11
#define MPI_LONG_LONG 12
#define MPI_LONG_LONG_INT MPI_LONG_LONG
#define MPI_UNSIGNED_LONG_LONG 13
#define MPI_FLOAT 14
#define MPI_DOUBLE 15
#define MPI_LONG_DOUBLE 16
#define MPI_INT8_T 17
#define MPI_INT16_T 18
#define MPI_INT32_T 19
#define MPI_INT64_T 20
#define MPI_UINT8_T 21
#define MPI_UINT16_T 22
#define MPI_UINT32_T 23
#define MPI_UINT64_T 24
#define MPI_C_COMPLEX 25
#define MPI_C_FLOAT_COMPLEX MPI_C_COMPLEX
#define MPI_C_DOUBLE_COMPLEX 26
#define MPI_AINT 27
#define MPI_LB 28 /* MPI-1 markers, only as members of MPI_Type_struct */
#define MPI_UB 29
#define MPI_2INT 30
#define MPI_DOUBLE_INT 31
#define PB2MPI_FIRST_DERIVED 64

#define MPI_COMBINER_NAMED 1
#define MPI_COMBINER_DUP 2
#define MPI_COMBINER_CONTIGUOUS 3
#define MPI_COMBINER_VECTOR 4
#define MPI_COMBINER_HVECTOR 5
#define MPI_COMBINER_INDEXED 6
#define MPI_COMBINER_INDEXED_BLOCK 7
#define MPI_COMBINER_STRUCT 8
#define MPI_COMBINER_RESIZED 9

#define MPI_MAX 1
#define MPI_MIN 2
#define MPI_SUM 3
#define MPI_MAXLOC 4
#define MPI_MINLOC 5
#define MPI_PROD 6
#define MPI_LAND 7
#define MPI_LOR 8
#define MPI_BAND 9
#define MPI_BOR 10

#define MPI_TAG_UB 1
#define MPI_ERRORS_RETURN 1
#define MPI_ERRORS_ARE_FATAL 2

/* ---- environment ---- */
int MPI_Init(int* argc, char*** argv);
int MPI_Init_thread(int* argc, char*** argv, int required, int* provided);
int MPI_Initialized(int* flag);
int MPI_Finalized(int* flag);
int MPI_Query_thread(int* provided);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm comm, int errorcode);
double MPI_Wtime(void);
int MPI_Get_processor_name(char* name, int* resultlen);
int MPI_Error_string(int errorcode, char* string, int* resultlen);

/* ---- communicators: every communicator of more than one rank is the world of the job ---- */
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_dup(MPI_Comm comm, MPI_Comm* newcomm);
int MPI_Comm_free(MPI_Comm* comm);
int MPI_Comm_compare(MPI_Comm a, MPI_Comm b, int* result);
int MPI_Comm_split_type(MPI_Comm comm, int split_type, int key, MPI_Info info, MPI_Comm* newcomm);
int MPI_Comm_set_errhandler(MPI_Comm comm, MPI_Errhandler e);
int MPI_Comm_get_attr(MPI_Comm comm, int keyval, void* attribute_val, int* flag);

/* ---- collectives (the ones the runtime and the test programs use) ---- */
int MPI_Barrier(MPI_Comm comm);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Reduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm comm);
int MPI_Bcast(void* buf, int count, MPI_Datatype type, int root, MPI_Comm comm);
int MPI_Allgather(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf, int recvcount,
                  MPI_Datatype recvtype, MPI_Comm comm);
int MPI_Gather(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf, int recvcount,
               MPI_Datatype recvtype, int root, MPI_Comm comm);
int MPI_Gatherv(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf, const int* recvcounts,
                const int* displs, MPI_Datatype recvtype, int root, MPI_Comm comm);

/* ---- datatypes ---- */
int MPI_Type_size(MPI_Datatype type, int* size);
int MPI_Type_get_extent(MPI_Datatype type, MPI_Aint* lb, MPI_Aint* extent);
int MPI_Type_extent(MPI_Datatype type, MPI_Aint* extent);
int MPI_Type_get_true_extent(MPI_Datatype type, MPI_Aint* true_lb, MPI_Aint* true_extent);
int MPI_Type_commit(MPI_Datatype* type);
int MPI_Type_free(MPI_Datatype* type);
int MPI_Type_dup(MPI_Datatype type, MPI_Datatype* newtype);
int MPI_Type_contiguous(int count, MPI_Datatype oldtype, MPI_Datatype* newtype);
int MPI_Type_vector(int count, int blocklength, int stride, MPI_Datatype oldtype, MPI_Datatype* newtype);
int MPI_Type_create_hvector(int count, int blocklength, MPI_Aint stride, MPI_Datatype oldtype, MPI_Datatype* newtype);
int MPI_Type_indexed(int count, const int* blocklengths, const int* displacements, MPI_Datatype oldtype,
                     MPI_Datatype* newtype);
int MPI_Type_create_indexed_block(int count, int blocklength, const int* displacements, MPI_Datatype oldtype,
                                  MPI_Datatype* newtype);
int MPI_Type_create_struct(int count, const int* blocklengths, const MPI_Aint* displacements, const MPI_Datatype* types,
                           MPI_Datatype* newtype);
int MPI_Type_struct(int count, int* blocklengths, MPI_Aint* displacements, MPI_Datatype* types, MPI_Datatype* newtype);
int MPI_Type_create_resized(MPI_Datatype oldtype, MPI_Aint lb, MPI_Aint extent, MPI_Datatype* newtype);
int MPI_Type_set_name(MPI_Datatype type, const char* name);
int MPI_Type_get_name(MPI_Datatype type, char* name, int* resultlen);
int MPI_Type_get_envelope(MPI_Datatype type, int* num_integers, int* num_addresses, int* num_datatypes, int* combiner);
int MPI_Pack_size(int incount, MPI_Datatype type, MPI_Comm comm, int* size);
int MPI_Pack(const void* inbuf, int incount, MPI_Datatype type, void* outbuf, int outsize, int* position, MPI_Comm comm);
int MPI_Unpack(const void* inbuf, int insize, int* position, void* outbuf, int outcount, MPI_Datatype type,
               MPI_Comm comm);
int MPI_Get_address(const void* location, MPI_Aint* address);

/* ---- not MPI: what overlay/comm_engine_b200.c needs from this layer ---- */
typedef struct {
    int64_t disp; /* bytes from the buffer address */
    int64_t len;  /* contiguous bytes */
} pb2mpi_segment_t;
/* the byte layout of ONE element of `type`, adjacent pieces merged, in packing order; valid until the type is freed */
int PB2MPI_Type_layout(MPI_Datatype type, const pb2mpi_segment_t** segments, int* nsegments, int64_t* extent,
                       int64_t* size);
/* copy `scount` elements of `stype` at src into `dcount` elements of `dtype` at dst (what a send-to-self does) */
int PB2MPI_Copy(const void* src, int scount, MPI_Datatype stype, void* dst, int dcount, MPI_Datatype dtype);
/* the transport of the job's world (a pb2_ce_t*, include/parsec_b200.h), NULL in a world of one */
void* PB2MPI_Transport(void);

#ifdef __cplusplus
}
#endif
#endif
