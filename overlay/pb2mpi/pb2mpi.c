/*
 * pb2mpi.c -- libpb2mpi.so: the MPI subset of overlay/pb2mpi/mpi.h for the ranks of one box.
 *
 * Datatypes are kept flattened: the byte layout of one element as a list of (displacement, length) pieces in packing
 * order plus lower bound and extent, built by the MPI-3.1 rules of each constructor (4.1.2-4.1.7).  That is all the
 * reference asks of a datatype: size / extent for its arenas (parsec/arena.c), Pack / Unpack of the activation
 * messages (remote_dep_mpi.c:1319-1368, :1848), and the piece list the comm engine turns into one vectored copy.
 * Communicators: handles only; every one is the job's world (or the process itself).  Collectives: pb2_ce_*.
 */
#define _GNU_SOURCE
#include "mpi.h"

#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include "parsec_b200.h"

typedef struct {
    int used, committed, combiner;
    int64_t size, lb, ub; /* ub - lb = extent */
    int nseg;
    pb2mpi_segment_t *seg;
    char name[MPI_MAX_OBJECT_NAME];
} type_t;

static const struct { int size; const char *name; } predefined[PB2MPI_FIRST_DERIVED] = {
    [MPI_BYTE] = {1, "MPI_BYTE"}, [MPI_PACKED] = {1, "MPI_PACKED"}, [MPI_CHAR] = {1, "MPI_CHAR"},
    [MPI_SIGNED_CHAR] = {1, "MPI_SIGNED_CHAR"}, [MPI_UNSIGNED_CHAR] = {1, "MPI_UNSIGNED_CHAR"},
    [MPI_SHORT] = {2, "MPI_SHORT"}, [MPI_UNSIGNED_SHORT] = {2, "MPI_UNSIGNED_SHORT"}, [MPI_INT] = {4, "MPI_INT"},
    [MPI_UNSIGNED] = {4, "MPI_UNSIGNED"}, [MPI_LONG] = {sizeof(long), "MPI_LONG"},
    [MPI_UNSIGNED_LONG] = {sizeof(long), "MPI_UNSIGNED_LONG"}, [MPI_LONG_LONG] = {8, "MPI_LONG_LONG"},
    [MPI_UNSIGNED_LONG_LONG] = {8, "MPI_UNSIGNED_LONG_LONG"}, [MPI_FLOAT] = {4, "MPI_FLOAT"},
    [MPI_DOUBLE] = {8, "MPI_DOUBLE"}, [MPI_LONG_DOUBLE] = {sizeof(long double), "MPI_LONG_DOUBLE"},
    [MPI_INT8_T] = {1, "MPI_INT8_T"}, [MPI_INT16_T] = {2, "MPI_INT16_T"}, [MPI_INT32_T] = {4, "MPI_INT32_T"},
    [MPI_INT64_T] = {8, "MPI_INT64_T"}, [MPI_UINT8_T] = {1, "MPI_UINT8_T"}, [MPI_UINT16_T] = {2, "MPI_UINT16_T"},
    [MPI_UINT32_T] = {4, "MPI_UINT32_T"}, [MPI_UINT64_T] = {8, "MPI_UINT64_T"}, [MPI_C_COMPLEX] = {8, "MPI_C_COMPLEX"},
    [MPI_C_DOUBLE_COMPLEX] = {16, "MPI_C_DOUBLE_COMPLEX"}, [MPI_AINT] = {sizeof(MPI_Aint), "MPI_AINT"},
    [MPI_LB] = {0, "MPI_LB"}, [MPI_UB] = {0, "MPI_UB"}, [MPI_2INT] = {8, "MPI_2INT"},
    [MPI_DOUBLE_INT] = {16, "MPI_DOUBLE_INT"},
};

#define MAX_TYPES 4096
static type_t types[MAX_TYPES];
static pb2mpi_segment_t predefined_seg[PB2MPI_FIRST_DERIVED];
static char predefined_name[PB2MPI_FIRST_DERIVED][MPI_MAX_OBJECT_NAME];
static pthread_mutex_t type_lock = PTHREAD_MUTEX_INITIALIZER;

/* ---- the job ---- */
static int initialized = 0, finalized = 0, thread_level = MPI_THREAD_SINGLE;
static int my_rank = 0, world_size = 1;
static pb2_ce_t *transport = NULL;
static int next_comm = 16;

/* A view of any datatype as (lb, ub, size, pieces).  Returns 0 for a handle that names nothing. */
static int view(MPI_Datatype t, type_t *out)
{
    if( t > 0 && t < PB2MPI_FIRST_DERIVED && (predefined[t].size > 0 || MPI_LB == t || MPI_UB == t) ) {
        memset(out, 0, sizeof(*out));
        out->used = out->committed = 1;
        out->combiner = MPI_COMBINER_NAMED;
        out->size = predefined[t].size;
        out->lb = 0;
        out->ub = predefined[t].size;
        predefined_seg[t].disp = 0;
        predefined_seg[t].len = predefined[t].size;
        out->nseg = predefined[t].size > 0 ? 1 : 0;
        out->seg = &predefined_seg[t];
        return 1;
    }
    if( t >= PB2MPI_FIRST_DERIVED && t < PB2MPI_FIRST_DERIVED + MAX_TYPES && types[t - PB2MPI_FIRST_DERIVED].used ) {
        *out = types[t - PB2MPI_FIRST_DERIVED];
        return 1;
    }
    return 0;
}

/* ---- builder: pieces appended in packing order, neighbours merged ---- */
typedef struct {
    pb2mpi_segment_t *seg;
    int n, cap;
    int64_t size, lb, ub;
    int any, lb_marked, ub_marked;
} builder_t;

static void b_piece(builder_t *b, int64_t disp, int64_t len)
{
    if( len <= 0 ) return;
    if( b->n && b->seg[b->n - 1].disp + b->seg[b->n - 1].len == disp ) { b->seg[b->n - 1].len += len; return; }
    if( b->n == b->cap ) {
        b->cap = b->cap ? 2 * b->cap : 16;
        b->seg = (pb2mpi_segment_t*)realloc(b->seg, (size_t)b->cap * sizeof(pb2mpi_segment_t));
    }
    b->seg[b->n].disp = disp;
    b->seg[b->n].len = len;
    b->n++;
}

/* `count` consecutive copies of `old` (spaced by its extent), the first at byte displacement `disp` */
static void b_place(builder_t *b, const type_t *old, int64_t disp, int64_t count)
{
    const int64_t ext = old->ub - old->lb;
    if( count <= 0 ) return;
    for( int64_t c = 0; c < count; c++ )
        for( int s = 0; s < old->nseg; s++ ) b_piece(b, disp + c * ext + old->seg[s].disp, old->seg[s].len);
    b->size += count * old->size;
    /* bounds of the block: with a negative extent the last copy is the lowest */
    int64_t lo = disp + old->lb + (ext < 0 ? (count - 1) * ext : 0);
    int64_t hi = disp + old->ub + (ext > 0 ? (count - 1) * ext : 0);
    if( !b->lb_marked && (!b->any || lo < b->lb) ) b->lb = lo;
    if( !b->ub_marked && (!b->any || hi > b->ub) ) b->ub = hi;
    b->any = 1;
}

static int b_finish(builder_t *b, int combiner, MPI_Datatype *newtype)
{
    int slot = -1;
    pthread_mutex_lock(&type_lock);
    for( int i = 0; i < MAX_TYPES; i++ ) if( !types[i].used ) { slot = i; break; }
    if( slot < 0 ) { pthread_mutex_unlock(&type_lock); free(b->seg); return MPI_ERR_OTHER; }
    type_t *t = &types[slot];
    memset(t, 0, sizeof(*t));
    t->used = 1;
    t->combiner = combiner;
    t->size = b->size;
    t->lb = b->any || b->lb_marked ? b->lb : 0;
    t->ub = b->any || b->ub_marked ? b->ub : 0;
    t->nseg = b->n;
    t->seg = b->seg;
    pthread_mutex_unlock(&type_lock);
    *newtype = PB2MPI_FIRST_DERIVED + slot;
    return MPI_SUCCESS;
}

/* ---- constructors ---- */
int MPI_Type_contiguous(int count, MPI_Datatype oldtype, MPI_Datatype *newtype)
{
    type_t old;
    builder_t b = {0};
    if( count < 0 || !view(oldtype, &old) ) return MPI_ERR_ARG;
    b_place(&b, &old, 0, count);
    return b_finish(&b, MPI_COMBINER_CONTIGUOUS, newtype);
}

int MPI_Type_create_hvector(int count, int blocklength, MPI_Aint stride, MPI_Datatype oldtype, MPI_Datatype *newtype)
{
    type_t old;
    builder_t b = {0};
    if( count < 0 || blocklength < 0 || !view(oldtype, &old) ) return MPI_ERR_ARG;
    for( int i = 0; i < count; i++ ) b_place(&b, &old, (int64_t)i * stride, blocklength);
    return b_finish(&b, MPI_COMBINER_HVECTOR, newtype);
}

int MPI_Type_vector(int count, int blocklength, int stride, MPI_Datatype oldtype, MPI_Datatype *newtype)
{
    type_t old;
    if( !view(oldtype, &old) ) return MPI_ERR_ARG;
    int rc = MPI_Type_create_hvector(count, blocklength, (MPI_Aint)stride * (old.ub - old.lb), oldtype, newtype);
    if( MPI_SUCCESS == rc ) types[*newtype - PB2MPI_FIRST_DERIVED].combiner = MPI_COMBINER_VECTOR;
    return rc;
}

int MPI_Type_indexed(int count, const int *blocklengths, const int *displacements, MPI_Datatype oldtype,
                     MPI_Datatype *newtype)
{
    type_t old;
    builder_t b = {0};
    if( count < 0 || !view(oldtype, &old) ) return MPI_ERR_ARG;
    for( int i = 0; i < count; i++ ) b_place(&b, &old, (int64_t)displacements[i] * (old.ub - old.lb), blocklengths[i]);
    return b_finish(&b, MPI_COMBINER_INDEXED, newtype);
}

int MPI_Type_create_indexed_block(int count, int blocklength, const int *displacements, MPI_Datatype oldtype,
                                  MPI_Datatype *newtype)
{
    type_t old;
    builder_t b = {0};
    if( count < 0 || !view(oldtype, &old) ) return MPI_ERR_ARG;
    for( int i = 0; i < count; i++ ) b_place(&b, &old, (int64_t)displacements[i] * (old.ub - old.lb), blocklength);
    return b_finish(&b, MPI_COMBINER_INDEXED_BLOCK, newtype);
}

int MPI_Type_create_struct(int count, const int *blocklengths, const MPI_Aint *displacements, const MPI_Datatype *tps,
                           MPI_Datatype *newtype)
{
    builder_t b = {0};
    if( count < 0 ) return MPI_ERR_ARG;
    for( int i = 0; i < count; i++ ) {   /* MPI-1 markers first: they fix a bound whatever the members say */
        if( MPI_LB == tps[i] ) { b.lb = displacements[i]; b.lb_marked = 1; }
        if( MPI_UB == tps[i] ) { b.ub = displacements[i]; b.ub_marked = 1; }
    }
    for( int i = 0; i < count; i++ ) {
        type_t old;
        if( MPI_LB == tps[i] || MPI_UB == tps[i] ) continue;
        if( !view(tps[i], &old) ) { free(b.seg); return MPI_ERR_ARG; }
        b_place(&b, &old, displacements[i], blocklengths[i]);
    }
    return b_finish(&b, MPI_COMBINER_STRUCT, newtype);
}

int MPI_Type_struct(int count, int *blocklengths, MPI_Aint *displacements, MPI_Datatype *tps, MPI_Datatype *newtype)
{
    return MPI_Type_create_struct(count, blocklengths, displacements, tps, newtype);
}

int MPI_Type_create_resized(MPI_Datatype oldtype, MPI_Aint lb, MPI_Aint extent, MPI_Datatype *newtype)
{
    type_t old;
    builder_t b = {0};
    if( !view(oldtype, &old) ) return MPI_ERR_ARG;
    for( int s = 0; s < old.nseg; s++ ) b_piece(&b, old.seg[s].disp, old.seg[s].len);
    b.size = old.size;
    b.lb = lb;
    b.ub = lb + extent;
    b.lb_marked = b.ub_marked = 1;
    return b_finish(&b, MPI_COMBINER_RESIZED, newtype);
}

int MPI_Type_dup(MPI_Datatype type, MPI_Datatype *newtype)
{
    type_t old;
    if( !view(type, &old) ) return MPI_ERR_ARG;
    int rc = MPI_Type_create_resized(type, old.lb, old.ub - old.lb, newtype);
    if( MPI_SUCCESS == rc ) types[*newtype - PB2MPI_FIRST_DERIVED].combiner = MPI_COMBINER_DUP;
    return rc;
}

int MPI_Type_commit(MPI_Datatype *type)
{
    type_t t;
    if( NULL == type || !view(*type, &t) ) return MPI_ERR_ARG;
    if( *type >= PB2MPI_FIRST_DERIVED ) types[*type - PB2MPI_FIRST_DERIVED].committed = 1;
    return MPI_SUCCESS;
}

int MPI_Type_free(MPI_Datatype *type)
{
    if( NULL == type || *type < PB2MPI_FIRST_DERIVED || *type >= PB2MPI_FIRST_DERIVED + MAX_TYPES ) return MPI_ERR_ARG;
    pthread_mutex_lock(&type_lock);
    type_t *t = &types[*type - PB2MPI_FIRST_DERIVED];
    if( t->used ) { free(t->seg); memset(t, 0, sizeof(*t)); }
    pthread_mutex_unlock(&type_lock);
    *type = MPI_DATATYPE_NULL;
    return MPI_SUCCESS;
}

/* ---- queries ---- */
int MPI_Type_size(MPI_Datatype type, int *size)
{
    type_t t;
    if( !view(type, &t) ) return MPI_ERR_ARG;
    *size = (int)t.size;
    return MPI_SUCCESS;
}

int MPI_Type_get_extent(MPI_Datatype type, MPI_Aint *lb, MPI_Aint *extent)
{
    type_t t;
    if( !view(type, &t) ) return MPI_ERR_ARG;
    if( lb ) *lb = (MPI_Aint)t.lb;
    if( extent ) *extent = (MPI_Aint)(t.ub - t.lb);
    return MPI_SUCCESS;
}

int MPI_Type_extent(MPI_Datatype type, MPI_Aint *extent) { return MPI_Type_get_extent(type, NULL, extent); }

int MPI_Type_get_true_extent(MPI_Datatype type, MPI_Aint *true_lb, MPI_Aint *true_extent)
{
    type_t t;
    if( !view(type, &t) ) return MPI_ERR_ARG;
    int64_t lo = 0, hi = 0;
    for( int s = 0; s < t.nseg; s++ ) {
        if( 0 == s || t.seg[s].disp < lo ) lo = t.seg[s].disp;
        if( 0 == s || t.seg[s].disp + t.seg[s].len > hi ) hi = t.seg[s].disp + t.seg[s].len;
    }
    if( true_lb ) *true_lb = (MPI_Aint)lo;
    if( true_extent ) *true_extent = (MPI_Aint)(hi - lo);
    return MPI_SUCCESS;
}

int MPI_Type_set_name(MPI_Datatype type, const char *name)
{
    if( type >= PB2MPI_FIRST_DERIVED && type < PB2MPI_FIRST_DERIVED + MAX_TYPES && types[type - PB2MPI_FIRST_DERIVED].used ) {
        snprintf(types[type - PB2MPI_FIRST_DERIVED].name, MPI_MAX_OBJECT_NAME, "%s", name);
        return MPI_SUCCESS;
    }
    if( type > 0 && type < PB2MPI_FIRST_DERIVED ) {
        snprintf(predefined_name[type], MPI_MAX_OBJECT_NAME, "%s", name);
        return MPI_SUCCESS;
    }
    return MPI_ERR_ARG;
}

int MPI_Type_get_name(MPI_Datatype type, char *name, int *resultlen)
{
    const char *src = "";
    if( type >= PB2MPI_FIRST_DERIVED && type < PB2MPI_FIRST_DERIVED + MAX_TYPES && types[type - PB2MPI_FIRST_DERIVED].used )
        src = types[type - PB2MPI_FIRST_DERIVED].name;
    else if( type > 0 && type < PB2MPI_FIRST_DERIVED )
        src = predefined_name[type][0] ? predefined_name[type] : (predefined[type].name ? predefined[type].name : "");
    else if( MPI_DATATYPE_NULL == type )
        src = "MPI_DATATYPE_NULL";
    else
        return MPI_ERR_ARG;
    int n = snprintf(name, MPI_MAX_OBJECT_NAME, "%s", src);
    if( resultlen ) *resultlen = n < MPI_MAX_OBJECT_NAME ? n : MPI_MAX_OBJECT_NAME - 1;
    return MPI_SUCCESS;
}

int MPI_Type_get_envelope(MPI_Datatype type, int *num_integers, int *num_addresses, int *num_datatypes, int *combiner)
{
    type_t t;
    if( !view(type, &t) ) return MPI_ERR_ARG;
    if( num_integers ) *num_integers = 0;   /* the constructor arguments are not kept: only the combiner is answered */
    if( num_addresses ) *num_addresses = 0;
    if( num_datatypes ) *num_datatypes = MPI_COMBINER_NAMED == t.combiner ? 0 : 1;
    if( combiner ) *combiner = t.combiner;
    return MPI_SUCCESS;
}

int MPI_Get_address(const void *location, MPI_Aint *address)
{
    *address = (MPI_Aint)(intptr_t)location;
    return MPI_SUCCESS;
}

/* ---- pack / unpack ---- */
int MPI_Pack_size(int incount, MPI_Datatype type, MPI_Comm comm, int *size)
{
    type_t t;
    (void)comm;
    if( incount < 0 || !view(type, &t) ) return MPI_ERR_ARG;
    *size = (int)(incount * t.size);
    return MPI_SUCCESS;
}

int MPI_Pack(const void *inbuf, int incount, MPI_Datatype type, void *outbuf, int outsize, int *position, MPI_Comm comm)
{
    type_t t;
    (void)comm;
    if( incount < 0 || !view(type, &t) || *position + incount * t.size > outsize ) return MPI_ERR_ARG;
    char *out = (char*)outbuf + *position;
    const int64_t ext = t.ub - t.lb;
    for( int i = 0; i < incount; i++ )
        for( int s = 0; s < t.nseg; s++ ) {
            memcpy(out, (const char*)inbuf + i * ext + t.seg[s].disp, (size_t)t.seg[s].len);
            out += t.seg[s].len;
        }
    *position += (int)(incount * t.size);
    return MPI_SUCCESS;
}

int MPI_Unpack(const void *inbuf, int insize, int *position, void *outbuf, int outcount, MPI_Datatype type, MPI_Comm comm)
{
    type_t t;
    (void)comm;
    if( outcount < 0 || !view(type, &t) || *position + outcount * t.size > insize ) return MPI_ERR_ARG;
    const char *in = (const char*)inbuf + *position;
    const int64_t ext = t.ub - t.lb;
    for( int i = 0; i < outcount; i++ )
        for( int s = 0; s < t.nseg; s++ ) {
            memcpy((char*)outbuf + i * ext + t.seg[s].disp, in, (size_t)t.seg[s].len);
            in += t.seg[s].len;
        }
    *position += (int)(outcount * t.size);
    return MPI_SUCCESS;
}

int PB2MPI_Type_layout(MPI_Datatype type, const pb2mpi_segment_t **segments, int *nsegments, int64_t *extent, int64_t *size)
{
    type_t t;
    if( !view(type, &t) ) return MPI_ERR_ARG;
    if( segments ) *segments = t.seg;
    if( nsegments ) *nsegments = t.nseg;
    if( extent ) *extent = t.ub - t.lb;
    if( size ) *size = t.size;
    return MPI_SUCCESS;
}

/* A cursor over the bytes of `count` elements of a type, in packing order. */
typedef struct { const type_t *t; char *base; int64_t ext; int count, elem, seg; int64_t off; } cursor_t;
static int64_t cursor_piece(cursor_t *c, char **p)
{
    while( c->elem < c->count ) {
        if( c->seg < c->t->nseg ) {
            *p = c->base + c->elem * c->ext + c->t->seg[c->seg].disp + c->off;
            return c->t->seg[c->seg].len - c->off;
        }
        c->seg = 0;
        c->elem++;
    }
    return 0;
}
static void cursor_advance(cursor_t *c, int64_t n)
{
    c->off += n;
    if( c->off == c->t->seg[c->seg].len ) { c->off = 0; c->seg++; }
}

int PB2MPI_Copy(const void *src, int scount, MPI_Datatype stype, void *dst, int dcount, MPI_Datatype dtype)
{
    type_t st, dt;
    if( !view(stype, &st) || !view(dtype, &dt) ) return MPI_ERR_ARG;
    if( (int64_t)scount * st.size > (int64_t)dcount * dt.size ) return MPI_ERR_ARG;   /* truncation */
    cursor_t s = {&st, (char*)src, st.ub - st.lb, scount, 0, 0, 0}, d = {&dt, (char*)dst, dt.ub - dt.lb, dcount, 0, 0, 0};
    char *sp, *dp;
    int64_t sn, dn;
    while( (sn = cursor_piece(&s, &sp)) > 0 && (dn = cursor_piece(&d, &dp)) > 0 ) {
        int64_t n = sn < dn ? sn : dn;
        memmove(dp, sp, (size_t)n);
        cursor_advance(&s, n);
        cursor_advance(&d, n);
    }
    return MPI_SUCCESS;
}

/* ---- environment ---- */
static const char *env_first(const char *a, const char *b)
{
    const char *v = getenv(a);
    return (NULL != v && v[0]) ? v : getenv(b);
}

int MPI_Init_thread(int *argc, char ***argv, int required, int *provided)
{
    (void)argc; (void)argv;
    if( initialized ) { if( provided ) *provided = thread_level; return MPI_SUCCESS; }
    const char *r = env_first("PB2_CE_RANK", "RANK"), *w = env_first("PB2_CE_WORLD", "WORLD_SIZE");
    const char *job = env_first("PB2_CE_JOB", "MASTER_PORT");
    my_rank = r ? atoi(r) : 0;
    world_size = w ? atoi(w) : 1;
    if( world_size < 1 || my_rank < 0 || my_rank >= world_size ) { my_rank = 0; world_size = 1; }
    if( world_size > 1 ) {
        const char *ring = getenv("PB2_CE_RING_BYTES");
        char name[96];
        const char *run = getenv("TORCHELASTIC_RUN_ID");   /* torchrun: tells two launches on the same port apart */
        if( NULL != getenv("PB2_CE_JOB") || NULL == run ) snprintf(name, sizeof(name), "%s", job ? job : "default");
        else snprintf(name, sizeof(name), "%.40s.%.40s", run, job ? job : "default");
        for( char *c = name; *c; c++ ) if( '/' == *c ) *c = '_';
        int rc = pb2_ce_attach(name, my_rank, world_size, ring ? (size_t)strtoull(ring, NULL, 0) : 0, &transport);
        if( PB2_SUCCESS != rc ) {
            fprintf(stderr, "pb2mpi: rank %d of %d could not join job '%s' (pb2_ce_attach: %d)\n", my_rank, world_size, name, rc);
            return MPI_ERR_OTHER;
        }
    }
    /* MPI_THREAD_MULTIPLE can be granted: the collectives here work on the job's shared segment, not on the rings, and
     * pb2_ce_send takes a lock per destination; only polling is for one thread (the comm thread of the runtime) */
    thread_level = required > MPI_THREAD_MULTIPLE ? MPI_THREAD_MULTIPLE : required;
    if( provided ) *provided = thread_level;
    initialized = 1;
    return MPI_SUCCESS;
}

int MPI_Init(int *argc, char ***argv) { return MPI_Init_thread(argc, argv, MPI_THREAD_SINGLE, NULL); }
int MPI_Initialized(int *flag) { *flag = initialized; return MPI_SUCCESS; }
int MPI_Finalized(int *flag) { *flag = finalized; return MPI_SUCCESS; }
int MPI_Query_thread(int *provided) { *provided = thread_level; return MPI_SUCCESS; }

int MPI_Finalize(void)
{
    if( NULL != transport ) { pb2_ce_barrier(transport); pb2_ce_detach(transport); transport = NULL; }
    finalized = 1;
    return MPI_SUCCESS;
}

int MPI_Abort(MPI_Comm comm, int errorcode)
{
    (void)comm;
    if( NULL != transport ) pb2_ce_mark_failed(transport);   /* the other ranks stop waiting for this one */
    fflush(NULL);
    _exit(errorcode ? errorcode : 1);
}

double MPI_Wtime(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int MPI_Get_processor_name(char *name, int *resultlen)
{
    if( 0 != gethostname(name, MPI_MAX_PROCESSOR_NAME) ) snprintf(name, MPI_MAX_PROCESSOR_NAME, "localhost");
    name[MPI_MAX_PROCESSOR_NAME - 1] = '\0';
    *resultlen = (int)strlen(name);
    return MPI_SUCCESS;
}

int MPI_Error_string(int errorcode, char *string, int *resultlen)
{
    *resultlen = snprintf(string, MPI_MAX_ERROR_STRING, "pb2mpi error %d", errorcode);
    return MPI_SUCCESS;
}

void *PB2MPI_Transport(void) { return transport; }

/* ---- communicators ---- */
static int comm_is_self(MPI_Comm c) { return MPI_COMM_SELF == c || (c >= 16 && (c & 1)); }   /* odd handles: dups of SELF */

int MPI_Comm_size(MPI_Comm comm, int *size)
{
    if( MPI_COMM_NULL == comm ) return MPI_ERR_ARG;
    *size = comm_is_self(comm) ? 1 : world_size;
    return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm comm, int *rank)
{
    if( MPI_COMM_NULL == comm ) return MPI_ERR_ARG;
    *rank = comm_is_self(comm) ? 0 : my_rank;
    return MPI_SUCCESS;
}

int MPI_Comm_dup(MPI_Comm comm, MPI_Comm *newcomm)
{
    if( MPI_COMM_NULL == comm ) return MPI_ERR_ARG;
    int id = __sync_fetch_and_add(&next_comm, 2);
    *newcomm = comm_is_self(comm) ? (id | 1) : (id & ~1);
    return MPI_SUCCESS;
}

int MPI_Comm_free(MPI_Comm *comm) { *comm = MPI_COMM_NULL; return MPI_SUCCESS; }

int MPI_Comm_compare(MPI_Comm a, MPI_Comm b, int *result)
{
    if( MPI_COMM_NULL == a || MPI_COMM_NULL == b ) return MPI_ERR_ARG;
    *result = (a == b) ? MPI_IDENT : (comm_is_self(a) == comm_is_self(b) ? MPI_CONGRUENT : MPI_UNEQUAL);
    return MPI_SUCCESS;
}

int MPI_Comm_split_type(MPI_Comm comm, int split_type, int key, MPI_Info info, MPI_Comm *newcomm)
{
    (void)split_type; (void)key; (void)info;
    return MPI_Comm_dup(comm, newcomm);   /* one box: the shared-memory group is the whole communicator */
}

int MPI_Comm_set_errhandler(MPI_Comm comm, MPI_Errhandler e) { (void)comm; (void)e; return MPI_SUCCESS; }

int MPI_Comm_get_attr(MPI_Comm comm, int keyval, void *attribute_val, int *flag)
{
    static int tag_ub = 0x7fffffff;
    (void)comm;
    if( MPI_TAG_UB == keyval ) { *(int**)attribute_val = &tag_ub; *flag = 1; return MPI_SUCCESS; }
    *flag = 0;
    return MPI_SUCCESS;
}

/* ---- collectives ---- */
int MPI_Barrier(MPI_Comm comm)
{
    if( comm_is_self(comm) || NULL == transport ) return MPI_SUCCESS;
    return PB2_SUCCESS == pb2_ce_barrier(transport) ? MPI_SUCCESS : MPI_ERR_OTHER;
}

/* every rank's `bytes` bytes, ordered by rank, into all (world_size * bytes); 128 bytes per rank per round */
static int allgather_bytes(const void *mine, size_t bytes, void *all)
{
    char *rounds = (char*)malloc((size_t)world_size * 128);
    for( size_t done = 0; done < bytes; done += 128 ) {
        const size_t n = bytes - done < 128 ? bytes - done : 128;
        char piece[128] = {0};
        memcpy(piece, (const char*)mine + done, n);
        if( PB2_SUCCESS != pb2_ce_allgather(transport, piece, 128, rounds) ) { free(rounds); return MPI_ERR_OTHER; }
        for( int r = 0; r < world_size; r++ ) memcpy((char*)all + (size_t)r * bytes + done, rounds + (size_t)r * 128, n);
    }
    free(rounds);
    return MPI_SUCCESS;
}

#define REDUCE_LOOP(T) do { T *o = (T*)out; const T *x = (const T*)in; for( int i = 0; i < count; i++ ) {                \
        switch( op ) { case MPI_MAX: if( x[i] > o[i] ) o[i] = x[i]; break; case MPI_MIN: if( x[i] < o[i] ) o[i] = x[i]; break; \
                       case MPI_SUM: o[i] += x[i]; break; case MPI_PROD: o[i] *= x[i]; break;                              \
                       case MPI_LAND: o[i] = (T)(o[i] && x[i]); break; case MPI_LOR: o[i] = (T)(o[i] || x[i]); break;    \
                       default: return 0; } } return 1; } while(0)
#define REDUCE_BITS(T) do { T *o = (T*)out; const T *x = (const T*)in; for( int i = 0; i < count; i++ ) {                \
        if( MPI_BAND == op ) o[i] &= x[i]; else o[i] |= x[i]; } return 1; } while(0)

/* out = out (op) in, element by element; 0 when the pair (type, op) is not served */
static int reduce_into(void *out, const void *in, int count, MPI_Datatype type, MPI_Op op)
{
    if( MPI_MAXLOC == op || MPI_MINLOC == op ) {
        if( MPI_2INT == type ) {
            int *o = (int*)out; const int *x = (const int*)in;
            for( int i = 0; i < count; i++ )
                if( (MPI_MAXLOC == op ? x[2*i] > o[2*i] : x[2*i] < o[2*i]) || (x[2*i] == o[2*i] && x[2*i+1] < o[2*i+1]) ) { o[2*i] = x[2*i]; o[2*i+1] = x[2*i+1]; }
            return 1;
        }
        if( MPI_DOUBLE_INT == type ) {
            struct di { double v; int i; };
            struct di *o = (struct di*)out; const struct di *x = (const struct di*)in;
            for( int i = 0; i < count; i++ )
                if( (MPI_MAXLOC == op ? x[i].v > o[i].v : x[i].v < o[i].v) || (x[i].v == o[i].v && x[i].i < o[i].i) ) o[i] = x[i];
            return 1;
        }
        return 0;
    }
    if( MPI_BAND == op || MPI_BOR == op ) {
        switch( type ) {
        case MPI_INT: case MPI_INT32_T: case MPI_UNSIGNED: case MPI_UINT32_T: REDUCE_BITS(uint32_t);
        case MPI_LONG: case MPI_LONG_LONG: case MPI_INT64_T: case MPI_UNSIGNED_LONG: case MPI_UNSIGNED_LONG_LONG: case MPI_UINT64_T: REDUCE_BITS(uint64_t);
        case MPI_CHAR: case MPI_BYTE: case MPI_UNSIGNED_CHAR: case MPI_UINT8_T: case MPI_INT8_T: REDUCE_BITS(uint8_t);
        default: return 0;
        }
    }
    switch( type ) {
    case MPI_INT: case MPI_INT32_T: REDUCE_LOOP(int32_t);
    case MPI_UNSIGNED: case MPI_UINT32_T: REDUCE_LOOP(uint32_t);
    case MPI_LONG: case MPI_LONG_LONG: case MPI_INT64_T: case MPI_AINT: REDUCE_LOOP(int64_t);
    case MPI_UNSIGNED_LONG: case MPI_UNSIGNED_LONG_LONG: case MPI_UINT64_T: REDUCE_LOOP(uint64_t);
    case MPI_SHORT: case MPI_INT16_T: REDUCE_LOOP(int16_t);
    case MPI_CHAR: case MPI_SIGNED_CHAR: case MPI_INT8_T: REDUCE_LOOP(int8_t);
    case MPI_UNSIGNED_CHAR: case MPI_BYTE: case MPI_UINT8_T: REDUCE_LOOP(uint8_t);
    case MPI_FLOAT: REDUCE_LOOP(float);
    case MPI_DOUBLE: REDUCE_LOOP(double);
    default: return 0;
    }
}

/* every rank gathers every contribution and reduces them in rank order: all ranks get the same bits */
int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm)
{
    int size;
    const void *in = (MPI_IN_PLACE == sendbuf) ? recvbuf : sendbuf;
    if( MPI_SUCCESS != MPI_Type_size(type, &size) || count < 0 ) return MPI_ERR_ARG;
    if( comm_is_self(comm) || NULL == transport ) {
        if( in != recvbuf ) memcpy(recvbuf, in, (size_t)count * size);
        return MPI_SUCCESS;
    }
    const size_t bytes = (size_t)count * size;
    if( 0 == bytes ) return MPI_Barrier(comm);
    char *all = (char*)malloc((size_t)world_size * bytes);
    int rc = allgather_bytes(in, bytes, all);
    if( MPI_SUCCESS == rc ) {
        memcpy(recvbuf, all, bytes);
        for( int r = 1; r < world_size && MPI_SUCCESS == rc; r++ )
            if( !reduce_into(recvbuf, all + (size_t)r * bytes, count, type, op) ) rc = MPI_ERR_ARG;
    }
    free(all);
    return rc;
}

int MPI_Allgather(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf, int recvcount,
                  MPI_Datatype recvtype, MPI_Comm comm)
{
    int size, rank;
    (void)recvcount; (void)recvtype;
    if( MPI_SUCCESS != MPI_Type_size(sendtype, &size) || sendcount < 0 ) return MPI_ERR_ARG;
    MPI_Comm_rank(comm, &rank);
    const size_t bytes = (size_t)sendcount * size;
    const void *in = (MPI_IN_PLACE == sendbuf) ? (const char*)recvbuf + (size_t)rank * bytes : sendbuf;
    if( comm_is_self(comm) || NULL == transport ) { if( in != recvbuf ) memcpy(recvbuf, in, bytes); return MPI_SUCCESS; }
    if( 0 == bytes ) return MPI_Barrier(comm);
    char *mine = (char*)malloc(bytes);
    memcpy(mine, in, bytes);
    int rc = allgather_bytes(mine, bytes, recvbuf);
    free(mine);
    return rc;
}

int MPI_Reduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm comm)
{
    int size, rank;
    if( MPI_SUCCESS != MPI_Type_size(type, &size) || count < 0 ) return MPI_ERR_ARG;
    MPI_Comm_rank(comm, &rank);
    void *tmp = malloc((size_t)count * size + 1);
    memcpy(tmp, (MPI_IN_PLACE == sendbuf) ? recvbuf : sendbuf, (size_t)count * size);
    int rc = MPI_Allreduce(MPI_IN_PLACE, tmp, count, type, op, comm);
    if( MPI_SUCCESS == rc && rank == root ) memcpy(recvbuf, tmp, (size_t)count * size);
    free(tmp);
    return rc;
}

int MPI_Bcast(void *buf, int count, MPI_Datatype type, int root, MPI_Comm comm)
{
    int size;
    if( MPI_SUCCESS != MPI_Type_size(type, &size) || count < 0 ) return MPI_ERR_ARG;
    if( comm_is_self(comm) || NULL == transport ) return MPI_SUCCESS;
    char *all = (char*)malloc((size_t)world_size * 128);
    for( size_t done = 0; done < (size_t)count * size; done += 128 ) {
        size_t n = (size_t)count * size - done < 128 ? (size_t)count * size - done : 128;
        char mine[128] = {0};
        if( my_rank == root ) memcpy(mine, (char*)buf + done, n);
        if( PB2_SUCCESS != pb2_ce_allgather(transport, mine, 128, all) ) { free(all); return MPI_ERR_OTHER; }
        memcpy((char*)buf + done, all + (size_t)root * 128, n);
    }
    free(all);
    return MPI_SUCCESS;
}

int MPI_Gather(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf, int recvcount,
               MPI_Datatype recvtype, int root, MPI_Comm comm)
{
    int size, rank, nranks;
    (void)recvcount; (void)recvtype;
    if( MPI_SUCCESS != MPI_Type_size(sendtype, &size) || sendcount < 0 ) return MPI_ERR_ARG;
    MPI_Comm_rank(comm, &rank);
    MPI_Comm_size(comm, &nranks);
    const size_t bytes = (size_t)sendcount * size;
    if( 1 == nranks || NULL == transport ) { memcpy(recvbuf, sendbuf, bytes); return MPI_SUCCESS; }
    if( bytes > 128 ) return MPI_ERR_ARG;   /* small contributions only */
    char *all = (char*)malloc((size_t)nranks * 128);
    char mine[128] = {0};
    memcpy(mine, sendbuf, bytes);
    int rc = pb2_ce_allgather(transport, mine, 128, all);
    if( PB2_SUCCESS == rc && rank == root )
        for( int r = 0; r < nranks; r++ ) memcpy((char*)recvbuf + (size_t)r * bytes, all + (size_t)r * 128, bytes);
    free(all);
    return PB2_SUCCESS == rc ? MPI_SUCCESS : MPI_ERR_OTHER;
}

int MPI_Gatherv(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf, const int *recvcounts,
                const int *displs, MPI_Datatype recvtype, int root, MPI_Comm comm)
{
    int size, rank, nranks;
    (void)recvtype;
    if( MPI_SUCCESS != MPI_Type_size(sendtype, &size) || sendcount < 0 ) return MPI_ERR_ARG;
    MPI_Comm_rank(comm, &rank);
    MPI_Comm_size(comm, &nranks);
    if( 1 == nranks || NULL == transport ) { memcpy((char*)recvbuf + (size_t)displs[0] * size, sendbuf, (size_t)sendcount * size); return MPI_SUCCESS; }
    /* rounds of 128 bytes per rank until the longest contribution is through; every rank knows its own length only,
     * so the number of rounds is agreed first */
    int64_t longest = (int64_t)sendcount * size;
    if( PB2_SUCCESS != pb2_ce_allreduce_i64(transport, &longest, 1, 0) ) return MPI_ERR_OTHER;
    char *all = (char*)malloc((size_t)nranks * 128);
    for( int64_t done = 0; done < longest; done += 128 ) {
        char mine[128] = {0};
        int64_t left = (int64_t)sendcount * size - done;
        if( left > 0 ) memcpy(mine, (const char*)sendbuf + done, (size_t)(left < 128 ? left : 128));
        if( PB2_SUCCESS != pb2_ce_allgather(transport, mine, 128, all) ) { free(all); return MPI_ERR_OTHER; }
        if( rank == root )
            for( int r = 0; r < nranks; r++ ) {
                int64_t rl = (int64_t)recvcounts[r] * size - done;
                if( rl > 0 ) memcpy((char*)recvbuf + (size_t)displs[r] * size + done, all + (size_t)r * 128, (size_t)(rl < 128 ? rl : 128));
            }
    }
    free(all);
    return MPI_SUCCESS;
}
