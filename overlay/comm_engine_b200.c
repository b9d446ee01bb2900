/*
 * comm_engine_b200.c -- the comm-engine seam (piece 3 / row a22) for the ranks of one box.
 *
 * Compiled INSIDE a copy of the reference tree in place of parsec/parsec_comm_engine.c and
 * parsec/parsec_mpi_funnelled.c (overlay/build_overlay.sh, the `dist` tree).  It defines the global `parsec_ce` and
 * parsec_comm_engine_init / _fini (parsec/parsec_comm_engine.h:161-182); everything above it -- remote_dep.c,
 * remote_dep_mpi.c with its ACTIVATE / GET_DATA / put protocol, the termination detectors -- is the reference's own
 * code, unchanged, and drives this file through the vtable exactly as it drives the MPI engine.
 *
 * Below it: libparsec_b200's pb2_ce_* transport (include/parsec_b200.h; one shared-memory ring per pair of ranks,
 * one-copy put into the peer's address space) and libpb2mpi's datatype layouts (overlay/pb2mpi/mpi.h).
 *
 *  - an active message is one ring record; a full ring parks the message in a local FIFO that progress() drains;
 *  - a registration (mem_register) is plain data that can travel inside messages, as the reference's callers
 *    require (remote_dep_mpi.c:2053-2065 copies it into the GET_DATA message, tests/dsl/dtd/dtd_test_ce.c:294-310
 *    forwards it twice): owner, address, datatype, count, and the layout a peer needs to address the memory
 *    directly -- `nblocks` pieces of `blocklen` bytes `stride` apart, which covers whole tiles and LAPACK-storage
 *    tiles; any other layout is served through a packed bounce buffer on the owner;
 *  - put: the origin gathers its own pieces and the peer's into one vectored copy, then sends the completion record
 *    that makes the target run the callback the target itself named (`r_tag` is the target's function pointer,
 *    remote_dep_mpi.c:1987-1991).  Never executed inside an AM callback: queued, run by progress();
 *  - get: the mirror image with process_vm_readv where the box allows it, request / reply through put otherwise.
 */
#include "parsec/parsec_config.h"
#include "parsec/parsec_comm_engine.h"
#include "parsec/parsec_internal.h"
#include "parsec/execution_stream.h"
#include "parsec/data_internal.h"
#include "parsec/utils/debug.h"

#include <mpi.h>
#include <pthread.h>
#include <sched.h>
#include <stdlib.h>
#include <stdio.h>
#include <time.h>
#include <string.h>

#include "parsec_b200.h"

parsec_comm_engine_t parsec_ce;

/* ---- registrations ---- */
typedef struct b200_reg_s {
    struct b200_reg_s *self;   /* the registration itself, on the rank that owns it (messages carry copies) */
    int32_t  owner;
    int32_t  count;
    uint64_t mem;              /* what was registered: address, datatype, count (mem_retrieve) */
    uint64_t bytes;            /* bytes it names */
    int32_t  datatype;
    int32_t  nblocks;          /* how a peer addresses it: nblocks pieces of blocklen bytes, stride bytes apart, */
    uint64_t target;           /* from this address in the owner's address space */
    int64_t  blocklen, stride;
    /* ---- owner only, not part of what travels ---- */
    void    *bounce;           /* packed image when the layout has no such description */
} b200_reg_t;
#define B200_REG_PUBLIC (offsetof(b200_reg_t, bounce))

/* ---- internal tags (64-bit, far from the reference's small integers) ---- */
#define B200_TAG_PUT_DONE  0x7062320000000001ull   /* {r_tag, target registration, cb bytes} after the data */
#define B200_TAG_GET_REQ   0x7062320000000002ull   /* no one-copy read on this box: please put it to me */
#define B200_TAG_GET_REPLY 0x7062320000000003ull
#define B200_MAX_TAGS      64
#define B200_MAX_CB_DATA   256

typedef struct {
    uint64_t r_tag;
    uint64_t target_self;      /* the target's b200_reg_t, for its bounce buffer */
    uint32_t cb_size;
    uint32_t pad;
    char     cb_data[B200_MAX_CB_DATA];
} b200_done_msg_t;

typedef struct {
    uint64_t op;               /* the origin's pending get */
    uint64_t r_tag;
    uint32_t cb_size, pad;
    char     origin_reg[B200_REG_PUBLIC];
    char     remote_reg[B200_REG_PUBLIC];
    char     cb_data[B200_MAX_CB_DATA];
} b200_get_req_t;

typedef struct b200_op_s {
    struct b200_op_s *next;
    int kind;                  /* 0 put, 1 get, 2 put on behalf of a peer's get */
    int remote;
    b200_reg_t *lreg;          /* ours (the real one) */
    char rreg[B200_REG_PUBLIC];/* copy of the peer's */
    ptrdiff_t ldispl, rdispl;
    parsec_ce_onesided_callback_t l_cb;
    void *l_cb_data;
    uint64_t r_tag;
    uint32_t cb_size;
    char cb_data[B200_MAX_CB_DATA];
    uint64_t peer_op;          /* kind 2: which get of the peer this answers */
    int status;                /* of the transfer, when a copier thread made it */
} b200_op_t;

typedef struct b200_parked_s {
    struct b200_parked_s *next;
    int dst;
    uint64_t tag;
    size_t len;
    char data[];
} b200_parked_t;

static struct {
    parsec_ce_am_callback_t cb;
    void *cb_data;
    size_t max_len;
} b200_tags[B200_MAX_TAGS];

static pb2_ce_t *b200_tp = NULL;
static int b200_rank = 0, b200_world = 1, b200_up = 0;
static b200_op_t *b200_ops_head = NULL, *b200_ops_tail = NULL;
static int b200_nb_ops = 0, b200_waiting_gets = 0;
static b200_parked_t *b200_parked_head = NULL, *b200_parked_tail = NULL;
static pthread_mutex_t b200_parked_lock = PTHREAD_MUTEX_INITIALIZER;
static int b200_events = 0;    /* counted by the dispatch below, reported by progress() */

/* PB2_CE_STATS=1: where the engine's time goes, printed by every rank when the engine shuts down */
static struct { uint64_t puts, put_bytes, put_ns, regs, ams, am_bytes, progress_calls, cb_ns; } b200_st;
static int b200_stats_on = 0;
static inline uint64_t b200_now_ns(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec;
}

/* ---- layouts ---- */
/* pieces of `count` elements of `datatype` at `base` (+displ); returns the number of pieces, -1 when out of memory */
static int b200_local_iov(const b200_reg_t *reg, ptrdiff_t displ, pb2_ce_iov_t **out)
{
    pb2_ce_iov_t *v;
    if( NULL != reg->bounce || reg->datatype == MPI_DATATYPE_NULL ) {   /* packed image, or plain bytes */
        v = (pb2_ce_iov_t*)malloc(sizeof(pb2_ce_iov_t));
        if( NULL == v ) return -1;
        v[0].base = (NULL != reg->bounce ? (char*)reg->bounce : (char*)(uintptr_t)reg->mem) + displ;
        v[0].len = reg->bytes;
        *out = v;
        return 1;
    }
    const pb2mpi_segment_t *seg;
    int nseg;
    int64_t extent, size;
    if( MPI_SUCCESS != PB2MPI_Type_layout(reg->datatype, &seg, &nseg, &extent, &size) ) return -1;
    v = (pb2_ce_iov_t*)malloc(sizeof(pb2_ce_iov_t) * (size_t)(reg->count * nseg + 1));
    if( NULL == v ) return -1;
    int n = 0;
    for( int i = 0; i < reg->count; i++ )
        for( int s = 0; s < nseg; s++ ) {
            char *p = (char*)(uintptr_t)reg->mem + displ + i * extent + seg[s].disp;
            if( n && (char*)v[n - 1].base + v[n - 1].len == p ) { v[n - 1].len += (size_t)seg[s].len; continue; }
            v[n].base = p;
            v[n].len = (size_t)seg[s].len;
            n++;
        }
    *out = v;
    return n;
}

/* pieces of a peer's registration, from the plain data that travelled */
static int b200_remote_iov(const b200_reg_t *pub, ptrdiff_t displ, pb2_ce_iov_t **out)
{
    pb2_ce_iov_t *v = (pb2_ce_iov_t*)malloc(sizeof(pb2_ce_iov_t) * (size_t)(pub->nblocks > 0 ? pub->nblocks : 1));
    if( NULL == v ) return -1;
    for( int i = 0; i < pub->nblocks; i++ ) {
        v[i].base = (char*)(uintptr_t)pub->target + displ + i * pub->stride;
        v[i].len = (size_t)pub->blocklen;
    }
    *out = v;
    return pub->nblocks;
}

/* ---- vtable: tags and active messages ---- */
static int b200_tag_register(parsec_ce_tag_t tag, parsec_ce_am_callback_t cb, void *cb_data, size_t msg_length)
{
    if( tag >= B200_MAX_TAGS || NULL == cb ) return PARSEC_ERR_BAD_PARAM;
    if( NULL != b200_tags[tag].cb ) return PARSEC_ERR_EXISTS;
    if( NULL != b200_tp && msg_length > pb2_ce_max_message(b200_tp) ) {
        parsec_warning("comm engine: tag %llu wants messages of up to %zu bytes, the rings of this job carry %zu "
                       "(PB2_CE_RING_BYTES is a quarter of that short)", (unsigned long long)tag, msg_length,
                       pb2_ce_max_message(b200_tp));
        return PARSEC_ERR_VALUE_OUT_OF_BOUNDS;
    }
    b200_tags[tag].cb = cb;
    b200_tags[tag].cb_data = cb_data;
    b200_tags[tag].max_len = msg_length;
    return PARSEC_SUCCESS;
}

static int b200_tag_unregister(parsec_ce_tag_t tag)
{
    if( tag >= B200_MAX_TAGS || NULL == b200_tags[tag].cb ) return PARSEC_ERR_NOT_FOUND;
    b200_tags[tag].cb = NULL;
    return PARSEC_SUCCESS;
}

/* a message that found the ring full waits here; order per destination is kept because everything behind it for the
 * same destination is parked too */
static int b200_send_or_park(int dst, uint64_t tag, const void *a, size_t alen)
{
    int rc = PB2_HOOK_RETURN_AGAIN;
    pthread_mutex_lock(&b200_parked_lock);
    int behind = 0;
    for( b200_parked_t *p = b200_parked_head; NULL != p; p = p->next ) if( p->dst == dst ) { behind = 1; break; }
    if( !behind ) rc = pb2_ce_send(b200_tp, dst, tag, a, alen);
    if( PB2_HOOK_RETURN_AGAIN == rc ) {
        b200_parked_t *p = (b200_parked_t*)malloc(sizeof(b200_parked_t) + alen);
        p->next = NULL;
        p->dst = dst;
        p->tag = tag;
        p->len = alen;
        if( alen ) memcpy(p->data, a, alen);
        if( NULL == b200_parked_tail ) b200_parked_head = p; else b200_parked_tail->next = p;
        b200_parked_tail = p;
        rc = PB2_SUCCESS;
    }
    pthread_mutex_unlock(&b200_parked_lock);
    return PB2_SUCCESS == rc ? PARSEC_SUCCESS : PARSEC_ERROR;
}

static int b200_flush_parked(void)
{
    int sent = 0;
    if( NULL == b200_parked_head ) return 0;
    pthread_mutex_lock(&b200_parked_lock);
    uint64_t stuck = 0;   /* destinations whose ring is still full: later messages to them stay behind */
    b200_parked_t **link = &b200_parked_head, *p, *last = NULL;
    while( NULL != (p = *link) ) {
        if( !(stuck & (1ull << (p->dst & 63))) ) {
            int rc = pb2_ce_send(b200_tp, p->dst, p->tag, p->data, p->len);
            if( PB2_SUCCESS == rc ) { *link = p->next; free(p); sent++; continue; }
            stuck |= 1ull << (p->dst & 63);
        }
        last = p;
        link = &p->next;
    }
    b200_parked_tail = last;
    pthread_mutex_unlock(&b200_parked_lock);
    return sent;
}

static int b200_send_am(parsec_comm_engine_t *ce, parsec_ce_tag_t tag, int remote, void *addr, size_t size)
{
    (void)ce;
    if( NULL == b200_tp || remote < 0 || remote >= b200_world ) return PARSEC_ERR_BAD_PARAM;
    return b200_send_or_park(remote, tag, addr, size);
}

/* ---- vtable: memory registration ---- */
static int b200_mem_register(void *mem, parsec_mem_type_t mem_type, size_t count, parsec_datatype_t datatype,
                             size_t mem_size, parsec_ce_mem_reg_handle_t *lreg, size_t *lreg_size)
{
    b200_reg_t *r = (b200_reg_t*)calloc(1, sizeof(b200_reg_t));
    if( NULL == r ) return PARSEC_ERR_OUT_OF_RESOURCE;
    r->self = r;
    r->owner = b200_rank;
    r->mem = (uint64_t)(uintptr_t)mem;
    r->target = r->mem;
    if( PARSEC_MEM_TYPE_CONTIGUOUS == mem_type ) {
        r->count = 1;
        r->datatype = MPI_DATATYPE_NULL;
        r->bytes = mem_size;
        r->nblocks = 1;
        r->blocklen = (int64_t)mem_size;
    } else {
        int64_t size;
        if( MPI_SUCCESS != PB2MPI_Type_layout(datatype, NULL, NULL, NULL, &size) ) { free(r); return PARSEC_ERR_BAD_PARAM; }
        r->count = (int32_t)count;
        r->datatype = datatype;
        r->bytes = (uint64_t)(size * (int64_t)count);
        /* regular?  look at the merged pieces of all `count` elements */
        pb2_ce_iov_t *v = NULL;
        int nb = b200_local_iov(r, 0, &v), regular = (nb > 0);
        int64_t first = 0, len = 0, stride = 0;
        if( nb < 0 ) { free(r); return PARSEC_ERR_OUT_OF_RESOURCE; }
        if( nb > 0 ) {
            first = (char*)v[0].base - (char*)mem;
            len = (int64_t)v[0].len;
            stride = nb > 1 ? (char*)v[1].base - (char*)v[0].base : 0;
            for( int i = 1; i < nb && regular; i++ )
                if( (int64_t)v[i].len != len || (char*)v[i].base - (char*)v[i - 1].base != stride ) regular = 0;
        }
        free(v);
        if( regular && nb > 0 ) {
            r->target = r->mem + (uint64_t)first;
            r->nblocks = nb;
            r->blocklen = len;
            r->stride = stride;
        } else if( r->bytes > 0 ) {   /* no short description: peers see a packed image */
            int pos = 0;
            r->bounce = malloc(r->bytes);
            if( NULL == r->bounce ) { free(r); return PARSEC_ERR_OUT_OF_RESOURCE; }
            MPI_Pack(mem, (int)count, datatype, r->bounce, (int)r->bytes, &pos, MPI_COMM_WORLD);
            r->target = (uint64_t)(uintptr_t)r->bounce;
            r->nblocks = 1;
            r->blocklen = (int64_t)r->bytes;
        }
    }
    b200_st.regs++;
    *lreg = (parsec_ce_mem_reg_handle_t)r;
    if( NULL != lreg_size ) *lreg_size = B200_REG_PUBLIC;
    return PARSEC_SUCCESS;
}

static int b200_mem_unregister(parsec_ce_mem_reg_handle_t *lreg)
{   /* *lreg may be a copy inside a message (dtd_test_ce.c:217-240): the registration is what `self` names */
    b200_reg_t *r = ((b200_reg_t*)*lreg)->self;
    if( NULL == r || r->owner != b200_rank ) return PARSEC_ERR_BAD_PARAM;
    free(r->bounce);
    free(r);
    *lreg = NULL;
    return PARSEC_SUCCESS;
}

static int b200_get_mem_handle_size(void) { return (int)B200_REG_PUBLIC; }

static int b200_mem_retrieve(parsec_ce_mem_reg_handle_t lreg, void **mem, parsec_datatype_t *datatype, int *count)
{
    const b200_reg_t *r = (const b200_reg_t*)lreg;
    *mem = (void*)(uintptr_t)r->mem;
    *datatype = r->datatype;
    *count = r->count;
    return PARSEC_SUCCESS;
}

/* ---- vtable: one-sided ---- */
static int b200_queue_op(int kind, parsec_ce_mem_reg_handle_t lreg, ptrdiff_t ldispl, parsec_ce_mem_reg_handle_t rreg,
                         ptrdiff_t rdispl, int remote, parsec_ce_onesided_callback_t l_cb, void *l_cb_data,
                         parsec_ce_tag_t r_tag, void *r_cb_data, size_t r_cb_data_size, uint64_t peer_op)
{
    if( NULL == b200_tp || remote < 0 || remote >= b200_world || r_cb_data_size > B200_MAX_CB_DATA ) return PARSEC_ERR_BAD_PARAM;
    b200_op_t *op = (b200_op_t*)calloc(1, sizeof(b200_op_t));
    if( NULL == op ) return PARSEC_ERR_OUT_OF_RESOURCE;
    op->kind = kind;
    op->remote = remote;
    op->lreg = ((b200_reg_t*)lreg)->self;          /* lreg itself may live in a message that is gone by then */
    memcpy(op->rreg, rreg, B200_REG_PUBLIC);
    op->ldispl = ldispl;
    op->rdispl = rdispl;
    op->l_cb = l_cb;
    op->l_cb_data = l_cb_data;
    op->r_tag = r_tag;
    op->cb_size = (uint32_t)r_cb_data_size;
    if( r_cb_data_size ) memcpy(op->cb_data, r_cb_data, r_cb_data_size);
    op->peer_op = peer_op;
    if( NULL == b200_ops_tail ) b200_ops_head = op; else b200_ops_tail->next = op;
    b200_ops_tail = op;
    b200_nb_ops++;
    return PARSEC_SUCCESS;
}

static int b200_put(parsec_comm_engine_t *ce, parsec_ce_mem_reg_handle_t lreg, ptrdiff_t ldispl,
                    parsec_ce_mem_reg_handle_t rreg, ptrdiff_t rdispl, size_t size, int remote,
                    parsec_ce_onesided_callback_t l_cb, void *l_cb_data,
                    parsec_ce_tag_t r_tag, void *r_cb_data, size_t r_cb_data_size)
{
    (void)ce; (void)size;   /* the registrations say how much (every caller passes 0) */
    return b200_queue_op(0, lreg, ldispl, rreg, rdispl, remote, l_cb, l_cb_data, r_tag, r_cb_data, r_cb_data_size, 0);
}

static int b200_get(parsec_comm_engine_t *ce, parsec_ce_mem_reg_handle_t lreg, ptrdiff_t ldispl,
                    parsec_ce_mem_reg_handle_t rreg, ptrdiff_t rdispl, size_t size, int remote,
                    parsec_ce_onesided_callback_t l_cb, void *l_cb_data,
                    parsec_ce_tag_t r_tag, void *r_cb_data, size_t r_cb_data_size)
{
    (void)ce; (void)size;
    return b200_queue_op(1, lreg, ldispl, rreg, rdispl, remote, l_cb, l_cb_data, r_tag, r_cb_data, r_cb_data_size, 0);
}

/* the target's half of a completed put (or the remote's half of a completed get): run the callback it named */
static void b200_serve_done(const b200_done_msg_t *m, int src)
{
    b200_reg_t *mine = (b200_reg_t*)(uintptr_t)m->target_self;
    if( NULL != mine && NULL != mine->bounce ) {   /* the bytes arrived packed: lay them out */
        int pos = 0;
        MPI_Unpack(mine->bounce, (int)mine->bytes, &pos, (void*)(uintptr_t)mine->mem, mine->count, mine->datatype, MPI_COMM_WORLD);
    }
    if( 0 != m->r_tag ) {
        parsec_ce_am_callback_t fn = (m->r_tag < B200_MAX_TAGS) ? b200_tags[m->r_tag].cb
                                                                 : (parsec_ce_am_callback_t)(uintptr_t)m->r_tag;
        void *cb_data = (m->r_tag < B200_MAX_TAGS) ? b200_tags[m->r_tag].cb_data : NULL;
        char copy[B200_MAX_CB_DATA];
        memcpy(copy, m->cb_data, m->cb_size);
        if( NULL != fn ) fn(&parsec_ce, m->r_tag, copy, m->cb_size, src, cb_data);
    }
}

static void b200_finish_local(b200_op_t *op, int unpack)
{
    if( unpack && NULL != op->lreg->bounce ) {
        int pos = 0;
        MPI_Unpack(op->lreg->bounce, (int)op->lreg->bytes, &pos, (void*)(uintptr_t)op->lreg->mem, op->lreg->count,
                   op->lreg->datatype, MPI_COMM_WORLD);
    }
    if( NULL != op->l_cb )
        op->l_cb(&parsec_ce, (parsec_ce_mem_reg_handle_t)op->lreg, op->ldispl, (parsec_ce_mem_reg_handle_t)op->rreg,
                 op->rdispl, (size_t)op->lreg->bytes, op->remote, op->l_cb_data);
}

static void b200_send_done(int dst, uint64_t r_tag, uint64_t target_self, const void *cb, uint32_t cb_size)
{
    b200_done_msg_t m;
    m.r_tag = r_tag;
    m.target_self = target_self;
    m.cb_size = cb_size;
    m.pad = 0;
    if( cb_size ) memcpy(m.cb_data, cb, cb_size);
    b200_send_or_park(dst, B200_TAG_PUT_DONE, &m, offsetof(b200_done_msg_t, cb_data) + cb_size);
}

static int b200_copy_now(b200_op_t *op);
static int b200_offload_put(b200_op_t *op);
static int b200_first_put_done;

static void b200_run_op(b200_op_t *op)
{
    const b200_reg_t *peer = (const b200_reg_t*)op->rreg;
    int rc;
    switch( op->kind ) {
    case 0:   /* put: ours -> theirs, then their callback, then ours */
    case 2:   /* the same on behalf of the peer's get: the reply finishes it over there */
        rc = b200_copy_now(op);
        b200_first_put_done = 1;
        if( PB2_SUCCESS != rc )
            parsec_fatal("comm engine: put of %llu bytes to rank %d failed (%d)", (unsigned long long)op->lreg->bytes, op->remote, rc);
        if( 0 == op->kind ) {
            b200_send_done(op->remote, op->r_tag, (uint64_t)(uintptr_t)peer->self, op->cb_data, op->cb_size);
            b200_finish_local(op, 0);
        } else {
            b200_send_or_park(op->remote, B200_TAG_GET_REPLY, &op->peer_op, sizeof(op->peer_op));
            b200_done_msg_t m;   /* and this side's notification, locally */
            m.r_tag = op->r_tag;
            m.target_self = 0;
            m.cb_size = op->cb_size;
            memcpy(m.cb_data, op->cb_data, op->cb_size);
            b200_serve_done(&m, op->remote);
        }
        free(op);
        break;
    }
}

/* ---- copier threads ----
 * A put of a tile is a copy of megabytes by whoever executes it; done by the communication thread it stops that
 * thread from answering anybody for as long (the MPI engine hands the same work to MPI_Isend and goes on).  Large
 * puts are therefore handed to a few copier threads (PB2_CE_COPY_THREADS = 0 .. 8, default 2; on the B200 box's host
 * two of them take a two-rank GEMM from 0.37 to 0.20 s per run, profiles/comm_shm_r2.md): a copier builds the
 * piece lists, makes the one-copy transfer, sends the completion record to the target, and gives the operation back;
 * the origin's callback still runs on the communication thread, from progress().  Only on the one-copy path: puts
 * staged through the rings need this rank's polling, which belongs to the communication thread. */
#define B200_COPY_MIN_BYTES (256u * 1024u)
static pthread_mutex_t b200_copy_lock = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t b200_copy_cond = PTHREAD_COND_INITIALIZER;
static b200_op_t *b200_copy_head = NULL, *b200_copy_tail = NULL;      /* waiting for a copier */
static b200_op_t *b200_copied_head = NULL, *b200_copied_tail = NULL;  /* copied, callback not run yet */
static pthread_t b200_copiers[8];
static int b200_nb_copiers = 0, b200_copiers_wanted = -1, b200_copy_inflight = 0, b200_copy_stop = 0, b200_first_put_done = 0;

static int b200_copy_now(b200_op_t *op)   /* the transfer itself; any thread */
{
    const b200_reg_t *peer = (const b200_reg_t*)op->rreg;
    pb2_ce_iov_t *lv = NULL, *rv = NULL;
    if( NULL != op->lreg->bounce ) {   /* refresh the packed image of what we send */
        int pos = 0;
        MPI_Pack((void*)(uintptr_t)op->lreg->mem, op->lreg->count, op->lreg->datatype, op->lreg->bounce,
                 (int)op->lreg->bytes, &pos, MPI_COMM_WORLD);
    }
    int nl = b200_local_iov(op->lreg, op->ldispl, &lv), nr = b200_remote_iov(peer, op->rdispl, &rv);
    const uint64_t t0 = b200_stats_on ? b200_now_ns() : 0;
    int rc = (nl < 0 || nr < 0) ? PB2_ERR_OUT_OF_RESOURCE : pb2_ce_put(b200_tp, op->remote, lv, nl, rv, nr);
    if( b200_stats_on ) {
        __atomic_fetch_add(&b200_st.put_ns, b200_now_ns() - t0, __ATOMIC_RELAXED);
        __atomic_fetch_add(&b200_st.puts, 1, __ATOMIC_RELAXED);
        __atomic_fetch_add(&b200_st.put_bytes, op->lreg->bytes, __ATOMIC_RELAXED);
    }
    free(lv);
    free(rv);
    return rc;
}

static void *b200_copier_main(void *arg)
{
    (void)arg;
    for( ;; ) {
        pthread_mutex_lock(&b200_copy_lock);
        while( NULL == b200_copy_head && !b200_copy_stop ) pthread_cond_wait(&b200_copy_cond, &b200_copy_lock);
        b200_op_t *op = b200_copy_head;
        if( NULL == op ) { pthread_mutex_unlock(&b200_copy_lock); return NULL; }
        b200_copy_head = op->next;
        if( NULL == b200_copy_head ) b200_copy_tail = NULL;
        pthread_mutex_unlock(&b200_copy_lock);
        op->next = NULL;
        op->status = b200_copy_now(op);
        if( PB2_SUCCESS == op->status )   /* the target may run its callback from now on; ours waits for progress() */
            b200_send_done(op->remote, op->r_tag, (uint64_t)(uintptr_t)((const b200_reg_t*)op->rreg)->self, op->cb_data, op->cb_size);
        pthread_mutex_lock(&b200_copy_lock);
        if( NULL == b200_copied_tail ) b200_copied_head = op; else b200_copied_tail->next = op;
        b200_copied_tail = op;
        pthread_mutex_unlock(&b200_copy_lock);
    }
}

/* communication thread: hand a put to the copiers; 0 when it has to be done here */
static int b200_offload_put(b200_op_t *op)
{
    if( b200_copiers_wanted < 0 ) {
        const char *e = getenv("PB2_CE_COPY_THREADS");
        b200_copiers_wanted = (NULL != e) ? atoi(e) : 2;
        if( b200_copiers_wanted > 8 ) b200_copiers_wanted = 8;
    }
    if( b200_copiers_wanted <= 0 || 0 != op->kind || op->lreg->bytes < B200_COPY_MIN_BYTES || op->remote == b200_rank ) return 0;
    if( !b200_first_put_done ) return 0;   /* the first put, done here, finds out whether this box allows one-copy puts */
    pb2_ce_stats_t st;
    pb2_ce_stats(b200_tp, &st);
    if( !st.single_copy ) return 0;
    while( b200_nb_copiers < b200_copiers_wanted ) {
        if( 0 != pthread_create(&b200_copiers[b200_nb_copiers], NULL, b200_copier_main, NULL) ) { b200_copiers_wanted = b200_nb_copiers; break; }
        b200_nb_copiers++;
    }
    if( 0 == b200_nb_copiers ) return 0;
    pthread_mutex_lock(&b200_copy_lock);
    op->next = NULL;
    if( NULL == b200_copy_tail ) b200_copy_head = op; else b200_copy_tail->next = op;
    b200_copy_tail = op;
    b200_copy_inflight++;
    pthread_cond_signal(&b200_copy_cond);
    pthread_mutex_unlock(&b200_copy_lock);
    return 1;
}

/* communication thread: run the callbacks of the puts the copiers have finished */
static int b200_reap_copied(void)
{
    if( 0 == b200_copy_inflight ) return 0;
    pthread_mutex_lock(&b200_copy_lock);
    b200_op_t *op = b200_copied_head;
    b200_copied_head = b200_copied_tail = NULL;
    pthread_mutex_unlock(&b200_copy_lock);
    int n = 0;
    while( NULL != op ) {
        b200_op_t *next = op->next;
        if( PB2_SUCCESS != op->status )
            parsec_fatal("comm engine: put of %llu bytes to rank %d failed (%d)", (unsigned long long)op->lreg->bytes, op->remote, op->status);
        b200_finish_local(op, 0);
        free(op);
        b200_copy_inflight--;
        n++;
        op = next;
    }
    return n;
}

static void b200_stop_copiers(void)
{
    pthread_mutex_lock(&b200_copy_lock);
    b200_copy_stop = 1;
    pthread_cond_broadcast(&b200_copy_cond);
    pthread_mutex_unlock(&b200_copy_lock);
    for( int i = 0; i < b200_nb_copiers; i++ ) pthread_join(b200_copiers[i], NULL);
    b200_nb_copiers = 0;
    b200_copy_stop = 0;
}

static void b200_dispatch(void *user, int src, uint64_t tag, const void *msg, size_t len)
{
    (void)user;
    b200_events++;
    b200_st.ams++;
    b200_st.am_bytes += len;
    if( tag < B200_MAX_TAGS ) {
        if( NULL == b200_tags[tag].cb ) {
            parsec_warning("comm engine: message with unregistered tag %llu from rank %d dropped", (unsigned long long)tag, src);
            return;
        }
        const uint64_t t0 = b200_stats_on ? b200_now_ns() : 0;
        b200_tags[tag].cb(&parsec_ce, tag, (void*)msg, len, src, b200_tags[tag].cb_data);
        if( b200_stats_on ) b200_st.cb_ns += b200_now_ns() - t0;
        return;
    }
    if( B200_TAG_PUT_DONE == tag ) {
        b200_done_msg_t m;
        memset(&m, 0, sizeof(m));
        memcpy(&m, msg, len < sizeof(m) ? len : sizeof(m));
        b200_serve_done(&m, src);
    } else if( B200_TAG_GET_REQ == tag ) {   /* a peer without one-copy reads wants our memory: put it there */
        const b200_get_req_t *q = (const b200_get_req_t*)msg;
        b200_queue_op(2, (parsec_ce_mem_reg_handle_t)q->remote_reg, 0, (parsec_ce_mem_reg_handle_t)q->origin_reg, 0, src,
                      NULL, NULL, q->r_tag, (void*)q->cb_data, q->cb_size, q->op);
    } else if( B200_TAG_GET_REPLY == tag ) {
        b200_op_t *op;
        memcpy(&op, msg, sizeof(op));
        b200_waiting_gets--;
        b200_finish_local(op, 1);
        free(op);
    }
}

/* Runs the queued one-sided operations.  Outside pb2_ce_poll: a callback may have queued them. */
static int b200_progress_ops(void)
{
    int done = 0;
    b200_op_t *op;
    while( NULL != (op = b200_ops_head) ) {
        b200_ops_head = op->next;
        if( NULL == b200_ops_head ) b200_ops_tail = NULL;
        op->next = NULL;
        b200_nb_ops--;
        done++;
        if( 1 != op->kind ) {
            if( !b200_offload_put(op) ) b200_run_op(op);   /* a large put goes to a copier thread */
            continue;
        }
        /* get */
        const b200_reg_t *peer = (const b200_reg_t*)op->rreg;
        pb2_ce_stats_t st;
        pb2_ce_stats(b200_tp, &st);
        pb2_ce_iov_t *lv = NULL, *rv = NULL;
        int nl = b200_local_iov(op->lreg, op->ldispl, &lv), nr = b200_remote_iov(peer, op->rdispl, &rv);
        int rc = (nl < 0 || nr < 0) ? PB2_ERR_OUT_OF_RESOURCE
                                    : (st.single_copy ? pb2_ce_get(b200_tp, op->remote, lv, nl, rv, nr) : PB2_ERR_NOT_SUPPORTED);
        free(lv);
        free(rv);
        if( PB2_SUCCESS == rc ) {   /* read straight out of the peer: tell it, finish here */
            b200_send_done(op->remote, op->r_tag, 0, op->cb_data, op->cb_size);
            b200_finish_local(op, 1);
            free(op);
        } else if( PB2_ERR_NOT_SUPPORTED == rc ) {   /* ask the peer to put it */
            b200_get_req_t q;
            memset(&q, 0, sizeof(q));
            q.op = (uint64_t)(uintptr_t)op;
            q.r_tag = op->r_tag;
            q.cb_size = op->cb_size;
            memcpy(q.origin_reg, op->lreg, B200_REG_PUBLIC);
            memcpy(q.remote_reg, op->rreg, B200_REG_PUBLIC);
            memcpy(q.cb_data, op->cb_data, op->cb_size);
            b200_send_or_park(op->remote, B200_TAG_GET_REQ, &q, offsetof(b200_get_req_t, cb_data) + op->cb_size);
            b200_waiting_gets++;   /* op is freed when the reply comes */
        } else {
            parsec_fatal("comm engine: get of %llu bytes from rank %d failed (%d)", (unsigned long long)op->lreg->bytes, op->remote, rc);
        }
    }
    return done;
}

static int b200_progress(parsec_comm_engine_t *ce)
{
    (void)ce;
    if( NULL == b200_tp ) return 0;
    b200_st.progress_calls++;
    int n = b200_flush_parked();
    b200_events = 0;
    pb2_ce_poll(b200_tp, b200_dispatch, NULL, 0);
    n += b200_events;
    n += b200_progress_ops();
    n += b200_reap_copied();
    return n;
}

static int b200_enable(parsec_comm_engine_t *ce) { (void)ce; return 1; }
static int b200_disable(parsec_comm_engine_t *ce) { (void)ce; return 1; }

static int b200_pack(parsec_comm_engine_t *ce, void *inbuf, int incount, parsec_datatype_t type, void *outbuf, int outsize, int *position)
{
    (void)ce;
    return MPI_SUCCESS == MPI_Pack(inbuf, incount, type, outbuf, outsize, position, MPI_COMM_WORLD) ? 1 : 0;
}

static int b200_pack_size(parsec_comm_engine_t *ce, int incount, parsec_datatype_t type, int *size)
{
    (void)ce;
    return MPI_SUCCESS == MPI_Pack_size(incount, type, MPI_COMM_WORLD, size) ? 1 : 0;
}

static int b200_unpack(parsec_comm_engine_t *ce, void *inbuf, int insize, int *position, void *outbuf, int outcount, parsec_datatype_t type)
{
    (void)ce;
    return MPI_SUCCESS == MPI_Unpack(inbuf, insize, position, outbuf, outcount, type, MPI_COMM_WORLD) ? 1 : 0;
}

/* a local copy between two layouts; the MPI engine sends to itself for this (parsec_mpi_funnelled.c:384-407) */
static int b200_reshape(parsec_comm_engine_t *ce, parsec_execution_stream_t *es,
                        parsec_data_copy_t *dst, int64_t displ_dst, parsec_datatype_t layout_dst, uint64_t count_dst,
                        parsec_data_copy_t *src, int64_t displ_src, parsec_datatype_t layout_src, uint64_t count_src)
{
    (void)ce; (void)es;
    return MPI_SUCCESS == PB2MPI_Copy((char*)PARSEC_DATA_COPY_GET_PTR(src) + displ_src, (int)count_src, layout_src,
                                      (char*)PARSEC_DATA_COPY_GET_PTR(dst) + displ_dst, (int)count_dst, layout_dst) ? 0 : -1;
}

static int b200_sync(parsec_comm_engine_t *ce)
{
    (void)ce;
    return (NULL == b200_tp || PB2_SUCCESS == pb2_ce_barrier(b200_tp)) ? 0 : -1;
}

/* the upper layer asks before it queues more one-sided work (remote_dep_mpi.c:1521-1530, :1582) */
static int b200_can_serve(parsec_comm_engine_t *ce)
{
    (void)ce;
    return b200_nb_ops + b200_waiting_gets + b200_copy_inflight < 64;
}

/* ---- the seam ---- */
parsec_comm_engine_t *parsec_comm_engine_init(parsec_context_t *context)
{
    int up = 0;
    MPI_Initialized(&up);
    if( !up ) {
        parsec_warning("comm engine: MPI_Init (overlay/pb2mpi) has not been called; it is what tells a rank who it is");
        return NULL;
    }
    if( b200_up ) { parsec_ce.parsec_context = context; return &parsec_ce; }
    { const char *e = getenv("PB2_CE_STATS"); b200_stats_on = (NULL != e && atoi(e)); }
    {   /* This runs on the communication thread, which a runtime built without hwloc has just pinned to core number
         * nb_workers -- the same core for every rank of the box (remote_dep.c:688-699).  A launcher that has given
         * this rank its own cores says so in PB2_CE_CORES=lo-hi, and the thread goes back to them. */
        const char *e = getenv("PB2_CE_CORES");
        int lo = -1, hi = -1;
        if( NULL != e && 2 == sscanf(e, "%d-%d", &lo, &hi) && lo >= 0 && hi >= lo && hi < CPU_SETSIZE ) {
            cpu_set_t mask;
            CPU_ZERO(&mask);
            for( int c = lo; c <= hi; c++ ) CPU_SET(c, &mask);
            (void)sched_setaffinity(0, sizeof(mask), &mask);
        }
    }
    MPI_Comm_rank(MPI_COMM_WORLD, &b200_rank);
    MPI_Comm_size(MPI_COMM_WORLD, &b200_world);
    b200_tp = (pb2_ce_t*)PB2MPI_Transport();   /* NULL in a world of one: nothing to talk to */
    memset(b200_tags, 0, sizeof(b200_tags));
    if( NULL != b200_tp ) pb2_ce_set_wait_poll(b200_tp, b200_dispatch, NULL);

    parsec_ce.parsec_context = context;
    parsec_ce.capabilites.sided = 2;   /* r_tag of put / get is called as a function on the target, as the callers of the
                                        * MPI engine expect (remote_dep_mpi.c:1987-1991, dtd_test_ce.c:468-472) */
    parsec_ce.capabilites.supports_noncontiguous_datatype = 1;
    parsec_ce.capabilites.multithreaded = 0;
    parsec_ce.tag_register = b200_tag_register;
    parsec_ce.tag_unregister = b200_tag_unregister;
    parsec_ce.mem_register = b200_mem_register;
    parsec_ce.mem_unregister = b200_mem_unregister;
    parsec_ce.get_mem_handle_size = b200_get_mem_handle_size;
    parsec_ce.mem_retrieve = b200_mem_retrieve;
    parsec_ce.put = b200_put;
    parsec_ce.get = b200_get;
    parsec_ce.progress = b200_progress;
    parsec_ce.enable = b200_enable;
    parsec_ce.disable = b200_disable;
    parsec_ce.pack = b200_pack;
    parsec_ce.pack_size = b200_pack_size;
    parsec_ce.unpack = b200_unpack;
    parsec_ce.reshape = b200_reshape;
    parsec_ce.sync = b200_sync;
    parsec_ce.can_serve = b200_can_serve;
    parsec_ce.send_am = b200_send_am;
    if( NULL != context ) {
        context->nb_nodes = b200_world;
        context->my_rank = b200_rank;
    }
    b200_up = 1;
    return &parsec_ce;
}

int parsec_comm_engine_fini(parsec_comm_engine_t *ce)
{
    if( !b200_up ) return 1;
    /* what is still queued goes out before the tags disappear */
    for( int spin = 0; spin < 1000 && (NULL != b200_parked_head || NULL != b200_ops_head); spin++ ) b200_progress(ce);
    while( b200_copy_inflight > 0 ) b200_progress(ce);   /* the copiers finish what they hold */
    b200_stop_copiers();
    if( NULL != b200_tp ) pb2_ce_set_wait_poll(b200_tp, NULL, NULL);
    if( b200_stats_on )
        fprintf(stderr, "[comm engine, rank %d] puts %llu (%.1f MiB, %.2f ms in the copies = %.2f GB/s), registrations %llu, "
                "active messages in %llu (%.1f KiB, %.2f ms in their callbacks), progress calls %llu\n",
                b200_rank, (unsigned long long)b200_st.puts, b200_st.put_bytes / 1048576.0, b200_st.put_ns * 1e-6,
                b200_st.put_ns ? (double)b200_st.put_bytes / (double)b200_st.put_ns : 0.0, (unsigned long long)b200_st.regs,
                (unsigned long long)b200_st.ams, b200_st.am_bytes / 1024.0, b200_st.cb_ns * 1e-6,
                (unsigned long long)b200_st.progress_calls);
    memset(b200_tags, 0, sizeof(b200_tags));
    if( NULL != ce->parsec_context ) ce->parsec_context->comm_ctx = -1;   /* as the MPI engine leaves it (:601-604) */
    b200_tp = NULL;   /* the transport belongs to MPI_Finalize */
    b200_up = 0;
    return 1;
}
