// Intra-box data plane for the P x Q block-cyclic tiled GEMM: tile broadcasts over NCCL / NVLink instead of the
// reference's host-bounced MPI sends.
//
// What it replaces: for every output flow with remote successors the reference sends one ACTIVATE active message,
// the receiver allocates a HOST arena tile and answers GET, the sender MPI_Isend's the host copy, and the consumer GPU
// later pays an H2D (remote_dep.c:454-591, remote_dep_mpi.c:1593-1726, :1963-2118, parsec_mpi_funnelled.c:793,896);
// broadcasts emerge from rank_bits walked along a chain/binomial tree (remote_dep.c:334-372).
// For GEMM the communication pattern is static: A(m,k) goes to every rank of process row m mod P from
// column k mod Q, B(k,n) to every rank of process column n mod Q from row k mod P.  Here each of those is ONE
// ncclBroadcast on a per-row (size Q) or per-column (size P) communicator, GPU to GPU, received straight into an HBM
// cache slot that the consumer's stage-in finds "already on the device, same version" (no host bounce, no H2D).
//
// Ordering: every rank issues its collectives in one canonical order -- for k ascending: the row broadcasts of step k
// (m ascending), then the column broadcasts of step k (n ascending) -- on one stream.  A rank's sequence is a
// subsequence of that global order restricted to its two communicators, so the earliest incomplete collective always
// is the next one on every participant's stream: no deadlock, no tree tuning (NVSwitch is uniform).
#include <dlfcn.h>
#include <nccl.h>  // types only: the library is bound at run time, see nccl_api()
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "pb2_internal.h"
#include "pb2_types.h"

// NCCL is bound lazily with dlopen instead of at link time: a host process may already carry its own libnccl.so.2
// (PyTorch bundles a newer one than the system's) and two different NCCLs under one soname cannot coexist.  The copy
// that is already loaded wins; otherwise the system library is loaded.
struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommInitRankConfig)(ncclComm_t*, int, ncclUniqueId, int, ncclConfig_t*);  // optional
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    const char* (*GetErrorString)(ncclResult_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*CommAbort)(ncclComm_t);
    bool ok = false;
};
static NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.ok ? &api : nullptr;
    tried = true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        fprintf(stderr, "[parsec_b200] cannot load libnccl.so.2: %s\n", dlerror());
        return nullptr;
    }
    api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))dlsym(h, "ncclCommInitRank");
    api.CommInitRankConfig = (decltype(api.CommInitRankConfig))dlsym(h, "ncclCommInitRankConfig");
    api.CommDestroy = (decltype(api.CommDestroy))dlsym(h, "ncclCommDestroy");
    api.Broadcast = (decltype(api.Broadcast))dlsym(h, "ncclBroadcast");
    api.GetErrorString = (decltype(api.GetErrorString))dlsym(h, "ncclGetErrorString");
    api.GroupStart = (decltype(api.GroupStart))dlsym(h, "ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))dlsym(h, "ncclGroupEnd");
    api.AllReduce = (decltype(api.AllReduce))dlsym(h, "ncclAllReduce");
    api.CommAbort = (decltype(api.CommAbort))dlsym(h, "ncclCommAbort");
    api.ok = api.AllReduce && api.CommAbort && api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.Broadcast && api.GetErrorString &&
             api.GroupStart && api.GroupEnd;
    return api.ok ? &api : nullptr;
}

struct pb2_comm_s {
    int P, Q, rank, myrow, mycol;
    ncclComm_t row = nullptr, col = nullptr;  // row: the Q ranks of my process row; col: the P ranks of my column
    cudaStream_t stream = nullptr;
    int cuda_index = 0;
    int* flag = nullptr;  // one device int for pb2_comm_agree
};
typedef struct pb2_comm_s pb2_comm_t;

extern "C" int pb2_comm_unique_id(void* out128) {
    if (!out128) return PB2_ERR_BAD_PARAM;
    NcclApi* nccl = nccl_api();
    if (!nccl) return PB2_ERR_NOT_FOUND;
    ncclUniqueId id;
    if (nccl->GetUniqueId(&id) != ncclSuccess) return PB2_ERR_DEVICE;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    memcpy(out128, &id, 128);
    return PB2_SUCCESS;
}

static void comm_free(pb2_comm_s* c) {  // whatever exists of a communicator (also the failure paths of pb2_comm_create)
    if (c->row) nccl_api()->CommDestroy(c->row);
    if (c->col) nccl_api()->CommDestroy(c->col);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->flag) cudaFree(c->flag);
    delete c;
}

// row_id: the unique id published by the rank in column 0 of my process row; col_id: by the rank in row 0 of my column.
extern "C" int pb2_comm_create(int cuda_index, int P, int Q, int rank, const void* row_id, const void* col_id,
                               pb2_comm_t** out) {
    if (!out || P <= 0 || Q <= 0 || rank < 0 || rank >= P * Q) return PB2_ERR_BAD_PARAM;
    NcclApi* nccl = nccl_api();
    if (!nccl) return PB2_ERR_NOT_FOUND;
    if (cudaSetDevice(cuda_index) != cudaSuccess) return PB2_ERR_DEVICE;
    pb2_comm_s* c = new pb2_comm_s();
    c->P = P; c->Q = Q; c->rank = rank; c->myrow = rank / Q; c->mycol = rank % Q; c->cuda_index = cuda_index;
    // highest priority: the exec streams always have tile kernels queued, and a broadcast kernel that has to wait for
    // SM slots behind them turns the data plane into the critical path
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess) { comm_free(c); return PB2_ERR_DEVICE; }
    // A tile broadcast needs ~10 GB/s per GPU while the FP64 tile kernel wants every one of the 148 SMs
    // (one CTA per SM, register bound): NCCL's default channel count took SMs away from it for no benefit (8 GPUs:
    // kernel roofline fraction 0.951 against 0.983 on one GPU).  The communicators are therefore created with a small
    // CTA budget; PB2_NCCL_MAX_CTAS overrides it (0 = NCCL's default).
    int max_ctas = 2;
    if (const char* e = getenv("PB2_NCCL_MAX_CTAS")) max_ctas = atoi(e);
    auto init = [&](ncclComm_t* comm, int n, const ncclUniqueId& id, int me) {
        if (max_ctas > 0 && nccl->CommInitRankConfig) {
            ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
            cfg.minCTAs = 1;
            cfg.maxCTAs = max_ctas;
            if (nccl->CommInitRankConfig(comm, n, id, me, &cfg) == ncclSuccess) return true;
            *comm = nullptr;
            return false;
        }
        return nccl->CommInitRank(comm, n, id, me) == ncclSuccess;
    };
    if (cudaMalloc((void**)&c->flag, sizeof(int)) != cudaSuccess) { comm_free(c); return PB2_ERR_DEVICE; }
    ncclUniqueId rid, cid;
    if (Q > 1) {
        if (!row_id) { comm_free(c); return PB2_ERR_BAD_PARAM; }
        memcpy(&rid, row_id, 128);
        if (!init(&c->row, Q, rid, c->mycol)) { comm_free(c); return PB2_ERR_DEVICE; }
    }
    if (P > 1) {
        if (!col_id) { comm_free(c); return PB2_ERR_BAD_PARAM; }
        memcpy(&cid, col_id, 128);
        if (!init(&c->col, P, cid, c->myrow)) { comm_free(c); return PB2_ERR_DEVICE; }
    }
    *out = c;
    return PB2_SUCCESS;
}

extern "C" int pb2_comm_destroy(pb2_comm_t* c) {
    if (!c) return PB2_ERR_BAD_PARAM;
    cudaSetDevice(c->cuda_index);
    cudaStreamSynchronize(c->stream);
    comm_free(c);
    return PB2_SUCCESS;
}

// ncclGroupStart / ncclGroupEnd around the broadcasts of one GEMM step: NCCL fuses them into one launch per
// communicator instead of one kernel per tile
extern "C" int pb2_comm_group_begin(pb2_comm_t* c) {
    return (c && nccl_api() && nccl_api()->GroupStart() == ncclSuccess) ? PB2_SUCCESS : PB2_ERR_DEVICE;
}
extern "C" int pb2_comm_group_end(pb2_comm_t* c) {
    return (c && nccl_api() && nccl_api()->GroupEnd() == ncclSuccess) ? PB2_SUCCESS : PB2_ERR_DEVICE;
}

// Every rank of the grid says whether it can go ahead (ok != 0); returns 1 only if all of them can.  Two tiny
// all-reduces (MIN), along the process row and then along the process column, reach the whole P x Q grid.  Used before
// a taskpool issues its first broadcast: a rank that cannot take part (its HBM heap cannot hold the operands of one
// lookahead window, say) must not leave the others blocked inside a collective.
extern "C" int pb2_comm_agree(pb2_comm_t* c, int ok) {
    if (!c || !c->flag) return 0;
    int v = ok ? 1 : 0;
    if (cudaMemcpyAsync(c->flag, &v, sizeof(int), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) return 0;
    for (ncclComm_t comm : {c->row, c->col})
        if (comm && nccl_api()->AllReduce(c->flag, c->flag, 1, ncclInt, ncclMin, comm, c->stream) != ncclSuccess) return 0;
    if (cudaMemcpyAsync(&v, c->flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess) return 0;
    if (cudaStreamSynchronize(c->stream) != cudaSuccess) return 0;
    return v;
}

// A rank that has to leave the collective sequence half-way (a CUDA or allocation failure after pb2_comm_agree) aborts
// both communicators: its peers then get an error out of NCCL instead of waiting for ever.
extern "C" int pb2_comm_abort(pb2_comm_t* c) {
    if (!c) return PB2_ERR_BAD_PARAM;
    if (c->row) { nccl_api()->CommAbort(c->row); c->row = nullptr; }
    if (c->col) { nccl_api()->CommAbort(c->col); c->col = nullptr; }
    return PB2_SUCCESS;
}

extern "C" void* pb2_comm_stream(pb2_comm_t* c) { return c ? (void*)c->stream : nullptr; }
extern "C" int pb2_comm_geometry(const pb2_comm_t* c, int* out /*[5]*/) {
    if (!c || !out) return PB2_ERR_BAD_PARAM;
    out[0] = c->P; out[1] = c->Q; out[2] = c->rank; out[3] = c->myrow; out[4] = c->mycol;
    return PB2_SUCCESS;
}

// one tile broadcast; which = 0: along my process row (A tiles), 1: along my process column (B tiles);
// root is the rank index INSIDE that communicator.  In place: the root sends from buf, the others receive into buf.
extern "C" int pb2_comm_bcast(pb2_comm_t* c, int which, void* buf, size_t bytes, int root) {
    if (!c || !buf) return PB2_ERR_BAD_PARAM;
    ncclComm_t comm = which == 0 ? c->row : c->col;
    if (!comm) return PB2_SUCCESS;  // communicator of size 1
    ncclResult_t r = nccl_api()->Broadcast(buf, buf, bytes, ncclChar, root, comm, c->stream);
    if (r != ncclSuccess) {
        fprintf(stderr, "[parsec_b200] ncclBroadcast failed: %s\n", nccl_api()->GetErrorString(r));
        return PB2_ERR_DEVICE;
    }
    return PB2_SUCCESS;
}

// ------------------------------------------------------------------------------------ the plan (host logic)
// The canonical sequence of collectives of ONE rank for C(MT x NT) += A(MT x KT) * B(KT x NT) on a P x Q grid.
// Pure integer logic, no CUDA: tests/test_dist_gloo.py checks it across ranks on the CPU.
// pb2_bcast_op_t: which (0 = row communicator / an A tile, 1 = column communicator / a B tile), k (GEMM step),
// idx (m for A(m,k), n for B(k,n)), root (rank inside the communicator), i_am_root.
// General form: C (and the rows of A) dealt in super-tiles of kp tile rows, C (and the columns of B) in super-tiles of
// kq tile columns, the k dimension in super-tiles of a_kq (columns of A) and b_kp (rows of B), process-grid origin
// (ip, jq): twoDBC_kcyclic_rank_of, two_dim_rectangle_cyclic.c:532-566.  A(m,k) travels along the process row that owns
// tile row m, from the process column that owns tile column k of A; B(k,n) along the process column that owns tile
// column n, from the process row that owns tile row k of B.
extern "C" int64_t pb2_comm_plan_ex(int P, int Q, int rank, int MT, int NT, int KT, int kp, int kq, int a_kq, int b_kp, int ip,
                                    int jq, pb2_bcast_op_t* out, int64_t max_ops) {
    if (P <= 0 || Q <= 0 || rank < 0 || rank >= P * Q || MT <= 0 || NT <= 0 || KT <= 0 || kp <= 0 || kq <= 0 || a_kq <= 0 ||
        b_kp <= 0 || ip < 0 || ip >= P || jq < 0 || jq >= Q)
        return PB2_ERR_BAD_PARAM;
    const int myrow = rank / Q, mycol = rank % Q;
    int64_t n = 0;
    for (int k = 0; k < KT; ++k) {
        const int a_col = ((k / a_kq) % Q + jq) % Q, b_row = ((k / b_kp) % P + ip) % P;
        if (Q > 1)
            for (int m = 0; m < MT; ++m) {
                if (((m / kp) % P + ip) % P != myrow) continue;
                if (out && n < max_ops) out[n] = pb2_bcast_op_t{0, k, m, a_col, a_col == mycol};
                ++n;
            }
        if (P > 1)
            for (int c = 0; c < NT; ++c) {
                if (((c / kq) % Q + jq) % Q != mycol) continue;
                if (out && n < max_ops) out[n] = pb2_bcast_op_t{1, k, c, b_row, b_row == myrow};
                ++n;
            }
    }
    return n;
}

// plain cyclic ownership: A(m,k) on (m mod P, k mod Q), B(k,n) on (k mod P, n mod Q)
extern "C" int64_t pb2_comm_plan(int P, int Q, int rank, int MT, int NT, int KT, pb2_bcast_op_t* out, int64_t max_ops) {
    return pb2_comm_plan_ex(P, Q, rank, MT, NT, KT, 1, 1, 1, 1, 0, 0, out, max_ops);
}
