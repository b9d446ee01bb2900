// Tiled GEMM taskpool: the PTG dgemm_NN DAG unfolded over block-cyclic A, B, C and run through the device
// module.  Host mirror of what the reference runtime does for a generated taskpool
// (parsec_context_add_taskpool / start / wait, parsec/scheduling.c:678-806; generated hook
// parsec/interfaces/ptg/ptg-compiler/jdf2c.c:6727-6904; complete_hook jdf2c.c:7102):
//
//   READ_A(m,k) : descA(m,k)   READ A -> A GEMM(m, 0..NT-1, k)          (empty body)
//   READ_B(k,n) : descB(k,n)   READ B -> B GEMM(0..MT-1, n, k)          (empty body)
//   GEMM(m,n,k) : descC(m,n)   READ A, READ B, RW C <- (k==0) ? descC : C GEMM(m,n,k-1)
//                                                 -> (k==KT-1) ? descC : C GEMM(m,n,k+1)
//                 BODY [type=CUDA dyld=parsec_b200_dgemm]
//
// Owner-computes: a rank runs the chains of the C tiles it owns (task affinity descC(m,n)); READ_A / READ_B run
// where their tile lives and their ranged output deps are what becomes the row / column broadcasts
// (comm_nccl.cu).  The C ABI keeps the reference's split between "the hook builds a parsec_gpu_task_t" and
// "the device module runs it": this file only builds tasks and reacts to completions, like the generated code.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>

#include "pb2_internal.h"
#include "pb2_types.h"

extern "C" {
int pb2_matrix_geometry(const pb2_matrix_block_cyclic_t* dc, int* out);
void* pb2_comm_stream(pb2_comm_t* c);
int pb2_comm_group_begin(pb2_comm_t* c);
int pb2_comm_group_end(pb2_comm_t* c);
int pb2_comm_agree(pb2_comm_t* c, int ok);
int pb2_comm_abort(pb2_comm_t* c);
size_t pb2_device_heap_capacity(pb2_device_t* d);
int pb2_comm_geometry(const pb2_comm_t* c, int* out);
int pb2_device_pin_for_read(pb2_device_t* d, pb2_data_t* data, void** dev_ptr, void* wait_stream);
int pb2_device_unpin(pb2_device_t* d, pb2_data_t* data);
int pb2_device_adopt_remote(pb2_device_t* d, pb2_data_t* data, uint32_t version, void** dev_ptr);
int pb2_device_set_ready_event(pb2_device_t* d, pb2_data_t* data, void* cuda_event);
int pb2_device_drop_remote(pb2_device_t* d, pb2_data_t* data);
}

struct Geo {
    int mtype, es, mb, nb, m, n, mt, nt, P, Q, rank, dc_id, kp, kq, ip, jq, i, j;
};
static Geo geo_of(const pb2_matrix_block_cyclic_t* dc) {
    int v[18];
    pb2_matrix_geometry(dc, v);
    return Geo{v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8], v[9], v[10], v[11], v[12], v[13], v[14], v[15], v[16], v[17]};
}

struct pb2_taskpool_s {
    pb2_matrix_block_cyclic_t *A, *B, *C;
    pb2_device_t* dev;
    pb2_comm_t* comm = nullptr;
    Geo gA, gB, gC;
    int MT, NT, KT, P, Q, myrow, mycol;
    // C-tile chains in flight (0 = all).  With every chain active the DAG advances k-outer in lockstep and all C
    // tiles finish, and are written back, together at the very end; a window lets finished tiles leave (D2H) while
    // later chains compute: +4 % resident, +7 % end to end at N=32768 (measured with 64..256, B200).
    // -1 = chosen in run(): 128 chains, or 16 when a whole chain computes in about the time its C tile takes to go
    // back over PCIe (tensor-core bodies): there the write-back is the critical path and has to start early
    // (bf16 N=32768 NB=2048: 658 -> 830 TFLOP/s resident), while the FP64 body wants the deeper stage-in lookahead.
    int64_t window = -1;
    std::vector<std::pair<int, int>> chains;
    size_t next_chain = 0;
    int64_t nb_gemm = 0, nb_read_a = 0, nb_read_b = 0;
    int status = 0;
    // tiles of A / B this rank's GEMMs read: local rows x KT and KT x local columns; remote ones are proxies
    // row_slot[m] / col_slot[n]: index of tile row m / tile column n among the ones this rank's process row / column
    // owns (-1: not mine), whatever the k-cyclicity and the grid origin
    std::vector<pb2_data_s*> a_tiles, b_tiles;
    std::vector<int> row_slot, col_slot;
    std::vector<pb2_data_s*> proxies, pinned;
    // consumers still to retire for each received (proxy) tile: A(m,k) feeds one GEMM per local tile column, B(k,n) one per
    // local tile row.  When the count reaches zero the HBM slot goes back to the heap, so that what a rank holds of the
    // other ranks' operands is bounded by the broadcast lookahead instead of growing with the whole taskpool.
    std::vector<int32_t> a_left, b_left;   // -1 = a tile this rank owns
    int64_t dropped = 0;
    std::vector<cudaEvent_t> comm_events;
    // the broadcast plan (canonical order, k ascending) and how far it has been issued.  Group k is issued no later than
    // the moment the first GEMM of step k - bcast_lookahead is spawned, so a root's stage-in of the tiles it
    // broadcasts interleaves, on the one h2d stream, with the stage-in the tasks themselves ask for; issuing the whole
    // plan up front (bcast_lookahead = 0) puts gigabytes of A / B tiles ahead of the first task's C tile and cost
    // 0.13 / 0.22 / 0.38 s per end-to-end step on 2 / 4 / 8 GPUs.
    std::vector<pb2_bcast_op_t> plan;
    size_t plan_pos = 0;
    int64_t bcast_lookahead = 4;
};

static pb2_data_s* a_tile(pb2_taskpool_s* tp, int m, int k) { return tp->a_tiles[(size_t)tp->row_slot[m] * tp->KT + k]; }
static pb2_data_s* b_tile(pb2_taskpool_s* tp, int k, int n) { return tp->b_tiles[(size_t)tp->col_slot[n] * tp->KT + k]; }

static int gemm_submit(pb2_device_t* dev, pb2_gpu_task_t* t, void* stream) {
    (void)dev;
    pb2_taskpool_s* tp = (pb2_taskpool_s*)t->complete_cookie;
    const int rm = (int)t->locals[3], rn = (int)t->locals[4], rk = (int)t->locals[5];
    int rc;
    if (tp->gC.mtype == PB2_MATRIX_DOUBLE) {
        rc = parsec_b200_dgemm('N', 'N', rm, rn, rk, 1.0, (const double*)t->flow_dev_ptr[0], tp->gA.mb,
                               (const double*)t->flow_dev_ptr[1], tp->gB.mb, 1.0, (double*)t->flow_dev_ptr[2], tp->gC.mb,
                               stream);
    } else if (tp->gA.mtype == PB2_MATRIX_FLOAT) {
        rc = parsec_b200_gemm_tf32_sm100('N', 'N', rm, rn, rk, 1.0f, (const float*)t->flow_dev_ptr[0], tp->gA.mb,
                                         (const float*)t->flow_dev_ptr[1], tp->gB.mb, 1.0f, (float*)t->flow_dev_ptr[2],
                                         tp->gC.mb, stream);
    } else {
        rc = parsec_b200_gemm_bf16_sm100('N', 'N', rm, rn, rk, 1.0f, t->flow_dev_ptr[0], tp->gA.mb, t->flow_dev_ptr[1],
                                         tp->gB.mb, 1.0f, (float*)t->flow_dev_ptr[2], tp->gC.mb, stream);
    }
    return rc == PB2_SUCCESS ? PB2_HOOK_RETURN_DONE : PB2_HOOK_RETURN_ERROR;
}

static void on_complete(pb2_gpu_task_t* t, int status, void* cookie);

static int issue_plan_upto(pb2_taskpool_s* tp, int k_limit);

static int spawn_gemm(pb2_taskpool_s* tp, int m, int n, int k) {
    if (tp->plan_pos < tp->plan.size()) {  // the tiles of step k (and of the lookahead) must have a place before the task is queued
        int prc = issue_plan_upto(tp, k + 1 + (int)tp->bcast_lookahead);
        if (prc != PB2_SUCCESS) return prc;
    }
    pb2_gpu_task_t* t = (pb2_gpu_task_t*)calloc(1, sizeof(pb2_gpu_task_t));
    if (!t) return PB2_ERR_OUT_OF_RESOURCE;
    t->task_type = 0;
    t->submit = gemm_submit;
    t->nb_flows = 3;
    t->flow_data[0] = a_tile(tp, m, k);
    t->flow_data[1] = b_tile(tp, k, n);
    t->flow_data[2] = pb2_matrix_data_of(tp->C, m, n);
    if (!t->flow_data[0] || !t->flow_data[1] || !t->flow_data[2]) {
        free(t);
        return PB2_ERR_NOT_FOUND;
    }
    t->flow_access[0] = PB2_FLOW_ACCESS_READ;
    t->flow_access[1] = PB2_FLOW_ACCESS_READ;
    t->flow_access[2] = PB2_FLOW_ACCESS_RW;
    t->flow_nb_elts[0] = (uint32_t)((size_t)tp->gA.mb * tp->gA.nb * tp->gA.es);
    t->flow_nb_elts[1] = (uint32_t)((size_t)tp->gB.mb * tp->gB.nb * tp->gB.es);
    t->flow_nb_elts[2] = (uint32_t)((size_t)tp->gC.mb * tp->gC.nb * tp->gC.es);
    t->trace_dc[0] = 0;
    t->trace_dc[1] = 1;
    t->trace_dc[2] = 2;
    // pushout: only the last task of the chain writes to memory (the JDF guard k == KT-1, jdf2c.c:6839-6893)
    t->pushout = (k == tp->KT - 1) ? (1 << 2) : 0;
    t->priority = tp->KT - k;
    t->locals[0] = m;
    t->locals[1] = n;
    t->locals[2] = k;
    t->locals[3] = (m == tp->MT - 1) ? tp->gC.m - m * tp->gC.mb : tp->gC.mb;
    t->locals[4] = (n == tp->NT - 1) ? tp->gC.n - n * tp->gC.nb : tp->gC.nb;
    t->locals[5] = (k == tp->KT - 1) ? tp->gA.n - k * tp->gA.nb : tp->gA.nb;
    t->complete = on_complete;
    t->complete_cookie = tp;
    int rc = pb2_kernel_scheduler(tp->dev, t);
    return rc == PB2_HOOK_RETURN_ASYNC ? PB2_SUCCESS : PB2_ERR_DEVICE;
}

static void on_complete(pb2_gpu_task_t* t, int status, void* cookie) {
    pb2_taskpool_s* tp = (pb2_taskpool_s*)cookie;
    if (status != PB2_HOOK_RETURN_DONE) {
        tp->status = status;
        return;
    }
    // generated complete_hook: data_out->version++ for every WRITE flow; data_out is the host copy
    {
        pb2_data_s* c = t->flow_data[2];
        std::lock_guard<std::mutex> g(c->lock);
        c->host_copy.version++;
    }
    tp->nb_gemm++;
    const int m = (int)t->locals[0], n = (int)t->locals[1], k = (int)t->locals[2];
    if (!tp->a_left.empty()) {  // P x Q: give received operands back once their last local reader has retired
        const size_t ia = (size_t)tp->row_slot[m] * tp->KT + k, ib = (size_t)tp->col_slot[n] * tp->KT + k;
        if (tp->a_left[ia] > 0 && --tp->a_left[ia] == 0 && pb2_device_drop_remote(tp->dev, tp->a_tiles[ia]) == PB2_SUCCESS) tp->dropped++;
        if (tp->b_left[ib] > 0 && --tp->b_left[ib] == 0 && pb2_device_drop_remote(tp->dev, tp->b_tiles[ib]) == PB2_SUCCESS) tp->dropped++;
    }
    if (tp->status) return;
    if (k + 1 < tp->KT) {
        int rc = spawn_gemm(tp, m, n, k + 1);  // release_deps: C GEMM(m,n,k) -> C GEMM(m,n,k+1)
        if (rc) tp->status = rc;
    } else if (tp->next_chain < tp->chains.size()) {
        auto mn = tp->chains[tp->next_chain++];
        int rc = spawn_gemm(tp, mn.first, mn.second, 0);
        if (rc) tp->status = rc;
    }
}

extern "C" pb2_taskpool_t* pb2_gemm_taskpool_new(pb2_matrix_block_cyclic_t* A, pb2_matrix_block_cyclic_t* B,
                                                 pb2_matrix_block_cyclic_t* C, pb2_device_t* dev) {
    if (!A || !B || !C || !dev) return nullptr;
    auto* tp = new pb2_taskpool_s();
    tp->A = A; tp->B = B; tp->C = C; tp->dev = dev;
    tp->gA = geo_of(A); tp->gB = geo_of(B); tp->gC = geo_of(C);
    // C (M x N) += A (M x K) * B (K x N): conforming tilings on one process grid
    bool ok = tp->gA.m == tp->gC.m && tp->gB.n == tp->gC.n && tp->gA.n == tp->gB.m && tp->gA.mb == tp->gC.mb &&
              tp->gB.nb == tp->gC.nb && tp->gA.nb == tp->gB.mb && tp->gA.P == tp->gC.P && tp->gA.Q == tp->gC.Q &&
              tp->gB.P == tp->gC.P && tp->gB.Q == tp->gC.Q && tp->gA.mtype == tp->gB.mtype &&
              tp->gA.rank == tp->gC.rank && tp->gB.rank == tp->gC.rank;
    bool types = (tp->gC.mtype == PB2_MATRIX_DOUBLE && tp->gA.mtype == PB2_MATRIX_DOUBLE) ||
                 (tp->gC.mtype == PB2_MATRIX_FLOAT && tp->gA.mtype == PB2_MATRIX_BF16) ||
                 (tp->gC.mtype == PB2_MATRIX_FLOAT && tp->gA.mtype == PB2_MATRIX_FLOAT);  // fp32 tiles: TF32 tensor cores
    // owner-computes with row / column broadcasts: A(m,k) must live in the process row of C(m,n) and B(k,n) in its
    // process column, i.e. A and C deal tile rows alike, B and C deal tile columns alike, on one grid origin; how the
    // k dimension is dealt (kq of A, kp of B) is free
    bool cyclic = tp->gA.kp == tp->gC.kp && tp->gB.kq == tp->gC.kq && tp->gA.ip == tp->gC.ip && tp->gB.ip == tp->gC.ip &&
                  tp->gA.jq == tp->gC.jq && tp->gB.jq == tp->gC.jq;
    // sub-matrix views (i, j != 0) shift the ownership of tile (0, 0); only whole matrices on a P x Q grid
    if (tp->gC.P * tp->gC.Q > 1 && (tp->gA.i || tp->gA.j || tp->gB.i || tp->gB.j || tp->gC.i || tp->gC.j)) cyclic = false;
    if (!ok || !types || !cyclic) {
        delete tp;
        return nullptr;
    }
    tp->MT = tp->gC.mt; tp->NT = tp->gC.nt; tp->KT = tp->gA.nt;
    tp->P = tp->gC.P; tp->Q = tp->gC.Q;
    tp->myrow = tp->gC.rank / tp->Q;
    tp->mycol = tp->gC.rank % tp->Q;
    int rows = 0, cols = 0;
    tp->row_slot.assign((size_t)tp->MT, -1);
    tp->col_slot.assign((size_t)tp->NT, -1);
    for (int m = 0; m < tp->MT; ++m)
        if (((m / tp->gC.kp) % tp->P + tp->gC.ip) % tp->P == tp->myrow) tp->row_slot[m] = rows++;
    for (int n = 0; n < tp->NT; ++n)
        if (((n / tp->gC.kq) % tp->Q + tp->gC.jq) % tp->Q == tp->mycol) tp->col_slot[n] = cols++;
    return tp;
}

extern "C" int pb2_gemm_taskpool_set(pb2_taskpool_t* tp, const char* key, int64_t value) {
    if (!tp || !key) return PB2_ERR_BAD_PARAM;
    if (!strcmp(key, "window")) tp->window = value;
    else if (!strcmp(key, "bcast_lookahead")) tp->bcast_lookahead = value < 0 ? 0 : value;  // 0 = whole plan up front
    else return PB2_ERR_NOT_FOUND;
    return PB2_SUCCESS;
}

// The communicator a P x Q taskpool moves its A / B tiles with (pb2_comm_create); must match the descriptors' grid.
extern "C" int pb2_gemm_taskpool_set_comm(pb2_taskpool_t* tp, pb2_comm_t* comm) {
    if (!tp) return PB2_ERR_BAD_PARAM;
    if (comm) {
        int g[5];
        pb2_comm_geometry(comm, g);
        if (g[0] != tp->P || g[1] != tp->Q || g[2] != tp->gC.rank) return PB2_ERR_BAD_PARAM;
    }
    tp->comm = comm;
    return PB2_SUCCESS;
}

extern "C" void pb2_gemm_taskpool_free(pb2_taskpool_t* tp) { delete tp; }

// Collect the A / B tiles this rank's GEMMs read and, on a P x Q grid, move the remote ones: the whole broadcast
// schedule is enqueued up front on the comm stream in canonical order; consumers wait on per-step events.
// Issues the broadcast groups with k < k_limit that have not been issued yet, in plan order (every rank walks the same
// plan, so each communicator sees the same sequence of collectives whatever the timing).  Manager thread only: called
// before the device is driven and from the completion callback.
static int issue_plan_upto(pb2_taskpool_s* tp, int k_limit) {
    if (tp->plan_pos >= tp->plan.size() || tp->plan[tp->plan_pos].k >= k_limit) return PB2_SUCCESS;
    cudaStream_t cs = (cudaStream_t)pb2_comm_stream(tp->comm);
    std::vector<pb2_data_s*> group;  // received tiles waiting for their group's event
    while (tp->plan_pos < tp->plan.size() && tp->plan[tp->plan_pos].k < k_limit) {
        const int k = tp->plan[tp->plan_pos].k;
        int rc = pb2_comm_group_begin(tp->comm);  // one fused launch per communicator for the whole GEMM step
        if (rc != PB2_SUCCESS) return rc;
        group.clear();
        for (; rc == PB2_SUCCESS && tp->plan_pos < tp->plan.size() && tp->plan[tp->plan_pos].k == k; ++tp->plan_pos) {
            const pb2_bcast_op_t& op = tp->plan[tp->plan_pos];
            pb2_data_s* d = op.which == 0 ? a_tile(tp, op.idx, op.k) : b_tile(tp, op.k, op.idx);
            void* ptr = nullptr;
            if (op.i_am_root) {
                rc = pb2_device_pin_for_read(tp->dev, d, &ptr, cs);  // H2D if needed; the comm stream waits for it
                if (rc == PB2_SUCCESS) tp->pinned.push_back(d);
            } else {
                rc = pb2_device_adopt_remote(tp->dev, d, 0, &ptr);   // sender's version: read-only operands are at 0
                if (rc == PB2_SUCCESS) group.push_back(d);
            }
            if (rc != PB2_SUCCESS) {
                fprintf(stderr, "[parsec_b200] cannot place tile (%s, k=%d, idx=%d) in HBM: %d\n", op.which ? "B" : "A", op.k,
                        op.idx, rc);
                break;
            }
            rc = pb2_comm_bcast(tp->comm, op.which, ptr, d->nb_elts, op.root);
        }
        {
            int erc = pb2_comm_group_end(tp->comm);
            if (rc == PB2_SUCCESS) rc = erc;
        }
        if (rc == PB2_SUCCESS && !group.empty()) {
            cudaEvent_t e;
            cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
            cudaEventRecord(e, cs);
            tp->comm_events.push_back(e);
            for (pb2_data_s* d : group) pb2_device_set_ready_event(tp->dev, d, e);
        }
        if (rc != PB2_SUCCESS) return rc;
    }
    return PB2_SUCCESS;
}

static int prepare_operands(pb2_taskpool_s* tp) {
    int rows = 0, cols = 0;
    for (int s : tp->row_slot) rows += s >= 0;
    for (int s : tp->col_slot) cols += s >= 0;
    tp->a_tiles.assign((size_t)rows * tp->KT, nullptr);
    tp->b_tiles.assign((size_t)cols * tp->KT, nullptr);
    const uint32_t a_bytes = (uint32_t)((size_t)tp->gA.mb * tp->gA.nb * tp->gA.es);
    const uint32_t b_bytes = (uint32_t)((size_t)tp->gB.mb * tp->gB.nb * tp->gB.es);
    for (int m = 0; m < tp->MT; ++m)
        for (int k = 0; k < tp->KT && tp->row_slot[m] >= 0; ++k) {
            pb2_data_s* d = pb2_matrix_data_of(tp->A, m, k);
            if (d) tp->nb_read_a++;  // READ_A(m,k) runs here
            else {
                d = pb2_data_create(pb2_matrix_data_key(tp->A, m, k), (uint32_t)tp->gA.dc_id, nullptr, a_bytes);
                tp->proxies.push_back(d);
            }
            tp->a_tiles[(size_t)tp->row_slot[m] * tp->KT + k] = d;
        }
    for (int n = 0; n < tp->NT; ++n)
        for (int k = 0; k < tp->KT && tp->col_slot[n] >= 0; ++k) {
            pb2_data_s* d = pb2_matrix_data_of(tp->B, k, n);
            if (d) tp->nb_read_b++;
            else {
                d = pb2_data_create(pb2_matrix_data_key(tp->B, k, n), (uint32_t)tp->gB.dc_id, nullptr, b_bytes);
                tp->proxies.push_back(d);
            }
            tp->b_tiles[(size_t)tp->col_slot[n] * tp->KT + k] = d;
        }
    tp->plan.clear();
    tp->plan_pos = 0;
    tp->a_left.clear();
    tp->b_left.clear();
    tp->dropped = 0;
    if (tp->P * tp->Q == 1) return PB2_SUCCESS;
    tp->a_left.assign(tp->a_tiles.size(), -1);
    tp->b_left.assign(tp->b_tiles.size(), -1);
    for (int m = 0; m < tp->MT; ++m)
        for (int k = 0; k < tp->KT && tp->row_slot[m] >= 0; ++k)
            if (!pb2_matrix_data_of(tp->A, m, k)) {  // a proxy: read by GEMM(m, n, k) for every local C tile of tile row m
                int readers = 0;
                for (int n = 0; n < tp->NT; ++n) readers += pb2_matrix_rank_of(tp->C, m, n) == tp->gC.rank;
                tp->a_left[(size_t)tp->row_slot[m] * tp->KT + k] = readers;
            }
    for (int n = 0; n < tp->NT; ++n)
        for (int k = 0; k < tp->KT && tp->col_slot[n] >= 0; ++k)
            if (!pb2_matrix_data_of(tp->B, k, n)) {
                int readers = 0;
                for (int m = 0; m < tp->MT; ++m) readers += pb2_matrix_rank_of(tp->C, m, n) == tp->gC.rank;
                tp->b_left[(size_t)tp->col_slot[n] * tp->KT + k] = readers;
            }
    // READ tasks of tiles that only feed remote GEMMs still execute on their owner: count the owned tiles outside the
    // rows / columns walked above (none on a grid where every rank owns C tiles in each of its rows and columns).
    int64_t nops = pb2_comm_plan_ex(tp->P, tp->Q, tp->gC.rank, tp->MT, tp->NT, tp->KT, tp->gC.kp, tp->gC.kq, tp->gA.kq,
                                    tp->gB.kp, tp->gC.ip, tp->gC.jq, nullptr, 0);
    if (nops < 0) return (int)nops;
    tp->plan.resize((size_t)nops);
    pb2_comm_plan_ex(tp->P, tp->Q, tp->gC.rank, tp->MT, tp->NT, tp->KT, tp->gC.kp, tp->gC.kq, tp->gA.kq, tp->gB.kp,
                     tp->gC.ip, tp->gC.jq, tp->plan.data(), nops);
    // Before the first collective every rank checks that it can hold what the schedule will ask of it -- the operands of
    // the k-steps in flight (lookahead + the steps the window of chains spans, or all of them with bcast_lookahead = 0)
    // next to one window of C tiles -- and the grid agrees: a rank that would fail half-way would otherwise leave its
    // peers blocked inside ncclBroadcast.
    {
        const int64_t steps = tp->bcast_lookahead > 0 ? std::min<int64_t>(tp->KT, tp->bcast_lookahead + 2) : tp->KT;
        const size_t c_bytes = (size_t)tp->gC.mb * tp->gC.nb * tp->gC.es;
        const size_t need = (size_t)steps * ((size_t)rows * a_bytes + (size_t)cols * b_bytes) + 4 * c_bytes;
        const bool fits = need <= pb2_device_heap_capacity(tp->dev);
        if (!fits)
            fprintf(stderr, "[parsec_b200] rank %d: the HBM heap (%zu bytes) cannot hold %lld k-steps of operands (%zu bytes)\n",
                    tp->gC.rank, pb2_device_heap_capacity(tp->dev), (long long)steps, need);
        if (!pb2_comm_agree(tp->comm, fits ? 1 : 0)) return PB2_ERR_OUT_OF_RESOURCE;
    }
    return issue_plan_upto(tp, tp->bcast_lookahead > 0 ? (int)std::min<int64_t>(tp->bcast_lookahead, tp->KT) : tp->KT);
}

static void release_operands(pb2_taskpool_s* tp) {
    if (tp->comm) cudaStreamSynchronize((cudaStream_t)pb2_comm_stream(tp->comm));
    for (pb2_data_s* d : tp->pinned) pb2_device_unpin(tp->dev, d);
    for (pb2_data_s* d : tp->proxies) pb2_data_destroy(d);  // frees the HBM slot of received tiles
    for (cudaEvent_t e : tp->comm_events) cudaEventDestroy(e);
    tp->pinned.clear();
    tp->proxies.clear();
    tp->comm_events.clear();
    tp->a_tiles.clear();
    tp->b_tiles.clear();
}

extern "C" int pb2_gemm_taskpool_run(pb2_taskpool_t* tp, pb2_gemm_stats_t* stats) {
    if (!tp) return PB2_ERR_BAD_PARAM;
    if (tp->P * tp->Q != 1 && !tp->comm) {
        fprintf(stderr, "[parsec_b200] a %d x %d taskpool needs a communicator (pb2_gemm_taskpool_set_comm)\n", tp->P, tp->Q);
        return PB2_ERR_BAD_PARAM;
    }
    auto t0 = std::chrono::steady_clock::now();
    tp->nb_gemm = tp->nb_read_a = tp->nb_read_b = 0;
    tp->status = 0;
    tp->chains.clear();
    tp->next_chain = 0;
    // chain activation order: column by column of C, so that finished tiles leave while later columns compute
    for (int n = 0; n < tp->NT; ++n)
        for (int m = 0; m < tp->MT; ++m)
            if (pb2_matrix_rank_of(tp->C, m, n) == tp->gC.rank) tp->chains.push_back({m, n});
    if (!pb2_device_hold(tp->dev)) return PB2_ERROR;  // somebody else is driving this device
    int64_t w = tp->window > 0 ? tp->window : (int64_t)tp->chains.size();
    if (tp->window < 0) {
        // chain time / write-back time of its C tile = 2 K pcie / (rate * sizeof(C element)), nominal rates
        const double rate = tp->gA.mtype == PB2_MATRIX_FLOAT ? 700e12 : 1100e12;
        const double ratio = 2.0 * (double)tp->gA.n * 50e9 / (rate * (double)tp->gC.es);
        w = (tp->gC.mtype != PB2_MATRIX_DOUBLE && ratio < 2.0) ? 16 : 128;
    }
    int rc = prepare_operands(tp);
    if (rc != PB2_SUCCESS) tp->status = rc;
    while (!tp->status && tp->next_chain < tp->chains.size() && (int64_t)tp->next_chain < w) {
        auto mn = tp->chains[tp->next_chain++];
        rc = spawn_gemm(tp, mn.first, mn.second, 0);
        if (rc) tp->status = rc;
    }
    rc = pb2_device_manage(tp->dev);  // add_taskpool + start + wait
    if (rc != PB2_SUCCESS && !tp->status) tp->status = rc;
    if (!tp->status && tp->plan_pos < tp->plan.size()) {  // a rank with no GEMM of its own still serves its broadcasts
        rc = issue_plan_upto(tp, tp->KT);
        if (rc != PB2_SUCCESS) tp->status = rc;
    }
    if (tp->status && tp->comm && tp->plan_pos < tp->plan.size()) {
        // this rank leaves the canonical sequence with broadcasts still owed: make its peers fail instead of hang
        fprintf(stderr, "[parsec_b200] rank %d aborts its communicators (status %d)\n", tp->gC.rank, tp->status);
        pb2_comm_abort(tp->comm);
    }
    release_operands(tp);
    auto t1 = std::chrono::steady_clock::now();
    if (stats) {
        stats->nb_gemm = tp->nb_gemm;
        stats->nb_read_a = tp->nb_read_a;
        stats->nb_read_b = tp->nb_read_b;
        stats->seconds = std::chrono::duration<double>(t1 - t0).count();
    }
    return tp->status ? PB2_ERROR : PB2_SUCCESS;
}
