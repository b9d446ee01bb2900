// A minimal DTD ("dynamic task discovery") front-end over the same device boundary: tasks are inserted in program
// order with per-flow access modes, dependencies are inferred from the data they touch, and every ready task goes
// through pb2_kernel_scheduler exactly like a PTG task.
//
// Mirrors, for GPU tasks only, parsec_dtd_insert_task / parsec_dtd_insert_task_with_task_class
// (parsec/interfaces/dtd/insert_function.c) and parsec_dtd_gpu_task_submit (:2359-2406): the gpu_task gets
// submit = the task class' GPU function, pushout from the per-flow PARSEC_PUSHOUT flag (insert_function.h:70), the
// default stage functions, and the body reads its device pointers through parsec_dtd_get_dev_ptr (:3599-3612), here
// pb2_gpu_task_t.flow_dev_ptr[i].  Fixture: tests/dsl/dtd/dtd_test_simple_gemm.c:260-330 (the only C += A*B program
// of the reference tree).
//
// Single host thread: insertion only queues (the taskpool holds the device's manager seat from creation),
// pb2_dtd_taskpool_wait drives the device until every inserted task has completed.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <unordered_map>
#include <vector>

#include "pb2_internal.h"
#include "pb2_types.h"

struct pb2_dtd_taskpool_s;

struct DtdTask {
    pb2_dtd_taskpool_s* tp;
    pb2_gpu_task_t proto;  // copied into a calloc'd pb2_gpu_task_t at submission (the module frees that one)
    int pending = 0;
    bool done = false;
    std::vector<DtdTask*> successors;
};

struct DtdDataState {
    DtdTask* last_writer = nullptr;
    std::vector<DtdTask*> readers;  // readers since the last write (write-after-read dependencies)
};

struct pb2_dtd_taskpool_s {
    pb2_device_t* dev;
    bool holding = false;
    std::unordered_map<pb2_data_s*, DtdDataState> state;
    std::vector<DtdTask*> tasks;
    int64_t inserted = 0, completed = 0;
    int status = 0;
};
typedef struct pb2_dtd_taskpool_s pb2_dtd_taskpool_t;

static void dtd_complete(pb2_gpu_task_t* t, int status, void* cookie);

static void dtd_submit(DtdTask* dt) {
    pb2_gpu_task_t* t = (pb2_gpu_task_t*)calloc(1, sizeof(pb2_gpu_task_t));
    *t = dt->proto;
    t->complete = dtd_complete;
    t->complete_cookie = dt;
    int rc = pb2_kernel_scheduler(dt->tp->dev, t);
    if (rc != PB2_HOOK_RETURN_ASYNC) dt->tp->status = rc;
}

static void dtd_complete(pb2_gpu_task_t* t, int status, void* cookie) {
    DtdTask* dt = (DtdTask*)cookie;
    pb2_dtd_taskpool_s* tp = dt->tp;
    if (status != PB2_HOOK_RETURN_DONE) tp->status = status;
    // the runtime's completion: version++ on the (host) data_out of every WRITE flow, then release the successors
    for (int i = 0; i < t->nb_flows; ++i)
        if ((t->flow_access[i] & PB2_FLOW_ACCESS_WRITE) && t->flow_data[i]) {
            std::lock_guard<std::mutex> g(t->flow_data[i]->lock);
            t->flow_data[i]->host_copy.version++;
        }
    dt->done = true;
    tp->completed++;
    for (DtdTask* s : dt->successors)
        if (--s->pending == 0 && !tp->status) dtd_submit(s);
}

// parsec_dtd_taskpool_new + parsec_context_add_taskpool + parsec_context_start
extern "C" pb2_dtd_taskpool_t* pb2_dtd_taskpool_new(pb2_device_t* dev) {
    if (!dev) return nullptr;
    if (!pb2_device_hold(dev)) return nullptr;  // another thread drives this device
    auto* tp = new pb2_dtd_taskpool_s();
    tp->dev = dev;
    tp->holding = true;
    return tp;
}

// parsec_dtd_insert_task_with_task_class: flows in argument order; access = PB2_FLOW_ACCESS_{READ,WRITE,RW}
// (PARSEC_INPUT / PARSEC_OUTPUT / PARSEC_INOUT), pushout_mask bit i = PARSEC_PUSHOUT on flow i; `locals` are the
// PARSEC_VALUE arguments the body unpacks.
extern "C" int pb2_dtd_insert_task(pb2_dtd_taskpool_t* tp, pb2_submit_fn gpu_body, int priority, int nb_flows,
                                   pb2_data_t* const* data, const uint8_t* access, int32_t pushout_mask,
                                   const int64_t* locals, int nb_locals) {
    if (!tp || !gpu_body || nb_flows < 0 || nb_flows > PB2_MAX_PARAM_COUNT || nb_locals < 0 || nb_locals > 8)
        return PB2_ERR_BAD_PARAM;
    for (int i = 0; i < nb_flows; ++i)
        if (!data[i] || !(access[i] & PB2_FLOW_ACCESS_RW)) return PB2_ERR_BAD_PARAM;
    DtdTask* dt = new DtdTask();
    dt->tp = tp;
    memset(&dt->proto, 0, sizeof(dt->proto));
    dt->proto.task_type = 0;
    dt->proto.submit = gpu_body;
    dt->proto.priority = priority;
    dt->proto.nb_flows = nb_flows;
    dt->proto.pushout = pushout_mask;
    for (int i = 0; i < nb_flows; ++i) {
        dt->proto.flow_data[i] = data[i];
        dt->proto.flow_access[i] = access[i];
        dt->proto.flow_nb_elts[i] = data[i]->nb_elts;
        dt->proto.trace_dc[i] = (int32_t)data[i]->dc_id;
    }
    for (int i = 0; i < nb_locals; ++i) dt->proto.locals[i] = locals[i];
    tp->tasks.push_back(dt);
    tp->inserted++;
    auto depend_on = [&](DtdTask* pred) {
        if (pred && pred != dt && !pred->done) {
            pred->successors.push_back(dt);
            dt->pending++;
        }
    };
    for (int i = 0; i < nb_flows; ++i) {
        DtdDataState& s = tp->state[data[i]];
        if (access[i] & PB2_FLOW_ACCESS_WRITE) {
            depend_on(s.last_writer);                      // write after write / read after write
            for (DtdTask* r : s.readers) depend_on(r);     // write after read
            s.last_writer = dt;
            s.readers.clear();
        } else {
            depend_on(s.last_writer);                      // read after write
            s.readers.push_back(dt);
        }
    }
    if (dt->pending == 0) dtd_submit(dt);
    return tp->status ? PB2_ERROR : PB2_SUCCESS;
}

// parsec_dtd_data_flush_all + parsec_dtd_taskpool_wait + parsec_context_wait
extern "C" int pb2_dtd_taskpool_wait(pb2_dtd_taskpool_t* tp, int64_t* nb_executed) {
    if (!tp) return PB2_ERR_BAD_PARAM;
    int rc = PB2_SUCCESS;
    if (tp->holding) {
        rc = pb2_device_manage(tp->dev);
        tp->holding = false;
    }
    if (nb_executed) *nb_executed = tp->completed;
    if (rc != PB2_SUCCESS || tp->status || tp->completed != tp->inserted) return PB2_ERROR;
    return PB2_SUCCESS;
}

extern "C" void pb2_dtd_taskpool_free(pb2_dtd_taskpool_t* tp) {
    if (!tp) return;
    if (tp->holding) pb2_device_manage(tp->dev);
    for (DtdTask* t : tp->tasks) delete t;
    delete tp;
}

// The GPU body of dtd_test_simple_gemm.c:165-216 (gemm_kernel_cuda): flows (A, B, C), locals (m, n, k, mb, nb, kb);
// cublasDgemm_v2(N, N, mb, nb, kb, 1.0, a, mb, b, kb, 1.0, c, mb) becomes the sm_100a tile kernel.
extern "C" int pb2_dtd_gemm_body(pb2_device_t* dev, pb2_gpu_task_t* t, void* stream) {
    (void)dev;
    const int mb = (int)t->locals[3], nb = (int)t->locals[4], kb = (int)t->locals[5];
    int rc = parsec_b200_dgemm('N', 'N', mb, nb, kb, 1.0, (const double*)t->flow_dev_ptr[0], mb,
                               (const double*)t->flow_dev_ptr[1], kb, 1.0, (double*)t->flow_dev_ptr[2], mb, stream);
    return rc == PB2_SUCCESS ? PB2_HOOK_RETURN_DONE : PB2_HOOK_RETURN_ERROR;
}
