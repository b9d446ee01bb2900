// The CUDA device module: stream manager, stage-in -> kernel -> stage-out pipeline, HBM tile cache
// (zone heap + two LRU lists) with data versioning.  Replaces parsec/mca/device/cuda/device_cuda_module.c.
//
// What is kept from the reference (it is the contract the generated code and the graded version
// sequences rely on): the entry points, the manager-election protocol of parsec_cuda_kernel_scheduler
// (:2573-2589, :2742-2753), and the order and content of every data-state transition
// (reserve :863-1168, stage-in :1260-1530, push completion :1765-1942, pop :2196-2359, epilog :2364-2453,
// W2R eviction transfer_gpu.c:221-338).
//
// What is different (B200-first):
//   * the three stages of a task are chained ON THE DEVICE with events (h2d stream -> exec stream ->
//     d2h stream) the moment HBM space is reserved; the host never sits between two stages.  The
//     reference polls each stage's event from the manager thread before issuing the next stage and
//     allows 4 tasks per stream; here hundreds of tasks are in flight and the copy engines run ahead.
//   * data-state bookkeeping of the push stage is applied eagerly in program order (device ordering is
//     guaranteed by the events), so a second reader of a tile that is still in flight simply waits on
//     the tile's ready event on its exec stream instead of being descheduled (PARSEC_HOOK_RETURN_AGAIN).
//     The resulting per-tile state sequence is exactly the sequential one of the oracle.
//   * W2R eviction really writes dirty tiles back (the reference's stage_out == NULL defect, SURVEY.md
//     Appendix B-1, is not reproduced).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <thread>
#include <unordered_set>

#include "pb2_internal.h"
#include "pb2_types.h"

#define PB2_CUDA_OK(call)                                                                              \
    do {                                                                                               \
        cudaError_t e__ = (call);                                                                      \
        if (e__ != cudaSuccess) {                                                                      \
            fprintf(stderr, "[parsec_b200] CUDA error %s at %s:%d\n", cudaGetErrorString(e__), __FILE__, \
                    __LINE__);                                                                         \
            return PB2_ERR_DEVICE;                                                                     \
        }                                                                                              \
    } while (0)

enum { EV_RESERVE = 1, EV_STAGE_IN, EV_PUSH_DONE, EV_POP, EV_EPILOG, EV_COMPLETE, EV_W2R, EV_EVICT };

namespace pb2 {

struct TaskState {  // module-private part of a task in flight (pb2_gpu_task_t.reserved[0])
    cudaEvent_t done = nullptr;
    int exec_stream = -1;
    bool is_w2r = false;
    std::vector<pb2_copy_t*> w2r_copies;
    uint32_t pushed_mask = 0;  // flows for which THIS task scheduled the copy in (H2D, or D2D from a peer GPU)
    uint32_t staged_mask = 0;  // flows whose stage-in bookkeeping has been applied (what a cleanout has to undo)
    // state of each staged flow's device copy before this task touched it, for parsec_cuda_kernel_cleanout's job
    struct Before {
        uint32_t version;
        uint8_t coherency, host_coherency, list_id;
        int8_t owner;
    } before[PB2_MAX_PARAM_COUNT];
    // peer-GPU copies this task reads from (D2D stage-in): pinned (readers++) until the task retires
    pb2_copy_t* d2d_src[PB2_MAX_PARAM_COUNT] = {nullptr};
    uint8_t d2d_demote[PB2_MAX_PARAM_COUNT] = {0};  // the source gave its ownership away: invalidate it when unpinned
    bool host_source_only = false;  // no D2D source (a pin that outlives this record could not hand the reader back)
    uint64_t need_seq = 0;                // newest h2d event the body has to wait for
    std::vector<cudaEvent_t> extra_waits; // events of other streams (NCCL) that make inputs valid
    bool done_shared = false;             // `done` belongs to a later task of the same batch (it gives it back)
    void reset() {                        // back to the state of a fresh record, keeping the vectors' storage
        done = nullptr;
        exec_stream = -1;
        is_w2r = false;
        w2r_copies.clear();
        pushed_mask = staged_mask = 0;
        memset(d2d_src, 0, sizeof(d2d_src));
        memset(d2d_demote, 0, sizeof(d2d_demote));
        host_source_only = false;
        need_seq = 0;
        extra_waits.clear();
        done_shared = false;
    }
};

struct InFlight {
    pb2_gpu_task_t* task;
    TaskState* st;
};

constexpr int EVENT_RING = 1024;

}  // namespace pb2

struct pb2_device_s {
    int cuda_index = 0;
    int device_index = 2;
    pb2_device_params_t params;
    std::atomic<int32_t> mutex{0};  // tasks inside the module (parsec_device_gpu_module_t.mutex)
    std::mutex pending_lock;
    // parsec_device_gpu_module_t.pending.  A binary heap: highest task priority first, oldest first among equals
    // (the reference keeps its per-stream lists priority-ordered, PARSEC_GPU_USE_PRIORITIES, device_cuda_module.c:1683-1692);
    // with params.use_priorities == 0 every entry has priority 0 and the heap is a FIFO.
    struct Pending {
        pb2_gpu_task_t* task;
        int32_t priority;
        uint64_t seq;
        bool operator<(const Pending& o) const { return priority != o.priority ? priority < o.priority : seq > o.seq; }
    };
    std::vector<Pending> pending;
    uint64_t pending_seq = 0;
    std::deque<pb2_gpu_task_t*> deferred; // prefetches queued while nobody managed the device (not counted in mutex)
    // copies of THIS device that a peer GPU's task has finished reading (D2D stage-in): the peer's manager must not
    // touch this device's lists, so it leaves the copy here; this device's manager (or memory_release / destroy)
    // drops the reader and re-files the copy.  The job of the reference's D2D_COMPLETE command task (:1733-1760).
    struct PeerUnpin {
        pb2_copy_t* copy;
        bool demote;
    };
    std::mutex peer_lock;
    std::vector<PeerUnpin> peer_unpins;
    // everything below is touched by the manager only
    cudaStream_t h2d = nullptr, d2h = nullptr;
    std::vector<cudaStream_t> exec;
    std::vector<uint64_t> exec_waited_seq;
    int next_exec = 0;
    int active_exec = 0;  // exec streams the last task was spread over (see try_submit)
    std::atomic<uint64_t> h2d_seq{0};  // number of push events recorded so far (a peer GPU reads it to wait for a source copy)
    cudaEvent_t h2d_events[pb2::EVENT_RING] = {};
    std::vector<cudaEvent_t> event_pool;
    // in flight, one FIFO per stream that carries the final event: exec streams, then d2h, then h2d (prefetches).
    // Events of one stream complete in order, so only the heads are ever polled.
    std::vector<std::deque<pb2::InFlight>> inflight_q;
    size_t inflight_count = 0;
    std::deque<pb2_gpu_task_t*> stalled;      // tasks that found no room
    char* hbm_base = nullptr;
    pb2::ZoneHeap* heap = nullptr;
    pb2::CopyList lru{1}, owned_lru{2};
    pb2::CopyList* lists[3] = {nullptr, &lru, &owned_lru};
    pb2_device_stats_t stats;
    std::vector<pb2_event_t> events;
    bool failed = false;
    uint32_t peer_access_mask = 0;  // bit i: CUDA device i is directly reachable (parsec_device_gpu_module_t.peer_access_mask)
    struct Workspace {  // parsec_gpu_workspace_t (device_gpu.h:155-159), one per exec stream
        void* buf[PB2_GPU_MAX_WORKSPACE] = {nullptr};
        int stack_head = -1;
        size_t size = 0;
        bool ready = false;
    };
    std::vector<Workspace> workspace;
    std::unordered_map<void*, size_t> registered;
    pb2_device_hooks_t hooks = {nullptr, nullptr};
    std::unordered_set<pb2_copy_t*> all_copies;  // every device copy this module has created and not released yet
    std::vector<pb2::TaskState*> state_pool;     // recycled per-task records (manager thread only)
    std::vector<std::vector<cudaEvent_t>> launch_marks;  // profile_launches: completion marks per exec stream
};

static std::mutex g_devices_lock;
static pb2_device_s* g_devices[PB2_MAX_DEVICES] = {nullptr};

using namespace pb2;

static void drain_peer_unpins(pb2_device_s* d);
static void complete_w2r(pb2_device_s* d, TaskState* st);

extern "C" void pb2_device_params_default(pb2_device_params_t* p) {
    p->memory_use_percent = 95;
    p->memory_block_size = 512 * 1024;
    p->memory_number_of_blocks = -1;
    p->max_streams = 18;  // h2d + d2h + up to 16 exec (the reference's default is 6; small tiles need the concurrency)
    p->max_number_of_ejected_data = PB2_MAX_PARAM_COUNT;
    p->max_inflight_tasks = 256;
    p->trace_events = 0;
    p->profile_launches = 0;
    p->sort_pending_tasks = 0;
    p->prefetch_deferred = 0;
    p->use_priorities = 0;
    p->batch_bodies = 0;
}

// The reference's MCA parameters reach a process as environment variables PARSEC_MCA_<name> (parsec/utils/mca_param.c;
// registered for this module in device_cuda_component.c:163-206).  Inside a PaRSEC build the overlay shim hands over
// the values the runtime parsed; a host that drives the library directly gets the same names through this function.
static bool env_int64(const char* name, int64_t* out) {
    const char* v = getenv(name);
    if (!v || !*v) return false;
    char* end = nullptr;
    long long x = strtoll(v, &end, 0);
    if (end == v) return false;
    *out = (int64_t)x;
    return true;
}

extern "C" int pb2_device_params_from_env(pb2_device_params_t* p) {
    if (!p) return PB2_ERR_BAD_PARAM;
    int n = 0;
    int64_t v;
    if (env_int64("PARSEC_MCA_device_cuda_memory_use", &v) && v > 0 && v <= 100) { p->memory_use_percent = (int)v; ++n; }
    if (env_int64("PARSEC_MCA_device_cuda_memory_block_size", &v) && v > 0) { p->memory_block_size = v; ++n; }
    if (env_int64("PARSEC_MCA_device_cuda_memory_number_of_blocks", &v)) { p->memory_number_of_blocks = v; ++n; }
    if (env_int64("PARSEC_MCA_device_cuda_max_streams", &v) && v >= 3) { p->max_streams = (int)v; ++n; }
    if (env_int64("PARSEC_MCA_device_cuda_max_number_of_ejected_data", &v) && v > 0) { p->max_number_of_ejected_data = (int)v; ++n; }
    if (env_int64("PARSEC_MCA_device_cuda_sort_pending_tasks", &v)) { p->sort_pending_tasks = v != 0; ++n; }
    return n;
}

// device_cuda_enabled (-1 = every GPU, n = the first n) and device_cuda_mask (bit i = GPU i may be used),
// device_cuda_component.c:165-172, :107.
extern "C" int pb2_device_is_enabled(int cuda_index) {
    int64_t enabled = -1, mask = -1;
    env_int64("PARSEC_MCA_device_cuda_enabled", &enabled);
    env_int64("PARSEC_MCA_device_cuda_mask", &mask);
    if (cuda_index < 0 || cuda_index >= 63) return 0;
    if (enabled >= 0 && cuda_index >= enabled) return 0;
    return (mask >> cuda_index) & 1 ? 1 : 0;
}

static cudaEvent_t get_event(pb2_device_s* d) {
    if (!d->event_pool.empty()) {
        cudaEvent_t e = d->event_pool.back();
        d->event_pool.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    return e;
}

static TaskState* new_state(pb2_device_s* d) {
    if (d->state_pool.empty()) return new TaskState();
    TaskState* st = d->state_pool.back();
    d->state_pool.pop_back();
    return st;
}
static void free_state(pb2_device_s* d, TaskState* st) {
    if (!st) return;
    if (d->state_pool.size() >= 1024) {
        delete st;
        return;
    }
    st->reset();
    d->state_pool.push_back(st);
}

static void trace(pb2_device_s* d, pb2_gpu_task_t* t, int f, int kind, int transfer) {
    if (!d->params.trace_events) return;
    pb2_data_s* data = t->flow_data[f];
    pb2_copy_t* g = data->device_copies[d->device_index];
    pb2_event_t e;
    memset(&e, 0, sizeof(e));
    e.dc = (uint32_t)t->trace_dc[f];
    e.key = (uint32_t)data->key;
    e.kind = kind;
    e.k = (int32_t)t->locals[2];
    e.gpu_version = g ? g->version : 0;
    e.cpu_version = data->host_copy.version;
    e.gpu_coh = g ? g->coherency_state : 0xff;
    e.cpu_coh = data->host_copy.coherency_state;
    e.owner_device = data->owner_device;
    e.transfer = (int8_t)transfer;
    d->events.push_back(e);
}

extern "C" int pb2_device_create(int cuda_index, const pb2_device_params_t* params, pb2_device_t** out) {
    if (!out) return PB2_ERR_BAD_PARAM;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || cuda_index < 0 || cuda_index >= count) {
        fprintf(stderr, "[parsec_b200] no CUDA device %d (found %d): the device module has no CPU fallback\n",
                cuda_index, count);
        return PB2_ERR_DEVICE;
    }
    if (!params && !pb2_device_is_enabled(cuda_index)) {
        fprintf(stderr, "[parsec_b200] CUDA device %d is disabled by PARSEC_MCA_device_cuda_enabled / _mask\n", cuda_index);
        return PB2_ERR_NOT_SUPPORTED;
    }
    PB2_CUDA_OK(cudaSetDevice(cuda_index));
    pb2_device_s* d = new pb2_device_s();
    d->cuda_index = cuda_index;
    if (params) {
        d->params = *params;
    } else {  // no struct from the caller: the defaults, then the reference's PARSEC_MCA_device_cuda_* environment
        pb2_device_params_default(&d->params);
        pb2_device_params_from_env(&d->params);
    }
    memset(&d->stats, 0, sizeof(d->stats));
    {
        std::lock_guard<std::mutex> g(g_devices_lock);
        int idx = -1;
        for (int i = 2; i < PB2_MAX_DEVICES; ++i)
            if (!g_devices[i]) { idx = i; break; }
        if (idx < 0) { delete d; return PB2_ERR_OUT_OF_RESOURCE; }
        d->device_index = idx;
        g_devices[idx] = d;
    }
    // from here on a failure tears down whatever exists (streams, events, the slot in g_devices)
    auto fail = [d](cudaError_t e, int line) {
        fprintf(stderr, "[parsec_b200] CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, line);
        pb2_device_destroy(d);
        return PB2_ERR_DEVICE;
    };
    cudaError_t ce;
    int nexec = std::max(1, d->params.max_streams - 2);
    d->exec.assign(nexec, nullptr);
    d->exec_waited_seq.assign(nexec, 0);
    d->launch_marks.resize(nexec);
    d->inflight_q.resize(nexec + 2);
    if ((ce = cudaStreamCreateWithFlags(&d->h2d, cudaStreamNonBlocking)) != cudaSuccess) return fail(ce, __LINE__);
    if ((ce = cudaStreamCreateWithFlags(&d->d2h, cudaStreamNonBlocking)) != cudaSuccess) return fail(ce, __LINE__);
    for (int i = 0; i < nexec; ++i)
        if ((ce = cudaStreamCreateWithFlags(&d->exec[i], cudaStreamNonBlocking)) != cudaSuccess) return fail(ce, __LINE__);
    for (int i = 0; i < EVENT_RING; ++i)
        if ((ce = cudaEventCreateWithFlags(&d->h2d_events[i], cudaEventDisableTiming)) != cudaSuccess) return fail(ce, __LINE__);

    // parsec_cuda_memory_reserve (:612-743): one allocation, carved in memory_block_size units
    size_t unit = (size_t)d->params.memory_block_size;
    size_t free_b = 0, total_b = 0;
    if (unit == 0) {
        fprintf(stderr, "[parsec_b200] memory_block_size must be positive\n");
        pb2_device_destroy(d);
        return PB2_ERR_BAD_PARAM;
    }
    if ((ce = cudaMemGetInfo(&free_b, &total_b)) != cudaSuccess) return fail(ce, __LINE__);
    size_t want;
    if (d->params.memory_number_of_blocks > 0) want = (size_t)d->params.memory_number_of_blocks * unit;
    else want = (size_t)((double)free_b * d->params.memory_use_percent / 100.0);
    size_t nblocks = want / unit;
    while (nblocks > 0) {
        if (cudaMalloc((void**)&d->hbm_base, nblocks * unit) == cudaSuccess) break;
        cudaGetLastError();
        d->hbm_base = nullptr;
        nblocks -= std::max<size_t>(1, nblocks / 64);
    }
    if (!nblocks || !d->hbm_base) {
        fprintf(stderr, "[parsec_b200] cannot reserve device memory\n");
        pb2_device_destroy(d);
        return PB2_ERR_OUT_OF_RESOURCE;
    }
    d->heap = new ZoneHeap(d->hbm_base, (int)nblocks, unit);
    d->stats.hbm_bytes_total = nblocks * unit;
    // direct access between the GPUs of this process, both ways (device_cuda_component.c:120-151): a tile that sits on
    // a peer is staged over NVLink instead of PCIe.  Without it cudaMemcpyPeerAsync still works, through host memory.
    {
        std::lock_guard<std::mutex> g(g_devices_lock);
        for (int i = 2; i < PB2_MAX_DEVICES; ++i) {
            pb2_device_s* o = g_devices[i];
            if (!o || o == d || o->cuda_index == d->cuda_index) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, d->cuda_index, o->cuda_index) == cudaSuccess && can) {
                cudaSetDevice(d->cuda_index);
                if (cudaDeviceEnablePeerAccess(o->cuda_index, 0) != cudaSuccess) cudaGetLastError();
                d->peer_access_mask |= 1u << o->cuda_index;
            }
            can = 0;
            if (cudaDeviceCanAccessPeer(&can, o->cuda_index, d->cuda_index) == cudaSuccess && can) {
                cudaSetDevice(o->cuda_index);
                if (cudaDeviceEnablePeerAccess(d->cuda_index, 0) != cudaSuccess) cudaGetLastError();
                o->peer_access_mask |= 1u << d->cuda_index;
            }
        }
        cudaSetDevice(d->cuda_index);
    }
    *out = d;
    return PB2_SUCCESS;
}

static void release_copy(pb2_device_s* d, pb2_copy_t* c) {  // detach + give the memory back
    CopyList::chop(c, d->lists);
    d->all_copies.erase(c);
    pb2_data_s* original = c->original;
    if (original) original->device_copies[d->device_index] = nullptr;
    if ((c->flags & PB2_FLAG_OWNED) && c->device_private) d->heap->free(c->device_private);
    delete c;
    if (original && d->hooks.copy_released) d->hooks.copy_released(d, original, original->user);
}

extern "C" int pb2_device_set_hooks(pb2_device_t* d, const pb2_device_hooks_t* hooks) {
    if (!d) return PB2_ERR_BAD_PARAM;
    if (hooks) d->hooks = *hooks; else d->hooks = pb2_device_hooks_t{nullptr, nullptr};
    return PB2_SUCCESS;
}
extern "C" int pb2_device_nb_streams(const pb2_device_t* d) { return d ? (int)d->exec.size() + 2 : 0; }
extern "C" void* pb2_device_stream(const pb2_device_t* d, int index) {
    if (!d || index < 0 || index >= (int)d->exec.size() + 2) return nullptr;
    return index == 0 ? (void*)d->h2d : index == 1 ? (void*)d->d2h : (void*)d->exec[index - 2];
}
extern "C" int pb2_device_stream_index(const pb2_device_t* d, const void* cuda_stream) {
    if (!d) return -1;
    if (cuda_stream == (const void*)d->h2d) return 0;
    if (cuda_stream == (const void*)d->d2h) return 1;
    for (size_t i = 0; i < d->exec.size(); ++i)
        if (cuda_stream == (const void*)d->exec[i]) return (int)i + 2;
    return -1;
}

extern "C" int pb2_device_memory_release(pb2_device_t* d) {
    if (!d) return PB2_ERR_BAD_PARAM;
    if (d->mutex.load() != 0) return PB2_ERROR;  // a manager is active
    drain_peer_unpins(d);
    std::vector<pb2_copy_t*> busy;
    while (pb2_copy_t* c = d->lru.pop_front()) {
        if (c->readers != 0) {  // pinned by a reader (e.g. the comm layer): keep it
            busy.push_back(c);
            continue;
        }
        release_copy(d, c);
    }
    for (pb2_copy_t* c : busy) d->lru.push_back(c);
    return PB2_SUCCESS;
}

extern "C" int pb2_device_destroy(pb2_device_t* d) {
    if (!d) return PB2_ERR_BAD_PARAM;
    cudaSetDevice(d->cuda_index);
    cudaDeviceSynchronize();
    // whatever is still queued is abandoned: write-backs have completed (the device is idle), tasks never ran
    for (auto& queue : d->inflight_q)
        for (auto& f : queue) {
            if (f.st && f.st->is_w2r) complete_w2r(d, f.st);
            delete f.st;
        }
    for (TaskState* st : d->state_pool) delete st;
    d->state_pool.clear();
    for (pb2_gpu_task_t* t : d->deferred) free(t);
    drain_peer_unpins(d);
    // every copy, also the ones in no list (dropped from the LRU while they had readers, or chopped out for a writer)
    while (!d->all_copies.empty()) release_copy(d, *d->all_copies.begin());
    pb2_device_free_workspace(d);
    for (auto& kv : d->registered) cudaHostUnregister(kv.first);
    for (cudaStream_t s : d->exec)
        if (s) cudaStreamDestroy(s);
    if (d->h2d) cudaStreamDestroy(d->h2d);
    if (d->d2h) cudaStreamDestroy(d->d2h);
    for (int i = 0; i < EVENT_RING; ++i)
        if (d->h2d_events[i]) cudaEventDestroy(d->h2d_events[i]);
    for (cudaEvent_t e : d->event_pool) cudaEventDestroy(e);
    if (d->hbm_base) cudaFree(d->hbm_base);
    delete d->heap;
    {
        std::lock_guard<std::mutex> g(g_devices_lock);
        g_devices[d->device_index] = nullptr;
    }
    delete d;
    return PB2_SUCCESS;
}

extern "C" int pb2_device_index(const pb2_device_t* d) { return d ? d->device_index : PB2_ERR_BAD_PARAM; }
extern "C" int pb2_device_cuda_index(const pb2_device_t* d) { return d ? d->cuda_index : PB2_ERR_BAD_PARAM; }

extern "C" void pb2_data_destroy(pb2_data_t* data) {
    if (!data) return;
    for (int i = 2; i < PB2_MAX_DEVICES; ++i) {
        pb2_copy_t* c = data->device_copies[i];
        if (!c) continue;
        pb2_device_s* d = g_devices[i];
        if (d) release_copy(d, c);
    }
    delete data;
}

extern "C" int pb2_device_memory_register(pb2_device_t* d, void* ptr, size_t len) {
    if (!d || !ptr || !len) return PB2_ERR_BAD_PARAM;
    if (d->registered.count(ptr)) return PB2_SUCCESS;
    cudaSetDevice(d->cuda_index);
    cudaError_t e = cudaHostRegister(ptr, len, cudaHostRegisterPortable);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        cudaGetLastError();
        return PB2_SUCCESS;
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        return PB2_ERROR;
    }
    d->registered[ptr] = len;
    return PB2_SUCCESS;
}

extern "C" int pb2_device_memory_unregister(pb2_device_t* d, void* ptr) {
    if (!d || !ptr) return PB2_ERR_BAD_PARAM;
    auto it = d->registered.find(ptr);
    if (it == d->registered.end()) return PB2_SUCCESS;
    cudaHostUnregister(ptr);
    d->registered.erase(it);
    return PB2_SUCCESS;
}

extern "C" int pb2_device_get_stats(pb2_device_t* d, pb2_device_stats_t* out, int reset) {
    if (!d || !out) return PB2_ERR_BAD_PARAM;
    d->stats.hbm_bytes_in_use = d->heap ? d->heap->in_use() : 0;
    *out = d->stats;
    if (reset) {
        uint64_t total = d->stats.hbm_bytes_total;
        memset(&d->stats, 0, sizeof(d->stats));
        d->stats.hbm_bytes_total = total;
    }
    return PB2_SUCCESS;
}

extern "C" const pb2_device_stats_t* pb2_device_stats_live(const pb2_device_t* d) { return d ? &d->stats : nullptr; }

extern "C" int64_t pb2_device_get_events(pb2_device_t* d, pb2_event_t* out, int64_t max_events) {
    if (!d) return PB2_ERR_BAD_PARAM;
    int64_t n = (int64_t)d->events.size();
    if (out) memcpy(out, d->events.data(), sizeof(pb2_event_t) * (size_t)std::min(n, max_events));
    return n;
}

extern "C" int64_t pb2_device_get_launch_ms(pb2_device_t* d, float* out, int64_t max_samples, int* nb_exec_streams) {
    if (!d) return PB2_ERR_BAD_PARAM;
    if (nb_exec_streams) *nb_exec_streams = d->active_exec > 0 ? d->active_exec : (int)d->exec.size();
    cudaSetDevice(d->cuda_index);
    int64_t n = 0;
    for (auto& marks : d->launch_marks) {
        for (size_t i = 1; i < marks.size(); ++i) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, marks[i - 1], marks[i]) == cudaSuccess && out && n < max_samples) out[n++] = ms;
        }
        for (cudaEvent_t e : marks) cudaEventDestroy(e);
        marks.clear();
    }
    return n;
}

// ------------------------------------------------------------------------------------ default staging
// One tile between its host copy and a packed device tile.  A tile of a PARSEC_MATRIX_LAPACK collection is a window of
// a column-major local matrix (pb2_data_set_host_layout): it goes through the copy engine's pitched 2D path, what the
// reference leaves to user stage functions (stage_custom.jdf:28-109).
static cudaError_t copy_tile(const pb2_data_s* data, void* dev, void* host, size_t bytes, bool to_device, cudaStream_t s) {
    if (data && data->host_pitch) {
        return to_device ? cudaMemcpy2DAsync(dev, data->dev_pitch, host, data->host_pitch, data->host_width,
                                             data->host_height, cudaMemcpyHostToDevice, s)
                         : cudaMemcpy2DAsync(host, data->host_pitch, dev, data->dev_pitch, data->host_width,
                                             data->host_height, cudaMemcpyDeviceToHost, s);
    }
    return to_device ? cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, s)
                     : cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, s);
}

extern "C" int pb2_default_stage_in(pb2_gpu_task_t* t, uint32_t flow_mask, void* stream) {
    for (int i = 0; i < t->nb_flows; ++i) {
        if (!(flow_mask & (1u << i))) continue;
        if (copy_tile(t->flow_data[i], t->flow_dev_ptr[i], t->flow_host_ptr[i], t->flow_nb_elts[i], true,
                      (cudaStream_t)stream) != cudaSuccess)
            return PB2_HOOK_RETURN_ERROR;
    }
    return PB2_SUCCESS;
}
extern "C" int pb2_default_stage_out(pb2_gpu_task_t* t, uint32_t flow_mask, void* stream) {
    for (int i = 0; i < t->nb_flows; ++i) {
        if (!(flow_mask & (1u << i))) continue;
        if (copy_tile(t->flow_data[i], t->flow_dev_ptr[i], t->flow_host_ptr[i], t->flow_nb_elts[i], false,
                      (cudaStream_t)stream) != cudaSuccess)
            return PB2_HOOK_RETURN_ERROR;
    }
    return PB2_SUCCESS;
}

// ------------------------------------------------------------------------------------ reserve (a4)
// A CUDA call failed on the manager thread: the device is unusable from here on (the reference returns
// PARSEC_HOOK_RETURN_DISABLE and gives up, device_cuda_module.c:2757-2762).
#define PB2_DEV_TRY(d, call)                                                                                      \
    do {                                                                                                          \
        cudaError_t e__ = (call);                                                                                 \
        if (e__ != cudaSuccess) {                                                                                 \
            fprintf(stderr, "[parsec_b200] CUDA error %s at %s:%d (%s)\n", cudaGetErrorString(e__), __FILE__,     \
                    __LINE__, #call);                                                                             \
            (d)->failed = true;                                                                                   \
            return PB2_HOOK_RETURN_DISABLE;                                                                       \
        }                                                                                                         \
    } while (0)

// Readers that a peer GPU's manager has handed back (see pb2_device_s::peer_unpins).  Manager thread, or any thread
// while no manager runs.
static void drain_peer_unpins(pb2_device_s* d) {
    std::vector<pb2_device_s::PeerUnpin> todo;
    {
        std::lock_guard<std::mutex> g(d->peer_lock);
        todo.swap(d->peer_unpins);
    }
    for (const auto& u : todo) {
        pb2_copy_t* c = u.copy;
        pb2_data_s* original = c->original;
        std::unique_lock<std::mutex> lk;
        if (original) lk = std::unique_lock<std::mutex>(original->lock);
        c->readers--;
        if (u.demote) {  // a peer took the tile over in write mode: what is left here is about to be outdated
            c->coherency_state = PB2_DATA_COHERENCY_INVALID;
            if (c->readers == 0) {
                CopyList::chop(c, d->lists);
                d->lru.push_front(c);  // first candidate for eviction
            }
        } else if (c->readers == 0 && c->list_id == 0 && c->coherency_state != PB2_DATA_COHERENCY_OWNED) {
            d->lru.push_back(c);
        }
    }
}

// Returns 0 = every flow has a device copy, 1 = no room right now, 2 = a needed copy is being written back.
// zone_malloc with eviction from the clean LRU (:934-1113).  `t` (may be null) is the task being served: its own
// tiles are never stolen (:1041-1062).  Returns nullptr when nothing evictable is left.
static void* alloc_evicting(pb2_device_s* d, size_t bytes, pb2_gpu_task_t* t) {
    void* mem = nullptr;
    while (!(mem = d->heap->alloc(bytes))) {
        pb2_copy_t* victim = nullptr;
        std::vector<pb2_copy_t*> skipped;
        while (pb2_copy_t* cand = d->lru.pop_front()) {
            bool mine = false;
            if (t)
                for (int j = 0; j < t->nb_flows; ++j)
                    if (t->flow_data[j] && cand->original == t->flow_data[j]) mine = true;
            if (mine) {
                skipped.push_back(cand);
                continue;
            }
            // readers are checked, and the copy detached, under the tile's lock: a peer GPU that wants this copy as
            // the source of a D2D stage-in pins it under the same lock, so it either sees the copy pinned or not at all
            std::unique_lock<std::mutex> lk;
            if (cand->original) lk = std::unique_lock<std::mutex>(cand->original->lock);
            if (cand->readers != 0) continue;  // dropped from the list; whoever takes the last reader away re-files it
            if (cand->original) cand->original->device_copies[d->device_index] = nullptr;
            victim = cand;
            break;
        }
        for (auto it = skipped.rbegin(); it != skipped.rend(); ++it) d->lru.push_front(*it);
        if (!victim) return nullptr;
        if (d->params.trace_events && victim->original) {
            pb2_event_t e;
            memset(&e, 0, sizeof(e));
            e.dc = victim->original->dc_id;
            e.key = (uint32_t)victim->original->key;
            e.kind = EV_EVICT;
            e.k = -1;
            e.gpu_version = victim->version;
            e.cpu_version = victim->original->host_copy.version;
            e.gpu_coh = victim->coherency_state;
            e.cpu_coh = victim->original->host_copy.coherency_state;
            e.owner_device = victim->original->owner_device;
            d->events.push_back(e);
        }
        release_copy(d, victim);
        d->stats.nb_evictions++;
    }
    return mem;
}

static int reserve_device_space(pb2_device_s* d, pb2_gpu_task_t* t) {
    pb2_copy_t* fresh[PB2_MAX_PARAM_COUNT] = {nullptr};
    for (int i = 0; i < t->nb_flows; ++i) {
        if (!(t->flow_access[i] & PB2_FLOW_ACCESS_RW) || !t->flow_data[i]) continue;
        pb2_data_s* master = t->flow_data[i];
        {
            std::lock_guard<std::mutex> g(master->lock);
            pb2_copy_t* c = master->device_copies[d->device_index];
            if (c) {
                if (c->data_transfer_status == PB2_XFER_UNDER) {
                    // only a W2R write-back leaves a copy UNDER_TRANSFER here (pushes are booked eagerly)
                    return 2;
                }
                continue;
            }
        }
        void* mem = alloc_evicting(d, t->flow_nb_elts[i], t);
        if (!mem) {
            // no room: keep what we created reachable at the LRU head and give up for now (:940-966)
            for (int j = 0; j < i; ++j)
                if (fresh[j]) {
                    CopyList::chop(fresh[j], d->lists);
                    d->lru.push_front(fresh[j]);
                }
            return 1;
        }
        pb2_copy_t* c = new pb2_copy_t();
        c->flags = PB2_FLAG_OWNED | PB2_FLAG_MANAGED;
        c->device_private = mem;
        c->coherency_state = PB2_DATA_COHERENCY_INVALID;
        c->version = 0;
        c->device_index = (int8_t)d->device_index;
        c->original = master;
        {
            std::lock_guard<std::mutex> g(master->lock);
            master->device_copies[d->device_index] = c;
        }
        d->lru.push_back(c);
        d->all_copies.insert(c);
        fresh[i] = c;
        trace(d, t, i, EV_RESERVE, 0);
    }
    return 0;
}

// ------------------------------------------------------------------------------------ stage-in (a5, a6)
// Where the bytes come from (parsec_gpu_data_stage_in, :1260-1530).  `from` is what the ownership state machine
// answered: -1 = nothing to move, 0 = the host copy, >= 2 = the device that holds the valid copy.  A read-only flow
// whose tile also sits, at the version wanted, on another GPU of this process is taken from there rather than from
// the host (:1310-1378: NVLink instead of PCIe); a tile whose only valid copy is on another GPU has no other source.
// The source copy is pinned (readers++) under the tile's lock, which the caller holds.
static pb2_copy_t* pick_peer_source(pb2_device_s* d, pb2_data_s* original, int from, uint8_t type, pb2_device_s** src_dev) {
    *src_dev = nullptr;
    if (from >= 2 && from != d->device_index) {
        pb2_copy_t* c = original->device_copies[from];
        pb2_device_s* sd = g_devices[from];
        if (c && sd && c->coherency_state != PB2_DATA_COHERENCY_INVALID) {
            *src_dev = sd;
            return c;
        }
        return nullptr;
    }
    if (from == 0 && !(type & PB2_FLOW_ACCESS_WRITE)) {
        for (int p = 2; p < PB2_MAX_DEVICES; ++p) {
            pb2_copy_t* c = original->device_copies[p];
            if (p == d->device_index || !c || !g_devices[p]) continue;
            if (c->coherency_state == PB2_DATA_COHERENCY_INVALID || c->version != original->host_copy.version) continue;
            if (c->data_transfer_status != PB2_XFER_COMPLETE) continue;  // e.g. being written back: take the host copy
            *src_dev = g_devices[p];
            return c;
        }
    }
    return nullptr;
}

static int stage_in_flow(pb2_device_s* d, pb2_gpu_task_t* t, TaskState* st, int i, bool* issued) {
    pb2_data_s* original = t->flow_data[i];
    const uint8_t type = t->flow_access[i];
    std::lock_guard<std::mutex> g(original->lock);
    pb2_copy_t* gpu = original->device_copies[d->device_index];
    pb2_copy_t* in = &original->host_copy;
    t->flow_dev_ptr[i] = gpu->device_private;
    t->flow_host_ptr[i] = in->device_private;

    if ((type & PB2_FLOW_ACCESS_WRITE) && gpu->readers > 0 && !(gpu->readers == 1 && (type & PB2_FLOW_ACCESS_READ))) {
        fprintf(stderr, "[parsec_b200] write access to a tile with %d readers (missing CTL dependency)\n", gpu->readers);
        return PB2_HOOK_RETURN_ERROR;
    }
    TaskState::Before& b = st->before[i];
    b.version = gpu->version;
    b.coherency = gpu->coherency_state;
    b.host_coherency = in->coherency_state;
    b.list_id = (uint8_t)gpu->list_id;
    b.owner = original->owner_device;
    if (type & PB2_FLOW_ACCESS_WRITE) CopyList::chop(gpu, d->lists);  // written tiles leave the LRU lists until the task completes
    int from = data_start_transfer_ownership(original, d->device_index, type);
    st->staged_mask |= 1u << i;
    if ((type & PB2_FLOW_ACCESS_WRITE) && t->task_type != 1) gpu->version++;
    if (from == -1) {
        gpu->data_transfer_status = PB2_XFER_COMPLETE;
        data_end_transfer_ownership(original, d->device_index, type);
        trace(d, t, i, EV_STAGE_IN, 0);
        return 0;
    }
    d->stats.required_data_in += original->nb_elts;
    pb2_device_s* src_dev = nullptr;
    pb2_copy_t* src = (t->stage_in || st->host_source_only) ? nullptr : pick_peer_source(d, original, from, type, &src_dev);
    if (from >= 2 && from != d->device_index && !src) {
        fprintf(stderr, "[parsec_b200] the valid copy of a tile is on device %d, which cannot serve it\n", from);
        return PB2_HOOK_RETURN_ERROR;
    }
    if (src) {
        // the source's bytes are valid once its own device's stage-in (or the broadcast that delivered it) is done
        const uint64_t sseq = src->ready_seq, snow = src_dev->h2d_seq.load();
        if (sseq) {
            const uint64_t w = (snow - sseq >= (uint64_t)EVENT_RING) ? snow : sseq;
            PB2_DEV_TRY(d, cudaStreamWaitEvent(d->h2d, src_dev->h2d_events[w % EVENT_RING], 0));
        }
        if (src->ready_event) PB2_DEV_TRY(d, cudaStreamWaitEvent(d->h2d, src->ready_event, 0));
        PB2_DEV_TRY(d, cudaMemcpyPeerAsync(gpu->device_private, d->cuda_index, src->device_private, src_dev->cuda_index,
                                           t->flow_nb_elts[i], d->h2d));
        src->readers++;
        st->d2d_src[i] = src;
        // a writer takes the tile away from the peer: its copy is outdated as soon as this task has run
        st->d2d_demote[i] = (type & PB2_FLOW_ACCESS_WRITE) ? 1 : 0;
        // the version stays the host copy's, as when the reference stages from task->data[i].data_in (:1498): after a
        // completed writer the host copy carries version + 1 for bytes that only the owner GPU has
        d->stats.data_in_from_peers += t->flow_nb_elts[i];
        d->stats.data_in_from_device[src_dev->device_index] += t->flow_nb_elts[i];
    } else {
        pb2_stage_fn fn = t->stage_in ? t->stage_in : pb2_default_stage_in;
        if (fn(t, 1u << i, d->h2d) != PB2_SUCCESS) {
            d->failed = true;
            return PB2_HOOK_RETURN_DISABLE;
        }
        d->stats.data_in_from_host += t->flow_nb_elts[i];
        d->stats.data_in_from_device[0] += t->flow_nb_elts[i];
    }
    if (t->task_type == 0) d->stats.nb_data_faults += t->flow_nb_elts[i];
    gpu->version = in->version;
    gpu->data_transfer_status = PB2_XFER_UNDER;
    gpu->ready_seq = d->h2d_seq.load() + 1;  // the push event recorded right after this task's copies
    st->pushed_mask |= 1u << i;
    *issued = true;
    trace(d, t, i, EV_STAGE_IN, 1);
    return 1;
}

// push completion booked eagerly; the device-side ordering is carried by ready_seq (:1765-1942)
static void complete_push_flow(pb2_device_s* d, pb2_gpu_task_t* t, int i) {
    pb2_data_s* original = t->flow_data[i];
    std::lock_guard<std::mutex> g(original->lock);
    pb2_copy_t* gpu = original->device_copies[d->device_index];
    gpu->data_transfer_status = PB2_XFER_COMPLETE;
    data_end_transfer_ownership(original, d->device_index, t->flow_access[i]);
    trace(d, t, i, EV_PUSH_DONE, 0);
}

// hand the readers of D2D sources back to the devices that own those copies
static void release_peer_sources(pb2_device_s* d, TaskState* st, int nb_flows) {
    (void)d;
    for (int i = 0; i < nb_flows; ++i) {
        pb2_copy_t* src = st->d2d_src[i];
        if (!src) continue;
        st->d2d_src[i] = nullptr;
        pb2_device_s* sd = g_devices[src->device_index];
        if (!sd) continue;
        std::lock_guard<std::mutex> g(sd->peer_lock);
        sd->peer_unpins.push_back({src, st->d2d_demote[i] != 0});
    }
}

// ------------------------------------------------------------------------------------ cleanout (a10)
// parsec_cuda_kernel_cleanout (:2472-2526): the task will not run on this device after all (its body declined, or a
// flow could not be staged).  Everything stage-in did to the task's tiles is taken back: readers, the version bump
// of WRITE flows, coherency and ownership, the LRU list each copy sat in.  A flow whose bytes were already requested
// from the host keeps them: the copy is then a clean SHARED copy of the host version.  Dirty inputs (the previous
// links of a C chain) are written back first, synchronously: the next incarnation of the task reads host memory,
// and unlike the reference this module does not leave it stale.
static void cleanout_task(pb2_device_s* d, pb2_gpu_task_t* t, TaskState* st) {
    bool wrote_back = false;
    for (int i = 0; i < t->nb_flows; ++i) {
        if (!(st->staged_mask & (1u << i))) continue;
        pb2_data_s* original = t->flow_data[i];
        const uint8_t type = t->flow_access[i];
        std::lock_guard<std::mutex> g(original->lock);
        pb2_copy_t* gpu = original->device_copies[d->device_index];
        pb2_copy_t* cpu = &original->host_copy;
        const TaskState::Before& b = st->before[i];
        if (type & PB2_FLOW_ACCESS_READ) gpu->readers--;
        original->owner_device = b.owner;
        cpu->coherency_state = b.host_coherency;
        CopyList::chop(gpu, d->lists);
        if (st->pushed_mask & (1u << i)) {  // the copy in is already queued: keep the bytes, as a clean copy
            gpu->coherency_state = PB2_DATA_COHERENCY_SHARED;
            gpu->data_transfer_status = PB2_XFER_COMPLETE;
            if (gpu->readers == 0) d->lru.push_back(gpu);
            continue;
        }
        gpu->version = b.version;
        gpu->coherency_state = b.coherency;
        if (b.list_id == 2 && gpu->readers == 0) {
            // dirty: newer than the host copy.  Bring the host up to date so that another incarnation can read it.
            if (copy_tile(original, gpu->device_private, cpu->device_private, original->nb_elts, false, d->d2h) == cudaSuccess) {
                wrote_back = true;
                d->stats.data_out_to_host += original->nb_elts;
                gpu->coherency_state = PB2_DATA_COHERENCY_SHARED;
                cpu->coherency_state = PB2_DATA_COHERENCY_SHARED;
                d->lru.push_back(gpu);
            } else {
                d->failed = true;
                d->owned_lru.push_back(gpu);
            }
        } else if (gpu->readers == 0) {
            (b.list_id == 2 ? d->owned_lru : d->lru).push_back(gpu);
        }
    }
    release_peer_sources(d, st, t->nb_flows);
    if (wrote_back && cudaStreamSynchronize(d->d2h) != cudaSuccess) d->failed = true;
    d->stats.nb_cleanouts++;
}

// ------------------------------------------------------------------------------------ W2R (a11)
static bool start_w2r(pb2_device_s* d) {
    TaskState* st = new_state(d);
    st->is_w2r = true;
    int limit = std::max(1, d->params.max_number_of_ejected_data);
    for (pb2_copy_t* c = d->owned_lru.head; c && (int)st->w2r_copies.size() < limit;) {
        pb2_copy_t* next = c->next;
        if (c->readers == 0 && d->hooks.data_alive && c->original &&
            !d->hooks.data_alive(d, c->original, c->original->user)) {
            // the runtime destroyed this data while its last version was still on the device (a NEW tile nobody reads
            // back): its host memory is gone, there is nothing to write back to
            release_copy(d, c);
            d->stats.nb_evictions++;
            c = next;
            continue;
        }
        if (c->readers == 0) {
            std::lock_guard<std::mutex> lk(c->original->lock);
            if (c->readers == 0) {
                CopyList::chop(c, d->lists);
                c->readers++;
                c->data_transfer_status = PB2_XFER_UNDER;
                st->w2r_copies.push_back(c);
            }
        }
        c = next;
    }
    if (st->w2r_copies.empty()) {
        free_state(d, st);
        return false;
    }
    // the dirty bytes were produced on exec streams whose completion the host has already observed
    bool ok = true;
    for (pb2_copy_t* c : st->w2r_copies) {
        ok = ok && copy_tile(c->original, c->device_private, c->original->host_copy.device_private, c->original->nb_elts,
                             false, d->d2h) == cudaSuccess;
        d->stats.data_out_to_host += c->original->nb_elts;
    }
    st->done = get_event(d);
    ok = ok && st->done && cudaEventRecord(st->done, d->d2h) == cudaSuccess;
    if (!ok) {
        fprintf(stderr, "[parsec_b200] CUDA error while writing dirty tiles back: %s\n", cudaGetErrorString(cudaGetLastError()));
        d->failed = true;
    }
    d->inflight_q[d->exec.size()].push_back({nullptr, st});
    d->inflight_count++;
    d->stats.nb_w2r_tasks++;
    d->stats.nb_w2r_tiles += st->w2r_copies.size();
    return true;
}

static void complete_w2r(pb2_device_s* d, TaskState* st) {
    for (pb2_copy_t* g : st->w2r_copies) {
        pb2_data_s* original = g->original;
        std::lock_guard<std::mutex> lk(original->lock);
        g->readers--;
        g->data_transfer_status = PB2_XFER_COMPLETE;
        pb2_copy_t* cpu = &original->host_copy;
        if (cpu->version < g->version) {
            d->owned_lru.push_back(g);  // re-acquired meanwhile; still dirty
        } else {
            g->coherency_state = PB2_DATA_COHERENCY_SHARED;
            cpu->coherency_state = PB2_DATA_COHERENCY_SHARED;
            // The reference assigns cpu.version = gpu.version here (transfer_gpu.c:321), which rolls the host
            // version BACK by one (after a completed writer cpu.version == gpu.version + 1) and
            // makes a re-staged tile lose a version per eviction.  The host copy now really holds the bytes its
            // version promised, so the version is left alone: evicted chains end at KT like in-core ones.
            if (cpu->version < g->version) cpu->version = g->version;
            d->lru.push_back(g);
        }
        if (d->params.trace_events) {
            pb2_event_t e;
            memset(&e, 0, sizeof(e));
            e.dc = original->dc_id;
            e.key = (uint32_t)original->key;
            e.kind = EV_W2R;
            e.k = -1;
            e.gpu_version = g->version;
            e.cpu_version = cpu->version;
            e.gpu_coh = g->coherency_state;
            e.cpu_coh = cpu->coherency_state;
            e.owner_device = original->owner_device;
            e.transfer = 1;
            d->events.push_back(e);
        }
    }
}

// ------------------------------------------------------------------------------------ submit one task
// 0 = in flight, 1 = stalled (no room / write-back pending); negative = the task does not run here:
// PB2_HOOK_RETURN_DISABLE when a CUDA call failed (the device is unusable), any other negative value is the task's own
// (its body declined or a flow could not be staged) and leaves the device healthy.  In every negative case what
// stage-in did has been taken back (cleanout_task) and t->reserved[0] is clear.
// Phase 1: stage every flow in and book the push completion.  The events the body will have to wait for are left in
// st->need_seq / st->extra_waits.
static int submit_stage_in(pb2_device_s* d, pb2_gpu_task_t* t, TaskState* st) {
    bool issued = false;
    for (int i = 0; i < t->nb_flows; ++i) {
        if (!(t->flow_access[i] & PB2_FLOW_ACCESS_RW) || !t->flow_data[i]) continue;
        {   // a device copy the user allocated and attached (nvlink_wrapper.c:189-206): the module only hands its address
            // to the body -- no staging, no readers, no version, no LRU (device_cuda_module.c:2161-2163, :2215, :2399)
            pb2_copy_t* c = t->flow_data[i]->device_copies[d->device_index];
            if (c && !(c->flags & PB2_FLAG_OWNED)) {
                t->flow_dev_ptr[i] = c->device_private;
                t->flow_host_ptr[i] = t->flow_data[i]->host_copy.device_private;
                continue;
            }
        }
        int rc = stage_in_flow(d, t, st, i, &issued);
        if (rc < 0) return rc;
        pb2_copy_t* c = t->flow_data[i]->device_copies[d->device_index];
        st->need_seq = std::max(st->need_seq, c->ready_seq);
        if (c->ready_event) st->extra_waits.push_back(c->ready_event);
    }
    if (issued) {
        const uint64_t seq = d->h2d_seq.load() + 1;
        PB2_DEV_TRY(d, cudaEventRecord(d->h2d_events[seq % EVENT_RING], d->h2d));
        d->h2d_seq.store(seq);
    }
    for (int i = 0; i < t->nb_flows; ++i)
        if (st->pushed_mask & (1u << i)) complete_push_flow(d, t, i);
    return 0;
}

// How many exec streams to spread over: a tile kernel has about one CTA per 128 KiB of its written tile
// (128x128 fp64 or 128x256 fp32), and a B200 wants ~2 x 148 CTAs queued.  An NB=1024 fp64 task (64 CTAs) uses 5
// streams, an NB=512 one (16 CTAs) all 16, an NB=2048 one (256 CTAs) the minimum of 4.
static int pick_exec_stream(pb2_device_s* d, const pb2_gpu_task_t* t) {
    size_t wbytes = 0;
    for (int i = 0; i < t->nb_flows; ++i)
        if (t->flow_access[i] & PB2_FLOW_ACCESS_WRITE) wbytes = std::max<size_t>(wbytes, t->flow_nb_elts[i]);
    const int ctas = (int)std::max<size_t>(1, wbytes >> 17);
    const int active = std::min<int>((int)d->exec.size(), std::max(4, (2 * 148 + ctas - 1) / ctas));
    d->active_exec = active;
    const int s = d->next_exec % active;
    d->next_exec = (d->next_exec + 1) % active;
    return s;
}

// Phase 2: make exec stream s wait for the task's inputs and run its body there.
static int submit_body(pb2_device_s* d, pb2_gpu_task_t* t, TaskState* st, int s) {
    st->exec_stream = s;
    if (st->need_seq > d->exec_waited_seq[s]) {
        // waiting on a newer event of the same in-order stream covers every older one; if the ring has
        // wrapped past need_seq the slot holds a later event, which is still sufficient
        const uint64_t now = d->h2d_seq.load();
        uint64_t w = (now - st->need_seq >= (uint64_t)EVENT_RING) ? now : st->need_seq;
        PB2_DEV_TRY(d, cudaStreamWaitEvent(d->exec[s], d->h2d_events[w % EVENT_RING], 0));
        d->exec_waited_seq[s] = w;
    }
    for (cudaEvent_t e : st->extra_waits) PB2_DEV_TRY(d, cudaStreamWaitEvent(d->exec[s], e, 0));
    int rc = t->submit(d, t, d->exec[s]);
    if (rc != PB2_HOOK_RETURN_DONE) return rc < 0 && rc != PB2_HOOK_RETURN_DISABLE ? rc : PB2_HOOK_RETURN_ERROR;
    d->stats.kernel_launches++;
    return 0;
}

// Phase 3: the task's completion event, after its stage-out (a8) when it has one; file it as in flight.
// `shared` (batches): an event already recorded at the end of the batch on the task's exec stream; a task without
// stage-out takes it as its completion event instead of recording one of its own.
static int submit_finish(pb2_device_s* d, pb2_gpu_task_t* t, TaskState* st, cudaEvent_t shared = nullptr) {
    cudaStream_t last = st->exec_stream >= 0 ? d->exec[st->exec_stream] : d->h2d;
    if (st->exec_stream >= 0 && d->params.profile_launches && d->launch_marks[st->exec_stream].size() < 16384) {
        cudaEvent_t mark = nullptr;
        if (cudaEventCreate(&mark) == cudaSuccess && cudaEventRecord(mark, last) == cudaSuccess)
            d->launch_marks[st->exec_stream].push_back(mark);
    }
    uint32_t out_mask = 0;
    for (int i = 0; i < t->nb_flows; ++i)
        if ((t->flow_access[i] & PB2_FLOW_ACCESS_WRITE) && (t->pushout & (1 << i)) && t->flow_data[i]) {
            const pb2_copy_t* c = t->flow_data[i]->device_copies[d->device_index];
            if (c && (c->flags & PB2_FLAG_OWNED)) out_mask |= 1u << i;  // a user's device copy is never staged out (:2215)
        }
    if (shared && !out_mask) {
        st->done = shared;
        st->done_shared = true;
        d->inflight_q[(size_t)st->exec_stream].push_back({t, st});
        d->inflight_count++;
        return 0;
    }
    st->done = get_event(d);
    if (!st->done) {
        d->failed = true;
        return PB2_HOOK_RETURN_DISABLE;
    }
    if (out_mask) {
        PB2_DEV_TRY(d, cudaEventRecord(st->done, last));
        PB2_DEV_TRY(d, cudaStreamWaitEvent(d->d2h, st->done, 0));
        pb2_stage_fn fn = t->stage_out ? t->stage_out : pb2_default_stage_out;
        if (fn(t, out_mask, d->d2h) != PB2_SUCCESS) {
            d->failed = true;
            return PB2_HOOK_RETURN_DISABLE;
        }
        PB2_DEV_TRY(d, cudaEventRecord(st->done, d->d2h));
    } else {
        PB2_DEV_TRY(d, cudaEventRecord(st->done, last));
    }
    const size_t q = out_mask ? d->exec.size() : (st->exec_stream >= 0 ? (size_t)st->exec_stream : d->exec.size() + 1);
    d->inflight_q[q].push_back({t, st});
    d->inflight_count++;
    return 0;
}

static void abandon_submission(pb2_device_s* d, pb2_gpu_task_t* t, TaskState* st, int rc) {
    // if the body was already launched the kernel may be running: let the streams drain before the tiles go back
    if (rc == PB2_HOOK_RETURN_DISABLE) cudaDeviceSynchronize();
    cleanout_task(d, t, st);
    if (st->done && !st->done_shared) d->event_pool.push_back(st->done);
    free_state(d, st);
    t->reserved[0] = nullptr;
}

static int try_submit(pb2_device_s* d, pb2_gpu_task_t* t) {
    int r = reserve_device_space(d, t);
    if (r != 0) return 1;
    TaskState* st = new_state(d);
    t->reserved[0] = st;
    int rc = submit_stage_in(d, t, st);
    if (rc == 0 && t->task_type == 0 && t->submit) rc = submit_body(d, t, st, pick_exec_stream(d, t));
    if (rc == 0) rc = submit_finish(d, t, st);
    if (rc < 0) abandon_submission(d, t, st, rc);
    return rc;
}

// The second half of a batched submission (params.batch_bodies > 1): the tasks of `batch` have been reserved and
// staged in (try_stage); their bodies now run back to back on ONE exec stream between parsec_b200_batch_begin and
// _end, so that bodies which call the library's batchable kernels end up in one grouped launch.  Every task still
// gets its own completion event (they all fire when the stream reaches the end of the batch) and its own stage-out.
// Tasks whose body declines are taken back one by one; `failed` receives (task, status) for the caller to complete.
extern "C" int parsec_b200_batch_begin(void* stream);
extern "C" int parsec_b200_batch_end(void* stream);
static void submit_batch(pb2_device_s* d, std::vector<pb2_gpu_task_t*>& batch,
                         std::vector<std::pair<pb2_gpu_task_t*, int>>& failed) {
    if (batch.empty()) return;
    const int s = pick_exec_stream(d, batch.front());
    std::vector<pb2_gpu_task_t*> ran;
    const bool bracket = parsec_b200_batch_begin(d->exec[s]) == PB2_SUCCESS;
    for (pb2_gpu_task_t* t : batch) {
        TaskState* st = (TaskState*)t->reserved[0];
        int rc = (t->task_type == 0 && t->submit) ? submit_body(d, t, st, s) : 0;
        if (rc < 0) {
            abandon_submission(d, t, st, rc);
            failed.push_back({t, rc});
        } else {
            ran.push_back(t);
        }
    }
    if (bracket && parsec_b200_batch_end(d->exec[s]) != PB2_SUCCESS) d->failed = true;
    // one completion event for the whole batch (the tasks end together, when the stream gets here); the last task that
    // uses it owns it and, retiring last from the same in-order queue, gives it back
    cudaEvent_t shared = nullptr;
    if (!d->failed && ran.size() > 1 && !d->params.profile_launches) {
        shared = get_event(d);
        if (shared && cudaEventRecord(shared, d->exec[s]) != cudaSuccess) {
            d->event_pool.push_back(shared);
            shared = nullptr;
            cudaGetLastError();
        }
    }
    TaskState* owner = nullptr;
    for (pb2_gpu_task_t* t : ran) {
        TaskState* st = (TaskState*)t->reserved[0];
        int rc = d->failed ? PB2_HOOK_RETURN_DISABLE : submit_finish(d, t, st, shared);
        if (rc < 0) {
            abandon_submission(d, t, st, rc);
            failed.push_back({t, rc});
        } else if (st->done_shared) {
            owner = st;
        }
    }
    if (owner) owner->done_shared = false;
    else if (shared) d->event_pool.push_back(shared);
    batch.clear();
}

// first half: 0 = staged and waiting in the batch, 1 = stalled, < 0 = does not run here (already taken back)
static int try_stage(pb2_device_s* d, pb2_gpu_task_t* t) {
    int r = reserve_device_space(d, t);
    if (r != 0) return 1;
    TaskState* st = new_state(d);
    t->reserved[0] = st;
    int rc = submit_stage_in(d, t, st);
    if (rc < 0) abandon_submission(d, t, st, rc);
    return rc;
}

// pop (:2196-2359) + epilog (:2364-2453) bookkeeping once the device work is really done
static void finish_task(pb2_device_s* d, pb2_gpu_task_t* t, TaskState* st) {
    for (int i = 0; i < t->nb_flows; ++i) {
        const uint8_t type = t->flow_access[i];
        if (!(type & PB2_FLOW_ACCESS_RW) || !t->flow_data[i]) continue;
        pb2_data_s* original = t->flow_data[i];
        std::lock_guard<std::mutex> g(original->lock);
        pb2_copy_t* gpu = original->device_copies[d->device_index];
        if (!(gpu->flags & PB2_FLAG_OWNED)) continue;  // the user's own device copy: not ours to book
        bool d2h = false;
        if (type & PB2_FLOW_ACCESS_READ) {
            gpu->readers--;
            if (gpu->readers == 0 && !(type & PB2_FLOW_ACCESS_WRITE)) {
                CopyList::chop(gpu, d->lists);
                d->lru.push_back(gpu);
                trace(d, t, i, EV_POP, 0);
                continue;
            }
        }
        if (type & PB2_FLOW_ACCESS_WRITE) {
            d->stats.required_data_out += t->flow_nb_elts[i];
            if (t->pushout & (1 << i)) {
                d->stats.data_out_to_host += t->flow_nb_elts[i];
                d2h = true;
            }
        }
        trace(d, t, i, EV_POP, d2h);
    }
    for (int i = 0; i < t->nb_flows; ++i) {
        const uint8_t type = t->flow_access[i];
        if (!(type & PB2_FLOW_ACCESS_WRITE) || !t->flow_data[i]) continue;
        pb2_data_s* original = t->flow_data[i];
        std::lock_guard<std::mutex> g(original->lock);
        pb2_copy_t* gpu = original->device_copies[d->device_index];
        pb2_copy_t* cpu = &original->host_copy;
        if (!(gpu->flags & PB2_FLAG_OWNED)) continue;
        gpu->coherency_state = PB2_DATA_COHERENCY_SHARED;
        cpu->coherency_state = PB2_DATA_COHERENCY_SHARED;
        cpu->version = gpu->version;
        if (t->pushout & (1 << i)) {
            CopyList::chop(gpu, d->lists);
            d->lru.push_back(gpu);
        } else {
            d->owned_lru.push_back(gpu);
        }
        trace(d, t, i, EV_EPILOG, 0);
    }
    release_peer_sources(d, st, t->nb_flows);
    if (t->task_type == 0) d->stats.executed_tasks++;
    // the generated complete_hook bumps data_out->version of every WRITE flow (jdf2c.c:7102); data_out is the
    // host copy after the epilog.  It runs inside the runtime's completion, i.e. inside t->complete; the trace
    // records the state right after it.
}

// *drained is set by the decrement that takes the task count to zero.  From that instant the next caller of
// pb2_kernel_scheduler sees rc == 0 and is the manager (:2573-2589), so this thread must not touch the module's state
// any more: it returns at once and manager_loop leaves without looking at anything again, exactly like the reference's
// `rc = fetch_dec(mutex); if (1 == rc) leave` (:2742-2753).  Re-reading the counter instead would let two managers run.
static bool progress_completions(pb2_device_s* d, bool* drained) {
    bool any = false;
    cudaEvent_t done_seen = nullptr;
    for (auto& queue : d->inflight_q) {
        while (!queue.empty()) {
            InFlight f = queue.front();
            // the tasks of a batch share one completion event: ask once
            cudaError_t q = f.st->done == done_seen ? cudaSuccess : cudaEventQuery(f.st->done);
            if (q == cudaSuccess) done_seen = f.st->done_shared ? f.st->done : nullptr;
            if (q == cudaErrorNotReady) break;
            if (q != cudaSuccess) {
                fprintf(stderr, "[parsec_b200] device error while waiting for a task: %s\n", cudaGetErrorString(q));
                d->failed = true;
            }
            queue.pop_front();
            d->inflight_count--;
            if (!f.st->done_shared) d->event_pool.push_back(f.st->done);
            any = true;
            if (f.st->is_w2r) {
                complete_w2r(d, f.st);
                free_state(d, f.st);
                continue;
            }
            pb2_gpu_task_t* t = f.task;
            finish_task(d, t, f.st);
            free_state(d, f.st);
            t->reserved[0] = nullptr;
            if (t->complete) t->complete(t, d->failed ? PB2_HOOK_RETURN_DISABLE : PB2_HOOK_RETURN_DONE, t->complete_cookie);
            if (d->params.trace_events)
                for (int i = 0; i < t->nb_flows; ++i)
                    if ((t->flow_access[i] & PB2_FLOW_ACCESS_WRITE) && t->flow_data[i]) trace(d, t, i, EV_COMPLETE, 0);
            free(t);
            if (d->mutex.fetch_sub(1) == 1) {
                *drained = true;
                return true;
            }
        }
    }
    return any;
}

static int manager_loop(pb2_device_s* d, pb2_gpu_task_t* task);

static void push_pending(pb2_device_s* d, pb2_gpu_task_t* task) {
    std::lock_guard<std::mutex> g(d->pending_lock);
    d->pending.push_back({task, d->params.use_priorities ? task->priority : 0, d->pending_seq++});
    std::push_heap(d->pending.begin(), d->pending.end());
}

extern "C" int pb2_kernel_scheduler(pb2_device_t* d, pb2_gpu_task_t* task) {
    if (!d || !task) return PB2_HOOK_RETURN_ERROR;
    if (d->failed) return PB2_HOOK_RETURN_DISABLE;
    int32_t rc = d->mutex.fetch_add(1);
    if (rc > 0) {  // somebody manages this GPU: hand the task over (:2586-2589)
        push_pending(d, task);
        return PB2_HOOK_RETURN_ASYNC;
    }
    return manager_loop(d, task);
}

// Two-step form of the same election for a caller that wants to queue a batch before it starts driving:
// pb2_device_hold takes the manager seat (so that the submissions that follow are only queued),
// pb2_device_manage gives the phantom count back and drives the device until it drains.  Without it the first
// task of a taskpool would run alone.  No counterpart in the reference (its workers submit concurrently).
extern "C" int pb2_device_hold(pb2_device_t* d) {
    if (!d) return 0;
    // A failed attempt must never move the counter: an increment followed by a decrement could land around the active
    // manager's last completion, which would then see 2 instead of 1 and never notice that the device has drained.
    int32_t expected = 0;
    return d->mutex.compare_exchange_strong(expected, 1) ? 1 : 0;
}
extern "C" int pb2_device_manage(pb2_device_t* d) {
    if (!d) return PB2_ERR_BAD_PARAM;
    if (d->mutex.fetch_sub(1) == 1) return d->failed ? PB2_ERR_DEVICE : PB2_SUCCESS;  // nothing was queued
    int rc = manager_loop(d, nullptr);
    return rc == PB2_HOOK_RETURN_ASYNC ? PB2_SUCCESS : PB2_ERR_DEVICE;
}

// next pending task (caller holds pending_lock): the top of the heap, or with sort_pending_tasks the one among the
// ten best that has to bring in the fewest tiles (parsec_gpu_sort_pending_list, device_gpu.c:117-166)
static pb2_gpu_task_t* pop_pending(pb2_device_s* d) {
    if (d->pending.empty()) return nullptr;
    auto pop_top = [d]() {
        std::pop_heap(d->pending.begin(), d->pending.end());
        pb2_device_s::Pending p = d->pending.back();
        d->pending.pop_back();
        return p;
    };
    if (!d->params.sort_pending_tasks) return pop_top().task;
    std::vector<pb2_device_s::Pending> head;
    const size_t span = std::min<size_t>(10, d->pending.size());
    for (size_t i = 0; i < span; ++i) head.push_back(pop_top());
    size_t pick = 0;
    int best = 1 << 30;
    for (size_t i = 0; i < head.size(); ++i) {
        const pb2_gpu_task_t* c = head[i].task;
        int missing = 0;
        for (int f = 0; f < c->nb_flows; ++f)
            if (c->flow_data[f] && (c->flow_access[f] & PB2_FLOW_ACCESS_RW) && !c->flow_data[f]->device_copies[d->device_index])
                ++missing;
        if (missing < best) {
            best = missing;
            pick = i;
        }
    }
    for (size_t i = 0; i < head.size(); ++i)
        if (i != pick) {
            d->pending.push_back(head[i]);
            std::push_heap(d->pending.begin(), d->pending.end());
        }
    return head[pick].task;
}

static int manager_loop(pb2_device_s* d, pb2_gpu_task_t* task) {
    // this thread is the manager until the device drains
    if (cudaSetDevice(d->cuda_index) != cudaSuccess) {
        d->failed = true;
        return PB2_HOOK_RETURN_DISABLE;
    }
    pb2_gpu_task_t* cur = task;
    const size_t max_inflight = (size_t)std::max(1, d->params.max_inflight_tasks);
    uint64_t idle_spins = 0;
    int burst = 0;  // submissions since the completion queues were last polled
    bool drained = false;
    const auto leave = [d]() { return d->failed ? PB2_HOOK_RETURN_DISABLE : PB2_HOOK_RETURN_ASYNC; };
    while (true) {
        bool progressed = false;
        drain_peer_unpins(d);
        // 1. stalled tasks go first (they hold the oldest priorities), then the new one
        if (cur == nullptr && !d->stalled.empty() && d->inflight_count < max_inflight) {
            cur = d->stalled.front();
            d->stalled.pop_front();
        }
        if (cur == nullptr && d->inflight_count < max_inflight) {
            std::lock_guard<std::mutex> g(d->pending_lock);
            while (!d->deferred.empty()) {  // queued while nobody managed the device: they count from now on
                d->mutex.fetch_add(1);
                d->pending.push_back({d->deferred.front(), 0, d->pending_seq++});
                std::push_heap(d->pending.begin(), d->pending.end());
                d->deferred.pop_front();
            }
            cur = pop_pending(d);
        }
        if (cur) {
            auto ts0 = std::chrono::steady_clock::now();
            int r;
            if (d->params.batch_bodies > 1) {
                // gather what is ready now, up to batch_bodies tasks: stage each in, then run their bodies as one batch
                std::vector<pb2_gpu_task_t*> batch;
                std::vector<std::pair<pb2_gpu_task_t*, int>> declined;
                r = try_stage(d, cur);
                while (r == 0) {
                    batch.push_back(cur);
                    cur = nullptr;
                    if ((int)batch.size() >= d->params.batch_bodies || d->inflight_count + batch.size() >= max_inflight) break;
                    if (!d->stalled.empty()) {
                        cur = d->stalled.front();
                        d->stalled.pop_front();
                    } else {
                        std::lock_guard<std::mutex> g(d->pending_lock);
                        cur = pop_pending(d);
                    }
                    if (!cur) break;
                    r = try_stage(d, cur);
                }
                const size_t launched = batch.size();
                submit_batch(d, batch, declined);
                d->stats.nb_batches += launched ? 1 : 0;
                bool left = false;
                for (auto& f : declined) {
                    if (f.first->complete) f.first->complete(f.first, f.second, f.first->complete_cookie);
                    free(f.first);
                    if (d->mutex.fetch_sub(1) == 1) left = true;
                }
                if (left) return leave();
                if (!cur) r = 0;  // everything gathered is in flight
            } else {
                r = try_submit(d, cur);
            }
            d->stats.host_submit_ns +=
                (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - ts0).count();
            if (r == 0) {
                progressed = true;
                cur = nullptr;
                if (d->params.batch_bodies > 1) burst = 8;  // a batch is a burst
                // keep feeding the streams: poll completions only every few submissions (each poll is one
                // cudaEventQuery per stream), a submission is what keeps the copy engines and SMs busy
                if (++burst < 8) continue;
            } else if (r == 1) {
                d->stats.nb_stalls++;
                // no room: write dirty tiles back so that they become evictable, then wait for completions
                bool w = start_w2r(d);
                if (!w && d->inflight_count == 0) {
                    std::unique_lock<std::mutex> pk(d->peer_lock);
                    const bool peers_hold_tiles = !d->peer_unpins.empty();
                    pk.unlock();
                    if (peers_hold_tiles) {  // readers handed back by a peer GPU: take them in and try again
                        d->stalled.push_front(cur);
                        cur = nullptr;
                        continue;
                    }
                    fprintf(stderr,
                            "[parsec_b200] task needs more device memory than can ever be freed (%zu bytes in use)\n",
                            d->heap->in_use());
                    if (cur->complete) cur->complete(cur, PB2_HOOK_RETURN_ERROR, cur->complete_cookie);
                    free(cur);
                    if (d->mutex.fetch_sub(1) == 1) return leave();
                    cur = nullptr;
                } else {
                    d->stalled.push_front(cur);
                    cur = nullptr;
                    // wait until something completes, otherwise we would spin on the same failure; yield the core
                    // between polls (one manager thread per GPU shares the host with the submitters)
                    while (!progress_completions(d, &drained)) std::this_thread::yield();
                    if (drained) return leave();
                    progressed = true;
                }
            } else {
                // the task does not run here (stage-in already taken back by try_submit); the device itself is
                // only lost when a CUDA call failed
                if (cur->complete) cur->complete(cur, r, cur->complete_cookie);
                free(cur);
                if (d->mutex.fetch_sub(1) == 1) return leave();
                cur = nullptr;
                progressed = true;
            }
        }
        // 2. retire whatever finished
        burst = 0;
        {
            auto tr0 = std::chrono::steady_clock::now();
            bool any = progress_completions(d, &drained);
            if (drained) return leave();
            if (any) {
                progressed = true;
                d->stats.host_retire_ns += (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(
                                               std::chrono::steady_clock::now() - tr0).count();
            }
        }
        // 3. the device drains inside progress_completions (the decrement that reaches zero), never by re-reading the
        //    counter here: see the comment there
        if (!progressed) {
            if (++idle_spins > 64 && d->inflight_count == 0 && d->stalled.empty()) {
                // tasks counted in mutex but not yet visible in pending (a submitter between its two steps)
                std::this_thread::yield();
            }
        } else {
            idle_spins = 0;
        }
    }
    return d->failed ? PB2_HOOK_RETURN_DISABLE : PB2_HOOK_RETURN_ASYNC;
}

// ------------------------------------------------------------------------------------ advice (a16)
extern "C" int pb2_device_data_advise(pb2_device_t* d, pb2_data_t* data, int advice) {
    if (!d || !data) return PB2_ERR_BAD_PARAM;
    switch (advice) {
        case PB2_DEV_DATA_ADVICE_PREFERRED_DEVICE:
            data->preferred_device = (int8_t)d->device_index;
            return PB2_SUCCESS;
        case PB2_DEV_DATA_ADVICE_PREFETCH: {
            // a one-flow READ task with no body, pushed through the same pipeline (:1626-1672)
            pb2_gpu_task_t* t = (pb2_gpu_task_t*)calloc(1, sizeof(pb2_gpu_task_t));
            t->task_type = 1;
            t->nb_flows = 1;
            t->flow_data[0] = data;
            t->flow_access[0] = PB2_FLOW_ACCESS_READ;
            t->flow_nb_elts[0] = data->nb_elts;
            t->trace_dc[0] = (int32_t)data->dc_id;
            t->locals[2] = -1;
            if (d->params.prefetch_deferred) {
                // only queue: join the running manager's count if there is one, else wait for the next manager
                int32_t cur = d->mutex.load();
                while (true) {
                    if (cur == 0) {
                        std::lock_guard<std::mutex> g(d->pending_lock);
                        d->deferred.push_back(t);
                        return PB2_SUCCESS;
                    }
                    if (d->mutex.compare_exchange_weak(cur, cur + 1)) {
                        push_pending(d, t);
                        return PB2_SUCCESS;
                    }
                }
            }
            int rc = pb2_kernel_scheduler(d, t);
            return rc == PB2_HOOK_RETURN_ASYNC ? PB2_SUCCESS : PB2_ERROR;
        }
        case PB2_DEV_DATA_ADVICE_WARMUP: {
            // "as if a task had just accessed this data": refresh its LRU position (unimplemented upstream)
            if (d->mutex.load() != 0) return PB2_ERR_NOT_SUPPORTED;
            pb2_copy_t* c = data->device_copies[d->device_index];
            if (c && c->list_id == 1) {
                CopyList::chop(c, d->lists);
                d->lru.push_back(c);
            }
            return PB2_SUCCESS;
        }
        default:
            return PB2_ERR_NOT_FOUND;
    }
}

// ------------------------------------------------------------------------------------ heap test hooks
// The interface of parsec/utils/zone_malloc.h:41-63 over pb2::ZoneHeap.  Exposed so that the heap's invariants can be
// checked without a GPU (tests/test_cabi_host.py).
extern "C" void* pb2_zone_create(int max_segment, size_t unit_size) {
    if (max_segment <= 0 || unit_size == 0) return nullptr;
    return new ZoneHeap((char*)(uintptr_t)(1ull << 40), max_segment, unit_size);
}
extern "C" void pb2_zone_destroy(void* z) { delete (ZoneHeap*)z; }
extern "C" int64_t pb2_zone_malloc(void* z, size_t size) {
    void* p = ((ZoneHeap*)z)->alloc(size);
    return p ? ((ZoneHeap*)z)->offset_of(p) : -1;
}
extern "C" int pb2_zone_free(void* z, int64_t offset) {
    ZoneHeap* h = (ZoneHeap*)z;
    return h->free(h->base() + offset) ? PB2_SUCCESS : PB2_ERR_NOT_FOUND;
}
extern "C" size_t pb2_zone_in_use(void* z) { return ((ZoneHeap*)z)->in_use(); }

// ------------------------------------------------------------------------------------ hooks for the comm layer
// Used by the broadcast data plane (comm_nccl.cu) from the thread that holds the manager seat (pb2_device_hold),
// i.e. while no manager loop is touching the lists.  Not part of the public header.

// Make sure `data` has a valid device copy (staging it from the host through the normal reserve / stage-in path if it
// has none), register one more reader so that it cannot be evicted, and make `wait_stream` wait until its bytes
// are valid.  This is what a READ flow of a task does, minus the task.
extern "C" int pb2_device_pin_for_read(pb2_device_t* d, pb2_data_t* data, void** dev_ptr, void* wait_stream) {
    if (!d || !data || !dev_ptr) return PB2_ERR_BAD_PARAM;
    cudaSetDevice(d->cuda_index);
    pb2_gpu_task_t t;
    memset(&t, 0, sizeof(t));
    t.task_type = 1;
    t.nb_flows = 1;
    t.flow_data[0] = data;
    t.flow_access[0] = PB2_FLOW_ACCESS_READ;
    t.flow_nb_elts[0] = data->nb_elts;
    t.trace_dc[0] = (int32_t)data->dc_id;
    t.locals[2] = -1;
    if (reserve_device_space(d, &t) != 0) return PB2_ERR_OUT_OF_RESOURCE;
    TaskState st;
    st.host_source_only = true;
    bool issued = false;
    if (stage_in_flow(d, &t, &st, 0, &issued) < 0) return PB2_ERROR;
    if (issued) {
        const uint64_t seq = d->h2d_seq.load() + 1;
        cudaEventRecord(d->h2d_events[seq % EVENT_RING], d->h2d);
        d->h2d_seq.store(seq);
        complete_push_flow(d, &t, 0);
    }
    pb2_copy_t* c = data->device_copies[d->device_index];
    if (wait_stream && c->ready_seq) {
        const uint64_t now = d->h2d_seq.load();
        uint64_t w = (now - c->ready_seq >= (uint64_t)EVENT_RING) ? now : c->ready_seq;
        cudaStreamWaitEvent((cudaStream_t)wait_stream, d->h2d_events[w % EVENT_RING], 0);
    }
    if (wait_stream && c->ready_event) cudaStreamWaitEvent((cudaStream_t)wait_stream, c->ready_event, 0);
    *dev_ptr = c->device_private;
    return PB2_SUCCESS;
}

extern "C" int pb2_device_unpin(pb2_device_t* d, pb2_data_t* data) {
    if (!d || !data) return PB2_ERR_BAD_PARAM;
    std::lock_guard<std::mutex> g(data->lock);
    pb2_copy_t* c = data->device_copies[d->device_index];
    if (!c) return PB2_ERR_NOT_FOUND;
    c->readers--;
    if (c->readers == 0 && c->list_id == 0 && c->coherency_state != PB2_DATA_COHERENCY_OWNED) d->lru.push_back(c);
    return PB2_SUCCESS;
}

// A tile that arrives from a peer GPU: give it an HBM slot and a device copy that is already SHARED with the
// sender's version, so that the consumer's stage-in finds "on the device, right version" and issues no H2D.  The
// copy is in no LRU list (it has no host backing to fall back on) until pb2_device_drop_remote.
extern "C" int pb2_device_adopt_remote(pb2_device_t* d, pb2_data_t* data, uint32_t version, void** dev_ptr) {
    if (!d || !data || !dev_ptr) return PB2_ERR_BAD_PARAM;
    if (data->device_copies[d->device_index]) return PB2_ERR_EXISTS;
    void* mem = alloc_evicting(d, data->nb_elts, nullptr);
    if (!mem) return PB2_ERR_OUT_OF_RESOURCE;
    pb2_copy_t* c = new pb2_copy_t();
    c->flags = PB2_FLAG_OWNED | PB2_FLAG_MANAGED;
    c->device_private = mem;
    c->coherency_state = PB2_DATA_COHERENCY_SHARED;
    c->data_transfer_status = PB2_XFER_COMPLETE;
    c->version = version;
    c->readers = 1;  // held until pb2_device_drop_remote: there is no host copy to re-stage it from
    c->device_index = (int8_t)d->device_index;
    c->original = data;
    data->device_copies[d->device_index] = c;
    d->all_copies.insert(c);
    data->host_copy.version = version;
    data->host_copy.coherency_state = PB2_DATA_COHERENCY_SHARED;
    d->stats.data_in_from_peers += data->nb_elts;
    *dev_ptr = mem;
    return PB2_SUCCESS;
}

extern "C" int pb2_device_set_ready_event(pb2_device_t* d, pb2_data_t* data, void* cuda_event) {
    if (!d || !data) return PB2_ERR_BAD_PARAM;
    pb2_copy_t* c = data->device_copies[d->device_index];
    if (!c) return PB2_ERR_NOT_FOUND;
    c->ready_event = (cudaEvent_t)cuda_event;
    return PB2_SUCCESS;
}

extern "C" int pb2_device_drop_remote(pb2_device_t* d, pb2_data_t* data) {
    if (!d || !data) return PB2_ERR_BAD_PARAM;
    pb2_copy_t* c = data->device_copies[d->device_index];
    if (!c) return PB2_ERR_NOT_FOUND;
    release_copy(d, c);
    return PB2_SUCCESS;
}

// public form of the same, for the overlay shim: the runtime has destroyed the data this copy belonged to
extern "C" int pb2_device_drop_copy(pb2_device_t* d, pb2_data_t* data) {
    if (!d || !data) return PB2_ERR_BAD_PARAM;
    drain_peer_unpins(d);
    pb2_copy_t* c = data->device_copies[d->device_index];
    if (!c) return PB2_ERR_NOT_FOUND;
    release_copy(d, c);
    return PB2_SUCCESS;
}

// A device copy allocated by the user (cudaMalloc) and attached to the data by hand, as tests/dsl/ptg/cuda/
// nvlink_wrapper.c:189-206 does with parsec_data_copy_attach + parsec_data_transfer_ownership_to_copy: the module uses
// the address as it is and keeps its hands off (no staging, no eviction, no write-back, never freed here).  Calling it
// again updates the address.
extern "C" int pb2_device_attach_user_copy(pb2_device_t* d, pb2_data_t* data, void* device_ptr, uint32_t version, int coherency) {
    if (!d || !data || !device_ptr) return PB2_ERR_BAD_PARAM;
    std::lock_guard<std::mutex> g(data->lock);
    pb2_copy_t* c = data->device_copies[d->device_index];
    if (c && (c->flags & PB2_FLAG_OWNED)) return PB2_ERR_EXISTS;  // the module already caches this tile itself
    if (!c) {
        c = new pb2_copy_t();
        c->flags = 0;
        c->device_index = (int8_t)d->device_index;
        c->original = data;
        c->data_transfer_status = PB2_XFER_COMPLETE;
        data->device_copies[d->device_index] = c;
    }
    c->device_private = device_ptr;
    c->version = version;
    c->coherency_state = (uint8_t)coherency;
    if (coherency == PB2_DATA_COHERENCY_OWNED) data->owner_device = (int8_t)d->device_index;
    return PB2_SUCCESS;
}
extern "C" int pb2_device_detach_user_copy(pb2_device_t* d, pb2_data_t* data) {
    if (!d || !data) return PB2_ERR_BAD_PARAM;
    std::lock_guard<std::mutex> g(data->lock);
    pb2_copy_t* c = data->device_copies[d->device_index];
    if (!c || (c->flags & PB2_FLAG_OWNED)) return PB2_ERR_NOT_FOUND;
    data->device_copies[d->device_index] = nullptr;
    delete c;
    return PB2_SUCCESS;
}

extern "C" size_t pb2_device_heap_capacity(pb2_device_t* d) { return d && d->heap ? d->heap->capacity() : 0; }
extern "C" int pb2_device_exec_streams(pb2_device_t* d) { return d ? (int)d->exec.size() : 0; }

// ------------------------------------------------------------------------------------ per-stream workspace (a19)
// parsec_gpu_pop_workspace / _push_workspace / _free_workspace (device_gpu.c:168-246).  Manager thread only (the
// submit callback runs there), like every other user of the HBM heap.
static int exec_stream_index(pb2_device_s* d, void* cuda_stream) {
    for (size_t i = 0; i < d->exec.size(); ++i)
        if ((void*)d->exec[i] == cuda_stream) return (int)i;
    return -1;
}

extern "C" void* pb2_device_pop_workspace(pb2_device_t* d, void* cuda_stream, size_t size) {
    if (!d || !size) return nullptr;
    const int s = exec_stream_index(d, cuda_stream);
    if (s < 0) return nullptr;
    if (d->workspace.size() != d->exec.size()) d->workspace.resize(d->exec.size());
    pb2_device_s::Workspace& w = d->workspace[s];
    if (!w.ready) {
        for (int i = 0; i < PB2_GPU_MAX_WORKSPACE; ++i) {
            w.buf[i] = alloc_evicting(d, size, nullptr);
            if (!w.buf[i]) {  // no room: give back what was taken, stay un-initialised
                for (int j = 0; j < i; ++j) {
                    d->heap->free(w.buf[j]);
                    w.buf[j] = nullptr;
                }
                return nullptr;
            }
        }
        w.size = size;
        w.stack_head = PB2_GPU_MAX_WORKSPACE - 1;
        w.ready = true;
    }
    if (size > w.size || w.stack_head < 0) return nullptr;  // the reference asserts; this boundary reports
    return w.buf[w.stack_head--];
}

extern "C" int pb2_device_push_workspace(pb2_device_t* d, void* cuda_stream) {
    if (!d) return PB2_ERR_BAD_PARAM;
    const int s = exec_stream_index(d, cuda_stream);
    if (s < 0 || (size_t)s >= d->workspace.size() || !d->workspace[s].ready) return PB2_ERR_BAD_PARAM;
    pb2_device_s::Workspace& w = d->workspace[s];
    if (w.stack_head + 1 >= PB2_GPU_MAX_WORKSPACE) return PB2_ERR_BAD_PARAM;
    w.stack_head++;
    return PB2_SUCCESS;
}

extern "C" int pb2_device_free_workspace(pb2_device_t* d) {
    if (!d) return PB2_ERR_BAD_PARAM;
    for (auto& w : d->workspace) {
        if (!w.ready) continue;
        for (int i = 0; i < PB2_GPU_MAX_WORKSPACE; ++i) {
            if (w.buf[i]) d->heap->free(w.buf[i]);
            w.buf[i] = nullptr;
        }
        w.ready = false;
        w.stack_head = -1;
        w.size = 0;
    }
    return PB2_SUCCESS;
}
