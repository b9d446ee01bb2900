/*
 * device_cuda_b200_shim.c -- the CUDA device module of a PaRSEC build, backed by libparsec_b200.so.
 *
 * Compiled INSIDE the reference tree in place of three of its files (overlay/build_overlay.sh):
 *     parsec/mca/device/cuda/device_cuda_module.c   the module itself
 *     parsec/mca/device/device_gpu.c                workspace, pending-list sort, debug output
 *     parsec/mca/device/transfer_gpu.c              W2R eviction tasks
 * Everything else of the reference is compiled unchanged: device_cuda_component.c still registers the
 * device_cuda_* MCA parameters and calls parsec_cuda_module_init / _fini (device_cuda_component.c:87-161,
 * :163-206, :271-325), device.c still owns the registry and parsec_get_best_device, data.c the copies, and
 * parsec-ptgpp still emits calls to parsec_cuda_kernel_scheduler and parsec_default_cuda_stage_in / _out by name
 * (jdf2c.c:6789-6795, :6902).  This file exports exactly those symbols and fills the reference's own structures
 * (parsec_device_cuda_module_t, parsec_cuda_exec_stream_t with per-stream infos, parsec_data_copy_t) so that
 * generated code, DTD and user bodies run unmodified.
 *
 * Division of labour: libparsec_b200.so (C ABI, include/parsec_b200.h) owns the streams, the HBM heap, the tile
 * cache with its LRU lists and the stage-in -> kernel -> stage-out chaining; it keeps a mirror (pb2_data_t) of
 * every parsec_data_t it has seen.  This shim translates a parsec_gpu_task_t into a pb2_gpu_task_t, keeps the
 * runtime's parsec_data_copy_t objects in step with the mirror (device_private, version, coherency_state, readers,
 * owner_device: the values of SURVEY.md Appendix C), and runs the epilog + __parsec_complete_execution where the
 * reference does (device_cuda_module.c:2716-2735).
 *
 * Plain C; the only CUDA calls are cudaSetDevice / cudaGetDeviceProperties / cudaHostRegister, as in the
 * reference's module.
 */
#include "parsec/parsec_config.h"
#include "parsec/parsec_internal.h"
#include "parsec/mca/device/cuda/device_cuda_internal.h"
#include "parsec/mca/device/cuda/device_cuda.h"
#include "parsec/mca/device/device_gpu.h"
#include "parsec/scheduling.h"
#include "parsec/execution_stream.h"
#include "parsec/constants.h"
#include "parsec/utils/argv.h"
#include "parsec/utils/debug.h"
#include "parsec/utils/mca_param.h"
#include "parsec/utils/output.h"
#include "parsec/class/list.h"
#include "parsec/data_internal.h"

#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "parsec_b200.h"

#if defined(PARSEC_PROF_TRACE)
#error "the B200 overlay is built with PARSEC_PROF_TRACE=OFF: kernel evidence comes from ncu / CUDA events, see DESIGN.md"
#endif

/* ------------------------------------------------------------------------------------------------------------
 * What device_gpu.c and transfer_gpu.c used to define.
 * ---------------------------------------------------------------------------------------------------------- */
int parsec_gpu_output_stream = -1;
int parsec_gpu_verbosity;
int32_t parsec_gpu_d2h_max_flows = MAX_PARAM_COUNT; /* device_cuda_max_number_of_ejected_data (device_cuda_component.c:191) */

void parsec_gpu_enable_debug(void)
{   /* device_gpu.c:106-115 */
    if( -1 == parsec_gpu_output_stream ) {
        parsec_gpu_output_stream = parsec_device_output;
        if( parsec_gpu_verbosity >= 0 ) {
            parsec_gpu_output_stream = parsec_output_open(NULL);
            parsec_output_set_verbosity(parsec_gpu_output_stream, parsec_gpu_verbosity);
        }
    }
}

/* ------------------------------------------------------------------------------------------------------------
 * Types
 * ---------------------------------------------------------------------------------------------------------- */
struct b200_module_s;

/* One pb2_data_t mirrors one parsec_data_t for ALL the GPUs of the process (it holds one device copy per GPU, which is
 * what lets a tile be staged from a peer GPU instead of the host); `copy` is indexed by the device index inside
 * libparsec_b200. */
struct b200_copy_s;
typedef struct b200_mirror_s {
    pb2_data_t         *data;
    struct b200_copy_s *copy[PB2_MAX_DEVICES];
    int                 nb_copies;
} b200_mirror_t;

/* The device copy the runtime sees in parsec_data_t.device_copies[device_index]: a parsec_data_copy_t plus the
 * mirror inside libparsec_b200.  It stays attached for as long as the parsec_data_t lives, INVALID with
 * device_private == NULL while the tile has no HBM (the state parsec_gpu_data_reserve_device_space creates,
 * device_cuda_module.c:1143-1151); parsec_data_destruct detaches it without releasing it (data.c:88-104, GPU
 * copies belong to their module), which is how the shim learns that the data is gone (original == NULL). */
typedef struct b200_copy_s {
    parsec_data_copy_t    super;     /* super.super is the list item that files it in b200_module_t.copies */
    b200_mirror_t        *mirror;
    struct b200_module_s *mod;
} b200_copy_t;
PARSEC_OBJ_CLASS_DECLARATION(b200_copy_t);
PARSEC_OBJ_CLASS_INSTANCE(b200_copy_t, parsec_data_copy_t, NULL, NULL);

/* A device copy the user allocated and attached (nvlink_wrapper.c:189-206): a plain parsec_data_copy_t without
 * PARSEC_DATA_FLAG_PARSEC_OWNED.  The module only hands its address to the bodies; the mirror exists so that the
 * library knows the tile "is on the device". */
typedef struct b200_foreign_s {
    parsec_data_copy_t    *copy;
    pb2_data_t            *data;
    struct b200_foreign_s *next;
} b200_foreign_t;

typedef struct b200_module_s {
    parsec_device_cuda_module_t super;
    pb2_device_t               *pb2;
    int                         pb2_index;   /* device index inside libparsec_b200 (pb2_data_state_t.owner_device) */
    parsec_list_t               copies;      /* every b200_copy_t of this module (under b200_lock) */
    struct b200_foreign_s      *foreign;     /* device copies the user attached by hand, and their mirrors */
    pb2_device_stats_t          folded;      /* what has been added to the parsec_device_module_t counters so far */
} b200_module_t;

/* creation and retirement of mirrors and copies, the modules' lists of copies: rare events, one lock for all GPUs */
static pthread_mutex_t b200_lock = PTHREAD_MUTEX_INITIALIZER;
static b200_module_t *b200_modules[PB2_MAX_DEVICES];   /* by pb2 device index */

/* one per task in flight; pb2 frees it (free()) after the completion callback, it is the head of the block */
typedef struct b200_task_s {
    pb2_gpu_task_t     pb2;
    parsec_gpu_task_t *gpu_task;
    b200_module_t     *mod;
    b200_copy_t       *copy[MAX_PARAM_COUNT];
    parsec_data_copy_t *foreign[MAX_PARAM_COUNT];   /* user-managed device copies among the flows */
} b200_task_t;

/* The execution stream of the thread that is inside parsec_cuda_kernel_scheduler.  Completions run on the thread
 * that manages the device, and __parsec_complete_execution must be given THAT thread's stream (it schedules the
 * successors on its queues), as in the reference where `es` is the manager's (device_cuda_module.c:2733). */
static __thread parsec_execution_stream_t *b200_tls_es = NULL;

/* ------------------------------------------------------------------------------------------------------------
 * Mirrors
 * ---------------------------------------------------------------------------------------------------------- */
static int b200_pb2_to_parsec_device(const b200_module_t *mod, int pb2_owner)
{
    (void)mod;
    if( pb2_owner <= 0 ) return pb2_owner;   /* 0 = host, -1 = nobody */
    if( pb2_owner < PB2_MAX_DEVICES && NULL != b200_modules[pb2_owner] )
        return b200_modules[pb2_owner]->super.super.super.device_index;
    return -1;
}

/* Write the device side of the mirror into the runtime's copy. */
static void b200_refresh_copy(b200_module_t *mod, b200_copy_t *bc, parsec_data_t *original)
{
    pb2_data_state_t st;
    if( PB2_SUCCESS != pb2_data_get_state(bc->mirror->data, mod->pb2_index, &st) ) return;
    bc->super.device_private  = st.dev_ptr;
    bc->super.version         = st.dev_version;
    bc->super.coherency_state = (parsec_data_coherency_t)st.dev_coherency;
    bc->super.readers         = st.readers;
    bc->super.data_transfer_status = st.has_copy ? PARSEC_DATA_STATUS_COMPLETE_TRANSFER : PARSEC_DATA_STATUS_NOT_TRANSFER;
    if( NULL != original ) {
        /* owner_device steers parsec_get_best_device (device.c:104-107): the mirror knows which GPU, if any, holds the
         * tile in write mode; a host-side owner (a CPU task wrote it) is the runtime's own business */
        int owner = b200_pb2_to_parsec_device(mod, st.owner_device);
        if( owner > 1 || original->owner_device > 1 ) original->owner_device = (int8_t)owner;
    }
}

/* pb2 gave the HBM of this tile back (LRU eviction, memory_release): the runtime's copy stays attached but holds
 * nothing, like a copy that reserve_device_space has just created.  Manager thread. */
static void b200_hook_copy_released(pb2_device_t *dev, pb2_data_t *data, void *user)
{
    b200_mirror_t *mirror = (b200_mirror_t*)user;
    int idx = pb2_device_index(dev);
    (void)data;
    if( NULL == mirror || idx < 0 || idx >= PB2_MAX_DEVICES || NULL == mirror->copy[idx] ) return;
    b200_copy_t *bc = mirror->copy[idx];
    bc->super.device_private  = NULL;
    bc->super.version         = 0;
    bc->super.coherency_state = PARSEC_DATA_COHERENCY_INVALID;
    bc->super.readers         = 0;
    bc->super.data_transfer_status = PARSEC_DATA_STATUS_NOT_TRANSFER;
    if( NULL != bc->super.original &&
        bc->super.original->owner_device == (int8_t)bc->mod->super.super.super.device_index )
        bc->super.original->owner_device = 0;
}

static int b200_hook_data_alive(pb2_device_t *dev, pb2_data_t *data, void *user)
{
    b200_mirror_t *mirror = (b200_mirror_t*)user;
    int idx = pb2_device_index(dev);
    (void)data;
    if( NULL == mirror || idx < 0 || idx >= PB2_MAX_DEVICES || NULL == mirror->copy[idx] ) return 1;
    return NULL != mirror->copy[idx]->super.original;
}

/* Find or create the device copy of `original` on this device, and the mirror all the GPUs share.  Any thread. */
static b200_copy_t *b200_copy_of(b200_module_t *mod, parsec_data_t *original, parsec_data_copy_t *host, uint32_t nb_elts)
{
    const uint8_t dev = mod->super.super.super.device_index;
    parsec_data_copy_t *c = original->device_copies[dev];
    if( NULL != c ) {
        if( c->super.super.obj_class == &b200_copy_t_class ) return (b200_copy_t*)c;
        return NULL;  /* a copy attached by the user (cudaMalloc + parsec_data_copy_attach): not handled */
    }
    pthread_mutex_lock(&b200_lock);
    c = original->device_copies[dev];
    if( NULL == c ) {
        b200_mirror_t *mirror = NULL;
        for( uint32_t i = 2; i < parsec_nb_devices; i++ ) {   /* another GPU of this process already knows the tile */
            parsec_data_copy_t *oc = original->device_copies[i];
            if( NULL != oc && oc->super.super.obj_class == &b200_copy_t_class ) { mirror = ((b200_copy_t*)oc)->mirror; break; }
        }
        if( NULL == mirror ) {
            mirror = (b200_mirror_t*)calloc(1, sizeof(b200_mirror_t));
            mirror->data = pb2_data_create((uint64_t)original->key, NULL != original->dc ? (uint32_t)original->dc->dc_id : 0,
                                           host->device_private, nb_elts > original->nb_elts ? nb_elts : original->nb_elts);
            pb2_data_set_user(mirror->data, mirror);
        }
        b200_copy_t *bc = PARSEC_OBJ_NEW(b200_copy_t);
        bc->mod = mod;
        bc->mirror = mirror;
        bc->super.flags = PARSEC_DATA_FLAG_PARSEC_OWNED | PARSEC_DATA_FLAG_PARSEC_MANAGED;
        bc->super.coherency_state = PARSEC_DATA_COHERENCY_INVALID;
        bc->super.version = 0;
        bc->super.dtt = host->dtt;
        if( PARSEC_SUCCESS != parsec_data_copy_attach(original, &bc->super, dev) ) {
            if( 0 == mirror->nb_copies ) {
                pb2_data_set_user(mirror->data, NULL);
                pb2_data_destroy(mirror->data);
                free(mirror);
            }
            PARSEC_OBJ_RELEASE(bc);
            pthread_mutex_unlock(&b200_lock);
            return NULL;
        }
        mirror->copy[mod->pb2_index] = bc;
        mirror->nb_copies++;
        parsec_list_nolock_push_back(&mod->copies, (parsec_list_item_t*)bc);
        c = &bc->super;
    }
    pthread_mutex_unlock(&b200_lock);
    return c->super.super.obj_class == &b200_copy_t_class ? (b200_copy_t*)c : NULL;
}

/* The mirror of a tile whose device copy belongs to the user; created on first sight, refreshed on every use. */
static pb2_data_t *b200_foreign_mirror(b200_module_t *mod, parsec_data_t *original, parsec_data_copy_t *c,
                                       parsec_data_copy_t *host)
{
    b200_foreign_t *f;
    pthread_mutex_lock(&b200_lock);
    for( f = mod->foreign; NULL != f && f->copy != c; f = f->next );
    if( NULL == f ) {
        f = (b200_foreign_t*)calloc(1, sizeof(b200_foreign_t));
        f->copy = c;
        f->data = pb2_data_create((uint64_t)original->key, NULL != original->dc ? (uint32_t)original->dc->dc_id : 0,
                                  NULL != host ? host->device_private : NULL, original->nb_elts);
        f->next = mod->foreign;
        mod->foreign = f;
    }
    pthread_mutex_unlock(&b200_lock);
    if( PB2_SUCCESS != pb2_device_attach_user_copy(mod->pb2, f->data, c->device_private, c->version, (int)c->coherency_state) )
        return NULL;
    return f->data;
}

/* Drop a copy (the data is gone, or the module shuts down), and the mirror with its last copy.  Only while this
 * device is idle (its manager seat is held by the caller); b200_lock held. */
static void b200_retire_copy(b200_module_t *mod, b200_copy_t *bc)
{
    b200_mirror_t *mirror = bc->mirror;
    if( NULL != bc->super.original )
        parsec_data_copy_detach(bc->super.original, &bc->super, bc->super.device_index);
    mirror->copy[mod->pb2_index] = NULL;       /* the release hook finds nothing to update */
    pb2_device_drop_copy(mod->pb2, mirror->data);   /* gives the HBM back */
    if( 0 == --mirror->nb_copies ) {
        pb2_data_set_user(mirror->data, NULL);
        pb2_data_destroy(mirror->data);
        free(mirror);
    }
    bc->mirror = NULL;
    bc->super.device_private = NULL;
    PARSEC_OBJ_RELEASE(bc);
}

/* Collect the copies whose parsec_data_t has been destroyed (data.c:88-104 detached them).  `all` != 0: every
 * copy (module shutdown).  Needs the device idle: it takes the manager seat and gives up if somebody holds it. */
static void b200_collect(b200_module_t *mod, int all)
{
    if( !pb2_device_hold(mod->pb2) ) return;
    pthread_mutex_lock(&b200_lock);
    parsec_list_item_t *it, *keep_head = NULL, *keep_tail = NULL;
    while( NULL != (it = parsec_list_nolock_pop_front(&mod->copies)) ) {
        b200_copy_t *bc = (b200_copy_t*)it;
        PARSEC_LIST_ITEM_SINGLETON(it);
        if( all || NULL == bc->super.original ) {
            b200_retire_copy(mod, bc);
        } else {   /* chain the survivors through list_next, re-filed below */
            it->list_next = NULL;
            if( NULL == keep_head ) keep_head = it; else keep_tail->list_next = it;
            keep_tail = it;
        }
    }
    while( NULL != (it = keep_head) ) {
        keep_head = (parsec_list_item_t*)it->list_next;
        PARSEC_LIST_ITEM_SINGLETON(it);
        parsec_list_nolock_push_back(&mod->copies, it);
    }
    pthread_mutex_unlock(&b200_lock);
    pb2_device_manage(mod->pb2);  /* nothing was queued meanwhile in the usual case: returns at once */
}

static void b200_fold_stats(b200_module_t *mod)
{
    parsec_device_module_t *device = &mod->super.super.super;
    const pb2_device_stats_t *live = pb2_device_stats_live(mod->pb2);
    if( NULL == live ) return;
    if( NULL != device->data_in_from_device ) {
        for( int i = 0; i < PB2_MAX_DEVICES; i++ ) {   /* by source: [0] = host, peers by their runtime device index */
            uint64_t delta = live->data_in_from_device[i] - mod->folded.data_in_from_device[i];
            int src = b200_pb2_to_parsec_device(mod, i);
            if( 0 != delta && src >= 0 && src < (int)device->data_in_array_size ) device->data_in_from_device[src] += delta;
        }
    }
    device->data_out_to_host  += live->data_out_to_host  - mod->folded.data_out_to_host;
    device->required_data_in  += live->required_data_in  - mod->folded.required_data_in;
    device->required_data_out += live->required_data_out - mod->folded.required_data_out;
    device->nb_data_faults    += live->nb_data_faults    - mod->folded.nb_data_faults;
    mod->folded = *live;
}

/* ------------------------------------------------------------------------------------------------------------
 * Task callbacks (manager thread, cudaSetDevice done by libparsec_b200)
 * ---------------------------------------------------------------------------------------------------------- */
/* Before the body or a user stage function runs: data_out of every flow is the device copy, with the address the
 * module has reserved (the reference sets it in reserve_device_space, device_cuda_module.c:1152-1161). */
static void b200_bind_flows(b200_task_t *bt, uint32_t mask)
{
    parsec_task_t *task = bt->gpu_task->ec;
    for( int i = 0; i < bt->pb2.nb_flows; i++ ) {
        b200_copy_t *bc = bt->copy[i];
        if( NULL != bt->foreign[i] && (mask & (1u << i)) ) task->data[i].data_out = bt->foreign[i];
        if( NULL == bc || !(mask & (1u << i)) ) continue;
        b200_refresh_copy(bt->mod, bc, NULL);
        bc->super.device_private = bt->pb2.flow_dev_ptr[i];
        task->data[i].data_out = &bc->super;
    }
}

static int b200_submit(pb2_device_t *dev, pb2_gpu_task_t *t, void *cuda_stream)
{
    b200_task_t *bt = (b200_task_t*)t;
    parsec_device_gpu_module_t *gpu_device = &bt->mod->super.super;
    int s = pb2_device_stream_index(dev, cuda_stream);
    if( s < 0 || s >= gpu_device->num_exec_streams ) return PB2_HOOK_RETURN_ERROR;
    b200_bind_flows(bt, ~0u);
    /* the generated cuda_kernel_submit_<jdf>_<TASK> (jdf2c.c:6608-6724), or DTD's gpu_func_ptr */
    int rc = bt->gpu_task->submit(gpu_device, bt->gpu_task, gpu_device->exec_stream[s]);
    gpu_device->exec_stream[s]->executed++;
    return rc == PARSEC_HOOK_RETURN_DONE ? PB2_HOOK_RETURN_DONE : (rc < 0 ? rc : PB2_HOOK_RETURN_ERROR);
}

/* user stage_in= / stage_out= functions of the BODY (stage_custom.jdf:28-109) */
static int b200_stage_in(pb2_gpu_task_t *t, uint32_t flow_mask, void *cuda_stream)
{
    b200_task_t *bt = (b200_task_t*)t;
    parsec_device_gpu_module_t *gpu_device = &bt->mod->super.super;
    int s = pb2_device_stream_index(bt->mod->pb2, cuda_stream);
    if( s < 0 ) return PB2_HOOK_RETURN_ERROR;
    b200_bind_flows(bt, flow_mask);
    return bt->gpu_task->stage_in(bt->gpu_task, flow_mask, gpu_device->exec_stream[s]) == PARSEC_SUCCESS
           ? PB2_SUCCESS : PB2_HOOK_RETURN_ERROR;
}

static int b200_stage_out(pb2_gpu_task_t *t, uint32_t flow_mask, void *cuda_stream)
{
    b200_task_t *bt = (b200_task_t*)t;
    parsec_device_gpu_module_t *gpu_device = &bt->mod->super.super;
    int s = pb2_device_stream_index(bt->mod->pb2, cuda_stream);
    if( s < 0 ) return PB2_HOOK_RETURN_ERROR;
    return bt->gpu_task->stage_out(bt->gpu_task, flow_mask, gpu_device->exec_stream[s]) == PARSEC_SUCCESS
           ? PB2_SUCCESS : PB2_HOOK_RETURN_ERROR;
}

/* complete_task of the reference (device_cuda_module.c:2716-2741): epilog on the runtime's copies, completion,
 * statistics, load accounting, free the gpu_task.  pb2 has already done the device-side bookkeeping. */
static void b200_complete(pb2_gpu_task_t *t, int status, void *cookie)
{
    b200_task_t *bt = (b200_task_t*)t;
    b200_module_t *mod = bt->mod;
    parsec_device_module_t *device = &mod->super.super.super;
    parsec_gpu_task_t *gpu_task = bt->gpu_task;
    parsec_task_t *task = gpu_task->ec;
    parsec_execution_stream_t *es = b200_tls_es;
    (void)cookie;

    for( int i = 0; i < t->nb_flows; i++ ) {
        b200_copy_t *bc = bt->copy[i];
        if( NULL == bc ) continue;
        parsec_data_t *original = bc->super.original;
        b200_refresh_copy(mod, bc, original);
        if( !(t->flow_access[i] & PB2_FLOW_ACCESS_WRITE) || NULL == original ) continue;
        /* epilog (device_cuda_module.c:2415-2434): gpu and cpu copies SHARED, cpu.version = gpu.version, and the
         * CPU copy is what the successors receive; the generated complete_hook then bumps its version */
        parsec_data_copy_t *cpu_copy = original->device_copies[0];
        if( NULL != cpu_copy ) {
            if( PB2_HOOK_RETURN_DONE == status ) {
                cpu_copy->coherency_state = PARSEC_DATA_COHERENCY_SHARED;
                cpu_copy->version = bc->super.version;
            }
            task->data[i].data_out = cpu_copy;
        }
    }
    if( PB2_HOOK_RETURN_DONE != status ) {
        /* The task did not run here.  libparsec_b200 has taken its stage-in back and written the dirty inputs back to
         * the host (parsec_cuda_kernel_cleanout's job, device_cuda_module.c:2472-2526); data_out of the written flows is
         * the CPU copy by now.  A device that lost a CUDA call is fatal, as in the reference (:2757-2762); a body that
         * declined sends the task to its next incarnation (:2644-2656 -- where the reference leaves chore_mask alone and
         * would come straight back here, the CUDA incarnation is struck out first). */
        b200_fold_stats(mod);
        parsec_device_load[device->device_index] -= parsec_device_sweight[device->device_index];
        free(gpu_task);
        bt->gpu_task = NULL;
        if( PB2_HOOK_RETURN_DISABLE == status )
            parsec_fatal("GPU[%s]: critical issue related to the GPU discovered. Giving up", device->name);
        for( int j = 0; NULL != task->task_class->incarnations[j].hook; j++ ) {
            if( (task->chore_mask & (1 << j)) && PARSEC_DEV_CUDA == task->task_class->incarnations[j].type ) {
                task->chore_mask &= ~(1 << j);
                break;
            }
        }
        /* its inputs are looked up already: a second prepare_input would forget which repository entries it consumed */
        task->status = PARSEC_TASK_STATUS_HOOK;
        PARSEC_LIST_ITEM_SINGLETON(task);
        __parsec_reschedule(es, task);
        return;
    }
    device->executed_tasks++;
    b200_fold_stats(mod);
    parsec_device_load[device->device_index] -= parsec_device_sweight[device->device_index];
    free(gpu_task);
    bt->gpu_task = NULL;
    __parsec_complete_execution(es, task);
}

/* ------------------------------------------------------------------------------------------------------------
 * Entry points the generated code and DTD call by name (device_cuda.h:91-122)
 * ---------------------------------------------------------------------------------------------------------- */
parsec_hook_return_t
parsec_cuda_kernel_scheduler( parsec_execution_stream_t *es, parsec_gpu_task_t *gpu_task, int which_gpu )
{
    b200_module_t *mod = (b200_module_t*)parsec_mca_device_get(which_gpu);
    parsec_task_t *task = gpu_task->ec;
    if( NULL == mod || PARSEC_DEV_CUDA != mod->super.super.super.type ) return PARSEC_HOOK_RETURN_DISABLE;

    b200_task_t *bt = (b200_task_t*)calloc(1, sizeof(b200_task_t));
    pb2_gpu_task_t *t = &bt->pb2;
    bt->gpu_task = gpu_task;
    bt->mod = mod;
    t->task_type = 0;
    t->pushout = gpu_task->pushout;
    t->submit = b200_submit;
    t->stage_in  = (NULL == gpu_task->stage_in  || gpu_task->stage_in  == parsec_default_cuda_stage_in)  ? NULL : b200_stage_in;
    t->stage_out = (NULL == gpu_task->stage_out || gpu_task->stage_out == parsec_default_cuda_stage_out) ? NULL : b200_stage_out;
    t->ec = task;
    t->priority = task->priority;
    t->nb_flows = task->task_class->nb_flows;
    t->complete = b200_complete;
    for( int i = 0; i < 8 && i < task->task_class->nb_locals; i++ ) t->locals[i] = task->locals[i].value;
    for( int i = 0; i < t->nb_flows; i++ ) {
        const parsec_flow_t *flow = gpu_task->flow[i];
        if( NULL == flow ) continue;
        uint8_t access = (uint8_t)(flow->flow_flags & PARSEC_FLOW_ACCESS_MASK);
        parsec_data_copy_t *in = task->data[i].data_in;
        if( PARSEC_FLOW_ACCESS_NONE == access || NULL == in || NULL == in->original ) continue;
        parsec_data_t *original = in->original;
        /* the host side of the tile: data_in itself, or the CPU copy when a predecessor left its device copy in
         * data_out (read-only flows are not rewritten by the epilog) */
        parsec_data_copy_t *host = (0 == in->device_index) ? in : original->device_copies[0];
        if( NULL == host ) { free(bt); return PARSEC_HOOK_RETURN_NEXT; }
        parsec_data_copy_t *mine = original->device_copies[mod->super.super.super.device_index];
        if( NULL != mine && mine->super.super.obj_class != &b200_copy_t_class ) {
            /* a device copy the user attached: its address goes to the body as it is (device_cuda_module.c:2161-2163) */
            pb2_data_t *fm = b200_foreign_mirror(mod, original, mine, host);
            if( NULL == fm ) { free(bt); return PARSEC_HOOK_RETURN_NEXT; }
            bt->foreign[i] = mine;
            t->flow_data[i] = fm;
            t->flow_access[i] = access;
            t->flow_nb_elts[i] = gpu_task->flow_nb_elts[i];
            continue;
        }
        b200_copy_t *bc = b200_copy_of(mod, original, host, gpu_task->flow_nb_elts[i]);
        if( NULL == bc ) { free(bt); return PARSEC_HOOK_RETURN_NEXT; }
        pb2_data_sync_host(bc->mirror->data, host->device_private, host->version,
                           PARSEC_DATA_COHERENCY_OWNED == host->coherency_state ? PB2_DATA_COHERENCY_OWNED : -1);
        bt->copy[i] = bc;
        t->flow_data[i] = bc->mirror->data;
        t->flow_access[i] = access;           /* PARSEC_FLOW_ACCESS_* and PB2_FLOW_ACCESS_* are the same bits */
        t->flow_nb_elts[i] = gpu_task->flow_nb_elts[i];
        t->trace_dc[i] = NULL != original->dc ? (int32_t)original->dc->dc_id : 0;
    }
    b200_tls_es = es;
    int rc = pb2_kernel_scheduler(mod->pb2, t);
    switch( rc ) {
    case PB2_HOOK_RETURN_ASYNC:   return PARSEC_HOOK_RETURN_ASYNC;
    case PB2_HOOK_RETURN_DISABLE: /* fallthrough */
    default:
        parsec_warning("Critical issue related to the GPU discovered. Giving up\n");
        return PARSEC_HOOK_RETURN_DISABLE;
    }
}

/* Called by name from generated code only to be stored in gpu_task->stage_in / stage_out; a task that carries them
 * is staged by the module's own copy-engine path.  Kept callable for user stage functions that delegate to them
 * (same contract as device_cuda_module.c:1179-1250: task->data[i].data_in -> data_out, flow_nb_elts[i] bytes). */
int parsec_default_cuda_stage_in(parsec_gpu_task_t *gtask, uint32_t flow_mask, parsec_gpu_exec_stream_t *gpu_stream)
{
    parsec_cuda_exec_stream_t *cuda_stream = (parsec_cuda_exec_stream_t*)gpu_stream;
    parsec_task_t *task = gtask->ec;
    for( int i = 0; i < task->task_class->nb_flows; i++ ) {
        if( !(flow_mask & (1U << i)) ) continue;
        parsec_data_copy_t *in = task->data[i].data_in, *out = task->data[i].data_out;
        size_t count = (in->original->nb_elts <= out->original->nb_elts) ? in->original->nb_elts : out->original->nb_elts;
        cudaError_t e = cudaMemcpyAsync(out->device_private, in->device_private, count,
                                        0 == in->device_index ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice,
                                        cuda_stream->cuda_stream);
        PARSEC_CUDA_CHECK_ERROR("cudaMemcpyAsync ", e, { return PARSEC_ERROR; });
    }
    return PARSEC_SUCCESS;
}

int parsec_default_cuda_stage_out(parsec_gpu_task_t *gtask, uint32_t flow_mask, parsec_gpu_exec_stream_t *gpu_stream)
{
    parsec_cuda_exec_stream_t *cuda_stream = (parsec_cuda_exec_stream_t*)gpu_stream;
    parsec_task_t *task = gtask->ec;
    for( int i = 0; i < task->task_class->nb_flows; i++ ) {
        if( !(flow_mask & (1U << i)) ) continue;
        parsec_data_copy_t *in = task->data[i].data_out;  /* the device copy */
        parsec_data_copy_t *out = in->original->device_copies[0];
        size_t count = (in->original->nb_elts <= out->original->nb_elts) ? in->original->nb_elts : out->original->nb_elts;
        cudaError_t e = cudaMemcpyAsync(out->device_private, in->device_private, count, cudaMemcpyDeviceToHost,
                                        cuda_stream->cuda_stream);
        PARSEC_CUDA_CHECK_ERROR("cudaMemcpyAsync ", e, { return PARSEC_ERROR; });
    }
    return PARSEC_SUCCESS;
}

/* ------------------------------------------------------------------------------------------------------------
 * Module vtable (device.h:115-148)
 * ---------------------------------------------------------------------------------------------------------- */
static int b200_memory_register(parsec_device_module_t *device, parsec_data_collection_t *desc, void *ptr, size_t length)
{   /* device_cuda_module.c:117-146 */
    b200_module_t *mod = (b200_module_t*)device;
    if( PARSEC_MEMORY_STATUS_REGISTERED == desc->memory_registration_status ) return PARSEC_SUCCESS;
    if( PB2_SUCCESS != pb2_device_memory_register(mod->pb2, ptr, length) ) return PARSEC_ERROR;
    desc->memory_registration_status = PARSEC_MEMORY_STATUS_REGISTERED;
    return PARSEC_SUCCESS;
}

static int b200_memory_unregister(parsec_device_module_t *device, parsec_data_collection_t *desc, void *ptr)
{   /* :148-172 */
    b200_module_t *mod = (b200_module_t*)device;
    if( PARSEC_MEMORY_STATUS_UNREGISTERED == desc->memory_registration_status ) return PARSEC_SUCCESS;
    pb2_device_memory_unregister(mod->pb2, ptr);
    desc->memory_registration_status = PARSEC_MEMORY_STATUS_UNREGISTERED;
    return PARSEC_SUCCESS;
}

/* dyld= lookup: "<name>_sm<cc>" with the device's own capability first (100 on a B200), then plain "<name>".  The
 * reference walks a table of older capabilities in between (device_cuda_module.c:175-240); an sm_100a module has no
 * use for a kernel built for an older part, so the table is gone. */
static void *b200_find_incarnation(b200_module_t *mod, const char *fname)
{
    char library_name[FILENAME_MAX], function_name[FILENAME_MAX];
    char **argv = NULL;
    const char *env = getenv("PARSEC_CUCORES_LIB");
    if( NULL != env ) argv = parsec_argv_split(env, ';');
    else if( NULL != parsec_cuda_lib_path ) argv = parsec_argv_split(parsec_cuda_lib_path, ';');
    snprintf(library_name, FILENAME_MAX, "%s.so", fname);
    snprintf(function_name, FILENAME_MAX, "%s_sm%d", fname, mod->super.major * 10 + mod->super.minor);
    void *fn = parsec_device_find_function(function_name, library_name, (const char**)argv);
    if( NULL == fn ) fn = parsec_device_find_function(fname, library_name, (const char**)argv);
    if( NULL != argv ) parsec_argv_free(argv);
    return fn;
}

static int b200_taskpool_register(parsec_device_module_t *device, parsec_taskpool_t *tp)
{   /* :248-291 */
    b200_module_t *mod = (b200_module_t*)device;
    int32_t rc = PARSEC_ERR_NOT_FOUND;
    for( uint32_t i = 0; i < tp->nb_task_classes; i++ ) {
        const parsec_task_class_t *tc = tp->task_classes_array[i];
        __parsec_chore_t *chores = (__parsec_chore_t*)tc->incarnations;
        for( uint32_t j = 0; NULL != chores[j].hook; j++ ) {
            if( chores[j].type != device->type ) continue;
            if( NULL != chores[j].dyld_fn ) return PARSEC_SUCCESS;  /* resolved for another device of this type */
            if( NULL == chores[j].dyld ) { rc = PARSEC_SUCCESS; continue; }
            void *fn = b200_find_incarnation(mod, chores[j].dyld);
            if( NULL != fn ) { chores[j].dyld_fn = fn; rc = PARSEC_SUCCESS; }
        }
    }
    if( PARSEC_SUCCESS != rc ) tp->devices_index_mask &= ~(1 << device->device_index);
    return rc;
}

static int b200_taskpool_unregister(parsec_device_module_t *device, parsec_taskpool_t *tp)
{
    (void)tp;
    b200_fold_stats((b200_module_t*)device);
    return PARSEC_SUCCESS;
}

static int b200_attach(parsec_device_module_t *device, parsec_context_t *context)
{
    return parsec_mca_device_add(context, device);
}

static int b200_detach(parsec_device_module_t *device, parsec_context_t *context)
{
    (void)context;
    return parsec_mca_device_remove(device);
}

/* memory_release = parsec_cuda_flush_lru (:803-820): start from an empty cache */
static int b200_memory_release(parsec_device_module_t *device)
{
    b200_module_t *mod = (b200_module_t*)device;
    b200_collect(mod, 0);
    int rc = pb2_device_memory_release(mod->pb2);
    b200_fold_stats(mod);
    return PB2_SUCCESS == rc ? PARSEC_SUCCESS : PARSEC_ERROR;
}

static int b200_data_advise(parsec_device_module_t *device, parsec_data_t *data, int advice)
{   /* :1611-1679 */
    b200_module_t *mod = (b200_module_t*)device;
    switch( advice ) {
    case PARSEC_DEV_DATA_ADVICE_PREFERRED_DEVICE:
        data->preferred_device = device->device_index;
        return PARSEC_SUCCESS;
    case PARSEC_DEV_DATA_ADVICE_PREFETCH:
    case PARSEC_DEV_DATA_ADVICE_WARMUP: {
        parsec_data_copy_t *host = data->device_copies[0];
        if( NULL == host ) return PARSEC_ERR_NOT_FOUND;
        b200_copy_t *bc = b200_copy_of(mod, data, host, data->nb_elts);
        if( NULL == bc ) return PARSEC_ERR_NOT_SUPPORTED;
        pb2_data_sync_host(bc->mirror->data, host->device_private, host->version,
                           PARSEC_DATA_COHERENCY_OWNED == host->coherency_state ? PB2_DATA_COHERENCY_OWNED : -1);
        int rc = pb2_device_data_advise(mod->pb2, bc->mirror->data, advice);  /* same advice values */
        return PB2_SUCCESS == rc ? PARSEC_SUCCESS : (PB2_ERR_NOT_SUPPORTED == rc ? PARSEC_ERR_NOT_IMPLEMENTED : PARSEC_ERROR);
    }
    default:
        return PARSEC_ERR_NOT_FOUND;
    }
}

/* ------------------------------------------------------------------------------------------------------------
 * Per-stream workspace (device_gpu.c:168-246), pending-list sort, debug dumps, W2R task entry points
 * ---------------------------------------------------------------------------------------------------------- */
void *parsec_gpu_pop_workspace(parsec_device_gpu_module_t *gpu_device, parsec_gpu_exec_stream_t *gpu_stream, size_t size)
{
    b200_module_t *mod = (b200_module_t*)gpu_device;
    return pb2_device_pop_workspace(mod->pb2, (void*)((parsec_cuda_exec_stream_t*)gpu_stream)->cuda_stream, size);
}

int parsec_gpu_push_workspace(parsec_device_gpu_module_t *gpu_device, parsec_gpu_exec_stream_t *gpu_stream)
{
    b200_module_t *mod = (b200_module_t*)gpu_device;
    return PB2_SUCCESS == pb2_device_push_workspace(mod->pb2, (void*)((parsec_cuda_exec_stream_t*)gpu_stream)->cuda_stream)
           ? 0 : PARSEC_ERROR;
}

int parsec_gpu_free_workspace(parsec_device_gpu_module_t *gpu_device)
{
    return PB2_SUCCESS == pb2_device_free_workspace(((b200_module_t*)gpu_device)->pb2) ? 0 : PARSEC_ERROR;
}

/* The pending list lives inside libparsec_b200 and is ordered there when device_cuda_sort_pending_tasks is set
 * (pb2_device_params_t.sort_pending_tasks); nothing to do from outside. */
int parsec_gpu_sort_pending_list(parsec_device_gpu_module_t *gpu_device)
{
    (void)gpu_device;
    return 0;
}

/* Write-backs of dirty tiles are issued and completed inside libparsec_b200 (start_w2r / complete_w2r); the
 * runtime-level D2H task of transfer_gpu.c:221-338 has no role left. */
parsec_gpu_task_t *parsec_gpu_create_w2r_task(parsec_device_gpu_module_t *gpu_device, parsec_execution_stream_t *es)
{
    (void)gpu_device; (void)es;
    return NULL;
}

int parsec_gpu_complete_w2r_task(parsec_device_gpu_module_t *gpu_device, parsec_gpu_task_t *w2r_task, parsec_execution_stream_t *es)
{
    (void)gpu_device; (void)w2r_task; (void)es;
    return 0;
}

void dump_exec_stream(parsec_gpu_exec_stream_t *exec_stream)
{
    parsec_debug_verbose(0, parsec_gpu_output_stream, "Dev: GPU stream %s: %d bodies submitted",
                         NULL != exec_stream->name ? exec_stream->name : "?", exec_stream->executed);
}

void dump_GPU_state(parsec_device_gpu_module_t *gpu_device)
{
    b200_module_t *mod = (b200_module_t*)gpu_device;
    pb2_device_stats_t s;
    if( PB2_SUCCESS != pb2_device_get_stats(mod->pb2, &s, 0) ) return;
    parsec_output(parsec_gpu_output_stream,
                  "Device %d:%d (%s): %llu tasks, HBM %llu / %llu bytes in use, %llu evictions, %llu write-back tasks (%llu tiles), %llu stalls\n",
                  gpu_device->super.device_index, mod->super.cuda_index, gpu_device->super.name,
                  (unsigned long long)s.executed_tasks, (unsigned long long)s.hbm_bytes_in_use,
                  (unsigned long long)s.hbm_bytes_total, (unsigned long long)s.nb_evictions,
                  (unsigned long long)s.nb_w2r_tasks, (unsigned long long)s.nb_w2r_tiles, (unsigned long long)s.nb_stalls);
    for( int i = 0; i < gpu_device->num_exec_streams; i++ ) dump_exec_stream(gpu_device->exec_stream[i]);
}

/* ------------------------------------------------------------------------------------------------------------
 * Module init / fini (called by device_cuda_component.c:110 and :285)
 * ---------------------------------------------------------------------------------------------------------- */
/* double / single / tensor / half flop per clock per SM; what the reference keeps in
 * parsec_cuda_device_lookup_cudamp_floprate (:40-115, stops at 8.0).  Only used for load balancing. */
static void b200_floprate(int major, int *drate, int *srate, int *trate, int *hrate)
{
    if( major >= 10 )      { *drate = 64;  *srate = 128; *trate = 4096; *hrate = 256; }  /* B200: DMMA 4 x 32 flop/clk/SM halved as FMA pairs */
    else if( major == 9 )  { *drate = 64;  *srate = 128; *trate = 1024; *hrate = 256; }
    else                   { *drate = 32;  *srate = 64;  *trate = 512;  *hrate = 256; }
}

int parsec_cuda_module_init( int dev_id, parsec_device_module_t **module )
{
    struct cudaDeviceProp prop;
    cudaError_t cudastatus;
    int show_caps_index, show_caps = 0;

    *module = NULL;
    show_caps_index = parsec_mca_param_find("device", NULL, "show_capabilities");
    if( 0 < show_caps_index ) parsec_mca_param_lookup_int(show_caps_index, &show_caps);

    cudastatus = cudaSetDevice(dev_id);
    PARSEC_CUDA_CHECK_ERROR( "cudaSetDevice ", cudastatus, { return PARSEC_ERROR; } );
    cudastatus = cudaGetDeviceProperties(&prop, dev_id);
    PARSEC_CUDA_CHECK_ERROR( "cudaGetDeviceProperties ", cudastatus, { return PARSEC_ERROR; } );

    /* the component frees the module with free() (device_cuda_component.c:301) */
    b200_module_t *mod = (b200_module_t*)calloc(1, sizeof(b200_module_t));
    parsec_device_cuda_module_t *cuda_device = &mod->super;
    parsec_device_gpu_module_t *gpu_device = &cuda_device->super;
    parsec_device_module_t *device = &gpu_device->super;
    PARSEC_OBJ_CONSTRUCT(cuda_device, parsec_device_cuda_module_t);
    cuda_device->cuda_index = (uint8_t)dev_id;
    cuda_device->major = (uint8_t)prop.major;
    cuda_device->minor = (uint8_t)prop.minor;
    if( -1 == asprintf(&device->name, "%s: cuda(%d)", prop.name, dev_id) ) { device->name = NULL; free(mod); return PARSEC_ERROR; }

    /* the unchanged MCA parameters (device_cuda_component.c:163-206) */
    pb2_device_params_t params;
    pb2_device_params_default(&params);
    params.memory_use_percent      = parsec_cuda_memory_percentage;
    params.memory_block_size       = parsec_cuda_memory_block_size;
    params.memory_number_of_blocks = parsec_cuda_memory_number_of_blocks;
    params.max_number_of_ejected_data = parsec_gpu_d2h_max_flows;
    params.sort_pending_tasks      = parsec_cuda_sort_pending;
    /* device_cuda_max_streams: the reference's default (6 = h2d + d2h + 4 exec) means "not set"; this module then
     * takes its own default (18), which is what small tiles need on 148 SMs.  Any other value is obeyed. */
    if( PARSEC_GPU_MAX_STREAMS != parsec_cuda_max_streams && parsec_cuda_max_streams >= 3 )
        params.max_streams = parsec_cuda_max_streams;
    params.prefetch_deferred = 1;  /* a thread that only gives advice never becomes the device's manager */
    params.use_priorities = 1;     /* PARSEC_GPU_USE_PRIORITIES (device_gpu.h:20): task priorities order the pending list */
    {   /* batched submissions for small tiles: not a reference parameter, so an environment name of this library */
        const char *b = getenv("PB2_BATCH_BODIES");
        if( NULL != b && atoi(b) > 1 ) params.batch_bodies = atoi(b);
    }

    if( PB2_SUCCESS != pb2_device_create(dev_id, &params, &mod->pb2) ) {
        free(device->name); free(mod);
        return PARSEC_ERROR;
    }
    mod->pb2_index = pb2_device_index(mod->pb2);
    PARSEC_OBJ_CONSTRUCT(&mod->copies, parsec_list_t);
    if( mod->pb2_index >= 0 && mod->pb2_index < PB2_MAX_DEVICES ) b200_modules[mod->pb2_index] = mod;
    pb2_device_hooks_t hooks = { b200_hook_copy_released, b200_hook_data_alive };
    pb2_device_set_hooks(mod->pb2, &hooks);

    /* parsec_gpu_exec_stream_t views of the module's CUDA streams, same numbering (0 = h2d, 1 = d2h, 2.. = exec),
     * each with its info-object array: bodies fetch per-stream state (cuBLAS handles...) with
     * parsec_info_get(&gpu_stream->infos, id) (nvlink.jdf:140, info.h:149-236) */
    int nstreams = pb2_device_nb_streams(mod->pb2);
    if( nstreams > 255 ) nstreams = 255;
    gpu_device->max_exec_streams = (uint8_t)nstreams;
    gpu_device->exec_stream = (parsec_gpu_exec_stream_t**)malloc(nstreams * sizeof(parsec_gpu_exec_stream_t*));
    gpu_device->exec_stream[0] = (parsec_gpu_exec_stream_t*)calloc(nstreams, sizeof(parsec_cuda_exec_stream_t));
    for( int j = 0; j < nstreams; j++ ) {
        parsec_cuda_exec_stream_t *cuda_stream = (parsec_cuda_exec_stream_t*)gpu_device->exec_stream[0] + j;
        parsec_gpu_exec_stream_t *exec_stream = &cuda_stream->super;
        gpu_device->exec_stream[j] = exec_stream;
        gpu_device->num_exec_streams++;
        cuda_stream->cuda_stream = (cudaStream_t)pb2_device_stream(mod->pb2, j);
        cuda_stream->events = NULL;        /* completion events live inside libparsec_b200 */
        exec_stream->tasks = NULL;
        exec_stream->max_events = 0;
        exec_stream->workspace = NULL;
        exec_stream->fifo_pending = NULL;
        PARSEC_OBJ_CONSTRUCT(&exec_stream->infos, parsec_info_object_array_t);
        parsec_info_object_array_init(&exec_stream->infos, &parsec_per_stream_infos, exec_stream);
        if( -1 == asprintf(&exec_stream->name, 0 == j ? "h2d_cuda(%d)" : (1 == j ? "d2h_cuda(%d)" : "cuda(%d)"), j) )
            exec_stream->name = NULL;
    }

    device->type                = PARSEC_DEV_CUDA;
    device->attach              = b200_attach;
    device->detach              = b200_detach;
    device->memory_register     = b200_memory_register;
    device->memory_unregister   = b200_memory_unregister;
    device->taskpool_register   = b200_taskpool_register;
    device->taskpool_unregister = b200_taskpool_unregister;
    device->data_advise         = b200_data_advise;
    device->memory_release      = b200_memory_release;

    int drate, srate, trate, hrate;
    float clock_ghz = prop.clockRate / 1e6f;
    b200_floprate(prop.major, &drate, &srate, &trate, &hrate);
    device->device_hweight = (float)prop.multiProcessorCount * (float)hrate * clock_ghz * 2.f;
    device->device_tweight = (float)prop.multiProcessorCount * (float)trate * clock_ghz * 2.f;
    device->device_sweight = (float)prop.multiProcessorCount * (float)srate * clock_ghz * 2.f;
    device->device_dweight = (float)prop.multiProcessorCount * (float)drate * clock_ghz * 2.f;

    PARSEC_OBJ_CONSTRUCT(&gpu_device->gpu_mem_lru, parsec_list_t);        /* kept constructed: dump tools walk them */
    PARSEC_OBJ_CONSTRUCT(&gpu_device->gpu_mem_owned_lru, parsec_list_t);
    PARSEC_OBJ_CONSTRUCT(&gpu_device->pending, parsec_fifo_t);
    gpu_device->sort_starting_p = NULL;
    gpu_device->peer_access_mask = 0;
    gpu_device->memory = NULL;             /* the HBM heap is inside libparsec_b200 */
    gpu_device->mem_block_size = (size_t)params.memory_block_size;
    {
        pb2_device_stats_t s;
        pb2_device_get_stats(mod->pb2, &s, 0);
        gpu_device->mem_nb_blocks = (int64_t)(s.hbm_bytes_total / (uint64_t)params.memory_block_size);
    }
    if( show_caps ) {
        parsec_inform("GPU Device %d (capability %d.%d): %s\n\tSM                 : %d\n\tclockRate (GHz)    : %2.2f\n"
                      "\tHBM heap           : %lld blocks of %zu bytes\n"
                      "\tpeak Gflops         : double %2.3f, single %2.3f tensor %2.3f half %2.3f\n",
                      cuda_device->cuda_index, cuda_device->major, cuda_device->minor, device->name,
                      prop.multiProcessorCount, clock_ghz, (long long)gpu_device->mem_nb_blocks, gpu_device->mem_block_size,
                      device->device_dweight, device->device_sweight, device->device_tweight, device->device_hweight);
    }
    *module = device;
    return PARSEC_SUCCESS;
}

int parsec_cuda_module_fini(parsec_device_module_t *device)
{
    b200_module_t *mod = (b200_module_t*)device;
    parsec_device_gpu_module_t *gpu_device = &mod->super.super;

    b200_fold_stats(mod);
    b200_collect(mod, 1);
    while( NULL != mod->foreign ) {
        b200_foreign_t *f = mod->foreign;
        mod->foreign = f->next;
        pb2_device_detach_user_copy(mod->pb2, f->data);
        pb2_data_destroy(f->data);
        free(f);
    }
    for( int j = 0; j < gpu_device->num_exec_streams; j++ ) {
        parsec_gpu_exec_stream_t *exec_stream = gpu_device->exec_stream[j];
        free(exec_stream->name);
        PARSEC_OBJ_DESTRUCT(&exec_stream->infos);   /* runs the destructors of the per-stream objects */
    }
    free(gpu_device->exec_stream[0]);
    free(gpu_device->exec_stream);
    gpu_device->exec_stream = NULL;
    if( mod->pb2_index >= 0 && mod->pb2_index < PB2_MAX_DEVICES ) b200_modules[mod->pb2_index] = NULL;
    pb2_device_destroy(mod->pb2);
    mod->pb2 = NULL;
    mod->super.cuda_index = (uint8_t)-1;
    PARSEC_OBJ_DESTRUCT(&gpu_device->pending);
    PARSEC_OBJ_DESTRUCT(&gpu_device->gpu_mem_lru);
    PARSEC_OBJ_DESTRUCT(&gpu_device->gpu_mem_owned_lru);
    PARSEC_OBJ_DESTRUCT(&mod->copies);
    return PARSEC_SUCCESS;
}
