/*
 * parsec_b200.h -- C ABI of libparsec_b200.so: a B200-native (sm_100a) implementation of PaRSEC's
 * CUDA device task-execution path for tiled GEMM.  Plain pointers and sizes only; no C++ or torch types.
 *
 * Every entry point names the reference interface it replaces (paths relative to the reference tree,
 * uiuc-hpc/parsec-1).  INTEGRATION.md shows the overlay shim that maps the reference's symbols onto
 * these.  Return codes are the reference's own values: 0 = PARSEC_SUCCESS, negative = PARSEC_ERR_*
 * (parsec/include/parsec/constants.h:13-25); task hooks return PARSEC_HOOK_RETURN_* values
 * (parsec/runtime.h:140-146).
 */
#ifndef PARSEC_B200_H
#define PARSEC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PB2_SUCCESS 0
#define PB2_ERROR (-1)
#define PB2_ERR_OUT_OF_RESOURCE (-2)
#define PB2_ERR_NOT_FOUND (-3)
#define PB2_ERR_BAD_PARAM (-4)
#define PB2_ERR_NOT_SUPPORTED (-7)
#define PB2_ERR_VALUE_OUT_OF_BOUNDS (-8)
#define PB2_ERR_DEVICE (-10)

#define PB2_HOOK_RETURN_DONE 0
#define PB2_HOOK_RETURN_AGAIN (-1)
#define PB2_HOOK_RETURN_NEXT (-2)
#define PB2_HOOK_RETURN_DISABLE (-3)
#define PB2_HOOK_RETURN_ASYNC (-4)
#define PB2_HOOK_RETURN_ERROR (-5)

/* parsec/include/parsec/parsec_description_structures.h:62-66 */
#define PB2_FLOW_ACCESS_NONE 0x00
#define PB2_FLOW_ACCESS_READ 0x04
#define PB2_FLOW_ACCESS_WRITE 0x08
#define PB2_FLOW_ACCESS_RW 0x0c

/* parsec/data.h coherency / transfer-status values */
#define PB2_DATA_COHERENCY_INVALID 0x0
#define PB2_DATA_COHERENCY_OWNED 0x1
#define PB2_DATA_COHERENCY_EXCLUSIVE 0x2
#define PB2_DATA_COHERENCY_SHARED 0x4

/* parsec/mca/device/device.h:74-76 */
#define PB2_DEV_DATA_ADVICE_PREFETCH 0x01
#define PB2_DEV_DATA_ADVICE_PREFERRED_DEVICE 0x02
#define PB2_DEV_DATA_ADVICE_WARMUP 0x03

#define PB2_MAX_PARAM_COUNT 20 /* MAX_PARAM_COUNT, parsec/include/parsec/parsec_config_bottom.h:159 */
#define PB2_MAX_DEVICES 10     /* device 0 = host, 1 = "recursive" pseudo device, 2.. = GPUs (device.c:843) */

/* ------------------------------------------------------------------------------------------------
 * Tile kernels (piece 2).  Resolved by the reference's dyld= lookup: "<name>_sm100" first, then "<name>"
 * (cuda_find_incarnation, parsec/mca/device/cuda/device_cuda_module.c:175-240).
 * ---------------------------------------------------------------------------------------------- */

/* Replaces the cublasDgemm_v2(handle, transa, transb, m, n, k, &alpha, A, lda, B, ldb, &beta, C, ldc) call of
 * the reference's CUDA bodies (tests/dsl/dtd/dtd_test_simple_gemm.c:197-202, tests/dsl/ptg/cuda/nvlink.jdf:142,
 * stress.jdf:135).  Column-major, device pointers, `stream` is the cudaStream_t of parsec_body.stream.
 * Launches asynchronously; returns 0 or PB2_ERR_BAD_PARAM / PB2_ERR_DEVICE. */
int parsec_b200_dgemm_sm100(char transa, char transb, int m, int n, int k, double alpha, const double* A, int lda,
                            const double* B, int ldb, double beta, double* C, int ldc, void* stream);
int parsec_b200_dgemm(char transa, char transb, int m, int n, int k, double alpha, const double* A, int lda,
                      const double* B, int ldb, double beta, double* C, int ldc, void* stream);

/* Replaces cublasDsyrk_v2(handle, uplo, trans, n, k, &alpha, A, lda, &beta, C, ldc), the rank-k update of a tiled
 * Cholesky (the HERK / SYRK body DPLASMA supplies to this runtime): trans = 'N': C = alpha A A^T + beta C with A n x k,
 * 'T': C = alpha A^T A + beta C with A k x n; only the `uplo` ('L' / 'U') triangle of C is read and written.  Same
 * TMA + DMMA tile code as parsec_b200_dgemm, blocks on the other side of the diagonal skipped. */
int parsec_b200_dsyrk_sm100(char uplo, char trans, int n, int k, double alpha, const double* A, int lda, double beta,
                            double* C, int ldc, void* stream);
int parsec_b200_dsyrk(char uplo, char trans, int n, int k, double alpha, const double* A, int lda, double beta, double* C,
                      int ldc, void* stream);

/* Replaces cublasDtrsm_v2(handle, side, uplo, trans, diag, m, n, &alpha, T, ldt, B, ldb): B <- alpha * B * op(T)^-1
 * (side 'R') or alpha * op(T)^-1 * B (side 'L'), T triangular ('L' / 'U'), diag 'N' / 'U'.  The 128 x 128 diagonal
 * blocks of T are inverted in one launch, everything else is parsec_b200_dgemm on the block columns (rows).  Needs
 * scratch in device memory: parsec_b200_dtrsm_workspace(side, m, n) bytes, which a CUDA body takes from
 * parsec_gpu_pop_workspace (parsec/mca/device/device_gpu.c:168-246; pb2_device_pop_workspace here). */
size_t parsec_b200_dtrsm_workspace(char side, int m, int n);
int parsec_b200_dtrsm_sm100(char side, char uplo, char trans, char diag, int m, int n, double alpha, const double* T,
                            int ldt, double* B, int ldb, void* workspace, size_t workspace_bytes, void* stream);
int parsec_b200_dtrsm(char side, char uplo, char trans, char diag, int m, int n, double alpha, const double* T, int ldt,
                      double* B, int ldb, void* workspace, size_t workspace_bytes, void* stream);
/* The two halves on their own, for a DAG that solves many tiles against the same triangular tile (every TRSM of a
 * Cholesky panel): invert the diagonal blocks once -- Tinv takes ceil(n / 128) blocks of 128 x 128 doubles -- and hand
 * them to each solve; the workspace of the solve then only holds the m x n scratch (any size that
 * parsec_b200_dtrsm_workspace returns is enough).  diag is a property of the inversion; the solve does not need it. */
int parsec_b200_dtrtri_diag_sm100(char uplo, char diag, int n, const double* T, int ldt, double* Tinv, void* stream);
int parsec_b200_dtrsm_inv_sm100(char side, char uplo, char trans, int m, int n, double alpha, const double* T, int ldt,
                                const double* Tinv, double* B, int ldb, void* workspace, size_t workspace_bytes,
                                void* stream);

/* Replaces cusolverDnDpotrf(handle, uplo, n, A, lda, work, lwork, devInfo): Cholesky factorisation of one tile in
 * place, A = L L^T ('L') or U^T U ('U'); the other triangle is not touched.  Right-looking over 128-column blocks: the
 * diagonal block is factored and its factor inverted by one CTA, the panel is a parsec_b200_dgemm with that inverse,
 * the trailing update a parsec_b200_dsyrk.  workspace: parsec_b200_dpotrf_workspace(n) bytes of device memory;
 * dev_info: device int, zeroed by the caller, receives the 1-based index of the first non-positive pivot (may be NULL). */
size_t parsec_b200_dpotrf_workspace(int n);
int parsec_b200_dpotrf_sm100(char uplo, int n, double* A, int lda, void* workspace, size_t workspace_bytes, int* dev_info,
                             void* stream);
int parsec_b200_dpotrf(char uplo, int n, double* A, int lda, void* workspace, size_t workspace_bytes, int* dev_info,
                       void* stream);

/* BF16-in / FP32-accumulate tile GEMM on tcgen05 + TMEM fed by TMA (BASELINE config 5).  Same argument meaning as
 * cublasGemmEx(CUDA_R_16BF, CUDA_R_16BF, CUDA_R_32F, CUBLAS_COMPUTE_32F): A, B are bf16, C is fp32, column-major.
 * The reference has no bf16 body; the call shape follows its cublasDgemm_v2 bodies.  transa = transb = 'N' only
 * (PB2_ERR_NOT_SUPPORTED otherwise); A and B 16-byte aligned with lda, ldb multiples of 8.  C may be any fp32
 * matrix; when its base is 16-byte aligned and ldc is a multiple of 4 it is moved by TMA, which is the fast path. */
int parsec_b200_gemm_bf16_sm100(char transa, char transb, int m, int n, int k, float alpha, const void* A, int lda,
                                const void* B, int ldb, float beta, float* C, int ldc, void* stream);

/* Same kernel with fp32 operands consumed as TF32 (kind::tf32; the low 13 mantissa bits of A and B are ignored) and
 * fp32 accumulation: the cublasSgemm shape under CUBLAS_COMPUTE_32F_FAST_TF32.  Leading dimensions: multiples of 4. */
int parsec_b200_gemm_tf32_sm100(char transa, char transb, int m, int n, int k, float alpha, const float* A, int lda,
                                const float* B, int ldb, float beta, float* C, int ldc, void* stream);

/* Batches.  The GEMM calls one thread makes on `stream` between parsec_b200_batch_begin and parsec_b200_batch_end are
 * recorded and launched together: the bf16 / tf32 ones as one grouped persistent tcgen05 kernel (at _end, or every 16
 * calls), the fp64 ones (untransposed, aligned operands) as one grid holding the tiles of all of them (at _end, or
 * every 64 calls; a 256^3 fp64 tile task is four CTAs and costs the host more to launch than the GPU to run): the tile tasks a task runtime has ready at the same time share one launch, whose CTA pairs walk the output
 * tiles of all of them with the epilogue of one tile under the MMAs of the next.  The reference has no counterpart
 * (its bodies launch one cuBLAS call per task, nvlink.jdf:142); the device module brackets the bodies it submits
 * together with these calls, so bodies need no change.  Results are those of the single calls. */
int parsec_b200_batch_begin(void* stream);
int parsec_b200_batch_end(void* stream);

/* HBM heap: the job of zone_malloc_init / zone_malloc / zone_free / zone_in_use (parsec/utils/zone_malloc.h:41-63) --
 * carve the module's one device allocation into tiles, in units of `unit_size` bytes -- with a policy of its own: a
 * freed extent waits on a stack of extents of its size and is handed to the next request of that size (O(1) both
 * ways: a tile cache asks for one or two sizes); other requests are carved from the lowest run of the coalesced
 * free space, into which the stacks are merged when a request cannot be served otherwise.  The device module owns
 * one per GPU; these handles let a test drive the heap without a GPU (the base address is symbolic). */
void* pb2_zone_create(int max_segment, size_t unit_size);
void pb2_zone_destroy(void* zone);
int64_t pb2_zone_malloc(void* zone, size_t size); /* byte offset from the base, -1 = no run of free units is long enough */
int pb2_zone_free(void* zone, int64_t offset);
size_t pb2_zone_in_use(void* zone);

/* ------------------------------------------------------------------------------------------------
 * Data and copies (boundary kept from parsec/data.c, parsec/data_internal.h:35-81).
 * ---------------------------------------------------------------------------------------------- */
typedef struct pb2_data_s pb2_data_t;     /* parsec_data_t */
typedef struct pb2_device_s pb2_device_t; /* parsec_device_cuda_module_t */

/* parsec_data_create (parsec/data.c:484-522): one host copy, coherency OWNED, version 0, owner device 0. */
pb2_data_t* pb2_data_create(uint64_t key, uint32_t dc_id, void* host_ptr, uint32_t nb_elts);
/* The host tile is a window of a larger column-major matrix (the job of the user stage functions of
 * tests/dsl/ptg/cuda/stage_custom.jdf:28-109, which read dc->llm): `height` columns of `width_bytes`, `host_pitch_bytes`
 * apart on the host and `dev_pitch_bytes` apart in the packed device tile (dev_pitch_bytes * height <= nb_elts).  The
 * default stage-in / stage-out and the write-back of evicted tiles then use pitched 2D copies.  host_pitch_bytes = 0
 * restores the packed layout.  pb2_matrix_block_cyclic_lapack_new sets this on the tiles it creates. */
int pb2_data_set_host_layout(pb2_data_t* data, uint32_t width_bytes, uint32_t height, size_t host_pitch_bytes,
                             size_t dev_pitch_bytes);
/* parsec_data_destroy (parsec/data.c) -- detaches device copies first. */
void pb2_data_destroy(pb2_data_t* data);
/* field readers used by tests / shims: parsec_data_copy_t.version, .coherency_state, .readers, parsec_data_t.owner_device */
int pb2_data_copy_version(const pb2_data_t* data, int device_index, uint32_t* version);
int pb2_data_copy_coherency(const pb2_data_t* data, int device_index);
int pb2_data_copy_readers(const pb2_data_t* data, int device_index);
int pb2_data_owner_device(const pb2_data_t* data);
void* pb2_data_copy_ptr(const pb2_data_t* data, int device_index); /* PARSEC_DATA_COPY_GET_PTR */

/* Overlay support (overlay/device_cuda_b200_shim.c, the file compiled in place of
 * parsec/mca/device/cuda/device_cuda_module.c): a parsec_data_t keeps a pb2 mirror.  `user` is the shim's
 * parsec_data_copy_t of the device; pb2_data_sync_host copies what the runtime's host copy says (address, version,
 * coherency or -1 = keep) into the mirror before a task is submitted, where the reference's stage-in reads
 * task->data[i].data_in directly (device_cuda_module.c:1260-1300); pb2_data_get_state reads the device side back so the
 * shim can write it into the real parsec_data_copy_t (version, coherency_state, readers, device_private) and
 * parsec_data_t.owner_device. */
void pb2_data_set_user(pb2_data_t* data, void* user);
void* pb2_data_user(const pb2_data_t* data);
int pb2_data_sync_host(pb2_data_t* data, void* host_ptr, uint32_t version, int coherency);
typedef struct pb2_data_state_s {
    void* dev_ptr;
    uint32_t dev_version, host_version;
    int32_t readers;
    uint8_t dev_coherency, host_coherency, has_copy;
    int8_t owner_device; /* pb2 device index, 0 = host, -1 = none */
} pb2_data_state_t;
int pb2_data_get_state(const pb2_data_t* data, int device_index, pb2_data_state_t* out);

/* The ownership / coherency state machine on bare copies (no device needed): parsec_data_copy_attach / _detach and
 * parsec_data_start / end_transfer_ownership_to_copy (parsec/data.c:163-202, :299-433).  A CPU test replays operation
 * sequences recorded from the reference's own data.c through these (tests/golden/ownership_trace.json). */
int pb2_data_test_attach(pb2_data_t* data, int device_index);
int pb2_data_test_detach(pb2_data_t* data, int device_index);
int pb2_data_test_start_transfer(pb2_data_t* data, int device_index, int access); /* source device or -1 */
void pb2_data_test_end_transfer(pb2_data_t* data, int device_index, int access);
void pb2_data_test_set(pb2_data_t* data, int device_index, int coherency, int version, int readers); /* -1 = keep */

/* ------------------------------------------------------------------------------------------------
 * Device module (piece 1): replaces parsec/mca/device/cuda/device_cuda_module.c and device_cuda_component.c.
 * ---------------------------------------------------------------------------------------------- */
typedef struct pb2_device_params_s {
    int memory_use_percent;       /* device_cuda_memory_use (95)            device_cuda_component.c:187 */
    int64_t memory_block_size;    /* device_cuda_memory_block_size (512 KiB) :183 */
    int64_t memory_number_of_blocks; /* device_cuda_memory_number_of_blocks (-1 = from percent) :191 */
    int max_streams;              /* device_cuda_max_streams :199 (reference default 6 = h2d + d2h + 4 exec; default here 18:
                                     a task is spread over as many exec streams as its tile size needs, 4..16) */
    int max_number_of_ejected_data; /* device_cuda_max_number_of_ejected_data (20) :195 */
    int max_inflight_tasks;       /* tasks enqueued on the GPU but not completed (reference: 4 per stream, device_gpu.h:22) */
    int trace_events;             /* record pb2_event_t for every data-state transition */
    int profile_launches;         /* time every body launch with CUDA events on its exec stream (bench.py roofline) */
    int sort_pending_tasks;       /* device_cuda_sort_pending_tasks (0): among the first 10 pending tasks take the one
                                     that needs the fewest new tiles (parsec_gpu_sort_pending_list, device_gpu.c:117-166) */
    int prefetch_deferred;        /* 1: a PREFETCH advice given while nobody manages the device is only queued, like the
                                     reference's push to `pending` (device_cuda_module.c:1668); the next manager runs it.
                                     0 (default): the adviser drives the device until the prefetch is done */
    int use_priorities;           /* 1: the pending list hands out the task with the highest priority first (oldest first
                                     among equals), as the reference's priority-ordered lists do (PARSEC_GPU_USE_PRIORITIES,
                                     device_cuda_module.c:1683-1692); 0 (default): first come, first served */
    int batch_bodies;             /* > 1: the manager gathers up to this many ready tasks, stages them in, and runs their
                                     bodies back to back on one exec stream between parsec_b200_batch_begin and _end, so
                                     that bodies calling the library's bf16 / tf32 GEMM share one grouped launch; every task
                                     keeps its own completion event and stage-out.  0 (default): one task at a time, spread
                                     over the exec streams (what the FP64 kernel wants) */
} pb2_device_params_t;
void pb2_device_params_default(pb2_device_params_t* p);
/* The reference's MCA parameter names, as they reach a process in its environment (PARSEC_MCA_<name>,
 * parsec/utils/mca_param.c): device_cuda_memory_use, _memory_block_size, _memory_number_of_blocks, _max_streams,
 * _max_number_of_ejected_data, _sort_pending_tasks (device_cuda_component.c:163-206) overwrite the matching fields;
 * returns how many were found.  pb2_device_create(idx, NULL, ..) applies the defaults, then these.  Inside a PaRSEC
 * build the overlay shim passes the values the runtime itself parsed (files, environment, --mca on the command line). */
int pb2_device_params_from_env(pb2_device_params_t* p);
/* device_cuda_enabled (-1 = all GPUs, n = the first n, 0 = none) and device_cuda_mask (bit i), same source:
 * 1 when GPU `cuda_index` may be used.  pb2_device_create(idx, NULL, ..) refuses a disabled GPU (PB2_ERR_NOT_SUPPORTED). */
int pb2_device_is_enabled(int cuda_index);

/* parsec_cuda_module_init (device_cuda_module.c:325-547): streams, events, one cudaMalloc carved by the zone
 * allocator.  The device gets index 2 + (number of devices created before it), as in parsec_mca_device_add. */
int pb2_device_create(int cuda_index, const pb2_device_params_t* params, pb2_device_t** out);
/* parsec_cuda_module_fini (:553-605) */
int pb2_device_destroy(pb2_device_t* dev);
int pb2_device_index(const pb2_device_t* dev);
int pb2_device_cuda_index(const pb2_device_t* dev);
/* parsec_cuda_memory_register / _unregister (:117-172): cudaHostRegister(ptr, len, Portable) */
int pb2_device_memory_register(pb2_device_t* dev, void* ptr, size_t len);
int pb2_device_memory_unregister(pb2_device_t* dev, void* ptr);
/* memory_release = parsec_cuda_flush_lru (:803-820): drop every clean copy from the LRU */
int pb2_device_memory_release(pb2_device_t* dev);
/* parsec_cuda_data_advise (:1611-1679) */
int pb2_device_data_advise(pb2_device_t* dev, pb2_data_t* data, int advice);

/* Overlay support.  copy_released: the module has just given back the HBM of `data`'s device copy (LRU eviction,
 * memory_release, destroy) -- where the reference detaches and releases the parsec_data_copy_t
 * (device_cuda_module.c:1090-1110); called on the manager thread.  data_alive: asked before a dirty tile is written
 * back to its host copy; 0 = the runtime has destroyed the data, the tile is dropped instead. */
typedef struct pb2_device_hooks_s {
    void (*copy_released)(pb2_device_t* dev, pb2_data_t* data, void* user);
    int (*data_alive)(pb2_device_t* dev, pb2_data_t* data, void* user);
} pb2_device_hooks_t;
int pb2_device_set_hooks(pb2_device_t* dev, const pb2_device_hooks_t* hooks);
/* A device copy the USER allocated and attached by hand (cudaMalloc + parsec_data_copy_attach +
 * parsec_data_transfer_ownership_to_copy, tests/dsl/ptg/cuda/nvlink_wrapper.c:189-206).  The module hands its address to
 * task bodies and otherwise leaves it alone, as the reference does for copies without PARSEC_DATA_FLAG_PARSEC_OWNED
 * (device_cuda_module.c:2161-2163, :2215, :2399): no staging, no eviction, no write-back, never freed by the module.
 * A second call updates address / version / coherency.  PB2_ERR_EXISTS when the module already caches the tile itself. */
int pb2_device_attach_user_copy(pb2_device_t* dev, pb2_data_t* data, void* device_ptr, uint32_t version, int coherency);
int pb2_device_detach_user_copy(pb2_device_t* dev, pb2_data_t* data);
/* Give back the HBM of `data`'s copy on this device (the runtime has destroyed the data; what the reference's LRU flush
 * does to detached copies, device_cuda_module.c:803-820).  The device must be idle or held by the caller (pb2_device_hold). */
int pb2_device_drop_copy(pb2_device_t* dev, pb2_data_t* data);
/* The module's CUDA streams, numbered like parsec_device_gpu_module_t.exec_stream[] (device_cuda_module.c:414-420):
 * 0 = h2d, 1 = d2h, 2.. = exec.  pb2_device_stream_index maps the cuda_stream a submit / stage callback received
 * back to that number (-1 = not one of ours). */
int pb2_device_nb_streams(const pb2_device_t* dev);
void* pb2_device_stream(const pb2_device_t* dev, int index);
int pb2_device_stream_index(const pb2_device_t* dev, const void* cuda_stream);

/* statistics of parsec_device_module_t (parsec/mca/device/device.h:132-137) plus cache counters */
typedef struct pb2_device_stats_s {
    uint64_t executed_tasks, data_in_from_host, data_in_from_peers, data_out_to_host;
    uint64_t required_data_in, required_data_out, nb_data_faults;
    uint64_t nb_evictions, nb_w2r_tasks, nb_w2r_tiles, nb_stalls, kernel_launches;
    uint64_t hbm_bytes_total, hbm_bytes_in_use;
    uint64_t host_submit_ns, host_retire_ns; /* manager-thread time spent submitting / retiring tasks */
    uint64_t nb_cleanouts;                   /* tasks whose stage-in was taken back (parsec_cuda_kernel_cleanout) */
    uint64_t nb_batches;                     /* batched submissions (params.batch_bodies > 1) */
    uint64_t data_in_from_device[PB2_MAX_DEVICES]; /* bytes staged in, by source device (parsec_device_module_t.data_in_from_device,
                                                device.h:130): [0] = host, [i >= 2] = the peer GPU with that device index */
} pb2_device_stats_t;
int pb2_device_get_stats(pb2_device_t* dev, pb2_device_stats_t* out, int reset);
/* the live counters, for a caller on the manager thread that folds them into its own after every task (the overlay
 * shim does, into parsec_device_module_t); hbm_bytes_in_use is only refreshed by pb2_device_get_stats */
const pb2_device_stats_t* pb2_device_stats_live(const pb2_device_t* dev);

/* Data-state trace (enabled by params.trace_events); same layout as the oracle's orc_event_t. */
typedef struct pb2_event_s {
    uint32_t dc, key;
    int32_t kind, k;
    uint32_t gpu_version, cpu_version;
    uint8_t gpu_coh, cpu_coh;
    int8_t owner_device, transfer;
} pb2_event_t;
int64_t pb2_device_get_events(pb2_device_t* dev, pb2_event_t* out, int64_t max_events);
/* params.profile_launches: per-launch durations in milliseconds, measured between consecutive completion events
 * of the same exec stream (a launch starts when its predecessor on that stream ends).  Returns the number of
 * samples and clears them; *nb_exec_streams receives how many launches are resident at a time. */
int64_t pb2_device_get_launch_ms(pb2_device_t* dev, float* out, int64_t max_samples, int* nb_exec_streams);

/* ------------------------------------------------------------------------------------------------
 * GPU tasks.  parsec_gpu_task_t (parsec/mca/device/device_gpu.h:79-113) as the generated hook fills it
 * (parsec/interfaces/ptg/ptg-compiler/jdf2c.c:6775-6897).  calloc'd by the caller, owned by the module after
 * submission and free()d by it at completion (device_cuda_module.c:2741).
 * ---------------------------------------------------------------------------------------------- */
typedef struct pb2_gpu_task_s pb2_gpu_task_t;
/* gpu_task->submit: enqueue the body on `cuda_stream`, return PB2_HOOK_RETURN_DONE (jdf2c.c:6608-6724) */
typedef int (*pb2_submit_fn)(pb2_device_t* dev, pb2_gpu_task_t* task, void* cuda_stream);
/* parsec_stage_in_function_t / parsec_stage_out_function_t (device_gpu.h:58-78) */
typedef int (*pb2_stage_fn)(pb2_gpu_task_t* task, uint32_t flow_mask, void* cuda_stream);
/* what the module calls where the reference calls __parsec_complete_execution(es, gpu_task->ec)
 * (device_cuda_module.c:2733-2734); status is 0 or a negative PB2_HOOK_RETURN_* */
typedef void (*pb2_complete_fn)(pb2_gpu_task_t* task, int status, void* cookie);

struct pb2_gpu_task_s {
    int32_t task_type;  /* 0 = kernel, 1 = prefetch (device_gpu.h PARSEC_GPU_TASK_TYPE_*) */
    int32_t pushout;    /* bit i: flow i goes back to the host when the task completes */
    pb2_submit_fn submit;
    pb2_stage_fn stage_in;  /* NULL = pb2_default_stage_in */
    pb2_stage_fn stage_out; /* NULL = pb2_default_stage_out */
    void* ec;               /* the runtime's task, opaque here */
    int32_t priority;
    int32_t nb_flows;
    pb2_data_t* flow_data[PB2_MAX_PARAM_COUNT];
    uint8_t flow_access[PB2_MAX_PARAM_COUNT];   /* PB2_FLOW_ACCESS_* */
    uint32_t flow_nb_elts[PB2_MAX_PARAM_COUNT]; /* bytes to stage, as gpu_task->flow_nb_elts */
    void* flow_dev_ptr[PB2_MAX_PARAM_COUNT];    /* set by the module: data_out->device_private */
    void* flow_host_ptr[PB2_MAX_PARAM_COUNT];   /* set by the module: data_in->device_private (host copy) */
    pb2_complete_fn complete;
    void* complete_cookie;
    int64_t locals[8]; /* task locals for the body (m, n, k, ...) */
    int32_t trace_dc[PB2_MAX_PARAM_COUNT]; /* collection id written to the event trace */
    void* reserved[4]; /* module private */
};

/* parsec_cuda_kernel_scheduler(es, gpu_task, which_gpu) (device_cuda_module.c:2536-2763).  Any thread may call;
 * the first caller becomes the device's manager and drives push -> exec -> pop until the device drains.
 * Returns PB2_HOOK_RETURN_ASYNC (completion is signalled through task->complete) or PB2_HOOK_RETURN_DISABLE. */
int pb2_kernel_scheduler(pb2_device_t* dev, pb2_gpu_task_t* task);
/* Two-step form of the manager election for a caller that queues a batch before it starts driving the device:
 * pb2_device_hold takes the manager seat (returns 1; 0 if another thread already manages), every
 * pb2_kernel_scheduler call that follows only queues, pb2_device_manage then drives the device until it
 * drains.  Same mutex protocol as device_cuda_module.c:2573-2589 / :2742-2753, split in two calls. */
int pb2_device_hold(pb2_device_t* dev);
int pb2_device_manage(pb2_device_t* dev);
/* Per-stream scratch for task bodies: parsec_gpu_pop_workspace / parsec_gpu_push_workspace /
 * parsec_gpu_free_workspace (parsec/mca/device/device_gpu.c:168-246).  To be called from a submit callback with the
 * cuda_stream it was handed.  The first pop on a stream carves PB2_GPU_MAX_WORKSPACE buffers of `size` bytes from the
 * HBM heap (evicting clean tiles if it must) and keeps them until pb2_device_free_workspace; a pop takes the top of
 * that stream's stack, a push gives the last one back.  Work queued on one stream runs in order, which is what makes a
 * buffer handed back right after the launch safe to hand to the next body on the same stream.  Returns NULL when the
 * stream is not one of the device's exec streams, the stack is empty, `size` exceeds the size of the first pop, or the
 * heap has no room. */
#define PB2_GPU_MAX_WORKSPACE 2
void* pb2_device_pop_workspace(pb2_device_t* dev, void* cuda_stream, size_t size);
int pb2_device_push_workspace(pb2_device_t* dev, void* cuda_stream);
int pb2_device_free_workspace(pb2_device_t* dev);
/* parsec_default_cuda_stage_in / _stage_out (device_cuda_module.c:1179-1209, :1220-1250) */
int pb2_default_stage_in(pb2_gpu_task_t* task, uint32_t flow_mask, void* cuda_stream);
int pb2_default_stage_out(pb2_gpu_task_t* task, uint32_t flow_mask, void* cuda_stream);

/* ------------------------------------------------------------------------------------------------
 * 2D block-cyclic tiled matrix: parsec_matrix_block_cyclic_t
 * (parsec/data_dist/matrix/two_dim_rectangle_cyclic.h:24-83); tile storage (PARSEC_MATRIX_TILE).
 * ---------------------------------------------------------------------------------------------- */
typedef struct pb2_matrix_block_cyclic_s pb2_matrix_block_cyclic_t;
#define PB2_MATRIX_DOUBLE 0 /* parsec_matrix_type_t PARSEC_MATRIX_DOUBLE */
#define PB2_MATRIX_FLOAT 1
#define PB2_MATRIX_BF16 2   /* extension: the reference enum has no 16-bit float (matrix.h:25-32) */

/* parsec_matrix_block_cyclic_init (two_dim_rectangle_cyclic.c:109-236): same argument order. */
pb2_matrix_block_cyclic_t* pb2_matrix_block_cyclic_new(int mtype, int myrank, int mb, int nb, int lm, int ln, int i,
                                                       int j, int m, int n, int P, int Q, int kp, int kq, int ip,
                                                       int jq);
/* parsec_matrix_block_cyclic_lapack_init (two_dim_rectangle_cyclic.c:61-107): PARSEC_MATRIX_LAPACK storage.  dc->mat is
 * a column-major local matrix of mloc x nloc elements (ld = mloc), not padded to whole tiles; a tile is a window of
 * it.  The device copy of a tile stays packed (ld = mb): the module's default stage-in / stage-out and its write-back
 * of evicted tiles move such tiles with pitched 2D copies, so no custom stage functions are needed (what
 * tests/dsl/ptg/cuda/stage_custom.jdf:28-109 does by hand), and short edge tiles move only the rows and columns the
 * matrix has.  pb2_matrix_fill_lcg* do not support this storage. */
pb2_matrix_block_cyclic_t* pb2_matrix_block_cyclic_lapack_new(int mtype, int myrank, int mb, int nb, int lm, int ln,
                                                              int i, int j, int m, int n, int P, int Q, int kp, int kq,
                                                              int ip, int jq, int mloc, int nloc);
void pb2_matrix_block_cyclic_free(pb2_matrix_block_cyclic_t* dc);
/* dc->mat = user memory holding nb_local_tiles * mb * nb elements */
int pb2_matrix_set_mat(pb2_matrix_block_cyclic_t* dc, void* mat);
int64_t pb2_matrix_nb_local_tiles(const pb2_matrix_block_cyclic_t* dc);
int pb2_matrix_mt(const pb2_matrix_block_cyclic_t* dc);
int pb2_matrix_nt(const pb2_matrix_block_cyclic_t* dc);
/* twoDBC_rank_of (:258-286), local tile position (:333-349), data_of (:368-412), data_key (matrix.c); with kp or
 * kq > 1 their k-cyclic counterparts twoDBC_kcyclic_rank_of (:532-566) and twoDBC_kcyclic_data_of (:634-680) */
int pb2_matrix_rank_of(const pb2_matrix_block_cyclic_t* dc, int m, int n);
int64_t pb2_matrix_position(const pb2_matrix_block_cyclic_t* dc, int m, int n);
uint64_t pb2_matrix_data_key(const pb2_matrix_block_cyclic_t* dc, int m, int n);
/* element offset of tile (m, n) from dc->mat, either storage: position * mb * nb, or (local_m * mb) + (local_n * nb) *
 * mloc for LAPACK storage (twoDBC_data_of :397-411) */
int64_t pb2_matrix_tile_offset(const pb2_matrix_block_cyclic_t* dc, int m, int n);
pb2_data_t* pb2_matrix_data_of(pb2_matrix_block_cyclic_t* dc, int m, int n);
/* twoDBC_memory_register (:39-49) -> pb2_device_memory_register over the local matrix */
int pb2_matrix_register_memory(pb2_matrix_block_cyclic_t* dc, pb2_device_t* dev);
int pb2_matrix_unregister_memory(pb2_matrix_block_cyclic_t* dc, pb2_device_t* dev);

/* Synthetic inputs: the reference's LCG tile initialiser (tests/dsl/dtd/dtd_test_simple_gemm.c:43-91), run on the
 * GPU for every local tile of `dc` and written to dc->mat (fp64) or, for PB2_MATRIX_BF16 / FLOAT, rounded from
 * the same values.  Used by bench.py; bit-exact with the CPU generator. */
int pb2_matrix_fill_lcg(pb2_matrix_block_cyclic_t* dc, unsigned seed, pb2_device_t* dev);
/* Same result as pb2_matrix_fill_lcg, for callers that must not allocate device memory: code that runs after
 * pb2_device_create has handed 95 % of HBM to the tile heap (a staging cudaMalloc may then fail), or while the
 * pipeline is busy (cudaMalloc / cudaFree synchronise the device).  Has no counterpart in the reference.
 * Tiles are staged through a 16 MiB buffer that is part of the library image on every GPU that loads it, so one tile
 * may be at most 16 MiB; larger tiles return PB2_ERR_VALUE_OUT_OF_BOUNDS (-8).  Plain, k-cyclic and origin-shifted
 * grids are handled; a sub-matrix view (i, j != 0, or fewer tiles than the parent stores) and LAPACK storage return
 * PB2_ERR_NOT_SUPPORTED (-7) and leave the matrix untouched.  The first call on a device resolves
 * the buffer's address and should come before any stream capture.  All calls on a device share that buffer (it keeps
 * the previous call's bytes; nothing needs it zeroed), so calls that overlap on one GPU race: serialise them. */
int pb2_matrix_fill_lcg_noalloc(pb2_matrix_block_cyclic_t* dc, unsigned seed, pb2_device_t* dev);

/* ------------------------------------------------------------------------------------------------
 * Tiled GEMM taskpool: the PTG dgemm_NN DAG (READ_A, READ_B, GEMM(m,n,k) with the C chain over k) unfolded over
 * block-cyclic A, B, C and executed through pb2_kernel_scheduler, i.e. what
 * parsec_context_add_taskpool + parsec_context_start + parsec_context_wait (parsec/scheduling.c:678-806) do
 * for the generated parsec_dgemm_NN_new() taskpool.
 * ---------------------------------------------------------------------------------------------- */
typedef struct pb2_taskpool_s pb2_taskpool_t;
typedef struct pb2_gemm_stats_s {
    int64_t nb_gemm, nb_read_a, nb_read_b; /* executed task counts per task class on this rank */
    double seconds;                        /* wall time of start -> wait */
} pb2_gemm_stats_t;

pb2_taskpool_t* pb2_gemm_taskpool_new(pb2_matrix_block_cyclic_t* A, pb2_matrix_block_cyclic_t* B,
                                      pb2_matrix_block_cyclic_t* C, pb2_device_t* dev);
/* knobs (before run): "window" = C-tile chains in flight (0 = all, -1 = default: 128, or 16 for tensor-core
 * bodies); "bcast_lookahead" = k-steps the broadcast plan is issued ahead of the GEMMs (default 4, 0 = whole plan at
 * start).  PB2_ERR_NOT_FOUND for an unknown key. */
int pb2_gemm_taskpool_set(pb2_taskpool_t* tp, const char* key, int64_t value);
/* ------------------------------------------------------------------------------------------------
 * Intra-box tile data plane (piece 3): NCCL broadcasts along the rows / columns of the P x Q grid.  Replaces, for
 * tile payloads, remote_dep_mpi_put_start / get_start / get_end_cb (parsec/remote_dep_mpi.c:1593, :1963, :2082) and
 * mpi_no_thread_put / get (parsec/parsec_mpi_funnelled.c:793, :896); the broadcast topology knob
 * runtime_comm_coll_bcast (parsec/remote_dep.c:272-291) has no equivalent because NVSwitch is uniform.
 * ---------------------------------------------------------------------------------------------- */
typedef struct pb2_comm_s pb2_comm_t;
/* ncclGetUniqueId: 128 bytes, to be published by column 0 of each process row and row 0 of each process column */
int pb2_comm_unique_id(void* out128);
/* one communicator per process row (size Q, my index = rank mod Q) and per process column (size P, rank / Q) */
int pb2_comm_create(int cuda_index, int P, int Q, int rank, const void* row_id, const void* col_id, pb2_comm_t** out);
int pb2_comm_destroy(pb2_comm_t* comm);
/* one in-place tile broadcast: which = 0 along my process row (A tiles), 1 along my process column (B tiles) */
int pb2_comm_bcast(pb2_comm_t* comm, int which, void* device_buf, size_t bytes, int root);
/* The canonical per-rank sequence of broadcasts for C(MT x NT) += A(MT x KT) * B(KT x NT) (host logic, no CUDA):
 * for k ascending, the row broadcasts A(m,k) of my process row (m ascending) then the column broadcasts B(k,n) of
 * my process column (n ascending); root = k mod Q resp. k mod P, the owner under twoDBC_rank_of.  Returns the
 * number of operations (call with out == NULL to size the buffer). */
typedef struct pb2_bcast_op_s {
    int32_t which, k, idx, root, i_am_root;
} pb2_bcast_op_t;
int64_t pb2_comm_plan(int P, int Q, int rank, int MT, int NT, int KT, pb2_bcast_op_t* out, int64_t max_ops);
/* The same for k-cyclic descriptors and a shifted grid origin (twoDBC_kcyclic_rank_of,
 * two_dim_rectangle_cyclic.c:532-566): tile rows of A and C dealt in super-tiles of kp, tile columns of B and C in
 * super-tiles of kq, the k dimension in super-tiles of a_kq (columns of A) and b_kp (rows of B), origin (ip, jq).
 * pb2_comm_plan is the case (1, 1, 1, 1, 0, 0). */
int64_t pb2_comm_plan_ex(int P, int Q, int rank, int MT, int NT, int KT, int kp, int kq, int a_kq, int b_kp, int ip,
                         int jq, pb2_bcast_op_t* out, int64_t max_ops);
/* give a P x Q taskpool its communicator (must match the descriptors' grid and rank) */
int pb2_gemm_taskpool_set_comm(pb2_taskpool_t* tp, pb2_comm_t* comm);

/* add_taskpool + start + wait; C is valid on the host afterwards */
int pb2_gemm_taskpool_run(pb2_taskpool_t* tp, pb2_gemm_stats_t* stats);
void pb2_gemm_taskpool_free(pb2_taskpool_t* tp);

/* ------------------------------------------------------------------------------------------------
 * Comm-engine transport for the ranks of one box (piece 3).
 * What the reference does with MPI inside parsec/parsec_mpi_funnelled.c -- active messages on persistent receives
 * (:621-703, :994-1008, :1103-1154), put/get as Isend/Irecv pairs plus a completion message (:793-991), the
 * barrier (:1200) -- done over one shared-memory segment per job: a single-producer / single-consumer ring per
 * ordered pair of ranks, and put as ONE copy by the origin into the target's address space (process_vm_writev;
 * through the ring where the kernel refuses that).  Per-pair FIFO order.  Host memory.  The parsec_comm_engine_t
 * vtable (parsec/parsec_comm_engine.h:161-182) over these is overlay/comm_engine_b200.c.
 * ---------------------------------------------------------------------------------------------- */
typedef struct pb2_ce_s pb2_ce_t;
typedef struct {
    void* base;
    size_t len;
} pb2_ce_iov_t;
typedef struct {
    uint64_t sent_msgs, recv_msgs; /* active messages */
    uint64_t put_bytes;            /* bytes moved by pb2_ce_put */
    uint64_t staged_bytes;         /* of which through the ring (0 when single-copy put is available) */
    int32_t single_copy;           /* 1: process_vm_writev in use */
} pb2_ce_stats_t;
typedef void (*pb2_ce_recv_fn)(void* user, int src, uint64_t tag, const void* msg, size_t len);

/* Collective over the `world` ranks of job `job` (any string unique to the launch; the segment is /dev/shm/pb2ce.<job>).
 * ring_bytes: capacity of each pair's ring, a power of two >= 4096, 0 = 1 MiB.  Returns when every rank has attached
 * (PB2_ERR_NOT_FOUND after 120 s). */
int pb2_ce_attach(const char* job, int rank, int world, size_t ring_bytes, pb2_ce_t** ce);
int pb2_ce_detach(pb2_ce_t* ce); /* the last rank out removes the segment */
int pb2_ce_rank(const pb2_ce_t* ce);
int pb2_ce_world(const pb2_ce_t* ce);
size_t pb2_ce_max_message(const pb2_ce_t* ce); /* largest len pb2_ce_send accepts (ring_bytes / 4 - 16) */
void pb2_ce_mark_failed(pb2_ce_t* ce);         /* a rank that aborts: the others' waits return PB2_ERROR */
/* send_am (parsec_comm_engine.h:122-125).  Any thread.  PB2_HOOK_RETURN_AGAIN (-1) when dst's ring is full: the
 * caller keeps the message and retries after dst has polled. */
int pb2_ce_send(pb2_ce_t* ce, int dst, uint64_t tag, const void* msg, size_t len);
/* progress (:127): delivers at most max_msgs (0 = all) pending messages to fn, in per-source FIFO order; msg is valid
 * during the call only.  One thread at a time; fn may call pb2_ce_send.  Returns the number delivered. */
int pb2_ce_poll(pb2_ce_t* ce, pb2_ce_recv_fn fn, void* user, int max_msgs);
/* While a put that goes through the ring waits for room it delivers this rank's incoming messages to fn (two ranks
 * putting to each other would otherwise wait for each other).  Call pb2_ce_put outside pb2_ce_poll callbacks. */
void pb2_ce_set_wait_poll(pb2_ce_t* ce, pb2_ce_recv_fn fn, void* user);
/* put (:97-105): copies local[0..nlocal) of this process into remote[0..nremote) of rank dst (addresses valid in dst's
 * address space; same total length).  Complete on return; a message sent to dst afterwards is seen after the data. */
int pb2_ce_put(pb2_ce_t* ce, int dst, const pb2_ce_iov_t* local, int nlocal, const pb2_ce_iov_t* remote, int nremote);
/* get (:107-120): the mirror image, remote[] of rank src into local[]; PB2_ERR_NOT_SUPPORTED (-7) where the box has no
 * cross-memory attach -- the caller then asks the peer to put (overlay/comm_engine_b200.c does). */
int pb2_ce_get(pb2_ce_t* ce, int src, const pb2_ce_iov_t* local, int nlocal, const pb2_ce_iov_t* remote, int nremote);
int pb2_ce_barrier(pb2_ce_t* ce);                                            /* sync (:146) */
int pb2_ce_allreduce_i64(pb2_ce_t* ce, int64_t* vals, int n, int op);        /* n <= 16; op 0 max, 1 sum, 2 min */
int pb2_ce_allgather(pb2_ce_t* ce, const void* mine, size_t bytes, void* all); /* bytes <= 128 per rank */
int pb2_ce_stats(const pb2_ce_t* ce, pb2_ce_stats_t* out);

/* ------------------------------------------------------------------------------------------------
 * DTD front-end: the second caller of the same device boundary (GPU tasks only).
 * parsec_dtd_taskpool_new + parsec_context_add_taskpool/start; parsec_dtd_insert_task_with_task_class
 * (parsec/interfaces/dtd/insert_function.c) with parsec_dtd_gpu_task_submit (:2359-2406);
 * parsec_dtd_data_flush_all + parsec_dtd_taskpool_wait + parsec_context_wait.
 * Dependencies are inferred from the insertion order and the access modes of the data each task touches.
 * ---------------------------------------------------------------------------------------------- */
typedef struct pb2_dtd_taskpool_s pb2_dtd_taskpool_t;
pb2_dtd_taskpool_t* pb2_dtd_taskpool_new(pb2_device_t* dev);
/* access[i]: PB2_FLOW_ACCESS_READ / _WRITE / _RW (PARSEC_INPUT / OUTPUT / INOUT); pushout_mask bit i: PARSEC_PUSHOUT on
 * flow i (insert_function.h:70); locals: the PARSEC_VALUE arguments (at most 8) copied into pb2_gpu_task_t.locals */
int pb2_dtd_insert_task(pb2_dtd_taskpool_t* tp, pb2_submit_fn gpu_body, int priority, int nb_flows,
                        pb2_data_t* const* data, const uint8_t* access, int32_t pushout_mask, const int64_t* locals,
                        int nb_locals);
int pb2_dtd_taskpool_wait(pb2_dtd_taskpool_t* tp, int64_t* nb_executed);
void pb2_dtd_taskpool_free(pb2_dtd_taskpool_t* tp);
/* gemm_kernel_cuda of tests/dsl/dtd/dtd_test_simple_gemm.c:165-216: flows (A, B, C), locals (m, n, k, mb, nb, kb) */
int pb2_dtd_gemm_body(pb2_device_t* dev, pb2_gpu_task_t* task, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif
