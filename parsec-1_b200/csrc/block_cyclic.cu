// parsec_matrix_block_cyclic_t mirror (tile storage) and the GPU version of the reference's synthetic
// input generator.  parsec/data_dist/matrix/matrix.c:56-125, grid_2Dcyclic.c:18-37,
// two_dim_rectangle_cyclic.c:109-236 (init), :258-286 (rank_of), :333-349 (position), :368-412 (data_of).
#include <cuda_bf16.h>
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>

#include "pb2_internal.h"
#include "pb2_types.h"

struct pb2_matrix_block_cyclic_s {
    int mtype, elem_size;
    int mb, nb, lm, ln, i, j, m, n, mt, nt, lmt, lnt, llm, lln, bsiz;
    int P, Q, kp, kq, ip, jq, rrank, crank, myrank;
    int nb_elem_r, nb_elem_c;
    int lapack = 0;  // 0 = PARSEC_MATRIX_TILE storage, 1 = PARSEC_MATRIX_LAPACK (column-major local matrix, ld = llm)
    int64_t nb_local_tiles;
    void* mat;
    uint32_t dc_id;
    std::vector<pb2_data_s*> data_map;
    std::mutex lock;
};

static std::atomic<uint32_t> g_next_dc_id{0};

static int owned_along(int first_block, int k, int procs, int total) {
    int have = 0;
    for (int at = first_block * k; at < total; at += procs * k) have += std::min(k, total - at);
    return have;
}

static pb2_matrix_block_cyclic_t* make_descriptor(int mtype, int myrank, int mb, int nb, int lm, int ln, int i, int j, int m,
                                                  int n, int P, int Q, int kp, int kq, int ip, int jq, int lapack, int mloc,
                                                  int nloc) {
    if (mb <= 0 || nb <= 0 || lm <= 0 || ln <= 0 || m <= 0 || n <= 0 || P <= 0 || Q <= 0 || kp <= 0 || kq <= 0 ||
        myrank < 0 || myrank >= P * Q || i < 0 || j < 0 || i + m > lm || j + n > ln)
        return nullptr;
    int es = mtype == PB2_MATRIX_DOUBLE ? 8 : mtype == PB2_MATRIX_FLOAT ? 4 : mtype == PB2_MATRIX_BF16 ? 2 : 0;
    if (!es) return nullptr;
    auto* dc = new pb2_matrix_block_cyclic_s();
    dc->mtype = mtype;
    dc->elem_size = es;
    dc->mb = mb; dc->nb = nb; dc->bsiz = mb * nb;
    dc->lmt = (lm + mb - 1) / mb;
    dc->lnt = (ln + nb - 1) / nb;
    // tile storage pads the matrix to whole tiles, LAPACK storage does not (matrix.c:103-107)
    dc->lm = lapack ? lm : dc->lmt * mb;
    dc->ln = lapack ? ln : dc->lnt * nb;
    dc->lapack = lapack;
    dc->i = i; dc->j = j; dc->m = m; dc->n = n;
    dc->mt = (i + m - 1) / mb - i / mb + 1;
    dc->nt = (j + n - 1) / nb - j / nb + 1;
    dc->P = P; dc->Q = Q; dc->kp = kp; dc->kq = kq; dc->ip = ip; dc->jq = jq;
    dc->myrank = myrank;
    dc->rrank = ((myrank / Q) + (P - ip)) % P;
    dc->crank = ((myrank % Q) + (Q - jq)) % Q;
    dc->nb_elem_r = owned_along(dc->rrank, kp, P, dc->lmt);
    dc->nb_elem_c = owned_along(dc->crank, kq, Q, dc->lnt);
    if (!dc->nb_elem_r || !dc->nb_elem_c) dc->nb_elem_r = dc->nb_elem_c = 0;
    dc->nb_local_tiles = (int64_t)dc->nb_elem_r * dc->nb_elem_c;
    dc->llm = dc->nb_elem_r * mb;
    dc->lln = dc->nb_elem_c * nb;
    if (lapack) {  // parsec_matrix_block_cyclic_lapack_init (:61-107): the caller says how large the local matrix is
        dc->llm = mloc;
        dc->lln = nloc;
    }
    dc->mat = nullptr;
    dc->dc_id = g_next_dc_id.fetch_add(1);
    dc->data_map.assign((size_t)dc->nb_local_tiles, nullptr);
    return dc;
}

extern "C" pb2_matrix_block_cyclic_t* pb2_matrix_block_cyclic_new(int mtype, int myrank, int mb, int nb, int lm,
                                                                  int ln, int i, int j, int m, int n, int P, int Q,
                                                                  int kp, int kq, int ip, int jq) {
    return make_descriptor(mtype, myrank, mb, nb, lm, ln, i, j, m, n, P, Q, kp, kq, ip, jq, 0, 0, 0);
}

extern "C" pb2_matrix_block_cyclic_t* pb2_matrix_block_cyclic_lapack_new(int mtype, int myrank, int mb, int nb, int lm,
                                                                         int ln, int i, int j, int m, int n, int P,
                                                                         int Q, int kp, int kq, int ip, int jq,
                                                                         int mloc, int nloc) {
    if (mloc <= 0 || nloc <= 0) return nullptr;
    return make_descriptor(mtype, myrank, mb, nb, lm, ln, i, j, m, n, P, Q, kp, kq, ip, jq, 1, mloc, nloc);
}

extern "C" void pb2_matrix_block_cyclic_free(pb2_matrix_block_cyclic_t* dc) {
    if (!dc) return;
    for (pb2_data_s* d : dc->data_map)
        if (d) pb2_data_destroy(d);
    delete dc;
}

extern "C" int pb2_matrix_set_mat(pb2_matrix_block_cyclic_t* dc, void* mat) {
    if (!dc) return PB2_ERR_BAD_PARAM;
    dc->mat = mat;
    return PB2_SUCCESS;
}
extern "C" int64_t pb2_matrix_nb_local_tiles(const pb2_matrix_block_cyclic_t* dc) { return dc ? dc->nb_local_tiles : -1; }
extern "C" int pb2_matrix_mt(const pb2_matrix_block_cyclic_t* dc) { return dc ? dc->mt : -1; }
extern "C" int pb2_matrix_nt(const pb2_matrix_block_cyclic_t* dc) { return dc ? dc->nt : -1; }

extern "C" int pb2_matrix_rank_of(const pb2_matrix_block_cyclic_t* dc, int m, int n) {
    if (!dc || m < 0 || n < 0 || m >= dc->mt || n >= dc->nt) return PB2_ERR_BAD_PARAM;
    m += dc->i / dc->mb;
    n += dc->j / dc->nb;
    // (m, n) lies in k-cyclic super-tile (m / kp, n / kq): twoDBC_kcyclic_rank_of (:532-566), which is twoDBC_rank_of
    // (:258-286) when kp = kq = 1
    return (((m / dc->kp) % dc->P + dc->ip) % dc->P) * dc->Q + ((n / dc->kq) % dc->Q + dc->jq) % dc->Q;
}

extern "C" int64_t pb2_matrix_position(const pb2_matrix_block_cyclic_t* dc, int m, int n) {
    if (!dc || m < 0 || n < 0 || m >= dc->mt || n >= dc->nt) return PB2_ERR_BAD_PARAM;
    m += dc->i / dc->mb;
    n += dc->j / dc->nb;
    // twoDBC_kcyclic_data_of (:652-664); twoDBC_coordinates_to_position (:333-349) when kp = kq = 1
    const int local_m = (m / (dc->kp * dc->P)) * dc->kp + (m % (dc->kp * dc->P)) % dc->kp;
    const int local_n = (n / (dc->kq * dc->Q)) * dc->kq + (n % (dc->kq * dc->Q)) % dc->kq;
    return (int64_t)dc->nb_elem_r * local_n + local_m;
}

// Element offset of tile (m, n) inside the local matrix, what data_of adds to dc->mat: tile storage keeps tile
// `position` at position * bsiz (:397-401); LAPACK storage addresses element (local_m * mb, local_n * nb) of the
// column-major local matrix with ld = llm (:402-407, :666-675).
extern "C" int64_t pb2_matrix_tile_offset(const pb2_matrix_block_cyclic_t* dc, int m, int n) {
    if (!dc || m < 0 || n < 0 || m >= dc->mt || n >= dc->nt) return PB2_ERR_BAD_PARAM;
    if (!dc->lapack) return pb2_matrix_position(dc, m, n) * dc->bsiz;
    m += dc->i / dc->mb;
    n += dc->j / dc->nb;
    const int64_t local_m = (m / (dc->kp * dc->P)) * dc->kp + (m % (dc->kp * dc->P)) % dc->kp;
    const int64_t local_n = (n / (dc->kq * dc->Q)) * dc->kq + (n % (dc->kq * dc->Q)) % dc->kq;
    return local_n * dc->nb * dc->llm + local_m * dc->mb;
}

extern "C" uint64_t pb2_matrix_data_key(const pb2_matrix_block_cyclic_t* dc, int m, int n) {
    m += dc->i / dc->mb;
    n += dc->j / dc->nb;
    return (uint64_t)n * (uint64_t)dc->lmt + (uint64_t)m;
}

extern "C" pb2_data_t* pb2_matrix_data_of(pb2_matrix_block_cyclic_t* dc, int m, int n) {
    if (!dc || pb2_matrix_rank_of(dc, m, n) != dc->myrank) return nullptr;
    int64_t pos = pb2_matrix_position(dc, m, n);
    std::lock_guard<std::mutex> g(dc->lock);
    if (!dc->data_map[(size_t)pos]) {
        char* ptr = dc->mat ? (char*)dc->mat + (size_t)pb2_matrix_tile_offset(dc, m, n) * dc->elem_size : nullptr;
        // key of the parsec_data_t: the data key, except that twoDBC_kcyclic_data_of builds it from the coordinates it
        // has just reduced modulo the super-tile period (:652-678); kept as the reference has it
        uint64_t key = pb2_matrix_data_key(dc, m, n);
        if (dc->kp > 1 || dc->kq > 1) {
            const int gm = (m + dc->i / dc->mb) % (dc->kp * dc->P), gn = (n + dc->j / dc->nb) % (dc->kq * dc->Q);
            key = (uint64_t)gn * (uint64_t)dc->lmt + (uint64_t)gm;
        }
        pb2_data_s* d = pb2_data_create(key, dc->dc_id, ptr, (uint32_t)((size_t)dc->bsiz * dc->elem_size));
        if (d && dc->lapack) {
            // a window of the local matrix: only the rows / columns the matrix really has (the last tile row / column of
            // an unpadded matrix is short; the reference's stage_custom functions copy mb x nb regardless and need N % NB == 0)
            const int gm = m + dc->i / dc->mb, gn = n + dc->j / dc->nb;
            const int rows = std::min(dc->mb, dc->lm - gm * dc->mb), cols = std::min(dc->nb, dc->ln - gn * dc->nb);
            pb2_data_set_host_layout(d, (uint32_t)(rows * dc->elem_size), (uint32_t)cols, (size_t)dc->llm * dc->elem_size,
                                     (size_t)dc->mb * dc->elem_size);
        }
        dc->data_map[(size_t)pos] = d;
    }
    return dc->data_map[(size_t)pos];
}

extern "C" int pb2_matrix_register_memory(pb2_matrix_block_cyclic_t* dc, pb2_device_t* dev) {
    if (!dc || !dev || !dc->mat) return PB2_ERR_BAD_PARAM;
    const size_t elems = dc->lapack ? (size_t)dc->llm * dc->lln : (size_t)dc->nb_local_tiles * dc->bsiz;
    return pb2_device_memory_register(dev, dc->mat, elems * dc->elem_size);
}
extern "C" int pb2_matrix_unregister_memory(pb2_matrix_block_cyclic_t* dc, pb2_device_t* dev) {
    if (!dc || !dev || !dc->mat) return PB2_ERR_BAD_PARAM;
    return pb2_device_memory_unregister(dev, dc->mat);
}

// accessors for the taskpool (same translation unit family; not part of the public header)
extern "C" int pb2_matrix_geometry(const pb2_matrix_block_cyclic_t* dc, int* out /*[18]*/) {
    if (!dc || !out) return PB2_ERR_BAD_PARAM;
    int v[18] = {dc->mtype, dc->elem_size, dc->mb, dc->nb, dc->m, dc->n, dc->mt, dc->nt, dc->P, dc->Q, dc->myrank, (int)dc->dc_id,
                 dc->kp, dc->kq, dc->ip, dc->jq, dc->i, dc->j};
    for (int i = 0; i < 18; ++i) out[i] = v[i];
    return PB2_SUCCESS;
}

// ------------------------------------------------------------------------------------ generator
// tests/dsl/dtd/dtd_test_simple_gemm.c:43-91.  One thread per tile column: jump ahead to the column start, then
// run the LCG down the column.  The value is formed in single precision with two roundings (no FMA), as the
// reference's `0.5f - ran * RndF_Mul` compiles on the host.
#define PB2_RND_A 6364136223846793005ULL
#define PB2_RND_C 1ULL

__device__ __forceinline__ unsigned long long lcg_jump(unsigned long long n, unsigned long long seed) {
    unsigned long long mul = PB2_RND_A, add = PB2_RND_C, x = seed;
    while (n) {
        if (n & 1ULL) x = mul * x + add;
        add *= (mul + 1ULL);
        mul *= mul;
        n >>= 1;
    }
    return x;
}

template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ double from_f32<double>(float v) { return (double)v; }
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// tiles[]: (m, n) tile coordinates of the batch; out: batch of tiles, tile t at out + t * mb * nb
template <typename T>
__global__ void lcg_fill_tiles(T* out, const int2* tiles, int mb, int nb, int M, unsigned seed) {
    const int tile = blockIdx.y;
    const int col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= nb) return;
    const int2 mn = tiles[tile];
    unsigned long long jump = (unsigned long long)mn.x + ((unsigned long long)mn.y + (unsigned long long)col) * (unsigned long long)M;
    unsigned long long x = lcg_jump(jump, seed);
    T* p = out + (size_t)tile * mb * nb + (size_t)col * mb;
    for (int row = 0; row < mb; ++row) {
        float scaled = __fmul_rn(__ull2float_rn(x), 5.4210108624275222e-20f);
        p[row] = from_f32<T>(__fsub_rn(0.5f, scaled));
        x = PB2_RND_A * x + PB2_RND_C;
    }
}

extern "C" int pb2_matrix_fill_lcg(pb2_matrix_block_cyclic_t* dc, unsigned seed, pb2_device_t* dev) {
    if (!dc || !dc->mat || !dev) return PB2_ERR_BAD_PARAM;
    if (dc->lapack) return PB2_ERR_NOT_SUPPORTED;  // the generator writes packed tiles (the reference's does too)
    if (cudaSetDevice(pb2_device_cuda_index(dev)) != cudaSuccess) return PB2_ERR_DEVICE;
    // local tiles in storage order
    std::vector<int2> coords((size_t)dc->nb_local_tiles);
    for (int n = 0; n < dc->nt; ++n)
        for (int m = 0; m < dc->mt; ++m)
            if (pb2_matrix_rank_of(dc, m, n) == dc->myrank) coords[(size_t)pb2_matrix_position(dc, m, n)] = make_int2(m, n);
    const size_t tile_bytes = (size_t)dc->bsiz * dc->elem_size;
    // at most 256 MiB of staging and 65535 tiles (gridDim.y) per launch
    const int batch = (int)std::max<size_t>(
        1, std::min<size_t>({(size_t)dc->nb_local_tiles, (size_t)(256u << 20) / tile_bytes, (size_t)65535}));
    char* dbuf = nullptr;
    int2* dcoords = nullptr;
    if (cudaMalloc((void**)&dbuf, tile_bytes * batch) != cudaSuccess) return PB2_ERR_OUT_OF_RESOURCE;
    if (cudaMalloc((void**)&dcoords, sizeof(int2) * coords.size()) != cudaSuccess) {
        cudaFree(dbuf);
        return PB2_ERR_OUT_OF_RESOURCE;
    }
    cudaMemcpy(dcoords, coords.data(), sizeof(int2) * coords.size(), cudaMemcpyHostToDevice);
    int rc = PB2_SUCCESS;
    // the column walk is strided by mb elements across threads: fine for a one-off initialiser
    for (int64_t t0 = 0; t0 < dc->nb_local_tiles && rc == PB2_SUCCESS; t0 += batch) {
        int nt = (int)std::min<int64_t>(batch, dc->nb_local_tiles - t0);
        dim3 grid((dc->nb + 127) / 128, nt);
        if (dc->mtype == PB2_MATRIX_DOUBLE)
            lcg_fill_tiles<double><<<grid, 128>>>((double*)dbuf, dcoords + t0, dc->mb, dc->nb, dc->m, seed);
        else if (dc->mtype == PB2_MATRIX_FLOAT)
            lcg_fill_tiles<float><<<grid, 128>>>((float*)dbuf, dcoords + t0, dc->mb, dc->nb, dc->m, seed);
        else
            lcg_fill_tiles<__nv_bfloat16><<<grid, 128>>>((__nv_bfloat16*)dbuf, dcoords + t0, dc->mb, dc->nb, dc->m, seed);
        if (cudaGetLastError() != cudaSuccess ||
            cudaMemcpy((char*)dc->mat + (size_t)t0 * tile_bytes, dbuf, tile_bytes * nt, cudaMemcpyDeviceToHost) != cudaSuccess)
            rc = PB2_ERR_DEVICE;
    }
    cudaFree(dbuf);
    cudaFree(dcoords);
    return rc;
}

// ------------------------------------------------------------------------------------ generator, no allocation
// pb2_matrix_fill_lcg above stages tiles through a cudaMalloc'd buffer.  That is a problem for two kinds of callers
// of this library:
//   * anything that runs after pb2_device_create: the device module has by then taken 95 % of the free HBM for its
//     tile heap, so a further cudaMalloc of a staging buffer can fail with out-of-memory on a perfectly healthy run;
//   * anything that runs while the pipeline is busy: cudaMalloc/cudaFree synchronise the whole device, which stalls
//     every stream the manager is driving.
// The variant below therefore owns its staging space statically.  The array lives in every GPU context that loads
// this library until the process exits, so it is kept modest: 16 MiB, i.e. tiles of at most 16 MiB (NB = 1024 fp64 is
// 8 MiB; NB = 2048 fp64 does not fit and is refused).
#define PB2_FILL_SCRATCH_BYTES (16u << 20)
static __device__ __align__(256) unsigned char pb2_fill_scratch[PB2_FILL_SCRATCH_BYTES];

// In host code `pb2_fill_scratch` is only a handle for the symbol, not a device address; the address has to be asked
// for once per device, with that device current.
static void* g_fill_scratch_addr[64] = {nullptr};

static int fill_scratch_address(int cuda_index, void** out) {
    if (cuda_index < 0 || cuda_index >= 64) return PB2_ERR_BAD_PARAM;
    if (!g_fill_scratch_addr[cuda_index]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, pb2_fill_scratch);
        if (e != cudaSuccess || !p) {
            fprintf(stderr, "[parsec_b200] cudaGetSymbolAddress(pb2_fill_scratch) failed on device %d: %s\n", cuda_index,
                    cudaGetErrorString(e));
            return PB2_ERR_DEVICE;
        }
        g_fill_scratch_addr[cuda_index] = p;
    }
    *out = g_fill_scratch_addr[cuda_index];
    return PB2_SUCCESS;
}

// Same arithmetic as lcg_fill_tiles; the tile coordinates are derived from the storage position instead of being read
// from a coordinate table (which would be one more buffer to allocate).
// pb2_matrix_position inverted: a rank stores its tiles column by column, nb_elem_r per local column; with k-cyclic
// super-tiles local row l is global row (l / kp) * kp * P + rrank * kp + l % kp (two_dim_rectangle_cyclic.c:652-664
// read backwards), and the same for columns with kq, Q, crank.  kp = kq = 1 gives l * P + rrank.
template <typename T>
__global__ void lcg_fill_tiles_by_position(T* out, long long first_pos, int nb_elem_r, int P, int Q, int kp, int kq,
                                           int rrank, int crank, int mb, int nb, int M, unsigned seed) {
    const int col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= nb) return;
    const long long pos = first_pos + blockIdx.y;
    const int lrow = (int)(pos % nb_elem_r), lcol = (int)(pos / nb_elem_r);
    const int m = (lrow / kp) * kp * P + rrank * kp + lrow % kp;
    const int n = (lcol / kq) * kq * Q + crank * kq + lcol % kq;
    unsigned long long jump = (unsigned long long)m + ((unsigned long long)n + (unsigned long long)col) * (unsigned long long)M;
    unsigned long long x = lcg_jump(jump, seed);
    T* p = out + (size_t)blockIdx.y * mb * nb + (size_t)col * mb;
    for (int row = 0; row < mb; ++row) {
        float scaled = __fmul_rn(__ull2float_rn(x), 5.4210108624275222e-20f);
        p[row] = from_f32<T>(__fsub_rn(0.5f, scaled));
        x = PB2_RND_A * x + PB2_RND_C;
    }
}

// Fills every local tile of `dc` like pb2_matrix_fill_lcg, without allocating anything (no cudaMalloc / cudaFree, no
// host heap, no library call that allocates): tiles are generated into the static scratch and copied to dc->mat.
//   * The first call on a device looks up the scratch address (cudaGetSymbolAddress); make that first call before
//     starting any stream capture.
//   * One tile must fit in the 16 MiB scratch; larger tiles return PB2_ERR_VALUE_OUT_OF_BOUNDS with a message.
//   * Every call on a device reuses the same scratch and finds the previous call's bytes in it.  Nothing here needs
//     zero-initialised scratch: each launch overwrites exactly the bytes that are then copied out.  Two calls that
//     overlap on the same GPU (different host threads) would race on it -- serialise them.
extern "C" int pb2_matrix_fill_lcg_noalloc(pb2_matrix_block_cyclic_t* dc, unsigned seed, pb2_device_t* dev) {
    if (!dc || !dc->mat || !dev) return PB2_ERR_BAD_PARAM;
    // sub-matrix views (an origin, or fewer tiles than the parent matrix stores) and LAPACK storage: pb2_matrix_fill_lcg
    if (dc->i != 0 || dc->j != 0 || dc->mt != dc->lmt || dc->nt != dc->lnt || dc->lapack) return PB2_ERR_NOT_SUPPORTED;
    const size_t tile_bytes = (size_t)dc->bsiz * dc->elem_size;
    if (tile_bytes > PB2_FILL_SCRATCH_BYTES) {
        fprintf(stderr,
                "[parsec_b200] pb2_matrix_fill_lcg_noalloc: a %d x %d tile is %zu bytes, the static scratch holds %u; "
                "use pb2_matrix_fill_lcg for tiles this large\n",
                dc->mb, dc->nb, tile_bytes, PB2_FILL_SCRATCH_BYTES);
        return PB2_ERR_VALUE_OUT_OF_BOUNDS;
    }
    const int cuda_index = pb2_device_cuda_index(dev);
    if (cudaSetDevice(cuda_index) != cudaSuccess) return PB2_ERR_DEVICE;
    void* scratch = nullptr;
    int rc = fill_scratch_address(cuda_index, &scratch);
    if (rc != PB2_SUCCESS) return rc;
    const long long batch = std::min<long long>((long long)(PB2_FILL_SCRATCH_BYTES / tile_bytes), 65535);  // gridDim.y
    for (long long t0 = 0; t0 < dc->nb_local_tiles; t0 += batch) {
        const int nt = (int)std::min<long long>(batch, dc->nb_local_tiles - t0);
        dim3 grid((dc->nb + 127) / 128, nt);
        if (dc->mtype == PB2_MATRIX_DOUBLE)
            lcg_fill_tiles_by_position<double><<<grid, 128>>>((double*)scratch, t0, dc->nb_elem_r, dc->P, dc->Q, dc->kp, dc->kq,
                                                              dc->rrank, dc->crank, dc->mb, dc->nb, dc->m, seed);
        else if (dc->mtype == PB2_MATRIX_FLOAT)
            lcg_fill_tiles_by_position<float><<<grid, 128>>>((float*)scratch, t0, dc->nb_elem_r, dc->P, dc->Q, dc->kp, dc->kq,
                                                             dc->rrank, dc->crank, dc->mb, dc->nb, dc->m, seed);
        else
            lcg_fill_tiles_by_position<__nv_bfloat16><<<grid, 128>>>((__nv_bfloat16*)scratch, t0, dc->nb_elem_r, dc->P,
                                                                     dc->Q, dc->kp, dc->kq, dc->rrank, dc->crank, dc->mb, dc->nb, dc->m,
                                                                     seed);
        if (cudaGetLastError() != cudaSuccess) return PB2_ERR_DEVICE;
        if (cudaMemcpy((char*)dc->mat + (size_t)t0 * tile_bytes, scratch, tile_bytes * nt, cudaMemcpyDeviceToHost) !=
            cudaSuccess)
            return PB2_ERR_DEVICE;
    }
    return PB2_SUCCESS;
}
