// dtrsm_sm100.cu -- parsec_b200_dtrsm: the triangular solve of a tile against a triangular tile, for the TRSM bodies of
// a tiled Cholesky / LU (what DPLASMA hands this runtime as cublasDtrsm_v2 calls).
//
// B200-first shape of the algorithm: almost all of the work is GEMM, so almost all of it goes to the DMMA tile kernel.
//   1. the 128 x 128 diagonal blocks of the triangular tile are inverted (one CTA each, a few microseconds, done
//      concurrently for all of them);
//   2. block column (row) j of the solution is ONE GEMM with the inverted block, and the blocks not yet solved are
//      updated with ONE GEMM of depth 128 -- parsec_b200_dgemm_sm100 with the transposes the call asks for.
// So a solve against an NB = 1024 tile is 8 inversions in one launch and 16 GEMM launches; substitution, the part
// that does not parallelise, is confined to the 128 x 128 blocks.  The scratch (inverted blocks + one block column)
// is the caller's: a CUDA body takes it from parsec_gpu_pop_workspace (device_gpu.c:168-246, here
// pb2_device_pop_workspace), parsec_b200_dtrsm_workspace says how much.
#include <cuda_runtime.h>
#include <stdint.h>

#include "parsec_b200.h"
#include "pb2_internal.h"

namespace pb2 {

constexpr int TB_NB = 128;                      // block size of the solve
constexpr int TB_LD = TB_NB + 1;                // padded shared-memory row
constexpr int TRTRI_THREADS = 256;
constexpr size_t TRTRI_SMEM = ((size_t)TB_NB * TB_LD + 8 * 16 * 17) * sizeof(double);

// The lower triangular 128 x 128 matrix in S (row-major, row length TB_LD, zeros above the diagonal) is replaced by its
// inverse.  All TRTRI_THREADS threads of the CTA; X0 is scratch for the 16 x 16 stage.  In the doubling steps a thread
// owns four neighbouring outputs of a row, so the four sums run as independent chains off one load of the left factor;
// the lanes of a warp take different ROWS: their loads of the left factor are TB_LD apart (no bank conflict) and their
// loads of the right factor are one address (broadcast).
__device__ __forceinline__ void invert_lower_in_smem(double* S, double* X0, int tid) {
    if (tid < TB_NB) {  // the eight 16 x 16 diagonal sub-blocks: sub-block b, column c
        const int b = tid >> 4, c = tid & 15;
        const double* L = S + (b * 16) * TB_LD + b * 16;
        double x[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] = 0.0;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            if (i == c) x[i] = 1.0 / L[i * TB_LD + i];
            if (i > c) {
                double acc = 0.0;
#pragma unroll
                for (int k = 0; k < 16; ++k)
                    if (k < i) acc = fma(L[i * TB_LD + k], x[k], acc);  // x[k] = 0 for k < c
                x[i] = -acc / L[i * TB_LD + i];
            }
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) X0[(b * 16 + i) * 17 + c] = x[i];
    }
    __syncthreads();
    for (int e = tid; e < 8 * 256; e += TRTRI_THREADS) {
        const int b = e >> 8, i = (e >> 4) & 15, c = e & 15;
        S[(b * 16 + i) * TB_LD + b * 16 + c] = X0[(b * 16 + i) * 17 + c];
    }
    __syncthreads();
    for (int sz = 16; sz < TB_NB; sz *= 2) {  // pairs of inverted sz x sz blocks -> inverted 2sz x 2sz blocks
        const int quads = (TB_NB / (2 * sz)) * sz * (sz / 4);
        for (int o = tid; o < quads; o += TRTRI_THREADS) {  // tmp = C A^-1, parked in the mirror block
            const int i = o % sz, j = ((o / sz) % (sz / 4)) * 4, base = (o / (sz * (sz / 4))) * 2 * sz;
            const double* Crow = S + (base + sz + i) * TB_LD + base;
            const double* Ainv = S + base * TB_LD + base;
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll 4
            for (int k = j; k < sz; ++k) {  // A^-1 is lower triangular: its entries above the diagonal are stored zeros
                const double c = Crow[k];
                const double* ar = Ainv + k * TB_LD + j;
                a0 = fma(c, ar[0], a0);
                a1 = fma(c, ar[1], a1);
                a2 = fma(c, ar[2], a2);
                a3 = fma(c, ar[3], a3);
            }
            double* out = S + (base + i) * TB_LD + base + sz + j;
            out[0] = a0; out[1] = a1; out[2] = a2; out[3] = a3;
        }
        __syncthreads();
        for (int o = tid; o < quads; o += TRTRI_THREADS) {  // C = -B^-1 tmp
            const int i = o % sz, j = ((o / sz) % (sz / 4)) * 4, base = (o / (sz * (sz / 4))) * 2 * sz;
            const double* Brow = S + (base + sz + i) * TB_LD + base + sz;
            const double* tmp = S + base * TB_LD + base + sz + j;
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll 4
            for (int k = 0; k <= i; ++k) {
                const double bv = Brow[k];
                const double* tr = tmp + k * TB_LD;
                a0 = fma(bv, tr[0], a0);
                a1 = fma(bv, tr[1], a1);
                a2 = fma(bv, tr[2], a2);
                a3 = fma(bv, tr[3], a3);
            }
            double* out = S + (base + sz + i) * TB_LD + base + j;
            out[0] = -a0; out[1] = -a1; out[2] = -a2; out[3] = -a3;
        }
        __syncthreads();
        for (int o = tid; o < quads; o += TRTRI_THREADS) {
            const int i = o % sz, j = ((o / sz) % (sz / 4)) * 4, base = (o / (sz * (sz / 4))) * 2 * sz;
            double* out = S + (base + i) * TB_LD + base + sz + j;
            out[0] = out[1] = out[2] = out[3] = 0.0;
        }
        __syncthreads();
    }
}

// Inverse of each diagonal block of the triangular matrix T (n x n, `lower` says which triangle is stored), written as
// full 128 x 128 blocks (ld 128) with zeros in the other triangle.  One CTA per block, all in shared memory:
//   1. the eight 16 x 16 diagonal sub-blocks by substitution (one thread per column: 16 short dependent steps);
//   2. three doubling steps 16 -> 32 -> 64 -> 128 with  inv [A 0; C B] = [A^-1 0; -B^-1 C A^-1  B^-1]:  two small
//      matrix products per step by all the threads, the product C A^-1 parked in the (empty) mirror block above the
//      diagonal.  No step depends on more than 16 serial operations, against 128 for plain substitution.
// An upper triangle is inverted as the transpose of a lower one.
__global__ void __launch_bounds__(TRTRI_THREADS) trtri_diag_blocks(int n, int lower, int unit, const double* __restrict__ T,
                                                                   int ldt, double* __restrict__ Tinv) {
    extern __shared__ double sm[];
    double* S = sm;                          // S[i * TB_LD + j]: the block, made lower triangular
    double* X0 = sm + TB_NB * TB_LD;         // X0[(b * 16 + i) * 17 + c]: inverses of the 16 x 16 sub-blocks
    const int blk = blockIdx.x, j0 = blk * TB_NB;
    const int nb = n - j0 < TB_NB ? n - j0 : TB_NB;
    const int tid = threadIdx.x;
    const double* t = T + (size_t)j0 + (size_t)j0 * ldt;
    for (int e = tid; e < TB_NB * TB_NB; e += TRTRI_THREADS) {
        const int r = e % TB_NB, c = e / TB_NB;  // element (r, c) of the stored block, column-major: coalesced reads
        double v = 0.0;
        if (r < nb && c < nb) {
            if (r == c) v = unit ? 1.0 : t[(size_t)r + (size_t)c * ldt];
            else if (lower ? r > c : r < c) v = t[(size_t)r + (size_t)c * ldt];
        } else if (r == c) {
            v = 1.0;                             // identity padding of a ragged last block
        }
        if (lower) S[r * TB_LD + c] = v; else S[c * TB_LD + r] = v;
    }
    __syncthreads();
    invert_lower_in_smem(S, X0, tid);
    double* out = Tinv + (size_t)blk * TB_NB * TB_NB;
    for (int e = tid; e < TB_NB * TB_NB; e += TRTRI_THREADS) {
        const int r = e % TB_NB, c = e / TB_NB;
        out[e] = lower ? S[r * TB_LD + c] : S[c * TB_LD + r];
    }
}

__global__ void scale_matrix(int m, int n, double alpha, double* __restrict__ B, int ldb) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
    if (i < m && j < n) {
        double* p = B + (size_t)i + (size_t)j * ldb;
        *p = alpha == 0.0 ? 0.0 : alpha * *p;
    }
}

static inline int flag(char c, char a, char b) {  // 1 for a, 0 for b, -1 otherwise (case-insensitive)
    if (c == a || c == a + 32) return 1;
    if (c == b || c == b + 32) return 0;
    return -1;
}

}  // namespace pb2

extern "C" size_t parsec_b200_dtrsm_workspace(char side, int m, int n) {
    using namespace pb2;
    const int right = flag(side, 'R', 'L');
    if (right < 0 || m < 0 || n < 0) return 0;
    const int tri = right ? n : m;
    const size_t nblk = (size_t)(tri + TB_NB - 1) / TB_NB;
    return (nblk * TB_NB * TB_NB + (size_t)((m + 1) & ~1) * (size_t)n) * sizeof(double) + 512;
}

namespace pb2 {
static int ensure_trtri_attr() {
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 64 && !attr_set[dev]) {
        if (cudaFuncSetAttribute(trtri_diag_blocks, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TRTRI_SMEM) != cudaSuccess)
            return PB2_ERR_CUDA;
        attr_set[dev] = true;
    }
    return PB2_SUCCESS;
}

// The solve proper, given the inverted diagonal blocks.  X is built in W (same shape as B, ld ldw) so that no step
// reads and writes the same memory, and copied over B at the end: per block column (row) j
//     W_j   =  B_j op(Tinv_jj)            (left:  op(Tinv_jj) B_j)
//     B_rest -= W_j U(j, rest)             (left:  U(rest, j) W_j),     U = op(T)
// two GEMM launches per block, nothing else in the chain.
static int solve_with_inverses(int right, int lower, int tr, int m, int n, double alpha, const double* T, int ldt,
                               const double* Tinv, double* B, int ldb, double* W, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (alpha != 1.0) scale_matrix<<<dim3((m + 127) / 128, n), 128, 0, st>>>(m, n, alpha, B, ldb);
    const int tri = right ? n : m;
    const int nblk = (tri + TB_NB - 1) / TB_NB;
    // U = op(T) is upper triangular when T is upper and not transposed, or lower and transposed
    const bool u_upper = (lower == 0) != (tr == 1);
    const char tc = tr ? 'T' : 'N';
    const bool forward = right ? u_upper : !u_upper;
    const int ldw = (m + 1) & ~1;
    for (int s = 0; s < nblk; ++s) {
        const int j = forward ? s : nblk - 1 - s;
        const int j0 = j * TB_NB, nbj = tri - j0 < TB_NB ? tri - j0 : TB_NB;
        const int r0 = forward ? j0 + nbj : 0;  // the blocks not solved yet: [r0, r0 + nrest)
        const int nrest = forward ? tri - r0 : j0;
        const double* inv = Tinv + (size_t)j * TB_NB * TB_NB;
        int rc;
        if (right) {
            double* Wj = W + (size_t)j0 * ldw;
            rc = parsec_b200_dgemm_sm100('N', tc, m, nbj, nbj, 1.0, B + (size_t)j0 * ldb, ldb, inv, TB_NB, 0.0, Wj, ldw, stream);
            if (rc != PB2_SUCCESS) return rc;
            if (nrest > 0) {  // U(j, rest) = T(j, rest) or T(rest, j)^T
                const double* Tp = tr ? T + (size_t)r0 + (size_t)j0 * ldt : T + (size_t)j0 + (size_t)r0 * ldt;
                rc = parsec_b200_dgemm_sm100('N', tc, m, nrest, nbj, -1.0, Wj, ldw, Tp, ldt, 1.0, B + (size_t)r0 * ldb, ldb, stream);
                if (rc != PB2_SUCCESS) return rc;
            }
        } else {
            double* Wj = W + j0;
            rc = parsec_b200_dgemm_sm100(tc, 'N', nbj, n, nbj, 1.0, inv, TB_NB, B + j0, ldb, 0.0, Wj, ldw, stream);
            if (rc != PB2_SUCCESS) return rc;
            if (nrest > 0) {  // U(rest, j) = T(rest, j) or T(j, rest)^T
                const double* Tp = tr ? T + (size_t)j0 + (size_t)r0 * ldt : T + (size_t)r0 + (size_t)j0 * ldt;
                rc = parsec_b200_dgemm_sm100(tc, 'N', nrest, n, nbj, -1.0, Tp, ldt, Wj, ldw, 1.0, B + r0, ldb, stream);
                if (rc != PB2_SUCCESS) return rc;
            }
        }
    }
    if (cudaMemcpy2DAsync(B, (size_t)ldb * 8, W, (size_t)ldw * 8, (size_t)m * 8, n, cudaMemcpyDeviceToDevice, st) != cudaSuccess)
        return PB2_ERR_CUDA;
    return cudaGetLastError() == cudaSuccess ? PB2_SUCCESS : PB2_ERR_CUDA;
}

static int parse(char side, char uplo, char trans, char diag, int* right, int* lower, int* tr, int* unit) {
    *right = flag(side, 'R', 'L');
    *lower = flag(uplo, 'L', 'U');
    *unit = flag(diag, 'U', 'N');
    *tr = (trans == 'N' || trans == 'n') ? 0 : (trans == 'T' || trans == 't' || trans == 'C' || trans == 'c') ? 1 : -1;
    return (*right < 0 || *lower < 0 || *unit < 0 || *tr < 0) ? PB2_ERR_BAD_PARAM : PB2_SUCCESS;
}
static inline double* align256(void* p) { return reinterpret_cast<double*>(((uintptr_t)p + 255) & ~(uintptr_t)255); }
}  // namespace pb2

// Inverses of the 128 x 128 diagonal blocks of the triangular n x n matrix T, into Tinv: ceil(n / 128) blocks of
// 128 x 128 doubles.  On its own so that a DAG can invert the diagonal tile of a panel ONCE and hand the result to
// every solve of the panel (parsec_b200_dtrsm_inv).
extern "C" int parsec_b200_dtrtri_diag_sm100(char uplo, char diag, int n, const double* T, int ldt, double* Tinv, void* stream) {
    using namespace pb2;
    const int lower = flag(uplo, 'L', 'U'), unit = flag(diag, 'U', 'N');
    if (lower < 0 || unit < 0 || n < 0 || ldt < (n > 1 ? n : 1) || (n && (!T || !Tinv))) return PB2_ERR_BAD_PARAM;
    if (n == 0) return PB2_SUCCESS;
    int rc = ensure_trtri_attr();
    if (rc != PB2_SUCCESS) return rc;
    trtri_diag_blocks<<<(n + TB_NB - 1) / TB_NB, TRTRI_THREADS, TRTRI_SMEM, (cudaStream_t)stream>>>(n, lower, unit, T, ldt, Tinv);
    return cudaGetLastError() == cudaSuccess ? PB2_SUCCESS : PB2_ERR_CUDA;
}

extern "C" int parsec_b200_dtrsm_inv_sm100(char side, char uplo, char trans, int m, int n, double alpha, const double* T,
                                            int ldt, const double* Tinv, double* B, int ldb, void* workspace,
                                            size_t workspace_bytes, void* stream) {
    using namespace pb2;
    int right, lower, tr, unit;
    if (parse(side, uplo, trans, 'N', &right, &lower, &tr, &unit) != PB2_SUCCESS || m < 0 || n < 0) return PB2_ERR_BAD_PARAM;
    const int tri = right ? n : m;
    if (ldt < (tri > 1 ? tri : 1) || ldb < (m > 1 ? m : 1)) return PB2_ERR_BAD_PARAM;
    if (m == 0 || n == 0) return PB2_SUCCESS;
    if (!workspace || !Tinv || workspace_bytes < (size_t)((m + 1) & ~1) * (size_t)n * sizeof(double) + 256) return PB2_ERR_BAD_PARAM;
    return solve_with_inverses(right, lower, tr, m, n, alpha, T, ldt, Tinv, B, ldb, align256(workspace), stream);
}

extern "C" int parsec_b200_dtrsm_sm100(char side, char uplo, char trans, char diag, int m, int n, double alpha,
                                        const double* T, int ldt, double* B, int ldb, void* workspace,
                                        size_t workspace_bytes, void* stream) {
    using namespace pb2;
    int right, lower, tr, unit;
    if (parse(side, uplo, trans, diag, &right, &lower, &tr, &unit) != PB2_SUCCESS || m < 0 || n < 0) return PB2_ERR_BAD_PARAM;
    const int tri = right ? n : m;
    if (ldt < (tri > 1 ? tri : 1) || ldb < (m > 1 ? m : 1)) return PB2_ERR_BAD_PARAM;
    if (m == 0 || n == 0) return PB2_SUCCESS;
    if (!workspace || workspace_bytes < parsec_b200_dtrsm_workspace(side, m, n)) return PB2_ERR_BAD_PARAM;
    const int nblk = (tri + TB_NB - 1) / TB_NB;
    double* Tinv = align256(workspace);
    double* W = Tinv + (size_t)nblk * TB_NB * TB_NB;
    int rc = parsec_b200_dtrtri_diag_sm100(uplo, diag, tri, T, ldt, Tinv, stream);
    if (rc != PB2_SUCCESS) return rc;
    return solve_with_inverses(right, lower, tr, m, n, alpha, T, ldt, Tinv, B, ldb, W, stream);
}

extern "C" int parsec_b200_dtrsm(char side, char uplo, char trans, char diag, int m, int n, double alpha, const double* T,
                                  int ldt, double* B, int ldb, void* workspace, size_t workspace_bytes, void* stream) {
    return parsec_b200_dtrsm_sm100(side, uplo, trans, diag, m, n, alpha, T, ldt, B, ldb, workspace, workspace_bytes, stream);
}

// ---------------------------------------------------------------------------------------------------------------
// parsec_b200_dpotrf: Cholesky factorisation of one tile on the GPU, for the POTRF body of the same DAG (what DPLASMA
// hands the runtime as a cusolverDnDpotrf call).  Right-looking over 128-column blocks, and again almost all GEMM:
//   potf2_trtri   the 128 x 128 diagonal block factored AND its factor inverted, one CTA, in shared memory
//   panel         L21 = A21 L11^-T = A21 * inv(L11)^T : parsec_b200_dgemm with the inverse just computed
//   trailing      A22 -= L21 L21^T                     : parsec_b200_dsyrk
namespace pb2 {

// Factor the nb x nb diagonal block (lower: A = L L^T; upper: A = U^T U, handled as the transpose) in place, and write
// inv(L) (resp. inv(U)) as a full 128 x 128 block to `inv`.  Thread i owns row i: column j of L is a dot product of
// length j per row, then one square root, then a scale.  *info = first non-positive pivot (1-based, offset by j0).
__global__ void __launch_bounds__(TRTRI_THREADS) potf2_trtri(int nb, int lower, double* __restrict__ A, int lda,
                                                             double* __restrict__ inv, int j0, int* __restrict__ info) {
    extern __shared__ double sm[];
    double* S = sm;
    double* X0 = sm + TB_NB * TB_LD;
    __shared__ int bad;
    const int tid = threadIdx.x;
    if (tid == 0) bad = 0;
    for (int e = tid; e < TB_NB * TB_NB; e += TRTRI_THREADS) {
        const int r = e % TB_NB, c = e / TB_NB;
        double v = (r == c) ? 1.0 : 0.0;  // identity padding of a ragged block
        if (r < nb && c < nb) v = (lower ? r >= c : r <= c) ? A[(size_t)r + (size_t)c * lda] : 0.0;
        if (lower) S[r * TB_LD + c] = v; else S[c * TB_LD + r] = v;
    }
    __syncthreads();
    // Right-looking, register blocked: the 128 x 128 block is a 16 x 16 grid of 8 x 8 sub-blocks, thread (ti, tj) keeps
    // sub-block (ti, tj) in registers for the whole factorisation.  Step j: the owner of the diagonal sub-block takes the
    // square root, the owners of block column j/8 scale their piece of column j and publish it (shared memory, two
    // buffers in turn), and everybody applies the rank-1 update to its own registers: two barriers and 64 independent
    // multiply-adds per step, nothing read-modify-written in shared memory.
    {
        const int ti = tid & 15, tj = tid >> 4;
        double* colj = X0;            // 2 x 128 doubles
        double* dsh = X0 + 2 * TB_NB;  // the pivot of the step
        double a[8][8];
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < 8; ++c) a[r][c] = S[(8 * ti + r) * TB_LD + 8 * tj + c];
        for (int jb = 0; jb < 16; ++jb) {
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) {
                const int j = 8 * jb + jj;
                double* col = colj + (j & 1) * TB_NB;
                if (ti == jb && tj == jb) {
                    double d = a[jj][jj];
                    if (!(d > 0.0)) {
                        if (!bad && j < nb) bad = j + 1;
                        d = 1.0;  // keep going with something finite; the caller reads info
                    }
                    d = sqrt(d);
                    a[jj][jj] = d;
                    *dsh = d;
                }
                __syncthreads();
                if (tj == jb && ti >= jb) {
                    const double d = *dsh;
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        if (ti > jb || r > jj) {
                            a[r][jj] /= d;
                            col[16 * r + ti] = a[r][jj];   /* row 8 ti + r, stored r-major: lanes differ in ti */
                        }
                    }
                }
                __syncthreads();
                if (ti >= tj && tj >= jb) {
                    double li[8], lc[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r) li[r] = col[16 * r + ti];
#pragma unroll
                    for (int c = 0; c < 8; ++c) lc[c] = col[16 * c + tj];
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        if (tj > jb || c > jj) {   // columns to the right of j only
#pragma unroll
                            for (int r = 0; r < 8; ++r)
                                if (ti > tj || r >= c) a[r][c] = fma(-li[r], lc[c], a[r][c]);   // lower triangle only
                        }
                    }
                }
            }
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < 8; ++c)
                if (ti > tj || (ti == tj && r >= c)) S[(8 * ti + r) * TB_LD + 8 * tj + c] = a[r][c];
        __syncthreads();
    }
    if (tid == 0 && bad && info && *info == 0) *info = j0 + bad;
    for (int e = tid; e < TB_NB * TB_NB; e += TRTRI_THREADS) {  // the factor goes back, the other triangle is left alone
        const int r = e % TB_NB, c = e / TB_NB;
        if (r < nb && c < nb && (lower ? r >= c : r <= c)) A[(size_t)r + (size_t)c * lda] = lower ? S[r * TB_LD + c] : S[c * TB_LD + r];
    }
    __syncthreads();
    invert_lower_in_smem(S, X0, tid);  // the inverse of the factor
    for (int e = tid; e < TB_NB * TB_NB; e += TRTRI_THREADS) {
        const int r = e % TB_NB, c = e / TB_NB;
        inv[e] = lower ? S[r * TB_LD + c] : S[c * TB_LD + r];
    }
}

}  // namespace pb2

extern "C" size_t parsec_b200_dpotrf_workspace(int n) {
    if (n < 0) return 0;
    return ((size_t)pb2::TB_NB * pb2::TB_NB + (size_t)((n + 1) & ~1) * pb2::TB_NB) * sizeof(double) + 512;
}

// A = L L^T ('L') or U^T U ('U') in place; the other triangle is not touched.  *dev_info (device memory, may be NULL;
// set it to 0 before the call) receives the index of the first non-positive pivot, as cusolverDnDpotrf's devInfo does.
extern "C" int parsec_b200_dpotrf_sm100(char uplo, int n, double* A, int lda, void* workspace, size_t workspace_bytes,
                                         int* dev_info, void* stream) {
    using namespace pb2;
    const int lower = flag(uplo, 'L', 'U');
    if (lower < 0 || n < 0 || lda < (n > 1 ? n : 1)) return PB2_ERR_BAD_PARAM;
    if (n == 0) return PB2_SUCCESS;
    if (!A || !workspace || workspace_bytes < parsec_b200_dpotrf_workspace(n)) return PB2_ERR_BAD_PARAM;
    cudaStream_t st = (cudaStream_t)stream;
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 64 && !attr_set[dev]) {
        if (cudaFuncSetAttribute(potf2_trtri, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TRTRI_SMEM) != cudaSuccess)
            return PB2_ERR_CUDA;
        attr_set[dev] = true;
    }
    double* inv = align256(workspace);
    double* W = inv + (size_t)TB_NB * TB_NB;
    const int ldw = (n + 1) & ~1;
    for (int j0 = 0; j0 < n; j0 += TB_NB) {
        const int nbj = n - j0 < TB_NB ? n - j0 : TB_NB;
        const int r0 = j0 + nbj, nrest = n - r0;
        double* Ajj = A + (size_t)j0 + (size_t)j0 * lda;
        potf2_trtri<<<1, TRTRI_THREADS, TRTRI_SMEM, st>>>(nbj, lower, Ajj, lda, inv, j0, dev_info);
        if (cudaGetLastError() != cudaSuccess) return PB2_ERR_CUDA;
        if (nrest == 0) break;
        int rc;
        if (lower) {  // L21 = A21 inv(L11)^T, then A22 -= L21 L21^T
            double* A21 = A + (size_t)r0 + (size_t)j0 * lda;
            rc = parsec_b200_dgemm_sm100('N', 'T', nrest, nbj, nbj, 1.0, A21, lda, inv, TB_NB, 0.0, W, ldw, stream);
            if (rc != PB2_SUCCESS) return rc;
            if (cudaMemcpy2DAsync(A21, (size_t)lda * 8, W, (size_t)ldw * 8, (size_t)nrest * 8, nbj, cudaMemcpyDeviceToDevice, st) != cudaSuccess)
                return PB2_ERR_CUDA;
            rc = parsec_b200_dsyrk_sm100('L', 'N', nrest, nbj, -1.0, W, ldw, 1.0, A + (size_t)r0 + (size_t)r0 * lda, lda, stream);
        } else {      // U12 = inv(U11)^T A12, then A22 -= U12^T U12
            double* A12 = A + (size_t)j0 + (size_t)r0 * lda;
            rc = parsec_b200_dgemm_sm100('T', 'N', nbj, nrest, nbj, 1.0, inv, TB_NB, A12, lda, 0.0, W, TB_NB, stream);
            if (rc != PB2_SUCCESS) return rc;
            if (cudaMemcpy2DAsync(A12, (size_t)lda * 8, W, (size_t)TB_NB * 8, (size_t)nbj * 8, nrest, cudaMemcpyDeviceToDevice, st) != cudaSuccess)
                return PB2_ERR_CUDA;
            rc = parsec_b200_dsyrk_sm100('U', 'T', nrest, nbj, -1.0, W, TB_NB, 1.0, A + (size_t)r0 + (size_t)r0 * lda, lda, stream);
        }
        if (rc != PB2_SUCCESS) return rc;
    }
    return cudaGetLastError() == cudaSuccess ? PB2_SUCCESS : PB2_ERR_CUDA;
}

extern "C" int parsec_b200_dpotrf(char uplo, int n, double* A, int lda, void* workspace, size_t workspace_bytes, int* dev_info,
                                   void* stream) {
    return parsec_b200_dpotrf_sm100(uplo, n, A, lda, workspace, workspace_bytes, dev_info, stream);
}
