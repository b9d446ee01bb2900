/*
 * CPU oracle (see pb2_oracle.h: test infrastructure only).  All citations are file:line under
 * /root/reference.
 */
#include "pb2_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ========================================================================================
 * Input generator.  tests/dsl/dtd/dtd_test_simple_gemm.c:43-48 (constants), :50-66 (jump-ahead),
 * :68-91 (initialize_tile).  64-bit LCG x <- a*x + c; element value is computed in *single*
 * precision (the reference multiplies by the float constant RndF_Mul and subtracts from 0.5f)
 * and then widened to double.  The jump uses the TILE coordinates (m + n*M), not element
 * coordinates: that quirk is the reference's and is kept.
 * ====================================================================================== */
#define ORC_RND_A 6364136223846793005ULL
#define ORC_RND_C 1ULL

uint64_t orc_rnd64_jump(uint64_t n, uint64_t seed) {
    uint64_t mul = ORC_RND_A, add = ORC_RND_C, x = seed;
    while (n) {
        if (n & 1ULL) x = mul * x + add;
        add *= (mul + 1ULL);
        mul *= mul;
        n >>= 1;
    }
    return x;
}

void orc_init_tile(double* data, int m, int n, int mb, int nb, int M, int ld, unsigned seed) {
    uint64_t jump = (uint64_t)m + (uint64_t)n * (uint64_t)M;
    for (int col = 0; col < nb; ++col, jump += (uint64_t)M) {
        uint64_t x = orc_rnd64_jump(jump, seed);
        double* out = data + (size_t)col * ld;
        for (int row = 0; row < mb; ++row) {
            volatile float scaled = (float)x * 5.4210108624275222e-20f; /* two roundings, no FMA */
            float v = 0.5f - scaled;
            out[row] = (double)v;
            x = ORC_RND_A * x + ORC_RND_C;
        }
    }
}

/* ========================================================================================
 * 2D block-cyclic tiled matrix.  parsec/data_dist/matrix/matrix.c:56-125 (tiled descriptor:
 * lmt/lnt, padded lm/ln, mt/nt, bsiz), grid_2Dcyclic.c:18-37 (rrank/crank),
 * two_dim_rectangle_cyclic.c:109-236 (local tile counts), :258-286 (rank_of),
 * :333-349 (local position), :368-412 (data_of: key = n*lmt + m, tile storage offset).
 * ====================================================================================== */
static int count_local(int first, int k, int procs, int total) {
    /* tiles owned along one dimension: k-blocks dealt round-robin starting at block `first` */
    int have = 0, at = first * k;
    while (at < total) {
        if (at + k < total) {
            have += k;
            at += procs * k;
            continue;
        }
        have += total - at;
        break;
    }
    return have;
}

void orc_bc_init(orc_bc_t* dc, int myrank, int mb, int nb, int lm, int ln, int i, int j, int m, int n, int P, int Q,
                 int kp, int kq, int ip, int jq) {
    memset(dc, 0, sizeof(*dc));
    dc->mb = mb; dc->nb = nb; dc->bsiz = mb * nb;
    dc->lmt = (lm + mb - 1) / mb;
    dc->lnt = (ln + nb - 1) / nb;
    dc->lm = dc->lmt * mb; /* tile storage pads to whole tiles */
    dc->ln = dc->lnt * nb;
    dc->i = i; dc->j = j; dc->m = m; dc->n = n;
    dc->mt = (i + m - 1) / mb - i / mb + 1;
    dc->nt = (j + n - 1) / nb - j / nb + 1;
    dc->P = P; dc->Q = Q; dc->kp = kp; dc->kq = kq; dc->ip = ip; dc->jq = jq;
    dc->myrank = myrank;
    dc->rrank = ((myrank / Q) + (P - ip)) % P;
    dc->crank = ((myrank % Q) + (Q - jq)) % Q;
    dc->nb_elem_r = count_local(dc->rrank, kp, P, dc->lmt);
    dc->nb_elem_c = count_local(dc->crank, kq, Q, dc->lnt);
    if (dc->nb_elem_r == 0) dc->nb_elem_c = 0;
    if (dc->nb_elem_c == 0) dc->nb_elem_r = 0;
    dc->nb_local_tiles = dc->nb_elem_r * dc->nb_elem_c;
    dc->llm = dc->nb_elem_r * mb;
    dc->lln = dc->nb_elem_c * nb;
}

/* parsec_matrix_block_cyclic_lapack_init (two_dim_rectangle_cyclic.c:61-107): the block-cyclic init with
 * PARSEC_MATRIX_LAPACK storage -- lm / ln are NOT padded to whole tiles (matrix.c:103-107) -- and then llm / lln
 * overwritten with the local matrix size the caller gives. */
void orc_bc_init_lapack(orc_bc_t* dc, int myrank, int mb, int nb, int lm, int ln, int i, int j, int m, int n, int P,
                        int Q, int kp, int kq, int ip, int jq, int mloc, int nloc) {
    orc_bc_init(dc, myrank, mb, nb, lm, ln, i, j, m, n, P, Q, kp, kq, ip, jq);
    dc->lm = lm;
    dc->ln = ln;
    dc->llm = mloc;
    dc->lln = nloc;
    dc->lapack = 1;
}

/* twoDBC_data_of (:397-411) / twoDBC_kcyclic_data_of (:666-675): tile storage puts tile `position` at position * bsiz;
 * LAPACK storage addresses element (local_m * mb, local_n * nb) of the column-major local matrix with ld = llm. */
int64_t orc_bc_offset(const orc_bc_t* dc, int m, int n) {
    if (!dc->lapack) return (int64_t)orc_bc_position(dc, m, n) * dc->bsiz;
    m += dc->i / dc->mb;
    n += dc->j / dc->nb;
    int64_t local_m = (m / (dc->kp * dc->P)) * dc->kp + (m % (dc->kp * dc->P)) % dc->kp;
    int64_t local_n = (n / (dc->kq * dc->Q)) * dc->kq + (n % (dc->kq * dc->Q)) % dc->kq;
    return local_n * dc->nb * dc->llm + local_m * dc->mb;
}

int orc_bc_rank_of(const orc_bc_t* dc, int m, int n) {
    m += dc->i / dc->mb;
    n += dc->j / dc->nb;
    /* (m, n) lies in k-cyclic super-tile (m / kp, n / kq): two_dim_rectangle_cyclic.c:532-566; with kp = kq = 1 this
     * is twoDBC_rank_of, :258-286 */
    int rr = ((m / dc->kp) % dc->P + dc->ip) % dc->P;
    int cr = ((n / dc->kq) % dc->Q + dc->jq) % dc->Q;
    return rr * dc->Q + cr;
}

int orc_bc_position(const orc_bc_t* dc, int m, int n) {
    m += dc->i / dc->mb;
    n += dc->j / dc->nb;
    /* twoDBC_kcyclic_data_of, :652-664 (twoDBC_coordinates_to_position, :333-349, when kp = kq = 1) */
    int local_m = (m / (dc->kp * dc->P)) * dc->kp + (m % (dc->kp * dc->P)) % dc->kp;
    int local_n = (n / (dc->kq * dc->Q)) * dc->kq + (n % (dc->kq * dc->Q)) % dc->kq;
    return dc->nb_elem_r * local_n + local_m;
}

uint64_t orc_bc_key(const orc_bc_t* dc, int m, int n) {
    m += dc->i / dc->mb;
    n += dc->j / dc->nb;
    return (uint64_t)n * (uint64_t)dc->lmt + (uint64_t)m;
}

/* The key stored in the parsec_data_t that data_of creates.  Plain cyclic: the data key.  k-cyclic:
 * twoDBC_kcyclic_data_of reduces m and n modulo the super-tile period while it computes the local position
 * (:652-664) and then builds the key from the REDUCED coordinates (:676-678), so tiles one period apart share a
 * key.  Reproduced as it is (the reference's own behaviour is the contract), see DESIGN.md. */
uint64_t orc_bc_data_of_key(const orc_bc_t* dc, int m, int n) {
    m += dc->i / dc->mb;
    n += dc->j / dc->nb;
    if (dc->kp > 1 || dc->kq > 1) {
        m %= dc->kp * dc->P;
        n %= dc->kq * dc->Q;
    }
    return (uint64_t)n * (uint64_t)dc->lmt + (uint64_t)m;
}

void orc_bc_key2coords(const orc_bc_t* dc, uint64_t key, int* m, int* n) {
    *m = (int)(key % (uint64_t)dc->lmt) - dc->i / dc->mb;
    *n = (int)(key / (uint64_t)dc->lmt) - dc->j / dc->nb;
}

/* create_initialize_matrix in the reference inserts one init task per (i, j) with
 * (m, n, mb, nb, M, ld) = (i, j, mb, nb, mat->m, mb): dtd_test_simple_gemm.c:118-131. */
void orc_bc_fill(const orc_bc_t* dc, double* mat, unsigned seed, int nthreads) {
    int total = dc->mt * dc->nt;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads > 0 ? nthreads : 1)
#endif
    for (int idx = 0; idx < total; ++idx) {
        int m = idx % dc->mt, n = idx / dc->mt;
        if (orc_bc_rank_of(dc, m, n) != dc->myrank) continue;
        double* tile = mat + (size_t)orc_bc_position(dc, m, n) * dc->bsiz;
        orc_init_tile(tile, m, n, dc->mb, dc->nb, dc->m, dc->mb, seed);
    }
    (void)nthreads;
}

double orc_matrix_inf_norm_tiled(const orc_bc_t* dc, const double* mat) {
    /* inf-norm = max row sum of |a_ij|; single rank only */
    double* rows = (double*)calloc((size_t)dc->m, sizeof(double));
    for (int n = 0; n < dc->nt; ++n)
        for (int m = 0; m < dc->mt; ++m) {
            const double* tile = mat + (size_t)orc_bc_position(dc, m, n) * dc->bsiz;
            int rm = (m == dc->mt - 1) ? dc->m - m * dc->mb : dc->mb;
            int rn = (n == dc->nt - 1) ? dc->n - n * dc->nb : dc->nb;
            for (int c = 0; c < rn; ++c)
                for (int r = 0; r < rm; ++r) rows[(size_t)m * dc->mb + r] += fabs(tile[(size_t)c * dc->mb + r]);
        }
    double best = 0;
    for (int r = 0; r < dc->m; ++r)
        if (rows[r] > best) best = rows[r];
    free(rows);
    return best;
}

/* ========================================================================================
 * Tile GEMM.  Shape of the call in every CUDA body of the reference: column-major, N/N,
 * (mb, nb, kb), lda = mb, ldb = kb, ldc = mb (tests/dsl/dtd/dtd_test_simple_gemm.c:197-202,
 * tests/dsl/ptg/cuda/nvlink.jdf:142-147).  The arithmetic itself lives in cuBLAS / CBLAS, which
 * the reference does not vendor; this is the textbook definition C <- alpha*A*B + beta*C with the
 * k-sum taken in ascending order.
 * ====================================================================================== */
/* The definition, one column at a time.  Kept as the statement of what orc_dgemm_tile computes (and checked against it
 * bit for bit in tests/test_oracle.py). */
void orc_dgemm_tile_simple(int mb, int nb, int kb, double alpha, const double* A, int lda, const double* B, int ldb,
                           double beta, double* C, int ldc) {
    double* col = (double*)malloc(sizeof(double) * (size_t)(mb > 0 ? mb : 1));
    for (int j = 0; j < nb; ++j) {
        for (int i = 0; i < mb; ++i) col[i] = 0.0;
        const double* bj = B + (size_t)j * ldb;
        for (int k = 0; k < kb; ++k) {
            const double bkj = bj[k];
            const double* ak = A + (size_t)k * lda;
            for (int i = 0; i < mb; ++i) col[i] += ak[i] * bkj;
        }
        double* cj = C + (size_t)j * ldc;
        if (beta == 0.0)
            for (int i = 0; i < mb; ++i) cj[i] = alpha * col[i];
        else
            for (int i = 0; i < mb; ++i) cj[i] = alpha * col[i] + beta * cj[i];
    }
    free(col);
}

/* Same values, bit for bit (every element is still sum_k a_ik * b_kj taken in ascending k with a separate multiply and
 * add, then alpha * sum + beta * c), arranged so that the host cores are not waiting on memory: 8 rows of A are packed
 * into a k-major panel once and swept against 4 columns of B with the 8 x 4 sums held in registers.  About 4x the
 * one-column-at-a-time form; still plain C without fused multiply-add, not a tuned BLAS. */
#include <immintrin.h>
void orc_dgemm_tile(int mb, int nb, int kb, double alpha, const double* A, int lda, const double* B, int ldb,
                    double beta, double* C, int ldc) {
    if (mb <= 0 || nb <= 0) return;
    double* Ap = (double*)aligned_alloc(64, sizeof(double) * 8 * (size_t)(kb > 0 ? kb : 1));
    for (int i0 = 0; i0 < mb; i0 += 8) {
        const int ib = mb - i0 < 8 ? mb - i0 : 8;
        for (int k = 0; k < kb; ++k)
            for (int r = 0; r < 8; ++r) Ap[(size_t)k * 8 + r] = r < ib ? A[(size_t)k * lda + i0 + r] : 0.0;
        for (int j0 = 0; j0 < nb; j0 += 4) {
            const int jb = nb - j0 < 4 ? nb - j0 : 4;
            double sum[4][8];
            if (jb == 4) {
                __m256d c00 = _mm256_setzero_pd(), c01 = c00, c10 = c00, c11 = c00, c20 = c00, c21 = c00, c30 = c00, c31 = c00;
                const double *b0 = B + (size_t)j0 * ldb, *b1 = b0 + ldb, *b2 = b1 + ldb, *b3 = b2 + ldb;
                for (int k = 0; k < kb; ++k) {
                    const __m256d a0 = _mm256_load_pd(Ap + (size_t)k * 8), a1 = _mm256_load_pd(Ap + (size_t)k * 8 + 4);
                    __m256d b = _mm256_broadcast_sd(b0 + k);
                    c00 = _mm256_add_pd(c00, _mm256_mul_pd(a0, b));
                    c01 = _mm256_add_pd(c01, _mm256_mul_pd(a1, b));
                    b = _mm256_broadcast_sd(b1 + k);
                    c10 = _mm256_add_pd(c10, _mm256_mul_pd(a0, b));
                    c11 = _mm256_add_pd(c11, _mm256_mul_pd(a1, b));
                    b = _mm256_broadcast_sd(b2 + k);
                    c20 = _mm256_add_pd(c20, _mm256_mul_pd(a0, b));
                    c21 = _mm256_add_pd(c21, _mm256_mul_pd(a1, b));
                    b = _mm256_broadcast_sd(b3 + k);
                    c30 = _mm256_add_pd(c30, _mm256_mul_pd(a0, b));
                    c31 = _mm256_add_pd(c31, _mm256_mul_pd(a1, b));
                }
                _mm256_storeu_pd(sum[0], c00); _mm256_storeu_pd(sum[0] + 4, c01);
                _mm256_storeu_pd(sum[1], c10); _mm256_storeu_pd(sum[1] + 4, c11);
                _mm256_storeu_pd(sum[2], c20); _mm256_storeu_pd(sum[2] + 4, c21);
                _mm256_storeu_pd(sum[3], c30); _mm256_storeu_pd(sum[3] + 4, c31);
            } else {
                for (int jj = 0; jj < jb; ++jj) {
                    for (int r = 0; r < 8; ++r) sum[jj][r] = 0.0;
                    const double* bj = B + (size_t)(j0 + jj) * ldb;
                    for (int k = 0; k < kb; ++k) {
                        const double bkj = bj[k];
                        for (int r = 0; r < 8; ++r) sum[jj][r] += Ap[(size_t)k * 8 + r] * bkj;
                    }
                }
            }
            for (int jj = 0; jj < jb; ++jj) {
                double* cj = C + (size_t)(j0 + jj) * ldc + i0;
                if (beta == 0.0)
                    for (int r = 0; r < ib; ++r) cj[r] = alpha * sum[jj][r];
                else
                    for (int r = 0; r < ib; ++r) cj[r] = alpha * sum[jj][r] + beta * cj[r];
            }
        }
    }
    free(Ap);
}

uint16_t orc_f32_to_bf16_rn(float x) {
    uint32_t u;
    memcpy(&u, &x, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40u); /* quiet NaN */
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
static float bf16_to_f32(uint16_t h) {
    uint32_t u = (uint32_t)h << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}
void orc_gemm_tile_bf16(int mb, int nb, int kb, float alpha, const uint16_t* A, int lda, const uint16_t* B, int ldb,
                        float beta, float* C, int ldc) {
    double* col = (double*)malloc(sizeof(double) * (size_t)(mb > 0 ? mb : 1));
    for (int j = 0; j < nb; ++j) {
        for (int i = 0; i < mb; ++i) col[i] = 0.0;
        for (int k = 0; k < kb; ++k) {
            const double bkj = (double)bf16_to_f32(B[(size_t)j * ldb + k]);
            for (int i = 0; i < mb; ++i) col[i] += (double)bf16_to_f32(A[(size_t)k * lda + i]) * bkj;
        }
        for (int i = 0; i < mb; ++i) {
            float* c = C + (size_t)j * ldc + i;
            *c = (float)((double)alpha * col[i] + (beta == 0.0f ? 0.0 : (double)beta * (double)*c));
        }
    }
    free(col);
}

/* ========================================================================================
 * Tiled GEMM DAG (the DPLASMA-style dgemm_NN the reference's runtime
 * executes; DAG shape follows tests/dsl/ptg/cuda/stress.jdf:111-127 / nvlink.jdf:115-130):
 *   READ_A(m,k), READ_B(k,n): empty bodies that forward the tile;
 *   GEMM(m,n,k): RW C chained over k, affinity descC(m,n).
 * The host-copy version of a C tile is bumped once per GEMM task by the generated
 * complete_hook (parsec/interfaces/ptg/ptg-compiler/jdf2c.c:7102), so it ends at KT.
 * ====================================================================================== */
int orc_tiled_dgemm(int M, int N, int K, int mb, int nb, int kb, const double* A, const double* B, double* C,
                    int nthreads, orc_counts_t* counts, uint32_t* c_versions) {
    if (M <= 0 || N <= 0 || K <= 0 || mb <= 0 || nb <= 0 || kb <= 0) return -4;
    const int MT = (M + mb - 1) / mb, NT = (N + nb - 1) / nb, KT = (K + kb - 1) / kb;
    int64_t n_gemm = 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : n_gemm) num_threads(nthreads > 0 ? nthreads : 1)
#endif
    for (int idx = 0; idx < MT * NT; ++idx) {
        const int m = idx % MT, n = idx / MT;
        const int rm = (m == MT - 1) ? M - m * mb : mb;
        const int rn = (n == NT - 1) ? N - n * nb : nb;
        double* c = C + ((size_t)n * MT + m) * (size_t)mb * nb;
        uint32_t version = 0;
        for (int k = 0; k < KT; ++k) {
            const int rk = (k == KT - 1) ? K - k * kb : kb;
            const double* a = A + ((size_t)k * MT + m) * (size_t)mb * kb;
            const double* b = B + ((size_t)n * KT + k) * (size_t)kb * nb;
            orc_dgemm_tile(rm, rn, rk, 1.0, a, mb, b, kb, 1.0, c, mb);
            ++version;
            ++n_gemm;
        }
        if (c_versions) c_versions[idx] = version;
    }
    if (counts) {
        counts->gemm = n_gemm;
        counts->read_a = (int64_t)MT * KT;
        counts->read_b = (int64_t)KT * NT;
    }
    (void)nthreads;
    return 0;
}

/* ========================================================================================
 * Data copies: coherency and versions.
 *   parsec_data_create                      parsec/data.c:484-522  (host copy OWNED, owner 0)
 *   start_transfer_ownership_to_copy        parsec/data.c:319-433
 *   end_transfer_ownership_to_copy          parsec/data.c:299-317
 * ====================================================================================== */
void orc_data_create(orc_data_t* d, uint64_t key, int nb_devices) {
    memset(d, 0, sizeof(*d));
    d->key = key;
    d->owner_device = 0;
    d->preferred_device = -1;
    d->nb_devices = nb_devices;
    d->copies[0].present = 1;
    d->copies[0].coherency = ORC_COH_OWNED;
}

int orc_data_start_transfer_ownership(orc_data_t* d, int device, int access) {
    orc_copy_t* mine = &d->copies[device];
    int need = 0, source = d->owner_device;
    switch (mine->coherency) {
        case ORC_COH_INVALID:
            need = 1;
            if (source == -1)
                for (int i = 0; i < d->nb_devices; ++i)
                    if (d->copies[i].present && d->copies[i].coherency != ORC_COH_INVALID) source = i;
            break;
        case ORC_COH_SHARED:
            for (int i = 0; i < d->nb_devices; ++i)
                if (d->copies[i].present && d->copies[i].coherency == ORC_COH_OWNED &&
                    d->copies[i].version > mine->version)
                    need = 1;
            break;
        default: /* EXCLUSIVE, OWNED: already the freshest */
            break;
    }
    if (access & ORC_ACCESS_READ) {
        for (int i = 0; i < d->nb_devices; ++i) {
            orc_copy_t* o = &d->copies[i];
            if (i == device || !o->present || o->coherency == ORC_COH_INVALID) continue;
            if (mine->coherency == ORC_COH_OWNED && !(access & ORC_ACCESS_WRITE)) {
                if (o->version < mine->version) o->coherency = ORC_COH_INVALID;
                d->owner_device = -1;
            }
            if (o->coherency == ORC_COH_EXCLUSIVE) o->coherency = ORC_COH_SHARED;
        }
    } else {
        need = 0; /* write-only: will be overwritten, nothing to fetch */
    }
    if (access & ORC_ACCESS_WRITE)
        for (int i = 0; i < d->nb_devices; ++i)
            if (d->copies[i].present && d->copies[i].coherency != ORC_COH_INVALID)
                d->copies[i].coherency = ORC_COH_SHARED;
    if (access & ORC_ACCESS_READ) mine->readers++;
    if (access & ORC_ACCESS_WRITE) d->owner_device = (int8_t)device;
    return need ? source : -1;
}

void orc_data_end_transfer_ownership(orc_data_t* d, int device, int access) {
    if (access & ORC_ACCESS_READ) d->copies[device].coherency = ORC_COH_SHARED;
    if (access & ORC_ACCESS_WRITE) d->copies[device].coherency = ORC_COH_OWNED;
}

/* ---- device-module steps on ONE flow ---------------------------------------------------
 * reserve:        parsec/mca/device/cuda/device_cuda_module.c:863-1168 (new copy: INVALID, version 0)
 *                 returns 1 = copy created, 0 = already there, -1 = copy still UNDER_TRANSFER (AGAIN)
 * stage_in:       :1260-1530.  WRITE bumps the device version (:1383-1390); when a transfer is scheduled the
 *                 device version is overwritten with the source's (:1498) and the copy is UNDER_TRANSFER.
 *                 returns 1 = H2D scheduled, 0 = no transfer, -1 = write with foreign readers.
 * complete_push:  :1765-1942 (COMPLETE_TRANSFER + end of ownership transfer; only for the pusher)
 * pop:            :2196-2359 (readers--, read-only copies back to gpu_mem_lru, D2H when pushout)
 *                 returns 1 = copy re-filed in gpu_mem_lru, 2 = D2H scheduled, 0 = nothing
 * epilog:         :2364-2453 (both copies SHARED, cpu.version = gpu.version; dirty copy -> owned LRU)
 *                 returns 1 = gpu_mem_lru, 2 = gpu_mem_owned_lru, 0 = not a write flow
 * complete_hook:  generated code, jdf2c.c:7102: data_out->version++ for WRITE flows; data_out is the
 *                 host copy after the epilog.
 * w2r:            parsec/mca/device/transfer_gpu.c:221-338 (the product really copies; the reference's
 *                 stage_out == NULL defect, SURVEY.md Appendix B-1, is not reproduced).
 */
int orc_gpu_reserve(orc_data_t* d, int device) {
    orc_copy_t* c = &d->copies[device];
    if (c->present) return c->transfer_status == ORC_XFER_UNDER ? -1 : 0;
    memset(c, 0, sizeof(*c));
    c->present = 1;
    c->coherency = ORC_COH_INVALID;
    return 1;
}

int orc_gpu_stage_in(orc_data_t* d, int device, int access) {
    orc_copy_t* c = &d->copies[device];
    if ((access & ORC_ACCESS_WRITE) && c->readers > 0 && !(c->readers == 1 && (access & ORC_ACCESS_READ))) return -1;
    int from = orc_data_start_transfer_ownership(d, device, access);
    if (access & ORC_ACCESS_WRITE) c->version++;
    if (from == -1) {
        c->transfer_status = ORC_XFER_COMPLETE;
        orc_data_end_transfer_ownership(d, device, access);
        return 0;
    }
    if (c->transfer_status != ORC_XFER_UNDER) {
        c->version = d->copies[0].version; /* data_in is always the host copy on this path */
        c->transfer_status = ORC_XFER_UNDER;
    }
    return 1;
}

void orc_gpu_complete_push(orc_data_t* d, int device, int access) {
    d->copies[device].transfer_status = ORC_XFER_COMPLETE;
    orc_data_end_transfer_ownership(d, device, access);
}

int orc_gpu_pop(orc_data_t* d, int device, int access, int pushout) {
    orc_copy_t* c = &d->copies[device];
    if (access & ORC_ACCESS_READ) {
        c->readers--;
        if (c->readers == 0 && !(access & ORC_ACCESS_WRITE)) return 1;
    }
    if ((access & ORC_ACCESS_WRITE) && pushout) return 2;
    return 0;
}

int orc_gpu_epilog(orc_data_t* d, int device, int access, int pushout) {
    if (!(access & ORC_ACCESS_WRITE)) return 0;
    d->copies[device].coherency = ORC_COH_SHARED;
    d->copies[0].coherency = ORC_COH_SHARED;
    d->copies[0].version = d->copies[device].version;
    return pushout ? 1 : 2;
}

void orc_complete_hook(orc_data_t* d, int access) {
    if (access & ORC_ACCESS_WRITE) d->copies[0].version++;
}

void orc_gpu_w2r_start(orc_data_t* d, int device) {
    d->copies[device].readers++;
    d->copies[device].transfer_status = ORC_XFER_UNDER;
}

int orc_gpu_w2r_complete(orc_data_t* d, int device) {
    orc_copy_t* g = &d->copies[device];
    g->readers--;
    g->transfer_status = ORC_XFER_COMPLETE;
    if (d->copies[0].version < g->version) return 0;
    g->coherency = ORC_COH_SHARED;
    d->copies[0].coherency = ORC_COH_SHARED;
    /* transfer_gpu.c:321 assigns cpu.version = gpu.version, i.e. rolls the host version back by one after a
     * completed writer (cpu == gpu + 1 at that point).  Together with the missing copy (Appendix B-1) that path
     * cannot produce the CPU result; the oracle keeps the host version, so an evicted chain still ends at KT,
     * which is what the CPU-body run of the reference yields. */
    return 1;
}

void orc_gpu_evict(orc_data_t* d, int device) { memset(&d->copies[device], 0, sizeof(orc_copy_t)); }

/* ---- in-core trace of the whole DAG ------------------------------------------------------ */
typedef struct {
    orc_event_t* ev;
    int64_t n, cap;
} evbuf_t;

static void emit(evbuf_t* b, uint32_t dc, const orc_data_t* d, int device, int kind, int k, int transfer) {
    if (b->ev && b->n < b->cap) {
        orc_event_t* e = &b->ev[b->n];
        memset(e, 0, sizeof(*e));
        e->dc = dc;
        e->key = (uint32_t)d->key;
        e->kind = kind;
        e->k = k;
        e->gpu_version = d->copies[device].version;
        e->cpu_version = d->copies[0].version;
        e->gpu_coh = d->copies[device].present ? d->copies[device].coherency : 0xff;
        e->cpu_coh = d->copies[0].coherency;
        e->owner_device = d->owner_device;
        e->transfer = (int8_t)transfer;
    }
    b->n++;
}

int64_t orc_trace_tiled_gemm(int MT, int NT, int KT, int device, orc_event_t* events, int64_t max_events) {
    evbuf_t buf = {events, 0, max_events};
    const int ndev = device + 1;
    orc_data_t* A = (orc_data_t*)malloc(sizeof(orc_data_t) * (size_t)MT * KT);
    orc_data_t* B = (orc_data_t*)malloc(sizeof(orc_data_t) * (size_t)KT * NT);
    orc_data_t* C = (orc_data_t*)malloc(sizeof(orc_data_t) * (size_t)MT * NT);
    for (int k = 0; k < KT; ++k)
        for (int m = 0; m < MT; ++m) orc_data_create(&A[(size_t)k * MT + m], (uint64_t)k * MT + m, ndev);
    for (int n = 0; n < NT; ++n)
        for (int k = 0; k < KT; ++k) orc_data_create(&B[(size_t)n * KT + k], (uint64_t)n * KT + k, ndev);
    for (int n = 0; n < NT; ++n)
        for (int m = 0; m < MT; ++m) orc_data_create(&C[(size_t)n * MT + m], (uint64_t)n * MT + m, ndev);

    for (int k = 0; k < KT; ++k)
        for (int n = 0; n < NT; ++n)
            for (int m = 0; m < MT; ++m) {
                orc_data_t* flow[3] = {&A[(size_t)k * MT + m], &B[(size_t)n * KT + k], &C[(size_t)n * MT + m]};
                const int access[3] = {ORC_ACCESS_READ, ORC_ACCESS_READ, ORC_ACCESS_RW};
                const int pushout[3] = {0, 0, k == KT - 1};
                int xfer[3];
                for (int f = 0; f < 3; ++f)
                    if (orc_gpu_reserve(flow[f], device) == 1) emit(&buf, f, flow[f], device, ORC_EV_RESERVE, k, 0);
                for (int f = 0; f < 3; ++f) {
                    xfer[f] = orc_gpu_stage_in(flow[f], device, access[f]);
                    emit(&buf, f, flow[f], device, ORC_EV_STAGE_IN, k, xfer[f] == 1);
                }
                for (int f = 0; f < 3; ++f)
                    if (xfer[f] == 1) {
                        orc_gpu_complete_push(flow[f], device, access[f]);
                        emit(&buf, f, flow[f], device, ORC_EV_PUSH_DONE, k, 0);
                    }
                for (int f = 0; f < 3; ++f) {
                    int r = orc_gpu_pop(flow[f], device, access[f], pushout[f]);
                    emit(&buf, f, flow[f], device, ORC_EV_POP, k, r == 2);
                }
                for (int f = 0; f < 3; ++f)
                    if (orc_gpu_epilog(flow[f], device, access[f], pushout[f]))
                        emit(&buf, f, flow[f], device, ORC_EV_EPILOG, k, 0);
                for (int f = 0; f < 3; ++f)
                    if (access[f] & ORC_ACCESS_WRITE) {
                        orc_complete_hook(flow[f], access[f]);
                        emit(&buf, f, flow[f], device, ORC_EV_COMPLETE, k, 0);
                    }
            }
    free(A);
    free(B);
    free(C);
    return buf.n;
}

/* ========================================================================================
 * Zone allocator.  parsec/utils/zone_malloc.c:24-60 (init), :71-122 (malloc: first fit scanning
 * from the last remembered free segment, wrapping once), :124-176 (free with coalescing; the
 * freed segment becomes the next starting point), :178-193 (in_use).  Segments are addressed
 * by their first unit; `before` is the size of the preceding segment.
 * ====================================================================================== */
typedef struct {
    int busy, units, before;
} orc_seg_t;
struct orc_zone_s {
    orc_seg_t* seg;
    int nseg, cursor;
    size_t unit;
};

static orc_seg_t* seg_at(const orc_zone_t* z, int id) { return (id < 0 || id >= z->nseg) ? NULL : &z->seg[id]; }

orc_zone_t* orc_zone_init(int max_segment, size_t unit_size) {
    orc_zone_t* z = (orc_zone_t*)malloc(sizeof(*z));
    z->seg = (orc_seg_t*)calloc((size_t)max_segment, sizeof(orc_seg_t));
    z->nseg = max_segment;
    z->unit = unit_size;
    z->cursor = 0;
    z->seg[0].busy = 0;
    z->seg[0].units = max_segment;
    z->seg[0].before = 1; /* so that "previous of 0" is out of range */
    return z;
}
void orc_zone_fini(orc_zone_t* z) {
    free(z->seg);
    free(z);
}

int64_t orc_zone_malloc(orc_zone_t* z, size_t size) {
    const int want = (int)((size + z->unit - 1) / z->unit);
    int id = z->cursor, wrapped = 0;
    do {
        orc_seg_t* s = seg_at(z, id);
        if (!s) {
            if (wrapped) return -1;
            wrapped = 1;
            id = 0;
            s = seg_at(z, 0);
        }
        if (!s->busy && s->units >= want) {
            s->busy = 1;
            if (s->units > want) {
                orc_seg_t* after = seg_at(z, id + s->units);
                if (after) after->before -= want;
                orc_seg_t* rest = seg_at(z, id + want);
                rest->busy = 0;
                rest->before = want;
                rest->units = s->units - want;
                z->cursor = id + want;
                s->units = want;
            }
            return (int64_t)id * (int64_t)z->unit;
        }
        id += s->units;
    } while (id != z->cursor);
    return -1;
}

int orc_zone_free(orc_zone_t* z, int64_t offset) {
    int id = (int)(offset / (int64_t)z->unit);
    orc_seg_t* s = seg_at(z, id);
    if (!s || !s->busy) return -1;
    s->busy = 0;
    int pid = id - s->before, nid = id + s->units;
    orc_seg_t* prev = seg_at(z, pid);
    orc_seg_t* next = seg_at(z, nid);
    if (prev && !prev->busy) {
        if (next) next->before += prev->units;
        prev->units += s->units;
        s = prev;
        id = pid;
    }
    z->cursor = id;
    if (next && !next->busy) {
        nid += next->units;
        s->units += next->units;
        next = seg_at(z, nid);
        if (next) next->before = s->units;
    }
    return 0;
}

size_t orc_zone_in_use(const orc_zone_t* z) {
    size_t used = 0;
    for (int id = 0; id < z->nseg;) {
        const orc_seg_t* s = &z->seg[id];
        if (s->busy) used += z->unit * (size_t)s->units;
        id += s->units;
    }
    return used;
}
