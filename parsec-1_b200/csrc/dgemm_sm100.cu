// FP64 tile DGEMM for sm_100a: C = alpha * op(A) * op(B) + beta * C, column-major.
//
// This is the kernel a `BODY [type=CUDA dyld=parsec_b200_dgemm]` task body launches on the
// exec stream the device module hands it.  It replaces the cublasDgemm_v2 call of the
// reference's CUDA bodies (tests/dsl/dtd/dtd_test_simple_gemm.c:197-202,
// tests/dsl/ptg/cuda/nvlink.jdf:142, stress.jdf:135): same argument order and meaning,
// column-major, leading dimensions in elements.
//
// Fast path (N,N; M%128 == N%128 == 0; K%16 == 0; 16-byte aligned operands, even lds):
//   CTA tile 128x128, BK = 16, 256 threads = 8 warps in a 2(m) x 4(n) grid, warp tile
//   64x32 = 8x4 DMMA.8x8x4 tiles (mma.sync.m8n8k4.f64 is the native FP64 tensor op on
//   sm_100a; wider shapes lower to the same SASS).  Operands are staged with 16-byte
//   cp.async into a 4-deep shared-memory ring in their *global* orientation
//   (A: m fastest, B: k fastest), so no transpose is ever needed:
//     - A fragment rows are permuted (DMMA tile i, fragment row g  <->  tile row 8g+i) so one
//       thread reads 8 consecutive doubles of a k-column with four LDS.128; the row pitch
//       BM+2 makes the quarter-warp accesses hit 8 distinct 16-byte bank groups.
//     - B fragments are LDS.64 with pitch BK+4 (conflict free per half warp).
//   The same row permutation makes every thread own 8 consecutive rows of 8 columns of C,
//   so the epilogue is 16-byte loads/stores.
// Everything else (ragged tiles, transposes, unaligned) goes through a bounds-checked
// 64x64 DMMA kernel; it exists for the edge tiles of N % NB != 0 matrices.
#include <cuda_runtime.h>
#include <stdint.h>
#include "pb2_internal.h"
#include "tma_utils.cuh"
#include <stdlib.h>
#include <string.h>
#include <vector>

namespace pb2 {

constexpr int BM = 128, BN = 128, BK = 16;
constexpr int STAGES = 4;
constexpr int LDA_S = BM + 2;          // doubles; (LDA_S/2) % 8 == 1
constexpr int LDB_S = BK + 4;          // doubles; LDB_S % 16 == 4
constexpr int A_STAGE = BK * LDA_S;    // doubles per stage
constexpr int B_STAGE = BN * LDB_S;
constexpr int DGEMM_THREADS = 256;
constexpr size_t DGEMM_SMEM = (size_t)STAGES * (A_STAGE + B_STAGE) * sizeof(double);

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(DGEMM_THREADS, 1)
dgemm_nn_128x128(int M, int N, int K, double alpha, const double* __restrict__ A, int lda,
                 const double* __restrict__ B, int ldb, double beta, double* __restrict__ C, int ldc) {
    extern __shared__ __align__(16) double smem[];
    double* As = smem;
    double* Bs = smem + STAGES * A_STAGE;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp & 1, wn = warp >> 1;
    const int tiles_m = M / BM;
    const int bm = blockIdx.x % tiles_m, bn = blockIdx.x / tiles_m;
    (void)N;

    const double* Ag = A + (size_t)bm * BM;
    const double* Bg = B + (size_t)bn * BN * ldb;

    auto load_stage = [&](int stage, int kt) {
        double* as = As + stage * A_STAGE;
        const double* ag = Ag + (size_t)kt * BK * lda;
#pragma unroll
        for (int i = 0; i < (BK * BM / 2) / DGEMM_THREADS; ++i) {
            int c = tid + i * DGEMM_THREADS;
            int kr = c / (BM / 2), mc = c % (BM / 2);
            cp_async16(as + kr * LDA_S + mc * 2, ag + (size_t)kr * lda + mc * 2);
        }
        double* bs = Bs + stage * B_STAGE;
        const double* bg = Bg + (size_t)kt * BK;
#pragma unroll
        for (int i = 0; i < (BN * BK / 2) / DGEMM_THREADS; ++i) {
            int c = tid + i * DGEMM_THREADS;
            int n = c / (BK / 2), kc = c % (BK / 2);
            cp_async16(bs + n * LDB_S + kc * 2, bg + (size_t)n * ldb + kc * 2);
        }
    };

    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    const int KT = K / BK;
#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < KT) load_stage(s, s);
        cp_async_commit();
    }

    for (int kt = 0; kt < KT; ++kt) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        {
            int nk = kt + STAGES - 1;
            if (nk < KT) load_stage(nk % STAGES, nk);
            cp_async_commit();
        }
        const double* as = As + (kt % STAGES) * A_STAGE + wm * 64 + 8 * g;
        const double* bs = Bs + (kt % STAGES) * B_STAGE + (wn * 32 + g) * LDB_S + t;
#pragma unroll
        for (int ks = 0; ks < BK / 4; ++ks) {
            double a[8], b[4];
            const double* ap = as + (ks * 4 + t) * LDA_S;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                double2 v = *reinterpret_cast<const double2*>(ap + 2 * q);
                a[2 * q] = v.x;
                a[2 * q + 1] = v.y;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = bs[j * 8 * LDB_S + ks * 4];
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
    }
    cp_async_wait<0>();

    // Epilogue: thread owns rows [8g, 8g+8) of columns 8j + 2t + e of its warp tile.
    const size_t row0 = (size_t)bm * BM + wm * 64 + 8 * g;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const size_t col = (size_t)bn * BN + wn * 32 + 8 * j + 2 * t + e;
            double2* cp = reinterpret_cast<double2*>(C + row0 + col * ldc);
            if (beta == 0.0) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    double2 v;
                    v.x = alpha * acc[2 * q][j][e];
                    v.y = alpha * acc[2 * q + 1][j][e];
                    cp[q] = v;
                }
            } else {
                double2 old[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) old[q] = cp[q];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    double2 v;
                    v.x = alpha * acc[2 * q][j][e] + beta * old[q].x;
                    v.y = alpha * acc[2 * q + 1][j][e] + beta * old[q].y;
                    cp[q] = v;
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// TMA + mbarrier warp-specialised variant (the default fast path).
//   warps 0-7: consumers (2 x 4 grid of 64x32 warp tiles, DMMA.8x8x4), warp 8: TMA producer
//   (warps 9-11 only exist so the producer is a whole warpgroup and can setmaxnreg.dec).
//   Stage = A tile as 8 boxes of [16 k][16 m] + B tile as one box [128 n][16 k], both with
//   128-byte rows and CU_TENSOR_MAP_SWIZZLE_128B, so every LDS below is bank-conflict free:
//     A: LDS.128, quarter warp = (g&1, t) -> chunk ((g&1)^(ks&1))*4 + (q^t)      (8 distinct)
//     B: LDS.64,  half warp    = (g&3, t) -> slot 4*(ks^(g&3)) + 2*((t>>1)^(j&1)) + (t&1) (16 distinct)
//   B columns are permuted inside a warp tile (DMMA tile j, fragment column g <-> column
//   16*(j>>1) + 2g + (j&1)) to make the hardware swizzle (which XORs with n&7) conflict free.
//   There is no CTA-wide barrier in the main loop: consumers wait on full[stage], arrive on
//   empty[stage]; the first fragments of tile kt+1 are fetched under the last DMMAs of tile kt.
//   Out-of-range rows/columns/k are zero-filled by TMA, so any m, n, k is accepted.
constexpr int T_STAGES = 6;
constexpr int T_A_BYTES = BK * BM * 8;  // 16 KiB
constexpr int T_B_BYTES = BK * BN * 8;  // 16 KiB
constexpr int T_STAGE_BYTES = T_A_BYTES + T_B_BYTES;
constexpr int T_CONSUMER_WARPS = 8;
constexpr int T_THREADS = (T_CONSUMER_WARPS + 4) * 32;  // 2 consumer warpgroups + 1 producer warpgroup
constexpr size_t T_SMEM = (size_t)T_STAGES * T_STAGE_BYTES + 1024 + 2 * T_STAGES * sizeof(uint64_t);

__device__ __forceinline__ void lds128(double& x, double& y, uint32_t addr) {
    asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];\n" : "=d"(x), "=d"(y) : "r"(addr));
}
__device__ __forceinline__ double lds64(uint32_t addr) {
    double x;
    asm volatile("ld.shared.f64 %0, [%1];\n" : "=d"(x) : "r"(addr));
    return x;
}
__device__ __forceinline__ void dmma884p(double& d0, double& d1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
        : "+d"(d0), "+d"(d1)
        : "d"(a), "d"(b));
}

// alpha * acc + beta * c with the roundings spelled out (one multiply, one fused multiply-add), so that every
// instantiation of the tile code gives the same bits whatever the compiler would have contracted on its own
__device__ __forceinline__ double axpby(double alpha, double acc, double beta, double c) {
    return __fma_rn(beta, c, __dmul_rn(alpha, acc));
}

// One 128 x 128 tile of C = alpha op(A) op(B) + beta C; `tile` is its index among the tiles of this GEMM (column-major
// over tiles).  Shared by the one-GEMM kernels and the grouped one.
//
// Transposed operands (TA / TB = 1) change which of the two shared-memory layouts an operand arrives in, nothing
// else: op(A) = A^T is stored with k contiguous, which is the layout B has in the N,N case (one box [128 rows][16 k]),
// and op(B) = B^T is stored with n contiguous, the layout A has in the N,N case (8 boxes [16 k][16 cols]).  Each
// operand therefore has two fragment loaders, both conflict free under the 128-byte hardware swizzle:
//   "contiguous along its own dimension" (A: N, B: T): LDS.128 of two neighbouring rows (columns) at one k,
//       quarter warp (g&1, t) -> chunk ((g&1)^(ks&1))*4 + (x^t), 8 distinct;
//   "contiguous along k" (A: T, B: N): LDS.64, half warp (g&3, t) -> slot 4*(ks^(g&3)) + 2*((t>>1)^(x&1)) + (t&1).
// What a loader implies is the order in which a thread's DMMA fragments map to rows (columns) of the warp tile; the
// epilogue applies the same map.
// TRI = 1 / 2: only the lower / upper triangle of C (diagonal included) is computed and written -- the rank-k update
// C = alpha A A^T + beta C (SYRK) is this code with B = A; blocks entirely on the other side leave at once.
template <int TA, int TB, int TRI = 0>
__device__ __forceinline__ void dgemm_tile_tma(const CUtensorMap* mapA_p, const CUtensorMap* mapB_p, int M, int N, int K,
                                               double alpha, double beta, double* __restrict__ C, int ldc, int tile) {
    const CUtensorMap& mapA = *mapA_p;
    const CUtensorMap& mapB = *mapB_p;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t sbase0 = (raw + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (sbase0 - raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + T_STAGES * T_STAGE_BYTES);
    uint64_t* empty = full + T_STAGES;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tiles_m = (M + BM - 1) / BM;
    const int bm = tile % tiles_m, bn = tile / tiles_m;
    const int m0 = bm * BM, n0 = bn * BN;
    const int KT = (K + BK - 1) / BK;
    if (TRI == 1 && bn > bm) return;  // BM == BN: block (bm, bn) lies above / below the diagonal as a whole
    if (TRI == 2 && bn < bm) return;

    if (tid == 0) {
        for (int s = 0; s < T_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], T_CONSUMER_WARPS);
        }
        fence_barrier_init();
        fence_proxy_async();
    }
    __syncthreads();

    if (warp >= T_CONSUMER_WARPS) {
        // producer warpgroup: hand its registers to the consumers (168*384 = 232*256 + 40*128)
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n" ::);
        if (warp == T_CONSUMER_WARPS && lane == 0) {
            tma_prefetch_desc(&mapA);
            tma_prefetch_desc(&mapB);
            int stage = 0;
            uint32_t phase = 0;
            for (int kt = 0; kt < KT; ++kt) {
                if (kt >= T_STAGES) mbar_wait(&empty[stage], phase ^ 1);
                uint8_t* sa = smem + stage * T_STAGE_BYTES;
                mbar_arrive_expect_tx(&full[stage], T_STAGE_BYTES);
                if (TA == 0) {
#pragma unroll
                    for (int b = 0; b < BM / 16; ++b) tma_load_2d(sa + b * 2048, &mapA, m0 + 16 * b, kt * BK, &full[stage]);
                } else {
                    tma_load_2d(sa, &mapA, kt * BK, m0, &full[stage]);
                }
                if (TB == 0) {
                    tma_load_2d(sa + T_A_BYTES, &mapB, kt * BK, n0, &full[stage]);
                } else {
#pragma unroll
                    for (int b = 0; b < BN / 16; ++b)
                        tma_load_2d(sa + T_A_BYTES + b * 2048, &mapB, n0 + 16 * b, kt * BK, &full[stage]);
                }
                if (++stage == T_STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
        return;
    }

    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;\n" ::);
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp & 1, wn = warp >> 1;

    if (beta != 0.0) {
        // pull the old C tile towards L2 while the main loop runs: lane c takes column c of the warp tile, whose
        // 64 rows are four 128-byte lines
        const int col = n0 + wn * 32 + lane;
        const int rowb = m0 + wm * 64;
        if (col < N && rowb < M) {
            const double* cp = C + (size_t)rowb + (size_t)col * ldc;
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (rowb + 16 * i < M) asm volatile("prefetch.global.L2 [%0];\n" ::"l"(cp + 16 * i));
        }
    }

    // per-thread shared-memory offsets (bytes, relative to the stage base)
    // A, contiguous along m (TA = 0): box 4wm + (g>>1), half (g&1) of its 16 rows -> rows 8g .. 8g+7 of the warp tile
    // A, contiguous along k (TA = 1): fragment i of this thread is row 16(i>>1) + 2g + (i&1)
    const uint32_t a_thread = TA == 0 ? (uint32_t)((4 * wm + (g >> 1)) * 2048 + t * 128)
                                      : (uint32_t)((64 * wm + 2 * g) * 128 + (t & 1) * 8);
    uint32_t a_off[2][4];
#pragma unroll
    for (int p = 0; p < 2; ++p)
#pragma unroll
        for (int q = 0; q < 4; ++q) a_off[p][q] = (uint32_t)(((((g & 1) ^ p) * 4) + (q ^ t)) * 16);
    // B, contiguous along k (TB = 0): fragment j is column 16(j>>1) + 2g + (j&1)
    // B, contiguous along n (TB = 1): box 2wn + ((g>>1)&1), chunks 4(g&1) + 2(g>>2) + h -> columns
    //                                 16((g>>1)&1) + 8(g&1) + 4(g>>2) + j
    const uint32_t b_thread = TB == 0 ? (uint32_t)(T_A_BYTES + (32 * wn + 2 * g) * 128 + (t & 1) * 8)
                                      : (uint32_t)(T_A_BYTES + (2 * wn + ((g >> 1) & 1)) * 2048 + t * 128);
    uint32_t b_x[4], b_par[2];
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) b_x[ks] = (uint32_t)((ks ^ (g & 3)) * 32);
    b_par[0] = (uint32_t)((t >> 1) * 16);
    b_par[1] = (uint32_t)(((t >> 1) ^ 1) * 16);
    uint32_t bt_off[2][2];  // TB = 1: [ks&1][h]
#pragma unroll
    for (int p = 0; p < 2; ++p)
#pragma unroll
        for (int h = 0; h < 2; ++h) bt_off[p][h] = (uint32_t)(((((g & 1) ^ p) * 4) + ((2 * (g >> 2) + h) ^ t)) * 16);

    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    double a[2][8], b[2][4];
    auto load_frags = [&](int buf, uint32_t sb, int ks) {
        if (TA == 0) {
            const uint32_t ab = sb + a_thread + ks * 512;
#pragma unroll
            for (int q = 0; q < 4; ++q) lds128(a[buf][2 * q], a[buf][2 * q + 1], ab + a_off[ks & 1][q]);
        } else {
            const uint32_t ab = sb + a_thread + b_x[ks];
#pragma unroll
            for (int i = 0; i < 8; ++i) a[buf][i] = lds64(ab + (16 * (i >> 1) + (i & 1)) * 128 + b_par[i & 1]);
        }
        if (TB == 0) {
            const uint32_t bb = sb + b_thread + b_x[ks];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[buf][j] = lds64(bb + (16 * (j >> 1) + (j & 1)) * 128 + b_par[j & 1]);
        } else {
            const uint32_t bb = sb + b_thread + ks * 512;
#pragma unroll
            for (int h = 0; h < 2; ++h) lds128(b[buf][2 * h], b[buf][2 * h + 1], bb + bt_off[ks & 1][h]);
        }
    };

    if (KT > 0) {
        int stage = 0;
        uint32_t phase = 0;
        mbar_wait(&full[0], 0);
        load_frags(0, sbase0, 0);
        for (int kt = 0; kt < KT; ++kt) {
            const uint32_t sb = sbase0 + stage * T_STAGE_BYTES;
            int nstage = stage + 1;
            uint32_t nphase = phase;
            if (nstage == T_STAGES) {
                nstage = 0;
                nphase ^= 1;
            }
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                if (ks < 3) {
                    load_frags((ks + 1) & 1, sb, ks + 1);
                } else if (kt + 1 < KT) {
                    mbar_wait(&full[nstage], nphase);
                    load_frags(0, sbase0 + nstage * T_STAGE_BYTES, 0);
                }
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) dmma884p(acc[i][j][0], acc[i][j][1], a[ks & 1][i], b[ks & 1][j]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
            stage = nstage;
            phase = nphase;
        }
    }

    // Epilogue.  acc[i][j][e] is row row_of(i), column col_of(j, e) of C; the maps follow the fragment loaders above:
    //   TA = 0: row = 8g + i                      TA = 1: row = 16(i>>1) + 2g + (i&1)
    //   TB = 0: col = 16(j>>1) + 4t + 2e + (j&1)  TB = 1: col = 16(t&1) + 8e + 4(t>>1) + j
    // Either way fragments 2q and 2q+1 are two neighbouring rows, so a thread moves double2's: pair q sits at
    // row0 + 2q (TA = 0) or row0 + 16q (TA = 1).
    const int row0 = m0 + wm * 64 + (TA == 0 ? 8 * g : 2 * g);
    constexpr int QS = TA == 0 ? 1 : 8;  // distance between pairs, in double2's
    auto col_of = [&](int j, int e) {
        return n0 + wn * 32 + (TB == 0 ? 16 * (j >> 1) + 4 * t + 2 * e + (j & 1) : 16 * (t & 1) + 8 * e + 4 * (t >> 1) + j);
    };
    auto row_of = [&](int i) { return TA == 0 ? row0 + i : row0 + 16 * (i >> 1) + (i & 1); };
    const bool rows_full = (row_of(7) < M) && ((((uintptr_t)C) & 15) == 0) && !(ldc & 1) && !(TRI != 0 && bm == bn);
    if (rows_full && beta != 0.0) {
        // C += A B (the task-graph case).  The old values of FOUR columns (16 x 16 bytes per thread) are requested
        // before the first one is used: two L2 round trips per thread instead of eight -- a load cannot move above the
        // store of the previous column on its own (same array), and 8 x ~0.7 us was ~4 % of a 1024^3 tile task.
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            double2 old[4][4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int j = 2 * h + (c >> 1), e = c & 1;
                const int col = col_of(j, e);
                const double2* cp = reinterpret_cast<const double2*>(C + (size_t)row0 + (size_t)min(col, N - 1) * ldc);
#pragma unroll
                for (int q = 0; q < 4; ++q) old[c][q] = cp[q * QS];  // a column past N re-reads column N-1 and is dropped below
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int j = 2 * h + (c >> 1), e = c & 1;
                const int col = col_of(j, e);
                if (col >= N) continue;
                double2* cp = reinterpret_cast<double2*>(C + (size_t)row0 + (size_t)col * ldc);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    double2 v;
                    v.x = axpby(alpha, acc[2 * q][j][e], beta, old[c][q].x);
                    v.y = axpby(alpha, acc[2 * q + 1][j][e], beta, old[c][q].y);
                    cp[q * QS] = v;
                }
            }
        }
        return;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int col = col_of(j, e);
            if (col >= N) continue;
            if (rows_full) {
                double2* cp = reinterpret_cast<double2*>(C + (size_t)row0 + (size_t)col * ldc);
                if (beta == 0.0) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        double2 v;
                        v.x = __dmul_rn(alpha, acc[2 * q][j][e]);
                        v.y = __dmul_rn(alpha, acc[2 * q + 1][j][e]);
                        cp[q * QS] = v;
                    }
                } else {
                    double2 old[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) old[q] = cp[q * QS];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        double2 v;
                        v.x = axpby(alpha, acc[2 * q][j][e], beta, old[q].x);
                        v.y = axpby(alpha, acc[2 * q + 1][j][e], beta, old[q].y);
                        cp[q * QS] = v;
                    }
                }
            } else {
                // element by element (ragged rows, unaligned C, or the diagonal block of a triangular update); the old
                // values of the column are all requested before the first store, one L2 round trip instead of eight
                double oldv[8];
                bool ok[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int row = row_of(i);
                    ok[i] = row < M && !(TRI == 1 && row < col) && !(TRI == 2 && row > col);
                    oldv[i] = (ok[i] && beta != 0.0) ? C[(size_t)row + (size_t)col * ldc] : 0.0;
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    if (ok[i])
                        C[(size_t)row_of(i) + (size_t)col * ldc] =
                            beta != 0.0 ? axpby(alpha, acc[i][j][e], beta, oldv[i]) : __dmul_rn(alpha, acc[i][j][e]);
                }
            }
        }
    }
}

__global__ void __launch_bounds__(T_THREADS, 1)
dgemm_nn_tma(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, int M, int N, int K,
             double alpha, double beta, double* __restrict__ C, int ldc) {
    dgemm_tile_tma<0, 0>(&mapA, &mapB, M, N, K, alpha, beta, C, ldc, (int)blockIdx.x);
}

// The same tile code for transposed operands (N,T / T,N / T,T): what cublasDgemm_v2 accepts, at the same speed.
template <int TA, int TB>
__global__ void __launch_bounds__(T_THREADS, 1)
dgemm_tma_trans(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, int M, int N, int K,
                double alpha, double beta, double* __restrict__ C, int ldc) {
    dgemm_tile_tma<TA, TB>(&mapA, &mapB, M, N, K, alpha, beta, C, ldc, (int)blockIdx.x);
}

// C = alpha A A^T + beta C (TRANS = 0, A is n x k) or alpha A^T A + beta C (TRANS = 1, A is k x n) on one triangle of C.
template <int TRANS, int TRI>
__global__ void __launch_bounds__(T_THREADS, 1)
dsyrk_tma(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, int N, int K, double alpha,
          double beta, double* __restrict__ C, int ldc) {
    dgemm_tile_tma<TRANS, 1 - TRANS, TRI>(&mapA, &mapB, N, N, K, alpha, beta, C, ldc, (int)blockIdx.x);
}

// Grouped launch: the tiles of several independent GEMMs (the tile tasks a task runtime has ready together) in ONE
// grid.  A 256 x 256 fp64 tile task is four CTAs and about a microsecond of tensor-core time; launched one by one such
// tasks cost the host more than the GPU (6.5 us of manager time per task against 1 us of work).  Here CTA b finds its
// GEMM by tile_begin and runs the same tile code; tensor maps travel in the kernel parameter space.
constexpr int DG_MAX_GROUP = 64;
struct DgemmTask {
    CUtensorMap mapA, mapB;
    double* C;
    double alpha, beta;
    int M, N, K, ldc;
    int tile_begin, pad_;
};
template <int MAXT>
struct DgemmGroup {
    DgemmTask t[MAXT];
    int ntasks;
};
template <int MAXT>
__global__ void __launch_bounds__(T_THREADS, 1) dgemm_nn_tma_grouped(const __grid_constant__ DgemmGroup<MAXT> g) {
    int t = 0;
    const int b = (int)blockIdx.x;
#pragma unroll 1
    while (t + 1 < g.ntasks && b >= g.t[t + 1].tile_begin) ++t;
    const DgemmTask& k = g.t[t];
    dgemm_tile_tma<0, 0>(&k.mapA, &k.mapB, k.M, k.N, k.K, k.alpha, k.beta, k.C, k.ldc, b - k.tile_begin);
}

// ---------------------------------------------------------------------------------------------
// General kernel: any M, N, K, any transposes, any alignment.  CTA tile 64x64, BK = 16,
// 128 threads = 4 warps (2x2), warp tile 32x32 = 4x4 DMMA tiles, predicated scalar staging.
constexpr int GM = 64, GN = 64, GK = 16;
constexpr int LDGA = GM + 4;  // As[k][m]
constexpr int LDGB = GK + 4;  // Bs[n][k]

__global__ void __launch_bounds__(128)
dgemm_generic_64x64(int transa, int transb, int M, int N, int K, double alpha,
                    const double* __restrict__ A, int lda, const double* __restrict__ B, int ldb,
                    double beta, double* __restrict__ C, int ldc) {
    __shared__ double As[GK * LDGA];
    __shared__ double Bs[GN * LDGB];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp & 1, wn = warp >> 1;
    const int m0 = blockIdx.x * GM, n0 = blockIdx.y * GN;

    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    for (int k0 = 0; k0 < K; k0 += GK) {
        // stage A: 64x16 elements, 128 threads -> 8 each; m fastest across threads when !transa
        for (int idx = tid; idx < GM * GK; idx += 128) {
            int m, k;
            if (!transa) { m = idx % GM; k = idx / GM; } else { k = idx % GK; m = idx / GK; }
            int gm = m0 + m, gk = k0 + k;
            double v = 0.0;
            if (gm < M && gk < K) v = transa ? A[(size_t)gk + (size_t)gm * lda] : A[(size_t)gm + (size_t)gk * lda];
            As[k * LDGA + m] = v;
        }
        for (int idx = tid; idx < GN * GK; idx += 128) {
            int n, k;
            if (!transb) { k = idx % GK; n = idx / GK; } else { n = idx % GN; k = idx / GN; }
            int gn = n0 + n, gk = k0 + k;
            double v = 0.0;
            if (gn < N && gk < K) v = transb ? B[(size_t)gn + (size_t)gk * ldb] : B[(size_t)gk + (size_t)gn * ldb];
            Bs[n * LDGB + k] = v;
        }
        __syncthreads();
#pragma unroll
        for (int ks = 0; ks < GK / 4; ++ks) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[(ks * 4 + t) * LDGA + wm * 32 + 8 * i + g];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = Bs[(wn * 32 + 8 * j + g) * LDGB + ks * 4 + t];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                int row = m0 + wm * 32 + 8 * i + g;
                int col = n0 + wn * 32 + 8 * j + 2 * t + e;
                if (row < M && col < N) {
                    double* p = C + (size_t)row + (size_t)col * ldc;
                    double r = alpha * acc[i][j][e];
                    if (beta != 0.0) r += beta * *p;
                    *p = r;
                }
            }
}

static bool g_attr_set[64] = {false};

static int ensure_smem_attr() {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return PB2_ERR_CUDA;
    if (dev < 64 && g_attr_set[dev]) return PB2_SUCCESS;
    e = cudaFuncSetAttribute(dgemm_nn_128x128, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DGEMM_SMEM);
    if (e != cudaSuccess) return PB2_ERR_CUDA;
    e = cudaFuncSetAttribute(dgemm_nn_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dgemm_tma_trans<0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dgemm_tma_trans<1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dgemm_tma_trans<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dsyrk_tma<0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dsyrk_tma<0, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dsyrk_tma<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dsyrk_tma<1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_SMEM);
    if (e != cudaSuccess) return PB2_ERR_CUDA;
    e = cudaFuncSetAttribute(dgemm_nn_tma_grouped<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_SMEM);
    if (e != cudaSuccess) return PB2_ERR_CUDA;
    e = cudaFuncSetAttribute(dgemm_nn_tma_grouped<DG_MAX_GROUP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_SMEM);
    if (e != cudaSuccess) return PB2_ERR_CUDA;
    if (dev < 64) g_attr_set[dev] = true;
    return PB2_SUCCESS;
}

// 0 = TMA/mbarrier kernel (default), 1 = cp.async kernel (PB2_DGEMM_IMPL=cpasync; kept for A/B runs)
static int dgemm_impl() {
    static int impl = -1;
    if (impl < 0) {
        const char* e = getenv("PB2_DGEMM_IMPL");
        impl = (e && !strcmp(e, "cpasync")) ? 1 : 0;
    }
    return impl;
}

// Tensor maps depend only on (base, dims, ld); tile pointers recur (HBM cache slots), so keep a
// small direct-mapped cache per thread.
struct TmapKey {
    const void* p;
    int d0, d1, ld, kind;
};
struct TmapEntry {
    TmapKey k;
    CUtensorMap m;
    bool valid;
};
static int get_tmap(CUtensorMap* out, const void* p, int d0, int d1, int ld, int kind) {
    // 16 Ki entries: a 24 GiB working set of 8 MiB tiles is 3 Ki distinct (pointer, shape) keys
    static thread_local std::vector<TmapEntry> cache;
    if (cache.empty()) cache.resize(16384);
    uintptr_t h = ((uintptr_t)p >> 9) * 0x9E3779B97F4A7C15ull + (uintptr_t)(d0 * 31 + d1 * 17 + ld * 7 + kind);
    TmapEntry& e = cache[(h >> 20) & 16383];
    if (e.valid && e.k.p == p && e.k.d0 == d0 && e.k.d1 == d1 && e.k.ld == ld && e.k.kind == kind) {
        *out = e.m;
        return 0;
    }
    int rc = kind == 0 ? encode_tmap_2d(&e.m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, p, d0, d1, ld, 16, BK,
                                        CU_TENSOR_MAP_SWIZZLE_128B)
                       : encode_tmap_2d(&e.m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, p, d0, d1, ld, BK, BN,
                                        CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) {
        e.valid = false;
        return rc;
    }
    e.k = {p, d0, d1, ld, kind};
    e.valid = true;
    *out = e.m;
    return 0;
}

// ---- batches (parsec_b200_batch_begin / _end, gemm_tcgen05.cu): DGEMM calls recorded on one stream by one thread
struct DgemmBatch {
    bool open = false;
    void* stream = nullptr;
    int status = PB2_SUCCESS;
    std::vector<DgemmTask> tasks;
};
static thread_local DgemmBatch g_dbatch;

static int dgemm_batch_launch(std::vector<DgemmTask>& tasks, cudaStream_t st) {
    if (tasks.empty()) return PB2_SUCCESS;
    int total = 0;
    for (DgemmTask& t : tasks) {
        t.tile_begin = total;
        total += ((t.M + BM - 1) / BM) * ((t.N + BN - 1) / BN);
    }
    const int n = (int)tasks.size();
    if (n <= 8) {
        DgemmGroup<8> g;
        memcpy(g.t, tasks.data(), sizeof(DgemmTask) * (size_t)n);
        g.ntasks = n;
        dgemm_nn_tma_grouped<8><<<dim3(total), T_THREADS, T_SMEM, st>>>(g);
    } else {
        DgemmGroup<DG_MAX_GROUP> g;
        memcpy(g.t, tasks.data(), sizeof(DgemmTask) * (size_t)n);
        g.ntasks = n;
        dgemm_nn_tma_grouped<DG_MAX_GROUP><<<dim3(total), T_THREADS, T_SMEM, st>>>(g);
    }
    tasks.clear();
    return cudaGetLastError() == cudaSuccess ? PB2_SUCCESS : PB2_ERR_CUDA;
}

void dgemm_batch_open(void* stream) {
    g_dbatch.open = true;
    g_dbatch.stream = stream;
    g_dbatch.status = PB2_SUCCESS;
}
int dgemm_batch_close(void* stream) {
    if (!g_dbatch.open || g_dbatch.stream != stream) return PB2_SUCCESS;
    g_dbatch.open = false;
    int rc = dgemm_batch_launch(g_dbatch.tasks, (cudaStream_t)stream);
    return g_dbatch.status != PB2_SUCCESS ? g_dbatch.status : rc;
}

static inline int trans_of(char c) {
    if (c == 'N' || c == 'n') return 0;
    if (c == 'T' || c == 't' || c == 'C' || c == 'c') return 1;
    return -1;
}

}  // namespace pb2

extern "C" int parsec_b200_dgemm_sm100(char transa, char transb, int m, int n, int k, double alpha,
                                        const double* A, int lda, const double* B, int ldb, double beta,
                                        double* C, int ldc, void* stream) {
    using namespace pb2;
    const int ta = trans_of(transa), tb = trans_of(transb);
    if (ta < 0 || tb < 0 || m < 0 || n < 0 || k < 0) return PB2_ERR_BAD_PARAM;
    if (lda < (ta ? (k > 1 ? k : 1) : (m > 1 ? m : 1)) || ldb < (tb ? (n > 1 ? n : 1) : (k > 1 ? k : 1)) ||
        ldc < (m > 1 ? m : 1))
        return PB2_ERR_BAD_PARAM;
    if (m == 0 || n == 0) return PB2_SUCCESS;
    cudaStream_t st = (cudaStream_t)stream;
    const bool aligned = (((uintptr_t)A | (uintptr_t)B | (uintptr_t)C) & 15) == 0 && !(lda & 1) && !(ldb & 1) &&
                         !(ldc & 1);
    const bool batching = g_dbatch.open && g_dbatch.stream == stream;
    if (aligned && k > 0 && dgemm_impl() == 0) {
        int rc = ensure_smem_attr();
        if (rc != PB2_SUCCESS) return rc;
        // op(A) is m x k: stored m x k (m contiguous, boxes [16 m][16 k]) or, transposed, k x m (k contiguous, one box
        // [16 k][128 m]); op(B) is k x n: stored k x n (k contiguous) or, transposed, n x k (n contiguous)
        CUtensorMap ma, mb;
        if ((ta ? get_tmap(&ma, A, k, m, lda, 1) : get_tmap(&ma, A, m, k, lda, 0)) ||
            (tb ? get_tmap(&mb, B, n, k, ldb, 0) : get_tmap(&mb, B, k, n, ldb, 1)))
            return PB2_ERR_CUDA;
        if (batching && !ta && !tb) {  // recorded; launched with the rest of the batch
            DgemmTask t;
            memset(&t, 0, sizeof(t));
            t.mapA = ma; t.mapB = mb; t.C = C; t.alpha = alpha; t.beta = beta; t.M = m; t.N = n; t.K = k; t.ldc = ldc;
            g_dbatch.tasks.push_back(t);
            if ((int)g_dbatch.tasks.size() == DG_MAX_GROUP) {
                rc = dgemm_batch_launch(g_dbatch.tasks, st);
                if (rc != PB2_SUCCESS) g_dbatch.status = rc;
                return rc;
            }
            return PB2_SUCCESS;
        }
        if (batching) {  // not a recordable call: what was recorded so far goes first, stream order is call order
            rc = dgemm_batch_launch(g_dbatch.tasks, st);
            if (rc != PB2_SUCCESS) return g_dbatch.status = rc;
        }
        dim3 grid(((m + BM - 1) / BM) * ((n + BN - 1) / BN));
        if (!ta && !tb) dgemm_nn_tma<<<grid, T_THREADS, T_SMEM, st>>>(ma, mb, m, n, k, alpha, beta, C, ldc);
        else if (!ta && tb) dgemm_tma_trans<0, 1><<<grid, T_THREADS, T_SMEM, st>>>(ma, mb, m, n, k, alpha, beta, C, ldc);
        else if (ta && !tb) dgemm_tma_trans<1, 0><<<grid, T_THREADS, T_SMEM, st>>>(ma, mb, m, n, k, alpha, beta, C, ldc);
        else dgemm_tma_trans<1, 1><<<grid, T_THREADS, T_SMEM, st>>>(ma, mb, m, n, k, alpha, beta, C, ldc);
        return cudaGetLastError() == cudaSuccess ? PB2_SUCCESS : PB2_ERR_CUDA;
    }
    if (batching) {
        int rc = dgemm_batch_launch(g_dbatch.tasks, st);
        if (rc != PB2_SUCCESS) return g_dbatch.status = rc;
    }
    if (!ta && !tb && aligned && k > 0 && m % BM == 0 && n % BN == 0 && k % BK == 0) {
        int rc = ensure_smem_attr();
        if (rc != PB2_SUCCESS) return rc;
        dim3 grid((m / BM) * (n / BN));
        dgemm_nn_128x128<<<grid, DGEMM_THREADS, DGEMM_SMEM, st>>>(m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);
    } else {
        dim3 grid((m + GM - 1) / GM, (n + GN - 1) / GN);
        dgemm_generic_64x64<<<grid, 128, 0, st>>>(ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);
    }
    return cudaGetLastError() == cudaSuccess ? PB2_SUCCESS : PB2_ERR_CUDA;
}

// Symmetric rank-k update on one triangle of C, for the SYRK / HERK bodies of a tiled Cholesky:
//   trans = 'N': C = alpha A A^T + beta C, A is n x k;  trans = 'T': C = alpha A^T A + beta C, A is k x n.
// The other triangle of C is neither read nor written.
__global__ void dsyrk_generic(int lower, int trans, int N, int K, double alpha, const double* __restrict__ A, int lda,
                              double beta, double* __restrict__ C, int ldc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= N || j >= N || (lower ? i < j : i > j)) return;
    double acc = 0.0;
    for (int l = 0; l < K; ++l)
        acc += trans ? A[(size_t)l + (size_t)i * lda] * A[(size_t)l + (size_t)j * lda]
                     : A[(size_t)i + (size_t)l * lda] * A[(size_t)j + (size_t)l * lda];
    double* c = C + (size_t)i + (size_t)j * ldc;
    *c = beta != 0.0 ? pb2::axpby(alpha, acc, beta, *c) : __dmul_rn(alpha, acc);
}

extern "C" int parsec_b200_dsyrk_sm100(char uplo, char trans, int n, int k, double alpha, const double* A, int lda,
                                        double beta, double* C, int ldc, void* stream) {
    using namespace pb2;
    const int tr = trans_of(trans);
    const int lower = (uplo == 'L' || uplo == 'l') ? 1 : (uplo == 'U' || uplo == 'u') ? 0 : -1;
    if (tr < 0 || lower < 0 || n < 0 || k < 0) return PB2_ERR_BAD_PARAM;
    if (lda < (tr ? (k > 1 ? k : 1) : (n > 1 ? n : 1)) || ldc < (n > 1 ? n : 1)) return PB2_ERR_BAD_PARAM;
    if (n == 0) return PB2_SUCCESS;
    cudaStream_t st = (cudaStream_t)stream;
    if (g_dbatch.open && g_dbatch.stream == stream) {  // stream order is call order
        int rc = dgemm_batch_launch(g_dbatch.tasks, st);
        if (rc != PB2_SUCCESS) return g_dbatch.status = rc;
    }
    const bool aligned = (((uintptr_t)A | (uintptr_t)C) & 15) == 0 && !(lda & 1) && !(ldc & 1);
    if (aligned && k > 0) {
        int rc = ensure_smem_attr();
        if (rc != PB2_SUCCESS) return rc;
        CUtensorMap ma, mb;  // the two roles of A: op(A) and op(A)^T
        if (tr ? (get_tmap(&ma, A, k, n, lda, 1) || get_tmap(&mb, A, k, n, lda, 1))
               : (get_tmap(&ma, A, n, k, lda, 0) || get_tmap(&mb, A, n, k, lda, 0)))
            return PB2_ERR_CUDA;
        const int tiles = (n + BM - 1) / BM;
        dim3 grid(tiles * tiles);
        if (!tr && lower) dsyrk_tma<0, 1><<<grid, T_THREADS, T_SMEM, st>>>(ma, mb, n, k, alpha, beta, C, ldc);
        else if (!tr) dsyrk_tma<0, 2><<<grid, T_THREADS, T_SMEM, st>>>(ma, mb, n, k, alpha, beta, C, ldc);
        else if (lower) dsyrk_tma<1, 1><<<grid, T_THREADS, T_SMEM, st>>>(ma, mb, n, k, alpha, beta, C, ldc);
        else dsyrk_tma<1, 2><<<grid, T_THREADS, T_SMEM, st>>>(ma, mb, n, k, alpha, beta, C, ldc);
    } else {
        dim3 block(16, 16), grid((n + 15) / 16, (n + 15) / 16);
        dsyrk_generic<<<grid, block, 0, st>>>(lower, tr, n, k, alpha, A, lda, beta, C, ldc);
    }
    return cudaGetLastError() == cudaSuccess ? PB2_SUCCESS : PB2_ERR_CUDA;
}
extern "C" int parsec_b200_dsyrk(char uplo, char trans, int n, int k, double alpha, const double* A, int lda, double beta,
                                  double* C, int ldc, void* stream) {
    return parsec_b200_dsyrk_sm100(uplo, trans, n, k, alpha, A, lda, beta, C, ldc, stream);
}

// dyld= lookup tries "<fn>_sm100" first and then "<fn>" (device_cuda_module.c:175-240).
extern "C" int parsec_b200_dgemm(char transa, char transb, int m, int n, int k, double alpha, const double* A,
                                  int lda, const double* B, int ldb, double beta, double* C, int ldc, void* stream) {
    return parsec_b200_dgemm_sm100(transa, transb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, stream);
}
