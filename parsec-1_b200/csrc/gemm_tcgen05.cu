// BF16-in / FP32-accumulate tile GEMM on the 5th-generation tensor cores: tcgen05.mma with the accumulator in
// TMEM, operands staged by TMA (128-byte swizzle), warp-specialised roles, mbarrier pipelines.
//
//   C (m x n, fp32, column-major) = alpha * A (m x k, bf16, column-major) * B (k x n, bf16, column-major) + beta * C
//
// This is the BASELINE config 5 body (NB = 2048 tiles): the reference has no bf16 body at all (its
// parsec_matrix_type_t has no 16-bit float, matrix.h:25-32); the call shape follows its cublasDgemm_v2 bodies.
//
// Operand orientation, no transposes anywhere:
//   A is column-major, i.e. M is contiguous  -> UMMA "MN-major" operand.  One stage holds two TMA boxes
//        [64 k][64 m] (128-byte rows of 64 bf16 along m): exactly the canonical SW128 MN-major layout
//        ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units, LBO = 8 KiB (next 64 rows of m), SBO = 1 KiB (next 8 k).
//   B is column-major k x n, i.e. K is contiguous -> UMMA "K-major" operand.  One TMA box [256 n][64 k]
//        (128-byte rows of 64 bf16 along k): canonical SW128 K-major layout, SBO = 1 KiB (next 8 rows of n).
//   D lives in TMEM: 128 lanes (rows of the tile) x 256 fp32 columns.  The epilogue reads 32 columns per
//        tcgen05.ld (one row per thread), so for a fixed column a warp holds 32 consecutive rows: column-major
//        stores to C are 128-byte coalesced.
//
// Roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (one elected lane),
// warps 2-5 = epilogue (warp w may only touch TMEM lanes 32*(w%4) .. +31, and 2,3,4,5 cover all four quarters).
//
// Two kernels share these pieces: gemm_bf16_tcgen05 (one CTA per 128 x 256 tile, cta_group::1) and the default
// gemm_umma_pair (a cluster of two CTAs per 256 x 256 tile, cta_group::2, see below).  Both finish through
// umma_epilogue_tma (C moved by TMA boxes through the idle operand ring) or, for a C that TMA cannot address,
// through the register epilogue.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>

#include "pb2_internal.h"
#include "tma_utils.cuh"

namespace pb2 {

constexpr int G_BM = 128, G_BN = 256;
// ring depth: 4 stages with one CTA per SM, or 2 stages with two CTAs per SM (2 x 256 TMEM columns), where one CTA's
// epilogue (the C tile streams in and out once per task) runs under the other CTA's main loop
constexpr int G_STAGES_1CTA = 4, G_STAGES_2CTA = 2;
constexpr int G_A_BYTES = G_BM * 128;  // 16 KiB: 128 rows of m by 128 bytes of k, as boxes [k][one 128-byte row of m]
constexpr int G_B_BYTES = G_BN * 128;  // 32 KiB: one box [256 n][128 bytes of k]
constexpr int G_STAGE_BYTES = G_A_BYTES + G_B_BYTES;
constexpr int G_THREADS = 192;
constexpr int G_TMEM_COLS = 256;
constexpr size_t gemm_smem(int stages) { return (size_t)stages * G_STAGE_BYTES + 1024 + 512; }

// instruction descriptor (cute/arch/mma_sm100_desc.hpp InstrDescriptor), built per element type in the kernel:
//   [4,6) c_format = 1 (F32); [7,10) a_format, [10,13) b_format = 1 (BF16) or 2 (TF32);
//   [15] a_major = 1 (MN-major); [16] b_major = 0 (K-major); [17,23) N >> 3; [24,29) M >> 4

// shared-memory matrix descriptor (SmemDescriptor): [0,14) start >> 4; [16,30) LBO >> 4; [32,46) SBO >> 4;
// [46,48) version = 1; [61,64) layout type = 2 (SWIZZLE_128B)
// layout: 2 = SWIZZLE_128B (16-byte chunks XOR row, 8-row atom), 1 = SWIZZLE_128B_BASE32B (32-byte chunks XOR row,
// 4-row atom) -- the only layout the hardware accepts for an MN-major operand with 32-bit elements (TF32)
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout = 2) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)layout << 61;
    return d;
}

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(smem_dst)), "r"(ncols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::);
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(ncols));
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrives on the mbarrier once every MMA issued so far by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(smem_u32(bar))
                 : "memory");
}
// 32 lanes x 32 columns of 32-bit: thread t of the warp receives columns [c, c+32) of lane (base lane + t)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
}

// -DPB2_UMMA_TRACE (tools/umma_trace.cu only): per-CTA globaltimer marks of gemm_umma_pair -- 0 entry, 1 set-up done
// (barriers, TMEM, cluster sync), 2 accumulator complete, 3 epilogue done -- plus the SM id; the slot index lives in the
// word after the epilogue barriers.  Compiled out of the library.
#ifdef PB2_UMMA_TRACE
__device__ unsigned long long pb2_trace_buf[5 * 65536];
__device__ unsigned int pb2_trace_n;
__device__ __forceinline__ unsigned long long pb2_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
    return t;
}
#define PB2_TRACE_MARK(cbar_, which_, cond_)                                                       \
    do {                                                                                           \
        if (cond_) pb2_trace_buf[5 * ((cbar_)[4 * E_NBUF] & 65535) + (which_)] = pb2_now();        \
    } while (0)
#else
#define PB2_TRACE_MARK(cbar_, which_, cond_) \
    do {                                     \
    } while (0)
#endif

// epilogue of one warp: TMEM -> registers -> global (column-major, 128-byte coalesced per column); `quarter` selects
// the 32 TMEM lanes (rows of the tile) this warp may touch
__device__ __forceinline__ void umma_epilogue(uint32_t tmem_base, int quarter, int lane, int m0, int n0, int M, int N, int KT,
                                              float alpha, float beta, float* __restrict__ C, int ldc, uint64_t* tmem_full) {
    const int row = m0 + quarter * 32 + lane;
    if (beta != 0.0f && m0 + quarter * 32 < M) {
        // pull this warp's 32-row strip of the old C tile towards L2 while the main loop runs: for one column the
        // strip is one 128-byte line; lane j takes columns j, j+32, ...
        const float* strip = C + (size_t)(m0 + quarter * 32);
        for (int c = lane; c < G_BN && n0 + c < N; c += 32)
            asm volatile("prefetch.global.L2 [%0];\n" ::"l"(strip + (size_t)(n0 + c) * ldc));
    }
    if (KT > 0) {
        mbar_wait(tmem_full, 0);
        tc_fence_after();
    }
#pragma unroll 1
    for (int c0 = 0; c0 < G_BN; c0 += 32) {
        if (n0 + c0 >= N) break;  // warp-uniform
        uint32_t v[32];
        if (KT > 0) {
            tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c0, v);
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
        if (row < M) {
            float* base = C + (size_t)row + (size_t)(n0 + c0) * ldc;
            const int ncols = min(32, N - (n0 + c0));
            if (beta != 0.0f) {
                // all 32 loads of the chunk are issued before the first use: one DRAM round trip per chunk
                // instead of one per column
                float old[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) old[j] = (j < ncols) ? __ldcs(base + (size_t)j * ldc) : 0.0f;
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (j < ncols) base[(size_t)j * ldc] = alpha * __uint_as_float(v[j]) + beta * old[j];
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (j < ncols) base[(size_t)j * ldc] = alpha * __uint_as_float(v[j]);
            }
        }
    }
}

// TMA-staged epilogue (used when C is 16-byte aligned with ldc % 4 == 0).  Once the accumulator is complete the operand
// ring is idle, so each epilogue warp turns its quarter of it into E_NBUF buffers of 32 rows x 32 columns of fp32
// (4 KiB, one TMA box of C).  beta != 0: the boxes of the old C strip are requested up front, E_NBUF in flight, instead
// of one dependent global-load round trip per 32 columns.  The result overwrites the box in place (each lane owns its
// row: bank = lane) and leaves through a TMA store, which also clips ragged edges; the warp never waits on a store
// except to recycle a buffer for columns >= 32 * E_NBUF.
constexpr int E_NBUF = 6, E_BOX_BYTES = 32 * 32 * 4;
constexpr int E_RING_BYTES = 4 * E_NBUF * E_BOX_BYTES;  // 96 KiB: every variant's operand ring is at least this large

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];\n" ::"l"(map),
                 "r"(smem_u32(src)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;\n" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory"); }

// ring: the CTA's operand ring (1024-byte aligned); cbar: 4 * E_NBUF mbarriers initialised to count 1
__device__ __forceinline__ void umma_epilogue_tma(uint32_t tmem_base, int quarter, int lane, int m0, int n0, int M, int N,
                                                  int KT, float alpha, float beta, const CUtensorMap* mapC, uint8_t* ring,
                                                  uint64_t* cbar, uint64_t* tmem_full, const float* __restrict__ C, int ldc) {
    const int row0 = m0 + quarter * 32;
    const bool live = row0 < M && n0 < N;  // warp-uniform
    const int nchunks = live ? (min(G_BN, N - n0) + 31) / 32 : 0;
    const bool readc = beta != 0.0f;
    if (live && readc) {
        const float* strip = C + (size_t)row0;
        for (int c = lane; c < G_BN && n0 + c < N; c += 32)
            asm volatile("prefetch.global.L2 [%0];\n" ::"l"(strip + (size_t)(n0 + c) * ldc));
    }
    if (KT > 0) {
        mbar_wait(tmem_full, 0);  // every MMA has completed: the operand ring is no longer read
        tc_fence_after();
    }
    PB2_TRACE_MARK(cbar, 2, quarter == 0 && lane == 0);
    if (!live) return;
    float* bufs = reinterpret_cast<float*>(ring + quarter * (E_NBUF * E_BOX_BYTES));
    uint64_t* bar = cbar + quarter * E_NBUF;
    if (readc && lane == 0) {
        tma_prefetch_desc(mapC);
        for (int c = 0; c < nchunks && c < E_NBUF; ++c) {
            mbar_arrive_expect_tx(&bar[c], E_BOX_BYTES);
            tma_load_2d(bufs + c * (E_BOX_BYTES / 4), mapC, row0, n0 + 32 * c, &bar[c]);
        }
    }
#pragma unroll 1
    for (int c = 0; c < nchunks; ++c) {
        const int b = c % E_NBUF;
        float* buf = bufs + b * (E_BOX_BYTES / 4);
        uint32_t v[32];
        if (KT > 0) {
            tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(32 * c), v);
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
        if (readc) {
            mbar_wait(&bar[b], (uint32_t)(c / E_NBUF) & 1u);
#pragma unroll
            for (int j = 0; j < 32; ++j) buf[j * 32 + lane] = alpha * __uint_as_float(v[j]) + beta * buf[j * 32 + lane];
        } else {
            if (c >= E_NBUF) {  // the store that last read this buffer must be done with it
                if (lane == 0) tma_store_wait_read<E_NBUF - 1>();
                __syncwarp();
            }
#pragma unroll
            for (int j = 0; j < 32; ++j) buf[j * 32 + lane] = alpha * __uint_as_float(v[j]);
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
            tma_store_2d(mapC, buf, row0, n0 + 32 * c);
            tma_store_commit();
            if (readc && c + E_NBUF < nchunks) {  // recycle this buffer for the box E_NBUF columns-chunks ahead
                tma_store_wait_read<0>();
                mbar_arrive_expect_tx(&bar[b], E_BOX_BYTES);
                tma_load_2d(buf, mapC, row0, n0 + 32 * (c + E_NBUF), &bar[b]);
            }
        }
    }
    if (lane == 0) tma_store_wait_all();
    __syncwarp();
}

// EB = operand element size: 2 = bf16 (kind::f16), 4 = fp32 consumed as TF32 (kind::tf32).  A stage is 128 bytes deep
// along K for B (64 bf16 or 32 fp32) and always 48 KiB; an UMMA consumes 32 bytes of K (16 bf16 / 8 tf32).
template <int G_STAGES, int EB>
__global__ void __launch_bounds__(G_THREADS, G_STAGES <= 2 ? 2 : 1)
gemm_bf16_tcgen05(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, int M, int N, int K,
                  float alpha, float beta, float* __restrict__ C, int ldc, const __grid_constant__ CUtensorMap mapC, int c_tma) {
    constexpr int BKE = 128 / EB;             // K elements per stage
    constexpr int MBOX = 128 / EB;            // m elements per A box (one 128-byte row)
    constexpr int NBOX = G_BM / MBOX;         // A boxes per stage
    constexpr int BOX_BYTES = BKE * 128;      // one A box: BKE k-rows of 128 bytes
    constexpr int UK = 32 / EB;               // K elements per UMMA
    constexpr uint32_t IDESC = (1u << 4) | ((EB == 2 ? 1u : 2u) << 7) | ((EB == 2 ? 1u : 2u) << 10) | (1u << 15) |
                               ((uint32_t)(G_BN >> 3) << 17) | ((uint32_t)(G_BM >> 4) << 24);
    static_assert(NBOX * BOX_BYTES == G_A_BYTES && G_BN * 128 == G_B_BYTES, "stage layout");
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t sbase0 = (raw + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (sbase0 - raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + G_STAGES * G_STAGE_BYTES);
    uint64_t* empty = full + G_STAGES;
    uint64_t* tmem_full = empty + G_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
    uint64_t* cbar = tmem_full + 2;  // 4 warps x E_NBUF barriers of the TMA-staged epilogue

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_m = (M + G_BM - 1) / G_BM;
    const int m0 = (blockIdx.x % tiles_m) * G_BM, n0 = (blockIdx.x / tiles_m) * G_BN;
    const int KT = (K + BKE - 1) / BKE;

    if (threadIdx.x == 0) {
        for (int s = 0; s < G_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        mbar_init(tmem_full, 1);
        for (int i = 0; i < 4 * E_NBUF; ++i) mbar_init(&cbar[i], 1);
        fence_barrier_init();
        fence_proxy_async();
    }
    if (warp == 1) tmem_alloc(tmem_slot, G_TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0 && KT > 0) {
            tma_prefetch_desc(&mapA);
            tma_prefetch_desc(&mapB);
            int stage = 0;
            uint32_t phase = 0;
            for (int kt = 0; kt < KT; ++kt) {
                if (kt >= G_STAGES) mbar_wait(&empty[stage], phase ^ 1);
                uint8_t* sa = smem + stage * G_STAGE_BYTES;
                mbar_arrive_expect_tx(&full[stage], G_STAGE_BYTES);
#pragma unroll
                for (int b = 0; b < NBOX; ++b)
                    tma_load_2d(sa + b * BOX_BYTES, &mapA, m0 + b * MBOX, kt * BKE, &full[stage]);
                tma_load_2d(sa + G_A_BYTES, &mapB, kt * BKE, n0, &full[stage]);
                if (++stage == G_STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int kt = 0; kt < KT; ++kt) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                const uint32_t sa = sbase0 + stage * G_STAGE_BYTES;
                const uint32_t sb = sa + G_A_BYTES;
#pragma unroll
                for (int ku = 0; ku < BKE / UK; ++ku) {
                    // A: UK k-rows of 128 B further per step (LBO = one box: the next 128 bytes of m);
                    // B: 32 bytes further inside the 128-byte swizzle row
                    const uint64_t adesc = EB == 2 ? umma_desc(sa + ku * UK * 128, BOX_BYTES, 1024, 2)
                                                   : umma_desc(sa + ku * UK * 128, BOX_BYTES, 512, 1);
                    const uint64_t bdesc = umma_desc(sb + ku * 32, 16, 1024);
                    if (EB == 2) umma_bf16(tmem_base, adesc, bdesc, IDESC, (kt | ku) != 0);
                    else umma_tf32(tmem_base, adesc, bdesc, IDESC, (kt | ku) != 0);
                }
                umma_commit(&empty[stage]);  // the stage may be refilled once these MMAs have read it
                if (++stage == G_STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            umma_commit(tmem_full);  // accumulator complete
        }
    } else {
        if (c_tma)
            umma_epilogue_tma(tmem_base, warp & 3, lane, m0, n0, M, N, KT, alpha, beta, &mapC, smem, cbar, tmem_full, C, ldc);
        else
            umma_epilogue(tmem_base, warp & 3, lane, m0, n0, M, N, KT, alpha, beta, C, ldc, tmem_full);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, G_TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// CTA-pair variant: a cluster of two CTAs (the two SMs of a TPC) computes one 256 x 256 tile with
// tcgen05.mma.cta_group::2 (M = 256, N = 256).  CTA r of the pair stages rows [128 r, 128 r + 128) of the A tile and
// rows [128 r, 128 r + 128) of the (n-major) B tile; the tensor cores of both SMs read both halves of B, so each SM
// pulls 32 KiB per 64 (bf16) k instead of 48 KiB from L2 and reads 8 KiB instead of 12 KiB of shared memory per UMMA.
// Its accumulator half (its 128 rows x 256 columns) sits in its own TMEM.
//   * both producers signal the LEADER's full barrier (TMA .cta_group::2 with the peer bit of the barrier address
//     cleared); the leader's producer posts the expected bytes of both CTAs;
//   * the leader's elected lane issues the MMAs and commits with .multicast::cluster to the empty / tmem_full barriers
//     of both CTAs (same offsets in both);
//   * TMEM is allocated and freed with .cta_group::2 by warp 1 of both CTAs, cluster barrier before the free.
constexpr int P_STAGES_2CTA = 3, P_STAGES_1CTA = 6;
#ifndef PB2_UMMA_DEFAULT_MODE
#define PB2_UMMA_DEFAULT_MODE 3  // see launch_umma_gemm: 2 = two independent CTAs per SM, 3 = CTA pairs
#endif
constexpr int P_BH_BYTES = 128 * 128;  // this CTA's half of B: box [128 n][128 bytes of k]
constexpr int P_STAGE_BYTES = G_A_BYTES + P_BH_BYTES;
constexpr size_t pair_smem(int stages) { return (size_t)stages * P_STAGE_BYTES + 1024 + 512; }
static_assert(P_STAGES_2CTA * P_STAGE_BYTES >= E_RING_BYTES && G_STAGES_2CTA * G_STAGE_BYTES >= E_RING_BYTES,
              "the TMA-staged epilogue borrows 96 KiB of the operand ring");

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
// TMA load issued by either CTA of the pair; completion bytes go to the barrier at `bar_cluster_addr` (a
// shared::cluster address: the leader's barrier when the peer bit is cleared)
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, int c0, int c1, uint32_t bar_cluster_addr) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
        : "memory");
}
template <int EB>
__device__ __forceinline__ void umma_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    if (EB == 2)
        asm volatile(
            "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
            "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d),
            "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
            : "memory");
    else
        asm volatile(
            "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
            "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d),
            "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
            : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    const uint16_t mask = 3;
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n" ::"r"(
                     smem_u32(bar)),
                 "h"(mask)
                 : "memory");
}

template <int STAGES, int EB>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(G_THREADS, STAGES <= 3 ? 2 : 1)
gemm_umma_pair(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapBh, int M, int N, int K,
               float alpha, float beta, float* __restrict__ C, int ldc, const __grid_constant__ CUtensorMap mapC, int c_tma) {
    constexpr int BKE = 128 / EB;
    constexpr int MBOX = 128 / EB;
    constexpr int NBOX = G_BM / MBOX;
    constexpr int BOX_BYTES = BKE * 128;
    constexpr int UK = 32 / EB;
    constexpr uint32_t IDESC = (1u << 4) | ((EB == 2 ? 1u : 2u) << 7) | ((EB == 2 ? 1u : 2u) << 10) | (1u << 15) |
                               ((uint32_t)(G_BN >> 3) << 17) | ((uint32_t)((2 * G_BM) >> 4) << 24);
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t sbase0 = (raw + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (sbase0 - raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * P_STAGE_BYTES);
    uint64_t* empty = full + STAGES;
    uint64_t* tmem_full = empty + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
    uint64_t* cbar = tmem_full + 2;  // 4 warps x E_NBUF barriers of the TMA-staged epilogue

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int pair = blockIdx.x >> 1;
    const int pairs_m = (M + 2 * G_BM - 1) / (2 * G_BM);
    const int m0 = (pair % pairs_m) * (2 * G_BM) + rank * G_BM;
    const int n0 = (pair / pairs_m) * G_BN;
    const int KT = (K + BKE - 1) / BKE;
#ifdef PB2_UMMA_TRACE
    if (threadIdx.x == 128) {
        const unsigned slot = atomicAdd(&pb2_trace_n, 1u);
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;\n" : "=r"(smid));
        cbar[4 * E_NBUF] = slot;
        pb2_trace_buf[5 * (slot & 65535) + 0] = pb2_now();
        pb2_trace_buf[5 * (slot & 65535) + 4] = ((unsigned long long)blockIdx.x << 32) | smid;
    }
#endif

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);   // only the leader's is used: its producer's arrive + the bytes of both CTAs
            mbar_init(&empty[s], 1);  // one multicast commit per refill
        }
        mbar_init(tmem_full, 1);
        for (int i = 0; i < 4 * E_NBUF; ++i) mbar_init(&cbar[i], 1);
        fence_barrier_init();
        fence_proxy_async();
    }
    if (warp == 1) {
#ifdef PB2_UMMA_TRACE
        const unsigned long long ta = pb2_now();
#endif
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)G_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::);
#ifdef PB2_UMMA_TRACE
        if (lane == 0) cbar[4 * E_NBUF + 1] = pb2_now() - ta;  // how long the TMEM allocation took
#endif
    }
    __syncwarp();
    tc_fence_before();
    cluster_sync_all();  // barriers of both CTAs initialised, TMEM of both allocated
    tc_fence_after();
    PB2_TRACE_MARK(cbar, 1, threadIdx.x == 128);
#ifdef PB2_UMMA_TRACE
    if (threadIdx.x == 128) pb2_trace_buf[5 * (cbar[4 * E_NBUF] & 65535) + 4] |= (cbar[4 * E_NBUF + 1] & 0xFFFFF) << 44;
#endif
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0 && KT > 0) {
            tma_prefetch_desc(&mapA);
            tma_prefetch_desc(&mapBh);
            int stage = 0;
            uint32_t phase = 0;
            for (int kt = 0; kt < KT; ++kt) {
                if (kt >= STAGES) mbar_wait(&empty[stage], phase ^ 1);
                uint8_t* sa = smem + stage * P_STAGE_BYTES;
                if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * P_STAGE_BYTES);
                const uint32_t lead_full = smem_u32(&full[stage]) & 0xFEFFFFFFu;
#pragma unroll
                for (int b = 0; b < NBOX; ++b)
                    tma_load_2d_pair(sa + b * BOX_BYTES, &mapA, m0 + b * MBOX, kt * BKE, lead_full);
                tma_load_2d_pair(sa + G_A_BYTES, &mapBh, kt * BKE, n0 + rank * (G_BN / 2), lead_full);
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int kt = 0; kt < KT; ++kt) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                const uint32_t sa = sbase0 + stage * P_STAGE_BYTES;
                const uint32_t sb = sa + G_A_BYTES;
#pragma unroll
                for (int ku = 0; ku < BKE / UK; ++ku) {
                    const uint64_t adesc = EB == 2 ? umma_desc(sa + ku * UK * 128, BOX_BYTES, 1024, 2)
                                                   : umma_desc(sa + ku * UK * 128, BOX_BYTES, 512, 1);
                    const uint64_t bdesc = umma_desc(sb + ku * 32, 16, 1024);
                    umma_pair<EB>(tmem_base, adesc, bdesc, IDESC, (kt | ku) != 0);
                }
                umma_commit_pair(&empty[stage]);
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            umma_commit_pair(tmem_full);
        }
    } else {
        if (c_tma)
            umma_epilogue_tma(tmem_base, warp & 3, lane, m0, n0, M, N, KT, alpha, beta, &mapC, smem, cbar, tmem_full, C, ldc);
        else
            umma_epilogue(tmem_base, warp & 3, lane, m0, n0, M, N, KT, alpha, beta, C, ldc, tmem_full);
        PB2_TRACE_MARK(cbar, 3, warp == 4 && lane == 0);
    }
    __syncwarp();
    tc_fence_before();
    cluster_sync_all();  // the peer's MMAs read this CTA's shared memory and TMEM is freed as a pair
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"((uint32_t)G_TMEM_COLS));
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Grouped persistent variant: ONE launch runs a group of independent GEMMs (the tile tasks the device module has ready
// at the same time), one CTA pair per TPC, each pair walking the 256 x 256 output tiles of the whole group.
//
// Why: a 2048^3 tile task is 64 output tiles for 74 CTA pairs, and with one launch per task the C tile of every task
// streams in and out while its CTAs' tensor cores idle (7 us of set-up + epilogue per 26 us CTA lifetime,
// profiles/umma_trace_r1.md).  Here
//   * the accumulator is double buffered in TMEM (2 x 256 columns = the whole TMEM of the SM): the MMA warp starts the
//     next tile while the four epilogue warps drain the previous one, whatever task either belongs to;
//   * the epilogue has its own shared-memory boxes (the operand ring is never idle any more), the old C boxes of the
//     next tile are requested while its MMAs run;
//   * set-up (barriers, TMEM allocation, cluster rendez-vous) is paid once per launch, not once per tile;
//   * tiles are walked in a grouped raster (8 tile rows at a time), so that the A and B panels the pairs touch together
//     fit in L2 for a large GEMM (the column-major walk of the per-tile kernel re-read them ~3x from DRAM at 8192^3).
// Tensor maps travel in the kernel parameter space (one __grid_constant__ struct), like the per-tile kernels' do.
#ifndef PB2_Q_STAGES
#define PB2_Q_STAGES 6
#endif
#ifndef PB2_Q_EBUF
#define PB2_Q_EBUF 2
#endif
constexpr int Q_STAGES = PB2_Q_STAGES;  // operand ring: 6 x 32 KiB (5 stages + 3 C boxes measured 2 % slower)
constexpr int Q_EBUF = PB2_Q_EBUF;      // C boxes (32 x 32 fp32) per epilogue warp
constexpr int Q_RASTER = 8;          // tile rows per raster group
constexpr int Q_TMEM_COLS = 512;     // two accumulator stages
constexpr size_t grouped_smem() {
    return (size_t)Q_STAGES * P_STAGE_BYTES + 4 * Q_EBUF * E_BOX_BYTES + 1024 + 512;
}

struct UmmaTask {
    CUtensorMap mapA, mapBh, mapC;
    float* C;
    int M, N, K, ldc;
    float alpha, beta;
    int pairs_m, tiles_n;   // 256-row and 256-column tiles of this GEMM
    int tile_begin;         // index of its first tile in the group's walk
    int pad_;
};
template <int MAXT>
struct UmmaGroup {
    UmmaTask t[MAXT];
    int ntasks, total_tiles;
};

struct TileRef {
    int task, m0, n0, kt;  // m0: first row of the PAIR's tile (this CTA adds rank * 128)
};
template <int MAXT, int EB>
__device__ __forceinline__ TileRef decode_tile(const UmmaGroup<MAXT>& g, int tile) {
    int t = 0;
#pragma unroll 1
    while (t + 1 < g.ntasks && tile >= g.t[t + 1].tile_begin) ++t;
    const UmmaTask& k = g.t[t];
    const int local = tile - k.tile_begin;
    const int per_group = Q_RASTER * k.tiles_n;
    const int first = (local / per_group) * Q_RASTER;
    const int rows = min(Q_RASTER, k.pairs_m - first);
    const int in_group = local % per_group;
    TileRef r;
    r.task = t;
    r.m0 = (first + in_group % rows) * (2 * G_BM);
    r.n0 = (in_group / rows) * G_BN;
    r.kt = (k.K + 128 / EB - 1) / (128 / EB);
    return r;
}

// arrive on the barrier at the same offset in CTA `target_rank` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t target_rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(remote) : "r"(smem_u32(bar)), "r"(target_rank));
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(remote) : "memory");
}

template <int MAXT, int EB>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(G_THREADS, 1)
gemm_umma_grouped(const __grid_constant__ UmmaGroup<MAXT> g) {
    constexpr int BKE = 128 / EB;
    constexpr int MBOX = 128 / EB;
    constexpr int NBOX = G_BM / MBOX;
    constexpr int BOX_BYTES = BKE * 128;
    constexpr int UK = 32 / EB;
    constexpr uint32_t IDESC = (1u << 4) | ((EB == 2 ? 1u : 2u) << 7) | ((EB == 2 ? 1u : 2u) << 10) | (1u << 15) |
                               ((uint32_t)(G_BN >> 3) << 17) | ((uint32_t)((2 * G_BM) >> 4) << 24);
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t sbase0 = (raw + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (sbase0 - raw);
    uint8_t* ebufs = smem + Q_STAGES * P_STAGE_BYTES;                   // 4 warps x Q_EBUF boxes
    uint64_t* full = reinterpret_cast<uint64_t*>(ebufs + 4 * Q_EBUF * E_BOX_BYTES);
    uint64_t* empty = full + Q_STAGES;
    uint64_t* acc_full = empty + Q_STAGES;     // [2] accumulator stage complete (multicast commit, both CTAs)
    uint64_t* acc_empty = acc_full + 2;        // [2] accumulator stage drained: 8 arrivals on the LEADER's barrier
    uint64_t* cbar = acc_empty + 2;            // [4][Q_EBUF] C boxes landed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(cbar + 4 * Q_EBUF);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < Q_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&acc_full[a], 1);
            mbar_init(&acc_empty[a], 8);
        }
        for (int i = 0; i < 4 * Q_EBUF; ++i) mbar_init(&cbar[i], 1);
        fence_barrier_init();
        fence_proxy_async();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)Q_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::);
    }
    __syncwarp();
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer (one lane per CTA)
        if (lane == 0) {
            int stage = 0, last_task = -1;
            uint32_t phase = 0;
            long long it = 0;
            for (int tile = pair; tile < g.total_tiles; tile += npairs) {
                const TileRef r = decode_tile<MAXT, EB>(g, tile);
                const UmmaTask& k = g.t[r.task];
                if (r.task != last_task) {
                    tma_prefetch_desc(&k.mapA);
                    tma_prefetch_desc(&k.mapBh);
                    last_task = r.task;
                }
                const int m0 = r.m0 + rank * G_BM;
                for (int kt = 0; kt < r.kt; ++kt, ++it) {
                    if (it >= Q_STAGES) mbar_wait(&empty[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * P_STAGE_BYTES;
                    if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * P_STAGE_BYTES);
                    const uint32_t lead_full = smem_u32(&full[stage]) & 0xFEFFFFFFu;
#pragma unroll
                    for (int b = 0; b < NBOX; ++b)
                        tma_load_2d_pair(sa + b * BOX_BYTES, &k.mapA, m0 + b * MBOX, kt * BKE, lead_full);
                    tma_load_2d_pair(sa + G_A_BYTES, &k.mapBh, kt * BKE, r.n0 + rank * (G_BN / 2), lead_full);
                    if (++stage == Q_STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (leader CTA, one lane)
        if (lane == 0 && rank == 0) {
            int stage = 0, i = 0;
            uint32_t phase = 0;
            for (int tile = pair; tile < g.total_tiles; tile += npairs, ++i) {
                const TileRef r = decode_tile<MAXT, EB>(g, tile);
                const int a = i & 1;
                if (i >= 2) {  // both CTAs' epilogues must have drained this accumulator stage
                    mbar_wait(&acc_empty[a], (uint32_t)((i >> 1) - 1) & 1u);
                    tc_fence_after();
                }
                const uint32_t tmem_d = tmem_base + (uint32_t)(a * G_TMEM_COLS);
                for (int kt = 0; kt < r.kt; ++kt) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = sbase0 + stage * P_STAGE_BYTES;
                    const uint32_t sb = sa + G_A_BYTES;
#pragma unroll
                    for (int ku = 0; ku < BKE / UK; ++ku) {
                        const uint64_t adesc = EB == 2 ? umma_desc(sa + ku * UK * 128, BOX_BYTES, 1024, 2)
                                                       : umma_desc(sa + ku * UK * 128, BOX_BYTES, 512, 1);
                        const uint64_t bdesc = umma_desc(sb + ku * 32, 16, 1024);
                        umma_pair<EB>(tmem_d, adesc, bdesc, IDESC, (kt | ku) != 0);
                    }
                    umma_commit_pair(&empty[stage]);
                    if (++stage == Q_STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit_pair(&acc_full[a]);
            }
        }
    } else {
        // ------------------------------------------------------------------ epilogue warps (4 per CTA)
        const int quarter = warp & 3;
        float* bufs = reinterpret_cast<float*>(ebufs + quarter * (Q_EBUF * E_BOX_BYTES));
        uint64_t* bar = cbar + quarter * Q_EBUF;
        uint32_t loads = 0;  // C boxes requested so far on each buffer decides the parity to wait for: count per buffer
        uint32_t nload[Q_EBUF];
#pragma unroll
        for (int b = 0; b < Q_EBUF; ++b) nload[b] = 0;
        (void)loads;
        int i = 0, last_task = -1;
        for (int tile = pair; tile < g.total_tiles; tile += npairs, ++i) {
            const TileRef r = decode_tile<MAXT, EB>(g, tile);
            const UmmaTask& k = g.t[r.task];
            const int a = i & 1;
            const int row0 = r.m0 + rank * G_BM + quarter * 32;
            const bool live = row0 < k.M;  // warp-uniform (n0 < N always: the tile exists)
            const int nchunks = live ? (min(G_BN, k.N - r.n0) + 31) / 32 : 0;
            const bool readc = k.beta != 0.0f;
            const float alpha = k.alpha, beta = k.beta;
            if (lane == 0 && r.task != last_task) tma_prefetch_desc(&k.mapC);
            last_task = r.task;
            // the old C boxes of this tile are requested now, while its MMAs still run; the stores of the previous tile
            // must be done reading the buffers first
            if (readc && nchunks > 0 && lane == 0) {
                tma_store_wait_read<0>();
                for (int c = 0; c < nchunks && c < Q_EBUF; ++c) {
                    mbar_arrive_expect_tx(&bar[c], E_BOX_BYTES);
                    tma_load_2d(bufs + c * (E_BOX_BYTES / 4), &k.mapC, row0, r.n0 + 32 * c, &bar[c]);
                }
            }
            mbar_wait(&acc_full[a], (uint32_t)(i >> 1) & 1u);
            tc_fence_after();
            const uint32_t tmem_acc = tmem_base + (uint32_t)(a * G_TMEM_COLS) + ((uint32_t)(quarter * 32) << 16);
            if (!live) {  // rows beyond M: nothing to write, but the accumulator stage still has to be handed back
                tc_fence_before();
                if (lane == 0) mbar_arrive_cluster(&acc_empty[a], 0);
                continue;
            }
#pragma unroll 1
            for (int c = 0; c < nchunks; ++c) {
                const int b = c % Q_EBUF;
                float* buf = bufs + b * (E_BOX_BYTES / 4);
                uint32_t v[32];
                tmem_ld32(tmem_acc + (uint32_t)(32 * c), v);
                if (c == nchunks - 1) {  // last read of this accumulator stage by this warp: the MMA warp may reuse it
                    tc_fence_before();
                    if (lane == 0) mbar_arrive_cluster(&acc_empty[a], 0);
                }
                if (readc) {
                    mbar_wait(&bar[b], nload[b] & 1u);
                    nload[b]++;
#pragma unroll
                    for (int j = 0; j < 32; ++j) buf[j * 32 + lane] = alpha * __uint_as_float(v[j]) + beta * buf[j * 32 + lane];
                } else {
                    if (lane == 0) tma_store_wait_read<Q_EBUF - 1>();  // the store that last read this buffer is done with it
                    __syncwarp();
#pragma unroll
                    for (int j = 0; j < 32; ++j) buf[j * 32 + lane] = alpha * __uint_as_float(v[j]);
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    tma_store_2d(&k.mapC, buf, row0, r.n0 + 32 * c);
                    tma_store_commit();
                    if (readc && c + Q_EBUF < nchunks) {  // recycle the buffer for the box Q_EBUF chunks ahead
                        tma_store_wait_read<0>();
                        mbar_arrive_expect_tx(&bar[b], E_BOX_BYTES);
                        tma_load_2d(buf, &k.mapC, row0, r.n0 + 32 * (c + Q_EBUF), &bar[b]);
                    }
                }
                // nload is per lane but only its parity is used, and every lane counts the same waits
            }
        }
        if (lane == 0) tma_store_wait_all();
        __syncwarp();
    }
    __syncwarp();
    tc_fence_before();
    cluster_sync_all();  // the peer's MMAs read this CTA's shared memory; TMEM is freed as a pair
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"((uint32_t)Q_TMEM_COLS));
    }
}

struct BfMapKey {
    const void* p;
    int d0, d1, ld, kind;
};
struct BfMapEntry {
    BfMapKey k;
    CUtensorMap m;
    bool valid;
};
// kind: 0 = A operand (boxes of one 128-byte row of m by one stage of k), 1 = B operand (one stage of k by G_BN columns)
static int umma_tmap(CUtensorMap* out, const void* p, int d0, int d1, int ld, int kind, int eb) {
    static thread_local std::vector<BfMapEntry> cache;
    if (cache.empty()) cache.resize(16384);
    const int tag = kind + 2 * eb;
    uintptr_t h = ((uintptr_t)p >> 9) * 0x9E3779B97F4A7C15ull + (uintptr_t)(d0 * 31 + d1 * 17 + ld * 7 + tag);
    BfMapEntry& e = cache[(h >> 20) & 16383];
    if (e.valid && e.k.p == p && e.k.d0 == d0 && e.k.d1 == d1 && e.k.ld == ld && e.k.kind == tag) {
        *out = e.m;
        return 0;
    }
    const CUtensorMapDataType dt = eb == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    const uint32_t row = 128 / eb;  // elements in one 128-byte swizzle row
    // the MN-major fp32 (TF32) operand needs the 32-byte-atom flavour of the 128-byte swizzle (see umma_desc)
    const CUtensorMapSwizzle swa = eb == 4 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
    int rc = kind == 3 ? encode_tmap_2d(&e.m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, p, d0, d1, ld, 32, 32, CU_TENSOR_MAP_SWIZZLE_NONE)
           : kind == 0 ? encode_tmap_2d(&e.m, dt, eb, p, d0, d1, ld, row, row, swa)
                       : encode_tmap_2d(&e.m, dt, eb, p, d0, d1, ld, row, kind == 2 ? G_BN / 2 : G_BN,
                                        CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) {
        e.valid = false;
        return rc;
    }
    e.k = {p, d0, d1, ld, tag};
    e.valid = true;
    *out = e.m;
    return 0;
}

// ---- host side of the grouped kernel -------------------------------------------------------------------------------
constexpr int Q_MAX_GROUP = 16;  // GEMMs per launch: 16 x 448 bytes of tasks in the kernel parameter space

// pairs that can be resident at once (one per TPC whose two SMs are both free for a 2-CTA cluster)
template <typename K>
static int resident_pairs(K kernel, int dev) {
    static int cached[64] = {0};
    if (dev >= 0 && dev < 64 && cached[dev]) return cached[dev];
    int pairs = 0;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) == cudaSuccess) pairs = prop.multiProcessorCount / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * (pairs > 0 ? pairs : 74));
    cfg.blockDim = dim3(G_THREADS);
    cfg.dynamicSmemBytes = grouped_smem();
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = 2;
    attr.val.clusterDim.y = 1;
    attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    int clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&clusters, kernel, &cfg) == cudaSuccess && clusters > 0) pairs = clusters;
    else cudaGetLastError();
    if (pairs <= 0) pairs = 74;
    if (dev >= 0 && dev < 64) cached[dev] = pairs;
    return pairs;
}

template <int MAXT, int EB>
static int launch_group_n(const UmmaTask* tasks, int ntasks, int total_tiles, int dev, cudaStream_t stream) {
    static bool attr_set[64] = {false};
    auto kernel = gemm_umma_grouped<MAXT, EB>;
    if (dev < 64 && !attr_set[dev]) {
        if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)grouped_smem()) != cudaSuccess)
            return PB2_ERR_CUDA;
        attr_set[dev] = true;
    }
    UmmaGroup<MAXT> g;
    memcpy(g.t, tasks, sizeof(UmmaTask) * (size_t)ntasks);
    g.ntasks = ntasks;
    g.total_tiles = total_tiles;
    const int pairs = std::min(total_tiles, resident_pairs(kernel, dev));
    kernel<<<dim3(2 * pairs), G_THREADS, grouped_smem(), stream>>>(g);
    return cudaGetLastError() == cudaSuccess ? PB2_SUCCESS : PB2_ERR_CUDA;
}

template <int EB>
static int launch_group(std::vector<UmmaTask>& tasks, cudaStream_t stream) {
    if (tasks.empty()) return PB2_SUCCESS;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return PB2_ERR_CUDA;
    int total = 0;
    for (UmmaTask& t : tasks) {
        t.tile_begin = total;
        total += t.pairs_m * t.tiles_n;
    }
    const int n = (int)tasks.size();
    int rc = n == 1   ? launch_group_n<1, EB>(tasks.data(), n, total, dev, stream)
             : n <= 4 ? launch_group_n<4, EB>(tasks.data(), n, total, dev, stream)
                      : launch_group_n<Q_MAX_GROUP, EB>(tasks.data(), n, total, dev, stream);
    tasks.clear();
    return rc;
}

// A batch: the GEMM calls one thread makes on one stream between parsec_b200_batch_begin and _end are collected and
// launched together (one grouped kernel per element type and per Q_MAX_GROUP calls).
struct UmmaBatch {
    bool open = false;
    void* stream = nullptr;
    int status = PB2_SUCCESS;
    std::vector<UmmaTask> bf16, tf32;
};
static thread_local UmmaBatch g_batch;

template <int EB>
static int launch_umma_gemm(char transa, char transb, int m, int n, int k, float alpha, const void* A, int lda, const void* B,
                            int ldb, float beta, float* C, int ldc, void* stream) {
    if ((transa != 'N' && transa != 'n') || (transb != 'N' && transb != 'n')) return PB2_ERR_NOT_SUPPORTED;
    if (m < 0 || n < 0 || k < 0 || lda < (m > 1 ? m : 1) || ldb < (k > 1 ? k : 1) || ldc < (m > 1 ? m : 1))
        return PB2_ERR_BAD_PARAM;
    if (m == 0 || n == 0) return PB2_SUCCESS;
    // TMA: 16-byte aligned bases and row pitches
    const int pitch_mask = 16 / EB - 1;
    if ((((uintptr_t)A | (uintptr_t)B) & 15) || (lda & pitch_mask) || (ldb & pitch_mask) || (((uintptr_t)C) & 3))
        return PB2_ERR_NOT_SUPPORTED;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return PB2_ERR_CUDA;
    static bool attr_set[64] = {false};
    if (dev < 64 && !attr_set[dev]) {
        if (cudaFuncSetAttribute(gemm_bf16_tcgen05<G_STAGES_1CTA, EB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)gemm_smem(G_STAGES_1CTA)) != cudaSuccess ||
            cudaFuncSetAttribute(gemm_bf16_tcgen05<G_STAGES_2CTA, EB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)gemm_smem(G_STAGES_2CTA)) != cudaSuccess)
            return PB2_ERR_CUDA;
        attr_set[dev] = true;
    }
    // PB2_GEMM_BF16_CTAS: 1 = deep single-CTA ring, one CTA per SM; 2 = two independent CTAs per SM;
    // pair = cta_group::2 CTA pairs, two pairs per TPC; pair1 = CTA pairs with the deep ring (kept for A/B runs)
    // grouped = the persistent grouped kernel (default): one launch per call, or per batch of calls
    static int ctas = -1;
    if (ctas < 0) {
        const char* e = getenv("PB2_GEMM_BF16_CTAS");
        ctas = !e ? 5 : !strcmp(e, "grouped") ? 5 : !strcmp(e, "1") ? 1 : !strcmp(e, "pair") ? 3 : !strcmp(e, "pair1") ? 4 : 2;
    }
    CUtensorMap ma, mb;
    if (k > 0 && (umma_tmap(&ma, A, m, k, lda, 0, EB) || umma_tmap(&mb, B, k, n, ldb, ctas >= 3 ? 2 : 1, EB)))
        return PB2_ERR_CUDA;
    if (k == 0) {  // nothing to stage: the kernel only scales C; the maps are never dereferenced
        memset(&ma, 0, sizeof(ma));
        memset(&mb, 0, sizeof(mb));
    }
    // C through TMA (boxes of 32 rows x 32 columns) when its base and pitch allow it; PB2_GEMM_EPILOGUE=direct keeps
    // the register epilogue for A/B runs
    static int epi_direct = -1;
    if (epi_direct < 0) {
        const char* e = getenv("PB2_GEMM_EPILOGUE");
        epi_direct = (e && !strcmp(e, "direct")) ? 1 : 0;
    }
    CUtensorMap mc;
    int c_tma = (!epi_direct && !(((uintptr_t)C) & 15) && !(ldc & 3)) ? 1 : 0;
    if (c_tma && umma_tmap(&mc, C, m, n, ldc, 3, 4)) return PB2_ERR_CUDA;
    if (!c_tma) memset(&mc, 0, sizeof(mc));
    int mode = ctas;
    if (mode == 5 && (k == 0 || !c_tma)) mode = 3;  // nothing to multiply, or a C that TMA cannot address: per-tile kernel
    if (mode == 5) {
        UmmaTask t;
        memset(&t, 0, sizeof(t));
        t.mapA = ma; t.mapBh = mb; t.mapC = mc;
        t.C = C; t.M = m; t.N = n; t.K = k; t.ldc = ldc; t.alpha = alpha; t.beta = beta;
        t.pairs_m = (m + 2 * G_BM - 1) / (2 * G_BM);
        t.tiles_n = (n + G_BN - 1) / G_BN;
        if (g_batch.open && g_batch.stream == stream) {  // collected; launched by parsec_b200_batch_end or when full
            std::vector<UmmaTask>& q = EB == 2 ? g_batch.bf16 : g_batch.tf32;
            q.push_back(t);
            if ((int)q.size() == Q_MAX_GROUP) {
                int rc = launch_group<EB>(q, (cudaStream_t)stream);
                if (rc != PB2_SUCCESS) g_batch.status = rc;
                return rc;
            }
            return PB2_SUCCESS;
        }
        // A single call with no more tiles than CTA pairs gains nothing from the persistent walk (one tile per pair, one
        // CTA per SM) and loses the co-residency of two per-tile CTAs per SM that lets launches on other streams overlap
        // (measured: a stream of unbatched 2048^3 tasks on 4 streams 1130 -> 638 TFLOP/s): it takes the per-tile kernel.
        if (t.pairs_m * t.tiles_n > 74) {
            std::vector<UmmaTask> one(1, t);
            return launch_group<EB>(one, (cudaStream_t)stream);
        }
        mode = 3;
    }
    if (mode >= 3) {
        static bool pair_attr_set[64] = {false};
        if (dev < 64 && !pair_attr_set[dev]) {
            if (cudaFuncSetAttribute(gemm_umma_pair<P_STAGES_2CTA, EB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)pair_smem(P_STAGES_2CTA)) != cudaSuccess ||
                cudaFuncSetAttribute(gemm_umma_pair<P_STAGES_1CTA, EB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)pair_smem(P_STAGES_1CTA)) != cudaSuccess)
                return PB2_ERR_CUDA;
            pair_attr_set[dev] = true;
        }
        dim3 pgrid(2 * ((m + 2 * G_BM - 1) / (2 * G_BM)) * ((n + G_BN - 1) / G_BN));
        if (mode == 3)
            gemm_umma_pair<P_STAGES_2CTA, EB><<<pgrid, G_THREADS, pair_smem(P_STAGES_2CTA), (cudaStream_t)stream>>>(
                ma, mb, m, n, k, alpha, beta, C, ldc, mc, c_tma);
        else
            gemm_umma_pair<P_STAGES_1CTA, EB><<<pgrid, G_THREADS, pair_smem(P_STAGES_1CTA), (cudaStream_t)stream>>>(
                ma, mb, m, n, k, alpha, beta, C, ldc, mc, c_tma);
        return cudaGetLastError() == cudaSuccess ? PB2_SUCCESS : PB2_ERR_CUDA;
    }
    dim3 grid(((m + G_BM - 1) / G_BM) * ((n + G_BN - 1) / G_BN));
    if (mode == 1)
        gemm_bf16_tcgen05<G_STAGES_1CTA, EB><<<grid, G_THREADS, gemm_smem(G_STAGES_1CTA), (cudaStream_t)stream>>>(
            ma, mb, m, n, k, alpha, beta, C, ldc, mc, c_tma);
    else
        gemm_bf16_tcgen05<G_STAGES_2CTA, EB><<<grid, G_THREADS, gemm_smem(G_STAGES_2CTA), (cudaStream_t)stream>>>(
            ma, mb, m, n, k, alpha, beta, C, ldc, mc, c_tma);
    return cudaGetLastError() == cudaSuccess ? PB2_SUCCESS : PB2_ERR_CUDA;
}

}  // namespace pb2

// Batches: the tile GEMMs a caller has ready at the same time go to the GPU as ONE grouped launch.  Between
// parsec_b200_batch_begin(stream) and parsec_b200_batch_end(stream) the bf16 / tf32 / fp64 GEMM calls this thread makes
// on `stream` are only recorded; _end (or the 16th tensor-core call, the 64th FP64 call) launches them.  The device module brackets the bodies of the
// tasks it submits together with these two calls, so a JDF body that calls parsec_b200_gemm_bf16 needs no change.
namespace pb2 {
void dgemm_batch_open(void* stream);   // dgemm_sm100.cu: the FP64 kernel's share of a batch
int dgemm_batch_close(void* stream);
}
extern "C" int parsec_b200_batch_begin(void* stream) {
    if (pb2::g_batch.open) return PB2_ERR_EXISTS;
    pb2::dgemm_batch_open(stream);
    pb2::g_batch.open = true;
    pb2::g_batch.stream = stream;
    pb2::g_batch.status = PB2_SUCCESS;
    return PB2_SUCCESS;
}
extern "C" int parsec_b200_batch_end(void* stream) {
    if (!pb2::g_batch.open || pb2::g_batch.stream != stream) return PB2_ERR_BAD_PARAM;
    pb2::g_batch.open = false;
    int rc = pb2::g_batch.status;
    int r1 = pb2::launch_group<2>(pb2::g_batch.bf16, (cudaStream_t)stream);
    int r2 = pb2::launch_group<4>(pb2::g_batch.tf32, (cudaStream_t)stream);
    int r3 = pb2::dgemm_batch_close(stream);
    return rc != PB2_SUCCESS ? rc : r1 != PB2_SUCCESS ? r1 : r2 != PB2_SUCCESS ? r2 : r3;
}

extern "C" int parsec_b200_gemm_bf16_sm100(char transa, char transb, int m, int n, int k, float alpha, const void* A,
                                           int lda, const void* B, int ldb, float beta, float* C, int ldc,
                                           void* stream) {
    return pb2::launch_umma_gemm<2>(transa, transb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, stream);
}

// fp32 operands consumed by the tensor cores as TF32 (the low 13 mantissa bits of A and B are ignored), fp32 accumulate:
// the cublasSgemm shape with CUBLAS_COMPUTE_32F_FAST_TF32.
extern "C" int parsec_b200_gemm_tf32_sm100(char transa, char transb, int m, int n, int k, float alpha, const float* A,
                                           int lda, const float* B, int ldb, float beta, float* C, int ldc,
                                           void* stream) {
    return pb2::launch_umma_gemm<4>(transa, transb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, stream);
}
