// Standalone B200 micro-measurements used to set the FP64 roofline denominator and to
// sanity-check the tile kernel before the device module is wired in.
//   1. DMMA.8x8x4 issue-rate peak (register-only loop)
//   2. cuBLAS DGEMM 8192^3 and 1024^3 (the reference bodies' cublasDgemm_v2 on this box)
//   3. parsec_b200_dgemm_sm100: correctness vs cuBLAS, then the same shapes
//   4. pinned H2D / D2H copy bandwidth at tile sizes
// Build: see parsec-1_b200/Makefile (target tools).  Not part of the shipped library.
#include <cublas_v2.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

extern "C" int parsec_b200_dgemm_sm100(char, char, int, int, int, double, const double*, int, const double*, int,
                                        double, double*, int, void*);

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e_ = (x);                                                         \
        if (e_ != cudaSuccess) {                                                      \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(2);                                                                  \
        }                                                                             \
    } while (0)

__global__ void __launch_bounds__(256, 1) dmma_peak(double* out, int iters, double a, double b) {
    double acc[16][2];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i][0] = acc[i][1] = 0.0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                         : "+d"(acc[i][0]), "+d"(acc[i][1])
                         : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i][0] + acc[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void dfma_peak(double* out, int iters, double a, double b) {
    double acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void fill(double* p, size_t n, unsigned long long seed) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        unsigned long long x = (i + 1) * 6364136223846793005ULL + seed;
        x ^= x >> 29;
        x *= 0xBF58476D1CE4E5B9ULL;
        x ^= x >> 32;
        p[i] = 0.5 - (double)(x >> 11) * (1.0 / 9007199254740992.0);
    }
}

__global__ void maxdiff(const double* a, const double* b, size_t n, double* out) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    double m = 0;
    for (; i < n; i += stride) m = fmax(m, fabs(a[i] - b[i]));
    unsigned long long* o = (unsigned long long*)out;
    atomicMax(o, (unsigned long long)__double_as_longlong(m));
}

static float time_ms(cudaEvent_t a, cudaEvent_t b) {
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    return ms;
}

int main(int argc, char** argv) {
    int quick = argc > 1 && atoi(argv[1]) == 1;
    int probe = argc > 1 && atoi(argv[1]) == 2;  // profiling mode: one cuBLAS + a few of ours at n=4096
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    printf("device %s sms=%d clock=%d kHz\n", prop.name, prop.multiProcessorCount, prop.clockRate);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));

    if (probe) {
        int n = argc > 2 ? atoi(argv[2]) : 4096;
        size_t ne = (size_t)n * n;
        double *A, *B, *C1;
        CK(cudaMalloc(&A, ne * 8));
        CK(cudaMalloc(&B, ne * 8));
        CK(cudaMalloc(&C1, ne * 8));
        fill<<<1024, 256>>>(A, ne, 1);
        fill<<<1024, 256>>>(B, ne, 2);
        fill<<<1024, 256>>>(C1, ne, 3);
        cublasHandle_t hh;
        cublasCreate(&hh);
        const double o = 1.0;
        cublasDgemm(hh, CUBLAS_OP_N, CUBLAS_OP_N, n, n, n, &o, A, n, B, n, &o, C1, n);
        for (int i = 0; i < 2; ++i) parsec_b200_dgemm_sm100('N', 'N', n, n, n, 1.0, A, n, B, n, 1.0, C1, n, 0);
        CK(cudaDeviceSynchronize());
        printf("PROBE DONE\n");
        return 0;
    }
    // 1. DMMA / DFMA peaks
    {
        double* out;
        CK(cudaMalloc(&out, sizeof(double) * 148 * 8 * 1024));
        for (int warps = 4; warps <= 32; warps *= 2) {
            int iters = 20000;
            dmma_peak<<<prop.multiProcessorCount, warps * 32>>>(out, 100, 1.0, 1.0);
            CK(cudaDeviceSynchronize());
            float best = 1e30f;
            for (int r = 0; r < 5; ++r) {
                CK(cudaEventRecord(e0));
                dmma_peak<<<prop.multiProcessorCount, warps * 32>>>(out, iters, 1.0, 1.0);
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                best = fminf(best, time_ms(e0, e1));
            }
            double flops = 2.0 * 8 * 8 * 4 * 16.0 * iters * warps * prop.multiProcessorCount;
            printf("DMMA peak: %2d warps/SM  %.2f TFLOP/s (%.3f ms)\n", warps, flops / best * 1e-9, best);
        }
        for (int warps = 8; warps <= 32; warps *= 2) {
            int iters = 20000;
            float best = 1e30f;
            for (int r = 0; r < 4; ++r) {
                CK(cudaEventRecord(e0));
                dfma_peak<<<prop.multiProcessorCount, warps * 32>>>(out, iters, 1.0000001, 1e-9);
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                best = fminf(best, time_ms(e0, e1));
            }
            double flops = 2.0 * 16.0 * iters * warps * 32 * prop.multiProcessorCount;
            printf("DFMA peak: %2d warps/SM  %.2f TFLOP/s (%.3f ms)\n", warps, flops / best * 1e-9, best);
        }
        CK(cudaFree(out));
    }

    cublasHandle_t h;
    cublasCreate(&h);
    const double one = 1.0;

    int sizes[] = {1024, 2048, 4096, 8192};
    for (int si = 0; si < (quick ? 2 : 4); ++si) {
        int n = sizes[si];
        size_t ne = (size_t)n * n;
        double *A, *B, *C1, *C2, *d;
        CK(cudaMalloc(&A, ne * 8));
        CK(cudaMalloc(&B, ne * 8));
        CK(cudaMalloc(&C1, ne * 8));
        CK(cudaMalloc(&C2, ne * 8));
        CK(cudaMalloc(&d, 8));
        fill<<<1024, 256>>>(A, ne, 1);
        fill<<<1024, 256>>>(B, ne, 2);
        fill<<<1024, 256>>>(C1, ne, 3);
        CK(cudaMemcpy(C2, C1, ne * 8, cudaMemcpyDeviceToDevice));
        CK(cudaMemset(d, 0, 8));
        cublasDgemm(h, CUBLAS_OP_N, CUBLAS_OP_N, n, n, n, &one, A, n, B, n, &one, C1, n);
        int rc = parsec_b200_dgemm_sm100('N', 'N', n, n, n, 1.0, A, n, B, n, 1.0, C2, n, 0);
        CK(cudaDeviceSynchronize());
        maxdiff<<<1024, 256>>>(C1, C2, ne, d);
        double hd;
        CK(cudaMemcpy(&hd, d, 8, cudaMemcpyDeviceToHost));
        printf("n=%d rc=%d max|ours-cublas| = %.3e (n*eps*0.25 = %.3e)\n", n, rc, hd, n * 1.1e-16 * 0.25);

        int reps = n >= 4096 ? 5 : 50;
        for (int which = 0; which < 2; ++which) {
            float best = 1e30f, tot = 0;
            for (int r = 0; r < 3; ++r) {
                CK(cudaEventRecord(e0));
                for (int i = 0; i < reps; ++i) {
                    if (which == 0)
                        cublasDgemm(h, CUBLAS_OP_N, CUBLAS_OP_N, n, n, n, &one, A, n, B, n, &one, C1, n);
                    else
                        parsec_b200_dgemm_sm100('N', 'N', n, n, n, 1.0, A, n, B, n, 1.0, C2, n, 0);
                }
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                float ms = time_ms(e0, e1) / reps;
                best = fminf(best, ms);
                tot += ms;
            }
            printf("  %-8s n=%5d  %.4f ms  %.2f TFLOP/s\n", which ? "ours" : "cublas", n, best,
                   2.0 * n * n * (double)n / best * 1e-9);
        }
        cudaFree(A); cudaFree(B); cudaFree(C1); cudaFree(C2); cudaFree(d);
    }

    // tile stream concurrency: many independent 1024^3 tile GEMMs over S streams
    {
        const int nb = 1024, ntile = 48;
        size_t ne = (size_t)nb * nb;
        double* pool;
        CK(cudaMalloc(&pool, ne * 8 * ntile * 3));
        fill<<<2048, 256>>>(pool, ne * ntile * 3, 7);
        CK(cudaDeviceSynchronize());
        for (int which = 0; which < 2; ++which) {
            for (int S = 1; S <= 8; S *= 2) {
                std::vector<cudaStream_t> st(S);
                std::vector<cublasHandle_t> hs(S);
                for (int s = 0; s < S; ++s) {
                    CK(cudaStreamCreateWithFlags(&st[s], cudaStreamNonBlocking));
                    cublasCreate(&hs[s]);
                    cublasSetStream(hs[s], st[s]);
                }
                float best = 1e30f;
                const int rounds = 8;
                for (int r = 0; r < 3; ++r) {
                    CK(cudaDeviceSynchronize());
                    CK(cudaEventRecord(e0, 0));
                    for (int s = 0; s < S; ++s) CK(cudaStreamWaitEvent(st[s], e0, 0));
                    for (int i = 0; i < ntile * rounds; ++i) {
                        int tix = i % ntile, s = i % S;
                        double* A = pool + (size_t)(3 * tix) * ne;
                        double* B = A + ne;
                        double* C = B + ne;
                        if (which == 0)
                            cublasDgemm(hs[s], CUBLAS_OP_N, CUBLAS_OP_N, nb, nb, nb, &one, A, nb, B, nb, &one, C, nb);
                        else
                            parsec_b200_dgemm_sm100('N', 'N', nb, nb, nb, 1.0, A, nb, B, nb, 1.0, C, nb, st[s]);
                    }
                    std::vector<cudaEvent_t> ev(S);
                    for (int s = 0; s < S; ++s) {
                        CK(cudaEventCreateWithFlags(&ev[s], cudaEventDisableTiming));
                        CK(cudaEventRecord(ev[s], st[s]));
                        CK(cudaStreamWaitEvent(0, ev[s], 0));
                    }
                    CK(cudaEventRecord(e1, 0));
                    CK(cudaEventSynchronize(e1));
                    for (int s = 0; s < S; ++s) cudaEventDestroy(ev[s]);
                    best = fminf(best, time_ms(e0, e1));
                }
                printf("tile stream %-6s S=%d: %d x 1024^3 in %.3f ms -> %.2f TFLOP/s\n", which ? "ours" : "cublas", S,
                       ntile * rounds, best, 2.0 * nb * nb * (double)nb * ntile * rounds / best * 1e-9);
                for (int s = 0; s < S; ++s) {
                    cublasDestroy(hs[s]);
                    cudaStreamDestroy(st[s]);
                }
            }
        }
        cudaFree(pool);
    }

    // 4. pinned copies
    {
        size_t sz = 256ull << 20;
        void *hbuf, *dbuf;
        CK(cudaHostAlloc(&hbuf, sz, cudaHostAllocPortable));
        CK(cudaMalloc(&dbuf, sz));
        memset(hbuf, 1, sz);
        size_t chunks[] = {512 << 10, 8 << 20, 32 << 20, 256 << 20};
        for (size_t c : chunks) {
            for (int dir = 0; dir < 2; ++dir) {
                float best = 1e30f;
                for (int r = 0; r < 3; ++r) {
                    CK(cudaEventRecord(e0));
                    for (size_t off = 0; off + c <= sz; off += c) {
                        if (dir == 0)
                            CK(cudaMemcpyAsync((char*)dbuf + off, (char*)hbuf + off, c, cudaMemcpyHostToDevice, 0));
                        else
                            CK(cudaMemcpyAsync((char*)hbuf + off, (char*)dbuf + off, c, cudaMemcpyDeviceToHost, 0));
                    }
                    CK(cudaEventRecord(e1));
                    CK(cudaEventSynchronize(e1));
                    best = fminf(best, time_ms(e0, e1));
                }
                printf("%s chunk %6zu KiB: %.1f GB/s\n", dir ? "D2H" : "H2D", c >> 10, sz / best * 1e-6);
            }
        }
        // bidirectional
        cudaStream_t s1, s2;
        CK(cudaStreamCreate(&s1));
        CK(cudaStreamCreate(&s2));
        void *h2, *d2;
        CK(cudaHostAlloc(&h2, sz, cudaHostAllocPortable));
        CK(cudaMalloc(&d2, sz));
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        CK(cudaStreamWaitEvent(s1, e0, 0));
        CK(cudaStreamWaitEvent(s2, e0, 0));
        for (int r = 0; r < 4; ++r) {
            CK(cudaMemcpyAsync(dbuf, hbuf, sz, cudaMemcpyHostToDevice, s1));
            CK(cudaMemcpyAsync(h2, d2, sz, cudaMemcpyDeviceToHost, s2));
        }
        cudaEvent_t ea, eb;
        CK(cudaEventCreate(&ea));
        CK(cudaEventCreate(&eb));
        CK(cudaEventRecord(ea, s1));
        CK(cudaEventRecord(eb, s2));
        CK(cudaStreamWaitEvent(0, ea, 0));
        CK(cudaStreamWaitEvent(0, eb, 0));
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        printf("bidirectional H2D+D2H: %.1f GB/s each way\n", 4.0 * sz / time_ms(e0, e1) * 1e-6);
    }
    printf("MICROBENCH DONE\n");
    return 0;
}
