// Where does a stream of 2048^3 bf16 tile tasks spend its time?  Builds the tcgen05 kernels with -DPB2_UMMA_TRACE (per-CTA
// globaltimer marks: entry, set-up done, accumulator complete, epilogue done) and prints the phase lengths and one
// SM's timeline.  Measurement tool only, not part of the library.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DPB2_UMMA_TRACE -Icsrc -I../include \
//        tools/umma_trace.cu build/tma_utils.o -o tools/umma_trace -lcudart
#include "../csrc/gemm_tcgen05.cu"

#include <algorithm>
#include <array>
#include <cstdio>

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e = (x);                                                          \
        if (e != cudaSuccess) {                                                       \
            printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__);    \
            return 1;                                                                 \
        }                                                                             \
    } while (0)

int main(int argc, char** argv) {
    const int nb = 2048, ntile = 24, S = argc > 1 ? atoi(argv[1]) : 4, ntask = 96;
    const float beta = argc > 2 ? (float)atof(argv[2]) : 1.0f;
    std::vector<void*> A(ntile), B(ntile);
    std::vector<float*> C(ntile);
    for (int i = 0; i < ntile; ++i) {
        CK(cudaMalloc(&A[i], (size_t)nb * nb * 2));
        CK(cudaMalloc(&B[i], (size_t)nb * nb * 2));
        CK(cudaMalloc(&C[i], (size_t)nb * nb * 4));
        CK(cudaMemset(A[i], 0, (size_t)nb * nb * 2));
        CK(cudaMemset(B[i], 0, (size_t)nb * nb * 2));
        CK(cudaMemset(C[i], 0, (size_t)nb * nb * 4));
    }
    std::vector<cudaStream_t> st(S);
    for (auto& s : st) CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int rep = 0; rep < 2; ++rep) {
        unsigned zero = 0;
        CK(cudaMemcpyToSymbol(pb2::pb2_trace_n, &zero, sizeof(zero)));
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0, 0));
        for (auto& s : st) CK(cudaStreamWaitEvent(s, e0, 0));
        for (int i = 0; i < ntask; ++i) {
            const int t = i % ntile;
            if (parsec_b200_gemm_bf16_sm100('N', 'N', nb, nb, nb, 1.0f, A[t], nb, B[t], nb, beta, C[t], nb, st[i % S])) return 2;
        }
        CK(cudaDeviceSynchronize());
    }
    unsigned n = 0;
    CK(cudaMemcpyFromSymbol(&n, pb2::pb2_trace_n, sizeof(n)));
    std::vector<unsigned long long> buf((size_t)5 * n);
    CK(cudaMemcpyFromSymbol(buf.data(), pb2::pb2_trace_buf, buf.size() * 8));
    unsigned long long tmin = ~0ull, tmax = 0;
    for (unsigned i = 0; i < n; ++i) {
        tmin = std::min(tmin, buf[5 * i]);
        tmax = std::max(tmax, buf[5 * i + 3]);
    }
    const double total_us = (tmax - tmin) * 1e-3;
    printf("S=%d beta=%.0f: %u CTAs, %d tasks in %.1f us -> %.1f TFLOP/s, %.2f us per task\n", S, beta, n, ntask, total_us,
           2.0 * nb * nb * nb * ntask / total_us * 1e-6, total_us / ntask);
    double p[3] = {0, 0, 0};
    std::vector<double> d[3];
    for (unsigned i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) {
            const double v = (buf[5 * i + k + 1] - buf[5 * i + k]) * 1e-3;
            p[k] += v;
            d[k].push_back(v);
        }
    const char* names[3] = {"set-up (barriers, TMEM alloc, cluster sync)", "main loop (first TMA to accumulator complete)",
                            "epilogue (C in, C out)"};
    for (int k = 0; k < 3; ++k) {
        std::sort(d[k].begin(), d[k].end());
        printf("  %-48s mean %7.2f us  p10 %7.2f  p50 %7.2f  p90 %7.2f\n", names[k], p[k] / n, d[k][n / 10], d[k][n / 2],
               d[k][n * 9 / 10]);
    }
    double alloc = 0;
    for (unsigned i = 0; i < n; ++i) alloc += (double)(buf[5 * i + 4] >> 44) * 1e-3;
    printf("  of the set-up, tcgen05.alloc took            mean %7.2f us\n", alloc / n);
    printf("  CTA lifetime mean %.2f us; ideal main loop alone: %.2f us at 8192 flop/clk/SM, 1.92 GHz\n", (p[0] + p[1] + p[2]) / n,
           2.0 * 128 * 256 * nb / 8192 / 1920.0);
    // one SM's timeline
    const unsigned sm = (unsigned)(buf[4] & 0xffffffffu);
    std::vector<std::array<double, 5>> rows;
    for (unsigned i = 0; i < n; ++i)
        if ((unsigned)(buf[5 * i + 4] & 0xffffffffu) == sm)
            rows.push_back({(buf[5 * i] - tmin) * 1e-3, (buf[5 * i + 1] - tmin) * 1e-3, (buf[5 * i + 2] - tmin) * 1e-3,
                            (buf[5 * i + 3] - tmin) * 1e-3, (double)((buf[5 * i + 4] >> 32) & 0xFFF)});
    std::sort(rows.begin(), rows.end());
    printf("  SM %u ran %zu CTAs; entry / set-up done / accumulator complete / done (us), block:\n", sm, rows.size());
    for (size_t i = 0; i < rows.size() && i < 24; ++i)
        printf("    %8.2f %8.2f %8.2f %8.2f   block %4.0f\n", rows[i][0], rows[i][1], rows[i][2], rows[i][3], rows[i][4]);
    return 0;
}
