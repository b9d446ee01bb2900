/* ce_bench.c -- latency and bandwidth of the pb2_ce_* transport between two processes of this box.
 *   gcc -O2 -I../../include ce_bench.c -o ce_bench -L.. -lparsec_b200 -Wl,-rpath,'$ORIGIN/..'
 *   ./ce_bench            (forks the second rank itself; PB2_CE_NO_CMA=1 for the staged put path)
 * Prints one JSON line: half round trip of an 8-byte and a 1 KiB active message, rate of 64-byte messages one way,
 * and put bandwidth for 1 MiB, 8 MiB (one fp64 tile at NB=1024) and 32 MiB (NB=2048) blocks. */
#define _GNU_SOURCE
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>
#include <sys/wait.h>
#include "parsec_b200.h"

static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

static volatile long got = 0;
static uint64_t peer_addr = 0;
static void on_msg(void *u, int src, uint64_t tag, const void *msg, size_t len)
{
    (void)u; (void)src; (void)len;
    if( 2 == tag ) memcpy(&peer_addr, msg, 8);
    got++;
}

static void send_all(pb2_ce_t *ce, int dst, uint64_t tag, const void *m, size_t n)
{
    while( PB2_SUCCESS != pb2_ce_send(ce, dst, tag, m, n) ) pb2_ce_poll(ce, on_msg, NULL, 0);
}

static void wait_for(pb2_ce_t *ce, long n) { while( got < n ) pb2_ce_poll(ce, on_msg, NULL, 0); }

int main(void)
{
    char job[64];
    snprintf(job, sizeof(job), "bench%d", (int)getpid());
    pid_t child = fork();
    const int rank = child ? 0 : 1, peer = 1 - rank;
    pb2_ce_t *ce;
    if( PB2_SUCCESS != pb2_ce_attach(job, rank, 2, 0, &ce) ) return 1;
    pb2_ce_set_wait_poll(ce, on_msg, NULL);
    char buf[1024] = {0};
    double lat[2];
    const size_t sizes[2] = {8, 1024};
    for( int s = 0; s < 2; s++ ) {   /* ping-pong */
        const int iters = 200000;
        pb2_ce_barrier(ce);
        got = 0;
        double t0 = now();
        for( int i = 0; i < iters; i++ ) {
            if( 0 == rank ) { send_all(ce, peer, 1, buf, sizes[s]); wait_for(ce, i + 1); }
            else            { wait_for(ce, i + 1); send_all(ce, peer, 1, buf, sizes[s]); }
        }
        lat[s] = (now() - t0) / iters / 2 * 1e6;
    }
    /* one-way message rate */
    const long burst = 2000000;
    pb2_ce_barrier(ce);
    got = 0;
    double t0 = now(), rate;
    if( 0 == rank ) { for( long i = 0; i < burst; i++ ) send_all(ce, peer, 1, buf, 64); pb2_ce_barrier(ce); }
    else { wait_for(ce, burst); pb2_ce_barrier(ce); }
    rate = burst / (now() - t0) * 1e-6;
    /* put bandwidth: rank 0 writes into rank 1 */
    const size_t blocks[3] = {1u << 20, 8u << 20, 32u << 20};
    double gbs[3];
    char *mem = (char*)malloc(32u << 20);
    memset(mem, rank + 1, 32u << 20);
    uint64_t mine = (uint64_t)(uintptr_t)mem;
    got = 0;
    send_all(ce, peer, 2, &mine, 8);
    wait_for(ce, 1);
    for( int b = 0; b < 3; b++ ) {
        const int reps = (int)((2048u << 20) / blocks[b]);
        pb2_ce_barrier(ce);
        double t1 = now();
        if( 0 == rank ) {
            pb2_ce_iov_t l = {mem, blocks[b]}, r = {(void*)(uintptr_t)peer_addr, blocks[b]};
            for( int i = 0; i < reps; i++ ) if( PB2_SUCCESS != pb2_ce_put(ce, peer, &l, 1, &r, 1) ) return 2;
            send_all(ce, peer, 1, buf, 8);
        } else {
            got = 0;
            wait_for(ce, 1);
        }
        pb2_ce_barrier(ce);
        gbs[b] = (double)blocks[b] * reps / (now() - t1) * 1e-9;
    }
    /* the same 8 MiB put, but never twice from or to the same memory: 32 source and 32 target tiles in turn (256 MiB
     * each side, far more than the caches hold), which is what a runtime moving distinct tiles does */
    double cold_gbs = 0.0, cold_memcpy_gbs = 0.0;
    {
        const size_t tile = 8u << 20, ntiles = 32;
        char *big = (char*)malloc(tile * ntiles);
        memset(big, rank + 3, tile * ntiles);
        uint64_t mine2 = (uint64_t)(uintptr_t)big;
        got = 0;
        send_all(ce, peer, 2, &mine2, 8);
        wait_for(ce, 1);
        pb2_ce_barrier(ce);
        double t1 = now();
        if( 0 == rank ) {
            for( int rep = 0; rep < 4; rep++ )
                for( size_t i = 0; i < ntiles; i++ ) {
                    pb2_ce_iov_t l = {big + i * tile, tile}, r = {(void*)(uintptr_t)(peer_addr + ((i * 7) % ntiles) * tile), tile};
                    if( PB2_SUCCESS != pb2_ce_put(ce, peer, &l, 1, &r, 1) ) return 2;
                }
            send_all(ce, peer, 1, buf, 8);
        } else {
            got = 0;
            wait_for(ce, 1);
        }
        pb2_ce_barrier(ce);
        cold_gbs = (double)tile * ntiles * 4 / (now() - t1) * 1e-9;
        char *other = (char*)malloc(tile * ntiles);
        memset(other, 1, tile * ntiles);
        pb2_ce_barrier(ce);
        t1 = now();
        if( 0 == rank ) for( size_t i = 0; i < ntiles; i++ ) memcpy(other + ((i * 7) % ntiles) * tile, big + i * tile, tile);
        cold_memcpy_gbs = (double)tile * ntiles / (now() - t1) * 1e-9;
        pb2_ce_barrier(ce);
        free(other);
        free(big);
    }
    pb2_ce_stats_t st;
    pb2_ce_stats(ce, &st);
    if( 0 == rank )
        printf("{\"am_half_rtt_us_8B\": %.3f, \"am_half_rtt_us_1KiB\": %.3f, \"am_rate_Mmsg_s_64B\": %.2f, "
               "\"put_GBs_1MiB\": %.2f, \"put_GBs_8MiB\": %.2f, \"put_GBs_32MiB\": %.2f, "
               "\"put_GBs_8MiB_distinct_tiles\": %.2f, \"memcpy_GBs_8MiB_distinct_tiles\": %.2f, \"single_copy\": %d}\n",
               lat[0], lat[1], rate, gbs[0], gbs[1], gbs[2], cold_gbs, cold_memcpy_gbs, st.single_copy);
    pb2_ce_barrier(ce);
    pb2_ce_detach(ce);
    if( child ) { int status; waitpid(child, &status, 0); }
    return 0;
}
