// comm_shm.cu -- pb2_ce_*: the transport under the comm-engine seam (SURVEY.md 8, row a22 / piece 3).
//
// What it replaces: the MPI calls inside parsec/parsec_mpi_funnelled.c -- persistent MPI_Recv_init + MPI_Testsome for
// active messages (:621-703, :1103-1154), MPI_Isend/MPI_Irecv pairs behind put/get (:793-991), MPI_Barrier (:1200) --
// for the case the north star names: the ranks of ONE box, one process per GPU, no MPI library.  The reference-side
// half (the parsec_comm_engine_t vtable of parsec/parsec_comm_engine.h:161-182) is overlay/comm_engine_b200.c.
//
// Design (one box, so no NIC and no wire format to respect):
//  * one POSIX shared-memory segment per job; inside it one single-producer / single-consumer byte ring per ordered
//    pair of ranks.  An active message is one record in the ring of (src -> dst): no system call, no matching, no
//    unexpected-message queue; per-pair FIFO order, which is the order MPI gives the reference on one communicator.
//  * put is ONE copy, done by the origin straight into the target's address space (process_vm_writev); the target is
//    not involved until the completion record that follows the data in the same ring.  Where the kernel refuses
//    cross-memory attach, the bytes travel through the ring in chunks and the target's poll lays them down; FIFO order
//    still puts them before the completion record.
//  * barrier and small all-reduce over counters in the segment header.
// Host memory only: the reference's remote-dependency protocol moves host tiles (remote_dep_mpi.c:1593-1726 reads
// PARSEC_DATA_COPY_GET_PTR of the CPU copy; the device module has pushed the tile out before).
#define _GNU_SOURCE 1
#include <errno.h>
#include <fcntl.h>
#include <sched.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/uio.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <mutex>
#include <new>
#include <vector>

#include "parsec_b200.h"

namespace {

constexpr uint64_t kMagic = 0x7062326365763031ull;  // "pb2cev01"
constexpr uint32_t kWrap = 0xffffffffu;
constexpr uint64_t kTagPutChunk = ~0ull;            // internal: payload = {remote address, bytes}
constexpr size_t kAlign = 16;
constexpr int kMaxRanks = 64;
constexpr int kReduceSlots = 16;

struct alignas(64) RingCtl {
    std::atomic<uint64_t> head;  // consumer position (bytes, monotonically increasing)
    char pad0[56];
    std::atomic<uint64_t> tail;  // producer position
    char pad1[56];
};

struct RecHdr {  // 16 bytes
    uint32_t total;    // header + payload, padded to kAlign; kWrap = "continue at offset 0"
    uint32_t payload;  // bytes of payload
    uint64_t tag;
};

struct alignas(64) SegHdr {
    std::atomic<uint64_t> magic;
    std::atomic<int32_t> attached;
    std::atomic<int32_t> detached;
    int32_t world;
    int32_t pad;
    uint64_t ring_bytes;
    std::atomic<int32_t> barrier_count;
    std::atomic<int32_t> barrier_sense;
    std::atomic<int32_t> failed;  // a rank that dies sets it so that the others stop waiting
    int32_t pids[kMaxRanks];
    int64_t reduce[kMaxRanks][kReduceSlots];
};

inline size_t pad_to(size_t v, size_t a) { return (v + a - 1) / a * a; }

double now_s() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

}  // namespace

struct pb2_ce_s {
    int rank = 0, world = 1;
    size_t ring_bytes = 0, seg_bytes = 0;
    char name[128] = {0};
    char* base = nullptr;
    SegHdr* hdr = nullptr;
    int local_sense = 0;
    int cma_ok = 1;  // process_vm_writev usable
    std::mutex send_lock[kMaxRanks];
    pb2_ce_recv_fn wait_fn = nullptr;
    void* wait_user = nullptr;
    uint64_t sent_msgs = 0, recv_msgs = 0, put_bytes = 0, staged_bytes = 0;

    RingCtl* ctl(int src, int dst) const {
        return reinterpret_cast<RingCtl*>(base + pad_to(sizeof(SegHdr), 4096) + ((size_t)src * world + dst) * sizeof(RingCtl));
    }
    char* ring(int src, int dst) const {
        return base + pad_to(sizeof(SegHdr), 4096) + pad_to((size_t)world * world * sizeof(RingCtl), 4096) +
               ((size_t)src * world + dst) * ring_bytes;
    }
};

static size_t segment_bytes(int world, size_t ring_bytes) {
    return pad_to(sizeof(SegHdr), 4096) + pad_to((size_t)world * world * sizeof(RingCtl), 4096) +
           (size_t)world * world * ring_bytes;
}

extern "C" int pb2_ce_attach(const char* job, int rank, int world, size_t ring_bytes, pb2_ce_t** out) {
    if (!job || !out || world < 1 || world > kMaxRanks || rank < 0 || rank >= world) return PB2_ERR_BAD_PARAM;
    if (ring_bytes == 0) ring_bytes = (size_t)1 << 20;
    if (ring_bytes < 4096 || (ring_bytes & (ring_bytes - 1))) return PB2_ERR_BAD_PARAM;  // power of two
    pb2_ce_s* ce = new (std::nothrow) pb2_ce_s();
    if (!ce) return PB2_ERR_OUT_OF_RESOURCE;
    ce->rank = rank;
    ce->world = world;
    ce->ring_bytes = ring_bytes;
    ce->seg_bytes = segment_bytes(world, ring_bytes);
    snprintf(ce->name, sizeof(ce->name), "/pb2ce.%s", job);
    int fd = shm_open(ce->name, O_CREAT | O_RDWR, 0600);
    if (fd < 0) {
        delete ce;
        return PB2_ERR_OUT_OF_RESOURCE;
    }
    if (ftruncate(fd, (off_t)ce->seg_bytes) != 0) {  // every rank asks for the same size; new pages read as zero
        close(fd);
        delete ce;
        return PB2_ERR_OUT_OF_RESOURCE;
    }
    void* p = mmap(nullptr, ce->seg_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) {
        delete ce;
        return PB2_ERR_OUT_OF_RESOURCE;
    }
    ce->base = static_cast<char*>(p);
    ce->hdr = reinterpret_cast<SegHdr*>(p);
    SegHdr* h = ce->hdr;
    uint64_t zero = 0;
    if (h->magic.compare_exchange_strong(zero, kMagic)) {
        h->world = world;
        h->ring_bytes = ring_bytes;
    } else if (zero != kMagic) {
        munmap(p, ce->seg_bytes);
        delete ce;
        return PB2_ERR_BAD_PARAM;  // something else lives under this name
    }
    h->pids[rank] = (int32_t)getpid();
    std::atomic_thread_fence(std::memory_order_seq_cst);
    if (h->attached.fetch_add(1) >= world) {  // more arrivals than ranks: the segment is what a dead job left behind
        fprintf(stderr, "pb2_ce_attach: /dev/shm%s belongs to an earlier job of the same name; remove it or pick another name\n",
                ce->name);
        munmap(p, ce->seg_bytes);
        delete ce;
        return PB2_ERR_BAD_PARAM;
    }
    const double limit = now_s() + 120.0;
    while (h->attached.load() < world) {  // rendezvous: everybody's pid and rings are in place after this
        if (now_s() > limit || h->failed.load()) {
            munmap(p, ce->seg_bytes);
            delete ce;
            return PB2_ERR_NOT_FOUND;
        }
        sched_yield();
    }
    if (h->world != world || h->ring_bytes != ring_bytes) {
        munmap(p, ce->seg_bytes);
        delete ce;
        return PB2_ERR_BAD_PARAM;
    }
    const char* env = getenv("PB2_CE_NO_CMA");
    if (env && atoi(env)) ce->cma_ok = 0;
    *out = ce;
    return PB2_SUCCESS;
}

extern "C" int pb2_ce_detach(pb2_ce_t* ce) {
    if (!ce) return PB2_ERR_BAD_PARAM;
    const int last = ce->hdr->detached.fetch_add(1) + 1 == ce->world;
    munmap(ce->base, ce->seg_bytes);
    if (last) shm_unlink(ce->name);
    delete ce;
    return PB2_SUCCESS;
}

extern "C" int pb2_ce_rank(const pb2_ce_t* ce) { return ce ? ce->rank : -1; }
extern "C" int pb2_ce_world(const pb2_ce_t* ce) { return ce ? ce->world : 0; }
extern "C" size_t pb2_ce_max_message(const pb2_ce_t* ce) { return ce ? ce->ring_bytes / 4 - sizeof(RecHdr) : 0; }
extern "C" void pb2_ce_mark_failed(pb2_ce_t* ce) {
    if (ce) ce->hdr->failed.store(1);
}

// One record into the ring (src = me -> dst).  PB2_HOOK_RETURN_AGAIN when the ring has no room right now.
static int ring_write(pb2_ce_s* ce, int dst, uint64_t tag, const void* a, size_t alen, const void* b, size_t blen) {
    const size_t payload = alen + blen;
    const size_t total = pad_to(sizeof(RecHdr) + payload, kAlign);
    if (total > ce->ring_bytes / 4) return PB2_ERR_VALUE_OUT_OF_BOUNDS;
    RingCtl* c = ce->ctl(ce->rank, dst);
    char* r = ce->ring(ce->rank, dst);
    const uint64_t mask = ce->ring_bytes - 1;
    uint64_t tail = c->tail.load(std::memory_order_relaxed);
    const uint64_t head = c->head.load(std::memory_order_acquire);
    size_t off = (size_t)(tail & mask);
    size_t to_end = ce->ring_bytes - off;
    size_t need = total + (to_end < total ? to_end : 0);  // a record never straddles the end
    if (ce->ring_bytes - (size_t)(tail - head) < need + sizeof(RecHdr)) return PB2_HOOK_RETURN_AGAIN;
    if (to_end < total) {
        RecHdr w{kWrap, 0, 0};
        memcpy(r + off, &w, sizeof(w));  // to_end >= 16 always: positions stay multiples of kAlign
        tail += to_end;
        off = 0;
    }
    RecHdr h{(uint32_t)total, (uint32_t)payload, tag};
    memcpy(r + off, &h, sizeof(h));
    if (alen) memcpy(r + off + sizeof(h), a, alen);
    if (blen) memcpy(r + off + sizeof(h) + alen, b, blen);
    c->tail.store(tail + total, std::memory_order_release);
    return PB2_SUCCESS;
}

extern "C" int pb2_ce_send(pb2_ce_t* ce, int dst, uint64_t tag, const void* msg, size_t len) {
    if (!ce || dst < 0 || dst >= ce->world || (len && !msg) || tag == kTagPutChunk) return PB2_ERR_BAD_PARAM;
    std::lock_guard<std::mutex> g(ce->send_lock[dst]);
    int rc = ring_write(ce, dst, tag, msg, len, nullptr, 0);
    if (rc == PB2_SUCCESS) ce->sent_msgs++;
    return rc;
}

extern "C" int pb2_ce_poll(pb2_ce_t* ce, pb2_ce_recv_fn fn, void* user, int max_msgs) {
    if (!ce || !fn) return PB2_ERR_BAD_PARAM;
    int delivered = 0;
    const uint64_t mask = ce->ring_bytes - 1;
    for (int src = 0; src < ce->world; ++src) {
        RingCtl* c = ce->ctl(src, ce->rank);
        char* r = ce->ring(src, ce->rank);
        uint64_t head = c->head.load(std::memory_order_relaxed);
        while (max_msgs <= 0 || delivered < max_msgs) {
            const uint64_t tail = c->tail.load(std::memory_order_acquire);
            if (head == tail) break;
            size_t off = (size_t)(head & mask);
            RecHdr h;
            memcpy(&h, r + off, sizeof(h));
            if (h.total == kWrap) {
                head += ce->ring_bytes - off;
                c->head.store(head, std::memory_order_release);
                continue;
            }
            const char* payload = r + off + sizeof(RecHdr);
            if (h.tag == kTagPutChunk) {  // staged put: {address in this process, bytes} then the bytes
                uint64_t where[2];
                memcpy(where, payload, sizeof(where));
                memcpy(reinterpret_cast<void*>((uintptr_t)where[0]), payload + sizeof(where), (size_t)where[1]);
            } else {
                // the callback may send (even to src): the record stays valid until head moves, below
                fn(user, src, h.tag, payload, h.payload);
                ++delivered;
                ce->recv_msgs++;
            }
            head += h.total;
            c->head.store(head, std::memory_order_release);
        }
    }
    return delivered;
}

extern "C" void pb2_ce_set_wait_poll(pb2_ce_t* ce, pb2_ce_recv_fn fn, void* user) {
    if (!ce) return;
    ce->wait_fn = fn;
    ce->wait_user = user;
}

// Origin-side copy of local[] into remote[] of rank dst (same total bytes on both sides).
extern "C" int pb2_ce_put(pb2_ce_t* ce, int dst, const pb2_ce_iov_t* local, int nlocal, const pb2_ce_iov_t* remote,
                          int nremote) {
    if (!ce || dst < 0 || dst >= ce->world || nlocal < 0 || nremote < 0 || (nlocal && !local) || (nremote && !remote))
        return PB2_ERR_BAD_PARAM;
    size_t lbytes = 0, rbytes = 0;
    for (int i = 0; i < nlocal; ++i) lbytes += local[i].len;
    for (int i = 0; i < nremote; ++i) rbytes += remote[i].len;
    if (lbytes != rbytes) return PB2_ERR_BAD_PARAM;
    if (lbytes == 0) return PB2_SUCCESS;
    // walk both lists in step, one contiguous piece at a time
    int li = 0, ri = 0;
    size_t lo = 0, ro = 0;
    std::vector<iovec> lv, rv;
    auto flush_cma = [&]() -> int {
        size_t want = 0;
        for (auto& v : lv) want += v.iov_len;
        size_t done = 0;
        size_t l0 = 0, r0 = 0;
        while (done < want) {
            ssize_t n = process_vm_writev((pid_t)ce->hdr->pids[dst], lv.data() + l0, lv.size() - l0, rv.data() + r0,
                                          rv.size() - r0, 0);
            if (n < 0) return -1;
            done += (size_t)n;
            if (done == want) break;
            size_t skip = (size_t)n;  // partial transfer: advance both lists by n bytes
            while (skip) {
                size_t t = lv[l0].iov_len < skip ? lv[l0].iov_len : skip;
                lv[l0].iov_base = (char*)lv[l0].iov_base + t;
                lv[l0].iov_len -= t;
                skip -= t;
                if (!lv[l0].iov_len) ++l0;
            }
            skip = (size_t)n;
            while (skip) {
                size_t t = rv[r0].iov_len < skip ? rv[r0].iov_len : skip;
                rv[r0].iov_base = (char*)rv[r0].iov_base + t;
                rv[r0].iov_len -= t;
                skip -= t;
                if (!rv[r0].iov_len) ++r0;
            }
        }
        lv.clear();
        rv.clear();
        return 0;
    };
    const size_t chunk = ce->ring_bytes / 8;
    while (li < nlocal && ri < nremote) {
        size_t n = local[li].len - lo;
        if (remote[ri].len - ro < n) n = remote[ri].len - ro;
        char* lp = (char*)local[li].base + lo;
        char* rp = (char*)remote[ri].base + ro;
        if (n) {
            if (dst == ce->rank) {
                memmove(rp, lp, n);
            } else if (ce->cma_ok) {
                lv.push_back(iovec{lp, n});
                rv.push_back(iovec{rp, n});
                if (lv.size() == 512) {
                    if (flush_cma() != 0) {
                        if (errno != EPERM && errno != ENOSYS) return PB2_ERROR;
                        ce->cma_ok = 0;  // refused on this box: nothing was written by the failing call; start over staged
                        return pb2_ce_put(ce, dst, local, nlocal, remote, nremote);
                    }
                }
            } else {
                for (size_t o = 0; o < n;) {  // through the ring; the target's poll copies each piece into place
                    size_t m = n - o < chunk ? n - o : chunk;
                    uint64_t where[2] = {(uint64_t)(uintptr_t)(rp + o), (uint64_t)m};
                    int rc;
                    {
                        std::lock_guard<std::mutex> g(ce->send_lock[dst]);
                        rc = ring_write(ce, dst, kTagPutChunk, where, sizeof(where), lp + o, m);
                    }
                    if (rc == PB2_HOOK_RETURN_AGAIN) {
                        if (ce->hdr->failed.load()) return PB2_ERROR;
                        // the target drains its ring from its own progress loop; meanwhile keep ours moving
                        if (ce->wait_fn) pb2_ce_poll(ce, ce->wait_fn, ce->wait_user, 0);
                        else sched_yield();
                        continue;
                    }
                    if (rc != PB2_SUCCESS) return rc;
                    o += m;
                    ce->staged_bytes += m;
                }
            }
        }
        lo += n;
        ro += n;
        if (lo == local[li].len) ++li, lo = 0;
        if (ro == remote[ri].len) ++ri, ro = 0;
    }
    if (!lv.empty() && flush_cma() != 0) {
        if (errno != EPERM && errno != ENOSYS) return PB2_ERROR;
        ce->cma_ok = 0;
        return pb2_ce_put(ce, dst, local, nlocal, remote, nremote);
    }
    ce->put_bytes += lbytes;
    return PB2_SUCCESS;
}

// Origin-side read of remote[] of rank src into local[]: the mirror image of put, one copy, only where the kernel
// allows cross-memory attach (PB2_ERR_NOT_SUPPORTED otherwise: the caller asks the peer to put instead).
extern "C" int pb2_ce_get(pb2_ce_t* ce, int src, const pb2_ce_iov_t* local, int nlocal, const pb2_ce_iov_t* remote,
                          int nremote) {
    if (!ce || src < 0 || src >= ce->world || nlocal < 0 || nremote < 0 || (nlocal && !local) || (nremote && !remote))
        return PB2_ERR_BAD_PARAM;
    size_t lbytes = 0, rbytes = 0;
    for (int i = 0; i < nlocal; ++i) lbytes += local[i].len;
    for (int i = 0; i < nremote; ++i) rbytes += remote[i].len;
    if (lbytes != rbytes) return PB2_ERR_BAD_PARAM;
    if (lbytes == 0) return PB2_SUCCESS;
    if (src != ce->rank && !ce->cma_ok) return PB2_ERR_NOT_SUPPORTED;
    int li = 0, ri = 0;
    size_t lo = 0, ro = 0;
    while (li < nlocal && ri < nremote) {  // one contiguous piece per call: tiles are a few large pieces
        size_t n = local[li].len - lo;
        if (remote[ri].len - ro < n) n = remote[ri].len - ro;
        char* lp = (char*)local[li].base + lo;
        char* rp = (char*)remote[ri].base + ro;
        if (n) {
            if (src == ce->rank) {
                memmove(lp, rp, n);
            } else {
                for (size_t done = 0; done < n;) {
                    iovec l{lp + done, n - done}, r{rp + done, n - done};
                    ssize_t got = process_vm_readv((pid_t)ce->hdr->pids[src], &l, 1, &r, 1, 0);
                    if (got <= 0) {
                        if (done == 0 && li == 0 && lo == 0 && (errno == EPERM || errno == ENOSYS)) {
                            ce->cma_ok = 0;
                            return PB2_ERR_NOT_SUPPORTED;
                        }
                        return PB2_ERROR;
                    }
                    done += (size_t)got;
                }
            }
        }
        lo += n;
        ro += n;
        if (lo == local[li].len) ++li, lo = 0;
        if (ro == remote[ri].len) ++ri, ro = 0;
    }
    ce->put_bytes += lbytes;
    return PB2_SUCCESS;
}

extern "C" int pb2_ce_barrier(pb2_ce_t* ce) {
    if (!ce) return PB2_ERR_BAD_PARAM;
    SegHdr* h = ce->hdr;
    const int sense = ce->local_sense ^= 1;
    if (h->barrier_count.fetch_add(1) + 1 == ce->world) {
        h->barrier_count.store(0);
        h->barrier_sense.store(sense, std::memory_order_release);
    } else {
        while (h->barrier_sense.load(std::memory_order_acquire) != sense) {
            if (h->failed.load()) return PB2_ERROR;
            sched_yield();
        }
    }
    return PB2_SUCCESS;
}

// op: 0 = max, 1 = sum, 2 = min; n <= 16 values, in place
extern "C" int pb2_ce_allreduce_i64(pb2_ce_t* ce, int64_t* vals, int n, int op) {
    if (!ce || !vals || n < 1 || n > kReduceSlots || op < 0 || op > 2) return PB2_ERR_BAD_PARAM;
    SegHdr* h = ce->hdr;
    for (int i = 0; i < n; ++i) h->reduce[ce->rank][i] = vals[i];
    int rc = pb2_ce_barrier(ce);
    if (rc != PB2_SUCCESS) return rc;
    for (int i = 0; i < n; ++i) {
        int64_t acc = h->reduce[0][i];
        for (int r = 1; r < ce->world; ++r) {
            const int64_t v = h->reduce[r][i];
            acc = op == 0 ? (v > acc ? v : acc) : op == 1 ? acc + v : (v < acc ? v : acc);
        }
        vals[i] = acc;
    }
    return pb2_ce_barrier(ce);  // nobody overwrites its slot before everybody has read
}

// every rank contributes `bytes` (<= 128) and receives world * bytes, ordered by rank
extern "C" int pb2_ce_allgather(pb2_ce_t* ce, const void* mine, size_t bytes, void* all) {
    if (!ce || !mine || !all || bytes == 0 || bytes > sizeof(int64_t) * kReduceSlots) return PB2_ERR_BAD_PARAM;
    SegHdr* h = ce->hdr;
    memcpy(h->reduce[ce->rank], mine, bytes);
    int rc = pb2_ce_barrier(ce);
    if (rc != PB2_SUCCESS) return rc;
    for (int r = 0; r < ce->world; ++r) memcpy((char*)all + (size_t)r * bytes, h->reduce[r], bytes);
    return pb2_ce_barrier(ce);
}

extern "C" int pb2_ce_stats(const pb2_ce_t* ce, pb2_ce_stats_t* out) {
    if (!ce || !out) return PB2_ERR_BAD_PARAM;
    out->sent_msgs = ce->sent_msgs;
    out->recv_msgs = ce->recv_msgs;
    out->put_bytes = ce->put_bytes;
    out->staged_bytes = ce->staged_bytes;
    out->single_copy = ce->cma_ok;
    return PB2_SUCCESS;
}
