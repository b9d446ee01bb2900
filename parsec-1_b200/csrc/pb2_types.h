// Internal C++ types behind the opaque handles of include/parsec_b200.h.
#ifndef PB2_TYPES_H
#define PB2_TYPES_H

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <deque>
#include <map>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "parsec_b200.h"

#define PB2_XFER_NONE 0
#define PB2_XFER_UNDER 1
#define PB2_XFER_COMPLETE 2

#define PB2_FLAG_OWNED 0x80   /* PARSEC_DATA_FLAG_PARSEC_OWNED: the module allocated the device memory */
#define PB2_FLAG_MANAGED 0x40 /* PARSEC_DATA_FLAG_PARSEC_MANAGED */

// parsec_data_copy_t (parsec/data_internal.h:57-81)
struct pb2_copy_t {
    pb2_copy_t* prev = nullptr;  // LRU list links (parsec_list_item_t super)
    pb2_copy_t* next = nullptr;
    int list_id = 0;             // 0 = none, 1 = gpu_mem_lru, 2 = gpu_mem_owned_lru
    int8_t device_index = 0;
    uint8_t flags = 0;
    uint8_t coherency_state = PB2_DATA_COHERENCY_INVALID;
    uint8_t data_transfer_status = PB2_XFER_NONE;
    int32_t readers = 0;
    uint32_t version = 0;
    pb2_data_s* original = nullptr;
    void* device_private = nullptr;
    uint64_t ready_seq = 0;      // h2d-stream sequence number after which the device bytes are valid
    cudaEvent_t ready_event = nullptr;  // non-null when validity comes from another stream (NCCL)
};

// parsec_data_t (parsec/data_internal.h:35-55)
struct pb2_data_s {
    std::mutex lock;
    uint64_t key = 0;
    uint32_t dc_id = 0;
    int8_t owner_device = 0;
    int8_t preferred_device = -1;
    uint32_t nb_elts = 0;
    // layout of the HOST tile when it is not packed (PARSEC_MATRIX_LAPACK storage: a window of a column-major local
    // matrix): `host_height` runs of `host_width` bytes, `host_pitch` bytes apart; on the device the tile is packed with
    // `dev_pitch` bytes per column.  host_pitch == 0: contiguous, nb_elts bytes (PARSEC_MATRIX_TILE).
    uint32_t host_width = 0, host_height = 0;
    size_t host_pitch = 0, dev_pitch = 0;
    pb2_copy_t host_copy;
    pb2_copy_t* device_copies[PB2_MAX_DEVICES] = {nullptr};
    void* user = nullptr;  // pb2_data_set_user: the overlay shim keeps its parsec_data_copy_t here
};

namespace pb2 {

// data.c:319-433 / :299-317 -- implemented in data.cu
int data_start_transfer_ownership(pb2_data_s* data, int device, uint8_t access);
void data_end_transfer_ownership(pb2_data_s* data, int device, uint8_t access);

// HBM heap over one cudaMalloc, in units of `unit` bytes.  Plays the part of parsec/utils/zone_malloc.c for the device
// module (one allocation carved into tiles) with a policy of its own, built for what a tile cache does: nearly every
// request is one of one or two tile sizes, and a freed tile is wanted again at once by the task that evicted it.
//   * free(p) pushes the extent on a LIFO stack of extents of exactly its size: O(1), no merging;
//   * alloc(size) pops that stack: O(1); if it is empty the extent is carved from the lowest-addressed run of the
//     coalesced free space (an ordered map offset -> length);
//   * only when that fails too are the stacks emptied into the map, where neighbours merge, and the carve retried --
//     the one place where extents of different sizes meet (a taskpool with other tile sizes has started).
// in_use() is a counter.
class ZoneHeap {
public:
    ZoneHeap(char* base, int nb_units, size_t unit);
    void* alloc(size_t size);
    bool free(void* p);
    size_t in_use() const { return used_units_ * unit_; }
    size_t capacity() const { return unit_ * (size_t)nunits_; }
    char* base() const { return base_; }
    int64_t offset_of(const void* p) const { return (const char*)p - base_; }

private:
    bool carve(uint32_t want, uint32_t* at);
    void merge_in(uint32_t at, uint32_t len);
    char* base_;
    uint32_t nunits_;
    size_t unit_;
    size_t used_units_ = 0;
    std::vector<uint32_t> len_at_;                                // units of the live extent starting here, 0 = none
    std::map<uint32_t, uint32_t> runs_;                           // coalesced free space: first unit -> units
    std::unordered_map<uint32_t, std::vector<uint32_t>> parked_;  // freed extents by size, not merged yet
};

struct CopyList {  // parsec_list_t of data copies, LRU order: front = oldest
    pb2_copy_t* head = nullptr;
    pb2_copy_t* tail = nullptr;
    int id;
    size_t count = 0;
    explicit CopyList(int id_) : id(id_) {}
    void push_back(pb2_copy_t* c);
    void push_front(pb2_copy_t* c);
    pb2_copy_t* pop_front();
    static void chop(pb2_copy_t* c, CopyList* lists[3]);
};

}  // namespace pb2

#endif
