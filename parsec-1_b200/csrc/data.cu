// Data masters, per-device copies, the ownership/coherency state machine, the HBM zone heap and the
// LRU lists.  Host-only code (compiled by nvcc with the rest of the library).
//
// Behavioural contract kept from the reference (the generated code and the graded version sequences
// depend on it):  parsec/data.c:484-522 (create), :319-433 (start of an ownership transfer), :299-317 (end).
// The heap only shares its job with parsec/utils/zone_malloc.c (carve one device allocation); its policy is its own.
#include <stdlib.h>
#include <string.h>

#include <iterator>

#include "pb2_internal.h"
#include "pb2_types.h"

namespace pb2 {

static inline pb2_copy_t* copy_of(pb2_data_s* d, int dev) { return dev == 0 ? &d->host_copy : d->device_copies[dev]; }

int data_start_transfer_ownership(pb2_data_s* data, int device, uint8_t access) {
    pb2_copy_t* mine = copy_of(data, device);
    bool fetch = false;
    int source = data->owner_device;

    if (mine->coherency_state == PB2_DATA_COHERENCY_INVALID) {
        fetch = true;
        if (source == -1) {
            for (int d = 0; d < PB2_MAX_DEVICES; ++d) {
                pb2_copy_t* c = copy_of(data, d);
                if (c && c->coherency_state != PB2_DATA_COHERENCY_INVALID) source = d;
            }
        }
    } else if (mine->coherency_state == PB2_DATA_COHERENCY_SHARED) {
        for (int d = 0; d < PB2_MAX_DEVICES; ++d) {
            pb2_copy_t* c = copy_of(data, d);
            if (c && c->coherency_state == PB2_DATA_COHERENCY_OWNED && c->version > mine->version) fetch = true;
        }
    }
    const bool rd = access & PB2_FLOW_ACCESS_READ, wr = access & PB2_FLOW_ACCESS_WRITE;
    if (rd) {
        for (int d = 0; d < PB2_MAX_DEVICES; ++d) {
            pb2_copy_t* c = copy_of(data, d);
            if (d == device || !c || c->coherency_state == PB2_DATA_COHERENCY_INVALID) continue;
            if (mine->coherency_state == PB2_DATA_COHERENCY_OWNED && !wr) {
                if (c->version < mine->version) c->coherency_state = PB2_DATA_COHERENCY_INVALID;
                data->owner_device = -1;
            }
            if (c->coherency_state == PB2_DATA_COHERENCY_EXCLUSIVE) c->coherency_state = PB2_DATA_COHERENCY_SHARED;
        }
    } else {
        fetch = false;  // pure write: content will be overwritten
    }
    if (wr) {
        for (int d = 0; d < PB2_MAX_DEVICES; ++d) {
            pb2_copy_t* c = copy_of(data, d);
            if (c && c->coherency_state != PB2_DATA_COHERENCY_INVALID) c->coherency_state = PB2_DATA_COHERENCY_SHARED;
        }
    }
    if (rd) mine->readers++;
    if (wr) data->owner_device = (int8_t)device;
    return fetch ? source : -1;
}

void data_end_transfer_ownership(pb2_data_s* data, int device, uint8_t access) {
    pb2_copy_t* mine = copy_of(data, device);
    if (access & PB2_FLOW_ACCESS_READ) mine->coherency_state = PB2_DATA_COHERENCY_SHARED;
    if (access & PB2_FLOW_ACCESS_WRITE) mine->coherency_state = PB2_DATA_COHERENCY_OWNED;
}

// ------------------------------------------------------------------------------------------ heap
ZoneHeap::ZoneHeap(char* base, int nb_units, size_t unit)
    : base_(base), nunits_((uint32_t)(nb_units > 0 ? nb_units : 0)), unit_(unit), len_at_((size_t)nunits_, 0u) {
    if (nunits_) runs_[0] = nunits_;
}

// lowest-addressed run that is long enough; the remainder stays a run
bool ZoneHeap::carve(uint32_t want, uint32_t* at) {
    for (auto it = runs_.begin(); it != runs_.end(); ++it) {
        if (it->second < want) continue;
        const uint32_t start = it->first, len = it->second;
        runs_.erase(it);
        if (len > want) runs_[start + want] = len - want;
        *at = start;
        return true;
    }
    return false;
}

// put a free extent into the map, merging it with the runs that touch it
void ZoneHeap::merge_in(uint32_t at, uint32_t len) {
    auto next = runs_.lower_bound(at);
    if (next != runs_.begin()) {
        auto prev = std::prev(next);
        if (prev->first + prev->second == at) {
            at = prev->first;
            len += prev->second;
            runs_.erase(prev);
        }
    }
    if (next != runs_.end() && at + len == next->first) {
        len += next->second;
        runs_.erase(next);
    }
    runs_[at] = len;
}

void* ZoneHeap::alloc(size_t size) {
    if (size == 0 || size > capacity()) return nullptr;
    const uint32_t want = (uint32_t)((size + unit_ - 1) / unit_);
    uint32_t at = 0;
    auto parked = parked_.find(want);
    if (parked != parked_.end() && !parked->second.empty()) {
        at = parked->second.back();
        parked->second.pop_back();
    } else if (!carve(want, &at)) {
        bool any = false;
        for (auto& kv : parked_) {
            for (uint32_t off : kv.second) merge_in(off, kv.first);
            any = any || !kv.second.empty();
            kv.second.clear();
        }
        if (!any || !carve(want, &at)) return nullptr;
    }
    len_at_[at] = want;
    used_units_ += want;
    return base_ + (size_t)at * unit_;
}

bool ZoneHeap::free(void* p) {
    const int64_t off = (char*)p - base_;
    if (off < 0 || off % (int64_t)unit_ || (uint64_t)off / unit_ >= nunits_) return false;
    const uint32_t at = (uint32_t)((uint64_t)off / unit_);
    const uint32_t len = len_at_[at];
    if (!len) return false;  // not the start of a live extent (double free included)
    len_at_[at] = 0;
    used_units_ -= len;
    parked_[len].push_back(at);
    return true;
}

// ------------------------------------------------------------------------------------------ lists
void CopyList::push_back(pb2_copy_t* c) {
    c->prev = tail;
    c->next = nullptr;
    if (tail) tail->next = c; else head = c;
    tail = c;
    c->list_id = id;
    ++count;
}
void CopyList::push_front(pb2_copy_t* c) {
    c->next = head;
    c->prev = nullptr;
    if (head) head->prev = c; else tail = c;
    head = c;
    c->list_id = id;
    ++count;
}
pb2_copy_t* CopyList::pop_front() {
    pb2_copy_t* c = head;
    if (!c) return nullptr;
    head = c->next;
    if (head) head->prev = nullptr; else tail = nullptr;
    c->prev = c->next = nullptr;
    c->list_id = 0;
    --count;
    return c;
}
void CopyList::chop(pb2_copy_t* c, CopyList* lists[3]) {
    if (!c->list_id) return;
    CopyList* l = lists[c->list_id];
    if (c->prev) c->prev->next = c->next; else l->head = c->next;
    if (c->next) c->next->prev = c->prev; else l->tail = c->prev;
    c->prev = c->next = nullptr;
    c->list_id = 0;
    --l->count;
}

}  // namespace pb2

// ------------------------------------------------------------------------------------------ C ABI
extern "C" pb2_data_t* pb2_data_create(uint64_t key, uint32_t dc_id, void* host_ptr, uint32_t nb_elts) {
    pb2_data_s* d = new pb2_data_s();
    d->key = key;
    d->dc_id = dc_id;
    d->nb_elts = nb_elts;
    d->owner_device = 0;
    d->preferred_device = -1;
    d->host_copy.device_index = 0;
    d->host_copy.original = d;
    d->host_copy.coherency_state = PB2_DATA_COHERENCY_OWNED;
    d->host_copy.device_private = host_ptr;
    d->host_copy.flags = PB2_FLAG_MANAGED;
    return d;
}

// A host tile that is a window of a larger column-major matrix (what stage_custom.jdf:28-109 handles with its own
// stage functions and dc->llm): the default stage-in / stage-out and the write-back of evicted tiles then move
// `height` columns of `width_bytes` with cudaMemcpy2DAsync, host pitch `host_pitch_bytes`, device pitch `dev_pitch_bytes`.
extern "C" int pb2_data_set_host_layout(pb2_data_t* data, uint32_t width_bytes, uint32_t height, size_t host_pitch_bytes,
                                        size_t dev_pitch_bytes) {
    if (!data) return PB2_ERR_BAD_PARAM;
    if (host_pitch_bytes == 0) {  // back to a packed tile
        data->host_width = data->host_height = 0;
        data->host_pitch = data->dev_pitch = 0;
        return PB2_SUCCESS;
    }
    if (!width_bytes || !height || host_pitch_bytes < width_bytes || dev_pitch_bytes < width_bytes ||
        (uint64_t)dev_pitch_bytes * height > data->nb_elts)
        return PB2_ERR_BAD_PARAM;
    data->host_width = width_bytes;
    data->host_height = height;
    data->host_pitch = host_pitch_bytes;
    data->dev_pitch = dev_pitch_bytes;
    return PB2_SUCCESS;
}

extern "C" int pb2_data_copy_version(const pb2_data_t* data, int dev, uint32_t* version) {
    if (!data || dev < 0 || dev >= PB2_MAX_DEVICES) return PB2_ERR_BAD_PARAM;
    const pb2_copy_t* c = dev == 0 ? &data->host_copy : data->device_copies[dev];
    if (!c) return PB2_ERR_NOT_FOUND;
    if (version) *version = c->version;
    return PB2_SUCCESS;
}
extern "C" int pb2_data_copy_coherency(const pb2_data_t* data, int dev) {
    if (!data || dev < 0 || dev >= PB2_MAX_DEVICES) return PB2_ERR_BAD_PARAM;
    const pb2_copy_t* c = dev == 0 ? &data->host_copy : data->device_copies[dev];
    return c ? c->coherency_state : PB2_ERR_NOT_FOUND;
}
extern "C" int pb2_data_copy_readers(const pb2_data_t* data, int dev) {
    if (!data || dev < 0 || dev >= PB2_MAX_DEVICES) return PB2_ERR_BAD_PARAM;
    const pb2_copy_t* c = dev == 0 ? &data->host_copy : data->device_copies[dev];
    return c ? c->readers : PB2_ERR_NOT_FOUND;
}
extern "C" int pb2_data_owner_device(const pb2_data_t* data) { return data ? data->owner_device : PB2_ERR_BAD_PARAM; }
extern "C" void* pb2_data_copy_ptr(const pb2_data_t* data, int dev) {
    if (!data || dev < 0 || dev >= PB2_MAX_DEVICES) return nullptr;
    const pb2_copy_t* c = dev == 0 ? &data->host_copy : data->device_copies[dev];
    return c ? c->device_private : nullptr;
}

// ------------------------------------------------------------------------------------------ overlay support
// What the shim that stands in for device_cuda_module.c inside a PaRSEC build (overlay/device_cuda_b200_shim.c) needs
// to keep a parsec_data_t and its pb2 mirror in step.
extern "C" void pb2_data_set_user(pb2_data_t* data, void* user) {
    if (data) data->user = user;
}
extern "C" void* pb2_data_user(const pb2_data_t* data) { return data ? data->user : nullptr; }

// The host side of a tile is owned by the runtime: before a task is handed to the module the shim copies what the
// runtime's host copy says (address, version, and OWNED when a CPU task wrote it since) into the mirror.
extern "C" int pb2_data_sync_host(pb2_data_t* data, void* host_ptr, uint32_t version, int coherency) {
    if (!data) return PB2_ERR_BAD_PARAM;
    std::lock_guard<std::mutex> g(data->lock);
    data->host_copy.device_private = host_ptr;
    data->host_copy.version = version;
    if (coherency >= 0 && coherency != data->host_copy.coherency_state) {
        data->host_copy.coherency_state = (uint8_t)coherency;
        if (coherency == PB2_DATA_COHERENCY_OWNED) {
            // a CPU writer took the tile: every device copy that is not in use falls back to SHARED (data.c:398-404)
            data->owner_device = 0;
            for (int d = 1; d < PB2_MAX_DEVICES; ++d)
                if (data->device_copies[d] && data->device_copies[d]->coherency_state == PB2_DATA_COHERENCY_OWNED &&
                    data->device_copies[d]->readers == 0)
                    data->device_copies[d]->coherency_state = PB2_DATA_COHERENCY_SHARED;
        }
    }
    return PB2_SUCCESS;
}

extern "C" int pb2_data_get_state(const pb2_data_t* data, int dev, pb2_data_state_t* out) {
    if (!data || !out || dev < 1 || dev >= PB2_MAX_DEVICES) return PB2_ERR_BAD_PARAM;
    const pb2_copy_t* c = data->device_copies[dev];
    out->has_copy = c != nullptr;
    out->dev_ptr = c ? c->device_private : nullptr;
    out->dev_version = c ? c->version : 0;
    out->dev_coherency = c ? c->coherency_state : (uint8_t)PB2_DATA_COHERENCY_INVALID;
    out->readers = c ? c->readers : 0;
    out->host_version = data->host_copy.version;
    out->host_coherency = data->host_copy.coherency_state;
    out->owner_device = data->owner_device;
    return PB2_SUCCESS;
}

// ------------------------------------------------------------------------------------------ state-machine hooks
// parsec_data_copy_attach / _detach, parsec_data_start_transfer_ownership_to_copy, parsec_data_end_transfer_...
// (parsec/data.c:163-202, :299-433) on bare copies, without a device: lets a CPU test drive the product's coherency
// state machine with the operation sequences recorded from the reference's own data.c (tests/golden/ownership_trace.json).
extern "C" int pb2_data_test_attach(pb2_data_t* data, int dev) {
    if (!data || dev < 1 || dev >= PB2_MAX_DEVICES || data->device_copies[dev]) return PB2_ERR_BAD_PARAM;
    pb2_copy_t* c = new pb2_copy_t();
    c->device_index = (int8_t)dev;
    c->original = data;
    c->coherency_state = PB2_DATA_COHERENCY_INVALID;
    data->device_copies[dev] = c;
    return PB2_SUCCESS;
}
extern "C" int pb2_data_test_detach(pb2_data_t* data, int dev) {
    if (!data || dev < 1 || dev >= PB2_MAX_DEVICES || !data->device_copies[dev]) return PB2_ERR_NOT_FOUND;
    delete data->device_copies[dev];
    data->device_copies[dev] = nullptr;
    return PB2_SUCCESS;
}
extern "C" int pb2_data_test_start_transfer(pb2_data_t* data, int dev, int access) {
    return pb2::data_start_transfer_ownership(data, dev, (uint8_t)access);
}
extern "C" void pb2_data_test_end_transfer(pb2_data_t* data, int dev, int access) {
    pb2::data_end_transfer_ownership(data, dev, (uint8_t)access);
}
extern "C" void pb2_data_test_set(pb2_data_t* data, int dev, int coherency, int version, int readers) {
    pb2_copy_t* c = dev == 0 ? &data->host_copy : data->device_copies[dev];
    if (!c) return;
    if (coherency >= 0) c->coherency_state = (uint8_t)coherency;
    if (version >= 0) c->version = (uint32_t)version;
    if (readers >= 0) c->readers = readers;
}
