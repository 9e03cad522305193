// mt_table.cu -- host runtime of libmt_b200.so: error state, validation, tile policy and the
// device-resident tensor-table cache.
//
// The reference re-packs pointers and sizes into a ~14 KB by-value kernel argument for EVERY launch of EVERY
// call (/root/reference/csrc/multi_tensor_apply.cuh:84-143; ~383 launches for a Llama-3-8B Adam step,
// >=1042 for 50k small tensors).  Here the table lives in device memory, is built once per distinct
// (pointer set, sizes, tile size) and is found again by hash, so a steady-state optimizer step uploads nothing
// and issues ONE launch per op.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "mt_host.h"

namespace mt {

std::atomic<uint64_t> g_launches{0};

static thread_local std::string tls_error;

int set_error(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    tls_error = buf;
    return code;
}

int device_info(int device, DeviceInfo* out) {
    static DeviceInfo info[64];
    static bool have[64];
    static std::mutex mu;
    if (device < 0 || device >= 64) return set_error(MT_EINVAL, "device ordinal %d out of range", device);
    std::lock_guard<std::mutex> lk(mu);
    if (!have[device]) {
        int sms = 0;
        MT_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        info[device].sms = sms;
        have[device] = true;
    }
    *out = info[device];
    return MT_OK;
}

int validate_lists(const MtTensorLists* tl, int lo, int hi, const char* who) {
    if (!tl) return set_error(MT_EINVAL, "%s: tensor lists are NULL", who);
    if (tl->depth < lo || tl->depth > hi) return set_error(MT_EINVAL, "%s: tensor_lists.size() != depth (got %d)", who, tl->depth);
    if (tl->ntensors <= 0) return set_error(MT_EINVAL, "%s: tensor_lists[0].size() is not > 0", who);
    if (tl->ntensors > 0x7fffffff) return set_error(MT_EINVAL, "%s: too many tensors", who);
    if (!tl->ptrs || !tl->numel) return set_error(MT_EINVAL, "%s: ptrs / numel arrays are NULL", who);
    for (int d = 0; d < tl->depth; d++)
        if (tl->dtype[d] < MT_F32 || tl->dtype[d] > MT_E5M2)
            return set_error(MT_EINVAL, "%s: list %d has unknown dtype tag %d", who, d, tl->dtype[d]);
    return MT_OK;
}

// Per-entry checks.  O(ntensors x depth): run when a table is BUILT (a pointer set that hits the cache was checked when
// it was inserted), and by callers that build their own table.
int validate_entries(const MtTensorLists* tl, const char* who) {
    for (int64_t t = 0; t < tl->ntensors; t++) {
        if (tl->numel[t] < 0) return set_error(MT_EINVAL, "%s: tensor %lld has negative numel", who, (long long)t);
        if (tl->numel[t] > 0)
            for (int d = 0; d < tl->depth; d++)
                if (!tl->ptrs[(size_t)d * tl->ntensors + t])
                    return set_error(MT_EINVAL, "%s: list %d tensor %lld is a NULL pointer", who, d, (long long)t);
    }
    return MT_OK;
}

int64_t total_elems(const MtTensorLists* tl) {
    int64_t s = 0;
    for (int64_t t = 0; t < tl->ntensors; t++) s += tl->numel[t];
    return s;
}

uint64_t count_tiles(const MtTensorLists* tl, int tile_elems) {
    uint64_t n = 0;
    for (int64_t t = 0; t < tl->ntensors; t++) n += (uint64_t)((tl->numel[t] + tile_elems - 1) / tile_elems);
    return n;
}

int choose_tile_elems(const MtTensorLists* tl, int device) {
    // MT_TILE_ELEMS: tuning override (multiple of 2048)
    static const int forced = [] {
        const char* e = getenv("MT_TILE_ELEMS");
        return e ? atoi(e) : 0;
    }();
    if (forced >= kRound && forced % kRound == 0) return forced;
    DeviceInfo di;
    if (device_info(device, &di) != MT_OK || di.sms <= 0) di.sms = 148;
    // want >= 4 waves of (4 CTAs/SM) so that the tail of the last wave is small; one pass counts the tiles for every
    // candidate size (16384, 8192, 4096)
    const uint64_t want = (uint64_t)di.sms * 4 * 4;
    uint64_t cnt[3] = {0, 0, 0};
    for (int64_t t = 0; t < tl->ntensors; t++) {
        const uint64_t n = (uint64_t)tl->numel[t];
        cnt[0] += (n + 16383) >> 14;
        cnt[1] += (n + 8191) >> 13;
        cnt[2] += (n + 4095) >> 12;
        if (cnt[0] >= want) return 16384;  // the common case ends after a few tensors
    }
    static_assert(kMaxTile == 16384 && kRound == 2048, "candidate tile sizes are hard-wired below");
    if (cnt[0] >= want) return 16384;
    if (cnt[1] >= want) return 8192;
    if (cnt[2] >= want) return 4096;
    return kRound;
}

// ---- table cache --------------------------------------------------------------------------------
//
// Ordering rules (all under the per-device mutex; none of them ever synchronises the device):
//   * upload   every Entry owns an event recorded right behind its cudaMemcpyAsync.  A hit from ANOTHER stream makes
//              that stream wait on the event (cudaStreamWaitEvent) until the event has been seen complete once; the
//              uploading stream itself is ordered by the stream.  A capturing stream cannot wait on an event recorded
//              outside its capture, so there the host waits for the upload instead (once).
//   * last use TableRef's destructor records a per-(entry, stream) event behind the call's last launch (kUseSlots
//              slots; when more distinct streams than slots use one table, the newcomer is first made to wait on the
//              slot it takes over, so that its own event covers the displaced use transitively).
//   * eviction least-recently-used first, and only entries whose upload and use events have all completed; an entry
//              that is still in flight (or referenced by a stream capture) is skipped, the cap is soft.  Blobs come
//              from the stream-ordered allocator (cudaMallocAsync / cudaFreeAsync): unlike cudaMalloc / cudaFree these
//              neither synchronise the device nor count as unsafe calls while another thread captures a graph.
namespace {

constexpr int kUseSlots = 4;

// Event queries / waits and pinned allocations are "potentially unsafe" calls: while a stream capture in global mode is
// open (on this thread or any other) the runtime refuses them AND invalidates the capture.  None of the calls made
// here touches captured work, so the cache's slow paths run with this thread's capture-interaction mode relaxed.
struct RelaxedCaptureMode {
    cudaStreamCaptureMode mode = cudaStreamCaptureModeRelaxed;
    RelaxedCaptureMode() { cudaThreadExchangeStreamCaptureMode(&mode); }
    ~RelaxedCaptureMode() { cudaThreadExchangeStreamCaptureMode(&mode); }
};

struct Entry {
    std::vector<uint8_t> host;  // exact blob, for collision-proof comparison
    void* dev = nullptr;
    size_t bytes = 0;
    uint32_t ntiles = 0;
    bool in_graph = false;  // referenced by a stream capture: never evicted
    uint64_t tick = 0;      // LRU clock value of the last hit
    cudaEvent_t upload_ev = nullptr;
    cudaStream_t upload_stream = nullptr;
    bool upload_done = false;
    cudaStream_t use_stream[kUseSlots] = {};
    cudaEvent_t use_ev[kUseSlots] = {};
    bool use_live[kUseSlots] = {};
    int use_next = 0;  // round-robin victim when every slot belongs to another stream
    int leased = 0;    // calls between get_table() and TableRef::release(): their launches are not covered by an event yet
};

struct DeviceCache {
    std::mutex mu;
    std::unordered_multimap<uint64_t, Entry> map;
    // Most-recently-used entries, checked before anything is hashed: a training loop cycles through a handful of
    // (op, lists) pairs, and comparing the caller's pointer / numel arrays straight against a cached blob is one
    // memcmp pass instead of blob build (a 64-bit division per tensor) + hash + memcmp.  Pointers into `map` nodes
    // (stable until erased; an evicted entry is removed from this list first).
    static constexpr int kMru = 16;
    Entry* mru[kMru] = {};
    int mru_n = 0;
    void* staging = nullptr;  // pinned
    size_t staging_cap = 0;
    cudaEvent_t staging_ev = nullptr;
    bool staging_pending = false;
    uint64_t clock = 0;
    MtCacheStats stats{};
};

DeviceCache g_cache[64];

// Soft caps (MT_TABLE_CACHE_MAX_ENTRIES / MT_TABLE_CACHE_MAX_BYTES override them; the tests use tiny values).
size_t env_or(const char* name, size_t dflt) {
    const char* e = getenv(name);
    if (!e || !*e) return dflt;
    const long long v = atoll(e);
    return v > 0 ? (size_t)v : dflt;
}
size_t max_entries() { static const size_t v = env_or("MT_TABLE_CACHE_MAX_ENTRIES", 1024); return v; }
size_t max_device_bytes() { static const size_t v = env_or("MT_TABLE_CACHE_MAX_BYTES", 1ull << 30); return v; }

inline uint64_t mix64(uint64_t h, uint64_t v) {
    h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
    h *= 0xff51afd7ed558ccdull;
    return h ^ (h >> 32);
}

uint64_t hash_blob(const uint8_t* p, size_t n) {
    // four independent lanes: the multiply chain of one lane would make a 1.8 MB table (50k tensors) cost ~0.5 ms
    uint64_t h[4] = {0x243f6a8885a308d3ull ^ n, 0x13198a2e03707344ull, 0xa4093822299f31d0ull, 0x082efa98ec4e6c89ull};
    size_t i = 0;
    for (; i + 32 <= n; i += 32) {
        uint64_t v[4];
        memcpy(v, p + i, 32);
        h[0] = mix64(h[0], v[0]);
        h[1] = mix64(h[1], v[1]);
        h[2] = mix64(h[2], v[2]);
        h[3] = mix64(h[3], v[3]);
    }
    for (; i + 8 <= n; i += 8) {
        uint64_t v;
        memcpy(&v, p + i, 8);
        h[0] = mix64(h[0], v);
    }
    uint64_t v = 0;
    if (i < n) memcpy(&v, p + i, n - i);
    h[0] = mix64(h[0], v);
    return mix64(mix64(h[0], h[1]), mix64(h[2], h[3]));
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

inline void mru_drop(DeviceCache& c, const Entry* e) {
    int w = 0;
    for (int i = 0; i < c.mru_n; i++)
        if (c.mru[i] != e) c.mru[w++] = c.mru[i];
    c.mru_n = w;
}

// Has every piece of GPU work that may still read this entry's blob finished?  (Queries only -- never blocks.)
bool entry_idle(Entry& e) {
    if (e.leased > 0) return false;
    if (!e.upload_done) {
        if (cudaEventQuery(e.upload_ev) != cudaSuccess) { (void)cudaGetLastError(); return false; }
        e.upload_done = true;
    }
    for (int i = 0; i < kUseSlots; i++)
        if (e.use_live[i]) {
            if (cudaEventQuery(e.use_ev[i]) != cudaSuccess) { (void)cudaGetLastError(); return false; }
            e.use_live[i] = false;
        }
    return true;
}

void destroy_entry(DeviceCache& c, Entry& e, cudaStream_t free_stream) {
    mru_drop(c, &e);
    if (e.dev) cudaFreeAsync(e.dev, free_stream);
    if (e.upload_ev) cudaEventDestroy(e.upload_ev);
    for (int i = 0; i < kUseSlots; i++)
        if (e.use_ev[i]) cudaEventDestroy(e.use_ev[i]);
    c.stats.device_bytes -= e.bytes;
}

// Make room for `incoming` more bytes: drop idle entries, oldest first, until the cache is back under 3/4 of its caps.
// Entries that are in flight or captured stay; if nothing can go the caps are simply exceeded for a while.
void evict_lru_locked(DeviceCache& c, size_t incoming, cudaStream_t stream) {
    const size_t want_entries = max_entries() * 3 / 4, want_bytes = max_device_bytes() * 3 / 4;
    std::vector<std::pair<uint64_t, decltype(c.map)::iterator>> order;
    order.reserve(c.map.size());
    for (auto it = c.map.begin(); it != c.map.end(); ++it)
        if (!it->second.in_graph) order.emplace_back(it->second.tick, it);
    std::sort(order.begin(), order.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    for (auto& o : order) {
        if (c.map.size() < want_entries && c.stats.device_bytes + incoming <= want_bytes) break;
        Entry& e = o.second->second;
        if (!entry_idle(e)) continue;
        destroy_entry(c, e, stream);
        c.map.erase(o.second);
        c.stats.evictions++;
    }
    c.stats.entries = c.map.size();
}

// Everything goes (mt_table_cache_clear: an explicit request, documented to wait for the device).
int evict_all_locked(DeviceCache& c) {
    MT_CUDA_OK(cudaDeviceSynchronize());
    for (auto it = c.map.begin(); it != c.map.end();) {
        if (it->second.in_graph) { ++it; continue; }
        destroy_entry(c, it->second, nullptr);
        it = c.map.erase(it);
    }
    c.stats.entries = c.map.size();
    return MT_OK;
}

inline void mru_touch(DeviceCache& c, Entry* e) {
    int pos = c.mru_n;
    for (int i = 0; i < c.mru_n; i++)
        if (c.mru[i] == e) { pos = i; break; }
    if (pos == c.mru_n) {
        if (c.mru_n < DeviceCache::kMru) c.mru_n++;
        pos = c.mru_n - 1;
    }
    for (int i = pos; i > 0; i--) c.mru[i] = c.mru[i - 1];
    c.mru[0] = e;
}

// A cache hit on `stream`: order the stream behind the upload, pick the use slot this call will record into.
int acquire_locked(DeviceCache& c, Entry* e, cudaStream_t stream, bool capturing, int* slot_out) {
    e->tick = ++c.clock;
    *slot_out = -1;
    if (!e->upload_done) {
        RelaxedCaptureMode relaxed;
        if (cudaEventQuery(e->upload_ev) == cudaSuccess) {
            e->upload_done = true;
        } else {
            (void)cudaGetLastError();
            if (capturing) {  // an event recorded outside the capture cannot be waited on inside it
                MT_CUDA_OK(cudaEventSynchronize(e->upload_ev));
                e->upload_done = true;
            } else if (stream != e->upload_stream) {
                MT_CUDA_OK(cudaStreamWaitEvent(stream, e->upload_ev, 0));
                c.stats.cross_stream_waits++;
            }
        }
    }
    if (capturing) {
        e->in_graph = true;  // pinned for good: replays may come at any time
        return MT_OK;
    }
    int slot = -1;
    for (int i = 0; i < kUseSlots && slot < 0; i++)
        if (e->use_ev[i] && e->use_stream[i] == stream) slot = i;
    for (int i = 0; i < kUseSlots && slot < 0; i++)
        if (!e->use_ev[i]) slot = i;
    if (slot < 0) {  // every slot belongs to another stream: take one over, inheriting what it guards
        RelaxedCaptureMode relaxed;
        for (int k = 0; k < kUseSlots && slot < 0; k++) {
            const int i = (e->use_next + k) % kUseSlots;
            if (!e->use_live[i] || cudaEventQuery(e->use_ev[i]) == cudaSuccess) { e->use_live[i] = false; slot = i; }
        }
        (void)cudaGetLastError();
        if (slot < 0) {
            slot = e->use_next;
            MT_CUDA_OK(cudaStreamWaitEvent(stream, e->use_ev[slot], 0));
            c.stats.cross_stream_waits++;
        }
        e->use_next = (slot + 1) % kUseSlots;
    }
    if (!e->use_ev[slot]) MT_CUDA_OK(cudaEventCreateWithFlags(&e->use_ev[slot], cudaEventDisableTiming));
    e->use_stream[slot] = stream;
    e->leased++;
    *slot_out = slot;
    return MT_OK;
}

inline void fill_out(Entry* e, int device, cudaStream_t stream, int slot, size_t nt, int depth, int tile_elems,
                     size_t off_tb, size_t off_numel, size_t off_ptrs, TableRef* out) {
    const uint8_t* d = static_cast<const uint8_t*>(e->dev);
    out->inl = 0;
    if (nt <= (size_t)DevTable::kInlineTensors && depth <= DevTable::kInlineLists) {
        // small table: a copy rides in the kernel arguments (DevTable::inl), taken from the entry's exact host blob
        const uint8_t* h = e->host.data();
        out->inl = 1;
        memset(out->inl_tile_begin, 0, sizeof(out->inl_tile_begin));
        memset(out->inl_numel, 0, sizeof(out->inl_numel));
        memset(out->inl_ptrs, 0, sizeof(out->inl_ptrs));
        memcpy(out->inl_tile_begin, h + off_tb, (nt + 1) * 4);
        memcpy(out->inl_numel, h + off_numel, nt * 8);
        for (int l = 0; l < depth; l++) memcpy(out->inl_ptrs[l], h + off_ptrs + (size_t)l * nt * 8, nt * 8);
    }
    out->tile_begin = reinterpret_cast<const uint32_t*>(d + off_tb);
    out->numel = reinterpret_cast<const int64_t*>(d + off_numel);
    out->ptrs = reinterpret_cast<const void* const*>(d + off_ptrs);
    out->ntensors = (int32_t)nt;
    out->ntiles = e->ntiles;
    out->tile_elems = tile_elems;
    out->tickets = reinterpret_cast<uint32_t*>(const_cast<uint8_t*>(d) + align_up(e->bytes, 16));
    out->entry_ = slot >= 0 ? e : nullptr;
    out->slot_ = slot;
    out->device_ = device;
    out->stream_ = stream;
}

}  // namespace

// End of a call that used a cached table: mark "last use" on its stream, behind everything the call launched.
void TableRef::release() {
    if (!entry_) return;
    Entry* e = static_cast<Entry*>(entry_);
    entry_ = nullptr;
    DeviceCache& c = g_cache[device_];
    std::lock_guard<std::mutex> lk(c.mu);
    // (the entry cannot have been evicted since get_table(): `leased` kept it off the eviction list)
    e->leased--;
    if (cudaEventRecord(e->use_ev[slot_], stream_) == cudaSuccess) e->use_live[slot_] = true;
    else (void)cudaGetLastError();
}


int get_table(const MtTensorLists* tl, int tile_elems, cudaStream_t stream, TableRef* out) {
    const int device = tl->device;
    if (device < 0 || device >= 64) return set_error(MT_EINVAL, "device ordinal %d out of range", device);
    const size_t nt = (size_t)tl->ntensors;
    const int depth = tl->depth;
    out->release();

    // blob layout (device and host identical):
    //   [0]                 header: depth, tile_elems (part of the key)
    //   [off_tb]            uint32 tile_begin[nt + 1]
    //   [off_numel]         int64  numel[nt]
    //   [off_ptrs]          void*  ptrs[depth][nt]
    const size_t off_tb = 16;
    const size_t off_numel = align_up(off_tb + (nt + 1) * 4, 16);
    const size_t off_ptrs = off_numel + nt * 8;
    const size_t bytes = off_ptrs + (size_t)depth * nt * 8;

    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    MT_CUDA_OK(cudaStreamIsCapturing(stream, &cap));
    const bool capturing = cap != cudaStreamCaptureStatusNone;
    DeviceCache& c = g_cache[device];
    {   // MRU fast path: no blob, no hash
        std::lock_guard<std::mutex> lk(c.mu);
        for (int i = 0; i < c.mru_n; i++) {
            Entry* e = c.mru[i];
            if (e->bytes != bytes) continue;
            const uint8_t* hb = e->host.data();
            const int32_t* hdr = reinterpret_cast<const int32_t*>(hb);
            if (hdr[0] != depth || hdr[1] != tile_elems) continue;
            if (memcmp(hb + off_ptrs, tl->ptrs, (size_t)depth * nt * 8) != 0) continue;
            if (memcmp(hb + off_numel, tl->numel, nt * 8) != 0) continue;
            c.stats.hits++;
            int slot;
            const int rc = acquire_locked(c, e, stream, capturing, &slot);
            if (rc) return rc;
            if (i) mru_touch(c, e);
            fill_out(e, device, stream, slot, nt, depth, tile_elems, off_tb, off_numel, off_ptrs, out);
            return MT_OK;
        }
    }

    {   // not seen recently: per-entry validation before anything is built from the arrays
        const int vrc = validate_entries(tl, "tensor lists");
        if (vrc) return vrc;
    }
    thread_local std::vector<uint8_t> blob;
    blob.resize(bytes);
    memset(blob.data(), 0, off_tb);  // header padding; the arrays below are fully overwritten (plus their alignment gap)
    if (off_numel > off_tb + (nt + 1) * 4) memset(blob.data() + off_tb + (nt + 1) * 4, 0, off_numel - (off_tb + (nt + 1) * 4));
    {
        int32_t hdr[4] = {depth, tile_elems, 0, 0};
        memcpy(blob.data(), hdr, sizeof hdr);
        uint32_t* tb = reinterpret_cast<uint32_t*>(blob.data() + off_tb);
        uint64_t acc = 0;
        for (size_t t = 0; t < nt; t++) {
            tb[t] = (uint32_t)acc;
            acc += (uint64_t)((tl->numel[t] + tile_elems - 1) / tile_elems);
            if (acc > 0x7fffffffull) return set_error(MT_EINVAL, "too many tiles (%llu)", (unsigned long long)acc);
        }
        tb[nt] = (uint32_t)acc;
        memcpy(blob.data() + off_numel, tl->numel, nt * 8);
        memcpy(blob.data() + off_ptrs, tl->ptrs, (size_t)depth * nt * 8);
    }
    const uint32_t ntiles = reinterpret_cast<uint32_t*>(blob.data() + off_tb)[nt];
    const uint64_t h = hash_blob(blob.data(), bytes);

    std::lock_guard<std::mutex> lk(c.mu);
    RelaxedCaptureMode relaxed;  // miss path: eviction queries, staging waits, pinned allocation
    Entry* e = nullptr;
    auto range = c.map.equal_range(h);
    for (auto it = range.first; it != range.second; ++it)
        if (it->second.bytes == bytes && memcmp(it->second.host.data(), blob.data(), bytes) == 0) {
            e = &it->second;
            break;
        }
    if (e) {
        c.stats.hits++;
    } else {
        c.stats.misses++;
        if (capturing)
            return set_error(MT_ECAPTURE,
                             "device tensor-table cache miss while the stream is being captured; run the same call "
                             "once eagerly (warm-up) before capture so the table is resident");
        if (c.map.size() >= max_entries() || c.stats.device_bytes + bytes > max_device_bytes())
            evict_lru_locked(c, bytes, stream);
        // staging: one pinned buffer per device, reused once the previous upload has drained
        if (!c.staging_ev) MT_CUDA_OK(cudaEventCreateWithFlags(&c.staging_ev, cudaEventDisableTiming));
        if (c.staging_pending) {
            MT_CUDA_OK(cudaEventSynchronize(c.staging_ev));
            c.staging_pending = false;
        }
        if (c.staging_cap < bytes) {
            if (c.staging) cudaFreeHost(c.staging);
            c.staging = nullptr;
            c.staging_cap = 0;
            const size_t cap_bytes = align_up(bytes * 2, 1 << 16);
            MT_CUDA_OK(cudaMallocHost(&c.staging, cap_bytes));
            c.staging_cap = cap_bytes;
        }
        Entry ne;
        MT_CUDA_OK(cudaEventCreateWithFlags(&ne.upload_ev, cudaEventDisableTiming));
        // device copy = blob + the completion counters of the reducing ops ([nt + 1] uint32, zero; see norm_tile_done)
        const size_t off_tickets = align_up(bytes, 16), dev_bytes = off_tickets + (nt + 1) * 4;
        cudaError_t ce = cudaMallocAsync(&ne.dev, dev_bytes, stream);
        if (ce == cudaSuccess) {
            memcpy(c.staging, blob.data(), bytes);
            ce = cudaMemcpyAsync(ne.dev, c.staging, bytes, cudaMemcpyHostToDevice, stream);
            if (ce == cudaSuccess)
                ce = cudaMemsetAsync(static_cast<uint8_t*>(ne.dev) + off_tickets, 0, dev_bytes - off_tickets, stream);
            if (ce == cudaSuccess) ce = cudaEventRecord(ne.upload_ev, stream);
            if (ce == cudaSuccess) ce = cudaEventRecord(c.staging_ev, stream);
        }
        if (ce != cudaSuccess) {
            if (ne.dev) cudaFreeAsync(ne.dev, stream);
            cudaEventDestroy(ne.upload_ev);
            return set_error(MT_ECUDA, "table upload failed: %s", cudaGetErrorString(ce));
        }
        c.staging_pending = true;
        ne.upload_stream = stream;
        ne.host = blob;
        ne.bytes = bytes;
        ne.ntiles = ntiles;
        c.stats.uploads_bytes += bytes;
        c.stats.device_bytes += bytes;  // (the few counter bytes behind the blob are not itemised)
        auto it = c.map.emplace(h, std::move(ne));
        c.stats.entries = c.map.size();
        e = &it->second;
    }
    int slot;
    const int rc = acquire_locked(c, e, stream, capturing, &slot);
    if (rc) return rc;
    mru_touch(c, e);
    fill_out(e, device, stream, slot, nt, depth, tile_elems, off_tb, off_numel, off_ptrs, out);
    return MT_OK;
}

}  // namespace mt

// ---- C ABI: runtime / introspection ---------------------------------------------------------------
extern "C" {

const char* mt_last_error(void) { return mt::tls_error.c_str(); }

const char* mt_version(void) { return "mt_b200 0.1 (sm_100a)"; }

uint64_t mt_launch_count(void) { return mt::g_launches.load(std::memory_order_relaxed); }

int mt_table_cache_stats(int device, MtCacheStats* out) {
    if (device < 0 || device >= 64 || !out) return mt::set_error(MT_EINVAL, "mt_table_cache_stats: bad argument");
    std::lock_guard<std::mutex> lk(mt::g_cache[device].mu);
    *out = mt::g_cache[device].stats;
    return MT_OK;
}

int mt_table_cache_clear(int device) {
    if (device < 0 || device >= 64) return mt::set_error(MT_EINVAL, "mt_table_cache_clear: bad device");
    mt::DeviceGuard g(device);
    if (!g.ok) return mt::set_error(MT_ECUDA, "mt_table_cache_clear: cannot select device %d", device);
    std::lock_guard<std::mutex> lk(mt::g_cache[device].mu);
    return mt::evict_all_locked(mt::g_cache[device]);
}

}  // extern "C"
