// mt_table.cu -- host runtime of libmt_b200.so: error state, validation, tile policy and the
// device-resident tensor-table cache.
//
// The reference re-packs pointers and sizes into a ~14 KB by-value kernel argument for EVERY launch of EVERY
// call (/root/reference/csrc/multi_tensor_apply.cuh:84-143; ~383 launches for a Llama-3-8B Adam step,
// >=1042 for 50k small tensors).  Here the table lives in device memory, is built once per distinct
// (pointer set, sizes, tile size) and is found again by hash, so a steady-state optimizer step uploads nothing
// and issues ONE launch per op.
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
        if (tl->dtype[d] < MT_F32 || tl->dtype[d] > MT_F64)
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
namespace {

struct Entry {
    std::vector<uint8_t> host;  // exact blob, for collision-proof comparison
    void* dev = nullptr;
    size_t bytes = 0;
    uint32_t ntiles = 0;
    bool in_graph = false;  // referenced by a stream capture: never evicted
};

struct DeviceCache {
    std::mutex mu;
    std::unordered_multimap<uint64_t, Entry> map;
    // Most-recently-used entries, checked before anything is hashed: a training loop cycles through a handful of
    // (op, lists) pairs, and comparing the caller's pointer / numel arrays straight against a cached blob is one
    // memcmp pass instead of blob build (a 64-bit division per tensor) + hash + memcmp.  Pointers into `map` nodes
    // (stable until erased; cleared on eviction).
    static constexpr int kMru = 16;
    Entry* mru[kMru] = {};
    int mru_n = 0;
    void* staging = nullptr;  // pinned
    size_t staging_cap = 0;
    cudaEvent_t staging_ev = nullptr;
    bool staging_pending = false;
    MtCacheStats stats{};
};

DeviceCache g_cache[64];

constexpr size_t kMaxEntries = 1024;
constexpr size_t kMaxDeviceBytes = 1ull << 30;

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

int evict_all_locked(DeviceCache& c) {
    // rare path: the cache outgrew its caps.  In-flight kernels may still read the tables -> full sync first.
    MT_CUDA_OK(cudaDeviceSynchronize());
    for (auto it = c.map.begin(); it != c.map.end();) {
        if (it->second.in_graph) { ++it; continue; }
        cudaFree(it->second.dev);
        c.stats.device_bytes -= it->second.bytes;
        it = c.map.erase(it);
    }
    c.mru_n = 0;
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

inline void fill_out(const Entry* e, size_t nt, int tile_elems, size_t off_tb, size_t off_numel, size_t off_ptrs,
                     DevTable* out) {
    const uint8_t* d = static_cast<const uint8_t*>(e->dev);
    out->tile_begin = reinterpret_cast<const uint32_t*>(d + off_tb);
    out->numel = reinterpret_cast<const int64_t*>(d + off_numel);
    out->ptrs = reinterpret_cast<const void* const*>(d + off_ptrs);
    out->ntensors = (int32_t)nt;
    out->ntiles = e->ntiles;
    out->tile_elems = tile_elems;
}

}  // namespace

int get_table(const MtTensorLists* tl, int tile_elems, cudaStream_t stream, DevTable* out) {
    const int device = tl->device;
    if (device < 0 || device >= 64) return set_error(MT_EINVAL, "device ordinal %d out of range", device);
    const size_t nt = (size_t)tl->ntensors;
    const int depth = tl->depth;

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
            if (capturing) e->in_graph = true;
            if (i) mru_touch(c, e);
            fill_out(e, nt, tile_elems, off_tb, off_numel, off_ptrs, out);
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
    Entry* e = nullptr;
    auto range = c.map.equal_range(h);
    for (auto it = range.first; it != range.second; ++it)
        if (it->second.bytes == bytes && memcmp(it->second.host.data(), blob.data(), bytes) == 0) {
            e = &it->second;
            break;
        }
    if (e) {
        c.stats.hits++;
        if (capturing) e->in_graph = true;
    } else {
        c.stats.misses++;
        if (capturing)
            return set_error(MT_ECAPTURE,
                             "device tensor-table cache miss while the stream is being captured; run the same call "
                             "once eagerly (warm-up) before capture so the table is resident");
        if (c.map.size() >= kMaxEntries || c.stats.device_bytes + bytes > kMaxDeviceBytes) {
            int rc = evict_all_locked(c);
            if (rc) return rc;
        }
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
        MT_CUDA_OK(cudaMalloc(&ne.dev, bytes));
        memcpy(c.staging, blob.data(), bytes);
        cudaError_t ce = cudaMemcpyAsync(ne.dev, c.staging, bytes, cudaMemcpyHostToDevice, stream);
        if (ce == cudaSuccess) ce = cudaEventRecord(c.staging_ev, stream);
        if (ce != cudaSuccess) {
            cudaFree(ne.dev);
            return set_error(MT_ECUDA, "table upload failed: %s", cudaGetErrorString(ce));
        }
        c.staging_pending = true;
        ne.host = blob;
        ne.bytes = bytes;
        ne.ntiles = ntiles;
        c.stats.uploads_bytes += bytes;
        c.stats.device_bytes += bytes;
        auto it = c.map.emplace(h, std::move(ne));
        c.stats.entries = c.map.size();
        e = &it->second;
    }
    mru_touch(c, e);
    fill_out(e, nt, tile_elems, off_tb, off_numel, off_ptrs, out);
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
