// mt_host.h -- host-side internals of libmt_b200.so shared by the per-family translation units:
// error state, device guard, the device-resident table cache and the table-driven launcher.
#pragma once

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <string>

#include "../../include/mt_b200.h"
#include "mt_common.cuh"

namespace mt {

// ---- error state (thread-local message, negative return codes) ---------------------------------
int set_error(int code, const char* fmt, ...);
#define MT_CUDA_OK(expr)                                                                          \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess)                                                                    \
            return ::mt::set_error(MT_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

extern std::atomic<uint64_t> g_launches;

struct DeviceGuard {
    int prev = -1, cur = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        cur = dev;
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() {
        if (ok && prev >= 0 && prev != cur) cudaSetDevice(prev);
    }
};

struct DeviceInfo {
    int sms = 0;
};
int device_info(int device, DeviceInfo* out);

// ---- validation ---------------------------------------------------------------------------------
// Mirrors the checks of /root/reference/csrc/multi_tensor_apply.cuh:53-72 that survive the loss of
// torch types (depth, list lengths and numel equality are structural in MtTensorLists).
int validate_lists(const MtTensorLists* tl, int expect_depth_lo, int expect_depth_hi, const char* who);  // O(1)
int validate_entries(const MtTensorLists* tl, const char* who);  // O(ntensors): negative numel, NULL pointers
int64_t total_elems(const MtTensorLists* tl);

// ---- tiles ----------------------------------------------------------------------------------------
// Element-wise ops ignore the caller's chunk_size and pick their own tile: 16384 elements (8 rounds of a 256-thread
// CTA) when that still yields several waves of CTAs, shrinking to one round (2048) for small problems.
int choose_tile_elems(const MtTensorLists* tl, int device);
uint64_t count_tiles(const MtTensorLists* tl, int tile_elems);

// A cached table as one call holds it: the DevTable the kernels take (pass it by value -- it slices) plus a lease on
// the cache entry.  release() (run by the destructor, i.e. behind the call's last launch) records the entry's
// "last use" event on the call's stream; eviction only ever takes entries whose events have completed.
struct TableRef : DevTable {
    TableRef() : DevTable{} {}
    TableRef(const TableRef&) = delete;
    TableRef& operator=(const TableRef&) = delete;
    ~TableRef() { release(); }
    void release();
    void* entry_ = nullptr;
    int slot_ = -1, device_ = 0;
    cudaStream_t stream_ = nullptr;
};

// Look the (lists, tile_elems) pair up in the per-device table cache; on a miss build the table, upload it
// (one cudaMemcpyAsync on `stream`) and cache it.  Steady-state steps with unchanged pointers upload nothing.
// A hit from a stream other than the uploading one is ordered behind the upload (see mt_table.cu).
int get_table(const MtTensorLists* tl, int tile_elems, cudaStream_t stream, TableRef* out);

// ---- launcher -------------------------------------------------------------------------------------
// One CTA per tile (see mt_common.cuh for the measurements behind this choice).  gridDim.x can hold 2^31-1 tiles,
// which get_table() enforces, so the in-kernel tile loop runs once.
template <class Op, int G, int U, int THREADS, int MINB>
int launch_table_kernel(const DevTable& tb, const typename Op::Params& prm, int32_t* noop, int device,
                        cudaStream_t stream) {
    (void)device;
    if (tb.ntiles == 0) return MT_OK;  // all tensors empty: nothing to do (the reference launches nothing either)
    mt_table_kernel<Op, G, U, THREADS, MINB><<<tb.ntiles, THREADS, 0, stream>>>(tb, prm, noop);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    MT_CUDA_OK(cudaGetLastError());
    return MT_OK;
}

// Launch shape shared by the streaming families: 256 threads, 4 CTAs/SM (<= 64 registers), 8 elements per thread per
// access (LDG.256 for fp32, LDG.128 for 16-bit), one group in flight per thread; tiles of up to 16384 elements.
constexpr int kG = 8, kU = 1, kThreads = 256, kMinBlocks = 4;
constexpr int kRound = kG * kThreads;   // elements one CTA moves per round
constexpr int kMaxTile = 16384;

// dtype-tag dispatch helpers
#define MT_CASE(TAG, CT, NAME, ...) \
    case TAG: {                     \
        using NAME = CT;            \
        __VA_ARGS__;                \
    } break;
#define MT_DISPATCH_FHB(tag, NAME, who, ...)                                          \
    switch (tag) {                                                                    \
        MT_CASE(MT_F32, float, NAME, __VA_ARGS__)                                     \
        MT_CASE(MT_F16, __half, NAME, __VA_ARGS__)                                    \
        MT_CASE(MT_BF16, __nv_bfloat16, NAME, __VA_ARGS__)                            \
        default:                                                                      \
            return ::mt::set_error(MT_EINVAL, "%s: unsupported dtype tag %d", who, (int)(tag)); \
    }

}  // namespace mt
