// mt_api_noalloc.cu -- mt_multi_tensor_l2norm_noalloc: the L2-norm op for hosts that cannot give us scratch memory.
//
// Who needs this.  mt_multi_tensor_l2norm wants two things from device memory besides its inputs and outputs: the
// cached tensor table (mt_table.cu cudaMalloc()s it on a cache miss) and a caller-provided scratch buffer for the
// per-chunk partial sums.  Some users of this library cannot supply either:
//   * a host written against the bare C ABI from another runtime (the cgo / JNI style stub of INTEGRATION.md) that
//     owns no device allocator -- it has the gradient buffers it was handed and nothing else;
//   * a trainer or server that reserves the whole HBM for one pool at start-up, so a later cudaMalloc simply fails;
//   * a step that has to stay free of cudaMalloc / cudaFree because both synchronise the whole device and would stall
//     the copy and collective streams running beside the optimizer.
// For them this file keeps ALL working memory of one l2norm call -- the tensor table and the partial sums -- in a
// fixed array that is part of the library image.  Every GPU that loads libmt_b200.so carries that array (kBytes, below)
// until the process exits, whether or not this entry point is ever called.
//
// Rules that follow from sharing one array:
//   * Sizes are capped (MT_NOALLOC_MAX_TENSORS tensors, MT_NOALLOC_MAX_CHUNKS chunks in total); larger requests are
//     refused with MT_EINVAL and a message, nothing is launched.
//   * The array keeps whatever the previous call left in it.  Each call first overwrites the table region, and the
//     reduction reads exactly the partials [0, nchunks) that the same call has just written (one per chunk, assigned
//     -- never accumulated), so a shorter call after a longer one cannot see stale partials.  The one region that must
//     start at zero is the block of per-tensor completion counters the kernel counts its finished chunks in: every
//     call clears the counters it is going to use (a cudaMemsetAsync on the caller's stream, no allocation) rather
//     than trusting the previous call to have left them at zero.
//   * Two calls whose GPU work overlaps (different streams, or different host threads) race on the array and produce
//     garbage.  Calls on one stream are ordered by the stream and are safe.
//   * The first call on a device looks up the array's device address (cudaGetSymbolAddress).  Make that first call
//     before starting any stream capture.
//   * No call may be issued on a stream that is being captured, the first or any later one: the table travels to the
//     GPU through one host buffer that the next call overwrites, so a replayed copy would deliver some other call's
//     table.  Such a call is refused with MT_ECAPTURE before anything is enqueued.
#include <cstring>
#include <mutex>

#include "mt_host.h"
#include "mt_ops.cuh"

namespace mt {

int launch_norm_finalize(const DevTable& tb, const float* partials, int nsets, float* ret_global, float* ret_pt,
                         int device, cudaStream_t stream, const int32_t* skip_if_set = nullptr);  // mt_api_norm.cu

namespace {

constexpr size_t kMaxTensors = MT_NOALLOC_MAX_TENSORS;
constexpr size_t kMaxChunks = MT_NOALLOC_MAX_CHUNKS;
// table region: tile_begin[(T+1) x u32, padded] + numel[T x i64] + ptrs[T x ptr]
constexpr size_t kTableBytes = ((kMaxTensors + 1) * 4 + 15) / 16 * 16 + kMaxTensors * 8 + kMaxTensors * 8;
constexpr size_t kTicketOff = (kTableBytes + 255) / 256 * 256;  // completion counters, [T + 1] x u32 (norm_tile_done)
constexpr size_t kPartialOff = (kTicketOff + (kMaxTensors + 1) * 4 + 255) / 256 * 256;
constexpr size_t kBytes = kPartialOff + kMaxChunks * sizeof(float);

// The working memory.  In host code this name is only a handle for the runtime, NOT a device address: it must go
// through cudaGetSymbolAddress before a kernel may be given it.
static __device__ __align__(256) unsigned char g_noalloc_mem[kBytes];

std::mutex g_mu;
unsigned char* g_dev_addr[64];        // per device: resolved address of g_noalloc_mem
unsigned char g_host_table[kTableBytes];  // host-side staging for the table (guarded by g_mu)

// Once per device, with that device current (the caller holds a DeviceGuard).
int resolve(int device, unsigned char** out) {
    if (!g_dev_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_mem);
        if (e != cudaSuccess || !p)
            return set_error(MT_ECUDA, "multi_tensor_l2norm_noalloc: cudaGetSymbolAddress failed on device %d: %s", device,
                             cudaGetErrorString(e));
        g_dev_addr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_dev_addr[device];
    return MT_OK;
}

}  // namespace
}  // namespace mt

extern "C" size_t mt_noalloc_resident_bytes(void) { return mt::kBytes; }

extern "C" int mt_multi_tensor_l2norm_noalloc(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, int per_tensor,
                                              float* ret, float* ret_per_tensor, void* stream_) {
    using namespace mt;
    const char* who = "multi_tensor_l2norm_noalloc";
    int rc = validate_lists(tl, 1, 1, who);
    if (!rc) rc = validate_entries(tl, who);
    if (rc) return rc;
    if (chunk_size <= 0) return set_error(MT_EINVAL, "%s: chunk_size must be positive", who);
    if (!noop_flag || !ret) return set_error(MT_EINVAL, "%s: noop_flag / ret is NULL", who);
    if (per_tensor && !ret_per_tensor) return set_error(MT_EINVAL, "%s: per_tensor requested but ret_per_tensor is NULL", who);
    const size_t nt = (size_t)tl->ntensors;
    if (nt > kMaxTensors)
        return set_error(MT_EINVAL, "%s: %zu tensors exceed the built-in limit of %zu (MT_NOALLOC_MAX_TENSORS); use "
                         "mt_multi_tensor_l2norm with caller-provided scratch instead", who, nt, kMaxTensors);
    const uint64_t nchunks = count_tiles(tl, chunk_size);
    if (nchunks > kMaxChunks)
        return set_error(MT_EINVAL, "%s: %llu chunks of %d elements exceed the built-in limit of %zu "
                         "(MT_NOALLOC_MAX_CHUNKS); use a larger chunk_size or mt_multi_tensor_l2norm", who,
                         (unsigned long long)nchunks, chunk_size, kMaxChunks);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (tl->device < 0 || tl->device >= 64) return set_error(MT_EINVAL, "%s: device ordinal %d out of range", who, tl->device);
    DeviceGuard guard(tl->device);
    if (!guard.ok) return set_error(MT_ECUDA, "%s: cannot select CUDA device %d", who, tl->device);

    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    MT_CUDA_OK(cudaStreamIsCapturing(stream, &cap));
    if (cap != cudaStreamCaptureStatusNone)
        return set_error(MT_ECAPTURE, "%s: the stream is being captured; this entry point stages its tensor table through a "
                         "shared host buffer and cannot be recorded into a graph -- call it outside the capture, or "
                         "capture mt_multi_tensor_l2norm with caller-provided scratch instead", who);

    std::lock_guard<std::mutex> lk(g_mu);
    unsigned char* dev = nullptr;
    rc = resolve(tl->device, &dev);
    if (rc) return rc;

    // table, laid out for this call's tensor count
    const size_t off_numel = ((nt + 1) * 4 + 15) / 16 * 16;
    const size_t off_ptrs = off_numel + nt * 8;
    const size_t table_bytes = off_ptrs + nt * 8;
    uint32_t* tb_host = reinterpret_cast<uint32_t*>(g_host_table);
    uint64_t acc = 0;
    for (size_t t = 0; t < nt; t++) {
        tb_host[t] = (uint32_t)acc;
        acc += (uint64_t)((tl->numel[t] + chunk_size - 1) / chunk_size);
    }
    tb_host[nt] = (uint32_t)acc;
    memcpy(g_host_table + off_numel, tl->numel, nt * 8);
    memcpy(g_host_table + off_ptrs, tl->ptrs, nt * 8);
    // pageable source: the runtime has consumed g_host_table when this returns, so the next call may overwrite it
    MT_CUDA_OK(cudaMemcpyAsync(dev, g_host_table, table_bytes, cudaMemcpyHostToDevice, stream));
    MT_CUDA_OK(cudaMemsetAsync(dev + kTicketOff, 0, (nt + 1) * 4, stream));  // the counters must start at zero

    DevTable tb;
    tb.inl = 0;
    tb.tile_begin = reinterpret_cast<const uint32_t*>(dev);
    tb.numel = reinterpret_cast<const int64_t*>(dev + off_numel);
    tb.ptrs = reinterpret_cast<const void* const*>(dev + off_ptrs);
    tb.ntensors = (int32_t)nt;
    tb.ntiles = (uint32_t)nchunks;
    tb.tile_elems = chunk_size;
    tb.tickets = reinterpret_cast<uint32_t*>(dev + kTicketOff);
    float* partials = reinterpret_cast<float*>(dev + kPartialOff);

    float* ret_pt = per_tensor ? ret_per_tensor : nullptr;
    if (nchunks == 0) return launch_norm_finalize(tb, partials, 1, ret, ret_pt, tl->device, stream);  // all empty: zeros
    constexpr int G = 8, THREADS = 256, MINB = 4;  // same kernel and launch shape as mt_multi_tensor_l2norm
    MT_DISPATCH_FHB(tl->dtype[0], Tx, who, {
        constexpr int U = sizeof(Tx) == 4 ? 2 : 4;
        using Op = L2NormOp<Tx, false>;
        typename Op::Params prm{NormOut{partials, ret, ret_pt, 1}};
        rc = launch_table_kernel<Op, G, U, THREADS, MINB>(tb, prm, noop_flag, tl->device, stream);
    })
    return rc;
}
