// mt_api_norm.cu -- multi_tensor_l2norm / multi_tensor_l2norm_scale entry points and the norm finalize kernel.
//
// ONE launch per call (table-driven grid, one tile == one reference chunk): per-tile sum of squares ->
// scratch partials[tile], flag raised when a tile's partial is non-finite (the reference's per-chunk definition,
// /root/reference/csrc/multi_tensor_l2norm_kernel.cu:103-104); the CTA that completes a tensor / the launch reduces
// the partials in a fixed order (norm_tile_done, mt_common.cuh) -> per-tensor norms and global norm.  Deterministic
// (no float atomics); replaces cleanup / cleanup_v3 (csrc/multi_tensor_l2norm_kernel.cu:198-237), which are separate
// launches in the reference.  norm_finalize_kernel is what remains of this library's own second launch: it serves the
// degenerate all-tensors-empty call only.
#include <cstdlib>

#include "mt_host.h"
#include "mt_ops.cuh"

namespace mt {

constexpr int kFinThreads = 512;

// partials: [nsets][ntiles].  ret_global[s] (may be NULL) = sqrt(sum over all tiles of set s), by block 0.
// ret_pt + s*ntensors (may be NULL) = per-tensor norms of set s: one warp per tensor, lanes stride the tensor's
// tiles, fixed shuffle tree.
__global__ void __launch_bounds__(kFinThreads)
norm_finalize_kernel(const float* __restrict__ partials, int nsets, uint32_t ntiles,
                     const uint32_t* __restrict__ tile_begin, int ntensors, float* __restrict__ ret_global,
                     float* __restrict__ ret_pt, const int32_t* __restrict__ skip_if_set) {
    __shared__ float smem[kFinThreads / 32];
    // "_mp" variants: leave the outputs untouched when the flag is set (csrc/multi_tensor_l2norm_kernel_mp.cu:121-123)
    if (skip_if_set && *reinterpret_cast<const volatile int32_t*>(skip_if_set) != 0) return;
    if (blockIdx.x == 0 && ret_global) {
        for (int s = 0; s < nsets; s++) {
            float acc = 0.f;
            for (uint32_t i = threadIdx.x; i < ntiles; i += kFinThreads) acc = fadd(acc, partials[(size_t)s * ntiles + i]);
            const float tot = block_sum<kFinThreads>(acc, smem);
            if (threadIdx.x == 0) ret_global[s] = fsqrt(tot);
        }
    }
    if (ret_pt) {
        const int lane = threadIdx.x & 31;
        const int warps = (gridDim.x * kFinThreads) >> 5;
        for (int w = (blockIdx.x * kFinThreads + threadIdx.x) >> 5; w < ntensors * nsets; w += warps) {
            const int s = w / ntensors, t = w - s * ntensors;
            const uint32_t b = tile_begin[t], e = tile_begin[t + 1];
            float acc = 0.f;
            for (uint32_t i = b + lane; i < e; i += 32) acc = fadd(acc, partials[(size_t)s * ntiles + i]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc = fadd(acc, __shfl_xor_sync(0xffffffffu, acc, o));
            if (lane == 0) ret_pt[(size_t)s * ntensors + t] = fsqrt(acc);
        }
    }
}

int launch_norm_finalize(const DevTable& tb, const float* partials, int nsets, float* ret_global, float* ret_pt,
                         int device, cudaStream_t stream, const int32_t* skip_if_set) {
    DeviceInfo di;
    int rc = device_info(device, &di);
    if (rc) return rc;
    int grid = 1;
    if (ret_pt) {
        const long long warps_needed = (long long)tb.ntensors * nsets;
        long long blocks = (warps_needed + (kFinThreads / 32) - 1) / (kFinThreads / 32);
        const long long cap = (long long)di.sms * 4;
        grid = (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
    }
    norm_finalize_kernel<<<grid, kFinThreads, 0, stream>>>(partials, nsets, tb.ntiles, tb.tile_begin, tb.ntensors,
                                                           ret_global, ret_pt, skip_if_set);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    MT_CUDA_OK(cudaGetLastError());
    return MT_OK;
}

// One tile per reference chunk by default.  Splitting a large chunk_size (FusedLAMB passes 65536) into several tiles
// gives the last wave more CTAs to balance with, but every tile pays a fence + two atomics + a barrier for its
// completion tickets: measured on GPT-2's 124 M fp16 gradients (profiles/r2_norm_tile_sweep.jsonl) 8192-element tiles
// took 118 us against 55 us for whole chunks.  MT_NORM_TILE=<elements> (a divisor of chunk_size) overrides, for sweeps;
// the flag keeps its per-chunk meaning either way (NormOut::sub).
int norm_tiles_per_chunk(int chunk_size) {
    static const int forced = [] {
        const char* e = getenv("MT_NORM_TILE");
        return e ? atoi(e) : 0;
    }();
    if (forced > 0 && chunk_size > forced && chunk_size % forced == 0) return chunk_size / forced;
    return 1;
}

static int norm_common(const char* who, int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, int depth,
                       bool mp, float scale, int per_tensor, float* ret, float* ret_per_tensor, void* scratch,
                       size_t scratch_bytes, cudaStream_t stream) {
    int rc = validate_lists(tl, depth, depth, who);
    if (rc) return rc;
    if (chunk_size <= 0) return set_error(MT_EINVAL, "%s: chunk_size must be positive", who);
    if (!noop_flag || !ret) return set_error(MT_EINVAL, "%s: noop_flag / ret is NULL", who);
    if (per_tensor && !ret_per_tensor) return set_error(MT_EINVAL, "%s: per_tensor requested but ret_per_tensor is NULL", who);
    DeviceGuard guard(tl->device);
    if (!guard.ok) return set_error(MT_ECUDA, "%s: cannot select CUDA device %d", who, tl->device);
    TableRef tb;
    const int sub = norm_tiles_per_chunk(chunk_size);
    rc = get_table(tl, chunk_size / sub, stream, &tb);
    if (rc) return rc;
    const size_t need = (size_t)tb.ntiles * sizeof(float);
    if (need > 0 && (!scratch || scratch_bytes < need))
        return set_error(MT_ESCRATCH, "%s: scratch too small (%zu < %zu bytes)", who, scratch_bytes, need);
    float* partials = static_cast<float*>(scratch);
    float* ret_pt = per_tensor ? ret_per_tensor : nullptr;
    if (tb.ntiles == 0)  // every tensor empty: zeros, by the small kernel (no tile exists to do it in-launch)
        return launch_norm_finalize(tb, partials, 1, ret, ret_pt, tl->device, stream, mp ? noop_flag : nullptr);
    const NormOut out{partials, ret, ret_pt, sub};
    // Bytes in flight per thread.  A norm pass has one input stream where Adam has five, so each thread must keep more
    // groups in flight to cover the loaded HBM latency: kDepth x (2 groups of 32 B for fp32 | 4 groups of 16 B for the
    // 16-bit types).  MT_NORM_DEPTH=1|2 selects the depth for sweeps (profiles/r2_norm_depth_sweep.jsonl).
    // MT_NORM_DEPTH=3 additionally widens the 16-bit accesses to 32 bytes (16 elements per load, LDG.256): measured equal
    // to depth 2 (49.2 vs 48.7 us on GPT-2's 249 MB of fp16 gradients) -- at that size the pass is at what a plain copy
    // of the same bytes attains (49.9 us), so depth 2 stays the default and 3 a knob.
    constexpr int G = 8, THREADS = 256, MINB = 4;
    static const int depth_knob = [] {
        const char* e = getenv("MT_NORM_DEPTH");
        const int v = e ? atoi(e) : 0;
        return (v >= 1 && v <= 3) ? v : 2;
    }();
    if (depth == 1) {
        MT_DISPATCH_FHB(tl->dtype[0], Tx, who, {
            constexpr int U1 = sizeof(Tx) == 4 ? 2 : 4;
            constexpr int GW = sizeof(Tx) == 4 ? 8 : 16;   // wide variant: 32-byte accesses for every dtype
            if (mp) {
                using Op = L2NormOp<Tx, true>;
                typename Op::Params prm{out};
                rc = depth_knob == 3 ? launch_table_kernel<Op, GW, 4, THREADS, MINB>(tb, prm, noop_flag, tl->device, stream)
                   : depth_knob == 2 ? launch_table_kernel<Op, G, 2 * U1, THREADS, MINB>(tb, prm, noop_flag, tl->device, stream)
                                     : launch_table_kernel<Op, G, U1, THREADS, MINB>(tb, prm, noop_flag, tl->device, stream);
            } else {
                using Op = L2NormOp<Tx, false>;
                typename Op::Params prm{out};
                rc = depth_knob == 3 ? launch_table_kernel<Op, GW, 4, THREADS, MINB>(tb, prm, noop_flag, tl->device, stream)
                   : depth_knob == 2 ? launch_table_kernel<Op, G, 2 * U1, THREADS, MINB>(tb, prm, noop_flag, tl->device, stream)
                                     : launch_table_kernel<Op, G, U1, THREADS, MINB>(tb, prm, noop_flag, tl->device, stream);
            }
        })
    } else {
        MT_DISPATCH_FHB(tl->dtype[0], Tin, who,
            MT_DISPATCH_FHB(tl->dtype[1], Tout, who, {
                using Op = L2NormScaleOp<Tin, Tout>;
                typename Op::Params prm{out, scale};
                rc = launch_table_kernel<Op, G, 2, THREADS, MINB>(tb, prm, noop_flag, tl->device, stream);
            }))
    }
    return rc;
}

}  // namespace mt

#ifdef MT_TRACE
extern "C" int mt_debug_set_trace(void* buf) {  // tools/norm_trace.py; not part of the product ABI
    unsigned long long* p = static_cast<unsigned long long*>(buf);
    return (int)cudaMemcpyToSymbol(mt::g_mt_trace, &p, sizeof(p));
}
#endif

extern "C" size_t mt_norm_scratch_bytes(int chunk_size, const MtTensorLists* tl) {
    if (!tl || !tl->numel || chunk_size <= 0) return 0;
    return (size_t)mt::count_tiles(tl, chunk_size / mt::norm_tiles_per_chunk(chunk_size)) * sizeof(float);
}

extern "C" int mt_multi_tensor_l2norm(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, int per_tensor,
                                      float* ret, float* ret_per_tensor, void* scratch, size_t scratch_bytes,
                                      void* stream) {
    return mt::norm_common("multi_tensor_l2norm", chunk_size, noop_flag, tl, 1, false, 0.f, per_tensor, ret, ret_per_tensor,
                           scratch, scratch_bytes, static_cast<cudaStream_t>(stream));
}

extern "C" int mt_multi_tensor_l2norm_scale(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float scale,
                                            int per_tensor, float* ret, float* ret_per_tensor, void* scratch,
                                            size_t scratch_bytes, void* stream) {
    return mt::norm_common("multi_tensor_l2norm_scale", chunk_size, noop_flag, tl, 2, false, scale, per_tensor, ret,
                           ret_per_tensor, scratch, scratch_bytes, static_cast<cudaStream_t>(stream));
}

extern "C" int mt_multi_tensor_l2norm_mp(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, int per_tensor,
                                         float* ret, float* ret_per_tensor, void* scratch, size_t scratch_bytes,
                                         void* stream) {
    return mt::norm_common("multi_tensor_l2norm_mp", chunk_size, noop_flag, tl, 1, true, 0.f, per_tensor, ret,
                           ret_per_tensor, scratch, scratch_bytes, static_cast<cudaStream_t>(stream));
}
