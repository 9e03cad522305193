// mt_api_distlamb.cu -- the two kernels behind apex.contrib.optimizers.DistributedFusedLAMB (SURVEY.md section 8f,
// rank 4): /root/reference/apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb.cpp:3-29 (signatures),
// multi_tensor_distopt_lamb_kernel.cu:119-321 (stage 1: update term), :325-433 (stage 2: weight update).
//
// Besides the reference's two entry points, stage 1 can hand back what the reference's Python gathers with two more
// multi_tensor_l2norm sweeps (distributed_fused_lamb.py:673-692): per parameter SEGMENT (a run of consecutive
// fragments that belong to one model parameter) the sums of round_model(p)^2 and u^2 over this rank's shard.  The
// caller all-reduces that small vector and takes the square roots.
#include "mt_host.h"
#include "mt_ops.cuh"

namespace mt {

constexpr int kSegThreads = 256;
// Stage 1 moves five lists and keeps two norm accumulators live across the IEEE-division slow-path calls: at 4 CTAs/SM
// (64 registers) ptxas spills up to 188 bytes per thread inside the element loop; 3 CTAs/SM (80 registers) is
// spill-free.  Measured on a 336 M-element shard (tools/bench_ops.py --configs c6, profiles/r1_ops_sharded_lamb.jsonl):
// 1.484 ms vs 1.502 ms -- no difference beyond noise, so the spill-free build is kept.
constexpr int kStage1MinBlocks = 3;

// partials: [2][ntiles] per-tile sums.  out: [2][nseg] sums (NOT square-rooted: they are all-reduced first).
// Segment s covers fragments [seg_begin[s], seg_begin[s+1]) = tiles [tile_begin[first], tile_begin[last+1]).
// One warp per (set, segment); lanes stride the tiles; fixed shuffle tree: deterministic.
__global__ void __launch_bounds__(kSegThreads)
seg_sq_finalize_kernel(const float* __restrict__ partials, uint32_t ntiles, const uint32_t* __restrict__ tile_begin,
                       const int32_t* __restrict__ seg_begin, int nseg, float* __restrict__ out,
                       const int32_t* __restrict__ skip_if_set) {
    if (*reinterpret_cast<const volatile int32_t*>(skip_if_set) != 0) return;
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * kSegThreads) >> 5;
    for (int w = (blockIdx.x * kSegThreads + threadIdx.x) >> 5; w < 2 * nseg; w += warps) {
        const int set = w / nseg, s = w - set * nseg;
        const uint32_t b = tile_begin[seg_begin[s]], e = tile_begin[seg_begin[s + 1]];
        float acc = 0.f;
        for (uint32_t i = b + lane; i < e; i += 32) acc = fadd(acc, partials[(size_t)set * ntiles + i]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc = fadd(acc, __shfl_xor_sync(0xffffffffu, acc, o));
        if (lane == 0) out[(size_t)set * nseg + s] = acc;
    }
}

}  // namespace mt

#define MT_PROLOGUE(who, lo, hi)                                                                    \
    using namespace mt;                                                                             \
    (void)chunk_size;                                                                               \
    int rc = validate_lists(tl, lo, hi, who);                                                       \
    if (rc) return rc;                                                                              \
    if (!noop_flag) return set_error(MT_EINVAL, who ": noop_flag is NULL");                         \
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);                                       \
    DeviceGuard guard(tl->device);                                                                  \
    if (!guard.ok) return set_error(MT_ECUDA, who ": cannot select CUDA device %d", tl->device);

extern "C" size_t mt_dist_lamb_scratch_bytes(const MtTensorLists* tl) {
    if (!tl || !tl->numel || tl->ntensors <= 0) return 0;
    const int tile = mt::choose_tile_elems(tl, tl->device);
    return 2 * (size_t)mt::count_tiles(tl, tile) * sizeof(float);
}

extern "C" int mt_multi_tensor_lamb_compute_update_term(
    int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, const float* per_tensor_beta1,
    const float* per_tensor_beta2, const float* per_tensor_beta3, const int32_t* per_tensor_bias_correction,
    const int32_t* step, const float* per_tensor_epsilon, int mode, const float* per_tensor_decay,
    const float* global_scale, const float* global_grad_norm, float max_grad_norm, const int32_t* seg_begin, int nseg,
    int model_dtype, float* seg_sq_out, void* scratch, size_t scratch_bytes, void* stream_) {
    MT_PROLOGUE("multi_tensor_lamb_compute_update_term", 5, 5)
    (void)per_tensor_beta3;  // the reference functor recomputes beta3 = 1 - beta1 and never reads it (:140)
    if (!per_tensor_beta1 || !per_tensor_beta2 || !per_tensor_bias_correction || !step || !per_tensor_epsilon ||
        !per_tensor_decay || !global_scale || !global_grad_norm)
        return set_error(MT_EINVAL, "multi_tensor_lamb_compute_update_term: hyper-parameter arrays, step, global_scale and "
                                    "global_grad_norm must be device pointers");
    for (int d = 1; d < 5; d++)
        if (tl->dtype[d] != MT_F32)
            return set_error(MT_EINVAL, "multi_tensor_lamb_compute_update_term: p, m, v and the update term must be float32 "
                                        "(the sharded optimizer keeps fp32 state; list %d has dtype tag %d)", d, tl->dtype[d]);
    const int need_norms = seg_sq_out ? 1 : 0;
    if (need_norms && (!seg_begin || nseg <= 0))
        return set_error(MT_EINVAL, "multi_tensor_lamb_compute_update_term: seg_sq_out needs seg_begin and nseg > 0");
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    float* partials = static_cast<float*>(scratch);
    if (need_norms) {
        const size_t need = 2 * (size_t)tb.ntiles * sizeof(float);
        if (!scratch || scratch_bytes < need)
            return set_error(MT_ESCRATCH, "multi_tensor_lamb_compute_update_term: scratch too small (%zu < %zu bytes)",
                             scratch_bytes, need);
    }
#define MT_DL1(TG, TN)                                                                                                \
    {                                                                                                                 \
        if (mode == 0) {                                                                                              \
            using Op = DistLambStage1Op<TG, TN, 0>;                                                                   \
            typename Op::Params prm{per_tensor_beta1, per_tensor_beta2, per_tensor_epsilon, per_tensor_decay,         \
                                    per_tensor_bias_correction, step, global_scale, global_grad_norm, max_grad_norm,  \
                                    partials, tb.ntiles, need_norms};                                                 \
            rc = launch_table_kernel<Op, kG, kU, kThreads, kStage1MinBlocks>(tb, prm, noop_flag, tl->device, stream); \
        } else {                                                                                                      \
            using Op = DistLambStage1Op<TG, TN, 1>;                                                                   \
            typename Op::Params prm{per_tensor_beta1, per_tensor_beta2, per_tensor_epsilon, per_tensor_decay,         \
                                    per_tensor_bias_correction, step, global_scale, global_grad_norm, max_grad_norm,  \
                                    partials, tb.ntiles, need_norms};                                                 \
            rc = launch_table_kernel<Op, kG, kU, kThreads, kStage1MinBlocks>(tb, prm, noop_flag, tl->device, stream); \
        }                                                                                                             \
    }
    // Tn only matters with need_norms; without it one instantiation per gradient type is enough
    const int mdt = need_norms ? model_dtype : MT_F32;
    if (mdt == MT_E5M2) {  // the model parameters are the e5m2-rounded masters (e5m2_allgather): norm them as such
        MT_DISPATCH_FHB(tl->dtype[0], Tg, "multi_tensor_lamb_compute_update_term", MT_DL1(Tg, e5m2_t))
    } else {
        MT_DISPATCH_FHB(tl->dtype[0], Tg, "multi_tensor_lamb_compute_update_term",
            MT_DISPATCH_FHB(mdt, Tn, "multi_tensor_lamb_compute_update_term", MT_DL1(Tg, Tn)))
    }
#undef MT_DL1
    if (rc) return rc;
    if (need_norms && tb.ntiles > 0) {
        const int blocks = (2 * nseg + (kSegThreads / 32) - 1) / (kSegThreads / 32);
        seg_sq_finalize_kernel<<<blocks > 592 ? 592 : blocks, kSegThreads, 0, stream>>>(
            partials, tb.ntiles, tb.tile_begin, seg_begin, nseg, seg_sq_out, noop_flag);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        MT_CUDA_OK(cudaGetLastError());
    }
    return MT_OK;
}

extern "C" int mt_multi_tensor_lamb_update_weights(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                                   const float* per_tensor_param_norm,
                                                   const float* per_tensor_update_norm,
                                                   const int64_t* update_norm_offset, const float* learning_rate,
                                                   const float* per_tensor_decay, const float* global_grad_norm,
                                                   int use_nvlamb, void* stream_) {
    MT_PROLOGUE("multi_tensor_lamb_update_weights", 3, 3)
    (void)global_grad_norm;  // carried for signature compatibility; the reference functor never reads it
    if (!per_tensor_param_norm || !per_tensor_update_norm || !update_norm_offset || !learning_rate || !per_tensor_decay)
        return set_error(MT_EINVAL, "multi_tensor_lamb_update_weights: norms, update_norm_offset, learning_rate and "
                                    "per_tensor_decay must be device pointers");
    if (tl->dtype[0] != MT_F32 || tl->dtype[1] != MT_F32)
        return set_error(MT_EINVAL, "multi_tensor_lamb_update_weights: the update term and p must be float32");
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    if (tl->dtype[2] == MT_E5M2) {  // reference DISPATCH_FLOAT_HALF_AND_BYTE (:489): the e5m2_allgather packing
        using Op = DistLambStage2Op<e5m2_t>;
        typename Op::Params prm{per_tensor_param_norm, per_tensor_update_norm, update_norm_offset, learning_rate,
                                per_tensor_decay, use_nvlamb};
        return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream);
    }
    MT_DISPATCH_FHB(tl->dtype[2], Tc, "multi_tensor_lamb_update_weights", {
        using Op = DistLambStage2Op<Tc>;
        typename Op::Params prm{per_tensor_param_norm, per_tensor_update_norm, update_norm_offset, learning_rate,
                                per_tensor_decay, use_nvlamb};
        rc = launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream);
    })
    return rc;
}
