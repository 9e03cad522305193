// mt_api_optim.cu -- optimizer entry points: Adam, mixed-precision (sharded) Adam, SGD, LAMB, legacy split LAMB.
#include <cmath>
#include <cstdlib>

#include "mt_host.h"
#include "mt_ops.cuh"

namespace mt {

// Host-side bias corrections: /root/reference/csrc/multi_tensor_adam.cu:145-149.  std::pow(float, int)
// promotes to double; the result is rounded to float.
static void bias_corrections(float beta1, float beta2, int step, int bias_correction, float* bc1, float* bc2) {
    *bc1 = 1.0f;
    *bc2 = 1.0f;
    if (bias_correction == 1) {
        *bc1 = (float)(1.0 - std::pow((double)beta1, (double)step));
        *bc2 = (float)(1.0 - std::pow((double)beta2, (double)step));
    }
}

// LAMB stage 1 keeps two norm accumulators live across the tile's rounds: under the 64-register cap of 4 CTAs/SM ptxas
// spills a few values (LDL/STL at round boundaries; profiles/r2_sass_summary.txt), at 80 registers (3 CTAs/SM) it does
// not.  Measured on BERT-large (profiles/r2_lamb_minb_ab.jsonl): multi_tensor_lamb 2.249 ms at 4 CTAs/SM vs 2.295 ms
// at 3 -- occupancy is worth more than the spills, so 4 stays the default for it.  MT_LAMB_STAGE1_MINB=3|4 forces
// either build of both stage-1 kernels for A/B timing.
static int lamb_minb_override() {
    static const int v = [] {
        const char* e = getenv("MT_LAMB_STAGE1_MINB");
        const int x = e ? atoi(e) : 0;
        return (x == 3 || x == 4) ? x : 0;
    }();
    return v;
}
static bool lamb_stage1_wide() { return lamb_minb_override() == 3; }
static bool lamb_mp_stage1_wide() { return lamb_minb_override() != 4; }

static inline size_t align256(size_t x) { return (x + 255) / 256 * 256; }

struct LambScratch {
    size_t off_partials, off_pnorm, off_unorm, total;
};
static LambScratch lamb_scratch_layout(uint64_t ntiles, int64_t ntensors) {
    LambScratch s;
    s.off_partials = 0;
    s.off_pnorm = align256(2 * (size_t)ntiles * sizeof(float));
    s.off_unorm = s.off_pnorm + (size_t)ntensors * sizeof(float);  // contiguous with pnorm: finalize writes [2][ntensors]
    s.total = align256(s.off_unorm + (size_t)ntensors * sizeof(float));
    return s;
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

// ---------------------------------------------------------------------------------------------------
extern "C" int mt_multi_tensor_adam(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float lr, float beta1,
                                    float beta2, float epsilon, int step, int mode, int bias_correction,
                                    float weight_decay, void* stream_) {
    MT_PROLOGUE("multi_tensor_adam", 4, 4)
    for (int d = 1; d < 4; d++)
        if (tl->dtype[d] != tl->dtype[0])
            return set_error(MT_EINVAL, "multi_tensor_adam: g, p, m, v must share one dtype (as the reference assumes)");
    float bc1, bc2;
    bias_corrections(beta1, beta2, step, bias_correction, &bc1, &bc2);
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    // (double lists: the 64-bit conversions need more registers than the 64 of 4 CTAs/SM leave -> 3 CTAs/SM, no spills)
#define MT_ADAM(CT, MINB)                                                                                 \
    {                                                                                                     \
        if (mode == 0) {                                                                                  \
            using Op = AdamOp<CT, 0>;                                                                     \
            typename Op::Params prm{beta1, beta2, bc1, bc2, epsilon, lr, weight_decay};                   \
            return launch_table_kernel<Op, kG, kU, kThreads, MINB>(tb, prm, noop_flag, tl->device, stream); \
        }                                                                                                 \
        using Op = AdamOp<CT, 1>;                                                                         \
        typename Op::Params prm{beta1, beta2, bc1, bc2, epsilon, lr, weight_decay};                       \
        return launch_table_kernel<Op, kG, kU, kThreads, MINB>(tb, prm, noop_flag, tl->device, stream);   \
    }
    switch (tl->dtype[0]) {
        case MT_F32: MT_ADAM(float, kMinBlocks)
        case MT_F16: MT_ADAM(__half, kMinBlocks)
        case MT_BF16: MT_ADAM(__nv_bfloat16, kMinBlocks)
        case MT_F64: MT_ADAM(double, 3)
        default: return set_error(MT_EINVAL, "multi_tensor_adam: unsupported dtype tag %d", tl->dtype[0]);
    }
#undef MT_ADAM
}

// ---------------------------------------------------------------------------------------------------
static int fused_adam_impl(const char* who, int32_t* noop_flag, const MtTensorLists* tl, const float* per_tensor_beta1,
                           const float* per_tensor_beta2, const int32_t* per_tensor_bias_correction,
                           const float* per_tensor_eps, const float* per_tensor_weight_decay, float lr, float grad_scale,
                           int step, int mode, int honour_noop, const float* lr_dev, const float* grad_scale_dev,
                           const int32_t* step_dev, void* stream_) {
    using namespace mt;
    int rc = validate_lists(tl, 4, 5, who);
    if (rc) return rc;
    if (!noop_flag) return set_error(MT_EINVAL, "%s: noop_flag is NULL", who);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    DeviceGuard guard(tl->device);
    if (!guard.ok) return set_error(MT_ECUDA, "%s: cannot select CUDA device %d", who, tl->device);
    if (!per_tensor_beta1 || !per_tensor_beta2 || !per_tensor_bias_correction || !per_tensor_eps || !per_tensor_weight_decay)
        return set_error(MT_EINVAL, "%s: per-tensor hyper-parameter arrays must be device pointers", who);
    for (int d = 0; d < 3; d++)
        if (tl->dtype[d] != MT_F32) return set_error(MT_EINVAL, "%s: p, m, v must be float32", who);
    if (tl->depth == 5 && tl->dtype[4] != tl->dtype[3])
        return set_error(MT_EINVAL, "%s: p_copy must have the gradient dtype", who);
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    const bool copy = tl->depth == 5;
#define MT_DADAM(TG, COPY, HN)                                                                            \
    {                                                                                                     \
        using Op = DistAdamOp<TG, COPY, HN>;                                                              \
        typename Op::Params prm{per_tensor_beta1, per_tensor_beta2, per_tensor_eps, per_tensor_weight_decay, \
                                per_tensor_bias_correction, lr, grad_scale, step, mode, lr_dev, grad_scale_dev, step_dev}; \
        return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream); \
    }
    MT_DISPATCH_FHB(tl->dtype[3], Tg, who, {
        if (copy) { if (honour_noop) MT_DADAM(Tg, true, true) else MT_DADAM(Tg, true, false) }
        else      { if (honour_noop) MT_DADAM(Tg, false, true) else MT_DADAM(Tg, false, false) }
    })
#undef MT_DADAM
    return MT_OK;
}

extern "C" int mt_multi_tensor_fused_adam(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                          const float* per_tensor_beta1, const float* per_tensor_beta2,
                                          const int32_t* per_tensor_bias_correction, const float* per_tensor_eps,
                                          const float* per_tensor_weight_decay, float lr, float grad_scale, int step,
                                          int mode, int honour_noop, void* stream_) {
    (void)chunk_size;
    return fused_adam_impl("multi_tensor_fused_adam", noop_flag, tl, per_tensor_beta1, per_tensor_beta2,
                           per_tensor_bias_correction, per_tensor_eps, per_tensor_weight_decay, lr, grad_scale, step, mode,
                           honour_noop, nullptr, nullptr, nullptr, stream_);
}

extern "C" int mt_multi_tensor_fused_adam_capturable(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                                     const float* per_tensor_beta1, const float* per_tensor_beta2,
                                                     const int32_t* per_tensor_bias_correction,
                                                     const float* per_tensor_eps, const float* per_tensor_weight_decay,
                                                     const float* lr, const float* grad_scale, const int32_t* step,
                                                     int mode, void* stream_) {
    (void)chunk_size;
    if (!lr || !grad_scale || !step)
        return mt::set_error(MT_EINVAL, "multi_tensor_fused_adam_capturable: lr, grad_scale, step must be device pointers");
    return fused_adam_impl("multi_tensor_fused_adam_capturable", noop_flag, tl, per_tensor_beta1, per_tensor_beta2,
                           per_tensor_bias_correction, per_tensor_eps, per_tensor_weight_decay, 0.f, 1.f, 0, mode, 1, lr,
                           grad_scale, step, stream_);
}

// ---------------------------------------------------------------------------------------------------
extern "C" int mt_multi_tensor_sgd(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float wd,
                                   float momentum, float dampening, float lr, int nesterov, int first_run,
                                   int wd_after_momentum, float scale, void* stream_) {
    MT_PROLOGUE("multi_tensor_sgd", 3, 4)
    const int gt = tl->dtype[0], wt = tl->dtype[1];
    if (tl->dtype[2] != wt) return set_error(MT_EINVAL, "multi_tensor_sgd: momentum dtype must equal the weight dtype");
    if (tl->depth == 4 && tl->dtype[3] != MT_F16)
        return set_error(MT_EINVAL, "Additional output tensors should always be fp16.");
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
#define MT_SGD(TG, TW, COPY)                                                                              \
    {                                                                                                     \
        using Op = SgdOp<TG, TW, COPY>;                                                                   \
        typename Op::Params prm{wd, momentum, dampening, lr, scale, nesterov, first_run, wd_after_momentum}; \
        return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream); \
    }
    // the reference's six combinations, /root/reference/csrc/multi_tensor_sgd_kernel.cu:165-314
    if (gt == MT_F16 && wt == MT_F16 && tl->depth == 3) MT_SGD(__half, __half, false)
    if (gt == MT_F32 && wt == MT_F32 && tl->depth == 3) MT_SGD(float, float, false)
    if (gt == MT_F16 && wt == MT_F32 && tl->depth == 4) MT_SGD(__half, float, true)
    if (gt == MT_F32 && wt == MT_F32 && tl->depth == 4) MT_SGD(float, float, true)
    if (gt == MT_BF16 && wt == MT_BF16 && tl->depth == 3) MT_SGD(__nv_bfloat16, __nv_bfloat16, false)
    if (gt == MT_BF16 && wt == MT_F32 && tl->depth == 4) MT_SGD(__nv_bfloat16, float, true)
#undef MT_SGD
    return set_error(MT_EINVAL,
                     "multi_tensor_sgd only supports some combinations of gradient & weight types. Given: "
                     "gradient: %d, weight: %d, num_lists: %d", gt, wt, tl->depth);
}

// ---------------------------------------------------------------------------------------------------
extern "C" size_t mt_lamb_scratch_bytes(int chunk_size, const MtTensorLists* tl) {
    (void)chunk_size;
    if (!tl || !tl->numel || tl->ntensors <= 0) return 0;
    const int tile = mt::choose_tile_elems(tl, tl->device);
    return mt::lamb_scratch_layout(mt::count_tiles(tl, tile), tl->ntensors).total;
}

extern "C" int mt_multi_tensor_lamb(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float lr, float beta1,
                                    float beta2, float epsilon, int step, int bias_correction, float weight_decay,
                                    int grad_averaging, int mode, const float* global_grad_norm, float max_grad_norm,
                                    int use_nvlamb, void* scratch, size_t scratch_bytes, void* stream_) {
    MT_PROLOGUE("multi_tensor_lamb", 4, 4)
    if (!global_grad_norm) return set_error(MT_EINVAL, "multi_tensor_lamb: global_grad_norm must be a device pointer");
    for (int d = 1; d < 4; d++)
        if (tl->dtype[d] != tl->dtype[0])
            return set_error(MT_EINVAL, "multi_tensor_lamb: g, p, m, v must share one dtype (as the reference assumes)");
    float bc1, bc2;
    bias_corrections(beta1, beta2, step, bias_correction, &bc1, &bc2);
    float beta3 = 1.0f;
    if (grad_averaging == 1) beta3 = 1 - beta1;  // csrc/multi_tensor_lamb.cu:363-364
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    const LambScratch lay = lamb_scratch_layout(tb.ntiles, tl->ntensors);
    if (!scratch || scratch_bytes < lay.total)
        return set_error(MT_ESCRATCH, "multi_tensor_lamb: scratch too small (%zu < %zu bytes)", scratch_bytes, lay.total);
    char* sc = static_cast<char*>(scratch);
    float* partials = reinterpret_cast<float*>(sc + lay.off_partials);
    float* pnorm = reinterpret_cast<float*>(sc + lay.off_pnorm);
    float* unorm = reinterpret_cast<float*>(sc + lay.off_unorm);
    // The per-tensor norms only feed the trust ratio, which stage 2 ignores unless use_nvlamb || decay != 0
    // (csrc/multi_tensor_lamb.cu:255-263): skip both norm passes otherwise.
    const int need_norms = (use_nvlamb || weight_decay != 0.0f) ? 1 : 0;

    MT_DISPATCH_FHB(tl->dtype[0], Tp, "multi_tensor_lamb", {
        if (mode == 0) {
            using Op1 = LambStage1Op<Tp, 0>;
            typename Op1::Params p1{beta1, beta2, beta3, bc1, bc2, epsilon, weight_decay, max_grad_norm,
                                    global_grad_norm, NormOut{partials, nullptr, pnorm, 1}, mode, need_norms};
            rc = lamb_stage1_wide() ? launch_table_kernel<Op1, kG, kU, kThreads, 3>(tb, p1, noop_flag, tl->device, stream)
                                    : launch_table_kernel<Op1, kG, kU, kThreads, kMinBlocks>(tb, p1, noop_flag, tl->device, stream);
        } else {
            using Op1 = LambStage1Op<Tp, 1>;
            typename Op1::Params p1{beta1, beta2, beta3, bc1, bc2, epsilon, weight_decay, max_grad_norm,
                                    global_grad_norm, NormOut{partials, nullptr, pnorm, 1}, mode, need_norms};
            rc = lamb_stage1_wide() ? launch_table_kernel<Op1, kG, kU, kThreads, 3>(tb, p1, noop_flag, tl->device, stream)
                                    : launch_table_kernel<Op1, kG, kU, kThreads, kMinBlocks>(tb, p1, noop_flag, tl->device, stream);
        }
        if (rc) return rc;
        // (the per-tensor norms were finalized inside the stage-1 launch: norm_tile_done)
        // stage 2 walks lists [u (= g), p] of the SAME table: list indices 0 and 1
        using Op2 = LambStage2Op<Tp>;
        typename Op2::Params p2{pnorm, unorm, lr, weight_decay, use_nvlamb};
        return launch_table_kernel<Op2, kG, kU, kThreads, kMinBlocks>(tb, p2, noop_flag, tl->device, stream);
    })
    return MT_OK;
}

// ---------------------------------------------------------------------------------------------------
extern "C" int mt_multi_tensor_lamb_mp(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, const float* lr,
                                       float beta1, float beta2, float epsilon, const int32_t* step, int bias_correction,
                                       float weight_decay, int grad_averaging, int mode, const float* global_grad_norm,
                                       const float* max_grad_norm, int use_nvlamb, const float* found_inf,
                                       const float* inv_scale, void* scratch, size_t scratch_bytes, void* stream_) {
    MT_PROLOGUE("multi_tensor_lamb_mp", 4, 5)
    (void)found_inf;  // carried for signature compatibility; the reference functor never reads it either
    if (!lr || !step || !global_grad_norm || !max_grad_norm || !inv_scale)
        return set_error(MT_EINVAL, "multi_tensor_lamb_mp: lr, step, global_grad_norm, max_grad_norm, inv_scale must be device pointers");
    const int gt = tl->dtype[0], pt = tl->dtype[1];
    if (tl->dtype[2] != pt || tl->dtype[3] != pt) return set_error(MT_EINVAL, "multi_tensor_lamb_mp: p, m, v must share one dtype");
    const bool copy = tl->depth == 5;
    if (copy && (pt != MT_F32 || tl->dtype[4] != gt))
        return set_error(MT_EINVAL, "multi_tensor_lamb_mp: with 5 lists p/m/v must be float32 masters and list 4 the gradient dtype");
    if (!copy && pt != gt) return set_error(MT_EINVAL, "multi_tensor_lamb_mp: with 4 lists all lists must share one dtype");
    float beta3 = 1.0f;
    if (grad_averaging == 1) beta3 = 1 - beta1;
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    const LambScratch lay = lamb_scratch_layout(tb.ntiles, tl->ntensors);
    if (!scratch || scratch_bytes < lay.total)
        return set_error(MT_ESCRATCH, "multi_tensor_lamb_mp: scratch too small (%zu < %zu bytes)", scratch_bytes, lay.total);
    char* sc = static_cast<char*>(scratch);
    float* partials = reinterpret_cast<float*>(sc + lay.off_partials);
    float* pnorm = reinterpret_cast<float*>(sc + lay.off_pnorm);
    float* unorm = reinterpret_cast<float*>(sc + lay.off_unorm);
    const int need_norms = (use_nvlamb || weight_decay != 0.0f) ? 1 : 0;
#define MT_LAMB_MP(TG, TP, COPY)                                                                                    \
    {                                                                                                               \
        if (mode == 0) {                                                                                            \
            using Op1 = LambMpStage1Op<TG, TP, 0>;                                                                  \
            typename Op1::Params p1{beta1, beta2, beta3, epsilon, weight_decay, step, global_grad_norm,             \
                                    max_grad_norm, inv_scale, NormOut{partials, nullptr, pnorm, 1}, mode, bias_correction, need_norms}; \
            rc = lamb_mp_stage1_wide() ? launch_table_kernel<Op1, kG, kU, kThreads, 3>(tb, p1, noop_flag, tl->device, stream) \
                                    : launch_table_kernel<Op1, kG, kU, kThreads, kMinBlocks>(tb, p1, noop_flag, tl->device, stream); \
        } else {                                                                                                    \
            using Op1 = LambMpStage1Op<TG, TP, 1>;                                                                  \
            typename Op1::Params p1{beta1, beta2, beta3, epsilon, weight_decay, step, global_grad_norm,             \
                                    max_grad_norm, inv_scale, NormOut{partials, nullptr, pnorm, 1}, mode, bias_correction, need_norms}; \
            rc = lamb_mp_stage1_wide() ? launch_table_kernel<Op1, kG, kU, kThreads, 3>(tb, p1, noop_flag, tl->device, stream) \
                                    : launch_table_kernel<Op1, kG, kU, kThreads, kMinBlocks>(tb, p1, noop_flag, tl->device, stream); \
        }                                                                                                           \
        if (rc) return rc;                                                                                          \
        using Op2 = LambMpStage2Op<TG, TP, COPY>;                                                                   \
        typename Op2::Params p2{pnorm, unorm, lr, weight_decay, use_nvlamb};                                        \
        return launch_table_kernel<Op2, kG, kU, kThreads, kMinBlocks>(tb, p2, noop_flag, tl->device, stream);       \
    }
    // reference dispatch: DISPATCH_FLOAT_AND_HALF on the gradient type; bf16 is accepted here as well
    if (copy) {
        if (gt == MT_F16) MT_LAMB_MP(__half, float, true)
        if (gt == MT_BF16) MT_LAMB_MP(__nv_bfloat16, float, true)
        if (gt == MT_F32) MT_LAMB_MP(float, float, true)
    } else {
        if (gt == MT_F16) MT_LAMB_MP(__half, __half, false)
        if (gt == MT_BF16) MT_LAMB_MP(__nv_bfloat16, __nv_bfloat16, false)
        if (gt == MT_F32) MT_LAMB_MP(float, float, false)
    }
#undef MT_LAMB_MP
    return set_error(MT_EINVAL, "multi_tensor_lamb_mp: unsupported gradient dtype tag %d", gt);
}

// ---------------------------------------------------------------------------------------------------
extern "C" int mt_multi_tensor_lamb_stage1(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                           const float* per_tensor_decay, int step, float beta1, float beta2,
                                           float epsilon, float global_grad_norm, float max_global_grad_norm,
                                           void* stream_) {
    MT_PROLOGUE("multi_tensor_lamb_stage1_cuda", 5, 5)
    if (!per_tensor_decay) return set_error(MT_EINVAL, "multi_tensor_lamb_stage1_cuda: per_tensor_decay is NULL");
    if (tl->dtype[2] != tl->dtype[1] || tl->dtype[3] != tl->dtype[1])
        return set_error(MT_EINVAL, "multi_tensor_lamb_stage1_cuda: p, m, v must share one dtype");
    // csrc/multi_tensor_lamb_stage_1.cu:126-130
    const float clip = global_grad_norm > max_global_grad_norm ? global_grad_norm / max_global_grad_norm : 1.0f;
    const float next_step = (float)(step + 1);
    const float bc1 = 1.0f - std::pow(beta1, next_step);
    const float bc2 = 1.0f - std::pow(beta2, next_step);
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    MT_DISPATCH_FHB(tl->dtype[0], Tg, "lamb_stage_1",
        MT_DISPATCH_FHB(tl->dtype[1], Tp, "lamb_stage_1",
            MT_DISPATCH_FHB(tl->dtype[4], Tu, "lamb_stage_1", {
                using Op = LambStage1LegacyOp<Tg, Tp, Tu>;
                typename Op::Params prm{per_tensor_decay, beta1, beta2, bc1, bc2, epsilon, clip};
                return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream);
            })))
    return MT_OK;
}

extern "C" int mt_multi_tensor_lamb_stage2(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                           const float* per_tensor_param_norm, const float* per_tensor_update_norm,
                                           float lr, float weight_decay, int use_nvlamb, void* stream_) {
    MT_PROLOGUE("multi_tensor_lamb_stage2_cuda", 2, 2)
    if (!per_tensor_param_norm || !per_tensor_update_norm)
        return set_error(MT_EINVAL, "multi_tensor_lamb_stage2_cuda: norm arrays are NULL");
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    MT_DISPATCH_FHB(tl->dtype[0], Tp, "lamb_stage_2",
        MT_DISPATCH_FHB(tl->dtype[1], Tu, "lamb_stage_2", {
            using Op = LambStage2LegacyOp<Tp, Tu>;
            typename Op::Params prm{per_tensor_param_norm, per_tensor_update_norm, lr, weight_decay, use_nvlamb};
            return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream);
        }))
    return MT_OK;
}
