// mt_api_more.cu -- the remaining optimizers exported by the reference's amp_C module:
// Adagrad, NovoGrad (with its per-tensor norm blend) and LARS.
#include <cmath>

#include "mt_host.h"
#include "mt_ops.cuh"

namespace mt {

constexpr int kBlendThreads = 256;

// cleanup_v2, per-tensor part (/root/reference/csrc/multi_tensor_l2norm_kernel.cu:272-298): blend this step's norm
// into the running per-tensor norm.  norm_type 0: out = a*out + b*max|g|;  else: out = sqrt(a*out^2 + b*sum g^2).
// One warp per tensor, lanes stride the tensor's chunk partials, fixed shuffle tree.
__global__ void __launch_bounds__(kBlendThreads)
norm_blend_kernel(const float* __restrict__ partials, const uint32_t* __restrict__ tile_begin, int ntensors,
                  float* __restrict__ out, int norm_type, float alpha, float beta) {
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * kBlendThreads) >> 5;
    for (int t = (blockIdx.x * kBlendThreads + threadIdx.x) >> 5; t < ntensors; t += warps) {
        const uint32_t b = tile_begin[t], e = tile_begin[t + 1];
        float acc = 0.f;
        if (norm_type == 0) {
            for (uint32_t i = b + lane; i < e; i += 32) acc = fmaxf(fabsf(acc), fabsf(partials[i]));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc = fmaxf(acc, __shfl_xor_sync(0xffffffffu, acc, o));
            if (lane == 0) out[t] = fadd(fmul(alpha, out[t]), fmul(beta, acc));
        } else {
            for (uint32_t i = b + lane; i < e; i += 32) acc = fadd(acc, partials[i]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc = fadd(acc, __shfl_xor_sync(0xffffffffu, acc, o));
            if (lane == 0) out[t] = fsqrt(fadd(fmul(fmul(alpha, out[t]), out[t]), fmul(beta, acc)));
        }
    }
}

}  // namespace mt

#define MT_PROLOGUE(who, lo, hi)                                                                    \
    using namespace mt;                                                                             \
    int rc = validate_lists(tl, lo, hi, who);                                                       \
    if (rc) return rc;                                                                              \
    if (!noop_flag) return set_error(MT_EINVAL, who ": noop_flag is NULL");                         \
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);                                       \
    DeviceGuard guard(tl->device);                                                                  \
    if (!guard.ok) return set_error(MT_ECUDA, who ": cannot select CUDA device %d", tl->device);

#define MT_DISPATCH_FHBD(tag, NAME, who, ...)                                         \
    switch (tag) {                                                                    \
        MT_CASE(MT_F32, float, NAME, __VA_ARGS__)                                     \
        MT_CASE(MT_F16, __half, NAME, __VA_ARGS__)                                    \
        MT_CASE(MT_BF16, __nv_bfloat16, NAME, __VA_ARGS__)                            \
        MT_CASE(MT_F64, double, NAME, __VA_ARGS__)                                    \
        default:                                                                      \
            return ::mt::set_error(MT_EINVAL, "%s: unsupported dtype tag %d", who, (int)(tag)); \
    }

extern "C" int mt_multi_tensor_adagrad(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float lr,
                                       float epsilon, int mode, float weight_decay, void* stream_) {
    (void)chunk_size;
    MT_PROLOGUE("multi_tensor_adagrad", 3, 3)
    if (tl->dtype[1] != tl->dtype[0] || tl->dtype[2] != tl->dtype[0])
        return set_error(MT_EINVAL, "multi_tensor_adagrad: g, p, h must share one dtype (as the reference assumes)");
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    MT_DISPATCH_FHBD(tl->dtype[0], Tp, "multi_tensor_adagrad", {
        using Op = AdagradOp<Tp>;
        typename Op::Params prm{epsilon, lr, weight_decay, mode};
        return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream);
    })
    return MT_OK;
}

extern "C" int mt_multi_tensor_novograd(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float* grad_norms,
                                        float lr, float beta1, float beta2, float epsilon, int step, int bias_correction,
                                        float weight_decay, int grad_averaging, int moment_mode, int norm_type,
                                        void* scratch, size_t scratch_bytes, void* stream_) {
    MT_PROLOGUE("multi_tensor_novograd", 3, 3)
    if (chunk_size <= 0) return set_error(MT_EINVAL, "multi_tensor_novograd: chunk_size must be positive");
    if (!grad_norms) return set_error(MT_EINVAL, "multi_tensor_novograd: grad_norms must be a device float[ntensors]");
    if (tl->dtype[1] != tl->dtype[0] || tl->dtype[2] != tl->dtype[0])
        return set_error(MT_EINVAL, "multi_tensor_novograd: g, p, m must share one dtype (as the reference assumes)");
    // host part, /root/reference/csrc/multi_tensor_novograd.cu:145-156
    float bc1 = 1.0f, bc2 = 1.0f;
    if (bias_correction == 1) {
        bc1 = (float)(1.0 - std::pow((double)beta1, (double)step));
        bc2 = (float)std::sqrt(1.0 - std::pow((double)beta2, (double)step));
    }
    float beta3 = 1;
    if (grad_averaging == 1) beta3 = 1 - beta1;

    // 1. multi_tensor_norm_out_cuda(grad_list, grad_norms, beta2, 1 - beta2, norm_type): per-chunk partials of the
    //    gradient list (tile == chunk, so the flag has the reference's granularity), then the per-tensor blend.
    MtTensorLists gl = *tl;
    gl.depth = 1;
    TableRef tn;
    rc = get_table(&gl, chunk_size, stream, &tn);
    if (rc) return rc;
    const size_t need = (size_t)tn.ntiles * sizeof(float);
    if (need > 0 && (!scratch || scratch_bytes < need))
        return set_error(MT_ESCRATCH, "multi_tensor_novograd: scratch too small (%zu < %zu bytes)", scratch_bytes, need);
    float* partials = static_cast<float*>(scratch);
    constexpr int G = 8, U = 2, THREADS = 256, MINB = 4;
    MT_DISPATCH_FHBD(tl->dtype[0], Tx, "multi_tensor_novograd", {
        if (norm_type == 0) {
            using Op = MaxNormOp<Tx>;
            typename Op::Params prm{partials};
            rc = launch_table_kernel<Op, G, U, THREADS, MINB>(tn, prm, noop_flag, tl->device, stream);
        } else {
            using Op = L2NormOp<Tx, false>;
            typename Op::Params prm{NormOut{partials, nullptr, nullptr, 1}};
            rc = launch_table_kernel<Op, G, U, THREADS, MINB>(tn, prm, noop_flag, tl->device, stream);
        }
    })
    if (rc) return rc;
    {
        DeviceInfo di;
        rc = device_info(tl->device, &di);
        if (rc) return rc;
        long long blocks = ((long long)tn.ntensors + (kBlendThreads / 32) - 1) / (kBlendThreads / 32);
        const long long cap = (long long)di.sms * 8;
        const int grid = (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
        norm_blend_kernel<<<grid, kBlendThreads, 0, stream>>>(partials, tn.tile_begin, tn.ntensors, grad_norms, norm_type,
                                                              beta2, 1.0f - beta2);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        MT_CUDA_OK(cudaGetLastError());
    }
    // 2. the update
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    MT_DISPATCH_FHBD(tl->dtype[0], Tp, "multi_tensor_novograd", {
        using Op = NovoGradOp<Tp>;
        typename Op::Params prm{beta1, beta2, beta3, bc1, bc2, epsilon, lr, weight_decay, moment_mode, grad_norms};
        return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream);
    })
    return MT_OK;
}

extern "C" int mt_multi_tensor_lars(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, const float* grad_norms,
                                    const float* param_norms, float lr, float trust_coefficient, float epsilon,
                                    float weight_decay, float momentum, float dampening, int nesterov, int first_run,
                                    int wd_after_momentum, float scale, int is_skipped, void* stream_) {
    (void)chunk_size; (void)dampening; (void)first_run; (void)wd_after_momentum; (void)scale;  // unused by the reference functor too
    MT_PROLOGUE("multi_tensor_lars", 3, 4)
    if (!is_skipped && (!grad_norms || !param_norms))
        return set_error(MT_EINVAL, "multi_tensor_lars: grad_norms / param_norms must be device float[ntensors]");
    const int gt = tl->dtype[0], wt = tl->dtype[1];
    if (tl->dtype[2] != wt) return set_error(MT_EINVAL, "multi_tensor_lars: momentum dtype must equal the weight dtype");
    if (tl->depth == 4 && tl->dtype[3] != MT_F16)
        return set_error(MT_EINVAL, "Additional output tensors should always be fp16.");
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
#define MT_LARS(TG, TW, COPY)                                                                                 \
    {                                                                                                         \
        using Op = LarsOp<TG, TW, COPY>;                                                                      \
        typename Op::Params prm{grad_norms, param_norms, lr, trust_coefficient, epsilon, weight_decay,        \
                                momentum, nesterov, is_skipped};                                              \
        return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream); \
    }
    // the reference's six combinations, /root/reference/csrc/multi_tensor_lars.cu:196-345
    if (gt == MT_F16 && wt == MT_F16 && tl->depth == 3) MT_LARS(__half, __half, false)
    if (gt == MT_F32 && wt == MT_F32 && tl->depth == 3) MT_LARS(float, float, false)
    if (gt == MT_F16 && wt == MT_F32 && tl->depth == 4) MT_LARS(__half, float, true)
    if (gt == MT_F32 && wt == MT_F32 && tl->depth == 4) MT_LARS(float, float, true)
    if (gt == MT_BF16 && wt == MT_BF16 && tl->depth == 3) MT_LARS(__nv_bfloat16, __nv_bfloat16, false)
    if (gt == MT_BF16 && wt == MT_F32 && tl->depth == 4) MT_LARS(__nv_bfloat16, float, true)
#undef MT_LARS
    return set_error(MT_EINVAL,
                     "multi_tensor_lars only supports some combinations of gradient & weight types. Given: "
                     "gradient: %d, weight: %d, num_lists: %d", gt, wt, tl->depth);
}
