// mt_api_scale.cu -- multi_tensor_scale and multi_tensor_axpby entry points.
#include "mt_host.h"
#include "mt_ops.cuh"

extern "C" int mt_multi_tensor_scale(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float scale,
                                     void* stream_) {
    using namespace mt;
    (void)chunk_size;
    int rc = validate_lists(tl, 2, 2, "multi_tensor_scale");
    if (rc) return rc;
    if (!noop_flag) return set_error(MT_EINVAL, "multi_tensor_scale: noop_flag is NULL");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    DeviceGuard guard(tl->device);
    if (!guard.ok) return set_error(MT_ECUDA, "multi_tensor_scale: cannot select CUDA device %d", tl->device);
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    MT_DISPATCH_FHB(tl->dtype[0], Tin, "multi_tensor_scale",
        MT_DISPATCH_FHB(tl->dtype[1], Tout, "multi_tensor_scale", {
            using Op = ScaleOp<Tin, Tout>;
            typename Op::Params prm{scale};
            return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream);
        }))
    return MT_OK;
}

extern "C" int mt_multi_tensor_maybe_cast(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, void* stream_) {
    using namespace mt;
    (void)chunk_size;
    int rc = validate_lists(tl, 2, 2, "maybe_cast_mt");
    if (rc) return rc;
    if (!noop_flag) return set_error(MT_EINVAL, "maybe_cast_mt: overflow_flag is NULL");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    DeviceGuard guard(tl->device);
    if (!guard.ok) return set_error(MT_ECUDA, "maybe_cast_mt: cannot select CUDA device %d", tl->device);
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    if (tl->dtype[0] == MT_E5M2) {  // unpack of the e5m2 all-gather (fused_adam_cuda_kernel.cu:417-463)
        MT_DISPATCH_FHB(tl->dtype[1], Tout, "maybe_cast_mt", {
            using Op = CastOp<e5m2_t, Tout>;
            typename Op::Params prm{0};
            return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream);
        })
    }
    MT_DISPATCH_FHB(tl->dtype[0], Tin, "maybe_cast_mt",
        MT_DISPATCH_FHB(tl->dtype[1], Tout, "maybe_cast_mt", {
            using Op = CastOp<Tin, Tout>;
            typename Op::Params prm{0};
            return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream);
        }))
    return MT_OK;
}

extern "C" int mt_multi_tensor_axpby(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float a, float b,
                                     int arg_to_check, void* stream_) {
    using namespace mt;
    (void)chunk_size;
    int rc = validate_lists(tl, 3, 3, "multi_tensor_axpby");
    if (rc) return rc;
    if (!noop_flag) return set_error(MT_EINVAL, "multi_tensor_axpby: noop_flag is NULL");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    DeviceGuard guard(tl->device);
    if (!guard.ok) return set_error(MT_ECUDA, "multi_tensor_axpby: cannot select CUDA device %d", tl->device);
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    MT_DISPATCH_FHB(tl->dtype[0], Tx, "multi_tensor_axpby",
        MT_DISPATCH_FHB(tl->dtype[1], Ty, "multi_tensor_axpby",
            MT_DISPATCH_FHB(tl->dtype[2], To, "multi_tensor_axpby", {
                using Op = AxpbyOp<Tx, Ty, To>;
                typename Op::Params prm{a, b, arg_to_check};
                return launch_table_kernel<Op, kG, kU, kThreads, kMinBlocks>(tb, prm, noop_flag, tl->device, stream);
            })))
    return MT_OK;
}
