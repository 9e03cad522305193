// mt_api_p2p.cu -- mt_multi_tensor_fused_adam_p2p: gradient reduce-scatter + sharded Adam + parameter all-gather in ONE
// kernel over NVLink peer memory, for the ZeRO-style step of SURVEY.md section 8(e).
//
// The NCCL route (reduce_scatter -> update -> all_gather, distributed_fused_adam.py in the reference, :668-742 and
// :977-1115) moves every gradient byte over NVLink into a local buffer, reads it back from HBM in the update, writes
// the updated parameters to HBM and reads them again to send them.  Here the update kernel itself loads the gradient
// shard straight from every peer (ld.global on peer-mapped addresses, summed in rank order) and stores the updated
// 16-bit parameters straight into every peer's buffer, tile by tile, so the link transfers overlap the HBM traffic of
// p / m / v and nothing is staged.  The caller provides the peer-mapped base addresses (symmetric allocations, e.g.
// torch.distributed._symmetric_memory) and a cross-GPU barrier before and after the launch.
#include <cstdlib>
#include <cstring>

#include "mt_host.h"
#include "mt_ops.cuh"

namespace mt {

template <class Tg, int WORLD, int kRoute, int MINB>
static int launch_p2p(const DevTable& tb, const typename DistAdamOp<Tg, true, true>::Params& base, int rank,
                      const void* const* grad_bases, void* const* param_bases, const void* grad_mc, void* param_mc,
                      int ld_p2p_of16, int st_p2p_of16, int32_t* noop, int device, cudaStream_t stream) {
    using Op = DistAdamP2POp<Tg, WORLD, kRoute>;
    typename Op::Params prm;
    prm.base = base;
    prm.gown = static_cast<const char*>(grad_bases[rank]);
    prm.pown = static_cast<char*>(param_bases[rank]);
    prm.gmc = static_cast<const char*>(grad_mc);
    prm.pmc = static_cast<char*>(param_mc);
    prm.ld_p2p_of16 = ld_p2p_of16;
    prm.st_p2p_of16 = st_p2p_of16;
    for (int r = 0; r < WORLD; r++) {
        prm.gbase[r] = static_cast<const char*>(grad_bases[r]);
        prm.pbase[r] = static_cast<char*>(param_bases[r]);
    }
    return launch_table_kernel<Op, kG, kU, kThreads, MINB>(tb, prm, noop, device, stream);
}

// Occupancy: the peer stream needs many bytes in flight (NVLink latency is ~3.5x HBM's), so the register budget shrinks
// with the number of peers loaded per thread: 4 CTAs/SM for multicast and 2 ranks, 3 for 4 ranks, 2 for 8.  The hybrid
// kernel carries the peer-to-peer path and takes its budget.  Launch shape does not matter for the multicast variant:
// 1 or 2 groups in flight, 2 / 3 / 4 CTAs per SM and 4K..32K tiles all give 34.37 +- 0.02 ms at N=4
// (profiles/r1_sweep_p2p_n4.log) -- the NVSwitch reduction / multicast path sets the time, not the SMs.
template <class Tg>
static int dispatch_p2p(const DevTable& tb, const typename DistAdamOp<Tg, true, true>::Params& base, int world, int rank,
                        const void* const* gb, void* const* pb, const void* gmc, void* pmc, int route, int ld16,
                        int st16, int32_t* noop, int device, cudaStream_t stream) {
#define MT_P2P(W, ROUTE, MB) \
    return launch_p2p<Tg, W, ROUTE, MB>(tb, base, rank, gb, pb, gmc, pmc, ld16, st16, noop, device, stream)
    if (route == kRouteMc) {
        if (world == 2) MT_P2P(2, kRouteMc, 4);
        if (world == 4) MT_P2P(4, kRouteMc, 4);
        MT_P2P(8, kRouteMc, 4);
    }
    if (route == kRouteHybrid) {
        if (world == 2) MT_P2P(2, kRouteHybrid, 4);
        if (world == 4) MT_P2P(4, kRouteHybrid, 3);
        MT_P2P(8, kRouteHybrid, 2);
    }
    if (world == 2) MT_P2P(2, kRouteP2P, 4);
    if (world == 4) MT_P2P(4, kRouteP2P, 3);
    MT_P2P(8, kRouteP2P, 2);
#undef MT_P2P
}

static int fused_adam_p2p_impl(const char* who, int32_t* noop_flag, const MtTensorLists* tl, const float* per_tensor_beta1,
                               const float* per_tensor_beta2, const int32_t* per_tensor_bias_correction,
                               const float* per_tensor_eps, const float* per_tensor_weight_decay, float lr,
                               float grad_scale, int step, int mode, int world, int rank, const void* const* grad_bases,
                               void* const* param_bases, const void* grad_multicast, void* param_multicast,
                               int ld_p2p_of16, int st_p2p_of16, void* stream_) {
    int rc = validate_lists(tl, 5, 5, who);
    if (rc) return rc;
    if (!noop_flag) return set_error(MT_EINVAL, "%s: noop_flag is NULL", who);
    if (!per_tensor_beta1 || !per_tensor_beta2 || !per_tensor_bias_correction || !per_tensor_eps || !per_tensor_weight_decay)
        return set_error(MT_EINVAL, "%s: per-tensor hyper-parameter arrays must be device pointers", who);
    if (world != 2 && world != 4 && world != 8) return set_error(MT_EINVAL, "%s: world must be 2, 4 or 8 (got %d)", who, world);
    if (rank < 0 || rank >= world) return set_error(MT_EINVAL, "%s: rank %d out of range", who, rank);
    if (!grad_bases || !param_bases) return set_error(MT_EINVAL, "%s: peer base arrays are NULL", who);
    for (int r = 0; r < world; r++)
        if (!grad_bases[r] || !param_bases[r]) return set_error(MT_EINVAL, "%s: peer %d base pointer is NULL", who, r);
    if ((grad_multicast == nullptr) != (param_multicast == nullptr))
        return set_error(MT_EINVAL, "%s: give both multicast addresses or neither", who);
    if (ld_p2p_of16 < 0 || ld_p2p_of16 > 16 || st_p2p_of16 < 0 || st_p2p_of16 > 16)
        return set_error(MT_EINVAL, "%s: the peer-to-peer shares are counted in sixteenths (0..16)", who);
    for (int d = 0; d < 3; d++)
        if (tl->dtype[d] != MT_F32) return set_error(MT_EINVAL, "%s: p, m, v must be float32", who);
    if (tl->dtype[4] != tl->dtype[3]) return set_error(MT_EINVAL, "%s: p_copy must have the gradient dtype", who);
    if (tl->dtype[3] != MT_BF16 && tl->dtype[3] != MT_F16)
        return set_error(MT_EINVAL, "%s: gradients must be bf16 or fp16", who);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    DeviceGuard guard(tl->device);
    if (!guard.ok) return set_error(MT_ECUDA, "%s: cannot select CUDA device %d", who, tl->device);
    TableRef tb;
    rc = get_table(tl, choose_tile_elems(tl, tl->device), stream, &tb);
    if (rc) return rc;
    // route: no multicast address -> peer-to-peer; shares 0/0 -> pure multicast; 16/16 -> pure peer-to-peer; else hybrid
    int route = kRouteP2P;
    if (grad_multicast != nullptr) {
        if (ld_p2p_of16 == 0 && st_p2p_of16 == 0) route = kRouteMc;
        else if (ld_p2p_of16 == 16 && st_p2p_of16 == 16) route = kRouteP2P;
        else route = kRouteHybrid;
    }
    if (tl->dtype[3] == MT_BF16) {
        typename DistAdamOp<__nv_bfloat16, true, true>::Params base{per_tensor_beta1, per_tensor_beta2, per_tensor_eps,
                                                                    per_tensor_weight_decay, per_tensor_bias_correction,
                                                                    lr, grad_scale, step, mode};
        return dispatch_p2p<__nv_bfloat16>(tb, base, world, rank, grad_bases, param_bases, grad_multicast, param_multicast,
                                           route, ld_p2p_of16, st_p2p_of16, noop_flag, tl->device, stream);
    }
    typename DistAdamOp<__half, true, true>::Params base{per_tensor_beta1, per_tensor_beta2, per_tensor_eps,
                                                         per_tensor_weight_decay, per_tensor_bias_correction, lr, grad_scale,
                                                         step, mode};
    return dispatch_p2p<__half>(tb, base, world, rank, grad_bases, param_bases, grad_multicast, param_multicast, route,
                                ld_p2p_of16, st_p2p_of16, noop_flag, tl->device, stream);
}

}  // namespace mt

extern "C" int mt_multi_tensor_fused_adam_p2p(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                              const float* per_tensor_beta1, const float* per_tensor_beta2,
                                              const int32_t* per_tensor_bias_correction, const float* per_tensor_eps,
                                              const float* per_tensor_weight_decay, float lr, float grad_scale, int step,
                                              int mode, int world, int rank, const void* const* grad_bases,
                                              void* const* param_bases, const void* grad_multicast,
                                              void* param_multicast, void* stream_) {
    (void)chunk_size;
    return mt::fused_adam_p2p_impl("multi_tensor_fused_adam_p2p", noop_flag, tl, per_tensor_beta1, per_tensor_beta2,
                                   per_tensor_bias_correction, per_tensor_eps, per_tensor_weight_decay, lr, grad_scale, step,
                                   mode, world, rank, grad_bases, param_bases, grad_multicast, param_multicast, 0, 0, stream_);
}

extern "C" int mt_multi_tensor_fused_adam_p2p_hybrid(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                                     const float* per_tensor_beta1, const float* per_tensor_beta2,
                                                     const int32_t* per_tensor_bias_correction, const float* per_tensor_eps,
                                                     const float* per_tensor_weight_decay, float lr, float grad_scale,
                                                     int step, int mode, int world, int rank,
                                                     const void* const* grad_bases, void* const* param_bases,
                                                     const void* grad_multicast, void* param_multicast, int ld_p2p_of16,
                                                     int st_p2p_of16, void* stream_) {
    (void)chunk_size;
    return mt::fused_adam_p2p_impl("multi_tensor_fused_adam_p2p_hybrid", noop_flag, tl, per_tensor_beta1, per_tensor_beta2,
                                   per_tensor_bias_correction, per_tensor_eps, per_tensor_weight_decay, lr, grad_scale, step,
                                   mode, world, rank, grad_bases, param_bases, grad_multicast, param_multicast, ld_p2p_of16,
                                   st_p2p_of16, stream_);
}
