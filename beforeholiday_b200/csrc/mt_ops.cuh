// mt_ops.cuh -- the per-element functors (the "what"); mt_common.cuh is the "how".
//
// Each Op restates one reference functor in IEEE fp32, in the reference's operation order, with
// non-contractable arithmetic (mt::fadd/fmul/...), so that it is bit-identical to oracle/mt_oracle.c.
// Citations are file:line in /root/reference.
#pragma once

#include "mt_common.cuh"

namespace mt {

template <class A, class B = void, class C = void, class D = void, class E = void> struct Types {
    template <int I> using at = std::conditional_t<I == 0, A, std::conditional_t<I == 1, B,
                                std::conditional_t<I == 2, C, std::conditional_t<I == 3, D, E>>>>;
};

// ------------------------------------------------------------------------------------------------
// ScaleFunctor: csrc/multi_tensor_scale_kernel.cu:68-69,108-109.   lists: in, out
template <class Tin, class Tout> struct ScaleOp : OpBase {
    static constexpr int kLists = 2;
    template <int D> using T = typename Types<Tin, Tout>::template at<D>;
    static constexpr unsigned kRead = 0b01, kWrite = 0b10;
    struct Params { float scale; };
    using State = FiniteState;
    Params p;
    __device__ explicit ScaleOp(const Params& q) : p(q) {}
    __device__ __forceinline__ void init(State& s) const { s.finite = true; }
    __device__ __forceinline__ void elem(float (&x)[2], State& s) const {
        s.finite = s.finite && finite_f(x[0]);
        x[1] = fmul(x[0], p.scale);
    }
    __device__ __forceinline__ void finish(State& s, int32_t* noop) const { finish_finite(s, noop); }
};

// MaybeCastFunctor: apex/contrib/csrc/optimizers/fused_adam_cuda_kernel.cu:734-783 (fused_adam_cuda.maybe_cast_mt, the
// flat-buffer -> model-parameter copy of the sharded optimizers).  A converting copy that does nothing when the flag
// is set and never writes it.   lists: in, out
template <class Tin, class Tout> struct CastOp : OpBase {
    static constexpr int kLists = 2;
    template <int D> using T = typename Types<Tin, Tout>::template at<D>;
    static constexpr unsigned kRead = 0b01, kWrite = 0b10;
    static constexpr bool kEarlyExit = true;
    struct Params { int unused; };
    __device__ explicit CastOp(const Params&) {}
    __device__ __forceinline__ void elem(float (&x)[2], State&) const { x[1] = x[0]; }
};

// AxpbyFunctor: csrc/multi_tensor_axpby_kernel.cu:73-79.   lists: x, y, out
template <class Tx, class Ty, class To> struct AxpbyOp : OpBase {
    static constexpr int kLists = 3;
    template <int D> using T = typename Types<Tx, Ty, To>::template at<D>;
    static constexpr unsigned kRead = 0b011, kWrite = 0b100;
    struct Params { float a, b; int arg_to_check; };
    using State = FiniteState;
    Params p;
    __device__ explicit AxpbyOp(const Params& q) : p(q) {}
    __device__ __forceinline__ void init(State& s) const { s.finite = true; }
    __device__ __forceinline__ void elem(float (&x)[3], State& s) const {
        if (p.arg_to_check == -1) s.finite = s.finite && (finite_f(x[0]) && finite_f(x[1]));
        if (p.arg_to_check == 0) s.finite = s.finite && finite_f(x[0]);
        if (p.arg_to_check == 1) s.finite = s.finite && finite_f(x[1]);
        x[2] = fadd(fmul(p.a, x[0]), fmul(p.b, x[1]));
    }
    __device__ __forceinline__ void finish(State& s, int32_t* noop) const { finish_finite(s, noop); }
};

// ------------------------------------------------------------------------------------------------
// AdamFunctor: csrc/multi_tensor_adam.cu:94-112.   lists: g, p, m, v (one dtype, fp32 math)
// kMode is a template parameter (0 = L2, 1 = AdamW) so that only one branch is compiled in: with both in one kernel
// the 64-register budget of 4 CTAs/SM spilled.
template <class Tp, int kMode> struct AdamOp : OpBase {
    static constexpr int kLists = 4;
    template <int D> using T = Tp;
    static constexpr unsigned kRead = 0b1111, kWrite = 0b1110;
    struct Params { float beta1, beta2, bc1, bc2, eps, lr, decay; };
    Params p;
    float omb1, omb2;
    __device__ explicit AdamOp(const Params& q) : p(q), omb1(fsub(1.f, q.beta1)), omb2(fsub(1.f, q.beta2)) {}
    __device__ __forceinline__ void elem(float (&x)[4], State&) const {
        float g = x[0], w = x[1], m = x[2], v = x[3];
        if constexpr (kMode == 0) {  // L2
            g = fadd(g, fmul(p.decay, w));
            m = fadd(fmul(p.beta1, m), fmul(omb1, g));
            v = fadd(fmul(p.beta2, v), fmul(fmul(omb2, g), g));
            const float denom = fadd(fsqrt(fdiv(v, p.bc2)), p.eps);
            const float upd = fdiv(fdiv(m, p.bc1), denom);
            w = fsub(w, fmul(p.lr, upd));
        } else {  // AdamW
            m = fadd(fmul(p.beta1, m), fmul(omb1, g));
            v = fadd(fmul(p.beta2, v), fmul(fmul(omb2, g), g));
            const float denom = fadd(fsqrt(fdiv(v, p.bc2)), p.eps);
            const float upd = fadd(fdiv(fdiv(m, p.bc1), denom), fmul(p.decay, w));
            w = fsub(w, fmul(p.lr, upd));
        }
        x[1] = w; x[2] = m; x[3] = v;
    }
};

// DistAdamFunctor: apex/contrib/csrc/optimizers/multi_tensor_distopt_adam_kernel.cu:52-63,103-116.
// lists: p, m, v (fp32), g (Tg) [, p_copy (Tg)].  Per-tensor hyper-parameters live in device arrays.
template <class Tg, bool kCopy, bool kHonourNoop> struct DistAdamOp : OpBase {
    static constexpr bool kInlineTable = false;  // shard updates are never small launches (OpBase)
    static constexpr int kLists = kCopy ? 5 : 4;
    template <int D> using T = typename Types<float, float, float, Tg, Tg>::template at<D>;
    static constexpr unsigned kRead = 0b01111, kWrite = kCopy ? 0b10111 : 0b00111;
    static constexpr bool kEarlyExit = kHonourNoop;
    struct Params {
        const float *beta1, *beta2, *eps, *decay;
        const int32_t* bias_correction;
        float lr, grad_scale;
        int step, mode;
        // Capturable form: when non-NULL these DEVICE scalars replace lr / grad_scale / step, so a captured CUDA graph
        // replays correctly while the host-side values move on.
        const float* lr_dev;
        const float* grad_scale_dev;
        const int32_t* step_dev;
    };
    Params p;
    float b1, b2, omb1, omb2, bc1, bc2, eps, decay;
    int bc_on;
    __device__ explicit DistAdamOp(const Params& q) : p(q), b1(-1.f), b2(-1.f), bc1(1.f), bc2(1.f), bc_on(-1) {
        if (q.lr_dev) p.lr = __ldg(q.lr_dev);
        if (q.grad_scale_dev) p.grad_scale = __ldg(q.grad_scale_dev);
        if (q.step_dev) p.step = __ldg(q.step_dev);
    }
    __device__ __forceinline__ void begin_tensor(int t) {
        const float nb1 = __ldg(p.beta1 + t), nb2 = __ldg(p.beta2 + t);
        const int on = __ldg(p.bias_correction + t) == 1;
        eps = __ldg(p.eps + t); decay = __ldg(p.decay + t);
        // The bias corrections cost two double pow(): recompute only when the hyper-parameters change between
        // consecutive tiles of this CTA (they are almost always identical for every tensor).
        if (nb1 != b1 || nb2 != b2 || on != bc_on) {
            b1 = nb1; b2 = nb2; bc_on = on;
            omb1 = fsub(1.f, b1); omb2 = fsub(1.f, b2);
            bc1 = 1.f; bc2 = 1.f;
            if (on) {
                bc1 = (float)(1.0 - pow((double)b1, (double)p.step));
                bc2 = (float)(1.0 - pow((double)b2, (double)p.step));
            }
        }
    }
    __device__ __forceinline__ void elem(float (&x)[kLists], State&) const {
        float w = x[0], m = x[1], v = x[2];
        const float sg = fdiv(x[3], p.grad_scale);
        m = fadd(fmul(b1, m), fmul(omb1, sg));
        v = fadd(fmul(b2, v), fmul(fmul(omb2, sg), sg));
        const float mu = fdiv(m, bc1), vu = fdiv(v, bc2);
        const float denom = p.mode == 0 ? fsqrt(fadd(vu, eps)) : fadd(fsqrt(vu), eps);
        const float upd = fadd(fdiv(mu, denom), fmul(decay, w));
        w = fsub(w, fmul(p.lr, upd));
        x[0] = w; x[1] = m; x[2] = v;
        if constexpr (kCopy) x[4] = w;
    }
};

// Fused reduce-scatter + DistAdam + all-gather over peer memory (NVLink P2P through NVSwitch).  Same arithmetic as
// DistAdamOp; list 3 (gradients) is READ FROM EVERY RANK'S buffer and summed in rank order, list 4 (updated 16-bit
// params) is WRITTEN TO EVERY RANK'S buffer.  The tensor table holds this rank's own pointers; a peer's address is
// base[r] + (own pointer - base[rank]) because the buffers are symmetric allocations.  The sum is rounded to Tg once,
// which is what a reduce-scatter producing Tg gradients delivers, so the result matches the NCCL route bit for bit
// whenever the inputs' sum is exact in Tg.
template <class Tg> struct Packed16;  // two 16-bit elements in one 32-bit word
template <> struct Packed16<__nv_bfloat16> {
    __device__ static __forceinline__ void unpack(uint32_t w, float& lo, float& hi) {
        lo = __uint_as_float(w << 16);
        hi = __uint_as_float(w & 0xffff0000u);
    }
};
template <> struct Packed16<__half> {
    __device__ static __forceinline__ void unpack(uint32_t w, float& lo, float& hi) {
        const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w));
        lo = t.x;
        hi = t.y;
    }
};

// NVSwitch multicast (multimem.*) accesses on a multicast address: one instruction reduces across / stores to every
// GPU of the multicast group inside the switch (NVLS), so each gradient byte crosses this GPU's link once instead of
// once per peer.
template <class Tg> struct Multimem;
template <> struct Multimem<__nv_bfloat16> {
    __device__ static __forceinline__ void ld_reduce(const void* mc, uint32_t (&w)[4]) {
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                     : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(mc) : "memory");
    }
};
template <> struct Multimem<__half> {
    __device__ static __forceinline__ void ld_reduce(const void* mc, uint32_t (&w)[4]) {
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                     : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(mc) : "memory");
    }
};
__device__ __forceinline__ void multimem_st_16B(void* mc, const uint32_t (&w)[4]) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                 "r"(w[3]) : "memory");
}

// kRoute = 0: peer-to-peer loads / stores on every rank's mapped buffer.
// kRoute = 1: the 8-element vector path goes through the multicast address (in-switch reduction and broadcast);
//             the scalar head / tail elements still use the per-peer pointers.
// kRoute = 2: hybrid -- both paths are compiled in and every 32 KB block of the flat buffers picks one, independently
//             for the gradient loads and the parameter stores: block i (of every 16 consecutive ones) goes peer-to-peer
//             when i < ld_p2p_of16 / st_p2p_of16, through the switch otherwise.  If the switch's reduction / multicast
//             path and the plain peer-to-peer path have separate ceilings, a mix moves more bytes per second than
//             either alone (tools/sweep_p2p.py measures the split).  Both paths deliver the same bits whenever the sum
//             is exact in Tg (and the same rounding: fp32 accumulate, one rounding to Tg).
constexpr int kRouteP2P = 0, kRouteMc = 1, kRouteHybrid = 2;
template <class Tg, int WORLD, int kRoute> struct DistAdamP2POp : DistAdamOp<Tg, true, true> {
    using Base = DistAdamOp<Tg, true, true>;
    static constexpr unsigned kCustomLoad = 0b01000, kCustomStore = 0b10000;
    struct Params {
        typename Base::Params base;
        const char* gbase[WORLD];  // every rank's flat gradient buffer, as mapped in THIS process
        char* pbase[WORLD];        // every rank's flat parameter buffer
        const char* gown;          // = gbase[rank], pbase[rank]: passed separately so that the arrays are only ever
        char* pown;                //   indexed by compile-time constants (a runtime index would spill them to local memory)
        const char* gmc;           // multicast addresses of the two buffers (kRoute != 0)
        char* pmc;
        int ld_p2p_of16, st_p2p_of16;  // hybrid split (kRoute == 2)
    };
    Params q;
    __device__ explicit DistAdamP2POp(const Params& prm) : Base(prm.base), q(prm) {}
    __device__ __forceinline__ static bool block_is_p2p(ptrdiff_t off, int of16) {
        return (int)(((uint64_t)off >> 15) & 15u) < of16;
    }
    template <int G>
    __device__ __forceinline__ void load_p2p_vec(ptrdiff_t off, float (&x)[G]) const {
        uint32_t w[WORLD][4];
#pragma unroll
        for (int r = 0; r < WORLD; r++) RawVec<4, true>::ld(q.gbase[r] + off, w[r]);  // all peers in flight
#pragma unroll
        for (int i = 0; i < 4; i++) {
            float lo, hi;
            Packed16<Tg>::unpack(w[0][i], lo, hi);
#pragma unroll
            for (int r = 1; r < WORLD; r++) {
                float l2, h2;
                Packed16<Tg>::unpack(w[r][i], l2, h2);
                lo = fadd(lo, l2);
                hi = fadd(hi, h2);
            }
            x[2 * i] = Elem<Tg>::round(lo);
            x[2 * i + 1] = Elem<Tg>::round(hi);
        }
    }
    template <int G>
    __device__ __forceinline__ void load_mc_vec(ptrdiff_t off, float (&x)[G]) const {
        uint32_t w[4];
        Multimem<Tg>::ld_reduce(q.gmc + off, w);  // sum over all ranks, fp32 accumulate, rounded once to Tg
#pragma unroll
        for (int i = 0; i < 4; i++) Packed16<Tg>::unpack(w[i], x[2 * i], x[2 * i + 1]);
    }
    template <int D, int G, bool kVec>
    __device__ __forceinline__ void custom_load(const Tg* p, float (&x)[G]) const {
        static_assert(D == 3, "only the gradient list is gathered");
        const ptrdiff_t off = reinterpret_cast<const char*>(p) - q.gown;
        if constexpr (kVec) {
            static_assert(G == 8, "vector path moves 8 x 16-bit = 16 bytes");
            if constexpr (kRoute == kRouteMc) load_mc_vec<G>(off, x);
            else if constexpr (kRoute == kRouteP2P) load_p2p_vec<G>(off, x);
            else {
                if (block_is_p2p(off, q.ld_p2p_of16)) load_p2p_vec<G>(off, x);
                else load_mc_vec<G>(off, x);
            }
        } else {
            float acc = Elem<Tg>::load(reinterpret_cast<const Tg*>(q.gbase[0] + off));
#pragma unroll
            for (int r = 1; r < WORLD; r++) acc = fadd(acc, Elem<Tg>::load(reinterpret_cast<const Tg*>(q.gbase[r] + off)));
            x[0] = Elem<Tg>::round(acc);
        }
    }
    template <int G>
    __device__ __forceinline__ void store_mc_vec(ptrdiff_t off, const float (&x)[G]) const {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            if constexpr (std::is_same<Tg, __nv_bfloat16>::value) {
                __nv_bfloat162 h = __floats2bfloat162_rn(x[2 * i], x[2 * i + 1]);
                w[i] = *reinterpret_cast<uint32_t*>(&h);
            } else {
                __half2 h = __floats2half2_rn(x[2 * i], x[2 * i + 1]);
                w[i] = *reinterpret_cast<uint32_t*>(&h);
            }
        }
        multimem_st_16B(q.pmc + off, w);
    }
    template <int D, int G, bool kVec>
    __device__ __forceinline__ void custom_store(Tg* p, const float (&x)[G]) const {
        static_assert(D == 4, "only the parameter copy is scattered");
        const ptrdiff_t off = reinterpret_cast<const char*>(p) - q.pown;
        if constexpr (kVec && kRoute != kRouteP2P) {
            bool mc = true;
            if constexpr (kRoute == kRouteHybrid) mc = !block_is_p2p(off, q.st_p2p_of16);
            if (mc) {
                store_mc_vec<G>(off, x);
                return;
            }
        }
#pragma unroll
        for (int r = 0; r < WORLD; r++) {
            Tg* dst = reinterpret_cast<Tg*>(q.pbase[r] + off);
            if constexpr (kVec) Group<Tg, G, false>::st(dst, x);
            else Elem<Tg>::store(dst, x[0]);
        }
    }
};

// ------------------------------------------------------------------------------------------------
// SGDFunctor: csrc/multi_tensor_sgd_kernel.cu:87-134.   lists: g, w, mom [, fp16 model copy]
template <class Tg, class Tw, bool kCopy> struct SgdOp : OpBase {
    static constexpr int kLists = kCopy ? 4 : 3;
    template <int D> using T = typename Types<Tg, Tw, Tw, __half>::template at<D>;
    static constexpr unsigned kRead = 0b0111, kWrite = kCopy ? 0b1110 : 0b0110;
    static constexpr bool kEarlyExit = true;  // :46
    struct Params { float wd, momentum, dampening, lr, scale; int nesterov, first_run, wd_after_momentum; };
    Params p;
    float omd, nlr;
    __device__ explicit SgdOp(const Params& q) : p(q), omd(fsub(1.f, q.dampening)), nlr(-q.lr) {}
    // The momentum buffer is only consumed when momentum != 0 and this is not the first run, and only
    // written when momentum != 0 (:113-132): skip the traffic otherwise.
    __device__ __forceinline__ bool load(int d) const { return d != 2 || (p.momentum != 0.f && !p.first_run); }
    __device__ __forceinline__ bool store(int d) const { return d != 2 || p.momentum != 0.f; }
    __device__ __forceinline__ void elem(float (&x)[kLists], State&) const {
        float g = fmul(x[0], p.scale);
        const float w = x[1];
        float mom = x[2];
        if (p.wd != 0.f && !p.wd_after_momentum) g = fadd(g, fmul(p.wd, w));
        if (p.momentum != 0.f) {
            if (!p.first_run) mom = fadd(fmul(mom, p.momentum), fmul(omd, g));
            else mom = g;
            if (p.nesterov) g = fadd(g, fmul(p.momentum, mom));
            else g = mom;
        }
        if (p.wd != 0.f && p.wd_after_momentum) g = fadd(g, fmul(p.wd, w));
        // weight_in[i] += (-lr * g): for a 16-bit T_weight the increment is rounded to T first (the only
        // viable overload is operator+=(T&, const T&), torch/headeronly/util/Half.h:529), then T + T.
        const float nw = fadd(w, Elem<Tw>::round(fmul(nlr, g)));
        x[1] = nw;
        x[2] = mom;
        if constexpr (kCopy) x[3] = Elem<Tw>::round(nw);  // cast from the STORED weight (:130)
    }
};

// ------------------------------------------------------------------------------------------------
// Norm accumulation state: U independent accumulators are not needed for correctness; one per thread.
struct SumState { float sum; };

// L2NormFunctor: csrc/multi_tensor_l2norm_kernel.cu:56-107.  One tile == one reference chunk, so the flag
// ("chunk partial sum is non-finite", :103-104) has the reference's granularity.   lists: x
// kMp = true is the reference's "_mp" variant (csrc/multi_tensor_l2norm_kernel_mp.cu:40-42): do nothing when the flag
// is already set.
template <class Tx, bool kMp = false> struct L2NormOp : OpBase {
    static constexpr int kLists = 1;
    static constexpr bool kEarlyExit = kMp;
    static constexpr bool kTickets = true;
    template <int D> using T = Tx;
    static constexpr unsigned kRead = 0b1, kWrite = 0b0;
    struct Params { NormOut out; };
    using State = SumState;
    Params p;
    __device__ explicit L2NormOp(const Params& q) : p(q) {}
    __device__ __forceinline__ void init(State& s) const { s.sum = 0.f; }
    __device__ __forceinline__ void elem(float (&x)[1], State& s) const { s.sum = fadd(s.sum, fmul(x[0], x[0])); }
    // vector path: groups stay packed while in flight and are unpacked here, in element order (same sum order as elem())
    static constexpr bool kRawReduce = sizeof(Tx) <= 4;
    template <int W> __device__ __forceinline__ void consume_raw(const uint32_t (&w)[W], State& s) const {
#pragma unroll
        for (int i = 0; i < W; i++) {
            if constexpr (sizeof(Tx) == 4) {
                const float x = __uint_as_float(w[i]);
                s.sum = fadd(s.sum, fmul(x, x));
            } else {
                float lo, hi;
                Packed16<Tx>::unpack(w[i], lo, hi);
                s.sum = fadd(s.sum, fmul(lo, lo));
                s.sum = fadd(s.sum, fmul(hi, hi));
            }
        }
    }
    template <int THREADS>
    __device__ __forceinline__ void end_tile_n(const DevTable& tb, const Tile& t, uint32_t tile, State& s, float* smem,
                                               int32_t* noop, bool skipped) const {
        if (!skipped) {
            const float tot = block_sum<THREADS>(s.sum, smem);
            if (threadIdx.x == 0) {
                p.out.partials[tile] = tot;
                if (!finite_f(tot)) *noop = 1;
            }
        }
        s.sum = 0.f;
        norm_tile_done<THREADS, 1, kMp>(tb, t, tile, p.out, smem, noop);
    }
};

// L2NormScaleFunctor: csrc/multi_tensor_l2norm_scale_kernel.cu:79-81,127-131.   lists: in, out
template <class Tin, class Tout> struct L2NormScaleOp : OpBase {
    static constexpr int kLists = 2;
    static constexpr bool kTickets = true;
    template <int D> using T = typename Types<Tin, Tout>::template at<D>;
    static constexpr unsigned kRead = 0b01, kWrite = 0b10;
    struct Params { NormOut out; float scale; };
    using State = SumState;
    Params p;
    __device__ explicit L2NormScaleOp(const Params& q) : p(q) {}
    __device__ __forceinline__ void init(State& s) const { s.sum = 0.f; }
    __device__ __forceinline__ void elem(float (&x)[2], State& s) const {
        s.sum = fadd(s.sum, fmul(x[0], x[0]));
        x[1] = fmul(x[0], p.scale);
    }
    template <int THREADS>
    __device__ __forceinline__ void end_tile_n(const DevTable& tb, const Tile& t, uint32_t tile, State& s, float* smem,
                                               int32_t* noop, bool) const {
        const float tot = block_sum<THREADS>(s.sum, smem);
        if (threadIdx.x == 0) {
            p.out.partials[tile] = tot;
            if (!finite_f(tot)) *noop = 1;
        }
        s.sum = 0.f;
        norm_tile_done<THREADS, 1, false>(tb, t, tile, p.out, smem, noop);
    }
};

// ------------------------------------------------------------------------------------------------
// LAMB stage 1: csrc/multi_tensor_lamb.cu:66,121-143, fused with BOTH per-tensor norm passes of
// multi_tensor_lamb_cuda (:370 ||p|| before the update, :394 ||u|| of the stored update): the per-tile
// partial sums of p^2 and (T-rounded u)^2 are written to scratch and reduced per tensor by
// lamb_norm_finalize_kernel.   lists: g (overwritten with u), p, m, v
struct Sum2State { float sp, su; };
template <class Tp, int kMode> struct LambStage1Op : OpBase {
    static constexpr int kLists = 4;
    template <int D> using T = Tp;
    static constexpr unsigned kRead = 0b1111, kWrite = 0b1101;
    struct Params {
        float beta1, beta2, beta3, bc1, bc2, eps, decay, max_grad_norm;
        const float* global_grad_norm;
        NormOut out;          // partials [2][ntiles]: p^2 then u^2; ret_pt [2][ntensors]: ||p|| then ||u||
        int mode, need_norms;
    };
    using State = Sum2State;
    static constexpr bool kTickets = true;
    Params p;
    float clip, omb2;
    __device__ explicit LambStage1Op(const Params& q) : p(q), omb2(fsub(1.f, q.beta2)) {
        const float gn = __ldg(q.global_grad_norm);
        clip = gn > q.max_grad_norm ? fdiv(gn, q.max_grad_norm) : 1.0f;  // :66
    }
    __device__ __forceinline__ void init(State& s) const { s.sp = 0.f; s.su = 0.f; }
    // p is only needed for the decay term (:102-103) or for its norm
    __device__ __forceinline__ bool load(int d) const { return d != 1 || p.decay != 0.f || p.need_norms; }
    __device__ __forceinline__ void elem(float (&x)[4], State& s) const {
        const float w_in = x[1];
        s.sp = fadd(s.sp, fmul(w_in, w_in));
        const float w = p.decay == 0.f ? 0.f : w_in;  // r_p = 0 when decay == 0 (:123-128)
        float m = x[2], v = x[3];
        float sg = fdiv(x[0], clip);
        if constexpr (kMode == 0) sg = fadd(sg, fmul(p.decay, w));
        m = fadd(fmul(m, p.beta1), fmul(p.beta3, sg));
        v = fadd(fmul(v, p.beta2), fmul(fmul(omb2, sg), sg));
        const float denom = fadd(fsqrt(fdiv(v, p.bc2)), p.eps);
        float u = fdiv(fdiv(m, p.bc1), denom);
        if constexpr (kMode != 0) u = fadd(u, fmul(p.decay, w));
        const float ur = Elem<Tp>::round(u);  // the reference norms the STORED update
        s.su = fadd(s.su, fmul(ur, ur));
        x[0] = u; x[2] = m; x[3] = v;
    }
    template <int THREADS>
    __device__ __forceinline__ void end_tile_n(const DevTable& tb, const Tile& t, uint32_t tile, State& s, float* smem,
                                               int32_t* noop, bool) const {
        if (p.need_norms) {  // uniform
            const float tp = block_sum<THREADS>(s.sp, smem);
            const float tu = block_sum<THREADS>(s.su, smem);
            if (threadIdx.x == 0) { p.out.partials[tile] = tp; p.out.partials[tb.ntiles + tile] = tu; }
            s.sp = 0.f; s.su = 0.f;
            norm_tile_done<THREADS, 2, false>(tb, t, tile, p.out, smem, noop);
        }
    }
};

// LAMB stage 2: csrc/multi_tensor_lamb.cu:255-263,289.   lists: u, p
template <class Tp> struct LambStage2Op : OpBase {
    static constexpr int kLists = 2;
    template <int D> using T = Tp;
    static constexpr unsigned kRead = 0b11, kWrite = 0b10;
    struct Params { const float *param_norm, *update_norm; float lr, decay; int use_nvlamb; };
    Params p;
    float ratio;
    __device__ explicit LambStage2Op(const Params& q) : p(q), ratio(q.lr) {}
    __device__ __forceinline__ void begin_tensor(int t) {
        ratio = p.lr;
        if (p.use_nvlamb || p.decay != 0.0f) {
            const float pn = p.param_norm[t], un = p.update_norm[t];  // plain loads: written by the previous kernel
            ratio = (un != 0.0f && pn != 0.0f) ? fmul(p.lr, fdiv(pn, un)) : p.lr;
        }
    }
    __device__ __forceinline__ void elem(float (&x)[2], State&) const { x[1] = fsub(x[1], fmul(ratio, x[0])); }
};

// ------------------------------------------------------------------------------------------------
// Mixed-precision / capturable LAMB ("_mp"): csrc/multi_tensor_lamb_mp.cu:40-232 (stage 1), :234-364 (stage 2).
// Differences from LambStage1Op: g has its own dtype Tg and is multiplied by *inv_scale first; p, m, v have dtype Tp
// (fp32 masters or Tg); step, max_grad_norm and the learning rate are read from device memory (so a captured graph
// replays across steps); the kernels return at once when the flag is set.
template <class Tg, class Tp, int kMode> struct LambMpStage1Op : OpBase {
    static constexpr int kLists = 4;
    template <int D> using T = typename Types<Tg, Tp, Tp, Tp>::template at<D>;
    static constexpr unsigned kRead = 0b1111, kWrite = 0b1101;
    static constexpr bool kEarlyExit = true;
    struct Params {
        float beta1, beta2, beta3, eps, decay;
        const int32_t* step;
        const float *global_grad_norm, *max_grad_norm, *inv_scale;
        NormOut out;  // partials [2][ntiles], ret_pt [2][ntensors]
        int mode, bias_correction, need_norms;
    };
    using State = Sum2State;
    static constexpr bool kTickets = true;
    Params p;
    float clip, omb2, bc1, bc2, inv_scale;
    __device__ explicit LambMpStage1Op(const Params& q) : p(q), omb2(fsub(1.f, q.beta2)), bc1(1.f), bc2(1.f) {
        const float gn = __ldg(q.global_grad_norm), mx = __ldg(q.max_grad_norm);
        clip = gn > mx ? fdiv(gn, mx) : 1.0f;
        inv_scale = __ldg(q.inv_scale);
        if (q.bias_correction == 1) {  // :64-70 (the reference uses a fast-math float pow; double here and in the oracle)
            const int step = __ldg(q.step);
            bc1 = (float)(1.0 - pow((double)q.beta1, (double)step));
            bc2 = (float)(1.0 - pow((double)q.beta2, (double)step));
        }
    }
    __device__ __forceinline__ void init(State& s) const { s.sp = 0.f; s.su = 0.f; }
    __device__ __forceinline__ bool load(int d) const { return d != 1 || p.decay != 0.f || p.need_norms; }
    __device__ __forceinline__ void elem(float (&x)[4], State& s) const {
        const float w_in = x[1];
        s.sp = fadd(s.sp, fmul(w_in, w_in));
        const float w = p.decay == 0.f ? 0.f : w_in;
        float m = x[2], v = x[3];
        float sg = fdiv(fmul(x[0], inv_scale), clip);
        if constexpr (kMode == 0) sg = fadd(sg, fmul(p.decay, w));
        m = fadd(fmul(m, p.beta1), fmul(p.beta3, sg));
        v = fadd(fmul(v, p.beta2), fmul(fmul(omb2, sg), sg));
        const float denom = fadd(fsqrt(fdiv(v, bc2)), p.eps);
        float u = fdiv(fdiv(m, bc1), denom);
        if constexpr (kMode != 0) u = fadd(u, fmul(p.decay, w));
        const float ur = Elem<Tg>::round(u);
        s.su = fadd(s.su, fmul(ur, ur));
        x[0] = u; x[2] = m; x[3] = v;
    }
    template <int THREADS>
    __device__ __forceinline__ void end_tile_n(const DevTable& tb, const Tile& t, uint32_t tile, State& s, float* smem,
                                               int32_t* noop, bool skipped) const {
        if (p.need_norms) {
            if (!skipped) {
                const float tp = block_sum<THREADS>(s.sp, smem);
                const float tu = block_sum<THREADS>(s.su, smem);
                if (threadIdx.x == 0) { p.out.partials[tile] = tp; p.out.partials[tb.ntiles + tile] = tu; }
            }
            s.sp = 0.f; s.su = 0.f;
            norm_tile_done<THREADS, 2, true>(tb, t, tile, p.out, smem, noop);
        }
    }
};

// lists: u (Tg), p (Tp) [, m, v untouched, p_model (Tg)]: with kCopy the op walks the caller's 5-list table and simply
// leaves lists 2 and 3 alone, so no second table is needed.
template <class Tg, class Tp, bool kCopy> struct LambMpStage2Op : OpBase {
    static constexpr int kLists = kCopy ? 5 : 2;
    template <int D> using T = typename Types<Tg, Tp, Tp, Tp, Tg>::template at<D>;
    static constexpr unsigned kRead = 0b00011, kWrite = kCopy ? 0b10010 : 0b00010;
    static constexpr bool kEarlyExit = true;
    struct Params { const float *param_norm, *update_norm, *lr; float decay; int use_nvlamb; };
    Params p;
    float lr, ratio;
    __device__ explicit LambMpStage2Op(const Params& q) : p(q), lr(__ldg(q.lr)), ratio(0.f) {}
    __device__ __forceinline__ void begin_tensor(int t) {
        ratio = lr;
        if (p.use_nvlamb || p.decay != 0.0f) {
            const float pn = p.param_norm[t], un = p.update_norm[t];
            ratio = (un != 0.0f && pn != 0.0f) ? fmul(lr, fdiv(pn, un)) : lr;
        }
    }
    __device__ __forceinline__ void elem(float (&x)[kLists], State&) const {
        const float w = fsub(x[1], fmul(ratio, x[0]));
        x[1] = w;
        if constexpr (kCopy) x[4] = Elem<Tp>::round(w);  // r_out_p = r_p: cast from the stored param_t value (:319-321)
    }
};

// ------------------------------------------------------------------------------------------------
// AdagradFunctor: csrc/multi_tensor_adagrad.cu:63-72.   lists: g, p, h (one dtype, fp32 math)
template <class Tp> struct AdagradOp : OpBase {
    static constexpr int kLists = 3;
    template <int D> using T = Tp;
    static constexpr unsigned kRead = 0b111, kWrite = 0b110;
    struct Params { float eps, lr, decay; int mode; };
    Params p;
    __device__ explicit AdagradOp(const Params& q) : p(q) {}
    __device__ __forceinline__ void elem(float (&x)[3], State&) const {
        float g = x[0], w = x[1], h = x[2];
        if (p.mode == 0) {  // L2
            g = fadd(g, fmul(p.decay, w));
            h = fadd(h, fmul(g, g));
            w = fsub(w, fmul(p.lr, fdiv(g, fadd(fsqrt(h), p.eps))));
        } else {  // AdamW-style
            h = fadd(h, fmul(g, g));
            w = fsub(w, fmul(p.lr, fadd(fdiv(g, fadd(fsqrt(h), p.eps)), fmul(p.decay, w))));
        }
        x[1] = w; x[2] = h;
    }
};

// NovoGradFunctor: csrc/multi_tensor_novograd.cu:90-107.   lists: g, p, m; per-tensor second moment (a norm) on device
template <class Tp> struct NovoGradOp : OpBase {
    static constexpr int kLists = 3;
    template <int D> using T = Tp;
    static constexpr unsigned kRead = 0b111, kWrite = 0b110;
    struct Params { float beta1, beta2, beta3, bc1, bc2, eps, lr, decay; int mode; const float* grad_norm; };
    Params p;
    float gn;
    __device__ explicit NovoGradOp(const Params& q) : p(q), gn(0.f) {}
    __device__ __forceinline__ void begin_tensor(int t) { gn = p.grad_norm[t]; }  // written by the previous kernel
    __device__ __forceinline__ void elem(float (&x)[3], State&) const {
        float g = x[0], w = x[1], m = x[2];
        const float denom = fadd(fdiv(gn, p.bc2), p.eps);
        if (p.mode == 0) {
            g = fadd(fdiv(g, denom), fmul(p.decay, w));
            m = fadd(fmul(p.beta1, m), fmul(p.beta3, g));
            w = fsub(w, fmul(p.lr, fdiv(m, p.bc1)));
        } else {
            m = fadd(fmul(p.beta1, m), fmul(p.beta3, g));
            const float upd = fadd(fdiv(fdiv(m, p.bc1), denom), fmul(p.decay, w));
            w = fsub(w, fmul(p.lr, upd));
        }
        x[1] = w; x[2] = m;
    }
};

// MaxNormFunctor: csrc/multi_tensor_l2norm_kernel.cu:113-195 -- per-chunk max |x| (for NovoGrad's L-inf norm).
struct MaxState { float mx; };
template <int THREADS> __device__ __forceinline__ float block_max(float v, float* smem) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    float r = 0.f;
    if (warp == 0) {
        r = lane < THREADS / 32 ? smem[lane] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) r = fmaxf(r, __shfl_xor_sync(0xffffffffu, r, o));
    }
    return r;
}
template <class Tx> struct MaxNormOp : OpBase {
    static constexpr int kLists = 1;
    template <int D> using T = Tx;
    static constexpr unsigned kRead = 0b1, kWrite = 0b0;
    struct Params { float* partials; };
    using State = MaxState;
    Params p;
    __device__ explicit MaxNormOp(const Params& q) : p(q) {}
    __device__ __forceinline__ void init(State& s) const { s.mx = 0.f; }
    __device__ __forceinline__ void elem(float (&x)[1], State& s) const { s.mx = fmaxf(fabsf(s.mx), fabsf(x[0])); }
    template <int THREADS>
    __device__ __forceinline__ void end_tile_n(const DevTable&, const Tile&, uint32_t tile, State& s, float* smem,
                                               int32_t* noop, bool) const {
        const float tot = block_max<THREADS>(s.mx, smem);
        if (threadIdx.x == 0) {
            p.partials[tile] = tot;
            if (!finite_f(tot)) *noop = 1;  // :187-188
        }
        s.mx = 0.f;
    }
};

// LARSFunctor: csrc/multi_tensor_lars.cu:58-139.   lists: g, w, mom [, fp16 model copy]
template <class Tg, class Tw, bool kCopy> struct LarsOp : OpBase {
    static constexpr int kLists = kCopy ? 4 : 3;
    template <int D> using T = typename Types<Tg, Tw, Tw, __half>::template at<D>;
    static constexpr unsigned kRead = 0b0111, kWrite = kCopy ? 0b1110 : 0b0110;
    static constexpr bool kEarlyExit = true;
    struct Params {
        const float *grad_norms, *param_norms;
        float lr, trust_coefficient, eps, weight_decay, momentum;
        int nesterov, is_skipped;
    };
    Params p;
    float scaled_lr;
    __device__ explicit LarsOp(const Params& q) : p(q), scaled_lr(q.lr) {}
    __device__ __forceinline__ void begin_tensor(int t) {
        scaled_lr = p.lr;
        if (!p.is_skipped) {
            const float pn = p.param_norms[t], gn = p.grad_norms[t];
            float trust = 1.0f;
            if (gn > 0.0f && pn > 0.0f)
                trust = fdiv(fmul(p.trust_coefficient, pn), fadd(fadd(gn, fmul(pn, p.weight_decay)), p.eps));
            scaled_lr = fmul(p.lr, trust);
        }
    }
    __device__ __forceinline__ void elem(float (&x)[kLists], State&) const {
        const float g = fadd(x[0], fmul(p.weight_decay, x[1]));
        float w = x[1];
        const float mom = fsub(fmul(x[2], p.momentum), fmul(scaled_lr, g));
        if (p.nesterov) w = fadd(w, fsub(fmul(mom, p.momentum), fmul(scaled_lr, g)));
        else w = fadd(w, mom);
        x[1] = w; x[2] = mom;
        if constexpr (kCopy) x[3] = Elem<Tw>::round(w);
    }
};

// ------------------------------------------------------------------------------------------------
// Legacy split LAMB, arithmetic "in T": every value assigned to a T variable is rounded to T.
// csrc/multi_tensor_lamb_stage_1.cu:81-100.   lists: g (Tg), p, m, v (Tp), update (Tu)
template <class Tg, class Tp, class Tu> struct LambStage1LegacyOp : OpBase {
    static constexpr int kLists = 5;
    template <int D> using T = typename Types<Tg, Tp, Tp, Tp, Tu>::template at<D>;
    static constexpr unsigned kRead = 0b01111, kWrite = 0b11100;
    struct Params { const float* per_tensor_decay; float beta1, beta2, bc1, bc2, eps, clip; };
    Params p;
    float decay, omb1, omb2;
    __device__ explicit LambStage1LegacyOp(const Params& q) : p(q), decay(0.f), omb1(fsub(1.f, q.beta1)), omb2(fsub(1.f, q.beta2)) {}
    __device__ __forceinline__ void begin_tensor(int t) { decay = __ldg(p.per_tensor_decay + t); }
    __device__ __forceinline__ void elem(float (&x)[5], State&) const {
        using R = Elem<Tp>;
        const float w = x[1];
        float m = x[2], v = x[3];
        const float sg = R::round(fdiv(x[0], p.clip));
        m = R::round(fadd(fmul(m, p.beta1), fmul(omb1, sg)));
        v = R::round(fadd(fmul(v, p.beta2), fmul(fmul(omb2, sg), sg)));
        const float mu = R::round(fdiv(m, p.bc1)), vu = R::round(fdiv(v, p.bc2));
        const float denom = R::round(fadd(fsqrt(vu), p.eps));
        const float q = R::round(fdiv(mu, denom));  // T / T returns T
        x[4] = R::round(fadd(q, fmul(decay, w)));
        x[2] = m; x[3] = v;
    }
};

// csrc/multi_tensor_lamb_stage_2.cu:41-47,77.   lists: p (Tp), update (Tu)
template <class Tp, class Tu> struct LambStage2LegacyOp : OpBase {
    static constexpr int kLists = 2;
    template <int D> using T = typename Types<Tp, Tu>::template at<D>;
    static constexpr unsigned kRead = 0b11, kWrite = 0b01;
    struct Params { const float *param_norm, *update_norm; float lr, decay; int use_nvlamb; };
    Params p;
    float ratio;
    __device__ explicit LambStage2LegacyOp(const Params& q) : p(q), ratio(q.lr) {}
    __device__ __forceinline__ void begin_tensor(int t) {
        ratio = p.lr;
        if (p.use_nvlamb || p.decay != 0.0f) {
            const float pn = p.param_norm[t], un = p.update_norm[t];
            ratio = (un != 0.0f && pn != 0.0f) ? fmul(p.lr, fdiv(pn, un)) : p.lr;
        }
    }
    __device__ __forceinline__ void elem(float (&x)[2], State&) const {
        x[0] = fsub(x[0], fmul(ratio, Elem<Tp>::round(x[1])));
    }
};

// ------------------------------------------------------------------------------------------------
// Sharded LAMB (the kernels behind apex.contrib.optimizers.DistributedFusedLAMB):
// apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb_kernel.cu:119-321 (stage 1), :325-433 (stage 2).
// Stage 1.  lists: g (Tg), p, m, v (fp32 shard state), u (fp32 update term).  Per-FRAGMENT hyper-parameters live in
// device arrays; step, global_scale and the global gradient norm are device scalars; beta3 is 1 - beta1 whatever
// per_tensor_beta3 holds (:140); the bias corrections are evaluated in the kernel (:142-148; double pow here and in
// the oracle, as for lamb_mp).  The clip factor is folded into the un-scaling divisor (:133-137).
// Extension (need_norms): the per-tile sums of round_Tn(p)^2 and u^2 are written to scratch, so neither of the two
// norm passes the reference's Python runs around this kernel (distributed_fused_lamb.py:673-692) needs its own sweep:
// the model parameter IS round_Tn(master) at every step boundary, so its norm can be taken from the master shard.
template <class Tg, class Tn, int kMode> struct DistLambStage1Op : OpBase {
    static constexpr int kLists = 5;
    template <int D> using T = typename Types<Tg, float, float, float, float>::template at<D>;
    static constexpr unsigned kRead = 0b01111, kWrite = 0b11100;
    static constexpr bool kEarlyExit = true;
    struct Params {
        const float *beta1, *beta2, *eps, *decay;
        const int32_t *bias_correction, *step;
        const float *global_scale, *global_grad_norm;
        float max_grad_norm;
        float* partials;  // [2][ntiles]: round_Tn(p)^2 then u^2
        uint32_t ntiles;
        int need_norms;
    };
    using State = Sum2State;
    Params p;
    float cscale, b1, b2, b3, omb2, bc1, bc2, eps, decay;
    int step, bc_on;
    __device__ explicit DistLambStage1Op(const Params& q) : p(q), b1(-1.f), b2(-1.f), bc1(1.f), bc2(1.f), bc_on(-1) {
        const float gs = __ldg(q.global_scale);
        cscale = gs;
        if (q.max_grad_norm > 0.f) {  // :134-137 -- the 1e-6 literal makes the quotient a double expression
            const float c = (float)((double)q.max_grad_norm / ((double)fdiv(__ldg(q.global_grad_norm), gs) + 1e-6));
            cscale = fdiv(gs, fminf(1.0f, c));
        }
        step = __ldg(q.step);
    }
    __device__ __forceinline__ void begin_tensor(int t) {
        const float nb1 = __ldg(p.beta1 + t), nb2 = __ldg(p.beta2 + t);
        const int on = __ldg(p.bias_correction + t) == 1;
        eps = __ldg(p.eps + t); decay = __ldg(p.decay + t);
        if (nb1 != b1 || nb2 != b2 || on != bc_on) {
            b1 = nb1; b2 = nb2; bc_on = on;
            b3 = fsub(1.f, b1); omb2 = fsub(1.f, b2);
            bc1 = 1.f; bc2 = 1.f;
            if (on) {
                bc1 = (float)(1.0 - pow((double)b1, (double)step));
                bc2 = (float)(1.0 - pow((double)b2, (double)step));
            }
        }
    }
    __device__ __forceinline__ void init(State& s) const { s.sp = 0.f; s.su = 0.f; }
    __device__ __forceinline__ bool load(int d) const { return d != 1 || decay != 0.f || p.need_norms; }
    __device__ __forceinline__ void elem(float (&x)[5], State& s) const {
        const float wn = Elem<Tn>::round(x[1]);
        s.sp = fadd(s.sp, fmul(wn, wn));
        const float w = decay == 0.f ? 0.f : x[1];  // :203-208
        float m = x[2], v = x[3];
        float sg = fdiv(x[0], cscale);
        if constexpr (kMode == 0) sg = fadd(sg, fmul(decay, w));
        m = fadd(fmul(m, b1), fmul(b3, sg));
        v = fadd(fmul(v, b2), fmul(fmul(omb2, sg), sg));
        const float denom = fadd(fsqrt(fdiv(v, bc2)), eps);
        float u = fdiv(fdiv(m, bc1), denom);
        if constexpr (kMode != 0) u = fadd(u, fmul(decay, w));
        s.su = fadd(s.su, fmul(u, u));
        x[4] = u; x[2] = m; x[3] = v;
    }
    template <int THREADS>
    __device__ __forceinline__ void end_tile_n(const DevTable&, const Tile&, uint32_t tile, State& s, float* smem,
                                               int32_t* noop, bool) const {
        if (p.need_norms) {
            const float tp = block_sum<THREADS>(s.sp, smem);
            const float tu = block_sum<THREADS>(s.su, smem);
            if (threadIdx.x == 0) {
                p.partials[tile] = tp; p.partials[p.ntiles + tile] = tu;
                // the reference's two multi_tensor_l2norm calls raise the flag on a non-finite chunk sum, which
                // makes stage 2 a no-op on this rank
                if (!finite_f(tp) || !finite_f(tu)) *noop = 1;
            }
        }
        s.sp = 0.f; s.su = 0.f;
    }
};

// Stage 2.  lists: u (fp32), p (fp32), p_copy (Tc).  update_norm is indexed through update_norm_offset (:349-350).
template <class Tc> struct DistLambStage2Op : OpBase {
    static constexpr int kLists = 3;
    template <int D> using T = typename Types<float, float, Tc>::template at<D>;
    static constexpr unsigned kRead = 0b011, kWrite = 0b110;
    static constexpr bool kEarlyExit = true;
    struct Params {
        const float *param_norm, *update_norm;
        const int64_t* update_norm_offset;
        const float *lr, *decay;
        int use_nvlamb;
    };
    Params p;
    float lr, ratio;
    __device__ explicit DistLambStage2Op(const Params& q) : p(q), lr(__ldg(q.lr)), ratio(0.f) {}
    __device__ __forceinline__ void begin_tensor(int t) {
        ratio = lr;
        if (p.use_nvlamb || __ldg(p.decay + t) != 0.0f) {
            const float pn = p.param_norm[t], un = p.update_norm[p.update_norm_offset[t]];
            ratio = (un != 0.0f && pn != 0.0f) ? fmul(lr, fdiv(pn, un)) : lr;
        }
    }
    __device__ __forceinline__ void elem(float (&x)[3], State&) const {
        const float w = fsub(x[1], fmul(ratio, x[0]));
        x[1] = w; x[2] = w;
    }
};

}  // namespace mt
