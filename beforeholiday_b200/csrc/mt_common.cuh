// mt_common.cuh -- device-side machinery shared by every multi-tensor kernel.
//
// Replaces the reference's chunk launcher (/root/reference/csrc/multi_tensor_apply.cuh:16-147): instead of
// packing <=36..110 tensors / <=2560 chunks into a ~14 KB by-value kernel argument and launching repeatedly,
// ONE launch walks a DEVICE-RESIDENT tensor table:
//
//   DevTable { ptrs[depth][ntensors], numel[ntensors], tile_begin[ntensors+1] }   (built + cached by mt_table.cu)
//
// Work unit = "tile": `tile_elems` consecutive elements of one tensor.  The grid is table-driven: CTA b takes tile b
// (and b+grid, ... when the grid is capped), so the hardware block scheduler is the dynamic tile scheduler.
// Measured on B200 (profiles/r1_tune_sweep*.log, 28 B/param Adam, 45 GB/step): one 256-thread CTA per 16384-element
// tile at 4 CTAs/SM streams 6.68 TB/s (= cudaMemcpy D2D on the same box); a resident grid of SMs x occupancy CTAs
// striding the tiles statically reached only 5.87 TB/s (slow SMs set the finish time), and a cp.async.bulk (TMA)
// staged ring 6.2 TB/s -- so the streaming kernels use plain 256-bit LDG/STG and many short-lived CTAs.
// The owning tensor is found with a binary search in tile_begin (uniform across the CTA -> broadcast loads).
// Inside a tile each thread moves G consecutive elements per access: 32 B (LDG.256/STG.256) for fp32 and 16 B
// for 16-bit types at G = 8, fully coalesced; a per-tile scalar head/tail peel handles arbitrary element
// offsets (shard fragments, odd sizes) without dropping the whole tile to scalar code the way the reference's
// is_aligned() test does (csrc/multi_tensor_scale_kernel.cu:59).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace mt {

struct DevTable {
    const void* const* ptrs;     // [depth][ntensors], list-major
    const int64_t* numel;        // [ntensors]
    const uint32_t* tile_begin;  // [ntensors + 1] exclusive prefix sum of per-tensor tile counts
    int32_t ntensors;
    uint32_t ntiles;
    int32_t tile_elems;
    // [ntensors + 1] completion counters of the reducing ops (per tensor; [ntensors] = whole launch).  Zero between
    // launches: the CTA that takes the last ticket finalizes and puts the counter back to zero (norm_tile_done below).
    uint32_t* tickets;
    // Small tables (<= kInlineTensors tensors, <= kInlineLists lists) travel in the kernel arguments as well: the lookup
    // at the head of every CTA then reads the parameter bank instead of chasing two or three dependent global loads
    // (cold after anything that swept L2), which is most of what a launch over a few small tensors costs beyond its
    // bytes (C1: 6 tensors, 27.6 MB).  `inl` = 0: the arrays below are unset, use the device blob.
    static constexpr int kInlineTensors = 8, kInlineLists = 5;
    int32_t inl;
    uint32_t inl_tile_begin[kInlineTensors + 1];
    int64_t inl_numel[kInlineTensors];
    const void* inl_ptrs[kInlineLists][kInlineTensors];
};

// ------------------------------------------------------------------------------------------------
// IEEE-754 binary32 arithmetic that the compiler may not contract into FMA.  The CPU oracle
// (oracle/mt_oracle.c) performs the same operations in the same order, so fp32 results are bit-identical.
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float fsqrt(float a) { return __fsqrt_rn(a); }
__device__ __forceinline__ bool finite_f(float a) { return (__float_as_uint(a) & 0x7f800000u) != 0x7f800000u; }

// ------------------------------------------------------------------------------------------------
// Element types.
template <class T> struct Elem;
template <> struct Elem<float> {
    static constexpr int kBytes = 4;
    __device__ static __forceinline__ float load(const float* p) { return *p; }
    __device__ static __forceinline__ void store(float* p, float v) { *p = v; }
    __device__ static __forceinline__ float round(float v) { return v; }
};
template <> struct Elem<double> {
    static constexpr int kBytes = 8;
    __device__ static __forceinline__ float load(const double* p) { return (float)*p; }
    __device__ static __forceinline__ void store(double* p, float v) { *p = (double)v; }
    __device__ static __forceinline__ float round(float v) { return v; }
};
template <> struct Elem<__half> {
    static constexpr int kBytes = 2;
    __device__ static __forceinline__ float load(const __half* p) { return __half2float(*p); }
    __device__ static __forceinline__ void store(__half* p, float v) { *p = __float2half_rn(v); }
    __device__ static __forceinline__ float round(float v) { return __half2float(__float2half_rn(v)); }
};
template <> struct Elem<__nv_bfloat16> {
    static constexpr int kBytes = 2;
    __device__ static __forceinline__ float load(const __nv_bfloat16* p) { return __bfloat162float(*p); }
    __device__ static __forceinline__ void store(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
    __device__ static __forceinline__ float round(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
};

// One-byte e5m2 packing of the reference's `e5m2_allgather` option
// (apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb_kernel.cu:34-66): the byte is the high byte of
// half(v + 2^floor(log2|v|) / 8) -- half an e5m2 ulp is added in magnitude before the fp16 rounding, then the low
// mantissa byte is dropped -- and unpacks as the high byte of an fp16 whose low byte is zero.
struct e5m2_t { uint8_t b; };
__device__ __forceinline__ uint8_t float_to_e5m2(float v) {
    const float pow2 = __uint_as_float(__float_as_uint(v) & 0xFF800000u);
    const __half h = __float2half_rn(fadd(v, fmul(pow2, 0.125f)));
    return (uint8_t)(__half_as_ushort(h) >> 8);
}
__device__ __forceinline__ float e5m2_to_float(uint8_t b) { return __half2float(__ushort_as_half((unsigned short)(b << 8))); }
template <> struct Elem<e5m2_t> {
    static constexpr int kBytes = 1;
    __device__ static __forceinline__ float load(const e5m2_t* p) { return e5m2_to_float(p->b); }
    __device__ static __forceinline__ void store(e5m2_t* p, float v) { p->b = float_to_e5m2(v); }
    __device__ static __forceinline__ float round(float v) { return e5m2_to_float(float_to_e5m2(v)); }
};

// ------------------------------------------------------------------------------------------------
// Raw vector accesses (N 32-bit words).  Streaming data is touched exactly once per step, so nothing is
// allocated in L1 (L1::no_allocate).  The non-coherent (.nc) path is deliberately NOT used even for input-only
// lists: in-place calls, where an output list aliases an input list, are legal for this API
// (/root/reference/tests/L0/run_amp/test_multi_tensor_scale.py:58-77).
// Cache-policy qualifiers of the streaming accesses.  Swept on B200 (profiles/r1_cache_hint_sweep.log); overridable at
// compile time for tools/tune_adam.cu.
#ifndef MT_LD_HINT
#define MT_LD_HINT ".L1::no_allocate"
#endif
#ifndef MT_ST_HINT
#define MT_ST_HINT ".L1::no_allocate"
#endif
#ifndef MT_LD_HINT256   // 256-bit accesses additionally accept .L2::evict_* (ptxas rejects it on narrower ones)
#define MT_LD_HINT256 MT_LD_HINT
#endif
#ifndef MT_ST_HINT256
#define MT_ST_HINT256 MT_ST_HINT
#endif
template <int N, bool RO> struct RawVec;

template <bool RO> struct RawVec<2, RO> {
    __device__ static __forceinline__ void ld(const void* p, uint32_t (&w)[2]) {
        asm volatile("ld.global" MT_LD_HINT ".v2.b32 {%0,%1}, [%2];" : "=r"(w[0]), "=r"(w[1]) : "l"(p) : "memory");
    }
    __device__ static __forceinline__ void st(void* p, const uint32_t (&w)[2]) {
        asm volatile("st.global" MT_ST_HINT ".v2.b32 [%0], {%1,%2};" ::"l"(p), "r"(w[0]), "r"(w[1]) : "memory");
    }
};
template <bool RO> struct RawVec<4, RO> {
    __device__ static __forceinline__ void ld(const void* p, uint32_t (&w)[4]) {
        asm volatile("ld.global" MT_LD_HINT ".v4.b32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p) : "memory");
    }
    __device__ static __forceinline__ void st(void* p, const uint32_t (&w)[4]) {
        asm volatile("st.global" MT_ST_HINT ".v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(w[0]), "r"(w[1]),
                     "r"(w[2]), "r"(w[3]) : "memory");
    }
};
// 256-bit global accesses are an sm_100 addition (SASS: LDG.E.256 / STG.E.256).
template <bool RO> struct RawVec<8, RO> {
    __device__ static __forceinline__ void ld(const void* p, uint32_t (&w)[8]) {
        asm volatile("ld.global" MT_LD_HINT256 ".v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                         : "l"(p) : "memory");
    }
    __device__ static __forceinline__ void st(void* p, const uint32_t (&w)[8]) {
        asm volatile("st.global" MT_ST_HINT256 ".v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(w[0]),
                     "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]) : "memory");
    }
};

// Typed group access: G consecutive elements <-> float[G].
template <class T, int G, bool RO> struct Group;

template <int G, bool RO> struct Group<float, G, RO> {
    __device__ static __forceinline__ void ld(const float* p, float (&f)[G]) {
        uint32_t w[G];
        RawVec<G, RO>::ld(p, w);
#pragma unroll
        for (int i = 0; i < G; i++) f[i] = __uint_as_float(w[i]);
    }
    __device__ static __forceinline__ void st(float* p, const float (&f)[G]) {
        uint32_t w[G];
#pragma unroll
        for (int i = 0; i < G; i++) w[i] = __float_as_uint(f[i]);
        RawVec<G, RO>::st(p, w);
    }
};
template <int G, bool RO> struct Group<__half, G, RO> {
    __device__ static __forceinline__ void ld(const __half* p, float (&f)[G]) {
        uint32_t w[G / 2];
        RawVec<G / 2, RO>::ld(p, w);
#pragma unroll
        for (int i = 0; i < G / 2; i++) {
            float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
            f[2 * i] = t.x;
            f[2 * i + 1] = t.y;
        }
    }
    __device__ static __forceinline__ void st(__half* p, const float (&f)[G]) {
        uint32_t w[G / 2];
#pragma unroll
        for (int i = 0; i < G / 2; i++) {
            __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
            w[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        RawVec<G / 2, RO>::st(p, w);
    }
};
template <int G, bool RO> struct Group<__nv_bfloat16, G, RO> {
    __device__ static __forceinline__ void ld(const __nv_bfloat16* p, float (&f)[G]) {
        uint32_t w[G / 2];
        RawVec<G / 2, RO>::ld(p, w);
#pragma unroll
        for (int i = 0; i < G / 2; i++) {  // bf16 -> f32 is a 16-bit shift
            f[2 * i] = __uint_as_float(w[i] << 16);
            f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    __device__ static __forceinline__ void st(__nv_bfloat16* p, const float (&f)[G]) {
        uint32_t w[G / 2];
#pragma unroll
        for (int i = 0; i < G / 2; i++) {
            __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
            w[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        RawVec<G / 2, RO>::st(p, w);
    }
};
template <int G, bool RO> struct Group<e5m2_t, G, RO> {  // 8 elements = 8 bytes
    static_assert(G == 8, "e5m2 lists move 8 elements per access");
    __device__ static __forceinline__ void ld(const e5m2_t* p, float (&f)[G]) {
        uint32_t w[2];
        RawVec<2, RO>::ld(p, w);
#pragma unroll
        for (int i = 0; i < 8; i++) f[i] = e5m2_to_float((uint8_t)(w[i >> 2] >> (8 * (i & 3))));
    }
    __device__ static __forceinline__ void st(e5m2_t* p, const float (&f)[G]) {
        uint32_t w[2] = {0u, 0u};
#pragma unroll
        for (int i = 0; i < 8; i++) w[i >> 2] |= (uint32_t)float_to_e5m2(f[i]) << (8 * (i & 3));
        RawVec<2, RO>::st(p, w);
    }
};
template <int G, bool RO> struct Group<double, G, RO> {  // rare (Adam dispatches on double); 32-B pieces
    __device__ static __forceinline__ void ld(const double* p, float (&f)[G]) {
#pragma unroll
        for (int k = 0; k < G / 4; k++) {
            uint32_t w[8];
            RawVec<8, RO>::ld(p + 4 * k, w);
#pragma unroll
            for (int i = 0; i < 4; i++)
                f[4 * k + i] = (float)__hiloint2double((int)w[2 * i + 1], (int)w[2 * i]);
        }
    }
    __device__ static __forceinline__ void st(double* p, const float (&f)[G]) {
#pragma unroll
        for (int k = 0; k < G / 4; k++) {
            uint32_t w[8];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                double d = (double)f[4 * k + i];
                w[2 * i] = (uint32_t)__double2loint(d);
                w[2 * i + 1] = (uint32_t)__double2hiint(d);
            }
            RawVec<8, RO>::st(p + 4 * k, w);
        }
    }
};

// ------------------------------------------------------------------------------------------------
// Tile lookup.
struct Tile {
    int32_t tensor;
    int32_t len;    // elements in this tile (<= tile_elems)
    int64_t start;  // element offset inside the tensor
};

template <bool kInline = true>
__device__ __forceinline__ Tile locate_tile(const DevTable& tb, uint32_t tile) {
    // upper_bound(tile_begin, tile) - 1 : last tensor whose first tile is <= tile.  Empty tensors have
    // tile_begin[t] == tile_begin[t+1] and are skipped naturally.
    // 33-ary search: the 32 lanes of a warp each probe one point of [lo, hi) and vote, so 291 tensors take 2 dependent
    // loads and 50 000 take 4 (a bisection: 9 and 16).  That latency sits at the head of every CTA; it is hidden while
    // other CTAs of the SM stream, but not at the start of a launch, which is what a 40-microsecond norm pass feels.
    // Called with the whole CTA converged (the tile index is uniform), every warp searching redundantly.
    if (kInline && tb.inl) {  // uniform: the table came with the kernel arguments
        int lo = 0;
#pragma unroll
        for (int k = 1; k < DevTable::kInlineTensors; k++)
            if (k < tb.ntensors && tb.inl_tile_begin[k] <= tile) lo = k;
        Tile t;
        t.tensor = lo;
        t.start = (int64_t)(tile - tb.inl_tile_begin[lo]) * tb.tile_elems;
        const int64_t rem = tb.inl_numel[lo] - t.start;
        t.len = rem < tb.tile_elems ? (int32_t)rem : tb.tile_elems;
        return t;
    }
    int lo = 0, hi = tb.ntensors;  // invariant: tile_begin[lo] <= tile < tile_begin[hi]
    const int lane = threadIdx.x & 31;
    while (hi - lo > 1) {
        const int span = hi - lo;
        const int probe = lo + (int)(((long long)span * (lane + 1)) / 33);
        const unsigned le = __ballot_sync(0xffffffffu, __ldg(tb.tile_begin + probe) <= tile);
        const int k = __popc(le);  // probes are non-decreasing and tile_begin is sorted: the true votes form a prefix
        const int nlo = k > 0 ? lo + (int)(((long long)span * k) / 33) : lo;
        const int nhi = k < 32 ? lo + (int)(((long long)span * (k + 1)) / 33) : hi;
        lo = nlo;
        hi = nhi > lo ? nhi : lo + 1;
    }
    Tile t;
    t.tensor = lo;
    uint32_t k = tile - __ldg(tb.tile_begin + lo);
    t.start = (int64_t)k * tb.tile_elems;
    int64_t rem = __ldg(tb.numel + lo) - t.start;
    t.len = rem < tb.tile_elems ? (int32_t)rem : tb.tile_elems;
    return t;
}

// ------------------------------------------------------------------------------------------------
// Block reductions (deterministic: fixed shuffle tree, fixed warp order).
template <int THREADS> __device__ __forceinline__ float block_sum(float v, float* smem /* >= THREADS/32 floats */) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fadd(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();  // protect smem reuse across calls
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    float r = 0.f;
    if (warp == 0) {
        r = lane < THREADS / 32 ? smem[lane] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) r = fadd(r, __shfl_xor_sync(0xffffffffu, r, o));
    }
    return r;  // valid in warp 0 (all lanes)
}

// ------------------------------------------------------------------------------------------------
// In-launch finalize of the reducing ops (replaces the reference's separate cleanup kernels,
// /root/reference/csrc/multi_tensor_l2norm_kernel.cu:198-237, and this library's former second launch).
//
// Every tile leaves NSETS partial sums in scratch, then hands in a ticket for its tensor and one for the launch
// (fence, atomicAdd).  The CTA holding a tensor's last ticket reduces that tensor's partials -- one warp, lanes striding
// the tiles, fixed shuffle tree: the order does not depend on which CTA does it or when, so the result is
// deterministic and bit-identical to the former finalize kernel -- and writes sqrt() to ret_pt.  The CTA holding the
// launch's last ticket reduces ALL partials in the order of a 512-thread block_sum (two virtual threads per thread)
// and writes ret_global.  Both put their counter back to zero, so the counters (DevTable::tickets) need no clearing
// between launches, captured graph replays included.  Tensors without elements own no tile: the first tile of the next
// tensor (or the last tile of the launch, for trailing ones) writes their zeros.
struct NormOut {
    float* partials;    // [NSETS][ntiles] scratch
    float* ret_global;  // [NSETS] or NULL
    float* ret_pt;      // [NSETS][ntensors] or NULL
    // Tiles per reference chunk (1 unless the l2norm entry points split a large chunk_size into several tiles to get
    // enough CTAs for an even last wave).  The overflow flag keeps the reference's per-CHUNK meaning
    // (csrc/multi_tensor_l2norm_kernel.cu:103-104: "this chunk's sum of squares is not finite"): the tensor's
    // finalizer rebuilds every chunk's sum from its `sub` consecutive tile partials and tests that.
    int sub;
};

template <int THREADS> constexpr int kSmemFloats = 512 / 32 + 4;

template <int THREADS, int NSETS, bool kQuietWhenFlagged>
__device__ __forceinline__ void norm_tile_done(const DevTable& tb, const Tile& t, uint32_t tile, const NormOut& o,
                                               float* smem, int32_t* noop) {
    static_assert(THREADS == 256, "the global reduction emulates a 512-thread block with two accumulators per thread");
    unsigned* sflag = reinterpret_cast<unsigned*>(smem + 512 / 32);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool per_tensor = o.ret_pt != nullptr || o.sub > 1;
    if (tid == 0) {
        unsigned r = 0;
        uint32_t b = 0, e = 0;
        if (per_tensor) {
            b = __ldg(tb.tile_begin + t.tensor);
            e = __ldg(tb.tile_begin + t.tensor + 1);
        }
        __threadfence();  // this tile's partials (and flag) before its tickets
        // both tickets go out before either answer is looked at: one round trip to L2 instead of two
        unsigned tk = 0, gk = 0;
        if (per_tensor) tk = atomicAdd(tb.tickets + t.tensor, 1u);
        if (o.ret_global) gk = atomicAdd(tb.tickets + tb.ntensors, 1u);
        if (per_tensor) {
            if (tk + 1u == e - b) r |= 1u;
            const bool quiet = kQuietWhenFlagged && *reinterpret_cast<const volatile int32_t*>(noop) != 0;
            if (o.ret_pt && !quiet) {
                if (t.start == 0)
                    for (int k = t.tensor - 1; k >= 0 && __ldg(tb.tile_begin + k) == b; k--)
                        for (int s = 0; s < NSETS; s++) o.ret_pt[(size_t)s * tb.ntensors + k] = 0.f;
                if (tile == tb.ntiles - 1)
                    for (int k = t.tensor + 1; k < tb.ntensors; k++)
                        for (int s = 0; s < NSETS; s++) o.ret_pt[(size_t)s * tb.ntensors + k] = 0.f;
            }
        }
        if (o.ret_global && gk + 1u == tb.ntiles) r |= 2u;
        if (r) __threadfence();  // everybody else's partials after the last ticket
        *sflag = r;
    }
    __syncthreads();
    const unsigned r = *sflag;
    if (r == 0) return;  // uniform across the CTA
    // "_mp" variants leave their outputs alone when the flag is up (csrc/multi_tensor_l2norm_kernel_mp.cu:121-123)
    const bool quiet = kQuietWhenFlagged && *reinterpret_cast<const volatile int32_t*>(noop) != 0;
    if ((r & 1u) && warp == 0) {
        const uint32_t b = __ldg(tb.tile_begin + t.tensor), e = __ldg(tb.tile_begin + t.tensor + 1);
        const uint32_t sub = (uint32_t)o.sub, nchunks = (e - b + sub - 1) / sub;
        bool bad = false;
#pragma unroll
        for (int s = 0; s < NSETS; s++) {
            const float* p = o.partials + (size_t)s * tb.ntiles;
            float acc = 0.f;
            if (sub == 1) {
                // four tiles per lane and trip: the loads go out together, the additions keep the lane's tile order
                for (uint32_t c0 = lane; c0 < nchunks; c0 += 4 * 32) {
                    float cs[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) cs[k] = c0 + 32 * k < nchunks ? __ldcg(p + b + c0 + 32 * k) : 0.f;
#pragma unroll
                    for (int k = 0; k < 4; k++)
                        if (c0 + 32 * k < nchunks) acc = fadd(acc, cs[k]);
                }
            } else {
                for (uint32_t c = lane; c < nchunks; c += 32) {
                    const uint32_t i0 = b + c * sub, i1 = min(i0 + sub, e);
                    float cs = __ldcg(p + i0);
                    for (uint32_t i = i0 + 1; i < i1; i++) cs = fadd(cs, __ldcg(p + i));
                    bad = bad || !finite_f(cs);
                    acc = fadd(acc, cs);
                }
            }
#pragma unroll
            for (int k = 16; k > 0; k >>= 1) acc = fadd(acc, __shfl_xor_sync(0xffffffffu, acc, k));
            if (lane == 0 && o.ret_pt && !quiet) o.ret_pt[(size_t)s * tb.ntensors + t.tensor] = fsqrt(acc);
        }
        if (sub > 1 && __any_sync(0xffffffffu, bad) && lane == 0) *noop = 1;
        if (lane == 0) tb.tickets[t.tensor] = 0u;
    }
    if (r & 2u) {
#pragma unroll
        for (int s = 0; s < NSETS; s++) {
            const float* p = o.partials + (size_t)s * tb.ntiles;
            float a0 = 0.f, a1 = 0.f;  // virtual threads tid and tid + 256 of a 512-thread block
            // eight loads per trip go out together (this runs once, at the very end of the launch, with nothing
            // else left to hide its latency); each accumulator still adds its partials in increasing order
            for (uint32_t i0 = tid; i0 < tb.ntiles; i0 += 4 * 512) {
                float x[4], y[4];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t i = i0 + 512 * k;
                    x[k] = i < tb.ntiles ? __ldcg(p + i) : 0.f;
                    y[k] = i + 256 < tb.ntiles ? __ldcg(p + i + 256) : 0.f;
                }
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t i = i0 + 512 * k;
                    if (i < tb.ntiles) a0 = fadd(a0, x[k]);
                    if (i + 256 < tb.ntiles) a1 = fadd(a1, y[k]);
                }
            }
#pragma unroll
            for (int k = 16; k > 0; k >>= 1) {
                a0 = fadd(a0, __shfl_xor_sync(0xffffffffu, a0, k));
                a1 = fadd(a1, __shfl_xor_sync(0xffffffffu, a1, k));
            }
            __syncthreads();
            if (lane == 0) { smem[warp] = a0; smem[warp + 8] = a1; }
            __syncthreads();
            if (warp == 0) {
                float v = lane < 16 ? smem[lane] : 0.f;
#pragma unroll
                for (int k = 16; k > 0; k >>= 1) v = fadd(v, __shfl_xor_sync(0xffffffffu, v, k));
                if (lane == 0 && !quiet) o.ret_global[s] = fsqrt(v);
            }
        }
        if (tid == 0) tb.tickets[tb.ntensors] = 0u;
    }
}

// ------------------------------------------------------------------------------------------------
// Op concept (all static / const, so everything inlines):
//
//   static constexpr int kLists;                 number of tensor lists
//   template <int D> using T = ...;              element type of list D
//   static constexpr unsigned kRead, kWrite;     bit d set: list d is loaded / stored
//   static constexpr bool kEarlyExit;            return at once if *noop != 0
//   struct Params;                               POD kernel argument
//   struct State;                                per-thread running state (flags, partial sums)
//   __device__ void begin_tensor(int t);         per-tile, uniform: fetch per-tensor scalars
//   __device__ bool load(int d) / store(int d);  runtime (uniform) refinement of kRead / kWrite
//   __device__ void elem(float (&x)[kLists], State&);   one element: inputs in x, outputs written to x
//   static constexpr unsigned kCustomLoad, kCustomStore;  + custom_load<D,G,kVec>() / custom_store<D,G,kVec>()
//                                                (lists the Op moves itself, e.g. over NVLink peer pointers)
//   template <int THREADS> __device__ void end_tile_n(const DevTable&, const Tile&, uint32_t tile, State&, float* smem,
//                                                     int32_t* noop, bool skipped);
//                                                (norm ops: block reduce + per-tile partial + flag + finalize ticket)
//   static constexpr bool kTickets;              the op finalizes in-launch (norm_tile_done): early-exiting CTAs still tick
//   __device__ void finish(State&, int32_t* noop);                  (flag ops)
//
template <class Op, int D> using ListT = typename Op::template T<D>;

template <class Op, int G, int D, bool kVec>
__device__ __forceinline__ void load_list(const Op& op, char* const (&base)[Op::kLists], int64_t idx,
                                          float (&x)[Op::kLists][G]) {
    using T = ListT<Op, D>;
    if constexpr (((Op::kRead >> D) & 1u) != 0) {
        if (op.load(D)) {
            const T* p = reinterpret_cast<const T*>(base[D]) + idx;
            constexpr bool RO = ((Op::kWrite >> D) & 1u) == 0;
            if constexpr (((Op::kCustomLoad >> D) & 1u) != 0) op.template custom_load<D, G, kVec>(p, x[D]);
            else if constexpr (kVec) Group<T, G, RO>::ld(p, x[D]);
            else x[D][0] = Elem<T>::load(p);
            return;
        }
    }
#pragma unroll
    for (int i = 0; i < G; i++) x[D][i] = 0.f;
}

template <class Op, int G, int D, bool kVec>
__device__ __forceinline__ void store_list(const Op& op, char* const (&base)[Op::kLists], int64_t idx,
                                           const float (&x)[Op::kLists][G]) {
    using T = ListT<Op, D>;
    if constexpr (((Op::kWrite >> D) & 1u) != 0) {
        if (op.store(D)) {
            T* p = reinterpret_cast<T*>(base[D]) + idx;
            if constexpr (((Op::kCustomStore >> D) & 1u) != 0) op.template custom_store<D, G, kVec>(p, x[D]);
            else if constexpr (kVec) Group<T, G, false>::st(p, x[D]);
            else Elem<T>::store(p, x[D][0]);
        }
    }
}

template <class Op, int G, bool kVec, int... Ds>
__device__ __forceinline__ void load_all(const Op& op, char* const (&base)[Op::kLists], int64_t idx,
                                         float (&x)[Op::kLists][G], std::integer_sequence<int, Ds...>) {
    (load_list<Op, G, Ds, kVec>(op, base, idx, x), ...);
}
template <class Op, int G, bool kVec, int... Ds>
__device__ __forceinline__ void store_all(const Op& op, char* const (&base)[Op::kLists], int64_t idx,
                                          const float (&x)[Op::kLists][G], std::integer_sequence<int, Ds...>) {
    (store_list<Op, G, Ds, kVec>(op, base, idx, x), ...);
}

template <class Op, int G>
__device__ __forceinline__ void compute_group(const Op& op, float (&x)[Op::kLists][G], typename Op::State& st) {
#pragma unroll
    for (int i = 0; i < G; i++) {
        float e[Op::kLists];
#pragma unroll
        for (int d = 0; d < Op::kLists; d++) e[d] = x[d][i];
        op.elem(e, st);
#pragma unroll
        for (int d = 0; d < Op::kLists; d++) x[d][i] = e[d];
    }
}

// One scalar element (head / tail peel and the mutually-misaligned fallback).
template <class Op>
__device__ __forceinline__ void scalar_elem(const Op& op, char* const (&base)[Op::kLists], int64_t idx,
                                            typename Op::State& st) {
    using Seq = std::make_integer_sequence<int, Op::kLists>;
    float x[Op::kLists][1];
    load_all<Op, 1, false>(op, base, idx, x, Seq{});
    compute_group<Op, 1>(op, x, st);
    store_all<Op, 1, false>(op, base, idx, x, Seq{});
}

// Common head: number of leading elements h in [0, G) after which EVERY list is aligned to its G-element
// vector (G * sizeof(T) bytes).  -1 if the lists are mutually misaligned (no such h).
template <class Op, int G, int... Ds>
__device__ __forceinline__ int common_head(char* const (&base)[Op::kLists], int64_t start,
                                           std::integer_sequence<int, Ds...>) {
    int mis[Op::kLists] = {
        (int)(((reinterpret_cast<uintptr_t>(base[Ds]) / sizeof(ListT<Op, Ds>)) + (uint64_t)start) % G)...};
    bool same = true;
#pragma unroll
    for (int d = 1; d < Op::kLists; d++) same = same && (mis[d] == mis[0]);
    return same ? ((G - mis[0]) % G) : -1;
}

template <class Op, int G, int U, int THREADS>
__device__ __forceinline__ void process_tile(const Op& op, char* const (&base)[Op::kLists], const Tile& t,
                                             typename Op::State& st) {
    using Seq = std::make_integer_sequence<int, Op::kLists>;
    const int tid = threadIdx.x;
    int h = common_head<Op, G>(base, t.start, Seq{});
    if (h < 0) {  // mutually misaligned lists: coalesced element-wise path for this tile
        for (int i = tid; i < t.len; i += THREADS) scalar_elem<Op>(op, base, t.start + i, st);
        return;
    }
    if (h > t.len) h = t.len;
    const int nvec = (t.len - h) / G;
    const int64_t v0 = t.start + h;

    if constexpr (Op::kRawReduce) {
        // One read-only list that is only reduced: keep the groups in flight as RAW words (4 registers per 16-bit
        // group instead of 8 floats) and unpack at consumption, so that twice as many loads fit the register budget.
        using T0 = ListT<Op, 0>;
        constexpr int W = G * (int)sizeof(T0) / 4;
        const T0* p0 = reinterpret_cast<const T0*>(base[0]) + v0;
        for (int j0 = 0; j0 < nvec; j0 += THREADS * U) {
            uint32_t w[U][W];
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int j = j0 + u * THREADS + tid;
                if (j < nvec) RawVec<W, true>::ld(p0 + (int64_t)j * G, w[u]);
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int j = j0 + u * THREADS + tid;
                if (j < nvec) op.consume_raw(w[u], st);
            }
        }
    } else
    for (int j0 = 0; j0 < nvec; j0 += THREADS * U) {
        float x[U][Op::kLists][G];
        if (j0 + THREADS * U <= nvec) {  // full round: all loads are issued before the first use
#pragma unroll
            for (int u = 0; u < U; u++)
                load_all<Op, G, true>(op, base, v0 + (int64_t)(j0 + u * THREADS + tid) * G, x[u], Seq{});
#pragma unroll
            for (int u = 0; u < U; u++) compute_group<Op, G>(op, x[u], st);
#pragma unroll
            for (int u = 0; u < U; u++)
                store_all<Op, G, true>(op, base, v0 + (int64_t)(j0 + u * THREADS + tid) * G, x[u], Seq{});
        } else {
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int j = j0 + u * THREADS + tid;
                if (j < nvec) load_all<Op, G, true>(op, base, v0 + (int64_t)j * G, x[u], Seq{});
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int j = j0 + u * THREADS + tid;
                if (j < nvec) {
                    compute_group<Op, G>(op, x[u], st);
                    store_all<Op, G, true>(op, base, v0 + (int64_t)j * G, x[u], Seq{});
                }
            }
        }
    }
    // scalar edges: [0, h) and [h + nvec*G, len): at most 2*(G-1) elements
    const int tail0 = h + nvec * G;
    const int nedge = h + (t.len - tail0);
    if (tid < nedge) scalar_elem<Op>(op, base, t.start + (tid < h ? tid : tail0 + (tid - h)), st);
}

template <class Op, int... Ds>
__device__ __forceinline__ void fetch_bases(const DevTable& tb, int tensor, char* (&base)[Op::kLists],
                                            std::integer_sequence<int, Ds...>) {
    if constexpr (Op::kInlineTable && Op::kLists <= DevTable::kInlineLists) {
        if (tb.inl) {
            ((base[Ds] = (char*)tb.inl_ptrs[Ds][tensor]), ...);
            return;
        }
    }
    ((base[Ds] = (char*)__ldg(reinterpret_cast<const unsigned long long*>(tb.ptrs) + (size_t)Ds * tb.ntensors + tensor)),
     ...);
}

#ifdef MT_TRACE
// Debug builds only (tools/norm_trace.py): per-CTA globaltimer stamps {entry, tile located, tile streamed, tile handed
// in, SM id} written by thread 0.  One copy per translation unit; the setter lives beside the kernels it traces.
static __device__ unsigned long long* g_mt_trace = nullptr;
__device__ __forceinline__ unsigned long long mt_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void mt_stamp(int slot) {
    if (threadIdx.x == 0 && g_mt_trace) g_mt_trace[(size_t)blockIdx.x * 8 + slot] = mt_now();
}
#define MT_STAMP(slot) mt_stamp(slot)
#else
#define MT_STAMP(slot)
#endif

// The table-driven kernel.  grid = ntiles (one CTA per tile); the loop only matters when the grid is capped.
template <class Op, int G, int U, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
mt_table_kernel(const DevTable tb, const typename Op::Params prm, int32_t* __restrict__ noop) {
    __shared__ float smem[kSmemFloats<THREADS>];
    bool skipped = false;
    if (Op::kEarlyExit) {
        if (*reinterpret_cast<volatile int32_t*>(noop) != 0) {
            // Ops that finalize through completion tickets must still hand their tickets in, or the counters would
            // not return to zero; everything else simply leaves.
            if (!Op::kTickets) return;
            skipped = true;
        }
    }
    Op op(prm);
    typename Op::State st;
    op.init(st);
    using Seq = std::make_integer_sequence<int, Op::kLists>;
    MT_STAMP(0);
    for (uint32_t tile = blockIdx.x; tile < tb.ntiles; tile += gridDim.x) {
        const Tile t = locate_tile<Op::kInlineTable>(tb, tile);
        if (!skipped) {
            char* base[Op::kLists];
            fetch_bases<Op>(tb, t.tensor, base, Seq{});
            op.begin_tensor(t.tensor);
            MT_STAMP(1);
            process_tile<Op, G, U, THREADS>(op, base, t, st);
            MT_STAMP(2);
        }
        op.template end_tile_n<THREADS>(tb, t, tile, st, smem, noop, skipped);
        MT_STAMP(3);
    }
#ifdef MT_TRACE
    if (threadIdx.x == 0 && g_mt_trace) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        g_mt_trace[(size_t)blockIdx.x * 8 + 4] = smid;
    }
#endif
    op.finish(st, noop);
}

// Ops that neither keep per-tensor scalars nor reduce inherit these no-ops.
struct OpBase {
    struct State {};
    static constexpr bool kEarlyExit = false;
    // bit d set: list d is moved by the Op's own custom_load / custom_store (peer-memory ops) instead of Group<>
    static constexpr unsigned kCustomLoad = 0, kCustomStore = 0;
    __device__ __forceinline__ void init(State&) const {}
    __device__ __forceinline__ void begin_tensor(int) {}
    __device__ __forceinline__ bool load(int) const { return true; }
    __device__ __forceinline__ bool store(int) const { return true; }
    static constexpr bool kTickets = false;  // finalizes inside the launch through DevTable::tickets (norm_tile_done)
    static constexpr bool kRawReduce = false;  // single read-only list consumed through consume_raw() (process_tile)
    // Small tables are read from the kernel arguments (DevTable::inl).  The sharded-update ops opt out: their launches
    // are never small, and the second lookup path costs them registers they do not have (hot-loop spills under the
    // 64-register cap, tools/sass_summary.py).
    static constexpr bool kInlineTable = true;
    template <int THREADS, class S>
    __device__ __forceinline__ void end_tile_n(const DevTable&, const Tile&, uint32_t, S&, float*, int32_t*, bool) const {}
    template <class S> __device__ __forceinline__ void finish(S&, int32_t*) const {}
};

// Flag state shared by scale / axpby: any thread that saw a non-finite input raises the flag once per CTA.
struct FiniteState {
    bool finite;
};
__device__ __forceinline__ void finish_finite(const FiniteState& st, int32_t* noop) {
    if (__syncthreads_or(!st.finite)) {
        if (threadIdx.x == 0) *noop = 1;
    }
}

}  // namespace mt
