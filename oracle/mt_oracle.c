/*
 * mt_oracle.c -- CPU restatement of the reference's multi-tensor-apply functors.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity checker for the CUDA
 * product in beforeholiday_b200/csrc; it is never linked into, imported by or
 * executed from the product path.  Only tests/, __graft_entry__.smoke(),
 * bench.py's cpu_baseline / --impl reference legs and tools/bench_ops.py (the same
 * baseline leg for the other configs) may load it.
 *
 * Arithmetic contract: every functor is restated in IEEE-754 binary32 with the
 * SAME operation order as the reference's device code and NO fused
 * multiply-add (build with -ffp-contract=off, no -ffast-math).  The product
 * kernels use the non-contractable __f{add,mul,div,sqrt}_rn intrinsics in the
 * same order, so fp32 results are expected to be bit-identical to this file;
 * against the reference's own CUDA build (--use_fast_math: approximate
 * div/sqrt, FTZ, FMA contraction) they agree to a few ulp.
 *
 * Parity pin (there is no separate README): (1) the known answers of the reference's unit tests
 * (tests/L0/run_amp/test_multi_tensor_*.py) and torch.optim / RefLAMB / Novograd trajectories on CPU
 * (tests/test_oracle_pins.py, fixtures tests/golden/ref_cpu.pt and ref_novograd.pt with their generating scripts);
 * (2) the reference's own CUDA kernels, compiled unmodified by oracle/build_ref.py into oracle/_ref/ and run beside
 * this file on the same inputs on a B200 (tests/test_gpu_vs_reference.py) -- live comparison, no stored CUDA goldens.
 *
 * dtype tags: 0 = float32, 1 = float16, 2 = bfloat16, 3 = float64, 4 = e5m2 byte packing.
 * Citations are file:line in /root/reference.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

enum { DT_F32 = 0, DT_F16 = 1, DT_BF16 = 2, DT_F64 = 3, DT_E5M2 = 4 };

/* ---------------------------------------------------------------- 16-bit conversions */

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* float -> IEEE half, round-to-nearest-even (what static_cast<at::Half>(float) does on device: __float2half_rn) */
static inline uint16_t f32_to_f16(float f) {
    uint32_t x = f2u(f);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t ax = x & 0x7fffffffu;
    if (ax > 0x7f800000u) return (uint16_t)(sign | 0x7fffu);            /* NaN (canonical, as CUDA) */
    if (ax >= 0x47800000u) return (uint16_t)(sign | 0x7c00u);           /* |f| >= 65536 -> inf (incl. inf) */
    if (ax >= 0x38800000u) {                                            /* normal half */
        uint32_t mant = ax & 0x7fffffu;
        uint32_t exp = (ax >> 23) - 112u;
        uint32_t h = (exp << 10) | (mant >> 13);
        uint32_t rem = mant & 0x1fffu;
        if (rem > 0x1000u || (rem == 0x1000u && (h & 1u))) h++;         /* may carry into exp / inf: correct */
        return (uint16_t)(sign | h);
    }
    if (ax < 0x33000000u) return (uint16_t)sign;                        /* < 2^-25 -> 0 */
    {                                                                   /* subnormal half */
        uint32_t mant = (ax & 0x7fffffu) | 0x800000u;
        int shift = 126 - (int)(ax >> 23);                              /* 14..24 */
        uint32_t h = mant >> shift;
        uint32_t rem = mant & ((1u << shift) - 1u);
        uint32_t half = 1u << (shift - 1);
        if (rem > half || (rem == half && (h & 1u))) h++;
        return (uint16_t)(sign | h);
    }
}

static inline float f16_to_f32(uint16_t h) {
    uint32_t sign = ((uint32_t)h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu;
    uint32_t mant = h & 0x3ffu;
    if (exp == 0x1f) return u2f(sign | 0x7f800000u | (mant << 13));
    if (exp == 0) {
        if (mant == 0) return u2f(sign);
        int e = -1;
        do { e++; mant <<= 1; } while (!(mant & 0x400u));
        mant &= 0x3ffu;
        return u2f(sign | ((uint32_t)(127 - 15 - e) << 23) | (mant << 13));
    }
    return u2f(sign | ((exp + 112u) << 23) | (mant << 13));
}

/* float -> bfloat16 round-to-nearest-even (__float2bfloat16_rn) */
static inline uint16_t f32_to_bf16(float f) {
    uint32_t x = f2u(f);
    if ((x & 0x7fffffffu) > 0x7f800000u) return (uint16_t)0x7fffu;      /* NaN (canonical, as CUDA) */
    uint32_t lsb = (x >> 16) & 1u;
    x += 0x7fffu + lsb;
    return (uint16_t)(x >> 16);
}
static inline float bf16_to_f32(uint16_t h) { return u2f((uint32_t)h << 16); }

/* e5m2 packing of the reference's e5m2_allgather (apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb_kernel.cu:34-66,
 * fused_adam_cuda_kernel.cu:396-428): the byte is the HIGH byte of the fp16 value of (v + 2^floor(log2|v|) / 8), i.e.
 * half an e5m2 ulp is added in magnitude before the fp16 rounding and the low mantissa byte is dropped; unpacking puts
 * the byte back as the high byte of an fp16 with a zero low byte. */
static inline uint8_t f32_to_e5m2(float v) {
    const float pow2 = u2f(f2u(v) & 0xFF800000u);
    return (uint8_t)(f32_to_f16(v + pow2 / 8.0f) >> 8);
}
static inline float e5m2_to_f32(uint8_t b) { return f16_to_f32((uint16_t)((uint16_t)b << 8)); }

static inline float ld(const void* p, int dt, int64_t i) {
    switch (dt) {
        case DT_E5M2: return e5m2_to_f32(((const uint8_t*)p)[i]);
        case DT_F32: return ((const float*)p)[i];
        case DT_F16: return f16_to_f32(((const uint16_t*)p)[i]);
        case DT_BF16: return bf16_to_f32(((const uint16_t*)p)[i]);
        default: return (float)((const double*)p)[i];
    }
}
static inline void st(void* p, int dt, int64_t i, float v) {
    switch (dt) {
        case DT_E5M2: ((uint8_t*)p)[i] = f32_to_e5m2(v); break;
        case DT_F32: ((float*)p)[i] = v; break;
        case DT_F16: ((uint16_t*)p)[i] = f32_to_f16(v); break;
        case DT_BF16: ((uint16_t*)p)[i] = f32_to_bf16(v); break;
        default: ((double*)p)[i] = (double)v; break;
    }
}
/* round a float through storage type dt (models arithmetic carried out "in T") */
static inline float rnd(float v, int dt) {
    switch (dt) {
        case DT_F16: return f16_to_f32(f32_to_f16(v));
        case DT_BF16: return bf16_to_f32(f32_to_bf16(v));
        case DT_E5M2: return e5m2_to_f32(f32_to_e5m2(v));
        default: return v;
    }
}

/* exported so the python side can test the converters against torch */
void orc_convert(const void* in, int in_dt, void* out, int out_dt, int64_t n) {
    for (int64_t i = 0; i < n; i++) st(out, out_dt, i, ld(in, in_dt, i));
}

/* ---------------------------------------------------------------- scale / axpby */

/* ScaleFunctor: csrc/multi_tensor_scale_kernel.cu:68-69 (value), :108-109 (flag).
 * out = float(in) * scale ; flag raised iff any INPUT element is non-finite. Returns the flag. */
int orc_scale(const void* in, int in_dt, void* out, int out_dt, int64_t n, float scale) {
    int flag = 0;
#pragma omp parallel for reduction(| : flag) schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float x = ld(in, in_dt, i);
        if (!isfinite(x)) flag |= 1;
        st(out, out_dt, i, x * scale);
    }
    return flag;
}

/* AxpbyFunctor: csrc/multi_tensor_axpby_kernel.cu:73-79.
 * out = a*float(x) + b*float(y); finite check on x (0), y (1) or both (-1). */
int orc_axpby(const void* x, int x_dt, const void* y, int y_dt, void* out, int out_dt, int64_t n,
              float a, float b, int arg_to_check) {
    int flag = 0;
#pragma omp parallel for reduction(| : flag) schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float xv = ld(x, x_dt, i), yv = ld(y, y_dt, i);
        float ax = a * xv;
        float by = b * yv;
        st(out, out_dt, i, ax + by);
        if (arg_to_check == -1) { if (!(isfinite(xv) && isfinite(yv))) flag |= 1; }
        if (arg_to_check == 0) { if (!isfinite(xv)) flag |= 1; }
        if (arg_to_check == 1) { if (!isfinite(yv)) flag |= 1; }
    }
    return flag;
}

/* ---------------------------------------------------------------- Adam */

/* Host-side bias corrections: csrc/multi_tensor_adam.cu:145-149.
 * std::pow(float, int) promotes to double; the result is rounded to float. */
void orc_bias_corrections(float beta1, float beta2, int step, int bias_correction, float* bc1, float* bc2) {
    *bc1 = 1.0f; *bc2 = 1.0f;
    if (bias_correction == 1) {
        *bc1 = (float)(1.0 - pow((double)beta1, (double)step));
        *bc2 = (float)(1.0 - pow((double)beta2, (double)step));
    }
}

/* AdamFunctor: csrc/multi_tensor_adam.cu:94-112.  All four lists share dtype dt; math is float
 * (MATH_T = float, :21) even for dt == double.  mode 0 = L2, mode 1 = AdamW. */
void orc_adam(void* g, void* p, void* m, void* v, int dt, int64_t n,
              float lr, float beta1, float beta2, float eps, int step, int mode, int bias_correction, float decay) {
    float bc1, bc2;
    orc_bias_corrections(beta1, beta2, step, bias_correction, &bc1, &bc2);
    const float omb1 = 1 - beta1, omb2 = 1 - beta2;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float rg = ld(g, dt, i), rp = ld(p, dt, i), rm = ld(m, dt, i), rv = ld(v, dt, i);
        if (mode == 0) {
            float dp = decay * rp;
            rg = rg + dp;
            float t1 = beta1 * rm, t2 = omb1 * rg;
            rm = t1 + t2;
            float t3 = beta2 * rv, t4 = omb2 * rg; t4 = t4 * rg;
            rv = t3 + t4;
            float mu = rm / bc1, vu = rv / bc2;
            float denom = sqrtf(vu) + eps;
            float upd = mu / denom;
            float lu = lr * upd;
            rp = rp - lu;
        } else {
            float t1 = beta1 * rm, t2 = omb1 * rg;
            rm = t1 + t2;
            float t3 = beta2 * rv, t4 = omb2 * rg; t4 = t4 * rg;
            rv = t3 + t4;
            float mu = rm / bc1, vu = rv / bc2;
            float denom = sqrtf(vu) + eps;
            float q = mu / denom;
            float dp = decay * rp;
            float upd = q + dp;
            float lu = lr * upd;
            rp = rp - lu;
        }
        st(p, dt, i, rp); st(m, dt, i, rm); st(v, dt, i, rv);
    }
}

/* DistAdamFunctor: apex/contrib/csrc/optimizers/multi_tensor_distopt_adam_kernel.cu:103-116.
 * p,m,v are float (acc type), g and the optional p_copy are g_dt.  mode 0: eps under the
 * square root, mode 1: eps outside.  Bias corrections are evaluated in float on the device
 * (:59-63, std::pow(float,int), which promotes to double: double pow() here and in the product).  If `noop` is non-NULL and *noop != 0 the
 * update is skipped (product extension; the reference ignores the flag). */
void orc_dist_adam(float* p, float* m, float* v, const void* g, int g_dt, void* p_copy, int64_t n,
                   float b1, float b2, int bias_correction, float eps, float decay,
                   float lr, float grad_scale, int step, int mode, const int* noop) {
    if (noop && *noop) return;
    float bc1 = 1.0f, bc2 = 1.0f;
    if (bias_correction == 1) {
        bc1 = (float)(1.0 - pow((double)b1, (double)step));
        bc2 = (float)(1.0 - pow((double)b2, (double)step));
    }
    const float omb1 = 1 - b1, omb2 = 1 - b2;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float rp = p[i], rm = m[i], rv = v[i];
        float sg = ld(g, g_dt, i) / grad_scale;
        float t1 = b1 * rm, t2 = omb1 * sg;
        rm = t1 + t2;
        float t3 = b2 * rv, t4 = omb2 * sg; t4 = t4 * sg;
        rv = t3 + t4;
        float mu = rm / bc1, vu = rv / bc2;
        float denom;
        if (mode == 0) denom = sqrtf(vu + eps); else denom = sqrtf(vu) + eps;
        float q = mu / denom;
        float dp = decay * rp;
        float upd = q + dp;
        float lu = lr * upd;
        rp = rp - lu;
        p[i] = rp; m[i] = rm; v[i] = rv;
        if (p_copy) st(p_copy, g_dt, i, rp);
    }
}

/* ---------------------------------------------------------------- SGD */

/* SGDFunctor: csrc/multi_tensor_sgd_kernel.cu:87-134.  Returns without touching anything when
 * *noop is set (:46).  model_out (fp16, may be NULL) is cast from the STORED weight (:130). */
void orc_sgd(const void* g, int g_dt, void* w, int w_dt, void* mom, void* model_out, int64_t n,
             float wd, float momentum, float dampening, float lr, int nesterov, int first_run,
             int wd_after_momentum, float scale, const int* noop) {
    if (noop && *noop) return;
    const float omd = 1.f - dampening;
    const float nlr = -lr;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float ig = ld(g, g_dt, i) * scale;
        float iw = ld(w, w_dt, i);
        float im = ld(mom, w_dt, i);
        if (wd != 0.f && !wd_after_momentum) { float t = wd * iw; ig = ig + t; }
        if (momentum != 0.f) {
            if (!first_run) { float a = im * momentum, b = omd * ig; im = a + b; }
            else im = ig;
            if (nesterov) { float t = momentum * im; ig = ig + t; }
            else ig = im;
        }
        if (wd != 0.f && wd_after_momentum) { float t = wd * iw; ig = ig + t; }
        float d = nlr * ig;
        /* weight_in[i] += (-lr * g).  For T_weight = float this is a float add.  For a 16-bit T_weight the only
         * viable overload is operator+=(T&, const T&) (torch/headeronly/util/Half.h:529, BFloat16 likewise), so the
         * float increment is first ROUNDED to T, then T+T is evaluated in float and rounded on the store. */
        float nw = iw + rnd(d, w_dt);
        st(w, w_dt, i, nw);
        if (model_out) st(model_out, DT_F16, i, ld(w, w_dt, i));
        if (momentum != 0.f) st(mom, w_dt, i, im);
    }
}

/* ---------------------------------------------------------------- L2 norm (faithful block simulation) */

#define L2_BLOCK 512
#define L2_ILP 4

/* reduce_block_into_lanes, lanes=1: csrc/type_shim.h:421-466 (blockSize 512). x has 512 entries. */
static float block_reduce_512(float* x) {
    for (int i = L2_BLOCK >> 1; i >= 64; i >>= 1)
        for (int t = 0; t < i; t++) x[t] = x[t] + x[t + i];
    float w[32];
    for (int t = 0; t < 32; t++) w[t] = x[t] + x[t + 32];
    for (int i = 16; i >= 1; i >>= 1)
        for (int t = 0; t < 32; t++) w[t] = w[t] + ((t + i < 32) ? w[t + i] : w[t]); /* shfl_down: OOR lanes read own */
    return w[0];
}

/* One chunk of L2NormFunctor: csrc/multi_tensor_l2norm_kernel.cu:56-98.
 * x points at the chunk start, n = elements of the tensor remaining from there.
 * `aligned` tells whether the reference would take its vector path (:67). */
static float l2norm_chunk(const void* x, int dt, int64_t off, int64_t n, int chunk_size, int aligned) {
    float part[L2_BLOCK];
    int64_t lim = n < chunk_size ? n : chunk_size;
    for (int tid = 0; tid < L2_BLOCK; tid++) {
        float vals[L2_ILP] = {0.f, 0.f, 0.f, 0.f};
        if (aligned) {
            for (int64_t is = tid; is * L2_ILP < lim; is += L2_BLOCK)
                for (int ii = 0; ii < L2_ILP; ii++) {
                    float nx = ld(x, dt, off + is * L2_ILP + ii);
                    float sq = nx * nx;
                    vals[ii] = vals[ii] + sq;
                }
        } else {
            for (int64_t is = 0; is < lim; is += (int64_t)L2_BLOCK * L2_ILP)
                for (int ii = 0; ii < L2_ILP; ii++) {
                    int64_t i = is + tid + (int64_t)ii * L2_BLOCK;
                    if (i < lim) {
                        float nx = ld(x, dt, off + i);
                        float sq = nx * nx;
                        vals[ii] = vals[ii] + sq;
                    }
                }
        }
        float val = 0.f;
        for (int i = 0; i < L2_ILP; i++) val = val + vals[i];
        part[tid] = val;
    }
    return block_reduce_512(part);
}

static int esize(int dt) { return dt == DT_F32 ? 4 : (dt == DT_F64 ? 8 : (dt == DT_E5M2 ? 1 : 2)); }

/* multi_tensor_l2norm_cuda: csrc/multi_tensor_l2norm_kernel.cu:301-364 with the 320-block / 110-tensor
 * launcher of csrc/multi_tensor_apply_base.cuh:16-17,84-143 and cleanup (:198-237).
 *   xs[t], numel[t]   tensors (byte address alignment of xs[t] decides the vector path, as on device)
 *   ret[1]            global norm
 *   ret_per_tensor[ntensors]  (may be NULL = per_tensor false)
 * Optional out_scale path (l2norm_scale, csrc/multi_tensor_l2norm_scale_kernel.cu:79-81): if outs != NULL,
 * outs[t][i] = float(x)*scale is also written (alignment of out also feeds the path decision there, :70).
 * Returns the noop flag (non-finite chunk partial, :103-104). */
int orc_l2norm(const void* const* xs, const int64_t* numel, int64_t ntensors, int dt, int chunk_size,
               float* ret, float* ret_per_tensor,
               void* const* outs, int out_dt, float scale) {
    enum { MAX_BLOCKS = 320, MAX_TENSORS_D1 = 110, MAX_TENSORS_D2 = 64 };
    const int max_tensors = outs ? MAX_TENSORS_D2 : MAX_TENSORS_D1;
    float output[MAX_BLOCKS];
    for (int i = 0; i < MAX_BLOCKS; i++) output[i] = 0.f;
    int flag = 0;
    int64_t max_chunks = 0;
    for (int64_t t = 0; t < ntensors; t++) {
        int64_t c = (numel[t] + chunk_size - 1) / chunk_size;
        if (c > max_chunks) max_chunks = c;
    }
    float* opt = NULL;
    if (ret_per_tensor) opt = (float*)calloc((size_t)(ntensors * (max_chunks > 0 ? max_chunks : 1)), sizeof(float));

    int loc_block = 0, loc_tensor = 0;
    for (int64_t t = 0; t < ntensors; t++) {
        if (numel[t] == 0) continue;
        loc_tensor++;
        int64_t chunks = (numel[t] + chunk_size - 1) / chunk_size;
        for (int64_t c = 0; c < chunks; c++) {
            int64_t off = c * (int64_t)chunk_size;
            int64_t n = numel[t] - off;
            uintptr_t addr = (uintptr_t)xs[t] + (uintptr_t)(off * esize(dt));
            int aligned = (n % L2_ILP == 0) && (chunk_size % L2_ILP == 0) && (addr % (L2_ILP * esize(dt)) == 0);
            if (outs) {
                uintptr_t oaddr = (uintptr_t)outs[t] + (uintptr_t)(off * esize(out_dt));
                aligned = aligned && (oaddr % (L2_ILP * esize(out_dt)) == 0);
                int64_t lim = n < chunk_size ? n : chunk_size;
                for (int64_t i = 0; i < lim; i++) st(outs[t], out_dt, off + i, ld(xs[t], dt, off + i) * scale);
            }
            float fin = l2norm_chunk(xs[t], dt, off, n, chunk_size, aligned);
            if (!isfinite(fin)) flag = 1;
            output[loc_block] = output[loc_block] + fin;
            /* NB: the reference indexes by (start_tensor_this_launch + tensor_loc), which equals t when no
             * tensor is empty; with empties it drifts.  t is used here. */
            if (opt) opt[t * max_chunks + c] = fin;
            loc_block++;
            int tensors_full = (loc_tensor == max_tensors && c == chunks - 1);
            int blocks_full = (loc_block == MAX_BLOCKS);
            int last = (t == ntensors - 1 && c == chunks - 1);
            if (tensors_full || blocks_full || last) {
                loc_block = 0;
                loc_tensor = (c == chunks - 1) ? 0 : 1;
            }
        }
    }
    /* cleanup, block 0 (:210-221) */
    {
        float x[L2_BLOCK];
        for (int t = 0; t < L2_BLOCK; t++) x[t] = t < MAX_BLOCKS ? output[t] : 0.f;
        ret[0] = (float)sqrt((double)block_reduce_512(x)); /* sqrt(float) in device code == sqrtf, correctly rounded either way */
    }
    if (ret_per_tensor) {
        for (int64_t t = 0; t < ntensors; t++) {
            float x[L2_BLOCK];
            for (int tid = 0; tid < L2_BLOCK; tid++) {
                float val = 0.f;
                for (int64_t i = tid; i < max_chunks; i += L2_BLOCK) val = val + opt[t * max_chunks + i];
                x[tid] = val;
            }
            ret_per_tensor[t] = sqrtf(block_reduce_512(x));
        }
        free(opt);
    }
    return flag;
}

/* ---------------------------------------------------------------- LAMB (fused op, csrc/multi_tensor_lamb.cu) */

/* LAMBStage1Functor: csrc/multi_tensor_lamb.cu:66,121-143.  Writes the update over g. */
void orc_lamb_stage1(void* g, const void* p, void* m, void* v, int dt, int64_t n,
                     float beta1, float beta2, float beta3, float bc1, float bc2, float eps,
                     int mode, float decay, float global_grad_norm, float max_global_grad_norm) {
    const float clip = global_grad_norm > max_global_grad_norm ? global_grad_norm / max_global_grad_norm : 1.0f;
    const float omb2 = 1 - beta2;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float rg = ld(g, dt, i), rm = ld(m, dt, i), rv = ld(v, dt, i);
        float rp = (decay == 0) ? 0.f : ld(p, dt, i);               /* :102-103, :123-128 */
        float sg = rg / clip;
        if (mode == 0) { float dp = decay * rp; sg = sg + dp; }
        float t1 = rm * beta1, t2 = beta3 * sg;
        rm = t1 + t2;
        float t3 = rv * beta2, t4 = omb2 * sg; t4 = t4 * sg;
        rv = t3 + t4;
        float mu = rm / bc1, vu = rv / bc2;
        float denom = sqrtf(vu) + eps;
        float q = mu / denom;
        float u;
        if (mode == 0) u = q; else { float dp = decay * rp; u = q + dp; }
        st(g, dt, i, u); st(m, dt, i, rm); st(v, dt, i, rv);
    }
}

/* LAMBStage2Functor: csrc/multi_tensor_lamb.cu:255-263,289. */
void orc_lamb_stage2(const void* u, void* p, int dt, int64_t n, float lr, float decay, int use_nvlamb,
                     float param_norm, float update_norm) {
    float ratio = lr;
    if (use_nvlamb || decay != 0.0f)
        ratio = (update_norm != 0.0f && param_norm != 0.0f) ? lr * (param_norm / update_norm) : lr;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float ru = ratio * ld(u, dt, i);
        st(p, dt, i, ld(p, dt, i) - ru);
    }
}

/* multi_tensor_lamb_cuda: csrc/multi_tensor_lamb.cu:332-413 (four phases on one list set). */
void orc_lamb(void* const* gs, void* const* ps, void* const* ms, void* const* vs, const int64_t* numel,
              int64_t ntensors, int dt, int chunk_size,
              float lr, float beta1, float beta2, float eps, int step, int bias_correction, float decay,
              int grad_averaging, int mode, float global_grad_norm, float max_grad_norm, int use_nvlamb,
              float* out_param_norms, float* out_update_norms) {
    float bc1, bc2;
    orc_bias_corrections(beta1, beta2, step, bias_correction, &bc1, &bc2);
    float beta3 = 1.0f;
    if (grad_averaging == 1) beta3 = 1 - beta1;
    float* pn = (float*)malloc(sizeof(float) * (size_t)(ntensors > 0 ? ntensors : 1));
    float* un = (float*)malloc(sizeof(float) * (size_t)(ntensors > 0 ? ntensors : 1));
    float tot;
    orc_l2norm((const void* const*)ps, numel, ntensors, dt, chunk_size, &tot, pn, NULL, 0, 0.f);
    for (int64_t t = 0; t < ntensors; t++)
        orc_lamb_stage1(gs[t], ps[t], ms[t], vs[t], dt, numel[t], beta1, beta2, beta3, bc1, bc2, eps, mode, decay,
                        global_grad_norm, max_grad_norm);
    orc_l2norm((const void* const*)gs, numel, ntensors, dt, chunk_size, &tot, un, NULL, 0, 0.f);
    for (int64_t t = 0; t < ntensors; t++)
        orc_lamb_stage2(gs[t], ps[t], dt, numel[t], lr, decay, use_nvlamb, pn[t], un[t]);
    if (out_param_norms) memcpy(out_param_norms, pn, sizeof(float) * (size_t)ntensors);
    if (out_update_norms) memcpy(out_update_norms, un, sizeof(float) * (size_t)ntensors);
    free(pn); free(un);
}

/* ---------------------------------------------------------------- "_mp" LAMB (csrc/multi_tensor_lamb_mp.cu) */

/* LAMBStage1Functor<T,param_t>: csrc/multi_tensor_lamb_mp.cu:58-75,118-151.  g (g_dt) is multiplied by inv_scale;
 * p, m, v are p_dt; the update is written over g in g_dt.  The reference evaluates 1 - pow(beta, step) in-kernel
 * with a fast-math float pow; oracle and product use double pow rounded to float. */
void orc_lamb_mp_stage1(void* g, int g_dt, const void* p, void* m, void* v, int p_dt, int64_t n,
                        float beta1, float beta2, float beta3, int step, int bias_correction, float eps, int mode,
                        float decay, float global_grad_norm, float max_global_grad_norm, float inv_scale) {
    const float clip = global_grad_norm > max_global_grad_norm ? global_grad_norm / max_global_grad_norm : 1.0f;
    float bc1 = 1.0f, bc2 = 1.0f;
    if (bias_correction == 1) {
        bc1 = (float)(1.0 - pow((double)beta1, (double)step));
        bc2 = (float)(1.0 - pow((double)beta2, (double)step));
    }
    const float omb2 = 1 - beta2;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float rg = ld(g, g_dt, i) * inv_scale, rm = ld(m, p_dt, i), rv = ld(v, p_dt, i);
        float rp = (decay == 0) ? 0.f : ld(p, p_dt, i);
        float sg = rg / clip;
        if (mode == 0) { float dp = decay * rp; sg = sg + dp; }
        float t1 = rm * beta1, t2 = beta3 * sg;
        rm = t1 + t2;
        float t3 = rv * beta2, t4 = omb2 * sg; t4 = t4 * sg;
        rv = t3 + t4;
        float mu = rm / bc1, vu = rv / bc2;
        float denom = sqrtf(vu) + eps;
        float q = mu / denom;
        float u;
        if (mode == 0) u = q; else { float dp = decay * rp; u = q + dp; }
        st(g, g_dt, i, u); st(m, p_dt, i, rm); st(v, p_dt, i, rv);
    }
}

/* LAMBStage2Functor<T,N,param_t>: csrc/multi_tensor_lamb_mp.cu:258-266,316-323. out (g_dt, may be NULL) = stored p. */
void orc_lamb_mp_stage2(const void* u, int g_dt, void* p, int p_dt, void* out, int64_t n, float lr, float decay,
                        int use_nvlamb, float param_norm, float update_norm) {
    float ratio = lr;
    if (use_nvlamb || decay != 0.0f)
        ratio = (update_norm != 0.0f && param_norm != 0.0f) ? lr * (param_norm / update_norm) : lr;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float ru = ratio * ld(u, g_dt, i);
        st(p, p_dt, i, ld(p, p_dt, i) - ru);
        if (out) st(out, g_dt, i, ld(p, p_dt, i));
    }
}

/* ---------------------------------------------------------------- sharded LAMB (apex/contrib/csrc/optimizers) */

/* DistOptLAMBStage1Functor<T=float, GRAD_T, MATH_T=float>:
 * apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb_kernel.cu:133-148 (scalars), :212-249 / :283-304 (element).
 * combined_scale folds the clip factor into the loss-scale divisor; the "+ 1e-6" literal makes that quotient a double
 * expression.  beta3 = 1 - beta1 regardless of the per-tensor beta3 array.  bias corrections: double pow rounded to
 * float (the reference's fast-math device pow differs in the last ulps, as for lamb_mp).
 * sq_out (may be NULL): [0] += sum round_to(model_dt)(p)^2, [1] += sum u^2 -- double accumulation; this is what the
 * product's fused norm extension is checked against (to 1e-5 relative). */
void orc_dist_lamb_stage1(const void* g, int g_dt, const float* p, float* m, float* v, float* u, int64_t n,
                          float beta1, float beta2, int bias_correction, int step, float eps, int mode, float decay,
                          float global_scale, float global_grad_norm, float max_grad_norm, int model_dt,
                          double* sq_out) {
    float cs = global_scale;
    if (max_grad_norm > 0) {
        float q = global_grad_norm / global_scale;
        cs = (float)((double)max_grad_norm / ((double)q + 1e-6));
        cs = global_scale / (1.0f < cs ? 1.0f : cs);
    }
    const float beta3 = 1 - beta1;
    float bc1 = 1.0f, bc2 = 1.0f;
    if (bias_correction == 1) {
        bc1 = (float)(1.0 - pow((double)beta1, (double)step));
        bc2 = (float)(1.0 - pow((double)beta2, (double)step));
    }
    const float omb2 = 1 - beta2;
    double sp = 0.0, su = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : sp, su)
    for (int64_t i = 0; i < n; i++) {
        float rg = ld(g, g_dt, i), rm = m[i], rv = v[i];
        float rp = (decay == 0) ? 0.f : p[i];
        float pm = rnd(p[i], model_dt);
        sp += (double)pm * (double)pm;
        float sg = rg / cs;
        if (mode == 0) { float dp = decay * rp; sg = sg + dp; }
        float t1 = rm * beta1, t2 = beta3 * sg;
        rm = t1 + t2;
        float t3 = rv * beta2, t4 = omb2 * sg; t4 = t4 * sg;
        rv = t3 + t4;
        float mu = rm / bc1, vu = rv / bc2;
        float denom = sqrtf(vu) + eps;
        float q = mu / denom;
        float uu;
        if (mode == 0) uu = q; else { float dp = decay * rp; uu = q + dp; }
        su += (double)uu * (double)uu;
        u[i] = uu; m[i] = rm; v[i] = rv;
    }
    if (sq_out) { sq_out[0] += sp; sq_out[1] += su; }
}

/* DistOptLAMBStage2Functor: multi_tensor_distopt_lamb_kernel.cu:342-351 (ratio), :376-383 / :407-419 (element).
 * p_copy (copy_dt) = convert(new p). */
void orc_dist_lamb_stage2(const float* u, float* p, void* p_copy, int copy_dt, int64_t n, float lr, float decay,
                          int use_nvlamb, float param_norm, float update_norm) {
    float ratio = lr;
    if (use_nvlamb || decay != 0.0f)
        ratio = (update_norm != 0.0f && param_norm != 0.0f) ? lr * (param_norm / update_norm) : lr;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float ru = ratio * u[i];
        float w = p[i] - ru;
        p[i] = w;
        st(p_copy, copy_dt, i, w);
    }
}

/* ---------------------------------------------------------------- Adagrad / NovoGrad / LARS */

/* AdagradFunctor: csrc/multi_tensor_adagrad.cu:63-72. */
void orc_adagrad(const void* g, void* p, void* h, int dt, int64_t n, float lr, float eps, int mode, float decay) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float rg = ld(g, dt, i), rp = ld(p, dt, i), rh = ld(h, dt, i);
        if (mode == 0) {
            float t = decay * rp; rg = rg + t;
            float sq = rg * rg; rh = rh + sq;
            float den = sqrtf(rh) + eps; float q = rg / den; float lq = lr * q;
            rp = rp - lq;
        } else {
            float sq = rg * rg; rh = rh + sq;
            float den = sqrtf(rh) + eps; float q = rg / den; float t = decay * rp; float u = q + t; float lu = lr * u;
            rp = rp - lu;
        }
        st(p, dt, i, rp); st(h, dt, i, rh);
    }
}

/* Per-tensor norm blend of multi_tensor_norm_out_cuda: csrc/multi_tensor_l2norm_kernel.cu:367-456 (host), :113-195
 * (MaxNormFunctor), :272-298 (cleanup_v2).  Chunk partials are max|x| or sum x^2 (plain sequential fp32 here; the
 * product's reduction order differs, tolerance 1e-5).  Returns the flag (non-finite chunk partial). */
int orc_norm_blend(const void* x, int dt, int64_t n, int chunk_size, float* out, int norm_type, float alpha, float beta) {
    int flag = 0;
    float acc = 0.f;
    for (int64_t c = 0; c < n; c += chunk_size) {
        int64_t e = c + chunk_size < n ? c + chunk_size : n;
        float part = 0.f;
        if (norm_type == 0) { for (int64_t i = c; i < e; i++) part = fmaxf(fabsf(part), fabsf(ld(x, dt, i))); }
        else { double d = 0; for (int64_t i = c; i < e; i++) { float xv = ld(x, dt, i); d += (double)(xv * xv); } part = (float)d; }
        if (!isfinite(part)) flag = 1;
        if (norm_type == 0) acc = fmaxf(fabsf(acc), fabsf(part)); else acc = acc + part;
    }
    if (norm_type == 0) { float a = alpha * (*out), b = beta * acc; *out = a + b; }
    else { float a = alpha * (*out); a = a * (*out); float b = beta * acc; *out = sqrtf(a + b); }
    return flag;
}

/* NovoGradFunctor: csrc/multi_tensor_novograd.cu:90-107; host part :145-156 (bc2 = sqrt(1 - beta2^step)). */
void orc_novograd(const void* g, void* p, void* m, int dt, int64_t n, float grad_norm, float lr, float beta1, float beta2,
                  float eps, int step, int bias_correction, float decay, int grad_averaging, int mode) {
    float bc1 = 1.0f, bc2 = 1.0f;
    if (bias_correction == 1) {
        bc1 = (float)(1.0 - pow((double)beta1, (double)step));
        bc2 = (float)sqrt(1.0 - pow((double)beta2, (double)step));
    }
    (void)beta2;
    float beta3 = 1; if (grad_averaging == 1) beta3 = 1 - beta1;
    const float denom = grad_norm / bc2 + eps;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float rg = ld(g, dt, i), rp = ld(p, dt, i), rm = ld(m, dt, i);
        if (mode == 0) {
            float q = rg / denom, t = decay * rp; rg = q + t;
            float a = beta1 * rm, b = beta3 * rg; rm = a + b;
            float mu = rm / bc1; float lm = lr * mu;
            rp = rp - lm;
        } else {
            float a = beta1 * rm, b = beta3 * rg; rm = a + b;
            float mu = rm / bc1; float q = mu / denom, t = decay * rp; float u = q + t; float lu = lr * u;
            rp = rp - lu;
        }
        st(p, dt, i, rp); st(m, dt, i, rm);
    }
}

/* LARSFunctor: csrc/multi_tensor_lars.cu:84-135.  model_out (fp16, may be NULL) is cast from the stored weight. */
void orc_lars(const void* g, int g_dt, void* w, int w_dt, void* mom, void* model_out, int64_t n, float grad_norm,
              float param_norm, float lr, float trust_coefficient, float eps, float weight_decay, float momentum,
              int nesterov, int is_skipped, const int* noop) {
    if (noop && *noop) return;
    float scaled_lr = lr;
    if (!is_skipped) {
        float trust = 1.0f;
        if (grad_norm > 0.0f && param_norm > 0.0f) {
            float num = trust_coefficient * param_norm; float pw = param_norm * weight_decay;
            float den = grad_norm + pw; den = den + eps; trust = num / den;
        }
        scaled_lr = lr * trust;
    }
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float ig = ld(g, g_dt, i), iw = ld(w, w_dt, i), im = ld(mom, w_dt, i);
        float t = weight_decay * iw; ig = ig + t;
        float a = im * momentum, b = scaled_lr * ig; im = a - b;
        if (nesterov) { float a2 = im * momentum, b2 = scaled_lr * ig; float d = a2 - b2; iw = iw + d; }
        else iw = iw + im;
        st(w, w_dt, i, iw);
        if (model_out) st(model_out, DT_F16, i, ld(w, w_dt, i));
        st(mom, w_dt, i, im);
    }
}

/* ---------------------------------------------------------------- legacy split LAMB (math carried out in T) */

/* LAMBStage1Functor<GRAD_T,T,UPD_T>: csrc/multi_tensor_lamb_stage_1.cu:81-100; host part :126-130
 * (clip factor read on the host, bias corrections with step+1 via std::pow(float,float)). */
void orc_lamb_stage1_legacy(const void* g, int g_dt, const void* p, void* m, void* v, int t_dt, void* upd, int u_dt,
                            int64_t n, float decay, int step, float beta1, float beta2, float eps,
                            float global_grad_norm, float max_global_grad_norm) {
    const float clip = global_grad_norm > max_global_grad_norm ? global_grad_norm / max_global_grad_norm : 1.0f;
    const float next_step = (float)(step + 1);
    const float bc1 = 1.0f - powf(beta1, next_step);
    const float bc2 = 1.0f - powf(beta2, next_step);
    const float omb1 = 1 - beta1, omb2 = 1 - beta2;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float rg = ld(g, g_dt, i), rp = ld(p, t_dt, i), rm = ld(m, t_dt, i), rv = ld(v, t_dt, i);
        /* every intermediate of type T is rounded to T (identity for T = float) */
        float sg = rnd(rg / clip, t_dt);
        float t1 = rm * beta1, t2 = omb1 * sg;
        rm = rnd(t1 + t2, t_dt);
        float t4 = omb2 * sg; t4 = t4 * sg;
        float t3 = rv * beta2;
        rv = rnd(t3 + t4, t_dt);
        float mu = rnd(rm / bc1, t_dt), vu = rnd(rv / bc2, t_dt);
        float denom = rnd(sqrtf(vu) + eps, t_dt);
        float q = rnd(mu / denom, t_dt);            /* T / T returns T (Half.h:513) */
        float dp = decay * rp;
        float u = rnd(q + dp, t_dt);
        st(upd, u_dt, i, u); st(m, t_dt, i, rm); st(v, t_dt, i, rv);
    }
}

/* LAMBStage2Functor<T,UPD_T>: csrc/multi_tensor_lamb_stage_2.cu:41-47,77. */
void orc_lamb_stage2_legacy(void* p, int t_dt, const void* upd, int u_dt, int64_t n, float lr, float decay,
                            int use_nvlamb, float param_norm, float update_norm) {
    float ratio = lr;
    if (use_nvlamb || decay != 0.0f)
        ratio = (update_norm != 0.0f && param_norm != 0.0f) ? lr * (param_norm / update_norm) : lr;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float ru = ratio * rnd(ld(upd, u_dt, i), t_dt);
        st(p, t_dt, i, ld(p, t_dt, i) - ru);
    }
}

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}
