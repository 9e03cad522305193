// tools/tune_adam.cu -- on-GPU design-space sweep for the mixed-precision Adam update kernel (28 B/param).
// Not part of the product: it instantiates the product's kernel template with different launch shapes and an
// experimental cp.async.bulk (TMA) staged variant, times each with CUDA events and prints achieved GB/s.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o tools/tune_adam tools/tune_adam.cu
//   ./tools/tune_adam [nelems]
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <vector>

#include "../beforeholiday_b200/csrc/mt_ops.cuh"

using namespace mt;
using bf16 = __nv_bfloat16;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__global__ void init_kernel(float* p, float* m, float* v, bf16* g, size_t n) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += st) {
        unsigned h = (unsigned)(i * 2654435761u) ^ (unsigned)(i >> 32);
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        float r = (float)(h & 0xffffff) / 16777216.f - 0.5f;
        p[i] = r * 0.04f; m[i] = r * 1e-3f; v[i] = r * r * 1e-6f; g[i] = __float2bfloat16(r * 2e-3f);
    }
}
__global__ void checksum_kernel(const float* p, const bf16* pc, size_t n, double* out) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
    double s = 0;
    for (; i < n; i += st) s += (double)p[i] * 3.0 + (double)__bfloat162float(pc[i]);
    atomicAdd(out, s);
}

// ---------------------------------------------------------------------------------------------- TMA staged variant
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n"
        ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

struct AdamScalars { float b1, b2, omb1, omb2, bc1, bc2, eps, decay, lr, grad_scale; };

__device__ __forceinline__ void adam_elem(const AdamScalars& a, float& w, float& m, float& v, float g) {
    const float sg = fdiv(g, a.grad_scale);
    m = fadd(fmul(a.b1, m), fmul(a.omb1, sg));
    v = fadd(fmul(a.b2, v), fmul(fmul(a.omb2, sg), sg));
    const float mu = fdiv(m, a.bc1), vu = fdiv(v, a.bc2);
    const float denom = fadd(fsqrt(vu), a.eps);
    const float upd = fadd(fdiv(mu, denom), fmul(a.decay, w));
    w = fsub(w, fmul(a.lr, upd));
}

// One CTA: warp 0 lane 0 = producer (bulk loads), CONSUMERS threads compute, consumer thread 0 issues bulk stores.
// Stage layout: [p E f32][m E f32][v E f32][g E bf16]; p/m/v updated in place, the bf16 copy overwrites g's slot.
template <int E, int S, int CONSUMERS>
__global__ void __launch_bounds__(CONSUMERS + 32, 1)
adam_tma_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, const bf16* __restrict__ g,
                bf16* __restrict__ pc, size_t nunits, AdamScalars a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int STAGE_BYTES = E * 14;
    __shared__ uint64_t full[S], empty[S];
    unsigned char* stages = smem_raw;
    const int tid = threadIdx.x;
    if (tid == 0) {
        for (int s = 0; s < S; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const size_t first = blockIdx.x, stride = gridDim.x;
    const size_t my_units = first < nunits ? (nunits - first + stride - 1) / stride : 0;

    if (tid < 32) {  // producer warp
        if (tid == 0) {
            for (size_t k = 0; k < my_units; k++) {
                const int s = (int)(k % S);
                const uint32_t ph = (uint32_t)((k / S) & 1);
                mbar_wait(&empty[s], ph ^ 1);  // first pass over the ring falls through
                const size_t off = (first + k * stride) * (size_t)E;
                unsigned char* st = stages + (size_t)s * STAGE_BYTES;
                mbar_expect_tx(&full[s], STAGE_BYTES);
                bulk_g2s(st, p + off, E * 4, &full[s]);
                bulk_g2s(st + E * 4, m + off, E * 4, &full[s]);
                bulk_g2s(st + E * 8, v + off, E * 4, &full[s]);
                bulk_g2s(st + E * 12, g + off, E * 2, &full[s]);
            }
        }
        return;
    }
    const int ct = tid - 32;  // consumer thread id
    constexpr int PER = E / CONSUMERS;  // elements per consumer thread per unit (4 or 8)
    static_assert(PER == 4 || PER == 8, "E / CONSUMERS must be 4 or 8");
    for (size_t k = 0; k < my_units; k++) {
        const int s = (int)(k % S);
        const uint32_t ph = (uint32_t)((k / S) & 1);
        unsigned char* st = stages + (size_t)s * STAGE_BYTES;
        mbar_wait(&full[s], ph);
        float* sp = reinterpret_cast<float*>(st) + ct * PER;
        float* sm = reinterpret_cast<float*>(st + E * 4) + ct * PER;
        float* sv = reinterpret_cast<float*>(st + E * 8) + ct * PER;
        bf16* sg = reinterpret_cast<bf16*>(st + E * 12) + ct * PER;
        float w[PER], mm[PER], vv[PER], gg[PER];
#pragma unroll
        for (int q = 0; q < PER / 4; q++) {
            float4 t;
            t = *reinterpret_cast<float4*>(sp + 4 * q); w[4 * q] = t.x; w[4 * q + 1] = t.y; w[4 * q + 2] = t.z; w[4 * q + 3] = t.w;
            t = *reinterpret_cast<float4*>(sm + 4 * q); mm[4 * q] = t.x; mm[4 * q + 1] = t.y; mm[4 * q + 2] = t.z; mm[4 * q + 3] = t.w;
            t = *reinterpret_cast<float4*>(sv + 4 * q); vv[4 * q] = t.x; vv[4 * q + 1] = t.y; vv[4 * q + 2] = t.z; vv[4 * q + 3] = t.w;
            uint2 gb = *reinterpret_cast<uint2*>(sg + 4 * q);
            gg[4 * q] = __uint_as_float(gb.x << 16); gg[4 * q + 1] = __uint_as_float(gb.x & 0xffff0000u);
            gg[4 * q + 2] = __uint_as_float(gb.y << 16); gg[4 * q + 3] = __uint_as_float(gb.y & 0xffff0000u);
        }
#pragma unroll
        for (int i = 0; i < PER; i++) adam_elem(a, w[i], mm[i], vv[i], gg[i]);
#pragma unroll
        for (int q = 0; q < PER / 4; q++) {
            *reinterpret_cast<float4*>(sp + 4 * q) = make_float4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
            *reinterpret_cast<float4*>(sm + 4 * q) = make_float4(mm[4 * q], mm[4 * q + 1], mm[4 * q + 2], mm[4 * q + 3]);
            *reinterpret_cast<float4*>(sv + 4 * q) = make_float4(vv[4 * q], vv[4 * q + 1], vv[4 * q + 2], vv[4 * q + 3]);
            __nv_bfloat162 lo = __floats2bfloat162_rn(w[4 * q], w[4 * q + 1]), hi = __floats2bfloat162_rn(w[4 * q + 2], w[4 * q + 3]);
            uint2 ob; ob.x = *reinterpret_cast<uint32_t*>(&lo); ob.y = *reinterpret_cast<uint32_t*>(&hi);
            *reinterpret_cast<uint2*>(sg + 4 * q) = ob;
        }
        fence_async_smem();                                         // generic-proxy writes -> visible to the bulk engine
        asm volatile("bar.sync 1, %0;" ::"n"(CONSUMERS) : "memory");  // all consumers done with this stage
        if (ct == 0) {
            const size_t off = (first + k * stride) * (size_t)E;
            bulk_s2g(p + off, st, E * 4);
            bulk_s2g(m + off, st + E * 4, E * 4);
            bulk_s2g(v + off, st + E * 8, E * 4);
            bulk_s2g(pc + off, st + E * 12, E * 2);
            bulk_commit();
            // free the PREVIOUS stage once its stores have finished reading shared memory (this unit's may still be pending)
            if (k > 0) { bulk_wait_read<1>(); mbar_arrive(&empty[(int)((k - 1) % S)]); }
        }
    }
    if (ct == 0) { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
}

// ---------------------------------------------------------------------------------------------- harness
struct Bufs { float *p, *m, *v; bf16 *g, *pc; size_t n; };

template <class F> float time_ms(F f, int warm = 2, int iters = 5) {
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    for (int i = 0; i < warm; i++) f();
    CK(cudaEventRecord(a));
    for (int i = 0; i < iters; i++) f();
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    CK(cudaGetLastError());
    return ms / iters;
}

double checksum(const Bufs& B) {
    double* d; CK(cudaMalloc(&d, 8)); CK(cudaMemset(d, 0, 8));
    checksum_kernel<<<1184, 256>>>(B.p, B.pc, B.n, d);
    double h; CK(cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost)); cudaFree(d);
    return h;
}
void reinit(const Bufs& B) { init_kernel<<<1184, 512>>>(B.p, B.m, B.v, B.g, B.n); CK(cudaDeviceSynchronize()); }

struct Table { DevTable tb; void* blob; };
Table make_table(const Bufs& B, int tile) {
    // [ptrs 5][numel 1][tile_begin 2]
    std::vector<unsigned char> h(5 * 8 + 8 + 8);
    void* ptrs[5] = {B.p, B.m, B.v, B.g, B.pc};
    memcpy(h.data(), ptrs, 40);
    int64_t n = (int64_t)B.n; memcpy(h.data() + 40, &n, 8);
    uint32_t tbv[2] = {0, (uint32_t)((B.n + tile - 1) / tile)}; memcpy(h.data() + 48, tbv, 8);
    Table t; CK(cudaMalloc(&t.blob, h.size())); CK(cudaMemcpy(t.blob, h.data(), h.size(), cudaMemcpyHostToDevice));
    unsigned char* d = (unsigned char*)t.blob;
    t.tb.ptrs = (const void* const*)d; t.tb.numel = (const int64_t*)(d + 40); t.tb.tile_begin = (const uint32_t*)(d + 48);
    t.tb.ntensors = 1; t.tb.ntiles = tbv[1]; t.tb.tile_elems = tile;
    return t;
}

struct HP { float *b1, *b2, *eps, *wd; int32_t* bc; };

template <int G, int U, int T, int MB>
void run_lib(const char* name, const Bufs& B, const HP& hp, int32_t* noop, int tile, bool persistent, int sms, double ref_sum) {
    using Op = DistAdamOp<bf16, true, true>;
    Table t = make_table(B, tile);
    typename Op::Params prm{hp.b1, hp.b2, hp.eps, hp.wd, hp.bc, 1e-4f, 1.0f, 3, 1};
    auto k = mt_table_kernel<Op, G, U, T, MB>;
    int occ = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, T, 0));
    unsigned grid = persistent ? (unsigned)(occ * sms) : t.tb.ntiles;
    if (grid > t.tb.ntiles) grid = t.tb.ntiles;
    reinit(B);
    k<<<grid, T>>>(t.tb, prm, noop); CK(cudaDeviceSynchronize());
    double cs = checksum(B);
    float ms = time_ms([&] { k<<<grid, T>>>(t.tb, prm, noop); });
    printf("%-44s tile=%6d grid=%7u occ=%d  %8.3f ms  %7.1f GB/s  %s\n", name, tile, grid, occ, ms, 28.0 * B.n / ms / 1e6,
           (fabs(cs - ref_sum) <= 1e-9 * fabs(ref_sum)) ? "ok" : "CHECKSUM-DIFF");
    fflush(stdout);
    cudaFree(t.blob);
}

template <int E, int S, int C>
void run_tma(const char* name, const Bufs& B, int ctas_per_sm, int sms, double ref_sum) {
    AdamScalars a;
    a.b1 = 0.9f; a.b2 = 0.95f; a.omb1 = 1.f - a.b1; a.omb2 = 1.f - a.b2;
    a.bc1 = (float)(1.0 - pow((double)a.b1, 3.0)); a.bc2 = (float)(1.0 - pow((double)a.b2, 3.0));
    a.eps = 1e-8f; a.decay = 0.1f; a.lr = 1e-4f; a.grad_scale = 1.0f;
    auto k = adam_tma_kernel<E, S, C>;
    size_t smem = (size_t)E * 14 * S;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    size_t nunits = B.n / E;
    unsigned grid = (unsigned)(ctas_per_sm * sms);
    reinit(B);
    k<<<grid, C + 32, smem>>>(B.p, B.m, B.v, B.g, B.pc, nunits, a); CK(cudaDeviceSynchronize());
    double cs = checksum(B);
    float ms = time_ms([&] { k<<<grid, C + 32, smem>>>(B.p, B.m, B.v, B.g, B.pc, nunits, a); });
    printf("%-44s E=%5d S=%2d C=%4d grid=%5u smem=%6zu  %8.3f ms  %7.1f GB/s  %s\n", name, E, S, C, grid, smem, ms,
           28.0 * B.n / ms / 1e6, (fabs(cs - ref_sum) <= 1e-9 * fabs(ref_sum)) ? "ok" : "CHECKSUM-DIFF");
    fflush(stdout);
}

int main(int argc, char** argv) {
    size_t n = argc > 1 ? (size_t)atoll(argv[1]) : (size_t)1536 * 1024 * 1024;
    n = n / (1 << 16) * (1 << 16);
    int dev = 0; CK(cudaSetDevice(dev));
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    printf("n=%zu (%.1f GB resident, %.1f GB traffic/step), SMs=%d\n", n, n * 16 / 1e9, n * 28 / 1e9, sms);
    Bufs B; B.n = n;
    CK(cudaMalloc(&B.p, n * 4)); CK(cudaMalloc(&B.m, n * 4)); CK(cudaMalloc(&B.v, n * 4)); CK(cudaMalloc(&B.g, n * 2)); CK(cudaMalloc(&B.pc, n * 2));
    HP hp; float h1 = 0.9f, h2 = 0.95f, he = 1e-8f, hw = 0.1f; int32_t hb = 1;
    CK(cudaMalloc(&hp.b1, 4)); CK(cudaMalloc(&hp.b2, 4)); CK(cudaMalloc(&hp.eps, 4)); CK(cudaMalloc(&hp.wd, 4)); CK(cudaMalloc(&hp.bc, 4));
    CK(cudaMemcpy(hp.b1, &h1, 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(hp.b2, &h2, 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(hp.eps, &he, 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(hp.wd, &hw, 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(hp.bc, &hb, 4, cudaMemcpyHostToDevice));
    int32_t* noop; CK(cudaMalloc(&noop, 4)); CK(cudaMemset(noop, 0, 4));

    // copy ceiling on this box, same definition as MEASURED_PEAKS.json (read + write bytes)
    {
        float ms = time_ms([&] { CK(cudaMemcpyAsync(B.m, B.p, n * 4, cudaMemcpyDeviceToDevice)); });
        printf("%-44s %8.3f ms  %7.1f GB/s (read+write)\n", "cudaMemcpy D2D fp32", ms, 8.0 * n / ms / 1e6);
    }
    // reference result checksum from the product's default shape
    double ref_sum;
    {
        using Op = DistAdamOp<bf16, true, true>;
        Table t = make_table(B, 16384);
        typename Op::Params prm{hp.b1, hp.b2, hp.eps, hp.wd, hp.bc, 1e-4f, 1.0f, 3, 1};
        reinit(B);
        mt_table_kernel<Op, 8, 1, 512, 2><<<296, 512>>>(t.tb, prm, noop); CK(cudaDeviceSynchronize());
        ref_sum = checksum(B);
        cudaFree(t.blob);
    }
    if (argc > 2) {  // quick mode (cache-hint sweep): only the product's launch shape, three times
        for (int rep = 0; rep < 3; rep++)
            run_lib<8, 1, 256, 4>("T256B4 ld" MT_LD_HINT "|" MT_LD_HINT256 " st" MT_ST_HINT "|" MT_ST_HINT256, B, hp, noop, 16384, false, sms, ref_sum);
        return 0;
    }
    run_lib<8, 1, 512, 2>("lib G8 U1 T512 B2 persistent", B, hp, noop, 16384, true, sms, ref_sum);
    for (int tile : {4096, 6144, 8192, 12288, 16384, 32768})
        run_lib<8, 1, 256, 4>("lib G8 U1 T256 B4 grid=ntiles", B, hp, noop, tile, false, sms, ref_sum);
    for (int tile : {2048, 4096, 8192})
        run_lib<8, 1, 128, 8>("lib G8 U1 T128 B8 grid=ntiles", B, hp, noop, tile, false, sms, ref_sum);
    for (int tile : {8192, 16384})
        run_lib<8, 1, 256, 5>("lib G8 U1 T256 B5 grid=ntiles", B, hp, noop, tile, false, sms, ref_sum);
    for (int tile : {8192, 16384})
        run_lib<8, 1, 256, 6>("lib G8 U1 T256 B6 grid=ntiles", B, hp, noop, tile, false, sms, ref_sum);
    for (int tile : {8192, 16384, 32768})
        run_lib<8, 2, 256, 3>("lib G8 U2 T256 B3 grid=ntiles", B, hp, noop, tile, false, sms, ref_sum);
    for (int tile : {8192, 16384, 32768})
        run_lib<8, 2, 256, 2>("lib G8 U2 T256 B2 grid=ntiles", B, hp, noop, tile, false, sms, ref_sum);
    for (int tile : {8192, 16384})
        run_lib<8, 1, 512, 2>("lib G8 U1 T512 B2 grid=ntiles", B, hp, noop, tile, false, sms, ref_sum);
    for (int tile : {8192, 16384})
        run_lib<8, 1, 384, 3>("lib G8 U1 T384 B3 grid=ntiles", B, hp, noop, tile, false, sms, ref_sum);
    for (int tile : {4096, 8192})
        run_lib<4, 2, 256, 4>("lib G4 U2 T256 B4 grid=ntiles", B, hp, noop, tile, false, sms, ref_sum);
    for (int tile : {4096, 8192})
        run_lib<4, 1, 256, 6>("lib G4 U1 T256 B6 grid=ntiles", B, hp, noop, tile, false, sms, ref_sum);
    for (int tile : {4096, 8192})
        run_lib<4, 1, 256, 8>("lib G4 U1 T256 B8 grid=ntiles", B, hp, noop, tile, false, sms, ref_sum);
    run_tma<1024, 6, 256>("tma (2 CTA/SM)", B, 2, sms, ref_sum);
    run_tma<2048, 4, 512>("tma", B, 1, sms, ref_sum);
    return 0;
}
