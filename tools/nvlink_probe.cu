// tools/nvlink_probe.cu -- micro-benchmark kernels that separate the two NVLink directions of the fused
// reduce-scatter + Adam + all-gather kernel (DistAdamP2POp, csrc/mt_ops.cuh): in-switch reduction loads only,
// multicast stores only, both in one kernel, and the same three with plain peer-to-peer accesses.
// Built by tools/nvlink_probe.py (nvcc -shared) and driven from there over symmetric memory.  Measurement tool only.
#include <cuda_runtime.h>
#include <stdint.h>

namespace {

__device__ __forceinline__ void mc_ld_reduce(const void* p, uint32_t (&w)[4]) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p) : "memory");
}
__device__ __forceinline__ void mc_st(void* p, const uint32_t (&w)[4]) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                 "r"(w[3]) : "memory");
}
__device__ __forceinline__ void ld16(const void* p, uint32_t (&w)[4]) {
    asm volatile("ld.global.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p) : "memory");
}
__device__ __forceinline__ void st16(void* p, const uint32_t (&w)[4]) {
    asm volatile("st.global.L1::no_allocate.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3])
                 : "memory");
}

struct Peers {
    const char* g[8];
    char* p[8];
};

// mode bit 0: load side on, bit 1: store side on; kMc: through the switch or peer-to-peer.  `off`/`bytes`: this rank's
// shard of the symmetric buffers.  U 16-byte accesses in flight per thread.
template <bool kMc, int WORLD, int U>
__global__ void __launch_bounds__(256) probe_kernel(Peers peers, const char* gmc, char* pmc, size_t off, size_t bytes,
                                                    int mode, uint32_t* sink) {
    const size_t nvec = bytes / 16;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    uint32_t acc = 0;
    for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * U) {
        uint32_t w[U][4];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const size_t i = i0 + u * stride;
            w[u][0] = (uint32_t)i; w[u][1] = w[u][2] = w[u][3] = 0x3f803f80u;
            if ((mode & 1) && i < nvec) {
                if (kMc) mc_ld_reduce(gmc + off + i * 16, w[u]);
                else {
                    uint32_t t[WORLD][4];
#pragma unroll
                    for (int r = 0; r < WORLD; r++) ld16(peers.g[r] + off + i * 16, t[r]);
#pragma unroll
                    for (int r = 0; r < WORLD; r++)
#pragma unroll
                        for (int k = 0; k < 4; k++) w[u][k] = (r ? w[u][k] : 0u) ^ t[r][k];
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const size_t i = i0 + u * stride;
            acc ^= w[u][0] ^ w[u][1] ^ w[u][2] ^ w[u][3];
            if ((mode & 2) && i < nvec) {
                if (kMc) mc_st(pmc + off + i * 16, w[u]);
                else {
#pragma unroll
                    for (int r = 0; r < WORLD; r++) st16(peers.p[r] + off + i * 16, w[u]);
                }
            }
        }
    }
    if (acc == 0x12345678u) *sink = acc;  // keeps the loads alive
}

template <bool kMc, int WORLD>
int launch(const Peers& peers, const void* gmc, void* pmc, size_t off, size_t bytes, int mode, int ctas, uint32_t* sink,
           cudaStream_t s) {
    probe_kernel<kMc, WORLD, 4><<<ctas, 256, 0, s>>>(peers, (const char*)gmc, (char*)pmc, off, bytes, mode, sink);
    return (int)cudaGetLastError();
}

}  // namespace

extern "C" int nvlink_probe(int world, int multicast, const void* const* grad_bases, void* const* param_bases,
                            const void* gmc, void* pmc, size_t off, size_t bytes, int mode, int ctas, void* sink,
                            void* stream) {
    Peers peers{};
    for (int r = 0; r < world && r < 8; r++) {
        peers.g[r] = (const char*)grad_bases[r];
        peers.p[r] = (char*)param_bases[r];
    }
    cudaStream_t s = (cudaStream_t)stream;
    uint32_t* sk = (uint32_t*)sink;
#define GO(MC, W) return launch<MC, W>(peers, gmc, pmc, off, bytes, mode, ctas, sk, s)
    if (multicast) {
        if (world == 2) GO(true, 2);
        if (world == 4) GO(true, 4);
        if (world == 8) GO(true, 8);
    } else {
        if (world == 2) GO(false, 2);
        if (world == 4) GO(false, 4);
        if (world == 8) GO(false, 8);
    }
#undef GO
    return -1;
}
