/*
 * mt_b200.h -- C ABI of libmt_b200.so: the B200 (sm_100a) multi-tensor-apply optimizer path.
 *
 * This is the drop-in boundary for the reference's `amp_C` pybind module
 * (/root/reference/csrc/amp_C_frontend.cpp:165-194).  Every entry point below replaces one
 * `*_cuda()` host function of the reference and keeps its argument meaning; torch types are
 * replaced by plain pointers and sizes (MtTensorLists).  The pybind module `amp_C` shipped in
 * beforeholiday_b200/csrc/amp_C_module.cpp is a thin adaptor over this header; INTEGRATION.md
 * shows the binding a maintainer of the reference would write.
 *
 * Conventions
 *   - All pointers inside MtTensorLists, `noop_flag`, norm outputs and scratch are DEVICE pointers on
 *     `tl->device`.  Tensors are dense (any memory format: the op runs over physical order, as
 *     /root/reference/csrc/multi_tensor_apply.cuh:64-68 allows).
 *   - Calls enqueue work on `stream` (a cudaStream_t) and return; no synchronisation.
 *   - Return value: 0 on success, a negative MT_E* code otherwise; mt_last_error() returns a
 *     thread-local message (the pybind layer raises RuntimeError with it, mirroring TORCH_CHECK).
 *   - `chunk_size` is accepted for signature compatibility.  It only has semantic meaning for the
 *     norm ops, where the overflow flag is defined per chunk (see mt_multi_tensor_l2norm).
 *   - dtype tags: MT_F32, MT_F16, MT_BF16, MT_F64 (and MT_E5M2 where an entry point says so).
 *   - There is no CPU fallback: on a machine without a CUDA device every compute entry point
 *     returns MT_ECUDA.
 */
#ifndef MT_B200_H_
#define MT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MT_MAX_DEPTH 5

enum { MT_F32 = 0, MT_F16 = 1, MT_BF16 = 2, MT_F64 = 3, MT_E5M2 = 4 };
/* MT_E5M2: one byte per element, the high byte of an fp16 value rounded as the reference's e5m2_allgather packing does
 * (apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb_kernel.cu:34-66).  Accepted only as the copy list of
 * mt_multi_tensor_lamb_update_weights and as the input list of mt_multi_tensor_maybe_cast. */

enum {
    MT_OK = 0,
    MT_EINVAL = -1,   /* bad argument (depth, dtype combination, null pointer, size mismatch) */
    MT_ECUDA = -2,    /* a CUDA runtime call failed (message has the cudaError string) */
    MT_ECAPTURE = -3, /* not possible on a capturing stream: a device-table cache miss (run one eager warm-up call first),
                         or an entry point that cannot be recorded at all (mt_multi_tensor_l2norm_noalloc) */
    MT_ESCRATCH = -4  /* scratch buffer missing or too small */
};

/* `depth` lists of `ntensors` tensors; replaces std::vector<std::vector<at::Tensor>>
 * (/root/reference/csrc/multi_tensor_apply.cuh:44-72).  ptrs is list-major: ptrs[d * ntensors + t].
 * numel[t] is shared by all lists (the reference checks numel equality, :70).  Empty tensors are legal. */
typedef struct MtTensorLists {
    int32_t depth;
    int32_t device;              /* CUDA device ordinal that owns every pointer */
    int64_t ntensors;
    const void* const* ptrs;
    const int64_t* numel;
    int32_t dtype[MT_MAX_DEPTH]; /* per-list dtype tag */
} MtTensorLists;

const char* mt_last_error(void);
const char* mt_version(void);

/* out = float(in) * scale; *noop_flag = 1 iff any input element is non-finite.
 * Lists: [in, out], any of f32/f16/bf16 each.
 * Replaces multi_tensor_scale_cuda, /root/reference/csrc/multi_tensor_scale_kernel.cu:113-136. */
int mt_multi_tensor_scale(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float scale, void* stream);

/* out = a*float(x) + b*float(y); finite check on x (arg_to_check 0), y (1) or both (-1).
 * Lists: [x, y, out], any of f32/f16/bf16 each.
 * Replaces multi_tensor_axpby_cuda, /root/reference/csrc/multi_tensor_axpby_kernel.cu:128-157. */
int mt_multi_tensor_axpby(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float a, float b,
                          int arg_to_check, void* stream);

/* Bytes of scratch the norm / LAMB entry points need for these lists and this chunk size. */
size_t mt_norm_scratch_bytes(int chunk_size, const MtTensorLists* tl);
size_t mt_lamb_scratch_bytes(int chunk_size, const MtTensorLists* tl);

/* ret[0] = ||all tensors||_2 ; ret_per_tensor[t] = ||tensor t||_2 (when per_tensor != 0).
 * *noop_flag = 1 iff the sum of squares of any `chunk_size`-element chunk of any tensor is non-finite
 * (the reference's definition, csrc/multi_tensor_l2norm_kernel.cu:103-104).  Lists: [x], f32/f16/bf16.
 * scratch: mt_norm_scratch_bytes() bytes of device memory, contents irrelevant on entry.
 * Replaces multi_tensor_l2norm_cuda + cleanup, /root/reference/csrc/multi_tensor_l2norm_kernel.cu:198-237,301-364. */
int mt_multi_tensor_l2norm(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, int per_tensor,
                           float* ret, float* ret_per_tensor, void* scratch, size_t scratch_bytes, void* stream);

/* As mt_multi_tensor_l2norm over list 0 and additionally out = float(in) * scale into list 1.
 * Lists: [in, out], f32/f16 in the reference; bf16 is accepted too.
 * Replaces multi_tensor_l2norm_scale_cuda, /root/reference/csrc/multi_tensor_l2norm_scale_kernel.cu:258-324. */
int mt_multi_tensor_l2norm_scale(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float scale,
                                 int per_tensor, float* ret, float* ret_per_tensor, void* scratch,
                                 size_t scratch_bytes, void* stream);

/* Adam (mode 0, L2) / AdamW (mode 1).  Lists: [g, p, m, v], one dtype for all four (f32/f16/bf16/f64),
 * fp32 math; bias corrections evaluated on the host in double and rounded to float.  The noop flag is
 * ignored, as in the reference.
 * Replaces multi_tensor_adam_cuda, /root/reference/csrc/multi_tensor_adam.cu:129-171. */
int mt_multi_tensor_adam(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float lr, float beta1,
                         float beta2, float epsilon, int step, int mode, int bias_correction, float weight_decay,
                         void* stream);

/* SGD with momentum / nesterov / dampening / weight decay, fused gradient scale, optional fp16 model copy.
 * Lists: [g, w, mom] or [g, w, mom, w_fp16]; dtype combinations as the reference's six cases.
 * Does nothing when *noop_flag != 0 (device-side test).
 * Replaces multi_tensor_sgd_cuda, /root/reference/csrc/multi_tensor_sgd_kernel.cu:141-322. */
int mt_multi_tensor_sgd(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float wd, float momentum,
                        float dampening, float lr, int nesterov, int first_run, int wd_after_momentum, float scale,
                        void* stream);

/* LAMB.  Lists: [g, p, m, v], one dtype (f32/f16/bf16).  global_grad_norm is a DEVICE float[1]; the trust
 * ratios are computed on the device.  g is overwritten with the update, as in the reference.
 * scratch: mt_lamb_scratch_bytes() bytes.
 * Replaces multi_tensor_lamb_cuda, /root/reference/csrc/multi_tensor_lamb.cu:332-413. */
int mt_multi_tensor_lamb(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float lr, float beta1,
                         float beta2, float epsilon, int step, int bias_correction, float weight_decay,
                         int grad_averaging, int mode, const float* global_grad_norm, float max_grad_norm,
                         int use_nvlamb, void* scratch, size_t scratch_bytes, void* stream);

/* "_mp" (mixed-precision, capturable) variants: the l2norm that does nothing once the flag is set, and LAMB with the
 * learning rate, step, max grad norm and inverse loss scale read from DEVICE memory.  lamb_mp lists: [g, p, m, v] (one
 * dtype) or [g, p32, m32, v32, p_model] (fp32 masters + a model copy in g's dtype).  g is multiplied by *inv_scale,
 * overwritten with the update.  All kernels return at once when *noop_flag != 0.  scratch: mt_lamb_scratch_bytes().
 * Replace multi_tensor_l2norm_mp_cuda (/root/reference/csrc/multi_tensor_l2norm_kernel_mp.cu:153-216) and
 * multi_tensor_lamb_mp_cuda (/root/reference/csrc/multi_tensor_lamb_mp.cu:367-496). */
int mt_multi_tensor_l2norm_mp(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, int per_tensor, float* ret,
                              float* ret_per_tensor, void* scratch, size_t scratch_bytes, void* stream);
int mt_multi_tensor_lamb_mp(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, const float* lr, float beta1,
                            float beta2, float epsilon, const int32_t* step, int bias_correction, float weight_decay,
                            int grad_averaging, int mode, const float* global_grad_norm, const float* max_grad_norm,
                            int use_nvlamb, const float* found_inf, const float* inv_scale, void* scratch,
                            size_t scratch_bytes, void* stream);

/* Adagrad (mode 0: L2, mode 1: decoupled weight decay).  Lists: [g, p, h], one dtype (f32/f16/bf16/f64), fp32 math.
 * Replaces multi_tensor_adagrad_cuda, /root/reference/csrc/multi_tensor_adagrad.cu:84-100. */
int mt_multi_tensor_adagrad(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float lr, float epsilon,
                            int mode, float weight_decay, void* stream);

/* NovoGrad.  Lists: [g, p, m], one dtype.  grad_norms: DEVICE float[ntensors], the running per-tensor second moment;
 * it is first blended with this step's gradient norm (norm_type 0: L-inf, a*old + b*max|g|; otherwise L2,
 * sqrt(a*old^2 + b*sum g^2), a = beta2, b = 1 - beta2), then used by the update.  scratch: mt_norm_scratch_bytes().
 * Replaces multi_tensor_novograd_cuda + multi_tensor_norm_out_cuda + cleanup_v2,
 * /root/reference/csrc/multi_tensor_novograd.cu:127-188, multi_tensor_l2norm_kernel.cu:239-299,367-456. */
int mt_multi_tensor_novograd(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, float* grad_norms, float lr,
                             float beta1, float beta2, float epsilon, int step, int bias_correction, float weight_decay,
                             int grad_averaging, int moment_mode, int norm_type, void* scratch, size_t scratch_bytes,
                             void* stream);

/* LARS.  Lists: [g, w, mom] or [g, w, mom, w_fp16]; grad_norms / param_norms: DEVICE float[ntensors].  Does nothing when
 * *noop_flag != 0.  dampening, first_run, wd_after_momentum and scale are accepted and ignored, as in the reference.
 * Replaces multi_tensor_lars_cuda, /root/reference/csrc/multi_tensor_lars.cu:141-354. */
int mt_multi_tensor_lars(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, const float* grad_norms,
                         const float* param_norms, float lr, float trust_coefficient, float epsilon, float weight_decay,
                         float momentum, float dampening, int nesterov, int first_run, int wd_after_momentum, float scale,
                         int is_skipped, void* stream);

/* Converting copy out = (Tout)in that does nothing when *noop_flag != 0 and never writes the flag.  Lists: [in, out],
 * each f32/f16/bf16 (the reference's e5m2-in-a-byte packing belongs to its deprecated fused_adam_cuda optimizer and is
 * not carried over).  Replaces fused_adam_cuda.maybe_cast_mt,
 * /root/reference/apex/contrib/csrc/optimizers/fused_adam_cuda_kernel.cu:734-783,939-959. */
int mt_multi_tensor_maybe_cast(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, void* stream);

/* Sharded LAMB, stage 1 (update term).  Lists: [g, p, m, v, u]; g f32/f16/bf16, the rest f32 shard state.  All
 * per_tensor_* arrays are DEVICE arrays of ntensors entries (one per parameter fragment); step, global_scale and
 * global_grad_norm are DEVICE scalars; per_tensor_beta3 is accepted and ignored (the reference recomputes 1 - beta1).
 * Does nothing when *noop_flag != 0.
 * Extension (seg_sq_out != NULL): seg_begin is a DEVICE int32[nseg+1] of fragment indices splitting the fragment list
 * into runs that belong to one model parameter; seg_sq_out (DEVICE float[2][nseg]) receives, per run, the sum of
 * round_to(model_dtype)(p)^2 and the sum of u^2 over this rank's shard -- the inputs of the two per-parameter norms
 * the reference computes with separate multi_tensor_l2norm sweeps (distributed_fused_lamb.py:673-692).  scratch:
 * mt_dist_lamb_scratch_bytes() bytes, only needed with the extension.
 * Replaces multi_tensor_lamb_compute_update_term_cuda,
 * /root/reference/apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb.cpp:3-17, _kernel.cu:119-321,435-474. */
size_t mt_dist_lamb_scratch_bytes(const MtTensorLists* tl);
int mt_multi_tensor_lamb_compute_update_term(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                             const float* per_tensor_beta1, const float* per_tensor_beta2,
                                             const float* per_tensor_beta3, const int32_t* per_tensor_bias_correction,
                                             const int32_t* step, const float* per_tensor_epsilon, int mode,
                                             const float* per_tensor_decay, const float* global_scale,
                                             const float* global_grad_norm, float max_grad_norm,
                                             const int32_t* seg_begin, int nseg, int model_dtype, float* seg_sq_out,
                                             void* scratch, size_t scratch_bytes, void* stream);

/* Sharded LAMB, stage 2 (weight update).  Lists: [u (f32), p (f32), p_copy (f32/f16/bf16)].  per_tensor_param_norm is
 * indexed by fragment, per_tensor_update_norm through update_norm_offset[fragment] (DEVICE int64[ntensors]);
 * learning_rate is a DEVICE scalar; global_grad_norm is accepted and ignored, as in the reference.  Does nothing when
 * *noop_flag != 0.  Replaces multi_tensor_lamb_update_weights_cuda,
 * /root/reference/apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb.cpp:19-29, _kernel.cu:325-433,476-506. */
int mt_multi_tensor_lamb_update_weights(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                        const float* per_tensor_param_norm, const float* per_tensor_update_norm,
                                        const int64_t* update_norm_offset, const float* learning_rate,
                                        const float* per_tensor_decay, const float* global_grad_norm, int use_nvlamb,
                                        void* stream);

/* Legacy split LAMB.  stage1 lists: [g, p, m, v, update]; per_tensor_decay is a DEVICE float[ntensors];
 * clipped_global_grad_norm is passed by value (the reference reads the norm on the host,
 * csrc/multi_tensor_lamb_stage_1.cu:126-127).  Arithmetic is carried out in the parameter dtype.
 * Replaces multi_tensor_lamb_stage1_cuda / stage2_cuda,
 * /root/reference/csrc/multi_tensor_lamb_stage_1.cu:112-151, multi_tensor_lamb_stage_2.cu:94-125. */
int mt_multi_tensor_lamb_stage1(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                const float* per_tensor_decay, int step, float beta1, float beta2, float epsilon,
                                float global_grad_norm, float max_global_grad_norm, void* stream);
int mt_multi_tensor_lamb_stage2(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                const float* per_tensor_param_norm, const float* per_tensor_update_norm, float lr,
                                float weight_decay, int use_nvlamb, void* stream);

/* Mixed-precision (sharded) Adam: fp32 p, m, v; 16- or 32-bit g; optional p_copy in g's dtype.
 * Lists: [p, m, v, g] or [p, m, v, g, p_copy].  Per-tensor hyper-parameters are DEVICE arrays of ntensors.
 * g is divided by grad_scale.  mode 0: eps inside the square root, mode 1: outside.
 * honour_noop != 0: do nothing when *noop_flag != 0 (extension: lets a fused overflow check skip the step
 * without a host round trip; the reference ignores the flag).
 * Replaces multi_tensor_fused_adam_cuda,
 * /root/reference/apex/contrib/csrc/optimizers/multi_tensor_distopt_adam_kernel.cu:169-228. */
int mt_multi_tensor_fused_adam(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                               const float* per_tensor_beta1, const float* per_tensor_beta2,
                               const int32_t* per_tensor_bias_correction, const float* per_tensor_eps,
                               const float* per_tensor_weight_decay, float lr, float grad_scale, int step, int mode,
                               int honour_noop, void* stream);

/* Capturable form of mt_multi_tensor_fused_adam: lr, grad_scale and step are DEVICE scalars read by the kernel, the
 * flag is always honoured.  A step captured in a CUDA graph therefore replays correctly while a host-side schedule
 * updates those scalars (and while the overflow check that precedes it may set the flag).  No counterpart in this
 * fork of the reference (upstream Apex has multi_tensor_adam_capturable); same arithmetic as DistAdamFunctor. */
int mt_multi_tensor_fused_adam_capturable(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                          const float* per_tensor_beta1, const float* per_tensor_beta2,
                                          const int32_t* per_tensor_bias_correction, const float* per_tensor_eps,
                                          const float* per_tensor_weight_decay, const float* lr,
                                          const float* grad_scale, const int32_t* step, int mode, void* stream);

/* Fused gradient reduce-scatter + sharded Adam + parameter all-gather in one kernel over NVLink peer memory.
 * Same arithmetic and lists as mt_multi_tensor_fused_adam with p_copy ([p, m, v, g, p_copy], honour_noop on), but list 3
 * is read from EVERY rank's gradient buffer (summed in rank order, rounded once to the gradient dtype) and list 4 is
 * written to EVERY rank's parameter buffer.  grad_bases[r] / param_bases[r] are rank r's flat buffers as mapped into
 * this process (symmetric allocations of identical layout); tl's lists 3 and 4 point into grad_bases[rank] /
 * param_bases[rank].  world is 2, 4 or 8.  grad_multicast / param_multicast: NVSwitch multicast addresses of the same
 * two buffers (both or neither; NULL = plain peer-to-peer).  With them the 16-byte vector path uses
 * multimem.ld_reduce / multimem.st, i.e. the sum and the broadcast happen inside the switch.  The caller runs a
 * cross-GPU barrier before (all gradients complete) and after (all parameter writes landed) the call.
 * Replaces the reduce_scatter -> multi_tensor_fused_adam -> all_gather sequence of
 * /root/reference/apex/contrib/optimizers/distributed_fused_adam.py:668-742,977-1115. */
int mt_multi_tensor_fused_adam_p2p(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                   const float* per_tensor_beta1, const float* per_tensor_beta2,
                                   const int32_t* per_tensor_bias_correction, const float* per_tensor_eps,
                                   const float* per_tensor_weight_decay, float lr, float grad_scale, int step, int mode,
                                   int world, int rank, const void* const* grad_bases, void* const* param_bases,
                                   const void* grad_multicast, void* param_multicast, void* stream);

/* Same kernel with a per-block choice of route (needs the multicast addresses): of every 16 consecutive 32 KB blocks of
 * the flat buffers the first ld_p2p_of16 take their gradients by peer-to-peer loads and the first st_p2p_of16 deliver
 * their parameters by peer-to-peer stores; the other blocks go through the switch (multimem).  0 / 0 is the plain
 * multicast kernel above, 16 / 16 the plain peer-to-peer one.  Results are bit-identical for every split whenever the
 * rank sum is exact in the gradient dtype.  Same reference lines as above. */
int mt_multi_tensor_fused_adam_p2p_hybrid(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl,
                                          const float* per_tensor_beta1, const float* per_tensor_beta2,
                                          const int32_t* per_tensor_bias_correction, const float* per_tensor_eps,
                                          const float* per_tensor_weight_decay, float lr, float grad_scale, int step,
                                          int mode, int world, int rank, const void* const* grad_bases,
                                          void* const* param_bases, const void* grad_multicast, void* param_multicast,
                                          int ld_p2p_of16, int st_p2p_of16, void* stream);

/* L2 norm for callers that cannot allocate.  No counterpart in the reference (its multi_tensor_l2norm_cuda allocates
 * four tensors per call through torch).
 *
 * Same result as mt_multi_tensor_l2norm (same kernels, same summation order), but it takes no scratch pointer and makes
 * no cudaMalloc / cudaFree / cudaMallocHost call of its own: the tensor table and the per-chunk partial sums live in a
 * fixed __device__ array inside the library, mt_noalloc_resident_bytes() bytes (about 1.4 MB) on every GPU that loads it.
 * Meant for hosts bound to the bare C ABI that own no device allocator, for processes that gave all of HBM to one pool
 * at start-up (cudaMalloc would fail), and for steps that must not incur cudaMalloc's device-wide synchronisation.
 *
 * Limits: at most MT_NOALLOC_MAX_TENSORS tensors and MT_NOALLOC_MAX_CHUNKS chunks (sum over tensors of
 * ceil(numel / chunk_size)); beyond either the call returns MT_EINVAL with a message and launches nothing.
 * The array is shared by all calls in the process and keeps the previous call's contents; a call reads back only
 * what it has itself just written, and clears the one part that has to start at zero (its per-tensor completion
 * counters) with a cudaMemsetAsync on `stream` first.  Calls whose GPU work overlaps -- different streams or threads -- race
 * on it; calls on one stream are safe.  The first call on each device resolves the array's address
 * (cudaGetSymbolAddress): issue it before any stream capture begins.  The table is uploaded from pageable host
 * memory, so the call may wait for earlier work on `stream`.
 * Never call it on a stream that is being captured -- not the first call, not a later one: the table is staged through
 * one host buffer shared by all calls, so a graph replay would upload whatever the most recent call left there.  A
 * capturing stream is refused with MT_ECAPTURE and nothing is enqueued (the capture itself stays valid). */
#define MT_NOALLOC_MAX_TENSORS 16384
#define MT_NOALLOC_MAX_CHUNKS 262144
int mt_multi_tensor_l2norm_noalloc(int chunk_size, int32_t* noop_flag, const MtTensorLists* tl, int per_tensor,
                                   float* ret, float* ret_per_tensor, void* stream);
size_t mt_noalloc_resident_bytes(void);

/* Device-table cache introspection (the table replaces the by-value TensorListMetadata kernel argument,
 * /root/reference/csrc/multi_tensor_apply.cuh:19-26). */
typedef struct MtCacheStats {
    uint64_t hits, misses, uploads_bytes, entries, device_bytes;
    uint64_t evictions;          /* entries dropped by the LRU (only idle ones: their events had completed) */
    uint64_t cross_stream_waits; /* hits that had to order a stream behind another stream's upload / last use */
} MtCacheStats;
int mt_table_cache_stats(int device, MtCacheStats* out);
int mt_table_cache_clear(int device);

/* Launch counter: number of kernels this library has launched since load (all devices). */
uint64_t mt_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* MT_B200_H_ */
