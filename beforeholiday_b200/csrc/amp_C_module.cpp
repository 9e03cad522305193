// amp_C_module.cpp -- the `amp_C` Python extension: a thin torch/pybind11 adaptor over the C ABI in
// include/mt_b200.h.  It keeps the exact operator names and positional signatures of the reference module
// (/root/reference/csrc/amp_C_frontend.cpp:165-194) so that apex.optimizers.* / amp's LossScaler / user code
// calling amp_C.multi_tensor_*(chunk_size, noop_flag, tensor_lists, ...) run unchanged.
//
// What lives here: argument validation with the reference's error messages
// (/root/reference/csrc/multi_tensor_apply.cuh:53-72), flattening of list[list[Tensor]] into MtTensorLists,
// allocation of return / scratch tensors through torch's caching allocator (stream-ordered, capture-safe),
// and translation of error codes into RuntimeError.  No arithmetic happens here.
#include <torch/extension.h>

#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>

#include <unistd.h>

#include <algorithm>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <tuple>
#include <vector>

#include "../../include/mt_b200.h"
#include "walk_pool.h"

namespace {

using mt_host::WalkPool;
using TensorLists = py::object;  // list[list[Tensor]], walked in place (see flatten())

int dtype_tag(at::ScalarType t, const char* who) {
    switch (t) {
        case at::ScalarType::Float: return MT_F32;
        case at::ScalarType::Half: return MT_F16;
        case at::ScalarType::BFloat16: return MT_BF16;
        case at::ScalarType::Double: return MT_F64;
        case at::ScalarType::Byte: return MT_E5M2;  // e5m2 packing; only the sharded-LAMB copy / unpack ops take it
        default: TORCH_CHECK(false, who, ": unsupported dtype ", t);
    }
    return -1;
}

struct Flat {
    MtTensorLists tl{};
    std::vector<const void*> ptrs;
    std::vector<int64_t> numel;
    std::vector<at::Tensor> keep;  // densified sparse inputs
    at::Tensor first;              // tensor_lists[0][0]: device / options for allocations
    c10::DeviceIndex device = 0;
    cudaStream_t stream = nullptr;
};

// RAII for PySequence_Fast
struct FastSeq {
    PyObject* o;
    FastSeq(PyObject* src, const char* msg) : o(PySequence_Fast(src, msg)) {
        if (!o) throw py::error_already_set();
    }
    ~FastSeq() { Py_XDECREF(o); }
    Py_ssize_t size() const { return PySequence_Fast_GET_SIZE(o); }
    PyObject* operator[](Py_ssize_t i) const { return PySequence_Fast_GET_ITEM(o, i); }
    FastSeq(const FastSeq&) = delete;
    FastSeq& operator=(const FastSeq&) = delete;
};

// torch.nn.Parameter instances are THPVariable objects of a subclass type: same layout, no __torch_function__ games.
// Optimizers pass lists of them, so they must stay on the fast path (an isinstance() check needs the Python API).
PyTypeObject* g_parameter_type = nullptr;
inline bool is_plain_tensor(PyObject* o) {
    PyTypeObject* t = Py_TYPE(o);
    return t == reinterpret_cast<PyTypeObject*>(THPVariableClass) || (g_parameter_type && t == g_parameter_type);
}

// One (list, tensor) entry: the reference's per-tensor checks (/root/reference/csrc/multi_tensor_apply.cuh:60-71).
// Returns NULL or a static error message; never throws and never calls into Python, so it may run on worker threads.
// *special is set for entries the fast path does not handle (tensor subclasses needing isinstance, sparse tensors).
inline const char* check_entry(PyObject* item, const c10::Device& ref_device, at::ScalarType st0, bool first_list,
                               int64_t* numel_slot, const void** ptr_slot, bool* special, bool grad = false) {
    if (!is_plain_tensor(item)) { *special = true; return nullptr; }
    const at::Tensor& x0 = THPVariable_Unpack(item);
    // grad = true: the entry is the .grad of the listed parameter (the optimizers' fast path: no Python-side list of
    // gradients is ever built).  Reading autograd metadata is a pure read; safe on the worker threads.
    // TensorImpl::grad(), not Tensor::grad(): the latter warns (through Python, needing the GIL) for non-leaf tensors
    const at::Tensor& x = grad ? x0.unsafeGetTensorImpl()->grad() : x0;
    if (grad) {
        if (!x.defined()) return "grads_of: a parameter has no gradient";
        if (x.is_sparse()) return "grads_of: sparse gradient";
        if (x.sizes() != x0.sizes() || x.strides() != x0.strides()) return "grads_of: gradient layout differs from its parameter";
    }
    if (x.is_sparse()) { *special = true; return nullptr; }
    if (!x.is_contiguous() && !x.is_contiguous(at::MemoryFormat::ChannelsLast) &&
        !x.is_contiguous(at::MemoryFormat::ChannelsLast3d))
        return "A tensor was not contiguous.";
    if (x.device() != ref_device) return "A tensor was not on the same device as the first tensor";
    const int64_t n = x.numel();
    if (first_list) *numel_slot = n;
    else if (n != *numel_slot) return "Size mismatch";
    if (x.scalar_type() != st0) return "all tensors of one list must share a dtype";
    *ptr_slot = x.data_ptr();
    return nullptr;
}

// Validation + flattening.  Same checks, same messages as multi_tensor_apply<depth>(), but walking the Python lists
// directly: the reference's signature (std::vector<std::vector<at::Tensor>> by value) costs a type-cast plus a
// reference-count round trip per tensor, which at 50k tensors x 3 lists is several milliseconds -- more than the
// kernel itself on a B200.  Long lists are checked by a few worker threads (pure reads of tensor metadata).
// grads_of >= 0 (the optimizers' fast path): an extra list 0 is synthesised from the .grad of the tensors of given list
// `grads_of`, so the caller never builds a Python list of gradients; with grads_only the given lists themselves are
// dropped (norm of the gradients of a parameter list).  Any entry the fast path cannot take raises a RuntimeError whose
// message starts with "grads_of:" -- the Python caller falls back to building the lists itself.
void flatten(const py::handle& lists_obj, size_t depth_lo, size_t depth_hi, const char* who, Flat& f, int grads_of = -1,
             bool grads_only = false) {
    FastSeq outer(lists_obj.ptr(), "tensor_lists must be a sequence of sequences of tensors");
    const size_t given = (size_t)outer.size();
    TORCH_CHECK(given >= 1, "tensor_lists.size() != depth");
    TORCH_CHECK(grads_of < (int)given, "grads_of: list index out of range");
    const size_t depth = grads_of < 0 ? given : (grads_only ? 1 : given + 1);
    TORCH_CHECK(depth >= depth_lo && depth <= depth_hi, "tensor_lists.size() != depth");
    std::vector<std::unique_ptr<FastSeq>> inner;
    inner.reserve(given);
    for (size_t l = 0; l < given; l++)
        inner.emplace_back(new FastSeq(outer[l], "tensor_lists must be a sequence of sequences of tensors"));
    const size_t len0 = (size_t)inner[0]->size();
    TORCH_CHECK(len0 > 0, "tensor_lists[0].size() is not > 0");
    for (size_t l = 0; l < given; l++) TORCH_CHECK((size_t)inner[l]->size() == len0, "Size mismatch among tensor lists");
    // list sources: where list l's entries come from
    struct Src { PyObject** items; bool grad; };
    Src src[MT_MAX_DEPTH + 1];
    {
        size_t k = 0;
        if (grads_of >= 0) src[k++] = Src{PySequence_Fast_ITEMS(inner[grads_of]->o), true};
        if (!(grads_of >= 0 && grads_only))
            for (size_t l = 0; l < given; l++) src[k++] = Src{PySequence_Fast_ITEMS(inner[l]->o), false};
    }
    TORCH_CHECK(THPVariable_Check((*inner[0])[0]), who, ": tensor_lists must contain tensors");
    f.first = THPVariable_Unpack((*inner[0])[0]);
    const c10::Device ref_device = f.first.device();
    TORCH_CHECK(ref_device.type() == at::kCUDA, "expected input to be on cuda");
    f.ptrs.assign(depth * len0, nullptr);
    f.numel.assign(len0, 0);
    at::ScalarType st0[MT_MAX_DEPTH];
    for (size_t l = 0; l < depth; l++) {
        PyObject* it0 = src[l].items[0];
        TORCH_CHECK(THPVariable_Check(it0), who, ": tensor_lists must contain tensors");
        const at::Tensor& x0 = THPVariable_Unpack(it0);
        const at::Tensor& g0 = x0.unsafeGetTensorImpl()->grad();
        if (src[l].grad) TORCH_CHECK(g0.defined(), "grads_of: a parameter has no gradient");
        st0[l] = src[l].grad ? g0.scalar_type() : x0.scalar_type();
        f.tl.dtype[l] = dtype_tag(st0[l], who);
    }

    // fast path over [t0, t1): list 0 first (it defines numel), then the others
    auto run_range = [&](size_t t0, size_t t1, const char** err, bool* special) {
        for (size_t l = 0; l < depth && !*err; l++) {
            PyObject** items = src[l].items;
            const bool grad = src[l].grad;
            for (size_t t = t0; t < t1; t++) {
                // The walk is bound by cache misses on objects scattered over the heap (PyObject -> TensorImpl ->
                // StorageImpl, one dependent miss each): run three prefetch stages ahead of the checks.
                if (t + 24 < t1) __builtin_prefetch(items[t + 24]);
                if (t + 16 < t1 && is_plain_tensor(items[t + 16])) {
                    const char* impl = reinterpret_cast<const char*>(THPVariable_Unpack(items[t + 16]).unsafeGetTensorImpl());
                    __builtin_prefetch(impl);
                    __builtin_prefetch(impl + 64);
                    __builtin_prefetch(impl + 128);
                }
                if (t + 8 < t1 && is_plain_tensor(items[t + 8]) && !grad)
                    __builtin_prefetch(THPVariable_Unpack(items[t + 8]).unsafeGetTensorImpl()->unsafe_storage().unsafeGetStorageImpl());
                const char* e = check_entry(items[t], ref_device, st0[l], l == 0, &f.numel[t], &f.ptrs[l * len0 + t], special, grad);
                if (e) { *err = e; break; }
            }
        }
    };
    const char* err = nullptr;
    bool special = false;
    const size_t entries = depth * len0;
    unsigned nthreads = 1;
    if (entries >= 32768) nthreads = WalkPool::get().workers() + 1;
    if (nthreads <= 1) {
        run_range(0, len0, &err, &special);
    } else {
        std::vector<const char*> errs(nthreads, nullptr);
        std::vector<char> specials(nthreads, 0);
        const size_t per = (len0 + nthreads - 1) / nthreads;
        WalkPool::get().run(nthreads, [&](unsigned w) {
            const size_t t0 = std::min(len0, (size_t)w * per), t1 = std::min(len0, t0 + per);
            bool sp = false;
            run_range(t0, t1, &errs[w], &sp);
            specials[w] = sp;
        });
        for (unsigned w = 0; w < nthreads; w++) {
            if (errs[w] && !err) err = errs[w];
            special = special || specials[w];
        }
    }
    if (special) {
        // slow path (rare): tensor subclasses that need isinstance(), sparse tensors (densified as
        // multi_tensor_apply.cuh:92-95 does).  Serial, with the GIL, may allocate.
        TORCH_CHECK(grads_of < 0, "grads_of: unsupported tensor kind (subclass or sparse) in the lists");
        err = nullptr;
        for (size_t l = 0; l < depth; l++)
            for (size_t t = 0; t < len0; t++) {
                PyObject* item = (*inner[l])[t];
                TORCH_CHECK(THPVariable_Check(item), who, ": tensor_lists must contain tensors");
                const at::Tensor& x = THPVariable_Unpack(item);
                if (!x.is_sparse()) {
                    bool sp2 = false;
                    PyObject* exact = item;
                    const char* e = is_plain_tensor(exact)
                        ? check_entry(item, ref_device, st0[l], l == 0, &f.numel[t], &f.ptrs[l * len0 + t], &sp2) : nullptr;
                    if (!is_plain_tensor(exact)) {  // subclass: same checks through the Tensor API
                        bool cm = x.is_contiguous() || x.is_contiguous(at::MemoryFormat::ChannelsLast) ||
                                  x.is_contiguous(at::MemoryFormat::ChannelsLast3d);
                        TORCH_CHECK(cm, "A tensor was not contiguous.");
                        TORCH_CHECK(x.device() == ref_device, "A tensor was not on the same device as the first tensor");
                        if (l == 0) f.numel[t] = x.numel(); else TORCH_CHECK(x.numel() == f.numel[t], "Size mismatch");
                        TORCH_CHECK(x.scalar_type() == st0[l], who, ": all tensors of one list must share a dtype");
                        f.ptrs[l * len0 + t] = x.data_ptr();
                    }
                    TORCH_CHECK(!e, e);
                    continue;
                }
                TORCH_CHECK(x._values().is_contiguous(), "A tensor was not contiguous.");
                TORCH_CHECK(x.device() == ref_device, "A tensor was not on the same device as the first tensor");
                if (l == 0) f.numel[t] = x.numel(); else TORCH_CHECK(x.numel() == f.numel[t], "Size mismatch");
                TORCH_CHECK(x.scalar_type() == st0[l], who, ": all tensors of one list must share a dtype");
                at::Tensor dst = at::zeros(x.sizes(), x.options().layout(at::kStrided));
                dst.add_(x);
                f.keep.push_back(dst);
                f.ptrs[l * len0 + t] = dst.data_ptr();
            }
    }
    TORCH_CHECK(!err, err);
    f.device = ref_device.index();
    f.tl.depth = (int32_t)depth;
    f.tl.device = (int32_t)f.device;
    f.tl.ntensors = (int64_t)len0;
    f.tl.ptrs = f.ptrs.data();
    f.tl.numel = f.numel.data();
    f.stream = at::cuda::getCurrentCUDAStream(f.device).stream();
}

int32_t* flag_ptr(const at::Tensor& noop_flag) {
    TORCH_CHECK(noop_flag.is_cuda() && noop_flag.scalar_type() == at::ScalarType::Int && noop_flag.numel() >= 1,
                "noop_flag must be a CUDA int32 tensor with at least one element");
    return noop_flag.data_ptr<int32_t>();
}

void check(int rc) { TORCH_CHECK(rc == MT_OK, mt_last_error()); }

// ---------------------------------------------------------------------------------------------------
void multi_tensor_scale(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists, float scale) {
    Flat f;
    flatten(tensor_lists, 2, 2, "multi_tensor_scale", f);
    check(mt_multi_tensor_scale(chunk_size, flag_ptr(noop_flag), &f.tl, scale, f.stream));
}

void multi_tensor_axpby(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists, float a, float b,
                        int arg_to_check) {
    Flat f;
    flatten(tensor_lists, 3, 3, "multi_tensor_axpby", f);
    check(mt_multi_tensor_axpby(chunk_size, flag_ptr(noop_flag), &f.tl, a, b, arg_to_check, f.stream));
}

std::tuple<at::Tensor, at::Tensor> l2norm_impl(int chunk_size, at::Tensor noop_flag, const py::object& tensor_lists,
                                               bool with_scale, float scale, at::optional<bool> per_tensor_python,
                                               bool mp = false, int grads_of = -1) {
    const bool per_tensor = per_tensor_python.has_value() ? per_tensor_python.value() : false;
    Flat f;
    flatten(tensor_lists, with_scale ? 2 : 1, with_scale ? 2 : 1, with_scale ? "multi_tensor_l2norm_scale" : "multi_tensor_l2norm", f,
            grads_of, /*grads_only=*/grads_of >= 0);
    const c10::cuda::CUDAGuard guard(f.device);
    auto fopt = f.first.options().dtype(at::kFloat).layout(at::kStrided);
    at::Tensor ret = at::empty({1}, fopt);
    at::Tensor ret_per_tensor = at::empty({per_tensor ? (int64_t)f.tl.ntensors : 0}, fopt);
    const size_t sbytes = mt_norm_scratch_bytes(chunk_size, &f.tl);
    at::Tensor scratch = at::empty({(int64_t)(sbytes ? sbytes : 4)}, fopt.dtype(at::kByte));
    int rc;
    if (with_scale)
        rc = mt_multi_tensor_l2norm_scale(chunk_size, flag_ptr(noop_flag), &f.tl, scale, per_tensor,
                                          ret.data_ptr<float>(), per_tensor ? ret_per_tensor.data_ptr<float>() : nullptr,
                                          scratch.data_ptr(), sbytes, f.stream);
    else if (mp)
        rc = mt_multi_tensor_l2norm_mp(chunk_size, flag_ptr(noop_flag), &f.tl, per_tensor, ret.data_ptr<float>(),
                                       per_tensor ? ret_per_tensor.data_ptr<float>() : nullptr, scratch.data_ptr(), sbytes,
                                       f.stream);
    else
        rc = mt_multi_tensor_l2norm(chunk_size, flag_ptr(noop_flag), &f.tl, per_tensor, ret.data_ptr<float>(),
                                    per_tensor ? ret_per_tensor.data_ptr<float>() : nullptr, scratch.data_ptr(), sbytes,
                                    f.stream);
    check(rc);
    return std::make_tuple(ret, ret_per_tensor);
}

// grads_of (extension, default -1): norm of the GRADIENTS of the parameters in tensor_lists[grads_of]
std::tuple<at::Tensor, at::Tensor> multi_tensor_l2norm(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists,
                                                       at::optional<bool> per_tensor_python, int grads_of) {
    return l2norm_impl(chunk_size, noop_flag, tensor_lists, false, 0.f, per_tensor_python, false, grads_of);
}

std::tuple<at::Tensor, at::Tensor> multi_tensor_l2norm_scale(int chunk_size, at::Tensor noop_flag,
                                                             TensorLists tensor_lists, float scale,
                                                             at::optional<bool> per_tensor_python) {
    return l2norm_impl(chunk_size, noop_flag, tensor_lists, true, scale, per_tensor_python);
}

std::tuple<at::Tensor, at::Tensor> multi_tensor_l2norm_mp(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists,
                                                          at::optional<bool> per_tensor_python) {
    return l2norm_impl(chunk_size, noop_flag, tensor_lists, false, 0.f, per_tensor_python, true);
}

void multi_tensor_lamb_mp(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists, at::Tensor lr, const float beta1,
                          const float beta2, const float epsilon, at::Tensor step, const int bias_correction,
                          const float weight_decay, const int grad_averaging, const int mode, at::Tensor global_grad_norm,
                          at::Tensor max_grad_norm, at::optional<bool> use_nvlamb_python, at::Tensor found_inf,
                          at::Tensor inv_scale) {
    const bool use_nvlamb = use_nvlamb_python.has_value() ? use_nvlamb_python.value() : false;
    Flat f;
    flatten(tensor_lists, 4, 5, "multi_tensor_lamb_mp", f);
    TORCH_CHECK(lr.is_cuda() && lr.scalar_type() == at::kFloat && step.is_cuda() && step.scalar_type() == at::kInt &&
                    global_grad_norm.is_cuda() && max_grad_norm.is_cuda() && inv_scale.is_cuda() && found_inf.is_cuda(),
                "multi_tensor_lamb_mp: lr (float), step (int), global_grad_norm, max_grad_norm, found_inf, inv_scale must be CUDA tensors");
    const c10::cuda::CUDAGuard guard(f.device);
    const size_t sbytes = mt_lamb_scratch_bytes(chunk_size, &f.tl);
    at::Tensor scratch = at::empty({(int64_t)(sbytes ? sbytes : 4)}, f.first.options().dtype(at::kByte).layout(at::kStrided));
    check(mt_multi_tensor_lamb_mp(chunk_size, flag_ptr(noop_flag), &f.tl, lr.data_ptr<float>(), beta1, beta2, epsilon,
                                  step.data_ptr<int32_t>(), bias_correction, weight_decay, grad_averaging, mode,
                                  global_grad_norm.data_ptr<float>(), max_grad_norm.data_ptr<float>(), use_nvlamb,
                                  found_inf.data_ptr<float>(), inv_scale.data_ptr<float>(), scratch.data_ptr(), sbytes,
                                  f.stream));
}

void multi_tensor_adam(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists, const float lr,
                       const float beta1, const float beta2, const float epsilon, const int step, const int mode,
                       const int bias_correction, const float weight_decay, int grads_of) {
    Flat f;
    flatten(tensor_lists, 4, 4, "multi_tensor_adam", f, grads_of);
    check(mt_multi_tensor_adam(chunk_size, flag_ptr(noop_flag), &f.tl, lr, beta1, beta2, epsilon, step, mode,
                               bias_correction, weight_decay, f.stream));
}

void multi_tensor_sgd(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists, float wd, float momentum,
                      float dampening, float lr, bool nesterov, bool first_run, bool wd_after_momentum, float scale,
                      int grads_of) {
    Flat f;
    flatten(tensor_lists, 3, 4, "multi_tensor_sgd", f, grads_of);
    if (f.tl.depth == 4) TORCH_CHECK(f.tl.dtype[3] == MT_F16, "Additional output tensors should always be fp16.");
    TORCH_CHECK(noop_flag.device() == f.first.device(),
                "expected noop flag to be on the same device as tensors");
    check(mt_multi_tensor_sgd(chunk_size, flag_ptr(noop_flag), &f.tl, wd, momentum, dampening, lr, nesterov, first_run,
                              wd_after_momentum, scale, f.stream));
}

void multi_tensor_lamb(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists, const float lr,
                       const float beta1, const float beta2, const float epsilon, const int step,
                       const int bias_correction, const float weight_decay, const int grad_averaging, const int mode,
                       at::Tensor global_grad_norm, const float max_grad_norm, at::optional<bool> use_nvlamb_python,
                       int grads_of) {
    const bool use_nvlamb = use_nvlamb_python.has_value() ? use_nvlamb_python.value() : false;
    Flat f;
    flatten(tensor_lists, 4, 4, "multi_tensor_lamb", f, grads_of);
    TORCH_CHECK(global_grad_norm.is_cuda() && global_grad_norm.scalar_type() == at::ScalarType::Float,
                "global_grad_norm must be a CUDA float tensor");
    const c10::cuda::CUDAGuard guard(f.device);
    const size_t sbytes = mt_lamb_scratch_bytes(chunk_size, &f.tl);
    at::Tensor scratch = at::empty({(int64_t)(sbytes ? sbytes : 4)},
                                   f.first.options().dtype(at::kByte).layout(at::kStrided));
    check(mt_multi_tensor_lamb(chunk_size, flag_ptr(noop_flag), &f.tl, lr, beta1, beta2, epsilon, step, bias_correction,
                               weight_decay, grad_averaging, mode, global_grad_norm.data_ptr<float>(), max_grad_norm,
                               use_nvlamb, scratch.data_ptr(), sbytes, f.stream));
}

void multi_tensor_lamb_stage1_cuda(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists,
                                   at::Tensor per_tensor_decay, const int step, const float beta1, const float beta2,
                                   const float epsilon, at::Tensor global_grad_norm,
                                   const float max_global_grad_norm) {
    Flat f;
    flatten(tensor_lists, 5, 5, "multi_tensor_lamb_stage1_cuda", f);
    // The reference dereferences the norm pointer on the host (csrc/multi_tensor_lamb_stage_1.cu:126-127), i.e. it
    // only works for a host-readable tensor; .item() accepts both host and device tensors.
    const float gnorm = global_grad_norm.item<float>();
    check(mt_multi_tensor_lamb_stage1(chunk_size, flag_ptr(noop_flag), &f.tl, per_tensor_decay.data_ptr<float>(), step,
                                      beta1, beta2, epsilon, gnorm, max_global_grad_norm, f.stream));
}

void multi_tensor_lamb_stage2_cuda(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists,
                                   at::Tensor per_tensor_param_norm, at::Tensor per_tensor_update_norm, const float lr,
                                   const float weight_decay, at::optional<bool> use_nvlamb_python) {
    const bool use_nvlamb = use_nvlamb_python.has_value() ? use_nvlamb_python.value() : false;
    Flat f;
    flatten(tensor_lists, 2, 2, "multi_tensor_lamb_stage2_cuda", f);
    check(mt_multi_tensor_lamb_stage2(chunk_size, flag_ptr(noop_flag), &f.tl, per_tensor_param_norm.data_ptr<float>(),
                                      per_tensor_update_norm.data_ptr<float>(), lr, weight_decay, use_nvlamb, f.stream));
}

void multi_tensor_adagrad(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists, const float lr,
                          const float epsilon, const int mode, const float weight_decay) {
    Flat f;
    flatten(tensor_lists, 3, 3, "multi_tensor_adagrad", f);
    check(mt_multi_tensor_adagrad(chunk_size, flag_ptr(noop_flag), &f.tl, lr, epsilon, mode, weight_decay, f.stream));
}

void multi_tensor_novograd(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists, at::Tensor grad_norms,
                           const float lr, const float beta1, const float beta2, const float epsilon, const int step,
                           const int bias_correction, const float weight_decay, const int grad_averaging,
                           const int mode, const int norm_type) {
    Flat f;
    flatten(tensor_lists, 3, 3, "multi_tensor_novograd", f);
    TORCH_CHECK(grad_norms.is_cuda() && grad_norms.scalar_type() == at::kFloat && grad_norms.numel() >= f.tl.ntensors,
                "grad_norms must be a CUDA float tensor with one entry per tensor");
    TORCH_CHECK(f.first.device() == noop_flag.device(), "noop flag should be on the same device as tensors");
    const c10::cuda::CUDAGuard guard(f.device);
    const size_t sbytes = mt_norm_scratch_bytes(chunk_size, &f.tl);
    at::Tensor scratch = at::empty({(int64_t)(sbytes ? sbytes : 4)}, f.first.options().dtype(at::kByte).layout(at::kStrided));
    check(mt_multi_tensor_novograd(chunk_size, flag_ptr(noop_flag), &f.tl, grad_norms.data_ptr<float>(), lr, beta1, beta2,
                                   epsilon, step, bias_correction, weight_decay, grad_averaging, mode, norm_type,
                                   scratch.data_ptr(), sbytes, f.stream));
}

void multi_tensor_lars(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists, at::Tensor grad_norms,
                       at::Tensor param_norms, float lr, float trust_coefficient, float epsilon, float weight_decay,
                       float momentum, float dampening, bool nesterov, bool first_run, bool wd_after_momentum, float scale,
                       const bool is_skipped) {
    Flat f;
    flatten(tensor_lists, 3, 4, "multi_tensor_lars", f);
    if (f.tl.depth == 4) TORCH_CHECK(f.tl.dtype[3] == MT_F16, "Additional output tensors should always be fp16.");
    TORCH_CHECK(noop_flag.device() == f.first.device(), "expected noop flag to be on the same device as tensors");
    TORCH_CHECK(grad_norms.is_cuda() && param_norms.is_cuda() && grad_norms.scalar_type() == at::kFloat &&
                    param_norms.scalar_type() == at::kFloat, "grad_norms / param_norms must be CUDA float tensors");
    check(mt_multi_tensor_lars(chunk_size, flag_ptr(noop_flag), &f.tl, grad_norms.data_ptr<float>(),
                               param_norms.data_ptr<float>(), lr, trust_coefficient, epsilon, weight_decay, momentum,
                               dampening, nesterov, first_run, wd_after_momentum, scale, is_skipped, f.stream));
}

// distributed_adam_cuda.multi_tensor_fused_adam
// (/root/reference/apex/contrib/csrc/optimizers/multi_tensor_distopt_adam.cpp:3-19) plus a trailing
// `honour_noop` (default false = reference behaviour).
void multi_tensor_fused_adam(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists,
                             at::Tensor per_tensor_beta1, at::Tensor per_tensor_beta2,
                             at::Tensor per_tensor_bias_correction, at::Tensor per_tensor_eps,
                             at::Tensor per_tensor_weight_decay, float lr, float grad_scale, int step, int mode,
                             bool honour_noop) {
    Flat f;
    TORCH_CHECK(py::len(tensor_lists) == 4 || py::len(tensor_lists) == 5, "expected tensor lists of size 4 or 5");
    flatten(tensor_lists, 4, 5, "multi_tensor_fused_adam", f);
    const int64_t n = f.tl.ntensors;
    TORCH_CHECK(per_tensor_beta1.numel() >= n && per_tensor_beta2.numel() >= n && per_tensor_bias_correction.numel() >= n &&
                    per_tensor_eps.numel() >= n && per_tensor_weight_decay.numel() >= n,
                "per-tensor hyper-parameter tensors are shorter than the tensor lists");
    TORCH_CHECK(per_tensor_bias_correction.scalar_type() == at::ScalarType::Int, "per_tensor_bias_correction must be int32");
    check(mt_multi_tensor_fused_adam(chunk_size, flag_ptr(noop_flag), &f.tl, per_tensor_beta1.data_ptr<float>(),
                                     per_tensor_beta2.data_ptr<float>(), per_tensor_bias_correction.data_ptr<int32_t>(),
                                     per_tensor_eps.data_ptr<float>(), per_tensor_weight_decay.data_ptr<float>(), lr,
                                     grad_scale, step, mode, honour_noop ? 1 : 0, f.stream));
}

// Capturable form: lr, grad_scale (= loss scale, g is divided by it) and step are CUDA scalar tensors.
void multi_tensor_fused_adam_capturable(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists,
                                        at::Tensor per_tensor_beta1, at::Tensor per_tensor_beta2,
                                        at::Tensor per_tensor_bias_correction, at::Tensor per_tensor_eps,
                                        at::Tensor per_tensor_weight_decay, at::Tensor lr, at::Tensor grad_scale,
                                        at::Tensor step, int mode) {
    Flat f;
    flatten(tensor_lists, 4, 5, "multi_tensor_fused_adam_capturable", f);
    TORCH_CHECK(lr.is_cuda() && lr.scalar_type() == at::kFloat && grad_scale.is_cuda() &&
                    grad_scale.scalar_type() == at::kFloat && step.is_cuda() && step.scalar_type() == at::kInt,
                "lr (float), grad_scale (float) and step (int32) must be CUDA tensors");
    TORCH_CHECK(per_tensor_bias_correction.scalar_type() == at::ScalarType::Int, "per_tensor_bias_correction must be int32");
    check(mt_multi_tensor_fused_adam_capturable(chunk_size, flag_ptr(noop_flag), &f.tl, per_tensor_beta1.data_ptr<float>(),
                                                per_tensor_beta2.data_ptr<float>(),
                                                per_tensor_bias_correction.data_ptr<int32_t>(),
                                                per_tensor_eps.data_ptr<float>(), per_tensor_weight_decay.data_ptr<float>(),
                                                lr.data_ptr<float>(), grad_scale.data_ptr<float>(),
                                                step.data_ptr<int32_t>(), mode, f.stream));
}

// fused_adam_cuda.maybe_cast_mt (/root/reference/apex/contrib/csrc/optimizers/fused_adam_cuda.cpp:13,85)
void maybe_cast_mt(int chunk_size, at::Tensor overflow_flag, TensorLists tensor_lists) {
    Flat f;
    TORCH_CHECK(py::len(tensor_lists) == 2, "expected tensor lists of size 2");
    flatten(tensor_lists, 2, 2, "maybe_cast_mt", f);
    check(mt_multi_tensor_maybe_cast(chunk_size, flag_ptr(overflow_flag), &f.tl, f.stream));
}

// distributed_lamb_cuda.multi_tensor_lamb_compute_update_term
// (/root/reference/apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb.cpp:3-17).  Two trailing optional
// arguments switch on the fused per-parameter norm sums: `seg_begin` (CUDA int32[nseg+1]) and `model_dtype`
// (torch dtype of the model parameters); the function then returns a CUDA float[2][nseg] tensor (else None).
py::object multi_tensor_lamb_compute_update_term(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists,
                                                 at::Tensor per_tensor_beta1, at::Tensor per_tensor_beta2,
                                                 at::Tensor per_tensor_beta3, at::Tensor per_tensor_bias_correction,
                                                 at::Tensor step, at::Tensor per_tensor_epsilon, const int mode,
                                                 at::Tensor per_tensor_decay, at::Tensor global_scale,
                                                 at::Tensor global_grad_norm, const float max_grad_norm,
                                                 c10::optional<at::Tensor> seg_begin,
                                                 c10::optional<at::ScalarType> model_dtype) {
    Flat f;
    flatten(tensor_lists, 5, 5, "multi_tensor_lamb_compute_update_term", f);
    const int64_t n = f.tl.ntensors;
    TORCH_CHECK(per_tensor_beta1.numel() >= n && per_tensor_beta2.numel() >= n && per_tensor_bias_correction.numel() >= n &&
                    per_tensor_epsilon.numel() >= n && per_tensor_decay.numel() >= n,
                "per-tensor hyper-parameter tensors are shorter than the tensor lists");
    TORCH_CHECK(per_tensor_beta1.scalar_type() == at::kFloat && per_tensor_beta2.scalar_type() == at::kFloat &&
                    per_tensor_epsilon.scalar_type() == at::kFloat && per_tensor_decay.scalar_type() == at::kFloat &&
                    global_scale.scalar_type() == at::kFloat && global_grad_norm.scalar_type() == at::kFloat,
                "per-tensor hyper-parameters, global_scale and global_grad_norm must be float32 (the update term's type)");
    TORCH_CHECK(per_tensor_bias_correction.scalar_type() == at::kInt && step.scalar_type() == at::kInt,
                "per_tensor_bias_correction and step must be int32");
    TORCH_CHECK(step.is_cuda() && global_scale.is_cuda() && global_grad_norm.is_cuda(),
                "step, global_scale and global_grad_norm must be CUDA tensors");
    const c10::cuda::CUDAGuard guard(f.device);
    at::Tensor out, scratch;
    const int32_t* seg = nullptr;
    int nseg = 0, mdt = MT_F32;
    size_t sbytes = 0;
    if (seg_begin.has_value()) {
        TORCH_CHECK(seg_begin->is_cuda() && seg_begin->scalar_type() == at::kInt && seg_begin->numel() >= 2,
                    "seg_begin must be a CUDA int32 tensor of nseg+1 fragment indices");
        nseg = (int)seg_begin->numel() - 1;
        seg = seg_begin->data_ptr<int32_t>();
        const at::ScalarType md = model_dtype.has_value() ? model_dtype.value() : at::kFloat;
        mdt = md == at::kHalf ? MT_F16 : md == at::kBFloat16 ? MT_BF16 : md == at::kByte ? MT_E5M2 : MT_F32;
        TORCH_CHECK(md == at::kHalf || md == at::kBFloat16 || md == at::kFloat || md == at::kByte,
                    "model_dtype must be float32, float16, bfloat16, or uint8 for parameters that travel e5m2-packed");
        out = at::zeros({2, nseg}, f.first.options().dtype(at::kFloat).layout(at::kStrided));
        sbytes = mt_dist_lamb_scratch_bytes(&f.tl);
        scratch = at::empty({(int64_t)(sbytes ? sbytes : 4)}, f.first.options().dtype(at::kByte).layout(at::kStrided));
    }
    check(mt_multi_tensor_lamb_compute_update_term(
        chunk_size, flag_ptr(noop_flag), &f.tl, per_tensor_beta1.data_ptr<float>(), per_tensor_beta2.data_ptr<float>(),
        per_tensor_beta3.defined() && per_tensor_beta3.numel() ? per_tensor_beta3.data_ptr<float>() : nullptr,
        per_tensor_bias_correction.data_ptr<int32_t>(), step.data_ptr<int32_t>(), per_tensor_epsilon.data_ptr<float>(),
        mode, per_tensor_decay.data_ptr<float>(), global_scale.data_ptr<float>(), global_grad_norm.data_ptr<float>(),
        max_grad_norm, seg, nseg, mdt, seg ? out.data_ptr<float>() : nullptr, seg ? scratch.data_ptr() : nullptr, sbytes,
        f.stream));
    if (seg) return py::cast(out);
    return py::none();
}

// distributed_lamb_cuda.multi_tensor_lamb_update_weights (multi_tensor_distopt_lamb.cpp:19-29)
void multi_tensor_lamb_update_weights(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists,
                                      at::Tensor per_tensor_param_norm, at::Tensor per_tensor_update_norm,
                                      at::Tensor update_norm_offset, at::Tensor learning_rate,
                                      at::Tensor per_tensor_decay, at::Tensor global_grad_norm, bool use_nvlamb) {
    Flat f;
    flatten(tensor_lists, 3, 3, "multi_tensor_lamb_update_weights", f);
    const int64_t n = f.tl.ntensors;
    TORCH_CHECK(per_tensor_param_norm.numel() >= n && update_norm_offset.numel() >= n && per_tensor_decay.numel() >= n,
                "per-tensor tensors are shorter than the tensor lists");
    TORCH_CHECK(per_tensor_param_norm.scalar_type() == at::kFloat && per_tensor_update_norm.scalar_type() == at::kFloat &&
                    learning_rate.scalar_type() == at::kFloat && per_tensor_decay.scalar_type() == at::kFloat,
                "norms, learning_rate and per_tensor_decay must be float32 (the update term's type)");
    TORCH_CHECK(update_norm_offset.scalar_type() == at::kLong, "update_norm_offset must be int64");
    TORCH_CHECK(learning_rate.is_cuda() && update_norm_offset.is_cuda(), "learning_rate and update_norm_offset must be CUDA tensors");
    check(mt_multi_tensor_lamb_update_weights(
        chunk_size, flag_ptr(noop_flag), &f.tl, per_tensor_param_norm.data_ptr<float>(),
        per_tensor_update_norm.data_ptr<float>(), update_norm_offset.data_ptr<int64_t>(), learning_rate.data_ptr<float>(),
        per_tensor_decay.data_ptr<float>(), global_grad_norm.defined() && global_grad_norm.numel() ? global_grad_norm.data_ptr<float>() : nullptr,
        use_nvlamb ? 1 : 0, f.stream));
}

// Fused reduce-scatter + Adam + all-gather over peer memory.  grad_ptrs / param_ptrs: every rank's flat buffer base
// address as mapped in this process (e.g. torch.distributed._symmetric_memory handle.buffer_ptrs).
void multi_tensor_fused_adam_p2p(int chunk_size, at::Tensor noop_flag, TensorLists tensor_lists,
                                 at::Tensor per_tensor_beta1, at::Tensor per_tensor_beta2,
                                 at::Tensor per_tensor_bias_correction, at::Tensor per_tensor_eps,
                                 at::Tensor per_tensor_weight_decay, float lr, float grad_scale, int step, int mode,
                                 int rank, std::vector<uint64_t> grad_ptrs, std::vector<uint64_t> param_ptrs,
                                 uint64_t grad_multicast_ptr, uint64_t param_multicast_ptr, int ld_p2p_of16,
                                 int st_p2p_of16) {
    Flat f;
    flatten(tensor_lists, 5, 5, "multi_tensor_fused_adam_p2p", f);
    TORCH_CHECK(grad_ptrs.size() == param_ptrs.size(), "grad_ptrs and param_ptrs must have one entry per rank");
    const int world = (int)grad_ptrs.size();
    std::vector<const void*> gb(world);
    std::vector<void*> pb(world);
    for (int r = 0; r < world; r++) {
        gb[r] = reinterpret_cast<const void*>(grad_ptrs[r]);
        pb[r] = reinterpret_cast<void*>(param_ptrs[r]);
    }
    TORCH_CHECK(per_tensor_bias_correction.scalar_type() == at::ScalarType::Int, "per_tensor_bias_correction must be int32");
    check(mt_multi_tensor_fused_adam_p2p_hybrid(
        chunk_size, flag_ptr(noop_flag), &f.tl, per_tensor_beta1.data_ptr<float>(), per_tensor_beta2.data_ptr<float>(),
        per_tensor_bias_correction.data_ptr<int32_t>(), per_tensor_eps.data_ptr<float>(),
        per_tensor_weight_decay.data_ptr<float>(), lr, grad_scale, step, mode, world, rank, gb.data(), pb.data(),
        reinterpret_cast<const void*>(grad_multicast_ptr), reinterpret_cast<void*>(param_multicast_ptr), ld_p2p_of16,
        st_p2p_of16, f.stream));
}

py::dict table_cache_stats(int device) {
    MtCacheStats s{};
    check(mt_table_cache_stats(device, &s));
    py::dict d;
    d["hits"] = s.hits;
    d["misses"] = s.misses;
    d["uploads_bytes"] = s.uploads_bytes;
    d["entries"] = s.entries;
    d["device_bytes"] = s.device_bytes;
    d["evictions"] = s.evictions;
    d["cross_stream_waits"] = s.cross_stream_waits;
    return d;
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    m.doc() = "B200-native amp_C: multi-tensor-apply optimizer path over libmt_b200.so";
    m.def("multi_tensor_scale", &multi_tensor_scale, "Fused overflow check + scale for a list of contiguous tensors");
    m.def("multi_tensor_sgd", &multi_tensor_sgd, "Fused SGD optimizer for list of contiguous tensors", py::arg("chunk_size"), py::arg("noop_flag"), py::arg("tensor_lists"), py::arg("wd"), py::arg("momentum"), py::arg("dampening"), py::arg("lr"), py::arg("nesterov"), py::arg("first_run"), py::arg("wd_after_momentum"), py::arg("scale"), py::arg("grads_of") = -1);
    m.def("multi_tensor_axpby", &multi_tensor_axpby, "out = a*x + b*y for a list of contiguous tensors");
    m.def("multi_tensor_l2norm", &multi_tensor_l2norm, "Computes L2 norm for a list of contiguous tensors", py::arg("chunk_size"), py::arg("noop_flag"), py::arg("tensor_lists"), py::arg("per_tensor") = py::none(), py::arg("grads_of") = -1);
    m.def("multi_tensor_l2norm_mp", &multi_tensor_l2norm_mp, "Computes L2 norm for a list of contiguous tensors");
    m.def("multi_tensor_lamb_mp", &multi_tensor_lamb_mp, "Computes and apply update for LAMB optimizer");
    m.def("multi_tensor_l2norm_scale", &multi_tensor_l2norm_scale,
          "Computes L2 norm for a list of contiguous tensors and does scaling");
    m.def("multi_tensor_lamb_stage1_cuda", &multi_tensor_lamb_stage1_cuda, "Computes update part of LAMB optimizer");
    m.def("multi_tensor_lamb_stage2_cuda", &multi_tensor_lamb_stage2_cuda,
          "Completes application of gradient to parameters for LAMB optimizer");
    m.def("multi_tensor_adam", &multi_tensor_adam, "Compute and apply gradient update to parameters for Adam optimizer", py::arg("chunk_size"), py::arg("noop_flag"), py::arg("tensor_lists"), py::arg("lr"), py::arg("beta1"), py::arg("beta2"), py::arg("epsilon"), py::arg("step"), py::arg("mode"), py::arg("bias_correction"), py::arg("weight_decay"), py::arg("grads_of") = -1);
    m.def("multi_tensor_lamb", &multi_tensor_lamb, "Computes and apply update for LAMB optimizer", py::arg("chunk_size"), py::arg("noop_flag"), py::arg("tensor_lists"), py::arg("lr"), py::arg("beta1"), py::arg("beta2"), py::arg("epsilon"), py::arg("step"), py::arg("bias_correction"), py::arg("weight_decay"), py::arg("grad_averaging"), py::arg("mode"), py::arg("global_grad_norm"), py::arg("max_grad_norm"), py::arg("use_nvlamb") = py::none(), py::arg("grads_of") = -1);
    m.def("multi_tensor_adagrad", &multi_tensor_adagrad, "Compute and apply gradient update to parameters for Adagrad optimizer");
    m.def("multi_tensor_novograd", &multi_tensor_novograd, "Compute and apply gradient update to parameters for NovoGrad optimizer");
    m.def("multi_tensor_lars", &multi_tensor_lars, "Fused LARS optimizer for list of contiguous tensors");
    m.def("multi_tensor_fused_adam", &multi_tensor_fused_adam, "Mixed-precision / sharded Adam (DistAdamFunctor)",
          py::arg("chunk_size"), py::arg("noop_flag"), py::arg("tensor_lists"), py::arg("per_tensor_beta1"),
          py::arg("per_tensor_beta2"), py::arg("per_tensor_bias_correction"), py::arg("per_tensor_eps"),
          py::arg("per_tensor_weight_decay"), py::arg("lr"), py::arg("grad_scale"), py::arg("step"), py::arg("mode"),
          py::arg("honour_noop") = false);
    m.def("multi_tensor_fused_adam_capturable", &multi_tensor_fused_adam_capturable,
          "Mixed-precision Adam with device-side lr / grad_scale / step (CUDA-graph capturable)");
    m.def("multi_tensor_fused_adam_p2p", &multi_tensor_fused_adam_p2p,
          "Fused gradient reduce-scatter + sharded Adam + parameter all-gather over NVLink peer memory",
          py::arg("chunk_size"), py::arg("noop_flag"), py::arg("tensor_lists"), py::arg("per_tensor_beta1"),
          py::arg("per_tensor_beta2"), py::arg("per_tensor_bias_correction"), py::arg("per_tensor_eps"),
          py::arg("per_tensor_weight_decay"), py::arg("lr"), py::arg("grad_scale"), py::arg("step"), py::arg("mode"),
          py::arg("rank"), py::arg("grad_ptrs"), py::arg("param_ptrs"), py::arg("grad_multicast_ptr") = 0,
          py::arg("param_multicast_ptr") = 0, py::arg("ld_p2p_of16") = 0, py::arg("st_p2p_of16") = 0);
    m.def("maybe_cast_mt", &maybe_cast_mt, "Converting multi-tensor copy, skipped when the overflow flag is set");
    m.def("multi_tensor_lamb_compute_update_term", &multi_tensor_lamb_compute_update_term,
          "Sharded LAMB stage 1: update term (optionally with fused per-parameter norm sums)", py::arg("chunk_size"),
          py::arg("noop_flag"), py::arg("tensor_lists"), py::arg("per_tensor_beta1"), py::arg("per_tensor_beta2"),
          py::arg("per_tensor_beta3"), py::arg("per_tensor_bias_correction"), py::arg("step"),
          py::arg("per_tensor_epsilon"), py::arg("mode"), py::arg("per_tensor_decay"), py::arg("global_scale"),
          py::arg("global_grad_norm"), py::arg("max_grad_norm"), py::arg("seg_begin") = py::none(),
          py::arg("model_dtype") = py::none());
    m.def("multi_tensor_lamb_update_weights", &multi_tensor_lamb_update_weights, "Sharded LAMB stage 2: weight update");
    {
        py::object param_cls = py::module_::import("torch.nn.parameter").attr("Parameter");
        g_parameter_type = reinterpret_cast<PyTypeObject*>(param_cls.release().ptr());  // kept for the life of the process
    }
    m.def("_flatten_only", [](py::object lists) { Flat f; flatten(lists, 1, MT_MAX_DEPTH, "_flatten_only", f); return f.tl.ntensors; },
          "host-path profiling aid: validate + flatten the lists, launch nothing");
    m.def("launch_count", []() { return mt_launch_count(); }, "kernels launched by libmt_b200.so since load");
    m.def("table_cache_stats", &table_cache_stats, "device tensor-table cache counters");
    m.def("table_cache_clear", [](int device) { check(mt_table_cache_clear(device)); });
    m.def("version", []() { return std::string(mt_version()); });
}
