// ops::Tensor — API mirror of the reference's core/tensor.h:78-248 (same public names and meaning), with a
// different storage model: the buffer lives in HBM and the host pointer handed out by data<T>() is a lazily
// synchronised mirror (D2H on first host read after a kernel wrote; H2D before the next kernel after a
// non-const host access). Kernels never see host memory; the host mirror never computes.
#pragma once
#include <cassert>
#include <functional>
#include <initializer_list>
#include <iostream>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "device.h"
#include "dtype.h"

namespace ops {

class Tensor;
namespace autograd {
class Node;
class Engine;
using NodePtr = std::shared_ptr<Node>;
}  // namespace autograd

using TensorPtr = std::shared_ptr<Tensor>;
using GradFn = std::function<std::vector<TensorPtr>(const TensorPtr&)>;

class TensorError : public std::exception {
public:
    explicit TensorError(const std::string& msg) : message_(msg) {}
    const char* what() const noexcept override { return message_.c_str(); }
private:
    std::string message_;
};

namespace detail {
// One HBM allocation + optional host mirror. `rows x ld` elements when pitched (ld >= last dim).
struct Storage {
    void* dev = nullptr;
    void* host = nullptr;
    size_t dev_bytes = 0;
    bool dev_valid = false, host_valid = false;
    bool host_external = false;      // wrap mode: the caller owns `host`
    bool in_arena = false;           // allocated from the per-step arena (dies at the next arena_reset)
    uint64_t version = 0;            // bumped on every write access (device or host)
    // bf16 copy written by the kernel that produced this buffer's current contents (version lowp_version). When
    // f32_pending is set that copy is the ONLY one: the fp32 buffer has not been written and is filled from it (one upcast
    // kernel) the first time anybody asks for fp32 data. Lives here, not on the Tensor, so reshaped aliases see it too.
    std::shared_ptr<Storage> lowp;
    int64_t lowp_ld = 0, lowp_cols = 0;
    uint64_t lowp_version = ~0ull;
    bool f32_pending = false;
    ~Storage();
};
}  // namespace detail

class Tensor : public std::enable_shared_from_this<Tensor> {
public:
    Tensor() = default;
    Tensor(const std::vector<int64_t>& shape, DType dtype = kFloat32, Device device = kCPU);
    Tensor(const std::vector<int64_t>& shape, const void* data, DType dtype = kFloat32, Device device = kCPU);
    Tensor(const std::vector<int64_t>& shape, void* external_data, DType dtype, Device device, bool wrap_external_flag);
    Tensor(const Tensor& other);
    Tensor& operator=(const Tensor& other);
    Tensor(Tensor&& other) noexcept;
    Tensor& operator=(Tensor&& other) noexcept;
    ~Tensor();

    const std::vector<int64_t>& shape() const { return shape_; }
    int64_t size(int dim) const { return shape_[static_cast<size_t>(dim < 0 ? dim + static_cast<int>(shape_.size()) : dim)]; }
    int64_t ndim() const { return static_cast<int64_t>(shape_.size()); }
    int64_t numel() const;
    DType dtype() const { return dtype_; }
    Device device() const { return device_; }

    // ---- host view (synchronising). Non-const access assumes the caller may write.
    void* data_ptr() { return host_rw(); }
    const void* data_ptr() const { return host_ro(); }
    template <typename T> T* data() { return static_cast<T*>(host_rw()); }
    template <typename T> const T* data() const { return static_cast<const T*>(host_ro()); }
    template <typename T> T& item() { assert(numel() == 1); return *static_cast<T*>(host_rw()); }
    template <typename T> const T& item() const { assert(numel() == 1); return *static_cast<const T*>(host_ro()); }
    float item() const { assert(numel() == 1); return dtype_ == kFloat32 ? *static_cast<const float*>(host_ro()) : 0.0f; }

    TensorPtr to(const Device& device) const;
    TensorPtr cpu() const { return to(kCPU); }
    TensorPtr cuda() const { return to(kCUDA); }

    TensorPtr reshape(const std::vector<int64_t>& new_shape) const;
    TensorPtr view(const std::vector<int64_t>& new_shape) const;
    TensorPtr squeeze(int dim = -1) const;
    TensorPtr unsqueeze(int dim) const;
    TensorPtr transpose(int dim0, int dim1) const;
    TensorPtr slice(int dim, int64_t start, int64_t end, int64_t step = 1) const;
    TensorPtr operator[](int64_t index) const;

    bool requires_grad() const { return requires_grad_; }
    void set_requires_grad(bool r) { requires_grad_ = r; }
    TensorPtr grad() const { return grad_; }
    void set_grad(const TensorPtr& g) { grad_ = g; }
    bool is_leaf() const { return grad_fn_ == nullptr; }
    void backward(const TensorPtr& gradient = nullptr);
    void zero_grad();
    void set_grad_fn(const GradFn& fn) { grad_fn_ = fn; }
    void retain_grad() { retain_grad_ = true; }
    bool retains_grad() const { return retain_grad_; }
    TensorPtr clone() const;
    TensorPtr detach() const;

    void print() const;
    std::string to_string() const;

    TensorPtr operator+(const Tensor& other) const;
    TensorPtr operator-(const Tensor& other) const;
    TensorPtr operator*(const Tensor& other) const;
    TensorPtr operator/(const Tensor& other) const;
    TensorPtr operator+(float scalar) const;
    TensorPtr operator-(float scalar) const;
    TensorPtr operator*(float scalar) const;
    TensorPtr operator/(float scalar) const;

    // ---- device view (B200 extension; used by the op implementations, never by the trainers)
    const void* dev() const;                 // read access: uploads the host mirror first if it is newer
    void* dev_mut();                         // write access: additionally invalidates the host mirror
    void* dev_out();                         // write-only access: no upload of stale host data
    const float* dev_f32() const { return static_cast<const float*>(dev()); }
    float* dev_f32_mut() { return static_cast<float*>(dev_mut()); }
    float* dev_f32_out() { return static_cast<float*>(dev_out()); }
    int64_t ld() const { return ld_; }       // leading dimension (elements) of the last axis; == shape.back() when dense
    bool is_dense() const { return shape_.empty() || ld_ == shape_.back(); }
    int64_t rows() const;                    // numel / shape.back()
    uint64_t version() const;                // changes whenever the contents may have changed
    bool shares_storage(const Tensor& o) const { return st_ == o.st_; }
    // Uninitialised device tensor (the reference zero-fills every allocation, core/tensor.cpp:215-218;
    // ops here overwrite their outputs completely, so the memset is skipped unless `zero`).
    static TensorPtr make(const std::vector<int64_t>& shape, DType dtype = kFloat32, bool zero = false, int64_t ld = 0);
    // Cached K-major shadow of a 2-D weight (its transpose), rebuilt when version() changes.
    TensorPtr transposed_shadow() const;
    // Cached bf16 copies for the bf16 tensor-core GEMM (device pointer + leading dimension in elements): same layout
    // viewed as [rows(), shape.back()], or the transpose of a 2-D tensor. Rebuilt when version() changes.
    struct Bf16View { const void* ptr; int64_t ld; };
    Bf16View bf16_view() const;
    Bf16View bf16_transposed_view() const;
    // For producers that write the bf16 copy themselves: call AFTER dev_*_out(); returns the shadow buffer (allocating it
    // on first use) and marks it valid for the tensor's current contents, so the next bf16_view() needs no cast kernel.
    struct Bf16Out { void* ptr; int64_t ld; };
    Bf16Out bf16_emit();
    // For producers whose result is only ever needed as a bf16 GEMM / attention operand: like bf16_emit(), but the fp32
    // buffer is NOT written (do not call dev_*_out()). Anything that later asks for the fp32 data (dev(), data<T>()) gets it
    // through one upcast kernel, so correctness never depends on the guess; only the saved HBM traffic does.
    Bf16Out bf16_only_out();
    bool f32_pending() const { return st_ && st_->f32_pending; }
    // For a refresh kernel that rewrites the bf16 shadows of a trained 2-D tensor itself (mft_refresh_bf16_shadows): the
    // shadow buffers that exist already (null = that shadow has never been asked for), and a way to declare them current.
    struct Bf16Shadows { void* plain; int64_t ld_plain; void* transposed; int64_t ld_transposed; };
    Bf16Shadows existing_bf16_shadows() const;
    void mark_bf16_shadows_current();
    // true when bf16_view() costs nothing (a valid bf16 copy of the current contents exists)
    bool has_bf16() const;
    // hint set by the producer of an activation whose gradient consumer reads bf16 (the GELU backward kernel): the op that
    // computes that gradient may then write it as bf16 only
    void set_grad_may_be_bf16(bool v) { grad_may_be_bf16_ = v; }
    bool grad_may_be_bf16() const { return grad_may_be_bf16_; }

private:
    std::vector<int64_t> shape_;
    DType dtype_ = kFloat32;
    Device device_ = kCPU;
    int64_t ld_ = 0;
    std::shared_ptr<detail::Storage> st_;

    bool requires_grad_ = false;
    TensorPtr grad_ = nullptr;
    GradFn grad_fn_ = nullptr;
    bool retain_grad_ = false;
    bool grad_may_be_bf16_ = false;
    autograd::NodePtr grad_node_ = nullptr;

    mutable TensorPtr shadow_t_;
    mutable uint64_t shadow_version_ = ~0ull;
    mutable std::shared_ptr<detail::Storage> bf16_, bf16_t_;
    mutable uint64_t bf16_version_ = ~0ull, bf16_t_version_ = ~0ull;
    mutable int64_t bf16_ld_ = 0, bf16_t_ld_ = 0;

    void allocate(bool zero, bool defer_device = false);
    void ensure_dev() const;   // device memory of a host-initialised tensor is allocated at its first device use
    void* host_rw();
    const void* host_ro() const;
    void ensure_host() const;
    size_t dense_bytes() const;

    friend class autograd::Engine;
    friend class autograd::Node;
};

TensorPtr zeros(const std::vector<int64_t>& shape, DType dtype = kFloat32, Device device = kCPU);
TensorPtr ones(const std::vector<int64_t>& shape, DType dtype = kFloat32, Device device = kCPU);
TensorPtr randn(const std::vector<int64_t>& shape, DType dtype = kFloat32, Device device = kCPU);
TensorPtr rand(const std::vector<int64_t>& shape, DType dtype = kFloat32, Device device = kCPU);
TensorPtr empty(const std::vector<int64_t>& shape, DType dtype = kFloat32, Device device = kCPU);
TensorPtr full(const std::vector<int64_t>& shape, float value, DType dtype = kFloat32, Device device = kCPU);
TensorPtr from_blob(void* data, const std::vector<int64_t>& shape, DType dtype = kFloat32, Device device = kCPU);
TensorPtr tensor(const std::vector<float>& data, DType dtype = kFloat32, Device device = kCPU);
TensorPtr tensor(std::initializer_list<float> data, DType dtype = kFloat32, Device device = kCPU);
TensorPtr cast(const TensorPtr& tensor, DType dtype);

TensorPtr operator+(const TensorPtr& a, const TensorPtr& b);
TensorPtr operator-(const TensorPtr& a, const TensorPtr& b);
TensorPtr operator*(const TensorPtr& a, const TensorPtr& b);
TensorPtr operator/(const TensorPtr& a, const TensorPtr& b);
TensorPtr operator+(const TensorPtr& t, float s);
TensorPtr operator-(const TensorPtr& t, float s);
TensorPtr operator*(const TensorPtr& t, float s);
TensorPtr operator/(const TensorPtr& t, float s);
TensorPtr operator+(float s, const TensorPtr& t);
TensorPtr operator-(float s, const TensorPtr& t);
TensorPtr operator*(float s, const TensorPtr& t);
TensorPtr operator/(float s, const TensorPtr& t);
std::ostream& operator<<(std::ostream& os, const Tensor& t);

}  // namespace ops
