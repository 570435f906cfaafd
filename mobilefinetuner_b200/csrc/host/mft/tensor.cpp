// ops::Tensor storage + factories. Reference semantics cited per function (core/tensor.cpp).
#include "tensor.h"

#include <cstdlib>
#include <cstring>
#include <numeric>
#include <random>
#include <sstream>

#include "autograd_engine.h"
#include "mft_b200.h"
#include "ops.h"
#include "runtime.h"

namespace ops {

namespace detail {
Storage::~Storage() {
    if (dev) mft_rt::dev_free(dev, dev_bytes);
    if (host && !host_external) std::free(host);
}
}  // namespace detail

static int64_t prod(const std::vector<int64_t>& s) {
    int64_t n = 1;
    for (auto d : s) n *= d;
    return n;
}

int64_t Tensor::numel() const { return prod(shape_); }
int64_t Tensor::rows() const { return shape_.empty() ? 1 : (shape_.back() == 0 ? 0 : numel() / shape_.back()); }
size_t Tensor::dense_bytes() const { return static_cast<size_t>(numel()) * DTypeUtils::size_of(dtype_); }
uint64_t Tensor::version() const { return st_ ? st_->version : 0; }

void Tensor::allocate(bool zero, bool defer_device) {
    for (auto d : shape_)
        if (d < 0) throw TensorError("Tensor: negative dimension");
    if (ld_ == 0) ld_ = shape_.empty() ? 1 : shape_.back();
    st_ = std::make_shared<detail::Storage>();
    const size_t es = DTypeUtils::size_of(dtype_);
    st_->dev_bytes = static_cast<size_t>(rows()) * static_cast<size_t>(ld_) * es;
    st_->in_arena = mft_rt::arena_active();
    if (defer_device) return;          // tensors that start life as host data (from_blob, copy-from-host, file loaders)
    st_->dev = mft_rt::dev_alloc(st_->dev_bytes);
    st_->dev_valid = true;
    if (zero) mft_rt::memset0(st_->dev, st_->dev_bytes);
}

// Host-initialised tensors get their HBM buffer when a kernel first needs them, so the host-side components (tokenizer,
// dataset, safetensors reader) work on a machine without a GPU and a batch that is only inspected never costs an upload.
// The buffer goes where the tensor was created: outside the step arena if it was created outside of it.
void Tensor::ensure_dev() const {
    if (st_->dev) return;
    if (!st_->in_arena && mft_rt::arena_active()) {
        mft_rt::PersistentScope keep;
        st_->dev = mft_rt::dev_alloc(st_->dev_bytes);
    } else st_->dev = mft_rt::dev_alloc(st_->dev_bytes);
}

// core/tensor.cpp:191-225 — the reference zero-fills every new tensor; the public constructor keeps that.
Tensor::Tensor(const std::vector<int64_t>& shape, DType dtype, Device device)
    : shape_(shape), dtype_(dtype), device_(device) {
    allocate(true);
}

// copy-from-host constructor: host mirror holds the data, uploaded before first device use
Tensor::Tensor(const std::vector<int64_t>& shape, const void* data, DType dtype, Device device)
    : shape_(shape), dtype_(dtype), device_(device) {
    const bool from_host = data && prod(shape) > 0;
    allocate(false, from_host);
    if (data && dense_bytes() > 0) {
        st_->host = std::malloc(dense_bytes());
        std::memcpy(st_->host, data, dense_bytes());
        st_->host_valid = true;
        st_->dev_valid = false;
    }
}

// wrap-external mode (core/tensor.cpp:110-121): the caller's buffer IS the host mirror
Tensor::Tensor(const std::vector<int64_t>& shape, void* external, DType dtype, Device device, bool wrap)
    : shape_(shape), dtype_(dtype), device_(device) {
    if (!wrap) throw TensorError("Tensor: wrap_external_flag must be true for the wrapping constructor");
    allocate(false, true);
    st_->host = external;
    st_->host_external = true;
    st_->host_valid = true;
    st_->dev_valid = false;
}

Tensor::Tensor(const Tensor& o)
    : std::enable_shared_from_this<Tensor>(), shape_(o.shape_), dtype_(o.dtype_), device_(o.device_), ld_(0),
      requires_grad_(o.requires_grad_) {
    allocate(false);
    if (numel() > 0) {
        if (o.is_dense()) mft_rt::d2d(st_->dev, o.dev(), dense_bytes());
        else {
            auto d = contiguous(std::const_pointer_cast<Tensor>(o.shared_from_this()));
            mft_rt::d2d(st_->dev, d->dev(), dense_bytes());
        }
    }
}
Tensor& Tensor::operator=(const Tensor& o) {
    if (this != &o) { Tensor tmp(o); *this = std::move(tmp); }
    return *this;
}
Tensor::Tensor(Tensor&& o) noexcept { *this = std::move(o); }
Tensor& Tensor::operator=(Tensor&& o) noexcept {
    shape_ = std::move(o.shape_); dtype_ = o.dtype_; device_ = o.device_; ld_ = o.ld_; st_ = std::move(o.st_);
    requires_grad_ = o.requires_grad_; grad_ = std::move(o.grad_); grad_fn_ = std::move(o.grad_fn_);
    retain_grad_ = o.retain_grad_; grad_node_ = std::move(o.grad_node_);
    return *this;
}
Tensor::~Tensor() = default;

TensorPtr Tensor::make(const std::vector<int64_t>& shape, DType dtype, bool zero, int64_t ld) {
    auto t = std::shared_ptr<Tensor>(new Tensor());
    t->shape_ = shape; t->dtype_ = dtype; t->ld_ = ld;
    // Op results take their HBM buffer at first device access: a result that is produced as bf16 only (bf16_only_out) and
    // never read as fp32 then never occupies its fp32 buffer in the step arena.
    t->allocate(zero, /*defer_device=*/!zero);
    if (!zero) t->st_->dev_valid = true;     // nothing on the host to upload
    return t;
}

// ---------------------------------------------------------------- host <-> device coherence
void Tensor::ensure_host() const {
    if (!st_) throw TensorError("Tensor: no storage");
    if (st_->f32_pending) (void)dev();          // the data exists as bf16 only: materialise the fp32 buffer first
    if (!st_->host) {
        st_->host = std::malloc(dense_bytes() ? dense_bytes() : 4);
        if (!st_->host) throw TensorError("Tensor: host mirror allocation failed");
    }
    if (!st_->host_valid) {
        const size_t es = DTypeUtils::size_of(dtype_);
        ensure_dev();
        if (numel() > 0) {
            if (is_dense()) mft_rt::d2h(st_->host, st_->dev, dense_bytes());
            else mft_rt::d2h_2d(st_->host, shape_.back() * es, st_->dev, ld_ * es, shape_.back() * es, rows());
        }
        st_->host_valid = true;
    }
}
const void* Tensor::host_ro() const { ensure_host(); return st_->host; }
void* Tensor::host_rw() {
    ensure_host();
    st_->dev_valid = false;   // the caller may write through the pointer
    ++st_->version;
    return st_->host;
}
const void* Tensor::dev() const {
    if (!st_) throw TensorError("Tensor: no storage");
    ensure_dev();
    if (!st_->dev_valid) {
        const size_t es = DTypeUtils::size_of(dtype_);
        if (numel() > 0) {
            if (is_dense()) mft_rt::h2d(st_->dev, st_->host, dense_bytes());
            else mft_rt::h2d_2d(st_->dev, ld_ * es, st_->host, shape_.back() * es, shape_.back() * es, rows());
        }
        st_->dev_valid = true;
    }
    if (st_->f32_pending) {                      // produced as bf16 only: fill the fp32 buffer from it, once
        st_->f32_pending = false;
        if (numel() > 0)
            mft_rt::check(mft_cast_from_bf16(st_->lowp->dev, st_->lowp_ld, static_cast<float*>(st_->dev), ld_, rows(), st_->lowp_cols, mft_rt::stream()),
                          "mft_cast_from_bf16");
    }
    return st_->dev;
}
void* Tensor::dev_mut() {
    (void)dev();
    st_->host_valid = false;
    ++st_->version;
    return st_->dev;
}
void* Tensor::dev_out() {
    if (!st_) throw TensorError("Tensor: no storage");
    ensure_dev();
    st_->dev_valid = true;
    st_->host_valid = false;
    st_->f32_pending = false;
    ++st_->version;
    return st_->dev;
}

TensorPtr Tensor::transposed_shadow() const {
    if (shape_.size() != 2) throw TensorError("transposed_shadow: 2-D tensors only");
    if (!shadow_t_ || shadow_version_ != version()) {
        mft_rt::PersistentScope keep;   // survives the step arena
        const int64_t r = shape_[0], c = shape_[1];
        const int64_t ld = (r % 4 == 0) ? r : ((r + 31) / 32) * 32;   // TMA needs 16-byte row pitch
        if (!shadow_t_) shadow_t_ = Tensor::make({c, r}, dtype_, false, ld);
        mft_rt::check(mft_transpose2d(dev_f32(), ld_, shadow_t_->dev_f32_out(), ld, r, c, 1, 0, 0, mft_rt::stream()),
                      "mft_transpose2d");
        shadow_version_ = version();
    }
    return shadow_t_;
}

namespace {
std::shared_ptr<detail::Storage> raw_buffer(size_t bytes, bool persistent) {
    auto s = std::make_shared<detail::Storage>();
    s->dev_bytes = bytes;
    if (persistent) { mft_rt::PersistentScope keep; s->dev = mft_rt::dev_alloc(bytes); }
    else s->dev = mft_rt::dev_alloc(bytes);
    s->dev_valid = true;
    return s;
}
}  // namespace

Tensor::Bf16View Tensor::bf16_view() const {
    if (dtype_ != kFloat32) throw TensorError("bf16_view: float32 tensors only");
    const int64_t cols = shape_.empty() ? 1 : shape_.back(), r = rows();
    const int64_t ld = (cols + 7) / 8 * 8;
    if (bf16_ && bf16_version_ == version()) return {bf16_->dev, bf16_ld_};
    if (st_->lowp && st_->lowp_version == version() && st_->lowp_cols == cols) {      // written by the producer (maybe through an alias)
        bf16_ = st_->lowp; bf16_version_ = version(); bf16_ld_ = st_->lowp_ld;
        return {bf16_->dev, bf16_ld_};
    }
    if (!bf16_ || bf16_ == st_->lowp) bf16_ = raw_buffer(size_t(r) * ld * 2, !st_->in_arena);
    mft_rt::check(mft_cast_bf16(dev_f32(), ld_, bf16_->dev, ld, r, cols, mft_rt::stream()), "mft_cast_bf16");
    bf16_version_ = version();
    bf16_ld_ = ld;
    return {bf16_->dev, bf16_ld_};
}

Tensor::Bf16Shadows Tensor::existing_bf16_shadows() const {
    Bf16Shadows o{nullptr, 0, nullptr, 0};
    if (bf16_) { o.plain = bf16_->dev; o.ld_plain = bf16_ld_; }
    if (bf16_t_) { o.transposed = bf16_t_->dev; o.ld_transposed = bf16_t_ld_; }
    return o;
}
void Tensor::mark_bf16_shadows_current() {
    if (bf16_) bf16_version_ = version();
    if (bf16_t_) bf16_t_version_ = version();
}

bool Tensor::has_bf16() const {
    if (!st_ || dtype_ != kFloat32) return false;
    if (bf16_ && bf16_version_ == version()) return true;
    const int64_t cols = shape_.empty() ? 1 : shape_.back();
    return st_->lowp && st_->lowp_version == version() && st_->lowp_cols == cols;
}

Tensor::Bf16Out Tensor::bf16_emit() {
    if (dtype_ != kFloat32) throw TensorError("bf16_emit: float32 tensors only");
    const int64_t cols = shape_.empty() ? 1 : shape_.back(), r = rows();
    const int64_t ld = (cols + 7) / 8 * 8;
    if (!bf16_) bf16_ = raw_buffer(size_t(r) * ld * 2, !st_->in_arena);
    bf16_version_ = version();
    bf16_ld_ = ld;
    st_->lowp = bf16_; st_->lowp_ld = ld; st_->lowp_cols = cols; st_->lowp_version = version();
    return {bf16_->dev, ld};
}

Tensor::Bf16Out Tensor::bf16_only_out() {
    if (dtype_ != kFloat32) throw TensorError("bf16_only_out: float32 tensors only");
    if (!st_) throw TensorError("Tensor: no storage");
    st_->dev_valid = true;           // the device side is authoritative (as bf16, until somebody needs fp32); no fp32 buffer yet
    st_->host_valid = false;
    ++st_->version;
    Bf16Out o = bf16_emit();
    st_->f32_pending = true;
    return o;
}

Tensor::Bf16View Tensor::bf16_transposed_view() const {
    if (dtype_ != kFloat32 || shape_.size() != 2) throw TensorError("bf16_transposed_view: 2-D float32 tensors only");
    const int64_t r = shape_[0], c = shape_[1];
    const int64_t ld = (r + 7) / 8 * 8;
    if (!bf16_t_ || bf16_t_version_ != version()) {
        if (!bf16_t_) bf16_t_ = raw_buffer(size_t(c) * ld * 2, !st_->in_arena);
        mft_rt::check(mft_transpose_cast_bf16(dev_f32(), ld_, bf16_t_->dev, ld, r, c, mft_rt::stream()), "mft_transpose_cast_bf16");
        bf16_t_version_ = version();
        bf16_t_ld_ = ld;
    }
    return {bf16_t_->dev, bf16_t_ld_};
}

// ---------------------------------------------------------------- shape ops
static std::vector<int64_t> resolve_shape(const std::vector<int64_t>& want, int64_t numel) {
    std::vector<int64_t> s = want;
    int64_t known = 1; int infer = -1;
    for (size_t i = 0; i < s.size(); ++i) {
        if (s[i] == -1) { if (infer >= 0) throw TensorError("reshape: more than one -1"); infer = static_cast<int>(i); }
        else known *= s[i];
    }
    if (infer >= 0) {
        if (known == 0 || numel % known) throw TensorError("reshape: cannot infer dimension");
        s[infer] = numel / known;
    } else if (known != numel) throw TensorError("reshape: element count mismatch");
    return s;
}

// core/tensor.cpp:274-300 copies; here a reshape aliases the same HBM buffer (tensors are never mutated in
// place by the ops), falling back to a dense copy when a pitched tensor's rows change.
TensorPtr Tensor::reshape(const std::vector<int64_t>& new_shape) const {
    auto s = resolve_shape(new_shape, numel());
    const bool keeps_rows = !s.empty() && !shape_.empty() && s.back() == shape_.back();
    if (!is_dense() && !keeps_rows) {
        auto d = contiguous(std::const_pointer_cast<Tensor>(shared_from_this()));
        return d->reshape(s);
    }
    auto t = std::shared_ptr<Tensor>(new Tensor());
    t->shape_ = s; t->dtype_ = dtype_; t->device_ = device_; t->st_ = st_;
    t->ld_ = is_dense() ? (s.empty() ? 1 : s.back()) : ld_;
    t->requires_grad_ = requires_grad_;
    return t;
}
TensorPtr Tensor::view(const std::vector<int64_t>& new_shape) const { return reshape(new_shape); }

TensorPtr Tensor::squeeze(int dim) const {
    std::vector<int64_t> s;
    const int nd = static_cast<int>(shape_.size());
    if (dim >= nd || dim < -nd - 0) { if (dim != -1) throw TensorError("squeeze: bad dim"); }
    for (int i = 0; i < nd; ++i) {
        bool drop = (dim == -1) ? (shape_[i] == 1) : (i == (dim < 0 ? dim + nd : dim) && shape_[i] == 1);
        if (!drop) s.push_back(shape_[i]);
    }
    return reshape(s);
}
TensorPtr Tensor::unsqueeze(int dim) const {
    const int nd = static_cast<int>(shape_.size());
    if (dim < 0) dim += nd + 1;
    if (dim < 0 || dim > nd) throw TensorError("unsqueeze: bad dim");
    std::vector<int64_t> s = shape_;
    s.insert(s.begin() + dim, 1);
    return reshape(s);
}
TensorPtr Tensor::transpose(int d0, int d1) const {
    return ops::transpose(std::const_pointer_cast<Tensor>(shared_from_this()), d0, d1);
}

// slice / operator[] are host-side conveniences (not on the training path): staged through the host mirror.
TensorPtr Tensor::slice(int dim, int64_t start, int64_t end, int64_t step) const {
    const int nd = static_cast<int>(shape_.size());
    if (dim < 0) dim += nd;
    if (dim < 0 || dim >= nd || step <= 0) throw TensorError("slice: bad arguments");
    if (start < 0) start += shape_[dim];
    if (end < 0) end += shape_[dim];
    if (end > shape_[dim]) end = shape_[dim];
    if (start < 0 || start > end) throw TensorError("slice: bad range");
    std::vector<int64_t> s = shape_;
    s[dim] = (end - start + step - 1) / step;
    auto out = Tensor::make(s, dtype_);
    const size_t es = DTypeUtils::size_of(dtype_);
    int64_t inner = 1, outer = 1;
    for (int i = dim + 1; i < nd; ++i) inner *= shape_[i];
    for (int i = 0; i < dim; ++i) outer *= shape_[i];
    const char* src = static_cast<const char*>(host_ro());
    char* dst = static_cast<char*>(out->host_rw());
    for (int64_t o = 0; o < outer; ++o)
        for (int64_t j = 0; j < s[dim]; ++j)
            std::memcpy(dst + ((o * s[dim] + j) * inner) * es, src + ((o * shape_[dim] + start + j * step) * inner) * es, inner * es);
    return out;
}
TensorPtr Tensor::operator[](int64_t index) const {
    if (shape_.empty()) throw TensorError("operator[]: 0-d tensor");
    auto s = slice(0, index, index + 1);
    std::vector<int64_t> ns(shape_.begin() + 1, shape_.end());
    if (ns.empty()) ns.push_back(1);
    return s->reshape(ns);
}

TensorPtr Tensor::to(const Device& device) const {
    auto t = clone();
    t->device_ = device;
    return t;
}

// core/tensor.cpp:525-535 — clone/detach copy the data; neither carries autograd history.
TensorPtr Tensor::clone() const {
    auto t = Tensor::make(shape_, dtype_);
    t->device_ = device_;
    if (numel() > 0) {
        if (is_dense()) mft_rt::d2d(t->dev_out(), dev(), dense_bytes());
        else {
            auto d = contiguous(std::const_pointer_cast<Tensor>(shared_from_this()));
            mft_rt::d2d(t->dev_out(), d->dev(), dense_bytes());
        }
    }
    t->requires_grad_ = requires_grad_;
    return t;
}
TensorPtr Tensor::detach() const {
    auto t = clone();
    t->requires_grad_ = false;
    return t;
}

// core/tensor.cpp:537-562 — delegate to the engine; non-scalar roots need an explicit gradient.
void Tensor::backward(const TensorPtr& gradient) {
    if (!requires_grad_) throw TensorError("backward() called on a tensor that does not require grad");
    std::vector<TensorPtr> grads;
    if (gradient) grads.push_back(gradient);
    autograd::Engine::instance().run_backward({shared_from_this()}, grads);
}
// core/tensor.cpp:598-603 — memset the existing gradient (it stays non-null)
void Tensor::zero_grad() {
    if (grad_ && grad_->numel() > 0) mft_rt::memset0(grad_->dev_out(), grad_->st_->dev_bytes);
}

std::string Tensor::to_string() const {
    std::ostringstream os;
    os << "Tensor(shape=[";
    for (size_t i = 0; i < shape_.size(); ++i) os << (i ? ", " : "") << shape_[i];
    os << "], dtype=" << DTypeUtils::to_string(dtype_) << ", device=" << device_.to_string() << ")";
    if (dtype_ == kFloat32 && numel() > 0 && numel() <= 16) {
        const float* d = static_cast<const float*>(host_ro());
        os << " [";
        for (int64_t i = 0; i < numel(); ++i) os << (i ? ", " : "") << d[i];
        os << "]";
    }
    return os.str();
}
void Tensor::print() const { std::cout << to_string() << std::endl; }
std::ostream& operator<<(std::ostream& os, const Tensor& t) { return os << t.to_string(); }

// ---------------------------------------------------------------- factories (core/tensor.cpp:655-770)
TensorPtr zeros(const std::vector<int64_t>& shape, DType dtype, Device device) {
    return std::make_shared<Tensor>(shape, dtype, device);
}
TensorPtr empty(const std::vector<int64_t>& shape, DType dtype, Device device) { return zeros(shape, dtype, device); }
TensorPtr full(const std::vector<int64_t>& shape, float value, DType dtype, Device device) {
    if (dtype != kFloat32) {
        auto t = zeros(shape, dtype, device);
        if (dtype == kInt32) { int32_t* p = t->data<int32_t>(); for (int64_t i = 0; i < t->numel(); ++i) p[i] = static_cast<int32_t>(value); }
        return t;
    }
    auto t = Tensor::make(shape, dtype);
    mft_rt::check(mft_fill(t->dev_f32_out(), t->rows() * t->ld(), value, mft_rt::stream()), "mft_fill");
    return t;
}
TensorPtr ones(const std::vector<int64_t>& shape, DType dtype, Device device) { return full(shape, 1.0f, dtype, device); }

static std::mt19937& host_rng() {
    static std::mt19937 g{std::random_device{}()};   // non-deterministic like the reference (Appendix B15)
    return g;
}
TensorPtr randn(const std::vector<int64_t>& shape, DType dtype, Device device) {
    auto t = zeros(shape, dtype, device);
    std::normal_distribution<float> d(0.f, 1.f);
    float* p = t->data<float>();
    for (int64_t i = 0; i < t->numel(); ++i) p[i] = d(host_rng());
    return t;
}
TensorPtr rand(const std::vector<int64_t>& shape, DType dtype, Device device) {
    auto t = zeros(shape, dtype, device);
    std::uniform_real_distribution<float> d(0.f, 1.f);
    float* p = t->data<float>();
    for (int64_t i = 0; i < t->numel(); ++i) p[i] = d(host_rng());
    return t;
}
TensorPtr from_blob(void* data, const std::vector<int64_t>& shape, DType dtype, Device device) {
    return std::make_shared<Tensor>(shape, data, dtype, device, true);
}
TensorPtr tensor(const std::vector<float>& data, DType dtype, Device device) {
    return std::make_shared<Tensor>(std::vector<int64_t>{static_cast<int64_t>(data.size())}, data.data(), dtype, device);
}
TensorPtr tensor(std::initializer_list<float> data, DType dtype, Device device) {
    return tensor(std::vector<float>(data), dtype, device);
}

// ---------------------------------------------------------------- operator sugar
TensorPtr Tensor::operator+(const Tensor& o) const { return add(std::const_pointer_cast<Tensor>(shared_from_this()), std::const_pointer_cast<Tensor>(o.shared_from_this())); }
TensorPtr Tensor::operator-(const Tensor& o) const { return sub(std::const_pointer_cast<Tensor>(shared_from_this()), std::const_pointer_cast<Tensor>(o.shared_from_this())); }
TensorPtr Tensor::operator*(const Tensor& o) const { return mul(std::const_pointer_cast<Tensor>(shared_from_this()), std::const_pointer_cast<Tensor>(o.shared_from_this())); }
TensorPtr Tensor::operator/(const Tensor& o) const { return div(std::const_pointer_cast<Tensor>(shared_from_this()), std::const_pointer_cast<Tensor>(o.shared_from_this())); }
TensorPtr Tensor::operator+(float s) const { return add(std::const_pointer_cast<Tensor>(shared_from_this()), s); }
TensorPtr Tensor::operator-(float s) const { return sub(std::const_pointer_cast<Tensor>(shared_from_this()), s); }
TensorPtr Tensor::operator*(float s) const { return mul(std::const_pointer_cast<Tensor>(shared_from_this()), s); }
TensorPtr Tensor::operator/(float s) const { return div(std::const_pointer_cast<Tensor>(shared_from_this()), s); }
TensorPtr operator+(const TensorPtr& a, const TensorPtr& b) { return add(a, b); }
TensorPtr operator-(const TensorPtr& a, const TensorPtr& b) { return sub(a, b); }
TensorPtr operator*(const TensorPtr& a, const TensorPtr& b) { return mul(a, b); }
TensorPtr operator/(const TensorPtr& a, const TensorPtr& b) { return div(a, b); }
TensorPtr operator+(const TensorPtr& t, float s) { return add(t, s); }
TensorPtr operator-(const TensorPtr& t, float s) { return sub(t, s); }
TensorPtr operator*(const TensorPtr& t, float s) { return mul(t, s); }
TensorPtr operator/(const TensorPtr& t, float s) { return div(t, s); }
TensorPtr operator+(float s, const TensorPtr& t) { return add(s, t); }
TensorPtr operator-(float s, const TensorPtr& t) { return sub(s, t); }
TensorPtr operator*(float s, const TensorPtr& t) { return mul(s, t); }
TensorPtr operator/(float s, const TensorPtr& t) { return div(s, t); }

}  // namespace ops
