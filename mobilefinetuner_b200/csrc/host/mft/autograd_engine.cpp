#include <memory>
#include "autograd_engine.h"
#include "ops_internal.h"

#include <algorithm>
#include <unordered_set>

#include "mft_b200.h"
#include "ops.h"
#include "runtime.h"

namespace ops {
namespace autograd {

Engine& Engine::instance() {
    static Engine e;
    return e;
}

NodePtr Engine::get_node(const TensorPtr& t) {
    if (!t) return nullptr;
    auto it = nodes_.find(t.get());
    if (it != nodes_.end()) return it->second;
    auto n = std::make_shared<Node>(t);
    nodes_.emplace(t.get(), n);
    return n;
}

// core/autograd_engine.cpp:31-73 — every non-null input gets an edge whether or not it requires grad.
NodePtr Engine::register_node(const TensorPtr& output, const std::vector<TensorPtr>& inputs, BackwardFunctionPtr fn) {
    if (!enabled_ || !output || !fn) return nullptr;
    NodePtr out = get_node(output);
    out->set_backward_fn(std::move(fn));
    output->grad_node_ = out;
    keepalive_[output.get()] = output;
    for (size_t i = 0; i < inputs.size(); ++i) {
        const TensorPtr& in = inputs[i];
        if (!in) continue;
        NodePtr n = get_node(in);
        in->grad_node_ = n;
        keepalive_[in.get()] = in;
        out->add_next_edge(std::make_shared<Edge>(n, static_cast<int>(i)));
        ++n->ref_count;
    }
    return out;
}

bool Engine::needs_grad(const TensorPtr& t) const {
    if (!t) return false;
    if (t->requires_grad()) return true;
    auto it = nodes_.find(t.get());
    return it != nodes_.end() && it->second->backward_fn != nullptr;
}

// Reverse post-order of an iterative DFS over the input edges (consumers before producers).
std::vector<NodePtr> Engine::order_from(const std::vector<NodePtr>& roots) const {
    std::vector<NodePtr> post;
    std::unordered_set<const Node*> done, open;
    struct Frame { NodePtr n; size_t next; };
    std::vector<Frame> stack;
    for (const auto& r : roots) {
        if (!r || done.count(r.get())) continue;
        stack.push_back({r, 0});
        open.insert(r.get());
        while (!stack.empty()) {
            Frame& f = stack.back();
            if (f.next < f.n->next_edges.size()) {
                const auto& e = f.n->next_edges[f.next++];
                if (!e || !e->input_node) continue;
                const Node* c = e->input_node.get();
                if (done.count(c)) continue;
                if (open.count(c)) throw std::runtime_error("Cycle detected in computation graph");
                open.insert(c);
                stack.push_back({e->input_node, 0});
            } else {
                open.erase(f.n.get());
                done.insert(f.n.get());
                post.push_back(f.n);
                stack.pop_back();
            }
        }
    }
    std::reverse(post.begin(), post.end());
    return post;
}

// Copy-on-write accumulation: the first gradient is shared; a second arrival allocates once (a + b).
void Engine::add_grad(const TensorPtr& target, const TensorPtr& g) {
    if (!target || !g) return;
    auto it = pending_.find(target.get());
    if (it == pending_.end()) { pending_[target.get()] = Pending{g, false}; return; }
    Pending& p = it->second;
    if (p.grad->numel() != g->numel()) throw TensorError("autograd: gradient size mismatch while accumulating");
    TensorPtr a = contiguous(p.grad), b = contiguous(g);
    if (!p.owned) {
        auto sum = Tensor::make(a->shape(), a->dtype());
        mft_rt::check(mft_binary(0, a->dev_f32(), b->dev_f32(), sum->dev_f32_out(), 1, 1, a->numel(), 0, 1, mft_rt::stream()), "mft_binary(add)");
        p.grad = sum;
        p.owned = true;
    } else {
        mft_rt::check(mft_accumulate(p.grad->dev_f32_mut(), b->dev_f32(), b->numel(), mft_rt::stream()), "mft_accumulate");
    }
}

// core/autograd_engine.cpp:142-254
void Engine::run_backward(const std::vector<TensorPtr>& outputs, const std::vector<TensorPtr>& output_grads) {
    if (!enabled_) return;
    std::vector<NodePtr> roots;
    for (size_t i = 0; i < outputs.size(); ++i) {
        const TensorPtr& o = outputs[i];
        if (!o || !o->requires_grad()) continue;
        NodePtr n = o->grad_node_ ? o->grad_node_ : get_node(o);
        o->grad_node_ = n;
        TensorPtr seed;
        if (i < output_grads.size() && output_grads[i]) seed = output_grads[i];
        else if (o->numel() == 1) seed = ones(o->shape(), o->dtype(), o->device());
        else throw std::runtime_error("Gradient must be provided for non-scalar outputs");
        keepalive_[o.get()] = o;
        add_grad(o, seed);
        roots.push_back(n);
    }
    if (roots.empty()) { clear_graph(); return; }

    // adapter-gradient products are recorded while the nodes run and launched together afterwards (one grouped kernel)
    struct DeferGuard {
        DeferGuard() { internal::begin_deferred_lora_tn(); }
        ~DeferGuard() { try { internal::flush_deferred_lora_tn(); } catch (...) {} }
    };
    std::unique_ptr<DeferGuard> defer;
    if (!leaf_hook_) defer = std::make_unique<DeferGuard>();
    const std::vector<NodePtr> order = order_from(roots);
    // leaf-ready hook: consumers still to run, per leaf that wants a gradient
    std::unordered_map<const Tensor*, int> consumers_left;
    auto is_hooked_leaf = [](const EdgePtr& e) {
        return e && e->input_node && e->input_node->tensor && !e->input_node->backward_fn && e->input_node->tensor->requires_grad();
    };
    if (leaf_hook_)
        for (const auto& node : order)
            if (node && node->backward_fn)
                for (const auto& e : node->next_edges) if (is_hooked_leaf(e)) ++consumers_left[e->input_node->tensor.get()];
    auto consumers_done = [&](const NodePtr& node) {
        if (!leaf_hook_) return;
        for (const auto& e : node->next_edges) {
            if (!is_hooked_leaf(e)) continue;
            const Tensor* key = e->input_node->tensor.get();
            if (--consumers_left[key] != 0) continue;
            auto pit = pending_.find(key);
            leaf_hook_(e->input_node->tensor, pit != pending_.end() ? pit->second.grad : nullptr);
        }
    };
    for (const auto& node : order) {
        if (!node || !node->tensor || !node->backward_fn) continue;
        auto it = pending_.find(node->tensor.get());
        if (it == pending_.end()) { consumers_done(node); continue; }     // nothing flowed into this node
        Pending* cur = &it->second;                         // element addresses survive the insertions below, iterators may not
        TensorPtr g = cur->grad;
        std::vector<TensorPtr> in_grads;
        std::vector<bool> absorbed;
        if (node->backward_fn->wants_pending()) {
            // hand the node the gradients already waiting at its inputs: it may add them while producing its own
            std::vector<TensorPtr> waiting;
            for (const auto& e : node->next_edges) {
                if (!e || !e->input_node || !e->input_node->tensor || e->input_idx < 0) continue;
                if (size_t(e->input_idx) >= waiting.size()) waiting.resize(size_t(e->input_idx) + 1);
                auto pit = pending_.find(e->input_node->tensor.get());
                if (pit != pending_.end()) waiting[size_t(e->input_idx)] = pit->second.grad;
            }
            absorbed.assign(waiting.size(), false);
            in_grads = node->backward_fn->apply_fused(g, waiting, absorbed);
        } else {
            in_grads = node->backward_fn->apply(g);
        }
        for (const auto& e : node->next_edges) {
            if (!e || !e->input_node || !e->input_node->tensor) continue;
            const int idx = e->input_idx;
            if (idx >= 0 && idx < static_cast<int>(in_grads.size()) && in_grads[idx]) {
                if (size_t(idx) < absorbed.size() && absorbed[size_t(idx)])
                    pending_[e->input_node->tensor.get()] = Pending{in_grads[idx], true};   // already contains what was waiting
                else
                    add_grad(e->input_node->tensor, in_grads[idx]);
            }
        }
        // interior gradients are dead once consumed (unless retained): release them early
        if (!node->tensor->retains_grad()) cur->grad.reset();
        consumers_done(node);
    }

    if (defer) {
        internal::flush_deferred_lora_tn();  // every recorded dA / dB product, now (errors propagate; the guard then finds nothing left)
        defer.reset();
    }

    // write-back: leaves (nodes without a backward function) and retain_grad tensors; `.grad` is overwritten
    for (auto& kv : pending_) {
        if (!kv.second.grad) continue;
        auto ka = keepalive_.find(kv.first);
        if (ka == keepalive_.end() || !ka->second) continue;
        const TensorPtr& t = ka->second;
        const bool leaf = !t->grad_node_ || t->grad_node_->backward_fn == nullptr;
        if (!(leaf || t->retains_grad())) continue;
        TensorPtr g = kv.second.grad;
        if (accumulate_ && t->grad() && t->grad()->numel() == g->numel() && t->grad()->is_dense()) {   // grad += g
            TensorPtr gc = contiguous(g);
            mft_rt::check(mft_accumulate(t->grad()->dev_f32_mut(), gc->dev_f32(), gc->numel(), mft_rt::stream()), "mft_accumulate");
            continue;
        }
        // `.grad` may be scaled in place by the caller (host clip), so it must not alias another tensor's gradient
        if (!g->is_dense()) g = contiguous(g);
        else if (!kv.second.owned && g.use_count() > 2) g = g->clone();
        t->set_grad(g);
    }
    pending_.clear();
    clear_graph();
}

// core/autograd_engine.cpp:256-268
void Engine::clear_graph() {
    for (auto& kv : keepalive_)
        if (kv.second) kv.second->grad_node_.reset();
    nodes_.clear();
    keepalive_.clear();
    pending_.clear();
}

}  // namespace autograd
}  // namespace ops
