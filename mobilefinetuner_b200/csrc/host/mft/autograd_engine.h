// Eager autograd graph — API mirror of the reference's core/autograd_engine.h:78-143 and
// core/backward_functions.h:12-18. Same observable semantics:
//   * ops register a node at call time; backward() runs nodes in reverse topological order;
//   * nodes no gradient flowed into are skipped; leaf / retain_grad tensors get `.grad` OVERWRITTEN;
//   * the graph is dropped after every run_backward (clear_graph), or explicitly for no-grad forwards.
// Differences that are not observable: gradients are shared copy-on-write instead of cloned on first
// arrival (the clone + elementwise += of core/autograd_engine.cpp:116-140 becomes at most one fused
// out-of-place add), and accumulation runs as a CUDA kernel.
#pragma once
#include <functional>
#include <memory>
#include <unordered_map>
#include <vector>

#include "tensor.h"

namespace ops {

class BackwardFunction {
public:
    virtual ~BackwardFunction() = default;
    // Returns one gradient per registered input (nullptr = no gradient for that input).
    virtual std::vector<TensorPtr> apply(const TensorPtr& grad_output) = 0;
    // B200 extension (optional): `pending[i]` is the gradient that has already been accumulated for input i, or null. A node
    // may fold it into the gradient it returns for that input — the add then happens where the new gradient is produced
    // instead of in a separate accumulation kernel — and reports so in `absorbed[i]`; the engine replaces the pending
    // gradient by the returned one for those inputs. The default ignores `pending`.
    virtual std::vector<TensorPtr> apply_fused(const TensorPtr& grad_output, const std::vector<TensorPtr>& pending, std::vector<bool>& absorbed) {
        (void)pending; (void)absorbed;
        return apply(grad_output);
    }
    virtual bool wants_pending() const { return false; }
};
using BackwardFunctionPtr = std::shared_ptr<BackwardFunction>;

namespace autograd {

class Node;
class Edge;
using NodePtr = std::shared_ptr<Node>;
using EdgePtr = std::shared_ptr<Edge>;

class Edge {
public:
    NodePtr input_node;
    int input_idx;
    Edge(NodePtr input, int idx) : input_node(std::move(input)), input_idx(idx) {}
};

class Node {
public:
    TensorPtr tensor;
    BackwardFunctionPtr backward_fn;
    std::vector<EdgePtr> next_edges;
    int ref_count = 0;
    bool ready = false;
    explicit Node(const TensorPtr& t) : tensor(t) {}
    void add_next_edge(const EdgePtr& e) { next_edges.push_back(e); }
    void set_backward_fn(BackwardFunctionPtr fn) { backward_fn = std::move(fn); }
};

class Engine {
public:
    static Engine& instance();
    void run_backward(const std::vector<TensorPtr>& outputs, const std::vector<TensorPtr>& output_grads = {});
    NodePtr register_node(const TensorPtr& output, const std::vector<TensorPtr>& inputs, BackwardFunctionPtr backward_fn);
    NodePtr get_node(const TensorPtr& tensor);
    void clear_graph();
    void set_enabled(bool enabled) { enabled_ = enabled; }
    bool is_enabled() const { return enabled_; }
    // True when `t` takes part in the graph being built (has a producer node) or is a leaf that wants a gradient:
    // ops use it to skip gradients nobody will read.
    bool needs_grad(const TensorPtr& t) const;
    // B200 extension, off by default: `.grad` of a leaf ACCUMULATES over run_backward calls (+=) instead of being
    // overwritten (core/autograd_engine.cpp:243 overwrites, which breaks gradient accumulation — SURVEY Appendix B2).
    void set_accumulate(bool on) { accumulate_ = on; }
    bool accumulates() const { return accumulate_; }
    // B200 extension, off by default: called during run_backward as soon as the LAST consumer of a leaf that requires a
    // gradient has run, with the leaf's final gradient (null if nothing flowed into it). The data-parallel step uses it to
    // start the exchange of a finished part of the gradient bucket while the rest of backward still runs. While a hook is
    // set the adapter-gradient products are launched where they occur (no deferred group), so "final" holds for them too.
    using LeafReadyHook = std::function<void(const TensorPtr& leaf, const TensorPtr& grad)>;
    void set_leaf_ready_hook(LeafReadyHook h) { leaf_hook_ = std::move(h); }

private:
    Engine() = default;
    struct Pending { TensorPtr grad; bool owned; };
    std::unordered_map<const Tensor*, NodePtr> nodes_;
    std::unordered_map<const Tensor*, TensorPtr> keepalive_;
    std::unordered_map<const Tensor*, Pending> pending_;
    bool enabled_ = true;
    bool accumulate_ = false;
    LeafReadyHook leaf_hook_;
    void add_grad(const TensorPtr& target, const TensorPtr& g);
    std::vector<NodePtr> order_from(const std::vector<NodePtr>& roots) const;
};

}  // namespace autograd
}  // namespace ops
