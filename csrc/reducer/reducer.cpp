// Gradient reducer: the C++ core behind DistributedDataParallel.
//
// It owns what must run on the autograd engine's threads without the GIL: one post-hook per
// parameter's AccumulateGrad node, per-bucket pending counters, the strict bucket launch order
// (bucket i is never launched before bucket i-1, so every rank issues the cross-GPU kernels in
// the same order and the flag protocol cannot deadlock), the record of the order in which
// gradients became ready (used to rebuild the buckets after the first iteration) and the
// end-of-backward callback.  What a "launch" is (the fused sm_100a all-reduce on a side stream,
// or a gloo / NCCL all-reduce on CPU) is decided by the Python wrapper through two callbacks.
//
// Counterpart in the reference's stack: c10d::Reducer, reached from
// /root/reference/ddp_gpus.py:32,38 (spec: torch/csrc/distributed/c10d/reducer.hpp).
#include <torch/extension.h>
#include <torch/csrc/autograd/engine.h>
#include <torch/csrc/autograd/utils/lambda_post_hook.h>

#include <atomic>
#include <chrono>
#include <mutex>
#include <vector>

namespace py = pybind11;
using torch::autograd::variable_list;

namespace {

// Same greedy size-capped assignment as c10d::compute_bucket_assignment_by_size: tensors are
// taken in the given order, grouped per (dtype, device); a group is closed as soon as it reaches
// the current cap, and every group advances through the cap list independently.
std::pair<std::vector<std::vector<int64_t>>, std::vector<int64_t>> assign_buckets(
    const std::vector<at::Tensor>& tensors, const std::vector<int64_t>& caps,
    const std::vector<int64_t>& tensor_indices) {
  TORCH_CHECK(!caps.empty(), "bucket size caps must not be empty");
  TORCH_CHECK(tensor_indices.empty() || tensor_indices.size() == tensors.size(),
              "tensor_indices must be empty or match tensors");
  struct Open {
    std::vector<int64_t> idx;
    int64_t bytes = 0;
    size_t cap_pos = 0;
  };
  std::vector<std::pair<std::pair<int, std::string>, Open>> groups;   // insertion-ordered
  std::vector<std::vector<int64_t>> out;
  std::vector<int64_t> out_caps;
  for (size_t i = 0; i < tensors.size(); ++i) {
    const auto& t = tensors[i];
    const int64_t id = tensor_indices.empty() ? static_cast<int64_t>(i) : tensor_indices[i];
    std::pair<int, std::string> key{static_cast<int>(t.scalar_type()), t.device().str()};
    Open* g = nullptr;
    for (auto& kv : groups)
      if (kv.first == key) { g = &kv.second; break; }
    if (g == nullptr) {
      groups.emplace_back(key, Open{});
      g = &groups.back().second;
    }
    g->idx.push_back(id);
    g->bytes += t.numel() * static_cast<int64_t>(t.element_size());
    const int64_t cap = caps[g->cap_pos];
    if (g->bytes >= cap) {
      out.push_back(std::move(g->idx));
      out_caps.push_back(cap);
      g->idx.clear();
      g->bytes = 0;
      if (g->cap_pos + 1 < caps.size()) ++g->cap_pos;
    }
  }
  for (auto& kv : groups)
    if (!kv.second.idx.empty()) {
      out.push_back(std::move(kv.second.idx));
      out_caps.push_back(caps[kv.second.cap_pos]);
    }
  if (tensor_indices.empty()) {
    // deterministic across ranks: order buckets by their smallest member
    std::vector<size_t> order(out.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](size_t a, size_t b) {
      return *std::min_element(out[a].begin(), out[a].end()) <
             *std::min_element(out[b].begin(), out[b].end());
    });
    std::vector<std::vector<int64_t>> s;
    std::vector<int64_t> sc;
    for (size_t i : order) { s.push_back(std::move(out[i])); sc.push_back(out_caps[i]); }
    out.swap(s);
    out_caps.swap(sc);
  }
  return {out, out_caps};
}

class Reducer {
 public:
  Reducer(std::vector<at::Tensor> params, std::vector<std::vector<int64_t>> buckets,
          py::function on_bucket_ready, py::function on_finalize)
      : params_(std::move(params)), on_bucket_ready_(std::move(on_bucket_ready)),
        on_finalize_(std::move(on_finalize)) {
    set_buckets(std::move(buckets));
    install_hooks();
  }

  // (Re-)register the post-hooks.  AccumulateGrad nodes remember the CUDA stream that was current
  // when they were created; calling this under a side stream (after remove_hooks() released the
  // old nodes) re-creates them there, which whole-step CUDA graph capture requires: a node bound
  // to the legacy default stream cannot take part in a capture.
  void install_hooks() {
    remove_hooks();
    for (size_t i = 0; i < params_.size(); ++i) {
      auto& p = params_[i];
      TORCH_CHECK(p.requires_grad(), "every reducer parameter must require grad");
      auto acc = torch::autograd::impl::grad_accumulator(p);
      const size_t index = i;
      auto key = acc->add_post_hook(std::make_unique<torch::autograd::utils::LambdaPostHook>(
          [this, index](const variable_list& outputs, const variable_list& /*inputs*/) {
            this->autograd_hook(index);
            return outputs;
          }));
      hooks_.emplace_back(key, acc);
    }
  }

  ~Reducer() { remove_hooks(); }

  void remove_hooks() {
    for (auto& h : hooks_) h.second->del_post_hook(h.first);
    hooks_.clear();
  }

  void set_buckets(std::vector<std::vector<int64_t>> buckets) {
    std::lock_guard<std::mutex> lock(mu_);
    buckets_ = std::move(buckets);
    bucket_of_.assign(params_.size(), -1);
    size_t covered = 0;
    for (size_t b = 0; b < buckets_.size(); ++b)
      for (int64_t i : buckets_[b]) {
        TORCH_CHECK(i >= 0 && static_cast<size_t>(i) < params_.size(), "bucket index out of range");
        TORCH_CHECK(bucket_of_[i] == -1, "parameter ", i, " is in two buckets");
        bucket_of_[i] = static_cast<int64_t>(b);
        ++covered;
      }
    TORCH_CHECK(covered == params_.size(), "buckets must cover every parameter exactly once");
    pending_.assign(buckets_.size(), 0);
    expecting_ = false;
  }

  // Called by the wrapper after forward when this backward must synchronise gradients.
  void prepare_for_backward() {
    std::lock_guard<std::mutex> lock(mu_);
    TORCH_CHECK(!expecting_ || launched_ == 0 || finalized_,
                "Expected to have finished reduction in the prior iteration before starting a new one. "
                "Some parameters did not receive a gradient (find_unused_parameters is not supported).");
    for (size_t b = 0; b < buckets_.size(); ++b) pending_[b] = static_cast<int64_t>(buckets_[b].size());
    ready_.assign(params_.size(), 0);
    next_bucket_ = 0;
    launched_ = 0;
    finalized_ = false;
    expecting_ = true;
    callback_queued_ = false;
    if (record_order_) ready_order_.clear();
    t_prepare_ = now_us();
    t_first_launch_ = t_last_launch_ = 0;
  }

  void set_require_sync(bool v) { require_sync_ = v; }
  void set_record_order(bool v) { record_order_ = v; }
  std::vector<int64_t> ready_order() const { return ready_order_; }
  std::vector<std::vector<int64_t>> buckets() const { return buckets_; }
  bool finalized() const { return finalized_; }

  py::dict stats() const {
    py::dict d;
    d["iterations"] = iterations_;
    d["hooks_fired"] = hooks_fired_.load();
    d["buckets"] = buckets_.size();
    d["last_first_launch_us"] = t_first_launch_ ? t_first_launch_ - t_prepare_ : 0;
    d["last_last_launch_us"] = t_last_launch_ ? t_last_launch_ - t_prepare_ : 0;
    d["last_finalize_us"] = t_finalize_ ? t_finalize_ - t_prepare_ : 0;
    return d;
  }

 private:
  static int64_t now_us() {
    return std::chrono::duration_cast<std::chrono::microseconds>(
               std::chrono::steady_clock::now().time_since_epoch()).count();
  }

  void autograd_hook(size_t index) {
    hooks_fired_.fetch_add(1, std::memory_order_relaxed);
    std::vector<int64_t> to_launch;
    bool queue_finalize = false;
    {
      std::lock_guard<std::mutex> lock(mu_);
      if (!require_sync_ || !expecting_) return;      // no_sync(), or a backward without forward
      if (ready_[index]) {
        // a parameter used twice in the graph fires once per accumulation; count it once
        return;
      }
      ready_[index] = 1;
      if (record_order_) ready_order_.push_back(static_cast<int64_t>(index));
      const int64_t b = bucket_of_[index];
      if (--pending_[b] == 0) {
        while (next_bucket_ < buckets_.size() && pending_[next_bucket_] == 0) {
          to_launch.push_back(static_cast<int64_t>(next_bucket_));
          ++next_bucket_;
        }
      }
      if (!callback_queued_) {
        callback_queued_ = true;
        queue_finalize = true;
      }
    }
    if (queue_finalize) {
      torch::autograd::Engine::get_default_engine().queue_callback([this] { this->finalize(); });
    }
    if (!to_launch.empty()) {
      py::gil_scoped_acquire gil;
      for (int64_t b : to_launch) {
        const int64_t t = now_us();
        if (t_first_launch_ == 0) t_first_launch_ = t;
        t_last_launch_ = t;
        on_bucket_ready_(b);
        ++launched_;
      }
    }
  }

  void finalize() {
    {
      std::lock_guard<std::mutex> lock(mu_);
      if (!expecting_) return;
      TORCH_CHECK(next_bucket_ == buckets_.size(),
                  "backward finished but only ", next_bucket_, " of ", buckets_.size(),
                  " gradient buckets were ready: some parameters took no part in the loss "
                  "(find_unused_parameters is not supported).");
      expecting_ = false;
      finalized_ = true;
      ++iterations_;
      t_finalize_ = now_us();
    }
    py::gil_scoped_acquire gil;
    on_finalize_();
  }

  std::vector<at::Tensor> params_;
  py::function on_bucket_ready_, on_finalize_;
  std::vector<std::pair<uintptr_t, std::shared_ptr<torch::autograd::Node>>> hooks_;
  std::mutex mu_;
  std::vector<std::vector<int64_t>> buckets_;
  std::vector<int64_t> bucket_of_, pending_, ready_order_;
  std::vector<char> ready_;
  size_t next_bucket_ = 0;
  int64_t launched_ = 0;
  bool expecting_ = false, finalized_ = true, callback_queued_ = false;
  bool require_sync_ = true, record_order_ = false;
  int64_t iterations_ = 0;
  std::atomic<int64_t> hooks_fired_{0};
  int64_t t_prepare_ = 0, t_first_launch_ = 0, t_last_launch_ = 0, t_finalize_ = 0;
};

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "b200 gradient reducer (autograd hooks, bucket bookkeeping, launch ordering)";
  m.def("assign_buckets", &assign_buckets, py::arg("tensors"), py::arg("caps"),
        py::arg("tensor_indices") = std::vector<int64_t>{},
        "Greedy size-capped bucket assignment; index-exact with torch's "
        "_compute_bucket_assignment_by_size for dense tensors.");
  py::class_<Reducer>(m, "Reducer")
      .def(py::init<std::vector<at::Tensor>, std::vector<std::vector<int64_t>>, py::function,
                    py::function>())
      .def("prepare_for_backward", &Reducer::prepare_for_backward)
      .def("set_buckets", &Reducer::set_buckets)
      .def("set_require_sync", &Reducer::set_require_sync)
      .def("set_record_order", &Reducer::set_record_order)
      .def("ready_order", &Reducer::ready_order)
      .def("buckets", &Reducer::buckets)
      .def("finalized", &Reducer::finalized)
      .def("remove_hooks", &Reducer::remove_hooks)
      .def("install_hooks", &Reducer::install_hooks)
      .def("stats", &Reducer::stats);
}
