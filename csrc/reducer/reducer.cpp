// Gradient reducer: the C++ core behind DistributedDataParallel.
//
// It owns what must run on the autograd engine's threads without the GIL: one post-hook per
// parameter's AccumulateGrad node, per-bucket pending counters, the strict bucket launch order
// (bucket i is never launched before bucket i-1, so every rank issues the cross-GPU kernels in
// the same order and the flag protocol cannot deadlock), the record of the order in which
// gradients became ready (used to rebuild the buckets after the first iteration) and the
// end-of-backward callback.  On GPUs the reducer also LAUNCHES: given a native plan (one descriptor per
// bucket, set_native_plan) it records the ready event, makes the side stream wait for it, refreshes the
// bucket's device-side pointer table from pinned staging when a gradient moved, and calls
// b200dt_coll_launch -- all from the autograd thread, without the GIL and without touching Python.  The
// finalize callback joins the side stream, re-points .grad at the arena views and reads the collectives'
// host-mapped status word, raising when a peer timed out.  Without a plan (CPU / gloo, the NCCL opt-out,
// debug handshake, tracing) "launch" and "finalize" are two Python callbacks.
//
// Counterpart in the reference's stack: c10d::Reducer, reached from
// /root/reference/ddp_gpus.py:32,38 (spec: torch/csrc/distributed/c10d/reducer.hpp).
#include "linear.h"
#include <torch/extension.h>
#include <torch/csrc/autograd/engine.h>
#include <torch/csrc/autograd/utils/lambda_post_hook.h>
#include <c10/cuda/CUDACachingAllocator.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <cuda_runtime.h>

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


// ---- native launch plan ---------------------------------------------------------------------
// Signature of b200dt_coll_launch (include/b200dt.h); the address comes from the ctypes handle.
typedef int (*CollLaunchFn)(int, const uint64_t*, const uint64_t*, uint64_t, int, int, int, const void*,
                            const void*, int, long long, int, int, float, int, int, int*, int, void*);

#define B200_CUDA_OK(expr)                                                                         \
  do {                                                                                             \
    cudaError_t e__ = (expr);                                                                      \
    TORCH_CHECK(e__ == cudaSuccess, #expr, " failed: ", cudaGetErrorString(e__));                  \
  } while (0)

struct PinnedSlot {
  int64_t* host = nullptr;
  cudaEvent_t done = nullptr;
  bool used = false;
};

struct NativeBucket {
  int mode = 0, n_chunks = 0, tensor_bf16 = 0, wire_bf16 = 0, sig_base = 0, grid = 1, in_place = 0;
  long long stage_base = 0;
  float scale = 1.f;
  uint64_t tensors_dev = 0, chunks_dev = 0;
  int64_t nbytes = 0;
  std::vector<int64_t> param_ids, numels;
  std::vector<at::Tensor> views;          // gradient views living in the arena (may be empty)
  std::vector<uint64_t> last_ptrs;
  std::vector<PinnedSlot> ring;
  size_t ring_pos = 0;
};

struct NativePlan {
  bool enabled = false;
  CollLaunchFn fn = nullptr;
  uint64_t bufs[8] = {0}, sigs[8] = {0};
  uint64_t mc = 0, status_dev = 0;
  volatile int* status_host = nullptr;
  int rank = 0, world = 1, device = 0;
  cudaStream_t side = nullptr;
  cudaEvent_t ready_ev = nullptr, done_ev = nullptr;
  std::vector<NativeBucket> buckets;
};

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

  ~Reducer() {
    remove_hooks();
    clear_native_plan();
  }

  // plan: {"launch_fn": addr, "bufs": [...], "sigs": [...], "mc": int, "status_dev": int, "status_host": addr,
  //        "rank", "world", "device", "side_stream": cudaStream_t as int,
  //        "buckets": [{"mode","n_chunks","tensor_bf16","wire_bf16","sig_base","grid","in_place","stage_base",
  //                     "scale","tensors_dev","chunks_dev","nbytes","param_ids","numels","views"}]}
  void set_native_plan(const py::dict& plan) {
    clear_native_plan();
    NativePlan np;
    np.fn = reinterpret_cast<CollLaunchFn>(plan["launch_fn"].cast<uintptr_t>());
    auto bufs = plan["bufs"].cast<std::vector<uint64_t>>();
    auto sigs = plan["sigs"].cast<std::vector<uint64_t>>();
    np.world = plan["world"].cast<int>();
    TORCH_CHECK(np.world >= 1 && np.world <= 8 && bufs.size() == (size_t)np.world && sigs.size() == (size_t)np.world,
                "native plan: bad world / pointer tables");
    for (int r = 0; r < np.world; ++r) { np.bufs[r] = bufs[r]; np.sigs[r] = sigs[r]; }
    np.mc = plan["mc"].cast<uint64_t>();
    np.status_dev = plan["status_dev"].cast<uint64_t>();
    np.status_host = reinterpret_cast<volatile int*>(plan["status_host"].cast<uintptr_t>());
    np.rank = plan["rank"].cast<int>();
    np.device = plan["device"].cast<int>();
    np.side = reinterpret_cast<cudaStream_t>(plan["side_stream"].cast<uintptr_t>());
    c10::cuda::CUDAGuard guard(static_cast<c10::DeviceIndex>(np.device));
    B200_CUDA_OK(cudaEventCreateWithFlags(&np.ready_ev, cudaEventDisableTiming));
    B200_CUDA_OK(cudaEventCreateWithFlags(&np.done_ev, cudaEventDisableTiming));
    for (auto item : plan["buckets"].cast<py::list>()) {
      py::dict d = item.cast<py::dict>();
      NativeBucket b;
      b.mode = d["mode"].cast<int>();
      b.n_chunks = d["n_chunks"].cast<int>();
      b.tensor_bf16 = d["tensor_bf16"].cast<int>();
      b.wire_bf16 = d["wire_bf16"].cast<int>();
      b.sig_base = d["sig_base"].cast<int>();
      b.grid = d["grid"].cast<int>();
      b.in_place = d["in_place"].cast<int>();
      b.stage_base = d["stage_base"].cast<long long>();
      b.scale = d["scale"].cast<float>();
      b.tensors_dev = d["tensors_dev"].cast<uint64_t>();
      b.chunks_dev = d["chunks_dev"].cast<uint64_t>();
      b.nbytes = d["nbytes"].cast<int64_t>();
      b.param_ids = d["param_ids"].cast<std::vector<int64_t>>();
      b.numels = d["numels"].cast<std::vector<int64_t>>();
      if (d.contains("views") && !d["views"].is_none()) b.views = d["views"].cast<std::vector<at::Tensor>>();
      TORCH_CHECK(b.numels.size() == b.param_ids.size(), "native plan: numels / param_ids mismatch");
      TORCH_CHECK(b.views.empty() || b.views.size() == b.param_ids.size(), "native plan: views / param_ids mismatch");
      b.ring.resize(4);
      for (auto& sl : b.ring) {
        B200_CUDA_OK(cudaHostAlloc(reinterpret_cast<void**>(&sl.host), sizeof(int64_t) * 2 * b.param_ids.size() + 16,
                                   cudaHostAllocDefault));
        B200_CUDA_OK(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
      }
      np.buckets.push_back(std::move(b));
    }
    {
      std::lock_guard<std::mutex> lock(mu_);
      TORCH_CHECK(np.buckets.size() == buckets_.size(), "native plan must describe every bucket");
      np.enabled = true;
      native_ = std::move(np);
    }
  }

  void clear_native_plan() {
    std::lock_guard<std::mutex> lock(mu_);
    if (!native_.enabled && native_.buckets.empty()) return;
    for (auto& b : native_.buckets)
      for (auto& sl : b.ring) {
        if (sl.done) { cudaEventSynchronize(sl.done); cudaEventDestroy(sl.done); }
        if (sl.host) cudaFreeHost(sl.host);
      }
    if (native_.ready_ev) cudaEventDestroy(native_.ready_ev);
    if (native_.done_ev) cudaEventDestroy(native_.done_ev);
    native_ = NativePlan{};
  }
  bool has_native_plan() const { return native_.enabled; }

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
    native_.enabled = false;        // a new layout needs a new plan (set_native_plan)
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
    d["native_launches"] = native_launches_;
    d["native_bytes"] = native_bytes_;
    d["native_table_uploads"] = native_uploads_;
    d["native_plan"] = native_.enabled;
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
      if (native_.enabled) {
        for (int64_t b : to_launch) {
          const int64_t t = now_us();
          if (t_first_launch_ == 0) t_first_launch_ = t;
          t_last_launch_ = t;
          launch_native(static_cast<size_t>(b));
          ++launched_;
        }
        return;
      }
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

  // One bucket: ready event -> side stream -> (pointer table refresh) -> fused collective.  No GIL.
  void launch_native(size_t bi) {
    NativePlan& np = native_;
    NativeBucket& b = np.buckets[bi];
    if (np.world == 1) return;
    c10::cuda::CUDAGuard guard(static_cast<c10::DeviceIndex>(np.device));
    cudaStream_t cur = c10::cuda::getCurrentCUDAStream(static_cast<c10::DeviceIndex>(np.device)).stream();
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    B200_CUDA_OK(cudaStreamIsCapturing(cur, &cap));
    const bool capturing = cap == cudaStreamCaptureStatusActive;
    B200_CUDA_OK(cudaEventRecord(np.ready_ev, cur));
    B200_CUDA_OK(cudaStreamWaitEvent(np.side, np.ready_ev, 0));
    const size_t n = b.param_ids.size();
    bool changed = b.last_ptrs.size() != n;
    if (changed) b.last_ptrs.assign(n, 0);
    auto side_stream = c10::cuda::getStreamFromExternal(np.side, static_cast<c10::DeviceIndex>(np.device));
    for (size_t k = 0; k < n; ++k) {
      const at::Tensor& g = params_[b.param_ids[k]].grad();
      TORCH_CHECK(g.defined(), "parameter ", b.param_ids[k], " has no gradient at bucket launch");
      TORCH_CHECK(g.is_contiguous(), "parameter ", b.param_ids[k], " has a non-contiguous .grad; the fused "
                  "all-reduce reads gradients in place (make the gradient contiguous or use allreduce='nccl')");
      TORCH_CHECK(g.numel() == b.numels[k], "gradient of parameter ", b.param_ids[k], " changed its size");
      const uint64_t ptr = reinterpret_cast<uint64_t>(g.data_ptr());
      if (ptr != b.last_ptrs[k]) { b.last_ptrs[k] = ptr; changed = true; }
      if (!capturing) c10::cuda::CUDACachingAllocator::recordStream(g.storage().data_ptr(), side_stream);
    }
    if (changed) upload_table(np, b, capturing);
    const int rc = np.fn(b.mode, np.bufs, np.sigs, (b.mode == 2 || b.mode == 3) ? np.mc : 0, np.rank, np.world, 0,
                         reinterpret_cast<const void*>(b.tensors_dev), reinterpret_cast<const void*>(b.chunks_dev),
                         b.n_chunks, b.stage_base, b.tensor_bf16, b.wire_bf16, b.scale, b.sig_base, b.grid,
                         reinterpret_cast<int*>(np.status_dev), b.in_place, np.side);
    TORCH_CHECK(rc == 0, "b200dt_coll_launch failed for bucket ", bi, ": rc ", rc,
                rc > 0 ? cudaGetErrorString(static_cast<cudaError_t>(rc)) : "");
    ++native_launches_;
    native_bytes_ += b.nbytes;
  }

  void upload_table(NativePlan& np, NativeBucket& b, bool capturing) {
    const size_t n = b.param_ids.size();
    const size_t bytes = sizeof(int64_t) * 2 * n;
    PinnedSlot* slot = nullptr;
    PinnedSlot frozen;
    if (capturing) {
      // the captured memcpy node re-reads its source on every replay: give it a buffer of its own that is
      // never written again (relaxed capture mode: cudaHostAlloc is not a stream operation)
      cudaStreamCaptureMode mode = cudaStreamCaptureModeRelaxed;
      B200_CUDA_OK(cudaThreadExchangeStreamCaptureMode(&mode));
      cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&frozen.host), bytes + 16, cudaHostAllocDefault);
      cudaThreadExchangeStreamCaptureMode(&mode);
      B200_CUDA_OK(e);
      slot = &frozen;
    } else {
      for (size_t tries = 0; tries < b.ring.size(); ++tries) {
        PinnedSlot& c = b.ring[(b.ring_pos + tries) % b.ring.size()];
        if (!c.used || cudaEventQuery(c.done) == cudaSuccess) { slot = &c; b.ring_pos = (b.ring_pos + tries + 1) % b.ring.size(); break; }
      }
      if (slot == nullptr) {
        slot = &b.ring[b.ring_pos];
        b.ring_pos = (b.ring_pos + 1) % b.ring.size();
        B200_CUDA_OK(cudaEventSynchronize(slot->done));
      }
    }
    for (size_t k = 0; k < n; ++k) {
      slot->host[2 * k] = static_cast<int64_t>(b.last_ptrs[k]);
      slot->host[2 * k + 1] = b.numels[k];
    }
    B200_CUDA_OK(cudaMemcpyAsync(reinterpret_cast<void*>(b.tensors_dev), slot->host, bytes, cudaMemcpyHostToDevice, np.side));
    if (!capturing) {
      B200_CUDA_OK(cudaEventRecord(slot->done, np.side));
      slot->used = true;
    }
    ++native_uploads_;
  }

  void finalize_native() {
    NativePlan& np = native_;
    if (np.world == 1) return;
    c10::cuda::CUDAGuard guard(static_cast<c10::DeviceIndex>(np.device));
    cudaStream_t cur = c10::cuda::getCurrentCUDAStream(static_cast<c10::DeviceIndex>(np.device)).stream();
    B200_CUDA_OK(cudaEventRecord(np.done_ev, np.side));
    B200_CUDA_OK(cudaStreamWaitEvent(cur, np.done_ev, 0));
    for (auto& b : np.buckets) {
      for (size_t k = 0; k < b.views.size(); ++k) {
        at::Tensor& p = params_[b.param_ids[k]];
        const at::Tensor& g = p.grad();
        // the reduced gradients are in the arena: point .grad at them (no copy back)
        if (!g.defined() || g.data_ptr() != b.views[k].data_ptr()) p.mutable_grad() = b.views[k];
      }
    }
    if (np.status_host != nullptr && np.status_host[0] != 0) {
      const int code = np.status_host[0], where = np.status_host[1], ms = np.status_host[2];
      TORCH_CHECK(false, "rank ", np.rank, ": peer-memory collective gave up after ", ms, " ms waiting for rank ",
                  where & 0xFF, " (CTA ", where >> 8, ", status ", code,
                  "): a rank died, hung, or issued its collectives in another order. The kernel trapped instead of "
                  "reducing partial data.");
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
    if (native_.enabled) {
      finalize_native();
      return;
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
  NativePlan native_;
  int64_t native_launches_ = 0, native_bytes_ = 0, native_uploads_ = 0;
};

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "b200 gradient reducer (autograd hooks, bucket bookkeeping, launch ordering) + C++ autograd Linear";
  m.def("linear_set_fns", &b200_linear_set_fns, "addresses of b200dt_gemm_tn/nn/nt_reduce_bf16 and b200dt_relu_bwd_bias");
  m.def("linear", &b200_linear, py::arg("x"), py::arg("weight"), py::arg("wb"), py::arg("bias"), py::arg("relu"),
        py::arg("out_fp32"), "y = x W^T + b (ReLU) with a C++ backward node (no GIL in backward)");
  m.def("linear_launches", &b200_linear_launches);
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
      .def("set_native_plan", &Reducer::set_native_plan)
      .def("clear_native_plan", &Reducer::clear_native_plan)
      .def("has_native_plan", &Reducer::has_native_plan)
      .def("stats", &Reducer::stats);
}
