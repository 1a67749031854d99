// hostaccel.cpp — native fast path of the PYTHON host mirror (agilerl_b200.components).
//
// Not part of the C-ABI: libb200xp.so stays torch-free.  This is a small CPython/torch extension
// that does, in one call, what the Python host does in ~15 interpreter round trips per buffer
// method: allocate the fresh output tensors (replay_buffer.py:126 `storage[idx]` copies), stage
// the host inputs through a pinned ring, and call the C-ABI entry points on torch's current
// stream.  The Python classes use it when it is built and the inputs are in canonical form, and
// fall back to their own (identical, slower) code otherwise; results are the same either way —
// every byte is still produced by libb200xp.so kernels.
#include <ATen/CPUGeneratorImpl.h>
#include <ATen/cuda/CUDAEvent.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/extension.h>

#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "b200xp.h"

namespace py = pybind11;

namespace {

using gather_fn = xp_status_t (*)(const xp_key *, int, const int64_t *, int64_t, int64_t, int *, void *);
using sample_fn = xp_status_t (*)(const xp_tree *, const float *, int64_t, int64_t, int64_t, double,
                                  int64_t *, float *, void *);
using update_fn = xp_status_t (*)(const xp_tree *, xp_per_state *, const int64_t *, const float *,
                                  int64_t, int64_t, double, int *, void *);
using sample_hostu_fn = xp_status_t (*)(const xp_tree *, const float *, int64_t, int64_t, double, int64_t *, float *,
                                        void *);
using copy_fn = xp_status_t (*)(void *, const void *, int64_t, void *);
using peer_fn = xp_status_t (*)(const void *const *, const void *const *, void *const *, int, int, uint32_t *,
                                const xp_peer_seg *, int, int *, void *);

using peer_pub_fn = xp_status_t (*)(const void *const *, const void *const *, void *const *, int, int, uint32_t *,
                                    const xp_peer_seg *, int, const void *const *, int, const float *, int *, void *);

struct Fns {
  gather_fn gather = nullptr;
  sample_fn per_sample = nullptr;
  update_fn per_update = nullptr;
  copy_fn copy_to_device = nullptr;
  peer_fn peer_allgather_graph = nullptr;
  sample_hostu_fn per_sample_hostu = nullptr;  // host-vector variants (B <= XP_HOST_VEC_MAX): no staging copy
  update_fn per_update_hostp = nullptr;
  peer_pub_fn peer_publish_allgather_graph = nullptr;  // block written by the exchange kernel itself (<= 32 KB)
  // agilerl_b200._native.check(status, what): raises the mapped exception.  A borrowed handle whose
  // reference is deliberately never released: this struct is destroyed by the C runtime's exit
  // handlers, after the interpreter has been finalised, where a Py_DECREF would touch freed memory.
  py::handle check;
} g;

void bind_peer(uint64_t peer_allgather_graph) {
  g.peer_allgather_graph = reinterpret_cast<peer_fn>(peer_allgather_graph);
}

void bind_host_vec(uint64_t per_sample_hostu, uint64_t per_update_hostp, uint64_t peer_publish_allgather_graph) {
  g.per_sample_hostu = reinterpret_cast<sample_hostu_fn>(per_sample_hostu);
  g.per_update_hostp = reinterpret_cast<update_fn>(per_update_hostp);
  g.peer_publish_allgather_graph = reinterpret_cast<peer_pub_fn>(peer_publish_allgather_graph);
}

void bind(uint64_t gather, uint64_t per_sample, uint64_t per_update, uint64_t copy_to_device,
          py::object check) {
  g.gather = reinterpret_cast<gather_fn>(gather);
  g.per_sample = reinterpret_cast<sample_fn>(per_sample);
  g.per_update = reinterpret_cast<update_fn>(per_update);
  g.copy_to_device = reinterpret_cast<copy_fn>(copy_to_device);
  g.check = check.release();  // leaked on purpose, see Fns::check
}

inline void ok(xp_status_t rc, const char *what) {
  if (rc != XP_OK) g.check((int)rc, what);  // raises in Python
}

// ---- pinned staging ring (same protocol as components/staging.py) ------------------------------
struct Slot {
  at::Tensor host, dev;
  char *hp = nullptr, *dp = nullptr;
  int64_t nbytes = 0;
};

class Ring {
 public:
  Ring(c10::DeviceIndex index, int slots) : index_(index), slots_(slots), half_(slots / 2) {}

  Slot &acquire(int64_t nbytes) {
    const int i = next_;
    if (i % half_ == 0) {
      const int h = i / half_;
      if (pending_[h]) {
        if (!ev_[h].query()) ev_[h].synchronize();
        pending_[h] = false;
      }
    }
    Slot &s = slots_[i];
    if (s.nbytes < nbytes) {
      if (s.nbytes) c10::cuda::getCurrentCUDAStream(index_).synchronize();
      const int64_t n = std::max<int64_t>(nbytes, 1 << 16);
      s.host = at::empty({n}, at::TensorOptions().dtype(at::kByte).pinned_memory(true));
      s.dev = at::empty({n}, at::TensorOptions().dtype(at::kByte).device(c10::Device(c10::kCUDA, index_)));
      s.hp = static_cast<char *>(s.host.data_ptr());
      s.dp = static_cast<char *>(s.dev.data_ptr());
      s.nbytes = n;
    }
    return s;
  }

  // host -> device of the first nbytes of the acquired slot, on the current stream
  void push(Slot &s, int64_t nbytes, c10::cuda::CUDAStream st) {
    if (have_last_ && st.stream() != last_) c10::cuda::device_synchronize();
    last_ = st.stream();
    have_last_ = true;
    ok(g.copy_to_device(s.dp, s.hp, nbytes, st.stream()), "xp_copy_to_device");
    advance(st);
  }

  // the slot just acquired has been consumed by a copy the caller enqueued on `st` itself
  void advance(c10::cuda::CUDAStream st) {
    const int i = next_;
    next_ = (i + 1) % (int)slots_.size();
    if ((i + 1) % half_ == 0) {
      const int h = i / half_;
      ev_[h].record(st);
      pending_[h] = true;
    }
  }

 private:
  c10::DeviceIndex index_;
  std::vector<Slot> slots_;
  int half_, next_ = 0;
  at::cuda::CUDAEvent ev_[2];
  bool pending_[2] = {false, false};
  cudaStream_t last_ = nullptr;
  bool have_last_ = false;
};

// ---- storage[idx] for every key of a table ------------------------------------------------------
struct KeyMeta {
  std::string name;
  std::vector<int64_t> tail;
  at::ScalarType dtype;
};

class FastGather {
 public:
  FastGather(std::vector<at::Tensor> leaves, std::vector<std::string> names, int64_t max_size,
             at::Tensor err)
      : leaves_(std::move(leaves)), max_size_(max_size), err_(std::move(err)) {
    TORCH_CHECK(leaves_.size() == names.size() && !leaves_.empty() && leaves_.size() <= XP_MAX_KEYS);
    index_ = leaves_[0].device().index();
    keys_.resize(leaves_.size());
    for (size_t i = 0; i < leaves_.size(); ++i) {
      const at::Tensor &t = leaves_[i];
      TORCH_CHECK(t.is_cuda() && t.is_contiguous() && t.dim() >= 1);
      keys_[i].src = t.data_ptr();
      keys_[i].dst = nullptr;
      int64_t per_row = 1;
      for (int64_t d = 1; d < t.dim(); ++d) per_row *= t.size(d);
      keys_[i].row_bytes = per_row * (int64_t)t.element_size();
      KeyMeta m;
      m.name = names[i];
      m.tail.assign(t.sizes().begin() + 1, t.sizes().end());
      m.dtype = t.scalar_type();
      meta_.push_back(std::move(m));
    }
  }

  // idx: int64 CUDA contiguous; lead: leading shape of the outputs (numel(lead) == idx.numel())
  py::dict gather(const at::Tensor &idx, const std::vector<int64_t> &lead) {
    py::dict out;
    gather_into(out, idx.data_ptr<int64_t>(), idx.numel(), lead, c10::cuda::getCurrentCUDAStream(index_));
    return out;
  }

  void gather_into(py::dict &out, const int64_t *idx, int64_t B, const std::vector<int64_t> &lead,
                   c10::cuda::CUDAStream st) {
    const auto dev = c10::Device(c10::kCUDA, index_);
    std::vector<int64_t> shape;
    for (size_t i = 0; i < keys_.size(); ++i) {
      shape.assign(lead.begin(), lead.end());
      shape.insert(shape.end(), meta_[i].tail.begin(), meta_[i].tail.end());
      at::Tensor o = at::empty(shape, at::TensorOptions().dtype(meta_[i].dtype).device(dev));
      keys_[i].dst = o.data_ptr();
      out[py::str(meta_[i].name)] = o;
    }
    if (B > 0)
      ok(g.gather(keys_.data(), (int)keys_.size(), idx, B, max_size_, err_.data_ptr<int>(), st.stream()),
         "xp_gather");
  }

  c10::DeviceIndex index() const { return index_; }
  int64_t max_size() const { return max_size_; }
  size_t n_keys() const { return keys_.size(); }

  // Allocate this table's fresh outputs into `out` and append its keys (dst filled in) to `keys`: lets a
  // caller move the rows of TWO tables (PER + n-step, same indices) with ONE xp_gather launch.
  void stage_keys(py::dict &out, std::vector<xp_key> &keys, const std::vector<int64_t> &lead) {
    const auto dev = c10::Device(c10::kCUDA, index_);
    std::vector<int64_t> shape;
    for (size_t i = 0; i < keys_.size(); ++i) {
      shape.assign(lead.begin(), lead.end());
      shape.insert(shape.end(), meta_[i].tail.begin(), meta_[i].tail.end());
      at::Tensor o = at::empty(shape, at::TensorOptions().dtype(meta_[i].dtype).device(dev));
      xp_key k = keys_[i];
      k.dst = o.data_ptr();
      keys.push_back(k);
      out[py::str(meta_[i].name)] = o;
    }
  }
  int *err_ptr() { return err_.data_ptr<int>(); }

 private:
  std::vector<at::Tensor> leaves_;
  std::vector<xp_key> keys_;
  std::vector<KeyMeta> meta_;
  int64_t max_size_;
  at::Tensor err_;
  c10::DeviceIndex index_;
};

// ---- PrioritizedReplayBuffer.sample / update_priorities (replay_buffer.py:331-428) --------------
class FastPer {
 public:
  FastPer(std::shared_ptr<FastGather> table, uint64_t tree_struct_addr, at::Tensor nodes,
          at::Tensor aux, at::Tensor state, at::Tensor tree_err, int64_t max_size, double alpha)
      : table_(std::move(table)), nodes_(std::move(nodes)), aux_(std::move(aux)),
        state_(std::move(state)), tree_err_(std::move(tree_err)), max_size_(max_size), alpha_(alpha),
        ring_(table_->index(), 32) {
    std::memcpy(&tree_, reinterpret_cast<const void *>(tree_struct_addr), sizeof(xp_tree));
  }

  // sample(batch_size, beta) with the torch CPU generator's draws (u == None) or given fp32
  // uniforms already on the device.  Returns {keys..., "weights" (B,1) f32, "idxs" (B,1) i64}.
  py::dict sample(int64_t B, double beta, int64_t size, const c10::optional<at::Tensor> &u_dev) {
    const auto index = table_->index();
    const auto dev = c10::Device(c10::kCUDA, index);
    auto st = c10::cuda::getCurrentCUDAStream(index);
    const float *u = nullptr;
    float u_host[XP_HOST_VEC_MAX];
    const bool in_launch = !u_dev.has_value() && B <= XP_HOST_VEC_MAX && g.per_sample_hostu;
    if (u_dev.has_value()) {
      u = u_dev->data_ptr<float>();
    } else if (in_launch) {
      // one agent's batch: the draws travel inside the launch as a kernel parameter (xp_per_sample_hostu)
      auto *gen = at::get_generator_or_default<at::CPUGeneratorImpl>(c10::nullopt, at::detail::getDefaultCPUGenerator());
      std::lock_guard<std::mutex> lock(gen->mutex_);
      for (int64_t i = 0; i < B; ++i) u_host[i] = (float)(gen->random() & ((1u << 24) - 1u)) * (1.0f / 16777216.0f);
    } else {
      // batch_size x torch.rand(1).item() == torch.rand(batch_size) == uniform_(0,1) on the
      // default CPU generator (replay_buffer.py:377), drawn straight into pinned memory
      Slot &s = ring_.acquire(B * 4);
      {
        // what at::uniform_real_distribution<float> does per element (ATen/core/DistributionsHelper.h): one 32-bit
        // word of the generator, low 24 bits scaled by 2^-24 — without a temporary tensor and a dispatcher trip
        auto *gen = at::get_generator_or_default<at::CPUGeneratorImpl>(c10::nullopt, at::detail::getDefaultCPUGenerator());
        std::lock_guard<std::mutex> lock(gen->mutex_);
        float *h = reinterpret_cast<float *>(s.hp);
        for (int64_t i = 0; i < B; ++i) h[i] = (float)(gen->random() & ((1u << 24) - 1u)) * (1.0f / 16777216.0f);
      }
      ring_.push(s, B * 4, st);
      u = reinterpret_cast<const float *>(s.dp);
    }
    at::Tensor idx = at::empty({B, 1}, at::TensorOptions().dtype(at::kLong).device(dev));
    at::Tensor w = at::empty({B, 1}, at::TensorOptions().dtype(at::kFloat).device(dev));
    if (in_launch)
      ok(g.per_sample_hostu(&tree_, u_host, B, size, beta, idx.data_ptr<int64_t>(), w.data_ptr<float>(), st.stream()),
         "xp_per_sample_hostu");
    else
      ok(g.per_sample(&tree_, u, B, 1, size, beta, idx.data_ptr<int64_t>(), w.data_ptr<float>(), st.stream()),
         "xp_per_sample");
    py::dict out;
    if (sibling_) {
      // linked n-step table (same capacity, index-aligned rows, train_off_policy.py:341-343): the rows of BOTH
      // tables leave in ONE xp_gather launch; the sibling's batch is handed back under "__sibling__" and the
      // Python host serves MultiStepReplayBuffer.sample_from_indices(batch["idxs"]) from it
      std::vector<xp_key> keys;
      keys.reserve(table_->n_keys() + sibling_->n_keys());
      py::dict sib;
      table_->stage_keys(out, keys, {B});
      sibling_->stage_keys(sib, keys, {B, 1});  // storage[idxs] with (B,1) indices -> [B,1,...] (replay_buffer.py:204)
      if (B > 0)
        ok(g.gather(keys.data(), (int)keys.size(), idx.data_ptr<int64_t>(), B, max_size_, table_->err_ptr(),
                    st.stream()), "xp_gather");
      out["__sibling__"] = sib;
    } else {
      table_->gather_into(out, idx.data_ptr<int64_t>(), B, {B}, st);
    }
    out["weights"] = w;
    out["idxs"] = idx;
    return out;
  }

  // link / unlink the n-step table whose rows are gathered together with this buffer's
  void set_sibling(std::shared_ptr<FastGather> sib) {
    if (sib) {
      TORCH_CHECK(sib->max_size() == table_->max_size() && sib->index() == table_->index(),
                  "linked buffers need the same capacity and device");
      TORCH_CHECK(sib->n_keys() + table_->n_keys() <= XP_MAX_KEYS, "too many keys for one gather launch");
    }
    sibling_ = std::move(sib);
  }

  // Replicated shared memory (one process per GPU): this rank's (idx, priority) pairs go into its peer-mapped
  // block, ONE kernel exchanges every rank's block over NVLink (device-resident epoch), and the update applies
  // all ranks' pairs in rank-major order — population.ReplicatedReplay.update_priorities in one native call.
  // block0/1: this rank's two blocks (device addresses); layout of a block: int64 idx[n] | float prio[n].
  void set_exchange(std::vector<uint64_t> bases0, std::vector<uint64_t> bases1, std::vector<uint64_t> flags,
                    int rank, at::Tensor epoch_dev, at::Tensor all_idx, at::Tensor all_prio, at::Tensor px_err,
                    int64_t n) {
    world_ = (int)bases0.size();
    TORCH_CHECK(world_ >= 1 && world_ <= XP_MAX_PEERS && bases1.size() == bases0.size() && flags.size() == bases0.size());
    for (int r = 0; r < world_; ++r) {
      xb0_[r] = reinterpret_cast<const void *>(bases0[r]);
      xb1_[r] = reinterpret_cast<const void *>(bases1[r]);
      xfl_[r] = reinterpret_cast<void *>(flags[r]);
    }
    rank_ = rank;
    x_epoch_ = std::move(epoch_dev);
    x_idx_ = std::move(all_idx);
    x_prio_ = std::move(all_prio);
    x_err_ = std::move(px_err);
    x_n_ = n;
    segs_[0] = {0, x_idx_.data_ptr(), 8 * n};
    segs_[1] = {8 * n, x_prio_.data_ptr(), 4 * n};
  }

  // `k`: the block the next exchange publishes (PeerExchange.current); idx int64 CUDA contiguous [n]; prio fp32
  // on the host (pinned or not) or the device.
  void update_replicated(int k, const at::Tensor &idx, const at::Tensor &prio) {
    TORCH_CHECK(world_ > 0 && g.peer_allgather_graph, "exchange not set");
    TORCH_CHECK(idx.numel() == x_n_ && prio.numel() == x_n_, "replicated update: every rank passes ", x_n_, " pairs");
    const auto index = table_->index();
    auto st = c10::cuda::getCurrentCUDAStream(index);
    if (!prio.is_cuda() && x_n_ <= XP_HOST_VEC_MAX && g.peer_publish_allgather_graph) {
      // one agent's batch: the exchange kernel writes this rank's block itself — the indices from the device tensor,
      // the priorities from the host vector inside the launch — so two copy calls fall away
      const void *src[2] = {idx.data_ptr<int64_t>(), nullptr};
      ok(g.peer_publish_allgather_graph(xb0_, xb1_, xfl_, rank_, world_, reinterpret_cast<uint32_t *>(x_epoch_.data_ptr()),
                                        segs_, 2, src, 1, prio.data_ptr<float>(), x_err_.data_ptr<int>(), st.stream()),
         "xp_peer_publish_allgather_graph");
      ok(g.per_update(&tree_, reinterpret_cast<xp_per_state *>(state_.data_ptr()), x_idx_.data_ptr<int64_t>(),
                      x_prio_.data_ptr<float>(), x_n_ * world_, max_size_, alpha_, tree_err_.data_ptr<int>(), st.stream()),
         "xp_per_update");
      return;
    }
    char *blk = const_cast<char *>(static_cast<const char *>((k ? xb1_ : xb0_)[rank_]));
    C10_CUDA_CHECK(cudaMemcpyAsync(blk, idx.data_ptr<int64_t>(), (size_t)x_n_ * 8, cudaMemcpyDeviceToDevice, st.stream()));
    if (prio.is_cuda()) {
      C10_CUDA_CHECK(cudaMemcpyAsync(blk + 8 * x_n_, prio.data_ptr<float>(), (size_t)x_n_ * 4, cudaMemcpyDeviceToDevice,
                                     st.stream()));
    } else {
      Slot &s = ring_.acquire(x_n_ * 4);
      std::memcpy(s.hp, prio.data_ptr<float>(), (size_t)x_n_ * 4);
      // straight from pinned memory into the peer-mapped block
      if (have_ring_stream_ && st.stream() != ring_stream_) c10::cuda::device_synchronize();
      ring_stream_ = st.stream();
      have_ring_stream_ = true;
      C10_CUDA_CHECK(cudaMemcpyAsync(blk + 8 * x_n_, s.hp, (size_t)x_n_ * 4, cudaMemcpyHostToDevice, st.stream()));
      ring_.advance(st);
    }
    ok(g.peer_allgather_graph(xb0_, xb1_, xfl_, rank_, world_, reinterpret_cast<uint32_t *>(x_epoch_.data_ptr()), segs_,
                              2, x_err_.data_ptr<int>(), st.stream()), "xp_peer_allgather_graph");
    ok(g.per_update(&tree_, reinterpret_cast<xp_per_state *>(state_.data_ptr()), x_idx_.data_ptr<int64_t>(),
                    x_prio_.data_ptr<float>(), x_n_ * world_, max_size_, alpha_, tree_err_.data_ptr<int>(), st.stream()),
       "xp_per_update");
  }

  // update_priorities(indices, priorities): idx int64 CUDA contiguous; prio fp32 contiguous on the
  // host (staged through the pinned ring) or on the device.
  void update(const at::Tensor &idx, const at::Tensor &prio) {
    const auto index = table_->index();
    auto st = c10::cuda::getCurrentCUDAStream(index);
    const int64_t n = std::min<int64_t>(idx.numel(), prio.numel());
    const float *p = nullptr;
    if (prio.is_cuda()) {
      p = prio.data_ptr<float>();
    } else if (n <= XP_HOST_VEC_MAX && g.per_update_hostp) {
      // one agent's batch: the priorities travel inside the launch (xp_per_update_hostp)
      ok(g.per_update_hostp(&tree_, reinterpret_cast<xp_per_state *>(state_.data_ptr()), idx.data_ptr<int64_t>(),
                            prio.data_ptr<float>(), n, max_size_, alpha_, tree_err_.data_ptr<int>(), st.stream()),
         "xp_per_update_hostp");
      return;
    } else {
      Slot &s = ring_.acquire(n * 4);
      std::memcpy(s.hp, prio.data_ptr<float>(), (size_t)n * 4);
      ring_.push(s, n * 4, st);
      p = reinterpret_cast<const float *>(s.dp);
    }
    ok(g.per_update(&tree_, reinterpret_cast<xp_per_state *>(state_.data_ptr()), idx.data_ptr<int64_t>(), p, n,
                    max_size_, alpha_, tree_err_.data_ptr<int>(), st.stream()),
       "xp_per_update");
  }

 private:
  std::shared_ptr<FastGather> table_;
  xp_tree tree_;
  at::Tensor nodes_, aux_, state_, tree_err_;  // keep the device memory behind tree_ alive
  int64_t max_size_;
  double alpha_;
  Ring ring_;
  std::shared_ptr<FastGather> sibling_;
  // replicated-update exchange (set_exchange)
  const void *xb0_[XP_MAX_PEERS] = {nullptr}, *xb1_[XP_MAX_PEERS] = {nullptr};
  void *xfl_[XP_MAX_PEERS] = {nullptr};
  xp_peer_seg segs_[2];
  at::Tensor x_epoch_, x_idx_, x_prio_, x_err_;
  int world_ = 0, rank_ = 0;
  int64_t x_n_ = 0;
  cudaStream_t ring_stream_ = nullptr;
  bool have_ring_stream_ = false;
};

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "native fast path of the agilerl_b200 Python host (calls libb200xp.so)";
  m.def("bind", &bind);
  m.def("bind_peer", &bind_peer);
  m.def("bind_host_vec", &bind_host_vec);
  py::class_<FastGather, std::shared_ptr<FastGather>>(m, "FastGather")
      .def(py::init<std::vector<at::Tensor>, std::vector<std::string>, int64_t, at::Tensor>())
      .def("gather", &FastGather::gather);
  py::class_<FastPer>(m, "FastPer")
      .def(py::init<std::shared_ptr<FastGather>, uint64_t, at::Tensor, at::Tensor, at::Tensor, at::Tensor,
                    int64_t, double>())
      .def("sample", &FastPer::sample)
      .def("update", &FastPer::update)
      .def("set_sibling", &FastPer::set_sibling)
      .def("set_exchange", &FastPer::set_exchange)
      .def("update_replicated", &FastPer::update_replicated);
}
