// One learner step of DQNAgent.update (hanabi_agents/rlax_dqn/rlax_rainbow.py:310-339).
//
// Single process (hb_learner_step): ONE replayed CUDA graph
//     sample + prios + importance weights + gather   one launch (hb_rb_sample_gather), keys and the
//                                                     step's scalars read in place from pinned memory
//     online forward (2B rows) | target forward (B rows) | transposes, clears, Adam tick
//     loss -> update_priorities (own stream)
//     dZ chain | per layer: dW GEMM, bias gradient, optix.adam of that layer
// with the kernels of the critical path chained by programmatic dependent launch.
// Data parallel (hb_learner_grads / hb_learner_apply): the same without the per-layer Adam, as two
// graphs around the caller's gradient all-reduce.
//
// What changes from step to step -- the stratified keys (or the uniform indices) and the scalars that
// are schedules in the reference (beta_is, learning_rate; params.py:5-20) -- lives in a ring of pinned
// staging slots that the kernels read, so nothing but shapes and pointers is frozen into a graph; every
// slot has its own graph (the slot address is a kernel argument).
//
// Built only on the public entry points of the other translation units (hb_rb_*, hb_net_*).
#include "common.cuh"
#include <new>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

struct hb_learner {
  int device = 0;
  hb_qnet* net = nullptr;
  hb_replay* rb = nullptr;
  int B = 0, stride = 0, prioritized = 0;
  cudaStream_t stream = nullptr, prio = nullptr;   // prio: priority update beside the backward pass
  cudaEvent_t ev_in = nullptr, ev_out = nullptr, ev_prio = nullptr;
  // staging (pinned): per slot [B keys | 8 scalars]; scalars = beta, discount, lr, b1, b2, eps, grad_scale
  static constexpr int kSlots = 2;   // each slot has its own captured graph; the host runs at most 2 steps ahead
  static constexpr int kDyn = 8;
  float* h_in = nullptr;         // [kSlots][B + kDyn]
  int64_t* h_idx = nullptr;      // [kSlots][B] (uniform replay)
  float* h_dynB = nullptr;       // [kSlots][kDyn] scalars of phase B (data-parallel path)
  cudaEvent_t slot_done[kSlots] = {nullptr, nullptr};
  cudaEvent_t slotB_done[kSlots] = {nullptr, nullptr};
  int slot = 0, slotB = 0;
  const float* d_in = nullptr;   // h_in as the device addresses it
  // the slot is copied to device memory on `copy` when it is staged (the host runs ahead of the device)
  // instead of being read over PCIe by the sample kernel: -0.6 us per update (HB_LEARNER_KEYS_H2D=0: in place)
  int keys_h2d = 1;
  float* d_keys = nullptr;       // [kSlots][B + kDyn]
  cudaStream_t copy = nullptr;
  cudaEvent_t ev_keys = nullptr;
  int64_t* d_idx = nullptr;
  float* d_p32 = nullptr;
  float* d_dyn = nullptr;        // [kDyn] this step's scalars on the device
  float* d_dynB = nullptr;
  uint32_t *o1 = nullptr, *o2 = nullptr;
  int32_t* act = nullptr;
  float* rew = nullptr;
  uint8_t* term = nullptr;
  float* loss = nullptr;         // [B]
  float* mean = nullptr;         // [1]
  float* w_is = nullptr;         // the network's importance-weight buffer
  // graphs: per staging slot; what they were captured with
  cudaGraphExec_t gStep[kSlots] = {nullptr, nullptr};
  cudaGraphExec_t gA[kSlots] = {nullptr, nullptr};
  cudaGraphExec_t gB = nullptr;
  int kA_mask = -1;              // mask_terminal the graphs were captured with
  const void* kA_ptr[4] = {nullptr, nullptr};   // the net's staging the graphs hold
  int64_t rb_gen = -1;           // generation of the replay shard's scratch the graphs were captured with
  int warm = 0;                  // eager steps run (the first one allocates workspaces)
  int recaptures = 0;
  int use_graph = 1;
  // update_priorities beside the backward pass (default) or after Adam (HB_LEARNER_PRIO_OVERLAP=0)
  int prio_overlap = 1;
  int pdl = 1;                   // HB_LEARNER_PDL=0: no programmatic dependent launch
  hb_comm* comm = nullptr;       // data parallel: per-layer all-reduce inside hb_learner_step
  int world = 1;
};

extern "C" const void* hb_net_kernel_handle(int which);   // qnet.cu

namespace {

void drop_graphs(hb_learner* L) {
  for (int i = 0; i < hb_learner::kSlots; ++i) {
    if (L->gStep[i]) { cudaGraphExecDestroy(L->gStep[i]); L->gStep[i] = nullptr; }
    if (L->gA[i]) { cudaGraphExecDestroy(L->gA[i]); L->gA[i] = nullptr; }
  }
  if (L->gB) { cudaGraphExecDestroy(L->gB); L->gB = nullptr; }
}

// sample / gather -> gradients (-> per-layer Adam when `fused_adam`) -> priorities.  `slot`: the staging
// slot holding this step's keys and scalars.
int phase_a(hb_learner* L, int slot, float beta, float discount, int mask_terminal, bool fused_adam,
            float lr, float b1, float b2, float eps, float grad_scale, cudaStream_t s) {
  int rc;
  const float* h_keys = (L->keys_h2d ? L->d_keys : L->d_in) + (size_t)slot * (L->B + hb_learner::kDyn);
  // (pinned host memory read in place, or its device copy)
  if (L->prioritized) {
    rc = hb_rb_sample_gather(L->rb, h_keys, L->B, h_keys + L->B, L->d_dyn, beta, L->d_idx, L->d_p32, nullptr,
                             L->w_is, L->o1, L->act, L->rew, L->o2, nullptr, L->term, 0, s);
  } else {   // d_idx and d_dyn were filled by the copies enqueued just before this phase
    rc = hb_rb_gather(L->rb, L->d_idx, L->B, L->o1, L->act, L->rew, L->o2, nullptr, L->term, s);
  }
  if (rc) return rc;
  hb_net_step_opts opts;
  memset(&opts, 0, sizeof(opts));
  opts.d_w_is = L->prioritized ? L->w_is : nullptr;       // uniform replay: prios = ones  (rlax_rainbow.py:318)
  opts.d_dyn = L->d_dyn;
  opts.pdl = L->pdl;
  opts.adam_per_layer = fused_adam ? 1 : 0;
  opts.lr = lr; opts.b1 = b1; opts.b2 = b2; opts.eps = eps; opts.grad_scale = grad_scale;
  opts.comm = fused_adam ? L->comm : nullptr;
  rc = hb_net_grads_ex(L->net, L->o1, L->act, L->rew, L->o2, L->term, nullptr, nullptr, L->B, beta, discount,
                       mask_terminal, L->loss, L->mean, &opts, s);
  if (rc || !L->prioritized) return rc;
  if (!L->prio_overlap) {
    if (fused_adam) return hb_rb_update_priorities(L->rb, L->d_idx, L->loss, L->B, s);
    return HB_OK;      // phase B does it
  }
  // update_priorities (rlax_rainbow.py:331-334) needs only the per-sample losses: it runs on its own stream
  // beside the backward GEMMs instead of after Adam (same values: nothing between the two reads the tree)
  HB_CUDA(cudaStreamWaitEvent(L->prio, (cudaEvent_t)hb_net_loss_event(L->net), 0));
  rc = hb_rb_update_priorities(L->rb, L->d_idx, L->loss, L->B, L->prio);
  HB_CUDA(cudaEventRecord(L->ev_prio, L->prio));
  HB_CUDA(cudaStreamWaitEvent(s, L->ev_prio, 0));
  return rc;
}

int phase_b(hb_learner* L, cudaStream_t s) {
  int rc = hb_net_adam_dyn(L->net, L->d_dynB, s);
  if (!rc && L->prioritized && !L->prio_overlap) rc = hb_rb_update_priorities(L->rb, L->d_idx, L->loss, L->B, s);
  return rc;
}

// The packed rows are gathered straight into the network's staging and the losses are read where the
// network leaves them (hb_net_learner_buffers): four device-to-device copies less on the critical path of
// every update.  Should the network ever reallocate them, every graph is captured again.
int borrow_net_buffers(hb_learner* L, int mask_terminal) {
  uint32_t* st = nullptr;
  float *lo = nullptr, *me = nullptr;
  int rc = hb_net_learner_buffers(L->net, L->B, &st, &lo, &me);
  if (rc) return rc;
  float* w = hb_net_is_weights(L->net, L->B);
  HB_REQUIRE(w != nullptr, "hb_learner: the network has no importance-weight buffer");
  const void* now[4] = {st, lo, me, w};
  // the replay scratch (sampled values / new priorities) is baked into the graphs as well: a call with a
  // larger n than the batch (sample_batch(512), a long update_priorities) moves it (hb_rb_info 6)
  const int64_t gen = hb_rb_info(L->rb, 6);
  if (memcmp(now, L->kA_ptr, sizeof(now)) != 0 || gen != L->rb_gen || mask_terminal != L->kA_mask) {
    L->rb_gen = gen;
    L->kA_mask = mask_terminal;
    drop_graphs(L);
    memcpy(L->kA_ptr, now, sizeof(now));
  }
  L->o1 = st;
  L->o2 = st + (size_t)L->B * L->stride;
  L->loss = lo;
  L->mean = me;
  L->w_is = w;
  return HB_OK;
}

// HB_GRAPH_DOT: the types of the edges into the update's two PDL kernels on stderr.  (Measured: stream
// capture makes EVERY incoming edge of a launch with the programmatic attribute programmatic -- type 1, port
// 1 -- the cross-stream ones included; the backward chain still starts 2 us after the loss has ended, every SM
// holding loss CTAs of 52 KB beside which a 213 KB chain CTA does not fit.)
int inspect_edges(cudaGraph_t g) {
  if (getenv("HB_GRAPH_DOT") == nullptr) return HB_OK;
  size_t n = 0;
  HB_CUDA(cudaGraphGetEdges_v2(g, nullptr, nullptr, nullptr, &n));
  if (n == 0) return HB_OK;
  cudaGraphNode_t* from = new cudaGraphNode_t[n];
  cudaGraphNode_t* to = new cudaGraphNode_t[n];
  cudaGraphEdgeData* data = new cudaGraphEdgeData[n];
  const cudaError_t e = cudaGraphGetEdges_v2(g, from, to, data, &n);
  auto name_of = [](cudaGraphNode_t node) -> const char* {
    cudaGraphNodeType t;
    if (cudaGraphNodeGetType(node, &t) != cudaSuccess || t != cudaGraphNodeTypeKernel) return "other";
    cudaKernelNodeParams kp;
    if (cudaGraphKernelNodeGetParams(node, &kp) != cudaSuccess) { cudaGetLastError(); return "kernel"; }
    if (kp.func == hb_net_kernel_handle(0)) return "chain";
    if (kp.func == hb_net_kernel_handle(1) || kp.func == hb_net_kernel_handle(2)) return "loss";
    return "kernel";
  };
  for (size_t i = 0; i < n && e == cudaSuccess; ++i) {
    const char* nt = name_of(to[i]);
    if (nt[0] == 'c' || nt[0] == 'l')
      fprintf(stderr, "[graph] edge -> %s from %s: type %d from_port %d\n", nt, name_of(from[i]),
              (int)data[i].type, (int)data[i].from_port);
  }
  delete[] from; delete[] to; delete[] data;
  if (e != cudaSuccess) return hb::cuda_fail(e, "inspect_edges", __FILE__, __LINE__);
  return HB_OK;
}

template <class F>
int capture(hb_learner* L, cudaGraphExec_t* exec, F body) {
  if (*exec) { cudaGraphExecDestroy(*exec); *exec = nullptr; }
  cudaGraph_t g = nullptr;
  HB_CUDA(cudaStreamBeginCapture(L->stream, cudaStreamCaptureModeThreadLocal));
  const int rc = body();
  const cudaError_t e = cudaStreamEndCapture(L->stream, &g);
  if (rc != HB_OK) { if (g) cudaGraphDestroy(g); return rc; }
  if (e != cudaSuccess) return hb::cuda_fail(e, "cudaStreamEndCapture", __FILE__, __LINE__);
  // HB_GRAPH_DOT=<path>: the captured update (nodes, edges and their types) as a dot file
  if (const char* dot = getenv("HB_GRAPH_DOT")) {
    static int n_dots = 0;
    char path[512];
    snprintf(path, sizeof(path), "%s.%d", dot, n_dots++);
    cudaGraphDebugDotPrint(g, path, cudaGraphDebugDotFlagsKernelNodeAttributes);
  }
  if (int prc = inspect_edges(g)) { cudaGraphDestroy(g); return prc; }
  const cudaError_t e2 = cudaGraphInstantiate(exec, g, 0);
  cudaGraphDestroy(g);
  if (e2 != cudaSuccess) return hb::cuda_fail(e2, "cudaGraphInstantiate", __FILE__, __LINE__);
  ++L->recaptures;
  return HB_OK;
}

// Stages this step's inputs in the next slot of the ring and orders L->stream behind the caller's stream.
int stage_inputs(hb_learner* L, const float* h_keys, const int64_t* h_idx, const float* dyn,
                 cudaStream_t cs, int* slot_out) {
  const int slot = L->slot;
  L->slot = (L->slot + 1) % hb_learner::kSlots;
  HB_CUDA(cudaEventSynchronize(L->slot_done[slot]));     // the step that last read this slot is through
  HB_CUDA(cudaEventRecord(L->ev_in, cs));
  HB_CUDA(cudaStreamWaitEvent(L->stream, L->ev_in, 0));
  float* in = L->h_in + (size_t)slot * (L->B + hb_learner::kDyn);
  memcpy(in + L->B, dyn, sizeof(float) * hb_learner::kDyn);
  if (L->prioritized) {
    memcpy(in, h_keys, (size_t)L->B * 4);                // read in place by the sample kernel
    if (L->keys_h2d) {
      const size_t words = (size_t)L->B + hb_learner::kDyn;
      HB_CUDA(cudaMemcpyAsync(L->d_keys + (size_t)slot * words, in, words * 4, cudaMemcpyHostToDevice, L->copy));
      HB_CUDA(cudaEventRecord(L->ev_keys, L->copy));
      HB_CUDA(cudaStreamWaitEvent(L->stream, L->ev_keys, 0));
    }
  } else {
    int64_t* hi = L->h_idx + (size_t)slot * L->B;
    memcpy(hi, h_idx, (size_t)L->B * 8);
    HB_CUDA(cudaMemcpyAsync(L->d_idx, hi, (size_t)L->B * 8, cudaMemcpyHostToDevice, L->stream));
    HB_CUDA(cudaMemcpyAsync(L->d_dyn, in + L->B, sizeof(float) * hb_learner::kDyn, cudaMemcpyHostToDevice,
                            L->stream));
  }
  *slot_out = slot;
  return HB_OK;
}

int run_phase_a(hb_learner* L, const float* h_keys, const int64_t* h_idx, float beta, float discount,
                int mask_terminal, bool fused_adam, float lr, float b1, float b2, float eps, float grad_scale,
                void* stream) {
  HB_REQUIRE(L, "hb_learner: NULL learner");
  HB_REQUIRE(L->prioritized ? h_keys != nullptr : h_idx != nullptr,
             "hb_learner: %s", L->prioritized ? "keys required" : "indices required");
  hb::DeviceGuard g(L->device);
  cudaStream_t cs = (cudaStream_t)stream;
  if (int brc = borrow_net_buffers(L, mask_terminal)) return brc;
  const float dyn[hb_learner::kDyn] = {beta, discount, lr, b1, b2, eps, grad_scale, 0.f};
  int slot = 0;
  if (int src = stage_inputs(L, h_keys, h_idx, dyn, cs, &slot)) return src;
  int rc = HB_OK;
  if (!L->use_graph || L->warm < 2) {
    rc = phase_a(L, slot, beta, discount, mask_terminal, fused_adam, lr, b1, b2, eps, grad_scale, L->stream);
    if (fused_adam) ++L->warm;
  } else {
    cudaGraphExec_t* exec = fused_adam ? &L->gStep[slot] : &L->gA[slot];
    if (!*exec)
      rc = capture(L, exec, [&] {
        return phase_a(L, slot, beta, discount, mask_terminal, fused_adam, lr, b1, b2, eps, grad_scale, L->stream);
      });
    if (!rc) HB_CUDA(cudaGraphLaunch(*exec, L->stream));
  }
  if (rc) return rc;
  HB_CUDA(cudaEventRecord(L->slot_done[slot], L->stream));
  HB_CUDA(cudaEventRecord(L->ev_out, L->stream));
  HB_CUDA(cudaStreamWaitEvent(cs, L->ev_out, 0));
  return HB_OK;
}

}  // namespace

extern "C" {

int hb_learner_destroy(hb_learner* L);

int hb_learner_create(hb_qnet* net, hb_replay* rb, int batch, int device, hb_learner** out) {
  HB_REQUIRE(net && rb && out && batch >= 1, "hb_learner_create: bad arguments");
  hb::DeviceGuard g(device);
  if (!g.ok) { hb::set_error("hb_learner_create: cannot select CUDA device %d", device); return HB_ERR_CUDA; }
  hb_learner* L = new (std::nothrow) hb_learner();
  HB_REQUIRE(L != nullptr, "hb_learner_create: out of host memory");
  L->device = device;
  L->net = net;
  L->rb = rb;
  L->B = batch;
  L->stride = (int)hb_rb_info(rb, 3);
  L->prioritized = (int)hb_rb_info(rb, 5);
  if (L->prioritized && batch > 1024) {
    delete L;
    hb::set_error("hb_learner_create: prioritized batches of at most 1024 rows (got %d)", batch);
    return HB_ERR_INVALID;
  }
  if (const char* v = getenv("HB_LEARNER_PRIO_OVERLAP")) L->prio_overlap = atoi(v) != 0;
  if (const char* v = getenv("HB_LEARNER_PDL")) L->pdl = atoi(v) != 0;
  if (const char* v = getenv("HB_LEARNER_KEYS_H2D")) L->keys_h2d = atoi(v) != 0;
  const size_t B = (size_t)batch;
  const int S = hb_learner::kSlots, D = hb_learner::kDyn;
  // the update's main chain runs at the highest priority, the priority write-back at the lowest (qnet.cu:
  // "Priorities")
  int prio_lo = 0, prio_hi = 0;
  cudaError_t e = cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
  if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&L->stream, cudaStreamNonBlocking, prio_hi);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&L->ev_in, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&L->ev_out, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&L->prio, cudaStreamNonBlocking, prio_lo);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&L->ev_prio, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaHostAlloc((void**)&L->h_in, S * (B + D) * 4, cudaHostAllocMapped | cudaHostAllocPortable);
  if (e == cudaSuccess) e = cudaHostGetDevicePointer((void**)&L->d_in, L->h_in, 0);
  if (e == cudaSuccess) e = cudaMallocHost(&L->h_idx, S * B * 8);
  if (e == cudaSuccess) e = cudaMallocHost(&L->h_dynB, S * D * 4);
  for (int i = 0; i < S && e == cudaSuccess; ++i)
    e = cudaEventCreateWithFlags(&L->slot_done[i], cudaEventDisableTiming);
  for (int i = 0; i < S && e == cudaSuccess; ++i)
    e = cudaEventCreateWithFlags(&L->slotB_done[i], cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaMalloc(&L->d_keys, S * (B + D) * 4);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&L->copy, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&L->ev_keys, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaMalloc(&L->d_idx, B * 8);
  if (e == cudaSuccess) e = cudaMalloc(&L->d_p32, B * 4);
  if (e == cudaSuccess) e = cudaMalloc(&L->d_dyn, D * 4);
  if (e == cudaSuccess) e = cudaMalloc(&L->d_dynB, D * 4);
  if (e == cudaSuccess) e = cudaMalloc(&L->act, B * 4);
  if (e == cudaSuccess) e = cudaMalloc(&L->rew, B * 4);
  if (e == cudaSuccess) e = cudaMalloc(&L->term, B);
  if (e != cudaSuccess) {
    hb_learner_destroy(L);
    return hb::cuda_fail(e, "hb_learner_create alloc", __FILE__, __LINE__);
  }
  if (int rc = borrow_net_buffers(L, -1)) { hb_learner_destroy(L); return rc; }
  *out = L;
  return HB_OK;
}

int hb_learner_destroy(hb_learner* L) {
  if (!L) return HB_OK;
  hb::DeviceGuard g(L->device);
  cudaDeviceSynchronize();
  drop_graphs(L);
  if (L->h_in) cudaFreeHost(L->h_in);
  if (L->h_idx) cudaFreeHost(L->h_idx);
  if (L->h_dynB) cudaFreeHost(L->h_dynB);
  void* dv[] = {L->d_idx, L->d_p32, L->d_dyn, L->d_dynB, L->act, L->rew, L->term};   // o1 / o2 / loss / mean / w_is are the net's
  for (void* p : dv)
    if (p) cudaFree(p);
  for (cudaEvent_t ev : L->slot_done)
    if (ev) cudaEventDestroy(ev);
  for (cudaEvent_t ev : L->slotB_done)
    if (ev) cudaEventDestroy(ev);
  if (L->ev_in) cudaEventDestroy(L->ev_in);
  if (L->ev_out) cudaEventDestroy(L->ev_out);
  if (L->ev_prio) cudaEventDestroy(L->ev_prio);
  if (L->ev_keys) cudaEventDestroy(L->ev_keys);
  if (L->copy) cudaStreamDestroy(L->copy);
  if (L->d_keys) cudaFree(L->d_keys);
  if (L->prio) cudaStreamDestroy(L->prio);
  if (L->stream) cudaStreamDestroy(L->stream);
  delete L;
  return HB_OK;
}

/* 0: disable graph replay (every step launched eagerly), 1: enable (default). */
int hb_learner_set_graph(hb_learner* L, int on) {
  HB_REQUIRE(L, "hb_learner_set_graph: NULL learner");
  L->use_graph = on ? 1 : 0;
  return HB_OK;
}

/* what: 0 batch, 1 graph captures so far, 2 eager steps run */
int64_t hb_learner_info(const hb_learner* L, int what) {
  if (!L) return -1;
  switch (what) { case 0: return L->B; case 1: return L->recaptures; case 2: return L->warm; default: return -1; }
}

float* hb_learner_loss(hb_learner* L) { return L ? L->loss : nullptr; }
float* hb_learner_mean_loss(hb_learner* L) { return L ? L->mean : nullptr; }
int64_t* hb_learner_indices(hb_learner* L) { return L ? L->d_idx : nullptr; }

/* The whole update in one call (single process): sample -> gather -> jax.grad(loss) -> optix.adam ->
 * update_priorities.  Ordered after everything already enqueued on `stream`; `stream` waits for it. */
int hb_learner_step(hb_learner* L, const float* h_keys, const int64_t* h_idx, float beta, float discount,
                    int mask_terminal, float lr, float b1, float b2, float eps, void* stream) {
  const float scale = L && L->comm ? 1.0f / (float)L->world : 1.0f;
  return run_phase_a(L, h_keys, h_idx, beta, discount, mask_terminal, true, lr, b1, b2, eps, scale, stream);
}

int hb_learner_set_comm(hb_learner* L, hb_comm* comm, int world) {
  HB_REQUIRE(L && world >= 1, "hb_learner_set_comm: bad arguments");
  hb::DeviceGuard g(L->device);
  HB_CUDA(cudaDeviceSynchronize());
  drop_graphs(L);                 // the communicator is part of what the graphs captured
  L->comm = world > 1 ? comm : nullptr;
  L->world = world;
  return HB_OK;
}

/* Phase A: sample (h_keys: B stratified keys as float32, prioritized shards; h_idx: B row indices,
 * uniform shards) -> gather -> forward / loss / backward.  Leaves the gradients in the network's
 * flat bucket and the per-sample losses / their weighted mean in hb_learner_loss / _mean_loss.
 * Ordered after everything already enqueued on `stream`; `stream` waits for its completion. */
int hb_learner_grads(hb_learner* L, const float* h_keys, const int64_t* h_idx, float beta, float discount,
                     int mask_terminal, void* stream) {
  return run_phase_a(L, h_keys, h_idx, beta, discount, mask_terminal, false, 0.f, 0.f, 0.f, 0.f, 1.f, stream);
}

/* Phase B: optix.adam on grad_scale * bucket (rlax_rainbow.py:215-216) and, for prioritized shards when
 * the overlap is off, update_priorities(indices, |loss|) (:333-334). */
int hb_learner_apply(hb_learner* L, float lr, float b1, float b2, float eps, float grad_scale,
                     void* stream) {
  HB_REQUIRE(L, "hb_learner_apply: NULL learner");
  hb::DeviceGuard g(L->device);
  cudaStream_t cs = (cudaStream_t)stream;
  HB_CUDA(cudaEventRecord(L->ev_in, cs));
  HB_CUDA(cudaStreamWaitEvent(L->stream, L->ev_in, 0));
  // the step's hyper-parameters: pinned ring slot -> device, outside the graph (the source moves)
  const int slot = L->slotB;
  L->slotB = (L->slotB + 1) % hb_learner::kSlots;
  float* hb = L->h_dynB + (size_t)slot * hb_learner::kDyn;
  HB_CUDA(cudaEventSynchronize(L->slotB_done[slot]));    // the copy that last read this slot is done
  const float dyn[hb_learner::kDyn] = {0.f, 0.f, lr, b1, b2, eps, grad_scale, 0.f};
  memcpy(hb, dyn, sizeof(dyn));
  HB_CUDA(cudaMemcpyAsync(L->d_dynB, hb, sizeof(dyn), cudaMemcpyHostToDevice, L->stream));
  HB_CUDA(cudaEventRecord(L->slotB_done[slot], L->stream));
  int rc = HB_OK;
  if (!L->use_graph || L->warm < 2) {
    rc = phase_b(L, L->stream);
    ++L->warm;
  } else {
    if (!L->gB) rc = capture(L, &L->gB, [&] { return phase_b(L, L->stream); });
    if (!rc) HB_CUDA(cudaGraphLaunch(L->gB, L->stream));
  }
  if (rc) return rc;
  HB_CUDA(cudaEventRecord(L->ev_out, L->stream));
  HB_CUDA(cudaStreamWaitEvent(cs, L->ev_out, 0));
  return HB_OK;
}

}  // extern "C"
