// One learner step of DQNAgent.update (hanabi_agents/rlax_dqn/rlax_rainbow.py:310-339) as TWO
// C calls around the caller's gradient all-reduce: stratified sample -> gather -> forward / loss /
// backward (hb_learner_grads) | Adam -> priority update (hb_learner_apply).  A phase is a fixed
// sequence of ~15-20 small launches, so each is captured once into a CUDA graph and replayed; what
// changes from step to step (the stratified keys, or the uniform indices) is copied from a 4-slot
// ring of pinned staging rows to the device right before the phase-A graph is launched.
//
// Built only on the public entry points of the other translation units (hb_rb_*, hb_net_*).
#include "common.cuh"
#include <new>
#include <stdlib.h>
#include <string.h>

struct hb_learner {
  int device = 0;
  hb_qnet* net = nullptr;
  hb_replay* rb = nullptr;
  int B = 0, stride = 0, prioritized = 0;
  cudaStream_t stream = nullptr, prio = nullptr;   // prio: priority update beside the backward pass
  cudaEvent_t ev_in = nullptr, ev_out = nullptr, ev_prio = nullptr;
  // staging (pinned) and device inputs
  static constexpr int kSlots = 4;
  float* h_keys = nullptr;       // [kSlots][B] ring of pinned staging rows
  int64_t* h_idx = nullptr;      // [kSlots][B] (uniform replay)
  cudaEvent_t slot_done[kSlots] = {nullptr, nullptr, nullptr, nullptr};
  int slot = 0;
  float* d_keys = nullptr;
  int64_t* d_idx = nullptr;
  float* d_p32 = nullptr;
  uint32_t *o1 = nullptr, *o2 = nullptr;
  int32_t* act = nullptr;
  float* rew = nullptr;
  uint8_t* term = nullptr;
  float* loss = nullptr;         // [B]
  float* mean = nullptr;         // [1]
  // graphs + the scalar arguments they were captured with
  cudaGraphExec_t gA = nullptr, gB = nullptr;
  float kA[3] = {0, 0, 0};       // beta, discount, mask_terminal
  const void* kA_ptr[3] = {nullptr, nullptr, nullptr};   // the net's staging the graph was captured with
  int64_t rb_gen = -1;           // generation of the replay shard's scratch the graphs were captured with
  float kB[5] = {0, 0, 0, 0, 0}; // lr, b1, b2, eps, grad_scale
  int warm = 0;                  // eager steps run (the first one allocates workspaces)
  int recaptures = 0;
  int use_graph = 1;
  // update_priorities beside the backward pass (default) or after Adam (HB_LEARNER_PRIO_OVERLAP=0): alone the
  // learner gains 7 % from the overlap, a self-play session tick with two updates 4 % (0.68 vs 0.70 ms)
  int prio_overlap = 1;
};

namespace {

int phase_a(hb_learner* L, float beta, float discount, int mask_terminal, cudaStream_t s) {
  int rc;
  if (L->prioritized) {   // d_keys / d_idx were filled by the copy enqueued just before this phase
    if ((rc = hb_rb_sample(L->rb, L->d_keys, L->B, L->d_idx, nullptr, L->d_p32, s))) return rc;
  }
  if ((rc = hb_rb_gather(L->rb, L->d_idx, L->B, L->o1, L->act, L->rew, L->o2, nullptr, L->term, s))) return rc;
  rc = hb_net_grads(L->net, L->o1, L->act, L->rew, L->o2, L->term, L->prioritized ? L->d_p32 : nullptr,
                    nullptr, L->B, beta, discount, mask_terminal, L->loss, L->mean, s);
  if (rc || !L->prioritized || !L->prio_overlap) return rc;
  // update_priorities (rlax_rainbow.py:331-334) needs only the per-sample losses: it runs on its own stream
  // beside the backward GEMMs instead of after Adam (same values: nothing between the two reads the tree)
  HB_CUDA(cudaStreamWaitEvent(L->prio, (cudaEvent_t)hb_net_loss_event(L->net), 0));
  rc = hb_rb_update_priorities(L->rb, L->d_idx, L->loss, L->B, L->prio);
  HB_CUDA(cudaEventRecord(L->ev_prio, L->prio));
  HB_CUDA(cudaStreamWaitEvent(s, L->ev_prio, 0));
  return rc;
}

int phase_b(hb_learner* L, float lr, float b1, float b2, float eps, float grad_scale, cudaStream_t s) {
  int rc = hb_net_adam(L->net, lr, b1, b2, eps, grad_scale, s);
  if (!rc && L->prioritized && !L->prio_overlap) rc = hb_rb_update_priorities(L->rb, L->d_idx, L->loss, L->B, s);
  return rc;
}

// The packed rows are gathered straight into the network's staging and the losses are read where the
// network leaves them (hb_net_learner_buffers): four device-to-device copies less on the critical path of
// every update.  Should the network ever reallocate them, both graphs are captured again.
int borrow_net_buffers(hb_learner* L) {
  uint32_t* st = nullptr;
  float *lo = nullptr, *me = nullptr;
  int rc = hb_net_learner_buffers(L->net, L->B, &st, &lo, &me);
  if (rc) return rc;
  const void* now[3] = {st, lo, me};
  // the replay scratch (sampled values / new priorities) is baked into both graphs as well: a call with a
  // larger n than the batch (sample_batch(512), a long update_priorities) moves it (hb_rb_info 6)
  const int64_t gen = hb_rb_info(L->rb, 6);
  if (memcmp(now, L->kA_ptr, sizeof(now)) != 0 || gen != L->rb_gen) {
    L->rb_gen = gen;
    if (L->gA) { cudaGraphExecDestroy(L->gA); L->gA = nullptr; }
    if (L->gB) { cudaGraphExecDestroy(L->gB); L->gB = nullptr; }
    memcpy(L->kA_ptr, now, sizeof(now));
  }
  L->o1 = st;
  L->o2 = st + (size_t)L->B * L->stride;
  L->loss = lo;
  L->mean = me;
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
  const cudaError_t e2 = cudaGraphInstantiate(exec, g, 0);
  cudaGraphDestroy(g);
  if (e2 != cudaSuccess) return hb::cuda_fail(e2, "cudaGraphInstantiate", __FILE__, __LINE__);
  ++L->recaptures;
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
  if (const char* v = getenv("HB_LEARNER_PRIO_OVERLAP")) L->prio_overlap = atoi(v) != 0;
  const size_t B = (size_t)batch;
  cudaError_t e = cudaStreamCreateWithFlags(&L->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&L->ev_in, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&L->ev_out, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&L->prio, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&L->ev_prio, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaMallocHost(&L->h_keys, hb_learner::kSlots * B * 4);
  if (e == cudaSuccess) e = cudaMallocHost(&L->h_idx, hb_learner::kSlots * B * 8);
  for (int i = 0; i < hb_learner::kSlots && e == cudaSuccess; ++i)
    e = cudaEventCreateWithFlags(&L->slot_done[i], cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaMalloc(&L->d_keys, B * 4);
  if (e == cudaSuccess) e = cudaMalloc(&L->d_idx, B * 8);
  if (e == cudaSuccess) e = cudaMalloc(&L->d_p32, B * 4);
  if (e == cudaSuccess) e = cudaMalloc(&L->act, B * 4);
  if (e == cudaSuccess) e = cudaMalloc(&L->rew, B * 4);
  if (e == cudaSuccess) e = cudaMalloc(&L->term, B);
  if (e != cudaSuccess) {
    hb_learner_destroy(L);
    return hb::cuda_fail(e, "hb_learner_create alloc", __FILE__, __LINE__);
  }
  if (int rc = borrow_net_buffers(L)) { hb_learner_destroy(L); return rc; }
  *out = L;
  return HB_OK;
}

int hb_learner_destroy(hb_learner* L) {
  if (!L) return HB_OK;
  hb::DeviceGuard g(L->device);
  cudaDeviceSynchronize();
  if (L->gA) cudaGraphExecDestroy(L->gA);
  if (L->gB) cudaGraphExecDestroy(L->gB);
  if (L->h_keys) cudaFreeHost(L->h_keys);
  if (L->h_idx) cudaFreeHost(L->h_idx);
  void* dv[] = {L->d_keys, L->d_idx, L->d_p32, L->act, L->rew, L->term};   // o1 / o2 / loss / mean are the net's
  for (void* p : dv)
    if (p) cudaFree(p);
  for (cudaEvent_t ev : L->slot_done)
    if (ev) cudaEventDestroy(ev);
  if (L->ev_in) cudaEventDestroy(L->ev_in);
  if (L->ev_out) cudaEventDestroy(L->ev_out);
  if (L->ev_prio) cudaEventDestroy(L->ev_prio);
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

/* Phase A: sample (h_keys: B stratified keys as float32, prioritized shards; h_idx: B row indices,
 * uniform shards) -> gather -> forward / loss / backward.  Leaves the gradients in the network's
 * flat bucket and the per-sample losses / their weighted mean in hb_learner_loss / _mean_loss.
 * Ordered after everything already enqueued on `stream`; `stream` waits for its completion. */
int hb_learner_grads(hb_learner* L, const float* h_keys, const int64_t* h_idx, float beta, float discount,
                     int mask_terminal, void* stream) {
  HB_REQUIRE(L, "hb_learner_grads: NULL learner");
  HB_REQUIRE(L->prioritized ? h_keys != nullptr : h_idx != nullptr,
             "hb_learner_grads: %s", L->prioritized ? "keys required" : "indices required");
  hb::DeviceGuard g(L->device);
  cudaStream_t cs = (cudaStream_t)stream;
  if (int brc = borrow_net_buffers(L)) return brc;
  // this step's keys / indices: pinned ring slot -> device (outside the graph: the source moves)
  const int slot = L->slot;
  L->slot = (L->slot + 1) % hb_learner::kSlots;
  HB_CUDA(cudaEventSynchronize(L->slot_done[slot]));     // the copy that last used this slot is done
  HB_CUDA(cudaEventRecord(L->ev_in, cs));
  HB_CUDA(cudaStreamWaitEvent(L->stream, L->ev_in, 0));
  if (L->prioritized) {
    float* hk = L->h_keys + (size_t)slot * L->B;
    memcpy(hk, h_keys, (size_t)L->B * 4);
    HB_CUDA(cudaMemcpyAsync(L->d_keys, hk, (size_t)L->B * 4, cudaMemcpyHostToDevice, L->stream));
  } else {
    int64_t* hi = L->h_idx + (size_t)slot * L->B;
    memcpy(hi, h_idx, (size_t)L->B * 8);
    HB_CUDA(cudaMemcpyAsync(L->d_idx, hi, (size_t)L->B * 8, cudaMemcpyHostToDevice, L->stream));
  }
  HB_CUDA(cudaEventRecord(L->slot_done[slot], L->stream));
  int rc = HB_OK;
  const float key[3] = {beta, discount, (float)mask_terminal};
  if (!L->use_graph || L->warm < 2) {
    rc = phase_a(L, beta, discount, mask_terminal, L->stream);      // eager (allocates on first use)
  } else {
    if (!L->gA || memcmp(key, L->kA, sizeof(key)) != 0) {
      rc = capture(L, &L->gA, [&] { return phase_a(L, beta, discount, mask_terminal, L->stream); });
      memcpy(L->kA, key, sizeof(key));
    }
    if (!rc) HB_CUDA(cudaGraphLaunch(L->gA, L->stream));
  }
  if (rc) return rc;
  HB_CUDA(cudaEventRecord(L->ev_out, L->stream));
  HB_CUDA(cudaStreamWaitEvent(cs, L->ev_out, 0));
  return HB_OK;
}

/* Phase B: optix.adam on grad_scale * bucket (rlax_rainbow.py:215-216) and, for prioritized shards,
 * update_priorities(indices, |loss|) (:333-334). */
int hb_learner_apply(hb_learner* L, float lr, float b1, float b2, float eps, float grad_scale,
                     void* stream) {
  HB_REQUIRE(L, "hb_learner_apply: NULL learner");
  hb::DeviceGuard g(L->device);
  cudaStream_t cs = (cudaStream_t)stream;
  HB_CUDA(cudaEventRecord(L->ev_in, cs));
  HB_CUDA(cudaStreamWaitEvent(L->stream, L->ev_in, 0));
  int rc = HB_OK;
  const float key[5] = {lr, b1, b2, eps, grad_scale};
  if (!L->use_graph || L->warm < 2) {
    rc = phase_b(L, lr, b1, b2, eps, grad_scale, L->stream);
    ++L->warm;
  } else {
    if (!L->gB || memcmp(key, L->kB, sizeof(key)) != 0) {
      rc = capture(L, &L->gB, [&] { return phase_b(L, lr, b1, b2, eps, grad_scale, L->stream); });
      memcpy(L->kB, key, sizeof(key));
    }
    if (!rc) HB_CUDA(cudaGraphLaunch(L->gB, L->stream));
  }
  if (rc) return rc;
  HB_CUDA(cudaEventRecord(L->ev_out, L->stream));
  HB_CUDA(cudaStreamWaitEvent(cs, L->ev_out, 0));
  return HB_OK;
}

}  // extern "C"
