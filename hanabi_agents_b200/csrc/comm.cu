// The learner's one exchange step over NVLink peer memory: an in-place sum of the flat fp32
// gradient bucket across the GPUs of one node ("ncclAllReduce(sum, fp32) of the flat
// gradient bucket once per update, fixed algorithm for run-to-run determinism").
//
// One process per GPU.  Every rank exports its bucket and one more allocation (barrier flags + a staging
// area as large as the bucket) with CUDA IPC and maps its peers'.  The kernel is a two-shot all-reduce made of
// PUSHES -- remote stores over NVLink, which are one-way, where the first version pulled with remote loads,
// each a round trip:
//   shot 1       rank r stores slice j of its bucket into slot r of rank j's staging area, for every j != r,
//                then raises flag[j][r] (release)
//   reduce       rank r waits for the W-1 flags, sums slice r -- its own bucket and the W-1 staged copies --
//                in fixed rank order 0..W-1 (every element is summed ONCE, by its owner: all ranks end up with
//                the same bits), and
//   shot 2       stores the result into slice r of EVERY rank's bucket, then raises flag2[j][r]
//   done         rank r waits for the W-1 flag2's: its bucket holds every reduced slice.
// Two one-way hops instead of three barriers of round trips, and no trailing barrier: nobody ever READS a
// peer's bucket, so a rank may overwrite its bucket as soon as its own call is through; a peer can only push
// into my staging area for call e+1 after it finished call e, which needed my flag2, which I raise after my
// reduction has read the staging area.  2 (W-1)/W of the bucket leaves each rank per call.
// Flags are per CTA: CTA b of every rank handles the same elements, so only CTAs with the same index on
// different GPUs wait for each other and no co-residency of a whole grid is assumed.  A wait for a late peer
// does NOT trap (a rank that is seconds late -- an evaluation on rank 0, a graph recapture, a first-use
// allocation -- is normal, and a trap would poison the CUDA context of every rank): it polls for up to ~2
// minutes, then raises a flag in host-mapped memory (hb_comm_status) and leaves so that the host can raise an
// error instead of hanging the device.  Ranks must call hb_comm_allreduce in lockstep, as with any collective.
#include "common.cuh"
#include <new>
#include <string.h>

namespace {
constexpr int kMaxWorld = 8;
constexpr int kCommBlocks = 144;      // most CTAs per rank (same on every rank; < 148 SMs, all resident)
constexpr int kCommThreads = 512;
// Independent CHANNELS: all-reduces of different channels may be in flight at the same time (the learner
// reduces every layer's gradient on that layer's stream, beside the backward pass); calls of one channel
// must be stream-ordered on every rank and made in the same order on all ranks.  Each channel has its own
// barrier flags and its own epoch counter IN DEVICE MEMORY (bumped by the kernel itself), so a call can sit
// in a replayed CUDA graph: nothing that changes from call to call is a kernel argument.
constexpr int kChannels = 8;
constexpr size_t kStageOffset = (size_t)1 << 20;   // flags (110 KB) and the staging area share one allocation
}  // namespace

struct hb_comm {
  int device = 0, rank = 0, world = 1;
  float* own = nullptr;               // this rank's bucket (not owned)
  int64_t n = 0;
  uint32_t* flags = nullptr;          // [kChannels][3 barriers][kCommBlocks][kMaxWorld] epoch counters (owned, exported)
  uint32_t* epochs = nullptr;         // [kChannels] {calls completed, CTAs done} pairs (device, local)
  float* peer_buf[kMaxWorld] = {nullptr};
  uint32_t* peer_flags[kMaxWorld] = {nullptr};   // staging of rank j = (float*)((char*)peer_flags[j] + kStageOffset)
  bool opened = false;
  uint32_t epoch = 0;
  int* h_timeout = nullptr;           // pinned + mapped: set by a barrier that gave up waiting
  int* d_timeout = nullptr;
};

namespace {

struct CommArgs {
  float* buf[kMaxWorld];
  uint32_t* flags[kMaxWorld];
  int rank, world;
  long long off4, n4;                 // range of the bucket, in float4 elements
  long long chan4;                    // float4 elements of staging per channel (bucket + slack)
  int channel;
  uint32_t* epoch;                    // this channel's {calls completed, CTAs done}
  int* timeout;                       // host-mapped flag
};

// peer data: system-scope loads that never come from this SM's L1
__device__ __forceinline__ float4 ld_peer(const float4* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_flag(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_flag(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void st_flag_relaxed(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_flag_relaxed(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Wait until peer `peer`'s flag of (stage, this CTA) has reached `epoch` (bounded, see the header).
__device__ __forceinline__ void wait_flag(const CommArgs& a, int stage, int peer, uint32_t epoch) {
  const size_t slot = (((size_t)a.channel * 3 + stage) * kCommBlocks + blockIdx.x) * kMaxWorld;
  const uint32_t* mine = a.flags[a.rank] + slot + peer;
  const long long t0 = clock64();
  unsigned spins = 0;
  while ((int32_t)(ld_flag_relaxed(mine) - epoch) < 0) {
    // every 2^16 polls: has the bound (~2 min at 1.9 GHz) passed, or has another CTA given up?
    // (the status word is host memory: one PCIe read, so it is not part of the hot poll)
    if ((++spins & 0xffffu) == 0 &&
        (clock64() - t0 > 240000000000ll || *(volatile int*)a.timeout != 0)) {
      *(volatile int*)a.timeout = 1;                    // the host reports it (hb_comm_status)
      break;
    }
  }
}
__device__ __forceinline__ void raise_flag(const CommArgs& a, int stage, int peer, uint32_t epoch) {
  const size_t slot = (((size_t)a.channel * 3 + stage) * kCommBlocks + blockIdx.x) * kMaxWorld;
  st_flag(a.flags[peer] + slot + a.rank, epoch);        // release: the CTA's stores before the barrier first
}
// every channel has a staging area of its own (calls of different channels overlap in time and the slots of
// a range reach up to W - 1 elements past its end)
__device__ __forceinline__ float4* stage_of(const CommArgs& a, int j) {
  return reinterpret_cast<float4*>(reinterpret_cast<char*>(a.flags[j]) + kStageOffset) + a.channel * a.chan4;
}

template <int W>
__global__ void __launch_bounds__(kCommThreads)
allreduce_push_kernel(CommArgs a) {
  const int r = a.rank;
  const long long per = (a.n4 + W - 1) / W;               // slice length in float4
  // this call's number on its channel: one more than the calls completed (the counter moves only when the
  // last CTA of a call leaves, and calls of a channel are stream-ordered)
  const uint32_t epoch = *(volatile uint32_t*)a.epoch + 1u;
  const long long stride = (long long)gridDim.x * kCommThreads;
  const long long t = (long long)blockIdx.x * kCommThreads + threadIdx.x;
  const long long end4 = a.off4 + a.n4;
  const float4* own = reinterpret_cast<const float4*>(a.buf[r]);
  // ---- shot 1: my copy of slice j -> slot r of rank j's staging area (element k of the slice at
  // off4 + r * per + k; W * per <= n4 + W, the staging area has that slack)
#pragma unroll 1
  for (int jj = 1; jj < W; ++jj) {
    const int j = (r + jj) % W;                            // spread the peers over the links
    const long long lo = a.off4 + per * j;
    const long long len = lo + per < end4 ? per : (end4 > lo ? end4 - lo : 0);
    float4* dst = stage_of(a, j) + a.off4 + per * r;
    for (long long k = t; k < len; k += 4 * stride) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (k + u * stride < len) v[u] = own[lo + k + u * stride];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (k + u * stride < len) dst[k + u * stride] = v[u];
    }
  }
  __syncthreads();
  if (threadIdx.x < W && threadIdx.x != r) raise_flag(a, 0, threadIdx.x, epoch);
  // ---- reduce slice r in rank order, store it into every rank's bucket
  if (threadIdx.x < W && threadIdx.x != r) wait_flag(a, 0, threadIdx.x, epoch);
  if (threadIdx.x < W) asm volatile("fence.acq_rel.sys;" ::: "memory");
  __syncthreads();
  {
    const long long lo = a.off4 + per * r;
    const long long len = lo + per < end4 ? per : (end4 > lo ? end4 - lo : 0);
    const float4* st = stage_of(a, r) + a.off4;
    for (long long k = t; k < len; k += stride) {
      float4 s;
#pragma unroll
      for (int p = 0; p < W; ++p) {
        // staged copies were written by remote stores: read them past the L1 (ld.volatile)
        const float4 v = p == r ? own[lo + k] : ld_peer(st + per * p + k);
        if (p == 0) s = v;
        else { s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
      }
#pragma unroll
      for (int j = 0; j < W; ++j) reinterpret_cast<float4*>(a.buf[j])[lo + k] = s;
    }
  }
  __syncthreads();
  if (threadIdx.x < W && threadIdx.x != r) raise_flag(a, 1, threadIdx.x, epoch);
  // ---- done when every other owner's slice has landed in my bucket
  if (threadIdx.x < W && threadIdx.x != r) wait_flag(a, 1, threadIdx.x, epoch);
  if (threadIdx.x < W) asm volatile("fence.acq_rel.sys;" ::: "memory");
  __syncthreads();
  // the last CTA to leave completes the call: CTAs-done back to 0, calls-completed + 1
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(a.epoch + 1, 1u) == gridDim.x - 1) {
      a.epoch[1] = 0u;
      __threadfence();
      *(volatile uint32_t*)a.epoch = epoch;
    }
  }
}

}  // namespace

extern "C" {

int hb_comm_destroy(hb_comm* c);

int hb_comm_create(int rank, int world, int device, float* d_bucket, int64_t n, hb_comm** out) {
  HB_REQUIRE(out != nullptr && d_bucket != nullptr, "hb_comm_create: NULL argument");
  HB_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, "hb_comm_create: rank %d of %d",
             rank, world);
  HB_REQUIRE(n > 0 && n % 4 == 0 && ((uintptr_t)d_bucket & 15) == 0,
             "hb_comm_create: the bucket must be 16-byte aligned with a multiple of 4 elements");
  hb::DeviceGuard g(device);
  HB_REQUIRE(g.ok, "hb_comm_create: cannot select device %d", device);
  hb_comm* c = new (std::nothrow) hb_comm();
  HB_REQUIRE(c != nullptr, "hb_comm_create: out of host memory");
  c->device = device; c->rank = rank; c->world = world; c->own = d_bucket; c->n = n;
  const size_t fbytes = (size_t)kChannels * 3 * kCommBlocks * kMaxWorld * sizeof(uint32_t);
  static_assert((size_t)kChannels * 3 * kCommBlocks * kMaxWorld * sizeof(uint32_t) <= kStageOffset, "flags overlap the staging area");
  // one block of its own: the IPC handle then maps exactly this allocation on the peers.  Flags first, the
  // staging area (the bucket's size + slack for the rounded-up slices) from kStageOffset on.
  const size_t total = kStageOffset + (size_t)kChannels * ((size_t)n + 16 * kMaxWorld) * sizeof(float);
  cudaError_t e = cudaMalloc(&c->flags, total < ((size_t)2 << 20) ? ((size_t)2 << 20) : total);
  if (e == cudaSuccess) e = cudaMemset(c->flags, 0, fbytes);
  if (e == cudaSuccess) e = cudaMalloc(&c->epochs, kChannels * 2 * sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaMemset(c->epochs, 0, kChannels * 2 * sizeof(uint32_t));
  if (e != cudaSuccess) { delete c; return hb::cuda_fail(e, "hb_comm_create alloc", __FILE__, __LINE__); }
  if (cudaHostAlloc((void**)&c->h_timeout, 64, cudaHostAllocMapped) != cudaSuccess ||
      cudaHostGetDevicePointer((void**)&c->d_timeout, c->h_timeout, 0) != cudaSuccess) {
    cudaFree(c->flags);
    delete c;
    hb::set_error("hb_comm_create: cannot allocate the mapped status word");
    return HB_ERR_CUDA;
  }
  *c->h_timeout = 0;
  c->peer_buf[rank] = d_bucket;
  c->peer_flags[rank] = c->flags;
  *out = c;
  return HB_OK;
}

/* 128 bytes: the IPC handles of the bucket and of the flag array, to be all-gathered by the caller. */
int hb_comm_export(hb_comm* c, void* h_out128) {
  HB_REQUIRE(c && h_out128, "hb_comm_export: NULL argument");
  hb::DeviceGuard g(c->device);
  cudaIpcMemHandle_t hb_, hf;
  HB_CUDA(cudaIpcGetMemHandle(&hb_, c->own));
  HB_CUDA(cudaIpcGetMemHandle(&hf, c->flags));
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  memcpy(h_out128, &hb_, 64);
  memcpy((char*)h_out128 + 64, &hf, 64);
  return HB_OK;
}

/* h_all: world x 128 bytes in rank order (what hb_comm_export produced on every rank). */
int hb_comm_open(hb_comm* c, const void* h_all) {
  HB_REQUIRE(c && h_all, "hb_comm_open: NULL argument");
  HB_REQUIRE(!c->opened, "hb_comm_open: already opened");
  hb::DeviceGuard g(c->device);
  for (int j = 0; j < c->world; ++j) {
    if (j == c->rank) continue;
    cudaIpcMemHandle_t hb_, hf;
    memcpy(&hb_, (const char*)h_all + (size_t)j * 128, 64);
    memcpy(&hf, (const char*)h_all + (size_t)j * 128 + 64, 64);
    HB_CUDA(cudaIpcOpenMemHandle((void**)&c->peer_buf[j], hb_, cudaIpcMemLazyEnablePeerAccess));
    HB_CUDA(cudaIpcOpenMemHandle((void**)&c->peer_flags[j], hf, cudaIpcMemLazyEnablePeerAccess));
  }
  c->opened = true;
  return HB_OK;
}

int hb_comm_destroy(hb_comm* c) {
  if (c == nullptr) return HB_OK;
  hb::DeviceGuard g(c->device);
  if (c->opened)
    for (int j = 0; j < c->world; ++j) {
      if (j == c->rank) continue;
      if (c->peer_buf[j]) cudaIpcCloseMemHandle(c->peer_buf[j]);
      if (c->peer_flags[j]) cudaIpcCloseMemHandle(c->peer_flags[j]);
    }
  if (c->flags) cudaFree(c->flags);
  if (c->epochs) cudaFree(c->epochs);
  if (c->h_timeout) cudaFreeHost(c->h_timeout);
  delete c;
  return HB_OK;
}

/* In-place sum of bucket[offset, offset + n) over all ranks on `channel` (0..7).  Calls of one channel must
 * be stream-ordered on every rank and come in the same order on all ranks; different channels are
 * independent and may overlap.  Stream ordered, capturable into a CUDA graph (no per-call state in the
 * arguments); the result is identical on every rank (fixed summation order). */
int hb_comm_allreduce_range(hb_comm* c, int channel, int64_t offset, int64_t n, void* stream) {
  HB_REQUIRE(c != nullptr, "hb_comm_allreduce: NULL communicator");
  HB_REQUIRE(channel >= 0 && channel < kChannels, "hb_comm_allreduce: channel %d (0..%d)", channel, kChannels - 1);
  HB_REQUIRE(offset >= 0 && n > 0 && offset + n <= c->n && offset % 4 == 0 && n % 4 == 0,
             "hb_comm_allreduce: range [%lld, +%lld) of a %lld-element bucket must be 16-byte aligned",
             (long long)offset, (long long)n, (long long)c->n);
  if (c->world == 1) return HB_OK;
  if (!c->opened) { hb::set_error("hb_comm_allreduce: peers are not mapped (hb_comm_open)"); return HB_ERR_STATE; }
  hb::DeviceGuard g(c->device);
  CommArgs a;
  for (int j = 0; j < kMaxWorld; ++j) { a.buf[j] = c->peer_buf[j]; a.flags[j] = c->peer_flags[j]; }
  a.rank = c->rank; a.world = c->world; a.off4 = offset / 4; a.n4 = n / 4;
  a.chan4 = (c->n + 16 * kMaxWorld) / 4;
  a.channel = channel;
  a.epoch = c->epochs + 2 * channel;
  a.timeout = c->d_timeout;
  ++c->epoch;
  if (*(volatile int*)c->h_timeout != 0) {
    hb::set_error("hb_comm_allreduce: an earlier all-reduce gave up waiting for a peer (ranks out of lockstep)");
    return HB_ERR_STATE;
  }
  // CTAs: enough for ~4 float4 per thread of this rank's slice, the same number on every rank
  const long long per = (a.n4 + c->world - 1) / c->world;
  long long want = (per + 4 * kCommThreads - 1) / (4 * kCommThreads);
  const unsigned blocks = (unsigned)(want < 16 ? 16 : (want > kCommBlocks ? kCommBlocks : want));
  cudaStream_t st = (cudaStream_t)stream;
  switch (c->world) {
    case 2: allreduce_push_kernel<2><<<blocks, kCommThreads, 0, st>>>(a); break;
    case 3: allreduce_push_kernel<3><<<blocks, kCommThreads, 0, st>>>(a); break;
    case 4: allreduce_push_kernel<4><<<blocks, kCommThreads, 0, st>>>(a); break;
    case 5: allreduce_push_kernel<5><<<blocks, kCommThreads, 0, st>>>(a); break;
    case 6: allreduce_push_kernel<6><<<blocks, kCommThreads, 0, st>>>(a); break;
    case 7: allreduce_push_kernel<7><<<blocks, kCommThreads, 0, st>>>(a); break;
    default: allreduce_push_kernel<8><<<blocks, kCommThreads, 0, st>>>(a); break;
  }
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

/* The whole bucket, on the last channel (the learner's per-layer calls use channels 0 .. layers - 1). */
int hb_comm_allreduce(hb_comm* c, void* stream) {
  HB_REQUIRE(c != nullptr, "hb_comm_allreduce: NULL communicator");
  return hb_comm_allreduce_range(c, kChannels - 1, 0, c->n, stream);
}

/* HB_OK, or HB_ERR_STATE when a barrier of an earlier all-reduce timed out waiting for a peer. */
int hb_comm_status(const hb_comm* c) {
  HB_REQUIRE(c != nullptr, "hb_comm_status: NULL communicator");
  if (c->h_timeout && *(volatile int*)c->h_timeout != 0) {
    hb::set_error("all-reduce: a peer did not arrive within the wait bound (ranks out of lockstep)");
    return HB_ERR_STATE;
  }
  return HB_OK;
}

int hb_comm_info(const hb_comm* c, int what) {
  if (c == nullptr) return -1;
  switch (what) {
    case 0: return c->rank;
    case 1: return c->world;
    case 2: return c->opened ? 1 : 0;
    case 3: return (int)c->epoch;
    default: return -1;
  }
}

}  // extern "C"
