#include "common.cuh"
#include <string.h>

namespace hb {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace hb

// ---- stream gate --------------------------------------------------------------------------------
struct hb_gate {
  int device = 0;
  volatile int* h_flag = nullptr;   // pinned, mapped
  int* d_flag = nullptr;            // the same word as the device sees it
};

namespace {
// One thread polls a word of mapped host memory; everything enqueued behind it on the stream waits.
// The wait is bounded (~20 s of polling) so that a host that never opens the gate cannot hang the
// device: the kernel then simply ends.
__global__ void gate_kernel(const int* flag) {
  for (long long spins = 0; spins < (1ll << 26); ++spins) {
    int v;
    asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
    if (v != 0) break;
    __nanosleep(256);
  }
}
}  // namespace

extern "C" {
int hb_gate_create(int device, hb_gate** out) {
  HB_REQUIRE(out != nullptr, "hb_gate_create: out is NULL");
  hb::DeviceGuard g(device);
  if (!g.ok) { hb::set_error("hb_gate_create: cannot select CUDA device %d", device); return HB_ERR_CUDA; }
  hb_gate* G = new hb_gate();
  G->device = device;
  void* h = nullptr;
  cudaError_t e = cudaHostAlloc(&h, 64, cudaHostAllocMapped);
  if (e == cudaSuccess) e = cudaHostGetDevicePointer((void**)&G->d_flag, h, 0);
  if (e != cudaSuccess) {
    if (h) cudaFreeHost(h);
    delete G;
    return hb::cuda_fail(e, "hb_gate_create", __FILE__, __LINE__);
  }
  G->h_flag = (volatile int*)h;
  *G->h_flag = 1;
  *out = G;
  return HB_OK;
}

int hb_gate_destroy(hb_gate* G) {
  if (!G) return HB_OK;
  hb::DeviceGuard g(G->device);
  *G->h_flag = 1;
  cudaDeviceSynchronize();
  cudaFreeHost((void*)G->h_flag);
  delete G;
  return HB_OK;
}

int hb_gate_close(hb_gate* G, void* stream) {
  HB_REQUIRE(G != nullptr, "hb_gate_close: NULL gate");
  hb::DeviceGuard g(G->device);
  *G->h_flag = 0;
  __sync_synchronize();
  gate_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(G->d_flag);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_gate_open(hb_gate* G) {
  HB_REQUIRE(G != nullptr, "hb_gate_open: NULL gate");
  __sync_synchronize();
  *G->h_flag = 1;
  __sync_synchronize();
  return HB_OK;
}

const char* hb_last_error(void) { return hb::g_err; }
int hb_version(void) { return 100; }
const char* hb_build_arch(void) { return "sm_100a"; }
}
