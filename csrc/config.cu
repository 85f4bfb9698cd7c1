// Library-wide switches.  The only one so far is the deterministic-reduction mode: a debugging /
// regression-testing aid with no counterpart call in the reference (which inherits
// torch.use_deterministic_algorithms from its substrate and never sets it).
#include <atomic>
#include <stdlib.h>
#include "common.cuh"
#include "../include/b200dt.h"

namespace b200 {

static std::atomic<int> g_det{-1};

int deterministic() {
  int v = g_det.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("B200DT_DETERMINISTIC");
    v = (e != nullptr && e[0] != '\0' && e[0] != '0') ? 1 : 0;
    g_det.store(v, std::memory_order_relaxed);
  }
  return v;
}

int bn_replicas() { return deterministic() ? B200DT_BN_REPLICAS_DET : B200DT_BN_REPLICAS; }

// Copies actually spread over by the atomics of a C-channel reduction: the contention per address falls with
// the channel count (the CTAs are divided among more column groups), and the last CTA's fold costs
// 2 * C * copies loads, so wide layers use few copies (2048 channels: one, and no fold at all).
int bn_replicas_used(int C) {
  if (deterministic()) return B200DT_BN_REPLICAS_DET;
  int r = 2048 / (C > 0 ? C : 1);
  if (r < 1) r = 1;
  return r > B200DT_BN_REPLICAS ? B200DT_BN_REPLICAS : r;
}

static std::atomic<int> g_pdl{-1};
int pdl_enabled() {
  int v = g_pdl.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("B200DT_PDL");
    v = (e != nullptr && e[0] != '\0') ? atoi(e) : 1;
    g_pdl.store(v, std::memory_order_relaxed);
  }
  return v;
}

static std::atomic<int> g_prezeroed{0};
int sums_prezeroed() { return g_prezeroed.load(std::memory_order_relaxed); }

}  // namespace b200

extern "C" void b200dt_set_deterministic(int on) { b200::g_det.store(on ? 1 : 0, std::memory_order_relaxed); }
extern "C" int b200dt_get_deterministic(void) { return b200::deterministic(); }
extern "C" int b200dt_bn_replicas(void) { return b200::bn_replicas(); }
extern "C" int b200dt_bn_replicas_used(int C) { return b200::bn_replicas_used(C); }
extern "C" void b200dt_set_sums_prezeroed(int on) { b200::g_prezeroed.store(on ? 1 : 0, std::memory_order_relaxed); }
extern "C" int b200dt_get_sums_prezeroed(void) { return b200::sums_prezeroed(); }
extern "C" unsigned long long b200dt_capture_id(void* stream) {
  cudaStreamCaptureStatus status = cudaStreamCaptureStatusNone;
  unsigned long long id = 0;
  if (cudaStreamGetCaptureInfo((cudaStream_t)stream, &status, &id) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return status == cudaStreamCaptureStatusActive ? id : 0;
}
