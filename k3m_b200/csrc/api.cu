// Error plumbing shared by every entry point of libk3m_sm100a.so (see include/k3m.h).
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void k3m_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int k3m_check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    k3m_set_error("%s: %s", what, cudaGetErrorString(e));
    return K3M_ERR_CUDA;
  }
  return K3M_OK;
}

extern "C" const char* k3m_last_error_string(void) { return g_err; }
extern "C" int k3m_get_version(void) { return 1; }

// Debug aid: cudaStreamIsCapturing status of `stream` (0 none, 1 active, 2 invalidated; < 0: the query itself failed).
extern "C" int k3m_debug_capture_status(void* stream) {
  cudaStreamCaptureStatus st;
  cudaError_t e = cudaStreamIsCapturing(reinterpret_cast<cudaStream_t>(stream), &st);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return -(int)e;
  }
  return (int)st;
}
