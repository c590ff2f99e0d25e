// Device runtime: error reporting, static arena, CUDA IPC import/export.
//
// Replaces the pieces of TensorFlow's runtime the reference configures at
// mpipy.py:69-71 (BFC allocator with growth): all shapes are static (mpipy.py:33-37),
// so one arena is allocated up front and carved by the Python side.  The IPC
// functions let peers map each other's arena so that collectives run as kernels
// over NVLink instead of host-staged MPI buffers (mpipy.py:121-127).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "common.cuh"

namespace {
thread_local std::string g_err;
}

int mtb_set_error(cudaError_t e, const char* expr, const char* file, int line) {
  char buf[512];
  snprintf(buf, sizeof(buf), "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, expr);
  g_err = buf;
  return MTB_E_CUDA;
}

int mtb_set_error_msg(int code, const char* msg, const char* file, int line) {
  char buf[512];
  snprintf(buf, sizeof(buf), "%s (%s:%d)", msg, file, line);
  g_err = buf;
  return code;
}

namespace mtb {
bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("MTB_PDL");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}
}  // namespace mtb

extern "C" {

const char* mtb_last_error(void) { return g_err.c_str(); }

int mtb_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem) {
  cudaDeviceProp p;
  MTB_CUDA_CHECK(cudaGetDeviceProperties(&p, device));
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (cc_major) *cc_major = p.major;
  if (cc_minor) *cc_minor = p.minor;
  if (total_mem) *total_mem = p.totalGlobalMem;
  return 0;
}

int mtb_arena_alloc(size_t bytes, void** out_ptr) {
  MTB_REQUIRE(out_ptr && bytes > 0, "arena_alloc: bad argument");
  MTB_CUDA_CHECK(cudaMalloc(out_ptr, bytes));
  MTB_CUDA_CHECK(cudaMemset(*out_ptr, 0, bytes));
  return 0;
}

int mtb_arena_free(void* ptr) {
  MTB_CUDA_CHECK(cudaFree(ptr));
  return 0;
}

int mtb_ipc_export(void* ptr, void* out_handle64) {
  cudaIpcMemHandle_t h;
  MTB_CUDA_CHECK(cudaIpcGetMemHandle(&h, ptr));
  static_assert(sizeof(h) == 64, "IPC handle is 64 bytes");
  memcpy(out_handle64, &h, 64);
  return 0;
}

int mtb_ipc_import(const void* handle64, void** out_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  MTB_CUDA_CHECK(cudaIpcOpenMemHandle(out_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int mtb_ipc_close(void* ptr) {
  MTB_CUDA_CHECK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

int mtb_enable_peer_access(int peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
  MTB_CUDA_CHECK(e);
  return 0;
}

}  // extern "C"
