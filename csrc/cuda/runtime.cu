// Runtime services of the CUDA library: error text, device queries, parameter-shard
// allocation in HBM, cross-process peer mapping (CUDA IPC), L2 flush, step counter.
//
// Replaces the TF1 worker-service resource manager that held ps variables in the ps
// process's host RAM (reference main.py:67-68): here a "ps task" owns a
// cudaMalloc'd shard, exports an IPC handle, and every worker maps it.
#include "common.cuh"
#include <stdarg.h>
#include <string.h>

static thread_local char g_err[1024] = "";

void dtfe_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

DTFE_API const char* dtfe_last_error() { return g_err; }

DTFE_API int dtfe_abi_version() { return 1; }

// sizeof of the structs that cross the C ABI, so a binding can verify its mirror of the layout:
// out[0] = dtfe_mlp2_launch, out[1] = dtfe_sync_exchange, out[2] = dtfe_gemm_tc.
DTFE_API int dtfe_struct_sizes(size_t* out, int n) {
  const size_t sizes[3] = {sizeof(dtfe_mlp2_launch), sizeof(dtfe_sync_exchange), sizeof(dtfe_gemm_tc)};
  for (int i = 0; i < n && i < 3; ++i) out[i] = sizes[i];
  return 3;
}

DTFE_API int dtfe_device_count(int* n) {
  DTFE_CUDA_CHECK(cudaGetDeviceCount(n));
  return 0;
}

// out[0]=major out[1]=minor out[2]=sm_count out[3]=max_smem_optin out[4]=l2_bytes out[5]=coop_launch
DTFE_API int dtfe_device_info(int dev, int* out) {
  cudaDeviceProp p;
  DTFE_CUDA_CHECK(cudaGetDeviceProperties(&p, dev));
  out[0] = p.major;
  out[1] = p.minor;
  out[2] = p.multiProcessorCount;
  out[3] = (int)p.sharedMemPerBlockOptin;
  out[4] = p.l2CacheSize;
  out[5] = p.cooperativeLaunch;
  return 0;
}

// ---- shard allocation ----------------------------------------------------------
DTFE_API int dtfe_shard_alloc(size_t bytes, void** out) {
  DTFE_REQUIRE(bytes > 0, "dtfe_shard_alloc: bytes must be > 0");
  DTFE_CUDA_CHECK(cudaMalloc(out, bytes));
  DTFE_CUDA_CHECK(cudaMemset(*out, 0, bytes));
  return 0;
}
DTFE_API int dtfe_shard_free(void* p) {
  DTFE_CUDA_CHECK(cudaFree(p));
  return 0;
}
// 64-byte opaque handle that another process on the same box can open.
DTFE_API int dtfe_ipc_export(void* p, void* handle64) {
  cudaIpcMemHandle_t h;
  DTFE_CUDA_CHECK(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return 0;
}
DTFE_API int dtfe_ipc_open(const void* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DTFE_CUDA_CHECK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
DTFE_API int dtfe_ipc_close(void* p) {
  DTFE_CUDA_CHECK(cudaIpcCloseMemHandle(p));
  return 0;
}
// In-process peer mapping (one process driving several GPUs, e.g. tests).
DTFE_API int dtfe_enable_peer(int dev, int peer) {
  int can = 0;
  DTFE_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, dev, peer));
  DTFE_REQUIRE(can, "device %d cannot access peer %d", dev, peer);
  int cur;
  DTFE_CUDA_CHECK(cudaGetDevice(&cur));
  DTFE_CUDA_CHECK(cudaSetDevice(dev));
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
  cudaSetDevice(cur);
  DTFE_CUDA_CHECK(e);
  return 0;
}

// Copy-engine pull: one asynchronous device-to-device copy (cudaMemcpyDefault: source may be a peer-mapped shard).
// Unlike dtfe_pull_f32 it takes no SM away from the GEMMs that run next to it (reference sync_op, worker.py:81-85).
DTFE_API int dtfe_memcpy_async(void* dst, const void* src, size_t bytes, void* stream) {
  DTFE_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, (cudaStream_t)stream));
  return 0;
}

// ---- L2 flush (bench hygiene): overwrite a buffer larger than L2 ----------------
__global__ void k_fill_u32(uint32_t* p, size_t n, uint32_t v) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) p[i] = v;
}
DTFE_API int dtfe_l2_flush(void* buf, size_t bytes, void* stream) {
  k_fill_u32<<<148 * 8, 256, 0, (cudaStream_t)stream>>>((uint32_t*)buf, bytes / 4, 0u);
  DTFE_LAUNCH_CHECK();
  return 0;
}

// ---- global_step (reference worker.py:29-32,141) ----------------------
// int32 scalar living in shard 0; +1 per applied update with a system-scope atomic,
// returning the pre-increment value (what the reference's worker fetches as `step`).
__global__ void k_step_add(int* gs, int inc, int* observed) {
  int old = dtfe::atom_add_sys_s32(gs, inc);
  if (observed) *observed = old;
}
DTFE_API int dtfe_global_step_add(int* gs, int inc, int* observed_or_null, void* stream) {
  k_step_add<<<1, 1, 0, (cudaStream_t)stream>>>(gs, inc, observed_or_null);
  DTFE_LAUNCH_CHECK();
  return 0;
}
