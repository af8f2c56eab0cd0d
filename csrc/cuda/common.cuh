// Shared device/host helpers for the dtfe_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#define DTFE_API extern "C" __attribute__((visibility("default")))

// The public C ABI; every entry point defined in csrc/cuda is declared there.
#pragma GCC visibility push(default)
#include "dtfe_b200.h"
#pragma GCC visibility pop

// Error plumbing: every entry point returns 0 on success; on failure the text is
// kept per thread and can be read back with dtfe_last_error().
void dtfe_set_error(const char* fmt, ...);

#define DTFE_CUDA_CHECK(expr)                                                        \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) {                                                         \
      dtfe_set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,            \
                     cudaGetErrorString(_e));                                        \
      return (int)_e;                                                                \
    }                                                                                \
  } while (0)

#define DTFE_REQUIRE(cond, ...)                                                      \
  do {                                                                               \
    if (!(cond)) {                                                                   \
      dtfe_set_error(__VA_ARGS__);                                                   \
      return -1;                                                                     \
    }                                                                                \
  } while (0)

#define DTFE_LAUNCH_CHECK() DTFE_CUDA_CHECK(cudaGetLastError())

namespace dtfe {

// ---- cache-bypassing accessors -------------------------------------------------
// Parameter shards can live in a peer GPU's HBM and are concurrently updated by
// other workers, so every read of shard memory bypasses L1 (L1 is not coherent
// across SMs, and for peer apertures it is the only local cache).
__device__ __forceinline__ float ld_relaxed_sys(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_relaxed_sys_v4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

// ---- the push primitive: fire-and-forget reductions into the owner shard -------
// red.global.add at .sys scope is performed by the L2 of the GPU that owns the
// address (the ps shard), so a worker's gradient push *is* the apply: no staging
// buffer, no ps-side kernel.  (Reference: worker.py:75-79 apply_gradients on the
// global variables; TF1 ApplyGradientDescent with use_locking=False.)
__device__ __forceinline__ void red_add_sys(float* p, float v) {
  asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" :: "l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void red_add_sys_v4(float* p, float4 v) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void red_add_sys_v2(float* p, float2 v) {
  asm volatile("red.relaxed.sys.global.add.v2.f32 [%0], {%1,%2};"
               :: "l"(p), "f"(v.x), "f"(v.y) : "memory");
}
// packed bf16x2 reduction (bf16 shards of the wide config)
__device__ __forceinline__ void red_add_sys_bf16x2(__nv_bfloat16* p, uint32_t packed) {
  asm volatile("red.relaxed.sys.global.add.noftz.bf16x2 [%0], %1;" :: "l"(p), "r"(packed) : "memory");
}
__device__ __forceinline__ void red_add_sys_v4_bf16x2(__nv_bfloat16* p, uint4 v) {
  asm volatile("red.relaxed.sys.global.add.noftz.v4.bf16x2 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ int atom_add_sys_s32(int* p, int v) {
  int old;
  asm volatile("atom.acq_rel.sys.global.add.s32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}
// Same counter bump with no ordering attached: later loads of the issuing thread do not wait for the
// atomic's round trip (over NVLink that is microseconds); the result is scoreboarded until first use.
__device__ __forceinline__ int atom_add_relaxed_sys_s32(int* p, int v) {
  int old;
  asm volatile("atom.relaxed.sys.global.add.s32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}
// 16-byte asynchronous global->shared copy that bypasses L1 (.cg).  No register (and no scoreboard
// slot) holds the data in flight, so a slow source (a peer GPU) cannot delay unrelated waits of the
// issuing warp.  Only `src_bytes` (<= 16) are read, the rest of the 16 bytes is zero-filled.
__device__ __forceinline__ void cp_async_cg16(void* smem_dst, const void* gsrc, int src_bytes = 16) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }
// Whole-CTA copy of n floats (both sides 16-byte aligned; the tail quad is read partially, never past n).
__device__ __forceinline__ void cta_copy_async(float* smem_dst, const float* gsrc, int n, int tid, int nthreads) {
  for (int f = tid; 4 * f < n; f += nthreads) {
    int rem = n - 4 * f;
    cp_async_cg16(smem_dst + 4 * f, gsrc + 4 * f, rem >= 4 ? 16 : 4 * rem);
  }
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ float sigmoidf_fast(float z) {
  return __fdividef(1.0f, 1.0f + __expf(-z));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace dtfe
