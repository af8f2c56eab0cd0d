// Parameter-server data movement over peer memory, as standalone kernels:
//   * async push + SGD apply  : shard[i] += alpha*grad[i] with fire-and-forget red.add.sys
//   * Adam                    : push accumulates into an owner-side slot; owner applies
//   * pull                    : shard -> local copy with L1-bypassing vector loads
//   * sync reduce-scatter+apply: every owner reduces its slice of all workers' gradient
//                                buffers straight out of peer HBM and applies it, with
//                                flag barriers in peer memory (no NCCL, no staging)
//
// Reference mechanism replaced: gRPC _Send/_Recv of 4 gradient tensors to the ps and a
// CPU ApplyGradientDescent there (worker.py:75-79), and the 4 Assign ops of the pull
// (worker.py:81-85).  The fused-epilogue forms live in dense_simt.cu / gemm_tc.cu /
// mlp2_step.cu; these flat kernels are the bandwidth-sweep subjects and the sync path.
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256) k_push_sgd_f32(float* __restrict__ shard,
                                                      const float* __restrict__ grad, size_t n,
                                                      float alpha, int atomic) {
  size_t nv = n / 4;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  const float4* g4 = reinterpret_cast<const float4*>(grad);
  for (size_t v = i; v < nv; v += stride) {
    float4 g = __ldg(g4 + v);
    g.x *= alpha; g.y *= alpha; g.z *= alpha; g.w *= alpha;
    if (atomic) {
      dtfe::red_add_sys_v4(shard + v * 4, g);
    } else {
      float4 w = dtfe::ld_relaxed_sys_v4(shard + v * 4);
      w.x += g.x; w.y += g.y; w.z += g.z; w.w += g.w;
      *reinterpret_cast<float4*>(shard + v * 4) = w;
    }
  }
  for (size_t e = nv * 4 + i; e < n; e += stride) {
    float g = alpha * grad[e];
    if (atomic) dtfe::red_add_sys(shard + e, g);
    else shard[e] = dtfe::ld_relaxed_sys(shard + e) + g;
  }
}

__global__ void __launch_bounds__(256) k_push_sgd_bf16(__nv_bfloat16* __restrict__ shard,
                                                       const float* __restrict__ grad, size_t n,
                                                       float alpha) {
  size_t nv = n / 8;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  const float4* g4 = reinterpret_cast<const float4*>(grad);
  for (size_t v = i; v < nv; v += stride) {
    float4 a = __ldg(g4 + 2 * v), b = __ldg(g4 + 2 * v + 1);
    uint4 p;
    p.x = dtfe::pack_bf16x2(alpha * a.x, alpha * a.y);
    p.y = dtfe::pack_bf16x2(alpha * a.z, alpha * a.w);
    p.z = dtfe::pack_bf16x2(alpha * b.x, alpha * b.y);
    p.w = dtfe::pack_bf16x2(alpha * b.z, alpha * b.w);
    dtfe::red_add_sys_v4_bf16x2(shard + v * 8, p);
  }
  for (size_t e = nv * 8 + i * 2; e + 1 < n; e += stride * 2)
    dtfe::red_add_sys_bf16x2(shard + e, dtfe::pack_bf16x2(alpha * grad[e], alpha * grad[e + 1]));
}

__global__ void __launch_bounds__(256) k_pull_f32(float* __restrict__ dst, const float* __restrict__ shard,
                                                  size_t n) {
  size_t nv = n / 4;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t v = i; v < nv; v += stride)
    *reinterpret_cast<float4*>(dst + v * 4) = dtfe::ld_relaxed_sys_v4(shard + v * 4);
  for (size_t e = nv * 4 + i; e < n; e += stride) dst[e] = dtfe::ld_relaxed_sys(shard + e);
}

// Adam: workers red.add raw gradients into `slot` (owner HBM); the owner applies and clears.
__global__ void __launch_bounds__(256) k_adam_apply(float* __restrict__ p, float* __restrict__ m,
                                                    float* __restrict__ v, float* __restrict__ slot,
                                                    size_t n, float lr_t, float b1, float b2, float eps,
                                                    float gscale) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    // swap the mailbox out in ONE atomic: a push that lands between a separate load and store would be lost
    float g = atomicExch_system(slot + i, 0.f) * gscale;
    float mi = b1 * m[i] + (1.f - b1) * g;
    float vi = b2 * v[i] + (1.f - b2) * g * g;
    m[i] = mi; v[i] = vi;
    p[i] -= lr_t * mi / (sqrtf(vi) + eps);
  }
}

// ---- synchronous mode: reduce-scatter fused with the update --------------------------
constexpr int kMaxRanks = 16;
struct SyncArgs {
  const float* grads[kMaxRanks];  // every rank's gradient buffer (peer-mapped), full model length
  unsigned* flags[kMaxRanks];     // every rank's flag array (peer-mapped): [2][kMaxRanks] u32
  // Up to kMaxSeg owned variables: flat-gradient range [begin,end) of each and where its parameters
  // live (this rank's shard, local HBM); parameter element i of segment s is param[s][i - begin[s]].
  int nseg;
  float* param[8];
  size_t begin[8], end[8];
  float alpha;                    // -lr / W
  int rank, world;
  unsigned epoch;                 // strictly increasing per call, starts at 1
  unsigned* local_sync;           // [4] u32 in local HBM: {go, done_count, _, _}
  int* global_step;               // non-null on the owner of shard 0
  const unsigned* abort_flag;     // optional: a peer-visible word that turns non-zero when a rank leaves
  unsigned long long timeout_ns;  // give up spinning after this long (0 = never)
};

__device__ __forceinline__ unsigned long long gtimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Spin until *flag >= epoch.  Returns false if the wait was abandoned: a rank announced it is leaving
// (abort_flag, e.g. the shard header's stop word set by a worker's signal path) or the timeout expired.
// An abandoned wait must never hang the GPU: the kernel then skips the update and reports it.
__device__ __forceinline__ bool spin_until(const unsigned* flag, unsigned epoch, const SyncArgs& a,
                                           unsigned long long t0) {
  unsigned it = 0;
  while (dtfe::ld_acquire_sys(flag) < epoch) {
    __nanosleep(20);
    if ((++it & 0x3ff) == 0) {
      if (a.abort_flag && dtfe::ld_acquire_sys(a.abort_flag) != 0) return false;
      if (a.timeout_ns && gtimer_ns() - t0 > a.timeout_ns) return false;
    }
  }
  return true;
}

__global__ void __launch_bounds__(256) k_sync_rs_apply(SyncArgs a) {
  // local_sync words: [0] go (epoch once all gradients are in, 0xffffffff on abort) [1] done count
  //                   [2] abort status of the last call (0 ok, 1 entry barrier, 2 exit barrier)
  const unsigned long long t0 = gtimer_ns();
  // epoch 0 = take it from device memory (last completed epoch + 1), so a captured CUDA graph can be
  // replayed step after step without a host-side argument changing.
  if (a.epoch == 0) a.epoch = *reinterpret_cast<volatile unsigned*>(a.local_sync + 3) + 1u;
  // (a) arrive: tell every peer "my gradients for this epoch are complete", then wait for all.
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    __threadfence_system();
    for (int p = 0; p < a.world; ++p) dtfe::st_release_sys(a.flags[p] + a.rank, a.epoch);
    bool ok = true;
    for (int q = 0; q < a.world && ok; ++q) ok = spin_until(a.flags[a.rank] + q, a.epoch, a, t0);
    a.local_sync[2] = ok ? 0u : 1u;
    asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(a.local_sync), "r"(ok ? a.epoch : 0xffffffffu) : "memory");
  }
  __shared__ unsigned s_go;
  if (threadIdx.x == 0) {
    unsigned v;
    do {
      asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a.local_sync) : "memory");
    } while (v < a.epoch);
    __threadfence();
    s_go = v;
  }
  __syncthreads();
  if (s_go == 0xffffffffu) {            // abandoned: leave parameters untouched
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned prev = atomicAdd(a.local_sync + 1, 1u);
      if (prev == gridDim.x - 1) { a.local_sync[1] = 0; a.local_sync[0] = a.epoch; a.local_sync[3] = a.epoch; }
    }
    return;
  }

  // (b) reduce my slices over all ranks' buffers (peer loads) and apply in place.
  size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int sg = 0; sg < a.nseg; ++sg) {
    const size_t begin = a.begin[sg], n = a.end[sg] - begin;
    float* param = a.param[sg];
    bool vec_ok = (begin % 4 == 0) && ((reinterpret_cast<uintptr_t>(param) & 15) == 0);
    size_t nv = vec_ok ? n / 4 : 0;
    for (size_t v = tid; v < nv; v += stride) {
      size_t i = begin + v * 4;
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int q = 0; q < a.world; ++q) {
        float4 g = dtfe::ld_relaxed_sys_v4(a.grads[q] + i);
        s.x += g.x; s.y += g.y; s.z += g.z; s.w += g.w;
      }
      float4* w = reinterpret_cast<float4*>(param + v * 4);
      float4 o = *w;
      o.x += a.alpha * s.x; o.y += a.alpha * s.y; o.z += a.alpha * s.z; o.w += a.alpha * s.w;
      *w = o;
    }
    for (size_t e = nv * 4 + tid; e < n; e += stride) {
      float s = 0.f;
      for (int q = 0; q < a.world; ++q) s += dtfe::ld_relaxed_sys(a.grads[q] + begin + e);
      param[e] += a.alpha * s;
    }
  }

  // (c) depart: once the whole grid has applied, publish and wait until every owner has too,
  // so that on kernel exit all shards hold the new parameters and no gradient buffer is still
  // being read by a peer.  One gpu-scope fence per CTA chains the CTAs through the done counter;
  // only the CTA that arrives last pays the system-scope fence before signalling the peers.
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned prev = atomicAdd(a.local_sync + 1, 1u);
    if (prev == gridDim.x - 1) {
      a.local_sync[1] = 0;
      a.local_sync[3] = a.epoch;
      __threadfence_system();
      if (a.global_step) dtfe::atom_add_sys_s32(a.global_step, 1);
      for (int p = 0; p < a.world; ++p)
        dtfe::st_release_sys(a.flags[p] + kMaxRanks + a.rank, a.epoch);
      bool ok = true;
      for (int q = 0; q < a.world && ok; ++q) ok = spin_until(a.flags[a.rank] + kMaxRanks + q, a.epoch, a, t0);
      if (!ok) a.local_sync[2] = 2u;
    }
  }
}

// Batch gather with a device-side cursor: rows idx[cursor*B .. +B) of the resident dataset -> a dense
// batch.  The cursor lives in device memory so the launch can sit in a replayed CUDA graph.
__global__ void __launch_bounds__(256) k_gather_rows(const float* __restrict__ X, long ldx,
                                                     const int* __restrict__ labels, const int* __restrict__ idx,
                                                     const int* __restrict__ cursor, int B, int D,
                                                     float* __restrict__ Xout, int* __restrict__ Yout) {
  const long base = (long)(*cursor) * B;
  const int r = blockIdx.x;
  const long row = idx[base + r];
  const float* src = X + row * ldx;
  float* dst = Xout + (long)r * D;
  if ((D & 3) == 0 && ((ldx & 3) == 0)) {
    for (int c = threadIdx.x; c < D / 4; c += blockDim.x)
      reinterpret_cast<float4*>(dst)[c] = __ldg(reinterpret_cast<const float4*>(src) + c);
  } else {
    for (int c = threadIdx.x; c < D; c += blockDim.x) dst[c] = __ldg(src + c);
  }
  if (threadIdx.x == 0) Yout[r] = labels[row];
}

inline int grid_for(size_t n, int per_thread) {
  size_t blocks = (n / per_thread + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  return (int)blocks;
}

}  // namespace

DTFE_API int dtfe_push_sgd_f32(float* shard, const float* grad, size_t n, float alpha, int atomic,
                               void* stream) {
  DTFE_REQUIRE(((uintptr_t)shard % 16 == 0) && ((uintptr_t)grad % 16 == 0),
               "dtfe_push_sgd_f32: shard and grad must be 16-byte aligned");
  k_push_sgd_f32<<<grid_for(n, 4), 256, 0, (cudaStream_t)stream>>>(shard, grad, n, alpha, atomic);
  DTFE_LAUNCH_CHECK();
  return 0;
}

DTFE_API int dtfe_push_sgd_bf16(void* shard_bf16, const float* grad, size_t n, float alpha, void* stream) {
  DTFE_REQUIRE(((uintptr_t)shard_bf16 % 16 == 0) && ((uintptr_t)grad % 16 == 0) && (n % 2 == 0),
               "dtfe_push_sgd_bf16: 16-byte alignment and even n required");
  k_push_sgd_bf16<<<grid_for(n, 8), 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)shard_bf16, grad, n, alpha);
  DTFE_LAUNCH_CHECK();
  return 0;
}

DTFE_API int dtfe_pull_f32(float* dst, const float* shard, size_t n, void* stream) {
  DTFE_REQUIRE(((uintptr_t)shard % 16 == 0) && ((uintptr_t)dst % 16 == 0),
               "dtfe_pull_f32: 16-byte alignment required");
  k_pull_f32<<<grid_for(n, 4), 256, 0, (cudaStream_t)stream>>>(dst, shard, n);
  DTFE_LAUNCH_CHECK();
  return 0;
}

// lr_t is the bias-corrected step size lr*sqrt(1-b2^t)/(1-b1^t) computed by the caller.
DTFE_API int dtfe_adam_apply_f32(float* param, float* m, float* v, float* slot, size_t n, float lr_t,
                                 float beta1, float beta2, float eps, float grad_scale, void* stream) {
  k_adam_apply<<<grid_for(n, 1), 256, 0, (cudaStream_t)stream>>>(param, m, v, slot, n, lr_t, beta1, beta2,
                                                                 eps, grad_scale);
  DTFE_LAUNCH_CHECK();
  return 0;
}

// One call per rank per step, all ranks, ONE RANK PER GPU (the kernel spins on peer flags).
// The rank reduces and applies the `nseg` variables it owns; ranks owning nothing pass nseg = 0 and
// only take part in the barriers.
DTFE_API int dtfe_sync_rs_apply_f32(const float* const* grads, unsigned* const* flags, int rank, int world,
                                    int nseg, float* const* seg_param, const size_t* seg_begin,
                                    const size_t* seg_end, float alpha, unsigned epoch, unsigned* local_sync,
                                    int* global_step_or_null, const unsigned* abort_flag_or_null,
                                    double timeout_s, void* stream) {
  DTFE_REQUIRE(world >= 1 && world <= kMaxRanks, "dtfe_sync_rs_apply_f32: world must be in [1,%d]", kMaxRanks);
  DTFE_REQUIRE(nseg >= 0 && nseg <= 8, "dtfe_sync_rs_apply_f32: nseg must be in [0,8]");
  SyncArgs a{};
  for (int i = 0; i < world; ++i) { a.grads[i] = grads[i]; a.flags[i] = flags[i]; }
  size_t total = 0;
  a.nseg = nseg;
  for (int s = 0; s < nseg; ++s) {
    DTFE_REQUIRE(seg_end[s] >= seg_begin[s], "dtfe_sync_rs_apply_f32: bad range in segment %d", s);
    a.param[s] = seg_param[s]; a.begin[s] = seg_begin[s]; a.end[s] = seg_end[s];
    total += seg_end[s] - seg_begin[s];
  }
  a.alpha = alpha; a.rank = rank; a.world = world; a.epoch = epoch; a.local_sync = local_sync;
  a.global_step = global_step_or_null;
  a.abort_flag = abort_flag_or_null;
  a.timeout_ns = timeout_s > 0 ? (unsigned long long)(timeout_s * 1e9) : 0ull;
  // All CTAs must be co-resident (they wait on block 0): cap at one wave and launch cooperatively.
  int grid = grid_for(total, 8);
  if (grid > 148) grid = 148;
  void* args[] = {&a};
  DTFE_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)k_sync_rs_apply, dim3(grid), dim3(256), args, 0,
                                              (cudaStream_t)stream));
  return 0;
}

DTFE_API int dtfe_sync_max_ranks() { return kMaxRanks; }

// Xout[r,:] = X[idx[cursor*B + r], :], Yout[r] = labels[...]; *cursor is read on the device.
DTFE_API int dtfe_gather_rows_f32(const float* X, long ldx, const int* labels, const int* idx, const int* cursor,
                                  int B, int D, float* Xout, int* Yout, void* stream) {
  k_gather_rows<<<B, 256, 0, (cudaStream_t)stream>>>(X, ldx, labels, idx, cursor, B, D, Xout, Yout);
  DTFE_LAUNCH_CHECK();
  return 0;
}
