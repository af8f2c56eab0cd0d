// tcgen05 GEMM for sm_100a: TMA -> 128B-swizzled shared memory -> tcgen05.mma (bf16 operands,
// fp32 accumulators in TMEM) -> tcgen05.ld epilogue, with the framework's fused epilogues:
//
//   EPI_STORE        D = acc                                   (plain GEMM, tests)
//   EPI_BIAS_ACT     D = act(acc + bias[n])                    dense forward  (reference worker.py:50-53)
//   EPI_ACTGRAD      D = acc * h*(1-h)                         dX with SigmoidGrad fused (worker.py:72-73)
//   EPI_PUSH         W[m,n] += alpha*acc  via red.global.add   dW with the gradient push + SGD apply fused
//                                                             (worker.py:75-79)
//
// Any operand may live in a peer GPU's HBM: the TMA descriptor is built on whatever device pointer
// the caller passes, so "B = weights in the owning ps shard" makes the forward GEMM read the shard
// over NVLink tile by tile - the pull fused with the GEMM (worker.py:81-85 + :50).
//
// Operand majors.  UMMA computes D[M,N] = A[M,K] * B[N,K]^T with each operand either K-major
// (K contiguous in memory) or MN-major (M resp. N contiguous):
//   forward  Z = X[B,in] . W[in,out]      : A K-major, B MN-major   (W is [K][N], N contiguous)
//   dX       dX = dY[B,out] . W[in,out]^T : A K-major, B K-major    (W is [N][K], K contiguous)
//   dW       dW = X[B,in]^T . dY[B,out]   : A MN-major, B MN-major  (both are [K][MN])
//
// Shared-memory tile layouts (bf16, BLOCK_K = 64 -> 128-byte swizzle rows):
//   K-major  operand, R rows : one TMA box {64 k, R rows}; row r at r*128 B; SBO = 1024 B; the k-th
//                              UMMA_K=16 slice starts 32 B further.
//   MN-major operand, R cols : R/64 TMA boxes {64 mn, 64 k-rows}, box b at b*8192 B; k-row at
//                              +128 B; LBO = 8192 B (next 64 mn), SBO = 1024 B (next 8 k-rows); the
//                              k-th UMMA_K slice starts 2048 B further.
// (Both verified against CuTe's make_umma_desc on the host; see DESIGN.md.)
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer,
// warps 2..5 = epilogue (TMEM lane quarter = warp_id % 4).
#include "common.cuh"
#include "tc_ptx.cuh"
#include <cuda.h>
#include <cuda_bf16.h>
#include <mutex>
#include <algorithm>
#include <stdlib.h>

namespace {

constexpr int BM = 128;
// Element traits.  Everything in shared memory is laid out in 128-byte rows, and one UMMA consumes 32 bytes
// of K per row, so bf16 (kind::f16) and fp32-as-tf32 (kind::tf32) differ only in these element counts.
template <int DT> struct Elem {
  static constexpr int ES = DT == DT_BF16 ? 2 : 4;       // bytes per operand element
  static constexpr int BKE = 128 / ES;                    // K elements per pipeline stage (64 / 32)
  static constexpr int UKE = 32 / ES;                     // K elements per tcgen05.mma (16 / 8)
  static constexpr int MNB = 128 / ES;                    // MN elements per MN-major TMA box (64 / 32)
  static constexpr uint32_t FMT = DT == DT_BF16 ? 1u : 2u;  // instruction-descriptor operand format
  // MN-major operands: bf16 uses the ordinary 128B swizzle (8-row atoms); tf32 must use the 32-byte-base
  // variant (4-row atoms) on both the TMA side (SWIZZLE_128B_ATOM_32B) and the descriptor side.
  static constexpr uint32_t MN_LAYOUT = DT == DT_BF16 ? 2u : 1u;
  static constexpr uint32_t MN_SBO = DT == DT_BF16 ? 1024u : 512u;
};
constexpr int kThreads = 192;

enum { EPI_STORE = 0, EPI_BIAS_ACT = 1, EPI_ACTGRAD = 2, EPI_PUSH = 3, EPI_PUSH_TMA = 4, EPI_BIAS_ACT_TMA = 5, EPI_ACTGRAD_TMA = 6 };
// the *_TMA forms (2-CTA kernel, bf16 output): the epilogue stages the bf16 tile in shared memory and hands it to the
// TMA engine - a reduce-add for the push, a plain store for the forward / dX outputs - instead of four row-strided
// 16-byte stores per thread and chunk.
__host__ __device__ constexpr bool epi_tma(int e) { return e == EPI_PUSH_TMA || e == EPI_BIAS_ACT_TMA || e == EPI_ACTGRAD_TMA; }
__host__ __device__ constexpr int epi_base(int e) { return e == EPI_PUSH_TMA ? EPI_PUSH : e == EPI_BIAS_ACT_TMA ? EPI_BIAS_ACT : e == EPI_ACTGRAD_TMA ? EPI_ACTGRAD : e; }
enum { OUT_F32 = 0, OUT_BF16 = 1 };

struct GemmTcArgs {
  int M, N, K;
  void* D; long ldd;               // output (or pushed-into shard) matrix, row-major [M][ldd]
  int out_dtype;                   // OUT_F32 | OUT_BF16
  const void* bias; int bias_dtype;  // [N]  (EPI_BIAS_ACT)
  const void* H; long ldh; int h_f32;  // [M][ldh] activation output being differentiated (EPI_ACTGRAD), bf16 or fp32
  float alpha;                     // EPI_PUSH scale (-lr)
  int act;                         // 0 none, 1 sigmoid
  float* G; long ldg;              // optional fp32 raw gradient copy (EPI_PUSH), may be null
  int nparts; int rows_per_part;   // EPI_PUSH into a row-partitioned variable: row m lives in Dparts[m / rows_per_part]
  void* Dparts[8];
};

// ---------------------------------------------------------------- epilogue helpers
__device__ __forceinline__ float ld_param(const void* p, int dtype, int i) {
  return dtype == OUT_F32 ? __ldg((const float*)p + i) : __bfloat162float(((const __nv_bfloat16*)p)[i]);
}

// The element-wise part of an epilogue: v[0..31] = accumulators of row m, columns n0..n0+31, transformed in place.
template <int EPI>
__device__ __forceinline__ void epilogue_transform(const GemmTcArgs& g, int m, int n0, int nvalid, float* v) {
  if (EPI == EPI_BIAS_ACT) {
    // The 32 bias values of this chunk are fetched up front with vector loads: one load per element inside the
    // guarded loop below made every element a serial load -> add -> exp -> rcp chain (ncu: the bias+sigmoid
    // epilogue of a 128x256 tile took ~30 us and bounded every forward GEMM).
    float bv[32];
    bool have = false;
    if (g.bias && nvalid == 32) {
      if (g.bias_dtype == OUT_BF16 && ((reinterpret_cast<uintptr_t>(g.bias) + 2 * (size_t)n0) & 15) == 0) {
        const uint4* b4 = reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(g.bias) + n0);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint4 pk = __ldg(b4 + q);
          const __nv_bfloat162* bp = reinterpret_cast<const __nv_bfloat162*>(&pk);
#pragma unroll
          for (int e = 0; e < 4; ++e) { const float2 f = __bfloat1622float2(bp[e]); bv[q * 8 + 2 * e] = f.x; bv[q * 8 + 2 * e + 1] = f.y; }
        }
        have = true;
      } else if (g.bias_dtype == OUT_F32 && ((reinterpret_cast<uintptr_t>(g.bias) + 4 * (size_t)n0) & 15) == 0) {
        const float4* b4 = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(g.bias) + n0);
#pragma unroll
        for (int q = 0; q < 8; ++q) { const float4 f = __ldg(b4 + q); bv[4 * q] = f.x; bv[4 * q + 1] = f.y; bv[4 * q + 2] = f.z; bv[4 * q + 3] = f.w; }
        have = true;
      }
    }
    if (have || !g.bias) {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float z = v[j] + (have ? bv[j] : 0.f);
        v[j] = g.act == 1 ? dtfe::sigmoidf_fast(z) : z;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        if (j < nvalid) {
          float z = v[j] + ld_param(g.bias, g.bias_dtype, n0 + j);
          v[j] = g.act == 1 ? dtfe::sigmoidf_fast(z) : z;
        }
      }
    }
  } else if (EPI == EPI_ACTGRAD) {
    if (g.h_f32) {
      const float* hf = (const float*)g.H + (long)m * g.ldh + n0;
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < nvalid) { float hv = __ldg(hf + j); v[j] *= hv * (1.f - hv); }
    } else {
    const __nv_bfloat16* h = (const __nv_bfloat16*)g.H + (long)m * g.ldh + n0;
    if (nvalid == 32 && ((reinterpret_cast<uintptr_t>(h) & 15) == 0)) {
      const uint4* h4 = reinterpret_cast<const uint4*>(h);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 pk = __ldg(h4 + q);
        const __nv_bfloat162* hp = reinterpret_cast<const __nv_bfloat162*>(&pk);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float2 hf = __bfloat1622float2(hp[e]);
          v[q * 8 + 2 * e] *= hf.x * (1.f - hf.x);
          v[q * 8 + 2 * e + 1] *= hf.y * (1.f - hf.y);
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < nvalid) { float hf = __bfloat162float(h[j]); v[j] *= hf * (1.f - hf); }
    }
    }
  }
}

template <int EPI>
__device__ __forceinline__ void epilogue_chunk(const GemmTcArgs& g, int m, int n0, const uint32_t* acc_bits) {
  // one row m, 32 consecutive columns n0..n0+31
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc_bits[j]);
  const int nvalid = min(32, g.N - n0);
  if (nvalid <= 0) return;
  epilogue_transform<EPI>(g, m, n0, nvalid, v);
  if (EPI == EPI_PUSH) {
    if (g.G) {
      float* gp = g.G + (long)m * g.ldg + n0;
#pragma unroll
      for (int j = 0; j < 32; ++j) if (j < nvalid) gp[j] = v[j];
    }
    // Row-partitioned variable: each block of rows is owned by a different ps shard (usually a different GPU)
    void* dbase = g.D;
    int mrow = m;
    if (g.nparts) {
      int part = m / g.rows_per_part;
      dbase = g.Dparts[part];
      mrow = m - part * g.rows_per_part;
    }
    if (dbase == nullptr) return;
    if (g.out_dtype == OUT_F32) {
      float* w = (float*)dbase + (long)mrow * g.ldd + n0;
      if (nvalid == 32 && ((reinterpret_cast<uintptr_t>(w) & 15) == 0)) {
#pragma unroll
        for (int q = 0; q < 8; ++q)
          dtfe::red_add_sys_v4(w + 4 * q, make_float4(g.alpha * v[4 * q], g.alpha * v[4 * q + 1],
                                                      g.alpha * v[4 * q + 2], g.alpha * v[4 * q + 3]));
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j < nvalid) dtfe::red_add_sys(w + j, g.alpha * v[j]);
      }
    } else {
      __nv_bfloat16* w = (__nv_bfloat16*)dbase + (long)mrow * g.ldd + n0;
      if (nvalid == 32 && ((reinterpret_cast<uintptr_t>(w) & 15) == 0)) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 p;
          p.x = dtfe::pack_bf16x2(g.alpha * v[8 * q + 0], g.alpha * v[8 * q + 1]);
          p.y = dtfe::pack_bf16x2(g.alpha * v[8 * q + 2], g.alpha * v[8 * q + 3]);
          p.z = dtfe::pack_bf16x2(g.alpha * v[8 * q + 4], g.alpha * v[8 * q + 5]);
          p.w = dtfe::pack_bf16x2(g.alpha * v[8 * q + 6], g.alpha * v[8 * q + 7]);
          dtfe::red_add_sys_v4_bf16x2(w + 8 * q, p);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j < nvalid) atomicAdd(w + j, __float2bfloat16(g.alpha * v[j]));
      }
    }
    return;
  }
  // plain stores
  if (g.out_dtype == OUT_F32) {
    float* d = (float*)g.D + (long)m * g.ldd + n0;
    if (nvalid == 32 && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
#pragma unroll
      for (int q = 0; q < 8; ++q)
        reinterpret_cast<float4*>(d)[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) if (j < nvalid) d[j] = v[j];
    }
  } else {
    __nv_bfloat16* d = (__nv_bfloat16*)g.D + (long)m * g.ldd + n0;
    if (nvalid == 32 && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 p;
        p.x = dtfe::pack_bf16x2(v[8 * q + 0], v[8 * q + 1]);
        p.y = dtfe::pack_bf16x2(v[8 * q + 2], v[8 * q + 3]);
        p.z = dtfe::pack_bf16x2(v[8 * q + 4], v[8 * q + 5]);
        p.w = dtfe::pack_bf16x2(v[8 * q + 6], v[8 * q + 7]);
        reinterpret_cast<uint4*>(d)[q] = p;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) if (j < nvalid) d[j] = __float2bfloat16(v[j]);
    }
  }
}

// ---------------------------------------------------------------- the kernel
// BN: tile width (128 or 256); A_MN/B_MN: operand majors; STAGES: smem ring depth.
template <int DT, int BN, int A_MN, int B_MN, int EPI, int STAGES>
__global__ void __launch_bounds__(kThreads, BN == 128 ? 2 : 1)
k_gemm_tc(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, GemmTcArgs g) {
  using E = Elem<DT>;
  constexpr int BK = E::BKE;
  constexpr uint32_t A_BYTES = BM * 128;      // 16 KB (128-byte rows)
  constexpr uint32_t B_BYTES = BN * 128;      // 16 / 32 KB
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t BOX_BYTES = 128 * BK;     // one MN-major box: 128 bytes of mn x BK k-rows
  constexpr uint32_t TMEM_COLS = BN;           // fp32 accumulator columns (power of two >= 32)

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t full_bar[STAGES];
  __shared__ __align__(8) uint64_t empty_bar[STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar;
  __shared__ uint32_t tmem_base_holder;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int num_kb = (g.K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
#pragma unroll
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(&tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_holder, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_holder;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + (size_t)stage * STAGE_BYTES;
        uint8_t* sb = sa + A_BYTES;
        mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
        const int k0 = kb * BK;
        if (A_MN) {
#pragma unroll
          for (int b = 0; b < BM / E::MNB; ++b) tma_load_2d(sa + b * BOX_BYTES, &tmap_a, &full_bar[stage], m0 + b * E::MNB, k0);
        } else {
          tma_load_2d(sa, &tmap_a, &full_bar[stage], k0, m0);
        }
        if (B_MN) {
#pragma unroll
          for (int b = 0; b < BN / E::MNB; ++b) tma_load_2d(sb + b * BOX_BYTES, &tmap_b, &full_bar[stage], n0 + b * E::MNB, k0);
        } else {
          tma_load_2d(sb, &tmap_b, &full_bar[stage], k0, n0);
        }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(BM, BN, A_MN, B_MN, E::FMT);
      int stage = 0; uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + (size_t)stage * STAGE_BYTES);
        const uint32_t sb = sa + A_BYTES;
#pragma unroll
        for (int k = 0; k < 4; ++k) {   // 128-byte rows / 32 bytes of K per MMA
          const uint64_t adesc = A_MN ? make_smem_desc(sa + k * (E::UKE * 128), BOX_BYTES, E::MN_SBO, E::MN_LAYOUT)
                                      : make_smem_desc(sa + k * 32, 16, 1024);
          const uint64_t bdesc = B_MN ? make_smem_desc(sb + k * (E::UKE * 128), BOX_BYTES, E::MN_SBO, E::MN_LAYOUT)
                                      : make_smem_desc(sb + k * 32, 16, 1024);
          umma_1sm<DT>(tmem_base, adesc, bdesc, idesc, (kb | k) != 0);
        }
        umma_commit(&empty_bar[stage]);   // frees this smem stage once the MMAs above have read it
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      umma_commit(&tmem_full_bar);        // accumulator complete
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int q = warp & 3;               // TMEM lane quarter this warp may access
    mbar_wait(&tmem_full_bar, 0);
    tc_fence_after();
    const int m = m0 + q * 32 + lane;
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
      uint32_t r[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, r);
      tmem_ld_wait();
      if (m < g.M && n0 + c < g.N) epilogue_chunk<EPI>(g, m, n0 + c, r);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ---------------------------------------------------------------- 2-CTA persistent kernel
// cta_group::2: a cluster of two CTAs (one TPC) computes a 256 x BN2 tile.  Each CTA stages its own 128
// rows of A and its own half of B (BN2/2 rows), the leader CTA issues tcgen05.mma.cta_group::2 which reads
// both CTAs' shared memory and writes 128 accumulator rows into each CTA's TMEM.  Shared-memory traffic
// per FLOP is half of the 1-CTA 128x128 kernel's.  The kernel is persistent (clusters stride over the
// tiles) and the accumulator is double-buffered in TMEM (2 x 256 columns), so the epilogue of tile i
// overlaps the MMAs of tile i+1.
constexpr int BN2 = 256;
constexpr int kStages2 = 6;
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;   // clears the CTA-rank bit of a shared::cluster address

// TMA reduce-store: one bulk tensor op adds a whole shared-memory box into global (possibly peer) memory.
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* map, const void* smem_src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
               :: "l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
               :: "l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

struct DMaps { CUtensorMap m[8]; };   // one tensor map per row block of a partitioned push target

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  // completes on the LEADER CTA's barrier (rank bit cleared), whichever CTA issues it
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      :: "r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1) : "memory");
}
template <int DT>
__device__ __forceinline__ void umma_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
  if (DT == DT_BF16) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
  } else {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}" :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
  }
}
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
  // arrives on the barrier at this offset in BOTH CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               :: "r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
  // arrive on the leader CTA's copy of `bar` from either CTA
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];"
               :: "r"(smem_u32(bar) & kPeerBitMask) : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_holder, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;"
               :: "r"(smem_u32(smem_holder)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" :: "r"(taddr), "r"(ncols) : "memory");
}

template <int DT, int A_MN, int B_MN, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
k_gemm_tc2(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
           const __grid_constant__ DMaps dmaps, GemmTcArgs g) {
  // EPI_PUSH_TMA: the epilogue stages the scaled bf16 tile in shared memory (4 swizzled 128x64 boxes) and
  // pushes it with TMA reduce-adds - 16 KB per operation instead of 16 B per red instruction, which is what
  // makes a push into a PEER shard run at NVLink speed.  It trades two pipeline stages for the staging tile.
  constexpr int kStages = epi_tma(EPI) ? 4 : kStages2;
  using E = Elem<DT>;
  constexpr int BK = E::BKE;
  constexpr uint32_t A_BYTES = BM * 128;               // this CTA's 128 rows of A
  constexpr uint32_t B_BYTES = (BN2 / 2) * 128;        // this CTA's half of B
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;  // 32 KB
  constexpr uint32_t STAGING_BOX = BM * 128;           // 128 rows x 64 bf16
  constexpr uint32_t BOX_BYTES = 128 * BK;
  constexpr uint32_t TMEM_COLS = 2 * BN2;              // two accumulator stages

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t full_bar[kStages];
  __shared__ __align__(8) uint64_t empty_bar[kStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_holder;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int num_kb = (g.K + BK - 1) / BK;
  const int tiles_m = (g.M + 2 * BM - 1) / (2 * BM), tiles_n = (g.N + BN2 - 1) / BN2;
  const int num_tiles = tiles_m * tiles_n;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
#pragma unroll
    for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
#pragma unroll
    for (int s = 0; s < 2; ++s) { mbar_init(&tmem_full_bar[s], 1); mbar_init(&tmem_empty_bar[s], 8); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_2sm(&tmem_base_holder, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();            // barriers of both CTAs initialised before any cross-CTA signal
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_holder;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const int m0 = (tile % tiles_m) * (2 * BM) + (int)rank * BM;      // this CTA's rows of A / D
        const int n0 = (tile / tiles_m) * BN2 + (int)rank * (BN2 / 2);    // this CTA's half of B
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + (size_t)stage * STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);   // both CTAs' bytes land on the leader
          const int k0 = kb * BK;
          if (A_MN) {
#pragma unroll
            for (int b = 0; b < BM / E::MNB; ++b) tma_load_2d_2sm(sa + b * BOX_BYTES, &tmap_a, &full_bar[stage], m0 + b * E::MNB, k0);
          } else {
            tma_load_2d_2sm(sa, &tmap_a, &full_bar[stage], k0, m0);
          }
          if (B_MN) {
#pragma unroll
            for (int b = 0; b < (BN2 / 2) / E::MNB; ++b) tma_load_2d_2sm(sb + b * BOX_BYTES, &tmap_b, &full_bar[stage], n0 + b * E::MNB, k0);
          } else {
            tma_load_2d_2sm(sb, &tmap_b, &full_bar[stage], k0, n0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && lane == 0) {
      constexpr uint32_t idesc = make_idesc(2 * BM, BN2, A_MN, B_MN, E::FMT);
      int stage = 0; uint32_t phase = 0;
      int as = 0; uint32_t aphase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        mbar_wait(&tmem_empty_bar[as], aphase ^ 1);      // both CTAs' epilogues drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN2);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)stage * STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t adesc = A_MN ? make_smem_desc(sa + k * (E::UKE * 128), BOX_BYTES, E::MN_SBO, E::MN_LAYOUT)
                                        : make_smem_desc(sa + k * 32, 16, 1024);
            const uint64_t bdesc = B_MN ? make_smem_desc(sb + k * (E::UKE * 128), BOX_BYTES, E::MN_SBO, E::MN_LAYOUT)
                                        : make_smem_desc(sb + k * 32, 16, 1024);
            umma_2sm<DT>(tmem_d, adesc, bdesc, idesc, (kb | k) != 0);
          }
          umma_commit_2sm(&empty_bar[stage]);            // frees the stage in both CTAs
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit_2sm(&tmem_full_bar[as]);             // accumulator ready, both CTAs' epilogues
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5, both CTAs) =====================
    const int q = warp & 3;
    int as = 0; uint32_t aphase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      const int m0 = (tile % tiles_m) * (2 * BM) + (int)rank * BM;
      const int n0 = (tile / tiles_m) * BN2;
      mbar_wait(&tmem_full_bar[as], aphase);
      tc_fence_after();
      const int m = m0 + q * 32 + lane;
      if (epi_tma(EPI)) {
        uint8_t* staging = smem + (size_t)kStages * STAGE_BYTES;
        const int row = q * 32 + lane;
        // the previous tile's bulk stores must have finished reading the staging tile
        if (warp == 2 && lane == 0) tma_store_wait_read();
        named_bar_sync(1, 128);
#pragma unroll 1
        for (int c = 0; c < BN2; c += 32) {
          uint32_t r[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN2 + c), r);
          tmem_ld_wait();
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
          if (EPI == EPI_PUSH_TMA) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] *= g.alpha;
          } else {
            // rows / columns beyond the matrix are clipped by the tensor map; their values only have to be harmless
            const int nvalid = min(32, g.N - (n0 + c));
            if (m < g.M && nvalid > 0) epilogue_transform<epi_base(EPI)>(g, m, n0 + c, nvalid, v);
          }
          // 32 columns = 4 chunks of 16 B inside the 128-byte row of box c/64, XOR-swizzled like the tensor map
          uint8_t* rowp = staging + (size_t)(c >> 6) * STAGING_BOX + (size_t)row * 128;
          const int chunk0 = (c & 32) ? 4 : 0;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            uint4 pk;
            pk.x = dtfe::pack_bf16x2(v[8 * u + 0], v[8 * u + 1]);
            pk.y = dtfe::pack_bf16x2(v[8 * u + 2], v[8 * u + 3]);
            pk.z = dtfe::pack_bf16x2(v[8 * u + 4], v[8 * u + 5]);
            pk.w = dtfe::pack_bf16x2(v[8 * u + 6], v[8 * u + 7]);
            *reinterpret_cast<uint4*>(rowp + (((chunk0 + u) ^ (row & 7)) << 4)) = pk;
          }
        }
        fence_proxy_async();               // make the generic-proxy writes visible to the TMA engine
        named_bar_sync(1, 128);
        if (warp == 2 && lane == 0) {
          int part = 0, mloc = m0;
          if (EPI == EPI_PUSH_TMA && g.nparts) { part = m0 / g.rows_per_part; mloc = m0 - part * g.rows_per_part; }
#pragma unroll
          for (int jb = 0; jb < BN2 / 64; ++jb)
            if (n0 + jb * 64 < g.N) {
              if (EPI == EPI_PUSH_TMA) tma_reduce_add_2d(&dmaps.m[part], staging + (size_t)jb * STAGING_BOX, n0 + jb * 64, mloc);
              else tma_store_2d(&dmaps.m[0], staging + (size_t)jb * STAGING_BOX, n0 + jb * 64, mloc);
            }
          tma_store_commit();
        }
      } else {
#pragma unroll 1
        for (int c = 0; c < BN2; c += 32) {
          uint32_t r[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN2 + c), r);
          tmem_ld_wait();
          if (m < g.M && n0 + c < g.N) epilogue_chunk<epi_base(EPI)>(g, m, n0 + c, r);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_leader(&tmem_empty_bar[as]);
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
  }

  if (epi_tma(EPI) && warp == 2 && lane == 0) tma_store_wait_all();   // bulk stores / pushes complete before exit
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();            // nobody exits (or frees TMEM) while its peer may still signal it
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, TMEM_COLS);
  }
}

// ---------------------------------------------------------------- host side
template <int DT, int BN, int A_MN, int B_MN, int EPI>
int launch_cfg(const CUtensorMap& ta, const CUtensorMap& tb, const GemmTcArgs& g, cudaStream_t stream) {
  constexpr int STAGES = BN == 128 ? 3 : 4;   // 128: 97 KB -> two CTAs per SM; 256: 193 KB -> one
  constexpr size_t smem = (size_t)STAGES * (BM * 128 + BN * 128) + 1024;
  auto kern = k_gemm_tc<DT, BN, A_MN, B_MN, EPI, STAGES>;
  static bool configured[64] = {};
  int dev;
  DTFE_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    DTFE_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev] = true;
  }
  dim3 grid((g.N + BN - 1) / BN, (g.M + BM - 1) / BM);
  kern<<<grid, kThreads, smem, stream>>>(ta, tb, g);
  DTFE_LAUNCH_CHECK();
  return 0;
}

template <int DT, int A_MN, int B_MN, int EPI>
int launch_cfg2(const CUtensorMap& ta, const CUtensorMap& tb, const DMaps& dm, const GemmTcArgs& g, cudaStream_t stream) {
  constexpr size_t smem = epi_tma(EPI)
      ? (size_t)4 * (BM * 128 + (BN2 / 2) * 128) + (size_t)(BN2 / 64) * BM * 128 + 1024
      : (size_t)kStages2 * (BM * 128 + (BN2 / 2) * 128) + 1024;
  auto kern = k_gemm_tc2<DT, A_MN, B_MN, EPI>;
  static bool configured[64] = {};
  static int sm_count[64] = {};
  int dev;
  DTFE_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    DTFE_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DTFE_CUDA_CHECK(cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev));
    configured[dev] = true;
  }
  int tiles = ((g.M + 2 * BM - 1) / (2 * BM)) * ((g.N + BN2 - 1) / BN2);
  int clusters = (dev < 64 ? sm_count[dev] : 148) / 2;
  if (clusters > tiles) clusters = tiles;
  kern<<<dim3(2 * clusters), kThreads, smem, stream>>>(ta, tb, dm, g);
  DTFE_LAUNCH_CHECK();
  return 0;
}

template <int DT, int EPI>
int launch_major2(int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const DMaps& dm,
                  const GemmTcArgs& g, cudaStream_t s) {
  if (!a_mn && b_mn) return launch_cfg2<DT, 0, 1, EPI>(ta, tb, dm, g, s);
  if (!a_mn && !b_mn) return launch_cfg2<DT, 0, 0, EPI>(ta, tb, dm, g, s);
  if (a_mn && b_mn) return launch_cfg2<DT, 1, 1, EPI>(ta, tb, dm, g, s);
  return launch_cfg2<DT, 1, 0, EPI>(ta, tb, dm, g, s);
}

template <int DT>
int launch_epi2(int epi, int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const DMaps& dm,
                const GemmTcArgs& g, cudaStream_t s) {
  switch (epi) {
    case EPI_STORE: return launch_major2<DT, EPI_STORE>(a_mn, b_mn, ta, tb, dm, g, s);
    case EPI_BIAS_ACT: return launch_major2<DT, EPI_BIAS_ACT>(a_mn, b_mn, ta, tb, dm, g, s);
    case EPI_ACTGRAD: return launch_major2<DT, EPI_ACTGRAD>(a_mn, b_mn, ta, tb, dm, g, s);
    case EPI_PUSH: return launch_major2<DT, EPI_PUSH>(a_mn, b_mn, ta, tb, dm, g, s);
    case EPI_PUSH_TMA:
      if (DT == DT_BF16) return launch_major2<DT_BF16, EPI_PUSH_TMA>(a_mn, b_mn, ta, tb, dm, g, s);
      break;
    case EPI_BIAS_ACT_TMA:
      if (DT == DT_BF16) return launch_major2<DT_BF16, EPI_BIAS_ACT_TMA>(a_mn, b_mn, ta, tb, dm, g, s);
      break;
    case EPI_ACTGRAD_TMA:
      if (DT == DT_BF16) return launch_major2<DT_BF16, EPI_ACTGRAD_TMA>(a_mn, b_mn, ta, tb, dm, g, s);
      break;
  }
  dtfe_set_error("gemm_tc: unknown epilogue %d", epi);
  return -1;
}

template <int DT, int BN, int EPI>
int launch_major(int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const GemmTcArgs& g,
                 cudaStream_t s) {
  if (!a_mn && b_mn) return launch_cfg<DT, BN, 0, 1, EPI>(ta, tb, g, s);
  if (!a_mn && !b_mn) return launch_cfg<DT, BN, 0, 0, EPI>(ta, tb, g, s);
  if (a_mn && b_mn) return launch_cfg<DT, BN, 1, 1, EPI>(ta, tb, g, s);
  return launch_cfg<DT, BN, 1, 0, EPI>(ta, tb, g, s);
}

template <int DT, int BN>
int launch_epi(int epi, int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const GemmTcArgs& g,
               cudaStream_t s) {
  switch (epi) {
    case EPI_STORE: return launch_major<DT, BN, EPI_STORE>(a_mn, b_mn, ta, tb, g, s);
    case EPI_BIAS_ACT: return launch_major<DT, BN, EPI_BIAS_ACT>(a_mn, b_mn, ta, tb, g, s);
    case EPI_ACTGRAD: return launch_major<DT, BN, EPI_ACTGRAD>(a_mn, b_mn, ta, tb, g, s);
    case EPI_PUSH: return launch_major<DT, BN, EPI_PUSH>(a_mn, b_mn, ta, tb, g, s);
  }
  dtfe_set_error("gemm_tc: unknown epilogue %d", epi);
  return -1;
}

// Shared implementation of the two public entry points.
template <int DT>
int gemm_tc_impl(const dtfe_gemm_tc* p, void* stream) {
  using E = Elem<DT>;
  DTFE_REQUIRE(p->M > 0 && p->N > 0 && p->K > 0, "gemm_tc: empty problem");
  DTFE_REQUIRE(p->tile_n == 128 || p->tile_n == 256 || p->tile_n == 0 || p->tile_n == 512,
               "gemm_tc: tile_n must be 0 (auto), 128, 256 (1-CTA kernels) or 512 (= 2-CTA 256x256 persistent kernel)");
  DTFE_REQUIRE(!(DT == DT_TF32 && p->tile_n == 256), "gemm_tc: the fp32/tf32 path has 128-wide and 2-CTA tiles only");
  // auto: the 2-CTA kernel whenever the problem has at least one full 256x256 tile per SM pair
  const bool two_cta = p->tile_n == 512 ||
                       (p->tile_n == 0 && p->M >= 512 && p->N >= 512 && p->K >= 1024 &&
                        (long)p->M * p->N >= 256L * 256L * 74);
  int bn = two_cta ? 128 : (p->tile_n ? p->tile_n : 128);   // B box rows: 128 = half of BN2
  CUtensorMap ta, tb;
  int rc;
  if (p->a_mn_major) rc = make_tmap(&ta, p->A, (uint64_t)p->M, (uint64_t)p->K, (uint64_t)p->lda, E::MNB, E::BKE, E::ES, DT == DT_TF32);
  else rc = make_tmap(&ta, p->A, (uint64_t)p->K, (uint64_t)p->M, (uint64_t)p->lda, E::BKE, BM, E::ES);
  if (rc) return rc;
  if (p->b_mn_major) rc = make_tmap(&tb, p->B, (uint64_t)p->N, (uint64_t)p->K, (uint64_t)p->ldb, E::MNB, E::BKE, E::ES, DT == DT_TF32);
  else rc = make_tmap(&tb, p->B, (uint64_t)p->K, (uint64_t)p->N, (uint64_t)p->ldb, E::BKE, (uint32_t)bn, E::ES);
  if (rc) return rc;
  GemmTcArgs g{};
  g.M = p->M; g.N = p->N; g.K = p->K;
  g.D = p->D; g.ldd = p->ldd; g.out_dtype = p->out_bf16 ? OUT_BF16 : OUT_F32;
  g.bias = p->bias; g.bias_dtype = p->bias_bf16 ? OUT_BF16 : OUT_F32;
  g.H = p->H; g.ldh = p->ldh; g.h_f32 = DT == DT_TF32;
  g.alpha = p->alpha; g.act = p->act; g.G = p->G; g.ldg = p->ldg;
  DTFE_REQUIRE(p->nparts >= 0 && p->nparts <= 8, "gemm_tc: nparts must be in [0,8]");
  if (p->nparts) {
    DTFE_REQUIRE(p->epilogue == EPI_PUSH && p->rows_per_part > 0 && (long)p->rows_per_part * p->nparts >= p->M,
                 "gemm_tc: D_parts is for the push epilogue and must cover M rows");
    g.nparts = p->nparts; g.rows_per_part = p->rows_per_part;
    for (int i = 0; i < p->nparts; ++i) g.Dparts[i] = p->D_parts[i];
  }
  if (p->epilogue == EPI_ACTGRAD) DTFE_REQUIRE(p->H != nullptr, "gemm_tc: EPI_ACTGRAD needs H");
  if (p->epilogue != EPI_PUSH) DTFE_REQUIRE(p->D != nullptr, "gemm_tc: D is null");
  if (two_cta) {
    // Push through TMA reduce-adds when the target is a bf16 matrix whose row blocks line up with the tiles
    static DMaps dm;   // zero-initialised; only read by the EPI_PUSH_TMA instantiation
    DMaps local{};
    int epi = p->epilogue;
    // (locally the per-thread red.add path keeps two more pipeline stages and is ~10 % faster; the bulk path
    // is for targets in peer HBM, where 16 KB operations instead of 16 B ones matter - callers ask for it)
    const bool want_tma = (p->tma_push || getenv("DTFE_TMA_PUSH") != nullptr) && getenv("DTFE_NO_TMA_PUSH") == nullptr;
    const bool tma_ok = DT == DT_BF16 && want_tma && epi == EPI_PUSH && p->out_bf16 && p->G == nullptr &&
                        (p->ldd * 2) % 16 == 0 && (p->nparts == 0 ? p->D != nullptr : p->rows_per_part % BM == 0);
    if (tma_ok) {
      const int np = p->nparts ? p->nparts : 1;
      bool ok = true;
      for (int i = 0; i < np && ok; ++i) {
        void* base = p->nparts ? p->D_parts[i] : p->D;
        long rows = p->nparts ? std::min<long>(p->rows_per_part, (long)p->M - (long)i * p->rows_per_part) : p->M;
        ok = base != nullptr && ((uintptr_t)base % 16) == 0 && rows > 0 &&
             make_tmap(&local.m[i], base, (uint64_t)p->N, (uint64_t)rows, (uint64_t)p->ldd, 64, BM) == 0;
      }
      if (ok) epi = EPI_PUSH_TMA;
    }
    // forward / dX outputs in bf16 can go through the TMA engine as well (DTFE_TMA_STORE=1).  Measured on 4096x8192x8192:
    // 358 / 354 us against 326 / 315 us with per-thread stores - the staging tile costs two of the six pipeline stages
    // and a CTA-wide handshake per tile, which outweighs the better store pattern when the target is local HBM.
    if (DT == DT_BF16 && (epi == EPI_BIAS_ACT || epi == EPI_ACTGRAD) && p->out_bf16 && (p->ldd * 2) % 16 == 0 &&
        ((uintptr_t)p->D % 16) == 0 && getenv("DTFE_TMA_STORE") != nullptr &&
        make_tmap(&local.m[0], p->D, (uint64_t)p->N, (uint64_t)p->M, (uint64_t)p->ldd, 64, BM) == 0)
      epi = epi == EPI_BIAS_ACT ? EPI_BIAS_ACT_TMA : EPI_ACTGRAD_TMA;
    return launch_epi2<DT>(epi, p->a_mn_major, p->b_mn_major, ta, tb, epi_tma(epi) ? local : dm, g, (cudaStream_t)stream);
  }
  if (DT == DT_BF16 && bn == 256)
    return launch_epi<DT_BF16, 256>(p->epilogue, p->a_mn_major, p->b_mn_major, ta, tb, g, (cudaStream_t)stream);
  return launch_epi<DT, 128>(p->epilogue, p->a_mn_major, p->b_mn_major, ta, tb, g, (cudaStream_t)stream);
}

}  // namespace

// D[M,N] (op)= A . B with bf16 operands and fp32 accumulation on the 5th-gen tensor cores.
//   a_mn_major = 0: A is [M][lda] (K contiguous);  1: A is [K][lda] (M contiguous)
//   b_mn_major = 0: B is [N][ldb] (K contiguous);  1: B is [K][ldb] (N contiguous)
DTFE_API int dtfe_gemm_tc_bf16(const dtfe_gemm_tc* p, void* stream) { return gemm_tc_impl<DT_BF16>(p, stream); }

// Same kernels with fp32 operands fed to the tensor cores as tf32 (kind::tf32; fp32 accumulation): lets
// fp32 parameter shards - the reference's dtype - be pulled by TMA straight into the MMA without a cast.
// H (EPI_ACTGRAD) and bias are fp32 here; row pitches must be multiples of 4 elements.
DTFE_API int dtfe_gemm_tc_tf32(const dtfe_gemm_tc* p, void* stream) { return gemm_tc_impl<DT_TF32>(p, stream); }
