// The output layer of a wide bf16 MLP (8192 -> 10 in the benchmark family) is a "thin" layer: its three
// products have one dimension of at most 16, so a 128-wide tensor-core tile wastes >= 87 % of its MMAs and the
// operations are bound by streaming the [M,K] activation matrix once.  These kernels do exactly that stream:
//
//   thin_fwd      logits[M,C] = act[M,K] . W[K,C] + b            (reference worker.py:53, tf.layers.dense)
//   thin_dx       dX[M,K] = (dZ[M,C] . W[K,C]^T) * H*(1-H)       (gradient of the layer input, sigmoid grad fused)
//   thin_dw_push  W[K,C] += alpha * act[M,K]^T . dZ[M,C]         (gradient push fused with the apply: red.add into
//                                                                 the shard, possibly peer HBM; worker.py:75-79)
//
// W is the ps shard's [K][ldw] bf16 matrix (ldw >= C, the padded pitch the store gives thin variables) or its
// local replica; C <= 16; K a multiple of 8; 16-byte aligned rows.
#include "common.cuh"
#include <cuda_bf16.h>

namespace {

constexpr int kMaxC = 16;

__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) { float2 t = __bfloat1622float2(p[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}

// ---- forward: a warp computes R rows at a time; a lane owns 8 consecutive k of every 256-wide chunk.  The W
//      chunk (256 x C) is staged class-major in shared memory once per CTA, so a lane reads the 8 k of a class as one
//      conflict-free LDS.128; the next chunk's activations and W rows are requested before the current chunk is
//      computed (one chunk of look-ahead hides the HBM latency that 8 warps per CTA cannot).
template <int R, int CN>
__global__ void __launch_bounds__(256, 2) k_thin_fwd(const __nv_bfloat16* __restrict__ act, long lda,
                                                     const __nv_bfloat16* __restrict__ W, long ldw,
                                                     const __nv_bfloat16* __restrict__ bias, float* __restrict__ out,
                                                     long ldo, int M, int K, int C) {
  __shared__ __align__(16) __nv_bfloat16 Wt[CN][256 + 8];            // +8: rows 528 bytes apart (16-byte aligned, banks shifted)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = (blockIdx.x * 8 + warp) * R;
  float acc[R][CN];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int c = 0; c < CN; ++c) acc[r][c] = 0.f;
  uint4 an[R], wn0 = make_uint4(0, 0, 0, 0), wn1 = wn0;
  auto request = [&](int kc) {
    const int k0 = kc + lane * 8;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      an[r] = make_uint4(0, 0, 0, 0);
      if (row0 + r < M && k0 < K) an[r] = __ldcs(reinterpret_cast<const uint4*>(act + (long)(row0 + r) * lda + k0));   // streamed once
    }
    const int k = kc + threadIdx.x;                     // thread = one W row of the chunk
    wn0 = wn1 = make_uint4(0, 0, 0, 0);
    if (k < K) {
      wn0 = __ldg(reinterpret_cast<const uint4*>(W + (long)k * ldw));
      if (CN > 8) wn1 = __ldg(reinterpret_cast<const uint4*>(W + (long)k * ldw + 8));
    }
  };
  request(0);
  for (int kc = 0; kc < K; kc += 256) {
    uint4 ac[R];
#pragma unroll
    for (int r = 0; r < R; ++r) ac[r] = an[r];
    __align__(16) __nv_bfloat16 w[16];
    *reinterpret_cast<uint4*>(w) = wn0; *reinterpret_cast<uint4*>(w + 8) = wn1;
    if (kc + 256 < K) request(kc + 256);
    __syncthreads();                                    // previous chunk's readers are done
#pragma unroll
    for (int c = 0; c < CN; ++c) Wt[c][threadIdx.x] = c < C ? w[c] : __float2bfloat16(0.f);
    __syncthreads();
    float a[R][8];
#pragma unroll
    for (int r = 0; r < R; ++r) unpack8(ac[r], a[r]);
#pragma unroll
    for (int c = 0; c < CN; ++c) {
      float wv[8];
      unpack8(*reinterpret_cast<const uint4*>(&Wt[c][lane * 8]), wv);
#pragma unroll
      for (int r = 0; r < R; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[r][c] = fmaf(a[r][j], wv[j], acc[r][c]);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int c = 0; c < CN; ++c) acc[r][c] += __shfl_xor_sync(0xffffffffu, acc[r][c], o);
  // lane c writes class c of every row
#pragma unroll
  for (int r = 0; r < R; ++r) {
    float v = 0.f;
#pragma unroll
    for (int c = 0; c < CN; ++c)
      if (lane == c) v = acc[r][c];
    if (lane < C && row0 + r < M) out[(long)(row0 + r) * ldo + lane] = v + (bias ? __bfloat162float(bias[lane]) : 0.f);
  }
}

// ---- dX: a thread owns 4 consecutive k (its W rows in registers) and walks over a chunk of batch rows, 8 rows of H
//      in flight at a time; a warp reads/writes 256 contiguous bytes per row.  CN = class count rounded up to 4
//      (compile time: the dZ row is read as CN/4 LDS.128 and the products unroll without branches).
constexpr int kRowsInFlight = 8;
template <int CN>
__global__ void __launch_bounds__(256) k_thin_dx(const __nv_bfloat16* __restrict__ dZ, long lddz,
                                                 const __nv_bfloat16* __restrict__ W, long ldw,
                                                 const __nv_bfloat16* __restrict__ Hact, long ldh,
                                                 __nv_bfloat16* __restrict__ dX, long lddx, int M, int K, int C,
                                                 int rows_per_cta) {
  __shared__ __align__(16) float dz_s[64][CN];
  const int k0 = (blockIdx.x * 256 + threadIdx.x) * 4;
  const int m_begin = blockIdx.y * rows_per_cta, m_end = min(M, m_begin + rows_per_cta);
  const bool live = k0 < K;
  float w[4][CN];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int c = 0; c < CN; ++c) w[j][c] = (live && c < C) ? __bfloat162float(__ldg(W + (long)(k0 + j) * ldw + c)) : 0.f;
  for (int mb = m_begin; mb < m_end; mb += 64) {
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * CN; i += 256) {
      const int r = i / CN, c = i % CN;
      dz_s[r][c] = (mb + r < m_end && c < C) ? __bfloat162float(dZ[(long)(mb + r) * lddz + c]) : 0.f;
    }
    __syncthreads();
    if (!live) continue;
    const int nr = min(64, m_end - mb);
    for (int r0 = 0; r0 < nr; r0 += kRowsInFlight) {
      uint2 hb[kRowsInFlight];
#pragma unroll
      for (int u = 0; u < kRowsInFlight; ++u)
        hb[u] = r0 + u < nr ? __ldcs(reinterpret_cast<const uint2*>(Hact + (long)(mb + r0 + u) * ldh + k0)) : make_uint2(0u, 0u);
#pragma unroll
      for (int u = 0; u < kRowsInFlight; ++u) {
        if (r0 + u < nr) {
          const float2 h01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&hb[u].x));
          const float2 h23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&hb[u].y));
          float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
          for (int c4 = 0; c4 < CN; c4 += 4) {
            const float4 z = *reinterpret_cast<const float4*>(&dz_s[r0 + u][c4]);      // zero beyond C
            s0 = fmaf(z.x, w[0][c4], fmaf(z.y, w[0][c4 + 1], fmaf(z.z, w[0][c4 + 2], fmaf(z.w, w[0][c4 + 3], s0))));
            s1 = fmaf(z.x, w[1][c4], fmaf(z.y, w[1][c4 + 1], fmaf(z.z, w[1][c4 + 2], fmaf(z.w, w[1][c4 + 3], s1))));
            s2 = fmaf(z.x, w[2][c4], fmaf(z.y, w[2][c4 + 1], fmaf(z.z, w[2][c4 + 2], fmaf(z.w, w[2][c4 + 3], s2))));
            s3 = fmaf(z.x, w[3][c4], fmaf(z.y, w[3][c4 + 1], fmaf(z.z, w[3][c4 + 2], fmaf(z.w, w[3][c4 + 3], s3))));
          }
          uint2 o;
          o.x = dtfe::pack_bf16x2(s0 * h01.x * (1.f - h01.x), s1 * h01.y * (1.f - h01.y));
          o.y = dtfe::pack_bf16x2(s2 * h23.x * (1.f - h23.x), s3 * h23.y * (1.f - h23.y));
          *reinterpret_cast<uint2*>(dX + (long)(mb + r0 + u) * lddx + k0) = o;
        }
      }
    }
  }
}

// ---- dW: a thread owns 4 consecutive k and accumulates over a chunk of batch rows (8 rows in flight); partial sums
//      of the row chunks meet in an fp32 scratch ([chunk][k][16]), the last CTA of every k-block to finish pushes
//      alpha * sum once per element ----
template <int CN>
__global__ void __launch_bounds__(256) k_thin_dw(const __nv_bfloat16* __restrict__ act, long lda,
                                                 const __nv_bfloat16* __restrict__ dZ, long lddz, float* __restrict__ partial,
                                                 unsigned* __restrict__ counters, __nv_bfloat16* __restrict__ Wshard, long ldw,
                                                 float alpha, int M, int K, int C, int rows_per_cta) {
  __shared__ __align__(16) float dz_s[64][CN];
  __shared__ unsigned s_last;
  const int k0 = (blockIdx.x * 256 + threadIdx.x) * 4;
  const int m_begin = blockIdx.y * rows_per_cta, m_end = min(M, m_begin + rows_per_cta);
  const bool live = k0 < K;
  float acc[4][CN];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int c = 0; c < CN; ++c) acc[j][c] = 0.f;
  for (int mb = m_begin; mb < m_end; mb += 64) {
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * CN; i += 256) {
      const int r = i / CN, c = i % CN;
      dz_s[r][c] = (mb + r < m_end && c < C) ? __bfloat162float(dZ[(long)(mb + r) * lddz + c]) : 0.f;
    }
    __syncthreads();
    if (!live) continue;
    const int nr = min(64, m_end - mb);
    for (int r0 = 0; r0 < nr; r0 += kRowsInFlight) {
      uint2 ab[kRowsInFlight];
#pragma unroll
      for (int u = 0; u < kRowsInFlight; ++u)
        ab[u] = r0 + u < nr ? __ldcs(reinterpret_cast<const uint2*>(act + (long)(mb + r0 + u) * lda + k0)) : make_uint2(0u, 0u);
#pragma unroll
      for (int u = 0; u < kRowsInFlight; ++u) {
        const float2 a01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ab[u].x));   // zero rows add nothing
        const float2 a23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ab[u].y));
        const int r = min(r0 + u, 63);
#pragma unroll
        for (int c4 = 0; c4 < CN; c4 += 4) {
          const float4 z = *reinterpret_cast<const float4*>(&dz_s[r][c4]);
          acc[0][c4] = fmaf(a01.x, z.x, acc[0][c4]); acc[0][c4 + 1] = fmaf(a01.x, z.y, acc[0][c4 + 1]);
          acc[0][c4 + 2] = fmaf(a01.x, z.z, acc[0][c4 + 2]); acc[0][c4 + 3] = fmaf(a01.x, z.w, acc[0][c4 + 3]);
          acc[1][c4] = fmaf(a01.y, z.x, acc[1][c4]); acc[1][c4 + 1] = fmaf(a01.y, z.y, acc[1][c4 + 1]);
          acc[1][c4 + 2] = fmaf(a01.y, z.z, acc[1][c4 + 2]); acc[1][c4 + 3] = fmaf(a01.y, z.w, acc[1][c4 + 3]);
          acc[2][c4] = fmaf(a23.x, z.x, acc[2][c4]); acc[2][c4 + 1] = fmaf(a23.x, z.y, acc[2][c4 + 1]);
          acc[2][c4 + 2] = fmaf(a23.x, z.z, acc[2][c4 + 2]); acc[2][c4 + 3] = fmaf(a23.x, z.w, acc[2][c4 + 3]);
          acc[3][c4] = fmaf(a23.y, z.x, acc[3][c4]); acc[3][c4 + 1] = fmaf(a23.y, z.y, acc[3][c4 + 1]);
          acc[3][c4 + 2] = fmaf(a23.y, z.z, acc[3][c4 + 2]); acc[3][c4 + 3] = fmaf(a23.y, z.w, acc[3][c4 + 3]);
        }
      }
    }
  }
  // partial[chunk][k][kMaxC]
  if (live) {
    float* p = partial + ((long)blockIdx.y * K + k0) * kMaxC;
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int c = 0; c < CN; c += 4)
        __stcg(reinterpret_cast<float4*>(p + j * kMaxC + c), make_float4(acc[j][c], acc[j][c + 1], acc[j][c + 2], acc[j][c + 3]));
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(counters + blockIdx.x, 1u) == gridDim.y - 1 ? 1u : 0u;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (threadIdx.x == 0) counters[blockIdx.x] = 0;           // ready for the next launch
  if (!live) return;
  // fixed summation order over the row chunks -> reproducible; one rounded update per element into the bf16 shard
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float s[CN];
#pragma unroll
    for (int c = 0; c < CN; ++c) s[c] = 0.f;
    for (int ch = 0; ch < (int)gridDim.y; ++ch) {
      const float* p = partial + ((long)ch * K + k0 + j) * kMaxC;
#pragma unroll
      for (int c = 0; c < CN; c += 4) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(p + c));
        s[c] += v.x; s[c + 1] += v.y; s[c + 2] += v.z; s[c + 3] += v.w;
      }
    }
    __nv_bfloat16* wr = Wshard + (long)(k0 + j) * ldw;
#pragma unroll
    for (int c = 0; c < CN; c += 2)
      if (c < C) dtfe::red_add_sys_bf16x2(wr + c, dtfe::pack_bf16x2(alpha * s[c], c + 1 < C ? alpha * s[c + 1] : 0.f));
  }
}

}  // namespace

DTFE_API int dtfe_thin_fwd_bf16(const void* act, long lda, const void* W, long ldw, const void* bias_or_null,
                                float* out, long ldo, int M, int K, int C, void* stream) {
  DTFE_REQUIRE(C >= 1 && C <= kMaxC && K % 8 == 0 && lda % 8 == 0 && ldw % 8 == 0 && (C <= 8 || ldw >= 16),
               "dtfe_thin_fwd_bf16: needs C <= 16, K and pitches multiples of 8 (ldw >= 16 when C > 8)");
  DTFE_REQUIRE((((uintptr_t)act | (uintptr_t)W) % 16) == 0, "dtfe_thin_fwd_bf16: operands must be 16-byte aligned");
  constexpr int R = 2;
  const int rows_per_cta = 8 * R;
#define DTFE_THIN_FWD(CN) k_thin_fwd<R, CN><<<(M + rows_per_cta - 1) / rows_per_cta, 256, 0, (cudaStream_t)stream>>>( \
      (const __nv_bfloat16*)act, lda, (const __nv_bfloat16*)W, ldw, (const __nv_bfloat16*)bias_or_null, out, ldo, M, K, C)
  if (C <= 4) DTFE_THIN_FWD(4); else if (C <= 8) DTFE_THIN_FWD(8); else if (C <= 12) DTFE_THIN_FWD(12); else DTFE_THIN_FWD(16);
#undef DTFE_THIN_FWD
  DTFE_LAUNCH_CHECK();
  return 0;
}

DTFE_API int dtfe_thin_dx_bf16(const void* dZ, long lddz, const void* W, long ldw, const void* Hact, long ldh, void* dX,
                               long lddx, int M, int K, int C, void* stream) {
  DTFE_REQUIRE(C >= 1 && C <= kMaxC && K % 8 == 0 && ldw % 8 == 0 && ldh % 8 == 0 && lddx % 8 == 0 && (C <= 8 || ldw >= 16),
               "dtfe_thin_dx_bf16: needs C <= 16, K and pitches multiples of 8 (ldw >= 16 when C > 8)");
  DTFE_REQUIRE((((uintptr_t)W | (uintptr_t)Hact | (uintptr_t)dX) % 16) == 0, "dtfe_thin_dx_bf16: operands must be 16-byte aligned");
  const int kblocks = (K / 4 + 255) / 256;
  int chunks = (148 * 4 + kblocks - 1) / kblocks;                 // about four CTAs per SM
  int rows_per_cta = ((M + chunks - 1) / chunks + 63) / 64 * 64;
  chunks = (M + rows_per_cta - 1) / rows_per_cta;
#define DTFE_THIN_DX(CN) k_thin_dx<CN><<<dim3(kblocks, chunks), 256, 0, (cudaStream_t)stream>>>( \
      (const __nv_bfloat16*)dZ, lddz, (const __nv_bfloat16*)W, ldw, (const __nv_bfloat16*)Hact, ldh, (__nv_bfloat16*)dX, lddx, M, K, C, rows_per_cta)
  if (C <= 4) DTFE_THIN_DX(4); else if (C <= 8) DTFE_THIN_DX(8); else if (C <= 12) DTFE_THIN_DX(12); else DTFE_THIN_DX(16);
#undef DTFE_THIN_DX
  DTFE_LAUNCH_CHECK();
  return 0;
}

// Scratch: dtfe_thin_dw_scratch_words(M, K) 4-byte words, zeroed once (the kernel leaves its counters zeroed).
static int thin_dw_chunks(int K) {
  const int kblocks = (K / 4 + 255) / 256;
  return (148 * 2 + kblocks - 1) / kblocks;                       // about two CTAs per SM
}
DTFE_API size_t dtfe_thin_dw_scratch_words(int M, int K) {
  (void)M;
  return (size_t)thin_dw_chunks(K) * K * kMaxC + (K / 4 + 255) / 256 + 64;
}

DTFE_API int dtfe_thin_dw_push_bf16(const void* act, long lda, const void* dZ, long lddz, void* Wshard, long ldw, float alpha,
                                    int M, int K, int C, float* scratch, void* stream) {
  DTFE_REQUIRE(C >= 1 && C <= kMaxC && K % 8 == 0 && lda % 8 == 0 && ldw % 2 == 0 && ldw >= ((C + 1) & ~1),
               "dtfe_thin_dw_push_bf16: needs C <= 16, K and lda multiples of 8, an even W pitch that holds C rounded up to 2");
  DTFE_REQUIRE(((uintptr_t)act % 16) == 0 && ((uintptr_t)Wshard % 4) == 0 && ((uintptr_t)scratch % 16) == 0,
               "dtfe_thin_dw_push_bf16: misaligned operand");
  const int kblocks = (K / 4 + 255) / 256;
  int chunks = thin_dw_chunks(K);
  int rows_per_cta = ((M + chunks - 1) / chunks + 63) / 64 * 64;
  chunks = (M + rows_per_cta - 1) / rows_per_cta;
  float* partial = scratch;
  unsigned* counters = reinterpret_cast<unsigned*>(scratch + (size_t)thin_dw_chunks(K) * K * kMaxC);
#define DTFE_THIN_DW(CN) k_thin_dw<CN><<<dim3(kblocks, chunks), 256, 0, (cudaStream_t)stream>>>( \
      (const __nv_bfloat16*)act, lda, (const __nv_bfloat16*)dZ, lddz, partial, counters, (__nv_bfloat16*)Wshard, ldw, alpha, M, K, C, rows_per_cta)
  if (C <= 4) DTFE_THIN_DW(4); else if (C <= 8) DTFE_THIN_DW(8); else if (C <= 12) DTFE_THIN_DW(12); else DTFE_THIN_DW(16);
#undef DTFE_THIN_DW
  DTFE_LAUNCH_CHECK();
  return 0;
}
