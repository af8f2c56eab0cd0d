// fp32 SIMT dense-layer kernels: exact-fp32 path used for parity runs against the
// numpy oracle, for shapes the tcgen05 path does not take (N=10 output layer), and as
// the evaluation forward.  One strided GEMM core with three fused epilogues:
//
//   EPI_BIAS_ACT : C = act(A.B + bias)            dense forward  (reference worker.py:50-53)
//   EPI_ACTGRAD  : C = (A.B) * h*(1-h)            dX with the sigmoid gradient fused
//   EPI_PUSH     : W += alpha * (A.B)  (red.add)  dW with the gradient push + SGD apply fused
//                                                 (reference worker.py:72-79)
#include "common.cuh"

namespace {

constexpr int BM = 64, BN = 64, BK = 16, TM = 4, TN = 4;  // 256 threads

enum { EPI_BIAS_ACT = 0, EPI_ACTGRAD = 1, EPI_PUSH = 2 };

struct GemmArgs {
  const float* A; long sam, sak;   // A(m,k) = A[m*sam + k*sak]
  const float* B; long sbk, sbn;   // B(k,n) = B[k*sbk + n*sbn]
  float* C; long ldc;              // C(m,n) = C[m*ldc + n]
  int M, N, K;
  const float* bias;               // [N] or null (EPI_BIAS_ACT)
  const float* Hprev; long ldh;    // activation output of the layer being differentiated (EPI_ACTGRAD)
  float* G; long ldg;              // optional raw-gradient output (EPI_PUSH), may be null
  float alpha;                     // -lr (EPI_PUSH)
  int act;                         // 0 identity, 1 sigmoid
  int shard_reads;                 // B (fwd) / B (dX) is shard memory: bypass L1
  int push_atomic;                 // 1: red.add (lossless Hogwild); 0: plain read-modify-write
};

template <int EPI>
__global__ void __launch_bounds__(256) k_gemm_simt(GemmArgs g) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int tx = tid % 16, ty = tid / 16;  // 16x16 threads, each 4x4
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  // Load mapping: pick the contiguous axis per operand so that a warp reads
  // consecutive addresses whichever transpose the caller asked for.
  const bool a_k_contig = (g.sak == 1);
  const bool b_n_contig = (g.sbn == 1);

  for (int k0 = 0; k0 < g.K; k0 += BK) {
#pragma unroll
    for (int it = 0; it < (BM * BK) / 256; ++it) {
      int e = tid + it * 256;
      int mm, kk;
      if (a_k_contig) { kk = e % BK; mm = e / BK; } else { mm = e % BM; kk = e / BM; }
      int m = m0 + mm, k = k0 + kk;
      float v = 0.f;
      if (m < g.M && k < g.K) v = __ldg(g.A + (long)m * g.sam + (long)k * g.sak);
      As[kk][mm] = v;
    }
#pragma unroll
    for (int it = 0; it < (BN * BK) / 256; ++it) {
      int e = tid + it * 256;
      int nn, kk;
      if (b_n_contig) { nn = e % BN; kk = e / BN; } else { kk = e % BK; nn = e / BK; }
      int n = n0 + nn, k = k0 + kk;
      float v = 0.f;
      if (n < g.N && k < g.K) {
        const float* p = g.B + (long)k * g.sbk + (long)n * g.sbn;
        v = g.shard_reads ? dtfe::ld_relaxed_sys(p) : __ldg(p);
      }
      Bs[kk][nn] = v;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) a[i] = As[kk][ty * TM + i];
#pragma unroll
      for (int j = 0; j < TN; ++j) b[j] = Bs[kk][tx * TN + j];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < TM; ++i) {
    int m = m0 + ty * TM + i;
    if (m >= g.M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      int n = n0 + tx * TN + j;
      if (n >= g.N) continue;
      float v = acc[i][j];
      if (EPI == EPI_BIAS_ACT) {
        if (g.bias) v += g.shard_reads ? dtfe::ld_relaxed_sys(g.bias + n) : __ldg(g.bias + n);
        if (g.act == 1) v = dtfe::sigmoidf_fast(v);
        g.C[(long)m * g.ldc + n] = v;
      } else if (EPI == EPI_ACTGRAD) {
        if (g.act == 1) {
          float h = g.Hprev[(long)m * g.ldh + n];
          v *= h * (1.f - h);
        }
        g.C[(long)m * g.ldc + n] = v;
      } else {
        if (g.G) g.G[(long)m * g.ldg + n] = v;
        if (g.C) {
          float* w = g.C + (long)m * g.ldc + n;
          if (g.push_atomic) dtfe::red_add_sys(w, g.alpha * v);
          else *w = dtfe::ld_relaxed_sys(w) + g.alpha * v;
        }
      }
    }
  }
}

template <int EPI>
int launch(const GemmArgs& g, cudaStream_t s) {
  dim3 grid((g.N + BN - 1) / BN, (g.M + BM - 1) / BM);
  k_gemm_simt<EPI><<<grid, 256, 0, s>>>(g);
  DTFE_LAUNCH_CHECK();
  return 0;
}

// Column sum with push: b[n] += alpha * sum_m D[m,n]   (TF1 BiasAddGrad + apply)
__global__ void __launch_bounds__(256) k_colsum_push(const float* D, long ldd, int M, int N,
                                                     float* b, float* gout, float alpha, int atomic) {
  // one warp per 32 columns, warps stride over rows; 8 warps per block cooperate via smem
  __shared__ float red[8][33];
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int n = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (n < N)
    for (int m = warp; m < M; m += 8) s += D[(long)m * ldd + n];
  red[warp][lane] = s;
  __syncthreads();
  if (warp == 0 && n < N) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[w][lane];
    if (gout) gout[n] = t;
    if (b) {
      if (atomic) dtfe::red_add_sys(b + n, alpha * t);
      else b[n] = dtfe::ld_relaxed_sys(b + n) + alpha * t;
    }
  }
}

// bf16 variant: lanes own column pairs (bf16x2 loads, 128 B per warp-row), rows split over gridDim.y;
// every (column block, row slice) pushes its partial sum - the push is additive, so no second pass.
__global__ void __launch_bounds__(256) k_colsum_push_bf16(const __nv_bfloat16* __restrict__ D, long ldd, int M, int N,
                                                          void* b, int b_bf16, float* gout, float alpha) {
  __shared__ float red[8][64 + 2];
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int n = blockIdx.x * 64 + lane * 2;
  int rows_per = (M + gridDim.y - 1) / gridDim.y;
  int r0 = blockIdx.y * rows_per, r1 = min(M, r0 + rows_per);
  float s0 = 0.f, s1 = 0.f;
  if (n + 1 < N) {
    for (int m = r0 + warp; m < r1; m += 8) {
      __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(D + (long)m * ldd + n);
      float2 f = __bfloat1622float2(v);
      s0 += f.x; s1 += f.y;
    }
  } else if (n < N) {
    for (int m = r0 + warp; m < r1; m += 8) s0 += __bfloat162float(D[(long)m * ldd + n]);
  }
  red[warp][lane * 2] = s0; red[warp][lane * 2 + 1] = s1;
  __syncthreads();
  if (warp == 0) {
    float t0 = 0.f, t1 = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { t0 += red[w][lane * 2]; t1 += red[w][lane * 2 + 1]; }
    if (gout) {
      if (n < N) atomicAdd(gout + n, t0);
      if (n + 1 < N) atomicAdd(gout + n + 1, t1);
    }
    if (b) {
      if (b_bf16) {
        __nv_bfloat16* bb = (__nv_bfloat16*)b;
        if (n + 1 < N) dtfe::red_add_sys_bf16x2(bb + n, dtfe::pack_bf16x2(alpha * t0, alpha * t1));   // one packed red per lane
        else if (n < N) atomicAdd(bb + n, __float2bfloat16(alpha * t0));
      } else {
        if (n < N) dtfe::red_add_sys((float*)b + n, alpha * t0);
        if (n + 1 < N) dtfe::red_add_sys((float*)b + n + 1, alpha * t1);
      }
    }
  }
}

}  // namespace

// C[M,N] = act(A[M,K] . W[K,N] + bias[N]).  W uses the reference's [in,out] kernel layout.
DTFE_API int dtfe_dense_fwd_f32(const float* A, long lda, const float* W, long ldw, const float* bias,
                                float* C, long ldc, int M, int N, int K, int act, int shard_reads,
                                void* stream) {
  GemmArgs g{};
  g.A = A; g.sam = lda; g.sak = 1;
  g.B = W; g.sbk = ldw; g.sbn = 1;
  g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K;
  g.bias = bias; g.act = act; g.shard_reads = shard_reads;
  return launch<EPI_BIAS_ACT>(g, (cudaStream_t)stream);
}

// dX[M,K] = (dY[M,N] . W[K,N]^T) * act'(Hprev[M,K]);  act: 0 none, 1 sigmoid (Hprev = sigmoid output)
DTFE_API int dtfe_dense_bwd_dx_f32(const float* dY, long lddy, const float* W, long ldw, const float* Hprev,
                                   long ldh, float* dX, long lddx, int M, int N, int K, int act,
                                   int shard_reads, void* stream) {
  GemmArgs g{};
  g.A = dY; g.sam = lddy; g.sak = 1;          // contraction over N
  g.B = W;  g.sbk = 1;    g.sbn = ldw;        // B(kc=n, out=k) = W[k*ldw + n]
  g.C = dX; g.ldc = lddx; g.M = M; g.N = K; g.K = N;
  g.Hprev = Hprev; g.ldh = ldh; g.act = act; g.shard_reads = shard_reads;
  return launch<EPI_ACTGRAD>(g, (cudaStream_t)stream);
}

// dW[K,N] = X[M,K]^T . dY[M,N];  W += alpha*dW (push fused; W may be peer memory), optional raw dW to G.
DTFE_API int dtfe_dense_bwd_dw_push_f32(const float* X, long ldx, const float* dY, long lddy, float* W,
                                        long ldw, float* G, long ldg, int M, int N, int K, float alpha,
                                        int atomic, void* stream) {
  GemmArgs g{};
  g.A = X;  g.sam = 1;    g.sak = ldx;        // A(out=k, kc=m) = X[m*ldx + k]
  g.B = dY; g.sbk = lddy; g.sbn = 1;
  g.C = W; g.ldc = ldw; g.G = G; g.ldg = ldg;
  g.M = K; g.N = N; g.K = M;
  g.alpha = alpha; g.push_atomic = atomic;
  return launch<EPI_PUSH>(g, (cudaStream_t)stream);
}

DTFE_API int dtfe_bias_grad_push_f32(const float* dY, long lddy, int M, int N, float* b, float* gout,
                                     float alpha, int atomic, void* stream) {
  k_colsum_push<<<(N + 31) / 32, 256, 0, (cudaStream_t)stream>>>(dY, lddy, M, N, b, gout, alpha, atomic);
  DTFE_LAUNCH_CHECK();
  return 0;
}

// bf16 activations-gradient in, bias (bf16 or fp32 shard) updated in place; gout (fp32, pre-zeroed) optional.
DTFE_API int dtfe_bias_grad_push_bf16(const void* dY_bf16, long lddy, int M, int N, void* b, int b_is_bf16,
                                      float* gout, float alpha, void* stream) {
  DTFE_REQUIRE(lddy % 2 == 0 && ((uintptr_t)dY_bf16 % 4) == 0, "dtfe_bias_grad_push_bf16: dY must be 4-byte aligned with even pitch");
  int ysplit = M >= 2048 ? 8 : 1;
  dim3 grid((N + 63) / 64, ysplit);
  k_colsum_push_bf16<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dY_bf16, lddy, M, N, b, b_is_bf16,
                                                             gout, alpha);
  DTFE_LAUNCH_CHECK();
  return 0;
}
