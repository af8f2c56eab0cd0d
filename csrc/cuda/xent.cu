// Softmax cross-entropy forward+backward, mean loss and accuracy in one pass
// (reference worker.py:59-68 loss, :87-90 accuracy; TF1 SoftmaxCrossEntropyWithLogits
// + Mean + mean-grad + ArgMax/Equal/Cast/Mean — SURVEY K7-K10, K17 / N7, N9).
//
// One warp per row; classes are spread over lanes (C <= 1024 handled by striding).
// argmax(softmax(z)) == argmax(z), so the softmax needed by the reference's accuracy
// op is never materialised.  Labels are class ids (int32) or, to keep the reference's
// one-hot/soft-label feed, a dense float [M,C] matrix.
#include "common.cuh"
#include <math.h>

namespace {

__global__ void __launch_bounds__(256) k_xent(const float* __restrict__ Z, long ldz,
                                              const int* __restrict__ ids,
                                              const float* __restrict__ dense, long ldl, int M, int C,
                                              float scale, float* __restrict__ dZ, long lddz,
                                              float* __restrict__ rowloss, int* __restrict__ rowcorrect,
                                              __nv_bfloat16* __restrict__ dZb, long lddzb) {
  int lane = threadIdx.x & 31;
  int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  const float* z = Z + (long)row * ldz;
  // max + argmax(logits) (first index wins on ties, like tf.argmax)
  float mx = -INFINITY; int amx = 0x7fffffff;
  for (int c = lane; c < C; c += 32) {
    float v = z[c];
    if (v > mx) { mx = v; amx = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float om = __shfl_xor_sync(0xffffffffu, mx, o);
    int oa = __shfl_xor_sync(0xffffffffu, amx, o);
    if (om > mx || (om == mx && oa < amx)) { mx = om; amx = oa; }
  }
  float se = 0.f;
  for (int c = lane; c < C; c += 32) se += __expf(z[c] - mx);
  se = dtfe::warp_sum(se);
  float lse = mx + __logf(se);
  float inv = 1.f / se;

  float loss = 0.f;
  int label_arg = 0x7fffffff; float label_max = -INFINITY;
  if (dense) {
    const float* y = dense + (long)row * ldl;
    float ysum = 0.f;
    for (int c = lane; c < C; c += 32) {
      float yc = y[c];
      ysum += yc;
      loss += yc * (lse - z[c]);
      if (yc > label_max) { label_max = yc; label_arg = c; }
    }
    ysum = dtfe::warp_sum(ysum);
    loss = dtfe::warp_sum(loss);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float om = __shfl_xor_sync(0xffffffffu, label_max, o);
      int oa = __shfl_xor_sync(0xffffffffu, label_arg, o);
      if (om > label_max || (om == label_max && oa < label_arg)) { label_max = om; label_arg = oa; }
    }
    if (dZ || dZb)
      for (int c = lane; c < C; c += 32) {
        float d = (__expf(z[c] - mx) * inv * ysum - y[c]) * scale;
        if (dZ) dZ[(long)row * lddz + c] = d;
        if (dZb) dZb[(long)row * lddzb + c] = __float2bfloat16(d);
      }
  } else {
    int yid = ids[row];
    label_arg = yid;
    loss = lse - z[yid];  // every lane computes the same value
    if (dZ || dZb)
      for (int c = lane; c < C; c += 32) {
        float d = (__expf(z[c] - mx) * inv - (c == yid ? 1.f : 0.f)) * scale;
        if (dZ) dZ[(long)row * lddz + c] = d;
        if (dZb) dZb[(long)row * lddzb + c] = __float2bfloat16(d);
      }
  }
  if (lane == 0) {
    if (rowloss) rowloss[row] = loss;
    if (rowcorrect) rowcorrect[row] = (amx == label_arg) ? 1 : 0;
  }
}

// Deterministic single-block tree reduction: out[0] = mean(rowloss), out[1] = mean(rowcorrect)
__global__ void __launch_bounds__(256) k_mean2(const float* rowloss, const int* rowcorrect, int M,
                                               float* out_loss, float* out_acc, int* out_count) {
  __shared__ float sl[256];
  __shared__ int sc[256];
  float l = 0.f; int c = 0;
  for (int i = threadIdx.x; i < M; i += 256) {
    if (rowloss) l += rowloss[i];
    if (rowcorrect) c += rowcorrect[i];
  }
  sl[threadIdx.x] = l; sc[threadIdx.x] = c;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) { sl[threadIdx.x] += sl[threadIdx.x + s]; sc[threadIdx.x] += sc[threadIdx.x + s]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    if (out_loss) *out_loss = sl[0] / (float)M;
    if (out_acc) *out_acc = (float)sc[0] / (float)M;
    if (out_count) *out_count = sc[0];
  }
}

}  // namespace

// logits [M,C] -> dZ = (softmax - y) * scale, rowloss[M], rowcorrect[M]; labels either ids or dense.
DTFE_API int dtfe_softmax_xent_f32(const float* logits, long ldz, const int* label_ids, const float* label_dense,
                                   long ldl, int M, int C, float scale, float* dZ, long lddz, float* rowloss,
                                   int* rowcorrect, void* stream) {
  DTFE_REQUIRE((label_ids != nullptr) != (label_dense != nullptr),
               "dtfe_softmax_xent_f32: pass exactly one of label_ids / label_dense");
  int rows_per_block = 8;
  k_xent<<<(M + rows_per_block - 1) / rows_per_block, 256, 0, (cudaStream_t)stream>>>(
      logits, ldz, label_ids, label_dense, ldl, M, C, scale, dZ, lddz, rowloss, rowcorrect, nullptr, 0);
  DTFE_LAUNCH_CHECK();
  return 0;
}

// Same, additionally writing a bf16 copy of dZ (the operand format of the tcgen05 backward GEMMs).
DTFE_API int dtfe_softmax_xent_f32_bf16grad(const float* logits, long ldz, const int* label_ids, int M, int C,
                                            float scale, float* dZ, long lddz, void* dZ_bf16, long lddzb,
                                            float* rowloss, int* rowcorrect, void* stream) {
  DTFE_REQUIRE(label_ids != nullptr, "dtfe_softmax_xent_f32_bf16grad: label_ids is null");
  k_xent<<<(M + 7) / 8, 256, 0, (cudaStream_t)stream>>>(logits, ldz, label_ids, nullptr, 0, M, C, scale, dZ, lddz,
                                                        rowloss, rowcorrect, (__nv_bfloat16*)dZ_bf16, lddzb);
  DTFE_LAUNCH_CHECK();
  return 0;
}

DTFE_API int dtfe_mean_loss_acc(const float* rowloss, const int* rowcorrect, int M, float* out_loss,
                                float* out_acc, int* out_count, void* stream) {
  k_mean2<<<1, 256, 0, (cudaStream_t)stream>>>(rowloss, rowcorrect, M, out_loss, out_acc, out_count);
  DTFE_LAUNCH_CHECK();
  return 0;
}
