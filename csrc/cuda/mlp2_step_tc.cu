// The fused MNIST training step on the 5th-generation tensor cores: the same step as mlp2_step.cu
// (pull -> forward -> loss/accuracy -> backward -> push+apply -> global_step, reference worker.py:129-141, any
// number of steps per launch), but the two contractions that carry the FLOPs - X.W1 and X^T.dA - run as
// tcgen05.mma kind::tf32 with fp32 accumulators in TMEM, on ONE thread-block cluster of 8 CTAs instead of a
// cooperative grid of 148.
//
//   CTA i of the cluster owns the features [i*KSL, (i+1)*KSL) (KSL = 100 for D = 784): its slice of the batch
//   (gathered rows, cp.async with the UMMA swizzles applied by hand, K-major for the forward and MN-major for
//   the backward product) and its slice of W1, which TMA pulls straight out of the ps shard (possibly a peer
//   GPU's HBM) in the 32-byte-base swizzle tcgen05 wants for MN-major tf32 operands.
//
//   forward   D1[b,h] = Xslice[b,:] . W1slice[:,h]      13 MMAs 128x128x8, accumulator in TMEM columns 0..127
//             -> the split-K partial goes to scratch (8 partials instead of the SIMT kernel's 98)
//   cluster barrier (0.2 us instead of a 1.2 us grid barrier)
//   phase B   a warp per batch row (rows i, i+8, ...): sum the 8 partials, bias, sigmoid, logits, softmax-xent
//             forward+backward, accuracy, dA
//   cluster barrier
//   backward  D2[f,h] = Xslice[:,f]^T . dA[:,h]         13 MMAs, accumulator in TMEM columns 128..255;
//             epilogue: tcgen05.ld -> red.add.v4 of -lr*dW1 into the shard (push = apply), and the same delta
//             added to the NEXT step's W1 slice, which TMA fetched before this CTA pushed (pull one step ahead;
//             own updates exact, other workers' one step late - mlp2_step.cu's contract).
//   The small gradients (dW2, db1, db2) are spread over the warps of all CTAs while the backward MMAs run.
//
// tf32 feeds the tensor cores 10 mantissa bits of X, W1 and dA (accumulation is fp32); everything else - the
// reductions, sigmoid, softmax, the updates - is fp32.  Limits: B <= 104, D <= 832, H <= 128 and H % 4 == 0,
// C <= 16; asynchronous lossless training only (other cases stay on mlp2_step.cu).
#include "common.cuh"
#include "tc_ptx.cuh"
#include <math.h>

namespace {

constexpr int kT = 512;                 // threads per CTA: 16 warps = 4 TMEM lane quarters x 4 column groups
constexpr int kLoadT = 13 * 32;         // warps 0..12 stage operands (and own a batch row each); warps 13..15 fetch
                                        // the small parameters, so that nobody who waits for local data ever waits
                                        // for a peer GPU (cp.async groups complete in order, per thread)
constexpr int kCl = 8;                  // CTAs in the cluster
constexpr int kRowsPad = 104;           // padded batch rows / slice width (multiple of 8)
constexpr int kChunks = kRowsPad / 4;   // 16-byte chunks per padded slice row (26)
constexpr uint32_t kBox = kRowsPad * 128;          // one 128-byte-wide box of 104 rows
constexpr uint32_t kBuf = 4 * kBox;                // four boxes: 128 K- or MN-elements wide
constexpr int kMaxH = 128, kMaxC = 16;

struct TcArgs {
  int B, D, H, C;
  float *W1, *b1, *W2, *b2;
  int* global_step;
  const float* X; long ldx;
  const int* labels;
  const int* batch_idx;
  long rows_per_step;
  float* part;              // [kCl][B][H] split-K partials
  float* Hbuf;              // [B][H]
  float* dA;                // [B][H]
  float* dZ;                // [B][C]
  float* rowloss;           // [B]
  int* rowcorrect;          // [B]
  float* delta;             // this worker's last small-parameter updates, [dWords]
  float* loss_out; float* acc_out; int* step_out;
  float lr;
  int nsteps;
  int remote;
  int ksl;                  // features per CTA (multiple of 4, <= 104)
  int* cursor;
  long long* dbg;           // optional [nsteps][kCl][16] globaltimer stamps
};

__device__ __forceinline__ long long gtime_tc() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define TC_STAMP(i) do { if (a.dbg && tid == 0) a.dbg[((long)s * kCl + rank) * 16 + (i)] = gtime_tc(); } while (0)

__device__ __forceinline__ void cluster_barrier() {
  __syncwarp();
  asm volatile("barrier.cluster.arrive.release;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire;" ::: "memory");
}
__device__ __forceinline__ uint32_t ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}

// Byte offset of the 16-byte chunk `c` (4 tf32 values) of row `r` inside a four-box operand buffer.
//   K-major (A of the forward product): box = 32 K values, rows are batch rows, 8-row 128B-swizzle atoms.
__device__ __forceinline__ uint32_t off_kmajor(int r, int c) {
  return (uint32_t)(c >> 3) * kBox + (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u +
         ((uint32_t)((c & 7) ^ (r & 7)) << 4);
}
//   MN-major (both operands of the backward product, B of the forward one): box = 32 MN values, rows are K
//   rows, 4-row atoms with the 32-byte-base swizzle (byte-address bits 5-6 ^= bits 7-8).
__device__ __forceinline__ uint32_t off_mnmajor(int krow, int c) {
  return (uint32_t)(c >> 3) * kBox + (uint32_t)(krow >> 2) * 512u + (uint32_t)(krow & 3) * 128u +
         ((uint32_t)(((c & 7) >> 1) ^ (krow & 3)) << 5) + ((uint32_t)(c & 1) << 4);
}

// CT: number of classes known at compile time (0: taken from the arguments).  With a run-time class count every
// per-class shuffle of the unrolled loops below sits in its own basic block and their latencies add up instead of
// overlapping (measured: 2.8 us for the logits of one row instead of 0.6).
template <int CT>
__global__ void __cluster_dims__(kCl, 1, 1) __launch_bounds__(kT, 1)
k_mlp2_step_tc(const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmX,
               const __grid_constant__ CUtensorMap tmDA, TcArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* Xk = smem;                 // forward A: [4 K-boxes][104 batch rows][128 B]
  uint8_t* Xmn = Xk + kBuf;           // backward A: [4 feature boxes][104 batch rows (K)][128 B]
  uint8_t* W1s = Xmn + kBuf;          // forward B: [4 h-boxes][104 feature rows (K)][128 B]   (TMA)
  uint8_t* dAs = W1s + kBuf;          // backward B: [4 h-boxes][104 batch rows (K)][128 B]
  const int B = a.B, D = a.D, H = a.H, C = CT ? CT : a.C;
  const int dOffB1 = (H * C + 3) & ~3, dOffB2 = dOffB1 + ((H + 3) & ~3), dWords = dOffB2 + ((C + 3) & ~3);
  float* cur = reinterpret_cast<float*>(dAs + kBuf);   // small parameters of this step: W2 | b1 | b2
  float* stg = cur + dWords;                           // the same, fetched for the next step
  float* dZs = stg + dWords;                           // [B][C]
  int* rows_cur = reinterpret_cast<int*>(dZs + kRowsPad * kMaxC);
  int* rows_nxt = rows_cur + kRowsPad;
  __shared__ __align__(8) uint64_t bar_w1, bar_x, bar_da, bar_f, bar_b;
  __shared__ uint32_t tmem_holder;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = (int)ctarank();
  const int k0 = rank * a.ksl;
  const int nk = max(0, min(a.ksl, D - k0));            // real features of this slice
  const float invB = 1.f / (float)B, nlr = -a.lr;
  const long sbase = a.cursor ? (long)*reinterpret_cast<volatile int*>(a.cursor) : 0;
  const int book = kCl - 1;
  // `cur` keeps W2 class-major ([C][H]: a lane reads the four hidden units it owns as one conflict-free float4),
  // then b1, then b2; `stg`, `delta` and the shard use the variable's own [H][C] order.
  float* W2t = cur; float* b1s = cur + dOffB1; float* b2s = cur + dOffB2;
  auto cur_slot = [&](int e) -> int { return e < H * C ? (e % C) * H + e / C : e; };

  if (tid == 0) {
    tma_prefetch_desc(&tmW1); tma_prefetch_desc(&tmX); tma_prefetch_desc(&tmDA);
    mbar_init(&bar_w1, 1); mbar_init(&bar_x, 1); mbar_init(&bar_da, 1); mbar_init(&bar_f, 1); mbar_init(&bar_b, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_holder, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_holder;
  constexpr uint32_t idesc_f = make_idesc(128, 128, 0, 1, 2u);   // X K-major, W1 MN-major, tf32
  constexpr uint32_t idesc_b = make_idesc(128, 128, 1, 1, 2u);   // X^T and dA both MN-major

  auto row_of = [&](int st, int b) -> int {
    return a.batch_idx ? __ldg(a.batch_idx + (sbase + st) * B + b) : (int)((sbase + st) * a.rows_per_step + b);
  };
  auto pull_w1 = [&]() {            // thread 0: this CTA's W1 slice, straight from the shard, in operand layout
    mbar_expect_tx(&bar_w1, 4u * (uint32_t)a.ksl * 128u);     // four boxes of ksl rows (the rows behind them stay zero)
#pragma unroll
    for (int j = 0; j < 4; ++j) tma_load_2d(W1s + j * kBox, &tmW1, &bar_w1, 32 * j, k0);
  };
  // The batch slice in the forward product's layout: thread b fetches row b of the batch as four 128-byte TMA
  // boxes, which land swizzled where the K-major descriptor expects them (columns behind the slice multiply the
  // zero rows of W1s; rows beyond B are never written and stay zero).  The backward layout is a permutation of
  // the 16-byte chunks, made in shared memory during phase B.  cp.async is not used for this: thousands of
  // 16-byte gathers per step queue up in the load/store unit and stall every shuffle and shared-memory access
  // behind them (measured: 2.1 us to issue, phase B three times slower).
  const int xgroups = (B + 3) / 4;                       // TMA gather4: four batch rows x 32 features per instruction
  const uint32_t xbytes = (uint32_t)xgroups * 4u * 512u;
  auto load_x = [&](const int* rows) {                   // lanes 0..6 of every warp: 112 >= 26 groups x 4 K-boxes
    const int t = warp * 7 + lane;
    if (lane < 7 && t < xgroups * 4) {
      const int g = t >> 2, kb = t & 3, r0 = 4 * g;
      const int i0 = rows[r0], i1 = rows[min(r0 + 1, B - 1)], i2 = rows[min(r0 + 2, B - 1)], i3 = rows[min(r0 + 3, B - 1)];
      tma_gather4_2d(Xk + kb * kBox + (uint32_t)(r0 >> 3) * 1024u + (uint32_t)(r0 & 7) * 128u, &tmX, &bar_x,
                     k0 + 32 * kb, i0, i1, i2, i3);
    }
  };

  // ---- prologue: rows, X and W1 of step 0; small parameters straight from the shard -------------------
  if (tid < B) rows_cur[tid] = row_of(0, tid);
  for (int e = tid; e < dWords; e += kT) {
    float v = 0.f;
    if (e < dOffB1) { if (e < H * C) v = dtfe::ld_relaxed_sys(a.W2 + e); }
    else if (e < dOffB2) { if (e - dOffB1 < H) v = dtfe::ld_relaxed_sys(a.b1 + (e - dOffB1)); }
    else if (e - dOffB2 < C) v = dtfe::ld_relaxed_sys(a.b2 + (e - dOffB2));
    cur[cur_slot(e)] = v;
  }
  // zero what no copy ever writes: batch rows beyond B in Xk, feature rows beyond the slice in W1s
  for (int i = tid; i < (int)(kBuf / 16); i += kT) {
    reinterpret_cast<float4*>(Xk)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    reinterpret_cast<float4*>(W1s)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  fence_proxy_async();
  __syncthreads();
  if (tid == 0) { mbar_expect_tx(&bar_x, xbytes); pull_w1(); }
  __syncthreads();
  load_x(rows_cur);
  int pending_step = 0;
  uint32_t ph = 0;                                      // mbarrier parity: every barrier completes once per step
  constexpr int kSm = (kMaxH * kMaxC + kMaxH + kMaxC + 8 + kT - 1) / kT;   // small-parameter words per thread

  for (int s = 0; s < a.nsteps; ++s, ph ^= 1) {
    TC_STAMP(0);
    // label of this warp's batch row and the rows of the next step: issued now, used later
    const int myrow = rank + kCl * warp;
    int ylab = 0;
    if (lane == 0 && myrow < B) ylab = __ldg(a.labels + rows_cur[myrow]);
    int rnext = 0;
    if (tid < B && s + 1 < a.nsteps) rnext = row_of(s + 1, tid);

    // ------------------------------ forward product ---------------------------------------------
    dtfe::cp_async_wait<0>();                           // (warps 13..15) the small parameters fetched a step ago have landed
    fence_proxy_async();                                // last step's update of W1s is visible to the tensor core's proxy
    __syncthreads();
    mbar_wait(&bar_x, ph);                              // the batch slice has landed (every thread: warps 13..15 read it too)
    if (tid == 0) {
      mbar_wait(&bar_w1, ph);                           // W1 slice (TMA) has landed
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < kRowsPad / 8; ++ks) {
        const uint64_t ad = make_smem_desc(smem_u32(Xk) + (ks >> 2) * kBox + (ks & 3) * 32, 16, 1024);
        const uint64_t bd = make_smem_desc(smem_u32(W1s) + ks * 1024, kBox, 512, 1);
        umma_1sm<DT_TF32>(tmem, ad, bd, idesc_f, ks != 0);
      }
      umma_commit(&bar_f);
    }
    mbar_wait(&bar_f, ph);
    tc_fence_after();
    TC_STAMP(1);
    if (tid == 0 && s + 1 < a.nsteps) {
      fence_proxy_async_all();                          // after last step's pushes (ordered before by the barriers since)
      pull_w1();                                        // next step's slice, BEFORE this CTA pushes to it
    }
    {
      // split-K partial of this CTA -> scratch: warp = (lane quarter, 32-column group)
      const int q = warp & 3, cg = warp >> 2, b = q * 32 + lane;
      uint32_t r[32];
      __syncwarp();
      tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(cg * 32), r);
      tmem_ld_wait();
      TC_STAMP(9);
      // A thread owns a row of the accumulator; stored directly, a warp's 32 rows are 400 bytes apart.  Stage the
      // tile in the backward operand buffer (dead until phase B rebuilds it) with the float4 index XOR-ed by the row,
      // then write whole rows: 25 lanes x 16 bytes, contiguous.
      if (b < kRowsPad) {
        float4* row = reinterpret_cast<float4*>(Xmn + (size_t)b * 512);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          row[(cg * 8 + j) ^ (b & 31)] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                                     __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
      }
    }
    tc_fence_before();
    __syncthreads();
    for (int b = warp; b < B; b += kT / 32) {
      if (4 * lane < H)
        __stcg(reinterpret_cast<float4*>(a.part + ((long)rank * B + b) * H) + lane,
               reinterpret_cast<const float4*>(Xmn + (size_t)b * 512)[lane ^ (b & 31)]);
    }
    TC_STAMP(2);
    cluster_barrier();                                  // all 8 partials are in scratch
    TC_STAMP(3);

    // ------------------------------ phase B: a warp per batch row --------------------------------
    if (rank == book && tid == 0 && s > 0) pending_step = dtfe::atom_add_relaxed_sys_s32(a.global_step, 1);
    // everything this phase reads from L2 is requested first: this worker's small-parameter updates of the
    // previous step and the 8 partials of this warp's row
    float dl[kSm];
#pragma unroll
    for (int i = 0; i < kSm; ++i) {
      const int e = tid + i * kT;
      dl[i] = (s > 0 && e < dWords) ? __ldcg(a.delta + e) : 0.f;
    }
    const int j0 = 4 * lane;
    const bool act = myrow < B && j0 < H;
    float4 pv[kCl];
#pragma unroll
    for (int i = 0; i < kCl; ++i)
      pv[i] = act ? __ldcg(reinterpret_cast<const float4*>(a.part + ((long)i * B + myrow) * H + j0)) : make_float4(0.f, 0.f, 0.f, 0.f);
    if (s > 0) {
      // small parameters = what was fetched a step ago + this worker's own updates since (exact, see header)
#pragma unroll
      for (int i = 0; i < kSm; ++i) {
        const int e = tid + i * kT;
        if (e < dWords) cur[cur_slot(e)] = __fadd_rn(stg[e], dl[i]);
      }
    }
    if (tid < B) rows_nxt[tid] = rnext;
    if (tid == 0 && s + 1 < a.nsteps) mbar_expect_tx(&bar_x, xbytes);   // next batch, issued after barrier 2
    __syncthreads();
    TC_STAMP(10);
    if (tid >= kLoadT) {
      if (s + 1 < a.nsteps) {
        // fetch the small parameters for the next step now: between the two barriers no CTA of this worker pushes
        for (int f = tid - kLoadT; 4 * f < dWords; f += kT - kLoadT) {
          const int e = 4 * f;
          const float* src; int n;
          if (e < dOffB1) { n = H * C - e; src = a.W2 + e; }
          else if (e < dOffB2) { n = H - (e - dOffB1); src = a.b1 + (e - dOffB1); }
          else { n = C - (e - dOffB2); src = a.b2 + (e - dOffB2); }
          dtfe::cp_async_cg16(stg + e, src, n >= 4 ? 16 : (n > 0 ? 4 * n : 0));
        }
      }
      dtfe::cp_async_commit();
      // the batch slice in the backward product's layout: a permutation of the 16-byte chunks of Xk
      for (int t = tid - kLoadT; t < kRowsPad * kChunks; t += kT - kLoadT) {
        const int b = t / kChunks, c = t - b * kChunks;
        *reinterpret_cast<float4*>(Xmn + off_mnmajor(b, c)) = *reinterpret_cast<const float4*>(Xk + off_kmajor(b, c));
      }
      fence_proxy_async();                              // visible to the tensor core once the cluster barrier is passed
    }
    if (myrow < B) {
      float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int i = 0; i < kCl; ++i) { z.x += pv[i].x; z.y += pv[i].y; z.z += pv[i].z; z.w += pv[i].w; }
      float h[4] = {0.f, 0.f, 0.f, 0.f};
      if (act) {
        h[0] = dtfe::sigmoidf_fast(z.x + b1s[j0]); h[1] = dtfe::sigmoidf_fast(z.y + b1s[j0 + 1]);
        h[2] = dtfe::sigmoidf_fast(z.z + b1s[j0 + 2]); h[3] = dtfe::sigmoidf_fast(z.w + b1s[j0 + 3]);
        __stcg(reinterpret_cast<float4*>(a.Hbuf + (long)myrow * H + j0), make_float4(h[0], h[1], h[2], h[3]));
      }
      TC_STAMP(11);
      // logits: all classes' partial sums first, then their butterflies interleaved; class c ends up in lane c
      float p[kMaxC];
#pragma unroll
      for (int c = 0; c < kMaxC; ++c) {
        p[c] = 0.f;
        if (c < C && act) {
          const float4 w = *reinterpret_cast<const float4*>(W2t + c * H + j0);
          p[c] = fmaf(h[0], w.x, fmaf(h[1], w.y, fmaf(h[2], w.z, h[3] * w.w)));
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int c = 0; c < kMaxC; ++c)
          if (c < C) p[c] += __shfl_xor_sync(0xffffffffu, p[c], o);
      }
      float zc = -INFINITY;
#pragma unroll
      for (int c = 0; c < kMaxC; ++c)
        if (c < C && lane == c) zc = p[c] + b2s[c];
      float mx = zc;
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      const int am = __ffs(__ballot_sync(0xffffffffu, lane < C && zc == mx)) - 1;      // first index wins ties
      const float ex = (lane < C) ? __expf(zc - mx) : 0.f;
      float se = ex;
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
      const int y = __shfl_sync(0xffffffffu, ylab, 0);
      const float zy = __shfl_sync(0xffffffffu, zc, y);
      float dz = 0.f;
      if (lane < C) {
        dz = (ex / se - (lane == y ? 1.f : 0.f)) * invB;
        __stcg(a.dZ + (long)myrow * C + lane, dz);
      }
      if (lane == 0) {
        __stcg(a.rowloss + myrow, mx + __logf(se) - zy);
        __stcg(a.rowcorrect + myrow, (am == y) ? 1 : 0);
      }
      TC_STAMP(12);
      float dh[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < kMaxC; ++c) {
        if (c < C) {
          const float dzc = __shfl_sync(0xffffffffu, dz, c);
          if (act) {
            const float4 w = *reinterpret_cast<const float4*>(W2t + c * H + j0);
            dh[0] = fmaf(dzc, w.x, dh[0]); dh[1] = fmaf(dzc, w.y, dh[1]);
            dh[2] = fmaf(dzc, w.z, dh[2]); dh[3] = fmaf(dzc, w.w, dh[3]);
          }
        }
      }
      if (act)
        __stcg(reinterpret_cast<float4*>(a.dA + (long)myrow * H + j0),
               make_float4(dh[0] * h[0] * (1.f - h[0]), dh[1] * h[1] * (1.f - h[1]),
                           dh[2] * h[2] * (1.f - h[2]), dh[3] * h[3] * (1.f - h[3])));
    }
    TC_STAMP(4);
    cluster_barrier();                                  // dA, dZ, H and the row metrics of all rows are in scratch
    TC_STAMP(5);

    // ------------------------------ backward product ---------------------------------------------
    // Warp 15 feeds the tensor core (dA by TMA, then the MMAs) while warps 0..12 stage dZ and go straight on to
    // the small gradients; nobody waits for the issuing thread.
    const bool small_all = (H + kCl - 1) / kCl > 13 || C > 13;    // more hidden units per CTA than loader warps: everybody helps
    if (warp == 15 && lane == 0) {                      // dA in operand layout: four TMA boxes (zeros beyond B rows / H columns)
      fence_proxy_async_all();                          // (written with ordinary stores by the other CTAs before the barrier)
      mbar_expect_tx(&bar_da, kBuf);
#pragma unroll
      for (int j = 0; j < 4; ++j) tma_load_2d(dAs + j * kBox, &tmDA, &bar_da, 32 * j, 0);
    }
    // next step's batch: Xk is free (forward MMAs done, backward layout copied out of it)
    if (s + 1 < a.nsteps) load_x(rows_nxt);
    if (tid < kLoadT) {
      dtfe::cta_copy_async(dZs, a.dZ, B * C, tid, kLoadT);
      dtfe::cp_async_commit();
    }
    if (rank == book && warp == 14) {                   // metrics of this step, fixed summation order
      float l = 0.f; int cnt = 0;
      for (int r = lane; r < B; r += 32) { l += __ldcg(a.rowloss + r); cnt += __ldcg(a.rowcorrect + r); }
      l = dtfe::warp_sum(l);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
      if (lane == 0) { a.loss_out[s] = l * invB; a.acc_out[s] = (float)cnt * invB; }
    }
    TC_STAMP(13);
    if (warp == 15 && lane == 0) {
      mbar_wait(&bar_da, ph);
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < kRowsPad / 8; ++ks) {
        const uint64_t ad = make_smem_desc(smem_u32(Xmn) + ks * 1024, kBox, 512, 1);
        const uint64_t bd = make_smem_desc(smem_u32(dAs) + ks * 1024, kBox, 512, 1);
        umma_1sm<DT_TF32>(tmem + 128u, ad, bd, idesc_b, ks != 0);
      }
      umma_commit(&bar_b);
    }
    if (tid < kLoadT) dtfe::cp_async_wait<0>();         // dZ has landed
    if (small_all) __syncthreads();
    else if (tid < kLoadT) named_bar_sync(1, kLoadT);
    TC_STAMP(6);
    // small gradients while the tensor core works: hidden unit tt = rank + 8*warp -> dW2[tt][:], db1[tt]
    {
      const int per = (H + kCl - 1) / kCl;                // hidden units per CTA, a contiguous block (shared sectors)
      const int tt = rank * per + warp;
      if (warp < per && tt < H) {
        float hc[4], dac = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int b = lane + 32 * i;
          hc[i] = b < B ? a.Hbuf[(long)b * H + tt] : 0.f;      // written before the cluster barrier, whose acquire
          if (b < B) dac += a.dA[(long)b * H + tt];            // drops L1: ordinary cached loads are coherent here
        }
        float v[kMaxC];
#pragma unroll
        for (int c = 0; c < kMaxC; ++c) {
          v[c] = 0.f;
          if (c < C) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int b = lane + 32 * i;
              if (b < B) v[c] = fmaf(hc[i], dZs[b * C + c], v[c]);
            }
          }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
          for (int c = 0; c < kMaxC; ++c)
            if (c < C) v[c] += __shfl_xor_sync(0xffffffffu, v[c], o);
          dac += __shfl_xor_sync(0xffffffffu, dac, o);
        }
        // lane c publishes class c, lane C the bias gradient
        float u = 0.f;
#pragma unroll
        for (int c = 0; c < kMaxC; ++c)
          if (c < C && lane == c) u = __fmul_rn(nlr, v[c]);
        if (lane < C) {
          __stcg(a.delta + (long)tt * C + lane, u);
          dtfe::red_add_sys(a.W2 + (long)tt * C + lane, u);
        } else if (lane == C) {
          u = __fmul_rn(nlr, dac);
          __stcg(a.delta + dOffB1 + tt, u);
          dtfe::red_add_sys(a.b1 + tt, u);
        }
      }
      if (rank == book && warp < C) {                   // db2[c], one warp per class
        float v = 0.f;
        for (int b = lane; b < B; b += 32) v += dZs[b * C + warp];
        v = dtfe::warp_sum(v);
        if (lane == 0) {
          const float u = __fmul_rn(nlr, v);
          __stcg(a.delta + dOffB2 + warp, u);
          dtfe::red_add_sys(a.b2 + warp, u);
        }
      }
    }
    TC_STAMP(14);
    mbar_wait(&bar_b, ph);
    tc_fence_after();
    TC_STAMP(7);
    if (s + 1 < a.nsteps) mbar_wait(&bar_w1, ph ^ 1);   // next step's W1 slice is in W1s: add this step's update to it
    {
      // dW1 rows of this slice: lane = feature, 32 hidden units per warp -> push (= apply) and own delta
      const int q = warp & 3, cg = warp >> 2, f = q * 32 + lane;
      uint32_t r[32];
      __syncwarp();
      tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + 128u + (uint32_t)(cg * 32), r);
      tmem_ld_wait();
      if (f < nk) {
        float* wrow = a.W1 + (long)(k0 + f) * H + cg * 32;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const int hcol = cg * 32 + j;
          if (hcol < H) {
            const float4 u = make_float4(__fmul_rn(nlr, __uint_as_float(r[j])), __fmul_rn(nlr, __uint_as_float(r[j + 1])),
                                         __fmul_rn(nlr, __uint_as_float(r[j + 2])), __fmul_rn(nlr, __uint_as_float(r[j + 3])));
            dtfe::red_add_sys_v4(wrow + j, u);
            if (s + 1 < a.nsteps) {
              float4* w = reinterpret_cast<float4*>(W1s + off_mnmajor(f, hcol >> 2));
              float4 o = *w;
              o.x = __fadd_rn(o.x, u.x); o.y = __fadd_rn(o.y, u.y); o.z = __fadd_rn(o.z, u.z); o.w = __fadd_rn(o.w, u.w);
              *w = o;
            }
          }
        }
      }
    }
    tc_fence_before();
    if (rank == book && tid == 0 && s > 0) a.step_out[s - 1] = pending_step;
    // rows of the next step become current
    __syncthreads();
    if (tid < B) rows_cur[tid] = rows_nxt[tid];
    __syncthreads();
    TC_STAMP(8);
  }

  // ---- exit: every push of this launch is issued; count the last step ------------------------------
  __threadfence();
  cluster_barrier();
  if (rank == book && tid == 0 && a.nsteps > 0) {
    if (a.remote) __threadfence_system();
    a.step_out[a.nsteps - 1] = dtfe::atom_add_sys_s32(a.global_step, 1);
    if (a.cursor) *a.cursor = (int)sbase + a.nsteps;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

size_t tc_smem_bytes(int H, int C) {
  const size_t dWords = (((size_t)H * C + 3) & ~(size_t)3) + (((size_t)H + 3) & ~(size_t)3) + (((size_t)C + 3) & ~(size_t)3);
  return 4 * (size_t)kBuf + 2 * dWords * 4 + (size_t)kRowsPad * kMaxC * 4 + 2 * (size_t)kRowsPad * 4 + 1024;
}

}  // namespace

// 1 when dtfe_mlp2_step_tc can run this shape on the current device, else 0 (the reason is in dtfe_last_error()).
DTFE_API int dtfe_mlp2_step_tc_supported(int B, int D, int H, int C) {
  if (!(B >= 1 && B <= kRowsPad && H >= 4 && H <= kMaxH && H % 4 == 0 && C >= 1 && C <= kMaxC && D >= 1 &&
        D <= kCl * kRowsPad)) {
    dtfe_set_error("mlp2_step_tc: needs B <= %d, D <= %d, H <= %d with H %% 4 == 0, C <= %d", kRowsPad, kCl * kRowsPad,
                   kMaxH, kMaxC);
    return 0;
  }
  if (tc_smem_bytes(H, C) + 1024 > 232448) {
    dtfe_set_error("mlp2_step_tc: H = %d, C = %d need %zu bytes of shared memory for the small parameters; the operand "
                   "buffers leave %zu", H, C, tc_smem_bytes(H, C) - 4 * (size_t)kBuf, (size_t)232448 - 1024 - 4 * (size_t)kBuf);
    return 0;
  }
  return 1;
}

// The tensor-core form of dtfe_mlp2_step: same launch struct, same workspace (dtfe_mlp2_step_workspace_words),
// same results.  Asynchronous training with the lossless push only (train = 1, atomic = 1, no gradient buffers, no
// sync_exchange); `ks` is ignored.  Parameters must be 16-byte aligned.
DTFE_API int dtfe_mlp2_step_tc(const dtfe_mlp2_launch* L, void* stream) {
  if (!dtfe_mlp2_step_tc_supported(L->B, L->D, L->H, L->C)) return -1;
  DTFE_REQUIRE(L->train && L->atomic && !L->gW1 && !L->sync_exchange,
               "mlp2_step_tc: asynchronous training with the lossless push only (use dtfe_mlp2_step otherwise)");
  DTFE_REQUIRE(L->nsteps >= 1, "mlp2_step_tc: nsteps must be at least 1 (got %d)", L->nsteps);
  DTFE_REQUIRE((((uintptr_t)L->W1 | (uintptr_t)L->W2 | (uintptr_t)L->b1 | (uintptr_t)L->b2 | (uintptr_t)L->workspace) % 16) == 0,
               "mlp2_step_tc: parameters and workspace must be 16-byte aligned");
  DTFE_REQUIRE(((uintptr_t)L->X % 16) == 0 && (L->ldx % 4) == 0, "mlp2_step_tc: X must be 16-byte aligned with a row pitch that is a multiple of 4");
  TcArgs a{};
  a.B = L->B; a.D = L->D; a.H = L->H; a.C = L->C;
  a.W1 = L->W1; a.b1 = L->b1; a.W2 = L->W2; a.b2 = L->b2; a.global_step = L->global_step;
  a.X = L->X; a.ldx = L->ldx; a.labels = L->labels; a.batch_idx = L->batch_idx; a.rows_per_step = L->rows_per_step;
  a.ksl = (((L->D + kCl - 1) / kCl) + 3) & ~3;
  DTFE_REQUIRE(a.ksl <= kRowsPad, "mlp2_step_tc: D = %d needs %d features per CTA (limit %d)", L->D, a.ksl, kRowsPad);
  float* w = L->workspace;
  auto carve = [&](size_t n) { float* p = w; w += (n + 3) & ~(size_t)3; return p; };
  a.part = carve((size_t)kCl * L->B * L->H);
  a.dA = carve((size_t)L->B * L->H);
  a.Hbuf = carve((size_t)L->B * L->H);
  a.dZ = carve((size_t)L->B * L->C);
  a.rowloss = carve(L->B);
  a.rowcorrect = (int*)carve(L->B);
  a.delta = carve((size_t)L->H * L->C + L->H + L->C + 12);
  a.loss_out = L->loss_out; a.acc_out = L->acc_out; a.step_out = L->step_out;
  a.lr = L->lr; a.nsteps = L->nsteps; a.remote = L->remote; a.cursor = L->cursor; a.dbg = L->dbg;
  CUtensorMap tm, tx, tda;
  // W1 [D][H]: boxes of 32 hidden units x ksl feature rows, 32-byte-base swizzle (MN-major tf32 operand)
  if (int rc = make_tmap(&tm, L->W1, (uint64_t)L->H, (uint64_t)L->D, (uint64_t)L->H, 32, (uint32_t)a.ksl, 4, true)) return rc;
  // X rows (any row of the data set): boxes of 32 features x 1 row, 128-byte swizzle (K-major operand)
  if (int rc = make_tmap(&tx, L->X, (uint64_t)L->D, (uint64_t)1 << 31, (uint64_t)L->ldx, 32, 1, 4, false)) return rc;
  // dA [B][H] (scratch): boxes of 32 hidden units x 104 batch rows, like W1
  if (int rc = make_tmap(&tda, a.dA, (uint64_t)L->H, (uint64_t)L->B, (uint64_t)L->H, 32, kRowsPad, 4, true)) return rc;
  const size_t smem = tc_smem_bytes(L->H, L->C);
  static size_t configured[64] = {};      // dynamic shared memory the kernels are currently allowed, per device
  int dev;
  DTFE_CUDA_CHECK(cudaGetDevice(&dev));
  DTFE_REQUIRE(dev >= 0 && dev < 64, "mlp2_step_tc: device ordinal %d not supported", dev);
  if (configured[dev] < smem) {
    cudaFuncAttributes fa;
    DTFE_CUDA_CHECK(cudaFuncGetAttributes(&fa, k_mlp2_step_tc<0>));
    int optin = 0;
    DTFE_CUDA_CHECK(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    DTFE_REQUIRE(smem + fa.sharedSizeBytes <= (size_t)optin,
                 "mlp2_step_tc: needs %zu + %zu bytes of shared memory, the device allows %d", smem,
                 (size_t)fa.sharedSizeBytes, optin);
    DTFE_CUDA_CHECK(cudaFuncSetAttribute(k_mlp2_step_tc<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DTFE_CUDA_CHECK(cudaFuncSetAttribute(k_mlp2_step_tc<10>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev] = smem;
  }
  if (L->C == 10) k_mlp2_step_tc<10><<<dim3(kCl), dim3(kT), smem, (cudaStream_t)stream>>>(tm, tx, tda, a);
  else k_mlp2_step_tc<0><<<dim3(kCl), dim3(kT), smem, (cudaStream_t)stream>>>(tm, tx, tda, a);
  DTFE_LAUNCH_CHECK();
  return 0;
}
