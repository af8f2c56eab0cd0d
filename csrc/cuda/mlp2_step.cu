// The whole parameter-server training step of the reference's MNIST MLP as ONE
// persistent cooperative kernel: pull -> forward -> loss/accuracy -> backward ->
// push+apply -> global_step, for `nsteps` consecutive steps without returning to
// the host.
//
// Reference step (worker.py:129-141): sess.run(sync_op) [4 Assign pulls over gRPC],
// sess.run([train_op, loss, summary, global_step]) [fwd/bwd on GPU, 4 gradient
// tensors over gRPC, ApplyGradientDescent on the ps CPU], sess.run(counter_op).
// Three RPC round trips and ~25 TF kernels.  Here:
//
//   phase A  each CTA owns a KS-row slice of W1 ([D,H], reference layout [in,out]).
//            It reads that slice *directly from the ps shard* (possibly a peer GPU's
//            HBM: this is the pull) and the matching KS columns of the batch, and
//            writes a split-K partial of X.W1 to scratch.
//   barrier
//   phase B  one batch row per CTA (the grid has max(slices, B+1) CTAs; the ones beyond the
//            slices only work here and on the small gradients): reduce the partials, +b1,
//            sigmoid, logits with W2/b2, softmax cross-entropy fwd+bwd, accuracy,
//            dA = (dZ.W2^T) * h(1-h).
//   barrier
//   phase C  each CTA computes dW1 for the SAME KS rows it pulled (X slice is still in
//            shared memory) and pushes -lr*dW1 into the shard with red.add.sys: the
//            push is the apply, there is no ps-side kernel.  dW2/db1/db2 rows are
//            spread over all CTAs.  The last CTA of the grid (which owns no batch row when
//            the grid is B+1 wide) keeps the books: metrics, db2, and the system-scope
//            global_step atomic whose pre-increment value is the `step` the reference fetches.
//
// The pull runs one step ahead ("ahead" mode, asynchronous training): a thread reads its W1
// elements BEFORE it pushes to them and adds its own update in registers; the small parameters
// are fetched (cp.async / registers) during phase B for the NEXT step and completed there with
// this worker's own deltas from a local buffer.  Own updates are therefore seen exactly (bitwise
// what the shard holds when no other worker interferes), other workers' updates one step later,
// and no load from a peer GPU is ever waited for on the critical path.
//
// Scratch (global memory beyond inputs/outputs): the split-K partials, H, dA, dZ,
// per-row loss/correct flags, the small-parameter deltas and four sync words.
#include "common.cuh"
#include <math.h>
#include <stdlib.h>
#include <time.h>

namespace {

constexpr int kThreads = 256;
constexpr int kMaxB = 128, kMaxH = 128, kMaxC = 16;

struct Mlp2Args {
  int B, D, H, C;
  float *W1, *b1, *W2, *b2;
  int* global_step;
  const float* X; long ldx;
  const int* labels;
  const int* batch_idx;     // [nsteps*B] dataset rows, or null
  long rows_per_step;       // used when batch_idx is null: row = s*rows_per_step + r
  float* part;              // [G][B][H]
  float* Hbuf;              // [B][H]
  float* dA;                // [B][H]
  float* dZ;                // [B][C]
  float* rowloss;           // [B]
  int* rowcorrect;          // [B]
  float* delta;             // [H*C + H + C (+pad)] this worker's last small-parameter updates (ahead mode)
  float* mirror;            // [2][same] the small parameters as the bookkeeper last fetched them (ahead mode)
  unsigned* sync;           // [4]: grid barrier, exit count, phase-D go words
  float* loss_out;          // [nsteps]
  float* acc_out;           // [nsteps]
  int* step_out;            // [nsteps]
  float lr;
  int nsteps;
  int atomic;
  int train;                // 0: forward/metrics only (no phase C)
  int nslices;              // G: CTAs that own a KS-row slice of W1 (the grid may be wider)
  int ahead;                // pull one step ahead + own deltas (see the file header)
  int remote;               // some shard lives on a peer GPU: completion fence must be .sys
  long long* dbg;           // optional [nsteps][grid][16] globaltimer stamps (profiling)
  int poison;               // debugging aid (DTFE_MLP2_POISON=1): scratch is overwritten with NaN as soon as its last
                            // legitimate reader is done, so a read that races with the next producer shows up as NaN
  float *gW1, *gb1, *gW2, *gb2;  // sync mode: raw gradients are stored here instead of being pushed
  int* cursor;              // optional device-side step cursor: batches start at *cursor (graph replay)
  // In-kernel synchronous exchange (phase D): reduce-scatter of all ranks' gradient buffers + apply,
  // with flag barriers in peer memory, once per step.  has_sx = 0: none.
  int has_sx;
  dtfe_sync_exchange sx;
};

// Release/acquire fences at GPU scope: __threadfence() is a sequentially consistent fence (MEMBAR.SC), which
// the arrive/poll pattern does not need; fence.acq_rel is MEMBAR.ALL.
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

__device__ __forceinline__ void grid_barrier(unsigned* ctr, unsigned& target, unsigned G) {
  __syncthreads();
  if (threadIdx.x == 0) {
    target += G;
    fence_acq_rel_gpu();
    asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" :: "l"(ctr) : "memory");
    unsigned v;
    do {
      asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
    } while (v < target);
    fence_acq_rel_gpu();
  }
  __syncthreads();
}

__device__ __forceinline__ long long gtime() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define DBG_STAMP(i) do { if (a.dbg && tid == 0) a.dbg[((long)s * NC + g) * 16 + (i)] = gtime(); } while (0)

__device__ __forceinline__ void push1(float* p, float v, int atomic) {
  if (atomic) dtfe::red_add_sys(p, v);
  else *p = dtfe::ld_relaxed_sys(p) + v;
}
// Low-latency synchronous exchange (dtfe_sync_exchange mode 2): where to send a gradient element.
struct LLSend {
  float* const* recv;       // every rank's receive buffer
  long slot;                // pair index of element 0 of THIS rank's area in the current copy: (copy * world + rank) * stride
  int world;
  unsigned epoch;
};
__device__ __forceinline__ void ll_send(const LLSend& ll, long idx, float g) {
  const float tag = __uint_as_float(ll.epoch);
  for (int p = 0; p < ll.world; ++p)       // one 8-byte store per rank (own buffer included): value + step number
    asm volatile("st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" :: "l"(ll.recv[p] + 2 * (ll.slot + idx)), "f"(g), "f"(tag) : "memory");
}
// Asynchronous mode: push -lr*g into the shard.  Synchronous mode: leave the raw gradient in this
// rank's exchange buffer for the reduce-scatter kernel (or send it, low-latency mode: gbuf then only
// carries the element's offset in the flat gradient).
__device__ __forceinline__ void emit(float* shard, float* gbuf, long i, float g, float nlr, int atomic,
                                     const LLSend* ll = nullptr, const float* gbase = nullptr) {
  if (ll) ll_send(*ll, (gbuf - gbase) + i, g);
  else if (gbuf) __stcg(gbuf + i, g);
  else push1(shard + i, __fmul_rn(nlr, g), atomic);
}
// Small parameters in ahead mode: the update also goes to the local delta buffer every CTA reads next step.
__device__ __forceinline__ void emit_small(float* shard, float* gbuf, float* delta, long i, float g, float nlr,
                                           int atomic, const LLSend* ll = nullptr, const float* gbase = nullptr) {
  if (delta) __stcg(delta + i, __fmul_rn(nlr, g));
  emit(shard, gbuf, i, g, nlr, atomic, ll, gbase);
}

__device__ __forceinline__ void st_relaxed_sys_u32(unsigned* p, unsigned v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
constexpr int kCtaFlagPitch = 160;      // per-CTA arrival words per rank (dtfe_sync_exchange.cta_flags)
constexpr unsigned kSxAbort = 0xffffffffu;

// Wait for a peer's flag; gives up (returns false) when the stop word is raised or the timeout expires, so a
// rank that left can never hang this GPU.
__device__ __forceinline__ bool sx_wait(const unsigned* flag, unsigned epoch, const dtfe_sync_exchange& sx,
                                        long long t0) {
  unsigned it = 0;
  while (dtfe::ld_acquire_sys(flag) < epoch) {
    if ((++it & 0xff) == 0) {
      if (sx.abort_flag && dtfe::ld_acquire_sys(sx.abort_flag) != 0) return false;
      if (sx.timeout_ns && (unsigned long long)(gtime() - t0) > sx.timeout_ns) return false;
    }
  }
  return true;
}

// FULL = false is the lean instantiation for plain asynchronous training and evaluation: no gradient buffers, no
// synchronous exchange (phase D), no debugging aids.  Keeping that code out of the headline path is worth 0.9 us per
// step (12.99 -> 12.11 us measured on one B200; an attempt to poll all ranks' pairs at once inside phase D had cost
// the asynchronous path 2.4 us through register pressure alone, which is what prompted the split).
template <int KS, bool FULL>
__global__ void __launch_bounds__(kThreads, 1) k_mlp2_step(Mlp2Args a) {
  static_assert(KS % 4 == 0 && KS <= 16, "KS must be 4, 8, 12 or 16");
  extern __shared__ __align__(16) float smem[];
  const int B = a.B, D = a.D, H = a.H, C = a.C;
  const int HP = H;  // row pitch of H-wide smem tiles
  float* Xs = smem;                       // [B][KS]
  float* Ws = Xs + kMaxB * KS;            // [KS][H]
  float* W2s = Ws + KS * kMaxH;           // [H][C]
  float* b1s = W2s + kMaxH * kMaxC;       // [H]
  float* b2s = b1s + kMaxH;               // [C]
  float* hs = b2s + kMaxC;                // [H] (2*kMaxH reserved)
  float* zs = hs + 2 * kMaxH;             // [8][C] per-warp partial logits
  float* dzs = zs + 8 * kMaxC;            // [C] (2*kMaxC reserved)
  float* red = dzs + 2 * kMaxC;           // [2][kMaxB] phase C: the H columns of this CTA's hidden units
  float* dZs = red + kThreads;            // [B][C]
  float* dAs = dZs + kMaxB * kMaxC;       // [B][H]  (phase C)
  float* red_part = dAs;                  // [16][kMaxH] phase-B scratch, aliases dAs (never live together)
  // ahead mode, fetching CTAs: the quads of the small parameters this CTA fetched from the shard (cp.async
  // destination, one slot per thread and round), republished in the mirror at the end of phase B
  float* stg = dAs + kMaxB * kMaxH;       // [kMaxH*kMaxC + kMaxH + kMaxC]
  float* dwp = stg + kMaxH * kMaxC + kMaxH + kMaxC;   // [8][KS][kMaxH] per-warp partial dW1 tiles (phase C)

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned NC = gridDim.x;          // CTAs in the grid
  const unsigned G = (unsigned)a.nslices; // the first G of them own a W1 slice
  const int g = blockIdx.x;
  const bool slice = g < (int)G;
  const int book = (int)NC - 1;           // the CTA that keeps the books (no batch row when NC > B)
  const int k0 = g * KS;
  unsigned bar_target = 0;
  const float invB = 1.f / (float)B;

  // ---- software prefetch state -------------------------------------------------------------------
  // Batch rows are data dependent (index -> row), i.e. two chained global loads.  Indices are fetched
  // two steps ahead and the X slice one step ahead, so neither latency is ever on the critical path.
  // device-side cursor: this launch processes batches cursor .. cursor+nsteps-1 of batch_idx
  const long sbase = a.cursor ? (long)*reinterpret_cast<volatile int*>(a.cursor) : 0;
  constexpr int kXPer = (kMaxB * KS + kThreads - 1) / kThreads;
  float xnext[kXPer];
  int rownext[kXPer];       // kept as loaded (int): converting would make the warp wait for the load right here
  int ylab_row = 0;         // dataset row of the first batch row this thread's slot handles in phase B
  auto fetch_rows = [&](int st) {
#pragma unroll
    for (int i = 0; i < kXPer; ++i) {
      int e = tid + i * kThreads;
      int row = 0;
      if (slice && e < B * KS && st < a.nsteps) {
        int r = e / KS;
        row = a.batch_idx ? __ldg(a.batch_idx + (sbase + st) * B + r) : (int)((sbase + st) * a.rows_per_step + r);
      }
      rownext[i] = row;
    }
  };
  auto fetch_x = [&](int st) {   // consumes rownext (rows of step st)
#pragma unroll
    for (int i = 0; i < kXPer; ++i) {
      int e = tid + i * kThreads;
      float v = 0.f;
      if (slice && e < B * KS && st < a.nsteps) {
        int k = k0 + e % KS;
        if (k < D) v = __ldg(a.X + (long)rownext[i] * a.ldx + k);
      }
      xnext[i] = v;
    }
  };
  fetch_rows(0);
  fetch_x(0);
  fetch_rows(1);

  const int t = tid;                                         // phase B: one row per CTA, thread = hidden unit

  // Ownership of the W1 slice inside the CTA (used by the phase-C dW1 tiling AND by the pull): a warp owns
  // 4 consecutive rows and a contiguous block of columns; a lane owns up to two columns of those rows.
  constexpr int kGroups = KS / 4, kParts = 8 / kGroups;
  const int wgrp = warp % kGroups, wpart = warp / kGroups;
  const int wcq = (H + kParts - 1) / kParts;            // columns per part (<= 64)
  const int wj0 = wpart * wcq + lane, wj1 = wj0 + 32;
  const bool wv0 = lane < wcq && wj0 < H, wv1 = lane + 32 < wcq && wj1 < H;
  float wnext0[4], wnext1[4];
  auto fetch_w1 = [&]() {                               // the pull: this thread's elements of the W1 slice
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      int k = k0 + wgrp * 4 + q;
      wnext0[q] = (wv0 && k < D) ? dtfe::ld_relaxed_sys(a.W1 + (long)k * H + wj0) : 0.f;
      wnext1[q] = (wv1 && k < D) ? dtfe::ld_relaxed_sys(a.W1 + (long)k * H + wj1) : 0.f;
    }
  };
  const bool ahead = a.ahead != 0;
  const bool has_row = g < B;                           // this CTA handles batch row g in phase B
  // delta / mirror / staging layout (each part starts on a float4): [H*C] W2 | [H] b1 | [C] b2
  const int dOffB1 = (H * C + 3) & ~3, dOffB2 = dOffB1 + ((H + 3) & ~3), dWords = dOffB2 + ((C + 3) & ~3);
  float* const delta = ahead ? a.delta : nullptr;
  const int nW1 = slice ? (min(D, k0 + KS) - k0) * H : 0;   // floats of this CTA's W1 slice (contiguous)
  // Ahead mode: who fetches the small parameters from the shard.  The CTAs without a batch row share the
  // quads (a few 16-byte requests per SM instead of hundreds from one); without such CTAs the bookkeeper
  // does it alone.
  const int nfetch = (int)NC > B ? (int)NC - B : 1;
  const int fidx = (int)NC > B ? g - B : (g == book ? 0 : -1);
  constexpr int kFetchSlots = (kMaxH * kMaxC + kMaxH + kMaxC + 4 * kThreads - 1) / (4 * kThreads);
  auto small_quad_src = [&](int e, int& n) -> const float* {   // element e of the W2|b1|b2 layout
    if (e < dOffB1) { n = H * C - e; return a.W2 + e; }
    if (e < dOffB2) { n = H - (e - dOffB1); return a.b1 + (e - dOffB1); }
    n = C - (e - dOffB2);
    return a.b2 + (e - dOffB2);
  };
  // Synchronous all-reduce mode: the parameters this kernel is given are this rank's own copies; the W1 slice
  // stays in shared memory for the whole launch and is updated in place in phase D.
  const bool sx1 = (FULL && a.has_sx) && a.sx.mode >= 1;      // all-reduce forms: flags + peer loads (1) or low-latency pairs (2)
  const bool sxll = (FULL && a.has_sx) && a.sx.mode == 2;
  if (ahead || sx1) {
    dtfe::cta_copy_async(Ws, a.W1 + (long)k0 * H, nW1, tid, kThreads);
    dtfe::cp_async_commit();
  } else {
    fetch_w1();
  }
  // offsets (floats) of the four gradients inside every rank's exchange buffer
  const long sxOffW1 = sx1 ? (long)(a.gW1 - a.sx.grads[a.sx.rank]) : 0, sxOffB1 = sx1 ? (long)(a.gb1 - a.sx.grads[a.sx.rank]) : 0;
  const long sxOffW2 = sx1 ? (long)(a.gW2 - a.sx.grads[a.sx.rank]) : 0, sxOffB2 = sx1 ? (long)(a.gb2 - a.sx.grads[a.sx.rank]) : 0;
  int pending_step = 0;                                 // book CTA / thread 0: global step observed for the previous step
  const unsigned sx_epoch0 = (FULL && a.has_sx) ? *reinterpret_cast<volatile unsigned*>(a.sx.epoch) : 0u;
  const unsigned long long sx_t0 = gtime();
  __shared__ unsigned s_go;
  int steps_done = 0;
  auto fetch_label_row = [&](int st) {
    int r = g;
    if (r < B && st < a.nsteps)
      ylab_row = a.batch_idx ? __ldg(a.batch_idx + (sbase + st) * B + r) : (int)((sbase + st) * a.rows_per_step + r);
  };
  fetch_label_row(0);

  for (int s = 0; s < a.nsteps; ++s) {
    // ------------------------------ phase A ------------------------------------
    DBG_STAMP(0);
    int ylab = 0;
    // all-reduce mode alternates between two copies of the gradient buffer (see phase D)
    const long goff = (sx1 && !sxll) ? (long)((sx_epoch0 + (unsigned)s + 1u) & 1u) * (long)a.sx.grad_stride : 0;
    float* const gW1 = (FULL && a.gW1) ? a.gW1 + goff : nullptr;
    float* const gb1 = (FULL && a.gb1) ? a.gb1 + goff : nullptr;
    float* const gW2 = (FULL && a.gW2) ? a.gW2 + goff : nullptr;
    float* const gb2 = (FULL && a.gb2) ? a.gb2 + goff : nullptr;
    LLSend llv;
    const LLSend* const ll = sxll ? &llv : nullptr;
    const float* const gbase = sxll ? a.sx.grads[a.sx.rank] : nullptr;
    if (sxll) {
      llv.recv = a.sx.ll_recv; llv.world = a.sx.world; llv.epoch = sx_epoch0 + (unsigned)s + 1u;
      llv.slot = ((long)(llv.epoch & 1u) * a.sx.world + a.sx.rank) * (long)a.sx.grad_stride;
    }
    if (slice) {
#pragma unroll
    for (int i = 0; i < kXPer; ++i) {
      int e = tid + i * kThreads;
      if (e < B * KS) Xs[e] = xnext[i];
    }
    if (ahead || sx1) {
      // W1 slice: already in Ws (fetched during the previous step's phase C, own update added there; all-reduce
      // mode: resident for the whole launch)
      dtfe::cp_async_wait<0>();              // only the first step of a launch still has copies in flight here
    } else {
      // W1 slice: every thread re-read, right after its own pushes in the previous step's phase C (same
      // thread, same addresses: program order guarantees it sees them), exactly the elements it owns.
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (wv0) Ws[(wgrp * 4 + q) * HP + wj0] = wnext0[q];
        if (wv1) Ws[(wgrp * 4 + q) * HP + wj1] = wnext1[q];
      }
    }
    // label of the row this CTA handles in phase B (latency hidden behind phase A)
    if (g < B) ylab = __ldg(a.labels + ylab_row);   // row index arrived a step ago
    __syncthreads();
    {
      // W slice in registers once; X rows as broadcast float4 loads: KS/4 + 0 shared loads per 4*KS FMAs
      float w[KS][4];
#pragma unroll
      for (int kk = 0; kk < KS; ++kk)
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          int j = lane + 32 * jj;
          w[kk][jj] = (j < H) ? Ws[kk * HP + j] : 0.f;
        }
      float* P = a.part + (long)g * B * H;
      for (int r = warp; r < B; r += 8) {
        float x[KS];
#pragma unroll
        for (int q = 0; q < KS / 4; ++q) {
          float4 v = *reinterpret_cast<const float4*>(Xs + r * KS + 4 * q);
          x[4 * q] = v.x; x[4 * q + 1] = v.y; x[4 * q + 2] = v.z; x[4 * q + 3] = v.w;
        }
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          float acc = 0.f;
#pragma unroll
          for (int kk = 0; kk < KS; ++kk) acc = fmaf(x[kk], w[kk][jj], acc);
          int j = lane + 32 * jj;
          if (j < H) __stcg(P + (long)r * H + j, acc);
        }
      }
    }
    } else if (g < B) {
      ylab = __ldg(a.labels + ylab_row);
    }
    DBG_STAMP(1);
    grid_barrier(a.sync, bar_target, NC);
    DBG_STAMP(2);
    if ((FULL && a.poison) && s > 0) {                 // phase C of the previous step (the only reader of these rows) is over
      for (int r = g; r < B; r += (int)NC) {
        for (int i = tid; i < H; i += kThreads) {
          __stcg(a.Hbuf + (long)r * H + i, __int_as_float(0x7fc00000));
          __stcg(a.dA + (long)r * H + i, __int_as_float(0x7fc00000));
        }
        if (tid < C) __stcg(a.dZ + (long)r * C + tid, __int_as_float(0x7fc00000));
        if (tid == 0) __stcg(a.rowloss + r, __int_as_float(0x7fc00000));
      }
    }
    if (sx1) {
      // a CTA that gave up on its peers in the previous step said so before arriving here: stop, all CTAs alike
      unsigned v;
      asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a.sync + 2) : "memory");
      if (v == kSxAbort) break;
    }

    // ------------------------------ phase B ------------------------------------
    // Next step's X slice and the rows after that: issued now, consumed a whole step later.
    fetch_x(s + 1);
    fetch_rows(s + 2);
    fetch_label_row(s + 1);
    // Every CTA has passed barrier 1, so every push of step s-1 has been issued and fenced (GPU scope): count
    // that update now.  The atomic is relaxed - a release at system scope costs this CTA ~6 us when the
    // counter lives on a peer GPU, and it would be late at barrier 2 for everybody - so WITHIN a launch an
    // observer on another GPU may see the counter a few in-flight NVLink writes ahead of the parameters (the
    // reference's global_step and variable updates are separate, unordered ops as well); the exit handshake
    // puts a system fence before the launch's last increment.  The result (the global step the previous step
    // observed, what the reference fetches next to train_op) is consumed at the end of phase C.
    if (g == book && tid == 0 && s > 0 && a.train && !(FULL && a.has_sx))
      pending_step = (FULL && a.gW1) ? *reinterpret_cast<volatile int*>(a.global_step)
                           : dtfe::atom_add_relaxed_sys_s32(a.global_step, 1);
    // Small parameters.  Direct mode (evaluation, synchronous mode, first step of a launch): read from the
    // shard after the barrier, so every CTA sees the pushes all CTAs made in the previous step's phase C.
    // Ahead mode: ONE CTA (the bookkeeper) fetches them from the shard during phase B - between the two grid
    // barriers no CTA of this worker pushes - and republishes them in a local mirror before barrier 2; a step
    // later every row CTA adds this worker's own deltas of the step in between (local buffer, written in
    // phase C).  Row CTAs therefore only ever load local memory here, and the peer sees one reader, not 100.
    constexpr int kW2Per = (kMaxH * kMaxC + kThreads - 1) / kThreads;
    constexpr int kSmallVec = (kMaxH * kMaxC + kMaxH + kMaxC + 4 * kThreads - 1) / (4 * kThreads);   // quads per thread
    const bool from_mirror = ahead && s > 0;
    const bool fetches = ahead && fidx >= 0 && s + 1 < a.nsteps;
    if (fetches) {
#pragma unroll
      for (int k = 0; k < kFetchSlots; ++k) {
        const int slot = tid + k * kThreads, e = 4 * (fidx + slot * nfetch);
        if (e < dWords) {
          int n;
          const float* src = small_quad_src(e, n);
          dtfe::cp_async_cg16(stg + 4 * slot, src, n >= 4 ? 16 : 4 * n);   // tails zero-filled
        }
      }
      dtfe::cp_async_commit();
    }
    float w2r[kW2Per];
    float b1r = 0.f, b2r = 0.f;
    float4 mq[kSmallVec], dq[kSmallVec];
    if (has_row) {
      if (!from_mirror) {
#pragma unroll
        for (int i = 0; i < kW2Per; ++i) {
          int e = tid + i * kThreads;
          w2r[i] = (e < H * C) ? dtfe::ld_relaxed_sys(a.W2 + e) : 0.f;
        }
        b1r = (tid < H) ? dtfe::ld_relaxed_sys(a.b1 + tid) : 0.f;
        b2r = (tid < C) ? dtfe::ld_relaxed_sys(a.b2 + tid) : 0.f;
      } else {
        const float4* m4 = reinterpret_cast<const float4*>(a.mirror + (s & 1) * dWords);
        const float4* d4 = reinterpret_cast<const float4*>(delta);
#pragma unroll
        for (int i = 0; i < kSmallVec; ++i) {
          int f = tid + i * kThreads;
          if (4 * f < dWords) { mq[i] = __ldcg(m4 + f); dq[i] = __ldcg(d4 + f); }
        }
      }
    }
    for (int r = g; r < B; r += (int)NC) {
      const bool first_round = r == g;
      float z = 0.f;
      const bool vec_reduce = (H & 3) == 0;
      const int nq4 = vec_reduce ? (H >> 2) : 1;      // float4 columns per row (25)
      const int nqg = min(kThreads / nq4, 16);        // partial groups (10)
      // After the reduction, hidden unit tu belongs to lane `lane` of warp `warp` (per = ceil(H/8) <= 16 units
      // per warp): its activation stays in that thread's register through logits and dA.
      const int per = (H + 7) / 8;
      const int tu = warp * per + lane;
      const bool own = lane < per && tu < H;
      if (vec_reduce) {
        // Cooperative float4 reduction of the G split-K partials of row r: thread = (column quad, partial
        // group); every thread sums ~G/nqg float4 loads in registers, the nqg group sums meet in smem.
        const int c4 = t % nq4, qg = t / nq4;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4* P4 = reinterpret_cast<const float4*>(a.part + (long)r * H) + c4;
        const long gs4 = ((long)B * H) >> 2;
        const bool mine = qg < nqg;
        float4 v[10];                                           // up to 10 loads in flight per thread
        auto load_batch = [&](unsigned q0) {
#pragma unroll
          for (int u = 0; u < 10; ++u) {
            unsigned q = q0 + u * nqg;
            v[u] = (mine && q < G) ? __ldcg(P4 + (long)q * gs4) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
        };
        auto add_batch = [&]() {
#pragma unroll
          for (int u = 0; u < 10; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
        };
        load_batch(qg);
        add_batch();
        for (unsigned q0 = qg + 10 * nqg; q0 < G; q0 += 10 * nqg) { load_batch(q0); add_batch(); }
        if (mine) reinterpret_cast<float4*>(red_part + qg * kMaxH)[c4] = acc;
      } else if (own) {
        const float* P = a.part + (long)r * H + tu;
        const long gs = (long)B * H;
        float z0 = 0.f, z1 = 0.f;
        unsigned q = 0;
        for (; q + 2 <= G; q += 2) { z0 += __ldcg(P + (long)q * gs); z1 += __ldcg(P + (long)(q + 1) * gs); }
        if (q < G) z0 += __ldcg(P + (long)q * gs);
        z = z0 + z1;
      }
      DBG_STAMP(8);
      if (first_round) {
        if (!from_mirror) {
#pragma unroll
          for (int i = 0; i < kW2Per; ++i) {
            int e = tid + i * kThreads;
            if (e < H * C) W2s[e] = w2r[i];
          }
          if (tid < H) b1s[tid] = b1r;
          if (tid < C) b2s[tid] = b2r;
        } else {
          // mirrored value + own update = what the shard holds now (bitwise, absent other writers)
#pragma unroll
          for (int i = 0; i < kSmallVec; ++i) {
            int e = 4 * (tid + i * kThreads);
            if (e < dWords) {
              float4 o = mq[i];
              o.x = __fadd_rn(o.x, dq[i].x); o.y = __fadd_rn(o.y, dq[i].y);
              o.z = __fadd_rn(o.z, dq[i].z); o.w = __fadd_rn(o.w, dq[i].w);
              float* dst = e < dOffB1 ? W2s + e : e < dOffB2 ? b1s + (e - dOffB1) : b2s + (e - dOffB2);
              *reinterpret_cast<float4*>(dst) = o;
            }
          }
        }
      }
      __syncthreads();
      // Activation of the owned unit, then this warp's share of the logits straight out of registers:
      // lane = class, the warp's activations are broadcast with shuffles (no smem round trip, no barrier).
      float h = 0.f;
      if (own) {
        if (vec_reduce)
          for (int qg = 0; qg < nqg; ++qg) z += red_part[qg * kMaxH + tu];
        h = dtfe::sigmoidf_fast(z + b1s[tu]);
        __stcg(a.Hbuf + (long)r * H + tu, h);
      }
      {
        float p = 0.f;
        for (int u = 0; u < per; ++u) {
          const float hu = __shfl_sync(0xffffffffu, h, u);       // 0 for units beyond H
          const int uu = warp * per + u;
          if (lane < C && uu < H) p = fmaf(hu, W2s[uu * C + lane], p);
        }
        if (lane < C) zs[warp * kMaxC + lane] = p;
      }
      __syncthreads();
      DBG_STAMP(9);
      DBG_STAMP(10);
      // Softmax cross-entropy of the row, computed by EVERY warp (same latency as by one, and dz stays in the
      // lanes 0..C-1 of each warp for the dA product below); warp 0 publishes dZ / loss / correct.
      float dz = 0.f;
      {
        float zc = -INFINITY;
        if (lane < C) {
          float p = b2s[lane];
#pragma unroll
          for (int w = 0; w < 8; ++w) p += zs[w * kMaxC + lane];
          zc = p;
        }
        float mx = zc;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));   // classes live in lanes 0..15
        // first index wins on ties, like tf.argmax
        const int am = __ffs(__ballot_sync(0xffffffffu, lane < C && zc == mx)) - 1;
        float ex = (lane < C) ? __expf(zc - mx) : 0.f;
        float se = ex;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
        int y = ylab;
        if (!first_round) {                  // later rounds (B > grid): fetch the label now
          long row = a.batch_idx ? (long)a.batch_idx[(sbase + s) * B + r] : (sbase + s) * a.rows_per_step + r;
          y = a.labels[row];
        }
        float zy = __shfl_sync(0xffffffffu, zc, y);
        if (lane < C) dz = (ex / se - (lane == y ? 1.f : 0.f)) * invB;
        if (warp == 0) {
          if (lane < C) __stcg(a.dZ + (long)r * C + lane, dz);
          if (lane == 0) {
            __stcg(a.rowloss + r, mx + __logf(se) - zy);
            __stcg(a.rowcorrect + r, (am == y) ? 1 : 0);
          }
        }
      }
      DBG_STAMP(11);
      {
        float dh = 0.f;
        for (int c = 0; c < C; ++c) {
          const float dzc = __shfl_sync(0xffffffffu, dz, c);
          if (own) dh = fmaf(dzc, W2s[tu * C + c], dh);
        }
        if (own) __stcg(a.dA + (long)r * H + tu, dh * h * (1.f - h));
      }
      if (r + (int)NC < B) __syncthreads();  // another round reuses red_part / zs
    }
    if (fetches) {
      // The fetch for the next step has landed before anyone pushes (phase C starts behind barrier 2);
      // republish it locally for every row CTA.
      dtfe::cp_async_wait<0>();
      float4* m4 = reinterpret_cast<float4*>(a.mirror + ((s + 1) & 1) * dWords);
#pragma unroll
      for (int k = 0; k < kFetchSlots; ++k) {       // every thread republishes the quads it fetched itself
        const int slot = tid + k * kThreads, f = fidx + slot * nfetch;
        if (4 * f < dWords) __stcg(m4 + f, reinterpret_cast<const float4*>(stg)[slot]);
      }
    }
    DBG_STAMP(3);
    grid_barrier(a.sync, bar_target, NC);
    DBG_STAMP(4);
    if ((FULL && a.poison) && slice) {                 // every reader of this CTA's split-K partial is behind the barrier
      float* P = a.part + (long)g * B * H;
      for (int i = tid; i < B * H; i += kThreads) __stcg(P + i, __int_as_float(0x7fc00000));
    }

    // ------------------------------ phase C ------------------------------------
    if (g == book && warp == 7) {
      // metrics of this step, fixed summation order -> bitwise reproducible
      float l = 0.f; int cnt = 0;
      for (int r = lane; r < B; r += 32) { l += __ldcg(a.rowloss + r); cnt += __ldcg(a.rowcorrect + r); }
      l = dtfe::warp_sum(l);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
      if (lane == 0) {
        a.loss_out[s] = l * invB;
        a.acc_out[s] = (float)cnt * invB;
      }
    }
    if (a.train) {
      const float nlr = -a.lr;
      // Everything this phase reads is requested up front with cp.async, in the order it is needed:
      //   groups 1..4  dA in four row chunks      -> dW1 starts on chunk 0 while the others are in flight
      //   group 5      dZ                         -> small gradients
      //   group 6      (ahead mode) the W1 slice for the next step: fetched BEFORE this CTA pushes to it (only
      //                this CTA of this worker does), its own update is added in shared memory below.
      constexpr int kChunks = 4;
      const int rb = (((B + kChunks - 1) / kChunks) + 7) & ~7;     // rows per chunk: 16-byte starts, whole rounds of 8 warps
      // Who computes the small gradients: the helper CTAs (the ones without a W1 slice) when there are enough
      // of them, in parallel with the slice CTAs' dW1; otherwise every CTA, after its dW1.
      const int nhelp = (int)NC - (int)G;
      const bool on_helpers = nhelp * 4 >= H;
      const bool does_small = on_helpers ? !slice : true;
      const int t0 = on_helpers ? g - (int)G : g, tstride = on_helpers ? nhelp : (int)NC;
      const int t1 = t0 + tstride;
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        const int b0 = min(B, c * rb), b1 = min(B, b0 + rb);
        if (slice) dtfe::cta_copy_async(dAs + b0 * HP, a.dA + (long)b0 * H, (b1 - b0) * H, tid, kThreads);
        dtfe::cp_async_commit();
      }
      if (does_small) dtfe::cta_copy_async(dZs, a.dZ, B * C, tid, kThreads);
      dtfe::cp_async_commit();
      // H and dA columns of this CTA's hidden units (local loads: issued before the fetch from the peer)
      // (column loads touch one 32-byte sector per row: one warp fetches a column for the whole CTA)
      float hreg[4] = {0.f, 0.f, 0.f, 0.f}, dacol0[4] = {0.f, 0.f, 0.f, 0.f}, dacol1[4] = {0.f, 0.f, 0.f, 0.f};
      if (does_small) {
        const int tw = warp == 0 ? t0 : t1;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          int b = lane + 32 * i;
          if (warp < 2 && b < B && tw < H) hreg[i] = __ldcg(a.Hbuf + (long)b * H + tw);
          if (warp == 5 && b < B && t0 < H) dacol0[i] = __ldcg(a.dA + (long)b * H + t0);
          if (warp == 5 && b < B && t1 < H) dacol1[i] = __ldcg(a.dA + (long)b * H + t1);
        }
      }
      if (ahead && s + 1 < a.nsteps) dtfe::cta_copy_async(Ws, a.W1 + (long)k0 * H, nW1, tid, kThreads);
      dtfe::cp_async_commit();
      // dW1 = Xslice^T . dA for rows [k0, k0+KS).  Warp w takes every 8th batch row: two broadcast float4 of X
      // and one dA row (4 columns per lane) feed 4*KS FMAs, like phase A.  The eight per-warp partial tiles
      // meet in shared memory, where every thread sums - in fixed order - the elements it owns and pushes.
      if (slice) {
        float acc[KS][4];
#pragma unroll
        for (int kk = 0; kk < KS; ++kk)
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) acc[kk][jj] = 0.f;
        for (int c = 0; c < kChunks; ++c) {
          switch (c) {                            // chunk c has landed: all but the (kChunks + 1 - c) newer groups
            case 0: dtfe::cp_async_wait<kChunks + 1>(); break;
            case 1: dtfe::cp_async_wait<kChunks>(); break;
            case 2: dtfe::cp_async_wait<kChunks - 1>(); break;
            default: dtfe::cp_async_wait<kChunks - 2>(); break;
          }
          __syncthreads();
          if (c == 0) DBG_STAMP(12);
          const int b0 = min(B, c * rb), bend = min(B, b0 + rb);
          for (int b = b0 + warp; b < bend; b += 8) {
            float x[KS], da[4];
#pragma unroll
            for (int q = 0; q < KS / 4; ++q) {
              float4 v = *reinterpret_cast<const float4*>(Xs + b * KS + 4 * q);
              x[4 * q] = v.x; x[4 * q + 1] = v.y; x[4 * q + 2] = v.z; x[4 * q + 3] = v.w;
            }
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
              int j = lane + 32 * jj;
              da[jj] = (j < H) ? dAs[b * HP + j] : 0.f;
            }
#pragma unroll
            for (int kk = 0; kk < KS; ++kk)
#pragma unroll
              for (int jj = 0; jj < 4; ++jj) acc[kk][jj] = fmaf(x[kk], da[jj], acc[kk][jj]);
          }
        }
        float* pt = dwp + warp * (KS * kMaxH);
#pragma unroll
        for (int kk = 0; kk < KS; ++kk)
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) pt[kk * kMaxH + lane + 32 * jj] = acc[kk][jj];
      }
      DBG_STAMP(13);
      if (does_small) {
        if (warp < 2) {
#pragma unroll
          for (int i = 0; i < 4; ++i) red[warp * kMaxB + lane + 32 * i] = hreg[i];   // H columns t0, t1
        }
        dtfe::cp_async_wait<1>();                 // dZ has landed (this thread's quads) ...
        __syncthreads();                          // ... and everybody else's
        if (!slice) DBG_STAMP(12);
        // small gradients: every warp takes one or two classes of the dW2 row; warp 5 also sums the dA column
        for (int tt = t0, rnd = 0; tt < H; tt += tstride, ++rnd) {
          float hc[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            int b = lane + 32 * i;
            hc[i] = rnd < 2 ? red[rnd * kMaxB + b] : (b < B ? __ldcg(a.Hbuf + (long)b * H + tt) : 0.f);
          }
          for (int c = warp; c < C; c += 8) {     // dW2[tt][c] = sum_b H[b][tt] * dZ[b][c]
            float v = 0.f;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              int b = lane + 32 * i;
              if (b < B) v = fmaf(hc[i], dZs[b * C + c], v);
            }
            v = dtfe::warp_sum(v);
            if (lane == 0) emit_small(a.W2, gW2, delta, (long)tt * C + c, v, nlr, a.atomic, ll, gbase);
          }
          if (warp == 5) {                        // db1[tt] = sum_b dA[b][tt]
            float v = 0.f;
            if (rnd == 0) v = (dacol0[0] + dacol0[1]) + (dacol0[2] + dacol0[3]);
            else if (rnd == 1) v = (dacol1[0] + dacol1[1]) + (dacol1[2] + dacol1[3]);
            else for (int b = lane; b < B; b += 32) v += __ldcg(a.dA + (long)b * H + tt);
            v = dtfe::warp_sum(v);
            if (lane == 0) emit_small(a.b1, gb1, delta ? delta + dOffB1 : nullptr, tt, v, nlr, a.atomic, ll, gbase);
          }
        }
        if (g == book) {                          // db2[c] = sum_b dZ[b][c], spread over the warps
          for (int c = warp; c < C; c += 8) {
            float v = 0.f;
            for (int b = lane; b < B; b += 32) v += dZs[b * C + c];
            v = dtfe::warp_sum(v);
            if (lane == 0) emit_small(a.b2, gb2, delta ? delta + dOffB2 : nullptr, c, v, nlr, a.atomic, ll, gbase);
          }
        }
      }
      DBG_STAMP(14);
      dtfe::cp_async_wait<0>();                   // (ahead) next step's W1 slice is in Ws before this CTA pushes to it
      __syncthreads();                            // ... and all eight partial dW1 tiles are in shared memory
      if (slice) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          int k = k0 + wgrp * 4 + q;
          if (k < D) {
            float g0 = 0.f, g1 = 0.f;
            const float* pq = dwp + (wgrp * 4 + q) * kMaxH;
#pragma unroll
            for (int w = 0; w < 8; ++w) {
              if (wv0) g0 += pq[w * (KS * kMaxH) + wj0];
              if (kParts < 4 && wv1) g1 += pq[w * (KS * kMaxH) + wj1];
            }
            if (wv0) emit(a.W1, gW1, (long)k * H + wj0, g0, nlr, a.atomic, ll, gbase);
            if (kParts < 4 && wv1) emit(a.W1, gW1, (long)k * H + wj1, g1, nlr, a.atomic, ll, gbase);
            if (ahead) {        // fetched value + own update = what the shard holds (bitwise, absent other writers)
              float* w0 = Ws + (wgrp * 4 + q) * HP + wj0;
              if (wv0) *w0 = __fadd_rn(*w0, __fmul_rn(nlr, g0));
              if (kParts < 4 && wv1) w0[32] = __fadd_rn(w0[32], __fmul_rn(nlr, g1));
            }
          }
        }
        // Direct mode: the same thread reads back exactly the elements it just pushed (synchronous mode:
        // the parameters only change in phase D, the read-back happens there)
        if (!ahead && !(FULL && a.has_sx)) fetch_w1();
      }
      DBG_STAMP(5);
      if (g == book && tid == 0 && s > 0 && !(FULL && a.has_sx)) a.step_out[s - 1] = pending_step;
      // No per-step completion handshake: the next grid barrier (or the exit handshake below) orders all
      // of this step's pushes before the global_step increment that counts them.
      __syncthreads();                            // dAs / Xs are rewritten by the next step
      DBG_STAMP(7);
      if (sx1) {
        // ------------------------------ phase D, all-reduce form -----------------------------------
        // This CTA's gradients are in this rank's exchange buffer (the __syncthreads above ordered every
        // thread's stores before thread 0's fence).  Tell the same CTA of every other rank, wait for theirs,
        // sum exactly the elements this CTA produced over all ranks in rank order, apply to the home copy.
        const unsigned epoch = sx_epoch0 + (unsigned)s + 1u;
        const int W = a.sx.world, me = a.sx.rank;
        bool ok = true;
        if (!sxll) {
          if (tid == 0 && W > 1) {
            __threadfence_system();
            for (int p = 0; p < W; ++p)
              if (p != me) st_relaxed_sys_u32(a.sx.cta_flags[p] + me * kCtaFlagPitch + g, epoch);
          }
          if (tid < W && tid != me) ok = sx_wait(a.sx.cta_flags[me] + tid * kCtaFlagPitch + g, epoch, a.sx, sx_t0);
          ok = __syncthreads_and(ok) != 0;
        }
        DBG_STAMP(6);
        // Source of rank q's copy of flat-gradient element i.  Modes: peer loads from q's exchange buffer (1), or
        // this rank's own receive buffer, where every rank stored {value, step number} pairs (2).
        const float* const llr = sxll ? a.sx.ll_recv[me] + 2 * ((long)(epoch & 1u) * W * (long)a.sx.grad_stride) : nullptr;
        const float etag = __uint_as_float(epoch);
        auto ll_pair = [&](int q, long i, bool& good) -> float {     // spins until rank q's element i of this step is here
          const float* p = llr + 2 * ((long)q * (long)a.sx.grad_stride + i);
          float v, t;
          unsigned it = 0;
          for (;;) {
            asm volatile("ld.relaxed.sys.global.v2.f32 {%0, %1}, [%2];" : "=f"(v), "=f"(t) : "l"(p) : "memory");
            if (__float_as_uint(t) == __float_as_uint(etag)) return v;
            if ((++it & 0x3ff) == 0) {
              if ((a.sx.abort_flag && dtfe::ld_acquire_sys(a.sx.abort_flag) != 0) ||
                  (a.sx.timeout_ns && (unsigned long long)(gtime() - sx_t0) > a.sx.timeout_ns)) { good = false; return 0.f; }
            }
          }
        };
        if (ok) {
          const float alpha = a.sx.alpha;
          if (slice) {
            const long base = sxOffW1 + goff + (long)k0 * H;
            if (sxll) {
              // 16-byte loads carry two pairs; every pair is validated on its own (8-byte stores are atomic)
              for (int f = tid; 4 * f < nW1; f += kThreads) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                const int nel = min(4, nW1 - 4 * f);
                for (int q = 0; q < W; ++q) {
                  const float* p = llr + 2 * ((long)q * (long)a.sx.grad_stride + base + 4 * f);
                  float4 lo = make_float4(0.f, etag, 0.f, etag), hi = lo;
                  const bool vec = nel == 4 && ((base & 1) == 0);
                  if (vec) {
                    unsigned it = 0;
                    for (;;) {
                      lo = dtfe::ld_relaxed_sys_v4(p); hi = dtfe::ld_relaxed_sys_v4(p + 4);
                      if (__float_as_uint(lo.y) == epoch && __float_as_uint(lo.w) == epoch &&
                          __float_as_uint(hi.y) == epoch && __float_as_uint(hi.w) == epoch) break;
                      if ((++it & 0x3ff) == 0) {
                        if ((a.sx.abort_flag && dtfe::ld_acquire_sys(a.sx.abort_flag) != 0) ||
                            (a.sx.timeout_ns && (unsigned long long)(gtime() - sx_t0) > a.sx.timeout_ns)) { ok = false; break; }
                      }
                    }
                  } else {
                    lo.x = ll_pair(q, base + 4 * f, ok);
                    if (nel > 1) lo.z = ll_pair(q, base + 4 * f + 1, ok);
                    if (nel > 2) hi.x = ll_pair(q, base + 4 * f + 2, ok);
                    if (nel > 3) hi.z = ll_pair(q, base + 4 * f + 3, ok);
                  }
                  if (!ok) break;
                  acc.x += lo.x; acc.y += lo.z; acc.z += hi.x; acc.w += hi.z;
                }
                if (!ok) break;
                float w4[4] = {acc.x, acc.y, acc.z, acc.w};
                for (int j = 0; j < nel; ++j) {
                  const float w = fmaf(alpha, w4[j], Ws[4 * f + j]);
                  Ws[4 * f + j] = w;
                  __stcg(a.W1 + (long)k0 * H + 4 * f + j, w);
                }
              }
            } else if ((H & 3) == 0 && (base & 3) == 0) {
              for (int f = tid; 4 * f < nW1; f += kThreads) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int q0 = 0; q0 < W; q0 += 8) {
                  float4 v[8];
#pragma unroll
                  for (int u = 0; u < 8; ++u)
                    if (q0 + u < W) v[u] = dtfe::ld_relaxed_sys_v4(a.sx.grads[q0 + u] + base + 4 * f);
#pragma unroll
                  for (int u = 0; u < 8; ++u)
                    if (q0 + u < W) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
                }
                float4 w = *reinterpret_cast<float4*>(Ws + 4 * f);
                w.x = fmaf(alpha, acc.x, w.x); w.y = fmaf(alpha, acc.y, w.y);
                w.z = fmaf(alpha, acc.z, w.z); w.w = fmaf(alpha, acc.w, w.w);
                *reinterpret_cast<float4*>(Ws + 4 * f) = w;
                __stcg(reinterpret_cast<float4*>(a.W1 + (long)k0 * H) + f, w);
              }
            } else {
              for (int e = tid; e < nW1; e += kThreads) {
                float acc = 0.f;
                for (int q = 0; q < W; ++q) acc += dtfe::ld_relaxed_sys(a.sx.grads[q] + base + e);
                const float w = fmaf(alpha, acc, Ws[e]);
                Ws[e] = w;
                __stcg(a.W1 + (long)k0 * H + e, w);
              }
            }
          }
          if (does_small) {
            const int nrows = t0 < H ? (H - t0 + tstride - 1) / tstride : 0, per = C + 1;
            for (int i = tid; i < nrows * per; i += kThreads) {
              const int tt = t0 + (i / per) * tstride, c = i % per;
              const long off = goff + (c < C ? sxOffW2 + (long)tt * C + c : sxOffB1 + tt);
              float acc = 0.f;
              for (int q = 0; q < W && ok; ++q) acc += sxll ? ll_pair(q, off, ok) : dtfe::ld_relaxed_sys(a.sx.grads[q] + off);
              float* home = c < C ? a.W2 + (long)tt * C + c : a.b1 + tt;
              if (ok) __stcg(home, fmaf(alpha, acc, __ldcg(home)));
            }
            if (g == book)
              for (int c = tid; c < C; c += kThreads) {
                float acc = 0.f;
                for (int q = 0; q < W && ok; ++q)
                  acc += sxll ? ll_pair(q, goff + sxOffB2 + c, ok) : dtfe::ld_relaxed_sys(a.sx.grads[q] + goff + sxOffB2 + c);
                if (ok) __stcg(a.b2 + c, fmaf(alpha, acc, __ldcg(a.b2 + c)));
              }
          }
          if (sxll) ok = __syncthreads_and(ok) != 0;
        }
        if (ok) {
          if (g == book && tid == 0) {        // every rank counts the aggregated update in its own copy
            const int old = *reinterpret_cast<volatile int*>(a.global_step);
            *reinterpret_cast<volatile int*>(a.global_step) = old + 1;
            a.step_out[s] = old;
          }
          ++steps_done;
        } else if (tid == 0) {
          a.sx.status[0] = 1u;
          asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(a.sync + 2), "r"(kSxAbort) : "memory");
        }
        __syncthreads();                          // the updated W1 slice is in Ws for every thread
      } else if ((FULL && a.has_sx)) {
        // ------------------------------ phase D (synchronous mode) -------------------------------
        // D1: every CTA's gradients are in this rank's exchange buffer
        grid_barrier(a.sync, bar_target, NC);
        const unsigned epoch = sx_epoch0 + (unsigned)s + 1u;
        const unsigned kAbort = 0xffffffffu;
        // D2: tell every peer, wait for every peer (CTA 0), then release the grid
        if (g == 0 && tid == 0) {
          __threadfence_system();
          for (int p = 0; p < a.sx.world; ++p) dtfe::st_release_sys(a.sx.flags[p] + a.sx.rank, epoch);
          bool ok = true;
          for (int q = 0; q < a.sx.world && ok; ++q) ok = sx_wait(a.sx.flags[a.sx.rank] + q, epoch, a.sx, sx_t0);
          asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(a.sync + 2), "r"(ok ? epoch : kAbort) : "memory");
        }
        if (tid == 0) {
          unsigned v;
          do { asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a.sync + 2) : "memory"); } while (v < epoch);
          __threadfence();
          s_go = v;
        }
        __syncthreads();
        const bool aborted = s_go == kAbort;
        // D3: reduce the slices of the variables this rank owns out of all ranks' buffers and apply in place
        if (!aborted) {
          const size_t gtid = (size_t)g * kThreads + tid, gstride = (size_t)NC * kThreads;
          for (int sg = 0; sg < a.sx.nseg; ++sg) {
            const size_t begin = a.sx.seg_begin[sg], n = a.sx.seg_end[sg] - begin;
            float* param = a.sx.seg_param[sg];
            const bool vec_ok = (begin % 4 == 0) && ((reinterpret_cast<uintptr_t>(param) & 15) == 0);
            const size_t nv = vec_ok ? n / 4 : 0;
            for (size_t v = gtid; v < nv; v += gstride) {
              float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
              for (int q = 0; q < a.sx.world; ++q) {
                float4 gq = dtfe::ld_relaxed_sys_v4(a.sx.grads[q] + begin + v * 4);
                acc.x += gq.x; acc.y += gq.y; acc.z += gq.z; acc.w += gq.w;
              }
              float4 w = dtfe::ld_relaxed_sys_v4(param + v * 4);
              w.x += a.sx.alpha * acc.x; w.y += a.sx.alpha * acc.y; w.z += a.sx.alpha * acc.z; w.w += a.sx.alpha * acc.w;
              *reinterpret_cast<float4*>(param + v * 4) = w;
            }
            for (size_t e = nv * 4 + gtid; e < n; e += gstride) {
              float acc = 0.f;
              for (int q = 0; q < a.sx.world; ++q) acc += dtfe::ld_relaxed_sys(a.sx.grads[q] + begin + e);
              param[e] = dtfe::ld_relaxed_sys(param + e) + a.sx.alpha * acc;
            }
          }
        }
        // D4: all of this rank's applies are done; wait until every rank's are (nobody may read a stale
        // parameter or overwrite a gradient buffer that a peer is still reducing)
        grid_barrier(a.sync, bar_target, NC);
        if (g == 0 && tid == 0) {
          bool ok = !aborted;
          if (ok) {
            __threadfence_system();
            if (a.sx.global_step) dtfe::atom_add_sys_s32(a.sx.global_step, 1);
            for (int p = 0; p < a.sx.world; ++p) dtfe::st_release_sys(a.sx.flags[p] + 16 + a.sx.rank, epoch);
            for (int q = 0; q < a.sx.world && ok; ++q) ok = sx_wait(a.sx.flags[a.sx.rank] + 16 + q, epoch, a.sx, sx_t0);
            if (ok) a.step_out[s] = (int)dtfe::ld_acquire_sys(reinterpret_cast<const unsigned*>(a.global_step)) - 1;
          }
          if (!ok) a.sx.status[0] = aborted ? 1u : 2u;
          asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(a.sync + 3), "r"(ok ? epoch : kAbort) : "memory");
        }
        if (tid == 0) {
          unsigned v;
          do { asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a.sync + 3) : "memory"); } while (v < epoch);
          __threadfence();
          s_go = v;
        }
        __syncthreads();
        if (s_go == kAbort) break;                // a rank left: stop after this step, on every CTA alike
        ++steps_done;
        fetch_w1();                               // the pull of the next step sees every owner's update
        __syncthreads();
      }
    } else {
      fetch_w1();
      if (g == 0 && tid == 0) a.step_out[s] = *reinterpret_cast<volatile int*>(a.global_step);
      __syncthreads();
    }
  }

  // Exit handshake: the last CTA to get here leaves the sync words zeroed for the next launch and counts
  // the final step (every CTA fenced its pushes before arriving).
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    unsigned prev = atomicAdd(a.sync + 1, 1u);
    if (prev == NC - 1) {
      a.sync[0] = 0; a.sync[1] = 0;
      if ((FULL && a.has_sx) && a.sx.mode >= 1) a.sync[2] = 0;
      if ((FULL && a.has_sx)) {
        *a.sx.epoch = sx_epoch0 + (unsigned)steps_done;
      } else if (a.train && a.nsteps > 0) {
        if (a.remote) __threadfence_system();
        int old = (FULL && a.gW1) ? *reinterpret_cast<volatile int*>(a.global_step) : dtfe::atom_add_sys_s32(a.global_step, 1);
        a.step_out[a.nsteps - 1] = old;
      }
      if (a.cursor) *a.cursor = (int)sbase + a.nsteps;   // every CTA read the cursor before arriving here
      __threadfence();
    }
  }
}

template <int KS>
constexpr size_t smem_floats() {
  return (size_t)kMaxB * KS + (size_t)KS * kMaxH + kMaxH * kMaxC + kMaxH + kMaxC + 2 * kMaxH + 8 * kMaxC +
         2 * kMaxC + kThreads + kMaxB * kMaxC + (size_t)kMaxB * kMaxH + (size_t)kMaxH * kMaxC + kMaxH + kMaxC +
         8 * (size_t)KS * kMaxH;
}

}  // namespace

// Number of W1 slices (= CTAs that own one; split-K factor of the first layer).
DTFE_API int dtfe_mlp2_step_grid(int D, int ks) { return (D + ks - 1) / ks; }

// Element counts (4-byte words) of the scratch regions, in the order the workspace is carved:
// part[G*B*H] | dA[B*H] | Hbuf[B*H] | dZ[B*C] | rowloss[B] | rowcorrect[B] | delta | mirror[2] | sync[4]
DTFE_API size_t dtfe_mlp2_step_workspace_words(int B, int D, int H, int C, int ks) {
  size_t G = (size_t)((D + ks - 1) / ks);
  if (G < 8) G = 8;          // the tensor-core form (dtfe_mlp2_step_tc) always keeps 8 split-K partials
  size_t w = G * B * H + 2 * (size_t)B * H + (size_t)B * C + 2 * (size_t)B + 3 * ((size_t)H * C + H + C + 12) + 4;
  return (w + 3) / 4 * 4 + 64;
}

using Mlp2Launch = dtfe_mlp2_launch;

static int mlp2_fill(const Mlp2Launch& L, Mlp2Args& a) {
  DTFE_REQUIRE(L.B >= 1 && L.B <= kMaxB, "mlp2_step: batch %d outside [1,%d]", L.B, kMaxB);
  DTFE_REQUIRE(L.H >= 1 && L.H <= kMaxH, "mlp2_step: hidden %d outside [1,%d]", L.H, kMaxH);
  DTFE_REQUIRE(L.C >= 1 && L.C <= kMaxC, "mlp2_step: classes %d outside [1,%d]", L.C, kMaxC);
  DTFE_REQUIRE(L.ks == 8 || L.ks == 16, "mlp2_step: ks must be 8 or 16");
  DTFE_REQUIRE(L.nsteps >= 1, "mlp2_step: nsteps must be at least 1 (got %d)", L.nsteps);
  DTFE_REQUIRE(((uintptr_t)L.workspace % 16) == 0, "mlp2_step: workspace must be 16-byte aligned");
  size_t G = (size_t)((L.D + L.ks - 1) / L.ks);
  DTFE_REQUIRE(G <= 148, "mlp2_step: D/ks = %zu CTAs exceeds one wave (148); use ks=16 or the layer path", G);
  a.B = L.B; a.D = L.D; a.H = L.H; a.C = L.C;
  a.W1 = L.W1; a.b1 = L.b1; a.W2 = L.W2; a.b2 = L.b2; a.global_step = L.global_step;
  a.X = L.X; a.ldx = L.ldx; a.labels = L.labels; a.batch_idx = L.batch_idx; a.rows_per_step = L.rows_per_step;
  float* w = L.workspace;
  a.part = w;               w += G * L.B * L.H;
  w = (float*)(((uintptr_t)w + 15) & ~(uintptr_t)15);
  a.dA = w;                 w += (size_t)L.B * L.H;
  w = (float*)(((uintptr_t)w + 15) & ~(uintptr_t)15);
  a.Hbuf = w;               w += (size_t)L.B * L.H;
  w = (float*)(((uintptr_t)w + 15) & ~(uintptr_t)15);
  a.dZ = w;                 w += (size_t)L.B * L.C;
  a.rowloss = w;            w += L.B;
  a.rowcorrect = (int*)w;   w += L.B;
  w = (float*)(((uintptr_t)w + 15) & ~(uintptr_t)15);
  const size_t small_words = (((size_t)L.H * L.C + 3) & ~(size_t)3) + (((size_t)L.H + 3) & ~(size_t)3) +
                             (((size_t)L.C + 3) & ~(size_t)3);   // W2 | b1 | b2, each part on a float4
  a.delta = w;              w += small_words;
  a.mirror = w;             w += 2 * small_words;
  a.sync = (unsigned*)w;
  a.nslices = (int)G;
  a.loss_out = L.loss_out; a.acc_out = L.acc_out; a.step_out = L.step_out;
  a.lr = L.lr; a.nsteps = L.nsteps; a.atomic = L.atomic; a.train = L.train; a.remote = L.remote; a.dbg = L.dbg;
  a.gW1 = L.gW1; a.gb1 = L.gb1; a.gW2 = L.gW2; a.gb2 = L.gb2; a.cursor = L.cursor;
  a.has_sx = 0;
  if (L.sync_exchange) {
    a.sx = *L.sync_exchange;
    a.has_sx = 1;
    DTFE_REQUIRE(a.gW1 != nullptr, "mlp2_step: sync_exchange needs the four gradient buffers (they are what is exchanged)");
    DTFE_REQUIRE(a.sx.world >= 1 && a.sx.world <= 16 && a.sx.nseg >= 0 && a.sx.nseg <= 8 && a.sx.epoch && a.sx.status,
                 "mlp2_step: malformed sync_exchange");
    DTFE_REQUIRE(a.sx.mode >= 0 && a.sx.mode <= 2, "mlp2_step: sync_exchange.mode must be 0, 1 or 2");
    if (a.sx.mode >= 1) {
      DTFE_REQUIRE(a.sx.rank >= 0 && a.sx.rank < a.sx.world && a.sx.grad_stride > 0,
                   "mlp2_step: all-reduce exchange needs rank < world and grad_stride");
      for (int r = 0; r < a.sx.world; ++r)
        DTFE_REQUIRE(a.sx.grads[r] && (a.sx.mode == 1 ? (void*)a.sx.cta_flags[r] : (void*)a.sx.ll_recv[r]),
                     "mlp2_step: all-reduce exchange: rank %d's buffers are missing", r);
      DTFE_REQUIRE(((uintptr_t)L.W1 % 16) == 0, "mlp2_step: all-reduce exchange needs a 16-byte aligned W1 copy");
      DTFE_REQUIRE(a.train, "mlp2_step: sync_exchange is for training launches");
    }
  }
  DTFE_REQUIRE((a.gW1 != nullptr) == (a.gb1 != nullptr) && (a.gW1 != nullptr) == (a.gW2 != nullptr) &&
               (a.gW1 != nullptr) == (a.gb2 != nullptr), "mlp2_step: pass all four gradient buffers or none");
  // Pull-ahead copies in 16-byte units (cp.async) and only applies where this kernel itself pushes the updates.
  static const bool ahead_off = getenv("DTFE_MLP2_NO_AHEAD") != nullptr;
  const bool aligned16 = (((uintptr_t)L.W1 | (uintptr_t)L.W2 | (uintptr_t)L.b1 | (uintptr_t)L.b2) % 16) == 0;
  a.ahead = (!ahead_off && a.train && !a.gW1 && aligned16) ? 1 : 0;
  const char* poison = getenv("DTFE_MLP2_POISON");     // read per call: tests switch it on and off
  a.poison = (poison && poison[0] == '1') ? 1 : 0;
  return 0;
}

// CTAs the cooperative grid can hold on the current device (1 per SM with this kernel's registers/smem).
template <int KS>
static int mlp2_max_ctas(int* out) {
  static int cached[64] = {};
  int dev;
  DTFE_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev >= 64 || cached[dev] == 0) {
    size_t smem = smem_floats<KS>() * sizeof(float);
    DTFE_CUDA_CHECK(cudaFuncSetAttribute(k_mlp2_step<KS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DTFE_CUDA_CHECK(cudaFuncSetAttribute(k_mlp2_step<KS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0, sms = 0;
    DTFE_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_mlp2_step<KS, true>, kThreads, smem));
    DTFE_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    int n = per_sm * sms;
    if (dev < 64) cached[dev] = n;
    *out = n;
    return 0;
  }
  *out = cached[dev];
  return 0;
}

// Grid width: the W1 slices, widened to one CTA per batch row plus one that keeps the books when the
// device can hold them all at once (cooperative launch).
template <int KS>
static int mlp2_grid_width(int B, int D, int H, int* out) {
  int G = (D + KS - 1) / KS, cap = 0;
  if (int rc = mlp2_max_ctas<KS>(&cap)) return rc;
  DTFE_REQUIRE(G <= cap, "mlp2_step: %d W1 slices need %d co-resident CTAs, this device holds %d", G, G, cap);
  // B+1: one CTA per batch row plus the bookkeeper; G + H/2: helper CTAs that take the small gradients
  // (two hidden units each) off the slice CTAs.
  int want = B + 1 > G + (H + 1) / 2 ? B + 1 : G + (H + 1) / 2;
  if (want > cap) want = cap;
  *out = G > want ? G : want;
  return 0;
}

template <int KS>
static int mlp2_launch(Mlp2Args& a, cudaStream_t stream) {
  size_t smem = smem_floats<KS>() * sizeof(float);
  int NC = 0;
  if (int rc = mlp2_grid_width<KS>(a.B, a.D, a.H, &NC)) return rc;   // also sets the kernel's smem attribute once
  void* args[] = {&a};
  const bool full = a.has_sx || a.poison || a.gW1 != nullptr;
  DTFE_CUDA_CHECK(cudaLaunchCooperativeKernel(full ? (void*)k_mlp2_step<KS, true> : (void*)k_mlp2_step<KS, false>, dim3(NC),
                                              dim3(kThreads), args, smem, stream));
  return 0;
}

// CTAs one launch uses on the current device (the first dimension of the optional `dbg` stamp array).
DTFE_API int dtfe_mlp2_step_ctas(int B, int D, int H, int ks) {
  int NC = 0;
  int rc = ks == 8 ? mlp2_grid_width<8>(B, D, H, &NC) : mlp2_grid_width<16>(B, D, H, &NC);
  return rc ? -1 : NC;
}

// Run `nsteps` full training (train=1) or metric-only (train=0) steps of the 2-layer MLP.
DTFE_API int dtfe_mlp2_step(const Mlp2Launch* L, void* stream) {
  Mlp2Args a{};
  if (int rc = mlp2_fill(*L, a)) return rc;
  return L->ks == 8 ? mlp2_launch<8>(a, (cudaStream_t)stream) : mlp2_launch<16>(a, (cudaStream_t)stream);
}

// ---- variant for callers that must not allocate ------------------------------------------------
// Same kernel, same arguments, but the scratch workspace is a fixed array that is part of this
// library's device image instead of memory the caller provides.  Meant for callers that cannot
// obtain device memory at the point where they train: a worker confined to a fixed slice of a
// shared GPU (the reference packs several workers onto one GPU by memory fraction, main.py:57-64),
// a host application whose own pool already owns all of HBM, or a step issued from a place where
// cudaMalloc's device-wide synchronisation is not acceptable.
//
// Cost: every GPU that loads libdtfe_cuda.so holds these 4 MiB until the process exits.
constexpr size_t kNoAllocScratchWords = size_t(1) << 20;  // 4-byte words -> 4 MiB
static __device__ __align__(256) float g_mlp2_noalloc_scratch[kNoAllocScratchWords];

DTFE_API size_t dtfe_mlp2_step_noalloc_max_words(void) { return kNoAllocScratchWords; }

// `launch->workspace` is ignored.  See include/dtfe_b200.h for the contract.
DTFE_API int dtfe_mlp2_step_noalloc(const dtfe_mlp2_launch* launch, void* stream) {
  // In host code the array's name is only a handle for the runtime; the address a kernel can use
  // has to be looked up, per device, with that device current.
  static float* scratch_on_device[64] = {};
  int dev = -1;
  DTFE_CUDA_CHECK(cudaGetDevice(&dev));
  DTFE_REQUIRE(dev >= 0 && dev < 64, "dtfe_mlp2_step_noalloc: device ordinal %d not supported", dev);
  DTFE_REQUIRE(launch->ks == 8 || launch->ks == 16, "dtfe_mlp2_step_noalloc: ks must be 8 or 16");
  DTFE_REQUIRE(launch->B >= 1 && launch->D >= 1 && launch->H >= 1 && launch->C >= 1,
               "dtfe_mlp2_step_noalloc: non-positive dimension");
  size_t need = dtfe_mlp2_step_workspace_words(launch->B, launch->D, launch->H, launch->C, launch->ks);
  DTFE_REQUIRE(need <= kNoAllocScratchWords,
               "dtfe_mlp2_step_noalloc: B=%d D=%d H=%d C=%d ks=%d needs %zu words of scratch but the built-in "
               "array holds %zu (4 MiB); use dtfe_mlp2_step with a caller-provided workspace",
               launch->B, launch->D, launch->H, launch->C, launch->ks, need, (size_t)kNoAllocScratchWords);
  if (scratch_on_device[dev] == nullptr) {
    float* p = nullptr;
    DTFE_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_mlp2_noalloc_scratch));
    DTFE_REQUIRE(p != nullptr, "dtfe_mlp2_step_noalloc: cudaGetSymbolAddress returned a null address");
    scratch_on_device[dev] = p;
  }
  Mlp2Launch L = *launch;
  L.workspace = scratch_on_device[dev];
  Mlp2Args a{};
  if (int rc = mlp2_fill(L, a)) return rc;
  // The array still holds whatever the previous call (finished or aborted) left behind.  Only the
  // sync words have to start at zero; every other region is written before it is read.
  DTFE_CUDA_CHECK(cudaMemsetAsync(a.sync, 0, 4 * sizeof(unsigned), (cudaStream_t)stream));
  return L.ks == 8 ? mlp2_launch<8>(a, (cudaStream_t)stream) : mlp2_launch<16>(a, (cudaStream_t)stream);
}

// ---- host-fed training loop (the reference's feed_dict path, worker.py:133-137) -----------------
// One call runs `nsteps` steps, each fed by its own batch in pinned host memory.  Up to
// `steps_per_launch` consecutive steps form a group that shares one kernel launch; groups go through a
// ring of staging slots:
//   copy stream     H2D of every image batch of the group (its own copy each) and ONE copy of the
//                   group's label vectors, gathered on the host (a copy costs its engine ~4 us before
//                   the first byte moves - scripts/h2d_probe.py - a third of a step);
//   caller's stream the step kernel, waiting on exactly one event (the group's copies);
//   back stream     D2H of the group's losses.
// Slots are recycled by the HOST (cudaEventSynchronize on events that completed kFeedRing groups ago), so
// the kernel stream waits on exactly one event per launch.  Doing the loop here instead of in Python keeps
// the per-step host cost at one driver call (measured 1.6 us per step).
namespace {
constexpr int kFeedMaxSteps = 16;
constexpr int kFeedRing = 4;
struct HostFeed {
  int B, D, S, dev;                    // S = training steps (host batches) fused into one kernel launch
  float* xs[kFeedRing];                // staging: S batches per slot
  int* ys[kFeedRing];
  int* ypin[kFeedRing];                // pinned host staging: the S label vectors of a group, gathered for ONE copy
  float* out[kFeedRing];               // per slot: loss[S] | acc[S] | step[S] (int)
  cudaStream_t copy, back;             // H2D and D2H side streams
  cudaEvent_t copied[kFeedRing], done[kFeedRing], readback[kFeedRing];
};
}  // namespace

DTFE_API int dtfe_hostfeed_create(int B, int D, int steps_per_launch, void** out) {
  DTFE_REQUIRE(B >= 1 && D >= 1, "dtfe_hostfeed_create: bad shape");
  DTFE_REQUIRE(steps_per_launch >= 1 && steps_per_launch <= kFeedMaxSteps,
               "dtfe_hostfeed_create: steps_per_launch must be in [1,%d]", kFeedMaxSteps);
  HostFeed* h = new HostFeed();
  h->B = B; h->D = D; h->S = steps_per_launch;
  DTFE_CUDA_CHECK(cudaGetDevice(&h->dev));
  for (int i = 0; i < kFeedRing; ++i) {
    DTFE_CUDA_CHECK(cudaMalloc(&h->xs[i], (size_t)h->S * B * D * sizeof(float)));
    DTFE_CUDA_CHECK(cudaMalloc(&h->ys[i], (size_t)h->S * B * sizeof(int)));
    DTFE_CUDA_CHECK(cudaHostAlloc((void**)&h->ypin[i], (size_t)h->S * B * sizeof(int), cudaHostAllocDefault));
    DTFE_CUDA_CHECK(cudaMalloc(&h->out[i], 3 * kFeedMaxSteps * sizeof(float)));
    DTFE_CUDA_CHECK(cudaEventCreateWithFlags(&h->copied[i], cudaEventDisableTiming));
    DTFE_CUDA_CHECK(cudaEventCreateWithFlags(&h->done[i], cudaEventDisableTiming));
    DTFE_CUDA_CHECK(cudaEventCreateWithFlags(&h->readback[i], cudaEventDisableTiming));
  }
  DTFE_CUDA_CHECK(cudaStreamCreateWithFlags(&h->copy, cudaStreamNonBlocking));
  DTFE_CUDA_CHECK(cudaStreamCreateWithFlags(&h->back, cudaStreamNonBlocking));
  *out = h;
  return 0;
}

DTFE_API int dtfe_hostfeed_destroy(void* handle) {
  HostFeed* h = static_cast<HostFeed*>(handle);
  if (!h) return 0;
  for (int i = 0; i < kFeedRing; ++i) {
    cudaFree(h->xs[i]); cudaFree(h->ys[i]); cudaFree(h->out[i]); cudaFreeHost(h->ypin[i]);
    cudaEventDestroy(h->copied[i]); cudaEventDestroy(h->done[i]); cudaEventDestroy(h->readback[i]);
  }
  cudaStreamDestroy(h->copy);
  cudaStreamDestroy(h->back);
  delete h;
  return 0;
}

// launch: B/D/H/C/ks, parameter pointers, workspace, lr, atomic, train, remote are used; X/labels/
// batch_idx/nsteps are supplied from the host batches.  Every step's batch is copied H2D from its own
// pinned buffer (x_host[i], y_host[i]) and every step's loss is copied back D2H into loss_host[i]
// (pinned).  The call returns after the last read-back has completed, so every loss_host[i] is valid.
DTFE_API int dtfe_mlp2_train_host(void* handle, const dtfe_mlp2_launch* launch, const float* const* x_host,
                                  const int* const* y_host, int nsteps, float* loss_host, void* stream) {
  HostFeed* h = static_cast<HostFeed*>(handle);
  DTFE_REQUIRE(h != nullptr, "dtfe_mlp2_train_host: null feed handle");
  DTFE_REQUIRE(launch->B == h->B && launch->D == h->D, "dtfe_mlp2_train_host: feed was created for B=%d D=%d", h->B, h->D);
  cudaStream_t st = (cudaStream_t)stream;
  Mlp2Launch L = *launch;
  L.batch_idx = nullptr; L.rows_per_step = h->B; L.ldx = h->D; L.dbg = nullptr; L.cursor = nullptr;
  Mlp2Args a{};
  const size_t xbytes = (size_t)h->B * h->D * sizeof(float), ybytes = (size_t)h->B * sizeof(int);
  int group = 0, last_slot = 0, last_n = 0;
  static const bool feed_timing = getenv("DTFE_FEED_TIMING") != nullptr;   // debugging aid: host issue cost
  struct timespec ts0, ts1;
  if (feed_timing) clock_gettime(CLOCK_MONOTONIC, &ts0);
  for (int i = 0; i < nsteps; i += h->S, ++group) {
    const int s = group % kFeedRing;
    const int n = nsteps - i < h->S ? nsteps - i : h->S;
    // The slot is free once the launch that last read it has finished and its losses have been read back
    // (kFeedRing groups ago: these return at once unless the host runs that far ahead of the GPU).
    DTFE_CUDA_CHECK(cudaEventSynchronize(h->done[s]));
    DTFE_CUDA_CHECK(cudaEventSynchronize(h->readback[s]));
    for (int j = 0; j < n; ++j) memcpy(h->ypin[s] + (size_t)j * h->B, y_host[i + j], ybytes);
    DTFE_CUDA_CHECK(cudaMemcpyAsync(h->ys[s], h->ypin[s], (size_t)n * ybytes, cudaMemcpyHostToDevice, h->copy));
    for (int j = 0; j < n; ++j)
      DTFE_CUDA_CHECK(cudaMemcpyAsync(h->xs[s] + (size_t)j * h->B * h->D, x_host[i + j], xbytes, cudaMemcpyHostToDevice, h->copy));
    DTFE_CUDA_CHECK(cudaEventRecord(h->copied[s], h->copy));
    DTFE_CUDA_CHECK(cudaStreamWaitEvent(st, h->copied[s], 0));
    L.X = h->xs[s]; L.labels = h->ys[s]; L.nsteps = n;
    L.loss_out = h->out[s]; L.acc_out = h->out[s] + kFeedMaxSteps;
    L.step_out = reinterpret_cast<int*>(h->out[s] + 2 * kFeedMaxSteps);
    if (int rc = mlp2_fill(L, a)) return rc;
    int rc = L.ks == 8 ? mlp2_launch<8>(a, st) : mlp2_launch<16>(a, st);
    if (rc) return rc;
    DTFE_CUDA_CHECK(cudaEventRecord(h->done[s], st));
    // D2H of these steps' losses on the read-back stream
    DTFE_CUDA_CHECK(cudaStreamWaitEvent(h->back, h->done[s], 0));
    DTFE_CUDA_CHECK(cudaMemcpyAsync(loss_host + i, h->out[s], (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, h->back));
    DTFE_CUDA_CHECK(cudaEventRecord(h->readback[s], h->back));
    last_slot = s; last_n = n;
  }
  if (feed_timing && nsteps > 0) {
    clock_gettime(CLOCK_MONOTONIC, &ts1);
    fprintf(stderr, "dtfe_mlp2_train_host: issued %d steps in %.1f us of host time (%.2f us/step)\n", nsteps,
            ((ts1.tv_sec - ts0.tv_sec) * 1e9 + (ts1.tv_nsec - ts0.tv_nsec)) * 1e-3,
            ((ts1.tv_sec - ts0.tv_sec) * 1e9 + (ts1.tv_nsec - ts0.tv_nsec)) * 1e-3 / nsteps);
  }
  if (nsteps > 0) {
    // keep the caller-visible output views (launch->loss_out etc.) in step with the last step
    const float* o = h->out[last_slot];
    DTFE_CUDA_CHECK(cudaMemcpyAsync(launch->loss_out, o + last_n - 1, sizeof(float), cudaMemcpyDeviceToDevice, st));
    DTFE_CUDA_CHECK(cudaMemcpyAsync(launch->acc_out, o + kFeedMaxSteps + last_n - 1, sizeof(float), cudaMemcpyDeviceToDevice, st));
    DTFE_CUDA_CHECK(cudaMemcpyAsync(launch->step_out, o + 2 * kFeedMaxSteps + last_n - 1, sizeof(int), cudaMemcpyDeviceToDevice, st));
    for (int k = 0; k < kFeedRing; ++k) DTFE_CUDA_CHECK(cudaEventSynchronize(h->readback[k]));
  }
  return 0;
}
