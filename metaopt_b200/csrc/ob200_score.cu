// a14-a17: fused cross-kernel + posterior mean/variance + acquisition + argmax.
//
// Replaces, per candidate x (sklearn/gaussian_process/_gpr.py:444-500, kernels.py:1720-1729 /
// :1569-1570, skopt.acquisition.gaussian_ei/pi/lcb, numpy.argmin):
//     k*   = amplitude * m(|| (x - X_j)/ls ||)            j = 1..n
//     mu   = y_std * (k* . alpha) + y_mean
//     v    = inv(L) k*            (== solve_triangular(L, k*))
//     var  = max(amplitude - sum v^2, 0) * y_std^2 ; sigma = sqrt(var)
//     EI / PI / -LCB, first-occurrence argmax
//
// One persistent CTA per SM owns tiles of SC_BN candidates (default 32) and runs three phases per tile:
//   phase 1  cross-kernel: the pairwise term cand.X^T is a DMMA (FP64 tensor core) contraction over
//            the D columns, with observed-point tiles (fragment-packed X/ls, ||X||^2, alpha) staged
//            into shared memory by bulk async copies (TMA engine) through an mbarrier ring; the
//            Matern/RBF nonlinearity and the k*.alpha mean are fused into its epilogue.  k* for the
//            tile (n_pad x SC_BN doubles) is written once, already in DMMA B-fragment order, to the
//            CTA's private scratch slice, which is pinned in L2 (persisting access-policy window).
//   phase 2  variance: V = inv(L) K*^T as a lower-triangular DMMA GEMM, SC_BM (192) rows x SC_BN
//            candidates per row block, inv(L) tiles (SC_BM x SC_BK doubles, 48 KB) and k* tiles
//            (SC_BK x SC_BN, 8 KB) streamed by a dedicated producer warp with cp.async.bulk into an
//            SC_STAGES-deep ring; row fragments above the diagonal / beyond n are skipped; the
//            result is never stored, only sum v^2.
//   phase 3  posterior + acquisition + running argmax in registers; per-CTA partials at the end.
// Tile geometry (SC_*) lives in ob200_internal.cuh; DESIGN.md section 3 has the measurements.
#include <math.h>
#include <stdlib.h>

#include "ob200_internal.cuh"

namespace {

struct best_t {
  double value;
  long long index;
  double mu, sigma;
};

__device__ __forceinline__ bool better(double v, long long i, double bv, long long bi) {
  return (v > bv) || (v == bv && i < bi);
}

struct score_args {
  const double* cand; long long ldc; long long M; unsigned long long first_index;
  const double* ls; const double* xs; const double* xx; const double* alpha; const double* linv;
  int n, n_pad, D, D_pad, kind;
  double amplitude, y_mean, y_std;
  int acq_mask; double y_opt, xi, kappa;
  double* kstar;
  double *mu_out, *sigma_out, *acq_out;
  best_t* partials;  // [grid][3]
  long long n_tiles;
  int accumulate;    // 1: fold the partials already stored (multi-chunk calls)
  int p1_stages;     // phase-1 ring depth (1..P1_STAGES, the deepest that fits in shared memory)
  int mean_only;     // 1: phase 1 only, write mu_out (pruned scoring, pass 1) and, if sigma_out is set, an upper bound of sigma
  double diag;       // term added to the diagonal of K (alpha + noise)
  const long long* row_map;    // optional: row r of this launch is row row_map[r] of `cand` (pruned scoring: survivors);
                               // global index = first_index + row_map[r]; dense outputs stay indexed by r
  const long long* M_dev;      // optional: the number of rows is read from device memory at kernel start
};

// ---- branch-free FP64 math for the cross-kernel epilogue ------------------------------------
// The library exp()/sqrt() cost ~150 issued instructions per kernel evaluation (special-case
// paths, integer fix-ups); phase 1 was issue-bound on them.  These straight-line versions are
// ~35 instructions, accurate to ~1-2 ulp on the domain that can occur here, and interleave
// freely across the 16 evaluations each lane owns.

// sqrt(x) for finite x >= 0: float rsqrt seed (23 bits), then two Heron corrections that reuse the
// seed as the (approximate) derivative: error 1.5 eps^2 after the first, 1.5 eps^3 after the second,
// i.e. below double rounding (<= 1 ulp measured).  x = 0 gives 0; x below 1e-30 gives a tiny finite
// value <= 1e-15, which the Matern polynomial maps to amplitude * (1 + O(1e-30)).
__device__ __forceinline__ double sqrt_pos(double x) {
  // seed: the clamp in FP32 (one FMNMX; an FP64 max is five instructions) and the bare MUFU.RSQ (rsqrtf() wraps it in a
  // denormal-input path of three more); same seed bits for every x the clamp lets through
  const float xf = fmaxf((float)x, 1e-30f);
  float yf;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(yf) : "f"(xf));
  const double y0 = (double)yf;
  // y0 / 2 by an exponent decrement (y0 is a normal double in [1e-5, 1e15]): one integer add instead of an FP32 multiply and
  // a second conversion through the quarter-rate pipe; the same value
  const double hy = __hiloint2double(__double2hiint(y0) - 0x00100000, __double2loint(y0));
  double s = x * y0;
  s = fma(fma(-s, s, x), hy, s);
  s = fma(fma(-s, s, x), hy, s);
  return s;
}

// max(x, 0) for the squared distance (rounding can leave it a few ulp below zero): an FP64 max costs seven
// instructions (DSETP.MAX, selects, NaN quieting); masking with the sign takes three and gives the same value for
// every non-NaN x (a NaN -- a NaN coordinate in a candidate -- now stays a NaN and can never win the argmax).
__device__ __forceinline__ double clamp_neg(double x) {
  const int hi = __double2hiint(x), m = ~(hi >> 31);
  return __hiloint2double(hi & m, __double2loint(x) & m);
}

__device__ __forceinline__ double kernel_from_d2(int kind, double d2, double amplitude, const double* __restrict__ tab) {
  if (kind == OB200_KERNEL_MATERN52) {
    const double k = sqrt_pos(d2) * ob_kc[2];  // sqrt(5) * dist
    return amplitude * (fma(k, fma(k, ob_kc[1], 1.0), 1.0) * ob_exp_tab(-k, tab));
  }
  return amplitude * ob_exp_tab(-0.5 * d2, tab);
}

__device__ __forceinline__ void acquisition(double mu, double sigma, double y_opt, double xi, double kappa, double out[3]) {
  double ei = 0.0, pi = 0.0;
  if (sigma > 0.0) {
    const double imp = y_opt - xi - mu;
    const double z = imp / sigma;
    const double cdf = 0.5 * erfc(-z * 0.7071067811865476);
    const double pdf = exp(-0.5 * z * z) * 0.3989422804014327;
    ei = imp * cdf + sigma * pdf;
    pi = cdf;
  }
  out[OB200_ACQ_EI] = ei;
  out[OB200_ACQ_PI] = pi;
  out[OB200_ACQ_LCB] = -(mu - kappa * sigma);
}

__device__ __forceinline__ void consumer_bar() {
  asm volatile("bar.sync 1, %0;" ::"n"(SC_CONSUMER_WARPS * 32) : "memory");
}

#ifndef P1_STAGES
#define P1_STAGES 6   // maximum depth of the phase-1 ring; the launcher picks the deepest that fits
#endif
static_assert(SC_BM % 64 == 0, "row block must be a multiple of the fit block");
static_assert(SC_BM % SC_P1_OBS == 0, "row block must be a multiple of the phase-1 chunk");
static_assert(SC_BN % 32 == 0 && SC_BN <= SC_CONSUMER_WARPS * 32, "candidate tile shape");
static_assert(SC_BK % 8 == 0, "stage depth");

__host__ __device__ inline size_t score_union_elems(int D_pad, int p1_stages, bool mean_only) {
  const size_t un = mean_only ? 0 : (size_t)SC_STAGES * (SC_BM * SC_BK + SC_BK * SC_BN);  // the mean-only pass has no phase 2
  const size_t p1 = (size_t)SC_BN * D_pad + (size_t)p1_stages * (SC_P1_OBS * D_pad + 2 * SC_P1_OBS);
  return p1 > un ? p1 : un;
}

#ifdef SC_PROFILE_CYCLES
// timing experiment only (tools/prof_phases.py): cycles spent by consumer thread 0 of every CTA in each part of a tile
__device__ unsigned long long g_sc_prof[1024 * 8];
#define SC_PROF_MARK(K) if (tid == 0) { const long long now_ = clock64(); prof_acc[K] += now_ - prof_t; prof_t = now_; }
#else
#define SC_PROF_MARK(K)
#endif

// MEAN_ONLY: phase 1 only (pruned scoring, pass 1) -- its own instantiation, so the dense kernel carries none of it
// (two CTAs per SM there: phase 1 is FP64-latency-bound and needs neither the phase-2 buffers nor its registers; no
// setmaxnreg either, so ONE producer warp is enough: 13 warps, 72 registers)
#define SC_MEAN_THREADS ((SC_CONSUMER_WARPS + 1) * 32)
template <bool MEAN_ONLY>
#ifndef SC_MEAN_MIN_CTAS
#define SC_MEAN_MIN_CTAS 2
#endif
__global__ void __launch_bounds__(MEAN_ONLY ? SC_MEAN_THREADS : SC_THREADS, MEAN_ONLY ? SC_MEAN_MIN_CTAS : 1) score_kernel(score_args a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int D_pad = a.D_pad, nd4 = D_pad >> 2;
  const int p1_stage_elems = SC_P1_OBS * D_pad + 2 * SC_P1_OBS;
  double* un = reinterpret_cast<double*>(smem_raw);
  // phase-1 view of the union
  double* Cs = un;                                // [SC_NCF][nd4][32]
  double* p1buf = un + SC_BN * D_pad;             // P1_STAGES x (Xs chunk | xx | alpha)
  // phase-2 view
  double* Abuf = un;                              // SC_STAGES x (SC_BM*SC_BK)
  double* Bbuf = un + SC_STAGES * SC_BM * SC_BK;  // SC_STAGES x (SC_BK*SC_BN)
  const int P1S = a.p1_stages;
  double* small = un + score_union_elems(D_pad, P1S, MEAN_ONLY);
  double* cc = small;                                   // [SC_BN]
  double* mured = cc + SC_BN;                           // [SC_CONSUMER_WARPS][SC_BN]
  double* ssred = mured + SC_CONSUMER_WARPS * SC_BN;    // [SC_WM][SC_BN]
  uint64_t* bars = reinterpret_cast<uint64_t*>(ssred + SC_WM * SC_BN);
  uint64_t* full1 = bars;                 // [P1_STAGES]
  uint64_t* empty1 = full1 + P1_STAGES;
  uint64_t* full2 = empty1 + P1_STAGES;   // [SC_STAGES]
  uint64_t* empty2 = full2 + SC_STAGES;
  double* exp_tab = reinterpret_cast<double*>(empty2 + SC_STAGES);  // 2^(j/32), j = 0..31

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  // batch size: a launch argument, or read from device memory (written by an earlier kernel of the stream)
  const long long n_rows = a.M_dev ? *a.M_dev : a.M;
  const long long n_tiles = a.M_dev ? (n_rows + SC_BN - 1) / SC_BN : a.n_tiles;
  const bool is_producer = (warp >= SC_CONSUMER_WARPS);   // last warpgroup; only its first lane issues copies
  const bool is_issuer = (warp == SC_CONSUMER_WARPS) && (lane == 0);

  if (tid < 32) exp_tab[tid] = exp2((double)tid * 0.03125);
  if (tid == 0) {
    for (int s = 0; s < P1_STAGES; ++s) { mbar_init(&full1[s], 1); mbar_init(&empty1[s], SC_CONSUMER_WARPS); }
    for (int s = 0; s < SC_STAGES; ++s) { mbar_init(&full2[s], 1); mbar_init(&empty2[s], SC_CONSUMER_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  double* kslice = a.kstar + (size_t)blockIdx.x * a.n_pad * SC_BN;
#ifdef SC_DEBUG_SKIP_PHASE1
  const int n_chunks1 = 0;  // timing experiment only: phase 2 + 3 on stale k*
#else
  const int n_chunks1 = (a.n + SC_P1_OBS - 1) / SC_P1_OBS;    // observed-point chunks that hold real rows
#endif
#ifdef SC_DEBUG_SKIP_PHASE2
  const int n_rb = 0;   // timing experiment only: phase 1 + 3
#else
  const int n_rb = (a.n + SC_BM - 1) / SC_BM;                 // row blocks of inv(L) that hold real rows
#endif
  const int k_lim = (int)(((long long)a.n + SC_BK - 1) / SC_BK);  // stages with real columns

  // running ring positions (continue across tiles)
  uint32_t it1 = 0, it2 = 0;

  if (is_producer) {
    // register rebalancing: each role keeps its own tile loop so the register budget is decided by
    // dominance (per SM sub-partition: consumers x SC_CONSUMER_REGS + 1 producer x 40 <= 512).
    if (!MEAN_ONLY) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      if (is_issuer) {
        // ---- phase 1 producer: observed-point chunks
        const uint32_t xs_bytes = (uint32_t)(SC_P1_OBS * D_pad * 8);
        for (int c = 0; c < n_chunks1; ++c, ++it1) {
          const int s = it1 % P1S;
          const uint32_t ph = (it1 / P1S) & 1;
          mbar_wait(&empty1[s], ph ^ 1);
          double* dst = p1buf + (size_t)s * p1_stage_elems;
          mbar_expect_tx(&full1[s], xs_bytes + 2 * SC_P1_OBS * 8);
          bulk_g2s(dst, a.xs + (size_t)c * SC_P1_OBS * D_pad, xs_bytes, &full1[s]);
          bulk_g2s(dst + SC_P1_OBS * D_pad, a.xx + (size_t)c * SC_P1_OBS, SC_P1_OBS * 8, &full1[s]);
          bulk_g2s(dst + SC_P1_OBS * D_pad + SC_P1_OBS, a.alpha + (size_t)c * SC_P1_OBS, SC_P1_OBS * 8, &full1[s]);
        }
      }
      __syncthreads();  // (A) consumers finished phase 1, k* slice written
      if (is_issuer && !MEAN_ONLY) {
        // ---- phase 2 producer: inv(L) tiles + k* tiles
        const int nJ = a.n_pad / SC_BK;
        for (int I = 0; I < n_rb; ++I) {
          int jend = (I + 1) * (SC_BM / SC_BK);
          if (jend > k_lim) jend = k_lim;
          for (int J = 0; J < jend; ++J, ++it2) {
            const int s = it2 % SC_STAGES;
            const uint32_t ph = (it2 / SC_STAGES) & 1;
            mbar_wait(&empty2[s], ph ^ 1);
            mbar_expect_tx(&full2[s], (SC_BM * SC_BK + SC_BK * SC_BN) * 8);
            bulk_g2s(Abuf + (size_t)s * SC_BM * SC_BK, a.linv + ((size_t)I * nJ + J) * (SC_BM * SC_BK), SC_BM * SC_BK * 8, &full2[s]);
            bulk_g2s(Bbuf + (size_t)s * SC_BK * SC_BN, kslice + (size_t)J * SC_BK * SC_BN, SC_BK * SC_BN * 8, &full2[s]);
          }
        }
      }
      __syncthreads();  // (B) end of tile
    }
    return;
  }

  // =============================== consumers ===============================
  if (!MEAN_ONLY) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(SC_CONSUMER_REGS));

  // running best per acquisition (threads 0..SC_BN-1 own one candidate column each)
  best_t best[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) { best[k].value = -INFINITY; best[k].index = 0x7fffffffffffffffLL; best[k].mu = 0.0; best[k].sigma = 0.0; }
  if (a.accumulate && tid == 0) {
#pragma unroll
    for (int k = 0; k < 3; ++k) best[k] = a.partials[(size_t)blockIdx.x * 3 + k];
  }

#ifdef SC_PROFILE_CYCLES
  long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long prof_t = clock64();
#endif
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long row0 = tile * SC_BN;
    SC_PROF_MARK(7)  // (loop overhead / before the first tile)
    // ---- stage the candidate tile: Cs = -2 cand / ls in A-fragment order (the factor makes the DMMA accumulator the
    // -2 c.x term of the squared distance as it stands -- exact, a power of two -- and saves the doubling per kernel
    // evaluation), cc = ||cand / ls||^2
    for (int e = tid; e < SC_BN * D_pad; e += SC_CONSUMER_WARPS * 32) {
      const int r = e / D_pad, d = e - r * D_pad;
      double v = 0.0;
      if (d < a.D && row0 + r < n_rows) v = a.cand[(size_t)(a.row_map ? a.row_map[row0 + r] : row0 + r) * a.ldc + d] / a.ls[d];
      Cs[((size_t)(r >> 3) * nd4 + (d >> 2)) * 32 + (r & 7) * 4 + (d & 3)] = -2.0 * v;
    }
    consumer_bar();
    if (tid < SC_BN) {
      const int r = tid;
      double acc = 0.0;
      for (int d = 0; d < D_pad; ++d) {
        const double v = Cs[((size_t)(r >> 3) * nd4 + (d >> 2)) * 32 + (r & 7) * 4 + (d & 3)];
        acc = fma(v, v, acc);
      }
      cc[r] = 0.25 * acc;
    }
    consumer_bar();

    SC_PROF_MARK(0)  // candidate staging
    // ---- phase 1: k* = kernel(cand, X) and mu partials; warp w owns obs8 block w of each chunk
    {
      double ccr[SC_NCF], mup[SC_NCF];
      // Mean-only pass: (a lower bound of) the largest kernel value against a real observation, kept as the HIGH WORD of
      // the bit pattern.  Kernel values are >= 0, so integer order is numeric order, one 32-bit max per value runs on
      // the integer pipe instead of the busy FP64 one, and dropping the low word can only make the value smaller
      // (by < 2^-20 relative), which keeps the sigma bound derived from it on the safe side.
      int kmx[SC_NCF];
#pragma unroll
      for (int cf = 0; cf < SC_NCF; ++cf) { ccr[cf] = cc[cf * 8 + g]; mup[cf] = 0.0; kmx[cf] = 0; }
      for (int c = 0; c < n_chunks1; ++c, ++it1) {
        const int s = it1 % P1S;
        const uint32_t ph = (it1 / P1S) & 1;
        mbar_wait(&full1[s], ph);
        const double* xs_s = p1buf + (size_t)s * p1_stage_elems + (size_t)warp * nd4 * 32;  // this warp's obs8 block
        const double* xx_s = p1buf + (size_t)s * p1_stage_elems + SC_P1_OBS * D_pad + warp * 8;
        const double* al_s = xx_s + SC_P1_OBS;
        double acc[SC_NCF][2];
#pragma unroll
        for (int cf = 0; cf < SC_NCF; ++cf) acc[cf][0] = acc[cf][1] = 0.0;
        for (int d4 = 0; d4 < nd4; ++d4) {
          const double b = xs_s[d4 * 32 + lane];
#pragma unroll
          for (int cf = 0; cf < SC_NCF; ++cf) dmma884(acc[cf][0], acc[cf][1], Cs[((size_t)cf * nd4 + d4) * 32 + lane], b);
        }
        const double xx0 = xx_s[2 * q], xx1 = xx_s[2 * q + 1];
        const double al0 = al_s[2 * q], al1 = al_s[2 * q + 1];
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty1[s]);
        const int ob = c * SC_CONSUMER_WARPS + warp;  // global obs8 block
        double2* dst = reinterpret_cast<double2*>(kslice + (size_t)ob * SC_NCF * 64) + lane;
#pragma unroll
        for (int cf = 0; cf < SC_NCF; ++cf) {
          const double d20 = clamp_neg(ccr[cf] + xx0 + acc[cf][0]);
          const double d21 = clamp_neg(ccr[cf] + xx1 + acc[cf][1]);
          double2 kv;
          kv.x = kernel_from_d2(a.kind, d20, a.amplitude, exp_tab);
          kv.y = kernel_from_d2(a.kind, d21, a.amplitude, exp_tab);
          mup[cf] = fma(kv.x, al0, mup[cf]);
          mup[cf] = fma(kv.y, al1, mup[cf]);
          if (!MEAN_ONLY) dst[cf * 32] = kv;
          else if (ob * 8 + 8 <= a.n) kmx[cf] = max(kmx[cf], max(__double2hiint(kv.x), __double2hiint(kv.y)));
          else kmx[cf] = max(kmx[cf], max(ob * 8 + 2 * q < a.n ? __double2hiint(kv.x) : 0,
                                           ob * 8 + 2 * q + 1 < a.n ? __double2hiint(kv.y) : 0));  // padded observations
        }
      }
      // mu: reduce over q (4 lanes) then across warps through shared memory
#pragma unroll
      for (int cf = 0; cf < SC_NCF; ++cf) {
        double v = mup[cf];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        if (q == 0) mured[warp * SC_BN + cf * 8 + g] = v;
#if SC_WN == 1
        if (MEAN_ONLY) {  // ssred is idle in this mode: it carries the per-warp maxima to phase 3
          int kx = kmx[cf];
          kx = max(kx, __shfl_xor_sync(0xffffffffu, kx, 1));
          kx = max(kx, __shfl_xor_sync(0xffffffffu, kx, 2));
          if (q == 0) ssred[warp * SC_BN + cf * 8 + g] = __hiloint2double(kx, 0);
        }
#endif
      }
    }
    SC_PROF_MARK(1)       // phase 1 loop of this warp
    fence_proxy_async();  // k* global writes (generic proxy) -> bulk reads (async proxy)
    __syncthreads();      // (A)
    SC_PROF_MARK(2)       // wait for the other warps at (A)

    // ---- phase 2: sum v^2 with v = inv(L) k*
    const int wm = warp / SC_WN, wn = warp % SC_WN;
    double ss[SC_FN][2];
#pragma unroll
    for (int j = 0; j < SC_FN; ++j) ss[j][0] = ss[j][1] = 0.0;
    for (int I = 0; I < (MEAN_ONLY ? 0 : n_rb); ++I) {
      double acc[SC_FM][SC_FN][2];
#pragma unroll
      for (int t = 0; t < SC_FM; ++t)
#pragma unroll
        for (int j = 0; j < SC_FN; ++j) acc[t][j][0] = acc[t][j][1] = 0.0;
      int jend = (I + 1) * (SC_BM / SC_BK);
      if (jend > k_lim) jend = k_lim;
      // first row of fragment t: row fragments are dealt back and forth over the warp rows (see pack_linv_kernel)
      const int rb_even = I * SC_BM + wm * 8, rb_odd = I * SC_BM + (SC_WM - 1 - wm) * 8;
      for (int J = 0; J < jend; ++J, ++it2) {
        const int s = it2 % SC_STAGES;
        const uint32_t ph = (it2 / SC_STAGES) & 1;
        mbar_wait(&full2[s], ph);
        const double2* As = reinterpret_cast<const double2*>(Abuf + (size_t)s * SC_BM * SC_BK) + (size_t)wm * SC_FM * (SC_BK / 8) * 32 + lane;
        const double2* Bs = reinterpret_cast<const double2*>(Bbuf + (size_t)s * SC_BK * SC_BN) + (size_t)wn * SC_FN * 32 + lane;
        // Stages strictly below the diagonal with every row fragment inside n need no predicates:
        // plain unrolled loads + DMMAs.  Diagonal / last-row-block stages predicate each fragment.
        const bool full = (J * SC_BK + SC_BK - 1 <= I * SC_BM) && (I * SC_BM + SC_BM <= a.n);
        if (full) {
#pragma unroll
          for (int k8 = 0; k8 < SC_BK / 8; ++k8) {
            double2 b[SC_FN], av[SC_FM];
#pragma unroll
            for (int j = 0; j < SC_FN; ++j) b[j] = Bs[(k8 * SC_NCF + j) * 32];
#pragma unroll
            for (int t = 0; t < SC_FM; ++t) av[t] = As[(t * (SC_BK / 8) + k8) * 32];
#pragma unroll
            for (int t = 0; t < SC_FM; ++t)
#pragma unroll
              for (int j = 0; j < SC_FN; ++j) dmma884(acc[t][j][0], acc[t][j][1], av[t].x, b[j].x);
#pragma unroll
            for (int t = 0; t < SC_FM; ++t)
#pragma unroll
              for (int j = 0; j < SC_FN; ++j) dmma884(acc[t][j][0], acc[t][j][1], av[t].y, b[j].y);
          }
        } else {
          // Row fragments of a warp are ordered by row, so the active ones form a range [t0, t1):
          // t >= t0 lies on/below the diagonal of this 8-column block, t < t1 lies inside n.  The
          // fall-through switch makes the skip a real (warp-uniform) branch; predicated-off DMMAs
          // would still occupy the tensor pipe.
#pragma unroll
          for (int k8 = 0; k8 < SC_BK / 8; ++k8) {
            const int k0 = J * SC_BK + k8 * 8;
            double2 b[SC_FN], av[SC_FM];
#pragma unroll
            for (int j = 0; j < SC_FN; ++j) b[j] = Bs[(k8 * SC_NCF + j) * 32];
#pragma unroll
            for (int t = 0; t < SC_FM; ++t) av[t] = As[(t * (SC_BK / 8) + k8) * 32];
            int t0 = SC_FM, t1 = 0;
#pragma unroll
            for (int t = SC_FM - 1; t >= 0; --t) {
              const int r0 = ((t & 1) ? rb_odd : rb_even) + 8 * SC_WM * t;
              if (r0 + 7 >= k0) t0 = t;
              if (r0 < a.n && t1 == 0) t1 = t + 1;
            }
#define SC_FRAG(T)                                                                                   \
  {                                                                                                  \
    _Pragma("unroll") for (int j = 0; j < SC_FN; ++j) dmma884(acc[T][j][0], acc[T][j][1], av[T].x, b[j].x); \
    _Pragma("unroll") for (int j = 0; j < SC_FN; ++j) dmma884(acc[T][j][0], acc[T][j][1], av[T].y, b[j].y); \
  }
            static_assert(SC_FM == 2 || SC_FM == 4 || SC_FM == 8, "fragment switch is written for 2, 4 or 8 row fragments per warp");
            if (t0 < t1) {
              switch (t0) {
                case 0: SC_FRAG(0) if (t1 == 1) break;
                case 1: SC_FRAG(1)
#if SC_FM >= 4
                  if (t1 == 2) break;
                case 2: SC_FRAG(2) if (t1 == 3) break;
                case 3: SC_FRAG(3)
#endif
#if SC_FM == 8
                  if (t1 == 4) break;
                case 4: SC_FRAG(4) if (t1 == 5) break;
                case 5: SC_FRAG(5) if (t1 == 6) break;
                case 6: SC_FRAG(6) if (t1 == 7) break;
                case 7: SC_FRAG(7)
#endif
                default: break;
              }
            }
#undef SC_FRAG
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty2[s]);
      }
#pragma unroll
      for (int t = 0; t < SC_FM; ++t)
#pragma unroll
        for (int j = 0; j < SC_FN; ++j) {
          ss[j][0] = fma(acc[t][j][0], acc[t][j][0], ss[j][0]);
          ss[j][1] = fma(acc[t][j][1], acc[t][j][1], ss[j][1]);
        }
    }
    SC_PROF_MARK(3)  // phase 2 (incl. pipeline fill)
    // reduce over the 8 row lanes (g), then across the SC_WM row warps via shared memory
    if (!MEAN_ONLY)
#pragma unroll
    for (int j = 0; j < SC_FN; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        double v = ss[j][e];
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        if (g == 0) ssred[wm * SC_BN + wn * SC_FN * 8 + j * 8 + 2 * q + e] = v;
      }
    consumer_bar();

    // ---- phase 3: posterior, acquisition, running argmax (one thread per candidate)
    if (tid < SC_BN && row0 + tid < n_rows) {
      double s2 = 0.0;
#pragma unroll
      for (int w = 0; w < SC_WM; ++w) s2 += ssred[w * SC_BN + tid];
      double m = 0.0;
#pragma unroll
      for (int w = 0; w < SC_CONSUMER_WARPS; ++w) m += mured[w * SC_BN + tid];
      const double mu = a.y_std * m + a.y_mean;
      const long long row = row0 + tid;
      if (MEAN_ONLY) {
        a.mu_out[row] = mu;
        if (a.sigma_out) {
          // Conditioning on ONE observation j already gives var <= amp - k_j^2 / (amp + diag), and more observations
          // only lower it: the nearest observation bounds sigma without any triangular product.  The slack covers the
          // rounding of the full computation it is compared with (amp - sum v^2 cancels ~cond(K) eps).
          double var_ub = a.amplitude;
#if SC_WN == 1
          double kx = 0.0;
#pragma unroll
          for (int w = 0; w < SC_WM; ++w) kx = fmax(kx, ssred[w * SC_BN + tid]);
          var_ub = fmin(fmax(a.amplitude - kx * kx / (a.amplitude + a.diag), 0.0) + 1e-6 * a.amplitude, a.amplitude);
#endif
          a.sigma_out[row] = a.y_std * sqrt(var_ub);
        }
      } else {
        double var = a.amplitude - s2;
        if (var < 0.0) var = 0.0;
        var = var * (a.y_std * a.y_std);
        const double sigma = sqrt(var);
        double acq[3];
        acquisition(mu, sigma, a.y_opt, a.xi, a.kappa, acq);
        if (a.mu_out) a.mu_out[row] = mu;
        if (a.sigma_out) a.sigma_out[row] = sigma;
        const long long gidx = (long long)a.first_index + (a.row_map ? a.row_map[row] : row);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          if (!(a.acq_mask & (1 << k))) continue;
          if (a.acq_out) a.acq_out[(size_t)k * a.M + row] = acq[k];
          if (better(acq[k], gidx, best[k].value, best[k].index)) { best[k].value = acq[k]; best[k].index = gidx; best[k].mu = mu; best[k].sigma = sigma; }
        }
      }
    }
    SC_PROF_MARK(4)  // reductions + phase 3
    fence_proxy_async();
    __syncthreads();  // (B) end of tile: smem union and the k* slice may be reused
    SC_PROF_MARK(5)  // wait at (B)
  }
#ifdef SC_PROFILE_CYCLES
  if (tid == 0 && blockIdx.x < 1024)
    for (int k = 0; k < 8; ++k) g_sc_prof[blockIdx.x * 8 + k] = (unsigned long long)prof_acc[k];
#endif

  // ---- per-CTA winner: warp-shuffle argmax inside the SC_BN/32 owner warps, then through smem
  if (MEAN_ONLY) return;   // no acquisition was evaluated: the partials (possibly an earlier batch's) stay untouched
  if (warp < SC_BN / 32) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      double v = best[k].value, m = best[k].mu, sg = best[k].sigma;
      long long i = best[k].index;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, v, off);
        const long long oi = __shfl_xor_sync(0xffffffffu, i, off);
        const double om = __shfl_xor_sync(0xffffffffu, m, off);
        const double os = __shfl_xor_sync(0xffffffffu, sg, off);
        if (better(ov, oi, v, i)) { v = ov; i = oi; m = om; sg = os; }
      }
      if (lane == 0) {
        double* slot = mured + (warp * 3 + k) * 4;
        slot[0] = v; slot[1] = __longlong_as_double(i); slot[2] = m; slot[3] = sg;
      }
    }
  }
  consumer_bar();
  if (tid < 3) {
    best_t b;
    b.value = -INFINITY; b.index = 0x7fffffffffffffffLL; b.mu = 0.0; b.sigma = 0.0;
    for (int w = 0; w < SC_BN / 32; ++w) {
      const double* sl = mured + (w * 3 + tid) * 4;
      const long long i1 = __double_as_longlong(sl[1]);
      if (better(sl[0], i1, b.value, b.index)) { b.value = sl[0]; b.index = i1; b.mu = sl[2]; b.sigma = sl[3]; }
    }
    a.partials[(size_t)blockIdx.x * 3 + tid] = b;
  }
}

// final reduction of per-CTA partials -> best[3]
__global__ void __launch_bounds__(256) reduce_best_kernel(const best_t* __restrict__ partials, int n_part, int acq_mask,
                                                          ob200_best* __restrict__ out) {
  __shared__ best_t sm[256];
  for (int k = 0; k < 3; ++k) {
    best_t b;
    b.value = -INFINITY; b.index = 0x7fffffffffffffffLL; b.mu = 0.0; b.sigma = 0.0;
    for (int p = threadIdx.x; p < n_part; p += 256) {
      const best_t o = partials[(size_t)p * 3 + k];
      if (better(o.value, o.index, b.value, b.index)) b = o;
    }
    sm[threadIdx.x] = b;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
      if (threadIdx.x < s) {
        const best_t o = sm[threadIdx.x + s];
        if (better(o.value, o.index, sm[threadIdx.x].value, sm[threadIdx.x].index)) sm[threadIdx.x] = o;
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      ob200_best r;
      const bool on = (acq_mask >> k) & 1;
      r.value = on ? sm[0].value : 0.0;
      r.index = on ? sm[0].index : -1;
      r.mu = on ? sm[0].mu : 0.0;
      r.sigma = on ? sm[0].sigma : 0.0;
      out[k] = r;
    }
    __syncthreads();
  }
}

// a16-a17 alone: acquisition + argmax over given mu/sigma; per-CTA partials in scratch
__global__ void __launch_bounds__(256) acq_partial_kernel(const double* __restrict__ mu, const double* __restrict__ sigma, long long M,
                                                          unsigned long long first_index, int acq_mask, double y_opt, double xi,
                                                          double kappa, double* __restrict__ acq_out, best_t* __restrict__ partials) {
  __shared__ best_t sm[256];
  best_t best[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) { best[k].value = -INFINITY; best[k].index = 0x7fffffffffffffffLL; best[k].mu = 0.0; best[k].sigma = 0.0; }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += (long long)gridDim.x * blockDim.x) {
    const double m = mu[i], s = sigma[i];
    double acq[3];
    acquisition(m, s, y_opt, xi, kappa, acq);
    const long long gi = (long long)first_index + i;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      if (!(acq_mask & (1 << k))) continue;
      if (acq_out) acq_out[(size_t)k * M + i] = acq[k];
      if (better(acq[k], gi, best[k].value, best[k].index)) { best[k].value = acq[k]; best[k].index = gi; best[k].mu = m; best[k].sigma = s; }
    }
  }
  for (int k = 0; k < 3; ++k) {
    sm[threadIdx.x] = best[k];
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
      if (threadIdx.x < s) {
        const best_t o = sm[threadIdx.x + s];
        if (better(o.value, o.index, sm[threadIdx.x].value, sm[threadIdx.x].index)) sm[threadIdx.x] = o;
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) partials[(size_t)blockIdx.x * 3 + k] = sm[0];
    __syncthreads();
  }
}

size_t score_smem_bytes(int D_pad, int p1_stages, bool mean_only = false) {
  return (score_union_elems(D_pad, p1_stages, mean_only) + SC_BN + SC_CONSUMER_WARPS * SC_BN + SC_WM * SC_BN) * sizeof(double) +
         (2 * P1_STAGES + 2 * SC_STAGES) * sizeof(uint64_t) + 32 * sizeof(double) + 128;
}

int score_launch(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, uint64_t first_index, int acq_mask,
                 double y_opt, double xi, double kappa, double* mu_dev, double* sigma_dev, double* acq_dev, int accumulate,
                 int* grid_out, int mean_only = 0, const long long* row_map = nullptr, const long long* M_dev = nullptr) {
  const ob200_fit_header& h = ctx->hdr;
  const ob200_blob_layout lay = ob_blob_layout(h.n_pad, h.D_pad);
  // phase-1 ring depth: the deepest that fits (observed-point chunks are small, so the bulk-copy latency has
  // to be covered by depth, not by chunk size); OB200_P1_STAGES overrides for tuning
  int p1_stages = P1_STAGES;
  if (const char* env = getenv("OB200_P1_STAGES")) { const int v = atoi(env); if (v >= 1 && v <= P1_STAGES) p1_stages = v; }
  size_t smem = score_smem_bytes(h.D_pad, p1_stages);
  while (smem > 227 * 1024 && p1_stages > 1) { --p1_stages; smem = score_smem_bytes(h.D_pad, p1_stages); }
  if (smem > 227 * 1024) OB_FAIL(ctx, OB200_EINVAL, "D=%d too wide for the score kernel's shared-memory tile", h.D);
  int ctas_per_sm = 1;
  if (mean_only) {  // phase 1 alone: two CTAs per SM when a ring of at least two stages fits in half the shared memory
    smem = score_smem_bytes(h.D_pad, p1_stages, true);
    int st2 = p1_stages;
    while (st2 > 2 && score_smem_bytes(h.D_pad, st2, true) > 112 * 1024) --st2;
    if (score_smem_bytes(h.D_pad, st2, true) <= 112 * 1024 && !getenv("OB200_MEAN_1CTA")) {
      p1_stages = st2;
      smem = score_smem_bytes(h.D_pad, st2, true);
      ctas_per_sm = 2;
    }
  }
  const int64_t n_tiles = (M + SC_BN - 1) / SC_BN;
  const int64_t max_grid = (int64_t)ctx->num_sms * ctas_per_sm;
  const int grid = (int)((n_tiles < max_grid && !M_dev) ? n_tiles : max_grid);   // device-sized batches: full persistent grid
  int rc;
  if ((rc = ob_reserve(ctx, &ctx->kstar_dev, &ctx->kstar_cap, (int64_t)ctx->num_sms * h.n_pad * SC_BN))) return rc;
  if ((rc = ob_reserve(ctx, &ctx->partial_dev, &ctx->partial_cap, (int64_t)4096 * 3 * (int64_t)sizeof(best_t)))) return rc;
  score_args a{};
  a.cand = cand_dev; a.ldc = ldc; a.M = M; a.first_index = first_index;
  a.ls = reinterpret_cast<const double*>(ctx->blob + lay.off_ls);
  a.xs = reinterpret_cast<const double*>(ctx->blob + lay.off_xs);
  a.xx = reinterpret_cast<const double*>(ctx->blob + lay.off_xx);
  a.alpha = reinterpret_cast<const double*>(ctx->blob + lay.off_alpha);
  a.linv = reinterpret_cast<const double*>(ctx->blob + lay.off_linv);
  a.n = h.n; a.n_pad = h.n_pad; a.D = h.D; a.D_pad = h.D_pad; a.kind = h.kernel_kind;
  a.amplitude = h.amplitude; a.y_mean = h.y_mean; a.y_std = h.y_std;
  a.acq_mask = acq_mask; a.y_opt = y_opt; a.xi = xi; a.kappa = kappa;
  a.kstar = ctx->kstar_dev;
  a.mu_out = mu_dev; a.sigma_out = sigma_dev; a.acq_out = acq_dev;
  a.partials = reinterpret_cast<best_t*>(ctx->partial_dev);
  a.n_tiles = n_tiles;
  a.accumulate = accumulate;
  a.p1_stages = p1_stages;
  a.mean_only = mean_only;
  a.diag = h.diag;
  a.row_map = row_map;
  a.M_dev = M_dev;
  auto kernel = mean_only ? score_kernel<true> : score_kernel<false>;
  OB_CUDA(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // k* slices -> persisting L2 lines for this launch (the window is a launch-time stream attribute)
  const size_t slices = (size_t)grid * h.n_pad * SC_BN * sizeof(double);
  // (not for launches of a wave or two -- the first batches of the pruned path: their slices are written and read once
  // while they are the newest lines in L2 anyway, and the two cudaDeviceSetLimit calls cost ~35 us per call)
  const bool persist = ctx->l2_persist_max > 0 && !mean_only && (M_dev == nullptr) && n_tiles > 2 * (long long)grid;
  if (persist) {
    ob_l2_acquire(ctx);
    cudaStreamAttrValue attr{};
    attr.accessPolicyWindow.base_ptr = ctx->kstar_dev;
    attr.accessPolicyWindow.num_bytes = slices < ctx->l2_window_max ? slices : ctx->l2_window_max;
    attr.accessPolicyWindow.hitRatio = ctx->l2_persist_max >= slices ? 1.0f : (float)ctx->l2_persist_max / (float)slices;
    attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    OB_CUDA(ctx, cudaStreamSetAttribute(ctx->stream, cudaStreamAttributeAccessPolicyWindow, &attr));
  }
  kernel<<<grid, mean_only ? SC_MEAN_THREADS : SC_THREADS, smem, ctx->stream>>>(a);
  OB_LAUNCH_CHECK(ctx);
  if (persist) {  // later launches on this (caller-owned) stream must not inherit the window
    cudaStreamAttrValue attr{};
    attr.accessPolicyWindow.num_bytes = 0;
    OB_CUDA(ctx, cudaStreamSetAttribute(ctx->stream, cudaStreamAttributeAccessPolicyWindow, &attr));
    ctx->l2_dirty = true;
  }
  *grid_out = grid;
  return OB200_OK;
}

// the persisting-L2 carve-out is handed back on every exit path of a synchronous entry point
struct l2_scope {
  ob200_ctx* ctx;
  explicit l2_scope(ob200_ctx* c) : ctx(c) {}
  ~l2_scope() { ob_l2_release(ctx); }
};

int check_score_args(ob200_ctx* ctx, int64_t M, int acq_mask, const void* best, double y_opt, double xi, double kappa) {
  if (!ctx->fitted) OB_FAIL(ctx, OB200_ESTATE, "ob200_score called before ob200_gp_fit / ob200_gp_import");
  if (!isfinite(y_opt) || !isfinite(xi) || !isfinite(kappa)) OB_FAIL(ctx, OB200_EINVAL, "y_opt, xi and kappa must be finite");
  if (M < 1) OB_FAIL(ctx, OB200_EINVAL, "M must be >= 1");
  if ((acq_mask & 7) == 0 || (acq_mask & ~7)) OB_FAIL(ctx, OB200_EINVAL, "acq_mask must select EI(1), PI(2) and/or LCB(4)");
  if (!best) OB_FAIL(ctx, OB200_EINVAL, "best is null");
  return OB200_OK;
}

}  // namespace

extern "C" {

#ifdef SC_PROFILE_CYCLES
int ob200_debug_score_cycles(unsigned long long* host, int n_cta) {
  return cudaMemcpyFromSymbol(host, g_sc_prof, (size_t)n_cta * 8 * sizeof(unsigned long long)) == cudaSuccess ? 0 : -2;
}
#endif

int ob200_score_async(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, uint64_t seed, uint64_t first_index,
                      int acq_mask, double y_opt, double xi, double kappa, ob200_best* best_dev, double* mu_dev,
                      double* sigma_dev, double* acq_dev) {
  if (!ctx) return OB200_EINVAL;
  int rc = check_score_args(ctx, M, acq_mask, best_dev, y_opt, xi, kappa);
  if (rc) return rc;
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  const int D = ctx->hdr.D;
  int grid = 0;
  if (cand_dev) {
    if (ldc < D) OB_FAIL(ctx, OB200_EINVAL, "ldc=%lld < D=%d", (long long)ldc, D);
    if ((rc = score_launch(ctx, cand_dev, ldc, M, first_index, acq_mask, y_opt, xi, kappa, mu_dev, sigma_dev, acq_dev, 0, &grid))) return rc;
  } else {
    // candidates from the registered space: generate (normalised) in chunks, fold the argmax
    if (ctx->n_slots <= 0) OB_FAIL(ctx, OB200_ESTATE, "cand_dev is NULL and no space was set");
    if (ctx->d_out != D) OB_FAIL(ctx, OB200_EINVAL, "space width %d != fitted D %d", ctx->d_out, D);
    int64_t chunk_rows = ((int64_t)1 << 31) / ((int64_t)D * 8) / SC_BN * SC_BN;  // <= 2 GiB of candidates at a time
    if (const char* env = getenv("OB200_CHUNK_ROWS")) {  // testing hook: exercise the multi-chunk path at small sizes
      const long long v = atoll(env);
      if (v >= SC_BN) chunk_rows = v / SC_BN * SC_BN;
    }
    if (acq_dev && M > chunk_rows) OB_FAIL(ctx, OB200_EINVAL, "dense acquisition output needs M <= %lld when candidates are generated", (long long)chunk_rows);
    const int64_t rows_cap = M < chunk_rows ? M : chunk_rows;
    if ((rc = ob_reserve(ctx, &ctx->cand_tmp, &ctx->cand_cap, rows_cap * D))) return rc;
    int grid_max = 0;
    for (int64_t r0 = 0; r0 < M; r0 += chunk_rows) {
      const int64_t rows = (M - r0 < chunk_rows) ? M - r0 : chunk_rows;
      if ((rc = ob200_sample_launch(ctx, seed, first_index + (uint64_t)r0, rows, 1, ctx->cand_tmp, D, nullptr))) return rc;
      // every chunk but the last is full-size, so the first launch has the widest grid and
      // initialises every partial the final reduction reads; later launches fold into their own.
      if ((rc = score_launch(ctx, ctx->cand_tmp, D, rows, first_index + (uint64_t)r0, acq_mask, y_opt, xi, kappa,
                             mu_dev ? mu_dev + r0 : nullptr, sigma_dev ? sigma_dev + r0 : nullptr, acq_dev, r0 == 0 ? 0 : 1, &grid)))
        return rc;
      if (grid > grid_max) grid_max = grid;
    }
    grid = grid_max;
  }
  reduce_best_kernel<<<1, 256, 0, ctx->stream>>>(reinterpret_cast<const best_t*>(ctx->partial_dev), grid, acq_mask, best_dev);
  OB_LAUNCH_CHECK(ctx);
  return OB200_OK;
}

int ob200_score(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, uint64_t seed, uint64_t first_index,
                int acq_mask, double y_opt, double xi, double kappa, ob200_best* best_host, double* mu_dev,
                double* sigma_dev, double* acq_dev) {
  if (!ctx) return OB200_EINVAL;
  if (!best_host) OB_FAIL(ctx, OB200_EINVAL, "best_host is null");
  l2_scope l2(ctx);
  int rc = ob200_score_async(ctx, cand_dev, ldc, M, seed, first_index, acq_mask, y_opt, xi, kappa, ctx->best_dev, mu_dev,
                             sigma_dev, acq_dev);
  if (rc) return rc;
  OB_CUDA(ctx, cudaMemcpyAsync(ctx->best_pinned, ctx->best_dev, 3 * sizeof(ob200_best), cudaMemcpyDeviceToHost, ctx->stream));
  OB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(best_host, ctx->best_pinned, 3 * sizeof(ob200_best));
  ob_l2_release(ctx);
  if (!cand_dev) return ob200_sample_status(ctx);
  return OB200_OK;
}

int ob200_score_host(ob200_ctx* ctx, const double* cand_host, int64_t ldc, int64_t M, int acq_mask, double y_opt, double xi,
                     double kappa, ob200_best* best_host, double* mu_host, double* sigma_host, double* acq_host) {
  if (!ctx) return OB200_EINVAL;
  if (!cand_host) OB_FAIL(ctx, OB200_EINVAL, "cand_host is null");
  int rc = check_score_args(ctx, M, acq_mask, best_host, y_opt, xi, kappa);
  if (rc) return rc;
  const int D = ctx->hdr.D;
  if (ldc < D) OB_FAIL(ctx, OB200_EINVAL, "ldc=%lld < D=%d", (long long)ldc, D);
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  l2_scope l2(ctx);
  // Taken here, before any copy is in flight: cudaDeviceSetLimit waits for the device to drain, and inside the first score
  // launch it used to wait for the tail's H2D copy -- the overlap below was lost (+2.9 ms per 10^6 candidates).
  if (M > 2 * (int64_t)ctx->num_sms * SC_BN) ob_l2_acquire(ctx);
  if ((rc = ob_reserve(ctx, &ctx->cand_tmp, &ctx->cand_cap, M * D))) return rc;
  const int64_t dense = (mu_host ? M : 0) + (sigma_host ? M : 0) + (acq_host ? 3 * M : 0);
  if (dense && (rc = ob_reserve(ctx, &ctx->dense_tmp, &ctx->dense_cap, dense))) return rc;
  double* p = ctx->dense_tmp;
  double* mu_dev = mu_host ? p : nullptr; if (mu_host) p += M;
  double* sg_dev = sigma_host ? p : nullptr; if (sigma_host) p += M;
  double* aq_dev = acq_host ? p : nullptr;
  const int64_t head = (int64_t)ctx->num_sms * SC_BN * 16;  // 16 tiles per CTA: its scoring outlasts the tail's H2D copy
  if (ldc == D && !dense && M >= 4 * head) {
    // Overlap: copy the head, score it while the tail travels on a second stream, then score the tail
    // into the same running argmax.
    if (!ctx->copy_stream) {
      OB_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
      OB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_main, cudaEventDisableTiming));
      OB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_copy, cudaEventDisableTiming));
    }
    OB_CUDA(ctx, cudaEventRecord(ctx->ev_main, ctx->stream));        // earlier users of cand_tmp are done
    OB_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_main, 0));
    OB_CUDA(ctx, cudaMemcpyAsync(ctx->cand_tmp, cand_host, (size_t)head * D * 8, cudaMemcpyHostToDevice, ctx->stream));
    OB_CUDA(ctx, cudaMemcpyAsync(ctx->cand_tmp + head * D, cand_host + head * D, (size_t)(M - head) * D * 8,
                                 cudaMemcpyHostToDevice, ctx->copy_stream));
    OB_CUDA(ctx, cudaEventRecord(ctx->ev_copy, ctx->copy_stream));
    int grid = 0, grid2 = 0;
    if ((rc = score_launch(ctx, ctx->cand_tmp, D, head, 0, acq_mask, y_opt, xi, kappa, nullptr, nullptr, nullptr, 0, &grid))) return rc;
    OB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_copy, 0));
    if ((rc = score_launch(ctx, ctx->cand_tmp + head * D, D, M - head, (uint64_t)head, acq_mask, y_opt, xi, kappa, nullptr, nullptr,
                           nullptr, 1, &grid2)))
      return rc;
    reduce_best_kernel<<<1, 256, 0, ctx->stream>>>(reinterpret_cast<const best_t*>(ctx->partial_dev), grid > grid2 ? grid : grid2,
                                                   acq_mask, ctx->best_dev);
    OB_LAUNCH_CHECK(ctx);
  } else {
    if (ldc == D)  // contiguous rows: one flat DMA (a 2-D copy of 8*D-byte rows is an order of magnitude slower)
      OB_CUDA(ctx, cudaMemcpyAsync(ctx->cand_tmp, cand_host, (size_t)M * D * 8, cudaMemcpyHostToDevice, ctx->stream));
    else
      OB_CUDA(ctx, cudaMemcpy2DAsync(ctx->cand_tmp, (size_t)D * 8, cand_host, (size_t)ldc * 8, (size_t)D * 8, (size_t)M,
                                     cudaMemcpyHostToDevice, ctx->stream));
    if ((rc = ob200_score_async(ctx, ctx->cand_tmp, D, M, 0, 0, acq_mask, y_opt, xi, kappa, ctx->best_dev, mu_dev, sg_dev, aq_dev))) return rc;
  }
  OB_CUDA(ctx, cudaMemcpyAsync(ctx->best_pinned, ctx->best_dev, 3 * sizeof(ob200_best), cudaMemcpyDeviceToHost, ctx->stream));
  if (mu_host) OB_CUDA(ctx, cudaMemcpyAsync(mu_host, mu_dev, (size_t)M * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (sigma_host) OB_CUDA(ctx, cudaMemcpyAsync(sigma_host, sg_dev, (size_t)M * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (acq_host) OB_CUDA(ctx, cudaMemcpyAsync(acq_host, aq_dev, (size_t)3 * M * 8, cudaMemcpyDeviceToHost, ctx->stream));
  OB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(best_host, ctx->best_pinned, 3 * sizeof(ob200_best));
  ob_l2_release(ctx);
  return OB200_OK;
}

int ob200_acq_argmax(ob200_ctx* ctx, const double* mu_dev, const double* sigma_dev, int64_t M, uint64_t first_index,
                     int acq_mask, double y_opt, double xi, double kappa, ob200_best* best_host, double* acq_dev) {
  if (!ctx) return OB200_EINVAL;
  if (!mu_dev || !sigma_dev || !best_host || M < 1) OB_FAIL(ctx, OB200_EINVAL, "bad mu/sigma/best/M");
  if ((acq_mask & 7) == 0 || (acq_mask & ~7)) OB_FAIL(ctx, OB200_EINVAL, "acq_mask must select EI(1), PI(2) and/or LCB(4)");
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc;
  if ((rc = ob_reserve(ctx, &ctx->partial_dev, &ctx->partial_cap, (int64_t)4096 * 3 * (int64_t)sizeof(best_t)))) return rc;
  int64_t blocks = (M + 255) / 256;
  const int grid = (int)(blocks < (int64_t)ctx->num_sms * 8 ? blocks : (int64_t)ctx->num_sms * 8);
  acq_partial_kernel<<<grid, 256, 0, ctx->stream>>>(mu_dev, sigma_dev, M, first_index, acq_mask, y_opt, xi, kappa, acq_dev,
                                                    reinterpret_cast<best_t*>(ctx->partial_dev));
  OB_LAUNCH_CHECK(ctx);
  reduce_best_kernel<<<1, 256, 0, ctx->stream>>>(reinterpret_cast<const best_t*>(ctx->partial_dev), grid, acq_mask, ctx->best_dev);
  OB_LAUNCH_CHECK(ctx);
  OB_CUDA(ctx, cudaMemcpyAsync(ctx->best_pinned, ctx->best_dev, 3 * sizeof(ob200_best), cudaMemcpyDeviceToHost, ctx->stream));
  OB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(best_host, ctx->best_pinned, 3 * sizeof(ob200_best));
  return OB200_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Exact pruned scoring.  EI, PI (for y_opt - xi - mu <= 0) and -LCB are increasing in sigma and
// decreasing in mu, and sigma^2 <= (amplitude - k_max^2 / (amplitude + diag)) * y_std^2, the variance left by the
// nearest observation alone (k_max = largest kernel value of the candidate against the observations; at worst the
// prior variance), so U(mu_lb, sigma_ub) is an upper bound of the utility that needs only the O(nD) cross-kernel.
//
//   pass 1   mu_lb, sigma_ub for every candidate: the FP32 / TF32 bound kernel (ob200_bound.cu, rigorous error
//            budget), or -- when that kernel cannot take the fit -- phase 1 of the FP64 score kernel;
//   batch A  the candidates with the largest bounds (at most one tile per CTA; the union over the requested
//            acquisitions) are scored in full -> incumbents;
//   batch B  every other candidate whose bound still reaches an incumbent; when pass 1 was FP32 their bounds are
//            first recomputed by the FP64 mean-only kernel (9 ns per candidate instead of 125) and only the ones that
//            still reach the incumbent are scored.
// Every candidate left out has utility <= bound < incumbent <= maximum, so the argmax (ties included) and the
// winner's values are bit-identical to the dense call.  The whole selection is device-resident: thresholds come
// from a two-level radix histogram of the bounds (the last block to finish a histogram kernel turns it into the
// threshold), batch sizes live in device memory and the score launches read them from there (persistent grid, idle
// CTAs fall through), so a call has ONE host synchronisation: the read-back of the winners.
namespace {

#define PR_BINS 4096   // 12 bits of the order-preserving key per histogram level

struct prune_ctl {
  unsigned long long count_a, count_b, count_b2, count_c;   // rows in batch A / B (pass-1 bounds) / B after the FP64 refinement / B after the subset bound
  unsigned long long tau_key[3];                   // batch A of acquisition k = {key(ub_k) >= tau_key[k]}
  long long rem[3];                                // quota left for level 2
  int b1[3];                                       // level-1 bin that holds the threshold
  unsigned int ticket1, ticket2;
  unsigned int hist1[3][PR_BINS];
  unsigned int hist2[3][PR_BINS];
};

// order-preserving map double -> uint64 (NaN -> largest key: a NaN bound is always selected)
__device__ __forceinline__ unsigned long long ord_key(double v) {
  if (!(v == v)) return ~0ull;
  const unsigned long long u = (unsigned long long)__double_as_longlong(v);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}

__device__ __forceinline__ double upper_bound_utility(int acq, double mu_lb, double sigma_ub, double y_opt, double xi, double kappa) {
  double u[3];
  acquisition(mu_lb, sigma_ub, y_opt, xi, kappa, u);
  if (acq == OB200_ACQ_PI && (y_opt - xi - mu_lb) > 0.0) return 1.0;  // PI decreases in sigma when the mean already improves
  return u[acq];
}

__device__ __forceinline__ void block_hist_flush(unsigned int* sh, unsigned int* __restrict__ gl) {
  __syncthreads();
  for (int b = threadIdx.x; b < PR_BINS; b += blockDim.x) { const unsigned int c = sh[b]; if (c) atomicAdd(&gl[b], c); sh[b] = 0; }
  __syncthreads();
}

// true in exactly one block: the last one to get here (all of its predecessors' global writes are visible)
__device__ __forceinline__ bool last_block(unsigned int* ticket) {
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (is_last) __threadfence();
  return is_last;
}

// From the top bin downwards: the largest set of whole bins holding at most `quota` entries.  Returns the lowest
// bin taken (PR_BINS if none) and the number of entries taken; *next = the bin below (the one that did not fit).
__device__ void scan_from_top(const unsigned int* __restrict__ hist, long long quota, bool force_first, int* lowest, long long* taken,
                              unsigned int* sh) {
  // 256 threads: chunk sums of 16 bins, then a serial walk over 256 + 16 values
  unsigned int s = 0;
  for (int b = 0; b < PR_BINS / 256; ++b) s += __ldcg(&hist[threadIdx.x * (PR_BINS / 256) + b]);
  sh[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long cum = 0;
    int low = PR_BINS;
    int c = 255;
    for (; c >= 0; --c) { if (cum + sh[c] > quota) break; cum += sh[c]; low = c * (PR_BINS / 256); }
    if (c >= 0)
      for (int b = c * (PR_BINS / 256) + PR_BINS / 256 - 1; b >= c * (PR_BINS / 256); --b) {
        const unsigned int h = __ldcg(&hist[b]);
        if (cum + h > quota && !(force_first && cum == 0 && h > 0)) break;   // force_first: never return an empty set
        cum += h; low = b;
        if (cum > quota) break;
      }
    *lowest = low; *taken = cum;
  }
  __syncthreads();
}

// utility bounds of the requested acquisitions + level-1 histogram (top 12 key bits); the last block picks, per
// acquisition, the level-1 bin that holds the threshold
__global__ void __launch_bounds__(256) ub_hist1_kernel(const double* __restrict__ mu, const double* __restrict__ sig, long long M, int acq_mask,
                                                       double y_opt, double xi, double kappa, double* __restrict__ ub3, long long quota,
                                                       prune_ctl* __restrict__ ctl) {
  __shared__ unsigned int sh[PR_BINS];
  for (int b = threadIdx.x; b < PR_BINS; b += 256) sh[b] = 0;
  __syncthreads();
  for (int k = 0; k < 3; ++k) {
    if (!(acq_mask & (1 << k))) continue;
    double* ub = ub3 + (size_t)k * M;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < M; i += (long long)gridDim.x * 256) {
      const double v = upper_bound_utility(k, mu[i], sig[i], y_opt, xi, kappa);
      ub[i] = v;
      atomicAdd(&sh[ord_key(v) >> 52], 1u);
    }
    block_hist_flush(sh, ctl->hist1[k]);
  }
  if (!last_block(&ctl->ticket1)) return;
  __shared__ int low;
  __shared__ long long taken;
  for (int k = 0; k < 3; ++k) {
    if (!(acq_mask & (1 << k))) continue;
    scan_from_top(ctl->hist1[k], quota, false, &low, &taken, sh);
    if (threadIdx.x == 0) { ctl->b1[k] = low - 1; ctl->rem[k] = quota - taken; }   // low - 1: the bin that did not fit (-1: all fit)
    __syncthreads();
  }
}

// level 2: the next 12 key bits inside the level-1 bin; the last block fixes tau_key
__global__ void __launch_bounds__(256) hist2_kernel(const double* __restrict__ ub3, long long M, int acq_mask, long long quota,
                                                    prune_ctl* __restrict__ ctl) {
  __shared__ unsigned int sh[PR_BINS];
  for (int b = threadIdx.x; b < PR_BINS; b += 256) sh[b] = 0;
  __syncthreads();
  for (int k = 0; k < 3; ++k) {
    if (!(acq_mask & (1 << k))) continue;
    const int b1 = ctl->b1[k];
    if (b1 >= 0) {
      const double* ub = ub3 + (size_t)k * M;
      for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < M; i += (long long)gridDim.x * 256) {
        const unsigned long long key = ord_key(ub[i]);
        if ((int)(key >> 52) == b1) atomicAdd(&sh[(key >> 40) & (PR_BINS - 1)], 1u);
      }
    }
    block_hist_flush(sh, ctl->hist2[k]);
  }
  if (!last_block(&ctl->ticket2)) return;
  __shared__ int low;
  __shared__ long long taken;
  for (int k = 0; k < 3; ++k) {
    if (!(acq_mask & (1 << k))) continue;
    const int b1 = ctl->b1[k];
    if (b1 < 0) { if (threadIdx.x == 0) ctl->tau_key[k] = 0ull; __syncthreads(); continue; }   // everything fits the quota
    scan_from_top(ctl->hist2[k], ctl->rem[k], ctl->rem[k] == quota, &low, &taken, sh);   // nothing taken above: batch A must not be empty
    if (threadIdx.x == 0) {
      unsigned long long key;
      if (low < PR_BINS) key = ((unsigned long long)b1 << 52) | ((unsigned long long)low << 40);
      else if (b1 + 1 < PR_BINS) key = (unsigned long long)(b1 + 1) << 52;   // nothing of this bin fits: only the bins above
      else key = ~0ull;
      ctl->tau_key[k] = key;
    }
    __syncthreads();
  }
}

// batch A (phase 0): rows with key(ub_k) >= tau_key[k] for some requested k.
// batch B (phase 1): rows NOT in batch A whose bound reaches the incumbent of some requested k.
__global__ void __launch_bounds__(256) compact_kernel(const double* __restrict__ ub3, long long M, int acq_mask, int phase,
                                                      const ob200_best* __restrict__ incumbent, prune_ctl* __restrict__ ctl,
                                                      long long* __restrict__ rows) {
  unsigned long long tau[3];
  double thr[3];
  for (int k = 0; k < 3; ++k) {
    tau[k] = ctl->tau_key[k];
    const double v = incumbent ? incumbent[k].value : 0.0;
    thr[k] = v - 1e-12 * fabs(v) - 1e-300;   // rounding margin between the bound's and the score kernel's arithmetic
  }
  unsigned long long* counter = phase == 0 ? &ctl->count_a : &ctl->count_b;
  for (long long i0 = (long long)blockIdx.x * 256; i0 < M; i0 += (long long)gridDim.x * 256) {
    const long long i = i0 + threadIdx.x;
    bool in_a = false, reach = false;
    if (i < M) {
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        if (!(acq_mask & (1 << k))) continue;
        const double v = ub3[(size_t)k * M + i];
        in_a |= ord_key(v) >= tau[k];
        reach |= v >= thr[k];
      }
    }
    const bool keep = phase == 0 ? in_a : (!in_a && reach);
    const unsigned mask = __ballot_sync(0xffffffffu, keep);
    if (mask) {
      const int lane = threadIdx.x & 31, leader = __ffs(mask) - 1;
      unsigned long long base = 0;
      if (lane == leader) base = atomicAdd(counter, (unsigned long long)__popc(mask));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (keep) rows[base + __popc(mask & ((1u << lane) - 1))] = i;
    }
  }
}

// batch B after the FP64 refinement: mu / sigma_ub of the count_b rows were recomputed exactly (local arrays); keep
// the rows whose refined bound still reaches an incumbent
__global__ void __launch_bounds__(256) compact_refined_kernel(const double* __restrict__ mu, const double* __restrict__ sig, int acq_mask,
                                                              double y_opt, double xi, double kappa, const ob200_best* __restrict__ incumbent,
                                                              prune_ctl* __restrict__ ctl, const long long* __restrict__ rows_in,
                                                              long long* __restrict__ rows_out, double* __restrict__ mu_out,
                                                              double* __restrict__ sig_out) {
  const long long n = (long long)ctl->count_b;
  double thr[3];
  for (int k = 0; k < 3; ++k) { const double v = incumbent[k].value; thr[k] = v - 1e-12 * fabs(v) - 1e-300; }
  for (long long i0 = (long long)blockIdx.x * 256; i0 < n; i0 += (long long)gridDim.x * 256) {
    const long long i = i0 + threadIdx.x;
    bool keep = false;
    double m = 0.0, s = 0.0;
    if (i < n) {
      m = mu[i]; s = sig[i];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        if (!(acq_mask & (1 << k))) continue;
        const double v = upper_bound_utility(k, m, s, y_opt, xi, kappa);
        keep |= !(v == v) || v >= thr[k];
      }
    }
    const unsigned mask = __ballot_sync(0xffffffffu, keep);
    if (mask) {
      const int lane = threadIdx.x & 31, leader = __ffs(mask) - 1;
      unsigned long long base = 0;
      if (lane == leader) base = atomicAdd(&ctl->count_b2, (unsigned long long)__popc(mask));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (keep) {   // the refined mean / sigma bound travel with the row (the subset-GP step reads them by position)
        const unsigned long long pos = base + __popc(mask & ((1u << lane) - 1));
        rows_out[pos] = rows_in[i];
        mu_out[pos] = m;
        sig_out[pos] = s;
      }
    }
  }
}

// Batch B with a second variance bound: sigma_s[r] = posterior std (normalised units) of row rows_in[r] under a GP on
// a SUBSET of the observations (conditioning on fewer points can only raise the variance).  mu / sig are the mean and
// the nearest-observation bound of the same rows: indexed by r (`local` = 1: the arrays of the FP64 refinement) or by
// rows_in[r] (`local` = 0: the arrays of an FP64 pass 1).
__global__ void __launch_bounds__(256) compact_subset_kernel(const double* __restrict__ mu, const double* __restrict__ sig,
                                                             const double* __restrict__ sigma_s, int local, double y_std, double slack,
                                                             int acq_mask, double y_opt, double xi, double kappa,
                                                             const ob200_best* __restrict__ incumbent, const unsigned long long* __restrict__ n_in,
                                                             unsigned long long* __restrict__ n_out, const long long* __restrict__ rows_in,
                                                             long long* __restrict__ rows_out) {
  const long long n = (long long)*n_in;
  double thr[3];
  for (int k = 0; k < 3; ++k) { const double v = incumbent[k].value; thr[k] = v - 1e-12 * fabs(v) - 1e-300; }
  for (long long i0 = (long long)blockIdx.x * 256; i0 < n; i0 += (long long)gridDim.x * 256) {
    const long long i = i0 + threadIdx.x;
    bool keep = false;
    if (i < n) {
      const long long src = local ? i : rows_in[i];
      const double m = mu[src];
      const double s = fmin(sig[src], y_std * sigma_s[i] + slack);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        if (!(acq_mask & (1 << k))) continue;
        const double v = upper_bound_utility(k, m, s, y_opt, xi, kappa);
        keep |= !(v == v) || v >= thr[k];
      }
    }
    const unsigned mask = __ballot_sync(0xffffffffu, keep);
    if (mask) {
      const int lane = threadIdx.x & 31, leader = __ffs(mask) - 1;
      unsigned long long base = 0;
      if (lane == leader) base = atomicAdd(n_out, (unsigned long long)__popc(mask));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (keep) rows_out[base + __popc(mask & ((1u << lane) - 1))] = rows_in[i];
    }
  }
}

// observation i of the subset = observation floor(i n / m) of the fit, un-scaled back from the blob's packed X / ls
__global__ void __launch_bounds__(256) gather_subset_kernel(const double* __restrict__ xs, const double* __restrict__ ls, int n, int D,
                                                            int D_pad, int m, double* __restrict__ out) {
  const int nd4 = D_pad >> 2;
  for (int e = blockIdx.x * 256 + threadIdx.x; e < m * D; e += gridDim.x * 256) {
    const int i = e / D, d = e - i * D;
    const int obs = (int)((long long)i * n / m);
    out[e] = xs[((size_t)(obs >> 3) * nd4 + (d >> 2)) * 32 + (obs & 7) * 4 + (d & 3)] * ls[d];
  }
}

#define PR_SUBSET_M SC_BM   // one row block of the score kernel (192 observations)

// GP with the fit's kernel on PR_SUBSET_M evenly spaced observations, in a context of its own (ctx->sub); built once
// per fit, when a pruned call first needs the second variance bound.
int ensure_subset_fit(ob200_ctx* ctx) {
  if (ctx->sub && ctx->sub_serial == ctx->fit_serial) return OB200_OK;
  const ob200_fit_header& h = ctx->hdr;
  const ob200_blob_layout lay = ob_blob_layout(h.n_pad, h.D_pad);
  if (!ctx->sub) {
    const int rc = ob200_ctx_create(ctx->device, &ctx->sub);
    if (rc) OB_FAIL(ctx, rc, "cannot create the context of the subset GP");
  }
  ob200_ctx* sub = ctx->sub;
  sub->stream = ctx->stream;
  const int m = h.n < PR_SUBSET_M ? h.n : PR_SUBSET_M, D = h.D;
  int rc;
  if ((rc = ob_reserve(ctx, &ctx->gather_dev, &ctx->gather_cap, (int64_t)m * D + m))) return rc;
  std::vector<double> ls((size_t)h.D_pad);
  OB_CUDA(ctx, cudaMemcpyAsync(ls.data(), ctx->blob + lay.off_ls, (size_t)h.D_pad * 8, cudaMemcpyDeviceToHost, ctx->stream));
  gather_subset_kernel<<<ctx->num_sms, 256, 0, ctx->stream>>>(reinterpret_cast<const double*>(ctx->blob + lay.off_xs),
                                                              reinterpret_cast<const double*>(ctx->blob + lay.off_ls), h.n, D, h.D_pad, m,
                                                              ctx->gather_dev);
  OB_LAUNCH_CHECK(ctx);
  OB_CUDA(ctx, cudaMemsetAsync(ctx->gather_dev + (size_t)m * D, 0, (size_t)m * 8, ctx->stream));   // y = 0: only the variance is used
  OB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ob200_kernel kern{};
  kern.kind = h.kernel_kind; kern.n_length_scale = D; kern.amplitude = h.amplitude; kern.length_scale_host = ls.data();
  rc = ob200_gp_fit(sub, ctx->gather_dev, ctx->gather_dev + (size_t)m * D, m, D, &kern, h.diag, 0);
  if (rc) OB_FAIL(ctx, rc, "subset GP: %s", sub->err.c_str());
  ctx->sub_serial = ctx->fit_serial;
  return OB200_OK;
}

}  // namespace

extern "C" int ob200_score_pruned_multi(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, uint64_t seed,
                                        uint64_t first_index, int acq_mask, double y_opt, double xi, double kappa,
                                        ob200_best* best_host, int64_t* n_scored_host) {
  if (!ctx) return OB200_EINVAL;
  int rc = check_score_args(ctx, M, acq_mask, best_host, y_opt, xi, kappa);
  if (rc) return rc;
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  const int D = ctx->hdr.D;
  const int64_t batch = (int64_t)ctx->num_sms * SC_BN;   // one tile per CTA
  for (int k = 0; k < 3; ++k) { best_host[k] = ob200_best{}; if (n_scored_host) n_scored_host[k] = 0; }
  // Dense path: small problems; candidate sets that do not fit in one resident buffer; -LCB with kappa < 0, which
  // DEcreases in sigma, so U(mu_lb, sigma_ub) would not bound it.
  const bool too_big = !cand_dev && (M * (int64_t)D * 8 > ((int64_t)1 << 31));
  const bool lcb_not_monotone = (acq_mask & (1 << OB200_ACQ_LCB)) && kappa < 0.0;
  // Up to one tile per CTA the dense kernel is one wave and nothing can beat it.  Beyond, the pruned path costs one
  // wave + ~0.10 ms (0.14 ms for a portfolio) whatever M, the dense kernel one wave per `batch` candidates, and a wave
  // costs ~125 ns per candidate at n_pad = 2112, growing with n_pad^2 (tools/small_m_perf.py, B200): dense while its
  // extra waves are cheaper than that overhead -- 1 wave from n ~ 1000 (n = 2000: 0.80 ms flat for 6 000 ... 50 000
  // candidates against 1.3 / 1.8 / 2.4 / 6.6 ms dense at 6 000 / 10 000 / 15 000 / 50 000; the plugin's default
  // n_points is 10 000), 2 at n = 700, 4 at n = 300, 17 at n = 50.
  int n_acq_req = 0;
  for (int k = 0; k < 3; ++k) n_acq_req += (acq_mask >> k) & 1;
  const double wave_ns = (double)batch * (125.0 * (ctx->hdr.n_pad / 2112.0) * (ctx->hdr.n_pad / 2112.0) + 0.3);
  int64_t dense_upto = batch * (1 + (int64_t)((n_acq_req > 1 ? 1.4e5 : 1.0e5) / wave_ns));
  if (const char* env = getenv("OB200_PRUNE_MIN_BATCHES")) { const int v = atoi(env); if (v >= 1 && v <= 64) dense_upto = v * batch; }   // (tuning hook)
  if (M <= dense_upto || too_big || lcb_not_monotone) {
    rc = ob200_score(ctx, cand_dev, ldc, M, seed, first_index, acq_mask, y_opt, xi, kappa, best_host, nullptr, nullptr, nullptr);
    if (rc) return rc;
    for (int k = 0; k < 3; ++k) if (n_scored_host && (acq_mask & (1 << k))) n_scored_host[k] = M;
    return OB200_OK;
  }
  l2_scope l2(ctx);
  if (cand_dev) {
    if (ldc < D) OB_FAIL(ctx, OB200_EINVAL, "ldc=%lld < D=%d", (long long)ldc, D);
  } else {
    if (ctx->n_slots <= 0) OB_FAIL(ctx, OB200_ESTATE, "cand_dev is NULL and no space was set");
    if (ctx->d_out != D) OB_FAIL(ctx, OB200_EINVAL, "space width %d != fitted D %d", ctx->d_out, D);
    if ((rc = ob_reserve(ctx, &ctx->cand_tmp, &ctx->cand_cap, M * D))) return rc;
    if ((rc = ob200_sample_launch(ctx, seed, first_index, M, 1, ctx->cand_tmp, D, nullptr))) return rc;
    cand_dev = ctx->cand_tmp;
    ldc = D;
  }
  if ((rc = ob_reserve(ctx, &ctx->ub_dev, &ctx->ub_cap, 7 * M))) return rc;
  if ((rc = ob_reserve(ctx, &ctx->idx_dev, &ctx->idx_cap, 2 * M))) return rc;
  if ((rc = ob_reserve(ctx, &ctx->prune_small_dev, &ctx->prune_small_cap, (int64_t)sizeof(prune_ctl)))) return rc;
  if (!ctx->ctl_pinned) OB_CUDA(ctx, cudaMallocHost((void**)&ctx->ctl_pinned, 64));
  double *mu = ctx->ub_dev, *sig = ctx->ub_dev + M, *ub3 = ctx->ub_dev + 2 * M, *mu_b = ctx->ub_dev + 5 * M, *sig_b = ctx->ub_dev + 6 * M;
  long long *rows0 = ctx->idx_dev, *rows1 = ctx->idx_dev + M;
  prune_ctl* ctl = reinterpret_cast<prune_ctl*>(ctx->prune_small_dev);
  cudaStream_t st = ctx->stream;
  const int nblk = ctx->num_sms * 8;
  int n_acq = 0, grid = 0;
  for (int k = 0; k < 3; ++k) n_acq += (acq_mask >> k) & 1;
  OB_CUDA(ctx, cudaMemsetAsync(ctl, 0, sizeof(prune_ctl), st));
  // pass 1 (shared by the acquisitions)
  int f32_done = 0;
  if ((rc = ob200_bound_f32_launch(ctx, cand_dev, ldc, M, mu, sig, &f32_done))) return rc;
  if (!f32_done && (rc = score_launch(ctx, cand_dev, ldc, M, first_index, acq_mask, y_opt, xi, kappa, mu, sig, nullptr, 0, &grid, 1))) return rc;
  // thresholds of the first batches
  ub_hist1_kernel<<<nblk, 256, 0, st>>>(mu, sig, M, acq_mask, y_opt, xi, kappa, ub3, batch / n_acq, ctl);
  OB_LAUNCH_CHECK(ctx);
  hist2_kernel<<<nblk, 256, 0, st>>>(ub3, M, acq_mask, batch / n_acq, ctl);
  OB_LAUNCH_CHECK(ctx);
  // batch A
  compact_kernel<<<nblk, 256, 0, st>>>(ub3, M, acq_mask, 0, nullptr, ctl, rows0);
  OB_LAUNCH_CHECK(ctx);
  if ((rc = score_launch(ctx, cand_dev, ldc, M, first_index, acq_mask, y_opt, xi, kappa, nullptr, nullptr, nullptr, 0, &grid, 0, rows0,
                         reinterpret_cast<const long long*>(&ctl->count_a)))) return rc;
  reduce_best_kernel<<<1, 256, 0, st>>>(reinterpret_cast<const best_t*>(ctx->partial_dev), grid, acq_mask, ctx->best_dev);
  OB_LAUNCH_CHECK(ctx);
  // batch B
  compact_kernel<<<nblk, 256, 0, st>>>(ub3, M, acq_mask, 1, ctx->best_dev, ctl, rows1);
  OB_LAUNCH_CHECK(ctx);
  const long long* rows_b = rows1;
  const unsigned long long* count_b = &ctl->count_b;
  if (f32_done) {  // exact bounds for the rows the FP32 bounds could not exclude
    int g2 = 0;
    if ((rc = score_launch(ctx, cand_dev, ldc, M, first_index, acq_mask, y_opt, xi, kappa, mu_b, sig_b, nullptr, 0, &g2, 1, rows1,
                           reinterpret_cast<const long long*>(&ctl->count_b)))) return rc;
    compact_refined_kernel<<<nblk, 256, 0, st>>>(mu_b, sig_b, acq_mask, y_opt, xi, kappa, ctx->best_dev, ctl, rows1, rows0, ub3, ub3 + M);
    OB_LAUNCH_CHECK(ctx);
    rows_b = rows0;
    count_b = &ctl->count_b2;
  }
  // Escalation: when many rows are still standing, the nearest-observation variance bound is the weak part (long
  // length scales: every kernel value is close to the amplitude).  A GP on a subset of the observations bounds the
  // variance from above as well, at (192 / n)^2 of the cost of the full posterior.
  long long* rows_c = (rows_b == rows0) ? rows1 : rows0;
  OB_CUDA(ctx, cudaMemcpyAsync(ctx->ctl_pinned, ctl, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
  OB_CUDA(ctx, cudaStreamSynchronize(st));
  {
    const unsigned long long* c = reinterpret_cast<const unsigned long long*>(ctx->ctl_pinned);
    const int64_t standing = (int64_t)(f32_done ? c[2] : c[1]);
    int64_t escalate_at = 4 * batch;
    if (const char* env = getenv("OB200_PRUNE_SUBSET_AT")) escalate_at = atoll(env);   // (testing hook; < 0 disables)
    ctx->last_prune_counts[3] = -1;
    if (escalate_at >= 0 && standing > escalate_at && ctx->hdr.n > PR_SUBSET_M) {
      if ((rc = ensure_subset_fit(ctx))) return rc;
      ob200_ctx* sub = ctx->sub;
      sub->stream = ctx->stream;   // (the caller may have moved the context to another stream since the subset fit was made)
      l2_scope l2s(sub);
      int gs = 0;
      double* sig_s = ub3 + 2 * M;   // (the utility bounds of pass 1 are not needed any more: ub3 = refined mu | sigma | this)
      if ((rc = score_launch(sub, cand_dev, ldc, M, first_index, 1, 0.0, 0.0, 0.0, nullptr, sig_s, nullptr, 0, &gs, 0, rows_b,
                             reinterpret_cast<const long long*>(count_b)))) OB_FAIL(ctx, rc, "subset GP: %s", sub->err.c_str());
      ctx->launches += 1;
      unsigned long long* count_c = &ctl->count_c;
      // slack: the rounding of the subset posterior (cancellation ~ cond(K_S) eps), as for the nearest-observation bound
      compact_subset_kernel<<<nblk, 256, 0, st>>>(f32_done ? ub3 : mu, f32_done ? ub3 + M : sig, sig_s, f32_done ? 1 : 0, ctx->hdr.y_std,
                                                  ctx->hdr.y_std * sqrt(1e-6 * ctx->hdr.amplitude), acq_mask, y_opt, xi, kappa,
                                                  ctx->best_dev, count_b, count_c, rows_b, rows_c);
      OB_LAUNCH_CHECK(ctx);
      rows_b = rows_c;
      count_b = count_c;
      ctx->last_prune_counts[3] = 0;
    }
  }
  int grid_b = 0;
  if ((rc = score_launch(ctx, cand_dev, ldc, M, first_index, acq_mask, y_opt, xi, kappa, nullptr, nullptr, nullptr, 1, &grid_b, 0, rows_b,
                         reinterpret_cast<const long long*>(count_b)))) return rc;
  reduce_best_kernel<<<1, 256, 0, st>>>(reinterpret_cast<const best_t*>(ctx->partial_dev), grid, acq_mask, ctx->best_dev);
  OB_LAUNCH_CHECK(ctx);
  OB_CUDA(ctx, cudaMemcpyAsync(ctx->best_pinned, ctx->best_dev, 3 * sizeof(ob200_best), cudaMemcpyDeviceToHost, st));
  OB_CUDA(ctx, cudaMemcpyAsync(ctx->ctl_pinned, ctl, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
  OB_CUDA(ctx, cudaStreamSynchronize(st));
  const unsigned long long* counts = reinterpret_cast<const unsigned long long*>(ctx->ctl_pinned);
  const bool escalated = ctx->last_prune_counts[3] >= 0;
  const int64_t n_scored = (int64_t)counts[0] + (int64_t)(escalated ? counts[3] : (f32_done ? counts[2] : counts[1]));
  ctx->last_prune_f32 = f32_done;
  ctx->last_prune_counts[0] = (int64_t)counts[0]; ctx->last_prune_counts[1] = (int64_t)counts[1]; ctx->last_prune_counts[2] = (int64_t)counts[2];
  if (escalated) ctx->last_prune_counts[3] = (int64_t)counts[3];
  for (int k = 0; k < 3; ++k) {
    if (!(acq_mask & (1 << k))) continue;
    best_host[k] = ctx->best_pinned[k];
    if (n_scored_host) n_scored_host[k] = n_scored;
  }
  if (cand_dev == ctx->cand_tmp) return ob200_sample_status(ctx);
  return OB200_OK;
}

extern "C" int ob200_score_pruned(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, uint64_t seed, uint64_t first_index,
                                  int acq, double y_opt, double xi, double kappa, ob200_best* best_host, int64_t* n_scored_host) {
  if (!ctx) return OB200_EINVAL;
  if (acq < 0 || acq > 2) OB_FAIL(ctx, OB200_EINVAL, "acq must be OB200_ACQ_EI, _PI or _LCB");
  if (!best_host) OB_FAIL(ctx, OB200_EINVAL, "best_host is null");
  ob200_best best[3];
  int64_t ns[3];
  const int rc = ob200_score_pruned_multi(ctx, cand_dev, ldc, M, seed, first_index, 1 << acq, y_opt, xi, kappa, best, ns);
  if (rc) return rc;
  best_host[0] = best[acq];
  if (n_scored_host) *n_scored_host = ns[acq];
  return OB200_OK;
}

// Pass 1 of the pruned path on its own (tests, tools): mu_lb[i] <= posterior mean and sigma_ub[i] >= posterior std of
// candidate i, from the FP32 bound kernel (*used_f32 = 1) or the FP64 mean-only kernel (*used_f32 = 0, mu_lb exact).
extern "C" int ob200_score_bounds(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, double* mu_lb_dev, double* sigma_ub_dev,
                                  int* used_f32) {
  if (!ctx) return OB200_EINVAL;
  if (!ctx->fitted) OB_FAIL(ctx, OB200_ESTATE, "ob200_score_bounds called before ob200_gp_fit / ob200_gp_import");
  if (!cand_dev || !mu_lb_dev || !sigma_ub_dev || M < 1 || ldc < ctx->hdr.D) OB_FAIL(ctx, OB200_EINVAL, "bad cand/mu/sigma/M/ldc");
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  int done = 0, grid = 0, rc;
  if ((rc = ob200_bound_f32_launch(ctx, cand_dev, ldc, M, mu_lb_dev, sigma_ub_dev, &done))) return rc;
  if (!done && (rc = score_launch(ctx, cand_dev, ldc, M, 0, 1, 0.0, 0.0, 0.0, mu_lb_dev, sigma_ub_dev, nullptr, 0, &grid, 1))) return rc;
  if (used_f32) *used_f32 = done;
  return OB200_OK;
}

// Counters of the last ob200_score_pruned[_multi] call that took the pruned path: used_f32, rows in batch A, rows the
// pass-1 bounds left for batch B, rows left after the FP64 refinement of those bounds, rows left after the subset-GP
// variance bound (-1: that step did not run).
extern "C" int ob200_prune_stats(const ob200_ctx* ctx, int64_t out[5]) {
  if (!ctx || !out) return OB200_EINVAL;
  out[0] = ctx->last_prune_f32; out[1] = ctx->last_prune_counts[0]; out[2] = ctx->last_prune_counts[1]; out[3] = ctx->last_prune_counts[2];
  out[4] = ctx->last_prune_counts[3];
  return OB200_OK;
}

// ---------------------------------------------------------------------------------------------
// Allocation-free acquisition + argmax.
//
// Some hosts of this library may not allocate device memory when they call into it: a real-time
// or latency-critical service that reserves all of its GPU memory at start-up and treats any
// later cudaMalloc as a fault (it can stall on the driver's allocator lock or fail under a full
// memory pool), and code that runs in a phase where allocation is simply forbidden, e.g. between
// a framework's "memory frozen" point and shutdown.  ob200_acq_argmax needs a context (whose
// creation allocates) and grows a scratch buffer for the per-block winners on first use; this
// variant needs neither.  Its scratch is the fixed array below, which lives in the library's
// device image: every GPU that loads libob200.so holds these 384 KiB until the process exits.
//
// The array's name is only meaningful to device code.  On the host it is a placeholder symbol,
// not a device address, so its address is looked up once per device with cudaGetSymbolAddress
// (with that device current) and the looked-up pointer is what the kernels receive.
#define OB200_STATIC_ITEMS_PER_BLOCK 4096                 /* 256 threads x 16 candidates */
#define OB200_STATIC_MAX_BLOCKS 4096
#define OB200_STATIC_MAX_M ((int64_t)OB200_STATIC_ITEMS_PER_BLOCK * OB200_STATIC_MAX_BLOCKS) /* 2^24 */
#define OB200_STATIC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_static_partials[OB200_STATIC_MAX_BLOCKS * 3 * sizeof(best_t)];

namespace {
best_t* g_static_addr[OB200_STATIC_MAX_DEVICES];  // per-device address of g_static_partials (0 = not looked up yet)

void static_fail(char* errbuf, int errbuf_len, const char* msg) {
  if (errbuf && errbuf_len > 0) snprintf(errbuf, (size_t)errbuf_len, "%s", msg);
}
}  // namespace

extern "C" int ob200_acq_argmax_static(const double* mu_dev, const double* sigma_dev, int64_t M, uint64_t first_index,
                                       int acq_mask, double y_opt, double xi, double kappa, ob200_best* best_dev,
                                       double* acq_dev, void* cuda_stream, char* errbuf, int errbuf_len) {
  char msg[256];
  if (!mu_dev || !sigma_dev || !best_dev || M < 1) {
    static_fail(errbuf, errbuf_len, "ob200_acq_argmax_static: mu_dev, sigma_dev and best_dev must be non-null and M >= 1");
    return OB200_EINVAL;
  }
  if ((acq_mask & 7) == 0 || (acq_mask & ~7)) {
    static_fail(errbuf, errbuf_len, "ob200_acq_argmax_static: acq_mask must select EI(1), PI(2) and/or LCB(4)");
    return OB200_EINVAL;
  }
  if (M > OB200_STATIC_MAX_M) {
    snprintf(msg, sizeof msg,
             "ob200_acq_argmax_static: M=%lld exceeds the fixed scratch (at most %lld candidates per call); "
             "split the call or use ob200_acq_argmax", (long long)M, (long long)OB200_STATIC_MAX_M);
    static_fail(errbuf, errbuf_len, msg);
    return OB200_EINVAL;
  }
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= OB200_STATIC_MAX_DEVICES) {
    static_fail(errbuf, errbuf_len, "ob200_acq_argmax_static: cannot identify the current CUDA device");
    return OB200_ECUDA;
  }
  best_t* scratch = g_static_addr[dev];
  if (!scratch) {  // first call on this device: resolve the symbol (do this before any stream capture)
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_partials);
    if (e != cudaSuccess || !p) {
      snprintf(msg, sizeof msg, "ob200_acq_argmax_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      static_fail(errbuf, errbuf_len, msg);
      return OB200_ECUDA;
    }
    scratch = static_cast<best_t*>(p);
    g_static_addr[dev] = scratch;
  }
  cudaStream_t st = (cudaStream_t)cuda_stream;
  const int grid = (int)((M + OB200_STATIC_ITEMS_PER_BLOCK - 1) / OB200_STATIC_ITEMS_PER_BLOCK);
  // Nothing in the scratch has to start at zero: each of the `grid` blocks overwrites its own three
  // records, and the reduction reads exactly those `grid` records, so whatever an earlier (possibly
  // larger) call left behind is never read.
  acq_partial_kernel<<<grid, 256, 0, st>>>(mu_dev, sigma_dev, M, first_index, acq_mask, y_opt, xi, kappa, acq_dev, scratch);
  e = cudaGetLastError();
  if (e == cudaSuccess) {
    reduce_best_kernel<<<1, 256, 0, st>>>(scratch, grid, acq_mask, best_dev);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) {
    snprintf(msg, sizeof msg, "ob200_acq_argmax_static: kernel launch failed: %s", cudaGetErrorString(e));
    static_fail(errbuf, errbuf_len, msg);
    return OB200_ECUDA;
  }
  return OB200_OK;
}
