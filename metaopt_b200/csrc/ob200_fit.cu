// a13: GP fit on the device -- y statistics, Gram matrix, blocked FP64 Cholesky, inv(L), alpha.
//
// Replaces GaussianProcessRegressor.fit with fixed hyper-parameters
// (sklearn/gaussian_process/_gpr.py:233-368): y normalisation :275-280 (population std,
// _handle_zeros_in_scale), K = kernel(X) :349 (Matern nu=2.5 kernels.py:1720-1729 via pdist-style
// direct differences, RBF :1569-1570), K[diag] += alpha :350, L = cholesky(K, lower) :352,
// alpha_ = cho_solve((L, True), y) :363-367, and log_marginal_likelihood value :541-656.
//
// B200 design: everything n x n is tiled in 64 x 64 blocks of FP64 and every O(n^3) step is a
// DMMA (mma.sync m8n8k4 f64 -> SASS DMMA.8x8x4) tile GEMM:
//   * right-looking Cholesky: 64x64 diagonal factor + its inverse in one CTA (shared memory),
//     panel = A21 * L11^-T by forward substitution (a GEMM with inv(L11) loses cond(L11) digits),
//     trailing A22 -= P P^T (DMMA GEMM, lower tiles only);
//   * inv(L) by recursive doubling: X21 = -inv(C) (B inv(A)), two batched GEMMs per level;
//   * alpha by blocked forward/back substitution in one CTA (n^2 work, bandwidth trivial).
// inv(L) is then written in the tile-packed layout the score kernel streams with bulk copies.
#include <math.h>
#include <stdlib.h>

#include "ob200_internal.cuh"

#define FB 64            // block size of the fit algorithms
#define GEMM_BK 16
#define OB200_FIT_MAGIC 0x0b200f18

namespace {

// ------------------------------------------------------------------------------------------
// y statistics + normalised y.  Single CTA, deterministic tree reductions.
// stats_out: [0]=mean, [1]=std (zero-scale guarded), [2]=min(y)
__global__ void __launch_bounds__(1024) ystats_kernel(const double* __restrict__ y, int n, int n_pad, int normalize,
                                                      double* __restrict__ y_norm, double* __restrict__ stats_out) {
  __shared__ double red[1024];
  __shared__ double s_mean, s_std;
  const int t = threadIdx.x;
  double acc = 0.0, mn = INFINITY;
  for (int i = t; i < n; i += 1024) { acc += y[i]; mn = fmin(mn, y[i]); }
  red[t] = acc;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) { if (t < s) red[t] += red[t + s]; __syncthreads(); }
  if (t == 0) s_mean = red[0] / (double)n;
  __syncthreads();
  const double mean = s_mean;
  acc = 0.0;
  for (int i = t; i < n; i += 1024) { double d = y[i] - mean; acc += d * d; }
  __syncthreads();
  red[t] = acc;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) { if (t < s) red[t] += red[t + s]; __syncthreads(); }
  if (t == 0) {
    double sd = sqrt(red[0] / (double)n);
    if (sd < 10.0 * 2.220446049250313e-16) sd = 1.0;  // sklearn _handle_zeros_in_scale
    s_std = sd;
  }
  __syncthreads();
  red[t] = mn;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) { if (t < s) red[t] = fmin(red[t], red[t + s]); __syncthreads(); }
  const double m_use = normalize ? mean : 0.0, s_use = normalize ? s_std : 1.0;
  if (t == 0) { stats_out[0] = m_use; stats_out[1] = s_use; stats_out[2] = red[0]; }
  for (int i = t; i < n_pad; i += 1024) y_norm[i] = i < n ? (y[i] - m_use) / s_use : 0.0;
}

// Xs = X / length_scale (row-major [n_pad x D_pad], zero padded) and xx = ||Xs_i||^2
__global__ void scale_x_kernel(const double* __restrict__ X, int n, int D, int n_pad, int D_pad,
                               const double* __restrict__ ls, int n_ls,
                               double* __restrict__ Xs, double* __restrict__ xx) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pad) return;
  double acc = 0.0;
  for (int d = 0; d < D_pad; ++d) {
    double v = 0.0;
    if (i < n && d < D) v = X[(size_t)i * D + d] / ls[n_ls == 1 ? 0 : d];
    Xs[(size_t)i * D_pad + d] = v;
    acc = fma(v, v, acc);
  }
  xx[i] = acc;
}

__device__ __forceinline__ double kernel_from_d2(int kind, double d2, double amplitude) {
  if (kind == OB200_KERNEL_MATERN52) {
    const double k = sqrt(d2) * 2.23606797749979;  // sqrt(5)
    return amplitude * ((1.0 + k + k * k / 3.0) * exp(-k));
  }
  return amplitude * exp(-0.5 * d2);
}

// Lower triangle of K (+ alpha on the diagonal) with pdist-style direct differences; identity pad.
__global__ void __launch_bounds__(256) gram_kernel(const double* __restrict__ Xs, int n, int n_pad, int D_pad, int kind,
                                                   double amplitude, double alpha, double* __restrict__ K) {
  const int ti = blockIdx.y, tj = blockIdx.x;
  if (tj > ti) {  // strictly upper tiles: zero (the buffer is reused across fits of different n_pad)
    for (int e = threadIdx.x; e < 1024; e += 256) K[(size_t)(ti * 32 + (e >> 5)) * n_pad + tj * 32 + (e & 31)] = 0.0;
    return;
  }
  extern __shared__ double sm[];
  double* xi = sm;                       // [32][D_pad+1]
  double* xj = sm + 32 * (D_pad + 1);
  const int ldx = D_pad + 1;
  for (int e = threadIdx.x; e < 32 * D_pad; e += 256) {
    const int r = e / D_pad, d = e - r * D_pad;
    xi[r * ldx + d] = Xs[(size_t)(ti * 32 + r) * D_pad + d];
    xj[r * ldx + d] = Xs[(size_t)(tj * 32 + r) * D_pad + d];
  }
  __syncthreads();
  for (int e = threadIdx.x; e < 1024; e += 256) {
    const int r = e >> 5, c = e & 31;
    const int i = ti * 32 + r, j = tj * 32 + c;
    double v;
    if (i >= n || j >= n) {
      v = (i == j) ? 1.0 : 0.0;
    } else if (j > i) {
      v = 0.0;
    } else {
      double d2 = 0.0;
      for (int d = 0; d < D_pad; ++d) { const double df = xi[r * ldx + d] - xj[c * ldx + d]; d2 = fma(df, df, d2); }
      v = kernel_from_d2(kind, d2, amplitude);
      if (i == j) v += alpha;
    }
    K[(size_t)i * n_pad + j] = v;
  }
}

// ------------------------------------------------------------------------------------------
// Generic DMMA tile GEMM for the fit:  C[z] = alpha * A[z] * op(B[z]) + beta * C[z]
//   tile 64x64 per CTA, 4 warps (2x2), warp tile 32x32 = 4x4 DMMA fragments, K step 16.
//   All extents are multiples of 64 (M,N) / 16 (K).  TRANS_B: B is [N x K] row-major ("NT").
//   lower_only: skip tiles strictly above the diagonal (SYRK).  m_last: M of the last batch entry.
struct gemm_args {
  const double *A, *B;
  double* C;
  int M, N, K, lda, ldb, ldc;
  long long strideA, strideB, strideC;
  int m_last, k_eq_m;  // k_eq_m: K follows the batch entry's M (TRTRI second product)
  double alpha, beta;
  int lower_only;
  int tri_k;  // bit 0: B is lower triangular, k starts at 64*tn; bit 1: A is lower triangular, k ends at 64*(tm+1); 3: both;
              // bit 2: A (M x K, row-major) is UPPER triangular, k starts at 64*tm
};

template <bool TRANS_B>
__global__ void __launch_bounds__(128) gemm64_kernel(gemm_args g) {
  const int z = blockIdx.z;
  int M = g.M, K = g.K;
  if (z == (int)gridDim.z - 1 && g.m_last > 0) { M = g.m_last; if (g.k_eq_m) K = g.m_last; }
  const int tm = blockIdx.y, tn = blockIdx.x;
  if (tm * 64 >= M) return;
  if (g.lower_only && tn > tm) return;
  const double* __restrict__ A = g.A + z * g.strideA + (size_t)tm * 64 * g.lda;
  const double* __restrict__ B = g.B + z * g.strideB + (TRANS_B ? (size_t)tn * 64 * g.ldb : (size_t)tn * 64);
  double* C = g.C + z * g.strideC + (size_t)tm * 64 * g.ldc + (size_t)tn * 64;

  __shared__ double As[64][GEMM_BK + 4];
  __shared__ double Bs[TRANS_B ? 64 : GEMM_BK][TRANS_B ? GEMM_BK + 4 : 64 + 4];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int gq = lane >> 2, q = lane & 3;
  const int wm = warp >> 1, wn = warp & 1;
  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

  const int k_begin = (g.tri_k & 4) ? tm * 64 : ((g.tri_k & 1) ? tn * 64 : 0);
  if (g.tri_k & 2) { const int k_end = (tm + 1) * 64; if (k_end < K) K = k_end; }
  // The next K slab is fetched into registers while the current one is multiplied: at the one or two CTAs per SM the
  // fit's GEMMs have, the loop is bound by the latency of these loads.
  const int ar = tid >> 1, ac = (tid & 1) * 8;                          // A tile 64 x 16: row tid/2, 8 contiguous doubles
  const int br = TRANS_B ? ar : (tid >> 3), bc = TRANS_B ? ac : (tid & 7) * 8;
  const double* a_src = A + (size_t)ar * g.lda + ac;
  const double* b_src = TRANS_B ? B + (size_t)br * g.ldb + bc : B + (size_t)br * g.ldb + bc;
  double2 ra[4], rb[4];
  auto fetch = [&](int k0) {
    const double2* pa = reinterpret_cast<const double2*>(a_src + k0);
    const double2* pb = reinterpret_cast<const double2*>(TRANS_B ? b_src + k0 : b_src + (size_t)k0 * g.ldb);
#pragma unroll
    for (int v = 0; v < 4; ++v) { ra[v] = pa[v]; rb[v] = pb[v]; }
  };
  if (k_begin < K) fetch(k_begin);
  for (int k0 = k_begin; k0 < K; k0 += GEMM_BK) {
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      As[ar][ac + 2 * v] = ra[v].x; As[ar][ac + 2 * v + 1] = ra[v].y;
      Bs[br][bc + 2 * v] = rb[v].x; Bs[br][bc + 2 * v + 1] = rb[v].y;
    }
    __syncthreads();
    if (k0 + GEMM_BK < K) fetch(k0 + GEMM_BK);
#pragma unroll
    for (int kk = 0; kk < GEMM_BK; kk += 4) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[wm * 32 + i * 8 + gq][kk + q];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = TRANS_B ? Bs[wn * 32 + j * 8 + gq][kk + q] : Bs[kk + q][wn * 32 + j * 8 + gq];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      double2* dst = reinterpret_cast<double2*>(C + (size_t)(wm * 32 + i * 8 + gq) * g.ldc + wn * 32 + j * 8 + 2 * q);
      double2 out;
      if (g.beta != 0.0) {
        double2 old = *dst;
        out.x = g.alpha * acc[i][j][0] + g.beta * old.x;
        out.y = g.alpha * acc[i][j][1] + g.beta * old.y;
      } else {
        out.x = g.alpha * acc[i][j][0];
        out.y = g.alpha * acc[i][j][1];
      }
      *dst = out;
    }
}

// ------------------------------------------------------------------------------------------
// Diagonal block: factor A[kb,kb] (64x64) in shared memory, write L back (upper zeroed), and
// write inv(L_kk) to Dinv[kb].  info: 0 ok, else 1-based index of the failing pivot.
// Forward substitution L x = b for ONE right-hand side per thread, L (64 x 64, lower) broadcast from shared memory,
// x in registers, both loops fully unrolled: the only serial chain is the 4-way split dot product (<= 16 FMAs)
// plus one multiply by the stored reciprocal of the pivot; no barrier, no division, no shared-memory round trip.
// LS = row stride of L in doubles (even: the compiler can pair the broadcast loads).
template <int LS>
__device__ __forceinline__ void fwd_subst64(const double* __restrict__ L, const double* __restrict__ rdiag, double (&x)[FB]) {
#pragma unroll
  for (int j = 0; j < FB; ++j) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
    for (int k = 0; k < j; ++k) {
      const double l = L[j * LS + k];
      if ((k & 3) == 0) a0 = fma(l, x[k], a0);
      else if ((k & 3) == 1) a1 = fma(l, x[k], a1);
      else if ((k & 3) == 2) a2 = fma(l, x[k], a2);
      else a3 = fma(l, x[k], a3);
    }
    x[j] = (x[j] - ((a0 + a1) + (a2 + a3))) * rdiag[j];
  }
}

#define FBS (FB + 2)  // shared-memory row stride of the 64 x 64 blocks (even: 16-byte aligned rows)

// Left-looking factorisation of a 64 x 64 block held in shared memory (lower triangle, upper part zero), by the first
// two warps: one thread per row, the own row in registers, fully unrolled; column j of row i is
// (a_ij - sum_{k<j} l_ik l_jk) / l_jj with row j read from shared memory (broadcast).  Per column: the split dot product,
// the pivot through shared memory, one sqrt, two 64-thread barriers.  Every thread of the CTA must call it.  On return
// S = L, rdiag[j] = 1 / l_jj; returns the 1-based index of the first non-positive pivot (0 = ok).
__device__ __forceinline__ int factor64(double* S, double* piv_s, double* rdiag, int* bad_s, int tid) {
  if (tid == 0) *bad_s = 0;
  __syncthreads();
  if (tid < FB) {
    const int i = tid;
    double l[FB];
#pragma unroll
    for (int k = 0; k < FB; ++k) l[k] = S[i * FBS + k];
#pragma unroll
    for (int j = 0; j < FB; ++j) {
      double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
      for (int k = 0; k < j; ++k) {
        const double r = S[j * FBS + k];
        if ((k & 3) == 0) a0 = fma(l[k], r, a0);
        else if ((k & 3) == 1) a1 = fma(l[k], r, a1);
        else if ((k & 3) == 2) a2 = fma(l[k], r, a2);
        else a3 = fma(l[k], r, a3);
      }
      const double v = l[j] - ((a0 + a1) + (a2 + a3));
      if (i == j) piv_s[j] = v;
      asm volatile("bar.sync 1, 64;" ::: "memory");
      const double piv = piv_s[j];
      if (!(piv > 0.0)) {  // also catches NaN (uniform: everybody reads the same pivot)
        if (tid == 0) *bad_s = j + 1;
        break;
      }
      const double d = sqrt(piv), rd = 1.0 / d;
      l[j] = (i == j) ? d : v * rd;
      if (i >= j) S[i * FBS + j] = l[j];
      if (i == j) rdiag[j] = rd;
      asm volatile("bar.sync 1, 64;" ::: "memory");  // column j is final before row j + 1 is read by everybody
    }
  }
  __syncthreads();
  return *bad_s;
}

// One step of the blocked Cholesky in ONE launch: block 0 factors the diagonal block A[kb,kb], inverts the factor
// (Dinv[kb], for the recursive-doubling inverse) and stores the factor in Lfac[kb]; block b >= 1 factors the SAME
// diagonal block redundantly in its own shared memory (6 us of arithmetic, identical bits) and solves its 64 rows of
// the panel P = A21 L11^-T by forward substitution (a GEMM with inv(L11) loses cond(L11) digits: 7e-6 instead of
// 3e-8 in the posterior variance at cond(K) = 1e12, tools/illcond_probe.py).  No block waits for another, so the
// diagonal factor costs no launch of its own.  The diagonal block itself stays untouched in A until the factor blocks
// are copied in after the last step (the panel CTAs of this launch still read the original values).
__global__ void __launch_bounds__(256) chol_step_kernel(double* __restrict__ A, int ld, int kb, double* __restrict__ Dinv,
                                                        double* __restrict__ Lfac, int* __restrict__ info) {
  extern __shared__ double diag_sm[];
  double* S = diag_sm;                 // [FB][FBS] the diagonal block, then its factor
  double* W = diag_sm + FB * FBS;      // block 0: [FB][FBS] inverse of the factor; others: [FB][FB + 1] panel tile, transposed
  __shared__ double piv_s[FB], rdiag[FB];
  __shared__ int bad;
  const double* blk = A + (size_t)kb * FB * ld + (size_t)kb * FB;
  const int tid = threadIdx.x;
  double* At = nullptr;
  for (int e = tid; e < FB * FB; e += 256) { const int r = e >> 6, c = e & 63; S[r * FBS + c] = (c <= r) ? blk[(size_t)r * ld + c] : 0.0; }
  if (blockIdx.x > 0) {
    At = A + (size_t)(kb + blockIdx.x) * FB * ld + (size_t)kb * FB;
    for (int e = tid; e < FB * FB; e += 256) { const int r = e >> 6, c = e & 63; W[c * (FB + 1) + r] = At[(size_t)r * ld + c]; }
  }
  __syncthreads();
  if (*info != 0) return;  // an earlier step already failed
  const int fail = factor64(S, piv_s, rdiag, &bad, tid);
  if (fail) { if (blockIdx.x == 0 && tid == 0) *info = kb * FB + fail; return; }
  if (blockIdx.x == 0) {
    // inverse of the factor: thread c solves L x = e_c (rows < c come out as exact zeros)
    if (tid < FB) {
      double x[FB];
#pragma unroll
      for (int k = 0; k < FB; ++k) x[k] = (k == tid) ? 1.0 : 0.0;
      fwd_subst64<FBS>(S, rdiag, x);
#pragma unroll
      for (int k = 0; k < FB; ++k) W[k * FBS + tid] = x[k];
    }
    __syncthreads();
    double* dinv = Dinv + (size_t)kb * FB * FB;
    double* lf = Lfac + (size_t)kb * FB * FB;
    for (int e = tid; e < FB * FB; e += 256) {
      const int r = e >> 6, c = e & 63;
      lf[e] = S[r * FBS + c];
      dinv[e] = W[r * FBS + c];
    }
  } else {
    // x L11^T = a, i.e. L11 x^T = a^T: one thread per row of the panel, true forward substitution, see fwd_subst64
    if (tid < FB) {
      double x[FB];
#pragma unroll
      for (int k = 0; k < FB; ++k) x[k] = W[k * (FB + 1) + tid];
      fwd_subst64<FBS>(S, rdiag, x);
#pragma unroll
      for (int k = 0; k < FB; ++k) W[k * (FB + 1) + tid] = x[k];
    }
    __syncthreads();
    for (int e = tid; e < FB * FB; e += 256) {
      const int r = e >> 6, c = e & 63;
      At[(size_t)r * ld + c] = W[c * (FB + 1) + r];
    }
  }
}

// ---- compact variant of the step kernel ---------------------------------------------------------------------------
// Same job as chol_step_kernel (which is 14.7 k instructions of straight-line code executed once per launch, i.e.
// instruction-fetch-bound: ~40 us whatever the panel height), written as loops over 16-column sub-blocks so that it
// stays in the instruction cache (~1.5 k instructions):
//   factor   for each 16-column sub-block of the diagonal block: (1) its 16 x 16 diagonal part by one warp (lane = row,
//            pivots by shuffle, no CTA barrier), (2) the rows below by forward substitution, one thread per row,
//            (3) the rank-16 update of the rest of the block by all threads;
//   panel    X L11^T = A21 by the same sub-blocks: (a) 16 unrolled substitution steps per row, (b) rank-16 update of
//            the columns to the right by all threads.  Block 0 runs the panel code on the identity: X = inv(L11)^T.
// The panel is still a true substitution (see the note on cond(K) above); only the summation order differs from the
// unrolled kernel.
#define SB 16                 // sub-block width
#define WS (FB + 1)           // row stride of the panel tile in shared memory (odd: conflict-free column access)

__device__ __forceinline__ double rsqrt_pos(double x) {
  // 1 / sqrt(x) for a positive double in float range: MUFU.RSQ seed, two Newton steps (23 -> 46 -> 92 bits before rounding)
  double y = (double)rsqrtf((float)x);
  y = fma(y, 0.5 * fma(-x * y, y, 1.0), y);
  y = fma(y, 0.5 * fma(-x * y, y, 1.0), y);
  return y;
}

// 16 forward-substitution steps for one row: x <- x * inv(Lsub)^T with Lsub = S[c0.., c0..] (lower) and rd = 1 / diag
__device__ __forceinline__ void subst16(const double* __restrict__ S, const double* __restrict__ rdiag, int c0, double (&x)[SB]) {
#pragma unroll
  for (int k = 0; k < SB; ++k) {
    double a0 = 0.0, a1 = 0.0;
#pragma unroll
    for (int m = 0; m < k; ++m) {
      const double l = S[(c0 + k) * FBS + c0 + m];
      if (m & 1) a1 = fma(l, x[m], a1); else a0 = fma(l, x[m], a0);
    }
    x[k] = (x[k] - (a0 + a1)) * rdiag[c0 + k];
  }
}

__global__ void __launch_bounds__(256) chol_step_compact_kernel(double* __restrict__ A, int ld, int kb, double* __restrict__ Dinv,
                                                                double* __restrict__ Lfac, int* __restrict__ info) {
  extern __shared__ double diag_sm[];
  double* S = diag_sm;                 // [FB][FBS] the diagonal block, then its factor
  double* W = diag_sm + FB * FBS;      // [FB][WS] panel tile (block 0: the identity, then inv(L11)^T)
  __shared__ double rdiag[FB];
  __shared__ int bad;
  const double* blk = A + (size_t)kb * FB * ld + (size_t)kb * FB;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  double* At = nullptr;
  for (int e = tid; e < FB * FB; e += 256) { const int r = e >> 6, c = e & 63; S[r * FBS + c] = (c <= r) ? blk[(size_t)r * ld + c] : 0.0; }
  if (blockIdx.x > 0) {
    At = A + (size_t)(kb + blockIdx.x) * FB * ld + (size_t)kb * FB;
    for (int e = tid; e < FB * FB; e += 256) { const int r = e >> 6, c = e & 63; W[r * WS + c] = At[(size_t)r * ld + c]; }
  } else {
    for (int e = tid; e < FB * FB; e += 256) { const int r = e >> 6, c = e & 63; W[r * WS + c] = (r == c) ? 1.0 : 0.0; }
  }
  if (tid == 0) bad = 0;
  __syncthreads();
  if (*info != 0) return;  // an earlier step already failed

  // ---- factor the diagonal block, sub-block by sub-block
  for (int c0 = 0; c0 < FB; c0 += SB) {
    if (warp == 0) {
      // (1) 16 x 16 diagonal part: lane r < 16 owns row c0 + r; left-looking inside the sub-block
      const int r = lane & 15;
      double a[SB];
#pragma unroll
      for (int k = 0; k < SB; ++k) a[k] = S[(c0 + r) * FBS + c0 + k];
      int fail = 0;
#pragma unroll
      for (int j = 0; j < SB; ++j) {
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int k = 0; k < j; ++k) {
          const double l = S[(c0 + j) * FBS + c0 + k];   // row j of the factor, written by lane j in earlier columns
          if (k & 1) s1 = fma(a[k], l, s1); else s0 = fma(a[k], l, s0);
        }
        const double v = a[j] - (s0 + s1);
        const double piv = __shfl_sync(0xffffffffu, v, j);
        if (!(piv > 0.0) && !fail) fail = c0 + j + 1;   // also catches NaN; uniform across the warp
        const double rd = (piv > 1e-30 && piv < 1e30) ? rsqrt_pos(piv) : 1.0 / sqrt(piv);
        a[j] = (r == j) ? piv * rd : v * rd;
        if (lane < SB && r >= j) S[(c0 + r) * FBS + c0 + j] = a[j];
        if (lane == j) rdiag[c0 + j] = rd;
        __syncwarp();
      }
      if (fail && lane == 0) bad = fail;
    }
    __syncthreads();
    if (bad) break;
    const int below = FB - (c0 + SB);   // rows under the sub-block
    // (2) rows below: P = A[i, c0:c0+16] * inv(Lsub)^T, one thread per row
    if (tid < below) {
      const int i = c0 + SB + tid;
      double x[SB];
#pragma unroll
      for (int k = 0; k < SB; ++k) x[k] = S[i * FBS + c0 + k];
      subst16(S, rdiag, c0, x);
#pragma unroll
      for (int k = 0; k < SB; ++k) S[i * FBS + c0 + k] = x[k];
    }
    __syncthreads();
    // (3) rank-16 update of the rest of the block (lower triangle): S[i][k] -= P[i] . P[k]
    for (int e = tid; e < below * below; e += 256) {
      const int ii = e / below, kk = e - ii * below;
      if (kk > ii) continue;
      const double* pi = S + (c0 + SB + ii) * FBS + c0;
      const double* pk = S + (c0 + SB + kk) * FBS + c0;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int c = 0; c < SB; c += 2) { s0 = fma(pi[c], pk[c], s0); s1 = fma(pi[c + 1], pk[c + 1], s1); }
      S[(c0 + SB + ii) * FBS + c0 + SB + kk] -= s0 + s1;
    }
    __syncthreads();
  }
  if (bad) { if (blockIdx.x == 0 && tid == 0) *info = kb * FB + bad; return; }

  // ---- panel: X L11^T = W, sub-block by sub-block
  for (int c0 = 0; c0 < FB; c0 += SB) {
    if (tid < FB) {   // (a) 16 substitution steps for row tid
      double x[SB];
#pragma unroll
      for (int k = 0; k < SB; ++k) x[k] = W[tid * WS + c0 + k];
      subst16(S, rdiag, c0, x);
#pragma unroll
      for (int k = 0; k < SB; ++k) W[tid * WS + c0 + k] = x[k];
    }
    __syncthreads();
    // (b) columns to the right: W[r][cc] -= X[r, c0:c0+16] . L11[cc, c0:c0+16]
    const int right = FB - (c0 + SB);
    for (int e = tid; e < FB * right; e += 256) {
      const int r = e & 63, cc = c0 + SB + (e >> 6);
      const double* xr = W + r * WS + c0;
      const double* lc = S + cc * FBS + c0;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int c = 0; c < SB; c += 2) { s0 = fma(xr[c], lc[c], s0); s1 = fma(xr[c + 1], lc[c + 1], s1); }
      W[r * WS + cc] -= s0 + s1;
    }
    __syncthreads();
  }
  if (blockIdx.x == 0) {
    double* dinv = Dinv + (size_t)kb * FB * FB;
    double* lf = Lfac + (size_t)kb * FB * FB;
    for (int e = tid; e < FB * FB; e += 256) {
      const int r = e >> 6, c = e & 63;
      lf[e] = S[r * FBS + c];
      dinv[e] = (c <= r) ? W[c * WS + r] : 0.0;   // inv(L11) = X^T, lower triangular
    }
  } else {
    for (int e = tid; e < FB * FB; e += 256) {
      const int r = e >> 6, c = e & 63;
      At[(size_t)r * ld + c] = W[r * WS + c];
    }
  }
}

// ---- the step split in two: diagonal block / panel ------------------------------------------------------------------
// The factorisation of the 64 x 64 diagonal block is a serial chain (~8 us) that every CTA of the fused step kernel
// repeats before it can touch its panel rows.  Split off, it runs as ONE CTA beside the GEMM that updates the rest of the
// next block column (it applies that update to the diagonal tile itself), and the panel kernel starts from the finished
// factor: the chain leaves the critical path of the step.
//
// chol_diag_kernel: S = A[kb, kb] - P P^T (P = the previous panel's 64 x 64 block in these rows; kb = 0: no update),
// factor as in chol_step_compact_kernel, factor -> Lfac[kb], reciprocal pivots -> Rd[kb].
__global__ void __launch_bounds__(256) chol_diag_kernel(const double* __restrict__ A, int ld, int kb, double* __restrict__ Lfac,
                                                        double* __restrict__ Rd, int* __restrict__ info) {
  extern __shared__ double diag_sm[];
  double* S = diag_sm;                 // [FB][FBS]
  double* W = diag_sm + FB * FBS;      // [FB][WS] P
  __shared__ double rdiag[FB];
  __shared__ int bad;
  const double* blk = A + (size_t)kb * FB * ld + (size_t)kb * FB;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int e = tid; e < FB * FB; e += 256) { const int r = e >> 6, c = e & 63; S[r * FBS + c] = (c <= r) ? blk[(size_t)r * ld + c] : 0.0; }
  if (kb > 0) {
    const double* P = blk - FB;
    for (int e = tid; e < FB * FB; e += 256) { const int r = e >> 6, c = e & 63; W[r * WS + c] = P[(size_t)r * ld + c]; }
  }
  if (tid == 0) bad = 0;
  __syncthreads();
  if (*info != 0) return;  // an earlier step already failed
#ifdef FIT_DEBUG_SKIP_DIAG_UPDATE   // timing experiment only
  if (false) {
#else
  if (kb > 0) {
#endif
    // thread (ri, cj) owns the elements (ri + 16 u, cj + 16 v), v <= u (the others lie above the diagonal): rows of P are
    // read at stride 16, conflict-free with the odd row stride WS
    const int ri = tid >> 4, cj = tid & 15;
    double acc[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int v = 0; v < 4; ++v) acc[u][v] = 0.0;
    for (int c = 0; c < FB; ++c) {
      double a[4], b[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) { a[u] = W[(ri + 16 * u) * WS + c]; b[u] = W[(cj + 16 * u) * WS + c]; }
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v <= u; ++v) acc[u][v] = fma(a[u], b[v], acc[u][v]);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int v = 0; v <= u; ++v)
        if (cj + 16 * v <= ri + 16 * u) S[(ri + 16 * u) * FBS + cj + 16 * v] -= acc[u][v];
    __syncthreads();
  }
#ifdef FIT_DEBUG_SKIP_DIAG_FACTOR   // timing experiment only
  if (tid < FB) rdiag[tid] = 1.0;
  for (int c0 = FB; c0 < FB; c0 += SB) {
#else
  for (int c0 = 0; c0 < FB; c0 += SB) {
#endif
    if (warp == 0) {
      const int r = lane & 15;
      double a[SB];
#pragma unroll
      for (int k = 0; k < SB; ++k) a[k] = S[(c0 + r) * FBS + c0 + k];
      int fail = 0;
#pragma unroll
      for (int j = 0; j < SB; ++j) {
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int k = 0; k < j; ++k) {
          const double l = S[(c0 + j) * FBS + c0 + k];
          if (k & 1) s1 = fma(a[k], l, s1); else s0 = fma(a[k], l, s0);
        }
        const double v = a[j] - (s0 + s1);
        const double piv = __shfl_sync(0xffffffffu, v, j);
        if (!(piv > 0.0) && !fail) fail = c0 + j + 1;
        const double rd = (piv > 1e-30 && piv < 1e30) ? rsqrt_pos(piv) : 1.0 / sqrt(piv);
        a[j] = (r == j) ? piv * rd : v * rd;
        if (lane < SB && r >= j) S[(c0 + r) * FBS + c0 + j] = a[j];
        if (lane == j) rdiag[c0 + j] = rd;
        __syncwarp();
      }
      if (fail && lane == 0) bad = fail;
    }
    __syncthreads();
    if (bad) break;
    const int below = FB - (c0 + SB);
    if (tid < below) {
      const int i = c0 + SB + tid;
      double x[SB];
#pragma unroll
      for (int k = 0; k < SB; ++k) x[k] = S[i * FBS + c0 + k];
      subst16(S, rdiag, c0, x);
#pragma unroll
      for (int k = 0; k < SB; ++k) S[i * FBS + c0 + k] = x[k];
    }
    __syncthreads();
    for (int e = tid; e < below * below; e += 256) {
      const int ii = e / below, kk = e - ii * below;
      if (kk > ii) continue;
      const double* pi = S + (c0 + SB + ii) * FBS + c0;
      const double* pk = S + (c0 + SB + kk) * FBS + c0;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int c = 0; c < SB; c += 2) { s0 = fma(pi[c], pk[c], s0); s1 = fma(pi[c + 1], pk[c + 1], s1); }
      S[(c0 + SB + ii) * FBS + c0 + SB + kk] -= s0 + s1;
    }
    __syncthreads();
  }
  if (bad) { if (tid == 0) *info = kb * FB + bad; return; }
  double* lf = Lfac + (size_t)kb * FB * FB;
  for (int e = tid; e < FB * FB; e += 256) lf[e] = S[(e >> 6) * FBS + (e & 63)];
  if (tid < FB) Rd[(size_t)kb * FB + tid] = rdiag[tid];
}

// chol_panel_kernel: X L11^T = A21 for the 64 rows of CTA blockIdx.x - 1, L11 and its reciprocal pivots from
// chol_diag_kernel; CTA 0 solves for the identity: inv(L11)^T -> Dinv[kb].  The panel code of chol_step_compact_kernel.
__global__ void __launch_bounds__(256) chol_panel_kernel(double* __restrict__ A, int ld, int kb, double* __restrict__ Dinv,
                                                         const double* __restrict__ Lfac, const double* __restrict__ Rd,
                                                         const int* __restrict__ info) {
  extern __shared__ double diag_sm[];
  double* S = diag_sm;                 // [FB][FBS] L11
  double* W = diag_sm + FB * FBS;      // [FB][WS] panel tile (CTA 0: the identity, then inv(L11)^T)
  __shared__ double rdiag[FB];
  const int tid = threadIdx.x;
  double* At = nullptr;
  const double* lf = Lfac + (size_t)kb * FB * FB;
  if (blockIdx.x > 0) {
    At = A + (size_t)(kb + blockIdx.x) * FB * ld + (size_t)kb * FB;
    for (int e = tid; e < FB * FB; e += 256) { const int r = e >> 6, c = e & 63; W[r * WS + c] = At[(size_t)r * ld + c]; }
  } else {
    for (int e = tid; e < FB * FB; e += 256) { const int r = e >> 6, c = e & 63; W[r * WS + c] = (r == c) ? 1.0 : 0.0; }
  }
  if (*info != 0) return;  // this or an earlier diagonal block failed (uniform: nothing below is reached by any thread)
  for (int e = tid; e < FB * FB; e += 256) S[(e >> 6) * FBS + (e & 63)] = lf[e];
  if (tid < FB) rdiag[tid] = Rd[(size_t)kb * FB + tid];
  __syncthreads();
#ifdef FIT_DEBUG_SKIP_PANEL_SOLVE   // timing experiment only
  for (int c0 = FB; c0 < FB; c0 += SB) {
#else
  for (int c0 = 0; c0 < FB; c0 += SB) {
#endif
    if (tid < FB) {
      double x[SB];
#pragma unroll
      for (int k = 0; k < SB; ++k) x[k] = W[tid * WS + c0 + k];
      subst16(S, rdiag, c0, x);
#pragma unroll
      for (int k = 0; k < SB; ++k) W[tid * WS + c0 + k] = x[k];
    }
    __syncthreads();
    const int right = FB - (c0 + SB);
    for (int e = tid; e < FB * right; e += 256) {
      const int r = e & 63, cc = c0 + SB + (e >> 6);
      const double* xr = W + r * WS + c0;
      const double* lc = S + cc * FBS + c0;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int c = 0; c < SB; c += 2) { s0 = fma(xr[c], lc[c], s0); s1 = fma(xr[c + 1], lc[c + 1], s1); }
      W[r * WS + cc] -= s0 + s1;
    }
    __syncthreads();
  }
  if (blockIdx.x == 0) {
    double* dinv = Dinv + (size_t)kb * FB * FB;
    for (int e = tid; e < FB * FB; e += 256) {
      const int r = e >> 6, c = e & 63;
      dinv[e] = (c <= r) ? W[c * WS + r] : 0.0;   // inv(L11) = X^T, lower triangular
    }
  } else {
    for (int e = tid; e < FB * FB; e += 256) {
      const int r = e >> 6, c = e & 63;
      At[(size_t)r * ld + c] = W[r * WS + c];
    }
  }
}

// the factored diagonal blocks (upper triangle zero) go into L after the last step
__global__ void __launch_bounds__(256) chol_copy_diag_kernel(double* __restrict__ A, int ld, const double* __restrict__ Lfac) {
  const int kb = blockIdx.x;
  double* blk = A + (size_t)kb * FB * ld + (size_t)kb * FB;
  const double* lf = Lfac + (size_t)kb * FB * FB;
  for (int e = threadIdx.x; e < FB * FB; e += 256) blk[(size_t)(e >> 6) * ld + (e & 63)] = lf[e];
}


// zero the strict upper triangle tiles left over from the Gram/SYRK (they hold garbage-free zeros
// already from gram_kernel; SYRK only touches lower tiles) -- nothing to do.  X (inverse) init:
// X = 0 except diagonal blocks = Dinv.
__global__ void trtri_init_kernel(double* __restrict__ X, int ld, const double* __restrict__ Dinv) {
  const int kb = blockIdx.x;
  double* blk = X + (size_t)kb * FB * ld + (size_t)kb * FB;
  const double* d = Dinv + (size_t)kb * FB * FB;
  for (int e = threadIdx.x; e < FB * FB; e += blockDim.x) blk[(size_t)(e >> 6) * ld + (e & 63)] = d[e];
}


__global__ void __launch_bounds__(256) eye_minus_lower_kernel(double* __restrict__ R, int n_pad) {
  const int ti = blockIdx.y, tj = blockIdx.x;
  for (int e = threadIdx.x; e < 1024; e += 256) {
    const int i = ti * 32 + (e >> 5), j = tj * 32 + (e & 31);
    double* p = R + (size_t)i * n_pad + j;
    *p = (j > i) ? 0.0 : ((i == j ? 1.0 : 0.0) - *p);
  }
}

// alpha = L^-T L^-1 y by blocked substitution (diagonal blocks by substitution too, like LAPACK potrs); also
// lml = -0.5 y^T alpha - sum log L_ii - n/2 log(2 pi).  One CTA of 1024 threads.
__global__ void __launch_bounds__(1024) chol_solve_kernel(const double* __restrict__ L, int ld, int n, int n_pad,
                                                          const double* __restrict__ y, double* __restrict__ tvec,
                                                          double* __restrict__ alpha_out, double* __restrict__ lml_out) {
  __shared__ double part[16][FB + 1];
  __shared__ double r[FB];
  __shared__ double shbuf[FB * (FB + 1)];  // diagonal block L_ii during the sweeps, reduction scratch at the end
  double (*Ld)[FB + 1] = reinterpret_cast<double (*)[FB + 1]>(shbuf);
  const int tid = threadIdx.x, lane = tid & 31;
  const int nb = n_pad / FB;
  // forward: L_ii t_i = y_i - L[i, :64 i] t[:64 i], the diagonal system by substitution (warp 0)
  for (int ib = 0; ib < nb; ++ib) {
    {
      const int row = tid >> 4, g = tid & 15;  // 64 rows x 16 column lanes
      const double* Lr = L + (size_t)(ib * FB + row) * ld;
      double acc = 0.0;
      for (int j = g; j < ib * FB; j += 16) acc = fma(Lr[j], tvec[j], acc);
      part[g][row] = acc;
    }
    for (int e = tid; e < FB * FB; e += 1024) Ld[e >> 6][e & 63] = L[(size_t)(ib * FB + (e >> 6)) * ld + ib * FB + (e & 63)];
    __syncthreads();
    if (tid < FB) {
      double s = 0.0;
      for (int g = 0; g < 16; ++g) s += part[g][tid];
      r[tid] = y[ib * FB + tid] - s;
    }
    __syncthreads();
    if (tid < 32) {
      for (int j = 0; j < FB; ++j) {
        double p = 0.0;
        if (lane < j) p = Ld[j][lane] * r[lane];
        if (lane + 32 < j) p = fma(Ld[j][lane + 32], r[lane + 32], p);
        for (int off = 16; off > 0; off >>= 1) p += __shfl_xor_sync(0xffffffffu, p, off);
        if (lane == 0) r[j] = (r[j] - p) / Ld[j][j];
        __syncwarp();
      }
      tvec[ib * FB + lane] = r[lane];
      tvec[ib * FB + lane + 32] = r[lane + 32];
    }
    __threadfence_block();
    __syncthreads();
  }
  // backward: L_ii^T alpha_i = t_i - sum_{k >= 64(i+1)} L[k, i-block]^T alpha_k
  for (int ib = nb - 1; ib >= 0; --ib) {
    {
      const int c = tid & 63, g = tid >> 6;  // 64 columns x 16 row lanes
      double acc = 0.0;
      for (int k = (ib + 1) * FB + g; k < n_pad; k += 16) acc = fma(L[(size_t)k * ld + ib * FB + c], alpha_out[k], acc);
      part[g][c] = acc;
    }
    for (int e = tid; e < FB * FB; e += 1024) Ld[e >> 6][e & 63] = L[(size_t)(ib * FB + (e >> 6)) * ld + ib * FB + (e & 63)];
    __syncthreads();
    if (tid < FB) {
      double s = 0.0;
      for (int g = 0; g < 16; ++g) s += part[g][tid];
      r[tid] = tvec[ib * FB + tid] - s;
    }
    __syncthreads();
    if (tid < 32) {
      for (int j = FB - 1; j >= 0; --j) {
        double p = 0.0;
        if (lane > j) p = Ld[lane][j] * r[lane];
        if (lane + 32 > j) p = fma(Ld[lane + 32][j], r[lane + 32], p);
        for (int off = 16; off > 0; off >>= 1) p += __shfl_xor_sync(0xffffffffu, p, off);
        if (lane == 0) r[j] = (r[j] - p) / Ld[j][j];
        __syncwarp();
      }
      alpha_out[ib * FB + lane] = (ib * FB + lane < n) ? r[lane] : 0.0;
      alpha_out[ib * FB + lane + 32] = (ib * FB + lane + 32 < n) ? r[lane + 32] : 0.0;
    }
    __threadfence_block();
    __syncthreads();
  }
  // lml
  double a = 0.0, b = 0.0;
  for (int i = tid; i < n; i += 1024) { a = fma(y[i], alpha_out[i], a); b += log(L[(size_t)i * ld + i]); }
  double* ra = shbuf;
  double* rb = shbuf + 1024;
  ra[tid] = a; rb[tid] = b;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) { if (tid < s) { ra[tid] += ra[tid + s]; rb[tid] += rb[tid + s]; } __syncthreads(); }
  if (tid == 0) *lml_out = -0.5 * ra[0] - rb[0] - 0.5 * (double)n * 1.8378770664093453;  // log(2 pi)
}

// Pack inv(L) (dense row-major X, ld = n_pad) into the score kernel's tile layout:
//   tile (I, J) = rows [I*SC_BM, +SC_BM) x cols [J*SC_BK, +SC_BK), stored contiguously as
//   [slot (SC_BM/8)][k8 (SC_BK/8)][8 rows][8 cols]; slot = wm*SC_FM + t holds the 8 rows of original
//   row fragment f = SC_WM*t + (t odd ? SC_WM-1-wm : wm): row fragments are dealt to the SC_WM warp rows
//   back and forth, so every warp row owns the same amount of work inside the triangular diagonal block.
// Rows/cols >= n are written as zeros.
__global__ void pack_linv_kernel(const double* __restrict__ X, int n, int n_pad, double* __restrict__ out) {
  const int64_t total = (int64_t)n_pad * n_pad;
  const int nJ = n_pad / SC_BK;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t rem = e;
    const int col8 = rem & 7; rem >>= 3;
    const int row8 = rem & 7; rem >>= 3;
    const int k8 = rem % (SC_BK / 8); rem /= (SC_BK / 8);
    const int slot = rem % (SC_BM / 8); rem /= (SC_BM / 8);
    const int J = rem % nJ;
    const int I = (int)(rem / nJ);
    const int wm = slot / SC_FM, t = slot % SC_FM;
    const int row = I * SC_BM + (SC_WM * t + ((t & 1) ? SC_WM - 1 - wm : wm)) * 8 + row8;
    const int col = J * SC_BK + k8 * 8 + col8;
    out[e] = (row < n && col < n && col <= row) ? X[(size_t)row * n_pad + col] : 0.0;
  }
}

// Pack Xs into DMMA fragment order: [obs8][d4][8 obs][4 dims]
__global__ void pack_xs_kernel(const double* __restrict__ Xs, int n_pad, int D_pad, double* __restrict__ out) {
  const int64_t total = (int64_t)n_pad * D_pad;
  const int nd4 = D_pad / 4;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t rem = e;
    const int dq = rem & 3; rem >>= 2;
    const int o8 = rem & 7; rem >>= 3;
    const int d4 = rem % nd4;
    const int ob = (int)(rem / nd4);
    out[e] = Xs[(size_t)(ob * 8 + o8) * D_pad + d4 * 4 + dq];
  }
}

__global__ void write_header_kernel(ob200_fit_header* hdr_dev, ob200_fit_header h, const double* stats, const double* lml,
                                    double* ls_out, const double* ls, int n_ls, int D, int D_pad) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    h.y_mean = stats[0]; h.y_std = stats[1]; h.y_min = stats[2]; h.lml = *lml;
    *hdr_dev = h;
  }
  for (int d = threadIdx.x; d < D_pad; d += blockDim.x) ls_out[d] = d < D ? ls[n_ls == 1 ? 0 : d] : 1.0;
}

// ------------------------------------------------------------------------------------------
// Gradient of the log marginal likelihood w.r.t. log(amplitude), log(length_scale_d) and the diagonal
// term (sklearn/gaussian_process/_gpr.py:541-656: 0.5 tr((alpha alpha^T - K^-1) dK/dtheta)).
//   W = alpha alpha^T - K^-1 ;  T_ij = W_ij * g(r_ij) with dk/dlog(ls_d) = g(r) * (dx_d/ls_d)^2
//   d/dlog(ls_d)  = sum_i t_i xs_id^2 - sum_i xs_id (T Xs)_id      (t = row sums of T, Xs = X/ls)
//   d/dlog(amp)   = 0.5 sum_ij W_ij k_f(r_ij) ;  d/d(diag) = 0.5 tr(W)
__global__ void transpose_kernel(const double* __restrict__ A, int n_pad, double* __restrict__ At) {
  __shared__ double t[32][33];
  const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += 8) t[r][threadIdx.x] = A[(size_t)(by + r) * n_pad + bx + threadIdx.x];
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += 8) At[(size_t)(bx + r) * n_pad + by + threadIdx.x] = t[threadIdx.x][r];
}

// one 32x32 tile (tj <= ti): T both triangles, per-tile partial of sum W k_f (off-diagonal tiles count twice)
__global__ void __launch_bounds__(256) lml_t_kernel(const double* __restrict__ Xs, const double* __restrict__ alpha,
                                                    const double* __restrict__ Kinv, int n, int n_pad, int D_pad, int kind,
                                                    double amplitude, double* __restrict__ T, double* __restrict__ amp_part) {
  const int ti = blockIdx.y, tj = blockIdx.x;
  const int tile_id = ti * gridDim.x + tj;
  __shared__ double red[256];
  if (tj > ti) { if (threadIdx.x == 0) amp_part[tile_id] = 0.0; return; }
  extern __shared__ double sm[];
  double* xi = sm;
  double* xj = sm + 32 * (D_pad + 1);
  const int ldx = D_pad + 1;
  for (int e = threadIdx.x; e < 32 * D_pad; e += 256) {
    const int r = e / D_pad, d = e - r * D_pad;
    xi[r * ldx + d] = Xs[(size_t)(ti * 32 + r) * D_pad + d];
    xj[r * ldx + d] = Xs[(size_t)(tj * 32 + r) * D_pad + d];
  }
  __syncthreads();
  double part = 0.0;
  for (int e = threadIdx.x; e < 1024; e += 256) {
    const int r = e >> 5, c = e & 31;
    const int i = ti * 32 + r, j = tj * 32 + c;
    double tv = 0.0;
    if (i < n && j < n && (ti != tj || j <= i)) {
      double d2 = 0.0;
      for (int d = 0; d < D_pad; ++d) { const double df = xi[r * ldx + d] - xj[c * ldx + d]; d2 = fma(df, df, d2); }
      const double w = alpha[i] * alpha[j] - Kinv[(size_t)i * n_pad + j];
      double kf, g;
      if (kind == OB200_KERNEL_MATERN52) {
        const double k5 = sqrt(d2) * 2.23606797749979, ex = exp(-k5);
        kf = amplitude * (1.0 + k5 + k5 * k5 / 3.0) * ex;
        g = amplitude * (5.0 / 3.0) * (1.0 + k5) * ex;
      } else {
        kf = amplitude * exp(-0.5 * d2);
        g = kf;
      }
      tv = w * g;
      part += (i == j ? 1.0 : 2.0) * w * kf;
    }
    if (i < n_pad && j < n_pad) {
      if (ti != tj || j <= i) {
        T[(size_t)i * n_pad + j] = tv;
        if (i != j) T[(size_t)j * n_pad + i] = tv;
      }
    }
  }
  red[threadIdx.x] = part;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) { if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s]; __syncthreads(); }
  if (threadIdx.x == 0) amp_part[tile_id] = red[0];
}

// one warp per row i: t_i = sum_j T_ij and P[i][d] = sum_j T_ij Xs[j][d]; then the row's contribution
// c[i][d] = t_i xs_id^2 - xs_id P_id (deterministic: fixed lane striding, shuffle tree)
__global__ void __launch_bounds__(256) lml_row_kernel(const double* __restrict__ T, const double* __restrict__ Xs, int n, int n_pad,
                                                      int D_pad, double* __restrict__ contrib) {
  const int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n) return;
  const double* Tr = T + (size_t)warp * n_pad;
  double t = 0.0;
  for (int j = lane; j < n; j += 32) t += Tr[j];
  for (int off = 16; off > 0; off >>= 1) t += __shfl_xor_sync(0xffffffffu, t, off);
  for (int d0 = 0; d0 < D_pad; d0 += 4) {
    double p0 = 0, p1 = 0, p2 = 0, p3 = 0;
    for (int j = lane; j < n; j += 32) {
      const double tv = Tr[j];
      const double* x = Xs + (size_t)j * D_pad + d0;
      p0 = fma(tv, x[0], p0); p1 = fma(tv, x[1], p1); p2 = fma(tv, x[2], p2); p3 = fma(tv, x[3], p3);
    }
    for (int off = 16; off > 0; off >>= 1) {
      p0 += __shfl_xor_sync(0xffffffffu, p0, off); p1 += __shfl_xor_sync(0xffffffffu, p1, off);
      p2 += __shfl_xor_sync(0xffffffffu, p2, off); p3 += __shfl_xor_sync(0xffffffffu, p3, off);
    }
    if (lane == 0) {
      const double* xi = Xs + (size_t)warp * D_pad + d0;
      double* c = contrib + (size_t)warp * D_pad + d0;
      c[0] = t * xi[0] * xi[0] - xi[0] * p0; c[1] = t * xi[1] * xi[1] - xi[1] * p1;
      c[2] = t * xi[2] * xi[2] - xi[2] * p2; c[3] = t * xi[3] * xi[3] - xi[3] * p3;
    }
  }
}

// out[0] = d/dlog(amp), out[1 + d] = d/dlog(ls_d) (d < D_pad), out[1 + D_pad] = d/d(diag)
__global__ void __launch_bounds__(256) lml_final_kernel(const double* __restrict__ contrib, const double* __restrict__ amp_part,
                                                        int n_tiles, const double* __restrict__ alpha, const double* __restrict__ Kinv,
                                                        int n, int n_pad, int D_pad, double* __restrict__ out) {
  __shared__ double red[256];
  const int which = blockIdx.x;  // 0: amp, 1..D_pad: ls, D_pad+1: diag
  double acc = 0.0;
  if (which == 0) {
    for (int t = threadIdx.x; t < n_tiles; t += 256) acc += amp_part[t];
    acc *= 0.5;
  } else if (which <= D_pad) {
    for (int i = threadIdx.x; i < n; i += 256) acc += contrib[(size_t)i * D_pad + which - 1];
  } else {
    for (int i = threadIdx.x; i < n; i += 256) acc += alpha[i] * alpha[i] - Kinv[(size_t)i * n_pad + i];
    acc *= 0.5;
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) { if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s]; __syncthreads(); }
  if (threadIdx.x == 0) out[which] = red[0];
}

// ---- posterior mean / std-dev gradients at a few points (acquisition refinement) ------------------------------
// k_j = amp f(r_j), dk_j/dx_d = -g_j u_jd / ls_d with u_jd = (x_d - X_jd) / ls_d;
// Matern 5/2: g = amp (5/3)(1 + sqrt5 r) exp(-sqrt5 r) (no 1/r singularity);  RBF: g = k.
__global__ void __launch_bounds__(256) pg_k_kernel(const double* __restrict__ x, int ldx, const double* __restrict__ ls,
                                                   const double* __restrict__ Xs, int n, int n_pad, int D, int D_pad, int kind,
                                                   double amplitude, double* __restrict__ xs_out, double* __restrict__ Kv,
                                                   double* __restrict__ Gv) {
  extern __shared__ double pg_xs[];
  const int p = blockIdx.y;
  for (int d = threadIdx.x; d < D_pad; d += blockDim.x) {
    const double v = d < D ? x[(size_t)p * ldx + d] / ls[d] : 0.0;  // the same quotient as scale_x_kernel
    pg_xs[d] = v;
    if (blockIdx.x == 0) xs_out[(size_t)p * D_pad + d] = v;
  }
  __syncthreads();
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_pad) return;
  double k = 0.0, g = 0.0;
  if (j < n) {
    double r2 = 0.0;
    for (int d = 0; d < D_pad; ++d) { const double u = pg_xs[d] - Xs[(size_t)j * D_pad + d]; r2 = fma(u, u, r2); }
    if (kind == OB200_KERNEL_MATERN52) {
      const double t = sqrt(r2) * 2.23606797749979, e = exp(-t);
      k = amplitude * ((1.0 + t + t * t / 3.0) * e);
      g = amplitude * (5.0 / 3.0) * (1.0 + t) * e;
    } else {
      k = amplitude * exp(-0.5 * r2);
      g = k;
    }
  }
  Kv[(size_t)p * n_pad + j] = k;
  Gv[(size_t)p * n_pad + j] = g;
}

// v = inv(L) k: one warp per row, lanes stride the columns (coalesced), fixed shuffle tree.
__global__ void __launch_bounds__(256) pg_v_kernel(const double* __restrict__ Xinv, const double* __restrict__ Kv, int n_pad,
                                                   double* __restrict__ V) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31, p = blockIdx.y;
  if (i >= n_pad) return;
  const double* row = Xinv + (size_t)i * n_pad;
  const double* kv = Kv + (size_t)p * n_pad;
  double acc = 0.0;
  for (int j = lane; j <= i; j += 32) acc = fma(row[j], kv[j], acc);
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) V[(size_t)p * n_pad + i] = acc;
}

// w = inv(L)^T v = K^-1 k: 8 columns per CTA (n_pad / 8 CTAs: the 32-column version left half the SMs idle and walked
// 264 dependent steps per thread), 32 row groups walking down the rows two at a time (each row segment is one 64 B
// read), fixed-order reduction over the groups.
#define PG_W_COLS 8
__global__ void __launch_bounds__(256) pg_w_kernel(const double* __restrict__ Xinv, const double* __restrict__ V, int n_pad,
                                                   double* __restrict__ Wv) {
  __shared__ double part[32][PG_W_COLS + 1];
  const int c = threadIdx.x & (PG_W_COLS - 1), g = threadIdx.x / PG_W_COLS, j0 = blockIdx.x * PG_W_COLS, j = j0 + c, p = blockIdx.y;
  const double* v = V + (size_t)p * n_pad;
  double acc0 = 0.0, acc1 = 0.0;
  if (j < n_pad) {
    int i = j0 + g;
    for (; i + 32 < n_pad; i += 64) {
      const double x0 = Xinv[(size_t)i * n_pad + j], x1 = Xinv[(size_t)(i + 32) * n_pad + j];
      if (i >= j) acc0 = fma(x0, v[i], acc0);
      acc1 = fma(x1, v[i + 32], acc1);   // i + 32 >= j0 + 32 > j
    }
    if (i < n_pad && i >= j) acc0 = fma(Xinv[(size_t)i * n_pad + j], v[i], acc0);
  }
  part[g][c] = acc0 + acc1;
  __syncthreads();
  if (g == 0 && j < n_pad) {
    double t = 0.0;
    for (int k = 0; k < 32; ++k) t += part[k][c];
    Wv[(size_t)p * n_pad + j] = t;
  }
}

__device__ __forceinline__ double pg_block_sum(double v, double* red) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
  for (int w = 0; w < 8; ++w) t += red[w];
  return t;
}

// out[p] = (mu, sigma, dmu/dx[D], dsigma/dx[D]) in the units of ob200_score's mu / sigma.
__global__ void __launch_bounds__(256) pg_final_kernel(const double* __restrict__ xs, const double* __restrict__ ls,
                                                       const double* __restrict__ Xs, const double* __restrict__ alpha,
                                                       const double* __restrict__ Kv, const double* __restrict__ Gv,
                                                       const double* __restrict__ V, const double* __restrict__ Wv, int n, int n_pad,
                                                       int D, int D_pad, double amplitude, double y_mean, double y_std,
                                                       double* __restrict__ out) {
  __shared__ double red[8];
  const int p = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const double *kv = Kv + (size_t)p * n_pad, *gv = Gv + (size_t)p * n_pad, *v = V + (size_t)p * n_pad, *wv = Wv + (size_t)p * n_pad;
  double m = 0.0, s2 = 0.0;
  for (int j = threadIdx.x; j < n_pad; j += 256) { m = fma(kv[j], alpha[j], m); s2 = fma(v[j], v[j], s2); }
  m = pg_block_sum(m, red);
  s2 = pg_block_sum(s2, red);
  double var = amplitude - s2;
  if (var < 0.0) var = 0.0;
  const double sd = sqrt(var);
  double* o = out + (size_t)p * (2 + 2 * D);
  if (threadIdx.x == 0) { o[0] = y_std * m + y_mean; o[1] = y_std * sd; }
  for (int d = warp; d < D; d += 8) {
    const double xd = xs[(size_t)p * D_pad + d];
    double a1 = 0.0, a2 = 0.0;
    for (int j = lane; j < n; j += 32) {
      const double gu = gv[j] * (xd - Xs[(size_t)j * D_pad + d]);
      a1 = fma(alpha[j], gu, a1);
      a2 = fma(wv[j], gu, a2);
    }
    for (int s = 16; s > 0; s >>= 1) { a1 += __shfl_xor_sync(0xffffffffu, a1, s); a2 += __shfl_xor_sync(0xffffffffu, a2, s); }
    if (lane == 0) {
      o[2 + d] = -y_std * a1 / ls[d];
      o[2 + D + d] = var > 0.0 ? y_std * (a2 / ls[d]) / sd : 0.0;  // d sqrt(var) = dvar / (2 sd), dvar = 2 a2 / ls
    }
  }
}

// ---- rank-1 append of one observation to the current fit (ob200_gp_append) --------------------------------------
// new row of K against the n old points (direct differences, like gram_kernel), X/ls row and its squared norm
__global__ void __launch_bounds__(256) append_k_kernel(const double* __restrict__ x_new, const double* __restrict__ ls, int D,
                                                       int D_pad, int n, int n_pad, int kind, double amplitude,
                                                       double* __restrict__ Xs, double* __restrict__ xx, double* __restrict__ kvec) {
  extern __shared__ double ap_xs[];
  for (int d = threadIdx.x; d < D_pad; d += blockDim.x) ap_xs[d] = d < D ? x_new[d] / ls[d] : 0.0;
  __syncthreads();
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n_pad) {
    double k = 0.0;
    if (j < n) {
      double d2 = 0.0;
      for (int d = 0; d < D_pad; ++d) { const double df = ap_xs[d] - Xs[(size_t)j * D_pad + d]; d2 = fma(df, df, d2); }
      k = kernel_from_d2(kind, d2, amplitude);
    }
    kvec[j] = k;
  }
  if (blockIdx.x == 0) {
    __syncthreads();
    double acc = 0.0;
    for (int d = threadIdx.x; d < D_pad; d += blockDim.x) Xs[(size_t)n * D_pad + d] = ap_xs[d];
    if (threadIdx.x == 0) {
      for (int d = 0; d < D_pad; ++d) acc = fma(ap_xs[d], ap_xs[d], acc);
      xx[n] = acc;
    }
  }
}

// pivot d = sqrt(k_nn + alpha - |l|^2), new row of L; scal[0] = d, scal[1] = 1/d
__global__ void __launch_bounds__(1024) append_diag_kernel(const double* __restrict__ lvec, int n, int n_pad, double knn,
                                                           double* __restrict__ L, double* __restrict__ scal, int* __restrict__ info) {
  __shared__ double red[1024];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += 1024) { const double v = lvec[i]; acc = fma(v, v, acc); L[(size_t)n * n_pad + i] = v; }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) { if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s]; __syncthreads(); }
  if (threadIdx.x == 0) {
    const double d2 = knn - red[0];
    if (!(d2 > 0.0)) { *info = n + 1; scal[0] = 1.0; scal[1] = 1.0; }
    else { const double d = sqrt(d2); scal[0] = d; scal[1] = 1.0 / d; L[(size_t)n * n_pad + n] = d; }
  }
}

// new row of inv(L): -(l^T inv(L)) / d, 1 / d on the diagonal
__global__ void __launch_bounds__(256) append_row_kernel(const double* __restrict__ wvec, const double* __restrict__ scal, int n,
                                                         int n_pad, double* __restrict__ Xinv) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j > n) return;
  Xinv[(size_t)n * n_pad + j] = j < n ? -wvec[j] * scal[1] : scal[1];
}

__global__ void __launch_bounds__(1024) append_lml_kernel(const double* __restrict__ L, int ld, int n, const double* __restrict__ y,
                                                          const double* __restrict__ alpha, double* __restrict__ lml_out) {
  __shared__ double ra[1024], rb[1024];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < n; i += 1024) { a = fma(y[i], alpha[i], a); b += log(L[(size_t)i * ld + i]); }
  ra[threadIdx.x] = a; rb[threadIdx.x] = b;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) { if (threadIdx.x < s) { ra[threadIdx.x] += ra[threadIdx.x + s]; rb[threadIdx.x] += rb[threadIdx.x + s]; } __syncthreads(); }
  if (threadIdx.x == 0) *lml_out = -0.5 * ra[0] - rb[0] - 0.5 * (double)n * 1.8378770664093453;  // log(2 pi)
}

// r = y - q ;  a += d   (vectors of n_pad entries)
__global__ void vec_sub_kernel(const double* __restrict__ y, const double* __restrict__ q, int n, double* __restrict__ r) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) r[i] = y[i] - q[i];
}
__global__ void vec_add_kernel(double* __restrict__ a, const double* __restrict__ d, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] += d[i];
}

// alpha = K^-1 y from the explicit inverse factor, a = inv(L)^T (inv(L) y), followed by `refine` steps of iterative
// refinement with the residual taken through the Cholesky factor itself, r = y - L (L^T a), a += inv(L)^T (inv(L) r).
// The products with inv(L) carry an error ~cond(L) eps; each refinement step multiplies it by that factor again, so
// two steps reach the accuracy of the substitution-based solve (cho_solve) at a fraction of its serial latency.
// Scratch: 3 vectors of n_pad doubles.  All reductions have a fixed order.
int solve_alpha(ob200_ctx* ctx, const double* Xinv, const double* L, const double* y, int n_pad, double* alpha, double* scratch,
                int refine) {
  cudaStream_t st = ctx->stream;
  double *t = scratch, *r = scratch + n_pad, *d = scratch + 2 * (size_t)n_pad;
  const dim3 gv((n_pad * 32 + 255) / 256, 1), gw((n_pad + PG_W_COLS - 1) / PG_W_COLS, 1);
  const int gb = (n_pad + 255) / 256;
  pg_v_kernel<<<gv, 256, 0, st>>>(Xinv, y, n_pad, t);
  OB_LAUNCH_CHECK(ctx);
  pg_w_kernel<<<gw, 256, 0, st>>>(Xinv, t, n_pad, alpha);
  OB_LAUNCH_CHECK(ctx);
  for (int it = 0; it < refine; ++it) {
    pg_w_kernel<<<gw, 256, 0, st>>>(L, alpha, n_pad, t);   // t = L^T a
    OB_LAUNCH_CHECK(ctx);
    pg_v_kernel<<<gv, 256, 0, st>>>(L, t, n_pad, d);       // d = L t = K a
    OB_LAUNCH_CHECK(ctx);
    vec_sub_kernel<<<gb, 256, 0, st>>>(y, d, n_pad, r);
    OB_LAUNCH_CHECK(ctx);
    pg_v_kernel<<<gv, 256, 0, st>>>(Xinv, r, n_pad, t);
    OB_LAUNCH_CHECK(ctx);
    pg_w_kernel<<<gw, 256, 0, st>>>(Xinv, t, n_pad, d);
    OB_LAUNCH_CHECK(ctx);
    vec_add_kernel<<<gb, 256, 0, st>>>(alpha, d, n_pad);
    OB_LAUNCH_CHECK(ctx);
  }
  return OB200_OK;
}

int launch_gemm(ob200_ctx* ctx, bool trans_b, const gemm_args& g, int batch, cudaStream_t stream = nullptr) {
  dim3 grid(g.N / 64, g.M / 64, batch);
  if (grid.x == 0 || grid.y == 0 || batch == 0) return OB200_OK;
  cudaStream_t st = stream ? stream : ctx->stream;
  if (trans_b) gemm64_kernel<true><<<grid, 128, 0, st>>>(g);
  else gemm64_kernel<false><<<grid, 128, 0, st>>>(g);
  OB_LAUNCH_CHECK(ctx);
  return OB200_OK;
}

}  // namespace

extern "C" {

int ob200_gp_state_bytes(ob200_ctx* ctx, int n, int D, int64_t* bytes) {
  if (!ctx || !bytes) return OB200_EINVAL;
  if (n < 1 || n > OB200_MAX_N || D < 1) OB_FAIL(ctx, OB200_EINVAL, "n must be in [1, %d], D >= 1", OB200_MAX_N);
  const int n_pad = (int)ob_round_up(n, SC_BM), D_pad = (int)ob_round_up(D, 4);
  *bytes = ob_blob_layout(n_pad, D_pad).total;
  return OB200_OK;
}

}  // extern "C"

namespace {
// search = 1: the fit of one log-marginal-likelihood evaluation of a hyper-parameter search -- L, inv(L), alpha and
// the LML are produced (everything ob200_gp_lml / ob200_gp_lml_grad read), the Newton-Schulz refinement of inv(L) and
// the tile packing for the score kernel are skipped, and the context is left NOT scoreable.
int fit_impl(ob200_ctx* ctx, const double* X_dev, const double* y_dev, int n, int D, const ob200_kernel* kernel,
             double alpha, int normalize_y, int search) {
  if (!ctx) return OB200_EINVAL;
  if (!X_dev || !y_dev || !kernel) OB_FAIL(ctx, OB200_EINVAL, "null argument");
  if (n < 1 || n > OB200_MAX_N) OB_FAIL(ctx, OB200_EINVAL, "n_obs=%d outside [1, %d]", n, OB200_MAX_N);
  if (D < 1 || D > 384) OB_FAIL(ctx, OB200_EINVAL, "D=%d outside [1, 384]", D);
  if (kernel->kind != OB200_KERNEL_MATERN52 && kernel->kind != OB200_KERNEL_RBF) OB_FAIL(ctx, OB200_EINVAL, "unknown kernel kind");
  if (kernel->n_length_scale != 1 && kernel->n_length_scale != D) OB_FAIL(ctx, OB200_EINVAL, "length_scale must have 1 or D entries");
  if (!kernel->length_scale_host) OB_FAIL(ctx, OB200_EINVAL, "length_scale_host is null");
  for (int i = 0; i < kernel->n_length_scale; ++i)
    if (!(kernel->length_scale_host[i] > 0.0)) OB_FAIL(ctx, OB200_EINVAL, "length_scale[%d] must be > 0", i);
  if (!(kernel->amplitude > 0.0) || !(alpha >= 0.0)) OB_FAIL(ctx, OB200_EINVAL, "amplitude must be > 0 and alpha >= 0");
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->fitted = false;
  ctx->have_local_fit = false;

  const int n_pad = (int)ob_round_up(n, SC_BM), D_pad = (int)ob_round_up(D, 4);
  const int nb = n_pad / FB;
  const ob200_blob_layout lay = ob_blob_layout(n_pad, D_pad);
  int rc;
  if ((rc = ob_reserve(ctx, &ctx->blob, &ctx->blob_cap, lay.total))) return rc;
  if ((rc = ob_reserve(ctx, &ctx->L_dev, &ctx->L_cap, (int64_t)n_pad * n_pad))) return rc;
  // workspace: X (n_pad^2) | T (n_pad^2 / 2) | Dinv (nb*64*64) | Xs (n_pad*D_pad) | y_norm | tvec | ls | stats(4) | info
  const int64_t w_X = 0, w_T = w_X + (int64_t)n_pad * n_pad, w_X2 = w_T + (int64_t)n_pad * n_pad, w_Dinv = w_X2 + (int64_t)n_pad * n_pad,
                w_Lf = w_Dinv + (int64_t)nb * FB * FB, w_Xs = w_Lf + (int64_t)nb * FB * FB, w_y = w_Xs + (int64_t)n_pad * D_pad, w_t = w_y + n_pad,
                w_ls = w_t + n_pad, w_stats = w_ls + D_pad, w_Rd = w_stats + 8, w_end = w_Rd + n_pad;
  if ((rc = ob_reserve(ctx, &ctx->work_dev, &ctx->work_cap, w_end))) return rc;
  double* W = ctx->work_dev;
  double *Xinv = W + w_X, *T = W + w_T, *X2 = W + w_X2, *Dinv = W + w_Dinv, *Lfac = W + w_Lf, *Xs = W + w_Xs, *ynorm = W + w_y, *tvec = W + w_t,
         *ls_dev = W + w_ls, *stats = W + w_stats, *Rd = W + w_Rd;
  int* info_dev = reinterpret_cast<int*>(stats + 6);
  double* lml_dev = stats + 4;
  cudaStream_t st = ctx->stream;
  char* blob = ctx->blob;
  double* b_ls = reinterpret_cast<double*>(blob + lay.off_ls);
  double* b_xs = reinterpret_cast<double*>(blob + lay.off_xs);
  double* b_xx = reinterpret_cast<double*>(blob + lay.off_xx);
  double* b_alpha = reinterpret_cast<double*>(blob + lay.off_alpha);
  double* b_linv = reinterpret_cast<double*>(blob + lay.off_linv);

  OB_CUDA(ctx, cudaMemcpyAsync(ls_dev, kernel->length_scale_host, (size_t)kernel->n_length_scale * 8, cudaMemcpyHostToDevice, st));
  OB_CUDA(ctx, cudaMemsetAsync(stats, 0, 8 * sizeof(double), st));

  // OB200_FIT_TIMING=1: CUDA-event time of each phase on stderr (tuning aid)
  const bool timing = getenv("OB200_FIT_TIMING") != nullptr;
  cudaEvent_t tev[9];
  int n_tev = 0;
  auto mark = [&]() { if (timing && n_tev < 8) { cudaEventCreate(&tev[n_tev]); cudaEventRecord(tev[n_tev++], st); } };
  mark();
  ystats_kernel<<<1, 1024, 0, st>>>(y_dev, n, n_pad, normalize_y, ynorm, stats);
  OB_LAUNCH_CHECK(ctx);
  scale_x_kernel<<<(n_pad + 127) / 128, 128, 0, st>>>(X_dev, n, D, n_pad, D_pad, ls_dev, kernel->n_length_scale, Xs, b_xx);
  OB_LAUNCH_CHECK(ctx);
  {
    const size_t smem = (size_t)2 * 32 * (D_pad + 1) * sizeof(double);
    OB_CUDA(ctx, cudaFuncSetAttribute(gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem > 49152 ? smem : 49152)));
    dim3 grid(n_pad / 32, n_pad / 32);
    gram_kernel<<<grid, 256, smem, st>>>(Xs, n, n_pad, D_pad, kernel->kind, kernel->amplitude, alpha, ctx->L_dev);
    OB_LAUNCH_CHECK(ctx);
  }
  mark();
  // ---- right-looking blocked Cholesky with a one-step look-ahead --------------------------------
  // Per 64-column step: ONE launch factors the diagonal block and solves the panel; the trailing update A22 -= P P^T is
  // split into (a) the next block column, on the main stream -- all the next step needs -- and (b) the rest, on a second
  // stream, where it overlaps the next step's factor + panel.  (a) of step k + 1 waits for (b) of step k (they
  // accumulate into the same tiles, in a fixed order).  (Fusing (a) into the next step's launch was tried: the update
  // inside the two-warp-wide step kernel cost more than the launch it saved, 1.66 -> 1.82 ms at n = 2000.)
  double* Lm = ctx->L_dev;
  const size_t diag_smem = (size_t)2 * FB * (FB + 2) * sizeof(double);
  OB_CUDA(ctx, cudaFuncSetAttribute(chol_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)diag_smem));
  const bool compact_step = getenv("OB200_FIT_UNROLLED_STEP") == nullptr;   // (testing hook: the fully unrolled step kernel of round 2a)
  // Step schedule: below ~3000 observations the fused step (diagonal block factored inside every CTA of the step, two
  // launches per step on the main stream) wins by a launch and two cross-stream hops per step (n = 2000: 1.61 against
  // 1.68 ms); above, where the block column's GEMM is long enough to hide the diagonal kernel, the split step does
  // (n = 4000: 3.57 against 3.79 ms, n = 8000: 12.3 against 13.6).  (Testing hooks force either.)
  bool split_step = compact_step && n_pad >= 3000;
  if (getenv("OB200_FIT_FUSED_STEP")) split_step = false;
  if (getenv("OB200_FIT_SPLIT_STEP")) split_step = compact_step;
  if (!ctx->fit_stream) {
    OB_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->fit_stream, cudaStreamNonBlocking));
    OB_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->fit_stream2, cudaStreamNonBlocking));
    OB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_panel, cudaEventDisableTiming));
    OB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_trail, cudaEventDisableTiming));
    OB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_diag, cudaEventDisableTiming));
  }
  cudaStream_t side = ctx->fit_stream, side2 = ctx->fit_stream2;
  const bool lookahead = getenv("OB200_FIT_NO_LOOKAHEAD") == nullptr;   // (testing hook: everything on one stream)
  bool trail_pending = false;
  const size_t step_smem = compact_step ? ((size_t)FB * (FB + 2) + (size_t)FB * (FB + 1)) * sizeof(double) : diag_smem;
  if (compact_step) {
    OB_CUDA(ctx, cudaFuncSetAttribute(chol_step_compact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)step_smem));
    OB_CUDA(ctx, cudaFuncSetAttribute(chol_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)step_smem));
    OB_CUDA(ctx, cudaFuncSetAttribute(chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)step_smem));
  }
  if (split_step) {
    // Per 64-column step: panel(kb) | then, side by side, diag(kb + 1) on its own stream (one CTA: the diagonal tile's
    // update and factorisation) and GEMM (a') = the rest of block column kb + 1 on the main stream | panel(kb + 1).
    // (b), the trailing update from block kb + 2 on, runs on the second stream as before; diag(kb + 1) and (a') of step
    // kb wait for (b) of step kb - 1, which was the last to write their tiles.
    chol_diag_kernel<<<1, 256, step_smem, st>>>(Lm, n_pad, 0, Lfac, Rd, info_dev);
    OB_LAUNCH_CHECK(ctx);
    for (int kb = 0; kb < nb; ++kb) {
      const int rem = n_pad - (kb + 1) * FB;
      chol_panel_kernel<<<1 + (rem > 0 ? rem / FB : 0), 256, step_smem, st>>>(Lm, n_pad, kb, Dinv, Lfac, Rd, info_dev);
      OB_LAUNCH_CHECK(ctx);
      if (rem <= 0) break;
      const double* panel = Lm + (size_t)(kb + 1) * FB * n_pad + (size_t)kb * FB;
      gemm_args a{};   // (a') block column kb + 1 below its diagonal tile: C[(kb+2):, kb+1] -= P[(kb+2):] P[kb+1]^T
      a.A = panel + (size_t)FB * n_pad; a.lda = n_pad; a.B = panel; a.ldb = n_pad;
      a.C = Lm + (size_t)(kb + 2) * FB * n_pad + (size_t)(kb + 1) * FB; a.ldc = n_pad;
      a.M = rem - FB; a.N = FB; a.K = FB; a.alpha = -1.0; a.beta = 1.0; a.lower_only = 0;
      gemm_args b = a;  // (b) the rest: rows and columns from block kb + 2
      b.B = b.A;
      b.C = Lm + (size_t)(kb + 2) * FB * n_pad + (size_t)(kb + 2) * FB;
      b.N = rem - FB; b.lower_only = 1;
      if (lookahead) {
        OB_CUDA(ctx, cudaEventRecord(ctx->ev_panel, st));
        OB_CUDA(ctx, cudaStreamWaitEvent(side2, ctx->ev_panel, 0));
        if (trail_pending) {   // (b) of the previous step
          OB_CUDA(ctx, cudaStreamWaitEvent(side2, ctx->ev_trail, 0));
          OB_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_trail, 0));
        }
        chol_diag_kernel<<<1, 256, step_smem, side2>>>(Lm, n_pad, kb + 1, Lfac, Rd, info_dev);
        OB_LAUNCH_CHECK(ctx);
        OB_CUDA(ctx, cudaEventRecord(ctx->ev_diag, side2));
        if ((rc = launch_gemm(ctx, true, a, 1))) return rc;
        if (b.M > 0) {
          OB_CUDA(ctx, cudaStreamWaitEvent(side, ctx->ev_panel, 0));
          if ((rc = launch_gemm(ctx, true, b, 1, side))) return rc;
          OB_CUDA(ctx, cudaEventRecord(ctx->ev_trail, side));
          trail_pending = true;
        } else {
          trail_pending = false;
        }
        OB_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_diag, 0));
      } else {
        chol_diag_kernel<<<1, 256, step_smem, st>>>(Lm, n_pad, kb + 1, Lfac, Rd, info_dev);
        OB_LAUNCH_CHECK(ctx);
        if ((rc = launch_gemm(ctx, true, a, 1))) return rc;
        if (b.M > 0 && (rc = launch_gemm(ctx, true, b, 1))) return rc;
      }
    }
  } else
  for (int kb = 0; kb < nb; ++kb) {
    const int rem = n_pad - (kb + 1) * FB;
    if (compact_step) chol_step_compact_kernel<<<1 + (rem > 0 ? rem / FB : 0), 256, step_smem, st>>>(Lm, n_pad, kb, Dinv, Lfac, info_dev);
    else chol_step_kernel<<<1 + (rem > 0 ? rem / FB : 0), 256, diag_smem, st>>>(Lm, n_pad, kb, Dinv, Lfac, info_dev);
    OB_LAUNCH_CHECK(ctx);
    if (rem <= 0) break;
    const double* panel = Lm + (size_t)(kb + 1) * FB * n_pad + (size_t)kb * FB;
    gemm_args a{};   // (a) block column kb + 1: C[(kb+1):, kb+1] -= P P[0:64]^T
    a.A = panel; a.lda = n_pad; a.B = panel; a.ldb = n_pad;
    a.C = Lm + (size_t)(kb + 1) * FB * n_pad + (size_t)(kb + 1) * FB; a.ldc = n_pad;
    a.M = rem; a.N = FB; a.K = FB; a.alpha = -1.0; a.beta = 1.0; a.lower_only = 1;
    gemm_args b = a;  // (b) the rest: rows and columns from block kb + 2
    b.A = panel + (size_t)FB * n_pad; b.B = b.A;
    b.C = Lm + (size_t)(kb + 2) * FB * n_pad + (size_t)(kb + 2) * FB;
    b.M = rem - FB; b.N = rem - FB;
    if (lookahead && b.M > 0) {
      OB_CUDA(ctx, cudaEventRecord(ctx->ev_panel, st));
      if (trail_pending) OB_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_trail, 0));   // (b) of the previous step
      if ((rc = launch_gemm(ctx, true, a, 1))) return rc;
      OB_CUDA(ctx, cudaStreamWaitEvent(side, ctx->ev_panel, 0));
      if ((rc = launch_gemm(ctx, true, b, 1, side))) return rc;
      OB_CUDA(ctx, cudaEventRecord(ctx->ev_trail, side));
      trail_pending = true;
    } else {
      if (trail_pending) { OB_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_trail, 0)); trail_pending = false; }
      if ((rc = launch_gemm(ctx, true, a, 1))) return rc;
      if (b.M > 0 && (rc = launch_gemm(ctx, true, b, 1))) return rc;
    }
  }
  if (trail_pending) OB_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_trail, 0));
  chol_copy_diag_kernel<<<nb, 256, 0, st>>>(Lm, n_pad, Lfac);
  OB_LAUNCH_CHECK(ctx);
  // The pivot check is read back with the header at the end of the fit (one host round trip per fit instead of two);
  // a failed factorisation only costs the NaN-filled remainder of a fit that is discarded anyway.
  int info = 0;
  OB_CUDA(ctx, cudaMemcpyAsync(&info, info_dev, sizeof(int), cudaMemcpyDeviceToHost, st));

  const bool serial_solve = getenv("OB200_SERIAL_SOLVE") != nullptr;  // (testing hook: the substitution-based solve)
  if (serial_solve) {
    chol_solve_kernel<<<1, 1024, 0, st>>>(Lm, n_pad, n, n_pad, ynorm, tvec, b_alpha, lml_dev);
    OB_LAUNCH_CHECK(ctx);
  }

  mark();
  // ---- inv(L) by recursive doubling ----------------------------------------------------------
  OB_CUDA(ctx, cudaMemsetAsync(Xinv, 0, (size_t)n_pad * n_pad * 8, st));
  trtri_init_kernel<<<nb, 256, 0, st>>>(Xinv, n_pad, Dinv);
  OB_LAUNCH_CHECK(ctx);
  for (int s = FB; s < n_pad; s *= 2) {
    int pairs = 0, c_last = 0;
    for (int lo = 0; lo + s < n_pad; lo += 2 * s) { ++pairs; c_last = (lo + 2 * s <= n_pad ? s : n_pad - lo - s); }
    if (pairs == 0) break;
    const long long stride = (long long)2 * s * (n_pad + 1);
    gemm_args g1{};  // T_z = L[mid:hi, lo:mid] * X[lo:mid, lo:mid]
    g1.A = Lm + (size_t)s * n_pad; g1.lda = n_pad; g1.strideA = stride;
    g1.B = Xinv; g1.ldb = n_pad; g1.strideB = stride;
    g1.C = T; g1.ldc = s; g1.strideC = (long long)s * s;
    g1.M = s; g1.N = s; g1.K = s; g1.m_last = c_last; g1.alpha = 1.0; g1.beta = 0.0; g1.tri_k = 1;   // X[lo:mid, lo:mid] is lower triangular
    if ((rc = launch_gemm(ctx, false, g1, pairs))) return rc;
    gemm_args g2{};  // X[mid:hi, lo:mid] = -X[mid:hi, mid:hi] * T_z
    g2.A = Xinv + (size_t)s * n_pad + s; g2.lda = n_pad; g2.strideA = stride;
    g2.B = T; g2.ldb = s; g2.strideB = (long long)s * s;
    g2.C = Xinv + (size_t)s * n_pad; g2.ldc = n_pad; g2.strideC = stride;
    g2.M = s; g2.N = s; g2.K = s; g2.m_last = c_last; g2.k_eq_m = 1; g2.alpha = -1.0; g2.beta = 0.0; g2.tri_k = 2;   // so is X[mid:hi, mid:hi]
    if ((rc = launch_gemm(ctx, false, g2, pairs))) return rc;
  }
  mark();
  // One Newton-Schulz step X <- X + X (I - L X): the recursive-doubling products compound rounding errors
  // level by level (residual ~1e-6 at cond(K) = 1e12); the residual I - L X is computed to cond(L) eps, so one
  // step brings inv(L) to the accuracy of a backward-stable triangular solve (tools/illcond_probe.py).
  if (!search && !getenv("OB200_NO_NS")) {  // (testing hook: measure what the refinement buys)
    gemm_args r{};  // T = L X  (lower x lower -> lower tiles, triangular k range)
    r.A = Lm; r.lda = n_pad; r.B = Xinv; r.ldb = n_pad; r.C = T; r.ldc = n_pad;
    r.M = n_pad; r.N = n_pad; r.K = n_pad; r.alpha = 1.0; r.beta = 0.0; r.lower_only = 1; r.tri_k = 3;
    if ((rc = launch_gemm(ctx, false, r, 1))) return rc;
    eye_minus_lower_kernel<<<dim3(n_pad / 32, n_pad / 32), 256, 0, st>>>(T, n_pad);
    OB_LAUNCH_CHECK(ctx);
    OB_CUDA(ctx, cudaMemcpyAsync(X2, Xinv, (size_t)n_pad * n_pad * 8, cudaMemcpyDeviceToDevice, st));
    gemm_args u{};  // X2 = X + X T
    u.A = Xinv; u.lda = n_pad; u.B = T; u.ldb = n_pad; u.C = X2; u.ldc = n_pad;
    u.M = n_pad; u.N = n_pad; u.K = n_pad; u.alpha = 1.0; u.beta = 1.0; u.lower_only = 1; u.tri_k = 3;
    if ((rc = launch_gemm(ctx, false, u, 1))) return rc;
    OB_CUDA(ctx, cudaMemcpyAsync(Xinv, X2, (size_t)n_pad * n_pad * 8, cudaMemcpyDeviceToDevice, st));
  }
  mark();
  // ---- alpha, lml (from the refined inverse factor; T is free again and serves as scratch) -----------------------
  if (!serial_solve) {
    // (a search evaluation takes one refinement step: the objective needs ~1e-10, not the last bits)
    if ((rc = solve_alpha(ctx, Xinv, Lm, ynorm, n_pad, b_alpha, T, search ? 1 : 2))) return rc;
    append_lml_kernel<<<1, 1024, 0, st>>>(Lm, n_pad, n, ynorm, b_alpha, lml_dev);
    OB_LAUNCH_CHECK(ctx);
  }
  if (!search) {
    pack_linv_kernel<<<ctx->num_sms * 8, 256, 0, st>>>(Xinv, n, n_pad, b_linv);
    OB_LAUNCH_CHECK(ctx);
    pack_xs_kernel<<<ctx->num_sms * 2, 256, 0, st>>>(Xs, n_pad, D_pad, b_xs);
    OB_LAUNCH_CHECK(ctx);
  }
  ob200_fit_header h{};
  h.magic = OB200_FIT_MAGIC; h.n = n; h.D = D; h.n_pad = n_pad; h.D_pad = D_pad; h.kernel_kind = kernel->kind;
  h.amplitude = kernel->amplitude;
  h.diag = alpha;
  write_header_kernel<<<1, 128, 0, st>>>(reinterpret_cast<ob200_fit_header*>(blob), h, stats, lml_dev, b_ls, ls_dev,
                                         kernel->n_length_scale, D, D_pad);
  OB_LAUNCH_CHECK(ctx);
  OB_CUDA(ctx, cudaMemcpyAsync(&ctx->hdr, blob, sizeof(ob200_fit_header), cudaMemcpyDeviceToHost, st));
  OB_CUDA(ctx, cudaStreamSynchronize(st));
  if (info != 0) {
    OB_FAIL(ctx, OB200_ENOTSPD,
            "The kernel is not returning a positive definite matrix (pivot %d of %d). Try gradually increasing the 'alpha' "
            "parameter of your GaussianProcessRegressor estimator.", info, n);
  }
  if (timing) {
    static const char* names[] = {"gram", "cholesky", "trtri", "newton-schulz", "alpha+lml", "pack"};
    float ms = 0.f;
    fprintf(stderr, "ob200_gp_fit n=%d:", n);
    cudaEventCreate(&tev[n_tev]);
    cudaEventRecord(tev[n_tev], st);
    cudaEventSynchronize(tev[n_tev]);
    for (int i = 0; i < n_tev; ++i) { cudaEventElapsedTime(&ms, tev[i], tev[i + 1]); fprintf(stderr, " %s %.3f ms", names[i], ms); }
    fprintf(stderr, "\n");
    for (int i = 0; i <= n_tev; ++i) cudaEventDestroy(tev[i]);
  }
  ctx->blob_bytes = lay.total;
  ctx->fitted = !search;
  ctx->have_local_fit = true;
  ctx->fit_serial++;
  ctx->w_Xs_off = w_Xs;
  ctx->w_y_off = w_y;
  ctx->w_t_off = w_t;
  ctx->w_stats_off = w_stats;
  ctx->fit_alpha = alpha;
  return OB200_OK;
}
}  // namespace

extern "C" {

int ob200_gp_fit(ob200_ctx* ctx, const double* X_dev, const double* y_dev, int n, int D, const ob200_kernel* kernel,
                 double alpha, int normalize_y) {
  return fit_impl(ctx, X_dev, y_dev, n, D, kernel, alpha, normalize_y, 0);
}

int ob200_gp_fit_search(ob200_ctx* ctx, const double* X_dev, const double* y_dev, int n, int D, const ob200_kernel* kernel,
                        double alpha, int normalize_y) {
  return fit_impl(ctx, X_dev, y_dev, n, D, kernel, alpha, normalize_y, 1);
}

int ob200_gp_fit_host(ob200_ctx* ctx, const double* X_host, const double* y_host, int n, int D, const ob200_kernel* kernel,
                      double alpha, int normalize_y) {
  if (!ctx) return OB200_EINVAL;
  if (!X_host || !y_host || n < 1 || D < 1) OB_FAIL(ctx, OB200_EINVAL, "bad X/y/n/D");
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = ob_reserve(ctx, &ctx->cand_tmp, &ctx->cand_cap, (int64_t)n * D + n);
  if (rc) return rc;
  OB_CUDA(ctx, cudaMemcpyAsync(ctx->cand_tmp, X_host, (size_t)n * D * 8, cudaMemcpyHostToDevice, ctx->stream));
  OB_CUDA(ctx, cudaMemcpyAsync(ctx->cand_tmp + (size_t)n * D, y_host, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
  return ob200_gp_fit(ctx, ctx->cand_tmp, ctx->cand_tmp + (size_t)n * D, n, D, kernel, alpha, normalize_y);
}

int ob200_gp_lml(ob200_ctx* ctx, double* lml_host) {
  if (!ctx || !lml_host) return OB200_EINVAL;
  if (!ctx->fitted && !ctx->have_local_fit) OB_FAIL(ctx, OB200_ESTATE, "no fit");
  *lml_host = ctx->hdr.lml;
  return OB200_OK;
}

int ob200_gp_lml_grad(ob200_ctx* ctx, double* lml_host, double* grad_host, int n_grad) {
  if (!ctx || !grad_host) return OB200_EINVAL;
  if (!ctx->have_local_fit) OB_FAIL(ctx, OB200_ESTATE, "the gradient needs a fit made on this context");
  const int n = ctx->hdr.n, n_pad = ctx->hdr.n_pad, D = ctx->hdr.D, D_pad = ctx->hdr.D_pad;
  if (n_grad != D + 2) OB_FAIL(ctx, OB200_EINVAL, "n_grad must be D + 2 = %d (amplitude, D length scales, diagonal)", D + 2);
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const ob200_blob_layout lay = ob_blob_layout(n_pad, D_pad);
  const double* alpha = reinterpret_cast<const double*>(ctx->blob + lay.off_alpha);
  const double* Xinv = ctx->work_dev;                 // dense inv(L), row-major (fit workspace, offset 0)
  const double* Xs = ctx->work_dev + ctx->w_Xs_off;   // X / ls, row-major [n_pad x D_pad]
  const int nt = n_pad / 32;
  const int64_t g_M = 0, g_Kinv = g_M + (int64_t)n_pad * n_pad, g_T = g_Kinv + (int64_t)n_pad * n_pad,
                g_contrib = g_T + (int64_t)n_pad * n_pad, g_part = g_contrib + (int64_t)n_pad * D_pad,
                g_out = g_part + (int64_t)nt * nt, g_end = g_out + D_pad + 8;
  int rc = ob_reserve(ctx, &ctx->grad_dev, &ctx->grad_cap, g_end);
  if (rc) return rc;
  double* G = ctx->grad_dev;
  transpose_kernel<<<dim3(nt, nt), dim3(32, 8), 0, st>>>(Xinv, n_pad, G + g_M);
  OB_LAUNCH_CHECK(ctx);
  gemm_args k{};  // K^-1 = inv(L)^T inv(L) = M M^T, lower tiles
  k.A = G + g_M; k.lda = n_pad; k.B = G + g_M; k.ldb = n_pad; k.C = G + g_Kinv; k.ldc = n_pad;
  k.M = n_pad; k.N = n_pad; k.K = n_pad; k.alpha = 1.0; k.beta = 0.0; k.lower_only = 1;
  k.tri_k = 4;   // M is upper triangular: row i is zero left of column i, and the lower tiles have i >= j
  if ((rc = launch_gemm(ctx, true, k, 1))) return rc;
  const size_t smem = (size_t)2 * 32 * (D_pad + 1) * sizeof(double);
  OB_CUDA(ctx, cudaFuncSetAttribute(lml_t_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem > 49152 ? smem : 49152)));
  lml_t_kernel<<<dim3(nt, nt), 256, smem, st>>>(Xs, alpha, G + g_Kinv, n, n_pad, D_pad, ctx->hdr.kernel_kind, ctx->hdr.amplitude,
                                                 G + g_T, G + g_part);
  OB_LAUNCH_CHECK(ctx);
  lml_row_kernel<<<(n * 32 + 255) / 256, 256, 0, st>>>(G + g_T, Xs, n, n_pad, D_pad, G + g_contrib);
  OB_LAUNCH_CHECK(ctx);
  lml_final_kernel<<<D_pad + 2, 256, 0, st>>>(G + g_contrib, G + g_part, nt * nt, alpha, G + g_Kinv, n, n_pad, D_pad, G + g_out);
  OB_LAUNCH_CHECK(ctx);
  std::vector<double> h((size_t)D_pad + 2);
  OB_CUDA(ctx, cudaMemcpyAsync(h.data(), G + g_out, (size_t)(D_pad + 2) * 8, cudaMemcpyDeviceToHost, st));
  OB_CUDA(ctx, cudaStreamSynchronize(st));
  grad_host[0] = h[0];
  for (int d = 0; d < D; ++d) grad_host[1 + d] = h[1 + d];
  grad_host[1 + D] = h[1 + D_pad];
  if (lml_host) *lml_host = ctx->hdr.lml;
  return OB200_OK;
}

int ob200_gp_predict_grad_host(ob200_ctx* ctx, const double* x_host, int64_t ldx, int m, double* out_host) {
  if (!ctx) return OB200_EINVAL;
  if (!x_host || !out_host) OB_FAIL(ctx, OB200_EINVAL, "null argument");
  if (!ctx->fitted || !ctx->have_local_fit) OB_FAIL(ctx, OB200_ESTATE, "the posterior gradient needs a fit made on this context");
  const int n = ctx->hdr.n, n_pad = ctx->hdr.n_pad, D = ctx->hdr.D, D_pad = ctx->hdr.D_pad;
  if (m < 1 || m > 64) OB_FAIL(ctx, OB200_EINVAL, "m=%d outside [1, 64]", m);
  if (ldx < D) OB_FAIL(ctx, OB200_EINVAL, "ldx=%lld < D=%d", (long long)ldx, D);
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const ob200_blob_layout lay = ob_blob_layout(n_pad, D_pad);
  const double* alpha = reinterpret_cast<const double*>(ctx->blob + lay.off_alpha);
  const double* ls = reinterpret_cast<const double*>(ctx->blob + lay.off_ls);
  const double* Xinv = ctx->work_dev;                 // dense inv(L), row-major (fit workspace, offset 0)
  const double* Xs = ctx->work_dev + ctx->w_Xs_off;   // X / ls, row-major [n_pad x D_pad]
  const int W = 2 + 2 * D;
  const int64_t p_x = 0, p_xs = p_x + (int64_t)m * D, p_K = p_xs + (int64_t)m * D_pad, p_G = p_K + (int64_t)m * n_pad,
                p_V = p_G + (int64_t)m * n_pad, p_W = p_V + (int64_t)m * n_pad, p_out = p_W + (int64_t)m * n_pad,
                p_end = p_out + (int64_t)m * W;
  int rc = ob_reserve(ctx, &ctx->pg_dev, &ctx->pg_cap, p_end);
  if (rc) return rc;
  double* P = ctx->pg_dev;
  if (ldx == D) {
    OB_CUDA(ctx, cudaMemcpyAsync(P + p_x, x_host, (size_t)m * D * 8, cudaMemcpyHostToDevice, st));
  } else {
    for (int p = 0; p < m; ++p)
      OB_CUDA(ctx, cudaMemcpyAsync(P + p_x + (size_t)p * D, x_host + (size_t)p * ldx, (size_t)D * 8, cudaMemcpyHostToDevice, st));
  }
  pg_k_kernel<<<dim3((n_pad + 255) / 256, m), 256, (size_t)D_pad * 8, st>>>(P + p_x, D, ls, Xs, n, n_pad, D, D_pad, ctx->hdr.kernel_kind,
                                                                          ctx->hdr.amplitude, P + p_xs, P + p_K, P + p_G);
  OB_LAUNCH_CHECK(ctx);
  pg_v_kernel<<<dim3((n_pad * 32 + 255) / 256, m), 256, 0, st>>>(Xinv, P + p_K, n_pad, P + p_V);
  OB_LAUNCH_CHECK(ctx);
  pg_w_kernel<<<dim3((n_pad + PG_W_COLS - 1) / PG_W_COLS, m), 256, 0, st>>>(Xinv, P + p_V, n_pad, P + p_W);
  OB_LAUNCH_CHECK(ctx);
  pg_final_kernel<<<m, 256, 0, st>>>(P + p_xs, ls, Xs, alpha, P + p_K, P + p_G, P + p_V, P + p_W, n, n_pad, D, D_pad,
                                     ctx->hdr.amplitude, ctx->hdr.y_mean, ctx->hdr.y_std, P + p_out);
  OB_LAUNCH_CHECK(ctx);
  OB_CUDA(ctx, cudaMemcpyAsync(out_host, P + p_out, (size_t)m * W * 8, cudaMemcpyDeviceToHost, st));
  OB_CUDA(ctx, cudaStreamSynchronize(st));
  return OB200_OK;
}

int ob200_gp_append(ob200_ctx* ctx, const double* x_new_dev, const double* y_dev, int n_total, int normalize_y) {
  if (!ctx) return OB200_EINVAL;
  if (!x_new_dev || !y_dev) OB_FAIL(ctx, OB200_EINVAL, "null argument");
  if (!ctx->fitted || !ctx->have_local_fit) OB_FAIL(ctx, OB200_ESTATE, "append needs a fit made on this context");
  const int n = ctx->hdr.n, n_pad = ctx->hdr.n_pad, D = ctx->hdr.D, D_pad = ctx->hdr.D_pad;
  if (n_total != n + 1) OB_FAIL(ctx, OB200_EINVAL, "n_total=%d, the current fit has %d observations", n_total, n);
  if (n + 1 > n_pad || n + 1 > OB200_MAX_N) OB_FAIL(ctx, OB200_ESTATE, "no room in the padded factor (n=%d, n_pad=%d): refit", n, n_pad);
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const ob200_blob_layout lay = ob_blob_layout(n_pad, D_pad);
  char* blob = ctx->blob;
  double* b_ls = reinterpret_cast<double*>(blob + lay.off_ls);
  double* b_xs = reinterpret_cast<double*>(blob + lay.off_xs);
  double* b_xx = reinterpret_cast<double*>(blob + lay.off_xx);
  double* b_alpha = reinterpret_cast<double*>(blob + lay.off_alpha);
  double* b_linv = reinterpret_cast<double*>(blob + lay.off_linv);
  double* W = ctx->work_dev;
  double *Xinv = W, *Xs = W + ctx->w_Xs_off, *ynorm = W + ctx->w_y_off, *stats = W + ctx->w_stats_off;
  int* info_dev = reinterpret_cast<int*>(stats + 6);
  double* lml_dev = stats + 4;
  // scratch: k | l | w | (d, 1/d)
  int rc = ob_reserve(ctx, &ctx->pg_dev, &ctx->pg_cap, (int64_t)6 * n_pad + 8);
  if (rc) return rc;
  double *kvec = ctx->pg_dev, *lvec = kvec + n_pad, *wvec = lvec + n_pad, *sol = wvec + n_pad, *scal = sol + 3 * (size_t)n_pad;
  ctx->fitted = false;  // a failure below leaves no usable fit
  ctx->have_local_fit = false;
  OB_CUDA(ctx, cudaMemsetAsync(info_dev, 0, sizeof(int), st));
  append_k_kernel<<<(n_pad + 255) / 256, 256, (size_t)D_pad * 8, st>>>(x_new_dev, b_ls, D, D_pad, n, n_pad, ctx->hdr.kernel_kind,
                                                                     ctx->hdr.amplitude, Xs, b_xx, kvec);
  OB_LAUNCH_CHECK(ctx);
  pg_v_kernel<<<dim3((n_pad * 32 + 255) / 256, 1), 256, 0, st>>>(Xinv, kvec, n_pad, lvec);  // l = inv(L) k
  OB_LAUNCH_CHECK(ctx);
  append_diag_kernel<<<1, 1024, 0, st>>>(lvec, n, n_pad, ctx->hdr.amplitude + ctx->fit_alpha, ctx->L_dev, scal, info_dev);
  OB_LAUNCH_CHECK(ctx);
  pg_w_kernel<<<dim3((n_pad + PG_W_COLS - 1) / PG_W_COLS, 1), 256, 0, st>>>(Xinv, lvec, n_pad, wvec);         // w = inv(L)^T l
  OB_LAUNCH_CHECK(ctx);
  append_row_kernel<<<(n + 256) / 256, 256, 0, st>>>(wvec, scal, n, n_pad, Xinv);
  OB_LAUNCH_CHECK(ctx);
  // y statistics change with every observation when normalize_y is on: alpha = inv(L)^T inv(L) y_norm in full
  ystats_kernel<<<1, 1024, 0, st>>>(y_dev, n + 1, n_pad, normalize_y, ynorm, stats);
  OB_LAUNCH_CHECK(ctx);
  if ((rc = solve_alpha(ctx, Xinv, ctx->L_dev, ynorm, n_pad, b_alpha, sol, 1))) return rc;
  append_lml_kernel<<<1, 1024, 0, st>>>(ctx->L_dev, n_pad, n + 1, ynorm, b_alpha, lml_dev);
  OB_LAUNCH_CHECK(ctx);
  pack_linv_kernel<<<ctx->num_sms * 8, 256, 0, st>>>(Xinv, n + 1, n_pad, b_linv);
  OB_LAUNCH_CHECK(ctx);
  pack_xs_kernel<<<ctx->num_sms * 2, 256, 0, st>>>(Xs, n_pad, D_pad, b_xs);
  OB_LAUNCH_CHECK(ctx);
  ob200_fit_header h = ctx->hdr;
  h.n = n + 1;
  write_header_kernel<<<1, 128, 0, st>>>(reinterpret_cast<ob200_fit_header*>(blob), h, stats, lml_dev, b_ls, b_ls, D_pad, D_pad, D_pad);
  OB_LAUNCH_CHECK(ctx);
  int info = 0;
  OB_CUDA(ctx, cudaMemcpyAsync(&info, info_dev, sizeof(int), cudaMemcpyDeviceToHost, st));
  OB_CUDA(ctx, cudaMemcpyAsync(&ctx->hdr, blob, sizeof(ob200_fit_header), cudaMemcpyDeviceToHost, st));
  OB_CUDA(ctx, cudaStreamSynchronize(st));
  if (info != 0)
    OB_FAIL(ctx, OB200_ENOTSPD,
            "The kernel is not returning a positive definite matrix (pivot %d of %d). Try gradually increasing the 'alpha' "
            "parameter of your GaussianProcessRegressor estimator.", info, n + 1);
  ctx->fitted = true;
  ctx->have_local_fit = true;
  ctx->fit_serial++;
  return OB200_OK;
}

int ob200_gp_get_ystats(ob200_ctx* ctx, double* y_mean, double* y_std, double* y_min) {
  if (!ctx) return OB200_EINVAL;
  if (!ctx->fitted) OB_FAIL(ctx, OB200_ESTATE, "no fit");
  if (y_mean) *y_mean = ctx->hdr.y_mean;
  if (y_std) *y_std = ctx->hdr.y_std;
  if (y_min) *y_min = ctx->hdr.y_min;
  return OB200_OK;
}

int ob200_gp_get_factor(ob200_ctx* ctx, double* L_host, double* alpha_host) {
  if (!ctx) return OB200_EINVAL;
  if (!ctx->fitted) OB_FAIL(ctx, OB200_ESTATE, "no fit");
  if (L_host && !ctx->have_local_fit) OB_FAIL(ctx, OB200_ESTATE, "L is not kept on contexts that imported their fit");
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  const int n = ctx->hdr.n, n_pad = ctx->hdr.n_pad;
  if (L_host)
    OB_CUDA(ctx, cudaMemcpy2DAsync(L_host, (size_t)n * 8, ctx->L_dev, (size_t)n_pad * 8, (size_t)n * 8, (size_t)n,
                                   cudaMemcpyDeviceToHost, ctx->stream));
  if (alpha_host) {
    const ob200_blob_layout lay = ob_blob_layout(n_pad, ctx->hdr.D_pad);
    OB_CUDA(ctx, cudaMemcpyAsync(alpha_host, ctx->blob + lay.off_alpha, (size_t)n * 8, cudaMemcpyDeviceToHost, ctx->stream));
  }
  OB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return OB200_OK;
}

int ob200_gp_export(ob200_ctx* ctx, void* blob_dev, int64_t bytes) {
  if (!ctx || !blob_dev) return OB200_EINVAL;
  if (!ctx->fitted) OB_FAIL(ctx, OB200_ESTATE, "no fit to export");
  if (bytes < ctx->blob_bytes) OB_FAIL(ctx, OB200_EINVAL, "blob too small: %lld < %lld", (long long)bytes, (long long)ctx->blob_bytes);
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  OB_CUDA(ctx, cudaMemcpyAsync(blob_dev, ctx->blob, (size_t)ctx->blob_bytes, cudaMemcpyDeviceToDevice, ctx->stream));
  return OB200_OK;
}

int ob200_gp_import(ob200_ctx* ctx, const void* blob_dev, int64_t bytes) {
  if (!ctx || !blob_dev) return OB200_EINVAL;
  if (bytes < 256) OB_FAIL(ctx, OB200_EINVAL, "blob too small");
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  ob200_fit_header h{};
  OB_CUDA(ctx, cudaMemcpyAsync(&h, blob_dev, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
  OB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (h.magic != OB200_FIT_MAGIC || h.n < 1 || h.n > OB200_MAX_N || h.n_pad % SC_BM || h.D_pad % 4)
    OB_FAIL(ctx, OB200_EINVAL, "not a fit blob");
  const ob200_blob_layout lay = ob_blob_layout(h.n_pad, h.D_pad);
  if (bytes < lay.total) OB_FAIL(ctx, OB200_EINVAL, "blob truncated");
  ctx->fitted = false;
  ctx->have_local_fit = false;
  int rc = ob_reserve(ctx, &ctx->blob, &ctx->blob_cap, lay.total);
  if (rc) return rc;
  OB_CUDA(ctx, cudaMemcpyAsync(ctx->blob, blob_dev, (size_t)lay.total, cudaMemcpyDeviceToDevice, ctx->stream));
  ctx->hdr = h;
  ctx->blob_bytes = lay.total;
  ctx->fitted = true;
  ctx->fit_serial++;
  return OB200_OK;
}

}  // extern "C"
