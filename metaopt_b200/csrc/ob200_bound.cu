// Bound pass of the exact pruned scoring: for every candidate, a LOWER bound of the posterior mean and an UPPER
// bound of the posterior standard deviation, from the cross-kernel alone, in FP32 on the TF32 tensor cores.
//
// Why it may be inexact: the pruned path (ob200_score.cu) only needs U(mu_lb, sigma_ub) >= U(mu, sigma); every
// candidate whose bound reaches the incumbent is re-scored by the FP64 score kernel, so the argmax and the
// winner's values stay bit-identical to the dense call however loose the bound is.  What the bound must be is
// RIGOROUS, so this kernel carries its own error budget:
//     mu~   = sum_j k~_j alpha_j              (k~ = FP32 kernel value, alpha rounded to FP32)
//     S~    = sum_j k~_j |alpha_j|
//     |mu - mu~| <= y_std * S~ * rel(c)       rel(c) = relative error bound of one kernel value for candidate c
//     k_max >= k~_max * (1 - rel(c))          (nearest observation -> sigma bound, see score_kernel<MEAN_ONLY>)
// rel(c) is assembled below (`bound_rel`) from: the 3xTF32 split of the distance contraction (dropped lo*lo term
// and split residuals), a deliberately generous model of the tensor core's FP32 accumulation, the FP32 roundings of
// ||c||^2, ||x||^2, the MUFU.SQRT / MUFU.EX2 approximations and the FP32 partial sums (flushed into FP64 every
// observation chunk).  tests/test_gpu_bound.py checks mu_lb <= mu and sigma_ub >= sigma against the FP64 kernel on
// every candidate of the fuzz cases and reports how much of the budget is actually used.
//
// Replaces (for the bound only) sklearn kernels.py:1720-1729 / :1569-1570 and _gpr.py:446-447.
//
// Work decomposition: CTA tile = CT candidates (128 / 64 / 32, the largest whose A fragments fit) x all observed
// points; WC x WO consumer warps + one producer warp.  D <= 32 (1..4 k-steps): 4 x 2 warps, each with a 32-candidate
// x 64-observation share of a chunk, its A fragments in registers for the whole tile, the chunk walked two n8 blocks
// at a time (MMAs of the next blocks beside the kernel evaluations of the previous ones in one unrolled stream).  Wider
// spaces: WC x 4 warps with 32 x 32 register tiles, A fragments re-read from shared memory per k-step.  Observed points
// arrive in chunks of 128, pre-split into TF32 hi/lo B fragments once per fit, staged by bulk async copies (TMA
// engine) from the producer warp through an mbarrier ring.
// The distance is in "T units": coordinates are pre-scaled so that Matern-5/2 is (1 + t + t^2/3) e^-t with
// t = sqrt(T) and RBF is e^-T.
#include <math.h>
#include <stdlib.h>

#include "ob200_internal.cuh"

namespace {

#define BD_OBC 128            // observed points per chunk
#define BD_STAGES_MAX 4
#define BD_PROBES 64          // observations at which the size of sum k |alpha| is sampled for the per-fit verdict
#ifdef BD_RSQ
#define BD_T_FLOOR 1e-30f     // t = T * rsqrt(T) needs T > 0; the shift is far below the error budget
#else
#define BD_T_FLOOR 0.f
#endif

struct bound_args {
  const double* cand; long long ldc; long long M;
  const double* ls;
  const float4* bp;     // [chunk][kslice][obs8 (16)][kstep in slice (ksl)][lane] {hi b0, hi b1, lo b0, lo b1}
  const float4* aux;    // [obs8][q] {T-space ||x||^2 of obs 2q, 2q+1, amplitude-free alpha of obs 2q, 2q+1}
  int n_chunks, D, nk, ksl, n_slices, kind, stages;
  double scale;         // coordinate scale: T = ||scale (c - x) / ls||^2
  double amplitude, y_mean, y_std, diag;
  const double* xx_max; // device scalar: the largest T-space ||x||^2 over the real observations
  double g_T;           // absolute error of T~ per unit of (sqrt(cc) + sqrt(xx_max))^2
  double c0, c1;        // function-evaluation error: c0 + c1 * t_max
  double tiny;          // absolute slack (flushed denormals)
  double* mu_lb; double* sigma_ub;
  long long n_tiles;
};

__device__ __forceinline__ uint32_t to_tf32(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ float sqrt_approx(float v) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ float ex2_approx(float v) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  return r;
}
// D(16x8, f32) += A(16x8, tf32, row) * B(8x8, tf32, col)
// lane (g = lane/4, q = lane%4): a0 = A[g][q], a1 = A[g+8][q], a2 = A[g][q+4], a3 = A[g+8][q+4];
// b0 = B[q][g], b1 = B[q+4][g]; c0 = C[g][2q], c1 = C[g][2q+1], c2 = C[g+8][2q], c3 = C[g+8][2q+1]
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t a0, const uint32_t a1, const uint32_t a2,
                                         const uint32_t a3, const uint32_t b0, const uint32_t b1) {
#ifdef BD_DBG_NOMMA
  c[0] += __uint_as_float(a0 ^ b0); return;    // timing experiment only
#endif
  asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__host__ __device__ inline size_t bound_smem_bytes(int ct, int nk, int ksl, int stages) {
  return (size_t)(ct / 16) * nk * 1024            // A fragments (hi plane + lo plane)
         + (size_t)stages * ksl * 8192            // B stages
         + (size_t)ct * 8                         // cc
         + (size_t)3 * 4 * ct * 8                 // cross-warp reduction (mu, S, kmax): WO (<= 4) x CT entries each
         + 2 * BD_STAGES_MAX * 8 + (size_t)nk * 64 + 128;
}

template <int KIND>
__device__ __forceinline__ float kernel_T(float T) {
  if (KIND == OB200_KERNEL_MATERN52) {
#if defined(BD_DBG_NOSQRT)
    const float t = T * 0.25f;                 // timing experiment only
#elif defined(BD_RSQ)
    const float t = T * rsqrt_approx(T);       // (T >= 1e-30 here)
#else
    const float t = sqrt_approx(T);
#endif
    const float p = fmaf(t, fmaf(t, 0.333333343f, 1.0f), 1.0f);
#if defined(BD_DBG_NOEX2)
    return p * fmaf(t, -0.01f, 1.0f);          // timing experiment only
#else
    return p * ex2_approx(t * -1.44269504f);
#endif
  }
  return ex2_approx(T * -1.44269504f);
}

// D = A * B (no accumulator input: the first product of a tile needs no zeroed registers)
__device__ __forceinline__ void mma_tf32_first(float (&c)[4], const uint32_t a0, const uint32_t a1, const uint32_t a2,
                                               const uint32_t a3, const uint32_t b0, const uint32_t b1) {
  asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
      : "=f"(c[0]), "=f"(c[1]), "=f"(c[2]), "=f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "f"(0.f));
}

// the producer's wait: back off between polls so that the spinning lane does not eat its sub-partition's issue slots
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0;
  for (;;) {
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    if (done) break;
    __nanosleep(200);
  }
}

// kernel values of one n8 block (32 candidates x 8 observations per warp): mean / |mean| partial sums, nearest observation
template <int KIND>
__device__ __forceinline__ void epilogue_nb(const float (&acc)[2][4], const float4 ax, const float (&ccr)[2][2], float (&mu32)[2][2],
                                            float (&s32)[2][2], float (&kmx)[2][2]) {
  const float xx0 = ax.x, xx1 = ax.y, al0 = ax.z, al1 = ax.w;
  const float ab0 = fabsf(al0), ab1 = fabsf(al1);
#pragma unroll
  for (int mb = 0; mb < 2; ++mb)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float T0 = fmaxf(fmaf(-2.f, acc[mb][2 * h], ccr[mb][h] + xx0), BD_T_FLOOR);
      const float T1 = fmaxf(fmaf(-2.f, acc[mb][2 * h + 1], ccr[mb][h] + xx1), BD_T_FLOOR);
      const float k0 = kernel_T<KIND>(T0), k1 = kernel_T<KIND>(T1);
      mu32[mb][h] = fmaf(k0, al0, mu32[mb][h]);
      mu32[mb][h] = fmaf(k1, al1, mu32[mb][h]);
      s32[mb][h] = fmaf(k0, ab0, s32[mb][h]);
      s32[mb][h] = fmaf(k1, ab1, s32[mb][h]);
      kmx[mb][h] = fmaxf(kmx[mb][h], fmaxf(k0, k1));
    }
}

// WC x WO consumer warps: WC along the candidates (CT = 32 WC), WO along the 128 observations of a chunk (NB = 16 / WO n8
// blocks per warp).  NKR > 0: the space has exactly NKR k-steps (D <= 8 NKR <= 32) -- the warp keeps its A fragments in
// registers for the whole tile and walks the chunk n8 block by n8 block, MMAs of block nb + 1 next to the kernel
// evaluations of block nb in one unrolled instruction stream, so the tensor, MUFU and FFMA pipes overlap inside a warp.
// NKR == 0: any width; A fragments re-read from shared memory per k-step, all MMAs of a chunk before its evaluations.
template <int WC, int WO, int KIND, int NKR>
__global__ void __launch_bounds__((WC * WO + 1) * 32, 1) bound_kernel(bound_args a) {
  constexpr int BD_WARPS = WC * WO, CT = 32 * WC, NB = (BD_OBC / 8) / WO;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int nk = NKR > 0 ? NKR : a.nk, ksl = NKR > 0 ? NKR : a.ksl;
  float4* As = reinterpret_cast<float4*>(smem_raw);                                   // [CT/16][nk][hi | lo][32]
  float4* Bs = As + (size_t)(CT / 16) * nk * 64;                                      // stages x [16][ksl][32]
  double* cc = reinterpret_cast<double*>(Bs + (size_t)a.stages * ksl * 512);          // [CT]
  double* red = cc + CT;                                                              // [3][WO][CT]
  uint64_t* full = reinterpret_cast<uint64_t*>(red + 3 * WO * CT);
  uint64_t* empty = full + BD_STAGES_MAX;
  double* sl = reinterpret_cast<double*>(empty + BD_STAGES_MAX);                      // [8 nk] scale / ls (0 beyond D)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  for (int d = tid; d < nk * 8; d += blockDim.x) sl[d] = d < a.D ? a.scale / a.ls[d] : 0.0;
  if (tid == 0) {
    for (int s = 0; s < BD_STAGES_MAX; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], BD_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  uint32_t it = 0;
  const uint32_t stage_bytes = (uint32_t)ksl * 8192u;
  if (warp == BD_WARPS) {   // ---- producer warp: B stages for every tile this CTA owns
    if (lane == 0) {
      for (long long tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x)
        for (int c = 0; c < a.n_chunks; ++c)
          for (int s = 0; s < a.n_slices; ++s, ++it) {
            const int st = it % a.stages;
            const uint32_t ph = (it / a.stages) & 1;
            mbar_wait_backoff(&empty[st], ph ^ 1);
            mbar_expect_tx(&full[st], stage_bytes);
            bulk_g2s(Bs + (size_t)st * ksl * 512, a.bp + ((size_t)c * a.n_slices + s) * ksl * 512, stage_bytes, &full[st]);
          }
    }
    return;
  }

  const int wc = warp / WO, wo = warp % WO;
  for (long long tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
    const long long row0 = tile * CT;
    // ---- stage the candidate tile: u = scale * c / ls -> TF32 hi/lo A fragments; cc = ||u||^2 in FP64.
    // (32 WO / CT = WO / WC ... ) TPR threads per row (columns d = part, part + TPR, ...).
    {
      constexpr int TPR = (BD_WARPS * 32) / CT;   // = WO: 2 or 4 threads per row
      const int r = tid / TPR, part = tid % TPR;
      const bool live = row0 + r < a.M;
      double acc = 0.0;
      const int mb = r >> 4, rr = r & 15, gg = rr & 7, hi8 = rr >> 3;
      const double* src = a.cand + (size_t)(row0 + r) * a.ldc;
      {  // the next tile of this CTA: pull its rows towards L2 / L1 while this one is processed
        const long long nrow = row0 + (long long)gridDim.x * CT + r;
        if (nrow < a.M && part == 0)
          for (int b = 0; b < a.D * 8; b += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(a.cand + (size_t)nrow * a.ldc) + b));
      }
#pragma unroll 4
      for (int d = part; d < nk * 8; d += TPR) {
        double u = 0.0;
        if (live && d < a.D) u = src[d] * sl[d];   // (one rounding away from scale * c / ls: far inside the budget)
        acc = fma(u, u, acc);
        const float uf = (float)u;
        const uint32_t hi = to_tf32(uf);
        const uint32_t lo = to_tf32(uf - __uint_as_float(hi));
        const int ks = d >> 3, kk = d & 7, qq = kk & 3, hi4 = kk >> 2;
        uint32_t* dst = reinterpret_cast<uint32_t*>(As + (size_t)(mb * nk + ks) * 64 + (gg * 4 + qq));
        dst[hi8 + 2 * hi4] = hi;
        dst[32 * 4 + hi8 + 2 * hi4] = lo;
      }
#pragma unroll
      for (int o = 1; o < TPR; o <<= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (part == 0) cc[r] = acc;
    }
    asm volatile("bar.sync 1, %0;" ::"n"(BD_WARPS * 32) : "memory");

    float ccr[2][2];
#pragma unroll
    for (int mb = 0; mb < 2; ++mb) { ccr[mb][0] = (float)cc[wc * 32 + mb * 16 + g]; ccr[mb][1] = (float)cc[wc * 32 + mb * 16 + g + 8]; }
    double mu64[2][2] = {{0.0, 0.0}, {0.0, 0.0}}, s64[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
    float kmx[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
    const float4* Aw = As + (size_t)(wc * 2) * nk * 64 + lane;

    if constexpr (NKR > 0) {
      uint32_t Ah[2][NKR][4], Al[2][NKR][4];
#pragma unroll
      for (int mb = 0; mb < 2; ++mb)
#pragma unroll
        for (int ks = 0; ks < NKR; ++ks) {
          const float4 h = Aw[(size_t)(mb * NKR + ks) * 64], l = Aw[(size_t)(mb * NKR + ks) * 64 + 32];
          Ah[mb][ks][0] = __float_as_uint(h.x); Ah[mb][ks][1] = __float_as_uint(h.y); Ah[mb][ks][2] = __float_as_uint(h.z); Ah[mb][ks][3] = __float_as_uint(h.w);
          Al[mb][ks][0] = __float_as_uint(l.x); Al[mb][ks][1] = __float_as_uint(l.y); Al[mb][ks][2] = __float_as_uint(l.z); Al[mb][ks][3] = __float_as_uint(l.w);
        }
      for (int c = 0; c < a.n_chunks; ++c, ++it) {
        const float4* axp = a.aux + ((size_t)(c * (BD_OBC / 8) + wo * NB)) * 4 + q;   // this lane's observation constants
        const int st = it % a.stages;
        const uint32_t ph = (it / a.stages) & 1;
        mbar_wait(&full[st], ph);
        const float4* Bw = Bs + (size_t)st * NKR * 512 + (size_t)(wo * NB) * NKR * 32 + lane;
        float mu32[2][2] = {{0.f, 0.f}, {0.f, 0.f}}, s32[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
        // Two n8 blocks at a time: four independent accumulators (2 x mb) take turns, so an HMMA never waits for the
        // one issued just before it (the chain into one accumulator is 3 NKR long); small terms first.
#pragma unroll
        for (int nb = 0; nb < NB; nb += 2) {
          float acc[2][2][4];
          uint32_t bh[2][NKR][2], bl[2][NKR][2];
#pragma unroll
          for (int e = 0; e < 2; ++e)
#pragma unroll
            for (int ks = 0; ks < NKR; ++ks) {
              const float4 b = Bw[(size_t)((nb + e) * NKR + ks) * 32];
              bh[e][ks][0] = __float_as_uint(b.x); bh[e][ks][1] = __float_as_uint(b.y); bl[e][ks][0] = __float_as_uint(b.z); bl[e][ks][1] = __float_as_uint(b.w);
            }
#pragma unroll
          for (int ks = 0; ks < NKR; ++ks)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
              for (int mb = 0; mb < 2; ++mb) {
                if (ks == 0) mma_tf32_first(acc[e][mb], Al[mb][ks][0], Al[mb][ks][1], Al[mb][ks][2], Al[mb][ks][3], bh[e][ks][0], bh[e][ks][1]);
                else mma_tf32(acc[e][mb], Al[mb][ks][0], Al[mb][ks][1], Al[mb][ks][2], Al[mb][ks][3], bh[e][ks][0], bh[e][ks][1]);
              }
#pragma unroll
          for (int ks = 0; ks < NKR; ++ks)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
              for (int mb = 0; mb < 2; ++mb) mma_tf32(acc[e][mb], Ah[mb][ks][0], Ah[mb][ks][1], Ah[mb][ks][2], Ah[mb][ks][3], bl[e][ks][0], bl[e][ks][1]);
#pragma unroll
          for (int ks = 0; ks < NKR; ++ks)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
              for (int mb = 0; mb < 2; ++mb) mma_tf32(acc[e][mb], Ah[mb][ks][0], Ah[mb][ks][1], Ah[mb][ks][2], Ah[mb][ks][3], bh[e][ks][0], bh[e][ks][1]);
#pragma unroll
          for (int e = 0; e < 2; ++e) epilogue_nb<KIND>(acc[e], __ldg(axp + (nb + e) * 4), ccr, mu32, s32, kmx);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[st]);
#pragma unroll
        for (int mb = 0; mb < 2; ++mb)
#pragma unroll
          for (int h = 0; h < 2; ++h) { mu64[mb][h] += (double)mu32[mb][h]; s64[mb][h] += (double)s32[mb][h]; }
      }
    } else {
      for (int c = 0; c < a.n_chunks; ++c) {
        float acc[NB][2][4];
        const float4* axp = a.aux + ((size_t)(c * (BD_OBC / 8) + wo * NB)) * 4 + q;
        for (int s = 0; s < a.n_slices; ++s, ++it) {
          const int st = it % a.stages;
          const uint32_t ph = (it / a.stages) & 1;
          mbar_wait(&full[st], ph);
          const float4* Bw = Bs + (size_t)st * ksl * 512 + (size_t)(wo * NB) * ksl * 32 + lane;
          const int jn = min(ksl, nk - s * ksl);
          for (int j = 0; j < jn; ++j) {
            const int ks = s * ksl + j;
            float4 ah[2], al[2];
#pragma unroll
            for (int mb = 0; mb < 2; ++mb) { ah[mb] = Aw[(size_t)(mb * nk + ks) * 64]; al[mb] = Aw[(size_t)(mb * nk + ks) * 64 + 32]; }
            // split products in the OUTER position: the three MMAs into one accumulator are 2 NB instructions apart
#pragma unroll
            for (int term = 0; term < 3; ++term)
#pragma unroll
              for (int nb = 0; nb < NB; ++nb) {
                const float4 b = Bw[(size_t)(nb * ksl + j) * 32];
                const uint32_t b0 = __float_as_uint(term == 1 ? b.z : b.x), b1 = __float_as_uint(term == 1 ? b.w : b.y);
#pragma unroll
                for (int mb = 0; mb < 2; ++mb) {
                  const float4 av = term == 0 ? al[mb] : ah[mb];
                  if (term == 0 && ks == 0)
                    mma_tf32_first(acc[nb][mb], __float_as_uint(av.x), __float_as_uint(av.y), __float_as_uint(av.z), __float_as_uint(av.w), b0, b1);
                  else
                    mma_tf32(acc[nb][mb], __float_as_uint(av.x), __float_as_uint(av.y), __float_as_uint(av.z), __float_as_uint(av.w), b0, b1);
                }
              }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&empty[st]);
        }
        float mu32[2][2] = {{0.f, 0.f}, {0.f, 0.f}}, s32[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) epilogue_nb<KIND>(acc[nb], __ldg(axp + nb * 4), ccr, mu32, s32, kmx);
#pragma unroll
        for (int mb = 0; mb < 2; ++mb)
#pragma unroll
          for (int h = 0; h < 2; ++h) { mu64[mb][h] += (double)mu32[mb][h]; s64[mb][h] += (double)s32[mb][h]; }
      }
    }

    // ---- reduce over the 4 column lanes, then over the WO observation warps through shared memory
#pragma unroll
    for (int mb = 0; mb < 2; ++mb)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        double m = mu64[mb][h], s = s64[mb][h];
        float k = kmx[mb][h];
        m += __shfl_xor_sync(0xffffffffu, m, 1); m += __shfl_xor_sync(0xffffffffu, m, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 1); s += __shfl_xor_sync(0xffffffffu, s, 2);
        k = fmaxf(k, __shfl_xor_sync(0xffffffffu, k, 1)); k = fmaxf(k, __shfl_xor_sync(0xffffffffu, k, 2));
        if (q == 0) {
          const int r = wc * 32 + mb * 16 + h * 8 + g;
          red[(0 * WO + wo) * CT + r] = m;
          red[(1 * WO + wo) * CT + r] = s;
          red[(2 * WO + wo) * CT + r] = (double)k;
        }
      }
    asm volatile("bar.sync 1, %0;" ::"n"(BD_WARPS * 32) : "memory");
    if (tid < CT && row0 + tid < a.M) {
      double m = 0.0, s = 0.0, k = 0.0;
#pragma unroll
      for (int w = 0; w < WO; ++w) { m += red[(0 * WO + w) * CT + tid]; s += red[(1 * WO + w) * CT + tid]; k = fmax(k, red[(2 * WO + w) * CT + tid]); }
      const double xm = *a.xx_max;
      const double rc = sqrt(cc[tid]) + sqrt(xm);   // >= sqrt(T) for every observation; rc^2 >= cc + xx + 2|G|
      const double dT = a.g_T * rc * rc;
      double rel = (KIND == OB200_KERNEL_MATERN52 ? dT * (1.0 / 6.0) : dT) * 1.02 + a.c0 + a.c1 * (KIND == OB200_KERNEL_MATERN52 ? rc : rc * rc);
      if (!(xm < 1e29)) rel = 1.0;   // coordinates beyond the FP32 range of this pass: vacuous (but valid) bounds
      // S~ itself carries the relative error rel: sum k|alpha| <= S~ / (1 - rel)
      const double eps = rel < 0.5 ? a.y_std * (a.amplitude * s * rel / (1.0 - rel) * 1.001 + a.tiny) : INFINITY;
      a.mu_lb[row0 + tid] = a.y_std * (a.amplitude * m) + a.y_mean - eps;
      const double klo = rel < 0.5 ? a.amplitude * k * (1.0 - rel) : 0.0;
      const double var_ub = fmin(fmax(a.amplitude - klo * klo / (a.amplitude + a.diag), 0.0) + 1e-6 * a.amplitude, a.amplitude);
      a.sigma_ub[row0 + tid] = a.y_std * sqrt(var_ub);
    }
    asm volatile("bar.sync 1, %0;" ::"n"(BD_WARPS * 32) : "memory");   // As / cc / red are rewritten by the next tile
  }
}

// Observed points of the fit blob (fragment-packed X/ls for the DMMA kernel) -> TF32 hi/lo B fragments in T units,
// plus per-observation {T-space ||x||^2, alpha}; padded observations get ||x||^2 = 1e30 (kernel value 0) and alpha 0.
// Also the largest real ||x||^2 (one block: the table is small).
__global__ void __launch_bounds__(256) bound_pack_kernel(const double* __restrict__ xs, const double* __restrict__ xx,
                                                         const double* __restrict__ alpha, int n, int n_pad, int D_pad, int nk,
                                                         int ksl, int n_slices, int n_chunks, double scale, int kind,
                                                         float4* __restrict__ bp, float4* __restrict__ aux,
                                                         double* __restrict__ xx_max_out) {
  const int nd4 = D_pad >> 2;
  const long long total = (long long)n_chunks * n_slices * 16 * ksl * 32;
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < total; e += (long long)gridDim.x * 256) {
    const int lane = (int)(e & 31);
    long long t = e >> 5;
    const int j = (int)(t % ksl); t /= ksl;
    const int o = (int)(t % 16); t /= 16;
    const int s = (int)(t % n_slices);
    const int c = (int)(t / n_slices);
    const int g = lane >> 2, q = lane & 3;
    const int obs = c * BD_OBC + o * 8 + g, ks = s * ksl + j;
    float v[2] = {0.f, 0.f};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int d = ks * 8 + q + 4 * h;
      if (obs < n && d < D_pad && ks < nk)
        v[h] = (float)(scale * xs[((size_t)(obs >> 3) * nd4 + (d >> 2)) * 32 + (obs & 7) * 4 + (d & 3)]);
    }
    const uint32_t h0 = to_tf32(v[0]), h1 = to_tf32(v[1]);
    float4 out;
    out.x = __uint_as_float(h0); out.y = __uint_as_float(h1);
    out.z = __uint_as_float(to_tf32(v[0] - __uint_as_float(h0)));
    out.w = __uint_as_float(to_tf32(v[1] - __uint_as_float(h1)));
    bp[e] = out;
  }
  const long long n_aux = (long long)n_chunks * (BD_OBC / 8) * 4;
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < n_aux; e += (long long)gridDim.x * 256) {
    const int q = (int)(e & 3);
    const long long ob = e >> 2;
    float4 out;
    const long long j0 = ob * 8 + 2 * q, j1 = j0 + 1;
    // round the squared norm UP-or-nearest is irrelevant: its rounding is part of the error budget (g_T)
    out.x = j0 < n ? (float)(scale * scale * xx[j0]) : 1e30f;
    out.y = j1 < n ? (float)(scale * scale * xx[j1]) : 1e30f;
    out.z = j0 < n ? (float)alpha[j0] : 0.f;
    out.w = j1 < n ? (float)alpha[j1] : 0.f;
    aux[e] = out;
  }
  if (blockIdx.x == 0) {   // xx_max_out[0] = largest T-space ||x||^2, [1] = ||alpha||_1
    __shared__ double sm[256], sa[256];
    double m = 0.0, l1 = 0.0;
    for (int j = threadIdx.x; j < n; j += 256) { m = fmax(m, xx[j]); l1 += fabs(alpha[j]); }
    sm[threadIdx.x] = m; sa[threadIdx.x] = l1;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
      if (threadIdx.x < s) { sm[threadIdx.x] = fmax(sm[threadIdx.x], sm[threadIdx.x + s]); sa[threadIdx.x] += sa[threadIdx.x + s]; }
      __syncthreads();
    }
    if (threadIdx.x == 0) { xx_max_out[0] = sm[0] * scale * scale; xx_max_out[1] = sa[0]; }
  }
  // blocks 1 .. BD_PROBES: S_p = sum_j k(x_p, x_j) |alpha_j| (kernel without the amplitude) at BD_PROBES evenly spaced
  // observations, in FP64 -- the size of the sum whose relative error the FP32 pass has to absorb (verdict per fit)
  if (blockIdx.x >= 1 && blockIdx.x <= BD_PROBES) {
    __shared__ double sp[256];
    const int pr = (int)((long long)(blockIdx.x - 1) * n / BD_PROBES);
    double acc = 0.0;
    for (int j = threadIdx.x; j < n; j += 256) {
      double d2 = 0.0;
      for (int d = 0; d < D_pad; ++d) {
        const double a = xs[((size_t)(pr >> 3) * nd4 + (d >> 2)) * 32 + (pr & 7) * 4 + (d & 3)];
        const double b = xs[((size_t)(j >> 3) * nd4 + (d >> 2)) * 32 + (j & 7) * 4 + (d & 3)];
        d2 = fma(a - b, a - b, d2);
      }
      const double T = scale * scale * d2, t = sqrt(T);
      const double k = kind == OB200_KERNEL_MATERN52 ? (1.0 + t + T * (1.0 / 3.0)) * exp(-t) : exp(-T);
      acc = fma(k, fabs(alpha[j]), acc);
    }
    sp[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) { if (threadIdx.x < s) sp[threadIdx.x] += sp[threadIdx.x + s]; __syncthreads(); }
    if (threadIdx.x == 0) xx_max_out[2 + blockIdx.x - 1] = sp[0];
  }
}

// =====================================================================================================================
// tcgen05 variant of the bound kernel (D <= 32): the 3xTF32 distance contraction runs on the 5th-generation tensor cores
// (tcgen05.mma kind::tf32, issued by one thread, accumulators in TMEM) instead of in the warps' own instruction stream
// (mma.sync = legacy HMMA, 2.25 of the 18.4 instructions per 32 kernel evaluations above, and a pipe that was 41 % busy).
//   CTA = 18 warps: warp 0 streams the observation chunks (bulk async copies), warp 1 allocates TMEM and issues the
//   MMAs, warps 2..17 evaluate: an accumulator block is 128 candidates (TMEM lanes) x 256 observations (TMEM columns),
//   double-buffered over the 512 columns; each evaluating thread owns ONE candidate (its TMEM lane) and 64 of the 256
//   columns, reads them 32 at a time with tcgen05.ld and keeps mu / S / k_max of its candidate in registers.
//   (Tried: staging the next candidate tile into a second A buffer two chunks before the end of the current one, so
//   that the MMA warp never drains -- 1.46 -> 1.68 ms at cfg2: the evaluating warps are the bottleneck and the
//   staging in mid-stream stalls both accumulator blocks; dropped.)
//   Operands are in the UMMA canonical K-major no-swizzle layout: 8-row x 16-byte core matrices, SBO = 128 B between
//   8-row groups, LBO = bytes between the two 16-byte K chunks of a k-step.
#define TC_CT 128             // candidates per tile = TMEM lanes
#define TC_OBC 256            // observations per chunk = accumulator columns
#define TC_STAGES 3
#ifndef TC_EPI_WARPS
#define TC_EPI_WARPS 16          // evaluating warps: 4 TMEM lane quarters x TC_CG column groups (8 warps: 1.58 ms at cfg2, 16: 1.46)
#endif
#define TC_CG (TC_EPI_WARPS / 4)   // column groups of an accumulator block
#define TC_CW (TC_OBC / TC_CG)     // columns per evaluating thread and chunk
#define TC_THREADS ((2 + TC_EPI_WARPS) * 32)

struct bound_tc_args {
  bound_args b;          // cand, ldc, M, ls, D, nk, kind, scale, amplitude, ..., mu_lb, sigma_ub, n_tiles (bp / aux unused)
  const float* pack;     // [chunk][k-slice]: [B hi | B lo] of ksl k-steps in the canonical layout, 2 * 2048 * ksl floats
  const float4* aux2;    // [chunk][obs pair (128)] {xx0, alpha0, xx1, alpha1}
  int n_chunks, stages, ksl, n_slices;   // a B stage = one k-slice (ksl <= 3 k-steps) of one chunk: 16 KB * ksl whatever D is
};

__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);   // version 1 (sm_100), no swizzle, base offset 0
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,"
      "%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n"
      "tcgen05.wait::ld.sync.aligned;\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

__host__ __device__ inline size_t bound_tc_smem_bytes(int nk, int ksl, int stages) {
  return 1024 + (size_t)nk * 8192 /* A hi + lo */ + (size_t)stages * ksl * 16384 /* B stages */ + TC_CT * 8 /* cc */ +
         (size_t)3 * TC_CG * TC_CT * 8 /* mu, S, kmax of the column groups */ + 256;
}

template <int KIND>
__global__ void __launch_bounds__(TC_THREADS, 1) bound_tc_kernel(bound_tc_args ta) {
  const bound_args& a = ta.b;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int nk = a.nk;
  float* Ahi = reinterpret_cast<float*>(smem_raw);          // [2 nk chunks][16 groups][8][4]
  float* Alo = Ahi + (size_t)nk * 1024;
  float* Bst = Alo + (size_t)nk * 1024;                     // stages x ([B hi | B lo] = 2 * 2048 ksl floats)
  const int n_st = ta.stages, ksl = ta.ksl, n_sl = ta.n_slices;
  double* cc = reinterpret_cast<double*>(Bst + (size_t)n_st * ksl * 4096);
  double* red = cc + TC_CT;                                 // [3][TC_CG][TC_CT]
  uint64_t* bars = reinterpret_cast<uint64_t*>(red + 3 * TC_CG * TC_CT);
  uint64_t* full = bars;                  // [TC_STAGES] B stage landed
  uint64_t* empty = full + TC_STAGES;     // [TC_STAGES] B stage consumed by the MMAs
  uint64_t* tfull = empty + TC_STAGES;    // [2] accumulator block complete
  uint64_t* tempty = tfull + 2;           // [2] accumulator block drained
  uint64_t* a_ready = tempty + 2;         // A tile staged
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(a_ready + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < TC_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], TC_EPI_WARPS); }
    mbar_init(a_ready, TC_EPI_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {   // the whole TMEM: two 256-column accumulator blocks (one CTA per SM by shared-memory size)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_base_s)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_s;
  const uint32_t stage_bytes = (uint32_t)ksl * 16384u;

  if (warp == 0) {
    // ---- producer: B chunks of every tile this CTA owns
    if (lane == 0) {
      uint32_t it = 0;
      for (long long tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x)
        for (int c = 0; c < ta.n_chunks * n_sl; ++c, ++it) {   // (chunk, k-slice) pairs are contiguous in the pack
          const int st = it % n_st;
          mbar_wait_backoff(&empty[st], ((it / n_st) & 1) ^ 1);
          mbar_expect_tx(&full[st], stage_bytes);
          bulk_g2s(Bst + (size_t)st * ksl * 4096, ta.pack + (size_t)c * ksl * 4096, stage_bytes, &full[st]);
        }
    }
  } else if (warp == 1) {
    // ---- MMA issuer (one thread): 9 x (M128 N256 K8) per chunk -- lo*hi, hi*lo, hi*hi for each k-step
    if (lane == 0) {
      // instruction descriptor: D = F32, A = B = TF32, both K-major, N = 256, M = 128
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TC_OBC >> 3) << 17) | ((uint32_t)(TC_CT >> 4) << 24);
      const uint32_t a_hi = smem_u32(Ahi), a_lo = smem_u32(Alo);
      uint32_t it = 0, ch = 0, tl = 0;
      for (long long tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, ++tl) {
        mbar_wait_backoff(a_ready, tl & 1);
        for (int c = 0; c < ta.n_chunks; ++c, ++ch) {
          const int as = ch & 1;
          mbar_wait_backoff(&tempty[as], ((ch >> 1) & 1) ^ 1);   // long wait (the evaluating warps are the bottleneck): do not spin
          const uint32_t d = tmem_base + (uint32_t)as * TC_OBC;
          uint32_t acc = 0;
          for (int sl = 0; sl < n_sl; ++sl, ++it) {
            const int st = it % n_st;
            mbar_wait(&full[st], (it / n_st) & 1);
            tc_fence_after();
            const uint32_t b_hi = smem_u32(Bst + (size_t)st * ksl * 4096), b_lo = b_hi + (uint32_t)ksl * 8192u;
            const int jn = min(ksl, nk - sl * ksl);
            for (int j = 0; j < jn; ++j) {
              // k-step = two 16-byte K chunks; A: 16 row groups per chunk (2048 B), B: 32 (4096 B)
              const uint32_t ka = (uint32_t)(sl * ksl + j) * 4096u, kb = (uint32_t)j * 8192u;
              const uint64_t dah = umma_desc(a_hi + ka, 2048u, 128u), dal = umma_desc(a_lo + ka, 2048u, 128u);
              const uint64_t dbh = umma_desc(b_hi + kb, 4096u, 128u), dbl = umma_desc(b_lo + kb, 4096u, 128u);
              umma_tf32(d, dal, dbh, idesc, acc); acc = 1;
              umma_tf32(d, dah, dbl, idesc, 1);
              umma_tf32(d, dah, dbh, idesc, 1);
            }
            umma_commit(&empty[st]);    // the stage is free once these MMAs have read it
          }
          umma_commit(&tfull[as]);      // the accumulator block is complete
        }
      }
    }
  } else {
    // ---- evaluating warps
    const int ew = warp - 2;                    // 0 .. TC_EPI_WARPS - 1
    const int q = warp & 3, half = ew >> 2;     // TMEM lane quarter of this warp (hardware: warp index mod 4), column group
    const int row = q * 32 + lane;              // candidate of this thread inside the tile
    const int et = ew * 32 + lane;
    uint32_t ch = 0;
    for (long long tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
      const long long row0 = tile * TC_CT;
      {  // stage the tile: TPR threads per candidate row (columns part, part + TPR, ...)
        constexpr int TPR = TC_EPI_WARPS * 32 / TC_CT;
        const int r = et / TPR, part = et % TPR;
        const bool live = row0 + r < a.M;
        const double* src = a.cand + (size_t)(row0 + r) * a.ldc;
        double acc = 0.0;
        for (int d = part; d < nk * 8; d += TPR) {
          double u = 0.0;
          if (live && d < a.D) u = a.scale * src[d] / a.ls[d];
          acc = fma(u, u, acc);
          const float uf = (float)u;
          const uint32_t hi = to_tf32(uf);
          const uint32_t lo = to_tf32(uf - __uint_as_float(hi));
          const int idx = (((d >> 2) * 16 + (r >> 3)) * 8 + (r & 7)) * 4 + (d & 3);
          Ahi[idx] = __uint_as_float(hi);
          Alo[idx] = __uint_as_float(lo);
        }
#pragma unroll
        for (int o = 1; o < TPR; o <<= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (part == 0) cc[r] = acc;
      }
      fence_proxy_async_smem();   // generic-proxy writes of A -> the tensor core's (async proxy) reads
      asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_WARPS * 32) : "memory");
      if (lane == 0) mbar_arrive(a_ready);
      const float ccr = (float)cc[row];
      double mu64 = 0.0, s64 = 0.0;
      float kmx = 0.f;
      for (int c = 0; c < ta.n_chunks; ++c, ++ch) {
        const int as = ch & 1;
        mbar_wait(&tfull[as], (ch >> 1) & 1);
        tc_fence_after();
        const float4* axp = ta.aux2 + (size_t)c * (TC_OBC / 2) + half * (TC_CW / 2);
        float mu32 = 0.f, s32 = 0.f;
#pragma unroll 1
        for (int sub = 0; sub < TC_CW / 32; ++sub) {
          float g[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * TC_OBC + half * TC_CW + sub * 32), g);
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            const float4 ax = __ldg(axp + sub * 16 + (j >> 1));   // {xx, alpha} of two observations: one address per warp
            const float T0 = fmaxf(fmaf(-2.f, g[j], ccr + ax.x), BD_T_FLOOR);
            const float T1 = fmaxf(fmaf(-2.f, g[j + 1], ccr + ax.z), BD_T_FLOOR);
            const float k0 = kernel_T<KIND>(T0), k1 = kernel_T<KIND>(T1);
            mu32 = fmaf(k0, ax.y, mu32);
            mu32 = fmaf(k1, ax.w, mu32);
            s32 = fmaf(k0, fabsf(ax.y), s32);
            s32 = fmaf(k1, fabsf(ax.w), s32);
            kmx = fmaxf(kmx, fmaxf(k0, k1));
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[as]);
        mu64 += (double)mu32;
        s64 += (double)s32;
      }
      red[(0 * TC_CG + half) * TC_CT + row] = mu64;
      red[(1 * TC_CG + half) * TC_CT + row] = s64;
      red[(2 * TC_CG + half) * TC_CT + row] = (double)kmx;
      asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_WARPS * 32) : "memory");
      if (half == 0 && row0 + row < a.M) {
        double m = 0.0, sS = 0.0, k = 0.0;
#pragma unroll
        for (int cg = 0; cg < TC_CG; ++cg) {
          m += red[(0 * TC_CG + cg) * TC_CT + row];
          sS += red[(1 * TC_CG + cg) * TC_CT + row];
          k = fmax(k, red[(2 * TC_CG + cg) * TC_CT + row]);
        }
        const double xm = *a.xx_max;
        const double rc = sqrt(cc[row]) + sqrt(xm);
        const double dT = a.g_T * rc * rc;
        double rel = (KIND == OB200_KERNEL_MATERN52 ? dT * (1.0 / 6.0) : dT) * 1.02 + a.c0 + a.c1 * (KIND == OB200_KERNEL_MATERN52 ? rc : rc * rc);
        if (!(xm < 1e29)) rel = 1.0;
        const double eps = rel < 0.5 ? a.y_std * (a.amplitude * sS * rel / (1.0 - rel) * 1.001 + a.tiny) : INFINITY;
        a.mu_lb[row0 + row] = a.y_std * (a.amplitude * m) + a.y_mean - eps;
        const double klo = rel < 0.5 ? a.amplitude * k * (1.0 - rel) : 0.0;
        const double var_ub = fmin(fmax(a.amplitude - klo * klo / (a.amplitude + a.diag), 0.0) + 1e-6 * a.amplitude, a.amplitude);
        a.sigma_ub[row0 + row] = a.y_std * sqrt(var_ub);
      }
      asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_WARPS * 32) : "memory");   // cc / red are rewritten by the next tile
    }
  }
  tc_fence_before();
  __syncthreads();
  __syncwarp();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
}

// observations -> UMMA canonical K-major layout (TF32 hi / lo) per (256-observation chunk, k-slice), and {xx, alpha} pairs
__global__ void __launch_bounds__(256) bound_pack_tc_kernel(const double* __restrict__ xs, const double* __restrict__ xx,
                                                            const double* __restrict__ alpha, int n, int D_pad, int ksl, int n_slices,
                                                            int n_chunks, double scale, float* __restrict__ pack, float4* __restrict__ aux2) {
  const int nd4 = D_pad >> 2;
  const long long per_slice = (long long)ksl * 2048;   // floats of one of hi / lo of a slice
  const long long total = (long long)n_chunks * n_slices * per_slice;
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < total; e += (long long)gridDim.x * 256) {
    const long long cs = e / per_slice;                 // chunk * n_slices + slice
    const int i = (int)(e - cs * per_slice);
    const int c = (int)(cs / n_slices), sl = (int)(cs - (long long)c * n_slices);
    const int el = i & 3, r0 = (i >> 2) & 7, g = (i >> 5) & 31, kc = i >> 10;   // [kc][g][r0][el] inside the slice
    const int obs = c * TC_OBC + g * 8 + r0, d = (sl * ksl * 2 + kc) * 4 + el;
    float v = 0.f;
    if (obs < n && d < D_pad) v = (float)(scale * xs[((size_t)(obs >> 3) * nd4 + (d >> 2)) * 32 + (obs & 7) * 4 + (d & 3)]);
    const uint32_t hi = to_tf32(v);
    pack[(size_t)cs * 2 * per_slice + i] = __uint_as_float(hi);
    pack[(size_t)cs * 2 * per_slice + per_slice + i] = __uint_as_float(to_tf32(v - __uint_as_float(hi)));
  }
  const long long n_aux = (long long)n_chunks * (TC_OBC / 2);
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < n_aux; e += (long long)gridDim.x * 256) {
    const long long j0 = 2 * e, j1 = j0 + 1;
    float4 out;
    out.x = j0 < n ? (float)(scale * scale * xx[j0]) : 1e30f;
    out.y = j0 < n ? (float)alpha[j0] : 0.f;
    out.z = j1 < n ? (float)(scale * scale * xx[j1]) : 1e30f;
    out.w = j1 < n ? (float)alpha[j1] : 0.f;
    aux2[e] = out;
  }
}

template <int WC, int WO, int NKR>
int launch_cfg(ob200_ctx* ctx, const bound_args& a, int grid, size_t smem) {
  auto kern = a.kind == OB200_KERNEL_MATERN52 ? bound_kernel<WC, WO, OB200_KERNEL_MATERN52, NKR> : bound_kernel<WC, WO, OB200_KERNEL_RBF, NKR>;
  OB_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<grid, (WC * WO + 1) * 32, smem, ctx->stream>>>(a);
  OB_LAUNCH_CHECK(ctx);
  return OB200_OK;
}

}  // namespace

// FP32 bound pass over `M` candidates: mu_lb[i] <= posterior mean, sigma_ub[i] >= posterior std.
// Returns OB200_OK and sets *done = 1 when the pass ran; *done = 0 when this fit cannot use it (D too wide for the
// shared-memory tile, OB200_BOUND_F32=0) and the caller must run the FP64 mean-only pass instead.
int ob200_bound_f32_launch(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, double* mu_lb_dev,
                           double* sigma_ub_dev, int* done) {
  *done = 0;
  if (const char* env = getenv("OB200_BOUND_F32")) if (env[0] == '0') return OB200_OK;
  const ob200_fit_header& h = ctx->hdr;
  const ob200_blob_layout lay = ob_blob_layout(h.n_pad, h.D_pad);
  const int nk = (h.D_pad + 7) / 8;
  const int n_slices = (nk + 3) / 4, ksl = (nk + n_slices - 1) / n_slices;
  const int n_chunks = (h.n + BD_OBC - 1) / BD_OBC;
  // tile shape: the most candidates per CTA whose A fragments + at least two B stages fit
  int wc = 4, stages = BD_STAGES_MAX;
  size_t smem = 0;
  for (;; ) {
    stages = BD_STAGES_MAX;
    while (stages > 2 && bound_smem_bytes(32 * wc, nk, ksl, stages) > 220 * 1024) --stages;
    smem = bound_smem_bytes(32 * wc, nk, ksl, stages);
    if (smem <= 220 * 1024) break;
    if (wc == 1) return OB200_OK;   // not even a 32-candidate tile fits: FP64 pass
    wc >>= 1;
  }
  const int ctas_per_sm = 1;   // one CTA of 4 WC + 1 warps per SM
  const int64_t bp_elems = (int64_t)n_chunks * n_slices * 16 * ksl * 32;   // float4
  const int64_t aux_elems = (int64_t)n_chunks * (BD_OBC / 8) * 4;
  int rc;
  if ((rc = ob_reserve(ctx, &ctx->b32_dev, &ctx->b32_cap, (bp_elems + aux_elems) * 4 + 2 * (2 + BD_PROBES) + 16))) return rc;
  float4* bp = reinterpret_cast<float4*>(ctx->b32_dev);
  float4* aux = bp + bp_elems;
  double* xx_max_dev = reinterpret_cast<double*>(aux + aux_elems);
  const double scale = h.kernel_kind == OB200_KERNEL_MATERN52 ? 2.23606797749979 : 0.7071067811865476;
  // ---- error budget (see the header comment).  u = 2^-24 (FP32 unit round-off), tf = 2^-11 (TF32).
  //   G~ = sum of 3 nk MMAs.  Split: c = ch + cl + rc, |rc| <= tf^2 |c| (+ 2u from the FP64 -> FP32 rounding), same
  //   for x; dropped cl*xl, rc*x, c*rx: <= (3 tf^2 + 4u) |c||x|.  Tensor-core accumulation: modelled, generously,
  //   as 2^-19 of sum |c_i||x_i| per MMA (published measurements put it near 2^-23 per added term).
  //   On T~ = cc + xx - 2 G~:  2 |dG| <= 2 (3 nk 2^-19 + 3 tf^2 + 4u) sqrt(cc xx) <= 0.5 (...) (sqrt cc + sqrt xx)^2,
  //   plus u (cc + xx) for the rounded norms and 2u (cc + xx + 2|G|) for the add and the fma.
  const double u = ldexp(1.0, -24), tf2 = ldexp(1.0, -22);
  const double g_T = 0.5 * (3.0 * nk * ldexp(1.0, -19) + 3.0 * tf2 + 4.0 * u) + 3.0 * u;
  //   function evaluation: sqrt.approx and ex2.approx (2^-21 each, generous), the argument product t * log2(e)
  //   (relative 2u of an exponent of size 1.4427 t -> relative ln2 * 1.4427 t * 2u = 2 t u of the result), the two
  //   polynomial FMAs and the final product (3u), alpha rounded to FP32 (u), the FMA into the partial sum (u), the
  //   FP32 partial sums (at most 2 * 128 / (4 WO) <= 64 terms per flush: 64u).  Matern: d ln k / dt <= 1, so the
  //   sqrt error 2^-21 t enters as 2^-21 t.
  const double c0 = 2.0 * ldexp(1.0, -21) + 70.0 * u, c1 = ldexp(1.0, -21) + 2.0 * u;
  if (ctx->b32_serial != ctx->fit_serial) {
    // once per fit: pack the observations, and decide whether the budget leaves a usable bound.  The mean's error is
    // y_std * amplitude * S * rel with S = sum_j k_j |alpha_j|; S is sampled (in FP64) at BD_PROBES observations and rel
    // taken at a candidate as far out as the farthest observation.  When that exceeds half a standard deviation of y
    // (tiny noise, long length scales, huge amplitude: |alpha| explodes while the mean stays O(1) by cancellation)
    // FP32 kernel values say nothing about the mean: FP64 pass.
    bound_pack_kernel<<<ctx->num_sms > BD_PROBES ? ctx->num_sms : BD_PROBES + 1, 256, 0, ctx->stream>>>(
        reinterpret_cast<const double*>(ctx->blob + lay.off_xs), reinterpret_cast<const double*>(ctx->blob + lay.off_xx),
        reinterpret_cast<const double*>(ctx->blob + lay.off_alpha), h.n, h.n_pad, h.D_pad, nk, ksl, n_slices, n_chunks, scale,
        h.kernel_kind, bp, aux, xx_max_dev);
    OB_LAUNCH_CHECK(ctx);
    double stats[2 + BD_PROBES];
    OB_CUDA(ctx, cudaMemcpyAsync(stats, xx_max_dev, sizeof stats, cudaMemcpyDeviceToHost, ctx->stream));
    OB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    double s_max = 0.0;
    for (int p = 0; p < BD_PROBES; ++p) s_max = stats[2 + p] > s_max ? stats[2 + p] : s_max;
    const double rc0 = 2.0 * sqrt(stats[0]);
    const bool matern = h.kernel_kind == OB200_KERNEL_MATERN52;
    const double rel0 = (matern ? g_T * rc0 * rc0 / 6.0 : g_T * rc0 * rc0) + c0 + c1 * (matern ? rc0 : rc0 * rc0);
    double limit = 0.5;
    if (const char* env = getenv("OB200_BOUND_F32_LIMIT")) limit = atof(env);   // (tuning / testing hook)
    ctx->b32_ok = stats[0] < 1e29 && s_max == s_max && h.amplitude * s_max * rel0 <= limit;
    ctx->b32_serial = ctx->fit_serial;
  }
  if (!ctx->b32_ok) return OB200_OK;

  bound_args a{};
  a.cand = cand_dev; a.ldc = ldc; a.M = M;
  a.ls = reinterpret_cast<const double*>(ctx->blob + lay.off_ls);
  a.bp = bp; a.aux = aux;
  a.n_chunks = n_chunks; a.D = h.D; a.nk = nk; a.ksl = ksl; a.n_slices = n_slices; a.kind = h.kernel_kind; a.stages = stages;
  a.scale = scale;
  a.amplitude = h.amplitude; a.y_mean = h.y_mean; a.y_std = h.y_std; a.diag = h.diag;
  a.xx_max = xx_max_dev;
  a.g_T = g_T; a.c0 = c0; a.c1 = c1;
  a.tiny = 1e-20;   // kernel values flushed to zero (< 2^-126) times any plausible ||alpha||_1
  a.mu_lb = mu_lb_dev; a.sigma_ub = sigma_ub_dev;
  const int ct = 32 * wc;
  a.n_tiles = (M + ct - 1) / ct;
  const int64_t max_grid = (int64_t)ctx->num_sms * ctas_per_sm;
  const int grid = (int)(a.n_tiles < max_grid ? a.n_tiles : max_grid);
  // tcgen05 variant: the contraction on the 5th-generation tensor cores, accumulators in TMEM.  A B stage is one k-slice
  // (<= 3 k-steps = 24 columns) of one 256-observation chunk, so its size does not grow with D; the candidate tile's own
  // fragments (8 KB per k-step) bound the width: D <= 120.
  const int n_sl_tc = (nk + 2) / 3, ksl_tc = (nk + n_sl_tc - 1) / n_sl_tc;
  int st_tc = TC_STAGES;
  while (st_tc > 2 && bound_tc_smem_bytes(nk, ksl_tc, st_tc) > 225 * 1024) --st_tc;
  bool use_tc = bound_tc_smem_bytes(nk, ksl_tc, st_tc) <= 225 * 1024;
  if (const char* env = getenv("OB200_BOUND_TC")) use_tc = use_tc && env[0] != '0';   // (OB200_BOUND_TC=0: the mma.sync kernel)
  if (use_tc) {
    const int n_chunks_tc = (h.n + TC_OBC - 1) / TC_OBC;
    const int64_t pack_floats = (int64_t)n_chunks_tc * n_sl_tc * 2 * ksl_tc * 2048, aux_floats = (int64_t)n_chunks_tc * (TC_OBC / 2) * 4;
    if ((rc = ob_reserve(ctx, &ctx->tc_dev, &ctx->tc_cap, pack_floats + aux_floats))) return rc;
    float4* aux2 = reinterpret_cast<float4*>(ctx->tc_dev + pack_floats);
    if (ctx->tc_serial != ctx->fit_serial) {
      bound_pack_tc_kernel<<<ctx->num_sms, 256, 0, ctx->stream>>>(
          reinterpret_cast<const double*>(ctx->blob + lay.off_xs), reinterpret_cast<const double*>(ctx->blob + lay.off_xx),
          reinterpret_cast<const double*>(ctx->blob + lay.off_alpha), h.n, h.D_pad, ksl_tc, n_sl_tc, n_chunks_tc, scale, ctx->tc_dev, aux2);
      OB_LAUNCH_CHECK(ctx);
      ctx->tc_serial = ctx->fit_serial;
    }
    bound_tc_args ta{};
    ta.b = a;
    ta.b.c0 = c0 + 70.0 * u;   // FP32 partial sums of up to 128 terms per flush here (64 above)
    ta.b.n_tiles = (M + TC_CT - 1) / TC_CT;
    ta.pack = ctx->tc_dev; ta.aux2 = aux2; ta.n_chunks = n_chunks_tc;
    ta.stages = st_tc; ta.ksl = ksl_tc; ta.n_slices = n_sl_tc;
    const size_t smem_tc = bound_tc_smem_bytes(nk, ksl_tc, st_tc);
    auto kern = a.kind == OB200_KERNEL_MATERN52 ? bound_tc_kernel<OB200_KERNEL_MATERN52> : bound_tc_kernel<OB200_KERNEL_RBF>;
    OB_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tc));
    const int grid_tc = (int)(ta.b.n_tiles < ctx->num_sms ? ta.b.n_tiles : ctx->num_sms);
    kern<<<grid_tc, TC_THREADS, smem_tc, ctx->stream>>>(ta);
    OB_LAUNCH_CHECK(ctx);
    *done = 1;
    return OB200_OK;
  }
  if (wc == 4 && nk == 1) rc = launch_cfg<4, 2, 1>(ctx, a, grid, smem);        // D <= 8
  else if (wc == 4 && nk == 2) rc = launch_cfg<4, 2, 2>(ctx, a, grid, smem);   // D <= 16
  else if (wc == 4 && nk == 3) rc = launch_cfg<4, 2, 3>(ctx, a, grid, smem);   // D <= 24 (BASELINE config 2)
  else if (wc == 4 && nk == 4) rc = launch_cfg<4, 2, 4>(ctx, a, grid, smem);   // D <= 32
  else if (wc == 4) rc = launch_cfg<4, 4, 0>(ctx, a, grid, smem);
  else if (wc == 2) rc = launch_cfg<2, 4, 0>(ctx, a, grid, smem);
  else rc = launch_cfg<1, 4, 0>(ctx, a, grid, smem);
  if (rc) return rc;
  *done = 1;
  return OB200_OK;
}
