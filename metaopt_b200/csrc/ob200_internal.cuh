// Internal definitions shared by the libob200 translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>

#include "ob200.h"

#define OB200_ABI_VERSION 1
#define OB200_MAX_N 8192      // observed points (n_obs) supported by the fit
#define OB200_NUM_SMS_FALLBACK 148

// ---- tiling constants of the score kernel (see DESIGN.md "score kernel") ------------------
// Consumer warps form a SC_WM x SC_WN grid; each owns SC_FM row fragments (8 rows of inv(L)) x
// SC_FN candidate fragments (8 candidates).  Defaults: 12x1 warps, 2x4 fragments -> 192 x 32 tile:
// 32 candidates per CTA keep the 148 live k* slices (n_pad x 32 doubles each; 80 MB at n = 2000)
// inside the 126 MB L2 together with inv(L) (L2 hit rate 97 % instead of 74 % with 64 candidates,
// 4x less DRAM traffic, same speed); 3 consumer warps per SM sub-partition keep the DMMA pipe fed.
#ifndef SC_WM
#define SC_WM 12
#endif
#ifndef SC_WN
#define SC_WN 1
#endif
#ifndef SC_FM
#define SC_FM 2
#endif
#ifndef SC_FN
#define SC_FN 4
#endif
#define SC_BM (SC_WM * SC_FM * 8)   // rows of inv(L) per row block
#define SC_BN (SC_WN * SC_FN * 8)   // candidates per CTA tile
#define SC_NCF (SC_BN / 8)          // candidate fragments per tile
#ifndef SC_BK
#define SC_BK 32         // k (observed-point) extent of one pipeline stage
#endif
#ifndef SC_STAGES
#define SC_STAGES 3
#endif
#define SC_CONSUMER_WARPS (SC_WM * SC_WN)
#define SC_PRODUCER_WARPS 4   // a full warpgroup so that setmaxnreg can rebalance registers
#define SC_THREADS ((SC_CONSUMER_WARPS + SC_PRODUCER_WARPS) * 32)
#ifndef SC_CONSUMER_REGS      // per-SMSP budget: (consumer warps / 4) * regs + 40 <= 512
#define SC_CONSUMER_REGS ((SC_CONSUMER_WARPS == 8) ? 232 : 152)
#endif
#define SC_P1_OBS (SC_CONSUMER_WARPS * 8)   // observed points per phase-1 chunk (one obs8 block per warp)

struct ob200_fit_header {   // first bytes of the exported fit blob
  int32_t magic, n, D, n_pad, D_pad, kernel_kind;
  double amplitude, y_mean, y_std, y_min, lml;
  double diag;   // alpha + noise: the term added to the diagonal of K (bounds the posterior variance from one observation)
};

struct ob200_ctx {
  int device = 0;
  int num_sms = OB200_NUM_SMS_FALLBACK;
  bool l2_dirty = false;   // persisting lines of a finished score launch have not been released yet
  bool l2_limit_on = false;  // the device-wide persisting-L2 carve-out is currently set (only around score launches)
  size_t l2_prev_limit = 0;  // what the host had set before ob_l2_acquire (restored by ob_l2_release)
  size_t l2_persist_max = 0, l2_window_max = 0;  // persisting-L2 set-aside for the k* slices (0 = unavailable/disabled)
  cudaStream_t stream = 0;
  std::string err;
  int64_t launches = 0;

  // space
  int n_slots = 0, d_out = 0, n_cdf = 0, n_dslots = 0, n_narrow = 0;
  ob200_slot* slots_dev = nullptr;   // device copy of the ABI table (kept for reference)
  char* dslots_dev = nullptr;        // packed 64-byte device slots used by the sampler kernel
  double* cdf_dev = nullptr;
  double* col_nlo_dev = nullptr;   // per output column normalisation bounds
  double* col_nhi_dev = nullptr;
  int* status_dev = nullptr;       // latched sampler status

  // fit state: one device blob [header | ls[D_pad] | Xs packed | xx | alpha | Linv packed]
  bool fitted = false;
  ob200_fit_header hdr{};
  char* blob = nullptr;
  int64_t blob_bytes = 0, blob_cap = 0;
  double* L_dev = nullptr;         // dense L (n_pad x n_pad) kept for read-back / lml
  int64_t L_cap = 0;
  double* work_dev = nullptr;      // fit workspace
  int64_t work_cap = 0;
  bool have_local_fit = false;     // work_dev / L_dev hold the factors of the current fit (not an imported one)
  int64_t w_Xs_off = 0, w_y_off = 0;  // offsets (doubles) of X/ls and normalised y inside work_dev
  int64_t w_t_off = 0, w_stats_off = 0;  // ... of the scratch vector and the statistics block
  double fit_alpha = 0.0;             // diagonal term of the current local fit (ob200_gp_append re-uses it)
  double* grad_dev = nullptr;      // LML-gradient workspace
  int64_t grad_cap = 0;
  double* pg_dev = nullptr;        // posterior-gradient workspace (ob200_gp_predict_grad_host)
  int64_t pg_cap = 0;

  // score scratch
  double* kstar_dev = nullptr;     // per-CTA K* slices
  int64_t kstar_cap = 0;
  double* cand_tmp = nullptr;      // candidates generated on device / staged from host
  int64_t cand_cap = 0;
  double* dense_tmp = nullptr;     // staged dense outputs for the host entry point
  int64_t dense_cap = 0;
  char* partial_dev = nullptr;     // per-CTA argmax partials (bytes)
  int64_t partial_cap = 0;
  ob200_best* best_dev = nullptr;  // [3]
  ob200_best* best_pinned = nullptr;
  double* ub_dev = nullptr;        // pruned scoring: mu / upper bound per candidate
  int64_t ub_cap = 0;
  long long* idx_dev = nullptr;    // pruned scoring: compacted survivor indices
  int64_t idx_cap = 0;
  double* gather_dev = nullptr;    // pruned scoring: gathered survivor rows
  int64_t gather_cap = 0;
  char* prune_small_dev = nullptr; // pruned scoring: device-resident control block (histograms, thresholds, counters)
  int64_t prune_small_cap = 0;
  void* ctl_pinned = nullptr;      // pinned read-back of the control block's counters
  int last_prune_f32 = 0;          // statistics of the last pruned call (ob200_prune_stats)
  int64_t last_prune_counts[4] = {0, 0, 0, -1};
  int64_t fit_serial = 0;          // bumped by every fit / append / import: keys the caches below
  int64_t b32_serial = -1;         // fit the FP32 pack (b32_dev) and its verdict belong to
  int b32_ok = 0;                  // this fit may use the FP32 bound pass (its error budget is not vacuous)
  ob200_ctx* sub = nullptr;        // pruned scoring: GP on a subset of the observations (variance upper bound)
  int64_t sub_serial = -1;
  float* tc_dev = nullptr;         // FP32 bound pass, tcgen05 variant: observations in the UMMA canonical layout
  int64_t tc_cap = 0, tc_serial = -1;
  float* b32_dev = nullptr;        // FP32 bound pass: TF32 hi/lo B fragments + per-observation constants
  int64_t b32_cap = 0;
  cudaStream_t fit_stream = nullptr;   // trailing updates of the Cholesky look-ahead
  cudaEvent_t ev_panel = nullptr, ev_trail = nullptr;
  cudaStream_t fit_stream2 = nullptr;  // the next diagonal block's update + factorisation, beside the block column's update
  cudaEvent_t ev_diag = nullptr;
  cudaStream_t copy_stream = nullptr;  // H2D of the candidate tail overlapped with scoring of the head
  cudaEvent_t ev_main = nullptr, ev_copy = nullptr;
  double* pinned = nullptr;        // pinned staging for host entry points
  int64_t pinned_cap = 0;
};

#define OB_FAIL(ctx, code, ...)                               \
  do {                                                        \
    char _b[512];                                             \
    snprintf(_b, sizeof _b, __VA_ARGS__);                     \
    (ctx)->err = _b;                                          \
    return (code);                                            \
  } while (0)

#define OB_CUDA(ctx, expr)                                                              \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess)                                                              \
      OB_FAIL(ctx, _e == cudaErrorMemoryAllocation ? OB200_ENOMEM : OB200_ECUDA,        \
              "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

#define OB_LAUNCH_CHECK(ctx)        \
  do {                              \
    (ctx)->launches++;              \
    OB_CUDA(ctx, cudaGetLastError()); \
  } while (0)

// grow-only device buffer helper
template <typename T>
static inline int ob_reserve(ob200_ctx* ctx, T** ptr, int64_t* cap, int64_t want_elems) {
  if (*cap >= want_elems && *ptr) return OB200_OK;
  if (*ptr) {
    OB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    OB_CUDA(ctx, cudaFree(*ptr));
    *ptr = nullptr;
    *cap = 0;
  }
  OB_CUDA(ctx, cudaMalloc((void**)ptr, (size_t)want_elems * sizeof(T)));
  *cap = want_elems;
  return OB200_OK;
}

static inline int64_t ob_round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

// blob layout (all offsets in bytes, 256-aligned)
struct ob200_blob_layout {
  int64_t off_ls, off_xs, off_xx, off_alpha, off_linv, total;
};
static inline ob200_blob_layout ob_blob_layout(int n_pad, int D_pad) {
  ob200_blob_layout l;
  int64_t o = 256;  // header
  l.off_ls = o; o += ob_round_up((int64_t)D_pad * 8, 256);
  l.off_xs = o;    o += ob_round_up((int64_t)n_pad * D_pad * 8, 256);
  l.off_xx = o;    o += ob_round_up((int64_t)n_pad * 8, 256);
  l.off_alpha = o; o += ob_round_up((int64_t)n_pad * 8, 256);
  l.off_linv = o;  o += (int64_t)n_pad * n_pad * 8;
  l.total = o;
  return l;
}

#ifdef __CUDACC__
// ---- device helpers -----------------------------------------------------------------------
// D(8x8) += A(8x4, row) * B(4x8, col), FP64 tensor core (SASS: DMMA.8x8x4).
// lane l: a = A[l/4][l%4], b = B[l%4][l/4], c0/c1 = C[l/4][2*(l%4) + {0,1}]
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D bulk async copy global -> shared (TMA engine; SASS: UBLKCP), completion on an mbarrier.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// 1-D bulk async copy shared -> global (TMA engine), bulk-group completion.
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_read1() {
  asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }
// The same for shared memory only (generic-proxy writes to this CTA's shared memory -> async-proxy reads of it): without the
// state space the fence covers global memory too and ptxas puts a MEMBAR.ALL.GPU in front of it.
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- Philox4x32-10 (Salmon et al. 2011); restated bit-for-bit in oracle/philox.py -----------
struct philox_out { uint32_t r0, r1, r2, r3; };
__device__ __forceinline__ philox_out philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                    uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return philox_out{c0, c1, c2, c3};
}
__device__ __forceinline__ double u53(uint32_t lo, uint32_t hi) {
  uint64_t x = ((uint64_t)hi << 32) | lo;
  return (double)(x >> 11) * 0x1.0p-53;
}
// exp(t) with a 32-entry table (any t whose result is a normal double; used with t <= 0 by the score kernel and
// with log-uniform exponents by the sampler): t = (32 n + j) ln2/32 + r, |r| <= ln2/64, so
// exp(t) = 2^n * 2^(j/32) * exp(r) and a degree-6 Taylor polynomial suffices (truncation 3.5e-18);
// Cody-Waite two-constant reduction, scaling by exponent-field addition.  ~2 ulp; below 2^-1021 -> 0.
// `tab` = 2^(j/32), j = 0..31, in shared memory.
// Constants whose low word is not zero cannot be FP64 immediates; as literals the compiler rebuilds them with two moves
// per use when registers are short (the mean-only pass runs at 72).  From the constant bank they are operands.
static __constant__ double ob_kc[3] = {46.16624130844683 /* 32 / ln2 */, 0.3333333333333333, 2.23606797749979 /* sqrt(5) */};

__device__ __forceinline__ double ob_exp_tab(double t, const double* __restrict__ tab) {
  const double magic = 6755399441055744.0;         // 1.5 * 2^52: round-to-nearest-integer trick
  const double mb = fma(t, ob_kc[0], magic);  // t * 32 / ln2
  const int m = __double2loint(mb);
  const double md = mb - magic;
  double r = fma(md, -2.16608493865351192653e-02, t);   // ln2/32, high part
  r = fma(md, -5.96317165397058692146e-12, r);          // ln2/32, low part
  double p = 1.0 / 720.0;
  p = fma(p, r, 1.0 / 120.0);
  p = fma(p, r, 1.0 / 24.0);
  p = fma(p, r, 1.0 / 6.0);
  p = fma(p, r, 0.5);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  const double v = tab[m & 31] * p;
  const int n = m >> 5;
  const double scaled = __hiloint2double(__double2hiint(v) + (n << 20), __double2loint(v));
  return (n >= -1021) ? scaled : 0.0;
}

// The persisting-L2 carve-out is a DEVICE-WIDE setting: while it is on, streaming writes of every kernel on the
// device drop from 7.4 to 2.8 TB/s (tools/l2_persist_probe.py).  It is therefore switched on for score launches
// only and handed back as soon as a score call has been synchronised.
inline void ob_l2_acquire(ob200_ctx* ctx) {
  if (ctx->l2_persist_max == 0 || ctx->l2_limit_on) return;
  size_t prev = 0;
  if (cudaDeviceGetLimit(&prev, cudaLimitPersistingL2CacheSize) != cudaSuccess) prev = 0;
  if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, ctx->l2_persist_max) == cudaSuccess) {
    ctx->l2_prev_limit = prev;   // the host framework's own set-aside (usually 0) comes back on release
    ctx->l2_limit_on = true;
  }
}
inline void ob_l2_release(ob200_ctx* ctx) {
  if (ctx->l2_dirty) { cudaCtxResetPersistingL2Cache(); ctx->l2_dirty = false; }
  if (ctx->l2_limit_on) { cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, ctx->l2_prev_limit); ctx->l2_limit_on = false; }
}

#endif  // __CUDACC__

// internal cross-TU entry points
int ob200_sample_launch(ob200_ctx* ctx, uint64_t seed, uint64_t first_index, int64_t count, int normalise,
                        double* out_dev, int64_t ld, const double* variates_dev);
int ob200_bound_f32_launch(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, double* mu_lb_dev,
                           double* sigma_ub_dev, int* done);
