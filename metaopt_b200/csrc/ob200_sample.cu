// a1-a12: Space.sample + TransformedSpace.transform + normalise, fused, HBM-write bound.
//
// Replaces the per-point Python loops of Space.sample (src/orion/algo/space.py:745-773),
// Real/_Discrete/Categorical.sample (:371-391, :410-423, :542-552) and the transformer chain of
// build_required_space(requires='real') (src/orion/core/worker/transformer.py:51-66, :219-221,
// :252-254, :269-282, :304-323) plus the skopt-side `(x-low)/(high-low)` normalisation.
//
// Layout: output is dense row-major [count x ld] float64.  A CTA produces a tile of 128 rows in
// shared memory; inside the tile a WARP owns one slot pair for 32 consecutive rows (lane = row), so
// control flow (kind / prior / number of categories) is warp-uniform and one Philox4x32-10 call
// yields the first-try uniforms of both slots.  The shared-memory row stride is the row width
// padded to 2 (mod 4) doubles: rows stay 16-byte aligned and the lane-per-row stores are at most
// 2-way bank conflicted.  The tile then leaves through the TMA engine: one bulk async copy per row
// (cp.async.bulk shared->global), or a single one for the whole tile when it is contiguous on both
// sides.  Odd widths / unaligned outputs fall back to coalesced scalar stores.  Two CTAs of 16
// warps fit per SM, so one CTA's generation overlaps the other's store phase.
//
// RNG contract (restated in oracle/philox.py + oracle/sampling.py, checked bit-exactly):
//   key = (seed_lo, seed_hi); counter = (index_lo, index_hi, c2, try)
//   RNG contract, version 2 (round 2; restated bit for bit in oracle/sampling.py + oracle/philox.py):
//   Slots are grouped by operation, table order inside a group.  NARROW groups come first -- uniform int, one-hot,
//   small categorical, each padded to a multiple of FOUR slots: such a slot needs no more than 32 random bits, so one
//   Philox call serves four of them (u = r_sel * 2^-32, exact in a double) -- then the WIDE groups -- uniform real,
//   loguniform real, loguniform int, normal real, normal int, each padded to an even size: two 53-bit uniforms per call.
//   q = position of the slot in that order; n_narrow = number of narrow positions (a multiple of 4).
//   item(q) = q >> 2 for narrow slots, n_narrow / 4 + (q - n_narrow) / 2 for wide ones; sel = q & 3 resp. q & 1.
//   uniform / loguniform / choice slot: c2 = item(q), component sel of the call
//   normal slot: c2 = 0x80000000|q, Box-Muller of the pair
//   try t = 0..3 (Real.sample's 4-try rejection, space.py:379-389); exhausted -> status latch.
#include "ob200_internal.cuh"

namespace {

#ifndef ST_TILE_BYTES
#define ST_TILE_BYTES 32768   // target size of a CTA's tile: rows per tile = as many 32-row groups as fit (>= 1); six such tiles
                              // stay resident per SM (the register limit), and 60-column spaces get 64 rows (+7 % over 32)
#endif
#ifndef ST_THREADS
#define ST_THREADS 128   // 4 warps share the slot pairs of the tile; many small independent CTAs per SM
#endif
#ifndef ST_NBUF
#define ST_NBUF 1        // tile buffers per CTA: with 2, a tile drains through the TMA engine while the next one is generated
#endif
#ifndef ST_MIN_CTAS
#define ST_MIN_CTAS 6   // 80 registers: with the quad (four narrow slots per Philox call) code the kernel spills 128 B at 72; measured
                        // on cfg5 3.74 TB/s with 7 CTAs / 72 registers, 4.04 with 6 / 80, 3.75 with 5 / 96
#endif

// Slot as the kernel reads it: 64 bytes = four 16-byte shared-memory loads, everything that only
// depends on the slot precomputed on the host (ob200_space_set).
struct __align__(16) dslot {
  double p0, p1;      // real/int: prior parameters; categorical: normalised values of 0.0 and 1.0
  double low, high;   // acceptance interval [low, high)
  double nlo, rspan;  // normalisation: (x - nlo) / span, rspan = RN(1 / span)
  double span;
  int col_src;        // out_col | (original slot index << 16)
  int op;             // OP_* below
};
static_assert(sizeof(dslot) == 64, "device slot must be 64 bytes");
struct dslot_cat { int ncat, cdf_off; };  // categorical slots reinterpret (low) as these two ints

enum { OP_UNIFORM = 0, OP_UNIFORM_INT = 1, OP_LOGUNIFORM = 2, OP_LOGUNIFORM_INT = 3, OP_NORMAL = 4, OP_NORMAL_INT = 5,
       OP_ONEHOT = 6, OP_CAT_SMALL = 7, OP_NONE = 8 };

// 2^(j/32), j = 0..31, correctly rounded (numpy.exp2): read through L1 (__ldg) so that the tile keeps all of the
// CTA's shared memory budget (8 CTAs per SM)
__device__ const double g_exp2_tab[32] = {
    0x1.0000000000000p+0, 0x1.059b0d3158574p+0, 0x1.0b5586cf9890fp+0, 0x1.11301d0125b51p+0,
    0x1.172b83c7d517bp+0, 0x1.1d4873168b9aap+0, 0x1.2387a6e756238p+0, 0x1.29e9df51fdee1p+0,
    0x1.306fe0a31b715p+0, 0x1.371a7373aa9cbp+0, 0x1.3dea64c123422p+0, 0x1.44e086061892dp+0,
    0x1.4bfdad5362a27p+0, 0x1.5342b569d4f82p+0, 0x1.5ab07dd485429p+0, 0x1.6247eb03a5585p+0,
    0x1.6a09e667f3bcdp+0, 0x1.71f75e8ec5f74p+0, 0x1.7a11473eb0187p+0, 0x1.82589994cce12p+0,
    0x1.8ace5422aa0dbp+0, 0x1.93737b0cdc5e5p+0, 0x1.9c49182a3f090p+0, 0x1.a5503b23e255cp+0,
    0x1.ae89f995ad3adp+0, 0x1.b7f76f2fb5e47p+0, 0x1.c199bdd85529cp+0, 0x1.cb720dcef9069p+0,
    0x1.d5818dcfba487p+0, 0x1.dfc97337b9b5fp+0, 0x1.ea4afa2a490dap+0, 0x1.f50765b6e4540p+0};

struct philox_keys { uint32_t k0[10], k1[10]; };  // round keys, precomputed on the host (constant bank operands)

__device__ __forceinline__ philox_out philox_rk(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, const philox_keys& rk) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ rk.k0[r], n2 = hi0 ^ c3 ^ rk.k1[r];
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
  }
  return philox_out{c0, c1, c2, c3};
}

// (hi:lo >> 11) * 2^-53 without a 64-bit integer conversion: both halves convert exactly and the
// fused sum of two exact products of powers of two is exact, so this equals u53() bit for bit.
__device__ __forceinline__ double u53_fast(uint32_t lo, uint32_t hi) {
  const uint32_t top = hi >> 11;                           // 21 bits
  const uint32_t bot = (hi << 21) | (lo >> 11);            // 32 bits
  return fma((double)top, 0x1.0p-21, (double)bot * 0x1.0p-53);
}

// r * 2^-32: 32 random bits as a uniform in [0, 1) (exact: one conversion and a power-of-two scale)
__device__ __forceinline__ double u32_unit(uint32_t r) { return (double)r * 0x1.0p-32; }

__device__ __forceinline__ double box_muller(philox_out r) {
  const double ua = u53_fast(r.r0, r.r1), ub = u53_fast(r.r2, r.r3);
  return __dmul_rn(sqrt(__dmul_rn(-2.0, log(__dsub_rn(1.0, ua)))), cos(__dmul_rn(6.283185307179586, ub)));
}

// a / b for a slot-constant divisor b with rb = RN(1/b) from the host: q0 = RN(a rb), exact residual
// r = a - q0 b (one FMA), q = RN(q0 + r rb).  With a correctly rounded reciprocal this final
// correction step returns the correctly rounded quotient (Markstein 1990), i.e. the same bits as the
// IEEE division numpy performs, at 3 instructions instead of ~21; tests compare against numpy.
__device__ __forceinline__ double div_const(double a, double b, double rb) {
  const double q0 = __dmul_rn(a, rb);
  const double r = fma(-q0, b, a);
  return fma(r, rb, q0);
}

// value of a real / integer slot; products and sums rounded separately like scipy (loc + scale * U)
template <int OP>
__device__ __forceinline__ double draw_op(const dslot& s, double u, double z, const double* __restrict__ exp_tab) {
  double x;
  if (OP == OP_UNIFORM || OP == OP_UNIFORM_INT) x = __dadd_rn(s.p0, __dmul_rn(s.p1, u));
  else if (OP == OP_LOGUNIFORM || OP == OP_LOGUNIFORM_INT) x = ob_exp_tab(__dadd_rn(s.p0, __dmul_rn(u, s.p1)), exp_tab);  // ~2 ulp, branch-free
  else x = __dadd_rn(s.p0, __dmul_rn(s.p1, z));
  if (OP & 1) x = floor(x);
  return x;
}

// Real.sample's retries (space.py:379-389), tries 1..3.  Rare: out of line.  NaN = all rejected.
__device__ __noinline__ double retry_slot(const dslot* s, int q, int n_narrow, uint64_t index, const philox_keys* rk, const double* exp_tab) {
  for (int t = 1; t < 4; ++t) {
    double u = 0.0, z = 0.0;
    if (s->op == OP_NORMAL || s->op == OP_NORMAL_INT) {
      z = box_muller(philox_rk((uint32_t)index, (uint32_t)(index >> 32), 0x80000000u | (uint32_t)q, (uint32_t)t, *rk));
    } else if (q < n_narrow) {   // narrow slot: 32 bits of call q >> 2
      const philox_out r = philox_rk((uint32_t)index, (uint32_t)(index >> 32), (uint32_t)(q >> 2), (uint32_t)t, *rk);
      const int sel = q & 3;
      u = u32_unit(sel == 0 ? r.r0 : (sel == 1 ? r.r1 : (sel == 2 ? r.r2 : r.r3)));
    } else {
      const philox_out r = philox_rk((uint32_t)index, (uint32_t)(index >> 32), (uint32_t)((n_narrow >> 2) + ((q - n_narrow) >> 1)),
                                     (uint32_t)t, *rk);
      u = (q & 1) ? u53_fast(r.r2, r.r3) : u53_fast(r.r0, r.r1);
    }
    double xt;
    switch (s->op) {
      case OP_UNIFORM: xt = draw_op<OP_UNIFORM>(*s, u, z, exp_tab); break;
      case OP_UNIFORM_INT: xt = draw_op<OP_UNIFORM_INT>(*s, u, z, exp_tab); break;
      case OP_LOGUNIFORM: xt = draw_op<OP_LOGUNIFORM>(*s, u, z, exp_tab); break;
      case OP_LOGUNIFORM_INT: xt = draw_op<OP_LOGUNIFORM_INT>(*s, u, z, exp_tab); break;
      case OP_NORMAL: xt = draw_op<OP_NORMAL>(*s, u, z, exp_tab); break;
      default: xt = draw_op<OP_NORMAL_INT>(*s, u, z, exp_tab); break;
    }
    if ((xt >= s->low) && (xt < s->high)) return xt;
  }
  return __longlong_as_double(0x7ff8000000000000LL);
}

// Kernel modes (template parameter): the hot generation paths know `normalise` at compile time and carry no
// per-row predicates; the parity mode (host-supplied variates) keeps the runtime flags.
enum { MODE_RAW = 0, MODE_NORM = 1, MODE_VARIATES = 2 };

// shared-memory store through a 32-bit shared address (the generic-pointer form re-derives the window base per store)
__device__ __forceinline__ void sts_f64(uint32_t addr, double v) {
  asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}

// Per-row context of one lane.  Rows past the end of a ragged last tile still generate (their smem rows exist and
// are never copied out); `active` only gates the status latch and the parity-mode loads.
struct row_ctx {
  uint64_t index;
  uint32_t row_sa;          // shared address of the row in the tile
  const double* var_row;    // parity mode only
  bool active;
};

template <int OP, int MODE>
__device__ __forceinline__ void real_slot(const dslot* __restrict__ sp, int q, int n_narrow, const row_ctx& rc, const philox_keys& rk, double u0,
                                          int normalise, int* __restrict__ status, const double* __restrict__ exp_tab) {
  const dslot s = *sp;  // four LDS.128, broadcast
  double x;
  bool ok;
  if (MODE == MODE_VARIATES) {  // parity mode: one host-supplied variate, no resampling
    const double v = rc.active ? rc.var_row[s.col_src >> 16] : 0.5;
    x = draw_op<OP>(s, v, v, exp_tab);
    ok = (x >= s.low) && (x < s.high);
  } else {
    double z = 0.0;
    if (OP == OP_NORMAL || OP == OP_NORMAL_INT)
      z = box_muller(philox_rk((uint32_t)rc.index, (uint32_t)(rc.index >> 32), 0x80000000u | (uint32_t)q, 0u, rk));
    x = draw_op<OP>(s, u0, z, exp_tab);
    ok = (x >= s.low) && (x < s.high);
    if (__any_sync(0xffffffffu, !ok)) {
      if (!ok) { x = retry_slot(sp, q, n_narrow, rc.index, &rk, exp_tab); ok = (x == x); }
    }
  }
  if (!ok) { if (rc.active) atomicOr(status, 1); x = 0.0; }
  if (MODE == MODE_NORM || (MODE == MODE_VARIATES && normalise)) x = div_const(__dsub_rn(x, s.nlo), s.span, s.rspan);
  sts_f64(rc.row_sa + (uint32_t)(s.col_src & 0xffff) * 8u, x);
}

// numpy searchsorted(cdf, u, side='right') clamped to k-1: number of entries <= u
__device__ __forceinline__ int cdf_count(const double* __restrict__ cdf, int k, double u) {
  int n = 0;
  if (k <= 4) {
#pragma unroll
    for (int j = 0; j < 4; ++j) n += (j < k && cdf[j] <= u) ? 1 : 0;
  } else {
    int lo = 0, hi = k;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (cdf[mid] <= u) lo = mid + 1; else hi = mid; }
    n = lo;
  }
  return n < k - 1 ? n : k - 1;
}

// one-hot group of a compile-time width: K - 1 compares (the clamp of cdf_count absorbs the last one), K selects
template <int K>
__device__ __forceinline__ void onehot_fixed(const double* __restrict__ cdf, double u, double v0, double v1, uint32_t dst_sa) {
  int hot = 0;
#pragma unroll
  for (int j = 0; j < K - 1; ++j) hot += (cdf[j] <= u) ? 1 : 0;
#pragma unroll
  for (int j = 0; j < K; ++j) sts_f64(dst_sa + 8u * j, (j == hot) ? v1 : v0);
}

template <bool ONEHOT, int MODE>
__device__ __forceinline__ void cat_slot(const dslot* __restrict__ sp, const double* __restrict__ cdf, double u0, const row_ctx& rc,
                                         int normalise) {
  const dslot s = *sp;
  const dslot_cat c = *reinterpret_cast<const dslot_cat*>(&sp->low);
  const double u = (MODE == MODE_VARIATES) ? (rc.active ? rc.var_row[s.col_src >> 16] : 0.0) : u0;
  const bool norm = MODE == MODE_NORM || (MODE == MODE_VARIATES && normalise);
  const double v0 = norm ? s.p0 : 0.0, v1 = norm ? s.p1 : 1.0;
  const uint32_t dst = rc.row_sa + (uint32_t)(s.col_src & 0xffff) * 8u;
  if (ONEHOT && c.ncat <= 4) {  // (one-hot means ncat >= 3) straight-line code for the common widths
    if (c.ncat == 4) onehot_fixed<4>(cdf + c.cdf_off, u, v0, v1, dst);
    else onehot_fixed<3>(cdf + c.cdf_off, u, v0, v1, dst);
    return;
  }
  const int hot = cdf_count(cdf + c.cdf_off, c.ncat, u);
  if (ONEHOT) {
    for (int j = 0; j < c.ncat; ++j) sts_f64(dst + 8u * j, (j == hot) ? v1 : v0);
  } else {
    sts_f64(dst, hot ? v1 : v0);  // k == 2: float(index); k == 1: 0.0
  }
}

// Two real / integer slots of one pair side by side: their dependency chains (draw, range test, exact division)
// are independent, so keeping them in one basic block lets the scheduler overlap them.  A padding second slot
// (all-zero parameters) is computed too and just not stored.
template <int OP, int MODE>
__device__ __forceinline__ void real_pair(const dslot* __restrict__ sp, int q, int n_narrow, const row_ctx& rc, const philox_keys& rk, double ua,
                                          double ub, int normalise, int* __restrict__ status, const double* __restrict__ exp_tab) {
  const dslot s0 = sp[0], s1 = sp[1];
  const bool two = s1.op != OP_NONE;
  double z0 = 0.0, z1 = 0.0;
  if (OP == OP_NORMAL || OP == OP_NORMAL_INT) {
    z0 = box_muller(philox_rk((uint32_t)rc.index, (uint32_t)(rc.index >> 32), 0x80000000u | (uint32_t)q, 0u, rk));
    z1 = box_muller(philox_rk((uint32_t)rc.index, (uint32_t)(rc.index >> 32), 0x80000000u | (uint32_t)(q + 1), 0u, rk));
  }
  double x0 = draw_op<OP>(s0, ua, z0, exp_tab), x1 = draw_op<OP>(s1, ub, z1, exp_tab);
  bool ok0 = (x0 >= s0.low) && (x0 < s0.high), ok1 = !two || ((x1 >= s1.low) && (x1 < s1.high));
  if (__any_sync(0xffffffffu, !(ok0 && ok1))) {
    if (!ok0) { x0 = retry_slot(sp, q, n_narrow, rc.index, &rk, exp_tab); ok0 = (x0 == x0); }
    if (!ok1) { x1 = retry_slot(sp + 1, q + 1, n_narrow, rc.index, &rk, exp_tab); ok1 = (x1 == x1); }
    if (!(ok0 && ok1)) {
      if (rc.active) atomicOr(status, 1);
      if (!ok0) x0 = 0.0;
      if (!ok1) x1 = 0.0;
    }
  }
  if (MODE == MODE_NORM) {
    x0 = div_const(__dsub_rn(x0, s0.nlo), s0.span, s0.rspan);
    x1 = div_const(__dsub_rn(x1, s1.nlo), s1.span, s1.rspan);
  }
  sts_f64(rc.row_sa + (uint32_t)(s0.col_src & 0xffff) * 8u, x0);
  if (two) sts_f64(rc.row_sa + (uint32_t)(s1.col_src & 0xffff) * 8u, x1);
}

// Two one-hot groups of width 4 side by side (the common case); anything else goes slot by slot.
template <int MODE>
__device__ __forceinline__ bool onehot4_pair(const dslot* __restrict__ sp, const double* __restrict__ cdf, double ua, double ub,
                                             const row_ctx& rc) {
  const dslot_cat c0 = *reinterpret_cast<const dslot_cat*>(&sp[0].low), c1 = *reinterpret_cast<const dslot_cat*>(&sp[1].low);
  if (c0.ncat != 4 || c1.ncat != 4 || sp[1].op == OP_NONE) return false;
  const double2 p0 = *reinterpret_cast<const double2*>(&sp[0].p0), p1 = *reinterpret_cast<const double2*>(&sp[1].p0);
  const double a0 = MODE == MODE_NORM ? p0.x : 0.0, a1 = MODE == MODE_NORM ? p0.y : 1.0;
  const double b0 = MODE == MODE_NORM ? p1.x : 0.0, b1 = MODE == MODE_NORM ? p1.y : 1.0;
  const double *ca = cdf + c0.cdf_off, *cb = cdf + c1.cdf_off;
  const int ha = (ca[0] <= ua ? 1 : 0) + (ca[1] <= ua ? 1 : 0) + (ca[2] <= ua ? 1 : 0);
  const int hb = (cb[0] <= ub ? 1 : 0) + (cb[1] <= ub ? 1 : 0) + (cb[2] <= ub ? 1 : 0);
  const uint32_t da = rc.row_sa + (uint32_t)(sp[0].col_src & 0xffff) * 8u, db = rc.row_sa + (uint32_t)(sp[1].col_src & 0xffff) * 8u;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    sts_f64(da + 8u * j, (j == ha) ? a1 : a0);
    sts_f64(db + 8u * j, (j == hb) ? b1 : b0);
  }
  return true;
}

// Both slots of a pair share the operation (groups are padded to even size), so one warp-uniform
// switch per pair selects straight-line code for the two of them.
template <int OP, int MODE>
__device__ __forceinline__ void pair_op(const dslot* __restrict__ sp, const double* __restrict__ cdf, int q, int n_narrow, const row_ctx& rc,
                                        const philox_keys& rk, double ua, double ub, int normalise, int* __restrict__ status,
                                        const double* __restrict__ exp_tab) {
  if (OP == OP_ONEHOT || OP == OP_CAT_SMALL) {
    if (OP == OP_ONEHOT && MODE != MODE_VARIATES && onehot4_pair<MODE>(sp, cdf, ua, ub, rc)) return;
    cat_slot<OP == OP_ONEHOT, MODE>(sp, cdf, ua, rc, normalise);
    if (sp[1].op != OP_NONE) cat_slot<OP == OP_ONEHOT, MODE>(sp + 1, cdf, ub, rc, normalise);
  } else if (MODE != MODE_VARIATES) {
    real_pair<OP, MODE>(sp, q, n_narrow, rc, rk, ua, ub, normalise, status, exp_tab);
  } else {
    real_slot<OP, MODE>(sp, q, n_narrow, rc, rk, ua, normalise, status, exp_tab);
    if (sp[1].op != OP_NONE) real_slot<OP, MODE>(sp + 1, q + 1, n_narrow, rc, rk, ub, normalise, status, exp_tab);
  }
}

template <int MODE>
__global__ void __launch_bounds__(ST_THREADS, ST_MIN_CTAS)
sample_transform_kernel(const dslot* __restrict__ slots_g, int n_dslots, int n_narrow, int n_slots, const double* __restrict__ cdf_g, int n_cdf,
                        int d_out, int stride, int tile_rows, const __grid_constant__ philox_keys rk, uint64_t first_index, int64_t count,
                        int normalise, double* __restrict__ out, int64_t ld, const double* __restrict__ variates,
                        int bulk_ok, int vec_lanes, int* __restrict__ status) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const size_t tile_elems = ((size_t)tile_rows * stride + 1) & ~(size_t)1;               // 16-byte aligned
  double* tile0 = reinterpret_cast<double*>(smem_raw);                                  // ST_NBUF x [tile_rows][stride]
  dslot* slots = reinterpret_cast<dslot*>(tile0 + ST_NBUF * tile_elems);
  double* cdf = reinterpret_cast<double*>(slots + n_dslots);
  const double* exp_tab = g_exp2_tab;
  for (int i = threadIdx.x; i < n_dslots * (int)(sizeof(dslot) / 8); i += ST_THREADS)
    reinterpret_cast<double*>(slots)[i] = reinterpret_cast<const double*>(slots_g)[i];
  for (int i = threadIdx.x; i < n_cdf; i += ST_THREADS) cdf[i] = cdf_g[i];
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_quads = n_narrow >> 2;
  const int n_items = n_quads + ((n_dslots - n_narrow) >> 1);   // one Philox call each: a quad of narrow slots or a pair of wide ones
  const int64_t n_tiles = (count + tile_rows - 1) / tile_rows;
  const uint32_t row_bytes = (uint32_t)d_out * 8u;

  int buf = 0;
  int q0 = warp;   // this warp's next item in the round-robin over (tile, row group, Philox call): carried across row groups and tiles
  for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, buf ^= (ST_NBUF - 1)) {
    double* tile = tile0 + (size_t)buf * tile_elems;
    const uint32_t tile_sa = (uint32_t)__cvta_generic_to_shared(tile);
    const int64_t row0 = t * tile_rows;
    const int rows = (int)min((int64_t)tile_rows, count - row0);
#ifdef ST_DEBUG_NO_COMPUTE  // timing experiment only: copy-out without generation
    const int n_rg = 0;
#else
    const int n_rg = (rows + 31) >> 5;
#endif
    // items = (row group, Philox call), dealt round-robin to the warps -- the narrow calls (four 32-bit slots each) and
    // the wide calls (two 53-bit slots each) in two loops of their own, so that each loop's switch holds only its
    // operations and the call's words die right after the conversion (one common loop cost 9 registers and spilled)
    const int n_wide = n_items - n_quads;
    for (int rg = 0; rg < n_rg; ++rg) {
      const int r = rg * 32 + lane;
      row_ctx rc;
      rc.active = r < rows;
      rc.index = first_index + (uint64_t)(row0 + r);
      rc.row_sa = tile_sa + (uint32_t)r * (uint32_t)stride * 8u;
      rc.var_row = (MODE == MODE_VARIATES) ? variates + (size_t)(row0 + (rc.active ? r : 0)) * n_slots : nullptr;
#define ST_NARROW(SP, Q, UA, UB)                                                                                              \
  switch ((SP)->op) {                                                                                                           \
    case OP_UNIFORM_INT: pair_op<OP_UNIFORM_INT, MODE>(SP, cdf, Q, n_narrow, rc, rk, UA, UB, normalise, status, exp_tab); break; \
    case OP_ONEHOT: pair_op<OP_ONEHOT, MODE>(SP, cdf, Q, n_narrow, rc, rk, UA, UB, normalise, status, exp_tab); break;           \
    default: pair_op<OP_CAT_SMALL, MODE>(SP, cdf, Q, n_narrow, rc, rk, UA, UB, normalise, status, exp_tab); break;               \
  }
      int p = q0;
      for (; p < n_quads; p += ST_THREADS / 32) {
        const dslot* sp = slots + 4 * p;
        double u0 = 0.0, u1 = 0.0, u2 = 0.0, u3 = 0.0;
        if (MODE != MODE_VARIATES) {
          const philox_out pr = philox_rk((uint32_t)rc.index, (uint32_t)(rc.index >> 32), (uint32_t)p, 0u, rk);
          u0 = u32_unit(pr.r0); u1 = u32_unit(pr.r1); u2 = u32_unit(pr.r2); u3 = u32_unit(pr.r3);
        }
        ST_NARROW(sp, 4 * p, u0, u1)
        if (sp[2].op != OP_NONE) { ST_NARROW(sp + 2, 4 * p + 2, u2, u3) }
      }
#undef ST_NARROW
      p -= n_quads;  // the round-robin continues into the wide calls, the next row group and the next tile: an even load
      for (; p < n_wide; p += ST_THREADS / 32) {
        const int q = n_narrow + 2 * p;
        const dslot* sp = slots + q;
        double ua = 0.0, ub = 0.0;
        if (MODE != MODE_VARIATES) {
          const philox_out pr = philox_rk((uint32_t)rc.index, (uint32_t)(rc.index >> 32), (uint32_t)(n_quads + p), 0u, rk);
          ua = u53_fast(pr.r0, pr.r1);
          ub = u53_fast(pr.r2, pr.r3);
        }
        switch (sp->op) {
          case OP_UNIFORM: pair_op<OP_UNIFORM, MODE>(sp, cdf, q, n_narrow, rc, rk, ua, ub, normalise, status, exp_tab); break;
          case OP_LOGUNIFORM: pair_op<OP_LOGUNIFORM, MODE>(sp, cdf, q, n_narrow, rc, rk, ua, ub, normalise, status, exp_tab); break;
          case OP_LOGUNIFORM_INT: pair_op<OP_LOGUNIFORM_INT, MODE>(sp, cdf, q, n_narrow, rc, rk, ua, ub, normalise, status, exp_tab); break;
          case OP_NORMAL: pair_op<OP_NORMAL, MODE>(sp, cdf, q, n_narrow, rc, rk, ua, ub, normalise, status, exp_tab); break;
          default: pair_op<OP_NORMAL_INT, MODE>(sp, cdf, q, n_narrow, rc, rk, ua, ub, normalise, status, exp_tab); break;
        }
      }
      q0 = p - n_wide;
    }
    if (bulk_ok) {
      fence_proxy_async_smem();  // generic-proxy smem writes -> async-proxy (TMA) reads
#if ST_NBUF == 2
      if (threadIdx.x == 0) bulk_wait_read0();  // the previous tile has left the other buffer (it had a whole tile's time)
#endif
      __syncthreads();
      if (warp == 0) {
        if (bulk_ok == 2) {
#ifndef ST_DEBUG_NO_STORE  // timing experiment only: generation without copy-out
          if (lane == 0) bulk_s2g(out + row0 * ld, tile, (uint32_t)rows * row_bytes);
#endif
        } else {
          for (int r = lane; r < rows; r += 32) bulk_s2g(out + (row0 + r) * ld, tile + (size_t)r * stride, row_bytes);
        }
        bulk_commit();
#if ST_NBUF == 1
        bulk_wait_read0();  // the tile is rewritten next iteration
#endif
      }
#if ST_NBUF == 1
      __syncthreads();
#endif
    } else if (vec_lanes) {
      __syncthreads();
      // 16-byte coalesced stores: a power-of-two group of lanes per row
      const int sub = lane / vec_lanes, l = lane - sub * vec_lanes, rows_per_iter = 32 / vec_lanes;
      for (int r = warp * rows_per_iter + sub; r < rows; r += (ST_THREADS / 32) * rows_per_iter) {
        const double2* src = reinterpret_cast<const double2*>(tile + (size_t)r * stride);
        double2* dst = reinterpret_cast<double2*>(out + (row0 + r) * ld);
        for (int c = l; c < (d_out >> 1); c += vec_lanes) dst[c] = src[c];
      }
#if ST_NBUF == 1
      __syncthreads();
#endif
    } else {
      __syncthreads();
      for (int r = warp; r < rows; r += ST_THREADS / 32) {
        const double* src = tile + (size_t)r * stride;
        double* dst = out + (row0 + r) * ld;
        for (int c = lane; c < d_out; c += 32) dst[c] = src[c];
      }
#if ST_NBUF == 1
      __syncthreads();
#endif
    }
  }
  if (bulk_ok && warp == 0) bulk_wait0();
}

// host: ABI slot -> device slot
dslot make_dslot(const ob200_slot& s, int src) {
  dslot d{};
  d.col_src = s.out_col | (src << 16);
  d.nlo = s.nlo;
  d.span = s.nhi - s.nlo;
  d.rspan = 1.0 / d.span;
  if (s.kind == OB200_KIND_CAT) {
    d.op = s.ncat > 2 ? OP_ONEHOT : OP_CAT_SMALL;
    d.p0 = (0.0 - s.nlo) / d.span;  // same IEEE operations as numpy's (x - low) / (high - low)
    d.p1 = (1.0 - s.nlo) / d.span;
    dslot_cat c{s.ncat, s.cdf_off};
    memcpy(&d.low, &c, sizeof c);
  } else {
    const int base = s.prior == OB200_PRIOR_UNIFORM ? OP_UNIFORM : (s.prior == OB200_PRIOR_LOGUNIFORM ? OP_LOGUNIFORM : OP_NORMAL);
    d.op = base + (s.kind == OB200_KIND_INT ? 1 : 0);
    d.p0 = s.p0; d.p1 = s.p1; d.low = s.low; d.high = s.high;
  }
  return d;
}

}  // namespace

int ob200_sample_launch(ob200_ctx* ctx, uint64_t seed, uint64_t first_index, int64_t count, int normalise,
                        double* out_dev, int64_t ld, const double* variates_dev) {
  if (ctx->n_slots <= 0) OB_FAIL(ctx, OB200_ESTATE, "ob200_space_set has not been called");
  if (count < 0) OB_FAIL(ctx, OB200_EINVAL, "count must be >= 0");
  if (count == 0) return OB200_OK;   // Space.sample(0) is an empty list; an empty tensor has no address
  if (!out_dev || ld < ctx->d_out) OB_FAIL(ctx, OB200_EINVAL, "bad out/ld (ld=%lld, D_out=%d)", (long long)ld, ctx->d_out);
  const int d_out = ctx->d_out;
  // 16-byte paths need an even width, an even leading dimension and a 16-byte aligned base.  Shared-memory
  // row stride: width padded to 2 (mod 4) doubles (16-byte aligned rows, <= 2-way conflicts); else odd stride.
  const int aligned = (d_out % 2 == 0) && (ld % 2 == 0) && ((reinterpret_cast<uintptr_t>(out_dev) & 15) == 0);
  // A dense output (ld == D_out) keeps the tile dense in shared memory too, so that it leaves as ONE bulk copy: for widths
  // that are a multiple of 4 doubles the column writes then take 8 shared-memory wavefronts instead of 4, which costs far
  // less than the cooperative copy-out loop it replaces (20 reals: 1.65 -> 2.16 TB/s, the cfg3 space: 3.51 -> 4.05).
  int stride = aligned ? ((d_out % 4 == 2 || ld == d_out) ? d_out : d_out + 2) : (d_out | 1);
  if (const char* env = getenv("OB200_ST_PADDED")) { if (env[0] != '0' && aligned && d_out % 4 == 0) stride = d_out + 2; }   // (testing hook)
  // copy-out: 2 = one bulk copy per tile (contiguous both sides), 1 = one bulk copy per row (rows >= ST_ROW_BULK_MIN),
  // 0 = cooperative stores (16-byte vectors when aligned)
  int bulk_ok = 0, vec_lanes = 0;
  if (aligned && stride == d_out && ld == d_out) bulk_ok = 2;
#ifndef ST_ROW_BULK_MIN
#define ST_ROW_BULK_MIN 256   // bytes: narrower rows leave through cooperative 16-byte stores (measured: 480 B rows +12 % with bulk copies, 160 B rows -3 %)
#endif
  else if (aligned && d_out * 8 >= ST_ROW_BULK_MIN) bulk_ok = 1;
  else if (aligned) { vec_lanes = 1; while (vec_lanes < 32 && vec_lanes < d_out / 2) vec_lanes <<= 1; }
  // rows per tile: narrow spaces get several 32-row groups per CTA (the slot pairs of ONE group do not divide evenly
  // among the warps and the per-tile barriers / copy-out would dominate), wide ones a single group
  int tile_rows = 32 * (int)(ST_TILE_BYTES / ((size_t)32 * stride * 8));
  if (tile_rows < 32) tile_rows = 32;
  if (tile_rows > 256) tile_rows = 256;
  if (const char* env = getenv("OB200_TILE_ROWS")) { const int v = atoi(env); if (v >= 32 && v <= 1024 && v % 32 == 0) tile_rows = v; }
  const size_t smem = (ST_NBUF * (((size_t)tile_rows * stride + 1) & ~(size_t)1) + (size_t)ctx->n_cdf) * sizeof(double) + (size_t)ctx->n_dslots * sizeof(dslot);
  if (smem > 227 * 1024) OB_FAIL(ctx, OB200_EINVAL, "D_out=%d too wide for the sampler tile", d_out);
  const int64_t n_tiles = (count + tile_rows - 1) / tile_rows;
  auto kernel = variates_dev ? sample_transform_kernel<MODE_VARIATES>
                             : (normalise ? sample_transform_kernel<MODE_NORM> : sample_transform_kernel<MODE_RAW>);
  OB_CUDA(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;  // resident CTAs per SM for this instantiation (registers and shared memory): one wave, grid-stride
  OB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, ST_THREADS, smem));
  if (per_sm < 1) per_sm = 1;
  const int64_t max_grid = (int64_t)ctx->num_sms * per_sm;
  const int grid = (int)(n_tiles < max_grid ? n_tiles : max_grid);
  philox_keys rk;
  {
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    for (int r = 0; r < 10; ++r) { rk.k0[r] = k0; rk.k1[r] = k1; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
  }
  kernel<<<grid, ST_THREADS, smem, ctx->stream>>>(reinterpret_cast<const dslot*>(ctx->dslots_dev), ctx->n_dslots, ctx->n_narrow, ctx->n_slots, ctx->cdf_dev, ctx->n_cdf, d_out,
                                                  stride, tile_rows, rk, first_index, count, normalise, out_dev, ld, variates_dev, bulk_ok, vec_lanes,
                                                  ctx->status_dev);
  OB_LAUNCH_CHECK(ctx);
  return OB200_OK;
}

extern "C" {

int ob200_space_set(ob200_ctx* ctx, const ob200_slot* slots_host, int n_slots, const double* cdf_host, int n_cdf) {
  if (!ctx) return OB200_EINVAL;
  if (!slots_host || n_slots <= 0 || n_cdf < 0 || (n_cdf > 0 && !cdf_host)) OB_FAIL(ctx, OB200_EINVAL, "bad slot table");
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  int d_out = 0;
  for (int i = 0; i < n_slots; ++i) {
    const ob200_slot& s = slots_host[i];
    if (s.kind < 0 || s.kind > 2 || s.prior < 0 || s.prior > 3 || s.width < 1 || s.out_col < 0)
      OB_FAIL(ctx, OB200_EINVAL, "slot %d: bad kind/prior/width/out_col", i);
    if (s.kind == OB200_KIND_CAT) {
      if (s.ncat < 1 || s.cdf_off < 0 || s.cdf_off + s.ncat > n_cdf) OB_FAIL(ctx, OB200_EINVAL, "slot %d: bad categorical table", i);
      if (s.width != (s.ncat > 2 ? s.ncat : 1)) OB_FAIL(ctx, OB200_EINVAL, "slot %d: width must be k for k>2 else 1", i);
      if (s.prior != OB200_PRIOR_CHOICE) OB_FAIL(ctx, OB200_EINVAL, "slot %d: categorical slots use the choice prior", i);
    } else {
      if (s.width != 1 || s.prior == OB200_PRIOR_CHOICE) OB_FAIL(ctx, OB200_EINVAL, "slot %d: bad real/integer slot", i);
      if (!(s.high > s.low)) OB_FAIL(ctx, OB200_EINVAL, "slot %d: Lower bound %g has to be less than upper bound %g", i, s.low, s.high);
    }
    if (!(s.nhi > s.nlo)) OB_FAIL(ctx, OB200_EINVAL, "slot %d: empty normalisation interval", i);
    if (s.out_col + s.width > d_out) d_out = s.out_col + s.width;
  }
  OB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  cudaFree(ctx->slots_dev); ctx->slots_dev = nullptr;
  cudaFree(ctx->dslots_dev); ctx->dslots_dev = nullptr;
  cudaFree(ctx->cdf_dev); ctx->cdf_dev = nullptr;
  cudaFree(ctx->col_nlo_dev); ctx->col_nlo_dev = nullptr;
  cudaFree(ctx->col_nhi_dev); ctx->col_nhi_dev = nullptr;
  ctx->n_slots = 0;
  OB_CUDA(ctx, cudaMalloc((void**)&ctx->slots_dev, (size_t)n_slots * sizeof(ob200_slot)));
  OB_CUDA(ctx, cudaMemcpy(ctx->slots_dev, slots_host, (size_t)n_slots * sizeof(ob200_slot), cudaMemcpyHostToDevice));
  int n_dslots = 0;
  {
    if (n_slots >= 32768 || d_out >= 65536) OB_FAIL(ctx, OB200_EINVAL, "space too wide (slots < 32768, columns < 65536)");
    // group by operation -- this order defines the RNG position of a slot (contract at the top of this file): the
    // narrow groups first, each padded with no-op slots to a multiple of four, then the wide ones, padded to even
    std::vector<dslot> ds;
    static const int narrow_ops[] = {OP_UNIFORM_INT, OP_ONEHOT, OP_CAT_SMALL};
    static const int wide_ops[] = {OP_UNIFORM, OP_LOGUNIFORM, OP_LOGUNIFORM_INT, OP_NORMAL, OP_NORMAL_INT};
    dslot pad{};
    pad.op = OP_NONE;
    for (int op : narrow_ops) {
      for (int i = 0; i < n_slots; ++i) {
        const dslot d = make_dslot(slots_host[i], i);
        if (d.op == op) ds.push_back(d);
      }
      while (ds.size() & 3) ds.push_back(pad);
    }
    ctx->n_narrow = (int)ds.size();
    for (int op : wide_ops) {
      for (int i = 0; i < n_slots; ++i) {
        const dslot d = make_dslot(slots_host[i], i);
        if (d.op == op) ds.push_back(d);
      }
      if (ds.size() & 1) ds.push_back(pad);
    }
    n_dslots = (int)ds.size();
    OB_CUDA(ctx, cudaMalloc((void**)&ctx->dslots_dev, ds.size() * sizeof(dslot)));
    OB_CUDA(ctx, cudaMemcpy(ctx->dslots_dev, ds.data(), ds.size() * sizeof(dslot), cudaMemcpyHostToDevice));
  }
  OB_CUDA(ctx, cudaMalloc((void**)&ctx->cdf_dev, (size_t)(n_cdf > 0 ? n_cdf : 1) * sizeof(double)));
  if (n_cdf > 0) OB_CUDA(ctx, cudaMemcpy(ctx->cdf_dev, cdf_host, (size_t)n_cdf * sizeof(double), cudaMemcpyHostToDevice));
  std::vector<double> lo(d_out, 0.0), hi(d_out, 1.0);
  for (int i = 0; i < n_slots; ++i)
    for (int j = 0; j < slots_host[i].width; ++j) {
      lo[slots_host[i].out_col + j] = slots_host[i].nlo;
      hi[slots_host[i].out_col + j] = slots_host[i].nhi;
    }
  OB_CUDA(ctx, cudaMalloc((void**)&ctx->col_nlo_dev, (size_t)d_out * sizeof(double)));
  OB_CUDA(ctx, cudaMalloc((void**)&ctx->col_nhi_dev, (size_t)d_out * sizeof(double)));
  OB_CUDA(ctx, cudaMemcpy(ctx->col_nlo_dev, lo.data(), (size_t)d_out * sizeof(double), cudaMemcpyHostToDevice));
  OB_CUDA(ctx, cudaMemcpy(ctx->col_nhi_dev, hi.data(), (size_t)d_out * sizeof(double), cudaMemcpyHostToDevice));
  ctx->n_slots = n_slots;
  ctx->n_dslots = n_dslots;
  ctx->n_cdf = n_cdf;
  ctx->d_out = d_out;
  return OB200_OK;
}

int ob200_sample_transform(ob200_ctx* ctx, uint64_t seed, uint64_t first_index, int64_t count, int normalise,
                           double* out_dev, int64_t ld, const double* variates_dev) {
  if (!ctx) return OB200_EINVAL;
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  return ob200_sample_launch(ctx, seed, first_index, count, normalise, out_dev, ld, variates_dev);
}

int ob200_sample_status(ob200_ctx* ctx) {
  if (!ctx) return OB200_EINVAL;
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  int st = 0;
  OB_CUDA(ctx, cudaMemcpyAsync(&st, ctx->status_dev, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  OB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (st) {
    OB_CUDA(ctx, cudaMemsetAsync(ctx->status_dev, 0, sizeof(int), ctx->stream));
    OB_FAIL(ctx, OB200_EBOUNDS, "Improbable bounds: a draw was rejected 4 times. Please make interval larger.");
  }
  return OB200_OK;
}

int ob200_sample_transform_host(ob200_ctx* ctx, uint64_t seed, uint64_t first_index, int64_t count, int normalise,
                                double* out_host, int64_t ld, const double* variates_host) {
  if (!ctx) return OB200_EINVAL;
  if (count < 0) OB_FAIL(ctx, OB200_EINVAL, "count must be >= 0");
  if (count == 0) return OB200_OK;
  if (!out_host || ld < ctx->d_out) OB_FAIL(ctx, OB200_EINVAL, "bad out/ld");
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  const int64_t out_elems = count * ctx->d_out, var_elems = variates_host ? count * ctx->n_slots : 0;
  int rc = ob_reserve(ctx, &ctx->cand_tmp, &ctx->cand_cap, out_elems + var_elems);
  if (rc) return rc;
  double* var_dev = nullptr;
  if (variates_host) {
    var_dev = ctx->cand_tmp + out_elems;
    OB_CUDA(ctx, cudaMemcpyAsync(var_dev, variates_host, (size_t)var_elems * 8, cudaMemcpyHostToDevice, ctx->stream));
  }
  rc = ob200_sample_launch(ctx, seed, first_index, count, normalise, ctx->cand_tmp, ctx->d_out, var_dev);
  if (rc) return rc;
  OB_CUDA(ctx, cudaMemcpy2DAsync(out_host, (size_t)ld * 8, ctx->cand_tmp, (size_t)ctx->d_out * 8, (size_t)ctx->d_out * 8,
                                 (size_t)count, cudaMemcpyDeviceToHost, ctx->stream));
  OB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return OB200_OK;
}

}  // extern "C"
