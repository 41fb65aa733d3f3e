// a1-a12: Space.sample + TransformedSpace.transform + normalise, fused, HBM-write bound.
//
// Replaces the per-point Python loops of Space.sample (src/orion/algo/space.py:745-773),
// Real/_Discrete/Categorical.sample (:371-391, :410-423, :542-552) and the transformer chain of
// build_required_space(requires='real') (src/orion/core/worker/transformer.py:51-66, :219-221,
// :252-254, :269-282, :304-323) plus the skopt-side `(x-low)/(high-low)` normalisation.
//
// Layout: output is dense row-major [count x ld] float64.  A CTA produces a tile of R rows into
// shared memory (one thread per (row, slot pair): one Philox4x32-10 call yields the first-try
// uniforms of two slots), then the whole contiguous tile leaves with ONE bulk async copy
// (cp.async.bulk shared->global, TMA engine) when ld == D_out, else with coalesced stores.
// Two tiles of shared memory are in flight so generation overlaps the store.
//
// RNG contract (restated in oracle/philox.py + oracle/sampling.py, checked bit-exactly):
//   key = (seed_lo, seed_hi); counter = (index_lo, index_hi, c2, try)
//   uniform / loguniform / choice slot s: c2 = s>>1, uniform = component (s&1) of the pair
//   normal slot s: c2 = 0x80000000|s, Box-Muller of the pair
//   try t = 0..3 (Real.sample's 4-try rejection, space.py:379-389); exhausted -> status latch.
#include "ob200_internal.cuh"

namespace {

__device__ __forceinline__ int cdf_search_right(const double* __restrict__ cdf, int k, double u) {
  // numpy searchsorted(cdf, u, side='right'): number of entries <= u
  int lo = 0, hi = k;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (cdf[mid] <= u) lo = mid + 1; else hi = mid;
  }
  return lo < k - 1 ? lo : k - 1;
}

__device__ __forceinline__ double draw_value(const ob200_slot& s, double u, double z) {
  // no FMA contraction: the reference rounds the product and the sum separately
  if (s.prior == OB200_PRIOR_UNIFORM) return __dadd_rn(s.p0, __dmul_rn(s.p1, u));
  if (s.prior == OB200_PRIOR_LOGUNIFORM) return exp(__dadd_rn(s.p0, __dmul_rn(u, s.p1)));
  return __dadd_rn(s.p0, __dmul_rn(s.p1, z));
}

__device__ __forceinline__ void emit_slot(const ob200_slot& s, double value, int normalise,
                                          double* __restrict__ row_out) {
  const double span = __dsub_rn(s.nhi, s.nlo);
  if (s.kind == OB200_KIND_CAT && s.ncat > 2) {
    const int hot = (int)value;
    for (int j = 0; j < s.width; ++j) {
      double v = (j == hot) ? 1.0 : 0.0;
      if (normalise) v = __ddiv_rn(__dsub_rn(v, s.nlo), span);
      row_out[s.out_col + j] = v;
    }
  } else {
    double v = (s.kind == OB200_KIND_CAT && s.ncat == 1) ? 0.0 : value;
    if (normalise) v = __ddiv_rn(__dsub_rn(v, s.nlo), span);
    row_out[s.out_col] = v;
  }
}

// one slot of one point; u0 = first-try uniform already drawn (non-normal priors)
__device__ __forceinline__ void do_slot(const ob200_slot& s, int si, const double* __restrict__ cdf,
                                        uint64_t index, uint32_t k0, uint32_t k1, double u0,
                                        const double* __restrict__ variate, int normalise,
                                        double* __restrict__ row_out, int* __restrict__ status) {
  double value = 0.0;
  bool done = false;
  const int tries = variate ? 1 : 4;
  for (int t = 0; t < tries && !done; ++t) {
    double u = u0, z = 0.0;
    if (variate) {
      u = z = *variate;
    } else if (s.prior == OB200_PRIOR_NORMAL) {
      philox_out r = philox4x32_10((uint32_t)index, (uint32_t)(index >> 32), 0x80000000u | (uint32_t)si,
                                   (uint32_t)t, k0, k1);
      double ua = u53(r.r0, r.r1), ub = u53(r.r2, r.r3);
      z = __dmul_rn(sqrt(__dmul_rn(-2.0, log(__dsub_rn(1.0, ua)))), cos(__dmul_rn(6.283185307179586, ub)));
    } else if (t > 0) {
      philox_out r = philox4x32_10((uint32_t)index, (uint32_t)(index >> 32), (uint32_t)(si >> 1),
                                   (uint32_t)t, k0, k1);
      u = (si & 1) ? u53(r.r2, r.r3) : u53(r.r0, r.r1);
    }
    double x;
    bool ok;
    if (s.kind == OB200_KIND_CAT) {
      x = (double)cdf_search_right(cdf + s.cdf_off, s.ncat, u);
      ok = true;
    } else {
      x = draw_value(s, u, z);
      if (s.kind == OB200_KIND_INT) x = floor(x);
      ok = (x >= s.low) && (x < s.high);
    }
    if (ok) { value = x; done = true; }
  }
  if (!done) atomicOr(status, 1);
  emit_slot(s, value, normalise, row_out);
}

__global__ void __launch_bounds__(256)
sample_transform_kernel(const ob200_slot* __restrict__ slots, int n_slots, const double* __restrict__ cdf,
                        int d_out, uint64_t seed, uint64_t first_index, int64_t count, int normalise,
                        double* __restrict__ out, int64_t ld, const double* __restrict__ variates,
                        int rows_per_tile, int use_bulk, int* __restrict__ status) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* tile[2];
  tile[0] = reinterpret_cast<double*>(smem_raw);
  tile[1] = tile[0] + (size_t)rows_per_tile * d_out;
  const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  const int n_pairs = (n_slots + 1) >> 1;
  const int64_t n_tiles = (count + rows_per_tile - 1) / rows_per_tile;
  int buf = 0;
  for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, buf ^= 1) {
    const int64_t row0 = t * rows_per_tile;
    const int rows = (int)min((int64_t)rows_per_tile, count - row0);
    double* sm = tile[buf];
    // the bulk store that last read this buffer (two iterations ago) must have drained
    if (use_bulk && threadIdx.x == 0) bulk_wait_read1();
    __syncthreads();
    const int items = rows * n_pairs;
    for (int it = threadIdx.x; it < items; it += blockDim.x) {
      const int r = it / n_pairs, p = it - r * n_pairs;
      const uint64_t index = first_index + (uint64_t)(row0 + r);
      double* row_out = sm + (size_t)r * d_out;
      const int s0 = 2 * p, s1 = 2 * p + 1;
      double ua = 0.0, ub = 0.0;
      if (!variates) {
        philox_out pr = philox4x32_10((uint32_t)index, (uint32_t)(index >> 32), (uint32_t)p, 0u, k0, k1);
        ua = u53(pr.r0, pr.r1);
        ub = u53(pr.r2, pr.r3);
      }
      const double* v0 = variates ? variates + (size_t)(row0 + r) * n_slots + s0 : nullptr;
      {
        const ob200_slot s = slots[s0];
        do_slot(s, s0, cdf, index, k0, k1, ua, v0, normalise, row_out, status);
      }
      if (s1 < n_slots) {
        const ob200_slot s = slots[s1];
        do_slot(s, s1, cdf, index, k0, k1, ub, v0 ? v0 + 1 : nullptr, normalise, row_out, status);
      }
    }
    const bool bulk_ok = use_bulk && ((((int64_t)rows * d_out) & 1) == 0);
    if (bulk_ok) {
      fence_proxy_async();  // make generic-proxy smem writes visible to the async proxy
      __syncthreads();
      if (threadIdx.x == 0) {
        bulk_s2g(out + row0 * ld, sm, (uint32_t)((size_t)rows * d_out * sizeof(double)));
        bulk_commit();
      }
    } else {
      __syncthreads();
      const int elems = rows * d_out;
      for (int e = threadIdx.x; e < elems; e += blockDim.x) {
        const int r = e / d_out, c = e - r * d_out;
        out[(row0 + r) * ld + c] = sm[e];
      }
    }
  }
  if (use_bulk && threadIdx.x == 0) bulk_wait0();
}

}  // namespace

int ob200_sample_launch(ob200_ctx* ctx, uint64_t seed, uint64_t first_index, int64_t count, int normalise,
                        double* out_dev, int64_t ld, const double* variates_dev) {
  if (ctx->n_slots <= 0) OB_FAIL(ctx, OB200_ESTATE, "ob200_space_set has not been called");
  if (count < 0 || !out_dev || ld < ctx->d_out) OB_FAIL(ctx, OB200_EINVAL, "bad count/out/ld (ld=%lld, D_out=%d)", (long long)ld, ctx->d_out);
  if (count == 0) return OB200_OK;
  const int d_out = ctx->d_out;
  int rows = (int)(49152 / ((int64_t)d_out * 8));  // 48 KiB per buffer, two buffers
  if (rows > 128) rows = 128;
  if (rows < 1) OB_FAIL(ctx, OB200_EINVAL, "D_out=%d too wide for the sampler tile", d_out);
  if (rows > 1) rows &= ~1;
  const size_t smem = (size_t)2 * rows * d_out * sizeof(double);
  const int use_bulk = (ld == d_out) && ((reinterpret_cast<uintptr_t>(out_dev) & 15) == 0) && (rows % 2 == 0 || d_out % 2 == 0);
  const int64_t n_tiles = (count + rows - 1) / rows;
  int grid = (int)(n_tiles < (int64_t)ctx->num_sms * 4 ? n_tiles : (int64_t)ctx->num_sms * 4);
  OB_CUDA(ctx, cudaFuncSetAttribute(sample_transform_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 49152));
  sample_transform_kernel<<<grid, 256, smem, ctx->stream>>>(ctx->slots_dev, ctx->n_slots, ctx->cdf_dev, d_out,
                                                            seed, first_index, count, normalise, out_dev, ld,
                                                            variates_dev, rows, use_bulk, ctx->status_dev);
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
  cudaFree(ctx->cdf_dev); ctx->cdf_dev = nullptr;
  cudaFree(ctx->col_nlo_dev); ctx->col_nlo_dev = nullptr;
  cudaFree(ctx->col_nhi_dev); ctx->col_nhi_dev = nullptr;
  ctx->n_slots = 0;
  OB_CUDA(ctx, cudaMalloc((void**)&ctx->slots_dev, (size_t)n_slots * sizeof(ob200_slot)));
  OB_CUDA(ctx, cudaMemcpy(ctx->slots_dev, slots_host, (size_t)n_slots * sizeof(ob200_slot), cudaMemcpyHostToDevice));
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
  if (!out_host || count < 0 || ld < ctx->d_out) OB_FAIL(ctx, OB200_EINVAL, "bad out/count/ld");
  OB_CUDA(ctx, cudaSetDevice(ctx->device));
  if (count == 0) return OB200_OK;
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
