"""GPU box: randomised parity sweep of the device path against the oracle (shapes, seeds and spaces the fixed
parametrisations of tests/ do not hit).  Stops at the first mismatch with the generating recipe; prints a summary.
usage: python tools/fuzz_parity.py [seconds] [seed]"""
import collections
import json
import os
import sys
import time
import traceback

import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import test_gpu_sample as tsamp  # noqa: E402
import test_gpu_score as tscore  # noqa: E402
from helpers import transformed  # noqa: E402
from metaopt_b200 import compiler  # noqa: E402
from metaopt_b200.engine import get_engine  # noqa: E402
from metaopt_b200.space import Categorical, Integer, Real, Space  # noqa: E402
from oracle import acquisition as oacq  # noqa: E402
from oracle import gp as ogp  # noqa: E402
from oracle import sampling as osamp  # noqa: E402

os.environ.setdefault("OB200_PRUNE_MIN_BATCHES", "1")   # the pruned legs want the pruned path at every n
BUDGET = float(sys.argv[1]) if len(sys.argv) > 1 else 240.0
rng = np.random.RandomState(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
eng = get_engine(0)
counts = collections.Counter({"sample": 0, "score": 0, "generated": 0, "pruned": 0, "grad": 0})


def random_space():
    s = Space()
    spec = []
    for i in range(rng.randint(1, 14)):
        kind = rng.randint(0, 6)
        name = "d%02d" % i
        shape = None if rng.rand() < 0.8 else int(rng.randint(1, 4))
        if kind == 0:
            lo = float(rng.uniform(-50, 50))
            args = (lo, lo + float(rng.uniform(0.1, 100)))
            s.register(Real(name, "uniform", args[0], args[1] - args[0], shape=shape))
            spec.append(("real-uniform", args, shape))
        elif kind == 1:
            a = float(10 ** rng.uniform(-6, 0))
            b = a * float(10 ** rng.uniform(0.5, 6))
            s.register(Real(name, "reciprocal", a, b, shape=shape))
            spec.append(("real-reciprocal", (a, b), shape))
        elif kind == 2:
            loc, scale = float(rng.uniform(-5, 5)), float(rng.uniform(0.5, 3))
            lo, hi = loc - float(rng.uniform(2.5, 5)) * scale, loc + float(rng.uniform(2.5, 5)) * scale
            s.register(Real(name, "norm", loc, scale, low=lo, high=hi, shape=shape))
            spec.append(("real-norm", (loc, scale, lo, hi), shape))
        elif kind == 3:
            lo = int(rng.randint(-20, 20))
            width = int(rng.randint(1, 1000))
            s.register(Integer(name, "uniform", lo, width, shape=shape))
            spec.append(("int-uniform", (lo, width), shape))
        else:
            k = int(rng.randint(1, 7))
            if shape is not None and k > 2:
                shape = None  # 2-D transformed points are rejected by Orion's flatten_dims and by the compiler
            cats = ["c%d" % j for j in range(k)]
            if rng.rand() < 0.5:
                p = rng.dirichlet(np.ones(k))
                p = np.round(p, 3)
                p[-1] = 1.0 - p[:-1].sum()
                if p.min() <= 0:
                    p = np.ones(k) / k
                s.register(Categorical(name, dict(zip(cats, p.tolist())), shape=shape))
                spec.append(("cat", p.tolist(), shape))
            else:
                s.register(Categorical(name, cats, shape=shape))
                spec.append(("cat", k, shape))
    return s, spec


def fuzz_sample():
    s, spec = random_space()
    comp = compiler.compile_space(transformed(s))
    compiler.upload(eng, comp)
    seed = int(rng.randint(0, 2 ** 31)) * int(rng.randint(1, 2 ** 31)) + int(rng.randint(0, 2 ** 31))
    first = int(rng.choice([0, 1, 31, 2 ** 32 - 5, 2 ** 40 + 17, int(rng.randint(0, 2 ** 31))]))
    count = int(rng.choice([1, 2, 31, 32, 33, 127, 1025, int(rng.randint(1, 5000))]))
    normalise = bool(rng.randint(0, 2))
    recipe = dict(kind="sample", spec=spec, seed=seed, first=first, count=count, normalise=normalise)
    dev = eng.sample(seed, first, count, normalise=normalise).cpu().numpy()
    ref, status = osamp.sample_transform(comp.slots, comp.cdf, seed, first, count, normalise=normalise)
    assert status == 0, recipe
    tsamp._check(dev, ref, comp)
    counts["sample"] += 1
    return recipe


def fuzz_score():
    n = int(rng.choice([1, 2, 7, 63, 64, 65, 191, 192, 193, int(rng.randint(1, 900))]))
    d = int(rng.choice([1, 2, 3, 4, 5, 19, 20, 21, int(rng.randint(1, 120))]))
    M = int(rng.choice([1, 2, 31, 32, 33, 4735, 4736, 4737, int(rng.randint(1, 6000))]))
    kind = int(rng.randint(0, 2))
    norm_y = bool(rng.randint(0, 2))
    amp = float(10 ** rng.uniform(-1, 1))
    ls = (rng.uniform(0.3, 2.0, size=d) * np.sqrt(d) / 2) if rng.rand() < 0.7 else float(rng.uniform(0.3, 2.0) * np.sqrt(d) / 2)
    alpha = float(10 ** rng.uniform(-8, -2))
    recipe = dict(kind="score", n=n, d=d, M=M, kernel=kind, normalize_y=norm_y, amp=amp, ls=np.asarray(ls).tolist(), alpha=alpha)
    rs = np.random.RandomState(int(rng.randint(0, 2 ** 31)))
    X = rs.uniform(size=(n, d))
    y = ((X - 0.5) ** 2).sum(axis=1) * float(10 ** rng.uniform(-2, 2)) + 0.01 * rs.randn(n) + float(rng.uniform(-10, 10))
    C = rs.uniform(size=(M, d))
    ls_vec = np.broadcast_to(np.asarray(ls, dtype=float), (d,)).copy()
    eng.fit(X, y, kind=kind, amplitude=amp, length_scale=ls, alpha=alpha, normalize_y=norm_y)
    sk = ogp.SklearnGP(kind, amp, ls_vec, alpha, norm_y).fit(X, y)
    mu_r, sd_r = sk.predict(C)
    y_opt = float(y.min())
    if rng.randint(0, 2):
        out = eng.score_host(C, acq_mask=7, y_opt=y_opt, dense=True)
        mu, sd, acq = out["mu"], out["sigma"], out["acq"]
    else:
        out = eng.score(torch.as_tensor(C).cuda(), acq_mask=7, y_opt=y_opt, dense=True)
        mu, sd, acq = out["mu"].cpu().numpy(), out["sigma"].cpu().numpy(), out["acq"].cpu().numpy()
    prior_sd = float(np.sqrt(amp) * np.ravel(sk.gpr._y_train_std)[0])
    np.testing.assert_allclose(mu, mu_r, rtol=1e-6, atol=1e-6 * max(np.abs(mu_r).max(), prior_sd))
    np.testing.assert_allclose(sd, sd_r, rtol=1e-6, atol=1e-6 * prior_sd)
    good = sd_r > 1e-3 * prior_sd  # below that sigma is cancellation noise in both implementations
    for k, name in ((0, "EI"), (1, "PI"), (2, "LCB")):
        util = oacq.utility(k, mu_r, sd_r, y_opt)
        if good.any():  # far tails (z << 0) amplify the 1e-9 posterior differences by ~z^2 (seen: 1.2e-4 on PI): 1e-3 there
            np.testing.assert_allclose(acq[k][good], util[good], rtol=1e-3, atol=1e-6 * max(np.abs(util[good]).max(), 1e-6 * prior_sd))
        got = out[name]
        assert got["index"] == int(np.argmax(acq[k])), (recipe, name)      # exact first-occurrence argmax of the device values
        assert got["value"] == acq[k][got["index"]] and got["mu"] == mu[got["index"]] and got["sigma"] == sd[got["index"]]
        if good[got["index"]] and good.all():
            assert util[got["index"]] >= util.max() - 1e-6 * max(abs(util.max()), prior_sd), (recipe, name)
    counts["score"] += 1
    # pruned == dense on the same fit
    y_opt = float(y.min())
    Cd = torch.as_tensor(C).cuda()
    dense = eng.score(Cd, acq_mask=7, y_opt=y_opt)
    for name in ("EI", "PI", "LCB"):
        one, n_scored = eng.score_pruned(Cd, acq=name, y_opt=y_opt)
        assert one[name] == dense[name], (recipe, name, one[name], dense[name])
        assert 0 <= n_scored <= M
    multi, _ = eng.score_pruned_multi(Cd, acq_mask=7, y_opt=y_opt)
    assert multi == {k: dense[k] for k in ("EI", "PI", "LCB")}, (recipe, multi, dense)
    counts["pruned"] += 1
    # posterior gradients on the same fit
    P = rs.uniform(size=(int(rng.randint(1, 9)), d))
    ref = ogp.NumpyGP(kind, amp, ls_vec, alpha, norm_y).fit(X, y)
    mu, sd, dmu, dsd = eng.predict_grad(P)
    mu0, sd0, dmu0, dsd0 = ref.predict_grad(P)
    np.testing.assert_allclose(mu, mu0, rtol=1e-6, atol=1e-6 * max(np.abs(mu0).max(), prior_sd))
    np.testing.assert_allclose(dmu, dmu0, rtol=1e-5, atol=1e-5 * max(np.abs(dmu0).max(), 1e-300))  # sum alpha_j dk_j cancels when K is ill-conditioned
    ok = sd0 > 1e-3 * np.sqrt(amp) * ref.y_std  # below that sigma^2 = amp - |v|^2 is cancellation noise
    # the gradient sums cancel (terms ~ sigma / ls each): absolute tolerance on that natural scale
    if ok.any():
        scale = max(np.abs(dsd0[ok]).max(), sd0.max() / ls_vec.min())
        np.testing.assert_allclose(dsd[ok], dsd0[ok], rtol=1e-5, atol=1e-6 * scale)
    counts["grad"] += 1
    return recipe


def fuzz_generated():
    s, spec = random_space()
    comp = compiler.compile_space(transformed(s))
    if not comp.normalisable:
        return None
    compiler.upload(eng, comp)
    d = comp.d_out
    n = int(rng.randint(1, 300))
    seed, first = int(rng.randint(0, 2 ** 31)), int(rng.choice([0, 2 ** 33 + 3, int(rng.randint(0, 2 ** 31))]))
    M = int(rng.choice([1, 33, 4737, int(rng.randint(1, 20000))]))
    recipe = dict(kind="generated", spec=spec, n=n, seed=seed, first=first, M=M)
    Xobs, _ = osamp.sample_transform(comp.slots, comp.cdf, seed + 1, 0, n, True)
    y = ((Xobs - 0.4) ** 2).sum(axis=1) + 0.01 * np.random.RandomState(seed).randn(n)
    ls = 0.5 * np.sqrt(d)
    eng.fit(Xobs, y, length_scale=ls, alpha=1e-4, normalize_y=True)
    # explicit rows come from the device sampler: log/normal priors differ from numpy's libm by an ulp (tests/test_gpu_sample.py)
    Cd = eng.sample(seed, first, M, normalise=True).clone()
    y_opt = float(y.min())
    a = eng.score(None, M=M, seed=seed, first_index=first, acq_mask=7, y_opt=y_opt)
    b = eng.score(Cd, acq_mask=7, y_opt=y_opt)
    for name in ("EI", "PI", "LCB"):
        rb = dict(b[name], index=b[name]["index"] + first)
        assert a[name] == rb, (recipe, name, a[name], rb)
        p, _ = eng.score_pruned(None, M=M, seed=seed, first_index=first, acq=name, y_opt=y_opt)
        assert p[name] == a[name], (recipe, name, p[name], a[name])
    counts["generated"] += 1
    return recipe


def fuzz_pruned():
    """The pruned path proper (M above the dense-fallback threshold): FP32 bound pass, FP64 refinement, subset-GP variance
    bound, device-resident selection.  Winners must equal the dense call bit for bit; the pass-1 bounds must be rigorous
    (mu_lb <= mu, sigma_ub >= sigma for every candidate).  alpha from 1e-10 to 1e-2, amplitudes to 1e3, length scales from
    a fraction of the cube to far beyond it (what a fitted GP on a smooth objective looks like)."""
    n = int(rng.choice([13, 192, 193, 500, int(rng.randint(2, 900))]))
    d = int(rng.choice([1, 2, 5, 8, 9, 20, 24, 25, 33, int(rng.randint(1, 70))]))
    M = int(rng.choice([4737, 4768, 9473, 18945, 40001, int(rng.randint(4737, 60000))]))   # above the single-wave dense fallback
    kind = int(rng.randint(0, 2))
    amp = float(10 ** rng.uniform(-1, 3))
    ls_scale = float(10 ** rng.uniform(-0.7, 1.3))
    ls = (rng.uniform(0.5, 1.5, size=d) * ls_scale * np.sqrt(d) / 2) if rng.rand() < 0.5 else ls_scale * np.sqrt(d) / 2
    alpha = float(10 ** rng.uniform(-10, -2))
    kappa = float(rng.choice([1.96, 0.0, 10.0, -1.0]))
    recipe = dict(kind="pruned", n=n, d=d, M=M, kernel=kind, amp=amp, ls=np.asarray(ls).tolist(), alpha=alpha, kappa=kappa)
    rs = np.random.RandomState(int(rng.randint(0, 2 ** 31)))
    X = rs.uniform(size=(n, d))
    y = ((X - 0.5) ** 2).sum(axis=1) * float(10 ** rng.uniform(-2, 2)) + 0.01 * rs.randn(n) + float(rng.uniform(-10, 10))
    try:
        eng.fit(X, y, kind=kind, amplitude=amp, length_scale=ls, alpha=alpha, normalize_y=True)
    except Exception as exc:  # not positive definite at alpha = 1e-10 with long length scales: sklearn raises too
        if "positive definite" in str(exc):
            counts["pruned_not_spd"] += 1
            return recipe
        raise
    Cd = torch.as_tensor(rs.uniform(size=(M, d))).cuda()
    y_opt = float(y.min())
    dense = eng.score(Cd, acq_mask=7, y_opt=y_opt, kappa=kappa, dense=True)
    if rng.rand() < 0.3:
        os.environ["OB200_PRUNE_SUBSET_AT"] = "0"      # force the subset-GP step
    if rng.rand() < 0.3:
        os.environ["OB200_BOUND_F32_LIMIT"] = "1e30"   # force the FP32 pass whatever its budget
    try:
        multi, _ = eng.score_pruned_multi(Cd, acq_mask=7, y_opt=y_opt, kappa=kappa)
        st = eng.prune_stats()
        assert multi == {k: dense[k] for k in ("EI", "PI", "LCB")}, (recipe, multi, {k: dense[k] for k in ("EI", "PI", "LCB")}, st)
        name = ("EI", "PI", "LCB")[int(rng.randint(0, 3))]
        one, _ = eng.score_pruned(Cd, acq=name, y_opt=y_opt, kappa=kappa)
        assert one[name] == dense[name], (recipe, name, one[name], dense[name], eng.prune_stats())
        mu_lb, sg_ub, used = eng.score_bounds(Cd)
        assert bool(torch.all(mu_lb <= dense["mu"])) and bool(torch.all(sg_ub >= dense["sigma"])), (recipe, used)
    finally:
        os.environ.pop("OB200_PRUNE_SUBSET_AT", None)
        os.environ.pop("OB200_BOUND_F32_LIMIT", None)
    counts["pruned_big"] += 1
    counts["pruned_big_f32"] += int(st["f32"])
    counts["pruned_big_subset"] += int(st["batch_b_subset"] >= 0)
    counts["pruned_big_scored"] += st["batch_a"] + max(st["batch_b_subset"], 0 if st["batch_b_subset"] >= 0 else (st["batch_b_refined"] if st["f32"] else st["batch_b"]))
    counts["pruned_big_candidates"] += M
    return recipe


def main():
    t0 = time.time()
    k = 0
    while time.time() - t0 < BUDGET:
        fn = (fuzz_sample, fuzz_score, fuzz_generated, fuzz_pruned)[k % 4]
        k += 1
        state = rng.get_state()
        try:
            fn()
        except Exception:  # noqa: BLE001
            traceback.print_exc()
            print(json.dumps(dict(failed=fn.__name__, iteration=k, counts=counts)))
            np.save("gpurun_out/fuzz_rng_state.npy", np.asarray(state, dtype=object), allow_pickle=True)
            sys.exit(1)
    print(json.dumps(dict(ok=True, seconds=round(time.time() - t0, 1), rng_seed=int(sys.argv[2]) if len(sys.argv) > 2 else 0, counts=counts)))


if __name__ == "__main__":
    main()
