"""Ensemble pieces (SURVEY 8(f)-1,2): the device fake-data generator and the Contour
estimator.  Generator output depends on the RNG (ROOT's in the reference, Philox here), so it
is checked through what the algorithm guarantees: exact counts without Poisson, Poisson moments
with it, events distributed as the sampled histogram, limits respected, reproducibility per
(seed, experiment).  The estimator is compared exactly with the numpy restatement."""
import numpy as np
import pytest

from oracle import port
from sxmc_b200 import workloads as W

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _env(oracle_built):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from sxmc_b200 import _lib
    _lib.check(_lib.lib().sxmc_init(0))


def _generator(fit):
    from sxmc_b200 import ensemble, mcmc
    hists, chain = mcmc.build_fit(fit)
    signals = [mcmc.Signal("sig%d" % j, hists[j], s.nexpected, s.n_mc, None) for j, s in enumerate(fit.signals)]
    systs = [mcmc.Systematic("y%d" % i, y.npars, list(y.means), list(y.sigmas), y.fixed)
             for i, y in enumerate(fit.systematics)]
    obs = [mcmc.Observable("o%d" % i, fit.lower[i], fit.upper[i]) for i in range(fit.nobs)]
    return ensemble.FakeDataGenerator(signals, systs, obs), hists, chain


@pytest.mark.parametrize("name,scale", [("B", 0.05), ("C", 0.002), ("D", 0.003)])
def test_generator_counts_efficiency_and_distribution(name, scale):
    fit = W.WORKLOADS[name](scale)
    gen, hists, chain = _generator(fit)
    bins, norms, _ = port.eval_all(fit, fit.parameter_means())
    eff = norms / np.asarray([s.n_mc for s in fit.signals], dtype=np.float64)
    assert np.array_equal(gen.efficiency(), eff)                       # get_efficiency, exact
    nexp = np.asarray([s.nexpected for s in fit.signals]) * eff

    # poisson = false: observed = nint(float(nexpected * eff))
    ev, obs = gen(False, 7, 0)
    want = np.rint(nexp.astype(np.float32)).astype(np.uint32)
    assert np.array_equal(obs, want)
    assert ev.shape == (int(want.sum()), fit.nobs + 1)
    assert np.all(ev[:, -1] == 0)                                      # dataset id column
    for o in range(fit.nobs):                                          # limits respected
        assert np.all(ev[:, o] >= np.float32(fit.lower[o])) and np.all(ev[:, o] <= np.float32(fit.upper[o]))

    # reproducible per (seed, experiment); different experiments differ
    ev2, obs2 = gen(False, 7, 0)
    assert np.array_equal(ev, ev2)
    ev3, _ = gen(False, 7, 1)
    assert not np.array_equal(ev, ev3)

    # events of signal 0 follow its histogram: chi2 over coarse groups of bins
    n0 = int(obs[0])
    e0 = ev[:n0, :fit.nobs].astype(np.float64)
    idx = np.zeros(n0, dtype=np.int64)
    stride = 1
    for o in reversed(range(fit.nobs)):
        k = np.floor((e0[:, o] - fit.lower[o]) / (fit.upper[o] - fit.lower[o]) * fit.nbins[o]).astype(np.int64)
        idx += np.clip(k, 0, fit.nbins[o] - 1) * stride
        stride *= fit.nbins[o]
    got = np.bincount(idx, minlength=fit.total_nbins).astype(np.float64)
    p = bins[0] / bins[0].sum()
    assert np.all(got[p == 0] == 0)                                    # never an empty bin
    order = np.argsort(-p)
    groups = np.array_split(order, 20)
    chi2 = sum((got[g].sum() - n0 * p[g].sum()) ** 2 / max(n0 * p[g].sum(), 1e-9) for g in groups
               if n0 * p[g].sum() > 5)
    assert chi2 < 60, chi2                                             # ~19 dof; P(>60) ~ 4e-6

    # poisson = true: mean and variance of the counts over many experiments
    draws = np.stack([gen(True, 11, i)[1] for i in range(400)]).astype(np.float64)
    m, v = draws.mean(axis=0), draws.var(axis=0)
    assert np.all(np.abs(m - nexp) < 5 * np.sqrt(nexp / 400) + 0.05), (m, nexp)
    assert np.all(np.abs(v / nexp - 1) < 0.35), (v, nexp)
    gen.close()
    chain.close()
    for h in hists:
        h.close()


def test_poisson_small_means_and_zero():
    fit = W.config_B(0.01)
    for s, ne in zip(fit.signals, (0.0, 0.7, 4.2)):
        s.nexpected = ne
    gen, hists, chain = _generator(fit)
    eff = gen.efficiency()
    draws = np.stack([gen(True, 3, i)[1] for i in range(3000)]).astype(np.float64)
    assert np.all(draws[:, 0] == 0)
    for j in (1, 2):
        mu = fit.signals[j].nexpected * eff[j]
        assert abs(draws[:, j].mean() - mu) < 5 * np.sqrt(mu / 3000)
        assert abs(draws[:, j].var() / mu - 1) < 0.15
    gen.close()
    chain.close()
    for h in hists:
        h.close()


def test_contour_matches_restatement_on_a_real_chain_and_on_edge_cases():
    from sxmc_b200 import ensemble, mcmc
    fit = W.config_B(0.02)
    hists, chain = mcmc.build_fit(fit)
    chain.set_seed(5, 0)
    ls = chain(fit.data, 3000, 0.1)
    rows = ls.get_samples()
    for cl in (0.9, 0.68, 0.99):
        delta = ensemble.contour_delta(cl)
        want = port.contour(rows, delta)
        iv, lmin = ensemble.contour_intervals(chain, cl)
        names = chain.parameter_names[:-1]
        assert np.array_equal(np.float32([iv[n].lower for n in names]), want[0])
        assert np.array_equal(np.float32([iv[n].upper for n in names]), want[1])
        assert np.array_equal(np.float32([iv[n].point_estimate for n in names]), want[2])
        assert np.float32(lmin) == np.float32(want[3])
        got = ensemble.rows_contour(rows, cl)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
        assert np.array_equal(got[2], want[2]) and got[4] == want[4]
    # synthetic rows: negative and zero parameters, huge |NLL| (6-digit rounding of -lmin
    # leaves NO row within 0.13 of the printed minimum: the threshold must widen x5)
    rng = np.random.default_rng(3)
    r = rng.normal(size=(5000, 4)).astype(np.float32)
    r[:, 1] = 0.0
    r[:, 3] = (-123456.8 + 3.0 * rng.random(5000)).astype(np.float32)   # printed offset: 123457
    got = ensemble.rows_contour(r, 0.9)
    want = port.contour(r, ensemble.contour_delta(0.9))
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    assert np.array_equal(got[2], want[2]) and got[4] == want[4]
    chain.close()
    for h in hists:
        h.close()


def test_ensemble_limits_are_reproducible_and_cover_the_truth():
    """A small ensemble: upper limits on source 0 (true rate 1.0 x nexpected) at 90 % CL; the
    batched launch, the one-chain-at-a-time loop and a rank of a 2-rank run agree exactly."""
    from sxmc_b200 import ensemble
    fit = W.config_B(0.02)
    a = ensemble.ensemble(fit, 12, "s0", 0.9, seed=42, nsteps=4000, burnin_fraction=0.1)
    s = ensemble.ensemble(fit, 12, "s0", 0.9, seed=42, nsteps=4000, burnin_fraction=0.1, batch=0)
    b = ensemble.ensemble(fit, 12, "s0", 0.9, seed=42, nsteps=4000, burnin_fraction=0.1, rank=1, world=2)
    assert a["mode"] == "batch" and s["mode"] == "single" and a["launches"] == 1
    assert a["experiments"] == list(range(12)) and b["experiments"] == list(range(1, 12, 2))
    lim = np.asarray(a["limits"])
    assert np.all(np.isfinite(lim))
    assert np.array_equal(lim, np.asarray(s["limits"]))
    assert np.array_equal(np.stack(a["observed"]), np.stack(s["observed"]))
    assert np.array_equal(lim[1::2], np.asarray(b["limits"]))
    assert np.mean(lim > 1.0) >= 0.7          # upper edge of a 90 % contour: mostly above truth
    assert np.all(lim < 2.0)


# ---------------------------------------------------------------------------------------
# chain batch: many chains over the same samples in one cooperative launch
# ---------------------------------------------------------------------------------------
def _with_data(fit, data):
    import copy
    f = copy.copy(fit)
    f.data = np.ascontiguousarray(data, dtype=np.float32)
    return f


@pytest.fixture(params=[False, True], ids=["mono", "nomono"])
def bin_path(request):
    """Both binning paths of the batch kernel: monotone endpoint evaluation (1-D shift/scale
    fits) and the sample-by-sample pass."""
    from sxmc_b200 import pdfz
    pdfz.set_bin_tuning(8192, 0, False, False, request.param)
    yield request.param
    pdfz.set_bin_tuning(8192, 0, False, False, False)


@pytest.mark.parametrize("name,scale,nchains", [("B", 0.005, 5), ("A", 0.004, 3), ("B", 0.0002, 3)])
def test_batch_tape_driven_chains_match_oracle_chains(name, scale, nchains, bin_path):
    """Every chain of a batch, fed its own tape and its own data events, reproduces the oracle
    chain of the same inputs: proposals and accepted rows bit-identical, NLL within 1e-6."""
    from sxmc_b200 import mcmc
    fit = W.WORKLOADS[name](scale)
    nsteps = 300
    hists, batch = mcmc.build_fit(fit, batch=nchains)
    batch.enable_trace()
    wants = []
    for k in range(nchains):
        rng = np.random.default_rng(100 + k)
        normals = rng.normal(size=(nsteps + 1) * fit.nparameters)
        uniforms = rng.random(nsteps)
        data = fit.data[k::nchains] if k else fit.data[:0]          # slot 0: no events at all
        wants.append(port.mcmc(_with_data(fit, data), nsteps, 0.1, normals, uniforms, trace=True))
        batch.set_tapes(k, normals, uniforms)
        batch.set_data(k, data)
    batch.run(nchains, nsteps, 0.1)
    for k in range(nchains):
        want = wants[k]
        got = batch.rows(k)
        nll_tr, prop_tr = batch.read_trace(k, nsteps)
        assert got.shape == want["chain"].shape
        assert batch.nrows(k)[1] == want["accepted"]
        assert np.array_equal(prop_tr, want["prop_trace"])
        assert np.allclose(nll_tr, want["nll_trace"], rtol=1e-6, atol=0)
        assert np.array_equal(got[:, :-1], want["chain"][:, :-1])
        assert np.allclose(got[:, -1], want["chain"][:, -1], rtol=1e-6)
    batch.close()
    for h in hists:
        h.close()


def _scale_ctscale_fit(special, reverse=False):
    """1-D fit with scale + ctscale + shift on the observable (a monotone program with a
    decreasing piece when 1+p < 0); `special` adds +-inf / NaN / -0 samples."""
    from sxmc_b200.workloads import CTSCALE, SCALE, SHIFT, FitSpec, SignalSpec, SystSpec
    rng = np.random.default_rng(77)
    sigs = []
    for j, mu in enumerate((-0.3, 0.4)):
        x = rng.normal(mu, 0.5, 40_000).astype(np.float32)
        if special:
            x[:6] = [np.inf, -0.0, 0.0, np.inf, 1.0, -1.0]
            if special == "nan":
                x[6:9] = [np.nan, -np.inf, np.nan]
        rows = np.stack([x, np.zeros_like(x)], axis=1)
        sigs.append(SignalSpec(samples=rows, nexpected=300.0, source=j))
    ev = np.clip(rng.normal(0, 0.5, 600), -0.999, 0.999)
    data = np.stack([ev, np.zeros_like(ev)], axis=1).astype(np.float32)
    systs = [SystSpec(type=SCALE, obs=0, means=[-2.0 if reverse else 0.0], sigmas=[0.3]),   # 1+p < 0: order reversed
             SystSpec(type=CTSCALE, obs=0, means=[0.0], sigmas=[0.05]),
             SystSpec(type=SHIFT, obs=0, means=[0.0], sigmas=[0.02])]
    return FitSpec(name="mono", lower=[-1.0], upper=[1.0], nbins=[200], source_means=[1.0, 1.0],
                   source_sigmas=[0.0, 0.0], source_fixed=[False, False], systematics=systs,
                   signals=sigs, data=data, nsteps=100)


@pytest.mark.parametrize("special,reverse", [(None, False), ("inf", False), ("inf", True), ("nan", True)])
def test_batch_monotone_program_with_negative_scale_and_special_values(special, reverse, bin_path):
    """Negative scale factors reverse the order of the transformed samples; +inf, -0 and (with
    "nan": NaN / -inf, which switch the endpoint path off) must be binned as the oracle does."""
    from sxmc_b200 import mcmc
    fit = _scale_ctscale_fit(special, reverse)
    nsteps = 250
    hists, batch = mcmc.build_fit(fit, batch=2)
    batch.enable_trace()
    wants = []
    for k in range(2):
        rng = np.random.default_rng(500 + k)
        normals = rng.normal(size=(nsteps + 1) * fit.nparameters) * (3.0 if k else 1.0)
        uniforms = rng.random(nsteps)
        wants.append(port.mcmc(fit, nsteps, 0.1, normals, uniforms, trace=True))
        batch.set_tapes(k, normals, uniforms)
        batch.set_data(k, fit.data)
    batch.run(2, nsteps, 0.1)
    for k in range(2):
        nll_tr, prop_tr = batch.read_trace(k, nsteps)
        assert np.array_equal(prop_tr, wants[k]["prop_trace"])
        assert np.all(1.0 + prop_tr[:, fit.nsources] < 0) == reverse
        ok = np.isfinite(wants[k]["nll_trace"])
        assert np.allclose(nll_tr[ok], wants[k]["nll_trace"][ok], rtol=1e-6, atol=0)
        assert np.array_equal(batch.rows(k)[:, :-1], wants[k]["chain"][:, :-1])
    batch.close()
    for h in hists:
        h.close()


def test_batch_fixed_systematics_and_debug_mode():
    from sxmc_b200 import mcmc
    fit = W.config_A(0.004)
    for y in fit.systematics:
        y.fixed = True
    nsteps = 200
    for debug in (False, True):
        hists, batch = mcmc.build_fit(fit, batch=2)
        wants = []
        for k in range(2):
            rng = np.random.default_rng(7 + k)
            normals = rng.normal(size=(nsteps + 1) * fit.nparameters)
            uniforms = rng.random(nsteps)
            wants.append(port.mcmc(fit, nsteps, 0.1, normals, uniforms, debug_mode=debug))
            batch.set_tapes(k, normals, uniforms)
            batch.set_data(k, fit.data)
        batch.run(2, nsteps, 0.1, debug)
        for k in range(2):
            got = batch.rows(k)
            assert got.shape == wants[k]["chain"].shape
            assert np.array_equal(got[:, :-1], wants[k]["chain"][:, :-1])
        batch.close()
        for h in hists:
            h.close()


def test_batch_philox_chain_equals_single_chain_and_full_width_batch():
    """148 chains at once; chain k of the batch = the one-chain path with the same (seed, stream)."""
    from sxmc_b200 import mcmc
    fit = W.config_B(0.01)
    nsteps = 500
    hists, batch = mcmc.build_fit(fit, batch=148)
    for k in range(148):
        batch.set_seed(k, 77, k)
        batch.set_data(k, fit.data[(k % 7)::7])
    batch.run(148, nsteps, 0.1)
    h2, chain = mcmc.build_fit(fit)
    chain.set_mode(2)            # the persistent one-chain kernel: a different binning path
    for k in (0, 1, 73, 147):
        chain.set_seed(77, k)
        want = chain(fit.data[(k % 7)::7], nsteps, 0.1).get_samples()
        got = batch.rows(k)
        assert got.shape == want.shape and got.shape[0] == nsteps - 2 * 50
        assert np.array_equal(got[:, :-1], want[:, :-1])
        assert np.allclose(got[:, -1], want[:, -1], rtol=1e-6)
        assert batch.nrows(k)[1] == chain.accepted()
    # different streams really are different chains
    assert not np.array_equal(batch.rows(3), batch.rows(10))
    chain.close()
    batch.close()
    for h in hists + h2:
        h.close()


def test_batch_rejects_fits_whose_samples_do_not_fit_on_chip():
    from sxmc_b200 import mcmc
    from sxmc_b200._lib import SxmcError
    fit = W.config_C(0.05)
    hists, batch = mcmc.build_fit(fit, batch=4)
    batch.set_data(0, fit.data)
    with pytest.raises(SxmcError):
        batch.run(1, 10, 0.1)
    batch.close()
    for h in hists:
        h.close()
