"""Pins the CPU oracle (oracle/sxmc_oracle.c) before anything trusts it:
  1. against the reference's own gtest golden vectors (tests/golden_vectors.py),
  2. against the unmodified reference sources compiled into oracle/_ref (when built),
  3. against committed fixtures generated from oracle/_ref (tests/golden/*.npz) so the pin
     also holds on a box without /root/reference.
No GPU needed."""
import os

import numpy as np
import pytest

import oracle
from oracle import port
from sxmc_b200 import workloads as W
from tests import golden_vectors as G

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module", autouse=True)
def _built(oracle_built):
    return oracle_built


def _ref():
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built (no /root/reference on this box)")
    from oracle import ref
    return ref


@pytest.mark.parametrize("case", G.CASES, ids=[c[0] for c in G.CASES])
def test_port_matches_reference_gtest_vectors(case):
    name, samples, geom, evalp, syst, value, norm, expect = case
    fit = G.case_fit(samples, geom, evalp, syst, value)
    pars = np.array([1.0] + ([value] if syst else []))
    bins, norms, lut = port.eval_all(fit, pars)
    assert int(norms[0]) == norm
    assert int(bins.sum()) == norm
    G.check_pdf_values(lut[0], expect)


@pytest.mark.parametrize("case", G.CASES, ids=[c[0] for c in G.CASES])
def test_ref_matches_reference_gtest_vectors(case):
    ref = _ref()
    name, samples, geom, evalp, syst, value, norm, expect = case
    fit = G.case_fit(samples, geom, evalp, syst, value)
    pars = np.array([1.0] + ([value] if syst else []))
    bins, norms, lut = ref.eval_all(fit, pars)
    assert int(norms[0]) == norm
    G.check_pdf_values(lut[0], expect)


def test_ref_offset_stride_vector():
    ref = _ref()
    fit = G.case_fit(G.SAMPLES_1D, G.GEOM_1D, G.EVAL_1D, None, 0.0)
    h = ref.make_hist(fit, 0)
    h.set_eval_points(fit.data)
    o = G.OFFSET_STRIDE
    pdf_io = np.full(20, -55.0, np.float32)
    pdf, norm = h.eval(np.array([1.0]), pdf_io=pdf_io, pdf_offset=o["pdf_offset"],
                       pdf_stride=o["pdf_stride"],
                       norm_io=np.array(o["norm_init"], np.uint32),
                       norm_offset=o["norm_offset"])
    assert list(norm) == o["norm_expect"]
    for i in range(20):
        if i in o["pdf_expect"]:
            e = o["pdf_expect"][i]
            assert np.isnan(pdf[i]) if e is None else G.float_eq_4ulp(pdf[i], e)
        else:
            assert pdf[i] == -55.0
    h.close()


@pytest.mark.parametrize("name,kw,msg", G.CTOR_ERRORS, ids=[c[0] for c in G.CTOR_ERRORS])
def test_ref_ctor_errors(name, kw, msg):
    ref = _ref()
    args = dict(samples=G.SAMPLES_1D, nfields=1, nobs=1, lower=[0.0], upper=[1.0], nbins=[2])
    args.update(kw)
    with pytest.raises(ref.RefError) as e:
        ref.Hist(**args)
    assert str(e.value) == msg


SMALL = [("A", 0.01), ("B", 0.02), ("C", 0.002), ("D", 0.003)]


def _pars(fit, seed):
    rng = np.random.default_rng(seed)
    p = fit.parameter_means().copy()
    p[:fit.nsources] = rng.uniform(0.5, 1.5, fit.nsources)
    sig = fit.parameter_sigmas()
    p[fit.nsources:] += rng.normal(0, 1, fit.nparameters - fit.nsources) * 3 * sig[fit.nsources:]
    return p


@pytest.mark.parametrize("name,scale", SMALL, ids=[s[0] for s in SMALL])
def test_port_equals_ref_on_workloads(name, scale):
    ref = _ref()
    fit = W.WORKLOADS[name](scale)
    for seed in (1, 2):
        pars = _pars(fit, seed)
        b0, n0, l0 = port.eval_all(fit, pars)
        b1, n1, l1 = ref.eval_all(fit, pars)
        assert np.array_equal(b0, b1) and np.array_equal(n0, n1)
        assert l0.tobytes() == l1.tobytes()
        v0, e0 = port.nll(fit, l0, pars, n0)
        v1, e1 = ref.nll(fit, l1, pars, n1)
        assert v0 == v1 and e0 == e1


def test_port_chain_equals_ref_chain():
    ref = _ref()
    fit = W.config_B(0.005)
    nsteps = 300
    rng = np.random.default_rng(7)
    normals = rng.normal(size=(nsteps + 1) * fit.nparameters)
    uniforms = rng.random(nsteps)
    a = port.mcmc(fit, nsteps, 0.1, normals, uniforms)
    b = ref.mcmc(fit, nsteps, 0.1, normals, uniforms)
    assert a["chain"].shape == b["chain"].shape
    assert a["chain"].tobytes() == b["chain"].tobytes()
    # burn-in resets at i==30 and i==60 leave steps [60, 300)
    assert a["chain"].shape[0] == nsteps - 60


def test_port_matches_committed_ref_fixtures():
    path = os.path.join(HERE, "golden", "ref_fixtures.npz")
    if not os.path.exists(path):
        pytest.skip("fixtures not generated")
    z = np.load(path)
    for name, scale in SMALL:
        fit = W.WORKLOADS[name](scale)
        pars = z[name + "_pars"]
        b, n, l = port.eval_all(fit, pars)
        assert np.array_equal(b, z[name + "_bins"])
        assert np.array_equal(n, z[name + "_norms"])
        assert l.tobytes() == z[name + "_lut"].tobytes()
        v, e = port.nll(fit, l, pars, n)
        assert v == float(z[name + "_nll"]) and e == float(z[name + "_evsum"])


@pytest.mark.parametrize("name", ["CT", "P2", "P3", "P4"])
def test_port_equals_ref_on_ctscale_and_polynomial_systematics(name):
    """CTSCALE and 2-, 3-, 4-coefficient series (src/pdfz.cpp:310-325): the restatement against
    the compiled reference, live and against the committed reference output."""
    from tests import cases
    fit = cases.POLY_CASES[name]()
    z = np.load(os.path.join(HERE, "golden", "ref_fixtures.npz"))
    pars = cases.off_nominal(fit, 31)
    assert np.array_equal(pars, z[name + "_pars"])
    b0, n0, l0 = port.eval_all(fit, pars)
    assert np.array_equal(b0, z[name + "_bins"]) and np.array_equal(n0, z[name + "_norms"])
    assert l0.tobytes() == z[name + "_lut"].tobytes()
    v0, e0 = port.nll(fit, l0, pars, n0)
    assert v0 == float(z[name + "_nll"]) and e0 == float(z[name + "_evsum"])
    assert int(n0.sum()) > 0 and int(n0.sum()) < fit.total_samples()      # the arms move samples across the edges
    if oracle.have_ref():
        from oracle import ref
        for seed in (31, 32):
            pars = cases.off_nominal(fit, seed)
            b0, n0, l0 = port.eval_all(fit, pars)
            b1, n1, l1 = ref.eval_all(fit, pars)
            assert np.array_equal(b0, b1) and np.array_equal(n0, n1) and l0.tobytes() == l1.tobytes()


def test_projection_estimator_walk_on_known_shapes():
    """sxmc::errors::Projection (src/error_estimators/projection.cpp:14-77), host restatement:
    a Gaussian column gives a central interval with about cl coverage around the mean; a column
    piled up at zero (a rate with no signal) gives a one-sided limit at the cl quantile."""
    from sxmc_b200 import ensemble
    rng = np.random.default_rng(3)
    g = rng.normal(5.0, 0.5, 200000).astype(np.float32)
    iv = ensemble.projection_interval(g, 0.683)
    assert not iv.one_sided and abs(iv.point_estimate - 5.0) < 0.01
    assert 0.683 <= iv.coverage < 0.75                      # whole bins: at least cl
    assert abs(iv.lower - 4.5) < 0.06 and abs(iv.upper - 5.5) < 0.06
    e = rng.exponential(1.0, 200000).astype(np.float32)     # mode at the lower edge
    iv = ensemble.projection_interval(e, 0.9)
    assert iv.one_sided and abs(iv.upper - np.quantile(e, 0.9)) < 0.15 and iv.coverage >= 0.9
    lo, hi, nb = ensemble._root_axis_limits(0.0, 9.7, 100)
    assert (lo, nb) == (0.0, 98) and abs(hi - 9.8) < 1e-12  # width 0.097 -> 0.1
    out = ensemble.projection_intervals(np.stack([g, e, g], axis=1), ["a", "b", "likelihood"], 0.683)
    assert set(out) == {"a", "b"}
