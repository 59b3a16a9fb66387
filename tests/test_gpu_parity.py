"""Parity tests proper: the CUDA path (through the C ABI) against the CPU oracle, the
reference's gtest golden vectors and the committed reference fixtures.  Integer work
(bins, norms, read_bins) and the float lut are compared bit-exactly; NLL within 1e-6
relative (the tolerance BASELINE.json's north_star states)."""
import os

import numpy as np
import pytest

from oracle import port
from sxmc_b200 import workloads as W
from tests import golden_vectors as G

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
NLL_RTOL = 1e-6   # north_star: 1e-6 on NLL per step


@pytest.fixture(scope="module", autouse=True)
def _env(oracle_built):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from sxmc_b200 import _lib
    _lib.check(_lib.lib().sxmc_init(0))
    yield
    from sxmc_b200 import pdfz
    pdfz.set_bin_tuning(8192, 0, False)


def _pars(fit, seed):
    rng = np.random.default_rng(seed)
    p = fit.parameter_means().copy()
    p[:fit.nsources] = rng.uniform(0.5, 1.5, fit.nsources)
    sig = fit.parameter_sigmas()
    p[fit.nsources:] += rng.normal(0, 1, fit.nparameters - fit.nsources) * 3 * sig[fit.nsources:]
    return p


def _gpu_eval_api(fit, pars):
    """Through the pdfz::EvalHist mirror, one PDF at a time (EvalAsync/EvalFinished)."""
    from sxmc_b200 import pdfz
    ns, ne = fit.nsignals, fit.nevents
    params = pdfz.Array.from_host(pars, np.float64)
    norms = pdfz.Array(ns, np.uint32)
    lut = pdfz.Array(ns * ne, np.float32)
    bins = []
    for j, sig in enumerate(fit.signals):
        h = pdfz.EvalHist(sig.samples, sig.samples.shape[1], fit.nobs, fit.lower, fit.upper,
                          fit.nbins, sig.dataset)
        for i in sig.systematics:
            y = fit.systematics[i]
            cls = {pdfz.SHIFT: pdfz.ShiftSystematic, pdfz.SCALE: pdfz.ScaleSystematic,
                   pdfz.CTSCALE: pdfz.CosThetaScaleSystematic}.get(y.type)
            pars_arr = pdfz.Array.from_host(y.pidx, np.int16)
            if cls is None:
                h.AddSystematic(pdfz.ResolutionScaleSystematic(y.obs, y.extra, pars_arr))
            else:
                h.AddSystematic(cls(y.obs, pars_arr))
        h.SetEvalPoints(fit.data)
        h.SetPDFValueBuffer(lut, j * ne, 1)
        h.SetNormalizationBuffer(norms, j)
        h.SetParameterBuffer(params, fit.nsources)
        h.EvalAsync()
        h.EvalFinished()
        bins.append(h.Bins())
        h.close()
    return (np.stack(bins), norms.readOnlyHostPtr().copy(),
            lut.readOnlyHostPtr().reshape(ns, ne).copy())


# ---------------------------------------------------------------------------------------
# the reference's own gtest vectors (test/test_pdfz*.cpp), through the EvalHist mirror
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", G.CASES, ids=[c[0] for c in G.CASES])
def test_reference_gtest_vectors(case):
    name, samples, geom, evalp, syst, value, norm, expect = case
    fit = G.case_fit(samples, geom, evalp, syst, value)
    pars = np.array([1.0] + ([value] if syst else []))
    bins, norms, lut = _gpu_eval_api(fit, pars)
    assert int(norms[0]) == norm
    assert int(bins.sum()) == norm
    G.check_pdf_values(lut[0], expect)


def test_offset_stride_vector():
    """test/test_pdfz.cpp:98-126: sentinel values around the written slots survive."""
    from sxmc_b200 import pdfz
    o = G.OFFSET_STRIDE
    h = pdfz.EvalHist(G.SAMPLES_1D, 1, 1, [0.0], [1.0], [2])
    h.SetEvalPoints(G.EVAL_1D)
    pdf = pdfz.Array.from_host(np.full(20, -55.0, np.float32), np.float32)
    norm = pdfz.Array.from_host(o["norm_init"], np.uint32)
    params = pdfz.Array.from_host([1.0], np.float64)
    h.SetPDFValueBuffer(pdf, o["pdf_offset"], o["pdf_stride"])
    h.SetNormalizationBuffer(norm, o["norm_offset"])
    h.SetParameterBuffer(params)
    h.EvalAsync()
    h.EvalFinished()
    assert list(norm.readOnlyHostPtr()) == o["norm_expect"]
    got = pdf.readOnlyHostPtr()
    for i in range(20):
        if i in o["pdf_expect"]:
            e = o["pdf_expect"][i]
            assert np.isnan(got[i]) if e is None else G.float_eq_4ulp(got[i], e)
        else:
            assert got[i] == -55.0
    h.close()


def test_eval_without_points_and_get_samples():
    from sxmc_b200 import pdfz
    fit = W.config_C(0.0005)
    sig = fit.signals[3]
    h = pdfz.EvalHist(sig.samples, sig.samples.shape[1], fit.nobs, fit.lower, fit.upper, fit.nbins)
    h.AddSystematic(pdfz.ScaleSystematic(0, [0]))
    norm = pdfz.Array(1, np.uint32)
    params = pdfz.Array.from_host([0.02], np.float64)
    h.SetNormalizationBuffer(norm, 0)
    h.SetParameterBuffer(params, 0)
    h.EvalAsync(False)       # src/pdfz.cpp:474: no eval points, histogram only
    h.EvalFinished()
    want_bins, want_norm = port.bin_signal(W.FitSpec(
        name="x", lower=fit.lower, upper=fit.upper, nbins=fit.nbins, source_means=[1.0],
        source_sigmas=[0.0], source_fixed=[False],
        systematics=[W.SystSpec(type=W.SCALE, obs=0)],
        signals=[W.SignalSpec(samples=sig.samples, nexpected=1.0, source=0)],
        data=np.zeros((0, fit.nobs + 1), np.float32)), 0, [0.02])
    assert np.array_equal(h.Bins(), want_bins) and int(norm.readOnlyHostPtr()[0]) == want_norm
    # GetSamples returns the caller's rows in the caller's order although the device
    # layout is bin-sorted and tiled
    assert h.GetSamples().tobytes() == sig.samples.tobytes()
    h.close()


# ---------------------------------------------------------------------------------------
# workloads A-D (scaled), CUDA vs oracle, several kernel configurations
# ---------------------------------------------------------------------------------------
SMALL = [("A", 0.01), ("B", 0.02), ("C", 0.002), ("D", 0.003)]
VARIANTS = [("smem", dict(smem_hist_max_bins=56 * 1024)), ("global", dict(smem_hist_max_bins=0)),
            ("dynamic", dict(force_dynamic=True)), ("dynamic_global", dict(force_dynamic=True, smem_hist_max_bins=0)),
            ("generic", dict(force_generic=True)), ("generic_global", dict(force_generic=True, smem_hist_max_bins=0)),
            # the measurement forms of the fast kernel's partition (kernel_select bits 8-9)
            ("contiguous", dict(cta_split=1)), ("contiguous_global", dict(cta_split=1, smem_hist_max_bins=0)),
            ("warp_ranges_global", dict(cta_split=2, smem_hist_max_bins=0)),
            ("segments_global", dict(cta_split=3, smem_hist_max_bins=0)),
            # four-row chunks (what launches of >= 384 MB take) forced onto the small cases, and the
            # two-row form
            ("four_row_chunks", dict(chunk_form=2)), ("four_row_chunks_global", dict(chunk_form=2, smem_hist_max_bins=0)),
            ("two_row_chunks", dict(chunk_form=1))]


@pytest.mark.parametrize("variant", VARIANTS, ids=[v[0] for v in VARIANTS])
@pytest.mark.parametrize("name,scale", SMALL, ids=[s[0] for s in SMALL])
def test_chain_eval_matches_oracle(name, scale, variant):
    from sxmc_b200 import mcmc, pdfz
    pdfz.set_bin_tuning(**dict(dict(smem_hist_max_bins=8192, blocks_per_sm=0, force_generic=False, force_dynamic=False),
                               **variant[1]))
    fit = W.WORKLOADS[name](scale)
    hists, chain = mcmc.build_fit(fit)
    chain.set_data(fit.data)
    for seed in (1, 2, 3):
        pars = _pars(fit, seed)
        nll, ev = chain.nll(pars)
        bins, norms, lut = port.eval_all(fit, pars)
        want, want_ev = port.nll(fit, lut, pars, norms)
        assert np.array_equal(chain.bins(), bins)
        assert np.array_equal(chain.norms(), norms)
        assert chain.lut().tobytes() == lut.tobytes()
        assert abs(ev - want_ev) <= NLL_RTOL * abs(want_ev)
        assert abs(nll - want) <= NLL_RTOL * abs(want)
    pdfz.set_bin_tuning(8192, 0, False)


def test_two_fits_in_one_process_keep_their_own_tuning():
    """The kernel-selection knobs are defaults that a handle copies when it is created: three chains
    over the same fit, created under three settings, keep three different kernels side by side (and
    agree on every histogram)."""
    from sxmc_b200 import mcmc, pdfz
    fit = W.config_C(0.003)
    chains = []
    for kw in (dict(), dict(force_dynamic=True), dict(force_generic=True)):
        pdfz.set_bin_tuning(**kw)
        chains.append(mcmc.build_fit(fit))
    pdfz.set_bin_tuning()
    pars = _pars(fit, 4)
    outs = []
    for hists, chain in chains:              # used only now, after the defaults went back to normal
        chain.set_data(fit.data)
        chain.nll(pars)
        outs.append((chain.bin_kernel_kind(), chain.bins(), chain.norms()))
    assert [o[0] for o in outs] == [2, 1, 0]
    assert all(np.array_equal(o[1], outs[0][1]) and np.array_equal(o[2], outs[0][2]) for o in outs)
    for hists, chain in chains:
        chain.close()
        for h in hists:
            h.close()


def test_unsorted_layout_gives_identical_histograms():
    from sxmc_b200 import mcmc
    fit = W.config_C(0.003)
    pars = _pars(fit, 5)
    out = []
    for sort in (True, False):
        hists, chain = mcmc.build_fit(fit, sort=sort)
        chain.set_data(fit.data)
        chain.nll(pars)
        out.append((chain.bins(), chain.norms()))
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])


def test_api_path_matches_oracle_and_read_bins():
    from sxmc_b200 import pdfz
    fit = W.config_D(0.002)
    pars = _pars(fit, 9)
    bins, norms, lut = _gpu_eval_api(fit, pars)
    b0, n0, l0 = port.eval_all(fit, pars)
    assert np.array_equal(bins, b0) and np.array_equal(norms, n0) and lut.tobytes() == l0.tobytes()
    sig = fit.signals[0]
    h = pdfz.EvalHist(sig.samples, sig.samples.shape[1], fit.nobs, fit.lower, fit.upper, fit.nbins, 0)
    # edge rows: outside the domain, other dataset, exactly on the bounds
    extra = np.array([[4.99, 1.0, 0.0, 0], [15.0, 1.0, 0.0, 0], [5.0, 0.0, -1.0, 0],
                      [7.0, 3.0, 0.5, 1], [20.0, 3.0, 0.5, 1]], dtype=np.float32)
    pts = np.concatenate([fit.data[:1000], extra])
    h.SetEvalPoints(pts)
    assert np.array_equal(h.ReadBins(), port.read_bins(fit, 0, pts))
    h.close()


def test_committed_reference_fixtures():
    """Against outputs of the UNMODIFIED reference (tests/golden/make_ref_fixtures.py)."""
    from sxmc_b200 import mcmc
    z = np.load(os.path.join(HERE, "golden", "ref_fixtures.npz"))
    for name, scale in SMALL:
        fit = W.WORKLOADS[name](scale)
        hists, chain = mcmc.build_fit(fit)
        chain.set_data(fit.data)
        pars = z[name + "_pars"]
        nll, ev = chain.nll(pars)
        assert np.array_equal(chain.bins(), z[name + "_bins"])
        assert np.array_equal(chain.norms(), z[name + "_norms"])
        assert chain.lut().tobytes() == z[name + "_lut"].tobytes()
        assert abs(nll - float(z[name + "_nll"])) <= NLL_RTOL * abs(float(z[name + "_nll"]))
        assert abs(ev - float(z[name + "_evsum"])) <= NLL_RTOL * abs(float(z[name + "_evsum"]))


@pytest.mark.parametrize("name,scale", SMALL + [("nan", 0.0)], ids=[s[0] for s in SMALL] + ["nan_and_guards"])
def test_standalone_nll_of_a_materialised_lut(name, scale):
    """sxmc_nll_eval = MCMC::nll (src/mcmc.cpp:390-415) over a caller-held lut: against the value the
    compiled reference's MCMC::nll gave for the same lut (fixture), the oracle, and the chain's own
    fused evaluation.  Extra case: NaN lut entries count as 0 (nll_kernels.cpp:107), events whose
    sum is not positive are skipped (:109), a negative source rate returns 1e18 (:162-165)."""
    from sxmc_b200 import mcmc, pdfz
    if name == "nan":
        fit = W.config_B(0.002)
        hists, chain = mcmc.build_fit(fit)
        rng = np.random.default_rng(4)
        lut = rng.random((fit.nsignals, fit.nevents)).astype(np.float32)
        lut[:, ::7] = np.nan            # whole events without a PDF value
        lut[1, ::3] = np.nan
        lut[:, 5] = 0.0
        norms = np.asarray([s.n_mc // 2 for s in fit.signals], np.uint32)
        pars = fit.parameter_means().copy()
        want, want_ev = port.nll(fit, lut, pars, norms)
        got, got_ev = chain.nll_lut(pdfz.Array.from_host(lut.ravel(), np.float32),
                                    pdfz.Array.from_host(pars, np.float64), pdfz.Array.from_host(norms, np.uint32))
        assert abs(got - want) <= NLL_RTOL * abs(want) and abs(got_ev - want_ev) <= NLL_RTOL * abs(want_ev)
        pars[0] = -0.5
        want, _ = port.nll(fit, lut, pars, norms)
        got, _ = chain.nll_lut(pdfz.Array.from_host(lut.ravel(), np.float32),
                               pdfz.Array.from_host(pars, np.float64), pdfz.Array.from_host(norms, np.uint32))
        assert want == 1e18 and got == 1e18
        chain.close()
        return
    z = np.load(os.path.join(HERE, "golden", "ref_fixtures.npz"))
    fit = W.WORKLOADS[name](scale)
    hists, chain = mcmc.build_fit(fit)
    chain.set_data(fit.data)
    pars = z[name + "_pars"]
    fused, fused_ev = chain.nll(pars)
    lut = pdfz.Array.from_host(z[name + "_lut"].ravel(), np.float32)
    got, got_ev = chain.nll_lut(lut, pdfz.Array.from_host(pars, np.float64),
                                pdfz.Array.from_host(z[name + "_norms"], np.uint32))
    ref_nll, ref_ev = float(z[name + "_nll"]), float(z[name + "_evsum"])
    assert abs(got - ref_nll) <= NLL_RTOL * abs(ref_nll) and abs(got_ev - ref_ev) <= NLL_RTOL * abs(ref_ev)
    assert abs(got - fused) <= NLL_RTOL * abs(fused) and abs(got_ev - fused_ev) <= NLL_RTOL * abs(fused_ev)
    chain.close()
    for h in hists:
        h.close()


def test_polynomial_and_ctscale_systematics():
    """Multi-coefficient series (p = sum p_i x^i, src/pdfz.cpp:310-314) and ctscale."""
    from sxmc_b200 import mcmc
    base = W.config_D(0.002)
    systs = [W.SystSpec(type=W.SHIFT, obs=0, means=[0.01, 0.002], sigmas=[0.01, 0.001]),
             W.SystSpec(type=W.SCALE, obs=0, means=[0.0, 0.001, 0.0001], sigmas=[0.01, 0.001, 0.0001]),
             W.SystSpec(type=W.CTSCALE, obs=2, means=[0.01], sigmas=[0.01]),
             W.SystSpec(type=W.RESOLUTION_SCALE, obs=0, extra=3, means=[0.02], sigmas=[0.01])]
    fit = W.FitSpec(name="poly", lower=base.lower, upper=base.upper, nbins=base.nbins,
                    source_means=base.source_means[:4], source_sigmas=base.source_sigmas[:4],
                    source_fixed=base.source_fixed[:4], systematics=systs,
                    signals=[W.SignalSpec(samples=s.samples, nexpected=s.nexpected, source=j)
                             for j, s in enumerate(base.signals[:4])], data=base.data[:5000])
    hists, chain = mcmc.build_fit(fit)
    chain.set_data(fit.data)
    pars = fit.parameter_means().copy()
    nll, ev = chain.nll(pars)
    bins, norms, lut = port.eval_all(fit, pars)
    want, _ = port.nll(fit, lut, pars, norms)
    assert np.array_equal(chain.bins(), bins) and np.array_equal(chain.norms(), norms)
    assert abs(nll - want) <= NLL_RTOL * abs(want)


@pytest.mark.parametrize("name", ["CT", "P2", "P3", "P4"])
def test_ctscale_and_polynomial_systematics_against_reference_output(name):
    """CUDA against what the UNMODIFIED reference produced (tests/golden/ref_fixtures.npz) for
    CTSCALE and 2-, 3-, 4-coefficient series.  CT, P2, P3: bit-exact.  P4 evaluates pow(x, 3):
    glibc's pow is within 1 ulp of, but not always equal to, the correctly rounded cube (0.17 % of
    arguments differ, measured on the host), and which glibc variant runs depends on the CPU
    (ifunc: FMA or not); the CUDA kernel computes the correctly rounded cube.  A sample changes
    bin only when that 1-ulp difference straddles a bin edge: the mismatch count is reported and
    bounded, everything else is exact."""
    from sxmc_b200 import mcmc
    from tests import cases
    fit = cases.POLY_CASES[name]()
    z = np.load(os.path.join(HERE, "golden", "ref_fixtures.npz"))
    pars = z[name + "_pars"]
    hists, chain = mcmc.build_fit(fit)
    chain.set_data(fit.data)
    nll, ev = chain.nll(pars)
    bins, norms = chain.bins(), chain.norms()
    want_nll = float(z[name + "_nll"])
    moved = int(np.abs(bins.astype(np.int64) - z[name + "_bins"].astype(np.int64)).sum()) // 2
    print("%s: %d of %d in-domain samples in a different bin than the reference" % (name, moved, int(norms.sum())))
    if name == "P4":
        assert moved <= max(2, int(1e-6 * fit.total_samples()))
        assert abs(int(norms.sum()) - int(z[name + "_norms"].sum())) <= 2
        assert abs(nll - want_nll) <= 1e-5 * abs(want_nll)
    else:
        assert moved == 0 and np.array_equal(bins, z[name + "_bins"])
        assert np.array_equal(norms, z[name + "_norms"])
        assert chain.lut().tobytes() == z[name + "_lut"].tobytes()
        assert abs(nll - want_nll) <= NLL_RTOL * abs(want_nll)
        assert abs(ev - float(z[name + "_evsum"])) <= NLL_RTOL * abs(float(z[name + "_evsum"]))
    chain.close()
    for h in hists:
        h.close()


@pytest.mark.parametrize("case", ["D_shift_ctscale", "C_four_ops", "B_ctscale"])
def test_programs_outside_the_build_table_run_nvrtc_compiled_and_bit_exact(case):
    """A fit whose systematic program the build did not instantiate gets bin_fast_kernel compiled for it by
    NVRTC (jit.cu): bins, norms, lut bit-exact against the oracle, the same through the run-time dispatching
    kernel (JIT switched off), and a short chain identical between the two."""
    from sxmc_b200 import mcmc, pdfz
    if case == "D_shift_ctscale":
        fit = W.config_D(0.01)
        fit.systematics = [W.SystSpec(type=W.SHIFT, obs=1, means=[0.0], sigmas=[0.01]),
                           W.SystSpec(type=W.CTSCALE, obs=2, means=[0.0], sigmas=[0.01])]
    elif case == "C_four_ops":
        fit = W.config_C(0.01)
        fit.systematics = [W.SystSpec(type=W.SCALE, obs=0, means=[0.0], sigmas=[0.01]),
                           W.SystSpec(type=W.RESOLUTION_SCALE, obs=0, extra=2, means=[0.0], sigmas=[0.001]),
                           W.SystSpec(type=W.SHIFT, obs=1, means=[0.0], sigmas=[0.01]),
                           W.SystSpec(type=W.SCALE, obs=1, means=[0.0], sigmas=[0.01])]
    else:
        fit = W.config_B(0.05)
        fit.systematics = [W.SystSpec(type=W.CTSCALE, obs=0, means=[0.0], sigmas=[0.01])]
    fit.finalize()
    for sig in fit.signals:
        sig.systematics = list(range(len(fit.systematics)))
    out = {}
    # (True, 2): the NVRTC-compiled kernel launched in the four-row-chunk form (12-warp CTAs for D's
    # four columns) that gigabyte launches take
    for jit, chunk_form in ((True, 0), (False, 0), (True, 2)):
        pdfz.set_bin_tuning(no_jit=not jit, chunk_form=chunk_form)
        hists, chain = mcmc.build_fit(fit)
        chain.set_data(fit.data)
        assert chain.bin_kernel_kind() == (3 if jit else 1)
        for seed in (1, 2):
            pars = _pars(fit, seed)
            nll, ev = chain.nll(pars)
            bins, norms, lut = port.eval_all(fit, pars)
            want, _ = port.nll(fit, lut, pars, norms)
            assert np.array_equal(chain.bins(), bins) and np.array_equal(chain.norms(), norms)
            assert chain.lut().tobytes() == lut.tobytes()
            assert abs(nll - want) <= NLL_RTOL * abs(want)
        chain.set_seed(7, 0)
        chain.set_mode(1)
        out[(jit, chunk_form)] = chain(fit.data, 200, 0.1).get_samples()
        chain.close()
        for h in hists:
            h.close()
    pdfz.set_bin_tuning()
    assert out[(True, 0)].shape[0] > 0
    assert out[(True, 0)].tobytes() == out[(False, 0)].tobytes() == out[(True, 2)].tobytes()


def test_pdfz_group_ragged_batch_matches_oracle():
    """29 PDFs from 100 to 3x10^5 samples in ONE batched launch (the reference's pdfz_group
    benchmark shape at 1/10 size): bins, norms and PDF values bit-exact."""
    from sxmc_b200 import mcmc, pdfz
    fit = W.pdfz_group_shape(0.1)
    pars = fit.parameter_means().copy()
    pars[fit.nsources] = 0.0173
    bins, norms, lut = port.eval_all(fit, pars)
    hists, chain = mcmc.build_fit(fit)
    for h in hists:
        h.SetEvalPoints(fit.data)
    got_norms, got_pdf = pdfz.eval_hists_host(hists, pars, fit.nsources)
    assert np.array_equal(got_norms, norms)
    assert got_pdf.tobytes() == lut.tobytes()
    assert np.array_equal(np.stack([h.Bins() for h in hists]), bins)
    chain.close()
    for h in hists:
        h.close()


def test_empty_and_ragged_inputs():
    from sxmc_b200 import mcmc
    base = W.config_B(0.001)
    sigs = [W.SignalSpec(samples=base.signals[0].samples[:0], nexpected=10.0, source=0, n_mc=5),
            W.SignalSpec(samples=base.signals[1].samples[:1], nexpected=10.0, source=1),
            W.SignalSpec(samples=base.signals[2].samples[:1025], nexpected=10.0, source=2)]
    fit = W.FitSpec(name="ragged", lower=base.lower, upper=base.upper, nbins=base.nbins,
                    source_means=[1.0] * 3, source_sigmas=[0.0] * 3, source_fixed=[False] * 3,
                    systematics=base.systematics, signals=sigs, data=base.data[:7])
    hists, chain = mcmc.build_fit(fit)
    chain.set_data(fit.data)
    pars = fit.parameter_means().copy()
    nll, ev = chain.nll(pars)
    bins, norms, lut = port.eval_all(fit, pars)
    want, want_ev = port.nll(fit, lut, pars, norms)
    assert np.array_equal(chain.bins(), bins) and np.array_equal(chain.norms(), norms)
    assert np.array_equal(np.isnan(chain.lut()), np.isnan(lut))   # empty PDF -> NaN values
    assert (nll == want) or abs(nll - want) <= NLL_RTOL * abs(want)


# ---------------------------------------------------------------------------------------
# chains
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4], ids=["auto", "stream", "persistent", "graph", "monotone"])
def test_tape_driven_chain_matches_oracle_chain(mode):
    """Same random numbers in => same accept/reject sequence and the same float rows out."""
    from sxmc_b200 import mcmc
    fit = W.config_B(0.005)
    nsteps = 400
    rng = np.random.default_rng(123)
    normals = rng.normal(size=(nsteps + 1) * fit.nparameters)
    uniforms = rng.random(nsteps)
    want = port.mcmc(fit, nsteps, 0.1, normals, uniforms, trace=True)
    hists, chain = mcmc.build_fit(fit)
    chain.set_mode(mode)
    chain.set_tapes(normals, uniforms)
    chain.enable_trace()
    ls = chain(fit.data, nsteps, 0.1)
    got = ls.get_samples()
    assert chain.last_path() == {0: 4, 1: 1, 2: 2, 3: 1, 4: 4}[mode]
    nll_tr, prop_tr = chain.read_trace(nsteps)
    assert got.shape == want["chain"].shape
    assert chain.accepted() == want["accepted"]
    assert np.array_equal(prop_tr, want["prop_trace"])          # proposals bit-identical
    assert np.allclose(nll_tr, want["nll_trace"], rtol=NLL_RTOL, atol=0)
    assert np.array_equal(got[:, :-1], want["chain"][:, :-1])
    assert np.allclose(got[:, -1], want["chain"][:, -1], rtol=1e-6)
    # and against the chain the UNMODIFIED reference produced from the same tape
    z = np.load(os.path.join(HERE, "golden", "ref_fixtures.npz"))
    assert np.array_equal(z["chainB_normals"], normals)
    assert np.array_equal(got[:, :-1], z["chainB_chain"][:, :-1])


def test_fixed_systematics_and_debug_mode_chain():
    from sxmc_b200 import mcmc
    fit = W.config_A(0.004)
    for y in fit.systematics:
        y.fixed = True
    nsteps = 300
    rng = np.random.default_rng(5)
    normals = rng.normal(size=(nsteps + 1) * fit.nparameters)
    uniforms = rng.random(nsteps)
    for debug in (False, True):
        want = port.mcmc(fit, nsteps, 0.2, normals, uniforms, debug_mode=debug)
        hists, chain = mcmc.build_fit(fit)
        chain.set_tapes(normals, uniforms)
        got = chain(fit.data, nsteps, 0.2, debug_mode=debug).get_samples()
        assert got.shape == want["chain"].shape
        assert np.array_equal(got[:, :-1], want["chain"][:, :-1])


def test_philox_chain_posterior_agrees_with_oracle_chain():
    """RNG streams differ by construction (TRandom vs Philox): compare posteriors.  Quantiles
    must agree within Monte-Carlo error (here: 5 x the batch-means standard error)."""
    from sxmc_b200 import mcmc
    # config B's shape with 40 bins: at 1000 bins and a few thousand samples per signal the
    # likelihood is jagged in the shift and the ORACLE's own acceptance varies 1-10 % with the
    # seed, which makes quantiles of two independent chains incomparable
    fit = W._make("Bq", 1, 3, 1_000_000, 10_000, 1, [(W.SHIFT, 0)], 100_000, 0.005,
                  nbins_override=[40])
    nsteps = 20000
    rng = np.random.default_rng(99)
    want = port.mcmc(fit, nsteps, 0.1, rng.normal(size=(nsteps + 1) * fit.nparameters),
                     rng.random(nsteps))["chain"]
    hists, chain = mcmc.build_fit(fit)
    chain.set_seed(2024, 0)
    got = chain(fit.data, nsteps, 0.1).get_samples()
    assert got.shape == want.shape
    acc = chain.accepted() / nsteps
    assert 0.05 < acc < 0.9

    def batch_se(x, q, nb=20):
        parts = np.array_split(x, nb)
        return np.std([np.quantile(p, q) for p in parts]) / np.sqrt(nb)

    for k in range(fit.nparameters):
        for q in (0.16, 0.5, 0.84):
            se = np.hypot(batch_se(got[:, k], q), batch_se(want[:, k], q))
            assert abs(np.quantile(got[:, k], q) - np.quantile(want[:, k], q)) <= 5 * se + 1e-9, (k, q)
    # two seeds give different chains; the same seed reproduces bit-identically
    chain.set_seed(2024, 0)
    again = chain(fit.data, 2000, 0.1).get_samples()
    chain.set_seed(2024, 0)
    again2 = chain(fit.data, 2000, 0.1).get_samples()
    chain.set_seed(2025, 0)
    other = chain(fit.data, 2000, 0.1).get_samples()
    assert again.tobytes() == again2.tobytes() and again.tobytes() != other.tobytes()


# ---------------------------------------------------------------------------------------
# full-size properties (BASELINE.json sizes; the oracle still finishes in seconds for B)
# ---------------------------------------------------------------------------------------
def test_full_size_config_B_bit_exact_and_properties():
    from sxmc_b200 import mcmc
    fit = W.config_B(1.0)
    hists, chain = mcmc.build_fit(fit)
    chain.set_data(fit.data)
    pars = _pars(fit, 21)
    nll, ev = chain.nll(pars)
    bins, norms, lut = port.eval_all(fit, pars)
    want, _ = port.nll(fit, lut, pars, norms)
    assert np.array_equal(chain.bins(), bins) and np.array_equal(chain.norms(), norms)
    assert chain.lut().tobytes() == lut.tobytes()
    assert abs(nll - want) <= NLL_RTOL * abs(want)
    # size-independent properties: counts conserve, a huge shift empties the histogram
    b = chain.bins()
    assert np.array_equal(b.sum(axis=1), chain.norms())
    assert np.all(chain.norms() <= [s.n_mc for s in fit.signals])
    far = pars.copy()
    far[fit.nsources] = 1e6
    chain.nll(far)
    assert chain.bins().sum() == 0 and chain.norms().sum() == 0


def test_config_C_tenth_scale_bit_exact():
    from sxmc_b200 import mcmc
    fit = W.config_C(0.1)      # 10 x 1M samples, 5000 bins, scale + resolution_scale
    hists, chain = mcmc.build_fit(fit)
    chain.set_data(fit.data)
    pars = _pars(fit, 22)
    nll, ev = chain.nll(pars)
    bins, norms, lut = port.eval_all(fit, pars)
    want, _ = port.nll(fit, lut, pars, norms)
    assert np.array_equal(chain.bins(), bins) and np.array_equal(chain.norms(), norms)
    assert abs(nll - want) <= NLL_RTOL * abs(want)


# ---------------------------------------------------------------------------------------
# BASELINE.json sizes: C (10 x 10M), D (30 x 3.33M) and the 200 000-bin D, bit-exact against the
# oracle evaluated on the same box AND against digests of the compiled reference's output
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["Cfull", "Dfull", "Dstress"])
def test_full_size_bit_exact_against_oracle_and_reference_digests(name):
    import json
    from sxmc_b200 import mcmc
    from tests import cases
    with open(os.path.join(HERE, "golden", "ref_full_size.json")) as f:
        gold = json.load(f)[name]
    fit = cases.FULL_CASES[name]()
    pars = cases.off_nominal(fit, 41)
    assert [float(v) for v in pars] == gold["pars"]
    hists, chain = mcmc.build_fit(fit)
    chain.set_data(fit.data)
    # the launch shapes these sizes take (capi_hist.cu plan_launch): C privatised histograms and
    # two-row chunks on 16 warps; D four-row chunks on 12 warps; the 200 000-bin D two-row chunks
    want_plan = {"Cfull": dict(warps=16, rows_per_chunk=2, stages=4, smem_hist=True),
                 "Dfull": dict(warps=12, rows_per_chunk=4, stages=2, smem_hist=False),
                 "Dstress": dict(warps=16, rows_per_chunk=2, stages=3, smem_hist=False)}[name]
    assert chain.bin_plan() == want_plan and chain.bin_kernel_kind() == 2
    nll, ev = chain.nll(pars)
    bins, norms, lut = chain.bins(), chain.norms(), chain.lut()
    # the compiled reference's output (tests/golden/make_ref_fixtures.py --full)
    assert [int(v) for v in norms] == gold["norms"]
    assert int(bins.sum()) == gold["bins_sum"] and int((bins > 0).sum()) == gold["nonzero_bins"]
    assert cases.digest(bins) == gold["bins_sha256"]
    assert cases.digest(lut) == gold["lut_sha256"]
    assert abs(nll - gold["nll"]) <= NLL_RTOL * abs(gold["nll"])
    assert abs(ev - gold["evsum"]) <= NLL_RTOL * abs(gold["evsum"])
    # the oracle on this box, array for array
    b0, n0, l0 = port.eval_all(fit, pars)
    assert np.array_equal(bins, b0) and np.array_equal(norms, n0) and lut.tobytes() == l0.tobytes()
    want, want_ev = port.nll(fit, l0, pars, n0)
    assert abs(nll - want) <= NLL_RTOL * abs(want) and abs(ev - want_ev) <= NLL_RTOL * abs(want_ev)
    # size-independent properties
    assert np.array_equal(bins.sum(axis=1), norms)
    assert np.all(norms <= [s.n_mc for s in fit.signals])
    # one device-resident chain step sequence on the full-size fit: rows come back, NLL finite
    chain.set_seed(5, 0)
    rows = chain(fit.data, 40, 0.1).get_samples()
    assert rows.shape == (40 - 2 * 4, fit.nparameters + 1) and np.all(np.isfinite(rows))
    chain.close()
    for h in hists:
        h.close()
