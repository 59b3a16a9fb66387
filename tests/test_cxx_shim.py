"""The C++ class API of the reference (pdfz::Eval/EvalHist, hemi::Array, Signal, MCMC,
LikelihoodSpace; cxx/include/sxmc/*.h) written over the C ABI: a test runner in C++
(cxx/tests/test_class_api.cpp) drives it the way the reference's gtest suite does
(test/test_pdfz*.cpp), on the same golden constants, and runs one fit whose chain must equal
the ctypes mirror's chain row for row (same library, same Philox seed)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from tests import golden_vectors as G

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RUNNER = os.path.join(ROOT, "cxx", "tests", "test_class_api")
SHIM = os.path.join(ROOT, "sxmc_b200", "libsxmc_cxx.so")


def _fmt(v):
    return " ".join("nan" if (x is None or (isinstance(x, float) and np.isnan(x))) else repr(float(x)) for x in v)


def write_cases(path):
    """golden_vectors.CASES / CTOR_ERRORS in the runner's record format."""
    with open(path, "w") as f:
        for name, samples, geom, evalp, syst, value, norm, expect in G.CASES:
            nobs = len(geom["nbins"])
            stype, sobs, sextra = syst if syst is not None else (-1, 0, 0)
            f.write("case %s %d %d %d\n%s\n%s %s %s\n%d\n%s\n%d %d %d %r %d\n%s\n" % (
                name, samples.shape[0], samples.shape[1], nobs, _fmt(samples.ravel()),
                _fmt(geom["lower"]), _fmt(geom["upper"]), " ".join(str(int(b)) for b in geom["nbins"]),
                evalp.shape[0], _fmt(evalp.ravel()), stype, sobs, sextra, float(value), norm,
                _fmt(expect + [None] * (evalp.shape[0] - len(expect)))))
        for name, kw, msg in G.CTOR_ERRORS:
            nfields, nobs = kw.get("nfields", 1), kw.get("nobs", 1)
            lower, upper, nbins = kw.get("lower", [0.0]), kw.get("upper", [1.0]), kw.get("nbins", [2])
            f.write("error %s 7 %d %d %d %d %d %s %s\n" % (name, nfields, nobs, len(lower), len(upper), len(nbins),
                                                          " ".join(str(b) for b in nbins), msg))


def write_fit(fit, outdir, nsteps, seed):
    """FitSpec -> fit.txt + raw float32 files, the runner's input."""
    nfields = fit.signals[0].samples.shape[1]
    lines = ["%d %d" % (fit.nobs, nfields)]
    for i in range(fit.nobs):
        lines.append("%r %r %d" % (float(fit.lower[i]), float(fit.upper[i]), int(fit.nbins[i])))
    lines.append(str(fit.nsources))
    for i in range(fit.nsources):
        lines.append("%r %r %d" % (float(fit.source_means[i]), float(fit.source_sigmas[i]), int(fit.source_fixed[i])))
    lines.append(str(len(fit.systematics)))
    for y in fit.systematics:
        lines.append("%d %d %d %d %d %s %s %s" % (y.type, y.obs, y.extra, y.npars, int(y.fixed), _fmt(y.means[:y.npars]),
                                                 _fmt(y.sigmas[:y.npars]), " ".join(str(i) for i in y.pidx)))
    lines.append(str(fit.nsignals))
    for j, sig in enumerate(fit.signals):
        assert sig.systematics == list(range(len(fit.systematics)))
        name = "samples_%d.f32" % j
        np.ascontiguousarray(sig.samples, dtype=np.float32).tofile(os.path.join(outdir, name))
        lines.append("%r %d %d %d %s" % (float(sig.nexpected), sig.source, sig.n_mc, sig.samples.shape[0], name))
    fit.data.tofile(os.path.join(outdir, "data.f32"))
    lines.append("%d data.f32" % fit.data.size)
    lines.append("%d 0.1 %d" % (nsteps, seed))
    path = os.path.join(outdir, "fit.txt")
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return path


def test_shim_library_and_runner_are_built():
    """CPU check: the class-API library links against the C ABI and loads; the runner refuses to
    run without a device (no CPU fallback anywhere above the C ABI either)."""
    assert os.path.exists(SHIM), "run `make -C cxx` (python __graft_entry__.py does)"
    ctypes.CDLL(os.path.join(ROOT, "sxmc_b200", "libsxmc_b200.so"), mode=ctypes.RTLD_GLOBAL)
    ctypes.CDLL(SHIM)
    assert os.access(RUNNER, os.X_OK)
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([RUNNER, "x", "y", "z"], capture_output=True, text=True, timeout=120)
        assert r.returncode == 3 and "no CPU fallback" in r.stderr


def test_cxx_projection_estimator_equals_the_python_host_estimator():
    """sxmc::errors::Projection (src/error_estimators/projection.cpp:14-77) in cxx/src/projection.cpp
    against sxmc_b200.ensemble.projection_interval: same axis, same bin walk; the Gaussian fits are two
    implementations (Levenberg-Marquardt here, scipy there) of the same chi-square minimum."""
    from sxmc_b200 import ensemble
    lib = ctypes.CDLL(SHIM)
    fn = lib.sxmc_cxx_projection_interval
    fn.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.c_long, ctypes.c_long, ctypes.c_float,
                   ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int)]
    rng = np.random.default_rng(3)
    cols = {"gauss": rng.normal(5.0, 0.5, 200000), "exp": rng.exponential(1.0, 200000),
            "narrow": rng.normal(1e-3, 2e-5, 50000), "skew": rng.gamma(3.0, 2.0, 100000),
            "const": np.full(1000, 2.5)}
    table = np.stack([v[:1000] for v in cols.values()], axis=1).astype(np.float32)      # strided access
    for k, (name, v) in enumerate(cols.items()):
        for cl in (0.683, 0.9):
            for arr, n, stride in ((np.ascontiguousarray(v, np.float32), len(v), 1),
                                   (table[:, k:].ravel(), 1000, table.shape[1] - k)):
                want = ensemble.projection_interval(arr[::stride][:n] if stride > 1 else arr, cl)
                out = (ctypes.c_float * 5)()
                one = ctypes.c_int(0)
                assert fn(arr.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), n, stride, cl, out, ctypes.byref(one)) == 0
                assert bool(one.value) == want.one_sided, (name, cl)
                assert abs(out[0] - want.point_estimate) <= 2e-4 * max(1e-6, abs(want.point_estimate)) + 1e-9, (name, cl)
                assert np.float32(out[1]) == np.float32(want.lower) and np.float32(out[2]) == np.float32(want.upper), (name, cl)
                assert abs(out[4] - want.coverage) < 1e-6 and abs(out[3] - cl) < 1e-6


def test_case_files_round_trip(tmp_path):
    write_cases(tmp_path / "cases.txt")
    text = (tmp_path / "cases.txt").read_text()
    assert text.count("case ") == len(G.CASES) and text.count("error ") == len(G.CTOR_ERRORS)


@pytest.mark.gpu
def test_class_api_runner_matches_golden_vectors_and_ctypes_chain(tmp_path):
    from sxmc_b200 import mcmc, workloads as W
    fit = W.config_C(0.002)
    nsteps, seed = 1500, 4242
    write_cases(tmp_path / "cases.txt")
    fit_txt = write_fit(fit, str(tmp_path), nsteps, seed)
    r = subprocess.run([RUNNER, str(tmp_path / "cases.txt"), fit_txt, str(tmp_path)], capture_output=True,
                       text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert " 0 failures" in r.stdout

    # the same fit through the ctypes mirror: identical rows, acceptance count, efficiencies
    hists, chain = mcmc.build_fit(fit)
    chain.set_seed(seed, 0)
    ls = chain(fit.data, nsteps, 0.1)
    rows = ls.get_samples()
    got = np.fromfile(tmp_path / "rows.bin", dtype=np.float32).reshape(-1, fit.nparameters + 1)
    assert got.shape == rows.shape and got.tobytes() == rows.tobytes()
    head = (tmp_path / "fit_out.txt").read_text().split("\n")
    nrows, accepted, _ = head[0].split()
    assert int(nrows) == rows.shape[0] and int(accepted) == chain.accepted()
    eff = np.asarray([float(x) for x in head[1:1 + fit.nsignals]])
    pars = fit.parameter_means()
    chain.nll(pars)
    want = chain.norms() / np.asarray([s.n_mc for s in fit.signals], dtype=np.float64)
    assert np.array_equal(eff, want)
    # the contour intervals bracket the point estimate
    for line in head[1 + fit.nsignals:1 + fit.nsignals + fit.nparameters]:
        _, pt, lo, hi = line.split()
        assert float(lo) <= float(pt) <= float(hi)
    chain.close()
    for h in hists:
        h.close()


SXMC_BIN = os.path.join(ROOT, "cxx", "bin", "sxmc")


def test_cxx_command_line_usage_and_no_cpu_fallback():
    """cxx/bin/sxmc keeps the reference's argument convention (src/sxmc.cpp:176-182) and refuses to
    run without a device."""
    assert os.access(SXMC_BIN, os.X_OK), "run `make -C cxx`"
    r = subprocess.run([SXMC_BIN], capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "fit_configuration.json output_path" in r.stderr
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([SXMC_BIN, "x.json", "/tmp/sxmc_no_gpu_out"], capture_output=True, text=True, timeout=60)
        assert r.returncode == 3 and "no CPU fallback" in r.stderr


@pytest.mark.gpu
def test_cxx_command_line_equals_python_command_line(tmp_path):
    """The same JSON configuration and .npy sample tables through the C++ main (FitConfig, Signal,
    make_fake_dataset, MCMC, LikelihoodSpace in cxx/) and through `python -m sxmc_b200`: every
    experiment's likelihood space bit for bit, the same limits."""
    import sys
    from tests.test_config import CONFIG
    rng = np.random.default_rng(5)
    for name, n, mu in (("signal1", 20000, 8.0), ("signal2", 40000, 11.0)):
        e = rng.normal(mu, 1.5, n).astype(np.float32)
        table = np.stack([(12 * rng.random(n) ** (1 / 3)).astype(np.float32),
                          (e - rng.normal(0, 0.3, n)).astype(np.float32), e, np.zeros(n, np.float32)], axis=1)
        np.save(tmp_path / (name + ".npy"), np.ascontiguousarray(table))
        (tmp_path / (name + ".npy.fields")).write_text("radius\nmc_energy\nenergy\nunused\n")
    cfg = tmp_path / "fit.json"
    cfg.write_text(CONFIG.replace(".npz", ".npy"))
    rc = subprocess.run([SXMC_BIN, str(cfg), str(tmp_path / "out_cxx")], capture_output=True, text=True, timeout=900)
    assert rc.returncode == 0, rc.stdout[-3000:] + rc.stderr[-3000:]
    rp = subprocess.run([sys.executable, "-m", "sxmc_b200", str(cfg), str(tmp_path / "out_py")], cwd=ROOT,
                        capture_output=True, text=True, timeout=900)
    assert rp.returncode == 0, rp.stdout[-3000:] + rp.stderr[-3000:]
    # ... and once more with the C++ main forced onto its one-fit-at-a-time loop (MCMC::operator())
    r1 = subprocess.run([SXMC_BIN, str(cfg), str(tmp_path / "out_cxx1")], capture_output=True, text=True, timeout=900,
                        env=dict(os.environ, SXMC_NO_BATCH="1"))
    assert r1.returncode == 0, r1.stdout[-3000:] + r1.stderr[-3000:]
    assert "Sampling fake dataset" in r1.stdout and "Sampling fake dataset" not in rc.stdout
    for i in range(3):
        a = np.load(tmp_path / "out_cxx" / ("example_%d.npy" % i))
        a1 = np.load(tmp_path / "out_cxx1" / ("example_%d.npy" % i))
        b = np.load(tmp_path / "out_py" / ("example_%d.npy" % i))
        assert a.shape == b.shape == (3000 - 2 * 600, 5) and a.tobytes() == b.tobytes(), i
        assert a1.tobytes() == b.tobytes(), i
    cols = (tmp_path / "out_cxx" / "example_columns.txt").read_text().split()
    assert cols == ["signal1", "signal2", "energy_resolution_0", "energy_scale_0", "likelihood"]

    def limits(text):
        line = [ln for ln in text.splitlines() if ln.startswith("sxmc: Upper limits:")][0]
        return [float(x) for x in line.split(":", 2)[2].split(",") if x.strip()]
    assert np.allclose(limits(rc.stdout), limits(rp.stdout), rtol=1e-5)
    assert np.allclose(limits(r1.stdout), limits(rp.stdout), rtol=1e-5)
    assert "sxmc: Median upper limit:" in rc.stdout

    # fit.debug_mode reaches every experiment's MCMC::operator() (src/sxmc.cpp:97) on both command
    # lines: every step accepted, no chain reset at burn-in -> nsteps rows, identical files
    import json as _json
    import re as _re
    text = CONFIG.replace(".npz", ".npy")
    assert '"debug_mode"' not in text
    text = _re.sub(r'"nsteps"\s*:', '"debug_mode": true, "nsteps":', text, count=1)
    cfgd = tmp_path / "fit_debug.json"
    cfgd.write_text(text)
    rd = subprocess.run([SXMC_BIN, str(cfgd), str(tmp_path / "dbg_cxx")], capture_output=True, text=True, timeout=900)
    assert rd.returncode == 0, rd.stdout[-3000:] + rd.stderr[-3000:]
    rq = subprocess.run([sys.executable, "-m", "sxmc_b200", str(cfgd), str(tmp_path / "dbg_py")], cwd=ROOT,
                        capture_output=True, text=True, timeout=900)
    assert rq.returncode == 0, rq.stdout[-3000:] + rq.stderr[-3000:]
    for i in range(3):
        a = np.load(tmp_path / "dbg_cxx" / ("example_%d.npy" % i))
        b = np.load(tmp_path / "dbg_py" / ("example_%d.npy" % i))
        assert a.shape == b.shape == (3000, 5) and a.tobytes() == b.tobytes(), i
