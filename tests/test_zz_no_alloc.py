"""sxmc_nll_eval_noalloc (include/sxmc_b200.h): MCMC::nll over a caller-held lut without any
allocation -- scratch in a fixed device array -- against sxmc_nll_eval, through the C ABI.
No counterpart in the reference; the file is
named to be collected last."""
import ctypes as C

import numpy as np
import pytest

NLL_RTOL = 1e-6   # the tolerance of the NLL parity tests (tests/test_gpu_parity.py)


@pytest.fixture(scope="module", autouse=True)
def _gpu():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from sxmc_b200 import _lib
    _lib.check(_lib.lib().sxmc_init(0))


def _case(seed, nevents, nsignals, nparameters):
    rng = np.random.default_rng(seed)
    lut = rng.random((nsignals, nevents)).astype(np.float32)
    lut[:, ::11] = np.nan                 # events without a PDF value (nll_kernels.cpp:107)
    lut[0, ::5] = 0.0
    return dict(
        lut=lut.ravel(), nevents=nevents, nsignals=nsignals, nparameters=nparameters, nsources=nsignals,
        pars=(0.5 + rng.random(nparameters)), means=np.ones(nparameters),
        sigmas=np.where(np.arange(nparameters) % 2 == 0, 0.1, 0.0),
        nexpected=(50.0 + 100.0 * rng.random(nsignals)), n_mc=rng.integers(1000, 5000, nsignals).astype(np.uint32),
        source_id=np.arange(nsignals).astype(np.int16), norms=rng.integers(500, 1000, nsignals).astype(np.uint32))


def _both(case):
    """(nll, event sum) from the existing entry point and from the no-allocation one."""
    from sxmc_b200 import _lib, pdfz
    L = _lib.lib()
    A = pdfz.Array.from_host
    arrs = [A(case["lut"], np.float32), A(case["pars"], np.float64), A(case["means"], np.float64),
            A(case["sigmas"], np.float64), A(case["nexpected"], np.float64), A(case["n_mc"], np.uint32),
            A(case["source_id"], np.int16), A(case["norms"], np.uint32)]
    lut, pars, means, sigmas, nexp, n_mc, sid, norms = [a.readOnlyDevicePtr() for a in arrs]
    sizes = (case["nevents"], case["nsignals"])
    v, ev = C.c_double(0), C.c_double(0)
    _lib.check(L.sxmc_nll_eval(lut, *sizes, pars, case["nparameters"], case["nsources"], means, sigmas, nexp, n_mc,
                               sid, norms, C.byref(v), C.byref(ev)))
    out = pdfz.Array(2, np.float64)            # the caller's output buffer (device)
    _lib.check(L.sxmc_nll_eval_noalloc(lut, *sizes, pars, case["nparameters"], case["nsources"], means, sigmas, nexp,
                                       n_mc, sid, norms, out.writeOnlyDevicePtr(), None))
    got = out.readOnlyHostPtr().copy()         # cudaMemcpy: ordered after the default stream's work
    for a in arrs + [out]:
        a.free()
    return (v.value, ev.value), (float(got[0]), float(got[1]))


@pytest.mark.gpu
def test_noalloc_nll_matches_nll_eval_on_two_calls_in_a_row():
    # first call: more events than 1184 blocks x 256 threads (grid at its cap, strided loop);
    # second call: a handful of blocks, so most of the scratch still holds the first call's partials
    for case in (_case(1, 400_000, 5, 9), _case(2, 1000, 3, 4)):
        want, got = _both(case)
        assert np.isfinite(want[0]) and want[0] != 1e18
        assert abs(got[0] - want[0]) <= NLL_RTOL * abs(want[0])
        assert abs(got[1] - want[1]) <= NLL_RTOL * abs(want[1])
    # a negative source rate: both return the reference's 1e18 guard (nll_kernels.cpp:162-165)
    case = _case(3, 1000, 3, 4)
    case["pars"][1] = -0.25
    want, got = _both(case)
    assert want[0] == 1e18 and got[0] == 1e18


@pytest.mark.gpu
def test_noalloc_nll_rejects_sizes_beyond_its_fixed_scratch():
    from sxmc_b200 import _lib, pdfz
    L = _lib.lib()
    dummy = pdfz.Array(4, np.float64)
    p = dummy.readOnlyDevicePtr()
    with pytest.raises(_lib.SxmcError) as e:
        _lib.check(L.sxmc_nll_eval_noalloc(p, 1, 129, p, 4, 4, p, p, p, p, p, p, p, None))
    assert "128 signals" in str(e.value)
    with pytest.raises(_lib.SxmcError) as e:
        _lib.check(L.sxmc_nll_eval_noalloc(p, (1 << 31) + 1, 2, p, 4, 2, p, p, p, p, p, p, p, None))
    assert "2^31" in str(e.value)
    dummy.free()
