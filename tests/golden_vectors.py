"""The reference's own gtest golden vectors for pdfz::EvalHist, transcribed as data.

Sources (paths relative to /root/reference): test/test_pdfz_fixtures.h:7-75,
test/test_pdfz_fixtures_2d.h:7-79, test/test_pdfz.cpp:79-126, test/test_pdfz_2d.cpp:54-104,
test/test_pdfz_syst.cpp:39-272.  The tests at that commit are stale against src/pdfz.h
(SURVEY.md section 4); the mechanical adaptation is: append a 0 dataset-id column to every
eval point, and give each systematic the one-element coefficient list [0] (the tests'
`int` parameter index 0).  All expected values are the tests' ASSERT_FLOAT_EQ constants
(4-ULP float compare); NaN is written as None.
"""
import numpy as np

from sxmc_b200.workloads import SHIFT, SCALE, RESOLUTION_SCALE

NAN = None

# test/test_pdfz_fixtures.h:7-30 (1-D: 7 samples, range [0,1), 2 bins)
SAMPLES_1D = np.array([0.1, 0.2, 0.3, 0.4, 0.5, 1.1, -0.1], dtype=np.float32).reshape(7, 1)
# test/test_pdfz_syst.cpp:168-177 (observable + constant truth column 0.7)
SAMPLES_1D_TRUTH = np.stack([SAMPLES_1D[:, 0], np.full(7, 0.7, np.float32)], axis=1)
# test/test_pdfz_fixtures.h:48-54, dataset column appended
EVAL_1D = np.array([[-0.1, 0], [0.0, 0], [0.25, 0], [0.5, 0], [0.75, 0], [1.0, 0]],
                   dtype=np.float32)
GEOM_1D = dict(lower=[0.0], upper=[1.0], nbins=[2])

# test/test_pdfz_fixtures_2d.h:12-31,48-56
SAMPLES_2D = np.array([[0.4, 10.5], [0.5, 11.0], [0.75, 11.0], [0.6, 11.5], [0.6, 11.8],
                       [0.9, 11.5], [0.4, 12.0]], dtype=np.float32)
EVAL_2D = np.array([[0.2, 10.2, 0], [0.7, 10.4, 0], [0.5, 11.0, 0], [0.25, 11.8, 0],
                    [0.9, 11.9, 0], [0.3, 12.0, 0], [0.3, 13.0, 0], [0.3, 5.0, 0]],
                   dtype=np.float32)
GEOM_2D = dict(lower=[0.0, 10.0], upper=[1.0, 12.0], nbins=[2, 3])
_N2D = 6 * (0.5 * 2.0 / 3.0)  # norm * bin volume, test/test_pdfz_2d.cpp:58

# (name, samples, geometry, eval points, systematic (type, obs, extra) or None,
#  parameter value, expected norm, expected pdf values)
CASES = [
    # test/test_pdfz.cpp:79-96
    ("1d_nosyst", SAMPLES_1D, GEOM_1D, EVAL_1D, None, 0.0, 5, [NAN, 1.6, 1.6, 0.4, 0.4, NAN]),
    # test/test_pdfz_syst.cpp:39-87
    ("shift_zero", SAMPLES_1D, GEOM_1D, EVAL_1D, (SHIFT, 0, 0), 0.0, 5,
     [NAN, 1.6, 1.6, 0.4, 0.4, NAN]),
    ("shift_neg", SAMPLES_1D, GEOM_1D, EVAL_1D, (SHIFT, 0, 0), -0.25, 4,
     [NAN, 1.5, 1.5, 0.5, 0.5, NAN]),
    ("shift_pos", SAMPLES_1D, GEOM_1D, EVAL_1D, (SHIFT, 0, 0), 0.25, 6,
     [NAN, 1.0, 1.0, 1.0, 1.0, NAN]),
    # test/test_pdfz_syst.cpp:109-157
    ("scale_zero", SAMPLES_1D, GEOM_1D, EVAL_1D, (SCALE, 0, 0), 0.0, 5,
     [NAN, 1.6, 1.6, 0.4, 0.4, NAN]),
    ("scale_neg", SAMPLES_1D, GEOM_1D, EVAL_1D, (SCALE, 0, 0), -0.1, 6,
     [NAN, 5.0 / 3, 5.0 / 3, 1.0 / 3, 1.0 / 3, NAN]),
    ("scale_pos", SAMPLES_1D, GEOM_1D, EVAL_1D, (SCALE, 0, 0), 1.0, 4,
     [NAN, 1.0, 1.0, 1.0, 1.0, NAN]),
    # test/test_pdfz_syst.cpp:224-272
    ("res_zero", SAMPLES_1D_TRUTH, GEOM_1D, EVAL_1D, (RESOLUTION_SCALE, 0, 1), 0.0, 5,
     [NAN, 1.6, 1.6, 0.4, 0.4, NAN]),
    ("res_neg", SAMPLES_1D_TRUTH, GEOM_1D, EVAL_1D, (RESOLUTION_SCALE, 0, 1), -0.30, 7,
     [NAN, 10.0 / 7, 10.0 / 7, 4.0 / 7, 4.0 / 7, NAN]),
    ("res_pos", SAMPLES_1D_TRUTH, GEOM_1D, EVAL_1D, (RESOLUTION_SCALE, 0, 1), 0.30, 4,
     [NAN, 2.0, 2.0, 0.0, 0.0, NAN]),
    # test/test_pdfz_2d.cpp:54-67 (indices 5..7 are all outside the domain: NaN)
    ("2d_nosyst", SAMPLES_2D, GEOM_2D, EVAL_2D, None, 0.0, 6,
     [1.0 / _N2D, 0.0, 2.0 / _N2D, 0.0, 3.0 / _N2D, NAN, NAN, NAN]),
]

# test/test_pdfz.cpp:98-126: pdf buffer offset 3 stride 2, norm slot 1 of {77,88,99}
OFFSET_STRIDE = dict(pdf_offset=3, pdf_stride=2, norm_offset=1, norm_init=[77, 88, 99],
                     norm_expect=[77, 5, 99],
                     pdf_expect={3: NAN, 5: 1.6, 7: 1.6, 9: 0.4, 11: 0.4, 13: NAN})

# test/test_pdfz.cpp:42-73: constructor argument errors -> pdfz::Error
# (kwargs override the 1-D fixture: samples has 7 floats, nfields=1, nobs=1)
CTOR_ERRORS = [
    ("wrong_sample_size", dict(nfields=2),
     "Length of samples array is not divisible by number of fields."),
    ("nobs_larger_than_nfields", dict(nobs=7),
     "Number of observables cannot be greater than number of fields."),
    ("wrong_lower_size", dict(lower=[0.0, 0.0]),
     "Number of lower bounds must be same as number of observables."),
    ("wrong_upper_size", dict(upper=[1.0, 0.0]),
     "Number of upper bounds must be same as number of observables."),
    ("wrong_nbins_size", dict(nbins=[2, 0]),
     "Size of nbins array must be same as number of observables."),
    ("zero_bins", dict(nbins=[0]), "Cannot make histogram with zero bins."),
]


def float_eq_4ulp(a, b):
    """gtest ASSERT_FLOAT_EQ: within 4 ULPs as float32."""
    a32, b32 = np.float32(a), np.float32(b)
    if np.isnan(a32) or np.isnan(b32):
        return False
    ia = np.array([a32], dtype=np.float32).view(np.int32)[0]
    ib = np.array([b32], dtype=np.float32).view(np.int32)[0]
    bias = lambda i: (np.int64(0x80000000) - np.int64(i)) if i < 0 else np.int64(i) + np.int64(0x80000000)
    return abs(int(bias(ia)) - int(bias(ib))) <= 4


def check_pdf_values(got, expect):
    assert len(got) >= len(expect)
    for i, e in enumerate(expect):
        if e is None:
            assert np.isnan(got[i]), (i, got[i])
        else:
            assert float_eq_4ulp(got[i], e), (i, got[i], e)


def case_fit(samples, geom, evalp, syst, value):
    """Wrap one golden case as a one-signal FitSpec (one source, coefficient index 0)."""
    from sxmc_b200.workloads import FitSpec, SignalSpec, SystSpec
    systs = []
    if syst is not None:
        systs = [SystSpec(type=syst[0], obs=syst[1], extra=syst[2], means=[value],
                          sigmas=[0.0])]
    return FitSpec(name="golden", lower=geom["lower"], upper=geom["upper"],
                   nbins=geom["nbins"], source_means=[1.0], source_sigmas=[0.0],
                   source_fixed=[False], systematics=systs,
                   signals=[SignalSpec(samples=samples, nexpected=5.0, source=0)],
                   data=evalp)
