"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/sxmc_b200.h declares, validates arguments with the reference's messages before it
touches a device, and refuses to compute without one (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import __graft_entry__ as graft
from tests import golden_vectors as G

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module", autouse=True)
def _built():
    graft.build()


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "sxmc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sxmc_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from sxmc_b200 import _lib
    L = C.CDLL(_lib.LIB_PATH)
    names = _declared_symbols()
    assert len(names) > 40
    for n in names:
        assert hasattr(L, n), "missing export: " + n
    # and the ctypes prototype table covers the same set
    assert sorted(_lib.SYMBOLS) == names


def test_library_is_sm100a_only():
    from sxmc_b200 import _lib
    import subprocess
    out = subprocess.run(["cuobjdump", "--list-elf", _lib.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = set(re.findall(r"sm_(\d+a?)", out.stdout))
    assert archs == {"100a"}, archs


@pytest.mark.parametrize("name,kw,msg", G.CTOR_ERRORS, ids=[c[0] for c in G.CTOR_ERRORS])
def test_ctor_errors_match_reference_messages(name, kw, msg):
    """test/test_pdfz.cpp:42-73 -- validation happens before any device work."""
    from sxmc_b200 import pdfz
    args = dict(samples=G.SAMPLES_1D, nfields=1, nobservables=1, lower=[0.0], upper=[1.0], nbins=[2])
    for k, v in kw.items():
        args["nobservables" if k == "nobs" else k] = v
    with pytest.raises(pdfz.Error) as e:
        pdfz.EvalHist(**args)
    assert e.value.msg == msg


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from sxmc_b200 import _lib, pdfz
    with pytest.raises(_lib.SxmcError) as e:
        pdfz.EvalHist(G.SAMPLES_1D, 1, 1, [0.0], [1.0], [2])
    assert e.value.code == 2 and "no CPU fallback" in e.value.msg


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under sxmc_b200/ may reference it."""
    pkg = os.path.join(ROOT, "sxmc_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, fn)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, fn
                assert "liboracle" not in txt and "libsxmc_ref" not in txt, fn


def test_nvrtc_compiles_a_systematic_program_outside_the_table():
    """jit.cu: bin_fast_kernel instantiated for a program the build does not carry -- compile only, no GPU.
    codes = 8 bits per systematic: type | slot_obs << 2 | slot_extra << 5 (src/pdfz.h:111-116 types)."""
    import ctypes as C
    from sxmc_b200 import _lib
    L = _lib.lib()
    SHIFT, SCALE, RES, CT = 0, 1, 2, 3
    n = C.c_size_t(0)
    programs = [(3, 3, 1, 2, (SHIFT | 1 << 2) | ((CT | 2 << 2) << 8)),                    # shift(radius), ctscale(costheta)
                (3, 3, 0, 2, (SHIFT | 1 << 2) | ((CT | 2 << 2) << 8)),                    # ... histograms in global memory
                (2, 3, 1, 4, SCALE | ((RES | 2 << 5) << 8) | ((SHIFT | 1 << 2) << 16) | ((SCALE | 1 << 2) << 24)),
                (1, 1, 1, 1, CT)]
    for nobs, ns, smem, nops, codes in programs:
        rc = L.sxmc_jit_compile_check(nobs, ns, smem, nops, codes, C.byref(n))
        assert rc == 0, L.sxmc_last_error().decode()
        assert n.value > 20000
    # a program that addresses a slot the kernel does not load must not compile (static_assert in apply_static)
    rc = L.sxmc_jit_compile_check(1, 1, 1, 1, RES | 0 << 2 | 3 << 5, C.byref(n))
    assert rc != 0 and "NVRTC could not compile" in L.sxmc_last_error().decode()
