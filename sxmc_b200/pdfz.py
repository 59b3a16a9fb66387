"""Host-side mirror of the reference's ``pdfz`` namespace (src/pdfz.h:93-574) over the C ABI.

Same names, argument meaning and error behaviour as the reference classes, so the parity
tests read like the reference's own gtest files (test/test_pdfz*.cpp):

    Error                       pdfz::Error            (src/pdfz.h:93-102)
    ShiftSystematic, ...        pdfz::*Systematic      (src/pdfz.h:109-233)
    Array                       hemi::Array<T>         (contrib/hemi/hemi/array.h:21-267)
    EvalHist                    pdfz::EvalHist         (src/pdfz.h:402-574)

All compute happens in libsxmc_b200.so on the GPU; nothing here does arithmetic."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import SxmcError, c_dp, c_fp, c_ip, c_sp, c_up, c_vp, check, lib, ptr

SHIFT, SCALE, RESOLUTION_SCALE, CTSCALE = 0, 1, 2, 3


class Error(Exception):
    """pdfz::Error: ``msg`` carries the reference's message (src/pdfz.h:93-102)."""

    def __init__(self, msg):
        super().__init__(msg)
        self.msg = msg


def _check(rc):
    try:
        check(rc)
    except SxmcError as e:
        if e.code == 1:      # SXMC_EINVAL == a pdfz::Error in the reference
            raise Error(e.msg) from None
        raise


class Array:
    """hemi::Array<T>: a host numpy mirror plus a device buffer with lazy, validity-tracked
    copies (contrib/hemi/hemi/array.h:106-174)."""

    def __init__(self, n, dtype, pinned=True):
        self.dtype = np.dtype(dtype)
        self.host = np.zeros(int(n), dtype=self.dtype)
        self._dev = c_vp()
        _check(lib().sxmc_buf_alloc(max(1, self.host.nbytes), C.byref(self._dev)))
        self.host_valid, self.dev_valid = True, False

    @classmethod
    def from_host(cls, values, dtype):
        a = cls(len(values), dtype)
        a.writeOnlyHostPtr()[:] = values
        return a

    def size(self):
        return self.host.size

    def copyFromHost(self, values):
        values = np.asarray(values, dtype=self.dtype)
        if values.size != self.host.size:
            self.free()
            self.__init__(values.size, self.dtype)
        self.host[:] = values
        self.host_valid, self.dev_valid = True, False

    # host side
    def readOnlyHostPtr(self):
        if not self.host_valid:
            _check(lib().sxmc_buf_download(self.host.ctypes.data_as(c_vp), self._dev, self.host.nbytes))
            self.host_valid = True
        return self.host

    def hostPtr(self):
        self.readOnlyHostPtr()
        self.dev_valid = False
        return self.host

    def writeOnlyHostPtr(self):
        self.host_valid, self.dev_valid = True, False
        return self.host

    # device side (raw address as int)
    def readOnlyDevicePtr(self):
        if not self.dev_valid:
            _check(lib().sxmc_buf_upload(self._dev, self.host.ctypes.data_as(c_vp), self.host.nbytes))
            self.dev_valid = True
        return self._dev

    def devicePtr(self):
        self.readOnlyDevicePtr()
        self.host_valid = False
        return self._dev

    def writeOnlyDevicePtr(self):
        self.dev_valid, self.host_valid = True, False
        return self._dev

    def free(self):
        if self._dev:
            lib().sxmc_buf_free(self._dev)
            self._dev = c_vp()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Systematic:
    def __init__(self, type_):
        self.type = type_


class ShiftSystematic(Systematic):
    """x' = x + p (src/pdfz.h:143-157)"""

    def __init__(self, obs, pars):
        super().__init__(SHIFT)
        self.obs, self.pars = obs, pars


class ScaleSystematic(Systematic):
    """x' = x * (1 + p) (src/pdfz.h:168-182)"""

    def __init__(self, obs, pars):
        super().__init__(SCALE)
        self.obs, self.pars = obs, pars


class CosThetaScaleSystematic(Systematic):
    """x' = 1 + (x - 1) * (1 + p) (src/pdfz.h:196-210)"""

    def __init__(self, obs, pars):
        super().__init__(CTSCALE)
        self.obs, self.pars = obs, pars


class ResolutionScaleSystematic(Systematic):
    """x' = x + p * (x - x_true) (src/pdfz.h:222-237)"""

    def __init__(self, obs, true_obs, pars):
        super().__init__(RESOLUTION_SCALE)
        self.obs, self.true_obs, self.pars = obs, true_obs, pars


def _pars_of(syst):
    p = syst.pars
    if isinstance(p, Array):
        return np.asarray(p.readOnlyHostPtr(), dtype=np.int16)
    return np.asarray(p, dtype=np.int16)


class EvalHist:
    """pdfz::EvalHist (src/pdfz.h:402-574; ctor src/pdfz.cpp:57-97,179-236)."""

    def __init__(self, samples, nfields, nobservables, lower, upper, nbins, dataset=0,
                 optimize=True):
        s = np.ascontiguousarray(samples, dtype=np.float32).ravel()
        lo = np.asarray(lower, dtype=np.float64).ravel()
        hi = np.asarray(upper, dtype=np.float64).ravel()
        nb = np.asarray(nbins, dtype=np.int32).ravel()
        pad = lambda a: a if a.size else np.zeros(1, a.dtype)
        self._h = c_vp()
        self.nobservables = nobservables
        self.nfields = nfields
        self._pdf = self._norm = self._params = None
        _check(lib().sxmc_hist_create(ptr(pad(s), c_fp), s.size, int(nfields), int(nobservables),
                                      ptr(pad(lo), c_dp), lo.size, ptr(pad(hi), c_dp), hi.size,
                                      ptr(pad(nb), c_ip), nb.size, int(dataset), C.byref(self._h)))

    def close(self):
        if self._h:
            lib().sxmc_hist_destroy(self._h)
            self._h = c_vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- Eval interface (src/pdfz.h:246-395)
    def AddSystematic(self, syst):
        extra = getattr(syst, "true_obs", 0)
        pars = _pars_of(syst)
        pp = pars if pars.size else np.zeros(1, np.int16)
        _check(lib().sxmc_hist_add_systematic(self._h, syst.type, syst.obs, extra, pars.size,
                                              ptr(pp, c_sp)))

    def SetEvalPoints(self, points):
        p = np.ascontiguousarray(points, dtype=np.float32).ravel()
        pp = p if p.size else np.zeros(1, np.float32)
        _check(lib().sxmc_hist_set_eval_points(self._h, ptr(pp, c_fp), p.size))

    def SetPDFValueBuffer(self, output, offset=0, stride=1):
        self._pdf = (output, offset, stride)

    def SetNormalizationBuffer(self, norm, offset=0):
        self._norm = (norm, offset)

    def SetParameterBuffer(self, params, offset=0, stride=1):
        self._params = (params, offset, stride)

    def _bind(self):
        L = lib()
        if self._params is not None:
            a, off, st = self._params
            _check(L.sxmc_hist_set_param_buffer(self._h, a.readOnlyDevicePtr(), off, st))
        if self._norm is not None:
            a, off = self._norm
            _check(L.sxmc_hist_set_norm_buffer(self._h, a.devicePtr(), off))
        if self._pdf is not None:
            a, off, st = self._pdf
            _check(L.sxmc_hist_set_pdf_buffer(self._h, a.devicePtr(), off, st))

    def EvalAsync(self, do_eval_pdf=True):
        self._bind()
        _check(lib().sxmc_hist_eval_async(self._h, int(do_eval_pdf)))

    def EvalFinished(self):
        _check(lib().sxmc_hist_eval_finished(self._h))

    # -- introspection
    def TotalBins(self):
        return lib().sxmc_hist_total_nbins(self._h)

    def BinVolume(self):
        return lib().sxmc_hist_bin_volume(self._h)

    def NSamples(self):
        return lib().sxmc_hist_nsamples(self._h)

    def Bins(self):
        out = np.zeros(self.TotalBins(), dtype=np.uint32)
        _check(lib().sxmc_hist_get_bins(self._h, ptr(out, c_up)))
        return out

    def ReadBins(self):
        n = lib().sxmc_hist_neval_points(self._h)
        out = np.zeros(max(1, n), dtype=np.int32)
        _check(lib().sxmc_hist_get_read_bins(self._h, ptr(out, c_ip)))
        return out[:n]

    def GetSamples(self):
        """src/pdfz.h:542-556: the samples in their original order and layout."""
        out = np.zeros(max(1, self.NSamples() * self.nfields), dtype=np.float32)
        _check(lib().sxmc_hist_get_samples(self._h, ptr(out, c_fp)))
        return out[:self.NSamples() * self.nfields]

    def SetLayout(self, sort=True):
        _check(lib().sxmc_hist_set_layout(self._h, int(sort)))

    def Prepare(self):
        _check(lib().sxmc_hist_prepare(self._h))


def eval_hists_host(hists, params, param_offset, want_pdf=True):
    """sxmc_hists_eval_host: host parameter vector in, (norms, pdf[ns][npoints]) out."""
    n = len(hists)
    harr = (c_vp * n)(*[h._h for h in hists])
    params = np.ascontiguousarray(params, dtype=np.float64)
    norms = np.zeros(n, dtype=np.uint32)
    npts = max([lib().sxmc_hist_neval_points(h._h) for h in hists] + [0])
    pdf = np.zeros((n, max(1, npts)), dtype=np.float32) if want_pdf and npts else None
    pp = params if params.size else np.zeros(1)
    _check(lib().sxmc_hists_eval_host(harr, n, ptr(pp, c_dp), params.size, param_offset,
                                      ptr(norms, c_up), ptr(pdf, c_fp) if pdf is not None else None))
    return norms, (pdf[:, :npts] if pdf is not None else None)


def set_bin_tuning(smem_hist_max_bins=-1, blocks_per_sm=0, force_generic=False,
                   force_dynamic=False, no_monotone=False, p2p_form=0, no_grouped_lookup=False, no_pdl=False, no_jit=False, cta_split=False, chunk_form=0):
    sel = (1 if force_generic else 0) | (2 if force_dynamic else 0) | (4 if no_monotone else 0) | ((int(p2p_form) & 3) << 3) | (32 if no_grouped_lookup else 0) | (64 if no_pdl else 0) | (128 if no_jit else 0) | ((int(cta_split) & 3) << 8) | ((int(chunk_form) & 3) << 10)
    _check(lib().sxmc_set_bin_tuning(smem_hist_max_bins, blocks_per_sm, sel))
