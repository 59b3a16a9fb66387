"""The JSON fit configuration of sxmc and its sample reader, without ROOT.

``FitConfig(filename)`` follows `src/config.cpp:19-297` key for key: the same JSON schema
(`config/example.json`), the same defaults, the same ordering rules (sources and systematics
are collected in the iteration order of the ``signals`` OBJECT, which jsoncpp keeps sorted by
key; sample columns = observables' fields, then truth fields of resolution systematics, then
DATASET, `:162-194`), the same ``rate`` / ``scale`` convention (`:213-221`,
`src/signal.cpp:31-36`) and the same cut semantics (`src/signal.cpp:50-108`: an event is kept
when every cut field lies inside [lower, upper], both ends included; ``n_mc`` counts the rows
BEFORE cuts).

The reference reads ROOT TTrees/TNtuples (`src/io/ttree_io.cpp:21-159`).  ROOT is not on this
image, so ``read_float_table`` is the drop-in for ``read_float_vector_ttree``: it returns the
same thing -- a row-major float32 table and its column names -- from

* ``.npz``  : one 1-D array per column (keys = field names), or ``data`` [n, k] + ``fields``;
* ``.npy``  : a structured array (named float fields), or a plain [n, k] array next to a
  ``<file>.fields`` text file with one column name per line;
* ``.csv`` / ``.txt`` : header line of column names, then numbers;
* ``.f32``  : flat little-endian float32, row-major, with the same ``.fields`` sidecar.

A ``.root`` path fails loudly with the conversion hint.  Nothing here computes on samples
beyond selecting rows and columns; the fit itself runs through ``sxmc_b200.mcmc`` /
``sxmc_b200.ensemble`` (CUDA only).
"""
import json
import os
import re

import numpy as np

from . import workloads as W

TYPE_OF = {"shift": W.SHIFT, "scale": W.SCALE, "ctscale": W.CTSCALE, "resolution_scale": W.RESOLUTION_SCALE}


class ConfigError(ValueError):
    """A failed `assert` / `throw` of src/config.cpp, with the offending key."""


def _need(obj, key, where):
    if key not in obj:
        raise ConfigError("%s: missing key '%s'" % (where, key))
    return obj[key]


def load_json(filename):
    """jsoncpp accepts // and /* */ comments where Python's json does not: they are stripped.
    Like the reference's reader (jsoncpp 0.6, pinned by tests/test_config.py against the compiled
    reference) a missing comma is a parse error -- `config/example.json:42-43` as shipped does not
    load in the reference either."""
    with open(filename) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"(^|[^:\"])//[^\n]*", r"\1", text)
    try:
        return json.loads(text)
    except json.JSONDecodeError as e:
        raise ConfigError("FitConfig::FitConfig: JSON parse error: %s" % e)    # src/config.cpp:28-33


def read_float_table(filename):
    """-> (float32 [nrows, nfields] C-contiguous, [field names]); see the module docstring."""
    ext = os.path.splitext(filename)[1].lower()
    if ext == ".root":
        raise ConfigError("%s: ROOT files cannot be read on this box (no ROOT); convert the tree to .npz "
                          "(one array per branch) and point 'filename' at it" % filename)
    if not os.path.exists(filename):
        raise ConfigError("%s: no such file" % filename)

    def sidecar():
        p = filename + ".fields"
        if not os.path.exists(p):
            raise ConfigError("%s: needs a '%s' file naming its columns" % (filename, p))
        with open(p) as f:
            return [ln.strip() for ln in f if ln.strip()]

    if ext == ".npz":
        z = np.load(filename)
        if "data" in z.files and "fields" in z.files:
            fields = [str(x) for x in z["fields"]]
            data = np.asarray(z["data"], dtype=np.float32)
        else:
            fields = list(z.files)
            cols = [np.asarray(z[k], dtype=np.float32).ravel() for k in fields]
            if len(set(c.size for c in cols)) > 1:
                raise ConfigError("%s: columns of different lengths" % filename)
            data = np.stack(cols, axis=1) if cols else np.zeros((0, 0), np.float32)
    elif ext == ".npy":
        a = np.load(filename)
        if a.dtype.names:
            fields = list(a.dtype.names)
            data = np.stack([np.asarray(a[k], dtype=np.float32) for k in fields], axis=1)
        else:
            fields = sidecar()
            data = np.asarray(a, dtype=np.float32)
    elif ext in (".csv", ".txt"):
        with open(filename) as f:
            fields = [t for t in re.split(r"[,\s]+", f.readline().strip().lstrip("#")) if t]
        data = np.loadtxt(filename, dtype=np.float32, delimiter="," if ext == ".csv" else None, skiprows=1, ndmin=2)
    elif ext == ".f32":
        fields = sidecar()
        data = np.fromfile(filename, dtype="<f4")
        if data.size % len(fields):
            raise ConfigError("%s: size is not a multiple of %d columns" % (filename, len(fields)))
        data = data.reshape(-1, len(fields))
    else:
        raise ConfigError("%s: unknown sample file type" % filename)
    if data.ndim != 2 or data.shape[1] != len(fields):
        raise ConfigError("%s: %s does not match %d column names" % (filename, data.shape, len(fields)))
    return np.ascontiguousarray(data, dtype=np.float32), fields


def read_dataset_to_samples(table, fields, dataset_id, sample_fields, cuts):
    """Signal::read_dataset_to_samples (src/signal.cpp:50-108): apply the cuts on the file's own
    columns, then pick the sample columns in `sample_fields` order and append the dataset tag."""
    keep = np.ones(table.shape[0], dtype=bool)
    for j, name in enumerate(fields):
        # the lookup table of :57-67 holds ONE cut per file column: the last one naming it wins
        cut = None
        for c in cuts:
            if c.field == name:
                cut = c
        if cut is not None:
            col = table[:, j]
            # float column against the Observable's float bounds widened to double (:57-58,86)
            keep &= ~((col.astype(np.float64) < float(np.float32(cut.lower))) |
                      (col.astype(np.float64) > float(np.float32(cut.upper))))
    idx = []
    for name in sample_fields[:-1]:
        if name not in fields:
            raise ConfigError("sample field '%s' is not a column of the file (%s)" % (name, ", ".join(fields)))
        idx.append(fields.index(name))
    out = np.empty((int(keep.sum()), len(sample_fields)), dtype=np.float32)
    out[:, :-1] = table[keep][:, idx]
    out[:, -1] = np.float32(dataset_id)
    return out


class ObservableConfig:
    """Observable::Observable(name, Json::Value) (src/observable.cpp:8-34)"""

    def __init__(self, name, cfg):
        where = "observable '%s'" % name
        self.name = name
        self.title = _need(cfg, "title", where)
        self.field = _need(cfg, "field", where)
        self.bins = int(_need(cfg, "bins", where))
        self.lower = float(np.float32(_need(cfg, "min", where)))
        self.upper = float(np.float32(_need(cfg, "max", where)))
        self.units = cfg.get("units", "")
        self.logscale = bool(cfg.get("logscale", False))
        self.yrange = [float(x) for x in cfg.get("yrange", [-1, -1])]
        self.field_index = -1


class SystematicConfig:
    """Systematic::Systematic(name, Json::Value) (src/systematic.cpp:10-66)"""

    def __init__(self, name, cfg):
        where = "systematic '%s'" % name
        self.name = name
        self.title = _need(cfg, "title", where)
        self.observable_field = _need(cfg, "observable_field", where)
        t = _need(cfg, "type", where)
        if t not in TYPE_OF:
            raise ConfigError("FitConfig::load_pdf_systematics: Unknown systematic type %s" % t)
        self.type = TYPE_OF[t]
        self.truth_field = _need(cfg, "truth_field", where) if self.type == W.RESOLUTION_SCALE else ""
        self.means = [float(x) for x in _need(cfg, "mean", where)]
        self.npars = len(self.means)
        self.sigmas = [float(x) for x in cfg["sigma"]] if "sigma" in cfg else [0.0] * self.npars
        if len(self.sigmas) != self.npars:
            raise ConfigError("%s: 'sigma' must have as many entries as 'mean'" % where)
        self.fixed = bool(cfg.get("fixed", False))
        self.pidx = []
        self.observable_field_index = -1
        self.truth_field_index = 0


class SourceConfig:
    """Source (src/source.h:15-34, src/source.cpp:7-12)"""

    def __init__(self, name, index, mean=1.0, sigma=0.0, fixed=False):
        self.name, self.index = name, int(index)
        self.mean, self.sigma = float(np.float32(mean)), float(np.float32(sigma))
        self.fixed = bool(fixed)


class SignalConfig:
    """Signal::Signal(name, title, filename, ...) (src/signal.cpp:11-47) minus build_pdfz."""

    def __init__(self, name, title, filename, dataset, source, nexpected, sample_fields, cuts, systematics,
                 base_dir=""):
        self.name, self.title, self.filename = name, title, filename
        self.dataset, self.source = int(dataset), source
        path = filename if os.path.isabs(filename) or not base_dir else os.path.join(base_dir, filename)
        table, fields = read_float_table(path)
        self.n_mc = int(table.shape[0])                                  # rank[0]: rows before cuts (:27)
        self.nexpected = float(nexpected)
        if self.nexpected < 0:                                           # 'scale' given instead of 'rate' (:32-36)
            self.nexpected *= -1.0 * self.n_mc
        self.samples = read_dataset_to_samples(table, fields, self.dataset, sample_fields, cuts)
        self.systematic_names = [s.name for s in systematics]


class FitConfig:
    """FitConfig::FitConfig(filename) (src/config.cpp:19-297)."""

    def __init__(self, filename):
        root = load_json(filename)
        base = os.path.dirname(os.path.abspath(filename))
        fit = _need(root, "fit", "config")
        obs_params = _need(_need(root, "pdfs", "config"), "observables", "pdfs")
        sys_params = root["pdfs"].get("systematics", {})
        sig_params = _need(root, "signals", "config")
        src_params = root.get("sources", {})

        self.nexperiments = int(_need(fit, "nexperiments", "fit"))                       # :43-44
        self.nsteps = int(_need(fit, "nsteps", "fit"))                                   # :47-48
        self.samples = fit.get("samples", "")
        self.error_type = fit.get("error_type", "contour")                               # :52-63
        if self.error_type not in ("contour", "projection"):
            raise ConfigError("FitConfig::FitConfig: Unknown error type %s" % self.error_type)
        self.burnin_fraction = float(np.float32(fit.get("burnin_fraction", 0.1)))
        self.debug_mode = bool(fit.get("debug_mode", False))
        self.output_prefix = fit.get("output_prefix", "lspace")
        self.plots = bool(fit.get("plots", True))
        self.seed = int(fit.get("seed", 0))
        self.confidence = float(np.float32(fit.get("confidence", 0.683)))
        self.signal_name = fit.get("signal_name", "")

        self.observables = []                                                            # :74-80
        for name in _need(fit, "observables", "fit"):
            if name not in obs_params:
                raise ConfigError("fit.observables: '%s' is not defined under pdfs.observables" % name)
            self.observables.append(ObservableConfig(name, obs_params[name]))
        self.cuts = []                                                                   # :82-91
        for name in fit.get("cuts", []):
            if name not in obs_params:
                raise ConfigError("fit.cuts: '%s' is not defined under pdfs.observables" % name)
            if any(o.name == name for o in self.observables):
                raise ConfigError("fit.cuts: '%s' is already an observable" % name)
            self.cuts.append(ObservableConfig(name, obs_params[name]))

        # systematics and sources: union over ALL configured signals, in jsoncpp's (sorted) key order (:93-151)
        self.systematics, self.sources = [], []
        pidx = 0
        for sname in sorted(sig_params.keys()):
            for yname in sig_params[sname].get("systematics", []):
                if yname not in sys_params:
                    raise ConfigError("signal '%s': systematic '%s' is not defined" % (sname, yname))
                if not any(y.name == yname for y in self.systematics):
                    y = SystematicConfig(yname, sys_params[yname])
                    y.pidx = list(range(pidx, pidx + y.npars))
                    pidx += y.npars
                    self.systematics.append(y)
            if "source" in sig_params[sname]:
                src = sig_params[sname]["source"]
                if src not in src_params:
                    raise ConfigError("signal '%s': source '%s' is not defined" % (sname, src))
                if not any(s.name == src for s in self.sources):
                    p = src_params[src]
                    self.sources.append(SourceConfig(src, len(self.sources), p.get("mean", 1.0), p.get("sigma", 0.0),
                                                     p.get("fixed", False)))
            else:                                   # the signal is a source for itself
                p = sig_params[sname]
                self.sources.append(SourceConfig(sname, len(self.sources), p.get("mean", 1.0), p.get("sigma", 0.0),
                                                 p.get("fixed", False)))

        # sample columns (:153-194)
        self.sample_fields = []

        def index_with_append(name):
            if name not in self.sample_fields:
                self.sample_fields.append(name)
            return self.sample_fields.index(name)

        for o in self.observables:
            o.field_index = index_with_append(o.field)
        for y in self.systematics:
            if y.observable_field not in self.sample_fields:
                raise ConfigError("systematic '%s': field '%s' is not an observable of the fit"
                                  % (y.name, y.observable_field))
            y.observable_field_index = self.sample_fields.index(y.observable_field)
            if y.type == W.RESOLUTION_SCALE:
                y.truth_field_index = index_with_append(y.truth_field)
        self.sample_fields.append("DATASET")

        # signals of the fit (:196-257)
        self.signals, self.datasets = [], set()
        for name in _need(fit, "signals", "fit"):
            if name not in sig_params:
                raise ConfigError("fit.signals: '%s' is not defined under signals" % name)
            cfg = sig_params[name]
            where = "signal '%s'" % name
            dataset = int(_need(cfg, "dataset", where))
            self.datasets.add(dataset)
            if ("rate" in cfg) == ("scale" in cfg):
                raise ConfigError("%s: exactly one of 'rate' and 'scale' is required" % where)
            # `float nexpected = -1.0 / config["scale"].asFloat()` (:213-221): the quotient is rounded to float
            nexpected = (float(np.float32(cfg["rate"])) if "rate" in cfg
                         else float(np.float32(-1.0 / float(np.float32(cfg["scale"])))))
            systs = [y for yname in cfg.get("systematics", []) for y in self.systematics if y.name == yname]
            source = next((s for s in self.sources if s.name == cfg.get("source", name)), None)
            if source is None:
                raise ConfigError("%s: no source" % where)
            print("FitConfig: Loading signal: %s" % name)
            self.signals.append(SignalConfig(name, _need(cfg, "title", where), _need(cfg, "filename", where), dataset,
                                             source, nexpected, self.sample_fields, self.cuts, systs, base))

        # data, per dataset id a list of files = one per experiment (:259-296)
        self.data = {}
        for ds_name, rows in root.get("data", {}).items():
            dataset = int(ds_name)
            cc = list(self.observables) + list(self.cuts)       # observables act as cuts: clip to the PDF domain
            for drow in rows:
                print("FitConfig: Loading data: %s" % drow["filename"])
                d = SignalConfig(drow.get("title", ""), drow.get("title", ""), drow["filename"], dataset, None, -1.0,
                                 self.sample_fields, cc, [], base)
                self.data.setdefault(dataset, []).append(d)

    # ------------------------------------------------------------------------------------
    def data_events(self, i):
        """The data of experiment i as rows of nobservables+1 floats (EvalHist::GetSamples over
        every dataset, src/sxmc.cpp:72-81, src/pdfz.h:542-556)."""
        nobs = len(self.observables)
        parts = []
        for dataset in sorted(self.data):
            d = self.data[dataset][i]
            rows = np.empty((d.samples.shape[0], nobs + 1), dtype=np.float32)
            rows[:, :nobs] = d.samples[:, :nobs]
            rows[:, nobs] = np.float32(dataset)
            parts.append(rows)
        return np.ascontiguousarray(np.concatenate(parts, axis=0)) if parts else np.zeros((0, nobs + 1), np.float32)

    def to_fitspec(self, data=None):
        """-> workloads.FitSpec (what mcmc.build_fit / ensemble.ensemble consume)."""
        nobs = len(self.observables)
        systs = [W.SystSpec(type=y.type, obs=y.observable_field_index, extra=y.truth_field_index,
                            means=list(y.means), sigmas=list(y.sigmas), fixed=y.fixed) for y in self.systematics]
        names = [y.name for y in self.systematics]
        signals = [W.SignalSpec(samples=s.samples, nexpected=s.nexpected, source=s.source.index, dataset=s.dataset,
                                systematics=[names.index(n) for n in s.systematic_names], n_mc=s.n_mc)
                   for s in self.signals]
        if data is None:
            data = np.zeros((0, nobs + 1), dtype=np.float32)
        fit = W.FitSpec(name="config", lower=[o.lower for o in self.observables],
                        upper=[o.upper for o in self.observables], nbins=[o.bins for o in self.observables],
                        source_means=[s.mean for s in self.sources], source_sigmas=[s.sigma for s in self.sources],
                        source_fixed=[s.fixed for s in self.sources], systematics=systs, signals=signals,
                        data=data, nsteps=self.nsteps, burnin_fraction=self.burnin_fraction)
        fit.names = {"sources": [s.name for s in self.sources], "systematics": names,
                     "signals": [s.name for s in self.signals], "observables": [o.name for o in self.observables]}
        return fit

    def print(self):
        """FitConfig::print (src/config.cpp:300-340), abridged."""
        print("Fit:\n  Fake experiments: %d\n  MCMC steps: %d\n  Burn-in fraction: %g\n  Signal name: %s\n"
              "  Output prefix: %s\n  Confidence level: %g" % (self.nexperiments, self.nsteps, self.burnin_fraction,
                                                               self.signal_name, self.output_prefix, self.confidence))
        print("Observables: " + ", ".join("%s [%g, %g) x%d" % (o.name, o.lower, o.upper, o.bins)
                                          for o in self.observables))
        print("Cuts: " + ", ".join("%s [%g, %g]" % (o.name, o.lower, o.upper) for o in self.cuts))
        print("Systematics: " + ", ".join("%s%s" % (y.name, " (fixed)" if y.fixed else "") for y in self.systematics))
        for s in self.signals:
            print("  %s: nexpected %g, n_mc %d, %d samples after cuts, source %s, dataset %d"
                  % (s.name, s.nexpected, s.n_mc, s.samples.shape[0], s.source.name, s.dataset))
