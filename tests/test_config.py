"""JSON configuration + sample reader + command line (src/config.cpp:19-297,
src/signal.cpp:11-108, src/sxmc.cpp:145-235).  CPU part: the parse follows the reference's
ordering and default rules on a config written in its schema (with a comment, which jsoncpp
tolerates), and equals the parse of the reference's OWN FitConfig (compiled into oracle/_ref,
fed the same .npy tables) key for key.  GPU part: `python -m sxmc_b200` runs
the ensemble and writes what it promises."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CONFIG = """{
  // same schema as config/example.json
  "fit": {
    "nexperiments": 3, "nsteps": 3000, "burnin_fraction": 0.2, "signal_name": "signal1",
    "output_prefix": "example", "seed": 7, "confidence": 0.9,
    "signals": ["signal2", "signal1"],
    "observables": ["energy"],
    "cuts": ["radius"]
  },
  "pdfs": {
    "observables": {
      "energy": {"title": "Energy (MeV)", "units": "MeV", "field": "energy", "bins": 10, "min": 5.0, "max": 15.0},
      "radius": {"title": "Radius (m)", "units": "m", "field": "radius", "bins": 10, "min": 0.0, "max": 10.0}
    },
    "systematics": {
      "energy_scale": {"title": "Energy scale", "type": "scale", "observable_field": "energy",
                       "mean": [0.0], "sigma": [1e-2]},
      "energy_resolution": {"title": "Energy resolution", "type": "resolution_scale", "observable_field": "energy",
                            "truth_field": "mc_energy", "mean": [0.0], "sigma": [1e-3]}
    }
  },
  "signals": {
    "signal2": {"title": "Signal 2", "filename": "signal2.npz", "systematics": ["energy_scale", "energy_resolution"],
                "dataset": 0, "scale": 20.0},
    "signal1": {"title": "Signal 1", "filename": "signal1.npz", "dataset": 0,
                "systematics": ["energy_resolution", "energy_scale"], "scale": 10.0, "sigma": 0.05}
  }
}
"""


def write_inputs(d, n1=20000, n2=40000):
    rng = np.random.default_rng(5)
    for name, n, mu in (("signal1", n1, 8.0), ("signal2", n2, 11.0)):
        e = rng.normal(mu, 1.5, n).astype(np.float32)
        np.savez(os.path.join(d, name + ".npz"), radius=(12 * rng.random(n) ** (1 / 3)).astype(np.float32),
                 mc_energy=(e - rng.normal(0, 0.3, n)).astype(np.float32), energy=e,
                 unused=np.zeros(n, np.float32))
    path = os.path.join(d, "fit.json")
    with open(path, "w") as f:
        f.write(CONFIG)
    return path


def test_fitconfig_follows_reference_ordering_and_defaults(tmp_path):
    from sxmc_b200 import config, workloads as W
    fc = config.FitConfig(write_inputs(str(tmp_path)))
    assert (fc.nexperiments, fc.nsteps, fc.signal_name, fc.output_prefix, fc.seed) == (3, 3000, "signal1", "example", 7)
    assert fc.burnin_fraction == float(np.float32(0.2)) and fc.confidence == float(np.float32(0.9))
    assert fc.error_type == "contour" and fc.debug_mode is False and fc.plots is True      # defaults, :52-68
    # sources / systematics are collected over the signals OBJECT in sorted key order (jsoncpp): signal1 first,
    # and within it in the order of its own "systematics" list
    assert [s.name for s in fc.sources] == ["signal1", "signal2"] and [s.index for s in fc.sources] == [0, 1]
    assert fc.sources[0].sigma == float(np.float32(0.05)) and fc.sources[1].sigma == 0.0 and fc.sources[1].mean == 1.0
    assert [y.name for y in fc.systematics] == ["energy_resolution", "energy_scale"]
    assert [y.pidx for y in fc.systematics] == [[0], [1]]
    # sample columns: observables' fields, truth fields, DATASET (:153-194)
    assert fc.sample_fields == ["energy", "mc_energy", "DATASET"]
    assert fc.systematics[0].type == W.RESOLUTION_SCALE and fc.systematics[0].truth_field_index == 1
    assert fc.systematics[1].observable_field_index == 0
    # fit.signals order is kept for the signals; each keeps ITS systematics order
    assert [s.name for s in fc.signals] == ["signal2", "signal1"]
    assert fc.signals[0].systematic_names == ["energy_scale", "energy_resolution"]
    z = np.load(tmp_path / "signal2.npz")
    keep = (z["radius"] >= 0.0) & (z["radius"] <= 10.0)
    s2 = fc.signals[0]
    assert s2.n_mc == 40000 and s2.samples.shape == (int(keep.sum()), 3)           # n_mc counts rows before cuts
    assert np.array_equal(s2.samples[:, 0], z["energy"][keep]) and np.array_equal(s2.samples[:, 1], z["mc_energy"][keep])
    assert np.all(s2.samples[:, 2] == 0)
    assert s2.nexpected == pytest.approx(40000 / 20.0) and fc.signals[1].nexpected == pytest.approx(20000 / 10.0)
    fit = fc.to_fitspec()
    assert fit.nparameters == 4 and fit.nbins == [10] and fit.signals[0].source == 1
    assert fit.signals[0].systematics == [1, 0] and fit.signals[1].systematics == [0, 1]
    assert fit.names["sources"] == ["signal1", "signal2"]


def test_sample_readers_agree(tmp_path):
    from sxmc_b200 import config
    rng = np.random.default_rng(1)
    a = rng.random((50, 3)).astype(np.float32)
    fields = ["x", "y", "z"]
    np.savez(tmp_path / "t.npz", data=a, fields=np.asarray(fields))
    np.save(tmp_path / "t.npy", a)
    (tmp_path / "t.npy.fields").write_text("\n".join(fields))
    a.tofile(tmp_path / "t.f32")
    (tmp_path / "t.f32.fields").write_text("\n".join(fields))
    np.savetxt(tmp_path / "t.csv", a, delimiter=",", header=",".join(fields), comments="", fmt="%.9g")
    rec = np.zeros(50, dtype=[(f, "<f4") for f in fields])
    for j, f in enumerate(fields):
        rec[f] = a[:, j]
    np.save(tmp_path / "r.npy", rec)
    for name in ("t.npz", "t.npy", "t.f32", "t.csv", "r.npy"):
        data, got = config.read_float_table(str(tmp_path / name))
        assert got == fields and np.array_equal(data, a), name
    with pytest.raises(config.ConfigError, match="ROOT"):
        config.read_float_table("/data/signal1.root")


def test_config_errors_name_the_key(tmp_path):
    from sxmc_b200 import config
    path = write_inputs(str(tmp_path))
    cfg = config.load_json(path)
    del cfg["signals"]["signal1"]["scale"]
    bad = tmp_path / "bad.json"
    bad.write_text(json.dumps(cfg))
    with pytest.raises(config.ConfigError, match="rate.*scale"):
        config.FitConfig(str(bad))
    cfg = config.load_json(path)
    cfg["pdfs"]["systematics"]["energy_scale"]["type"] = "wobble"
    bad.write_text(json.dumps(cfg))
    with pytest.raises(config.ConfigError, match="Unknown systematic type"):
        config.FitConfig(str(bad))


# a second configuration: three sources (one shared by two signals), three systematics (one with a
# two-coefficient series), 'rate' and 'scale' signals, two cuts of which two name the SAME field
# (the reference keeps the last one, src/signal.cpp:59-67), a fixed source, a fixed systematic
CONFIG3 = """{
  "fit": {
    "nexperiments": 2, "nsteps": 500, "signal_name": "b8", "error_type": "projection", "debug_mode": true,
    "signals": ["tl208", "b8", "bi214"],
    "observables": ["radius", "energy"],
    "cuts": ["fiducial", "fiducial_tight"]
  },
  "pdfs": {
    "observables": {
      "energy": {"title": "E", "units": "MeV", "field": "energy", "bins": 12, "min": 5.1, "max": 14.7},
      "radius": {"title": "R", "units": "m", "field": "radius", "bins": 4, "min": 0.0, "max": 11.3},
      "fiducial": {"title": "z", "units": "m", "field": "z", "bins": 1, "min": -0.9, "max": 0.9},
      "fiducial_tight": {"title": "z", "units": "m", "field": "z", "bins": 1, "min": -0.7, "max": 0.3}
    },
    "systematics": {
      "escale": {"title": "s", "type": "scale", "observable_field": "energy", "mean": [0.001, 0.0002],
                 "sigma": [0.01, 0.001]},
      "eres": {"title": "r", "type": "resolution_scale", "observable_field": "energy", "truth_field": "mc_energy",
               "mean": [0.0], "sigma": [0.02], "fixed": true},
      "rshift": {"title": "d", "type": "shift", "observable_field": "radius", "mean": [0.1]}
    }
  },
  "sources": {
    "thorium": {"mean": 1.1, "sigma": 0.3},
    "solar": {"mean": 0.9, "sigma": 0.0, "fixed": true}
  },
  "signals": {
    "tl208": {"title": "Tl", "filename": "tl208.npy", "dataset": 0, "source": "thorium",
              "systematics": ["rshift", "escale"], "rate": 123.4},
    "bi214": {"title": "Bi", "filename": "bi214.npy", "dataset": 0, "scale": 3.3, "mean": 1.3, "sigma": 0.2,
              "systematics": ["eres"]},
    "b8": {"title": "B8", "filename": "b8.npy", "dataset": 0, "source": "solar", "scale": 7.0,
           "systematics": ["escale", "eres", "rshift"]},
    "unused": {"title": "x", "filename": "none.npy", "dataset": 0, "source": "thorium", "rate": 1.0,
               "systematics": ["eres"]}
  }
}
"""


def write_npy_inputs(d, which):
    """.npy tables + .fields sidecars (what all three readers take) and the JSON file."""
    rng = np.random.default_rng(5)
    if which == "A":
        text = CONFIG.replace(".npz", ".npy")
        for name, n, mu in (("signal1", 3000, 8.0), ("signal2", 5000, 11.0)):
            e = rng.normal(mu, 1.5, n).astype(np.float32)
            table = np.stack([(12 * rng.random(n) ** (1 / 3)).astype(np.float32),
                              (e - rng.normal(0, 0.3, n)).astype(np.float32), e, np.zeros(n, np.float32)], axis=1)
            np.save(os.path.join(d, name + ".npy"), np.ascontiguousarray(table))
            with open(os.path.join(d, name + ".npy.fields"), "w") as f:
                f.write("radius\nmc_energy\nenergy\nunused\n")
    else:
        text = CONFIG3
        for name, n, mu in (("tl208", 3001, 8.0), ("bi214", 1777, 10.0), ("b8", 4242, 12.0)):
            e = rng.normal(mu, 1.5, n).astype(np.float32)
            table = np.stack([rng.uniform(-1, 1, n).astype(np.float32), e,
                              (12 * rng.random(n) ** (1 / 3)).astype(np.float32),
                              (e - rng.normal(0, 0.3, n)).astype(np.float32)], axis=1)
            np.save(os.path.join(d, name + ".npy"), np.ascontiguousarray(table))
            with open(os.path.join(d, name + ".npy.fields"), "w") as f:
                f.write("z\nenergy\nradius\nmc_energy\n")
    for name in ("signal1", "signal2", "tl208", "bi214", "b8"):
        text = text.replace('"%s.npy"' % name, '"%s"' % os.path.join(d, name + ".npy"))
    path = os.path.join(d, "fit_%s.json" % which)
    with open(path, "w") as f:
        f.write(text)
    return path


def _checksum(rows):
    bits = np.ascontiguousarray(rows, dtype=np.float32).ravel().view(np.uint32).astype(np.uint64)
    k = np.arange(1, bits.size + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        return str(int((k * bits).sum(dtype=np.uint64)))


def python_dump(fc):
    """sxmc_b200.config.FitConfig in the schema of oracle/ref_glue.cpp::ref_config_dump."""
    from sxmc_b200 import workloads as W
    return {
        "nexperiments": fc.nexperiments, "nsteps": fc.nsteps, "confidence": fc.confidence,
        "burnin_fraction": fc.burnin_fraction, "debug_mode": fc.debug_mode, "plots": fc.plots, "seed": fc.seed,
        "signal_name": fc.signal_name, "output_prefix": fc.output_prefix, "error_type": fc.error_type,
        "samples": fc.samples, "sample_fields": list(fc.sample_fields), "datasets": sorted(fc.datasets),
        "observables": [{"name": o.name, "field": o.field, "field_index": o.field_index, "bins": o.bins,
                         "lower": o.lower, "upper": o.upper} for o in fc.observables],
        "cuts": [{"name": o.name, "field": o.field, "lower": o.lower, "upper": o.upper} for o in fc.cuts],
        "sources": [{"name": s.name, "index": s.index, "mean": s.mean, "sigma": s.sigma, "fixed": s.fixed}
                    for s in fc.sources],
        "systematics": [{"name": y.name, "type": y.type, "observable_field": y.observable_field,
                         "truth_field": y.truth_field, "observable_field_index": y.observable_field_index,
                         "truth_field_index": y.truth_field_index if y.type == W.RESOLUTION_SCALE else -1,
                         "npars": y.npars, "fixed": y.fixed, "means": list(y.means), "sigmas": list(y.sigmas),
                         "pidx": list(y.pidx)} for y in fc.systematics],
        "signals": [{"name": g.name, "filename": g.filename, "dataset": g.dataset, "source": g.source.name,
                     "source_index": g.source.index, "nexpected": g.nexpected, "n_mc": g.n_mc,
                     "nsample_floats": int(g.samples.size), "sample_checksum": _checksum(g.samples),
                     "systematic_names": list(g.systematic_names)} for g in fc.signals],
    }


def _strip_dirs(dump):
    for g in dump["signals"]:
        g["filename"] = os.path.basename(g["filename"])
    return dump


GOLDEN_DUMPS = os.path.join(ROOT, "tests", "golden", "ref_fitconfig_dumps.json")


@pytest.mark.parametrize("which", ["A", "B3"])
def test_fitconfig_equals_the_reference_fitconfig(tmp_path, which):
    """Parameter order, signal order, sample-column order, cuts, n_mc, nexpected (float rounding of
    1/scale included) and the sample rows themselves: sxmc_b200.config.FitConfig against the
    UNMODIFIED reference FitConfig/Signal (src/config.cpp:19-297, src/signal.cpp:11-109) compiled into
    oracle/_ref and fed the same files, and against its committed output."""
    import oracle
    from sxmc_b200 import config
    path = write_npy_inputs(str(tmp_path), which)
    got = _strip_dirs(python_dump(config.FitConfig(path)))
    with open(GOLDEN_DUMPS) as f:
        gold = json.load(f)[which]
    assert got == gold
    if oracle.have_ref():
        from oracle import ref
        assert _strip_dirs(ref.config_dump(path)) == gold


def test_missing_comma_is_a_parse_error_as_in_the_reference(tmp_path):
    """jsoncpp 0.6 rejects `config/example.json:42-43` as shipped ("Missing ',' or '}' in object
    declaration"); so do the drop-in readers."""
    import oracle
    from sxmc_b200 import config
    path = write_npy_inputs(str(tmp_path), "A")
    text = open(path).read().replace('"sigma": [1e-2]},', '"sigma": [1e-2]}')
    bad = tmp_path / "bad.json"
    bad.write_text(text)
    with pytest.raises(config.ConfigError, match="JSON parse error"):
        config.FitConfig(str(bad))
    if oracle.have_ref():
        from oracle import ref
        with pytest.raises(ref.RefError):
            ref.config_dump(str(bad))


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["A", "B3"])
def test_cxx_fitconfig_equals_the_reference_fitconfig(tmp_path, which):
    """The same comparison for the C++ FitConfig (cxx/src/config.cpp) through `sxmc --dump-config`."""
    path = write_npy_inputs(str(tmp_path), which)
    exe = os.path.join(ROOT, "cxx", "bin", "sxmc")
    r = subprocess.run([exe, "--dump-config", path], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    with open(GOLDEN_DUMPS) as f:
        gold = json.load(f)[which]
    assert _strip_dirs(json.loads(r.stdout)) == gold
    bad = tmp_path / "bad.json"
    bad.write_text(open(path).read().replace('"bins": 10, "min": 5.0, "max": 15.0},', '"bins": 10, "min": 5.0, "max": 15.0}')
                   if which == "A" else open(path).read().replace('"sigma": [0.01, 0.001]},', '"sigma": [0.01, 0.001]}'))
    r = subprocess.run([exe, "--dump-config", str(bad)], capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "Missing ','" in r.stderr


@pytest.mark.gpu
def test_command_line_runs_the_ensemble(tmp_path):
    from sxmc_b200 import config, ensemble
    path = write_inputs(str(tmp_path))
    out = tmp_path / "out"
    r = subprocess.run([sys.executable, "-m", "sxmc_b200", path, str(out)], cwd=ROOT, capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "sxmc: Median upper limit:" in r.stdout
    cols = json.loads((out / "example_columns.json").read_text())
    assert cols == ["signal1", "signal2", "energy_resolution_0", "energy_scale_0", "likelihood"]
    fc = config.FitConfig(path)
    want = ensemble.ensemble(fc.to_fitspec(), 3, signal_name="signal1", confidence=fc.confidence, seed=7,
                             keep_chains=True)
    for i in range(3):
        rows = np.load(out / ("example_%d.npy" % i))
        assert rows.shape == (3000 - 2 * 600, 5)
        assert rows.tobytes() == want["chains"][i].tobytes()
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("sxmc: Upper limits:")][0]
    got = [float(x) for x in line.split(":", 2)[2].split(",") if x.strip()]
    assert np.allclose(got, want["limits"], rtol=1e-5)
