"""JSON configuration + sample reader + command line (src/config.cpp:19-297,
src/signal.cpp:11-108, src/sxmc.cpp:145-235).  CPU part: the parse follows the reference's
ordering and default rules on a config written in its schema (with the shipped example's
missing comma and a comment, which jsoncpp tolerates).  GPU part: `python -m sxmc_b200` runs
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
                       "mean": [0.0], "sigma": [1e-2]}
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
