"""Generates tests/golden/ref_fixtures.npz from the UNMODIFIED reference sources compiled
into oracle/_ref/libsxmc_ref.so (run in the build container, where /root/reference exists):

    python tests/golden/make_ref_fixtures.py

The fixtures hold, for scaled-down A/B/C/D workloads at a seeded off-nominal parameter
vector: bins, norms, lut, NLL and event sum as computed by the reference's own
bin_samples / eval_pdf / nll_event_chunks / nll_total; plus one reference chain driven by a
seeded random tape.  They travel to the GPU box, where /root/reference does not exist."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from oracle import ref  # noqa: E402
from sxmc_b200 import workloads as W  # noqa: E402
from tests.test_oracle import SMALL, _pars  # noqa: E402
from tests import cases  # noqa: E402


def main():
    oracle.build()
    out = {}
    for name, scale in SMALL:
        fit = W.WORKLOADS[name](scale)
        pars = _pars(fit, 11)
        bins, norms, lut = ref.eval_all(fit, pars)
        nll, ev = ref.nll(fit, lut, pars, norms)
        out[name + "_pars"] = pars
        out[name + "_bins"] = bins
        out[name + "_norms"] = norms
        out[name + "_lut"] = lut
        out[name + "_nll"] = np.float64(nll)
        out[name + "_evsum"] = np.float64(ev)
    # a tape-driven reference chain (config B, tiny) for the chain parity test
    fit = W.config_B(0.005)
    nsteps = 400
    rng = np.random.default_rng(123)
    normals = rng.normal(size=(nsteps + 1) * fit.nparameters)
    uniforms = rng.random(nsteps)
    r = ref.mcmc(fit, nsteps, 0.1, normals, uniforms)
    out["chainB_normals"] = normals
    out["chainB_uniforms"] = uniforms
    out["chainB_chain"] = r["chain"]
    # CTSCALE and 2-, 3-, 4-coefficient series (src/pdfz.cpp:310-325), reference output
    for name, make in cases.POLY_CASES.items():
        fit = make()
        pars = cases.off_nominal(fit, 31)
        bins, norms, lut = ref.eval_all(fit, pars)
        nll, ev = ref.nll(fit, lut, pars, norms)
        out[name + "_pars"] = pars
        out[name + "_bins"] = bins
        out[name + "_norms"] = norms
        out[name + "_lut"] = lut
        out[name + "_nll"] = np.float64(nll)
        out[name + "_evsum"] = np.float64(ev)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_fixtures.npz"), **out)
    print("wrote", {k: v.shape for k, v in out.items()})
    if "--full" in sys.argv:
        # full-size C, D and the 200 000-bin D: SHA-256 of the reference's bins and lut, norms, NLL
        import json
        full = {}
        for name, make in cases.FULL_CASES.items():
            fit = make()
            pars = cases.off_nominal(fit, 41)
            bins, norms, lut = ref.eval_all(fit, pars)
            nll, ev = ref.nll(fit, lut, pars, norms)
            full[name] = {"pars": [float(v) for v in pars], "bins_sha256": cases.digest(bins),
                          "lut_sha256": cases.digest(lut), "norms": [int(v) for v in norms],
                          "nll": float(nll), "evsum": float(ev), "bins_sum": int(bins.sum()),
                          "nonzero_bins": int((bins > 0).sum())}
            print(name, full[name]["bins_sha256"][:16], full[name]["nll"])
            del fit, bins, lut
        with open(os.path.join(ROOT, "tests", "golden", "ref_full_size.json"), "w") as f:
            json.dump(full, f, indent=1)


if __name__ == "__main__":
    main()
