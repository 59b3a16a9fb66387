"""``python -m sxmc_b200 fit_configuration.json output_path`` -- the command line of the
reference's ``bin/sxmc`` (`src/sxmc.cpp:145-235`): load the JSON configuration, run the
ensemble of fits (`:44-141`), print the upper limits and their median.

Differences forced by the missing ROOT: sample files are .npz/.npy/.csv/.f32 tables
(``sxmc_b200.config.read_float_table``); each experiment's likelihood space is written as
``<output_path>/<output_prefix>_<i>.npy`` (float32 rows, one column per parameter plus
"likelihood") with the column names in ``<output_prefix>_columns.json`` instead of a TNtuple in
a .root file; no plots; the projection estimator is a host restatement (ensemble.projection_interval).

Under ``torch.distributed.run`` (one process per GPU) experiment ``i`` runs on rank
``i mod N`` and rank 0 prints the gathered limits."""
import json
import os
import sys

import numpy as np


def main(argv=None):
    argv = sys.argv if argv is None else argv
    if len(argv) != 3:
        sys.stderr.write("Usage: %s fit_configuration.json output_path\n" % argv[0])
        return 1
    from . import _lib, config, ensemble, mcmc, sharding

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _lib.check(_lib.lib().sxmc_init(local))      # fails loudly without a CUDA device: there is no CPU path
    say = print if rank == 0 else (lambda *a, **k: None)

    output_path = argv[2] if argv[2].endswith("/") else argv[2] + "/"
    if os.path.exists(output_path) and not os.path.isdir(output_path):
        sys.stderr.write("Output path exists and is not a directory\nError with output path.\n")
        return 2
    os.makedirs(output_path, exist_ok=True)

    say("sxmc: Loading configuration...")
    say("sxmc: Configuration: %s" % argv[1])
    fc = config.FitConfig(argv[1])
    if rank == 0:
        fc.print()
    if fc.samples:
        raise config.ConfigError("fit.samples (re-reading a saved .root likelihood space) needs ROOT")

    say("sxmc: Running ensemble...")
    mine = sharding.experiments_for_rank(fc.nexperiments, rank, world)
    limits = []
    columns = None
    if not fc.data:
        # fake data sets drawn on the device, the fits of up to 148 experiments in one launch
        fit = fc.to_fitspec()
        r = ensemble.ensemble(fit, fc.nexperiments, signal_name=fc.signal_name, confidence=fc.confidence,
                              seed=fc.seed, rank=rank, world=world, keep_chains=True,
                              debug_mode=fc.debug_mode)
        limits = r["limits"]
        columns = fit.names["sources"] + ["%s_%d" % (y.name, j) for y in fc.systematics for j in range(y.npars)]
        if fc.error_type == "projection":          # ERROR_PROJECTION: host estimator on the chain rows
            limits = []
            for rows in r["chains"]:
                best = ensemble.projection_intervals(rows, columns + ["likelihood"], fc.confidence)
                limits.append(best[fc.signal_name].upper if fc.signal_name in best else float("nan"))
        for i, rows, n in zip(r["experiments"], r["chains"], r["observed"]):
            print("Experiment %d / %d: %d fake events, %d chain rows, limit %g"
                  % (i + 1, fc.nexperiments, int(np.sum(n)), rows.shape[0], limits[r["experiments"].index(i)]))
            np.save("%s%s_%d.npy" % (output_path, fc.output_prefix, i), rows)
        say("ensemble: %s mode, %.2f s" % (r["mode"], r["seconds"]))
    else:
        for i in mine:
            print("Experiment %d / %d" % (i + 1, fc.nexperiments))
            data = fc.data_events(i)
            fit = fc.to_fitspec(data)
            hists, chain = mcmc.build_fit(fit)
            chain.set_seed(fc.seed, i)
            ls = chain(data, fc.nsteps, fc.burnin_fraction, fc.debug_mode)
            best_fit, lmin = ensemble.contour_intervals(chain, fc.confidence)
            columns = chain.parameter_names[:-1]
            if fc.error_type == "projection":
                best_fit = ensemble.projection_intervals(ls.get_samples(), chain.parameter_names, fc.confidence)
            print("-- Best fit --")
            for name in columns:
                iv = best_fit[name]
                print(" %s: %g -%g +%g" % (name, iv.point_estimate, iv.point_estimate - iv.lower,
                                           iv.upper - iv.point_estimate))
            if fc.signal_name in best_fit:
                bfi = best_fit[fc.signal_name]
                print("ensemble: lower = %g, upper = %g" % (bfi.lower, bfi.upper))
                limits.append(bfi.upper)
            else:
                limits.append(float("nan"))
            np.save("%s%s_%d.npy" % (output_path, fc.output_prefix, i), ls.get_samples())
            chain.close()
            for h in hists:
                h.close()
    full = sharding.gather_limits(mine, limits, fc.nexperiments)
    if rank == 0:
        with open("%s%s_columns.json" % (output_path, fc.output_prefix), "w") as f:
            json.dump((columns or []) + ["likelihood"], f)
        have = [float(x) for x in full if not np.isnan(x)] if fc.signal_name else []
        print("sxmc: Upper limits: " + "".join("%g, " % x for x in have))
        print("sxmc: Median upper limit: %g" % (sharding.median(have) if have else -1))
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
