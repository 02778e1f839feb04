#!/usr/bin/env python
"""Re-fit the alpha-beta link model from measured sweeps (BASELINE.json: "the alpha-beta cost model is
re-fit from measured NVLink latency and bandwidth on the box").

  python bench/fit_cost_model.py profiles/raw/sweep_n8.jsonl profiles/raw/sweep_n2_tma.jsonl ... \
         [--out profiles/link_model_b200.json] [--report profiles/cost_model_validation.md]

Every flextree row of a sweep file is one observation (algorithm, widths, N, bytes, registered, us). The
native predictor (csrc/planner/cost_model.cc through ft_predict_us) is evaluated with trial parameters and
scipy least-squares minimises the log-time residuals, so small and large messages weigh the same. The
report lists predicted vs measured time for every algorithm and width vector - including all
factorizations of 8 (8, 2x4, 4x2, 2x2x2: BASELINE config #3) - and whether the model's arg-min is the
measured arg-min. The reference's cost model (cost_model/CostModel.h) has hand-set constants and is
never compared with measurements; its formula is kept only as an oracle (ft.legacy_cost).
"""
import argparse
import json
import math
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flextree as ft  # noqa: E402

PARAMS = ["alpha_launch_us", "alpha_sync_us", "alpha_peer_us", "alpha_ll_us", "alpha_nvls_us", "beta_p2p_GBps",
          "beta_p2p_1peer_GBps", "beta_nvls_GBps", "beta_ll_GBps", "beta_ll_self_GBps", "half_bw_bytes"]
LOWER = [2.0, 0.5, 0.0, 0.5, 0.0, 100, 100, 100, 50, 20, 1024]
UPPER = [15.0, 10.0, 3.0, 10.0, 20.0, 900, 900, 900, 2000, 2000, 4 << 20]


def load(paths):
    obs = []
    for p in paths:
        for line in open(p):
            line = line.strip()
            if not line.startswith("{"):
                continue
            r = json.loads(line)
            if r.get("impl") != "flextree" or "error" in r or r.get("err"):
                continue
            if r.get("requested") == "auto":
                continue  # duplicates of a forced row
            algo = ft.ALGOS[r["algo"]]
            obs.append({"algo": algo, "name": r["algo"], "widths": r["widths"], "n": r["n"], "bytes": r["bytes"],
                        "registered": r.get("mode", "registered") == "registered", "us": r["us"]})
    return obs


def predict(model, o):
    w = o["widths"] if o["algo"] == ft.ALGO_STAGES else [o["n"]]
    return model.predict_us(o["algo"], w, o["n"], o["bytes"], o["registered"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("files", nargs="+")
    ap.add_argument("--out", default="")
    ap.add_argument("--report", default="")
    ap.add_argument("--no-fit", action="store_true", help="only validate the current/built-in constants")
    ap.add_argument("--exclude", default="", help="comma list of N:algo pairs to leave out, e.g. 2:nvls,4:stages")
    args = ap.parse_args()
    drop = {tuple(x.split(":")) for x in args.exclude.split(",") if x}
    obs = [o for o in load(args.files) if (str(o["n"]), o["name"]) not in drop and o["name"] != "ring"]
    model = ft.LinkModel().set(has_nvls=1)
    x0 = np.array([model.dump()[k] for k in PARAMS], dtype=float)

    def resid(x):
        model.set(**{k: float(v) for k, v in zip(PARAMS, x)})
        return np.array([math.log(max(predict(model, o), 1e-3) / o["us"]) for o in obs])

    before = float(np.sqrt(np.mean(resid(x0) ** 2)))
    x = x0
    if not args.no_fit:
        from scipy.optimize import least_squares
        sol = least_squares(resid, np.clip(x0, LOWER, UPPER), bounds=(LOWER, UPPER), loss="soft_l1", f_scale=0.2)
        x = sol.x
    after = float(np.sqrt(np.mean(resid(x) ** 2)))
    model.set(**{k: float(v) for k, v in zip(PARAMS, x)})
    fitted = model.dump()
    print(f"observations: {len(obs)}   rms log error: {before:.3f} -> {after:.3f}  (x{math.exp(after):.2f} typical ratio)")
    for k in PARAMS:
        print(f"  {k:22s} {x0[PARAMS.index(k)]:12.4g} -> {fitted[k]:12.4g}")
    if args.out:
        json.dump(fitted, open(args.out, "w"), indent=1)
    if args.report:
        lines = ["# Link-model validation: predicted vs measured", "",
                 f"{len(obs)} observations from {', '.join(os.path.basename(f) for f in args.files)}; "
                 f"rms log error {after:.3f} (typical ratio x{math.exp(after):.2f}).", "",
                 "Constants: `" + json.dumps({k: round(fitted[k], 4) for k in PARAMS}) + "`", ""]
        for n in sorted({o["n"] for o in obs}):
            lines += [f"## N = {n}", "", "| bytes | algorithm | widths | measured us | predicted us | ratio |", "|---:|---|---|---:|---:|---:|"]
            sizes = sorted({o["bytes"] for o in obs if o["n"] == n})
            agree = 0
            for sz in sizes:
                rows = [o for o in obs if o["n"] == n and o["bytes"] == sz and o["registered"]]
                for o in rows:
                    p = predict(model, o)
                    lines.append(f"| {sz} | {o['name']} | {'x'.join(map(str, o['widths']))} | {o['us']:.1f} | {p:.1f} | {p / o['us']:.2f} |")
                if rows:
                    best_meas = min(rows, key=lambda o: o["us"])
                    best_pred = min(rows, key=lambda o: predict(model, o))
                    slack = best_pred["us"] / best_meas["us"]
                    agree += slack <= 1.10
                    lines.append(f"| {sz} | **model picks** {best_pred['name']} {'x'.join(map(str, best_pred['widths']))} | "
                                 f"**measured best** {best_meas['name']} {'x'.join(map(str, best_meas['widths']))} | | | "
                                 f"pick is {slack:.2f}x the best |")
            lines += ["", f"Model's pick within 10 % of the measured best at {agree} of {len(sizes)} sizes.", ""]
        open(args.report, "w").write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
