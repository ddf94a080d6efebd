#!/usr/bin/env python
"""ncu recipe for the chain-batched kernels (eight schools: chain_kernel<0>, stochastic
volatility: chain_kernel<1|2>), the counterpart of tools/ncu_traffic.py:

    python tools/ncu_chain.py capture eight_schools|sv   (GPU box, from the repo root)
    python tools/ncu_chain.py parse   eight_schools|sv   -> profiles/r02_<name>_ncu_summary.txt,
                                                            profiles/<name>_ncu.json (read by bench.py)
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
KEYS = ["gpu__time_duration.sum", "sm__inst_executed.sum", "smsp__issue_active.avg.pct", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"]


def paths(name):
    base = os.path.join(ROOT, "gpurun_out", f"chain_{name}_r02")
    return base, base + "_raw.csv", base + "_workload.json"


def workload(name):
    import numpy as np

    import __graft_entry__ as g

    g.build()
    import bench
    from numpyro_b200.infer import MCMC, NUTS
    from numpyro_b200.random import PRNGKey, split

    model, margs, mkw, kkw, chains, desc = bench.make_workload(name, 0)
    k = NUTS(model, **kkw)
    # ncu replays every launch ~40 times: the CTA-per-chain SV launches are seconds long
    nw, ns_ = (20, 10) if name == "sv" else (100, 30)
    m = MCMC(k, num_warmup=nw, num_samples=ns_, num_chains=chains, chain_method="vectorized")
    m.warmup(split(PRNGKey(0), chains), *margs, **mkw)
    g0 = int(np.asarray(k.grad_evals()).sum())
    m.run(m.post_warmup_state.rng_key, *margs, extra_fields=("num_steps",), **mkw)
    ns = int(m.get_extra_fields()["num_steps"].sum())
    json.dump({"name": name, "chains": chains, "launches": [{"what": f"init + {nw} warm-up transitions", "grad_evals": g0},
                                                            {"what": f"{ns_} sampling transitions", "grad_evals": ns}]},
              open(paths(name)[2], "w"))
    print(g0, ns)


def capture(name):
    base, raw, _ = paths(name)
    os.makedirs(os.path.dirname(base), exist_ok=True)
    subprocess.check_call([sys.executable, __file__, "workload", name])
    subprocess.check_call(["ncu", "--set", "full", "--clock-control", "none", "--import-source", "on",
                           "-k", "regex:chain_kernel", "-c", "3", "-f", "-o", base, sys.executable, __file__,
                           "workload", name], stdout=open(base + "_ncu.log", "w"), stderr=subprocess.STDOUT)
    with open(raw, "w") as f:
        subprocess.check_call(["ncu", "-i", base + ".ncu-rep", "--page", "raw", "--csv"], stdout=f)
    os.remove(base + ".ncu-rep")
    print("wrote", raw)


def parse(name):
    base, raw, wl = paths(name)
    rows = list(csv.reader(open(raw)))
    head = rows[0]
    col = {h: i for i, h in enumerate(head)}
    data = [r for r in rows[2:] if len(r) == len(head)]
    work = json.load(open(wl))
    off = len(work["launches"]) - len(data)
    out = [f"# ncu --set full --clock-control none (tools/ncu_chain.py {name}), {work['chains']} chains"]
    last = None
    for li, r in enumerate(data):
        wi = li + off
        w = work["launches"][wi] if 0 <= wi < len(work["launches"]) else None
        out.append(f"launch {li}: {r[col['Kernel Name']][:50]} {w}")
        vals = {}
        for k in KEYS:
            if k in col:
                vals[k] = r[col[k]]
                out.append(f"  {k} = {r[col[k]]} {rows[1][col[k]]}")
        ok = "nan" not in str(vals.get("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "nan")).lower()
        if w and ok:  # a launch whose counters ncu failed to collect (nan) is listed, not used
            last = (vals, w)
    commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    open(os.path.join(ROOT, "profiles", f"r02_{name}_ncu_summary.txt"), "w").write("\n".join(out) + "\n")
    if last:
        vals, w = last
        f = lambda k: float(vals[k].replace(",", "")) if k in vals else None  # noqa: E731
        js = {"kernel": "chain_kernel", "workload": name, "chains": work["chains"], "launch": w,
              "issue_active_pct": f("sm__issue_active.avg.pct_of_peak_sustained_elapsed"),
              "active_threads_per_inst": f("smsp__thread_inst_executed_per_inst_executed.ratio"),
              "warps_active_pct": f("sm__warps_active.avg.pct_of_peak_sustained_active"),
              "dram_bytes_per_grad_eval": ((f("dram__bytes_read.sum") or 0) + (f("dram__bytes_write.sum") or 0)) * 1e9
              / max(1, w["grad_evals"]) if rows[1][col["dram__bytes_read.sum"]] == "Gbyte" else None,
              "fma_pipe_pct": f("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
              "xu_pipe_pct": f("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
              "commit": commit, "recipe": f"python tools/ncu_chain.py capture {name} && python tools/ncu_chain.py parse {name}"}
        json.dump(js, open(os.path.join(ROOT, "profiles", f"{name}_ncu.json"), "w"), indent=1)
        print(json.dumps(js))


if __name__ == "__main__":
    {"workload": workload, "capture": capture, "parse": parse}[sys.argv[1]](sys.argv[2])
