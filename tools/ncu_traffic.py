#!/usr/bin/env python
"""The committed ncu recipe behind bench.py's `roofline.traffic` (profiles/traffic.json).

Run on the GPU box from the repo root (one GPU; ncu replays each kernel ~40 times):

    python tools/ncu_traffic.py capture  ->  gpurun_out/glm_r02.ncu-rep + gpurun_out/glm_r02_raw.csv
    python tools/ncu_traffic.py parse    ->  profiles/traffic.json, profiles/r02_glm_ncu_summary.txt

`capture` first runs the workload WITHOUT ncu (must exit 0), then under
`ncu --set full --clock-control none --import-source on -k regex:glm_kernel -c 4`.
`parse` needs only the csv, so it also runs in the build container.  traffic.json records the
commit of the tree that was profiled; bench.py passes that through as `roofline.traffic_source`.
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REP = os.path.join(ROOT, "gpurun_out", "glm_r02")
RAW = REP + "_raw.csv"
WORK = os.path.join(ROOT, "gpurun_out", "glm_r02_workload.json")
METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed.sum", "smsp__issue_active.avg.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "launch__registers_per_thread",
           "launch__shared_mem_per_block_dynamic", "smsp__average_warp_latency_issue_stalled_barrier.ratio",
           "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"]


def workload():
    """One adaptation launch + sampling launches of the headline kernel; writes the gradient
    evaluations of every glm launch so that `parse` can divide."""
    import numpy as np

    import __graft_entry__ as g

    g.build()
    import bench
    from numpyro_b200 import models
    from numpyro_b200.infer import MCMC, NUTS
    from numpyro_b200.random import PRNGKey

    chains = int(os.environ.get("NCU_CHAINS", "1"))
    X, y = bench.make_covtype_shape()
    k = NUTS(models.logistic_regression)
    m = MCMC(k, num_warmup=100, num_samples=50, num_chains=chains, chain_method="vectorized")
    launches = []
    m.warmup(PRNGKey(1), X, y, extra_fields=("num_steps",), collect_warmup=True)
    g0 = np.asarray(k.grad_evals())
    launches.append({"what": "init + warm-up (100 transitions)", "grad_evals": int(g0.max()), "per_chain": g0.tolist()})
    for _ in range(2):
        m.run(m.post_warmup_state.rng_key, X, y, extra_fields=("num_steps",))
        ns = m.get_extra_fields(group_by_chain=True)["num_steps"].sum(1)
        launches.append({"what": "50 sampling transitions", "grad_evals": int(ns.max()), "per_chain": ns.tolist()})
    with open(WORK, "w") as f:
        json.dump({"chains": chains, "launches": launches}, f)
    print(json.dumps(launches))


def capture():
    os.makedirs(os.path.dirname(REP), exist_ok=True)
    subprocess.check_call([sys.executable, __file__, "workload"])  # must exit 0 without ncu first
    kernel = "regex:glm_mc_kernel" if int(os.environ.get("NCU_CHAINS", "1")) > 1 else "regex:glm_kernel"
    subprocess.check_call(["ncu", "--set", "full", "--clock-control", "none", "--import-source", "on",
                           "-k", kernel, "-c", "4", "-f", "-o", REP, sys.executable, __file__, "workload"],
                          stdout=open(REP + "_ncu.log", "w"), stderr=subprocess.STDOUT)
    with open(RAW, "w") as f:
        subprocess.check_call(["ncu", "-i", REP + ".ncu-rep", "--page", "raw", "--csv"], stdout=f)
    print("wrote", RAW)


def parse():
    rows = list(csv.reader(open(RAW)))
    head = rows[0]
    data = [r for r in rows[2:] if len(r) == len(head)]
    work = json.load(open(WORK))
    name = "_mc" if work["chains"] > 1 else ""
    col = {h: i for i, h in enumerate(head)}
    out = [f"# ncu --set full --clock-control none (tools/ncu_traffic.py), chains={work['chains']}"]
    per = []
    unit_scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    # the launches of the capture and of the workload list are aligned from the END (the capture
    # may start with the short init launch, which the list folds into the warm-up entry)
    off = len(work["launches"]) - len(data)
    for li, r in enumerate(data):
        wi = li + off
        w = work["launches"][wi] if 0 <= wi < len(work["launches"]) else None
        out.append(f"launch {li}: {r[col['Kernel Name']][:40]} {w}")
        vals = {}
        for m in METRICS:
            if m in col:
                vals[m] = r[col[m]]
                out.append(f"  {m} = {r[col[m]]} {rows[1][col[m]]}")
        if w and "dram__bytes_read.sum" in vals:
            byt = sum(float(vals[m].replace(",", "")) * unit_scale.get(rows[1][col[m]], 1)
                      for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            per.append((li, byt / max(1, w["grad_evals"]), float(vals.get("lts__t_sector_hit_rate.pct", "nan")),
                        w["what"]))
            out.append(f"  => DRAM bytes per pass over X: {per[-1][1]:.0f}")
    commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    with open(os.path.join(ROOT, "profiles", f"r02_glm{name}_ncu_summary.txt"), "w") as f:
        f.write("\n".join(out) + "\n")
    if per and work["chains"] == 1:
        li, b, hit, _ = per[-1]  # a sampling launch
        tj = {"glm_kernel_bytes_per_grad_eval": b, "algorithmic_bytes_per_grad_eval": 128403652,
              "lts_sector_hit_rate_pct": hit, "commit": commit, "captured": "round 2",
              "recipe": "python tools/ncu_traffic.py capture && python tools/ncu_traffic.py parse "
                        "((dram__bytes_read.sum + dram__bytes_write.sum) / grad evals of the launch)"}
        with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
            json.dump(tj, f, indent=1)
        print(json.dumps(tj))


if __name__ == "__main__":
    {"workload": workload, "capture": capture, "parse": parse}[sys.argv[1]]()
