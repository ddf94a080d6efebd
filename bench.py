#!/usr/bin/env python
"""Benchmark of the HMC/NUTS hot path on the headline workload.

Workload (BASELINE.json configs[2]): Bayesian logistic regression NUTS on synthetic
covtype-shape data (N=581,012 rows, D=55 columns, float32 X, uint8 labels), one chain per
GPU.  A "step" is one pass of the hot path over one batch: ``--transitions`` NUTS
transitions (after an untimed adaptation warm-up) executed by ONE persistent kernel
launch; the metric is gradient evaluations (= leapfrogs, ``num_steps``) per second, the
quantity the reference's own notebook divides by
(notebooks/source/logistic_regression.ipynb:309-310).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl own|reference]

N > 1 (torchrun, one rank per GPU): BASELINE configs[2]'s own partitioning -- ONE chain, the
rows of X sharded over the ranks, the D+1 partial sums exchanged per leapfrog over NVLink peer
memory inside the kernel ("strong" scaling); time is the max over ranks.  The line also carries
a `replicas` sub-object: one independent chain per GPU, X replicated, no collective ("weak").
`--sharding chains` makes the replicas the headline instead.

``--impl reference`` times the CPU restatement of the reference's algorithm (oracle/,
C potential with OpenMP over all host cores): jax/numpyro cannot be imported in this image
(no wheel, no network), so the oracle port stands in for NumPyro's JAX-CPU path.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_ROWS, N_COLS = 581012, 55
ALGO_BYTES_PER_GRAD = N_ROWS * N_COLS * 4 + N_ROWS  # X once as fp32 + 1-byte labels
METRIC = "grad_evals_per_sec"
UNIT = "grad-evals/s"


def make_covtype_shape(seed=0, n=N_ROWS, d=N_COLS):
    """standardised N(0,1) features + intercept column (covtype.py:47-49)."""
    rng = np.random.default_rng(seed)
    X0 = rng.standard_normal((n, d - 1), dtype=np.float32)
    X0 = (X0 - X0.mean(0)) / X0.std(0)
    X = np.hstack([X0, np.ones((n, 1), np.float32)]).astype(np.float32)
    theta = (rng.standard_normal(d) * 0.5).astype(np.float32)
    y = (rng.random(n) < 1.0 / (1.0 + np.exp(-(X @ theta)))).astype(np.uint8)
    return np.ascontiguousarray(X), y


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index, period_ms=100):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", str(period_ms)],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1]))
            except ValueError:
                continue
            for nm, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ---------------------------------------------------------------------------- CPU legs
def cpu_nuts_sample(X, y, budget_s, max_transitions, cores_note=True):
    """NUTS transitions of the oracle (reference restatement, C potential on all host cores)
    from the init state with adaptation on, until ``budget_s`` seconds or
    ``max_transitions``.  Returns (grad_evals, seconds, transitions, cores)."""
    from oracle import cpotentials, hmc as ohmc, threefry as T
    from oracle.hmc_util import Config

    cpotentials.build()
    cpotentials.use_all_cores()
    fam = cpotentials.CLogistic(X, y)
    kern = ohmc.HMCKernel(fam, cfg=Config(np.float32, True))
    state = kern.init(T.PRNGKey(1), 1000)
    fam.potential_and_grad(np.zeros(X.shape[1], np.float32))  # touch pages / spin up threads
    g0, t0, n = kern.num_grad_evals, time.perf_counter(), 0
    while n < max_transitions and time.perf_counter() - t0 < budget_s:
        state = kern.sample(state)
        n += 1
    dt = time.perf_counter() - t0
    return kern.num_grad_evals - g0, dt, n, fam.cores


def base_config(args, world, rows):
    """The workload description shared VERBATIM by both arms (the driver compares them)."""
    return {"workload": "covtype-shape logistic NUTS N=581012 D=55 (configs[2])",
            "chains": 1 if rows or world == 1 else world,
            "transitions_per_step": args.transitions,
            "adapt_transitions_untimed": args.adapt, "max_tree_depth": 10,
            "sampler_key": "PRNGKey(1)", "threefry_partitionable": True,
            "parallelism": (f"rows x{world} (one chain, per-leapfrog NVLink exchange)" if rows
                            else f"chains x{world}")}


def run_reference(args):
    """The reference arm: the SAME run as the GPU arm -- `--adapt` untimed adaptation transitions
    from PRNGKey(1), then steps of `--transitions` post-warm-up NUTS transitions at the adapted
    step size and mass matrix -- by the oracle port (C potential, OpenMP over all host cores).
    A step is cut short only if the whole run would pass `--ref-budget` seconds."""
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    X, y = make_covtype_shape()
    from oracle import cpotentials, hmc as ohmc, threefry as T
    from oracle.hmc_util import Config

    cpotentials.build()
    cpotentials.use_all_cores()  # torchrun exports OMP_NUM_THREADS=1
    fam = cpotentials.CLogistic(X, y)
    kern = ohmc.HMCKernel(fam, cfg=Config(np.float32, True))
    t_begin = time.perf_counter()
    state = kern.init(T.PRNGKey(1), args.adapt)
    for _ in range(args.adapt):
        state = kern.sample(state)
    adapt_s, adapt_grads = time.perf_counter() - t_begin, kern.num_grad_evals
    per_step = args.ref_transitions or args.transitions
    budget = args.ref_budget
    cut = [False]

    def step(state):
        g0, t0, n = kern.num_grad_evals, time.perf_counter(), 0
        for _ in range(per_step):
            state = kern.sample(state)
            n += 1
            if time.perf_counter() - t_begin > budget:
                cut[0] = True
                break
        return state, kern.num_grad_evals - g0, time.perf_counter() - t0, n

    for _ in range(args.warmup):
        state, _, _, _ = step(state)
    grads, secs, ntr = 0, 0.0, 0
    for _ in range(args.steps):
        state, g, dt, n = step(state)
        grads += g
        secs += dt
        ntr += n
    value = grads / secs
    sample = (f"{args.steps} steps x {per_step} post-warm-up NUTS transitions ({ntr} transitions, "
              f"{grads} grad evals, {secs:.1f} s) at the step size / mass matrix adapted over "
              f"{args.adapt} untimed transitions ({adapt_grads} grad evals, {adapt_s:.1f} s), full "
              f"N={N_ROWS} D={N_COLS} data" + ("; steps cut short by --ref-budget" if cut[0] else ""))
    rows = args.sharding == "rows" and args.gpus > 1
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
        "higher_is_better": True, "scaling": "strong" if rows else "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": base_config(args, args.gpus, rows),
        "reference_kind": "CPU restatement of numpyro HMC/NUTS (oracle/, pinned bit for bit to the "
                          "reference's sampler source), C potential with OpenMP; jax is not "
                          "importable in this image, so NumPyro itself cannot run",
        "adapted_step_size": float(state.adapt_state.step_size),
        "mean_tree_size": grads / max(1, ntr),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": fam.cores, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------- GPU arm
def run_own(args):
    import ctypes as C
    import gc

    rank, world, local = dist_env()
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    import __graft_entry__ as g

    g.build()
    from numpyro_b200 import _lib, models
    from numpyro_b200 import distributed as nbd
    from numpyro_b200.infer import MCMC, NUTS
    from numpyro_b200.random import PRNGKey, split

    lib = _lib.lib()
    _lib.require_device()
    _lib.check(lib.b200mc_set_device(local))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist

        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        _lib.check(lib.b200mc_stream_synchronize(None))
        if dist is not None:
            dist.barrier()
        _lib.check(lib.b200mc_stream_synchronize(None))

    def over_ranks(x, op):
        if dist is None:
            return x
        import torch

        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
        return float(t.item())

    def pinned_copy(a):
        ptr = lib.b200mc_host_alloc(a.nbytes)
        if not ptr:
            raise _lib.B200Error(lib.b200mc_last_error().decode())
        buf = (C.c_char * a.nbytes).from_address(ptr)
        out = np.frombuffer(buf, dtype=a.dtype).reshape(a.shape)
        out[...] = a
        return out

    Xfull, yfull = make_covtype_shape()
    # the e2e leg copies from PINNED host memory
    Xfull, yfull = pinned_copy(Xfull), pinned_copy(yfull)
    allgather = None
    if world > 1:
        from tools import torch_transport as TT

        allgather = TT.allgather_bytes()
    T_ = args.transitions
    peak_hbm, peak_src = peaks()

    def layout_of(kern):
        out = (C.c_int * 8)()
        _lib.check(lib.b200mc_glm_layout(kern._handle, out))
        return dict(zip(("ctas", "stages", "pinned_per_cta", "tile_bytes", "tiles", "smem_bytes",
                         "chains_per_launch"), list(out)[:7]))

    def device_leg(kind, X, y, key, chains=1, steps=None, warmup=None, kernel_kw=None, adapt=None,
                   sample_clocks=False):
        """`adapt` untimed adaptation transitions, then `steps` timed steps of T_ transitions, each
        ONE b200mc_sample call on inputs resident in HBM, CUDA events on the launching stream.
        kind: "single" (this GPU), "rows" (one chain over all ranks), "replicas" (a chain per rank)."""
        steps = args.steps if steps is None else steps
        warmup = args.warmup if warmup is None else warmup
        adapt = args.adapt if adapt is None else adapt
        rows = kind == "rows"
        kern = NUTS(models.logistic_regression, **(kernel_kw or {})).shard_rows(
            rank, world if rows else 1, allgather)
        mc = MCMC(kern, num_warmup=adapt, num_samples=T_, num_chains=chains, chain_method="vectorized")
        t0 = time.perf_counter()
        mc.warmup(key, X, y, extra_fields=("num_steps",))
        adapt_s = time.perf_counter() - t0
        D = X.shape[1]
        ns_buf = _lib.DeviceArray((T_, chains), np.int32)
        z_buf = _lib.DeviceArray((T_, chains, D), np.float32)
        col = _lib.Collect(n_keep=T_, z_dev=z_buf.ptr, num_steps_dev=ns_buf.ptr)
        ev0, ev1 = lib.b200mc_event_create(), lib.b200mc_event_create()
        draws = []

        def device_step():
            _lib.check(lib.b200mc_event_record(ev0, None))
            _lib.check(lib.b200mc_sample(kern._handle, T_, 0, 1, C.byref(col), None))
            _lib.check(lib.b200mc_event_record(ev1, None))
            ms = C.c_float()
            _lib.check(lib.b200mc_event_elapsed_ms(ev0, ev1, C.byref(ms)))
            draws.append(z_buf.numpy())
            return float(ms.value), int(ns_buf.numpy().sum())

        for _ in range(warmup):
            device_step()
        prof = (C.c_uint64 * 16)()
        _lib.check(lib.b200mc_get_profile(kern._handle, prof, 1))  # reset the cycle counters
        launches0 = lib.b200mc_kernel_launches()
        clocks = ClockSampler(local) if (sample_clocks and rank == 0) else None
        barrier()
        step_ms, step_grads = [], []
        draws.clear()
        for _ in range(steps):
            ms, gsum = device_step()
            step_ms.append(ms)
            step_grads.append(gsum)
        barrier()
        clk = clocks.stop() if clocks is not None else None
        launches = lib.b200mc_kernel_launches() - launches0
        _lib.check(lib.b200mc_get_profile(kern._handle, prof, 0))
        passes = max(1, int(prof[4]))
        breakdown = {k: float(prof[i]) / passes for i, k in
                     enumerate(("stream", "arrival_wait", "grid_reduce", "sampler_step"))}
        breakdown["passes"] = passes
        total_ms = over_ranks(sum(step_ms), "MAX")
        local_grads = float(sum(step_grads))
        total_grads = local_grads if kind != "replicas" else over_ranks(local_grads, "SUM")
        # second half of the metric: min bulk ESS over the timed draws of this leg's chains
        from numpyro_b200.diagnostics import (bulk_effective_sample_size, effective_sample_size,
                                              split_gelman_rubin)

        x = np.concatenate(draws, axis=0).transpose(1, 0, 2)  # [chain, draw, D]
        ess = None
        if x.shape[1] >= 8:
            bulk = float(np.min(bulk_effective_sample_size(x)))
            tot = bulk if kind != "replicas" else over_ranks(bulk, "SUM")  # independent chains add
            ess = {"metric": "min_bulk_ess_per_sec", "value": tot / (total_ms * 1e-3), "unit": "ESS/s",
                   "min_bulk_ess": tot, "min_n_eff": float(np.min(effective_sample_size(x))),
                   "chains": chains if kind != "replicas" else chains * world,
                   "draws_per_chain": int(x.shape[1]),
                   "max_split_r_hat": (float(np.max(split_gelman_rubin(x.astype(np.float64))))
                                       if x.shape[1] >= 4 else None),
                   "note": "device-timed sampling phase, rank-normalised split-chain bulk ESS "
                           "(SURVEY 8d); min_n_eff is the reference's estimator"}
        st = mc.last_state if False else kern._last_state
        return {"kind": kind, "kern": kern, "value": total_grads / (total_ms * 1e-3),
                "total_ms": total_ms, "local_ms": sum(step_ms), "local_grads": local_grads,
                "total_grads": total_grads, "step_ms": step_ms, "step_grads": step_grads,
                "ess": ess, "breakdown": breakdown, "launches": int(launches), "clocks": clk,
                "adapt_s": adapt_s, "state": st,
                "adapted_step": np.atleast_1d(np.asarray(st.adapt_state.step_size, np.float64)),
                "imm": np.asarray(next(iter(st.adapt_state.inverse_mass_matrix.values())), np.float32),
                "layout": layout_of(kern)}

    rows = args.sharding == "rows" and world > 1
    if rows:
        lo, hi = nbd.shard_bounds(N_ROWS, rank, world)
        Xh, yh = Xfull[lo:hi], yfull[lo:hi]
        head = device_leg("rows", Xh, yh, PRNGKey(1), sample_clocks=True)
    elif world > 1:
        Xh, yh = Xfull, yfull
        head = device_leg("replicas", Xh, yh, split(PRNGKey(1), world)[rank], sample_clocks=True)
    else:
        Xh, yh = Xfull, yfull
        head = device_leg("single", Xh, yh, PRNGKey(1), sample_clocks=True)
    kern = head["kern"]
    adapted_step = float(head["adapted_step"][0])
    imm = head["imm"].reshape(-1, N_COLS)[0]

    # ---- roofline of the dominant (only) kernel on this rank.  The re-tiled design matrix
    #      (130.7 MB, minus the tiles pinned in shared memory) fits the 126 MB L2 and the ping-pong
    #      sweep keeps it there, so the roof that binds is the L2 -> SM feed of this access pattern,
    #      measured here with the kernel's own tile size and ring depth (b200mc_membench); the HBM
    #      figures sit beside it: DRAM bytes per gradient from the committed ncu capture, and a run
    #      on 2x the rows (out of L2) against the measured HBM copy peak.
    lay = head["layout"]
    my_bytes = ALGO_BYTES_PER_GRAD / world if rows else ALGO_BYTES_PER_GRAD
    my_gbs = my_bytes * head["local_grads"] / (head["local_ms"] * 1e-3) / 1e9
    streamed = (lay["tiles"] - lay["pinned_per_cta"] * lay["ctas"]) * lay["tile_bytes"]
    l2_peak = None
    if rank == 0:
        gbs = C.c_float()
        stages = lay["stages"]  # the kernel's own ring depth (a multiple of its consumer-warp count)
        _lib.check(lib.b200mc_membench(max(streamed, 8 << 20), lay["tile_bytes"], stages, 40, 1,
                                       C.byref(gbs), None))
        l2_peak = float(gbs.value)
    traffic, traffic_src = None, None
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        with open(tr) as f:
            tj = json.load(f)
        traffic = tj.get("glm_kernel_bytes_per_grad_eval")
        traffic_src = {k: tj.get(k) for k in ("commit", "recipe", "captured") if k in tj}
    roofline = {"bound": "l2", "achieved": my_gbs, "peak": l2_peak, "unit": "GB/s",
                "frac": (my_gbs / l2_peak) if l2_peak else None,
                "peak_source": "measured in this run: b200mc_membench, the kernel's access pattern "
                               "(7.2 KB TMA tiles, same ring, ping-pong) on an L2-resident buffer of the "
                               "streamed size",
                "kernel": "glm_kernel", "bytes_per_grad_eval": ALGO_BYTES_PER_GRAD,
                "streamed_bytes_per_pass": int(streamed),
                "grad_evals_per_launch": head["local_grads"] / max(1, len(head["step_grads"])),
                "us_per_grad_eval": 1e3 * head["local_ms"] / max(1.0, head["local_grads"]),
                "traffic": traffic, "traffic_source": traffic_src,
                "hbm": {"peak": peak_hbm, "peak_source": peak_src,
                        "algorithmic_frac_of_hbm_peak": my_gbs / peak_hbm,
                        "dram_frac": (traffic * head["local_grads"] / (head["local_ms"] * 1e-3) / 1e9 / peak_hbm
                                      if traffic else None),
                        "note": "algorithmic bytes / HBM peak exceeds 1 because ~90 % of every pass is "
                                "served by L2; dram_frac is the DRAM traffic actually measured (ncu)"}}

    # ---- out-of-L2 companion (1 GPU): the same model on 2x the rows (261 MB re-tiled, nothing
    #      reusable across passes), one chain, against the measured HBM copy peak
    out_of_l2 = None
    if world == 1 and not args.no_out_of_l2:
        X2 = np.concatenate([Xfull, Xfull], axis=0)
        y2 = np.concatenate([yfull, yfull], axis=0)
        big = device_leg("single", X2, y2, PRNGKey(1), steps=max(2, args.steps // 4), warmup=2, adapt=150)
        b2 = 2 * ALGO_BYTES_PER_GRAD
        gbs2 = b2 * big["local_grads"] / (big["local_ms"] * 1e-3) / 1e9
        # the roof of a READ-only stream: MEASURED_PEAKS' copy figure counts read + write traffic
        # of a copy and sits below what a pure read reaches, so the same access pattern is
        # measured out of L2 here (1 GiB buffer, forward sweeps) and the larger of the two is used
        gb = C.c_float()
        bl = big["layout"]
        _lib.check(lib.b200mc_membench(1 << 30, bl["tile_bytes"], bl["stages"], 6, 0, C.byref(gb), None))
        read_peak = float(gb.value)
        hbm_roof = max(peak_hbm, read_peak)
        # bytes that actually leave memory per pass: the tiles pinned in shared memory (2 % here) are
        # read once per launch, so the algorithmic rate can exceed the streaming roof by that share;
        # `frac` is taken on the streamed bytes and can therefore not pass 1 for that reason
        streamed2 = (bl["tiles"] - bl["pinned_per_cta"] * bl["ctas"]) * bl["tile_bytes"]
        gbs2_streamed = streamed2 * big["local_grads"] / (big["local_ms"] * 1e-3) / 1e9
        out_of_l2 = {"rows": int(X2.shape[0]), "bytes_per_grad_eval": b2,
                     "grad_evals_per_sec": big["value"],
                     "us_per_grad_eval": 1e3 * big["local_ms"] / max(1.0, big["local_grads"]),
                     "roofline": {"bound": "hbm", "achieved": gbs2, "peak": hbm_roof, "unit": "GB/s",
                                  "frac": gbs2_streamed / hbm_roof, "achieved_streamed": gbs2_streamed,
                                  "streamed_bytes_per_pass": streamed2,
                                  "peak_source": ("HBM read roof measured in this run (b200mc_membench, 1 GiB, "
                                                  "the kernel's TMA pattern)" if read_peak >= peak_hbm
                                                  else "MEASURED_PEAKS.json copy bandwidth"),
                                  "copy_peak": peak_hbm, "read_peak_measured": read_peak},
                     "cta0_cycles_per_grad_eval": big["breakdown"], "layout": big["layout"]}
        big["kern"]._destroy()
        del X2, y2, big

    # ---- parity check on the full data, after the timed region: U and grad of the CUDA path at
    #      the chain's last position against one float64 oracle evaluation
    parity = None
    z_last = np.asarray(head["state"].z["coefs"], np.float32).reshape(-1, N_COLS)[:1]
    Ud, gd = kern.potential_and_grad(z_last)  # every rank takes part when the rows are sharded
    if rank == 0:
        from oracle import potentials as opot

        U0, g0 = opot.Logistic(Xfull, yfull, np.float64).potential_and_grad(z_last[0].astype(np.float64))
        parity = {"max_rel_err_U": float(abs(Ud[0] - U0) / abs(U0)),
                  "max_rel_err_grad": float(np.max(np.abs(gd[0] - g0)) / np.max(np.abs(g0))),
                  "at": "last position of the timed chain, full N=581012 D=55 data",
                  "oracle": "oracle.potentials.Logistic float64", "tolerance": 1e-5}
        parity["ok"] = bool(parity["max_rel_err_U"] < 1e-5 and parity["max_rel_err_grad"] < 1e-5)

    # ---- end-to-end leg: public API (MCMC.run) with HOST buffers every step: X, y are copied
    #      host->device inside the timed region (pinned memory), re-tiled, the chain initialised,
    #      T_ transitions run and the draws read back.  (No clock sampler here: nvidia-smi's NVML
    #      queries take driver locks the many small driver calls of this leg wait on.)
    e2e_grads, e2e_s, e2e_ms = 0, 0.0, []
    z_last_d = {k: np.asarray(v) for k, v in head["state"].z.items()}
    e2e_key = PRNGKey(1) if head["kind"] != "replicas" else split(PRNGKey(1), world)[rank]
    gc.collect()
    gc.disable()  # no collector pauses inside the timed host-side steps
    barrier()
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        k2 = NUTS(models.logistic_regression, step_size=adapted_step, inverse_mass_matrix=imm,
                  adapt_step_size=False, adapt_mass_matrix=False).shard_rows(
                      rank, world if rows else 1, allgather)
        m2 = MCMC(k2, num_warmup=0, num_samples=T_)
        m2.run(e2e_key, Xh, yh, init_params=z_last_d, extra_fields=("num_steps",))
        ns = m2.get_extra_fields()["num_steps"]
        _ = m2.get_samples()["coefs"].mean()
        dt = time.perf_counter() - t0
        k2._destroy()
        if i >= args.warmup:
            e2e_grads += int(ns.sum())
            e2e_s += dt
            e2e_ms.append(round(dt * 1e3, 3))
    gc.enable()
    barrier()
    e2e_s = over_ranks(e2e_s, "MAX")
    e2e_grads = float(e2e_grads) if head["kind"] != "replicas" else over_ranks(float(e2e_grads), "SUM")
    e2e = {"value": e2e_grads / e2e_s, "unit": UNIT,
           "h2d_bytes_per_step": int(Xh.nbytes + yh.nbytes + 4 * N_COLS * 2 + 8),
           "d2h_bytes_per_step": int(T_ * (N_COLS * 4 + 4 + 4)),
           "ms_per_step": e2e_ms, "grad_evals_per_step": e2e_grads / max(1, len(e2e_ms))}
    clk = head["clocks"]
    if clk is not None and rank == 0:
        probe = ClockSampler(local, 100)  # one reading right after the leg (never during it)
        time.sleep(0.35)
        clk2 = probe.stop()
        clk["e2e_leg"] = clk2
        clk["reasons"] = sorted(set(clk["reasons"]) | set(clk2["reasons"]))

    # ---- N > 1: the other partitioning as a sub-object (replicas when rows are the headline)
    replicas = None
    if rows:
        rep = device_leg("replicas", Xfull, yfull, split(PRNGKey(1), world)[rank],
                         steps=max(2, args.steps // 2), warmup=2)
        replicas = {"scaling": "weak", "parallelism": f"chains x{world} (one chain per GPU, X replicated, "
                    "no data-path collective)", "value": rep["value"], "unit": UNIT,
                    "ms_per_step": rep["total_ms"] / len(rep["step_ms"]), "ess": rep["ess"]}
        rep["kern"]._destroy()

    # ---- 4 chains on ONE stream of X (glm_mc_kernel): the ESS half of the metric on one GPU
    multi = None
    if world == 1 and not args.no_multi_chain:
        m4 = device_leg("single", Xfull, yfull, PRNGKey(1), chains=4, steps=max(3, args.steps // 2), warmup=2)
        multi = {"chains": 4, "kernel": "glm_mc_kernel (4 chains per pass over X)",
                 "grad_evals_per_sec": m4["value"], "vs_one_chain": m4["value"] / head["value"],
                 "ms_per_step": m4["total_ms"] / len(m4["step_ms"]), "ess": m4["ess"],
                 "cta0_cycles_per_pass": m4["breakdown"], "layout": m4["layout"],
                 "gpu_launches": m4["launches"],
                 "adapted_step_sizes": [float(v) for v in m4["adapted_step"]]}
        m4["kern"]._destroy()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        gr, dt, ntr, cores = cpu_nuts_sample(Xfull, yfull, args.cpu_budget, 10**6)
        cpu = {"value": gr / dt, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{ntr} NUTS transitions ({gr} grad evals, {dt:.1f} s) of the oracle "
                         f"from the init state on the full N={N_ROWS} D={N_COLS} data"}
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        line = {
            "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["total_ms"] / args.steps, "higher_is_better": True,
            "scaling": "strong" if rows else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": base_config(args, world, rows),
            "config_detail": {"adapted_step_size": adapted_step, "layout": lay,
                              "l2": "re-tiled inputs 130.7 MB per pass; the tiles not pinned in shared "
                                    "memory (%.1f MB) stay L2-resident under the ping-pong sweep: DRAM "
                                    "traffic is ~0.09x the algorithmic bytes (roofline.traffic)"
                                    % (streamed / 1e6)},
            "roofline": roofline, "out_of_l2": out_of_l2, "parity_check": parity,
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": head["launches"],
            "clocks": clk, "adapt_seconds": head["adapt_s"], "ess": head["ess"],
            "cta0_cycles_per_grad_eval": head["breakdown"],
            "mean_tree_size": float(np.mean(head["step_grads"])) / T_,
        }
        if rows:
            bd = head["breakdown"]
            line["exchange"] = {"note": "per-leapfrog serial chain at this rank count, SM cycles of CTA 0 "
                                        "per gradient: grid arrival wait, fixed-order grid reduction, and the "
                                        "sampler step that contains the NVLink (value, seq) exchange",
                                "cycles": {k: bd[k] for k in ("arrival_wait", "grid_reduce", "sampler_step")}}
            line["replicas"] = replicas
        if multi is not None:
            line["multi_chain"] = multi
        if world == 1 and not args.no_many_chain:
            line["many_chain"] = many_chain_leg()
        print(json.dumps(line))
    return 0


def many_chain_leg():
    """The same data with 256 vectorized chains through the tensor-core lock-step engine
    (north_star: from 64 chains on X.Theta is a dense contraction), run as a separate process
    (`--workload logistic_many`) after the headline legs and reported next to them: grad-evals/s
    and, above all, min-bulk-ESS/s, which is what many chains buy.  Never fatal for the line."""
    try:
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "--workload", "logistic_many",
                              "--chains", "256", "--steps", "3", "--warmup", "2", "--no-cpu"],
                             capture_output=True, text=True, timeout=300)
        d = json.loads(out.stdout.strip().splitlines()[-1])
        return {"chains": 256, "grad_evals_per_sec": d["value"], "ms_per_step": d["ms_per_step"],
                "transitions_per_step": d["config"]["transitions_per_step"],
                "min_bulk_ess_per_sec": (d.get("ess") or {}).get("value"),
                "max_split_r_hat": (d.get("ess") or {}).get("max_split_r_hat"),
                "roofline": d.get("roofline"), "clocks": d.get("clocks"),
                "gpu_launches": d.get("gpu_launches")}
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}


# ---------------------------------------------------------------------------- other configs
def make_workload(name, chains):
    """BASELINE.json configs[1], [3], [4] on synthetic data."""
    from numpyro_b200 import models

    if name == "eight_schools":
        y = np.array([28.0, 8.0, -3.0, 7.0, -1.0, 1.0, 18.0, 12.0])
        sg = np.array([15.0, 10.0, 16.0, 11.0, 9.0, 11.0, 10.0, 18.0])
        return (models.eight_schools_noncentered, (8, sg), {"y": y}, {}, chains or 16384,
                "eight schools non-centred NUTS, vectorized chains (configs[1])")
    if name == "sv":
        rng = np.random.default_rng(21)
        T_ = 3000
        s = np.cumsum(rng.normal(size=T_) * 0.02) - 4.5
        r = (np.exp(s) * rng.standard_t(10, size=T_)).astype(np.float32)
        return (models.stochastic_volatility, (r,), {}, {}, chains or 1024,
                "stochastic volatility random walk T=3000 NUTS (configs[4])")
    if name == "logistic_many":
        X, y = make_covtype_shape()
        return (models.logistic_regression, (X, y), {}, {}, chains or 256,
                "covtype-shape logistic NUTS N=581012 D=55, vectorized chains through the "
                "tensor-core lock-step engine (configs[2] data, many chains)")
    if name == "horseshoe":
        rng = np.random.RandomState(0)
        N, D = 10000, 1000
        X = rng.randn(N, D).astype(np.float32)
        Y = X[:, :3] @ np.array([2.0, -1.0, 0.5])
        Y = (Y - Y.mean() + 0.05 * rng.randn(N)).astype(np.float32)
        return (models.horseshoe_normal, (X, Y), {}, {"dense_mass": True}, chains or 256,
                "horseshoe regression D=1000 N=10000 dense mass NUTS (configs[3])")
    raise SystemExit(f"unknown workload {name}")


def run_chain_workload(args):
    """Chain-batched configs: a step = `--transitions` NUTS transitions of ALL chains in one
    persistent launch; metric = grad evals/s summed over chains (device timed)."""
    import ctypes as C

    rank, world, local = dist_env()
    import __graft_entry__ as g

    g.build()
    from numpyro_b200 import _lib
    from numpyro_b200.infer import MCMC, NUTS
    from numpyro_b200.random import PRNGKey

    lib = _lib.lib()
    _lib.require_device()
    _lib.check(lib.b200mc_set_device(local))
    dist = None
    if world > 1:  # chains sharded over the ranks, data replicated, no data-path collective
        import torch
        import torch.distributed as dist

        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def reduce_over_ranks(x, op):
        if dist is None:
            return x
        import torch

        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
        return float(t.item())

    def barrier():
        _lib.check(lib.b200mc_stream_synchronize(None))
        if dist is not None:
            dist.barrier()

    model, margs, mkw, kkw, chains, desc = make_workload(args.workload, args.chains)
    total_chains = chains
    from numpyro_b200 import distributed as nbd
    from numpyro_b200.random import split

    keys = nbd.shard_keys(split(PRNGKey(0), total_chains), rank, world)  # mcmc.py:661-662
    chains = int(keys.shape[0])
    if args.fixed_step > 0:  # profiling aid: no adaptation, trees at a given step size
        kkw = dict(kkw, step_size=args.fixed_step, adapt_step_size=False, adapt_mass_matrix=False)
    if args.max_tree_depth:
        kkw = dict(kkw, max_tree_depth=args.max_tree_depth)
    kern = NUTS(model, **kkw)
    mc = MCMC(kern, num_warmup=args.adapt, num_samples=args.transitions, num_chains=chains,
              chain_method="vectorized")
    t0 = time.perf_counter()
    mc.warmup(keys if total_chains > 1 else PRNGKey(0), *margs, **mkw)
    adapt_s = time.perf_counter() - t0
    T_ = args.transitions
    ns_buf = _lib.DeviceArray((T_, chains), np.int32)
    col = _lib.Collect(n_keep=T_, num_steps_dev=ns_buf.ptr)
    # second half of BASELINE's metric (min bulk ESS / s): keep the timed draws when they are
    # small enough to copy out between steps (outside the timed events)
    D_all = int(kern._bound.dim)
    z_buf, draws = None, []
    if T_ * chains * D_all * 4 <= (512 << 20) and world == 1:
        z_buf = _lib.DeviceArray((T_, chains, D_all), np.float32)
        col.z_dev = z_buf.ptr
    ev0, ev1 = lib.b200mc_event_create(), lib.b200mc_event_create()
    tot_ms, tot_g = 0.0, 0
    launches0 = None
    # The lock-step engine is HOST-driven (three launches per global step): nvidia-smi's start-up
    # and NVML queries take driver locks those launches wait on.  The sampler therefore starts
    # before the warm-up steps (its start-up is over when the timed region begins) and polls
    # every 500 ms instead of every 100 ms.
    clocks = ClockSampler(local, 500)
    for i in range(args.warmup + args.steps):
        if i == args.warmup:
            launches0 = lib.b200mc_kernel_launches()
            barrier()
        _lib.check(lib.b200mc_event_record(ev0, None))
        _lib.check(lib.b200mc_sample(kern._handle, T_, 0, 1, C.byref(col), None))
        _lib.check(lib.b200mc_event_record(ev1, None))
        ms = C.c_float()
        _lib.check(lib.b200mc_event_elapsed_ms(ev0, ev1, C.byref(ms)))
        if i >= args.warmup:
            tot_ms += float(ms.value)
            tot_g += int(ns_buf.numpy().sum())
            if z_buf is not None:
                draws.append(z_buf.numpy())
    barrier()
    clk = clocks.stop()
    local_g, local_ms = tot_g, tot_ms
    tot_ms = reduce_over_ranks(tot_ms, "MAX")
    tot_g = int(reduce_over_ranks(float(tot_g), "SUM"))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0
    # step-level roofline of the secondary workloads: whole-step time, not one kernel
    roofline = None
    D_flat = int(kern._bound.dim)
    if args.workload == "horseshoe" and kkw.get("dense_mass"):
        peak, src = peaks()
        per = D_flat * D_flat * 4  # one pass over M^-1 per chain and leapfrog
        ach = per * local_g / (local_ms * 1e-3) / 1e9  # this rank's GPU
        roofline = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                    "traffic": None, "peak_source": src, "kernel": "batch_matvec_kernel (whole step timed)",
                    "bytes_per_grad_eval": per}
    elif args.workload == "sv":
        # SURVEY 8d's state-traffic figure (six D-vectors per chain and leapfrog, 72 KB) is NOT what
        # bounds this kernel: the committed ncu capture (tools/ncu_chain.py sv) measures ~11 KB of
        # DRAM traffic per gradient (the leaf vectors live in shared memory, the tree slots in L2)
        # at ~2 % of the HBM peak; the kernel is issue / latency bound, so the roof reported is
        # issue-slot utilisation, with the HBM figures beside it.
        peak, src = peaks()
        per = 6 * D_flat * 4
        ach = per * local_g / (local_ms * 1e-3) / 1e9
        roofline = {"bound": "issue_slots", "achieved": None, "peak": 100.0, "unit": "% of issue slots",
                    "frac": None, "traffic": None, "kernel": "chain_kernel<2> (CTA per chain)",
                    "hbm": {"state_traffic_bytes_per_grad_eval": per, "state_traffic_gbs": ach, "peak": peak,
                            "peak_source": src, "state_traffic_frac_of_hbm_peak": ach / peak}}
        pj = os.path.join(ROOT, "profiles", "sv_ncu.json")
        if os.path.exists(pj):
            with open(pj) as f:
                nj = json.load(f)
            roofline.update({"achieved": nj.get("issue_active_pct"), "frac": (nj.get("issue_active_pct") or 0) / 100.0,
                             "traffic": nj.get("dram_bytes_per_grad_eval"),
                             "fma_pipe_pct": nj.get("fma_pipe_pct"), "xu_pipe_pct": nj.get("xu_pipe_pct"),
                             "source": {k: nj.get(k) for k in ("commit", "recipe", "launch")}})
    elif args.workload == "eight_schools":
        # no meaningful HBM roofline (state in registers, ~200 flops per gradient): issue-slot
        # utilisation from the committed ncu capture + the collection write bandwidth of this run
        roofline = {"bound": "issue_slots", "achieved": None, "peak": 100.0, "unit": "% of issue slots",
                    "frac": None, "traffic": None, "kernel": "chain_kernel<0> (thread per chain)",
                    "collection_write_gbs": (T_ * chains * D_flat * 4 * args.steps) / (local_ms * 1e-3) / 1e9}
        pj = os.path.join(ROOT, "profiles", "eight_schools_ncu.json")
        if os.path.exists(pj):
            with open(pj) as f:
                nj = json.load(f)
            roofline.update({"achieved": nj.get("issue_active_pct"), "frac": (nj.get("issue_active_pct") or 0) / 100.0,
                             "active_threads_per_inst": nj.get("active_threads_per_inst"),
                             "source": {k: nj.get(k) for k in ("commit", "recipe", "launch")}})
    elif args.workload == "logistic_many":
        # two contractions of 2*N*D*C flops each, three TF32 passes; dense TF32 peak 1130 TFLOP/s
        # (blackwell guide; MEASURED_PEAKS.json holds bf16 only)
        flops = 3 * 2 * 2.0 * N_ROWS * N_COLS * local_g
        ach = flops / (local_ms * 1e-3) / 1e12  # this rank's GPU
        roofline = {"bound": "tensor", "achieved": ach, "peak": 1130.0, "unit": "TFLOP/s",
                    "frac": ach / 1130.0, "traffic": None, "peak_source": "nominal TF32 dense",
                    "kernel": "gemm3x_kernel x2 (whole step timed; the residual images' HBM "
                              "round trip and the transcendental epilogue dominate, DESIGN 4.3)"}
    ess = None
    if draws and args.steps * T_ >= 8:
        from numpyro_b200.diagnostics import bulk_effective_sample_size, device_ess_rhat

        xs = np.concatenate(draws, axis=0)  # [draw, chain, D], the collection layout
        # the reference's estimators (Geyer ESS, split R-hat) over ALL chains and ALL parameters on
        # the device (csrc/diag_kernel.cuh)
        t0 = time.perf_counter()
        xd = _lib.DeviceArray.from_numpy(xs)
        dd = device_ess_rhat(xd, layout="draw_chain")
        diag_s = time.perf_counter() - t0
        del xd
        # rank-normalised bulk ESS (harness metric, host): the pooled ranking of chains x draws
        # values per parameter is the expensive part, so at most 16 evenly spaced parameters
        x = xs.transpose(1, 0, 2)
        if x.shape[2] > 16:
            x = x[:, :, np.unique(np.linspace(0, x.shape[2] - 1, 16).astype(int))]
        bulk = np.array([bulk_effective_sample_size(x[:, :, d:d + 1])[0] for d in range(x.shape[2])])
        ess = {"min_bulk_ess": float(np.min(bulk)), "draws_per_chain": int(x.shape[1]),
               "chains": int(x.shape[0]), "params_checked_bulk": int(x.shape[2]),
               "min_n_eff_all_params_device": float(np.min(dd["n_eff"])),
               "min_n_eff_per_sec": float(np.min(dd["n_eff"])) / (tot_ms * 1e-3),
               "max_split_r_hat": float(np.max(dd["split_r_hat"])), "params_checked_device": int(xs.shape[2]),
               "device_diagnostics_seconds": diag_s,
               "metric": "min_bulk_ess_per_sec", "value": float(np.min(bulk)) / (tot_ms * 1e-3),
               "unit": "ESS/s", "note": "device-timed sampling phase; n_eff and split R-hat (the "
               "reference's estimators) over all chains and all parameters on the device; "
               "rank-normalised bulk ESS (SURVEY 8d) on the host over <= 16 parameters"}
    line = {"metric": METRIC, "value": tot_g / (tot_ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "ess": ess,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": desc, "chains": total_chains, "chains_per_gpu": chains,
                       "parallelism": f"chains x{world}", "transitions_per_step": T_,
                       "adapt_transitions_untimed": args.adapt, **{k: str(v) for k, v in kkw.items()}},
            "gpu_launches": int(lib.b200mc_kernel_launches() - launches0), "clocks": clk,
            "adapt_seconds": adapt_s, "mean_tree_size": tot_g / (args.steps * T_ * total_chains),
            "roofline": roofline}
    print(json.dumps(line))
    return 0


def run_ecs_workload(args):
    """HMCECS on the covtype shape (the default algorithm of examples/covtype.py:154-165,223):
    NUTS on a 1000-row subsample with the second-order Taylor proxy, 100 blocks.  One GPU; a step
    is `--transitions` iterations of { Gibbs step; NUTS transition } in one launch."""
    rank, world, local = dist_env()
    if rank != 0:
        return 0
    sys.path.insert(0, ROOT)
    import ctypes as C

    import __graft_entry__ as g

    g.build()
    from numpyro_b200 import _lib, models
    from numpyro_b200.diagnostics import bulk_effective_sample_size, split_gelman_rubin
    from numpyro_b200.infer import HMCECS, NUTS
    from numpyro_b200.random import PRNGKey, split

    lib = _lib.lib()
    _lib.require_device()
    X, y = make_covtype_shape()
    m, nb = 1000, 100
    # reference parameters: the posterior mode by Newton steps on the full data, float64 (the
    # example gets them from SVI, covtype.py:137-152)
    X64, y64, ref = X.astype(np.float64), y.astype(np.float64), np.zeros(N_COLS)
    for _ in range(8):
        sg = 1.0 / (1.0 + np.exp(-X64 @ ref))
        gr = ref + X64.T @ (sg - y64)
        H = np.eye(N_COLS) + (X64.T * (sg * (1 - sg))) @ X64
        ref = ref - np.linalg.solve(H, gr)
    del X64
    chains = args.chains or 1
    kern = HMCECS(NUTS(models.logistic_regression), num_blocks=nb,
                  proxy=HMCECS.taylor_proxy({"coefs": ref.astype(np.float32)}))
    keys = split(PRNGKey(1), chains) if chains > 1 else PRNGKey(1)
    margs, mkw = (X, y), {"subsample_size": m}
    t0 = time.perf_counter()
    state = kern.init(keys, args.adapt, None, margs, mkw)
    _, state = kern.run_loop(state, args.adapt, 0, 1, 0, [], margs, mkw)
    adapt_s = time.perf_counter() - t0
    ev0, ev1 = lib.b200mc_event_create(), lib.b200mc_event_create()
    clocks = ClockSampler(local)
    tot_ms, tot_g, draws, accs, launches0 = 0.0, 0, [], [], None
    T_ = args.transitions
    for it in range(args.warmup + args.steps):
        if it == args.warmup:
            launches0 = lib.b200mc_kernel_launches()
        _lib.check(lib.b200mc_event_record(ev0, None))
        raw, state = kern.run_loop(state, T_, 0, 1, T_, ["z", "accept_prob", "hmc_state.num_steps"],
                                   margs, mkw)
        _lib.check(lib.b200mc_event_record(ev1, None))
        ms = C.c_float()
        _lib.check(lib.b200mc_event_elapsed_ms(ev0, ev1, C.byref(ms)))
        if it >= args.warmup:
            tot_ms += float(ms.value)
            tot_g += int(raw["hmc_state.num_steps"].sum())
            draws.append(raw["z"])
            accs.append(raw["accept_prob"])
    clk = clocks.stop()
    launches = lib.b200mc_kernel_launches() - launches0
    x = np.concatenate(draws, axis=0).transpose(1, 0, 2)
    bulk = float(np.min(bulk_effective_sample_size(x)))
    line = {
        "metric": "grad_evals_per_sec", "value": tot_g / (tot_ms * 1e-3), "unit": "grad-evals/s",
        "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "covtype-shape logistic HMCECS(NUTS), subsample 1000 of 581012 rows, "
                               "100 blocks, second-order Taylor proxy (examples/covtype.py:154-165)",
                   "chains": chains, "transitions_per_step": T_, "adapt_transitions_untimed": args.adapt,
                   "note": "a gradient evaluation here touches 1000 rows, not 581012"},
        "transitions_per_sec": args.steps * T_ * chains / (tot_ms * 1e-3),
        "mean_subsample_accept_prob": float(np.mean(np.concatenate(accs))),
        "ess": {"metric": "min_bulk_ess_per_sec", "value": bulk / (tot_ms * 1e-3), "unit": "ESS/s",
                "min_bulk_ess": bulk, "draws_per_chain": int(x.shape[1]), "chains": chains,
                "max_split_r_hat": float(np.max(split_gelman_rubin(x.astype(np.float64))))},
        "gpu_launches": int(launches), "clocks": clk, "adapt_seconds": adapt_s,
        "posterior_mean_minus_mode_max_sd": float(np.max(
            np.abs(x.reshape(-1, N_COLS).mean(0) - ref) / x.reshape(-1, N_COLS).std(0))),
    }
    if not args.no_cpu:  # the oracle's HMCECS on one core, same data / proxy / blocks
        from oracle import ecs as oecs
        from oracle.threefry import PRNGKey as OKey

        fam = oecs.LogisticECS(X, y, m, 2, ref=ref.astype(np.float32))
        ok = oecs.HMCECSKernel(fam, num_blocks=nb)
        st = ok.init(OKey(1), 50)
        t0, n_tr, g0 = time.perf_counter(), 0, None
        while time.perf_counter() - t0 < args.cpu_budget:
            if n_tr == 50:
                g0, t1 = ok.inner.num_grad_evals, time.perf_counter()
            st = ok.sample(st)
            n_tr += 1
        if g0 is not None and n_tr > 60:
            dt = time.perf_counter() - t1
            line["cpu_baseline"] = {"value": (ok.inner.num_grad_evals - g0) / dt, "unit": "grad-evals/s",
                                    "cores": 1, "kind": "port",
                                    "sample": f"{n_tr - 50} HMCECS transitions after 50 adaptive ones "
                                              "(oracle/ecs.py, NumPy)",
                                    "transitions_per_sec": (n_tr - 50) / dt}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="logistic",
                    choices=["logistic", "eight_schools", "sv", "horseshoe", "logistic_many", "ecs"])
    ap.add_argument("--chains", type=int, default=0)
    ap.add_argument("--sharding", default="rows", choices=["chains", "rows"],
                    help="N>1 headline: one row-sharded chain (strong, configs[2]'s partitioning; "
                         "default) or independent chains per GPU (weak)")
    ap.add_argument("--no-out-of-l2", action="store_true", help="skip the 2x-rows HBM-bound companion run")
    ap.add_argument("--no-multi-chain", action="store_true", help="skip the 4-chains-on-one-stream leg")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--transitions", type=int, default=100, help="NUTS transitions per step")
    ap.add_argument("--adapt", type=int, default=300, help="untimed adaptation transitions")
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU baseline work")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-many-chain", action="store_true",
                    help="skip the 256-chain leg appended to the default line")
    ap.add_argument("--fixed-step", type=float, default=0.0,
                    help="chain workloads: fixed step size, no adaptation (profiling aid)")
    ap.add_argument("--max-tree-depth", type=int, default=0)
    ap.add_argument("--ref-transitions", type=int, default=0,
                    help="reference arm: transitions per step (0 = --transitions, the GPU arm's)")
    ap.add_argument("--ref-budget", type=float, default=240.0,
                    help="reference arm: seconds after which steps are cut short")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "ecs":
        return run_ecs_workload(args)
    if args.workload != "logistic":
        return run_chain_workload(args)
    return run_own(args)


if __name__ == "__main__":
    sys.exit(main())
