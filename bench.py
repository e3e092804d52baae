#!/usr/bin/env python
"""bench.py — throughput of the celerite hot path on B200 (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload cfg2|cfg2m|cfg3|cfg4|cfg5|cfg1] [--no-extras]

A "step" is one pass of the hot path over one batch of synthetic irregularly-sampled series.
The headline line is config 2 of BASELINE.json (65,536 series x N=2,000, one SHO term, J=2,
fused factorisation + log-likelihood); metric = point-terms/s = B*N*J / time, whole job over all
ranks (weak scaling: every rank sweeps its own batch of the named size; no data-path collective,
SURVEY.md section 8e).  The same JSON line carries, under "extras", one entry per remaining
config, each measured in the same process right after the headline:

    cfg1   single GP, RealTerm+ComplexTerm (J=3), N=4,096: compute + log_likelihood +
           grad_log_likelihood through the drop-in solver class, wall clock per call, the CPU
           reference beside it                                             (N=1 only)
    cfg2m  cfg2's batch through the materialising `compute` (phi,u,W,D written): the HBM-bound
           member, (3J+3)*8 B/point                                        (N=1 only)
    cfg3   1,024 series x N=100,000 with gradient at J = 2, 8, 32          (N=1 only)
    cfg4   one series N=10^7, J=8: apply_inverse + dot_L as chunked scans; at N>1 the time axis
           is split over the ranks (STRONG scaling) and the span carries are exchanged with NCCL
    cfg5   10^6 series x N=1,000, J=4 as 125,000 series per GPU (weak scaling)

Keys beyond the base contract (headline and every extra):
    roofline      algorithmic HBM bytes per launch / CUDA-event time of the kernel, against
                  MEASURED_PEAKS.json; traffic = ncu DRAM bytes of the committed capture
    cpu_baseline  the UNMODIFIED reference (oracle/_ref) timed on this box's host cores on a
                  bounded sample of the same workload, with the pure relative difference between
                  its results and the GPU's on that sample
    e2e           same metric through the C ABI with HOST (pinned) buffers: H2D of every input
                  and D2H of every result inside the timed region; h2d_roof = what a bare
                  cudaMemcpyAsync of the same pinned buffers reaches at the same moment
`--impl reference` times the reference's own CPU implementation (all host threads, one series
per task) on a bounded sample of the headline config.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "GP log-likelihood point-terms/sec (N*J*batch / time)"
METRIC_GRAD = "GP log-likelihood + gradient point-terms/sec (N*J*batch / time)"
METRIC_COMPUTE = "GP batched factorisation (materialised) point-terms/sec (N*J*batch / time)"
METRIC_CFG4 = "GP apply_inverse + dot_L point-terms/sec (2*N*J / time)"
UNIT = "point-terms/s"

WORKLOADS = {
    # name: (B per GPU, N, Jr, Jc)
    "cfg2": dict(B=65536, N=2000, Jr=0, Jc=1, desc="cfg2: 65,536 series x N=2,000, one SHOTerm (J=2), log_likelihood"),
    "cfg5": dict(B=125000, N=1000, Jr=0, Jc=2, desc="cfg5: 10^6 series x N=1,000 over 8 GPUs (125,000 per GPU), two SHOTerms (J=4), log_likelihood"),
    # the materialising batched factorisation (compute, factor written in the reference layout) on cfg2's
    # batch: the HBM-bound member of the family, (3J+3)*8 = 72 B/point at J=2
    "cfg2m": dict(B=65536, N=2000, Jr=0, Jc=1, materialise=True,
                  desc="cfg2m: 65,536 series x N=2,000, one SHOTerm (J=2), batched compute writing phi,u,W,D"),
    # cfg3 is a sweep over J in BASELINE.json; --jc picks the number of SHO terms (J = 2*jc)
    "cfg3": dict(B=1024, N=100000, Jr=0, Jc=4, grad=True, desc="cfg3: 1,024 series x N=100,000, J=2*jc (SHOTerms), grad_log_likelihood"),
}


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


def ncu_traffic(workload):
    """dram bytes per launch of the dominant kernel from the committed ncu summary, if any."""
    p = os.path.join(ROOT, "profiles", "ncu_summary.json")
    try:
        return json.load(open(p))[workload]["dram_bytes_per_launch"]
    except Exception:
        return None


def pure_rel(got, ref):
    got, ref = np.asarray(got, dtype=float), np.asarray(ref, dtype=float)
    return float(np.max(np.abs(got - ref) / np.maximum(np.abs(ref), 1e-300)))


def grad_rel(g, ref):
    """per-component |d| / max(|ref|, 1e-12 ||ref||_inf), maximised."""
    g, ref = np.asarray(g, dtype=float), np.asarray(ref, dtype=float)
    floor = 1e-12 * np.max(np.abs(ref), axis=-1, keepdims=True)
    return float(np.max(np.abs(g - ref) / np.maximum(np.abs(ref), np.maximum(floor, 1e-300))))


# ------------------------------------------------------------------------------------------------
# --impl reference: the reference's CPU solver on all host threads, bounded sample of cfg2
# ------------------------------------------------------------------------------------------------
def run_reference(args, wl):
    rank = env_int("RANK", 0)
    if rank != 0:
        return 0
    import oracle
    from celerite_b200 import synth
    chk = oracle.best()
    cores = os.cpu_count() or 1
    N, Jr, Jc = wl["N"], wl["Jr"], wl["Jc"]
    J = Jr + 2 * Jc
    # about 0.1 us per point-term per core: size each step to roughly 1 s of wall time
    sample = int(min(wl["B"], max(256, cores * int(1.0e7 / (N * J)))))
    b = synth.make_batch(sample, N, Jr, Jc, seed=42)
    P = b["coeffs"].shape[1]
    call = lambda: chk.batched_loglik(sample, N, N, Jr, Jc, b["coeffs"], P, b["t"], b["diag"], b["y"],  # noqa: E731
                                      nthreads=cores)
    for _ in range(args.warmup):
        call()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = call()
    dt = (time.perf_counter() - t0) / args.steps
    assert np.all(out[:, 2] == 0)
    value = sample * N * J / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value,
        "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl["desc"], "sample": "%d of %d series per step" % (sample, wl["B"])},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": chk.kind,
                         "sample": "%d series x N=%d per step, %d threads, one series per task" % (sample, N, cores)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
# shared plumbing of the GPU arm
# ------------------------------------------------------------------------------------------------
class Ctx:
    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank, self.world, self.local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device (celerite_b200 has no CPU fallback)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.stream = torch.cuda.current_stream(self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def timed(self, fn, steps):
        """ms per step: CUDA events on the launching stream, barrier + synchronize on both sides,
        max over ranks."""
        torch = self.torch
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        for _ in range(steps):
            fn()
        e1.record(self.stream)
        self.barrier()
        ms = e0.elapsed_time(e1)
        if self.world > 1:
            tms = torch.tensor([ms], dtype=torch.float64, device=self.dev)
            self.dist.all_reduce(tms, op=self.dist.ReduceOp.MAX)
            ms = float(tms.item())
        return ms / steps

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def h2d_roof(ctx, host_tensors, reps=2):
    """GB/s a bare cudaMemcpyAsync of these pinned buffers reaches right now (all ranks copying at
    once): the roof the end-to-end number can be read against."""
    torch = ctx.torch
    devs = [torch.empty_like(h, device=ctx.dev) for h in host_tensors]
    nbytes = sum(h.numel() * h.element_size() for h in host_tensors)

    def copy():
        for h, d in zip(host_tensors, devs):
            d.copy_(h, non_blocking=True)
    copy()
    ms = ctx.timed(copy, reps)
    del devs
    return nbytes / (ms * 1e-3) / 1e9


# ------------------------------------------------------------------------------------------------
# batched workloads: cfg2 (headline), cfg2m, cfg3, cfg5
# ------------------------------------------------------------------------------------------------
class BatchData:
    """Synthetic batch generated straight into pinned host buffers, then copied to HBM."""

    def __init__(self, ctx, B, N, seed):
        torch = ctx.torch
        from celerite_b200 import synth
        self.B, self.N = B, N
        self.ht = torch.empty((B, N), dtype=torch.float64).pin_memory()
        self.hd = torch.empty((B, N), dtype=torch.float64).pin_memory()
        self.hy = torch.empty((B, N), dtype=torch.float64).pin_memory()
        self.rng = np.random.Generator(np.random.PCG64(seed))
        synth.fill_series(self.rng, self.ht.numpy(), self.hd.numpy(), self.hy.numpy())
        self.dt, self.dd, self.dy = self.ht.to(ctx.dev), self.hd.to(ctx.dev), self.hy.to(ctx.dev)


def bench_batched(ctx, wl, steps, warmup, e2e_steps, data=None, cpu=True, with_grad_extra=False, clocks=False,
                  workload_key="cfg2", coeff_seed=None):
    """One batched workload on every rank; returns the JSON line as a dict (rank 0) or None."""
    torch = ctx.torch
    from celerite_b200 import synth, _capi
    from celerite_b200.batched import BatchedSolver
    dev, stream, world, rank = ctx.dev, ctx.stream, ctx.world, ctx.rank
    B, N, Jr, Jc = wl["B"], wl["N"], wl["Jr"], wl["Jc"]
    J, P = Jr + 2 * Jc, 1 + 2 * Jr + 4 * Jc
    if data is None:
        data = BatchData(ctx, B, N, 42 + rank)
    ht, hd, hy, dt_, dd_, dy_ = data.ht, data.hd, data.hy, data.dt, data.dd, data.dy
    crng = np.random.Generator(np.random.PCG64((coeff_seed if coeff_seed is not None else 4242) + rank))
    coeffs_np = synth.make_coeffs(crng, B, Jr, Jc)
    hc = torch.from_numpy(coeffs_np).pin_memory()
    dc_ = hc.to(dev)
    h_out = torch.empty((3, B), dtype=torch.float64).pin_memory()
    h_status = torch.empty(B, dtype=torch.int32).pin_memory()
    d_out = torch.empty((3, B), dtype=torch.float64, device=dev)
    d_status = torch.empty(B, dtype=torch.int32, device=dev)
    d_gval = torch.empty(B, dtype=torch.float64, device=dev)
    d_grad = torch.empty((B, P), dtype=torch.float64, device=dev)

    bs = BatchedSolver(ctx.local)
    lib, h = bs._lib, bs._h
    v = lambda x: C.c_void_p(x)  # noqa: E731
    lib.cel_set_stream(h, C.c_void_p(stream.cuda_stream or 1))

    def step_ll():
        bs.raw_log_likelihood(B, N, N, Jr, Jc, dc_.data_ptr(), P, dt_.data_ptr(), dd_.data_ptr(), dy_.data_ptr(),
                              d_out[0].data_ptr(), d_out[1].data_ptr(), d_out[2].data_ptr(), d_status.data_ptr(),
                              flags=0, stream=stream.cuda_stream)

    def step_ll_host():
        bs.raw_log_likelihood(B, N, N, Jr, Jc, hc.data_ptr(), P, ht.data_ptr(), hd.data_ptr(), hy.data_ptr(),
                              h_out[0].data_ptr(), h_out[1].data_ptr(), h_out[2].data_ptr(), h_status.data_ptr(),
                              flags=_capi.HOST_POINTERS, stream=stream.cuda_stream)

    def step_grad():
        rc = lib.cel_batched_grad_log_likelihood(h, B, N, N, Jr, Jc, v(dc_.data_ptr()), P, v(dt_.data_ptr()),
                                                 v(dd_.data_ptr()), v(dy_.data_ptr()), v(d_gval.data_ptr()),
                                                 v(d_grad.data_ptr()), v(d_status.data_ptr()), 0)
        _capi.check(rc, h)

    is_grad, materialise = bool(wl.get("grad")), bool(wl.get("materialise"))
    step_device = step_grad if is_grad else step_ll
    if materialise:
        m_phi = torch.empty((B, N - 1, J), dtype=torch.float64, device=dev)
        m_u = torch.empty((B, N - 1, J), dtype=torch.float64, device=dev)
        m_W = torch.empty((B, N, J), dtype=torch.float64, device=dev)
        m_D = torch.empty((B, N), dtype=torch.float64, device=dev)

        def step_device():  # noqa: F811
            rc = lib.cel_batched_compute(h, B, N, N, Jr, Jc, v(dc_.data_ptr()), P, v(dt_.data_ptr()),
                                         v(dd_.data_ptr()), v(m_phi.data_ptr()), v(m_u.data_ptr()), v(m_W.data_ptr()),
                                         v(m_D.data_ptr()), v(d_out[0].data_ptr()), v(d_status.data_ptr()))
            _capi.check(rc, h)

    # ---- device-resident metric -------------------------------------------------------------
    for _ in range(warmup):
        step_device()
    torch.cuda.synchronize(dev)
    assert int(d_status.sum().item()) == 0, "a synthetic series failed to factorise"
    launches0 = bs.launch_count()
    sampler = ClockSampler(ctx.local) if (clocks and rank == 0) else None
    if sampler:
        sampler.start()
    ms_step = ctx.timed(step_device, steps)
    clk = sampler.stop() if sampler else None
    launches = bs.launch_count() - launches0
    value = world * B * N * J / (ms_step * 1e-3)
    peak, peak_src = measured_peak_gbs()

    line = {
        "metric": METRIC_GRAD if is_grad else (METRIC_COMPUTE if materialise else METRIC),
        "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl["desc"], "series_per_gpu": B, "N": N, "J": J,
                   "sharding": "batch index, no data-path collective"},
        "gpu_launches": launches,
    }
    if clk is not None:
        line["clocks"] = clk

    if materialise:
        alg_bytes = B * N * 16 + B * (2 * (N - 1) * J + N * J + N) * 8 + B * 12
        achieved = alg_bytes / (ms_step * 1e-3) / 1e9
        line["config"]["inputs"] = ("%.2f GB read + %.2f GB written per step per GPU, larger than the 126 MB L2"
                                    % (2 * B * N * 8 / 1e9, (alg_bytes - 2 * B * N * 8) / 1e9))
        line["roofline"] = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                            "traffic": ncu_traffic(workload_key), "peak_source": peak_src,
                            "kernel": "k_compute_lockstep", "algorithmic_bytes_per_launch": alg_bytes}
        line["e2e"] = {"value": None, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                       "note": "device-resident API only (the factor is 7.3 GB)"}
        if cpu and rank == 0:
            # parity of the written factor against the reference's, first series
            import oracle
            chk = oracle.best()
            j, ar, cr, ac, bc, cc, dcc = synth.split_row(coeffs_np[0], Jr, Jc)
            f = chk.factor(oracle.Problem(j, ar, cr, ac, bc, cc, dcc, ht.numpy()[0], hd.numpy()[0]))
            line["cpu_baseline"] = {"kind": chk.kind, "sample": "factor of series 0",
                                    "max_rel_diff_vs_gpu": max(pure_rel(m_D[0].cpu().numpy(), f.D),
                                                               pure_rel(m_W[0].cpu().numpy(), np.asarray(f.W).reshape(N, J)))}
            if hasattr(chk, "release"):
                chk.release(f)
        return line if rank == 0 else None

    line["config"]["inputs"] = "%.2f GB per step per GPU, larger than the 126 MB L2 (no flush needed)" % (3 * B * N * 8 / 1e9)
    # ---- end-to-end through the C ABI with host buffers -------------------------------------
    step_host = step_ll_host
    if is_grad:
        h_gval = torch.empty(B, dtype=torch.float64).pin_memory()
        h_grad = torch.empty((B, P), dtype=torch.float64).pin_memory()

        def step_host():  # noqa: F811
            rc = lib.cel_batched_grad_log_likelihood(h, B, N, N, Jr, Jc, v(hc.data_ptr()), P, v(ht.data_ptr()),
                                                     v(hd.data_ptr()), v(hy.data_ptr()), v(h_gval.data_ptr()),
                                                     v(h_grad.data_ptr()), v(h_status.data_ptr()), _capi.HOST_POINTERS)
            _capi.check(rc, h)
    step_host()  # warm-up (allocates the staging buffers)
    e2e_ms = ctx.timed(step_host, max(1, e2e_steps))
    e2e_value = world * B * N * J / (e2e_ms * 1e-3)
    torch.cuda.synchronize(dev)
    # the e2e path must give the same answer as the device-resident path
    same = bool(torch.equal(h_grad.to(dev), d_grad)) if is_grad else bool(torch.equal(h_out.to(dev), d_out))
    roof = h2d_roof(ctx, [ht, hd, hy])
    h2d = 3 * B * N * 8 + B * P * 8

    # ---- the same batch with the gradient (the "log-lik+grad" half of the metric's name) ----
    grad_extra = None
    if with_grad_extra and not is_grad:
        step_grad()
        torch.cuda.synchronize(dev)
        g_ms = ctx.timed(step_grad, 3)
        alg_g = B * N * 24 + B * (P + 1) * 8
        grad_extra = {"value": world * B * N * J / (g_ms * 1e-3), "unit": UNIT, "ms_per_step": g_ms,
                      "roofline": {"achieved": alg_g / (g_ms * 1e-3) / 1e9, "frac": alg_g / (g_ms * 1e-3) / 1e9 / peak,
                                   "traffic": ncu_traffic(workload_key + "_grad"), "algorithmic_bytes_per_launch": alg_g,
                                   "kernel": ("k_loglik_lockstep2<CKPT> + " + ("k_grad2_reverse" if J == 2 else "k_gradN_reverse")) if J <= 4 else "k_grad_coop"},
                      "what": "cel_batched_grad_log_likelihood on the same batch (value + 1+2Jr+4Jc derivatives per series)"}

    if rank != 0:
        return None
    if is_grad:
        alg_bytes = B * N * 24 + B * (P + 1) * 8
        kernel, d2h = ("k_grad_lockstep" if (J <= 4 and B >= 8192) else "k_grad_coop"), B * (P + 1) * 8 + B * 4
    else:
        alg_bytes = B * N * 24 + B * 28          # t, diag, y in; logdet, quad, ll, status out
        kernel, d2h = ("k_loglik_lockstep" if J <= 6 else "k_factor_coop"), B * 28
    achieved = alg_bytes / (ms_step * 1e-3) / 1e9
    line["roofline"] = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                        "traffic": ncu_traffic(workload_key), "peak_source": peak_src, "kernel": kernel,
                        "algorithmic_bytes_per_launch": alg_bytes}
    if is_grad:
        line["roofline"]["note"] = "fp64-pipe/latency bound (1,024 independent recurrences); HBM fraction is not the limiter"
    line["e2e"] = {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": h2d,
                   "d2h_bytes_per_step": d2h, "matches_device_path": same,
                   "h2d_roof_gbs": roof, "h2d_achieved_gbs": h2d / (e2e_ms * 1e-3) / 1e9,
                   "frac_of_h2d_roof": h2d / (e2e_ms * 1e-3) / 1e9 / roof}
    if grad_extra:
        line["with_grad"] = grad_extra
    if cpu:
        import oracle
        chk = oracle.best()
        cores = os.cpu_count() or 1
        if is_grad:
            # the AD reference costs ~(1+P) factorisations per series: bound the sample in B and in N, and
            # run the GPU on the same truncated series for the parity figure
            sample = int(min(B, cores))
            Nc = int(min(N, max(2000, 1.6e6 / (J * J))))
            tN, dN, yN = (np.ascontiguousarray(a.numpy()[:sample, :Nc]) for a in (ht, hd, hy))
            t0 = time.perf_counter()
            gv, gg, gs = chk.batched_grad(sample, Nc, Nc, Jr, Jc, coeffs_np[:sample], P, tN, dN, yN, nthreads=cores)
            cdt = time.perf_counter() - t0
            ov, og, ost = bs.grad_log_likelihood(coeffs_np[:sample], tN, dN, yN, Jr, Jc)
            line["cpu_baseline"] = {"value": sample * Nc * J / cdt, "unit": UNIT, "cores": cores, "kind": chk.kind,
                                    "sample": "first %d series truncated to N=%d, forward-mode AD gradient "
                                              "(solver.cpp:347-463), %d threads" % (sample, Nc, cores),
                                    "max_rel_diff_vs_gpu": pure_rel(ov - 0.5 * np.pi * np.log(Nc), gv),   # solver.cpp:415 constant
                                    "max_rel_diff_grad_vs_gpu": grad_rel(og, gg)}
        else:
            sample = int(min(B, max(256, cores * 1024)))
            call = lambda: chk.batched_loglik(sample, N, N, Jr, Jc, coeffs_np[:sample], P,  # noqa: E731
                                              ht.numpy()[:sample], hd.numpy()[:sample], hy.numpy()[:sample],
                                              nthreads=cores)
            call()
            t0 = time.perf_counter()
            reps = 3
            for _ in range(reps):
                ref = call()
            cdt = (time.perf_counter() - t0) / reps
            got = d_out[:, :sample].cpu().numpy()
            line["cpu_baseline"] = {"value": sample * N * J / cdt, "unit": UNIT, "cores": cores, "kind": chk.kind,
                                    "sample": "first %d of %d series, %d threads, one series per task" % (sample, B, cores),
                                    "max_rel_diff_vs_gpu": max(pure_rel(got[0], ref[:, 0]), pure_rel(got[1], ref[:, 1]))}
            try:   # the reference's own round-off floor on the same sample: FMA-contracted build against the plain-C restatement
                alt = oracle.Port().batched_loglik(sample, N, N, Jr, Jc, coeffs_np[:sample], P, ht.numpy()[:sample],
                                                   hd.numpy()[:sample], hy.numpy()[:sample], nthreads=cores)
                line["cpu_baseline"]["max_rel_diff_cpu_vs_cpu"] = max(pure_rel(alt[:, 0], ref[:, 0]), pure_rel(alt[:, 1], ref[:, 1]))
            except Exception:   # noqa: BLE001
                pass
    return line


# ------------------------------------------------------------------------------------------------
# cfg4: one long series
# ------------------------------------------------------------------------------------------------
def cfg4_series(N, jc=4, seed=42):
    """The one long series of config 4: N points, jc SHO terms (J = 2*jc)."""
    from celerite_b200 import synth
    rng = np.random.Generator(np.random.PCG64(seed))
    coeffs = synth.make_coeffs(rng, 1, 0, jc)
    t = np.empty((1, N)); d = np.empty((1, N)); y = np.empty((1, N))
    synth.fill_series(rng, t, d, y)
    z = rng.standard_normal(N)
    return coeffs[0], t[0], d[0], y[0], z


def cfg4_cpu(chk, coeffs, t, d, y, z, jc, n_sample):
    """reference solve + dot_L on ONE host thread (a single series is inherently serial on the
    CPU), on the first n_sample points; returns seconds for (solve + dot_L), compute untimed."""
    import oracle
    from celerite_b200 import synth
    j, ar, cr, ac, bc, cc, dc = synth.split_row(coeffs, 0, jc)
    p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, t[:n_sample], d[:n_sample])
    f = chk.factor(p)
    t0 = time.perf_counter()
    chk.solve(f, y[:n_sample])
    chk.dot_L(f, z[:n_sample])
    dt = time.perf_counter() - t0
    if hasattr(chk, "release"):
        chk.release(f)
    return dt


def cfg4_reference(args):
    import oracle
    N, jc = (args.batch or 10_000_000), (args.jc or 4)
    J = 2 * jc
    chk = oracle.best()
    n_sample = min(N, 1_000_000)
    coeffs, t, d, y, z = cfg4_series(n_sample, jc)
    for _ in range(max(0, min(args.warmup, 1))):
        cfg4_cpu(chk, coeffs, t, d, y, z, jc, n_sample)
    steps = max(1, min(args.steps, 5))
    dt = sum(cfg4_cpu(chk, coeffs, t, d, y, z, jc, n_sample) for _ in range(steps)) / steps
    value = 2.0 * n_sample * J / dt
    desc = "cfg4: single series N=%d, J=%d (%d SHOTerms), apply_inverse + dot_L via chunked scan" % (N, J, jc)
    print(json.dumps({
        "impl": "reference", "metric": METRIC_CFG4, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "sample": "first %d points" % n_sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": chk.kind,
                         "sample": "first %d of %d points, 1 thread (one series is serial on the CPU)" % (n_sample, N)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))
    return 0


def bench_cfg4(ctx, steps, warmup, N=10_000_000, jc=4, cpu=True, graph=False, clocks=False, blocks=1):
    """config 4: one series, N = 1e7, J = 8: apply_inverse + dot_L as chunked scans; at world > 1 the
    time axis is split over the ranks (strong scaling) and the span carries travel by NCCL."""
    torch = ctx.torch
    from celerite_b200 import longseries as ls, solver, synth
    J = 2 * jc
    rank, world, local, dev, stream = ctx.rank, ctx.world, ctx.local, ctx.dev, ctx.stream
    desc = "cfg4: single series N=%d, J=%d (%d SHOTerms), apply_inverse + dot_L via chunked scan" % (N, J, jc)
    coeffs, t, d, y, z = cfg4_series(N, jc)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(coeffs, 0, jc)
    halo = ls.safe_halo(float(np.min(cc)), float(np.mean(np.diff(t[:100000]))), tol=1e-18, cap=N // (2 * world))
    if world > 1:
        eng = ls.DistributedLongSeries(local)
        eng.compute(j, ar, cr, ac, bc, cc, dc, t, d, halo)
        lo, hi = eng.lo, eng.hi
        loc = eng.local
    else:
        eng = ls.LongSeries(local)
        eng.compute(j, ar, cr, ac, bc, cc, dc, t, d)
        lo, hi = 0, N
        loc = eng
    yb = torch.from_numpy(y[lo:hi]).to(dev)
    zb = torch.from_numpy(z[lo:hi]).to(dev)

    graphed = False
    if graph and world == 1:
        try:
            g_solve = ls.CapturedSweep(eng, "solve", yb)
            g_dotl = ls.CapturedSweep(eng, "dot_L", zb)
            graphed = True
        except Exception as e:   # noqa: BLE001
            print("cuda graph capture failed (%s); timing plain calls" % e, file=sys.stderr)
            torch.cuda.synchronize(dev)

    def step():
        if graphed:
            return g_solve(yb), g_dotl(zb)
        return eng.solve(yb), eng.dot_L(zb)

    for _ in range(max(3, warmup)):
        a_inv, lz = step()
    l0 = loc.launch_count()
    sampler = ClockSampler(local) if (clocks and rank == 0) else None
    if sampler:
        sampler.start()
    out = {}

    def timed_step():
        out["r"] = step()
    ms = ctx.timed(timed_step, steps)
    for _ in range(blocks - 1):       # as an extra of the default run: the faster of `blocks` blocks of K steps
        l0 = loc.launch_count()       # (a single block has been seen 2.7x slow once on a shared box)
        ms = min(ms, ctx.timed(timed_step, steps))
    a_inv, lz = out["r"]
    clk = sampler.stop() if sampler else None
    launches = (g_solve.launches + g_dotl.launches) * steps if graphed else (loc.launch_count() - l0)
    value = 2.0 * N * J / (ms * 1e-3)
    # the dominant kernel alone: one chunk sweep (pass 1 of the forward solve; the call also runs
    # the two one-warp walk kernels of the middle stage, ~15 us)
    cols = yb[None, :].contiguous()
    loc.summarize(ls.OP_SOLVE_FORWARD, cols)
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    k0.record(stream)
    for _ in range(5):
        loc.summarize(ls.OP_SOLVE_FORWARD, cols)
    k1.record(stream)
    torch.cuda.synchronize(dev)
    sweep_ms = k0.elapsed_time(k1) / 5

    line = None
    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        alg = (hi - lo) * ((6 * J + 5) * 8 + (3 * J + 3) * 8)     # solve + dot_L, SURVEY.md section 8d
        alg_sweep = (hi - lo) * (3 * J + 2) * 8                   # one forward sweep: phi, u, W rows + D + rhs
        line = {
            "metric": METRIC_CFG4, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps,
            "warmup": max(3, warmup), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": desc, "N": N, "J": J, "halo": int(halo) if world > 1 else 0, "cuda_graph": graphed,
                       "timed_blocks": blocks,
                       "sharding": "time axis split into %d spans; one all-gather of (J^2+J) doubles per sweep" % world,
                       "collectives": getattr(eng, "collective_backend", "none") if world > 1 else "none",
                       "inputs": "factor %.2f GB per rank, larger than the 126 MB L2" % ((hi - lo) * (3 * J + 1) * 8 / 1e9)},
            "roofline": {"bound": "hbm", "achieved": alg_sweep / (sweep_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg_sweep / (sweep_ms * 1e-3) / 1e9 / peak, "traffic": ncu_traffic("cfg4"),
                         "peak_source": peak_src, "kernel": "k_chunk_forward (one sweep; a step launches six)",
                         "algorithmic_bytes_per_launch": alg_sweep, "kernel_ms": sweep_ms,
                         "whole_step": {"algorithmic_bytes": alg, "achieved": alg / (ms * 1e-3) / 1e9,
                                        "frac": alg / (ms * 1e-3) / 1e9 / peak,
                                        "note": "one streaming pass of solve + dot_L as the serial algorithm "
                                                "would read it; the two-pass scan reads the factor twice"}},
            "gpu_launches": launches,
        }
        if clk is not None:
            line["clocks"] = clk
    # e2e through the drop-in host API (single GPU only: cel_solve / cel_dot_L with host arrays)
    if world == 1:
        s = solver.CholeskySolver(local)
        E0, E2 = np.empty(0), np.empty((0, 0))
        s.compute(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, t, d)
        s.solve(y); s.dot_L(z)
        t0 = time.perf_counter()
        reps = 2
        for _ in range(reps):
            zi = s.solve(y)
            lzi = s.dot_L(z)  # noqa: F841
        edt = (time.perf_counter() - t0) / reps
        line["e2e"] = {"value": 2.0 * N * J / edt, "unit": UNIT, "ms_per_step": edt * 1e3,
                       "h2d_bytes_per_step": 2 * N * 8, "d2h_bytes_per_step": 2 * N * 8,
                       "matches_device_path": bool(np.allclose(zi[:, 0], a_inv.cpu().numpy(), rtol=1e-12, atol=0)),
                       "note": "cel_solve + cel_dot_L with pageable host arrays, wall clock"}
        del s
    elif rank == 0:
        line["e2e"] = {"value": None, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                       "note": "the drop-in host API is single-GPU; multi-GPU spans are device-resident"}
    if rank == 0 and cpu:
        import oracle
        chk = oracle.best()
        n_sample = min(N, 1_000_000)
        cdt = cfg4_cpu(chk, coeffs, t, d, y, z, jc, n_sample)
        p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, t[:n_sample], d[:n_sample])
        f = chk.factor(p)
        zr = chk.solve(f, y[:n_sample])[:, 0]
        lr = chk.dot_L(f, z[:n_sample])[:, 0]
        m = min(n_sample, hi - lo) // 2      # away from the truncated end of the CPU sample
        got, gotl = a_inv[:m].cpu().numpy(), lz[:m].cpu().numpy()
        line["cpu_baseline"] = {"value": 2.0 * n_sample * J / cdt, "unit": UNIT, "cores": 1, "kind": chk.kind,
                                "sample": "first %d of %d points, 1 thread (one series is serial on the CPU)" % (n_sample, N),
                                "max_rel_diff_vs_gpu": max(float(np.max(np.abs(got - zr[:m])) / np.max(np.abs(zr[:m]))),
                                                           float(np.max(np.abs(gotl - lr[:m])) / np.max(np.abs(lr[:m])))),
                                "rel_diff_is": "||gpu - ref||_inf / ||ref||_inf over the first %d points, solve and dot_L" % m}
        if hasattr(chk, "release"):
            chk.release(f)
    return line


# ------------------------------------------------------------------------------------------------
# cfg1: one GP through the drop-in solver class, per-call wall clock
# ------------------------------------------------------------------------------------------------
def bench_cfg1(ctx, reps=30):
    """config 1: RealTerm + ComplexTerm (J=3), N=4,096: compute + log_likelihood + grad_log_likelihood,
    the calls celerite.GP makes (celerite.py:144-157, 211-216, 270-284), host arrays in and out,
    examples/benchmark/run.py style (wall clock per call, median of repeats)."""
    import oracle
    from celerite_b200 import synth
    try:
        from celerite_b200 import solver_ext as mod
        binding = "celerite_b200.solver_ext (pybind11 over the C ABI)"
    except ImportError:
        from celerite_b200 import solver as mod
        binding = "celerite_b200.solver (ctypes over the C ABI)"
    N, Jr, Jc = 4096, 1, 1
    J = Jr + 2 * Jc
    b = synth.make_batch(1, N, Jr, Jc, seed=42)
    j, ar, cr, ac, bc, cc, dc = synth.split_row(b["coeffs"][0], Jr, Jc)
    t, d, y = b["t"][0], b["diag"][0], b["y"][0]
    E0, E2 = np.empty(0), np.empty((0, 0))
    s = mod.CholeskySolver()

    def median_of(fn, n):
        fn()
        ts = []
        for _ in range(n):
            t0 = time.perf_counter()
            r = fn()
            ts.append(time.perf_counter() - t0)
        return float(np.median(ts)), r
    g_comp, _ = median_of(lambda: s.compute(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, t, d), reps)
    g_ll, gll = median_of(lambda: (s.dot_solve(y), s.log_determinant()), reps)
    g_grad, gg = median_of(lambda: s.grad_log_likelihood(j, ar, cr, ac, bc, cc, dc, E0, E2, E2, t, y, d), reps)
    chk = oracle.best()
    p = oracle.Problem(j, ar, cr, ac, bc, cc, dc, t, d)
    fs = []

    def cpu_factor():
        f = chk.factor(p)
        fs.append(f)
        return f
    c_comp, f = median_of(cpu_factor, reps)
    c_ll, cll = median_of(lambda: (chk.dot_solve(f, y), f.log_det), reps)
    c_grad, cg = median_of(lambda: chk.grad(p, y), 5)
    cval, cgrad = cg[1], cg[2]          # (status, value, gradient)
    gval = gg[0]                          # both carry the reference's -0.5*pi*log(N) constant (solver.cpp:415)
    gpu_total, cpu_total = g_comp + g_ll + g_grad, c_comp + c_ll + c_grad
    res = {
        "workload": "cfg1: single GP RealTerm+ComplexTerm (J=3), N=4,096: compute + log_likelihood + grad_log_likelihood per call",
        "binding": binding, "metric": "GP compute+log-lik+grad point-terms/sec (N*J / wall time of the three calls)",
        "value": N * J / gpu_total, "unit": UNIT, "n_gpus": 1,
        "gpu_ms_per_call": {"compute": g_comp * 1e3, "log_likelihood": g_ll * 1e3, "grad_log_likelihood": g_grad * 1e3},
        "cpu_ms_per_call": {"compute": c_comp * 1e3, "log_likelihood": c_ll * 1e3, "grad_log_likelihood": c_grad * 1e3},
        "cpu_baseline": {"value": N * J / cpu_total, "unit": UNIT, "cores": 1, "kind": chk.kind,
                         "sample": "the same series, 1 thread, median of repeated calls",
                         "max_rel_diff_vs_gpu": max(pure_rel(gll[0], cll[0]), pure_rel(gll[1], cll[1]), pure_rel(gval, cval)),
                         "max_rel_diff_grad_vs_gpu": grad_rel(gg[1], cgrad)},
        "e2e": {"value": N * J / gpu_total, "unit": UNIT, "h2d_bytes_per_step": (2 + 3) * N * 8, "d2h_bytes_per_step": 8 * (2 + 8),
                "note": "the measurement itself is end to end: host arrays in, host results out, wall clock"},
        "note": "one series of 4,096 points is a chain of 4,096 dependent fp64 steps: latency, not throughput "
                "(DESIGN.md states where a single CPU core wins)",
        "gpu_launches": int(s.launch_count()) if hasattr(s, "launch_count") else None,
    }
    if hasattr(chk, "release"):
        for ff in fs:
            chk.release(ff)
    return res


def compact(line):
    """An extras entry: the headline's keys minus the boilerplate."""
    if line is None:
        return None
    drop = ("higher_is_better", "vs_baseline", "dtype", "data", "clocks")
    out = {k: v for k, v in line.items() if k not in drop}
    if "config" in out:
        out["workload"] = out["config"].get("workload")
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="all", choices=["all", "cfg1", "cfg4"] + sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="override series per GPU / series length of cfg4 (debug)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline config only")
    ap.add_argument("--jc", type=int, default=0, help="cfg3: number of SHO terms (J = 2*jc)")
    ap.add_argument("--no-grad", action="store_true", help="skip the extra log-lik+grad measurement")
    ap.add_argument("--graph", action="store_true", help="cfg4, one GPU: time CUDA-graph replays of the two calls")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    name = "cfg2" if args.workload == "all" else args.workload

    if args.impl == "reference":
        if name == "cfg4":
            return cfg4_reference(args) if env_int("RANK", 0) == 0 else 0
        return run_reference(args, dict(WORKLOADS[name if name in WORKLOADS else "cfg2"]))

    ctx = Ctx()
    cpu = not args.no_cpu_baseline
    if name == "cfg4":
        line = bench_cfg4(ctx, args.steps, args.warmup, N=args.batch or 10_000_000, jc=args.jc or 4, cpu=cpu,
                          graph=args.graph, clocks=True)
        if ctx.rank == 0:
            print(json.dumps(line))
        ctx.close()
        return 0
    if name == "cfg1":
        if ctx.rank == 0:
            print(json.dumps(bench_cfg1(ctx)))
        ctx.close()
        return 0

    wl = dict(WORKLOADS[name])
    if args.batch:
        wl["B"] = args.batch
        wl["desc"] += " [batch overridden to %d]" % args.batch
    if args.jc:
        wl["Jc"] = args.jc
        wl["desc"] += " [jc=%d]" % args.jc
    data = BatchData(ctx, wl["B"], wl["N"], 42 + ctx.rank)
    line = bench_batched(ctx, wl, args.steps, args.warmup, args.e2e_steps, data=data, cpu=cpu,
                         with_grad_extra=not args.no_grad, clocks=True, workload_key=name)

    # ---- the other configs of BASELINE.json, same process, same line -------------------------
    if args.workload == "all" and not args.no_extras:
        torch = ctx.torch
        extras = {}

        def attempt(key, fn):
            try:
                t0 = time.perf_counter()
                r = fn()
                if ctx.rank == 0 and r is not None:
                    r["bench_wall_s"] = round(time.perf_counter() - t0, 2)
                    extras[key] = r
            except Exception as e:   # noqa: BLE001  an extra must never cost the headline
                if ctx.rank == 0:
                    extras[key] = {"error": "%s: %s" % (type(e).__name__, e)}
            ctx.barrier()
            torch.cuda.empty_cache()

        if ctx.world == 1:
            attempt("cfg2m", lambda: compact(bench_batched(ctx, dict(WORKLOADS["cfg2m"]), 10, 3, 0, data=data, cpu=cpu,
                                                           workload_key="cfg2m")))
        del data
        torch.cuda.empty_cache()
        wl5 = dict(WORKLOADS["cfg5"])
        wl5["desc"] = ("cfg5: 10^6 series x N=1,000 sharded by batch index, 125,000 series per GPU (%d of 8 GPUs here), "
                       "two SHOTerms (J=4), log_likelihood" % ctx.world)
        attempt("cfg5", lambda: compact(bench_batched(ctx, wl5, 10, 3, 2, cpu=cpu, with_grad_extra=not args.no_grad,
                                                      workload_key="cfg5", coeff_seed=5151)))
        attempt("cfg4", lambda: compact(bench_cfg4(ctx, 10, 3, cpu=cpu, blocks=2)))
        if ctx.world == 1:
            holder = {}

            def cfg3(jc):
                if "d" not in holder:
                    holder["d"] = BatchData(ctx, 1024, 100000, 333)
                wl3 = dict(WORKLOADS["cfg3"])
                wl3["Jc"] = jc
                wl3["desc"] = "cfg3: 1,024 series x N=100,000, J=%d (%d SHOTerms), grad_log_likelihood" % (2 * jc, jc)
                return compact(bench_batched(ctx, wl3, 3, 3, 1, data=holder["d"], cpu=cpu, workload_key="cfg3_J%d" % (2 * jc),
                                             coeff_seed=3000 + jc))
            for jc in (1, 4, 16):
                attempt("cfg3_J%d" % (2 * jc), lambda jc=jc: cfg3(jc))
            holder.clear()
            attempt("cfg1", lambda: bench_cfg1(ctx))
        if ctx.rank == 0:
            line["extras"] = extras
            if ctx.world > 1:
                line["extras"]["skipped"] = "cfg1, cfg2m, cfg3: single-GPU entries, reported by the --gpus 1 run"
    if ctx.rank == 0:
        print(json.dumps(line))
    ctx.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
