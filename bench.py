#!/usr/bin/env python
"""bench.py — throughput of the batched GP log-likelihood hot path on B200 (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2]

A "step" is one pass of the hot path (fused factorisation + log-likelihood) over one batch of
synthetic irregularly-sampled series.  Workloads (BASELINE.json `configs`):
    cfg2  65,536 series x N=2,000, one SHO term (J=2), log_likelihood      <- default, the metric's config
    cfg5  per-GPU slice of 10^6 series x N=1,000, two SHO terms (J=4)
metric = point-terms/s = B*N*J / device time, whole job over all ranks (weak scaling: every rank
sweeps its own batch of the named size; no data-path collective).

Keys beyond the base contract:
    roofline      algorithmic HBM bytes (24 B/point in, 32 B/series out) / CUDA-event time of the
                  kernel, against MEASURED_PEAKS.json
    cpu_baseline  the UNMODIFIED reference (oracle/_ref) timed on this box's host cores on a
                  bounded sample of the same workload
    e2e           same metric through the C ABI with HOST (pinned) buffers: H2D of every input
                  and D2H of every result inside the timed region
`--impl reference` times the reference's own CPU implementation (all host threads, one series
per task) on a bounded sample of the same config.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

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


def run_reference(args, wl):
    """--impl reference: the reference's CPU solver on all host threads, bounded sample."""
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
        "impl": "reference", "metric": "GP log-likelihood point-terms/sec (N*J*batch / time)", "value": value,
        "unit": "point-terms/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl["desc"], "sample": "%d of %d series per step" % (sample, wl["B"])},
        "cpu_baseline": {"value": value, "unit": "point-terms/s", "cores": cores, "kind": chk.kind,
                         "sample": "%d series x N=%d per step, %d threads, one series per task" % (sample, N, cores)},
        "e2e": {"value": value, "unit": "point-terms/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


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


def run_cfg4(args):
    """config 4: one series, N = 1e7, J = 8: apply_inverse + dot_L as chunked scans; the time
    axis is split over the ranks (strong scaling), carries exchanged with one all-gather per sweep."""
    N, jc = (args.batch or 10_000_000), (args.jc or 4)
    J = 2 * jc
    desc = "cfg4: single series N=%d, J=%d (%d SHOTerms), apply_inverse + dot_L via chunked scan" % (N, J, jc)
    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    metric = "GP apply_inverse + dot_L point-terms/sec (2*N*J / device time)"
    if args.impl == "reference":
        if rank != 0:
            return 0
        import oracle
        chk = oracle.best()
        n_sample = min(N, 1_000_000)
        coeffs, t, d, y, z = cfg4_series(n_sample, jc)
        for _ in range(max(0, min(args.warmup, 1))):
            cfg4_cpu(chk, coeffs, t, d, y, z, jc, n_sample)
        steps = max(1, min(args.steps, 5))
        dt = sum(cfg4_cpu(chk, coeffs, t, d, y, z, jc, n_sample) for _ in range(steps)) / steps
        value = 2.0 * n_sample * J / dt
        print(json.dumps({
            "impl": "reference", "metric": metric, "value": value, "unit": "point-terms/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": desc, "sample": "first %d points" % n_sample},
            "cpu_baseline": {"value": value, "unit": "point-terms/s", "cores": 1, "kind": chk.kind,
                             "sample": "first %d of %d points, 1 thread (one series is serial on the CPU)" % (n_sample, N)},
            "e2e": {"value": value, "unit": "point-terms/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}))
        return 0

    import torch
    import torch.distributed as dist
    from celerite_b200 import longseries as ls, solver, synth
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
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
    stream = torch.cuda.current_stream(dev)

    # optional: the two calls recorded as CUDA graphs (one GPU only; measured 2.54 against 2.51 ms:
    # nothing to gain when the sweeps themselves take 2 ms)
    graphed = False
    if getattr(args, "graph", False) and world == 1:
        try:
            g_solve = ls.CapturedSweep(eng, "solve", yb)
            g_dotl = ls.CapturedSweep(eng, "dot_L", zb)
            graphed = True
        except Exception as e:   # noqa: BLE001
            if rank == 0:
                print("cuda graph capture failed (%s); timing plain calls" % e, file=sys.stderr)
            torch.cuda.synchronize(dev)

    def step():
        if graphed:
            return g_solve(yb), g_dotl(zb)
        return eng.solve(yb), eng.dot_L(zb)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(3, args.warmup)):
        a_inv, lz = step()
    l0 = loc.launch_count()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        a_inv, lz = step()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    if world > 1:
        tms = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms.item())
    clocks = sampler.stop() if rank == 0 else None
    launches = (g_solve.launches + g_dotl.launches) * args.steps if graphed else (loc.launch_count() - l0)
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

    # parity on this rank's span against the reference on the first points (rank 0) ------------
    line = None
    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        alg = (hi - lo) * ((6 * J + 5) * 8 + (3 * J + 3) * 8)     # solve + dot_L, SURVEY.md §8d
        alg_sweep = (hi - lo) * (3 * J + 2) * 8                   # one forward sweep: phi, u, W rows + D + rhs
        line = {
            "metric": metric, "value": value, "unit": "point-terms/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": desc, "N": N, "J": J, "halo": int(halo) if world > 1 else 0, "cuda_graph": graphed,
                       "sharding": "time axis split into %d spans; one all-gather of (J^2+J) doubles per sweep" % world,
                       "inputs": "factor %.2f GB per rank, larger than the 126 MB L2" % ((hi - lo) * (3 * J + 1) * 8 / 1e9)},
            "roofline": {"bound": "hbm", "achieved": alg_sweep / (sweep_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg_sweep / (sweep_ms * 1e-3) / 1e9 / peak, "traffic": ncu_traffic("cfg4"),
                         "peak_source": peak_src, "kernel": "k_chunk_forward (one sweep; a step launches six)",
                         "algorithmic_bytes_per_launch": alg_sweep, "kernel_ms": sweep_ms,
                         "whole_step": {"algorithmic_bytes": alg, "achieved": alg / (ms * 1e-3) / 1e9,
                                        "frac": alg / (ms * 1e-3) / 1e9 / peak,
                                        "note": "one streaming pass of solve + dot_L as the serial algorithm "
                                                "would read it; the two-pass scan reads the factor twice"}},
            "gpu_launches": launches, "clocks": clocks,
        }
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
            lzi = s.dot_L(z)
        edt = (time.perf_counter() - t0) / reps
        line["e2e"] = {"value": 2.0 * N * J / edt, "unit": "point-terms/s", "ms_per_step": edt * 1e3,
                       "h2d_bytes_per_step": 2 * N * 8, "d2h_bytes_per_step": 2 * N * 8,
                       "matches_device_path": bool(np.allclose(zi[:, 0], a_inv.cpu().numpy(), rtol=1e-12, atol=0)),
                       "note": "cel_solve + cel_dot_L with pageable host arrays, wall clock"}
    elif rank == 0:
        line["e2e"] = {"value": None, "unit": "point-terms/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                       "note": "the drop-in host API is single-GPU; multi-GPU spans are device-resident"}
    if rank == 0 and not args.no_cpu_baseline:
        import oracle
        chk = oracle.best()
        n_sample = min(N, 1_000_000)
        cdt = cfg4_cpu(chk, coeffs, t, d, y, z, jc, n_sample)
        jj, a_r, c_r, a_c, b_c, c_c, d_c = synth.split_row(coeffs, 0, jc)
        p = oracle.Problem(jj, a_r, c_r, a_c, b_c, c_c, d_c, t[:n_sample], d[:n_sample])
        f = chk.factor(p)
        zr = chk.solve(f, y[:n_sample])[:, 0]
        m = min(n_sample, hi - lo) // 2      # away from the truncated end of the CPU sample
        got = a_inv[:m].cpu().numpy()
        line["cpu_baseline"] = {"value": 2.0 * n_sample * J / cdt, "unit": "point-terms/s", "cores": 1,
                                "kind": chk.kind,
                                "sample": "first %d of %d points, 1 thread (one series is serial on the CPU)" % (n_sample, N),
                                "max_rel_diff_vs_gpu": float(np.max(np.abs(got - zr[:m])) / np.max(np.abs(zr[:m])))}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS) + ["cfg4"])
    ap.add_argument("--batch", type=int, default=0, help="override series per GPU (debug)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--jc", type=int, default=0, help="cfg3: number of SHO terms (J = 2*jc)")
    ap.add_argument("--no-grad", action="store_true", help="skip the extra log-lik+grad measurement")
    ap.add_argument("--graph", action="store_true", help="cfg4, one GPU: time CUDA-graph replays of the two calls")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.workload == "cfg4":
        return run_cfg4(args)
    wl = dict(WORKLOADS[args.workload])
    if args.batch:
        wl["B"] = args.batch
        wl["desc"] += " [batch overridden to %d]" % args.batch
    if args.jc:
        wl["Jc"] = args.jc
        wl["desc"] += " [jc=%d]" % args.jc

    if args.impl == "reference":
        return run_reference(args, wl)

    import torch
    import torch.distributed as dist
    from celerite_b200 import synth
    from celerite_b200.batched import BatchedSolver
    from celerite_b200 import _capi

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (celerite_b200 has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    B, N, Jr, Jc = wl["B"], wl["N"], wl["Jr"], wl["Jc"]
    J, P = Jr + 2 * Jc, 1 + 2 * Jr + 4 * Jc

    # ---- synthetic inputs: generated straight into pinned host buffers, then copied to HBM ----
    ht = torch.empty((B, N), dtype=torch.float64).pin_memory()
    hd = torch.empty((B, N), dtype=torch.float64).pin_memory()
    hy = torch.empty((B, N), dtype=torch.float64).pin_memory()
    rng = np.random.Generator(np.random.PCG64(42 + rank))
    coeffs_np = synth.make_coeffs(rng, B, Jr, Jc)
    synth.fill_series(rng, ht.numpy(), hd.numpy(), hy.numpy())
    hc = torch.from_numpy(coeffs_np).pin_memory()
    h_out = torch.empty((3, B), dtype=torch.float64).pin_memory()
    h_status = torch.empty(B, dtype=torch.int32).pin_memory()
    dt_, dd_, dy_, dc_ = ht.to(dev), hd.to(dev), hy.to(dev), hc.to(dev)
    d_out = torch.empty((3, B), dtype=torch.float64, device=dev)
    d_status = torch.empty(B, dtype=torch.int32, device=dev)

    bs = BatchedSolver(local)
    stream = torch.cuda.current_stream(dev)

    def step_device():
        bs.raw_log_likelihood(B, N, N, Jr, Jc, dc_.data_ptr(), P, dt_.data_ptr(), dd_.data_ptr(), dy_.data_ptr(),
                              d_out[0].data_ptr(), d_out[1].data_ptr(), d_out[2].data_ptr(), d_status.data_ptr(),
                              flags=0, stream=stream.cuda_stream)

    def step_host():
        bs.raw_log_likelihood(B, N, N, Jr, Jc, hc.data_ptr(), P, ht.data_ptr(), hd.data_ptr(), hy.data_ptr(),
                              h_out[0].data_ptr(), h_out[1].data_ptr(), h_out[2].data_ptr(), h_status.data_ptr(),
                              flags=_capi.HOST_POINTERS, stream=stream.cuda_stream)

    d_gval = torch.empty(B, dtype=torch.float64, device=dev)
    d_grad = torch.empty((B, P), dtype=torch.float64, device=dev)

    def step_grad():
        import ctypes as C
        lib = bs._lib
        lib.cel_set_stream(bs._h, C.c_void_p(stream.cuda_stream or 1))
        v = lambda x: C.c_void_p(x)  # noqa: E731
        rc = lib.cel_batched_grad_log_likelihood(bs._h, B, N, N, Jr, Jc, v(dc_.data_ptr()), P, v(dt_.data_ptr()),
                                                 v(dd_.data_ptr()), v(dy_.data_ptr()), v(d_gval.data_ptr()),
                                                 v(d_grad.data_ptr()), v(d_status.data_ptr()), 0)
        _capi.check(rc, bs._h)

    primary_is_grad = bool(wl.get("grad"))
    if primary_is_grad:
        step_device = step_grad  # noqa: F811
    materialise = bool(wl.get("materialise"))
    if materialise:
        m_phi = torch.empty((B, N - 1, J), dtype=torch.float64, device=dev)
        m_u = torch.empty((B, N - 1, J), dtype=torch.float64, device=dev)
        m_W = torch.empty((B, N, J), dtype=torch.float64, device=dev)
        m_D = torch.empty((B, N), dtype=torch.float64, device=dev)

        def step_device():  # noqa: F811
            import ctypes as C
            lib = bs._lib
            lib.cel_set_stream(bs._h, C.c_void_p(stream.cuda_stream or 1))
            v = lambda x: C.c_void_p(x)  # noqa: E731
            rc = lib.cel_batched_compute(bs._h, B, N, N, Jr, Jc, v(dc_.data_ptr()), P, v(dt_.data_ptr()),
                                         v(dd_.data_ptr()), v(m_phi.data_ptr()), v(m_u.data_ptr()), v(m_W.data_ptr()),
                                         v(m_D.data_ptr()), v(d_out[0].data_ptr()), v(d_status.data_ptr()))
            _capi.check(rc, bs._h)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            tms = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ms = float(tms.item())
        return ms / steps

    # ---- device-resident metric -------------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    torch.cuda.synchronize(dev)
    assert int(d_status.sum().item()) == 0, "a synthetic series failed to factorise"
    launches0 = bs.launch_count()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_step = timed(step_device, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    launches = bs.launch_count() - launches0
    value = world * B * N * J / (ms_step * 1e-3)

    if materialise:
        # no host-buffer variant of the materialising call exists (the factor is 7.3 GB): report device numbers only
        if rank == 0:
            peak, peak_src = measured_peak_gbs()
            alg_bytes = B * N * 16 + B * (2 * (N - 1) * J + N * J + N) * 8 + B * 12
            achieved = alg_bytes / (ms_step * 1e-3) / 1e9
            print(json.dumps({
                "metric": "GP batched factorisation (materialised) point-terms/sec (N*J*batch / device time)",
                "value": value, "unit": "point-terms/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": wl["desc"], "series_per_gpu": B, "N": N, "J": J,
                           "inputs": "%.2f GB read + %.2f GB written per step per GPU, larger than the 126 MB L2"
                                     % (2 * B * N * 8 / 1e9, (alg_bytes - 2 * B * N * 8) / 1e9)},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": ncu_traffic(args.workload), "peak_source": peak_src,
                             "kernel": "k_compute_lockstep", "algorithmic_bytes_per_launch": alg_bytes},
                "e2e": {"value": None, "unit": "point-terms/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                        "note": "device-resident API only"},
                "gpu_launches": launches, "clocks": clocks}))
        if world > 1:
            dist.destroy_process_group()
        return 0
    # ---- end-to-end through the C ABI with host buffers -------------------------------------
    if primary_is_grad:
        h_gval = torch.empty(B, dtype=torch.float64).pin_memory()
        h_grad = torch.empty((B, P), dtype=torch.float64).pin_memory()

        def step_host():  # noqa: F811
            import ctypes as C
            lib = bs._lib
            lib.cel_set_stream(bs._h, C.c_void_p(stream.cuda_stream or 1))
            v = lambda x: C.c_void_p(x)  # noqa: E731
            rc = lib.cel_batched_grad_log_likelihood(bs._h, B, N, N, Jr, Jc, v(hc.data_ptr()), P, v(ht.data_ptr()),
                                                     v(hd.data_ptr()), v(hy.data_ptr()), v(h_gval.data_ptr()),
                                                     v(h_grad.data_ptr()), v(h_status.data_ptr()), _capi.HOST_POINTERS)
            _capi.check(rc, bs._h)
    step_host()  # warm-up (allocates the staging buffers)
    e2e_ms = timed(step_host, max(1, args.e2e_steps))
    e2e_value = world * B * N * J / (e2e_ms * 1e-3)
    # the e2e path must give the same answer as the device-resident path
    torch.cuda.synchronize(dev)
    if primary_is_grad:
        same = bool(torch.equal(h_grad.to(dev), d_grad))
    else:
        same = bool(torch.equal(h_out.to(dev), d_out))
    # ---- the same batch with the gradient (the "log-lik+grad" half of the metric's name) ----
    grad_extra = None
    if not primary_is_grad and not args.no_grad:
        step_grad()
        torch.cuda.synchronize(dev)
        g_ms = timed(step_grad, 3)
        grad_extra = {"value": world * B * N * J / (g_ms * 1e-3), "unit": "point-terms/s", "ms_per_step": g_ms,
                      "what": "cel_batched_grad_log_likelihood on the same batch (value + 1+2Jr+4Jc derivatives per series)"}

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        alg_bytes = B * N * 24 + B * 28          # t, diag, y in; logdet, quad, ll, status out
        achieved = alg_bytes / (ms_step * 1e-3) / 1e9
        line = {
            "metric": "GP log-likelihood point-terms/sec (N*J*batch / time)",
            "value": value, "unit": "point-terms/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": wl["desc"], "series_per_gpu": B, "N": N, "J": J,
                       "inputs": "%.2f GB per step per GPU, larger than the 126 MB L2 (no flush needed)"
                                 % (3 * B * N * 8 / 1e9),
                       "sharding": "batch index, no data-path collective"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic(args.workload), "peak_source": peak_src,
                         "kernel": "k_loglik_lockstep" if J <= 6 else "k_factor_coop",
                         "algorithmic_bytes_per_launch": alg_bytes},
            "e2e": {"value": e2e_value, "unit": "point-terms/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": 3 * B * N * 8 + B * P * 8, "d2h_bytes_per_step": B * 28,
                    "matches_device_path": same},
            "gpu_launches": launches,
            "clocks": clocks,
        }
        if grad_extra:
            line["with_grad"] = grad_extra
        if primary_is_grad:
            line["metric"] = "GP log-likelihood + gradient point-terms/sec (N*J*batch / device time)"
            line["roofline"]["kernel"] = "k_grad_coop"
            line["roofline"]["algorithmic_bytes_per_launch"] = B * N * 24 + B * (P + 1) * 8
            line["roofline"]["achieved"] = line["roofline"]["algorithmic_bytes_per_launch"] / (ms_step * 1e-3) / 1e9
            line["roofline"]["frac"] = line["roofline"]["achieved"] / peak
            line["roofline"]["note"] = "fp64-pipe/latency bound (1,024 independent recurrences); HBM fraction is not the limiter"
            line["e2e"]["d2h_bytes_per_step"] = B * (P + 1) * 8 + B * 4
        if not args.no_cpu_baseline:
            import oracle
            chk = oracle.best()
            cores = os.cpu_count() or 1
            sample = int(min(B, max(256, cores * 1024)))
            if primary_is_grad:
                sample = int(min(B, cores))
                Nc = min(N, 20000)   # the AD reference takes ~1.7 s per 1e5 points at J=8: bound the sample in N too
                tN, dN, yN = (np.ascontiguousarray(a.numpy()[:sample, :Nc]) for a in (ht, hd, hy))
                t0 = time.perf_counter()
                gv, gg, gs = chk.batched_grad(sample, Nc, Nc, Jr, Jc, coeffs_np[:sample], P, tN, dN, yN, nthreads=cores)
                cdt = time.perf_counter() - t0
                line["cpu_baseline"] = {"value": sample * Nc * J / cdt, "unit": "point-terms/s", "cores": cores,
                                        "kind": chk.kind,
                                        "sample": "first %d series truncated to N=%d, forward-mode AD gradient "
                                                  "(solver.cpp:347-463), %d threads" % (sample, Nc, cores)}
                print(json.dumps(line))
                if world > 1:
                    dist.destroy_process_group()
                return 0
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
            par = float(max(np.max(np.abs(got[0] - ref[:, 0]) / np.abs(ref[:, 0])),
                            np.max(np.abs(got[1] - ref[:, 1]) / np.abs(ref[:, 1]))))
            line["cpu_baseline"] = {"value": sample * N * J / cdt, "unit": "point-terms/s", "cores": cores,
                                    "kind": chk.kind,
                                    "sample": "first %d of %d series, %d threads, one series per task" % (sample, B, cores),
                                    "max_rel_diff_vs_gpu": par}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
