#!/usr/bin/env python
"""bench.py -- ILU factor+solve block-rows/s (fp64, blockw=32) on B200, with roofline and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg3]

A "step" is one fused factor + solve of the whole batch (BASELINE.json configs[2]: blockw=32,
nblocks=8192, ndiag=2, 64 independent systems, 1 RHS -- the config the metric is quoted on) with the
inputs already resident in HBM.  The same JSON line carries sub-records for the other configurations SURVEY 8(d)
names (cfg1 and cfg2 also replayed from a CUDA graph, cfg4 with its solve-only figure), the matvec, the
create-once / solve-many reuse pattern through the host API, and `cfg5_sharded`: the 4096-system ensemble of
BASELINE configs[4] SHARDED over the GPUs of the run (strong scaling; `--workload cfg5` makes it the headline).  The factor works in place, so every step starts from a pristine copy
of the batch restored by a device-to-device copy that is NOT timed: the timed quantity is the sum of
the K per-step CUDA-event intervals on the launching stream (max over ranks).  The batch (4.8 GB) is
38x larger than L2, so no explicit L2 flush is needed.

N > 1 (torchrun): every rank owns its own slab of 64 systems (weak scaling), no data-path
collective; NCCL only all-gathers the per-system residual norms after the timed region.

--impl reference times the reference's own CPU implementation (oracle/_ref: the untouched header +
LAPACK dgetrf, OpenMP over systems with all host threads) on a bounded sample of the same workload.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")  # the header's own advice (block_diag_ilu.hpp:450)

WORKLOADS = {
    # name: (blockw, nblocks, ndiag, nsys, nrhs)
    "cfg1": (4, 64, 1, 1, 1),
    "cfg2": (16, 1024, 1, 1, 1),
    "cfg3": (32, 8192, 2, 64, 1),
    "cfg4": (64, 2048, 3, 1, 256),
    "cfg5": (8, 512, 2, 4096, 1),
}
METRIC = "ILU factor+solve block-rows/s (fp64, blockw=32)"
UNIT = "block-rows/s"


def algorithmic_bytes(w, N, d, nrhs):
    """SURVEY.md section 8(d), ld = w, per system."""
    W_blk = N * w * w
    W_diag = (d * N - d * (d + 1) // 2) * w
    Nw = N * w
    return {
        "factor": 8 * (2 * W_blk + 2 * W_diag) + 4 * Nw,
        "solve": 8 * (W_blk + 2 * W_diag) + 8 * Nw + 16 * Nw * nrhs,
        "fused": 8 * (2 * W_blk + W_blk * (w + 1) / (2 * w) + 3 * W_diag) + 4 * Nw + 16 * Nw * nrhs,
    }


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def host_threads():
    """Every host core this process may run on.  Not omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def ncu_traffic(kernel, nsys):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed `ncu --set full`
    capture of this workload (profiles/traffic.json, written by tools/ncu_traffic.py); None when there is none or
    when it was taken on another batch size."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        e = next(v for k, v in t["kernels"].items() if k.startswith(kernel))
        return float(e["dram_bytes"]) if int(e["nsys"]) == int(nsys) else None
    except Exception:
        return None


def bind_host_to_gpu_numa(index):
    """Run this rank's host thread (and so its pinned allocations) on the NUMA node the GPU hangs off.

    The e2e leg of every rank streams ~5 GB per step from pinned host memory; with 8 ranks on one host the copies of
    a rank whose buffers sit on the far socket cross the inter-socket link.  Returns the node, or None when the
    topology cannot be read (then nothing is changed)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dom, rest = bus.split(":", 1)
        dev = "%s:%s" % (dom[-4:], rest)
        with open("/sys/bus/pci/devices/%s/numa_node" % dev.lower()) as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        cpus = set()
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            for part in f.read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed regions run."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self.stop_flag:
            try:
                util = nv.nvmlDeviceGetUtilizationRates(self.h).gpu
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                if util > 0:
                    self.samples.append(mhz)
                    for bit, name in names.items():
                        if r & bit:
                            self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


SEED0 = 3_000_000  # system s of workload cfgK uses numpy seed SEED0 + 100_000 * K + s in BOTH arms


def workload_string(name):
    w, N, d, nsys, nrhs = WORKLOADS[name]
    return "%s: blockw=%d nblocks=%d ndiag=%d nsys=%d nrhs=%d" % (name, w, N, d, nsys, nrhs)


def host_batch(name, lo, hi, nrhs=None):
    """Systems [lo, hi) of workload `name`, generated on the host with the seeds both arms share."""
    from block_diag_ilu_b200 import synth
    w, N, d, nsys, nr = WORKLOADS[name]
    return synth.make_batch(hi - lo, N, w, d, "D", seed0=SEED0 + 100_000 * int(name[3:]) + lo, nrhs=nrhs or nr)


def stats_ms(ts):
    s = sorted(ts)
    return {"mean": sum(s) / len(s), "best": s[0], "median": s[len(s) // 2], "n": len(s)}


def run_reference(args, rank, world):
    """CPU arm: the reference's own implementation on the box's host cores, same workload, same seeds."""
    if rank != 0:
        return
    import numpy as np
    from oracle import loader
    w, N, d, nsys, nrhs = WORKLOADS[args.workload]
    orc = loader.best()
    threads = host_threads()
    # bounded sample: one system per host thread (at least 8), at most the full batch; cfg3 costs about
    # 0.11 s of one core per system, so K+W = 13 steps stay well under a minute of wall time.
    ns = min(nsys, max(8, threads))
    nr = min(nrhs, 8)
    data, b = host_batch(args.workload, 0, ns, nrhs=nr)
    times = []
    for it in range(args.warmup + args.steps):
        work = data.copy()
        r = orc.factor_solve_batch(work, ns, N, w, d, 0, w, b, nrhs=nr, nthreads=threads, mode=1)
        assert int(np.abs(r["infos"]).max()) == 0
        if it >= args.warmup:
            times.append(r["seconds"])
    t = sum(times) / len(times)
    value = ns * N / t
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True,
        "scaling": "strong" if args.workload == "cfg5" else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_string(args.workload), "family": "D (diagonally dominant, SURVEY 8d)",
                   "seed0": SEED0},
        "sample": {"systems": ns, "rhs": nr, "of_systems": nsys, "of_rhs": nrhs,
                   "note": "block-rows/s of the sample; the first systems of the same seeded batch the GPU arm runs"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": orc.kind,
                         "sample": "%d of %d systems, %d rhs each, OpenMP over systems (%d threads), "
                                   "single-threaded header inside" % (ns, nsys, nr, threads)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def cpu_baseline_leg(w, N, d, nsys, nrhs, host_data, host_b):
    """Rank 0, N=1 only: the reference (or the port) timed on the host cores on a bounded sample."""
    import numpy as np
    from oracle import loader
    orc = loader.best()
    threads = host_threads()
    ns = min(nsys, max(8, threads))
    nr = min(nrhs, 8)
    work = np.ascontiguousarray(host_data[:ns]).copy()
    b = np.ascontiguousarray(host_b[:ns, :nr])
    best = None
    for _ in range(2):
        scratch = work.copy()
        r = orc.factor_solve_batch(scratch, ns, N, w, d, 0, w, b, nrhs=nr, nthreads=threads, mode=1)
        best = r["seconds"] if best is None else min(best, r["seconds"])
    out = {"value": ns * N / best, "unit": UNIT, "cores": threads, "kind": orc.kind,
           "sample": "%d of %d systems x %d rhs, OpenMP over systems (%d threads), single-threaded header inside, "
                     "best of 2" % (ns, nsys, nr, threads), "x": r["x"], "ns": ns, "nr": nr}
    if orc.kind == "reference":
        scratch = work[:min(ns, 4)].copy()
        r0 = orc.factor_solve_batch(scratch, scratch.shape[0], N, w, d, 0, w, b[:scratch.shape[0]], nrhs=nr,
                                    nthreads=threads, mode=0)
        out["as_shipped"] = {"value": scratch.shape[0] * N / r0["seconds"], "unit": UNIT,
                             "note": "systems one after another, header's OpenMP loop over blocks (%d threads), "
                                     "serial solve; %d systems" % (threads, scratch.shape[0])}
    return out


class DeviceCase:
    """One workload slab resident in HBM: pristine batch, work copy, outputs, and the three timed calls."""

    def __init__(self, name, lo, hi, device, host=None):
        import torch
        from block_diag_ilu_b200 import _lib, device as dev
        self.torch, self._lib, self.dev = torch, _lib, dev
        self.L = _lib.lib()
        self.name = name
        self.w, self.N, self.d, self.nsys_total, self.nrhs = WORKLOADS[name]
        self.nsys = hi - lo
        self.n = self.N * self.w
        data, b = host if host is not None else host_batch(name, lo, hi)
        self.host_data, self.host_b = data, b
        self.pristine = torch.from_numpy(data).to(device)
        self.b = torch.from_numpy(b).to(device)
        self.work = torch.empty_like(self.pristine)
        self.shape = dev.Shape(self.N, self.w, self.d)
        self.ipiv = torch.empty((self.nsys, self.n), dtype=torch.int32, device=device)
        self.rbc, self.cbr = torch.empty_like(self.ipiv), torch.empty_like(self.ipiv)
        self.info = torch.empty((self.nsys,), dtype=torch.int32, device=device)
        self.flags = torch.empty((self.nsys, self.nrhs), dtype=torch.int32, device=device)
        self.x = torch.empty_like(self.b)
        self.AB = algorithmic_bytes(self.w, self.N, self.d, self.nrhs)

    def fused(self):
        P, st = self._lib.ptr, self.torch.cuda.current_stream()
        self._lib.check(self.L.bdilu_factor_solve(P(self.work), self.work.stride(0), self.nsys, *self.shape.dims(),
                                                  P(self.ipiv), P(self.rbc), P(self.cbr), P(self.info), P(self.b), P(self.x),
                                                  self.nrhs, self.n, self.n * self.nrhs, P(self.flags), st.cuda_stream))

    def factor_only(self):
        P, st = self._lib.ptr, self.torch.cuda.current_stream()
        self._lib.check(self.L.bdilu_factor(P(self.work), self.work.stride(0), self.nsys, *self.shape.dims(), P(self.ipiv),
                                            P(self.rbc), P(self.cbr), P(self.info), st.cuda_stream))

    def solve_only(self):
        P, st = self._lib.ptr, self.torch.cuda.current_stream()
        self._lib.check(self.L.bdilu_solve(P(self.work), self.work.stride(0), self.nsys, *self.shape.dims(), P(self.rbc),
                                           P(self.cbr), P(self.b), P(self.x), self.nrhs, self.n, self.n * self.nrhs, None,
                                           P(self.flags), st.cuda_stream))

    def timed(self, fn, reps, restore=True, flush=None):
        torch = self.torch
        st = torch.cuda.current_stream()
        evs = []
        for _ in range(reps):
            if restore:
                self.work.copy_(self.pristine)
            if flush is not None:
                flush.add_(1.0)  # inputs smaller than L2: evict them (a buffer larger than L2 is rewritten)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            fn()
            e1.record(st)
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return [a.elapsed_time(c) for a, c in evs]

    def check_clean(self):
        assert int(self.info.abs().max().item()) == 0 and int(self.flags.abs().max().item()) == 0

    def rel_residual(self):
        norms = self.dev.residual_norm2(self.pristine, self.b, self.x, self.N, self.w, self.d)
        return (norms.sqrt() / self.b.norm(dim=2)).reshape(-1)


def kernel_names(w, workload, nsys, nrhs):
    """kernel-name prefixes as `ncu` prints them (tools/ncu_traffic.py keys profiles/traffic.json the same way)"""
    wp = 2
    while wp < w:
        wp *= 2
    fk = "factor_kernel<%d," % wp if w <= 32 else "factor_wide_kernel"
    if w > 32:
        sk = "solve_mrhs_wave_kernel<64,"
    elif w == 32 and workload == "cfg3":
        sk = "solve_quad32_kernel<"
    else:
        sk = "pf::solve_chain_pf_kernel<%d," % wp     # prefetched narrow-block chain kernel (csrc/bdilu_solve_pf.cu)
    return fk, sk


def sub_record(case, peak, steps, flush, solve_only=False):
    """A compact record of one extra workload: fused factor+solve (and solve-only) with its own roofline."""
    t = case.timed(case.fused, 2, flush=flush)  # warm-up
    L = case.L
    L.bdilu_launch_count_reset()
    t = stats_ms(case.timed(case.fused, steps, flush=flush))
    launches = int(L.bdilu_launch_count()) // steps
    case.check_clean()
    rec = {"workload": workload_string(case.name), "nsys_here": case.nsys, "ms": t,
           "value": case.nsys * case.N / (t["mean"] * 1e-3), "unit": UNIT, "launches_per_step": launches,
           "roofline": {"bound": "hbm", "algorithmic_bytes_per_step": case.nsys * case.AB["fused"],
                        "achieved": case.nsys * case.AB["fused"] / (t["mean"] * 1e-3) / 1e9, "peak": peak, "unit": "GB/s"},
           "max_rel_residual": float(case.rel_residual().max().item())}
    rec["roofline"]["frac"] = rec["roofline"]["achieved"] / peak
    if solve_only:
        case.work.copy_(case.pristine)
        case.factor_only()
        ts = stats_ms(case.timed(case.solve_only, steps, restore=False, flush=flush))
        gbs = case.nsys * case.AB["solve"] / (ts["mean"] * 1e-3) / 1e9
        rec["solve_only"] = {"ms": ts, "value": case.nrhs * case.nsys * case.N / (ts["mean"] * 1e-3),
                             "unit": "vector block-rows/s (nrhs * nsys * nblocks / t_solve, SURVEY 8d)",
                             "roofline": {"achieved": gbs, "frac": gbs / peak,
                                          "algorithmic_bytes": case.nsys * case.AB["solve"]}}
    return rec


def graph_replay_record(case, replays):
    """cfg1-sized work is launch bound: the fused call captured once in a CUDA graph and replayed back to back."""
    torch = case.torch
    g = case.dev.GraphedFactorSolve(case.pristine, case.b, case.N, case.w, case.d)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(replays):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / replays
    ok = bool(torch.equal(g.x, case.x)) if case.x is not None else None
    return {"replays": replays, "us_per_replay": us, "value": case.nsys * case.N / (us * 1e-6), "unit": UNIT,
            "note": "restore copy + factor + solve captured in one CUDA graph (device.GraphedFactorSolve)",
            "bit_identical_to_eager": ok}


def matvec_record(case, peak, steps):
    """dot_vec on the unfactored cfg3 batch: every matrix element read once (SURVEY 8d B_dot_vec)."""
    torch = case.torch
    out = torch.empty_like(case.b)
    st = torch.cuda.current_stream()
    ts = []
    for it in range(steps + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        case.dev.dot_vec(case.pristine, case.b, case.N, case.w, case.d, out=out)
        e1.record(st)
        torch.cuda.synchronize()
        if it >= 2:
            ts.append(e0.elapsed_time(e1))
    t = stats_ms(ts)
    W_blk, W_diag, Nw = case.N * case.w ** 2, (case.d * case.N - case.d * (case.d + 1) // 2) * case.w, case.n
    bytes_ = case.nsys * (8 * (W_blk + 2 * W_diag) + 16 * Nw)
    gbs = bytes_ / (t["mean"] * 1e-3) / 1e9
    return {"workload": workload_string(case.name) + " dot_vec", "ms": t, "value": case.nsys * case.N / (t["mean"] * 1e-3),
            "unit": UNIT, "roofline": {"bound": "hbm", "algorithmic_bytes": bytes_, "achieved": gbs, "peak": peak,
                                       "frac": gbs / peak, "unit": "GB/s"}}


def factor_pivoting_record(case, steps):
    """K1 on a batch that DOES pivot (uniform random blocks: ~87 % non-identity pivots at blockw 32).  The timed family D
    never pivots and takes the factor kernel's optimistic path; this is what the exact fallback costs (disclosure, not
    part of the metric).  The batch is generated over the pristine buffer and family D is restored afterwards."""
    torch = case.torch
    keep = case.pristine.clone()
    nblk = case.N * case.w * case.w
    g = torch.Generator(device=case.pristine.device)
    g.manual_seed(777)
    case.pristine[:, :nblk].uniform_(-1.0, 1.0, generator=g)
    st = torch.cuda.current_stream()
    ts = []
    nonid = 0.0
    for it in range(steps + 2):
        work = case.pristine.clone()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        ilu = case.dev.BatchedILU(work, case.N, case.w, case.d, check=False)
        e1.record(st)
        torch.cuda.synchronize()
        if it >= 2:
            ts.append(e0.elapsed_time(e1))
        ident = torch.arange(1, case.w + 1, device=work.device, dtype=torch.int32).repeat(case.N)
        nonid = float((ilu.ipiv != ident).float().mean())
        del work, ilu
    case.pristine.copy_(keep)
    del keep
    return {"workload": workload_string(case.name) + " factor only, uniform random blocks (every block pivots)",
            "ms": stats_ms(ts), "non_identity_pivots": nonid,
            "note": "optimistic diagonal-pivot attempt abandoned at its first vote, then the general elimination; includes the "
                    "allocation of the index arrays by the wrapper"}


def e2e_reuse_record(case, nsolves):
    """What an unmodified caller of the reference does with one Jacobian (anyode_iterative.hpp:64-88): one ILU, many
    solves -- bdilu_ilu_create once on host data, then bdilu_ilu_solve from host vectors, all copies inside."""
    import ctypes as C
    import numpy as np
    torch, _lib, L = case.torch, case._lib, case.L
    P = _lib.ptr
    h_data = torch.empty(case.pristine.shape, dtype=torch.float64, pin_memory=True)
    h_b = torch.empty(case.b.shape, dtype=torch.float64, pin_memory=True)
    h_x = torch.empty(case.b.shape, dtype=torch.float64, pin_memory=True)
    h_data.copy_(case.pristine)
    h_b.copy_(case.b)
    torch.cuda.synchronize()
    h_info = np.zeros(case.nsys, np.int32)
    h_flags = np.zeros(case.nsys * case.nrhs, np.int32)
    handle = C.c_void_p()
    t0 = time.perf_counter()
    _lib.check(L.bdilu_ilu_create(C.byref(handle), P(h_data), h_data.stride(0), case.nsys, *case.shape.dims(), 0,
                                  None, None, None, P(h_info)))
    t_create = time.perf_counter() - t0
    assert not h_info.any()
    _lib.check(L.bdilu_ilu_solve(handle, P(h_b), P(h_x), case.nrhs, P(h_flags)))
    t0 = time.perf_counter()
    for _ in range(nsolves):
        _lib.check(L.bdilu_ilu_solve(handle, P(h_b), P(h_x), case.nrhs, P(h_flags)))
    t_solve = (time.perf_counter() - t0) / nsolves
    L.bdilu_ilu_destroy(handle)
    assert not h_flags.any()
    diff = float((h_x.to(case.x.device) - case.x).abs().max().item())
    total = t_create + nsolves * t_solve
    return {"api": "bdilu_ilu_create(copy_back=0) once + %d x bdilu_ilu_solve, host pointers (pinned), copies inside" % nsolves,
            "create_ms": 1e3 * t_create, "solve_ms": 1e3 * t_solve, "nsolves": nsolves,
            "value": case.nsys * case.N * nsolves / total, "unit": "block-rows/s per solve, factorisation amortised",
            "h2d_bytes_create": int(h_data.numel() * 8), "h2d_bytes_per_solve": int(h_b.numel() * 8),
            "d2h_bytes_per_solve": int(h_x.numel() * 8 + 4 * h_flags.size),
            "max_abs_diff_vs_device_path": diff}


def cfg5_sharded_record(world, rank, device, peak, steps, dist):
    """BASELINE configs[4]: the 4096-system ensemble SHARDED over the ranks (strong scaling): rank r factors and solves
    its contiguous slab, no data-path collective; time = max over ranks."""
    import torch
    from block_diag_ilu_b200 import sharding
    w, N, d, nsys, nrhs = WORKLOADS["cfg5"]
    lo, hi = sharding.slab_bounds(nsys, world, rank)
    case = DeviceCase("cfg5", lo, hi, device)
    flush = torch.zeros(40 * 1024 * 1024, dtype=torch.float64, device=device) if case.pristine.numel() * 8 < 2.5e8 else None
    case.timed(case.fused, 3, flush=flush)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ts = case.timed(case.fused, steps, flush=flush)
    if world > 1:
        dist.barrier()
    case.check_clean()
    t = stats_ms(ts)
    mean = t["mean"]
    if world > 1:
        tt = torch.tensor([t["mean"], t["best"], t["median"]], dtype=torch.float64, device=device)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        mean, t = float(tt[0].item()), {"mean": float(tt[0].item()), "best": float(tt[1].item()),
                                        "median": float(tt[2].item()), "n": steps}
    rel = case.rel_residual()
    worst = float(rel.max().item())
    if world > 1:
        worst = sharding.max_over_ranks(worst, device=device)
    gbs = nsys * case.AB["fused"] / (mean * 1e-3) / 1e9
    return {"workload": workload_string("cfg5"), "scaling": "strong", "n_gpus": world,
            "systems_per_gpu": sharding.slab_sizes(nsys, world), "ms": t, "value": nsys * N / (mean * 1e-3), "unit": UNIT,
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_step": nsys * case.AB["fused"], "achieved": gbs,
                         "peak": peak * world, "frac": gbs / (peak * world), "unit": "GB/s",
                         "note": "whole-job bytes / max-over-ranks time against n_gpus x the measured peak"},
            "l2": ("slab smaller than L2: a 320 MB buffer is rewritten between timed steps" if flush is not None
                   else "slab larger than L2"),
            "max_rel_residual": worst}


def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    from block_diag_ilu_b200 import _lib, sharding

    w, N, d, nsys_total, nrhs = WORKLOADS[args.workload]
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    L = _lib.lib()
    _lib.require_device()
    peak, peak_src = measured_peak_gbs()
    P = _lib.ptr

    # cfg5 is the ensemble BASELINE shards over the GPUs (strong scaling); every other workload gives each rank its own
    # batch of the named size (weak scaling: independent replicas with different seeds, no data-path collective)
    strong = args.workload == "cfg5"
    if strong:
        lo, hi = sharding.slab_bounds(nsys_total, world, rank)
    else:
        lo, hi = rank * nsys_total, (rank + 1) * nsys_total
    case = DeviceCase(args.workload, lo, hi, device)
    nsys = case.nsys
    big = case.pristine.numel() * 8 > 2.5e8
    flush = None if big else torch.zeros(40 * 1024 * 1024, dtype=torch.float64, device=device)

    sampler = ClockSampler(local_rank)
    sampler.start()
    case.timed(case.fused, args.warmup, flush=flush)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    barrier()
    L.bdilu_launch_count_reset()
    t_steps = case.timed(case.fused, args.steps, flush=flush)
    barrier()
    launches = int(L.bdilu_launch_count())
    st = stats_ms(t_steps)
    if world > 1:
        tt = torch.tensor([st["mean"], st["best"], st["median"]], dtype=torch.float64, device=device)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        st = {"mean": float(tt[0].item()), "best": float(tt[1].item()), "median": float(tt[2].item()), "n": args.steps}
    ms_per_step = st["mean"]
    total_systems = nsys_total if strong else world * nsys_total
    value = total_systems * N / (ms_per_step * 1e-3)

    # sanity on what the timed steps produced + the residual norms (NCCL gather only here)
    case.check_clean()
    rel = case.rel_residual()  # ||b - A x|| / ||b|| per (system, vector): ILU is approximate
    if world > 1:
        rel = sharding.gather_norms(rel, nsys=total_systems * nrhs if strong else None)  # the path's only collective
    rel_res = float(rel.max().item())

    # per-kernel durations, measured live on the launching stream
    reps = max(3, min(args.steps, 10))
    t_factor = stats_ms(case.timed(case.factor_only, reps, flush=flush))
    case.work.copy_(case.pristine)
    case.factor_only()
    t_solve = stats_ms(case.timed(case.solve_only, reps, restore=False, flush=flush))
    kf, ks = t_factor["mean"], t_solve["mean"]
    AB = case.AB
    fk, sk = kernel_names(w, args.workload, nsys, nrhs)
    dom = "factor" if kf >= ks else "solve"
    fused_gbs = total_systems * AB["fused"] / (ms_per_step * 1e-3) / 1e9
    # The metric is quoted on the fused step: that is the top-level figure.  The kernels that make it up follow.
    roofline = {"bound": "hbm", "what": "fused factor+solve step (B_fused of SURVEY 8d over the step time)",
                "achieved": fused_gbs, "peak": peak * world, "unit": "GB/s", "frac": fused_gbs / (peak * world),
                "traffic": None, "peak_source": peak_src + (" x %d GPUs" % world if world > 1 else ""),
                "algorithmic_bytes_per_step": total_systems * AB["fused"],
                "frac_of_nominal_8TBs": fused_gbs / (8000.0 * world),
                "dominant_kernel": dom,
                "kernels": {
                    "factor": {"name": fk, "ms": t_factor, "algorithmic_bytes_per_launch": nsys * AB["factor"],
                               "achieved": nsys * AB["factor"] / (kf * 1e-3) / 1e9,
                               "frac": nsys * AB["factor"] / (kf * 1e-3) / 1e9 / peak, "traffic": ncu_traffic(fk, nsys)},
                    "solve": {"name": sk, "ms": t_solve, "algorithmic_bytes_per_launch": nsys * AB["solve"],
                              "achieved": nsys * AB["solve"] / (ks * 1e-3) / 1e9,
                              "frac": nsys * AB["solve"] / (ks * 1e-3) / 1e9 / peak, "traffic": ncu_traffic(sk, nsys)}}}
    tf_, ts_ = roofline["kernels"]["factor"]["traffic"], roofline["kernels"]["solve"]["traffic"]
    if tf_ is not None and ts_ is not None:
        roofline["traffic"] = tf_ + ts_  # DRAM bytes of the step = its two launches (ncu --set full, profiles/traffic.json)

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "ms_best": st["best"], "ms_median": st["median"],
            "higher_is_better": True, "scaling": "strong" if strong else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_string(args.workload), "family": "D (diagonally dominant, SURVEY 8d)",
                       "seed0": SEED0, "call": "fused factor+solve (bdilu_factor_solve), device-resident",
                       "l2": ("inputs (%.2f GB/step) larger than L2" % (case.pristine.numel() * 8 / 1e9) if big else
                              "inputs smaller than L2: a 320 MB buffer is rewritten between timed steps"),
                       "timing": "per-step CUDA-event intervals on the launching stream (mean = value; best and median "
                                 "beside it); the device-to-device restore of the in-place-factored batch between "
                                 "steps is untimed",
                       "parallelism": ("%d systems sharded over %d GPU(s) in contiguous slabs %s, no data-path collective"
                                       % (nsys_total, world, sharding.slab_sizes(nsys_total, world)) if strong else
                                       "%d systems per GPU (independent replicas), no data-path collective" % nsys_total)},
            "roofline": roofline, "gpu_launches": launches, "max_rel_residual": rel_res}

    # e2e through the host-pointer C ABI on every rank's slab: pinned host buffers, H2D + fused + D2H inside the timed
    # region; barrier on both sides, max over ranks (each GPU has its own PCIe link, the host memory is shared)
    numa_node = bind_host_to_gpu_numa(local_rank) if world > 1 else None
    h_data = torch.empty(case.pristine.shape, dtype=torch.float64, pin_memory=True)
    h_b = torch.empty(case.b.shape, dtype=torch.float64, pin_memory=True)
    h_x = torch.empty(case.b.shape, dtype=torch.float64, pin_memory=True)
    h_data.copy_(case.pristine)
    h_b.copy_(case.b)
    torch.cuda.synchronize()
    h_info = np.zeros(nsys, np.int32)
    h_flags = np.zeros(nsys * nrhs, np.int32)

    def e2e_call():
        _lib.check(L.bdilu_host_factor_solve(P(h_data), h_data.stride(0), nsys, *case.shape.dims(), P(h_b), P(h_x),
                                             nrhs, P(h_info), P(h_flags)))
    e2e_steps = max(2, min(args.steps, 5))
    e2e_call()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_call()
    t_e2e = (time.perf_counter() - t0) / e2e_steps
    if world > 1:
        t_e2e = sharding.max_over_ranks(t_e2e, device=device)
    assert not h_info.any() and not h_flags.any()
    e2e_diff = float((h_x.to(device) - case.x).abs().max().item())
    assert e2e_diff <= 1e-12 * float(case.x.abs().max().item()), "host-pointer path disagrees with the device path"
    per_rank_in = h_data.numel() * 8 + h_b.numel() * 8
    per_rank_out = h_x.numel() * 8 + 4 * nsys + 4 * nsys * nrhs
    if world > 1 and strong:
        tt = torch.tensor([per_rank_in, per_rank_out], dtype=torch.float64, device=device)
        dist.all_reduce(tt)
        tot_in, tot_out = int(tt[0].item()), int(tt[1].item())
    else:
        tot_in, tot_out = world * per_rank_in, world * per_rank_out
    line["e2e"] = {"value": total_systems * N / t_e2e, "unit": UNIT, "h2d_bytes_per_step": tot_in,
                   "d2h_bytes_per_step": tot_out, "ms_per_step": 1e3 * t_e2e,
                   "steps": e2e_steps, "n_gpus": world, "bit_identical_to_device_path": e2e_diff == 0.0,
                   "host_numa_node_rank0": numa_node,
                   "api": "bdilu_host_factor_solve (host-pointer C ABI; pinned host buffers; slabs of systems pipelined "
                          "over 3 streams; one call per rank on its own slab of the batch, max over ranks)"}
    del h_data, h_x
    sampler.stop_flag = True
    sampler.join(timeout=1.0)
    line["clocks"] = sampler.summary()

    # the ensemble of BASELINE configs[4], sharded over however many GPUs this run has (every rank takes part)
    extra = not args.no_extra
    if extra and not strong:
        line["cfg5_sharded"] = cfg5_sharded_record(world, rank, device, peak, max(5, min(args.steps, 10)), dist)
    # the other configurations SURVEY 8(d) names, the matvec and the reuse pattern: one GPU, rank 0, a few seconds
    if extra and world == 1 and args.workload == "cfg3":
        sub = {}
        steps = max(5, min(args.steps, 10))
        fl = flush if flush is not None else torch.zeros(40 * 1024 * 1024, dtype=torch.float64, device=device)
        line["matvec"] = matvec_record(case, peak, steps)
        line["e2e_reuse"] = e2e_reuse_record(case, 32)
        line["factor_pivoting"] = factor_pivoting_record(case, steps)
        for name in ("cfg1", "cfg2", "cfg4"):
            c = DeviceCase(name, 0, WORKLOADS[name][3], device)
            sub[name] = sub_record(c, peak, steps, fl, solve_only=(name == "cfg4"))
            if name in ("cfg1", "cfg2"):
                sub[name]["cuda_graph"] = graph_replay_record(c, 10000 if name == "cfg1" else 1000)
            del c
        line["configs"] = sub

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline_leg(w, N, d, nsys, nrhs, case.host_data, case.host_b)
        xs = cb.pop("x")
        ns, nr = cb.pop("ns"), cb.pop("nr")
        gx = case.x[:ns, :nr].cpu().numpy()
        cb["max_rel_diff_vs_gpu"] = float((np.abs(gx - xs).max(axis=2) / np.abs(xs).max(axis=2)).max())
        line["cpu_baseline"] = cb
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_JSON_OUT = None


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the sub-records (cfg1/2/4, cfg5 sharded, matvec, reuse)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly one JSON line: whatever libraries print there (NCCL's version banner under torchrun)
    # is sent to stderr for the duration of the run
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
