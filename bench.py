#!/usr/bin/env python
"""bench.py -- ILU factor+solve block-rows/s (fp64, blockw=32) on B200, with roofline and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg3]

A "step" is one fused factor + solve of the whole batch (BASELINE.json configs[2]: blockw=32,
nblocks=8192, ndiag=2, 64 independent systems, 1 RHS -- the config the metric is quoted on) with the
inputs already resident in HBM.  The factor works in place, so every step starts from a pristine copy
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


def run_reference(args, rank, world):
    """CPU arm: the reference's own implementation on the box's host cores."""
    if rank != 0:
        return
    import numpy as np
    from oracle import loader
    from block_diag_ilu_b200 import synth
    w, N, d, nsys, nrhs = WORKLOADS[args.workload]
    orc = loader.best()
    threads = host_threads()
    # bounded sample: one system per host thread (at least 8), at most the full batch; cfg3 costs about
    # 0.11 s of one core per system, so K+W = 13 steps stay well under a minute of wall time.
    ns = min(nsys, max(8, threads))
    nr = min(nrhs, 8)
    data, b = synth.make_batch(ns, N, w, d, "D", seed0=3_000_000, nrhs=nr)
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
        "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "%s: blockw=%d nblocks=%d ndiag=%d nsys=%d nrhs=%d" % (args.workload, w, N, d, nsys, nrhs),
                   "family": "D (diagonally dominant)", "sample_systems": ns, "sample_rhs": nr},
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


def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    from block_diag_ilu_b200 import _lib, device as dev, synth

    w, N, d, nsys, nrhs = WORKLOADS[args.workload]
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    L = _lib.lib()
    _lib.require_device()
    n = N * w
    AB = algorithmic_bytes(w, N, d, nrhs)
    peak, peak_src = measured_peak_gbs()

    pristine, b = synth.make_batch_device(nsys, N, w, d, device, seed=3 + rank, nrhs=nrhs)
    work = torch.empty_like(pristine)
    shape = dev.Shape(N, w, d)
    ipiv = torch.empty((nsys, n), dtype=torch.int32, device=device)
    rbc, cbr = torch.empty_like(ipiv), torch.empty_like(ipiv)
    info = torch.empty((nsys,), dtype=torch.int32, device=device)
    flags = torch.empty((nsys, nrhs), dtype=torch.int32, device=device)
    x = torch.empty_like(b)
    st = torch.cuda.current_stream()
    P = _lib.ptr

    def fused():
        _lib.check(L.bdilu_factor_solve(P(work), work.stride(0), nsys, *shape.dims(), P(ipiv), P(rbc), P(cbr), P(info),
                                        P(b), P(x), nrhs, n, n * nrhs, P(flags), st.cuda_stream))

    def factor_only():
        _lib.check(L.bdilu_factor(P(work), work.stride(0), nsys, *shape.dims(), P(ipiv), P(rbc), P(cbr), P(info),
                                  st.cuda_stream))

    def solve_only():
        _lib.check(L.bdilu_solve(P(work), work.stride(0), nsys, *shape.dims(), P(rbc), P(cbr), P(b), P(x), nrhs, n,
                                 n * nrhs, None, P(flags), st.cuda_stream))

    def timed(fn, reps, restore=True):
        evs = []
        for _ in range(reps):
            if restore:
                work.copy_(pristine)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            fn()
            e1.record(st)
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return [a.elapsed_time(c) for a, c in evs]

    sampler = ClockSampler(local_rank)
    sampler.start()
    timed(fused, args.warmup)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    barrier()
    L.bdilu_launch_count_reset()
    t_steps = timed(fused, args.steps)
    barrier()
    launches = int(L.bdilu_launch_count())
    t_total_ms = sum(t_steps)
    if world > 1:
        tt = torch.tensor([t_total_ms], dtype=torch.float64, device=device)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_total_ms = float(tt.item())
    ms_per_step = t_total_ms / args.steps
    value = world * nsys * N / (ms_per_step * 1e-3)

    # sanity on what the timed steps produced + the residual norms (NCCL gather only here)
    assert int(info.abs().max().item()) == 0 and int(flags.abs().max().item()) == 0
    norms = dev.residual_norm2(pristine, b, x, N, w, d)
    rel = (norms.sqrt() / b.norm(dim=2)).reshape(-1)  # ||b - A x|| / ||b|| per (system, vector): ILU is approximate
    if world > 1:
        gathered = [torch.empty_like(rel) for _ in range(world)]
        dist.all_gather(gathered, rel)  # the path's only collective: a few KB of norms over NCCL
        rel = torch.cat(gathered)
    rel_res = float(rel.max().item())

    # per-kernel durations, measured live on the launching stream (roofline of the dominant kernel)
    reps = max(3, min(args.steps, 10))
    t_factor = timed(factor_only, reps)
    work.copy_(pristine)
    factor_only()
    t_solve = timed(solve_only, reps, restore=False)
    kf, ks = sum(t_factor) / reps, sum(t_solve) / reps
    dom = "factor" if kf >= ks else "solve"
    dom_ms = max(kf, ks)
    achieved = nsys * AB[dom] / (dom_ms * 1e-3) / 1e9
    fused_gbs = nsys * AB["fused"] / (ms_per_step * 1e-3) / 1e9
    # kernel-name prefixes as `ncu` prints them (tools/ncu_traffic.py keys profiles/traffic.json the same way)
    wp = 2
    while wp < w:
        wp *= 2
    if dom == "factor":
        kname = "factor_kernel<%d," % wp if w <= 32 else "factor_wide_kernel"
    elif w > 32:
        kname = "solve_mrhs_wave_kernel<64,"
    elif w == 32 and args.workload == "cfg3":
        kname = "solve_quad32_kernel<"
    else:
        kname = "solve_wave_kernel" if w in (4, 16) and nsys * nrhs <= 8 * 148 else "solve_chain_kernel<%d>" % wp
    roofline = {"bound": "hbm", "kernel": kname,
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic(kname, nsys),
                "peak_source": peak_src, "algorithmic_bytes_per_launch": nsys * AB[dom],
                "kernel_ms": {"factor": kf, "solve": ks},
                "fused_step": {"achieved": fused_gbs, "frac": fused_gbs / peak,
                               "algorithmic_bytes_per_step": nsys * AB["fused"]}}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "%s: blockw=%d nblocks=%d ndiag=%d nsys=%d/GPU nrhs=%d fused factor+solve"
                                   % (args.workload, w, N, d, nsys, nrhs),
                       "family": "D (diagonally dominant, SURVEY 8d)", "l2": "inputs (%.2f GB/step) larger than L2"
                       % (pristine.numel() * 8 / 1e9), "timing": "sum of per-step CUDA-event intervals; the "
                       "device-to-device restore of the in-place-factored batch between steps is untimed",
                       "parallelism": "systems sharded %d/GPU, no data-path collective" % nsys},
            "roofline": roofline, "gpu_launches": launches, "max_rel_residual": rel_res}

    # e2e through the host-pointer C ABI on every rank's slab: pinned host buffers, H2D + fused + D2H inside the timed
    # region; barrier on both sides, max over ranks (each GPU has its own PCIe link, the host memory is shared)
    numa_node = bind_host_to_gpu_numa(local_rank) if world > 1 else None
    h_data = torch.empty(pristine.shape, dtype=torch.float64, pin_memory=True)
    h_b = torch.empty(b.shape, dtype=torch.float64, pin_memory=True)
    h_x = torch.empty(b.shape, dtype=torch.float64, pin_memory=True)
    h_data.copy_(pristine)
    h_b.copy_(b)
    torch.cuda.synchronize()
    h_info = np.zeros(nsys, np.int32)
    h_flags = np.zeros(nsys * nrhs, np.int32)

    def e2e_call():
        _lib.check(L.bdilu_host_factor_solve(P(h_data), h_data.stride(0), nsys, *shape.dims(), P(h_b), P(h_x),
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
        tt = torch.tensor([t_e2e], dtype=torch.float64, device=device)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_e2e = float(tt.item())
    assert not h_info.any() and not h_flags.any()
    e2e_diff = float((h_x.to(device) - x).abs().max().item())
    assert e2e_diff <= 1e-12 * float(x.abs().max().item()), "host-pointer path disagrees with the device path"
    line["e2e"] = {"value": world * nsys * N / t_e2e, "unit": UNIT,
                   "h2d_bytes_per_step": int(world * (h_data.numel() * 8 + h_b.numel() * 8)),
                   "d2h_bytes_per_step": int(world * (h_x.numel() * 8 + 4 * nsys + 4 * nsys * nrhs)), "ms_per_step": 1e3 * t_e2e,
                   "steps": e2e_steps, "n_gpus": world, "bit_identical_to_device_path": e2e_diff == 0.0,
                   "host_numa_node_rank0": numa_node,
                   "api": "bdilu_host_factor_solve (host-pointer C ABI; pinned host buffers; slabs of systems pipelined "
                          "over 3 streams; one call per rank on its own slab of the batch, max over ranks)"}
    sampler.stop_flag = True
    sampler.join(timeout=1.0)
    line["clocks"] = sampler.summary()

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline_leg(w, N, d, nsys, nrhs, h_data.numpy(), h_b.numpy())
        xs = cb.pop("x")
        ns, nr = cb.pop("ns"), cb.pop("nr")
        gx = x[:ns, :nr].cpu().numpy()
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
