#!/usr/bin/env python
"""
bench.py - the driver-facing benchmark of the pycc_b200 hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload o,v]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Metric (BASELINE.json): CCSD seconds per iteration (+ (T) seconds and FP64 tensor TFLOP/s vs peak)
on synthetic, symmetry-correct RHF integrals; default workload = BASELINE configs[2], CCSD(T) at
o=30 v=250 FP64 - the largest named FP64 configuration whose <ab|ef> block (31 GB) fits one GPU, and
the one named for 1 and 8 GPUs (configs[1], o=20 v=150, is ``--workload 20,150``).
One "step" = one pass of the solve_cc loop body (pycc/ccwfn.py:274-329): residuals, amplitude
update + norm, energy, DIIS push + extrapolation.  Prints ONE JSON line (rank 0).

  value        seconds per CCSD iteration, amplitudes and integrals resident in HBM
  e2e          the same iteration with the amplitudes living in pinned HOST memory: H2D of
               (F, t1, t2) and D2H of the updated (t1, t2) + energy inside the timed region
  roofline     the o^2 v^4 particle-particle ladder kernel (contract_f64_kernel): algorithmic
               2 o^2 v^3 v_local flops / CUDA-event time of that launch, against the FP64
               tensor (DMMA) issue-rate peak measured on this pool (profiles/r01_peaks.json)
  cpu_baseline the NumPy oracle (a port of the reference's algorithm) timed on the host cores
  triples      (T) seconds (t_tjl over i>=j>=k) and its TFLOP/s, timed after the K steps

``--impl reference`` times the reference's CPU algorithm (the oracle port; the reference itself
is pure Python + psi4 and cannot be imported on the GPU box) on the same workload.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

DMMA_PEAK_FALLBACK = 37.0   # TFLOP/s, profiles/r01_peaks.json (measured on this pool's B200)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="30,250", help="o,v (default: BASELINE.json configs[2], CCSD(T) o=30 v=250)")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--precision", default="DP", choices=["DP", "SP"],
                    help="SP = mixed mode: FP32 contractions (3xTF32), FP64 energy/norm/DIIS")
    ap.add_argument("--no-triples", action="store_true", help="skip the (T) timing")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--path", default="ccsd", choices=["ccsd", "lambda"],
                    help="ccsd: the north-star CCSD iteration (default); lambda: HBAR build + Lambda iteration "
                         "(SURVEY.md 8(f) rank 2)")
    return ap.parse_args()


def flops_ccsd_iteration(o, v):
    """Algorithmic flops of one reference residuals() call."""
    return 2 * (o**2 * v**4 + 9 * o**3 * v**3 + 2 * o**4 * v**2 + 7 * o**2 * v**3 + 8 * o**3 * v**2
                + o * v**3 + 2 * o**4 * v + 12 * o**2 * v**2 + o**3 * v + 3 * o * v**2 + 3 * o**2 * v)


def flops_triples(o, v):
    return (o * (o + 1) * (o + 2) // 6) * (12 * v**4 + 12 * o * v**3)


def dmma_peak():
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r01_peaks.json")))
        return float(d["dmma_w8_c2_tflops"]), "measured DMMA issue-rate peak, profiles/r01_peaks.json"
    except Exception:
        return DMMA_PEAK_FALLBACK, "fallback 37.0 TFLOP/s (earlier DMMA measurement on this pool)"


class ClockSampler(object):
    """Samples SM clock and clock-event reasons of one GPU every 100 ms while running."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake", 0x2: "applications_clocks_setting"}

    def __init__(self, index):
        self.samples, self.mask, self.max_mhz, self.ok = [], 0, None, False
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False
        self.thread = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                try:
                    self.mask |= int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    self.mask |= int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        if self.ok:
            self.thread.start()

    def stop(self):
        self._stop.set()
        if self.ok:
            self.thread.join(timeout=2)
        s = sorted(self.samples)
        med = s[len(s) // 2] if s else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz,
                "reasons": [n for b, n in self.REASONS.items() if self.mask & b], "samples": len(s)}


_OUT_FD = None


def emit(obj):
    """The ONE JSON line, on the process's real stdout (see main)."""
    line = (json.dumps(obj) + "\n").encode()
    if _OUT_FD is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        os.write(_OUT_FD, line)


def make_factor(o, v, seed):
    """Seeded factor B (scaled so that E_MP2 = -0.10 Eh), Fock matrix F."""
    from pycc_b200.synthetic import make_factor as mf, calibrated_scale
    import numpy as np
    B, eps, P = mf(o, v, seed)
    s = calibrated_scale(B, eps, o, -0.10)
    return B * s, np.diag(eps).copy(), s


def set_host_threads():
    """All host cores for the NumPy oracle (torchrun exports OMP_NUM_THREADS=1)."""
    n = os.cpu_count() or 1
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=n)
    except Exception:
        pass
    return n


def warm_numpy():
    """Tiny oracle iteration so that NumPy/BLAS thread pools exist before anything is timed."""
    import numpy as np
    from pycc_b200.synthetic import SyntheticSystem
    s = SyntheticSystem(3, 6, seed=1)
    oracle_iteration(s.F, s.blocks, np.zeros((3, 6)), 0.01 * s.blocks["oovv"])


def oracle_iteration(F, blocks, t1, t2):
    """One reference iteration on the CPU: residuals + update + energy (pycc/ccwfn.py:274-296)."""
    import numpy as np
    from oracle import ccsd_oracle as oc
    blk = blocks if isinstance(blocks, oc.Blocks) else oc.Blocks(blocks)
    Dia, Dijab = oc.denominators(F, blk.no)
    r1, r2 = oc.residuals(F, blk, t1, t2)
    i1, i2 = r1 / Dia, r2 / Dijab
    t1n, t2n = t1 + i1, t2 + i2
    rms = float(np.sqrt((i1 * i1).sum() + (i2 * i2).sum()))
    return r1, r2, t1n, t2n, oc.cc_energy(F, blk, t1n, t2n), rms



def workload_config(o, v, sp, world, seed, scale):
    """The ``config`` object of the JSON line - identical in both arms (``--impl ours|reference``)."""
    from pycc_b200.parallel import virtual_slice
    vl = max(hi - lo for lo, hi in (virtual_slice(v, r, world) for r in range(world)))
    return {"workload": "RHF-CCSD(T): CCSD iteration + (T), o=%d v=%d %s" % (o, v, "FP32 contractions / FP64 energy+DIIS" if sp else "FP64"),
            "o": o, "v": v,
            "parallelism": ("virtual-index sharding x%d: local symmetrise + in-place reduce-scatter(r2 by (i,j) blocks) -> split "
                            "update/energy/DIIS -> in-place allgather(t2); Fmi/Wmnij sliced over the contracted index; "
                            "(T) pairs dealt over ranks" % world) if world > 1 else "1 GPU",
            "l2": "inputs larger than L2 (vvvv rows %.2f GB per rank)" % (vl * v**3 * (4 if sp else 8) / 1e9),
            "seed": seed, "eri_scale": round(scale, 6), "diis": 8}


ORACLE_FITS_V = 160      # NumPy oracle on the host: full <ab|ef> block up to v = 160 (5.2 GB)
TRIPLET_SAMPLE = lambda o: [(o - 1, o // 2, 0), (o - 1, o - 1, o - 2), (o // 2, 1, 1), (3, 2, 1)]


class FactorChecker(object):
    """Independent full-size checker (SURVEY.md 8(d) "Correctness gates run with every timing"): the
    reference's formulas (oracle/torch_oracle.py, pinned to the golden vectors) evaluated with
    torch.einsum = cuBLAS FP64 on the device, on integrals REGENERATED from the seeded factor B row by
    row - neither the engine's kernels nor its stored blocks are involved.  With several ranks each one
    checks the ladder rows it owns (the pieces are summed), everything else is evaluated on every rank."""

    def __init__(self, B, F, o, dev, vslice):
        import torch
        from oracle import torch_oracle as to
        self.to, self.torch = to, torch
        Bd = torch.as_tensor(B).to(device=dev, dtype=torch.float64)
        n = Bd.shape[1]
        v = n - o
        Bo, Bv = Bd[:, :o, :], Bd[:, o:, :]
        self.Boo, self.Bov, self.Bvv = Bo[:, :, :o].contiguous(), Bo[:, :, o:].contiguous(), Bv[:, :, o:].contiguous()
        ein = torch.einsum
        small = {"oooo": ein('Qmi,Qnj->mnij', self.Boo, self.Boo), "ooov": ein('Qmi,Qne->mnie', self.Boo, self.Bov),
                 "oovv": ein('Qia,Qjb->ijab', self.Bov, self.Bov), "ovov": ein('Qmj,Qbe->mbje', self.Boo, self.Bvv),
                 "ovvv": lambda m: ein('Qe,Qbf->bef', self.Bov[:, m, :], self.Bvv),
                 "vvvv": lambda lo, hi: ein('Qae,Qbf->abef', self.Bvv[:, lo:hi, :], self.Bvv)}
        chunk = max(1, int((3 << 30) // (8 * v**3)))          # <= 3 GB of regenerated <ab|ef> rows at a time
        self.blk = to.Blocks(small, device=dev, vv_rows=vslice, chunk=chunk)
        self.F = torch.as_tensor(F).to(device=dev, dtype=torch.float64)

    def residuals(self, t1, t2, comm):
        """(r1, r2) of the reference algorithm, identical on every rank."""
        torch, to = self.torch, self.to
        r1, rest = to.residuals(self.F, self.blk, t1, t2, skip_ladder=True)
        _, lad = to.residuals(self.F, self.blk, t1, t2, ladder_only=True)
        comm.allreduce_(lad)
        return r1, rest + lad + lad.permute(1, 0, 3, 2)

    def energy(self, t1, t2):
        return self.to.cc_energy(self.F, self.blk, t1, t2)

    def et_triplet(self, t1, t2, i, j, k):
        return self.to.et_triplet(self.F, self.blk, t1, t2, i, j, k)


def run_ours(args):
    import numpy as np
    import torch
    import contextlib
    import io
    o, v = [int(x) for x in args.workload.split(",")]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the pycc_b200 engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    group = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        group = "world"
    import pycc_b200
    import pycc_b200.cctriples as cct
    from pycc_b200.parallel import virtual_slice
    from pycc_b200.synthetic import device_blocks
    from pycc_b200.utils import helper_diis

    B, F, scale = make_factor(o, v, args.seed)
    vslice = virtual_slice(v, rank, world)
    want_t = not args.no_triples
    sp = args.precision == "SP"
    blocks = device_blocks(B, o, dev, vslice=vslice, keep_full_ovvv=want_t or world == 1,
                           dtype=torch.float32 if sp else torch.float64)
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):
        w = pycc_b200.ccwfn.from_arrays(F, blocks, o, model="CCSD(T)" if want_t else "CCSD", group=group,
                                        precision=args.precision)
    eng, be = w.engine, w.backend
    diis = w.make_diis(8)

    def barrier():
        torch.cuda.synchronize()
        w.comm.barrier()
        torch.cuda.synchronize()

    # ---- warm-up + K timed steps (amplitudes and integrals resident in HBM) ------------------
    niter = 0
    for _ in range(max(args.warmup, 0)):
        niter += 1
        w._iteration(diis, niter, 1)
    clocks = ClockSampler(local_rank)
    barrier()
    clocks.start()
    launches0 = be.launch_count()
    eng.ladder_events = []
    w.phase_events = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    last = None
    for _ in range(args.steps):
        niter += 1
        last = w._iteration(diis, niter, 1)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    pe, w.phase_events = w.phase_events, None
    nst = max(len(pe) // 4, 1)
    phases = {"residuals_incl_exchange_ms": round(sum(pe[4 * q].elapsed_time(pe[4 * q + 1]) for q in range(len(pe) // 4)) / nst, 4),
              "update_energy_ms": round(sum(pe[4 * q + 1].elapsed_time(pe[4 * q + 2]) for q in range(len(pe) // 4)) / nst, 4),
              "diis_incl_allgather_ms": round(sum(pe[4 * q + 2].elapsed_time(pe[4 * q + 3]) for q in range(len(pe) // 4)) / nst, 4),
              "note": "rank 0, CUDA events on the launch stream at the phase boundaries of the timed steps (idle gaps included)"}
    clk = clocks.stop()
    launches = be.launch_count() - launches0
    ladder_ms = [a.elapsed_time(b) for a, b in eng.ladder_events]
    eng.ladder_events = None
    tt = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_step = float(tt.item()) / max(args.steps, 1)
    fi = torch.tensor([float(eng.flops_issued())], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(fi, op=dist.ReduceOp.SUM)
    flops_issued = float(fi.item())

    # ---- roofline of the dominant kernel (ladder launch on this rank) ------------------------
    peak, peak_src = dmma_peak()
    if sp:
        # the SP ladder runs on the tcgen05 kernel: TF32 issues at half the dense bf16 rate, three TF32
        # products per FP32-accurate product
        try:
            mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            peak = float(mp["bf16_tflops"]) / 2.0 / 3.0
            peak_src = "MEASURED_PEAKS.json dense bf16 burst %.1f TFLOP/s / 2 (TF32) / 3 (3xTF32 split)" % float(mp["bf16_tflops"])
        except Exception:
            peak = 1634.5 / 6.0
            peak_src = "fallback: dense bf16 1634.5 TFLOP/s (this pool's B200, earlier MEASURED_PEAKS.json) / 2 (TF32) / 3 (3xTF32 split)"
    vl = vslice[1] - vslice[0]
    lad_flops = 2.0 * o * o * vl * v**3
    lad_ms = sum(ladder_ms) / max(len(ladder_ms), 1)
    lad_tf = lad_flops / (lad_ms * 1e-3) / 1e12 if lad_ms > 0 else 0.0
    roofline = {"bound": "tensor", "kernel": "%s (ladder r2 += 1/2 tau <ab|ef>)" % ("contract_tf32x3_tc05_2sm_kernel (+ contract_tf32x3_tc05_kernel for an odd tile column)" if sp else "contract_f64_tma_kernel"),
                "achieved": round(lad_tf, 3), "peak": peak, "unit": "TFLOP/s", "frac": round(lad_tf / peak, 4),
                "traffic": None, "launch_ms": round(lad_ms, 4), "flops_per_launch": lad_flops,
                "peak_source": peak_src + (" (cuBLAS SGEMM on the same box: 63.9 TFLOP/s; the mma.sync 3xTF32 kernel this replaced: 66.7)" if sp else
                                           " (FP64 has no entry in MEASURED_PEAKS.json; cuBLAS DGEMM on the "
                                           "same box: 35.5 TFLOP/s)")}

    try:        # DRAM bytes of that launch from the committed ncu capture of the same configuration, if any
        tr = json.load(open(os.path.join(ROOT, "profiles", "r01_ncu_traffic.json")))["%d,%d,%d,%s" % (o, v, world, args.precision)]
        roofline["traffic"] = tr["traffic_bytes"]
        roofline["traffic_unit"] = "bytes/launch (algorithmic %d)" % tr["algorithmic_bytes"]
        roofline["traffic_source"] = tr["source"]
    except Exception:
        pass

    # ---- e2e: amplitudes live in pinned host memory between iterations -------------------------
    e2e = None
    if not args.no_e2e:
        # The amplitude vector lives in pinned host memory ONCE: rank g holds its shard [lo, hi) of the
        # flattened (t1, t2) (the whole vector on one GPU).  Every step copies the Fock matrix and the
        # shard H2D, rebuilds the replicated amplitudes (allgather over NVLink when N > 1), runs the
        # iteration, and copies the updated shard + the energy / norm back D2H.
        F_pin = torch.as_tensor(F).pin_memory()
        lo_, hi_ = w.amplitude_shard()
        amp_pin = torch.empty(hi_ - lo_, dtype=w.t2.dtype).pin_memory()
        w.store_amplitude_shard(amp_pin)
        torch.cuda.current_stream().synchronize()
        w_ = amp_pin.element_size()
        h2d = F_pin.numel() * 8 * world + (w.t1.numel() + w.t2.numel()) * w_
        d2h = (w.t1.numel() + w.t2.numel()) * w_ + 16 * world

        def e2e_step(n):
            eng.F.copy_(F_pin, non_blocking=True)
            eng.eps.copy_(torch.diagonal(eng.F))
            w.load_amplitude_shard(amp_pin)
            ecc, rms = w._iteration(diis, n, 1)          # energy and norm come back as host scalars
            w.store_amplitude_shard(amp_pin)
            torch.cuda.current_stream().synchronize()
            return ecc

        niter += 1
        e2e_step(niter)
        barrier()
        ksteps = max(min(args.steps, 5), 1)
        t0 = time.perf_counter()
        for _ in range(ksteps):
            niter += 1
            e2e_step(niter)
        barrier()
        dt = torch.tensor([(time.perf_counter() - t0) / ksteps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": round(float(dt.item()), 6), "unit": "s/iteration", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "steps": ksteps,
               "bytes_scope": "whole job (sum over ranks): the amplitude vector is sharded over the ranks' "
                              "host buffers, F is copied by every rank"}

    # ---- (T): timed once after the CCSD steps (all ranks; max over ranks) ---------------------
    triples = None
    if want_t:
        tri = cct._engine(w)
        tri.energy(tri.my_pairs()[:1])                 # warm-up: one occupied pair
        barrier()
        t0 = time.perf_counter()
        et = cct.t_tjl(w)
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        tsec = float(dt.item())
        triples = {"seconds": round(tsec, 4), "E_T": et, "flops": flops_triples(o, v),
                   "tflops": round(flops_triples(o, v) / tsec / 1e12, 3),
                   "frac_of_peak": round(flops_triples(o, v) / tsec / 1e12 / (peak * world), 4)}

    # ---- parity of the timed engine, every workload and every N (gate: no parity, no number) ---------
    # amplitudes pushed off the fixed point (at convergence every residual is ~0 and the check is void)
    t1h = (1.1 * w.t1.double() + 0.01)
    t2h = 1.1 * w.t2.double()
    r1g, r2g = w.residuals(F, t1h, t2h)               # the engine's answer (all ranks take part)
    e_eng = eng.energy()
    chk = FactorChecker(B, F, o, dev, vslice)
    r1c, r2c = chk.residuals(t1h, t2h, w.comm)
    tol_r = 2e-5 if sp else 1e-10
    parity = {"max_abs_dr1": float((r1g.double() - r1c).abs().max()), "max_abs_dr2": float((r2g.double() - r2c).abs().max()),
              "max_abs_r2": float(r2c.abs().max()),
              "abs_dE_same_amplitudes": abs(e_eng - chk.energy(w.t1, w.t2)),
              "tolerance": tol_r,
              "checked": "residuals() of the timed engine on perturbed (t1, t2) and its energy, o=%d v=%d, %d GPU(s), against "
                         "oracle/torch_oracle.py (reference formulas, torch.einsum/cuBLAS FP64 on the device) on integrals "
                         "regenerated from the factor B; each rank checks the <ab|ef> rows it owns" % (o, v, world)}
    # the ITERATION path (reduce-scatter of r2 -> update of this rank's (i,j) blocks -> fused energy -> allgather):
    # one solver step from the same perturbed amplitudes must give t + r/D, its norm and the energy of the
    # checker's residuals (pycc/ccwfn.py:287-296)
    keep1, keep2 = w.t1.clone(), w.t2.clone()
    w._load(F, t1h, t2h)
    e_it, rms_it = w._iteration(None, 0, 1, extrapolate=False)
    eng.sync_amplitudes()
    epsd = chk.F.diagonal()
    eo, ev = epsd[:o], epsd[o:]
    i1 = r1c / (eo[:, None] - ev[None, :])
    r2c /= (eo[:, None, None, None] + eo[None, :, None, None] - ev[None, None, :, None] - ev[None, None, None, :])
    rms_c = float(torch.sqrt((i1 * i1).sum() + (r2c * r2c).sum()))
    i1 += t1h
    r2c += t2h                                           # r2c now holds the expected new t2
    parity["iteration_max_abs_dt1"] = float((w.t1.double() - i1).abs().max())
    parity["iteration_max_abs_dt2"] = float((w.t2.double() - r2c).abs().max())
    parity["iteration_abs_dE"] = abs(e_it - chk.energy(i1, r2c))
    parity["iteration_rel_drms"] = abs(rms_it - rms_c) / max(rms_c, 1e-300)
    # mixed precision: residuals come from FP32 (3xTF32) products and the increments are stored in FP32, so the
    # increment norm carries FP32 rounding (observed 5e-5 relative at (40,300)); the north-star bars are the
    # energy (1e-6 Eh) and, in FP64, the amplitudes (1e-9)
    parity["tolerance_iteration"] = {"dt": 1e-5 if sp else 1e-9, "dE": 1e-6 if sp else 1e-10, "rel_rms": 5e-4 if sp else 1e-9}
    it_ok = (max(parity["iteration_max_abs_dt1"], parity["iteration_max_abs_dt2"]) <= parity["tolerance_iteration"]["dt"]
             and parity["iteration_abs_dE"] <= parity["tolerance_iteration"]["dE"]
             and parity["iteration_rel_drms"] <= parity["tolerance_iteration"]["rel_rms"])
    w.t1.copy_(keep1)
    w.t2.copy_(keep2)
    eng.set_fock(eng.F0)
    del r1c, r2c, r1g, r2g, i1, keep1, keep2
    if want_t:
        tri = cct._engine(w)
        sample = [t for t in TRIPLET_SAMPLE(o) if t[0] < o and t[1] <= t[0] and t[2] <= t[1]]
        pairs = sorted(set((i, j) for (i, j, k) in sample))
        _, per, trip = tri.energy(pairs)
        got = dict(zip(trip, per.tolist()))
        dev_t = [abs(got[t] - chk.et_triplet(w.t1.double(), w.t2.double(), *t)) for t in sample]
        parity["max_abs_dE_T_triplet"] = max(dev_t)
        parity["triplets_checked"] = sample
        parity["tolerance_triplet"] = 1e-7 if sp else 1e-12
    ok = parity["max_abs_dr1"] <= tol_r and parity["max_abs_dr2"] <= tol_r and parity["abs_dE_same_amplitudes"] <= (1e-6 if sp else 1e-10)
    ok = ok and it_ok
    if want_t:
        ok = ok and parity["max_abs_dE_T_triplet"] <= parity["tolerance_triplet"]
    parity["passed"] = bool(ok)
    del chk
    torch.cuda.empty_cache()

    # ---- CPU baseline (rank 0, N=1 only): the NumPy oracle on the host cores, a bounded sample ----------
    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        from oracle import ccsd_oracle as oc
        cores = set_host_threads()
        t1n, t2n = t1h.cpu().numpy(), t2h.cpu().numpy()
        hb = {k: t.double().cpu().numpy() for k, t in w.H.blocks.items() if k != "vvvv"}
        warm_numpy()
        if v <= ORACLE_FITS_V:
            hb["vvvv"] = w.H.blocks["vvvv"].double().cpu().numpy()
            blk = oc.Blocks(hb)
            t0 = time.perf_counter()
            r1c, r2c, _, _, ecpu, _ = oracle_iteration(F, blk, t1n, t2n)
            cpu_s = time.perf_counter() - t0
            r1g, r2g = w.residuals(F, t1h, t2h)
            parity["numpy_oracle_max_abs_dr2"] = float(np.abs(r2g.double().cpu().numpy() - r2c).max())
            sample_txt = ("1 full CCSD iteration (residuals + update + energy) of the NumPy oracle "
                          "(port of pycc/ccwfn.py:331-366) at o=%d v=%d, after a tiny warm-up" % (o, v))
            extrap = False
        else:
            # bounded: every builder except the ladder at full size + the ladder einsum on a slice of rows
            hb["vvvv"] = None
            blk = oc.Blocks(hb)
            t0 = time.perf_counter()
            oracle_iteration(F, blk, t1n, t2n)
            rest_s = time.perf_counter() - t0
            c = max(2, min(v, int(8e11 / (2.0 * o * o * v**3))))          # ~0.8 TFLOP of ladder work
            rows = eng.vvvv_loc[:c].double().cpu().numpy()
            tt = oc.tau(t1n, t2n)
            t0 = time.perf_counter()
            oc._ein('ijef,abef->ijab', tt, rows)
            lad_s = (time.perf_counter() - t0) * v / c
            cpu_s = rest_s + lad_s
            sample_txt = ("NumPy oracle at o=%d v=%d: all builders, update and energy at full size (%.1f s) + the o^2 v^4 "
                          "ladder einsum on %d of %d <ab|ef> rows, scaled by %d/%d (%.1f s)" % (o, v, rest_s, c, v, v, c, lad_s))
            extrap = True
        cpu_baseline = {"value": round(cpu_s, 4), "unit": "s/iteration", "cores": cores, "kind": "port",
                        "extrapolated": extrap, "sample": sample_txt}
        if want_t:
            from oracle import triples_oracle as ot
            sample = [t for t in TRIPLET_SAMPLE(o) if t[0] < o and t[1] <= t[0] and t[2] <= t[1]]
            t1o, t2o = w.t1.double().cpu().numpy(), w.t2.double().cpu().numpy()
            t0 = time.perf_counter()
            ecpu_t = [ot.et_triplet(F, blk, t1o, t2o, i, j, k) for (i, j, k) in sample]
            per = (time.perf_counter() - t0) / len(sample)
            ntrip = o * (o + 1) * (o + 2) // 6
            cpu_baseline["triples"] = {"value": round(per * ntrip, 2), "unit": "s", "extrapolated": True,
                                       "sample": "%d triplets timed (vectorised t_tjl body, NumPy), x %d triplets" % (len(sample), ntrip)}
            parity["numpy_oracle_max_abs_dE_T_triplet"] = max(abs(a - got[t]) for a, t in zip(ecpu_t, sample))

    hb = torch.tensor([torch.cuda.max_memory_allocated(dev) / 1e9], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(hb, op=dist.ReduceOp.MAX)
    hbm_peak_gb = round(float(hb.item()), 2)
    if rank == 0:
        fl = flops_ccsd_iteration(o, v)
        out = {"metric": "ccsd_seconds_per_iteration", "value": round(ms_step * 1e-3, 6), "unit": "s/iteration",
               "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 4),
               "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f32" if sp else "f64",
               "data": "synthetic (seeded factorised 8-fold-symmetric ERIs, diagonal Fock; SURVEY.md 8(d))",
               "config": workload_config(o, v, sp, world, args.seed, scale),
               "tflops": round(fl / (ms_step * 1e-3) / 1e12, 3),
               "tflops_frac_of_peak": round(fl / (ms_step * 1e-3) / 1e12 / (peak * world), 4),
               "flops_per_iteration": fl, "flops_issued": flops_issued,
               "tflops_issued": round(flops_issued / (ms_step * 1e-3) / 1e12, 3),
               "tflops_issued_frac_of_peak": round(flops_issued / (ms_step * 1e-3) / 1e12 / (peak * world), 4),
               "energy": last[0] if last else None, "rms": last[1] if last else None,
               "gpu_launches": int(launches), "clocks": clk, "roofline": roofline, "e2e": e2e, "phases": phases,
               "hbm_peak_gb_per_rank": hbm_peak_gb,
               "triples": triples, "cpu_baseline": cpu_baseline, "parity": parity}
        if not parity["passed"]:
            # a number without a green parity check is not a number: print the evidence, not the metric
            out = {"error": "parity check failed - timing withheld", "parity": parity, "config": out["config"]}
        emit(out)
    if world > 1:
        dist.destroy_process_group()
    if not parity["passed"]:
        sys.exit(3)


def run_reference(args):
    """The reference's CPU algorithm for the same workload, on all host cores.  The reference itself
    (pure Python + psi4 + opt_einsum) is not importable on the GPU box, so this is the oracle port
    (kind='port').  Each step is one full CCSD iteration incl. DIIS; the number of steps is cut down when
    the first one shows that K+W of them would not finish in ~4 minutes.  Shapes whose <ab|ef> block does
    not fit the host (v > 160) run the o^2 v^4 ladder chunk by chunk over rows manufactured from the factor
    on the fly - the same einsum per chunk; the seconds spent manufacturing rows are set-up and not timed."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    from oracle import ccsd_oracle as oc
    from pycc_b200.synthetic import _blk
    o, v = [int(x) for x in args.workload.split(",")]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = set_host_threads()
    B, F, scale = make_factor(o, v, args.seed)
    oo, vv = slice(0, o), slice(o, o + v)
    blocks = {"oooo": _blk(B, oo, oo, oo, oo), "ooov": _blk(B, oo, oo, oo, vv), "oovv": _blk(B, oo, oo, vv, vv),
              "ovov": _blk(B, oo, vv, oo, vv), "ovvv": _blk(B, oo, vv, vv, vv)}
    streamed = v > ORACLE_FITS_V
    if streamed:
        blocks["vvvv"] = lambda lo, hi: _blk(B, slice(o + lo, o + hi), vv, vv, vv)
        blocks["vvvv_chunk"] = max(1, int((2 << 30) // (8 * v**3)))
    else:
        blocks["vvvv"] = _blk(B, vv, vv, vv, vv)
    blk = oc.Blocks(blocks)
    Dia, Dijab = oc.denominators(F, o)
    t1 = np.zeros((o, v))
    t2 = blocks["oovv"] / Dijab
    diis = oc.Diis(t1, t2, 8)
    warm_numpy()

    def step(t1, t2):
        """Returns (t1, t2, energy, timed seconds of this step)."""
        g0 = blk.gen_seconds
        t0 = time.perf_counter()
        _, _, t1, t2, e, rms = oracle_iteration(F, blk, t1, t2)
        diis.add_error_vector(t1, t2)
        t1, t2 = diis.extrapolate(t1, t2)
        return t1, t2, e, (time.perf_counter() - t0) - (blk.gen_seconds - g0)

    w0 = time.perf_counter()
    t1, t2, e, first = step(t1, t2)               # first (warm-up) step, also the calibration
    wall_first = time.perf_counter() - w0
    budget = 200.0         # wall-clock seconds for all steps (row manufacturing included, although untimed)
    warm = max(min(args.warmup - 1, int((0.25 * budget) / max(wall_first, 1e-3))), 0)
    for _ in range(warm):
        t1, t2, e, _s = step(t1, t2)
    steps = max(min(args.steps, int((budget - (1 + warm) * wall_first) / max(wall_first, 1e-3))), 1)
    tot = 0.0
    for _ in range(steps):
        t1, t2, e, dt = step(t1, t2)
        tot += dt
    sec = tot / steps
    fl = flops_ccsd_iteration(o, v)
    out = {"impl": "reference", "metric": "ccsd_seconds_per_iteration", "value": round(sec, 4), "unit": "s/iteration",
           "n_gpus": world, "steps": steps, "warmup": 1 + warm,
           "ms_per_step": round(sec * 1e3, 2), "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic (same seeded system as the GPU arm)",
           "config": workload_config(o, v, args.precision == "SP", world, args.seed, scale),
           "tflops": round(fl / sec / 1e12, 4), "energy": float(e), "gpu_launches": 0,
           "cpu_baseline": {"value": round(sec, 4), "unit": "s/iteration", "cores": cores, "kind": "port",
                            "sample": "%d full CCSD iterations (residuals + update + energy + DIIS) of the NumPy "
                                      "oracle at o=%d v=%d (requested %d; bounded to ~4 min)%s"
                                      % (steps, o, v, args.steps,
                                         "; <ab|ef> rows manufactured chunk by chunk from the factor, untimed" if streamed else "")},
           "e2e": {"value": round(sec, 4), "unit": "s/iteration", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(out)


def run_lambda(args):
    """HBAR build + Lambda iterations (pycc/cchbar.py, pycc/cclambda.py:141-260) on one GPU.  A step is
    one pass of the solve_lambda loop body: Goo/Gvv, r_L1, r_L2, l += r/D with the norm, pseudoenergy,
    DIIS push + extrapolation.  Same JSON layout as the CCSD line; the roofline entry is the Lambda
    ladder r2 += 1/2 l2[ijef] Hvvvv[efab]; cpu_baseline is the NumPy oracle (oracle/lambda_oracle.py)."""
    import numpy as np
    import torch
    import contextlib
    import io
    o, v = [int(x) for x in args.workload.split(",")]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the pycc_b200 engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    group = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        group = "world"
    import pycc_b200
    from pycc_b200.parallel import virtual_slice
    from pycc_b200.synthetic import device_blocks
    from pycc_b200.utils import helper_diis
    B, F, scale = make_factor(o, v, args.seed)
    blocks = device_blocks(B, o, dev, vslice=virtual_slice(v, rank, world), keep_full_ovvv=True)
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(F, blocks, o, model="CCSD", group=group, full_ovvv=True)
        d0 = helper_diis(w.t1, w.t2, 8, "DP", backend=w.backend, comm=w.comm)
        for it in range(6):                      # amplitudes a few DIIS iterations from the MP2 guess
            w._iteration(d0, it + 1, 1)
        del d0
        hb = pycc_b200.cchbar(w)
        lam = pycc_b200.cclambda(w, hb)
    be, leng = w.backend, lam.engine

    def sync():
        torch.cuda.synchronize()
        w.comm.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # HBAR build, timed on its own (once per Lambda solve)
    sync()
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0.record()
    for _ in range(3):
        hb.engine.build()
    h1.record()
    sync()
    hbar_ms = max_over_ranks(h0.elapsed_time(h1) / 3)

    diis = helper_diis(lam.l1, lam.l2, 8, "DP", backend=be, comm=w.comm)

    def step(n):
        lecc, rms = lam._iteration()
        diis.add_error_vector(lam.l1, lam.l2)
        lam.l1, lam.l2 = diis.extrapolate(lam.l1, lam.l2)
        return lecc, rms

    niter = 0
    for _ in range(max(args.warmup, 0)):
        niter += 1
        step(niter)
    clocks = ClockSampler(local_rank)
    sync()
    clocks.start()
    launches0 = be.launch_count()
    leng.ladder_events = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    last = None
    for _ in range(args.steps):
        niter += 1
        last = step(niter)
    e1.record()
    sync()
    ms_step = max_over_ranks(e0.elapsed_time(e1)) / max(args.steps, 1)
    clk = clocks.stop()
    launches = be.launch_count() - launches0
    ladder_ms = [a.elapsed_time(b) for a, b in leng.ladder_events]
    leng.ladder_events = None
    peak, peak_src = dmma_peak()
    lad_flops = 2.0 * o * o * float(w.engine.vl) * float(v)**3        # this rank's rows of Hvvvv
    lad_ms = sum(ladder_ms) / max(len(ladder_ms), 1)
    lad_tf = lad_flops / (lad_ms * 1e-3) / 1e12 if lad_ms > 0 else 0.0
    roofline = {"bound": "tensor", "kernel": "contract_f64_tma_kernel (Lambda ladder r2 += 1/2 l2 Hvvvv)",
                "achieved": round(lad_tf, 3), "peak": peak, "unit": "TFLOP/s", "frac": round(lad_tf / peak, 4),
                "traffic": None, "launch_ms": round(lad_ms, 4), "flops_per_launch": lad_flops, "peak_source": peak_src}

    e2e = None
    if not args.no_e2e:
        l1_pin, l2_pin = lam.l1.detach().cpu().pin_memory(), lam.l2.detach().cpu().pin_memory()
        nb = (l1_pin.numel() + l2_pin.numel()) * 8

        def e2e_step(n):
            leng.l1.copy_(l1_pin, non_blocking=True)
            leng.l2.copy_(l2_pin, non_blocking=True)
            lecc, rms = step(n)
            l1_pin.copy_(leng.l1, non_blocking=True)
            l2_pin.copy_(leng.l2, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return lecc

        niter += 1
        e2e_step(niter)
        sync()
        ksteps = max(min(args.steps, 5), 1)
        t0 = time.perf_counter()
        for _ in range(ksteps):
            niter += 1
            e2e_step(niter)
        sync()
        e2e = {"value": round(max_over_ranks((time.perf_counter() - t0) / ksteps), 6), "unit": "s/iteration",
               "h2d_bytes_per_step": nb, "d2h_bytes_per_step": nb + 16, "steps": ksteps}

    cpu_baseline, parity = None, None
    if world == 1 and not args.no_cpu_baseline:
        from oracle import ccsd_oracle as oc
        from oracle import lambda_oracle as ol
        cores = set_host_threads()
        warm_numpy()
        blk = oc.Blocks({k: t.double().cpu().numpy() for k, t in w.H.blocks.items()})
        t1h, t2h = w.t1.cpu().numpy(), w.t2.cpu().numpy()
        l1h = 1.1 * lam.l1.cpu().numpy() + 0.01
        l2h = 1.1 * lam.l2.cpu().numpy()
        r1g, r2g = lam.residuals(F, t1h, t2h, l1h, l2h)          # rebuilds HBAR on the GPU as well
        t0 = time.perf_counter()
        H = ol.build_hbar(F, blk, t1h, t2h)
        cpu_hbar = time.perf_counter() - t0
        t0 = time.perf_counter()
        r1c, r2c = ol.lambda_residuals(blk, H, t2h, l1h, l2h)
        Dia, Dijab = oc.denominators(F, o)
        l2n = l2h + r2c / Dijab
        ol.pseudoenergy(blk, l2n)
        cpu_s = time.perf_counter() - t0
        parity = {"max_abs_dr1": float(np.abs(r1g.cpu().numpy() - r1c).max()),
                  "max_abs_dr2": float(np.abs(r2g.cpu().numpy() - r2c).max()), "max_abs_r2": float(np.abs(r2c).max()),
                  "max_abs_dHvvvo": float(np.abs(hb.Hvvvo.cpu().numpy() - H["Hvvvo"]).max()),
                  "checked": "HBAR + Lambda residuals of the timed engine vs the oracle on the same perturbed "
                             "(l1, l2), o=%d v=%d" % (o, v)}
        cpu_baseline = {"value": round(cpu_s, 4), "unit": "s/iteration", "cores": cores, "kind": "port",
                        "hbar_seconds": round(cpu_hbar, 3),
                        "sample": "1 Lambda iteration (Goo/Gvv, r_L1, r_L2, update, pseudoenergy) of the NumPy oracle "
                                  "(port of pycc/cclambda.py:450-543) at o=%d v=%d; HBAR build timed separately" % (o, v)}

    fi = float(leng.flops_issued())
    hfi = float(hb.engine.flops_issued())
    if world > 1:
        t = torch.tensor([fi, hfi], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        fi, hfi = float(t[0].item()), float(t[1].item())
    if rank != 0:
        dist.destroy_process_group()
        return
    out = {"metric": "lambda_seconds_per_iteration", "value": round(ms_step * 1e-3, 6), "unit": "s/iteration",
           "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 4),
           "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic (seeded factorised 8-fold-symmetric ERIs, diagonal Fock; SURVEY.md 8(d))",
           "config": {"workload": "RHF-CCSD Lambda iteration o=%d v=%d FP64" % (o, v), "o": o, "v": v,
                      "parallelism": "1 GPU" if world == 1 else "Hvvvv/Hvvvo rows sharded x%d + allreduce(r1,r2)" % world,
                      "l2": "inputs larger than L2 (Hvvvv rows %.2f GB per rank)" % (w.engine.vl * v**3 * 8 / 1e9),
                      "seed": args.seed, "eri_scale": round(scale, 6), "diis": 8},
           "flops_issued": fi, "tflops_issued": round(fi / (ms_step * 1e-3) / 1e12, 3),
           "tflops_issued_frac_of_peak": round(fi / (ms_step * 1e-3) / 1e12 / (peak * world), 4),
           "hbar": {"build_ms": round(hbar_ms, 3), "flops_issued": hfi,
                    "tflops_issued": round(hfi / (hbar_ms * 1e-3) / 1e12, 3)},
           "pseudoenergy": last[0] if last else None, "rms": last[1] if last else None,
           "gpu_launches": int(launches), "clocks": clk, "roofline": roofline, "e2e": e2e,
           "cpu_baseline": cpu_baseline, "parity": parity}
    emit(out)
    if world > 1:
        dist.destroy_process_group()


def main():
    global _OUT_FD
    args = parse_args()
    # stdout carries exactly one JSON line: whatever libraries print there (NCCL's version banner
    # on communicator creation, ...) is sent to stderr instead
    sys.stdout.flush()
    _OUT_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    elif args.path == "lambda":
        run_lambda(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
