"""bench.py — RK3 steps/s of the KMM channel-flow time-stepper on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--config re180|re550|re1000|re2000] [--precision 8|4]

One "step" is one full RK3 step (three substeps) of the named grid from a synthetic, self-generated field (seeded
smooth random modes on a laminar mean profile; run time does not depend on the data for a given dt).  Prints ONE JSON
line.  Workload: BASELINE.json configs[2], Re_tau=1000, 1024x385x768 physical (NX x NY x NZ = 1024x385x385), fp64
state — the largest configuration that fits one GPU and the one BASELINE.json lists for 1/2/4/8 GPUs.

value      whole-job RK3 steps/s, state resident in HBM, timed on the device with CUDA events on the library's stream
           (chan_get_stage_ms stage 4), max over ranks.
e2e        the same through the public call a user makes (pinned host slabs -> H2D -> rk_init -> K steps, each followed
           by a D2H read of the step scalars -> D2H of the slabs), wall clock, max over ranks.
roofline   the kernel with the largest share of the step against the measured HBM copy peak (MEASURED_PEAKS.json):
           algorithmic bytes of one substep / device time of that kernel's launches in one substep.
stages     per-substep device time of every kernel, algorithmic GB/s, and the residual (step minus its stages).
cpu_baseline  oracle/kmm.py (numpy port) on a bounded sample of the same workload (rank 0, N=1 only).
parity_vs_golden  N>1 only: before timing, every rank steps the 64x21x9 fp64 golden case (tests/golden, dumped from the
           reference build) with the kernels the size heuristics pick and reports its worst error / tolerance.
--impl reference  times oracle/_ref/ref_<grid>_f64, the reference's own CUDA sources built by oracle/build_ref.sh (the
           reference has no CPU path), one rank on one GPU whatever --gpus says.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    "re180": dict(nx=128, ny=97, nz=65, name="Re_tau=180 channel, 128x97x128"),
    "re550": dict(nx=384, ny=193, nz=193, name="Re_tau=550 channel, 384x193x384"),
    "re1000": dict(nx=1024, ny=385, nz=385, name="Re_tau=1000 channel, 1024x385x768"),
    "re2000": dict(nx=2048, ny=633, nz=769, name="Re_tau=2000 channel, 2048x633x1536"),
}
# kernels that only return early in -DCHAN_ABLATE builds: a run with one of them set measures nothing
ABLATION_KNOBS = ("CHAN_TUNE_Z_STOP", "CHAN_TUNE_X_SKIP", "CHAN_TUNE_POST_STOP")


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def sizes(cfg, prec):
    nx, ny, nz = cfg["nx"], cfg["ny"], cfg["nz"]
    Kx, Kz = 2 * (nx // 3) + 1, (2 * nz - 2) // 3 + 1
    c = 2 * prec
    n_act = Kx * Kz * ny
    return dict(Kx=Kx, Kz=Kz, n_act=n_act, P=n_act * c, Qc=nx * Kz * ny * c, mpts=nx * ny * (2 * nz - 2) / 1e6)


def mesh(ny):
    """y of the reference's grid (channel.h:21,50-51): float product j*DELTA_Y, double tanh."""
    dy = np.float32(2.0) / np.float32(ny - 1)
    x = (np.arange(ny, dtype=np.float32) * dy).astype(np.float64) - 1.0
    return np.tanh(2.0 * x) / np.tanh(2.0)


def synthetic_fields(nx, ny, nz, cplx, amp=0.12, seed=20261106, rows=None):
    """Smooth random (phi, g) on the active modes + laminar mean (simplified).  `rows`: the reference-layout
    rows i to generate (a rank's slab); every row has its own seed, so a slab is the same whoever generates it."""
    y = mesh(ny)
    n2 = nx * (2 * nz - 2)
    rows = np.arange(nx) if rows is None else np.asarray(rows)
    kx = np.where(rows < nx // 2, rows, rows - nx).astype(np.float64); kz = np.arange(nz, dtype=np.float64)
    a = np.exp(-(kx[:, None] / (nx / 9.0)) ** 2 - (kz[None, :] / ((2 * nz - 2) / 9.0)) ** 2)
    a[(np.abs(kx)[:, None] > nx // 3) | (kz[None, :] > (2 * nz - 2) // 3)] = 0
    shapes = np.stack([1 - y ** 2, y * (1 - y ** 2), (1 - y ** 2) ** 2])

    def field(scale, salt):
        out = np.empty((len(rows), nz, ny), cplx)
        for r, i in enumerate(rows):
            rng = np.random.default_rng([seed, salt, int(i)])
            xi = rng.standard_normal((nz, 3)) + 1j * rng.standard_normal((nz, 3))
            out[r] = (xi @ shapes) * a[r, :, None] * (scale * n2)
        return out

    return field(amp * 6.0, 1), field(amp * 2.0, 2), (1.35 * (1 - y ** 2) * n2).astype(cplx)


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML from a thread every 10 ms
    (the timed region is a fraction of a second, `nvidia-smi -lms` does not even start that fast);
    nvidia-smi one-shot queries are the fall-back when NVML cannot be loaded."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu=0):
        self.gpu = gpu; self.sm = []; self.mx = None; self.reasons = set(); self.run = False; self.t = None; self.how = None

    def _nvml_loop(self):
        import pynvml as nv
        bits = {}
        for name, attr in (("hw_slowdown", "nvmlClocksEventReasonHwSlowdown"), ("hw_thermal_slowdown", "nvmlClocksEventReasonHwThermalSlowdown"),
                           ("sw_thermal_slowdown", "nvmlClocksEventReasonSwThermalSlowdown"), ("sw_power_cap", "nvmlClocksEventReasonSwPowerCap")):
            v = getattr(nv, attr, None) or getattr(nv, attr.replace("ClocksEventReason", "ClocksThrottleReason"), None)
            if v is not None:
                bits[name] = v
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while self.run:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.hd, nv.NVML_CLOCK_SM)))
                r = get_reasons(self.hd)
                for name, bit in bits.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.01)

    def _smi_loop(self):
        while self.run:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                self.sm.append(float(f[0])); self.mx = float(f[1])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                    if v.lower().startswith("active"):
                        self.reasons.add(name)
            except Exception:
                pass

    def start(self):
        self.run = True
        try:
            import pynvml as nv
            nv.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            idx = self.gpu
            if vis and all(v.strip().isdigit() for v in vis.split(",")):
                idx = int(vis.split(",")[self.gpu])
            self.hd = nv.nvmlDeviceGetHandleByIndex(idx)
            self.mx = float(nv.nvmlDeviceGetMaxClockInfo(self.hd, nv.NVML_CLOCK_SM))
            self.how = "nvml, 10 ms"
            self.t = threading.Thread(target=self._nvml_loop, daemon=True)
        except Exception:
            self.how = "nvidia-smi one-shot queries"
            self.t = threading.Thread(target=self._smi_loop, daemon=True)
        self.t.start()

    def stop(self):
        self.run = False
        if self.t:
            self.t.join(timeout=6)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.mx, "reasons": ["no clock sample"], "samples": 0, "how": self.how}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": self.mx, "reasons": sorted(self.reasons), "samples": len(self.sm), "how": self.how}


def cpu_baseline_sample(cfg, prec):
    """oracle/kmm.py on a bounded sample of the workload: the pencil work of one substep on a
    1/16 kx slab and the plane work on 8 y-planes, each scaled to the full grid."""
    from oracle import kmm
    nx, ny, nz = cfg["nx"], cfg["ny"], cfg["nz"]
    real = np.float64 if prec == 8 else np.float32
    nxl = max(8, nx // 16); nyl = 8
    Gs = kmm.Grid(nx, ny, nz, real, i_range=(0, nxl)); Gf = kmm.Grid(nx, ny, nz, real)
    rng = np.random.default_rng(1)
    mk = lambda s: (rng.standard_normal(s) + 1j * rng.standard_normal(s)).astype(Gs.cplx)
    f = [mk((nxl, nz, ny)) for _ in range(6)]
    dt = 1e-3; n = 1
    t0 = time.perf_counter()
    pw = kmm.rk_step_1(Gs, f[0], kmm.deriv_yy(Gs, f[0]), f[2], dt, n)
    gw = kmm.rk_step_1(Gs, f[1], kmm.deriv_yy(Gs, f[1]), f[3], dt, n)
    u, w = kmm.calc_uw(Gs, f[4], f[1])
    wz = kmm.deriv_y(Gs, u); wx = kmm.deriv_y(Gs, w)
    kmm.calc_omega(Gs, wx, wz, u, f[5], w)
    hv, hg = kmm.calc_hvg(Gs, f[2], f[3], f[4])
    pw = kmm.rk_step_2(Gs, pw, hv, dt, n); gw = kmm.rk_step_2(Gs, gw, hg, dt, n)
    kmm.implicit(Gs, gw, kmm.BETHA[n], dt)
    kmm.bilaplacian(Gs, pw, kmm.BETHA[n], dt)
    t_pen = (time.perf_counter() - t0) * nx / nxl
    p = [mk((nx, nz, nyl)) for _ in range(6)]
    t0 = time.perf_counter()
    ph = [kmm.to_physical(Gf, a) for a in p]
    m = kmm.rotor(Gf, *ph)
    for a in m:
        kmm.dealias(Gf, kmm.to_spectral(Gf, a))
    t_pl = (time.perf_counter() - t0) * ny / nyl
    step = 3.0 * (t_pen + t_pl)
    return {"value": 1.0 / step, "unit": "steps/s", "cores": 1, "kind": "port",
            "sample": "oracle/kmm.py (numpy): pencil work of one substep on a %d/%d kx slab + plane work on %d/%d planes, scaled to the grid, x3 substeps" % (nxl, nx, nyl, ny)}


def run_reference(args, cfg):
    """The reference's own sources (oracle/_ref, built by oracle/build_ref.sh) on the same grid: one rank on one GPU."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    prec = 64 if args.precision == 8 else 32
    b = os.path.join(ROOT, "oracle", "_ref", "ref_%dx%dx%d_f%d" % (cfg["nx"], cfg["ny"], cfg["nz"], prec))
    sz = sizes(cfg, args.precision)
    # a bounded sample: the reference needs seconds per step at the larger grids (host-staged transposes); cap the
    # steps so that the whole run ends within a few minutes
    est = {"re180": 0.1, "re550": 1.6, "re1000": 18.0, "re2000": 130.0}.get(args.config, 1.0)
    steps = max(1, min(args.steps, int(120.0 / est))); warm = max(1, min(args.warmup, 2 if est > 5 else args.warmup))
    base = {"impl": "reference", "metric": "RK3 steps/s", "unit": "steps/s", "n_gpus": 1, "steps": steps, "warmup": warm,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64" if args.precision == 8 else "f32",
            "data": "the reference's own initial field (io.c:76-94)",
            "config": {"workload": cfg["name"] + (", fp64 state" if args.precision == 8 else ", fp32 state") + ", 1 GPU", "grid": [cfg["nx"], cfg["ny"], cfg["nz"]],
                       "ranks": 1, "requested_gpus": args.gpus,
                       "note": "reference CUDA build (cuFFT/cuSPARSE/cuBLAS, host-staged transposes), ONE rank on one GPU whatever --gpus says: its slab "
                               "code needs NY%%size==0 (channel.h:72-73; every BASELINE NY is odd) and no MPI exists on the box, so at N>1 this "
                               "arm is a one-GPU number; steps timed: %d of the %d asked for (bounded sample)" % (steps, args.steps)}}
    if not os.path.exists(b):
        base.update({"unavailable": "reference binary %s not built (oracle/build_ref.sh needs /root/reference)" % os.path.basename(b)})
        print(json.dumps(base)); return
    try:
        r = subprocess.run([b, "time", str(steps), str(warm)], capture_output=True, text=True, timeout=1500)
        line = [l for l in r.stderr.splitlines() if l.startswith("REF_TIME")]
        err = (r.stderr[-200:] or r.stdout[-200:]).replace("\n", " ")
    except subprocess.TimeoutExpired:
        line, err = [], "timed out after 1500 s"
    if not line:
        base.update({"unavailable": "reference run failed: " + err})
        print(json.dumps(base)); return
    ms = float(line[0].split("ms_per_step=")[1])
    v = 1e3 / ms
    base.update({"value": v, "ms_per_step": ms, "mpts_step_per_s": v * sz["mpts"],
                 "cpu_baseline": {"value": v, "unit": "steps/s", "cores": 1, "kind": "reference",
                                  "sample": "%d steps after %d warm-up steps of the full grid; the reference has no CPU implementation, this is its CUDA build driven by one host thread" % (steps, warm)},
                 "e2e": {"value": v, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    print(json.dumps(base))


def golden_parity(cg, rank, world, local, comm):
    """Every rank steps its slab of the 64x21x9 fp64 golden case (two steps, dumps of the reference build in tests/golden)
    with the kernels the library picks at this rank count; returns worst (error / tolerance) over the six fields."""
    from tests.util import Golden
    g = Golden("64x21x9_f64")
    ddv, gg, um = g.inputs()
    ch = cg.Channel(g.nx, g.ny, g.nz, precision=8, device=local, rank=rank, size=world, nccl_comm=comm)
    if world > 1 and os.environ.get("CHAN_PEER", "1") != "0":
        ch.enable_peer_exchange()
    ch.RKstep(ddv, gg, 2, fixed_dt=g.dt, umean=um)
    _, _, ii = ch.slab()
    worst = 0.0
    for name in ("ddv", "g", "v", "dv", "R_ddv", "R_g"):
        full = g.full("steps2/" + name)
        got = ch.get_field(name)[ii]
        e = float(np.linalg.norm((got - full[ii]).ravel()) / np.linalg.norm(full.ravel()))
        worst = max(worst, e / 1e-11)
    ch.close()
    return worst


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--config", default="re1000")
    ap.add_argument("--precision", type=int, default=8)
    ap.add_argument("--plane-batch", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile", action="store_true", help="device-timed leg only (for ncu runs and sweeps)")
    ap.add_argument("--grid", default="", help="NX,NY,NZ override for kernel experiments (profile runs only)")
    ap.add_argument("--allow-knobs", action="store_true", help="accept the timing-ablation knobs (needs a -DCHAN_ABLATE build)")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    if args.grid:
        gx, gy, gz = (int(v) for v in args.grid.split(","))
        cfg = dict(nx=gx, ny=gy, nz=gz, name="experiment %dx%dx%d" % (gx, gy, 2 * gz - 2))
        args.profile = True
    if args.impl == "reference":
        return run_reference(args, cfg)
    knobs = {k: v for k, v in os.environ.items() if k.startswith("CHAN_TUNE_") or k == "CHAN_PEER"}
    bad = [k for k in ABLATION_KNOBS if knobs.get(k, "0") not in ("0", "")]
    if bad and not args.allow_knobs:
        print(json.dumps({"error": "refusing to run: %s set (kernels would return early and the number would mean nothing); "
                                   "ablation runs need a -DCHAN_ABLATE build and --allow-knobs" % ",".join(bad)}))
        sys.exit(2)

    import torch
    import channel_gpu_b200 as cg
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    comm = None
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        from channel_gpu_b200 import nccl_boot
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        comm = nccl_boot.init_comm(rank, world, local)
    W = max(3, args.warmup); K = args.steps
    nx, ny, nz = cfg["nx"], cfg["ny"], cfg["nz"]
    prec = args.precision
    cplx = np.complex128 if prec == 8 else np.complex64
    sz = sizes(cfg, prec)

    parity = None
    if world > 1 and not args.profile:
        p = torch.tensor([golden_parity(cg, rank, world, local, comm)], device="cuda", dtype=torch.float64)
        dist.all_reduce(p, op=dist.ReduceOp.MAX)
        parity = float(p[0])

    ch = cg.Channel(nx, ny, nz, precision=prec, device=local, plane_batch=args.plane_batch, rank=rank, size=world, nccl_comm=comm)
    if world > 1 and os.environ.get("CHAN_PEER", "1") != "0":
        ch.enable_peer_exchange()
    # every rank generates and pins only the rows of its own kx slab (at Re_tau=2000 the dense array is 15 GB per field)
    _, kxl, rows = ch.slab()
    ddv, g, um = synthetic_fields(nx, ny, nz, cplx, rows=rows)
    slab_bytes = ddv.nbytes
    def pinned(a):
        t = torch.from_numpy(a).pin_memory()
        return t, t.numpy()
    ddv_t, ddv_p = pinned(ddv); g_t, g_p = pinned(g)
    del ddv, g
    ch.set_field_slab("ddv", ddv_p); ch.set_field_slab("g", g_p); ch.set_umean(um)
    ch.rk_init()
    ch.rk_step(W); ch.sync()

    # ---- device-resident timing --------------------------------------------------------
    clocks = ClockSampler(local); clocks.start()
    ch.set_timing(True); ch.stage_ms()
    l0 = ch.launch_count()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    ch.rk_step(K)
    ch.sync()
    wall = time.perf_counter() - t0
    ms, cnt = ch.stage_ms()
    if world > 1:      # max over ranks of the device-timed step and of every stage
        t = torch.tensor(ms + [wall], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = [float(v) for v in t[:-1]]; wall = float(t[-1])
    launches = ch.launch_count() - l0
    ch.set_timing(False)
    if cnt[4] != K:
        raise SystemExit("bench: %d of the %d timed steps were recorded by the stage timers" % (cnt[4], K))
    clk = clocks.stop()
    step_ms = ms[4] / max(cnt[4], 1)
    value = 1e3 / step_ms
    sc = ch.scalars()
    finite = bool(np.isfinite(sc["dt_next"]) and np.isfinite(sc["umax"][0]))

    # ---- per-substep stage times and the roofline of the dominant kernel ------------------------------
    # algorithmic bytes per SUBSTEP (P = one compact field, Qc = one x-padded field); a stage that
    # runs in several plane batches has several launches per substep: times are summed over the launches of a substep
    peak, peak_src = measured_peaks()
    P = sz["P"] / world; Qc = sz["Qc"] / world
    nsub = 3 * max(cnt[4], 1)
    kern = {"pencil_pre": (0, (8 if world == 1 else 9) * P), "pencil_post": (3, 13 * P), "xfft_inv": (5, 6 * P + 6 * Qc),
            "zstage": (6, 9 * Qc), "xfft_fwd": (7, 3 * Qc + 3 * P)}
    stages = {}
    for name, (sid, nbytes) in kern.items():
        t = ms[sid] / nsub
        stages[name] = {"ms_per_substep": t, "launches_per_substep": cnt[sid] / nsub, "alg_GB_per_substep": nbytes / 1e9,
                        "GBps": nbytes / 1e6 / t if t > 0 else None, "frac_of_hbm_peak": nbytes / 1e6 / t / peak if t > 0 else None}
    if world > 1:
        stages["peer_copy_out"] = {"ms_per_substep": ms[8] / nsub, "note": "pencil -> plane: 6 fields, bulk NVLink copies (between pencil_pre and the plane stage)"}
        stages["peer_copy_back"] = {"ms_per_substep": ms[9] / nsub, "note": "plane -> pencil: 3 fields (inside plane_stage_total)"}
    stages["scalar"] = {"ms_per_substep": ms[2] / nsub, "note": "side stream, beside pencil_post"}
    stages["plane_stage_total"] = {"ms_per_substep": ms[1] / nsub, "note": "xfft_inv + zstage + xfft_fwd + (N>1) the exchange barriers"}
    stages["residual_ms_per_step"] = step_ms - 3 * (ms[0] + ms[1] + ms[3] + ms[8]) / nsub      # launch gaps, side-stream joins, the end-of-step kernel
    if args.profile:
        if rank == 0:
            print(json.dumps({"profile_run": True, "n_gpus": world, "ms_per_step": step_ms, "stage_ms": ms, "stage_cnt": cnt, "launches": launches, "stages": stages, "knobs": knobs}))
        if world > 1:
            ch.close(); nccl_boot.destroy_comm(comm); dist.destroy_process_group()
        return
    dom = max(kern, key=lambda k: stages[k]["ms_per_substep"])
    traffic = None
    pj = os.path.join(ROOT, "profiles", "ncu_summary.json")
    if os.path.exists(pj) and world == 1:
        js = json.load(open(pj)).get(args.config, {}).get(dom, {})
        traffic = js.get("dram_bytes_per_launch")
        if traffic and js.get("planes_in_launch"):         # captured on a full plane batch: scale to the average launch of a substep
            traffic = int(traffic * (ny / max(stages[dom]["launches_per_substep"], 1e-9)) / js["planes_in_launch"])
    lps = max(stages[dom]["launches_per_substep"], 1e-9)
    ach = stages[dom]["GBps"]
    roofline = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                "peak_source": peak_src, "alg_bytes_per_launch": kern[dom][1] / lps, "launches_per_substep": lps,
                "ms_per_launch": stages[dom]["ms_per_substep"] / lps,
                "note": "algorithmic bytes of one substep / CUDA-event time of the kernel's launches in one substep (timed steps, the library's own stream); "
                        "traffic = dram bytes of one ncu --set full launch (profiles/ncu_summary.json), null when not captured for this config",
                "step_frac": (3 * (31 * sz["P"] + 18 * sz["Qc"]) / world / 1e6 / step_ms) / peak}

    # ---- end to end through the public API with host buffers ----------------------------
    big = slab_bytes > (1 << 30)
    out_ddv = ddv_p if big else torch.empty_like(ddv_t).pin_memory().numpy()      # large slabs: read back into the same pinned buffers
    out_g = g_p if big else torch.empty_like(g_t).pin_memory().numpy()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    ch.set_field_slab("ddv", ddv_p); ch.set_field_slab("g", g_p); ch.set_umean(um)
    t1 = time.perf_counter()
    ch.rk_init()
    d2h_scalars = 0
    for _ in range(K):
        ch.rk_step(1)
        ch.scalars(); d2h_scalars += 104
    t2 = time.perf_counter()
    ch.get_field_slab("ddv", out=out_ddv); ch.get_field_slab("g", out=out_g)
    e2e_s = time.perf_counter() - t0
    e2e_parts = {"set_fields_ms": 1e3 * (t1 - t0), "init_and_steps_ms": 1e3 * (t2 - t1), "get_fields_ms": 1e3 * (t0 + e2e_s - t2)}
    if world > 1:
        t = torch.tensor([e2e_s] + list(e2e_parts.values()), device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t[0]); e2e_parts = dict(zip(e2e_parts, [float(v) for v in t[1:]]))
    e2e = {"value": K / e2e_s, "unit": "steps/s", "h2d_bytes_per_step": int(world * 2 * slab_bytes / K), "d2h_bytes_per_step": int(world * (2 * slab_bytes + d2h_scalars) / K),
           "note": "per rank: pinned slab H2D + relayout (set_field_slab x2) + rk_init + K x (rk_step + D2H of the step scalars) + relayout + pinned D2H (get_field_slab x2); "
                   "field copies amortised over the K steps; bytes are totals over the ranks", "parts_max_over_ranks": e2e_parts}

    out = {"metric": "RK3 steps/s", "value": value, "unit": "steps/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": step_ms,
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64" if prec == 8 else "f32",
           "data": "synthetic (seeded smooth random modes on a laminar mean profile, free-running CFL dt)",
           "mpts_step_per_s": value * sz["mpts"],
           "config": {"workload": cfg["name"] + (", fp64 state" if prec == 8 else ", fp32 state") + ", %d GPU" % world,
                      "grid": [nx, ny, nz], "active_modes": sz["n_act"], "plane_batch_arg": args.plane_batch, "knobs": knobs,
                      "l2": "working set per step (>= 1 GB per GPU) exceeds the 126 MB L2; no flush needed" if sz["P"] / world > 2e8 else
                            "working set per GPU is %.0f MB per field: L2-resident pieces are part of the workload at this size" % (sz["P"] / world / 1e6),
                      "finite": finite},
           "clocks": clk, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "stages": stages,
           "wall_ms_per_step": 1e3 * wall / K}
    if world > 1:
        S = world
        xbytes = 9 * (P) * (S - 1) / S                      # per GPU per substep, each way (only active modes cross the wire)
        xtime = stages["peer_copy_out"]["ms_per_substep"] + stages["peer_copy_back"]["ms_per_substep"]
        if xtime <= 0:                                      # in-kernel stores (CHAN_TUNE_PRE_DIRECT / FWD_DIRECT): charge the producing kernels
            xtime = stages["pencil_pre"]["ms_per_substep"] + stages["xfft_fwd"]["ms_per_substep"]
        out["nvlink"] = {"bytes_per_gpu_per_step_each_way": int(3 * xbytes),
                         "achieved_GBps": xbytes / 1e6 / xtime if xtime > 0 else None, "peak_GBps": 900.0,
                         "frac": xbytes / 1e6 / xtime / 900.0 if xtime > 0 else None,
                         "note": "bytes leaving a GPU per substep / device time of the two bulk peer-copy kernels (CUDA events); the copies include the 1/N local part, which does not cross the wire and is not counted",
                         "exchange": "NCCL grouped send/recv" if os.environ.get("CHAN_PEER", "1") == "0" else "bulk peer copies over NVLink (CUDA IPC mappings, destinations rotated by rank) + 2 stream-ordered barriers per substep"}
        out["parity_vs_golden"] = {"worst_error_over_tolerance": parity, "case": "64x21x9 fp64, 2 steps vs the reference build's dump (tests/golden), tolerance 1e-11, max over ranks and fields",
                                   "ok": bool(parity is not None and parity < 1.0)}
        out["reference_arm_note"] = "--impl reference runs ONE rank on one GPU at every N (see its config.note): the N>1 ratio is N GPUs against one"
    if rank == 0:
        if not args.no_cpu_baseline and world == 1:
            out["cpu_baseline"] = cpu_baseline_sample(cfg, prec)
        print(json.dumps(out))
    if world > 1:
        ch.close()
        nccl_boot.destroy_comm(comm)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
