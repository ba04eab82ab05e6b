#!/usr/bin/env python
"""Benchmark of the Melissa VB hot path (one JSON line on stdout, rank 0).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's algorithm on host cores

A "step" is one VB iteration (reference R/melissa_vb.R:309-517): every covered
(cell, region) pair is updated once.  Metric: cell*region updates per second =
pairs * steps / time.  Workload: BASELINE.json configs[3], the configuration the
metric is quoted on -- synthetic 20,000 cells x 5,000 regions, K = 10, RBF M = 4,
~25 CpGs per pair -- cell-sharded over N GPUs (strong scaling: the job is the same
at every N).  `--workload c3` selects configs[2] (5,000 x 1,000, K = 4).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "c4": dict(name="synthetic 20,000 cells x 5,000 regions, K=10, RBF M=4, ~25 CpGs/pair (BASELINE configs[3])",
               N=20000, M=5000, K=10, n_basis=4, seed=2),
    "c3": dict(name="synthetic 5,000 cells x 1,000 regions, K=4, RBF M=4, ~25 CpGs/pair (BASELINE configs[2])",
               N=5000, M=1000, K=4, n_basis=4, seed=1),
    "c5": dict(name="synthetic 100,000 cells x 10,000 regions, K=20, RBF M=4, ~25 CpGs/pair (BASELINE configs[4]; needs 8 GPUs)",
               N=100000, M=10000, K=20, n_basis=4, seed=3),
    "tiny": dict(name="synthetic 500 cells x 200 regions, K=3 (development only)", N=500, M=200, K=3, n_basis=4, seed=3),
}
METRIC = "VB cell-region updates/s"
UNIT = "cell*region updates/s"
LB_EVERY = 10
CHUNK = 25


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=2)
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v == "Active"})
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        pw = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        return dict(sm_mhz=(sm[len(sm) // 2] if sm else None), sm_max_mhz=(max(mx) if mx else None),
                    power_w_max=(max(pw) if pw else None), samples=len(self.rows), reasons=reasons)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def algorithmic_bytes(nnz, P, D):
    """9*I + 16*D + 8 bytes per cell*region update."""
    return 9 * nnz + P * (16 * D + 8)


def host_cores():
    """Cores this process may use.  torchrun exports OMP_NUM_THREADS=1; the CPU arm must not inherit that."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_port_run(sample, K, basis, w0, steps, warmup, threads=0):
    """The plain-C restatement of the reference loop (oracle/), timed on host cores."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import c_oracle
    threads = threads or host_cores()
    delta_0 = np.full(K, 0.51)
    if warmup > 0:
        c_oracle.vb_run(sample, K, basis, w0, delta_0, vb_max_iter=warmup + 1, n_threads=threads)
    o = c_oracle.vb_run(sample, K, basis, w0, delta_0, vb_max_iter=steps + 1, epsilon_conv=0.0, n_threads=threads)
    P = int(sample["pair_cell"].shape[0])
    iters = o["n_iter"] - 1
    return dict(value=P * iters / o["loop_seconds"], seconds=o["loop_seconds"], iters=iters, cores=o["threads"], P=P)


def cpu_sample(wl, n_cells):
    """The first n_cells cells of the workload's recipe as host arrays + initial w (drawn on the GPU when there is one:
    the NumPy generator needs minutes for a sample of this size)."""
    from melissa_b200 import synth
    try:
        import torch
        if torch.cuda.is_available():
            d, tr = synth.generate_device(n_cells, wl["M"], wl["K"], n_basis=wl["n_basis"], seed=wl["seed"], N_total=n_cells,
                                          chunk_cells=CHUNK)
            return host_sample(d, n_cells), tr
    except Exception:
        pass
    d, tr = synth.generate(n_cells, wl["M"], wl["K"], n_basis=wl["n_basis"], seed=wl["seed"], N_total=n_cells)
    return d.as_dict(), tr


def default_sample_cells(wl):
    """~5e6 pairs: 13 iterations of the C restatement take 10-30 s on 16-32 cores, its design matrices ~1 GB."""
    return max(40, min(wl["N"], int(5.0e6 / wl["M"])))


def host_sample(data_t, n_cells):
    """First n_cells cells of a device data set as host NumPy arrays."""
    import numpy as np
    import torch
    pc = data_t["pair_cell"]
    p1 = int(torch.searchsorted(pc, torch.tensor([n_cells], device=pc.device, dtype=pc.dtype))[0])
    c1 = int(data_t["pair_ptr"][p1])
    return dict(N=n_cells, M=data_t["M"], pair_ptr=data_t["pair_ptr"][:p1 + 1].cpu().numpy(),
                pair_cell=pc[:p1].cpu().numpy(), pair_region=data_t["pair_region"][:p1].cpu().numpy(),
                x=data_t["x"][:c1].cpu().numpy(), y=data_t["y"][:c1].cpu().numpy())


def run_single_process(args, wl, config):
    """`--gpus N` without torchrun: ONE process drives all N GPUs through the C ABI (melissa_multi_create /
    melissa_vb_run_multi: a host thread and a stream per GPU inside the library, partial sums added over NVLink peer
    loads).  This is the path an R `.Call` binds; the torchrun arm (one process per GPU, NCCL) stays for the driver."""
    import numpy as np
    import torch
    import melissa_b200 as mb
    from melissa_b200 import synth
    G, steps, warmup = args.gpus, args.steps, max(args.warmup, 0)
    K, D = wl["K"], wl["n_basis"] + 1
    if torch.cuda.device_count() < G:
        raise SystemExit("need %d GPUs, found %d" % (G, torch.cuda.device_count()))
    per = -(-wl["N"] // G)
    per = -(-per // CHUNK) * CHUNK
    engines, host_parts, tr, c_true = [], [], None, []
    P_tot = nnz_tot = 0
    t_gen = t_pack = 0.0
    keep_host = not args.no_e2e
    for g in range(G):
        lo, hi = min(g * per, wl["N"]), min((g + 1) * per, wl["N"])
        torch.cuda.set_device(g)
        t0 = time.time()
        data_t, tr = synth.generate_device(wl["N"], wl["M"], K, n_basis=wl["n_basis"], seed=wl["seed"], cell_lo=lo,
                                           cell_hi=hi, N_total=wl["N"], device=g, chunk_cells=CHUNK)
        torch.cuda.synchronize(g)
        t_gen += time.time() - t0
        c_true.append(np.asarray(tr["C_true"]))
        P_tot += int(data_t["pair_cell"].shape[0])
        nnz_tot += int(data_t["x"].shape[0])
        t0 = time.time()
        engines.append(mb.Engine(data_t, tr["basis"], device=g, allow_empty_regions=True))
        t_pack += time.time() - t0
        if keep_host:
            host_parts.append({k: data_t[k].cpu() for k in ("pair_ptr", "pair_cell", "pair_region", "x", "y")})
        del data_t
        torch.cuda.empty_cache()
    delta_0 = np.full(K, 0.51)
    me = mb.MultiEngine(engines)
    if warmup:
        me.run(K, tr["w0"], delta_0, vb_max_iter=warmup + 1, epsilon_conv=0.0, lb_every=LB_EVERY)
    sampler = ClockSampler(0)
    sampler.start()
    time.sleep(0.15)
    launches0 = me.launch_count()
    engines[0].profile(True)
    # one bound in `steps` iterations, like the stepping arm (the run always takes a bound at vb_max_iter, so a second
    # one at a multiple of 10 is switched off when the run is that short)
    lb_every_run = LB_EVERY if steps > LB_EVERY else steps + 1
    out = me.run(K, tr["w0"], delta_0, vb_max_iter=steps + 1, epsilon_conv=0.0, lb_every=lb_every_run)
    ms = me.last_loop_ms()
    launches = me.launch_count() - launches0
    prof = engines[0].profile_read()
    engines[0].profile(False)
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < 1.0:                      # keep the GPUs busy so the clock sampler sees the loaded state
        me.run(K, tr["w0"], delta_0, vb_max_iter=steps + 1, epsilon_conv=0.0, lb_every=LB_EVERY)
    clocks = sampler.finish()
    info0 = engines[0].layout_info()
    sweep_ms = prof["sweep"][0] / max(1, prof["sweep"][1])
    peak, peak_src = measured_peak()
    alg0 = algorithmic_bytes(info0["cpgs"], info0["pairs"], D)
    value = P_tot * steps / (ms * 1e-3)
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=G, steps=steps, warmup=warmup, ms_per_step=ms / steps,
                higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f64", data="synthetic",
                config=dict(config, sharding="cells over %d GPU(s), one process (melissa_vb_run_multi)" % G))
    line["roofline"] = dict(bound="hbm", kernel="sweep3_kernel", achieved=alg0 / (sweep_ms * 1e-3) / 1e9, peak=peak, unit="GB/s",
                            frac=alg0 / (sweep_ms * 1e-3) / 1e9 / peak, traffic=None, peak_source=peak_src,
                            algorithmic_bytes_per_launch=alg0, kernel_ms=sweep_ms, note="GPU 0's shard")
    line["kernels_ms_per_step"] = {k: v[0] / steps for k, v in prof.items()}
    line["gpu_launches"] = launches
    line["clocks"] = clocks
    line["pairs"], line["cpgs"] = P_tot, nnz_tot
    line["setup_s"] = dict(generate=round(t_gen, 2), pack=round(t_pack, 2))
    line["check"] = dict(ari_vs_truth=mb.cluster_ari(np.concatenate(c_true), out["r_nk"]), last_lb=float(out["lb"][-1]))
    t0 = time.perf_counter()
    conv = me.run(K, tr["w0"], delta_0, vb_max_iter=300, epsilon_conv=1e-5, lb_every=LB_EVERY)
    line["time_to_converge"] = dict(seconds=time.perf_counter() - t0, iterations=int(conv["n_iter"]) - 1, flags=int(conv["flags"]),
                                    lb=float(conv["lb"][-1]), rule="abs(LB - LB_prev) < 1e-5 every 10th iteration, vb_max_iter = 300")
    me.close()
    for e in engines:
        e.close()
    torch.cuda.empty_cache()
    e2e = None
    if keep_host:
        try:
            # the whole job as ONE set of pinned host arrays, the way a caller of melissa_pack_multi holds it
            tot = {k: sum(int(h[k].shape[0]) for h in host_parts) for k in ("pair_cell", "x")}
            hp = torch.empty(tot["pair_cell"] + 1, dtype=torch.int64, pin_memory=True)
            hc = torch.empty(tot["pair_cell"], dtype=torch.int32, pin_memory=True)
            hr = torch.empty(tot["pair_cell"], dtype=torch.int32, pin_memory=True)
            hx = torch.empty(tot["x"], dtype=torch.float64, pin_memory=True)
            hy = torch.empty(tot["x"], dtype=torch.uint8, pin_memory=True)
            p0 = i0 = c0 = 0
            hp[0] = 0
            for g, h in enumerate(host_parts):
                n_p, n_i = int(h["pair_cell"].shape[0]), int(h["x"].shape[0])
                hp[p0 + 1:p0 + n_p + 1] = h["pair_ptr"][1:] + i0
                hc[p0:p0 + n_p] = h["pair_cell"] + c0
                hr[p0:p0 + n_p] = h["pair_region"]
                hx[i0:i0 + n_i] = h["x"]
                hy[i0:i0 + n_i] = h["y"]
                p0, i0, c0 = p0 + n_p, i0 + n_i, c0 + int(h["pair_cell"][-1]) + 1
            host_parts.clear()
            hd = mb.MelissaData(N=wl["N"], M=wl["M"], pair_ptr=hp.numpy(), pair_cell=hc.numpy(), pair_region=hr.numpy(),
                                x=hx.numpy(), y=hy.numpy())
            t0 = time.perf_counter()
            with mb.MultiEngine.pack(hd, tr["basis"], G) as me2:
                o2 = me2.run(K, tr["w0"], delta_0, vb_max_iter=steps + 1, epsilon_conv=0.0, lb_every=LB_EVERY)
                dt = time.perf_counter() - t0              # (the one-GPU arm also stops its clock before the free)
            h2d = hp.numel() * 8 + hc.numel() * 4 + hr.numel() * 4 + hx.numel() * 8 + hy.numel() + tr["w0"].nbytes
            d2h = sum(np.asarray(o2[k]).nbytes for k in ("W", "W_Sigma", "r_nk", "delta", "beta", "pi_k"))
            e2e = dict(value=P_tot * steps / dt, unit=UNIT, h2d_bytes_per_step=h2d / steps, d2h_bytes_per_step=d2h / steps, seconds=dt,
                       what="melissa_pack_multi + melissa_vb_run_multi(%d iterations) + result fetch, pinned host buffers, one process" % steps)
        except Exception as exc:  # reported, never hidden
            e2e = dict(value=None, unit=UNIT, error=repr(exc))
    line["e2e"] = e2e
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-sample-cells", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    steps, warmup = args.steps, max(args.warmup, 0)
    K, D = wl["K"], wl["n_basis"] + 1
    config = dict(workload=wl["name"], cells=wl["N"], regions=wl["M"], K=K, basis="rbf M=%d" % wl["n_basis"],
                  lb_every=LB_EVERY, sharding="cells over %d GPU(s)" % world,
                  l2="inputs larger than L2 (no flush needed)")

    import numpy as np
    if args.impl == "ours" and args.gpus > 1 and world == 1:
        return run_single_process(args, wl, config)
    if args.impl == "reference":
        # The reference is interpreted R (no R on this box): its algorithm as restated in oracle/ (plain C,
        # every host thread), on a bounded sample of the same workload.  Rank 0 only.
        if rank != 0:
            return
        n_s = args.cpu_sample_cells or default_sample_cells(wl)
        d, tr = cpu_sample(wl, n_s)
        r = cpu_port_run(d, K, tr["basis"], tr["w0"], steps, warmup)
        sample = "%d cells x %d regions (%d pairs) of the same recipe, %d iterations" % (n_s, wl["M"], r["P"], r["iters"])
        # the CPU arm runs a bounded SAMPLE of the workload named above (same regions, K, basis; fewer cells): say so
        # where the driver compares configs
        config = dict(config, sample_cells=n_s, sample_pairs=r["P"], sharding="host cores (%d threads)" % r["cores"])
        print(json.dumps(dict(impl="reference", metric=METRIC, value=r["value"], unit=UNIT, n_gpus=args.gpus,
                              steps=steps, warmup=warmup, ms_per_step=1e3 * r["seconds"] / max(1, r["iters"]),
                              higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f64",
                              data="synthetic", config=config,
                              cpu_baseline=dict(value=r["value"], unit=UNIT, cores=r["cores"], kind="port", sample=sample),
                              e2e=dict(value=r["value"], unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))))
        return

    import torch
    import torch.distributed as dist
    import melissa_b200 as mb
    from melissa_b200 import synth
    torch.cuda.set_device(local_rank)
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        import datetime
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"),
                                timeout=datetime.timedelta(seconds=180))
        group = dist.group.WORLD
    # ---- this rank's cell shard, generated on the device
    per = -(-wl["N"] // world)
    per = -(-per // CHUNK) * CHUNK
    lo, hi = min(rank * per, wl["N"]), min((rank + 1) * per, wl["N"])
    t0 = time.time()
    data_t, tr = synth.generate_device(wl["N"], wl["M"], K, n_basis=wl["n_basis"], seed=wl["seed"], cell_lo=lo,
                                       cell_hi=hi, N_total=wl["N"], device=local_rank, chunk_cells=CHUNK)
    torch.cuda.synchronize()
    t_gen = time.time() - t0
    P_loc, nnz_loc = int(data_t["pair_cell"].shape[0]), int(data_t["x"].shape[0])
    t0 = time.time()
    eng = mb.Engine(data_t, tr["basis"], device=local_rank, allow_empty_regions=world > 1, K_hint=K)
    t_pack = time.time() - t0
    delta_0 = np.full(K, 0.51)
    eng.begin(K, tr["w0"], delta_0, n_cells_total=wl["N"])
    tot = torch.tensor([P_loc, nnz_loc], dtype=torch.float64, device="cuda")
    if group is not None:
        dist.all_reduce(tot, group=group)
    P_tot, nnz_tot = int(tot[0]), int(tot[1])
    it = [1]

    def step():
        it[0] += 1
        eng.phase_a()
        if group is not None:
            dist.all_reduce(eng.reduce_tensor(), group=group)
        with_lb = it[0] % LB_EVERY == 0
        eng.phase_b(with_lb)
        if with_lb:
            if group is not None:
                dist.all_reduce(eng.lb_tensor(), group=group)
            return eng.lower_bound()
        return None

    def sync():
        torch.cuda.synchronize()
        if group is not None:
            dist.barrier(group=group)
            torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    sync()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.15)
    launches0 = eng.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync()
    e0.record()
    lbs = [step() for _ in range(steps)]
    e1.record()
    sync()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if group is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
    ms = float(ms[0])
    launches = eng.launch_count() - launches0
    # keep the GPU busy ~1 s longer so the clock sampler sees the loaded state; the count is derived from
    # the rank-agreed time so that every rank issues the same collectives
    for _ in range(int(min(400, max(1, 1000.0 / max(ms / steps, 1e-3))))):
        step()
    torch.cuda.synchronize()
    if sampler:
        clocks = sampler.finish()
    # ---- per-kernel durations (CUDA events on the launching stream), separate pass
    eng.profile(True)
    for _ in range(steps):
        step()
    prof = eng.profile_read()
    eng.profile(False)
    sweep_ms = prof["sweep"][0] / max(1, prof["sweep"][1])
    peak, peak_src = measured_peak()
    alg_bytes = algorithmic_bytes(nnz_loc, P_loc, D)
    achieved = alg_bytes / (sweep_ms * 1e-3) / 1e9
    value = P_tot * steps / (ms * 1e-3)
    res = eng.fetch()
    eng_info = eng.layout_info()
    ari = mb.cluster_ari(tr["C_true"], res["r_nk"])
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=steps, warmup=warmup,
                ms_per_step=ms / steps, higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f64",
                data="synthetic", config=config)
    if rank == 0:
        # dram__bytes_read + write of one sweep launch from the last `ncu --set full` capture of this kernel
        # (profiles/sweep_traffic.json names the capture it came from); null when the capture is of another kernel
        traffic = None
        tp = os.path.join(ROOT, "profiles", "sweep_traffic.json")
        if os.path.exists(tp):
            tj = json.load(open(tp))
            if tj.get("kernel") == "sweep3_kernel":
                traffic = tj.get(args.workload)
            if traffic is not None:
                traffic = traffic / world          # captured on one GPU holding the whole job; a shard moves 1/world of it
        line["roofline"] = dict(bound="hbm", kernel="sweep3_kernel", achieved=achieved, peak=peak, unit="GB/s",
                                frac=achieved / peak, traffic=traffic, peak_source=peak_src,
                                algorithmic_bytes_per_launch=alg_bytes, kernel_ms=sweep_ms,
                                whole_step_frac=(algorithmic_bytes(nnz_tot, P_tot, D) / world) / (ms / steps * 1e-3) / 1e9 / peak)
        # second view of the same kernel: it is FP64-issue bound, not HBM bound.  52 FP64
        # operations per CpG in the row loop (24 basis, 4 dot, 19 truncated-normal mean, 5 accumulate) plus 30 K per
        # pair (5 K mean, 20 K E-step terms, 5 K fixed-point conversion); peak = DFMA rate measured on this pool's
        # B200 with tools/fp64_peak.cu (profiles/fp64_peak.json).
        fp64_peak, fp64_src = 18.1, "round-1 measurement (profiles/fp64_peak.json missing)"
        fpp = os.path.join(ROOT, "profiles", "fp64_peak.json")
        if os.path.exists(fpp):
            fp64_peak, fp64_src = float(json.load(open(fpp))["dfma_T_per_s"]), "profiles/fp64_peak.json (tools/fp64_peak.cu on this pool)"
        fp64_ops = 52.0 * nnz_loc + 30.0 * K * P_loc
        line["fp64"] = dict(achieved=fp64_ops / (sweep_ms * 1e-3) / 1e12, peak=fp64_peak, unit="T FP64 op/s",
                            frac=fp64_ops / (sweep_ms * 1e-3) / 1e12 / fp64_peak, ops_per_cpg=52, ops_per_pair=30 * K,
                            peak_source=fp64_src)
        line["layout"] = eng_info
        line["kernels_ms_per_step"] = {k: v[0] / steps for k, v in prof.items()}
        line["gpu_launches"] = launches
        line["clocks"] = clocks
        line["pairs"], line["cpgs"] = P_tot, nnz_tot
        line["setup_s"] = dict(generate=round(t_gen, 2), pack=round(t_pack, 2))
        line["check"] = dict(ari_vs_truth_rank0=ari, last_lb=[v for v in lbs if v is not None][-1:] or None)
    # ---- time-to-converge under the reference's own stopping rule (|dLB| < 1e-5 tested every 10th iteration,
    #      cap 300: R/melissa_vb.R:309, :458, :499), data resident; every rank runs the same loop
    eng.begin(K, tr["w0"], delta_0, n_cells_total=wl["N"])
    sync()
    t0 = time.perf_counter()
    lb_c, it_c, flags_c = mb.vb_loop(eng, 300, 1e-5, LB_EVERY, group=group)
    sync()
    t_conv = time.perf_counter() - t0
    if rank == 0:
        line["time_to_converge"] = dict(seconds=t_conv, iterations=int(it_c) - 1, flags=int(flags_c), lb=float(lb_c[-1]),
                                        rule="abs(LB - LB_prev) < 1e-5 every 10th iteration, vb_max_iter = 300")
    # ---- end to end through the public API from HOST buffers (pack + H2D + K iterations + D2H of the result)
    e2e = None
    if not args.no_e2e:
        try:
            import psutil
            need = nnz_loc * 9 + P_loc * 16 + 4096
            fits = torch.tensor([1.0 if psutil.virtual_memory().available > 2.5 * need * world else 0.0],
                                dtype=torch.float64, device="cuda")
            if group is not None:                                  # all ranks take the same branch
                dist.all_reduce(fits, op=dist.ReduceOp.MIN, group=group)
            if float(fits[0]) == 0.0:
                raise MemoryError("host RAM too small for a pinned copy of the shards")
            host = {}
            for k in ("pair_ptr", "pair_cell", "pair_region", "x", "y"):
                t = torch.empty(data_t[k].shape, dtype=data_t[k].dtype, pin_memory=True)
                t.copy_(data_t[k])
                host[k] = t
            torch.cuda.synchronize()
            hd = mb.MelissaData(N=hi - lo, M=wl["M"], **{k: v.numpy() for k, v in host.items()})
            eng.close()
            del data_t
            torch.cuda.empty_cache()
            sync()
            t0 = time.perf_counter()
            eng2 = mb.Engine(hd, tr["basis"], device=local_rank, allow_empty_regions=world > 1, K_hint=K)
            if group is None:
                out = eng2.run(K, tr["w0"], delta_0, vb_max_iter=steps + 1, epsilon_conv=0.0, lb_every=LB_EVERY)
            else:
                eng2.begin(K, tr["w0"], delta_0, n_cells_total=wl["N"])
                mb.vb_loop(eng2, steps + 1, 0.0, LB_EVERY, group=group)
                out = eng2.fetch()
            sync()
            dt = time.perf_counter() - t0
            dtt = torch.tensor([dt], dtype=torch.float64, device="cuda")
            if group is not None:
                dist.all_reduce(dtt, op=dist.ReduceOp.MAX, group=group)
            h2d = sum(v.numel() * v.element_size() for v in host.values()) + tr["w0"].nbytes
            d2h = sum(np.asarray(out[k]).nbytes for k in ("W", "W_Sigma", "r_nk", "delta", "beta", "pi_k"))
            e2e = dict(value=P_tot * steps / float(dtt[0]), unit=UNIT, h2d_bytes_per_step=h2d / steps,
                       d2h_bytes_per_step=d2h / steps, seconds=float(dtt[0]),
                       what="melissa_pack + melissa_vb_run(%d iterations) + result fetch, pinned host buffers" % steps)
            eng2.close()
        except Exception as exc:  # reported, never hidden
            e2e = dict(value=None, unit=UNIT, error=repr(exc))
    if rank == 0:
        line["e2e"] = e2e
        if world == 1 and not args.no_cpu:
            n_s = args.cpu_sample_cells or default_sample_cells(wl)
            d, trs = cpu_sample(wl, n_s)
            r = cpu_port_run(d, K, trs["basis"], trs["w0"], 10, 0)
            if "time_to_converge" in line:      # the same number of iterations on the host cores, from the sampled rate
                line["time_to_converge"]["cpu_port_seconds_extrapolated"] = \
                    P_tot * line["time_to_converge"]["iterations"] / r["value"]
            line["cpu_baseline"] = dict(value=r["value"], unit=UNIT, cores=r["cores"], kind="port",
                                        sample="%d cells x %d regions (%d pairs) of the same recipe, %d iterations, "
                                               "loop only" % (n_s, wl["M"], r["P"], r["iters"]))
        print(json.dumps(line))
    if group is not None:
        dist.barrier(group=group)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
