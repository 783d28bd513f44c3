#!/usr/bin/env python
"""tools/bench_regimes.py -- the BASELINE.json configs that are not bench.py's headline line:

  --regime C5   low-latency: IndexIVFFlat nlist=16384 nprobe=32 L2 on synthetic 10Mx128, query batches
                1..64: p50 / p99 latency per search call (host wall clock around
                the C-ABI call with host buffers = what a serving thread sees, and the device-event time),
                QPS, and the list scan's GB/s (distinct list bytes / scan time: no reuse to exploit here).
  --regime C4s  one GPU's shard of C4 (1Bx96 L2, nlist 262144 over 8 GPUs): 125Mx96 with 32768 lists
                (the same list density, 48 GB slab), 10k-query batches, nprobe sweep. The HBM-bound
                regime of the list scan (a few queries per list, slab >> L2).
  --regime sweep  C1 / C2 / C3 at their sizes, stage times per nprobe.

One JSON line per measurement on stdout. Data, centroids and helpers are bench.py's.
"""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

import bench as B  # noqa: E402


def pct(v, p):
    v = sorted(v)
    return v[min(len(v) - 1, int(round(p / 100.0 * (len(v) - 1))))]


def unique_rows(torch, g, xq, nprobe, nlist):
    _, Iq = g.quantize(xq, nprobe)
    torch.cuda.synchronize()  # (the handle may be on its own stream: quantize is asynchronous)
    u = torch.unique(Iq.reshape(-1))
    u = u[u >= 0].cpu().numpy()
    sizes = g.list_sizes() if hasattr(g, "list_sizes") else None
    if sizes is None:
        return int(sum(g.get_list_size(int(l)) for l in u))
    return int(sizes[u].sum())


def build(torch, b2, w, dev, niter=10, chunk_add=8_000_000):
    """index built chunk by chunk so that the row-major copy of the base never sits in HBM whole"""
    d, nlist, metric = w["d"], w["nlist"], w["metric"]
    t0 = time.perf_counter()
    nt = min(w["nb"], 256 * nlist)
    xt, proj = B.gen_synthetic(torch, nt, d, B.SEED_TRAIN, w["normalize"], dev)
    cent = B.train_centroids(torch, xt, nlist, metric, niter=niter)
    del xt
    torch.cuda.empty_cache()  # the slab is allocated by the library, not by torch's caching allocator
    g = b2.IndexIVFFlat(d, nlist, metric, device=dev.index)
    g.set_centroids(cent)
    g.set_stream(torch.cuda.current_stream().cuda_stream)
    t1 = time.perf_counter()
    done = 0
    gen = torch.Generator(device=dev)
    gen.manual_seed(B.SEED_BASE)
    while done < w["nb"]:
        ni = min(chunk_add, w["nb"] - done)
        z = torch.randn((ni, B.D_LATENT), generator=gen, device=dev, dtype=torch.float32)
        x = torch.sin(z @ proj)
        if w["normalize"]:
            x /= x.norm(dim=1, keepdim=True).clamp_min(1e-20)
        ids = torch.arange(done, done + ni, device=dev, dtype=torch.int64)
        g.add_with_ids(x, ids)
        done += ni
        del x, z, ids
        torch.cuda.empty_cache()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    xq, _ = B.gen_synthetic(torch, w["nq"], d, B.SEED_QUERY, w["normalize"], dev, proj=proj)
    return g, xq, cent, {"train_s": round(t1 - t0, 1), "add_s": round(t2 - t1, 1)}


def regime_c5(args, torch, b2, dev):
    w = dict(desc="C5: IVFFlat nlist=16384 nprobe=32 L2, synthetic 10Mx128, batches 1..64", d=128,
             nb=args.nb or 10_000_000, nlist=16384, metric=B.METRIC_L2, nq=4096, k=10,
             normalize=False)
    g, xq, cent, tb = build(torch, b2, w, dev)
    g.nprobe = 32
    # a serving thread has no torch work to be ordered with: the handle's own (non-blocking) stream,
    # on which small-batch searches are replayed as one CUDA graph
    torch.cuda.synchronize()
    g.set_stream(None)
    peak, _ = B.measured_hbm_peak()
    xq_h = torch.empty_like(xq, device="cpu").pin_memory()
    xq_h.copy_(xq)
    xq_np = xq_h.numpy()
    if args.profile_batch:
        # launch list for ncu (--profile-from-start off): three searches of one batch size
        bs = args.profile_batch
        for r in range(6):
            if r == 3:
                torch.cuda.cudart().cudaProfilerStart()
            g.search(xq_np[r * bs:(r + 1) * bs], w["k"])
        torch.cuda.cudart().cudaProfilerStop()
        return
    # the reference on this box's host cores, same index, same batches (BASELINE.md section 4;
    # timing model: tutorial/cpp/benchmark-ivf/benchmark_ivf.cpp:309, one search call per batch)
    cpu = None
    if not args.no_cpu:
        lists = ((l,) + tuple(reversed(g.get_list(l))) for l in range(w["nlist"]))
        idx, kind, po = B.make_cpu_index(w, cent.cpu().numpy(), lists)
        threads = os.cpu_count() or 1
        if kind == "reference":
            po.ref_set_threads(threads)
        cpu = (idx, kind, po, threads)
    for bs in (1, 2, 4, 8, 16, 32, 64):
        D = np.empty((bs, w["k"]), np.float32)
        I = np.empty((bs, w["k"]), np.int64)
        Dd = torch.empty((bs, w["k"]), dtype=torch.float32, device=dev)
        Id = torch.empty((bs, w["k"]), dtype=torch.int64, device=dev)
        reps = args.reps
        wall, devt, scan, coarse = [], [], [], []
        rows = []
        # stage times exist only for eager launches: a short pass with the graph replay off
        g.set_option("use_graph", 0)
        for r in range(40):
            q0 = (r * bs) % (w["nq"] - bs + 1)
            g.search(xq_np[q0:q0 + bs], w["k"], D=D, I=I)
            if r >= 10:
                t = g.last_timings()
                scan.append(t["scan"] * 1e3)
                coarse.append(t["coarse"] * 1e3)
        g.set_option("use_graph", 1)
        for r in range(reps + 20):
            q0 = (r * bs) % (w["nq"] - bs + 1)
            t0 = time.perf_counter()
            g.search(xq_np[q0:q0 + bs], w["k"], D=D, I=I)
            t1 = time.perf_counter()
            if r >= 20:
                wall.append((t1 - t0) * 1e6)
                t = g.last_timings()
                devt.append(t["total"] * 1e3)
        # device-resident calls (no PCIe), same batches
        dwall = []
        for r in range(reps + 20):
            q0 = (r * bs) % (w["nq"] - bs + 1)
            xs = xq[q0:q0 + bs]
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            g.search(xs, w["k"], D=Dd, I=Id)
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            if r >= 20:
                dwall.append((t1 - t0) * 1e6)
        for r in range(0, 8):
            q0 = (r * bs) % (w["nq"] - bs + 1)
            rows.append(unique_rows(torch, g, xq[q0:q0 + bs].contiguous(), 32, w["nlist"]))
        ub = statistics.mean(rows) * w["d"] * 4
        cpu_out = None
        if cpu is not None:
            idx, kind, po, threads = cpu
            cw, same, slots = [], 0, 0
            creps = max(20, min(reps, 200))
            for r in range(creps + 5):
                q0 = (r * bs) % (w["nq"] - bs + 1)
                xs = np.ascontiguousarray(xq_np[q0:q0 + bs])
                t0 = time.perf_counter()
                Dc, Ic = B.cpu_search(idx, kind, po, xs, w["k"], 32, threads)
                t1 = time.perf_counter()
                if r >= 5:
                    cw.append((t1 - t0) * 1e6)
                if r < 8:
                    g.search(xs, w["k"], D=D, I=I)
                    same += int((np.asarray(Ic) == I).sum())
                    slots += I.size
            cpu_out = {"kind": kind, "cores": threads, "reps": creps,
                       "call_us": {"p50": pct(cw, 50), "p99": pct(cw, 99),
                                   "mean": statistics.mean(cw)},
                       "qps": bs / (statistics.mean(cw) * 1e-6),
                       "id_slots_equal_to_gpu": [same, slots]}
        out = {"regime": "C5", "workload": w["desc"], "nb": w["nb"], "batch": bs, "reps": reps,
               "host_call_us": {"p50": pct(wall, 50), "p99": pct(wall, 99), "mean": statistics.mean(wall)},
               "device_resident_call_us": {"p50": pct(dwall, 50), "p99": pct(dwall, 99)},
               "device_event_us": {"p50": pct(devt, 50), "p99": pct(devt, 99)},
               "coarse_us_p50": pct(coarse, 50), "scan_us_p50": pct(scan, 50),
               "qps_host_calls": bs / (statistics.mean(wall) * 1e-6),
               "scan_unique_bytes": ub,
               "scan_gbs": ub / (pct(scan, 50) * 1e-6) / 1e9,
               "scan_frac_of_hbm_peak": ub / (pct(scan, 50) * 1e-6) / 1e9 / peak,
               "centroid_table_bytes": w["nlist"] * w["d"] * 4,
               "cpu_baseline": cpu_out,
               "build": tb}
        print(json.dumps(out), flush=True)


def regime_c4s(args, torch, b2, dev):
    w = dict(desc="C4 shard: IVFFlat L2 125Mx96 (1/8 of 1Bx96), 32768 of the 262144 lists, "
                  "10k-query batch", d=96, nb=args.nb or 125_000_000, nlist=32768,
             metric=B.METRIC_L2, nq=10_000, k=10, normalize=False)
    if args.nb:
        w["nlist"] = max(256, int(args.nb / 3815))
    g, xq, cent, tb = build(torch, b2, w, dev, niter=args.niter)
    peak, _ = B.measured_hbm_peak()
    k = w["k"]
    # ground truth on a query subset against the index's own vectors: exhaustive = nprobe all is too
    # slow at this size, so recall is measured against nprobe = 256 (saturated) and said so
    ngt = 500
    g.nprobe = 256
    _, Igt = g.search(xq[:ngt].contiguous(), k)
    Igt = Igt.cpu().numpy()
    Dd = torch.empty((w["nq"], k), dtype=torch.float32, device=dev)
    Id = torch.empty((w["nq"], k), dtype=torch.int64, device=dev)
    for p, opts in [(int(s), o) for s in args.nprobes.split(",") for o in option_grid(args.options)]:
        g.nprobe = p
        for name, v in opts.items():
            g.set_option(name, v)
        _, I = g.search(xq[:ngt].contiguous(), k)
        rec = B.recall_at_k(I.cpu().numpy(), Igt)
        ur = unique_rows(torch, g, xq, p, w["nlist"])
        for _ in range(3):
            g.search(xq, k, D=Dd, I=Id)
        g.reset_stats()
        st, tot = [], []
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        torch.cuda.profiler.start()  # ncu --profile-from-start off sees only the timed steps
        e0.record()
        for _ in range(args.steps):
            g.search(xq, k, D=Dd, I=Id)
            t = g.last_timings()
            st.append(t)
        e1.record()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        ms = e0.elapsed_time(e1) / args.steps
        scan = statistics.mean(t["scan"] for t in st)
        ndis = g.get_stats().ndis / args.steps
        out = {"regime": "C4s", "workload": w["desc"], "nb": w["nb"], "nlist": w["nlist"],
               "nprobe": p, "options": opts, "recall_at_10_vs_nprobe256": rec, "ms_per_step": ms,
               "qps": w["nq"] / (ms * 1e-3),
               "stage_ms": {s: statistics.mean(t[s] for t in st) for s in st[0]},
               "scan_unique_bytes": ur * w["d"] * 4.0, "ndis": ndis,
               "queries_per_probed_list": ndis / max(ur, 1),
               "scan_gbs": ur * w["d"] * 4.0 / (scan * 1e-3) / 1e9,
               "scan_frac_of_hbm_peak": ur * w["d"] * 4.0 / (scan * 1e-3) / 1e9 / peak,
               "exact_redo_total": g.exact_redo_count(),
               "build": tb}
        print(json.dumps(out), flush=True)


def option_grid(spec):
    """'a=1:2,b=3' -> [{}, ...] every combination ({} alone when spec is empty)"""
    grid = [{}]
    for part in [p for p in spec.split(",") if p]:
        name, vals = part.split("=")
        grid = [dict(g, **{name: float(v)}) for g in grid for v in vals.split(":")]
    return grid


def regime_sweep(args, torch, b2, dev):
    peak, _ = B.measured_hbm_peak()
    for name in args.workloads.split(","):
        w = B.WORKLOADS[name]
        g, xq, cent, tb = build(torch, b2, w, dev)
        k = w["k"]
        Dd = torch.empty((w["nq"], k), dtype=torch.float32, device=dev)
        Id = torch.empty((w["nq"], k), dtype=torch.int64, device=dev)
        nprobes = ([int(t) for t in args.sweep_nprobes.split(",")] if args.sweep_nprobes
                   else w["nprobes"])
        for p, opts in [(p, o) for p in nprobes for o in option_grid(args.options)]:
            g.nprobe = p
            for oname, v in opts.items():
                g.set_option(oname, v)
            ur = unique_rows(torch, g, xq, p, w["nlist"])
            for _ in range(3):
                g.search(xq, k, D=Dd, I=Id)
            g.reset_stats()
            st = []
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(args.steps):
                g.search(xq, k, D=Dd, I=Id)
                st.append(g.last_timings())
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.steps
            scan = statistics.mean(t["scan"] for t in st)
            ndis = g.get_stats().ndis / args.steps
            out = {"regime": "sweep", "workload": name, "nprobe": p, "options": opts,
                   "exact_redo_total": g.exact_redo_count(),
                   "ms_per_step": ms,
                   "qps": w["nq"] / (ms * 1e-3),
                   "stage_ms": {s: round(statistics.mean(t[s] for t in st), 4) for s in st[0]},
                   "queries_per_probed_row": ndis / max(ur, 1),
                   "scan_gbs_unique": ur * w["d"] * 4.0 / (scan * 1e-3) / 1e9,
                   "scan_frac_of_hbm_peak": ur * w["d"] * 4.0 / (scan * 1e-3) / 1e9 / peak,
                   "scan_useful_tflops": 2.0 * w["d"] * ndis / (scan * 1e-3) / 1e12}
            print(json.dumps(out), flush=True)
        del g, xq
        torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--regime", required=True, choices=["C5", "C4s", "sweep"])
    ap.add_argument("--nb", type=int, default=0, help="override the base size (development)")
    ap.add_argument("--reps", type=int, default=300)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--niter", type=int, default=10)
    ap.add_argument("--no-cpu", action="store_true", help="C5: skip the reference's latency")
    ap.add_argument("--profile-batch", type=int, default=0,
                    help="C5: only run three searches of this batch size between cudaProfilerStart/Stop")
    ap.add_argument("--nprobes", default="1,2,4,8,16,32,64")
    ap.add_argument("--workloads", default="C1,C2,C3")
    ap.add_argument("--sweep-nprobes", default="", help="sweep: these instead of the workload's list")
    ap.add_argument("--options", default="",
                    help="A/B runs: name=v1:v2:...,name2=... every combination of set_option values "
                         "is measured at every nprobe (e.g. simt_max_avg_q=1.5:4:8)")
    args = ap.parse_args()
    import torch
    import hnsw_ivf_b200 as b2
    assert torch.cuda.is_available(), "needs a GPU: the B200 path has no CPU fallback"
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    {"C5": regime_c5, "C4s": regime_c4s, "sweep": regime_sweep}[args.regime](args, torch, b2, dev)


if __name__ == "__main__":
    main()
