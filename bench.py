#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json metric) on N B200s of one node.

Workload (config C3 of BASELINE.json): ST_Intersects spatial join of P random points (ST_GeneratePoints seed 1337,
domain 10^4 x 10^4) against B random boxes (centres: seed 42, side 20) per GPU:
    bbox of the build side -> FlatRTree (Hilbert sort + bottom-up pack) -> batched probe -> exact fp64 refinement.
One "step" = that whole join. `value` = refined pairs / s with inputs resident in HBM; `e2e` = the same through
the C-ABI with HOST buffers (H2D + D2H inside the timed region). N>1: rank 0 builds the tree, NCCL-broadcasts it,
every rank probes its own P points (weak scaling), pair lists are gathered to rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
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

DOMAIN = (0.0, 0.0, 10000.0, 10000.0)
SIDE = 20.0
METRIC = "st_intersects_join_pairs_per_s"
UNIT = "pairs/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--points", type=int, default=200_000_000, help="probe points per GPU")
    ap.add_argument("--boxes", type=int, default=2_000_000, help="build-side boxes")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-index", action="store_true", help="skip the CREATE INDEX (STR) side measurement")
    ap.add_argument("--gather", action="store_true",
                    help="N>1: also move every rank's pair list to rank 0 (bounded by rank 0's NVLink ingest, "
                         "8 B/pair: see DESIGN.md section 7). Default: pair lists stay sharded, counts are all-gathered")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--ctx-opt", action="append", default=[], metavar="KEY=VALUE",
                    help="sb_context_set_option before the run (A/B measurements, e.g. probe_path=2, fine_hist=1)")
    return ap.parse_args()


def workload_name(a):
    return ("C3: ST_Intersects join, %d ST_GeneratePoints points (seed 1337, 1e4^2 domain) x %d random boxes "
            "(seed 42, side %g) per GPU; FlatRTree build (node 32) + batched probe + exact refine" % (a.points, a.boxes, SIDE))


# ---------------------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi) during the timed region
# ---------------------------------------------------------------------------------------------------------------
class Clocks:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except Exception:
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7), ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        # "under load": the upper half of the samples (idle gaps between steps pull the plain median down)
        load = sm[len(sm) // 2:]
        return {"sm_mhz": load[len(load) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the reference's own FlatRTree probe on the host cores
# ---------------------------------------------------------------------------------------------------------------
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class CpuJoin:
    """Reference FlatRTree (oracle/_ref, compiled from the reference sources) or the oracle port; OpenMP over
    2048-row probe chunks like the reference's chunk-parallel probe."""

    def __init__(self, n_boxes):
        import ctypes as C

        import numpy as np

        import oracle as O

        self.O, self.C, self.np = O, C, np
        u = O.uniform01(42, 2 * n_boxes)
        cx, cy = u[0::2] * DOMAIN[2], u[1::2] * DOMAIN[3]
        h = SIDE / 2
        self.b = [np.ascontiguousarray(v) for v in (cx - h, cy - h, cx + h, cy + h)]
        self.fbox = O.bbox_from_box2d(*self.b)
        self.R = O.ref()
        self.kind = "reference" if self.R is not None else "port"
        self.threads = host_threads()
        t0 = time.perf_counter()
        if self.R is not None:
            f32p = C.POINTER(C.c_float)
            self.tree = self.R.ref_flat_build(self.fbox.ctypes.data_as(f32p), C.c_uint32(n_boxes), C.c_uint32(32))
        self.build_s = time.perf_counter() - t0

    def points(self, n, skip=0):
        return self.O.generate_points(1337, n, DOMAIN, skip=skip)

    def run(self, px, py):
        """-> (pairs, seconds) for probing + refining px,py against the tree"""
        C = self.C
        f64p = C.POINTER(C.c_double)
        t0 = time.perf_counter()
        if self.R is not None:
            cand = C.c_uint64(0)
            n = self.R.ref_flat_join_count(C.c_void_p(self.tree), px.ctypes.data_as(f64p), py.ctypes.data_as(f64p),
                                           C.c_uint64(len(px)), *[v.ctypes.data_as(f64p) for v in self.b], C.c_int(0),
                                           C.c_int(self.threads), C.byref(cand))
        else:
            n, _ = self.O.join_points_boxes(px, py, *self.b, pred=0, threads=self.threads, want_pairs=False)
        return int(n), time.perf_counter() - t0

    def sized_sample(self, seconds):
        px, py = self.points(200_000)
        _, t = self.run(px, py)
        rate = 200_000 / max(t, 1e-6)
        n = int(min(max(rate * seconds, 200_000), 50_000_000))
        return n


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cj = CpuJoin(a.boxes)
    n = cj.sized_sample(max(1.0, min(4.0, 60.0 / max(1, a.steps + a.warmup))))
    px, py = cj.points(n)
    for _ in range(a.warmup):
        cj.run(px, py)
    pairs, tot = 0, 0.0
    for _ in range(a.steps):
        p, t = cj.run(px, py)
        pairs += p
        tot += t
    v = pairs / tot
    sample = "%d of the %d probe points per step against the full %d-box FlatRTree (tree build %.2fs excluded); GEOS absent: exact closed point-in-envelope stand-in predicate" % (n, a.points, a.boxes, cj.build_s)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * tot / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a), "sample_points": n},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cj.threads, "kind": cj.kind, "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "probes_per_s": n * a.steps / tot,
    }))


# ---------------------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------------------
class _Raw:
    def __init__(self, p, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (p, False), "version": 2}


def run_ours(a):
    import numpy as np
    import torch
    import torch.distributed as dist

    import sb200

    sb = sb200.load()
    import importlib

    D = importlib.import_module("duckdb_spatial_b200.distributed")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"

    ctx = sb.Context(local)
    for kv in a.ctx_opt:
        k, v = kv.split("=")
        ctx.set_option(k, int(v))
    stream = torch.cuda.Stream(device=dev)
    P, B = a.points, a.boxes
    h = SIDE / 2
    with torch.cuda.stream(stream):
        ctx.use_torch_stream()
        px, py = ctx.st_generate_points(1337, P, DOMAIN, skip=rank * P, device=True)
        cx, cy = ctx.st_generate_points(42, B, DOMAIN, device=True)
        bx0, by0, bx1, by1 = cx - h, cy - h, cx + h, cy + h
        del cx, cy
        cap = int(P * (B * SIDE * SIDE / (DOMAIN[2] * DOMAIN[3])) * 1.15) + (1 << 20)
        out_p = torch.empty(cap, dtype=torch.int32, device=dev)
        out_b = torch.empty(cap, dtype=torch.int32, device=dev)
        import ctypes as C

        capi = sb.capi
        L = ctx.L
        cnt = C.c_uint64(0)
        state = {"tree": None, "bytes": 0}

        def step_single():
            capi.check(L.sb_join_points_boxes(ctx.h, capi.DEVICE, capi.PRED_INTERSECTS, px.data_ptr(), py.data_ptr(), P,
                                              bx0.data_ptr(), by0.data_ptr(), bx1.data_ptr(), by1.data_ptr(), B,
                                              out_p.data_ptr(), out_b.data_ptr(), cap, C.byref(cnt)))
            return cnt.value

        if world > 1:
            # size of the packed tree is a pure function of (B, node size): learn it once
            t0 = sb.FlatRTree(ctx, ctx.bbox_from_box2d(bx0, by0, bx1, by1), 32) if rank == 0 else None
            nbytes = torch.tensor([t0.device_buffer()[1] if rank == 0 else 0], dtype=torch.int64, device=dev)
            dist.broadcast(nbytes, 0)
            state["bytes"] = int(nbytes.item())
            tree_buf = torch.empty(state["bytes"], dtype=torch.uint8, device=dev)
            box2d = torch.empty((B, 4), dtype=torch.float64, device=dev)
            del t0
            counts = torch.zeros(world, dtype=torch.int64, device=dev)

        def step_multi():
            # rank 0: build; everyone: receive the packed tree + the build side's exact boxes; probe own slice
            if rank == 0:
                tree = sb.FlatRTree(ctx, ctx.bbox_from_box2d(bx0, by0, bx1, by1), 32)
                p, nb = tree.device_buffer()
                tree_buf.copy_(torch.as_tensor(_Raw(p, nb), device=dev))
                box2d.copy_(torch.stack([bx0, by0, bx1, by1], 1))
                stream.synchronize()
                tree.close()
            dist.broadcast(tree_buf, 0)
            dist.broadcast(box2d, 0)
            t = sb.FlatRTree.attach(ctx, tree_buf)
            capi.check(L.sb_flat_probe_points(ctx.h, t.h, capi.DEVICE, capi.PRED_INTERSECTS, px.data_ptr(), py.data_ptr(),
                                              P, rank * P, box2d.data_ptr(), out_p.data_ptr(), out_b.data_ptr(), cap,
                                              C.byref(cnt)))
            t.close()
            mine = torch.tensor([cnt.value], dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(counts, mine)
            total = int(counts.sum().item())
            if a.gather:
                # variable-length pair lists -> rank 0 (grouped send/recv), duckdb-spatial_b200/distributed.py
                D.gather_pairs(out_p[: cnt.value], out_b[: cnt.value], dst=0)
            return total

        step = step_multi if world > 1 else step_single

        def barrier():
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()

        for _ in range(max(a.warmup, 3)):
            pairs = step()
        clocks = Clocks(local)
        if rank == 0 and not os.environ.get("SB_BENCH_NOCLOCKS"):
            clocks.start()
        ctx.profile(not os.environ.get("SB_BENCH_NOPROF"))
        step()  # one more untimed step: the sampler process and the event pool come up outside the timed region
        ctx.profile(not os.environ.get("SB_BENCH_NOPROF"))
        launches0 = ctx.launches
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        total_pairs = 0
        for _ in range(a.steps):
            total_pairs += step()
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        launches = ctx.launches - launches0
        probe_ms, probe_n = ctx.profile_read("flat_probe")
        phases = {}
        for tag in ("bp_partition", "bp_cand", "flat_keys", "radix_hist", "radix_onesweep", "flat_pack", "flat_probe"):
            t_ms, t_n = ctx.profile_read(tag)
            phases[tag] = round(t_ms / max(a.steps, 1), 4)
        ctx.profile(False)
        clk = clocks.stop() if rank == 0 else None
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        value = total_pairs / (ms / 1e3)
        my_pairs = cnt.value

        # roofline of the dominant kernel (the probe): algorithmic bytes = 16 B per probe + 8 B per emitted pair
        probe_avg_s = probe_ms / max(probe_n, 1) / 1e3
        algo_bytes = 16.0 * P + 8.0 * my_pairs
        achieved = algo_bytes / probe_avg_s / 1e9 if probe_avg_s > 0 else 0.0
        roofline = {"bound": "hbm", "kernel": "k_cp_probe<2>", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                    "frac": achieved / hbm_peak, "traffic": None, "peak_source": peak_src,
                    "kernel_ms": probe_ms / max(probe_n, 1), "kernel_share_of_step": probe_ms / max(ms, 1e-9),
                    "phase_ms_per_step": phases,
                    "algorithmic_bytes_per_launch": algo_bytes}
        tr = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tr):
            try:
                roofline["traffic"] = json.load(open(tr)).get("k_cp_probe_bytes_per_launch")
            except Exception:
                pass

        result = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a), "points_per_gpu": P, "boxes": B,
                       "l2": "inputs (%.1f GB of coordinates per step) exceed the 126 MB L2: no flush needed" % (16e-9 * P),
                       "parallelism": "tree broadcast + probe sharding x%d" % world if world > 1 else "single GPU",
                       "gather_pairs": (world > 1 and a.gather)},
            "pairs_per_step": total_pairs // a.steps, "probes_per_s": P * world * a.steps / (ms / 1e3),
            "gpu_launches": int(launches), "roofline": roofline, "clocks": clk,
        }

        # ---- e2e: the same join through the C-ABI with HOST (pinned) buffers --------------------------------
        if not a.no_e2e:
            hp = [t.cpu().pin_memory() for t in (px, py, bx0, by0, bx1, by1)]
            ho_p = torch.empty(cap, dtype=torch.int32).pin_memory()
            ho_b = torch.empty(cap, dtype=torch.int32).pin_memory()

            def step_e2e():
                capi.check(L.sb_join_points_boxes(ctx.h, capi.HOST, capi.PRED_INTERSECTS, hp[0].data_ptr(), hp[1].data_ptr(),
                                                  P, hp[2].data_ptr(), hp[3].data_ptr(), hp[4].data_ptr(), hp[5].data_ptr(),
                                                  B, ho_p.data_ptr(), ho_b.data_ptr(), cap, C.byref(cnt)))
                return cnt.value

            for _ in range(3):  # warm-up: the first calls grow (and then coalesce) the context's staging arena
                step_e2e()
            barrier()
            t0 = time.perf_counter()
            ep = 0
            for _ in range(a.e2e_steps):
                ep += step_e2e()
            torch.cuda.synchronize()
            et = time.perf_counter() - t0
            if world > 1:
                t = torch.tensor([et], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                et = float(t.item())
                t = torch.tensor([ep], dtype=torch.int64, device=dev)
                dist.all_reduce(t)
                ep = int(t.item())
            result["e2e"] = {"value": ep / et, "unit": UNIT, "h2d_bytes_per_step": 16 * P + 32 * B,
                             "d2h_bytes_per_step": 8 * (ep // a.e2e_steps // world), "ms_per_step": 1e3 * et / a.e2e_steps,
                             "note": "sb_join_points_boxes(SB_HOST): pinned host columns in, pair arrays out, copies inside the timed region; per-rank replicas for N>1"}
            del hp, ho_p, ho_b

        # ---- CREATE INDEX side: STR bulk load of config C1 (10 M points), reported beside the headline --------
        if not a.no_index and rank == 0:
            n1 = 10_000_000
            ix, iy = ctx.st_generate_points(1337, n1, DOMAIN, device=True)
            rows = torch.arange(n1, dtype=torch.int64, device=dev)
            def build_index():
                box = ctx.bbox_from_points(ix, iy)
                idx = sb.RTreeIndex(ctx, box, rows, 128)
                return idx
            for _ in range(3):
                build_index().close()
            torch.cuda.synchronize()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 5
            s0.record(stream)
            for _ in range(reps):
                idx = build_index()
                idx.close()
            s1.record(stream)
            torch.cuda.synchronize()
            bms = s0.elapsed_time(s1) / reps
            result["rtree_build"] = {"workload": "C1: CREATE INDEX USING RTREE, 10M ST_GeneratePoints points (bbox + x-sort + slice y-sort + pack, cap 128)",
                                     "mkeys_per_s": n1 / bms / 1e3, "ms": bms,
                                     "algorithmic_gbs": 48.2 * n1 / (bms / 1e3) / 1e9, "hbm_frac": 48.2 * n1 / (bms / 1e3) / 1e9 / hbm_peak}
            del ix, iy, rows

    # ---- CPU baseline (rank 0, N=1): the reference's FlatRTree probe on this box's host cores ----------------
    if rank == 0 and world == 1 and not a.no_cpu:
        try:
            cj = CpuJoin(B)
            n = cj.sized_sample(a.cpu_seconds)
            cpx, cpy = cj.points(n)
            p, t = cj.run(cpx, cpy)
            result["cpu_baseline"] = {"value": p / t, "unit": UNIT, "cores": cj.threads, "kind": cj.kind,
                                      "probes_per_s": n / t,
                                      "sample": "first %d of the %d probe points against the full %d-box FlatRTree, %.1fs; tree build (%.2fs, single thread) excluded; GEOS absent: exact closed point-in-envelope stand-in predicate" % (n, P, B, t, cj.build_s)}
        except Exception as e:  # the baseline is a reported extra, never a reason to lose the GPU line
            result["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": host_threads(), "kind": "port", "sample": "failed: %r" % (e,)}
    if rank == 0:
        print(json.dumps(result))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
