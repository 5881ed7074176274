#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json metric) on N B200s of one node.

Workload (config C3 of BASELINE.json): ST_Intersects spatial join of P random points (ST_GeneratePoints seed 1337,
domain 10^4 x 10^4) against B random boxes (centres: seed 42, side 20) per GPU:
    bbox of the build side -> FlatRTree (Hilbert sort + bottom-up pack) -> batched probe -> exact fp64 refinement.
One "step" = that whole join. `value` = refined pairs / s with inputs resident in HBM; `e2e` = the same through
the C-ABI with HOST buffers (H2D + D2H inside the timed region).

N > 1 (one process per GPU, torch.distributed / NCCL): rank 0 builds the tree and NCCL-broadcasts the packed tree buffer,
every rank probes its own P points (weak scaling, global row ids) and the pair lists are gathered to rank 0 (`value`);
`value_sharded` is the same step with the pair lists left on their GPUs, `strong` the fixed P x B join split N ways.

Parity is part of the measurement: after the timed loop every rank compares count AND an order-independent 64-bit
checksum of its pairs over a prefix of its probe rows with the reference's own FlatRTree (oracle/_ref/libref.so) run
on the host, the C1 index build with the reference's known answers, and the line carries `parity`; a mismatch exits 1.

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
C1_LEVELS = [10_000_000, 78_125, 611, 5, 1]  # benchmark/rtree_points_index.benchmark, SURVEY 8(a) A7
C1_SCAN = 3986  # box (450,450,650,650), benchmark/rtree_points_index.benchmark:11-19


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
    ap.add_argument("--no-extras", action="store_true", help="skip the other BASELINE configs (C1 scan, C2, C4, C5)")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-gather", action="store_true", help="N>1: skip the gathered variant of the step")
    ap.add_argument("--parity-probes", type=int, default=0, help="probe rows per rank checked against the reference (0 = auto)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--_worker", default=None, help=argparse.SUPPRESS)  # internal: SQL reference work in a child process
    ap.add_argument("--_n", type=int, default=0, help=argparse.SUPPRESS)
    ap.add_argument("--_n2", type=int, default=0, help=argparse.SUPPRESS)
    ap.add_argument("--_threads", type=int, default=0, help=argparse.SUPPRESS)
    ap.add_argument("--ctx-opt", action="append", default=[], metavar="KEY=VALUE",
                    help="sb_context_set_option before the run (A/B measurements, e.g. probe_path=2)")
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
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def mark(self):
        return time.perf_counter()

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ts, r in self.rows:
            if t0 is not None and not (t0 <= ts <= t1 + 0.05):
                continue
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
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                "window": "samples taken inside the timed regions (device-resident steps + e2e steps)"}


# ---------------------------------------------------------------------------------------------------------------
# reference arm / CPU baseline / parity checker: the reference's own code on the host cores
# ---------------------------------------------------------------------------------------------------------------
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def make_boxes(O, n_boxes):
    import numpy as np

    u = O.uniform01(42, 2 * n_boxes)
    cx, cy = u[0::2] * DOMAIN[2], u[1::2] * DOMAIN[3]
    h = SIDE / 2
    return [np.ascontiguousarray(v) for v in (cx - h, cy - h, cx + h, cy + h)]


class CpuJoin:
    """The reference's FlatRTree class (oracle/_ref/libref.so, compiled from spatial_join_physical.cpp) in a count / hash
    harness with the exact closed point-in-envelope predicate; OpenMP over 2048-row probe chunks. Falls back to the
    oracle port when libref.so was never built."""

    def __init__(self, n_boxes, threads=None):
        import ctypes as C

        import numpy as np

        import oracle as O

        self.O, self.C, self.np = O, C, np
        self.b = make_boxes(O, n_boxes)
        self.fbox = O.bbox_from_box2d(*self.b)
        self.R = O.ref()
        self.kind = "reference" if self.R is not None else "port"
        self.threads = threads or host_threads()
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

    def checksum(self, px, py, probe_base):
        """-> (pairs, hash) of the reference pair set {(probe_base + i, build row)}"""
        C = self.C
        f64p = C.POINTER(C.c_double)
        if self.R is not None:
            h = C.c_uint64(0)
            n = self.R.ref_flat_join_hash(C.c_void_p(self.tree), px.ctypes.data_as(f64p), py.ctypes.data_as(f64p),
                                          C.c_uint64(len(px)), *[v.ctypes.data_as(f64p) for v in self.b], C.c_int(0),
                                          C.c_int(self.threads), C.c_uint32(probe_base), C.byref(h))
            return int(n), int(h.value)
        pairs, _ = self.O.join_points_boxes(px, py, *self.b, pred=0, threads=self.threads)
        return self.O.pair_checksum(pairs[:, 0].astype(self.np.uint64) + probe_base, pairs[:, 1])

    def sized_sample(self, seconds, cap=50_000_000):
        px, py = self.points(200_000)
        _, t = self.run(px, py)
        rate = 200_000 / max(t, 1e-6)
        return int(min(max(rate * seconds, 200_000), cap))


class SqlRef:
    """The reference's real operators through SQL (oracle/_ref/libref_duckdb.so: DuckDB core + the unmodified reference
    sources): SPATIAL_JOIN with the exact stand-in predicate (GEOS absent) and CREATE INDEX ... USING RTREE."""

    def __init__(self, threads=0):
        from oracle import refdb

        self.refdb = refdb
        self.threads = threads or host_threads()  # never DuckDB's default: it sizes its pool from the HOST, not the cgroup
        self.db = refdb.RefDB(threads=self.threads)

    @staticmethod
    def available():
        try:
            from oracle import refdb

            return refdb.available()
        except Exception:
            return False

    def load_boxes(self, b):
        import numpy as np

        self.db.create_table("b", x0=b[0], y0=b[1], x1=b[2], y1=b[3], k=np.arange(len(b[0])))
        self.db.execute("CREATE OR REPLACE TABLE bg AS SELECT k, ST_MakeEnvelope(x0, y0, x1, y1) AS g FROM b")

    def load_points(self, px, py):
        import numpy as np

        self.db.create_table("p", x=px, y=py, k=np.arange(len(px)))
        self.db.execute("CREATE OR REPLACE TABLE pg AS SELECT k, ST_Point(x, y) AS g FROM p")

    def join_count(self):
        """one full pass of the reference's join operator: build-side sink + FlatRTree build + probe + predicate"""
        t0 = time.perf_counter()
        n = self.db.scalar("SELECT count(*) FROM pg JOIN bg ON ST_Intersects(pg.g, bg.g)")
        return int(n), time.perf_counter() - t0

    def create_index(self, n, threads_note=""):
        """CREATE INDEX ... USING RTREE over n ST_GeneratePoints points (seed 1337) -> (seconds, levels, scan count)"""
        self.db.execute("CREATE OR REPLACE TABLE t1 AS SELECT point::GEOMETRY AS geom FROM st_generatepoints({min_x: 0, min_y: 0, "
                        "max_x: 10000, max_y: 10000}::BOX_2D, %d, 1337)" % n)
        t0 = time.perf_counter()
        self.db.execute("CREATE INDEX c1_idx ON t1 USING RTREE (geom)")
        dt = time.perf_counter() - t0
        lv = self.db.query("SELECT level, count(*) FROM rtree_index_dump('c1_idx') GROUP BY level ORDER BY level DESC", "ii")
        scan = self.db.scalar("SELECT count(*) FROM t1 WHERE ST_Within(geom, ST_MakeEnvelope(450, 450, 650, 650))")
        self.db.execute("DROP INDEX c1_idx")
        return dt, [int(c) for c in lv[1]], int(scan)


def sql_worker(a):
    """child process: the reference's SQL operators (a crash or an assertion inside DuckDB must not take the bench line
    with it). Prints one JSON line."""
    import oracle as O

    sq = SqlRef(threads=a._threads)
    out = {"threads": sq.threads}
    if a._worker == "join":
        sq.load_boxes(make_boxes(O, a.boxes))
        runs = []
        for n in [v for v in (a._n2, a._n) if v]:
            px, py = O.generate_points(1337, n, DOMAIN)
            sq.load_points(px, py)
            for _ in range(a.warmup):
                sq.join_count()
            times, pairs = [], 0
            for _ in range(max(a.steps, 1)):
                pairs, t = sq.join_count()
                times.append(t)
            runs.append({"n": n, "pairs": pairs, "times": times})
        out["runs"] = runs
    else:
        dt, lv, scan = sq.create_index(a._n)
        out.update({"seconds": dt, "levels": lv, "scan": scan, "n": a._n})
    sq.db.close()
    print(json.dumps(out))


def sql_call(kind, a, n, n2=0, steps=1, warmup=0, threads=0, timeout=600):
    """run sql_worker in a child process -> dict, or {"error": ...}"""
    cmd = [sys.executable, os.path.abspath(__file__), "--_worker", kind, "--_n", str(n), "--_n2", str(n2), "--steps", str(steps),
           "--warmup", str(warmup), "--_threads", str(threads), "--boxes", str(a.boxes)]
    # The reference's PhysicalSpatialJoin::Combine appends to the shared global collection without a lock
    # (spatial_join_physical.cpp:529-540), so with a parallel build-side scan its own `item_count == current_position`
    # assertion in FlatRTree::Build fires now and then (seen on the 16-core GPU box, not on 8 cores): retry, then halve
    # the thread count.
    err = None
    for attempt in range(4):
        try:
            if attempt == 3 and kind == "join":
                cmd[cmd.index("--_threads") + 1] = str(max(1, (threads or host_threads()) // 4))
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
            lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
            if r.returncode == 0 and lines:
                out = json.loads(lines[-1])
                out["attempts"] = attempt + 1
                return out
            err = "rc=%d %s" % (r.returncode, (r.stderr or "").strip()[-300:])
        except Exception as e:
            err = repr(e)
    return {"error": err}


def run_reference(a):
    """--impl reference: the reference's own CPU implementation of the path on this box's host cores. With
    libref_duckdb.so: the real SPATIAL_JOIN operator (build side 2 M envelope polygons, a bounded sample of the probe
    points per step); otherwise the FlatRTree count harness."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle as O

    steps, warm = max(a.steps, 1), max(a.warmup, 0)
    budget = 150.0  # seconds for the whole run
    sql = None
    if SqlRef.available():
        cal = sql_call("join", a, 1_600_000, n2=400_000)
        if "error" not in cal:
            t_small, t_big = cal["runs"][0]["times"][0], cal["runs"][1]["times"][0]
            per_probe = max((t_big - t_small) / 1_200_000, 1e-9)
            fixed = max(t_small - 400_000 * per_probe, 0.0)
            per_step = budget / (steps + warm)
            n = int(min(max((per_step - fixed) / per_probe, 400_000), 20_000_000))
            sql = sql_call("join", a, n, steps=steps, warmup=warm, timeout=900)
        else:
            sql = cal
    if sql is not None and "error" not in sql:
        run = sql["runs"][-1]
        tot = sum(run["times"])
        pairs = run["pairs"] * steps
        v = pairs / tot
        kind, cores = "reference", sql["threads"]
        sample = ("the reference's SPATIAL_JOIN operator through SQL (DuckDB %d threads): %d of the %d probe points per step as "
                  "POINT geometries x the full %d envelope polygons, count(*) of the join; every step includes the operator's "
                  "build side (sink + single-threaded FlatRTree build, ~%.2fs); GEOS absent: ST_Intersects(GEOMETRY, GEOMETRY) is "
                  "the exact stand-in of oracle/ref_duckdb.cpp" % (cores, n, a.points, a.boxes, fixed))
        extra = {"probes_per_s": n * steps / tot, "fixed_build_s_per_step": fixed, "marginal_probes_per_s": 1.0 / per_probe}
    else:
        cj = CpuJoin(a.boxes)
        n = cj.sized_sample(max(1.0, min(4.0, 60.0 / max(1, steps + warm))))
        px, py = cj.points(n)
        for _ in range(warm):
            cj.run(px, py)
        pairs, tot = 0, 0.0
        for _ in range(steps):
            p, t = cj.run(px, py)
            pairs += p
            tot += t
        v = pairs / tot
        kind, cores = cj.kind, cj.threads
        sample = ("%d of the %d probe points per step against the full %d-box FlatRTree, count-only harness (tree build %.2fs "
                  "excluded); GEOS absent: exact closed point-in-envelope stand-in predicate" % (n, a.points, a.boxes, cj.build_s))
        extra = {"probes_per_s": n * steps / tot}
        if sql is not None:
            extra["sql_operator_error"] = sql["error"]
    out = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": 1e3 * tot / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a), "sample_points": n},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    out.update(extra)
    print(json.dumps(out))


# ---------------------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------------------
class _Raw:
    def __init__(self, p, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (p, False), "version": 2}


def run_ours(a):
    import ctypes as C
    import importlib

    import numpy as np
    import torch
    import torch.distributed as dist

    import sb200

    sb = sb200.load()
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
    capi = sb.capi
    L = ctx.L
    failures = []
    with torch.cuda.stream(stream):
        ctx.use_torch_stream()
        px, py = ctx.st_generate_points(1337, P, DOMAIN, skip=rank * P, device=True)
        cx, cy = ctx.st_generate_points(42, B, DOMAIN, device=True)
        bx0, by0, bx1, by1 = cx - h, cy - h, cx + h, cy + h
        del cx, cy
        cap = int(P * (B * SIDE * SIDE / (DOMAIN[2] * DOMAIN[3])) * 1.15) + (1 << 20)
        out_p = torch.empty(cap, dtype=torch.int32, device=dev)
        out_b = torch.empty(cap, dtype=torch.int32, device=dev)
        cnt = C.c_uint64(0)

        def barrier():
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()

        def step_single():
            capi.check(L.sb_join_points_boxes(ctx.h, capi.DEVICE, capi.PRED_INTERSECTS, px.data_ptr(), py.data_ptr(), P,
                                              bx0.data_ptr(), by0.data_ptr(), bx1.data_ptr(), by1.data_ptr(), B,
                                              out_p.data_ptr(), out_b.data_ptr(), cap, C.byref(cnt)))
            return cnt.value

        if world > 1:
            # the packed tree's size is a pure function of (B, node size): learn it once. The build side's exact boxes
            # are an INPUT every rank holds (generated from the same seed), only the tree travels.
            t0 = sb.FlatRTree(ctx, ctx.bbox_from_box2d(bx0, by0, bx1, by1), 32) if rank == 0 else None
            nbytes = torch.tensor([t0.device_buffer()[1] if rank == 0 else 0], dtype=torch.int64, device=dev)
            dist.broadcast(nbytes, 0)
            tree_buf = torch.empty(int(nbytes.item()), dtype=torch.uint8, device=dev)
            box2d = torch.stack([bx0, by0, bx1, by1], 1).contiguous()
            if t0 is not None:
                ctx.sync()
                t0.close()
            del t0
            mine = torch.zeros(1, dtype=torch.int64, device=dev)
            counts = torch.zeros(world, dtype=torch.int64, device=dev)
            gath = {"p": None, "b": None}

        def probe_multi(n_points, x, y, base):
            # rank 0: build; everyone: receive the packed tree (one NCCL broadcast), probe own slice, global row ids
            if rank == 0:
                tree = sb.FlatRTree(ctx, ctx.bbox_from_box2d(bx0, by0, bx1, by1), 32)
                p, nb = tree.device_buffer()
                tree_buf.copy_(torch.as_tensor(_Raw(p, nb), device=dev))
            dist.broadcast(tree_buf, 0)
            if rank == 0:
                tree.close()  # stream-ordered free on the context's stream
            t = sb.FlatRTree.attach(ctx, tree_buf)
            capi.check(L.sb_flat_probe_points(ctx.h, t.h, capi.DEVICE, capi.PRED_INTERSECTS, x.data_ptr(), y.data_ptr(),
                                              n_points, base, box2d.data_ptr(), out_p.data_ptr(), out_b.data_ptr(), cap,
                                              C.byref(cnt)))
            t.close()
            return cnt.value

        def step_sharded():
            probe_multi(P, px, py, rank * P)
            return cnt.value

        def step_gather():
            gath["p"] = gath["b"] = None  # rank 0 holds N x 12.8 GB of gathered pairs: free the previous step's first
            n = probe_multi(P, px, py, rank * P)
            gp, gb, cl = D.gather_pairs(out_p[:n], out_b[:n], dst=0)
            gath["p"], gath["b"] = gp, gb
            return n

        sb0, sb1 = D.shard_range(P, rank, world) if world > 1 else (0, P)
        spx = spy = None

        def step_strong():
            gath["p"] = gath["b"] = None
            n = probe_multi(sb1 - sb0, spx, spy, sb0)
            gp, gb, cl = D.gather_pairs(out_p[:n], out_b[:n], dst=0)
            gath["p"], gath["b"] = gp, gb
            return n

        def timed(step, steps, prof=False):
            """-> (ms over `steps` steps, max over ranks; this rank's pairs of the last step)"""
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            n = 0
            for _ in range(steps):
                n = step()
            e1.record(stream)
            barrier()
            ms = e0.elapsed_time(e1)
            if world > 1:
                t = torch.tensor([ms], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            return ms, n

        def all_sum(v):
            if world == 1:
                return int(v)
            t = torch.tensor([int(v)], dtype=torch.int64, device=dev)
            dist.all_reduce(t)
            return int(t.item())

        main_step = step_sharded if world > 1 else step_single
        for _ in range(max(a.warmup, 3)):
            main_step()
        clocks = Clocks(local)
        if rank == 0 and not os.environ.get("SB_BENCH_NOCLOCKS"):
            clocks.start()
        ctx.profile(not os.environ.get("SB_BENCH_NOPROF"))
        main_step()  # one more untimed step: the sampler process and the event pool come up outside the timed region
        ctx.profile(not os.environ.get("SB_BENCH_NOPROF"))
        launches0 = ctx.launches
        tw0 = clocks.mark()
        ms, my_pairs = timed(main_step, a.steps)
        tw1 = clocks.mark()
        launches = ctx.launches - launches0
        probe_ms, probe_n = ctx.profile_read("flat_probe")
        phases = {}
        for tag in ("bp_partition", "bp_cand", "flat_keys", "radix_hist", "radix_onesweep", "flat_pack", "flat_probe"):
            t_ms, t_n = ctx.profile_read(tag)
            phases[tag] = round(t_ms / max(a.steps, 1), 4)
        ctx.profile(False)
        pairs_per_step = all_sum(my_pairs)
        value_sharded = pairs_per_step * a.steps / (ms / 1e3)

        # roofline of the dominant kernel (the probe): algorithmic bytes = 16 B per probe + 8 B per emitted pair
        probe_avg_s = probe_ms / max(probe_n, 1) / 1e3
        algo_bytes = 16.0 * P + 8.0 * my_pairs
        achieved = algo_bytes / probe_avg_s / 1e9 if probe_avg_s > 0 else 0.0
        roofline = {"bound": "hbm", "kernel": "bulk probe kernel (flat_probe profile scope)", "achieved": achieved,
                    "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": None, "peak_source": peak_src,
                    "kernel_ms": probe_ms / max(probe_n, 1), "kernel_share_of_step": probe_ms / max(ms, 1e-9),
                    "phase_ms_per_step": phases, "algorithmic_bytes_per_launch": algo_bytes,
                    "step_algorithmic_frac": algo_bytes / (ms / a.steps / 1e3) / 1e9 / hbm_peak}
        tr = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tr):
            try:
                tj = json.load(open(tr))
                roofline["traffic"] = tj.get("probe_bytes_per_launch", tj.get("k_cp_probe_bytes_per_launch"))
                roofline["kernel"] = tj.get("kernel", roofline["kernel"])
            except Exception:
                pass

        result = {
            "metric": METRIC, "value": value_sharded, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a), "points_per_gpu": P, "boxes": B,
                       "l2": "inputs (%.1f GB of coordinates per step) exceed the 126 MB L2: no flush needed" % (16e-9 * P),
                       "parallelism": "tree broadcast + probe sharding x%d" % world if world > 1 else "single GPU",
                       "gather_pairs": False},
            "pairs_per_step": pairs_per_step, "probes_per_s": P * world * a.steps / (ms / 1e3),
            "gpu_launches": int(launches), "roofline": roofline,
        }

        # ---- parity at the measured size, on every rank -------------------------------------------------------
        ref_prefix = None
        if not a.no_parity:
            np_rows = a.parity_probes or (50_000_000 if world == 1 else 12_500_000)
            np_rows = min(np_rows, P)
            thr = max(1, host_threads() // world)
            gc, gh = ctx.pairs_checksum(out_p, out_b, rank * P, rank * P + np_rows, count=my_pairs)
            tp0 = time.perf_counter()
            cj = CpuJoin(B, threads=thr)
            cpx, cpy = cj.points(np_rows, skip=rank * P)
            rc, rh = cj.checksum(cpx, cpy, rank * P)
            ok = (gc == rc) and (gh == rh)
            ref_prefix = (np_rows, rc, rh)
            tot_c, tot_h = ctx.pairs_checksum(out_p, out_b, 0, 0xFFFFFFFF, count=my_pairs)
            ok = ok and tot_c == my_pairs
            allok = all_sum(0 if ok else 1) == 0
            result["parity"] = {"ok": bool(allok), "checked_probes": np_rows * world, "pairs": all_sum(gc), "ref_pairs": all_sum(rc),
                                "rank0": {"gpu": [gc, "%016x" % gh], "reference": [rc, "%016x" % rh]},
                                "checker": ("reference FlatRTree (oracle/_ref/libref.so)" if cj.kind == "reference" else "oracle port")
                                + ", count + order-independent 64-bit checksum of the (probe row, build row) set over the first "
                                "%d probe rows of every rank's slice; %.1fs" % (np_rows, time.perf_counter() - tp0)}
            if not allok:
                failures.append("C3 pair set differs from the reference on the checked prefix")
            del cj, cpx, cpy

        # ---- N>1: the step with the pair lists gathered to rank 0, and the strong-scaling split of ONE P x B join ----
        if world > 1 and not a.no_gather:
            gsteps = max(2, min(a.steps, 5))
            step_gather()
            gms, n = timed(step_gather, gsteps)
            gbytes = 8.0 * (pairs_per_step - (my_pairs if rank == 0 else 0))
            gb0 = all_sum(int(gbytes) if rank == 0 else 0)
            floor_ms = gb0 / 750e9 * 1e3
            value_g = pairs_per_step * gsteps / (gms / 1e3)
            if rank == 0 and not a.no_parity:
                # the gathered set on rank 0 must be the union of the ranks' sets: counts + checksum of rank 1's prefix
                c_all, _ = ctx.pairs_checksum(gath["p"], gath["b"], 0, 0xFFFFFFFF)
                if c_all != pairs_per_step:
                    failures.append("gathered pair count %d != %d" % (c_all, pairs_per_step))
            result["gathered"] = {"value": value_g, "ms_per_step": gms / gsteps, "steps": gsteps,
                                  "bytes_into_rank0": gb0, "predicted_floor_ms": floor_ms,
                                  "floor_note": "8 B/pair x the other ranks' pairs into ONE GPU at ~750 GB/s NVLink ingest",
                                  "gather_ms": gms / gsteps - ms / a.steps,
                                  "achieved_fraction_of_floor": floor_ms / max(gms / gsteps - ms / a.steps, 1e-9)}
            gath["p"] = gath["b"] = None
            # strong scaling: rows [sb0, sb1) of the SAME P-point probe column (the stream's first P points) on every rank
            spx, spy = ctx.st_generate_points(1337, sb1 - sb0, DOMAIN, skip=sb0, device=True)
            step_strong()
            sms, n = timed(step_strong, gsteps)
            sp = all_sum(n)
            result["strong"] = {"workload": "ONE %d x %d join split over %d GPUs, pairs gathered to rank 0" % (P, B, world),
                                "value": sp * gsteps / (sms / 1e3), "ms_per_step": sms / gsteps, "pairs_per_step": sp}
            gath["p"] = gath["b"] = None
            result["value_sharded"] = value_sharded
            result["value"] = value_g
            result["ms_per_step_sharded"] = ms / a.steps
            result["ms_per_step"] = gms / gsteps
            result["config"]["gather_pairs"] = True

        # ---- e2e: the same join through the C-ABI with HOST (pinned) buffers --------------------------------
        if not a.no_e2e:
            if not run_e2e(a, result, ctx, capi, L, rank, world, dev, px, py, (bx0, by0, bx1, by1), cap, barrier, all_sum,
                           out_p, out_b, ref_prefix):
                failures.append("e2e (host-pointer) pair set differs from the reference on the checked prefix")
        # ---- CREATE INDEX side: STR bulk load of config C1 (10 M points), reported beside the headline --------
        if not a.no_index and rank == 0:
            n1 = 10_000_000
            ix, iy = ctx.st_generate_points(1337, n1, DOMAIN, device=True)
            rows = torch.arange(n1, dtype=torch.int64, device=dev)

            def build_index():
                # POINT_2D column -> index: the outward-rounded float boxes are computed inside the build
                # (sb_rtree_build_from_points = sb_bbox_from_points fused into sb_rtree_build_str, same tree)
                return sb.RTreeIndex.from_points(ctx, ix, iy, rows, 128)

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
            rb = {"workload": "C1: CREATE INDEX USING RTREE, 10M ST_GeneratePoints points (bbox fused into the build + x-slice select / partition + per-slice node select / partition + pack, cap 128)",
                  "mkeys_per_s": n1 / bms / 1e3, "ms": bms, "algorithmic_gbs": 48.2 * n1 / (bms / 1e3) / 1e9,
                  "hbm_frac": 48.2 * n1 / (bms / 1e3) / 1e9 / hbm_peak}
            if not a.no_parity:
                idx = build_index()
                levels = [int(c) for c in idx.level_counts()]
                q = ctx.bbox_from_box2d(*[torch.tensor([v], dtype=torch.float64, device=dev) for v in (450.0, 450.0, 650.0, 650.0)])
                _, got_rows = idx.scan(q)
                rr = got_rows.long()
                inside = int(((ix[rr] >= 450) & (ix[rr] <= 650) & (iy[rr] >= 450) & (iy[rr] <= 650)).sum().item())
                idx.close()
                okb = sorted(levels, reverse=True) == C1_LEVELS and inside == C1_SCAN
                rb["parity"] = {"ok": bool(okb), "levels": levels, "expected_levels": C1_LEVELS, "scan_rows_refined": inside,
                                "expected_scan": C1_SCAN, "candidates": int(len(got_rows))}
                if not okb:
                    failures.append("C1 index build differs from the reference's known answers")
            result["rtree_build"] = rb
            if not a.no_extras:
                import bench_extras as X

                idx = build_index()
                result["c1_scans"] = X.run_c1_scans(ctx, sb, torch, dev, stream, idx, ctx.bbox_from_points(ix, iy))
                idx.close()
                if not result["c1_scans"]["parity"]["ok"]:
                    failures.append("C1 batched scans differ from brute force")
            del ix, iy, rows

        # ---- the other BASELINE configs under the same clock record: C2 (rank 0), C4 (weak, all ranks), C5 (strong) ----
        if not a.no_extras:
            import bench_extras as X

            import oracle as O

            def all_max(v):
                if world == 1:
                    return float(v)
                t = torch.tensor([float(v)], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                return float(t.item())

            del px, py, out_p, out_b
            if world > 1:
                del tree_buf, box2d
            ctx.sync()
            torch.cuda.empty_cache()
            if rank == 0:
                result["c2"] = X.run_c2(ctx, sb, torch, dev, stream, hbm_peak)
                if not result["c2"]["parity"]["ok"]:
                    failures.append("C2 extents / scans differ")
                torch.cuda.empty_cache()
            barrier()
            result["c4"], ok4 = X.run_c4(ctx, sb, torch, dev, stream, O, rank, world, all_sum, all_max)
            if not ok4:
                failures.append("C4 pair sets differ (route agreement / oracle prefix)")
            torch.cuda.empty_cache()
            barrier()
            result["c5"], ok5 = X.run_c5(ctx, sb, torch, dev, stream, rank, world, all_sum, all_max)
            if not ok5:
                failures.append("C5 radius join differs from brute force / tolerance")
        result["clocks"] = clocks.stop(tw0, clocks.mark()) if rank == 0 else None

    # ---- CPU baselines (rank 0, N=1): the reference's own code on this box's host cores ---------------------------
    if rank == 0 and world == 1 and not a.no_cpu:
        try:
            cj = CpuJoin(B)
            n = cj.sized_sample(a.cpu_seconds / 2)
            cpx, cpy = cj.points(n)
            p, t = cj.run(cpx, cpy)
            harness = {"value": p / t, "unit": UNIT, "cores": cj.threads, "kind": cj.kind, "probes_per_s": n / t,
                       "sample": "count-only FlatRTree harness: first %d of the %d probe points against the full %d-box FlatRTree, %.1fs; "
                                 "tree build (%.2fs, single thread) excluded; exact closed point-in-envelope stand-in predicate"
                                 % (n, P, B, t, cj.build_s)}
            result["cpu_baseline"] = harness
            if SqlRef.available():
                m, m2 = 2_000_000, 500_000
                sj = sql_call("join", a, m, n2=m2)
                if "error" in sj:
                    result["cpu_baseline"]["sql_operator_error"] = sj["error"]
                else:
                    (c2, t2), (c1, t1) = [(r["pairs"], r["times"][0]) for r in sj["runs"]]
                    marg = (c1 - c2) / max(t1 - t2, 1e-9)
                    rc, _ = cj.checksum(cpx[:m], cpy[:m], 0)
                    result["cpu_baseline"] = {
                        "value": c1 / t1, "unit": UNIT, "cores": sj["threads"], "kind": "reference", "probes_per_s": m / t1,
                        "marginal_pairs_per_s": marg,
                        "sample": "the reference's SPATIAL_JOIN operator through SQL (oracle/_ref/libref_duckdb.so, DuckDB %d threads): first %d of "
                                  "the %d probe points as POINT geometries x the full %d envelope polygons, count(*) = %d pairs in %.1fs incl. the "
                                  "operator's build side; marginal rate from a second run on %d points; GEOS absent: exact stand-in predicate"
                                  % (sj["threads"], m, P, B, c1, t1, m2),
                        "operator_pairs_match_flatrtree_harness": bool(c1 == rc),
                        "harness": harness}
                    if c1 != rc:
                        failures.append("reference SPATIAL_JOIN count %d != FlatRTree harness %d" % (c1, rc))
                if "rtree_build" in result:
                    si = sql_call("index", a, 10_000_000)
                    s1 = sql_call("index", a, 2_000_000, threads=1)
                    if "error" in si or "error" in s1:
                        result["rtree_build"]["cpu_baseline_error"] = si.get("error") or s1.get("error")
                    else:
                        result["rtree_build"]["cpu_baseline"] = {
                            "value": 10.0 / si["seconds"], "unit": "Mkeys/s", "cores": si["threads"], "kind": "reference",
                            "seconds": si["seconds"], "single_thread_mkeys_per_s": 2.0 / s1["seconds"],
                            "sample": "the reference's CREATE INDEX ... USING RTREE through SQL on the same 10M points (all threads), and on 2M "
                                      "points with threads=1; levels %s, ST_Within box query -> %d rows" % (si["levels"], si["scan"])}
                        if si["levels"] != C1_LEVELS[:-1] or si["scan"] != C1_SCAN:
                            failures.append("reference CREATE INDEX answers %s / %d differ from the known answers" % (si["levels"], si["scan"]))
        except Exception as e:  # the baseline is a reported extra, never a reason to lose the GPU line
            result.setdefault("cpu_baseline", {"value": None, "unit": UNIT, "cores": host_threads(), "kind": "port", "sample": "failed: %r" % (e,)})
            result["cpu_baseline_error"] = repr(e)
    if rank == 0:
        if failures:
            result["parity_failures"] = failures
        print(json.dumps(result))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if failures:
        sys.exit(1)


def run_e2e(a, result, ctx, capi, L, rank, world, dev, px, py, boxes, cap, barrier, all_sum, out_p, out_b, ref_prefix):
    """e2e through the C-ABI with HOST (pinned) buffers, copies inside the timed region.
    N = 1: sb_join_points_boxes(SB_HOST). N > 1: rank 0 alone drives all N devices through sb_join_points_boxes_multi --
    ONE P x B join (its own probe column) split over N PCIe links, one pair list back (strong scaling); the other
    ranks wait on the host. Returns False when the host-pointer result fails the parity check."""
    import ctypes as C

    import torch
    import torch.distributed as dist

    import sb200

    sb = sb200.load()
    P, B = a.points, a.boxes
    cnt = C.c_uint64(0)
    ok = True
    store = dist.distributed_c10d._get_default_store() if world > 1 else None
    if rank == 0:
        hp = [t.cpu().pin_memory() for t in (px, py) + tuple(boxes)]
        ho_p = torch.empty(cap, dtype=torch.int32).pin_memory()
        ho_b = torch.empty(cap, dtype=torch.int32).pin_memory()
        extra = [sb.Context(d) for d in range(world) if d != ctx.device] if world > 1 else []
        ctxs = [ctx] + extra
        hs = (C.c_void_p * len(ctxs))(*[c.h for c in ctxs])

        def step_e2e():
            if world == 1:
                capi.check(L.sb_join_points_boxes(ctx.h, capi.HOST, capi.PRED_INTERSECTS, hp[0].data_ptr(), hp[1].data_ptr(),
                                                  P, hp[2].data_ptr(), hp[3].data_ptr(), hp[4].data_ptr(), hp[5].data_ptr(),
                                                  B, ho_p.data_ptr(), ho_b.data_ptr(), cap, C.byref(cnt)))
            else:
                capi.check(L.sb_join_points_boxes_multi(hs, len(ctxs), capi.PRED_INTERSECTS, hp[0].data_ptr(), hp[1].data_ptr(),
                                                        P, hp[2].data_ptr(), hp[3].data_ptr(), hp[4].data_ptr(),
                                                        hp[5].data_ptr(), B, ho_p.data_ptr(), ho_b.data_ptr(), cap,
                                                        C.byref(cnt)))
            return cnt.value

        for _ in range(3):  # warm-up: the first calls grow (and then coalesce) the contexts' staging arenas
            step_e2e()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ep = 0
        for _ in range(a.e2e_steps):
            ep += step_e2e()
        for d in range(world):
            torch.cuda.synchronize(d)
        et = time.perf_counter() - t0
        e2e = {"value": ep / et, "unit": UNIT, "h2d_bytes_per_step": 16 * P + 32 * B,
               "d2h_bytes_per_step": 8 * (ep // a.e2e_steps), "ms_per_step": 1e3 * et / a.e2e_steps,
               "note": ("sb_join_points_boxes(SB_HOST): pinned host columns in, pair arrays out, copies inside the timed region" if world == 1 else
                        "sb_join_points_boxes_multi over %d devices from ONE process (rank 0): one %d x %d join, pinned host columns in, one host "
                        "pair list out, tree replicated by peer copies, every device streams its probe slice over its own PCIe link; "
                        "strong scaling of the fixed job" % (world, P, B))}
        if ref_prefix is not None:
            # the host-pointer result of the last step against the reference prefix (rank 0's slice starts at row 0)
            n = cnt.value
            out_p[:n].copy_(ho_p[:n])
            out_b[:n].copy_(ho_b[:n])
            np_rows, rc, rh = ref_prefix
            c, hsh = ctx.pairs_checksum(out_p, out_b, 0, np_rows, count=n)
            ok = (c == rc) and (hsh == rh)
            e2e["parity"] = {"ok": bool(ok), "checked_probes": np_rows, "pairs": c, "ref_pairs": rc}
        result["e2e"] = e2e
        for c in extra:
            c.close()
        del hp, ho_p, ho_b
        if store is not None:
            store.set("sb_e2e_done", "1")
    elif store is not None:
        store.wait(["sb_e2e_done"])  # host-side wait: no spinning NCCL kernel on this rank's GPU while rank 0 measures
    if world > 1:
        dist.barrier()
    return ok


def main():
    a = parse()
    if a._worker:
        sql_worker(a)
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
