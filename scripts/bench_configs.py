#!/usr/bin/env python
"""Secondary configs of BASELINE.json (C1, C2, C4, C5) on one B200, device-resident inputs, CUDA-event timing.
bench.py measures the headline (C3); this script is the evidence for the other rows of SURVEY.md section 8(d).

    python scripts/bench_configs.py [--c1 N] [--c2 N] [--c4 POINTS] [--c4-polys N] [--c5 NA NB] [--reps R]
prints one JSON object per config.
"""
import argparse
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
DOMAIN = (0.0, 0.0, 10000.0, 10000.0)


def timed(stream, fn, reps, warm=2):
    import torch

    for _ in range(warm):
        r = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        r = fn()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, r


def star_polygons(ctx, torch, seed, npoly, k, rmin, rmax, dev):
    """npoly closed rings of k+1 vertices: centre uniform in DOMAIN, sorted random angles, per-vertex radius
    U(rmin, rmax) (non-self-intersecting by construction). Uniforms come from the pcg32 stream of `seed`."""
    per = 2 + 2 * k
    ux, uy = ctx.st_generate_points(seed, npoly * per // 2, (0.0, 0.0, 1.0, 1.0), device=True)
    u = torch.stack([ux, uy], 1).reshape(npoly, per)
    cx, cy = u[:, 0:1] * DOMAIN[2], u[:, 1:2] * DOMAIN[3]
    ang = torch.sort(u[:, 2:2 + k], dim=1).values * (2 * math.pi)
    rad = rmin + u[:, 2 + k:2 + 2 * k] * (rmax - rmin)
    xs, ys = cx + rad * torch.cos(ang), cy + rad * torch.sin(ang)
    vx = torch.cat([xs, xs[:, :1]], 1).contiguous()
    vy = torch.cat([ys, ys[:, :1]], 1).contiguous()
    voff = (torch.arange(npoly + 1, device=dev, dtype=torch.int64) * (k + 1)).contiguous()
    return vx, vy, voff


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--c1", type=int, default=10_000_000)
    ap.add_argument("--c2", type=int, default=5_000_000)
    ap.add_argument("--c4", type=int, default=1_000_000_000)
    ap.add_argument("--c4-polys", type=int, default=20_000)
    ap.add_argument("--c4-verts", type=int, default=500)
    ap.add_argument("--c5", type=int, nargs=2, default=[100_000_000, 10_000_000])
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--only", default="")
    ap.add_argument("--ctx-opt", action="append", default=[], metavar="KEY=VALUE", help="sb_context_set_option, for A/B runs")
    a = ap.parse_args()
    import numpy as np
    import torch

    import sb200

    sb = sb200.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    ctx = sb.Context(0)
    for kv in a.ctx_opt:
        k, v = kv.split("=")
        ctx.set_option(k, int(v))
    stream = torch.cuda.Stream(device=dev)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    only = set(a.only.split(",")) if a.only else None

    def want(c):
        return only is None or c in only

    with torch.cuda.stream(stream):
        ctx.use_torch_stream()
        # ---- C1: CREATE INDEX USING RTREE on random POINT_2D + the known-answer box query + 10^4 batched queries ----
        if want("c1") and a.c1:
            n = a.c1
            x, y = ctx.st_generate_points(1337, n, DOMAIN, device=True)
            rows = torch.arange(n, dtype=torch.int64, device=dev)

            def build():
                idx = sb.RTreeIndex(ctx, ctx.bbox_from_points(x, y), rows, 128)
                idx.close()

            ms, _ = timed(stream, build, a.reps)
            idx = sb.RTreeIndex(ctx, ctx.bbox_from_points(x, y), rows, 128)
            q1 = ctx.bbox_from_box2d(*[torch.tensor([v], dtype=torch.float64, device=dev) for v in (450.0, 450.0, 650.0, 650.0)])
            _, r1 = idx.scan(q1)
            qx, qy = ctx.st_generate_points(7, 10000, DOMAIN, device=True)
            sx, sy = ctx.st_generate_points(8, 10000, (0.0, 0.0, 200.0, 200.0), device=True)
            qb = ctx.bbox_from_box2d(qx, qy, qx + sx, qy + sy)
            cap = 1 << 26
            sms, (oq, orow) = timed(stream, lambda: idx.scan(qb, capacity=cap), a.reps)
            pms, (pinfo, pleaf, pbranch, _) = timed(stream, lambda: idx.pages(device=True), a.reps)
            pbytes = 24.0 * sum(idx.level_counts()) + pleaf.numel() + pbranch.numel()
            print(json.dumps({"config": "C1 rtree_points_index", "keys": n, "build_ms": ms, "build_mkeys_per_s": n / ms / 1e3,
                              "build_algorithmic_gbs": 48.2 * n / ms / 1e6, "build_hbm_frac": 48.2 * n / ms / 1e6 / hbm,
                              "levels": idx.level_counts(), "known_answer_(450,450,650,650)": int(len(r1)),
                              "batched_scan_queries": 10000, "batched_scan_ms": sms, "batched_scan_rows": int(len(orow)),
                              "scan_rows_per_s": len(orow) / sms * 1e3,
                              "export_pages_ms": pms, "export_pages": [pinfo["leaf_pages"], pinfo["branch_pages"]],
                              "export_pages_gbs": pbytes / pms / 1e6, "export_pages_hbm_frac": pbytes / pms / 1e6 / hbm}),
                  flush=True)
            idx.close()
            del x, y, rows, oq, orow, pleaf, pbranch
        # ---- C2: RTREE over polygon / box GEOMETRY blobs + range scans ---------------------------------------------
        if want("c2") and a.c2:
            n = a.c2
            # n envelope polygons (ST_MakeEnvelope layout: 5-vertex ring, cached float bbox), built on the device
            cx, cy = ctx.st_generate_points(2024, n, DOMAIN, device=True)
            sx, sy = ctx.st_generate_points(2025, n, (1.0, 1.0, 50.0, 50.0), device=True)
            x0, y0, x1, y1 = cx - sx / 2, cy - sy / 2, cx + sx / 2, cy + sy / 2
            fb = ctx.bbox_from_box2d(x0, y0, x1, y1)
            blob = torch.zeros((n, 120), dtype=torch.uint8, device=dev)  # 8 hdr + 16 bbox + 8 + 8 + 5*16
            blob[:, 0] = 2  # POLYGON
            blob[:, 1] = 4  # has bbox
            blob[:, 8:24] = fb.view(torch.uint8).reshape(n, 16)
            i32 = blob.view(torch.int32).reshape(n, 30)
            i32[:, 6] = 2  # type
            i32[:, 7] = 1  # ring count
            i32[:, 8] = 5  # vertices in ring 0
            f64 = blob.view(torch.float64).reshape(n, 15)
            for k, (vx_, vy_) in enumerate(((x0, y0), (x0, y1), (x1, y1), (x1, y0), (x0, y0))):
                f64[:, 5 + 2 * k] = vx_
                f64[:, 6 + 2 * k] = vy_
            data = blob.reshape(-1)
            off = (torch.arange(n + 1, dtype=torch.int64, device=dev) * 120).contiguous()
            rows = torch.arange(n, dtype=torch.int64, device=dev)
            del cx, cy, sx, sy

            def build():
                box, valid = ctx.st_extent_approx(data, off)
                idx = sb.RTreeIndex(ctx, box, rows, 128)
                idx.close()

            ms, _ = timed(stream, build, a.reps)
            box, valid = ctx.st_extent_approx(data, off)
            assert bool((box.view(torch.int32) == fb.view(torch.int32)).all()) and bool(valid.all())
            ems, _ = timed(stream, lambda: ctx.st_extent(data, off), a.reps)
            idx = sb.RTreeIndex(ctx, box, rows, 128)
            qx, qy = ctx.st_generate_points(7, 10000, DOMAIN, device=True)
            s2x, s2y = ctx.st_generate_points(8, 10000, (0.0, 0.0, 200.0, 200.0), device=True)
            qb = ctx.bbox_from_box2d(qx, qy, qx + s2x, qy + s2y)
            cap = 1 << 26
            sms, (oq, orow) = timed(stream, lambda: idx.scan(qb, capacity=cap), a.reps)
            print(json.dumps({"config": "C2 rtree_index (envelope-polygon GEOMETRY blobs, 120 B each)", "keys": n,
                              "build_from_blobs_ms": ms, "build_mkeys_per_s": n / ms / 1e3,
                              "st_extent_exact_ms": ems, "st_extent_gbs": 120.0 * n / ems / 1e6,
                              "levels": idx.level_counts(), "batched_scan_queries": 10000, "batched_scan_ms": sms,
                              "batched_scan_rows": int(len(orow)), "scan_rows_per_s": len(orow) / sms * 1e3}), flush=True)
            idx.close()
            del blob, data, off, rows, box, fb, oq, orow, x0, y0, x1, y1
        # ---- C4: ST_Within(point, polygon) join: points x star polygons ---------------------------------------------
        if want("c4") and a.c4:
            P, NP, K = a.c4, a.c4_polys, a.c4_verts
            vx, vy, voff = star_polygons(ctx, torch, 99, NP, K, 20.0, 60.0, dev)
            poly_off = torch.arange(NP + 1, dtype=torch.int64, device=dev)
            px, py = ctx.st_generate_points(1337, P, DOMAIN, device=True)
            fvx, fvy = vx.reshape(-1), vy.reshape(-1)
            if hasattr(ctx, "join_points_polygons"):
                cap = int(P * 1.3) + (1 << 20)
                out = (torch.empty(cap, dtype=torch.int32, device=dev), torch.empty(cap, dtype=torch.int32, device=dev))
                ms, (gp, gb) = timed(stream, lambda: ctx.join_points_polygons(px, py, fvx, fvy, voff, poly_off, out=out), a.reps, warm=1)
                res = {"path": "sb_join_points_polygons", "ms": ms, "pairs": int(len(gp))}
            else:
                ext = (vx.min(1).values, vy.min(1).values, vx.max(1).values, vy.max(1).values)
                fbox = ctx.bbox_from_box2d(*[e.contiguous() for e in ext])
                cap = int(P * 1.6) + (1 << 20)
                out = (torch.empty(cap, dtype=torch.int32, device=dev), torch.empty(cap, dtype=torch.int32, device=dev))

                def join():
                    tree = sb.FlatRTree(ctx, fbox, 32)
                    cp, cb = tree.probe_points(px, py, out=out)
                    tree.close()
                    return ctx.st_within_point_polygon(px, py, fvx, fvy, voff, poly_off, cp, cb), len(cp)

                ctx.profile(True)
                ms, ((gp, gb), ncand) = timed(stream, join, a.reps, warm=1)
                pm, pn = ctx.profile_read("refine_pip")
                fm, fn_ = ctx.profile_read("flat_probe")
                ctx.profile(False)
                res = {"path": "probe(candidates) + sb_refine_point_polygon", "ms": ms, "pairs": int(len(gp)), "candidates": int(ncand),
                       "refine_ms": pm / max(pn, 1), "probe_ms": fm / max(fn_, 1)}
            res.update({"config": "C4 ST_Within(point, polygon)", "points": P, "polygons": NP, "vertices": K,
                        "points_per_s": P / res["ms"] * 1e3, "pairs_per_s": res["pairs"] / res["ms"] * 1e3})
            print(json.dumps(res), flush=True)
            del px, py, vx, vy, out
        # ---- C5: ST_DWithin radius join ------------------------------------------------------------------------------
        if want("c5") and a.c5[0]:
            na, nb = a.c5
            ax, ay = ctx.st_generate_points(1337, na, DOMAIN, device=True)
            bx, by = ctx.st_generate_points(4242, nb, DOMAIN, device=True)
            r = 1.0
            cap = int(na * (nb * math.pi * r * r / 1e8) * 1.2) + (1 << 20)
            ms, (ga, gb, gd) = timed(stream, lambda: ctx.join_dwithin(ax, ay, bx, by, r, capacity=cap), a.reps, warm=1)
            # tolerance check on a sample: returned distance vs fp64 recomputation in torch (sqrt(dx*dx+dy*dy))
            s = slice(0, min(len(ga), 5_000_000))
            ia, ib = ga[s].long(), gb[s].long()
            dd = torch.sqrt((ax[ia] - bx[ib]) ** 2 + (ay[ia] - by[ib]) ** 2)
            rel = float(((gd[s] - dd).abs() / dd.clamp_min(1e-300)).max()) if len(ga) else 0.0
            print(json.dumps({"config": "C5 ST_DWithin radius join", "a": na, "b": nb, "radius": r, "ms": ms, "pairs": int(len(ga)),
                              "pairs_per_s": len(ga) / ms * 1e3, "probes_per_s": na / ms * 1e3, "max_rel_dist_err_sample": rel,
                              "all_within_r": bool((gd <= r).all())}), flush=True)


if __name__ == "__main__":
    main()
