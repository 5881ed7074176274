"""The other configs of BASELINE.json (C1 scans, C2, C4, C5), run by bench.py after the headline so that they share its
clock record, each with its own parity check and roofline figure. Device-resident inputs, CUDA-event timing.

  C1  rtree_points_index : 10^4 batched range scans over the 10 M-point index (known answer 3986 is checked in bench.py)
  C2  rtree_index        : 5 M star polygons (k = U{4..64} vertices, variable-length GEOMETRY blobs) + 5 M envelope boxes
                           -> ST_Extent_Approx / exact ST_Extent / CREATE INDEX from blobs / 10^4 range scans
  C4  ST_Within / ST_Contains point-in-polygon : points x 20 000 star polygons of 500 vertices, native winding contract
                           AND GEOS (exact) contract; sharded by points over the ranks (weak)
  C5  ST_DWithin radius join 100 M x 10 M, r = 1 : sharded by probe points over the ranks (strong)
"""
import math
import time

DOMAIN = (0.0, 0.0, 10000.0, 10000.0)


def timed(stream, fn, reps, warm=1):
    import torch

    r = None
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


def _range_queries(ctx, nq=10000):
    qx, qy = ctx.st_generate_points(7, nq, DOMAIN, device=True)
    sx, sy = ctx.st_generate_points(8, nq, (0.0, 0.0, 200.0, 200.0), device=True)
    return ctx.bbox_from_box2d(qx, qy, qx + sx, qy + sy)


def _scan_parity(torch, qb, fbox, oq, orow, nsample=64):
    """batched scan rows of the first `nsample` queries == brute force over all float boxes (Box::Intersects)"""
    ok = True
    q = qb[:nsample]
    for j in range(nsample):
        a = q[j]
        hit = ~((fbox[:, 0] > a[2]) | (fbox[:, 2] < a[0]) | (fbox[:, 1] > a[3]) | (fbox[:, 3] < a[1]))
        exp = torch.nonzero(hit).flatten()
        got = torch.sort(orow[oq == j].long()).values
        ok = ok and exp.numel() == got.numel() and bool((exp == got).all())
    return ok


def run_c1_scans(ctx, sb, torch, dev, stream, idx, fbox, reps=3):
    qb = _range_queries(ctx)
    cap = 1 << 24  # result slots of the caller (9.7 M rows come back): sized like a caller that knows its selectivity
    ms, (oq, orow) = timed(stream, lambda: idx.scan(qb, capacity=cap), reps)
    rows = int(len(orow))
    ok = _scan_parity(torch, qb, fbox, oq, orow)
    return {"workload": "C1: 10^4 batched ST_Intersects range scans (boxes of side U(0,200), seed 7/8) over the 10M-point RTREE",
            "ms": ms, "rows": rows, "rows_per_s": rows / ms * 1e3, "queries_per_s": 1e4 / ms * 1e3,
            "parity": {"ok": bool(ok), "check": "rows of the first 64 queries == brute force over all 10M float boxes"}}


def make_c2_blobs(ctx, torch, dev, n_poly, n_box):
    """C2 of SURVEY 8(d): n_poly star polygons (centres seed 2024, k = U{4..64} vertices, radius U(1,50)) as GEOMETRY blobs
    with a cached float bbox + n_box envelope polygons (side U(1,50)); returns data, offsets, exact extents [n,4] f64."""
    K = 64
    n = n_poly
    per = 4 + 2 * K
    ux, uy = ctx.st_generate_points(2024, n * per // 2, (0.0, 0.0, 1.0, 1.0), device=True)
    u = torch.stack([ux, uy], 1).reshape(n, per)
    del ux, uy
    cx, cy = u[:, 0] * DOMAIN[2], u[:, 1] * DOMAIN[3]
    k = (4 + torch.floor(u[:, 2] * 61.0)).clamp(4, 64).long()  # U{4..64}
    lane = torch.arange(K, device=dev)[None, :]
    live = lane < k[:, None]
    ang = torch.where(live, u[:, 4:4 + K], torch.full_like(u[:, 4:4 + K], 2.0))  # dead lanes sort last
    ang = torch.sort(ang, dim=1).values * (2 * math.pi)
    rad = 1.0 + u[:, 4 + K:4 + 2 * K] * 49.0
    xs = cx[:, None] + rad * torch.cos(ang)
    ys = cy[:, None] + rad * torch.sin(ang)
    del u, ang, rad
    inf = float("inf")
    ext_p = torch.stack([torch.where(live, xs, inf).min(1).values, torch.where(live, ys, inf).min(1).values,
                         torch.where(live, xs, -inf).max(1).values, torch.where(live, ys, -inf).max(1).values], 1)
    # envelope boxes
    bx, by = ctx.st_generate_points(2025, n_box, DOMAIN, device=True)
    sx, sy = ctx.st_generate_points(2026, n_box, (1.0, 1.0, 50.0, 50.0), device=True)
    ext_b = torch.stack([bx - sx / 2, by - sy / 2, bx + sx / 2, by + sy / 2], 1)
    del bx, by, sx, sy
    # blob sizes: 8 header + 16 bbox + 8 (type, ring count) + 8 (ring size + pad) + 16 * (k + 1)
    nv = torch.cat([k + 1, torch.full((n_box,), 5, dtype=torch.long, device=dev)])
    size = 40 + 16 * nv
    off = torch.zeros(n + n_box + 1, dtype=torch.int64, device=dev)
    off[1:] = torch.cumsum(size, 0)
    total = int(off[-1].item())
    data = torch.zeros(total, dtype=torch.uint8, device=dev)
    w32 = data.view(torch.int32)
    f32 = data.view(torch.float32)
    f64 = data.view(torch.float64)
    ext = torch.cat([ext_p, ext_b], 0)
    fb = ctx.bbox_from_box2d(*[ext[:, j].contiguous() for j in range(4)])
    o4 = off[:-1] // 4
    w32[o4] = 2 | (4 << 8)  # type POLYGON, flags = has bbox
    for j in range(4):
        f32[o4 + 2 + j] = fb[:, j]
    w32[o4 + 6] = 2
    w32[o4 + 7] = 1
    w32[o4 + 8] = nv.int()
    # vertices: polygons from the padded matrices (closing vertex = vertex 0), boxes from their corners
    o8 = off[:-1] // 8 + 5
    vcount = k + 1
    vbase = torch.zeros(n + 1, dtype=torch.int64, device=dev)
    vbase[1:] = torch.cumsum(vcount, 0)
    tot_v = int(vbase[-1].item())
    pid = torch.repeat_interleave(torch.arange(n, device=dev), vcount)
    j = torch.arange(tot_v, device=dev) - vbase[pid]
    jj = torch.where(j == k[pid], torch.zeros_like(j), j)
    dst = o8[pid] + 2 * j
    f64[dst] = xs[pid, jj]
    f64[dst + 1] = ys[pid, jj]
    del pid, j, jj, dst, xs, ys
    ob = o8[n:]
    for c, (ix, iy) in enumerate(((0, 1), (0, 3), (2, 3), (2, 1), (0, 1))):
        f64[ob + 2 * c] = ext_b[:, ix]
        f64[ob + 2 * c + 1] = ext_b[:, iy]
    return data, off, ext, fb, total


def run_c2(ctx, sb, torch, dev, stream, hbm, n_poly=5_000_000, n_box=5_000_000, reps=3):
    data, off, ext, fb, total = make_c2_blobs(ctx, torch, dev, n_poly, n_box)
    n = n_poly + n_box
    ams, (box, valid) = timed(stream, lambda: ctx.st_extent_approx(data, off), reps)
    ok_a = bool((box.view(torch.int32) == fb.view(torch.int32)).all()) and bool(valid.all())
    ems, (ebox, evalid) = timed(stream, lambda: ctx.st_extent(data, off), reps)
    ok_e = bool((ebox.view(torch.int64) == ext.view(torch.int64)).all()) and bool(evalid.all())

    def build():
        idx = sb.RTreeIndex.from_blobs(ctx, data, off, None, 128)
        idx.close()

    bms, _ = timed(stream, build, reps)
    idx = sb.RTreeIndex.from_blobs(ctx, data, off, None, 128)
    qb = _range_queries(ctx)
    sms, (oq, orow) = timed(stream, lambda: idx.scan(qb, capacity=1 << 25), reps)
    ok_s = _scan_parity(torch, qb, fb, oq, orow, 32) and idx.indexed_rows == n
    levels = [int(c) for c in idx.level_counts()]
    idx.close()
    return {"workload": "C2: RTREE over %d star-polygon GEOMETRY blobs (k = U{4..64} vertices) + %d envelope polygons (%.2f GB of blobs); "
                        "ST_Extent_Approx, exact ST_Extent, CREATE INDEX from blobs, 10^4 range scans" % (n_poly, n_box, total / 1e9),
            "rows": n, "blob_bytes": total,
            "st_extent_approx_ms": ams, "st_extent_approx_gbs": (40.0 * n) / ams / 1e6,
            "st_extent_exact_ms": ems, "st_extent_exact_gbs": total / ems / 1e6, "st_extent_exact_hbm_frac": total / ems / 1e6 / hbm,
            "build_from_blobs_ms": bms, "build_mkeys_per_s": n / bms / 1e3, "levels": levels,
            "batched_scan_ms": sms, "batched_scan_rows": int(len(orow)), "scan_rows_per_s": len(orow) / sms * 1e3,
            "parity": {"ok": bool(ok_a and ok_e and ok_s), "approx_bits": ok_a, "exact_extent_bits": ok_e, "scan_rows": ok_s,
                       "check": "float boxes bit-identical to the outward-rounded generator extents, exact extents bit-identical to min/max "
                                "over the generator's vertices, scan rows of 32 queries == brute force"}}


def run_c4(ctx, sb, torch, dev, stream, O, rank, world, all_sum, all_max, points=1_000_000_000, npoly=20_000, nverts=500, reps=2):
    import numpy as np

    vx, vy, voff = star_polygons(ctx, torch, 99, npoly, nverts, 20.0, 60.0, dev)
    poly_off = torch.arange(npoly + 1, dtype=torch.int64, device=dev)
    px, py = ctx.st_generate_points(1337, points, DOMAIN, skip=rank * points, device=True)
    fvx, fvy = vx.reshape(-1), vy.reshape(-1)
    cap = int(points * 1.3) + (1 << 20)
    out = (torch.empty(cap, dtype=torch.int32, device=dev), torch.empty(cap, dtype=torch.int32, device=dev))
    res = {"workload": "C4: ST_Within point-in-polygon join, %d points per GPU (seed 1337) x %d star polygons of %d vertices (seed 99), "
                       "polygon-centric kernel; weak scaling over %d GPU(s)" % (points, npoly, nverts, world),
           "points_per_gpu": points, "polygons": npoly}
    checks = {}
    for name, geos in (("native_winding", None), ("geos_exact_within", "within")):
        ctx.profile(True)
        ms, (gp, gq) = timed(stream, lambda: ctx.join_points_polygons(px, py, fvx, fvy, voff, poly_off, out=out, geos=geos), reps)
        kms, kn = ctx.profile_read("pip_join")
        ctx.profile(False)
        ms = all_max(ms)
        pairs = all_sum(int(len(gp)))
        res[name] = {"ms": ms, "pairs": pairs, "points_per_s": points * world / ms * 1e3, "pairs_per_s": pairs / ms * 1e3,
                     "pip_join_kernel_ms": kms / max(kn, 1)}
        # parity 1: a prefix of the points through the OTHER route of the product (candidate list + refine kernel)
        m = min(points, 4_000_000)
        c0, h0 = ctx.pairs_checksum(gp, gq, 0, m)
        ctx.set_option("pip_path", 1)
        try:
            cp, cq = ctx.join_points_polygons(px[:m], py[:m], fvx, fvy, voff, poly_off, geos=geos)
        finally:
            ctx.set_option("pip_path", 0)
        c1, h1 = ctx.pairs_checksum(cp, cq, 0, m)
        # parity 2: a small prefix against the oracle on the host (bbox prefilter + the oracle's predicate)
        ms_ = 20_000
        hx, hy = px[:ms_].cpu().numpy(), py[:ms_].cpu().numpy()
        x0, y0, x1, y1 = [t.cpu().numpy() for t in (vx.min(1).values, vy.min(1).values, vx.max(1).values, vy.max(1).values)]
        hvx, hvy = fvx.cpu().numpy(), fvy.cpu().numpy()
        ro = voff.cpu().numpy().astype(np.uint64)
        po = poly_off.cpu().numpy().astype(np.uint64)
        exp_c, exp_p, exp_q = 0, [], []
        inb = (hx[:, None] >= x0[None, :]) & (hx[:, None] <= x1[None, :]) & (hy[:, None] >= y0[None, :]) & (hy[:, None] <= y1[None, :])
        pi, qi = np.nonzero(inb)
        if geos is None:
            keep = np.array([O.pip_native(float(hx[a]), float(hy[a]), hvx, hvy, ro, int(b), int(b) + 1) for a, b in zip(pi, qi)], np.uint8)
        else:
            keep = O.point_polygon_exact(hx, hy, pi.astype(np.uint32), qi.astype(np.uint32), hvx, hvy, ro, po, pred=1)
        oc, oh = O.pair_checksum(pi[keep == 1], qi[keep == 1])
        gc, gh = ctx.pairs_checksum(gp, gq, 0, ms_)
        ok = (c0, h0) == (c1, h1) and (oc, oh) == (gc, gh)
        checks[name] = {"ok": bool(all_sum(0 if ok else 1) == 0), "routes_agree_on_first": m, "oracle_agrees_on_first": ms_}
        del cp, cq
    res["parity"] = checks
    res["fp64_note"] = "k_pip_join is fp64-pipe / issue bound (profiles/): the kernel time is reported for the roofline discussion"
    del px, py, out
    return res, all(c["ok"] for c in checks.values())


def run_c5(ctx, sb, torch, dev, stream, rank, world, all_sum, all_max, na=100_000_000, nb=10_000_000, r=1.0, reps=2):
    beg = rank * (na // world)
    mine = na // world if rank < world - 1 else na - beg
    ax, ay = ctx.st_generate_points(1337, mine, DOMAIN, skip=beg, device=True)
    bx, by = ctx.st_generate_points(4242, nb, DOMAIN, device=True)
    cap = int(mine * (nb * math.pi * r * r / 1e8) * 1.2) + (1 << 20)
    ms, (ga, gb, gd) = timed(stream, lambda: ctx.join_dwithin(ax, ay, bx, by, r, capacity=cap), reps)
    ms = all_max(ms)
    pairs = all_sum(int(len(ga)))
    # tolerance: returned distance vs fp64 recomputation (sqrt(dx*dx + dy*dy)), 1e-12 relative (north_star)
    s = slice(0, min(len(ga), 5_000_000))
    ia, ib = ga[s].long(), gb[s].long()
    dd = torch.sqrt((ax[ia] - bx[ib]) ** 2 + (ay[ia] - by[ib]) ** 2)
    rel = float(((gd[s] - dd).abs() / dd.clamp_min(1e-300)).max()) if len(ga) else 0.0
    # completeness: brute force for the first 2000 probe rows against all of b
    m = 2000
    exp = 0
    exp_h = 0
    for c0 in range(0, m, 250):
        d2 = (ax[c0:c0 + 250, None] - bx[None, :]) ** 2 + (ay[c0:c0 + 250, None] - by[None, :]) ** 2
        hit = torch.sqrt(d2) <= r
        exp += int(hit.sum().item())
        del d2, hit
    got = int((ga.long() < m).sum().item())
    ok = got == exp and rel <= 1e-12 and bool((gd <= r).all())
    ok = all_sum(0 if ok else 1) == 0
    return {"workload": "C5: ST_DWithin radius join, %d x %d points (seeds 1337 / 4242), r = %g; probe side split over %d GPU(s) (strong)"
                        % (na, nb, r, world),
            "ms": ms, "pairs": pairs, "pairs_per_s": pairs / ms * 1e3, "probes_per_s": na / ms * 1e3,
            "max_rel_dist_err_sample": rel,
            "parity": {"ok": bool(ok), "check": "pairs of the first 2000 probe rows == brute force over all build points; distances within 1e-12 "
                                                 "relative of sqrt(dx*dx+dy*dy); all <= r"}}, ok
