"""Generates tests/golden/ref_vectors.npz from the REFERENCE'S OWN code (oracle/_ref/libref.so, compiled from
/root/reference by oracle/Makefile). Run in the build container (where /root/reference exists):

    python scripts/make_golden.py

The fixtures travel with the repo, so parity against the real reference is checkable where /root/reference is absent
(the GPU box): tests/test_golden_fixtures.py compares the oracle (CPU) and the CUDA path (-m gpu) against them."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402

R = O.ref()
assert R is not None, "oracle/_ref/libref.so missing: needs /root/reference"
f32p, f64p, u32p, u8p = (C.POINTER(t) for t in (C.c_float, C.c_double, C.c_uint32, C.c_uint8))
rng = np.random.default_rng(20261104)
out = {}

# MathUtil::DoubleToFloatDown/Up (util/math.hpp:19-47)
d = np.concatenate([rng.uniform(-1e4, 1e4, 1500), rng.uniform(-1, 1, 400) * 10.0 ** rng.integers(-40, 38, 400),
                    [0.0, -0.0, 1.0, 1 + 2**-30, 1 - 2**-30, 3.4028234663852886e38, 3.4028235e38, 3.5e38, -3.5e38,
                     -3.4028234663852886e38, 1e-46, -1e-46, 1e-40, 16777217.0, -16777217.0, 0.1, -0.1, 1e300, -1e300]])
out["round_in"] = d
out["round_down"] = np.array([R.ref_double_to_float_down(float(v)) for v in d], np.float32)
out["round_up"] = np.array([R.ref_double_to_float_up(float(v)) for v in d], np.float32)

# sgl::util::hilbert_encode(16, x, y) (sgl.hpp:1799-1848) incl. 32-bit inputs; hilbert_f32_to_u32 (:1850-1862)
hx = np.concatenate([rng.integers(0, 2**16, 1500), rng.integers(0, 2**32, 1500)]).astype(np.uint32)
hy = np.concatenate([rng.integers(0, 2**16, 1500), rng.integers(0, 2**32, 1500)]).astype(np.uint32)
out["hil_x"], out["hil_y"] = hx, hy
out["hil_out"] = np.array([R.ref_hilbert_encode(16, int(a), int(b)) for a, b in zip(hx, hy)], np.uint32)
ff = np.concatenate([rng.uniform(-1e5, 1e5, 500), [0.0, -0.0, np.inf, -np.inf]]).astype(np.float32)
out["f32_in"] = ff
out["f32_order"] = np.array([R.ref_hilbert_f32_to_u32(C.c_float(v)) for v in ff], np.uint32)

# pcg32 / ST_GeneratePoints (spatial_functions_table.cpp:82-99)
n = 4000
gx, gy = np.empty(n), np.empty(n)
R.ref_generate_points(C.c_uint64(1337), C.c_uint64(n), C.c_double(0), C.c_double(0), C.c_double(10000), C.c_double(10000),
                      gx.ctypes.data_as(f64p), gy.ctypes.data_as(f64p))
out["gen_x"], out["gen_y"] = gx, gy

# FlatRTree (spatial_join_physical.cpp:76-302): arrays of a 6000-box tree with Hilbert-key ties + Lookup results
m = 6000
cx, cy, w = rng.uniform(0, 1000, m), rng.uniform(0, 1000, m), rng.uniform(0, 20, m)
cx[: m // 4], cy[: m // 4] = cx[m // 4 : 2 * (m // 4)], cy[m // 4 : 2 * (m // 4)]
b2d = np.stack([cx - w, cy - w, cx + w, cy + w], 1)
box = np.stack([[R.ref_double_to_float_down(float(v[0])), R.ref_double_to_float_down(float(v[1])),
                 R.ref_double_to_float_up(float(v[2])), R.ref_double_to_float_up(float(v[3]))] for v in b2d]).astype(np.float32)
h = R.ref_flat_build(box.ctypes.data_as(f32p), C.c_uint32(m), C.c_uint32(32))
nodes, nl = R.ref_flat_nodes(h), R.ref_flat_nlayers(h)
fb, fi, fl = np.zeros((nodes, 4), np.float32), np.zeros(nodes, np.uint32), np.zeros(nl, np.uint32)
R.ref_flat_export(h, fb.ctypes.data_as(f32p), fi.ctypes.data_as(u32p), fl.ctypes.data_as(u32p))
out["flat_box2d"], out["flat_in"], out["flat_boxes"], out["flat_idx"], out["flat_layers"] = b2d, box, fb, fi, fl
nq = 400
qx, qy, qw = rng.uniform(-20, 1020, nq), rng.uniform(-20, 1020, nq), rng.uniform(0, 40, nq)
q = np.stack([qx - qw, qy - qw, qx + qw, qy + qw], 1).astype(np.float32)
res, off = [], [0]
buf = np.zeros(m, np.uint32)
for i in range(nq):
    c = R.ref_flat_lookup(h, q[i].ctypes.data_as(f32p), buf.ctypes.data_as(u32p), C.c_uint64(m))
    res.append(np.sort(buf[:c]))
    off.append(off[-1] + c)
out["flat_q"], out["flat_q_off"], out["flat_q_hits"] = q, np.array(off, np.int64), np.concatenate(res)

# probe + exact stand-in predicate through the reference FlatRTree: pair counts for both predicates
npt = 30000
px, py = rng.uniform(-10, 1010, npt), rng.uniform(-10, 1010, npt)
px[:200] = b2d[:200, 0]  # on box edges: closed vs strict differ
py[:200] = (b2d[:200, 1] + b2d[:200, 3]) / 2
cols = [np.ascontiguousarray(b2d[:, k]) for k in range(4)]
cnts = []
for pred in (0, 1):
    cand = C.c_uint64(0)
    tot = R.ref_flat_join_count(C.c_void_p(h), px.ctypes.data_as(f64p), py.ctypes.data_as(f64p), C.c_uint64(npt),
                                *[v.ctypes.data_as(f64p) for v in cols], C.c_int(pred), C.c_int(1), C.byref(cand))
    cnts.append((tot, cand.value))
out["join_px"], out["join_py"], out["join_counts"] = px, py, np.array(cnts, np.int64)
R.ref_flat_free(h)

path = os.path.join(ROOT, "tests", "golden", "ref_vectors.npz")
np.savez_compressed(path, **out)
print("wrote", path, os.path.getsize(path), "bytes;", {k: v.shape for k, v in out.items()})
