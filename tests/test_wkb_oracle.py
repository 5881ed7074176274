"""CPU: the oracle's restatement of ST_Extent(WKB_BLOB) (wkb_reader_try_parse_stats, src/sgl/sgl.cpp:903-1058) against
the reference's own function compiled into oracle/_ref/libref.so, on every WKB flavour the reader accepts and on the
inputs it rejects. The reference has no SQL test for ST_Extent(WKB_BLOB); the compiled function is the pin."""
import struct

import numpy as np
import pytest

import geomblob


def random_geom(rng, depth=0):
    def ring():
        k = int(rng.integers(3, 40))
        return [(float(rng.uniform(-1e3, 1e3)), float(rng.uniform(-1e3, 1e3))) for _ in range(k)]

    kind = int(rng.integers(0, 7 if depth < 3 else 3))
    if kind == 0:
        return ("POINT", [ring()[0]] if rng.random() < 0.8 else [])
    if kind == 1:
        return ("LINESTRING", ring()[: int(rng.integers(0, 12))])
    if kind == 2:
        return ("POLYGON", [ring() for _ in range(int(rng.integers(0, 4)))])
    if kind == 3:
        return ("MULTIPOINT", [("POINT", [ring()[0]]) for _ in range(int(rng.integers(0, 5)))])
    if kind == 4:
        return ("MULTILINESTRING", [("LINESTRING", ring()) for _ in range(int(rng.integers(0, 4)))])
    if kind == 5:
        return ("MULTIPOLYGON", [("POLYGON", [ring(), ring()]) for _ in range(int(rng.integers(0, 3)))])
    return ("GEOMETRYCOLLECTION", [random_geom(rng, depth + 1) for _ in range(int(rng.integers(0, 4)))])


def wkb_cases(rng, n=400):
    blobs = []
    for _ in range(n):
        g = random_geom(rng)
        kw = dict(le=bool(rng.integers(0, 2)), z=bool(rng.integers(0, 2)), m=bool(rng.integers(0, 2)),
                  ewkb=bool(rng.integers(0, 2)))
        if kw["ewkb"] and rng.random() < 0.5:
            kw["srid"] = 4326
        blobs.append(geomblob.wkb(g, **kw))
    # NaN vertices inside a linestring never win; an all-NaN point is empty
    nan = float("nan")
    blobs.append(geomblob.wkb(("LINESTRING", [(nan, 1.0), (2.0, nan), (3.0, 4.0)])))
    blobs.append(geomblob.wkb(("POINT", [(nan, nan)])))
    blobs.append(geomblob.wkb(("POINT", [(nan, 5.0)])))
    # rejected inputs: unsupported type, truncations, nesting deeper than 128
    blobs.append(struct.pack("<BI", 1, 9))
    blobs.append(struct.pack("<BI", 1, 17) + b"\0" * 16)
    blobs.append(struct.pack("<BI", 1, 257) + struct.pack("<dd", 1.5, 2.5))  # low byte of 257 is POINT
    good = geomblob.wkb(("POLYGON", [geomblob.SQ, geomblob.SQ2]))
    for cut in (0, 1, 4, 5, 8, 9, 13, 20, len(good) - 1):
        blobs.append(good[:cut])
    deep = b"".join(struct.pack("<BII", 1, 7, 1) for _ in range(129)) + geomblob.wkb(("POINT", [(1, 2)]))
    blobs.append(deep)
    ok_deep = b"".join(struct.pack("<BII", 1, 7, 1) for _ in range(128)) + geomblob.wkb(("POINT", [(1, 2)]))
    blobs.append(ok_deep)
    return blobs


def test_wkb_extent_matches_the_compiled_reference(O):
    if O.ref() is None:
        pytest.skip("oracle/_ref/libref.so not built")
    rng = np.random.default_rng(8)
    blobs = wkb_cases(rng)
    data, off = O.concat_blobs(blobs)
    box, valid, status = O.extent_from_wkb(data, off)
    seen = set()
    for i, b in enumerate(blobs):
        rc, rbox, _ = O.ref_wkb_extent(b)
        seen.add(rc)
        assert valid[i] == (rc == 1), i
        assert status[i] == (rc if rc < 0 else 0), (i, status[i], rc)
        if rc == 1:
            assert (box[i].view(np.uint64) == rbox.view(np.uint64)).all(), i
    assert seen == {1, 0, -1, -2, -3}


def test_wkb_extent_known_answers(O):
    # the ST_Extent expectations of test/sql/geometry/st_extent.test:22-39 hold for the WKB of the same geometries
    # wherever a polygon has no hole outside its shell (ExecuteWKB scans every ring, Execute only the shell)
    blobs = [geomblob.wkb(g) for g, _ in geomblob.EXTENT_CASES]
    data, off = O.concat_blobs(blobs)
    box, valid, status = O.extent_from_wkb(data, off)
    assert (status == 0).all()
    for i, (_, exp) in enumerate(geomblob.EXTENT_CASES):
        assert valid[i] == (exp is not None)
        if exp is not None:
            assert tuple(box[i]) == tuple(map(float, exp))
