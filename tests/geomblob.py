"""Builds serialized GEOMETRY blobs for arbitrary geometry kinds (test helper).

Layout restated from Serde::Serialize, src/spatial/geometry/geometry_serialization.cpp:132-242.
A geometry is a nested tuple: ("POINT", [(x,y)] or []), ("LINESTRING", [(x,y),...]),
("POLYGON", [ring, ring, ...]), ("MULTIPOINT"|"MULTILINESTRING"|"MULTIPOLYGON"|"GEOMETRYCOLLECTION", [geom, ...]).
"""
import struct

TYPES = ["POINT", "LINESTRING", "POLYGON", "MULTIPOINT", "MULTILINESTRING", "MULTIPOLYGON", "GEOMETRYCOLLECTION"]


def _all_vertices(g):
    t, body = g
    if t in ("POINT", "LINESTRING"):
        return list(body)
    if t == "POLYGON":
        return [v for ring in body for v in ring]
    return [v for part in body for v in _all_vertices(part)]


def _is_empty(g):
    t, body = g
    return len(body) == 0


def _payload(g):
    t, body = g
    out = struct.pack("<II", TYPES.index(t), len(body))
    if t in ("POINT", "LINESTRING"):
        for x, y in body:
            out += struct.pack("<dd", x, y)
    elif t == "POLYGON":
        for ring in body:
            out += struct.pack("<I", len(ring))
        if len(body) % 2 == 1:
            out += b"\0\0\0\0"
        for ring in body:
            for x, y in ring:
                out += struct.pack("<dd", x, y)
    else:
        for part in body:
            out += _payload(part)
    return out


def blob(g, O):
    t, body = g
    has_bbox = t != "POINT" and not _is_empty(g)
    flags = 0x04 if has_bbox else 0
    out = struct.pack("<BBHI", TYPES.index(t), flags, 0, 0)
    if has_bbox:
        vs = _all_vertices(g)
        L = O.lib()
        if vs:
            xs = [v[0] for v in vs]
            ys = [v[1] for v in vs]
            bb = (L.orc_double_to_float_down(min(xs)), L.orc_double_to_float_down(min(ys)),
                  L.orc_double_to_float_up(max(xs)), L.orc_double_to_float_up(max(ys)))
        else:  # non-empty collection of empties: box_xyzm::smallest() rounded
            big = 1.7976931348623157e308
            bb = (L.orc_double_to_float_down(big), L.orc_double_to_float_down(big),
                  L.orc_double_to_float_up(-big), L.orc_double_to_float_up(-big))
        out += struct.pack("<ffff", *bb)
    out += _payload(g)
    assert len(out) % 8 == 0
    return out


SQ = [(0, 0), (1, 0), (1, 1), (0, 1), (0, 0)]
SQ2 = [(2, 2), (3, 2), (3, 3), (2, 3), (2, 2)]

# the 15 kinds of test/sql/geometry/st_extent.test:6-20 and the expected ST_Extent (:22-39)
EXTENT_CASES = [
    (("POINT", []), None),
    (("POINT", [(0, 0)]), (0, 0, 0, 0)),
    (("LINESTRING", []), None),
    (("LINESTRING", [(0, 0), (1, 1)]), (0, 0, 1, 1)),
    (("POLYGON", []), None),
    (("POLYGON", [SQ]), (0, 0, 1, 1)),
    (("MULTIPOINT", []), None),
    (("MULTIPOINT", [("POINT", [(0, 0)]), ("POINT", [(1, 1)])]), (0, 0, 1, 1)),
    (("MULTILINESTRING", []), None),
    (("MULTILINESTRING", [("LINESTRING", [(0, 0), (1, 1)]), ("LINESTRING", [(2, 2), (3, 3)])]), (0, 0, 3, 3)),
    (("MULTIPOLYGON", []), None),
    (("MULTIPOLYGON", [("POLYGON", [SQ]), ("POLYGON", [SQ2])]), (0, 0, 3, 3)),
    (("GEOMETRYCOLLECTION", []), None),
    (("GEOMETRYCOLLECTION", [("POINT", [(0, 0)]), ("LINESTRING", [(0, 0), (1, 1)])]), (0, 0, 1, 1)),
    (("GEOMETRYCOLLECTION", [("GEOMETRYCOLLECTION", [("POINT", [(0, 0)]), ("LINESTRING", [(0, 0), (1, 1)])])]), (0, 0, 1, 1)),
]


# ---- WKB (ISO and EWKB flavours, both byte orders) for ST_Extent(WKB_BLOB) ---------------------------------------
def wkb(g, le=True, z=False, m=False, ewkb=False, srid=None, extra=7.5):
    """WKB bytes of a nested-tuple geometry. ISO: type + 1000 (Z) / 2000 (M) / 3000 (ZM); EWKB: 0x80000000 / 0x40000000
    flags, optional SRID (0x20000000) on this part. Z / M ordinates are filled with `extra`."""
    t, body = g
    e = "<" if le else ">"
    code = TYPES.index(t) + 1
    if ewkb:
        code |= (0x80000000 if z else 0) | (0x40000000 if m else 0) | (0x20000000 if srid is not None else 0)
    else:
        code += 1000 * ((1 if z else 0) + (2 if m else 0))
    out = struct.pack("B", 1 if le else 0) + struct.pack(e + "I", code)
    if ewkb and srid is not None:
        out += struct.pack(e + "I", srid)
    nd = (1 if z else 0) + (1 if m else 0)

    def vert(v):
        return struct.pack(e + "dd", float(v[0]), float(v[1])) + struct.pack(e + "d", extra) * nd

    if t == "POINT":
        out += vert(body[0]) if body else struct.pack(e + "d", float("nan")) * (2 + nd)
    elif t == "LINESTRING":
        out += struct.pack(e + "I", len(body)) + b"".join(vert(v) for v in body)
    elif t == "POLYGON":
        out += struct.pack(e + "I", len(body))
        for ring in body:
            out += struct.pack(e + "I", len(ring)) + b"".join(vert(v) for v in ring)
    else:
        out += struct.pack(e + "I", len(body))
        for part in body:
            out += wkb(part, le=le, z=z, m=m, ewkb=ewkb, extra=extra)
    return out
