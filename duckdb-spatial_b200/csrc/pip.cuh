// pip.cuh -- ST_Contains::Operation (POLYGON_2D x POINT_2D winding number), spatial_functions_scalar.cpp:1938-2051,
// reached through ST_Within(POINT_2D, POLYGON_2D) :7775-7781. Literal control flow, plain fp64 (no FMA): an on-edge
// hit sets contains = false and only leaves the coordinate loop (:2011-2016); the ring is then judged by the winding
// number accumulated so far -- the shell overwrites `contains`, a hole that is not "in_ring" leaves it false (so a
// point on a hole's boundary is outside; pinned on the reference's own ST_Within, tests/test_oracle_vs_refdb.py).
#pragma once
#include "common.cuh"

// one edge of the coordinate loop; returns true when the loop must stop (point on the edge)
__device__ __forceinline__ bool sb_pip_edge(double x, double y, double x1, double y1, double x2, double y2, int &wn) {
	if (!(x1 == x2 && y1 == y2)) {
		const double y_min = y2 < y1 ? y2 : y1;
		const double y_max = y1 < y2 ? y2 : y1;
		if (!(y > y_max || y < y_min)) {
			const double side_v =
			    __dsub_rn(__dmul_rn(__dsub_rn(x, x1), __dsub_rn(y2, y1)), __dmul_rn(__dsub_rn(x2, x1), __dsub_rn(y, y1)));
			if (side_v == 0.0 &&
			    (((x1 <= x && x < x2) || (x1 >= x && x > x2)) || ((y1 <= y && y < y2) || (y1 >= y && y > y2)))) {
				return true;
			} else if (side_v < 0.0 && (y1 < y && y <= y2)) {
				wn++;
			} else if (side_v > 0.0 && (y2 <= y && y < y1)) {
				wn--;
			}
		}
	}
	return false;
}

// polygon = rings [r0, r1) of ring_off; ring r = vertices [ring_off[r], ring_off[r+1]) of vx / vy
__device__ __forceinline__ bool sb_pip_native(double x, double y, const double *__restrict__ vx,
                                              const double *__restrict__ vy, const uint64_t *__restrict__ ring_off,
                                              uint64_t r0, uint64_t r1) {
	bool contains = false, first = true;
	for (uint64_t r = r0; r < r1; r++) {
		const uint64_t o = ring_off[r], e = ring_off[r + 1];
		double x1 = 0.0, y1 = 0.0;
		if (o < e) { // an empty last ring must not read past the vertex columns
			x1 = vx[o], y1 = vy[o];
		}
		int wn = 0;
		for (uint64_t c = o + 1; c < e; c++) {
			const double x2 = vx[c], y2 = vy[c];
			if (sb_pip_edge(x, y, x1, y1, x2, y2, wn)) {
				contains = false;
				break;
			}
			x1 = x2;
			y1 = y2;
		}
		const bool in_ring = wn != 0;
		if (first) {
			if (!in_ring) {
				contains = false;
				break;
			}
			contains = true;
		} else if (in_ring) {
			contains = false;
			break;
		}
		first = false;
	}
	return contains;
}
