// exact.cuh -- the GEOMETRY-argument point x polygon predicates (the GEOS route of the reference: ST_Intersects
// geos_module.cpp:1128-1153, ST_Contains :386-415, ST_Within :1885-1908). GEOS is a third-party dependency that is not
// in the reference tree; what is implemented is its published point-in-area algorithm (algorithm::RayCrossingCounter::
// countSegment + locate::SimplePointInAreaLocator::locatePointInPolygon) with an EXACT orientation sign: an fp64
// filter with Shewchuk's static error bound decides almost every edge, the few ambiguous ones are decided by
// expansion arithmetic (error-free products through FMA, exact summation). Same control flow as the oracle's
// orc_point_polygon_locate, so the two agree on every input, including NaNs.
// location: 0 exterior, 1 boundary, 2 interior. intersects <=> location != 0; within / contains <=> location == 2.
#pragma once
#include "common.cuh"

// exact sign of (bx-ax)*(cy-ay) - (by-ay)*(cx-ax) = ax*by - ax*cy - ay*bx + ay*cx + bx*cy - by*cx, each product exact
// as (p, fma(a, b, -p)), the twelve terms summed as a non-overlapping expansion (grow-expansion). Rare path.
static __device__ __noinline__ int sb_orient2d_exact(double ax, double ay, double bx, double by, double cx, double cy) {
	const double pa[6] = {ax, -ax, -ay, ay, bx, -by};
	const double pb[6] = {by, cy, bx, cx, cy, cx};
	double e[12];
	int n = 0;
#pragma unroll 1
	for (int k = 0; k < 6; k++) {
		const double p = __dmul_rn(pa[k], pb[k]);
		const double r = __fma_rn(pa[k], pb[k], -p);
#pragma unroll 1
		for (int t = 0; t < 2; t++) {
			double q = t == 0 ? r : p;
			for (int i = 0; i < n; i++) {
				// two_sum(q, e[i]) -> (s, err): err stays at position i, s carries on
				const double s = __dadd_rn(q, e[i]);
				const double bv = __dsub_rn(s, q);
				const double av = __dsub_rn(s, bv);
				e[i] = __dadd_rn(__dsub_rn(q, av), __dsub_rn(e[i], bv));
				q = s;
			}
			e[n++] = q;
		}
	}
	for (int i = n - 1; i >= 0; i--) {
		if (e[i] > 0.0) {
			return 1;
		}
		if (e[i] < 0.0) {
			return -1;
		}
	}
	return 0;
}

// +1: c is to the left of a->b (counter-clockwise), -1: right, 0: collinear
__device__ __forceinline__ int sb_orient2d_sign(double ax, double ay, double bx, double by, double cx, double cy) {
	const double dl = __dmul_rn(__dsub_rn(bx, ax), __dsub_rn(cy, ay));
	const double dr = __dmul_rn(__dsub_rn(by, ay), __dsub_rn(cx, ax));
	const double det = __dsub_rn(dl, dr);
	const double eps = 1.1102230246251565e-16;
	const double bound = __dmul_rn((3.0 + 16.0 * eps) * eps, __dadd_rn(fabs(dl), fabs(dr)));
	if (det > bound) {
		return 1;
	}
	if (-det > bound) {
		return -1;
	}
	return sb_orient2d_exact(ax, ay, bx, by, cx, cy);
}

// RayCrossingCounter::countSegment for the segment (x1,y1)->(x2,y2) and the ray from (x,y) to +x.
// returns true when the point lies on the segment; otherwise cr is incremented when the ray crosses it.
__device__ __forceinline__ bool sb_geos_edge(double x, double y, double x1, double y1, double x2, double y2, int &cr) {
	if (x1 < x && x2 < x) {
		return false;
	}
	if (x == x2 && y == y2) {
		return true;
	}
	if (y1 == y && y2 == y) {
		const double mn = x1 < x2 ? x1 : x2, mx = x1 < x2 ? x2 : x1;
		return x >= mn && x <= mx;
	}
	if ((y1 > y && y2 <= y) || (y2 > y && y1 <= y)) {
		int s = sb_orient2d_sign(x1, y1, x2, y2, x, y);
		if (s == 0) {
			return true;
		}
		if (y2 < y1) {
			s = -s;
		}
		if (s > 0) {
			cr++;
		}
	}
	return false;
}

__device__ __forceinline__ int sb_ring_locate_exact(double x, double y, const double *__restrict__ vx,
                                                    const double *__restrict__ vy, uint64_t o, uint64_t e) {
	int cr = 0;
	for (uint64_t i = o + 1; i < e; i++) {
		if (sb_geos_edge(x, y, vx[i - 1], vy[i - 1], vx[i], vy[i], cr)) {
			return 1;
		}
	}
	return (cr & 1) ? 2 : 0;
}

// polygon = rings [r0, r1) of ring_off (ring r0 = shell); an empty polygon has no interior
__device__ __forceinline__ int sb_point_polygon_locate(double x, double y, const double *__restrict__ vx,
                                                       const double *__restrict__ vy,
                                                       const uint64_t *__restrict__ ring_off, uint64_t r0, uint64_t r1) {
	if (r0 >= r1 || x != x || y != y) {
		return 0;
	}
	const int shell = sb_ring_locate_exact(x, y, vx, vy, ring_off[r0], ring_off[r0 + 1]);
	if (shell != 2) {
		return shell;
	}
	for (uint64_t r = r0 + 1; r < r1; r++) {
		const int hole = sb_ring_locate_exact(x, y, vx, vy, ring_off[r], ring_off[r + 1]);
		if (hole == 2) {
			return 0;
		}
		if (hole == 1) {
			return 1;
		}
	}
	return 2;
}
