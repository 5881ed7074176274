// cells.cuh -- the quantiser grid over a tree box shared by the bulk probes (bucket_probe.cu) and the polygon join
// (pip_join.cu): a monotone 20-bit quantiser per axis = 32 x 32 buckets of 32 x 32 sub-cells of 1024 quanta.
#pragma once
#include "internal.cuh"

#define BP_G 32                  // buckets per axis
#define BP_NB (BP_G * BP_G)      // 1024 buckets (+1 for points with NaN coordinates)
#define BP_QBITS 20              // quantiser bits per axis
#define BP_QMAX ((1 << BP_QBITS) - 1)
#ifndef BP_SLICE
#define BP_SLICE 65536u          // points per work unit
#endif
#define BP_UNIT_TILES (BP_SLICE / 16384u) // 16384-record tiles per unit (pass 2 of the cell partition, k_pip_gather)

struct bp_grid {
	double minx, miny, sx, sy; // Q(v) = clamp(floor((v - min) * s), 0, BP_QMAX)
	float tb[4];               // tree box
};

__device__ __forceinline__ int bp_q(double v, double mn, double s) {
	const int c = __double2int_rd((v - mn) * s); // saturating; NaN -> 0
	return min(max(c, 0), BP_QMAX);
}

// Quantised extent of a leaf float box under the CANDIDATE semantics of the reference: a point x can produce a
// float box [down(x), up(x)] that intersects [b.x, b.z] only if pred(b.x) < x < succ(b.z) (pred / succ = the
// neighbouring floats), so [Q(pred(b.x)), Q(succ(b.z))] covers every such point. NaN bounds never exclude
// (Box::Intersects, bbox.hpp:27-29, is written with negated comparisons) => open-ended.
__device__ __forceinline__ void bp_leaf_range(const float4 b, const bp_grid &g, int &x0, int &y0, int &x1, int &y1) {
	const float ax = nextafterf(b.x, -INFINITY), ay = nextafterf(b.y, -INFINITY);
	const float cx = nextafterf(b.z, INFINITY), cy = nextafterf(b.w, INFINITY);
	x0 = ax != ax ? 0 : bp_q((double)ax, g.minx, g.sx);
	y0 = ay != ay ? 0 : bp_q((double)ay, g.miny, g.sy);
	x1 = cx != cx ? BP_QMAX : bp_q((double)cx, g.minx, g.sx);
	y1 = cy != cy ? BP_QMAX : bp_q((double)cy, g.miny, g.sy);
}


// ---- offsets, cursors and the work-unit table (written by k_bp_scan) ------------------------------------------------
struct bp_unit {
	uint32_t bucket, pbeg, pend;
};
// tables: [0] pt_hist (1025) [1] pt_off (1026) [2] pt_cursor (1025) [3] cand_cnt (1024) [4] cand_off (1025)
//         [5] cand_cursor (1024); totals: {points kept (no NaN), candidates, units, NaN points}
struct bp_tables {
	uint32_t pt_hist[BP_NB + 1];
	uint32_t pt_off[BP_NB + 2];
	uint32_t pt_cursor[BP_NB + 1];
	uint32_t cand_cnt[BP_NB];
	uint32_t cand_off[BP_NB + 1];
	uint32_t cand_cursor[BP_NB];
	uint32_t ticket;
	uint32_t pad;
	uint64_t totals[4];
};

