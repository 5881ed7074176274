// pip_join.cu -- ST_Within(point, polygon) / ST_Contains(polygon, point) spatial join, polygon-centric (config C4:
// very many probe points x few large polygons).
//
// Reference path per probe row (spatial_join_physical.cpp:748-849): TryGetCachedBounds of the point -> FlatRTree scan
// against the polygons' cached float boxes -> exact predicate per candidate = the winding-number routine
// ST_Contains::Operation (spatial_functions_scalar.cpp:1938-2051, pip.cuh). A pair (point, polygon) is emitted iff the
// two float boxes intersect (Box::Intersects, bbox.hpp:27-29) AND the routine returns true. The candidate-list route
// (probe -> pair list -> sb_refine_point_polygon) writes and re-reads 2.7 candidate pairs per point at C4 and walks
// all ~500 edges per pair. This kernel instead:
//
//   1. sorts the points into the 1024 x 1024 cells of the quantiser grid (the v3 front end, bucket_probe.cu) and
//      gathers their coordinates into that order;
//   2. gives every polygon to one CTA, which builds an index of the polygon's edges in shared memory: a 32 x 32 grid
//      over the polygon's exact extent; per grid cell (a) W0 = the winding contribution of all edges that lie wholly to
//      the right of the cell and span its row of the grid, (b) the list of the edges that still have to be evaluated;
//   3. streams the points of the grid cells under the polygon's float box (contiguous record ranges per cell row),
//      one point per lane: float-box candidate test, then wn = W0 + the reference's edge evaluation over the cell's
//      list (same arithmetic, same operands), `contains = wn != 0`.
//
// Why the result is bit-identical to walking every edge in order:
//   * cells come from monotone maps fx(x) = clamp(floor((x - minx) * sx)), fy likewise, applied to points and to edge
//     bounds alike. Rows: an edge with y-range [eymin, eymax] covers rows [fy(eymin), fy(eymax)]; a point in another
//     row has y outside [eymin, eymax] and the reference skips the edge (`y > y_max || y < y_min`). A point in a row
//     strictly between them has eymin < y < eymax, so the half-open y test of the crossing rule holds.
//   * columns: with lo = exmin - m, hi = exmax + m, m = max(|dx| 2^-40, 2^-500) (rounded outward), a point in a column
//     left of fx(lo) has x < lo, one right of fx(hi) has x > hi. For such a point every factor of side_v =
//     (x-x1)(y2-y1) - (x2-x1)(y-y1) is far from cancellation: |exact| >= |dy| min(|x-x1|, |x-x2|) >= |dy| m while the
//     rounding error is below 3u(|t1| + |t2|) <= |dy| m * 7e-4 (+ 2^-1073 of underflow, excluded by the eligibility
//     bounds below), so the computed sign is the exact sign and never 0: an edge wholly to the LEFT of the point
//     contributes nothing, one wholly to the RIGHT contributes +1 (upward) / -1 (downward) when it spans the row.
//   * the winding number is a sum, so the order of evaluation only matters when a point lies ON an edge (side_v == 0
//     inside the edge's span): the reference then leaves the loop with the count so far. Such a hit can only involve a
//     listed edge; it is detected and the point is re-evaluated with the literal in-order loop.
// Polygons that are not eligible for the index (holes, > 1024 edges, non-finite or extreme coordinates, list
// overflow) are evaluated with the literal loop for every candidate point. NaN probes never match (every comparison
// of the routine is false), they are dropped by the partition.
#include "cells.cuh"
#include "pip.cuh"
#include "exact.cuh"

#define PIP_G 32 // (the row scans below assume one warp = one row set: PIP_G == 32)
#define PIP_NCELL (PIP_G * PIP_G)
#define PIP_MAXE 1024
#define PIP_LMAX 12288
#define PIP_THREADS 256
#define PIP_WARPS (PIP_THREADS / 32)
#define PIP_STAGE 256

struct pip_poly {
	double minx, miny, maxx, maxy; // exact extent over all vertices
	float4 fbox;                   // the blob's cached bbox: extent rounded outward (geometry_serialization.cpp:226-230)
	uint32_t flags;                // 0 = no vertices (never a candidate), 1 = literal loop, 2 = index eligible
	uint32_t pad[3];
};

__device__ __forceinline__ uint32_t pip_f32_ord(float f) {
	const uint32_t r = __float_as_uint(f);
	return (r & 0x80000000u) ? ~r : (r | 0x80000000u);
}
__device__ __forceinline__ float pip_ord_f32(uint32_t k) {
	return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}

// one warp per polygon: extent, float box, eligibility; the union of the float boxes (= the FlatRTree's tree box)
__global__ void __launch_bounds__(256) k_pip_prepare(const double *__restrict__ vx, const double *__restrict__ vy,
                                                     const uint64_t *__restrict__ ring_off,
                                                     const uint64_t *__restrict__ poly_off, uint64_t n_polys,
                                                     pip_poly *__restrict__ polys, float4 *__restrict__ boxes,
                                                     uint32_t *__restrict__ tb_ord) {
	const unsigned lane = sb_lane();
	const uint64_t wid = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((uint64_t)gridDim.x * blockDim.x) >> 5;
	for (uint64_t p = wid; p < n_polys; p += nw) {
		const uint64_t r0 = poly_off[p], r1 = poly_off[p + 1];
		const uint64_t v0 = r0 < r1 ? ring_off[r0] : 0, v1 = r0 < r1 ? ring_off[r1] : 0;
		double mnx = INFINITY, mny = INFINITY, mxx = -INFINITY, mxy = -INFINITY;
		bool ok = (r1 - r0 == 1) && (v1 - v0 >= 2) && (v1 - v0 - 1 <= PIP_MAXE);
		for (uint64_t i = v0 + lane; i < v1; i += 32) {
			const double x = vx[i], y = vy[i];
			mnx = x < mnx ? x : mnx, mxx = x > mxx ? x : mxx;
			mny = y < mny ? y : mny, mxy = y > mxy ? y : mxy;
			ok = ok && fabs(x) <= 0x1p500 && fabs(y) <= 0x1p500; // false for NaN
			if (ok && i + 1 < v1) {
				const double dx = fabs(vx[i + 1] - x), dy = fabs(vy[i + 1] - y);
				ok = (dx == 0.0 || dx >= 0x1p-500) && (dy == 0.0 || dy >= 0x1p-500);
			}
		}
#pragma unroll
		for (int s = 16; s; s >>= 1) {
			const double a = __shfl_xor_sync(0xFFFFFFFFu, mnx, s), b = __shfl_xor_sync(0xFFFFFFFFu, mny, s);
			const double c = __shfl_xor_sync(0xFFFFFFFFu, mxx, s), d = __shfl_xor_sync(0xFFFFFFFFu, mxy, s);
			mnx = a < mnx ? a : mnx, mny = b < mny ? b : mny, mxx = c > mxx ? c : mxx, mxy = d > mxy ? d : mxy;
		}
		ok = __all_sync(0xFFFFFFFFu, ok);
		if (lane == 0) {
			pip_poly P;
			P.minx = mnx, P.miny = mny, P.maxx = mxx, P.maxy = mxy;
			P.pad[0] = P.pad[1] = P.pad[2] = 0;
			if (v1 == v0 || !(mnx <= mxx) || !(mny <= mxy)) {
				// no vertices (or only NaN ones): TryGetCachedBounds fails, the row is never a candidate.
				// The marker box maps to an empty range of the quantiser grid.
				P.flags = 0;
				P.fbox = make_float4(FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX);
			} else {
				P.flags = ok ? 2u : 1u;
				P.fbox = make_float4(sb_f64_to_f32_down(mnx), sb_f64_to_f32_down(mny), sb_f64_to_f32_up(mxx), sb_f64_to_f32_up(mxy));
				atomicMin(&tb_ord[0], pip_f32_ord(P.fbox.x));
				atomicMin(&tb_ord[1], pip_f32_ord(P.fbox.y));
				atomicMax(&tb_ord[2], pip_f32_ord(P.fbox.z));
				atomicMax(&tb_ord[3], pip_f32_ord(P.fbox.w));
			}
			polys[p] = P;
			boxes[p] = P.fbox;
		}
	}
}

__global__ void k_pip_header(const uint32_t *__restrict__ tb_ord, uint32_t n_polys, sb_flat_header *__restrict__ hdr) {
	if (threadIdx.x == 0) {
		hdr->magic = SB_FLAT_MAGIC;
		hdr->item_count = n_polys;
		hdr->node_size = 32;
		hdr->nodes = n_polys;
		hdr->nlayers = 1;
		for (int i = 0; i < 4; i++) {
			hdr->tree_box[i] = pip_ord_f32(tb_ord[i]);
		}
	}
}

// coordinates of the cell-sorted records, in record order (only the buckets that have work units are sorted). The
// source positions of a 16384-record tile of pass 2 lie inside that tile (a 256 KB window of the bucket-ordered
// coordinates): every fetched sector is used completely. (Gathering from the caller's columns through the probe id
// instead cost 254 B of DRAM reads per point: profiles r1f_full_keys_pip_join.txt.)
__global__ void __launch_bounds__(512) k_pip_gather(const bp_tables *__restrict__ T, const bp_unit *__restrict__ units,
                                                    const uint32_t *__restrict__ src, const double2 *__restrict__ xy_in,
                                                    double2 *__restrict__ xy) {
	const uint32_t n_units = (uint32_t)T->totals[2];
	for (uint32_t w = blockIdx.x; w < BP_UNIT_TILES * n_units; w += gridDim.x) {
		const bp_unit unit = units[w / BP_UNIT_TILES];
		const uint32_t beg = unit.pbeg + (w % BP_UNIT_TILES) * 16384u, end = min(unit.pend, beg + 16384u);
		for (uint32_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
			xy[i] = xy_in[src[i]];
		}
	}
}

struct pip_smem {
	double vx[PIP_MAXE + 1], vy[PIP_MAXE + 1];
	uchar4 rng[PIP_MAXE]; // per edge: x = first row, y = last row, z = first column (of lo), w = last column (of hi)
	int w0[PIP_NCELL];    // difference array per row, then W0 per cell
	uint32_t cnt[PIP_NCELL];  // cell lists: counts, then fill cursors
	uint32_t rcnt[PIP_NCELL]; // row lists: entries per (row, first column), then fill cursors
	uint16_t off[PIP_NCELL + 8];
	uint16_t npre[PIP_NCELL]; // per cell: how many entries of its row's list (sorted by first column, descending) lie right of it
	uint16_t rowoff[PIP_G + 8];
	uint16_t rowlist[2 * PIP_MAXE];
	uint16_t list[PIP_LMAX];
	uint32_t stage[PIP_WARPS][PIP_STAGE];
	uint32_t scan[PIP_WARPS];
	uint32_t poly, n_list;
};

__device__ __forceinline__ int pip_cell(double v, double mn, double s) {
	const int c = __double2int_rd(__dmul_rn(__dsub_rn(v, mn), s));
	return min(max(c, 0), PIP_G - 1);
}

__device__ __forceinline__ void pip_flush(uint32_t *stage, uint32_t cnt, uint32_t poly, uint32_t *__restrict__ out_point,
                                          uint32_t *__restrict__ out_poly, uint64_t capacity,
                                          unsigned long long *__restrict__ counter) {
	const unsigned lane = sb_lane();
	unsigned long long base = 0;
	if (lane == 0) {
		base = atomicAdd(counter, (unsigned long long)cnt);
	}
	base = __shfl_sync(0xFFFFFFFFu, base, 0);
	__syncwarp();
	for (uint32_t j = lane; j < cnt; j += 32) {
		if (base + j < capacity) {
			out_point[base + j] = stage[j];
			out_poly[base + j] = poly;
		}
	}
	__syncwarp();
}

// EXACT = 0: the native winding routine (ST_Within(POINT_2D, POLYGON_2D)). EXACT = 1: the GEOS-route predicates on
// GEOMETRY arguments (exact.cuh): the same index serves the ray-crossing count -- an edge wholly to the right of the
// point's cell is crossed iff the half-open y rule holds (its orientation sign is known), the parity of W0 equals the
// parity of the number of spanning edges, and a point can only lie ON an edge that overlaps its cell, i.e. a listed
// edge, which gets the exact evaluation. pred: SB_PRED_INTERSECTS (location != exterior) or SB_PRED_WITHIN (interior).
template <int EXACT>
__global__ void __launch_bounds__(PIP_THREADS, 3)
    k_pip_join(int pred, const pip_poly *__restrict__ polys, uint32_t n_polys, const double *__restrict__ gvx,
               const double *__restrict__ gvy, const uint64_t *__restrict__ ring_off, const uint64_t *__restrict__ poly_off,
               const bp_grid *__restrict__ gp, const uint2 *__restrict__ rec, const double2 *__restrict__ xy,
               const uint32_t *__restrict__ cell_off, uint32_t *__restrict__ ticket, uint32_t *__restrict__ out_point,
               uint32_t *__restrict__ out_poly, uint64_t capacity, unsigned long long *__restrict__ counter) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	pip_smem &S = *reinterpret_cast<pip_smem *>(smem_raw);
	const bp_grid g = *gp;
	const unsigned tid = threadIdx.x, lane = sb_lane(), warp = tid >> 5, lt = sb_lanemask_lt();
	while (true) {
		__syncthreads();
		if (tid == 0) {
			S.poly = atomicAdd(ticket, 1u);
			S.n_list = 0;
		}
		__syncthreads();
		const uint32_t p = S.poly;
		if (p >= n_polys) {
			break;
		}
		const pip_poly P = polys[p];
		if (P.flags == 0) {
			continue;
		}
		// cells of the quantiser grid under the polygon's float box (the candidate range of bp_leaf_range)
		int qx0, qy0, qx1, qy1;
		bp_leaf_range(P.fbox, g, qx0, qy0, qx1, qy1);
		const int gx0 = qx0 >> 10, gy0 = qy0 >> 10, gx1 = qx1 >> 10, gy1 = qy1 >> 10;
		if (gx0 > gx1 || gy0 > gy1) {
			continue;
		}
		const uint64_t r0 = poly_off[p];
		const uint64_t v0 = ring_off[r0];
		const uint32_t ne = (uint32_t)(ring_off[r0 + 1] - v0) - 1; // edges of the (first) ring
		bool indexed = P.flags == 2;
		const double fsx = (P.maxx - P.minx) > 0 ? (double)PIP_G / (P.maxx - P.minx) : 0.0;
		const double fsy = (P.maxy - P.miny) > 0 ? (double)PIP_G / (P.maxy - P.miny) : 0.0;
		if (indexed) {
			// ---- build the edge index ----
			for (uint32_t i = tid; i <= ne; i += PIP_THREADS) {
				S.vx[i] = gvx[v0 + i];
				S.vy[i] = gvy[v0 + i];
			}
			for (uint32_t c = tid; c < PIP_NCELL; c += PIP_THREADS) {
				S.w0[c] = 0;
				S.cnt[c] = 0;
				S.rcnt[c] = 0;
			}
			__syncthreads();
			for (uint32_t e = tid; e < ne; e += PIP_THREADS) {
				const double x1 = S.vx[e], y1 = S.vy[e], x2 = S.vx[e + 1], y2 = S.vy[e + 1];
				uchar4 rg = make_uchar4(1, 0, 0, 0); // first row > last row: the edge is in no list (zero length)
				if (EXACT || !(x1 == x2 && y1 == y2)) { // (exact mode: a zero-length edge can still carry the point)
					const double exmin = x1 < x2 ? x1 : x2, exmax = x1 < x2 ? x2 : x1;
					const double eymin = y1 < y2 ? y1 : y2, eymax = y1 < y2 ? y2 : y1;
					const double m = fmax(__dmul_ru(exmax - exmin, 0x1p-40), 0x1p-500);
					const int c0 = pip_cell(__dsub_rd(exmin, m), P.minx, fsx), c1 = pip_cell(__dadd_ru(exmax, m), P.minx, fsx);
					const int ry0 = pip_cell(eymin, P.miny, fsy), ry1 = pip_cell(eymax, P.miny, fsy);
					rg = make_uchar4((unsigned char)ry0, (unsigned char)ry1, (unsigned char)c0, (unsigned char)c1);
					const int dir = y2 > y1 ? 1 : -1;
					for (int r = ry0; r <= ry1; r++) {
						if (c0 > 0) {
							if (r > ry0 && r < ry1) {
								atomicAdd(&S.w0[r * PIP_G + c0], dir); // columns < c0 of a spanned row get `dir` (suffix sum below)
							} else {
								atomicAdd(&S.rcnt[r * PIP_G + c0], 1u); // end row: cells left of c0 apply the y rule only
							}
						}
						for (int c = c0; c <= c1; c++) {
							atomicAdd(&S.cnt[r * PIP_G + c], 1u);
						}
					}
					atomicAdd(&S.n_list, (uint32_t)((ry1 - ry0 + 1) * (c1 - c0 + 1)));
				}
				S.rng[e] = rg;
			}
			__syncthreads();
			if (S.n_list > PIP_LMAX) {
				indexed = false; // too many list entries for shared memory: literal loop for this polygon
			}
		}
		if (indexed) {
			// per row: W0[r][c] = sum of the difference entries of the columns right of c; npre[r][c] = number of row-list
			// entries whose first column is right of c; the row list is ordered by first column, descending
			if (tid < PIP_G) {
				int run = 0;
				uint32_t nrun = 0;
				for (int c = PIP_G - 1; c >= 0; c--) {
					const int d = S.w0[tid * PIP_G + c];
					const uint32_t n = S.rcnt[tid * PIP_G + c];
					S.w0[tid * PIP_G + c] = run;
					S.npre[tid * PIP_G + c] = (uint16_t)nrun;
					S.rcnt[tid * PIP_G + c] = nrun; // position of the group (row, first column c) inside the row list
					run += d;
					nrun += n;
				}
				// exclusive scan of the row totals over the 32 rows of this warp
				uint32_t inc = nrun;
#pragma unroll
				for (int sft = 1; sft < 32; sft <<= 1) {
					const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, sft);
					if (lane >= (unsigned)sft) {
						inc += o;
					}
				}
				S.rowoff[tid] = (uint16_t)(inc - nrun);
			}
			{
				// exclusive scan of cnt over 1024 cells with 256 threads (4 consecutive cells per thread)
				uint32_t c4[4], sum = 0;
#pragma unroll
				for (int k = 0; k < 4; k++) {
					c4[k] = S.cnt[tid * 4 + k];
					sum += c4[k];
				}
				uint32_t inc = sum;
#pragma unroll
				for (int sft = 1; sft < 32; sft <<= 1) {
					const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, sft);
					if (lane >= (unsigned)sft) {
						inc += o;
					}
				}
				if (lane == 31) {
					S.scan[warp] = inc;
				}
				__syncthreads();
				uint32_t wb = 0;
				for (unsigned w = 0; w < warp; w++) {
					wb += S.scan[w];
				}
				uint32_t ex = wb + inc - sum;
#pragma unroll
				for (int k = 0; k < 4; k++) {
					S.off[tid * 4 + k] = (uint16_t)ex;
					S.cnt[tid * 4 + k] = ex;
					ex += c4[k];
				}
				if (tid == PIP_THREADS - 1) {
					S.off[PIP_NCELL] = (uint16_t)ex;
				}
			}
			__syncthreads();
			for (uint32_t e = tid; e < ne; e += PIP_THREADS) {
				const uchar4 rg = S.rng[e];
				for (int r = rg.x; r <= (int)rg.y; r++) {
					if (rg.z > 0 && !(r > (int)rg.x && r < (int)rg.y)) {
						S.rowlist[S.rowoff[r] + atomicAdd(&S.rcnt[r * PIP_G + rg.z], 1u)] = (uint16_t)e;
					}
					for (int c = rg.z; c <= (int)rg.w; c++) {
						S.list[atomicAdd(&S.cnt[r * PIP_G + c], 1u)] = (uint16_t)e;
					}
				}
			}
			__syncthreads();
		}

		// ---- stream the candidate points: (row of grid cells, bucket column) segments, one per warp at a time ----
		const int nseg = (gx1 >> 5) - (gx0 >> 5) + 1;
		const int n_items = (gy1 - gy0 + 1) * nseg;
		uint32_t *stage = S.stage[warp];
		uint32_t nst = 0;
		for (int it = 0; it < n_items; it++) {
			const int gy = gy0 + it / nseg, bx = (gx0 >> 5) + it % nseg;
			const int xa = max(gx0, bx << 5), xb = min(gx1, (bx << 5) + 31);
			const uint32_t cell_a = ((uint32_t)((gy >> 5) * BP_G + bx) << 10) | ((uint32_t)(gy & 31) << 5) | (uint32_t)(xa & 31);
			const uint32_t cell_b = cell_a + (uint32_t)(xb - xa);
			const uint32_t p0 = __ldg(cell_off + cell_a), p1 = __ldg(cell_off + cell_b + 1);
			// the warps of the CTA interleave 32-point chunks of every segment (balanced to the chunk)
			for (uint32_t i0 = p0 + warp * 32; i0 < p1; i0 += 32 * PIP_WARPS) {
				const uint32_t i = i0 + lane;
				bool hit = false;
				uint32_t id = 0;
				if (i < p1) {
					const double2 q = xy[i];
					id = rec[i].y;
					const double x = q.x, y = q.y;
					const float4 pb = make_float4(sb_f64_to_f32_down(x), sb_f64_to_f32_down(y), sb_f64_to_f32_up(x), sb_f64_to_f32_up(y));
					if (sb_box_intersects(pb, P.fbox)) {
						if (indexed && EXACT) {
							const int row = pip_cell(y, P.miny, fsy);
							const int c = row * PIP_G + pip_cell(x, P.minx, fsx);
							int cr = S.w0[c]; // parity of the signed sum == parity of the number of spanning edges
							const uint16_t *rl = S.rowlist + S.rowoff[row];
							const uint32_t np = S.npre[c];
							for (uint32_t k = 0; k < np; k++) {
								const uint32_t e = rl[k];
								const double y1 = S.vy[e], y2 = S.vy[e + 1];
								cr += ((y1 > y && y2 <= y) || (y2 > y && y1 <= y)) ? 1 : 0;
							}
							bool on_edge = false;
							const uint32_t le = S.off[c + 1];
							for (uint32_t k = S.off[c]; k < le && !on_edge; k++) {
								const uint32_t e = S.list[k];
								on_edge = sb_geos_edge(x, y, S.vx[e], S.vy[e], S.vx[e + 1], S.vy[e + 1], cr);
							}
							const int loc = on_edge ? 1 : ((cr & 1) ? 2 : 0);
							hit = pred == SB_PRED_INTERSECTS ? loc != 0 : loc == 2;
						} else if (EXACT) {
							const int loc = sb_point_polygon_locate(x, y, gvx, gvy, ring_off, r0, poly_off[p + 1]);
							hit = pred == SB_PRED_INTERSECTS ? loc != 0 : loc == 2;
						} else if (indexed) {
							const int row = pip_cell(y, P.miny, fsy);
							const int c = row * PIP_G + pip_cell(x, P.minx, fsx);
							int wn = S.w0[c];
							// end-row edges wholly right of the cell: the sign of side_v is known (< 0 upward, > 0 downward,
							// never 0), only the half-open y rule of the crossing test is left
							const uint16_t *rl = S.rowlist + S.rowoff[row];
							const uint32_t np = S.npre[c];
							for (uint32_t k = 0; k < np; k++) {
								const uint32_t e = rl[k];
								const double y1 = S.vy[e], y2 = S.vy[e + 1];
								wn += (y1 < y && y <= y2) ? 1 : 0;
								wn -= (y2 <= y && y < y1) ? 1 : 0;
							}
							bool on_edge = false;
							const uint32_t le = S.off[c + 1];
							for (uint32_t k = S.off[c]; k < le; k++) {
								const uint32_t e = S.list[k];
								if (sb_pip_edge(x, y, S.vx[e], S.vy[e], S.vx[e + 1], S.vy[e + 1], wn)) {
									on_edge = true;
									break;
								}
							}
							if (on_edge) {
								// the reference leaves its loop at the first edge the point lies on: literal in-order loop
								wn = 0;
								for (uint32_t e = 0; e < ne; e++) {
									if (sb_pip_edge(x, y, S.vx[e], S.vy[e], S.vx[e + 1], S.vy[e + 1], wn)) {
										break;
									}
								}
							}
							hit = wn != 0;
						} else {
							hit = sb_pip_native(x, y, gvx, gvy, ring_off, r0, poly_off[p + 1]);
						}
					}
				}
				const uint32_t m = __ballot_sync(0xFFFFFFFFu, hit);
				if (m) {
					const uint32_t c = __popc(m);
					if (nst + c > PIP_STAGE) {
						pip_flush(stage, nst, p, out_point, out_poly, capacity, counter);
						nst = 0;
					}
					if (hit) {
						stage[nst + __popc(m & lt)] = id;
					}
					nst += c;
				}
			}
		}
		if (nst) {
			pip_flush(stage, nst, p, out_point, out_poly, capacity, counter);
		}
	}
}

int sbi_join_points_polygons(sb_context *ctx, const double *px, const double *py, uint64_t n_points, const double *vx,
                             const double *vy, const uint64_t *ring_off, const uint64_t *poly_off, uint64_t n_polys,
                             uint32_t *out_point, uint32_t *out_poly, uint64_t capacity, uint64_t *count, int exact,
                             int pred) {
	*count = 0;
	if (n_points == 0 || n_polys == 0) {
		return SB_OK;
	}
	SB_ARG(n_points <= 0xFFFFFFFFull && n_polys <= 0xFFFFFFFFull, "row ids must fit 32 bits");
	pip_poly *d_polys = nullptr;
	char *d_tree = nullptr;
	uint32_t *d_tb = nullptr, *d_ticket = nullptr;
	unsigned long long *d_counter = nullptr;
	double2 *d_xy = nullptr;
	const uint64_t tree_bytes = sizeof(sb_flat_header) + sizeof(float4) * n_polys;
	SB_TRY(ctx->scratch_t(n_polys, &d_polys));
	SB_TRY(ctx->scratch_t(tree_bytes, &d_tree));
	SB_TRY(ctx->scratch_t(4, &d_tb));
	SB_TRY(ctx->scratch_t(1, &d_ticket));
	SB_TRY(ctx->scratch_t(1, &d_counter));
	SB_TRY(ctx->scratch_t(n_points, &d_xy));
	SB_CUDA(cudaMemsetAsync(d_tb, 0xFF, 2 * sizeof(uint32_t), ctx->stream));
	SB_CUDA(cudaMemsetAsync(d_tb + 2, 0, 2 * sizeof(uint32_t), ctx->stream));
	SB_CUDA(cudaMemsetAsync(d_ticket, 0, sizeof(uint32_t), ctx->stream));
	SB_CUDA(cudaMemsetAsync(d_counter, 0, sizeof(*d_counter), ctx->stream));
	// a leaf-level-only FlatRTree view over the polygons' float boxes: the partition front end needs the tree box
	// (quantiser grid) and the leaf boxes (which buckets hold candidates)
	sb_flat tree;
	tree.device = ctx->device;
	tree.buffer = d_tree;
	tree.bytes = tree_bytes;
	tree.owns = false;
	memset(&tree.h, 0, sizeof(tree.h));
	tree.h.magic = SB_FLAT_MAGIC;
	tree.h.item_count = (uint32_t)n_polys;
	tree.h.node_size = 32;
	tree.h.nodes = (uint32_t)n_polys;
	tree.h.nlayers = 1;
	{
		sb_prof_scope ps(ctx, "pip_prepare");
		k_pip_prepare<<<sb_grid(n_polys, 8, 8), 256, 0, ctx->stream>>>(
		    vx, vy, ring_off, poly_off, n_polys, d_polys, reinterpret_cast<float4 *>(d_tree + sizeof(sb_flat_header)), d_tb);
		SB_LAUNCH_CHECK(ctx);
		k_pip_header<<<1, 32, 0, ctx->stream>>>(d_tb, (uint32_t)n_polys, reinterpret_cast<sb_flat_header *>(d_tree));
		SB_LAUNCH_CHECK(ctx);
	}
	sb_cell_partition part;
	SB_TRY(sbi_cell_partition(ctx, &tree, px, py, n_points, 0, &part, /*want_xy=*/true));
	if (part.n_units == 0) {
		return SB_OK;
	}
	SB_SMEM_OPTIN(ctx, k_pip_join<0>, sizeof(pip_smem));
	SB_SMEM_OPTIN(ctx, k_pip_join<1>, sizeof(pip_smem));
	{
		sb_prof_scope ps(ctx, "pip_gather");
		k_pip_gather<<<sb_grid(BP_UNIT_TILES * part.n_units, 1, 4), 512, 0, ctx->stream>>>(
		    static_cast<const bp_tables *>(part.tables), static_cast<const bp_unit *>(part.units), part.src,
		    static_cast<const double2 *>(part.xy_bucketed), d_xy);
		SB_LAUNCH_CHECK(ctx);
	}
	{
		sb_prof_scope ps(ctx, "pip_join");
		const unsigned grid = (unsigned)(n_polys < 3 * SB_NUM_SMS ? n_polys : 3 * SB_NUM_SMS);
		if (exact) {
			k_pip_join<1><<<grid, PIP_THREADS, sizeof(pip_smem), ctx->stream>>>(
			    pred, d_polys, (uint32_t)n_polys, vx, vy, ring_off, poly_off, static_cast<const bp_grid *>(part.grid),
			    static_cast<const uint2 *>(part.rec), d_xy, part.cell_off, d_ticket, out_point, out_poly, capacity, d_counter);
		} else {
			k_pip_join<0><<<grid, PIP_THREADS, sizeof(pip_smem), ctx->stream>>>(
			    pred, d_polys, (uint32_t)n_polys, vx, vy, ring_off, poly_off, static_cast<const bp_grid *>(part.grid),
			    static_cast<const uint2 *>(part.rec), d_xy, part.cell_off, d_ticket, out_point, out_poly, capacity, d_counter);
		}
		SB_LAUNCH_CHECK(ctx);
	}
	SB_TRY(sbi_read_u64(ctx, reinterpret_cast<const uint64_t *>(d_counter), count));
	if (*count > capacity) {
		sb_set_error("polygon join: %llu pairs, capacity %llu", (unsigned long long)*count, (unsigned long long)capacity);
		return SB_ERR_CAPACITY;
	}
	return SB_OK;
}

// float boxes of the polygons (the cached bbox of their blobs) for the candidate-list route
int sbi_pip_boxes(sb_context *ctx, const double *vx, const double *vy, const uint64_t *ring_off, const uint64_t *poly_off,
                  uint64_t n_polys, float *out_box4) {
	pip_poly *d_polys = nullptr;
	uint32_t *d_tb = nullptr;
	SB_TRY(ctx->scratch_t(n_polys, &d_polys));
	SB_TRY(ctx->scratch_t(4, &d_tb));
	SB_CUDA(cudaMemsetAsync(d_tb, 0, 4 * sizeof(uint32_t), ctx->stream));
	sb_prof_scope ps(ctx, "pip_prepare");
	k_pip_prepare<<<sb_grid(n_polys, 8, 8), 256, 0, ctx->stream>>>(vx, vy, ring_off, poly_off, n_polys, d_polys,
	                                                               reinterpret_cast<float4 *>(out_box4), d_tb);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}
