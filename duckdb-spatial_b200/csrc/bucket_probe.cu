// bucket_probe.cu -- K5 v2: the bulk point-probe of the spatial join (points x FlatRTree), bucketed.
//
// Replaces the per-row loop of PhysicalSpatialJoin::ExecuteInternal (spatial_join_physical.cpp:748-849:
// TryGetCachedBounds -> FlatRTree::InitScan/Scan per probe row -> exact predicate) for a whole probe column.
// The candidate set of a probe row is fixed by the reference: every leaf entry whose float box intersects the
// probe's float box (a leaf hit implies all its ancestors hit, because parent boxes are exact fmin/fmax unions).
// v1 (flat_rtree.cu) walks the tree per probe and is instruction- and L2-bound (0.8 % of the HBM roofline:
// profiles/README.md r1a). v2 makes the probes coherent first and then shares all per-node work between them:
//
//   1. a monotone 20-bit quantiser Q(v) = floor((v - tree_min) * 2^20 / tree_extent) per axis slices the tree box
//      into 32x32 buckets (Q >> 15) of 32x32 sub-cells ((Q >> 10) & 31) with a 15-bit in-bucket coordinate
//      (Q & 32767). Because Q is monotone, a <= x <= c implies Q(a) <= Q(x) <= Q(c): quantised comparisons are
//      conservative, equality is the only ambiguous case, and it is resolved with the exact predicate.
//   2. k_bp_hist / k_bp_scatter partition the probe points into the 1024 buckets (one pass over the points,
//      16 B read + 20 B written per point). Points whose float box misses the tree box are dropped here
//      (the root test of FlatRTree::Scan).
//   3. k_bp_cand_* bin the leaf level of the packed tree (Hilbert order, 16 B boxes) into the same buckets.
//   4. k_bp_probe: one CTA per (bucket, slice of its points): stages the bucket's leaf entries in shared memory
//      as 8-byte quantised boxes, builds a 32x32 sub-cell directory (CSR) over them in shared memory, then every
//      thread streams points: look up the sub-cell list (~18 entries for config C3 instead of ~200 box tests of a
//      root-to-leaf walk), test with 4 integer instructions per entry, resolve quantisation ties with the exact
//      fp64 (or float-box) predicate, and emit pairs through a shared-memory staging buffer so that the global
//      writes are coalesced and the global pair counter sees one atomic per ~4000 pairs.
//
// Algorithmic DRAM bytes per probe: 16 (coordinates) + 8 per emitted pair; the partition adds 36 B of real
// traffic per probe.
#include "internal.cuh"

#include "cells.cuh"

__global__ void k_bp_grid_init(const sb_flat_header *__restrict__ hdr, bp_grid *__restrict__ g) {
	if (threadIdx.x == 0) {
		const double x0 = hdr->tree_box[0], y0 = hdr->tree_box[1], x1 = hdr->tree_box[2], y1 = hdr->tree_box[3];
		const double w = x1 - x0, h = y1 - y0;
		g->minx = x0;
		g->miny = y0;
		g->sx = (w > 0 && w < 1.7e308) ? (double)(1 << BP_QBITS) / w : 0.0;
		g->sy = (h > 0 && h < 1.7e308) ? (double)(1 << BP_QBITS) / h : 0.0;
		g->tb[0] = hdr->tree_box[0], g->tb[1] = hdr->tree_box[1], g->tb[2] = hdr->tree_box[2], g->tb[3] = hdr->tree_box[3];
	}
}

// bucket of a probe point: -1 = its float box misses the tree box (no candidates), BP_NB = NaN coordinate
__device__ __forceinline__ int bp_bucket(double x, double y, const bp_grid &g) {
	if (x != x || y != y) {
		return BP_NB;
	}
	const float4 q = make_float4(sb_f64_to_f32_down(x), sb_f64_to_f32_down(y), sb_f64_to_f32_up(x), sb_f64_to_f32_up(y));
	if (!sb_box_intersects(q, make_float4(g.tb[0], g.tb[1], g.tb[2], g.tb[3]))) {
		return -1;
	}
	const int qx = bp_q(x, g.minx, g.sx), qy = bp_q(y, g.miny, g.sy);
	return (qy >> 15) * BP_G + (qx >> 15);
}

// ---- 2. partition of the probe points ------------------------------------------------------------------------
#define BPH_THREADS 512
__global__ void __launch_bounds__(BPH_THREADS) k_bp_hist(const double *__restrict__ px, const double *__restrict__ py,
                                                         uint64_t n, const bp_grid *__restrict__ gp,
                                                         uint32_t *__restrict__ hist, uint32_t *__restrict__ fine) {
	__shared__ uint32_t s_hist[BP_NB + 1];
	for (int i = threadIdx.x; i <= BP_NB; i += BPH_THREADS) {
		s_hist[i] = 0;
	}
	const bp_grid g = *gp;
	__syncthreads();
	const uint64_t stride = (uint64_t)gridDim.x * BPH_THREADS;
	for (uint64_t i = (uint64_t)blockIdx.x * BPH_THREADS + threadIdx.x; i < n; i += stride) {
		const double x = px[i], y = py[i];
		const int b = bp_bucket(x, y, g);
		if (b >= 0) {
			atomicAdd(&s_hist[b], 1u);
			if (fine && b < BP_NB) {
				const int qx = bp_q(x, g.minx, g.sx), qy = bp_q(y, g.miny, g.sy);
				atomicAdd(&fine[(b << 10) | (((qy >> 10) & 31) << 5) | ((qx >> 10) & 31)], 1u);
			}
		}
	}
	__syncthreads();
	for (int i = threadIdx.x; i <= BP_NB; i += BPH_THREADS) {
		const uint32_t c = s_hist[i];
		if (c) {
			atomicAdd(&hist[i], c);
		}
	}
}

// One tile = 4096 points: rank inside the tile's bucket with shared atomics, reserve the tile's share of every
// bucket with one global atomic per non-empty bucket, reorder the tile by bucket in shared memory, and write it out
// so that consecutive threads write consecutive addresses (runs of one bucket): whole sectors, no partial-sector
// read-modify-write in L2 (the v1 of this kernel scattered 8/16-byte stores: 396 M partial sectors, 9.7 ms).
#define BPS_THREADS 512
#define BPS_ITEMS 8
#define BPS_TILE (BPS_THREADS * BPS_ITEMS)
struct bps_smem {
	double2 xy[BPS_TILE];
	uint32_t idx[BPS_TILE];
	uint16_t bkt[BPS_TILE];
	uint32_t cnt[BP_NB + 8];   // counts, then (global base - tile offset) per bucket
	uint32_t off[BP_NB + 8];   // tile-local exclusive offsets
	uint32_t scan[16];
};
__global__ void __launch_bounds__(BPS_THREADS, 2)
    k_bp_scatter(const double *__restrict__ px, const double *__restrict__ py, uint64_t n, uint32_t index_base,
                 const bp_grid *__restrict__ gp, uint32_t *__restrict__ cursor, double2 *__restrict__ oxy,
                 uint32_t *__restrict__ oidx) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	bps_smem &S = *reinterpret_cast<bps_smem *>(smem_raw);
	const bp_grid g = *gp;
	const unsigned tid = threadIdx.x, lane = sb_lane(), warp = tid >> 5;
	const uint64_t tiles = (n + BPS_TILE - 1) / BPS_TILE;
	for (uint64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
		for (int i = tid; i <= BP_NB; i += BPS_THREADS) {
			S.cnt[i] = 0;
		}
		__syncthreads();
		double x[BPS_ITEMS], y[BPS_ITEMS];
		int b[BPS_ITEMS];
		uint32_t r[BPS_ITEMS];
		const uint64_t base = tile * BPS_TILE + tid;
#pragma unroll
		for (int k = 0; k < BPS_ITEMS; k++) {
			const uint64_t i = base + (uint64_t)k * BPS_THREADS;
			b[k] = -1;
			if (i < n) {
				x[k] = px[i];
				y[k] = py[i];
			}
		}
#pragma unroll
		for (int k = 0; k < BPS_ITEMS; k++) {
			if (base + (uint64_t)k * BPS_THREADS < n) {
				b[k] = bp_bucket(x[k], y[k], g);
			}
			r[k] = b[k] >= 0 ? atomicAdd(&S.cnt[b[k]], 1u) : 0u;
		}
		__syncthreads();
		// exclusive scan over the 1025 bins (2 per thread + the NaN bin) and the global reservations
		{
			const uint32_t c0 = S.cnt[2 * tid], c1 = S.cnt[2 * tid + 1];
			uint32_t inc = c0 + c1;
#pragma unroll
			for (int s = 1; s < 32; s <<= 1) {
				const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
				if (lane >= (unsigned)s) {
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
			const uint32_t e0 = wb + inc - (c0 + c1), e1 = e0 + c0;
			S.off[2 * tid] = e0;
			S.off[2 * tid + 1] = e1;
			if (c0) {
				S.cnt[2 * tid] = atomicAdd(&cursor[2 * tid], c0) - e0;
			}
			if (c1) {
				S.cnt[2 * tid + 1] = atomicAdd(&cursor[2 * tid + 1], c1) - e1;
			}
			if (tid == BPS_THREADS - 1) {
				const uint32_t cn = S.cnt[BP_NB], en = e1 + c1;
				S.off[BP_NB] = en;
				S.off[BP_NB + 1] = en + cn;
				if (cn) {
					S.cnt[BP_NB] = atomicAdd(&cursor[BP_NB], cn) - en;
				}
			}
		}
		__syncthreads();
#pragma unroll
		for (int k = 0; k < BPS_ITEMS; k++) {
			if (b[k] >= 0) {
				const uint32_t slot = S.off[b[k]] + r[k];
				S.xy[slot] = make_double2(x[k], y[k]);
				S.idx[slot] = (uint32_t)(base + (uint64_t)k * BPS_THREADS) + index_base;
				S.bkt[slot] = (uint16_t)b[k];
			}
		}
		__syncthreads();
		const uint32_t kept = S.off[BP_NB + 1];
		for (uint32_t t = tid; t < kept; t += BPS_THREADS) {
			const uint32_t dst = S.cnt[S.bkt[t]] + t;
			oxy[dst] = S.xy[t];
			oidx[dst] = S.idx[t];
		}
		__syncthreads();
	}
}

// ---- 3. the tree's leaf level binned into the same buckets ----------------------------------------------------
// The leaf level is in Hilbert order, so the 32 leaves of a warp almost always fall into the same bucket: one atomic
// per group of equal buckets (match.any) instead of one per leaf -- the per-leaf version spent 0.3 ms per launch at C3
// on same-address atomics. Leaves that straddle buckets take the plain loop.
template <bool FILL>
__global__ void __launch_bounds__(256) k_bp_cand(const float4 *__restrict__ boxes, uint32_t items,
                                                 const bp_grid *__restrict__ gp, uint32_t *__restrict__ counter,
                                                 uint32_t *__restrict__ cand_pos) {
	const bp_grid g = *gp;
	const uint32_t stride = gridDim.x * blockDim.x;
	const unsigned lane = sb_lane(), lt = sb_lanemask_lt();
	const uint32_t items_round = (items + 31u) & ~31u;
	for (uint32_t pos = blockIdx.x * blockDim.x + threadIdx.x; pos < items_round; pos += stride) {
		int x0 = 1, y0 = 1, x1 = 0, y1 = 0;
		if (pos < items) {
			bp_leaf_range(boxes[pos], g, x0, y0, x1, y1);
		}
		const int bx0 = x0 >> 15, bx1 = x1 >> 15, by0 = y0 >> 15, by1 = y1 >> 15;
		const bool live = pos < items && x0 <= x1 && y0 <= y1;
		const bool single = live && bx0 == bx1 && by0 == by1;
		const uint32_t key = single ? (uint32_t)(by0 * BP_G + bx0) : 0xFFFFFFFFu - lane; // non-single lanes: unique keys
		const uint32_t peers = __match_any_sync(0xFFFFFFFFu, key);
		if (single) {
			const int leader = __ffs(peers) - 1;
			uint32_t base = 0;
			if ((int)lane == leader) {
				base = atomicAdd(&counter[key], (uint32_t)__popc(peers));
			}
			base = __shfl_sync(peers, base, leader);
			if (FILL) {
				cand_pos[base + __popc(peers & lt)] = pos;
			}
		} else if (live) {
			for (int by = by0; by <= by1; by++) {
				for (int bx = bx0; bx <= bx1; bx++) {
					const uint32_t slot = atomicAdd(&counter[by * BP_G + bx], 1u);
					if (FILL) {
						cand_pos[slot] = pos;
					}
				}
			}
		}
	}
}

// exclusive scan over the CTA (WARPS x 32 threads, WARPS <= 32)
template <int WARPS>
__device__ __forceinline__ uint32_t bp_block_excl_scan(uint32_t v, uint32_t *s_warp, uint32_t *total) {
	const unsigned lane = sb_lane(), warp = threadIdx.x >> 5;
	uint32_t inc = v;
#pragma unroll
	for (int s = 1; s < 32; s <<= 1) {
		const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
		if (lane >= (unsigned)s) {
			inc += o;
		}
	}
	__syncthreads();
	if (lane == 31) {
		s_warp[warp] = inc;
	}
	__syncthreads();
	if (warp == 0) {
		uint32_t w = lane < WARPS ? s_warp[lane] : 0u;
#pragma unroll
		for (int s = 1; s < 32; s <<= 1) {
			const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, w, s);
			if (lane >= (unsigned)s) {
				w += o;
			}
		}
		s_warp[lane] = w; // inclusive over warps
	}
	__syncthreads();
	const uint32_t base = warp ? s_warp[warp - 1] : 0u;
	*total = s_warp[31];
	return base + inc - v;
}

__device__ __forceinline__ uint32_t bp_block_excl_scan_1024(uint32_t v, uint32_t *s_warp, uint32_t *total) {
	// blockDim.x == 1024
	const unsigned lane = sb_lane(), warp = threadIdx.x >> 5;
	uint32_t inc = v;
#pragma unroll
	for (int s = 1; s < 32; s <<= 1) {
		const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
		if (lane >= (unsigned)s) {
			inc += o;
		}
	}
	__syncthreads();
	if (lane == 31) {
		s_warp[warp] = inc;
	}
	__syncthreads();
	if (warp == 0) {
		uint32_t w = s_warp[lane];
#pragma unroll
		for (int s = 1; s < 32; s <<= 1) {
			const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, w, s);
			if (lane >= (unsigned)s) {
				w += o;
			}
		}
		s_warp[lane] = w; // inclusive over warps
	}
	__syncthreads();
	const uint32_t base = warp ? s_warp[warp - 1] : 0u;
	*total = s_warp[31];
	return base + inc - v;
}

__global__ void __launch_bounds__(1024) k_bp_scan(bp_tables *__restrict__ T, bp_unit *__restrict__ units) {
	__shared__ uint32_t s_warp[32];
	const unsigned b = threadIdx.x;
	uint32_t total;
	// points (the NaN bucket goes last)
	const uint32_t np = T->pt_hist[b];
	const uint32_t po = bp_block_excl_scan_1024(np, s_warp, &total);
	T->pt_off[b] = po;
	T->pt_cursor[b] = po;
	if (b == 0) {
		const uint32_t nan = T->pt_hist[BP_NB];
		T->pt_off[BP_NB] = total;
		T->pt_cursor[BP_NB] = total;
		T->pt_off[BP_NB + 1] = total + nan;
		T->totals[0] = total;
		T->totals[3] = nan;
	}
	// candidates
	const uint32_t nc = T->cand_cnt[b];
	const uint32_t co = bp_block_excl_scan_1024(nc, s_warp, &total);
	T->cand_off[b] = co;
	T->cand_cursor[b] = co;
	if (b == 0) {
		T->cand_off[BP_NB] = total;
		T->totals[1] = total;
	}
	// work units: slices of at most BP_SLICE points of a bucket that has both points and candidates
	const uint32_t nu = (np && nc) ? (np + BP_SLICE - 1) / BP_SLICE : 0u;
	const uint32_t uo = bp_block_excl_scan_1024(nu, s_warp, &total);
	for (uint32_t k = 0; k < nu; k++) {
		bp_unit u;
		u.bucket = b;
		u.pbeg = po + k * BP_SLICE;
		u.pend = min(po + np, u.pbeg + BP_SLICE);
		units[uo + k] = u;
	}
	if (b == 0) {
		T->totals[2] = total;
		T->ticket = 0;
	}
}

// ---- 4. the probe kernel ---------------------------------------------------------------------------------------
#define BPP_THREADS 1024
#define BPP_CMAX 4096   // staged leaf entries per chunk
#define BPP_EMAX 40960  // sub-cell directory entries per chunk (u16 offsets => < 65536)
#define BPP_WMAX 512    // entries covering more than BPP_WIDE_T sub-cells: tested by every point instead
#define BPP_WIDE_T 64
#define BPP_C 0x80008000u
#define BPP_K 0x00010001u

// MODE 1: candidate semantics (float boxes, Box::Intersects). MODE 2: fused exact predicate on the build's BOX_2D.
#define BPP_GROUPS 4                      // independent 256-thread groups per CTA (named barriers 1..4)
#define BPP_GTHREADS (BPP_THREADS / BPP_GROUPS)
template <int MODE>
struct bpp_smem {
	// pairs staged per group and round before the coalesced flush (a round = 256 points, ~2200 pairs at C3)
	static constexpr int STAGE = MODE == 1 ? 2304 : 2816;
	uint2 qbox[BPP_CMAX];     // x = (ly0 << 16) | lx0, y = ((ly1 << 16) | lx1) | BPP_C
	uint32_t cidx[BPP_CMAX];  // build-side index (FlatRTree idx_array value)
	uint32_t cpos[MODE == 1 ? BPP_CMAX : 1]; // position in the tree's leaf level (float box for the tie test)
	uint32_t cnt[BP_NB];      // sub-cell counts, then fill cursors
	uint16_t off[BP_NB + 8];  // CSR offsets
	uint16_t ent[BPP_EMAX];   // CSR entries: staged-entry numbers
	uint16_t wide[BPP_WMAX];
	// per sub-cell and 2x2 micro-cell: x = entries (of the first 32 of the list) that contain the whole micro-cell
	// strictly (every point of it is an exact hit), y = entries that only overlap it (tested per point)
	uint2 mask[BP_NB * 4];
	// staged pairs, 4 bytes each: (thread of the group << 16) | staged-entry number; the copy-out expands them
	// through rid[] (probe row id of each thread in this round) and cidx[]
	uint32_t stage[BPP_GROUPS][STAGE];
	uint32_t rid[BPP_GROUPS][BPP_GTHREADS];
	uint32_t scan[32];
	uint32_t unit, n_ent, n_wide;
	uint32_t n_stage[BPP_GROUPS], stage_limit[BPP_GROUPS];
	unsigned long long gbase[BPP_GROUPS];
};

// barriers of one 256-thread group: the four groups of a CTA stream their points independently, so that one group's
// flush (global atomic round trip + stores) overlaps the other groups' tests
__device__ __forceinline__ void bpp_gsync(unsigned gid) {
	asm volatile("bar.sync %0, %1;" ::"r"(gid + 1), "n"(BPP_GTHREADS) : "memory");
}
__device__ __forceinline__ bool bpp_gor(unsigned gid, bool p) {
	unsigned r;
	asm volatile("{\n\t.reg .pred q, r;\n\tsetp.ne.u32 q, %2, 0;\n\tbar.red.or.pred r, %1, %3, q;\n\tselp.u32 %0, 1, 0, r;\n\t}"
	             : "=r"(r)
	             : "r"(gid + 1), "r"((unsigned)p), "n"(BPP_GTHREADS)
	             : "memory");
	return r != 0;
}

template <int MODE>
__device__ __forceinline__ bool bp_exact(double x, double y, uint32_t pos, uint32_t bi, const float4 *__restrict__ boxes,
                                         const double4 *__restrict__ build_box2d, int pred) {
	if (MODE == 1) {
		const float4 q = make_float4(sb_f64_to_f32_down(x), sb_f64_to_f32_down(y), sb_f64_to_f32_up(x), sb_f64_to_f32_up(y));
		return sb_box_intersects(q, __ldg(boxes + pos));
	} else {
		const double4 b = build_box2d[bi];
		return pred == SB_PRED_INTERSECTS ? (b.x <= x && x <= b.z && b.y <= y && y <= b.w)
		                                  : (b.x < x && x < b.z && b.y < y && y < b.w);
	}
}

// Three-state test of a point (PqC = packed 15-bit in-bucket coordinates | BPP_C) against a quantised box.
// Each 16-bit half of t1 / t2 is 0x8000 + (p - lo) resp. 0x8000 + (hi - p), always in [1, 0xFFFF] (no borrow
// between halves): bit 15 set <=> that side holds in the closed sense; >= 0x8001 <=> it holds strictly.
// Strictly inside on all four sides => the exact predicate is true; outside on any side => false;
// otherwise (a tie in quantised space) the exact predicate decides.
#define BPP_TEST(j, bit, mask, tmask)                                                                                \
	do {                                                                                                             \
		const uint2 qb_ = S.qbox[j];                                                                                 \
		const uint32_t t1_ = PqC - qb_.x, t2_ = qb_.y - Pq;                                                          \
		if (((t1_ - BPP_K) & (t2_ - BPP_K) & BPP_C) == BPP_C) {                                                      \
			mask |= (bit);                                                                                           \
		} else if ((t1_ & t2_ & BPP_C) == BPP_C) {                                                                   \
			tmask |= (bit);                                                                                          \
		}                                                                                                            \
	} while (0)

// Flush protocol of one pass (per 256-thread group): every thread holds a 64-bit hit mask over its entry list `ep`.
// Warp scan of the hit counts -> one shared atomic per warp reserves staging space -> pairs go to the group's staging
// arrays -> one global atomic per group reserves output space -> copy-out with 16-byte stores. A warp whose pairs do
// not fit the staging buffer (dense inputs) writes them straight to global memory.
template <int MODE>
__device__ __forceinline__ void bpp_emit(bpp_smem<MODE> &S, unsigned gid, uint32_t mlo, uint32_t mhi, const uint16_t *ep,
                                         uint32_t id, uint32_t *__restrict__ out_probe, uint32_t *__restrict__ out_build,
                                         uint64_t capacity, unsigned long long *__restrict__ counter) {
	constexpr uint32_t STAGE = bpp_smem<MODE>::STAGE;
	const unsigned gtid = threadIdx.x & (BPP_GTHREADS - 1), lane = sb_lane();
	uint32_t *st = S.stage[gid];
	const uint32_t tag = gtid << 16;
	const uint32_t cntk = (uint32_t)(__popc(mlo) + __popc(mhi));
	uint32_t inc = cntk;
#pragma unroll
	for (int s = 1; s < 32; s <<= 1) {
		const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
		if (lane >= (unsigned)s) {
			inc += o;
		}
	}
	const uint32_t wtotal = __shfl_sync(0xFFFFFFFFu, inc, 31);
	uint32_t wbase = 0;
	if (lane == 31 && wtotal) {
		wbase = atomicAdd(&S.n_stage[gid], wtotal);
	}
	wbase = __shfl_sync(0xFFFFFFFFu, wbase, 31);
	if (wbase + wtotal <= STAGE) {
		uint32_t d = wbase + (inc - cntk);
		while (mlo) {
			const int k = 31 - __clz(mlo);
			mlo ^= 1u << k;
			st[d++] = tag | ep[k];
		}
		while (mhi) {
			const int k = 31 - __clz(mhi);
			mhi ^= 1u << k;
			st[d++] = tag | ep[32 + k];
		}
	} else if (wtotal) {
		if (lane == 31) {
			atomicMin(&S.stage_limit[gid], wbase);
		}
		unsigned long long gb = 0;
		if (lane == 31) {
			gb = atomicAdd(counter, (unsigned long long)wtotal);
		}
		gb = __shfl_sync(0xFFFFFFFFu, gb, 31) + (inc - cntk);
		while (mlo | mhi) {
			int k;
			if (mlo) {
				k = 31 - __clz(mlo);
				mlo ^= 1u << k;
			} else {
				k = 31 - __clz(mhi);
				mhi ^= 1u << k;
				k += 32;
			}
			if (gb < capacity) {
				out_probe[gb] = id;
				out_build[gb] = S.cidx[ep[k]];
			}
			gb++;
		}
	}
	bpp_gsync(gid);
	const uint32_t nst = min(S.n_stage[gid], S.stage_limit[gid]);
	if (gtid == 0) {
		S.gbase[gid] = nst ? atomicAdd(counter, (unsigned long long)nst) : 0ull;
	}
	bpp_gsync(gid);
	const unsigned long long gbase = S.gbase[gid];
	uint32_t *op = out_probe + gbase, *ob = out_build + gbase;
	// copy-out: one conflict-free shared load per pair, expansion through rid[] / cidx[], two coalesced 4-byte stores
	// (the shared-memory pipe, not the instruction count, limits this kernel: profiles/README.md r1e)
	const uint32_t *rid = S.rid[gid];
	if (gbase + nst <= capacity) {
		for (uint32_t t = gtid; t < nst; t += BPP_GTHREADS) {
			const uint32_t v = st[t];
			op[t] = rid[v >> 16];
			ob[t] = S.cidx[v & 0xFFFFu];
		}
	} else {
		for (uint32_t t = gtid; t < nst; t += BPP_GTHREADS) {
			if (gbase + t < capacity) {
				const uint32_t v = st[t];
				op[t] = rid[v >> 16];
				ob[t] = S.cidx[v & 0xFFFFu];
			}
		}
	}
	bpp_gsync(gid);
	if (gtid == 0) {
		S.n_stage[gid] = 0;
		S.stage_limit[gid] = 0xFFFFFFFFu;
	}
}

#define BPP_RESOLVE_TIES(ep, tlo_, thi_)                                                                             \
	while (tlo_ | thi_) {                                                                                            \
		int k_;                                                                                                      \
		if (tlo_) {                                                                                                  \
			k_ = 31 - __clz(tlo_);                                                                                   \
			tlo_ ^= 1u << k_;                                                                                        \
		} else {                                                                                                     \
			k_ = 31 - __clz(thi_);                                                                                   \
			thi_ ^= 1u << k_;                                                                                        \
			k_ += 32;                                                                                                \
		}                                                                                                            \
		const uint32_t j_ = (ep)[k_];                                                                                \
		if (bp_exact<MODE>(x, y, MODE == 1 ? S.cpos[j_] : 0u, S.cidx[j_], boxes, build_box2d, pred)) {               \
			if (k_ < 32) {                                                                                           \
				mlo |= 1u << k_;                                                                                     \
			} else {                                                                                                 \
				mhi |= 1u << (k_ - 32);                                                                              \
			}                                                                                                        \
		}                                                                                                            \
	}

// first pass over a sub-cell list with the micro-cell masks: `mk.x` are hits without any test, only the entries in
// `mk.y` are tested; entries 32..63 of long lists are tested one by one
#define BPP_PASS_MASKED(ep, n, mk)                                                                                   \
	do {                                                                                                             \
		uint32_t tlo_ = 0, thi_ = 0;                                                                                 \
		uint32_t amb_ = (mk).y;                                                                                      \
		mlo = (mk).x;                                                                                                \
		while (amb_) {                                                                                               \
			const int k_ = 31 - __clz(amb_);                                                                         \
			const uint32_t bit_ = 1u << k_;                                                                          \
			amb_ ^= bit_;                                                                                            \
			BPP_TEST((ep)[k_], bit_, mlo, tlo_);                                                                     \
		}                                                                                                            \
		if ((n) > 32u) {                                                                                             \
			uint32_t bit_ = 1;                                                                                       \
			for (uint32_t k_ = 32; k_ < (uint32_t)(n); k_++, bit_ <<= 1) {                                           \
				BPP_TEST((ep)[k_], bit_, mhi, thi_);                                                                 \
			}                                                                                                        \
		}                                                                                                            \
		BPP_RESOLVE_TIES(ep, tlo_, thi_);                                                                            \
	} while (0)

// tests up to 64 entries ep[0..n) for one point; ties are resolved with the exact predicate
#define BPP_PASS(ep, n)                                                                                              \
	do {                                                                                                             \
		uint32_t tlo_ = 0, thi_ = 0;                                                                                 \
		const uint32_t n0_ = min((uint32_t)(n), 32u);                                                                \
		uint32_t bit_ = 1;                                                                                           \
		for (uint32_t k_ = 0; k_ < n0_; k_++, bit_ <<= 1) {                                                          \
			BPP_TEST((ep)[k_], bit_, mlo, tlo_);                                                                     \
		}                                                                                                            \
		if ((n) > 32u) {                                                                                             \
			bit_ = 1;                                                                                                \
			for (uint32_t k_ = 32; k_ < (uint32_t)(n); k_++, bit_ <<= 1) {                                           \
				BPP_TEST((ep)[k_], bit_, mhi, thi_);                                                                 \
			}                                                                                                        \
		}                                                                                                            \
		BPP_RESOLVE_TIES(ep, tlo_, thi_);                                                                            \
	} while (0)

template <int MODE>
__global__ void __launch_bounds__(BPP_THREADS, 1)
    k_bp_probe(const float4 *__restrict__ boxes, const uint32_t *__restrict__ idx,
               const double4 *__restrict__ build_box2d, int pred, const bp_grid *__restrict__ gp,
               bp_tables *__restrict__ T, const bp_unit *__restrict__ units, const uint32_t *__restrict__ cand_pos,
               const double2 *__restrict__ pxy, const uint32_t *__restrict__ pidx, uint32_t *__restrict__ out_probe,
               uint32_t *__restrict__ out_build, uint64_t capacity, unsigned long long *__restrict__ counter) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	bpp_smem<MODE> &S = *reinterpret_cast<bpp_smem<MODE> *>(smem_raw);
	const bp_grid g = *gp;
	const unsigned tid = threadIdx.x;
	const uint32_t n_units = (uint32_t)T->totals[2];

	while (true) {
		__syncthreads();
		if (tid == 0) {
			S.unit = atomicAdd(&T->ticket, 1u);
		}
		__syncthreads();
		const uint32_t u = S.unit;
		if (u >= n_units) {
			break;
		}
		const bp_unit unit = units[u];
		const int ox = (int)(unit.bucket % BP_G) << 15, oy = (int)(unit.bucket / BP_G) << 15;
		const uint32_t cbeg = T->cand_off[unit.bucket], cend = T->cand_off[unit.bucket + 1];
		uint32_t c0 = cbeg;
		while (c0 < cend) {
			// ---- (a) stage a chunk of leaf entries as quantised in-bucket boxes, count directory entries ----
			uint32_t ntry = min((uint32_t)BPP_CMAX, cend - c0);
			while (true) {
				__syncthreads();
				if (tid == 0) {
					S.n_ent = 0;
					S.n_wide = 0;
				}
				if (tid < BPP_GROUPS) {
					S.n_stage[tid] = 0;
					S.stage_limit[tid] = 0xFFFFFFFFu;
				}
				S.cnt[tid] = 0; // BPP_THREADS == BP_NB
				__syncthreads();
				for (uint32_t j = tid; j < ntry; j += BPP_THREADS) {
					const uint32_t pos = cand_pos[c0 + j];
					const uint32_t bi = __ldg(idx + pos);
					int x0, y0, x1, y1;
					bool empty = false;
					if (MODE == 1) {
						bp_leaf_range(__ldg(boxes + pos), g, x0, y0, x1, y1);
						S.cpos[j] = pos;
					} else {
						const double4 e = build_box2d[bi];
						empty = (e.x != e.x) || (e.y != e.y) || (e.z != e.z) || (e.w != e.w);
						x0 = bp_q(e.x, g.minx, g.sx), y0 = bp_q(e.y, g.miny, g.sy);
						x1 = bp_q(e.z, g.minx, g.sx), y1 = bp_q(e.w, g.miny, g.sy);
					}
					empty = empty || x1 < ox || x0 > ox + 32767 || y1 < oy || y0 > oy + 32767 || x0 > x1 || y0 > y1;
					const uint32_t lx0 = (uint32_t)(max(x0, ox) - ox), ly0 = (uint32_t)(max(y0, oy) - oy);
					const uint32_t lx1 = (uint32_t)(min(x1, ox + 32767) - ox), ly1 = (uint32_t)(min(y1, oy + 32767) - oy);
					uint2 qb;
					qb.x = empty ? 0x7FFF7FFFu : ((ly0 << 16) | lx0);
					qb.y = (empty ? 0u : ((ly1 << 16) | lx1)) | BPP_C;
					S.qbox[j] = qb;
					S.cidx[j] = bi;
					if (!empty) {
						const uint32_t cells = ((lx1 >> 10) - (lx0 >> 10) + 1) * ((ly1 >> 10) - (ly0 >> 10) + 1);
						if (cells > BPP_WIDE_T) {
							atomicAdd(&S.n_wide, 1u);
						} else {
							atomicAdd(&S.n_ent, cells);
							for (uint32_t cy = ly0 >> 10; cy <= (ly1 >> 10); cy++) {
								for (uint32_t cx = lx0 >> 10; cx <= (lx1 >> 10); cx++) {
									atomicAdd(&S.cnt[cy * 32 + cx], 1u);
								}
							}
						}
					}
				}
				__syncthreads();
				if (S.n_ent <= BPP_EMAX && S.n_wide <= BPP_WMAX) {
					break;
				}
				ntry = (ntry + 1) / 2; // rare: the directory would overflow shared memory -> smaller chunk
			}
			// ---- (b) CSR offsets over the 1024 sub-cells (one per thread) ----
			{
				uint32_t total;
				const uint32_t c = S.cnt[tid];
				const uint32_t ex = bp_block_excl_scan_1024(c, S.scan, &total);
				S.off[tid] = (uint16_t)ex;
				S.cnt[tid] = ex;
				if (tid == BPP_THREADS - 1) {
					S.off[BP_NB] = (uint16_t)(ex + c);
				}
				if (tid == 0) {
					S.n_wide = 0;
				}
			}
			__syncthreads();
			// ---- (c) fill the directory ----
			for (uint32_t j = tid; j < ntry; j += BPP_THREADS) {
				const uint2 qb = S.qbox[j];
				const uint32_t hi = qb.y & ~BPP_C;
				const uint32_t lx0 = qb.x & 0xFFFFu, ly0 = qb.x >> 16, lx1 = hi & 0xFFFFu, ly1 = hi >> 16;
				if (lx0 > lx1 || ly0 > ly1) {
					continue; // empty
				}
				const uint32_t cells = ((lx1 >> 10) - (lx0 >> 10) + 1) * ((ly1 >> 10) - (ly0 >> 10) + 1);
				if (cells > BPP_WIDE_T) {
					S.wide[atomicAdd(&S.n_wide, 1u)] = (uint16_t)j;
				} else {
					for (uint32_t cy = ly0 >> 10; cy <= (ly1 >> 10); cy++) {
						for (uint32_t cx = lx0 >> 10; cx <= (lx1 >> 10); cx++) {
							S.ent[atomicAdd(&S.cnt[cy * 32 + cx], 1u)] = (uint16_t)j;
						}
					}
				}
			}
			__syncthreads();
			const uint32_t n_wide = S.n_wide;
			// ---- (c') micro-cell masks: one thread per sub-cell classifies the first 32 entries of its list against
			// the four 512x512-quantum quarters of the sub-cell
			{
				const uint32_t lb = S.off[tid], ln = min((uint32_t)S.off[tid + 1] - lb, 32u);
				const uint32_t sx = (tid & 31u) << 10, sy = (tid >> 5) << 10;
				uint32_t df[4] = {0, 0, 0, 0}, am[4] = {0, 0, 0, 0};
				for (uint32_t e = 0; e < ln; e++) {
					const uint2 qb = S.qbox[S.ent[lb + e]];
					const uint32_t hi = qb.y & ~BPP_C;
					const uint32_t lx0 = qb.x & 0xFFFFu, ly0 = qb.x >> 16, lx1 = hi & 0xFFFFu, ly1 = hi >> 16;
#pragma unroll
					for (int m = 0; m < 4; m++) {
						const uint32_t mx0 = sx + (m & 1) * 512u, mx1 = mx0 + 511u, my0 = sy + (m >> 1) * 512u, my1 = my0 + 511u;
						if (!(lx1 < mx0 || lx0 > mx1 || ly1 < my0 || ly0 > my1)) {
							if (lx0 < mx0 && mx1 < lx1 && ly0 < my0 && my1 < ly1) {
								df[m] |= 1u << e;
							} else {
								am[m] |= 1u << e;
							}
						}
					}
				}
#pragma unroll
				for (int m = 0; m < 4; m++) {
					S.mask[tid * 4 + m] = make_uint2(df[m], am[m]);
				}
			}
			__syncthreads();

			// ---- (d) stream the unit's points: each 256-thread group takes every fourth round of 256 points ----
			const unsigned gid = tid / BPP_GTHREADS, gtid = tid & (BPP_GTHREADS - 1);
			for (uint32_t p0 = unit.pbeg + gid * BPP_GTHREADS; p0 < unit.pend; p0 += BPP_THREADS) {
				const uint32_t i = p0 + gtid;
				const bool valid = i < unit.pend;
				double x = 0, y = 0;
				uint32_t id = 0, PqC = BPP_C, Pq = 0, beg = 0, lenc = 0, msel = 0;
				if (valid) {
					id = pidx[i];
				}
				S.rid[gid][gtid] = id; // read by this round's copy-outs (after the group barrier)
				if (i + BPP_THREADS < unit.pend) {
					// this group's next round: pull the lines into L2 now, the loads then cost an L2 hit
					asm volatile("prefetch.global.L2 [%0];" ::"l"(pxy + i + BPP_THREADS));
					if ((gtid & 7) == 0) {
						asm volatile("prefetch.global.L2 [%0];" ::"l"(pidx + i + BPP_THREADS));
					}
				}
				if (valid) {
					const double2 xy = pxy[i];
					x = xy.x, y = xy.y;
					const uint32_t lx = (uint32_t)(bp_q(x, g.minx, g.sx) - ox), ly = (uint32_t)(bp_q(y, g.miny, g.sy) - oy);
					Pq = (ly << 16) | lx;
					PqC = Pq | BPP_C;
					const uint32_t sc = (ly >> 10) * 32 + (lx >> 10);
					beg = S.off[sc];
					lenc = S.off[sc + 1] - beg;
					msel = sc * 4 + ((ly >> 9) & 1u) * 2 + ((lx >> 9) & 1u);
				}
				for (uint32_t e0 = 0; bpp_gor(gid, lenc > e0); e0 += 64) {
					uint32_t mlo = 0, mhi = 0;
					const uint16_t *ep = S.ent + beg + e0;
					if (lenc > e0) {
						const uint32_t n = min(lenc - e0, 64u);
						if (e0 == 0) {
							const uint2 mk = S.mask[msel];
							BPP_PASS_MASKED(ep, n, mk);
						} else {
							BPP_PASS(ep, n);
						}
					}
					bpp_emit<MODE>(S, gid, mlo, mhi, ep, id, out_probe, out_build, capacity, counter);
				}
				for (uint32_t w0 = 0; w0 < n_wide; w0 += 64) {
					uint32_t mlo = 0, mhi = 0;
					const uint16_t *ep = S.wide + w0;
					if (valid) {
						const uint32_t n = min(n_wide - w0, 64u);
						BPP_PASS(ep, n);
					}
					bpp_gsync(gid);
					bpp_emit<MODE>(S, gid, mlo, mhi, ep, id, out_probe, out_build, capacity, counter);
				}
			}
			c0 += ntry;
		}
	}
}

__global__ void __launch_bounds__(256) k_bp_unpack(const double2 *__restrict__ pxy, uint32_t n, double *__restrict__ x,
                                                   double *__restrict__ y) {
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		x[i] = pxy[i].x;
		y[i] = pxy[i].y;
	}
}

// ---- host side ---------------------------------------------------------------------------------------------------
int sbi_flat_probe_bucketed(sb_context *ctx, const sb_flat *tree, const sb_probe_args &a, uint32_t *out_probe,
                            uint32_t *out_build, uint64_t capacity, uint64_t *count) {
	*count = 0;
	const uint64_t n = a.m;
	const uint32_t items = tree->h.item_count;
	if (n == 0 || items == 0) {
		return SB_OK;
	}
	SB_ARG(a.px != nullptr && a.py != nullptr, "bucketed probe takes points");
	SB_ARG(n + a.index_base <= 0xFFFFFFFFull, "probe indices must fit 32 bits");
	const int mode = a.pred >= 0 ? 2 : 1;
	SB_ARG(mode == 1 || a.build_box2d != nullptr, "build_box2d");

	bp_grid *d_grid = nullptr;
	bp_tables *d_tab = nullptr;
	bp_unit *d_units = nullptr;
	double2 *d_pxy = nullptr;
	uint32_t *d_pidx = nullptr, *d_cand = nullptr;
	unsigned long long *d_counter = nullptr;
	uint32_t *d_fine = nullptr; // (bucket, sub-cell) histogram for a second partition level: measured +0.7 ms, unused
	const uint64_t max_units = n / BP_SLICE + BP_NB + 1;
	SB_TRY(ctx->scratch_t(1, &d_grid));
	SB_TRY(ctx->scratch_t(1, &d_tab));
	SB_TRY(ctx->scratch_t(max_units, &d_units));
	SB_TRY(ctx->scratch_t(1, &d_counter));
	SB_TRY(ctx->scratch_t(n, &d_pxy));
	SB_TRY(ctx->scratch_t(n, &d_pidx));
	SB_TRY(sbi_zero(ctx, d_tab, sizeof(bp_tables)));
	SB_TRY(sbi_zero(ctx, d_counter, sizeof(*d_counter)));
	const sb_flat_header *d_hdr = static_cast<const sb_flat_header *>(tree->buffer);
	SB_TRACE(ctx, "bp: scratch + memsets");
	{
		sb_prof_scope ps(ctx, "bp_partition");
		k_bp_grid_init<<<1, 32, 0, ctx->stream>>>(d_hdr, d_grid);
		SB_LAUNCH_CHECK(ctx);
		k_bp_hist<<<sb_grid(n, BPH_THREADS * 8, 4), BPH_THREADS, 0, ctx->stream>>>(a.px, a.py, n, d_grid, d_tab->pt_hist, d_fine);
		SB_LAUNCH_CHECK(ctx);
	}
	{
		sb_prof_scope ps(ctx, "bp_cand");
		k_bp_cand<false><<<sb_grid(items, 256, 8), 256, 0, ctx->stream>>>(tree->boxes(), items, d_grid, d_tab->cand_cnt,
		                                                                 nullptr);
		SB_LAUNCH_CHECK(ctx);
		k_bp_scan<<<1, 1024, 0, ctx->stream>>>(d_tab, d_units);
		SB_LAUNCH_CHECK(ctx);
	}
	SB_TRACE(ctx, "bp: hist + cand count + scan");
	uint64_t totals[4];
	SB_TRY(sbi_read_words(ctx, d_tab->totals, 4, totals));
	const uint64_t n_cand = totals[1], n_units = totals[2], n_nan = totals[3];
	SB_ARG(n_cand < 0xFFFFFFFFull, "bucketed probe: candidate directory too large");
	SB_TRY(ctx->scratch_t(n_cand ? n_cand : 1, &d_cand));
	{
		sb_prof_scope ps(ctx, "bp_partition");
		SB_SMEM_OPTIN(ctx, k_bp_scatter, sizeof(bps_smem));
		k_bp_scatter<<<sb_grid(n, BPS_TILE, 2), BPS_THREADS, sizeof(bps_smem), ctx->stream>>>(
		    a.px, a.py, n, a.index_base, d_grid, d_tab->pt_cursor, d_pxy, d_pidx);
		SB_LAUNCH_CHECK(ctx);
	}
	{
		sb_prof_scope ps(ctx, "bp_cand");
		k_bp_cand<true><<<sb_grid(items, 256, 8), 256, 0, ctx->stream>>>(tree->boxes(), items, d_grid, d_tab->cand_cursor,
		                                                                d_cand);
		SB_LAUNCH_CHECK(ctx);
	}
	SB_TRACE(ctx, "bp: scatter + cand fill");
	if (n_units) {
		SB_SMEM_OPTIN(ctx, k_bp_probe<1>, sizeof(bpp_smem<1>));
		SB_SMEM_OPTIN(ctx, k_bp_probe<2>, sizeof(bpp_smem<2>));
		const unsigned grid = (unsigned)(n_units < SB_NUM_SMS ? n_units : SB_NUM_SMS);
		sb_prof_scope ps(ctx, "flat_probe");
		if (mode == 1) {
			k_bp_probe<1><<<grid, BPP_THREADS, sizeof(bpp_smem<1>), ctx->stream>>>(
			    tree->boxes(), tree->idx(), nullptr, -1, d_grid, d_tab, d_units, d_cand, d_pxy, d_pidx, out_probe, out_build,
			    capacity, d_counter);
		} else {
			k_bp_probe<2><<<grid, BPP_THREADS, sizeof(bpp_smem<2>), ctx->stream>>>(
			    tree->boxes(), tree->idx(), a.build_box2d, a.pred, d_grid, d_tab, d_units, d_cand, d_pxy, d_pidx, out_probe,
			    out_build, capacity, d_counter);
		}
		SB_LAUNCH_CHECK(ctx);
	}
	uint64_t c1 = 0;
	SB_TRY(sbi_read_u64(ctx, reinterpret_cast<const uint64_t *>(d_counter), &c1));
	SB_TRACE(ctx, "bp: probe kernel");
	*count = c1;
	if (n_nan && mode == 1) {
		// NaN coordinates: Box::Intersects never rejects on a NaN axis. Such rows take the per-probe tree walk.
		double *nx = nullptr, *ny = nullptr;
		SB_TRY(ctx->scratch_t(n_nan, &nx));
		SB_TRY(ctx->scratch_t(n_nan, &ny));
		const uint64_t off = totals[0];
		k_bp_unpack<<<sb_grid(n_nan, 256, 8), 256, 0, ctx->stream>>>(d_pxy + off, (uint32_t)n_nan, nx, ny);
		SB_LAUNCH_CHECK(ctx);
		sb_probe_args b;
		b.px = nx;
		b.py = ny;
		b.m = n_nan;
		b.ids = d_pidx + off;
		b.force_walk = true;
		uint64_t c2 = 0;
		const uint64_t done = c1 < capacity ? c1 : capacity;
		int s = sbi_flat_probe(ctx, tree, b, out_probe + done, out_build + done, capacity - done, &c2);
		*count = c1 + c2;
		if (s != SB_OK && s != SB_ERR_CAPACITY) {
			return s;
		}
	}
	if (*count > capacity) {
		sb_set_error("probe: %llu pairs, capacity %llu", (unsigned long long)*count, (unsigned long long)capacity);
		return SB_ERR_CAPACITY;
	}
	return SB_OK;
}

// ==================================================================================================================
// K5 v3: cell-sorted probe. The v2 kernel above is issue-bound (33 warp instructions per probe, profiles r1e): every
// lane walks its own sub-cell list, so the test and emit loops run max-over-lanes iterations and every pair passes
// through a shared-memory staging buffer. v3 sorts the probes one level further -- by sub-cell inside the bucket --
// so that the 32 lanes of a warp always share ONE sub-cell and therefore one entry list:
//
//   pass 1  k_cp_scatter1   points -> 8-byte records {15+15-bit in-bucket quantised coordinate, probe id}, grouped
//                           by bucket (16 B read, 8 B written per point instead of 20 B), + the (bucket, sub-cell)
//                           histogram
//   scan    k_cp_scan2      per-bucket exclusive scan of the 1024 sub-cell counts -> cell_off[1024 x 1024 + 1]
//   pass 2  k_cp_scatter2   records of one bucket -> grouped by sub-cell (8 B read + 8 B written per point)
//   probe   k_cp_probe      a warp takes one sub-cell: the entry list is warp-uniform (broadcast shared loads, no
//                           divergence); entries whose quantised box strictly covers the sub-cell hit every point
//                           without a test (32 pairs per store instruction, straight to global memory); the others
//                           are tested by all lanes at once, the hit ballots are kept in registers (lane e keeps
//                           entry e's masks for up to 8 rounds of 32 points), one global atomic per warp and batch
//                           (<= 256 points x 32 entries) reserves the output range, and the pairs are written
//                           entry-major, ballot-compacted, with no staging buffer and no CTA barrier.
//
// The exact coordinates are only needed for quantisation ties (Q(p) equal to a box bound): they are fetched from the
// caller's columns through the probe id. Result set: identical to v2 and to the tree walk (same three-state test).
// ==================================================================================================================
#define CP_TILE1 8192u
#define CP_TILE2 16384u
#define CP_NCELL (BP_NB * 1024u)

__device__ __forceinline__ uint32_t cp_subcell(uint32_t pq) {
	return ((pq >> 26) << 5) | ((pq >> 10) & 31u); // (ly >> 10) * 32 + (lx >> 10)
}

template <bool XY>
struct cps1_smem {
	uint2 rec[CP_TILE1];
	uint16_t bkt[CP_TILE1];
	uint32_t cnt[BP_NB + 8];
	uint32_t off[BP_NB + 8];
	uint32_t lim[BP_NB + 8]; // SAMPLED: end of every bucket's region (exclusive)
	uint32_t scan[32];
	double2 xy[XY ? CP_TILE1 : 1]; // XY: the exact coordinates travel with the records (polygon join)
};

// SAMPLED: the bucket regions were sized from a sample (k_bp_hist_sample); region_end[b] = first slot after bucket b's
// region, records that do not fit are not written (the overflow is detected afterwards by k_cp_actual)
template <bool XY, bool SAMPLED = false, bool FINE = true>
__global__ void __launch_bounds__(1024, 1)
    k_cp_scatter1(const double *__restrict__ px, const double *__restrict__ py, uint64_t n, uint32_t index_base,
                  const bp_grid *__restrict__ gp, uint32_t *__restrict__ cursor, uint32_t *__restrict__ fine,
                  uint2 *__restrict__ orec, double2 *__restrict__ oxy, const uint32_t *__restrict__ region_end = nullptr) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	cps1_smem<XY> &S = *reinterpret_cast<cps1_smem<XY> *>(smem_raw);
	const bp_grid g = *gp;
	const unsigned tid = threadIdx.x;
	const uint64_t tiles = (n + CP_TILE1 - 1) / CP_TILE1;
	constexpr int ITEMS = CP_TILE1 / 1024;
	if (SAMPLED) {
		S.lim[tid] = region_end[tid];
		if (tid == 0) {
			S.lim[BP_NB] = region_end[BP_NB];
		}
	}
	// One 1024-thread CTA per SM (shared memory): nothing else hides the latency of a tile's coordinate loads, so the
	// next tile's loads are issued as soon as the registers are free (after the ranking atomics) and complete behind
	// the scan / reorder / write-out of the current tile. XY keeps the coordinates until the reorder: no prefetch.
	double x[ITEMS], y[ITEMS];
	if (!XY && blockIdx.x < tiles) {
#pragma unroll
		for (int k = 0; k < ITEMS; k++) {
			const uint64_t i = (uint64_t)blockIdx.x * CP_TILE1 + tid + (uint64_t)k * 1024;
			if (i < n) {
				x[k] = px[i];
				y[k] = py[i];
			}
		}
	}
	for (uint64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
		S.cnt[tid] = 0;
		if (tid == 0) {
			S.cnt[BP_NB] = 0;
		}
		__syncthreads();
		const uint64_t base = tile * CP_TILE1 + tid;
		if (XY) {
#pragma unroll
			for (int k = 0; k < ITEMS; k++) {
				const uint64_t i = base + (uint64_t)k * 1024;
				if (i < n) {
					x[k] = px[i];
					y[k] = py[i];
				}
			}
		}
		uint32_t pq[ITEMS], br[ITEMS];
#pragma unroll
		for (int k = 0; k < ITEMS; k++) {
			const uint64_t i = base + (uint64_t)k * 1024;
			br[k] = 0xFFFFFFFFu;
			pq[k] = 0;
			if (i < n) {
				const int b = bp_bucket(x[k], y[k], g);
				if (b >= 0) {
					const int qx = bp_q(x[k], g.minx, g.sx), qy = bp_q(y[k], g.miny, g.sy);
					pq[k] = ((uint32_t)(qy & 32767) << 16) | (uint32_t)(qx & 32767);
					br[k] = ((uint32_t)b << 16) | atomicAdd(&S.cnt[b], 1u);
				}
			}
		}
		if (!XY && tile + gridDim.x < tiles) {
			const uint64_t nbase = (tile + gridDim.x) * CP_TILE1 + tid;
#pragma unroll
			for (int k = 0; k < ITEMS; k++) {
				const uint64_t i = nbase + (uint64_t)k * 1024;
				if (i < n) {
					x[k] = px[i];
					y[k] = py[i];
				}
			}
		}
		__syncthreads();
		{
			uint32_t total;
			const uint32_t c = S.cnt[tid];
			const uint32_t ex = bp_block_excl_scan_1024(c, S.scan, &total);
			S.off[tid] = ex;
			if (c) {
				S.cnt[tid] = atomicAdd(&cursor[tid], c) - ex;
			}
			if (tid == 1023) {
				const uint32_t cn = S.cnt[BP_NB];
				S.off[BP_NB] = total;
				S.off[BP_NB + 1] = total + cn;
				if (cn) {
					S.cnt[BP_NB] = atomicAdd(&cursor[BP_NB], cn) - total;
				}
			}
		}
		__syncthreads();
#pragma unroll
		for (int k = 0; k < ITEMS; k++) {
			if (br[k] != 0xFFFFFFFFu) {
				const uint32_t b = br[k] >> 16;
				const uint32_t slot = S.off[b] + (br[k] & 0xFFFFu);
				S.rec[slot] = make_uint2(pq[k], (uint32_t)(base + (uint64_t)k * 1024) + index_base);
				S.bkt[slot] = (uint16_t)b;
				if (XY) {
					S.xy[slot] = make_double2(x[k], y[k]);
				}
			}
		}
		__syncthreads();
		const uint32_t kept = S.off[BP_NB + 1];
		for (uint32_t t = tid; t < kept; t += 1024) {
			const uint32_t b = S.bkt[t];
			const uint2 r = S.rec[t];
			if (!SAMPLED || S.cnt[b] + t < S.lim[b]) {
				orec[S.cnt[b] + t] = r;
			}
			if (XY) {
				oxy[S.cnt[b] + t] = S.xy[t];
			}
			if (FINE && b < BP_NB) {
				atomicAdd(&fine[(b << 10) | cp_subcell(r.x)], 1u);
			}
		}
		__syncthreads();
	}
}

// The (bucket, sub-cell) histogram from the bucket-ordered records instead of one global RED per record inside pass 1
// (200 M scattered REDs = 0.9 ms of LSU time there): a 16384-record tile of ONE bucket is counted in shared memory and
// leaves as at most 1024 global atomics.
__global__ void __launch_bounds__(1024, 2)
    k_cp_finehist(const bp_tables *__restrict__ T, const bp_unit *__restrict__ units, const uint2 *__restrict__ irec,
                  uint32_t *__restrict__ fine) {
	__shared__ uint32_t s_cnt[1024];
	const unsigned tid = threadIdx.x;
	const uint32_t n_units = (uint32_t)T->totals[2];
	for (uint32_t w = blockIdx.x; w < BP_UNIT_TILES * n_units; w += gridDim.x) {
		const bp_unit u = units[w / BP_UNIT_TILES];
		const uint32_t tbeg = u.pbeg + (w % BP_UNIT_TILES) * CP_TILE2;
		const uint32_t tend = min(u.pend, tbeg + CP_TILE2);
		if (tbeg >= tend) {
			continue; // uniform
		}
		s_cnt[tid] = 0;
		__syncthreads();
		for (uint32_t i = tbeg + tid; i < tend; i += 1024) {
			atomicAdd(&s_cnt[cp_subcell(irec[i].x)], 1u);
		}
		__syncthreads();
		const uint32_t c = s_cnt[tid];
		if (c) {
			atomicAdd(&fine[(u.bucket << 10) | tid], c);
		}
		__syncthreads();
	}
}

// ---- sampled bucket histogram (round 2) --------------------------------------------------------------------------
// k_bp_hist reads all coordinates once only to size the 1024 bucket regions of pass 1 (0.69 ms of the 9.9 ms C3 step).
// For large inputs the regions are sized from every CP_SAMPLE-th 4096-point tile instead (est = 1.25 x scaled count +
// 4096), pass 1 never writes past a region's end, and the exact counts are read off the region cursors afterwards; if
// any bucket did not fit (adversarial / periodic input) the caller repeats the partition with the exact histogram.
#define CP_SAMPLE 16u
#define CP_SAMPLE_MIN (32ull << 20) // inputs below this take the exact histogram
__global__ void __launch_bounds__(BPH_THREADS) k_bp_hist_sample(const double *__restrict__ px, const double *__restrict__ py,
                                                                uint64_t n, const bp_grid *__restrict__ gp,
                                                                uint32_t *__restrict__ hist) {
	__shared__ uint32_t s_hist[BP_NB + 1];
	for (int i = threadIdx.x; i <= BP_NB; i += BPH_THREADS) {
		s_hist[i] = 0;
	}
	const bp_grid g = *gp;
	__syncthreads();
	const uint64_t tiles = (n + 4095) / 4096;
	for (uint64_t tile = (uint64_t)blockIdx.x * CP_SAMPLE; tile < tiles; tile += (uint64_t)gridDim.x * CP_SAMPLE) {
#pragma unroll
		for (int k = 0; k < 8; k++) {
			const uint64_t i = tile * 4096 + (uint64_t)k * BPH_THREADS + threadIdx.x;
			if (i < n) {
				const int b = bp_bucket(px[i], py[i], g);
				if (b >= 0) {
					atomicAdd(&s_hist[b], 1u);
				}
			}
		}
	}
	__syncthreads();
	for (int i = threadIdx.x; i <= BP_NB; i += BPH_THREADS) {
		const uint32_t c = s_hist[i];
		if (c) {
			atomicAdd(&hist[i], c);
		}
	}
}

// sample counts -> region sizes
__global__ void __launch_bounds__(1024) k_bp_inflate(bp_tables *__restrict__ T) {
	for (uint32_t b = threadIdx.x; b <= BP_NB; b += 1024) {
		const uint64_t scaled = (uint64_t)T->pt_hist[b] * CP_SAMPLE;
		T->pt_hist[b] = (uint32_t)min((uint64_t)0x7FFFFFFFull, scaled + scaled / 4 + 4096);
	}
}

// after pass 1 with sampled regions: exact counts from the region cursors, overflow flag (T->pad), the work units and the
// totals the host reads ({offset of the NaN region, candidates, units, NaN points})
__global__ void __launch_bounds__(1024) k_cp_actual(bp_tables *__restrict__ T, bp_unit *__restrict__ units) {
	__shared__ uint32_t s_warp[32];
	__shared__ uint32_t s_over;
	const unsigned b = threadIdx.x;
	if (b == 0) {
		s_over = 0;
	}
	__syncthreads();
	const uint32_t po = T->pt_off[b], room = T->pt_off[b + 1] - po;
	uint32_t np = T->pt_cursor[b] - po;
	if (np > room) {
		atomicOr(&s_over, 1u);
		np = room;
	}
	T->pt_hist[b] = np;
	const uint32_t nc = T->cand_cnt[b];
	const uint32_t nu = (np && nc) ? (np + BP_SLICE - 1) / BP_SLICE : 0u;
	uint32_t total;
	const uint32_t uo = bp_block_excl_scan_1024(nu, s_warp, &total);
	for (uint32_t k = 0; k < nu; k++) {
		bp_unit u;
		u.bucket = b;
		u.pbeg = po + k * BP_SLICE;
		u.pend = min(po + np, u.pbeg + BP_SLICE);
		units[uo + k] = u;
	}
	__syncthreads();
	if (b == 0) {
		const uint32_t nan_room = T->pt_off[BP_NB + 1] - T->pt_off[BP_NB];
		uint32_t nan = T->pt_cursor[BP_NB] - T->pt_off[BP_NB];
		if (nan > nan_room) {
			s_over = 1;
			nan = nan_room;
		}
		T->pt_hist[BP_NB] = nan;
		T->totals[0] = T->pt_off[BP_NB];
		T->totals[2] = total;
		T->totals[3] = nan;
		T->ticket = 0;
		T->pad = s_over;
	}
}

// one CTA per bucket: cell_off[b * 1024 + s] = pt_off[b] + exclusive scan of the bucket's sub-cell counts;
// `fine` becomes the pass-2 cursor array
__global__ void __launch_bounds__(1024) k_cp_scan2(const bp_tables *__restrict__ T, uint32_t *__restrict__ fine,
                                                   uint32_t *__restrict__ cell_off) {
	__shared__ uint32_t s_warp[32];
	const uint32_t b = blockIdx.x, i = b * 1024u + threadIdx.x;
	uint32_t total;
	const uint32_t ex = bp_block_excl_scan_1024(fine[i], s_warp, &total) + T->pt_off[b];
	cell_off[i] = ex;
	fine[i] = ex;
	if (b == BP_NB - 1 && threadIdx.x == 0) {
		cell_off[CP_NCELL] = T->pt_off[BP_NB];
	}
}

template <bool XY>
struct cps2_smem {
	uint2 rec[CP_TILE2];
	uint32_t src[XY ? CP_TILE2 : 1]; // XY: position of the record in bucket order (where its coordinates are)
	uint32_t cnt[BP_NB];
	uint32_t off[BP_NB];
	uint32_t scan[32];
};

template <bool XY>
__global__ void __launch_bounds__(1024, 1)
    k_cp_scatter2(const bp_tables *__restrict__ T, const bp_unit *__restrict__ units, const uint2 *__restrict__ irec,
                  uint32_t *__restrict__ cursor2, uint2 *__restrict__ orec, uint32_t *__restrict__ osrc) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	cps2_smem<XY> &S = *reinterpret_cast<cps2_smem<XY> *>(smem_raw);
	const unsigned tid = threadIdx.x;
	const uint32_t n_units = (uint32_t)T->totals[2];
	constexpr int ITEMS = CP_TILE2 / 1024;
	// work items: tiles of the probe units (a unit has at most BP_SLICE = BP_UNIT_TILES x CP_TILE2 records of one bucket)
	static_assert(BP_SLICE == BP_UNIT_TILES * CP_TILE2, "unit = whole tiles");
	for (uint32_t w = blockIdx.x; w < BP_UNIT_TILES * n_units; w += gridDim.x) {
		const bp_unit u = units[w / BP_UNIT_TILES];
		const uint32_t tbeg = u.pbeg + (w % BP_UNIT_TILES) * CP_TILE2;
		const uint32_t tend = min(u.pend, tbeg + CP_TILE2);
		if (tbeg >= tend) {
			continue; // uniform
		}
		S.cnt[tid] = 0;
		__syncthreads();
		uint2 r[ITEMS];
		uint32_t rk[ITEMS];
#pragma unroll
		for (int k = 0; k < ITEMS; k++) {
			const uint32_t i = tbeg + k * 1024 + tid;
			if (i < tend) {
				r[k] = irec[i];
			}
		}
#pragma unroll
		for (int k = 0; k < ITEMS; k++) {
			const uint32_t i = tbeg + k * 1024 + tid;
			rk[k] = i < tend ? atomicAdd(&S.cnt[cp_subcell(r[k].x)], 1u) : 0u;
		}
		__syncthreads();
		{
			uint32_t total;
			const uint32_t c = S.cnt[tid];
			const uint32_t ex = bp_block_excl_scan_1024(c, S.scan, &total);
			S.off[tid] = ex;
			if (c) {
				S.cnt[tid] = atomicAdd(&cursor2[(u.bucket << 10) | tid], c) - ex;
			}
		}
		__syncthreads();
#pragma unroll
		for (int k = 0; k < ITEMS; k++) {
			const uint32_t i = tbeg + k * 1024 + tid;
			if (i < tend) {
				const uint32_t slot = S.off[cp_subcell(r[k].x)] + rk[k];
				S.rec[slot] = r[k];
				if (XY) {
					S.src[slot] = i;
				}
			}
		}
		__syncthreads();
		const uint32_t nt = tend - tbeg;
		for (uint32_t t = tid; t < nt; t += 1024) {
			const uint2 v = S.rec[t];
			const uint32_t dst = S.cnt[cp_subcell(v.x)] + t;
			orec[dst] = v;
			if (XY) {
				osrc[dst] = S.src[t];
			}
		}
		__syncthreads();
	}
}

#ifndef CPP_RB
#define CPP_RB 8 // rounds of 32 points per batch (8 or 4)
#endif
#ifndef CPP_THREADS
#define CPP_THREADS 1024 // threads per probe CTA: 1024 (one CTA per SM) or 512 (two CTAs per SM, smaller directory)
#endif
#ifndef CPP_CTAS
#define CPP_CTAS (1024 / CPP_THREADS) // resident probe CTAs per SM
#endif
#ifndef CPP_EMAX
#define CPP_EMAX (CPP_THREADS == 1024 ? BPP_EMAX : 24576)
#endif
#ifndef CPP_CMAX
#define CPP_CMAX BPP_CMAX
#endif
template <int MODE>
struct cpp_smem {
	uint2 qbox[CPP_CMAX];
	uint32_t cidx[CPP_CMAX];
	uint32_t cpos[MODE == 1 ? CPP_CMAX : 1];
	uint32_t cnt[BP_NB];
	uint16_t off[BP_NB + 8];
	uint16_t ent[CPP_EMAX];
	uint16_t wide[BPP_WMAX];
	uint32_t scan[32];
	uint32_t unit, n_ent, n_wide, next_sc, sc_first, sc_last;
};

// two predicated 4-byte stores sharing one predicate (no branch around the store pair)
__device__ __forceinline__ void cpp_st2_if(uint32_t *p0, uint32_t v0, uint32_t *p1, uint32_t v1, uint32_t pred) {
	asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %4, 0;\n\t@q st.global.u32 [%0], %1;\n\t@q st.global.u32 [%2], %3;\n\t}"
	             :
	             : "l"(p0), "r"(v0), "l"(p1), "r"(v1), "r"(pred)
	             : "memory");
}

// writes the pairs of one batch: entry-major; definite entries emit every point of the batch, tested entries their
// ballot-compacted hits. `base` is warp-uniform; offsets inside the batch are 32-bit.
template <bool GUARD, int NR>
__device__ __forceinline__ void cpp_write(const uint32_t (&id)[NR], const uint32_t (&mk)[NR], uint32_t defm,
                                          const uint16_t *ep, uint32_t ne, uint32_t npts, const uint32_t *cidx,
                                          unsigned long long base, uint64_t capacity, uint32_t *__restrict__ out_probe,
                                          uint32_t *__restrict__ out_build) {
	unsigned lane = sb_lane(), lt = sb_lanemask_lt(), eq = 1u << lane;
	uint32_t *pp = out_probe + base, *pb = out_build + base;
	// keep the lane id and the two batch pointers in registers: offsets inside the batch stay 32-bit (one IMAD.WIDE
	// per store instead of a re-derived 64-bit base + offset sum)
	asm volatile("" : "+r"(lane), "+r"(lt), "+r"(eq), "+l"(pp), "+l"(pb));
	const uint32_t room = GUARD ? (uint32_t)(capacity > base ? min((unsigned long long)0xFFFFFFFFu, capacity - base) : 0ull)
	                            : 0xFFFFFFFFu;
	uint32_t off = 0;
	for (uint32_t e = 0; e < ne; e++) {
		const uint32_t bi = cidx[ep[e]];
		if ((defm >> e) & 1u) {
#pragma unroll
			for (int r = 0; r < NR; r++) {
				const uint32_t o = off + (unsigned)r * 32u + lane;
				cpp_st2_if(pp + o, id[r], pb + o, bi,
				           (r < NR - 1 || (unsigned)r * 32u + lane < npts) && (!GUARD || o < room));
			}
			off += npts;
		} else {
#pragma unroll
			for (int r = 0; r < NR; r++) {
				const uint32_t m = __shfl_sync(0xFFFFFFFFu, mk[r], e);
				const uint32_t o = off + __popc(m & lt);
				cpp_st2_if(pp + o, id[r], pb + o, bi, GUARD ? ((m & eq) && o < room) : (m & eq));
				off += __popc(m);
			}
		}
	}
}

struct cpp_args {
	const float4 *boxes;
	const double4 *build_box2d;
	int pred;
	const uint2 *rec;
	const double *px, *py;
	uint32_t index_base;
	uint32_t *out_probe, *out_build;
	uint64_t capacity;
	unsigned long long *counter;
};

// One batch = the points [pb, pb + npts) of one sub-cell (NR = ceil(npts / 32) rounds of 32 lanes, only the last round
// may be partial) against the sub-cell's entry list and the bucket's wide list, 32 entries at a time.
template <int MODE, int NR>
__device__ __forceinline__ void cpp_batch(cpp_smem<MODE> &S, const cpp_args &A, unsigned wlane, uint32_t pb, uint32_t npts,
                                          uint32_t lb, uint32_t n0, uint32_t n_wide, uint32_t sx0, uint32_t sy0) {
	uint32_t PqC[NR], id[NR];
#pragma unroll
	for (int r = 0; r < NR; r++) {
		PqC[r] = BPP_C;
		id[r] = 0;
		if (r < NR - 1 || (unsigned)r * 32u + wlane < npts) {
			const uint2 v = A.rec[pb + r * 32 + wlane];
			PqC[r] = v.x | BPP_C;
			id[r] = v.y;
		}
	}
	const uint32_t last_mask = npts - (NR - 1) * 32u >= 32u ? 0xFFFFFFFFu : ((1u << (npts - (NR - 1) * 32u)) - 1u);
	for (int pass = 0; pass < 2; pass++) {
		const uint16_t *list = pass == 0 ? S.ent + lb : S.wide;
		const uint32_t nl = pass == 0 ? n0 : n_wide;
		for (uint32_t e0 = 0; e0 < nl; e0 += 32) {
			const uint16_t *ep = list + e0;
			const uint32_t ne = min(nl - e0, 32u);
			uint32_t mk[NR];
#pragma unroll
			for (int r = 0; r < NR; r++) {
				mk[r] = 0;
			}
			uint32_t defm = 0;
			for (uint32_t e = 0; e < ne; e++) {
				const uint32_t j = ep[e];
				const uint2 qb = S.qbox[j];
				const uint32_t hi = qb.y & ~BPP_C;
				const uint32_t lx0 = qb.x & 0xFFFFu, ly0 = qb.x >> 16, lx1 = hi & 0xFFFFu, ly1 = hi >> 16;
				if (lx0 < sx0 && sx0 + 1023u < lx1 && ly0 < sy0 && sy0 + 1023u < ly1) {
					defm |= 1u << e; // strictly covers the sub-cell: every point is an exact hit
					continue;
				}
				// per 16-bit half: t1 = 0x8000 + (p - lo), t2 = 0x8000 + (hi - p); s1 / s2 = the same minus 1.
				// bit 15 of t <=> that side holds (closed), of s <=> it holds strictly. hiC2 - PqC == qb.y - Pq.
				const uint32_t lo = qb.x, hiC2 = qb.y + BPP_C, loK = lo + BPP_K, hiK = hiC2 - BPP_K;
				const bool mine = wlane == e;
#pragma unroll
				for (int r = 0; r < NR; r++) {
					const uint32_t s1 = PqC[r] - loK, s2 = hiK - PqC[r];
					uint32_t ms = __ballot_sync(0xFFFFFFFFu, (s1 & s2 & BPP_C) == BPP_C);
					const uint32_t t1 = PqC[r] - lo, t2 = hiC2 - PqC[r];
					const bool closed = (t1 & t2 & BPP_C) == BPP_C;
					const uint32_t mc = __ballot_sync(0xFFFFFFFFu, closed);
					if (mc != ms) {
						// some lane sits exactly on a quantised bound of this entry: the exact predicate decides
						bool h = false;
						if (closed && !((ms >> wlane) & 1u) && (r < NR - 1 || (unsigned)r * 32u + wlane < npts)) {
							const uint32_t o = id[r] - A.index_base;
							h = bp_exact<MODE>(A.px[o], A.py[o], MODE == 1 ? S.cpos[j] : 0u, S.cidx[j], A.boxes, A.build_box2d,
							                   A.pred);
						}
						ms |= __ballot_sync(0xFFFFFFFFu, h);
					}
					if (mine) {
						mk[r] = ms;
					}
				}
			}
			// lanes past the end of the batch carry PqC = 0|C: drop whatever they matched
			mk[NR - 1] &= last_mask;
			uint32_t cnt = 0;
#pragma unroll
			for (int r = 0; r < NR; r++) {
				cnt += __popc(mk[r]);
			}
			cnt = __reduce_add_sync(0xFFFFFFFFu, cnt) + __popc(defm) * npts;
			if (cnt == 0) {
				continue;
			}
			unsigned long long base = 0;
			if (wlane == 0) {
				base = atomicAdd(A.counter, (unsigned long long)cnt);
			}
			base = __shfl_sync(0xFFFFFFFFu, base, 0);
			// a batch that does not fit is counted but not written: on SB_ERR_CAPACITY the caller gets the needed size and
			// unspecified array contents (a guarded write variant per round count was a third of the kernel's code)
			if (base + cnt <= A.capacity) {
				cpp_write<false, NR>(id, mk, defm, ep, ne, npts, S.cidx, base, A.capacity, A.out_probe, A.out_build);
			}
		}
	}
}

template <int MODE>
__global__ void __launch_bounds__(CPP_THREADS, CPP_CTAS)
    k_cp_probe(const float4 *__restrict__ boxes, const uint32_t *__restrict__ idx,
               const double4 *__restrict__ build_box2d, int pred, const bp_grid *__restrict__ gp,
               bp_tables *__restrict__ T, const bp_unit *__restrict__ units, const uint32_t *__restrict__ cand_pos,
               const uint2 *__restrict__ rec, const uint32_t *__restrict__ cell_off, const double *__restrict__ px,
               const double *__restrict__ py, uint32_t index_base, uint32_t *__restrict__ out_probe,
               uint32_t *__restrict__ out_build, uint64_t capacity, unsigned long long *__restrict__ counter) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	cpp_smem<MODE> &S = *reinterpret_cast<cpp_smem<MODE> *>(smem_raw);
	const bp_grid g = *gp;
	const unsigned tid = threadIdx.x, lane = sb_lane();
	const uint32_t n_units = (uint32_t)T->totals[2];

	while (true) {
		__syncthreads();
		if (tid == 0) {
			S.unit = atomicAdd(&T->ticket, 1u);
		}
		__syncthreads();
		const uint32_t u = S.unit;
		if (u >= n_units) {
			break;
		}
		const bp_unit unit = units[u];
		const uint32_t *coff = cell_off + ((size_t)unit.bucket << 10);
		if (tid == 0) {
			// sub-cells of this bucket that own points of [pbeg, pend): first = last s with coff[s] <= pbeg,
			// last = last s with coff[s] < pend
			uint32_t lo = 0, hi = 1023;
			while (lo < hi) {
				const uint32_t mid = (lo + hi + 1) >> 1;
				if (coff[mid] <= unit.pbeg) {
					lo = mid;
				} else {
					hi = mid - 1;
				}
			}
			S.sc_first = lo;
			lo = 0, hi = 1023;
			while (lo < hi) {
				const uint32_t mid = (lo + hi + 1) >> 1;
				if (coff[mid] < unit.pend) {
					lo = mid;
				} else {
					hi = mid - 1;
				}
			}
			S.sc_last = lo;
		}
		__syncthreads();
		const uint32_t sc_first = S.sc_first, sc_last = S.sc_last;
		const int ox = (int)(unit.bucket % BP_G) << 15, oy = (int)(unit.bucket / BP_G) << 15;
		const uint32_t cbeg = T->cand_off[unit.bucket], cend = T->cand_off[unit.bucket + 1];
		uint32_t c0 = cbeg;
		while (c0 < cend) {
			// ---- (a) stage a chunk of leaf entries as quantised in-bucket boxes, count directory entries ----
			uint32_t ntry = min((uint32_t)CPP_CMAX, cend - c0);
			while (true) {
				__syncthreads();
				if (tid == 0) {
					S.n_ent = 0;
					S.n_wide = 0;
				}
				for (uint32_t sc = tid; sc < BP_NB; sc += CPP_THREADS) {
					S.cnt[sc] = 0;
				}
				__syncthreads();
				for (uint32_t j = tid; j < ntry; j += CPP_THREADS) {
					const uint32_t pos = cand_pos[c0 + j];
					const uint32_t bi = __ldg(idx + pos);
					int x0, y0, x1, y1;
					bool empty = false;
					if (MODE == 1) {
						bp_leaf_range(__ldg(boxes + pos), g, x0, y0, x1, y1);
						S.cpos[j] = pos;
					} else {
						const double4 e = build_box2d[bi];
						empty = (e.x != e.x) || (e.y != e.y) || (e.z != e.z) || (e.w != e.w);
						x0 = bp_q(e.x, g.minx, g.sx), y0 = bp_q(e.y, g.miny, g.sy);
						x1 = bp_q(e.z, g.minx, g.sx), y1 = bp_q(e.w, g.miny, g.sy);
					}
					empty = empty || x1 < ox || x0 > ox + 32767 || y1 < oy || y0 > oy + 32767 || x0 > x1 || y0 > y1;
					const uint32_t lx0 = (uint32_t)(max(x0, ox) - ox), ly0 = (uint32_t)(max(y0, oy) - oy);
					const uint32_t lx1 = (uint32_t)(min(x1, ox + 32767) - ox), ly1 = (uint32_t)(min(y1, oy + 32767) - oy);
					// rows of sub-cells outside [sc_first, sc_last] hold no point of this unit
					if (!empty && ((ly1 >> 10) < (sc_first >> 5) || (ly0 >> 10) > (sc_last >> 5))) {
						empty = true;
					}
					uint2 qb;
					qb.x = empty ? 0x7FFF7FFFu : ((ly0 << 16) | lx0);
					qb.y = (empty ? 0u : ((ly1 << 16) | lx1)) | BPP_C;
					S.qbox[j] = qb;
					S.cidx[j] = bi;
					if (!empty) {
						const uint32_t cells = ((lx1 >> 10) - (lx0 >> 10) + 1) * ((ly1 >> 10) - (ly0 >> 10) + 1);
						if (cells > BPP_WIDE_T) {
							atomicAdd(&S.n_wide, 1u);
						} else {
							uint32_t added = 0;
							for (uint32_t cy = ly0 >> 10; cy <= (ly1 >> 10); cy++) {
								for (uint32_t cx = lx0 >> 10; cx <= (lx1 >> 10); cx++) {
									const uint32_t s = cy * 32 + cx;
									if (s >= sc_first && s <= sc_last) {
										atomicAdd(&S.cnt[s], 1u);
										added++;
									}
								}
							}
							atomicAdd(&S.n_ent, added);
						}
					}
				}
				__syncthreads();
				if (S.n_ent <= CPP_EMAX && S.n_wide <= BPP_WMAX) {
					break;
				}
				ntry = (ntry + 1) / 2; // rare: the directory would overflow shared memory -> smaller chunk
			}
			// ---- (b) CSR offsets over the 1024 sub-cells (one per thread) ----
			{
				// each thread owns PER consecutive sub-cells
				constexpr int PER = BP_NB / CPP_THREADS;
				uint32_t total, c[PER], sum = 0;
#pragma unroll
				for (int q = 0; q < PER; q++) {
					c[q] = S.cnt[tid * PER + q];
					sum += c[q];
				}
				uint32_t ex = bp_block_excl_scan<CPP_THREADS / 32>(sum, S.scan, &total);
#pragma unroll
				for (int q = 0; q < PER; q++) {
					S.off[tid * PER + q] = (uint16_t)ex;
					S.cnt[tid * PER + q] = ex;
					ex += c[q];
				}
				if (tid == CPP_THREADS - 1) {
					S.off[BP_NB] = (uint16_t)ex;
				}
				if (tid == 0) {
					S.n_wide = 0;
					S.next_sc = sc_first;
				}
			}
			__syncthreads();
			// ---- (c) fill the directory ----
			for (uint32_t j = tid; j < ntry; j += CPP_THREADS) {
				const uint2 qb = S.qbox[j];
				const uint32_t hi = qb.y & ~BPP_C;
				const uint32_t lx0 = qb.x & 0xFFFFu, ly0 = qb.x >> 16, lx1 = hi & 0xFFFFu, ly1 = hi >> 16;
				if (lx0 > lx1 || ly0 > ly1) {
					continue; // empty
				}
				const uint32_t cells = ((lx1 >> 10) - (lx0 >> 10) + 1) * ((ly1 >> 10) - (ly0 >> 10) + 1);
				if (cells > BPP_WIDE_T) {
					S.wide[atomicAdd(&S.n_wide, 1u)] = (uint16_t)j;
				} else {
					for (uint32_t cy = ly0 >> 10; cy <= (ly1 >> 10); cy++) {
						for (uint32_t cx = lx0 >> 10; cx <= (lx1 >> 10); cx++) {
							const uint32_t s = cy * 32 + cx;
							if (s >= sc_first && s <= sc_last) {
								S.ent[atomicAdd(&S.cnt[s], 1u)] = (uint16_t)j;
							}
						}
					}
				}
			}
			__syncthreads();
			const uint32_t n_wide = S.n_wide;

			// ---- (d) warps take sub-cells; all 32 lanes share the sub-cell's entry list ----
			unsigned wlane = lane;
			asm volatile("" : "+r"(wlane)); // a register copy of the lane id (otherwise re-read from SR_TID per use)
			const cpp_args A = {boxes, build_box2d, pred, rec, px, py, index_base, out_probe, out_build, capacity, counter};
			while (true) {
				uint32_t sc = 0;
				if (lane == 0) {
					sc = atomicAdd(&S.next_sc, 1u);
				}
				sc = __shfl_sync(0xFFFFFFFFu, sc, 0);
				if (sc > sc_last) {
					break;
				}
				const uint32_t p0 = max(__ldg(coff + sc), unit.pbeg);
				const uint32_t p1 = min(__ldg(coff + sc + 1), unit.pend); // cell_off is contiguous across buckets
				const uint32_t lb = S.off[sc], n0 = (uint32_t)S.off[sc + 1] - lb;
				if (p0 >= p1 || n0 + n_wide == 0) {
					continue;
				}
				const uint32_t sx0 = (sc & 31u) << 10, sy0 = (sc >> 5) << 10;
				for (uint32_t pb = p0; pb < p1; pb += 32 * CPP_RB) {
					const uint32_t npts = min(p1 - pb, (uint32_t)(32 * CPP_RB));
					switch ((npts + 31u) >> 5) { // one fully unrolled instance per round count: no per-round branches
					case 1: cpp_batch<MODE, 1>(S, A, wlane, pb, npts, lb, n0, n_wide, sx0, sy0); break;
					case 2: cpp_batch<MODE, 2>(S, A, wlane, pb, npts, lb, n0, n_wide, sx0, sy0); break;
					case 3: cpp_batch<MODE, 3>(S, A, wlane, pb, npts, lb, n0, n_wide, sx0, sy0); break;
#if CPP_RB == 4
					default: cpp_batch<MODE, 4>(S, A, wlane, pb, npts, lb, n0, n_wide, sx0, sy0); break;
#else
					case 4: cpp_batch<MODE, 4>(S, A, wlane, pb, npts, lb, n0, n_wide, sx0, sy0); break;
					case 5: cpp_batch<MODE, 5>(S, A, wlane, pb, npts, lb, n0, n_wide, sx0, sy0); break;
					case 6: cpp_batch<MODE, 6>(S, A, wlane, pb, npts, lb, n0, n_wide, sx0, sy0); break;
					case 7: cpp_batch<MODE, 7>(S, A, wlane, pb, npts, lb, n0, n_wide, sx0, sy0); break;
					default: cpp_batch<MODE, 8>(S, A, wlane, pb, npts, lb, n0, n_wide, sx0, sy0); break;
#endif
					}
				}
			}
			c0 += ntry;
		}
	}
}

// NaN probes (bucket BP_NB) back to coordinate columns + ids for the tree walk
__global__ void __launch_bounds__(256) k_cp_unpack(const uint2 *__restrict__ rec, uint32_t n, const double *__restrict__ px,
                                                   const double *__restrict__ py, uint32_t index_base,
                                                   double *__restrict__ x, double *__restrict__ y,
                                                   uint32_t *__restrict__ ids) {
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		const uint32_t id = rec[i].y;
		x[i] = px[id - index_base];
		y[i] = py[id - index_base];
		ids[i] = id;
	}
}

// Front end shared by the cell-sorted probe and the point-in-polygon join (pip_join.cu): partitions the probe points
// into (bucket, sub-cell) cells of the tree box's quantiser grid and bins the tree's leaf level into the buckets.
#define SB_ERR_RETRY_EXACT (-1000)
static int cell_partition_impl(sb_context *ctx, const sb_flat *tree, const double *px, const double *py, uint64_t n,
                               uint32_t index_base, sb_cell_partition *P, bool want_xy, bool sampled);

int sbi_cell_partition(sb_context *ctx, const sb_flat *tree, const double *px, const double *py, uint64_t n,
                       uint32_t index_base, sb_cell_partition *P, bool want_xy) {
	const bool sampled = !want_xy && ctx->hist_sample && n >= CP_SAMPLE_MIN;
	int s = cell_partition_impl(ctx, tree, px, py, n, index_base, P, want_xy, sampled);
	if (s == SB_ERR_RETRY_EXACT) {
		// a bucket outgrew the region sized from the sample: once more with the exact histogram
		s = cell_partition_impl(ctx, tree, px, py, n, index_base, P, want_xy, false);
	}
	return s;
}

static int cell_partition_impl(sb_context *ctx, const sb_flat *tree, const double *px, const double *py, uint64_t n,
                               uint32_t index_base, sb_cell_partition *P, bool want_xy, bool sampled) {
	const uint32_t items = tree->h.item_count;
	bp_grid *d_grid = nullptr;
	bp_tables *d_tab = nullptr;
	bp_unit *d_units = nullptr;
	uint2 *d_rec_a = nullptr, *d_rec_b = nullptr;
	uint32_t *d_cand = nullptr, *d_fine = nullptr, *d_cell_off = nullptr;
	const uint64_t max_units = n / BP_SLICE + BP_NB + 1;
	// sampled regions: every bucket gets 1.25 x its scaled sample count + 4096 slots (the sample covers whole 4096-point
	// tiles, so the scaled total is at most n + CP_SAMPLE tiles)
	const uint64_t rec_cap = sampled ? (n + CP_SAMPLE * 4096ull) / 4 * 5 + (uint64_t)(BP_NB + 2) * 4096 : n;
	SB_ARG(rec_cap < 0xFFFFFFFFull, "cell probe: too many probe rows for 32-bit record positions");
	SB_TRY(ctx->scratch_t(1, &d_grid));
	SB_TRY(ctx->scratch_t(1, &d_tab));
	SB_TRY(ctx->scratch_t(max_units, &d_units));
	SB_TRY(ctx->scratch_t(rec_cap, &d_rec_a));
	SB_TRY(ctx->scratch_t(rec_cap, &d_rec_b));
	double2 *d_xy_a = nullptr;
	uint32_t *d_src = nullptr;
	if (want_xy) {
		SB_TRY(ctx->scratch_t(n, &d_xy_a));
		SB_TRY(ctx->scratch_t(n, &d_src));
	}
	SB_TRY(ctx->scratch_t(CP_NCELL, &d_fine));
	SB_TRY(ctx->scratch_t(CP_NCELL + 1, &d_cell_off));
	SB_TRY(sbi_zero(ctx, d_tab, sizeof(bp_tables)));
	SB_TRY(sbi_zero(ctx, d_fine, sizeof(uint32_t) * CP_NCELL));
	const sb_flat_header *d_hdr = static_cast<const sb_flat_header *>(tree->buffer);
	{
		sb_prof_scope ps(ctx, "bp_partition");
		k_bp_grid_init<<<1, 32, 0, ctx->stream>>>(d_hdr, d_grid);
		SB_LAUNCH_CHECK(ctx);
		if (sampled) {
			k_bp_hist_sample<<<sb_grid((n + 4095) / 4096 / CP_SAMPLE + 1, 1, 4), BPH_THREADS, 0, ctx->stream>>>(px, py, n, d_grid,
			                                                                                                 d_tab->pt_hist);
			SB_LAUNCH_CHECK(ctx);
			k_bp_inflate<<<1, 1024, 0, ctx->stream>>>(d_tab);
		} else {
			k_bp_hist<<<sb_grid(n, BPH_THREADS * 8, 4), BPH_THREADS, 0, ctx->stream>>>(px, py, n, d_grid, d_tab->pt_hist, nullptr);
		}
		SB_LAUNCH_CHECK(ctx);
	}
	{
		sb_prof_scope ps(ctx, "bp_cand");
		k_bp_cand<false><<<sb_grid(items, 256, 8), 256, 0, ctx->stream>>>(tree->boxes(), items, d_grid, d_tab->cand_cnt,
		                                                                 nullptr);
		SB_LAUNCH_CHECK(ctx);
		k_bp_scan<<<1, 1024, 0, ctx->stream>>>(d_tab, d_units);
		SB_LAUNCH_CHECK(ctx);
	}
	SB_TRACE(ctx, "cp: hist + cand count + scan");
	uint64_t totals[4];
	SB_TRY(sbi_read_words(ctx, d_tab->totals, 4, totals));
	const uint64_t n_cand = totals[1];
	uint64_t n_units = totals[2];
	SB_ARG(n_cand < 0xFFFFFFFFull, "cell probe: candidate directory too large");
	SB_TRY(ctx->scratch_t(n_cand ? n_cand : 1, &d_cand));
	SB_SMEM_OPTIN(ctx, k_cp_scatter1<false>, sizeof(cps1_smem<false>));
	SB_SMEM_OPTIN(ctx, k_cp_scatter2<false>, sizeof(cps2_smem<false>));
	SB_SMEM_OPTIN(ctx, k_cp_scatter1<true>, sizeof(cps1_smem<true>));
	SB_SMEM_OPTIN(ctx, k_cp_scatter2<true>, sizeof(cps2_smem<true>));
	{
		sb_prof_scope ps(ctx, "bp_partition");
		if (want_xy) {
			k_cp_scatter1<true><<<sb_grid(n, CP_TILE1, 1), 1024, sizeof(cps1_smem<true>), ctx->stream>>>(
			    px, py, n, index_base, d_grid, d_tab->pt_cursor, d_fine, d_rec_a, d_xy_a);
		} else if (sampled) {
			SB_SMEM_OPTIN(ctx, (k_cp_scatter1<false, true, false>), sizeof(cps1_smem<false>));
			k_cp_scatter1<false, true, false><<<sb_grid(n, CP_TILE1, 1), 1024, sizeof(cps1_smem<false>), ctx->stream>>>(
			    px, py, n, index_base, d_grid, d_tab->pt_cursor, d_fine, d_rec_a, nullptr, d_tab->pt_off + 1);
			SB_LAUNCH_CHECK(ctx);
			// exact counts, units and totals from the region cursors; did everything fit?
			k_cp_actual<<<1, 1024, 0, ctx->stream>>>(d_tab, d_units);
		} else {
			k_cp_scatter1<false><<<sb_grid(n, CP_TILE1, 1), 1024, sizeof(cps1_smem<false>), ctx->stream>>>(
			    px, py, n, index_base, d_grid, d_tab->pt_cursor, d_fine, d_rec_a, nullptr);
		}
		SB_LAUNCH_CHECK(ctx);
		if (sampled) {
			static_assert(offsetof(bp_tables, pad) % 8 == 0 && offsetof(bp_tables, totals) == offsetof(bp_tables, pad) + 8,
			              "pad (overflow flag) and totals are read as five aligned 64-bit words");
			uint64_t w[5]; // pad (overflow flag; the struct was cleared, so the compiler's padding word is 0), totals[0..3]
			SB_TRY(sbi_read_words(ctx, &d_tab->pad, 5, w));
			if ((uint32_t)w[0] != 0) {
				return SB_ERR_RETRY_EXACT;
			}
			totals[0] = w[1];
			n_units = totals[2] = w[3];
			totals[3] = w[4];
			if (n_units) {
				k_cp_finehist<<<sb_grid(BP_UNIT_TILES * n_units, 1, 2), 1024, 0, ctx->stream>>>(d_tab, d_units, d_rec_a, d_fine);
				SB_LAUNCH_CHECK(ctx);
			}
		}
		// (cell_off is needed even without units: consumers index it for every cell)
		k_cp_scan2<<<BP_NB, 1024, 0, ctx->stream>>>(d_tab, d_fine, d_cell_off);
		SB_LAUNCH_CHECK(ctx);
		if (n_units) {
			if (want_xy) {
				k_cp_scatter2<true><<<sb_grid(BP_UNIT_TILES * n_units, 1, 1), 1024, sizeof(cps2_smem<true>), ctx->stream>>>(
				    d_tab, d_units, d_rec_a, d_fine, d_rec_b, d_src);
			} else {
				k_cp_scatter2<false><<<sb_grid(BP_UNIT_TILES * n_units, 1, 1), 1024, sizeof(cps2_smem<false>), ctx->stream>>>(
				    d_tab, d_units, d_rec_a, d_fine, d_rec_b, nullptr);
			}
			SB_LAUNCH_CHECK(ctx);
		}
	}
	{
		sb_prof_scope ps(ctx, "bp_cand");
		k_bp_cand<true><<<sb_grid(items, 256, 8), 256, 0, ctx->stream>>>(tree->boxes(), items, d_grid, d_tab->cand_cursor,
		                                                                d_cand);
		SB_LAUNCH_CHECK(ctx);
	}
	SB_TRACE(ctx, "cp: scatter passes + cand fill");
	P->grid = d_grid;
	P->tables = d_tab;
	P->units = d_units;
	P->cand_pos = d_cand;
	P->rec_bucketed = d_rec_a;
	P->rec = d_rec_b;
	P->xy_bucketed = d_xy_a;
	P->src = d_src;
	P->cell_off = d_cell_off;
	P->n_kept = totals[0];
	P->n_cand = n_cand;
	P->n_units = n_units;
	P->n_nan = totals[3];
	return SB_OK;
}

int sbi_flat_probe_cells(sb_context *ctx, const sb_flat *tree, const sb_probe_args &a, uint32_t *out_probe,
                         uint32_t *out_build, uint64_t capacity, uint64_t *count) {
	*count = 0;
	const uint64_t n = a.m;
	const uint32_t items = tree->h.item_count;
	if (n == 0 || items == 0) {
		return SB_OK;
	}
	SB_ARG(a.px != nullptr && a.py != nullptr, "cell probe takes points");
	SB_ARG(n + a.index_base <= 0xFFFFFFFFull, "probe indices must fit 32 bits");
	const int mode = a.pred >= 0 ? 2 : 1;
	SB_ARG(mode == 1 || a.build_box2d != nullptr, "build_box2d");

	unsigned long long *d_counter = nullptr;
	SB_TRY(ctx->scratch_t(1, &d_counter));
	SB_TRY(sbi_zero(ctx, d_counter, sizeof(*d_counter)));
	sb_cell_partition P;
	SB_TRY(sbi_cell_partition(ctx, tree, a.px, a.py, n, a.index_base, &P));
	const bp_grid *d_grid = static_cast<const bp_grid *>(P.grid);
	bp_tables *d_tab = static_cast<bp_tables *>(P.tables);
	const bp_unit *d_units = static_cast<const bp_unit *>(P.units);
	const uint32_t *d_cand = P.cand_pos, *d_cell_off = P.cell_off;
	const uint2 *d_rec_a = static_cast<const uint2 *>(P.rec_bucketed), *d_rec_b = static_cast<const uint2 *>(P.rec);
	const uint64_t n_units = P.n_units, n_nan = P.n_nan;
	uint64_t totals[1] = {P.n_kept};
	SB_SMEM_OPTIN(ctx, k_cp_probe<1>, sizeof(cpp_smem<1>));
	SB_SMEM_OPTIN(ctx, k_cp_probe<2>, sizeof(cpp_smem<2>));
	if (n_units) {
		const uint64_t ctas = (uint64_t)SB_NUM_SMS * CPP_CTAS;
		const unsigned grid = (unsigned)(n_units < ctas ? n_units : ctas);
		sb_prof_scope ps(ctx, "flat_probe");
		if (mode == 1) {
			k_cp_probe<1><<<grid, CPP_THREADS, sizeof(cpp_smem<1>), ctx->stream>>>(
			    tree->boxes(), tree->idx(), nullptr, -1, d_grid, d_tab, d_units, d_cand, d_rec_b, d_cell_off, a.px, a.py,
			    a.index_base, out_probe, out_build, capacity, d_counter);
		} else {
			k_cp_probe<2><<<grid, CPP_THREADS, sizeof(cpp_smem<2>), ctx->stream>>>(
			    tree->boxes(), tree->idx(), a.build_box2d, a.pred, d_grid, d_tab, d_units, d_cand, d_rec_b, d_cell_off, a.px,
			    a.py, a.index_base, out_probe, out_build, capacity, d_counter);
		}
		SB_LAUNCH_CHECK(ctx);
	}
	uint64_t c1 = 0;
	SB_TRY(sbi_read_u64(ctx, reinterpret_cast<const uint64_t *>(d_counter), &c1));
	SB_TRACE(ctx, "cp: probe kernel");
	*count = c1;
	if (n_nan && mode == 1) {
		// NaN coordinates: Box::Intersects never rejects on a NaN axis. Such rows take the per-probe tree walk.
		double *nx = nullptr, *ny = nullptr;
		uint32_t *nid = nullptr;
		SB_TRY(ctx->scratch_t(n_nan, &nx));
		SB_TRY(ctx->scratch_t(n_nan, &ny));
		SB_TRY(ctx->scratch_t(n_nan, &nid));
		k_cp_unpack<<<sb_grid(n_nan, 256, 8), 256, 0, ctx->stream>>>(d_rec_a + totals[0], (uint32_t)n_nan, a.px, a.py,
		                                                            a.index_base, nx, ny, nid);
		SB_LAUNCH_CHECK(ctx);
		sb_probe_args b;
		b.px = nx;
		b.py = ny;
		b.m = n_nan;
		b.ids = nid;
		b.force_walk = true;
		uint64_t c2 = 0;
		const uint64_t done = c1 < capacity ? c1 : capacity;
		int s = sbi_flat_probe(ctx, tree, b, out_probe + done, out_build + done, capacity - done, &c2);
		*count = c1 + c2;
		if (s != SB_OK && s != SB_ERR_CAPACITY) {
			return s;
		}
	}
	if (*count > capacity) {
		sb_set_error("probe: %llu pairs, capacity %llu", (unsigned long long)*count, (unsigned long long)capacity);
		return SB_ERR_CAPACITY;
	}
	return SB_OK;
}
