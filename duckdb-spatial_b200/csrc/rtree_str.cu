// rtree_str.cu -- K4a: STR bulk load of the persistent RTREE + batched index scan.
//
// Reference algorithm (see include/spatial_b200.h for file:line): ORDER BY x-centre, then per layer: runs of
// `slice` entries re-sorted by fp32 y-centre, packed `cap` per node, leaf nodes sorted by row id, node box = union.
// B200 formulation --
//   x-order : 32-bit key  = order-preserving bits of float(min_x + max_x)   (x0.5 in fp64 is monotone and exact). The
//             layer is never sorted by it: the slices of the x order come from a radix select of the slice-boundary keys
//             and one stable partition of the 32-byte records (str_partition.cu)
//   y-order : 32-bit key  = order-preserving bits of float((min_y+max_y)/2), sorted per slice of the x order by the
//             segmented radix sort (segment = `slice` consecutive positions of the layer); equal y keys of a slice
//             stay in (x bin, input position) order
// followed by a pack kernel (one CTA per node: gather, rank-sort by row id in shared memory, union, coalesced
// 24-byte entry stores). Level sizes are known on the host from n and cap, so the whole build is enqueued
// without a single device->host synchronisation.
#include "internal.cuh"

#include <math.h>

// x keys for the ORDER BY and, from the same read of the boxes, the y keys of the leaf level in input order: the
// leaf level later picks its y key with a 4-byte gather from this 4n-byte array (L2-resident at 10 M keys) instead
// of a 16-byte gather from the boxes.
// hist1 (may be NULL): the 2048-bin histogram of the x keys' top 11 bits, level 1 of the cut-key select (str_partition.cu).
template <bool POINTS>
__global__ void __launch_bounds__(256) k_str_xkeys(const sb_boxsrc box, uint64_t n,
                                                   uint32_t *__restrict__ keys, uint32_t *__restrict__ ykeys,
                                                   uint32_t *__restrict__ hist1) {
	__shared__ uint32_t s_h[4 * 2048]; // 4 replicas (lane & 3): float keys share their top bits
	if (hist1) {
		for (uint32_t i = threadIdx.x; i < 4 * 2048; i += 256) {
			s_h[i] = 0;
		}
		__syncthreads();
	}
	uint32_t *mine = s_h + (threadIdx.x & 3u) * 2048;
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		const float4 b = box.template load<POINTS>(i);
		// ST_Centroid::ExecuteBox<float>: (float)min_x + (float)max_x in fp32, spatial_functions_scalar.cpp:1490-1494
		const uint32_t xk = sb_f32_sort_key(__fadd_rn(b.x, b.z));
		keys[i] = xk;
		// bounds.Center().y = (min.y + max.y) / 2 in fp32, bbox.hpp:78-80, rtree_index_create_physical.cpp:172-174
		ykeys[i] = sb_f32_sort_key(__fdiv_rn(__fadd_rn(b.y, b.w), 2.0f));
		if (hist1) {
			atomicAdd(&mine[xk >> 21], 1u);
		}
	}
	if (hist1) {
		__syncthreads();
		for (uint32_t d = threadIdx.x; d < 2048; d += 256) {
			const uint32_t c = s_h[d] + s_h[2048 + d] + s_h[4096 + d] + s_h[6144 + d];
			if (c) {
				atomicAdd(&hist1[d], c);
			}
		}
	}
}

// level 0: entries are addressed through the x-order permutation
__global__ void __launch_bounds__(256) k_str_ykeys_leaf(const uint32_t *__restrict__ ykeys_in,
                                                        const uint32_t *__restrict__ perm, uint64_t n,
                                                        uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) {
		const uint32_t src = perm[j];
		keys[j] = ykeys_in[src]; // the slice is the segment of the segmented sort, not part of the key
		vals[j] = src;
	}
}

__global__ void __launch_bounds__(256) k_str_ykeys_upper(const sb_entry *__restrict__ cur, uint64_t n,
                                                         uint32_t *__restrict__ keys) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) {
		const float yc = __fdiv_rn(__fadd_rn(cur[j].box[1], cur[j].box[3]), 2.0f);
		keys[j] = sb_f32_sort_key(yc);
	}
}

// Upper levels, one CTA per node: gather already-built entries in sorted order, write the node's entries (24 B each,
// staged in shared memory, stored as coalesced 8-byte words) and one parent entry {(type<<56)|node index, union box}.
__global__ void __launch_bounds__(256)
    k_str_pack_upper(const sb_entry *__restrict__ cur, const uint32_t *__restrict__ perm, uint64_t count, uint32_t cap,
                     sb_entry *__restrict__ level_out, sb_entry *__restrict__ parent_out, uint64_t parent_type) {
	__shared__ uint64_t s_words[256 * 3];
	__shared__ float s_red[8][4];
	const uint64_t node = blockIdx.x;
	const uint64_t beg = node * cap;
	const uint32_t cnt = (uint32_t)min((uint64_t)cap, count - beg);
	const unsigned t = threadIdx.x;
	sb_entry e;
	e.ptr = 0;
	e.box[0] = e.box[1] = FLT_MAX;
	e.box[2] = e.box[3] = -FLT_MAX;
	if (t < cnt) {
		e = cur[perm[beg + t]];
		s_words[3 * t + 0] = e.ptr;
		s_words[3 * t + 1] = ((uint64_t)__float_as_uint(e.box[1]) << 32) | __float_as_uint(e.box[0]);
		s_words[3 * t + 2] = ((uint64_t)__float_as_uint(e.box[3]) << 32) | __float_as_uint(e.box[2]);
	}
	// union of the node's boxes (RTreeNode::GetBounds, rtree_node.hpp:92-99): fp32 min/max
	float u0 = e.box[0], u1 = e.box[1], u2 = e.box[2], u3 = e.box[3];
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		u0 = fminf(u0, __shfl_xor_sync(0xFFFFFFFFu, u0, s));
		u1 = fminf(u1, __shfl_xor_sync(0xFFFFFFFFu, u1, s));
		u2 = fmaxf(u2, __shfl_xor_sync(0xFFFFFFFFu, u2, s));
		u3 = fmaxf(u3, __shfl_xor_sync(0xFFFFFFFFu, u3, s));
	}
	if (sb_lane() == 0) {
		s_red[t >> 5][0] = u0, s_red[t >> 5][1] = u1, s_red[t >> 5][2] = u2, s_red[t >> 5][3] = u3;
	}
	__syncthreads();
	uint64_t *out_words = reinterpret_cast<uint64_t *>(level_out + beg);
	for (uint32_t w = t; w < cnt * 3; w += blockDim.x) {
		out_words[w] = s_words[w];
	}
	if (t == 0) {
		const unsigned nw = (blockDim.x + 31) >> 5;
		for (unsigned w = 1; w < nw; w++) {
			u0 = fminf(u0, s_red[w][0]);
			u1 = fminf(u1, s_red[w][1]);
			u2 = fmaxf(u2, s_red[w][2]);
			u3 = fmaxf(u3, s_red[w][3]);
		}
		sb_entry p;
		p.ptr = (parent_type << 56) | node;
		p.box[0] = u0, p.box[1] = u1, p.box[2] = u2, p.box[3] = u3;
		parent_out[node] = p;
	}
}

// Leaf pack, one WARP per node (lane l holds entries l, l + 32, ... of the node): the CTA-per-node form above spends
// ~570 warp instructions per 32 entries on block-wide reductions, barriers and address arithmetic that do not grow
// with the entries a lane holds (178 M warp instructions at C1 = instruction bound at 234 us for 572 MB of traffic).
// Same result: entries gathered through `perm` (from the partitioned records `rec` when given, else from the input
// columns), ordered by row id (RTreeNode::SortEntriesByRowId, rtree_node.hpp:151-155) with the bucketed rank described
// above, stored as coalesced 8-byte words; one parent entry {(LEAF_PAGE << 56) | node, union box} per node.
#define PKW_WARPS 4
#define PKW_CNT(b) ((b) + ((b) >> 3))
#ifndef PKW_MIN_CTAS
#define PKW_MIN_CTAS 8
#endif
template <int PER>
__global__ void __launch_bounds__(PKW_WARPS * 32, PKW_MIN_CTAS)
    k_str_pack_leaf(const float4 *__restrict__ box, const int64_t *__restrict__ row_ids, const sb_rec32 *__restrict__ rec,
                    const uint32_t *__restrict__ perm, uint64_t count, uint32_t cap, uint64_t nodes,
                    sb_entry *__restrict__ level_out, sb_entry *__restrict__ parent_out, uint64_t node_base) {
	// per warp: the node's output words; the ranking scratch (bucket counters padded so that lane * 8 + q falls in 32
	// different banks, keys and sources by bucket) lives in the same storage -- it is dead when the words are written
	__shared__ uint64_t s_words_all[PKW_WARPS][PER * 32 * 3];
	static_assert((PKW_CNT(256) + 1 + PER * 32 + PER * 8) * 4 <= PER * 32 * 3 * 8, "ranking scratch fits the output words");
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
	const uint64_t node = (uint64_t)blockIdx.x * PKW_WARPS + warp;
	if (node >= nodes) {
		return; // warps are independent: no block barrier below
	}
	uint64_t *s_words = s_words_all[warp];
	int64_t *s_row = reinterpret_cast<int64_t *>(s_words); // only read before the words are written (slow rank path)
	uint32_t *s_cnt = reinterpret_cast<uint32_t *>(s_words), *s_key = s_cnt + PKW_CNT(256) + 1;
	uint8_t *s_src = reinterpret_cast<uint8_t *>(s_key + PER * 32);
	const uint64_t beg = node * cap;
	const uint32_t cnt = (uint32_t)min((uint64_t)cap, count - beg);
	float4 bx[PER];
	int64_t row[PER];
#pragma unroll
	for (int j = 0; j < PER; j++) {
		const uint32_t i = j * 32 + lane;
		bx[j] = make_float4(FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX);
		row[j] = 0;
		if (i < cnt) {
			const uint32_t src = perm[beg + i];
			if (rec) { // partitioned records: `src` is a position inside this node's slice (local gather)
				const float4 *r = reinterpret_cast<const float4 *>(rec + src);
				bx[j] = r[0];
				const float4 hi = r[1];
				row[j] = (int64_t)(((uint64_t)__float_as_uint(hi.y) << 32) | __float_as_uint(hi.x));
			} else {
				bx[j] = box[src];
				row[j] = row_ids[src];
			}
			row[j] = (int64_t)((uint64_t)row[j] << 8) >> 8; // RTree::MakeRowId keeps 56 bits (type ROW_ID = 0)
		}
	}
	// node minimum / maximum of the row ids
	unsigned long long mn = ~0ull, mx = 0ull;
#pragma unroll
	for (int j = 0; j < PER; j++) {
		if (j * 32 + lane < cnt) {
			const unsigned long long u = (unsigned long long)row[j] ^ 0x8000000000000000ull;
			mn = u < mn ? u : mn;
			mx = u > mx ? u : mx;
		}
	}
	{
		// warp minimum / maximum: when every lane's values share their upper 32 bits (row ids of a node are almost always
		// within 2^32 of each other) the hardware 32-bit warp reductions do it in two instructions, else 64-bit shuffles
		const uint32_t hi0 = __shfl_sync(0xFFFFFFFFu, (uint32_t)(mn >> 32), 0); // lane 0 always holds entry 0
		const bool same_hi = (cnt <= lane || ((uint32_t)(mn >> 32) == hi0 && (uint32_t)(mx >> 32) == hi0));
		if (__all_sync(0xFFFFFFFFu, same_hi)) {
			const uint32_t lo_mn = __reduce_min_sync(0xFFFFFFFFu, cnt <= lane ? 0xFFFFFFFFu : (uint32_t)mn);
			const uint32_t lo_mx = __reduce_max_sync(0xFFFFFFFFu, cnt <= lane ? 0u : (uint32_t)mx);
			mn = ((unsigned long long)hi0 << 32) | lo_mn;
			mx = ((unsigned long long)hi0 << 32) | lo_mx;
		} else {
#pragma unroll
			for (int s = 16; s > 0; s >>= 1) {
				const unsigned long long a = __shfl_xor_sync(0xFFFFFFFFu, mn, s), b = __shfl_xor_sync(0xFFFFFFFFu, mx, s);
				mn = a < mn ? a : mn;
				mx = b > mx ? b : mx;
			}
		}
	}
	for (uint32_t i = lane; i < PKW_CNT(256) + 1; i += 32) {
		s_cnt[i] = 0;
	}
	__syncwarp();
	const unsigned long long range = mx - mn;
	const bool fast = range < 0xFFFFFFFFull;
	uint32_t k32[PER], bs[PER]; // bs = bucket | slot << 8
	bool crowded = false;
	if (fast) {
		const float scale = __fdividef(256.0f, __ull2float_rz(range) + 1.0f); // any positive factor keeps the map monotone
#pragma unroll
		for (int j = 0; j < PER; j++) {
			k32[j] = 0;
			bs[j] = 0;
			if (j * 32 + lane < cnt) {
				k32[j] = (uint32_t)(((unsigned long long)row[j] ^ 0x8000000000000000ull) - mn);
				const uint32_t b = min(255u, __float2uint_rz(__uint2float_rz(k32[j]) * scale));
				const uint32_t slot = atomicAdd(&s_cnt[PKW_CNT(b)], 1u);
				bs[j] = b | (slot << 8);
				crowded |= slot >= 8;
			}
		}
	}
	crowded = __any_sync(0xFFFFFFFFu, crowded);
	uint32_t rank[PER];
	if (fast && !crowded) {
		{ // exclusive scan of the 256 counts: 8 per lane
			uint32_t c[8], sum = 0;
#pragma unroll
			for (int q = 0; q < 8; q++) {
				c[q] = s_cnt[lane * 9 + q];
				sum += c[q];
			}
			uint32_t inc = sum;
#pragma unroll
			for (int s = 1; s < 32; s <<= 1) {
				const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
				if (lane >= (unsigned)s) {
					inc += o;
				}
			}
			uint32_t run = inc - sum;
			__syncwarp();
#pragma unroll
			for (int q = 0; q < 8; q++) {
				s_cnt[lane * 9 + q] = run;
				run += c[q];
			}
			if (lane == 31) {
				s_cnt[PKW_CNT(256)] = run;
			}
		}
		__syncwarp();
#pragma unroll
		for (int j = 0; j < PER; j++) {
			if (j * 32 + lane < cnt) {
				const uint32_t pos = s_cnt[PKW_CNT(bs[j] & 255u)] + (bs[j] >> 8);
				s_key[pos] = k32[j];
				s_src[pos] = (uint8_t)(j * 32 + lane);
			}
		}
		__syncwarp();
#pragma unroll
		for (int j = 0; j < PER; j++) {
			rank[j] = 0;
			if (j * 32 + lane < cnt) {
				const uint32_t b = bs[j] & 255u, me = j * 32 + lane;
				const uint32_t b0 = s_cnt[PKW_CNT(b)], b1 = s_cnt[PKW_CNT(b + 1)];
				uint32_t r = b0;
				for (uint32_t q = b0; q < b1; q++) {
					const uint32_t kk = s_key[q];
					r += (kk < k32[j]) || (kk == k32[j] && s_src[q] < me);
				}
				rank[j] = r;
			}
		}
	} else {
		// row ids spread over 2^32 or crowded into a bucket: 64-bit all-pairs count, equal row ids by position
#pragma unroll
		for (int j = 0; j < PER; j++) {
			if (j * 32 + lane < cnt) {
				s_row[j * 32 + lane] = row[j];
			}
		}
		__syncwarp();
#pragma unroll
		for (int j = 0; j < PER; j++) {
			rank[j] = 0;
			if (j * 32 + lane < cnt) {
				const uint32_t me = j * 32 + lane;
				uint32_t r = 0;
				for (uint32_t q = 0; q < cnt; q++) {
					const int64_t o = s_row[q];
					r += (o < row[j]) || (o == row[j] && q < me);
				}
				rank[j] = r;
			}
		}
	}
	__syncwarp();
	float u0 = FLT_MAX, u1 = FLT_MAX, u2 = -FLT_MAX, u3 = -FLT_MAX;
#pragma unroll
	for (int j = 0; j < PER; j++) {
		if (j * 32 + lane < cnt) {
			s_words[3 * rank[j] + 0] = (uint64_t)row[j] & 0x00FFFFFFFFFFFFFFull;
			s_words[3 * rank[j] + 1] = ((uint64_t)__float_as_uint(bx[j].y) << 32) | __float_as_uint(bx[j].x);
			s_words[3 * rank[j] + 2] = ((uint64_t)__float_as_uint(bx[j].w) << 32) | __float_as_uint(bx[j].z);
		}
		// union of the node's boxes (RTreeNode::GetBounds, rtree_node.hpp:92-99): fp32 min/max; absent entries are neutral
		u0 = fminf(u0, bx[j].x), u1 = fminf(u1, bx[j].y), u2 = fmaxf(u2, bx[j].z), u3 = fmaxf(u3, bx[j].w);
	}
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		u0 = fminf(u0, __shfl_xor_sync(0xFFFFFFFFu, u0, s));
		u1 = fminf(u1, __shfl_xor_sync(0xFFFFFFFFu, u1, s));
		u2 = fmaxf(u2, __shfl_xor_sync(0xFFFFFFFFu, u2, s));
		u3 = fmaxf(u3, __shfl_xor_sync(0xFFFFFFFFu, u3, s));
	}
	__syncwarp();
	uint64_t *out_words = reinterpret_cast<uint64_t *>(level_out + beg);
	for (uint32_t w = lane; w < cnt * 3; w += 32) {
		out_words[w] = s_words[w];
	}
	if (lane == 0) {
		sb_entry p;
		p.ptr = (1ull << 56) | (node_base + node);
		p.box[0] = u0, p.box[1] = u1, p.box[2] = u2, p.box[3] = u3;
		parent_out[node] = p;
	}
}

// Small upper levels (<= STR_SMALL entries) in ONE launch instead of eight: y keys, the sort inside the slices and the
// pack of a whole level by a single CTA. The order is the one the segmented radix sort produces -- by (slice, y key),
// ties in input order -- because the index is the last part of a unique 64-bit key: slice (11 bits) | y key (32) |
// index (11), sorted with a bitonic network in shared memory. One thread per node then streams its entries out and
// takes the union box in entry order (fp32 min / max are order-free).
#define STR_SMALL 2048
__global__ void __launch_bounds__(1024)
    k_str_small_level(const sb_entry *__restrict__ cur, uint32_t count, uint32_t cap, uint64_t slice,
                      sb_entry *__restrict__ level_out, sb_entry *__restrict__ parent_out, uint64_t parent_type) {
	__shared__ uint64_t s_key[STR_SMALL];
	const unsigned t = threadIdx.x;
	uint32_t sz = 2; // the network only spans the next power of two (a 5-entry level is not a 2048-entry sort)
	while (sz < count) {
		sz <<= 1;
	}
	for (uint32_t i = t; i < sz; i += blockDim.x) {
		uint64_t k = ~0ull; // padding sorts last
		if (i < count) {
			const float yc = __fdiv_rn(__fadd_rn(cur[i].box[1], cur[i].box[3]), 2.0f);
			k = ((uint64_t)(i / slice) << 43) | ((uint64_t)sb_f32_sort_key(yc) << 11) | i;
		}
		s_key[i] = k;
	}
	__syncthreads();
	for (uint32_t size = 2; size <= sz; size <<= 1) {
		for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
			for (uint32_t i = t; i < sz / 2; i += blockDim.x) {
				const uint32_t lo = 2 * i - (i & (stride - 1));
				const uint32_t hi = lo + stride;
				const bool up = (lo & size) == 0;
				const uint64_t a = s_key[lo], b = s_key[hi];
				if ((a > b) == up) {
					s_key[lo] = b;
					s_key[hi] = a;
				}
			}
			__syncthreads();
		}
	}
	// entries out in sorted order (all threads), then one warp per node takes the union box (fp32 min / max are order-free)
	for (uint32_t j = t; j < count; j += blockDim.x) {
		level_out[j] = cur[s_key[j] & 2047u];
	}
	__syncthreads();
	const uint32_t nodes = (count + cap - 1) / cap;
	for (uint32_t node = t >> 5; node < nodes; node += blockDim.x >> 5) {
		const uint32_t beg = node * cap, end = min(beg + cap, count);
		float u0 = FLT_MAX, u1 = FLT_MAX, u2 = -FLT_MAX, u3 = -FLT_MAX;
		for (uint32_t j = beg + sb_lane(); j < end; j += 32) {
			const sb_entry e = level_out[j];
			u0 = fminf(u0, e.box[0]), u1 = fminf(u1, e.box[1]), u2 = fmaxf(u2, e.box[2]), u3 = fmaxf(u3, e.box[3]);
		}
#pragma unroll
		for (int s = 16; s > 0; s >>= 1) {
			u0 = fminf(u0, __shfl_xor_sync(0xFFFFFFFFu, u0, s));
			u1 = fminf(u1, __shfl_xor_sync(0xFFFFFFFFu, u1, s));
			u2 = fmaxf(u2, __shfl_xor_sync(0xFFFFFFFFu, u2, s));
			u3 = fmaxf(u3, __shfl_xor_sync(0xFFFFFFFFu, u3, s));
		}
		if (sb_lane() == 0) {
			sb_entry p;
			p.ptr = (parent_type << 56) | node;
			p.box[0] = u0, p.box[1] = u1, p.box[2] = u2, p.box[3] = u3;
			parent_out[node] = p;
		}
	}
}

// level sizes, offsets and the slice size are pure functions of (n, cap)
static void str_shape(sb_rtree *t, uint64_t n, uint32_t cap) {
	// Finalize: slice = ceil(sqrt(ceil(n / cap))) * cap, rtree_index_create_physical.cpp:349-351
	t->slice = (uint64_t)ceil(sqrt((double)((n + cap - 1) / cap))) * cap;
	uint64_t cnt = n;
	int nl = 0;
	uint64_t total = 0;
	while (true) {
		t->counts[nl] = cnt;
		t->offsets[nl] = total;
		total += cnt;
		nl++;
		if (cnt == 1 && nl > 1) {
			break;
		}
		cnt = (cnt + cap - 1) / cap; // n == 1: a lone row id is wrapped into a leaf page (:241-246)
	}
	t->nlevels = nl;
	t->total = total;
}

// Leaf level of (a shard of) the layer: slices of the x order (select + stable partition, str_partition.cu), y-sort
// inside runs of `slice` positions, pack `cap` per node with the entries ordered by row id. level0 receives the n leaf
// entries, parents one entry per node whose pointer is (LEAF_PAGE << 56) | (node_base + node). With node_base = 0 and the
// whole layer this is the first stage of the build; a shard that holds a contiguous, slice-aligned range of the global x
// order (rows in ascending global input position, x_key_before / x_key_after = the global cut keys at its two ends,
// -1 at the ends of the layer) produces exactly its part of the global leaf level (multi-GPU build).
static int str_leaf_level(sb_context *ctx, sb_boxsrc src, const int64_t *row_ids, uint64_t n, uint32_t cap,
                          uint64_t slice, uint64_t node_base, int64_t x_key_before, int64_t x_key_after, sb_entry *level0,
                          sb_entry *parents, uint32_t **yk_out, uint32_t **ykb_out, uint32_t **yv_out, uint32_t **yvb_out) {
	// the one-pass record partition reads every box once, in input order: a POINT_2D source is rounded on the fly there.
	// The bin-id sort + gather form gathers boxes at random: it gets a materialised box column.
	const uint32_t ncut_host = (uint32_t)((n + slice - 1) / slice - 1) + (x_key_before >= 0 ? 1u : 0u) + (x_key_after >= 0 ? 1u : 0u);
	const bool one_pass = 2 * (uint64_t)ncut_host + 1 <= SP_MAXBINS && ctx->str_path != 1;
	sb_boxsrc box = src;
	if (!box.box && !one_pass) {
		float4 *mat = nullptr;
		SB_TRY(ctx->scratch_t(n, &mat));
		SB_TRY(sbi_bbox_points(ctx, src.x, src.y, n, reinterpret_cast<float *>(mat)));
		box.box = mat;
	}
	uint32_t *xk = nullptr, *yk_in = nullptr;
	uint32_t *yk = nullptr, *ykb = nullptr, *yv = nullptr, *yvb = nullptr;
	SB_TRY(ctx->scratch_t(n, &xk));
	SB_TRY(ctx->scratch_t(n, &yk_in));
	SB_TRY(ctx->scratch_t(n, &yk));
	SB_TRY(ctx->scratch_t(n, &ykb));
	SB_TRY(ctx->scratch_t(n, &yv));
	SB_TRY(ctx->scratch_t(n, &yvb));
	// --- slices of the x order: x / y keys (and the first histogram of the select), then the cut keys ------------
	sb_str_select sel;
	SB_TRY(sbi_str_select_begin(ctx, n, slice, x_key_before, x_key_after, &sel));
	{
		sb_prof_scope ps(ctx, "str_keys");
		if (box.box) {
			k_str_xkeys<false><<<sb_grid(n, 256 * 8, 8), 256, 0, ctx->stream>>>(box, n, xk, yk_in, sel.hist1);
		} else {
			k_str_xkeys<true><<<sb_grid(n, 256 * 8, 8), 256, 0, ctx->stream>>>(box, n, xk, yk_in, sel.hist1);
		}
		SB_LAUNCH_CHECK(ctx);
	}
	SB_TRY(sbi_str_select_finish(ctx, xk, n, &sel));
	uint32_t *cuts = sel.cuts;
	const uint32_t ncut = sel.ncut;
	const sb_rec32 *rec = nullptr;
	int alt = 0;
	bool iota = true;
	if (one_pass) {
		// one pass: the records themselves go to their place in the slice order, y keys beside them
		sb_rec32 *r = nullptr;
		SB_TRY(ctx->scratch_t(n, &r));
		SB_TRY(sbi_str_partition_rec(ctx, box, row_ids, xk, yk_in, n, cuts, ncut, r, yk));
		rec = r;
	} else {
		// many slices: the bin ids are sorted as small integers (stable, 8 bits per pass) and the entries are addressed
		// through the permutation, as a sorted layer would be
		int bits = 0;
		while ((1ull << bits) < 2ull * ncut + 1) {
			bits++;
		}
		SB_TRY(sbi_str_bin_ids(ctx, xk, n, cuts, ncut, yk));
		SB_TRY(sbi_radix_sort_u32(ctx, yk, ykb, yv, yvb, n, 0, bits, true, &alt));
		const uint32_t *xperm = alt ? yvb : yv;
		uint32_t *other = alt ? yv : yvb;
		{
			sb_prof_scope ps(ctx, "str_keys");
			k_str_ykeys_leaf<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(yk_in, xperm, n, yk, other);
			SB_LAUNCH_CHECK(ctx);
		}
		yv = other;
		yvb = const_cast<uint32_t *>(xperm);
		iota = false;
	}
	// --- y order inside slices, pack leaves ---------------------------------------------------------------------
	const uint32_t *perm;
	if (rec && slice % cap == 0 && slice / cap <= 512 && ctx->str_path != 2) {
		// leaf nodes are re-ordered by row id when packed: only the node an entry falls in matters, not the y order inside
		// it -- the y keys at the node boundaries of every slice + one stable partition of the slice's positions
		SB_TRY(sbi_str_node_partition(ctx, yk, n, slice, cap, yv));
		perm = yv;
	} else {
		// segmented sort: every slice of the x order is sorted by y on its own (4 passes over 32-bit keys; a 64-bit
		// (slice, y) key sort needed 6)
		SB_TRY(sbi_radix_sort_u32_segmented(ctx, yk, ykb, yv, yvb, n, slice, 0, 32, iota, &alt));
		perm = alt ? yvb : yv;
	}
	const uint64_t nodes = (n + cap - 1) / cap;
	{
		sb_prof_scope ps(ctx, "str_pack");
		const unsigned grid = (unsigned)((nodes + PKW_WARPS - 1) / PKW_WARPS);
		if (cap <= 128) {
			k_str_pack_leaf<4><<<grid, PKW_WARPS * 32, 0, ctx->stream>>>(box.box, row_ids, rec, perm, n, cap, nodes, level0, parents,
			                                                            node_base);
		} else {
			k_str_pack_leaf<8><<<grid, PKW_WARPS * 32, 0, ctx->stream>>>(box.box, row_ids, rec, perm, n, cap, nodes, level0, parents,
			                                                            node_base);
		}
		SB_LAUNCH_CHECK(ctx);
	}
	*yk_out = yk, *ykb_out = ykb, *yv_out = yv, *yvb_out = yvb;
	return SB_OK;
}

// Levels >= 1 from the node entries of the leaf level (in node order): per level y-sort inside runs of `slice`, pack.
// parents / parents_b: ping-pong buffers (parents holds counts[1] entries on entry); yk .. yvb: sort scratch of at
// least counts[1] words each.
static int str_upper_levels(sb_context *ctx, sb_rtree *t, sb_entry *parents, sb_entry *parents_b, uint32_t *yk,
                            uint32_t *ykb, uint32_t *yv, uint32_t *yvb) {
	const uint32_t cap = t->cap;
	const unsigned block = ((cap + 31) / 32) * 32;
	int alt = 0;
	for (int level = 1; level < t->nlevels; level++) {
		const uint64_t c = t->counts[level];
		if (c == 1) {
			SB_CUDA(cudaMemcpyAsync(t->entries + t->offsets[level], parents, sizeof(sb_entry), cudaMemcpyDeviceToDevice,
			                        ctx->stream));
			break;
		}
		if (c <= STR_SMALL) {
			sb_prof_scope ps(ctx, "str_pack");
			k_str_small_level<<<1, 1024, 0, ctx->stream>>>(parents, (uint32_t)c, cap, t->slice,
			                                               t->entries + t->offsets[level], parents_b, 2ull);
			SB_LAUNCH_CHECK(ctx);
			sb_entry *tmp = parents;
			parents = parents_b;
			parents_b = tmp;
			continue;
		}
		{
			sb_prof_scope ps(ctx, "str_keys");
			k_str_ykeys_upper<<<sb_grid(c, 256, 16), 256, 0, ctx->stream>>>(parents, c, yk);
			SB_LAUNCH_CHECK(ctx);
		}
		SB_TRY(sbi_radix_sort_u32_segmented(ctx, yk, ykb, yv, yvb, c, t->slice, 0, 32, true, &alt));
		const uint32_t *p = alt ? yvb : yv;
		const uint64_t nn = (c + cap - 1) / cap;
		{
			sb_prof_scope ps(ctx, "str_pack");
			k_str_pack_upper<<<(unsigned)nn, block, 0, ctx->stream>>>(parents, p, c, cap, t->entries + t->offsets[level],
			                                                         parents_b, 2ull);
			SB_LAUNCH_CHECK(ctx);
		}
		sb_entry *tmp = parents;
		parents = parents_b;
		parents_b = tmp;
	}
	return SB_OK;
}

int sbi_rtree_build_str(sb_context *ctx, const float *box4, const int64_t *row_ids, uint64_t n, uint32_t cap,
                        sb_rtree **out, const double *px, const double *py) {
	SB_ARG(cap >= 4 && cap <= 255, "max_node_capacity must be in [4, 255]");
	SB_ARG(n < ((uint64_t)1 << 30), "n must be < 2^30");
	sb_rtree *t = new sb_rtree();
	t->device = ctx->device;
	t->cap = cap;
	*out = t;
	if (n == 0) {
		return SB_OK;
	}
	str_shape(t, n, cap);
	SB_CUDA(sb_pool_alloc(ctx, (void **)&t->entries, t->total * sizeof(sb_entry)));
	t->owner = ctx;
	const uint64_t nodes = (n + cap - 1) / cap;
	sb_entry *parents = nullptr, *parents_b = nullptr;
	SB_TRY(ctx->scratch_t(nodes, &parents));
	SB_TRY(ctx->scratch_t((nodes + cap - 1) / cap + 1, &parents_b));
	uint32_t *yk, *ykb, *yv, *yvb;
	const sb_boxsrc src = {reinterpret_cast<const float4 *>(box4), px, py};
	SB_TRY(str_leaf_level(ctx, src, row_ids, n, cap, t->slice, 0, -1, -1, t->entries + t->offsets[0], parents, &yk, &ykb,
	                      &yv, &yvb));
	return str_upper_levels(ctx, t, parents, parents_b, yk, ykb, yv, yvb);
}

// ---- multi-GPU build: the leaf level is built shard by shard, the upper levels from the gathered nodes ----
// shard = the rows of a contiguous range of the GLOBAL x order (ties by global input position) that starts at a multiple
// of `slice` positions, presented in ascending global input position; slice = the global slice size
// (sb_rtree_slice_for); node_base = global index of the shard's first node; x_key_before / x_key_after = x keys of the
// global order at the shard's first rank and at the first rank after it (-1 at the two ends of the layer).
int sbi_rtree_leaf_shard(sb_context *ctx, const float *box4, const int64_t *row_ids, uint64_t n, uint32_t cap,
                         uint64_t slice, uint64_t node_base, int64_t x_key_before, int64_t x_key_after, sb_entry *level0,
                         sb_entry *parents) {
	SB_ARG(cap >= 4 && cap <= 255, "max_node_capacity must be in [4, 255]");
	SB_ARG(n < ((uint64_t)1 << 30), "n must be < 2^30");
	SB_ARG(slice >= cap && slice % cap == 0, "slice must be a multiple of the node capacity");
	if (n == 0) {
		return SB_OK;
	}
	uint32_t *yk, *ykb, *yv, *yvb;
	SB_ARG(x_key_before <= 0xFFFFFFFFll && x_key_after <= 0xFFFFFFFFll, "x keys are 32-bit (-1 = none)");
	const sb_boxsrc src = {reinterpret_cast<const float4 *>(box4), nullptr, nullptr};
	return str_leaf_level(ctx, src, row_ids, n, cap, slice, node_base, x_key_before, x_key_after, level0, parents, &yk, &ykb,
	                      &yv, &yvb);
}

// the whole tree from its leaf entries (global order) and the leaf nodes' entries (node order)
int sbi_rtree_from_leaf_level(sb_context *ctx, const sb_entry *level0, uint64_t n, const sb_entry *nodes_in, uint32_t cap,
                              sb_rtree **out) {
	SB_ARG(cap >= 4 && cap <= 255, "max_node_capacity must be in [4, 255]");
	SB_ARG(n > 0 && n < ((uint64_t)1 << 30), "n must be in (0, 2^30)");
	sb_rtree *t = new sb_rtree();
	t->device = ctx->device;
	t->cap = cap;
	*out = t;
	str_shape(t, n, cap);
	SB_CUDA(sb_pool_alloc(ctx, (void **)&t->entries, t->total * sizeof(sb_entry)));
	t->owner = ctx;
	const uint64_t nodes = (n + cap - 1) / cap;
	SB_CUDA(cudaMemcpyAsync(t->entries, level0, n * sizeof(sb_entry), cudaMemcpyDeviceToDevice, ctx->stream));
	sb_entry *parents = nullptr, *parents_b = nullptr;
	uint32_t *yk = nullptr, *ykb = nullptr, *yv = nullptr, *yvb = nullptr;
	SB_TRY(ctx->scratch_t(nodes, &parents));
	SB_TRY(ctx->scratch_t((nodes + cap - 1) / cap + 1, &parents_b));
	SB_TRY(ctx->scratch_t(nodes, &yk));
	SB_TRY(ctx->scratch_t(nodes, &ykb));
	SB_TRY(ctx->scratch_t(nodes, &yv));
	SB_TRY(ctx->scratch_t(nodes, &yvb));
	SB_CUDA(cudaMemcpyAsync(parents, nodes_in, nodes * sizeof(sb_entry), cudaMemcpyDeviceToDevice, ctx->stream));
	return str_upper_levels(ctx, t, parents, parents_b, yk, ykb, yv, yvb);
}

// ---------------------------------------------------------------------------------------------------------------
// Batched index scan. One warp per query, depth-first with an explicit stack in shared memory, so the row ids of
// a query come out in exactly the reference's DFS order (rtree_scanner.hpp:38-83). Two launches: count, fill.
// ---------------------------------------------------------------------------------------------------------------
#define SCAN_WARPS 4
#define SCAN_STACK 1536 // >= (levels-1) * cap + 1 for cap <= 255 and up to 7 levels

struct scan_levels {
	const sb_entry *lvl[16];
	uint64_t cnt[16];
	int nlevels;
	uint32_t cap;
};

__device__ __forceinline__ void load_entry(const sb_entry *e, uint64_t &ptr, float4 &b) {
	const uint64_t *w = reinterpret_cast<const uint64_t *>(e);
	ptr = w[0];
	const uint64_t a = w[1], c = w[2];
	b = make_float4(__uint_as_float((uint32_t)a), __uint_as_float((uint32_t)(a >> 32)), __uint_as_float((uint32_t)c),
	                __uint_as_float((uint32_t)(c >> 32)));
}

__global__ void __launch_bounds__(SCAN_WARPS * 32)
    k_rtree_scan(scan_levels L, const float4 *__restrict__ queries, uint32_t nq, uint64_t *__restrict__ counts,
                 const uint64_t *__restrict__ offsets, uint32_t *__restrict__ out_query,
                 int64_t *__restrict__ out_row, uint64_t capacity) {
	__shared__ uint32_t s_stack[SCAN_WARPS][SCAN_STACK];
	if (offsets && offsets[nq] > capacity) {
		return; // fill pass: the results do not fit the caller's arrays (the host reports SB_ERR_CAPACITY from the same word)
	}
	const unsigned warp = threadIdx.x >> 5, lane = sb_lane();
	const unsigned lt = sb_lanemask_lt();
	uint32_t *stack = s_stack[warp];
	const uint32_t nwarps = gridDim.x * SCAN_WARPS;
	for (uint32_t q = blockIdx.x * SCAN_WARPS + warp; q < nq; q += nwarps) {
		const float4 qb = queries[q];
		uint64_t emitted = 0;
		const uint64_t base = offsets ? offsets[q] : 0;
		int sp = 0;
		{
			// RTreeIndex::InitializeScan root pre-check, rtree_index.cpp:110
			uint64_t rp;
			float4 rb;
			load_entry(L.lvl[L.nlevels - 1], rp, rb);
			if (sb_box_intersects(qb, rb)) {
				if (lane == 0) {
					stack[0] = ((uint32_t)(L.nlevels - 2) << 28) | 0u;
				}
				sp = 1;
			}
		}
		__syncwarp();
		while (sp > 0) {
			const uint32_t top = stack[sp - 1];
			sp--;
			__syncwarp();
			const int level = (int)(top >> 28);
			const uint64_t node = top & 0x0FFFFFFFu;
			const uint64_t beg = node * L.cap;
			const uint32_t cnt = (uint32_t)min((uint64_t)L.cap, L.cnt[level] - beg);
			if (level == 0) {
				for (uint32_t c0 = 0; c0 < cnt; c0 += 32) {
					const uint32_t i = c0 + lane;
					bool hit = false;
					uint64_t ptr = 0;
					if (i < cnt) {
						float4 b;
						load_entry(L.lvl[0] + beg + i, ptr, b);
						hit = sb_box_intersects(qb, b);
					}
					const unsigned m = __ballot_sync(0xFFFFFFFFu, hit);
					if (hit && out_row) {
						const uint64_t o = base + emitted + __popc(m & lt);
						out_row[o] = (int64_t)(ptr << 8) >> 8;
						out_query[o] = q;
					}
					emitted += __popc(m);
				}
			} else {
				// push hit children in reverse entry order so that the first hit is popped first (pre-order DFS)
				const uint32_t chunks = (cnt + 31) / 32;
				for (int c = (int)chunks - 1; c >= 0; c--) {
					const uint32_t i = (uint32_t)c * 32 + lane;
					bool hit = false;
					uint64_t ptr = 0;
					if (i < cnt) {
						float4 b;
						load_entry(L.lvl[level] + beg + i, ptr, b);
						hit = sb_box_intersects(qb, b);
					}
					const unsigned m = __ballot_sync(0xFFFFFFFFu, hit);
					const int nh = __popc(m);
					if (hit) {
						// lanes above me are pushed first (deeper in the stack)
						const int above = __popc(m & ~lt & ~(1u << lane));
						const int slot = sp + above;
						if (slot < SCAN_STACK) {
							stack[slot] = ((uint32_t)(level - 1) << 28) | (uint32_t)(ptr & 0x0FFFFFFFull);
						}
					}
					sp += nh;
				}
				__syncwarp();
			}
		}
		if (lane == 0 && counts) {
			counts[q] = emitted;
		}
	}
}

// exclusive scan of the per-query result counts, in place is not needed: off[q] = first result slot of query q,
// off[nq] = total. One CTA (nq is a batch of queries, not a column).
__global__ void __launch_bounds__(1024) k_scan_offsets(const uint64_t *__restrict__ cnt, uint32_t nq, uint64_t *__restrict__ off) {
	__shared__ unsigned long long s_w[33];
	const unsigned lane = sb_lane(), warp = threadIdx.x >> 5;
	unsigned long long carry = 0;
	for (uint32_t q0 = 0; q0 < nq; q0 += 1024) {
		const uint32_t q = q0 + threadIdx.x;
		const unsigned long long v = q < nq ? cnt[q] : 0ull;
		unsigned long long inc = v;
#pragma unroll
		for (int s = 1; s < 32; s <<= 1) {
			const unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
			if (lane >= (unsigned)s) {
				inc += o;
			}
		}
		if (lane == 31) {
			s_w[warp] = inc;
		}
		__syncthreads();
		if (warp == 0) {
			const unsigned long long w = s_w[lane];
			unsigned long long wi = w;
#pragma unroll
			for (int s = 1; s < 32; s <<= 1) {
				const unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, wi, s);
				if (lane >= (unsigned)s) {
					wi += o;
				}
			}
			s_w[lane] = wi - w;
			if (lane == 31) {
				s_w[32] = wi;
			}
		}
		__syncthreads();
		if (q < nq) {
			off[q] = carry + s_w[warp] + inc - v;
		}
		carry += s_w[32];
		__syncthreads();
	}
	if (threadIdx.x == 0) {
		off[nq] = carry;
	}
}

int sbi_rtree_scan(sb_context *ctx, const sb_rtree *tree, const float *qbox4, uint32_t nq, uint32_t *out_query,
                   int64_t *out_row, uint64_t capacity, uint64_t *count) {
	*count = 0;
	if (nq == 0 || tree->nlevels == 0) {
		return SB_OK;
	}
	SB_ARG((uint64_t)(tree->nlevels - 1) * tree->cap + 1 <= SCAN_STACK, "tree too deep for the scan stack");
	scan_levels L;
	L.nlevels = tree->nlevels;
	L.cap = tree->cap;
	for (int i = 0; i < tree->nlevels; i++) {
		L.lvl[i] = tree->entries + tree->offsets[i];
		L.cnt[i] = tree->counts[i];
	}
	uint64_t *d_counts = nullptr, *d_offsets = nullptr;
	SB_TRY(ctx->scratch_t(nq, &d_counts));
	SB_TRY(ctx->scratch_t((size_t)nq + 1, &d_offsets));
	const unsigned grid = sb_grid(nq, SCAN_WARPS, 8);
	const float4 *q = reinterpret_cast<const float4 *>(qbox4);
	// count, offsets and fill are enqueued back to back: the fill pass checks the capacity on the device, the host only
	// reads the total when everything is in flight (one synchronisation per call instead of two + two pageable copies)
	{
		sb_prof_scope ps(ctx, "rtree_scan");
		k_rtree_scan<<<grid, SCAN_WARPS * 32, 0, ctx->stream>>>(L, q, nq, d_counts, nullptr, nullptr, nullptr, 0);
		SB_LAUNCH_CHECK(ctx);
		k_scan_offsets<<<1, 1024, 0, ctx->stream>>>(d_counts, nq, d_offsets);
		SB_LAUNCH_CHECK(ctx);
		if (capacity) {
			k_rtree_scan<<<grid, SCAN_WARPS * 32, 0, ctx->stream>>>(L, q, nq, nullptr, d_offsets, out_query, out_row, capacity);
			SB_LAUNCH_CHECK(ctx);
		}
	}
	uint64_t total = 0;
	SB_TRY(sbi_read_words(ctx, d_offsets + nq, 1, &total));
	*count = total;
	if (total > capacity) {
		sb_set_error("rtree scan: %llu results, capacity %llu", (unsigned long long)total, (unsigned long long)capacity);
		return SB_ERR_CAPACITY;
	}
	return SB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Persistent-index export: the tree as the block images DuckDB's FixedSizeAllocator holds.
//
// The reference allocates one page per node while packing (tree.MakePage, rtree_index_create_physical.cpp:181-184):
// leaf pages from `leaf_allocator`, branch pages from `node_allocator` (rtree.cpp:28-35), in node order, level by
// level. A fresh allocator hands out segments sequentially (FixedSizeAllocator::New + FixedSizeBuffer::GetOffset,
// duckdb/src/execution/index/fixed_size_allocator.cpp:43-80, fixed_size_buffer.cpp:154-165): page k lives in buffer
// k / segments_per_buffer at offset k % segments_per_buffer, so every IndexPointer is a pure function of the node
// index and the images can be written by the device in one pass:
//   buffer image = [bitmask_count validity words: bit = 1 for a FREE segment][segments_per_buffer segments][slack]
//   segment      = RTreeNode {u32 count, 20 unused bytes} + max_node_capacity RTreeEntry (rtree_node.hpp:66-75, 183-189)
//   pointer      = metadata (node type) << 56 | offset << 32 | buffer id (index_pointer.hpp:19-24)
// One CTA per page; 8-byte words, consecutive threads write consecutive words.
// ---------------------------------------------------------------------------------------------------------------
int sbi_fsa_layout(uint64_t segment_size, uint64_t block_size, uint64_t *segments_per_buffer, uint64_t *bitmask_count) {
	SB_ARG(segment_size >= 8 && (segment_size & 7) == 0, "segment size must be a multiple of 8");
	SB_ARG((block_size & 7) == 0 && segment_size + 8 <= block_size, "block size: multiple of 8, >= segment + bitmask word");
	// one validity word per 64 segments; a word is added whenever the words so far cannot cover one more segment
	uint64_t used = 0, words = 0, segs = 0;
	while (used < block_size) {
		if (words == 0 || (words * 64) % segs == 0) {
			words++;
			used += 8;
		}
		if (used >= block_size) {
			break;
		}
		const uint64_t fit = (block_size - used) / segment_size;
		const uint64_t take = fit < 64 ? fit : 64;
		if (take == 0) {
			break;
		}
		segs += take;
		used += take * segment_size;
	}
	*segments_per_buffer = segs;
	*bitmask_count = words;
	return SB_OK;
}

struct page_levels {
	const sb_entry *ent[16]; // entries the pages of this node level are cut from
	uint64_t cnt[16];        // number of such entries
	uint64_t page_base[16];  // first page of the node level inside this allocator
	uint64_t child_base[16]; // first page of the children inside THEIR allocator
	int n;
	uint32_t cap;
	uint32_t rewrite; // 0: leaf pages (row ids stay as they are), 1: branch pages
	uint64_t pages, spb, bitmask_words, seg_words, block_words;
};

__device__ __forceinline__ uint64_t page_pointer(uint64_t type_and_child, uint64_t child_base, uint64_t spb) {
	const uint64_t c = (type_and_child & 0x00FFFFFFFFFFFFFFull) + child_base;
	return (type_and_child & 0xFF00000000000000ull) | ((c % spb) << 32) | (c / spb);
}

__global__ void __launch_bounds__(128) k_rtree_pages(const page_levels P, uint64_t *__restrict__ img) {
	const uint64_t page = blockIdx.x;
	int l = 0;
	while (l + 1 < P.n && page >= P.page_base[l + 1]) {
		l++;
	}
	const uint64_t node = page - P.page_base[l];
	const uint64_t first = node * P.cap;
	const uint64_t left = P.cnt[l] - first;
	const uint32_t count = left < P.cap ? (uint32_t)left : P.cap;
	const uint64_t *src = reinterpret_cast<const uint64_t *>(P.ent[l] + first);
	const uint64_t buf = page / P.spb, slot = page % P.spb;
	uint64_t *bufp = img + buf * P.block_words;
	uint64_t *dst = bufp + P.bitmask_words + slot * P.seg_words;
	for (uint32_t w = threadIdx.x; w < P.seg_words; w += blockDim.x) {
		uint64_t v = 0;
		if (w == 0) {
			v = count; // RTreeNode::count, then 20 unused bytes
		} else if (w >= 3 && w - 3 < 3 * count) {
			v = src[w - 3];
			if (P.rewrite && (w % 3) == 0) {
				v = page_pointer(v, P.child_base[l], P.spb);
			}
		}
		dst[w] = v;
	}
	if (slot == 0) {
		// validity words (1 = free) and the slack behind the last segment
		const uint64_t rest = P.pages - buf * P.spb;
		const uint64_t used = rest < P.spb ? rest : P.spb;
		for (uint32_t w = threadIdx.x; w < P.bitmask_words; w += blockDim.x) {
			uint64_t m = 0;
			for (uint32_t b = 0; b < 64; b++) {
				const uint64_t s = (uint64_t)w * 64 + b;
				m |= (uint64_t)(s >= used && s < P.spb) << b;
			}
			bufp[w] = m;
		}
		for (uint64_t w = P.bitmask_words + P.spb * P.seg_words + threadIdx.x; w < P.block_words; w += blockDim.x) {
			bufp[w] = 0;
		}
	}
}

int sbi_rtree_pages_info(const sb_rtree *tree, uint64_t block_size, sb_rtree_pages_info *info) {
	memset(info, 0, sizeof(*info));
	info->segment_size = 24 + 24 * (uint64_t)tree->cap; // RTreeConfig::GetNodeByteSize == GetLeafByteSize, rtree.hpp:21-26
	SB_TRY(sbi_fsa_layout(info->segment_size, block_size, &info->segments_per_buffer, &info->bitmask_count));
	info->bitmask_offset = info->bitmask_count * 8;
	if (tree->nlevels >= 2) {
		info->leaf_pages = tree->counts[1];
		for (int l = 2; l < tree->nlevels; l++) {
			info->branch_pages += tree->counts[l];
		}
	}
	const uint64_t spb = info->segments_per_buffer;
	info->leaf_buffers = (info->leaf_pages + spb - 1) / spb;
	info->branch_buffers = (info->branch_pages + spb - 1) / spb;
	return SB_OK;
}

// leaf_img / branch_img: device, leaf_buffers / branch_buffers blocks of block_size bytes. root_entry: host, 24 bytes.
int sbi_rtree_export_pages(sb_context *ctx, const sb_rtree *tree, uint64_t block_size, void *leaf_img,
                           void *branch_img, void *root_entry) {
	sb_rtree_pages_info info;
	SB_TRY(sbi_rtree_pages_info(tree, block_size, &info));
	uint64_t root[3] = {0, 0, 0};
	if (tree->nlevels >= 2) {
		SB_ARG(info.leaf_pages < (1ull << 31) && info.segments_per_buffer < (1ull << 24), "too many pages");
		page_levels P;
		memset(&P, 0, sizeof(P));
		P.cap = tree->cap;
		P.spb = info.segments_per_buffer;
		P.bitmask_words = info.bitmask_count;
		P.seg_words = info.segment_size / 8;
		P.block_words = block_size / 8;
		for (int pass = 0; pass < 2; pass++) {
			uint8_t *img = static_cast<uint8_t *>(pass == 0 ? leaf_img : branch_img);
			const uint64_t pages = pass == 0 ? info.leaf_pages : info.branch_pages;
			if (pages == 0) {
				continue;
			}
			SB_ARG(img != nullptr, "page image buffer");
			P.pages = pages;
			P.rewrite = (uint32_t)pass;
			if (pass == 0) {
				P.n = 1;
				P.ent[0] = tree->entries + tree->offsets[0];
				P.cnt[0] = tree->counts[0];
				P.page_base[0] = 0;
				P.child_base[0] = 0;
			} else {
				// node level L >= 1 is cut from the level-L entries; its children are the pages of node level L-1
				P.n = tree->nlevels - 2;
				uint64_t base = 0;
				for (int L = 1; L <= P.n; L++) {
					P.ent[L - 1] = tree->entries + tree->offsets[L];
					P.cnt[L - 1] = tree->counts[L];
					P.page_base[L - 1] = base;
					P.child_base[L - 1] = L == 1 ? 0 : P.page_base[L - 2];
					base += tree->counts[L + 1];
				}
			}
			// segments of the last buffer that stay free
			const uint64_t nbuf = (pages + P.spb - 1) / P.spb;
			const uint64_t used_last = pages - (nbuf - 1) * P.spb;
			SB_TRY(sbi_zero(ctx, img + (nbuf - 1) * block_size + info.bitmask_offset + used_last * info.segment_size,
			                (P.spb - used_last) * info.segment_size));
			sb_prof_scope ps(ctx, "rtree_pages");
			k_rtree_pages<<<(unsigned)pages, 128, 0, ctx->stream>>>(P, reinterpret_cast<uint64_t *>(img));
			SB_LAUNCH_CHECK(ctx);
		}
		// the root entry: pointer of the single top-level entry, rewritten like any other page pointer
		SB_TRY(sbi_read_words(ctx, tree->entries + tree->offsets[tree->nlevels - 1], 3, root));
		const uint64_t top_base = tree->nlevels == 2 ? 0 : info.branch_pages - 1; // the top node is the last branch page
		const uint64_t c = (root[0] & 0x00FFFFFFFFFFFFFFull) + top_base;
		root[0] = (root[0] & 0xFF00000000000000ull) | ((c % info.segments_per_buffer) << 32) | (c / info.segments_per_buffer);
	}
	memcpy(root_entry, root, 24);
	return SB_OK;
}

// ---- rtree_index_dump ------------------------------------------------------------------------------------------
// (level, bounds, row_id) of every entry below the root entry in the DFS order of the reference's pragma
// (rtree_index_pragmas.cpp:137-227 over RTreeScanner, rtree_scanner.hpp:38-83: the root node's entries are level 0, the
// root entry itself is not emitted, an entry is followed by its whole subtree). Node c of stored level l is the
// entries [c * cap, (c + 1) * cap) of that level; the entry of level l + 1 that owns it carries c in its pointer (upper
// levels are re-sorted by STR, so it is NOT entry c). Three data-parallel sweeps instead of a recursion:
//   sizes   bottom-up: size(l, i) = 1 + sum of the sizes of the entries of its child node
//   places  top-down : the entries of a node follow their parent entry, each after its elder siblings' subtrees
//   write   every entry to its place
struct dump_levels {
	const sb_entry *lvl[16];
	uint64_t count[16];
	uint64_t off[16]; // entry offset of the level inside the size / place arrays
	int top;          // stored level of the root entry (not emitted)
	uint32_t cap;
};

__global__ void __launch_bounds__(256) k_dump_sizes(dump_levels L, int l, uint64_t *__restrict__ size) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < L.count[l]; i += stride) {
		uint64_t sz = 1;
		if (l > 0) {
			const uint64_t c = L.lvl[l][i].ptr & 0x00FFFFFFFFFFFFFFull;
			const uint64_t b = c * L.cap, e = min(b + L.cap, L.count[l - 1]);
			for (uint64_t k = b; k < e; k++) {
				sz += size[L.off[l - 1] + k];
			}
		}
		size[L.off[l] + i] = sz;
	}
}

// parents = entries of level l + 1 (the single root entry when l + 1 == top, whose children start at place 0)
__global__ void __launch_bounds__(256) k_dump_places(dump_levels L, int l, const uint64_t *__restrict__ size,
                                                     uint64_t *__restrict__ place) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < L.count[l + 1]; j += stride) {
		uint64_t at = l + 1 == L.top ? 0 : place[L.off[l + 1] + j] + 1;
		const uint64_t c = L.lvl[l + 1][j].ptr & 0x00FFFFFFFFFFFFFFull;
		const uint64_t b = c * L.cap, e = min(b + L.cap, L.count[l]);
		for (uint64_t k = b; k < e; k++) {
			place[L.off[l] + k] = at;
			at += size[L.off[l] + k];
		}
	}
}

__global__ void __launch_bounds__(256) k_dump_write(dump_levels L, int l, const uint64_t *__restrict__ place,
                                                    int32_t *__restrict__ out_level, float4 *__restrict__ out_box,
                                                    int64_t *__restrict__ out_row) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < L.count[l]; i += stride) {
		const uint64_t pos = place[L.off[l] + i];
		uint64_t ptr;
		float4 b;
		load_entry(L.lvl[l] + i, ptr, b);
		out_level[pos] = L.top - 1 - l;
		out_box[pos] = b;
		out_row[pos] = l == 0 ? (int64_t)(ptr & 0x00FFFFFFFFFFFFFFull) : -1; // RTreePointer::GetRowId of a ROW_ID pointer
	}
}

int sbi_rtree_dump(sb_context *ctx, const sb_rtree *tree, int32_t *out_level, float *out_box4, int64_t *out_row,
                   uint64_t capacity, uint64_t *count) {
	dump_levels L;
	memset(&L, 0, sizeof(L));
	L.top = tree->nlevels - 1;
	L.cap = tree->cap;
	uint64_t total = 0;
	for (int l = 0; l < tree->nlevels; l++) {
		L.lvl[l] = tree->entries + tree->offsets[l];
		L.count[l] = tree->counts[l];
		L.off[l] = tree->offsets[l];
		if (l < L.top) {
			total += tree->counts[l];
		}
	}
	*count = total;
	if (total > capacity) {
		sb_set_error("rtree dump: %llu rows, capacity %llu", (unsigned long long)total, (unsigned long long)capacity);
		return SB_ERR_CAPACITY;
	}
	if (total == 0) {
		return SB_OK;
	}
	uint64_t *d_size = nullptr, *d_place = nullptr;
	SB_TRY(ctx->scratch_t(tree->total, &d_size));
	SB_TRY(ctx->scratch_t(tree->total, &d_place));
	for (int l = 0; l < L.top; l++) {
		k_dump_sizes<<<sb_grid(L.count[l], 256, 16), 256, 0, ctx->stream>>>(L, l, d_size);
		SB_LAUNCH_CHECK(ctx);
	}
	for (int l = L.top - 1; l >= 0; l--) {
		k_dump_places<<<sb_grid(L.count[l + 1], 256, 16), 256, 0, ctx->stream>>>(L, l, d_size, d_place);
		SB_LAUNCH_CHECK(ctx);
	}
	for (int l = 0; l < L.top; l++) {
		k_dump_write<<<sb_grid(L.count[l], 256, 16), 256, 0, ctx->stream>>>(L, l, d_place, out_level,
		                                                                      reinterpret_cast<float4 *>(out_box4), out_row);
		SB_LAUNCH_CHECK(ctx);
	}
	return SB_OK;
}

// ---- CREATE INDEX straight from the GEOMETRY column --------------------------------------------------------------
// rows without bounds (empty geometries, malformed blobs) never reach the index: the reference plans a filter on
// the key's cached bbox below the sort (rtree_index_create_logical.cpp:42-72). Selection = tile counts + scan + gather,
// order preserving.
#define SEL_TILE 2048
__global__ void __launch_bounds__(256) k_sel_count(const uint8_t *__restrict__ valid, uint64_t n, uint32_t *__restrict__ tile_cnt) {
	__shared__ uint32_t s_cnt;
	for (uint64_t tile = blockIdx.x; tile * SEL_TILE < n; tile += gridDim.x) {
		if (threadIdx.x == 0) {
			s_cnt = 0;
		}
		__syncthreads();
		uint32_t c = 0;
		for (uint32_t k = threadIdx.x; k < SEL_TILE; k += 256) {
			const uint64_t i = tile * SEL_TILE + k;
			c += (i < n && valid[i]) ? 1u : 0u;
		}
		c = __reduce_add_sync(0xFFFFFFFFu, c);
		if (sb_lane() == 0 && c) {
			atomicAdd(&s_cnt, c);
		}
		__syncthreads();
		if (threadIdx.x == 0) {
			tile_cnt[tile] = s_cnt;
		}
		__syncthreads();
	}
}
// single CTA: exclusive scan of the tile counts (in place), total to *total
__global__ void __launch_bounds__(1024) k_sel_scan(uint32_t *__restrict__ tile_cnt, uint64_t tiles, unsigned long long *__restrict__ total) {
	__shared__ unsigned long long s_carry;
	__shared__ uint32_t s_w[32];
	if (threadIdx.x == 0) {
		s_carry = 0;
	}
	__syncthreads();
	for (uint64_t base = 0; base < tiles; base += 1024) {
		const uint64_t i = base + threadIdx.x;
		const uint32_t v = i < tiles ? tile_cnt[i] : 0u;
		uint32_t inc = v;
#pragma unroll
		for (int sft = 1; sft < 32; sft <<= 1) {
			const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, sft);
			if (sb_lane() >= (unsigned)sft) {
				inc += o;
			}
		}
		if (sb_lane() == 31) {
			s_w[threadIdx.x >> 5] = inc;
		}
		__syncthreads();
		uint32_t wb = 0;
		for (unsigned w = 0; w < (threadIdx.x >> 5); w++) {
			wb += s_w[w];
		}
		const unsigned long long carry = s_carry;
		if (i < tiles) {
			tile_cnt[i] = (uint32_t)(carry + wb + inc - v);
		}
		__syncthreads();
		if (threadIdx.x == 1023) {
			s_carry = carry + wb + inc;
		}
		__syncthreads();
	}
	if (threadIdx.x == 0) {
		*total = s_carry;
	}
}
__global__ void __launch_bounds__(256) k_sel_gather(const uint8_t *__restrict__ valid, uint64_t n, const uint32_t *__restrict__ tile_off,
                                                    const float4 *__restrict__ box, const int64_t *__restrict__ rows,
                                                    float4 *__restrict__ obox, int64_t *__restrict__ orows) {
	__shared__ uint32_t s_w[8];
	const unsigned lane = sb_lane(), warp = threadIdx.x >> 5;
	for (uint64_t tile = blockIdx.x; tile * SEL_TILE < n; tile += gridDim.x) {
		uint32_t run = tile_off[tile];
		for (uint32_t k0 = 0; k0 < SEL_TILE; k0 += 256) {
			const uint64_t i = tile * SEL_TILE + k0 + threadIdx.x;
			const bool v = i < n && valid[i];
			const uint32_t m = __ballot_sync(0xFFFFFFFFu, v);
			if (lane == 0) {
				s_w[warp] = __popc(m);
			}
			__syncthreads();
			uint32_t wb = 0, tot = 0;
			for (unsigned w = 0; w < 8; w++) {
				wb += w < warp ? s_w[w] : 0u;
				tot += s_w[w];
			}
			if (v) {
				const uint32_t o = run + wb + __popc(m & sb_lanemask_lt());
				obox[o] = box[i];
				orows[o] = rows ? rows[i] : (int64_t)i;
			}
			run += tot;
			__syncthreads();
		}
	}
}

int sbi_select_valid(sb_context *ctx, const uint8_t *valid, uint64_t n, const float *box4, const int64_t *rows,
                     float *obox4, int64_t *orows, uint64_t *count) {
	*count = 0;
	if (n == 0) {
		return SB_OK;
	}
	const uint64_t tiles = (n + SEL_TILE - 1) / SEL_TILE;
	uint32_t *d_tiles = nullptr;
	unsigned long long *d_total = nullptr;
	SB_TRY(ctx->scratch_t(tiles, &d_tiles));
	SB_TRY(ctx->scratch_t(1, &d_total));
	k_sel_count<<<sb_grid(tiles, 1, 16), 256, 0, ctx->stream>>>(valid, n, d_tiles);
	SB_LAUNCH_CHECK(ctx);
	k_sel_scan<<<1, 1024, 0, ctx->stream>>>(d_tiles, tiles, d_total);
	SB_LAUNCH_CHECK(ctx);
	k_sel_gather<<<sb_grid(tiles, 1, 16), 256, 0, ctx->stream>>>(valid, n, d_tiles, reinterpret_cast<const float4 *>(box4), rows,
	                                                            reinterpret_cast<float4 *>(obox4), orows);
	SB_LAUNCH_CHECK(ctx);
	return sbi_read_u64(ctx, reinterpret_cast<const uint64_t *>(d_total), count);
}
