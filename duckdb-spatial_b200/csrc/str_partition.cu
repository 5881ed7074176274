// str_partition.cu -- K4b: the ORDER BY of CREATE INDEX ... USING RTREE without a sort.
//
// BuildRTreeBottomUp (rtree_index_create_physical.cpp:149-236) consumes the x-sorted layer in runs of `slice` entries
// and re-sorts every run by y: of the x order it only uses WHICH slice an entry falls in. So the x sort (4 radix passes)
// is replaced by
//   1. the exact x keys at ranks slice, 2*slice, ... ("cut keys"): a radix SELECT of all ranks at once -- three
//      histogram levels (11 + 9 + 12 key bits), each level only counting keys that fall in a bucket holding a rank.
//      Float keys share their top bits, so level 2 still sees nearly every key: it is counted in shared memory (one
//      global atomic per key made it 138 us for 10 M keys); level 3 only sees the few keys next to a rank;
//   2. a stable PARTITION of the entries (in input order) into bins: bin(key) = #{cut < key} + #{cut <= key}, i.e. the
//      open intervals between consecutive cut keys (even bins) and the runs of keys EQUAL to a cut key (odd bins). Bins
//      are laid out in key order and keep the input order inside, so the positions [j*slice, (j+1)*slice) of the result
//      hold exactly the entries of slice j of the (x key, input position) order -- equal-key runs that straddle a slice
//      boundary are cut by input position, as a stable sort would cut them.
// The partition moves the whole 32-byte record (box, row id, y key), so the one random permute the leaf level needs
// happens here as writes in runs per bin, and everything after it (y sort inside slices, pack) is local to a slice.
// Order inside a slice after the y sort: (y key, bin, input position).
#include "internal.cuh"

#define SEL_L1 2048u // key bits 31..21
#define SEL_L2 512u  // key bits 20..12
#define SEL_L3 4096u // key bits 11..0
#define SEL_SH2 12
#define SEL_SMEM_C1 96u // level 2 is counted in shared memory (192 KB) when at most this many level-1 buckets hold a rank

// exclusive block scan of one value per thread; *total = sum over the block. s_w: 33 words.
template <int THREADS>
__device__ __forceinline__ uint32_t sp_block_scan(uint32_t v, uint32_t *s_w, uint32_t *total) {
	const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	uint32_t inc = v;
#pragma unroll
	for (int s = 1; s < 32; s <<= 1) {
		const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
		if (lane >= (unsigned)s) {
			inc += o;
		}
	}
	if (lane == 31) {
		s_w[warp] = inc;
	}
	__syncthreads();
	if (warp == 0) {
		const uint32_t w = lane < THREADS / 32 ? s_w[lane] : 0u;
		uint32_t wi = w;
#pragma unroll
		for (int s = 1; s < 32; s <<= 1) {
			const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, wi, s);
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
	const uint32_t r = s_w[warp] + inc - v;
	*total = s_w[32];
	__syncthreads();
	return r;
}

// first index in [0, cnt] whose prefix exceeds r, minus one: the bucket that holds rank r (pre[cnt] > r is guaranteed)
__device__ __forceinline__ uint32_t sp_bucket_of(const uint32_t *pre, uint32_t cnt, uint32_t r) {
	uint32_t lo = 0, hi = cnt;
	while (lo < hi) {
		const uint32_t mid = (lo + hi) >> 1;
		if (pre[mid] > r) {
			hi = mid;
		} else {
			lo = mid + 1;
		}
	}
	return lo - 1;
}

// ---- level 1: the histogram of the top 11 bits is counted by the kernel that makes the keys (rtree_str.cu k_str_xkeys)

// per splitter k (rank (k + 1) * slice): sp[0] = level-1 bucket, sp[1] = rank inside it, sp[2] = index of the bucket among
// the distinct buckets that hold a rank, sp[3..5] = the same for level 2. map1[bucket] = that index + 1 (0 = no rank).
__global__ void __launch_bounds__(1024) k_sel_cut1(const uint32_t *__restrict__ hist1, uint32_t nsp, uint64_t slice,
                                                   uint32_t *__restrict__ sp, uint32_t *__restrict__ map1,
                                                   uint32_t *__restrict__ ncut1) {
	__shared__ uint32_t s_pre[SEL_L1 + 1];
	__shared__ uint32_t s_w[33];
	const uint32_t t = threadIdx.x;
	const uint32_t a = hist1[2 * t], b = hist1[2 * t + 1];
	uint32_t total;
	const uint32_t ex = sp_block_scan<1024>(a + b, s_w, &total);
	s_pre[2 * t] = ex;
	s_pre[2 * t + 1] = ex + a;
	if (t == 0) {
		s_pre[SEL_L1] = total;
	}
	__syncthreads();
	uint32_t *b1 = sp, *rem1 = sp + nsp, *c1 = sp + 2 * (size_t)nsp;
	for (uint32_t k = t; k < nsp; k += 1024) {
		const uint32_t r = (uint32_t)((uint64_t)(k + 1) * slice);
		const uint32_t bk = sp_bucket_of(s_pre, SEL_L1, r);
		b1[k] = bk;
		rem1[k] = r - s_pre[bk];
	}
	__syncthreads();
	uint32_t base = 0;
	for (uint32_t k0 = 0; k0 < nsp; k0 += 1024) {
		const uint32_t k = k0 + t;
		const uint32_t flag = (k < nsp && (k == 0 || b1[k] != b1[k - 1])) ? 1u : 0u;
		uint32_t tot;
		const uint32_t e = sp_block_scan<1024>(flag, s_w, &tot);
		if (k < nsp) {
			const uint32_t c = base + e + flag - 1;
			c1[k] = c;
			if (flag) {
				map1[b1[k]] = c + 1;
			}
		}
		base += tot;
	}
	if (t == 0) {
		*ncut1 = base;
	}
}

// ---- levels 2 and 3: histograms of the next bits of the keys whose prefix holds a rank ----------------------------
// count one key into hist[a]; a warp whose keys all count into one word (constant columns, grids) issues one atomic
__device__ __forceinline__ void sel_count(uint32_t *hist, uint32_t a, unsigned lane) {
	const unsigned act = __ballot_sync(0xFFFFFFFFu, a != 0xFFFFFFFFu);
	if (!act) {
		return;
	}
	const int leader = __ffs(act) - 1;
	const uint32_t a0 = __shfl_sync(0xFFFFFFFFu, a, leader);
	const unsigned same = __ballot_sync(0xFFFFFFFFu, a == a0);
	if (same == act) {
		if ((int)lane == leader) {
			atomicAdd(&hist[a0], (uint32_t)__popc(act));
		}
	} else if (a != 0xFFFFFFFFu) {
		atomicAdd(&hist[a], 1u);
	}
}

// level 2. Dynamic shared memory: SEL_SMEM_C1 * SEL_L2 counters, used when *ncut1 <= SEL_SMEM_C1 (then flushed).
__global__ void __launch_bounds__(1024) k_sel_hist2(const uint32_t *__restrict__ keys, uint64_t n,
                                                    const uint32_t *__restrict__ map1, const uint32_t *__restrict__ ncut1,
                                                    uint32_t *__restrict__ hist2) {
	extern __shared__ uint32_t s_tab[];
	__shared__ uint16_t s_map1[SEL_L1];
	const uint32_t nc = *ncut1;
	const bool in_smem = nc <= SEL_SMEM_C1;
	for (uint32_t i = threadIdx.x; i < SEL_L1; i += 1024) {
		s_map1[i] = (uint16_t)map1[i];
	}
	if (in_smem) {
		for (uint32_t i = threadIdx.x; i < nc * SEL_L2; i += 1024) {
			s_tab[i] = 0;
		}
	}
	__syncthreads();
	uint32_t *tab = in_smem ? s_tab : hist2;
	const unsigned lane = threadIdx.x & 31u;
	const uint64_t stride = (uint64_t)gridDim.x * 1024;
	for (uint64_t wbase = (uint64_t)blockIdx.x * 1024 + (threadIdx.x & ~31u); wbase < n; wbase += stride) {
		const uint64_t i = wbase + lane;
		uint32_t a = 0xFFFFFFFFu;
		if (i < n) {
			const uint32_t key = keys[i];
			const uint32_t m1 = s_map1[key >> 21];
			if (m1) {
				a = (m1 - 1) * SEL_L2 + ((key >> SEL_SH2) & (SEL_L2 - 1));
			}
		}
		sel_count(tab, a, lane);
	}
	if (in_smem) {
		__syncthreads();
		for (uint32_t i = threadIdx.x; i < nc * SEL_L2; i += 1024) {
			const uint32_t c = s_tab[i];
			if (c) {
				atomicAdd(&hist2[i], c);
			}
		}
	}
}

__global__ void __launch_bounds__(256) k_sel_hist3(const uint32_t *__restrict__ keys, uint64_t n,
                                                   const uint32_t *__restrict__ map1, const uint32_t *__restrict__ map2,
                                                   uint32_t *__restrict__ hist3) {
	__shared__ uint16_t s_map1[SEL_L1];
	for (uint32_t i = threadIdx.x; i < SEL_L1; i += 256) {
		s_map1[i] = (uint16_t)map1[i];
	}
	__syncthreads();
	const unsigned lane = threadIdx.x & 31u;
	const uint64_t stride = (uint64_t)gridDim.x * 256;
	for (uint64_t wbase = (uint64_t)blockIdx.x * 256 + (threadIdx.x & ~31u); wbase < n; wbase += stride) {
		const uint64_t i = wbase + lane;
		uint32_t a = 0xFFFFFFFFu;
		if (i < n) {
			const uint32_t key = keys[i];
			const uint32_t m1 = s_map1[key >> 21];
			if (m1) {
				const uint32_t m2 = map2[(size_t)(m1 - 1) * SEL_L2 + ((key >> SEL_SH2) & (SEL_L2 - 1))];
				if (m2) {
					a = (m2 - 1) * SEL_L3 + (key & (SEL_L3 - 1));
				}
			}
		}
		sel_count(hist3, a, lane);
	}
}

// one CTA per level-1 bucket that holds a rank: the level-2 bucket of each of its ranks
__global__ void __launch_bounds__(SEL_L2 / 2) k_sel_cut2(const uint32_t *__restrict__ hist2, uint32_t nsp,
                                                          uint32_t *__restrict__ sp, const uint32_t *__restrict__ ncut1) {
	__shared__ uint32_t s_pre[SEL_L2 + 1];
	__shared__ uint32_t s_w[33];
	const uint32_t c = blockIdx.x;
	if (c >= *ncut1) {
		return;
	}
	const uint32_t t = threadIdx.x;
	const uint32_t *h = hist2 + (size_t)c * SEL_L2;
	const uint32_t a = h[2 * t], b = h[2 * t + 1];
	uint32_t total;
	const uint32_t ex = sp_block_scan<SEL_L2 / 2>(a + b, s_w, &total);
	s_pre[2 * t] = ex;
	s_pre[2 * t + 1] = ex + a;
	if (t == 0) {
		s_pre[SEL_L2] = total;
	}
	__syncthreads();
	const uint32_t *rem1 = sp + nsp, *c1 = sp + 2 * (size_t)nsp;
	uint32_t *b2 = sp + 3 * (size_t)nsp, *rem2 = sp + 4 * (size_t)nsp;
	for (uint32_t k = t; k < nsp; k += SEL_L2 / 2) {
		if (c1[k] == c) {
			const uint32_t r = rem1[k];
			const uint32_t bk = sp_bucket_of(s_pre, SEL_L2, r);
			b2[k] = bk;
			rem2[k] = r - s_pre[bk];
		}
	}
}

// distinct (level-1, level-2) prefixes that hold a rank: sp[5] and map2[c1 * SEL_L2 + b2] = index + 1
__global__ void __launch_bounds__(1024) k_sel_cut2b(uint32_t nsp, uint32_t *__restrict__ sp, uint32_t *__restrict__ map2) {
	__shared__ uint32_t s_w[33];
	const uint32_t t = threadIdx.x;
	const uint32_t *c1 = sp + 2 * (size_t)nsp, *b2 = sp + 3 * (size_t)nsp;
	uint32_t *c2 = sp + 5 * (size_t)nsp;
	uint32_t base = 0;
	for (uint32_t k0 = 0; k0 < nsp; k0 += 1024) {
		const uint32_t k = k0 + t;
		const uint32_t flag = (k < nsp && (k == 0 || c1[k] != c1[k - 1] || b2[k] != b2[k - 1])) ? 1u : 0u;
		uint32_t tot;
		const uint32_t e = sp_block_scan<1024>(flag, s_w, &tot);
		if (k < nsp) {
			const uint32_t c = base + e + flag - 1;
			c2[k] = c;
			if (flag) {
				map2[(size_t)c1[k] * SEL_L2 + b2[k]] = c + 1;
			}
		}
		base += tot;
	}
}

// one CTA per splitter: the low bits from its level-3 histogram -> the cut key. Block 0 also stores the boundary keys
// a shard of a larger layer brings along (the cut keys of the global order just outside the shard).
__global__ void __launch_bounds__(256) k_sel_cut3(const uint32_t *__restrict__ hist3, uint32_t nsp, const uint32_t *__restrict__ sp,
                                                  uint32_t *__restrict__ cuts, uint32_t ncut, int64_t key_lo, int64_t key_hi) {
	__shared__ uint32_t s_pre[SEL_L3 + 1];
	__shared__ uint32_t s_w[33];
	const uint32_t k = blockIdx.x, t = threadIdx.x;
	if (k == 0 && t == 0) {
		if (key_lo >= 0) {
			cuts[0] = (uint32_t)key_lo;
		}
		if (key_hi >= 0) {
			cuts[ncut - 1] = (uint32_t)key_hi;
		}
	}
	if (k >= nsp) {
		return;
	}
	const uint32_t *h = hist3 + (size_t)sp[5 * (size_t)nsp + k] * SEL_L3;
	constexpr int PER = SEL_L3 / 256;
	uint32_t v[PER], sum = 0;
#pragma unroll
	for (int j = 0; j < PER; j++) {
		v[j] = h[PER * t + j];
		sum += v[j];
	}
	uint32_t total;
	uint32_t run = sp_block_scan<256>(sum, s_w, &total);
#pragma unroll
	for (int j = 0; j < PER; j++) {
		s_pre[PER * t + j] = run;
		run += v[j];
	}
	if (t == 0) {
		s_pre[SEL_L3] = total;
	}
	__syncthreads();
	if (t == 0) {
		const uint32_t low = sp_bucket_of(s_pre, SEL_L3, sp[4 * (size_t)nsp + k]);
		cuts[(key_lo >= 0 ? 1 : 0) + k] = (sp[k] << 21) | (sp[3 * (size_t)nsp + k] << SEL_SH2) | low;
	}
}

// begin: scratch + zeroing (sel->hist1 is then filled by the key kernel); finish: levels 2, 3 and the cut keys
int sbi_str_select_begin(sb_context *ctx, uint64_t n, uint64_t slice, int64_t key_lo, int64_t key_hi, sb_str_select *sel) {
	const uint32_t nsp = (uint32_t)((n + slice - 1) / slice - 1);
	sel->nsp = nsp;
	sel->slice = slice;
	sel->key_lo = key_lo;
	sel->key_hi = key_hi;
	sel->ncut = nsp + (key_lo >= 0 ? 1u : 0u) + (key_hi >= 0 ? 1u : 0u);
	sel->cuts = nullptr;
	sel->hist1 = nullptr;
	sel->temp = nullptr;
	if (sel->ncut == 0) {
		return SB_OK;
	}
	SB_TRY(ctx->scratch_t(sel->ncut, &sel->cuts));
	if (nsp == 0) {
		return SB_OK;
	}
	const uint32_t c1max = nsp < SEL_L1 ? nsp : SEL_L1;
	// [hist1 2048][map1 2048][ncut1 (padded to 64)][sp 6 * nsp][hist2 c1max * 512][map2 c1max * 512][hist3 nsp * 4096]
	const size_t words = 2 * (size_t)SEL_L1 + 64 + 6 * (size_t)nsp + 2 * (size_t)c1max * SEL_L2 + (size_t)nsp * SEL_L3;
	SB_TRY(ctx->scratch_t(words, &sel->temp));
	SB_CUDA(cudaMemsetAsync(sel->temp, 0, words * sizeof(uint32_t), ctx->stream));
	sel->hist1 = sel->temp;
	return SB_OK;
}

int sbi_str_select_finish(sb_context *ctx, const uint32_t *xkeys, uint64_t n, sb_str_select *sel) {
	if (sel->ncut == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "str_select");
	const uint32_t nsp = sel->nsp;
	const uint32_t c1max = nsp < SEL_L1 ? nsp : SEL_L1;
	uint32_t *hist1 = sel->temp, *map1 = hist1 + SEL_L1, *ncut1 = map1 + SEL_L1, *sp = ncut1 + 64;
	uint32_t *hist2 = sp + 6 * (size_t)nsp, *map2 = hist2 + (size_t)c1max * SEL_L2, *hist3 = map2 + (size_t)c1max * SEL_L2;
	if (nsp) {
		const unsigned grid = sb_grid(n, 256 * 8, 8);
		k_sel_cut1<<<1, 1024, 0, ctx->stream>>>(hist1, nsp, sel->slice, sp, map1, ncut1);
		SB_LAUNCH_CHECK(ctx);
		SB_SMEM_OPTIN(ctx, k_sel_hist2, SEL_SMEM_C1 * SEL_L2 * sizeof(uint32_t));
		k_sel_hist2<<<sb_grid(n, 1024 * 16, 1), 1024, SEL_SMEM_C1 * SEL_L2 * sizeof(uint32_t), ctx->stream>>>(xkeys, n, map1, ncut1,
		                                                                                                    hist2);
		SB_LAUNCH_CHECK(ctx);
		k_sel_cut2<<<c1max, SEL_L2 / 2, 0, ctx->stream>>>(hist2, nsp, sp, ncut1);
		SB_LAUNCH_CHECK(ctx);
		k_sel_cut2b<<<1, 1024, 0, ctx->stream>>>(nsp, sp, map2);
		SB_LAUNCH_CHECK(ctx);
		k_sel_hist3<<<grid, 256, 0, ctx->stream>>>(xkeys, n, map1, map2, hist3);
		SB_LAUNCH_CHECK(ctx);
	}
	k_sel_cut3<<<nsp ? nsp : 1, 256, 0, ctx->stream>>>(hist3, nsp, sp, sel->cuts, sel->ncut, sel->key_lo, sel->key_hi);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// ---- the partition ----------------------------------------------------------------------------------------------
// bin of a key among the sorted cut keys (duplicates allowed): #{cut < key} + #{cut <= key}
__device__ __forceinline__ uint32_t sp_bin(uint32_t key, const uint32_t *s_cuts, uint32_t ncut) {
	uint32_t lo = 0, hi = ncut;
	while (lo < hi) {
		const uint32_t mid = (lo + hi) >> 1;
		if (s_cuts[mid] < key) {
			lo = mid + 1;
		} else {
			hi = mid;
		}
	}
	uint32_t ub = lo;
	if (lo < ncut && s_cuts[lo] == key) {
		ub = lo + 1;
		hi = ncut;
		while (ub < hi) {
			const uint32_t mid = (ub + hi) >> 1;
			if (s_cuts[mid] <= key) {
				ub = mid + 1;
			} else {
				hi = mid;
			}
		}
	}
	return lo + ub;
}

// A 10-step binary search per key made k_part_count instruction bound (58 us at C1). The search is narrowed first by a
// table over 1024 equal key intervals of [first cut, last cut]: lut[c] = #{cut < start of interval c}, so that the
// lower bound of a key of interval c lies in [lut[c], lut[c + 1]] -- usually zero or one step.
// lut buffer: [0] = first cut, [1] = shift, [2 .. 2 + 1024] = the table (lut[1024] = ncut).
#define SP_LUT 1024u
__global__ void __launch_bounds__(1024) k_part_lut(const uint32_t *__restrict__ cuts, uint32_t ncut, uint32_t *__restrict__ lut) {
	const uint32_t lo = cuts[0], hi = cuts[ncut - 1];
	const uint32_t range = hi - lo;
	const int bits = 32 - __clz(range | 1u);
	const int shift = bits > 10 ? bits - 10 : 0;
	const uint32_t c = threadIdx.x;
	const unsigned long long start = (unsigned long long)lo + ((unsigned long long)c << shift);
	uint32_t a = 0, b = ncut;
	while (a < b) {
		const uint32_t mid = (a + b) >> 1;
		if ((unsigned long long)cuts[mid] < start) {
			a = mid + 1;
		} else {
			b = mid;
		}
	}
	lut[2 + c] = a;
	if (c == 0) {
		lut[0] = lo;
		lut[1] = (uint32_t)shift;
		lut[2 + SP_LUT] = ncut;
	}
}

// is_run (optional): the key EQUALS a cut key (its bin is a run of equal keys; with d equal cut keys the bin id is
// 2 * lb + d, so the parity of the id does not tell)
__device__ __forceinline__ uint32_t sp_bin_lut(uint32_t key, const uint32_t *s_cuts, uint32_t ncut, const uint16_t *s_lut,
                                               uint32_t lo, uint32_t shift, bool *is_run = nullptr) {
	if (is_run) {
		*is_run = false;
	}
	if (key < lo) {
		return 0;
	}
	const uint32_t c = min(SP_LUT - 1, (key - lo) >> shift);
	uint32_t a = s_lut[c], b = s_lut[c + 1];
	if (b - a <= 1) { // the usual case: at most one cut key starts in the interval
		a += (a < b && s_cuts[a] < key) ? 1u : 0u;
	} else {
		while (a < b) {
			const uint32_t mid = (a + b) >> 1;
			if (s_cuts[mid] < key) {
				a = mid + 1;
			} else {
				b = mid;
			}
		}
	}
	uint32_t ub = a;
	if (a < ncut && s_cuts[a] == key) {
		if (is_run) {
			*is_run = true;
		}
		ub = a + 1;
		uint32_t hi = ncut;
		while (ub < hi) {
			const uint32_t mid = (ub + hi) >> 1;
			if (s_cuts[mid] <= key) {
				ub = mid + 1;
			} else {
				hi = mid;
			}
		}
	}
	return a + ub;
}

// generic form: bin ids as 32-bit sort keys (any number of cut keys that fits shared memory)
__global__ void __launch_bounds__(256) k_part_bins(const uint32_t *__restrict__ xkeys, uint64_t n,
                                                   const uint32_t *__restrict__ cuts, uint32_t ncut,
                                                   uint32_t *__restrict__ bins) {
	extern __shared__ uint32_t s_cuts_dyn[];
	for (uint32_t i = threadIdx.x; i < ncut; i += 256) {
		s_cuts_dyn[i] = cuts[i];
	}
	__syncthreads();
	const uint64_t stride = (uint64_t)gridDim.x * 256;
	for (uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
		bins[i] = sp_bin(xkeys[i], s_cuts_dyn, ncut);
	}
}

int sbi_str_bin_ids(sb_context *ctx, const uint32_t *xkeys, uint64_t n, const uint32_t *cuts, uint32_t ncut,
                    uint32_t *bins) {
	SB_ARG(ncut <= 12000, "STR partition: too many slices for one device build");
	sb_prof_scope ps(ctx, "str_partition");
	k_part_bins<<<sb_grid(n, 256 * 8, 8), 256, ncut * sizeof(uint32_t), ctx->stream>>>(xkeys, n, cuts, ncut, bins);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// fast form (<= SP_MAXBINS bins): count per (tile, bin) -> exclusive positions -> stable scatter of the records
#define PT_THREADS 256
#define PT_ITEMS 8
#define PT_TILE (PT_THREADS * PT_ITEMS) // 2048 records

__global__ void __launch_bounds__(PT_THREADS) k_part_count(const uint32_t *__restrict__ xkeys, uint64_t n,
                                                           const uint32_t *__restrict__ cuts, uint32_t ncut, uint32_t nb,
                                                           const uint32_t *__restrict__ lut, uint16_t *__restrict__ bins,
                                                           uint32_t *__restrict__ tile_hist) {
	__shared__ uint32_t s_cuts[SP_MAXBINS / 2];
	__shared__ uint32_t s_hist[SP_MAXBINS];
	__shared__ uint16_t s_lut[SP_LUT + 1];
	for (uint32_t i = threadIdx.x; i < ncut; i += PT_THREADS) {
		s_cuts[i] = cuts[i];
	}
	for (uint32_t i = threadIdx.x; i < SP_LUT + 1; i += PT_THREADS) {
		s_lut[i] = (uint16_t)lut[2 + i];
	}
	const uint32_t lut_lo = lut[0], lut_shift = lut[1];
	for (uint32_t i = threadIdx.x; i < nb; i += PT_THREADS) {
		s_hist[i] = 0;
	}
	__syncthreads();
	const uint64_t base = (uint64_t)blockIdx.x * PT_TILE;
#pragma unroll
	for (int j = 0; j < PT_ITEMS; j++) {
		const uint64_t i = base + (uint64_t)j * PT_THREADS + threadIdx.x;
		if (i < n) {
			const uint32_t b = sp_bin_lut(xkeys[i], s_cuts, ncut, s_lut, lut_lo, lut_shift);
			bins[i] = (uint16_t)b;
			atomicAdd(&s_hist[b], 1u);
		}
	}
	__syncthreads();
	// bin-major table [bin][tile]: its exclusive scan, read in memory order, IS the first position of every (tile, bin)
	for (uint32_t i = threadIdx.x; i < nb; i += PT_THREADS) {
		tile_hist[(size_t)i * gridDim.x + blockIdx.x] = s_hist[i];
	}
}

// exclusive scan of a u32 array in three launches: sums of 2048-element blocks, scan of the sums (one CTA), apply
#define SC_BLOCK 2048
__global__ void __launch_bounds__(256) k_scan_sums(const uint32_t *__restrict__ v, uint64_t n, uint32_t *__restrict__ sums) {
	__shared__ uint32_t s_w[33];
	const uint64_t base = (uint64_t)blockIdx.x * SC_BLOCK + threadIdx.x * 8;
	uint32_t a = 0;
#pragma unroll
	for (int j = 0; j < 8; j++) {
		a += base + j < n ? v[base + j] : 0u;
	}
	uint32_t tot;
	sp_block_scan<256>(a, s_w, &tot);
	if (threadIdx.x == 0) {
		sums[blockIdx.x] = tot;
	}
}
__global__ void __launch_bounds__(1024) k_scan_top(uint32_t *__restrict__ sums, uint32_t nblk) {
	__shared__ uint32_t s_w[33];
	uint32_t carry = 0;
	for (uint32_t b0 = 0; b0 < nblk; b0 += 1024) {
		const uint32_t b = b0 + threadIdx.x;
		const uint32_t a = b < nblk ? sums[b] : 0u;
		uint32_t tot;
		const uint32_t e = sp_block_scan<1024>(a, s_w, &tot);
		if (b < nblk) {
			sums[b] = carry + e;
		}
		carry += tot;
	}
}
__global__ void __launch_bounds__(256) k_scan_apply(uint32_t *__restrict__ v, uint64_t n, const uint32_t *__restrict__ sums) {
	__shared__ uint32_t s_w[33];
	const uint64_t base = (uint64_t)blockIdx.x * SC_BLOCK + threadIdx.x * 8;
	uint32_t x[8], a = 0;
	if (base + 8 <= n) { // scratch arrays are 256-byte aligned and base is a multiple of 8: two 16-byte loads
		const uint4 lo = *reinterpret_cast<const uint4 *>(v + base), hi = *reinterpret_cast<const uint4 *>(v + base + 4);
		x[0] = lo.x, x[1] = lo.y, x[2] = lo.z, x[3] = lo.w, x[4] = hi.x, x[5] = hi.y, x[6] = hi.z, x[7] = hi.w;
	} else {
#pragma unroll
		for (int j = 0; j < 8; j++) {
			x[j] = base + j < n ? v[base + j] : 0u;
		}
	}
#pragma unroll
	for (int j = 0; j < 8; j++) {
		a += x[j];
	}
	uint32_t tot;
	uint32_t run = sums[blockIdx.x] + sp_block_scan<256>(a, s_w, &tot);
	uint32_t o[8];
#pragma unroll
	for (int j = 0; j < 8; j++) {
		o[j] = run;
		run += x[j];
	}
	if (base + 8 <= n) {
		*reinterpret_cast<uint4 *>(v + base) = make_uint4(o[0], o[1], o[2], o[3]);
		*reinterpret_cast<uint4 *>(v + base + 4) = make_uint4(o[4], o[5], o[6], o[7]);
	} else {
#pragma unroll
		for (int j = 0; j < 8; j++) {
			if (base + j < n) {
				v[base + j] = o[j];
			}
		}
	}
}

// stable scatter of a tile: rank of every record among the tile's records of its bin (warp multi-split over the bin
// bits + per-warp counters, warps in order), then the records go straight from the ranking threads to their place.
// A record is one 32-byte sector but a thread stores 16 bytes at a time: lanes work in pairs so that ONE store
// instruction carries both halves of a record and L2 receives whole sectors. Only the y keys (4-byte items, runs of
// ~7 per bin and tile) are staged in shared memory in tile order and leave as runs.
// Measured at C1 (10 M records, 559 bins, ncu): plain 16-byte stores + 4-byte y-key stores 332 us (29.5 M store sectors
// for 12.7 M of payload); whole tile staged in shared memory and written out in runs 264 us (minimal sectors, but 100 KB
// of shared memory = 2 CTAs / SM and 26 M shared-memory wavefronts); this form 179 us.
#define PS_KBINS ((SP_MAXBINS + PT_THREADS - 1) / PT_THREADS) // bins per thread in the per-bin steps
struct pd_smem {
	uint32_t yk[PT_TILE];
	uint32_t gl[SP_MAXBINS + 3];
	uint16_t wh[PT_THREADS / 32][SP_MAXBINS + 1];
	uint16_t doff[SP_MAXBINS + 3];
	uint16_t tot[PS_KBINS * PT_THREADS];
	uint16_t bin[PT_TILE];
	uint32_t scan[33];
};
__device__ __forceinline__ void pd_store16(void *p, float x, float y, float z, float w) {
	asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
// measured at C1 (partition scope): unroll 4 / 4 CTAs 0.261 ms, 4 / 5 0.252 (kept), 2 / 5 0.254, 8 / 3 0.253, 8 / 2 0.282
#ifndef PS_MIN_CTAS
#define PS_MIN_CTAS 5
#endif
#ifndef PS_UNROLL
#define PS_UNROLL 4
#endif
// NBITS: bits of a bin id (compile-time: the ballot loop has no per-bit predicate)
template <bool POINTS, int NBITS>
__global__ void __launch_bounds__(PT_THREADS, POINTS ? 4 : PS_MIN_CTAS) k_part_scatter(const uint16_t *__restrict__ bins, uint64_t n, uint32_t nb,
                                                                       const uint32_t *__restrict__ tile_pos,
                                                                       const sb_boxsrc box,
                                                                       const int64_t *__restrict__ rows,
                                                                       const uint32_t *__restrict__ ykeys,
                                                                       sb_rec32 *__restrict__ rec_out,
                                                                       uint32_t *__restrict__ ykey_out) {
	__shared__ pd_smem S;
	{
		const uint32_t words = (nb + 2) / 2;
		for (uint32_t i = threadIdx.x; i < (PT_THREADS / 32) * words; i += PT_THREADS) {
			reinterpret_cast<uint32_t *>(S.wh[i / words])[i % words] = 0;
		}
	}
	__syncthreads();
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
	const unsigned lt_mask = sb_lanemask_lt();
	const uint64_t tile_base = (uint64_t)blockIdx.x * PT_TILE;
	const uint32_t tile_count = (uint32_t)min((uint64_t)PT_TILE, n - tile_base);
	const uint64_t first = tile_base + (uint64_t)warp * (32 * PT_ITEMS) + lane;
	uint32_t br[PT_ITEMS];
	uint16_t *my = S.wh[warp];
	uint32_t r_pos[PS_KBINS];
#pragma unroll
	for (int i = 0; i < PT_ITEMS; i++) {
		const uint64_t gi = first + (uint64_t)i * 32;
		br[i] = gi < n ? bins[gi] : 0u;
	}
#pragma unroll
	for (int k = 0; k < PS_KBINS; k++) {
		const uint32_t b = threadIdx.x * PS_KBINS + k;
		r_pos[k] = b < nb ? tile_pos[(size_t)b * gridDim.x + blockIdx.x] : 0u;
	}
#pragma unroll
	for (int i = 0; i < PT_ITEMS; i++) {
		const bool valid = first + (uint64_t)i * 32 < n;
		const uint32_t b = br[i];
		unsigned peers = __ballot_sync(0xFFFFFFFFu, valid);
#pragma unroll
		for (int k = 0; k < NBITS; k++) {
			const bool bit = (b >> k) & 1u;
			const unsigned m = __ballot_sync(0xFFFFFFFFu, bit);
			peers &= bit ? m : ~m;
		}
		const int leader = __ffs(peers) - 1;
		uint32_t prev = 0;
		if ((int)lane == leader && valid) {
			prev = my[b];
			my[b] = (uint16_t)(prev + __popc(peers));
		}
		prev = __shfl_sync(0xFFFFFFFFu, prev, leader < 0 ? 0 : leader);
		br[i] = b | ((prev + __popc(peers & lt_mask)) << 16);
		__syncwarp();
	}
	__syncthreads();
	for (uint32_t b = threadIdx.x; b < PS_KBINS * PT_THREADS; b += PT_THREADS) {
		uint32_t run = 0;
		if (b < nb) {
#pragma unroll
			for (int w = 0; w < PT_THREADS / 32; w++) {
				const uint32_t c = S.wh[w][b];
				S.wh[w][b] = (uint16_t)run;
				run += c;
			}
		}
		S.tot[b] = (uint16_t)run;
	}
	__syncthreads();
	{
		uint32_t tot[PS_KBINS], sum = 0;
#pragma unroll
		for (int k = 0; k < PS_KBINS; k++) {
			tot[k] = S.tot[threadIdx.x * PS_KBINS + k];
			sum += tot[k];
		}
		uint32_t all;
		uint32_t run = sp_block_scan<PT_THREADS>(sum, S.scan, &all);
#pragma unroll
		for (int k = 0; k < PS_KBINS; k++) {
			const uint32_t b = threadIdx.x * PS_KBINS + k;
			if (b < nb) {
				S.doff[b] = (uint16_t)run;
				S.gl[b] = r_pos[k] - run;
			}
			run += tot[k];
		}
	}
	__syncthreads();
	const bool odd = lane & 1u;
	char *out = reinterpret_cast<char *>(rec_out);
	constexpr int kUnroll = PS_UNROLL;
#pragma unroll kUnroll
	for (int i = 0; i < PT_ITEMS; i++) {
		const uint64_t gi = first + (uint64_t)i * 32;
		uint32_t g = 0xFFFFFFFFu;
		float4 lo = make_float4(0.f, 0.f, 0.f, 0.f), hi = lo;
		if (gi < n) {
			const uint32_t b = br[i] & 0xFFFFu;
			const uint32_t p = (uint32_t)S.doff[b] + my[b] + (br[i] >> 16);
			g = S.gl[b] + p;
			lo = box.template load<POINTS>(gi);
			const int64_t r = rows[gi];
			const uint32_t yk = ykeys[gi];
			hi = make_float4(__int_as_float((int)(uint32_t)(uint64_t)r), __int_as_float((int)(uint32_t)((uint64_t)r >> 32)),
			                 __uint_as_float(yk), 0.0f);
			S.yk[p] = yk;
			S.bin[p] = (uint16_t)b;
		}
		// the partner gets the half of MY record it stores (even lane: my upper half, odd lane: my lower half) + its position
		const float gx = __shfl_xor_sync(0xFFFFFFFFu, odd ? lo.x : hi.x, 1);
		const float gy = __shfl_xor_sync(0xFFFFFFFFu, odd ? lo.y : hi.y, 1);
		const float gz = __shfl_xor_sync(0xFFFFFFFFu, odd ? lo.z : hi.z, 1);
		const float gw = __shfl_xor_sync(0xFFFFFFFFu, odd ? lo.w : hi.w, 1);
		const uint32_t pg = __shfl_xor_sync(0xFFFFFFFFu, g, 1);
		// round 1: the even lane's record (even lane: own lower half, odd lane: the even lane's upper half)
		const uint32_t t1 = odd ? pg : g;
		if (t1 != 0xFFFFFFFFu) {
			pd_store16(out + (size_t)t1 * 32 + (odd ? 16 : 0), odd ? gx : lo.x, odd ? gy : lo.y, odd ? gz : lo.z, odd ? gw : lo.w);
		}
		// round 2: the odd lane's record (even lane: the odd lane's lower half, odd lane: own upper half)
		const uint32_t t2 = odd ? g : pg;
		if (t2 != 0xFFFFFFFFu) {
			pd_store16(out + (size_t)t2 * 32 + (odd ? 16 : 0), odd ? hi.x : gx, odd ? hi.y : gy, odd ? hi.z : gz, odd ? hi.w : gw);
		}
	}
	__syncthreads();
	for (uint32_t p = threadIdx.x; p < tile_count; p += PT_THREADS) {
		ykey_out[S.gl[S.bin[p]] + p] = S.yk[p];
	}
}

int sbi_str_partition_rec(sb_context *ctx, sb_boxsrc box, const int64_t *rows, const uint32_t *xkeys,
                          const uint32_t *ykeys, uint64_t n, const uint32_t *cuts, uint32_t ncut, sb_rec32 *rec_out,
                          uint32_t *ykey_out) {
	const uint32_t nb = 2 * ncut + 1;
	SB_ARG(nb <= SP_MAXBINS, "STR partition: bin count");
	int nbits = 0;
	while ((1u << nbits) < nb) {
		nbits++;
	}
	const uint32_t tiles = (uint32_t)((n + PT_TILE - 1) / PT_TILE);
	const uint64_t cells = (uint64_t)tiles * nb;
	const uint32_t nblk = (uint32_t)((cells + SC_BLOCK - 1) / SC_BLOCK);
	uint16_t *bins = nullptr;
	uint32_t *tile_hist = nullptr, *sums = nullptr;
	SB_TRY(ctx->scratch_t(n, &bins));
	SB_TRY(ctx->scratch_t(cells, &tile_hist));
	SB_TRY(ctx->scratch_t(nblk, &sums));
	uint32_t *lut = nullptr;
	SB_TRY(ctx->scratch_t(SP_LUT + 3, &lut));
	sb_prof_scope ps(ctx, "str_partition");
	if (ncut) {
		k_part_lut<<<1, SP_LUT, 0, ctx->stream>>>(cuts, ncut, lut);
		SB_LAUNCH_CHECK(ctx);
	} else {
		SB_CUDA(cudaMemsetAsync(lut, 0, (SP_LUT + 3) * sizeof(uint32_t), ctx->stream)); // no cuts: every key -> bin 0
	}
	k_part_count<<<tiles, PT_THREADS, 0, ctx->stream>>>(xkeys, n, cuts, ncut, nb, lut, bins, tile_hist);
	SB_LAUNCH_CHECK(ctx);
	k_scan_sums<<<nblk, 256, 0, ctx->stream>>>(tile_hist, cells, sums);
	SB_LAUNCH_CHECK(ctx);
	k_scan_top<<<1, 1024, 0, ctx->stream>>>(sums, nblk);
	SB_LAUNCH_CHECK(ctx);
	k_scan_apply<<<nblk, 256, 0, ctx->stream>>>(tile_hist, cells, sums);
	SB_LAUNCH_CHECK(ctx);
#define PS_LAUNCH(P, B)                                                                                                \
	k_part_scatter<P, B><<<tiles, PT_THREADS, 0, ctx->stream>>>(bins, n, nb, tile_hist, box, rows, ykeys, rec_out, ykey_out)
#define PS_CASE(B)                                                                                                     \
	case B:                                                                                                            \
		if (box.box) {                                                                                                 \
			PS_LAUNCH(false, B);                                                                                       \
		} else {                                                                                                       \
			PS_LAUNCH(true, B);                                                                                        \
		}                                                                                                              \
		break
	switch (nbits) {
		PS_CASE(0);
		PS_CASE(1);
		PS_CASE(2);
		PS_CASE(3);
		PS_CASE(4);
		PS_CASE(5);
		PS_CASE(6);
		PS_CASE(7);
		PS_CASE(8);
		PS_CASE(9);
		PS_CASE(10);
	default:
		PS_CASE(11);
	}
#undef PS_CASE
#undef PS_LAUNCH
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// =====================================================================================================================
// The y order of the leaf level without a sort. A leaf node's entries are re-ordered by row id when the node is packed
// (RTreeNode::SortEntriesByRowId), so of the y order inside a slice the leaf level only uses WHICH node an entry falls
// in: the same select + stable partition as for the x order, once per slice -- the y keys at ranks cap, 2 * cap, ...
// of every slice (k_ysel_slice: one CTA per slice, all histogram levels in shared memory) and a partition of the
// slice's positions into the bins of those keys (k_ypart_count / scan / k_ypart_scatter). Equal y keys that straddle a
// node boundary are cut by their order in the slice -- (x bin, input position) -- exactly as the stable segmented sort
// cut them, so the nodes are the same entry for entry. Replaces four segmented radix passes + their histogram pass.
// =====================================================================================================================
#ifndef YS_THREADS
#define YS_THREADS 1024 // 512 measured slower (0.336 vs 0.318 ms for both y phases at C1); must be >= YS_NSP
#endif
#define YS_NSP 512u     // ranks per slice (slice / cap - 1 <= 511)
#define YS_L2 512u      // key bits 20..12
#define YS_L34 64u      // key bits 11..6 and 5..0
#define YS_CHUNK 24u    // level-1 buckets whose level-2 rows fit the table at once
#define YS_REPL 4u      // level-1 histogram replicas (float keys share their top bits)

// Shared memory is carved by the number of ranks of a full slice (nspc) so that two CTAs fit an SM at C1 (105 KB):
// one wave of slices instead of two, and twice the loads in flight (the kernel is latency bound: four sweeps over the
// slice's keys). tab holds, in turn: YS_REPL x 2048 level-1 counters, YS_CHUNK x 512 level-2 counters, nspc x 64 level-3 /
// level-4 counters.
struct ys_smem {
	uint32_t *tab;
	uint32_t *pre1;   // [2049]
	uint32_t *bitmap; // [nspc * 16] per level-1 cut bucket: which level-2 buckets hold a rank
	uint32_t *rem;    // [nspc] rank inside the current prefix
	uint32_t *p2, *p3; // [nspc] 20-bit / 26-bit key prefix of every rank
	uint32_t *scan;   // [33]
	uint16_t *map1;   // [2048]
	uint16_t *b1, *c1, *c2, *c3; // [nspc]
	uint16_t *wpre;   // [nspc * 16] set bits of the bitmap before each word
	uint16_t *firstk; // [nspc + 1] first rank of each distinct 20-bit prefix
};
__host__ __device__ inline uint32_t ys_tab_words(uint32_t nspc) {
	uint32_t w = YS_REPL * SEL_L1;
	w = w > YS_CHUNK * YS_L2 ? w : YS_CHUNK * YS_L2;
	return w > nspc * YS_L34 ? w : nspc * YS_L34;
}
__host__ __device__ inline size_t ys_smem_bytes(uint32_t nspc) {
	// pre1 (level 1 only) and wpre / firstk (levels 3, 4) share their storage
	const size_t late = sizeof(uint16_t) * ((size_t)nspc * (YS_L2 / 32) + nspc + 2);
	const size_t early = sizeof(uint32_t) * (SEL_L1 + 1);
	return sizeof(uint32_t) * ((size_t)ys_tab_words(nspc) + (size_t)nspc * (YS_L2 / 32) + 3 * (size_t)nspc + 33 + 3) +
	       sizeof(uint16_t) * (SEL_L1 + 4 * (size_t)nspc + 8) + (late > early ? late : early) + 16;
}
__device__ __forceinline__ ys_smem ys_carve(unsigned char *raw, uint32_t nspc) {
	ys_smem S;
	uint32_t *w = reinterpret_cast<uint32_t *>(raw);
	S.tab = w, w += ys_tab_words(nspc);
	S.bitmap = w, w += (size_t)nspc * (YS_L2 / 32);
	S.rem = w, w += nspc;
	S.p2 = w, w += nspc;
	S.p3 = w, w += nspc;
	S.scan = w, w += 33;
	uint16_t *h = reinterpret_cast<uint16_t *>(w);
	S.map1 = h, h += SEL_L1;
	S.b1 = h, h += nspc;
	S.c1 = h, h += nspc;
	S.c2 = h, h += nspc;
	S.c3 = h, h += nspc;
	h += (reinterpret_cast<uintptr_t>(h) & 2u) ? 1 : 0; // 4-byte alignment for pre1
	S.pre1 = reinterpret_cast<uint32_t *>(h);
	S.wpre = h, h += (size_t)nspc * (YS_L2 / 32);
	S.firstk = h;
	return S;
}

// the bucket of `row` (32 * PER counters) that holds rank `rem`, by the whole warp; returns bucket, *left = rank inside it
template <int PER>
__device__ __forceinline__ uint32_t ys_warp_find(const uint32_t *row, uint32_t rem, unsigned lane, uint32_t *left) {
	uint32_t v[PER], sum = 0;
#pragma unroll
	for (int j = 0; j < PER; j++) {
		v[j] = row[lane * PER + j];
		sum += v[j];
	}
	uint32_t inc = sum;
#pragma unroll
	for (int s = 1; s < 32; s <<= 1) {
		const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
		if (lane >= (unsigned)s) {
			inc += o;
		}
	}
	const uint32_t excl = inc - sum;
	const unsigned holders = __ballot_sync(0xFFFFFFFFu, excl <= rem);
	const int L = 31 - __clz(holders);
	uint32_t bucket = 0, r = rem - excl;
	if ((int)lane == L) {
#pragma unroll
		for (int j = 0; j < PER; j++) {
			if (r < v[j]) {
				bucket = lane * PER + j;
				break;
			}
			r -= v[j];
		}
	}
	*left = __shfl_sync(0xFFFFFFFFu, r, L);
	return __shfl_sync(0xFFFFFFFFu, bucket, L);
}

// first index of the sorted array whose value is >= v
__device__ __forceinline__ uint32_t ys_lower_bound(const uint32_t *a, uint32_t n, uint32_t v) {
	uint32_t lo = 0, hi = n;
	while (lo < hi) {
		const uint32_t mid = (lo + hi) >> 1;
		if (a[mid] < v) {
			lo = mid + 1;
		} else {
			hi = mid;
		}
	}
	return lo;
}

// index of each rank's prefix among the distinct prefixes (p sorted, one per rank): c[k]; returns the distinct count
__device__ __forceinline__ uint32_t ys_distinct(const uint32_t *p, uint32_t nsp, uint16_t *c, uint32_t *s_scan) {
	const uint32_t k = threadIdx.x;
	const uint32_t flag = (k < nsp && (k == 0 || p[k] != p[k - 1])) ? 1u : 0u;
	uint32_t tot;
	const uint32_t e = sp_block_scan<YS_THREADS>(flag, s_scan, &tot);
	if (k < nsp) {
		c[k] = (uint16_t)(e + flag - 1);
	}
	__syncthreads();
	return tot;
}

__global__ void __launch_bounds__(YS_THREADS, 2) k_ysel_slice(const uint32_t *__restrict__ ykeys, uint64_t n, uint64_t slice,
                                                              uint32_t cap, uint32_t stride, uint32_t *__restrict__ ycuts,
                                                              uint32_t *__restrict__ yluts, uint32_t *__restrict__ perm,
                                                              uint16_t *__restrict__ bins) {
	extern __shared__ __align__(16) unsigned char ys_raw[];
	const ys_smem S = ys_carve(ys_raw, stride);
	const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	const uint64_t beg = (uint64_t)blockIdx.x * slice;
	const uint32_t m = (uint32_t)min(slice, n - beg);
	const uint32_t nsp = (m + cap - 1) / cap - 1;
	if (nsp == 0) {
		if (perm) { // one node: any order
			for (uint32_t i = tid; i < m; i += YS_THREADS) {
				perm[beg + i] = (uint32_t)(beg + i);
			}
		}
		return;
	}
	const uint32_t *keys = ykeys + beg;
	uint32_t *cuts = ycuts + (size_t)blockIdx.x * stride;

	// ---- level 1: top 11 bits, 8 replicas (float keys share their top bits) ----
	for (uint32_t i = tid; i < YS_REPL * SEL_L1; i += YS_THREADS) {
		S.tab[i] = 0;
	}
	__syncthreads();
	{
		uint32_t *mine = S.tab + (lane & (YS_REPL - 1)) * SEL_L1;
#pragma unroll 4
		for (uint32_t i = tid; i < m; i += YS_THREADS) {
			atomicAdd(&mine[keys[i] >> 21], 1u);
		}
	}
	__syncthreads();
	{
		constexpr int PERB = SEL_L1 / YS_THREADS; // consecutive level-1 bins per thread
		uint32_t v[PERB], sum = 0;
#pragma unroll
		for (int j = 0; j < PERB; j++) {
			v[j] = 0;
#pragma unroll
			for (int r = 0; r < (int)YS_REPL; r++) {
				v[j] += S.tab[r * SEL_L1 + PERB * tid + j];
			}
			sum += v[j];
		}
		uint32_t total;
		uint32_t run = sp_block_scan<YS_THREADS>(sum, S.scan, &total);
#pragma unroll
		for (int j = 0; j < PERB; j++) {
			S.pre1[PERB * tid + j] = run;
			S.map1[PERB * tid + j] = 0;
			run += v[j];
		}
		if (tid == 0) {
			S.pre1[SEL_L1] = total;
		}
	}
	__syncthreads();
	if (tid < nsp) {
		const uint32_t r = (tid + 1) * cap;
		const uint32_t bk = sp_bucket_of(S.pre1, SEL_L1, r);
		S.b1[tid] = (uint16_t)bk;
		S.rem[tid] = r - S.pre1[bk];
		S.p2[tid] = bk; // level-1 prefix for now
	}
	__syncthreads();
	const uint32_t ncut1 = ys_distinct(S.p2, nsp, S.c1, S.scan);
	if (tid < nsp && (tid == 0 || S.b1[tid] != S.b1[tid - 1])) {
		S.map1[S.b1[tid]] = (uint16_t)(S.c1[tid] + 1);
	}
	for (uint32_t i = tid; i < ncut1 * (YS_L2 / 32); i += YS_THREADS) {
		S.bitmap[i] = 0;
	}
	__syncthreads();

	// ---- level 2: 9 bits, YS_CHUNK level-1 buckets at a time ----
	for (uint32_t c0 = 0; c0 < ncut1; c0 += YS_CHUNK) {
		const uint32_t nc = min(YS_CHUNK, ncut1 - c0);
		for (uint32_t i = tid; i < nc * YS_L2; i += YS_THREADS) {
			S.tab[i] = 0;
		}
		__syncthreads();
#pragma unroll 4
		for (uint32_t i = tid; i < m; i += YS_THREADS) {
			const uint32_t key = keys[i];
			const uint32_t c = (uint32_t)S.map1[key >> 21] - 1u - c0; // map1 == 0 wraps to a large value
			if (c < nc) {
				atomicAdd(&S.tab[c * YS_L2 + ((key >> SEL_SH2) & (YS_L2 - 1))], 1u);
			}
		}
		__syncthreads();
		for (uint32_t k = warp; k < nsp; k += YS_THREADS / 32) {
			const uint32_t c = (uint32_t)S.c1[k] - c0;
			if (c < nc) { // warp-uniform
				uint32_t left;
				const uint32_t b2 = ys_warp_find<YS_L2 / 32>(S.tab + c * YS_L2, S.rem[k], lane, &left);
				if (lane == 0) {
					S.rem[k] = left;
					S.p2[k] = ((uint32_t)S.b1[k] << 9) | b2;
					atomicOr(&S.bitmap[(uint32_t)S.c1[k] * (YS_L2 / 32) + (b2 >> 5)], 1u << (b2 & 31u));
				}
			}
		}
		__syncthreads();
	}
	const uint32_t ncut2 = ys_distinct(S.p2, nsp, S.c2, S.scan);

	// The 20-bit prefixes that hold a rank are the set bits of the bitmap, in row-major = ascending order: the index of a
	// prefix among them (= c2 of its ranks) is a popcount rank, so a key finds its counter row without a search
	// (a binary search per passing key in divergent code was 53 % of this kernel's instructions).
	{
		const uint32_t nw = ncut1 * (YS_L2 / 32);
		uint32_t carry = 0;
		for (uint32_t w0 = 0; w0 < nw; w0 += YS_THREADS) {
			const uint32_t w = w0 + tid;
			const uint32_t pc = w < nw ? __popc(S.bitmap[w]) : 0u;
			uint32_t tot;
			const uint32_t e = sp_block_scan<YS_THREADS>(pc, S.scan, &tot);
			if (w < nw) {
				S.wpre[w] = (uint16_t)(carry + e);
			}
			carry += tot;
		}
		if (tid < nsp && (tid == 0 || S.p2[tid] != S.p2[tid - 1])) {
			S.firstk[S.c2[tid]] = (uint16_t)tid;
		}
		if (tid == 0) {
			S.firstk[ncut2] = (uint16_t)nsp;
		}
	}
	__syncthreads();

	// ---- levels 3 and 4: 6 bits each, only keys whose 20-bit prefix holds a rank ----
	for (int level = 3; level <= 4; level++) {
		const uint16_t *cidx = level == 3 ? S.c2 : S.c3;
		for (uint32_t i = tid; i < nsp * YS_L34; i += YS_THREADS) {
			S.tab[i] = 0;
		}
		__syncthreads();
#pragma unroll 4
		for (uint32_t i = tid; i < m; i += YS_THREADS) {
			const uint32_t key = keys[i];
			const uint32_t m1 = S.map1[key >> 21];
			if (m1) {
				const uint32_t mid = (key >> SEL_SH2) & (YS_L2 - 1);
				const uint32_t w = (m1 - 1) * (YS_L2 / 32) + (mid >> 5);
				const uint32_t bits = S.bitmap[w];
				if ((bits >> (mid & 31u)) & 1u) {
					const uint32_t c2 = (uint32_t)S.wpre[w] + __popc(bits & ((1u << (mid & 31u)) - 1u));
					if (level == 3) {
						atomicAdd(&S.tab[c2 * YS_L34 + ((key >> 6) & (YS_L34 - 1))], 1u);
					} else {
						const uint32_t want = key >> 6;
						for (uint32_t k = S.firstk[c2]; k < S.firstk[c2 + 1]; k++) {
							if (S.p3[k] == want) { // ranks with one 26-bit prefix are adjacent and share c3
								atomicAdd(&S.tab[(uint32_t)S.c3[k] * YS_L34 + (key & (YS_L34 - 1))], 1u);
								break;
							}
						}
					}
				}
			}
		}
		__syncthreads();
		for (uint32_t k = warp; k < nsp; k += YS_THREADS / 32) {
			uint32_t left;
			const uint32_t b = ys_warp_find<YS_L34 / 32>(S.tab + (uint32_t)cidx[k] * YS_L34, S.rem[k], lane, &left);
			if (lane == 0) {
				S.rem[k] = left;
				if (level == 3) {
					S.p3[k] = (S.p2[k] << 6) | b;
				} else {
					cuts[k] = S.p2[k] = (S.p3[k] << 6) | b; // p2 is free by now: the slice's cut keys stay in shared memory
				}
			}
		}
		__syncthreads();
		if (level == 3) {
			ys_distinct(S.p3, nsp, S.c3, S.scan);
		}
	}
	// the slice's table for the bin search of the partition (see k_part_lut): [0] = first cut, [1] = shift, [2 ..] = table
	{
		uint32_t *lut = yluts + (size_t)blockIdx.x * (SP_LUT + 3);
		const uint32_t lo = S.p2[0], hi = S.p2[nsp - 1];
		const int bits = 32 - __clz((hi - lo) | 1u);
		const int shift = bits > 10 ? bits - 10 : 0;
		constexpr int PERL = SP_LUT / YS_THREADS; // table entries per thread
		uint32_t la[PERL];
#pragma unroll
		for (int j = 0; j < PERL; j++) {
			const uint32_t c = tid + j * YS_THREADS;
			const unsigned long long start = (unsigned long long)lo + ((unsigned long long)c << shift);
			uint32_t a = 0, b = nsp;
			while (a < b) {
				const uint32_t mid = (a + b) >> 1;
				if ((unsigned long long)S.p2[mid] < start) {
					a = mid + 1;
				} else {
					b = mid;
				}
			}
			la[j] = a;
			lut[2 + c] = a;
		}
		if (tid == 0) {
			lut[0] = lo;
			lut[1] = (uint32_t)shift;
			lut[2 + SP_LUT] = nsp;
		}
		if (!perm) {
			return;
		}
		// ---- the slice's positions in node order, by the same CTA (slices of <= 65535 entries) -----------------------
		// Warp w owns the w-th 32nd of the slice. Pass A counts its keys per bin, a scan turns the counts into the first
		// place of (bin, warp), pass B ranks each row of 32 keys with ballots against the warp's running cursors: the
		// result is stable in (bin, position), no count table, bin array or second launch goes through global memory.
		uint16_t *s_lut = reinterpret_cast<uint16_t *>(S.pre1); // wpre / firstk / pre1 are dead: their storage holds the table
		__syncthreads();
#pragma unroll
		for (int j = 0; j < PERL; j++) {
			s_lut[tid + j * YS_THREADS] = (uint16_t)la[j];
		}
		if (tid == 0) {
			s_lut[SP_LUT] = (uint16_t)nsp;
		}
		constexpr uint32_t NW = YS_THREADS / 32;
		const uint32_t nb = 2 * nsp + 1, nbp = (nb + 2) & ~1u; // row length of the per-warp counters (even)
		uint16_t *wh = reinterpret_cast<uint16_t *>(S.tab);     // [NW][nbp]
		uint32_t *base = S.tab + (NW / 2) * nbp;                 // [nb] first place of the bin
		for (uint32_t i = tid; i < (NW / 2) * nbp; i += YS_THREADS) {
			S.tab[i] = 0;
		}
		__syncthreads();
		const uint32_t chunk = ((m + YS_THREADS - 1) / YS_THREADS) * 32u; // positions per warp, whole rows
		const uint32_t w0 = warp * chunk;
		uint32_t *wrow32 = S.tab + (warp * nbp) / 2;
		// pass A also leaves every key's bin (bit 15: the key equals a cut key) for pass B
#pragma unroll 4
		for (uint32_t r = 0; r < chunk; r += 32) {
			const uint32_t i = w0 + r + lane;
			if (i < m) {
				bool run;
				const uint32_t bin = sp_bin_lut(keys[i], S.p2, nsp, s_lut, lo, (uint32_t)shift, &run);
				bins[beg + i] = (uint16_t)(bin | (run ? 0x8000u : 0u));
				atomicAdd(&wrow32[bin >> 1], 1u << (16u * (bin & 1u)));
			}
		}
		__syncthreads();
		{
			uint32_t carry = 0;
			for (uint32_t b0 = 0; b0 < nb; b0 += YS_THREADS) {
				const uint32_t b = b0 + tid;
				uint32_t run = 0;
				if (b < nb) {
#pragma unroll 8
					for (uint32_t w = 0; w < NW; w++) {
						const uint32_t c = wh[w * nbp + b];
						wh[w * nbp + b] = (uint16_t)run;
						run += c;
					}
				}
				uint32_t tot;
				const uint32_t e = sp_block_scan<YS_THREADS>(run, S.scan, &tot);
				if (b < nb) {
					base[b] = carry + e;
				}
				carry += tot;
			}
		}
		__syncthreads();
		const unsigned lt_mask = sb_lanemask_lt();
		uint16_t *my = wh + warp * nbp;
		// four rows at a time: the key loads and bin searches of a batch overlap, only the cursor updates are sequential
		for (uint32_t r0 = 0; r0 < chunk; r0 += 128) {
			uint32_t bins4[4];
#pragma unroll
			for (int q = 0; q < 4; q++) {
				const uint32_t i = w0 + r0 + q * 32 + lane;
				bins4[q] = (r0 + q * 32 < chunk && i < m) ? bins[beg + i] : 0xFFFFFFFFu;
			}
#pragma unroll
			for (int q = 0; q < 4; q++) {
				const uint32_t i = w0 + r0 + q * 32 + lane;
				const bool valid = bins4[q] != 0xFFFFFFFFu;
				const bool run = valid && (bins4[q] & 0x8000u);
				const uint32_t bin = valid ? bins4[q] & 0x7FFFu : 0u;
				unsigned peers = __ballot_sync(0xFFFFFFFFu, valid);
				if (peers == 0) {
					continue; // warp-uniform: an empty row
				}
				// Keys strictly between two cut keys lie inside ONE node, whose entries are re-ordered by row id anyway:
				// their order is free, a shared-memory atomic on the warp's cursor places them. Keys EQUAL to a cut key
				// (one key in 128 is one, so a quarter of the rows hold some) form runs that may straddle a node
				// boundary and are cut by position: the lanes of each such bin are ranked by one ballot, in lane order,
				// against the same cursor. (A 10-ballot multi-split of every row was 2/3 of this phase's instructions.)
				if (valid && !run) {
					const uint32_t sh16 = 16u * (bin & 1u);
					const uint32_t old = atomicAdd(reinterpret_cast<uint32_t *>(my) + (bin >> 1), 1u << sh16);
					perm[beg + base[bin] + ((old >> sh16) & 0xFFFFu)] = (uint32_t)(beg + i);
				}
				__syncwarp();
				unsigned todo = __ballot_sync(0xFFFFFFFFu, valid && run);
				while (todo) {
					const int ld = __ffs(todo) - 1;
					const uint32_t b = __shfl_sync(0xFFFFFFFFu, bin, ld);
					const bool mine = valid && run && bin == b;
					const unsigned same = __ballot_sync(0xFFFFFFFFu, mine);
					uint32_t prev = 0;
					if ((int)lane == ld) {
						prev = my[b];
						my[b] = (uint16_t)(prev + __popc(same));
					}
					prev = __shfl_sync(0xFFFFFFFFu, prev, ld);
					if (mine) {
						perm[beg + base[b] + prev + __popc(same & lt_mask)] = (uint32_t)(beg + i);
					}
					todo &= ~same;
				}
				__syncwarp();
			}
		}
	}
}

// bins of a slice's positions among the slice's y cuts + the [slice][bin][tile] count table (tiles do not straddle slices)
__global__ void __launch_bounds__(PT_THREADS) k_ypart_count(const uint32_t *__restrict__ ykeys, uint64_t n, uint64_t slice,
                                                            uint32_t cap, const uint32_t *__restrict__ ycuts, uint32_t stride,
                                                            const uint32_t *__restrict__ yluts, uint32_t nb,
                                                            uint16_t *__restrict__ bins, uint32_t *__restrict__ tile_hist) {
	__shared__ uint32_t s_cuts[SP_MAXBINS / 2];
	__shared__ uint32_t s_hist[SP_MAXBINS];
	__shared__ uint16_t s_lut[SP_LUT + 1];
	const uint32_t sl = blockIdx.y, tile = blockIdx.x, tps = gridDim.x;
	const uint64_t sbeg = (uint64_t)sl * slice;
	const uint32_t m = (uint32_t)min(slice, n - sbeg);
	const uint32_t ncut = (m + cap - 1) / cap - 1;
	for (uint32_t i = threadIdx.x; i < ncut; i += PT_THREADS) {
		s_cuts[i] = ycuts[(size_t)sl * stride + i];
	}
	for (uint32_t i = threadIdx.x; i < nb; i += PT_THREADS) {
		s_hist[i] = 0;
	}
	// a slice without cuts (a short last slice of one node) has no table: lo = 0, shift = 0, all entries 0 -> bin 0
	const uint32_t *lut = yluts + (size_t)sl * (SP_LUT + 3);
	for (uint32_t i = threadIdx.x; i < SP_LUT + 1; i += PT_THREADS) {
		s_lut[i] = ncut ? (uint16_t)lut[2 + i] : (uint16_t)0;
	}
	const uint32_t lut_lo = ncut ? lut[0] : 0u, lut_shift = ncut ? lut[1] : 0u;
	__syncthreads();
	const uint32_t t0 = tile * PT_TILE;
#pragma unroll
	for (int j = 0; j < PT_ITEMS; j++) {
		const uint32_t i = t0 + (uint32_t)j * PT_THREADS + threadIdx.x;
		if (i < m) {
			const uint32_t b = sp_bin_lut(ykeys[sbeg + i], s_cuts, ncut, s_lut, lut_lo, lut_shift);
			bins[sbeg + i] = (uint16_t)b;
			atomicAdd(&s_hist[b], 1u);
		}
	}
	__syncthreads();
	uint32_t *tab = tile_hist + (size_t)sl * nb * tps;
	for (uint32_t i = threadIdx.x; i < nb; i += PT_THREADS) {
		tab[(size_t)i * tps + tile] = s_hist[i];
	}
}

// stable scatter of a slice tile's POSITIONS into node order: perm[first position of (slice, bin, tile) + rank] = position
struct yp_smem {
	uint32_t val[PT_TILE];
	uint32_t gl[SP_MAXBINS + 3];
	uint16_t wh[PT_THREADS / 32][SP_MAXBINS + 1];
	uint16_t doff[SP_MAXBINS + 3];
	uint16_t tot[PS_KBINS * PT_THREADS];
	uint16_t bin[PT_TILE];
	uint32_t scan[33];
};
__global__ void __launch_bounds__(PT_THREADS, 5) k_ypart_scatter(const uint16_t *__restrict__ bins, uint64_t n, uint64_t slice,
                                                                 uint32_t nb, int nbits, const uint32_t *__restrict__ tile_pos,
                                                                 uint32_t *__restrict__ perm) {
	__shared__ yp_smem S;
	const uint32_t sl = blockIdx.y, tile = blockIdx.x, tps = gridDim.x;
	const uint64_t sbeg = (uint64_t)sl * slice;
	const uint32_t m = (uint32_t)min(slice, n - sbeg);
	const uint32_t t0 = tile * PT_TILE;
	if (t0 >= m) {
		return;
	}
	{
		const uint32_t words = (nb + 2) / 2;
		for (uint32_t i = threadIdx.x; i < (PT_THREADS / 32) * words; i += PT_THREADS) {
			reinterpret_cast<uint32_t *>(S.wh[i / words])[i % words] = 0;
		}
	}
	__syncthreads();
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
	const unsigned lt_mask = sb_lanemask_lt();
	const uint32_t tile_count = min((uint32_t)PT_TILE, m - t0);
	const uint32_t first = t0 + warp * (32 * PT_ITEMS) + lane; // position inside the slice
	uint32_t br[PT_ITEMS];
	uint16_t *my = S.wh[warp];
	uint32_t r_pos[PS_KBINS];
#pragma unroll
	for (int i = 0; i < PT_ITEMS; i++) {
		const uint32_t gi = first + (uint32_t)i * 32;
		br[i] = gi < m ? bins[sbeg + gi] : 0u;
	}
	const uint32_t *tab = tile_pos + (size_t)sl * nb * tps;
#pragma unroll
	for (int k = 0; k < PS_KBINS; k++) {
		const uint32_t b = threadIdx.x * PS_KBINS + k;
		r_pos[k] = b < nb ? tab[(size_t)b * tps + tile] : 0u;
	}
#pragma unroll
	for (int i = 0; i < PT_ITEMS; i++) {
		const bool valid = first + (uint32_t)i * 32 < m;
		const uint32_t b = br[i];
		unsigned peers = __ballot_sync(0xFFFFFFFFu, valid);
#pragma unroll
		for (int k = 0; k < 11; k++) {
			if (k < nbits) {
				const bool bit = (b >> k) & 1u;
				const unsigned mm = __ballot_sync(0xFFFFFFFFu, bit);
				peers &= bit ? mm : ~mm;
			}
		}
		const int leader = __ffs(peers) - 1;
		uint32_t prev = 0;
		if ((int)lane == leader && valid) {
			prev = my[b];
			my[b] = (uint16_t)(prev + __popc(peers));
		}
		prev = __shfl_sync(0xFFFFFFFFu, prev, leader < 0 ? 0 : leader);
		br[i] = b | ((prev + __popc(peers & lt_mask)) << 16);
		__syncwarp();
	}
	__syncthreads();
	for (uint32_t b = threadIdx.x; b < PS_KBINS * PT_THREADS; b += PT_THREADS) {
		uint32_t run = 0;
		if (b < nb) {
#pragma unroll
			for (int w = 0; w < PT_THREADS / 32; w++) {
				const uint32_t c = S.wh[w][b];
				S.wh[w][b] = (uint16_t)run;
				run += c;
			}
		}
		S.tot[b] = (uint16_t)run;
	}
	__syncthreads();
	{
		uint32_t tot[PS_KBINS], sum = 0;
#pragma unroll
		for (int k = 0; k < PS_KBINS; k++) {
			tot[k] = S.tot[threadIdx.x * PS_KBINS + k];
			sum += tot[k];
		}
		uint32_t all;
		uint32_t run = sp_block_scan<PT_THREADS>(sum, S.scan, &all);
#pragma unroll
		for (int k = 0; k < PS_KBINS; k++) {
			const uint32_t b = threadIdx.x * PS_KBINS + k;
			if (b < nb) {
				S.doff[b] = (uint16_t)run;
				S.gl[b] = r_pos[k] - run;
			}
			run += tot[k];
		}
	}
	__syncthreads();
#pragma unroll
	for (int i = 0; i < PT_ITEMS; i++) {
		const uint32_t gi = first + (uint32_t)i * 32;
		if (gi < m) {
			const uint32_t b = br[i] & 0xFFFFu;
			const uint32_t p = (uint32_t)S.doff[b] + my[b] + (br[i] >> 16);
			S.val[p] = (uint32_t)(sbeg + gi);
			S.bin[p] = (uint16_t)b;
		}
	}
	__syncthreads();
	for (uint32_t p = threadIdx.x; p < tile_count; p += PT_THREADS) {
		perm[S.gl[S.bin[p]] + p] = S.val[p];
	}
}

// perm[j] = position (in the x-partitioned order) of the j-th entry of the leaf level's node order, up to the order of the
// entries inside a node. ykeys: y keys in x-partitioned order. Requires slice / cap - 1 <= 511.
int sbi_str_node_partition(sb_context *ctx, const uint32_t *ykeys, uint64_t n, uint64_t slice, uint32_t cap, uint32_t *perm) {
	const uint32_t nsp_max = (uint32_t)(slice / cap) - 1; // ranks of a full slice
	SB_ARG(slice % cap == 0 && nsp_max < YS_NSP, "STR node partition: slice / cap");
	const uint32_t nslices = (uint32_t)((n + slice - 1) / slice);
	const uint32_t stride = nsp_max ? nsp_max : 1;
	const uint32_t nb = 2 * nsp_max + 1;
	int nbits = 0;
	while ((1u << nbits) < nb) {
		nbits++;
	}
	const uint32_t tps = (uint32_t)((slice + PT_TILE - 1) / PT_TILE);
	const uint64_t cells = (uint64_t)nslices * nb * tps;
	const uint32_t nblk = (uint32_t)((cells + SC_BLOCK - 1) / SC_BLOCK);
	uint32_t *ycuts = nullptr, *yluts = nullptr, *tile_hist = nullptr, *sums = nullptr;
	uint16_t *bins = nullptr;
	SB_TRY(ctx->scratch_t((size_t)nslices * stride, &ycuts));
	SB_TRY(ctx->scratch_t((size_t)nslices * (SP_LUT + 3), &yluts));
	SB_TRY(ctx->scratch_t(n, &bins));
	SB_TRY(ctx->scratch_t(cells, &tile_hist));
	SB_TRY(ctx->scratch_t(nblk, &sums));
	const bool fused = slice <= 65535 && nsp_max > 0 && ctx->str_path != 3; // 16-bit cursors inside a slice
	{
		sb_prof_scope ps(ctx, "str_select");
		if (nsp_max) {
			SB_SMEM_OPTIN(ctx, k_ysel_slice, ys_smem_bytes(YS_NSP));
			k_ysel_slice<<<nslices, YS_THREADS, ys_smem_bytes(stride), ctx->stream>>>(ykeys, n, slice, cap, stride, ycuts, yluts,
			                                                                         fused ? perm : nullptr, bins);
			SB_LAUNCH_CHECK(ctx);
		}
	}
	if (fused) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "str_partition");
	// every (slice, tile) CTA writes its whole column of the table (zeros past the end of a short last slice)
	const dim3 grid(tps, nslices);
	k_ypart_count<<<grid, PT_THREADS, 0, ctx->stream>>>(ykeys, n, slice, cap, ycuts, stride, yluts, nb, bins, tile_hist);
	SB_LAUNCH_CHECK(ctx);
	k_scan_sums<<<nblk, 256, 0, ctx->stream>>>(tile_hist, cells, sums);
	SB_LAUNCH_CHECK(ctx);
	k_scan_top<<<1, 1024, 0, ctx->stream>>>(sums, nblk);
	SB_LAUNCH_CHECK(ctx);
	k_scan_apply<<<nblk, 256, 0, ctx->stream>>>(tile_hist, cells, sums);
	SB_LAUNCH_CHECK(ctx);
	k_ypart_scatter<<<grid, PT_THREADS, 0, ctx->stream>>>(bins, n, slice, nb, nbits, tile_hist, perm);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}
