// radix_sort.cu -- K3: stable LSD radix sort of (key, u32 value) pairs, "onesweep" style:
//   one histogram kernel reads the keys once and builds the digit histograms of ALL passes;
//   each pass is then ONE kernel that ranks a tile with warp-cooperative match/ballot multi-split, resolves the
//   tile's global digit offsets with a decoupled look-back over per-tile status words (no separate scan pass),
//   and scatters through shared memory so the global writes are runs of same-digit items.
// Traffic per pass: n*(key+value) read + n*(key+value) write, plus 1 KB of status words per 4096-item tile.
// Replaces DuckDB's PhysicalOrder in the CREATE INDEX plan and FlatRTree::Sort (see include/spatial_b200.h).
#include "internal.cuh"

#define RS_THREADS 256
#define RS_WARPS (RS_THREADS / 32)
#define RS_BINS 256
#define RS_FLAG_AGG 0x40000000u
#define RS_FLAG_INC 0x80000000u
#define RS_VALUE_MASK 0x3FFFFFFFu
#define RS_LOOK 8 // predecessor status words fetched per look-back round
// onesweep pass kernel shape (tile = OS_THREADS * items keys). Measured on the C1 build (16 launches, 10 M keys, sum of the
// onesweep scopes): 256 threads x 16 keys, 3 CTAs / SM (80 registers, spills) 0.873 ms; 4 CTAs / SM with late value loads
// and packed ranks 1.00 - 1.06 ms; 5 CTAs 4.3 ms (spills); 512 threads x 8 keys, 2 CTAs / SM 0.801 ms (kept);
// 512 x 8 with 3 CTAs 0.81 - 0.88 ms. Occupancy is not what bounds the pass.
#ifndef OS_THREADS
#define OS_THREADS 512
#endif
#ifndef OS_ITEMS32
#define OS_ITEMS32 8
#endif
#ifndef OS_ITEMS64
#define OS_ITEMS64 6
#endif
#ifndef OS_MIN_CTAS
#define OS_MIN_CTAS 2
#endif
#ifndef OS_LATE_VALS
#define OS_LATE_VALS 0 // 1: values are loaded after the ranking (not live across it), 2: at the scatter
#endif
#ifndef OS_PACK_POS
#define OS_PACK_POS 0 // 1: two 16-bit tile ranks per register
#endif
#define OS_WARPS (OS_THREADS / 32)

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t *p) {
	uint32_t v;
	asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t *p, uint32_t v) {
	asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Lanes of the warp whose digit equals this lane's, from 8 ballots (one per digit bit). `match.any` gives the same
// mask in one instruction but issues at ~200 cycles per warp on sm_100 (measured: 4 of them per key made the histogram
// kernel 225 us for 10 M keys, 16 per thread a third of a onesweep pass); the ballots cost ~25 issue slots.
__device__ __forceinline__ unsigned rs_peers8(uint32_t d, bool valid) {
	unsigned peers = __ballot_sync(0xFFFFFFFFu, valid);
#pragma unroll
	for (int b = 0; b < 8; b++) {
		const bool bit = (d >> b) & 1u;
		const unsigned m = __ballot_sync(0xFFFFFFFFu, bit);
		peers &= bit ? m : ~m;
	}
	return peers;
}

// ---- histogram of every pass in one read of the keys --------------------------------------------------------
template <typename KeyT, int MAX_PASSES>
__global__ void __launch_bounds__(RS_THREADS) k_radix_hist(const KeyT *__restrict__ keys, uint64_t n, int begin_bit,
                                                           int end_bit, int passes, uint32_t *__restrict__ hist) {
	// 4 replicas per pass (replica = lane & 3): a digit that is constant over the warp (the top bytes of float keys)
	// costs an 8-way instead of a 32-way same-address conflict, uniform digits pay nothing extra
	__shared__ uint32_t s_hist[MAX_PASSES * 4 * RS_BINS];
	for (int i = threadIdx.x; i < passes * 4 * RS_BINS; i += RS_THREADS) {
		s_hist[i] = 0;
	}
	__syncthreads();
	const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
	uint32_t *mine = s_hist + (threadIdx.x & 3u) * RS_BINS;
	for (uint64_t i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride) {
		const KeyT k = keys[i];
#pragma unroll
		for (int p = 0; p < MAX_PASSES; p++) {
			if (p < passes) {
				const int shift = begin_bit + 8 * p;
				const int bits = min(8, end_bit - shift);
				const uint32_t d = (uint32_t)((k >> shift) & (KeyT)((1u << bits) - 1));
				atomicAdd(&mine[p * 4 * RS_BINS + d], 1u);
			}
		}
	}
	__syncthreads();
	for (int i = threadIdx.x; i < passes * RS_BINS; i += RS_THREADS) {
		const int p = i / RS_BINS, d = i % RS_BINS;
		const uint32_t *h = s_hist + p * 4 * RS_BINS + d;
		const uint32_t c = h[0] + h[RS_BINS] + h[2 * RS_BINS] + h[3 * RS_BINS];
		if (c) {
			atomicAdd(&hist[i], c);
		}
	}
}

// exclusive scan of each pass's 256 bins, in place. One block per pass.
__global__ void __launch_bounds__(RS_BINS) k_radix_scan(uint32_t *__restrict__ hist) {
	__shared__ uint32_t s_warp[RS_BINS / 32];
	uint32_t *h = hist + blockIdx.x * RS_BINS;
	const uint32_t v = h[threadIdx.x];
	uint32_t inc = v;
#pragma unroll
	for (int s = 1; s < 32; s <<= 1) {
		const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
		if (sb_lane() >= (unsigned)s) {
			inc += o;
		}
	}
	if (sb_lane() == 31) {
		s_warp[threadIdx.x >> 5] = inc;
	}
	__syncthreads();
	uint32_t base = 0;
	for (int w = 0; w < (int)(threadIdx.x >> 5); w++) {
		base += s_warp[w];
	}
	h[threadIdx.x] = base + inc - v;
}

// segmented variant: one CTA per onesweep tile (tiles do not straddle segments), hist[(pass * nseg + seg) * 256 + digit]
template <typename KeyT, int MAX_PASSES, int TILE>
__global__ void __launch_bounds__(RS_THREADS) k_radix_hist_seg(const KeyT *__restrict__ keys, uint64_t n, int begin_bit,
                                                               int end_bit, int passes, uint64_t seg_size,
                                                               uint32_t tiles_per_seg, uint32_t nseg,
                                                               uint32_t *__restrict__ hist) {
	__shared__ uint32_t s_hist[MAX_PASSES * 4 * RS_BINS];
	for (int i = threadIdx.x; i < passes * 4 * RS_BINS; i += RS_THREADS) {
		s_hist[i] = 0;
	}
	__syncthreads();
	const uint32_t seg = blockIdx.x / tiles_per_seg;
	const uint64_t beg = (uint64_t)seg * seg_size + (uint64_t)(blockIdx.x % tiles_per_seg) * TILE;
	const uint64_t end = min(min(n, (uint64_t)(seg + 1) * seg_size), beg + TILE);
	uint32_t *mine = s_hist + (threadIdx.x & 3u) * RS_BINS;
	for (uint64_t i = beg + threadIdx.x; i < end; i += RS_THREADS) {
		const KeyT k = keys[i];
#pragma unroll
		for (int p = 0; p < MAX_PASSES; p++) {
			if (p < passes) {
				const int shift = begin_bit + 8 * p;
				const int bits = min(8, end_bit - shift);
				const uint32_t d = (uint32_t)((k >> shift) & (KeyT)((1u << bits) - 1));
				atomicAdd(&mine[p * 4 * RS_BINS + d], 1u);
			}
		}
	}
	__syncthreads();
	for (int i = threadIdx.x; i < passes * RS_BINS; i += RS_THREADS) {
		const int p = i / RS_BINS, d = i % RS_BINS;
		const uint32_t *h = s_hist + p * 4 * RS_BINS + d;
		const uint32_t c = h[0] + h[RS_BINS] + h[2 * RS_BINS] + h[3 * RS_BINS];
		if (c) {
			atomicAdd(&hist[((size_t)p * nseg + seg) * RS_BINS + d], c);
		}
	}
}

// exclusive scan of the 256 bins of every (pass, segment) + the segment's first position. One block per row.
__global__ void __launch_bounds__(RS_BINS) k_radix_scan_seg(uint32_t *__restrict__ hist, uint32_t nseg, uint64_t seg_size) {
	__shared__ uint32_t s_warp[RS_BINS / 32];
	uint32_t *h = hist + (size_t)blockIdx.x * RS_BINS;
	const uint32_t seg = blockIdx.x % nseg;
	const uint32_t v = h[threadIdx.x];
	uint32_t inc = v;
#pragma unroll
	for (int s = 1; s < 32; s <<= 1) {
		const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
		if (sb_lane() >= (unsigned)s) {
			inc += o;
		}
	}
	if (sb_lane() == 31) {
		s_warp[threadIdx.x >> 5] = inc;
	}
	__syncthreads();
	uint32_t base = (uint32_t)((uint64_t)seg * seg_size);
	for (int w = 0; w < (int)(threadIdx.x >> 5); w++) {
		base += s_warp[w];
	}
	h[threadIdx.x] = base + inc - v;
}

// ---- one onesweep pass ---------------------------------------------------------------------------------------
// SEG: segmented sort -- the input is a sequence of `seg_size`-item segments (the last one may be shorter) and items
// never leave their segment: tiles do not straddle segments (`tiles_per_seg` tiles each), the first tile of a segment
// starts a new look-back chain, and bin_base holds one row of 256 bases per segment (segment start + digit prefix).
template <typename KeyT, int ITEMS, bool SEG>
__global__ void __launch_bounds__(OS_THREADS, OS_MIN_CTAS)
    k_onesweep(const KeyT *__restrict__ keys_in, KeyT *__restrict__ keys_out, const uint32_t *__restrict__ vals_in,
               uint32_t *__restrict__ vals_out, uint64_t n, int shift, uint32_t digit_mask,
               const uint32_t *__restrict__ bin_base, uint32_t *__restrict__ tile_state, uint32_t *__restrict__ ticket,
               uint64_t seg_size, uint32_t tiles_per_seg) {
	constexpr int TILE = OS_THREADS * ITEMS;
	extern __shared__ __align__(16) unsigned char smem_raw[];
	KeyT *s_keys = reinterpret_cast<KeyT *>(smem_raw);
	uint32_t *s_vals = reinterpret_cast<uint32_t *>(smem_raw + sizeof(KeyT) * TILE);
	uint32_t *s_warp_hist = s_vals + TILE;          // [OS_WARPS][RS_BINS]
	uint32_t *s_digit_off = s_warp_hist + OS_WARPS * RS_BINS; // [RS_BINS] tile-local exclusive scan over digits
	uint32_t *s_global = s_digit_off + RS_BINS;     // [RS_BINS] global base - digit_off
	uint32_t *s_scan = s_global + RS_BINS;          // [RS_BINS / 32]
	__shared__ uint32_t s_tile;

	if (threadIdx.x == 0) {
		s_tile = atomicAdd(ticket, 1u); // tiles are claimed in order => a tile's predecessors are always running
	}
	for (int i = threadIdx.x; i < OS_WARPS * RS_BINS; i += OS_THREADS) {
		s_warp_hist[i] = 0;
	}
	__syncthreads();
	const uint32_t tile = s_tile;
	const uint32_t seg = SEG ? tile / tiles_per_seg : 0u;
	const uint32_t chain_first = SEG ? seg * tiles_per_seg : 0u; // first tile of this tile's look-back chain
	const uint64_t tile_base = SEG ? (uint64_t)seg * seg_size + (uint64_t)(tile - chain_first) * TILE : (uint64_t)tile * TILE;
	if (SEG) {
		n = min(n, (uint64_t)(seg + 1) * seg_size); // end of the segment: everything below is bounded by it
		bin_base += (size_t)seg * RS_BINS;
	}
	const unsigned warp = threadIdx.x >> 5, lane = sb_lane();
	const unsigned lt_mask = sb_lanemask_lt();
	const uint32_t tile_count = tile_base < n ? (uint32_t)min((uint64_t)TILE, n - tile_base) : 0u;

	KeyT key[ITEMS];
	uint32_t val[ITEMS];
#if OS_PACK_POS
	uint32_t pos2[(ITEMS + 1) / 2]; // tile ranks < TILE <= 65536: two per register
#else
	uint32_t pos[ITEMS];
#endif
	const uint64_t first = tile_base + (uint64_t)warp * (32 * ITEMS) + lane;
#pragma unroll
	for (int i = 0; i < ITEMS; i++) {
		const uint64_t gi = first + (uint64_t)i * 32;
		const bool valid = gi < n;
		key[i] = valid ? keys_in[gi] : KeyT(0);
#if !OS_LATE_VALS
		val[i] = valid ? (vals_in ? vals_in[gi] : (uint32_t)gi) : 0u;
#endif
	}
	// warp multi-split: stable rank of every key among the warp's keys with the same digit
	uint32_t *my_hist = s_warp_hist + warp * RS_BINS;
#pragma unroll
	for (int i = 0; i < ITEMS; i++) {
		const bool valid = first + (uint64_t)i * 32 < n;
		const uint32_t d = valid ? ((uint32_t)(key[i] >> shift) & digit_mask) : 0u;
		const unsigned peers = rs_peers8(d, valid);
		const int leader = __ffs(peers) - 1;
		uint32_t prev = 0;
		if ((int)lane == leader && valid) {
			prev = my_hist[d];
			my_hist[d] = prev + __popc(peers);
		}
		prev = __shfl_sync(0xFFFFFFFFu, prev, leader);
#if OS_PACK_POS
		if (i & 1) {
			pos2[i >> 1] |= (prev + __popc(peers & lt_mask)) << 16;
		} else {
			pos2[i >> 1] = prev + __popc(peers & lt_mask);
		}
#else
		pos[i] = prev + __popc(peers & lt_mask);
#endif
		__syncwarp();
	}
#if OS_LATE_VALS == 1
#pragma unroll
	for (int i = 0; i < ITEMS; i++) {
		const uint64_t gi = first + (uint64_t)i * 32;
		val[i] = gi < n ? (vals_in ? vals_in[gi] : (uint32_t)gi) : 0u;
	}
#endif
	__syncthreads();

	// thread d < 256 owns digit d: exclusive offsets of the warps, tile total, look-back
	const unsigned d = threadIdx.x & (RS_BINS - 1);
	const bool owner = OS_THREADS == RS_BINS || threadIdx.x < RS_BINS;
	uint32_t total = 0;
	uint32_t *my_state = tile_state + (uint64_t)tile * RS_BINS + d;
	if (owner) {
#pragma unroll
		for (int w = 0; w < OS_WARPS; w++) {
			const uint32_t c = s_warp_hist[w * RS_BINS + d];
			s_warp_hist[w * RS_BINS + d] = total;
			total += c;
		}
		st_volatile_u32(my_state, (tile == chain_first ? RS_FLAG_INC : RS_FLAG_AGG) | total);
	}

	// block exclusive scan of `total` over the 256 digits
	uint32_t inc = total;
#pragma unroll
	for (int s = 1; s < 32; s <<= 1) {
		const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
		if (lane >= (unsigned)s) {
			inc += o;
		}
	}
	if (lane == 31 && owner) {
		s_scan[warp] = inc;
	}
	__syncthreads();
	uint32_t dbase = 0;
	for (unsigned w = 0; w < (warp & (RS_BINS / 32 - 1)); w++) {
		dbase += s_scan[w];
	}
	const uint32_t digit_off = dbase + inc - total;
	if (owner) {
		s_digit_off[d] = digit_off;
	}

	uint32_t excl = 0;
	if (owner && tile > chain_first) {
		// look back over the predecessors' status words, RS_LOOK loads in flight at a time (a one-at-a-time walk made the
		// first wave of tiles pay one L2 round trip per predecessor: the whole pass was latency bound at 10 M keys)
		int64_t look = (int64_t)tile - 1;
		bool done = false;
		while (!done) {
			uint32_t st[RS_LOOK];
#pragma unroll
			for (int j = 0; j < RS_LOOK; j++) {
				st[j] = look - j >= (int64_t)chain_first ? ld_volatile_u32(tile_state + (uint64_t)(look - j) * RS_BINS + d)
				                                         : RS_FLAG_INC;
			}
#pragma unroll
			for (int j = 0; j < RS_LOOK; j++) {
				if (!done) {
					if ((st[j] & (RS_FLAG_AGG | RS_FLAG_INC)) == 0) {
						break; // this predecessor has not published yet: poll again from here
					}
					excl += st[j] & RS_VALUE_MASK;
					look--;
					done = (st[j] & RS_FLAG_INC) != 0;
				}
			}
		}
		st_volatile_u32(my_state, RS_FLAG_INC | (excl + total));
	}
	if (owner) {
		s_global[d] = bin_base[d] + excl - digit_off;
	}
	__syncthreads();

	// scatter into shared memory in tile-sorted order
#pragma unroll
	for (int i = 0; i < ITEMS; i++) {
		if (first + (uint64_t)i * 32 < n) {
			const uint32_t dg = (uint32_t)(key[i] >> shift) & digit_mask;
#if OS_PACK_POS
			const uint32_t rank = (i & 1) ? pos2[i >> 1] >> 16 : pos2[i >> 1] & 0xFFFFu;
#else
			const uint32_t rank = pos[i];
#endif
			const uint32_t p = s_digit_off[dg] + s_warp_hist[warp * RS_BINS + dg] + rank;
			s_keys[p] = key[i];
#if OS_LATE_VALS == 2
			const uint64_t gi = first + (uint64_t)i * 32;
			s_vals[p] = vals_in ? vals_in[gi] : (uint32_t)gi;
#else
			s_vals[p] = val[i];
#endif
		}
	}
	__syncthreads();
	// coalesced runs out to global memory
#pragma unroll
	for (int i = 0; i < ITEMS; i++) {
		const uint32_t p = threadIdx.x + i * OS_THREADS;
		if (p < tile_count) {
			const KeyT k = s_keys[p];
			const uint32_t dg = (uint32_t)(k >> shift) & digit_mask;
			const uint32_t g = s_global[dg] + p;
			keys_out[g] = k;
			vals_out[g] = s_vals[p];
		}
	}
}

template <typename KeyT>
struct rs_cfg {
	static constexpr int ITEMS = sizeof(KeyT) == 4 ? OS_ITEMS32 : OS_ITEMS64;
	static constexpr int TILE = OS_THREADS * ITEMS;
	static constexpr size_t SMEM =
	    (sizeof(KeyT) + 4) * TILE + sizeof(uint32_t) * (OS_WARPS * RS_BINS + 2 * RS_BINS + RS_BINS / 32);
	static constexpr int MAX_PASSES = sizeof(KeyT);
};

template <typename KeyT>
int sbi_radix_sort_impl(sb_context *ctx, KeyT *keys, KeyT *keys_alt, uint32_t *vals, uint32_t *vals_alt, uint64_t n,
                        int begin_bit, int end_bit, bool vals_iota, int *result_in_alt) {
	using cfg = rs_cfg<KeyT>;
	*result_in_alt = 0;
	if (n == 0) {
		return SB_OK;
	}
	SB_ARG(n < ((uint64_t)1 << 30), "radix sort: n must be < 2^30");
	SB_ARG(begin_bit >= 0 && end_bit <= (int)sizeof(KeyT) * 8 && begin_bit <= end_bit, "radix sort bit range");
	const int passes = (end_bit - begin_bit + 7) / 8;
	if (passes == 0) {
		if (vals_iota) { // still materialise the identity permutation
			SB_TRY(sbi_iota_u32(ctx, vals, n));
		}
		return SB_OK;
	}
	const uint64_t tiles = (n + cfg::TILE - 1) / cfg::TILE;
	// temp: [passes*256 hist][passes tickets (padded to 64)][passes * tiles * 256 status words]
	const size_t words = (size_t)passes * RS_BINS + 64 + (size_t)passes * tiles * RS_BINS;
	uint32_t *temp = nullptr;
	SB_TRY(ctx->scratch_t<uint32_t>(words, &temp));
	SB_CUDA(cudaMemsetAsync(temp, 0, words * sizeof(uint32_t), ctx->stream));
	uint32_t *hist = temp;
	uint32_t *tickets = temp + (size_t)passes * RS_BINS;
	uint32_t *state = tickets + 64;

	SB_SMEM_OPTIN(ctx, (k_onesweep<KeyT, cfg::ITEMS, false>), cfg::SMEM);
	{
		sb_prof_scope ps(ctx, "radix_hist");
		k_radix_hist<KeyT, cfg::MAX_PASSES><<<sb_grid(n, RS_THREADS * 8, 8), RS_THREADS, 0, ctx->stream>>>(
		    keys, n, begin_bit, end_bit, passes, hist);
		SB_LAUNCH_CHECK(ctx);
		k_radix_scan<<<passes, RS_BINS, 0, ctx->stream>>>(hist);
		SB_LAUNCH_CHECK(ctx);
	}
	KeyT *kin = keys, *kout = keys_alt;
	uint32_t *vin = vals, *vout = vals_alt;
	for (int p = 0; p < passes; p++) {
		const int shift = begin_bit + 8 * p;
		const int bits = end_bit - shift < 8 ? end_bit - shift : 8;
		sb_prof_scope ps(ctx, "radix_onesweep");
		k_onesweep<KeyT, cfg::ITEMS, false><<<(unsigned)tiles, OS_THREADS, cfg::SMEM, ctx->stream>>>(
		    kin, kout, (p == 0 && vals_iota) ? nullptr : vin, vout, n, shift, (1u << bits) - 1, hist + p * RS_BINS,
		    state + (size_t)p * tiles * RS_BINS, tickets + p, 0, 0);
		SB_LAUNCH_CHECK(ctx);
		KeyT *tk = kin;
		kin = kout;
		kout = tk;
		uint32_t *tv = vin;
		vin = vout;
		vout = tv;
	}
	*result_in_alt = (passes & 1);
	return SB_OK;
}

// Segmented sort of 32-bit keys: every `seg_size`-item segment of the input (the last one may be shorter) is sorted on
// its own, stably, in (end_bit - begin_bit) / 8 passes -- the segment index costs no pass (the STR y-sort inside
// vertical slices was a 64-bit (slice, y) sort of 6 passes before; it is 4 passes over 32-bit keys now).
int sbi_radix_sort_u32_segmented(sb_context *ctx, uint32_t *keys, uint32_t *keys_alt, uint32_t *vals, uint32_t *vals_alt,
                                 uint64_t n, uint64_t seg_size, int begin_bit, int end_bit, bool vals_iota,
                                 int *result_in_alt) {
	using cfg = rs_cfg<uint32_t>;
	*result_in_alt = 0;
	if (n == 0) {
		return SB_OK;
	}
	SB_ARG(n < ((uint64_t)1 << 30), "radix sort: n must be < 2^30");
	SB_ARG(seg_size >= 1, "segment size");
	SB_ARG(begin_bit >= 0 && end_bit <= 32 && begin_bit <= end_bit, "radix sort bit range");
	const int passes = (end_bit - begin_bit + 7) / 8;
	if (passes == 0) {
		if (vals_iota) {
			SB_TRY(sbi_iota_u32(ctx, vals, n));
		}
		return SB_OK;
	}
	const uint64_t nseg = (n + seg_size - 1) / seg_size;
	const uint64_t tps = (seg_size + cfg::TILE - 1) / cfg::TILE;
	const uint64_t tiles = nseg * tps;
	SB_ARG(tiles < 0x7FFFFFFFull && nseg < 0x7FFFFFFFull, "segmented sort: too many segments");
	const size_t hist_words = (size_t)passes * nseg * RS_BINS;
	const size_t words = hist_words + 64 + (size_t)passes * tiles * RS_BINS;
	uint32_t *temp = nullptr;
	SB_TRY(ctx->scratch_t<uint32_t>(words, &temp));
	SB_CUDA(cudaMemsetAsync(temp, 0, words * sizeof(uint32_t), ctx->stream));
	uint32_t *hist = temp;
	uint32_t *tickets = temp + hist_words;
	uint32_t *state = tickets + 64;
	SB_SMEM_OPTIN(ctx, (k_onesweep<uint32_t, cfg::ITEMS, true>), cfg::SMEM);
	{
		sb_prof_scope ps(ctx, "radix_hist");
		k_radix_hist_seg<uint32_t, cfg::MAX_PASSES, cfg::TILE><<<(unsigned)tiles, RS_THREADS, 0, ctx->stream>>>(
		    keys, n, begin_bit, end_bit, passes, seg_size, (uint32_t)tps, (uint32_t)nseg, hist);
		SB_LAUNCH_CHECK(ctx);
		k_radix_scan_seg<<<(unsigned)(passes * nseg), RS_BINS, 0, ctx->stream>>>(hist, (uint32_t)nseg, seg_size);
		SB_LAUNCH_CHECK(ctx);
	}
	uint32_t *kin = keys, *kout = keys_alt;
	uint32_t *vin = vals, *vout = vals_alt;
	for (int p = 0; p < passes; p++) {
		const int shift = begin_bit + 8 * p;
		const int bits = end_bit - shift < 8 ? end_bit - shift : 8;
		sb_prof_scope ps(ctx, "radix_onesweep");
		k_onesweep<uint32_t, cfg::ITEMS, true><<<(unsigned)tiles, OS_THREADS, cfg::SMEM, ctx->stream>>>(
		    kin, kout, (p == 0 && vals_iota) ? nullptr : vin, vout, n, shift, (1u << bits) - 1,
		    hist + (size_t)p * nseg * RS_BINS, state + (size_t)p * tiles * RS_BINS, tickets + p, seg_size, (uint32_t)tps);
		SB_LAUNCH_CHECK(ctx);
		uint32_t *tk = kin;
		kin = kout;
		kout = tk;
		uint32_t *tv = vin;
		vin = vout;
		vout = tv;
	}
	*result_in_alt = (passes & 1);
	return SB_OK;
}

int sbi_radix_sort_u32(sb_context *ctx, uint32_t *keys, uint32_t *keys_alt, uint32_t *vals, uint32_t *vals_alt,
                       uint64_t n, int begin_bit, int end_bit, bool vals_iota, int *result_in_alt) {
	return sbi_radix_sort_impl<uint32_t>(ctx, keys, keys_alt, vals, vals_alt, n, begin_bit, end_bit, vals_iota,
	                                     result_in_alt);
}

int sbi_radix_sort_u64(sb_context *ctx, uint64_t *keys, uint64_t *keys_alt, uint32_t *vals, uint32_t *vals_alt,
                       uint64_t n, int begin_bit, int end_bit, bool vals_iota, int *result_in_alt) {
	return sbi_radix_sort_impl<uint64_t>(ctx, keys, keys_alt, vals, vals_alt, n, begin_bit, end_bit, vals_iota,
	                                     result_in_alt);
}

__global__ void __launch_bounds__(256) k_iota_u32(uint32_t *out, uint64_t n) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		out[i] = (uint32_t)i;
	}
}

int sbi_iota_u32(sb_context *ctx, uint32_t *out, uint64_t n) {
	if (n == 0) {
		return SB_OK;
	}
	k_iota_u32<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(out, n);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}
