// sort.cu — BATsort (gdk_batop.c:2342-2810) and the plain form of BATfirstn (gdk_firstn.c:1280) on the device.
//
// The reference sorts with GDKqsort / GDKssort (timsort) / GDKrsort (gdk_batop.c:2265-2303); none of those map to
// a GPU.  Here every sort is ONE stable least-significant-digit radix sort over an order-preserving unsigned
// image of the tail value:
//   integers (and date / daytime / timestamp / oid)   v ^ sign bit, nil (the type minimum) -> 0;
//   flt / dbl                                         IEEE total order image, -0 folded onto +0, NaN (nil) -> 0;
//   descending                                        the complement of the image;
//   nils last when ascending / first when descending  one more stable pass over a nil flag, more significant
//                                                     than the value (inside each run of g).
// A stable sort satisfies both the `stable` and the `!stable` contract (the latter leaves the order of equal
// values open).  BATsort's refinement arguments are part of the key: `o` first rearranges b, and each run of
// equal consecutive values of `g` is sorted on its own — the run number (a prefix count of run starts) is
// sorted as the most significant key, and because run numbers already ascend with the position the stable
// passes leave the runs where they are.  Digit passes whose digit is the same for every row (the high bytes
// of small-range columns, most of the run-number bytes) are found by one histogram over all digits and skipped.
//
// Per pass: (1) per-tile digit histogram, (2) one scan over (digit, tile), (3) stable scatter — inside a tile
// rows are ranked warp by warp with __match_any_sync, so equal digits keep their order.
// Outputs follow the reference: `sorted` = the values in order, `order` = their oids (b's hseqbase based, or
// o's values), `groups` = dense ids of equal (run, value) combinations in sorted order, usable as o / g of the
// next BATsort call of a multi-column ORDER BY.
#include "common.cuh"
#include "keys.cuh"
#include <type_traits>

#define RS_THREADS 256
#define RS_ITEMS 8
#define RS_TILE (RS_THREADS * RS_ITEMS)
#define RS_WARPS (RS_THREADS / 32)

// ---------------------------------------------------------------- key images
template <int W> struct UKey;
template <> struct UKey<1> { typedef uint8_t type; };
template <> struct UKey<2> { typedef uint16_t type; };
template <> struct UKey<4> { typedef uint32_t type; };
template <> struct UKey<8> { typedef u64 type; };
template <> struct UKey<16> { typedef uhge type; };

template <typename T> __device__ __forceinline__ typename UKey<sizeof(T)>::type sort_image(T v, bool reverse, bool nilslast)
{
	typedef typename UKey<sizeof(T)>::type U;
	const U top = (U) 1 << (sizeof(T) * 8 - 1);
	U img;
	bool nil;
	if constexpr (Atom<T>::is_float) {
		nil = v != v;
		U bits;
		if (v == 0)
			v = 0;		/* -0 == +0 */
		memcpy(&bits, &v, sizeof(T));
		img = (bits & top) ? (U) ~bits : (U) (bits | top);
	} else {
		nil = Atom<T>::is_nil(v);
		img = (U) v ^ top;
	}
	(void) nilslast;
	if (nil)
		img = 0;	/* nil is the smallest value; no other value has image 0 (nil is the type minimum / NaN) */
	return reverse ? (U) ~img : img;
}

// build the key image column (and the initial position column) of the rearranged input
template <typename T>
__global__ void __launch_bounds__(256)
sort_keys_kernel(const T *__restrict__ vals, const oid *__restrict__ o, oid hseq, uint64_t n, bool reverse, bool nilslast,
		 typename UKey<sizeof(T)>::type *__restrict__ keys, uint32_t *__restrict__ pos)
{
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
		const T v = o ? vals[o[i] - hseq] : vals[i];
		keys[i] = sort_image<T>(v, reverse, nilslast);
		pos[i] = (uint32_t) i;
	}
}

// run numbers of g: prefix count of positions where g changes (two-level scan: per-tile counts, then offsets)
__global__ void __launch_bounds__(RS_THREADS)
run_count_kernel(const oid *__restrict__ g, uint64_t n, uint32_t *__restrict__ tilecnt)
{
	__shared__ unsigned tot;
	if (threadIdx.x == 0)
		tot = 0;
	__syncthreads();
	const uint64_t base = (uint64_t) blockIdx.x * RS_TILE;
	unsigned c = 0;
	for (unsigned j = threadIdx.x; j < RS_TILE; j += RS_THREADS) {
		const uint64_t i = base + j;
		if (i < n && i > 0 && g[i] != g[i - 1])
			c++;
	}
	for (int s = 16; s > 0; s >>= 1)
		c += __shfl_down_sync(0xffffffffu, c, s);
	if ((threadIdx.x & 31) == 0 && c)
		atomicAdd(&tot, c);
	__syncthreads();
	if (threadIdx.x == 0)
		tilecnt[blockIdx.x] = tot;
}

// exclusive scan of up to 2^32 32-bit counts by one CTA (in place), total -> *total
__global__ void __launch_bounds__(1024)
scan32_kernel(uint32_t *__restrict__ a, uint64_t n, unsigned long long *__restrict__ total)
{
	__shared__ unsigned long long wsum[32];
	__shared__ unsigned long long carry_s;
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x == 0)
		carry_s = 0;
	__syncthreads();
	for (uint64_t base = 0; base < n; base += 1024) {
		const uint64_t i = base + threadIdx.x;
		const unsigned long long c = i < n ? a[i] : 0;
		unsigned long long inc = c;
		for (int s = 1; s < 32; s <<= 1) {
			const unsigned long long t = __shfl_up_sync(0xffffffffu, inc, s);
			if (lane >= s)
				inc += t;
		}
		if (lane == 31)
			wsum[warp] = inc;
		__syncthreads();
		if (warp == 0) {
			const unsigned long long w = wsum[lane];
			unsigned long long winc = w;
			for (int s = 1; s < 32; s <<= 1) {
				const unsigned long long t = __shfl_up_sync(0xffffffffu, winc, s);
				if (lane >= s)
					winc += t;
			}
			wsum[lane] = winc - w;
		}
		__syncthreads();
		const unsigned long long excl = carry_s + wsum[warp] + inc - c;
		if (i < n)
			a[i] = (uint32_t) excl;
		__syncthreads();
		if (threadIdx.x == 1023)
			carry_s = excl + c;
		__syncthreads();
	}
	if (threadIdx.x == 0 && total)
		*total = carry_s;
}

__global__ void __launch_bounds__(RS_THREADS)
run_assign_kernel(const oid *__restrict__ g, uint64_t n, const uint32_t *__restrict__ tileoff, uint32_t *__restrict__ runs)
{
	__shared__ unsigned wtot[RS_WARPS];
	const uint64_t base = (uint64_t) blockIdx.x * RS_TILE;
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	/* thread t owns RS_ITEMS consecutive rows */
	const uint64_t i0 = base + (uint64_t) threadIdx.x * RS_ITEMS;
	unsigned f[RS_ITEMS], mine = 0;
#pragma unroll
	for (int j = 0; j < RS_ITEMS; j++) {
		const uint64_t i = i0 + j;
		f[j] = (i < n && i > 0 && g[i] != g[i - 1]) ? 1u : 0u;
		mine += f[j];
	}
	unsigned inc = mine;
	for (int s = 1; s < 32; s <<= 1) {
		const unsigned t = __shfl_up_sync(0xffffffffu, inc, s);
		if (lane >= s)
			inc += t;
	}
	if (lane == 31)
		wtot[warp] = inc;
	__syncthreads();
	unsigned run = tileoff[blockIdx.x] + inc - mine;
	for (unsigned w = 0; w < warp; w++)
		run += wtot[w];
#pragma unroll
	for (int j = 0; j < RS_ITEMS; j++) {
		const uint64_t i = i0 + j;
		run += f[j];
		if (i < n)
			runs[i] = run;
	}
}

// ---------------------------------------------------------------- radix passes
// digit occupancy of every byte of the key in one read: hist[byte][digit]
template <typename U>
__global__ void __launch_bounds__(256)
digit_survey_kernel(const U *__restrict__ keys, uint64_t n, unsigned long long *__restrict__ hist)
{
	__shared__ unsigned sh[sizeof(U)][256];
	for (unsigned i = threadIdx.x; i < sizeof(U) * 256; i += blockDim.x)
		(&sh[0][0])[i] = 0;
	__syncthreads();
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
		const U k = keys[i];
#pragma unroll
		for (int b = 0; b < (int) sizeof(U); b++)
			atomicAdd(&sh[b][(unsigned) (k >> (8 * b)) & 255u], 1u);
	}
	__syncthreads();
	for (unsigned i = threadIdx.x; i < sizeof(U) * 256; i += blockDim.x) {
		const unsigned v = (&sh[0][0])[i];
		if (v)
			atomicAdd(&hist[i], (unsigned long long) v);
	}
}

template <typename U>
__global__ void __launch_bounds__(RS_THREADS)
radix_hist_kernel(const U *__restrict__ keys, uint64_t n, int shift, uint32_t *__restrict__ hist /* [256][ntiles] */, uint64_t ntiles)
{
	__shared__ unsigned sh[256];
	sh[threadIdx.x] = 0;
	__syncthreads();
	const uint64_t base = (uint64_t) blockIdx.x * RS_TILE;
#pragma unroll
	for (int j = 0; j < RS_ITEMS; j++) {
		const uint64_t i = base + (uint64_t) j * RS_THREADS + threadIdx.x;
		if (i < n)
			atomicAdd(&sh[(unsigned) (keys[i] >> shift) & 255u], 1u);
	}
	__syncthreads();
	hist[(uint64_t) threadIdx.x * ntiles + blockIdx.x] = sh[threadIdx.x];
}

// stable scatter of one tile: warp w owns rows [w*256, w*256+256) of the tile as 8 rows of 32 lanes; lanes with
// the same digit are ranked by __match_any_sync in lane order, rows in row order, warps through a per-digit
// scan of the warp totals
template <typename U, bool HAS_AUX>
__global__ void __launch_bounds__(RS_THREADS)
radix_scatter_kernel(const U *__restrict__ keys, const uint32_t *__restrict__ pos, const uint32_t *__restrict__ aux, uint64_t n,
		     int shift, const uint32_t *__restrict__ offs /* scanned [256][ntiles] */, uint64_t ntiles, U *__restrict__ keys_out,
		     uint32_t *__restrict__ pos_out, uint32_t *__restrict__ aux_out)
{
	__shared__ unsigned wcnt[RS_WARPS][256];
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	for (unsigned i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS)
		(&wcnt[0][0])[i] = 0;
	__syncthreads();
	const uint64_t wbase = (uint64_t) blockIdx.x * RS_TILE + (uint64_t) warp * (RS_TILE / RS_WARPS);
	U k[RS_ITEMS];
	unsigned rank[RS_ITEMS], dig[RS_ITEMS];
	bool in[RS_ITEMS];
#pragma unroll
	for (int j = 0; j < RS_ITEMS; j++) {
		const uint64_t i = wbase + (uint64_t) j * 32 + lane;
		in[j] = i < n;
		k[j] = in[j] ? keys[i] : (U) 0;
		dig[j] = in[j] ? (unsigned) (k[j] >> shift) & 255u : 256u + lane;	/* out-of-range lanes match nobody */
		const unsigned m = __match_any_sync(0xffffffffu, dig[j]);
		const unsigned before = __popc(m & ((1u << lane) - 1));
		unsigned base = 0;
		if (in[j]) {
			if (before == 0)
				base = wcnt[warp][dig[j]];
		}
		base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
		rank[j] = base + before;
		if (in[j] && before == 0)
			wcnt[warp][dig[j]] = base + __popc(m);
		__syncwarp();
	}
	__syncthreads();
	/* per digit: exclusive scan over the warps, plus the tile's global offset */
	{
		const unsigned d = threadIdx.x;		/* RS_THREADS == 256 digits */
		unsigned run = offs[(uint64_t) d * ntiles + blockIdx.x];
#pragma unroll
		for (int w = 0; w < RS_WARPS; w++) {
			const unsigned c = wcnt[w][d];
			wcnt[w][d] = run;
			run += c;
		}
	}
	__syncthreads();
#pragma unroll
	for (int j = 0; j < RS_ITEMS; j++) {
		if (!in[j])
			continue;
		const uint64_t i = wbase + (uint64_t) j * 32 + lane;
		const unsigned dst = wcnt[warp][dig[j]] + rank[j];
		keys_out[dst] = k[j];
		pos_out[dst] = pos[i];
		if (HAS_AUX)
			aux_out[dst] = aux[i];
	}
}

// one LSD pass over the 8-bit digit at `shift` of `keys` (any unsigned width), carrying pos (and aux)
struct SortBufs {
	void *keys[2] = {nullptr, nullptr};
	uint32_t *pos[2] = {nullptr, nullptr};
	uint32_t *aux[2] = {nullptr, nullptr};	/* the run numbers while the value key is being sorted, or the value-sorted keys' companion */
	uint32_t *hist = nullptr;
	uint64_t ntiles = 0;
	int cur = 0;
};

template <typename U, bool HAS_AUX> static bool radix_pass(SortBufs &sb, uint64_t n, int shift, const U *keys_in, U *keys_out)
{
	Ctx *c = ctx();
	const uint64_t nt = sb.ntiles;
	{
		KTimer kt("sort_pass", n * (2 * sizeof(U) + 8 + sizeof(U)));
		radix_hist_kernel<U><<<(unsigned) nt, RS_THREADS, 0, c->stream>>>(keys_in, n, shift, sb.hist, nt);
		scan32_kernel<<<1, 1024, 0, c->stream>>>(sb.hist, 256 * nt, nullptr);
		radix_scatter_kernel<U, HAS_AUX><<<(unsigned) nt, RS_THREADS, 0, c->stream>>>(
			keys_in, sb.pos[sb.cur], HAS_AUX ? sb.aux[sb.cur] : nullptr, n, shift, sb.hist, nt, keys_out, sb.pos[sb.cur ^ 1],
			HAS_AUX ? sb.aux[sb.cur ^ 1] : nullptr);
	}
	COUNT_LAUNCH();
	COUNT_LAUNCH();
	COUNT_LAUNCH();
	sb.cur ^= 1;
	return cuda_ok(cudaGetLastError(), "radix pass");
}

// which byte positions of the key column actually differ between rows (one histogram read + one sync)
template <typename U> static bool survey_digits(const U *keys, uint64_t n, bool (&need)[16])
{
	Ctx *c = ctx();
	const size_t bytes = sizeof(U) * 256 * 8;
	unsigned long long *hist = (unsigned long long *) dev_alloc(bytes);
	if (hist == nullptr || !cuda_ok(cudaMemsetAsync(hist, 0, bytes, c->stream), "memset"))
		return false;
	digit_survey_kernel<U><<<grid_for(n, 256 * 16, NSM * 4), 256, 0, c->stream>>>(keys, n, hist);
	COUNT_LAUNCH();
	unsigned long long *h = (unsigned long long *) malloc(bytes);
	bool ok = h != nullptr && cuda_ok(cudaMemcpyAsync(h, hist, bytes, cudaMemcpyDeviceToHost, c->stream), "survey read") && stream_sync();
	if (ok)
		for (size_t b = 0; b < sizeof(U); b++) {
			need[b] = true;
			for (int d = 0; d < 256; d++)
				if (h[b * 256 + d] == n)
					need[b] = false;	/* every row has this digit: the pass would be the identity */
		}
	free(h);
	dev_free(hist);
	return ok;
}

// nil-flag digit of the row at each current position (reverse != nilslast: nils leave their natural end)
template <typename T>
__global__ void __launch_bounds__(256)
nilflag_kernel(const T *__restrict__ vals, const oid *__restrict__ o, oid hseq, const uint32_t *__restrict__ pos, uint64_t n, bool nilslast,
	       uint8_t *__restrict__ flag)
{
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
		const uint32_t p = pos[i];
		const T v = o ? vals[o[p] - hseq] : vals[p];
		const bool nil = Atom<T>::is_nil(v);
		flag[i] = (uint8_t) (nil == nilslast ? 1 : 0);
	}
}

// ---------------------------------------------------------------- outputs
template <typename T>
__global__ void __launch_bounds__(256)
sort_gather_kernel(const T *__restrict__ vals, const oid *__restrict__ o, oid hseq, const uint32_t *__restrict__ pos, uint64_t n,
		   T *__restrict__ sorted, oid *__restrict__ order)
{
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
		const uint32_t p = pos[i];
		const oid id = o ? o[p] : hseq + p;
		if (order)
			order[i] = id;
		if (sorted)
			sorted[i] = vals[id - hseq];
	}
}

// group ids: prefix count of boundaries between different (run, key image) neighbours — two-level like the runs
template <typename U>
__global__ void __launch_bounds__(RS_THREADS)
grp_count_kernel(const U *__restrict__ keys, const uint32_t *__restrict__ runs, uint64_t n, uint32_t *__restrict__ tilecnt)
{
	__shared__ unsigned tot;
	if (threadIdx.x == 0)
		tot = 0;
	__syncthreads();
	const uint64_t base = (uint64_t) blockIdx.x * RS_TILE;
	unsigned c = 0;
	for (unsigned j = threadIdx.x; j < RS_TILE; j += RS_THREADS) {
		const uint64_t i = base + j;
		if (i < n && i > 0 && (keys[i] != keys[i - 1] || (runs && runs[i] != runs[i - 1])))
			c++;
	}
	for (int s = 16; s > 0; s >>= 1)
		c += __shfl_down_sync(0xffffffffu, c, s);
	if ((threadIdx.x & 31) == 0 && c)
		atomicAdd(&tot, c);
	__syncthreads();
	if (threadIdx.x == 0)
		tilecnt[blockIdx.x] = tot;
}

template <typename U>
__global__ void __launch_bounds__(RS_THREADS)
grp_assign_ids_kernel(const U *__restrict__ keys, const uint32_t *__restrict__ runs, uint64_t n, const uint32_t *__restrict__ tileoff,
		      oid *__restrict__ gids)
{
	__shared__ unsigned wtot[RS_WARPS];
	const uint64_t base = (uint64_t) blockIdx.x * RS_TILE;
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint64_t i0 = base + (uint64_t) threadIdx.x * RS_ITEMS;
	unsigned f[RS_ITEMS], mine = 0;
#pragma unroll
	for (int j = 0; j < RS_ITEMS; j++) {
		const uint64_t i = i0 + j;
		f[j] = (i < n && i > 0 && (keys[i] != keys[i - 1] || (runs && runs[i] != runs[i - 1]))) ? 1u : 0u;
		mine += f[j];
	}
	unsigned inc = mine;
	for (int s = 1; s < 32; s <<= 1) {
		const unsigned t = __shfl_up_sync(0xffffffffu, inc, s);
		if (lane >= s)
			inc += t;
	}
	if (lane == 31)
		wtot[warp] = inc;
	__syncthreads();
	unsigned g = tileoff[blockIdx.x] + inc - mine;
	for (unsigned w = 0; w < warp; w++)
		g += wtot[w];
#pragma unroll
	for (int j = 0; j < RS_ITEMS; j++) {
		const uint64_t i = i0 + j;
		g += f[j];
		if (i < n)
			gids[i] = g;
	}
}

// positions from a cached order index (oids based on idx_hseq)
__global__ void __launch_bounds__(256)
oidx_pos_kernel(const oid *__restrict__ idx, oid idx_hseq, uint64_t n, uint32_t *__restrict__ pos)
{
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x)
		pos[i] = (uint32_t) (idx[i] - idx_hseq);
}
/* set by b200_BATsort when the column has an order index that answers the request: the digit passes are skipped */
static thread_local const oid *tl_sort_index = nullptr;
static thread_local oid tl_sort_index_hseq = 0;

// ---------------------------------------------------------------- the typed driver
struct SortOut {
	b200_bat *sorted = nullptr, *order = nullptr, *groups = nullptr;
	uint64_t ngroups = 0;
};

template <typename T>
static bool sort_typed(const b200_bat *b, const oid *o, const oid *g, uint64_t n, bool reverse, bool nilslast, bool want_sorted,
		       bool want_order, bool want_groups, oid order_hseq, SortOut &out)
{
	typedef typename UKey<sizeof(T)>::type U;
	Ctx *c = ctx();
	SortBufs sb;
	sb.ntiles = (n + RS_TILE - 1) / RS_TILE;
	bool ok = false;
	uint32_t *tile = nullptr;
	unsigned long long *tot_d = nullptr;
	bool need[16] = {false}, need_run[16] = {false};
	const T *vals = (const T *) b->base;
	const oid hseq = b->hseqbase;
	U *kout = nullptr;
	const U *kin = nullptr;
	uint32_t *runs_sorted = nullptr;
	sb.keys[0] = dev_alloc(n * sizeof(U));
	sb.keys[1] = dev_alloc(n * sizeof(U));
	sb.pos[0] = (uint32_t *) dev_alloc(n * 4);
	sb.pos[1] = (uint32_t *) dev_alloc(n * 4);
	sb.hist = (uint32_t *) dev_alloc(256 * sb.ntiles * 4);
	tile = (uint32_t *) dev_alloc((sb.ntiles + 1) * 4);
	tot_d = (unsigned long long *) dev_alloc(8);
	if (!sb.keys[0] || !sb.keys[1] || !sb.pos[0] || !sb.pos[1] || !sb.hist || !tile || !tot_d)
		goto done;
	if (g) {
		sb.aux[0] = (uint32_t *) dev_alloc(n * 4);
		sb.aux[1] = (uint32_t *) dev_alloc(n * 4);
		if (!sb.aux[0] || !sb.aux[1])
			goto done;
		run_count_kernel<<<(unsigned) sb.ntiles, RS_THREADS, 0, c->stream>>>(g, n, tile);
		scan32_kernel<<<1, 1024, 0, c->stream>>>(tile, sb.ntiles, nullptr);
		run_assign_kernel<<<(unsigned) sb.ntiles, RS_THREADS, 0, c->stream>>>(g, n, tile, sb.aux[0]);
		COUNT_LAUNCH();
		COUNT_LAUNCH();
		COUNT_LAUNCH();
	}
	if (tl_sort_index != nullptr && o == nullptr && g == nullptr && !reverse && !nilslast) {
		/* the order index IS the answer (gdk_batop.c:2489-2520): no key images, no digit passes */
		oidx_pos_kernel<<<grid_for(n, 256, NSM * 32), 256, 0, c->stream>>>(tl_sort_index, tl_sort_index_hseq, n, sb.pos[0]);
		COUNT_LAUNCH();
		goto passes_done;
	}
	sort_keys_kernel<T><<<grid_for(n, 256, NSM * 32), 256, 0, c->stream>>>(vals, o, hseq, n, reverse, nilslast, (U *) sb.keys[0], sb.pos[0]);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "sort keys") || !survey_digits<U>((const U *) sb.keys[0], n, need))
		goto done;
	if (g && !survey_digits<uint32_t>(sb.aux[0], n, need_run))
		goto done;
	/* value digits, least significant first; the run numbers ride along as aux */
	{
		int kc = 0;
		for (int byte = 0; byte < (int) sizeof(U); byte++) {
			if (!need[byte])
				continue;
			kin = (const U *) sb.keys[kc];
			kout = (U *) sb.keys[kc ^ 1];
			if (!(g ? radix_pass<U, true>(sb, n, byte * 8, kin, kout) : radix_pass<U, false>(sb, n, byte * 8, kin, kout)))
				goto done;
			kc ^= 1;
		}
		if (reverse != nilslast && !b->tnonil) {
			/* nils to the end they do not own naturally: a stable pass over the nil flag */
			uint8_t *f0 = (uint8_t *) dev_alloc(n), *f1 = (uint8_t *) dev_alloc(n);
			bool okf = f0 != nullptr && f1 != nullptr;
			if (okf) {
				nilflag_kernel<T><<<grid_for(n, 256, NSM * 32), 256, 0, c->stream>>>(vals, o, hseq, sb.pos[sb.cur], n, nilslast, f0);
				COUNT_LAUNCH();
				okf = g ? radix_pass<uint8_t, true>(sb, n, 0, f0, f1) : radix_pass<uint8_t, false>(sb, n, 0, f0, f1);
			}
			dev_free(f0);
			dev_free(f1);
			if (!okf)
				goto done;
		}
		/* then the run numbers as the key (stable: value order survives inside a run); the value images ride as
		 * aux only when they are 4 bytes wide, otherwise they are re-derived from pos for the group ids */
		if (g) {
			bool any = false;
			for (int byte = 0; byte < 4; byte++)
				any |= need_run[byte];
			if (any) {
				/* sort (run, pos) pairs by run; afterwards rebuild the value images in the final order */
				for (int byte = 0; byte < 4; byte++) {
					if (!need_run[byte])
						continue;
					/* keys = aux (runs), aux slot unused */
					Ctx *cc = ctx();
					const uint64_t nt = sb.ntiles;
					radix_hist_kernel<uint32_t><<<(unsigned) nt, RS_THREADS, 0, cc->stream>>>(sb.aux[sb.cur], n, byte * 8, sb.hist, nt);
					scan32_kernel<<<1, 1024, 0, cc->stream>>>(sb.hist, 256 * nt, nullptr);
					radix_scatter_kernel<uint32_t, false><<<(unsigned) nt, RS_THREADS, 0, cc->stream>>>(
						sb.aux[sb.cur], sb.pos[sb.cur], nullptr, n, byte * 8, sb.hist, nt, sb.aux[sb.cur ^ 1], sb.pos[sb.cur ^ 1], nullptr);
					COUNT_LAUNCH();
					COUNT_LAUNCH();
					COUNT_LAUNCH();
					sb.cur ^= 1;
				}
			}
			runs_sorted = sb.aux[sb.cur];
		}
	}
passes_done:
	if (!cuda_ok(cudaGetLastError(), "sort passes"))
		goto done;
	/* outputs */
	{
		const uint32_t *pos = sb.pos[sb.cur];
		if (want_sorted) {
			out.sorted = bat_new(order_hseq, b->ttype, n);
			if (out.sorted == nullptr)
				goto done;
		}
		if (want_order) {
			out.order = bat_new(order_hseq, B200_oid, n);
			if (out.order == nullptr)
				goto done;
		}
		if (want_sorted || want_order) {
			sort_gather_kernel<T><<<grid_for(n, 256, NSM * 32), 256, 0, c->stream>>>(
				vals, o, hseq, pos, n, want_sorted ? (T *) out.sorted->base : nullptr, want_order ? (oid *) out.order->base : nullptr);
			COUNT_LAUNCH();
		}
		if (want_groups) {
			/* the value images in final order: regenerate them from the final positions */
			U *fk = (U *) sb.keys[0];
			/* keys[0] / keys[1] hold ping-pong images sorted by value only; after the run passes they are stale,
			 * so rebuild: image of the value at each final position */
			{
				/* reuse sort_keys_kernel through an oid view of the final order */
				oid *ord = want_order ? (oid *) out.order->base : (oid *) dev_alloc(n * 8);
				if (ord == nullptr)
					goto done;
				if (!want_order) {
					sort_gather_kernel<T><<<grid_for(n, 256, NSM * 32), 256, 0, c->stream>>>(vals, o, hseq, pos, n, nullptr, ord);
					COUNT_LAUNCH();
				}
				sort_keys_kernel<T><<<grid_for(n, 256, NSM * 32), 256, 0, c->stream>>>(vals, ord, hseq, n, reverse, nilslast, fk, sb.pos[sb.cur ^ 1]);
				COUNT_LAUNCH();
				if (!want_order)
					dev_free(ord);
			}
			out.groups = bat_new(order_hseq, B200_oid, n);
			if (out.groups == nullptr)
				goto done;
			grp_count_kernel<U><<<(unsigned) sb.ntiles, RS_THREADS, 0, c->stream>>>(fk, runs_sorted, n, tile);
			scan32_kernel<<<1, 1024, 0, c->stream>>>(tile, sb.ntiles, tot_d);
			grp_assign_ids_kernel<U><<<(unsigned) sb.ntiles, RS_THREADS, 0, c->stream>>>(fk, runs_sorted, n, tile, (oid *) out.groups->base);
			COUNT_LAUNCH();
			COUNT_LAUNCH();
			COUNT_LAUNCH();
			if (!cuda_ok(cudaMemcpyAsync(&c->slot_h->v[0], tot_d, 8, cudaMemcpyDeviceToHost, c->stream), "group count"))
				goto done;
		}
	}
	if (!cuda_ok(cudaGetLastError(), "sort outputs") || !stream_sync())
		goto done;
	out.ngroups = want_groups ? c->slot_h->v[0] + 1 : 0;
	ok = true;
done:
	dev_free(sb.keys[0]);
	dev_free(sb.keys[1]);
	dev_free(sb.pos[0]);
	dev_free(sb.pos[1]);
	dev_free(sb.aux[0]);
	dev_free(sb.aux[1]);
	dev_free(sb.hist);
	dev_free(tile);
	dev_free(tot_d);
	if (!ok) {
		b200_BATfree(out.sorted);
		b200_BATfree(out.order);
		b200_BATfree(out.groups);
		out = SortOut();
	}
	return ok;
}

extern "C" b200_return b200_BATsort(b200_bat **sorted, b200_bat **order, b200_bat **groups, const b200_bat *b, const b200_bat *o,
				    const b200_bat *g, bool reverse, bool nilslast, bool stable)
{
	b200_clrerr();
	if (sorted)
		*sorted = nullptr;
	if (order)
		*order = nullptr;
	if (groups)
		*groups = nullptr;
	if (b == nullptr) {
		set_err("BATsort: b must exist\n");
		return B200_FAIL;
	}
	if (stable && reverse != nilslast) {
		set_err("stable sort cannot have reverse != nilslast\n");	/* gdk_batop.c:2362-2364 */
		return B200_FAIL;
	}
	if (sorted == nullptr && order == nullptr && groups == nullptr)
		return B200_SUCCEED;	/* "If they're all NULL, this function does nothing" */
	Staged sb_, so, sg;
	if (!stage(b, sb_) || !stage(o, so) || !stage(g, sg))
		return B200_FAIL;
	b = sb_.b;
	o = so.b;
	g = sg.b;
	const uint64_t n = b->count;
	if ((o && o->count != n) || (g && g->count != n)) {
		set_err("BATsort: o and g must be aligned with b\n");
		return B200_FAIL;
	}
	if (b->ccand != B200_CAND_NONE || (o && o->ccand != B200_CAND_NONE) || (g && g->ccand != B200_CAND_NONE) || b->ttype == B200_msk) {
		set_err("BATsort: complex candidate encodings / msk are not on the device path\n");
		return B200_FAIL;
	}
	if (n >= 0xffffffffULL) {
		set_err("BATsort: more than 2^32-1 rows in one device call\n");
		return B200_FAIL;
	}
	const oid ohseq = b->hseqbase;
	/* materialise dense / void o and g (rare: a previous BATsort returns materialised ones) */
	oid *otmp = nullptr, *gtmp = nullptr, *btmp = nullptr;
	const oid *op = nullptr, *gp = nullptr;
	b200_return rc = B200_FAIL;
	SortOut out;
	b200_bat bv;
	const b200_bat *bb = b;
	if (n == 0) {
		if (sorted)
			*sorted = bat_new(ohseq, b->ttype == B200_void ? B200_oid : b->ttype, 0);
		if (order)
			*order = bat_dense(ohseq, ohseq, 0);
		if (groups)
			*groups = bat_dense(ohseq, 0, 0);
		return B200_SUCCEED;
	}
	if (o) {
		if (o->ttype == B200_void) {
			if (o->tseqbase != OID_NIL && o->tseqbase == ohseq)
				op = nullptr;		/* the identity */
			else {
				otmp = (oid *) dev_alloc(n * 8);
				if (otmp == nullptr)
					goto done;
				launch_iota(otmp, n, o->tseqbase);
				op = otmp;
			}
		} else
			op = (const oid *) o->base;
	}
	if (g) {
		if (g->ttype == B200_void) {
			if (g->tseqbase == OID_NIL)
				gp = nullptr;		/* one run */
			else {
				gtmp = (oid *) dev_alloc(n * 8);
				if (gtmp == nullptr)
					goto done;
				launch_iota(gtmp, n, g->tseqbase);
				gp = gtmp;
			}
		} else
			gp = (const oid *) g->base;
	}
	if (b->ttype == B200_void) {
		/* a void column: sort its virtual oids */
		btmp = (oid *) dev_alloc(n * 8);
		if (btmp == nullptr)
			goto done;
		if (b->tseqbase == OID_NIL) {
			set_err("BATsort: void-nil columns are not on the device path\n");
			goto done;
		}
		launch_iota(btmp, n, b->tseqbase);
		bv = *b;
		bv.ttype = B200_oid;
		bv.base = btmp;
		bb = &bv;
	}
	ctx()->algo = "sort: cuda stable LSD radix sort";
	{
		bool ok = false;
		const bool ws = sorted != nullptr, wo = order != nullptr, wg = groups != nullptr;
		Shadow oix;
		tl_sort_index = nullptr;
		if (op == nullptr && gp == nullptr && !reverse && !nilslast && b->ttype != B200_void && b->device >= 0 &&
		    shadow_get(b->base, SHADOW_ORDERIDX, n, oix)) {
			tl_sort_index = (const oid *) oix.p;
			tl_sort_index_hseq = oix.a0;
			ctx()->algo = "sort: order index";
		}
		struct Reset {
			~Reset() { tl_sort_index = nullptr; }
		} reset_index;
		switch (basetype(bb->ttype)) {
		case B200_bte: ok = sort_typed<bte>(bb, op, gp, n, reverse, nilslast, ws, wo, wg, ohseq, out); break;
		case B200_sht: ok = sort_typed<sht>(bb, op, gp, n, reverse, nilslast, ws, wo, wg, ohseq, out); break;
		case B200_int: ok = sort_typed<int>(bb, op, gp, n, reverse, nilslast, ws, wo, wg, ohseq, out); break;
		case B200_lng: ok = sort_typed<lng>(bb, op, gp, n, reverse, nilslast, ws, wo, wg, ohseq, out); break;
		case B200_hge: ok = sort_typed<hge>(bb, op, gp, n, reverse, nilslast, ws, wo, wg, ohseq, out); break;
		case B200_flt: ok = sort_typed<flt>(bb, op, gp, n, reverse, nilslast, ws, wo, wg, ohseq, out); break;
		case B200_dbl: ok = sort_typed<dbl>(bb, op, gp, n, reverse, nilslast, ws, wo, wg, ohseq, out); break;
		default:
			set_err("BATsort: type %s is not on the device path (fixed-width atoms only)\n", atomname(b->ttype));
		}
		if (!ok)
			goto done;
	}
	if (sorted) {
		b200_bat *s = out.sorted;
		s->count = n;
		s->ttype = b->ttype == B200_void ? B200_oid : b->ttype;
		/* gdk_batop.c:2714-2715, 2749-2750: order bits only claimed for the natural nil position */
		s->tsorted = (!reverse && !nilslast && gp == nullptr) || n <= 1;
		s->trevsorted = (reverse && nilslast && gp == nullptr) || n <= 1;
		s->tkey = (b->tkey && gp == nullptr) || n <= 1;
		s->tnonil = b->tnonil;
		s->tnil = b->tnil;
		*sorted = s;
	}
	if (order) {
		b200_bat *x = out.order;
		x->count = n;
		x->tsorted = x->trevsorted = n <= 1;
		x->tkey = 1;
		x->tnonil = 1;
		x->tnil = 0;
		*order = x;
	}
	if (groups) {
		b200_bat *x = out.groups;
		x->count = n;
		x->tsorted = 1;		/* group ids ascend with the sorted position */
		x->trevsorted = n <= 1 || out.ngroups <= 1;
		x->tkey = out.ngroups == n;
		x->tnonil = 1;
		x->tnil = 0;
		x->oidrange_known = 1;
		x->first_oid = 0;
		x->last_oid = out.ngroups - 1;
		*groups = x;
	}
	rc = B200_SUCCEED;
done:
	dev_free(otmp);
	dev_free(gtmp);
	dev_free(btmp);
	return rc;
}

// ---------------------------------------------------------------- BATfirstn, plain form
// gdk_firstn.c:1280: the oids of the n smallest (asc) / largest values of b restricted to s, nils first unless
// nilslast; returned as a candidate list.  Which of several equal boundary values make it is open in the
// reference (its heap decides); here the earliest rows do.  The grouped / distinct / gids forms — the refinement steps
// of a multi-column top-N — are in firstn.cu.
extern "C" b200_return b200_BATfirstn(b200_bat **topn, b200_bat **gids, const b200_bat *b, const b200_bat *s, const b200_bat *g,
				      uint64_t n, bool asc, bool nilslast, bool distinct)
{
	b200_clrerr();
	if (topn)
		*topn = nullptr;
	if (gids)
		*gids = nullptr;
	if (b == nullptr || topn == nullptr) {
		set_err("BATfirstn: b and topn required\n");
		return B200_FAIL;
	}
	if (g != nullptr || gids != nullptr || distinct)
		return firstn_forms(topn, gids, b, s, g, n, asc, nilslast, distinct);	/* firstn.cu */
	/* values of the candidates, their sort order, the first n oids, as an ascending candidate list */
	b200_bat *vals = nullptr, *ord = nullptr, *head = nullptr, *picked = nullptr, *cl = nullptr;
	b200_return rc = B200_FAIL;
	b200_bat *cand = s ? nullptr : bat_dense(0, b->hseqbase, b->count);
	const b200_bat *sc = s ? s : cand;
	if (sc == nullptr)
		return B200_FAIL;
	vals = b200_BATproject(sc, b);
	if (vals == nullptr)
		goto done;
	if (n > vals->count)
		n = vals->count;
	if (n == 0) {
		*topn = bat_dense(0, 0, 0);
		rc = *topn ? B200_SUCCEED : B200_FAIL;
		goto done;
	}
	/* nilslast means the same in both: nils at the end of the conceptual order (gdk_firstn.c:23-27) */
	if (b200_BATsort(nullptr, &ord, nullptr, vals, nullptr, nullptr, !asc, nilslast, false) != B200_SUCCEED)
		goto done;
	head = bat_dense(0, ord->hseqbase, n);		/* positions 0..n-1 of the order */
	if (head == nullptr)
		goto done;
	picked = b200_BATproject(head, ord);		/* positions (in vals) of the first n */
	if (picked == nullptr)
		goto done;
	{
		/* positions in vals -> oids of b through the candidate list, then sort them ascending */
		b200_bat *oids = b200_BATproject(picked, sc);
		if (oids == nullptr)
			goto done;
		b200_return r2 = b200_BATsort(&cl, nullptr, nullptr, oids, nullptr, nullptr, false, false, false);
		b200_BATfree(oids);
		if (r2 != B200_SUCCEED)
			goto done;
	}
	cl->hseqbase = 0;
	cl->tsorted = cl->tkey = cl->tnonil = 1;
	cl->trevsorted = n <= 1;
	*topn = cl;
	cl = nullptr;
	ctx()->algo = "firstn: cuda radix sort + head";
	rc = B200_SUCCEED;
done:
	b200_BATfree(cand);
	b200_BATfree(vals);
	b200_BATfree(ord);
	b200_BATfree(head);
	b200_BATfree(picked);
	b200_BATfree(cl);
	return rc;
}

// ---------------------------------------------------------------- order index (gdk_orderidx.c:74 BATcheckorderidx, :184 BATorderidx)
// The reference keeps a column's stable ascending sort order as a heap beside the tail (b->torderidx) and uses it in
// BATsort (gdk_batop.c:2489-2520) and in range selects (gdk_select.c:1716-1900).  Here it is a twin of the tail heap
// (heaps this library owns), built by the radix sort, dropped with the heap or by b200_BATinvalidate.
extern "C" int b200_BATcheckorderidx(const b200_bat *b)
{
	Shadow sh;
	return b != nullptr && b->device >= 0 && b->base != nullptr && shadow_get(b->base, SHADOW_ORDERIDX, b->count, sh) ? 1 : 0;
}

extern "C" b200_return b200_BATorderidx(const b200_bat *b, bool stable)
{
	b200_clrerr();
	if (b == nullptr) {
		set_err("BATorderidx: BAT required\n");
		return B200_FAIL;
	}
	if (b->ttype == B200_void) {
		set_err("No order index on void type bats\n");	/* gdk_orderidx.c:187 */
		return B200_FAIL;
	}
	if (b200_BATcheckorderidx(b))
		return B200_SUCCEED;
	if (b->device < 0 || !b->owns || b->parent != nullptr || b->ccand != B200_CAND_NONE || !shadow_enabled()) {
		set_err("BATorderidx: an order index can only be kept for a device column whose heap this library owns\n");
		return B200_FAIL;
	}
	b200_bat *on = nullptr;
	if (b200_BATsort(nullptr, &on, nullptr, b, nullptr, nullptr, false, false, stable) != B200_SUCCEED)
		return B200_FAIL;
	if (on->ttype == B200_void || on->base == nullptr || b->count == 0) {
		b200_BATfree(on);	/* nothing to remember */
		return B200_SUCCEED;
	}
	Shadow s;
	s.kind = SHADOW_ORDERIDX;
	s.p = on->base;		/* the order heap changes hands */
	s.n = b->count;
	s.a0 = b->hseqbase;
	s.a1 = 1;		/* always stable: the radix sort is */
	on->base = nullptr;
	on->owns = 0;
	b200_BATfree(on);
	shadow_put(b->base, s);
	ctx()->algo = "create order index";
	return B200_SUCCEED;
}
