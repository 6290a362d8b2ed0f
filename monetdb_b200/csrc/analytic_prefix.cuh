// analytic_prefix.cuh — O(1)-per-row framed SUM for the integer -> hge case.
//
// When the result type is hge and the input at most 64 bits wide no sum of fewer than
// 2^63 rows can overflow, so the reference's checked tree walk (gdk_analytic_func.c:1762-1796)
// and a difference of two running totals give the same value:  sum[fs,fe) = P[fe] - P[fs],
// with P the exclusive prefix sum of the non-nil values over the WHOLE column (partition
// boundaries cancel in the difference) and C the matching prefix count of non-nil rows (only
// built when the column may hold nils; an all-nil frame yields nil).
// Three streaming kernels build P (tile sums, one-CTA scan of the tile sums, tile rescan);
// the frame kernel then reads two prefix entries per row instead of walking ~40 tree nodes.
#pragma once
#include "analytic.cuh"

#define AP_THREADS 256
#define AP_ITEMS 8
#define AP_TILE (AP_THREADS * AP_ITEMS)

#ifdef __CUDACC__
__device__ __forceinline__ uhge ap_shfl_up(uhge v, int d)
{
	const unsigned long long lo = __shfl_up_sync(0xffffffffu, (unsigned long long) v, d);
	const unsigned long long hi = __shfl_up_sync(0xffffffffu, (unsigned long long) (v >> 64), d);
	return ((uhge) hi << 64) | lo;
}

// CTA-wide exclusive scan of one (sum, count) pair per thread; returns the CTA totals too
__device__ __forceinline__ void ap_block_scan(uhge &sum, unsigned &cnt, uhge &tot_sum, unsigned &tot_cnt)
{
	__shared__ uhge wsum[AP_THREADS / 32];
	__shared__ unsigned wcnt[AP_THREADS / 32];
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	uhge is = sum;
	unsigned ic = cnt;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uhge ts = ap_shfl_up(is, o);
		const unsigned tc = __shfl_up_sync(0xffffffffu, ic, o);
		if (lane >= o) {
			is += ts;
			ic += tc;
		}
	}
	if (lane == 31) {
		wsum[warp] = is;
		wcnt[warp] = ic;
	}
	__syncthreads();
	uhge bs = 0, ts = 0;
	unsigned bc = 0, tc = 0;
#pragma unroll
	for (int w = 0; w < AP_THREADS / 32; w++) {
		if (w < (int) warp) {
			bs += wsum[w];
			bc += wcnt[w];
		}
		ts += wsum[w];
		tc += wcnt[w];
	}
	sum = bs + is - sum;
	cnt = bc + ic - cnt;
	tot_sum = ts;
	tot_cnt = tc;
	__syncthreads();
}

// A CTA takes AP_SUB consecutive tiles (one entry of tsum / tcnt per CTA): the single-CTA scan of those entries is the
// serial part of the build, and a CTA that only reduces one 16 KB tile spends most of its life in the reduction's barriers.
#define AP_SUB 4
#define AP_SUPER (AP_TILE * AP_SUB)
template <typename T1>
__global__ void __launch_bounds__(AP_THREADS)
an_prefix_tile_kernel(const T1 *__restrict__ bp, uint64_t n, uhge *__restrict__ tsum, unsigned long long *__restrict__ tcnt)
{
	const uint64_t base = (uint64_t) blockIdx.x * AP_SUPER;
	uhge s = 0;
	unsigned c = 0;
	bool done = false;
	if constexpr (sizeof(T1) == 8) {
		/* full super-tile of 8-byte values: 16-byte loads (base is a multiple of 8192 rows, so bp + base keeps
		 * bp's alignment) */
		if (base + AP_SUPER <= n && ((uintptr_t) bp & 15) == 0) {
			const ulonglong2 *vp = reinterpret_cast<const ulonglong2 *>(bp + base);
#pragma unroll
			for (int i = 0; i < AP_ITEMS * AP_SUB / 2; i++) {
				const ulonglong2 w = __ldcs(vp + i * AP_THREADS + threadIdx.x);
				const T1 v0 = (T1) w.x, v1 = (T1) w.y;
				if (!Atom<T1>::is_nil(v0)) {
					s += (uhge) (hge) v0;
					c++;
				}
				if (!Atom<T1>::is_nil(v1)) {
					s += (uhge) (hge) v1;
					c++;
				}
			}
			done = true;
		}
	}
	if (!done) {
#pragma unroll
		for (int i = 0; i < AP_ITEMS * AP_SUB; i++) {
			const uint64_t k = base + (uint64_t) i * AP_THREADS + threadIdx.x;
			if (k < n) {
				const T1 v = ld_stream(bp + k);
				if (!Atom<T1>::is_nil(v)) {
					s += (uhge) (hge) v;
					c++;
				}
			}
		}
	}
	uhge ts;
	unsigned tc;
	ap_block_scan(s, c, ts, tc);
	if (threadIdx.x == 0) {
		tsum[blockIdx.x] = ts;
		tcnt[blockIdx.x] = tc;
	}
}

// exclusive scan of the tile totals by one CTA (each thread owns a contiguous run)
static __global__ void __launch_bounds__(AP_THREADS)
an_prefix_scan_kernel(uhge *__restrict__ tsum, unsigned long long *__restrict__ tcnt, uint64_t ntiles)
{
	const uint64_t items = (ntiles + AP_THREADS - 1) / AP_THREADS;
	const uint64_t lo = threadIdx.x * items, hi = lo + items < ntiles ? lo + items : ntiles;
	uhge s = 0;
	unsigned long long c = 0;
	for (uint64_t i = lo; i < hi; i++) {
		s += tsum[i];
		c += tcnt[i];
	}
	/* counts can exceed 32 bits across the column: scan them as the low half of a second uhge */
	uhge cs = c, tots, totc;
	unsigned dummy = 0, dt;
	ap_block_scan(s, dummy, tots, dt);
	dummy = 0;
	ap_block_scan(cs, dummy, totc, dt);
	unsigned long long cc = (unsigned long long) cs;
	for (uint64_t i = lo; i < hi; i++) {
		const uhge t = tsum[i];
		const unsigned long long u = tcnt[i];
		tsum[i] = s;
		tcnt[i] = cc;
		s += t;
		cc += u;
	}
}

// P[k] = sum of the non-nil values of rows [0, k), k = 0..n; C likewise (optional)
template <typename T1>
__global__ void __launch_bounds__(AP_THREADS)
an_prefix_write_kernel(const T1 *__restrict__ bp, uint64_t n, const uhge *__restrict__ tsum,
		       const unsigned long long *__restrict__ tcnt, uhge *__restrict__ P, unsigned long long *__restrict__ C)
{
	/* a thread's 8 running totals are 128 consecutive bytes: written straight from the registers every store
	 * instruction touches 32 different lines (ncu: LSU-bound at 39 % of DRAM peak).  They are staged in shared memory
	 * (one element of padding per thread keeps the 16-byte accesses off each other's banks) and leave the CTA as
	 * fully coalesced rows. */
	__shared__ uhge sp[AP_TILE + AP_THREADS];
	uhge carry = tsum[blockIdx.x];			/* totals in front of the tile being written */
	unsigned long long carryc = tcnt[blockIdx.x];
	if (blockIdx.x == 0 && threadIdx.x == 0 && n > 0) {
		P[0] = 0;
		if (C)
			C[0] = 0;
	}
	for (int sub = 0; sub < AP_SUB; sub++) {
		const uint64_t tbase = (uint64_t) blockIdx.x * AP_SUPER + (uint64_t) sub * AP_TILE;
		if (tbase >= n)
			break;
		const uint64_t base = tbase + (uint64_t) threadIdx.x * AP_ITEMS;
		T1 v[AP_ITEMS];
		uhge s = 0;
		unsigned c = 0;
#pragma unroll
		for (int i = 0; i < AP_ITEMS; i++) {
			const uint64_t k = base + i;
			v[i] = k < n ? bp[k] : Atom<T1>::nil();
			if (!Atom<T1>::is_nil(v[i])) {
				s += (uhge) (hge) v[i];
				c++;
			}
		}
		uhge ts;
		unsigned tc;
		ap_block_scan(s, c, ts, tc);
		uhge run = carry + s;
		unsigned long long rc = carryc + c;
		const unsigned sbase = threadIdx.x * (AP_ITEMS + 1);
#pragma unroll
		for (int i = 0; i < AP_ITEMS; i++) {
			const uint64_t k = base + i;
			if (k < n) {
				if (!Atom<T1>::is_nil(v[i])) {
					run += (uhge) (hge) v[i];
					rc++;
				}
				if (C)
					C[k + 1] = rc;
			}
			sp[sbase + i] = run;
		}
		__syncthreads();
#pragma unroll
		for (int i = 0; i < AP_ITEMS; i++) {
			const unsigned j = i * AP_THREADS + threadIdx.x;
			const uint64_t k = tbase + j;
			if (k < n)
				P[k + 1] = sp[j + j / AP_ITEMS];
		}
		carry += ts;
		carryc += tc;
		__syncthreads();	/* sp is rewritten by the next tile */
	}
}

// frames as in an_frame_kernel (analytic_tree.cuh); result hge, nil when the frame holds no value
template <bool PEERS>
__global__ void __launch_bounds__(AN_THREADS)
an_frame_prefix_kernel(const oid *__restrict__ S, uint64_t cnt, const oid *__restrict__ O, uint64_t ocnt, uint64_t n,
		       const oid *__restrict__ fstart, const oid *__restrict__ fend, int frame_type,
		       const uhge *__restrict__ P, const unsigned long long *__restrict__ C, hge *__restrict__ rb,
		       unsigned long long *__restrict__ state)
{
	__shared__ TileSegs ts;
	__shared__ TileSegs to;
	const uint64_t base = (uint64_t) blockIdx.x * AN_TILE;
	const uint64_t last = (base + AN_TILE < n ? base + AN_TILE : n) - 1;
	const bool need_part = frame_type != 6;
	if (need_part)
		ts.load(S, cnt, base, last);
	if (PEERS)
		to.load(O, ocnt, base, last);
	unsigned nils = 0;
	const bool framed = frame_type < 3 || frame_type > 6;
	/* frame of row k, given its fstart / fend entries (framed types only) -> the sum, or nil */
	auto row = [&](uint64_t k, oid fs, oid fe) -> hge {
		oid ps = 0, pe = n;
		if (need_part)
			ts.locate(k, ps, pe);
		switch (frame_type) {
		case 3:
		case 4: {
			oid os = ps, oe = pe;
			if (PEERS)
				to.locate(k, os, oe);
			fs = frame_type == 3 ? ps : os;
			fe = frame_type == 3 ? oe : pe;
			break;
		}
		case 5: fs = ps; fe = pe; break;
		case 6: fs = k; fe = k + 1; break;
		default:
			if (fs < ps)
				fs = ps;
			if (fe > pe)
				fe = pe;
			break;
		}
		if (fs < fe && (C == nullptr || C[fe] != C[fs]))
			return (hge) (P[fe] - P[fs]);
		nils++;
		return Atom<hge>::nil();
	};
	/* two consecutive rows per thread: 16-byte loads of the frame columns (8-byte loads per lane stream at little
	 * more than half of that) and 32 contiguous bytes of results */
	const bool vec = framed && (((uintptr_t) fstart | (uintptr_t) fend) & 15) == 0;
	for (unsigned i = threadIdx.x * 2; i < AN_TILE; i += AN_THREADS * 2) {
		const uint64_t k = base + i;
		if (k >= n)
			break;
		if (k + 1 < n) {
			ulonglong2 fsv = make_ulonglong2(0, 0), fev = make_ulonglong2(0, 0);
			if (vec) {
				fsv = *reinterpret_cast<const ulonglong2 *>(fstart + k);	/* k is even */
				fev = *reinterpret_cast<const ulonglong2 *>(fend + k);
			} else if (framed) {
				fsv = make_ulonglong2(fstart[k], fstart[k + 1]);
				fev = make_ulonglong2(fend[k], fend[k + 1]);
			}
			const hge r0 = row(k, fsv.x, fev.x), r1 = row(k + 1, fsv.y, fev.y);
			rb[k] = r0;
			rb[k + 1] = r1;
		} else {
			rb[k] = row(k, framed ? fstart[k] : 0, framed ? fend[k] : 0);
		}
	}
	nils = __reduce_add_sync(0xffffffffu, nils);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(&state[1], (unsigned long long) nils);
}

// ---------------------------------------------------------------- the running totals of a value column
// Built by the first window aggregate over a column and kept as the column's SHADOW_PREFIX twin (heaps this library
// owns only; at most AP_CACHE_ROWS rows — 24 bytes per row stay allocated until the column is freed or rewritten):
// a window plan runs several aggregates over the same column (sum and avg, ROWS and RANGE frames), which then share
// one build.
#define AP_CACHE_ROWS ((uint64_t) 256 << 20)
struct PrefixArrays {
	uhge *P = nullptr;
	unsigned long long *C = nullptr;
	bool cached = false;
	~PrefixArrays()
	{
		if (!cached) {
			if (P)
				dev_free(P);
			if (C)
				dev_free(C);
		}
	}
};

// the three build kernels over a plain device array: P[0..n] (and C when given) must hold n + 1 entries
template <typename T1> static bool an_prefix_build_raw(const T1 *bp, uint64_t n, uhge *P, unsigned long long *C, const char *kname)
{
	Ctx *c = ctx();
	if (n == 0)
		return cuda_ok(cudaMemsetAsync(P, 0, 16, c->stream), "prefix zero") &&
		       (C == nullptr || cuda_ok(cudaMemsetAsync(C, 0, 8, c->stream), "prefix zero"));
	const uint64_t ntiles = (n + AP_SUPER - 1) / AP_SUPER;
	uhge *tsum = (uhge *) dev_alloc(ntiles * 16);
	unsigned long long *tcnt = (unsigned long long *) dev_alloc(ntiles * 8);
	bool ok = tsum && tcnt;
	if (ok) {
		KTimer kt(kname, n * (2 * sizeof(T1) + 16 + (C ? 8 : 0)));
		an_prefix_tile_kernel<T1><<<(unsigned) ntiles, AP_THREADS, 0, c->stream>>>(bp, n, tsum, tcnt);
		COUNT_LAUNCH();
		an_prefix_scan_kernel<<<1, AP_THREADS, 0, c->stream>>>(tsum, tcnt, ntiles);
		COUNT_LAUNCH();
		an_prefix_write_kernel<T1><<<(unsigned) ntiles, AP_THREADS, 0, c->stream>>>(bp, n, tsum, tcnt, P, C);
		COUNT_LAUNCH();
		ok = cuda_ok(cudaGetLastError(), "prefix launch");
	}
	if (tsum)
		dev_free(tsum);
	if (tcnt)
		dev_free(tcnt);
	return ok;
}

template <typename T1> static bool an_prefix_get(const b200_bat *b, uint64_t n, PrefixArrays &out)
{
	Ctx *c = ctx();
	const bool want_c = !b->tnonil;
	const bool cacheable = b->device >= 0 && b->owns && b->parent == nullptr && n <= AP_CACHE_ROWS && shadow_enabled();
	Shadow sh;
	if (cacheable && shadow_get(b->base, SHADOW_PREFIX, n, sh) && (sh.p2 != nullptr || !want_c)) {
		out.P = (uhge *) sh.p;
		out.C = (unsigned long long *) sh.p2;
		out.cached = true;
		return true;
	}
	const uint64_t ntiles = (n + AP_SUPER - 1) / AP_SUPER;
	uhge *tsum = (uhge *) dev_alloc(ntiles * 16);
	unsigned long long *tcnt = (unsigned long long *) dev_alloc(ntiles * 8);
	out.P = (uhge *) dev_alloc((n + 1) * 16);
	out.C = want_c ? (unsigned long long *) dev_alloc((n + 1) * 8) : nullptr;
	bool ok = tsum && tcnt && out.P && (!want_c || out.C);
	if (ok) {
		const T1 *bp = (const T1 *) b->base;
		{
			KTimer kt("an_prefix_tile", n * sizeof(T1));
			an_prefix_tile_kernel<T1><<<(unsigned) ntiles, AP_THREADS, 0, c->stream>>>(bp, n, tsum, tcnt);
			COUNT_LAUNCH();
		}
		{
			KTimer kt("an_prefix_scan", ntiles * 48);
			an_prefix_scan_kernel<<<1, AP_THREADS, 0, c->stream>>>(tsum, tcnt, ntiles);
			COUNT_LAUNCH();
		}
		KTimer kt("an_prefix", n * (sizeof(T1) + 16 + (out.C ? 8 : 0)));
		an_prefix_write_kernel<T1><<<(unsigned) ntiles, AP_THREADS, 0, c->stream>>>(bp, n, tsum, tcnt, out.P, out.C);
		COUNT_LAUNCH();
		ok = cuda_ok(cudaGetLastError(), "prefix launch");
	}
	if (tsum)
		dev_free(tsum);
	if (tcnt)
		dev_free(tcnt);
	if (!ok)
		return false;	/* ~PrefixArrays releases P / C */
	if (cacheable && stream_sync()) {	/* published to other threads' streams only once it is complete */
		Shadow s;
		s.kind = SHADOW_PREFIX;
		s.p = out.P;
		s.p2 = out.C;
		s.n = n;
		shadow_put(b->base, s);
		out.cached = true;
	}
	return true;
}
#endif
