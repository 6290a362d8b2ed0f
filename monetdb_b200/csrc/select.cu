// select.cu — BATselect / BATthetaselect on the device.
//
// Host side: the reference's argument normalisation, verbatim in meaning
// (gdk_select.c:1342-1560: lval/hval/lnil/equi/anti handling, anti-range swap,
// the empty / everything shortcuts; scanfunc's closed-range normalisation
// :300-446).  Device side: ONE streaming pass over the column (or a gather
// through a materialised candidate list) evaluating the closed-range predicate,
// with warp-scan + decoupled look-back compaction so the oids come out ascending,
// exactly as scanloop's `dst[cnt] = o; cnt += TEST` (:239-245) produces them.
// The result is virtualised to a dense void BAT when the hits are contiguous
// (virtualize, :31-89).
#include "common.cuh"
#include "compact.cuh"
#include <type_traits>

// closed-range predicate after normalisation:
//   nil(v) ? nilres : (anti ? (v <= vl || v >= vh) : (v >= vl && v <= vh))
template <typename T> struct RangePred {
	T vl, vh;
	bool anti, nilres;
	__device__ __forceinline__ bool operator()(T v) const
	{
		if (Atom<T>::is_nil(v))
			return nilres;
		return anti ? (v <= vl || v >= vh) : (v >= vl && v <= vh);
	}
};

// ---------------------------------------------------------------- dense scan kernel
// Step 1 of the compaction (compact.cuh) specialised for a dense candidate range:
// 128-bit streaming loads, VEC predicate bits per thread, and the 32-row mask
// words assembled from G = 32/VEC neighbouring lanes with one warp-level OR
// reduction.  A CTA covers one CP_CHUNK of rows; all of a thread's loads are
// independent, so 32/VEC of them are in flight per thread.
template <typename T>
__global__ void __launch_bounds__(CP_THREADS)
select_dense_kernel(const T *__restrict__ src,	/* 16-byte aligned; row r of the scan = src[r] */
		    uint64_t skip,		/* leading rows before the first candidate (alignment padding) */
		    uint64_t n,			/* rows to scan including skip */
		    RangePred<T> pred, uint32_t *__restrict__ words, uint32_t *__restrict__ counts)
{
	constexpr int VEC = 16 / sizeof(T);		/* values per 128-bit load */
	constexpr int G = 32 / VEC;			/* lanes that share one mask word */
	constexpr int NV = CP_CHUNK / (CP_THREADS * VEC);	/* loads per thread: 2/4/8/16/32 */
	__shared__ unsigned total;
	if (threadIdx.x == 0)
		total = 0;
	__syncthreads();
	const unsigned lane = threadIdx.x & 31;
	const uint64_t cbase = (uint64_t) blockIdx.x * CP_CHUNK;
	unsigned cnt = 0;
	constexpr int NB = NV > 8 ? 8 : NV;		/* loads issued back to back before any is consumed */
	for (int jb = 0; jb < NV; jb += NB) {
		int4 q[NB];
#pragma unroll
		for (int j = 0; j < NB; j++) {
			const uint64_t r0 = cbase + ((uint64_t) (jb + j) * CP_THREADS + threadIdx.x) * VEC;
			if (r0 + VEC <= n)
				q[j] = __ldcs(reinterpret_cast<const int4 *>(src + r0));
		}
#pragma unroll
		for (int j = 0; j < NB; j++) {
			const uint64_t r0 = cbase + ((uint64_t) (jb + j) * CP_THREADS + threadIdx.x) * VEC;
			unsigned m = 0;
			if (r0 + VEC <= n) {
				union { int4 q; T v[VEC]; } u;
				u.q = q[j];
#pragma unroll
				for (int e = 0; e < VEC; e++)
					m |= (unsigned) pred(u.v[e]) << e;
			} else {
				for (int e = 0; e < VEC; e++)
					if (r0 + e < n)
						m |= (unsigned) pred(src[r0 + e]) << e;
			}
			if (r0 < skip) {
#pragma unroll
				for (int e = 0; e < VEC; e++)
					if (r0 + e < skip)
						m &= ~(1u << e);
			}
			const unsigned gmask = (G == 32 ? 0xffffffffu : ((1u << (G & 31)) - 1)) << (lane & ~(G - 1));
			const unsigned word = __reduce_or_sync(gmask, m << (VEC * (lane & (G - 1))));
			if ((lane & (G - 1)) == 0) {
				words[r0 >> 5] = word;
				cnt += __popc(word);
			}
		}
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
		cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
	if (lane == 0 && cnt)
		atomicAdd(&total, cnt);
	__syncthreads();
	if (threadIdx.x == 0)
		counts[blockIdx.x] = total;
}

// Dense scan for the 4/8/16-byte atoms: lane l of a warp tests row 32j + l, so the
// ballot of step j IS mask word j — no cross-lane assembly, ~7 instructions per 32
// rows.  A warp covers 32 words (1024 rows); all 32 loads of a thread are independent
// and issued back to back (128-512 B in flight per thread).  Integer ranges without
// nil matching use one unsigned compare: (U)(v - vl) <= (U)(vh - vl), which a nil
// (the type minimum, below every valid vl) can never satisfy.
template <typename T> struct FastRange {
	typedef typename std::conditional<sizeof(T) == 4, uint32_t, typename std::conditional<sizeof(T) == 8, uint64_t, uhge>::type>::type U;
	U lo, span;
	__device__ __forceinline__ bool operator()(T v) const { return (U) ((U) v - lo) <= span; }
};

template <typename T, typename P>
__global__ void __launch_bounds__(CP_THREADS)
select_dense32_kernel(const T *__restrict__ src, uint64_t n, P pred, uint32_t *__restrict__ words, uint32_t *__restrict__ counts)
{
	__shared__ unsigned total;
	if (threadIdx.x == 0)
		total = 0;
	__syncthreads();
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint64_t wrow = (uint64_t) blockIdx.x * CP_CHUNK + warp * 1024;	/* first row of this warp */
	constexpr int NB = sizeof(T) == 4 ? 32 : sizeof(T) == 8 ? 16 : 8;	/* loads in flight per thread */
	unsigned cnt = 0, mine = 0;
	if (wrow + 1024 <= n) {
		const T *p = src + wrow + lane;
#pragma unroll
		for (int jb = 0; jb < 32; jb += NB) {
			T v[NB];
#pragma unroll
			for (int j = 0; j < NB; j++)
				v[j] = ld_stream(p + (jb + j) * 32);
#pragma unroll
			for (int j = 0; j < NB; j++) {
				const unsigned bal = __ballot_sync(0xffffffffu, pred(v[j]));
				cnt += __popc(bal);
				mine = lane == jb + j ? bal : mine;
			}
		}
	} else if (wrow < n) {
		for (int j = 0; j < 32; j++) {
			const uint64_t r = wrow + j * 32 + lane;
			const bool h = r < n && pred(src[r]);
			const unsigned bal = __ballot_sync(0xffffffffu, h);
			cnt += __popc(bal);
			mine = lane == j ? bal : mine;
		}
	}
	words[(wrow >> 5) + lane] = mine;	/* the scratch covers whole chunks: rows past n get zero words */
	if (lane == 0 && cnt)
		atomicAdd(&total, cnt);
	__syncthreads();
	if (threadIdx.x == 0)
		counts[blockIdx.x] = total;
}

// Keep the scan's bitmask and chunk offsets as a twin of the oid list about to be returned
// (SHADOW_SELMASK): bit i <-> oid bit0 + i.  Only for large results keeping at least a quarter of the
// scanned rows (below that a gather through the oid list moves fewer bytes than streaming the column).
static void keep_selmask(const CompactScratch &s, uint64_t nbits, oid bit0, const oid *dst, const CompactResult &res)
{
	if (!shadow_enabled() || res.count < (1u << 16) || res.count * 4 < nbits || res.last - res.first + 1 == res.count)
		return;
	Ctx *c = ctx();
	const size_t wbytes = s.nchunks * (s.chunk / 32) * 4, obytes = (s.nchunks + 1) * 8;
	const size_t used = (size_t) ((const char *) s.offs - (const char *) s.words) + obytes;
	Shadow sh;
	sh.kind = SHADOW_SELMASK;
	if ((const void *) s.words == scratch_get(0) && scratch_size() <= 4 * used) {
		/* the mask lives at the start of the thread's scratch block: keep the block itself */
		sh.p = scratch_steal(nullptr);
		sh.a3 = (uint64_t) ((const char *) s.offs - (const char *) s.words);
	} else {
		sh.p = dev_alloc(wbytes + obytes);
		sh.a3 = wbytes;
		if (sh.p == nullptr ||
		    !cuda_ok(cudaMemcpyAsync(sh.p, s.words, wbytes, cudaMemcpyDeviceToDevice, c->stream), "selmask words") ||
		    !cuda_ok(cudaMemcpyAsync((char *) sh.p + wbytes, s.offs, obytes, cudaMemcpyDeviceToDevice, c->stream), "selmask offs")) {
			dev_free(sh.p);
			b200_clrerr();
			return;
		}
	}
	sh.n = res.count;
	sh.a0 = bit0;
	sh.a1 = s.nchunks;
	sh.a2 = s.chunk;
	shadow_put(dst, sh);
}

template <typename T>
static bool run_select_dense(const T *col, uint64_t first_row, uint64_t ncand, oid seq, RangePred<T> pred, oid *dst, CompactResult &res)
{
	Ctx *c = ctx();
	CompactScratch s;
	if constexpr (sizeof(T) >= 4) {
		const T *p = col + first_row;
		if (!compact_scratch(ncand, s, CP_CHUNK))
			return false;
		{
			KTimer kt("select_dense", ncand * sizeof(T));
			bool fast = false;
			if constexpr (!Atom<T>::is_float) {
				if (!pred.anti && !pred.nilres && pred.vl <= pred.vh && pred.vl >= Atom<T>::minval()) {
					typedef typename FastRange<T>::U U;
					FastRange<T> fp{(U) pred.vl, (U) ((U) pred.vh - (U) pred.vl)};
					select_dense32_kernel<T, FastRange<T>><<<(unsigned) s.nchunks, CP_THREADS, 0, c->stream>>>(p, ncand, fp, s.words, s.counts);
					fast = true;
				}
			}
			if (!fast)
				select_dense32_kernel<T, RangePred<T>><<<(unsigned) s.nchunks, CP_THREADS, 0, c->stream>>>(p, ncand, pred, s.words, s.counts);
		}
		COUNT_LAUNCH();
		if (!compact_finish(s, DenseOid{seq}, dst, res))
			return false;
		keep_selmask(s, ncand, seq, dst, res);
		return true;
	} else {
		const T *p = col + first_row;
		uint64_t skip = ((uintptr_t) p & 15) / sizeof(T);
		p -= skip;
		uint64_t n = ncand + skip;
		if (!compact_scratch(n, s, CP_CHUNK))
			return false;
		{
			KTimer kt("select_dense", n * sizeof(T));
			select_dense_kernel<T><<<(unsigned) s.nchunks, CP_THREADS, 0, c->stream>>>(p, skip, n, pred, s.words, s.counts);
		}
		COUNT_LAUNCH();
		if (!compact_finish(s, DenseOid{seq - skip}, dst, res))
			return false;
		keep_selmask(s, n, seq - skip, dst, res);
		return true;
	}
}

// ---------------------------------------------------------------- candidate-list scan
template <typename T> struct GatherPredF {
	const T *col;		/* col[o - hseq] */
	const oid *cand;
	oid hseq;
	RangePred<T> pred;
	__device__ __forceinline__ bool operator()(uint64_t idx, oid &o) const
	{
		o = __ldcs(cand + idx);
		return pred(col[o - hseq]);
	}
	__device__ __forceinline__ oid at(uint64_t idx) const { return cand[idx]; }
	/* two-level form (compact_mask2_kernel) */
	__device__ __forceinline__ oid fetch(uint64_t idx) const { return __ldcs(cand + idx); }
	__device__ __forceinline__ T gather(oid o) const { return col[o - hseq]; }
	__device__ __forceinline__ bool test(T v) const { return pred(v); }
};

// ---------------------------------------------------------------- range select through the order index
// The index lists the oids in ascending value order, nils first: the rows with vl <= v <= vh are one contiguous run of
// it.  Two binary searches find the run; its oids, sorted ascending again, are the result.  Worth it only when the run
// is short against a scan that streams the whole column at HBM speed: the column must hold at least
// B200_OIDX_SELECT_MIN_BYTES (default 256 MiB) and the run at most 1/64 of the rows, otherwise the scan runs.
template <typename T>
__global__ void oidx_bounds_kernel(const oid *__restrict__ idx, oid idx_hseq, const T *__restrict__ vals, uint64_t n, T vl, T vh,
				   unsigned long long *__restrict__ out)
{
	/* thread 0: first position whose value is not below vl; thread 1: first position whose value is above vh */
	const bool upper = threadIdx.x == 1;
	uint64_t lo = 0, hi = n;
	while (lo < hi) {
		const uint64_t mid = (lo + hi) >> 1;
		const T v = vals[idx[mid] - idx_hseq];
		const bool before = Atom<T>::is_nil(v) || (upper ? v <= vh : v < vl);
		if (before)
			lo = mid + 1;
		else
			hi = mid;
	}
	out[threadIdx.x] = lo;
}

__global__ void oidx_slice_kernel(const oid *__restrict__ idx, oid idx_hseq, oid hseq, uint64_t lo, uint64_t k, oid *__restrict__ dst)
{
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < k; i += (uint64_t) gridDim.x * blockDim.x)
		dst[i] = idx[lo + i] - idx_hseq + hseq;
}

// 1: *out holds the result; -1: error; 0: no index / not worth it, scan instead
template <typename T> static int select_by_orderidx(const b200_bat *b, const RangePred<T> &pred, b200_bat **out)
{
	static const uint64_t min_bytes = [] {
		const char *e = getenv("B200_OIDX_SELECT_MIN_BYTES");
		return e ? (uint64_t) strtoull(e, nullptr, 10) : (uint64_t) 256 << 20;
	}();
	const uint64_t n = b->count;
	Shadow oix;
	if (b->device < 0 || n * sizeof(T) < min_bytes || !shadow_get(b->base, SHADOW_ORDERIDX, n, oix))
		return 0;
	Ctx *c = ctx();
	unsigned long long *st = (unsigned long long *) scratch_get(64);
	if (st == nullptr)
		return -1;
	const oid *idx = (const oid *) oix.p;
	oidx_bounds_kernel<T><<<1, 2, 0, c->stream>>>(idx, (oid) oix.a0, (const T *) b->base, n, pred.vl, pred.vh, st);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "orderidx bounds") ||
	    !cuda_ok(cudaMemcpyAsync(c->slot_h->v, st, 16, cudaMemcpyDeviceToHost, c->stream), "read") || !stream_sync())
		return -1;
	const uint64_t lo = c->slot_h->v[0], hi = c->slot_h->v[1];
	if (hi <= lo) {
		*out = bat_dense(0, 0, 0);
		ctx()->algo = "select: orderidx";
		return *out ? 1 : -1;
	}
	const uint64_t k = hi - lo;
	if (k > n / 64)
		return 0;
	b200_bat *tmp = bat_new(0, B200_oid, k);
	if (tmp == nullptr)
		return -1;
	oidx_slice_kernel<<<grid_for(k, 256, NSM * 16), 256, 0, c->stream>>>(idx, (oid) oix.a0, (oid) b->hseqbase, lo, k, (oid *) tmp->base);
	COUNT_LAUNCH();
	tmp->count = k;
	tmp->tnonil = 1;
	b200_bat *srt = nullptr;
	const b200_return rc = b200_BATsort(&srt, nullptr, nullptr, tmp, nullptr, nullptr, false, false, false);	/* "output must be sorted" */
	b200_BATfree(tmp);
	if (rc != B200_SUCCEED)
		return -1;
	oid ends[2];
	if (!cuda_ok(cudaMemcpyAsync(&c->slot_h->v[0], srt->base, 8, cudaMemcpyDeviceToHost, c->stream), "first") ||
	    !cuda_ok(cudaMemcpyAsync(&c->slot_h->v[1], (const oid *) srt->base + (k - 1), 8, cudaMemcpyDeviceToHost, c->stream), "last") ||
	    !stream_sync()) {
		b200_BATfree(srt);
		return -1;
	}
	ends[0] = c->slot_h->v[0];
	ends[1] = c->slot_h->v[1];
	CompactResult r;
	r.count = k;
	r.first = ends[0];
	r.last = ends[1];
	oid *dst = (oid *) srt->base;	/* the sorted heap changes hands */
	srt->base = nullptr;
	srt->owns = 0;
	b200_BATfree(srt);
	*out = oidlist_to_bat(dst, k, r);
	ctx()->algo = "select: orderidx";
	return *out ? 1 : -1;
}

// ---------------------------------------------------------------- typed driver
struct SelArgs {
	const b200_bat *b;
	const Cand *ci;
	const void *tl, *th;
	bool li, hi, equi, anti, nil_matches, lval, hval, lnil;
};

// returns false on error; on success *out is the result (possibly empty dense)
template <typename T>
static bool select_typed(const SelArgs &a, b200_bat **out)
{
	/* scanfunc normalisation to closed ranges, gdk_select.c:300-446 */
	T vl = *(const T *) a.tl, vh = *(const T *) a.th;
	bool li = a.li, hi = a.hi, lval = a.lval, hval = a.hval, anti = a.anti;
	const T minval = Atom<T>::minval(), maxval = Atom<T>::maxval();
	const bool nonil = a.b->tnonil;
	if (anti && li) {
		if (vl == minval) {
			anti = false;
			vl = vh;
			li = !hi;
			hval = false;
		} else {
			vl = Atom<T>::prev(vl);
			li = false;
		}
	}
	if (anti && hi) {
		if (vh == maxval) {
			anti = false;
			vh = vl;
			hi = !li;
			lval = false;
		} else {
			vh = Atom<T>::next(vh);
			hi = false;
		}
	}
	if (!anti) {
		if (lval) {
			if (!li) {
				if (vl == maxval) { *out = bat_dense(0, 0, 0); return *out != nullptr; }
				vl = Atom<T>::next(vl);
				li = true;
			}
		} else {
			vl = minval;
			li = true;
			lval = true;
		}
		if (hval) {
			if (!hi) {
				if (vh == minval) { *out = bat_dense(0, 0, 0); return *out != nullptr; }
				vh = Atom<T>::prev(vh);
				hi = true;
			}
		} else {
			vh = maxval;
			hi = true;
			hval = true;
		}
		if (vl > vh) { *out = bat_dense(0, 0, 0); return *out != nullptr; }
	}
	RangePred<T> pred;
	const char *algo;
	if (a.equi) {
		if (a.lnil) {
			/* is_nil(v): empty range for non-nil values, nil matches */
			pred.vl = maxval; pred.vh = minval; pred.anti = false; pred.nilres = true;
			algo = "select: cuda scan is_nil(v)";
		} else {
			pred.vl = vl; pred.vh = vl; pred.anti = false; pred.nilres = false;
			algo = "select: cuda scan v == vl";
		}
	} else if (anti) {
		pred.vl = vl; pred.vh = vh; pred.anti = true;
		pred.nilres = !nonil && a.nil_matches;
		algo = "select: cuda scan (v <= vl || v >= vh)";
	} else if (nonil && vl == minval) {
		/* v <= vh (also accepts -inf for floating point, as the reference's test does) */
		pred.vl = Atom<T>::is_float ? (T) -INFINITY : vl; pred.vh = vh; pred.anti = false; pred.nilres = false;
		algo = "select: cuda scan v <= vh";
	} else if (vh == maxval) {
		/* v >= vl (also accepts +inf for floating point) */
		pred.vl = vl; pred.vh = Atom<T>::is_float ? (T) INFINITY : vh; pred.anti = false; pred.nilres = false;
		algo = "select: cuda scan v >= vl";
	} else {
		pred.vl = vl; pred.vh = vh; pred.anti = false; pred.nilres = false;
		algo = "select: cuda scan v >= vl && v <= vh";
	}
	ctx()->algo = algo;
	const Cand &ci = *a.ci;
	if (!pred.anti && !pred.nilres && ci.dense && ci.ncand == a.b->count && a.b->count > 0) {
		/* gdk_select.c:1716-1900: a column with an order index answers a range select from the index */
		int used = select_by_orderidx<T>(a.b, pred, out);
		if (used != 0)
			return used > 0;
	}
	oid *dst = (oid *) dev_alloc(ci.ncand * sizeof(oid));
	if (dst == nullptr)
		return false;
	CompactResult r;
	bool ok;
	if (ci.dense) {
		ok = run_select_dense<T>((const T *) a.b->base, ci.seq - ci.hseq, ci.ncand, ci.seq, pred, dst, r);
	} else {
		GatherPredF<T> f{(const T *) a.b->base, ci.oids, ci.hseq, pred};
		ok = run_compact_gather(ci.ncand, f, dst, r);
	}
	if (!ok) {
		dev_free(dst);
		return false;
	}
	*out = oidlist_to_bat(dst, ci.ncand, r);
	if (*out == nullptr)
		shadow_drop(dst);	/* the list was freed: its bitmask twin must not outlive it */
	return *out != nullptr;
}

template <typename T> static int cmp_typed(const void *a, const void *b)
{
	/* ATOMcmp: nil is smaller than everything, nil == nil */
	T x = *(const T *) a, y = *(const T *) b;
	bool xn = Atom<T>::is_nil(x), yn = Atom<T>::is_nil(y);
	if (xn || yn)
		return xn && yn ? 0 : (xn ? -1 : 1);
	return x < y ? -1 : (x > y ? 1 : 0);
}
static int atomcmp(int t, const void *a, const void *b)
{
	switch (t) {
	case B200_bte: return cmp_typed<bte>(a, b);
	case B200_sht: return cmp_typed<sht>(a, b);
	case B200_int: return cmp_typed<int>(a, b);
	case B200_lng: return cmp_typed<lng>(a, b);
	case B200_hge: return cmp_typed<hge>(a, b);
	case B200_flt: return cmp_typed<flt>(a, b);
	case B200_dbl: return cmp_typed<dbl>(a, b);
	}
	return 0;
}
static const void *nilptr(int t)
{
	static const bte nb = Atom<bte>::nil();
	static const sht ns = Atom<sht>::nil();
	static const int ni = Atom<int>::nil();
	static const lng nl = Atom<lng>::nil();
	static const hge nh = Atom<hge>::nil();
	static const flt nf = NAN;
	static const dbl nd = NAN;
	switch (t) {
	case B200_bte: return &nb;
	case B200_sht: return &ns;
	case B200_int: return &ni;
	case B200_lng: return &nl;
	case B200_hge: return &nh;
	case B200_flt: return &nf;
	case B200_dbl: return &nd;
	}
	return nullptr;
}

static b200_bat *select_impl(const b200_bat *b, const b200_bat *s, const void *tl, const void *th,
			     bool li, bool hi, bool anti, bool nil_matches)
{
	if (b == nullptr) {
		set_err("BATselect: BAT required\n");
		return nullptr;
	}
	if (tl == nullptr) {
		set_err("BATselect: tl value required\n");
		return nullptr;
	}
	if (s && s->ttype != B200_msk && !s->tsorted && s->ccand == B200_CAND_NONE && s->ttype != B200_void) {
		set_err("BATselect: invalid argument: s must be sorted.\n");
		return nullptr;
	}
	Staged sb, ss;
	if (!stage(s, ss) || !stage_sparse(b, sb, s != nullptr && s->ttype != B200_void && s->ccand == B200_CAND_NONE ? s->count : UINT64_MAX))
		return nullptr;
	b = sb.b;
	s = ss.b;
	Cand ci;
	if (!cand_init(ci, b, s))
		return nullptr;
	if (ci.ncand == 0) {
		ctx()->algo = "select: trivially empty";
		return bat_dense(0, 0, 0);
	}
	int t = basetype(b->ttype);
	if (b->ttype == B200_void) {
		set_err("BATselect: void/dense tail columns are not on the device path\n");
		return nullptr;
	}
	const void *nil = nilptr(t);
	if (nil == nullptr) {
		set_err("BATselect: unsupported tail type %s (fixed-width atoms only)\n", atomname(b->ttype));
		return nullptr;
	}
	/* gdk_select.c:1394-1409 */
	bool lnil = atomcmp(t, tl, nil) == 0;
	bool lval = !lnil || th == nullptr;
	bool equi = th == nullptr || (lval && atomcmp(t, tl, th) == 0);
	bool hval;
	if (lnil && nil_matches && (th == nullptr || atomcmp(t, th, nil) == 0)) {
		equi = true;
		lval = true;
	}
	if (equi) {
		if (th == nullptr)
			hi = li;
		th = tl;
		hval = true;
		if (!anti && (!li || !hi)) {
			ctx()->algo = "select: empty interval";
			return bat_dense(0, 0, 0);
		}
	} else {
		nil_matches = false;
		hval = atomcmp(t, th, nil) != 0;
	}
	bool notnil_select = false;	/* "everything except nil" (:1480-1514) */
	if (anti) {
		if (lval != hval) {
			/* one end point is nil: swap sub-ranges (:1443-1465) */
			const void *tv = tl;
			bool ti = li;
			li = !hi;
			hi = !ti;
			tl = th;
			th = tv;
			ti = lval;
			lval = hval;
			hval = ti;
			lnil = atomcmp(t, tl, nil) == 0;
			anti = false;
		} else if (!lval && !hval) {
			ctx()->algo = "select: anti: nil-nil range, nonil";
			return bat_dense(0, 0, 0);
		} else if ((equi && (lnil || !(li && hi))) || atomcmp(t, tl, th) > 0) {
			if (equi && !lnil && nil_matches && !(li && hi)) {
				ctx()->algo = "select: anti, equi, open, nil_matches";
				return cand_to_bat(ci);
			}
			notnil_select = true;
		} else {
			/* antiequi handled by the anti predicate below */
			equi = false;
		}
	}
	if (!notnil_select) {
		if (hval && (equi ? !li || !hi : atomcmp(t, tl, th) > 0)) {
			ctx()->algo = "select: empty range";
			return bat_dense(0, 0, 0);
		}
		if (equi && lnil && b->tnonil) {
			ctx()->algo = "select: equi-nil, nonil";
			return bat_dense(0, 0, 0);
		}
		if (!equi && !lval && !hval && lnil && b->tnonil) {
			ctx()->algo = "select: everything, nonil";
			return cand_to_bat(ci);
		}
	} else if (b->tnonil) {
		ctx()->algo = "select: everything, nonil";
		return cand_to_bat(ci);
	}
	SelArgs a;
	a.b = b;
	a.ci = &ci;
	if (notnil_select) {
		/* x != nil  ==  nil-nil range, non-anti (:1296 of the table) */
		a.tl = nil; a.th = nil; a.li = a.hi = true;
		a.equi = false; a.anti = false; a.nil_matches = false;
		a.lval = a.hval = false; a.lnil = true;
	} else {
		a.tl = tl; a.th = th; a.li = li; a.hi = hi;
		a.equi = equi; a.anti = anti; a.nil_matches = nil_matches;
		a.lval = lval; a.hval = hval; a.lnil = lnil;
	}
	b200_bat *bn = nullptr;
	bool ok = false;
	switch (t) {
	case B200_bte: ok = select_typed<bte>(a, &bn); break;
	case B200_sht: ok = select_typed<sht>(a, &bn); break;
	case B200_int: ok = select_typed<int>(a, &bn); break;
	case B200_lng: ok = select_typed<lng>(a, &bn); break;
	case B200_hge: ok = select_typed<hge>(a, &bn); break;
	case B200_flt: ok = select_typed<flt>(a, &bn); break;
	case B200_dbl: ok = select_typed<dbl>(a, &bn); break;
	}
	return ok ? bn : nullptr;
}

extern "C" b200_bat *b200_BATselect(const b200_bat *b, const b200_bat *s, const void *tl, const void *th,
				    bool li, bool hi, bool anti, bool nil_matches)
{
	b200_clrerr();
	if (b != nullptr && b->ttype == B200_STR_TYPE)
		return str_select_entry(b, s, tl, th, li, hi, anti, nil_matches);
	return select_impl(b, s, tl, th, li, hi, anti, nil_matches);
}

// gdk_select.c:2102-2154
extern "C" b200_bat *b200_BATthetaselect(const b200_bat *b, const b200_bat *s, const void *val, const char *op)
{
	b200_clrerr();
	if (b == nullptr || val == nullptr || op == nullptr) {
		set_err("BATthetaselect: BAT, value and operator required\n");
		return nullptr;
	}
	if (b->ttype == B200_STR_TYPE)
		return str_thetaselect_entry(b, s, val, op);
	int t = basetype(b->ttype);
	const void *nil = nilptr(t);
	if (nil == nullptr) {
		set_err("BATthetaselect: unsupported tail type %s\n", atomname(b->ttype));
		return nullptr;
	}
	/* eq/ne can be used for "is" nil-handling */
	if (strcmp(op, "eq") == 0)
		return select_impl(b, s, val, nullptr, true, true, false, true);
	if (strcmp(op, "ne") == 0)
		return select_impl(b, s, val, nullptr, true, true, true, true);
	if (atomcmp(t, val, nil) == 0)
		return bat_dense(0, 0, 0);
	if (op[0] == '=' && ((op[1] == '=' && op[2] == 0) || op[1] == 0))
		return select_impl(b, s, val, nullptr, true, true, false, false);
	if (op[0] == '!' && op[1] == '=' && op[2] == 0)
		return select_impl(b, s, val, nullptr, true, true, true, false);
	if (op[0] == '<') {
		if (op[1] == 0)
			return select_impl(b, s, nil, val, false, false, false, false);
		if (op[1] == '=' && op[2] == 0)
			return select_impl(b, s, nil, val, false, true, false, false);
		if (op[1] == '>' && op[2] == 0)
			return select_impl(b, s, val, nullptr, true, true, true, false);
	}
	if (op[0] == '>') {
		if (op[1] == 0)
			return select_impl(b, s, val, nil, false, false, false, false);
		if (op[1] == '=' && op[2] == 0)
			return select_impl(b, s, val, nil, true, false, false, false);
	}
	set_err("BATthetaselect: unknown operator.\n");
	return nullptr;
}
