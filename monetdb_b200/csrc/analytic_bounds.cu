// analytic_bounds.cu — GDKanalyticaldiff and GDKanalyticalwindowbounds
// (gdk_analytic_bounds.c:95-175 and :1440-1456; ROWS :855-991, RANGE :994-1291,
// GROUPS :1295-1437, UNBOUNDED :614-651, peers :654-852).
//
// The reference processes one partition after the other and one row after the other;
// here every row is independent once it knows its partition [start, end), which it
// gets from the staged segment list of analytic.cuh.
#include <atomic>
#include "analytic.cuh"

// ---------------------------------------------------------------- shared helpers
unsigned long long *an_state_init()
{
	unsigned long long *st = (unsigned long long *) dev_alloc(16);
	if (st == nullptr)
		return nullptr;
	if (!cuda_ok(cudaMemsetAsync(st, 0, 16, ctx()->stream), "state init")) {
		dev_free(st);
		return nullptr;
	}
	return st;
}

// reads {error bits, nil count}, frees the state
bool an_read_state(unsigned long long *state, unsigned long long out[2])
{
	Ctx *c = ctx();
	bool ok = cuda_ok(cudaGetLastError(), "analytic launch") &&
		  cuda_ok(cudaMemcpyAsync(c->slot_h->v, state, 16, cudaMemcpyDeviceToHost, c->stream), "state read") &&
		  stream_sync();
	dev_free(state);
	if (!ok)
		return false;
	out[0] = c->slot_h->v[0];
	out[1] = c->slot_h->v[1];
	return true;
}

bool an_check_bit(const b200_bat *x, const char *what, uint64_t n)
{
	if (x == nullptr)
		return true;
	if (basetype(x->ttype) != B200_bte || x->base == nullptr || x->count < n) {
		set_err("analytic: %s must be a bit column aligned with the input\n", what);
		return false;
	}
	return true;
}

// A window plan hands the same partition column to its bounds and to every aggregate (sql_rank.c); the list of
// partition starts is therefore kept as a twin of that column (a heap this library owns, e.g. GDKanalyticaldiff's
// result) and rebuilt only when the column is rewritten (every GDKanalytical* entry point drops the twin of its r).
bool an_partitions(const b200_bat *p, const b200_bat *o, uint64_t n, SegList &out)
{
	const bool cacheable = p != nullptr && o == nullptr && n > 0 && p->device >= 0 && p->owns && p->parent == nullptr && shadow_enabled();
	Shadow sh;
	if (cacheable && shadow_get(p->base, SHADOW_SEGLIST, n, sh)) {
		out.S = (oid *) sh.p;
		out.cnt = sh.a0;
		out.n = n;
		out.cached = true;
		out.token = sh.a1;
		out.maxlen1 = sh.a2;
		out.pbase = p->base;
		return true;
	}
	FlagStart f{p ? (const bte *) p->base : nullptr, o ? (const bte *) o->base : nullptr};
	if (!build_seglist(n, f, out))
		return false;
	if (cacheable) {
		Shadow s;
		s.kind = SHADOW_SEGLIST;
		s.p = out.S;
		s.n = n;
		s.a0 = out.cnt;
		static std::atomic<uint64_t> next_token{1};
		s.a1 = out.token = next_token.fetch_add(1);
		shadow_put(p->base, s);
		out.cached = true;
		out.pbase = p->base;
	}
	return true;
}

// ---------------------------------------------------------------- GDKanalyticaldiff
template <typename T> __device__ __forceinline__ bool an_differs(T a, T b)
{
	if constexpr (Atom<T>::is_float)
		return a != b && !(Atom<T>::is_nil(a) && Atom<T>::is_nil(b));	/* nil == nil */
	else
		return a != b;
}

template <typename T>
__global__ void __launch_bounds__(256)
an_diff_kernel(const T *__restrict__ b, const bte *__restrict__ np, int mode /* 0 none, 1 np, 2 npbit */, bte npb,
	       uint64_t n, bte *__restrict__ r)
{
	uint64_t done = 0;
	if constexpr (sizeof(T) == 4 || sizeof(T) == 8) {
		/* four consecutive rows per thread: 16-byte loads of the values (one-row-per-lane loads and byte stores
		 * reach about half the bandwidth), the four flags leave as one 32-bit store */
		if ((((uintptr_t) b) & 15) == 0 && (((uintptr_t) r | (uintptr_t) np) & 3) == 0) {
			const uint64_t nquad = n / 4;
			for (uint64_t q = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; q < nquad; q += (uint64_t) gridDim.x * blockDim.x) {
				const uint64_t i0 = q * 4;
				T v[5];
				if constexpr (sizeof(T) == 8) {
					const uint4 w0 = reinterpret_cast<const uint4 *>(b + i0)[0], w1 = reinterpret_cast<const uint4 *>(b + i0)[1];
					memcpy(&v[1], &w0, 16);
					memcpy(&v[3], &w1, 16);
				} else {
					const uint4 w0 = reinterpret_cast<const uint4 *>(b + i0)[0];
					memcpy(&v[1], &w0, 16);
				}
				v[0] = i0 > 0 ? b[i0 - 1] : v[1];
				unsigned keep4 = mode == 1 ? *reinterpret_cast<const unsigned *>(np + i0) : mode == 2 ? 0x01010101u * (unsigned) (unsigned char) npb : 0u;
				unsigned out = 0;
#pragma unroll
				for (int j = 0; j < 4; j++) {
					const bool d = (i0 + j > 0) && an_differs<T>(v[j + 1], v[j]);
					const unsigned kb = (keep4 >> (8 * j)) & 0xffu;
					out |= (d ? 1u : kb) << (8 * j);
				}
				*reinterpret_cast<unsigned *>(r + i0) = out;
			}
			done = nquad * 4;
		}
	}
	for (uint64_t i = done + (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
		const bte keep = mode == 1 ? np[i] : mode == 2 ? npb : 0;
		r[i] = (i > 0 && an_differs<T>(b[i], b[i - 1])) ? (bte) 1 : keep;
	}
}

template <typename T>
static void diff_launch(const b200_bat *b, const b200_bat *p, const int8_t *npbit, uint64_t n, bte *r)
{
	Ctx *c = ctx();
	KTimer kt("an_diff", n * (sizeof(T) + 1 + (p ? 1 : 0)));
	an_diff_kernel<T><<<grid_for(sizeof(T) == 4 || sizeof(T) == 8 ? n / 4 + 1 : n, 256, NSM * 32), 256, 0, c->stream>>>(
		(const T *) b->base, p ? (const bte *) p->base : nullptr, p ? 1 : npbit ? 2 : 0, npbit ? *npbit : 0, n, r);
	COUNT_LAUNCH();
}

extern "C" b200_return b200_GDKanalyticaldiff(b200_bat *r, const b200_bat *b, const b200_bat *p, const int8_t *npbit, int tpe)
{
	b200_clrerr();
	if (r == nullptr || b == nullptr) {
		set_err("GDKanalyticaldiff: BAT required\n");
		return B200_FAIL;
	}
	Staged sb, sp;
	if (!stage(b, sb) || !stage(p, sp))
		return B200_FAIL;
	b = sb.b;
	p = sp.b;
	const uint64_t n = b->count;
	if (r->device < 0 || basetype(r->ttype) != B200_bte || r->capacity < n) {
		set_err("GDKanalyticaldiff: result must be a device bit BAT with capacity >= BATcount(b)\n");
		return B200_FAIL;
	}
	shadow_drop(r->base);	/* r is rewritten: what was cached about its old contents goes */
	if (!an_check_bit(p, "p", n))
		return B200_FAIL;
	if (n > 0) {
		ctx()->algo = "analyticaldiff: cuda streaming";
		switch (basetype(tpe)) {
		case B200_bte: diff_launch<bte>(b, p, npbit, n, (bte *) r->base); break;
		case B200_sht: diff_launch<sht>(b, p, npbit, n, (bte *) r->base); break;
		case B200_int: diff_launch<int>(b, p, npbit, n, (bte *) r->base); break;
		case B200_lng: diff_launch<lng>(b, p, npbit, n, (bte *) r->base); break;
		case B200_hge: diff_launch<hge>(b, p, npbit, n, (bte *) r->base); break;
		case B200_flt: diff_launch<flt>(b, p, npbit, n, (bte *) r->base); break;
		case B200_dbl: diff_launch<dbl>(b, p, npbit, n, (bte *) r->base); break;
		default:
			set_err("GDKanalyticaldiff: type %s is not on the device path (fixed-width atoms only)\n", atomname(tpe));
			return B200_FAIL;
		}
		if (!cuda_ok(cudaGetLastError(), "diff launch"))
			return B200_FAIL;
	}
	r->count = n;
	r->tnonil = 1;
	r->tnil = 0;
	r->tsorted = r->trevsorted = n <= 1;
	r->tkey = n <= 1;
	return B200_SUCCEED;
}

// ---------------------------------------------------------------- window bounds
#define OID_MAX (OID_NIL - 1)	/* GDK_oid_max */

// limit of row k as the kernel's compute type LT, from a dynamic column of DT or the static value
template <typename LT, typename DT>
__device__ __forceinline__ bool an_limit(const DT *__restrict__ lim, LT slim, uint64_t k, LT &out)
{
	if (lim) {
		const DT v = lim[k];
		if (Atom<DT>::is_nil(v) || v < 0)
			return false;
		out = (LT) v;
	} else {
		out = slim;
	}
	return true;
}
template <typename LT> __device__ __forceinline__ oid an_limit_oid(LT v)
{
	if constexpr (sizeof(LT) == 16)
		return v > (LT) OID_MAX ? OID_MAX : (oid) v;
	else
		return (oid) v;
}

// UNBOUNDED PRECEDING / FOLLOWING (allbounds) and peer bounds: start or end of the segment
__global__ void __launch_bounds__(AN_THREADS)
an_segbound_kernel(const oid *__restrict__ S, uint64_t cnt, uint64_t n, bool preceding, oid *__restrict__ rb)
{
	__shared__ TileSegs ts;
	const uint64_t base = (uint64_t) blockIdx.x * AN_TILE;
	const uint64_t last = (base + AN_TILE < n ? base + AN_TILE : n) - 1;
	ts.load(S, cnt, base, last);
	for (unsigned i = threadIdx.x; i < AN_TILE; i += AN_THREADS) {
		const uint64_t k = base + i;
		if (k >= n)
			break;
		oid s, e;
		ts.locate(k, s, e);
		rb[k] = preceding ? s : e;
	}
}

// ROWS (gdk_analytic_bounds.c:177-198)
template <typename DT>
__global__ void __launch_bounds__(AN_THREADS)
an_rows_kernel(const oid *__restrict__ S, uint64_t cnt, uint64_t n, const DT *__restrict__ lim, DT slim, bool preceding,
	       oid second_half, oid *__restrict__ rb, unsigned long long *__restrict__ state)
{
	__shared__ TileSegs ts;
	const uint64_t base = (uint64_t) blockIdx.x * AN_TILE;
	const uint64_t last = (base + AN_TILE < n ? base + AN_TILE : n) - 1;
	ts.load(S, cnt, base, last);
	bool bad = false;
	for (unsigned i = threadIdx.x; i < AN_TILE; i += AN_THREADS) {
		const uint64_t k = base + i;
		if (k >= n)
			break;
		oid j, e;
		ts.locate(k, j, e);
		DT ol;
		if (!an_limit<DT, DT>(lim, slim, k, ol)) {
			bad = true;
			rb[k] = k;
			continue;
		}
		oid rl = an_limit_oid<DT>(ol);
		if (preceding) {
			rb[k] = rl > k - j ? j : k - rl + second_half;
		} else {
			rl += second_half;
			rb[k] = rl > e - k ? e : k + rl;
		}
	}
	if (bad)
		atomicOr(&state[0], (unsigned long long) AN_ERR_BOUND);
}

// GROUPS (gdk_analytic_bounds.c:217-254): G = positions of the set bits of the peer column
template <typename DT>
__global__ void __launch_bounds__(AN_THREADS)
an_groups_kernel(const oid *__restrict__ S, uint64_t cnt, const oid *__restrict__ G, uint64_t gcnt, uint64_t n,
		 const DT *__restrict__ lim, DT slim, bool preceding, oid *__restrict__ rb, unsigned long long *__restrict__ state)
{
	__shared__ TileSegs ts, tg;
	const uint64_t base = (uint64_t) blockIdx.x * AN_TILE;
	const uint64_t last = (base + AN_TILE < n ? base + AN_TILE : n) - 1;
	ts.load(S, cnt, base, last);
	tg.load(G, gcnt, base, last);
	bool bad = false;
	for (unsigned i = threadIdx.x; i < AN_TILE; i += AN_THREADS) {
		const uint64_t k = base + i;
		if (k >= n)
			break;
		oid m, e, gs, ge;
		ts.locate(k, m, e);
		const uint64_t c = tg.locate(k, gs, ge);	/* set bits at positions <= k */
		DT ol;
		if (!an_limit<DT, DT>(lim, slim, k, ol)) {
			bad = true;
			rb[k] = k;
			continue;
		}
		const oid rl = an_limit_oid<DT>(ol);
		if (preceding) {
			oid pos = m;
			if (c >= 1 && rl <= c - 1) {
				const oid g = G[c - 1 - rl];
				pos = g > m ? g : m;
			}
			rb[k] = pos;
		} else {
			oid pos = e;
			if (rl < gcnt - c) {
				const oid g = G[c + rl];
				pos = g < e ? g : e;
			}
			rb[k] = pos;
		}
	}
	if (bad)
		atomicOr(&state[0], (unsigned long long) AN_ERR_BOUND);
}

// RANGE (gdk_analytic_bounds.c:273-369): nearest position whose order value differs from
// b[k] by more than the limit; nils form their own peer run; the subtraction is overflow-checked
template <typename T, typename DT>
__global__ void __launch_bounds__(AN_THREADS)
an_range_kernel(const oid *__restrict__ S, uint64_t cnt, uint64_t n, const T *__restrict__ bp, bool nonil,
		const DT *__restrict__ lim, DT slim, bool preceding, oid *__restrict__ rb, unsigned long long *__restrict__ state)
{
	typedef typename std::conditional<Atom<T>::is_float, T, hge>::type LT;
	__shared__ TileSegs ts;
	const uint64_t base = (uint64_t) blockIdx.x * AN_TILE;
	const uint64_t last = (base + AN_TILE < n ? base + AN_TILE : n) - 1;
	ts.load(S, cnt, base, last);
	unsigned err = 0;
	for (unsigned i = threadIdx.x; i < AN_TILE; i += AN_THREADS) {
		const uint64_t k = base + i;
		if (k >= n)
			break;
		oid m, e;
		ts.locate(k, m, e);
		LT ol;
		if (!an_limit<LT, DT>(lim, (LT) slim, k, ol)) {
			err |= AN_ERR_BOUND;
			rb[k] = k;
			continue;
		}
		const T v = bp[k];
		const bool vnil = !nonil && Atom<T>::is_nil(v);
		oid j;
		if (preceding) {
			for (j = k;; j--) {
				const T w = bp[j];
				if (vnil) {
					if (!Atom<T>::is_nil(w)) {
						j++;
						break;
					}
				} else {
					if (!nonil && Atom<T>::is_nil(w)) {
						j++;
						break;
					}
					T calc;
					if (!sub_with_check<T>(v, w, calc)) {
						err |= AN_ERR_OVERFLOW;
						break;
					}
					if ((LT) (calc < 0 ? -calc : calc) > ol) {
						j++;
						break;
					}
				}
				if (j == m)
					break;
			}
		} else {
			for (j = k + 1; j < e; j++) {
				const T w = bp[j];
				if (vnil) {
					if (!Atom<T>::is_nil(w))
						break;
				} else {
					if (!nonil && Atom<T>::is_nil(w))
						break;
					T calc;
					if (!sub_with_check<T>(v, w, calc)) {
						err |= AN_ERR_OVERFLOW;
						break;
					}
					if ((LT) (calc < 0 ? -calc : calc) > ol)
						break;
				}
			}
		}
		rb[k] = j;
	}
	if (err)
		atomicOr(&state[0], (unsigned long long) err);
}

// RANGE on an order column that is ascending (raw order, nil = smallest) inside every
// partition — what ORDER BY produces: the reference's linear walk stops at the first row whose
// distance exceeds the limit, and on sorted data that row is found by a galloping + binary
// search from k (O(log frame) instead of O(frame) loads per row).  The overflow verdict is kept:
// the walk only ever subtracts rows up to and including the first failing one, so the checked
// subtraction can only fail there.
template <typename T>
__global__ void __launch_bounds__(256)
an_sorted_check_kernel(const T *__restrict__ bp, const bte *__restrict__ np, uint64_t n, unsigned long long *__restrict__ state)
{
	bool bad = false;
	for (uint64_t k = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x + 1; k < n; k += (uint64_t) gridDim.x * blockDim.x) {
		if (np && np[k])
			continue;
		const T a = bp[k - 1], b = bp[k];
		if constexpr (Atom<T>::is_float)
			bad |= !(a <= b);	/* NaN anywhere -> not usable */
		else
			bad |= b < a;
	}
	if (bad)
		atomicOr(&state[1], 1ULL);
}

template <typename T, typename DT>
__global__ void __launch_bounds__(AN_THREADS)
an_range_sorted_kernel(const oid *__restrict__ S, uint64_t cnt, uint64_t n, const T *__restrict__ bp, bool nonil,
		       const DT *__restrict__ lim, DT slim, bool preceding, oid *__restrict__ rb,
		       unsigned long long *__restrict__ state)
{
	typedef typename std::conditional<Atom<T>::is_float, T, hge>::type LT;
	constexpr unsigned RPT = AN_TILE / AN_THREADS;	/* consecutive rows per thread */
	__shared__ TileSegs ts;
	const uint64_t base = (uint64_t) blockIdx.x * AN_TILE;
	const uint64_t last = (base + AN_TILE < n ? base + AN_TILE : n) - 1;
	ts.load(S, cnt, base, last);
	unsigned err = 0;
	/* A thread takes RPT CONSECUTIVE rows: on an ascending partition with one static limit the frame edge of row
	 * k + 1 is never in front of row k's, so after the first row's galloping search the edge only has to be
	 * advanced — one or two probes per row instead of a dozen. */
	bool have_prev = false;
	oid pm = 0, pe = 0, pgood = 0;
	for (unsigned jrow = 0; jrow < RPT; jrow++) {
		const uint64_t k = base + (uint64_t) threadIdx.x * RPT + jrow;
		if (k >= n)
			break;
		oid m, e;
		ts.locate(k, m, e);
		LT ol;
		if (!an_limit<LT, DT>(lim, (LT) slim, k, ol)) {
			err |= AN_ERR_BOUND;
			rb[k] = k;
			have_prev = false;
			continue;
		}
		const T v = bp[k];
		const bool vnil = !nonil && Atom<T>::is_nil(v);
		bool olbig = false;
		unsigned long long ol64 = 0;
		if constexpr (!Atom<T>::is_float) {
			olbig = ol > (LT) 0xffffffffffffffffULL;	/* wider than any 64-bit distance */
			ol64 = (unsigned long long) ol;
		}
		/* ok(j): row j is still inside the frame of row k */
		auto ok = [&](oid j) -> bool {
			const T w = bp[j];
			if (vnil)
				return Atom<T>::is_nil(w);
			if (!nonil && Atom<T>::is_nil(w))
				return false;
			if constexpr (Atom<T>::is_float) {
				const T d = v - w;
				return (d < 0 ? -d : d) <= ol;
			} else {
				/* |v - w| of two values of at most 64 bits is exact in unsigned 64-bit arithmetic */
				const unsigned long long d = v >= w ? (unsigned long long) v - (unsigned long long) w
								    : (unsigned long long) w - (unsigned long long) v;
				return olbig || d <= ol64;
			}
		};
		oid good = k, bad = 0;
		bool have_bad = false;
		const bool hint = have_prev && lim == nullptr && !vnil && pm == m && pe == e;
		if (preceding) {
			if (hint) {
				/* the start can only move forward: first row from the previous start on that is inside */
				oid g = pgood;
				while (g < k && !ok(g))
					g++;
				good = g;
				if (good > m) {
					bad = good - 1;
					have_bad = true;
				}
			} else {
				for (oid step = 1;; step <<= 1) {
					if (good - m < step) {
						if (good != m) {
							if (ok(m))
								good = m;
							else {
								bad = m;
								have_bad = true;
							}
						}
						break;
					}
					const oid pos = good - step;
					if (ok(pos))
						good = pos;
					else {
						bad = pos;
						have_bad = true;
						break;
					}
				}
				if (have_bad)
					while (good - bad > 1) {
						const oid mid = bad + (good - bad) / 2;
						if (ok(mid))
							good = mid;
						else
							bad = mid;
					}
			}
			rb[k] = good;
		} else {
			if (hint) {
				/* the end can only move forward: extend from the previous last row (or from k) */
				oid g = pgood > k ? pgood : k;
				while (g + 1 < e && ok(g + 1))
					g++;
				good = g;
				if (good + 1 < e) {
					bad = good + 1;
					have_bad = true;
				}
			} else {
				for (oid step = 1;; step <<= 1) {
					if (e - 1 - good < step) {
						if (good != e - 1) {
							if (ok(e - 1))
								good = e - 1;
							else {
								bad = e - 1;
								have_bad = true;
							}
						}
						break;
					}
					const oid pos = good + step;
					if (ok(pos))
						good = pos;
					else {
						bad = pos;
						have_bad = true;
						break;
					}
				}
				if (have_bad)
					while (bad - good > 1) {
						const oid mid = good + (bad - good) / 2;
						if (ok(mid))
							good = mid;
						else
							bad = mid;
					}
			}
			rb[k] = good + 1;
		}
		if (have_bad && !vnil) {
			const T w = bp[bad];
			T calc;
			if (!(!nonil && Atom<T>::is_nil(w)) && !sub_with_check<T>(v, w, calc))
				err |= AN_ERR_OVERFLOW;
		}
		have_prev = !vnil;
		pm = m;
		pe = e;
		pgood = good;
	}
	if (err)
		atomicOr(&state[0], (unsigned long long) err);
}

// RANGE with ONE static non-negative integer limit on a nil-free order column that is ascending inside every partition
// (ORDER BY output, the usual case): the tile's order values plus a halo of AN_HALO rows on the side the frame
// extends to are staged in shared memory once, with coalesced loads, and every row finds its edge by a binary search in
// shared memory — uniform trip count across a warp, no dependent global loads.  A row whose edge lies beyond the halo
// (or beyond the partition part that is staged) continues with the galloping search in global memory.  The kernel
// also verifies the precondition on the rows it stages — b[k-1] <= b[k] wherever k is not a partition start — and
// raises state[1] when it does not hold; the caller then discards the result and runs the per-row walk, so the
// separate an_sorted_check pass (and its host round trip) is not needed on this path.
#define AN_HALO 1024
// continuation of the staged search in global memory: row `good` is inside the frame of the row with value v and
// lies at the edge of the staged rows; returns the outermost inside row of the partition [m, e)
template <typename T>
__device__ __noinline__ oid an_gallop_back(const T *__restrict__ bp, oid m, oid good, T v, unsigned long long ol64)
{
	auto inside = [&](oid j) -> bool { return (unsigned long long) v - (unsigned long long) bp[j] <= ol64; };
	oid bad = 0;
	bool have_bad = false;
	for (oid step = 1;; step <<= 1) {
		if (good - m < step) {
			if (good != m) {
				if (inside(m))
					good = m;
				else {
					bad = m;
					have_bad = true;
				}
			}
			break;
		}
		const oid pos = good - step;
		if (inside(pos))
			good = pos;
		else {
			bad = pos;
			have_bad = true;
			break;
		}
	}
	if (have_bad)
		while (good - bad > 1) {
			const oid mid = bad + (good - bad) / 2;
			if (inside(mid))
				good = mid;
			else
				bad = mid;
		}
	return good;
}
template <typename T>
__device__ __noinline__ oid an_gallop_fwd(const T *__restrict__ bp, oid e, oid good, T v, unsigned long long ol64)
{
	auto inside = [&](oid j) -> bool { return (unsigned long long) bp[j] - (unsigned long long) v <= ol64; };
	oid bad = 0;
	bool have_bad = false;
	for (oid step = 1;; step <<= 1) {
		if (e - 1 - good < step) {
			if (good != e - 1) {
				if (inside(e - 1))
					good = e - 1;
				else {
					bad = e - 1;
					have_bad = true;
				}
			}
			break;
		}
		const oid pos = good + step;
		if (inside(pos))
			good = pos;
		else {
			bad = pos;
			have_bad = true;
			break;
		}
	}
	if (have_bad)
		while (bad - good > 1) {
			const oid mid = good + (bad - good) / 2;
			if (inside(mid))
				good = mid;
			else
				bad = mid;
		}
	return good;
}

template <typename T>
__global__ void __launch_bounds__(AN_THREADS, 6)
an_range_staged_kernel(const oid *__restrict__ S, uint64_t cnt, uint64_t n, const T *__restrict__ bp, unsigned long long ol64,
		       bool preceding, oid *__restrict__ rb, unsigned long long *__restrict__ state)
{
	constexpr unsigned RPT = AN_TILE / AN_THREADS;	/* consecutive rows per thread */
	const unsigned long long TMAX = (unsigned long long) Atom<T>::maxval();
	__shared__ TileSegs ts;
	/* a thread works on 4 consecutive rows, so a warp reads the staged rows with a stride of 4 elements: one pad
	 * element after every 16 keeps those reads (and the unit-stride staging writes) free of bank conflicts */
	__shared__ T shp[AN_TILE + AN_HALO + 1 + (AN_TILE + AN_HALO + 1) / 16 + 1];
	auto sh = [&](unsigned i) -> T & { return shp[i + (i >> 4)]; };
	const uint64_t base = (uint64_t) blockIdx.x * AN_TILE;
	const uint64_t last = (base + AN_TILE < n ? base + AN_TILE : n) - 1;
	ts.load(S, cnt, base, last);
	/* staged rows [wlo, whi]; always includes row base - 1 for the order check of row base */
	const uint64_t wlo = preceding ? (base > AN_HALO ? base - AN_HALO : 0) : (base ? base - 1 : 0);
	const uint64_t whi = preceding ? last : (last + AN_HALO < n ? last + AN_HALO : n - 1);
	for (unsigned i = threadIdx.x; i <= (unsigned) (whi - wlo); i += AN_THREADS)
		sh(i) = bp[wlo + i];
	__syncthreads();
	bool overflow = false, unsorted = false;
	const uint64_t k0 = base + (uint64_t) threadIdx.x * RPT;
	if (k0 <= last) {
		/* On ascending data the edge of row k + 1 is never in front of row k's: the thread's first row searches
		 * (binary, 32-bit indices into the staged rows), the others only advance the edge.  With data that is
		 * not ascending everything below stays inside the staged rows / the partition and the result is void. */
		oid m, e, out[RPT];
		ts.locate(k0, m, e);
		unsigned g = 0;
		bool have = false;
#pragma unroll
		for (unsigned j = 0; j < RPT; j++) {
			const uint64_t k = k0 + j;
			out[j] = 0;
			if (k > last)
				break;
			if (k >= e) {
				ts.locate(k, m, e);
				have = false;
			}
			const unsigned ik = (unsigned) (k - wlo);
			const T v = sh(ik);
			if (k > m)
				unsorted |= sh(ik - 1) > v;
			if (preceding) {
				/* inside(w): v - w <= limit, exact in unsigned 64-bit arithmetic for w <= v */
				const unsigned ilo = m > wlo ? (unsigned) (m - wlo) : 0;
				unsigned a;
				if (have) {
					a = g;
					while (a < ik && (unsigned long long) v - (unsigned long long) sh(a) > ol64)
						a++;
				} else {
					a = ilo;
					unsigned b = ik;	/* smallest staged row of [ilo, ik] that is inside; row k is */
					while (a < b) {
						const unsigned mid = (a + b) >> 1;
						if ((unsigned long long) v - (unsigned long long) sh(mid) <= ol64)
							b = mid;
						else
							a = mid + 1;
					}
				}
				g = a;
				have = true;
				oid good = wlo + a;
				if (a == ilo && good > m)	/* the frame reaches in front of the staged rows */
					good = an_gallop_back<T>(bp, m, good, v, ol64);
				if (good > m) {
					/* the reference's walk subtracts the first row outside the frame too, with the
					 * overflow check: v - w must not exceed the type's maximum */
					const T w = good - 1 >= wlo ? sh(good - 1 - wlo) : bp[good - 1];
					overflow |= (unsigned long long) v - (unsigned long long) w > TMAX;
				}
				out[j] = good;
			} else {
				const oid hi = e - 1 < whi ? e - 1 : whi;
				const unsigned ihi = (unsigned) (hi - wlo);
				unsigned a;
				if (have) {
					a = g > ik ? g : ik;
					while (a < ihi && (unsigned long long) sh(a + 1) - (unsigned long long) v <= ol64)
						a++;
				} else {
					a = ik;
					unsigned b = ihi;	/* largest staged row of [ik, ihi] that is inside */
					while (a < b) {
						const unsigned mid = (a + b + 1) >> 1;
						if ((unsigned long long) sh(mid) - (unsigned long long) v <= ol64)
							a = mid;
						else
							b = mid - 1;
					}
				}
				g = a;
				have = true;
				oid good = wlo + a;
				if (a == ihi && hi < e - 1)
					good = an_gallop_fwd<T>(bp, e, good, v, ol64);
				if (good + 1 < e) {
					const T w = good + 1 <= whi ? sh(good + 1 - wlo) : bp[good + 1];
					overflow |= (unsigned long long) w - (unsigned long long) v > TMAX;
				}
				out[j] = good + 1;
			}
		}
		if (k0 + RPT - 1 <= last) {
			static_assert(RPT == 4, "two 16-byte stores per thread");
			ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(rb + k0);	/* k0 is a multiple of 4: 32-byte aligned */
			dst[0] = make_ulonglong2(out[0], out[1]);
			dst[1] = make_ulonglong2(out[2], out[3]);
		} else {
			for (unsigned j = 0; j < RPT && k0 + j <= last; j++)
				rb[k0 + j] = out[j];
		}
	}
	if (overflow)
		atomicOr(&state[0], (unsigned long long) AN_ERR_OVERFLOW);
	if (__any_sync(0xffffffffu, unsorted) && (threadIdx.x & 31) == 0)
		atomicOr(&state[1], 1ULL);
}

// ---------------------------------------------------------------- temporal RANGE frames
// gdk_analytic_bounds.c:458-611: the order column is a date / daytime / timestamp, the bound an interval in
// months (int) or milliseconds (lng).  Per row vmin = SUB(v, limit), vmax = ADD(v, limit) with the calendar
// arithmetic of gdk_time.c (a nil from an out-of-range result switches that side of the test off), then the
// frame extends from k while vmin <= b[j] <= vmax.  The calendar functions are restated from gdk_time.c:
// date = ((year + 4712) * 12 + month - 1) << 5 | day (:17-31), timestamp = date << 37 | usec of day (:42-50).
namespace mtime {
constexpr int YEAR_MIN = -4712, YEAR_OFFSET = 4712, MONTH_WIDTH = 21, DAY_WIDTH = 5;
constexpr int YEAR_MAX = YEAR_MIN + (1 << MONTH_WIDTH) / 12 - 1;
constexpr lng DAY_USEC = 24LL * 60 * 60 * 1000000;
__device__ __forceinline__ bool leap(int y) { return y % 4 == 0 && (y % 100 != 0 || y % 400 == 0); }
__device__ __forceinline__ int monthdays(int y, int m)
{
	const int l = (m == 2) ? 29 : ((m == 4 || m == 6 || m == 9 || m == 11) ? 30 : 31);
	return l - (m == 2 && !leap(y));
}
__device__ __forceinline__ int d_day(int dt) { return (int) ((unsigned) dt & 31u); }
__device__ __forceinline__ int d_month(int dt) { return (int) ((((unsigned) dt >> DAY_WIDTH) & ((1u << MONTH_WIDTH) - 1)) % 12 + 1); }
__device__ __forceinline__ int d_year(int dt) { return (int) ((((unsigned) dt >> DAY_WIDTH) & ((1u << MONTH_WIDTH) - 1)) / 12) - YEAR_OFFSET; }
__device__ __forceinline__ int mkdate(int y, int m, int d) { return (int) (((unsigned) ((y + YEAR_OFFSET) * 12 + m - 1) << DAY_WIDTH) | (unsigned) d); }

// date_add_month, gdk_time.c:155-183
__device__ inline int date_add_month(int dt, int months)
{
	if (Atom<int>::is_nil(dt) || Atom<int>::is_nil(months))
		return Atom<int>::nil();
	if ((months < 0 ? -(long long) months : (long long) months) >= (1LL << MONTH_WIDTH))
		return Atom<int>::nil();
	int y = d_year(dt), m = d_month(dt), d = d_day(dt);
	m += months;
	if (m <= 0) {
		y -= (12 - m) / 12;
		if (y < YEAR_MIN)
			return Atom<int>::nil();
		m = 12 - (-m % 12);
	} else if (m > 12) {
		y += (m - 1) / 12;
		if (y > YEAR_MAX)
			return Atom<int>::nil();
		m = (m - 1) % 12 + 1;
	}
	const int md = monthdays(y, m);
	if (d > md)
		d = md;
	return mkdate(y, m, d);
}

// days since 0000-03-01 of the proleptic Gregorian calendar and back (the closed forms of the month-by-month
// walk in date_add_day, gdk_time.c:122-153: same calendar, same results, O(1) instead of O(months))
__device__ inline long long days_from_civil(int y, int m, int d)
{
	y -= m <= 2;
	const long long era = (y >= 0 ? y : y - 399) / 400;
	const unsigned yoe = (unsigned) (y - era * 400);
	const unsigned doy = (153u * (unsigned) (m + (m > 2 ? -3 : 9)) + 2) / 5 + (unsigned) d - 1;
	const unsigned doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
	return era * 146097 + (long long) doe;
}
__device__ inline void civil_from_days(long long z, long long &y, int &m, int &d)
{
	const long long era = (z >= 0 ? z : z - 146096) / 146097;
	const unsigned doe = (unsigned) (z - era * 146097);
	const unsigned yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
	y = (long long) yoe + era * 400;
	const unsigned doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
	const unsigned mp = (5 * doy + 2) / 153;
	d = (int) (doy - (153 * mp + 2) / 5 + 1);
	m = (int) (mp < 10 ? mp + 3 : mp - 9);
	y += m <= 2;
}
__device__ inline int date_add_day(int dt, int days)
{
	if (Atom<int>::is_nil(dt) || Atom<int>::is_nil(days))
		return Atom<int>::nil();
	if ((days < 0 ? -(long long) days : (long long) days) >= (1LL << (DAY_WIDTH + MONTH_WIDTH)))
		return Atom<int>::nil();
	long long y;
	int m, d;
	civil_from_days(days_from_civil(d_year(dt), d_month(dt), d_day(dt)) + days, y, m, d);
	if (y < YEAR_MIN || y > YEAR_MAX)
		return Atom<int>::nil();
	return mkdate((int) y, m, d);
}
__device__ __forceinline__ lng ts_time(lng ts) { return (lng) ((unsigned long long) ts & ((1ULL << 37) - 1)); }
__device__ __forceinline__ int ts_date(lng ts) { return (int) (((unsigned long long) ts >> 37) & ((1u << (DAY_WIDTH + MONTH_WIDTH)) - 1)); }
__device__ __forceinline__ lng mktimestamp(int d, lng t) { return (lng) (((unsigned long long) (unsigned) d << 37) | (unsigned long long) t); }
// timestamp_add_usec, gdk_time.c:436-457
__device__ inline lng timestamp_add_usec(lng t, lng usec)
{
	if (Atom<lng>::is_nil(t) || Atom<lng>::is_nil(usec))
		return Atom<lng>::nil();
	lng tm = ts_time(t);
	int dt = ts_date(t);
	tm = (lng) ((unsigned long long) tm + (unsigned long long) usec);
	if (tm < 0) {
		const int add = (int) ((DAY_USEC - 1 - tm) / DAY_USEC);
		tm += add * DAY_USEC;
		dt = date_add_day(dt, -add);
	} else if (tm >= DAY_USEC) {
		dt = date_add_day(dt, (int) (tm / DAY_USEC));
		tm %= DAY_USEC;
	}
	if (Atom<int>::is_nil(dt))
		return Atom<lng>::nil();
	return mktimestamp(dt, tm);
}
__device__ inline lng timestamp_add_month(lng t, int m)
{
	if (Atom<lng>::is_nil(t) || Atom<int>::is_nil(m))
		return Atom<lng>::nil();
	const int dt = date_add_month(ts_date(t), m);
	if (Atom<int>::is_nil(dt))
		return Atom<lng>::nil();
	return mktimestamp(dt, ts_time(t));
}
__device__ inline lng daytime_add_usec(lng t, lng usec)
{
	if (Atom<lng>::is_nil(t) || Atom<lng>::is_nil(usec))
		return Atom<lng>::nil();
	if ((usec < 0 ? -usec : usec) >= DAY_USEC)
		return Atom<lng>::nil();
	t += usec;
	if (t < 0 || t >= DAY_USEC)
		return Atom<lng>::nil();
	return t;
}
enum { DATE_MONTH, TS_MONTH, DATE_MSEC, DAYTIME_MSEC, TS_MSEC };
// SUB / ADD of gdk_analytic_bounds.c:455-466 for each (order type, interval unit)
template <int KIND, typename T, typename DT> __device__ __forceinline__ void minmax(T v, DT lim, T &vmin, T &vmax)
{
	if constexpr (KIND == DATE_MONTH) {
		vmin = date_add_month(v, -(int) lim);
		vmax = date_add_month(v, (int) lim);
	} else if constexpr (KIND == TS_MONTH) {
		vmin = timestamp_add_month(v, -(int) lim);
		vmax = timestamp_add_month(v, (int) lim);
	} else if constexpr (KIND == DATE_MSEC) {
		vmin = date_add_day(v, (int) (-(lng) lim / (24LL * 60 * 60 * 1000)));
		vmax = date_add_day(v, (int) ((lng) lim / (24LL * 60 * 60 * 1000)));
	} else if constexpr (KIND == DAYTIME_MSEC) {
		vmin = daytime_add_usec(v, (lng) ((unsigned long long) -1000LL * (unsigned long long) (lng) lim));
		vmax = daytime_add_usec(v, (lng) (1000ULL * (unsigned long long) (lng) lim));
	} else {
		vmin = timestamp_add_usec(v, (lng) ((unsigned long long) -(lng) lim * 1000ULL));
		vmax = timestamp_add_usec(v, (lng) ((unsigned long long) (lng) lim * 1000ULL));
	}
}
}	// namespace mtime

template <typename T, typename DT, int KIND, bool SORTED>
__global__ void __launch_bounds__(AN_THREADS)
an_range_mtime_kernel(const oid *__restrict__ S, uint64_t cnt, uint64_t n, const T *__restrict__ bp, bool nonil,
		      const DT *__restrict__ lim, DT slim, bool preceding, oid *__restrict__ rb, unsigned long long *__restrict__ state)
{
	__shared__ TileSegs ts;
	const uint64_t base = (uint64_t) blockIdx.x * AN_TILE;
	const uint64_t last = (base + AN_TILE < n ? base + AN_TILE : n) - 1;
	ts.load(S, cnt, base, last);
	unsigned err = 0;
	for (unsigned i = threadIdx.x; i < AN_TILE; i += AN_THREADS) {
		const uint64_t k = base + i;
		if (k >= n)
			break;
		oid m, e;
		ts.locate(k, m, e);
		const DT rl = lim ? lim[k] : slim;
		if (Atom<DT>::is_nil(rl) || rl < 0) {
			err |= AN_ERR_BOUND;
			rb[k] = k;
			continue;
		}
		const T v = bp[k];
		const bool vnil = !nonil && Atom<T>::is_nil(v);
		T vmin = 0, vmax = 0;
		if (!vnil)
			mtime::minmax<KIND, T, DT>(v, rl, vmin, vmax);
		const bool usemin = !Atom<T>::is_nil(vmin), usemax = !Atom<T>::is_nil(vmax);
		/* in(j): row j still belongs to the frame of row k */
		auto in = [&](oid j) -> bool {
			const T w = bp[j];
			if (vnil)
				return Atom<T>::is_nil(w);
			if (!nonil && Atom<T>::is_nil(w))
				return false;
			return !((usemin && w < vmin) || (usemax && w > vmax));
		};
		if constexpr (!SORTED) {
			oid j;
			if (preceding) {
				for (j = k;; j--) {
					if (!in(j)) {
						j++;
						break;
					}
					if (j == m)
						break;
				}
			} else {
				for (j = k + 1; j < e; j++)
					if (!in(j))
						break;
			}
			rb[k] = j;
		} else {
			/* ascending inside the partition: membership is monotone along the walk -> gallop + bisect */
			oid good = k, bad = 0;
			bool have_bad = false;
			if (preceding) {
				for (oid step = 1;; step <<= 1) {
					if (good - m < step) {
						if (good != m) {
							if (in(m))
								good = m;
							else {
								bad = m;
								have_bad = true;
							}
						}
						break;
					}
					const oid pos = good - step;
					if (in(pos))
						good = pos;
					else {
						bad = pos;
						have_bad = true;
						break;
					}
				}
				if (have_bad)
					while (good - bad > 1) {
						const oid mid = bad + (good - bad) / 2;
						if (in(mid))
							good = mid;
						else
							bad = mid;
					}
				rb[k] = good;
			} else {
				for (oid step = 1;; step <<= 1) {
					if (e - 1 - good < step) {
						if (good != e - 1) {
							if (in(e - 1))
								good = e - 1;
							else {
								bad = e - 1;
								have_bad = true;
							}
						}
						break;
					}
					const oid pos = good + step;
					if (in(pos))
						good = pos;
					else {
						bad = pos;
						have_bad = true;
						break;
					}
				}
				if (have_bad)
					while (bad - good > 1) {
						const oid mid = good + (bad - good) / 2;
						if (in(mid))
							good = mid;
						else
							bad = mid;
					}
				rb[k] = good + 1;
			}
		}
	}
	if (err)
		atomicOr(&state[0], (unsigned long long) err);
}

// ---------------------------------------------------------------- host side
// peer-run starts: row 0, partition starts and every row whose order value differs
// from its predecessor (nil == nil), gdk_analytic_bounds.c:654-852
template <typename T> struct PeerStart {
	const bte *p;
	const T *b;
	__device__ __forceinline__ bool operator()(uint64_t idx, oid &ov) const
	{
		return idx == 0 || (p && p[idx]) || an_differs<T>(b[idx], b[idx - 1]);
	}
	__device__ __forceinline__ oid at(uint64_t idx) const { return idx; }
};

static inline unsigned an_tiles(uint64_t n)
{
	return (unsigned) ((n + AN_TILE - 1) / AN_TILE);
}

template <typename T, typename DT, int KIND>
static void range_mtime_launch(const SegList &sl, uint64_t n, const b200_bat *b, const void *lim, DT slim, bool preceding, bool sorted,
			       oid *rb, unsigned long long *state)
{
	KTimer kt("an_range_mtime", n * (9 + sizeof(T) + (lim ? sizeof(DT) : 0)));
	if (sorted)
		an_range_mtime_kernel<T, DT, KIND, true><<<an_tiles(n), AN_THREADS, 0, ctx()->stream>>>(
			sl.S, sl.cnt, n, (const T *) b->base, b->tnonil != 0, (const DT *) lim, slim, preceding, rb, state);
	else
		an_range_mtime_kernel<T, DT, KIND, false><<<an_tiles(n), AN_THREADS, 0, ctx()->stream>>>(
			sl.S, sl.cnt, n, (const T *) b->base, b->tnonil != 0, (const DT *) lim, slim, preceding, rb, state);
	COUNT_LAUNCH();
}

static bool segbound_run(const SegList &sl, uint64_t n, bool preceding, oid *rb)
{
	Ctx *c = ctx();
	if (n == 0)
		return true;
	KTimer kt("an_segbound", n * 9);
	an_segbound_kernel<<<an_tiles(n), AN_THREADS, 0, c->stream>>>(sl.S, sl.cnt, n, preceding, rb);
	COUNT_LAUNCH();
	return cuda_ok(cudaGetLastError(), "segbound launch");
}

template <typename T>
static bool peers_run(const b200_bat *b, const b200_bat *p, uint64_t n, bool preceding, oid *rb)
{
	SegList pl;
	PeerStart<T> f{p ? (const bte *) p->base : nullptr, (const T *) b->base};
	return build_seglist(n, f, pl) && segbound_run(pl, n, preceding, rb);
}

static bool peers_dispatch(const b200_bat *b, const b200_bat *p, uint64_t n, bool preceding, oid *rb)
{
	ctx()->algo = "windowbounds: peers, cuda segment list";
	switch (basetype(b->ttype)) {
	case B200_bte: return peers_run<bte>(b, p, n, preceding, rb);
	case B200_sht: return peers_run<sht>(b, p, n, preceding, rb);
	case B200_int: return peers_run<int>(b, p, n, preceding, rb);
	case B200_lng: return peers_run<lng>(b, p, n, preceding, rb);
	case B200_hge: return peers_run<hge>(b, p, n, preceding, rb);
	case B200_flt: return peers_run<flt>(b, p, n, preceding, rb);
	case B200_dbl: return peers_run<dbl>(b, p, n, preceding, rb);
	}
	set_err("windowbounds: type %s is not on the device path (fixed-width atoms only)\n", atomname(b->ttype));
	return false;
}

// value of a static integer bound as lng per the reference's casts (:936-968): nil -> lng_nil,
// hge above lng max -> lng max
static bool static_limit_lng(const void *bound, int tp2, lng &limit)
{
	switch (tp2) {
	case B200_bte: limit = Atom<bte>::is_nil(*(const bte *) bound) ? Atom<lng>::nil() : (lng) *(const bte *) bound; return true;
	case B200_sht: limit = Atom<sht>::is_nil(*(const sht *) bound) ? Atom<lng>::nil() : (lng) *(const sht *) bound; return true;
	case B200_int: limit = Atom<int>::is_nil(*(const int *) bound) ? Atom<lng>::nil() : (lng) *(const int *) bound; return true;
	case B200_lng: limit = *(const lng *) bound; return true;
	case B200_hge: {
		hge v;
		memcpy(&v, bound, 16);
		limit = Atom<hge>::is_nil(v) ? Atom<lng>::nil() : v > (hge) Atom<lng>::maxval() ? Atom<lng>::maxval() : (lng) v;
		return true;
	}
	}
	return false;
}

template <typename DT>
static void rows_launch(const SegList &sl, uint64_t n, const void *lim, DT slim, bool preceding, oid second_half, oid *rb,
			unsigned long long *state)
{
	KTimer kt("an_rows", n * (9 + (lim ? sizeof(DT) : 0)));
	an_rows_kernel<DT><<<an_tiles(n), AN_THREADS, 0, ctx()->stream>>>(sl.S, sl.cnt, n, (const DT *) lim, slim, preceding, second_half, rb, state);
	COUNT_LAUNCH();
}
template <typename DT>
static void groups_launch(const SegList &sl, const SegList &gl, uint64_t n, const void *lim, DT slim, bool preceding, oid *rb,
			  unsigned long long *state)
{
	KTimer kt("an_groups", n * (10 + (lim ? sizeof(DT) : 0)));
	an_groups_kernel<DT><<<an_tiles(n), AN_THREADS, 0, ctx()->stream>>>(sl.S, sl.cnt, gl.S, gl.cnt, n, (const DT *) lim, slim, preceding, rb, state);
	COUNT_LAUNCH();
}
static thread_local bool tl_range_sorted = false;

template <typename T, typename DT>
static void range_launch(const SegList &sl, uint64_t n, const b200_bat *b, const void *lim, DT slim, bool preceding, oid *rb,
			 unsigned long long *state)
{
	KTimer kt("an_range", n * (9 + sizeof(T) + (lim ? sizeof(DT) : 0)));
	if constexpr (sizeof(T) <= 8) {
		if (tl_range_sorted) {
			an_range_sorted_kernel<T, DT><<<an_tiles(n), AN_THREADS, 0, ctx()->stream>>>(sl.S, sl.cnt, n, (const T *) b->base, b->tnonil != 0,
												 (const DT *) lim, slim, preceding, rb, state);
			COUNT_LAUNCH();
			return;
		}
	}
	an_range_kernel<T, DT><<<an_tiles(n), AN_THREADS, 0, ctx()->stream>>>(sl.S, sl.cnt, n, (const T *) b->base, b->tnonil != 0,
									       (const DT *) lim, slim, preceding, rb, state);
	COUNT_LAUNCH();
}
// dynamic limit column type x order column type
template <typename T>
static bool range_dyn(int tp2, const SegList &sl, uint64_t n, const b200_bat *b, const void *lim, bool preceding, oid *rb,
		      unsigned long long *state)
{
	if constexpr (Atom<T>::is_float) {
		return false;
	} else {
		switch (tp2) {
		case B200_bte: if (sizeof(T) <= 8) { range_launch<T, bte>(sl, n, b, lim, 0, preceding, rb, state); return true; } break;
		case B200_sht: if (sizeof(T) <= 8) { range_launch<T, sht>(sl, n, b, lim, 0, preceding, rb, state); return true; } break;
		case B200_int: if (sizeof(T) <= 8) { range_launch<T, int>(sl, n, b, lim, 0, preceding, rb, state); return true; } break;
		case B200_lng: if (sizeof(T) <= 8) { range_launch<T, lng>(sl, n, b, lim, 0, preceding, rb, state); return true; } break;
		case B200_hge: range_launch<T, hge>(sl, n, b, lim, 0, preceding, rb, state); return true;
		}
		return false;
	}
}

static b200_return bounds_done(b200_bat *r, uint64_t n, bool nonil)
{
	r->count = n;
	r->tnonil = nonil;
	r->tnil = 0;
	r->tsorted = r->trevsorted = r->tkey = n <= 1;
	r->tseqbase = OID_NIL;
	return B200_SUCCEED;
}

extern "C" b200_return b200_GDKanalyticalwindowbounds(b200_bat *r, const b200_bat *b, const b200_bat *p, const b200_bat *l,
						      const void *bound, int tp1, int tp2, int unit, bool preceding,
						      b200_oid second_half)
{
	static const char *unitname[] = {"row", "range", "groups"};
	b200_clrerr();
	if (r == nullptr || b == nullptr || (l == nullptr) == (bound == nullptr) || unit < 0 || unit > 2) {
		set_err("GDKanalyticalwindowbounds: needs r, b and exactly one of l / bound; unit 0..2\n");
		return B200_FAIL;
	}
	Staged sb, sp, sl_;
	if (!stage(b, sb) || !stage(p, sp) || !stage(l, sl_))
		return B200_FAIL;
	b = sb.b;
	p = sp.b;
	l = sl_.b;
	const uint64_t n = b->count;
	if (r->device >= 0)
		shadow_drop(r->base);
	if (r->device < 0 || basetype(r->ttype) != B200_lng || r->capacity < n) {
		set_err("GDKanalyticalwindowbounds: result must be a device oid BAT with capacity >= BATcount(b)\n");
		return B200_FAIL;
	}
	if (!an_check_bit(p, "p", n))
		return B200_FAIL;
	oid *rb = (oid *) r->base;
	const int t2 = tp2;	/* bound types are plain numeric atoms */
	if (l != nullptr) {
		if (l->tnil) {
			set_err("42000!%s frame bound must be non negative and non null.\n", unitname[unit]);
			return B200_FAIL;
		}
		if (l->base == nullptr || l->count < n || atomsize(l->ttype) != atomsize(t2)) {
			set_err("GDKanalyticalwindowbounds: bound column must be a materialised %s column aligned with b\n", atomname(t2));
			return B200_FAIL;
		}
	}
	SegList S;
	if (!an_partitions(p, nullptr, n, S))
		return B200_FAIL;

	if (unit == 1) {
		/* ---- RANGE ---- */
		const bool temporal = tp1 == B200_date || tp1 == B200_daytime || tp1 == B200_timestamp;
		if (temporal && t2 != B200_int && t2 != B200_lng) {
			set_err("42000!range frame bound type %s not supported.\n", atomname(t2));	/* :1003-1004 */
			return B200_FAIL;
		}
		const int t1 = basetype(tp1);
		if (b->base == nullptr || atomsize(b->ttype) != atomsize(t1) || atomsize(t1) <= 0) {
			set_err("42000!type %s not supported for %s frame bound type.\n", atomname(tp1), atomname(t2));
			return B200_FAIL;
		}
		bool type_ok = false;
		switch (t2) {
		case B200_bte: case B200_sht: case B200_int: case B200_lng:
			type_ok = t1 == B200_bte || t1 == B200_sht || t1 == B200_int || t1 == B200_lng;
			break;
		case B200_hge:
			type_ok = t1 == B200_bte || t1 == B200_sht || t1 == B200_int || t1 == B200_lng || t1 == B200_hge;
			break;
		case B200_flt: type_ok = t1 == B200_flt; break;
		case B200_dbl: type_ok = t1 == B200_dbl; break;
		default:
			set_err("42000!range frame bound type %s not supported.\n", atomname(t2));
			return B200_FAIL;
		}
		hge ilim = 0;
		flt flim = 0;
		dbl dlim = 0;
		if (l == nullptr) {
			/* static bound: max -> UNBOUNDED, 0 -> peers, nil / negative -> error (:1094-1262) */
			bool unbounded = false, zero = false, invalid = false;
			switch (t2) {
			case B200_bte: { bte v = *(const bte *) bound; unbounded = v == Atom<bte>::maxval(); zero = v == 0; invalid = Atom<bte>::is_nil(v) || v < 0; ilim = v; break; }
			case B200_sht: { sht v = *(const sht *) bound; unbounded = v == Atom<sht>::maxval(); zero = v == 0; invalid = Atom<sht>::is_nil(v) || v < 0; ilim = v; break; }
			case B200_int: { int v = *(const int *) bound; unbounded = v == Atom<int>::maxval(); zero = v == 0; invalid = Atom<int>::is_nil(v) || v < 0; ilim = v; break; }
			case B200_lng: { lng v = *(const lng *) bound; unbounded = v == Atom<lng>::maxval(); zero = v == 0; invalid = Atom<lng>::is_nil(v) || v < 0; ilim = v; break; }
			case B200_hge: { hge v; memcpy(&v, bound, 16); invalid = Atom<hge>::is_nil(v) || v < 0; unbounded = !invalid && v == Atom<hge>::maxval(); zero = v == 0; ilim = v; break; }
			case B200_flt: { flt v = *(const flt *) bound; invalid = Atom<flt>::is_nil(v) || v < 0; unbounded = !invalid && v == Atom<flt>::maxval(); zero = v == 0; flim = v; break; }
			case B200_dbl: { dbl v = *(const dbl *) bound; invalid = Atom<dbl>::is_nil(v) || v < 0; unbounded = !invalid && v == Atom<dbl>::maxval(); zero = v == 0; dlim = v; break; }
			}
			if (unbounded) {
				ctx()->algo = "windowbounds: unbounded, cuda segment list";
				return segbound_run(S, n, preceding, rb) ? bounds_done(r, n, false) : B200_FAIL;
			}
			if (zero)
				return peers_dispatch(b, p, n, preceding, rb) ? bounds_done(r, n, false) : B200_FAIL;
			if (invalid) {
				set_err("42000!range frame bound must be non negative and non null.\n");
				return B200_FAIL;
			}
		}
		if (temporal)	/* month intervals: date / timestamp; msec intervals: date / daytime / timestamp (:578-611) */
			type_ok = t2 == B200_lng || tp1 != B200_daytime;
		if (!type_ok) {
			set_err("42000!type %s not supported for %s frame bound type.\n", atomname(tp1), atomname(t2));
			return B200_FAIL;
		}
		if (n == 0)
			return bounds_done(r, n, true);
		unsigned long long *state = an_state_init();
		if (state == nullptr)
			return B200_FAIL;
		/* ascending inside every partition?  then search instead of walking */
		tl_range_sorted = false;
		bool known_unsorted = false;
		if (!temporal && l == nullptr && b->tnonil && (t1 == B200_bte || t1 == B200_sht || t1 == B200_int || t1 == B200_lng)) {
			/* one static limit, no nils: optimistic staged search that checks the order itself */
			hge tmax = 0;
			switch (t1) {
			case B200_bte: tmax = Atom<bte>::maxval(); break;
			case B200_sht: tmax = Atom<sht>::maxval(); break;
			case B200_int: tmax = Atom<int>::maxval(); break;
			default: tmax = Atom<lng>::maxval(); break;
			}
			if (ilim <= tmax) {
				const unsigned long long ol64 = (unsigned long long) ilim;
				{
					KTimer kt("an_range", n * (9 + atomsize(t1)));
					switch (t1) {
					case B200_bte: an_range_staged_kernel<bte><<<an_tiles(n), AN_THREADS, 0, ctx()->stream>>>(S.S, S.cnt, n, (const bte *) b->base, ol64, preceding, rb, state); break;
					case B200_sht: an_range_staged_kernel<sht><<<an_tiles(n), AN_THREADS, 0, ctx()->stream>>>(S.S, S.cnt, n, (const sht *) b->base, ol64, preceding, rb, state); break;
					case B200_int: an_range_staged_kernel<int><<<an_tiles(n), AN_THREADS, 0, ctx()->stream>>>(S.S, S.cnt, n, (const int *) b->base, ol64, preceding, rb, state); break;
					default: an_range_staged_kernel<lng><<<an_tiles(n), AN_THREADS, 0, ctx()->stream>>>(S.S, S.cnt, n, (const lng *) b->base, ol64, preceding, rb, state); break;
					}
				}
				COUNT_LAUNCH();
				unsigned long long st[2];
				if (!an_read_state(state, st))
					return B200_FAIL;
				if (st[1] == 0) {
					ctx()->algo = "windowbounds: range, cuda shared-memory staged search (sorted partitions)";
					if (st[0] & AN_ERR_OVERFLOW) {
						set_err("22003!overflow in calculation.\n");
						return B200_FAIL;
					}
					return bounds_done(r, n, true);
				}
				/* not ascending: the result is void, walk row by row */
				known_unsorted = true;
				state = an_state_init();
				if (state == nullptr)
					return B200_FAIL;
			}
		}
		if (atomsize(t1) <= 8 && !known_unsorted) {
			const bte *np = p ? (const bte *) p->base : nullptr;
			const unsigned g = grid_for(n, 256, NSM * 16);
			switch (t1) {
			case B200_bte: an_sorted_check_kernel<bte><<<g, 256, 0, ctx()->stream>>>((const bte *) b->base, np, n, state); break;
			case B200_sht: an_sorted_check_kernel<sht><<<g, 256, 0, ctx()->stream>>>((const sht *) b->base, np, n, state); break;
			case B200_int: an_sorted_check_kernel<int><<<g, 256, 0, ctx()->stream>>>((const int *) b->base, np, n, state); break;
			case B200_lng: an_sorted_check_kernel<lng><<<g, 256, 0, ctx()->stream>>>((const lng *) b->base, np, n, state); break;
			case B200_flt: an_sorted_check_kernel<flt><<<g, 256, 0, ctx()->stream>>>((const flt *) b->base, np, n, state); break;
			case B200_dbl: an_sorted_check_kernel<dbl><<<g, 256, 0, ctx()->stream>>>((const dbl *) b->base, np, n, state); break;
			}
			COUNT_LAUNCH();
			Ctx *c = ctx();
			if (!cuda_ok(cudaMemcpyAsync(&c->slot_h->v[8], state + 1, 8, cudaMemcpyDeviceToHost, c->stream), "sorted flag") ||
			    !stream_sync() || !cuda_ok(cudaMemsetAsync(state + 1, 0, 8, c->stream), "state reset")) {
				dev_free(state);
				return B200_FAIL;
			}
			tl_range_sorted = c->slot_h->v[8] == 0;
		}
		ctx()->algo = tl_range_sorted ? "windowbounds: range, cuda galloping search (sorted partitions)"
					      : "windowbounds: range, cuda per-row walk";
		bool launched = true;
		if (temporal) {
			const void *lim = l ? l->base : nullptr;
			const bool so = tl_range_sorted;
			ctx()->algo = so ? "windowbounds: temporal range, cuda galloping search (sorted partitions)"
					 : "windowbounds: temporal range, cuda per-row walk";
			if (t2 == B200_int) {
				const int il = (int) ilim;
				if (tp1 == B200_date)
					range_mtime_launch<int, int, mtime::DATE_MONTH>(S, n, b, lim, il, preceding, so, rb, state);
				else
					range_mtime_launch<lng, int, mtime::TS_MONTH>(S, n, b, lim, il, preceding, so, rb, state);
			} else {
				const lng ll = (lng) ilim;
				if (tp1 == B200_date)
					range_mtime_launch<int, lng, mtime::DATE_MSEC>(S, n, b, lim, ll, preceding, so, rb, state);
				else if (tp1 == B200_daytime)
					range_mtime_launch<lng, lng, mtime::DAYTIME_MSEC>(S, n, b, lim, ll, preceding, so, rb, state);
				else
					range_mtime_launch<lng, lng, mtime::TS_MSEC>(S, n, b, lim, ll, preceding, so, rb, state);
			}
		} else if (l != nullptr) {
			const void *lim = l->base;
			switch (t1) {
			case B200_bte: launched = range_dyn<bte>(t2, S, n, b, lim, preceding, rb, state); break;
			case B200_sht: launched = range_dyn<sht>(t2, S, n, b, lim, preceding, rb, state); break;
			case B200_int: launched = range_dyn<int>(t2, S, n, b, lim, preceding, rb, state); break;
			case B200_lng: launched = range_dyn<lng>(t2, S, n, b, lim, preceding, rb, state); break;
			case B200_hge: launched = range_dyn<hge>(t2, S, n, b, lim, preceding, rb, state); break;
			case B200_flt: range_launch<flt, flt>(S, n, b, lim, 0, preceding, rb, state); break;
			case B200_dbl: range_launch<dbl, dbl>(S, n, b, lim, 0, preceding, rb, state); break;
			}
		} else {
			switch (t1) {
			case B200_bte: range_launch<bte, hge>(S, n, b, nullptr, ilim, preceding, rb, state); break;
			case B200_sht: range_launch<sht, hge>(S, n, b, nullptr, ilim, preceding, rb, state); break;
			case B200_int: range_launch<int, hge>(S, n, b, nullptr, ilim, preceding, rb, state); break;
			case B200_lng: range_launch<lng, hge>(S, n, b, nullptr, ilim, preceding, rb, state); break;
			case B200_hge: range_launch<hge, hge>(S, n, b, nullptr, ilim, preceding, rb, state); break;
			case B200_flt: range_launch<flt, flt>(S, n, b, nullptr, flim, preceding, rb, state); break;
			case B200_dbl: range_launch<dbl, dbl>(S, n, b, nullptr, dlim, preceding, rb, state); break;
			}
		}
		unsigned long long st[2];
		if (!an_read_state(state, st))
			return B200_FAIL;
		if (!launched) {
			set_err("42000!type %s not supported for %s frame bound type.\n", atomname(tp1), atomname(t2));
			return B200_FAIL;
		}
		if (st[0] & AN_ERR_BOUND) {
			set_err("42000!range frame bound must be non negative and non null.\n");
			return B200_FAIL;
		}
		if (st[0] & AN_ERR_OVERFLOW) {
			set_err("22003!overflow in calculation.\n");
			return B200_FAIL;
		}
		return bounds_done(r, n, true);
	}

	/* ---- ROWS (unit 0) and GROUPS (unit 2) ---- */
	if (t2 != B200_bte && t2 != B200_sht && t2 != B200_int && t2 != B200_lng && t2 != B200_hge) {
		set_err("42000!%s frame bound type %s not supported.\n", unit == 0 ? "rows" : "groups", atomname(t2));
		return B200_FAIL;
	}
	lng limit = 0;
	if (l == nullptr) {
		static_limit_lng(bound, t2, limit);
		if (limit == Atom<lng>::maxval()) {
			ctx()->algo = "windowbounds: unbounded, cuda segment list";
			return segbound_run(S, n, preceding, rb) ? bounds_done(r, n, false) : B200_FAIL;
		}
		if (Atom<lng>::is_nil(limit) || limit < 0) {
			set_err("42000!%s frame bound must be non negative and non null.\n", unitname[unit]);
			return B200_FAIL;
		}
	}
	if (n == 0)
		return bounds_done(r, n, true);
	SegList G;
	if (unit == 2) {
		if (basetype(b->ttype) != B200_bte || b->base == nullptr) {
			set_err("GDKanalyticalwindowbounds: GROUPS frames take the peer-start bit column as b\n");
			return B200_FAIL;
		}
		FlagBits fb{(const bte *) b->base};
		if (!build_seglist(n, fb, G))
			return B200_FAIL;
	}
	unsigned long long *state = an_state_init();
	if (state == nullptr)
		return B200_FAIL;
	ctx()->algo = unit == 0 ? "windowbounds: rows, cuda segment list" : "windowbounds: groups, cuda segment list";
	const void *lim = l ? l->base : nullptr;
	const int dt = l ? t2 : B200_lng;
#define AN_RG(DT, SL)                                                                          \
	do {                                                                                   \
		if (unit == 0)                                                                 \
			rows_launch<DT>(S, n, lim, (DT) (SL), preceding, second_half, rb, state);   \
		else                                                                           \
			groups_launch<DT>(S, G, n, lim, (DT) (SL), preceding, rb, state);        \
	} while (0)
	switch (dt) {
	case B200_bte: AN_RG(bte, 0); break;
	case B200_sht: AN_RG(sht, 0); break;
	case B200_int: AN_RG(int, 0); break;
	case B200_lng: AN_RG(lng, limit); break;
	case B200_hge: AN_RG(hge, 0); break;
	}
#undef AN_RG
	unsigned long long st[2];
	if (!an_read_state(state, st))
		return B200_FAIL;
	if (st[0] & AN_ERR_BOUND) {
		set_err("42000!%s frame bound must be non negative and non null.\n", unitname[unit]);
		return B200_FAIL;
	}
	return bounds_done(r, n, true);
}
