// aggr.cu — BATgroupsum / BATsum / BATgroupavg / BATgroupcount on the device
// (gdk_aggr.c:65-146 BATgroupaggrinit, :900-974 BATgroupsum, :1018-1183 BATsum,
// :1801-1985 BATgroupavg, :3069-3202 BATgroupcount).  Kernels in aggr.cuh.
#include "aggr.cuh"

// min / max of the non-nil group ids when the caller gives no extents BAT
// (the "complete scan" of BATgroupaggrinit, gdk_aggr.c:119-128)
__global__ void gid_minmax_kernel(const oid *__restrict__ gids, uint64_t n, u64 *__restrict__ out)
{
	u64 mn = ~0ULL, mx = 0;
	bool any = false;
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
		oid g = gids[i];
		if (g != OID_NIL) {
			mn = g < mn ? g : mn;
			mx = g > mx ? g : mx;
			any = true;
		}
	}
	if (any) {
		atomicMin(&out[0], mn);
		atomicMax(&out[1], mx);
	}
}

// largest biased exponent of the finite non-nil candidate values (fixed-point window)
template <typename T>
__global__ void fexp_kernel(AggrIn in, int *__restrict__ out)
{
	int best = 0;
	for (uint64_t k = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; k < in.n; k += (uint64_t) gridDim.x * blockDim.x) {
		T x;
		oid gid;
		if (!aggr_fetch<T>(in, k, x, gid) || Atom<T>::is_nil(x) || isinf(x))
			continue;
		int e = (int) (((u64) __double_as_longlong((double) x) >> 52) & 0x7ff);
		best = e > best ? e : best;
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		int t = __shfl_down_sync(0xffffffffu, best, o);
		best = t > best ? t : best;
	}
	if ((threadIdx.x & 31) == 0 && best)
		atomicMax(out, best);
}

// BATgroupaggrinit.  g == NULL: one group (BATsum).  Returns false with the error set.
bool aggr_init(AggrSetup &a, const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s)
{
	if (b == nullptr) {
		set_err("b must exist\n");
		return false;
	}
	if (!stage(b, a.sb) || !stage(g, a.sg) || !stage(e, a.se) || !stage(s, a.ss))
		return false;
	a.b = a.sb.b;
	a.g = a.sg.b;
	a.e = a.se.b;
	if (!cand_init(a.ci, a.b, a.ss.b))
		return false;
	if (a.g) {
		if (a.ci.ncand != a.g->count || (a.ci.ncand != 0 && a.ci.seq != a.g->hseqbase)) {
			set_err("b with s and g must be aligned\n");
			return false;
		}
	}
	if (a.g == nullptr) {
		a.min = a.max = 0;
		a.ngrp = 1;
	} else if (a.e == nullptr) {
		oid mn = OID_NIL, mx = 0;
		if (bat_is_dense(a.g)) {
			mn = a.g->tseqbase;
			mx = a.g->tseqbase + a.g->count - 1;
		} else if (a.g->count > 0 && a.g->ttype != B200_void) {
			Ctx *c = ctx();
			u64 *st = (u64 *) scratch_get(64);
			static const u64 init[2] = {~0ULL, 0};
			if (st == nullptr || !cuda_ok(cudaMemcpyAsync(st, init, 16, cudaMemcpyHostToDevice, c->stream), "init"))
				return false;
			gid_minmax_kernel<<<grid_for(a.g->count, 256, NSM * 8), 256, 0, c->stream>>>((const oid *) a.g->base, a.g->count, st);
			COUNT_LAUNCH();
			if (!cuda_ok(cudaMemcpyAsync(c->slot_h->v, st, 16, cudaMemcpyDeviceToHost, c->stream), "read") || !stream_sync())
				return false;
			if (c->slot_h->v[0] != ~0ULL) {
				mn = c->slot_h->v[0];
				mx = c->slot_h->v[1];
			}
		}
		a.min = mn;
		a.max = mx;
		a.ngrp = mx < mn ? 0 : mx - mn + 1;
	} else {
		a.ngrp = a.e->count;
		a.min = a.e->hseqbase;
		a.max = a.e->hseqbase + a.ngrp - 1;
	}
	return true;
}

struct AggrAccum {
	u64 *cells = nullptr;
	u64 *state = nullptr;		/* device: [0] flags [1] nil groups [2] exact-check groups [3] nil rows seen */
	int nd = 0, ncell = 0, fexp = 0;
	bool check = false, niltrack = false;
	int pending_bits = 0;		/* fast form launched without waiting: the caller checks state[5..7] against this width */
	bool pending_narrow = false;
	AggrIn in;
	bool state_in_cells = false;	/* state is the tail of the cells allocation */
	~AggrAccum() { release(); }
	void release()
	{
		if (cells) dev_free(cells);
		if (state && !state_in_cells) dev_free(state);
		cells = state = nullptr;
		state_in_cells = false;
	}
};

// fast register accumulation (aggr_reg_fast_kernel).  On return `done` tells whether acc holds the
// result; it is false when an overflow check was asked for and rows x max|x| does not fit
// check_bits, in which case acc has been released for the exact form.
template <typename T1>
static bool aggr_fast(const AggrSetup &a, int check_bits, AggrAccum &acc, bool &done, bool allow_narrow = true, bool defer = false)
{
	Ctx *c = ctx();
	constexpr int ND = Digits<T1>::ND;
	/* hge values that all fit 64 bits take the two-word accumulators; the host learns whether they
	 * did from the readback it needs anyway for the overflow bound */
	const bool narrow = sizeof(T1) == 16 && check_bits != 0 && allow_narrow;
	const unsigned ngrp = (unsigned) a.ngrp;
	acc.check = false;
	const AggrLayout L(ND, false, false);
	acc.ncell = L.ncell;
	const size_t bytes = (size_t) L.ncell * ngrp * 8;
	/* cells and state in one allocation, cleared by one memset */
	acc.cells = (u64 *) dev_alloc(bytes + 64);
	if (acc.cells == nullptr || !cuda_ok(cudaMemsetAsync(acc.cells, 0, bytes + 64, c->stream), "memset cells"))
		return false;
	acc.state = acc.cells + bytes / 8;
	acc.state_in_cells = true;
	AggrIn &in = acc.in;
	memset(&in, 0, sizeof(in));
	in.n = a.ci.ncand;
	in.hseq = a.ci.hseq;
	in.vals = (const T1 *) a.b->base + (a.ci.seq - a.ci.hseq);
	in.gmin = a.min;
	in.gmax = a.max;
	in.gconst = a.g == nullptr;
	in.gids = a.g ? (const oid *) a.g->base : nullptr;
	Shadow gsh;
	if (a.g && ((uintptr_t) in.vals & 15) == 0 && shadow_get(a.g->base, SHADOW_GID8, a.g->count, gsh))
		in.g8 = (const unsigned char *) gsh.p;
	in.nonil = a.b->tnonil;
	in.skip_nils = true;
	const unsigned grid = grid_for(in.n, AGGR_THREADS * 8, NSM * 8);
	{
		KTimer kt("aggr", in.n * (sizeof(T1) + (in.gids ? 8 : 0)));	/* algorithmic: the API's 8-byte ids */
		const T1 *vals = (const T1 *) in.vals;
		u64 *mx = acc.state + 5;
#define FAST_LAUNCH(NG, HG, NRW, GT, GP) \
	aggr_reg_fast_kernel<T1, NG, HG, NRW, GT><<<grid, AGGR_THREADS, 0, c->stream>>>(vals, GP, in.gmin, in.gmax - in.gmin, in.n, in.nonil, acc.cells, ngrp, mx)
#define FAST_GID(NG, NRW)                                                                    \
	do {                                                                                 \
		if (in.g8)                                                                   \
			FAST_LAUNCH(NG, true, NRW, unsigned char, in.g8);                    \
		else if (in.gids)                                                            \
			FAST_LAUNCH(NG, true, NRW, oid, in.gids);                            \
		else                                                                         \
			FAST_LAUNCH(NG, false, NRW, oid, (const oid *) nullptr);             \
	} while (0)
#define FAST_CASE(NG)                                                                        \
	do {                                                                                 \
		if constexpr (sizeof(T1) == 16) {                                            \
			if (narrow) {                                                        \
				FAST_GID(NG, true);                                          \
				break;                                                       \
			}                                                                    \
		}                                                                            \
		FAST_GID(NG, false);                                                         \
	} while (0)
		if (ngrp <= 1) FAST_CASE(1);
		else if (ngrp <= 2) FAST_CASE(2);
		else if (ngrp <= 4) FAST_CASE(4);
		else FAST_CASE(8);
#undef FAST_GID
#undef FAST_LAUNCH
#undef FAST_CASE
	}
	COUNT_LAUNCH();
	ctx()->algo = "aggr: cuda wide register accumulators";
	if (!cuda_ok(cudaGetLastError(), "aggr launch"))
		return false;
	done = true;
	if (check_bits != 0 && defer) {
		acc.pending_bits = check_bits;
		acc.pending_narrow = narrow;
		return true;
	}
	if (check_bits != 0) {
		if (!cuda_ok(cudaMemcpyAsync(&c->slot_h->v[5], acc.state + 5, 24, cudaMemcpyDeviceToHost, c->stream), "maxabs") || !stream_sync())
			return false;
		if (narrow && c->slot_h->v[7] != 0) {
			/* a value needs more than 64 bits: run the three-word form instead */
			acc.release();
			return aggr_fast<T1>(a, check_bits, acc, done, false);
		}
		const uhge mx = ((uhge) c->slot_h->v[6] << 64) | c->slot_h->v[5];
		const uhge lim = ((uhge) 1 << (check_bits - 1)) - 1;
		if (mx != 0 && (uhge) in.n > lim / mx) {
			acc.release();
			done = false;
		}
	}
	return true;
}

template <typename T1>
static bool aggr_accumulate_typed(const AggrSetup &a, bool skip_nils, int check_bits, bool defer, AggrAccum &acc)
{
	Ctx *c = ctx();
	constexpr int ND = Digits<T1>::ND;
	const unsigned ngrp = (unsigned) a.ngrp;
	bool check = check_bits != 0;
	acc.nd = ND;
	acc.niltrack = !skip_nils && !a.b->tnonil;
	/* fast register form: integers, dense values, few groups; the overflow check becomes the
	 * rows x max|x| bound, with the exact digit form as the fallback */
	bool fast = false;
	if constexpr (!Atom<T1>::is_float)
		fast = !acc.niltrack && ngrp <= 8 && a.ci.dense && (a.g == nullptr || a.g->ttype != B200_void);
	if (fast) {
		if (!aggr_fast<T1>(a, check_bits, acc, fast, true, defer))
			return false;
		if (fast)
			return true;
		/* bound failed: acc was reset, run the exact form */
	}
	acc.check = check;
	const AggrLayout L(ND, check, acc.niltrack);
	acc.ncell = L.ncell;
	const size_t bytes = (size_t) L.ncell * ngrp * 8;
	acc.cells = (u64 *) dev_alloc(bytes);
	acc.state = (u64 *) dev_alloc(64);
	if (acc.cells == nullptr || acc.state == nullptr ||
	    !cuda_ok(cudaMemsetAsync(acc.cells, 0, bytes, c->stream), "memset cells") ||
	    !cuda_ok(cudaMemsetAsync(acc.state, 0, 64, c->stream), "memset state"))
		return false;
	if (acc.niltrack &&
	    !cuda_ok(cudaMemsetAsync(acc.cells + (size_t) L.c_first * ngrp, 0xff, (size_t) ngrp * 8, c->stream), "memset first"))
		return false;
	AggrIn &in = acc.in;
	memset(&in, 0, sizeof(in));
	in.n = a.ci.ncand;
	in.hseq = a.ci.hseq;
	if (a.ci.dense) {
		in.vals = (const T1 *) a.b->base + (a.ci.seq - a.ci.hseq);
		in.cand = nullptr;
	} else {
		in.vals = a.b->base;
		in.cand = a.ci.oids;
	}
	in.gmin = a.min;
	in.gmax = a.max;
	if (a.g == nullptr) {
		in.gconst = true;
	} else if (a.g->ttype == B200_void) {
		in.gdense = a.g->tseqbase;
	} else {
		in.gids = (const oid *) a.g->base;
		Shadow gsh;
		if (shadow_get(a.g->base, SHADOW_GID8, a.g->count, gsh))
			in.g8 = (const unsigned char *) gsh.p;
	}
	in.nonil = a.b->tnonil;
	in.skip_nils = skip_nils;
	in.check = check;
	unsigned *flags = (unsigned *) acc.state;
	if constexpr (Atom<T1>::is_float) {
		int *fe = (int *) (acc.state + 4);
		fexp_kernel<T1><<<grid_for(in.n, 256, NSM * 8), 256, 0, c->stream>>>(in, fe);
		COUNT_LAUNCH();
		if (!cuda_ok(cudaMemcpyAsync(&c->slot_h->v[4], fe, 4, cudaMemcpyDeviceToHost, c->stream), "fexp") || !stream_sync())
			return false;
		int be = (int) (unsigned) c->slot_h->v[4];
		acc.fexp = in.fexp = (be ? be : 1) - 1023;
	}
	const unsigned grid = grid_for(in.n, AGGR_THREADS * 4, NSM * 8);
	KTimer kt("aggr", in.n * (sizeof(T1) + (in.gids ? 8 : 0) + (in.cand ? 8 : 0)));
	if (!acc.niltrack && ngrp <= 8) {
#define REG_CASE(NG)                                                                                       \
	if (check)                                                                                         \
		aggr_reg_kernel<T1, NG, true><<<grid, AGGR_THREADS, 0, c->stream>>>(in, acc.cells, ngrp, flags); \
	else                                                                                               \
		aggr_reg_kernel<T1, NG, false><<<grid, AGGR_THREADS, 0, c->stream>>>(in, acc.cells, ngrp, flags);
		if constexpr (Atom<T1>::is_float) {
			/* no overflow-check variant for floating point */
			if (ngrp <= 1) aggr_reg_kernel<T1, 1, false><<<grid, AGGR_THREADS, 0, c->stream>>>(in, acc.cells, ngrp, flags);
			else if (ngrp <= 2) aggr_reg_kernel<T1, 2, false><<<grid, AGGR_THREADS, 0, c->stream>>>(in, acc.cells, ngrp, flags);
			else if (ngrp <= 4) aggr_reg_kernel<T1, 4, false><<<grid, AGGR_THREADS, 0, c->stream>>>(in, acc.cells, ngrp, flags);
			else aggr_reg_kernel<T1, 8, false><<<grid, AGGR_THREADS, 0, c->stream>>>(in, acc.cells, ngrp, flags);
		} else {
			if (ngrp <= 1) { REG_CASE(1) }
			else if (ngrp <= 2) { REG_CASE(2) }
			else if (ngrp <= 4) { REG_CASE(4) }
			else { REG_CASE(8) }
		}
#undef REG_CASE
		ctx()->algo = "aggr: cuda register accumulators";
	} else if (bytes <= 48 * 1024 && !acc.niltrack) {
		aggr_tab32_kernel<T1><<<grid, AGGR_THREADS, bytes, c->stream>>>(in, acc.cells, ngrp, flags);
		ctx()->algo = "aggr: cuda shared-memory table, 32-bit halves";
	} else if (bytes <= 48 * 1024) {
		aggr_tab_kernel<T1, true><<<grid, AGGR_THREADS, bytes, c->stream>>>(in, acc.cells, ngrp, acc.niltrack, flags);
		ctx()->algo = "aggr: cuda shared-memory table";
	} else {
		aggr_tab_kernel<T1, false><<<grid, AGGR_THREADS, 0, c->stream>>>(in, acc.cells, ngrp, acc.niltrack, flags);
		ctx()->algo = "aggr: cuda global table";
	}
	COUNT_LAUNCH();
	return cuda_ok(cudaGetLastError(), "aggr launch");
}

static bool aggr_accumulate(const AggrSetup &a, bool skip_nils, int check_bits, AggrAccum &acc, bool defer = false)
{
	if (a.ci.ncand >= (1ULL << 32)) {
		set_err("b200: aggregates over 2^32 or more rows in one call are not supported\n");
		return false;
	}
	switch (basetype(a.b->ttype)) {
	case B200_bte: return aggr_accumulate_typed<bte>(a, skip_nils, check_bits, defer, acc);
	case B200_sht: return aggr_accumulate_typed<sht>(a, skip_nils, check_bits, defer, acc);
	case B200_int: return aggr_accumulate_typed<int>(a, skip_nils, check_bits, defer, acc);
	case B200_lng: return aggr_accumulate_typed<lng>(a, skip_nils, check_bits, defer, acc);
	case B200_hge: return aggr_accumulate_typed<hge>(a, skip_nils, check_bits, defer, acc);
	case B200_flt: return aggr_accumulate_typed<flt>(a, skip_nils, 0, false, acc);
	case B200_dbl: return aggr_accumulate_typed<dbl>(a, skip_nils, 0, false, acc);
	}
	set_err("b200: aggregate over type %s is not on the device path\n", atomname(a.b->ttype));
	return false;
}

// ---------------------------------------------------------------- sum
static int int_rank(int t)
{
	switch (t) {
	case B200_bte: return 0;
	case B200_sht: return 1;
	case B200_int: return 2;
	case B200_lng: return 3;
	case B200_hge: return 4;
	}
	return -1;
}

template <typename T2>
static void launch_finish_sum(const AggrAccum &acc, unsigned ngrp, bool nil_if_empty, bool float_in, void *out, unsigned *exactlist)
{
	aggr_finish_sum_kernel<T2><<<(ngrp + 127) / 128, 128, 0, ctx()->stream>>>(acc.cells, ngrp, acc.nd, acc.check, acc.niltrack,
										  nil_if_empty, float_in, acc.fexp, (T2 *) out,
										  acc.state, exactlist);
	COUNT_LAUNCH();
}

// dosum (gdk_aggr.c:708-896): results (device, ngrp x tp2) must be allocated; returns
// the number of nil results, or BUN_NONE with the error set
static uint64_t do_sum(const AggrSetup &a, int tp2, bool skip_nils, bool nil_if_empty, void *results, const char *func, bool defer = true)
{
	Ctx *c = ctx();
	const int t1 = basetype(a.b->ttype), t2 = basetype(tp2);
	const int r1 = int_rank(t1), r2 = int_rank(t2);
	bool ok = false, float_in = false, check = false;
	if (r2 >= 0) {
		ok = r1 >= 0 && r1 <= r2;
		if (ok) {
			int diff = atomsize(t2) - atomsize(t1);
			bool noovl = diff >= 8 || a.ci.ncand < (1ULL << (8 * diff));
			check = !noovl;
		}
	} else if (t2 == B200_flt) {
		ok = float_in = t1 == B200_flt;
	} else if (t2 == B200_dbl) {
		ok = float_in = t1 == B200_flt || t1 == B200_dbl;
	}
	if (!ok) {
		set_err("%s: type combination (sum(%s)->%s) not supported.\n", func, atomname(a.b->ttype), atomname(tp2));
		return B200_BUN_NONE;
	}
	AggrAccum acc;
	if (!aggr_accumulate(a, skip_nils, check ? atomsize(t2) * 8 : 0, acc, defer))
		return B200_BUN_NONE;
	check = acc.check;		/* the fast form proves "no overflow possible" with a bound instead */
	const unsigned ngrp = (unsigned) a.ngrp;
	unsigned *exactlist = nullptr;
	if (check) {
		exactlist = (unsigned *) dev_alloc((size_t) ngrp * 4);
		if (exactlist == nullptr)
			return B200_BUN_NONE;
	}
	switch (t2) {
	case B200_bte: launch_finish_sum<bte>(acc, ngrp, nil_if_empty, float_in, results, exactlist); break;
	case B200_sht: launch_finish_sum<sht>(acc, ngrp, nil_if_empty, float_in, results, exactlist); break;
	case B200_int: launch_finish_sum<int>(acc, ngrp, nil_if_empty, float_in, results, exactlist); break;
	case B200_lng: launch_finish_sum<lng>(acc, ngrp, nil_if_empty, float_in, results, exactlist); break;
	case B200_hge: launch_finish_sum<hge>(acc, ngrp, nil_if_empty, float_in, results, exactlist); break;
	case B200_flt: launch_finish_sum<flt>(acc, ngrp, nil_if_empty, float_in, results, exactlist); break;
	case B200_dbl: launch_finish_sum<dbl>(acc, ngrp, nil_if_empty, float_in, results, exactlist); break;
	}
	uint64_t ret = B200_BUN_NONE;
	if (!cuda_ok(cudaGetLastError(), "finish launch") ||
	    !cuda_ok(cudaMemcpyAsync(c->slot_h->v, acc.state, 64, cudaMemcpyDeviceToHost, c->stream), "state read") ||
	    !stream_sync())
		goto done;
	if (acc.pending_bits != 0) {
		/* the fast form ran unchecked: its rows x max|x| bound (and, for the two-word hge form,
		 * "every value fits 64 bits") is verified here, after the one sync this call needs anyway */
		const uhge mx = ((uhge) c->slot_h->v[6] << 64) | c->slot_h->v[5];
		const uhge lim = ((uhge) 1 << (acc.pending_bits - 1)) - 1;
		if ((acc.pending_narrow && c->slot_h->v[7] != 0) || (mx != 0 && (uhge) acc.in.n > lim / mx)) {
			if (exactlist)
				dev_free(exactlist);
			return do_sum(a, tp2, skip_nils, nil_if_empty, results, func, false);
		}
	}
	if (((unsigned) c->slot_h->v[0] & (AGGR_F_SPECIAL | AGGR_F_OVERFLOW)) == 0 && c->slot_h->v[2] != 0) {
		/* some group's positive or negative part alone exceeds the result type:
		 * decide exactly whether the running sum overflows */
		const unsigned nex = (unsigned) c->slot_h->v[2];
		const int bits = atomsize(t2) * 8;
		switch (t1) {
		case B200_bte: aggr_exact_check_kernel<bte><<<nex, 256, 0, c->stream>>>(acc.in, exactlist, bits, acc.state); break;
		case B200_sht: aggr_exact_check_kernel<sht><<<nex, 256, 0, c->stream>>>(acc.in, exactlist, bits, acc.state); break;
		case B200_int: aggr_exact_check_kernel<int><<<nex, 256, 0, c->stream>>>(acc.in, exactlist, bits, acc.state); break;
		case B200_lng: aggr_exact_check_kernel<lng><<<nex, 256, 0, c->stream>>>(acc.in, exactlist, bits, acc.state); break;
		case B200_hge: aggr_exact_check_kernel<hge><<<nex, 256, 0, c->stream>>>(acc.in, exactlist, bits, acc.state); break;
		}
		COUNT_LAUNCH();
		if (!cuda_ok(cudaGetLastError(), "exact check launch") ||
		    !cuda_ok(cudaMemcpyAsync(c->slot_h->v, acc.state, 32, cudaMemcpyDeviceToHost, c->stream), "state read") ||
		    !stream_sync())
			goto done;
	}
	if ((unsigned) c->slot_h->v[0] & (AGGR_F_SPECIAL | AGGR_F_OVERFLOW)) {
		set_err("22003!overflow in sum aggregate.\n");
		goto done;
	}
	ret = c->slot_h->v[1];
	if (acc.niltrack && c->slot_h->v[3] != 0 && ret == 0)
		ret = 1;	/* the reference reports the nils it met even when a later value reset the group */
done:
	if (exactlist)
		dev_free(exactlist);
	return ret;
}

static void agg_props(b200_bat *bn, uint64_t nils)
{
	bn->tkey = bn->tsorted = bn->trevsorted = bn->count <= 1;
	bn->tnil = nils != 0;
	bn->tnonil = nils == 0;
}

static bool singleton_groups(const AggrSetup &a)
{
	/* every candidate row is its own group (gdk_aggr.c:929-935) */
	return (a.e == nullptr || (a.e->count == a.ci.ncand && a.e->hseqbase == a.ci.rhseq)) &&
	       (bat_is_dense(a.g) || (a.g->tkey && a.g->tnonil));
}

extern "C" b200_bat *b200_BATgroupsum(const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s, int tp, bool skip_nils)
{
	b200_clrerr();
	AggrSetup a;
	if (!aggr_init(a, b, g, e, s))
		return nullptr;
	if (a.g == nullptr) {
		set_err("b and g must be aligned\n");
		return nullptr;
	}
	if (a.ci.ncand == 0 || a.ngrp == 0)
		return bat_constant(a.ngrp == 0 ? 0 : a.min, tp, nullptr, a.ngrp);
	if (singleton_groups(a))
		return b200_BATconvert(b, s, tp, 0, 0, 0);
	b200_bat *bn = bat_new(a.min, tp, a.ngrp);
	if (bn == nullptr)
		return nullptr;
	uint64_t nils = do_sum(a, tp, skip_nils, true, bn->base, "BATgroupsum");
	if (nils == B200_BUN_NONE) {
		b200_BATfree(bn);
		return nullptr;
	}
	bn->count = a.ngrp;
	agg_props(bn, nils);
	return bn;
}

extern "C" b200_return b200_BATsum(void *res, int tp, const b200_bat *b, const b200_bat *s, bool skip_nils, bool nil_if_empty)
{
	b200_clrerr();
	AggrSetup a;
	if (!aggr_init(a, b, nullptr, nullptr, s))
		return B200_FAIL;
	const int t1 = basetype(a.b->ttype), t2 = basetype(tp);
	const int w2 = atomsize(t2);
	if (w2 <= 0 || t2 == B200_void) {
		set_err("type combination (sum(%s)->%s) not supported.\n", atomname(a.b->ttype), atomname(tp));
		return B200_FAIL;
	}
	Ctx *c = ctx();
	if ((t2 == B200_flt || t2 == B200_dbl) && int_rank(t1) >= 0) {
		/* integer -> floating point: exact average times count (gdk_aggr.c:1112-1154) */
		dbl avg = NAN;
		uint64_t cnt = 0;
		if (a.ci.ncand > 0) {
			AggrAccum acc;
			if (!aggr_accumulate(a, true, 0, acc))
				return B200_FAIL;
			dbl *out = (dbl *) dev_alloc(16);
			if (out == nullptr)
				return B200_FAIL;
			lng *cn = (lng *) (out + 1);
			switch (t1) {
			case B200_bte: aggr_finish_avg_kernel<bte><<<1, 32, 0, c->stream>>>(acc.cells, 1, false, 0, 1.0, out, cn, acc.state, true); break;
			case B200_sht: aggr_finish_avg_kernel<sht><<<1, 32, 0, c->stream>>>(acc.cells, 1, false, 0, 1.0, out, cn, acc.state, true); break;
			case B200_int: aggr_finish_avg_kernel<int><<<1, 32, 0, c->stream>>>(acc.cells, 1, false, 0, 1.0, out, cn, acc.state, true); break;
			case B200_lng: aggr_finish_avg_kernel<lng><<<1, 32, 0, c->stream>>>(acc.cells, 1, false, 0, 1.0, out, cn, acc.state, true); break;
			case B200_hge: aggr_finish_avg_kernel<hge><<<1, 32, 0, c->stream>>>(acc.cells, 1, false, 0, 1.0, out, cn, acc.state, true); break;
			}
			COUNT_LAUNCH();
			bool ok = cuda_ok(cudaMemcpyAsync(&c->slot_h->v[8], out, 16, cudaMemcpyDeviceToHost, c->stream), "read") && stream_sync();
			dev_free(out);
			if (!ok)
				return B200_FAIL;
			memcpy(&avg, &c->slot_h->v[8], 8);
			cnt = c->slot_h->v[9];
		}
		if (cnt == 0)
			avg = nil_if_empty ? NAN : 0;
		if (cnt < a.ci.ncand && !skip_nils)
			avg = NAN;
		if (t2 == B200_flt) {
			if (avg != avg)
				*(flt *) res = NAN;
			else if (cnt > 0 && Atom<flt>::maxval() / cnt < fabs(avg)) {
				set_err("22003!overflow in sum aggregate.\n");
				return B200_FAIL;
			} else
				*(flt *) res = (flt) avg * cnt;
		} else {
			if (avg != avg)
				*(dbl *) res = NAN;
			else if (cnt > 0 && Atom<dbl>::maxval() / cnt < fabs(avg)) {
				set_err("22003!overflow in sum aggregate.\n");
				return B200_FAIL;
			} else
				*(dbl *) res = avg * cnt;
		}
		return B200_SUCCEED;
	}
	/* default result for an empty input */
	unsigned char buf[16];
	if (nil_if_empty)
		atom_nil_bits(t2, buf);
	else
		memset(buf, 0, sizeof(buf));
	memcpy(res, buf, w2);
	if (a.ci.ncand == 0) {
		bool okc = (int_rank(t2) >= 0 && int_rank(t1) >= 0 && int_rank(t1) <= int_rank(t2)) ||
			   (t2 == B200_flt && t1 == B200_flt) || (t2 == B200_dbl && (t1 == B200_flt || t1 == B200_dbl));
		if (!okc && int_rank(t2) < 0 && t2 != B200_flt && t2 != B200_dbl) {
			set_err("type combination (sum(%s)->%s) not supported.\n", atomname(a.b->ttype), atomname(tp));
			return B200_FAIL;
		}
		return B200_SUCCEED;
	}
	/* the finishing kernel stores the scalar straight into the mapped result slot; do_sum's own
	 * sync makes it visible */
	if (do_sum(a, tp, skip_nils, nil_if_empty, &c->slot_d->v[8], "BATsum") == B200_BUN_NONE)
		return B200_FAIL;
	memcpy(res, &c->slot_h->v[8], w2);
	return B200_SUCCEED;
}

// ---------------------------------------------------------------- count
// histogram of group ids when the values need not be looked at (gdk_aggr.c:3118-3132)
template <bool USE_SHM, typename GT = oid>
__global__ void __launch_bounds__(256)
gid_count_kernel(const GT *__restrict__ gids, uint64_t n, oid gmin, oid gmax, unsigned ngrp, u64 *__restrict__ cnts)
{
	extern __shared__ u64 sh_cnt[];
	const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
	if (ngrp <= 8) {
		unsigned c[8] = {0, 0, 0, 0, 0, 0, 0, 0};
		constexpr int U = 8;	/* independent loads in flight per thread */
		for (uint64_t k0 = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; k0 < n; k0 += stride * U) {
			oid g[U];
#pragma unroll
			for (int u = 0; u < U; u++) {
				const uint64_t k = k0 + u * stride;
				g[u] = k < n ? (oid) __ldcs(gids + k) : OID_NIL;
			}
#pragma unroll
			for (int u = 0; u < U; u++) {
				const u64 go = g[u] - gmin;
				const unsigned gi = go <= gmax - gmin ? (unsigned) go : 0xffffffffu;
#pragma unroll
				for (int j = 0; j < 8; j++)
					c[j] += gi == j;
			}
		}
#pragma unroll
		for (int j = 0; j < 8; j++) {
			unsigned v = c[j];
#pragma unroll
			for (int o = 16; o > 0; o >>= 1)
				v += __shfl_down_sync(0xffffffffu, v, o);
			if ((threadIdx.x & 31) == 0 && v && j < ngrp)
				atomicAdd(&cnts[j], (u64) v);
		}
		return;
	}
	u64 *tab = cnts;
	if (USE_SHM) {
		tab = sh_cnt;
		for (unsigned i = threadIdx.x; i < ngrp; i += blockDim.x)
			tab[i] = 0;
		__syncthreads();
	}
	for (uint64_t k = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
		oid g = __ldcs(gids + k);
		if (g >= gmin && g <= gmax)
			atomicAdd(&tab[g - gmin], 1ULL);
	}
	if (USE_SHM) {
		__syncthreads();
		for (unsigned i = threadIdx.x; i < ngrp; i += blockDim.x)
			if (tab[i])
				atomicAdd(&cnts[i], tab[i]);
	}
}

// one-byte id twin, at most 8 groups (Q1's count(*)): 16 ids per 16-byte load, one SIMD byte compare per
// group and word; byte-wide loads ran this at 0.74 TB/s (profiles/r2_q1_sf100_launches.md)
__global__ void __launch_bounds__(256)
gid8_count_kernel(const unsigned char *__restrict__ gids, uint64_t n, unsigned gmin, unsigned ngrp, u64 *__restrict__ cnts)
{
	unsigned c[8] = {0, 0, 0, 0, 0, 0, 0, 0};
	const uint64_t nvec = n / 16, stride = (uint64_t) gridDim.x * blockDim.x;
	const uint4 *v = reinterpret_cast<const uint4 *>(gids);
	constexpr int U = 4;
	for (uint64_t i0 = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * U) {
		uint4 q[U];
#pragma unroll
		for (int u = 0; u < U; u++) {
			const uint64_t i = i0 + u * stride;
			q[u] = i < nvec ? __ldcs(v + i) : make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
		}
#pragma unroll
		for (int u = 0; u < U; u++) {
#pragma unroll
			for (int j = 0; j < 8; j++) {
				const unsigned pat = (gmin + j) * 0x01010101u;
				c[j] += (__popc(__vcmpeq4(q[u].x, pat)) + __popc(__vcmpeq4(q[u].y, pat)) +
					 __popc(__vcmpeq4(q[u].z, pat)) + __popc(__vcmpeq4(q[u].w, pat))) >> 3;
			}
		}
	}
	if (blockIdx.x == 0 && threadIdx.x < n % 16) {
		const unsigned go = (unsigned) gids[nvec * 16 + threadIdx.x] - gmin;
#pragma unroll
		for (int j = 0; j < 8; j++)
			c[j] += go == (unsigned) j;
	}
#pragma unroll
	for (int j = 0; j < 8; j++) {
		unsigned x = c[j];
#pragma unroll
		for (int o = 16; o > 0; o >>= 1)
			x += __shfl_down_sync(0xffffffffu, x, o);
		if ((threadIdx.x & 31) == 0 && x && j < (int) ngrp)
			atomicAdd(&cnts[j], (u64) x);
	}
}

extern "C" b200_bat *b200_BATgroupcount(const b200_bat *b, const b200_bat *g, const b200_bat *e, const b200_bat *s, int tp, bool skip_nils)
{
	b200_clrerr();
	(void) tp;
	AggrSetup a;
	if (!aggr_init(a, b, g, e, s))
		return nullptr;
	if (a.g == nullptr) {
		set_err("b and g must be aligned\n");
		return nullptr;
	}
	Ctx *c = ctx();
	if (a.ci.ncand == 0 || a.ngrp == 0) {
		lng zero = 0;
		return bat_constant(a.ngrp == 0 ? 0 : a.min, B200_lng, &zero, a.ngrp);
	}
	const unsigned ngrp = (unsigned) a.ngrp;
	if (!skip_nils || a.b->tnonil) {
		if (a.g->ttype == B200_void) {
			/* dense g: cnts[i] = 1 for every candidate row */
			lng one = 1;
			b200_bat *bn = bat_constant(a.min, B200_lng, &one, a.ngrp);
			if (bn && a.ci.ncand < a.ngrp &&
			    !cuda_ok(cudaMemsetAsync((lng *) bn->base + a.ci.ncand, 0, (a.ngrp - a.ci.ncand) * 8, c->stream), "memset")) {
				b200_BATfree(bn);
				return nullptr;
			}
			if (bn)
				bn->tsorted = bn->trevsorted = bn->tkey = bn->count <= 1;
			return bn;
		}
		b200_bat *bn = bat_new(a.min, B200_lng, a.ngrp);
		if (bn == nullptr)
			return nullptr;
		if (!cuda_ok(cudaMemsetAsync(bn->base, 0, (size_t) ngrp * 8, c->stream), "memset")) {
			b200_BATfree(bn);
			return nullptr;
		}
		const unsigned grid = grid_for(a.ci.ncand, 256 * 8, NSM * 8);
		Shadow gsh;
		KTimer kt("gid_count", a.ci.ncand * 8);		/* algorithmic: the API's 8-byte ids */
		if (shadow_get(a.g->base, SHADOW_GID8, a.g->count, gsh) && ngrp <= 8 && a.min + 8 <= 255 && ((uintptr_t) gsh.p & 15) == 0)
			gid8_count_kernel<<<grid_for(a.ci.ncand / 16 + 1, 256 * 4, NSM * 8), 256, 0, c->stream>>>(
				(const unsigned char *) gsh.p, a.ci.ncand, (unsigned) a.min, ngrp, (u64 *) bn->base);
		else if (shadow_get(a.g->base, SHADOW_GID8, a.g->count, gsh) && (size_t) ngrp * 8 <= 48 * 1024)
			gid_count_kernel<true, unsigned char><<<grid, 256, ngrp > 8 ? (size_t) ngrp * 8 : 0, c->stream>>>((const unsigned char *) gsh.p, a.ci.ncand, a.min, a.max, ngrp, (u64 *) bn->base);
		else if ((size_t) ngrp * 8 <= 48 * 1024)
			gid_count_kernel<true><<<grid, 256, ngrp > 8 ? (size_t) ngrp * 8 : 0, c->stream>>>((const oid *) a.g->base, a.ci.ncand, a.min, a.max, ngrp, (u64 *) bn->base);
		else
			gid_count_kernel<false><<<grid, 256, 0, c->stream>>>((const oid *) a.g->base, a.ci.ncand, a.min, a.max, ngrp, (u64 *) bn->base);
		COUNT_LAUNCH();
		kt.stop();
		ctx()->algo = "count: cuda gid histogram";
		if (!cuda_ok(cudaGetLastError(), "count launch")) {
			b200_BATfree(bn);
			return nullptr;
		}
		bn->count = a.ngrp;
		agg_props(bn, 0);
		return bn;
	}
	/* count the non-nil values per group */
	AggrAccum acc;
	if (!aggr_accumulate(a, true, 0, acc))
		return nullptr;
	b200_bat *bn = bat_new(a.min, B200_lng, a.ngrp);
	if (bn == nullptr)
		return nullptr;
	const AggrLayout L(acc.nd, false, false);
	aggr_finish_count_kernel<<<(ngrp + 127) / 128, 128, 0, c->stream>>>(acc.cells + (size_t) L.c_cnt * ngrp, ngrp, (lng *) bn->base);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "count finish")) {
		b200_BATfree(bn);
		return nullptr;
	}
	bn->count = a.ngrp;
	agg_props(bn, 0);
	return bn;
}

// ---------------------------------------------------------------- avg
extern "C" b200_return b200_BATgroupavg(b200_bat **bnp, b200_bat **cntsp, const b200_bat *b, const b200_bat *g,
					const b200_bat *e, const b200_bat *s, int tp, bool skip_nils, int scale)
{
	b200_clrerr();
	(void) tp;
	if (bnp == nullptr) {
		set_err("BATgroupavg: result pointer required\n");
		return B200_FAIL;
	}
	*bnp = nullptr;
	if (cntsp)
		*cntsp = nullptr;
	AggrSetup a;
	if (!aggr_init(a, b, g, e, s))
		return B200_FAIL;
	if (a.g == nullptr) {
		set_err("b and g must be aligned\n");
		return B200_FAIL;
	}
	Ctx *c = ctx();
	b200_bat *bn = nullptr, *cn = nullptr;
	if (a.ci.ncand == 0 || a.ngrp == 0) {
		bn = bat_constant(a.ngrp == 0 ? 0 : a.min, B200_dbl, nullptr, a.ngrp);
		if (bn == nullptr)
			return B200_FAIL;
		if (cntsp) {
			lng zero = 0;
			cn = bat_constant(a.ngrp == 0 ? 0 : a.min, B200_lng, &zero, a.ngrp);
			if (cn == nullptr) {
				b200_BATfree(bn);
				return B200_FAIL;
			}
			*cntsp = cn;
		}
		*bnp = bn;
		return B200_SUCCEED;
	}
	if ((!skip_nils || cntsp == nullptr || a.b->tnonil) &&
	    (a.e == nullptr || (a.e->count == a.ci.ncand && a.e->hseqbase == a.b->hseqbase)) &&
	    (bat_is_dense(a.g) || (a.g->tkey && a.g->tnonil))) {
		bn = b200_BATconvert(b, s, B200_dbl, 0, 0, 0);
		if (bn == nullptr)
			return B200_FAIL;
		if (cntsp) {
			lng one = 1;
			cn = bat_constant(a.min, B200_lng, &one, a.ngrp);
			if (cn == nullptr) {
				b200_BATfree(bn);
				return B200_FAIL;
			}
			*cntsp = cn;
		}
		*bnp = bn;
		return B200_SUCCEED;
	}
	AggrAccum acc;
	if (!aggr_accumulate(a, skip_nils, 0, acc))
		return B200_FAIL;
	const unsigned ngrp = (unsigned) a.ngrp;
	bn = bat_new(a.min, B200_dbl, a.ngrp);
	if (cntsp)
		cn = bat_new(a.min, B200_lng, a.ngrp);
	if (bn == nullptr || (cntsp && cn == nullptr)) {
		b200_BATfree(bn);
		b200_BATfree(cn);
		return B200_FAIL;
	}
	const double fac = scale != 0 ? pow(10.0, (double) scale) : 1.0;
	lng *cp = cn ? (lng *) cn->base : nullptr;
	const unsigned gr = (ngrp + 127) / 128;
#define AVG_CASE(T) aggr_finish_avg_kernel<T><<<gr, 128, 0, c->stream>>>(acc.cells, ngrp, acc.niltrack, acc.fexp, fac, (dbl *) bn->base, cp, acc.state)
	switch (basetype(a.b->ttype)) {
	case B200_bte: AVG_CASE(bte); break;
	case B200_sht: AVG_CASE(sht); break;
	case B200_int: AVG_CASE(int); break;
	case B200_lng: AVG_CASE(lng); break;
	case B200_hge: AVG_CASE(hge); break;
	case B200_flt: AVG_CASE(flt); break;
	case B200_dbl: AVG_CASE(dbl); break;
	}
#undef AVG_CASE
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "avg finish") ||
	    !cuda_ok(cudaMemcpyAsync(c->slot_h->v, acc.state, 32, cudaMemcpyDeviceToHost, c->stream), "state read") ||
	    !stream_sync()) {
		b200_BATfree(bn);
		b200_BATfree(cn);
		return B200_FAIL;
	}
	bn->count = a.ngrp;
	agg_props(bn, c->slot_h->v[1]);
	if (cn) {
		cn->count = a.ngrp;
		agg_props(cn, 0);
		*cntsp = cn;
	}
	*bnp = bn;
	return B200_SUCCEED;
}

// ---------------------------------------------------------------- avg3 / avg3combine / calcavg
static void free3(b200_bat *x, b200_bat *y, b200_bat *z)
{
	b200_BATfree(x);
	b200_BATfree(y);
	b200_BATfree(z);
}

// gdk_aggr.c:1996-2290.  Integer columns only (the reference's switch has no flt/dbl case).
extern "C" b200_return b200_BATgroupavg3(b200_bat **avgp, b200_bat **remp, b200_bat **cntp, const b200_bat *b, const b200_bat *g,
					 const b200_bat *e, const b200_bat *s, bool skip_nils)
{
	b200_clrerr();
	if (avgp == nullptr || remp == nullptr || cntp == nullptr) {
		set_err("BATgroupavg3: result pointers required\n");
		return B200_FAIL;
	}
	*avgp = *remp = *cntp = nullptr;
	AggrSetup a;
	if (!aggr_init(a, b, g, e, s))
		return B200_FAIL;
	const int t = basetype(a.b->ttype);
	if (int_rank(t) < 0) {
		set_err("BATgroupavg3: type %s is not an integer type\n", atomname(a.b->ttype));
		return B200_FAIL;
	}
	Ctx *c = ctx();
	b200_bat *bn, *rn, *cn;
	const lng zero = 0;
	if (a.ci.ncand == 0 || a.ngrp == 0) {
		const uint64_t hs = a.ngrp == 0 ? 0 : a.min;
		bn = bat_constant(hs, a.b->ttype, nullptr, a.ngrp);
		rn = bat_constant(hs, B200_lng, nullptr, a.ngrp);
		cn = bat_constant(hs, B200_lng, &zero, a.ngrp);
		if (bn == nullptr || rn == nullptr || cn == nullptr) {
			free3(bn, rn, cn);
			return B200_FAIL;
		}
		*avgp = bn;
		*remp = rn;
		*cntp = cn;
		return B200_SUCCEED;
	}
	AggrAccum acc;
	if (!aggr_accumulate(a, skip_nils, 0, acc))
		return B200_FAIL;
	const unsigned ngrp = (unsigned) a.ngrp;
	bn = bat_new(a.min, a.b->ttype, a.ngrp);
	rn = bat_new(a.min, B200_lng, a.ngrp);
	cn = bat_new(a.min, B200_lng, a.ngrp);
	if (bn == nullptr || rn == nullptr || cn == nullptr) {
		free3(bn, rn, cn);
		return B200_FAIL;
	}
	const unsigned gr = (ngrp + 127) / 128;
#define AVG3_CASE(T) aggr_finish_avg3_kernel<T><<<gr, 128, 0, c->stream>>>(acc.cells, ngrp, acc.niltrack, (T *) bn->base, (lng *) rn->base, (lng *) cn->base, acc.state)
	switch (t) {
	case B200_bte: AVG3_CASE(bte); break;
	case B200_sht: AVG3_CASE(sht); break;
	case B200_int: AVG3_CASE(int); break;
	case B200_lng: AVG3_CASE(lng); break;
	case B200_hge: AVG3_CASE(hge); break;
	}
#undef AVG3_CASE
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "avg3 finish") ||
	    !cuda_ok(cudaMemcpyAsync(c->slot_h->v, acc.state, 32, cudaMemcpyDeviceToHost, c->stream), "state read") ||
	    !stream_sync()) {
		free3(bn, rn, cn);
		return B200_FAIL;
	}
	if ((unsigned) c->slot_h->v[0] & AGGR_F_OVERFLOW) {
		free3(bn, rn, cn);
		set_err("BATgroupavg3: group sum does not fit the device path (2^32 or more rows of hge values)\n");
		return B200_FAIL;
	}
	bn->count = rn->count = cn->count = a.ngrp;
	agg_props(bn, c->slot_h->v[1]);
	agg_props(rn, c->slot_h->v[2]);
	agg_props(cn, c->slot_h->v[2]);
	*avgp = bn;
	*remp = rn;
	*cntp = cn;
	return B200_SUCCEED;
}

// gdk_aggr.c:2634-2900
extern "C" b200_bat *b200_BATgroupavg3combine(const b200_bat *avg, const b200_bat *rem, const b200_bat *cnt, const b200_bat *g,
					      const b200_bat *e, bool skip_nils)
{
	b200_clrerr();
	AggrSetup a;
	if (!aggr_init(a, avg, g, e, nullptr))
		return nullptr;
	Staged sr, sc;
	if (rem == nullptr || cnt == nullptr || !stage(rem, sr) || !stage(cnt, sc))
		return nullptr;
	if (a.ci.ncand != sr.b->count || a.ci.ncand != sc.b->count) {
		set_err("input bats not aligned");
		return nullptr;
	}
	if (a.ci.ncand == 0 || a.ngrp == 0)
		return bat_constant(a.ngrp == 0 ? 0 : a.min, a.b->ttype, nullptr, a.ngrp);
	const int t = basetype(a.b->ttype);
	if (int_rank(t) < 0 || t == B200_hge || basetype(sr.b->ttype) != B200_lng || basetype(sc.b->ttype) != B200_lng) {
		set_err("BATgroupavg3combine: (avg %s, rem %s, cnt %s) is not on the device path\n", atomname(a.b->ttype),
			atomname(sr.b->ttype), atomname(sc.b->ttype));
		return nullptr;
	}
	Ctx *c = ctx();
	const unsigned ngrp = (unsigned) a.ngrp;
	const size_t bytes = (size_t) 6 * ngrp * 8;
	u64 *cells = (u64 *) dev_alloc(bytes), *state = (u64 *) dev_alloc(64);
	b200_bat *bn = bat_new(a.min, a.b->ttype, a.ngrp);
	bool ok = cells && state && bn && cuda_ok(cudaMemsetAsync(cells, 0, bytes, c->stream), "memset cells") &&
		  cuda_ok(cudaMemsetAsync(state, 0, 64, c->stream), "memset state");
	if (ok) {
		const uint64_t n = a.ci.ncand;
		const oid *gids = a.g && a.g->ttype != B200_void ? (const oid *) a.g->base : nullptr;
		const bool gconst = a.g == nullptr;
		const oid gdense = a.g && a.g->ttype == B200_void ? a.g->tseqbase : 0;
		const unsigned grid = grid_for(n, 256, NSM * 8), gr = (ngrp + 127) / 128;
#define COMB_CASE(T)                                                                                                         \
	avg3_combine_kernel<T><<<grid, 256, 0, c->stream>>>((const T *) a.b->base, (const lng *) sr.b->base, (const lng *) sc.b->base, \
							     gids, a.min, a.max, gdense, gconst, n, skip_nils, cells, ngrp);            \
	COUNT_LAUNCH();                                                                                                      \
	avg3_combine_finish_kernel<T><<<gr, 128, 0, c->stream>>>(cells, ngrp, (T *) bn->base, state);                       \
	COUNT_LAUNCH()
		switch (t) {
		case B200_bte: COMB_CASE(bte); break;
		case B200_sht: COMB_CASE(sht); break;
		case B200_int: COMB_CASE(int); break;
		case B200_lng: COMB_CASE(lng); break;
		}
#undef COMB_CASE
		ok = cuda_ok(cudaGetLastError(), "avg3combine") &&
		     cuda_ok(cudaMemcpyAsync(c->slot_h->v, state, 32, cudaMemcpyDeviceToHost, c->stream), "state read") && stream_sync();
		if (ok && ((unsigned) c->slot_h->v[0] & AGGR_F_OVERFLOW)) {
			set_err("BATgroupavg3combine: combined count does not fit a lng\n");
			ok = false;
		}
	}
	if (cells) dev_free(cells);
	if (state) dev_free(state);
	if (!ok) {
		b200_BATfree(bn);
		return nullptr;
	}
	bn->count = a.ngrp;
	agg_props(bn, c->slot_h->v[1]);
	ctx()->algo = "avg3combine: cuda 128-bit digit cells";
	return bn;
}

// gdk_aggr.c:2987-3046.  Integers: (dbl) sum / n while the sum fits a hge, else floor average plus
// (dbl) rem / n; floats: exactly rounded sum / n (the reference's running mean differs in the last ulps).
extern "C" b200_return b200_BATcalcavg(const b200_bat *b, const b200_bat *s, double *avg, uint64_t *vals, int scale)
{
	b200_clrerr();
	AggrSetup a;
	if (avg == nullptr || !aggr_init(a, b, nullptr, nullptr, s))
		return B200_FAIL;
	const int t = basetype(a.b->ttype);
	if (int_rank(t) < 0 && t != B200_flt && t != B200_dbl) {
		set_err("average of type %s unsupported.\n", atomname(a.b->ttype));
		return B200_FAIL;
	}
	Ctx *c = ctx();
	dbl r = NAN;
	uint64_t n = 0;
	if (a.ci.ncand > 0) {
		AggrAccum acc;
		if (!aggr_accumulate(a, true, 0, acc))
			return B200_FAIL;
		dbl *out = (dbl *) dev_alloc(16);
		if (out == nullptr)
			return B200_FAIL;
		lng *cn = (lng *) (out + 1);
#define CAVG_CASE(T) aggr_finish_avg_kernel<T><<<1, 32, 0, c->stream>>>(acc.cells, 1, false, acc.fexp, 1.0, out, cn, acc.state, true)
		switch (t) {
		case B200_bte: CAVG_CASE(bte); break;
		case B200_sht: CAVG_CASE(sht); break;
		case B200_int: CAVG_CASE(int); break;
		case B200_lng: CAVG_CASE(lng); break;
		case B200_hge: CAVG_CASE(hge); break;
		case B200_flt: CAVG_CASE(flt); break;
		case B200_dbl: CAVG_CASE(dbl); break;
		}
#undef CAVG_CASE
		COUNT_LAUNCH();
		bool ok = cuda_ok(cudaMemcpyAsync(&c->slot_h->v[8], out, 16, cudaMemcpyDeviceToHost, c->stream), "read") && stream_sync();
		dev_free(out);
		if (!ok)
			return B200_FAIL;
		memcpy(&r, &c->slot_h->v[8], 8);
		n = c->slot_h->v[9];
	}
	if (n == 0)
		r = NAN;
	else if (scale != 0)
		r /= pow(10.0, (double) scale);
	*avg = r;
	if (vals)
		*vals = n;
	return B200_SUCCEED;
}

// ---------------------------------------------------------------- average from merged (sum, count)
// Phase 2 of the two-phase average (mergetable splits avg into sum + count per slice and recombines,
// opt_mergetable.c:1204-1215).  To be bit-identical with ONE BATgroupavg over all rows the recombination
// uses that function's own closing formula (gdk_aggr.c:1741-1748, 1952-1957): with a = floor(S / n) and
// rem = S - a * n (the invariant AVERAGE_ITER maintains, 0 <= rem < n),  avg = a + (dbl) rem / n, then
// / 10^scale.  S: lng or hge sums, n: lng counts; n == 0 / nil or S nil gives nil.
template <typename TS>
__global__ void avg_from_sum_kernel(const TS *__restrict__ sum, const lng *__restrict__ cnt, uint64_t n, double fac,
				    dbl *__restrict__ out, unsigned long long *__restrict__ nils)
{
	const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	const TS S = sum[i];
	const lng c = cnt[i];
	if (Atom<TS>::is_nil(S) || Atom<lng>::is_nil(c) || c <= 0) {
		out[i] = NAN;
		atomicAdd(nils, 1ULL);
		return;
	}
	const hge s = (hge) S;
	hge a = s / c, rem = s % c;	/* C truncates towards zero */
	if (rem < 0) {
		a -= 1;
		rem += c;
	}
	double r = (double) a + (double) (lng) rem / (double) c;
	if (fac != 1.0)
		r /= fac;
	out[i] = r;
}

extern "C" b200_bat *b200_BATcalcavgsum(const b200_bat *sum, const b200_bat *cnt, int scale)
{
	b200_clrerr();
	if (sum == nullptr || cnt == nullptr) {
		set_err("BATcalcavgsum: sum and cnt required\n");
		return nullptr;
	}
	Staged ss, sc;
	if (!stage(sum, ss) || !stage(cnt, sc))
		return nullptr;
	sum = ss.b;
	cnt = sc.b;
	const int ts = basetype(sum->ttype);
	if ((ts != B200_lng && ts != B200_hge) || basetype(cnt->ttype) != B200_lng || sum->ttype == B200_oid || cnt->ttype == B200_oid) {
		set_err("BATcalcavgsum: sums must be lng or hge and counts lng\n");
		return nullptr;
	}
	if (sum->count != cnt->count) {
		set_err("inputs not the same size.\n");
		return nullptr;
	}
	Ctx *c = ctx();
	const uint64_t n = sum->count;
	b200_bat *bn = bat_new(sum->hseqbase, B200_dbl, n);
	if (bn == nullptr)
		return nullptr;
	unsigned long long nils = 0;
	if (n > 0) {
		unsigned long long *dn = (unsigned long long *) scratch_get(64);
		const double fac = scale ? pow(10.0, (double) scale) : 1.0;
		if (dn == nullptr || !cuda_ok(cudaMemsetAsync(dn, 0, 8, c->stream), "memset")) {
			b200_BATfree(bn);
			return nullptr;
		}
		const unsigned grid = (unsigned) ((n + 255) / 256);
		if (ts == B200_hge)
			avg_from_sum_kernel<hge><<<grid, 256, 0, c->stream>>>((const hge *) sum->base, (const lng *) cnt->base, n, fac,
									    (dbl *) bn->base, dn);
		else
			avg_from_sum_kernel<lng><<<grid, 256, 0, c->stream>>>((const lng *) sum->base, (const lng *) cnt->base, n, fac,
									    (dbl *) bn->base, dn);
		COUNT_LAUNCH();
		if (!cuda_ok(cudaGetLastError(), "avg launch") ||
		    !cuda_ok(cudaMemcpyAsync(&c->slot_h->v[0], dn, 8, cudaMemcpyDeviceToHost, c->stream), "nil count") || !stream_sync()) {
			b200_BATfree(bn);
			return nullptr;
		}
		nils = c->slot_h->v[0];
	}
	ctx()->algo = "avg: from merged (sum, count)";
	bn->count = n;
	bn->tnil = nils != 0;
	bn->tnonil = nils == 0;
	bn->tsorted = bn->trevsorted = bn->tkey = n <= 1;
	return bn;
}
