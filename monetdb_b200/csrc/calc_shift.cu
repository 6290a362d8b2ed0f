// calc_shift.cu — BATcalclsh / BATcalcrsh and their cst forms (gdk_calc.c:3246-3420, 3596-3770; the typeswitch
// loops :3075-3240, 3425-3590).  Any integral left operand, any integral shift count; the result has the
// type of the LEFT operand.  nil in, nil out.  The reference refuses a row (GDK error, whole call fails) when
//   SHIFT_CHECK      the count is negative or >= the width of the left operand;
//   LSH_CHECK_<t>    (left shifts of bte/sht/int/lng by bte/sht/int/lng) additionally when the left operand is
//                    negative or the result would not fit:  a < 0 || a > (max >> b);
//   NO_SHIFT_CHECK   hge << bte: nothing is checked (gdk_calc.c:3208-3210) — reproduced: a count outside
//                    0..127 is undefined behaviour there, we mask it like the x86 shift the compiler emits.
// The error text names the function and the offending pair: "<func>: shift operand too large in LSH(a,b).".
#include "common.cuh"
#include "operand.cuh"
#include <type_traits>

enum { SH_LSH = 0, SH_RSH = 1 };
enum { CK_SHIFT = 0, CK_LSH = 1, CK_NONE = 2 };
#define SH_THREADS 256

template <int OP, typename T1, typename T2> __host__ __device__ constexpr int shift_check_kind()
{
	if (OP == SH_RSH)
		return CK_SHIFT;
	if (sizeof(T1) == 16)
		return sizeof(T2) == 1 ? CK_NONE : CK_SHIFT;
	return sizeof(T2) == 16 ? CK_SHIFT : CK_LSH;
}

template <int OP, typename T1, typename T2>
__global__ void __launch_bounds__(SH_THREADS)
shift_kernel(CmpOperand<T1> l, CmpOperand<T2> r, uint64_t n, T1 *__restrict__ dst, unsigned long long *__restrict__ state)
{
	typedef typename std::conditional<sizeof(T1) == 16, uhge, unsigned long long>::type UW;
	constexpr int CK = shift_check_kind<OP, T1, T2>();
	constexpr int BITS = 8 * (int) sizeof(T1);
	unsigned long long fail = ~0ULL;
	unsigned nils = 0;
	for (uint64_t k = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (uint64_t) gridDim.x * blockDim.x) {
		const T1 a = l.get(k);
		const T2 b = r.get(k);
		T1 z = Atom<T1>::nil();
		if (Atom<T1>::is_nil(a) || Atom<T2>::is_nil(b)) {
			nils++;
		} else {
			bool bad = false;
			if (CK != CK_NONE)
				bad = b < 0 || (hge) b >= (hge) BITS;
			if (CK == CK_LSH && !bad)
				bad = a < 0 || a > (T1) (Atom<T1>::maxval() >> (int) b);
			if (bad) {
				if (fail == ~0ULL)
					fail = k;
			} else {
				const int s = (int) b & (BITS - 1);
				if (OP == SH_LSH)
					z = (T1) ((UW) a << s);		/* bits beyond the type fall off, as in the reference's promoted shift + store */
				else
					z = (T1) (a >> s);		/* arithmetic shift of the signed value */
			}
		}
		dst[k] = z;
	}
	if (fail != ~0ULL)
		atomicMin(&state[0], fail);
#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
		nils += __shfl_down_sync(0xffffffffu, nils, o);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(&state[1], (unsigned long long) nils);
}

template <typename T> static void sh_fmt(char *buf, size_t n, T v)
{
	/* FMTbte..FMThge, gdk_calc_private.h:300-308 */
	if constexpr (sizeof(T) == 16)
		snprintf(buf, n, "%.40Lg (approx. value)", (long double) v);
	else if constexpr (sizeof(T) == 8)
		snprintf(buf, n, "%lld", (long long) v);
	else
		snprintf(buf, n, "%d", (int) v);
}

template <typename T> static bool sh_fetch(const CmpOperand<T> &o, uint64_t k, T &v)
{
	if (o.cst) {
		v = o.cval;
		return true;
	}
	uint64_t i = k;
	if (o.cand) {
		oid id;
		if (!cuda_ok(cudaMemcpy(&id, o.cand + k, 8, cudaMemcpyDeviceToHost), "fetch"))
			return false;
		i = id - o.hseq;
	}
	return cuda_ok(cudaMemcpy(&v, o.vals + i, sizeof(T), cudaMemcpyDeviceToHost), "fetch");
}

template <int OP, typename T1, typename T2>
static bool shift_run(const CmpSide &s1, const CmpSide &s2, uint64_t n, void *dst, uint64_t &nils, const char *func)
{
	Ctx *c = ctx();
	unsigned long long *state = (unsigned long long *) scratch_get(64);
	static const unsigned long long init[2] = {~0ULL, 0};
	if (state == nullptr || !cuda_ok(cudaMemcpyAsync(state, init, 16, cudaMemcpyHostToDevice, c->stream), "init"))
		return false;
	const CmpOperand<T1> l = make_operand<T1>(s1);
	const CmpOperand<T2> r = make_operand<T2>(s2);
	{
		KTimer kt("calc_shift", n * ((s1.b ? sizeof(T1) : 0) + (s2.b ? sizeof(T2) : 0) + sizeof(T1)));
		shift_kernel<OP, T1, T2><<<grid_for(n, SH_THREADS, NSM * 32), SH_THREADS, 0, c->stream>>>(l, r, n, (T1 *) dst, state);
	}
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "shift launch") ||
	    !cuda_ok(cudaMemcpyAsync(c->slot_h->v, state, 16, cudaMemcpyDeviceToHost, c->stream), "read") || !stream_sync())
		return false;
	nils = c->slot_h->v[1];
	const uint64_t fail = c->slot_h->v[0];
	if (fail != ~0ULL) {
		T1 a;
		T2 b;
		if (!sh_fetch(l, fail, a) || !sh_fetch(r, fail, b))
			return false;
		char fa[64], fb[64];
		sh_fmt(fa, sizeof(fa), a);
		sh_fmt(fb, sizeof(fb), b);
		set_err("%s: shift operand too large in %s(%s,%s).\n", func, OP == SH_LSH ? "LSH" : "RSH", fa, fb);
		return false;
	}
	return true;
}

template <int OP, typename T1>
static int shift_dispatch2(int t2, const CmpSide &s1, const CmpSide &s2, uint64_t n, void *dst, uint64_t &nils, const char *func)
{
	switch (t2) {
	case B200_bte: return shift_run<OP, T1, bte>(s1, s2, n, dst, nils, func) ? 1 : 0;
	case B200_sht: return shift_run<OP, T1, sht>(s1, s2, n, dst, nils, func) ? 1 : 0;
	case B200_int: return shift_run<OP, T1, int>(s1, s2, n, dst, nils, func) ? 1 : 0;
	case B200_lng: return shift_run<OP, T1, lng>(s1, s2, n, dst, nils, func) ? 1 : 0;
	case B200_hge: return shift_run<OP, T1, hge>(s1, s2, n, dst, nils, func) ? 1 : 0;
	}
	return -1;
}

template <int OP>
static int shift_dispatch(int t1, int t2, const CmpSide &s1, const CmpSide &s2, uint64_t n, void *dst, uint64_t &nils, const char *func)
{
	switch (t1) {
	case B200_bte: return shift_dispatch2<OP, bte>(t2, s1, s2, n, dst, nils, func);
	case B200_sht: return shift_dispatch2<OP, sht>(t2, s1, s2, n, dst, nils, func);
	case B200_int: return shift_dispatch2<OP, int>(t2, s1, s2, n, dst, nils, func);
	case B200_lng: return shift_dispatch2<OP, lng>(t2, s1, s2, n, dst, nils, func);
	case B200_hge: return shift_dispatch2<OP, hge>(t2, s1, s2, n, dst, nils, func);
	}
	return -1;
}

// mode 0: b1 OP b2; mode 1: b1 OP cst; mode 2: cst OP b1 (b1 is the shift-count column)
static b200_bat *shift_impl(int op, int mode, const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2,
			    const void *cst, int csttype, const char *func)
{
	b200_clrerr();
	if (b1 == nullptr || (mode == 0 ? b2 == nullptr : cst == nullptr)) {
		set_err("%s: BAT required\n", func);
		return nullptr;
	}
	Staged sb1, sb2, ss1, ss2;
	if (!stage(b1, sb1) || !stage(b2, sb2) || !stage(s1, ss1) || !stage(s2, ss2))
		return nullptr;
	b1 = sb1.b;
	b2 = sb2.b;
	Cand ci1, ci2;
	if (!cand_init(ci1, b1, ss1.b))
		return nullptr;
	if (mode == 0) {
		if (!cand_init(ci2, b2, ss2.b))
			return nullptr;
		if (ci1.ncand != ci2.ncand || ci1.rhseq != ci2.rhseq) {
			set_err("inputs not the same size.\n");	/* gdk_calc.c:3260-3263 */
			return nullptr;
		}
	}
	const uint64_t n = ci1.ncand;
	const int rtype = mode == 2 ? csttype : b1->ttype;	/* the type of the left operand */
	b200_bat *bn = bat_new(ci1.rhseq, rtype, n);
	if (bn == nullptr || n == 0)
		return bn;
	if (b1->ttype == B200_void || (mode == 0 && b2->ttype == B200_void)) {
		b200_BATfree(bn);
		set_err("%s: void columns are not on the device path\n", func);
		return nullptr;
	}
	CmpSide colside{b1, &ci1, nullptr, basetype(b1->ttype)};
	CmpSide l = colside, r = colside;
	if (mode == 0)
		r = CmpSide{b2, &ci2, nullptr, basetype(b2->ttype)};
	else if (mode == 1)
		r = CmpSide{nullptr, nullptr, cst, basetype(csttype)};
	else
		l = CmpSide{nullptr, nullptr, cst, basetype(csttype)};
	ctx()->algo = "calc: cuda streaming shift";
	uint64_t nils = 0;
	const int rc = op == SH_LSH ? shift_dispatch<SH_LSH>(l.type, r.type, l, r, n, bn->base, nils, func)
				    : shift_dispatch<SH_RSH>(l.type, r.type, l, r, n, bn->base, nils, func);
	if (rc <= 0) {
		b200_BATfree(bn);
		if (rc < 0)
			set_err("%s: bad input types %s,%s.\n", func, atomname(l.type), atomname(r.type));	/* gdk_calc.c:3236-3239 */
		return nullptr;
	}
	/* gdk_calc.c:3287-3293 */
	bn->count = n;
	bn->tsorted = bn->trevsorted = n <= 1 || nils == n;
	bn->tkey = n <= 1;
	bn->tnil = nils != 0;
	bn->tnonil = nils == 0;
	return bn;
}

#define SHIFTOP(NAME, OP)                                                                                            \
	extern "C" b200_bat *b200_BATcalc##NAME(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1,         \
						const b200_bat *s2)                                                   \
	{                                                                                                            \
		return shift_impl(OP, 0, b1, b2, s1, s2, nullptr, 0, "BATcalc" #NAME);                               \
	}                                                                                                            \
	extern "C" b200_bat *b200_BATcalc##NAME##cst(const b200_bat *b, const void *v, int vtype, const b200_bat *s) \
	{                                                                                                            \
		return shift_impl(OP, 1, b, nullptr, s, nullptr, v, vtype, "BATcalc" #NAME "cst");                   \
	}                                                                                                            \
	extern "C" b200_bat *b200_BATcalccst##NAME(const void *v, int vtype, const b200_bat *b, const b200_bat *s)   \
	{                                                                                                            \
		return shift_impl(OP, 2, b, nullptr, s, nullptr, v, vtype, "BATcalccst" #NAME);                      \
	}
SHIFTOP(lsh, SH_LSH)
SHIFTOP(rsh, SH_RSH)
