// compare.cu — BATcalc{lt,le,gt,ge,eq,ne} (+cst forms) and BATcalcisnil on the device
// (gdk_calc_compare.h template re-included by gdk_calc_compare_{lt,le,gt,ge,eq,ne}.c;
// loops BINARY_3TYPE_FUNC / _nilmatch / _nonil, gdk_calc_private.h:391-535;
// BATcalcisnil gdk_calc.c:802-913).
//
// dst[k] = bit(lft[i_k] OP rgt[j_k]); a nil operand gives bit_nil unless
// nil_matches (eq/ne only), where nil compares equal to nil.  Mixed operand types
// compare as C does after the usual arithmetic conversions (any dbl -> dbl, else
// any flt -> flt, else the wider integer).  One streaming pass, 4 rows per
// thread so the bit column is written 4 bytes at a time.
#include "common.cuh"
#include "operand.cuh"

enum { CMP_LT = 0, CMP_LE, CMP_GT, CMP_GE, CMP_EQ, CMP_NE };
#define BIT_NIL ((int8_t) -128)
#define CMP_THREADS 256

template <typename A, typename B> struct Common {
	typedef typename std::conditional<
		std::is_same<A, dbl>::value || std::is_same<B, dbl>::value, dbl,
		typename std::conditional<
			std::is_same<A, flt>::value || std::is_same<B, flt>::value, flt,
			typename std::conditional<(sizeof(A) >= sizeof(B)), A, B>::type>::type>::type type;
};

template <typename C> __device__ __forceinline__ int8_t cmp_apply(int op, C a, C b)
{
	switch (op) {
	case CMP_LT: return a < b;
	case CMP_LE: return a <= b;
	case CMP_GT: return a > b;
	case CMP_GE: return a >= b;
	case CMP_EQ: return a == b;
	default: return a != b;
	}
}

template <typename T1, typename T2>
__global__ void __launch_bounds__(CMP_THREADS)
cmp_kernel(CmpOperand<T1> l, CmpOperand<T2> r, uint64_t n, int op, bool nonil, bool nil_matches, int8_t *__restrict__ dst,
	   unsigned long long *__restrict__ nils_out)
{
	typedef typename Common<T1, T2>::type C;
	unsigned nils = 0;
	const uint64_t nquads = (n + 3) / 4;
	for (uint64_t q = (uint64_t) blockIdx.x * CMP_THREADS + threadIdx.x; q < nquads; q += (uint64_t) gridDim.x * CMP_THREADS) {
		int8_t res[4];
#pragma unroll
		for (int e = 0; e < 4; e++) {
			const uint64_t k = 4 * q + e;
			res[e] = 0;
			if (k < n) {
				const T1 a = l.get(k);
				const T2 b = r.get(k);
				const bool an = !nonil && Atom<T1>::is_nil(a), bn = !nonil && Atom<T2>::is_nil(b);
				if (an || bn) {
					if (nil_matches) {
						res[e] = cmp_apply<int>(op, (int) an, (int) bn);
					} else {
						res[e] = BIT_NIL;
						nils++;
					}
				} else {
					res[e] = cmp_apply<C>(op, (C) a, (C) b);
				}
			}
		}
		if (4 * q + 3 < n) {
			char4 v = make_char4(res[0], res[1], res[2], res[3]);
			*reinterpret_cast<char4 *>(dst + 4 * q) = v;
		} else {
			for (int e = 0; e < 4 && 4 * q + e < n; e++)
				dst[4 * q + e] = res[e];
		}
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
		nils += __shfl_down_sync(0xffffffffu, nils, o);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(nils_out, (unsigned long long) nils);
}

template <typename T>
__global__ void __launch_bounds__(CMP_THREADS)
isnil_kernel(CmpOperand<T> l, uint64_t n, int8_t *__restrict__ dst, int8_t notnil)
{
	const uint64_t nquads = (n + 3) / 4;
	for (uint64_t q = (uint64_t) blockIdx.x * CMP_THREADS + threadIdx.x; q < nquads; q += (uint64_t) gridDim.x * CMP_THREADS) {
		int8_t res[4] = {0, 0, 0, 0};
#pragma unroll
		for (int e = 0; e < 4; e++)
			if (4 * q + e < n)
				res[e] = (int8_t) Atom<T>::is_nil(l.get(4 * q + e)) ^ notnil;
		if (4 * q + 3 < n)
			*reinterpret_cast<char4 *>(dst + 4 * q) = make_char4(res[0], res[1], res[2], res[3]);
		else
			for (int e = 0; e < 4 && 4 * q + e < n; e++)
				dst[4 * q + e] = res[e];
	}
}

template <typename T1, typename T2>
static bool cmp_run(const CmpSide &s1, const CmpSide &s2, uint64_t n, int op, bool nonil, bool nil_matches, int8_t *dst,
		    unsigned long long *nils_d)
{
	Ctx *c = ctx();
	CmpOperand<T1> l = make_operand<T1>(s1);
	CmpOperand<T2> r = make_operand<T2>(s2);
	KTimer kt("compare", n * ((s1.b ? sizeof(T1) : 0) + (s2.b ? sizeof(T2) : 0) + 1));
	cmp_kernel<T1, T2><<<grid_for((n + 3) / 4, CMP_THREADS, NSM * 16), CMP_THREADS, 0, c->stream>>>(l, r, n, op, nonil, nil_matches, dst, nils_d);
	COUNT_LAUNCH();
	return cuda_ok(cudaGetLastError(), "compare launch");
}

template <typename T1>
static int cmp_dispatch2(const CmpSide &s1, const CmpSide &s2, uint64_t n, int op, bool nonil, bool nil_matches, int8_t *dst,
			 unsigned long long *nils_d)
{
	switch (s2.type) {
	case B200_bte: return cmp_run<T1, bte>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_sht: return cmp_run<T1, sht>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_int: return cmp_run<T1, int>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_lng: return cmp_run<T1, lng>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_hge: return cmp_run<T1, hge>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_flt: return cmp_run<T1, flt>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_dbl: return cmp_run<T1, dbl>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	}
	return -1;
}

static int cmp_dispatch(const CmpSide &s1, const CmpSide &s2, uint64_t n, int op, bool nonil, bool nil_matches, int8_t *dst,
			unsigned long long *nils_d)
{
	switch (s1.type) {
	case B200_bte: return cmp_dispatch2<bte>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_sht: return cmp_dispatch2<sht>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_int: return cmp_dispatch2<int>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_lng: return cmp_dispatch2<lng>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_hge: return cmp_dispatch2<hge>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_flt: return cmp_dispatch2<flt>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	case B200_dbl: return cmp_dispatch2<dbl>(s1, s2, n, op, nonil, nil_matches, dst, nils_d);
	}
	return -1;
}

// b2 == NULL: compare with the constant (cst, csttype)
static b200_bat *compare_impl(int op, const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2,
			      const void *cst, int csttype, bool nil_matches, const char *func)
{
	if (b1 == nullptr || (b2 == nullptr && cst == nullptr)) {
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
	if (b2) {
		if (!cand_init(ci2, b2, ss2.b))
			return nullptr;
		if (ci1.ncand != ci2.ncand || ci1.rhseq != ci2.rhseq) {
			set_err("inputs not the same size.\n");	/* gdk_calc_compare.h:841-844 */
			return nullptr;
		}
	}
	b200_bat *bn = bat_new(ci1.rhseq, B200_bit, ci1.ncand);
	if (bn == nullptr || ci1.ncand == 0)
		return bn;
	if (b1->ttype == B200_void || (b2 && b2->ttype == B200_void)) {
		b200_BATfree(bn);
		set_err("%s: void columns are not on the device path\n", func);
		return nullptr;
	}
	Ctx *c = ctx();
	CmpSide l{b1, &ci1, nullptr, basetype(b1->ttype)};
	CmpSide r = b2 ? CmpSide{b2, &ci2, nullptr, basetype(b2->ttype)} : CmpSide{nullptr, nullptr, cst, basetype(csttype)};
	const bool nonil = b1->tnonil && (b2 ? (bool) b2->tnonil : !cst_is_nil(r.type, cst));
	unsigned long long *nils_d = (unsigned long long *) dev_alloc(8);
	uint64_t nils = 0;
	bool ok = nils_d != nullptr && cuda_ok(cudaMemsetAsync(nils_d, 0, 8, c->stream), "memset nils");
	if (ok) {
		c->algo = "compare: cuda streaming";
		int rc = cmp_dispatch(l, r, ci1.ncand, op, nonil, nil_matches, (int8_t *) bn->base, nils_d);
		if (rc < 0) {
			set_err("%s: bad input types %s,%s.\n", func, atomname(b1->ttype), atomname(b2 ? b2->ttype : csttype));
			ok = false;
		} else
			ok = rc != 0 && cuda_ok(cudaMemcpyAsync(c->slot_h->v, nils_d, 8, cudaMemcpyDeviceToHost, c->stream), "read nils") &&
			     stream_sync();
	}
	if (nils_d)
		dev_free(nils_d);
	if (!ok) {
		b200_BATfree(bn);
		return nullptr;
	}
	nils = c->slot_h->v[0];
	const uint64_t n = ci1.ncand;
	bn->count = n;
	bn->tsorted = bn->trevsorted = n <= 1 || nils == n;	/* gdk_calc_compare.h:818-822 */
	bn->tkey = n <= 1;
	bn->tnil = nils != 0;
	bn->tnonil = nils == 0;
	return bn;
}

#define CMPOP(NAME, OP)                                                                                              \
	extern "C" b200_bat *b200_BATcalc##NAME(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1,          \
						const b200_bat *s2)                                                   \
	{                                                                                                            \
		b200_clrerr();                                                                                       \
		return compare_impl(OP, b1, b2, s1, s2, nullptr, 0, false, "BATcalc" #NAME);                         \
	}                                                                                                            \
	extern "C" b200_bat *b200_BATcalc##NAME##cst(const b200_bat *b, const void *v, int vtype, const b200_bat *s) \
	{                                                                                                            \
		b200_clrerr();                                                                                       \
		return compare_impl(OP, b, nullptr, s, nullptr, v, vtype, false, "BATcalc" #NAME "cst");             \
	}
CMPOP(lt, CMP_LT)
CMPOP(le, CMP_LE)
CMPOP(gt, CMP_GT)
CMPOP(ge, CMP_GE)

#define CMPOP_NM(NAME, OP)                                                                                           \
	extern "C" b200_bat *b200_BATcalc##NAME(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1,          \
						const b200_bat *s2, bool nil_matches)                                 \
	{                                                                                                            \
		b200_clrerr();                                                                                       \
		return compare_impl(OP, b1, b2, s1, s2, nullptr, 0, nil_matches, "BATcalc" #NAME);                   \
	}                                                                                                            \
	extern "C" b200_bat *b200_BATcalc##NAME##cst(const b200_bat *b, const void *v, int vtype, const b200_bat *s, \
						     bool nil_matches)                                               \
	{                                                                                                            \
		b200_clrerr();                                                                                       \
		return compare_impl(OP, b, nullptr, s, nullptr, v, vtype, nil_matches, "BATcalc" #NAME "cst");       \
	}
CMPOP_NM(eq, CMP_EQ)
CMPOP_NM(ne, CMP_NE)

// BATcalccst<op>(v, b, s) (BATcalccstop, gdk_calc_compare.h:928): the constant on the left is the mirrored
// operator with the constant on the right — the type matrix is symmetric and C converts both operands to
// their common type either way, so the bits agree.
#define CSTCMPOP(NAME, MIRROR)                                                                                       \
	extern "C" b200_bat *b200_BATcalccst##NAME(const void *v, int vtype, const b200_bat *b, const b200_bat *s)   \
	{                                                                                                            \
		b200_clrerr();                                                                                       \
		return compare_impl(MIRROR, b, nullptr, s, nullptr, v, vtype, false, "BATcalccst" #NAME);            \
	}
CSTCMPOP(lt, CMP_GT)
CSTCMPOP(le, CMP_GE)
CSTCMPOP(gt, CMP_LT)
CSTCMPOP(ge, CMP_LE)
#define CSTCMPOP_NM(NAME, OP)                                                                                        \
	extern "C" b200_bat *b200_BATcalccst##NAME(const void *v, int vtype, const b200_bat *b, const b200_bat *s,   \
						   bool nil_matches)                                                 \
	{                                                                                                            \
		b200_clrerr();                                                                                       \
		return compare_impl(OP, b, nullptr, s, nullptr, v, vtype, nil_matches, "BATcalccst" #NAME);          \
	}
CSTCMPOP_NM(eq, CMP_EQ)
CSTCMPOP_NM(ne, CMP_NE)

// ---------------------------------------------------------------- BATcalcisnil (gdk_calc.c:802-913)
template <typename T> static bool isnil_run(const CmpSide &s, uint64_t n, int8_t *dst, bool notnil)
{
	Ctx *c = ctx();
	CmpOperand<T> l = make_operand<T>(s);
	KTimer kt("isnil", n * (sizeof(T) + 1));
	isnil_kernel<T><<<grid_for((n + 3) / 4, CMP_THREADS, NSM * 16), CMP_THREADS, 0, c->stream>>>(l, n, dst, notnil ? 1 : 0);
	COUNT_LAUNCH();
	return cuda_ok(cudaGetLastError(), "isnil launch");
}

// BATcalcisnil_implementation(b, s, notnil), gdk_calc.c:802-913
static b200_bat *isnil_impl(const b200_bat *b, const b200_bat *s, bool notnil)
{
	b200_clrerr();
	if (b == nullptr) {
		set_err("%s: BAT required\n", notnil ? "BATcalcisnotnil" : "BATcalcisnil");
		return nullptr;
	}
	Staged sb, ss;
	if (!stage(b, sb) || !stage(s, ss))
		return nullptr;
	b = sb.b;
	Cand ci;
	if (!cand_init(ci, b, ss.b))
		return nullptr;
	const int8_t zero = notnil ? 1 : 0, one = notnil ? 0 : 1;
	if (b->tnonil || (b->ttype == B200_void && b->tseqbase != OID_NIL))
		return bat_constant(ci.rhseq, B200_bit, &zero, ci.ncand);
	if (b->ttype == B200_void)
		return bat_constant(ci.rhseq, B200_bit, &one, ci.ncand);
	b200_bat *bn = bat_new(ci.rhseq, B200_bit, ci.ncand);
	if (bn == nullptr)
		return nullptr;
	CmpSide l{b, &ci, nullptr, basetype(b->ttype)};
	bool ok = false;
	if (ci.ncand == 0)
		ok = true;
	else
		switch (l.type) {
		case B200_bte: ok = isnil_run<bte>(l, ci.ncand, (int8_t *) bn->base, notnil); break;
		case B200_sht: ok = isnil_run<sht>(l, ci.ncand, (int8_t *) bn->base, notnil); break;
		case B200_int: ok = isnil_run<int>(l, ci.ncand, (int8_t *) bn->base, notnil); break;
		case B200_lng: ok = isnil_run<lng>(l, ci.ncand, (int8_t *) bn->base, notnil); break;
		case B200_hge: ok = isnil_run<hge>(l, ci.ncand, (int8_t *) bn->base, notnil); break;
		case B200_flt: ok = isnil_run<flt>(l, ci.ncand, (int8_t *) bn->base, notnil); break;
		case B200_dbl: ok = isnil_run<dbl>(l, ci.ncand, (int8_t *) bn->base, notnil); break;
		default: set_err("BATcalcisnil: unsupported tail type %s\n", atomname(b->ttype));
		}
	if (!ok) {
		b200_BATfree(bn);
		return nullptr;
	}
	ctx()->algo = "isnil: cuda streaming";
	bn->count = ci.ncand;
	/* nils sort first: sorted input gives 1...0 for isnil, 0...1 for isnotnil (gdk_calc.c:886-896) */
	bn->tsorted = notnil ? b->tsorted : b->trevsorted;
	bn->trevsorted = notnil ? b->trevsorted : b->tsorted;
	bn->tnil = 0;
	bn->tnonil = 1;
	bn->tkey = ci.ncand <= 1;
	return bn;
}

extern "C" b200_bat *b200_BATcalcisnil(const b200_bat *b, const b200_bat *s) { return isnil_impl(b, s, false); }
extern "C" b200_bat *b200_BATcalcisnotnil(const b200_bat *b, const b200_bat *s) { return isnil_impl(b, s, true); }
