// calc_logic.cu — BATcalcand / or / xor and their constant forms (gdk_calc.c:2901,2967 / 2670,2736 /
// 2439,2506; loops and_typeswitchloop :2836, or_typeswitchloop :2600, xor_typeswitchloop :2376):
// three-valued logic on `bit` (and3 :2826, or3 :2590, XORBIT :2374), bitwise on the integer atoms
// with nil in, nil out; AND and XOR fail when two non-nil operands combine to the nil pattern
// (ON_OVERFLOW, gdk_calc_private.h:346).  One streaming pass; operands as in operand.cuh.
#include "common.cuh"
#include "operand.cuh"

#define LOGIC_THREADS 256
#define BIT_NIL ((int8_t) -128)

enum { L_AND = 0, L_OR = 1, L_XOR = 2 };
static const char *logic_name[] = {"AND", "OR", "XOR"};

template <int OP, typename T, bool ISBIT>
__global__ void __launch_bounds__(LOGIC_THREADS)
logic_kernel(CmpOperand<T> l, CmpOperand<T> r, uint64_t n, T *__restrict__ dst,
	     unsigned long long *__restrict__ state /* [0] nils, [1] first failing row */)
{
	unsigned nils = 0;
	unsigned long long fail = ~0ULL;
	for (uint64_t k = (uint64_t) blockIdx.x * LOGIC_THREADS + threadIdx.x; k < n; k += (uint64_t) gridDim.x * LOGIC_THREADS) {
		const T a = l.get(k), b = r.get(k);
		const bool n1 = Atom<T>::is_nil(a), n2 = Atom<T>::is_nil(b);
		T res;
		if constexpr (ISBIT && OP == L_AND) {
			res = (a == 0 || b == 0) ? 0 : (n1 || n2) ? BIT_NIL : 1;
			nils += res == BIT_NIL;
		} else if constexpr (ISBIT && OP == L_OR) {
			res = (a == 1 || b == 1) ? 1 : (n1 || n2) ? BIT_NIL : 0;
			nils += res == BIT_NIL;
		} else if (n1 || n2) {
			res = Atom<T>::nil();
			nils++;
		} else if constexpr (ISBIT) {
			res = (T) ((a == 0) != (b == 0));
		} else {
			res = OP == L_AND ? (T) (a & b) : OP == L_OR ? (T) (a | b) : (T) (a ^ b);
			if (OP != L_OR && Atom<T>::is_nil(res) && fail == ~0ULL)
				fail = k;
		}
		dst[k] = res;
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
		nils += __shfl_down_sync(0xffffffffu, nils, o);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(state, (unsigned long long) nils);
	if (fail != ~0ULL)
		atomicMin(state + 1, fail);
}

template <typename T> static void logic_fmt(char *buf, size_t sz, T v)
{
	/* FMTbte..FMThge, gdk_calc_private.h:300-306 */
	if constexpr (sizeof(T) == 16)
		snprintf(buf, sz, "%.40Lg (approx. value)", (long double) v);
	else if constexpr (sizeof(T) == 8)
		snprintf(buf, sz, "%lld", (long long) v);
	else
		snprintf(buf, sz, "%d", (int) v);
}

template <typename T> static bool logic_fetch(const CmpOperand<T> &o, uint64_t k, T &v)
{
	if (o.cst) {
		v = o.cval;
		return true;
	}
	uint64_t i = k;
	if (o.cand) {
		oid c;
		if (!cuda_ok(cudaMemcpy(&c, o.cand + k, sizeof(oid), cudaMemcpyDeviceToHost), "fetch candidate"))
			return false;
		i = c - o.hseq;
	}
	return cuda_ok(cudaMemcpy(&v, o.vals + i, sizeof(T), cudaMemcpyDeviceToHost), "fetch operand");
}

// returns false on failure (error text set); nils out
template <int OP, typename T, bool ISBIT>
static bool logic_run(const CmpSide &s1, const CmpSide &s2, uint64_t n, void *dst, uint64_t &nils)
{
	Ctx *c = ctx();
	CmpOperand<T> l = make_operand<T>(s1), r = make_operand<T>(s2);
	unsigned long long *state = (unsigned long long *) dev_alloc(16);
	static const unsigned long long init[2] = {0, ~0ULL};
	bool ok = state != nullptr && cuda_ok(cudaMemcpyAsync(state, init, 16, cudaMemcpyHostToDevice, c->stream), "logic state init");
	if (ok) {
		KTimer kt("calc_logic", n * ((s1.b ? 1 : 0) + (s2.b ? 1 : 0) + 1) * sizeof(T));
		logic_kernel<OP, T, ISBIT><<<grid_for(n, LOGIC_THREADS, NSM * 32), LOGIC_THREADS, 0, c->stream>>>(l, r, n, (T *) dst, state);
		COUNT_LAUNCH();
	}
	ok = ok && cuda_ok(cudaGetLastError(), "logic launch") &&
	     cuda_ok(cudaMemcpyAsync(c->slot_h->v, state, 16, cudaMemcpyDeviceToHost, c->stream), "read state") && stream_sync();
	dev_free(state);
	if (!ok)
		return false;
	nils = c->slot_h->v[0];
	const uint64_t fail = c->slot_h->v[1];
	if (fail != ~0ULL) {
		T a, b;
		if (!logic_fetch(l, fail, a) || !logic_fetch(r, fail, b))
			return false;
		char fa[64], fb[64];
		logic_fmt(fa, sizeof(fa), a);
		logic_fmt(fb, sizeof(fb), b);
		set_err("22003!overflow in calculation %s%s%s.\n", fa, logic_name[OP], fb);
		return false;
	}
	return true;
}

template <int OP>
static bool logic_dispatch(int ttype, const CmpSide &s1, const CmpSide &s2, uint64_t n, void *dst, uint64_t &nils, const char *func)
{
	switch (basetype(ttype)) {
	case B200_bte:
		return ttype == B200_bit ? logic_run<OP, bte, true>(s1, s2, n, dst, nils) : logic_run<OP, bte, false>(s1, s2, n, dst, nils);
	case B200_sht: return logic_run<OP, sht, false>(s1, s2, n, dst, nils);
	case B200_int: return logic_run<OP, int, false>(s1, s2, n, dst, nils);
	case B200_lng: return logic_run<OP, lng, false>(s1, s2, n, dst, nils);
	case B200_hge: return logic_run<OP, hge, false>(s1, s2, n, dst, nils);
	}
	set_err("%s: bad input type %s.\n", func, atomname(ttype));	/* gdk_calc.c:2428 */
	return false;
}

// mode 0: b1 OP b2; mode 1: b1 OP cst
static b200_bat *logic_impl(int op, int mode, const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2,
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
	if (basetype(b1->ttype) != basetype(mode == 0 ? b2->ttype : csttype)) {
		set_err("incompatible input types.\n");	/* gdk_calc.c:2451,2518 */
		return nullptr;
	}
	Cand ci1, ci2;
	if (!cand_init(ci1, b1, ss1.b))
		return nullptr;
	if (mode == 0) {
		if (!cand_init(ci2, b2, ss2.b))
			return nullptr;
		if (ci1.ncand != ci2.ncand || ci1.rhseq != ci2.rhseq) {
			set_err("inputs not the same size.\n");	/* gdk_calc.c:2459 */
			return nullptr;
		}
	}
	const uint64_t n = ci1.ncand;
	if (mode == 1 && b1->ttype == B200_bit && csttype == B200_bit) {
		/* true OR anything, false AND anything: a constant column (gdk_calc.c:2755, 2986) */
		const bte v = *(const bte *) cst;
		if ((op == L_OR && v == 1) || (op == L_AND && v == 0))
			return bat_constant(ci1.rhseq, B200_bit, &v, n);
	}
	b200_bat *bn = bat_new(ci1.rhseq, b1->ttype == B200_void ? B200_oid : b1->ttype, n);
	if (bn == nullptr || n == 0)
		return bn;
	if (b1->ttype == B200_void || (mode == 0 && b2->ttype == B200_void)) {
		b200_BATfree(bn);
		set_err("%s: void columns are not on the device path\n", func);
		return nullptr;
	}
	CmpSide l{b1, &ci1, nullptr, basetype(b1->ttype)};
	CmpSide r = mode == 0 ? CmpSide{b2, &ci2, nullptr, basetype(b2->ttype)} : CmpSide{nullptr, nullptr, cst, basetype(csttype)};
	ctx()->algo = "calc: cuda streaming and/or/xor";
	uint64_t nils = 0;
	bool ok;
	switch (op) {
	case L_AND: ok = logic_dispatch<L_AND>(b1->ttype, l, r, n, bn->base, nils, func); break;
	case L_OR: ok = logic_dispatch<L_OR>(b1->ttype, l, r, n, bn->base, nils, func); break;
	default: ok = logic_dispatch<L_XOR>(b1->ttype, l, r, n, bn->base, nils, func);
	}
	if (!ok) {
		b200_BATfree(bn);
		return nullptr;
	}
	/* gdk_calc.c:2484-2488 */
	bn->count = n;
	bn->tsorted = bn->trevsorted = n <= 1 || nils == n;
	bn->tkey = n <= 1;
	bn->tnil = nils != 0;
	bn->tnonil = nils == 0;
	return bn;
}

#define LOGICOP(NAME, OP)                                                                                            \
	extern "C" b200_bat *b200_BATcalc##NAME(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1,         \
						const b200_bat *s2)                                                   \
	{                                                                                                            \
		return logic_impl(OP, 0, b1, b2, s1, s2, nullptr, 0, "BATcalc" #NAME);                               \
	}                                                                                                            \
	extern "C" b200_bat *b200_BATcalc##NAME##cst(const b200_bat *b, const void *v, int vtype, const b200_bat *s) \
	{                                                                                                            \
		return logic_impl(OP, 1, b, nullptr, s, nullptr, v, vtype, "BATcalc" #NAME "cst");                   \
	}                                                                                                            \
	extern "C" b200_bat *b200_BATcalccst##NAME(const void *v, int vtype, const b200_bat *b, const b200_bat *s)   \
	{                                                                                                            \
		/* BATcalccst<op> forwards to BATcalc<op>cst: the operators commute (gdk_calc.c:2565,2801,3032) */   \
		return logic_impl(OP, 1, b, nullptr, s, nullptr, v, vtype, "BATcalc" #NAME "cst");                   \
	}
LOGICOP(and, L_AND)
LOGICOP(or, L_OR)
LOGICOP(xor, L_XOR)
