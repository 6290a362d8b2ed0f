// calc_misc.cu — the element-wise BATcalc operators SQL expressions need beside the arithmetic:
//   BATcalcmin / max / min_no_nil / max_no_nil   (gdk_calc.c:976,1194,1714,1890; loops MINMAX_TYPE :940-971,
//                                                  MINMAX_NONIL_TYPE :1150-1190)
//   BATcalcbetween + batcst / cstbat / cstcst      (gdk_calc.c:3968-4290; the BETWEEN macro :3801-3813 with
//                                                  its three-valued and3/or3/not3 :2590,2826,3795)
//   BATcalcifthenelse + cst forms                  (gdk_calc.c:4376-4800; loop IFTHENELSELOOP)
// One streaming pass each; operands are columns (dense or gathered through an oid candidate list) or
// constants (operand.cuh).  All three operands of an operator have the same storage type, as the
// reference requires.
#include "common.cuh"
#include "operand.cuh"

#define MISC_THREADS 256
#define BIT_NIL ((int8_t) -128)

// ---------------------------------------------------------------- min / max
template <typename T, bool MAX, bool NONIL>
__global__ void __launch_bounds__(MISC_THREADS)
minmax_kernel(CmpOperand<T> l, CmpOperand<T> r, uint64_t n, T *__restrict__ dst, unsigned *__restrict__ nilseen)
{
	bool nils = false;
	for (uint64_t k = (uint64_t) blockIdx.x * MISC_THREADS + threadIdx.x; k < n; k += (uint64_t) gridDim.x * MISC_THREADS) {
		const T p1 = l.get(k), p2 = r.get(k);
		const bool n1 = Atom<T>::is_nil(p1), n2 = Atom<T>::is_nil(p2);
		T res;
		if (NONIL) {
			/* a nil operand yields the other one (MINMAX_NONIL_TYPE) */
			if (n1) {
				res = p2;
				nils |= n2;
			} else {
				res = !n2 && (MAX ? p2 > p1 : p2 < p1) ? p2 : p1;
			}
		} else if (n1 || n2) {
			res = Atom<T>::nil();
			nils = true;
		} else {
			res = (MAX ? p1 > p2 : p1 < p2) ? p1 : p2;
		}
		dst[k] = res;
	}
	if (__any_sync(0xffffffffu, nils) && (threadIdx.x & 31) == 0)
		*nilseen = 1;
}

template <typename T>
static bool minmax_run(const CmpSide &s1, const CmpSide &s2, uint64_t n, bool max, bool nonil, void *dst, unsigned *nilseen)
{
	Ctx *c = ctx();
	CmpOperand<T> l = make_operand<T>(s1), r = make_operand<T>(s2);
	const unsigned g = grid_for(n, MISC_THREADS, NSM * 32);
	KTimer kt("calc_minmax", n * 3 * sizeof(T));
	if (max && nonil)
		minmax_kernel<T, true, true><<<g, MISC_THREADS, 0, c->stream>>>(l, r, n, (T *) dst, nilseen);
	else if (max)
		minmax_kernel<T, true, false><<<g, MISC_THREADS, 0, c->stream>>>(l, r, n, (T *) dst, nilseen);
	else if (nonil)
		minmax_kernel<T, false, true><<<g, MISC_THREADS, 0, c->stream>>>(l, r, n, (T *) dst, nilseen);
	else
		minmax_kernel<T, false, false><<<g, MISC_THREADS, 0, c->stream>>>(l, r, n, (T *) dst, nilseen);
	COUNT_LAUNCH();
	return cuda_ok(cudaGetLastError(), "min/max launch");
}

// cst != NULL: the BATcalc{min,max}cst[_no_nil] forms (gdk_calc.c:1403,1563,2086,2219), b2 unused
static b200_bat *minmax_impl(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2, bool max,
			     bool nonil, const char *func, const void *cst = nullptr, int csttype = 0)
{
	b200_clrerr();
	if (b1 == nullptr || (cst == nullptr && b2 == nullptr)) {
		set_err("%s: BAT required\n", func);
		return nullptr;
	}
	Staged sb1, sb2, ss1, ss2;
	if (!stage(b1, sb1) || !stage(b2, sb2) || !stage(s1, ss1) || !stage(s2, ss2))
		return nullptr;
	b1 = sb1.b;
	b2 = sb2.b;
	if (cst ? (b1->ttype == B200_void ? B200_oid : b1->ttype) != csttype : basetype(b1->ttype) != basetype(b2->ttype)) {
		set_err("inputs have incompatible types\n");	/* gdk_calc.c:993,1423 */
		return nullptr;
	}
	Cand ci1, ci2;
	if (!cand_init(ci1, b1, ss1.b) || (!cst && !cand_init(ci2, b2, ss2.b)))
		return nullptr;
	if (!cst && (ci1.ncand != ci2.ncand || ci1.rhseq != ci2.rhseq)) {
		set_err("inputs not the same size.\n");	/* gdk_calc.c:1000 */
		return nullptr;
	}
	const uint64_t n = ci1.ncand;
	if (cst && (n == 0 || (!nonil && cst_is_nil(basetype(csttype), cst))))
		return bat_constant(ci1.rhseq, b1->ttype, nullptr, n);	/* BATconstantV(nil), gdk_calc.c:1429-1432 */
	b200_bat *bn = bat_new(ci1.rhseq, b1->ttype == B200_void ? B200_oid : b1->ttype, n);
	if (bn == nullptr)
		return nullptr;
	bool nils = false;
	if (n > 0) {
		if (b1->ttype == B200_void || (!cst && b2->ttype == B200_void)) {
			b200_BATfree(bn);
			set_err("%s: void columns are not on the device path\n", func);
			return nullptr;
		}
		Ctx *c = ctx();
		CmpSide l{b1, &ci1, nullptr, basetype(b1->ttype)};
		CmpSide r = cst ? CmpSide{nullptr, nullptr, cst, basetype(csttype)} : CmpSide{b2, &ci2, nullptr, basetype(b2->ttype)};
		unsigned *flag = (unsigned *) dev_alloc(4);
		bool ok = flag != nullptr && cuda_ok(cudaMemsetAsync(flag, 0, 4, c->stream), "memset flag");
		if (ok) {
			c->algo = "calc: cuda streaming min/max";
			switch (l.type) {
			case B200_bte: ok = minmax_run<bte>(l, r, n, max, nonil, bn->base, flag); break;
			case B200_sht: ok = minmax_run<sht>(l, r, n, max, nonil, bn->base, flag); break;
			case B200_int: ok = minmax_run<int>(l, r, n, max, nonil, bn->base, flag); break;
			case B200_lng: ok = minmax_run<lng>(l, r, n, max, nonil, bn->base, flag); break;
			case B200_hge: ok = minmax_run<hge>(l, r, n, max, nonil, bn->base, flag); break;
			case B200_flt: ok = minmax_run<flt>(l, r, n, max, nonil, bn->base, flag); break;
			case B200_dbl: ok = minmax_run<dbl>(l, r, n, max, nonil, bn->base, flag); break;
			default:
				set_err("%s: unsupported tail type %s (fixed-width atoms only)\n", func, atomname(b1->ttype));
				ok = false;
			}
			ok = ok && cuda_ok(cudaMemcpyAsync(c->slot_h->v, flag, 4, cudaMemcpyDeviceToHost, c->stream), "read flag") &&
			     stream_sync();
		}
		dev_free(flag);
		if (!ok) {
			b200_BATfree(bn);
			return nullptr;
		}
		nils = (unsigned) c->slot_h->v[0] != 0;
	}
	/* gdk_calc.c:1117-1130 */
	bn->count = n;
	bn->tnil = nils;
	bn->tnonil = !nils;
	bn->tsorted = bn->trevsorted = bn->tkey = n <= 1;
	return bn;
}

extern "C" b200_bat *b200_BATcalcmin(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2)
{
	return minmax_impl(b1, b2, s1, s2, false, false, "BATcalcmin");
}
extern "C" b200_bat *b200_BATcalcmin_no_nil(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2)
{
	return minmax_impl(b1, b2, s1, s2, false, true, "BATcalcmin_no_nil");
}
extern "C" b200_bat *b200_BATcalcmax(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2)
{
	return minmax_impl(b1, b2, s1, s2, true, false, "BATcalcmax");
}
extern "C" b200_bat *b200_BATcalcmax_no_nil(const b200_bat *b1, const b200_bat *b2, const b200_bat *s1, const b200_bat *s2)
{
	return minmax_impl(b1, b2, s1, s2, true, true, "BATcalcmax_no_nil");
}

#define MINMAXCST(NAME, MAX, NONIL)                                                                                   \
	extern "C" b200_bat *b200_BATcalc##NAME(const b200_bat *b, const void *v, int vtype, const b200_bat *s)       \
	{                                                                                                             \
		if (v == nullptr) {                                                                                   \
			b200_clrerr();                                                                                \
			set_err("BATcalc" #NAME ": value required\n");                                                \
			return nullptr;                                                                               \
		}                                                                                                     \
		return minmax_impl(b, nullptr, s, nullptr, MAX, NONIL, "BATcalc" #NAME, v, vtype);                    \
	}
MINMAXCST(mincst, false, false)
MINMAXCST(mincst_no_nil, false, true)
MINMAXCST(maxcst, true, false)
MINMAXCST(maxcst_no_nil, true, true)

// ---------------------------------------------------------------- between
// three-valued logic on bit {0, 1, bit_nil} (gdk_calc.c:2590 or3, :2826 and3, :3795 not3)
__device__ __forceinline__ int8_t d_and3(int8_t a, int8_t b) { return a == 0 || b == 0 ? 0 : (a == BIT_NIL || b == BIT_NIL ? BIT_NIL : 1); }
__device__ __forceinline__ int8_t d_or3(int8_t a, int8_t b) { return a == 1 || b == 1 ? 1 : (a == BIT_NIL || b == BIT_NIL ? BIT_NIL : 0); }
__device__ __forceinline__ int8_t d_not3(int8_t a) { return a == BIT_NIL ? BIT_NIL : !a; }
template <typename T> __device__ __forceinline__ int8_t d_less3(T a, T b, bool inc)
{
	return Atom<T>::is_nil(a) || Atom<T>::is_nil(b) ? BIT_NIL : (int8_t) (a < b || (inc && a == b));
}
template <typename T> __device__ __forceinline__ int8_t d_grtr3(T a, T b, bool inc)
{
	return Atom<T>::is_nil(a) || Atom<T>::is_nil(b) ? BIT_NIL : (int8_t) (b < a || (inc && a == b));
}
template <typename T> __device__ __forceinline__ int8_t d_between3(T v, T lo, bool linc, T hi, bool hinc)
{
	return d_and3(d_grtr3<T>(v, lo, linc), d_less3<T>(v, hi, hinc));
}

enum { BW_SYMMETRIC = 1, BW_LINC = 2, BW_HINC = 4, BW_NILS_FALSE = 8, BW_ANTI = 16 };

template <typename T>
__global__ void __launch_bounds__(MISC_THREADS)
between_kernel(CmpOperand<T> b, CmpOperand<T> lo, CmpOperand<T> hi, uint64_t n, unsigned mode, int8_t *__restrict__ dst,
	       unsigned long long *__restrict__ nils_out)
{
	const bool sym = mode & BW_SYMMETRIC, linc = mode & BW_LINC, hinc = mode & BW_HINC, anti = mode & BW_ANTI;
	unsigned nils = 0;
	const uint64_t nquads = (n + 3) / 4;
	for (uint64_t q = (uint64_t) blockIdx.x * MISC_THREADS + threadIdx.x; q < nquads; q += (uint64_t) gridDim.x * MISC_THREADS) {
		int8_t res[4] = {0, 0, 0, 0};
#pragma unroll
		for (int e = 0; e < 4; e++) {
			const uint64_t k = 4 * q + e;
			if (k >= n)
				break;
			const T v = b.get(k), l = lo.get(k), h = hi.get(k);
			int8_t r;
			if (Atom<T>::is_nil(v)) {
				r = (mode & BW_NILS_FALSE) ? 0 : BIT_NIL;
			} else {
				r = d_between3<T>(v, l, linc, h, hinc);
				if (sym)
					r = d_or3(r, d_between3<T>(v, h, hinc, l, linc));
				if (anti)
					r = d_not3(r);
			}
			nils += r == BIT_NIL;
			res[e] = r;
		}
		if (4 * q + 3 < n)
			*reinterpret_cast<char4 *>(dst + 4 * q) = make_char4(res[0], res[1], res[2], res[3]);
		else
			for (int e = 0; e < 4 && 4 * q + e < n; e++)
				dst[4 * q + e] = res[e];
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
		nils += __shfl_down_sync(0xffffffffu, nils, o);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(nils_out, (unsigned long long) nils);
}

template <typename T>
static bool between_run(const CmpSide &sb, const CmpSide &sl, const CmpSide &sh, uint64_t n, unsigned mode, int8_t *dst,
			unsigned long long *nils_d)
{
	Ctx *c = ctx();
	KTimer kt("calc_between", n * (sizeof(T) * (1 + (sl.b != nullptr) + (sh.b != nullptr)) + 1));
	between_kernel<T><<<grid_for((n + 3) / 4, MISC_THREADS, NSM * 16), MISC_THREADS, 0, c->stream>>>(
		make_operand<T>(sb), make_operand<T>(sl), make_operand<T>(sh), n, mode, dst, nils_d);
	COUNT_LAUNCH();
	return cuda_ok(cudaGetLastError(), "between launch");
}

// lo / hi: a BAT (with optional candidates) or, when the BAT is NULL, the constant (cst, csttype)
static b200_bat *between_impl(const b200_bat *b, const b200_bat *lo, const b200_bat *hi, const b200_bat *s, const b200_bat *slo,
			      const b200_bat *shi, const void *clo, int clotype, const void *chi, int chitype, unsigned mode,
			      const char *func)
{
	b200_clrerr();
	if (b == nullptr || (lo == nullptr && clo == nullptr) || (hi == nullptr && chi == nullptr)) {
		set_err("%s: BAT required\n", func);
		return nullptr;
	}
	Staged sb, slb, shb, ss, sslo, sshi;
	if (!stage(b, sb) || !stage(lo, slb) || !stage(hi, shb) || !stage(s, ss) || !stage(slo, sslo) || !stage(shi, sshi))
		return nullptr;
	b = sb.b;
	lo = slb.b;
	hi = shb.b;
	const int t = basetype(b->ttype);
	if (t != basetype(lo ? lo->ttype : clotype) || t != basetype(hi ? hi->ttype : chitype)) {
		set_err("incompatible input types.\n");	/* gdk_calc.c:4073, 4141, 4221 */
		return nullptr;
	}
	Cand ci, cilo, cihi;
	if (!cand_init(ci, b, ss.b) || (lo && !cand_init(cilo, lo, sslo.b)) || (hi && !cand_init(cihi, hi, sshi.b)))
		return nullptr;
	if ((lo && (ci.ncand != cilo.ncand || ci.rhseq != cilo.rhseq)) || (hi && (ci.ncand != cihi.ncand || ci.rhseq != cihi.rhseq))) {
		set_err("inputs not the same size.\n");	/* gdk_calc.c:3986 */
		return nullptr;
	}
	const uint64_t n = ci.ncand;
	b200_bat *bn = bat_new(ci.rhseq, B200_bit, n);
	if (bn == nullptr || n == 0)
		return bn;
	if (b->ttype == B200_void || (lo && lo->ttype == B200_void) || (hi && hi->ttype == B200_void)) {
		b200_BATfree(bn);
		set_err("%s: void columns are not on the device path\n", func);
		return nullptr;
	}
	Ctx *c = ctx();
	CmpSide vb{b, &ci, nullptr, t};
	CmpSide vl = lo ? CmpSide{lo, &cilo, nullptr, t} : CmpSide{nullptr, nullptr, clo, t};
	CmpSide vh = hi ? CmpSide{hi, &cihi, nullptr, t} : CmpSide{nullptr, nullptr, chi, t};
	unsigned long long *nils_d = (unsigned long long *) dev_alloc(8);
	bool ok = nils_d != nullptr && cuda_ok(cudaMemsetAsync(nils_d, 0, 8, c->stream), "memset nils");
	if (ok) {
		c->algo = "calc: cuda streaming between";
		int8_t *dst = (int8_t *) bn->base;
		switch (t) {
		case B200_bte: ok = between_run<bte>(vb, vl, vh, n, mode, dst, nils_d); break;
		case B200_sht: ok = between_run<sht>(vb, vl, vh, n, mode, dst, nils_d); break;
		case B200_int: ok = between_run<int>(vb, vl, vh, n, mode, dst, nils_d); break;
		case B200_lng: ok = between_run<lng>(vb, vl, vh, n, mode, dst, nils_d); break;
		case B200_hge: ok = between_run<hge>(vb, vl, vh, n, mode, dst, nils_d); break;
		case B200_flt: ok = between_run<flt>(vb, vl, vh, n, mode, dst, nils_d); break;
		case B200_dbl: ok = between_run<dbl>(vb, vl, vh, n, mode, dst, nils_d); break;
		default:
			set_err("%s: bad input type %s.\n", func, atomname(b->ttype));	/* gdk_calc.c:3928 */
			ok = false;
		}
		ok = ok && cuda_ok(cudaMemcpyAsync(c->slot_h->v, nils_d, 8, cudaMemcpyDeviceToHost, c->stream), "read nils") && stream_sync();
	}
	dev_free(nils_d);
	if (!ok) {
		b200_BATfree(bn);
		return nullptr;
	}
	const uint64_t nils = c->slot_h->v[0];
	/* gdk_calc.c:3947-3953 */
	bn->count = n;
	bn->tsorted = bn->trevsorted = n <= 1 || nils == n;
	bn->tkey = n <= 1;
	bn->tnil = nils != 0;
	bn->tnonil = nils == 0;
	return bn;
}

static inline unsigned bw_mode(bool symmetric, bool linc, bool hinc, bool nils_false, bool anti)
{
	return (symmetric ? BW_SYMMETRIC : 0) | (linc ? BW_LINC : 0) | (hinc ? BW_HINC : 0) | (nils_false ? BW_NILS_FALSE : 0) |
	       (anti ? BW_ANTI : 0);
}

extern "C" b200_bat *b200_BATcalcbetween(const b200_bat *b, const b200_bat *lo, const b200_bat *hi, const b200_bat *s,
					 const b200_bat *slo, const b200_bat *shi, bool symmetric, bool linc, bool hinc,
					 bool nils_false, bool anti)
{
	if (lo == nullptr || hi == nullptr) {
		b200_clrerr();
		set_err("BATcalcbetween: BAT required\n");
		return nullptr;
	}
	return between_impl(b, lo, hi, s, slo, shi, nullptr, 0, nullptr, 0, bw_mode(symmetric, linc, hinc, nils_false, anti), "BATcalcbetween");
}
extern "C" b200_bat *b200_BATcalcbetweencstcst(const b200_bat *b, const void *lo, int lotype, const void *hi, int hitype,
					       const b200_bat *s, bool symmetric, bool linc, bool hinc, bool nils_false, bool anti)
{
	return between_impl(b, nullptr, nullptr, s, nullptr, nullptr, lo, lotype, hi, hitype,
			    bw_mode(symmetric, linc, hinc, nils_false, anti), "BATcalcbetweencstcst");
}
extern "C" b200_bat *b200_BATcalcbetweenbatcst(const b200_bat *b, const b200_bat *lo, const void *hi, int hitype, const b200_bat *s,
					       const b200_bat *slo, bool symmetric, bool linc, bool hinc, bool nils_false, bool anti)
{
	return between_impl(b, lo, nullptr, s, slo, nullptr, nullptr, 0, hi, hitype, bw_mode(symmetric, linc, hinc, nils_false, anti),
			    "BATcalcbetweenbatcst");
}
extern "C" b200_bat *b200_BATcalcbetweencstbat(const b200_bat *b, const void *lo, int lotype, const b200_bat *hi, const b200_bat *s,
					       const b200_bat *shi, bool symmetric, bool linc, bool hinc, bool nils_false, bool anti)
{
	return between_impl(b, nullptr, hi, s, nullptr, shi, lo, lotype, nullptr, 0, bw_mode(symmetric, linc, hinc, nils_false, anti),
			    "BATcalcbetweencstbat");
}

// ---------------------------------------------------------------- ifthenelse
// dst[i] = (b[i] != 0 && b[i] != bit_nil) ? then[i] : else[i]  (IFTHENELSELOOP, gdk_calc.c:4340-4373); values are
// moved by width, never interpreted, so one instantiation per atom size serves every type.
template <typename W>
__global__ void __launch_bounds__(MISC_THREADS)
ifthenelse_kernel(const int8_t *__restrict__ cond, CmpOperand<W> t, CmpOperand<W> e, uint64_t n, W *__restrict__ dst)
{
	for (uint64_t k = (uint64_t) blockIdx.x * MISC_THREADS + threadIdx.x; k < n; k += (uint64_t) gridDim.x * MISC_THREADS) {
		const int8_t c = ld_stream(cond + k);
		/* both sides are read: the loads are independent of the condition and stay coalesced */
		const W a = t.get(k), b = e.get(k);
		dst[k] = (c != 0 && c != BIT_NIL) ? a : b;
	}
}

template <typename W>
static bool ifthenelse_run(const int8_t *cond, const CmpSide &s1, const CmpSide &s2, uint64_t n, void *dst)
{
	Ctx *c = ctx();
	KTimer kt("calc_ifthenelse", n * (1 + sizeof(W) * (1 + (s1.b != nullptr) + (s2.b != nullptr))));
	ifthenelse_kernel<W><<<grid_for(n, MISC_THREADS, NSM * 32), MISC_THREADS, 0, c->stream>>>(cond, make_operand<W>(s1), make_operand<W>(s2), n,
												      (W *) dst);
	COUNT_LAUNCH();
	return cuda_ok(cudaGetLastError(), "ifthenelse launch");
}

// b1 / b2: a BAT or, when NULL, the constant (c1/c2 with its type)
static b200_bat *ifthenelse_impl(const b200_bat *b, const b200_bat *b1, const b200_bat *b2, const void *c1, int c1type,
				 const void *c2, int c2type, const char *func)
{
	b200_clrerr();
	if (b == nullptr || (b1 == nullptr && c1 == nullptr) || (b2 == nullptr && c2 == nullptr)) {
		set_err("%s: BAT required\n", func);
		return nullptr;
	}
	Staged sb, sb1, sb2;
	if (!stage(b, sb) || !stage(b1, sb1) || !stage(b2, sb2))
		return nullptr;
	b = sb.b;
	b1 = sb1.b;
	b2 = sb2.b;
	if ((b1 && b1->count != b->count) || (b2 && b2->count != b->count)) {
		set_err("%s: inputs not the same size.\n", func);	/* checkbats, gdk_calc.c:25-31 */
		return nullptr;
	}
	const int t1 = b1 ? b1->ttype : c1type, t2 = b2 ? b2->ttype : c2type;
	if (b->ttype != B200_bit || basetype(t1) != basetype(t2)) {
		set_err("\"then\" and \"else\" BATs have different types.\n");	/* gdk_calc.c:4676 */
		return nullptr;
	}
	const uint64_t n = b->count;
	b200_bat *bn = bat_new(b->hseqbase, t1 == B200_void ? B200_oid : t1, n);
	if (bn == nullptr || n == 0)
		return bn;
	if ((b1 && b1->ttype == B200_void) || (b2 && b2->ttype == B200_void) || b->base == nullptr) {
		b200_BATfree(bn);
		set_err("%s: void columns are not on the device path\n", func);
		return nullptr;
	}
	const int st = basetype(t1);
	Cand ci1, ci2;
	if ((b1 && !cand_init(ci1, b1, nullptr)) || (b2 && !cand_init(ci2, b2, nullptr))) {
		b200_BATfree(bn);
		return nullptr;
	}
	CmpSide s1 = b1 ? CmpSide{b1, &ci1, nullptr, st} : CmpSide{nullptr, nullptr, c1, st};
	CmpSide s2 = b2 ? CmpSide{b2, &ci2, nullptr, st} : CmpSide{nullptr, nullptr, c2, st};
	const int8_t *cond = (const int8_t *) b->base;
	bool ok;
	ctx()->algo = "calc: cuda streaming ifthenelse";
	switch (atomsize(st)) {
	case 1: ok = ifthenelse_run<bte>(cond, s1, s2, n, bn->base); break;
	case 2: ok = ifthenelse_run<sht>(cond, s1, s2, n, bn->base); break;
	case 4: ok = ifthenelse_run<int>(cond, s1, s2, n, bn->base); break;
	case 8: ok = ifthenelse_run<lng>(cond, s1, s2, n, bn->base); break;
	case 16: ok = ifthenelse_run<hge>(cond, s1, s2, n, bn->base); break;
	default:
		set_err("%s: unsupported tail type %s (fixed-width atoms only)\n", func, atomname(t1));
		ok = false;
	}
	if (!ok) {
		b200_BATfree(bn);
		return nullptr;
	}
	/* gdk_calc.c:4636-4643 */
	bn->count = n;
	bn->tsorted = bn->trevsorted = bn->tkey = n <= 1;
	bn->tnil = 0;
	bn->tnonil = (b1 ? (bool) b1->tnonil : !cst_is_nil(st, c1)) && (b2 ? (bool) b2->tnonil : !cst_is_nil(st, c2));
	return bn;
}

extern "C" b200_bat *b200_BATcalcifthenelse(const b200_bat *b, const b200_bat *b1, const b200_bat *b2)
{
	if (b1 == nullptr || b2 == nullptr) {
		b200_clrerr();
		set_err("BATcalcifthenelse: BAT required\n");
		return nullptr;
	}
	return ifthenelse_impl(b, b1, b2, nullptr, 0, nullptr, 0, "BATcalcifthenelse");
}
extern "C" b200_bat *b200_BATcalcifthenelsecst(const b200_bat *b, const b200_bat *b1, const void *c2, int c2type)
{
	return ifthenelse_impl(b, b1, nullptr, nullptr, 0, c2, c2type, "BATcalcifthenelsecst");
}
extern "C" b200_bat *b200_BATcalcifthencstelse(const b200_bat *b, const void *c1, int c1type, const b200_bat *b2)
{
	return ifthenelse_impl(b, nullptr, b2, c1, c1type, nullptr, 0, "BATcalcifthencstelse");
}
extern "C" b200_bat *b200_BATcalcifthencstelsecst(const b200_bat *b, const void *c1, int c1type, const void *c2, int c2type)
{
	return ifthenelse_impl(b, nullptr, nullptr, c1, c1type, c2, c2type, "BATcalcifthencstelsecst");
}

// ---------------------------------------------------------------- unary operators
// BATcalcnot / negate / absolute / iszero / sign (gdk_calc.c:41,233,371,513,652; loops UNARY_2TYPE_FUNC[_nilcheck],
// gdk_calc_private.h:354-388): nil in, nil out; NOT on an integer fails when ~x is the nil pattern.
enum { U_NOT = 0, U_NEGATE, U_ABSOLUTE, U_ISZERO, U_SIGN };

template <int OP, typename TI, typename TO, bool ISBIT>
__global__ void __launch_bounds__(MISC_THREADS)
unary_kernel(CmpOperand<TI> src, uint64_t n, TO *__restrict__ dst, unsigned long long *__restrict__ state /* [0] nils, [1] overflow */)
{
	unsigned nils = 0;
	bool ovf = false;
	for (uint64_t k = (uint64_t) blockIdx.x * MISC_THREADS + threadIdx.x; k < n; k += (uint64_t) gridDim.x * MISC_THREADS) {
		const TI x = src.get(k);
		TO r;
		if (Atom<TI>::is_nil(x)) {
			r = Atom<TO>::nil();
			nils++;
		} else if constexpr (OP == U_NOT) {
			if constexpr (ISBIT) {
				r = (TO) !x;
			} else {
				r = (TO) ~x;
				ovf |= Atom<TO>::is_nil(r);
			}
		} else if constexpr (OP == U_NEGATE) {
			r = (TO) -x;
		} else if constexpr (OP == U_ABSOLUTE) {
			/* abs / llabs / ABSOLUTE on integers, fabsf / fabs on floats: -0.0 becomes +0.0 (gdk_calc.c:400-420) */
			if constexpr (std::is_same<TI, flt>::value)
				r = (TO) fabsf(x);
			else if constexpr (std::is_same<TI, dbl>::value)
				r = (TO) fabs(x);
			else
				r = (TO) (x < 0 ? -x : x);
		} else if constexpr (OP == U_ISZERO) {
			r = (TO) (x == 0);
		} else {
			r = (TO) (x < 0 ? -1 : x > 0);
		}
		dst[k] = r;
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
		nils += __shfl_down_sync(0xffffffffu, nils, o);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(state, (unsigned long long) nils);
	if (ovf)
		state[1] = 1;
}

template <int OP, typename TI, typename TO, bool ISBIT = false>
static bool unary_run(const CmpSide &s, uint64_t n, void *dst, unsigned long long *state)
{
	Ctx *c = ctx();
	KTimer kt("calc_unary", n * (sizeof(TI) + sizeof(TO)));
	unary_kernel<OP, TI, TO, ISBIT><<<grid_for(n, MISC_THREADS, NSM * 32), MISC_THREADS, 0, c->stream>>>(make_operand<TI>(s), n, (TO *) dst, state);
	COUNT_LAUNCH();
	return cuda_ok(cudaGetLastError(), "unary launch");
}

template <int OP> static int unary_dispatch(const CmpSide &s, int ttype, uint64_t n, void *dst, unsigned long long *state)
{
	if constexpr (OP == U_NOT) {
		switch (s.type) {
		case B200_bte: return ttype == B200_bit ? unary_run<OP, bte, bte, true>(s, n, dst, state) : unary_run<OP, bte, bte>(s, n, dst, state);
		case B200_sht: return unary_run<OP, sht, sht>(s, n, dst, state);
		case B200_int: return unary_run<OP, int, int>(s, n, dst, state);
		case B200_lng: return unary_run<OP, lng, lng>(s, n, dst, state);
		case B200_hge: return unary_run<OP, hge, hge>(s, n, dst, state);
		}
		return -1;
	} else if constexpr (OP == U_NEGATE || OP == U_ABSOLUTE) {
		switch (s.type) {
		case B200_bte: return unary_run<OP, bte, bte>(s, n, dst, state);
		case B200_sht: return unary_run<OP, sht, sht>(s, n, dst, state);
		case B200_int: return unary_run<OP, int, int>(s, n, dst, state);
		case B200_lng: return unary_run<OP, lng, lng>(s, n, dst, state);
		case B200_hge: return unary_run<OP, hge, hge>(s, n, dst, state);
		case B200_flt: return unary_run<OP, flt, flt>(s, n, dst, state);
		case B200_dbl: return unary_run<OP, dbl, dbl>(s, n, dst, state);
		}
		return -1;
	} else {
		switch (s.type) {
		case B200_bte: return unary_run<OP, bte, bte>(s, n, dst, state);
		case B200_sht: return unary_run<OP, sht, bte>(s, n, dst, state);
		case B200_int: return unary_run<OP, int, bte>(s, n, dst, state);
		case B200_lng: return unary_run<OP, lng, bte>(s, n, dst, state);
		case B200_hge: return unary_run<OP, hge, bte>(s, n, dst, state);
		case B200_flt: return unary_run<OP, flt, bte>(s, n, dst, state);
		case B200_dbl: return unary_run<OP, dbl, bte>(s, n, dst, state);
		}
		return -1;
	}
}

static b200_bat *unary_impl(int op, const b200_bat *b, const b200_bat *s, const char *func)
{
	b200_clrerr();
	if (b == nullptr) {
		set_err("%s: BAT required\n", func);
		return nullptr;
	}
	Staged sb, ss;
	if (!stage(b, sb) || !stage(s, ss))
		return nullptr;
	b = sb.b;
	Cand ci;
	if (!cand_init(ci, b, ss.b))
		return nullptr;
	const uint64_t n = ci.ncand;
	const int st = basetype(b->ttype);
	const int rt = op == U_ISZERO ? B200_bit : op == U_SIGN ? B200_bte : (b->ttype == B200_void ? B200_oid : b->ttype);
	b200_bat *bn = bat_new(ci.rhseq, rt, n);
	if (bn == nullptr || n == 0)
		return bn;
	if (b->ttype == B200_void) {
		b200_BATfree(bn);
		set_err("%s: void columns are not on the device path\n", func);
		return nullptr;
	}
	Ctx *c = ctx();
	CmpSide side{b, &ci, nullptr, st};
	unsigned long long *state = (unsigned long long *) dev_alloc(16);
	bool ok = state != nullptr && cuda_ok(cudaMemsetAsync(state, 0, 16, c->stream), "memset state");
	if (ok) {
		c->algo = "calc: cuda streaming unary";
		int rc;
		switch (op) {
		case U_NOT: rc = unary_dispatch<U_NOT>(side, b->ttype, n, bn->base, state); break;
		case U_NEGATE: rc = unary_dispatch<U_NEGATE>(side, b->ttype, n, bn->base, state); break;
		case U_ABSOLUTE: rc = unary_dispatch<U_ABSOLUTE>(side, b->ttype, n, bn->base, state); break;
		case U_ISZERO: rc = unary_dispatch<U_ISZERO>(side, b->ttype, n, bn->base, state); break;
		default: rc = unary_dispatch<U_SIGN>(side, b->ttype, n, bn->base, state);
		}
		if (rc < 0) {
			if (op == U_ISZERO || op == U_SIGN)
				set_err("bad input type %s.\n", atomname(b->ttype));	/* gdk_calc.c:563,702 */
			else
				set_err("type %s not supported.\n", atomname(b->ttype));	/* gdk_calc.c:118,284 */
			ok = false;
		} else {
			ok = rc != 0 && cuda_ok(cudaMemcpyAsync(c->slot_h->v, state, 16, cudaMemcpyDeviceToHost, c->stream), "read state") &&
			     stream_sync();
		}
	}
	dev_free(state);
	if (ok && c->slot_h->v[1] != 0) {
		/* ON_OVERFLOW1(TYPE, "NOT"): ~x is the nil pattern only for x = the type's maximum */
		static const char *maxs[] = {"127", "32767", "2147483647", "9223372036854775807", "170141183460469231731687303715884105727"};
		const int w = atomsize(st);
		set_err("22003!overflow in calculation NOT(%s).\n", maxs[w == 1 ? 0 : w == 2 ? 1 : w == 4 ? 2 : w == 8 ? 3 : 4]);
		ok = false;
	}
	if (!ok) {
		b200_BATfree(bn);
		return nullptr;
	}
	const uint64_t nils = c->slot_h->v[0];
	bn->count = n;
	bn->tnil = nils != 0;
	bn->tnonil = nils == 0;
	if (op == U_NOT || op == U_NEGATE) {
		/* reverses the order, but nils mess it up (gdk_calc.c:124-130, 290-296) */
		bn->tsorted = nils == 0 && b->trevsorted;
		bn->trevsorted = nils == 0 && b->tsorted;
		bn->tkey = b->tkey && nils <= 1;
	} else if (op == U_SIGN) {
		bn->tsorted = b->tsorted || n <= 1 || nils == n;	/* gdk_calc.c:711-713 */
		bn->trevsorted = b->trevsorted || n <= 1 || nils == n;
		bn->tkey = n <= 1;
	} else {
		bn->tsorted = bn->trevsorted = n <= 1 || nils == n;	/* gdk_calc.c:433-435, 569-571 */
		bn->tkey = n <= 1;
	}
	return bn;
}

extern "C" b200_bat *b200_BATcalcnot(const b200_bat *b, const b200_bat *s) { return unary_impl(U_NOT, b, s, "BATcalcnot"); }
extern "C" b200_bat *b200_BATcalcnegate(const b200_bat *b, const b200_bat *s) { return unary_impl(U_NEGATE, b, s, "BATcalcnegate"); }
extern "C" b200_bat *b200_BATcalcabsolute(const b200_bat *b, const b200_bat *s) { return unary_impl(U_ABSOLUTE, b, s, "BATcalcabsolute"); }
extern "C" b200_bat *b200_BATcalciszero(const b200_bat *b, const b200_bat *s) { return unary_impl(U_ISZERO, b, s, "BATcalciszero"); }
extern "C" b200_bat *b200_BATcalcsign(const b200_bat *b, const b200_bat *s) { return unary_impl(U_SIGN, b, s, "BATcalcsign"); }
