// convert.cu — BATconvert for the fixed-width numeric atoms, including the decimal rescaling of SQL
// casts: scale1 / scale2 / precision (gdk_calc_convert.c:1415-1540; kernels convertimpl :262-363
// with its DIVIDE rounding :256-260, convertimpl_enlarge_float :98-146, convertimpl_reduce_float
// :365-424 whose `(long double) v * mul` + roundl is reproduced with exact 64-bit-mantissa integer
// arithmetic — the device has no x87).
// Used directly by SQL casts and by the singleton-group shortcut of
// BATgroupsum / BATgroupavg (gdk_aggr.c:929-935,1853-1860).
#include "common.cuh"
#include <type_traits>

enum { CV_OK = 0, CV_OVERFLOW = 1 };

// the reference's scales[] table (gdk_calc_convert.c:32-72): 10^0 .. 10^38
struct ConvScale {
	hge div;	/* 10^(scale1 - scale2) when scale1 > scale2, else 1 (int -> int); 10^scale1 (int -> float) */
	hge mul;	/* 10^(scale2 - scale1) when scale2 > scale1, else 1 (int -> int); 10^scale2 (float -> int) */
	hge prec;	/* 10^precision, 0 when precision == 0 */
};

// (long double) v * mul, then roundl (round half away from zero), as x87 computes it: mul is first rounded to a
// 64-bit mantissa, the product again; the result is returned as magnitude + sign, `big` when it cannot fit a hge
__host__ __device__ static inline void x87_mul_round(double v, uhge mul, uhge &mag, bool &neg, bool &big)
{
	neg = v < 0;
	big = false;
	mag = 0;
	if (v == 0 || mul == 0)
		return;
	/* v = m * 2^e with an integer m < 2^53 */
	unsigned long long bits;
	memcpy(&bits, &v, 8);
	int ex = (int) ((bits >> 52) & 0x7ff);
	unsigned long long m = bits & ((1ULL << 52) - 1);
	if (ex == 0)
		ex = 1;			/* subnormal */
	else
		m |= 1ULL << 52;
	int e = ex - 1075;
	/* mul as a 64-bit mantissa M2 * 2^e2, round to nearest even */
	int l2 = 0;
	for (uhge t = mul; t; t >>= 1)
		l2++;
	unsigned long long M2;
	int e2 = 0;
	if (l2 <= 64) {
		M2 = (unsigned long long) mul;
	} else {
		const int sh = l2 - 64;
		const uhge rem = mul & (((uhge) 1 << sh) - 1), half = (uhge) 1 << (sh - 1);
		uhge q = mul >> sh;
		if (rem > half || (rem == half && (q & 1)))
			q++;
		if (q >> 64) {
			q >>= 1;
			e2 = 1;
		}
		M2 = (unsigned long long) q;
		e2 += sh;
	}
	/* product, rounded to a 64-bit mantissa M3 * 2^e3 */
	uhge P = (uhge) m * (uhge) M2;
	int e3 = e + e2, lp = 0;
	for (uhge t = P; t; t >>= 1)
		lp++;
	if (lp > 64) {
		const int sh = lp - 64;
		const uhge rem = P & (((uhge) 1 << sh) - 1), half = (uhge) 1 << (sh - 1);
		uhge q = P >> sh;
		if (rem > half || (rem == half && (q & 1)))
			q++;
		if (q >> 64) {
			q >>= 1;
			e3++;
		}
		P = q;
		e3 += sh;
	}
	/* roundl */
	if (e3 >= 0) {
		if (e3 > 63 || (P >> (127 - e3)) != 0) {
			big = true;
			return;
		}
		mag = P << e3;
	} else {
		const int k = -e3;
		if (k > 64)
			return;		/* |value| < 1/2 */
		if (k == 64) {
			mag = (P >> 63) ? 1 : 0;	/* value in [1/2, 1) exactly when the top mantissa bit is set */
			return;
		}
		const uhge ip = P >> k, frac = P & (((uhge) 1 << k) - 1), half = (uhge) 1 << (k - 1);
		mag = ip + (frac >= half ? 1 : 0);
	}
}

template <typename T1, typename T2>
__host__ __device__ static inline int convert_one(T1 v, T2 &dst, const ConvScale &sc)
{
	if constexpr (Atom<T1>::is_float && !Atom<T2>::is_float) {
		/* convertimpl_reduce_float: range check against the UNSCALED type range, scale, round half away
		 * from zero, nil check, precision check */
		if (v < (T1) Atom<T2>::minval() || v > (T1) Atom<T2>::maxval())
			return CV_OVERFLOW;
		T2 r;
		if (sc.mul == 1) {
			double rr = round((double) v);
			const double lim = ldexp(1.0, (int) sizeof(T2) * 8 - 1);
			if (rr >= lim || rr <= -lim)
				return CV_OVERFLOW;	/* the C cast would have produced the nil pattern */
			r = (T2) rr;
		} else {
			uhge mag;
			bool neg, big;
			x87_mul_round((double) v, (uhge) (T2) sc.mul, mag, neg, big);
			if (big || mag > (uhge) Atom<T2>::maxval())
				return CV_OVERFLOW;
			r = neg ? (T2) -(T2) mag : (T2) mag;
		}
		if (sc.prec != 0) {
			const T2 prec = (T2) sc.prec;
			if (r >= prec || r <= (T2) -prec)
				return CV_OVERFLOW;
		}
		dst = r;
		return CV_OK;
	} else if constexpr (!Atom<T1>::is_float && Atom<T2>::is_float) {
		/* convertimpl_enlarge_float: (TYPE2) v / div, div = (TYPE1) scales[scale1] */
		dst = sc.div == 1 ? (T2) v : (T2) v / (T2) (T1) sc.div;
		return CV_OK;
	} else if constexpr (Atom<T1>::is_float) {
		/* flt -> dbl, flt -> flt, dbl -> dbl: enlarge_float with div = (TYPE1) scales[scale1];
		 * dbl -> flt: reduce_float without rounding, mul = (flt) scales[scale2] */
		if constexpr (sizeof(T1) > sizeof(T2)) {
			if (v < (T1) Atom<T2>::minval() || v > (T1) Atom<T2>::maxval())
				return CV_OVERFLOW;
			dst = sc.mul == 1 ? (T2) v : (T2) ((double) v * (double) (T2) sc.mul);
			if (dst != dst)
				return CV_OVERFLOW;
			return CV_OK;
		} else {
			dst = sc.div == 1 ? (T2) v : (T2) v / (T2) (T1) sc.div;
			return CV_OK;
		}
	} else {
		/* convertimpl: round-half-away division by 10^(scale1-scale2), range / precision check on the
		 * quotient against the target range divided by mul, then multiplication by 10^(scale2-scale1) */
		typedef typename std::conditional<sizeof(T1) == 16, uhge, unsigned long long>::type UW;
		T1 q = v;
		const T1 div = (T1) sc.div;
		if (div != 1) {
			const UW a = v < 0 ? (UW) 0 - (UW) v : (UW) v;
			const UW r = (a + ((UW) div >> 1)) / (UW) div;
			q = v < 0 ? (T1) -(T1) r : (T1) r;
		}
		const T2 mul = (T2) sc.mul;
		const T2 mn = (T2) (Atom<T2>::minval() / mul), mx = (T2) (Atom<T2>::maxval() / mul);
		typedef typename std::conditional<(sizeof(T1) > sizeof(T2)), T1, T2>::type W;
		if ((W) q < (W) mn || (W) q > (W) mx)
			return CV_OVERFLOW;
		if (sc.prec != 0) {
			const T2 prec = (T2) ((T2) sc.prec / mul);
			if ((W) q >= (W) prec || (W) q <= (W) (T2) -prec)
				return CV_OVERFLOW;
		}
		dst = (T2) ((T2) q * mul);
		return CV_OK;
	}
}

template <typename T1, typename T2>
__global__ void __launch_bounds__(256)
convert_kernel(const T1 *__restrict__ src, const oid *__restrict__ cand, oid hseq, uint64_t n, T2 *__restrict__ dst,
	       unsigned long long *__restrict__ state, ConvScale sc)
{
	unsigned long long fail = ~0ULL;
	unsigned nils = 0;
	for (uint64_t k = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (uint64_t) gridDim.x * blockDim.x) {
		T1 v = cand ? src[__ldcs(cand + k) - hseq] : ld_stream(src + k);
		T2 z = Atom<T2>::nil();
		if (Atom<T1>::is_nil(v))
			nils++;
		else if (convert_one<T1, T2>(v, z, sc) != CV_OK && fail == ~0ULL)
			fail = k;
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

template <typename T> static void fmt_conv(char *buf, size_t n, T v)
{
	if constexpr (std::is_same<T, hge>::value)
		snprintf(buf, n, "%.40Lg (approx. value)", (long double) v);
	else if constexpr (std::is_same<T, lng>::value)
		snprintf(buf, n, "%lld", (long long) v);
	else if constexpr (std::is_same<T, flt>::value)
		snprintf(buf, n, "%.9g", (double) v);
	else if constexpr (std::is_same<T, dbl>::value)
		snprintf(buf, n, "%.17g", v);
	else
		snprintf(buf, n, "%d", (int) v);
}

// 1 ok, 0 failure (error set)
template <typename T1, typename T2>
static int convert_run(const b200_bat *b, const Cand &ci, void *dst, int tp, uint64_t &nils, const ConvScale &sc, int scale2, int precision)
{
	Ctx *c = ctx();
	unsigned long long *state = (unsigned long long *) scratch_get(64);
	static const unsigned long long init[2] = {~0ULL, 0};
	if (state == nullptr || !cuda_ok(cudaMemcpyAsync(state, init, 16, cudaMemcpyHostToDevice, c->stream), "init"))
		return 0;
	const T1 *src = (const T1 *) b->base;
	const oid *cand = nullptr;
	if (ci.dense)
		src += ci.seq - ci.hseq;
	else
		cand = ci.oids;
	convert_kernel<T1, T2><<<grid_for(ci.ncand, 256, NSM * 32), 256, 0, c->stream>>>(src, cand, ci.hseq, ci.ncand, (T2 *) dst, state, sc);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "convert launch") ||
	    !cuda_ok(cudaMemcpyAsync(c->slot_h->v, state, 16, cudaMemcpyDeviceToHost, c->stream), "read") || !stream_sync())
		return 0;
	nils = c->slot_h->v[1];
	uint64_t fail = c->slot_h->v[0];
	if (fail != ~0ULL) {
		uint64_t i = fail;
		if (cand) {
			oid o;
			if (!cuda_ok(cudaMemcpy(&o, cand + fail, 8, cudaMemcpyDeviceToHost), "fetch"))
				return 0;
			i = o - ci.hseq;
		}
		T1 v;
		if (!cuda_ok(cudaMemcpy(&v, src + i, sizeof(T1), cudaMemcpyDeviceToHost), "fetch"))
			return 0;
		char buf[64];
		fmt_conv<T1>(buf, sizeof(buf), v);
		const bool decimal_text = precision > 0 && !(Atom<T2>::is_float);	/* CONV_OVERFLOW_PREC, :156-166 */
		if (decimal_text)
			set_err("22003!overflow in conversion to DECIMAL(%d,%d).\n", precision, scale2);
		else
			set_err("22003!overflow in conversion of %s to %s.\n", buf, atomname(basetype(tp)));
		return 0;
	}
	return 1;
}

template <typename T1>
static int convert_dispatch2(int t2, const b200_bat *b, const Cand &ci, void *dst, int tp, uint64_t &nils, const ConvScale &sc,
			     int scale2, int precision)
{
	switch (t2) {
	case B200_bte: return convert_run<T1, bte>(b, ci, dst, tp, nils, sc, scale2, precision);
	case B200_sht: return convert_run<T1, sht>(b, ci, dst, tp, nils, sc, scale2, precision);
	case B200_int: return convert_run<T1, int>(b, ci, dst, tp, nils, sc, scale2, precision);
	case B200_lng: return convert_run<T1, lng>(b, ci, dst, tp, nils, sc, scale2, precision);
	case B200_hge: return convert_run<T1, hge>(b, ci, dst, tp, nils, sc, scale2, precision);
	case B200_flt: return convert_run<T1, flt>(b, ci, dst, tp, nils, sc, scale2, precision);
	case B200_dbl: return convert_run<T1, dbl>(b, ci, dst, tp, nils, sc, scale2, precision);
	}
	return -1;
}

static hge pow10_hge(int k)
{
	hge r = 1;
	for (int i = 0; i < k; i++)
		r *= 10;
	return r;
}

extern "C" b200_bat *b200_BATconvert(const b200_bat *b, const b200_bat *s, int tp, uint8_t scale1, uint8_t scale2, uint8_t precision)
{
	b200_clrerr();
	if (b == nullptr) {
		set_err("BATconvert: BAT required\n");
		return nullptr;
	}
	if (tp == B200_void)
		tp = B200_oid;
	Staged sb, ss;
	if (!stage(b, sb) || !stage(s, ss))
		return nullptr;
	b = sb.b;
	Cand ci;
	if (!cand_init(ci, b, ss.b))
		return nullptr;
	if (ci.ncand == 0 || (b->ttype == B200_void && b->tseqbase == OID_NIL))
		return bat_constant(ci.rhseq, tp, nullptr, ci.ncand);
	if (tp == B200_bit || tp == B200_oid || tp == B200_msk ||
	    b->ttype == B200_void || b->ttype == B200_msk || b->ttype == B200_oid) {
		set_err("BATconvert: bit/oid/msk conversions are not on the device path (convert(%s)->%s)\n", atomname(b->ttype),
			atomname(tp));
		return nullptr;
	}
	const int t1 = basetype(b->ttype), t2 = basetype(tp);
	const bool f1 = t1 == B200_flt || t1 == B200_dbl, f2 = t2 == B200_flt || t2 == B200_dbl;
	if ((!f1 && scale1 > 38) || (!f2 && scale2 > 38) || precision > 38) {
		set_err("BATconvert: scale / precision beyond 38 digits\n");	/* the reference asserts (:986-987) */
		return nullptr;
	}
	/* which of scale1 / scale2 / precision a conversion family looks at (see convert_one) */
	ConvScale sc{1, 1, 0};
	if (!f1 && !f2) {
		sc.div = pow10_hge(scale1 > scale2 ? scale1 - scale2 : 0);
		sc.mul = pow10_hge(scale2 > scale1 ? scale2 - scale1 : 0);
		sc.prec = precision ? pow10_hge(precision) : 0;
	} else if (!f1 && f2) {
		sc.div = pow10_hge(scale1);
	} else if (f1 && !f2) {
		sc.mul = pow10_hge(scale2);
		sc.prec = precision ? pow10_hge(precision) : 0;
	}	/* floating point -> floating point: the reference passes scale 0 (gdk_calc_convert.c:1346-1405) */
	if (atomsize(t1) <= 0 || atomsize(t2) <= 0) {
		set_err("type combination (convert(%s)->%s) not supported.\n", atomname(b->ttype), atomname(tp));
		return nullptr;
	}
	{
		/* the powers of ten must fit the types they are cast to (the reference divides by a wrapped 0 otherwise) */
		auto fits = [](hge p, int t) {
			switch (t) {
			case B200_bte: return p <= 100;
			case B200_sht: return p <= 10000;
			case B200_int: return p <= 1000000000;
			case B200_lng: return p <= (hge) 1000000000000000000LL;
			default: return true;
			}
		};
		if ((!f2 && (!fits(sc.mul, t2) || !fits(sc.prec, t2))) || (!f1 && !fits(sc.div, t1))) {
			set_err("BATconvert: scale / precision does not fit the types (convert(%s)->%s, %d, %d, %d)\n", atomname(b->ttype),
				atomname(tp), scale1, scale2, precision);
			return nullptr;
		}
	}
	b200_bat *bn = bat_new(ci.rhseq, tp, ci.ncand);
	if (bn == nullptr)
		return nullptr;
	uint64_t nils = 0;
	int rc = -1;
	ctx()->algo = "convert: cuda streaming";
	switch (t1) {
	case B200_bte: rc = convert_dispatch2<bte>(t2, b, ci, bn->base, tp, nils, sc, scale2, precision); break;
	case B200_sht: rc = convert_dispatch2<sht>(t2, b, ci, bn->base, tp, nils, sc, scale2, precision); break;
	case B200_int: rc = convert_dispatch2<int>(t2, b, ci, bn->base, tp, nils, sc, scale2, precision); break;
	case B200_lng: rc = convert_dispatch2<lng>(t2, b, ci, bn->base, tp, nils, sc, scale2, precision); break;
	case B200_hge: rc = convert_dispatch2<hge>(t2, b, ci, bn->base, tp, nils, sc, scale2, precision); break;
	case B200_flt: rc = convert_dispatch2<flt>(t2, b, ci, bn->base, tp, nils, sc, scale2, precision); break;
	case B200_dbl: rc = convert_dispatch2<dbl>(t2, b, ci, bn->base, tp, nils, sc, scale2, precision); break;
	}
	if (rc <= 0) {
		b200_BATfree(bn);
		if (rc < 0)
			set_err("type combination (convert(%s)->%s) not supported.\n", atomname(b->ttype), atomname(tp));
		return nullptr;
	}
	bn->count = ci.ncand;
	bn->tnil = nils != 0;
	bn->tnonil = nils == 0;
	bn->tsorted = (nils == 0 && b->tsorted) || ci.ncand <= 1;
	bn->trevsorted = (nils == 0 && b->trevsorted) || ci.ncand <= 1;
	/* "reduce": several source values may map to one destination value */
	const bool reduce = (f1 && t2 != B200_dbl && t2 != t1) || (!f1 && !f2 && sc.div > 1) ||
			    ((t2 == B200_flt && atomsize(t1) * 8 > 24) || (t2 == B200_dbl && atomsize(t1) * 8 > 53 && !f1));
	bn->tkey = ci.ncand <= 1 || (!reduce && b->tkey && nils <= 1);
	return bn;
}
