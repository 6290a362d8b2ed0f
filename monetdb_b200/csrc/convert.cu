// convert.cu — BATconvert for the fixed-width numeric atoms, scale/precision 0
// (gdk_calc_convert.c:1415-1540; kernels convertimpl :262-363,
// convertimpl_enlarge_float :98-146, convertimpl_reduce_float :365-424).
// Used directly by SQL casts and by the singleton-group shortcut of
// BATgroupsum / BATgroupavg (gdk_aggr.c:929-935,1853-1860).
#include "common.cuh"
#include <type_traits>

enum { CV_OK = 0, CV_OVERFLOW = 1 };

template <typename T1, typename T2>
__host__ __device__ static inline int convert_one(T1 v, T2 &dst)
{
	if constexpr (Atom<T1>::is_float && !Atom<T2>::is_float) {
		/* convertimpl_reduce_float: range check, round half away from zero, nil check */
		if (v < (T1) Atom<T2>::minval() || v > (T1) Atom<T2>::maxval())
			return CV_OVERFLOW;
		double r = round((double) v);
		const double lim = ldexp(1.0, (int) sizeof(T2) * 8 - 1);
		if (r >= lim || r <= -lim)
			return CV_OVERFLOW;	/* the C cast would have produced the nil pattern */
		dst = (T2) r;
		return CV_OK;
	} else if constexpr (!Atom<T1>::is_float && Atom<T2>::is_float) {
		dst = (T2) v;		/* convertimpl_enlarge_float */
		return CV_OK;
	} else if constexpr (Atom<T1>::is_float) {
		/* flt -> dbl widens, dbl -> flt checks the range (convertimpl) */
		if constexpr (sizeof(T1) > sizeof(T2)) {
			if (v < (T1) Atom<T2>::minval() || v > (T1) Atom<T2>::maxval())
				return CV_OVERFLOW;
		}
		dst = (T2) v;
		return CV_OK;
	} else {
		if constexpr (sizeof(T1) > sizeof(T2)) {
			if (v < (T1) Atom<T2>::minval() || v > (T1) Atom<T2>::maxval())
				return CV_OVERFLOW;
		}
		dst = (T2) v;
		return CV_OK;
	}
}

template <typename T1, typename T2>
__global__ void __launch_bounds__(256)
convert_kernel(const T1 *__restrict__ src, const oid *__restrict__ cand, oid hseq, uint64_t n, T2 *__restrict__ dst,
	       unsigned long long *__restrict__ state)
{
	unsigned long long fail = ~0ULL;
	unsigned nils = 0;
	for (uint64_t k = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (uint64_t) gridDim.x * blockDim.x) {
		T1 v = cand ? src[__ldcs(cand + k) - hseq] : ld_stream(src + k);
		T2 z = Atom<T2>::nil();
		if (Atom<T1>::is_nil(v))
			nils++;
		else if (convert_one<T1, T2>(v, z) != CV_OK && fail == ~0ULL)
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
static int convert_run(const b200_bat *b, const Cand &ci, void *dst, int tp, uint64_t &nils)
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
	convert_kernel<T1, T2><<<grid_for(ci.ncand, 256, NSM * 32), 256, 0, c->stream>>>(src, cand, ci.hseq, ci.ncand, (T2 *) dst, state);
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
		set_err("22003!overflow in conversion of %s to %s.\n", buf, atomname(basetype(tp)));
		return 0;
	}
	return 1;
}

template <typename T1>
static int convert_dispatch2(int t2, const b200_bat *b, const Cand &ci, void *dst, int tp, uint64_t &nils)
{
	switch (t2) {
	case B200_bte: return convert_run<T1, bte>(b, ci, dst, tp, nils);
	case B200_sht: return convert_run<T1, sht>(b, ci, dst, tp, nils);
	case B200_int: return convert_run<T1, int>(b, ci, dst, tp, nils);
	case B200_lng: return convert_run<T1, lng>(b, ci, dst, tp, nils);
	case B200_hge: return convert_run<T1, hge>(b, ci, dst, tp, nils);
	case B200_flt: return convert_run<T1, flt>(b, ci, dst, tp, nils);
	case B200_dbl: return convert_run<T1, dbl>(b, ci, dst, tp, nils);
	}
	return -1;
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
	if (scale1 != 0 || scale2 != 0 || precision != 0 || tp == B200_bit || tp == B200_oid || tp == B200_msk ||
	    b->ttype == B200_void || b->ttype == B200_msk || b->ttype == B200_oid) {
		set_err("BATconvert: decimal rescaling and bit/oid/msk conversions are not on the device path "
			"(convert(%s)->%s)\n", atomname(b->ttype), atomname(tp));
		return nullptr;
	}
	const int t1 = basetype(b->ttype), t2 = basetype(tp);
	if (atomsize(t1) <= 0 || atomsize(t2) <= 0) {
		set_err("type combination (convert(%s)->%s) not supported.\n", atomname(b->ttype), atomname(tp));
		return nullptr;
	}
	b200_bat *bn = bat_new(ci.rhseq, tp, ci.ncand);
	if (bn == nullptr)
		return nullptr;
	uint64_t nils = 0;
	int rc = -1;
	ctx()->algo = "convert: cuda streaming";
	switch (t1) {
	case B200_bte: rc = convert_dispatch2<bte>(t2, b, ci, bn->base, tp, nils); break;
	case B200_sht: rc = convert_dispatch2<sht>(t2, b, ci, bn->base, tp, nils); break;
	case B200_int: rc = convert_dispatch2<int>(t2, b, ci, bn->base, tp, nils); break;
	case B200_lng: rc = convert_dispatch2<lng>(t2, b, ci, bn->base, tp, nils); break;
	case B200_hge: rc = convert_dispatch2<hge>(t2, b, ci, bn->base, tp, nils); break;
	case B200_flt: rc = convert_dispatch2<flt>(t2, b, ci, bn->base, tp, nils); break;
	case B200_dbl: rc = convert_dispatch2<dbl>(t2, b, ci, bn->base, tp, nils); break;
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
	const bool reduce = (Atom<flt>::is_float && (t1 == B200_flt || t1 == B200_dbl) && t2 != B200_dbl && t2 != t1) ||
			    ((t2 == B200_flt && atomsize(t1) * 8 > 24) || (t2 == B200_dbl && atomsize(t1) * 8 > 53 && t1 != B200_flt && t1 != B200_dbl));
	bn->tkey = ci.ncand <= 1 || (!reduce && b->tkey && nils <= 1);
	return bn;
}
