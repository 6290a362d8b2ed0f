// calc.cuh — element-wise BATcalc arithmetic on the device.
//
// dst[k] = lft[i_k] OP rgt[j_k] over two candidate iterators, with the
// reference's nil / overflow / division-by-zero rules:
//   add/sub  gdk_calc_addsub.c:22-127 (ADD_3TYPE / ADD_3TYPE_enlarge), checks
//            gdk_calc_private.h:40-122
//   mul      gdk_calc_mul.c:22-345 (MUL_4TYPE, MUL_3TYPE_enlarge, MUL_2TYPE_lng/hge,
//            MUL_2TYPE_float), checks gdk_calc_private.h:124-229
//   div      gdk_calc_div.c:20-140 (DIV_3TYPE, DIV_3TYPE_float)
// A nil operand gives nil (counted); the first row (in candidate order) whose
// result does not fit aborts the whole operation with the reference's message
// (ON_OVERFLOW, gdk_calc_private.h:346-352): the kernels record the smallest
// failing row with atomicMin and the host formats the text from the operands
// of that row.  One streaming pass, 128-bit loads/stores when operands are
// dense and aligned; candidate lists are consumed as gathers.
#pragma once
#include "common.cuh"
#include <type_traits>

enum { OP_ADD = 0, OP_SUB = 1, OP_MUL = 2, OP_DIV = 3, OP_MOD = 4 };
enum { ST_OK = 0, ST_OVERFLOW = 1, ST_DIVZERO = 2, ST_DIVRANGE = 3 };

template <typename T> struct Rank;
template <> struct Rank<bte> { static constexpr int v = 0; typedef sht wider; };
template <> struct Rank<sht> { static constexpr int v = 1; typedef int wider; };
template <> struct Rank<int> { static constexpr int v = 2; typedef lng wider; };
template <> struct Rank<lng> { static constexpr int v = 3; typedef hge wider; };
template <> struct Rank<hge> { static constexpr int v = 4; typedef hge wider; };
template <> struct Rank<flt> { static constexpr int v = 5; typedef dbl wider; };
template <> struct Rank<dbl> { static constexpr int v = 6; typedef dbl wider; };

template <typename T> constexpr bool is_fl = Atom<T>::is_float;

// which (T1,T2)->T3 combinations run on the device (the reference's matrices minus
// the int x float -> int decimal-rescale variants that use x87 long double)
template <int OP, typename T1, typename T2, typename T3> constexpr bool calc_supported()
{
	constexpr int r1 = Rank<T1>::v, r2 = Rank<T2>::v, r3 = Rank<T3>::v;
	constexpr int rm = r1 > r2 ? r1 : r2;
	if (OP == OP_MOD) {
		/* gdk_calc_mod.c:121-290: an integer result holds the narrower input; a
		 * floating-point one exists only in the widest input type */
		if (!is_fl<T1> && !is_fl<T2>)
			return !is_fl<T3> && r3 >= (r1 < r2 ? r1 : r2);
		return r3 == rm;
	}
	if (!is_fl<T3>) {
		if (is_fl<T1> || is_fl<T2>)
			return false;
		return OP == OP_DIV ? r3 >= r1 : r3 >= rm;
	}
	/* float result: flt only when no dbl input */
	if (OP == OP_DIV && is_fl<T1> && !is_fl<T2>)
		return r3 >= r1;	/* DIV_3TYPE(flt,int,flt|dbl), (dbl,int,dbl) */
	return r3 >= rm;
}

// 128-bit helpers -------------------------------------------------------------
__host__ __device__ static inline bool hge_add_ovf(hge a, hge b, hge &r)
{
	uhge ur = (uhge) a + (uhge) b;
	r = (hge) ur;
	/* overflow iff operands share a sign that the result does not; the type minimum is nil */
	return (((a ^ r) & (b ^ r)) < 0) || r == Atom<hge>::nil();
}
__host__ __device__ static inline bool hge_sub_ovf(hge a, hge b, hge &r)
{
	uhge ur = (uhge) a - (uhge) b;
	r = (hge) ur;
	return (((a ^ b) & (a ^ r)) < 0) || r == Atom<hge>::nil();
}
__host__ __device__ static inline bool hge_mul_ovf(hge a, hge b, hge &r)
{
	/* HGEMUL_CHECK's generic form, gdk_calc_private.h:196-226 */
	bool neg = false;
	if (a < 0) { neg = !neg; a = -a; }
	if (b < 0) { neg = !neg; b = -b; }
	uint64_t a1 = (uint64_t) ((uhge) a >> 64), a2 = (uint64_t) a;
	uint64_t b1 = (uint64_t) ((uhge) b >> 64), b2 = (uint64_t) b;
	if (a1 != 0 && b1 != 0)
		return true;
	uhge c = (uhge) a1 * b2 + (uhge) a2 * b1;
	if (c >> 63)
		return true;
	uhge lo = (uhge) a2 * b2;
	c = (c << 64) + lo;
	if (c < lo || (c >> 127))
		return true;
	r = neg ? -(hge) c : (hge) c;
	return false;
}

template <typename T> __host__ __device__ static inline T absval(T v) { return v < 0 ? -v : v; }

// ---------------------------------------------------------------- the per-row rule
template <int OP, typename T1, typename T2, typename T3> struct Calc {
	static constexpr bool plain = std::is_same<T3, T1>::value || std::is_same<T3, T2>::value;

	__host__ __device__ static inline int apply(T1 a, T2 b, T3 &dst)
	{
		const T3 max = Atom<T3>::maxval();
		if constexpr (OP == OP_MOD) {
			if (b == 0)
				return ST_DIVZERO;
			if constexpr (is_fl<T3>) {
				/* FMOD_3TYPE: fmodf / fmod on both operands converted to the result type */
				if constexpr (std::is_same<T3, flt>::value)
					dst = fmodf((flt) a, (flt) b);
				else
					dst = fmod((dbl) a, (dbl) b);
			} else {
				/* MOD_3TYPE: `(TYPE3) lft % rgt` — the cast binds to lft alone, so a
				 * result type narrower than lft truncates it first, as the reference does */
				const T3 l = (T3) a;
				if (b == (T2) -1)
					dst = 0;	/* the remainder is 0; avoids min % -1 after a truncating cast */
				else
					dst = (T3) (l % b);
			}
			return ST_OK;
		} else if constexpr (!is_fl<T3>) {
			/* integer result */
			if constexpr (OP == OP_DIV) {
				if (b == 0)
					return ST_DIVZERO;
				dst = (T3) (a / b);
				if (dst < -max || dst > max)
					return ST_DIVRANGE;
				return ST_OK;
			} else if constexpr (std::is_same<T3, hge>::value) {
				if constexpr (OP == OP_MUL) {
					if constexpr (std::is_same<T1, hge>::value || std::is_same<T2, hge>::value) {
						return hge_mul_ovf((hge) a, (hge) b, dst) ? ST_OVERFLOW : ST_OK;
					} else {
						dst = (hge) a * (hge) b;	/* <= 64 x 64 bits: cannot overflow */
						return ST_OK;
					}
				} else if constexpr (std::is_same<T1, hge>::value || std::is_same<T2, hge>::value) {
					bool o = OP == OP_ADD ? hge_add_ovf((hge) a, (hge) b, dst) : hge_sub_ovf((hge) a, (hge) b, dst);
					return o ? ST_OVERFLOW : ST_OK;
				} else {
					dst = OP == OP_ADD ? (hge) a + (hge) b : (hge) a - (hge) b;
					return ST_OK;
				}
			} else if constexpr (Rank<T3>::v > Rank<T1>::v && Rank<T3>::v > Rank<T2>::v) {
				/* result type wider than both inputs: cannot overflow (couldoverflow == false) */
				dst = OP == OP_ADD ? (T3) a + (T3) b : OP == OP_SUB ? (T3) a - (T3) b : (T3) a * (T3) b;
				return ST_OK;
			} else {
				/* compute exactly in the next wider type, then range-check against [-max,max] */
				typedef typename Rank<T3>::wider W;
				if constexpr (sizeof(W) == 16 && sizeof(T3) == 8 && OP != OP_MUL) {
					/* lng +- lng -> lng: the 128-bit detour made this kernel issue-bound (78 % SM,
					 * 60 % DRAM, profiles/r2_q1_75m_full.md); the wrapped 64-bit result and the
					 * sign rule give the same verdict: overflow iff the operands' signs allow it and
					 * the result's sign differs, or the result is the nil pattern (-max - 1) */
					const unsigned long long ua = (unsigned long long) (lng) a, ub = (unsigned long long) (lng) b;
					const unsigned long long ur = OP == OP_ADD ? ua + ub : ua - ub;
					const unsigned long long ov = OP == OP_ADD ? ((ua ^ ur) & (ub ^ ur)) : ((ua ^ ub) & (ua ^ ur));
					if ((long long) ov < 0 || ur == 0x8000000000000000ULL)
						return ST_OVERFLOW;
					dst = (T3) (long long) ur;
					return ST_OK;
				}
				W r = OP == OP_ADD ? (W) a + (W) b : OP == OP_SUB ? (W) a - (W) b : (W) a * (W) b;
				if (r < -(W) max || r > (W) max)
					return ST_OVERFLOW;
				dst = (T3) r;
				return ST_OK;
			}
		} else {
			/* floating-point result: the reference's expressions with C's usual conversions */
			const T3 max1 = (T3) Atom<T1>::maxval(), max2 = (T3) Atom<T2>::maxval();
			if constexpr (OP == OP_ADD) {
				const bool check = plain || max < max1 + max2;
				if (check) {
					if (b < 1) {
						if (-max - b > a)
							return ST_OVERFLOW;
					} else if (max - b < a)
						return ST_OVERFLOW;
				}
				dst = (T3) a + b;
				return ST_OK;
			} else if constexpr (OP == OP_SUB) {
				const bool check = plain || max < max1 + max2;
				if (check) {
					if (b < 1) {
						if (max + b < a)
							return ST_OVERFLOW;
					} else if (-max + b > a)
						return ST_OVERFLOW;
				}
				dst = (T3) a - b;
				return ST_OK;
			} else if constexpr (OP == OP_MUL) {
				if constexpr (plain) {
					dst = (T3) (a * b);
					if (isinf(dst) || absval(dst) > max)
						return ST_OVERFLOW;
					return ST_OK;
				} else {
					T3 c = (T3) a * b;
					if (max < max1 * max2 && (c < -max || c > max))
						return ST_OVERFLOW;
					dst = c;
					return ST_OK;
				}
			} else {
				if constexpr (is_fl<T2>) {
					/* DIV_3TYPE_float */
					if (b == 0)
						return ST_DIVZERO;
					if (absval(b) < 1 && max * absval(b) < absval(a))
						return ST_OVERFLOW;
					dst = (T3) a / b;
				} else {
					if (b == 0)
						return ST_DIVZERO;
					dst = (T3) (a / b);
				}
				if (dst < -max || dst > max)
					return ST_DIVRANGE;
				return ST_OK;
			}
		}
	}
};

// ---------------------------------------------------------------- kernels
template <typename T1, typename T2, typename T3> struct CalcArgs {
	const T1 *lft;		/* dense: already offset to the first candidate row; cand: column base */
	const T2 *rgt;
	const oid *c1, *c2;	/* candidate oids (NULL = dense / constant) */
	oid off1, off2;		/* hseqbase to subtract from candidate oids */
	T1 cst1;
	T2 cst2;
	bool incr1, incr2;	/* false: constant operand */
	T3 *dst;
	uint64_t n;
	unsigned long long *state;	/* [0] first failing row (init ~0), [1] nil count */
};

template <typename T, int K> struct VecIO {
	static constexpr int BYTES = K * sizeof(T);
	static constexpr int CHUNK = BYTES >= 16 ? 16 : BYTES;
	typedef typename std::conditional<CHUNK == 16, uint4,
		typename std::conditional<CHUNK == 8, uint2,
		typename std::conditional<CHUNK == 4, uint32_t,
		typename std::conditional<CHUNK == 2, uint16_t, uint8_t>::type>::type>::type>::type chunk_t;
	static constexpr int NCH = BYTES / CHUNK;
	__device__ static inline void load(const T *p, T (&v)[K])
	{
		union { chunk_t c[NCH]; T t[K]; } u;
#pragma unroll
		for (int i = 0; i < NCH; i++)
			u.c[i] = __ldcs(reinterpret_cast<const chunk_t *>(p) + i);
#pragma unroll
		for (int i = 0; i < K; i++)
			v[i] = u.t[i];
	}
	__device__ static inline void store(T *p, const T (&v)[K])
	{
		union { chunk_t c[NCH]; T t[K]; } u;
#pragma unroll
		for (int i = 0; i < K; i++)
			u.t[i] = v[i];
#pragma unroll
		for (int i = 0; i < NCH; i++)
			__stcs(reinterpret_cast<chunk_t *>(p) + i, u.c[i]);
	}
};
template <typename T> constexpr int calc_k() { return sizeof(T) >= 8 ? 2 : (sizeof(T) == 4 ? 4 : 8); }

__device__ __forceinline__ void calc_finish(unsigned long long fail, unsigned nils, unsigned long long *state)
{
	if (fail != ~0ULL)
		atomicMin(&state[0], fail);
#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
		nils += __shfl_down_sync(0xffffffffu, nils, o);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(&state[1], (unsigned long long) nils);
}

// dense / constant operands, aligned: K consecutive rows per thread
template <int OP, typename T1, typename T2, typename T3>
__global__ void __launch_bounds__(256, 4) calc_vec_kernel(CalcArgs<T1, T2, T3> a)
{
	constexpr int K = calc_k<T3>();
	/* results of up to 8 bytes: four independent 16-byte chunks per operand in flight per thread
	 * (one chunk leaves too few bytes in flight to saturate HBM); hge results already stream 32 bytes */
	constexpr int UNR = sizeof(T3) <= 8 && sizeof(T1) <= 8 && sizeof(T2) <= 8 ? 4 : 1;
	const uint64_t nchunks = a.n / K;
	const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
	unsigned long long fail = ~0ULL;
	unsigned nils = 0;
	for (uint64_t c0 = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; c0 < nchunks; c0 += stride * UNR) {
		T1 x[UNR][K];
		T2 y[UNR][K];
#pragma unroll
		for (int u = 0; u < UNR; u++) {
			const uint64_t c = c0 + u * stride;
			if (c < nchunks) {
				if (a.incr1)
					VecIO<T1, K>::load(a.lft + c * K, x[u]);
				if (a.incr2)
					VecIO<T2, K>::load(a.rgt + c * K, y[u]);
			}
		}
#pragma unroll
		for (int u = 0; u < UNR; u++) {
			const uint64_t c = c0 + u * stride;
			if (c >= nchunks)
				break;
			const uint64_t k0 = c * K;
			T3 z[K];
#pragma unroll
			for (int e = 0; e < K; e++) {
				const T1 xv = a.incr1 ? x[u][e] : a.cst1;
				const T2 yv = a.incr2 ? y[u][e] : a.cst2;
				z[e] = Atom<T3>::nil();
				if (Atom<T1>::is_nil(xv) || Atom<T2>::is_nil(yv)) {
					nils++;
				} else if (Calc<OP, T1, T2, T3>::apply(xv, yv, z[e]) != ST_OK && fail == ~0ULL) {
					fail = k0 + e;
				}
			}
			VecIO<T3, K>::store(a.dst + k0, z);
		}
	}
	/* tail rows */
	const uint64_t k = nchunks * K + (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (k < a.n) {
		T1 x = a.incr1 ? a.lft[k] : a.cst1;
		T2 y = a.incr2 ? a.rgt[k] : a.cst2;
		T3 z = Atom<T3>::nil();
		if (Atom<T1>::is_nil(x) || Atom<T2>::is_nil(y))
			nils++;
		else if (Calc<OP, T1, T2, T3>::apply(x, y, z) != ST_OK && fail == ~0ULL)
			fail = k;
		a.dst[k] = z;
	}
	calc_finish(fail, nils, a.state);
}

// general form: one row per thread step, operands through candidate lists or unaligned
template <int OP, typename T1, typename T2, typename T3>
__global__ void __launch_bounds__(256) calc_gen_kernel(CalcArgs<T1, T2, T3> a)
{
	unsigned long long fail = ~0ULL;
	unsigned nils = 0;
	for (uint64_t k = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; k < a.n; k += (uint64_t) gridDim.x * blockDim.x) {
		T1 x = !a.incr1 ? a.cst1 : a.c1 ? a.lft[__ldcs(a.c1 + k) - a.off1] : a.lft[k];
		T2 y = !a.incr2 ? a.cst2 : a.c2 ? a.rgt[__ldcs(a.c2 + k) - a.off2] : a.rgt[k];
		T3 z = Atom<T3>::nil();
		if (Atom<T1>::is_nil(x) || Atom<T2>::is_nil(y))
			nils++;
		else if (Calc<OP, T1, T2, T3>::apply(x, y, z) != ST_OK && fail == ~0ULL)
			fail = k;
		a.dst[k] = z;
	}
	calc_finish(fail, nils, a.state);
}

// ---------------------------------------------------------------- host driver
struct CalcOperand {
	const b200_bat *b;	/* NULL: constant */
	const Cand *ci;
	const void *cst;	/* constant value (already of type `type`) */
	int type;		/* storage type */
};
struct CalcOut {
	uint64_t nils = 0;
	uint64_t fail = ~0ULL;	/* first failing row */
	int status = ST_OK;	/* status of that row */
	char l[64], r[64];	/* operands of that row, formatted as the reference does */
};

template <typename T> static void fmt_val(char *buf, size_t n, T v)
{
	/* FMTbte..FMTdbl, gdk_calc_private.h:300-308 */
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

template <int OP, typename T1, typename T2, typename T3>
static bool calc_run(const CalcOperand &o1, const CalcOperand &o2, void *dst, uint64_t n, CalcOut &out)
{
	Ctx *c = ctx();
	unsigned long long *state = (unsigned long long *) scratch_get(64);
	if (state == nullptr)
		return false;
	static const unsigned long long init[2] = {~0ULL, 0};
	if (!cuda_ok(cudaMemcpyAsync(state, init, 16, cudaMemcpyHostToDevice, c->stream), "calc state init"))
		return false;
	CalcArgs<T1, T2, T3> a;
	memset(&a, 0, sizeof(a));
	a.incr1 = o1.b != nullptr;
	a.incr2 = o2.b != nullptr;
	a.dst = (T3 *) dst;
	a.n = n;
	a.state = state;
	bool dense = true;
	if (a.incr1) {
		a.lft = (const T1 *) o1.b->base;
		if (o1.ci->dense)
			a.lft += o1.ci->seq - o1.ci->hseq;
		else {
			a.c1 = o1.ci->oids;
			a.off1 = o1.ci->hseq;
			dense = false;
		}
	} else
		a.cst1 = *(const T1 *) o1.cst;
	if (a.incr2) {
		a.rgt = (const T2 *) o2.b->base;
		if (o2.ci->dense)
			a.rgt += o2.ci->seq - o2.ci->hseq;
		else {
			a.c2 = o2.ci->oids;
			a.off2 = o2.ci->hseq;
			dense = false;
		}
	} else
		a.cst2 = *(const T2 *) o2.cst;
	constexpr int K = calc_k<T3>();
	bool aligned = dense &&
		(!a.incr1 || ((uintptr_t) a.lft % VecIO<T1, K>::CHUNK) == 0) &&
		(!a.incr2 || ((uintptr_t) a.rgt % VecIO<T2, K>::CHUNK) == 0) &&
		((uintptr_t) a.dst % VecIO<T3, K>::CHUNK) == 0;
	KTimer kt("calc", n * ((a.incr1 ? sizeof(T1) : 0) + (a.incr2 ? sizeof(T2) : 0) + sizeof(T3)));
	if (aligned) {
		unsigned grid = grid_for(n / K + 256, 256, NSM * 32);
		calc_vec_kernel<OP, T1, T2, T3><<<grid, 256, 0, c->stream>>>(a);
	} else {
		unsigned grid = grid_for(n, 256, NSM * 32);
		calc_gen_kernel<OP, T1, T2, T3><<<grid, 256, 0, c->stream>>>(a);
	}
	COUNT_LAUNCH();
	kt.stop();
	if (!cuda_ok(cudaGetLastError(), "calc launch") ||
	    !cuda_ok(cudaMemcpyAsync(c->slot_h->v, state, 16, cudaMemcpyDeviceToHost, c->stream), "calc state read") ||
	    !stream_sync())
		return false;
	out.fail = c->slot_h->v[0];
	out.nils = c->slot_h->v[1];
	if (out.fail != ~0ULL) {
		/* fetch the operands of the first failing row and classify it on the host */
		T1 x;
		T2 y;
		if (a.incr1) {
			uint64_t i = out.fail;
			if (a.c1) {
				oid o;
				if (!cuda_ok(cudaMemcpy(&o, a.c1 + out.fail, 8, cudaMemcpyDeviceToHost), "fetch"))
					return false;
				i = o - a.off1;
			}
			if (!cuda_ok(cudaMemcpy(&x, a.lft + i, sizeof(T1), cudaMemcpyDeviceToHost), "fetch"))
				return false;
		} else
			x = a.cst1;
		if (a.incr2) {
			uint64_t j = out.fail;
			if (a.c2) {
				oid o;
				if (!cuda_ok(cudaMemcpy(&o, a.c2 + out.fail, 8, cudaMemcpyDeviceToHost), "fetch"))
					return false;
				j = o - a.off2;
			}
			if (!cuda_ok(cudaMemcpy(&y, a.rgt + j, sizeof(T2), cudaMemcpyDeviceToHost), "fetch"))
				return false;
		} else
			y = a.cst2;
		T3 z;
		out.status = Calc<OP, T1, T2, T3>::apply(x, y, z);
		fmt_val<T1>(out.l, sizeof(out.l), x);
		fmt_val<T2>(out.r, sizeof(out.r), y);
	}
	return true;
}

// returns 1 ok, 0 cuda failure, -1 unsupported type combination
template <int OP, typename T1, typename T2>
static int calc_dispatch3(int t3, const CalcOperand &o1, const CalcOperand &o2, void *dst, uint64_t n, CalcOut &out)
{
#define CASE3(ID, T3)                                                                     \
	case ID:                                                                          \
		if constexpr (calc_supported<OP, T1, T2, T3>())                           \
			return calc_run<OP, T1, T2, T3>(o1, o2, dst, n, out) ? 1 : 0;     \
		else                                                                      \
			return -1;
	switch (t3) {
		CASE3(B200_bte, bte)
		CASE3(B200_sht, sht)
		CASE3(B200_int, int)
		CASE3(B200_lng, lng)
		CASE3(B200_hge, hge)
		CASE3(B200_flt, flt)
		CASE3(B200_dbl, dbl)
	}
#undef CASE3
	return -1;
}
template <int OP, typename T1>
static int calc_dispatch2(int t2, int t3, const CalcOperand &o1, const CalcOperand &o2, void *dst, uint64_t n, CalcOut &out)
{
	switch (t2) {
	case B200_bte: return calc_dispatch3<OP, T1, bte>(t3, o1, o2, dst, n, out);
	case B200_sht: return calc_dispatch3<OP, T1, sht>(t3, o1, o2, dst, n, out);
	case B200_int: return calc_dispatch3<OP, T1, int>(t3, o1, o2, dst, n, out);
	case B200_lng: return calc_dispatch3<OP, T1, lng>(t3, o1, o2, dst, n, out);
	case B200_hge: return calc_dispatch3<OP, T1, hge>(t3, o1, o2, dst, n, out);
	case B200_flt: return calc_dispatch3<OP, T1, flt>(t3, o1, o2, dst, n, out);
	case B200_dbl: return calc_dispatch3<OP, T1, dbl>(t3, o1, o2, dst, n, out);
	}
	return -1;
}
template <int OP>
static int calc_dispatch(int t1, int t2, int t3, const CalcOperand &o1, const CalcOperand &o2, void *dst, uint64_t n, CalcOut &out)
{
	switch (t1) {
	case B200_bte: return calc_dispatch2<OP, bte>(t2, t3, o1, o2, dst, n, out);
	case B200_sht: return calc_dispatch2<OP, sht>(t2, t3, o1, o2, dst, n, out);
	case B200_int: return calc_dispatch2<OP, int>(t2, t3, o1, o2, dst, n, out);
	case B200_lng: return calc_dispatch2<OP, lng>(t2, t3, o1, o2, dst, n, out);
	case B200_hge: return calc_dispatch2<OP, hge>(t2, t3, o1, o2, dst, n, out);
	case B200_flt: return calc_dispatch2<OP, flt>(t2, t3, o1, o2, dst, n, out);
	case B200_dbl: return calc_dispatch2<OP, dbl>(t2, t3, o1, o2, dst, n, out);
	}
	return -1;
}

// per-op entry points, one translation unit each (calc_add.cu ... calc_div.cu)
typedef int (*calc_fn)(int, int, int, const CalcOperand &, const CalcOperand &, void *, uint64_t, CalcOut &);
int calc_add_entry(int, int, int, const CalcOperand &, const CalcOperand &, void *, uint64_t, CalcOut &);
int calc_sub_entry(int, int, int, const CalcOperand &, const CalcOperand &, void *, uint64_t, CalcOut &);
int calc_mul_entry(int, int, int, const CalcOperand &, const CalcOperand &, void *, uint64_t, CalcOut &);
int calc_mod_entry(int, int, int, const CalcOperand &, const CalcOperand &, void *, uint64_t, CalcOut &);
int calc_div_entry(int, int, int, const CalcOperand &, const CalcOperand &, void *, uint64_t, CalcOut &);
