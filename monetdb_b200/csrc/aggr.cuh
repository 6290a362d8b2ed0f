// aggr.cuh — grouped / scalar SUM, COUNT and AVG state on the device
// (gdk_aggr.c: AGGR_SUM :429-705, dosum :708-896, dofsum :183-427, AGGR_AVG
// :1717-1753, BATgroupcount :3069-3202).
//
// The reference adds row by row into sums[gid] and fails on the first prefix
// that leaves [-max,max] of the result type.  Here every value is split into
// 32-bit digits that are summed independently in 64-bit cells (exact for fewer
// than 2^32 rows, order independent, no carries between cells); per group we
// keep the digit sums of all values (S), of the positive values only (P, when an
// overflow check is required) and the count of non-nil rows.  A finishing kernel
// rebuilds the exact wide integer, and
//   * if P and S-P both fit the result type no prefix can overflow -> result S;
//   * otherwise an ordered prefix-extreme reduction decides exactly whether the
//     reference's running sum would have overflowed.
// flt/dbl values are first converted to a 128-bit fixed-point integer relative to
// the column's largest exponent, which makes floating-point sums order
// independent and (for values within 2^73 of the largest) exactly rounded like
// the reference's msum partials.
//
// Three accumulation layouts, chosen by the number of groups:
//   REG   ngrp <= 8     per-thread register accumulators, no atomics at all
//   SHM   ngrp <= ~4K   per-block shared-memory table, flushed with global REDs
//   GLB   otherwise     global REDs straight into the cell table
#pragma once
#include "common.cuh"
#include "aggr_init.cuh"
#include <type_traits>

typedef unsigned long long u64;
typedef long long i64;

// ---------------------------------------------------------------- digit split
template <typename T> struct Digits;
#define SMALL_DIGITS(T)                                                                  \
	template <> struct Digits<T> {                                                   \
		static constexpr int ND = 1;                                             \
		__device__ static inline void split(T x, int, i64 (&d)[1]) { d[0] = (i64) x; } \
	};
SMALL_DIGITS(bte)
SMALL_DIGITS(sht)
SMALL_DIGITS(int)
template <> struct Digits<lng> {
	static constexpr int ND = 2;
	__device__ static inline void split(lng x, int, i64 (&d)[2])
	{
		d[0] = (i64) (uint32_t) x;
		d[1] = x >> 32;
	}
};
__device__ static inline void split_hge(hge x, i64 (&d)[4])
{
	u64 lo = (u64) x;
	i64 hi = (i64) (x >> 64);
	d[0] = (i64) (uint32_t) lo;
	d[1] = (i64) (lo >> 32);
	d[2] = (i64) (uint32_t) hi;
	d[3] = hi >> 32;
}
template <> struct Digits<hge> {
	static constexpr int ND = 4;
	__device__ static inline void split(hge x, int, i64 (&d)[4]) { split_hge(x, d); }
};
// fixed point: v = trunc(x * 2^(125 - E)), E = unbiased exponent of the largest |x|
__device__ static inline hge dbl_to_fixed(double x, int E)
{
	u64 bits = (u64) __double_as_longlong(x);
	int e = (int) ((bits >> 52) & 0x7ff);
	u64 m = bits & 0xfffffffffffffULL;
	if (e == 0)
		e = 1;		/* subnormal: no implicit bit */
	else
		m |= 1ULL << 52;
	/* x = m * 2^(e - 1075); v = m * 2^(e - 1075 + 125 - E) */
	int sh = e - 1075 + 125 - E;
	uhge v;
	if (sh >= 0)
		v = sh < 128 ? (uhge) m << sh : 0;
	else
		v = -sh < 64 ? (uhge) (m >> -sh) : 0;
	return (bits >> 63) ? -(hge) v : (hge) v;
}
template <> struct Digits<dbl> {
	static constexpr int ND = 4;
	__device__ static inline void split(dbl x, int E, i64 (&d)[4]) { split_hge(dbl_to_fixed(x, E), d); }
};
template <> struct Digits<flt> {
	static constexpr int ND = 4;
	__device__ static inline void split(flt x, int E, i64 (&d)[4]) { split_hge(dbl_to_fixed((double) x, E), d); }
};

// ---------------------------------------------------------------- shared descriptors
struct AggrIn {
	const void *vals;	/* dense: offset to the first candidate row; gather: column base */
	const oid *cand;	/* candidate oids or NULL */
	oid hseq;		/* subtracted from candidate oids */
	const oid *gids;	/* group id per candidate row, or NULL */
	const unsigned char *g8;	/* one-byte twin of gids (SHADOW_GID8), preferred when present */
	oid gmin, gmax;		/* valid group-id range; cell index = gid - gmin */
	oid gdense;		/* gids == NULL: gid of row k = gdense + k (ngrp == 1: gmin) */
	bool gconst;		/* gids == NULL and every row is in group gmin */
	uint64_t n;
	int fexp;		/* fixed-point exponent (flt/dbl) */
	bool nonil, skip_nils, check;
};

// cell table layout (cell-major): cells[c * ngrp + g]
//   c in [0,ND)        digit sums of all non-nil values
//   c in [ND,2ND)      digit sums of the positive values (only when check)
//   next               count of non-nil rows
//   next 2             (niltrack) first non-nil row (min), last nil row + 1 (max)
struct AggrLayout {
	int nd, ncell, c_pos, c_cnt, c_first, c_lastnil;
	__host__ __device__ AggrLayout(int ND, bool check, bool niltrack)
	{
		nd = ND;
		c_pos = check ? ND : -1;
		c_cnt = check ? 2 * ND : ND;
		c_first = niltrack ? c_cnt + 1 : -1;
		c_lastnil = niltrack ? c_cnt + 2 : -1;
		ncell = c_cnt + 1 + (niltrack ? 2 : 0);
	}
};

template <typename T1>
__device__ __forceinline__ bool aggr_fetch(const AggrIn &in, uint64_t k, T1 &x, oid &gid)
{
	const T1 *vals = (const T1 *) in.vals;
	x = in.cand ? vals[__ldcs(in.cand + k) - in.hseq] : ld_stream(vals + k);
	gid = in.g8 ? (oid) in.g8[k] : in.gids ? __ldcs(in.gids + k) : (in.gconst ? in.gmin : in.gdense + k);
	return gid >= in.gmin && gid <= in.gmax;
}

template <typename T> __device__ __forceinline__ bool is_special(T) { return false; }
template <> __device__ __forceinline__ bool is_special<flt>(flt v) { return isinf(v); }
template <> __device__ __forceinline__ bool is_special<dbl>(dbl v) { return isinf(v); }

// ---------------------------------------------------------------- REG: ngrp <= NG <= 8
#define AGGR_THREADS 256
template <typename T1, int NG, bool CHECK>
__global__ void __launch_bounds__(AGGR_THREADS)
aggr_reg_kernel(AggrIn in, u64 *__restrict__ cells, unsigned ngrp, unsigned *__restrict__ flags)
{
	constexpr int ND = Digits<T1>::ND;
	constexpr int NA = CHECK ? 2 * ND : ND;
	i64 acc[NG][NA];
	unsigned cnt[NG];
#pragma unroll
	for (int g = 0; g < NG; g++) {
		cnt[g] = 0;
#pragma unroll
		for (int d = 0; d < NA; d++)
			acc[g][d] = 0;
	}
	bool special = false;
	constexpr int U = 4;
	const uint64_t stride = (uint64_t) gridDim.x * AGGR_THREADS;
	for (uint64_t k0 = (uint64_t) blockIdx.x * AGGR_THREADS + threadIdx.x; k0 < in.n; k0 += stride * U) {
		T1 x[U];
		oid gid[U];
		bool ok[U];
#pragma unroll
		for (int u = 0; u < U; u++) {
			const uint64_t k = k0 + u * stride;
			ok[u] = k < in.n;
			if (ok[u])
				ok[u] = aggr_fetch<T1>(in, k, x[u], gid[u]);
		}
#pragma unroll
		for (int u = 0; u < U; u++) {
			if (!ok[u] || Atom<T1>::is_nil(x[u]))
				continue;
			special |= is_special(x[u]);
			i64 d[ND];
			Digits<T1>::split(x[u], in.fexp, d);
			const unsigned gi = (unsigned) (gid[u] - in.gmin);
			const bool pos = x[u] > 0;
#pragma unroll
			for (int g = 0; g < NG; g++) {
				const bool m = gi == g;
				cnt[g] += m;
#pragma unroll
				for (int e = 0; e < ND; e++) {
					acc[g][e] += m ? d[e] : 0;
					if (CHECK)
						acc[g][ND + e] += (m && pos) ? d[e] : 0;
				}
			}
		}
	}
	/* warp reduce, then one RED per cell per warp-leader through shared memory */
	__shared__ i64 wsum[AGGR_THREADS / 32][NG * (NA + 1)];
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
	for (int g = 0; g < NG; g++) {
#pragma unroll
		for (int e = 0; e <= NA; e++) {
			i64 v = e < NA ? acc[g][e < NA ? e : 0] : (i64) cnt[g];
#pragma unroll
			for (int o = 16; o > 0; o >>= 1)
				v += __shfl_down_sync(0xffffffffu, v, o);
			if (lane == 0)
				wsum[warp][g * (NA + 1) + e] = v;
		}
	}
	if (__any_sync(0xffffffffu, special) && lane == 0)
		atomicOr(flags, 1u);
	__syncthreads();
	for (unsigned i = threadIdx.x; i < NG * (NA + 1); i += AGGR_THREADS) {
		const unsigned g = i / (NA + 1), e = i % (NA + 1);
		if (g >= ngrp)
			continue;
		i64 v = 0;
#pragma unroll
		for (int w = 0; w < AGGR_THREADS / 32; w++)
			v += wsum[w][i];
		if (v != 0)
			atomicAdd(&cells[(size_t) e * ngrp + g], (u64) v);	/* cell order: S digits, P digits, count */
	}
}

// ---------------------------------------------------------------- REG, fast form
// Integer values, dense value column, ngrp <= NG <= 8, no nil tracking.  Each thread
// keeps ONE wide two's-complement accumulator per group (64 bits for <= 4-byte
// values, 128 for lng, 192 for hge) that a row updates with predicated adds — a third
// of the instructions of the digit form above — plus the largest |x| it has seen:
// if rows x max|x| fits the result type no prefix of any group can overflow, which
// replaces the separate positive-part sums in all practical cases (otherwise the
// host falls back to the exact form).  The wide sums are cut into the same 32-bit
// digit cells at the end, so the finishing kernels are shared.
template <typename T1> struct WideAcc {
	static constexpr int NW = sizeof(T1) <= 4 ? 1 : sizeof(T1) == 8 ? 2 : 3;	/* 64-bit words */
};

// NARROW (hge only): every value is expected to fit 64 bits (DECIMAL(38) columns holding ordinary
// magnitudes): the accumulators are two words like the lng form's, and a value that does not fit
// raises maxabs_out[2] so that the host reruns the three-word form.
template <typename T1, int NG, bool HASGID, bool NARROW = false, typename GT = oid>
__global__ void __launch_bounds__(AGGR_THREADS)
aggr_reg_fast_kernel(const T1 *__restrict__ vals, const GT *__restrict__ gids, oid gmin, oid gspan /* gmax - gmin */,
		     uint64_t n, bool nonil, u64 *__restrict__ cells, unsigned ngrp, u64 *__restrict__ maxabs_out)
{
	static_assert(!NARROW || sizeof(T1) == 16, "NARROW is the hge form");
	constexpr int NW = NARROW ? 2 : WideAcc<T1>::NW;
	constexpr int ND = Digits<T1>::ND;
	typedef typename std::conditional<sizeof(T1) == 16 && !NARROW, uhge, u64>::type UA;
	bool anywide = false;
	u64 acc[NG][NW];
	unsigned cnt[NG];
#pragma unroll
	for (int g = 0; g < NG; g++) {
		cnt[g] = 0;
#pragma unroll
		for (int w = 0; w < NW; w++)
			acc[g][w] = 0;
	}
	UA maxabs = 0;
	u64 ormax = 0;
	/* one-byte ids: a thread takes 8 CONSECUTIVE rows, their ids are one 8-byte load and their
	 * values 16-byte loads (vals must be 16-byte aligned; byte-wide LDGs measured far slower) */
	constexpr bool G8 = HASGID && sizeof(GT) == 1;
	constexpr int U = G8 ? 8 : (sizeof(T1) == 16 && !NARROW ? 4 : 8);
	const uint64_t stride = (uint64_t) gridDim.x * AGGR_THREADS;
	const uint64_t kstep = G8 ? stride : stride * U;
	const uint64_t kend = G8 ? (n + 7) / 8 : n;
	for (uint64_t k0 = (uint64_t) blockIdx.x * AGGR_THREADS + threadIdx.x; k0 < kend; k0 += kstep) {
		T1 x[U];
		oid gid[U];
		if constexpr (G8) {
			const uint64_t r0 = k0 * 8;
			if (r0 + 8 <= n) {
				union {
					T1 t[8];
					uint4 q[sizeof(T1) >= 2 ? sizeof(T1) / 2 : 1];
					u64 w;
				} v;
				if constexpr (sizeof(T1) == 1) {
					v.w = __ldcs(reinterpret_cast<const u64 *>(vals + r0));
				} else {
#pragma unroll
					for (int i = 0; i < (int) sizeof(T1) / 2; i++)
						v.q[i] = __ldg(reinterpret_cast<const uint4 *>(vals + r0) + i);
				}
				const u64 gw = __ldcs(reinterpret_cast<const u64 *>(gids + r0));
#pragma unroll
				for (int u = 0; u < 8; u++) {
					x[u] = v.t[u];
					gid[u] = (gw >> (8 * u)) & 0xff;
				}
			} else {
#pragma unroll
				for (int u = 0; u < 8; u++) {
					const bool in = r0 + u < n;
					x[u] = in ? vals[r0 + u] : Atom<T1>::nil();
					gid[u] = in ? (oid) gids[r0 + u] : gmin;
				}
			}
		} else {
#pragma unroll
			for (int u = 0; u < U; u++) {
				const uint64_t k = k0 + u * stride;
				if (k < n) {
					x[u] = ld_stream(vals + k);
					gid[u] = HASGID ? (oid) __ldcs(gids + k) : gmin;
				} else {
					x[u] = Atom<T1>::nil();
					gid[u] = gmin;
				}
			}
		}
#pragma unroll
		for (int u = 0; u < U; u++) {
			const u64 go = gid[u] - gmin;
			const bool ok = !Atom<T1>::is_nil(x[u]) && go <= gspan;		/* a tail slot holds nil */
			const unsigned gi = ok ? (unsigned) go : 0xffffffffu;
			if constexpr (sizeof(UA) == 8) {
				/* |x| (or |x| - 1) OR-ed together: at most twice max |x|, in 3 instructions instead of the
				 * 9 of abs + compare + select — the bound check only needs an upper estimate */
				i64 lo;
				if constexpr (NARROW) {
					lo = (i64) (u64) x[u];
					anywide |= ok && (i64) ((uhge) x[u] >> 64) != (lo >> 63);
				} else {
					lo = (i64) x[u];
				}
				ormax |= ok ? (u64) (lo ^ (lo >> 63)) : 0;
			} else {
				const UA ax = x[u] < 0 ? (UA) 0 - (UA) x[u] : (UA) x[u];
				if (ok && ax > maxabs)
					maxabs = ax;
			}
			/* sign-extended words of x */
			u64 xw[NW];
			if constexpr (NARROW) {
				xw[0] = (u64) x[u];
				xw[1] = (u64) ((i64) xw[0] >> 63);
			} else if constexpr (sizeof(T1) == 16) {
				xw[0] = (u64) x[u];
				xw[1] = (u64) ((uhge) x[u] >> 64);
				xw[2] = (u64) ((i64) xw[1] >> 63);
			} else if constexpr (sizeof(T1) == 8) {
				xw[0] = (u64) x[u];
				xw[1] = (u64) ((i64) x[u] >> 63);
			} else {
				xw[0] = (u64) (i64) x[u];
			}
#pragma unroll
			for (int g = 0; g < NG; g++) {
				if (gi == g) {
					cnt[g]++;
					if constexpr (NW == 1) {
						acc[g][0] += xw[0];
					} else if constexpr (NW == 2) {
						const uhge t = (((uhge) acc[g][1] << 64) | acc[g][0]) + (((uhge) xw[1] << 64) | xw[0]);
						acc[g][0] = (u64) t;
						acc[g][1] = (u64) (t >> 64);
					} else {
						const uhge t = (uhge) acc[g][0] + xw[0];
						acc[g][0] = (u64) t;
						const uhge t2 = (((uhge) acc[g][2] << 64) | acc[g][1]) + (((uhge) xw[2] << 64) | xw[1]) + (u64) (t >> 64);
						acc[g][1] = (u64) t2;
						acc[g][2] = (u64) (t2 >> 64);
					}
				}
			}
		}
	}
	if constexpr (sizeof(UA) == 8)
		maxabs = (UA) (ormax + (ormax != 0));	/* >= max |x|; 0 when no value was seen */
	/* wide sum -> ND digit cells: digits 0..ND-2 are the low 32-bit pieces, the top
	 * digit is the arithmetic remainder (fits 64 bits: |sum| < 2^(32 ND + 31)) */
	__shared__ i64 wsum[AGGR_THREADS / 32][NG * (ND + 1)];
	__shared__ u64 wmax[AGGR_THREADS / 32][2];
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
	for (int g = 0; g < NG; g++) {
		i64 d[ND];
		if constexpr (ND == 1) {
			d[0] = (i64) acc[g][0];
		} else if constexpr (ND == 2) {
			d[0] = (i64) (uint32_t) acc[g][0];
			d[1] = (i64) ((acc[g][0] >> 32) | (acc[g][1] << 32));
		} else {
			const u64 top = NARROW ? (u64) ((i64) acc[g][1] >> 63) : acc[g][NW - 1];
			d[0] = (i64) (uint32_t) acc[g][0];
			d[1] = (i64) (acc[g][0] >> 32);
			d[2] = (i64) (uint32_t) acc[g][1];
			d[3] = (i64) ((acc[g][1] >> 32) | (top << 32));
		}
#pragma unroll
		for (int e = 0; e <= ND; e++) {
			i64 v = e < ND ? d[e < ND ? e : 0] : (i64) cnt[g];
#pragma unroll
			for (int o = 16; o > 0; o >>= 1)
				v += __shfl_down_sync(0xffffffffu, v, o);
			if (lane == 0)
				wsum[warp][g * (ND + 1) + e] = v;
		}
	}
	{
		u64 mlo = (u64) maxabs, mhi = sizeof(UA) == 16 ? (u64) ((uhge) maxabs >> 64) : 0;
		if (NARROW && anywide)
			maxabs_out[2] = 1;
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) {
			const u64 olo = __shfl_down_sync(0xffffffffu, mlo, o), ohi = __shfl_down_sync(0xffffffffu, mhi, o);
			if (ohi > mhi || (ohi == mhi && olo > mlo)) {
				mhi = ohi;
				mlo = olo;
			}
		}
		if (lane == 0) {
			wmax[warp][0] = mlo;
			wmax[warp][1] = mhi;
		}
	}
	__syncthreads();
	for (unsigned i = threadIdx.x; i < NG * (ND + 1); i += AGGR_THREADS) {
		const unsigned g = i / (ND + 1), e = i % (ND + 1);
		if (g >= ngrp)
			continue;
		i64 v = 0;
#pragma unroll
		for (int w = 0; w < AGGR_THREADS / 32; w++)
			v += wsum[w][i];
		if (v != 0)
			atomicAdd(&cells[(size_t) e * ngrp + g], (u64) v);	/* cell order: S digits, count */
	}
	if (threadIdx.x == 0) {
		u64 mlo = 0, mhi = 0;
		for (int w = 0; w < AGGR_THREADS / 32; w++)
			if (wmax[w][1] > mhi || (wmax[w][1] == mhi && wmax[w][0] > mlo)) {
				mhi = wmax[w][1];
				mlo = wmax[w][0];
			}
		/* (hi, lo) maximum with two ordered atomics; the bound only needs an upper estimate */
		const u64 oldhi = atomicMax(&maxabs_out[1], mhi);
		if (mhi >= oldhi)
			atomicMax(&maxabs_out[0], mlo);		/* a stale larger lo only makes the bound more careful */
	}
}

// ---------------------------------------------------------------- SHM / GLB: table of cells
// USE_SHM: the block accumulates into a shared-memory copy of the table and flushes it
// once; otherwise rows go straight to the global table with REDs.
template <typename T1, bool USE_SHM>
__global__ void __launch_bounds__(AGGR_THREADS)
aggr_tab_kernel(AggrIn in, u64 *__restrict__ cells, unsigned ngrp, bool niltrack, unsigned *__restrict__ flags)
{
	constexpr int ND = Digits<T1>::ND;
	extern __shared__ u64 sh_cells[];
	const AggrLayout L(ND, in.check, niltrack);
	u64 *tab = cells;
	if (USE_SHM) {
		tab = sh_cells;
		for (unsigned i = threadIdx.x; i < (unsigned) L.ncell * ngrp; i += AGGR_THREADS)
			tab[i] = (L.c_first >= 0 && i / ngrp == (unsigned) L.c_first) ? ~0ULL : 0;
		__syncthreads();
	}
	bool special = false;
	const uint64_t stride = (uint64_t) gridDim.x * AGGR_THREADS;
	for (uint64_t k = (uint64_t) blockIdx.x * AGGR_THREADS + threadIdx.x; k < in.n; k += stride) {
		T1 x;
		oid gid;
		if (!aggr_fetch<T1>(in, k, x, gid))
			continue;
		const size_t gi = (size_t) (gid - in.gmin);
		if (Atom<T1>::is_nil(x)) {
			if (niltrack)
				atomicMax(&tab[(size_t) L.c_lastnil * ngrp + gi], (u64) k + 1);
			continue;
		}
		special |= is_special(x);
		i64 d[ND];
		Digits<T1>::split(x, in.fexp, d);
#pragma unroll
		for (int e = 0; e < ND; e++)
			if (d[e] != 0)
				atomicAdd(&tab[(size_t) e * ngrp + gi], (u64) d[e]);
		if (in.check && x > 0) {
#pragma unroll
			for (int e = 0; e < ND; e++)
				if (d[e] != 0)
					atomicAdd(&tab[(size_t) (ND + e) * ngrp + gi], (u64) d[e]);
		}
		atomicAdd(&tab[(size_t) L.c_cnt * ngrp + gi], 1ULL);
		if (niltrack)
			atomicMin(&tab[(size_t) L.c_first * ngrp + gi], (u64) k);
	}
	if (__any_sync(0xffffffffu, special) && (threadIdx.x & 31) == 0)
		atomicOr(flags, 1u);
	if (USE_SHM) {
		__syncthreads();
		for (unsigned i = threadIdx.x; i < (unsigned) L.ncell * ngrp; i += AGGR_THREADS) {
			const int c = i / ngrp;
			const u64 v = tab[i];
			if (c == L.c_first) {
				if (v != ~0ULL)
					atomicMin(&cells[i], v);
			} else if (c == L.c_lastnil) {
				if (v)
					atomicMax(&cells[i], v);
			} else if (v)
				atomicAdd(&cells[i], v);
		}
	}
}

// ---------------------------------------------------------------- SHM32: shared table of split cells
// 64-bit shared-memory atomics compile to a compare-and-swap loop (SASS ATOMS.CAST.SPIN.64);
// the native shared atomic is 32 bits wide.  Without nil tracking every cell is a plain
// 64-bit sum, so it is kept as two 32-bit halves in separate arrays: a row adds the low
// half of its digit with one ATOMS.ADD, and touches the high half only when that add
// wrapped (the returned old value tells) or the digit is negative.  The number of wraps
// counted is exactly floor(sum of low halves / 2^32), so lo | hi << 32 at the flush is
// the same 64-bit sum modulo 2^64 the 64-bit table holds.
__device__ static inline void shm32_add(unsigned *lo, unsigned *hi, unsigned idx, u64 v)
{
	const unsigned vl = (unsigned) v;
	unsigned vh = (unsigned) (v >> 32);
	if (vl != 0) {
		const unsigned old = atomicAdd(&lo[idx], vl);
		vh += (old + vl) < old;
	}
	if (vh != 0)
		atomicAdd(&hi[idx], vh);
}

template <typename T1>
__global__ void __launch_bounds__(AGGR_THREADS)
aggr_tab32_kernel(AggrIn in, u64 *__restrict__ cells, unsigned ngrp, unsigned *__restrict__ flags)
{
	constexpr int ND = Digits<T1>::ND;
	extern __shared__ u64 sh_cells[];
	const AggrLayout L(ND, in.check, false);
	const unsigned ncells = (unsigned) L.ncell * ngrp;
	unsigned *lo = (unsigned *) sh_cells, *hi = lo + ncells;
	for (unsigned i = threadIdx.x; i < 2 * ncells; i += AGGR_THREADS)
		lo[i] = 0;
	__syncthreads();
	bool special = false;
	const uint64_t stride = (uint64_t) gridDim.x * AGGR_THREADS;
	constexpr int U = 4;	/* rows in flight per thread */
	for (uint64_t k0 = (uint64_t) blockIdx.x * AGGR_THREADS + threadIdx.x; k0 < in.n; k0 += stride * U) {
		T1 x[U];
		oid gid[U];
		bool ok[U];
#pragma unroll
		for (int u = 0; u < U; u++) {
			const uint64_t k = k0 + u * stride;
			ok[u] = k < in.n && aggr_fetch<T1>(in, k, x[u], gid[u]);
		}
#pragma unroll
		for (int u = 0; u < U; u++) {
			if (!ok[u] || Atom<T1>::is_nil(x[u]))
				continue;
			const unsigned gi = (unsigned) (gid[u] - in.gmin);
			special |= is_special(x[u]);
			i64 d[ND];
			Digits<T1>::split(x[u], in.fexp, d);
#pragma unroll
			for (int e = 0; e < ND; e++)
				if (d[e] != 0)
					shm32_add(lo, hi, e * ngrp + gi, (u64) d[e]);
			if (in.check && x[u] > 0) {
#pragma unroll
				for (int e = 0; e < ND; e++)
					if (d[e] != 0)
						shm32_add(lo, hi, (ND + e) * ngrp + gi, (u64) d[e]);
			}
			shm32_add(lo, hi, L.c_cnt * ngrp + gi, 1ULL);
		}
	}
	if (__any_sync(0xffffffffu, special) && (threadIdx.x & 31) == 0)
		atomicOr(flags, 1u);
	__syncthreads();
	for (unsigned i = threadIdx.x; i < ncells; i += AGGR_THREADS) {
		const u64 v = (u64) lo[i] | (u64) hi[i] << 32;
		if (v)
			atomicAdd(&cells[i], v);
	}
}

// ---------------------------------------------------------------- wide integers
struct I192 {		/* two's complement, little endian words */
	u64 w[3];
	__host__ __device__ bool neg() const { return (i64) w[2] < 0; }
};
__host__ __device__ static inline void i192_add_shifted(I192 &a, u64 v, bool sign_extend, int digit)
{
	/* a += (sign-extended or zero-extended v) << (32 * digit) */
	u64 ext = (sign_extend && (i64) v < 0) ? ~0ULL : 0;
	u64 add[4] = {0, 0, 0, 0};	/* 256-bit scratch, little endian */
	int word = digit >> 1;
	if (digit & 1) {
		add[word] = v << 32;
		if (word + 1 < 4) add[word + 1] = (v >> 32) | (ext << 32);
		for (int i = word + 2; i < 4; i++) add[i] = ext;
	} else {
		add[word] = v;
		for (int i = word + 1; i < 4; i++) add[i] = ext;
	}
	unsigned carry = 0;
	for (int i = 0; i < 3; i++) {
		u64 s = a.w[i] + add[i];
		unsigned c1 = s < a.w[i];
		u64 s2 = s + carry;
		unsigned c2 = s2 < s;
		a.w[i] = s2;
		carry = c1 | c2;
	}
}
__host__ __device__ static inline I192 i192_from_cells(const u64 *cells, size_t stride, int nd)
{
	I192 a = {{0, 0, 0}};
	for (int e = 0; e < nd; e++)
		i192_add_shifted(a, cells[(size_t) e * stride], e == nd - 1, e);	/* only the top digit is signed */
	return a;
}
__host__ __device__ static inline I192 i192_neg(I192 a)
{
	I192 r;
	unsigned carry = 1;
	for (int i = 0; i < 3; i++) {
		u64 s = ~a.w[i] + carry;
		carry = carry && s == 0;
		r.w[i] = s;
	}
	return r;
}
__host__ __device__ static inline I192 i192_sub(I192 a, I192 b)
{
	I192 nb = i192_neg(b), r;
	unsigned carry = 0;
	for (int i = 0; i < 3; i++) {
		u64 s = a.w[i] + nb.w[i];
		unsigned c1 = s < a.w[i];
		u64 s2 = s + carry;
		unsigned c2 = s2 < s;
		r.w[i] = s2;
		carry = c1 | c2;
	}
	return r;
}
// does a fit in [-max,max] of a signed integer type of `bits` bits (8..128)?
__host__ __device__ static inline bool i192_fits(const I192 &a, int bits)
{
	I192 m = a.neg() ? i192_neg(a) : a;	/* magnitude; |min192| cannot occur here */
	/* need m <= 2^(bits-1) - 1, i.e. m >> (bits-1) == 0 */
	int sh = bits - 1;
	for (int i = 0; i < 3; i++) {
		int lo = i * 64;
		if (sh <= lo) {
			if (m.w[i]) return false;
		} else if (sh < lo + 64) {
			if (m.w[i] >> (sh - lo)) return false;
		}
	}
	return true;
}
__host__ __device__ static inline hge i192_to_hge(const I192 &a)
{
	return (hge) (((uhge) a.w[1] << 64) | a.w[0]);
}
// V * 2^exp2 rounded to nearest-even double (may return +-inf)
__device__ static inline double i192_to_double(const I192 &a, int exp2)
{
	bool neg = a.neg();
	I192 m = neg ? i192_neg(a) : a;
	int top = -1;
	for (int i = 2; i >= 0; i--)
		if (m.w[i]) { top = i * 64 + 63 - __clzll((long long) m.w[i]); break; }
	if (top < 0)
		return 0.0;
	u64 mant;
	int shift = top - 52;		/* bits dropped */
	if (shift <= 0) {
		mant = m.w[0] << -shift;	/* top < 53: value lives in w[0] */
	} else {
		/* extract bits [shift, shift+53) and round on the rest */
		int wi = shift >> 6, bi = shift & 63;
		u64 lo = m.w[wi] >> bi;
		if (bi && wi + 1 < 3)
			lo |= m.w[wi + 1] << (64 - bi);
		mant = lo & ((1ULL << 53) - 1);
		/* round bit and sticky */
		int rb = shift - 1;
		bool round = (m.w[rb >> 6] >> (rb & 63)) & 1;
		bool sticky = false;
		for (int i = 0; i < 3; i++) {
			int lo_bit = i * 64;
			if (rb <= lo_bit)
				break;
			u64 mask = rb - lo_bit >= 64 ? ~0ULL : ((1ULL << (rb - lo_bit)) - 1);
			if (m.w[i] & mask)
				sticky = true;
		}
		if (round && (sticky || (mant & 1)))
			mant++;		/* may become 2^53: still exactly representable */
	}
	double r = ldexp((double) mant, (shift > 0 ? shift : shift) + exp2);
	return neg ? -r : r;
}

// ---------------------------------------------------------------- finishing kernels
// state: [0] flag bits (1 = +-inf input, 2 = overflow), [1] nil groups, [2] groups needing the exact check
#define AGGR_F_SPECIAL 1u
#define AGGR_F_OVERFLOW 2u

template <typename T2> struct OutBits { static constexpr int bits = sizeof(T2) * 8; };

template <typename T2>
__global__ void aggr_finish_sum_kernel(const u64 *__restrict__ cells, unsigned ngrp, int nd, bool check, bool niltrack,
				       bool nil_if_empty, bool float_in, int fexp, T2 *__restrict__ out,
				       u64 *__restrict__ state, unsigned *__restrict__ exactlist)
{
	const unsigned g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= ngrp)
		return;
	const AggrLayout L(nd, check, niltrack);
	const u64 cnt = cells[(size_t) L.c_cnt * ngrp + g];
	bool isnil = false;
	if (niltrack) {
		const u64 first = cells[(size_t) L.c_first * ngrp + g], lastnil = cells[(size_t) L.c_lastnil * ngrp + g];
		/* single group: any nil; several groups: a nil after the group's first value
		 * (AGGR_SUM resets sums[gid] when it sees the first value, gdk_aggr.c:516-520) */
		if (lastnil != 0)
			isnil = float_in /* dofsum: any nil ends the group, gdk_aggr.c:253-264 */ ||
				ngrp == 1 || cnt == 0 || lastnil - 1 > first;
		if (lastnil != 0)
			atomicAdd(&state[3], 1ULL);	/* the reference counts nils as it meets them */
	}
	T2 res = Atom<T2>::nil();
	if (cnt == 0) {
		if (!isnil && !nil_if_empty)
			res = 0;
		else
			isnil = true;
	} else {
		const I192 S = i192_from_cells(cells + g, ngrp, nd);
		if constexpr (Atom<T2>::is_float) {
			if (!isnil) {
				double r = i192_to_double(S, fexp - 125);
				if constexpr (sizeof(T2) == 4) {
					res = (flt) r;
					if (isinf(res) || isnan(res))
						atomicOr((unsigned *) &state[0], AGGR_F_OVERFLOW);
				} else {
					res = r;
					if (isinf(r) || isnan(r))
						atomicOr((unsigned *) &state[0], AGGR_F_OVERFLOW);
				}
			}
		} else {
			constexpr int bits = OutBits<T2>::bits;
			if (!isnil && !i192_fits(S, bits)) {
				atomicOr((unsigned *) &state[0], AGGR_F_OVERFLOW);
			} else if (check) {
				/* also for groups that end up nil: the reference adds (and checks) the
				 * values it meets before the nil that ends the group */
				const I192 P = i192_from_cells(cells + (size_t) nd * ngrp + g, ngrp, nd);
				const I192 M = i192_sub(P, S);
				if (!i192_fits(P, bits) || !i192_fits(M, bits)) {
					unsigned slot = (unsigned) atomicAdd(&state[2], 1ULL);
					exactlist[slot] = g;
				}
			}
			if (!isnil) {
				if constexpr (sizeof(T2) == 16)
					res = i192_to_hge(S);
				else
					res = (T2) (i64) S.w[0];
			}
		}
	}
	if (isnil) {
		res = Atom<T2>::nil();
		atomicAdd(&state[1], 1ULL);
	}
	out[g] = res;
}

// integer / float average as dbl, plus counts (gdk_aggr.c:1717-1787, 1952-1957)
template <typename T1>
__global__ void aggr_finish_avg_kernel(const u64 *__restrict__ cells, unsigned ngrp, bool niltrack, int fexp, double fac,
				       dbl *__restrict__ out, lng *__restrict__ cnts, u64 *__restrict__ state,
				       bool plain = false /* BATcalcavg: (dbl) sum / n while the sum fits a hge (gdk_aggr.c:2924) */)
{
	const unsigned g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= ngrp)
		return;
	constexpr int ND = Digits<T1>::ND;
	const AggrLayout L(ND, false, niltrack);
	u64 n = cells[(size_t) L.c_cnt * ngrp + g];
	if (niltrack && cells[(size_t) L.c_lastnil * ngrp + g] != 0)
		n = 0;			/* cnts[gid] = lng_nil is sticky */
	if (cnts)
		cnts[g] = (lng) n;
	if (n == 0) {
		out[g] = NAN;
		atomicAdd(&state[1], 1ULL);
		return;
	}
	const I192 S = i192_from_cells(cells + g, ngrp, ND);
	double r;
	if constexpr (Atom<T1>::is_float) {
		r = i192_to_double(S, fexp - 125) / (double) n;
	} else {
		if (plain && i192_fits(S, 128)) {
			out[g] = (double) i192_to_hge(S) / (double) (lng) n;
			return;
		}
		/* a = floor(S / n), rem = S - a * n  (0 <= rem < n), n < 2^32 */
		const bool neg = S.neg();
		I192 m = neg ? i192_neg(S) : S;
		uint32_t dig[6];
		for (int i = 0; i < 3; i++) {
			dig[2 * i] = (uint32_t) m.w[i];
			dig[2 * i + 1] = (uint32_t) (m.w[i] >> 32);
		}
		u64 rem = 0;
		for (int i = 5; i >= 0; i--) {
			u64 cur = (rem << 32) | dig[i];
			dig[i] = (uint32_t) (cur / n);
			rem = cur % n;
		}
		I192 q;
		for (int i = 0; i < 3; i++)
			q.w[i] = (u64) dig[2 * i] | ((u64) dig[2 * i + 1] << 32);
		if (neg) {
			q = i192_neg(q);
			if (rem != 0) {
				I192 one = {{1, 0, 0}};
				q = i192_sub(q, one);
				rem = n - rem;
			}
		}
		double a;
		if constexpr (sizeof(T1) == 16)
			a = (double) i192_to_hge(q);
		else
			a = (double) (T1) (i64) q.w[0];
		r = a + (double) (lng) rem / (double) (lng) n;
	}
	if (fac != 1.0)
		r /= fac;
	out[g] = r;
}

__global__ void aggr_finish_count_kernel(const u64 *__restrict__ cnt, unsigned ngrp, lng *__restrict__ out)
{
	const unsigned g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g < ngrp)
		out[g] = (lng) cnt[g];
}

// ---------------------------------------------------------------- exact running-sum check
// For the (rare) groups whose positive or negative part alone exceeds the result
// type, decide exactly whether some prefix of the row-ordered running sum leaves
// [-max,max] (ADDI_WITH_CHECK inside AGGR_SUM).  (sum, max prefix, min prefix)
// triples compose associatively, so an ordered tree reduction gives the answer.
struct Pfx { I192 s, mx, mn; };
__device__ static inline I192 i192_add(I192 a, I192 b)
{
	I192 r;
	unsigned carry = 0;
	for (int i = 0; i < 3; i++) {
		u64 s = a.w[i] + b.w[i];
		unsigned c1 = s < a.w[i];
		u64 s2 = s + carry;
		unsigned c2 = s2 < s;
		r.w[i] = s2;
		carry = c1 | c2;
	}
	return r;
}
__device__ static inline bool i192_less(const I192 &a, const I192 &b)
{
	if (a.w[2] != b.w[2]) return (i64) a.w[2] < (i64) b.w[2];
	if (a.w[1] != b.w[1]) return a.w[1] < b.w[1];
	return a.w[0] < b.w[0];
}
__device__ static inline Pfx pfx_compose(const Pfx &a, const Pfx &b)
{
	Pfx r;
	r.s = i192_add(a.s, b.s);
	I192 t = i192_add(a.s, b.mx);
	r.mx = i192_less(a.mx, t) ? t : a.mx;
	t = i192_add(a.s, b.mn);
	r.mn = i192_less(t, a.mn) ? t : a.mn;
	return r;
}
// With skip_nils == false a nil ends a group's summation: in the single-group loops
// the first nil (gdk_aggr.c:447-452,481-486), in the grouped loops the first nil
// that follows the group's first value (:506-525); only the values met before
// that point are added and checked.  seen / lead / stopped carry that through the
// ordered reduction.
struct PfxN {
	Pfx p;
	unsigned char seen, lead, stopped;
};
__device__ static inline PfxN pfxn_compose(const PfxN &a, const PfxN &b, bool single)
{
	if (a.stopped)
		return a;
	if (b.lead && (a.seen || single)) {
		PfxN r = a;
		r.stopped = 1;
		return r;
	}
	PfxN r;
	r.p = pfx_compose(a.p, b.p);
	r.seen = a.seen | b.seen;
	r.lead = a.lead | (!a.seen && b.lead);
	r.stopped = b.stopped;
	return r;
}
template <typename T1>
__global__ void __launch_bounds__(256)
aggr_exact_check_kernel(AggrIn in, const unsigned *__restrict__ exactlist, int bits, u64 *__restrict__ state)
{
	constexpr int R = 8;
	__shared__ PfxN sh[256];
	const oid want = in.gmin + exactlist[blockIdx.x];
	const bool single = in.gmin == in.gmax, stop_on_nil = !in.skip_nils;
	const PfxN zero = {{{{0, 0, 0}}, {{0, 0, 0}}, {{0, 0, 0}}}, 0, 0, 0};
	PfxN run = zero;
	for (uint64_t base = 0; base < in.n; base += 256 * R) {
		PfxN q = zero;
		for (int r = 0; r < R; r++) {
			const uint64_t k = base + (uint64_t) threadIdx.x * R + r;
			T1 x;
			oid gid;
			if (k >= in.n || !aggr_fetch<T1>(in, k, x, gid) || gid != want || q.stopped)
				continue;
			if (Atom<T1>::is_nil(x)) {
				if (stop_on_nil) {
					if (q.seen || single)
						q.stopped = 1;
					if (!q.seen)
						q.lead = 1;
				}
				continue;
			}
			I192 v;
			hge hx = (hge) x;
			v.w[0] = (u64) hx;
			v.w[1] = (u64) (hx >> 64);
			v.w[2] = hx < 0 ? ~0ULL : 0;
			q.p.s = i192_add(q.p.s, v);
			if (i192_less(q.p.mx, q.p.s)) q.p.mx = q.p.s;
			if (i192_less(q.p.s, q.p.mn)) q.p.mn = q.p.s;
			q.seen = 1;
		}
		sh[threadIdx.x] = q;
		__syncthreads();
		for (int st = 1; st < 256; st <<= 1) {
			if ((threadIdx.x % (2 * st)) == 0)
				sh[threadIdx.x] = pfxn_compose(sh[threadIdx.x], sh[threadIdx.x + st], single);
			__syncthreads();
		}
		if (threadIdx.x == 0)
			run = pfxn_compose(run, sh[0], single);
		__syncthreads();
	}
	if (threadIdx.x == 0 && (!i192_fits(run.p.mx, bits) || !i192_fits(run.p.mn, bits)))
		atomicOr((unsigned *) &state[0], AGGR_F_OVERFLOW);
}

// ---------------------------------------------------------------- avg3 / avg3combine
// floor(S / n) and S - floor(S / n) * n for a 192-bit S and 0 < n < 2^32 (long division on
// 32-bit digits); larger n goes through the 128-bit divide when S fits a hge.
__device__ static inline bool i192_floordiv(const I192 &S, u64 n, I192 &q, u64 &rem)
{
	if (n >= (1ULL << 32)) {
		if (!i192_fits(S, 128))
			return false;
		const hge s = i192_to_hge(S);
		hge hq = s / (hge) n, hr = s % (hge) n;
		if (hr < 0) {
			hq--;
			hr += (hge) n;
		}
		q.w[0] = (u64) hq;
		q.w[1] = (u64) (hq >> 64);
		q.w[2] = hq < 0 ? ~0ULL : 0;
		rem = (u64) hr;
		return true;
	}
	const bool neg = S.neg();
	const I192 m = neg ? i192_neg(S) : S;
	uint32_t dig[6];
	for (int i = 0; i < 3; i++) {
		dig[2 * i] = (uint32_t) m.w[i];
		dig[2 * i + 1] = (uint32_t) (m.w[i] >> 32);
	}
	rem = 0;
	for (int i = 5; i >= 0; i--) {
		u64 cur = (rem << 32) | dig[i];
		dig[i] = (uint32_t) (cur / n);
		rem = cur % n;
	}
	for (int i = 0; i < 3; i++)
		q.w[i] = (u64) dig[2 * i] | ((u64) dig[2 * i + 1] << 32);
	if (neg) {
		q = i192_neg(q);
		if (rem != 0) {
			I192 one = {{1, 0, 0}};
			q = i192_sub(q, one);
			rem = n - rem;
		}
	}
	return true;
}

// BATgroupavg3's result triple (gdk_aggr.c:1996-2290): avg = floor(S/n) rounded half away from the
// floor as the reference does (rem may become negative), rem, cnt.  state[1] counts nil avgs,
// state[2] the groups that a nil turned nil altogether (skip_nils == false).
template <typename T1>
__global__ void aggr_finish_avg3_kernel(const u64 *__restrict__ cells, unsigned ngrp, bool niltrack, T1 *__restrict__ avgs,
					lng *__restrict__ rems, lng *__restrict__ cnts, u64 *__restrict__ state)
{
	const unsigned g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= ngrp)
		return;
	constexpr int ND = Digits<T1>::ND;
	const AggrLayout L(ND, false, niltrack);
	const u64 n = cells[(size_t) L.c_cnt * ngrp + g];
	if (niltrack && cells[(size_t) L.c_lastnil * ngrp + g] != 0) {
		avgs[g] = Atom<T1>::nil();
		rems[g] = Atom<lng>::nil();
		cnts[g] = Atom<lng>::nil();
		atomicAdd(&state[1], 1ULL);
		atomicAdd(&state[2], 1ULL);
		return;
	}
	cnts[g] = (lng) n;
	if (n == 0) {
		avgs[g] = Atom<T1>::nil();
		rems[g] = 0;
		atomicAdd(&state[1], 1ULL);
		return;
	}
	const I192 S = i192_from_cells(cells + g, ngrp, ND);
	I192 q;
	u64 urem;
	if (!i192_floordiv(S, n, q, urem)) {
		atomicOr((unsigned *) &state[0], AGGR_F_OVERFLOW);
		return;
	}
	T1 a;
	if constexpr (sizeof(T1) == 16)
		a = i192_to_hge(q);
	else
		a = (T1) (i64) q.w[0];
	lng r = (lng) urem;
	if (r > 0) {
		if (a < 0 ? 2 * r > (lng) n : 2 * r >= (lng) n) {
			a++;
			r -= (lng) n;
		}
	}
	avgs[g] = a;
	rems[g] = r;
}

// BATgroupavg3combine (gdk_aggr.c:2634-2900): every input row is a partial (avg, rem, cnt); its
// total avg*cnt + rem (after undoing the rounding when rem < 0, combine_averages_T :2391-2632)
// is added to the group's 128-bit sum as four 32-bit digits.  cells: [0,4) digits, 4 cnt, 5 nil flag.
template <typename T>
__global__ void avg3_combine_kernel(const T *__restrict__ avg, const lng *__restrict__ rem, const lng *__restrict__ cnt,
				    const oid *__restrict__ gids, oid gmin, oid gmax, oid gdense, bool gconst, uint64_t n,
				    bool skip_nils, u64 *__restrict__ cells, unsigned ngrp)
{
	for (uint64_t k = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (uint64_t) gridDim.x * blockDim.x) {
		const oid gid = gids ? gids[k] : (gconst ? gmin : gdense + k);
		if (gid < gmin || gid > gmax)
			continue;
		const unsigned g = (unsigned) (gid - gmin);
		T a = avg[k];
		if (Atom<T>::is_nil(a)) {
			if (!skip_nils)
				atomicOr((unsigned long long *) &cells[(size_t) 5 * ngrp + g], 1ULL);
			continue;
		}
		lng r = rem[k];
		const lng c = cnt[k];
		if (r < 0) {
			a--;
			r += c;
		}
		i64 d[4];
		split_hge((hge) a * (hge) c + (hge) r, d);
#pragma unroll
		for (int e = 0; e < 4; e++)
			if (d[e])
				atomicAdd((unsigned long long *) &cells[(size_t) e * ngrp + g], (unsigned long long) d[e]);
		atomicAdd((unsigned long long *) &cells[(size_t) 4 * ngrp + g], (unsigned long long) c);
	}
}

template <typename T>
__global__ void avg3_combine_finish_kernel(const u64 *__restrict__ cells, unsigned ngrp, T *__restrict__ out, u64 *__restrict__ state)
{
	const unsigned g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= ngrp)
		return;
	const lng n = (lng) cells[(size_t) 4 * ngrp + g];
	if (cells[(size_t) 5 * ngrp + g] != 0 || n == 0) {
		out[g] = Atom<T>::nil();
		atomicAdd(&state[1], 1ULL);
		return;
	}
	const I192 S = i192_from_cells(cells + g, ngrp, 4);
	I192 q;
	u64 urem;
	if (n < 0 || !i192_floordiv(S, (u64) n, q, urem)) {
		atomicOr((unsigned *) &state[0], AGGR_F_OVERFLOW);
		return;
	}
	T a = (T) (i64) q.w[0];
	const lng r = (lng) urem;
	if (r > 0 && (a < 0 ? 2 * r > n : 2 * r >= n))
		a++;
	out[g] = a;
}
