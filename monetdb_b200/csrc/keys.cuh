// keys.cuh — canonical key bits and hashing shared by grouping and join
// (the device counterpart of hash_TYPE / mix_*, gdk_hash.h:159-242; bucket
// choice never shows in results, so the mixers are our own).
#pragma once
#include "common.cuh"

typedef unsigned long long u64;

// ---------------------------------------------------------------- keys
template <typename T> struct KeyBits;		/* canonical bit pattern used for equality + hashing */
#define KEYBITS_INT(T, U)                                                         \
	template <> struct KeyBits<T> {                                           \
		typedef U type;                                                   \
		__device__ static inline U get(T v) { return (U) v; }            \
	};
KEYBITS_INT(bte, uint8_t)
KEYBITS_INT(sht, uint16_t)
KEYBITS_INT(int, uint32_t)
KEYBITS_INT(lng, uint64_t)
KEYBITS_INT(hge, uhge)
template <> struct KeyBits<flt> {
	typedef uint32_t type;
	__device__ static inline uint32_t get(flt v)
	{
		if (v != v)
			return 0x7fc00000u;	/* every NaN is nil, nil == nil (gdk_group.c:179-180) */
		if (v == 0)
			return 0;		/* -0 == +0 */
		return __float_as_uint(v);
	}
};
template <> struct KeyBits<dbl> {
	typedef uint64_t type;
	__device__ static inline uint64_t get(dbl v)
	{
		if (v != v)
			return 0x7ff8000000000000ULL;
		if (v == 0)
			return 0;
		return (uint64_t) __double_as_longlong(v);
	}
};

__device__ __forceinline__ u64 mix64(u64 x)
{
	x ^= x >> 33;
	x *= 0xff51afd7ed558ccdULL;
	x ^= x >> 33;
	x *= 0xc4ceb9fe1a85ec53ULL;
	x ^= x >> 33;
	return x;
}
template <typename U> __device__ __forceinline__ u64 hash_key(U v, u64 g)
{
	if constexpr (sizeof(U) == 16)
		return mix64((u64) v ^ mix64((u64) (v >> 64) + 0x9E3779B97F4A7C15ULL) ^ (g * 0xD6E8FEB86659FD93ULL));
	else
		return mix64((u64) v ^ (g * 0xD6E8FEB86659FD93ULL));
}


// a column seen through a candidate list: candidate position k -> value / oid
template <typename T> struct ColSrc {
	const T *vals;		/* dense: value of position k is vals[k]; else vals[cand[k] - hseq] */
	const oid *cand;
	oid hseq, seq;
	uint64_t n;
	__device__ __forceinline__ T value(uint64_t k) const { return cand ? vals[cand[k] - hseq] : vals[k]; }
	__device__ __forceinline__ oid rowoid(uint64_t k) const { return cand ? cand[k] : seq + k; }
};
template <typename T> static inline ColSrc<T> col_src(const b200_bat *b, const Cand &ci)
{
	ColSrc<T> s;
	s.hseq = ci.hseq;
	s.seq = ci.seq;
	s.n = ci.ncand;
	if (ci.dense) {
		s.vals = (const T *) b->base + (ci.seq - ci.hseq);
		s.cand = nullptr;
	} else {
		s.vals = (const T *) b->base;
		s.cand = ci.oids;
	}
	return s;
}
