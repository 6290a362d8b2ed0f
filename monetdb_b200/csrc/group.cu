// group.cu — BATgroup on the device (gdk_group.c:657-1351).
//
// Semantics (gdk_group.c:21-72): candidate rows are partitioned by equal
// (prior group id, value); group ids are handed out in order of FIRST OCCURRENCE
// (GRPnotfound, :74-100), extents = oid of each group's first row (ascending),
// histo = rows per group.  nil is an ordinary value (flt/dbl: nil == nil).
//
// The reference walks the rows serially and numbers groups as it meets them.
// Here the numbering is derived without any serial step:
//   A  one streaming pass records, per distinct key, the smallest candidate
//      position holding it ("first row"), in a DIRECT table when the key domain
//      is small (1-byte values, optionally x <= 256 prior groups; 2-byte values —
//      the cases the reference serves with GRP_small_values, :607-654,921-939 — and,
//      beyond the reference, 4- / 8-byte integers whose values span a small RANGE,
//      found by one min/max pass: entry = value - min)
//      or in an open-addressing HASH table whose slots hold that first row
//      itself (keys are compared by reading the column at the slot's row, as the
//      reference's link-array hash does, so any key width works lock-free);
//   B  the first rows are marked in a bitmap over candidate positions and a
//      popcount prefix over the bitmap turns "first row" into "rank among first
//      rows" = the reference's group id; extents fall out in ascending order;
//   C  a second streaming pass maps every row to its group id, builds the
//      histogram (privatised in shared memory for few groups) and checks
//      whether the ids are non-decreasing (gn->tsorted).
#include "common.cuh"
#include "scan.cuh"
#include "keys.cuh"

#define GRP_EMPTY (~0ULL)
#define GRP_THREADS 256
#define GRP_SHM_DOMAIN 4096	/* direct tables up to this many entries are privatised per CTA */
#define GRP_SHM_HISTO 2048
#define GRP_RANGED_MIN_ROWS ((uint64_t) 1 << 18)	/* below this the hash path is cheap and the extra range pass + sync is not */
#define GRP_RANGED_MAX_ENTRIES ((u64) 1 << 20)	/* largest ranged direct table (8 MiB of first rows) */

template <typename T> struct GrpSrc {
	const T *vals;		/* dense: value of candidate position r is vals[r]; else vals[cand[r] - hseq] */
	const oid *cand;	/* materialised candidates or NULL */
	oid hseq;
	oid seq;		/* dense: oid of candidate position 0 */
	const oid *g;		/* prior group ids per candidate position, or NULL */
	const unsigned char *g8;	/* their one-byte twin (SHADOW_GID8) when the prior BATgroup left one */
	uint64_t n;
	/* RANGED direct table (4- and 8-byte integers whose values span a small range): entry = value - lo, nil -> span - 1 */
	unsigned span;		/* 0: not ranged */
	T lo;
	__device__ __forceinline__ oid gval(uint64_t r) const { return g8 ? (oid) g8[r] : g[r]; }
	__device__ __forceinline__ T value(uint64_t r) const { return cand ? vals[cand[r] - hseq] : vals[r]; }
	__device__ __forceinline__ oid rowoid(uint64_t r) const { return cand ? cand[r] : seq + r; }
};

// ---------------------------------------------------------------- pass A, direct table
// entry index = value | prior gid << (8 * sizeof(T))   (the GV expressions at gdk_group.c:921-939)
template <typename T, bool HASG> __device__ __forceinline__ unsigned direct_index(const GrpSrc<T> &src, uint64_t r)
{
	typedef typename KeyBits<T>::type U;
	if constexpr (sizeof(T) >= 4) {
		if constexpr (Atom<T>::is_float || sizeof(T) > 8) {
			return 0;	/* never direct */
		} else {
			const T v = src.value(r);
			unsigned k = Atom<T>::is_nil(v) ? src.span - 1 : (unsigned) ((u64) v - (u64) src.lo);
			if (HASG)
				k += src.span * (unsigned) src.gval(r);
			return k;
		}
	} else {
		unsigned k = (unsigned) (U) src.value(r);
		if (HASG)
			k |= (unsigned) src.gval(r) << (8 * sizeof(T));
		return k;
	}
}

// smallest / largest non-nil value of the candidates as order-preserving unsigned images; out[2] = a nil was seen
template <typename T>
__global__ void __launch_bounds__(GRP_THREADS)
grp_minmax_kernel(GrpSrc<T> src, unsigned long long *__restrict__ out)
{
	u64 mn = ~0ULL, mx = 0;
	bool nil = false;
	const uint64_t stride = (uint64_t) gridDim.x * GRP_THREADS;
	constexpr int U8 = 8;	/* independent loads in flight per thread */
	uint64_t rstart = (uint64_t) blockIdx.x * GRP_THREADS + threadIdx.x;
	if (src.cand == nullptr && ((uintptr_t) src.vals & 15) == 0) {
		/* no candidate list: 16-byte loads (4-byte loads stream at well under half of what 16-byte ones reach);
		 * the rows behind the last whole vector go through the scalar loop below */
		constexpr int VEC = 16 / sizeof(T), UV = 4;
		const uint64_t nvec = src.n / VEC;
		const uint4 *vp = reinterpret_cast<const uint4 *>(src.vals);
		for (uint64_t i0 = rstart; i0 < nvec; i0 += stride * UV) {
			uint4 w[UV];
#pragma unroll
			for (int u = 0; u < UV; u++)
				w[u] = i0 + u * stride < nvec ? __ldcs(vp + i0 + u * stride) : make_uint4(0, 0, 0, 0);
#pragma unroll
			for (int u = 0; u < UV; u++) {
				if (i0 + u * stride >= nvec)
					continue;
				T e[VEC];
				memcpy(e, &w[u], 16);
#pragma unroll
				for (int j = 0; j < VEC; j++) {
					if (Atom<T>::is_nil(e[j])) {
						nil = true;
						continue;
					}
					const u64 img = (u64) (long long) e[j] ^ (1ULL << 63);
					mn = img < mn ? img : mn;
					mx = img > mx ? img : mx;
				}
			}
		}
		rstart += nvec * VEC;
	}
	for (uint64_t r0 = rstart; r0 < src.n; r0 += stride * U8) {
		T v[U8];
#pragma unroll
		for (int u = 0; u < U8; u++) {
			const uint64_t r = r0 + u * stride;
			v[u] = r < src.n ? src.value(r) : Atom<T>::nil();
		}
		bool tail = false;
#pragma unroll
		for (int u = 0; u < U8; u++) {
			if (Atom<T>::is_nil(v[u])) {
				tail |= r0 + u * stride < src.n;
				continue;
			}
			const u64 img = (u64) (long long) v[u] ^ (1ULL << 63);
			mn = img < mn ? img : mn;
			mx = img > mx ? img : mx;
		}
		nil |= tail;
	}
	for (int o = 16; o > 0; o >>= 1) {
		const u64 a = __shfl_down_sync(0xffffffffu, mn, o), b = __shfl_down_sync(0xffffffffu, mx, o);
		mn = a < mn ? a : mn;
		mx = b > mx ? b : mx;
	}
	if ((threadIdx.x & 31) == 0) {
		if (mn != ~0ULL) {
			atomicMin(&out[0], mn);
			atomicMax(&out[1], mx);
		}
	}
	if (nil)
		out[2] = 1;
}

template <typename T, bool HASG, bool SHM>
__global__ void __launch_bounds__(GRP_THREADS)
grp_direct_first_kernel(GrpSrc<T> src, u64 *__restrict__ first, unsigned domain)
{
	extern __shared__ u64 sfirst[];
	if (SHM) {
		for (unsigned i = threadIdx.x; i < domain; i += GRP_THREADS)
			sfirst[i] = GRP_EMPTY;
		__syncthreads();
	}
	const uint64_t stride = (uint64_t) gridDim.x * GRP_THREADS;
	constexpr int U4 = 4;
	uint64_t rstart = (uint64_t) blockIdx.x * GRP_THREADS + threadIdx.x;
	if constexpr (!HASG && !Atom<T>::is_float && (sizeof(T) == 4 || sizeof(T) == 8)) {
		if (src.cand == nullptr && src.span != 0 && ((uintptr_t) src.vals & 15) == 0) {
			/* ranged keys without candidates or prior groups: 16-byte loads, a thread's rows still ascend */
			constexpr int VEC = 16 / sizeof(T), UV = 2;
			const uint64_t nvec = src.n / VEC;
			const uint4 *vp = reinterpret_cast<const uint4 *>(src.vals);
			for (uint64_t i0 = rstart; i0 < nvec; i0 += stride * UV) {
				uint4 w[UV];
#pragma unroll
				for (int u = 0; u < UV; u++)
					w[u] = i0 + u * stride < nvec ? __ldcs(vp + i0 + u * stride) : make_uint4(0, 0, 0, 0);
#pragma unroll
				for (int u = 0; u < UV; u++) {
					if (i0 + u * stride >= nvec)
						continue;
					T e[VEC];
					memcpy(e, &w[u], 16);
#pragma unroll
					for (int j = 0; j < VEC; j++) {
						const uint64_t r = (i0 + u * stride) * VEC + j;
						const unsigned k = Atom<T>::is_nil(e[j]) ? src.span - 1 : (unsigned) ((u64) e[j] - (u64) src.lo);
						if (SHM) {
							if (r < sfirst[k])
								atomicMin(&sfirst[k], (u64) r);
						} else {
							if (r < __ldcg(first + k))
								atomicMin(first + k, (u64) r);
						}
					}
				}
			}
			rstart += nvec * VEC;
		}
	}
	for (uint64_t r0 = rstart; r0 < src.n; r0 += stride * U4) {
		unsigned k[U4];
#pragma unroll
		for (int u = 0; u < U4; u++) {
			const uint64_t r = r0 + u * stride;
			k[u] = r < src.n ? direct_index<T, HASG>(src, r) : 0xffffffffu;
		}
#pragma unroll
		for (int u = 0; u < U4; u++) {
			const uint64_t r = r0 + u * stride;
			if (k[u] == 0xffffffffu)
				continue;
			if (SHM) {
				/* rows arrive in ascending order per thread: after the first few
				 * iterations the plain read filters out every atomic */
				if (r < sfirst[k[u]])
					atomicMin(&sfirst[k[u]], (u64) r);
			} else {
				if (r < __ldcg(first + k[u]))
					atomicMin(first + k[u], (u64) r);
			}
		}
	}
	if (SHM) {
		__syncthreads();
		for (unsigned i = threadIdx.x; i < domain; i += GRP_THREADS) {
			u64 v = sfirst[i];
			if (v != GRP_EMPTY && v < __ldcg(first + i))
				atomicMin(first + i, v);
		}
	}
}

// 1-byte values without a candidate list: 16 rows per thread from one 16-byte load (and eight
// 16-byte loads of the prior group ids); the table lives in shared memory.
template <bool HASG>
__global__ void __launch_bounds__(GRP_THREADS)
grp_byte_first_kernel(const unsigned char *__restrict__ vals, const unsigned char *__restrict__ g8, uint64_t n, u64 *__restrict__ first, unsigned domain)
{
	extern __shared__ u64 sfirst[];
	for (unsigned i = threadIdx.x; i < domain; i += GRP_THREADS)
		sfirst[i] = GRP_EMPTY;
	__syncthreads();
	const uint64_t nvec = n / 16;
	for (uint64_t v = (uint64_t) blockIdx.x * GRP_THREADS + threadIdx.x; v < nvec; v += (uint64_t) gridDim.x * GRP_THREADS) {
		const uint4 w = __ldcs(reinterpret_cast<const uint4 *>(vals) + v);
		const unsigned ww[4] = {w.x, w.y, w.z, w.w};
		unsigned gw[4] = {0, 0, 0, 0};
		if (HASG) {
			const uint4 q = __ldcs(reinterpret_cast<const uint4 *>(g8) + v);
			gw[0] = q.x;
			gw[1] = q.y;
			gw[2] = q.z;
			gw[3] = q.w;
		}
#pragma unroll
		for (int i = 0; i < 16; i++) {
			unsigned k = (ww[i >> 2] >> (8 * (i & 3))) & 0xff;
			if (HASG)
				k |= ((gw[i >> 2] >> (8 * (i & 3))) & 0xff) << 8;
			const u64 r = v * 16 + i;
			if (r < sfirst[k])
				atomicMin(&sfirst[k], r);
		}
	}
	/* tail rows */
	const uint64_t r = nvec * 16 + (uint64_t) blockIdx.x * GRP_THREADS + threadIdx.x;
	if (r < n) {
		unsigned k = vals[r];
		if (HASG)
			k |= (unsigned) g8[r] << 8;
		atomicMin(&sfirst[k], (u64) r);
	}
	__syncthreads();
	for (unsigned i = threadIdx.x; i < domain; i += GRP_THREADS) {
		u64 v = sfirst[i];
		if (v != GRP_EMPTY && v < __ldcg(first + i))
			atomicMin(first + i, v);
	}
}

// ---------------------------------------------------------------- pass A, hash table
// slots[h] = smallest candidate position seen so far of the key that owns slot h.
// A slot's value only ever changes to another row with the SAME key, so comparing
// against the key of whatever row is read from the slot is always valid.
template <typename T, bool HASG>
__global__ void __launch_bounds__(GRP_THREADS)
grp_hash_first_kernel(GrpSrc<T> src, u64 *__restrict__ slots, u64 mask, uint32_t *__restrict__ slotof,
		      unsigned *__restrict__ state /* [0] overflow flag, [1] inserted keys */)
{
	typedef typename KeyBits<T>::type U;
	__shared__ unsigned sins;
	if (threadIdx.x == 0)
		sins = 0;
	__syncthreads();
	unsigned inserted = 0;
	const uint64_t stride = (uint64_t) gridDim.x * GRP_THREADS;
	/* keys of up to 4 bytes: a direct-mapped shared-memory cache key -> global slot (one 64-bit word per entry,
	 * so an entry is always consistent).  A CTA meets its rows in ascending order, so once it has settled a key
	 * in the global table every later row with that key only needs the slot number: a hit costs one
	 * shared-memory load instead of two scattered global loads (the kernel was L1TEX-bound on those: 69 % of
	 * the L1 pipe, 70 % excess sectors). */
	constexpr bool CACHED = !HASG && sizeof(U) <= 4;
	constexpr unsigned CSZ = 2048;
	__shared__ unsigned long long kcache[CACHED ? CSZ : 1];
	if constexpr (CACHED) {
		for (unsigned i = threadIdx.x; i < CSZ; i += GRP_THREADS)
			kcache[i] = 0;
		__syncthreads();
	}
	constexpr int B = 4;		/* rows in flight per thread */
	/* the trip count is uniform over the CTA (barriers inside): rows of iteration i+1 are all larger than the
	 * rows of iteration i of this CTA */
	for (uint64_t rb = (uint64_t) blockIdx.x * GRP_THREADS; rb < src.n; rb += stride * B) {
		const uint64_t r0 = rb + threadIdx.x;
		/* The common case of a low-cardinality key — the slot at the key's home position already holds an
		 * EARLIER row with the same key — needs three dependent loads (key, slot, key of the slot's row) and
		 * no atomic.  One row at a time that chain left the kernel latency-bound (37 % + 27 % of the stall
		 * samples on the two loads, profiles/r2_cfg1_full.md); B rows' chains are issued together here and
		 * only rows that miss this fast test enter the probing loop. */
		U v[B];
		u64 gv[B], h[B], cur[B];
		bool live[B], done[B];
#pragma unroll
		for (int j = 0; j < B; j++) {
			const uint64_t r = r0 + (uint64_t) j * stride;
			live[j] = r < src.n;
			v[j] = live[j] ? KeyBits<T>::get(src.value(r)) : (U) 0;
			gv[j] = (HASG && live[j]) ? src.gval(r) : 0;
			h[j] = hash_key<U>(v[j], gv[j]) & mask;
		}
		if constexpr (CACHED) {
#pragma unroll
			for (int j = 0; j < B; j++) {
				if (!live[j])
					continue;
				const unsigned long long e = kcache[(unsigned) (h[j] >> 7) & (CSZ - 1)];
				if ((e >> 32) != 0 && (uint32_t) e == (uint32_t) v[j]) {
					slotof[r0 + (uint64_t) j * stride] = (uint32_t) (e >> 32) - 1;
					live[j] = false;	/* settled from the cache */
				}
			}
			/* entries read above were installed by EARLIER iterations, i.e. by smaller rows whose
			 * global update is done; this iteration's installs must not be seen by this iteration's
			 * lookups (a smaller row of the same iteration would skip its atomicMin) */
			__syncthreads();
		}
#pragma unroll
		for (int j = 0; j < B; j++)
			cur[j] = live[j] ? __ldcg(slots + h[j]) : GRP_EMPTY;
#pragma unroll
		for (int j = 0; j < B; j++) {
			const uint64_t r = r0 + (uint64_t) j * stride;
			done[j] = false;
			if (live[j] && cur[j] != GRP_EMPTY && cur[j] <= r)
				done[j] = cur[j] == r || (KeyBits<T>::get(src.value(cur[j])) == v[j] && (!HASG || src.gval(cur[j]) == gv[j]));
		}
#pragma unroll
		for (int j = 0; j < B; j++) {
			const uint64_t r = r0 + (uint64_t) j * stride;
			if (!live[j])
				continue;
			u64 hh = h[j];
			if (!done[j]) {
				unsigned probes = 0;
				for (;;) {
					u64 c = __ldcg(slots + hh);
					if (c == GRP_EMPTY) {
						c = atomicCAS(slots + hh, GRP_EMPTY, (u64) r);
						if (c == GRP_EMPTY) {
							inserted++;
							break;
						}
					}
					if (c == r)
						break;
					if (KeyBits<T>::get(src.value(c)) == v[j] && (!HASG || src.gval(c) == gv[j])) {
						if (r < c)
							atomicMin(slots + hh, (u64) r);
						break;
					}
					hh = (hh + 1) & mask;
					if (++probes > 8192) {
						state[0] = 1;	/* table too full: the host retries with a larger one */
						break;
					}
				}
			}
			slotof[r] = (uint32_t) hh;
			if constexpr (CACHED)
				kcache[(unsigned) (h[j] >> 7) & (CSZ - 1)] = (unsigned long long) (uint32_t) v[j] | ((unsigned long long) ((uint32_t) hh + 1) << 32);
		}
		if constexpr (CACHED)
			__syncthreads();
	}
	if (inserted)
		atomicAdd(&sins, inserted);
	__syncthreads();
	if (threadIdx.x == 0 && sins)
		atomicAdd(&state[1], sins);
}

// ---------------------------------------------------------------- pass B: first rows -> ranks
__global__ void grp_mark_kernel(const u64 *__restrict__ table, uint64_t nentries, u64 *__restrict__ bits)
{
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < nentries; i += (uint64_t) gridDim.x * blockDim.x) {
		u64 r = __ldcs(table + i);
		if (r != GRP_EMPTY)
			atomicOr(bits + (r >> 6), 1ULL << (r & 63));
	}
}

// exclusive popcount prefix per 64-bit word; total -> *total_out.  Tile = 256 x 8 words.
#define BP_ITEMS 8
__global__ void __launch_bounds__(GRP_THREADS)
grp_bitprefix_kernel(const u64 *__restrict__ bits, uint64_t nwords, u64 *__restrict__ pref, TileScan ts,
		     unsigned ntiles, unsigned long long *total_out)
{
	__shared__ unsigned wtot[GRP_THREADS / 32];
	__shared__ unsigned long long slot;
	__shared__ unsigned tslot;
	const unsigned tile = tile_ticket(ts, &tslot);
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint64_t base = (uint64_t) tile * (GRP_THREADS * BP_ITEMS) + (uint64_t) threadIdx.x * BP_ITEMS;
	unsigned c[BP_ITEMS], tsum = 0;
#pragma unroll
	for (int j = 0; j < BP_ITEMS; j++) {
		c[j] = base + j < nwords ? __popcll(bits[base + j]) : 0;
		tsum += c[j];
	}
	unsigned inc = tsum;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
		if (lane >= o)
			inc += t;
	}
	if (lane == 31)
		wtot[warp] = inc;
	__syncthreads();
	unsigned wexcl = 0, ttot = 0;
#pragma unroll
	for (int w = 0; w < GRP_THREADS / 32; w++) {
		unsigned t = wtot[w];
		wexcl += w < warp ? t : 0;
		ttot += t;
	}
	const unsigned long long texcl = tile_exclusive_prefix(ts, tile, ttot, &slot);
	unsigned long long run = texcl + wexcl + inc - tsum;
#pragma unroll
	for (int j = 0; j < BP_ITEMS; j++) {
		if (base + j < nwords)
			pref[base + j] = run;
		run += c[j];
	}
	if (tile == ntiles - 1 && threadIdx.x == 0)
		*total_out = texcl + ttot;
}

// table entry (first row) -> group id; extents[gid] = oid of that row
template <typename T>
__global__ void grp_rank_kernel(const u64 *__restrict__ table, uint64_t nentries, const u64 *__restrict__ bits,
				const u64 *__restrict__ pref, GrpSrc<T> src, uint32_t *__restrict__ gidtab,
				oid *__restrict__ exts)
{
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < nentries; i += (uint64_t) gridDim.x * blockDim.x) {
		u64 r = __ldcs(table + i);
		if (r == GRP_EMPTY)
			continue;
		u64 w = bits[r >> 6];
		u64 gid = pref[r >> 6] + __popcll(w & ((1ULL << (r & 63)) - 1));
		gidtab[i] = (uint32_t) gid;
		if (exts)
			exts[gid] = src.rowoid(r);
	}
}

// ---------------------------------------------------------------- pass C: rows -> group ids
// MODE 0: direct table, gidtab copied to shared memory (domain <= GRP_SHM_DOMAIN)
// MODE 1: direct table in global memory
// MODE 2: hash table, entry of row r saved in slotof[r] by pass A
template <typename T, bool HASG, int MODE>
__device__ __forceinline__ uint32_t grp_entry(const GrpSrc<T> &src, const uint32_t *slotof, uint64_t r)
{
	if constexpr (MODE == 2)
		return __ldcs(slotof + r);
	else
		return direct_index<T, HASG>(src, r);
}

template <typename T, bool HASG, int MODE, bool SHISTO>
__global__ void __launch_bounds__(GRP_THREADS)
grp_assign_kernel(GrpSrc<T> src, const uint32_t *__restrict__ slotof, const uint32_t *__restrict__ gidtab, unsigned domain,
		  oid *__restrict__ gids, unsigned char *__restrict__ g8out, u64 *__restrict__ cnts, unsigned ngrp,
		  unsigned *__restrict__ unsorted)
{
	extern __shared__ uint32_t sh[];
	uint32_t *sgid = sh;					/* MODE 0: domain entries */
	uint32_t *shist = sh + (MODE == 0 ? domain : 0);	/* SHISTO: ngrp counters */
	if (MODE == 0)
		for (unsigned i = threadIdx.x; i < domain; i += GRP_THREADS)
			sgid[i] = gidtab[i];
	if (SHISTO)
		for (unsigned i = threadIdx.x; i < ngrp; i += GRP_THREADS)
			shist[i] = 0;
	if (MODE == 0 || SHISTO)
		__syncthreads();
	bool bad = false;
	/* a thread handles two consecutive rows per step so the id stores are 16 bytes wide */
	const uint64_t npairs = (src.n + 1) / 2;
	const uint64_t stride = (uint64_t) gridDim.x * GRP_THREADS;
	for (uint64_t p = (uint64_t) blockIdx.x * GRP_THREADS + threadIdx.x; p < npairs; p += stride) {
		const uint64_t r = 2 * p;
		const bool two = r + 1 < src.n;
		uint32_t e0 = grp_entry<T, HASG, MODE>(src, slotof, r);
		uint32_t e1 = two ? grp_entry<T, HASG, MODE>(src, slotof, r + 1) : e0;
		uint32_t ep = r ? grp_entry<T, HASG, MODE>(src, slotof, r - 1) : e0;
		uint32_t g0, g1, gp;
		if (MODE == 0) {
			g0 = sgid[e0];
			g1 = sgid[e1];
			gp = sgid[ep];
		} else {
			g0 = __ldg(gidtab + e0);
			g1 = __ldg(gidtab + e1);
			gp = __ldg(gidtab + ep);
		}
		bad |= (r && g0 < gp) || (two && g1 < g0);
		if (two) {
			ulonglong2 v;
			v.x = g0;
			v.y = g1;
			__stcs(reinterpret_cast<ulonglong2 *>(gids + r), v);
			if (g8out)
				*reinterpret_cast<unsigned short *>(g8out + r) = (unsigned short) (g0 | (g1 << 8));
		} else {
			gids[r] = g0;
			if (g8out)
				g8out[r] = (unsigned char) g0;
		}
		if (cnts) {
			if (SHISTO && ngrp <= 32) {
				/* few groups: warp-aggregated, one shared atomic per distinct id per warp */
				const unsigned act = __activemask();
				unsigned m0 = __match_any_sync(act, g0);
				if ((unsigned) (__ffs(m0) - 1) == (threadIdx.x & 31))
					atomicAdd(&shist[g0], (unsigned) __popc(m0));
				unsigned m1 = __match_any_sync(act, two ? g1 : 0xffffffffu);
				if (two && (unsigned) (__ffs(m1) - 1) == (threadIdx.x & 31))
					atomicAdd(&shist[g1], (unsigned) __popc(m1));
			} else if (SHISTO) {
				/* many groups: lanes rarely collide, and MATCH.ANY cost a third of this kernel's issue slots
				 * at 500 groups (profiles/r2_cfg1_full.md) — plain shared-memory atomics */
				atomicAdd(&shist[g0], 1u);
				if (two)
					atomicAdd(&shist[g1], 1u);
			} else {
				atomicAdd(cnts + g0, 1ULL);
				if (two)
					atomicAdd(cnts + g1, 1ULL);
			}
		}
	}
	if (__syncthreads_or(bad) && threadIdx.x == 0)
		*unsorted = 1;
	if (SHISTO && cnts) {
		for (unsigned i = threadIdx.x; i < ngrp; i += GRP_THREADS)
			if (shist[i])
				atomicAdd(cnts + i, (u64) shist[i]);
	}
}

// 1-byte values without a candidate list, table in shared memory, at most 8 groups: a warp takes
// 512 consecutive rows per step; lane l handles rows 2l, 2l+1 of each of eight 64-row slices, so
// the value loads (2 bytes), prior-id loads and id stores (16 bytes) are all fully coalesced.  The
// histogram is a packed 8 x 8-bit counter per thread, flushed every 8 steps (<= 128 rows).
template <bool HASG>
__global__ void __launch_bounds__(GRP_THREADS)
grp_byte_assign_kernel(const unsigned char *__restrict__ vals, const unsigned char *__restrict__ g8, uint64_t n, const uint32_t *__restrict__ gidtab,
		       unsigned domain, oid *__restrict__ gids, unsigned char *__restrict__ g8out, u64 *__restrict__ cnts, unsigned ngrp,
		       unsigned *__restrict__ unsorted)
{
	extern __shared__ uint32_t sh[];
	uint32_t *sgid = sh, *shist = sh + domain;
	for (unsigned i = threadIdx.x; i < domain; i += GRP_THREADS)
		sgid[i] = gidtab[i];
	if (threadIdx.x < 8)
		shist[threadIdx.x] = 0;
	__syncthreads();
	const unsigned lane = threadIdx.x & 31;
	const uint64_t warp = ((uint64_t) blockIdx.x * GRP_THREADS + threadIdx.x) >> 5;
	const uint64_t nwarps = ((uint64_t) gridDim.x * GRP_THREADS) >> 5;
	bool bad = false;
	u64 packed = 0;
	unsigned steps = 0;
	for (uint64_t base = warp * 512; base < n; base += nwarps * 512) {
		unsigned short v[8], pg[HASG ? 8 : 1];
#pragma unroll
		for (int j = 0; j < 8; j++) {
			const uint64_t r = base + j * 64 + 2 * lane;
			v[j] = 0;
			if (r + 1 < n)
				v[j] = __ldcs(reinterpret_cast<const unsigned short *>(vals + r));
			else if (r < n)
				v[j] = vals[r];
			if (HASG) {
				pg[j] = 0;
				if (r + 1 < n)
					pg[j] = __ldcs(reinterpret_cast<const unsigned short *>(g8 + r));
				else if (r < n)
					pg[j] = g8[r];
			}
		}
		/* id of the row just before this warp's 512 rows (sortedness check) */
		uint32_t carry = 0;
		if (lane == 0 && base > 0) {
			unsigned k = vals[base - 1];
			if (HASG)
				k |= (unsigned) g8[base - 1] << 8;
			carry = sgid[k];
		}
#pragma unroll
		for (int j = 0; j < 8; j++) {
			const uint64_t r = base + j * 64 + 2 * lane;
			unsigned k0 = v[j] & 0xff, k1 = v[j] >> 8;
			if (HASG) {
				k0 |= (unsigned) (pg[j] & 0xff) << 8;
				k1 |= (unsigned) (pg[j] >> 8) << 8;
			}
			const bool in0 = r < n, in1 = r + 1 < n;
			const uint32_t g0 = in0 ? sgid[k0] : 0, g1 = in1 ? sgid[k1] : g0;
			uint32_t gp = __shfl_up_sync(0xffffffffu, g1, 1);
			if (lane == 0)
				gp = carry;
			carry = __shfl_sync(0xffffffffu, g1, 31);
			bad |= (in0 && r > 0 && g0 < gp) || (in1 && g1 < g0);
			if (in1) {
				ulonglong2 o;
				o.x = g0;
				o.y = g1;
				__stcs(reinterpret_cast<ulonglong2 *>(gids + r), o);
				if (g8out)
					*reinterpret_cast<unsigned short *>(g8out + r) = (unsigned short) (g0 | (g1 << 8));
			} else if (in0) {
				gids[r] = g0;
				if (g8out)
					g8out[r] = (unsigned char) g0;
			}
			packed += (in0 ? 1ULL << (8 * g0) : 0) + (in1 ? 1ULL << (8 * g1) : 0);
		}
		if (++steps == 8) {
#pragma unroll
			for (int q = 0; q < 8; q++)
				if ((packed >> (8 * q)) & 0xff)
					atomicAdd(&shist[q], (unsigned) ((packed >> (8 * q)) & 0xff));
			packed = 0;
			steps = 0;
		}
	}
#pragma unroll
	for (int q = 0; q < 8; q++)
		if ((packed >> (8 * q)) & 0xff)
			atomicAdd(&shist[q], (unsigned) ((packed >> (8 * q)) & 0xff));
	if (__syncthreads_or(bad) && threadIdx.x == 0)
		*unsorted = 1;
	if (cnts && threadIdx.x < ngrp && shist[threadIdx.x])
		atomicAdd(cnts + threadIdx.x, (u64) shist[threadIdx.x]);
}

// ---------------------------------------------------------------- host side
__global__ void grp_max_kernel(const oid *__restrict__ g, uint64_t n, u64 *__restrict__ out)
{
	u64 mx = 0;
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
		oid v = __ldcs(g + i);
		if (v != OID_NIL && v > mx)
			mx = v;
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		u64 t = __shfl_down_sync(0xffffffffu, mx, o);
		mx = t > mx ? t : mx;
	}
	if ((threadIdx.x & 31) == 0 && mx)
		atomicMax(out, mx);
}

__global__ void grp_finish_kernel(const oid *exts, u64 ngrp, const unsigned *unsorted, unsigned long long *slot)
{
	slot[0] = *unsorted;
	slot[1] = exts && ngrp ? exts[0] : 0;
	slot[2] = exts && ngrp ? exts[ngrp - 1] : 0;
}

struct GrpTemps {
	u64 *table = nullptr, *bits = nullptr, *pref = nullptr;
	uint32_t *slotof = nullptr, *gidtab = nullptr;
	unsigned *state = nullptr;
	~GrpTemps()
	{
		dev_free(table);
		dev_free(bits);
		dev_free(pref);
		dev_free(slotof);
		dev_free(gidtab);
		dev_free(state);
	}
};

static inline u64 next_pow2(u64 x)
{
	u64 p = 1;
	while (p < x)
		p <<= 1;
	return p;
}

template <typename T, bool HASG>
static bool group_run(GrpSrc<T> src, u64 maxgrp, b200_bat *gn, b200_bat **enp, b200_bat **hnp, bool want_e, bool want_h)
{
	Ctx *c = ctx();
	const uint64_t n = src.n;
	const unsigned sgrid = grid_for(n, GRP_THREADS * 4, NSM * 8);
	constexpr bool DIRECT_OK = !Atom<T>::is_float && sizeof(T) <= 8;
	GrpTemps t;
	int mode;		/* 0 direct/shared, 1 direct/global, 2 hash */
	u64 nentries;
	src.span = 0;
	src.lo = 0;
	if (sizeof(T) == 1 && (!HASG || maxgrp < 256)) {
		nentries = HASG ? 256 * (maxgrp + 1) : 256;
		mode = nentries <= GRP_SHM_DOMAIN ? 0 : 1;
	} else if (sizeof(T) == 2 && !HASG) {
		nentries = 65536;
		mode = 1;
	} else {
		mode = 2;
		nentries = 0;
	}
	if constexpr (DIRECT_OK && sizeof(T) >= 4) {
		/* wide integers that only take a small range of values (dimension keys, codes, dates of a few years): one
		 * pass finds the range; when it — times the prior groups — fits a direct table, no hashing is needed */
		if (n >= GRP_RANGED_MIN_ROWS && (!HASG || maxgrp < GRP_RANGED_MAX_ENTRIES)) {
			unsigned long long *mm = (unsigned long long *) scratch_get(64);
			static const unsigned long long init[3] = {~0ULL, 0, 0};
			if (mm == nullptr || !cuda_ok(cudaMemcpyAsync(mm, init, 24, cudaMemcpyHostToDevice, c->stream), "init minmax"))
				return false;
			{
				KTimer kt("group_minmax", n * sizeof(T));
				grp_minmax_kernel<T><<<grid_for(n, GRP_THREADS * 8, NSM * 8), GRP_THREADS, 0, c->stream>>>(src, mm);
			}
			COUNT_LAUNCH();
			if (!cuda_ok(cudaMemcpyAsync(c->slot_h->v, mm, 24, cudaMemcpyDeviceToHost, c->stream), "read minmax") || !stream_sync())
				return false;
			const u64 imn = c->slot_h->v[0], imx = c->slot_h->v[1];
			const u64 range = imn == ~0ULL ? 1 : imx - imn + 1;	/* images are order-preserving: their distance is the values' */
			if (range != 0 && range < GRP_RANGED_MAX_ENTRIES) {
				const u64 span = range + 1;	/* + the nil entry */
				const u64 ent = span * (HASG ? maxgrp + 1 : 1);
				if (ent <= GRP_RANGED_MAX_ENTRIES) {
					src.span = (unsigned) span;
					src.lo = imn == ~0ULL ? (T) 0 : (T) (long long) (imn ^ (1ULL << 63));
					nentries = ent;
					mode = nentries <= GRP_SHM_DOMAIN ? 0 : 1;
				}
			}
		}
	}
	t.state = (unsigned *) dev_alloc(16);
	if (t.state == nullptr)
		return false;
	if (mode != 2) {
		t.table = (u64 *) dev_alloc(nentries * 8);
		if (t.table == nullptr || !cuda_ok(cudaMemsetAsync(t.table, 0xff, nentries * 8, c->stream), "memset group table"))
			return false;
		if constexpr (DIRECT_OK) {
			KTimer kt(HASG ? "group_first_g" : "group_first", n * (sizeof(T) + (HASG ? 8 : 0)));
			bool fastb = false;
			if constexpr (sizeof(T) == 1)
				fastb = mode == 0 && src.cand == nullptr && ((uintptr_t) src.vals & 15) == 0 &&
					(!HASG || (src.g8 != nullptr && ((uintptr_t) src.g8 & 15) == 0));	/* 8-byte prior ids: the row-per-lane form is already coalesced */
			if (fastb)
				grp_byte_first_kernel<HASG><<<grid_for(n / 16 + GRP_THREADS, GRP_THREADS, NSM * 8), GRP_THREADS, nentries * 8, c->stream>>>(
					(const unsigned char *) src.vals, src.g8, n, t.table, (unsigned) nentries);
			else if (mode == 0)
				grp_direct_first_kernel<T, HASG, true><<<sgrid, GRP_THREADS, nentries * 8, c->stream>>>(src, t.table, (unsigned) nentries);
			else
				grp_direct_first_kernel<T, HASG, false><<<sgrid, GRP_THREADS, 0, c->stream>>>(src, t.table, (unsigned) nentries);
		}
		COUNT_LAUNCH();
		ctx()->algo = mode == 0 ? "group: cuda direct table (shared)" : "group: cuda direct table";
	} else {
		t.slotof = (uint32_t *) dev_alloc(n * 4);
		if (t.slotof == nullptr)
			return false;
		const u64 capmax = next_pow2(2 * n);
		u64 cap = next_pow2(n / 4);
		if (cap < (1u << 16))
			cap = 1u << 16;
		if (cap > (1u << 22))
			cap = 1u << 22;
		if (cap > capmax)
			cap = capmax;
		for (;;) {
			t.table = (u64 *) dev_alloc(cap * 8);
			if (t.table == nullptr || !cuda_ok(cudaMemsetAsync(t.table, 0xff, cap * 8, c->stream), "memset group hash") ||
			    !cuda_ok(cudaMemsetAsync(t.state, 0, 16, c->stream), "memset state"))
				return false;
			{
				KTimer kt("group_hash", n * (sizeof(T) + (HASG ? 8 : 0) + 4));
				grp_hash_first_kernel<T, HASG><<<grid_for(n, GRP_THREADS, NSM * 8), GRP_THREADS, 0, c->stream>>>(src, t.table, cap - 1, t.slotof, t.state);
			}
			COUNT_LAUNCH();
			if (!cuda_ok(cudaMemcpyAsync(c->slot_h->v, t.state, 8, cudaMemcpyDeviceToHost, c->stream), "read state") || !stream_sync())
				return false;
			const unsigned *st = (const unsigned *) c->slot_h->v;
			if (cap >= capmax || (st[0] == 0 && (u64) st[1] * 2 <= cap))
				break;
			/* too full for comfortable probing: retry with a table sized for the worst case seen */
			dev_free(t.table);
			t.table = nullptr;
			cap = cap * 16 > capmax ? capmax : cap * 16;
		}
		nentries = cap;
		ctx()->algo = "group: cuda hash table";
	}
	/* pass B */
	const u64 nwords = (n + 63) / 64;
	t.bits = (u64 *) dev_alloc(nwords * 8);
	t.pref = (u64 *) dev_alloc(nwords * 8);
	t.gidtab = (uint32_t *) dev_alloc(nentries * 4);
	if (t.bits == nullptr || t.pref == nullptr || t.gidtab == nullptr ||
	    !cuda_ok(cudaMemsetAsync(t.bits, 0, nwords * 8, c->stream), "memset bits") ||
	    !cuda_ok(cudaMemsetAsync(t.state, 0, 16, c->stream), "memset state"))
		return false;
	grp_mark_kernel<<<grid_for(nentries, 256, NSM * 8), 256, 0, c->stream>>>(t.table, nentries, t.bits);
	COUNT_LAUNCH();
	const u64 ntiles = (nwords + GRP_THREADS * BP_ITEMS - 1) / (GRP_THREADS * BP_ITEMS);
	char *sp = (char *) scratch_get(16 + tilescan_bytes(ntiles));
	if (sp == nullptr || !cuda_ok(cudaMemsetAsync(sp, 0, 16 + tilescan_bytes(ntiles), c->stream), "memset tile state"))
		return false;
	TileScan ts;
	ts.ticket = (unsigned int *) (sp + 16);
	ts.status = (unsigned long long *) (sp + 32);
	grp_bitprefix_kernel<<<(unsigned) ntiles, GRP_THREADS, 0, c->stream>>>(t.bits, nwords, t.pref, ts, (unsigned) ntiles, (unsigned long long *) sp);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaMemcpyAsync(c->slot_h->v, sp, 8, cudaMemcpyDeviceToHost, c->stream), "read ngrp") || !stream_sync())
		return false;
	const u64 ngrp = c->slot_h->v[0];
	b200_bat *en = nullptr, *hn = nullptr;
	if (want_e) {
		en = bat_new(0, B200_oid, ngrp);
		if (en == nullptr)
			return false;
		*enp = en;
	}
	if (want_h) {
		hn = bat_new(0, B200_lng, ngrp);
		if (hn == nullptr || !cuda_ok(cudaMemsetAsync(hn->base, 0, ngrp * 8, c->stream), "memset histo"))
			return false;
		*hnp = hn;
	}
	grp_rank_kernel<T><<<grid_for(nentries, 256, NSM * 8), 256, 0, c->stream>>>(t.table, nentries, t.bits, t.pref, src, t.gidtab,
										       en ? (oid *) en->base : nullptr);
	COUNT_LAUNCH();
	/* pass C */
	const bool shisto = ngrp <= GRP_SHM_HISTO;
	oid *gids = (oid *) gn->base;
	/* few groups: leave a one-byte twin of the ids for the aggregates / the next BATgroup level */
	unsigned char *g8out = shadow_enabled() && ngrp <= 256 && n >= (1u << 16) ? (unsigned char *) dev_alloc(n) : nullptr;
	u64 *cnts = hn ? (u64 *) hn->base : nullptr;
	const unsigned agrid = grid_for((n + 1) / 2, GRP_THREADS * 4, NSM * 8);
	const size_t shm = (mode == 0 ? nentries * 4 : 0) + (shisto ? ngrp * 4 : 0);
	{
		KTimer kt(HASG ? "group_assign_g" : "group_assign", n * ((mode == 2 ? 4 : sizeof(T) + (HASG ? 8 : 0)) + 8));
#define ASSIGN(MODE, SH) grp_assign_kernel<T, HASG, MODE, SH><<<agrid, GRP_THREADS, shm, c->stream>>>( \
		src, t.slotof, t.gidtab, (unsigned) nentries, gids, g8out, cnts, (unsigned) ngrp, t.state)
		bool fastb = false;
		if constexpr (sizeof(T) == 1)
			fastb = mode == 0 && ngrp <= 8 && src.cand == nullptr && ((uintptr_t) src.vals & 1) == 0 &&
				(!HASG || (src.g8 != nullptr && ((uintptr_t) src.g8 & 1) == 0));	/* 8-byte prior ids: measured no faster */
		if (fastb) {
			grp_byte_assign_kernel<HASG><<<grid_for(n, GRP_THREADS * 16, NSM * 8), GRP_THREADS, nentries * 4 + 32, c->stream>>>(
				(const unsigned char *) src.vals, src.g8, n, t.gidtab, (unsigned) nentries, gids, g8out, cnts, (unsigned) ngrp, t.state);
		} else if (mode == 2) {
			if (shisto) ASSIGN(2, true); else ASSIGN(2, false);
		} else if constexpr (DIRECT_OK) {
			if (mode == 0) {
				if (shisto) ASSIGN(0, true); else ASSIGN(0, false);
			} else {
				if (shisto) ASSIGN(1, true); else ASSIGN(1, false);
			}
		}
#undef ASSIGN
	}
	COUNT_LAUNCH();
	grp_finish_kernel<<<1, 1, 0, c->stream>>>(en ? (const oid *) en->base : nullptr, ngrp, t.state, c->slot_d->v);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "group launch") || !stream_sync()) {
		dev_free(g8out);
		return false;
	}
	if (g8out) {
		Shadow sh;
		sh.kind = SHADOW_GID8;
		sh.p = g8out;
		sh.n = n;
		sh.a0 = ngrp;
		shadow_put(gn->base, sh);
	}
	const bool unsorted = c->slot_h->v[0] != 0;
	const oid efirst = c->slot_h->v[1], elast = c->slot_h->v[2];
	/* result properties, gdk_group.c:1283-1315 */
	gn->count = n;
	gn->tsorted = !unsorted;
	gn->trevsorted = ngrp == 1 || n <= 1;
	gn->tkey = ngrp == n;
	gn->tnonil = 1;
	gn->tnil = 0;
	gn->tseqbase = OID_NIL;
	if (en) {
		en->count = ngrp;
		en->tkey = en->tsorted = en->tnonil = 1;
		en->trevsorted = ngrp == 1;
		en->tnil = 0;
		en->tseqbase = OID_NIL;
		en->oidrange_known = 1;
		en->first_oid = efirst;
		en->last_oid = elast;
		if (ngrp > 0 && elast - efirst + 1 == ngrp) {	/* virtualize (gdk_select.c:31-89) */
			b200_bat *d = bat_dense(0, efirst, ngrp);
			if (d == nullptr)
				return false;
			b200_BATfree(en);
			*enp = d;
		}
	}
	if (hn) {
		hn->count = ngrp;
		const bool triv = ngrp == n || ngrp == 1;
		hn->tkey = ngrp == 1;
		hn->tsorted = hn->trevsorted = triv;
		hn->tnonil = 1;
		hn->tnil = 0;
	}
	return true;
}

template <typename T>
static bool group_typed(const b200_bat *b, const Cand &ci, const b200_bat *g, u64 maxgrp, b200_bat *gn,
			b200_bat **enp, b200_bat **hnp, bool want_e, bool want_h)
{
	GrpSrc<T> src;
	src.hseq = ci.hseq;
	src.seq = ci.seq;
	src.n = ci.ncand;
	if (ci.dense) {
		src.vals = (const T *) b->base + (ci.seq - ci.hseq);
		src.cand = nullptr;
	} else {
		src.vals = (const T *) b->base;
		src.cand = ci.oids;
	}
	src.g = g ? (const oid *) g->base : nullptr;
	src.g8 = nullptr;
	Shadow sh;
	if (g && shadow_get(g->base, SHADOW_GID8, g->count, sh))
		src.g8 = (const unsigned char *) sh.p;
	if (g)
		return group_run<T, true>(src, maxgrp, gn, enp, hnp, want_e, want_h);
	return group_run<T, false>(src, maxgrp, gn, enp, hnp, want_e, want_h);
}

extern "C" b200_return b200_BATgroup(b200_bat **groups, b200_bat **extents, b200_bat **histo,
				     const b200_bat *b, const b200_bat *s, const b200_bat *g,
				     const b200_bat *e, const b200_bat *h)
{
	b200_clrerr();
	if (groups)
		*groups = nullptr;
	if (extents)
		*extents = nullptr;
	if (histo)
		*histo = nullptr;
	if (b == nullptr) {
		set_err("b must exist\n");
		return B200_FAIL;
	}
	if (groups == nullptr) {
		set_err("BATgroup: groups output required\n");
		return B200_FAIL;
	}
	Staged sb, ss, sg;
	if (!stage(b, sb) || !stage(s, ss) || !stage(g, sg))
		return B200_FAIL;
	b = sb.b;
	g = sg.b;
	b200_bat strview;
	if (b->ttype == B200_STR_TYPE) {
		/* a duplicate-free string heap is grouped by offset, as the reference does (gdk_group.c:897-919) */
		if (!str_as_offsets(b, strview, true, "BATgroup"))
			return B200_FAIL;
		b = &strview;
	}
	Cand ci;
	if (!cand_init(ci, b, ss.b))
		return B200_FAIL;
	if (g && g->count != ci.ncand) {
		set_err("BATgroup: g must be aligned with b and s\n");
		return B200_FAIL;
	}
	Ctx *c = ctx();
	const oid hseqb = ci.ncand == 0 ? 0 : ci.seq;
	b200_bat *gn = nullptr, *en = nullptr, *hn = nullptr;
	const bool bkey = b->tkey || (b->ttype == B200_void && b->tseqbase != OID_NIL);
	if (bkey || ci.ncand <= 1 || (g && (g->tkey || bat_is_dense(g)))) {
		/* grouping is trivial: 1 element per group (gdk_group.c:712-727) */
		c->algo = "group: 1 element per group";
		gn = bat_dense(hseqb, 0, b->count);
		if (gn == nullptr)
			goto bailout;
		if (extents && (en = cand_to_bat(ci)) == nullptr)
			goto bailout;
		if (histo) {
			const lng one = 1;
			if ((hn = bat_constant(0, B200_lng, &one, ci.ncand)) == nullptr)
				goto bailout;
		}
	} else {
		u64 maxgrp = 0;
		if (g) {
			if (g->ttype == B200_void) {	/* void-nil: every row in group nil; not a valid grouping */
				set_err("BATgroup: g must be a materialised oid column\n");
				goto bailout;
			}
			Shadow gsh;
			if (e != nullptr && e->count > 0) {
				maxgrp = e->count - 1;
			} else if (shadow_get(g->base, SHADOW_GID8, g->count, gsh) && gsh.a0 > 0) {
				maxgrp = gsh.a0 - 1;	/* the BATgroup call that made g left its group count */
			} else {
				u64 *st = (u64 *) scratch_get(64);
				if (st == nullptr || !cuda_ok(cudaMemsetAsync(st, 0, 8, c->stream), "memset"))
					goto bailout;
				grp_max_kernel<<<grid_for(g->count, 256, NSM * 8), 256, 0, c->stream>>>((const oid *) g->base, g->count, st);
				COUNT_LAUNCH();
				if (!cuda_ok(cudaMemcpyAsync(c->slot_h->v, st, 8, cudaMemcpyDeviceToHost, c->stream), "read") || !stream_sync())
					goto bailout;
				maxgrp = c->slot_h->v[0];
			}
			if (maxgrp == 0)
				g = nullptr;	/* single group (gdk_group.c:761-762) */
		}
		if (ci.ncand >= (1ULL << 32)) {
			set_err("BATgroup: more than 2^32 candidate rows in one device call\n");
			goto bailout;
		}
		gn = bat_new(hseqb, B200_oid, ci.ncand);
		if (gn == nullptr)
			goto bailout;
		bool ok = false;
		switch (basetype(b->ttype)) {
		case B200_bte: ok = group_typed<bte>(b, ci, g, maxgrp, gn, &en, &hn, extents != nullptr, histo != nullptr); break;
		case B200_sht: ok = group_typed<sht>(b, ci, g, maxgrp, gn, &en, &hn, extents != nullptr, histo != nullptr); break;
		case B200_int: ok = group_typed<int>(b, ci, g, maxgrp, gn, &en, &hn, extents != nullptr, histo != nullptr); break;
		case B200_lng: ok = group_typed<lng>(b, ci, g, maxgrp, gn, &en, &hn, extents != nullptr, histo != nullptr); break;
		case B200_hge: ok = group_typed<hge>(b, ci, g, maxgrp, gn, &en, &hn, extents != nullptr, histo != nullptr); break;
		case B200_flt: ok = group_typed<flt>(b, ci, g, maxgrp, gn, &en, &hn, extents != nullptr, histo != nullptr); break;
		case B200_dbl: ok = group_typed<dbl>(b, ci, g, maxgrp, gn, &en, &hn, extents != nullptr, histo != nullptr); break;
		default:
			set_err("BATgroup: unsupported tail type %s (fixed-width atoms only)\n", atomname(b->ttype));
		}
		if (!ok)
			goto bailout;
	}
	*groups = gn;
	if (extents)
		*extents = en;
	if (histo)
		*histo = hn;
	return B200_SUCCEED;
bailout:
	if (gn)
		b200_BATfree(gn);
	if (en)
		b200_BATfree(en);
	if (hn)
		b200_BATfree(hn);
	return B200_FAIL;
}

// BATunique (gdk_unique.c:30-): candidate list of the first row of every distinct tail value —
// "essentially the same output as the extents output of BATgroup" (gdk_unique.c:19-22), so it is
// the first-row table + bitmap passes of BATgroup; the ids are dropped.
extern "C" b200_bat *b200_BATunique(const b200_bat *b, const b200_bat *s)
{
	b200_bat *g = nullptr, *e = nullptr;
	if (b200_BATgroup(&g, &e, nullptr, b, s, nullptr, nullptr, nullptr) != B200_SUCCEED)
		return nullptr;
	b200_BATfree(g);
	ctx()->algo = "unique: cuda first-row table (BATgroup extents)";
	return e;
}
