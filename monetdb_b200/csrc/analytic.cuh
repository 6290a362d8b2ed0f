// analytic.cuh — shared machinery of the window-function kernels
// (gdk_analytic_bounds.c, gdk_analytic_func.c, gdk_analytic_statistics.c).
//
// The reference walks every partition sequentially (`for (; i < cnt; i++) if (np[i]) {...}`,
// e.g. gdk_analytic_bounds.c:201-215).  On the device a partitioning is a SEGMENT
// LIST: the ascending row positions where a new partition (or peer group) starts,
// produced by the ordered compaction of compact.cuh, closed by a sentinel equal to
// the row count.  Every kernel processes tiles of AN_TILE rows; a tile stages the
// (usually one or two) segment boundaries that intersect it in shared memory, so a
// row finds its partition [start, end) and partition number without touching
// global memory again.
//
// Framed aggregates use the reference's fan-out-16 segment tree
// (gdk_analytic.h:52-130), built per partition and aligned to the partition start
// exactly as populate_segment_tree does, so that the order of floating-point
// additions — and the reference's "average of node averages" behaviour for framed
// integer averages (gdk_analytic_statistics.c:163-185) — are reproduced bit for bit.
// Level 0 is the input column itself (never materialised); level l >= 1 of
// partition q lives at slot  (S[q] >> 4l) + q + t  of one array per level.
#pragma once
#include "common.cuh"
#include "compact.cuh"

#define AN_TILE 1024
#define AN_THREADS 256
#define AN_FANOUT 16
#define AN_MAXLEV 16

enum { AN_ERR_OVERFLOW = 1, AN_ERR_BOUND = 2 };

struct SegList {
	oid *S = nullptr;	/* cnt starts + sentinel S[cnt] = n */
	uint64_t cnt = 0, n = 0;
	bool cached = false;	/* S belongs to the flag column's SHADOW_SEGLIST twin */
	uint64_t token = 0;	/* cached: identity of this list (segment trees built over it are keyed by it) */
	uint64_t maxlen1 = 0;	/* cached: longest partition + 1 once some aggregate measured it, else 0 */
	const void *pbase = nullptr;	/* cached: the heap the twin hangs on */
	~SegList() { if (S && !cached) dev_free(S); }
};

static __global__ void an_set_u64(oid *p, oid v)
{
	*p = v;
}

// partition starts: row 0 and every row whose p (or o) flag is set
struct FlagStart {
	const bte *p, *o;
	__device__ __forceinline__ bool operator()(uint64_t idx, oid &ov) const
	{
		return idx == 0 || (p && p[idx]) || (o && o[idx]);
	}
	__device__ __forceinline__ oid at(uint64_t idx) const { return idx; }
};
// plain set bits of a bit column (peer-group starts for GROUPS frames)
struct FlagBits {
	const bte *b;
	__device__ __forceinline__ bool operator()(uint64_t idx, oid &ov) const { return b[idx] != 0; }
	__device__ __forceinline__ oid at(uint64_t idx) const { return idx; }
};

#ifdef __CUDACC__
// the mask pass of the compaction for a plain byte-flag column: a thread turns 32 consecutive flags (two 16-byte
// loads) into one mask word, instead of one byte load per lane and a ballot
__device__ __forceinline__ uint32_t flag_nibble(uint32_t x)
{
	const uint32_t m = ((((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u) >> 7;	/* bit 8k = byte k is non-zero */
	return (m | m >> 7 | m >> 14 | m >> 21) & 0xfu;
}
template <int CHUNK>
__global__ void __launch_bounds__(CP_THREADS)
flag_mask_kernel(uint64_t n, const bte *__restrict__ p, uint32_t *__restrict__ words, uint32_t *__restrict__ counts)
{
	__shared__ unsigned total;
	if (threadIdx.x == 0)
		total = 0;
	__syncthreads();
	unsigned cnt = 0;
	if (threadIdx.x < CHUNK / 32) {
		const uint64_t idx = (uint64_t) blockIdx.x * CHUNK + (uint64_t) threadIdx.x * 32;
		uint32_t word = 0;
		if (idx + 32 <= n) {
			const uint4 a = reinterpret_cast<const uint4 *>(p + idx)[0], b = reinterpret_cast<const uint4 *>(p + idx)[1];
			word = flag_nibble(a.x) | flag_nibble(a.y) << 4 | flag_nibble(a.z) << 8 | flag_nibble(a.w) << 12 |
			       flag_nibble(b.x) << 16 | flag_nibble(b.y) << 20 | flag_nibble(b.z) << 24 | flag_nibble(b.w) << 28;
		} else {
			for (unsigned k = 0; idx + k < n; k++)
				word |= (uint32_t) (p[idx + k] != 0) << k;
		}
		if (idx == 0)
			word |= 1u;	/* row 0 starts a partition */
		words[idx >> 5] = word;
		cnt = __popc(word);
	}
	cnt = __reduce_add_sync(0xffffffffu, cnt);
	if ((threadIdx.x & 31) == 0 && cnt)
		atomicAdd(&total, cnt);
	__syncthreads();
	if (threadIdx.x == 0)
		counts[blockIdx.x] = total;
}
#endif

// ordered compaction of the row positions for which f(idx) holds into an exactly
// sized list (mask -> scan -> read the total -> allocate -> expand)
template <typename F> static bool build_seglist(uint64_t n, F f, SegList &out)
{
	Ctx *c = ctx();
	out.n = n;
	out.cnt = 0;
	if (n == 0) {
		out.S = (oid *) dev_alloc(8);
		if (out.S == nullptr)
			return false;
		an_set_u64<<<1, 1, 0, c->stream>>>(out.S, 0);
		COUNT_LAUNCH();
		return true;
	}
	CompactScratch s;
	if (!compact_scratch(n, s))
		return false;
	{
		KTimer kt("an_seg_mask", n);
		bool done = false;
		if constexpr (std::is_same<F, FlagStart>::value) {
			if (f.p != nullptr && f.o == nullptr && ((uintptr_t) f.p & 15) == 0) {
				if (s.chunk == CP_CHUNK)
					flag_mask_kernel<CP_CHUNK><<<(unsigned) s.nchunks, CP_THREADS, 0, c->stream>>>(n, f.p, s.words, s.counts);
				else
					flag_mask_kernel<CP_CHUNK_SMALL><<<(unsigned) s.nchunks, CP_THREADS, 0, c->stream>>>(n, f.p, s.words, s.counts);
				COUNT_LAUNCH();
				done = true;
			}
		}
		if (!done)
			compact_launch_mask(s, n, f);
	}
	compact_launch_scan(s);
	if (!cuda_ok(cudaGetLastError(), "seglist launch") || !stream_sync())
		return false;
	out.cnt = c->slot_h->v[0];
	out.S = (oid *) dev_alloc((out.cnt + 1) * 8);
	if (out.S == nullptr)
		return false;
	if (out.cnt) {
		compact_launch_expand(s, FunctorOid<F>{f}, out.S);
	}
	an_set_u64<<<1, 1, 0, c->stream>>>(out.S + out.cnt, n);
	COUNT_LAUNCH();
	return cuda_ok(cudaGetLastError(), "seglist expand");
}

#ifdef __CUDACC__
// number of entries of S[0..cnt) that are <= k
__device__ __forceinline__ uint64_t seg_upper(const oid *__restrict__ S, uint64_t cnt, oid k)
{
	uint64_t lo = 0, hi = cnt;
	while (lo < hi) {
		uint64_t mid = (lo + hi) >> 1;
		if (S[mid] <= k)
			lo = mid + 1;
		else
			hi = mid;
	}
	return lo;
}

// the same count found by one whole warp: 32 probes per step narrow [lo, hi] about 33-fold, so a list of 2^14
// partitions takes 3 dependent loads instead of 14 (every tile kernel starts with two of these searches, with the
// rest of the CTA waiting).  The answer must lie in [lo, hi]; all 32 lanes call it and get the result.
__device__ __forceinline__ uint64_t seg_upper_warp(const oid *__restrict__ S, uint64_t lo, uint64_t hi, oid k)
{
	const unsigned lane = threadIdx.x & 31;
	while (lo < hi) {
		const uint64_t step = (hi - lo) / 33 + 1;
		const uint64_t pos = lo + step * (lane + 1) - 1;
		const bool le = pos < hi && S[pos] <= k;	/* monotone over the lanes */
		const unsigned t = __popc(__ballot_sync(0xffffffffu, le));
		if (t < 32) {
			const uint64_t cap = lo + step * (t + 1) - 1;	/* first probe that was > k (or beyond hi) */
			hi = cap < hi ? cap : hi;
		}
		lo += step * t;
	}
	return lo;
}

// The slice of a segment list that intersects one tile, staged in shared memory.
struct TileSegs {
	oid sm[AN_TILE + 2];
	uint64_t q0;		/* index in S of sm[0] */
	unsigned m;		/* sm[0..m] valid; sm[m] > last row of the tile */

	// all threads of the CTA must call this
	__device__ __forceinline__ void load(const oid *__restrict__ S, uint64_t cnt, uint64_t first, uint64_t last)
	{
		if (threadIdx.x < 32) {		/* tile kernels run AN_THREADS (>= 32) threads */
			const uint64_t u0 = seg_upper_warp(S, 0, cnt, first);
			/* at most last - first starts lie inside the tile, and S is strictly ascending */
			const uint64_t u1 = seg_upper_warp(S, u0, u0 + (last - first) < cnt ? u0 + (last - first) : cnt, last);
			if (threadIdx.x == 0) {
				q0 = u0 ? u0 - 1 : 0;
				m = (unsigned) (u1 - q0);
			}
		}
		__syncthreads();
		for (unsigned i = threadIdx.x; i <= m; i += blockDim.x)
			sm[i] = S[q0 + i];
		__syncthreads();
	}
	// number of list entries <= k (global count); start/end of the segment holding k
	__device__ __forceinline__ uint64_t locate(oid k, oid &start, oid &end) const
	{
		unsigned lo = 0, hi = m;
		while (lo < hi) {
			unsigned mid = (lo + hi) >> 1;
			if (sm[mid] <= k)
				lo = mid + 1;
			else
				hi = mid;
		}
		start = lo ? sm[lo - 1] : 0;
		end = sm[lo];
		return q0 + lo;
	}
};

// ADD_WITH_CHECK / SUB_WITH_CHECK (gdk_calc_private.h:53-102), usable for every atom
template <typename T> __device__ __forceinline__ bool add_with_check(T lft, T rgt, T &dst)
{
	const T mx = Atom<T>::maxval();
	if (rgt < 1) {
		if (-mx - rgt > lft)
			return false;
	} else {
		if (mx - rgt < lft)
			return false;
	}
	dst = lft + rgt;
	return true;
}
template <typename T> __device__ __forceinline__ bool sub_with_check(T lft, T rgt, T &dst)
{
	const T mx = Atom<T>::maxval();
	if (rgt < 1) {
		if (mx + rgt < lft)
			return false;
	} else {
		if (-mx + rgt > lft)
			return false;
	}
	dst = lft - rgt;
	return true;
}
#endif

// shared host helpers (analytic_bounds.cu)
bool an_partitions(const b200_bat *p, const b200_bat *o, uint64_t n, SegList &out);
bool an_check_bit(const b200_bat *x, const char *what, uint64_t n);
bool an_read_state(unsigned long long *state, unsigned long long out[2]);
unsigned long long *an_state_init();
