// join.cu — BAThash and the inner equi hash join BATjoin on the device
// (gdk_hash.c:717-1039 BAThash_impl/BAThash; gdk_join.c:4451-4623 BATjoin,
// :2900-3332 hashjoin, probe loop HASHJOIN :2781-2895).
//
// Build: a chained hash over the candidate positions of r — `head[bucket]` and
// `next[position]`, the device form of the reference's bucket + link arrays
// (gdk_hash.h:16-28).  Rows are pushed with one atomicExch each, so the build is a
// single streaming pass with no locks.
// Probe: the reference appends (lo, ro) pairs while it walks l in candidate order
// and grows the result BATs as it goes (maybeextend, gdk_join.c:257-292).  Here the
// result size is computed first: pass 1 walks each left row's chain and records
// its match count and first match, a scan turns the counts into output offsets,
// pass 2 writes the pairs — rows with a single match (every row of a foreign-key
// join) are written straight from pass 1's record without a second walk.
// Output is left-major in candidate order, like the reference's unswapped
// hashjoin; the order of several right matches of one left row follows the chain
// and is not specified (BATjoin promises no order, gdk_join.c:56-61).
#include "common.cuh"
#include "keys.cuh"
#include <mutex>
#include <unordered_map>

#define JN_THREADS 256
#define JN_ROWS 4				/* left rows per thread */
#define JN_CHUNK (JN_THREADS * JN_ROWS)
#define JN_END 0xffffffffu

struct HashIdx {
	uint32_t *head = nullptr, *next = nullptr;
	u64 mask = 0;
	uint64_t n = 0;
};

template <typename T>
__global__ void __launch_bounds__(JN_THREADS)
join_build_kernel(ColSrc<T> r, bool skip_nil, uint32_t *__restrict__ head, uint32_t *__restrict__ next, u64 mask)
{
	typedef typename KeyBits<T>::type U;
	for (uint64_t i = (uint64_t) blockIdx.x * JN_THREADS + threadIdx.x; i < r.n; i += (uint64_t) gridDim.x * JN_THREADS) {
		const T v = r.value(i);
		if (skip_nil && Atom<T>::is_nil(v)) {
			next[i] = JN_END;
			continue;
		}
		const u64 h = hash_key<U>(KeyBits<T>::get(v), 0) & mask;
		next[i] = atomicExch(head + h, (uint32_t) i);
	}
}

// pass 1: matches per left row (+ the first one met), total per chunk
template <typename T>
__global__ void __launch_bounds__(JN_THREADS)
join_count_kernel(ColSrc<T> l, ColSrc<T> r, bool nil_matches, const uint32_t *__restrict__ head,
		  const uint32_t *__restrict__ next, u64 mask, uint32_t *__restrict__ cnt, uint32_t *__restrict__ first,
		  unsigned long long *__restrict__ chunktot)
{
	typedef typename KeyBits<T>::type U;
	__shared__ unsigned long long total;
	if (threadIdx.x == 0)
		total = 0;
	__syncthreads();
	unsigned long long mine = 0;
#pragma unroll
	for (int j = 0; j < JN_ROWS; j++) {
		const uint64_t k = (uint64_t) blockIdx.x * JN_CHUNK + j * JN_THREADS + threadIdx.x;
		if (k >= l.n)
			break;
		const T v = l.value(k);
		unsigned c = 0, f = JN_END;
		if (nil_matches || !Atom<T>::is_nil(v)) {
			const U kv = KeyBits<T>::get(v);
			for (uint32_t p = __ldg(head + (hash_key<U>(kv, 0) & mask)); p != JN_END; p = __ldg(next + p)) {
				if (KeyBits<T>::get(r.value(p)) == kv) {
					if (c == 0)
						f = p;
					c++;
				}
			}
		}
		cnt[k] = c;
		first[k] = f;
		mine += c;
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
		mine += __shfl_xor_sync(0xffffffffu, mine, o);
	if ((threadIdx.x & 31) == 0 && mine)
		atomicAdd(&total, mine);
	__syncthreads();
	if (threadIdx.x == 0)
		chunktot[blockIdx.x] = total;
}

// exclusive scan of the chunk totals by one CTA; grand total -> offs[nchunks] and slot[0]
__global__ void __launch_bounds__(1024)
join_scan_kernel(const unsigned long long *__restrict__ tot, uint64_t nchunks, unsigned long long *__restrict__ offs,
		 unsigned long long *__restrict__ slot)
{
	__shared__ unsigned long long wsum[32];
	__shared__ unsigned long long itertotal;
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	unsigned long long carry = 0;
	for (uint64_t base = 0; base < nchunks; base += 1024) {
		const uint64_t i = base + threadIdx.x;
		const unsigned long long c = i < nchunks ? tot[i] : 0;
		unsigned long long inc = c;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o);
			if (lane >= o)
				inc += t;
		}
		if (lane == 31)
			wsum[warp] = inc;
		__syncthreads();
		if (warp == 0) {
			const unsigned long long w = wsum[lane];
			unsigned long long winc = w;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				unsigned long long t = __shfl_up_sync(0xffffffffu, winc, o);
				if (lane >= o)
					winc += t;
			}
			wsum[lane] = winc - w;
			if (lane == 31)
				itertotal = winc;
		}
		__syncthreads();
		if (i < nchunks)
			offs[i] = carry + wsum[warp] + inc - c;
		carry += itertotal;
		__syncthreads();
	}
	if (threadIdx.x == 0) {
		offs[nchunks] = carry;
		slot[0] = carry;
	}
}

// pass 2: write the (left oid, right oid) pairs of a chunk at its output offset
template <typename T>
__global__ void __launch_bounds__(JN_THREADS)
join_fill_kernel(ColSrc<T> l, ColSrc<T> r, const uint32_t *__restrict__ next, const uint32_t *__restrict__ cnt,
		 const uint32_t *__restrict__ first, const unsigned long long *__restrict__ offs, oid *__restrict__ r1,
		 oid *__restrict__ r2, unsigned *__restrict__ flags /* [0] a left row matched more than once */)
{
	typedef typename KeyBits<T>::type U;
	__shared__ unsigned long long wtot[JN_THREADS / 32];
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	/* thread t owns the JN_ROWS consecutive rows t*JN_ROWS.. of the chunk, so the scan order is row order */
	const uint64_t k0 = (uint64_t) blockIdx.x * JN_CHUNK + (uint64_t) threadIdx.x * JN_ROWS;
	unsigned c[JN_ROWS];
	unsigned long long mine = 0;
#pragma unroll
	for (int j = 0; j < JN_ROWS; j++) {
		c[j] = k0 + j < l.n ? cnt[k0 + j] : 0;
		mine += c[j];
	}
	unsigned long long inc = mine;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o);
		if (lane >= o)
			inc += t;
	}
	if (lane == 31)
		wtot[warp] = inc;
	__syncthreads();
	unsigned long long pos = offs[blockIdx.x] + inc - mine;
#pragma unroll
	for (int w = 0; w < JN_THREADS / 32; w++)
		pos += w < warp ? wtot[w] : 0;
	bool multi = false;
#pragma unroll
	for (int j = 0; j < JN_ROWS; j++) {
		if (c[j] == 0)
			continue;
		const uint64_t k = k0 + j;
		const oid lo = l.rowoid(k);
		uint32_t p = first[k];
		if (c[j] == 1) {
			r1[pos] = lo;
			if (r2)
				r2[pos] = r.rowoid(p);
			pos++;
			continue;
		}
		multi = true;
		const U kv = KeyBits<T>::get(l.value(k));
		for (; p != JN_END; p = __ldg(next + p)) {
			if (KeyBits<T>::get(r.value(p)) == kv) {
				r1[pos] = lo;
				if (r2)
					r2[pos] = r.rowoid(p);
				pos++;
			}
		}
	}
	if (__syncthreads_or(multi) && threadIdx.x == 0)
		flags[0] = 1;
}
