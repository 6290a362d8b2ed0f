// compact.cuh — ordered stream compaction into oid lists: the building block
// behind every operator that emits a candidate list (scanselect's
// `dst[cnt] = o; cnt += TEST` loop, gdk_select.c:219-248; candidate set algebra,
// gdk_cand.c:46-400; mask expansion, gdk_cand.c:1464).
//
// Three dependency-free steps (a first version used a single-pass decoupled
// look-back; ncu showed 60 % of its warp samples parked on the look-back barrier,
// see profiles/r1_select_dense_v0.md):
//   1. MASK    one streaming pass evaluates the predicate and leaves ONE BIT per
//              input index (32 indexes per word, LSB first — the layout of the
//              reference's msk candidates, gdk.h:835-860) plus a hit count per
//              chunk of CP_CHUNK indexes;
//   2. SCAN    a single CTA turns the chunk counts into output offsets;
//   3. EXPAND  one warp per 32 mask words writes the oids of the set bits,
//              256 contiguous bytes per store instruction.
// The input column is read once; the only extra traffic is the bitmask (n/8 bytes
// written and read back, L2-resident for all but the largest inputs).
#pragma once
#include "common.cuh"

#define CP_THREADS 256
#define CP_CHUNK 8192			/* indexes per CTA for large inputs: 256 mask words */
#define CP_CHUNK_SMALL 2048		/* ... and for inputs too small to fill the 148 SMs with 8192-index chunks */
#define CP_WORDS (CP_CHUNK / 32)

struct CompactScratch {
	uint32_t *words;		/* nchunks * chunk / 32 */
	uint32_t *counts;		/* nchunks */
	unsigned long long *offs;	/* nchunks + 1: exclusive offsets, then the total */
	uint64_t nchunks;
	unsigned chunk;			/* CP_CHUNK or CP_CHUNK_SMALL */
};

static inline bool compact_scratch(uint64_t n, CompactScratch &s, unsigned force_chunk = 0)
{
	s.chunk = force_chunk ? force_chunk : (n < (uint64_t) NSM * 8 * CP_CHUNK ? CP_CHUNK_SMALL : CP_CHUNK);
	s.nchunks = (n + s.chunk - 1) / s.chunk;
	size_t wbytes = s.nchunks * (s.chunk / 32) * 4, cbytes = (s.nchunks * 4 + 15) & ~(size_t) 15;
	char *p = (char *) scratch_get(wbytes + cbytes + (s.nchunks + 1) * 8);
	if (p == nullptr)
		return false;
	s.words = (uint32_t *) p;
	s.counts = (uint32_t *) (p + wbytes);
	s.offs = (unsigned long long *) (p + wbytes + cbytes);
	return true;
}

// ---- step 1 (generic form): F is a functor with
//        bool operator()(uint64_t idx, oid &o)   the test
//        oid  at(uint64_t idx)                   the oid to emit for a hit
template <int CHUNK, typename F>
__global__ void __launch_bounds__(CP_THREADS)
compact_mask_kernel(uint64_t n, F f, uint32_t *__restrict__ words, uint32_t *__restrict__ counts)
{
	__shared__ unsigned total;
	if (threadIdx.x == 0)
		total = 0;
	__syncthreads();
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint64_t cbase = (uint64_t) blockIdx.x * CHUNK;
	unsigned cnt = 0;		/* warp-uniform */
#pragma unroll 8
	for (int j = 0; j < CHUNK / CP_THREADS; j++) {
		/* warp w evaluates 32 consecutive indexes per step: the ballot IS the mask word */
		const uint64_t idx = cbase + (uint64_t) (j * (CP_THREADS / 32) + warp) * 32 + lane;
		oid o;
		const bool h = idx < n && f(idx, o);
		const unsigned bal = __ballot_sync(0xffffffffu, h);
		if (lane == 0)
			words[idx >> 5] = bal;
		cnt += __popc(bal);
	}
	if (lane == 0 && cnt)
		atomicAdd(&total, cnt);
	__syncthreads();
	if (threadIdx.x == 0)
		counts[blockIdx.x] = total;
}

// ---- step 1 for two-level functors (a candidate-list gather): F additionally has
//        K    fetch(uint64_t idx)      first-level load (the candidate oid)
//        V    gather(K k)              second-level load (the column value)
//        bool test(V v)                the predicate
// All first-level loads of a thread's U steps are issued before the second-level ones and those
// before any ballot/store, so U dependent load chains are in flight per thread instead of one
// (the store to words[] otherwise orders every iteration behind the previous one).
template <int CHUNK, typename F>
__global__ void __launch_bounds__(CP_THREADS)
compact_mask2_kernel(uint64_t n, F f, uint32_t *__restrict__ words, uint32_t *__restrict__ counts)
{
	__shared__ unsigned total;
	if (threadIdx.x == 0)
		total = 0;
	__syncthreads();
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint64_t cbase = (uint64_t) blockIdx.x * CHUNK;
	unsigned cnt = 0;		/* warp-uniform */
	constexpr int U = 8;
	static_assert((CHUNK / CP_THREADS) % U == 0, "chunk steps must be a multiple of the unroll");
	for (int j0 = 0; j0 < CHUNK / CP_THREADS; j0 += U) {
		decltype(f.fetch(0)) k[U];
		decltype(f.gather(k[0])) v[U];
		uint64_t idx[U];
#pragma unroll
		for (int u = 0; u < U; u++) {
			idx[u] = cbase + (uint64_t) ((j0 + u) * (CP_THREADS / 32) + warp) * 32 + lane;
			if (idx[u] < n)
				k[u] = f.fetch(idx[u]);
		}
#pragma unroll
		for (int u = 0; u < U; u++)
			if (idx[u] < n)
				v[u] = f.gather(k[u]);
#pragma unroll
		for (int u = 0; u < U; u++) {
			const bool h = idx[u] < n && f.test(v[u]);
			const unsigned bal = __ballot_sync(0xffffffffu, h);
			if (lane == 0)
				words[idx[u] >> 5] = bal;
			cnt += __popc(bal);
		}
	}
	if (lane == 0 && cnt)
		atomicAdd(&total, cnt);
	__syncthreads();
	if (threadIdx.x == 0)
		counts[blockIdx.x] = total;
}

// ---- step 2: exclusive scan of the chunk counts by one CTA; total -> slot[0].
// Every thread owns a contiguous run of counts, so all its loads are independent.
static __global__ void __launch_bounds__(1024)
compact_scan_kernel(const uint32_t *__restrict__ counts, uint64_t nchunks, unsigned long long *__restrict__ offs,
		    unsigned long long *__restrict__ slot)
{
	__shared__ unsigned long long wsum[32];
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint64_t items = (nchunks + 1023) / 1024;
	const uint64_t lo = threadIdx.x * items, hi = lo + items < nchunks ? lo + items : nchunks;
	unsigned long long mine = 0;
#pragma unroll 8
	for (uint64_t i = lo; i < hi; i++)
		mine += counts[i];
	unsigned long long inc = mine;
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
		wsum[lane] = winc - w;		/* exclusive over warps */
		if (lane == 31) {
			offs[nchunks] = winc;
			slot[0] = winc;
			slot[1] = 0;
			slot[2] = 0;
		}
	}
	__syncthreads();
	unsigned long long run = wsum[warp] + inc - mine;
	for (uint64_t i = lo; i < hi; i++) {
		offs[i] = run;
		run += counts[i];
	}
}

// ---- step 3: mask words -> oids.  OidOf gives the oid of input index idx.
// One lane per mask word.  Each lane drops the positions of its set bits into a
// per-warp shared-memory stage at its exclusive offset; the warp then streams the
// stage out, 32 consecutive oids (256 contiguous bytes) per store instruction.
struct DenseOid {
	oid seq;	/* oid of index 0 */
	__device__ __forceinline__ oid at(uint64_t idx) const { return seq + idx; }
};

template <int WORDS, typename OidOf>
__global__ void __launch_bounds__(WORDS)
compact_expand_kernel(const uint32_t *__restrict__ words, const uint32_t *__restrict__ counts,
		      const unsigned long long *__restrict__ offs, OidOf of, oid *__restrict__ dst,
		      unsigned long long *__restrict__ slot)
{
	if (counts[blockIdx.x] == 0)
		return;
	__shared__ unsigned short stage[WORDS * 32];
	__shared__ unsigned wtot[WORDS / 32];
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint64_t wbase = (uint64_t) blockIdx.x * WORDS + warp * 32;
	unsigned w = words[wbase + lane];
	const unsigned c = __popc(w);
	unsigned inc = c;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
		if (lane >= o)
			inc += t;
	}
	const unsigned T = __shfl_sync(0xffffffffu, inc, 31);
	if (lane == 31)
		wtot[warp] = inc;
	unsigned short *st = stage + warp * 1024;
	unsigned pos = inc - c;
	const unsigned short lbase = (unsigned short) (lane * 32);
	while (w) {
		const unsigned b = __ffs(w) - 1;
		w &= w - 1;
		st[pos++] = lbase + (unsigned short) b;
	}
	__syncthreads();
	unsigned long long base = offs[blockIdx.x];
#pragma unroll
	for (int k = 0; k < WORDS / 32; k++)
		base += k < warp ? wtot[k] : 0;
	const unsigned long long total = offs[gridDim.x];
	for (unsigned r = lane; r < T; r += 32) {
		const unsigned long long p = base + r;
		const oid o = of.at(wbase * 32 + st[r]);
		dst[p] = o;
		if (p == 0)
			slot[1] = o;
		if (p == total - 1)
			slot[2] = o;
	}
}

// Expand-with-gather: the same per-chunk structure as compact_expand_kernel, but instead of the
// oid of every set bit it emits src[bit index] — BATproject through a select result whose scan
// bitmask is still at hand (SHADOW_SELMASK): the column is streamed once and no oid is read.
template <int WORDS, typename VT>
__global__ void __launch_bounds__(WORDS)
compact_gather_kernel(const uint32_t *__restrict__ words, const unsigned long long *__restrict__ offs,
		      const VT *__restrict__ src, VT *__restrict__ dst)
{
	const unsigned long long cbase = offs[blockIdx.x];
	if (offs[blockIdx.x + 1] == cbase)
		return;
	__shared__ unsigned short stage[WORDS * 32];
	__shared__ unsigned wtot[WORDS / 32];
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint64_t wbase = (uint64_t) blockIdx.x * WORDS + warp * 32;
	unsigned w = words[wbase + lane];
	const unsigned c = __popc(w);
	unsigned inc = c;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
		if (lane >= o)
			inc += t;
	}
	const unsigned T = __shfl_sync(0xffffffffu, inc, 31);
	if (lane == 31)
		wtot[warp] = inc;
	unsigned short *st = stage + warp * 1024;
	unsigned pos = inc - c;
	const unsigned short lbase = (unsigned short) (lane * 32);
	while (w) {
		const unsigned b = __ffs(w) - 1;
		w &= w - 1;
		st[pos++] = lbase + (unsigned short) b;
	}
	__syncthreads();
	unsigned long long base = cbase;
#pragma unroll
	for (int k = 0; k < WORDS / 32; k++)
		base += k < warp ? wtot[k] : 0;
	const VT *sp = src + wbase * 32;
	for (unsigned r = lane; r < T; r += 32)
		dst[base + r] = sp[st[r]];
}

struct CompactResult {
	uint64_t count = 0;
	oid first = 0, last = 0;
};

template <typename OidOf>
static inline void compact_launch_expand(const CompactScratch &s, OidOf of, oid *dst, KTimer *ktp = nullptr)
{
	Ctx *c = ctx();
	KTimer own(ktp ? nullptr : "compact_expand", s.nchunks * (s.chunk / 32) * 4);
	if (s.chunk == CP_CHUNK)
		compact_expand_kernel<CP_CHUNK / 32, OidOf><<<(unsigned) s.nchunks, CP_CHUNK / 32, 0, c->stream>>>(
			s.words, s.counts, s.offs, of, dst, c->slot_d->v);
	else
		compact_expand_kernel<CP_CHUNK_SMALL / 32, OidOf><<<(unsigned) s.nchunks, CP_CHUNK_SMALL / 32, 0, c->stream>>>(
			s.words, s.counts, s.offs, of, dst, c->slot_d->v);
	COUNT_LAUNCH();
}

template <typename F> static inline void compact_launch_mask(const CompactScratch &s, uint64_t n, F f)
{
	Ctx *c = ctx();
	if (s.chunk == CP_CHUNK)
		compact_mask_kernel<CP_CHUNK, F><<<(unsigned) s.nchunks, CP_THREADS, 0, c->stream>>>(n, f, s.words, s.counts);
	else
		compact_mask_kernel<CP_CHUNK_SMALL, F><<<(unsigned) s.nchunks, CP_THREADS, 0, c->stream>>>(n, f, s.words, s.counts);
	COUNT_LAUNCH();
}

static inline void compact_launch_scan(const CompactScratch &s)
{
	Ctx *c = ctx();
	compact_scan_kernel<<<1, 1024, 0, c->stream>>>(s.counts, s.nchunks, s.offs, c->slot_d->v);
	COUNT_LAUNCH();
}

// steps 2 + 3 and the read-back of (count, first oid, last oid)
template <typename OidOf>
static bool compact_finish(const CompactScratch &s, OidOf of, oid *dst, CompactResult &res)
{
	Ctx *c = ctx();
	compact_launch_scan(s);
	KTimer kt("compact_expand", s.nchunks * (s.chunk / 32) * 4);	/* mask words read ... */
	compact_launch_expand(s, of, dst, &kt);
	kt.stop();
	if (!cuda_ok(cudaGetLastError(), "compact launch") || !stream_sync())
		return false;
	kt.add_bytes(c->slot_h->v[0] * sizeof(oid));			/* ... + the oids written */
	res.count = c->slot_h->v[0];
	res.first = c->slot_h->v[1];
	res.last = c->slot_h->v[2];
	return true;
}

template <typename F> struct FunctorOid {
	F f;
	__device__ __forceinline__ oid at(uint64_t idx) const { return f.at(idx); }
};

// driver for two-level (gather) functors
template <typename F>
static bool run_compact_gather(uint64_t n, F f, oid *dst, CompactResult &res)
{
	Ctx *c = ctx();
	if (n == 0) {
		res = CompactResult();
		return true;
	}
	CompactScratch s;
	if (!compact_scratch(n, s))
		return false;
	{
		KTimer kt("compact", n * 8);
		if (s.chunk == CP_CHUNK)
			compact_mask2_kernel<CP_CHUNK, F><<<(unsigned) s.nchunks, CP_THREADS, 0, c->stream>>>(n, f, s.words, s.counts);
		else
			compact_mask2_kernel<CP_CHUNK_SMALL, F><<<(unsigned) s.nchunks, CP_THREADS, 0, c->stream>>>(n, f, s.words, s.counts);
		COUNT_LAUNCH();
	}
	return compact_finish(s, FunctorOid<F>{f}, dst, res);
}

// generic driver (the ITEMS parameter of the first version is kept for the call sites)
template <int ITEMS, typename F>
static bool run_compact(uint64_t n, F f, oid *dst, CompactResult &res)
{
	Ctx *c = ctx();
	if (n == 0) {
		res = CompactResult();
		return true;
	}
	CompactScratch s;
	if (!compact_scratch(n, s))
		return false;
	{
		KTimer kt("compact", n * 8);
		compact_launch_mask(s, n, f);
	}
	return compact_finish(s, FunctorOid<F>{f}, dst, res);
}

// Turn a compacted oid array into a candidate-list BAT with the reference's result
// properties (gdk_select.c:1067-1074) and dense detection (virtualize, :31-89).
// Takes ownership of dst.
static inline b200_bat *oidlist_to_bat(oid *dst, uint64_t cap, const CompactResult &r)
{
	if (r.count == 0 || r.last - r.first + 1 == r.count) {
		dev_free(dst);
		return bat_dense(0, r.count ? r.first : 0, r.count);
	}
	b200_bat *bn = bat_alloc_desc();
	if (bn == nullptr) {
		dev_free(dst);
		return nullptr;
	}
	bn->ttype = B200_oid;
	bn->hseqbase = 0;
	bn->base = dst;
	bn->capacity = cap;
	bn->count = r.count;
	bn->owns = 1;
	bn->tsorted = bn->tkey = bn->tnonil = 1;
	bn->trevsorted = r.count <= 1;
	bn->tnil = 0;
	bn->tseqbase = OID_NIL;
	bn->oidrange_known = 1;
	bn->first_oid = r.first;
	bn->last_oid = r.last;
	return bn;
}
