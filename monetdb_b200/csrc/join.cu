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

// ---------------------------------------------------------------- hash index cache (BAThash)
// The reference caches the hash on the BAT (b->thash under thashlock, gdk_hash.c:989-1039).
// Our descriptor is a plain struct, so the cache lives beside it, keyed by the tail heap.
static std::mutex g_hash_mu;
static std::unordered_map<const void *, HashIdx> g_hash_cache;

void hash_cache_drop(const void *base)
{
	if (base == nullptr)
		return;
	std::lock_guard<std::mutex> lk(g_hash_mu);
	auto it = g_hash_cache.find(base);
	if (it != g_hash_cache.end()) {
		dev_free(it->second.head);
		dev_free(it->second.next);
		g_hash_cache.erase(it);
	}
}

static inline u64 hash_pow2(u64 x)
{
	u64 p = 1024;
	while (p < x)
		p <<= 1;
	return p;
}

template <typename T> static bool hash_build(const ColSrc<T> &r, bool skip_nil, HashIdx &h)
{
	Ctx *c = ctx();
	h.n = r.n;
	h.mask = hash_pow2(r.n) - 1;		/* load factor <= 1, like HASHmask (gdk_hash.c:42-55) */
	h.head = (uint32_t *) dev_alloc((h.mask + 1) * 4);
	h.next = (uint32_t *) dev_alloc((r.n ? r.n : 1) * 4);
	if (h.head == nullptr || h.next == nullptr ||
	    !cuda_ok(cudaMemsetAsync(h.head, 0xff, (h.mask + 1) * 4, c->stream), "memset hash heads"))
		return false;
	if (r.n == 0)
		return true;
	KTimer kt("hash_build", r.n * (sizeof(T) + 8));
	join_build_kernel<T><<<grid_for(r.n, JN_THREADS, NSM * 16), JN_THREADS, 0, c->stream>>>(r, skip_nil, h.head, h.next, h.mask);
	COUNT_LAUNCH();
	return cuda_ok(cudaGetLastError(), "hash build launch");
}

static bool hash_build_any(int t, const b200_bat *b, const Cand &ci, bool skip_nil, HashIdx &h)
{
	switch (t) {
	case B200_bte: return hash_build<bte>(col_src<bte>(b, ci), skip_nil, h);
	case B200_sht: return hash_build<sht>(col_src<sht>(b, ci), skip_nil, h);
	case B200_int: return hash_build<int>(col_src<int>(b, ci), skip_nil, h);
	case B200_lng: return hash_build<lng>(col_src<lng>(b, ci), skip_nil, h);
	case B200_hge: return hash_build<hge>(col_src<hge>(b, ci), skip_nil, h);
	case B200_flt: return hash_build<flt>(col_src<flt>(b, ci), skip_nil, h);
	case B200_dbl: return hash_build<dbl>(col_src<dbl>(b, ci), skip_nil, h);
	}
	set_err("BAThash: unsupported tail type %s (fixed-width atoms only)\n", atomname(t));
	return false;
}

extern "C" b200_return b200_BAThash(const b200_bat *b)
{
	b200_clrerr();
	if (b == nullptr) {
		set_err("BAThash: BAT required\n");
		return B200_FAIL;
	}
	if (b->device < 0 || b->ttype == B200_void || b->base == nullptr) {
		set_err("BAThash: a device-resident materialised column is required\n");
		return B200_FAIL;
	}
	if (b->count >= JN_END) {
		set_err("BAThash: more than 2^32-1 rows in one device call\n");
		return B200_FAIL;
	}
	{
		std::lock_guard<std::mutex> lk(g_hash_mu);
		auto it = g_hash_cache.find(b->base);
		if (it != g_hash_cache.end() && it->second.n == b->count)
			return B200_SUCCEED;
	}
	hash_cache_drop(b->base);
	Cand ci;
	if (!cand_init(ci, b, nullptr))
		return B200_FAIL;
	HashIdx h;
	ctx()->algo = "hash: cuda chained build";
	/* nils are indexed too (the reference hashes every BUN); probes skip them unless nil_matches */
	if (!hash_build_any(basetype(b->ttype), b, ci, false, h) || !stream_sync()) {
		dev_free(h.head);
		dev_free(h.next);
		return B200_FAIL;
	}
	std::lock_guard<std::mutex> lk(g_hash_mu);
	g_hash_cache[b->base] = h;
	return B200_SUCCEED;
}

// ---------------------------------------------------------------- BATjoin
template <typename T>
static bool join_typed(const b200_bat *l, const b200_bat *r, const Cand &lci, const Cand &rci, bool nil_matches,
		       const HashIdx &h, b200_bat **r1p, b200_bat **r2p)
{
	Ctx *c = ctx();
	const ColSrc<T> ls = col_src<T>(l, lci), rs = col_src<T>(r, rci);
	const uint64_t nl = ls.n, nchunks = (nl + JN_CHUNK - 1) / JN_CHUNK;
	uint32_t *cnt = (uint32_t *) dev_alloc(nl * 4), *first = (uint32_t *) dev_alloc(nl * 4);
	unsigned long long *tot = (unsigned long long *) dev_alloc(nchunks * 8);
	unsigned long long *offs = (unsigned long long *) dev_alloc((nchunks + 1) * 8);
	unsigned *flags = (unsigned *) dev_alloc(16);
	b200_bat *r1 = nullptr, *r2 = nullptr;
	bool ok = false;
	uint64_t total = 0;
	if (cnt == nullptr || first == nullptr || tot == nullptr || offs == nullptr || flags == nullptr ||
	    !cuda_ok(cudaMemsetAsync(flags, 0, 16, c->stream), "memset flags"))
		goto done;
	{
		KTimer kt("join_probe", nl * (sizeof(T) + 4 + 4 + sizeof(T) + 8));
		join_count_kernel<T><<<(unsigned) nchunks, JN_THREADS, 0, c->stream>>>(ls, rs, nil_matches, h.head, h.next, h.mask, cnt, first, tot);
	}
	COUNT_LAUNCH();
	join_scan_kernel<<<1, 1024, 0, c->stream>>>(tot, nchunks, offs, c->slot_d->v);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "join probe launch") || !stream_sync())
		goto done;
	total = c->slot_h->v[0];
	r1 = bat_new(0, B200_oid, total);
	if (r1 == nullptr)
		goto done;
	if (r2p) {
		r2 = bat_new(0, B200_oid, total);
		if (r2 == nullptr)
			goto done;
	}
	if (total > 0) {
		{
			KTimer kt("join_fill", nl * 8 + total * (r2 ? 16 : 8));
			join_fill_kernel<T><<<(unsigned) nchunks, JN_THREADS, 0, c->stream>>>(ls, rs, h.next, cnt, first, offs, (oid *) r1->base,
											      r2 ? (oid *) r2->base : nullptr, flags);
		}
		COUNT_LAUNCH();
		if (!cuda_ok(cudaMemcpyAsync(c->slot_h->v, flags, 8, cudaMemcpyDeviceToHost, c->stream), "read flags") || !stream_sync())
			goto done;
	}
	{
		const bool multi = total > 0 && (unsigned) c->slot_h->v[0] != 0;
		/* result properties (gdk_join.c:2877-2893,3260-3296): left-major output in candidate order */
		r1->count = total;
		r1->tsorted = 1;
		r1->trevsorted = total <= 1;
		r1->tkey = !multi;
		r1->tnonil = 1;
		r1->tnil = 0;
		r1->tseqbase = total == 0 ? 0 : OID_NIL;
		if (r2) {
			r2->count = total;
			r2->tsorted = r2->trevsorted = total <= 1;
			r2->tkey = total <= 1;
			r2->tnonil = 1;
			r2->tnil = 0;
			r2->tseqbase = total == 0 ? 0 : OID_NIL;
		}
	}
	ok = true;
done:
	dev_free(cnt);
	dev_free(first);
	dev_free(tot);
	dev_free(offs);
	dev_free(flags);
	if (!ok) {
		if (r1)
			b200_BATfree(r1);
		if (r2)
			b200_BATfree(r2);
		return false;
	}
	*r1p = r1;
	if (r2p)
		*r2p = r2;
	return true;
}

extern "C" b200_return b200_BATjoin(b200_bat **r1p, b200_bat **r2p, const b200_bat *l, const b200_bat *r,
				    const b200_bat *sl, const b200_bat *sr, bool nil_matches, uint64_t estimate)
{
	(void) estimate;	/* a size hint for the reference's growing result; sizes are exact here */
	b200_clrerr();
	if (r1p)
		*r1p = nullptr;
	if (r2p)
		*r2p = nullptr;
	if (r1p == nullptr || l == nullptr || r == nullptr) {
		set_err("BATjoin: l, r and r1p required\n");
		return B200_FAIL;
	}
	Staged stl, str_, stsl, stsr;
	if (!stage(l, stl) || !stage(r, str_) || !stage(sl, stsl) || !stage(sr, stsr))
		return B200_FAIL;
	const b200_bat *lb = stl.b, *rb = str_.b;
	const int t = basetype(lb->ttype);
	if (t != basetype(rb->ttype)) {
		/* joininitresults / BATjoin: ATOMtype(l->ttype) == ATOMtype(r->ttype) asserted (gdk_join.c:3890) */
		set_err("BATjoin: inputs not compatible (%s vs %s).\n", atomname(lb->ttype), atomname(rb->ttype));
		return B200_FAIL;
	}
	Cand lci, rci;
	if (!cand_init(lci, lb, stsl.b) || !cand_init(rci, rb, stsr.b))
		return B200_FAIL;
	if (lci.ncand == 0 || rci.ncand == 0) {
		/* nomatch (gdk_join.c:302-360) */
		ctx()->algo = "join: nomatch";
		*r1p = bat_dense(0, 0, 0);
		if (r2p)
			*r2p = bat_dense(0, 0, 0);
		if (*r1p == nullptr || (r2p && *r2p == nullptr))
			return B200_FAIL;
		return B200_SUCCEED;
	}
	if (lb->ttype == B200_void || rb->ttype == B200_void) {
		set_err("BATjoin: void columns are not on the device path\n");
		return B200_FAIL;
	}
	if (lci.ncand >= JN_END || rci.ncand >= JN_END) {
		set_err("BATjoin: more than 2^32-1 rows in one device call\n");
		return B200_FAIL;
	}
	/* a cached whole-BAT hash (BAThash) can serve when r is used without candidates */
	HashIdx h;
	bool owned = true;
	if (stsr.b == nullptr && rb == r) {
		std::lock_guard<std::mutex> lk(g_hash_mu);
		auto it = g_hash_cache.find(rb->base);
		if (it != g_hash_cache.end() && it->second.n == rb->count) {
			h = it->second;
			owned = false;
		}
	}
	ctx()->algo = owned ? "join: cuda hashjoin (build + probe)" : "join: cuda hashjoin (existing hash)";
	bool ok = true;
	if (owned)
		ok = hash_build_any(t, rb, rci, !nil_matches, h);
	if (ok) {
		switch (t) {
		case B200_bte: ok = join_typed<bte>(lb, rb, lci, rci, nil_matches, h, r1p, r2p); break;
		case B200_sht: ok = join_typed<sht>(lb, rb, lci, rci, nil_matches, h, r1p, r2p); break;
		case B200_int: ok = join_typed<int>(lb, rb, lci, rci, nil_matches, h, r1p, r2p); break;
		case B200_lng: ok = join_typed<lng>(lb, rb, lci, rci, nil_matches, h, r1p, r2p); break;
		case B200_hge: ok = join_typed<hge>(lb, rb, lci, rci, nil_matches, h, r1p, r2p); break;
		case B200_flt: ok = join_typed<flt>(lb, rb, lci, rci, nil_matches, h, r1p, r2p); break;
		case B200_dbl: ok = join_typed<dbl>(lb, rb, lci, rci, nil_matches, h, r1p, r2p); break;
		default: ok = false;
		}
	}
	if (owned) {
		dev_free(h.head);
		dev_free(h.next);
	}
	return ok ? B200_SUCCEED : B200_FAIL;
}
