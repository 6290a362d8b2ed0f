// join.cu — BAThash and the inner equi hash join BATjoin on the device
// (gdk_hash.c:717-1039 BAThash_impl/BAThash; gdk_join.c:4451-4623 BATjoin,
// :2900-3332 hashjoin, probe loop HASHJOIN :2781-2895).
//
// Build: an OPEN-ADDRESSING table over the candidate positions of r (load factor <= 1/2, linear
// probing).  A slot holds the key's canonical bits AND the row position side by side — 8 bytes for
// keys up to 4 bytes, 16 for 8-byte keys, 32 for hge — so a probe costs one aligned vector load per
// slot visited and never touches r's column; the first version's chained index (head[] / next[] /
// key in the column: three dependent random reads per probe) ran the 60 M x 15 M probe at 0.07 of
// HBM peak.  Rows claim a slot with one atomicCAS on its row word and then store the key; probes run
// in later launches, so they always see complete slots.
// A table of up to OA_SMEM_BYTES is STAGED IN SHARED MEMORY by the probe kernel: one elected thread
// arms an mbarrier with the byte count and issues cp.async.bulk (TMA, SASS UBLKCP) copies of the
// table tiles; every probe of the CTA's chunks then stays on chip (dimension-table joins).
// Probe: the reference appends (lo, ro) pairs while it walks l in candidate order
// and grows the result BATs as it goes (maybeextend, gdk_join.c:257-292).  Here the
// result size is computed first: pass 1 probes each left row and records
// its match count and first match, a scan turns the counts into output offsets,
// pass 2 writes the pairs — rows with a single match (every row of a foreign-key
// join) are written straight from pass 1's record without a second probe.
// Output is left-major in candidate order, like the reference's unswapped
// hashjoin; the order of several right matches of one left row follows the probe
// sequence and is not specified (BATjoin promises no order, gdk_join.c:56-61).
#include "common.cuh"
#include "keys.cuh"
#include <mutex>
#include <unordered_map>

#define JN_THREADS 256
#define JN_ROWS 4				/* left rows per thread */
#define JN_CHUNK (JN_THREADS * JN_ROWS)
#define JN_END 0xffffffffu
#define OA_SMEM_BYTES (160u * 1024u)		/* largest table staged in shared memory */
#define OA_TILE (32u * 1024u)			/* bytes per cp.async.bulk */

// one slot: key bits + row position, padded to a power of two so it never straddles a 32-byte sector
template <int KW> struct OaSlot;
template <> struct alignas(8) OaSlot<4> { uint32_t key; uint32_t row; };
template <> struct alignas(16) OaSlot<8> { u64 key; uint32_t row; uint32_t pad; };
template <> struct alignas(32) OaSlot<16> { uhge key; uint32_t row; uint32_t pad[3]; };
template <typename T> struct OaOf {
	typedef typename KeyBits<T>::type U;
	static constexpr int KW = sizeof(U) <= 4 ? 4 : (int) sizeof(U);
	typedef OaSlot<KW> slot;
	typedef decltype(slot::key) K;		/* key bits widened to the slot's key word */
	__device__ static inline K key(T v) { return (K) KeyBits<T>::get(v); }
};

struct HashIdx {
	void *slots = nullptr;	/* OaSlot<KW>[mask + 1] */
	u64 mask = 0;
	uint64_t n = 0;
	unsigned slot_bytes = 0;
};

__global__ void oa_clear_kernel(uint4 *__restrict__ p, uint64_t n16)
{
	const uint4 e = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (uint64_t) gridDim.x * blockDim.x)
		p[i] = e;
}

template <typename T>
__global__ void __launch_bounds__(JN_THREADS)
join_build_kernel(ColSrc<T> r, bool skip_nil, typename OaOf<T>::slot *__restrict__ slots, u64 mask,
		  unsigned *__restrict__ nilseen /* optional: set when r holds a nil (NOT IN, gdk_join.c:3027-3060) */)
{
	typedef typename KeyBits<T>::type U;
	for (uint64_t i = (uint64_t) blockIdx.x * JN_THREADS + threadIdx.x; i < r.n; i += (uint64_t) gridDim.x * JN_THREADS) {
		const T v = r.value(i);
		if (nilseen && Atom<T>::is_nil(v))
			*nilseen = 1;
		if (skip_nil && Atom<T>::is_nil(v))
			continue;
		const typename OaOf<T>::K kv = OaOf<T>::key(v);
		for (u64 h = hash_key<U>(KeyBits<T>::get(v), 0) & mask;; h = (h + 1) & mask) {
			if constexpr (OaOf<T>::KW == 4) {
				/* key and row are one 64-bit word: a single CAS from the all-ones empty pattern fills the slot */
				const unsigned long long want = (unsigned long long) kv | ((unsigned long long) (uint32_t) i << 32);
				if (atomicCAS(reinterpret_cast<unsigned long long *>(slots + h), ~0ULL, want) == ~0ULL)
					break;
			} else if (atomicCAS(&slots[h].row, JN_END, (uint32_t) i) == JN_END) {
				slots[h].key = kv;
				break;
			}
		}
	}
}

// one aligned vector load of a slot (global through the read-only path, or shared memory)
template <int KW> __device__ __forceinline__ OaSlot<KW> oa_load(const OaSlot<KW> *p)
{
	OaSlot<KW> s;
	if constexpr (KW == 4) {
		const uint2 q = __ldg(reinterpret_cast<const uint2 *>(p));
		s.key = q.x;
		s.row = q.y;
	} else if constexpr (KW == 8) {
		const uint4 q = __ldg(reinterpret_cast<const uint4 *>(p));
		s.key = (u64) q.x | ((u64) q.y << 32);
		s.row = q.z;
	} else {
		const uint4 a = reinterpret_cast<const uint4 *>(p)[0];
		s.key = (uhge) ((u64) a.x | ((u64) a.y << 32)) | ((uhge) ((u64) a.z | ((u64) a.w << 32)) << 64);
		s.row = reinterpret_cast<const uint32_t *>(p)[4];
	}
	return s;
}

// ---- TMA bulk copy global -> shared with mbarrier completion (cp.async.bulk, SASS UBLKCP)
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, uint64_t *bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
		     ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned phase)
{
	unsigned done;
	do {
		asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
			     : "=r"(done) : "r"(smem_u32(bar)), "r"(phase) : "memory");
	} while (!done);
}

// The flags of the reference's leftjoin (gdk_join.c:4033-4052) as one word.
enum { JM_NIL_MATCHES = 1, JM_NIL_ON_MISS = 2, JM_SEMI = 4, JM_ONLY_MISSES = 8, JM_NOT_IN = 16, JM_MAX_ONE = 32, JM_MIN_ONE = 64 };
/* flags[] words written by the kernels */
enum { JF_MULTI = 0 /* a left row produced more than one pair */, JF_ERR = 1 /* 1: more than one match, 2: not enough matches */,
       JF_MISS = 2 /* an outer-join row without match (r2 holds a nil) */, JF_RNIL = 3 /* build side holds a nil */ };

// pass 1: result rows per left row (+ the first match met), total per chunk.  One row of l is
// the body of the reference's HASHJOIN loop (gdk_join.c:2781-2895): nil handling, the walk of
// the chain, then the semi / only_misses / nil_on_miss / max_one / min_one rules on the count.
template <typename T, bool SMEM>
__global__ void __launch_bounds__(JN_THREADS)
join_count_kernel(ColSrc<T> l, unsigned mode, const typename OaOf<T>::slot *__restrict__ gslots, u64 mask,
		  uint32_t *__restrict__ cnt, uint32_t *__restrict__ first, unsigned long long *__restrict__ chunktot,
		  uint64_t nchunks, unsigned *__restrict__ flags)
{
	typedef typename KeyBits<T>::type U;
	typedef typename OaOf<T>::slot Slot;
	constexpr int KW = OaOf<T>::KW;
	extern __shared__ __align__(128) unsigned char oa_smem[];
	__shared__ unsigned long long total;
	__shared__ __align__(8) uint64_t bar;
	const Slot *slots = gslots;
	if constexpr (SMEM) {
		/* stage the whole table in shared memory: one thread arms the barrier with the byte count and
		 * issues the bulk copies, everybody waits for the transaction count to drain */
		const unsigned bytes = (unsigned) ((mask + 1) * sizeof(Slot));
		if (threadIdx.x == 0)
			mbar_init(&bar, 1);
		__syncthreads();
		if (threadIdx.x == 0) {
			mbar_expect_tx(&bar, bytes);
			for (unsigned off = 0; off < bytes; off += OA_TILE)
				bulk_g2s(oa_smem + off, (const unsigned char *) gslots + off, bytes - off < OA_TILE ? bytes - off : OA_TILE, &bar);
		}
		mbar_wait(&bar, 0);
		slots = reinterpret_cast<const Slot *>(oa_smem);
	}
	const bool stop_at_first = (mode & JM_ONLY_MISSES) || ((mode & JM_SEMI) && !(mode & JM_MAX_ONE));
	unsigned err = 0;
	bool missed = false;
	for (uint64_t chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
		__syncthreads();
		if (threadIdx.x == 0)
			total = 0;
		__syncthreads();
		unsigned long long mine = 0;
#pragma unroll
		for (int j = 0; j < JN_ROWS; j++) {
			const uint64_t k = chunk * JN_CHUNK + j * JN_THREADS + threadIdx.x;
			if (k >= l.n)
				break;
			const T v = l.value(k);
			unsigned c = 0, f = JN_END;
			const bool nilrow = (!(mode & JM_NIL_MATCHES) || (mode & JM_NOT_IN)) && Atom<T>::is_nil(v);
			if (!nilrow) {
				const typename OaOf<T>::K kv = OaOf<T>::key(v);
				for (u64 h = hash_key<U>(KeyBits<T>::get(v), 0) & mask;; h = (h + 1) & mask) {
					const Slot e = SMEM ? slots[h] : oa_load<KW>(slots + h);
					if (e.row == JN_END)
						break;
					if (e.key == kv) {
						if (c == 0)
							f = e.row;
						c++;
						if (stop_at_first)
							break;
					}
				}
			}
			if (mode & (JM_NIL_ON_MISS | JM_SEMI | JM_ONLY_MISSES | JM_MAX_ONE | JM_MIN_ONE)) {
				if (c > 1 && (mode & JM_MAX_ONE))
					err |= 1;
				if (mode & JM_ONLY_MISSES) {
					c = (c == 0 && !(nilrow && (mode & JM_NOT_IN))) ? 1 : 0;
					f = JN_END;
				} else if (c == 0) {
					if (mode & JM_NIL_ON_MISS) {
						c = 1;
						missed = true;
					} else if (mode & JM_MIN_ONE) {
						err |= 2;
					}
				} else if (mode & JM_SEMI) {
					c = 1;
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
			chunktot[chunk] = total;
	}
	if (err)
		atomicOr(flags + JF_ERR, err);
	if (missed)
		flags[JF_MISS] = 1;
}

// exclusive scan of the chunk totals by one CTA; grand total -> offs[nchunks] and slot[0]
__global__ void __launch_bounds__(1024)
join_scan_kernel(const unsigned long long *__restrict__ tot, uint64_t nchunks, unsigned long long *__restrict__ offs,
		 unsigned long long *__restrict__ slot, const unsigned *__restrict__ flags)
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
		slot[1] = flags[JF_ERR];
		slot[2] = flags[JF_MISS];
	}
}

// pass 2: write the (left oid, right oid) pairs of a chunk at its output offset
template <typename T>
__global__ void __launch_bounds__(JN_THREADS)
join_fill_kernel(ColSrc<T> l, ColSrc<T> r, const typename OaOf<T>::slot *__restrict__ slots, u64 mask, const uint32_t *__restrict__ cnt,
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
				r2[pos] = p == JN_END ? OID_NIL : r.rowoid(p);	/* nil_on_miss */
			pos++;
			continue;
		}
		multi = true;
		/* several matches: probe again from the key's home slot and emit every equal key */
		const T v = l.value(k);
		const typename OaOf<T>::K kv = OaOf<T>::key(v);
		unsigned left = c[j];
		for (u64 h = hash_key<U>(KeyBits<T>::get(v), 0) & mask; left > 0; h = (h + 1) & mask) {
			const typename OaOf<T>::slot e = oa_load<OaOf<T>::KW>(slots + h);
			if (e.row == JN_END)
				break;
			if (e.key == kv) {
				r1[pos] = lo;
				if (r2)
					r2[pos] = r.rowoid(e.row);
				pos++;
				left--;
			}
		}
	}
	if (__syncthreads_or(multi) && threadIdx.x == 0)
		flags[JF_MULTI] = 1;
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
		dev_free(it->second.slots);
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

template <typename T> static bool hash_build(const ColSrc<T> &r, bool skip_nil, HashIdx &h, unsigned *nilseen = nullptr)
{
	Ctx *c = ctx();
	typedef typename OaOf<T>::slot Slot;
	h.n = r.n;
	h.mask = hash_pow2(2 * r.n) - 1;		/* load factor <= 1/2 */
	h.slot_bytes = sizeof(Slot);
	const size_t bytes = (size_t) (h.mask + 1) * sizeof(Slot);
	h.slots = dev_alloc(bytes);
	if (h.slots == nullptr)
		return false;
	oa_clear_kernel<<<grid_for(bytes / 16, 256, NSM * 16), 256, 0, c->stream>>>((uint4 *) h.slots, bytes / 16);
	COUNT_LAUNCH();
	if (r.n == 0)
		return cuda_ok(cudaGetLastError(), "hash clear launch");
	KTimer kt("hash_build", r.n * (sizeof(T) + sizeof(Slot)));
	join_build_kernel<T><<<grid_for(r.n, JN_THREADS, NSM * 16), JN_THREADS, 0, c->stream>>>(r, skip_nil, (Slot *) h.slots, h.mask, nilseen);
	COUNT_LAUNCH();
	return cuda_ok(cudaGetLastError(), "hash build launch");
}

static bool hash_build_any(int t, const b200_bat *b, const Cand &ci, bool skip_nil, HashIdx &h, unsigned *nilseen = nullptr)
{
	switch (t) {
	case B200_bte: return hash_build<bte>(col_src<bte>(b, ci), skip_nil, h, nilseen);
	case B200_sht: return hash_build<sht>(col_src<sht>(b, ci), skip_nil, h, nilseen);
	case B200_int: return hash_build<int>(col_src<int>(b, ci), skip_nil, h, nilseen);
	case B200_lng: return hash_build<lng>(col_src<lng>(b, ci), skip_nil, h, nilseen);
	case B200_hge: return hash_build<hge>(col_src<hge>(b, ci), skip_nil, h, nilseen);
	case B200_flt: return hash_build<flt>(col_src<flt>(b, ci), skip_nil, h, nilseen);
	case B200_dbl: return hash_build<dbl>(col_src<dbl>(b, ci), skip_nil, h, nilseen);
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
	ctx()->algo = "hash: cuda open-addressing build";
	/* nils are indexed too (the reference hashes every BUN); probes skip them unless nil_matches */
	if (!hash_build_any(basetype(b->ttype), b, ci, false, h) || !stream_sync()) {
		dev_free(h.slots);
		return B200_FAIL;
	}
	std::lock_guard<std::mutex> lk(g_hash_mu);
	g_hash_cache[b->base] = h;
	return B200_SUCCEED;
}

// ---------------------------------------------------------------- BATjoin and the leftjoin family
template <typename T>
static bool join_typed(const b200_bat *l, const b200_bat *r, const Cand &lci, const Cand &rci, unsigned mode,
		       const HashIdx &h, unsigned *flags, b200_bat **r1p, b200_bat **r2p, const char *func)
{
	Ctx *c = ctx();
	const ColSrc<T> ls = col_src<T>(l, lci), rs = col_src<T>(r, rci);
	const uint64_t nl = ls.n, nchunks = (nl + JN_CHUNK - 1) / JN_CHUNK;
	uint32_t *cnt = (uint32_t *) dev_alloc(nl * 4), *first = (uint32_t *) dev_alloc(nl * 4);
	unsigned long long *tot = (unsigned long long *) dev_alloc(nchunks * 8);
	unsigned long long *offs = (unsigned long long *) dev_alloc((nchunks + 1) * 8);
	b200_bat *r1 = nullptr, *r2 = nullptr;
	bool ok = false, missed = false;
	uint64_t total = 0;
	if (cnt == nullptr || first == nullptr || tot == nullptr || offs == nullptr)
		goto done;
	{
		typedef typename OaOf<T>::slot Slot;
		const size_t tbytes = (size_t) (h.mask + 1) * sizeof(Slot);
		KTimer kt("join_probe", nl * (sizeof(T) + sizeof(Slot) + 8));	/* key + about one slot + count/first per probe row */
		if (tbytes <= OA_SMEM_BYTES && nchunks >= 4) {
			/* small build side: every CTA stages the table in shared memory (cp.async.bulk) and probes on chip */
			static bool attr_set = false;
			if (!attr_set) {
				cudaFuncSetAttribute(join_count_kernel<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) OA_SMEM_BYTES);
				attr_set = true;
			}
			const unsigned grid = (unsigned) (nchunks < (uint64_t) NSM ? nchunks : NSM);
			join_count_kernel<T, true><<<grid, JN_THREADS, tbytes, c->stream>>>(ls, mode, (const Slot *) h.slots, h.mask, cnt, first, tot,
											 nchunks, flags);
			c->algo = "join: cuda hashjoin, table staged in shared memory by cp.async.bulk";
		} else {
			join_count_kernel<T, false><<<(unsigned) nchunks, JN_THREADS, 0, c->stream>>>(ls, mode, (const Slot *) h.slots, h.mask, cnt, first,
												       tot, nchunks, flags);
		}
	}
	COUNT_LAUNCH();
	join_scan_kernel<<<1, 1024, 0, c->stream>>>(tot, nchunks, offs, c->slot_d->v, flags);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "join probe launch") || !stream_sync())
		goto done;
	total = c->slot_h->v[0];
	if (c->slot_h->v[1] & 1) {
		set_err("%s: more than one match\n", func);	/* HASHLOOPBODY, gdk_join.c:2759-2762 */
		goto done;
	}
	if (c->slot_h->v[1] & 2) {
		set_err("%s: not enough matches\n", func);	/* gdk_join.c:2869-2871 */
		goto done;
	}
	missed = c->slot_h->v[2] != 0;
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
			join_fill_kernel<T><<<(unsigned) nchunks, JN_THREADS, 0, c->stream>>>(ls, rs, (const typename OaOf<T>::slot *) h.slots, h.mask, cnt, first, offs, (oid *) r1->base,
											      r2 ? (oid *) r2->base : nullptr, flags);
		}
		COUNT_LAUNCH();
		if (!cuda_ok(cudaMemcpyAsync(c->slot_h->v, flags, 8, cudaMemcpyDeviceToHost, c->stream), "read flags") ||
		    !cuda_ok(cudaMemcpyAsync(c->slot_h->v + 1, r1->base, 8, cudaMemcpyDeviceToHost, c->stream), "read first oid") ||
		    !cuda_ok(cudaMemcpyAsync(c->slot_h->v + 2, (oid *) r1->base + (total - 1), 8, cudaMemcpyDeviceToHost, c->stream), "read last oid") ||
		    !stream_sync())
			goto done;
	}
	{
		const bool multi = total > 0 && (unsigned) c->slot_h->v[0] != 0;
		const oid lo = total ? c->slot_h->v[1] : 0, hi = total ? c->slot_h->v[2] : 0;
		if (!multi && (mode & (JM_SEMI | JM_ONLY_MISSES)) && r2 == nullptr && (total == 0 || hi - lo + 1 == total)) {
			/* virtualize(*r1p) for semi / only_misses results (gdk_join.c:4311-4312) */
			b200_BATfree(r1);
			r1 = bat_dense(0, lo, total);
			if (r1 == nullptr)
				goto done;
		} else {
			/* result properties (gdk_join.c:2877-2893,3260-3296): left-major output in candidate order */
			r1->count = total;
			r1->tsorted = 1;
			r1->trevsorted = total <= 1;
			r1->tkey = !multi;
			r1->tnonil = 1;
			r1->tnil = 0;
			r1->tseqbase = total == 0 ? 0 : OID_NIL;
			if (total > 0) {
				r1->oidrange_known = 1;
				r1->first_oid = lo;
				r1->last_oid = hi;
			}
		}
		if (r2) {
			r2->count = total;
			r2->tsorted = r2->trevsorted = total <= 1;
			r2->tkey = total <= 1;
			r2->tnonil = !missed;
			r2->tnil = missed;
			r2->tseqbase = total == 0 ? 0 : OID_NIL;
		}
	}
	ok = true;
done:
	dev_free(cnt);
	dev_free(first);
	dev_free(tot);
	dev_free(offs);
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

// nomatch (gdk_join.c:302-360): nothing on one side.  Inner / semi joins are empty; an outer join
// returns every candidate of l beside a nil, a difference returns every candidate of l.
static b200_return join_nomatch(b200_bat **r1p, b200_bat **r2p, const Cand &lci, unsigned mode)
{
	ctx()->algo = "join: nomatch";
	if (lci.ncand == 0 || !(mode & (JM_NIL_ON_MISS | JM_ONLY_MISSES))) {
		*r1p = bat_dense(0, 0, 0);
		if (r2p)
			*r2p = bat_dense(0, 0, 0);
	} else {
		*r1p = cand_to_bat(lci);
		if (r2p)
			*r2p = bat_dense(0, OID_NIL, lci.ncand);	/* BATconstant(0, TYPE_void, &oid_nil, ncand) */
		if (r2p && *r2p) {
			(*r2p)->tsorted = (*r2p)->trevsorted = 1;
			(*r2p)->tkey = lci.ncand <= 1;
			(*r2p)->tnonil = 0;
			(*r2p)->tnil = 1;
			(*r2p)->oidrange_known = 0;
		}
	}
	if (*r1p == nullptr || (r2p && *r2p == nullptr)) {
		if (*r1p)
			b200_BATfree(*r1p);
		*r1p = nullptr;
		if (r2p && *r2p) {
			b200_BATfree(*r2p);
			*r2p = nullptr;
		}
		return B200_FAIL;
	}
	return B200_SUCCEED;
}

// The shared driver: BATjoin (gdk_join.c:4451) and leftjoin (:4049) both end in hashjoin with a
// flag set; on the device every flag combination runs the same build + count/scan/fill plan.
static b200_return join_driver(b200_bat **r1p, b200_bat **r2p, const b200_bat *l, const b200_bat *r, const b200_bat *sl,
			       const b200_bat *sr, unsigned mode, const char *func)
{
	b200_clrerr();
	if (r1p)
		*r1p = nullptr;
	if (r2p)
		*r2p = nullptr;
	if (r1p == nullptr || l == nullptr || r == nullptr) {
		set_err("%s: l, r and r1p required\n", func);
		return B200_FAIL;
	}
	Staged stl, str_, stsl, stsr;
	if (!stage(l, stl) || !stage(r, str_) || !stage(sl, stsl) || !stage(sr, stsr))
		return B200_FAIL;
	const b200_bat *lb = stl.b, *rb = str_.b;
	/* ATOMtype: a void column is an oid column */
	const int t = lb->ttype == B200_void ? B200_lng : basetype(lb->ttype);
	if (t != (rb->ttype == B200_void ? B200_lng : basetype(rb->ttype))) {
		/* joinparamcheck: ATOMtype(l->ttype) == ATOMtype(r->ttype) (gdk_join.c:3890) */
		set_err("%s: inputs not compatible (%s vs %s).\n", func, atomname(lb->ttype), atomname(rb->ttype));
		return B200_FAIL;
	}
	Cand lci, rci;
	if (!cand_init(lci, lb, stsl.b) || !cand_init(rci, rb, stsr.b))
		return B200_FAIL;
	if (lci.ncand == 0 || rci.ncand == 0)
		return join_nomatch(r1p, r2p, lci, mode);
	/* dense (void) sides: mergejoin_void when r and its candidates are dense (gdk_join.c:4155-4160, 4556-4560), else
	 * the dense side becomes a real oid column and joins like any other */
	struct TmpBat {
		b200_bat *b = nullptr;
		~TmpBat() { if (b) b200_BATfree(b); }
	} lmat, rmat;
	if (lb->ttype == B200_void || rb->ttype == B200_void) {
		if (lb->ccand != B200_CAND_NONE || rb->ccand != B200_CAND_NONE ||
		    (lb->ttype == B200_void && lb->tseqbase == OID_NIL) || (rb->ttype == B200_void && rb->tseqbase == OID_NIL)) {
			set_err("%s: void-nil columns and candidate-list encodings as join columns are not on the device path\n", func);
			return B200_FAIL;
		}
		if (rb->ttype == B200_void && rci.dense && lb->ttype != B200_msk) {
			if (!join_dense_right(r1p, r2p, lb, rb, lci, rci, mode & (JM_NIL_ON_MISS | JM_ONLY_MISSES)))
				return B200_FAIL;
			return B200_SUCCEED;
		}
		if (lb->ttype == B200_void) {
			if ((lmat.b = join_materialise_dense(lb)) == nullptr)
				return B200_FAIL;
			lb = lmat.b;
		}
		if (rb->ttype == B200_void) {
			if ((rmat.b = join_materialise_dense(rb)) == nullptr)
				return B200_FAIL;
			rb = rmat.b;
		}
	}
	if (lci.ncand >= JN_END || rci.ncand >= JN_END) {
		set_err("%s: more than 2^32-1 rows in one device call\n", func);
		return B200_FAIL;
	}
	const bool nil_matches = (mode & JM_NIL_MATCHES) != 0, not_in = (mode & JM_NOT_IN) != 0;
	/* a cached whole-BAT hash (BAThash) can serve when r is used without candidates */
	HashIdx h;
	bool owned = true;
	if (stsr.b == nullptr && rb == r && !not_in) {
		std::lock_guard<std::mutex> lk(g_hash_mu);
		auto it = g_hash_cache.find(rb->base);
		if (it != g_hash_cache.end() && it->second.n == rb->count) {
			h = it->second;
			owned = false;
		}
	}
	ctx()->algo = owned ? "join: cuda hashjoin (build + probe)" : "join: cuda hashjoin (existing hash)";
	unsigned *flags = (unsigned *) dev_alloc(16);
	bool ok = flags != nullptr && cuda_ok(cudaMemsetAsync(flags, 0, 16, ctx()->stream), "memset flags");
	bool empty = false;
	if (ok && owned)
		ok = hash_build_any(t, rb, rci, !nil_matches || not_in, h, not_in ? flags + JF_RNIL : nullptr);
	if (ok && not_in) {
		/* NOT IN with a nil on the right: no row qualifies (gdk_join.c:3027-3060) */
		Ctx *c = ctx();
		ok = cuda_ok(cudaMemcpyAsync(c->slot_h->v, flags + JF_RNIL, 4, cudaMemcpyDeviceToHost, c->stream), "read nil flag") && stream_sync();
		empty = ok && (unsigned) c->slot_h->v[0] != 0;
	}
	b200_return rc = B200_FAIL;
	if (ok && empty) {
		rc = join_nomatch(r1p, r2p, lci, 0);
	} else if (ok) {
		switch (t) {
		case B200_bte: ok = join_typed<bte>(lb, rb, lci, rci, mode, h, flags, r1p, r2p, func); break;
		case B200_sht: ok = join_typed<sht>(lb, rb, lci, rci, mode, h, flags, r1p, r2p, func); break;
		case B200_int: ok = join_typed<int>(lb, rb, lci, rci, mode, h, flags, r1p, r2p, func); break;
		case B200_lng: ok = join_typed<lng>(lb, rb, lci, rci, mode, h, flags, r1p, r2p, func); break;
		case B200_hge: ok = join_typed<hge>(lb, rb, lci, rci, mode, h, flags, r1p, r2p, func); break;
		case B200_flt: ok = join_typed<flt>(lb, rb, lci, rci, mode, h, flags, r1p, r2p, func); break;
		case B200_dbl: ok = join_typed<dbl>(lb, rb, lci, rci, mode, h, flags, r1p, r2p, func); break;
		default: ok = false;
		}
		rc = ok ? B200_SUCCEED : B200_FAIL;
	}
	dev_free(flags);
	if (owned)
		dev_free(h.slots);
	return rc;
}

extern "C" b200_return b200_BATjoin(b200_bat **r1p, b200_bat **r2p, const b200_bat *l, const b200_bat *r,
				    const b200_bat *sl, const b200_bat *sr, bool nil_matches, uint64_t estimate)
{
	(void) estimate;	/* a size hint for the reference's growing result; sizes are exact here */
	return join_driver(r1p, r2p, l, r, sl, sr, nil_matches ? JM_NIL_MATCHES : 0, "BATjoin");
}

/* gdk_join.c:4320: same pairs as BATjoin, r1 guaranteed in l's order (ours always is) */
extern "C" b200_return b200_BATleftjoin(b200_bat **r1p, b200_bat **r2p, const b200_bat *l, const b200_bat *r,
					const b200_bat *sl, const b200_bat *sr, bool nil_matches, uint64_t estimate)
{
	(void) estimate;
	return join_driver(r1p, r2p, l, r, sl, sr, nil_matches ? JM_NIL_MATCHES : 0, "BATleftjoin");
}

/* gdk_join.c:4334: nil_on_miss; match_one = exactly one match per left row or GDKerror */
extern "C" b200_return b200_BATouterjoin(b200_bat **r1p, b200_bat **r2p, const b200_bat *l, const b200_bat *r,
					 const b200_bat *sl, const b200_bat *sr, bool nil_matches, bool match_one, uint64_t estimate)
{
	(void) estimate;
	if (r2p == nullptr) {
		set_err("BATouterjoin: r2p required\n");	/* assert(!nil_on_miss || r2p != NULL), gdk_join.c:4065 */
		return B200_FAIL;
	}
	return join_driver(r1p, r2p, l, r, sl, sr,
			   (nil_matches ? JM_NIL_MATCHES : 0) | JM_NIL_ON_MISS | (match_one ? JM_MAX_ONE | JM_MIN_ONE : 0), "BATouterjoin");
}

/* gdk_join.c:4347: at most one pair per left row */
extern "C" b200_return b200_BATsemijoin(b200_bat **r1p, b200_bat **r2p, const b200_bat *l, const b200_bat *r,
					const b200_bat *sl, const b200_bat *sr, bool nil_matches, bool max_one, uint64_t estimate)
{
	(void) estimate;
	return join_driver(r1p, r2p, l, r, sl, sr, (nil_matches ? JM_NIL_MATCHES : 0) | JM_SEMI | (max_one ? JM_MAX_ONE : 0), "BATsemijoin");
}

/* gdk_join.c:4378: candidate list of the rows of l whose value occurs in r */
extern "C" b200_bat *b200_BATintersect(const b200_bat *l, const b200_bat *r, const b200_bat *sl, const b200_bat *sr,
				       bool nil_matches, bool max_one, uint64_t estimate)
{
	(void) estimate;
	b200_bat *bn = nullptr;
	if (join_driver(&bn, nullptr, l, r, sl, sr, (nil_matches ? JM_NIL_MATCHES : 0) | JM_SEMI | (max_one ? JM_MAX_ONE : 0),
			"BATintersect") != B200_SUCCEED)
		return nullptr;
	return bn;
}

/* gdk_join.c:4395: candidate list of the rows of l whose value does not occur in r (anti-semi-join);
 * not_in = SQL's NOT IN: a nil on either side never qualifies */
extern "C" b200_bat *b200_BATdiff(const b200_bat *l, const b200_bat *r, const b200_bat *sl, const b200_bat *sr,
				  bool nil_matches, bool not_in, uint64_t estimate)
{
	(void) estimate;
	b200_bat *bn = nullptr;
	if (join_driver(&bn, nullptr, l, r, sl, sr, (nil_matches ? JM_NIL_MATCHES : 0) | JM_ONLY_MISSES | (not_in ? JM_NOT_IN : 0),
			"BATdiff") != B200_SUCCEED)
		return nullptr;
	return bn;
}
