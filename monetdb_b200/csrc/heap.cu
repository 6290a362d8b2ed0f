// heap.cu — runtime of the B200 GDK path: per-thread CUDA context, the
// device-resident BAT heap tier (the tier the north star adds beside
// HEAPalloc/HEAPfree, gdk_heap.c:141-225,626), BAT descriptor life cycle
// (COLnew gdk_bat.c:238-298, BATdense :298, BBPreclaim gdk_bbp.c:3161), and the
// candidate iterator set-up (canditer_init, gdk_cand.c:407-712).
#include "common.cuh"
#include <unordered_map>
#include "scan.cuh"
#include <mutex>
#include <vector>
#include <map>
#include <iterator>
#include <stdlib.h>

std::atomic<unsigned long long> g_launches{0};
static thread_local Ctx *tl_ctx = nullptr;
static std::mutex g_pool_mu;
static bool g_pool_ready[64] = {false};
struct CachedBlock { void *p; uint32_t ctx; };
static std::mutex g_cache_mu;
static std::multimap<size_t, CachedBlock> g_cache[64];
static std::unordered_map<const void *, size_t> g_block_size;	/* live + cached blocks handed out by dev_alloc */
static size_t g_cached_bytes[64] = {0};
static size_t g_cache_limit[64] = {0};	/* block cache bound per device (see dev_alloc) */
static std::mutex g_ctx_mu;
static std::vector<Ctx *> g_ctxs;		/* every context ever created; index = id - 1 (kept for their events) */

void set_err(const char *fmt, ...)
{
	Ctx *c = tl_ctx;
	va_list ap;
	va_start(ap, fmt);
	if (c)
		vsnprintf(c->err, sizeof(c->err), fmt, ap);
	else
		vfprintf(stderr, fmt, ap);
	va_end(ap);
}

bool cuda_ok(cudaError_t e, const char *what)
{
	if (e == cudaSuccess)
		return true;
	set_err("CUDA error %s: %s (%s)\n", cudaGetErrorName(e), cudaGetErrorString(e), what);
	return false;
}

static Ctx *ctx_create(int device)
{
	Ctx *c = new Ctx();
	tl_ctx = c;
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess || n == 0) {
		snprintf(c->err, sizeof(c->err),
			 "b200: no usable CUDA device (%s); this library has no CPU fallback\n",
			 e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
		c->device = -1;
		return c;
	}
	if (device < 0 || device >= n) {
		snprintf(c->err, sizeof(c->err), "b200: invalid device %d (have %d)\n", device, n);
		c->device = -1;
		return c;
	}
	if (!cuda_ok(cudaSetDevice(device), "cudaSetDevice")) {
		c->device = -1;
		return c;
	}
	c->device = device;
	{
		std::lock_guard<std::mutex> lk(g_pool_mu);
		if (!g_pool_ready[device]) {
			cudaMemPool_t pool;
			if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
				uint64_t thr = UINT64_MAX;	/* keep freed heaps cached in the pool */
				cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
			}
			/* B200_L2_FETCH = 32 | 64 | 128: cudaLimitMaxL2FetchGranularity (a hint; the sparse gathers of
			 * the candidate-list selects move one DRAM atom per touched row) */
			if (const char *e = getenv("B200_L2_FETCH")) {
				const int g = atoi(e);
				if (g == 32 || g == 64 || g == 128)
					cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t) g);
			}
			size_t fr = 0, tot = 0;
			cudaMemGetInfo(&fr, &tot);
			g_cache_limit[device] = tot / 10 * 6;
			if (const char *e = getenv("B200_HEAP_CACHE_GB"))
				g_cache_limit[device] = (size_t) atoll(e) << 30;
			g_pool_ready[device] = true;
		}
	}
	if (!cuda_ok(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking), "stream create") ||
	    !cuda_ok(cudaHostAlloc((void **) &c->slot_h, sizeof(ResultSlot), cudaHostAllocMapped), "slot alloc") ||
	    !cuda_ok(cudaHostGetDevicePointer((void **) &c->slot_d, c->slot_h, 0), "slot map")) {
		c->device = -1;
		return c;
	}
	memset(c->slot_h, 0, sizeof(ResultSlot));
	if (!cuda_ok(cudaEventCreateWithFlags(&c->ev, cudaEventDisableTiming), "event create")) {
		c->device = -1;
		return c;
	}
	{
		std::lock_guard<std::mutex> lk(g_ctx_mu);
		g_ctxs.push_back(c);
		c->id = (uint32_t) g_ctxs.size();
	}
	return c;
}

// the calling thread rebinds to another device: give back what the old context held.  The struct and its
// event stay registered because BATs it produced may still name it as their producer.
static void ctx_retire(Ctx *c)
{
	if (c == nullptr || c->device < 0)
		return;
	cudaSetDevice(c->device);
	cudaStreamSynchronize(c->stream);
	if (c->scratch) {
		{
			std::lock_guard<std::mutex> lk(g_cache_mu);
			g_block_size.erase(c->scratch);
		}
		cudaFree(c->scratch);
	}
	c->scratch = nullptr;
	c->scratch_bytes = 0;
	if (c->slot_h)
		cudaFreeHost(c->slot_h);
	c->slot_h = c->slot_d = nullptr;
	cudaStreamDestroy(c->stream);
	c->stream = nullptr;
	c->device = -1;
}

// Every descriptor is tagged with the context that made it (bat_alloc_desc).  A consumer on another
// thread records an event at the producer stream's current tail — everything that wrote the BAT was
// submitted before the BAT was handed over — and orders its own stream behind it.
static void order_after_ctx(uint32_t producer)
{
	Ctx *c = tl_ctx;
	if (producer == 0 || c == nullptr || c->device < 0 || producer == c->id)
		return;
	std::lock_guard<std::mutex> lk(g_ctx_mu);
	if (producer > g_ctxs.size())
		return;
	Ctx *p = g_ctxs[producer - 1];
	if (p->stream == nullptr || p->device != c->device)
		return;		/* retired (it synchronised on the way out) or another device */
	if (cudaEventRecord(c->ev, p->stream) == cudaSuccess)
		cudaStreamWaitEvent(c->stream, c->ev, 0);
}

void order_after(const b200_bat *b)
{
	if (b != nullptr)
		order_after_ctx(b->producer);
}

Ctx *ctx()
{
	if (tl_ctx == nullptr)
		ctx_create(0);
	return tl_ctx;
}

static bool ctx_usable()
{
	Ctx *c = ctx();
	if (c->device < 0) {
		if (c->err[0] == 0)
			snprintf(c->err, sizeof(c->err), "b200: no usable CUDA device; this library has no CPU fallback\n");
		return false;
	}
	return true;
}

// ---------------------------------------------------------------- the device heap tier's block cache
// cudaMallocAsync keeps freed memory in its pool, but a query plan whose intermediates are multi-GB blocks of
// several sizes makes the pool map and unmap physical memory again and again (measured on Q1 SF100: 8 ms of a
// 40 ms step; single BATgroup / BATcalcmul calls stalled 20-30 ms in the allocator).  A plan allocates the same
// sizes every time it runs, so freed heaps are kept in a size-keyed cache instead and handed out again without
// a driver call — GDK's own HEAPcache idea (gdk_heap.c:60-139) applied to device heaps.  Reuse is safe in stream
// order; a block freed by one thread context and reused by another orders the new owner's stream behind the
// old one's tail first.  The cache is trimmed when it exceeds B200_HEAP_CACHE_GB (default: 60 % of the device)
// and emptied when an allocation fails.

static size_t round_block(size_t bytes)
{
	if (bytes < 512)
		return 512;
	if (bytes < ((size_t) 1 << 20))
		return (bytes + 511) & ~(size_t) 511;
	return (bytes + ((size_t) 2 << 20) - 1) & ~(((size_t) 2 << 20) - 1);
}

static void order_after_ctx(uint32_t producer);

// give cached blocks back to the driver until at most `keep` bytes stay cached (caller holds no lock)
static void cache_trim(int device, size_t keep)
{
	std::vector<void *> drop;
	{
		std::lock_guard<std::mutex> lk(g_cache_mu);
		auto &m = g_cache[device];
		while (g_cached_bytes[device] > keep && !m.empty()) {
			auto it = std::prev(m.end());		/* largest first */
			g_cached_bytes[device] -= it->first;
			g_block_size.erase(it->second.p);
			drop.push_back(it->second.p);
			m.erase(it);
		}
	}
	if (drop.empty())
		return;
	/* blocks may have been used last on any stream of the device: settle everything, then free */
	cudaDeviceSynchronize();
	for (void *p : drop)
		cudaFree(p);
}

void *dev_alloc(size_t bytes)
{
	if (!ctx_usable())
		return nullptr;
	Ctx *c = ctx();
	const size_t size = round_block(bytes ? bytes : 16);
	{
		std::unique_lock<std::mutex> lk(g_cache_mu);
		auto &m = g_cache[c->device];
		auto it = m.lower_bound(size);
		if (it != m.end() && it->first <= size + size / 8) {
			CachedBlock b = it->second;
			g_cached_bytes[c->device] -= it->first;
			m.erase(it);
			lk.unlock();
			if (b.ctx != c->id)
				order_after_ctx(b.ctx);
			return b.p;
		}
	}
	void *p = nullptr;
	cudaError_t e = cudaMallocAsync(&p, size, c->stream);
	if (e != cudaSuccess) {
		cudaGetLastError();
		cache_trim(c->device, 0);		/* out of memory: the cache gives everything back, then retry */
		e = cudaMallocAsync(&p, size, c->stream);
	}
	if (!cuda_ok(e, "device heap alloc"))
		return nullptr;
	std::lock_guard<std::mutex> lk(g_cache_mu);
	g_block_size[p] = size;
	return p;
}

void dev_free(void *p)
{
	if (p == nullptr)
		return;
	Ctx *c = tl_ctx;
	const bool have_ctx = c && c->device >= 0;
	size_t size = 0;
	bool over = false;
	if (have_ctx) {
		std::lock_guard<std::mutex> lk(g_cache_mu);
		auto it = g_block_size.find(p);
		if (it != g_block_size.end()) {
			size = it->second;
			g_cache[c->device].insert({size, CachedBlock{p, c->id}});
			g_cached_bytes[c->device] += size;
			over = g_cached_bytes[c->device] > g_cache_limit[c->device];
		}
	}
	if (size != 0) {
		if (over)
			cache_trim(c->device, g_cache_limit[c->device] / 2);
		return;
	}
	{
		std::lock_guard<std::mutex> lk(g_cache_mu);
		g_block_size.erase(p);
	}
	if (have_ctx)
		cudaFreeAsync(p, c->stream);
	else
		cudaFree(p);		/* a thread without a context (or after ctx_retire): synchronous release */
}

extern "C" void b200_heap_cache_trim(void)
{
	if (tl_ctx && tl_ctx->device >= 0)
		cache_trim(tl_ctx->device, 0);
}

extern "C" uint64_t b200_heap_cache_bytes(void)
{
	if (tl_ctx == nullptr || tl_ctx->device < 0)
		return 0;
	std::lock_guard<std::mutex> lk(g_cache_mu);
	return g_cached_bytes[tl_ctx->device];
}

void *scratch_get(size_t bytes)
{
	Ctx *c = ctx();
	if (c->device < 0)
		return nullptr;
	if (bytes > c->scratch_bytes) {
		if (c->scratch)
			dev_free(c->scratch);
		size_t nb = bytes < (1u << 20) ? (1u << 20) : bytes * 2;
		c->scratch_bytes = 0;
		c->scratch = dev_alloc(nb);
		if (c->scratch == nullptr)
			return nullptr;
		c->scratch_bytes = nb;
	}
	return c->scratch;
}

// hand the scratch block over to the caller (who frees it with dev_free); the next scratch_get allocates afresh
void *scratch_steal(size_t *bytes)
{
	Ctx *c = ctx();
	void *p = c->scratch;
	if (bytes)
		*bytes = c->scratch_bytes;
	c->scratch = nullptr;
	c->scratch_bytes = 0;
	return p;
}

size_t scratch_size()
{
	return ctx()->scratch_bytes;
}

bool stream_sync()
{
	return cuda_ok(cudaStreamSynchronize(ctx()->stream), "stream sync");
}

// ---------------------------------------------------------------- C-ABI: runtime
extern "C" int b200_init(int device)
{
	if (tl_ctx != nullptr && tl_ctx->device == device)
		return 0;
	if (tl_ctx != nullptr) {
		stage_cache_retire();
		ctx_retire(tl_ctx);
	}
	Ctx *c = ctx_create(device);
	return c->device >= 0 ? 0 : -1;
}

extern "C" int b200_device_count(void)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess)
		return 0;
	return n;
}

extern "C" const char *b200_errbuf(void) { return tl_ctx ? tl_ctx->err : ""; }
extern "C" void b200_clrerr(void) { if (tl_ctx) tl_ctx->err[0] = 0; }
extern "C" const char *b200_algorithm(void) { return tl_ctx ? tl_ctx->algo : ""; }
extern "C" uint64_t b200_kernel_launches(void) { return g_launches.load(); }
extern "C" void b200_sync(void) { if (tl_ctx && tl_ctx->device >= 0) cudaStreamSynchronize(tl_ctx->stream); }
extern "C" void *b200_stream(void) { return ctx()->stream; }

extern "C" void *b200_pinned_alloc(size_t bytes)
{
	if (!ctx_usable())
		return nullptr;
	void *p = nullptr;
	if (!cuda_ok(cudaHostAlloc(&p, bytes ? bytes : 16, cudaHostAllocDefault), "pinned alloc"))
		return nullptr;
	return p;
}
extern "C" void b200_pinned_free(void *p) { if (p) cudaFreeHost(p); }

// ---------------------------------------------------------------- descriptors
b200_bat *bat_alloc_desc()
{
	b200_bat *b = (b200_bat *) calloc(1, sizeof(b200_bat));
	if (b == nullptr) {
		set_err("b200: out of host memory for a BAT descriptor\n");
		return nullptr;
	}
	b->tseqbase = OID_NIL;
	b->device = tl_ctx ? tl_ctx->device : -1;
	b->producer = tl_ctx ? tl_ctx->id : 0;
	return b;
}

b200_bat *bat_new(uint64_t hseq, int tt, uint64_t cap)
{
	if (!ctx_usable())
		return nullptr;
	int w = atomsize(tt);
	if (w < 0 && tt != B200_msk) {
		set_err("b200: COLnew: unsupported tail type %d (fixed-width atoms only)\n", tt);
		return nullptr;
	}
	b200_bat *b = bat_alloc_desc();
	if (b == nullptr)
		return nullptr;
	b->ttype = tt;
	b->hseqbase = hseq;
	b->capacity = cap;
	b->owns = 1;
	if (tt == B200_void) {
		b->tseqbase = 0;
		b->tsorted = b->tkey = b->tnonil = 1;
		return b;
	}
	size_t bytes = tt == B200_msk ? ((cap + 31) / 32) * 4 : (size_t) cap * w;
	b->base = dev_alloc(bytes);
	if (b->base == nullptr) {
		free(b);
		return nullptr;
	}
	/* an empty BAT is trivially sorted/key (BATsetcount, gdk_bat.c:2066) */
	b->tsorted = b->trevsorted = b->tkey = b->tnonil = 1;
	return b;
}

b200_bat *bat_dense(uint64_t hseq, uint64_t tseq, uint64_t cnt)
{
	if (!ctx_usable())
		return nullptr;
	b200_bat *b = bat_alloc_desc();
	if (b == nullptr)
		return nullptr;
	b->ttype = B200_void;
	b->hseqbase = hseq;
	b->tseqbase = tseq;
	b->count = cnt;
	b->tsorted = b->tkey = b->tnonil = 1;
	b->trevsorted = cnt <= 1;
	b->owns = 1;
	b->oidrange_known = cnt > 0;
	b->first_oid = tseq;
	b->last_oid = tseq + cnt - 1;
	return b;
}

extern "C" b200_bat *b200_COLnew(uint64_t hseq, int tt, uint64_t cap) { b200_clrerr(); return bat_new(hseq, tt, cap); }
extern "C" b200_bat *b200_BATdense(uint64_t hseq, uint64_t tseq, uint64_t cnt) { b200_clrerr(); return bat_dense(hseq, tseq, cnt); }

// BATextend / HEAPextend (gdk_bat.c:529, gdk_heap.c:246): grow the tail heap of a device BAT we own,
// keeping the first count items; a capacity that is already large enough is left alone.
extern "C" b200_return b200_BATextend(b200_bat *b, uint64_t newcap)
{
	b200_clrerr();
	if (b == nullptr || !ctx_usable())
		return B200_FAIL;
	if (newcap <= b->capacity || b->ttype == B200_void)
		return B200_SUCCEED;
	if (!b->owns || b->device < 0 || b->device != tl_ctx->device) {
		set_err("BATextend: only device heaps owned by the descriptor (on the calling thread's device) can grow\n");
		return B200_FAIL;
	}
	const int w = atomsize(b->ttype);
	const size_t nbytes = b->ttype == B200_msk ? ((newcap + 31) / 32) * 4 : (size_t) newcap * w;
	const size_t obytes = b->ttype == B200_msk ? ((b->count + 31) / 32) * 4 : (size_t) b->count * w;
	void *nb = dev_alloc(nbytes);
	if (nb == nullptr)
		return B200_FAIL;
	if (obytes && !cuda_ok(cudaMemcpyAsync(nb, b->base, obytes, cudaMemcpyDeviceToDevice, tl_ctx->stream), "BATextend copy")) {
		dev_free(nb);
		return B200_FAIL;
	}
	if (b->base) {
		hash_cache_drop(b->base);
		shadow_drop(b->base);
		dev_free(b->base);		/* stream-ordered: runs after the copy */
	}
	b->base = nb;
	b->capacity = newcap;
	return B200_SUCCEED;
}

extern "C" b200_bat *b200_BATfromhost(const void *src, int tt, uint64_t cnt, uint64_t hseq)
{
	b200_clrerr();
	b200_bat *b = bat_new(hseq, tt, cnt);
	if (b == nullptr)
		return nullptr;
	size_t bytes = tt == B200_msk ? ((cnt + 31) / 32) * 4 : (size_t) cnt * atomsize(tt);
	if (bytes && !cuda_ok(cudaMemcpyAsync(b->base, src, bytes, cudaMemcpyHostToDevice, ctx()->stream), "upload")) {
		b200_BATfree(b);
		return nullptr;
	}
	b->count = cnt;
	b->tsorted = b->trevsorted = b->tkey = cnt <= 1;
	b->tnonil = cnt == 0;
	b->tnil = 0;
	if (basetype(tt) == B200_lng && tt == B200_oid && cnt > 0) {
		/* callers that upload candidate lists tell us sortedness via BATsetprops */
	}
	/* the source buffer may be pageable: be done with it before returning */
	if (!stream_sync()) {
		b200_BATfree(b);
		return nullptr;
	}
	return b;
}

extern "C" b200_bat *b200_BATwrap(void *devptr, int tt, uint64_t cnt, uint64_t hseq)
{
	b200_clrerr();
	if (!ctx_usable())
		return nullptr;
	b200_bat *b = bat_alloc_desc();
	if (b == nullptr)
		return nullptr;
	b->ttype = tt;
	b->hseqbase = hseq;
	b->count = b->capacity = cnt;
	b->base = devptr;
	b->owns = 0;
	b->tsorted = b->trevsorted = b->tkey = cnt <= 1;
	b->tnonil = cnt == 0;
	return b;
}

extern "C" void b200_BATsetprops(b200_bat *b, bool sorted, bool revsorted, bool key, bool nonil, bool nil)
{
	b->tsorted = sorted || b->count <= 1;
	b->trevsorted = revsorted || b->count <= 1;
	b->tkey = key || b->count <= 1;
	b->tnonil = nonil;
	b->tnil = nil;
}

// ---------------------------------------------------------------- narrow twins (common.cuh)
static std::mutex g_shadow_mu;
static std::unordered_multimap<const void *, Shadow> g_shadows;	/* one entry per (heap, kind) */

bool shadow_enabled()
{
	static const bool on = [] {
		const char *e = getenv("B200_NO_SHADOW");
		return !(e && e[0] == '1');
	}();
	return on;
}

void shadow_put(const void *base, const Shadow &s)
{
	Shadow old;
	{
		std::lock_guard<std::mutex> lk(g_shadow_mu);
		auto r = g_shadows.equal_range(base);
		for (auto it = r.first; it != r.second; ++it)
			if (it->second.kind == s.kind) {
				old = it->second;
				g_shadows.erase(it);
				break;
			}
		g_shadows.emplace(base, s);
	}
	dev_free(old.p);
	dev_free(old.p2);
}

bool shadow_get(const void *base, int kind, uint64_t n, Shadow &out)
{
	if (base == nullptr)
		return false;
	std::lock_guard<std::mutex> lk(g_shadow_mu);
	auto r = g_shadows.equal_range(base);
	for (auto it = r.first; it != r.second; ++it)
		if (it->second.kind == kind && it->second.n == n) {
			out = it->second;
			return true;
		}
	return false;
}

void shadow_set_a2(const void *base, int kind, uint64_t n, uint64_t a2)
{
	std::lock_guard<std::mutex> lk(g_shadow_mu);
	auto r = g_shadows.equal_range(base);
	for (auto it = r.first; it != r.second; ++it)
		if (it->second.kind == kind && it->second.n == n)
			it->second.a2 = a2;
}

void shadow_drop(const void *base)
{
	if (base == nullptr)
		return;
	Shadow gone[8];
	int k = 0;
	{
		std::lock_guard<std::mutex> lk(g_shadow_mu);
		auto r = g_shadows.equal_range(base);
		for (auto it = r.first; it != r.second && k < 8; ++it)
			gone[k++] = it->second;
		g_shadows.erase(base);
	}
	for (int i = 0; i < k; i++) {
		dev_free(gone[i].p);
		dev_free(gone[i].p2);
	}
}

// ---------------------------------------------------------------- heap share counts (views)
struct HeapShare { uint64_t views = 0; bool orphan = false; };
static std::mutex g_share_mu;
static std::unordered_map<const void *, HeapShare> g_shares;

void heap_share(const void *root)
{
	std::lock_guard<std::mutex> lk(g_share_mu);
	g_shares[root].views++;
}

bool heap_unshare(const void *root)
{
	std::lock_guard<std::mutex> lk(g_share_mu);
	auto it = g_shares.find(root);
	if (it == g_shares.end())
		return false;
	if (--it->second.views > 0)
		return false;
	const bool orphan = it->second.orphan;
	g_shares.erase(it);
	return orphan;
}

bool heap_owner_release(const void *root)
{
	std::lock_guard<std::mutex> lk(g_share_mu);
	auto it = g_shares.find(root);
	if (it == g_shares.end())
		return true;
	it->second.orphan = true;
	return false;
}

static void heap_release(void *p, int device)
{
	if (p == nullptr)
		return;
	if (tl_ctx && tl_ctx->device == device) {
		dev_free(p);
	} else {
		/* freed from a thread bound to another device (or to none): synchronous path */
		{
			std::lock_guard<std::mutex> lk(g_cache_mu);
			g_block_size.erase(p);
		}
		cudaFree(p);
	}
}

extern "C" void b200_BATfree(b200_bat *b)
{
	if (b == nullptr)
		return;
	if (b->device >= 0)
		order_after(b);		/* a free on this thread's stream must not overtake the producer's kernels */
	if (b->device >= 0 && b->base != nullptr && b->owns) {
		/* what the library cached about this heap dies with its owner — never with a view or a
		 * wrapped alias (owns == 0), which would discard the parent's BAThash index */
		hash_cache_drop(b->base);
		shadow_drop(b->base);
	}
	if (b->parent != nullptr) {
		if (heap_unshare(b->parent))
			heap_release(b->parent, b->device);
		if (b->ttype == B200_STR_TYPE && b->device >= 0)
			heap_release(b->vheap, b->device);	/* a str view shares the offsets, its string heap is its own copy */
	} else if (b->owns) {
		if (b->device >= 0) {
			if (b->base == nullptr || heap_owner_release(b->base))
				heap_release(b->base, b->device);
			heap_release(b->vheap, b->device);
		} else {
			free(b->base);
			free(b->vheap);
		}
	}
	free(b);
}

extern "C" void b200_BATinvalidate(const b200_bat *b)
{
	if (b == nullptr || b->base == nullptr)
		return;
	if (b->device >= 0) {
		hash_cache_drop(b->base);
		shadow_drop(b->base);
	} else {
		stage_cache_forget(b->base);
	}
}

extern "C" b200_bat *b200_BATnegcands_fromhost(uint64_t seq, uint64_t cnt, const b200_oid *exceptions, uint64_t nexc)
{
	b200_clrerr();
	/* BATnegcands (gdk_cand.c:1296): only exceptions inside [seq, seq+cnt) count */
	uint64_t lo = 0, hi = nexc;
	while (lo < nexc && exceptions[lo] < seq) lo++;
	hi = lo;
	while (hi < nexc && exceptions[hi] < seq + cnt) hi++;
	b200_bat *b = bat_dense(0, seq, cnt);
	if (b == nullptr || lo == hi)
		return b;
	if (hi - lo == cnt) {
		b->count = 0;
		return b;
	}
	b->vheap = dev_alloc((hi - lo) * sizeof(oid));
	if (b->vheap == nullptr ||
	    !cuda_ok(cudaMemcpyAsync(b->vheap, exceptions + lo, (hi - lo) * sizeof(oid), cudaMemcpyHostToDevice, ctx()->stream), "upload") ||
	    !stream_sync()) {
		b200_BATfree(b);
		return nullptr;
	}
	b->ccand = B200_CAND_NEGOID;
	b->nvheap = hi - lo;
	b->count = cnt - (hi - lo);
	b->oidrange_known = 0;
	return b;
}

extern "C" b200_bat *b200_BATmaskedcands_fromhost(uint64_t hseq, uint64_t nrows, const uint32_t *words, bool selected)
{
	b200_clrerr();
	/* BATmaskedcands (gdk_cand.c:1366) for a mask of exactly nrows bits */
	uint64_t nmask = (nrows + 31) / 32;
	uint32_t *w = (uint32_t *) malloc(nmask * 4 + 4);
	if (w == nullptr) {
		set_err("b200: out of host memory\n");
		return nullptr;
	}
	for (uint64_t i = 0; i < nmask; i++)
		w[i] = selected ? words[i] : ~words[i];
	if (nrows % 32)
		w[nmask - 1] &= (1U << (nrows % 32)) - 1;
	uint64_t cnt = 0, firstbit = 0;
	for (uint64_t i = 0; i < nmask; i++) {
		if (cnt == 0 && w[i] != 0)
			firstbit = __builtin_ctz(w[i]) + i * 32;
		cnt += __builtin_popcount(w[i]);
	}
	b200_bat *b = bat_dense(hseq, hseq, 0);
	if (b == nullptr || cnt == 0) {
		free(w);
		return b;
	}
	b->vheap = dev_alloc(nmask * 4);
	if (b->vheap == nullptr ||
	    !cuda_ok(cudaMemcpyAsync(b->vheap, w, nmask * 4, cudaMemcpyHostToDevice, ctx()->stream), "upload") ||
	    !stream_sync()) {
		free(w);
		b200_BATfree(b);
		return nullptr;
	}
	free(w);
	b->ccand = B200_CAND_MSK;
	b->nvheap = nmask;
	b->firstbit = (uint32_t) firstbit;
	b->tseqbase = hseq + firstbit;
	b->count = cnt;
	b->oidrange_known = 0;
	return b;
}

// ---------------------------------------------------------------- BATconstant
void atom_nil_bits(int t, void *out)
{
	switch (basetype(t)) {
	case B200_bte: { bte v = Atom<bte>::nil(); memcpy(out, &v, 1); break; }
	case B200_sht: { sht v = Atom<sht>::nil(); memcpy(out, &v, 2); break; }
	case B200_int: { int v = Atom<int>::nil(); memcpy(out, &v, 4); break; }
	case B200_lng: { lng v = Atom<lng>::nil(); memcpy(out, &v, 8); break; }
	case B200_hge: { hge v = Atom<hge>::nil(); memcpy(out, &v, 16); break; }
	case B200_flt: { flt v = NAN; memcpy(out, &v, 4); break; }
	case B200_dbl: { dbl v = NAN; memcpy(out, &v, 8); break; }
	}
}

__global__ void fill16_kernel(uint4 *dst, uint64_t n16, uint4 v)
{
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (uint64_t) gridDim.x * blockDim.x)
		dst[i] = v;
}

b200_bat *bat_constant(uint64_t hseq, int tt, const void *val, uint64_t n)
{
	unsigned char pat[16];
	int w = atomsize(tt == B200_void ? B200_oid : tt);
	if (w <= 0) {
		set_err("b200: BATconstant: unsupported type %d\n", tt);
		return nullptr;
	}
	if (val)
		memcpy(pat, val, w);
	else
		atom_nil_bits(tt == B200_void ? B200_oid : tt, pat);
	unsigned char nilpat[16];
	atom_nil_bits(tt == B200_void ? B200_oid : tt, nilpat);
	bool isnil = memcmp(pat, nilpat, w) == 0;
	if (tt == B200_flt) { flt f; memcpy(&f, pat, 4); isnil = f != f; }
	if (tt == B200_dbl) { dbl d; memcpy(&d, pat, 8); isnil = d != d; }
	if ((tt == B200_oid || tt == B200_void) && (n == 0 || isnil)) {
		b200_bat *bn = bat_dense(hseq, n ? OID_NIL : 0, n);
		if (bn && n) {
			bn->tnonil = 0;
			bn->tnil = 1;
			bn->tkey = n <= 1;
			bn->trevsorted = 1;
			bn->oidrange_known = 0;
		}
		return bn;
	}
	b200_bat *bn = bat_new(hseq, tt, n);
	if (bn == nullptr)
		return nullptr;
	for (int i = w; i < 16; i++)
		pat[i] = pat[i % w];
	uint4 v;
	memcpy(&v, pat, 16);
	uint64_t n16 = ((uint64_t) n * w + 15) / 16;
	/* the pool hands out >= 16-byte granules, so rounding the fill up is safe */
	if (n16) {
		fill16_kernel<<<grid_for(n16, 256), 256, 0, ctx()->stream>>>((uint4 *) bn->base, n16, v);
		COUNT_LAUNCH();
	}
	bn->count = n;
	bn->tnonil = n == 0 || !isnil;
	bn->tnil = n > 0 && isnil;
	bn->tsorted = bn->trevsorted = 1;
	bn->tkey = n <= 1;
	return bn;
}

// ---------------------------------------------------------------- kernel timing
#include <string>
struct KRec { const char *name; uint64_t bytes; cudaEvent_t a, b; };
static thread_local std::vector<KRec> *tl_krecs = nullptr;
static std::atomic<int> g_timing{0};

KTimer::KTimer(const char *name, uint64_t bytes)
{
	if (name == nullptr || !g_timing.load(std::memory_order_relaxed) || tl_ctx == nullptr || tl_ctx->device < 0)
		return;
	if (tl_krecs == nullptr)
		tl_krecs = new std::vector<KRec>();
	KRec r{name, bytes, nullptr, nullptr};
	if (cudaEventCreate(&r.a) != cudaSuccess || cudaEventCreate(&r.b) != cudaSuccess)
		return;
	cudaEventRecord(r.a, tl_ctx->stream);
	tl_krecs->push_back(r);
	idx = (int) tl_krecs->size() - 1;
}
void KTimer::stop()
{
	if (idx >= 0 && !stopped)
		cudaEventRecord((*tl_krecs)[idx].b, tl_ctx->stream);
	stopped = true;
}
void KTimer::add_bytes(uint64_t more)
{
	if (idx >= 0 && tl_krecs && (size_t) idx < tl_krecs->size())
		(*tl_krecs)[idx].bytes += more;
}

extern "C" void b200_timing_enable(int on)
{
	g_timing.store(on);
}

extern "C" int b200_timing_read(char *buf, size_t bufsize)
{
	if (buf == nullptr || bufsize == 0)
		return -1;
	buf[0] = 0;
	if (tl_krecs == nullptr || tl_ctx == nullptr)
		return 0;
	cudaStreamSynchronize(tl_ctx->stream);
	struct Tot { std::string name; uint64_t n = 0, bytes = 0; double ms = 0; };
	std::vector<Tot> tots;
	for (KRec &r : *tl_krecs) {
		float ms = 0;
		cudaEventElapsedTime(&ms, r.a, r.b);
		cudaEventDestroy(r.a);
		cudaEventDestroy(r.b);
		Tot *t = nullptr;
		for (Tot &x : tots)
			if (x.name == r.name)
				t = &x;
		if (t == nullptr) {
			tots.push_back(Tot());
			t = &tots.back();
			t->name = r.name;
		}
		t->n++;
		t->bytes += r.bytes;
		t->ms += ms;
	}
	tl_krecs->clear();
	size_t off = 0;
	for (Tot &t : tots) {
		int w = snprintf(buf + off, bufsize - off, "%s %llu %.6f %llu\n", t.name.c_str(), (unsigned long long) t.n, t.ms,
				 (unsigned long long) t.bytes);
		if (w < 0 || (size_t) w >= bufsize - off)
			break;
		off += w;
	}
	return (int) tots.size();
}

extern "C" b200_bat *b200_BAThostwrap(const void *hostptr, int tt, uint64_t cnt, uint64_t hseq)
{
	b200_clrerr();
	b200_bat *b = bat_alloc_desc();
	if (b == nullptr)
		return nullptr;
	b->ttype = tt;
	b->device = -1;
	b->hseqbase = hseq;
	b->count = b->capacity = cnt;
	b->base = (void *) hostptr;
	b->owns = 0;
	b->tsorted = b->trevsorted = b->tkey = cnt <= 1;
	b->tnonil = cnt == 0;
	return b;
}
