// cand.cu — candidate lists on the device (gdk_cand.h:16-70, gdk_cand.c).
// canditer_init's normal form (dense range or ascending oid array, clipped to
// the BAT's oid range, gdk_cand.c:407-712), device expansion of the negative and
// bitmask encodings (gdk_cand.c:1296-1602), the candidate set algebra
// BATmergecand/BATintersectcand/BATdiffcand (gdk_cand.c:46,184,259) and the
// host<->device staging of BAT heaps.
#include "common.cuh"
#include <vector>
#include "compact.cuh"
#include <stdlib.h>

// ---------------------------------------------------------------- staging
// Optional per-thread cache of staged host heaps, keyed by (host base, bytes): a host column
// that several operators of one query touch (l_discount in Q6: select, then project) crosses
// PCIe once.  The owner of the host heaps decides how long a staged copy stays valid:
// b200_stage_cache(1) switches the cache on, b200_stage_cache_clear() drops every copy (the
// reference binding calls it when a heap is written or freed; bench.py calls it at the start
// of every e2e step so each step still uploads each input column).
struct StageEntry { const void *host; size_t bytes; void *dev; };
static thread_local std::vector<StageEntry> *tl_stage = nullptr;
static thread_local bool tl_stage_on = false;

extern "C" void b200_stage_cache_clear(void)
{
	if (tl_stage) {
		for (StageEntry &e : *tl_stage)
			dev_free(e.dev);
		tl_stage->clear();
	}
}

// the thread rebinds to another device (b200_init): the copies belong to the old one
void stage_cache_retire(void)
{
	b200_stage_cache_clear();
}

// the owner rewrote one host heap (b200_BATinvalidate)
void stage_cache_forget(const void *host)
{
	if (tl_stage == nullptr)
		return;
	for (size_t i = 0; i < tl_stage->size();) {
		if ((*tl_stage)[i].host == host) {
			dev_free((*tl_stage)[i].dev);
			tl_stage->erase(tl_stage->begin() + i);
		} else {
			i++;
		}
	}
}

extern "C" void b200_stage_cache(int on)
{
	tl_stage_on = on != 0;
	if (!on)
		b200_stage_cache_clear();
}

// A host heap that an operator will only touch sparsely (a gather through a short candidate
// list) is not staged at all when it lives in page-locked, device-mapped host memory: the
// kernel dereferences it over PCIe and moves one 32-64 byte atom per touched row instead of the
// whole column.  Falls back to stage() for pageable memory, dense access or a cached copy.
static unsigned sparse_div()
{
	/* a gather touching fewer than count / div rows reads the pinned host heap in place */
	static const unsigned d = [] {
		const char *e = getenv("B200_SPARSE_DIV");
		const int v = e ? atoi(e) : 0;
		return (unsigned) (v > 0 ? v : 16);
	}();
	return d;
}

bool stage_sparse(const b200_bat *in, Staged &out, uint64_t touch_rows)
{
	if (in != nullptr && in->device < 0 && in->base != nullptr && in->ccand == B200_CAND_NONE && in->ttype != B200_void &&
	    in->ttype != B200_msk && touch_rows < in->count / sparse_div() && ctx()->device >= 0) {
		bool cached = false;
		if (tl_stage_on && tl_stage)
			for (StageEntry &e : *tl_stage)
				cached |= e.host == in->base;
		cudaPointerAttributes at;
		if (!cached && cudaPointerGetAttributes(&at, in->base) == cudaSuccess && at.type == cudaMemoryTypeHost &&
		    at.devicePointer != nullptr) {
			b200_bat *t = bat_alloc_desc();
			if (t == nullptr)
				return false;
			*t = *in;
			t->device = ctx()->device;
			t->owns = 0;
			t->base = at.devicePointer;
			out.tmp = t;
			out.b = t;
			return true;
		}
		cudaGetLastError();	/* pageable memory: attributes query may have set an error */
	}
	return stage(in, out);
}

bool stage(const b200_bat *in, Staged &out)
{
	out.b = in;
	if (in == nullptr)
		return true;
	if (in->device >= 0) {
		order_after(in);	/* produced on another thread's stream without a sync? wait for it */
		return true;
	}
	Ctx *c = ctx();
	if (c->device < 0) {
		set_err("b200: no usable CUDA device; this library has no CPU fallback\n");
		return false;
	}
	if (tl_stage_on && in->base != nullptr && in->ttype != B200_void && in->ttype != B200_msk && in->ccand == B200_CAND_NONE) {
		const size_t bytes = (size_t) in->count * atomsize(in->ttype);
		if (tl_stage == nullptr)
			tl_stage = new std::vector<StageEntry>();
		void *dev = nullptr;
		for (StageEntry &e : *tl_stage)
			if (e.host == in->base && e.bytes == bytes)
				dev = e.dev;
		if (dev == nullptr) {
			dev = dev_alloc(bytes);
			if (dev == nullptr ||
			    !cuda_ok(cudaMemcpyAsync(dev, in->base, bytes, cudaMemcpyHostToDevice, c->stream), "stage BAT heap"))
				return false;
			tl_stage->push_back(StageEntry{in->base, bytes, dev});
		}
		b200_bat *t = bat_alloc_desc();
		if (t == nullptr)
			return false;
		*t = *in;
		t->device = c->device;
		t->owns = 0;		/* the cache owns the device copy */
		t->base = dev;
		t->capacity = in->count;
		out.tmp = t;
		out.b = t;
		return true;
	}
	b200_bat *t = bat_alloc_desc();
	if (t == nullptr)
		return false;
	*t = *in;
	t->device = c->device;
	t->owns = 1;
	t->base = nullptr;
	t->vheap = nullptr;
	out.tmp = t;
	if (in->base != nullptr && in->ttype != B200_void) {
		size_t bytes = in->ttype == B200_msk ? ((in->count + 31) / 32) * 4 : (size_t) in->count * atomsize(in->ttype);
		t->base = dev_alloc(bytes);
		if (t->base == nullptr ||
		    !cuda_ok(cudaMemcpyAsync(t->base, in->base, bytes, cudaMemcpyHostToDevice, c->stream), "stage BAT heap"))
			return false;
		t->capacity = in->count;
	}
	if (in->vheap != nullptr && in->ccand != B200_CAND_NONE) {
		size_t bytes = in->ccand == B200_CAND_NEGOID ? in->nvheap * 8 : in->nvheap * 4;
		t->vheap = dev_alloc(bytes);
		if (t->vheap == nullptr ||
		    !cuda_ok(cudaMemcpyAsync(t->vheap, in->vheap, bytes, cudaMemcpyHostToDevice, c->stream), "stage cand heap"))
			return false;
	}
	out.b = t;
	return true;
}

// ---------------------------------------------------------------- expansion kernels
// negative candidates: the i-th survivor of [seq, ...) minus the sorted exceptions
__global__ void negcand_expand_kernel(oid seq, const oid *__restrict__ exc, uint64_t nexc, uint64_t n, oid *__restrict__ dst)
{
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
		/* smallest k with exc[k] - seq - k > i  (k exceptions precede survivor i) */
		uint64_t lo = 0, hi = nexc;
		while (lo < hi) {
			uint64_t mid = (lo + hi) >> 1;
			if (exc[mid] - seq - mid > i)
				hi = mid;
			else
				lo = mid + 1;
		}
		dst[i] = seq + i + lo;
	}
}

struct MaskBitF {
	const uint32_t *words;
	oid base;		/* oid of bit 0 */
	__device__ bool operator()(uint64_t idx, oid &o) const
	{
		o = base + idx;
		return (words[idx >> 5] >> (idx & 31)) & 1u;
	}
	__device__ oid at(uint64_t idx) const { return base + idx; }
};

// first index with oids[idx] >= lo, first index with oids[idx] >= hi, plus the
// first/last oid of the clipped range
__global__ void oid_bounds_kernel(const oid *__restrict__ oids, uint64_t n, oid lo, oid hi, unsigned long long *slot)
{
	uint64_t a = 0, b = n;
	while (a < b) {
		uint64_t m = (a + b) >> 1;
		if (oids[m] < lo) a = m + 1; else b = m;
	}
	uint64_t first = a;
	b = n;
	while (a < b) {
		uint64_t m = (a + b) >> 1;
		if (oids[m] < hi) a = m + 1; else b = m;
	}
	slot[0] = first;
	slot[1] = a;
	slot[2] = first < a ? oids[first] : 0;
	slot[3] = first < a ? oids[a - 1] : 0;
}

// ---------------------------------------------------------------- cand_init
bool cand_init(Cand &ci, const b200_bat *b, const b200_bat *s)
{
	Ctx *c = ctx();
	const oid blo = b ? b->hseqbase : 0;
	const oid bhi = b ? b->hseqbase + b->count : ~(oid) 0;
	ci.hseq = blo;
	ci.rhseq = blo;
	ci.dense = true;
	ci.oids = nullptr;
	if (s == nullptr) {
		ci.seq = blo;
		ci.ncand = b ? b->count : 0;
		return true;
	}
	/* ci->hseq of the reference: s->hseqbase + leading candidates clipped away
	 * (gdk_cand.c:440-457,711); 0 when nothing overlaps (:603-611) */
	ci.rhseq = s->hseqbase;
	if (s->count == 0 || (b != nullptr && b->count == 0)) {
		ci.seq = blo;
		ci.ncand = 0;
		return true;
	}
	const oid *oids = nullptr;
	uint64_t n = s->count;
	bool range_known = false;
	oid first = 0, last = 0;
	if (s->ccand == B200_CAND_NEGOID) {
		oid *tmp = (oid *) dev_alloc(n * sizeof(oid));
		if (tmp == nullptr)
			return false;
		ci.owned = tmp;
		negcand_expand_kernel<<<grid_for(n, 256), 256, 0, c->stream>>>(s->tseqbase, (const oid *) s->vheap, s->nvheap, n, tmp);
		COUNT_LAUNCH();
		oids = tmp;
	} else if (s->ccand == B200_CAND_MSK) {
		oid *tmp = (oid *) dev_alloc(n * sizeof(oid));
		if (tmp == nullptr)
			return false;
		ci.owned = tmp;
		MaskBitF f{(const uint32_t *) s->vheap, s->tseqbase - s->firstbit};
		CompactResult r;
		if (!run_compact<8>(s->nvheap * 32, f, tmp, r))
			return false;
		n = r.count;
		oids = tmp;
		first = r.first;
		last = r.last;
		range_known = true;
	} else if (s->ttype == B200_void || s->base == nullptr) {
		if (s->tseqbase == OID_NIL) {
			ci.seq = blo;
			ci.ncand = 0;
			return true;
		}
		oid lo = s->tseqbase > blo ? s->tseqbase : blo;
		oid hi = s->tseqbase + n < bhi ? s->tseqbase + n : bhi;
		ci.seq = lo;
		ci.ncand = hi > lo ? hi - lo : 0;
		if (ci.ncand == 0) {
			ci.seq = blo;
			ci.rhseq = 0;
		} else
			ci.rhseq = s->hseqbase + (lo - s->tseqbase);
		return true;
	} else {
		if (basetype(s->ttype) != B200_lng) {
			set_err("b200: candidate list must be an oid or void BAT\n");
			return false;
		}
		oids = (const oid *) s->base;
		if (s->oidrange_known) {
			first = s->first_oid;
			last = s->last_oid;
			range_known = true;
		}
	}
	if (n == 0) {
		ci.seq = blo;
		ci.ncand = 0;
		return true;
	}
	if (!range_known || first < blo || last >= bhi) {
		oid_bounds_kernel<<<1, 1, 0, c->stream>>>(oids, n, blo, bhi, c->slot_d->v);
		COUNT_LAUNCH();
		if (!cuda_ok(cudaGetLastError(), "bounds launch") || !stream_sync())
			return false;
		uint64_t lo = c->slot_h->v[0], hi = c->slot_h->v[1];
		first = c->slot_h->v[2];
		last = c->slot_h->v[3];
		oids += lo;
		n = hi - lo;
		if (s->ccand != B200_CAND_MSK)
			ci.rhseq = s->hseqbase + lo;
	}
	if (n == 0) {
		ci.seq = blo;
		ci.ncand = 0;
		ci.rhseq = 0;
		return true;
	}
	ci.ncand = n;
	if (last - first + 1 == n) {	/* the oids are consecutive: demote to dense (gdk_cand.c:498-520) */
		ci.dense = true;
		ci.seq = first;
		return true;
	}
	ci.dense = false;
	ci.seq = first;
	ci.oids = oids;
	return true;
}

b200_bat *cand_to_bat(const Cand &ci)
{
	if (ci.dense || ci.ncand == 0)
		return bat_dense(0, ci.ncand ? ci.seq : 0, ci.ncand);
	oid *dst = (oid *) dev_alloc(ci.ncand * sizeof(oid));
	if (dst == nullptr)
		return nullptr;
	if (!cuda_ok(cudaMemcpyAsync(dst, ci.oids, ci.ncand * sizeof(oid), cudaMemcpyDeviceToDevice, ctx()->stream), "copy cands")) {
		dev_free(dst);
		return nullptr;
	}
	b200_bat *bn = bat_alloc_desc();
	if (bn == nullptr) {
		dev_free(dst);
		return nullptr;
	}
	bn->ttype = B200_oid;
	bn->base = dst;
	bn->capacity = bn->count = ci.ncand;
	bn->owns = 1;
	bn->tsorted = bn->tkey = bn->tnonil = 1;
	bn->trevsorted = ci.ncand <= 1;
	return bn;
}

// BATunmask (gdk_cand.c:1464): a mask (or negative) candidate list as its plain form — a dense
// void BAT when the oids are consecutive, else a materialised ascending oid[] tail.
extern "C" b200_bat *b200_BATunmask(const b200_bat *b)
{
	b200_clrerr();
	if (b == nullptr) {
		set_err("BATunmask: BAT required\n");
		return nullptr;
	}
	Staged sb;
	if (!stage(b, sb))
		return nullptr;
	Cand ci;
	if (!cand_init(ci, nullptr, sb.b))
		return nullptr;
	b200_bat *bn = cand_to_bat(ci);
	if (bn)
		bn->hseqbase = b->hseqbase;
	return bn;
}

// ---------------------------------------------------------------- download
extern "C" int b200_BATtohost(const b200_bat *b, void *dst)
{
	b200_clrerr();
	Ctx *c = ctx();
	if (b->count == 0)
		return 0;
	if (b->ttype == B200_void || b->ccand != B200_CAND_NONE) {
		oid *o = (oid *) dst;
		if (b->ccand == B200_CAND_NONE) {
			for (uint64_t i = 0; i < b->count; i++)
				o[i] = b->tseqbase == OID_NIL ? OID_NIL : b->tseqbase + i;
			return 0;
		}
		Staged sb;
		if (!stage(b, sb))
			return -1;
		Cand ci;
		if (!cand_init(ci, nullptr, sb.b))
			return -1;
		if (ci.dense) {
			for (uint64_t i = 0; i < ci.ncand; i++)
				o[i] = ci.seq + i;
			return 0;
		}
		if (!cuda_ok(cudaMemcpyAsync(dst, ci.oids, ci.ncand * sizeof(oid), cudaMemcpyDeviceToHost, c->stream), "download") ||
		    !stream_sync())
			return -1;
		return 0;
	}
	size_t bytes = b->ttype == B200_msk ? ((b->count + 31) / 32) * 4 : (size_t) b->count * atomsize(b->ttype);
	if (b->ttype == B200_STR_TYPE)
		bytes = (size_t) b->count * b->pad;	/* the offsets, in their stored width (str.cu) */
	if (b->device < 0) {
		memcpy(dst, b->base, bytes);
		return 0;
	}
	if (!cuda_ok(cudaMemcpyAsync(dst, b->base, bytes, cudaMemcpyDeviceToHost, c->stream), "download") || !stream_sync())
		return -1;
	return 0;
}

// ---------------------------------------------------------------- set algebra
// Both inputs are candidate lists (sorted, unique).  Each element of one list is
// looked up in the other by binary search and the survivors are compacted in
// order; merge additionally interleaves the two lists by rank.
struct Lookup {
	bool dense;
	oid seq;
	const oid *oids;
	uint64_t n;
	__device__ bool contains(oid o) const
	{
		if (dense)
			return o >= seq && o < seq + n;
		uint64_t a = 0, b = n;
		while (a < b) {
			uint64_t m = (a + b) >> 1;
			if (oids[m] < o) a = m + 1; else b = m;
		}
		return a < n && oids[a] == o;
	}
	__device__ oid at(uint64_t i) const { return dense ? seq + i : oids[i]; }
	/* number of elements < o */
	__device__ uint64_t rank(oid o) const
	{
		if (dense)
			return o <= seq ? 0 : (o - seq < n ? o - seq : n);
		uint64_t a = 0, b = n;
		while (a < b) {
			uint64_t m = (a + b) >> 1;
			if (oids[m] < o) a = m + 1; else b = m;
		}
		return a;
	}
};
struct SetOpF {
	Lookup a, b;
	bool want;	/* keep a[i] iff (a[i] in b) == want */
	__device__ bool operator()(uint64_t idx, oid &o) const
	{
		o = a.at(idx);
		return b.contains(o) == want;
	}
	__device__ oid at(uint64_t idx) const { return a.at(idx); }
};
static Lookup lookup_of(const Cand &c) { return Lookup{c.dense, c.seq, c.oids, c.ncand}; }

static b200_bat *setop(const b200_bat *a, const b200_bat *b, bool want)
{
	Staged sa, sb;
	if (!stage(a, sa) || !stage(b, sb))
		return nullptr;
	Cand ca, cb;
	if (!cand_init(ca, nullptr, sa.b) || !cand_init(cb, nullptr, sb.b))
		return nullptr;
	if (ca.ncand == 0)
		return bat_dense(0, 0, 0);
	if (cb.ncand == 0)
		return want ? bat_dense(0, 0, 0) : cand_to_bat(ca);
	oid *dst = (oid *) dev_alloc(ca.ncand * sizeof(oid));
	if (dst == nullptr)
		return nullptr;
	SetOpF f{lookup_of(ca), lookup_of(cb), want};
	CompactResult r;
	if (!run_compact<8>(ca.ncand, f, dst, r)) {
		dev_free(dst);
		return nullptr;
	}
	return oidlist_to_bat(dst, ca.ncand, r);
}

extern "C" b200_bat *b200_BATintersectcand(const b200_bat *a, const b200_bat *b)
{
	b200_clrerr();
	ctx()->algo = "intersectcand: cuda binary-search compaction";
	return setop(a, b, true);
}

extern "C" b200_bat *b200_BATdiffcand(const b200_bat *a, const b200_bat *b)
{
	b200_clrerr();
	ctx()->algo = "diffcand: cuda binary-search compaction";
	return setop(a, b, false);
}

__global__ void firstlast_kernel(const oid *dst, uint64_t n, unsigned long long *slot)
{
	slot[0] = n;
	slot[1] = n ? dst[0] : 0;
	slot[2] = n ? dst[n - 1] : 0;
}

// merge: out position of a[i] = i + |{b < a[i]}|; of b[j] (not in a) = j + |{a < b[j]}| - dup-adjust
__global__ void mergecand_kernel(Lookup a, Lookup b, oid *__restrict__ dst, const oid *__restrict__ bonly, uint64_t nbonly)
{
	uint64_t total = a.n + nbonly;
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t) gridDim.x * blockDim.x) {
		if (i < a.n) {
			oid o = a.at(i);
			/* elements of b\a smaller than o */
			uint64_t lo = 0, hi = nbonly;
			while (lo < hi) {
				uint64_t m = (lo + hi) >> 1;
				if (bonly[m] < o) lo = m + 1; else hi = m;
			}
			dst[i + lo] = o;
		} else {
			uint64_t j = i - a.n;
			oid o = bonly[j];
			dst[j + a.rank(o)] = o;
		}
	}
}

extern "C" b200_bat *b200_BATmergecand(const b200_bat *a, const b200_bat *b)
{
	b200_clrerr();
	ctx()->algo = "mergecand: cuda rank merge";
	Staged sa, sb;
	if (!stage(a, sa) || !stage(b, sb))
		return nullptr;
	Cand ca, cb;
	if (!cand_init(ca, nullptr, sa.b) || !cand_init(cb, nullptr, sb.b))
		return nullptr;
	if (ca.ncand == 0)
		return cand_to_bat(cb);
	if (cb.ncand == 0)
		return cand_to_bat(ca);
	/* b \ a, in order */
	oid *bonly = (oid *) dev_alloc(cb.ncand * sizeof(oid));
	if (bonly == nullptr)
		return nullptr;
	SetOpF f{lookup_of(cb), lookup_of(ca), false};
	CompactResult r;
	if (!run_compact<8>(cb.ncand, f, bonly, r)) {
		dev_free(bonly);
		return nullptr;
	}
	uint64_t total = ca.ncand + r.count;
	oid *dst = (oid *) dev_alloc(total * sizeof(oid));
	if (dst == nullptr) {
		dev_free(bonly);
		return nullptr;
	}
	mergecand_kernel<<<grid_for(total, 256), 256, 0, ctx()->stream>>>(lookup_of(ca), lookup_of(cb), dst, bonly, r.count);
	COUNT_LAUNCH();
	dev_free(bonly);
	if (!cuda_ok(cudaGetLastError(), "mergecand launch"))
		return nullptr;
	firstlast_kernel<<<1, 1, 0, ctx()->stream>>>(dst, total, ctx()->slot_d->v);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "firstlast launch") || !stream_sync()) {
		dev_free(dst);
		return nullptr;
	}
	CompactResult m;
	m.count = total;
	m.first = ctx()->slot_h->v[1];
	m.last = ctx()->slot_h->v[2];
	return oidlist_to_bat(dst, total, m);
}
