// compact.cuh — ordered stream compaction into oid lists: the building block
// behind every operator that emits a candidate list (scanselect's
// `dst[cnt] = o; cnt += TEST` loop, gdk_select.c:219-248; candidate set algebra,
// gdk_cand.c:46-400; mask expansion, gdk_cand.c:1464).  One pass over the input,
// results in input order, output staged through shared memory so the global
// writes are contiguous.
#pragma once
#include "common.cuh"
#include "scan.cuh"

#define CP_THREADS 256

// Generic form: F is a functor `bool f(uint64_t idx, oid &out)` evaluated once per
// input index; hits are written to dst in index order.
template <int ITEMS, typename F>
__global__ void __launch_bounds__(CP_THREADS)
compact_kernel(uint64_t n, F f, oid *__restrict__ dst, TileScan ts, unsigned ntiles, unsigned long long *total_out)
{
	__shared__ oid stage[CP_THREADS * ITEMS];
	__shared__ unsigned wtot[CP_THREADS / 32];
	__shared__ unsigned long long slot;
	__shared__ unsigned tslot;
	const unsigned tile = tile_ticket(ts, &tslot);
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint64_t wbase = (uint64_t) tile * (CP_THREADS * ITEMS) + (uint64_t) warp * (32 * ITEMS);
	oid vals[ITEMS];
	unsigned rank[ITEMS];
	unsigned hitmask = 0, wcount = 0;
#pragma unroll
	for (int j = 0; j < ITEMS; j++) {
		uint64_t idx = wbase + j * 32 + lane;
		oid o = 0;
		bool h = idx < n && f(idx, o);
		vals[j] = o;
		unsigned bal = __ballot_sync(0xffffffffu, h);
		rank[j] = wcount + __popc(bal & ((1u << lane) - 1));
		wcount += __popc(bal);
		hitmask |= (unsigned) h << j;
	}
	if (lane == 0)
		wtot[warp] = wcount;
	__syncthreads();
	unsigned wexcl = 0, ttot = 0;
#pragma unroll
	for (int w = 0; w < CP_THREADS / 32; w++) {
		unsigned t = wtot[w];
		wexcl += w < warp ? t : 0;
		ttot += t;
	}
	unsigned long long excl = tile_exclusive_prefix(ts, tile, ttot, &slot);
#pragma unroll
	for (int j = 0; j < ITEMS; j++)
		if (hitmask & (1u << j))
			stage[wexcl + rank[j]] = vals[j];
	__syncthreads();
	for (unsigned i = threadIdx.x; i < ttot; i += CP_THREADS)
		dst[excl + i] = stage[i];
	if (tile == ntiles - 1 && threadIdx.x == 0)
		*total_out = excl + ttot;
}

// count / first / last of a freshly compacted oid list -> mapped host slot
static __global__ void compact_finish_kernel(const oid *dst, const unsigned long long *total, unsigned long long *slot)
{
	unsigned long long c = *total;
	slot[0] = c;
	slot[1] = c ? dst[0] : 0;
	slot[2] = c ? dst[c - 1] : 0;
}

struct CompactResult {
	uint64_t count = 0;
	oid first = 0, last = 0;
};

// device scratch: [total (16 B)][tile scan state]
template <int ITEMS, typename F>
static bool run_compact(uint64_t n, F f, oid *dst, CompactResult &res)
{
	Ctx *c = ctx();
	uint64_t ntiles = (n + CP_THREADS * ITEMS - 1) / (CP_THREADS * ITEMS);
	if (ntiles == 0) {
		res = CompactResult();
		return true;
	}
	char *p = (char *) scratch_get(16 + tilescan_bytes(ntiles));
	if (p == nullptr)
		return false;
	if (!cuda_ok(cudaMemsetAsync(p, 0, 16 + tilescan_bytes(ntiles), c->stream), "memset tile state"))
		return false;
	unsigned long long *total = (unsigned long long *) p;
	TileScan ts;
	ts.ticket = (unsigned int *) (p + 16);
	ts.status = (unsigned long long *) (p + 32);
	{
		KTimer kt("compact", n * 8);
		compact_kernel<ITEMS, F><<<(unsigned) ntiles, CP_THREADS, 0, c->stream>>>(n, f, dst, ts, (unsigned) ntiles, total);
	}
	COUNT_LAUNCH();
	compact_finish_kernel<<<1, 1, 0, c->stream>>>(dst, total, c->slot_d->v);
	COUNT_LAUNCH();
	if (!cuda_ok(cudaGetLastError(), "compact launch") || !stream_sync())
		return false;
	res.count = c->slot_h->v[0];
	res.first = c->slot_h->v[1];
	res.last = c->slot_h->v[2];
	return true;
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
