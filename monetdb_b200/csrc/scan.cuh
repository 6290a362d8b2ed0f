// scan.cuh — single-pass ordered tile prefix ("decoupled look-back") shared by
// every compaction/scan kernel on the path (select -> oid lists, mask expansion,
// join output offsets, partition sizing).  Tiles take their id from a global
// ticket counter, so a tile only ever waits on tiles that are already running:
// forward progress does not depend on the hardware's CTA dispatch order.
#pragma once
#include "common.cuh"

#define TS_FLAG_A (1ULL << 62)			/* tile aggregate available */
#define TS_FLAG_P (2ULL << 62)			/* inclusive prefix available */
#define TS_VAL(x) ((x) & ((1ULL << 62) - 1))

struct TileScan {
	unsigned long long *status;		/* ntiles entries, zeroed before launch */
	unsigned int *ticket;			/* zeroed before launch */
};

// scratch layout: [ticket (16 B)] [status ntiles x 8 B]
static inline size_t tilescan_bytes(uint64_t ntiles) { return 16 + ntiles * 8; }
static inline bool tilescan_prepare(TileScan &ts, uint64_t ntiles)
{
	char *p = (char *) scratch_get(tilescan_bytes(ntiles));
	if (p == nullptr)
		return false;
	ts.ticket = (unsigned int *) p;
	ts.status = (unsigned long long *) (p + 16);
	return cuda_ok(cudaMemsetAsync(p, 0, tilescan_bytes(ntiles), ctx()->stream), "memset tile state");
}

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long *p)
{
	unsigned long long v;
	asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long *p, unsigned long long v)
{
	asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// block-wide: take the next tile id
__device__ __forceinline__ unsigned tile_ticket(const TileScan &ts, unsigned *sh_slot)
{
	if (threadIdx.x == 0)
		*sh_slot = atomicAdd(ts.ticket, 1u);
	__syncthreads();
	unsigned t = *sh_slot;
	__syncthreads();
	return t;
}

// block-wide: publish this tile's total and return the sum of all earlier tiles.
// Must be called by every thread of the block (contains __syncthreads).
__device__ __forceinline__ unsigned long long
tile_exclusive_prefix(const TileScan &ts, unsigned tile, unsigned long long total, unsigned long long *sh_slot)
{
	if (threadIdx.x < 32) {
		const unsigned lane = threadIdx.x;
		unsigned long long excl = 0;
		if (tile == 0) {
			if (lane == 0)
				st_volatile_u64(&ts.status[0], TS_FLAG_P | total);
		} else {
			if (lane == 0)
				st_volatile_u64(&ts.status[tile], TS_FLAG_A | total);
			long long base = (long long) tile - 1;
			for (;;) {
				long long idx = base - lane;
				unsigned long long s = idx >= 0 ? ld_volatile_u64(&ts.status[idx]) : TS_FLAG_P;
				unsigned flag = (unsigned) (s >> 62);
				unsigned inval = __ballot_sync(0xffffffffu, flag == 0);
				unsigned pmask = __ballot_sync(0xffffffffu, flag == 2);
				unsigned firstP = pmask ? (unsigned) __ffs(pmask) - 1 : 32;
				unsigned need = firstP >= 31 ? 0xffffffffu : ((2u << firstP) - 1);
				if (inval & need)
					continue;	/* a needed predecessor has not published yet */
				unsigned long long v = (lane <= firstP) ? TS_VAL(s) : 0;
#pragma unroll
				for (int o = 16; o > 0; o >>= 1)
					v += __shfl_down_sync(0xffffffffu, v, o);
				excl += v;
				if (firstP != 32)
					break;
				base -= 32;
			}
			if (lane == 0)
				st_volatile_u64(&ts.status[tile], TS_FLAG_P | (excl + total));
		}
		if (lane == 0)
			*sh_slot = excl;
	}
	__syncthreads();
	unsigned long long r = *sh_slot;
	__syncthreads();
	return r;
}
