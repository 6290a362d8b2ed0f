// partition.cu — hash repartitioning of (key, global oid) tuples for the multi-GPU hash join
// and high-cardinality GROUP BY ("partition tuples by hash(key) mod G into G send
// buffers, NCCL all-to-all, then purely local build/probe").  This is the device half of the one
// exchange step of the path; the reference has no counterpart because mitosis never repartitions
// (opt_mitosis.c slices by row range only) — on one box it joins against the whole build side.
//
//   pass 1  histogram of the destination of every candidate row (warp-aggregated atomics on G
//           counters), read back by the host: these are the all-to-all split sizes;
//   pass 2  scatter: per destination a warp reserves a contiguous run with ONE atomicAdd on the
//           destination's cursor and its lanes store (key, oid) into the send buffers.
// nil keys never match in an inner join (nil_matches = false) and are dropped here unless asked for.
// Tuple order inside a destination is not deterministic; BATjoin promises no order (gdk_join.c:56-61).
#include "common.cuh"
#include "keys.cuh"

#define PT_THREADS 256
#define PT_MAXPARTS 64

template <typename T> __device__ __forceinline__ unsigned part_of(T v, unsigned nparts)
{
	typedef typename KeyBits<T>::type U;
	/* the high half of the mixed hash: independent of the low bits the local hash tables use */
	return (unsigned) ((hash_key<U>(KeyBits<T>::get(v), 0x5bd1e995ULL) >> 32) % nparts);
}

template <typename T>
__global__ void __launch_bounds__(PT_THREADS)
partition_count_kernel(ColSrc<T> src, unsigned nparts, bool keep_nil, unsigned long long *__restrict__ counts)
{
	__shared__ unsigned sh[PT_MAXPARTS];
	if (threadIdx.x < PT_MAXPARTS)
		sh[threadIdx.x] = 0;
	__syncthreads();
	for (uint64_t k = (uint64_t) blockIdx.x * PT_THREADS + threadIdx.x; k < src.n; k += (uint64_t) gridDim.x * PT_THREADS) {
		const T v = src.value(k);
		if (!keep_nil && Atom<T>::is_nil(v))
			continue;
		const unsigned d = part_of<T>(v, nparts);
		const unsigned m = __match_any_sync(__activemask(), d);
		if ((unsigned) (__ffs(m) - 1) == (threadIdx.x & 31))
			atomicAdd(&sh[d], (unsigned) __popc(m));
	}
	__syncthreads();
	if (threadIdx.x < nparts && sh[threadIdx.x])
		atomicAdd(&counts[threadIdx.x], (unsigned long long) sh[threadIdx.x]);
}

template <typename T>
__global__ void __launch_bounds__(PT_THREADS)
partition_scatter_kernel(ColSrc<T> src, unsigned nparts, bool keep_nil, unsigned long long *__restrict__ cursor /* start offsets */,
			 T *__restrict__ keys_out, oid *__restrict__ oids_out)
{
	for (uint64_t k = (uint64_t) blockIdx.x * PT_THREADS + threadIdx.x; k < src.n; k += (uint64_t) gridDim.x * PT_THREADS) {
		const T v = src.value(k);
		if (!keep_nil && Atom<T>::is_nil(v))
			continue;
		const unsigned d = part_of<T>(v, nparts);
		const unsigned m = __match_any_sync(__activemask(), d);
		const unsigned lane = threadIdx.x & 31, leader = __ffs(m) - 1;
		unsigned long long base = 0;
		if (lane == leader)
			base = atomicAdd(&cursor[d], (unsigned long long) __popc(m));
		base = __shfl_sync(m, base, leader);
		const unsigned long long pos = base + __popc(m & ((1u << lane) - 1));
		keys_out[pos] = v;
		oids_out[pos] = src.rowoid(k);
	}
}

template <typename T>
static bool partition_typed(const b200_bat *b, const Cand &ci, unsigned nparts, bool keep_nil, void *keys_out, oid *oids_out,
			    uint64_t *counts, b200_bat **keysp = nullptr, b200_bat **oidsp = nullptr)
{
	Ctx *c = ctx();
	ColSrc<T> src = col_src<T>(b, ci);
	unsigned long long *dev = (unsigned long long *) dev_alloc(2 * PT_MAXPARTS * 8);
	if (dev == nullptr || !cuda_ok(cudaMemsetAsync(dev, 0, 2 * PT_MAXPARTS * 8, c->stream), "partition memset"))
		return false;
	const unsigned grid = grid_for(ci.ncand, PT_THREADS * 4, NSM * 8);
	bool ok = true;
	if (ci.ncand > 0) {
		{
			KTimer kt("partition_count", ci.ncand * sizeof(T));
			partition_count_kernel<T><<<grid, PT_THREADS, 0, c->stream>>>(src, nparts, keep_nil, dev);
		}
		COUNT_LAUNCH();
	}
	unsigned long long h[PT_MAXPARTS];
	ok = cuda_ok(cudaGetLastError(), "partition count") &&
	     cuda_ok(cudaMemcpyAsync(c->slot_h->v, dev, (nparts < 16 ? nparts : 16) * 8, cudaMemcpyDeviceToHost, c->stream), "partition counts") &&
	     (nparts <= 16 || cuda_ok(cudaMemcpyAsync(h, dev, nparts * 8, cudaMemcpyDeviceToHost, c->stream), "partition counts")) &&
	     stream_sync();
	if (ok) {
		if (nparts <= 16)
			for (unsigned i = 0; i < nparts; i++)
				h[i] = c->slot_h->v[i];
		unsigned long long off[PT_MAXPARTS], run = 0;
		for (unsigned i = 0; i < nparts; i++) {
			counts[i] = h[i];
			off[i] = run;
			run += h[i];
		}
		if (keysp != nullptr) {
			/* b200_BATpartition: the send runs are BATs sized by the histogram just read */
			*keysp = bat_new(0, b->ttype, run);
			*oidsp = bat_new(0, B200_oid, run);
			if (*keysp == nullptr || *oidsp == nullptr) {
				dev_free(dev);
				return false;
			}
			(*keysp)->count = (*oidsp)->count = run;
			(*keysp)->tnonil = !keep_nil || b->tnonil;
			(*oidsp)->tnonil = 1;
			(*keysp)->tsorted = (*keysp)->trevsorted = (*keysp)->tkey = run <= 1;
			(*oidsp)->tsorted = (*oidsp)->trevsorted = (*oidsp)->tkey = run <= 1;
			keys_out = (*keysp)->base;
			oids_out = (oid *) (*oidsp)->base;
		}
		if (keys_out != nullptr && run > 0) {
			ok = cuda_ok(cudaMemcpyAsync(dev + PT_MAXPARTS, off, nparts * 8, cudaMemcpyHostToDevice, c->stream), "partition cursors");
			if (ok) {
				{
					KTimer kt("partition_scatter", ci.ncand * (2 * sizeof(T) + 8));
					partition_scatter_kernel<T><<<grid, PT_THREADS, 0, c->stream>>>(src, nparts, keep_nil, dev + PT_MAXPARTS,
													  (T *) keys_out, oids_out);
				}
				COUNT_LAUNCH();
				/* the cursor upload reads a host stack array: finish before it goes out of scope */
				ok = cuda_ok(cudaGetLastError(), "partition scatter") && stream_sync();
			}
		}
	}
	dev_free(dev);
	return ok;
}

// b200_hash_partition: counts[d] (d < nparts) = tuples of (b restricted to s) whose key hashes to
// destination d.  With keys_out/oids_out == NULL only the counts are computed (size the buffers, then
// call again); otherwise destination d's tuples are written at offset sum(counts[0..d)) of keys_out
// (b's atom type) and oids_out (the rows' oids: hseqbase + position, i.e. global oids when every
// rank's shard carries its global hseqbase).
extern "C" b200_return b200_hash_partition(const b200_bat *b, const b200_bat *s, int nparts, bool keep_nil, void *keys_out,
					   b200_oid *oids_out, uint64_t *counts)
{
	b200_clrerr();
	if (b == nullptr || counts == nullptr || nparts < 1 || nparts > PT_MAXPARTS) {
		set_err("b200_hash_partition: needs b, counts and 1 <= nparts <= %d\n", PT_MAXPARTS);
		return B200_FAIL;
	}
	Staged sb, ss;
	if (!stage(b, sb) || !stage(s, ss))
		return B200_FAIL;
	b = sb.b;
	Cand ci;
	if (!cand_init(ci, b, ss.b))
		return B200_FAIL;
	if (b->ttype == B200_void || b->base == nullptr) {
		set_err("b200_hash_partition: key column must be materialised\n");
		return B200_FAIL;
	}
	ctx()->algo = "partition: cuda hash scatter";
	bool ok = false;
	switch (basetype(b->ttype)) {
	case B200_bte: ok = partition_typed<bte>(b, ci, nparts, keep_nil, keys_out, oids_out, counts); break;
	case B200_sht: ok = partition_typed<sht>(b, ci, nparts, keep_nil, keys_out, oids_out, counts); break;
	case B200_int: ok = partition_typed<int>(b, ci, nparts, keep_nil, keys_out, oids_out, counts); break;
	case B200_lng: ok = partition_typed<lng>(b, ci, nparts, keep_nil, keys_out, oids_out, counts); break;
	case B200_hge: ok = partition_typed<hge>(b, ci, nparts, keep_nil, keys_out, oids_out, counts); break;
	case B200_flt: ok = partition_typed<flt>(b, ci, nparts, keep_nil, keys_out, oids_out, counts); break;
	case B200_dbl: ok = partition_typed<dbl>(b, ci, nparts, keep_nil, keys_out, oids_out, counts); break;
	default:
		set_err("b200_hash_partition: type %s is not on the device path (fixed-width atoms only)\n", atomname(b->ttype));
		return B200_FAIL;
	}
	return ok ? B200_SUCCEED : B200_FAIL;
}

// One call form of the above for the library's own exchange step (b200_comm_exchange): *keysp / *oidsp are new
// BATs (b's atom type / oid) holding the tuples of destination d at offset sum(counts[0..d)).
extern "C" b200_return b200_BATpartition(b200_bat **keysp, b200_bat **oidsp, const b200_bat *b, const b200_bat *s, int nparts,
					 bool keep_nil, uint64_t *counts)
{
	b200_clrerr();
	if (b == nullptr || counts == nullptr || keysp == nullptr || oidsp == nullptr || nparts < 1 || nparts > PT_MAXPARTS) {
		set_err("b200_BATpartition: needs b, outputs, counts and 1 <= nparts <= %d\n", PT_MAXPARTS);
		return B200_FAIL;
	}
	*keysp = *oidsp = nullptr;
	Staged sb, ss;
	if (!stage(b, sb) || !stage(s, ss))
		return B200_FAIL;
	b = sb.b;
	Cand ci;
	if (!cand_init(ci, b, ss.b))
		return B200_FAIL;
	if (b->ttype == B200_void || b->base == nullptr) {
		set_err("b200_BATpartition: key column must be materialised\n");
		return B200_FAIL;
	}
	ctx()->algo = "partition: cuda hash scatter";
	bool ok = false;
	switch (basetype(b->ttype)) {
	case B200_bte: ok = partition_typed<bte>(b, ci, nparts, keep_nil, nullptr, nullptr, counts, keysp, oidsp); break;
	case B200_sht: ok = partition_typed<sht>(b, ci, nparts, keep_nil, nullptr, nullptr, counts, keysp, oidsp); break;
	case B200_int: ok = partition_typed<int>(b, ci, nparts, keep_nil, nullptr, nullptr, counts, keysp, oidsp); break;
	case B200_lng: ok = partition_typed<lng>(b, ci, nparts, keep_nil, nullptr, nullptr, counts, keysp, oidsp); break;
	case B200_hge: ok = partition_typed<hge>(b, ci, nparts, keep_nil, nullptr, nullptr, counts, keysp, oidsp); break;
	case B200_flt: ok = partition_typed<flt>(b, ci, nparts, keep_nil, nullptr, nullptr, counts, keysp, oidsp); break;
	case B200_dbl: ok = partition_typed<dbl>(b, ci, nparts, keep_nil, nullptr, nullptr, counts, keysp, oidsp); break;
	default:
		set_err("b200_BATpartition: type %s is not on the device path (fixed-width atoms only)\n", atomname(b->ttype));
		return B200_FAIL;
	}
	if (!ok) {
		b200_BATfree(*keysp);
		b200_BATfree(*oidsp);
		*keysp = *oidsp = nullptr;
	}
	return ok ? B200_SUCCEED : B200_FAIL;
}
