// noalloc.cu — b200_select_lng_noalloc: a range select on a lng column that ALLOCATES NOTHING.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of
// the ballot-mask compaction behind b200_BATselect (select.cu / compact.cuh: mask, scan, expand), whose scratch —
// one mask bit per row, one count and one offset per 2048-row chunk — normally comes from the device heap tier.
// Here the scratch is the file-scope device array below: B200_NOALLOC_MAX_ROWS = 2^24 rows at most, about
// 2.1 MiB, held by every GPU that loads the library until the process exits.
//
//   * the caller passes the input column and both outputs (device pointers); nothing is returned in new memory;
//   * `noalloc_scratch` names the array for DEVICE code.  In host code that name is only a host-side stand-in,
//     not a device address, so the host never passes it to a kernel: at set-up, once per device, with that
//     device current, cudaGetSymbolAddress gives the real address, which is what the kernels receive;
//   * no cudaMalloc / cudaFree / cudaMallocAsync / pool call, no thread context (ctx() would create a stream and
//     pinned slots): three kernel launches on the caller's stream and nothing else.  The first call on a device
//     does the cudaGetSymbolAddress set-up and therefore has to happen outside stream capture;
//   * every call shares the one array, which keeps the previous call's contents.  Nothing in it has to start at
//     zero: each call overwrites every mask word and every chunk count of its own n before reading them, and
//     never reads beyond them.  Calls that overlap on the GPU (two streams, two threads) RACE on the array —
//     serialise them (one stream, or events).
#include "common.cuh"
#include <mutex>

#define NA_MAX_ROWS ((uint64_t) 1 << 24)
#define NA_CHUNK 2048u			/* rows per CTA */
#define NA_THREADS 256u
#define NA_WORDS_PER_CHUNK (NA_CHUNK / 32)
#define NA_MAX_CHUNKS (NA_MAX_ROWS / NA_CHUNK)	/* 8192 */
#define NA_MAX_DEVICES 64

struct NoallocScratch {
	uint32_t words[NA_MAX_ROWS / 32];	/* the select's bitmask: 2 MiB */
	uint32_t counts[NA_MAX_CHUNKS];		/* selected rows per chunk */
	uint64_t offsets[NA_MAX_CHUNKS];	/* exclusive prefix of counts */
};

static __device__ __align__(256) NoallocScratch noalloc_scratch;

// (1) one CTA per chunk: test 2048 rows, the warp ballot IS the mask word; count the chunk
__global__ void __launch_bounds__(NA_THREADS)
na_mask_kernel(const lng *__restrict__ vals, uint64_t n, lng lo, lng hi, NoallocScratch *__restrict__ s)
{
	__shared__ unsigned wcnt[NA_THREADS / 32];
	const uint64_t base = (uint64_t) blockIdx.x * NA_CHUNK;
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	unsigned cnt = 0;
#pragma unroll
	for (unsigned it = 0; it < NA_CHUNK / NA_THREADS; it++) {
		const uint64_t r = base + it * NA_THREADS + threadIdx.x;
		bool hit = false;
		if (r < n) {
			const lng v = ld_stream(vals + r);
			hit = !Atom<lng>::is_nil(v) && v >= lo && v <= hi;
		}
		const unsigned w = __ballot_sync(0xffffffffu, hit);
		if (lane == 0) {
			s->words[(base >> 5) + it * (NA_THREADS / 32) + warp] = w;	/* words past n are written as 0 */
			cnt += __popc(w);
		}
	}
	if (lane == 0)
		wcnt[warp] = cnt;
	__syncthreads();
	if (threadIdx.x == 0) {
		unsigned t = 0;
		for (unsigned i = 0; i < NA_THREADS / 32; i++)
			t += wcnt[i];
		s->counts[blockIdx.x] = t;
	}
}

// (2) one CTA: exclusive scan of the chunk counts (at most 8192), total -> *count_out
__global__ void __launch_bounds__(1024)
na_scan_kernel(NoallocScratch *__restrict__ s, unsigned nchunks, unsigned long long *__restrict__ count_out)
{
	__shared__ unsigned long long wsum[32];
	__shared__ unsigned long long carry;
	const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x == 0)
		carry = 0;
	__syncthreads();
	for (unsigned b = 0; b < nchunks; b += 1024) {
		const unsigned i = b + threadIdx.x;
		const unsigned long long c = i < nchunks ? s->counts[i] : 0;
		unsigned long long inc = c;
		for (int d = 1; d < 32; d <<= 1) {
			const unsigned long long t = __shfl_up_sync(0xffffffffu, inc, d);
			if (lane >= d)
				inc += t;
		}
		if (lane == 31)
			wsum[warp] = inc;
		__syncthreads();
		if (warp == 0) {
			const unsigned long long w = wsum[lane];
			unsigned long long winc = w;
			for (int d = 1; d < 32; d <<= 1) {
				const unsigned long long t = __shfl_up_sync(0xffffffffu, winc, d);
				if (lane >= d)
					winc += t;
			}
			wsum[lane] = winc - w;
		}
		__syncthreads();
		const unsigned long long excl = carry + wsum[warp] + inc - c;
		if (i < nchunks)
			s->offsets[i] = excl;
		__syncthreads();
		if (threadIdx.x == 1023)
			carry = excl + c;
		__syncthreads();
	}
	if (threadIdx.x == 0)
		*count_out = carry;
}

// (3) one CTA per chunk: every set bit becomes its oid, in row order
__global__ void __launch_bounds__(NA_THREADS)
na_expand_kernel(const NoallocScratch *__restrict__ s, uint64_t n, oid hseq, oid *__restrict__ out)
{
	__shared__ unsigned w[NA_WORDS_PER_CHUNK];
	__shared__ unsigned pre[NA_WORDS_PER_CHUNK];
	if (s->counts[blockIdx.x] == 0)
		return;
	const uint64_t base = (uint64_t) blockIdx.x * NA_CHUNK;
	if (threadIdx.x < NA_WORDS_PER_CHUNK)
		w[threadIdx.x] = s->words[(base >> 5) + threadIdx.x];
	__syncthreads();
	if (threadIdx.x == 0) {
		unsigned run = 0;
		for (unsigned i = 0; i < NA_WORDS_PER_CHUNK; i++) {
			pre[i] = run;
			run += __popc(w[i]);
		}
	}
	__syncthreads();
	const uint64_t off = s->offsets[blockIdx.x];
#pragma unroll
	for (unsigned it = 0; it < NA_CHUNK / NA_THREADS; it++) {
		const unsigned r = it * NA_THREADS + threadIdx.x;
		const unsigned word = w[r >> 5], bit = r & 31;
		if ((word >> bit) & 1u)
			out[off + pre[r >> 5] + __popc(word & ((1u << bit) - 1u))] = hseq + base + r;
	}
	(void) n;
}

// set-up, once per device, with that device current: the device address of the scratch array
static NoallocScratch *na_scratch_for_current_device(void)
{
	static std::mutex mu;
	static NoallocScratch *ptr[NA_MAX_DEVICES] = {nullptr};
	int dev = -1;
	if (!cuda_ok(cudaGetDevice(&dev), "cudaGetDevice"))
		return nullptr;
	if (dev < 0 || dev >= NA_MAX_DEVICES) {
		set_err("b200_select_lng_noalloc: device ordinal %d out of range\n", dev);
		return nullptr;
	}
	std::lock_guard<std::mutex> lk(mu);
	if (ptr[dev] == nullptr) {
		void *p = nullptr;
		if (!cuda_ok(cudaGetSymbolAddress(&p, noalloc_scratch), "cudaGetSymbolAddress(noalloc_scratch)") || p == nullptr)
			return nullptr;
		ptr[dev] = (NoallocScratch *) p;
	}
	return ptr[dev];
}

extern "C" int b200_select_lng_noalloc(const int64_t *vals, uint64_t n, uint64_t hseqbase, int64_t lo, int64_t hi, uint64_t *oids_out,
				       uint64_t *count_out, void *stream)
{
	if (n > NA_MAX_ROWS) {
		set_err("b200_select_lng_noalloc: %llu rows exceed the static scratch (B200_NOALLOC_MAX_ROWS = %llu); use b200_BATselect\n",
			(unsigned long long) n, (unsigned long long) NA_MAX_ROWS);
		return -2;
	}
	if (count_out == nullptr || (n > 0 && (vals == nullptr || oids_out == nullptr))) {
		set_err("b200_select_lng_noalloc: input column and both outputs required\n");
		return -1;
	}
	if (lo == Atom<lng>::nil() || hi == Atom<lng>::nil()) {
		set_err("b200_select_lng_noalloc: nil bounds are not supported (closed range lo <= v <= hi only)\n");
		return -1;
	}
	NoallocScratch *s = na_scratch_for_current_device();
	if (s == nullptr)
		return -3;
	cudaStream_t st = (cudaStream_t) stream;
	const unsigned nchunks = (unsigned) ((n + NA_CHUNK - 1) / NA_CHUNK);
	if (nchunks > 0) {
		na_mask_kernel<<<nchunks, NA_THREADS, 0, st>>>((const lng *) vals, n, (lng) lo, (lng) hi, s);
		COUNT_LAUNCH();
	}
	na_scan_kernel<<<1, 1024, 0, st>>>(s, nchunks, (unsigned long long *) count_out);
	COUNT_LAUNCH();
	if (nchunks > 0) {
		na_expand_kernel<<<nchunks, NA_THREADS, 0, st>>>(s, n, (oid) hseqbase, (oid *) oids_out);
		COUNT_LAUNCH();
	}
	return cuda_ok(cudaGetLastError(), "b200_select_lng_noalloc launch") ? 0 : -3;
}
