// io.cu — the data-format edges of the device heap tier: how columns get ONTO the tier
// without passing through a host BAT first.
//
//   b200_bincopy_decode / b200_bincopy_encode
//       the fixed-width records of binary COPY INTO / COPY INTO BINARY (sql/backends/monet5/sql_bincopyconvert.c:
//       type table :590-616, byteswap_* :61-165, decode_date/time/timestamp :168-280, encode_* :186-330,
//       validate_bit :31-42, VALIDATE_DECIMAL :45-58; record layouts common/utils/copybinary.h:15-33).  The raw
//       records are uploaded as they come off the wire and converted by one kernel: byte swapping for foreign
//       endianness, date_create / daytime_create / timestamp_create (gdk_time.c:62,281,401) for the temporal
//       records — an impossible date or time becomes nil exactly as there —, the boolean and decimal range
//       validations with the reference's SQLSTATE texts.
//   b200_HEAPload / b200_HEAPsave
//       a tail heap file as GDK keeps it on disk (the raw array of `count` values, gdk_heap.c:729 HEAPload,
//       :860 HEAPsave -> GDKsave) read straight into / written straight out of device memory through two
//       page-locked staging buffers, the read of chunk i+1 overlapping the copy of chunk i.
// Variable-width records (str / blob loaders) stay with the reference: they build the string heap.
#include "common.cuh"
#include <type_traits>
#include <errno.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

// ---------------------------------------------------------------- record layouts (copybinary.h:15-33)
struct BinDate {
	uint8_t day, month;
	int16_t year;
};
struct BinTime {
	uint32_t ms;	/* microseconds, in spite of the name */
	uint8_t seconds, minutes, hours, padding;
};
struct BinTimestamp {
	BinTime time;
	BinDate date;
};
static_assert(sizeof(BinDate) == 4 && sizeof(BinTime) == 8 && sizeof(BinTimestamp) == 12, "copybinary.h layouts");

namespace bincopy {
constexpr int YEAR_MIN = -4712, YEAR_OFFSET = 4712, MONTH_WIDTH = 21, DAY_WIDTH = 5;
constexpr int YEAR_MAX = YEAR_MIN + (1 << MONTH_WIDTH) / 12 - 1;
__host__ __device__ __forceinline__ bool leap(int y) { return y % 4 == 0 && (y % 100 != 0 || y % 400 == 0); }
__host__ __device__ __forceinline__ int monthdays(int y, int m)
{
	const int l = (m == 2) ? 29 : ((m == 4 || m == 6 || m == 9 || m == 11) ? 30 : 31);
	return l - (m == 2 && !leap(y));
}
// date_create, gdk_time.c:62 (isdate :27, mkdate :28)
__host__ __device__ __forceinline__ int date_create(int y, int m, int d)
{
	if (!(m > 0 && m <= 12 && d > 0 && y >= YEAR_MIN && y <= YEAR_MAX && d <= monthdays(y, m)))
		return Atom<int>::nil();
	return (int) (((unsigned) ((y + YEAR_OFFSET) * 12 + m - 1) << DAY_WIDTH) | (unsigned) d);
}
// daytime_create, gdk_time.c:281 (istime :34 allows second 60, mkdaytime :35)
__host__ __device__ __forceinline__ lng daytime_create(int h, int m, int s, int u)
{
	if (!(h >= 0 && h < 24 && m >= 0 && m < 60 && s >= 0 && s <= 60 && u >= 0 && u < 1000000))
		return Atom<lng>::nil();
	return ((((lng) h * 60 + m) * 60) + s) * 1000000LL + u;
}
// timestamp_create, gdk_time.c:401 (mktimestamp :47: date << 37 | daytime)
__host__ __device__ __forceinline__ lng timestamp_create(int dt, lng tm)
{
	if (Atom<int>::is_nil(dt) || Atom<lng>::is_nil(tm))
		return Atom<lng>::nil();
	return (lng) (((unsigned long long) (unsigned) dt << 37) | (unsigned long long) tm);
}
__host__ __device__ __forceinline__ uint16_t swap16(uint16_t v) { return (uint16_t) ((v >> 8) | (v << 8)); }
__host__ __device__ __forceinline__ uint32_t swap32(uint32_t v)
{
	return (v >> 24) | ((v >> 8) & 0xff00u) | ((v << 8) & 0xff0000u) | (v << 24);
}
__host__ __device__ __forceinline__ unsigned long long swap64(unsigned long long v)
{
	return ((unsigned long long) swap32((uint32_t) v) << 32) | swap32((uint32_t) (v >> 32));
}
}	// namespace bincopy

// ---------------------------------------------------------------- decode kernels
// state: [0] nil count  [1] first offending row (~0 = none)
#define IO_THREADS 256

template <typename U> __device__ __forceinline__ U swap_any(U v);
template <> __device__ __forceinline__ uint8_t swap_any(uint8_t v) { return v; }
template <> __device__ __forceinline__ uint16_t swap_any(uint16_t v) { return bincopy::swap16(v); }
template <> __device__ __forceinline__ uint32_t swap_any(uint32_t v) { return bincopy::swap32(v); }
template <> __device__ __forceinline__ unsigned long long swap_any(unsigned long long v) { return bincopy::swap64(v); }
template <> __device__ __forceinline__ uhge swap_any(uhge v)
{
	return ((uhge) bincopy::swap64((unsigned long long) v) << 64) | bincopy::swap64((unsigned long long) (v >> 64));
}

// plain atoms: optional byte swap, optional decimal range check (|v| < 10^width), nil count.
// T = the atom, U = the unsigned carrier of its width.  flt/dbl nils are counted after the swap.
template <typename T, typename U>
__global__ void __launch_bounds__(IO_THREADS)
bin_atom_kernel(const U *src, uint64_t n, bool byteswap, bool check, T bound, T *dst,	/* may alias: converted in place */
		unsigned long long *__restrict__ state)
{
	unsigned nils = 0;
	unsigned long long bad = ~0ULL;
	for (uint64_t i = (uint64_t) blockIdx.x * IO_THREADS + threadIdx.x; i < n; i += (uint64_t) gridDim.x * IO_THREADS) {
		U u = src[i];
		if (byteswap)
			u = swap_any<U>(u);
		T v;
		memcpy(&v, &u, sizeof(T));
		if (Atom<T>::is_nil(v))
			nils++;
		else if constexpr (!Atom<T>::is_float) {
			if (check && (v >= bound || v <= -bound) && bad == ~0ULL)
				bad = i;
		}
		dst[i] = v;
	}
	if (bad != ~0ULL)
		atomicMin(&state[1], bad);
	for (int o = 16; o > 0; o >>= 1)
		nils += __shfl_down_sync(0xffffffffu, nils, o);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(&state[0], (unsigned long long) nils);
}

// bit: 0, 1 or 0x80 (nil); anything else is refused (validate_bit)
__global__ void __launch_bounds__(IO_THREADS)
bin_bit_kernel(const uint8_t *src, uint64_t n, bte *dst /* may alias src */, unsigned long long *__restrict__ state)
{
	unsigned nils = 0;
	unsigned long long bad = ~0ULL;
	for (uint64_t i = (uint64_t) blockIdx.x * IO_THREADS + threadIdx.x; i < n; i += (uint64_t) gridDim.x * IO_THREADS) {
		const uint8_t v = src[i];
		if (v == 0x80)
			nils++;
		else if (v > 1 && bad == ~0ULL)
			bad = i;
		dst[i] = (bte) v;
	}
	if (bad != ~0ULL)
		atomicMin(&state[1], bad);
	for (int o = 16; o > 0; o >>= 1)
		nils += __shfl_down_sync(0xffffffffu, nils, o);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(&state[0], (unsigned long long) nils);
}

// KIND 0 date, 1 daytime, 2 timestamp.  Records are read as 32-bit words (4 / 8 / 12 bytes, 4-byte aligned).
template <int KIND, typename T>
__global__ void __launch_bounds__(IO_THREADS)
bin_temporal_kernel(const uint32_t *__restrict__ src, uint64_t n, bool byteswap, T *__restrict__ dst, unsigned long long *__restrict__ state)
{
	unsigned nils = 0;
	for (uint64_t i = (uint64_t) blockIdx.x * IO_THREADS + threadIdx.x; i < n; i += (uint64_t) gridDim.x * IO_THREADS) {
		int dt = 0;
		lng tm = 0;
		if (KIND != 0) {
			const uint32_t *r = src + i * (KIND == 1 ? 2 : 3);
			uint32_t ms = r[0];
			const uint32_t hms = r[1];	/* seconds | minutes << 8 | hours << 16 | padding << 24 (little endian host) */
			if (byteswap)
				ms = bincopy::swap32(ms);
			tm = bincopy::daytime_create((int) ((hms >> 16) & 0xff), (int) ((hms >> 8) & 0xff), (int) (hms & 0xff), (int) ms);
		}
		if (KIND != 1) {
			const uint32_t d = src[i * (KIND == 0 ? 1 : 3) + (KIND == 0 ? 0 : 2)];	/* day | month << 8 | year << 16 */
			uint16_t y = (uint16_t) (d >> 16);
			if (byteswap)
				y = bincopy::swap16(y);
			dt = bincopy::date_create((int) (int16_t) y, (int) ((d >> 8) & 0xff), (int) (d & 0xff));
		}
		T v;
		if (KIND == 0)
			v = (T) dt;
		else if (KIND == 1)
			v = (T) tm;
		else
			v = (T) bincopy::timestamp_create(dt, tm);
		if (Atom<T>::is_nil(v))
			nils++;
		dst[i] = v;
	}
	for (int o = 16; o > 0; o >>= 1)
		nils += __shfl_down_sync(0xffffffffu, nils, o);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(&state[0], (unsigned long long) nils);
}

// the way back (encode_date / encode_time / encode_timestamp): nil -> all-ones record
template <int KIND, typename T>
__global__ void __launch_bounds__(IO_THREADS)
bin_temporal_encode_kernel(const T *__restrict__ src, uint64_t n, bool byteswap, uint32_t *__restrict__ dst)
{
	for (uint64_t i = (uint64_t) blockIdx.x * IO_THREADS + threadIdx.x; i < n; i += (uint64_t) gridDim.x * IO_THREADS) {
		const T v = src[i];
		uint32_t *r = dst + i * (KIND == 0 ? 1 : (KIND == 1 ? 2 : 3));
		if (Atom<T>::is_nil(v)) {
			r[0] = 0xffffffffu;
			if (KIND != 0)
				r[1] = 0xffffffffu;
			if (KIND == 2)
				r[2] = 0xffffffffu;
			continue;
		}
		int dt = 0;
		lng tm = 0;
		if (KIND == 0)
			dt = (int) v;
		else if (KIND == 1)
			tm = (lng) v;
		else {
			dt = (int) ((unsigned long long) v >> 37);
			tm = (lng) ((unsigned long long) v & ((1ULL << 37) - 1));
		}
		if (KIND != 1) {
			const unsigned mon = ((unsigned) dt >> bincopy::DAY_WIDTH) & ((1u << bincopy::MONTH_WIDTH) - 1);
			uint16_t y = (uint16_t) (int16_t) ((int) (mon / 12) - bincopy::YEAR_OFFSET);
			if (byteswap)
				y = bincopy::swap16(y);
			r[KIND == 0 ? 0 : 2] = ((unsigned) dt & 31u) | ((mon % 12 + 1) << 8) | ((uint32_t) y << 16);
		}
		if (KIND != 0) {
			uint32_t ms = (uint32_t) (tm % 1000000);
			if (byteswap)
				ms = bincopy::swap32(ms);
			const unsigned s = (unsigned) ((tm / 1000000) % 60), m = (unsigned) ((tm / 60000000) % 60), h = (unsigned) (tm / 3600000000LL);
			r[0] = ms;
			r[1] = s | (m << 8) | (h << 16);	/* padding 0 */
		}
	}
}

// ---------------------------------------------------------------- host side: binary COPY records
static size_t bin_record_size(int type)
{
	switch (type) {
	case B200_date: return sizeof(BinDate);
	case B200_daytime: return sizeof(BinTime);
	case B200_timestamp: return sizeof(BinTimestamp);
	default: {
		const int w = atomsize(type);
		return w > 0 ? (size_t) w : 0;
	}
	}
}

template <typename T, typename U>
static void bin_atom_launch(const void *raw, uint64_t n, bool byteswap, int width, void *dst, unsigned long long *state)
{
	T bound = 0;
	bool check = false;
	if constexpr (!Atom<T>::is_float) {
		if (width > 0) {	/* VALIDATE_DECIMAL: m = 10^width in the column's own type */
			check = true;
			bound = 1;
			for (int i = 0; i < width; i++)
				bound = (T) (bound * 10);
		}
	}
	bin_atom_kernel<T, U><<<grid_for(n, IO_THREADS, NSM * 32), IO_THREADS, 0, ctx()->stream>>>((const U *) raw, n, byteswap, check, bound, (T *) dst, state);
	COUNT_LAUNCH();
}

extern "C" b200_bat *b200_bincopy_decode(int type, const void *src, uint64_t count, bool byteswap, int width, const char *filename,
					 uint64_t hseq)
{
	b200_clrerr();
	const size_t rs = bin_record_size(type);
	if (rs == 0 || type == B200_void || type == B200_msk || type == B200_oid || type == B200_ptr) {
		set_err("bincopy: no fixed-width binary record for type %s (str / blob loaders stay with the reference)\n", atomname(type));
		return nullptr;
	}
	if (count > 0 && src == nullptr) {
		set_err("bincopy: source buffer required\n");
		return nullptr;
	}
	if (filename == nullptr)
		filename = "<buffer>";
	b200_bat *bn = bat_new(hseq, type, count);
	if (bn == nullptr || count == 0) {
		if (bn) {
			bn->tnonil = 1;
			bn->tsorted = bn->trevsorted = bn->tkey = 1;
		}
		return bn;
	}
	Ctx *c = ctx();
	const int st = basetype(type);
	const bool temporal = type == B200_date || type == B200_daytime || type == B200_timestamp;
	const bool direct = !temporal && type != B200_bit && !(byteswap && rs > 1) && width <= 0 && st != B200_flt && st != B200_dbl;
	unsigned long long *state = (unsigned long long *) scratch_get(64);
	static const unsigned long long init[2] = {0, ~0ULL};
	void *raw = nullptr;
	bool ok = state != nullptr && cuda_ok(cudaMemcpyAsync(state, init, 16, cudaMemcpyHostToDevice, c->stream), "init");
	if (ok && !temporal) {
		/* same width on both sides: upload into the result and convert in place */
		ok = cuda_ok(cudaMemcpyAsync(bn->base, src, count * rs, cudaMemcpyHostToDevice, c->stream), "bincopy upload");
		raw = bn->base;
	} else if (ok) {
		raw = dev_alloc(count * rs);
		ok = raw != nullptr && cuda_ok(cudaMemcpyAsync(raw, src, count * rs, cudaMemcpyHostToDevice, c->stream), "bincopy upload");
	}
	if (ok) {
		KTimer kt("bincopy_decode", count * (rs + (size_t) atomsize(type)));
		const unsigned grid = grid_for(count, IO_THREADS, NSM * 32);
		if (type == B200_date) {
			bin_temporal_kernel<0, int><<<grid, IO_THREADS, 0, c->stream>>>((const uint32_t *) raw, count, byteswap, (int *) bn->base, state);
			COUNT_LAUNCH();
		} else if (type == B200_daytime) {
			bin_temporal_kernel<1, lng><<<grid, IO_THREADS, 0, c->stream>>>((const uint32_t *) raw, count, byteswap, (lng *) bn->base, state);
			COUNT_LAUNCH();
		} else if (type == B200_timestamp) {
			bin_temporal_kernel<2, lng><<<grid, IO_THREADS, 0, c->stream>>>((const uint32_t *) raw, count, byteswap, (lng *) bn->base, state);
			COUNT_LAUNCH();
		} else if (type == B200_bit) {
			bin_bit_kernel<<<grid, IO_THREADS, 0, c->stream>>>((const uint8_t *) raw, count, (bte *) bn->base, state);
			COUNT_LAUNCH();
		} else {
			/* one pass also when nothing needs converting: it counts the nils for the properties */
			(void) direct;
			switch (st) {
			case B200_bte: bin_atom_kernel<bte, uint8_t><<<grid, IO_THREADS, 0, c->stream>>>((const uint8_t *) raw, count, false, false, (bte) 0, (bte *) bn->base, state); COUNT_LAUNCH(); break;
			case B200_sht: bin_atom_launch<sht, uint16_t>(raw, count, byteswap, width, bn->base, state); break;
			case B200_int: bin_atom_launch<int, uint32_t>(raw, count, byteswap, width, bn->base, state); break;
			case B200_lng: bin_atom_launch<lng, unsigned long long>(raw, count, byteswap, width, bn->base, state); break;
			case B200_hge: bin_atom_launch<hge, uhge>(raw, count, byteswap, width, bn->base, state); break;
			case B200_flt: bin_atom_launch<flt, uint32_t>(raw, count, byteswap, 0, bn->base, state); break;
			case B200_dbl: bin_atom_launch<dbl, unsigned long long>(raw, count, byteswap, 0, bn->base, state); break;
			default: ok = false; set_err("bincopy: unsupported type %s\n", atomname(type));
			}
		}
	}
	ok = ok && cuda_ok(cudaGetLastError(), "bincopy launch") &&
	     cuda_ok(cudaMemcpyAsync(c->slot_h->v, state, 16, cudaMemcpyDeviceToHost, c->stream), "read") && stream_sync();
	if (temporal && raw)
		dev_free(raw);
	if (!ok) {
		b200_BATfree(bn);
		return nullptr;
	}
	const uint64_t nils = c->slot_h->v[0], bad = c->slot_h->v[1];
	if (bad != ~0ULL) {
		if (type == B200_bit) {
			unsigned char v = 0;
			cudaMemcpy(&v, (const char *) bn->base + bad, 1, cudaMemcpyDeviceToHost);
			set_err("22003!invalid boolean byte value %d in %s\n", (int) v, filename);	/* validate_bit */
		} else {
			set_err("22003!decimal out of range in %s\n", filename);	/* VALIDATE_DECIMAL */
		}
		b200_BATfree(bn);
		return nullptr;
	}
	bn->count = count;
	bn->tnil = nils != 0;
	bn->tnonil = nils == 0;
	bn->tsorted = bn->trevsorted = bn->tkey = count <= 1;
	return bn;
}

// dump_binary_column (sql_bincopyconvert.c:640-700): the column as binary COPY records in `dst` (host, count * record size)
extern "C" int b200_bincopy_encode(const b200_bat *b, void *dst, bool byteswap)
{
	b200_clrerr();
	if (b == nullptr || (b->count > 0 && dst == nullptr)) {
		set_err("bincopy: BAT and destination buffer required\n");
		return -1;
	}
	Staged sb;
	if (!stage(b, sb))
		return -1;
	b = sb.b;
	const size_t rs = bin_record_size(b->ttype);
	if (rs == 0 || b->ttype == B200_void || b->ttype == B200_msk || b->ccand != B200_CAND_NONE) {
		set_err("bincopy: no fixed-width binary record for type %s\n", atomname(b->ttype));
		return -1;
	}
	const uint64_t n = b->count;
	if (n == 0)
		return 0;
	Ctx *c = ctx();
	const int st = basetype(b->ttype);
	const bool temporal = b->ttype == B200_date || b->ttype == B200_daytime || b->ttype == B200_timestamp;
	if (!temporal && !(byteswap && rs > 1))
		return cuda_ok(cudaMemcpyAsync(dst, b->base, n * rs, cudaMemcpyDeviceToHost, c->stream), "bincopy download") && stream_sync() ? 0 : -1;
	void *raw = dev_alloc(n * rs);
	unsigned long long *state = (unsigned long long *) scratch_get(64);
	if (raw == nullptr || state == nullptr)
		return -1;
	const unsigned grid = grid_for(n, IO_THREADS, NSM * 32);
	{
		KTimer kt("bincopy_encode", n * (rs + (size_t) atomsize(b->ttype)));
		if (b->ttype == B200_date)
			bin_temporal_encode_kernel<0, int><<<grid, IO_THREADS, 0, c->stream>>>((const int *) b->base, n, byteswap, (uint32_t *) raw);
		else if (b->ttype == B200_daytime)
			bin_temporal_encode_kernel<1, lng><<<grid, IO_THREADS, 0, c->stream>>>((const lng *) b->base, n, byteswap, (uint32_t *) raw);
		else if (b->ttype == B200_timestamp)
			bin_temporal_encode_kernel<2, lng><<<grid, IO_THREADS, 0, c->stream>>>((const lng *) b->base, n, byteswap, (uint32_t *) raw);
		else if (st == B200_sht)
			bin_atom_kernel<sht, uint16_t><<<grid, IO_THREADS, 0, c->stream>>>((const uint16_t *) b->base, n, true, false, (sht) 0, (sht *) raw, state);
		else if (st == B200_int || st == B200_flt)
			bin_atom_kernel<int, uint32_t><<<grid, IO_THREADS, 0, c->stream>>>((const uint32_t *) b->base, n, true, false, 0, (int *) raw, state);
		else if (st == B200_lng || st == B200_dbl)
			bin_atom_kernel<lng, unsigned long long><<<grid, IO_THREADS, 0, c->stream>>>((const unsigned long long *) b->base, n, true, false, (lng) 0, (lng *) raw, state);
		else
			bin_atom_kernel<hge, uhge><<<grid, IO_THREADS, 0, c->stream>>>((const uhge *) b->base, n, true, false, (hge) 0, (hge *) raw, state);
		COUNT_LAUNCH();
	}
	const bool ok = cuda_ok(cudaGetLastError(), "bincopy launch") &&
			cuda_ok(cudaMemcpyAsync(dst, raw, n * rs, cudaMemcpyDeviceToHost, c->stream), "bincopy download") && stream_sync();
	dev_free(raw);
	return ok ? 0 : -1;
}

// ---------------------------------------------------------------- tail heap files <-> device heaps
#define IO_CHUNK ((size_t) 32 << 20)

struct IoStage {
	char *buf[2] = {nullptr, nullptr};
	cudaEvent_t ev[2] = {nullptr, nullptr};
	bool ok = false;
	IoStage()
	{
		ok = cuda_ok(cudaHostAlloc((void **) &buf[0], IO_CHUNK, cudaHostAllocDefault), "pinned staging") &&
		     cuda_ok(cudaHostAlloc((void **) &buf[1], IO_CHUNK, cudaHostAllocDefault), "pinned staging") &&
		     cuda_ok(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming), "event") &&
		     cuda_ok(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming), "event");
	}
	~IoStage()
	{
		for (int i = 0; i < 2; i++) {
			if (buf[i])
				cudaFreeHost(buf[i]);
			if (ev[i])
				cudaEventDestroy(ev[i]);
		}
	}
};

// HEAPload (gdk_heap.c:729) for the device tier: the first count * width bytes of the file become the tail heap.
// A file shorter than that is refused the way GDKload refuses it; a longer one is fine (heaps are over-allocated).
extern "C" b200_bat *b200_HEAPload(const char *path, int type, uint64_t count, uint64_t hseq)
{
	b200_clrerr();
	const int w = atomsize(type);
	if (path == nullptr || w <= 0) {
		set_err("HEAPload: path and a fixed-width tail type required\n");
		return nullptr;
	}
	const int fd = open(path, O_RDONLY | O_CLOEXEC);
	if (fd < 0) {
		set_err("HEAPload: cannot open: name=%s: %s\n", path, strerror(errno));
		return nullptr;
	}
	struct stat sb;
	const size_t want = (size_t) count * (size_t) w;
	if (fstat(fd, &sb) < 0 || (size_t) sb.st_size < want) {
		set_err("HEAPload: heap file %s holds %lld bytes, %zu needed for %llu %s values\n", path, (long long) sb.st_size, want,
			(unsigned long long) count, atomname(type));
		close(fd);
		return nullptr;
	}
	b200_bat *bn = bat_new(hseq, type, count);
	if (bn == nullptr || count == 0) {
		close(fd);
		if (bn)
			bn->tsorted = bn->trevsorted = bn->tkey = bn->tnonil = 1;
		return bn;
	}
	Ctx *c = ctx();
	IoStage stg;
	bool ok = stg.ok;
	size_t off = 0;
	for (int k = 0; ok && off < want; k ^= 1) {
		const size_t len = want - off < IO_CHUNK ? want - off : IO_CHUNK;
		/* the copy that last used this buffer must have drained it */
		ok = cuda_ok(cudaEventSynchronize(stg.ev[k]), "staging wait");
		size_t got = 0;
		while (ok && got < len) {
			const ssize_t r = pread(fd, stg.buf[k] + got, len - got, (off_t) (off + got));
			if (r <= 0) {
				if (r < 0 && errno == EINTR)
					continue;
				set_err("HEAPload: read of %s failed at %zu: %s\n", path, off + got, r < 0 ? strerror(errno) : "unexpected end of file");
				ok = false;
			} else
				got += (size_t) r;
		}
		ok = ok && cuda_ok(cudaMemcpyAsync((char *) bn->base + off, stg.buf[k], len, cudaMemcpyHostToDevice, c->stream), "heap upload") &&
		     cuda_ok(cudaEventRecord(stg.ev[k], c->stream), "staging record");
		off += len;
	}
	close(fd);
	ok = ok && stream_sync();
	if (!ok) {
		b200_BATfree(bn);
		return nullptr;
	}
	bn->count = count;
	/* nothing is known about a heap that was just read: the BBP directory carries the properties (b200_BATsetprops) */
	bn->tsorted = bn->trevsorted = bn->tkey = count <= 1;
	bn->tnonil = bn->tnil = 0;
	return bn;
}

// HEAPsave (gdk_heap.c:860 -> GDKsave): the tail heap's used bytes (count * width) to `path`, created or truncated;
// dosync = fdatasync before returning (the reference's dosync)
extern "C" int b200_HEAPsave(const b200_bat *b, const char *path, bool dosync)
{
	b200_clrerr();
	if (b == nullptr || path == nullptr) {
		set_err("HEAPsave: BAT and path required\n");
		return -1;
	}
	if (b->ttype == B200_void || b->ccand != B200_CAND_NONE || atomsize(b->ttype) <= 0) {
		set_err("no heap to save\n");	/* gdk_heap.c:868 */
		return -1;
	}
	const size_t want = (size_t) b->count * (size_t) atomsize(b->ttype);
	const int fd = open(path, O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0600);
	if (fd < 0) {
		set_err("HEAPsave: cannot create: name=%s: %s\n", path, strerror(errno));
		return -1;
	}
	bool ok = true;
	if (want > 0 && b->device < 0) {
		size_t put = 0;
		while (ok && put < want) {
			const ssize_t r = write(fd, (const char *) b->base + put, want - put);
			if (r < 0 && errno == EINTR)
				continue;
			if (r <= 0) {
				set_err("HEAPsave: write of %s failed: %s\n", path, strerror(errno));
				ok = false;
			} else
				put += (size_t) r;
		}
	} else if (want > 0) {
		order_after(b);
		Ctx *c = ctx();
		IoStage stg;
		ok = stg.ok;
		size_t off = 0, pend_off[2] = {0, 0}, pend_len[2] = {0, 0};
		auto flush = [&](int k) {
			if (!ok || pend_len[k] == 0)
				return;
			ok = cuda_ok(cudaEventSynchronize(stg.ev[k]), "staging wait");
			size_t put = 0;
			while (ok && put < pend_len[k]) {
				const ssize_t r = pwrite(fd, stg.buf[k] + put, pend_len[k] - put, (off_t) (pend_off[k] + put));
				if (r < 0 && errno == EINTR)
					continue;
				if (r <= 0) {
					set_err("HEAPsave: write of %s failed: %s\n", path, strerror(errno));
					ok = false;
				} else
					put += (size_t) r;
			}
			pend_len[k] = 0;
		};
		for (int k = 0; ok && off < want; k ^= 1) {
			flush(k);	/* the buffer's previous chunk goes to disk while the other chunk's copy runs */
			const size_t len = want - off < IO_CHUNK ? want - off : IO_CHUNK;
			ok = ok && cuda_ok(cudaMemcpyAsync(stg.buf[k], (const char *) b->base + off, len, cudaMemcpyDeviceToHost, c->stream), "heap download") &&
			     cuda_ok(cudaEventRecord(stg.ev[k], c->stream), "staging record");
			pend_off[k] = off;
			pend_len[k] = len;
			off += len;
		}
		flush(0);
		flush(1);
	}
	if (ok && dosync && fdatasync(fd) < 0) {
		set_err("HEAPsave: fdatasync of %s failed: %s\n", path, strerror(errno));
		ok = false;
	}
	close(fd);
	return ok ? 0 : -1;
}
