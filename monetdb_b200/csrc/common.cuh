// common.cuh — runtime plumbing shared by every operator of the B200 GDK path:
// per-thread CUDA context (stream, mapped result slots, error buffer), the
// device heap tier (stream-ordered pool), BAT descriptor helpers, candidate
// views (gdk_cand.h:47-70 analogue) and the nil encodings (gdk_atoms.h:149-218).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include <math.h>
#include <atomic>
#include "../../include/gdk_b200.h"

typedef uint64_t oid;
typedef int8_t bte;
typedef int16_t sht;
typedef int64_t lng;
typedef __int128 hge;
typedef unsigned __int128 uhge;
typedef float flt;
typedef double dbl;

#define OID_NIL ((oid) 1 << 63)
#define NSM 148			/* B200: 148 SMs */

// ---------------------------------------------------------------- nil / limits
template <typename T> struct Atom;
#define INT_ATOM(T, U, BITS)                                                              \
	template <> struct Atom<T> {                                                      \
		static constexpr bool is_float = false;                                   \
		__host__ __device__ static inline T nil() { return (T) ((U) 1 << (BITS - 1)); } \
		__host__ __device__ static inline T minval() { return nil() + 1; }        \
		__host__ __device__ static inline T maxval() { return (T) ~((U) 1 << (BITS - 1)); } \
		__host__ __device__ static inline bool is_nil(T v) { return v == nil(); } \
		__host__ __device__ static inline T prev(T v) { return v - 1; }           \
		__host__ __device__ static inline T next(T v) { return v + 1; }           \
	};
INT_ATOM(bte, uint8_t, 8)
INT_ATOM(sht, uint16_t, 16)
INT_ATOM(int, uint32_t, 32)
INT_ATOM(lng, uint64_t, 64)
INT_ATOM(hge, uhge, 128)
template <> struct Atom<flt> {
	static constexpr bool is_float = true;
	__host__ __device__ static inline flt nil() { return NAN; }
	__host__ __device__ static inline flt minval() { return -3.402823466e+38F; }
	__host__ __device__ static inline flt maxval() { return 3.402823466e+38F; }
	__host__ __device__ static inline bool is_nil(flt v) { return v != v; }
	__host__ __device__ static inline flt prev(flt v) { return nextafterf(v, -3.402823466e+38F); }
	__host__ __device__ static inline flt next(flt v) { return nextafterf(v, 3.402823466e+38F); }
};
template <> struct Atom<dbl> {
	static constexpr bool is_float = true;
	__host__ __device__ static inline dbl nil() { return NAN; }
	__host__ __device__ static inline dbl minval() { return -1.7976931348623157e+308; }
	__host__ __device__ static inline dbl maxval() { return 1.7976931348623157e+308; }
	__host__ __device__ static inline bool is_nil(dbl v) { return v != v; }
	__host__ __device__ static inline dbl prev(dbl v) { return nextafter(v, -1.7976931348623157e+308); }
	__host__ __device__ static inline dbl next(dbl v) { return nextafter(v, 1.7976931348623157e+308); }
};

// storage ("base") type of an atom, ATOMbasetype gdk_atoms.h:285-289
static inline int basetype(int t)
{
	switch (t) {
	case B200_bit: return B200_bte;
	case B200_date: return B200_int;
	case B200_oid: case B200_daytime: case B200_timestamp: case B200_ptr: return B200_lng;
	default: return t;
	}
}
static inline int atomsize(int t)
{
	switch (basetype(t)) {
	case B200_void: return 0;
	case B200_bte: return 1;
	case B200_sht: return 2;
	case B200_int: case B200_flt: return 4;
	case B200_lng: case B200_dbl: return 8;
	case B200_hge: return 16;
	default: return -1;
	}
}
static inline const char *atomname(int t)
{
	static const char *n[] = {"void", "msk", "bit", "bte", "sht", "int", "oid", "ptr", "flt", "dbl",
				  "lng", "hge", "date", "daytime", "timestamp"};
	return (t >= 0 && t <= 14) ? n[t] : "?";
}

// ---------------------------------------------------------------- thread context
struct ResultSlot {		/* mapped pinned memory the finishing kernels write into */
	unsigned long long v[64];
};
struct Ctx {
	int device = -1;
	cudaStream_t stream = nullptr;
	ResultSlot *slot_h = nullptr, *slot_d = nullptr;	/* host / device view of the same page */
	void *scratch = nullptr;	/* small device scratch (tile state, counters) */
	size_t scratch_bytes = 0;
	char err[1024] = {0};
	const char *algo = "";
	uint32_t id = 0;		/* index + 1 in the context registry (b200_bat.producer) */
	cudaEvent_t ev = nullptr;	/* scratch event for order_after() */
};
Ctx *ctx();				/* lazily initialises device 0 if b200_init was not called */
extern std::atomic<unsigned long long> g_launches;
#define COUNT_LAUNCH() (g_launches.fetch_add(1, std::memory_order_relaxed))

// optional per-kernel timing (b200_timing_enable): CUDA events on the launching stream
// around the kernels that dominate an operator; read back with b200_timing_read.
struct KTimer {
	int idx = -1;
	bool stopped = false;
	KTimer(const char *name, uint64_t bytes);	/* bytes = algorithmic bytes of this launch */
	~KTimer() { stop(); }
	void stop();				/* record the end event now (idempotent) */
	void add_bytes(uint64_t more);		/* bytes only known after the result count was read */
};

void set_err(const char *fmt, ...);	/* GDKerror analogue */
bool cuda_ok(cudaError_t e, const char *what);
#define CK(x) do { if (!cuda_ok((x), #x)) goto bailout; } while (0)
#define CKN(x) do { if (!cuda_ok((x), #x)) return nullptr; } while (0)

void *dev_alloc(size_t bytes);		/* stream-ordered pool allocation, NULL + error on failure */
void dev_free(void *p);
// Cross-thread ordering (every thread has its own non-blocking stream): each descriptor names the
// context that made it; stage() — which every entry point runs on every input — and b200_BATfree()
// order the calling thread's stream behind the producer stream's tail when the BAT comes from another
// context of the same device.  Same-thread use needs nothing (stream order).
void order_after(const b200_bat *b);
void stage_cache_retire(void);			/* cand.cu: the thread rebinds to another device */
void stage_cache_forget(const void *host);	/* cand.cu: one host heap was rewritten (b200_BATinvalidate) */
// Share counts of heaps that have views (BATproject with dense candidates returns a slice view).
void heap_share(const void *root);
bool heap_unshare(const void *root);		/* true: the last view of an orphaned heap went away, free it */
bool heap_owner_release(const void *root);	/* true: no views, free now; false: orphaned, the last view frees */
void hash_cache_drop(const void *base);	/* join.cu: forget a cached BAThash index */

// Narrow twins of columns this library produced itself, kept beside the descriptor and keyed by
// the tail heap like the hash cache (heap.cu).  The oid[] tail the API mandates is always written
// in full; a later operator that finds a twin reads it instead of the 8-byte form.
enum { SHADOW_GID8 = 1 /* group ids < 256 as one byte per row */,
       SHADOW_SELMASK = 2 /* select result: the scan's bitmask + chunk offsets */,
       SHADOW_SEGLIST = 3 /* a partition-flag column (GDKanalyticaldiff result): the list of partition starts */,
       SHADOW_PREFIX = 4 /* an integer value column used by window aggregates: running sums (+ non-nil counts) */,
       SHADOW_ORDERIDX = 5 /* the order index of a column (gdk_orderidx.c): its oids in ascending stable order; a0 = hseqbase
			      the oids are based on */,
       SHADOW_TREE = 6 /* a value column used by framed window aggregates: the inner levels of the fan-out-16 segment
			  trees of one aggregate over one partition list (a0 = the list's token, a1 = aggregate, a2 = levels) */ };
struct Shadow {
	int kind = 0;
	void *p = nullptr, *p2 = nullptr;	/* owned device buffers */
	uint64_t n = 0;				/* rows of the BAT the twin describes */
	uint64_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;	/* kind-specific */
};
void *scratch_steal(size_t *bytes);	/* heap.cu */
size_t scratch_size();
void shadow_put(const void *base, const Shadow &s);
bool shadow_get(const void *base, int kind, uint64_t n, Shadow &out);
void shadow_drop(const void *base);
void shadow_set_a2(const void *base, int kind, uint64_t n, uint64_t a2);	/* update one field of an existing twin */
bool shadow_enabled();
void *scratch_get(size_t bytes);	/* per-thread reusable device scratch */
bool stream_sync();			/* sync + error check */

// ---------------------------------------------------------------- BAT helpers
b200_bat *bat_alloc_desc();
b200_bat *bat_new(uint64_t hseq, int tt, uint64_t cap);		/* device tail, count 0 */
b200_bat *bat_dense(uint64_t hseq, uint64_t tseq, uint64_t cnt);
/* BATconstant (gdk_batop.c): n copies of *val (NULL = the type's nil); oid nil -> void */
b200_bat *bat_constant(uint64_t hseq, int tt, const void *val, uint64_t n);
void atom_nil_bits(int t, void *out);	/* nil bit pattern of a storage type */
static inline bool bat_is_dense(const b200_bat *b)
{
	return b->ccand == B200_CAND_NONE && (b->ttype == B200_void || b->count == 0) && b->tseqbase != OID_NIL;
}
// make sure b is usable by a kernel on the current device: device BATs are returned
// as is; host BATs (device == -1) are staged through the heap tier (tmp owns the copy)
struct Staged {
	const b200_bat *b = nullptr;
	b200_bat *tmp = nullptr;
	bool cached = false;	/* tmp belongs to the thread's stage cache (b200_stage_cache) */
	~Staged() { if (tmp && !cached) b200_BATfree(tmp); }
};
bool stage(const b200_bat *in, Staged &out);
bool stage_sparse(const b200_bat *in, Staged &out, uint64_t touch_rows);	/* zero-copy for sparse gathers */

void launch_iota(uint64_t *dst, uint64_t n, uint64_t start);	/* project.cu: dst[i] = start + i */

// ---------------------------------------------------------------- joins with a dense side (join_void.cu)
struct Cand;
bool join_dense_right(b200_bat **r1p, b200_bat **r2p, const b200_bat *lb, const b200_bat *rb, const Cand &lci, const Cand &rci,
		      unsigned mode);
b200_bat *join_materialise_dense(const b200_bat *b);

b200_return firstn_forms(b200_bat **topn, b200_bat **gids, const b200_bat *b, const b200_bat *s, const b200_bat *g, uint64_t n,
			 bool asc, bool nilslast, bool distinct);	/* firstn.cu */

// ---------------------------------------------------------------- dictionary strings (str.cu)
#define B200_STR_TYPE B200_str
b200_bat *str_select_entry(const b200_bat *b, const b200_bat *s, const void *tl, const void *th, bool li, bool hi, bool anti,
			   bool nil_matches);
b200_bat *str_thetaselect_entry(const b200_bat *b, const b200_bat *s, const void *val, const char *op);
bool str_as_offsets(const b200_bat *b, b200_bat &view, bool need_elim, const char *func);
bool str_adopt_heap(b200_bat *bn, const b200_bat *src);

// ---------------------------------------------------------------- candidates
// canditer_init (gdk_cand.c:407-712) restated for the device: after clipping to
// b's oid range a candidate list is either dense [seq, seq+ncand) or a device
// array of ascending oids.  Negative and mask encodings are expanded on device.
struct Cand {
	bool dense = true;
	oid seq = 0;			/* first candidate oid (dense) */
	const oid *oids = nullptr;	/* device array (non-dense) */
	uint64_t ncand = 0;
	oid hseq = 0;			/* b->hseqbase: row index = oid - hseq */
	oid rhseq = 0;			/* the reference's ci->hseq: hseqbase of a result aligned with the candidates */
	void *owned = nullptr;		/* temp expansion to free */
	~Cand() { if (owned) dev_free(owned); }
};
bool cand_init(Cand &ci, const b200_bat *b, const b200_bat *s);	/* b may be NULL (no clipping) */
// result of slicing the whole candidate iterator (canditer_slice(&ci,0,ncand), gdk_cand.c:1102)
b200_bat *cand_to_bat(const Cand &ci);

static inline unsigned grid_for(uint64_t n, unsigned per_block, unsigned max_blocks = NSM * 16)
{
	uint64_t g = (n + per_block - 1) / per_block;
	if (g < 1) g = 1;
	return (unsigned) (g > max_blocks ? max_blocks : g);
}

// streaming (evict-first) load of any fixed-width atom
#ifdef __CUDACC__
template <typename T> __device__ __forceinline__ T ld_stream(const T *p)
{
	if constexpr (sizeof(T) == 16) {
		ulonglong2 q = __ldcs(reinterpret_cast<const ulonglong2 *>(p));
		union { ulonglong2 q; T t; } u;
		u.q = q;
		return u.t;
	} else if constexpr (sizeof(T) == 8) {
		unsigned long long q = __ldcs(reinterpret_cast<const unsigned long long *>(p));
		union { unsigned long long q; T t; } u;
		u.q = q;
		return u.t;
	} else if constexpr (sizeof(T) == 4) {
		unsigned q = __ldcs(reinterpret_cast<const unsigned *>(p));
		union { unsigned q; T t; } u;
		u.q = q;
		return u.t;
	} else if constexpr (sizeof(T) == 2) {
		unsigned short q = __ldcs(reinterpret_cast<const unsigned short *>(p));
		union { unsigned short q; T t; } u;
		u.q = q;
		return u.t;
	} else {
		unsigned char q = __ldcs(reinterpret_cast<const unsigned char *>(p));
		union { unsigned char q; T t; } u;
		u.q = q;
		return u.t;
	}
}
#endif
