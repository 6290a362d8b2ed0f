// str.cu — dictionary-encoded string columns on the path (first slice).
//
// A GDK str BAT is a tail of fixed-width OFFSETS (1, 2, 4 or 8 bytes; the 1- and 2-byte forms are stored minus
// GDK_VAROFFSET = 8192, gdk_atoms.h:370-377,420-432) into a string heap of NUL-terminated values; the nil string
// is "\200".  While the heap is below 64 KiB GDK keeps it duplicate-free (GDK_ELIMDOUBLES), and the reference
// then works on the offsets alone: BATgroup groups the offsets as integers of that width (gdk_group.c:897-919),
// the equality select compares offsets after ONE strLocate of the constant (fullscan_str, gdk_select.c:608-656),
// BATproject gathers offsets and shares the heap (gdk_project.c:681-695).  That is what runs here:
//   b200_BATstrfromhost   the column as it sits in GDK: offsets + heap bytes (+ the elim-doubles flag)
//   BATgroup              on a duplicate-free heap: the offsets go through the integer grouping kernels
//   BATselect / theta     equality and anti-equality: every row's string is compared with the constant on the
//                         device (a duplicate-free heap needs no special case: equal strings have equal offsets,
//                         and the compare of a few cached heap bytes is not what bounds the scan)
//   BATproject            offsets gathered by the existing kernels; the result gets its own copy of the heap
//                         (a dictionary heap is small by definition)
// Everything else on str columns (ordering, joins, non-unique heaps in BATgroup) fails loudly; the shim
// forwards such calls to the reference.
#include "common.cuh"
#include "compact.cuh"

#define STR_VAROFFSET 8192u

static inline int str_width(const b200_bat *b) { return b->pad; }			/* twidth */
static inline bool str_elim(const b200_bat *b) { return (b->pad2 & 1u) != 0; }	/* GDK_ELIMDOUBLES(vheap) */

extern "C" b200_bat *b200_BATstrfromhost(const void *offsets, int width, uint64_t cnt, uint64_t hseq, const void *heap,
					 uint64_t heapbytes, bool elim_doubles)
{
	b200_clrerr();
	if ((width != 1 && width != 2 && width != 4 && width != 8) || (cnt && offsets == nullptr) || heap == nullptr) {
		set_err("BATstrfromhost: offsets of width 1/2/4/8 and a string heap required\n");
		return nullptr;
	}
	Ctx *c = ctx();
	if (c->device < 0) {
		set_err("b200: no usable CUDA device; this library has no CPU fallback\n");
		return nullptr;
	}
	b200_bat *b = bat_alloc_desc();
	if (b == nullptr)
		return nullptr;
	b->ttype = B200_str;
	b->hseqbase = hseq;
	b->count = b->capacity = cnt;
	b->pad = (uint8_t) width;
	b->pad2 = elim_doubles ? 1u : 0u;
	b->owns = 1;
	b->base = dev_alloc(cnt * width);
	b->vheap = dev_alloc(heapbytes + 16);
	b->nvheap = heapbytes;
	if (b->base == nullptr || b->vheap == nullptr ||
	    (cnt && !cuda_ok(cudaMemcpyAsync(b->base, offsets, cnt * width, cudaMemcpyHostToDevice, c->stream), "upload offsets")) ||
	    !cuda_ok(cudaMemsetAsync((char *) b->vheap + heapbytes, 0, 16, c->stream), "heap guard") ||
	    !cuda_ok(cudaMemcpyAsync(b->vheap, heap, heapbytes, cudaMemcpyHostToDevice, c->stream), "upload heap") || !stream_sync()) {
		b200_BATfree(b);
		return nullptr;
	}
	b->tsorted = b->trevsorted = b->tkey = cnt <= 1;
	b->tnonil = cnt == 0;
	return b;
}

extern "C" int b200_BATstrheap_tohost(const b200_bat *b, void *dst)
{
	if (b == nullptr || b->ttype != B200_str || b->vheap == nullptr)
		return -1;
	return cuda_ok(cudaMemcpyAsync(dst, b->vheap, b->nvheap, cudaMemcpyDeviceToHost, ctx()->stream), "heap download") && stream_sync() ? 0 : -1;
}

extern "C" int b200_BATstrwidth(const b200_bat *b) { return b && b->ttype == B200_str ? str_width(b) : -1; }

// the offsets as an integer column of their width: what BATgroup / BATproject work on
bool str_as_offsets(const b200_bat *b, b200_bat &view, bool need_elim, const char *func)
{
	if (need_elim && !str_elim(b)) {
		set_err("%s: string heap with duplicates (>= 64 KiB): comparing by offset is not possible; not on the device path\n", func);
		return false;
	}
	view = *b;		/* owns / parent stay: BATproject's slice view shares the offsets heap through them */
	view.vheap = nullptr;
	view.nvheap = 0;
	switch (str_width(b)) {
	case 1: view.ttype = B200_bte; break;
	case 2: view.ttype = B200_sht; break;
	case 4: view.ttype = B200_int; break;
	case 8: view.ttype = B200_lng; break;
	default:
		set_err("%s: bad offset width %d\n", func, str_width(b));
		return false;
	}
	/* offsets are plain bit patterns: none of them is a nil of the integer type as far as equality goes, but the
	 * integer kernels must not skip or special-case any value — BATgroup and BATproject do not */
	view.tnonil = 1;
	view.tnil = 0;
	view.tsorted = view.trevsorted = view.tkey = b->count <= 1;
	return true;
}

// give a result BAT of offsets the string type and its own copy of the heap
bool str_adopt_heap(b200_bat *bn, const b200_bat *src)
{
	Ctx *c = ctx();
	void *h = dev_alloc(src->nvheap + 16);
	if (h == nullptr || !cuda_ok(cudaMemcpyAsync(h, src->vheap, src->nvheap + 16, cudaMemcpyDeviceToDevice, c->stream), "heap copy")) {
		dev_free(h);
		return false;
	}
	bn->ttype = B200_str;
	bn->pad = src->pad;
	bn->pad2 = src->pad2;
	bn->vheap = h;
	bn->nvheap = src->nvheap;
	bn->ccand = B200_CAND_NONE;
	if (bn->parent == nullptr)
		bn->owns = 1;
	return true;
}

// ---------------------------------------------------------------- equality select
template <int W> struct StrEq {
	const unsigned char *offs;	/* W bytes per row */
	const oid *cand;		/* NULL: dense */
	oid seq, hseq;
	const char *heap;
	const char *val;		/* device copy of the constant, NUL-terminated */
	unsigned vlen;
	bool anti, valnil, nil_matches;
	__device__ __forceinline__ size_t offset_of(uint64_t row) const
	{
		if (W == 1)
			return (size_t) offs[row] + STR_VAROFFSET;
		if (W == 2)
			return (size_t) reinterpret_cast<const unsigned short *>(offs)[row] + STR_VAROFFSET;
		if (W == 4)
			return (size_t) reinterpret_cast<const unsigned *>(offs)[row];
		return (size_t) reinterpret_cast<const unsigned long long *>(offs)[row];
	}
	__device__ __forceinline__ bool operator()(uint64_t idx, oid &o) const
	{
		o = cand ? __ldcs(cand + idx) : seq + idx;
		const char *s = heap + offset_of(o - hseq);
		const bool isnil = (unsigned char) s[0] == 0x80 && s[1] == 0;
		bool eq = true;
		for (unsigned i = 0; i <= vlen && eq; i++)
			eq = s[i] == val[i];
		if (!anti)
			return eq && (!isnil || valnil);	/* v = nil selects the nils (equi-nil select) */
		/* anti: everything that differs; nils only when asked for (gdk_select.c:1303-1339) */
		if (isnil)
			return nil_matches && !valnil;
		return !eq;
	}
	__device__ __forceinline__ oid at(uint64_t idx) const { return cand ? cand[idx] : seq + idx; }
};

template <int W>
static bool str_select_run(const b200_bat *b, const Cand &ci, const char *dval, unsigned vlen, bool anti, bool valnil, bool nil_matches,
			   oid *dst, CompactResult &res)
{
	StrEq<W> f;
	f.offs = (const unsigned char *) b->base;
	f.cand = ci.dense ? nullptr : ci.oids;
	f.seq = ci.seq;
	f.hseq = ci.hseq;
	f.heap = (const char *) b->vheap;
	f.val = dval;
	f.vlen = vlen;
	f.anti = anti;
	f.valnil = valnil;
	f.nil_matches = nil_matches;
	return run_compact<1, StrEq<W>>(ci.ncand, f, dst, res);
}

// called by b200_BATselect / b200_BATthetaselect for str columns: equality (anti == false) or inequality
b200_bat *str_select(const b200_bat *b, const b200_bat *s, const char *val, bool anti, bool nil_matches)
{
	Ctx *c = ctx();
	Cand ci;
	if (!cand_init(ci, b, s))
		return nullptr;
	if (ci.ncand == 0)
		return bat_dense(0, 0, 0);
	const unsigned vlen = (unsigned) strlen(val);
	const bool valnil = (unsigned char) val[0] == 0x80 && val[1] == 0;
	char *dval = (char *) dev_alloc(vlen + 1);
	oid *dst = (oid *) dev_alloc(ci.ncand * sizeof(oid));
	b200_bat *bn = nullptr;
	CompactResult res;
	bool ok = dval != nullptr && dst != nullptr &&
		  cuda_ok(cudaMemcpyAsync(dval, val, vlen + 1, cudaMemcpyHostToDevice, c->stream), "upload constant");
	if (ok) {
		switch (str_width(b)) {
		case 1: ok = str_select_run<1>(b, ci, dval, vlen, anti, valnil, nil_matches, dst, res); break;
		case 2: ok = str_select_run<2>(b, ci, dval, vlen, anti, valnil, nil_matches, dst, res); break;
		case 4: ok = str_select_run<4>(b, ci, dval, vlen, anti, valnil, nil_matches, dst, res); break;
		default: ok = str_select_run<8>(b, ci, dval, vlen, anti, valnil, nil_matches, dst, res);
		}
	}
	dev_free(dval);
	if (!ok) {
		dev_free(dst);
		return nullptr;
	}
	ctx()->algo = anti ? "select: cuda string anti-equi scan" : "select: cuda string equi scan";
	bn = oidlist_to_bat(dst, ci.ncand, res);
	return bn;
}

// BATselect on a str column: only the equality forms are on the device path (tl == th or th == NULL, both bounds
// inclusive; anti gives the inequality)
b200_bat *str_select_entry(const b200_bat *b, const b200_bat *s, const void *tl, const void *th, bool li, bool hi, bool anti,
			   bool nil_matches)
{
	if (tl == nullptr || !li || !hi || (th != nullptr && strcmp((const char *) tl, (const char *) th) != 0)) {
		set_err("BATselect: only equality / inequality selects on str columns are on the device path\n");
		return nullptr;
	}
	Staged sb, ss;
	if (!stage(s, ss))
		return nullptr;
	if (b->device < 0) {
		set_err("BATselect: str columns must be device-resident (b200_BATstrfromhost)\n");
		return nullptr;
	}
	order_after(b);
	const char *v = (const char *) tl;
	if ((unsigned char) v[0] == 0x80 && v[1] == 0 && th != nullptr && !nil_matches) {
		/* tl == th == nil without nil_matches is the nil-nil RANGE (gdk_select.c:1400-1407): every non-nil
		 * value, and its complement is empty (:1461-1476) */
		if (anti)
			return bat_dense(0, 0, 0);
		return str_select(b, ss.b, v, true, false);
	}
	return str_select(b, ss.b, v, anti, nil_matches);
}

b200_bat *str_thetaselect_entry(const b200_bat *b, const b200_bat *s, const void *val, const char *op)
{
	const char *v = (const char *) val;
	const bool valnil = (unsigned char) v[0] == 0x80 && v[1] == 0;
	if (strcmp(op, "eq") == 0)
		return str_select_entry(b, s, val, nullptr, true, true, false, true);
	if (strcmp(op, "ne") == 0)
		return str_select_entry(b, s, val, nullptr, true, true, true, true);
	if (valnil)
		return bat_dense(0, 0, 0);	/* gdk_select.c:2126-2127 */
	if (strcmp(op, "=") == 0 || strcmp(op, "==") == 0)
		return str_select_entry(b, s, val, nullptr, true, true, false, false);
	if (strcmp(op, "!=") == 0 || strcmp(op, "<>") == 0)
		return str_select_entry(b, s, val, nullptr, true, true, true, false);
	set_err("BATthetaselect: ordering selects on str columns are not on the device path\n");
	return nullptr;
}
