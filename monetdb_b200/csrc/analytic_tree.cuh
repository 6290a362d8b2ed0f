// analytic_tree.cuh — the reference's fan-out-16 segment tree (gdk_analytic.h:52-130)
// rebuilt for the device: all partitions' trees are built at once, one array per
// level; level l >= 1 of partition q occupies slots (S[q] >> 4l) + q + t, t = 0 ..
// ceil(len_q / 16^l) - 1 (slots of different partitions never collide because
// floor(s/B) + ceil(len/B) <= floor((s+len)/B) + 1).  Nodes are aligned to the
// partition start and children are combined left to right, as
// populate_segment_tree does, and a query adds nodes in compute_on_segment_tree's
// order, so floating-point results carry the reference's rounding.
//
// An aggregate is an `Op`:
//   In, Node, Acc, Out          types
//   leaf(In) -> Node            COMPUTE_LEVEL0_*
//   init(Acc&)                  INIT_AGGREGATE_*
//   add(Acc&, Node, err&)       COMPUTE_LEVELN_*
//   node(Acc) -> Node           value stored for an inner node
//   result(Acc, isnil&) -> Out  FINALIZE_AGGREGATE_*
#pragma once
#include "analytic.cuh"

template <typename Node> struct TreeDev {
	const Node *lev[AN_MAXLEV];	/* lev[0] unused */
	const uhge *pl[AN_MAXLEV];	/* running totals of a level's node values over its slots, or NULL (an_tree_query_p0) */
	int nlev;			/* highest built level */
};

template <typename Node> struct Tree {
	Node *lev[AN_MAXLEV] = {nullptr};	/* all inside block */
	uint64_t size[AN_MAXLEV] = {0};
	int nlev = 0;
	Node *block = nullptr;
	uhge *plev[AN_MAXLEV] = {nullptr};	/* all inside pblock (optional) */
	uhge *pblock = nullptr;
	bool cached = false;	/* block / pblock belong to the value column's SHADOW_TREE twin */
	~Tree()
	{
		if (!cached) {
			if (block)
				dev_free(block);
			if (pblock)
				dev_free(pblock);
		}
	}
	// entries of the per-level running totals: size[level] + 1 each, every level's start rounded to 16 entries
	uint64_t prefix_total() const
	{
		uint64_t total = 0;
		for (int level = 1; level <= nlev; level++)
			total += (size[level] + 1 + 15) & ~(uint64_t) 15;
		return total;
	}
	void place_prefix(uhge *b)
	{
		pblock = b;
		uint64_t off = 0;
		for (int level = 1; level <= nlev; level++) {
			plev[level] = b + off;
			off += (size[level] + 1 + 15) & ~(uint64_t) 15;
		}
	}
	// levels needed for partitions of up to maxlen rows (to_root: build until one node is left, as the checked
	// sums need; else stop when fewer than a full group is left, like GDKrebuild_segment_tree's callers use it)
	// and their slot counts; returns the total number of slots (each level's start rounded to 16 slots)
	uint64_t plan(uint64_t n, uint64_t nparts, uint64_t maxlen, bool to_root)
	{
		uint64_t total = 0;
		nlev = 0;
		for (int level = 1; level < AN_MAXLEV; level++) {
			const uint64_t below = (maxlen + ((uint64_t) 1 << (4 * (level - 1))) - 1) >> (4 * (level - 1));
			if (to_root ? below <= 1 : below < AN_FANOUT)
				break;
			size[level] = (n >> (4 * level)) + nparts + 1;
			total += (size[level] + 15) & ~(uint64_t) 15;
			nlev = level;
		}
		return total;
	}
	void place(Node *b)
	{
		block = b;
		uint64_t off = 0;
		for (int level = 1; level <= nlev; level++) {
			lev[level] = b + off;
			off += (size[level] + 15) & ~(uint64_t) 15;
		}
	}
	TreeDev<Node> dev() const
	{
		TreeDev<Node> d;
		for (int i = 0; i < AN_MAXLEV; i++) {
			d.lev[i] = lev[i];
			d.pl[i] = plev[i];
		}
		d.nlev = nlev;
		return d;
	}
};

#ifdef __CUDACC__
// build level `level` (>= 1): one thread per slot
template <typename Op>
__global__ void __launch_bounds__(256)
an_tree_build_kernel(const oid *__restrict__ S, uint64_t cnt, const typename Op::In *__restrict__ bp, int level,
		     const typename Op::Node *__restrict__ prev, typename Op::Node *__restrict__ out, uint64_t nslots,
		     unsigned long long *__restrict__ state)
{
	const uint64_t c = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= nslots)
		return;
	const int sh = 4 * level;
	/* partition q owning slot c: the last q with (S[q] >> sh) + q <= c */
	uint64_t lo = 0, hi = cnt;
	while (lo < hi) {
		const uint64_t mid = (lo + hi) >> 1;
		if ((S[mid] >> sh) + mid <= c)
			lo = mid + 1;
		else
			hi = mid;
	}
	typename Op::Acc acc;
	Op::init(acc);
	if (lo == 0) {
		out[c] = Op::node(acc);
		return;
	}
	const uint64_t q = lo - 1;
	const oid ps = S[q], len = S[q + 1] - ps;
	const uint64_t t = c - ((ps >> sh) + q);
	const uint64_t below = (len + ((oid) 1 << (sh - 4)) - 1) >> (sh - 4);	/* nodes of q one level down */
	if (t * AN_FANOUT >= below) {
		out[c] = Op::node(acc);	/* gap slot: the empty node, so that running totals over the slots are defined */
		return;
	}
	const uint64_t x0 = t * AN_FANOUT;
	const uint64_t x1 = x0 + AN_FANOUT < below ? x0 + AN_FANOUT : below;
	unsigned err = 0;
	if (level == 1) {
		for (uint64_t x = x0; x < x1; x++)
			Op::add(acc, Op::leaf(bp[ps + x]), err);
	} else {
		const uint64_t pbase = (ps >> (sh - 4)) + q;
		for (uint64_t x = x0; x < x1; x++)
			Op::add(acc, prev[pbase + x], err);
	}
	out[c] = Op::node(acc);
	if (err)
		atomicOr(&state[0], (unsigned long long) err);
}

// the values of one level's nodes as a lng column (nil = empty node or gap slot), the input of that level's running totals
template <typename Node>
__global__ void __launch_bounds__(256)
an_level_vals_kernel(const Node *__restrict__ lev, uint64_t nslots, lng *__restrict__ out)
{
	const uint64_t c = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (c < nslots) {
		const Node x = lev[c];
		out[c] = x.has ? (lng) x.a : Atom<lng>::nil();
	}
}

template <typename Op>
__device__ __forceinline__ typename Op::Node an_tree_node(const TreeDev<typename Op::Node> &t, const typename Op::In *__restrict__ bp,
							  int level, oid ps, uint64_t q, uint64_t pos)
{
	if (level == 0)
		return Op::leaf(bp[ps + pos]);
	return t.lev[level][(ps >> (4 * level)) + q + pos];
}

// compute_on_segment_tree (gdk_analytic.h:93-128) for the frame [begin, tend) of partition q
template <typename Op>
__device__ __forceinline__ void an_tree_query(const TreeDev<typename Op::Node> &t, const typename Op::In *__restrict__ bp, oid ps,
					      uint64_t q, uint64_t begin, uint64_t tend, typename Op::Acc &acc, unsigned &err)
{
	if (begin >= tend)
		return;
	for (int level = 0; level <= t.nlev; level++) {
		uint64_t parent_begin = begin / AN_FANOUT, parent_end = tend / AN_FANOUT;
		if (parent_begin == parent_end) {
			for (uint64_t pos = begin; pos < tend; pos++)
				Op::add(acc, an_tree_node<Op>(t, bp, level, ps, q, pos), err);
			break;
		}
		const uint64_t group_begin = parent_begin * AN_FANOUT;
		if (begin != group_begin) {
			const uint64_t limit = group_begin + AN_FANOUT;
			for (uint64_t pos = begin; pos < limit; pos++)
				Op::add(acc, an_tree_node<Op>(t, bp, level, ps, q, pos), err);
			parent_begin++;
		}
		const uint64_t group_end = parent_end * AN_FANOUT;
		if (tend != group_end) {
			for (uint64_t pos = group_end; pos < tend; pos++)
				Op::add(acc, an_tree_node<Op>(t, bp, level, ps, q, pos), err);
		}
		begin = parent_begin;
		tend = parent_end;
	}
}

// The same walk with LEVEL 0 TAKEN FROM PREFIX SUMS: the two partial runs of leaves a frame touches (up to 15 rows each,
// and of different lengths in every lane of a warp) become two differences of running totals.  For ops whose leaf
// contribution is (sum of the non-nil values, their number) — the integer averages: AVERAGE_ITER keeps a * n + rr equal
// to the exact sum of what it was fed — this gives the identical accumulator.  P[k] / C[k] = sum / number of the non-nil
// values of rows [0, k) of the whole column (C == NULL: no nils).
template <typename Op>
__device__ __forceinline__ void an_tree_query_p0(const TreeDev<typename Op::Node> &t, const typename Op::In *__restrict__ bp,
						 const uhge *__restrict__ P, const unsigned long long *__restrict__ C, oid ps, uint64_t q,
						 uint64_t begin, uint64_t tend, typename Op::Acc &acc, unsigned &err)
{
	if (begin >= tend)
		return;
	auto leaves = [&](uint64_t a, uint64_t b) {
		const uint64_t ra = ps + a, rb = ps + b;
		Op::add_leaves(acc, (hge) (P[rb] - P[ra]), C ? (lng) (C[rb] - C[ra]) : (lng) (b - a));
	};
	{
		uint64_t parent_begin = begin / AN_FANOUT, parent_end = tend / AN_FANOUT;
		if (parent_begin == parent_end) {
			leaves(begin, tend);
			return;
		}
		const uint64_t group_begin = parent_begin * AN_FANOUT;
		if (begin != group_begin) {
			leaves(begin, group_begin + AN_FANOUT);
			parent_begin++;
		}
		const uint64_t group_end = parent_end * AN_FANOUT;
		if (tend != group_end)
			leaves(group_end, tend);
		begin = parent_begin;
		tend = parent_end;
	}
	for (int level = 1; level <= t.nlev; level++) {
		/* the inner levels the same way when the column has no nils (every node of a run then counts as one
		 * value): a run of up to 15 nodes, of a different length in every lane, is one difference */
		const uhge *__restrict__ PL = t.pl[level] ? t.pl[level] + (ps >> (4 * level)) + q : nullptr;
		auto nodes = [&](uint64_t a, uint64_t b) {
			if (PL)
				Op::add_leaves(acc, (hge) (PL[b] - PL[a]), (lng) (b - a));
			else
				for (uint64_t pos = a; pos < b; pos++)
					Op::add(acc, an_tree_node<Op>(t, bp, level, ps, q, pos), err);
		};
		uint64_t parent_begin = begin / AN_FANOUT, parent_end = tend / AN_FANOUT;
		if (parent_begin == parent_end) {
			nodes(begin, tend);
			break;
		}
		const uint64_t group_begin = parent_begin * AN_FANOUT;
		if (begin != group_begin) {
			nodes(begin, group_begin + AN_FANOUT);
			parent_begin++;
		}
		const uint64_t group_end = parent_end * AN_FANOUT;
		if (tend != group_end)
			nodes(group_end, tend);
		begin = parent_begin;
		tend = parent_end;
	}
}

// frame of row k (sql_catalog.h:104-113): 0-2 explicit start/end columns, 3 unbounded ->
// current peer group, 4 current peer group -> unbounded, 5 whole partition, 6 current row.
// DIRECT ops (count(*)) need no tree: the answer is the frame length.
template <typename Op, bool PEERS, bool P0 = false>
__global__ void __launch_bounds__(AN_THREADS, 6)
an_frame_kernel(const oid *__restrict__ S, uint64_t cnt, const oid *__restrict__ O, uint64_t ocnt, uint64_t n,
		const typename Op::In *__restrict__ bp, const oid *__restrict__ fstart, const oid *__restrict__ fend,
		int frame_type, TreeDev<typename Op::Node> tree, typename Op::Out *__restrict__ rb,
		unsigned long long *__restrict__ state, const uhge *__restrict__ P = nullptr,
		const unsigned long long *__restrict__ C = nullptr)
{
	__shared__ TileSegs ts;
	__shared__ TileSegs to;		/* peer groups (frame types 3 and 4 only) */
	const uint64_t base = (uint64_t) blockIdx.x * AN_TILE;
	const uint64_t last = (base + AN_TILE < n ? base + AN_TILE : n) - 1;
	ts.load(S, cnt, base, last);
	if (PEERS)
		to.load(O, ocnt, base, last);
	unsigned err = 0, nils = 0;
	for (unsigned i = threadIdx.x; i < AN_TILE; i += AN_THREADS) {
		const uint64_t k = base + i;
		if (k >= n)
			break;
		oid ps, pe;
		const uint64_t q = ts.locate(k, ps, pe) - 1;
		oid fs, fe;
		switch (frame_type) {
		case 3:
		case 4: {
			oid os = ps, oe = pe;
			if (PEERS)
				to.locate(k, os, oe);
			fs = frame_type == 3 ? ps : os;
			fe = frame_type == 3 ? oe : pe;
			break;
		}
		case 5: fs = ps; fe = pe; break;
		case 6: fs = k; fe = k + 1; break;
		default: fs = fstart[k]; fe = fend[k]; break;
		}
		bool isnil = false;
		typename Op::Out res;
		if constexpr (Op::direct) {
			res = fe > fs ? (typename Op::Out) (fe - fs) : 0;
		} else {
			if (fs < ps)
				fs = ps;
			if (fe > pe)
				fe = pe;
			typename Op::Acc acc;
			Op::init(acc);
			if (fs < fe) {
				if constexpr (P0)
					an_tree_query_p0<Op>(tree, bp, P, C, ps, q, fs - ps, fe - ps, acc, err);
				else
					an_tree_query<Op>(tree, bp, ps, q, fs - ps, fe - ps, acc, err);
			}
			res = Op::result(acc, isnil);
		}
		rb[k] = res;
		nils += isnil;
	}
	if (err)
		atomicOr(&state[0], (unsigned long long) err);
	nils = __reduce_add_sync(0xffffffffu, nils);
	if ((threadIdx.x & 31) == 0 && nils)
		atomicAdd(&state[1], (unsigned long long) nils);
}
#endif
