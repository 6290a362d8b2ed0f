// analytic_aggr.cuh — aggregate definitions for the window functions and the host
// driver that builds the per-partition segment trees and runs the frame kernel.
#pragma once
#include "analytic_tree.cuh"
#include "analytic_prefix.cuh"
#include <type_traits>

// ---- sum (gdk_analytic_func.c:1762-1796)
// Ops whose add() can raise an error (a checked sum) set `checked`: their tree is then built up to the single root
// node, as GDKrebuild_segment_tree does (gdk_analytic_func.c:34-45), so that a node no frame uses — the partial
// last node of a level, the root — fails the call on exactly the inputs the reference fails on.  For all other
// ops a level is only built when some frame can use it (>= 16 nodes below).
template <typename Op, typename = void> struct op_checked : std::false_type {};
template <typename Op> struct op_checked<Op, std::void_t<decltype(Op::checked)>> : std::integral_constant<bool, Op::checked> {};

template <typename Op, typename = void> struct op_prefix_leaves : std::false_type {};
template <typename Op> struct op_prefix_leaves<Op, std::void_t<decltype(Op::prefix_leaves)>> : std::integral_constant<bool, Op::prefix_leaves> {};

template <typename T1, typename T2> struct SumOp {
	static constexpr bool checked = true;
	typedef T1 In;
	typedef T2 Node;
	typedef T2 Acc;
	typedef T2 Out;
	static constexpr bool direct = false;
	__device__ static __forceinline__ Node leaf(In v) { return Atom<T1>::is_nil(v) ? Atom<T2>::nil() : (T2) v; }
	__device__ static __forceinline__ void init(Acc &a) { a = Atom<T2>::nil(); }
	__device__ static __forceinline__ void add(Acc &a, Node v, unsigned &err)
	{
		if (!Atom<T2>::is_nil(v)) {
			if (Atom<T2>::is_nil(a))
				a = v;
			else if (!add_with_check<T2>(v, a, a))
				err |= AN_ERR_OVERFLOW;
		}
	}
	__device__ static __forceinline__ Node node(const Acc &a) { return a; }
	__device__ static __forceinline__ Out result(const Acc &a, bool &isnil)
	{
		isnil = Atom<T2>::is_nil(a);
		return a;
	}
};

// ---- min / max (gdk_analytic_func.c:971-992)
template <typename T, bool ISMIN> struct MinMaxOp {
	typedef T In;
	typedef T Node;
	typedef T Acc;
	typedef T Out;
	static constexpr bool direct = false;
	__device__ static __forceinline__ Node leaf(In v) { return v; }
	__device__ static __forceinline__ void init(Acc &a) { a = Atom<T>::nil(); }
	__device__ static __forceinline__ void add(Acc &a, Node v, unsigned &err)
	{
		if (!Atom<T>::is_nil(v)) {
			if (Atom<T>::is_nil(a))
				a = v;
			else if (ISMIN)
				a = a < v ? a : v;
			else
				a = a > v ? a : v;
		}
	}
	__device__ static __forceinline__ Node node(const Acc &a) { return a; }
	__device__ static __forceinline__ Out result(const Acc &a, bool &isnil)
	{
		isnil = Atom<T>::is_nil(a);
		return a;
	}
};

// ---- count of non-nil values (gdk_analytic_func.c:1413-1428); count(*) is `direct`
template <typename T> struct CountOp {
	typedef T In;
	typedef lng Node;
	typedef lng Acc;
	typedef lng Out;
	static constexpr bool direct = false;
	__device__ static __forceinline__ Node leaf(In v) { return Atom<T>::is_nil(v) ? 0 : 1; }
	__device__ static __forceinline__ void init(Acc &a) { a = 0; }
	__device__ static __forceinline__ void add(Acc &a, Node v, unsigned &err) { a += v; }
	__device__ static __forceinline__ Node node(const Acc &a) { return a; }
	__device__ static __forceinline__ Out result(const Acc &a, bool &isnil)
	{
		isnil = false;
		return a;
	}
};
struct CountAllOp {
	typedef bte In;
	typedef lng Node;
	typedef lng Acc;
	typedef lng Out;
	static constexpr bool direct = true;
	__device__ static __forceinline__ Node leaf(In v) { return 1; }
	__device__ static __forceinline__ void init(Acc &a) { a = 0; }
	__device__ static __forceinline__ void add(Acc &a, Node v, unsigned &err) { a += v; }
	__device__ static __forceinline__ Node node(const Acc &a) { return a; }
	__device__ static __forceinline__ Out result(const Acc &a, bool &isnil)
	{
		isnil = false;
		return a;
	}
};

// ---- average, frame types 3-6: (dbl) sum / n with the sum kept exactly in a hge
// (gdk_analytic_statistics.c:28-147); floats: sum of doubles / n
template <typename T> struct AvgExactOp {
	typedef typename std::conditional<Atom<T>::is_float, dbl, hge>::type SumT;
	struct NodeT {
		SumT sum;
		lng n;
	};
	typedef T In;
	typedef NodeT Node;
	typedef NodeT Acc;
	typedef dbl Out;
	static constexpr bool direct = false;
	__device__ static __forceinline__ Node leaf(In v)
	{
		Node x;
		if (Atom<T>::is_nil(v)) {
			x.sum = 0;
			x.n = 0;
		} else {
			x.sum = (SumT) v;
			x.n = 1;
		}
		return x;
	}
	__device__ static __forceinline__ void init(Acc &a)
	{
		a.sum = 0;
		a.n = 0;
	}
	__device__ static __forceinline__ void add(Acc &a, const Node &v, unsigned &err)
	{
		if (v.n == 0)
			return;
		if (a.n == 0)
			a.sum = v.sum;		/* keeps the sign of a lone -0.0 */
		else
			a.sum += v.sum;
		a.n += v.n;
	}
	__device__ static __forceinline__ Node node(const Acc &a) { return a; }
	__device__ static __forceinline__ Out result(const Acc &a, bool &isnil)
	{
		isnil = a.n == 0;
		if (isnil)
			return NAN;
		return (dbl) a.sum / (dbl) a.n;
	}
};

// ---- average, frame types 0-2, integers: the reference's tree of avg_num_deltas
// (gdk_analytic_statistics.c:158-193).  COMPUTE_LEVELN_AVG_NUM feeds a child's `a` to
// AVERAGE_ITER as ONE value, so an inner node holds floor(mean of its children's a) and a
// frame's result is the AVERAGE_ITER mean (a + rr/n) over the nodes the query touches.
// AVERAGE_ITER keeps a*n + rr == sum(values) exactly, so (sum, count) in a hge gives
// the identical a and rr whatever the order.
template <typename T> struct AvgTreeIntOp {
	struct NodeT {
		T a;
		bte has;
	};
	struct AccT {
		hge sum;
		lng n;
	};
	typedef T In;
	typedef NodeT Node;
	typedef AccT Acc;
	typedef dbl Out;
	static constexpr bool direct = false;
	__device__ static __forceinline__ Node leaf(In v)
	{
		Node x;
		x.has = !Atom<T>::is_nil(v);
		x.a = x.has ? v : (T) 0;
		return x;
	}
	__device__ static __forceinline__ void init(Acc &a)
	{
		a.sum = 0;
		a.n = 0;
	}
	__device__ static __forceinline__ void add(Acc &a, const Node &v, unsigned &err)
	{
		if (v.has) {
			a.sum += (hge) v.a;
			a.n++;
		}
	}
	static constexpr bool prefix_leaves = true;	/* level 0 may come from prefix sums (an_tree_query_p0) */
	__device__ static __forceinline__ void add_leaves(Acc &a, hge sum, lng cnt)
	{
		a.sum += sum;
		a.n += cnt;
	}
	__device__ static __forceinline__ void floor_div(const Acc &a, hge &q, lng &rr)
	{
		if ((hge) (lng) a.sum == a.sum) {
			/* the usual case: the sum fits 64 bits — one hardware-sized division instead of __divti3 */
			const lng s = (lng) a.sum;
			if ((lng) (int) s == s && (lng) (int) a.n == a.n) {
				/* ... and usually 32: a native-width division */
				const int s32 = (int) s, n32 = (int) a.n;
				int q32 = s32 / n32, r32 = s32 - q32 * n32;
				if (r32 < 0) {
					q32--;
					r32 += n32;
				}
				q = q32;
				rr = r32;
				return;
			}
			lng q64 = s / a.n, r64 = s - q64 * a.n;
			if (r64 < 0) {
				q64--;
				r64 += a.n;
			}
			q = q64;
			rr = r64;
			return;
		}
		q = a.sum / (hge) a.n;
		hge r = a.sum - q * (hge) a.n;
		if (r < 0) {
			q--;
			r += a.n;
		}
		rr = (lng) r;
	}
	__device__ static __forceinline__ Node node(const Acc &a)
	{
		Node x;
		x.has = a.n != 0;
		x.a = 0;
		if (x.has) {
			hge q;
			lng rr;
			floor_div(a, q, rr);
			x.a = (T) q;
		}
		return x;
	}
	__device__ static __forceinline__ Out result(const Acc &a, bool &isnil)
	{
		isnil = a.n == 0;
		if (isnil)
			return NAN;
		hge q;
		lng rr;
		floor_div(a, q, rr);
		/* rr == 0 (one frame in n): x + 0.0 / n is x, and a zero numerator sends the fp64 division down its
		 * out-of-line slow path (ncu: taken by ~2 lanes of nearly every warp, 9 % of the kernel's instructions) */
		if (rr == 0)
			return (dbl) (T) q;
		return (dbl) (T) q + (dbl) rr / (dbl) a.n;
	}
};

// ---- integer-typed average (GDKanalyticalavginteger, gdk_analytic_statistics.c:449-700):
// AVERAGE_ITER's (avg = floor(S/n), rem) followed by round-half-away-from... the reference's
// ANALYTICAL_AVERAGE_INT_CALC_FINALIZE (:434-446): rem > 0 and (avg < 0 ? 2 rem > n : 2 rem >= n)
// bumps avg by one.  EXACT = frame types 3-5 (plain running AVERAGE_ITER over the frame's values);
// otherwise the fan-out-16 tree of avg_int_deltas, whose inner nodes hold their children's
// UN-finalised floor averages and count as one value each (COMPUTE_LEVELN_AVG_INT, :538-542).
template <typename T, bool EXACT> struct AvgIntegerOp {
	struct NodeExact {
		hge sum;
		lng n;
	};
	struct NodeTree {
		T a;
		bte has;
	};
	struct AccT {
		hge sum;
		lng n;
	};
	typedef T In;
	typedef typename std::conditional<EXACT, NodeExact, NodeTree>::type Node;
	typedef AccT Acc;
	typedef T Out;
	static constexpr bool direct = false;
	__device__ static __forceinline__ Node leaf(In v)
	{
		Node x;
		const bool has = !Atom<T>::is_nil(v);
		if constexpr (EXACT) {
			x.sum = has ? (hge) v : (hge) 0;
			x.n = has;
		} else {
			x.a = has ? v : (T) 0;
			x.has = has;
		}
		return x;
	}
	__device__ static __forceinline__ void init(Acc &a)
	{
		a.sum = 0;
		a.n = 0;
	}
	__device__ static __forceinline__ void add(Acc &a, const Node &v, unsigned &err)
	{
		if constexpr (EXACT) {
			a.sum += v.sum;
			a.n += v.n;
		} else if (v.has) {
			a.sum += (hge) v.a;
			a.n++;
		}
	}
	static constexpr bool prefix_leaves = !EXACT;
	__device__ static __forceinline__ void add_leaves(Acc &a, hge sum, lng cnt)
	{
		a.sum += sum;
		a.n += cnt;
	}
	__device__ static __forceinline__ void floor_div(const Acc &a, hge &q, lng &rr)
	{
		if ((hge) (lng) a.sum == a.sum) {
			/* the usual case: the sum fits 64 bits — one hardware-sized division instead of __divti3 */
			const lng s = (lng) a.sum;
			if ((lng) (int) s == s && (lng) (int) a.n == a.n) {
				/* ... and usually 32: a native-width division */
				const int s32 = (int) s, n32 = (int) a.n;
				int q32 = s32 / n32, r32 = s32 - q32 * n32;
				if (r32 < 0) {
					q32--;
					r32 += n32;
				}
				q = q32;
				rr = r32;
				return;
			}
			lng q64 = s / a.n, r64 = s - q64 * a.n;
			if (r64 < 0) {
				q64--;
				r64 += a.n;
			}
			q = q64;
			rr = r64;
			return;
		}
		q = a.sum / (hge) a.n;
		hge r = a.sum - q * (hge) a.n;
		if (r < 0) {
			q--;
			r += a.n;
		}
		rr = (lng) r;
	}
	__device__ static __forceinline__ Node node(const Acc &a)
	{
		Node x;
		if constexpr (EXACT) {
			x.sum = a.sum;
			x.n = a.n;
		} else {
			x.has = a.n != 0;
			x.a = 0;
			if (x.has) {
				hge q;
				lng rr;
				floor_div(a, q, rr);
				x.a = (T) q;
			}
		}
		return x;
	}
	__device__ static __forceinline__ Out result(const Acc &a, bool &isnil)
	{
		isnil = a.n == 0;
		if (isnil)
			return Atom<T>::nil();
		hge q;
		lng rr;
		floor_div(a, q, rr);
		T avg = (T) q;
		if (rr > 0) {
			if (avg < 0) {
				if (2 * rr > a.n)
					avg++;
			} else if (2 * rr >= a.n) {
				avg++;
			}
		}
		return avg;
	}
};

// ---- average, frame types 0-2, floats: running AVERAGE_ITER_FLOAT over the touched nodes
// (gdk_analytic_statistics.c:268-303, gdk_calc_private.h:277-289), order-dependent and
// therefore evaluated in the reference's order
template <typename T> struct AvgTreeFltOp {
	struct NodeT {
		T a;
		lng n;
	};
	typedef T In;
	typedef NodeT Node;
	typedef NodeT Acc;
	typedef dbl Out;
	static constexpr bool direct = false;
	__device__ static __forceinline__ Node leaf(In v)
	{
		Node x;
		x.n = Atom<T>::is_nil(v) ? 0 : 1;
		x.a = x.n ? v : (T) 0;
		return x;
	}
	__device__ static __forceinline__ void init(Acc &a)
	{
		a.a = 0;
		a.n = 0;
	}
	__device__ static __forceinline__ void add(Acc &c, const Node &v, unsigned &err)
	{
		if (v.n) {
			c.n++;
			if ((c.a > 0) == (v.a > 0))
				c.a += (v.a - c.a) / (T) c.n;
			else
				c.a = c.a - c.a / (T) c.n + v.a / (T) c.n;
		}
	}
	__device__ static __forceinline__ Node node(const Acc &a) { return a; }
	__device__ static __forceinline__ Out result(const Acc &a, bool &isnil)
	{
		isnil = a.n == 0;
		return isnil ? (dbl) NAN : (dbl) a.a;
	}
};

// ---------------------------------------------------------------- host driver
static __global__ void an_maxlen_kernel(const oid *__restrict__ S, uint64_t cnt, unsigned long long *__restrict__ out)
{
	unsigned long long m = 0;
	for (uint64_t q = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; q < cnt; q += (uint64_t) gridDim.x * blockDim.x) {
		const unsigned long long len = S[q + 1] - S[q];
		m = len > m ? len : m;
	}
	if (m)
		atomicMax(out, m);
}

struct FrameArgs {
	const b200_bat *p, *o, *b, *s, *e;
	int frame_type;
	uint64_t n;
};

// runs Op over every row's frame into rb; st = {error bits, nil count}
template <typename Op>
static bool an_run(const FrameArgs &fa, typename Op::Out *rb, unsigned long long st[2], const char *kname,
		   const uhge *P = nullptr, const unsigned long long *C = nullptr)
{
	Ctx *c = ctx();
	const uint64_t n = fa.n;
	const bool peers = fa.frame_type == 3 || fa.frame_type == 4;
	const bool framed = fa.frame_type < 3 || fa.frame_type > 6;
	SegList S, O;
	if (!an_partitions(fa.p, nullptr, n, S))
		return false;
	if (peers && !an_partitions(fa.p, fa.o, n, O))
		return false;
	unsigned long long *state = an_state_init();
	if (state == nullptr)
		return false;
	Tree<typename Op::Node> tree;
	if (!Op::direct && fa.frame_type != 6) {
		uint64_t maxlen;
		if (S.maxlen1 != 0) {
			maxlen = S.maxlen1 - 1;		/* measured by an earlier aggregate over this partition list */
		} else {
			an_maxlen_kernel<<<grid_for(S.cnt, 256, NSM * 8), 256, 0, c->stream>>>(S.S, S.cnt, state + 1);
			COUNT_LAUNCH();
			if (!cuda_ok(cudaMemcpyAsync(c->slot_h->v, state + 1, 8, cudaMemcpyDeviceToHost, c->stream), "maxlen") ||
			    !stream_sync() || !cuda_ok(cudaMemsetAsync(state + 1, 0, 8, c->stream), "state reset")) {
				dev_free(state);
				return false;
			}
			maxlen = c->slot_h->v[0];
			if (S.cached)
				shadow_set_a2(S.pbase, SHADOW_SEGLIST, n, maxlen + 1);
		}
		const uint64_t total = tree.plan(n, S.cnt, maxlen, op_checked<Op>::value);
		/* The inner levels only depend on (value column, partition list, aggregate): a plan that runs the same
		 * aggregate over several frames of one column (ROWS and RANGE, say) builds them once.  Kept as a twin of the
		 * value column when this library owns it, keyed by the partition list's token; aggregates whose build can
		 * fail (the overflow-checked sums) are not kept. */
		static char op_tag;	/* one per instantiation: identifies the aggregate */
		/* integer averages over a nil-free column: the inner levels get running totals too (an_tree_query_p0) */
		bool level_totals = false;
		if constexpr (op_prefix_leaves<Op>::value)
			level_totals = P != nullptr && C == nullptr && !peers;
		const bool keep = !op_checked<Op>::value && S.cached && S.token != 0 && fa.b->device >= 0 && fa.b->owns &&
				  fa.b->parent == nullptr && shadow_enabled() && n <= ((uint64_t) 1 << 28);
		Shadow sh;
		if (total != 0 && keep && shadow_get(fa.b->base, SHADOW_TREE, n, sh) && sh.a0 == S.token &&
		    sh.a1 == (uint64_t) (uintptr_t) &op_tag && sh.a2 == (uint64_t) tree.nlev && sh.a3 == total &&
		    (sh.p2 != nullptr || !level_totals)) {
			tree.place((typename Op::Node *) sh.p);
			if (level_totals)
				tree.place_prefix((uhge *) sh.p2);
			tree.cached = true;
		} else if (total != 0) {
			typename Op::Node *blk = (typename Op::Node *) dev_alloc(total * sizeof(typename Op::Node));
			if (blk == nullptr) {
				dev_free(state);
				return false;
			}
			tree.place(blk);
			for (int level = 1; level <= tree.nlev; level++) {
				const uint64_t nslots = tree.size[level];
				KTimer kt("an_tree_build", level == 1 ? n * sizeof(typename Op::In) : nslots * sizeof(typename Op::Node) * 17);
				an_tree_build_kernel<Op><<<(unsigned) ((nslots + 255) / 256), 256, 0, c->stream>>>(
					S.S, S.cnt, (const typename Op::In *) fa.b->base, level, tree.lev[level - 1], tree.lev[level], nslots, state);
				COUNT_LAUNCH();
			}
			if constexpr (op_prefix_leaves<Op>::value) {
				if (level_totals) {
					uhge *pb = (uhge *) dev_alloc(tree.prefix_total() * 16);
					lng *vals = (lng *) dev_alloc(tree.size[1] * 8);	/* level 1 is the largest */
					if (pb == nullptr || vals == nullptr) {
						dev_free(pb);
						dev_free(vals);
						dev_free(state);
						return false;
					}
					tree.place_prefix(pb);
					bool ok = true;
					for (int level = 1; level <= tree.nlev && ok; level++) {
						const uint64_t nslots = tree.size[level];
						an_level_vals_kernel<typename Op::Node><<<(unsigned) ((nslots + 255) / 256), 256, 0, c->stream>>>(tree.lev[level], nslots, vals);
						COUNT_LAUNCH();
						ok = an_prefix_build_raw<lng>(vals, nslots, tree.plev[level], nullptr, "an_tree_prefix");
					}
					dev_free(vals);
					if (!ok) {
						dev_free(state);
						return false;
					}
				}
			}
			if (keep && stream_sync()) {	/* published to other threads' streams only once it is complete */
				Shadow put;
				put.kind = SHADOW_TREE;
				put.p = blk;
				put.p2 = tree.pblock;
				put.n = n;
				put.a0 = S.token;
				put.a1 = (uint64_t) (uintptr_t) &op_tag;
				put.a2 = (uint64_t) tree.nlev;
				put.a3 = total;
				shadow_put(fa.b->base, put);
				tree.cached = true;
			}
		}
	}
	{
		KTimer kt(kname, n * (sizeof(typename Op::In) + sizeof(typename Op::Out) + (framed ? 16 : 0)));
		const unsigned grid = (unsigned) ((n + AN_TILE - 1) / AN_TILE);
		const oid *fs = framed ? (const oid *) fa.s->base : nullptr, *fe = framed ? (const oid *) fa.e->base : nullptr;
		if constexpr (op_prefix_leaves<Op>::value) {
			if (P != nullptr && !peers) {
				an_frame_kernel<Op, false, true><<<grid, AN_THREADS, 0, c->stream>>>(S.S, S.cnt, nullptr, 0, n, (const typename Op::In *) fa.b->base,
												      fs, fe, fa.frame_type, tree.dev(), rb, state, P, C);
				COUNT_LAUNCH();
				kt.stop();
				return an_read_state(state, st);
			}
		}
		if (peers)
			an_frame_kernel<Op, true><<<grid, AN_THREADS, 0, c->stream>>>(S.S, S.cnt, O.S, O.cnt, n, (const typename Op::In *) fa.b->base,
										   fs, fe, fa.frame_type, tree.dev(), rb, state);
		else
			an_frame_kernel<Op, false><<<grid, AN_THREADS, 0, c->stream>>>(S.S, S.cnt, nullptr, 0, n, (const typename Op::In *) fa.b->base,
										    fs, fe, fa.frame_type, tree.dev(), rb, state);
		COUNT_LAUNCH();
	}
	return an_read_state(state, st);
}

// argument checks shared by the five entry points; stages host BATs
struct FrameStage {
	Staged p, o, b, s, e;
};
bool an_frame_prepare(const char *fn, b200_bat *r, const b200_bat *p, const b200_bat *o, const b200_bat *b, const b200_bat *s,
		      const b200_bat *e, int rtype, int frame_type, FrameStage &fs, FrameArgs &fa);
b200_return an_frame_done(b200_bat *r, uint64_t n, bool has_nils);
