// comm.cu — the multi-GPU half of the path inside the library: one process per GPU, NCCL over
// NVLink/NVSwitch on the calling thread's stream.
//
// Two primitives carry every plan that is not embarrassingly parallel:
//
//   b200_comm_pack      mat.pack across the ranks (the packing step of mergetable's two-phase plans,
//                       opt_mergetable.c:1204-1215,1674-): out[i] = the rank-order concatenation of every
//                       rank's in[i].  All k small BATs of a partial-aggregate table travel in ONE
//                       fixed-size record per rank through ONE ncclAllGather; a single kernel unpacks
//                       every record into the k result BATs and leaves their counts in the mapped result
//                       slot, so the whole step costs two kernels, one collective and one sync.  Phase 2
//                       is then ordinary GDK operators (BATgroup / BATgroupsum / BATsum) on the packed
//                       BATs — which also reproduces BATgroup's first-occurrence numbering, because rank
//                       r's rows all precede rank r+1's and each rank's partial table is in local
//                       first-occurrence order.
//   b200_comm_exchange  the one all-to-all of the repartitioning plans (hash join, high-cardinality
//                       GROUP BY): k parallel arrays, each split into `world` contiguous runs by the same
//                       per-destination counts (b200_BATpartition), are exchanged with one grouped
//                       ncclSend/ncclRecv; the count matrix travels once through an ncclAllGather.
//
// NCCL is resolved at run time (dlopen "libnccl.so.2": the copy the process already maps — e.g. the one
// bundled with torch in bench.py — or the system one), so a single-GPU user of the library needs no NCCL.
#include "common.cuh"
#include <dlfcn.h>
#include <nccl.h>
#include <mutex>
#include <vector>
#include <stdlib.h>

#define PACK_MAXBATS 16
#define PACK_HDR (PACK_MAXBATS * 2)	/* per BAT: count, flags (bit 0 = tnonil) */

struct NcclApi {
	void *h = nullptr;
	ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
	ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
	ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
	ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*GroupStart)() = nullptr;
	ncclResult_t (*GroupEnd)() = nullptr;
	const char *(*GetErrorString)(ncclResult_t) = nullptr;
	ncclResult_t (*GetVersion)(int *) = nullptr;
};
static NcclApi g_nccl;
static std::mutex g_comm_mu;
static ncclComm_t g_comm = nullptr;
static int g_rank = 0, g_world = 1, g_comm_device = -1;

static bool nccl_load()
{
	if (g_nccl.h != nullptr)
		return true;
	const char *env = getenv("B200_NCCL_LIB");
	const char *names[] = {env, "libnccl.so.2", "libnccl.so", nullptr};
	void *h = nullptr;
	for (int i = 0; i < 3 && h == nullptr; i++)
		if (names[i] && names[i][0])
			h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
	if (h == nullptr) {
		set_err("b200_comm: cannot load NCCL (libnccl.so.2): %s\n", dlerror());
		return false;
	}
#define SYM(field, name) \
	do { *(void **) &g_nccl.field = dlsym(h, name); \
	     if (g_nccl.field == nullptr) { set_err("b200_comm: NCCL lacks %s\n", name); return false; } } while (0)
	SYM(GetUniqueId, "ncclGetUniqueId");
	SYM(CommInitRank, "ncclCommInitRank");
	SYM(CommDestroy, "ncclCommDestroy");
	SYM(AllGather, "ncclAllGather");
	SYM(Send, "ncclSend");
	SYM(Recv, "ncclRecv");
	SYM(GroupStart, "ncclGroupStart");
	SYM(GroupEnd, "ncclGroupEnd");
	SYM(GetErrorString, "ncclGetErrorString");
	SYM(GetVersion, "ncclGetVersion");
#undef SYM
	g_nccl.h = h;
	return true;
}

static bool nccl_ok(ncclResult_t r, const char *what)
{
	if (r == ncclSuccess)
		return true;
	set_err("NCCL error %s (%s)\n", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?", what);
	return false;
}
#define NC(x) nccl_ok((x), #x)

extern "C" int b200_comm_unique_id(void *out128)
{
	b200_clrerr();
	if (out128 == nullptr || ctx()->device < 0 || !nccl_load())
		return -1;
	static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
	ncclUniqueId id;
	if (!NC(g_nccl.GetUniqueId(&id)))
		return -1;
	memcpy(out128, &id, sizeof(id));
	return 0;
}

extern "C" int b200_comm_init(int rank, int world, const void *id128)
{
	b200_clrerr();
	Ctx *c = ctx();
	if (c->device < 0) {
		set_err("b200_comm_init: no usable CUDA device; this library has no CPU fallback\n");
		return -1;
	}
	if (world < 1 || rank < 0 || rank >= world || (world > 1 && id128 == nullptr)) {
		set_err("b200_comm_init: bad rank %d / world %d\n", rank, world);
		return -1;
	}
	std::lock_guard<std::mutex> lk(g_comm_mu);
	if (g_comm != nullptr) {
		set_err("b200_comm_init: already initialised (rank %d of %d)\n", g_rank, g_world);
		return -1;
	}
	g_rank = rank;
	g_world = world;
	g_comm_device = c->device;
	if (world == 1)
		return 0;
	if (!nccl_load())
		return -1;
	ncclUniqueId id;
	memcpy(&id, id128, sizeof(id));
	if (!NC(g_nccl.CommInitRank(&g_comm, world, id, rank))) {
		g_comm = nullptr;
		g_world = 1;
		g_rank = 0;
		return -1;
	}
	return 0;
}

extern "C" int b200_comm_rank(void) { return g_rank; }
extern "C" int b200_comm_world(void) { return g_world; }
extern "C" int b200_comm_nccl_version(void)
{
	int v = 0;
	if (g_nccl.h && g_nccl.GetVersion)
		g_nccl.GetVersion(&v);
	return v;
}

extern "C" void b200_comm_destroy(void)
{
	std::lock_guard<std::mutex> lk(g_comm_mu);
	if (g_comm != nullptr) {
		if (ctx()->device >= 0)
			cudaStreamSynchronize(ctx()->stream);
		g_nccl.CommDestroy(g_comm);
		g_comm = nullptr;
	}
	g_world = 1;
	g_rank = 0;
}

static bool comm_ready(const char *fn)
{
	Ctx *c = ctx();
	if (c->device < 0) {
		set_err("%s: no usable CUDA device; this library has no CPU fallback\n", fn);
		return false;
	}
	if (g_world > 1 && (g_comm == nullptr || c->device != g_comm_device)) {
		set_err("%s: call b200_comm_init on this device first\n", fn);
		return false;
	}
	return true;
}

// ---------------------------------------------------------------- pack
struct PackSrc {
	const void *base[PACK_MAXBATS];		/* NULL: void tail, oids from tseq */
	unsigned long long tseq[PACK_MAXBATS];
	unsigned long long count[PACK_MAXBATS];
	unsigned long long flags[PACK_MAXBATS];
	unsigned width[PACK_MAXBATS];
	unsigned long long off[PACK_MAXBATS];	/* byte offset of BAT i's region inside a record */
	int k;
};
struct PackDst {
	void *base[PACK_MAXBATS];
	unsigned width[PACK_MAXBATS];
	unsigned long long off[PACK_MAXBATS];
	unsigned long long recbytes;
	int k, world;
};

// one CTA per BAT: copy its tail into this rank's record (bytes; the tables are tiny)
__global__ void pack_record_kernel(PackSrc s, unsigned char *__restrict__ rec)
{
	const int i = blockIdx.x;
	unsigned long long *hdr = (unsigned long long *) rec;
	if (threadIdx.x == 0) {
		hdr[2 * i] = s.count[i];
		hdr[2 * i + 1] = s.flags[i];
	}
	unsigned char *dst = rec + s.off[i];
	if (s.base[i] != nullptr) {
		const unsigned char *src = (const unsigned char *) s.base[i];
		const unsigned long long nb = s.count[i] * s.width[i];
		for (unsigned long long b = threadIdx.x; b < nb; b += blockDim.x)
			dst[b] = src[b];
	} else {
		oid *d = (oid *) dst;
		for (unsigned long long r = threadIdx.x; r < s.count[i]; r += blockDim.x)
			d[r] = s.tseq[i] == OID_NIL ? OID_NIL : s.tseq[i] + r;
	}
}

// one CTA per BAT: concatenate the ranks' regions in rank order; totals and the AND of the flags -> slot
__global__ void pack_unpack_kernel(PackDst d, const unsigned char *__restrict__ recs, unsigned long long *__restrict__ slot)
{
	const int i = blockIdx.x;
	unsigned long long run = 0, flags = ~0ULL;
	unsigned char *dst = (unsigned char *) d.base[i];
	for (int r = 0; r < d.world; r++) {
		const unsigned char *rec = recs + (unsigned long long) r * d.recbytes;
		const unsigned long long *hdr = (const unsigned long long *) rec;
		const unsigned long long cnt = hdr[2 * i], nb = cnt * d.width[i];
		flags &= hdr[2 * i + 1];
		const unsigned char *src = rec + d.off[i];
		for (unsigned long long b = threadIdx.x; b < nb; b += blockDim.x)
			dst[run * d.width[i] + b] = src[b];
		run += cnt;
	}
	if (threadIdx.x == 0) {
		slot[i] = run;
		slot[PACK_MAXBATS + i] = flags;
	}
}

static_assert(sizeof(ResultSlot) >= 2 * PACK_MAXBATS * 8, "the result slot holds a count and a flag word per packed BAT");

extern "C" b200_return b200_comm_pack(int k, const b200_bat *const *in, b200_bat **out, uint64_t max_rows)
{
	b200_clrerr();
	if (!comm_ready("b200_comm_pack"))
		return B200_FAIL;
	if (k < 1 || k > PACK_MAXBATS || in == nullptr || out == nullptr || max_rows == 0) {
		set_err("b200_comm_pack: 1 <= k <= %d BATs and max_rows > 0 required\n", PACK_MAXBATS);
		return B200_FAIL;
	}
	Ctx *c = ctx();
	Staged st[PACK_MAXBATS];
	PackSrc s;
	PackDst d;
	memset(&s, 0, sizeof(s));
	memset(&d, 0, sizeof(d));
	unsigned long long off = PACK_HDR * 8;
	int types[PACK_MAXBATS];
	for (int i = 0; i < k; i++) {
		out[i] = nullptr;
		if (in[i] == nullptr || !stage(in[i], st[i]))
			return B200_FAIL;
		const b200_bat *b = st[i].b;
		if (b->ccand != B200_CAND_NONE || b->ttype == B200_msk) {
			set_err("b200_comm_pack: complex candidate lists / msk tails are not packed; BATunmask them first\n");
			return B200_FAIL;
		}
		if (b->count > max_rows) {
			set_err("b200_comm_pack: BAT %d has %llu rows, more than max_rows = %llu (a partial table this large "
				"needs the repartition plan, b200_BATpartition + b200_comm_exchange)\n", i,
				(unsigned long long) b->count, (unsigned long long) max_rows);
			return B200_FAIL;
		}
		types[i] = b->ttype == B200_void ? B200_oid : b->ttype;
		const int w = atomsize(types[i]);
		s.base[i] = b->ttype == B200_void ? nullptr : b->base;
		s.tseq[i] = b->tseqbase;
		s.count[i] = b->count;
		s.flags[i] = (b->tnonil || b->count == 0 || (b->ttype == B200_void && b->tseqbase != OID_NIL)) ? 1 : 0;
		s.width[i] = d.width[i] = (unsigned) w;
		s.off[i] = d.off[i] = off;
		off += (max_rows * w + 15) & ~15ULL;
	}
	s.k = d.k = k;
	d.world = g_world;
	d.recbytes = off;
	unsigned char *recs = (unsigned char *) dev_alloc((size_t) off * (g_world + 1));
	if (recs == nullptr)
		return B200_FAIL;
	unsigned char *mine = recs + (size_t) off * g_world;	/* send record (allgather is not in place here) */
	bool ok = true;
	for (int i = 0; i < k && ok; i++) {
		out[i] = bat_new(0, types[i], max_rows * g_world);
		ok = out[i] != nullptr;
		if (ok)
			d.base[i] = out[i]->base;
	}
	if (ok) {
		KTimer kt("comm_pack", off);
		pack_record_kernel<<<k, 128, 0, c->stream>>>(s, g_world > 1 ? mine : recs);
		COUNT_LAUNCH();
		if (g_world > 1)
			ok = NC(g_nccl.AllGather(mine, recs, off, ncclUint8, g_comm, c->stream));
		if (ok) {
			pack_unpack_kernel<<<k, 128, 0, c->stream>>>(d, recs, c->slot_d->v);
			COUNT_LAUNCH();
		}
	}
	ok = ok && cuda_ok(cudaGetLastError(), "comm_pack launch") && stream_sync();
	dev_free(recs);
	if (!ok) {
		for (int i = 0; i < k; i++) {
			b200_BATfree(out[i]);
			out[i] = nullptr;
		}
		return B200_FAIL;
	}
	ctx()->algo = g_world > 1 ? "pack: nccl allgather of one record per rank" : "pack: single rank";
	for (int i = 0; i < k; i++) {
		b200_bat *bn = out[i];
		bn->count = c->slot_h->v[i];
		bn->tnonil = (c->slot_h->v[PACK_MAXBATS + i] & 1) != 0;
		bn->tnil = 0;
		bn->tsorted = bn->trevsorted = bn->tkey = bn->count <= 1;
	}
	return B200_SUCCEED;
}

// ---------------------------------------------------------------- exchange
// receive arenas for the peer-memory exchange
struct PeerArena {
	cudaIpcMemHandle_t handle;
	uint64_t gen = 0;
	void *mapped = nullptr;
};
struct ArenaRec {		/* what every rank publishes per exchange */
	uint64_t gen;
	uint64_t bytes;
	cudaIpcMemHandle_t handle;
};
static void *g_arena = nullptr;
static size_t g_arena_bytes = 0;
static uint64_t g_arena_gen = 0;
static cudaIpcMemHandle_t g_arena_handle;
static std::vector<PeerArena> g_peer;
static bool g_p2p_ok = true;
static unsigned long long *g_token = nullptr;	/* [1 + world] device words for the completion all-gather */
static ArenaRec *g_rec_dev = nullptr;		/* [1 + world] */

// make sure my arena holds `need` bytes, publish (gen, handle), open the peers' current arenas.
// Collective: every rank calls it once per exchange.  false -> fall back to ncclSend/ncclRecv (all ranks agree,
// because the decision is taken from the gathered records).
static bool arena_exchange_handles(size_t need)
{
	Ctx *c = ctx();
	const int W = g_world;
	static bool env_checked = false;
	if (!env_checked) {
		env_checked = true;
		const char *e = getenv("B200_EXCHANGE");
		if (e && strcmp(e, "nccl") == 0)
			g_p2p_ok = false;
	}
	if ((int) g_peer.size() != W)
		g_peer.assign(W, PeerArena());
	bool mine_ok = g_p2p_ok;
	if (mine_ok && need > g_arena_bytes) {
		cudaStreamSynchronize(c->stream);
		if (g_arena)
			cudaFree(g_arena);
		g_arena = nullptr;
		g_arena_bytes = 0;
		size_t nb = need + need / 2;
		nb = (nb + ((size_t) 2 << 20) - 1) & ~(((size_t) 2 << 20) - 1);
		if (cudaMalloc(&g_arena, nb) != cudaSuccess || cudaIpcGetMemHandle(&g_arena_handle, g_arena) != cudaSuccess) {
			cudaGetLastError();
			if (g_arena)
				cudaFree(g_arena);
			g_arena = nullptr;
			mine_ok = false;
		} else {
			g_arena_bytes = nb;
			g_arena_gen++;
		}
	}
	if (g_token == nullptr) {
		if (cudaMalloc(&g_token, (size_t) (W + 1) * 8) != cudaSuccess || cudaMalloc(&g_rec_dev, (size_t) (W + 1) * sizeof(ArenaRec)) != cudaSuccess) {
			cudaGetLastError();
			g_p2p_ok = false;
			return false;	/* every rank fails the same allocation pattern or the next gather disagrees; be safe: */
		}
		cudaMemsetAsync(g_token, 0, (size_t) (W + 1) * 8, c->stream);
	}
	ArenaRec mine;
	memset(&mine, 0, sizeof(mine));
	mine.gen = mine_ok ? g_arena_gen : 0;	/* gen 0 = "I cannot take part" */
	mine.bytes = g_arena_bytes;
	if (mine_ok)
		mine.handle = g_arena_handle;
	std::vector<ArenaRec> all(W);
	if (!cuda_ok(cudaMemcpyAsync(g_rec_dev, &mine, sizeof(mine), cudaMemcpyHostToDevice, c->stream), "arena rec up") ||
	    !NC(g_nccl.AllGather(g_rec_dev, g_rec_dev + 1, sizeof(ArenaRec), ncclUint8, g_comm, c->stream)) ||
	    !cuda_ok(cudaMemcpyAsync(all.data(), g_rec_dev + 1, (size_t) W * sizeof(ArenaRec), cudaMemcpyDeviceToHost, c->stream), "arena rec down") ||
	    !stream_sync())
		return false;
	bool everyone = true;
	for (int r = 0; r < W; r++)
		everyone &= all[r].gen != 0;
	if (!everyone) {
		g_p2p_ok = false;	/* the same verdict on every rank: the records are identical everywhere */
		return false;
	}
	/* (re)open the arenas whose generation changed; every rank sees the same set of changes (seen[] is kept for
	 * all ranks, this one included), so every rank takes the confirmation round below or none does */
	static std::vector<uint64_t> seen;
	if ((int) seen.size() != W)
		seen.assign(W, 0);
	bool anychange = false, opened = true;
	for (int r = 0; r < W; r++) {
		if (seen[r] == all[r].gen)
			continue;
		anychange = true;
		seen[r] = all[r].gen;
		if (r == g_rank)
			continue;
		if (g_peer[r].mapped)
			cudaIpcCloseMemHandle(g_peer[r].mapped);
		g_peer[r].mapped = nullptr;
		if (cudaIpcOpenMemHandle(&g_peer[r].mapped, all[r].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
			cudaGetLastError();
			g_peer[r].mapped = nullptr;
			opened = false;
		}
	}
	if (anychange) {
		/* confirmation: one word per rank — did it map every peer?  Any failure switches EVERY rank to the NCCL
		 * path for good, in this very call */
		unsigned long long okw = opened ? 1 : 0;
		std::vector<unsigned long long> oks(W);
		if (!cuda_ok(cudaMemcpyAsync(g_token, &okw, 8, cudaMemcpyHostToDevice, c->stream), "confirm up") ||
		    !NC(g_nccl.AllGather(g_token, g_token + 1, 1, ncclUint64, g_comm, c->stream)) ||
		    !cuda_ok(cudaMemcpyAsync(oks.data(), g_token + 1, (size_t) W * 8, cudaMemcpyDeviceToHost, c->stream), "confirm down") ||
		    !stream_sync())
			return false;
		for (int r = 0; r < W; r++)
			if (oks[r] == 0)
				g_p2p_ok = false;
		if (!g_p2p_ok) {
			for (int r = 0; r < W; r++)
				seen[r] = 0;
			return false;
		}
	}
	return true;
}

extern "C" b200_return b200_comm_exchange(int k, const b200_bat *const *in, const uint64_t *send_counts, b200_bat **out,
					  uint64_t *recv_counts)
{
	b200_clrerr();
	if (!comm_ready("b200_comm_exchange"))
		return B200_FAIL;
	if (k < 1 || k > PACK_MAXBATS || in == nullptr || out == nullptr || send_counts == nullptr) {
		set_err("b200_comm_exchange: 1 <= k <= %d BATs and send_counts required\n", PACK_MAXBATS);
		return B200_FAIL;
	}
	Ctx *c = ctx();
	const int W = g_world;
	uint64_t total_send = 0;
	for (int r = 0; r < W; r++)
		total_send += send_counts[r];
	for (int i = 0; i < k; i++) {
		out[i] = nullptr;
		if (in[i] == nullptr || in[i]->device < 0 || in[i]->ttype == B200_void || in[i]->ccand != B200_CAND_NONE ||
		    in[i]->count != total_send) {
			set_err("b200_comm_exchange: BAT %d must be a materialised device column of sum(send_counts) rows\n", i);
			return B200_FAIL;
		}
		order_after(in[i]);
	}
	// the count matrix: every rank learns how much every rank sends to everyone (one all-gather)
	std::vector<uint64_t> mat((size_t) W * W, 0);
	if (W == 1) {
		mat[0] = send_counts[0];
	} else {
		unsigned long long *dev = (unsigned long long *) dev_alloc((size_t) (W + 1) * W * 8);
		if (dev == nullptr)
			return B200_FAIL;
		bool ok = cuda_ok(cudaMemcpyAsync(dev + (size_t) W * W, send_counts, W * 8, cudaMemcpyHostToDevice, c->stream), "counts up") &&
			  NC(g_nccl.AllGather(dev + (size_t) W * W, dev, W, ncclUint64, g_comm, c->stream)) &&
			  cuda_ok(cudaMemcpyAsync(mat.data(), dev, (size_t) W * W * 8, cudaMemcpyDeviceToHost, c->stream), "counts down") &&
			  stream_sync();
		dev_free(dev);
		if (!ok)
			return B200_FAIL;
	}
	uint64_t total_recv = 0;
	std::vector<uint64_t> roff(W + 1, 0), soff(W + 1, 0);
	for (int r = 0; r < W; r++) {
		const uint64_t n = mat[(size_t) r * W + g_rank];	/* what rank r sends to me */
		if (recv_counts)
			recv_counts[r] = n;
		roff[r + 1] = roff[r] + n;
		soff[r + 1] = soff[r] + send_counts[r];
		total_recv += n;
	}
	bool ok = true;
	for (int i = 0; i < k && ok; i++) {
		out[i] = bat_new(0, in[i]->ttype, total_recv);
		ok = out[i] != nullptr;
	}
	bool pushed = false;
	if (ok && W > 1 && g_p2p_ok) {
		/* ---- direct peer-memory exchange: every rank owns a receive arena (cudaMalloc, exported with CUDA IPC);
		 * the senders WRITE their runs straight into the destination arenas over NVLink (copy engines, no
		 * staging through NCCL's channel buffers — grouped ncclSend/ncclRecv moved these runs at 120 GB/s per GPU
		 * on NV18 links), a token all-gather tells every rank that its arena is complete, and each rank copies
		 * its arena into the result BATs.  Stream order protects the arena: a peer can only start the next
		 * exchange after this rank has contributed to that exchange's count gather, which is queued behind the
		 * copies out of the arena. */
		size_t colbase[PACK_MAXBATS + 1];
		auto layout = [&](uint64_t rows, size_t *base) {
			size_t off = 0;
			for (int i = 0; i < k; i++) {
				base[i] = off;
				off += ((size_t) rows * atomsize(in[i]->ttype) + 255) & ~(size_t) 255;
			}
			base[k] = off;
		};
		layout(total_recv, colbase);
		pushed = arena_exchange_handles(colbase[k]);
		if (pushed) {
			uint64_t bytes = 0;
			for (int i = 0; i < k; i++)
				bytes += (total_send + total_recv) * atomsize(in[i]->ttype);
			KTimer kt("comm_exchange", bytes);
			/* fork: the writes to different peers run on side streams so several copy engines / NVLink
			 * ports work at once (one stream served the 7 peers of an 8-GPU exchange at 180 GB/s) */
			constexpr int NSIDE = 4;
			static cudaStream_t side[NSIDE] = {nullptr, nullptr, nullptr, nullptr};
			static cudaEvent_t ev_fork = nullptr, ev_join[NSIDE] = {nullptr, nullptr, nullptr, nullptr};
			if (side[0] == nullptr) {
				for (int q = 0; q < NSIDE; q++) {
					cudaStreamCreateWithFlags(&side[q], cudaStreamNonBlocking);
					cudaEventCreateWithFlags(&ev_join[q], cudaEventDisableTiming);
				}
				cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming);
			}
			cudaEventRecord(ev_fork, c->stream);
			for (int q = 0; q < NSIDE; q++)
				cudaStreamWaitEvent(side[q], ev_fork, 0);
			int nth = 0;
			for (int rr = 1; rr < W && ok; rr++) {
				const int r = (g_rank + rr) % W;	/* staggered: rank g starts with peer g+1, so the ranks do not all hit the same arena first */
				cudaStream_t st = side[nth++ % NSIDE];
				uint64_t peer_total = 0, my_off = 0;
				for (int q = 0; q < W; q++) {
					peer_total += mat[(size_t) q * W + r];
					if (q < g_rank)
						my_off += mat[(size_t) q * W + r];
				}
				size_t pbase[PACK_MAXBATS + 1];
				layout(peer_total, pbase);
				for (int i = 0; i < k && ok; i++) {
					const int w = atomsize(in[i]->ttype);
					if (send_counts[r])
						ok = cuda_ok(cudaMemcpyAsync((char *) g_peer[r].mapped + pbase[i] + my_off * w,
									     (const char *) in[i]->base + soff[r] * w, send_counts[r] * w,
									     cudaMemcpyDefault, st), "exchange peer write");
				}
			}
			/* join */
			for (int q = 0; q < NSIDE; q++) {
				cudaEventRecord(ev_join[q], side[q]);
				cudaStreamWaitEvent(c->stream, ev_join[q], 0);
			}
			/* completion token: after this collective every peer's writes into my arena have finished */
			ok = ok && NC(g_nccl.AllGather(g_token, g_token + 1, 1, ncclUint64, g_comm, c->stream));
			for (int i = 0; i < k && ok; i++) {
				const int w = atomsize(in[i]->ttype);
				for (int r = 0; r < W && ok; r++) {
					const uint64_t nr = roff[r + 1] - roff[r];
					if (nr == 0)
						continue;
					const char *src = r == g_rank ? (const char *) in[i]->base + soff[r] * w
								      : (const char *) g_arena + colbase[i] + roff[r] * w;
					ok = cuda_ok(cudaMemcpyAsync((char *) out[i]->base + roff[r] * w, src, nr * w, cudaMemcpyDeviceToDevice, c->stream),
						     "exchange arena copy");
				}
			}
		}
	}
	if (ok && !pushed) {
		uint64_t bytes = 0;
		for (int i = 0; i < k; i++)
			bytes += (total_send + total_recv) * atomsize(in[i]->ttype);
		KTimer kt("comm_exchange", bytes);
		if (W > 1)
			ok = NC(g_nccl.GroupStart());
		for (int i = 0; i < k && ok; i++) {
			const int w = atomsize(in[i]->ttype);
			for (int r = 0; r < W && ok; r++) {
				const char *sp = (const char *) in[i]->base + soff[r] * w;
				char *rp = (char *) out[i]->base + roff[r] * w;
				const uint64_t ns = send_counts[r], nr = roff[r + 1] - roff[r];
				if (r == g_rank) {
					if (ns)
						ok = cuda_ok(cudaMemcpyAsync(rp, sp, ns * w, cudaMemcpyDeviceToDevice, c->stream), "exchange self copy");
					continue;
				}
				if (ns)
					ok = NC(g_nccl.Send(sp, ns * w, ncclUint8, r, g_comm, c->stream));
				if (ok && nr)
					ok = NC(g_nccl.Recv(rp, nr * w, ncclUint8, r, g_comm, c->stream));
			}
		}
		if (W > 1)
			ok = NC(g_nccl.GroupEnd()) && ok;
	}
	if (!ok) {
		for (int i = 0; i < k; i++) {
			b200_BATfree(out[i]);
			out[i] = nullptr;
		}
		return B200_FAIL;
	}
	ctx()->algo = W == 1 ? "exchange: single rank" : pushed ? "exchange: peer-memory writes over NVLink (CUDA IPC arenas)" : "exchange: nccl grouped send/recv";
	for (int i = 0; i < k; i++) {
		b200_bat *bn = out[i];
		bn->count = total_recv;
		bn->tnonil = in[i]->tnonil;
		bn->tnil = 0;
		bn->tsorted = bn->trevsorted = bn->tkey = total_recv <= 1;
	}
	return B200_SUCCEED;
}
