// graph.cuh -- sm_100a device code of the step right after the k-mer table:
// the abundance filter of getUnitigGraph, HashList::filter and the unitig chains.
//
//   getCoveredEdges                      MBG/MBG.cpp:367-388   edges with coverage >= -a in both directions
//   getUnitigHashes / checkUnitigHashes  MBG/MBG.cpp:305-344   maximal non-branching chain; kept iff mean coverage >= -u
//   getUnitigGraph                       MBG/MBG.cpp:629-760   which chains exist (keepGaps = false)
//   removeOnecovNodes                    MBG/MBG.cpp:527-627   oneCovHeuristic
//   HashList::filter                     MBG/HashList.cpp:159-232
//
// The reference walks every chain sequentially.  A chain is a connected component of the LINK relation
//     x -> y   iff   out(x) = {y},  out(rev y) = {rev x}  and  y is not x's own node
// over oriented nodes x = 2*id + fw (rev x = x ^ 1; out() = covered successors).  Links are mirrored (x -> y  <=>
// rev y -> rev x), every oriented node has at most one link out and one in, so components are simple paths or cycles and
// the walk order of the reference does not matter for which nodes share a chain.  Here every oriented node ranks itself
// against the end of its chain by pointer doubling (Wyllie list ranking): after ceil(log2 L) rounds it knows the coverage
// sum and node count from itself to the chain end and which oriented node that end is; a node's chain total is the sum
// of its two directions minus itself.  Cycles never finish; they are detected by the round limit, cut in front of their
// smallest node id (where the reference, which visits nodes in id order, starts them) and ranked again as paths.
// All of it is 32-byte random access: bound by DRAM sector throughput like the table itself.
#pragma once
#include "kernels.cuh"

namespace mbgfe_dev {

constexpr uint32_t NO_LINK = 0xffffffffu;

// one oriented node while ranking: everything from itself up to (not including) nxt
struct __align__(32) ChainState
{
	unsigned long long acc;   // coverage sum
	uint32_t cnt;             // nodes
	uint32_t nxt;             // next oriented node not yet folded in, NO_LINK = reached the chain end
	uint32_t last;            // last oriented node folded in
	uint32_t lab;             // smallest node id folded in
	uint32_t pad[2];
};

// getCoveredEdges (MBG/MBG.cpp:367-388): out-degree and one successor of every oriented node.  The table holds each edge
// once, in canonical form; its mirror image is the same edge unless it is its own mirror (x -> rev x).
__global__ void __launch_bounds__(256) k_graph_degrees(const EdgeRec* __restrict__ edges, uint64_t n_edges, uint32_t min_cov, uint32_t* __restrict__ deg, uint32_t* __restrict__ nb)
{
	for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_edges; e += (uint64_t)gridDim.x * blockDim.x)
	{
		const EdgeRec r = edges[e];
		if (r.coverage < min_cov) continue;
		const uint32_t x = 2 * r.from + r.from_fw, y = 2 * r.to + r.to_fw;
		atomicAdd(deg + x, 1u);
		nb[x] = y;
		if ((y ^ 1) != x || (x ^ 1) != y)
		{
			atomicAdd(deg + (y ^ 1), 1u);
			nb[y ^ 1] = x ^ 1;
		}
	}
}

// the link relation (the continuation test of getUnitigHashes, MBG/MBG.cpp:311-317; its `newPos == start` test only ends
// cycles and is replaced by the cut below)
__global__ void __launch_bounds__(256) k_graph_links(const uint32_t* __restrict__ deg, const uint32_t* __restrict__ nb, uint64_t n_oriented, uint32_t* __restrict__ link)
{
	for (uint64_t x = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; x < n_oriented; x += (uint64_t)gridDim.x * blockDim.x)
	{
		uint32_t l = NO_LINK;
		if (deg[x] == 1)
		{
			const uint32_t y = nb[x];
			if (deg[y ^ 1] == 1 && (y >> 1) != (uint32_t)(x >> 1)) l = y;
		}
		link[x] = l;
	}
}

__global__ void __launch_bounds__(256) k_chain_init(const uint32_t* __restrict__ link, const uint32_t* __restrict__ cov, uint64_t n_oriented, ChainState* __restrict__ st)
{
	for (uint64_t x = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; x < n_oriented; x += (uint64_t)gridDim.x * blockDim.x)
	{
		ChainState s;
		s.acc = cov[x >> 1]; s.cnt = 1; s.nxt = link[x]; s.last = (uint32_t)x; s.lab = (uint32_t)(x >> 1); s.pad[0] = 0; s.pad[1] = 0;
		st[x] = s;
	}
}

// one doubling round, out of place (a state is one 32-byte sector: one load of the own state, one random load)
__global__ void __launch_bounds__(256) k_chain_jump(const ChainState* __restrict__ in, ChainState* __restrict__ out, uint64_t n_oriented, uint32_t* __restrict__ unfinished)
{
	bool open = false;
	for (uint64_t x = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; x < n_oriented; x += (uint64_t)gridDim.x * blockDim.x)
	{
		ChainState s = in[x];
		if (s.nxt != NO_LINK)
		{
			const ChainState t = in[s.nxt];
			s.acc += t.acc; s.cnt += t.cnt; s.last = t.last; s.lab = min(s.lab, t.lab); s.nxt = t.nxt;
			open |= s.nxt != NO_LINK;
		}
		out[x] = s;
	}
	if (open) *unfinished = 1;      // every writer stores the same value
}

// after the round limit only cycle members are unfinished, and lab is the smallest node id of their cycle.  The
// reference starts a cycle at the forward orientation of its first unvisited node (MBG/MBG.cpp:682-690, 746-760), i.e.
// of its smallest id: remove the link that enters (lab, fw) and its mirror image, the link that leaves (lab, bw).
__global__ void __launch_bounds__(256) k_chain_cut(const ChainState* __restrict__ st, uint64_t n_oriented, uint32_t* __restrict__ link)
{
	for (uint64_t x = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; x < n_oriented; x += (uint64_t)gridDim.x * blockDim.x)
	{
		const ChainState s = st[x];
		if (s.nxt == NO_LINK) continue;
		if (link[x] == 2 * s.lab + 1 || (uint32_t)x == 2 * s.lab) link[x] = NO_LINK;
	}
}

// checkUnitigHashes (MBG/MBG.cpp:322-344): kept iff (double) coverage sum / (double) chain length >= minUnitigCoverage
__global__ void __launch_bounds__(256) k_chain_kept(const ChainState* __restrict__ st, const uint32_t* __restrict__ cov, uint64_t n_nodes, double min_unitig_cov, uint32_t* __restrict__ kept)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_nodes; i += (uint64_t)gridDim.x * blockDim.x)
	{
		const ChainState b = st[2 * i], f = st[2 * i + 1];
		const double total = (double)(f.acc + b.acc - cov[i]);
		const double n = (double)(f.cnt + b.cnt - 1);
		kept[i] = (total / n >= min_unitig_cov) ? 1u : 0u;
	}
}

// ---- removeOnecovNodes (MBG/MBG.cpp:527-627), the oneCovHeuristic passes getUnitigGraph runs first under Verkko's flags.
// A RUN is a maximal stretch of coverage-1 k-mers along links (extendOneCoverageKmers: the link test plus coverage(next) == 1;
// its `next == start` test only stops circular runs, which the neighbour test below never removes).  The run goes when
// both of its ends are "replaceable": every k-mer an end leads to is also entered from some k-mer of coverage >= 2
// (MBG.cpp:583-612; the end itself has coverage 1, so "another predecessor with coverage >= 2" is "any").  With onlyTips
// exactly one end must have no successor at all.  The reference walks the runs one by one; here
//   strong[y] = 1   some predecessor of oriented node y has coverage >= 2          (edge pass)
//   bad[x]    = 1   some successor of x has no strong predecessor                  (edge pass)
//   run_end[x]      last oriented node of x's run in x's direction                 (pointer jumping, RUN_END = final)
// and every coverage-1 node reads the verdict of its own run from its two ends.
constexpr uint32_t RUN_END = 0x80000000u;

__global__ void __launch_bounds__(256) k_onecov_strong(const EdgeRec* __restrict__ edges, uint64_t n_edges, const uint32_t* __restrict__ cov, uint32_t* __restrict__ strong)
{
	for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_edges; e += (uint64_t)gridDim.x * blockDim.x)
	{
		const EdgeRec r = edges[e];
		if (r.coverage < 1) continue;
		const uint32_t x = 2 * r.from + r.from_fw, y = 2 * r.to + r.to_fw;
		if (cov[r.from] >= 2) strong[y] = 1;            // x -> y
		if (cov[r.to] >= 2) strong[x ^ 1] = 1;          // its mirror image rev y -> rev x
	}
}

__global__ void __launch_bounds__(256) k_onecov_bad(const EdgeRec* __restrict__ edges, uint64_t n_edges, const uint32_t* __restrict__ strong, uint32_t* __restrict__ bad)
{
	for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_edges; e += (uint64_t)gridDim.x * blockDim.x)
	{
		const EdgeRec r = edges[e];
		if (r.coverage < 1) continue;
		const uint32_t x = 2 * r.from + r.from_fw, y = 2 * r.to + r.to_fw;
		if (!strong[y]) bad[x] = 1;
		if (!strong[x ^ 1]) bad[y ^ 1] = 1;
	}
}

__global__ void __launch_bounds__(256) k_onecov_init(const uint32_t* __restrict__ link, const uint32_t* __restrict__ cov, uint64_t n_oriented, uint32_t* __restrict__ run_end)
{
	for (uint64_t x = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; x < n_oriented; x += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint32_t l = link[x];
		const bool on = cov[x >> 1] == 1 && l != NO_LINK && cov[l >> 1] == 1;
		run_end[x] = on ? l : ((uint32_t)x | RUN_END);
	}
}

// one jumping round, in place: a pointer only ever moves further along its own run, so reading a neighbour's pointer
// before or after that neighbour's update of the same round are both right
__global__ void __launch_bounds__(256) k_onecov_jump(uint32_t* run_end, uint64_t n_oriented, uint32_t* __restrict__ unfinished)
{
	bool open = false;
	for (uint64_t x = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; x < n_oriented; x += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint32_t p = run_end[x];
		if (p & RUN_END) continue;
		const uint32_t q = ((volatile uint32_t*)run_end)[p];
		run_end[x] = q;
		open |= !(q & RUN_END);
	}
	if (open) *unfinished = 1;
}

// kept = 0 for every node of a run that goes.  A pointer that is not final after the round limit belongs to a circular
// run: kept.
__global__ void __launch_bounds__(256) k_onecov_verdict(const uint32_t* __restrict__ run_end, const uint32_t* __restrict__ deg, const uint32_t* __restrict__ bad,
	const uint32_t* __restrict__ cov, uint64_t n_nodes, uint32_t only_tips, uint32_t* __restrict__ kept)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_nodes; i += (uint64_t)gridDim.x * blockDim.x)
	{
		uint32_t keep = 1;
		if (cov[i] == 1)
		{
			const uint32_t f = run_end[2 * i + 1], b = run_end[2 * i];
			if (f & b & RUN_END)
			{
				const uint32_t fw_node = f & ~RUN_END, bw_node = b & ~RUN_END;
				bool go = !bad[fw_node] && !bad[bw_node];
				if (only_tips) go = go && ((deg[fw_node] == 0) != (deg[bw_node] == 0));
				if (go) keep = 0;
			}
		}
		kept[i] = keep;
	}
}

// kept flags of the input table from the survivors' original indices
__global__ void __launch_bounds__(256) k_mark_kept(const uint32_t* __restrict__ orig, uint64_t n_kept, uint8_t* __restrict__ kept8)
{
	for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_kept; j += (uint64_t)gridDim.x * blockDim.x) kept8[orig[j]] = 1;
}

// HashList::filter (MBG/HashList.cpp:159-232), nodes: kept nodes keep their order; new_id = exclusive scan of kept
__global__ void __launch_bounds__(256) k_filter_nodes(const uint32_t* __restrict__ kept, const uint64_t* __restrict__ new_id, uint64_t n_nodes, const uint4* __restrict__ hash,
	const uint32_t* __restrict__ cov, const uint8_t* __restrict__ tip, const uint32_t* __restrict__ orig, uint4* __restrict__ f_hash, uint32_t* __restrict__ f_cov, uint8_t* __restrict__ f_tip,
	uint32_t* __restrict__ f_orig)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_nodes; i += (uint64_t)gridDim.x * blockDim.x)
	{
		if (!kept[i]) continue;
		const uint64_t j = new_id[i];
		f_hash[j] = hash[i]; f_cov[j] = cov[i]; f_tip[j] = tip[i];
		f_orig[j] = orig ? orig[i] : (uint32_t)i;      // where the node sat in the table the caller exported (orig == nullptr: this is that table)
	}
}

// edges between kept nodes (every coverage), flags for the scan
__global__ void __launch_bounds__(256) k_filter_edge_flags(const EdgeRec* __restrict__ edges, uint64_t n_edges, const uint32_t* __restrict__ kept, uint32_t* __restrict__ flag)
{
	for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_edges; e += (uint64_t)gridDim.x * blockDim.x)
		flag[e] = (kept[edges[e].from] && kept[edges[e].to]) ? 1u : 0u;
}

// renumbering is monotone, so canonical form and order of the edge list survive
__global__ void __launch_bounds__(256) k_filter_edges(const EdgeRec* __restrict__ edges, uint64_t n_edges, const uint32_t* __restrict__ flag, const uint64_t* __restrict__ at,
	const uint64_t* __restrict__ new_id, EdgeRec* __restrict__ out)
{
	for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_edges; e += (uint64_t)gridDim.x * blockDim.x)
	{
		if (!flag[e]) continue;
		EdgeRec r = edges[e];
		r.from = (uint32_t)new_id[r.from]; r.to = (uint32_t)new_id[r.to];
		out[at[e]] = r;
	}
}

// Unitigs of the filtered list (getUnitigGraph's second block, MBG/MBG.cpp:689-760): the chains again, over nodes with
// coverage >= -a.  The reference numbers and orients them by its walk order; here a unitig is read from its end with the
// smaller node id (a cut cycle: from its smallest node, forward), and unitigs are numbered by the id of that first node.
//   first[i]  oriented first node of i's unitig (NO_LINK when coverage < -a)
//   start_len[i] = unitig length when i is that first node, else 0   (scanned into unitig_begin)
__device__ __forceinline__ void unitig_place(const ChainState& b, const ChainState& f, uint32_t& first, uint32_t& pos, uint32_t& same)
{
	// in the direction of (i, fw) the unitig reads  A ... i ... B
	const uint32_t A = b.last ^ 1, B = f.last;
	same = (A >> 1) <= (B >> 1) ? 1u : 0u;
	first = same ? A : (B ^ 1);
	pos = same ? b.cnt - 1 : f.cnt - 1;
}

__global__ void __launch_bounds__(256) k_unitig_starts(const ChainState* __restrict__ st, const uint32_t* __restrict__ cov, uint64_t n_nodes, uint32_t min_cov,
	uint32_t* __restrict__ is_start, uint32_t* __restrict__ start_len)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_nodes; i += (uint64_t)gridDim.x * blockDim.x)
	{
		uint32_t s = 0, len = 0;
		if (cov[i] >= min_cov)
		{
			const ChainState b = st[2 * i], f = st[2 * i + 1];
			uint32_t first, pos, same;
			unitig_place(b, f, first, pos, same);
			if ((first >> 1) == (uint32_t)i) { s = 1; len = f.cnt + b.cnt - 1; }
		}
		is_start[i] = s; start_len[i] = len;
	}
}

// unitig_id = exclusive scan of is_start, unitig_at = exclusive scan of start_len (both indexed by node)
__global__ void __launch_bounds__(256) k_unitig_write(const ChainState* __restrict__ st, const uint32_t* __restrict__ cov, uint64_t n_nodes, uint32_t min_cov,
	const uint32_t* __restrict__ is_start, const uint64_t* __restrict__ unitig_id, const uint64_t* __restrict__ unitig_at,
	uint64_t* __restrict__ unitig_begin, uint32_t* __restrict__ unitig_nodes, uint32_t* __restrict__ node_unitig, uint32_t* __restrict__ node_ufw, uint32_t* __restrict__ node_upos)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_nodes; i += (uint64_t)gridDim.x * blockDim.x)
	{
		if (cov[i] < min_cov) { node_unitig[i] = NO_LINK; continue; }
		const ChainState b = st[2 * i], f = st[2 * i + 1];
		uint32_t first, pos, same;
		unitig_place(b, f, first, pos, same);
		const uint32_t head = first >> 1;
		const uint64_t at = unitig_at[head];
		unitig_nodes[at + pos] = 2 * (uint32_t)i + same;
		node_unitig[i] = (uint32_t)unitig_id[head];
		node_ufw[i] = same;
		node_upos[i] = pos;
		if (is_start[i]) unitig_begin[unitig_id[i]] = at;
	}
}
// UnitigGraph::edges / edgeCov (the last loop of getUnitigGraph, MBG/MBG.cpp:746-760): every covered edge of the filtered
// list that is not a link inside a chain leaves the last k-mer of one unitig and enters the first k-mer of another (or of
// the same one: a circular unitig closes through the link the ranking cut).  flag / at = compaction as for the edges.
__device__ __forceinline__ bool unitig_edge_of(const EdgeRec& r, uint32_t min_cov, const uint32_t* __restrict__ link, const uint32_t* __restrict__ node_unitig,
	const uint32_t* __restrict__ node_ufw, EdgeRec& out)
{
	if (r.coverage < min_cov) return false;
	const uint32_t x = 2 * r.from + r.from_fw, y = 2 * r.to + r.to_fw;
	if (link[x] == y) return false;
	uint32_t fu = node_unitig[r.from], ffw = r.from_fw == node_ufw[r.from] ? 1u : 0u;
	uint32_t tu = node_unitig[r.to], tfw = r.to_fw == node_ufw[r.to] ? 1u : 0u;
	if (tu < fu || (tu == fu && !tfw && !ffw)) { const uint32_t a = fu, b = ffw; fu = tu; ffw = tfw ^ 1; tu = a; tfw = b ^ 1; }   // canon(), MBG/MBGCommon.cpp:78-89
	out.from = fu; out.from_fw = ffw; out.to = tu; out.to_fw = tfw; out.coverage = r.coverage; out.overlap = 0;
	return true;
}

__global__ void __launch_bounds__(256) k_unitig_edge_flags(const EdgeRec* __restrict__ edges, uint64_t n_edges, uint32_t min_cov, const uint32_t* __restrict__ link,
	const uint32_t* __restrict__ node_unitig, const uint32_t* __restrict__ node_ufw, uint32_t* __restrict__ flag)
{
	for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_edges; e += (uint64_t)gridDim.x * blockDim.x)
	{
		EdgeRec o;
		flag[e] = unitig_edge_of(edges[e], min_cov, link, node_unitig, node_ufw, o) ? 1u : 0u;
	}
}

__global__ void __launch_bounds__(256) k_unitig_edge_write(const EdgeRec* __restrict__ edges, uint64_t n_edges, uint32_t min_cov, const uint32_t* __restrict__ link,
	const uint32_t* __restrict__ node_unitig, const uint32_t* __restrict__ node_ufw, const uint64_t* __restrict__ at, EdgeRec* __restrict__ out)
{
	for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_edges; e += (uint64_t)gridDim.x * blockDim.x)
	{
		EdgeRec o;
		if (unitig_edge_of(edges[e], min_cov, link, node_unitig, node_ufw, o)) out[at[e]] = o;
	}
}


// ---- read threading: getKmerLocator (MBG/KmerMatcher.cpp:3-31) and iterateReadPaths
// (MBG/KmerMatcher.h:12-144) as a batched lookup over the selected k-mer instances that are still resident.
// The reference walks a part k-mer by k-mer with a small state (last k-mer, last unitig position); every decision it takes
// at a k-mer depends only on that k-mer and the previous one that is in the filtered list, so each instance classifies
// itself:
//   0  not in the filtered list                    (KmerMatcher.h:29: skipped, the state does not change)
//   1  in no unitig                                (:30-47: ends the open path; readPoses is NOT cleared)
//   2  first located k-mer after the part start or after a 1   (:55-67: starts a path, readPoses appended to)
//   3  starts a path after a break                 (:68-93 no sequence overlap / wrong distance, :121-141 no unitig edge)
//   4  next k-mer of the same unitig               (:95-104)
//   5  first k-mer of the next unitig of the path  (:105-120)
// inst_unitig = 2*unitig + orientation as the read sees it, inst_upos = position in that orientation, inst_urest = k-mers
// behind it (leftClip / rightClip of a path are inst_upos of its first and inst_urest of its last k-mer).
__global__ void __launch_bounds__(256) k_invert_orig(const uint32_t* __restrict__ orig, uint64_t n_kept, uint32_t* __restrict__ final_of)
{
	for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_kept; j += (uint64_t)gridDim.x * blockDim.x) final_of[orig[j]] = (uint32_t)j;
}

// thin != nullptr: the list is the one filterKmersToUnitigKmers leaves (MBG's default): k-mers it dropped are unknown to getNodeOrNull,
// positions and sizes count the k-mers that stayed (node_tpos, tscan = scan of the marks over the unitig positions)
__global__ void __launch_bounds__(256) k_thread_locate(const uint32_t* __restrict__ inst_slot, const uint64_t* __restrict__ inst_hash, uint64_t n_inst, const NodeSlot* __restrict__ table,
	const uint32_t* __restrict__ final_of, const uint32_t* __restrict__ node_unitig, const uint32_t* __restrict__ node_ufw, const uint32_t* __restrict__ node_upos,
	const uint64_t* __restrict__ unitig_begin, const uint8_t* __restrict__ thin, const uint32_t* __restrict__ node_tpos, const uint64_t* __restrict__ tscan,
	uint32_t* __restrict__ inst_node, uint32_t* __restrict__ inst_unitig, uint32_t* __restrict__ inst_upos, uint32_t* __restrict__ inst_urest)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_inst; i += (uint64_t)gridDim.x * blockDim.x)
	{
		uint32_t node = NO_LINK, unitig = NO_LINK, upos = 0, urest = 0;
		const uint32_t slot = inst_slot[i];
		if (slot != NO_LINK)
		{
			const uint32_t fid = final_of[(uint32_t)table[slot].rank];       // getNodeOrNull (MBG/HashList.cpp:85-101) on the filtered list
			if (fid != NO_LINK && (!thin || thin[fid]))
			{
				const uint64_t lo = inst_hash[2 * i], hi = inst_hash[2 * i + 1];
				uint64_t key_lo, key_hi; uint32_t fw;
				canonical_key(lo, hi, key_lo, key_hi, fw);
				node = 2 * fid + fw;
				const uint32_t u = node_unitig[fid];
				if (u != NO_LINK)
				{
					const uint32_t size = thin ? (uint32_t)(tscan[unitig_begin[u + 1]] - tscan[unitig_begin[u]]) : (uint32_t)(unitig_begin[u + 1] - unitig_begin[u]);
					const uint32_t o = node_ufw[fid] == fw ? 1u : 0u;
					const uint32_t j = thin ? node_tpos[fid] : node_upos[fid];
					unitig = 2 * u + o;
					upos = o ? j : size - 1 - j;
					urest = size - 1 - upos;
				}
			}
		}
		inst_node[i] = node; inst_unitig[i] = unitig; inst_upos[i] = upos; inst_urest[i] = urest;
	}
}

__device__ __forceinline__ bool edge_lookup(const EdgeSlot* __restrict__ table, uint64_t mask, unsigned long long key, uint32_t& count, uint32_t& overlap)
{
	uint64_t slot = slot_hash(key, key >> 17) & mask;
	for (uint32_t probes = 0; probes < MAX_PROBES; probes++)
	{
		const unsigned long long cur = table[slot].key;
		if (cur == key) { count = table[slot].count; overlap = (uint32_t)(table[slot].ordered & 0xffffff); return true; }
		if (cur == ~0ull) return false;
		slot = (slot + 1) & mask;
	}
	return false;
}

__global__ void __launch_bounds__(256) k_thread_steps(const uint32_t* __restrict__ inst_node, const uint32_t* __restrict__ inst_unitig, const uint32_t* __restrict__ inst_upos,
	const uint32_t* __restrict__ inst_urest, const uint8_t* __restrict__ inst_flag, const uint32_t* __restrict__ inst_pos, const uint32_t* __restrict__ inst_slot,
	const uint64_t* __restrict__ inst_hash, uint64_t n_inst, const EdgeSlot* __restrict__ edge_table, uint64_t edge_mask, uint32_t k, uint32_t min_cov,
	const uint32_t* __restrict__ node_off, uint8_t* __restrict__ step)
{
	// node_off != nullptr: thinned list.  Neighbours of a thinned unitig have the sequence overlap filterKmersToUnitigKmers gave them,
	// k minus the distance of their starts inside the unitig (it replaces an older one of the same pair, HashList::addSequenceOverlap);
	// every other pair has the overlap of its edge in the table, as before (edges between kept k-mers survive HashList::filter)
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_inst; i += (uint64_t)gridDim.x * blockDim.x)
	{
		uint32_t cls;
		if (inst_node[i] == NO_LINK) cls = 0;
		else if (inst_unitig[i] == NO_LINK) cls = 1;
		else
		{
			// the previous k-mer of the part that is in the filtered list
			uint64_t j = i;
			bool have = false;
			while (!(inst_flag[j] & 1)) { j--; if (inst_node[j] != NO_LINK) { have = true; break; } }
			if (!have || inst_unitig[j] == NO_LINK) cls = 2;
			else
			{
				cls = 3;
				uint32_t count = 0, overlap = 0;
				if (node_off && inst_unitig[i] == inst_unitig[j] && inst_upos[i] == inst_upos[j] + 1)
				{
					const uint32_t oi = node_off[inst_node[i] >> 1], oj = node_off[inst_node[j] >> 1];
					const uint32_t dist = oi > oj ? oi - oj : oj - oi;
					if (inst_pos[i] - inst_pos[j] == dist) cls = 4;
					step[i] = (uint8_t)cls;
					continue;
				}
				const unsigned long long key = canonical_edge_key(inst_hash[2 * j], inst_hash[2 * j + 1], inst_hash[2 * i], inst_hash[2 * i + 1], inst_slot[j], inst_slot[i]);
				if (edge_lookup(edge_table, edge_mask, key, count, overlap) && inst_pos[i] - inst_pos[j] == k - overlap)
				{
					if (inst_unitig[i] == inst_unitig[j] && inst_upos[i] == inst_upos[j] + 1) cls = 4;
					else if (count >= min_cov && inst_upos[i] == 0 && inst_urest[j] == 0) cls = 5;
				}
			}
		}
		step[i] = (uint8_t)cls;
	}
}


// ---- filterKmersToUnitigKmers (MBG/MBG.cpp:1065-1185, filterWithinUnitig = true, MBG's default): which k-mers of a unitig
// stay.  The reference walks a unitig from its smaller-hash end (the orientation the unitigs already have here) and keeps
// a k-mer when it is an end, a tip k-mer, or when dropping it would leave the next k-mer without overlap to the last kept
// one.  Which k-mer is kept next depends only on the last kept one: next(p) = first t > p that is the last k-mer, a tip
// k-mer, or has pos(t + 1) >= pos(p) + k.  The kept set is the orbit of the unitig head under next(), marked by pointer
// doubling: after round r every k-mer within 2^r steps of the head is marked.
__device__ __forceinline__ uint32_t list_overlap(const EdgeRec* __restrict__ edges, uint64_t n_edges, uint32_t x, uint32_t y)
{
	uint32_t from = x >> 1, ffw = x & 1, to = y >> 1, tfw = y & 1;
	if (to < from || (to == from && !tfw && !ffw)) { const uint32_t a = from, b = ffw; from = to; ffw = !tfw; to = a; tfw = !b; }    // canon(), MBG/MBGCommon.cpp:78-89
	EdgeRec want; want.from = from; want.from_fw = ffw; want.to = to; want.to_fw = tfw; want.coverage = 0; want.overlap = 0;
	const uint64_t key = edge_order_key(want);
	uint64_t lo = 0, hi = n_edges;
	while (lo < hi) { const uint64_t mid = (lo + hi) >> 1; if (edge_order_key(edges[mid]) < key) lo = mid + 1; else hi = mid; }
	return (lo < n_edges && edge_order_key(edges[lo]) == key) ? edges[lo].overlap : 0u;
}

// step of every unitig position: k - overlap with the k-mer before it (0 at a unitig head)
__global__ void __launch_bounds__(256) k_thin_delta(const uint32_t* __restrict__ unitig_nodes, uint64_t n_pos, const uint32_t* __restrict__ node_unitig, const uint64_t* __restrict__ unitig_begin,
	const EdgeRec* __restrict__ edges, uint64_t n_edges, uint32_t k, uint32_t* __restrict__ delta)
{
	for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pos; p += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint32_t x = unitig_nodes[p];
		const uint64_t b = unitig_begin[node_unitig[x >> 1]];
		delta[p] = p == b ? 0u : k - list_overlap(edges, n_edges, unitig_nodes[p - 1], x);
	}
}

// next(p), the head marks, and the position of every k-mer inside its unitig (currentPos of the reference's walk)
__global__ void __launch_bounds__(256) k_thin_next(const uint32_t* __restrict__ unitig_nodes, uint64_t n_pos, const uint32_t* __restrict__ node_unitig, const uint64_t* __restrict__ unitig_begin,
	const uint8_t* __restrict__ tip, const uint32_t* __restrict__ delta, const uint64_t* __restrict__ scan, uint32_t k, uint32_t* __restrict__ next, uint32_t* __restrict__ mark,
	uint32_t* __restrict__ node_pos)
{
	for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pos; p += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint32_t x = unitig_nodes[p];
		const uint32_t u = node_unitig[x >> 1];
		const uint64_t b = unitig_begin[u], last = unitig_begin[u + 1] - 1;
		const uint64_t base = scan[b];                                  // delta[b] == 0
		const uint64_t here = scan[p] + delta[p] - base;
		node_pos[x >> 1] = (uint32_t)here;
		mark[p] = p == b ? 1u : 0u;
		uint64_t t = p;
		if (p < last)
			for (t = p + 1; t < last; t++)
			{
				if (tip[unitig_nodes[t] >> 1]) break;
				if (scan[t + 1] + delta[t + 1] - base >= here + k) break;
			}
		next[p] = (uint32_t)t;
	}
}

__global__ void __launch_bounds__(256) k_thin_mark(const uint32_t* __restrict__ jump, uint64_t n_pos, uint32_t* mark)
{
	for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pos; p += (uint64_t)gridDim.x * blockDim.x)
		if (((volatile uint32_t*)mark)[p]) mark[jump[p]] = 1;          // whatever gets marked is on the orbit
}

__global__ void __launch_bounds__(256) k_thin_double(const uint32_t* __restrict__ in, uint64_t n_pos, uint32_t* __restrict__ out)
{
	for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pos; p += (uint64_t)gridDim.x * blockDim.x) out[p] = in[in[p]];
}

__global__ void __launch_bounds__(256) k_thin_flags(const uint32_t* __restrict__ unitig_nodes, uint64_t n_pos, const uint32_t* __restrict__ mark, uint8_t* __restrict__ thin)
{
	for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pos; p += (uint64_t)gridDim.x * blockDim.x) thin[unitig_nodes[p] >> 1] = (uint8_t)mark[p];
}


// ---- applying the thinning (MBG/MBG.cpp:1166-1184): reads.filter(kept), the unitigs renumbered, and a sequence overlap for
// every pair of neighbours of a thinned unitig.  tscan = exclusive scan of the marks over the unitig positions.
__global__ void __launch_bounds__(256) k_thin_kept32(const uint8_t* __restrict__ thin, uint64_t n, uint32_t* __restrict__ kept32)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) kept32[i] = thin[i] ? 1u : 0u;
}

// thinned position of every kept k-mer inside its unitig, and the thinned unitigs as flat arrays: nodes (new ids), offsets in codes
// and the unitig each thinned position belongs to
__global__ void __launch_bounds__(256) k_thin_compact(const uint32_t* __restrict__ unitig_nodes, uint64_t n_pos, const uint32_t* __restrict__ mark, const uint64_t* __restrict__ tscan,
	const uint32_t* __restrict__ node_unitig, const uint64_t* __restrict__ unitig_begin, const uint64_t* __restrict__ newid, const uint32_t* __restrict__ node_off,
	uint32_t* __restrict__ node_tpos, uint32_t* __restrict__ t_nodes, uint32_t* __restrict__ t_off, uint32_t* __restrict__ t_unitig)
{
	for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pos; p += (uint64_t)gridDim.x * blockDim.x)
	{
		if (!mark[p]) continue;
		const uint32_t x = unitig_nodes[p], fid = x >> 1;
		const uint32_t u = node_unitig[fid];
		const uint64_t tq = tscan[p];
		node_tpos[fid] = (uint32_t)(tq - tscan[unitig_begin[u]]);
		t_nodes[tq] = 2 * (uint32_t)newid[fid] + (x & 1);
		t_off[tq] = node_off[fid];
		t_unitig[tq] = u;
	}
}

__global__ void __launch_bounds__(256) k_thin_begins(const uint64_t* __restrict__ unitig_begin, uint64_t n_unitigs, const uint64_t* __restrict__ tscan, uint64_t* __restrict__ t_begin)
{
	for (uint64_t u = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; u <= n_unitigs; u += (uint64_t)gridDim.x * blockDim.x) t_begin[u] = tscan[unitig_begin[u]];
}

// newOverlaps (MBG/MBG.cpp:1136-1141): neighbours of a thinned unitig overlap by k minus the distance of their starts; canon() form
__global__ void __launch_bounds__(256) k_thin_overlaps(const uint32_t* __restrict__ t_nodes, const uint32_t* __restrict__ t_off, const uint32_t* __restrict__ t_unitig,
	const uint64_t* __restrict__ t_begin, uint64_t n_tpos, uint32_t k, EdgeRec* __restrict__ out)
{
	for (uint64_t tq = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; tq < n_tpos; tq += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint32_t u = t_unitig[tq];
		if (tq == t_begin[u]) continue;
		const uint32_t x = t_nodes[tq - 1], y = t_nodes[tq];
		uint32_t from = x >> 1, ffw = x & 1, to = y >> 1, tfw = y & 1;
		if (to < from || (to == from && !tfw && !ffw)) { const uint32_t a = from, b = ffw; from = to; ffw = !tfw; to = a; tfw = !b; }
		EdgeRec r; r.from = from; r.from_fw = ffw; r.to = to; r.to_fw = tfw; r.coverage = 0; r.overlap = k - (t_off[tq] - t_off[tq - 1]);
		out[tq - (u + 1)] = r;
	}
}

}
