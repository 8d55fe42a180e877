/* oracle/mbg_unitig_oracle.c -- TEST INFRASTRUCTURE ONLY, never linked into or called by the product.
 *
 * Plain-C sequential restatement of the step right after the k-mer table in MBG:
 *   getCoveredEdges                      MBG/MBG.cpp:367-388   edges with coverage >= -a, both directions, deduplicated
 *                                        (SparseEdgeContainer::addEdge has set semantics, MBG/SparseEdgeContainer.cpp:12-24)
 *   getUnitigHashes / checkUnitigHashes  MBG/MBG.cpp:305-344   maximal non-branching chain; kept iff mean coverage >= -u
 *   getUnitigGraph, first block          MBG/MBG.cpp:629-696   which chains are visited (keepGaps = false)
 *   removeOnecovNodes                    MBG/MBG.cpp:527-627   oneCovHeuristic (on under Verkko's flags: -u 2, -r > 0, guesswork)
 *   HashList::filter                     MBG/HashList.cpp:159-232  kept nodes renumbered in order, edges between kept nodes
 *   getUnitigGraph, second block         MBG/MBG.cpp:689-760   unitigs = the same chains on the filtered list, coverage >= -a
 * Input is the table mbg_oracle.c builds (node order = ascending canonical hash).  The reference runs this on the
 * arrival-ordered list; the kept set and the chains do not depend on the order, the unitig numbering does (tests compare
 * order-normalised).  Pinned against the reference's own getUnitigGraph in tests/test_oracle.py.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "mbg_oracle.h"

typedef struct {
	uint64_t n;            /* nodes */
	uint64_t* begin;       /* 2n + 1: successors of oriented node x = 2*id + fw (SparseEdgeContainer::pairToInt) */
	uint32_t* to;
} adjacency;

static int cmp_u64(const void* a, const void* b) { uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b; return x < y ? -1 : x > y; }

/* getCoveredEdges: edges = u32 x6 records (from, from_fw, to, to_fw, coverage, overlap) */
static adjacency covered_edges(uint64_t n_nodes, const uint32_t* edges, uint64_t n_edges, uint32_t min_cov)
{
	adjacency a;
	a.n = n_nodes;
	uint64_t* pairs = (uint64_t*)malloc((2 * n_edges + 1) * 8);
	uint64_t np = 0;
	for (uint64_t e = 0; e < n_edges; e++)
	{
		const uint32_t* r = edges + 6 * e;
		if (r[4] < min_cov) continue;
		uint64_t x = 2ull * r[0] + r[1], y = 2ull * r[2] + r[3];
		pairs[np++] = (x << 32) | y;
		pairs[np++] = ((y ^ 1) << 32) | (x ^ 1);
	}
	qsort(pairs, np, 8, cmp_u64);
	uint64_t nu = 0;
	for (uint64_t i = 0; i < np; i++) if (i == 0 || pairs[i] != pairs[nu - 1]) pairs[nu++] = pairs[i];
	a.begin = (uint64_t*)calloc(2 * n_nodes + 2, 8);
	a.to = (uint32_t*)malloc((nu + 1) * 4);
	for (uint64_t i = 0; i < nu; i++) { a.begin[(pairs[i] >> 32) + 1]++; a.to[i] = (uint32_t)pairs[i]; }
	for (uint64_t x = 0; x < 2 * n_nodes; x++) a.begin[x + 1] += a.begin[x];
	free(pairs);
	return a;
}
static void adjacency_free(adjacency* a) { free(a->begin); free(a->to); }
static uint64_t deg(const adjacency* a, uint32_t x) { return a->begin[x + 1] - a->begin[x]; }
static uint32_t first(const adjacency* a, uint32_t x) { return a->to[a->begin[x]]; }

/* getUnitigHashes (MBG/MBG.cpp:305-320): returns the chain length, nodes written to out */
static uint64_t unitig_hashes(const adjacency* a, uint32_t start, uint32_t* out)
{
	uint64_t n = 0;
	out[n++] = start;
	for (;;)
	{
		uint32_t back = out[n - 1];
		if (deg(a, back) != 1) break;
		uint32_t next = first(a, back);
		if (deg(a, next ^ 1) != 1) break;
		if (next == start) break;
		if ((next >> 1) == (back >> 1)) break;
		out[n++] = next;
	}
	return n;
}

/* checkUnitigHashes (MBG/MBG.cpp:322-344) */
static void check_unitig(const adjacency* a, uint32_t start, uint8_t* checked, const uint32_t* cov, double min_unitig_cov, uint8_t* kept, uint32_t* scratch)
{
	uint64_t n = unitig_hashes(a, start, scratch);
	double total = 0;
	for (uint64_t i = 0; i < n; i++) { checked[scratch[i] >> 1] = 1; total += cov[scratch[i] >> 1]; }
	total /= (double)n;
	if (total >= min_unitig_cov) for (uint64_t i = 0; i < n; i++) kept[scratch[i] >> 1] = 1;
}

struct orc_unitigs {
	uint64_t n_nodes, n_kept, n_fedges, n_unitigs, n_unitig_nodes;
	uint8_t* kept;          /* per node of the input table */
	uint64_t* f_hash;       /* lo, hi per kept node */
	uint32_t* f_cov;
	uint8_t* f_tip;
	uint32_t* f_edges;      /* u32 x6 with renumbered ids */
	uint64_t* unitig_begin; /* n_unitigs + 1 */
	uint32_t* unitig_nodes; /* 2*filtered id + fw, in traversal order */
	uint64_t n_uedges;
	uint32_t* uedges;       /* u32 x6: from unitig, from_fw, to unitig, to_fw, coverage, 0 -- canonical (MBG/MBGCommon.cpp:78-89) */
};

void orc_unitigs_free(orc_unitigs* u)
{
	if (!u) return;
	free(u->kept); free(u->f_hash); free(u->f_cov); free(u->f_tip); free(u->f_edges); free(u->unitig_begin); free(u->unitig_nodes); free(u->uedges);
	free(u);
}

/* getUnitigGraph's second block: which oriented node starts a unitig, in the reference's loop order */
static void start_unitig(const adjacency* a, uint32_t start, uint8_t* belongs, orc_unitigs* u, uint32_t* scratch)
{
	uint64_t n = unitig_hashes(a, start, scratch);
	for (uint64_t i = 0; i < n; i++) { belongs[scratch[i] >> 1] = 1; u->unitig_nodes[u->n_unitig_nodes++] = scratch[i]; }
	u->unitig_begin[++u->n_unitigs] = u->n_unitig_nodes;
}

/* a HashList as this step sees it: nodes in order, edges as u32 x6 records, and where each node sat in the input table */
typedef struct { uint64_t n, m; uint64_t* hash; uint32_t* cov; uint8_t* tip; uint32_t* edges; uint32_t* orig; } hlist;

static hlist list_copy(uint64_t n, uint64_t m, const uint64_t* hash, const uint32_t* cov, const uint8_t* tip, const uint32_t* edges)
{
	hlist l;
	l.n = n; l.m = m;
	l.hash = (uint64_t*)malloc((n + 1) * 16); memcpy(l.hash, hash, n * 16);
	l.cov = (uint32_t*)malloc((n + 1) * 4); memcpy(l.cov, cov, n * 4);
	l.tip = (uint8_t*)malloc(n + 1); memcpy(l.tip, tip, n);
	l.edges = (uint32_t*)malloc((m + 1) * 24); memcpy(l.edges, edges, m * 24);
	l.orig = (uint32_t*)malloc((n + 1) * 4);
	for (uint64_t i = 0; i < n; i++) l.orig[i] = (uint32_t)i;
	return l;
}

/* HashList::filter (MBG/HashList.cpp:159-232): kept nodes renumbered in order, edges between kept nodes.  In place. */
static void list_filter(hlist* l, const uint8_t* kept)
{
	uint32_t* rank = (uint32_t*)malloc((l->n + 1) * 4);
	uint64_t nk = 0;
	for (uint64_t i = 0; i < l->n; i++) { rank[i] = (uint32_t)nk; if (kept[i]) nk++; }
	for (uint64_t i = 0; i < l->n; i++) if (kept[i])
	{
		const uint32_t j = rank[i];
		l->hash[2 * j] = l->hash[2 * i]; l->hash[2 * j + 1] = l->hash[2 * i + 1];
		l->cov[j] = l->cov[i]; l->tip[j] = l->tip[i]; l->orig[j] = l->orig[i];
	}
	uint64_t mk = 0;
	for (uint64_t e = 0; e < l->m; e++)
	{
		const uint32_t* r = l->edges + 6 * e;
		if (!kept[r[0]] || !kept[r[2]]) continue;
		uint32_t* w = l->edges + 6 * mk++;
		memmove(w, r, 24);
		w[0] = rank[w[0]]; w[2] = rank[w[2]];
	}
	l->n = nk; l->m = mk;
	free(rank);
}

/* extendOneCoverageKmers (MBG/MBG.cpp:527-546) */
static uint32_t extend_onecov(const adjacency* a, const uint32_t* cov, uint32_t pos)
{
	const uint32_t start = pos;
	for (;;)
	{
		if (deg(a, pos) != 1) return pos;
		const uint32_t next = first(a, pos);
		if (deg(a, next ^ 1) != 1) return pos;
		if ((next >> 1) == (pos >> 1)) return pos;
		if ((next >> 1) == (start >> 1)) return pos;
		if (cov[next >> 1] != 1) return pos;
		pos = next;
	}
}

/* the neighbour test of removeOnecovNodes (MBG/MBG.cpp:583-612): everything this end leads to is also reached from a
 * k-mer of coverage >= 2 */
static int onecov_end_replaceable(const adjacency* a, const uint32_t* cov, uint32_t end)
{
	for (uint64_t j = a->begin[end]; j < a->begin[end + 1]; j++)
	{
		const uint32_t back = a->to[j] ^ 1;
		int other = 0;
		for (uint64_t j2 = a->begin[back]; j2 < a->begin[back + 1]; j2++)
		{
			const uint32_t e2 = a->to[j2];
			if (e2 == (end ^ 1)) continue;
			if (cov[e2 >> 1] >= 2) other = 1;
		}
		if (!other) return 0;
	}
	return 1;
}

/* removeOnecovNodes (MBG/MBG.cpp:548-627) */
static void remove_onecov(hlist* l, int only_tips)
{
	const uint64_t n = l->n;
	adjacency a = covered_edges(n, l->edges, l->m, 1);
	uint8_t* kept = (uint8_t*)malloc(n + 1);
	uint8_t* checked = (uint8_t*)calloc(n + 1, 1);
	memset(kept, 1, n + 1);
	int removed_any = 0;
	for (uint64_t i = 0; i < n; i++)
	{
		if (l->cov[i] != 1 || checked[i]) continue;
		const uint32_t fw_node = extend_onecov(&a, l->cov, (uint32_t)(2 * i + 1));
		const uint32_t bw_node = extend_onecov(&a, l->cov, (uint32_t)(2 * i));
		for (uint32_t c = bw_node ^ 1; c != fw_node; c = first(&a, c)) checked[c >> 1] = 1;
		checked[fw_node >> 1] = 1;
		if (l->cov[fw_node >> 1] != 1 || l->cov[bw_node >> 1] != 1) continue;
		if (only_tips)
		{
			if (deg(&a, fw_node) != 0 && deg(&a, bw_node) != 0) continue;
			if (deg(&a, fw_node) == 0 && deg(&a, bw_node) == 0) continue;
		}
		if (!onecov_end_replaceable(&a, l->cov, fw_node)) continue;
		if (!onecov_end_replaceable(&a, l->cov, bw_node)) continue;
		uint32_t c = bw_node ^ 1;
		for (; c != fw_node; c = first(&a, c)) kept[c >> 1] = 0;
		kept[c >> 1] = 0;
		removed_any = 1;
	}
	adjacency_free(&a);
	if (removed_any) list_filter(l, kept);
	free(kept); free(checked);
}

orc_unitigs* orc_build_unitigs_ex(const orc* o, uint32_t min_cov, double min_unitig_cov, int one_cov_heuristic)
{
	const uint64_t n_in = orc_count(o, 4);
	hlist l = list_copy(n_in, orc_count(o, 5), (const uint64_t*)orc_data(o, 5), (const uint32_t*)orc_data(o, 6), (const uint8_t*)orc_data(o, 7), (const uint32_t*)orc_data(o, 8));
	orc_unitigs* u = (orc_unitigs*)calloc(1, sizeof(orc_unitigs));
	u->n_nodes = n_in;
	u->kept = (uint8_t*)calloc(n_in + 1, 1);
	/* ---- MBG/MBG.cpp:631-635 */
	if (one_cov_heuristic) { remove_onecov(&l, 1); remove_onecov(&l, 0); }
	const uint64_t n = l.n, m = l.m;
	const uint32_t* cov = l.cov;
	const uint32_t* edges = l.edges;
	uint8_t* kept = (uint8_t*)calloc(n + 1, 1);
	uint8_t* checked = (uint8_t*)calloc(n + 1, 1);
	uint32_t* scratch = (uint32_t*)malloc((n + 1) * 4);
	/* ---- first block, MBG/MBG.cpp:637-690 */
	adjacency a = covered_edges(n, edges, m, min_cov);
	for (uint64_t i = 0; i < n; i++)
	{
		uint32_t fw = (uint32_t)(2 * i + 1), bw = (uint32_t)(2 * i);
		if (deg(&a, bw) != 1 || first(&a, bw) == fw)
		{
			if (!checked[i]) check_unitig(&a, fw, checked, cov, min_unitig_cov, kept, scratch);
			for (uint64_t j = a.begin[bw]; j < a.begin[bw + 1]; j++)
				if (!checked[a.to[j] >> 1]) check_unitig(&a, a.to[j], checked, cov, min_unitig_cov, kept, scratch);
		}
		if (deg(&a, fw) != 1 || first(&a, fw) == bw)
		{
			if (!checked[i]) check_unitig(&a, bw, checked, cov, min_unitig_cov, kept, scratch);
			for (uint64_t j = a.begin[fw]; j < a.begin[fw + 1]; j++)
				if (!checked[a.to[j] >> 1]) check_unitig(&a, a.to[j], checked, cov, min_unitig_cov, kept, scratch);
		}
	}
	for (uint64_t i = 0; i < n; i++) if (!checked[i]) check_unitig(&a, (uint32_t)(2 * i + 1), checked, cov, min_unitig_cov, kept, scratch);
	adjacency_free(&a);
	/* ---- HashList::filter, MBG/HashList.cpp:159-232 */
	list_filter(&l, kept);
	free(kept);
	for (uint64_t i = 0; i < l.n; i++) u->kept[l.orig[i]] = 1;
	u->n_kept = l.n; u->n_fedges = l.m;
	u->f_hash = l.hash; u->f_cov = l.cov; u->f_tip = l.tip; u->f_edges = l.edges;
	free(l.orig);
	/* ---- second block, MBG/MBG.cpp:689-745, on the filtered list */
	const uint64_t nk = u->n_kept;
	a = covered_edges(nk, u->f_edges, u->n_fedges, min_cov);
	uint8_t* belongs = (uint8_t*)calloc(nk + 1, 1);
	u->unitig_begin = (uint64_t*)calloc(nk + 2, 8);
	u->unitig_nodes = (uint32_t*)malloc((nk + 1) * 4);
	for (uint64_t i = 0; i < nk; i++)
	{
		if (u->f_cov[i] < min_cov) continue;
		uint32_t fw = (uint32_t)(2 * i + 1), bw = (uint32_t)(2 * i);
		if (deg(&a, bw) != 1 || first(&a, bw) == fw)
		{
			if (!belongs[i]) start_unitig(&a, fw, belongs, u, scratch);
			for (uint64_t j = a.begin[bw]; j < a.begin[bw + 1]; j++) if (!belongs[a.to[j] >> 1]) start_unitig(&a, a.to[j], belongs, u, scratch);
		}
		if (deg(&a, fw) != 1 || first(&a, fw) == bw)
		{
			if (!belongs[i]) start_unitig(&a, bw, belongs, u, scratch);
			for (uint64_t j = a.begin[fw]; j < a.begin[fw + 1]; j++) if (!belongs[a.to[j] >> 1]) start_unitig(&a, a.to[j], belongs, u, scratch);
		}
	}
	for (uint64_t i = 0; i < nk; i++)
	{
		if (belongs[i] || u->f_cov[i] < min_cov) continue;
		start_unitig(&a, (uint32_t)(2 * i + 1), belongs, u, scratch);
	}
	/* ---- unitig edges, MBG/MBG.cpp:746-760: from every unitig tip along the covered edges of its end k-mer */
	{
		uint32_t* tip_of = (uint32_t*)malloc((2 * nk + 2) * 4);      /* oriented node -> 2*unitig + fw, or none */
		for (uint64_t x = 0; x < 2 * nk + 2; x++) tip_of[x] = 0xffffffffu;
		for (uint64_t t = 0; t < u->n_unitigs; t++)
		{
			uint32_t first_node = u->unitig_nodes[u->unitig_begin[t]], last_node = u->unitig_nodes[u->unitig_begin[t + 1] - 1];
			tip_of[last_node] = (uint32_t)(2 * t + 1);
			tip_of[first_node ^ 1] = (uint32_t)(2 * t);
		}
		uint64_t cap = 2 * u->n_fedges + 1;
		uint64_t* keys = (uint64_t*)malloc(cap * 8);
		uint32_t* covs = (uint32_t*)malloc(cap * 4);
		uint64_t nkey = 0;
		for (uint64_t e = 0; e < u->n_fedges; e++)
		{
			const uint32_t* r = u->f_edges + 6 * e;
			if (r[4] < min_cov) continue;
			/* the edge as stored and its mirror image: each is seen from the tip it leaves */
			for (int side = 0; side < 2; side++)
			{
				uint32_t x = 2 * r[0] + r[1], y = 2 * r[2] + r[3];
				if (side) { uint32_t t = x; x = y ^ 1; y = t ^ 1; }
				if (tip_of[x] == 0xffffffffu) continue;
				uint32_t from = tip_of[x], to = tip_of[y ^ 1] ^ 1;
				uint32_t fu = from >> 1, ffw = from & 1, tu = to >> 1, tfw = to & 1;
				if (tu < fu || (tu == fu && !tfw && !ffw)) { uint32_t a1 = fu, b1 = ffw; fu = tu; ffw = !tfw; tu = a1; tfw = !b1; }
				keys[nkey] = ((uint64_t)fu << 34) | ((uint64_t)ffw << 33) | ((uint64_t)tu << 1) | tfw;
				covs[nkey++] = r[4];
			}
		}
		/* set semantics: sort, drop duplicates (both tips report the same unitig edge) */
		uint64_t* order = (uint64_t*)malloc((nkey + 1) * 16);
		for (uint64_t i = 0; i < nkey; i++) { order[2 * i] = keys[i]; order[2 * i + 1] = covs[i]; }
		qsort(order, nkey, 16, cmp_u64);
		u->uedges = (uint32_t*)malloc((nkey + 1) * 24);
		for (uint64_t i = 0; i < nkey; i++)
		{
			if (i && order[2 * i] == order[2 * (i - 1)]) continue;
			uint32_t* w = u->uedges + 6 * u->n_uedges++;
			w[0] = (uint32_t)(order[2 * i] >> 34); w[1] = (uint32_t)((order[2 * i] >> 33) & 1);
			w[2] = (uint32_t)((order[2 * i] >> 1) & 0xffffffffu); w[3] = (uint32_t)(order[2 * i] & 1);
			w[4] = (uint32_t)order[2 * i + 1]; w[5] = 0;
		}
		free(order); free(keys); free(covs); free(tip_of);
	}
	adjacency_free(&a);
	free(belongs); free(checked); free(scratch);
	return u;
}

orc_unitigs* orc_build_unitigs(const orc* o, uint32_t min_cov, double min_unitig_cov) { return orc_build_unitigs_ex(o, min_cov, min_unitig_cov, 0); }

/* what: 0 nodes of the input table, 1 kept nodes, 2 filtered edges, 3 unitigs, 4 unitig nodes */
uint64_t orc_unitigs_count(const orc_unitigs* u, int what)
{
	switch (what) { case 0: return u->n_nodes; case 1: return u->n_kept; case 2: return u->n_fedges; case 3: return u->n_unitigs; case 4: return u->n_unitig_nodes; case 5: return u->n_uedges; }
	return 0;
}
/* what: 0 kept(u8 per input node) 1 filtered hashes(u64 lo,hi) 2 filtered coverage(u32) 3 filtered tip(u8)
 *       4 filtered edges(u32 x6) 5 unitig_begin(u64, n+1) 6 unitig nodes(u32 = 2*filtered id + fw) */
const void* orc_unitigs_data(const orc_unitigs* u, int what)
{
	switch (what) { case 0: return u->kept; case 1: return u->f_hash; case 2: return u->f_cov; case 3: return u->f_tip; case 4: return u->f_edges; case 5: return u->unitig_begin; case 6: return u->unitig_nodes; case 7: return u->uedges; }
	return 0;
}
