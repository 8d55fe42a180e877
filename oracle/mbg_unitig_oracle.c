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
	uint32_t* orig;         /* input-table index of every kept node */
	uint64_t n_overlaps;    /* HashList::sequenceOverlap of the filtered list: u32 x5 from, from_fw, to, to_fw, overlap (canonical) */
	uint32_t* overlaps;
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
	free(u->kept); free(u->f_hash); free(u->f_cov); free(u->f_tip); free(u->f_edges); free(u->unitig_begin); free(u->unitig_nodes); free(u->uedges); free(u->orig); free(u->overlaps);
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
	u->orig = l.orig;
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
	u->n_overlaps = u->n_fedges;
	u->overlaps = (uint32_t*)malloc((u->n_fedges + 1) * 20);
	for (uint64_t e = 0; e < u->n_fedges; e++) { memcpy(u->overlaps + 5 * e, u->f_edges + 6 * e, 16); u->overlaps[5 * e + 4] = u->f_edges[6 * e + 5]; }
	free(belongs); free(checked); free(scratch);
	return u;
}

orc_unitigs* orc_build_unitigs(const orc* o, uint32_t min_cov, double min_unitig_cov) { return orc_build_unitigs_ex(o, min_cov, min_unitig_cov, 0); }

/* what: 0 nodes of the input table, 1 kept nodes, 2 filtered edges, 3 unitigs, 4 unitig nodes */
uint64_t orc_unitigs_count(const orc_unitigs* u, int what)
{
	switch (what) { case 0: return u->n_nodes; case 1: return u->n_kept; case 2: return u->n_fedges; case 3: return u->n_unitigs; case 4: return u->n_unitig_nodes; case 5: return u->n_uedges; case 6: return u->n_overlaps; }
	return 0;
}
/* what: 0 kept(u8 per input node) 1 filtered hashes(u64 lo,hi) 2 filtered coverage(u32) 3 filtered tip(u8)
 *       4 filtered edges(u32 x6) 5 unitig_begin(u64, n+1) 6 unitig nodes(u32 = 2*filtered id + fw) */
const void* orc_unitigs_data(const orc_unitigs* u, int what)
{
	switch (what) { case 0: return u->kept; case 1: return u->f_hash; case 2: return u->f_cov; case 3: return u->f_tip; case 4: return u->f_edges; case 5: return u->unitig_begin; case 6: return u->unitig_nodes; case 7: return u->uedges; case 8: return u->overlaps; }
	return 0;
}

/* ---- filterKmersToUnitigKmers (MBG/MBG.cpp:1065-1185) with filterWithinUnitig = true, MBG's default: inside a unitig only
 * its ends, tip k-mers and the k-mers needed to keep consecutive kept k-mers overlapping stay; the list is filtered again
 * (HashList::filter) and consecutive kept k-mers get a sequence overlap of their own.  Unitig edges are untouched (ends stay).
 * canon_key / edge_index / edge_find are defined with the read threading below. */
static uint64_t canon_key(uint32_t from, uint32_t ffw, uint32_t to, uint32_t tfw);
static uint64_t* overlap_index(const uint32_t* ov, uint64_t m);
static int edge_find(const uint64_t* ix, uint64_t m, uint64_t key, uint64_t* value);

void orc_thin_unitigs(orc_unitigs* u, uint32_t k)
{
	const uint64_t n = u->n_kept;
	uint64_t* ix = overlap_index(u->overlaps, u->n_overlaps);
	uint8_t* kept = (uint8_t*)calloc(n + 1, 1);
	uint32_t* new_nodes = (uint32_t*)malloc((u->n_unitig_nodes + 1) * 4);
	uint64_t* new_begin = (uint64_t*)calloc(u->n_unitigs + 2, 8);
	uint32_t* add = (uint32_t*)malloc((u->n_unitig_nodes + 1) * 20);     /* new overlaps: from, from_fw, to, to_fw, overlap (old ids) */
	uint64_t n_add = 0, n_new = 0;
	uint32_t* seq = (uint32_t*)malloc((u->n_unitig_nodes + 1) * 4);
	for (uint64_t t = 0; t < u->n_unitigs; t++)
	{
		const uint64_t b = u->unitig_begin[t], len = u->unitig_begin[t + 1] - b;
		/* walked from the end whose k-mer has the smaller hash (MBG.cpp:1069-1095); ids here are in hash order */
		const int rev = (u->unitig_nodes[b + len - 1] >> 1) < (u->unitig_nodes[b] >> 1);
		for (uint64_t j = 0; j < len; j++) seq[j] = rev ? (u->unitig_nodes[b + len - 1 - j] ^ 1) : u->unitig_nodes[b + j];
		uint64_t last_start = 0, pos = 0, last_kmer = ~0ull;
		const uint64_t out0 = n_new;
		for (uint64_t j = 0; j < len; j++)
		{
			uint64_t ov = 0;
			if (j > 0) { edge_find(ix, u->n_overlaps, canon_key(seq[j - 1] >> 1, seq[j - 1] & 1, seq[j] >> 1, seq[j] & 1), &ov); pos += k - ov; }
			int skip = 1;
			if (u->f_tip[seq[j] >> 1]) skip = 0;
			else if (j == 0 || j == len - 1) skip = 0;
			else
			{
				uint64_t ov2 = 0;
				edge_find(ix, u->n_overlaps, canon_key(seq[j] >> 1, seq[j] & 1, seq[j + 1] >> 1, seq[j + 1] & 1), &ov2);
				if (pos + (k - ov2) >= last_start + k) skip = 0;
			}
			if (skip) continue;
			kept[seq[j] >> 1] = 1;
			if (last_kmer != ~0ull)
			{
				uint32_t* w = add + 5 * n_add++;
				w[0] = (uint32_t)(last_kmer >> 1); w[1] = (uint32_t)(last_kmer & 1); w[2] = seq[j] >> 1; w[3] = seq[j] & 1; w[4] = (uint32_t)(k - (pos - last_start));
			}
			new_nodes[n_new++] = seq[j];
			last_start = pos; last_kmer = seq[j];
		}
		if (rev) for (uint64_t i = out0, j = n_new - 1; i < j; i++, j--) { uint32_t x = new_nodes[i] ^ 1; new_nodes[i] = new_nodes[j] ^ 1; new_nodes[j] = x; }
		if (rev && ((n_new - out0) & 1)) new_nodes[out0 + (n_new - out0) / 2] ^= 1;
		new_begin[t + 1] = n_new;
	}
	free(ix); free(seq);
	/* ranks, HashList::filter, renumbering */
	uint32_t* rank = (uint32_t*)malloc((n + 1) * 4);
	uint64_t nk = 0;
	for (uint64_t i = 0; i < n; i++) { rank[i] = (uint32_t)nk; if (kept[i]) nk++; }
	for (uint64_t i = 0; i < n; i++) if (kept[i])
	{
		const uint32_t j = rank[i];
		u->f_hash[2 * j] = u->f_hash[2 * i]; u->f_hash[2 * j + 1] = u->f_hash[2 * i + 1]; u->f_cov[j] = u->f_cov[i]; u->f_tip[j] = u->f_tip[i]; u->orig[j] = u->orig[i];
	}
	memset(u->kept, 0, u->n_nodes);
	for (uint64_t j = 0; j < nk; j++) u->kept[u->orig[j]] = 1;
	uint64_t mk = 0;
	for (uint64_t e = 0; e < u->n_fedges; e++)
	{
		uint32_t* r = u->f_edges + 6 * e;
		if (!kept[r[0]] || !kept[r[2]]) continue;
		uint32_t* w = u->f_edges + 6 * mk++;
		memmove(w, r, 24); w[0] = rank[w[0]]; w[2] = rank[w[2]];
	}
	u->n_fedges = mk;
	/* sequence overlaps: the surviving old ones, then the new ones (set semantics on the canonical pair: a new value replaces) */
	uint64_t mo = 0;
	uint32_t* ov_all = (uint32_t*)malloc((u->n_overlaps + n_add + 1) * 20);
	for (uint64_t e = 0; e < u->n_overlaps; e++)
	{
		const uint32_t* r = u->overlaps + 5 * e;
		if (!kept[r[0]] || !kept[r[2]]) continue;
		uint32_t* w = ov_all + 5 * mo++;
		memcpy(w, r, 20); w[0] = rank[w[0]]; w[2] = rank[w[2]];
	}
	{
		const uint64_t mo_old = mo;
		uint64_t* old_ix = overlap_index(ov_all, mo_old);      /* sorted (key, value); find to replace */
		for (uint64_t e = 0; e < n_add; e++)
		{
			const uint32_t* r = add + 5 * e;
			uint32_t from = rank[r[0]], ffw = r[1], to = rank[r[2]], tfw = r[3];
			const uint64_t key = canon_key(from, ffw, to, tfw);
			int replaced = 0;
			uint64_t have = 0;
			if (edge_find(old_ix, mo_old, key, &have) && have == r[4]) continue;      /* neighbours that were neighbours before */
			if (edge_find(old_ix, mo_old, key, 0))
				for (uint64_t q = 0; q < mo_old && !replaced; q++)
					if (canon_key(ov_all[5 * q], ov_all[5 * q + 1], ov_all[5 * q + 2], ov_all[5 * q + 3]) == key) { ov_all[5 * q + 4] = r[4]; replaced = 1; }
			if (replaced) continue;
			uint32_t* w = ov_all + 5 * mo++;
			w[0] = (uint32_t)(key >> 34); w[1] = (uint32_t)((key >> 33) & 1); w[2] = (uint32_t)((key >> 1) & 0xffffffffu); w[3] = (uint32_t)(key & 1); w[4] = r[4];
		}
		free(old_ix);
	}
	free(u->overlaps);
	u->overlaps = ov_all; u->n_overlaps = mo;
	for (uint64_t i = 0; i < n_new; i++) new_nodes[i] = 2 * rank[new_nodes[i] >> 1] + (new_nodes[i] & 1);
	free(u->unitig_nodes); free(u->unitig_begin);
	u->unitig_nodes = new_nodes; u->unitig_begin = new_begin; u->n_unitig_nodes = n_new; u->n_kept = nk;
	free(kept); free(rank); free(add);
}

/* ---- read threading: getKmerLocator (MBG/KmerMatcher.cpp:3-31) and iterateReadPaths
 * (MBG/KmerMatcher.h:12-144) over every hashed part, in input order (getReadPaths, MBG/MBG.cpp:1235-1251).
 * Sequential, statement by statement; note that readPoses is NOT cleared when a path ends at a k-mer that sits in no
 * unitig (KmerMatcher.h:31-47), so the next path of that part starts with the old entries. */
struct orc_paths {
	uint64_t n_paths, n_path, n_poses, cap_paths, cap_path, cap_poses;
	uint64_t* head;      /* x6: read index, part offset, path entries, readPoses entries, leftClip, rightClip */
	uint32_t* path;      /* 2*unitig + fw */
	uint32_t* poses;
};

void orc_paths_free(orc_paths* p) { if (!p) return; free(p->head); free(p->path); free(p->poses); free(p); }

typedef struct { uint32_t* path; uint64_t n_path, cap_path; uint32_t* poses; uint64_t n_poses, cap_poses; uint64_t left, right; } cur_path;

static void emit_path(orc_paths* out, const orc_part* part, const cur_path* c)
{
	if (out->n_paths + 1 > out->cap_paths) { out->cap_paths = 2 * out->cap_paths + 64; out->head = (uint64_t*)realloc(out->head, out->cap_paths * 48); }
	if (out->n_path + c->n_path > out->cap_path) { out->cap_path = 2 * (out->n_path + c->n_path) + 64; out->path = (uint32_t*)realloc(out->path, out->cap_path * 4); }
	if (out->n_poses + c->n_poses > out->cap_poses) { out->cap_poses = 2 * (out->n_poses + c->n_poses) + 64; out->poses = (uint32_t*)realloc(out->poses, out->cap_poses * 4); }
	uint64_t* h = out->head + 6 * out->n_paths++;
	h[0] = part->read_idx; h[1] = part->part_offset; h[2] = c->n_path; h[3] = c->n_poses; h[4] = c->left; h[5] = c->right;
	memcpy(out->path + out->n_path, c->path, c->n_path * 4); out->n_path += c->n_path;
	memcpy(out->poses + out->n_poses, c->poses, c->n_poses * 4); out->n_poses += c->n_poses;
}
static void push_path(cur_path* c, uint32_t v) { if (c->n_path == c->cap_path) { c->cap_path = 2 * c->cap_path + 16; c->path = (uint32_t*)realloc(c->path, c->cap_path * 4); } c->path[c->n_path++] = v; }
static void push_pos(cur_path* c, uint32_t v) { if (c->n_poses == c->cap_poses) { c->cap_poses = 2 * c->cap_poses + 16; c->poses = (uint32_t*)realloc(c->poses, c->cap_poses * 4); } c->poses[c->n_poses++] = v; }

/* canon() (MBG/MBGCommon.cpp:78-89) of an oriented pair, packed */
static uint64_t canon_key(uint32_t from, uint32_t ffw, uint32_t to, uint32_t tfw)
{
	if (to < from || (to == from && !tfw && !ffw)) { uint32_t a = from, b = ffw; from = to; ffw = !tfw; to = a; tfw = !b; }
	return ((uint64_t)from << 34) | ((uint64_t)ffw << 33) | ((uint64_t)to << 1) | tfw;
}
static int cmp_pair(const void* a, const void* b) { return cmp_u64(a, b); }
/* sorted (key, value) pairs from u32 x6 edge records */
static uint64_t* edge_index(const uint32_t* edges, uint64_t m, int value_col)
{
	uint64_t* ix = (uint64_t*)malloc((m + 1) * 16);
	for (uint64_t e = 0; e < m; e++) { const uint32_t* r = edges + 6 * e; ix[2 * e] = canon_key(r[0], r[1], r[2], r[3]); ix[2 * e + 1] = r[value_col]; }
	qsort(ix, m, 16, cmp_pair);
	return ix;
}
/* the same from u32 x5 overlap records */
static uint64_t* overlap_index(const uint32_t* ov, uint64_t m)
{
	uint64_t* ix = (uint64_t*)malloc((m + 1) * 16);
	for (uint64_t e = 0; e < m; e++) { const uint32_t* r = ov + 5 * e; ix[2 * e] = canon_key(r[0], r[1], r[2], r[3]); ix[2 * e + 1] = r[4]; }
	qsort(ix, m, 16, cmp_pair);
	return ix;
}
static int edge_find(const uint64_t* ix, uint64_t m, uint64_t key, uint64_t* value)
{
	uint64_t lo = 0, hi = m;
	while (lo < hi) { uint64_t mid = (lo + hi) / 2; if (ix[2 * mid] < key) lo = mid + 1; else hi = mid; }
	if (lo < m && ix[2 * lo] == key) { if (value) *value = ix[2 * lo + 1]; return 1; }
	return 0;
}

orc_paths* orc_read_paths(const orc* o, const orc_unitigs* u, uint32_t k)
{
	const uint64_t NONE = ~0ull;
	orc_paths* out = (orc_paths*)calloc(1, sizeof(orc_paths));
	const uint64_t nk = u->n_kept;
	/* getKmerLocator: node -> (unitig, position, fw) */
	uint64_t* loc_unitig = (uint64_t*)malloc((nk + 1) * 8);
	uint64_t* loc_pos = (uint64_t*)malloc((nk + 1) * 8);
	uint8_t* loc_fw = (uint8_t*)malloc(nk + 1);
	for (uint64_t i = 0; i < nk; i++) { loc_unitig[i] = NONE; loc_pos[i] = NONE; loc_fw[i] = 1; }
	for (uint64_t t = 0; t < u->n_unitigs; t++)
	{
		const uint64_t b = u->unitig_begin[t], size = u->unitig_begin[t + 1] - b;
		for (uint64_t j = 0; j < size; j++)
		{
			const uint32_t x = u->unitig_nodes[b + j];
			loc_unitig[x >> 1] = t;
			if (x & 1) { loc_pos[x >> 1] = j; loc_fw[x >> 1] = 1; } else { loc_pos[x >> 1] = size - 1 - j; loc_fw[x >> 1] = 0; }
		}
	}
	uint64_t* overlap_ix = overlap_index(u->overlaps, u->n_overlaps);
	uint64_t* uedge_ix = edge_index(u->uedges, u->n_uedges, 4);
	const orc_part* parts = (const orc_part*)orc_data(o, 0);
	const uint64_t n_parts = orc_count(o, 0);
	const uint32_t* positions = (const uint32_t*)orc_data(o, 3);
	const uint64_t* hashes = (const uint64_t*)orc_data(o, 4);
	cur_path cur;
	memset(&cur, 0, sizeof(cur));
	for (uint64_t pi = 0; pi < n_parts; pi++)
	{
		const orc_part* part = parts + pi;
		if (!part->hashed) continue;
		cur.n_path = 0; cur.n_poses = 0; cur.left = 0; cur.right = 0;      /* a fresh ReadPath per call */
		uint64_t last_read_pos = NONE, last_kmer = NONE, last_kmer_fw = 1;
		uint64_t last_unitig = NONE, last_upos = NONE, last_ufw = 1;
		for (uint64_t i = 0; i < part->n_sel; i++)
		{
			const uint64_t read_pos = positions[part->sel_begin + i];
			const uint64_t f_lo = hashes[2 * (part->sel_begin + i)], f_hi = hashes[2 * (part->sel_begin + i) + 1];
			/* getNodeOrNull (MBG/HashList.cpp:85-101): bwHash = halves swapped */
			const uint64_t b_lo = f_hi, b_hi = f_lo;
			const int fw = (f_hi < b_hi) || (f_hi == b_hi && f_lo < b_lo);
			const uint64_t c_lo = fw ? f_lo : b_lo, c_hi = fw ? f_hi : b_hi;
			uint64_t lo = 0, hi = nk;
			while (lo < hi) { uint64_t mid = (lo + hi) / 2; if (u->f_hash[2 * mid + 1] < c_hi || (u->f_hash[2 * mid + 1] == c_hi && u->f_hash[2 * mid] < c_lo)) lo = mid + 1; else hi = mid; }
			if (!(lo < nk && u->f_hash[2 * lo] == c_lo && u->f_hash[2 * lo + 1] == c_hi)) continue;
			const uint64_t kmer = lo;
			if (loc_unitig[kmer] == NONE)
			{
				if (cur.n_path > 0) emit_path(out, part, &cur);
				cur.n_path = 0; cur.left = 0; cur.right = 0;
				last_unitig = NONE; last_upos = NONE; last_ufw = 1;
				last_read_pos = read_pos; last_kmer = kmer; last_kmer_fw = fw;
				continue;
			}
			uint64_t p_unitig = loc_unitig[kmer], p_pos = loc_pos[kmer], p_fw = loc_fw[kmer];
			const uint64_t usize = u->unitig_begin[p_unitig + 1] - u->unitig_begin[p_unitig];
			if (!fw) { p_fw = !p_fw; p_pos = usize - 1 - p_pos; }
			if (last_unitig == NONE)
			{
				push_path(&cur, (uint32_t)(2 * p_unitig + p_fw));
				push_pos(&cur, (uint32_t)read_pos);
				cur.right = usize - 1 - p_pos; cur.left = p_pos;
				last_unitig = p_unitig; last_upos = p_pos; last_ufw = p_fw; last_read_pos = read_pos; last_kmer = kmer; last_kmer_fw = fw;
				continue;
			}
			int restart = 0;
			if (last_kmer != NONE)
			{
				uint64_t ov = 0;
				const int has = edge_find(overlap_ix, u->n_overlaps, canon_key((uint32_t)last_kmer, (uint32_t)last_kmer_fw, (uint32_t)kmer, (uint32_t)fw), &ov);
				if (!has || read_pos - last_read_pos != k - ov) restart = 1;
			}
			if (!restart)
			{
				if (p_unitig == last_unitig && p_fw == last_ufw && p_pos == last_upos + 1)
				{
					last_unitig = p_unitig; last_upos = p_pos; last_ufw = p_fw; last_read_pos = read_pos; last_kmer = kmer; last_kmer_fw = fw;
					cur.right -= 1;
					push_pos(&cur, (uint32_t)read_pos);
					continue;
				}
				const uint64_t last_size = u->unitig_begin[last_unitig + 1] - u->unitig_begin[last_unitig];
				if (edge_find(uedge_ix, u->n_uedges, canon_key((uint32_t)last_unitig, (uint32_t)last_ufw, (uint32_t)p_unitig, (uint32_t)p_fw), 0) && p_pos == 0 && last_upos == last_size - 1)
				{
					push_path(&cur, (uint32_t)(2 * p_unitig + p_fw));
					push_pos(&cur, (uint32_t)read_pos);
					cur.right = usize - 1;
					last_unitig = p_unitig; last_upos = p_pos; last_ufw = p_fw; last_read_pos = read_pos; last_kmer = kmer; last_kmer_fw = fw;
					continue;
				}
			}
			emit_path(out, part, &cur);
			cur.n_path = 0; cur.n_poses = 0;
			push_path(&cur, (uint32_t)(2 * p_unitig + p_fw));
			push_pos(&cur, (uint32_t)read_pos);
			cur.left = p_pos; cur.right = usize - 1 - p_pos;
			last_unitig = p_unitig; last_upos = p_pos; last_ufw = p_fw; last_read_pos = read_pos; last_kmer = kmer; last_kmer_fw = fw;
		}
		if (cur.n_path > 0) emit_path(out, part, &cur);
	}
	free(cur.path); free(cur.poses);
	free(loc_unitig); free(loc_pos); free(loc_fw); free(overlap_ix); free(uedge_ix);
	return out;
}

/* what: 0 paths, 1 path entries, 2 readPoses entries */
uint64_t orc_paths_count(const orc_paths* p, int what) { switch (what) { case 0: return p->n_paths; case 1: return p->n_path; case 2: return p->n_poses; } return 0; }
/* what: 0 heads (u64 x6) 1 path (u32 = 2*unitig + fw) 2 readPoses (u32) */
const void* orc_paths_data(const orc_paths* p, int what) { switch (what) { case 0: return p->head; case 1: return p->path; case 2: return p->poses; } return 0; }
