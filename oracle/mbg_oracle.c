/* oracle/mbg_oracle.c
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product (verkko_b200/csrc, libmbgfe.so) never does.
 *
 * A plain, sequential C restatement of the front end of MBG's graph build as vendored in the reference at
 * /root/reference/src/hifioverlapper/MBG/src (commit 7bef84dd; called MBG/ below):
 *   read -> [collapse] -> run-length split at non-ACGT -> [microsatellite coding] -> rolling s-mer hash ->
 *   closed-syncmer selection -> 128-bit k-mer hash -> node / edge table with coverages, tips and overlaps.
 *
 * Third-party arithmetic that is NOT under /root/reference: the k-mer hash is libstdc++'s
 * std::hash<std::string_view> == _Hash_bytes (64-bit MurmurHash64A variant, seed 0xc70f6907; GCC 13.3 here,
 * unchanged since GCC 4.x on LP64 little-endian), called from MBG/MBGCommon.cpp:20-21,32-33.  Its published
 * algorithm is restated in orc_murmur64a().
 *
 * Parity pinning: the reference holds no tests or golden vectors for this path.  This oracle is
 * pinned instead against (1) outputs of the reference itself, compiled here by oracle/Makefile into
 * oracle/_ref/libmbgref.so and compared in tests/test_oracle.py, and (2) fixtures generated from that
 * library and committed under tests/golden/ (generator: tools/make_golden.py).
 */
#include "mbg_oracle.h"
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------------ vectors */

typedef struct { void* p; uint64_t n, cap; uint32_t esz; } vec;
static void vec_init(vec* v, uint32_t esz) { v->p = 0; v->n = 0; v->cap = 0; v->esz = esz; }
static void* vec_push(vec* v)
{
	if (v->n == v->cap)
	{
		v->cap = v->cap ? v->cap * 2 : 1024;
		v->p = realloc(v->p, v->cap * v->esz);
	}
	return (char*)v->p + (v->n++) * v->esz;
}
static void push_u16(vec* v, uint16_t x) { *(uint16_t*)vec_push(v) = x; }
static void push_u32(vec* v, uint32_t x) { *(uint32_t*)vec_push(v) = x; }
static void push_u64(vec* v, uint64_t x) { *(uint64_t*)vec_push(v) = x; }

/* ------------------------------------------------------------------------------- microsatellite alphabet */

/* codes below CODES_BEFORE[m+1] have motif length <= m; sum_{i<m} i*4^i (MBG/ErrorMaskHelper.cpp:9-25) */
static const uint32_t CODES_BEFORE[8] = { 0, 0, 4, 36, 228, 1252, 6372, 30948 };

uint32_t orc_max_code(void) { return CODES_BEFORE[7]; }

/* MBG/ErrorMaskHelper.cpp:32-41 */
uint32_t orc_code_motif_length(uint16_t code)
{
	uint32_t m = 1;
	while (m < 6 && CODES_BEFORE[m+1] <= code) m++;
	return m;
}

/* Reverse complement of one code (MBG/ErrorMaskHelper.cpp:43-73): a code is (motif length m, motif as m base-4
 * digits, first symbol most significant, overhang = run length mod m).  Read backwards the run starts with the
 * overhang, so the motif is rotated left by `overhang` digits, then every digit is complemented and the digit order
 * reversed; m and overhang are kept. */
uint16_t orc_complement(uint16_t code)
{
	uint32_t m = orc_code_motif_length(code);
	uint32_t v = code - CODES_BEFORE[m];
	uint32_t motif = v / m, overhang = v % m;
	uint32_t digits[6];
	for (uint32_t i = 0; i < m; i++) digits[i] = (motif >> (2 * (m - 1 - i))) & 3;
	uint32_t out = 0;
	for (uint32_t i = 0; i < m; i++)
	{
		uint32_t src = digits[(overhang + (m - 1 - i)) % m]; /* rotated, then read back to front */
		out = (out << 2) | (3 - src);
	}
	return (uint16_t)(CODES_BEFORE[m] + out * m + overhang);
}

/* ------------------------------------------------------------------------------------------------- hashes */

static uint64_t rotl64(uint64_t v, uint32_t r) { r &= 63; return r ? (v << r) | (v >> (64 - r)) : v; }
static uint64_t rotr64(uint64_t v, uint32_t r) { r &= 63; return r ? (v >> r) | (v << (64 - r)) : v; }

/* per-code 64-bit mix (MBG/FastHasher.cpp:33-43; Thomas Wang's 64-bit integer hash) */
uint64_t orc_char_hash(uint64_t key)
{
	key = (~key) + (key << 21);
	key ^= key >> 24;
	key *= 265;
	key ^= key >> 14;
	key *= 21;
	key ^= key >> 28;
	key += key << 31;
	return key;
}

/* libstdc++ _Hash_bytes, 64-bit (libstdc++-v3/libsupc++/hash_bytes.cc; not in /root/reference) */
uint64_t orc_murmur64a(const void* bytes, uint64_t len)
{
	const uint64_t mul = 0xc6a4a7935bd1e995ull;
	const unsigned char* p = (const unsigned char*)bytes;
	uint64_t h = 0xc70f6907ull ^ (len * mul);
	uint64_t blocks = len / 8;
	for (uint64_t b = 0; b < blocks; b++)
	{
		uint64_t wd;
		memcpy(&wd, p + 8 * b, 8);
		wd *= mul;
		wd ^= wd >> 47;
		wd *= mul;
		h ^= wd;
		h *= mul;
	}
	if (len & 7)
	{
		uint64_t tail = 0;
		for (uint32_t i = (uint32_t)(len & 7); i-- > 0; ) tail = (tail << 8) | p[8 * blocks + i];
		h ^= tail;
		h *= mul;
	}
	h ^= h >> 47;
	h *= mul;
	h ^= h >> 47;
	return h;
}

/* 128-bit k-mer hash (MBG/MBGCommon.cpp:25-35, MBG/VectorView.h:33-36): low half = Murmur of the raw bytes of the
 * first (k+1)/2 codes, high half = Murmur of the first (k+1)/2 codes of the reverse complement. */
void orc_kmer_hash(const uint16_t* codes, uint32_t k, uint64_t* lo, uint64_t* hi)
{
	uint32_t half = (k + 1) / 2;
	uint16_t* rc = (uint16_t*)malloc(half * sizeof(uint16_t));
	for (uint32_t i = 0; i < half; i++) rc[i] = orc_complement(codes[k - 1 - i]);
	*lo = orc_murmur64a(codes, (uint64_t)half * 2);
	*hi = orc_murmur64a(rc, (uint64_t)half * 2);
	free(rc);
}

/* Rolling canonical s-mer hash (MBG/FastHasher.h:16-32, .cpp:45-74).  Rotation amounts are taken mod 64, which is
 * what the reference's shifts evaluate to on x86-64 when s mod 64 is 0. */
typedef struct { uint64_t fw, bw; uint32_t s; } roll;
static void roll_add(roll* r, uint16_t c)
{
	r->fw = rotl64(r->fw, 1) ^ orc_char_hash(c);
	r->bw = rotr64(r->bw, 1) ^ rotl64(orc_char_hash(orc_complement(c)), (r->s + 63) & 63);
}
static void roll_remove(roll* r, uint16_t c)
{
	r->fw ^= rotl64(orc_char_hash(c), r->s & 63);
	r->bw ^= rotr64(orc_char_hash(orc_complement(c)), 1);
}
static uint64_t roll_value(const roll* r) { return r->fw < r->bw ? r->fw : r->bw; }

void orc_rolling_hash(uint32_t s, const uint16_t* codes, uint64_t n, uint64_t* fw, uint64_t* bw)
{
	roll r = { 0, 0, s };
	for (uint64_t i = 0; i < n; i++)
	{
		roll_add(&r, codes[i]);
		if (i >= s) roll_remove(&r, codes[i - s]);
	}
	*fw = r.fw;
	*bw = r.bw;
}

/* ------------------------------------------------------------------------------ microsatellite run coding */

typedef struct { uint64_t start, end; uint32_t m; } msat_run;

typedef struct {
	const uint16_t* sym;
	const uint64_t* poses;
	uint16_t* out_codes;
	uint64_t* out_poses;
	uint64_t n_out;
	/* state of the overlap resolver (MBG/ErrorMaskHelper.cpp:250-305) */
	int have_pending;
	msat_run pending;
	uint64_t singles_until;
} msat_state;

/* one output code for the stretch [a,b) cut from a run of motif length m (MBG/ErrorMaskHelper.cpp:107-127,307-329) */
static void msat_emit(msat_state* st, uint64_t a, uint64_t b, uint32_t m)
{
	if (m > b - a) m = (uint32_t)(b - a);
	uint32_t motif = 0;
	for (uint32_t i = 0; i < m; i++) motif = (motif << 2) | st->sym[a + i];
	st->out_codes[st->n_out] = (uint16_t)(CODES_BEFORE[m] + motif * m + (uint32_t)((b - a) % m));
	st->out_poses[st->n_out] = st->poses[a];
	st->n_out++;
}

/* Runs arrive with strictly increasing start and end.  Where two runs overlap, the shared symbols are written as
 * single-symbol codes; what is left of each run on either side keeps the run's motif length. */
static void msat_take_run(msat_state* st, msat_run r)
{
	if (!st->have_pending)
	{
		st->pending = r;
		st->have_pending = 1;
		return;
	}
	msat_run p = st->pending;
	if (st->singles_until > r.start)
	{
		for (uint64_t j = st->singles_until; j < p.end; j++) msat_emit(st, j, j + 1, 1);
		st->singles_until = p.end;
	}
	else if (r.start >= p.end)
	{
		uint64_t lo = st->singles_until > p.start ? st->singles_until : p.start;
		msat_emit(st, lo, p.end, p.m);
	}
	else
	{
		uint64_t lo = st->singles_until > p.start ? st->singles_until : p.start;
		if (lo < r.start) msat_emit(st, lo, r.start, p.m);
		st->singles_until = p.end;
		for (uint64_t j = r.start; j < p.end; j++) msat_emit(st, j, j + 1, 1);
	}
	st->pending = r;
}

/* multiRLECompress (MBG/ErrorMaskHelper.cpp:331-342) = greedy left-to-right search for the longest tandem run
 * (motif 2..max_motif, at least two copies) starting at each position (iterateRuns, :169-248), the overlap
 * resolver above, and the re-coding.  sym holds values 0..3, poses has n+1 entries.  Returns the code count;
 * out_poses receives count+1 entries. */
uint64_t orc_msat(const uint16_t* sym, const uint64_t* poses, uint64_t n, uint32_t max_motif, uint16_t* out_codes, uint64_t* out_poses)
{
	msat_state st;
	memset(&st, 0, sizeof(st));
	st.sym = sym; st.poses = poses; st.out_codes = out_codes; st.out_poses = out_poses;
	uint64_t last_end = 0;
	for (uint64_t i = 0; i < n; i++)
	{
		if (i + 7 < last_end) continue; /* :189 positions deep inside the previous run are not examined */
		msat_run best = { i, i + 1, 1 };
		for (uint32_t m = 2; m <= max_motif; m++)
		{
			if (i + 2 * (uint64_t)m > n) break;
			if (memcmp(sym + i, sym + i + m, m * sizeof(uint16_t)) != 0) continue;
			uint64_t e = i + 2 * (uint64_t)m; /* extend while the period holds: whole copies, then the overhang */
			while (e < n && sym[e] == sym[e - m]) e++;
			if (e <= last_end) continue;
			if (e > best.end) { best.end = e; best.m = m; }
		}
		if (best.end > last_end)
		{
			msat_take_run(&st, best);
			last_end = best.end;
		}
	}
	/* flush (:294-304) */
	if (st.singles_until > st.pending.start)
	{
		if (st.singles_until != st.pending.end) msat_emit(&st, st.singles_until, st.pending.end, st.pending.m);
	}
	else
	{
		msat_emit(&st, st.pending.start, st.pending.end, st.pending.m);
	}
	out_poses[st.n_out] = poses[st.pending.end];
	return st.n_out;
}

/* The same function in the scan-parallel form the CUDA kernel uses (kernel K2):
 *   cand(i)   = max over m of the end of a period-m run that STARTS at i (eq_m run of length >= m beginning at i)
 *   emitted(i)= cand(i) > max_{j<i} cand(j)                      (exclusive prefix maximum)
 *   cover(p)  = number of emitted runs containing p              (0, 1 or >= 2)
 *   a code starts at p iff p == 0, cover(p) != 1, cover(p-1) != 1, or an emitted run starts at p
 *   stretches with cover == 1 take the motif length of the latest emitted run starting at or before them.
 * Valid when no two neighbouring symbols are equal (always true after run-length / collapse), where the reference's
 * "i + 7 < lastRunEnd" skip can never hide a longer run. */
uint64_t orc_msat_scanform(const uint16_t* sym, const uint64_t* poses, uint64_t n, uint32_t max_motif, uint16_t* out_codes, uint64_t* out_poses)
{
	uint64_t* cand = (uint64_t*)calloc(n + 1, sizeof(uint64_t));
	uint8_t* cand_m = (uint8_t*)calloc(n + 1, 1);
	int32_t* delta = (int32_t*)calloc(n + 2, sizeof(int32_t));
	uint8_t* run_start = (uint8_t*)calloc(n + 1, 1);
	for (uint32_t m = 2; m <= max_motif; m++)
	{
		/* eq(j) = sym[j] == sym[j+m]; a maximal stretch of eq == 1 of length L >= m starting at a is a run [a, a+L+m) */
		uint64_t j = 0;
		while (j + m < n)
		{
			if (sym[j] != sym[j + m]) { j++; continue; }
			uint64_t a = j;
			while (j + m < n && sym[j] == sym[j + m]) j++;
			if (j - a >= m && j + m > cand[a]) { cand[a] = j + m; cand_m[a] = (uint8_t)m; }
		}
	}
	uint64_t prefix_max = 0, last_start = 0, last_prev_end = 0;
	uint32_t last_m = 1;
	(void)last_start; (void)last_prev_end; (void)last_m;
	/* pass 1: emitted runs and coverage deltas */
	uint64_t* emitted_prev_end = (uint64_t*)calloc(n + 1, sizeof(uint64_t));
	for (uint64_t i = 0; i < n; i++)
	{
		if (cand[i] > prefix_max)
		{
			run_start[i] = 1;
			emitted_prev_end[i] = prefix_max;
			delta[i] += 1;
			delta[cand[i]] -= 1;
			prefix_max = cand[i];
		}
	}
	/* pass 2: segmentation */
	uint64_t n_out = 0;
	int32_t cover = 0, prev_cover = 0;
	uint64_t stretch_start = 0; uint32_t stretch_m = 1; int in_stretch = 0;
	uint32_t cur_m = 1;
	for (uint64_t p = 0; p <= n; p++)
	{
		if (p < n) cover += delta[p];
		int starts_here = p < n && run_start[p];
		if (starts_here) cur_m = cand_m[p];
		int boundary = (p == n) || p == 0 || cover != 1 || prev_cover != 1 || starts_here;
		if (boundary && in_stretch)
		{
			uint32_t m = stretch_m;
			uint64_t len = p - stretch_start;
			if (m > len) m = (uint32_t)len;
			uint32_t motif = 0;
			for (uint32_t i = 0; i < m; i++) motif = (motif << 2) | sym[stretch_start + i];
			out_codes[n_out] = (uint16_t)(CODES_BEFORE[m] + motif * m + (uint32_t)(len % m));
			out_poses[n_out] = poses[stretch_start];
			n_out++;
			in_stretch = 0;
		}
		if (p == n) break;
		if (cover == 1) { if (boundary) { in_stretch = 1; stretch_start = p; stretch_m = cur_m; } }
		else
		{
			out_codes[n_out] = sym[p];
			out_poses[n_out] = poses[p];
			n_out++;
		}
		prev_cover = cover;
	}
	out_poses[n_out] = poses[n];
	free(cand); free(cand_m); free(delta); free(run_start); free(emitted_prev_end);
	return n_out;
}

/* ------------------------------------------------------------------------------------- syncmer selection */

/* Closed syncmers with ties (MBG/ReadHelper.h:31-79): k-mer p is selected when the first or the last of its
 * w = k-s+1 s-mer hashes equals the window minimum.  Monotone queue, popping only strictly greater values. */
static uint64_t select_syncmers(const uint16_t* codes, uint64_t n, uint32_t k, uint32_t w, vec* out_positions)
{
	if (n < k) return 0;
	uint32_t s = k - w + 1;
	uint64_t n_smers = n - s + 1;
	uint64_t* qpos = (uint64_t*)malloc((w + 1) * sizeof(uint64_t));
	uint64_t* qval = (uint64_t*)malloc((w + 1) * sizeof(uint64_t));
	uint64_t head = 0, tail = 0, count = 0; /* ring buffer [head, tail) */
	roll r = { 0, 0, s };
	for (uint32_t i = 0; i < s; i++) roll_add(&r, codes[i]);
	for (uint64_t i = 0; i < n_smers; i++)
	{
		if (i > 0)
		{
			roll_add(&r, codes[s + i - 1]);
			roll_remove(&r, codes[i - 1]);
		}
		uint64_t h = roll_value(&r);
		if (i >= w && tail > head && qpos[head % (w + 1)] + w <= i) head++;
		while (tail > head && qval[(tail - 1) % (w + 1)] > h) tail--;
		qpos[tail % (w + 1)] = i; qval[tail % (w + 1)] = h; tail++;
		if (i + 1 >= w)
		{
			uint64_t p = i + 1 - w;
			if (qpos[head % (w + 1)] == p || qval[head % (w + 1)] == h)
			{
				push_u32(out_positions, (uint32_t)p);
				count++;
			}
		}
	}
	free(qpos); free(qval);
	return count;
}

/* brute-force statement of the same rule, for cross-checking */
uint64_t orc_syncmers_bruteforce(const uint16_t* codes, uint64_t n, uint32_t k, uint32_t w, uint32_t* out_positions)
{
	if (n < k) return 0;
	uint32_t s = k - w + 1;
	uint64_t n_smers = n - s + 1, count = 0;
	uint64_t* h = (uint64_t*)malloc(n_smers * sizeof(uint64_t));
	for (uint64_t i = 0; i < n_smers; i++)
	{
		uint64_t fw, bw;
		orc_rolling_hash(s, codes + i, s, &fw, &bw);
		h[i] = fw < bw ? fw : bw;
	}
	for (uint64_t p = 0; p + k <= n; p++)
	{
		uint64_t mn = h[p];
		for (uint32_t j = 1; j < w; j++) if (h[p + j] < mn) mn = h[p + j];
		if (h[p] == mn || h[p + w - 1] == mn) out_positions[count++] = (uint32_t)p;
	}
	free(h);
	return count;
}

/* --------------------------------------------------------------------------------------------- front end */

struct orc {
	int mode;
	uint32_t k, w;
	uint32_t n_reads;
	vec parts, codes, poses, positions, hashes;
	vec node_hash, node_cov, node_tip, edges;
	uint64_t total_selected;
	/* hpc variants (ReadpartIterator::hpcVariants, MBG/ReadHelper.h:250): entries grouped by hash, pairs sorted by
	 * offset, lengths sorted (what addHpcVariants leaves, MBG/ReadHelper.cpp:63-70) */
	uint64_t n_var;
	uint64_t* var_hash;      /* lo, hi per entry, entries sorted by (hi, lo, offset) */
	uint32_t* var_offset;
	uint64_t* var_len_begin; /* n_var + 1 */
	uint32_t* var_len;
};

orc* orc_create(int mode, uint32_t k, uint32_t w)
{
	orc* o = (orc*)calloc(1, sizeof(orc));
	o->mode = mode; o->k = k; o->w = w;
	vec_init(&o->parts, sizeof(orc_part));
	vec_init(&o->codes, 2); vec_init(&o->poses, 4); vec_init(&o->positions, 4); vec_init(&o->hashes, 16);
	vec_init(&o->node_hash, 16); vec_init(&o->node_cov, 4); vec_init(&o->node_tip, 1); vec_init(&o->edges, 24);
	return o;
}

void orc_free(orc* o)
{
	if (!o) return;
	free(o->parts.p); free(o->codes.p); free(o->poses.p); free(o->positions.p); free(o->hashes.p);
	free(o->node_hash.p); free(o->node_cov.p); free(o->node_tip.p); free(o->edges.p);
	free(o->var_hash); free(o->var_offset); free(o->var_len_begin); free(o->var_len);
	free(o);
}

static int base_code(char c)
{
	switch (c)
	{
		case 'A': case 'a': return 0;
		case 'C': case 'c': return 1;
		case 'G': case 'g': return 2;
		case 'T': case 't': return 3;
	}
	return -1;
}

/* ---- hpc variants (pass 3 of runMBG): MBG/ReadHelper.h:680-737, MBG/MBGCommon.h:33-45, MBG/ReadHelper.cpp:63-70 */

typedef struct { uint64_t lo, hi; uint32_t offset; uint64_t idx; } var_key;
static int cmp_var_key(const void* a, const void* b)
{
	const var_key* x = (const var_key*)a; const var_key* y = (const var_key*)b;
	if (x->hi != y->hi) return x->hi < y->hi ? -1 : 1;
	if (x->lo != y->lo) return x->lo < y->lo ? -1 : 1;
	if (x->offset != y->offset) return x->offset < y->offset ? -1 : 1;
	return x->idx < y->idx ? -1 : x->idx > y->idx;
}
static int cmp_u32(const void* a, const void* b) { uint32_t x = *(const uint32_t*)a, y = *(const uint32_t*)b; return x < y ? -1 : x > y; }

/* entry e == one addHpcVariants(hash e, offsets[e], lengths[len_begin[e] .. len_begin[e+1])) call */
void orc_set_variants(orc* o, const uint64_t* hashes_lo_hi, const uint32_t* offsets, const uint64_t* len_begin, const uint32_t* lengths, uint64_t n)
{
	free(o->var_hash); free(o->var_offset); free(o->var_len_begin); free(o->var_len);
	o->n_var = n;
	o->var_hash = (uint64_t*)malloc((n + 1) * 16); o->var_offset = (uint32_t*)malloc((n + 1) * 4);
	o->var_len_begin = (uint64_t*)malloc((n + 1) * 8); o->var_len = (uint32_t*)malloc((len_begin[n] + 1) * 4);
	var_key* keys = (var_key*)malloc((n + 1) * sizeof(var_key));
	for (uint64_t e = 0; e < n; e++) { keys[e].lo = hashes_lo_hi[2 * e]; keys[e].hi = hashes_lo_hi[2 * e + 1]; keys[e].offset = offsets[e]; keys[e].idx = e; }
	qsort(keys, n, sizeof(var_key), cmp_var_key);
	uint64_t at = 0;
	for (uint64_t e = 0; e < n; e++)
	{
		const uint64_t src = keys[e].idx, cnt = len_begin[src + 1] - len_begin[src];
		o->var_hash[2 * e] = keys[e].lo; o->var_hash[2 * e + 1] = keys[e].hi; o->var_offset[e] = keys[e].offset;
		o->var_len_begin[e] = at;
		memcpy(o->var_len + at, lengths + len_begin[src], cnt * 4);
		qsort(o->var_len + at, cnt, 4, cmp_u32);
		at += cnt;
	}
	o->var_len_begin[n] = at;
	free(keys);
}

/* first entry with hash >= (lo, hi); entries of one hash are contiguous */
static uint64_t var_lower_bound(const orc* o, uint64_t lo, uint64_t hi)
{
	uint64_t a = 0, b = o->n_var;
	while (a < b)
	{
		uint64_t m = (a + b) / 2;
		uint64_t mlo = o->var_hash[2 * m], mhi = o->var_hash[2 * m + 1];
		if (mhi < hi || (mhi == hi && mlo < lo)) a = m + 1; else b = m;
	}
	return a;
}

/* std::hash<const std::vector<size_t>&> (MBG/MBGCommon.h:33-45); std::hash<size_t> is the identity in libstdc++ */
static uint64_t var_vec_hash_step(uint64_t h, uint64_t x) { h ^= x; return (h << 3) + (h >> 61); }

/* the XOR perturbation of one selected k-mer at part position `pos` (MBG/ReadHelper.h:684-730) */
static void apply_variants(const orc* o, const uint32_t* poses, uint32_t pos, uint64_t* lo, uint64_t* hi)
{
	if (o->n_var == 0) return;
	const uint32_t k = o->k;
	uint64_t first = 0, second = 0;
	uint64_t e = var_lower_bound(o, *lo, *hi);
	if (e < o->n_var && o->var_hash[2 * e] == *lo && o->var_hash[2 * e + 1] == *hi)
	{
		/* forward: first = offsets <= k/2 in ascending order, second = offsets >= k/2 reversed */
		uint64_t end = e;
		while (end < o->n_var && o->var_hash[2 * end] == *lo && o->var_hash[2 * end + 1] == *hi) end++;
		for (uint64_t i = e; i < end; i++)
		{
			uint32_t off = o->var_offset[i];
			if (off > k / 2) continue;
			uint32_t here = poses[pos + off + 1] - poses[pos + off];
			const uint32_t* l = o->var_len + o->var_len_begin[i]; uint64_t nl = o->var_len_begin[i + 1] - o->var_len_begin[i];
			uint64_t j = 0; while (j + 1 < nl && l[j] < here) j++;
			first = var_vec_hash_step(first, l[j]);
		}
		for (uint64_t i = end; i-- > e; )
		{
			uint32_t off = o->var_offset[i];
			if (off < k / 2) continue;
			uint32_t here = poses[pos + off + 1] - poses[pos + off];
			const uint32_t* l = o->var_len + o->var_len_begin[i]; uint64_t nl = o->var_len_begin[i + 1] - o->var_len_begin[i];
			uint64_t j = 0; while (j + 1 < nl && l[j] < here) j++;
			second = var_vec_hash_step(second, l[j]);
		}
	}
	else
	{
		/* the map holds the reverse complement: same walk with mirrored offsets and swapped roles */
		e = var_lower_bound(o, *hi, *lo);
		if (!(e < o->n_var && o->var_hash[2 * e] == *hi && o->var_hash[2 * e + 1] == *lo)) return;
		uint64_t end = e;
		while (end < o->n_var && o->var_hash[2 * end] == *hi && o->var_hash[2 * end + 1] == *lo) end++;
		for (uint64_t i = end; i-- > e; )
		{
			uint32_t off = o->var_offset[i];
			if (off < k / 2) continue;
			uint32_t here = poses[pos + k - off] - poses[pos + k - off - 1];
			const uint32_t* l = o->var_len + o->var_len_begin[i]; uint64_t nl = o->var_len_begin[i + 1] - o->var_len_begin[i];
			uint64_t j = 0; while (j + 1 < nl && l[j] < here) j++;
			first = var_vec_hash_step(first, l[j]);
		}
		for (uint64_t i = e; i < end; i++)
		{
			uint32_t off = o->var_offset[i];
			if (off > k / 2) continue;
			uint32_t here = poses[pos + k - off] - poses[pos + k - off - 1];
			const uint32_t* l = o->var_len + o->var_len_begin[i]; uint64_t nl = o->var_len_begin[i + 1] - o->var_len_begin[i];
			uint64_t j = 0; while (j + 1 < nl && l[j] < here) j++;
			second = var_vec_hash_step(second, l[j]);
		}
	}
	*lo ^= first; *hi ^= second;
}

/* k-mer selection + hashing + palindrome rule for one part (MBG/ReadHelper.h:740-799, 1075-1102) */
static int hash_part(orc* o, orc_part* part, const uint16_t* codes)
{
	uint32_t k = o->k;
	uint64_t n = part->n_codes;
	if (n < k) return 0;
	part->hashed = 1;
	vec cand;
	vec_init(&cand, 4);
	uint64_t n_cand = select_syncmers(codes, n, k, o->w, &cand);
	const uint32_t* cp = (const uint32_t*)cand.p;
	part->sel_begin = o->positions.n;
	int rc = 0;
	for (uint64_t i = 0; i < n_cand && rc == 0; i++)
	{
		const uint16_t* km = codes + cp[i];
		uint64_t lo, hi;
		orc_kmer_hash(km, k, &lo, &hi);
		if (lo == hi)
		{
			int palindrome = 1;
			for (uint32_t j = 0; j < k / 2; j++) if (km[j] != orc_complement(km[k - 1 - j])) { palindrome = 0; break; }
			if (!palindrome) { rc = -2; break; }
			if (i == 0 || i == n_cand - 1 || cp[i + 1] - cp[i - 1] < k) continue; /* dropped without leaving a gap */
			rc = -1;
			break;
		}
		push_u32(&o->positions, cp[i]);
		apply_variants(o, (const uint32_t*)o->poses.p + part->pose_begin, cp[i], &lo, &hi);
		uint64_t* h = (uint64_t*)vec_push(&o->hashes);
		h[0] = lo; h[1] = hi;
		part->n_sel++;
	}
	free(cand.p);
	return rc;
}

/* one maximal ACGT stretch of a (possibly collapsed) read: sym/starts hold its run-length symbols and the position
 * of each run in `coordinates of the string handed to the run-length step`; end_pos is the closing sentinel */
static int finish_part(orc* o, uint32_t read_idx, uint32_t raw_len, uint32_t part_offset, const uint16_t* sym, const uint64_t* starts, uint64_t n_sym)
{
	int msat = o->mode == ORC_DINUC || o->mode == ORC_MSAT || o->mode == ORC_COLLAPSE_DINUC || o->mode == ORC_COLLAPSE_MSAT;
	uint32_t max_motif = (o->mode == ORC_DINUC || o->mode == ORC_COLLAPSE_DINUC) ? 2 : 6;
	const uint16_t* codes = sym;
	const uint64_t* poses = starts;
	uint64_t n_codes = n_sym;
	uint16_t* mc = 0; uint64_t* mp = 0;
	if (msat)
	{
		if (n_sym < 32) return 0; /* MBG/ReadHelper.h:847,858,882,893 */
		mc = (uint16_t*)malloc(n_sym * sizeof(uint16_t));
		mp = (uint64_t*)malloc((n_sym + 1) * sizeof(uint64_t));
		n_codes = orc_msat(sym, starts, n_sym, max_motif, mc, mp);
		codes = mc; poses = mp;
	}
	orc_part part;
	memset(&part, 0, sizeof(part));
	part.read_idx = read_idx; part.part_offset = part_offset; part.raw_len = raw_len;
	part.n_codes = (uint32_t)n_codes;
	part.code_begin = o->codes.n; part.pose_begin = o->poses.n;
	for (uint64_t i = 0; i < n_codes; i++) push_u16(&o->codes, codes[i]);
	for (uint64_t i = 0; i <= n_codes; i++) push_u32(&o->poses, (uint32_t)poses[i]);
	int rc = hash_part(o, &part, codes);
	*(orc_part*)vec_push(&o->parts) = part;
	free(mc); free(mp);
	return rc;
}

int orc_add_read(orc* o, const char* raw, uint64_t raw_len)
{
	uint32_t read_idx = o->n_reads++;
	if (raw_len < 32) return 0; /* MBG/ReadHelper.h:838 */
	int collapse = o->mode == ORC_COLLAPSE || o->mode == ORC_COLLAPSE_DINUC || o->mode == ORC_COLLAPSE_MSAT;
	char* tmp = 0;
	const char* seq = raw;
	uint64_t len = raw_len;
	if (collapse)
	{
		/* MBG/ReadHelper.h:864-876: drop every character equal to its predecessor (any character, also N) */
		tmp = (char*)malloc(raw_len);
		len = 0;
		for (uint64_t i = 0; i < raw_len; i++) if (i == 0 || raw[i] != raw[i - 1]) tmp[len++] = raw[i];
		seq = tmp;
	}
	uint16_t* sym = (uint16_t*)malloc((len + 1) * sizeof(uint16_t));
	uint64_t* starts = (uint64_t*)malloc((len + 2) * sizeof(uint64_t));
	uint64_t n_sym = 0;
	int rc = 0;
	if (o->mode == ORC_NO)
	{
		/* MBG/ReadHelper.h:1019-1074: every base is a code; the reported part offset is the index of the LAST
		 * non-ACGT character in front of the part (0 for the first part) -- kept as the reference has it */
		uint64_t offset = 0;
		for (uint64_t i = 0; i <= len && rc == 0; i++)
		{
			int c = i < len ? base_code(seq[i]) : -1;
			if (c >= 0) { sym[n_sym] = (uint16_t)c; starts[n_sym] = i; n_sym++; continue; }
			if (n_sym > 0)
			{
				starts[n_sym] = i;
				rc = finish_part(o, read_idx, (uint32_t)raw_len, (uint32_t)offset, sym, starts, n_sym);
			}
			n_sym = 0;
			offset = i;
		}
	}
	else
	{
		/* MBG/ReadHelper.h:899-1018: homopolymer runs -> one symbol + start position; parts split at non-ACGT */
		uint64_t offset = 0;
		for (uint64_t i = 0; i <= len && rc == 0; i++)
		{
			int c = i < len ? base_code(seq[i]) : -1;
			if (c >= 0)
			{
				if (n_sym == 0) offset = i;
				if (n_sym == 0 || sym[n_sym - 1] != c) { sym[n_sym] = (uint16_t)c; starts[n_sym] = i; n_sym++; }
				continue;
			}
			if (n_sym > 0)
			{
				starts[n_sym] = i;
				rc = finish_part(o, read_idx, (uint32_t)raw_len, (uint32_t)offset, sym, starts, n_sym);
			}
			n_sym = 0;
		}
	}
	free(sym); free(starts); free(tmp);
	return rc;
}

int orc_add_reads(orc* o, const char* bases, const uint64_t* offsets, uint64_t n_reads)
{
	for (uint64_t r = 0; r < n_reads; r++)
	{
		int rc = orc_add_read(o, bases + offsets[r], offsets[r + 1] - offsets[r]);
		if (rc) return rc;
	}
	return 0;
}

/* -------------------------------------------------------------------------------------------------- table */

typedef struct { uint64_t hi, lo; } key128;
static int cmp_key128(const void* a, const void* b)
{
	const key128* x = (const key128*)a; const key128* y = (const key128*)b;
	if (x->hi != y->hi) return x->hi < y->hi ? -1 : 1;
	if (x->lo != y->lo) return x->lo < y->lo ? -1 : 1;
	return 0;
}
typedef struct { uint32_t from, from_fw, to, to_fw; uint64_t order; uint32_t overlap; } edge_obs;
static int cmp_edge(const void* a, const void* b)
{
	const edge_obs* x = (const edge_obs*)a; const edge_obs* y = (const edge_obs*)b;
	if (x->from != y->from) return x->from < y->from ? -1 : 1;
	if (x->from_fw != y->from_fw) return x->from_fw > y->from_fw ? -1 : 1; /* forward side first */
	if (x->to != y->to) return x->to < y->to ? -1 : 1;
	if (x->to_fw != y->to_fw) return x->to_fw < y->to_fw ? -1 : 1;
	if (x->order != y->order) return x->order < y->order ? -1 : 1;
	return 0;
}

/* What loadReadsAsHashesMultithread (MBG/MBG.cpp:44-76) leaves in the HashList after HashList::sortByHash
 * (MBG/HashList.cpp:245-353), for one worker thread: node ids are ranks of the canonical hash; coverage counts every
 * instance; the first and last selected k-mer of each part are tips; each adjacent pair adds one to the coverage of
 * its canonical edge (MBG/MBGCommon.cpp:78-89) and (last writer wins) sets its overlap. */
void orc_build_table(orc* o)
{
	const orc_part* parts = (const orc_part*)o->parts.p;
	const uint32_t* positions = (const uint32_t*)o->positions.p;
	const uint64_t* hashes = (const uint64_t*)o->hashes.p;
	uint64_t n_inst = o->positions.n;
	o->node_hash.n = o->node_cov.n = o->node_tip.n = o->edges.n = 0;
	o->total_selected = n_inst;
	if (n_inst == 0) return;
	key128* keys = (key128*)malloc(n_inst * sizeof(key128));
	for (uint64_t i = 0; i < n_inst; i++)
	{
		uint64_t lo = hashes[2*i], hi = hashes[2*i+1];
		/* forward hash = hi:lo, backward hash = lo:hi; canonical = the smaller 128-bit value */
		int fw = hi < lo || (hi == lo);
		keys[i].hi = fw ? hi : lo;
		keys[i].lo = fw ? lo : hi;
	}
	key128* sorted = (key128*)malloc(n_inst * sizeof(key128));
	memcpy(sorted, keys, n_inst * sizeof(key128));
	qsort(sorted, n_inst, sizeof(key128), cmp_key128);
	uint64_t n_nodes = 0;
	for (uint64_t i = 0; i < n_inst; i++) if (i == 0 || cmp_key128(&sorted[i], &sorted[n_nodes - 1]) != 0) sorted[n_nodes++] = sorted[i];
	uint32_t* cov = (uint32_t*)calloc(n_nodes, sizeof(uint32_t));
	uint8_t* tip = (uint8_t*)calloc(n_nodes, 1);
	uint32_t* inst_node = (uint32_t*)malloc(n_inst * sizeof(uint32_t));
	for (uint64_t i = 0; i < n_inst; i++)
	{
		key128* f = (key128*)bsearch(&keys[i], sorted, n_nodes, sizeof(key128), cmp_key128);
		inst_node[i] = (uint32_t)(f - sorted);
		cov[inst_node[i]]++;
	}
	uint64_t n_obs = 0;
	edge_obs* obs = (edge_obs*)malloc((n_inst + 1) * sizeof(edge_obs));
	for (uint64_t pi = 0; pi < o->parts.n; pi++)
	{
		const orc_part* p = &parts[pi];
		for (uint32_t j = 0; j < p->n_sel; j++)
		{
			uint64_t i = p->sel_begin + j;
			if (j == 0 || j == p->n_sel - 1) tip[inst_node[i]] = 1;
			if (j == 0) continue;
			uint32_t from = inst_node[i-1], to = inst_node[i];
			uint32_t from_fw = hashes[2*(i-1)+1] < hashes[2*(i-1)] ? 1 : 0; /* fwHash < bwHash  <=>  hi < lo */
			uint32_t to_fw = hashes[2*i+1] < hashes[2*i] ? 1 : 0;
			if (to < from || (to == from && !to_fw && !from_fw))
			{
				uint32_t t = from; from = to; to = t;
				uint32_t f = from_fw; from_fw = !to_fw; to_fw = !f;
			}
			edge_obs* e = &obs[n_obs++];
			e->from = from; e->from_fw = from_fw; e->to = to; e->to_fw = to_fw;
			e->order = i;
			e->overlap = positions[i-1] + o->k - positions[i];
		}
	}
	qsort(obs, n_obs, sizeof(edge_obs), cmp_edge);
	for (uint64_t i = 0; i < n_nodes; i++)
	{
		uint64_t* h = (uint64_t*)vec_push(&o->node_hash);
		h[0] = sorted[i].lo; h[1] = sorted[i].hi;
		push_u32(&o->node_cov, cov[i]);
		*(uint8_t*)vec_push(&o->node_tip) = tip[i];
	}
	for (uint64_t i = 0; i < n_obs; )
	{
		uint64_t j = i;
		while (j + 1 < n_obs && obs[j+1].from == obs[i].from && obs[j+1].from_fw == obs[i].from_fw && obs[j+1].to == obs[i].to && obs[j+1].to_fw == obs[i].to_fw) j++;
		uint32_t* e = (uint32_t*)vec_push(&o->edges);
		e[0] = obs[i].from; e[1] = obs[i].from_fw; e[2] = obs[i].to; e[3] = obs[i].to_fw;
		e[4] = (uint32_t)(j - i + 1);
		e[5] = obs[j].overlap; /* highest order = last writer */
		i = j + 1;
	}
	free(keys); free(sorted); free(cov); free(tip); free(inst_node); free(obs);
}

uint64_t orc_count(const orc* o, int what)
{
	switch (what)
	{
		case 0: return o->parts.n;
		case 1: return o->codes.n;
		case 2: return o->poses.n;
		case 3: return o->positions.n;
		case 4: return o->node_cov.n;
		case 5: return o->edges.n;
		case 6: return o->total_selected;
	}
	return 0;
}

const void* orc_data(const orc* o, int what)
{
	switch (what)
	{
		case 0: return o->parts.p;
		case 1: return o->codes.p;
		case 2: return o->poses.p;
		case 3: return o->positions.p;
		case 4: return o->hashes.p;
		case 5: return o->node_hash.p;
		case 6: return o->node_cov.p;
		case 7: return o->node_tip.p;
		case 8: return o->edges.p;
	}
	return 0;
}
