/* oracle/mbg_oracle.h -- TEST INFRASTRUCTURE ONLY (see mbg_oracle.c header). */
#ifndef MBG_ORACLE_H
#define MBG_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* error-masking modes, numbered like the reference enum (MBG/ReadHelper.h:20-29) */
enum { ORC_NO = 0, ORC_HPC = 1, ORC_COLLAPSE = 2, ORC_DINUC = 3, ORC_MSAT = 4, ORC_COLLAPSE_DINUC = 5, ORC_COLLAPSE_MSAT = 6 };

/* one read part = one callback of the reference's iterateParts / iterateHashes (48 bytes, no padding) */
typedef struct {
	uint32_t read_idx;    /* index of the read in the order it was added */
	uint32_t part_offset; /* ReadInfo.readName.second */
	uint32_t raw_len;     /* ReadInfo.readLength */
	uint32_t n_codes;     /* ReadInfo.readLengthHpc == number of codes of the part */
	uint64_t code_begin;  /* first code in codes[] */
	uint64_t pose_begin;  /* first of the n_codes+1 entries in poses[] */
	uint64_t sel_begin;   /* first selected k-mer in positions[] / hashes[] */
	uint32_t n_sel;       /* selected k-mers of the part */
	uint32_t hashed;      /* 1 when n_codes >= k (the reference calls the hash callback for it) */
} orc_part;

typedef struct orc orc;

orc* orc_create(int mode, uint32_t k, uint32_t w);
void orc_free(orc* o);
/* returns 0, or -1 palindromic k-mer that cannot be dropped / -2 unhashable k-mer (the reference aborts) */
int orc_add_read(orc* o, const char* seq, uint64_t len);
int orc_add_reads(orc* o, const char* bases, const uint64_t* offsets, uint64_t n_reads);
void orc_build_table(orc* o);
/* hpc variants of pass 3 (MBG/ReadHelper.h:680-737): entry e == one ReadpartIterator::addHpcVariants(hash e, offsets[e],
 * lengths[len_begin[e] .. len_begin[e+1])) call; call before orc_add_read(s).  hashes: lo, hi per entry */
void orc_set_variants(orc* o, const uint64_t* hashes_lo_hi, const uint32_t* offsets, const uint64_t* len_begin, const uint32_t* lengths, uint64_t n);
/* what: 0 parts 1 codes 2 poses 3 selected 4 nodes 5 edges 6 total selected k-mer instances in the table */
uint64_t orc_count(const orc* o, int what);
/* what: 0 parts 1 codes(u16) 2 poses(u32) 3 positions(u32) 4 hashes(u64 lo,hi) 5 node hashes(u64 lo,hi)
 *       6 node coverage(u32) 7 node tip(u8) 8 edges(u32 x6: from,from_fw,to,to_fw,coverage,overlap) */
const void* orc_data(const orc* o, int what);

/* The step after the table (mbg_unitig_oracle.c): getUnitigGraph's coverage filter (edges >= min_cov, chains with mean
 * coverage >= min_unitig_cov; MBG/MBG.cpp:629-696, oneCovHeuristic = keepGaps = false), HashList::filter
 * (MBG/HashList.cpp:159-232) and the unitig chains of the filtered list (MBG/MBG.cpp:689-745).  Needs orc_build_table. */
typedef struct orc_unitigs orc_unitigs;
orc_unitigs* orc_build_unitigs(const orc* o, uint32_t min_cov, double min_unitig_cov);
/* one_cov_heuristic: run removeOnecovNodes (tips, then all) first, as getUnitigGraph does under Verkko's flags (MBG/MBG.cpp:631-635, 1573) */
orc_unitigs* orc_build_unitigs_ex(const orc* o, uint32_t min_cov, double min_unitig_cov, int one_cov_heuristic);
void orc_unitigs_free(orc_unitigs* u);
/* filterKmersToUnitigKmers (MBG/MBG.cpp:1065-1185) with filterWithinUnitig = true (MBG's default), in place: thins the k-mers
 * inside every unitig, filters the list again, adds the sequence overlaps of the new neighbours (array 8 / count 6) */
void orc_thin_unitigs(orc_unitigs* u, uint32_t k);
/* what: 0 nodes of the input table, 1 kept nodes, 2 filtered edges, 3 unitigs, 4 unitig nodes, 5 unitig edges */
uint64_t orc_unitigs_count(const orc_unitigs* u, int what);
/* what: 0 kept(u8 per input node) 1 filtered hashes(u64 lo,hi) 2 filtered coverage(u32) 3 filtered tip(u8)
 *       4 filtered edges(u32 x6) 5 unitig_begin(u64, n+1) 6 unitig nodes(u32 = 2*filtered id + fw)
 *       7 unitig edges(u32 x6: from unitig, from_fw, to unitig, to_fw, coverage, 0; UnitigGraph::edges / edgeCov) */
const void* orc_unitigs_data(const orc_unitigs* u, int what);

/* Read threading (row f2): getKmerLocator + iterateReadPaths (MBG/KmerMatcher.cpp:3-31, KmerMatcher.h:12-144) of every
 * hashed part through the unitigs of `u`, in input order (getReadPaths, MBG/MBG.cpp:1235-1251) */
typedef struct orc_paths orc_paths;
orc_paths* orc_read_paths(const orc* o, const orc_unitigs* u, uint32_t k);
void orc_paths_free(orc_paths* p);
uint64_t orc_paths_count(const orc_paths* p, int what);   /* 0 paths, 1 path entries, 2 readPoses entries */
const void* orc_paths_data(const orc_paths* p, int what); /* 0 heads (u64 x6: read, part offset, n path, n poses, leftClip, rightClip) 1 path (u32 = 2*unitig + fw) 2 readPoses (u32) */

/* primitives, exported for the known-answer tests */
uint64_t orc_char_hash(uint64_t code);
uint16_t orc_complement(uint16_t code);
uint32_t orc_code_motif_length(uint16_t code);
uint32_t orc_max_code(void);
uint64_t orc_murmur64a(const void* bytes, uint64_t len);
void orc_kmer_hash(const uint16_t* codes, uint32_t k, uint64_t* lo, uint64_t* hi);
void orc_rolling_hash(uint32_t s, const uint16_t* codes, uint64_t n, uint64_t* fw, uint64_t* bw);
uint64_t orc_msat(const uint16_t* sym, const uint64_t* poses, uint64_t n, uint32_t max_motif, uint16_t* out_codes, uint64_t* out_poses);
/* the scan-parallel formulation the CUDA kernels implement (DESIGN.md "K2"), restated sequentially so that its
 * equivalence with orc_msat can be fuzzed on the CPU */
uint64_t orc_msat_scanform(const uint16_t* sym, const uint64_t* poses, uint64_t n, uint32_t max_motif, uint16_t* out_codes, uint64_t* out_poses);
uint64_t orc_syncmers_bruteforce(const uint16_t* codes, uint64_t n, uint32_t k, uint32_t w, uint32_t* out_positions);

#ifdef __cplusplus
}
#endif
#endif
