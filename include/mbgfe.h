/* mbgfe.h -- C ABI of libmbgfe: the B200 (sm_100a) front end of MBG's graph build.
 *
 * This is the drop-in boundary for ONE path of marbl/verkko: what MBG does between "a read arrives" and "the k-mer
 * table is built" in Verkko's buildGraph rule.  Reference = the MBG tree vendored at
 * /root/reference/src/hifioverlapper/MBG/src (commit 7bef84dd), cited below as MBG/.  Each entry point names the
 * reference interface it replaces.  Plain C types only; every array handed out is owned by the library and stays
 * valid until the next mbgfe_process()/mbgfe_run()/mbgfe_reset() on the handle or mbgfe_destroy().
 *
 * Errors: functions return 0 on success, a negative MBGFE_E_* otherwise; mbgfe_last_error() describes the failure.
 * The reference reports the same conditions with exit(1)/abort() (MBG/main.cpp:158-198, MBG/ReadHelper.h:775-783);
 * a caller that wants MBG's process-level behaviour maps the codes back.  There is NO CPU fallback:
 * without a usable sm_100 device mbgfe_create() fails with MBGFE_E_CUDA.
 */
#ifndef MBGFE_H
#define MBGFE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MBGFE_ABI_VERSION 2

/* --error-masking, numbered like the reference enum ErrorMasking (MBG/ReadHelper.h:20-29) */
enum {
	MBGFE_MASK_NO = 0, MBGFE_MASK_HPC = 1, MBGFE_MASK_COLLAPSE = 2, MBGFE_MASK_DINUC = 3, MBGFE_MASK_MSAT = 4,
	MBGFE_MASK_COLLAPSE_DINUC = 5, MBGFE_MASK_COLLAPSE_MSAT = 6
};

enum {
	MBGFE_OK = 0,
	MBGFE_E_PARAM = -1,      /* MBG/main.cpp:158-198 validation: k odd >= 31, 1 <= w <= k-30, ... */
	MBGFE_E_CUDA = -2,       /* no device / CUDA runtime failure */
	MBGFE_E_IO = -3,         /* unreadable input (MBG/MBG.cpp:1542-1549) */
	MBGFE_E_PALINDROME = -4, /* palindromic k-mer that cannot be dropped (MBG/ReadHelper.h:775-777) */
	MBGFE_E_UNHASHABLE = -5, /* fwHash == bwHash on a non-palindrome (MBG/ReadHelper.h:782-783) */
	MBGFE_E_STATE = -6,      /* call order (results requested before a run, ...) */
	MBGFE_E_LIMIT = -7       /* input exceeds a documented limit (read longer than 2^32-1 bases, k too large, ...) */
};

/* flags */
enum {
	MBGFE_KEEP_CODES = 1,   /* keep the per-part code arrays for mbgfe_data(MBGFE_CODES)         (iterateParts)  */
	MBGFE_KEEP_POSES = 2,   /* also compute/keep the per-code positions (poses)                   (iterateParts)  */
	MBGFE_NO_TABLE   = 4,   /* stop after k-mer hashing: no node/edge table                       (iterateHashes) */
	MBGFE_NO_EDGES   = 8,   /* count k-mers (nodes, tips) but skip the adjacency pass                              */
	MBGFE_INCLUDE_END_KMERS = 16 /* --include-end-kmers (MBG/ReadHelper.cpp:20-30, 38-61): NOT built (Verkko never passes it);
	                                mbgfe_create() rejects it with MBGFE_E_PARAM instead of silently ignoring it        */
};

typedef struct {
	uint32_t abi_version;     /* MBGFE_ABI_VERSION */
	uint32_t k;               /* -k  (MBG/main.cpp:14) */
	uint32_t w;               /* -w  (MBG/main.cpp:15); s-mer size is k-w+1 (MBG/ReadHelper.h:1085) */
	uint32_t error_masking;   /* --error-masking (MBG/main.cpp:18) */
	uint32_t kmer_abundance;  /* -a  (MBG/main.cpp:16): minimum coverage of nodes and edges in the filtered tables */
	uint32_t flags;
	int32_t  device;          /* CUDA device ordinal */
	uint32_t rank;            /* this process's shard of the k-mer space, 0 <= rank < world (single GPU: 0 of 1) */
	uint32_t world;
	uint32_t piece_bytes;     /* mbgfe_run: size of one pipelined H2D piece (0 = 48 MiB); small values exist for tests */
	uint64_t table_slots;     /* initial node-table capacity in slots (0 = sized from the input); it grows on demand */
} mbgfe_params;

/* One read part == one callback of ReadpartIterator::iterateParts / iterateHashes (MBG/ReadHelper.h:167-235);
 * fields follow ReadInfo (MBG/ReadHelper.h:81-88).  Parts are listed in (read, offset) order. 48 bytes. */
typedef struct {
	uint32_t read_idx;     /* index of the read in the order the reads were added */
	uint32_t part_offset;  /* ReadInfo.readName.second */
	uint32_t raw_len;      /* ReadInfo.readLength */
	uint32_t n_codes;      /* ReadInfo.readLengthHpc == number of codes in the part */
	uint64_t code_begin;   /* index of the part's first code in MBGFE_CODES */
	uint64_t pose_begin;   /* index of the first of its n_codes+1 entries in MBGFE_POSES */
	uint64_t sel_begin;    /* index of its first selected k-mer in MBGFE_POSITIONS / MBGFE_HASHES */
	uint32_t n_sel;        /* number of selected k-mers (positions.size() in the reference callback) */
	uint32_t hashed;       /* 1 when n_codes >= k, i.e. the reference would call the iterateHashes callback */
} mbgfe_part;

typedef struct { uint64_t lo, hi; } mbgfe_hash128;     /* HashType == unsigned __int128 (MBG/MBGCommon.h:10) */

/* HashList after loadReadsAsHashesMultithread + sortByHash (MBG/MBG.cpp:44-76, MBG/HashList.cpp:245-353) */
typedef struct {
	uint32_t from, from_fw, to, to_fw; /* canonical form, MBG/MBGCommon.cpp:78-89; ids = rank of the canonical hash */
	uint32_t coverage;                 /* HashList::getEdgeCoverage */
	uint32_t overlap;                  /* HashList::getOverlap, in codes */
} mbgfe_edge;

/* result arrays for mbgfe_count()/mbgfe_data() */
enum {
	MBGFE_PARTS = 0,      /* mbgfe_part[]                                                                    */
	MBGFE_CODES = 1,      /* uint16_t[]  (needs MBGFE_KEEP_CODES)  SequenceCharType of each part             */
	MBGFE_POSES = 2,      /* uint32_t[]  (needs MBGFE_KEEP_POSES)  SequenceLengthType of each part           */
	MBGFE_POSITIONS = 3,  /* uint32_t[]  selected k-mer start positions, part-local, in codes                */
	MBGFE_HASHES = 4,     /* mbgfe_hash128[] forward hashes of the selected k-mers                           */
	MBGFE_NODE_HASHES = 5,/* mbgfe_hash128[] canonical hashes, ascending == node id order after sortByHash   */
	MBGFE_NODE_COVERAGE = 6, /* uint32_t[] HashList::coverage                                                */
	MBGFE_NODE_TIP = 7,   /* uint8_t[]   HashList::isTipKmer                                                 */
	MBGFE_EDGES = 8,      /* mbgfe_edge[] sorted by (from, from_fw descending, to, to_fw)                     */
	/* after mbgfe_build_unitigs(): */
	MBGFE_NODE_KEPT = 9,        /* uint8_t[]  per node of the table: survives the abundance filter (RankBitvector kept) */
	MBGFE_F_NODE_HASHES = 10,   /* mbgfe_hash128[] the HashList after HashList::filter: kept nodes, order unchanged   */
	MBGFE_F_NODE_COVERAGE = 11, /* uint32_t[]                                                                          */
	MBGFE_F_NODE_TIP = 12,      /* uint8_t[]                                                                           */
	MBGFE_F_EDGES = 13,         /* mbgfe_edge[] edges between kept nodes (every coverage), renumbered, same order      */
	MBGFE_UNITIG_BEGIN = 14,    /* uint64_t[n_unitigs + 1] offsets into MBGFE_UNITIG_NODES                             */
	MBGFE_UNITIG_NODES = 15,    /* uint32_t[] 2 * filtered node id + fw, in unitig order (UnitigGraph::unitigs)        */
	MBGFE_F_NODE_UNITIG = 16,   /* uint32_t[] unitig of each filtered node, 0xffffffff when its coverage is below -a   */
	MBGFE_UNITIG_EDGES = 17,    /* mbgfe_edge[] between unitig ends (UnitigGraph::edges / edgeCov, MBG/MBG.cpp:746-760):
	                               from / to = unitig, *_fw = read forward, coverage of the k-mer edge, overlap 0          */
	/* after mbgfe_thread_reads(): one entry per selected k-mer instance, in MBGFE_POSITIONS order */
	MBGFE_INST_UNITIG = 18,     /* uint32_t[] 2 * unitig + orientation as the read sees it; 0xffffffff = not located      */
	MBGFE_INST_UNITIG_POS = 19, /* uint32_t[] position inside that oriented unitig (leftClip of a path = its first entry)   */
	MBGFE_INST_UNITIG_REST = 20,/* uint32_t[] k-mers of the unitig behind it (rightClip of a path = its last entry)         */
	MBGFE_INST_STEP = 21,       /* uint8_t[]  what iterateReadPaths does at this k-mer, see mbgfe_thread_reads             */
	/* after mbgfe_thin_unitigs(): one entry per filtered node */
	MBGFE_F_NODE_THIN = 22,     /* uint8_t[]  1 when filterKmersToUnitigKmers keeps the k-mer                             */
	MBGFE_F_NODE_UNITIG_OFFSET = 23,/* uint32_t[] start of the k-mer inside its unitig, in codes (currentPos of that walk)  */
	/* the thinning applied (MBG/MBG.cpp:1166-1184): the HashList after reads.filter(kept), ids = ranks among the k-mers that stay */
	MBGFE_T_NODE_HASHES = 24,   /* mbgfe_hash128[]                                                                    */
	MBGFE_T_NODE_COVERAGE = 25, /* uint32_t[]                                                                         */
	MBGFE_T_NODE_TIP = 26,      /* uint8_t[]                                                                          */
	MBGFE_T_EDGES = 27,         /* mbgfe_edge[] edges (coverage + sequence overlap) between k-mers that stay, renumbered */
	MBGFE_T_UNITIG_BEGIN = 28,  /* uint64_t[n_unitigs + 1] the unitigs of MBGFE_UNITIG_*, thinned                       */
	MBGFE_T_UNITIG_NODES = 29,  /* uint32_t[] 2 * thinned node id + fw                                                 */
	MBGFE_T_NEW_OVERLAPS = 30   /* mbgfe_edge[] one per pair of neighbours of a thinned unitig, canon() form, coverage 0:
	                               the (from, to, overlap) triples the reference hands to addSequenceOverlap (MBG.cpp:1178-1183),
	                               which replace the overlap an edge of MBGFE_T_EDGES may hold for the same pair             */
};

typedef struct {
	uint64_t reads, raw_bases;           /* R */
	uint64_t segments;                   /* maximal ACGT stretches of reads with >= 32 bases */
	uint64_t collapsed_symbols;          /* C1: run-length / collapsed symbols */
	uint64_t codes;                      /* C2: final codes of all parts */
	uint64_t parts, hashed_parts;
	uint64_t selected;                   /* S: selected k-mer instances ("total selected k-mers in reads") */
	uint64_t distinct;                   /* D: table entries ("distinct selected k-mers in reads") */
	uint64_t distinct_edges;
	uint64_t table_slots, table_grows;
	uint64_t nonacgt;                    /* characters outside ACGTacgt */
	uint64_t h2d_bytes, d2h_bytes;       /* copied by the last upload / process / result download */
	/* device time of the last run, milliseconds, CUDA events on the library's stream */
	float ms_upload, ms_pack, ms_collapse, ms_msat, ms_select, ms_hash, ms_nodes, ms_edges, ms_process, ms_export;
	uint32_t launches;                   /* kernels launched by the last mbgfe_process() / mbgfe_build_unitigs() */
	float ms_unitigs;                    /* device time of the last mbgfe_build_unitigs() */
	float ms_exchange;                   /* device time of the last mbgfe_exchange(), first kernel to last */
	uint32_t exchange_launches;          /* kernels it launched */
	uint64_t exchange_bytes;             /* bytes this rank sent (node records + id replies + edge records) */
} mbgfe_stats;

typedef struct mbgfe mbgfe;

/* ReadpartIterator{k, w, errorMasking, ...} + HashList{k}  (MBG/ReadHelper.cpp:8-31, MBG/HashList.cpp:6-10) */
int mbgfe_create(const mbgfe_params* params, mbgfe** out);
void mbgfe_destroy(mbgfe* h);
/* last error text of the handle (h == NULL: of the last failed mbgfe_create in this thread) */
const char* mbgfe_last_error(const mbgfe* h);

/* Pinned host memory for input buffers (so the H2D copies inside mbgfe_upload run at full PCIe rate). */
void* mbgfe_alloc_pinned(uint64_t bytes);
void mbgfe_free_pinned(void* p);

/* Input.  Reads are ASCII (any case; everything outside ACGTacgt splits parts like the reference does):
 * read r is bases[offsets[r] .. offsets[r+1]).  The buffers are NOT copied and must stay valid until the next
 * mbgfe_upload()/mbgfe_run() returns.   Replaces iterateReadsMultithreaded + FastQ (MBG/ReadHelper.h:102-160). */
int mbgfe_add_reads(mbgfe* h, const char* bases, const uint64_t* offsets, uint64_t n_reads);
/* FASTA/FASTQ, optionally .gz, type chosen by suffix exactly like MBG/fastqloader.h:106-147 (unknown suffix == no
 * reads); sequences are upper-cased, names kept for mbgfe_read_name().  Unreadable file -> MBGFE_E_IO. */
int mbgfe_add_file(mbgfe* h, const char* path);
const char* mbgfe_read_name(const mbgfe* h, uint64_t read_idx);

/* hpc variants for pass 3 of runMBG (MBG/MBG.cpp:1564-1570): replaces ReadpartIterator::addHpcVariants
 * (MBG/ReadHelper.cpp:63-70), all entries in one call -- entry e == addHpcVariants(hashes[e], offsets[e],
 * lengths[len_begin[e] .. len_begin[e+1])).  From then on every selected k-mer whose forward or swapped hash is
 * registered gets the hash of its clustered run lengths XORed in (MBG/ReadHelper.h:680-737) before MBGFE_HASHES is
 * reported and before it is counted.  n_entries == 0 clears the set.  Survives mbgfe_reset(). */
int mbgfe_set_variants(mbgfe* h, const mbgfe_hash128* hashes, const uint32_t* offsets, const uint64_t* len_begin,
                       const uint32_t* lengths, uint64_t n_entries);

/* Forget reads and results, keep parameters and device buffers. */
int mbgfe_reset(mbgfe* h);

/* mbgfe_upload: host -> HBM copy of the pending reads and 2-bit packing on the device (reads become resident).
 * mbgfe_process: the whole front end on the resident reads: collapse / run-length -> microsatellite coding ->
 *   rolling s-mer hash + closed-syncmer selection -> 128-bit k-mer hashes -> node table (coverage, tips) -> edge
 *   table (coverage, overlaps).  May be called repeatedly; each call starts from an empty table.
 *   == one pass of loadReadsAsHashesMultithread (MBG/MBG.cpp:44-76) over ReadpartIterator::iterateHashes.
 * mbgfe_run = upload + process. */
int mbgfe_upload(mbgfe* h);
int mbgfe_process(mbgfe* h);
int mbgfe_run(mbgfe* h);

/* Input larger than one resident batch (the reference streams any volume through its queue into ONE HashList,
 * MBG/ReadHelper.h:102-160): mbgfe_run_append() = mbgfe_run() on the pending reads, but the selected k-mers of the batches before
 * it stay in the instance lists (25 bytes each on the device; MBGFE_POSITIONS / MBGFE_HASHES then cover all batches, MBGFE_PARTS /
 * MBGFE_CODES the last one) and no table is built.  When every batch is in: mbgfe_build_table() counts all of them into one node /
 * edge table (single GPU), or mbgfe_exchange() into the sharded one.  mbgfe_reset() / mbgfe_run() start over. */
int mbgfe_run_append(mbgfe* h);
int mbgfe_upload_append(mbgfe* h);     /* the same in two steps: upload the next batch ...                              */
int mbgfe_process_append(mbgfe* h);    /* ... and run the front end on it (repeatable: replaces this batch's instances) */
int mbgfe_build_table(mbgfe* h);

/* Results (downloaded and, for the tables, ordered on first request). */
uint64_t mbgfe_count(mbgfe* h, int what);
const void* mbgfe_data(mbgfe* h, int what);
int mbgfe_get_stats(const mbgfe* h, mbgfe_stats* out);

/* The step after the table: getUnitigGraph (MBG/MBG.cpp:629-760) with oneCovHeuristic = keepGaps = false, i.e. what
 * runMBG does in pass 1 (MBG/MBG.cpp:1558); mbgfe_build_unitigs_ex below adds the heuristic.  On the device, from the ordered table of the last run:
 *   1. covered edges = edges with coverage >= kmer_abundance (getCoveredEdges, MBG/MBG.cpp:367-388)
 *   2. every maximal non-branching chain (getUnitigHashes, MBG/MBG.cpp:305-320) whose mean k-mer coverage is
 *      >= unitig_abundance keeps its nodes (checkUnitigHashes, MBG/MBG.cpp:322-344)       -> MBGFE_NODE_KEPT
 *   3. HashList::filter (MBG/HashList.cpp:159-232)                                        -> MBGFE_F_*
 *   4. unitigs = the chains of the filtered list over nodes with coverage >= kmer_abundance (MBG/MBG.cpp:689-760)
 *      -> MBGFE_UNITIG_BEGIN / MBGFE_UNITIG_NODES / MBGFE_F_NODE_UNITIG, and the edges between unitig ends (every covered
 *      edge that is not inside a chain) -> MBGFE_UNITIG_EDGES.  The reference numbers and orients unitigs in
 *      the order its sequential walk meets them (arrival order of the k-mers); here a unitig starts at its end with the
 *      smaller node id (a circular one at its smallest node, forward) and unitigs are numbered by that first node.
 * kmer_abundance == 0 takes mbgfe_params.kmer_abundance.  Needs the edge table and world == 1. */
int mbgfe_build_unitigs(mbgfe* h, uint32_t kmer_abundance, double unitig_abundance);

/* getUnitigGraph's oneCovHeuristic argument (MBG/MBG.cpp:629-635): removeOnecovNodes (MBG/MBG.cpp:527-627) runs twice on
 * the list before step 1 -- first over tips only, then over every run.  A run is a maximal stretch of coverage-1 k-mers
 * along non-branching edges (all edges, coverage >= 1); it is dropped when everything its two ends lead to is also entered
 * from a k-mer of coverage >= 2 (tips pass: and exactly one end leads nowhere).  runMBG sets it for its final graph when
 * -u >= 2, -r > 0 and --do-unsafe-guesswork-resolutions (MBG/MBG.cpp:1573) -- Verkko's flags (Snakefiles/1-buildGraph.sm).
 * MBGFE_NODE_KEPT then marks the survivors of all three filters. */
#define MBGFE_UNITIG_ONECOV_HEURISTIC 1u
int mbgfe_build_unitigs_ex(mbgfe* h, uint32_t kmer_abundance, double unitig_abundance, uint32_t flags);

/* Read threading: getReadPaths (MBG/MBG.cpp:1235-1251) = getKmerLocator (MBG/KmerMatcher.cpp:3-31) + iterateReadPaths
 * (MBG/KmerMatcher.h:12-144) as one batched lookup.  The selected k-mers of the last run are still resident (the
 * reference re-reads and re-hashes the input for this pass); each one is looked up in the filtered list and the unitigs
 * of the last mbgfe_build_unitigs[_ex] and classified.  MBGFE_INST_STEP says what iterateReadPaths does there:
 *   0  k-mer not in the filtered list: skipped, state unchanged (KmerMatcher.h:29)
 *   1  k-mer in no unitig: the open path, if any, ends; its readPoses are NOT cleared (KmerMatcher.h:30-47)
 *   2  first located k-mer of the part or after a 1: a path starts, readPoses is appended to (KmerMatcher.h:55-67)
 *   3  a path starts after a break: readPoses cleared (KmerMatcher.h:68-93, 121-141)
 *   4  next k-mer of the same unitig: readPoses gets the position, rightClip shrinks (KmerMatcher.h:95-104)
 *   5  first k-mer of the next unitig: path gets MBGFE_INST_UNITIG (KmerMatcher.h:105-120)
 * A ReadPath is a run 2|3, (4|5)*; path = MBGFE_INST_UNITIG of its 2|3|5 entries, readPoses = MBGFE_POSITIONS of all its
 * entries, leftClip = MBGFE_INST_UNITIG_POS of the first, rightClip = MBGFE_INST_UNITIG_REST of the last.
 * Unitigs are numbered as mbgfe_build_unitigs numbers them.  world == 1.
 * After mbgfe_thin_unitigs() the reads are threaded through the THINNED list, which is what runMBG does by default (pass 4 follows
 * filterKmersToUnitigKmers, MBG/MBG.cpp:1580-1586): k-mers the thinning dropped are step 0, positions and clips count the k-mers
 * that stay (MBGFE_T_UNITIG_*), and neighbours of a thinned unitig are tested against their new overlap (KmerMatcher.h:71). */
int mbgfe_thread_reads(mbgfe* h);

/* filterKmersToUnitigKmers (MBG/MBG.cpp:1065-1185) with filterWithinUnitig = true, MBG's default: inside a unitig only its
 * two ends, tip k-mers and the k-mers needed to keep neighbours overlapping stay.  Computes, for the unitigs of the last
 * mbgfe_build_unitigs[_ex], MBGFE_F_NODE_THIN (the RankBitvector `kept` of that function) and MBGFE_F_NODE_UNITIG_OFFSET (two
 * neighbours of the thinned unitig overlap by k minus the difference of their offsets: the newOverlaps of MBG.cpp:1136-1141),
 * and applies them the way the reference does (reads.filter(kept), renumbered unitigs, MBG.cpp:1166-1177): MBGFE_T_*. */
int mbgfe_thin_unitigs(mbgfe* h);

/* Device-side stopwatch on the library's stream (CUDA events): start, any number of calls, stop -> milliseconds. */
int mbgfe_timer_start(mbgfe* h);
int mbgfe_timer_stop(mbgfe* h, float* ms);

/* Multi-GPU sharding by k-mer owner (the reference has no multi-device path -- its HashList is one
 * process-wide map behind a mutex, MBG/HashList.h:15-48).  One handle per GPU, reads sharded by rank in contiguous blocks,
 * canonical hashes range-partitioned in equal shares (owner = floor(world * F(hash.hi)); F(x) = x + x (1 - x) is the distribution of
 * the canonical high word, the smaller of two uniform words), so that concatenating the ranks' sorted partitions IS the global
 * sortByHash order.  After mbgfe_process() / mbgfe_run() with MBGFE_NO_TABLE on every rank, ONE collective call does the rest inside
 * the library:
 *   1. one kernel counts, per owner, the node records AND the edge records this rank holds (the owner of an edge is the owner of one
 *      of its two nodes -- the canonical `from` when the low bits of the two canonical hashes agree, else `to`: decided from the hashes
 *      alone, the same for both strands, and uniform over the ranks); one all-gather gives every rank the whole count matrix
 *   2. node records {canonical hash lo, hi, tip, reserved} (32 bytes) go to their owners (all-to-all-v); owners count them into their
 *      partition and remember the slot of every record
 *   3. partition sizes are all-gathered; global node id = nodes owned by lower ranks + rank inside the sorted partition, read from the
 *      slot (no second probe); ids go back to the senders (reverse all-to-all-v, 4 bytes each)
 *   4. edge records {canonical (from id, to id), (global instance index << 24) | overlap} (16 bytes) go to the edge's owner
 *      (keeps "last observation in input order" global, MBG/HashList.cpp:67-72); owners build their edge partition.  The partitions
 *      are each sorted; together (not concatenated) they are the edge table
 * Afterwards MBGFE_NODE_* / MBGFE_EDGES return this rank's partition with GLOBAL node ids (limit 2^31-1 nodes).
 * Transports: NCCL between processes (grouped ncclSend / ncclRecv; libnccl.so.2 is loaded on first use), peer copies between the
 * handles of one process (mbgfe_exchange_group), or the caller's own collectives on the library's device buffers (callbacks). */
typedef int (*mbgfe_allgather_fn)(void* ctx, const void* device_send, void* device_recv, uint64_t bytes_per_rank);
typedef int (*mbgfe_alltoallv_fn)(void* ctx, const void* device_send, const uint64_t* send_bytes, void* device_recv, const uint64_t* recv_bytes);
int mbgfe_comm_unique_id(void* id128);                              /* ncclGetUniqueId: rank 0 makes it, the caller hands it to every rank */
int mbgfe_comm_init_nccl(mbgfe* h, const void* id128);              /* ncclCommInitRank(world, rank) on the handle's device (collective) */
int mbgfe_comm_init_callbacks(mbgfe* h, mbgfe_alltoallv_fn alltoallv, mbgfe_allgather_fn allgather, void* ctx);
int mbgfe_exchange(mbgfe* h);                                       /* collective over the ranks of the communicator */
int mbgfe_exchange_group(mbgfe** handles, uint32_t n);              /* handles[r] = rank r of world n, all in this process */

/* Known-answer access to the device primitives (each runs a tiny kernel): per-code hash of the rolling hasher
 * (MBG/FastHasher.cpp:33-43), code complement (MBG/ErrorMaskHelper.cpp:43-100), k-mer hash (MBG/MBGCommon.cpp:25-35). */
int mbgfe_kat_tables(mbgfe* h, uint64_t* char_hash_30948, uint16_t* complement_30948);
int mbgfe_kat_kmer_hash(mbgfe* h, const uint16_t* codes, uint32_t k, mbgfe_hash128* out);

/* mbgfe_kat_kmer_hash without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, same result.  Its device scratch (the code slot, the one-instance tables and the result) is a static
 * __device__ array of about 2 * MBGFE_NOALLOC_MAX_K + 512 bytes that every GPU loading the library holds until exit; the
 * call makes no cudaMalloc / cudaFree / allocating library call.  k > MBGFE_NOALLOC_MAX_K is rejected with MBGFE_E_LIMIT.
 * Every call on a device shares that one array: calls that overlap on the GPU (two handles on one device, two threads) race;
 * serialise them.  The call clears what it needs zeroed, so results do not depend on the previous call. */
#define MBGFE_NOALLOC_MAX_K 16384u
int mbgfe_kat_kmer_hash_noalloc(mbgfe* h, const uint16_t* codes, uint32_t k, mbgfe_hash128* out);

#ifdef __cplusplus
}
#endif
#endif
