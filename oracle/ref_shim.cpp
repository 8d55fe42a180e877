// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE, not product code.
//
// A thin extern "C" shim (written for this repo) over the UNMODIFIED reference MBG objects compiled by
// oracle/Makefile from /root/reference/src/hifioverlapper/MBG/src (commit 7bef84dd). It drives the reference's own
// front end one read at a time and flattens what its callbacks receive, so that tests can compare the CUDA path,
// the plain-C restatement (mbg_oracle.c) and the reference on identical inputs.
//
// Reference entry points used (all public in the reference):
//   ReadpartIterator::iteratePartsOfRead / iterateHashesOfRead   ReadHelper.h:215-235
//   ReadpartIterator::setMemoryReads                              ReadHelper.cpp:96-113
//   loadReadsAsHashesMultithread                                  MBG.cpp:44-76   (external linkage)
//   HashList::sortByHash / getEdgeCoverages / getSequenceOverlaps HashList.cpp:245-353,42-45,62-65
//   FastHasher, getHash-derived tables, complement(), maxCode()   FastHasher.h, ErrorMaskHelper.cpp
#include <cstdint>
#include <chrono>
#include <cstring>
#include <cstdio>
#include <sstream>
#include <string>
#include <vector>
#include <fstream>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <algorithm>
#include <phmap.h>
#include <zstr.hpp>
// the hpc-variant table of ReadpartIterator (ReadHelper.h:250) is private and has no getter; the T4 test has to read it back
// after the reference's own passes 1 and 2 (refshim_hpc_variants), so this translation unit -- test infrastructure -- opens it
#define private public
#include "ReadHelper.h"
#undef private
#include "HashList.h"
#include "FastHasher.h"
#include "ErrorMaskHelper.h"
#include "MBGCommon.h"

uint64_t getHash(uint64_t key); // FastHasher.cpp:33
void loadReadsAsHashesMultithread(HashList& result, const size_t kmerSize, const ReadpartIterator& partIterator, const size_t numThreads, std::ostream& log);
#include "UnitigGraph.h"
#include "KmerMatcher.h"
UnitigGraph getUnitigGraph(HashList& hashlist, const size_t minCoverage, const double minUnitigCoverage, const bool keepGaps, const bool oneCovHeuristic); // MBG.cpp:629
UnitigGraph getUnitigs(const UnitigGraph& oldgraph); // MBG.cpp:782
void filterKmersToUnitigKmers(UnitigGraph& unitigs, HashList& reads, const size_t kmerSize, const bool filterWithinUnitig); // MBG.cpp:1065
void sortKmersByHashes(UnitigGraph& unitigs, HashList& reads); // MBG.cpp:1059
#include "UnitigResolver.h"
#include "HPCConsensus.h"
#include "StringIndex.h"
#include "DumbSelect.h"
#include "CompressedSequence.h"
struct AssemblyStats { size_t nodes, edges, size, N50, approxKmers; };     // MBG.cpp:34-42 (defined in the .cpp only)
std::vector<ReadPath> getReadPaths(const UnitigGraph& graph, const HashList& hashlist, const size_t numThreads, const ReadpartIterator& partIterator, const size_t kmerSize); // MBG.cpp:1235
void verifyEdgeConsistency(const UnitigGraph& unitigs, const HashList& hashlist, const StringIndex& stringIndex, const std::vector<CompressedSequenceType>& unitigSequences, const size_t kmerSize); // MBG.cpp:1220
std::vector<double> getRawKmerCoverages(const UnitigGraph& unitigs, const std::vector<CompressedSequenceType>& unitigSequences, const HashList& reads, const size_t kmerSize); // MBG.cpp:1253
AssemblyStats writeGraph(const UnitigGraph& unitigs, const std::string& filename, const HashList& hashlist, const std::vector<CompressedSequenceType>& unitigSequences, const StringIndex& stringIndex, const size_t kmerSize, const std::vector<double>& unitigRawKmerCoverages, const std::string& nodeNamePrefix); // MBG.cpp:941
std::vector<DumbSelect> getUnitigExpandedPoses(const HashList& hashlist, const UnitigGraph& unitigs, const std::vector<CompressedSequenceType>& unitigSequences, const StringIndex& stringIndex); // MBG.cpp:1514
void sortPaths(std::vector<ReadPath>& readPaths); // MBG.cpp:1398
void writePaths(const HashList& hashlist, const UnitigGraph& unitigs, const std::vector<CompressedSequenceType>& unitigSequences, const StringIndex& stringIndex, const std::vector<DumbSelect>& unitigExpandedPoses, const std::vector<ReadPath>& readPaths, const size_t kmerSize, const std::string& outputSequencePaths, const std::string& nodeNamePrefix); // MBG.cpp:115
void printUnitigKmerCount(const UnitigGraph& unitigs); // MBG.cpp:1049

namespace {

struct Part {
	uint32_t read_idx, part_offset, raw_len, n_codes;
	uint64_t code_begin, pose_begin, sel_begin;
	uint32_t n_sel, hashed;
};

struct Run {
	std::vector<Part> parts;
	std::vector<uint16_t> codes;
	std::vector<uint32_t> poses;
	std::vector<uint32_t> positions;
	std::vector<uint64_t> hashes; // lo, hi interleaved
	// table
	std::vector<uint64_t> node_hash; // lo, hi interleaved, ascending canonical hash
	std::vector<uint32_t> node_cov;
	std::vector<uint8_t> node_tip;
	std::vector<uint32_t> edges; // from_id, from_fw, to_id, to_fw, coverage, overlap
	std::vector<uint32_t> uedges; // unitig graph edges: from unitig, from_fw, to unitig, to_fw, coverage, overlap -- canonical
	// read paths (iterateReadPaths): per path read index, part offset, path entries, readPoses entries, leftClip, rightClip
	std::vector<uint64_t> rp_head;
	std::vector<uint32_t> rp_path;  // 2*unitig + fw
	std::vector<uint32_t> rp_poses;
	std::vector<uint32_t> overlaps; // HashList::sequenceOverlap, canonical pairs: from, from_fw, to, to_fw, overlap
	uint64_t total_selected = 0;
};

}

extern "C" {

// mode numbering = reference enum ErrorMasking (ReadHelper.h:20-29): 0 no, 1 hpc, 2 collapse, 3 dinuc, 4 msat,
// 5 collapse-dinuc, 6 collapse-msat.
void* refshim_run(int mode, uint64_t k, uint64_t w, const char* bases, const uint64_t* offsets, uint64_t n_reads, int build_table)
{
	Run* run = new Run;
	std::vector<std::string> nofiles;
	ReadpartIterator iterator { (size_t)k, (size_t)w, (ErrorMasking)mode, 1, nofiles, false, "" };
	std::vector<std::pair<std::string, std::string>> memoryReads;
	for (uint64_t r = 0; r < n_reads; r++)
	{
		std::string seq { bases + offsets[r], bases + offsets[r+1] };
		if (seq.size() < 32) continue; // ReadHelper.h:838 (file path drops short reads before masking)
		std::string name = std::to_string(r);
		size_t firstPart = run->parts.size();
		iterator.iteratePartsOfRead(name, seq, [run, r](const ReadInfo& read, const SequenceCharType& codes, const SequenceLengthType& poses, const std::string& raw)
		{
			Part p;
			p.read_idx = (uint32_t)r;
			p.part_offset = (uint32_t)read.readName.second;
			p.raw_len = (uint32_t)read.readLength;
			p.n_codes = (uint32_t)codes.size();
			p.code_begin = run->codes.size();
			p.pose_begin = run->poses.size();
			p.sel_begin = 0;
			p.n_sel = 0;
			p.hashed = 0;
			run->codes.insert(run->codes.end(), codes.begin(), codes.end());
			for (size_t v : poses) run->poses.push_back((uint32_t)v);
			run->parts.push_back(p);
		});
		iterator.iterateHashesOfRead(name, seq, [run, firstPart](const ReadInfo& read, const SequenceCharType& codes, const SequenceLengthType& poses, const std::string& raw, const std::vector<size_t>& positions, const std::vector<HashType>& hashes)
		{
			for (size_t i = firstPart; i < run->parts.size(); i++)
			{
				Part& p = run->parts[i];
				if (p.part_offset != read.readName.second) continue;
				p.hashed = 1;
				p.sel_begin = run->positions.size();
				p.n_sel = (uint32_t)positions.size();
				for (size_t j = 0; j < positions.size(); j++)
				{
					run->positions.push_back((uint32_t)positions[j]);
					run->hashes.push_back((uint64_t)hashes[j]);
					run->hashes.push_back((uint64_t)(hashes[j] >> 64));
				}
				return;
			}
		});
		if (build_table) memoryReads.emplace_back(name, seq);
	}
	if (build_table)
	{
		ReadpartIterator tableIterator { (size_t)k, (size_t)w, (ErrorMasking)mode, 1, nofiles, false, "" };
		tableIterator.setMemoryReads(memoryReads);
		HashList list { (size_t)k };
		std::ostringstream log;
		if (memoryReads.size() > 0) loadReadsAsHashesMultithread(list, k, tableIterator, 1, log);
		list.sortByHash();
		size_t n = list.size();
		run->node_hash.resize(2*n);
		run->node_cov.resize(n);
		run->node_tip.resize(n);
		for (auto pair : list.hashToNode)
		{
			run->node_hash[2*pair.second] = (uint64_t)pair.first;
			run->node_hash[2*pair.second+1] = (uint64_t)(pair.first >> 64);
		}
		for (size_t i = 0; i < n; i++)
		{
			run->node_cov[i] = (uint32_t)list.coverage.get(i);
			run->node_tip[i] = list.isTipKmer(i) ? 1 : 0;
			run->total_selected += list.coverage.get(i);
			for (int fw = 1; fw >= 0; fw--)
			{
				std::pair<size_t, bool> from { i, fw == 1 };
				for (auto edge : list.getEdgeCoverages(from))
				{
					run->edges.push_back((uint32_t)i);
					run->edges.push_back((uint32_t)fw);
					run->edges.push_back((uint32_t)edge.first.first);
					run->edges.push_back(edge.first.second ? 1 : 0);
					run->edges.push_back((uint32_t)edge.second);
					run->edges.push_back((uint32_t)list.getOverlap(from, edge.first));
				}
			}
		}
	}
	return run;
}

// Pass 3 of runMBG (MBG.cpp:1570): the file-driven iterator with hpc variants registered (ReadpartIterator::addHpcVariants,
// ReadHelper.cpp:63-70; applied in iterateHashesFromFiles, ReadHelper.h:672-737 -- only the file path applies them, so
// the reads are written to a scratch FASTA first).  Entry e == one addHpcVariants(hash e, offsets[e], lengths[...]) call.
// One worker thread => callbacks arrive in file order.  Fills parts (hashed ones only), positions, hashes and the table.
void* refshim_run_variants(int mode, uint64_t k, uint64_t w, const char* bases, const uint64_t* offsets, uint64_t n_reads, const char* scratch_fasta,
	const uint64_t* var_hashes_lo_hi, const uint32_t* var_offsets, const uint64_t* var_len_begin, const uint32_t* var_lengths, uint64_t n_var)
{
	Run* run = new Run;
	{
		FILE* f = fopen(scratch_fasta, "w");
		if (!f) return run;
		for (uint64_t r = 0; r < n_reads; r++)
		{
			fprintf(f, ">%llu\n", (unsigned long long)r);
			fwrite(bases + offsets[r], 1, offsets[r+1] - offsets[r], f);
			fputc('\n', f);
		}
		fclose(f);
	}
	std::vector<std::string> files { std::string(scratch_fasta) };
	ReadpartIterator iterator { (size_t)k, (size_t)w, (ErrorMasking)mode, 1, files, false, "" };
	for (uint64_t e = 0; e < n_var; e++)
	{
		HashType h = ((HashType)var_hashes_lo_hi[2*e+1] << 64) + (HashType)var_hashes_lo_hi[2*e];
		std::vector<size_t> lens { var_lengths + var_len_begin[e], var_lengths + var_len_begin[e+1] };
		iterator.addHpcVariants(h, var_offsets[e], lens);
	}
	iterator.iterateHashes([run](const ReadInfo& read, const SequenceCharType& codes, const SequenceLengthType& poses, const std::string& raw, const std::vector<size_t>& positions, const std::vector<HashType>& hashes)
	{
		Part p;
		p.read_idx = (uint32_t)std::stoull(read.readName.first);
		p.part_offset = (uint32_t)read.readName.second;
		p.raw_len = (uint32_t)read.readLength;
		p.n_codes = (uint32_t)codes.size();
		p.code_begin = 0; p.pose_begin = 0;
		p.sel_begin = run->positions.size();
		p.n_sel = (uint32_t)positions.size();
		p.hashed = 1;
		for (size_t j = 0; j < positions.size(); j++)
		{
			run->positions.push_back((uint32_t)positions[j]);
			run->hashes.push_back((uint64_t)hashes[j]);
			run->hashes.push_back((uint64_t)(hashes[j] >> 64));
		}
		run->parts.push_back(p);
	});
	HashList list { (size_t)k };
	std::ostringstream log;
	loadReadsAsHashesMultithread(list, k, iterator, 1, log);
	list.sortByHash();
	size_t n = list.size();
	run->node_hash.resize(2*n);
	run->node_cov.resize(n);
	run->node_tip.resize(n);
	for (auto pair : list.hashToNode)
	{
		run->node_hash[2*pair.second] = (uint64_t)pair.first;
		run->node_hash[2*pair.second+1] = (uint64_t)(pair.first >> 64);
	}
	for (size_t i = 0; i < n; i++)
	{
		run->node_cov[i] = (uint32_t)list.coverage.get(i);
		run->node_tip[i] = list.isTipKmer(i) ? 1 : 0;
		run->total_selected += list.coverage.get(i);
		for (int fw = 1; fw >= 0; fw--)
		{
			std::pair<size_t, bool> from { i, fw == 1 };
			for (auto edge : list.getEdgeCoverages(from))
			{
				run->edges.push_back((uint32_t)i);
				run->edges.push_back((uint32_t)fw);
				run->edges.push_back((uint32_t)edge.first.first);
				run->edges.push_back(edge.first.second ? 1 : 0);
				run->edges.push_back((uint32_t)edge.second);
				run->edges.push_back((uint32_t)list.getOverlap(from, edge.first));
			}
		}
	}
	remove(scratch_fasta);
	return run;
}

// The step after the table exactly as runMBG takes it in pass 1 (MBG.cpp:1557-1558): loadReadsAsHashesMultithread, then
// getUnitigGraph(reads, minCoverage, minUnitigCoverage, keepGaps = false, oneCovHeuristic = false) on the ARRIVAL-ordered
// list; it filters the HashList in place (HashList::filter) and returns the unitigs.  Output, keyed by hash so that it
// does not depend on node numbering: the filtered list in the Run's node_* / edges arrays (ids = filtered arrival
// order), and every unitig as a run of (node id, fw) pairs in `positions` (2*id + fw) with `parts[u].sel_begin/n_sel`.
void* refshim_unitigs_ex(int mode, uint64_t k, uint64_t w, const char* bases, const uint64_t* offsets, uint64_t n_reads, uint64_t min_cov, double min_unitig_cov, int one_cov_heuristic);
void* refshim_unitigs(int mode, uint64_t k, uint64_t w, const char* bases, const uint64_t* offsets, uint64_t n_reads, uint64_t min_cov, double min_unitig_cov)
{
	return refshim_unitigs_ex(mode, k, w, bases, offsets, n_reads, min_cov, min_unitig_cov, 0);
}

// one_cov_heuristic: getUnitigGraph's last argument (MBG.cpp:629-635; runMBG sets it when -u >= 2, -r > 0 and guesswork, MBG.cpp:1573)
void* refshim_unitigs_ex(int mode, uint64_t k, uint64_t w, const char* bases, const uint64_t* offsets, uint64_t n_reads, uint64_t min_cov, double min_unitig_cov, int one_cov_heuristic)
{
	Run* run = new Run;
	std::vector<std::string> nofiles;
	std::vector<std::pair<std::string, std::string>> memoryReads;
	for (uint64_t r = 0; r < n_reads; r++)
	{
		std::string seq { bases + offsets[r], bases + offsets[r+1] };
		if (seq.size() < 32) continue;
		memoryReads.emplace_back(std::to_string(r), seq);
	}
	ReadpartIterator tableIterator { (size_t)k, (size_t)w, (ErrorMasking)mode, 1, nofiles, false, "" };
	tableIterator.setMemoryReads(memoryReads);
	HashList list { (size_t)k };
	std::ostringstream log;
	if (memoryReads.size() > 0) loadReadsAsHashesMultithread(list, k, tableIterator, 1, log);
	UnitigGraph unitigs = getUnitigGraph(list, min_cov, min_unitig_cov, false, (one_cov_heuristic & 1) != 0);
	if (one_cov_heuristic & 2)
	{
		// the rest of runMBG's pass 3 up to the read paths (MBG.cpp:1575-1581) with MBG's defaults
		if (min_unitig_cov > (double)min_cov) unitigs = getUnitigs(unitigs.filterUnitigsByCoverage(min_unitig_cov, false));
		filterKmersToUnitigKmers(unitigs, list, (size_t)k, true);
		sortKmersByHashes(unitigs, list);
	}
	size_t n = list.size();
	run->node_hash.resize(2*n);
	run->node_cov.resize(n);
	run->node_tip.resize(n);
	for (auto pair : list.hashToNode)
	{
		run->node_hash[2*pair.second] = (uint64_t)pair.first;
		run->node_hash[2*pair.second+1] = (uint64_t)(pair.first >> 64);
	}
	for (size_t i = 0; i < n; i++)
	{
		run->node_cov[i] = (uint32_t)list.coverage.get(i);
		run->node_tip[i] = list.isTipKmer(i) ? 1 : 0;
		for (int fw = 1; fw >= 0; fw--)
		{
			std::pair<size_t, bool> from { i, fw == 1 };
			for (auto edge : list.getEdgeCoverages(from))
			{
				run->edges.push_back((uint32_t)i);
				run->edges.push_back((uint32_t)fw);
				run->edges.push_back((uint32_t)edge.first.first);
				run->edges.push_back(edge.first.second ? 1 : 0);
				run->edges.push_back((uint32_t)edge.second);
				run->edges.push_back((uint32_t)list.getOverlap(from, edge.first));
			}
		}
	}
	for (size_t u = 0; u < unitigs.unitigs.size(); u++)
	{
		Part p;
		memset(&p, 0, sizeof(p));
		p.read_idx = (uint32_t)u;
		p.sel_begin = run->positions.size();
		p.n_sel = (uint32_t)unitigs.unitigs[u].size();
		for (auto node : unitigs.unitigs[u]) run->positions.push_back((uint32_t)(2 * node.first + (node.second ? 1 : 0)));
		run->parts.push_back(p);
	}
	// UnitigGraph::edges / edgeCov (set by getUnitigGraph's last loop, MBG.cpp:746-760), every edge once in canon() form
	for (size_t u = 0; u < unitigs.unitigs.size(); u++)
	{
		for (int fw = 1; fw >= 0; fw--)
		{
			std::pair<size_t, bool> from { u, fw == 1 };
			for (auto to : unitigs.edges[from])
			{
				auto c = canon(from, to);
				if (c.first != from || c.second != to) continue;       // listed from its mirror image
				run->uedges.push_back((uint32_t)from.first); run->uedges.push_back(from.second ? 1 : 0);
				run->uedges.push_back((uint32_t)to.first); run->uedges.push_back(to.second ? 1 : 0);
				run->uedges.push_back((uint32_t)unitigs.edgeCoverage(from, to));
				run->uedges.push_back((uint32_t)unitigs.edgeOverlap(from, to));
			}
		}
	}
	for (size_t i = 0; i < n; i++)
	{
		for (int fw = 1; fw >= 0; fw--)
		{
			std::pair<size_t, bool> from { i, fw == 1 };
			for (auto pair : list.getSequenceOverlaps(from))
			{
				run->overlaps.push_back((uint32_t)i); run->overlaps.push_back((uint32_t)fw);
				run->overlaps.push_back((uint32_t)pair.first.first); run->overlaps.push_back(pair.first.second ? 1 : 0);
				run->overlaps.push_back((uint32_t)pair.second);
			}
		}
	}
	// getReadPaths (MBG.cpp:1235-1251): every part threaded through the unitig graph (KmerMatcher.h:12-144), in input order
	if (memoryReads.size() > 0)
	{
		std::vector<std::tuple<size_t, size_t, bool>> kmerLocator = getKmerLocator(unitigs);
		tableIterator.iterateOnlyHashes([run, &kmerLocator, k, &unitigs, &list](const ReadInfo& read, const std::vector<size_t>& positions, const std::vector<HashType>& hashes)
		{
			iterateReadPaths(unitigs, list, (size_t)k, kmerLocator, read, positions, hashes, [run](ReadPath path)
			{
				run->rp_head.push_back(std::stoull(path.readName.first));
				run->rp_head.push_back(path.readName.second);
				run->rp_head.push_back(path.path.size());
				run->rp_head.push_back(path.readPoses.size());
				run->rp_head.push_back(path.leftClip);
				run->rp_head.push_back(path.rightClip);
				for (auto node : path.path) run->rp_path.push_back((uint32_t)(2 * node.id() + (node.forward() ? 1 : 0)));
				for (auto pos : path.readPoses) run->rp_poses.push_back(pos);
			});
		});
	}
	return run;
}

void refshim_free(void* h) { delete (Run*)h; }

uint64_t refshim_count(void* h, int what)
{
	Run* r = (Run*)h;
	switch (what)
	{
		case 0: return r->parts.size();
		case 1: return r->codes.size();
		case 2: return r->poses.size();
		case 3: return r->positions.size();
		case 4: return r->node_cov.size();
		case 5: return r->edges.size() / 6;
		case 6: return r->total_selected;
		case 7: return r->uedges.size() / 6;
		case 8: return r->rp_head.size() / 6;
		case 9: return r->rp_path.size();
		case 10: return r->rp_poses.size();
		case 11: return r->overlaps.size() / 5;
	}
	return 0;
}

const void* refshim_data(void* h, int what)
{
	Run* r = (Run*)h;
	switch (what)
	{
		case 0: return r->parts.data();
		case 1: return r->codes.data();
		case 2: return r->poses.data();
		case 3: return r->positions.data();
		case 4: return r->hashes.data();
		case 5: return r->node_hash.data();
		case 6: return r->node_cov.data();
		case 7: return r->node_tip.data();
		case 8: return r->edges.data();
		case 9: return r->uedges.data();
		case 10: return r->rp_head.data();
		case 11: return r->rp_path.data();
		case 12: return r->rp_poses.data();
		case 13: return r->overlaps.data();
	}
	return nullptr;
}

// The reference's own hot path, timed: ReadpartIterator over a FASTA/FASTQ file with `threads` worker threads feeding
// loadReadsAsHashesMultithread (exactly what `MBG -t threads` runs in its "selecting k-mers and building graph
// topology" phase before unitig construction, MBG/MBG.cpp:1570).  Returns seconds; fills selected / distinct counts.
double refshim_time_file(int mode, uint64_t k, uint64_t w, const char* path, uint64_t threads, uint64_t* total_selected, uint64_t* distinct)
{
	std::vector<std::string> files { std::string(path) };
	ReadpartIterator iterator { (size_t)k, (size_t)w, (ErrorMasking)mode, (size_t)threads, files, false, "" };
	HashList list { (size_t)k };
	std::ostringstream log;
	auto t0 = std::chrono::steady_clock::now();
	loadReadsAsHashesMultithread(list, k, iterator, threads, log);
	auto t1 = std::chrono::steady_clock::now();
	uint64_t total = 0;
	for (size_t i = 0; i < list.size(); i++) total += list.coverage.get(i);
	if (total_selected) *total_selected = total;
	if (distinct) *distinct = list.size();
	return std::chrono::duration<double>(t1 - t0).count();
}


// Order-free digest of a HashList (independent of node numbering and of which mirror image of an edge is stored), so
// that a table built elsewhere -- the CUDA path, on the same reads -- can be compared with the reference's without
// moving the table: out[0] = sum of coverage (selected k-mers), [1] = distinct k-mers, [2] = distinct edges,
// [3] = XOR of mix(hash) * (2 cov + 1), [4] = sum of mix(hash) * cov, [5] = tip k-mers, [6] = sum of edge coverage,
// [7] = sum of sym(edge) * cov, [8] = XOR of sym(edge) * (2 cov + 1), [9] = sum of sym(edge) * overlap, all mod 2^64, with
// mix(h) = lo * 0x9E3779B97F4A7C15 + hi * 0xC2B2AE3D27D4EB4F, v(node, fw) = mix * (fw ? 3 : 5),
// e(a, b) = v(a) * 7 + v(b) * 11 and sym = e(from, to) + e(reverse to, reverse from).
// Tips ([5]) and overlaps ([9]) are only deterministic with one worker thread.
// tests/oracle_lib.py:table_digest computes the same from exported arrays.
static void hashlist_digest(HashList& list, uint64_t* out)
{
	const size_t n = list.size();
	std::vector<uint64_t> mix(n);
	for (auto pair : list.hashToNode) mix[pair.second] = (uint64_t)pair.first * 0x9E3779B97F4A7C15ull + (uint64_t)(pair.first >> 64) * 0xC2B2AE3D27D4EB4Full;
	for (int i = 0; i < 10; i++) out[i] = 0;
	out[1] = n;
	auto v = [&](std::pair<size_t, bool> node) { return mix[node.first] * (node.second ? 3ull : 5ull); };
	for (size_t i = 0; i < n; i++)
	{
		const uint64_t cov = list.coverage.get(i);
		out[0] += cov;
		out[3] ^= mix[i] * (2 * cov + 1);
		out[4] += mix[i] * cov;
		out[5] += list.isTipKmer(i) ? 1 : 0;
		for (int fw = 1; fw >= 0; fw--)
		{
			std::pair<size_t, bool> from { i, fw == 1 };
			for (auto edge : list.getEdgeCoverages(from))
			{
				const std::pair<size_t, bool> to = edge.first;
				const std::pair<size_t, bool> rto { to.first, !to.second }, rfrom { from.first, !from.second };
				const uint64_t sym = (v(from) * 7 + v(to) * 11) + (v(rto) * 7 + v(rfrom) * 11);
				const uint64_t ecov = edge.second;
				out[2] += 1;
				out[6] += ecov;
				out[7] += sym * ecov;
				out[8] ^= sym * (2 * ecov + 1);
				out[9] += sym * (uint64_t)list.getOverlap(from, to);
			}
		}
	}
}

// refshim_time_file + the digest of the table it built (10 words, see hashlist_digest)
double refshim_time_file_digest(int mode, uint64_t k, uint64_t w, const char* path, uint64_t threads, uint64_t* digest10)
{
	std::vector<std::string> files { std::string(path) };
	ReadpartIterator iterator { (size_t)k, (size_t)w, (ErrorMasking)mode, (size_t)threads, files, false, "" };
	HashList list { (size_t)k };
	std::ostringstream log;
	auto t0 = std::chrono::steady_clock::now();
	loadReadsAsHashesMultithread(list, k, iterator, threads, log);
	auto t1 = std::chrono::steady_clock::now();
	if (digest10) hashlist_digest(list, digest10);
	return std::chrono::duration<double>(t1 - t0).count();
}


// ---------------------------------------------------------------------------------------------------------------------------
// T4: the tables of libmbgfe handed to the reference's OWN back end -- the binding INTEGRATION.md 1 describes,
// compiled.  The reference's containers are filled through their public members only (HashList::addNode / coverage.set /
// setTipKmer / setEdgeCoverage / addSequenceOverlap, MBG/HashList.h:15-48; UnitigGraph's public vectors and setters,
// MBG/UnitigGraph.h:13-38) and the rest of runMBG (MBG/MBG.cpp:1573-1628) runs unchanged, function by function, writing the GFA
// and the GAF.  Nodes arrive in ascending canonical-hash order (MBGFE_NODE_* / MBGFE_T_*), i.e. already as sortByHash leaves them.
//
//   stage 0  replaces loadReadsAsHashesMultithread (MBG.cpp:1570): HashList <- node table + edge table; the reference continues
//            from getUnitigGraph (MBG.cpp:1573)
//   stage 1  replaces everything up to sortKmersByHashes (MBG.cpp:1570-1581): HashList <- thinned list (+ new overlaps),
//            UnitigGraph <- thinned unitigs + unitig edges; the reference continues from getReadPaths (MBG.cpp:1586)
// Passes 4 and 5 (read paths, unitig sequences) read the FASTA through the reference's own iterator.
static void fill_hashlist(HashList& reads, const uint64_t* node_hash, const uint32_t* node_cov, const uint8_t* node_tip, uint64_t n_nodes, const uint32_t* edges6, uint64_t n_edges)
{
	for (uint64_t i = 0; i < n_nodes; i++)
	{
		const HashType hsh = ((HashType)node_hash[2 * i + 1] << 64) + (HashType)node_hash[2 * i];      // canonical: smaller than its swapped form
		const std::pair<size_t, bool> node = reads.addNode(hsh);
		if (node.first != i || !node.second) abort();
		reads.coverage.set(i, node_cov[i]);
		if (node_tip[i]) reads.setTipKmer(i);
	}
	for (uint64_t e = 0; e < n_edges; e++)
	{
		const uint32_t* r = edges6 + 6 * e;
		const std::pair<size_t, bool> from { r[0], r[1] != 0 }, to { r[2], r[3] != 0 };
		reads.addSequenceOverlap(from, to, r[5]);
		reads.setEdgeCoverage(from, to, r[4]);
	}
}

// Passes 1 and 2 of runMBG under --hpc-variant-onecopy-coverage (MBG.cpp:1554-1567), run by the reference itself: the variant table
// it leaves in its iterator, flattened the way mbgfe_set_variants takes it (entry e: hash, offset, lengths[len_begin[e] .. len_begin[e+1])).
// Row f3 (consensus + variant discovery) is not built, so this table is an INPUT of the CUDA path in the T4 test.
static std::vector<uint64_t> g_var_hash; static std::vector<uint32_t> g_var_offset, g_var_len; static std::vector<uint64_t> g_var_begin;
uint64_t refshim_hpc_variants(int mode, uint64_t k, uint64_t w, const char* fasta, uint64_t min_cov, double min_unitig_cov, double onecopy_coverage)
{
	std::vector<std::string> files { std::string(fasta) };
	ReadpartIterator partIterator { (size_t)k, (size_t)w, (ErrorMasking)mode, 1, files, false, "" };
	HashList reads { (size_t)k };
	std::ostringstream log;
	loadReadsAsHashesMultithread(reads, k, partIterator, 1, log);
	auto unitigs = getUnitigGraph(reads, min_cov, min_unitig_cov, false, false);
	if (min_unitig_cov > (double)min_cov) unitigs = getUnitigs(unitigs.filterUnitigsByCoverage(min_unitig_cov, false));
	sortKmersByHashes(unitigs, reads);
	getHpcVariantsAndReadPaths(reads, unitigs, k, partIterator, 1, onecopy_coverage * 1.5, onecopy_coverage * 0.5);
	g_var_hash.clear(); g_var_offset.clear(); g_var_len.clear(); g_var_begin.assign(1, 0);
	for (const auto& pair : partIterator.hpcVariants)
		for (const auto& entry : pair.second)
		{
			g_var_hash.push_back((uint64_t)pair.first); g_var_hash.push_back((uint64_t)(pair.first >> 64));
			g_var_offset.push_back((uint32_t)entry.first);
			for (size_t len : entry.second) g_var_len.push_back((uint32_t)len);
			g_var_begin.push_back(g_var_len.size());
		}
	return g_var_offset.size();
}
const void* refshim_hpc_variants_data(int what)
{
	switch (what) { case 0: return g_var_hash.data(); case 1: return g_var_offset.data(); case 2: return g_var_begin.data(); case 3: return g_var_len.data(); }
	return nullptr;
}
uint64_t refshim_hpc_variants_lengths() { return g_var_len.size(); }

int refshim_backend_variants(int stage, int mode, uint64_t k, uint64_t w, const char* fasta, uint64_t min_cov, double min_unitig_cov, uint64_t max_resolve, uint64_t max_uncond_resolve,
	int guesswork, int copycount_heuristic, int only_local_resolve, int do_cleaning,
	const uint64_t* node_hash, const uint32_t* node_cov, const uint8_t* node_tip, uint64_t n_nodes, const uint32_t* edges6, uint64_t n_edges,
	const uint32_t* new_overlaps6, uint64_t n_new_overlaps, const uint64_t* unitig_begin, const uint32_t* unitig_nodes, uint64_t n_unitigs, const uint32_t* unitig_edges6, uint64_t n_unitig_edges,
	const char* out_gfa, const char* out_gaf,
	const uint64_t* var_hash, const uint32_t* var_offset, const uint64_t* var_len_begin, const uint32_t* var_len, uint64_t n_var);

int refshim_backend(int stage, int mode, uint64_t k, uint64_t w, const char* fasta, uint64_t min_cov, double min_unitig_cov, uint64_t max_resolve, uint64_t max_uncond_resolve,
	int guesswork, int copycount_heuristic, int only_local_resolve, int do_cleaning,
	const uint64_t* node_hash, const uint32_t* node_cov, const uint8_t* node_tip, uint64_t n_nodes, const uint32_t* edges6, uint64_t n_edges,
	const uint32_t* new_overlaps6, uint64_t n_new_overlaps, const uint64_t* unitig_begin, const uint32_t* unitig_nodes, uint64_t n_unitigs, const uint32_t* unitig_edges6, uint64_t n_unitig_edges,
	const char* out_gfa, const char* out_gaf)
{
	return refshim_backend_variants(stage, mode, k, w, fasta, min_cov, min_unitig_cov, max_resolve, max_uncond_resolve, guesswork, copycount_heuristic, only_local_resolve, do_cleaning,
		node_hash, node_cov, node_tip, n_nodes, edges6, n_edges, new_overlaps6, n_new_overlaps, unitig_begin, unitig_nodes, n_unitigs, unitig_edges6, n_unitig_edges, out_gfa, out_gaf,
		nullptr, nullptr, nullptr, nullptr, 0);
}

// var_*: the hpc-variant table of pass 2 (what runMBG's iterator holds from MBG.cpp:1564 on); passes 4 and 5 need it too
int refshim_backend_variants(int stage, int mode, uint64_t k, uint64_t w, const char* fasta, uint64_t min_cov, double min_unitig_cov, uint64_t max_resolve, uint64_t max_uncond_resolve,
	int guesswork, int copycount_heuristic, int only_local_resolve, int do_cleaning,
	const uint64_t* node_hash, const uint32_t* node_cov, const uint8_t* node_tip, uint64_t n_nodes, const uint32_t* edges6, uint64_t n_edges,
	const uint32_t* new_overlaps6, uint64_t n_new_overlaps, const uint64_t* unitig_begin, const uint32_t* unitig_nodes, uint64_t n_unitigs, const uint32_t* unitig_edges6, uint64_t n_unitig_edges,
	const char* out_gfa, const char* out_gaf,
	const uint64_t* var_hash, const uint32_t* var_offset, const uint64_t* var_len_begin, const uint32_t* var_len, uint64_t n_var)
{
	std::vector<std::string> files { std::string(fasta) };
	ReadpartIterator partIterator { (size_t)k, (size_t)w, (ErrorMasking)mode, 1, files, false, "" };
	for (uint64_t e = 0; e < n_var; e++)
	{
		const HashType hsh = ((HashType)var_hash[2 * e + 1] << 64) + (HashType)var_hash[2 * e];
		std::vector<size_t> lens { var_len + var_len_begin[e], var_len + var_len_begin[e + 1] };
		partIterator.addHpcVariants(hsh, var_offset[e], lens);
	}
	HashList reads { (size_t)k };
	std::ostringstream log;
	fill_hashlist(reads, node_hash, node_cov, node_tip, n_nodes, edges6, n_edges);
	UnitigGraph unitigs;
	if (stage == 0)
	{
		unitigs = getUnitigGraph(reads, min_cov, min_unitig_cov, false, (min_unitig_cov >= 2) && (max_resolve > 0) && guesswork);      // MBG.cpp:1573
		if (min_unitig_cov > (double)min_cov) unitigs = getUnitigs(unitigs.filterUnitigsByCoverage(min_unitig_cov, false));
		filterKmersToUnitigKmers(unitigs, reads, k, true);
		sortKmersByHashes(unitigs, reads);
	}
	else
	{
		for (uint64_t e = 0; e < n_new_overlaps; e++)      // MBG.cpp:1178-1183
		{
			const uint32_t* r = new_overlaps6 + 6 * e;
			reads.addSequenceOverlap(std::make_pair((size_t)r[0], r[1] != 0), std::make_pair((size_t)r[2], r[3] != 0), r[5]);
		}
		unitigs.unitigs.resize(n_unitigs);
		unitigs.unitigCoverage.resize(n_unitigs);
		unitigs.edges.resize(n_unitigs);
		unitigs.edgeCov.resize(n_unitigs);
		unitigs.edgeOvlp.resize(n_unitigs);
		unitigs.leftClip.resize(n_unitigs, 0);       // as startUnitig leaves them (MBG.cpp:356-357)
		unitigs.rightClip.resize(n_unitigs, 0);
		for (uint64_t u = 0; u < n_unitigs; u++)
			for (uint64_t j = unitig_begin[u]; j < unitig_begin[u + 1]; j++)
			{
				unitigs.unitigs[u].emplace_back(unitig_nodes[j] >> 1, (unitig_nodes[j] & 1) != 0);
				unitigs.unitigCoverage[u].push_back(node_cov[unitig_nodes[j] >> 1]);
			}
		for (uint64_t e = 0; e < n_unitig_edges; e++)      // as the last loop of getUnitigGraph leaves them (MBG.cpp:764-779)
		{
			const uint32_t* r = unitig_edges6 + 6 * e;
			const std::pair<size_t, bool> from { r[0], r[1] != 0 }, to { r[2], r[3] != 0 };
			unitigs.edges.addEdge(from, to);
			unitigs.edges.addEdge(reverse(to), reverse(from));
			unitigs.setEdgeCoverage(from, to, r[4]);
			unitigs.setEdgeOverlap(from, to, 0);
		}
	}
	if (const char* dump = getenv("REFSHIM_DUMP_UNITIGS"))
	{
		FILE* f = fopen(dump, "w");
		for (size_t u = 0; u < unitigs.unitigs.size(); u++)
		{
			fprintf(f, "U %zu n=%zu:", u, unitigs.unitigs[u].size());
			for (auto nd : unitigs.unitigs[u]) fprintf(f, " %zu%c", (size_t)nd.first, nd.second ? '+' : '-');
			fprintf(f, "\n");
			for (int fw = 1; fw >= 0; fw--) for (auto to : unitigs.edges[std::make_pair(u, fw == 1)]) fprintf(f, "  E %zu%c -> %zu%c cov %zu\n", u, fw ? '+' : '-', (size_t)to.first, to.second ? '+' : '-', unitigs.edgeCoverage(std::make_pair(u, fw == 1), to));
		}
		for (size_t i = 0; i < reads.size(); i++)
		{
			fprintf(f, "N %zu cov %zu tip %d\n", i, (size_t)reads.coverage.get(i), reads.isTipKmer(i) ? 1 : 0);
			for (int fw = 1; fw >= 0; fw--) for (auto pr : reads.getSequenceOverlaps(std::make_pair(i, fw == 1))) fprintf(f, "  O %zu%c -> %zu%c ov %zu\n", i, fw ? '+' : '-', (size_t)pr.first.first, pr.first.second ? '+' : '-', (size_t)pr.second);
			for (int fw = 1; fw >= 0; fw--) for (auto pr : reads.getEdgeCoverages(std::make_pair(i, fw == 1))) fprintf(f, "  C %zu%c -> %zu%c cov %zu\n", i, fw ? '+' : '-', (size_t)pr.first.first, pr.first.second ? '+' : '-', (size_t)pr.second);
		}
		fclose(f);
	}
	std::vector<ReadPath> readPaths = getReadPaths(unitigs, reads, 1, partIterator, k);      // MBG.cpp:1586
	if (max_resolve > 0)
		std::tie(unitigs, readPaths) = resolveUnitigs(unitigs, reads, readPaths, partIterator, min_unitig_cov, k, max_resolve, max_uncond_resolve, false, guesswork != 0,
			copycount_heuristic != 0, only_local_resolve != 0, do_cleaning != 0, log);
	std::vector<CompressedSequenceType> unitigSequences;
	StringIndex stringIndex;
	std::tie(unitigSequences, stringIndex) = getHPCUnitigSequences(reads, unitigs, readPaths, k, partIterator, 1);
	verifyEdgeConsistency(unitigs, reads, stringIndex, unitigSequences, k);
	std::vector<double> rawCov = getRawKmerCoverages(unitigs, unitigSequences, reads, k);
	writeGraph(unitigs, out_gfa, reads, unitigSequences, stringIndex, k, rawCov, "");
	if (out_gaf && out_gaf[0])
	{
		std::vector<DumbSelect> expandedPoses = getUnitigExpandedPoses(reads, unitigs, unitigSequences, stringIndex);
		sortPaths(readPaths);
		writePaths(reads, unitigs, unitigSequences, stringIndex, expandedPoses, readPaths, k, out_gaf, "");
	}
	return 0;
}

// Known-answer helpers (SURVEY.md 8c "tables", "FastHasher", "hash(vector)", "msat" rows).
uint64_t refshim_max_code() { return maxCode(); }
uint16_t refshim_complement(uint16_t c) { return complement(c); }
uint64_t refshim_code_motif_length(uint16_t c) { return codeMotifLength(c); }
uint64_t refshim_char_hash(uint64_t c) { return getHash(c); }
void refshim_fasthasher(uint64_t s, const uint16_t* codes, uint64_t n, uint64_t* fw, uint64_t* bw)
{
	FastHasher hasher { (size_t)s };
	for (uint64_t i = 0; i < n; i++)
	{
		hasher.addChar(codes[i]);
		if (i >= s) hasher.removeChar(codes[i-s]);
	}
	*fw = hasher.getFwHash();
	*bw = hasher.getBwHash();
}
void refshim_kmer_hash(const uint16_t* codes, uint64_t n, uint64_t* lo, uint64_t* hi)
{
	std::vector<uint16_t> v { codes, codes + n };
	HashType h = hash(v);
	*lo = (uint64_t)h;
	*hi = (uint64_t)(h >> 64);
}
// multiRLECompress on an arbitrary symbol string (values 0..3), poses supplied by the caller (n+1 entries).
uint64_t refshim_msat(const uint16_t* sym, const uint64_t* poses, uint64_t n, uint64_t max_motif, uint16_t* out_codes, uint64_t* out_poses)
{
	SequenceCharType s { sym, sym + n };
	SequenceLengthType p { poses, poses + n + 1 };
	auto result = multiRLECompress(s, p, max_motif);
	for (size_t i = 0; i < result.first.size(); i++) out_codes[i] = result.first[i];
	for (size_t i = 0; i < result.second.size(); i++) out_poses[i] = result.second[i];
	return result.first.size();
}

}
