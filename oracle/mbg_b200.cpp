// oracle/mbg_b200.cpp -- INTEGRATION BUILD / TEST INFRASTRUCTURE, not product code.
//
// The outer (process) boundary of SURVEY.md 8(b): an MBG-compatible executable.  What a Verkko maintainer would do to MBG is
// replace the body of runMBG (MBG/MBG.cpp:1538-1658) by calls into libmbgfe; this file IS that replacement, written for this repo
// and compiled by oracle/Makefile against the UNMODIFIED reference sources where they lie: MBG's own main.cpp (option parsing,
// validation, parameter echo: MBG/main.cpp:6-222) and every other object stay as they are, only the symbol `runMBG` of MBG.o is
// renamed to `runMBG_reference` (objcopy) so that the definition below takes its place.  Result: oracle/_ref/MBG_b200, same argv,
// same GFA / GAF files, same stderr lines as MBG -- with passes 1 and 3 (FASTA parsing, compression, syncmer selection, k-mer
// hashing, HashList, sortByHash), getUnitigGraph + filterKmersToUnitigKmers (pass 3) and getReadPaths (pass 4) on the B200.
// Still the reference's own code: pass 2 (consensus votes + hpc-variant discovery, row f3), resolveUnitigs, pass 5
// (getHPCUnitigSequences) and the writers.  Options the library does not cover (--blunt, --keep-gaps, --include-end-kmers,
// --sequence-cache-file, --no-kmer-filter-inside-unitig) go to runMBG_reference unchanged.
// The product (verkko_b200/, include/) never links or loads this; it links the product.
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <algorithm>
#include <vector>
#include <phmap.h>
#include <zstr.hpp>
// the hpc-variant table of ReadpartIterator (ReadHelper.h:250) is private and has no getter: pass 2 stays the reference's, its
// result has to reach mbgfe_set_variants
#define private public
#include "ReadHelper.h"
#undef private
#include "HashList.h"
#include "UnitigGraph.h"
#include "UnitigResolver.h"
#include "HPCConsensus.h"
#include "StringIndex.h"
#include "DumbSelect.h"
#include "CompressedSequence.h"
#include "KmerMatcher.h"
#include "MBGCommon.h"
#include "../include/mbgfe.h"

struct AssemblyStats { size_t nodes, edges, size, N50, approxKmers; };     // MBG.cpp:34-42 (defined in the .cpp only)
void runMBG_reference(const std::vector<std::string>& inputReads, const std::string& outputGraph, const size_t kmerSize, const size_t windowSize, const size_t minCoverage, const double minUnitigCoverage, const ErrorMasking errorMasking, const size_t numThreads, const bool includeEndKmers, const std::string& outputSequencePaths, const size_t maxResolveLength, const bool blunt, const size_t maxUnconditionalResolveLength, const std::string& nodeNamePrefix, const std::string& sequenceCacheFile, const bool keepGaps, const double hpcVariantOnecopyCoverage, const bool guesswork, const bool copycountFilterHeuristic, const bool onlyLocalResolve, const std::string& outputHomologyMap, const bool filterWithinUnitig, const bool doCleaning);
UnitigGraph getUnitigGraph(HashList& hashlist, const size_t minCoverage, const double minUnitigCoverage, const bool keepGaps, const bool oneCovHeuristic);
UnitigGraph getUnitigs(const UnitigGraph& oldgraph);
void sortKmersByHashes(UnitigGraph& unitigs, HashList& reads);
std::vector<ReadPath> getReadPaths(const UnitigGraph& graph, const HashList& hashlist, const size_t numThreads, const ReadpartIterator& partIterator, const size_t kmerSize);
void verifyEdgeConsistency(const UnitigGraph& unitigs, const HashList& hashlist, const StringIndex& stringIndex, const std::vector<CompressedSequenceType>& unitigSequences, const size_t kmerSize);
std::vector<double> getRawKmerCoverages(const UnitigGraph& unitigs, const std::vector<CompressedSequenceType>& unitigSequences, const HashList& reads, const size_t kmerSize);
AssemblyStats writeGraph(const UnitigGraph& unitigs, const std::string& filename, const HashList& hashlist, const std::vector<CompressedSequenceType>& unitigSequences, const StringIndex& stringIndex, const size_t kmerSize, const std::vector<double>& unitigRawKmerCoverages, const std::string& nodeNamePrefix);
std::vector<DumbSelect> getUnitigExpandedPoses(const HashList& hashlist, const UnitigGraph& unitigs, const std::vector<CompressedSequenceType>& unitigSequences, const StringIndex& stringIndex);
void sortPaths(std::vector<ReadPath>& readPaths);
void writePaths(const HashList& hashlist, const UnitigGraph& unitigs, const std::vector<CompressedSequenceType>& unitigSequences, const StringIndex& stringIndex, const std::vector<DumbSelect>& unitigExpandedPoses, const std::vector<ReadPath>& readPaths, const size_t kmerSize, const std::string& outputSequencePaths, const std::string& nodeNamePrefix);
void writeHomologyMap(const HashList& hashlist, const UnitigGraph& unitigs, const size_t kmerSize, const std::vector<CompressedSequenceType>& unitigSequences, const StringIndex& stringIndex, const std::vector<DumbSelect>& unitigExpandedPoses, const std::string& outputHomologyMap);
void printUnitigKmerCount(const UnitigGraph& unitigs);

namespace {

auto now() { return std::chrono::steady_clock::now(); }
std::string took(std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b)
{
	// formatTime (MBG/MBG.cpp:1000-1004): "seconds,milliseconds s"
	const size_t ms = std::chrono::duration_cast<std::chrono::milliseconds>(b - a).count();
	return std::to_string(ms / 1000) + "," + std::to_string(ms % 1000) + " s";
}

void die(mbgfe* fe, int rc)
{
	std::cerr << "libmbgfe: " << mbgfe_last_error(fe) << std::endl;
	// MBG aborts on non-droppable palindromes / unhashable k-mers (ReadHelper.h:775-783) and exits 1 on unreadable input (MBG.cpp:1542-1549)
	if (rc == MBGFE_E_IO) std::exit(1);
	std::abort();
}

// HashList <- node + edge arrays in ascending canonical-hash order, through its public members (MBG/HashList.h:15-48)
void fill_hashlist(HashList& reads, const mbgfe_hash128* hashes, const uint32_t* cov, const uint8_t* tip, uint64_t n, const mbgfe_edge* e, uint64_t m)
{
	for (uint64_t i = 0; i < n; i++)
	{
		const HashType hsh = ((HashType)hashes[i].hi << 64) + (HashType)hashes[i].lo;
		const std::pair<size_t, bool> node = reads.addNode(hsh);
		if (node.first != i || !node.second) std::abort();
		reads.coverage.set(i, cov[i]);
		if (tip[i]) reads.setTipKmer(i);
	}
	for (uint64_t i = 0; i < m; i++)
	{
		const std::pair<size_t, bool> from { e[i].from, e[i].from_fw != 0 }, to { e[i].to, e[i].to_fw != 0 };
		reads.addSequenceOverlap(from, to, e[i].overlap);
		reads.setEdgeCoverage(from, to, e[i].coverage);
	}
}

// one pass of loadReadsAsHashesMultithread (MBG.cpp:44-76) on the device, with its two stderr lines
void load_reads(mbgfe* fe, const std::vector<std::string>& files)
{
	mbgfe_reset(fe);
	for (const std::string& f : files) if (int rc = mbgfe_add_file(fe, f.c_str())) die(fe, rc);
	if (int rc = mbgfe_run(fe)) die(fe, rc);
	mbgfe_stats st;
	mbgfe_get_stats(fe, &st);
	std::cerr << st.selected << " total selected k-mers in reads" << std::endl;
	std::cerr << st.distinct << " distinct selected k-mers in reads" << std::endl;
}

}

void runMBG(const std::vector<std::string>& inputReads, const std::string& outputGraph, const size_t kmerSize, const size_t windowSize, const size_t minCoverage, const double minUnitigCoverage, const ErrorMasking errorMasking, const size_t numThreads, const bool includeEndKmers, const std::string& outputSequencePaths, const size_t maxResolveLength, const bool blunt, const size_t maxUnconditionalResolveLength, const std::string& nodeNamePrefix, const std::string& sequenceCacheFile, const bool keepGaps, const double hpcVariantOnecopyCoverage, const bool guesswork, const bool copycountFilterHeuristic, const bool onlyLocalResolve, const std::string& outputHomologyMap, const bool filterWithinUnitig, const bool doCleaning)
{
	if (blunt || keepGaps || includeEndKmers || sequenceCacheFile != "" || !filterWithinUnitig || getenv("MBG_B200_OFF"))
	{
		std::cerr << "MBG_b200: option outside the B200 front end, running the reference path" << std::endl;
		runMBG_reference(inputReads, outputGraph, kmerSize, windowSize, minCoverage, minUnitigCoverage, errorMasking, numThreads, includeEndKmers, outputSequencePaths, maxResolveLength, blunt,
			maxUnconditionalResolveLength, nodeNamePrefix, sequenceCacheFile, keepGaps, hpcVariantOnecopyCoverage, guesswork, copycountFilterHeuristic, onlyLocalResolve, outputHomologyMap, filterWithinUnitig, doCleaning);
		return;
	}
	auto beforeReading = now();
	for (const std::string& name : inputReads)      // MBG.cpp:1541-1549
	{
		std::ifstream file { name };
		if (!file.good())
		{
			std::cerr << "Input file " << name << " can't be read!" << std::endl;
			std::exit(1);
		}
	}
	ReadpartIterator partIterator { kmerSize, windowSize, errorMasking, numThreads, inputReads, includeEndKmers, sequenceCacheFile };   // passes 2 and 5 (reference code)
	HashList reads { kmerSize };
	mbgfe_params prm;
	memset(&prm, 0, sizeof(prm));
	prm.abi_version = MBGFE_ABI_VERSION; prm.k = (uint32_t)kmerSize; prm.w = (uint32_t)windowSize; prm.error_masking = (uint32_t)errorMasking;
	prm.kmer_abundance = (uint32_t)minCoverage; prm.device = getenv("MBG_B200_DEVICE") ? atoi(getenv("MBG_B200_DEVICE")) : 0; prm.rank = 0; prm.world = 1;
	mbgfe* fe = nullptr;
	if (int rc = mbgfe_create(&prm, &fe)) { std::cerr << "libmbgfe: " << mbgfe_last_error(nullptr) << std::endl; (void)rc; std::abort(); }
	auto beforeVariants = now();
	if (hpcVariantOnecopyCoverage != 0)
	{
		std::cerr << "Collecting hpc variant k-mers" << std::endl;
		load_reads(fe, inputReads);                                               // pass 1 on the device
		fill_hashlist(reads, (const mbgfe_hash128*)mbgfe_data(fe, MBGFE_NODE_HASHES), (const uint32_t*)mbgfe_data(fe, MBGFE_NODE_COVERAGE), (const uint8_t*)mbgfe_data(fe, MBGFE_NODE_TIP),
			mbgfe_count(fe, MBGFE_NODE_HASHES), (const mbgfe_edge*)mbgfe_data(fe, MBGFE_EDGES), mbgfe_count(fe, MBGFE_EDGES));
		auto unitigs = getUnitigGraph(reads, minCoverage, minUnitigCoverage, keepGaps, false);
		if (minUnitigCoverage > minCoverage) unitigs = getUnitigs(unitigs.filterUnitigsByCoverage(minUnitigCoverage, keepGaps));
		sortKmersByHashes(unitigs, reads);
		getHpcVariantsAndReadPaths(reads, unitigs, kmerSize, partIterator, numThreads, hpcVariantOnecopyCoverage * 1.5, hpcVariantOnecopyCoverage * 0.5);   // pass 2: row f3, reference code
		std::vector<mbgfe_hash128> vh; std::vector<uint32_t> vo, vl; std::vector<uint64_t> vb { 0 };
		for (const auto& pair : partIterator.hpcVariants)
			for (const auto& entry : pair.second)
			{
				vh.push_back({ (uint64_t)pair.first, (uint64_t)(pair.first >> 64) });
				vo.push_back((uint32_t)entry.first);
				for (size_t len : entry.second) vl.push_back((uint32_t)len);
				vb.push_back(vl.size());
			}
		if (int rc = mbgfe_set_variants(fe, vh.data(), vo.data(), vb.data(), vl.data(), vo.size())) die(fe, rc);
		reads.clear();
		partIterator.clearCacheHashes();
	}
	auto beforeKmers = now();
	std::cerr << "Collecting selected k-mers" << std::endl;
	load_reads(fe, inputReads);                                                   // pass 3 on the device
	auto beforeUnitigs = now();
	std::cerr << "Unitigifying" << std::endl;
	const bool oneCov = (minUnitigCoverage >= 2) && (maxResolveLength > 0) && guesswork;      // MBG.cpp:1573
	if (int rc = mbgfe_build_unitigs_ex(fe, (uint32_t)minCoverage, minUnitigCoverage, oneCov ? MBGFE_UNITIG_ONECOV_HEURISTIC : 0)) die(fe, rc);
	auto beforeFilter = now();
	if (minUnitigCoverage > minCoverage) std::cerr << "Filtering by unitig coverage" << std::endl;      // nothing left to remove after the chain filter
	if (int rc = mbgfe_thin_unitigs(fe)) die(fe, rc);                              // filterKmersToUnitigKmers, applied
	UnitigGraph unitigs;
	{
		const uint64_t n = mbgfe_count(fe, MBGFE_T_NODE_HASHES);
		const uint32_t* cov = (const uint32_t*)mbgfe_data(fe, MBGFE_T_NODE_COVERAGE);
		fill_hashlist(reads, (const mbgfe_hash128*)mbgfe_data(fe, MBGFE_T_NODE_HASHES), cov, (const uint8_t*)mbgfe_data(fe, MBGFE_T_NODE_TIP), n,
			(const mbgfe_edge*)mbgfe_data(fe, MBGFE_T_EDGES), mbgfe_count(fe, MBGFE_T_EDGES));
		const mbgfe_edge* no = (const mbgfe_edge*)mbgfe_data(fe, MBGFE_T_NEW_OVERLAPS);
		for (uint64_t i = 0, m = mbgfe_count(fe, MBGFE_T_NEW_OVERLAPS); i < m; i++)      // MBG.cpp:1178-1183
			reads.addSequenceOverlap(std::make_pair((size_t)no[i].from, no[i].from_fw != 0), std::make_pair((size_t)no[i].to, no[i].to_fw != 0), no[i].overlap);
		const uint64_t nu = mbgfe_count(fe, MBGFE_T_UNITIG_BEGIN) - 1;
		const uint64_t* ub = (const uint64_t*)mbgfe_data(fe, MBGFE_T_UNITIG_BEGIN);
		const uint32_t* un = (const uint32_t*)mbgfe_data(fe, MBGFE_T_UNITIG_NODES);
		unitigs.unitigs.resize(nu); unitigs.unitigCoverage.resize(nu); unitigs.edges.resize(nu); unitigs.edgeCov.resize(nu); unitigs.edgeOvlp.resize(nu);
		unitigs.leftClip.resize(nu, 0); unitigs.rightClip.resize(nu, 0);
		for (uint64_t u = 0; u < nu; u++)
			for (uint64_t j = ub[u]; j < ub[u + 1]; j++)
			{
				unitigs.unitigs[u].emplace_back(un[j] >> 1, (un[j] & 1) != 0);
				unitigs.unitigCoverage[u].push_back(cov[un[j] >> 1]);
			}
		const mbgfe_edge* ue = (const mbgfe_edge*)mbgfe_data(fe, MBGFE_UNITIG_EDGES);
		for (uint64_t i = 0, m = mbgfe_count(fe, MBGFE_UNITIG_EDGES); i < m; i++)      // as the last loop of getUnitigGraph leaves them (MBG.cpp:764-779)
		{
			const std::pair<size_t, bool> from { ue[i].from, ue[i].from_fw != 0 }, to { ue[i].to, ue[i].to_fw != 0 };
			unitigs.edges.addEdge(from, to);
			unitigs.edges.addEdge(reverse(to), reverse(from));
			unitigs.setEdgeCoverage(from, to, ue[i].coverage);
			unitigs.setEdgeOverlap(from, to, 0);
		}
	}
	printUnitigKmerCount(unitigs);
	auto beforePaths = now();
	std::vector<ReadPath> readPaths;
	std::cerr << "Getting read paths" << std::endl;
	if (getenv("MBG_B200_REFERENCE_PATHS")) readPaths = getReadPaths(unitigs, reads, numThreads, partIterator, kmerSize);
	else
	{
		// pass 4 on the device: mbgfe_thread_reads on the thinned list; a ReadPath is a run 2|3, (4|5)* of MBGFE_INST_STEP (include/mbgfe.h)
		if (int rc = mbgfe_thread_reads(fe)) die(fe, rc);
		const mbgfe_part* parts = (const mbgfe_part*)mbgfe_data(fe, MBGFE_PARTS);
		const uint64_t n_parts = mbgfe_count(fe, MBGFE_PARTS);
		const uint32_t* positions = (const uint32_t*)mbgfe_data(fe, MBGFE_POSITIONS);
		const uint32_t* iu = (const uint32_t*)mbgfe_data(fe, MBGFE_INST_UNITIG);
		const uint32_t* ipos = (const uint32_t*)mbgfe_data(fe, MBGFE_INST_UNITIG_POS);
		const uint32_t* irest = (const uint32_t*)mbgfe_data(fe, MBGFE_INST_UNITIG_REST);
		const uint8_t* step = (const uint8_t*)mbgfe_data(fe, MBGFE_INST_STEP);
		for (uint64_t p = 0; p < n_parts; p++)
		{
			if (!parts[p].hashed) continue;
			ReadPath cur;
			auto fresh = [&]()
			{
				cur.path.clear(); cur.leftClip = 0; cur.rightClip = 0;
				cur.readName = ReadName { std::string(mbgfe_read_name(fe, parts[p].read_idx)), parts[p].part_offset };
				cur.readLength = parts[p].raw_len; cur.readLengthHPC = parts[p].n_codes;
			};
			fresh();
			for (uint64_t i = parts[p].sel_begin; i < parts[p].sel_begin + parts[p].n_sel; i++)
			{
				const Node node { (size_t)(iu[i] >> 1), (iu[i] & 1) != 0 };
				switch (step[i])
				{
					case 0: break;
					case 1: if (cur.path.size() > 0) readPaths.push_back(cur); fresh(); break;      // KmerMatcher.h:30-47: readPoses stays
					case 2: cur.path.push_back(node); cur.readPoses.push_back(positions[i]); cur.leftClip = ipos[i]; cur.rightClip = irest[i]; break;
					case 3: readPaths.push_back(cur); fresh(); cur.readPoses.clear(); cur.path.push_back(node); cur.readPoses.push_back(positions[i]); cur.leftClip = ipos[i]; cur.rightClip = irest[i]; break;
					case 4: cur.readPoses.push_back(positions[i]); cur.rightClip = irest[i]; break;
					default: cur.path.push_back(node); cur.readPoses.push_back(positions[i]); cur.rightClip = irest[i]; break;
				}
			}
			if (cur.path.size() > 0) readPaths.push_back(cur);
		}
	}
	mbgfe_destroy(fe);
	auto beforeResolve = now();
	if (maxResolveLength > 0)
	{
		std::cerr << "Resolving unitigs" << std::endl;
		std::cerr << unitigs.unitigs.size() << " unitigs before resolving" << std::endl;
		std::tie(unitigs, readPaths) = resolveUnitigs(unitigs, reads, readPaths, partIterator, minUnitigCoverage, kmerSize, maxResolveLength, maxUnconditionalResolveLength, keepGaps, guesswork, copycountFilterHeuristic, onlyLocalResolve, doCleaning, std::cerr);
		std::cerr << unitigs.unitigs.size() << " unitigs after resolving" << std::endl;
	}
	auto beforeSequences = now();
	std::cerr << "Building unitig sequences" << std::endl;
	std::vector<CompressedSequenceType> unitigSequences;
	StringIndex stringIndex;
	std::tie(unitigSequences, stringIndex) = getHPCUnitigSequences(reads, unitigs, readPaths, kmerSize, partIterator, numThreads);      // pass 5, reference code
	auto beforeConsistency = now();
	verifyEdgeConsistency(unitigs, reads, stringIndex, unitigSequences, kmerSize);
	auto beforeWrite = now();
	std::vector<double> unitigRawKmerCoverages = getRawKmerCoverages(unitigs, unitigSequences, reads, kmerSize);
	std::cerr << "Writing graph to " << outputGraph << std::endl;
	AssemblyStats stats = writeGraph(unitigs, outputGraph, reads, unitigSequences, stringIndex, kmerSize, unitigRawKmerCoverages, nodeNamePrefix);
	auto afterWrite = now();
	std::vector<DumbSelect> unitigExpandedPoses;
	if (outputSequencePaths != "" || outputHomologyMap != "") unitigExpandedPoses = getUnitigExpandedPoses(reads, unitigs, unitigSequences, stringIndex);
	if (outputSequencePaths != "")
	{
		std::cerr << "Writing paths to " << outputSequencePaths << std::endl;
		sortPaths(readPaths);
		writePaths(reads, unitigs, unitigSequences, stringIndex, unitigExpandedPoses, readPaths, kmerSize, outputSequencePaths, nodeNamePrefix);
	}
	auto afterPaths = now();
	if (outputHomologyMap != "")
	{
		std::cerr << "Writing homology map to " << outputHomologyMap << std::endl;
		writeHomologyMap(reads, unitigs, kmerSize, unitigSequences, stringIndex, unitigExpandedPoses, outputHomologyMap);
	}
	auto afterHomologyMap = now();
	(void)beforeReading;
	if (hpcVariantOnecopyCoverage != 0) std::cerr << "selecting hpc variant k-mers took " << took(beforeVariants, beforeKmers) << std::endl;
	std::cerr << "selecting k-mers and building graph topology took " << took(beforeKmers, beforeUnitigs) << std::endl;
	std::cerr << "unitigifying took " << took(beforeUnitigs, beforeFilter) << std::endl;
	std::cerr << "filtering unitigs took " << took(beforeFilter, beforePaths) << std::endl;
	std::cerr << "getting read paths took " << took(beforePaths, beforeResolve) << std::endl;
	if (maxResolveLength > 0) std::cerr << "resolving unitigs took " << took(beforeResolve, beforeSequences) << std::endl;
	std::cerr << "building unitig sequences took " << took(beforeSequences, beforeConsistency) << std::endl;
	if (errorMasking != ErrorMasking::No && errorMasking != ErrorMasking::Collapse) std::cerr << "forcing edge consistency took " << took(beforeConsistency, beforeWrite) << std::endl;
	std::cerr << "writing the graph and calculating stats took " << took(beforeWrite, afterWrite) << std::endl;
	if (outputSequencePaths != "") std::cerr << "writing sequence paths took " << took(afterWrite, afterPaths) << std::endl;
	if (outputHomologyMap != "") std::cerr << "writing homology map took " << took(afterPaths, afterHomologyMap) << std::endl;
	std::cerr << "nodes: " << stats.nodes << std::endl;
	std::cerr << "edges: " << stats.edges << std::endl;
	std::cerr << "assembly size " << stats.size << " bp, N50 " << stats.N50 << std::endl;
	if (stats.approxKmers > 0) std::cerr << "approximate number of k-mers ~ " << stats.approxKmers << std::endl;
}
