// mbgfe.cu -- host side of libmbgfe: C ABI (include/mbgfe.h), batch upload, kernel sequencing, result export.
//
// Reference being replaced (MBG/ = /root/reference/src/hifioverlapper/MBG/src, commit 7bef84dd):
//   ReadpartIterator::iterateHashes / iterateParts (MBG/ReadHelper.h:167-235, 792-841) and the loop body of
//   loadReadsAsHashesMultithread (MBG/MBG.cpp:44-76) with its HashList (MBG/HashList.cpp).
// The compute path is CUDA only (kernels.cuh).  There is no host implementation of any stage in this file: the host
// parses files, builds the (tiny) segment table, sequences kernels and orders the exported table.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <fcntl.h>
#include <functional>
#include <mutex>
#include <string>
#include <sys/stat.h>
#include <unistd.h>
#include <vector>
#include <zlib.h>

#include "kernels.cuh"
#include "graph.cuh"
#include "hostpack.h"
#include "../../include/mbgfe.h"
#ifndef MBGFE_EMU_BUILD
#include <dlfcn.h>
#include <nccl.h>
#endif

static_assert(sizeof(mbgfe_dev::EdgeRec) == sizeof(mbgfe_edge), "device edge record is the ABI edge record");

#ifdef MBGFE_EMU_BUILD
thread_local uint3 threadIdx;
thread_local uint3 blockIdx;
thread_local dim3 blockDim;
thread_local dim3 gridDim;
#endif

using namespace mbgfe_dev;

namespace {

thread_local std::string g_create_error;

struct DevBuf
{
	void* p = nullptr;
	size_t cap = 0;
	template <typename T> T* as() const { return (T*)p; }
};

struct HostBuf
{
	void* p = nullptr;
	size_t cap = 0;
	template <typename T> T* as() const { return (T*)p; }
};

struct HostChunk { const char* bases; const uint64_t* offsets; uint64_t n_reads; };

struct Timer
{
	cudaEvent_t a = nullptr, b = nullptr;
};

}


struct mbgfe
{
	mbgfe_params prm;
	std::string err;
	bool collapse = false, msat = false;
	uint32_t max_motif = 6;
	cudaStream_t stream = nullptr, copy_stream = nullptr;
	cudaStream_t front_stream[2] = { nullptr, nullptr };   // mbgfe_run: K0..K3a of consecutive pieces alternate between these
	std::vector<cudaEvent_t> front_events, pack_events;   // per piece: K3a done / K0 done
	std::vector<cudaEvent_t> piece_events;            // one per H2D piece of the pipelined upload
	int sm_count = 148;
	size_t max_smem = 0;
	SelectPlan plan;

	// code tables
	DevBuf tab_hash, tab_pair, tab_comp, tab_pext8, tab_in, tab_out;
	DevBuf d_sketch;                                                // distinct-count bitmaps (size_tables)

	// pending input
	std::vector<HostChunk> pending;
	std::vector<std::string> names;
	std::vector<std::vector<char>*> owned_bases;       // file-loaded reads
	std::vector<std::vector<uint64_t>*> owned_offsets;

	// resident batch (host metadata)
	bool resident = false, processed = false, exported_parts = false, exported_table = false, exported_codes = false, have_table = false;
	uint64_t n_reads = 0, n_bases = 0;
	std::vector<uint64_t> read_off;                   // n_reads + 1
	std::vector<uint64_t> read_end;                   // empty: read r ends where read r + 1 begins; else n_reads ends (file images: headers sit in between)
	// plain FASTA files waiting for mbgfe_run's streaming loader (row a1); base = batch position of the file's first byte
	struct DeferredFile { std::string path; int fd; uint64_t size, base; };
	std::vector<DeferredFile> files;
	std::vector<uint64_t> name_off; std::vector<uint32_t> name_len, name_file;   // names of file-image reads: fetched on demand
	std::vector<DeferredFile> image_files;            // the files the resident image was made from (kept open for names / rare re-reads)
	std::vector<uint64_t> seg_in, seg_out;            // per segment
	std::vector<uint32_t> seg_len, seg_read, seg_rawoff, seg_nonacgt_before;
	uint64_t total_out = 0;
	uint64_t nonacgt_count = 0;

	// device buffers
	DevBuf d_ascii, d_packed, d_nonacgt, d_counters;
	DevBuf d_seg_in, d_seg_out, d_seg_len, d_seg_c1, d_seg_c2, d_seg_nsel, d_sel_begin;
	DevBuf d_sym, d_codes, d_selmask, d_pos1, d_pos2;
	DevBuf d_inst_pos, d_inst_seg, d_inst_flag, d_inst_hash, d_inst_slot, d_pal_list;
	DevBuf d_nodes, d_edges, d_node_out, d_edge_out, d_records, d_recv;
	DevBuf d_inst_packpos, d_ids_out, d_id_begin, d_edge_records, d_xbuf[4];   // sharded table: id replies and edge records
	DevBuf d_xrow, d_xmatrix, d_rec_slot;                                    // exchange: this rank's counts, everybody's, slot of every received record
	int comm_kind = 0;                                                       // 0 none, 1 NCCL, 2 group (one process), 3 caller's callbacks
	void* nccl_comm = nullptr;
	mbgfe_alltoallv_fn cb_alltoallv = nullptr; mbgfe_allgather_fn cb_allgather = nullptr; void* cb_ctx = nullptr;
	std::vector<uint64_t> id_begin;                                          // first global node id of every rank (world + 1)
	const void* x_packed_ids = nullptr;                                      // ids of this rank's packed records, as received
	uint64_t x_inst_base = 0;
	std::vector<uint64_t> x_counts;                                          // per-owner record counts of the last mbgfe_exchange_counts
	uint32_t id_base = 0;                                                    // nodes owned by lower ranks
	bool nodes_ranked = false, edges_global = false;
	bool table_sharded = false;                                              // the node table holds this rank's owner range only (built by mbgfe_exchange_insert)
	DevBuf d_var_hash, d_var_offset, d_var_len_begin, d_var_len;       // hpc variants (mbgfe_set_variants), sorted
	// mbgfe_build_unitigs: graph scratch (degrees, successor, link, two ranking states, flags, scans) and results
	DevBuf d_g_deg, d_g_nb, d_g_link, d_g_st[2], d_g_kept, d_g_newid, d_g_flag, d_g_at, d_g_flag2, d_g_at2, d_g_flag_byte;
	DevBuf d_f_hash, d_f_cov, d_f_tip, d_f_edges, d_f_orig, d_t_hash, d_t_cov, d_t_tip, d_t_edges, d_t_orig, d_g_kept8, d_u_pos, d_th_final, d_th_node, d_th_unitig, d_th_upos, d_th_urest, d_th_step, d_thin, d_thin_pos, d_u_begin, d_u_nodes, d_u_of_node, d_u_fw, d_u_edges;
	bool have_unitigs = false;
	uint64_t n_kept = 0, n_fedges = 0, n_unitigs = 0, n_unitig_nodes = 0, n_uedges = 0;
	std::vector<mbgfe_edge> h_u_edges;
	std::vector<uint8_t> h_kept, h_f_tip, h_th_step;
	std::vector<uint32_t> h_th_unitig, h_th_upos, h_th_urest, h_thin_pos;
	std::vector<uint8_t> h_thin;
	bool have_thin = false;
	// the list after reads.filter(kept) of filterKmersToUnitigKmers, its unitigs and the new sequence overlaps
	DevBuf d_thin_scan, d_tpos, d_tu_nodes, d_tu_off, d_tu_unitig, d_tu_begin, d_t_newov;
	uint64_t n_thin = 0, n_tedges = 0, n_tpos = 0;
	std::vector<mbgfe_hash128> h_t_hash; std::vector<uint32_t> h_t_cov, h_tu_nodes; std::vector<uint8_t> h_t_tip; std::vector<mbgfe_edge> h_t_edges, h_t_newov; std::vector<uint64_t> h_tu_begin;
	uint32_t unitig_min_cov = 1;
	bool have_threads = false;
	std::vector<mbgfe_hash128> h_f_hash;
	std::vector<uint32_t> h_f_cov, h_u_nodes, h_u_of_node;
	std::vector<mbgfe_edge> h_f_edges;
	std::vector<uint64_t> h_u_begin;
	uint64_t n_var = 0;
	DevBuf d_bucket_count, d_bucket_begin, d_scan_tot, d_scan_off;      // export: bucket tables and scan scratch
	DevBuf d_node_hash_sorted, d_node_cov, d_node_tip, d_edges_sorted;    // export: final layout, device resident
	uint64_t node_slots = 0, edge_slots = 0;
	uint64_t n_inst = 0;
	uint64_t inst_cap = 0;                            // capacity of the instance lists (entries)
	// mbgfe_run_append: the selected k-mers of earlier batches stay in the instance lists, this batch's follow them
	bool appending = false;
	uint64_t append_base = 0;                         // instances of earlier batches
	DevBuf d_append_base, d_range;

	// host results
	std::vector<mbgfe_part> parts;
	std::vector<uint16_t> h_codes;
	std::vector<uint32_t> h_poses, h_positions;
	std::vector<mbgfe_hash128> h_hashes;
	HostBuf p_stage;                                                  // pinned staging of the host-packed pieces
	HostBuf p_node_hashes, p_node_cov, p_node_tip, p_edges;           // pinned: the table arrives by DMA at full rate
	uint64_t n_nodes_out = 0, n_edges_out = 0;
	std::vector<uint32_t> h_seg_c1, h_seg_c2, h_seg_nsel;
	std::vector<uint64_t> h_sel_begin;

	mbgfe_stats stats;
	cudaEvent_t ev[16];
};

namespace {

int fail(mbgfe* h, int code, const char* fmt, ...)
{
	char buf[512];
	va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
	if (h) h->err = buf; else g_create_error = buf;
	return code;
}

#define CU(h, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(h, MBGFE_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

int ensure(mbgfe* h, DevBuf& b, size_t bytes)
{
	bytes = (bytes + 255) & ~(size_t)255;
	if (b.cap >= bytes) return 0;
	if (b.p) CU(h, cudaFree(b.p));
	b.p = nullptr; b.cap = 0;
	CU(h, cudaMalloc(&b.p, bytes));
	b.cap = bytes;
	return 0;
}

void release(DevBuf& b) { if (b.p) cudaFree(b.p); b.p = nullptr; b.cap = 0; }

int ensure_host(mbgfe* h, HostBuf& b, size_t bytes)
{
	bytes = (bytes + 4095) & ~(size_t)4095;
	if (b.cap >= bytes) return 0;
	if (b.p) CU(h, cudaFreeHost(b.p));
	b.p = nullptr; b.cap = 0;
	CU(h, cudaMallocHost(&b.p, bytes));
	b.cap = bytes;
	return 0;
}

void release_host(HostBuf& b) { if (b.p) cudaFreeHost(b.p); b.p = nullptr; b.cap = 0; }

// poses are kept on request, and whenever hpc variants are registered (the re-hash reads run lengths from them)
bool want_poses_of(const mbgfe* h) { return (h->prm.flags & MBGFE_KEEP_POSES) != 0 || h->n_var != 0; }

uint64_t read_end_of(const mbgfe* h, uint64_t r) { return h->read_end.empty() ? h->read_off[r + 1] : h->read_end[r]; }

uint64_t next_pow2(uint64_t v) { uint64_t p = 1; while (p < v) p <<= 1; return p; }

enum { CTR_NONACGT = 0, CTR_PALINDROME = 1, CTR_DISTINCT = 2, CTR_EDGES = 3, CTR_CURSOR = 4 /* and 5 */, CTR_OVERFLOW = 6, CTR_OWNER = 8 /* .. 8+world */ };
constexpr uint64_t PAL_CAP = 1 << 20;

#ifndef MBGFE_NT_SELECT
#define MBGFE_NT_SELECT 64
#endif
constexpr int NT_MSAT = 128, NT_SELECT = MBGFE_NT_SELECT, NT_LIST = 128;   // k_select: 64 / 128 / 256 threads = 5.75 / 6.03 / 7.24 ms on the bench

SegTable seg_table(const mbgfe* h)
{
	SegTable t;
	t.in_start = h->d_seg_in.as<uint64_t>();
	t.len = h->d_seg_len.as<uint32_t>();
	t.out_base = h->d_seg_out.as<uint64_t>();
	t.n = (uint32_t)h->seg_len.size();
	return t;
}

// Segments of the resident reads: reads shorter than 32 bases are ignored (MBG/ReadHelper.h:838); the others are cut
// at every character outside ACGTacgt (positions found by k_pack_ascii, sorted ascending in `bad`).
void build_segments(mbgfe* h, const std::vector<uint64_t>& bad, const std::function<const char*(uint64_t)>& read_ptr)
{
	h->seg_in.clear(); h->seg_out.clear(); h->seg_len.clear(); h->seg_read.clear(); h->seg_rawoff.clear(); h->seg_nonacgt_before.clear();
	uint64_t out = 0;
	size_t bi = 0;
	auto add = [&](uint64_t r, uint64_t start, uint64_t len, uint32_t nonacgt_before)
	{
		h->seg_in.push_back(start);
		h->seg_len.push_back((uint32_t)len);
		h->seg_read.push_back((uint32_t)r);
		h->seg_rawoff.push_back((uint32_t)(start - h->read_off[r]));
		h->seg_nonacgt_before.push_back(nonacgt_before);
		h->seg_out.push_back(out);
		out += (len + 1 + 63) & ~(uint64_t)63;
	};
	for (uint64_t r = 0; r < h->n_reads; r++)
	{
		const uint64_t a = h->read_off[r], b = read_end_of(h, r);
		while (bi < bad.size() && bad[bi] < a) bi++;
		if (b - a < 32) { while (bi < bad.size() && bad[bi] < b) bi++; continue; }
		if (bi >= bad.size() || bad[bi] >= b) { add(r, a, b - a, 0); continue; }
		// read with other characters: walk its non-ACGT positions
		const char* raw = read_ptr(r);
		uint64_t cur = a;
		uint32_t collapsed_other = 0; // characters outside ACGT that survive the collapse step, so far
		while (cur < b)
		{
			uint64_t stop = (bi < bad.size() && bad[bi] < b) ? bad[bi] : b;
			if (stop > cur) add(r, cur, stop - cur, collapsed_other);
			if (stop >= b) break;
			// run of non-ACGT characters starting at stop
			uint64_t e = stop;
			while (bi < bad.size() && bad[bi] == e && e < b)
			{
				const uint64_t i = e - a;
				// the reference compares the (already upper-cased) characters exactly (MBG/ReadHelper.h:873)
				char c = raw[i], p = i > 0 ? raw[i - 1] : 0;
				if (c >= 'a' && c <= 'z') c = (char)(c - 32);
				if (p >= 'a' && p <= 'z') p = (char)(p - 32);
				if (i == 0 || c != p) collapsed_other++;
				e++; bi++;
			}
			cur = e;
		}
	}
	h->total_out = out;
}

int upload_segments(mbgfe* h)
{
	const size_t n = h->seg_len.size();
	if (int rc = ensure(h, h->d_seg_in, (n + 1) * 8)) return rc;
	if (int rc = ensure(h, h->d_seg_out, (n + 1) * 8)) return rc;
	if (int rc = ensure(h, h->d_seg_len, (n + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_seg_c1, (n + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_seg_c2, (n + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_seg_nsel, (n + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_sel_begin, (n + 2) * 8)) return rc;
	if (n)
	{
		CU(h, cudaMemcpyAsync(h->d_seg_in.p, h->seg_in.data(), n * 8, cudaMemcpyHostToDevice, h->stream));
		CU(h, cudaMemcpyAsync(h->d_seg_out.p, h->seg_out.data(), n * 8, cudaMemcpyHostToDevice, h->stream));
		CU(h, cudaMemcpyAsync(h->d_seg_out.as<uint64_t>() + n, &h->total_out, 8, cudaMemcpyHostToDevice, h->stream));   // out_base[n]: end of the last slot
		CU(h, cudaMemcpyAsync(h->d_seg_len.p, h->seg_len.data(), n * 4, cudaMemcpyHostToDevice, h->stream));
		h->stats.h2d_bytes += n * 20;
	}
	return 0;
}

SelectPlan make_plan(uint32_t k, uint32_t w, size_t max_smem, bool& ok, uint32_t nt = NT_SELECT)
{
	SelectPlan p;
	const uint32_t s = k - w + 1, a = (w - 1) >> 3, T = nt * 8;
	p.LH = (s - 1 + 7) & ~7u;
	p.LW = 8 * (a + 1);
	p.NBP = (p.LH + T) / 8 + 2;
	p.NBH = nt + p.LW / 8 + 2;
	const size_t bytes = (size_t)p.NBP * 32 + (size_t)(p.NBP + 1) * 16 + (size_t)p.NBH * 10 * 8 + (size_t)p.NBH * 16 + BM_MIRROR * 8 + 64;
	p.smem_bytes = (uint32_t)bytes;
	ok = bytes + 2048 <= max_smem;
	return p;
}

void tick(mbgfe* h, int i) { cudaEventRecord(h->ev[i], h->stream); }
float elapsed(mbgfe* h, int a, int b) { float ms = 0; cudaEventElapsedTime(&ms, h->ev[a], h->ev[b]); return ms; }

int read_counter(mbgfe* h, int idx, uint64_t* out)
{
	unsigned long long v = 0;
	CU(h, cudaMemcpyAsync(&v, h->d_counters.as<unsigned long long>() + idx, 8, cudaMemcpyDeviceToHost, h->stream));
	CU(h, cudaStreamSynchronize(h->stream));
	h->stats.d2h_bytes += 8;
	*out = v;
	return 0;
}

template <typename T>
int download(mbgfe* h, std::vector<T>& dst, const DevBuf& src, size_t count, size_t first = 0)
{
	dst.resize(count);
	if (count == 0) return 0;
	CU(h, cudaMemcpyAsync(dst.data(), src.as<T>() + first, count * sizeof(T), cudaMemcpyDeviceToHost, h->stream));
	CU(h, cudaStreamSynchronize(h->stream));
	h->stats.d2h_bytes += count * sizeof(T);
	return 0;
}

void clear_results(mbgfe* h)
{
	h->processed = false; h->exported_parts = false; h->exported_table = false; h->exported_codes = false; h->have_table = false;
	h->nodes_ranked = false; h->edges_global = false; h->table_sharded = false; h->id_base = 0; h->x_packed_ids = nullptr; h->x_counts.clear();
	h->parts.clear(); h->h_codes.clear(); h->h_poses.clear(); h->h_positions.clear(); h->h_hashes.clear();
	h->n_nodes_out = 0; h->n_edges_out = 0;
	h->n_inst = 0;
	h->have_unitigs = false; h->have_thin = false; h->have_threads = false;
	h->h_thin.clear(); h->h_thin_pos.clear(); h->h_th_unitig.clear(); h->h_th_upos.clear(); h->h_th_urest.clear(); h->h_th_step.clear();
}

// ---------------------------------------------------------------------------------------------- file input

// FASTA / FASTQ reader with the reference's conventions (MBG/fastqloader.h): type by suffix after stripping ".gz",
// unknown suffix reads nothing; name = header line without its first character and trailing '\r'; FASTA sequence =
// concatenation of the following non-empty lines; a-z upper-cased.
int load_file(mbgfe* h, const char* path)
{
	std::string name = path;
	std::string base = name;
	if (base.size() >= 3 && base.compare(base.size() - 3, 3, ".gz") == 0) base.resize(base.size() - 3);
	auto ends = [&](const char* suf) { size_t n = strlen(suf); return base.size() >= n && base.compare(base.size() - n, n, suf) == 0; };
	const bool fastq = ends(".fastq") || ends(".fq");
	const bool fasta = ends(".fasta") || ends(".fa");
	gzFile f = gzopen(path, "rb");
	if (!f) return fail(h, MBGFE_E_IO, "cannot read %s", path);
	if (!fastq && !fasta) { gzclose(f); return 0; }
	gzbuffer(f, 1 << 20);
	auto* bases = new std::vector<char>();
	auto* offsets = new std::vector<uint64_t>();
	offsets->push_back(0);
	std::vector<char> buf(1 << 22);
	std::string line;
	bool in_record = false;
	int fastq_state = 0; // 0 header, 1 sequence, 2 plus, 3 quality
	auto finish_line = [&]()
	{
		if (!line.empty() && line.back() == '\r') line.pop_back();
		if (fasta)
		{
			if (line.empty()) return;
			if (line[0] == '>')
			{
				if (in_record) offsets->push_back(bases->size());
				h->names.push_back(line.substr(1));
				in_record = true;
				return;
			}
			if (!in_record) return;
			for (char c : line) bases->push_back((c >= 'a' && c <= 'z') ? (char)(c - 32) : c);
		}
		else
		{
			switch (fastq_state)
			{
				case 0:
					if (line.empty() || line[0] != '@') return;
					h->names.push_back(line.substr(1));
					fastq_state = 1;
					return;
				case 1:
					for (char c : line) bases->push_back((c >= 'a' && c <= 'z') ? (char)(c - 32) : c);
					offsets->push_back(bases->size());
					fastq_state = 2;
					return;
				case 2: fastq_state = 3; return;
				case 3: fastq_state = 0; return;
			}
		}
	};
	while (true)
	{
		int got = gzread(f, buf.data(), (unsigned)buf.size());
		if (got < 0) { gzclose(f); delete bases; delete offsets; return fail(h, MBGFE_E_IO, "read error in %s", path); }
		if (got == 0) break;
		const char* p = buf.data();
		const char* end = p + got;
		while (p < end)
		{
			const char* nl = (const char*)memchr(p, '\n', end - p);
			if (!nl) { line.append(p, end - p); break; }
			line.append(p, nl - p);
			finish_line();
			line.clear();
			p = nl + 1;
		}
	}
	if (!line.empty()) finish_line();
	if (fasta && in_record) offsets->push_back(bases->size());
	if (fastq && fastq_state == 1) { h->names.pop_back(); }
	gzclose(f);
	h->owned_bases.push_back(bases);
	h->owned_offsets.push_back(offsets);
	HostChunk c = { bases->data(), offsets->data(), offsets->size() - 1 };
	if (c.n_reads) h->pending.push_back(c);
	return 0;
}

}

namespace {
void destroy_comm(mbgfe* h);
int materialize_files(mbgfe* h);
int run_files(mbgfe* h);
void close_files(std::vector<mbgfe::DeferredFile>& v) { for (auto& f : v) if (f.fd >= 0) close(f.fd); v.clear(); }
int prepare_process(mbgfe* h);
uint32_t launch_front(mbgfe* h, uint32_t g0, uint32_t cnt, bool timed, cudaStream_t on = nullptr);
int process_tail(mbgfe* h, uint32_t launches);
}

// ====================================================================================================== C ABI

extern "C" {

const char* mbgfe_last_error(const mbgfe* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int mbgfe_create(const mbgfe_params* prm, mbgfe** out)
{
	if (!prm || !out) return fail(nullptr, MBGFE_E_PARAM, "null argument");
	*out = nullptr;
	if (prm->abi_version != MBGFE_ABI_VERSION) return fail(nullptr, MBGFE_E_PARAM, "ABI version %u, library has %u", prm->abi_version, MBGFE_ABI_VERSION);
	// MBG/main.cpp:158-198
	if (prm->k % 2 == 0) return fail(nullptr, MBGFE_E_PARAM, "k must be odd");
	if (prm->k < 31) return fail(nullptr, MBGFE_E_PARAM, "k must be at least 31");
	if (prm->w == 0) return fail(nullptr, MBGFE_E_PARAM, "w must be at least 1");
	if (prm->w > prm->k - 30) return fail(nullptr, MBGFE_E_PARAM, "w must be at most k-30");
	if (prm->error_masking > MBGFE_MASK_COLLAPSE_MSAT) return fail(nullptr, MBGFE_E_PARAM, "unknown error masking mode %u", prm->error_masking);
	if (prm->flags & MBGFE_INCLUDE_END_KMERS) return fail(nullptr, MBGFE_E_PARAM, "--include-end-kmers (MBG/ReadHelper.cpp:38-61) is not implemented");
	if (prm->world > MAX_WORLD) return fail(nullptr, MBGFE_E_PARAM, "world %u: at most %u ranks", prm->world, MAX_WORLD);
	if (prm->world == 0 || prm->rank >= prm->world) return fail(nullptr, MBGFE_E_PARAM, "rank %u outside world %u", prm->rank, prm->world);
	int n_dev = 0;
	cudaError_t e = cudaGetDeviceCount(&n_dev);
	if (e != cudaSuccess || n_dev == 0) return fail(nullptr, MBGFE_E_CUDA, "no CUDA device: libmbgfe has no CPU path (%s)", e != cudaSuccess ? cudaGetErrorString(e) : "0 devices");
	if (prm->device < 0 || prm->device >= n_dev) return fail(nullptr, MBGFE_E_CUDA, "device %d of %d", prm->device, n_dev);
	mbgfe* h = new mbgfe();
	h->prm = *prm;
	memset(&h->stats, 0, sizeof(h->stats));
	const uint32_t em = prm->error_masking;
	h->collapse = em == MBGFE_MASK_COLLAPSE || em == MBGFE_MASK_COLLAPSE_DINUC || em == MBGFE_MASK_COLLAPSE_MSAT;
	h->msat = em == MBGFE_MASK_DINUC || em == MBGFE_MASK_MSAT || em == MBGFE_MASK_COLLAPSE_DINUC || em == MBGFE_MASK_COLLAPSE_MSAT;
	h->max_motif = (em == MBGFE_MASK_DINUC || em == MBGFE_MASK_COLLAPSE_DINUC) ? 2 : 6;
	auto bail = [&](int rc) { g_create_error = h->err; mbgfe_destroy(h); return rc; };
	if (cudaSetDevice(prm->device) != cudaSuccess) return bail(fail(h, MBGFE_E_CUDA, "cudaSetDevice(%d) failed", prm->device));
	cudaDeviceProp prop;
	if (cudaGetDeviceProperties(&prop, prm->device) != cudaSuccess) return bail(fail(h, MBGFE_E_CUDA, "cudaGetDeviceProperties failed"));
	if (prop.major < 10) return bail(fail(h, MBGFE_E_CUDA, "device %s is sm_%d%d; libmbgfe is built for sm_100a only", prop.name, prop.major, prop.minor));
	h->sm_count = prop.multiProcessorCount;
	h->max_smem = prop.sharedMemPerBlockOptin;
	bool ok = false;
	h->plan = make_plan(prm->k, prm->w, h->max_smem, ok);
	if (!ok) return bail(fail(h, MBGFE_E_LIMIT, "k=%u needs more shared memory than one SM has (limit about k=11000)", prm->k));
	if (cudaStreamCreate(&h->stream) != cudaSuccess || cudaStreamCreate(&h->copy_stream) != cudaSuccess || cudaStreamCreate(&h->front_stream[0]) != cudaSuccess ||
		cudaStreamCreate(&h->front_stream[1]) != cudaSuccess) return bail(fail(h, MBGFE_E_CUDA, "cudaStreamCreate failed"));
	for (auto& e2 : h->ev) if (cudaEventCreate(&e2) != cudaSuccess) return bail(fail(h, MBGFE_E_CUDA, "cudaEventCreate failed"));
	if (ensure(h, h->tab_hash, MAX_CODE * 8) || ensure(h, h->tab_pair, MAX_CODE * 16) || ensure(h, h->tab_comp, MAX_CODE * 2) ||
		ensure(h, h->d_counters, (CTR_OWNER + prm->world + 8) * 8)) return bail(MBGFE_E_CUDA);
	MBGFE_LAUNCH(k_build_tables, (MAX_CODE + 255) / 256, 256, 0, h->stream, h->tab_hash.as<uint64_t>(), h->tab_pair.as<uint4>(), h->tab_comp.as<uint16_t>());
	if (ensure(h, h->tab_pext8, 65536)) return bail(MBGFE_E_CUDA);
	MBGFE_LAUNCH(k_build_pext8, 256, 256, 0, h->stream, h->tab_pext8.as<uint8_t>());
	cudaFuncSetAttribute(k_collapse_lut, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(65536 + COLLAPSE_LUT_WARPS * 2 * COLLAPSE_STAGE * 4));
	if (ensure(h, h->tab_in, MAX_CODE * 16) || ensure(h, h->tab_out, MAX_CODE * 16)) return bail(MBGFE_E_CUDA);
	MBGFE_LAUNCH(k_build_roll_tables, (MAX_CODE + 255) / 256, 256, 0, h->stream, h->tab_pair.as<uint4>(), h->tab_in.as<uint4>(), h->tab_out.as<uint4>(), prm->k - prm->w + 1);
	cudaFuncSetAttribute(k_select<NT_SELECT, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->plan.smem_bytes);
	cudaFuncSetAttribute(k_select<NT_SELECT, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->plan.smem_bytes);
	cudaFuncSetAttribute(k_select<NT_SELECT, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->plan.smem_bytes);
	cudaFuncSetAttribute(k_select<NT_SELECT, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->plan.smem_bytes);
	cudaFuncSetAttribute(k_select<NT_SELECT, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->plan.smem_bytes);
	cudaFuncSetAttribute(k_select<NT_SELECT, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->plan.smem_bytes);
	cudaFuncSetAttribute(k_select<NT_SELECT, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->plan.smem_bytes);
	cudaFuncSetAttribute(k_select<NT_SELECT, 7>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->plan.smem_bytes);
	if (cudaStreamSynchronize(h->stream) != cudaSuccess || cudaGetLastError() != cudaSuccess)
		return bail(fail(h, MBGFE_E_CUDA, "table kernel failed: is this an sm_100 device?"));
	*out = h;
	return 0;
}

void mbgfe_destroy(mbgfe* h)
{
	if (!h) return;
	for (DevBuf* b : { &h->tab_hash, &h->tab_pair, &h->tab_comp, &h->tab_pext8, &h->tab_in, &h->tab_out, &h->d_sketch, &h->d_ascii, &h->d_packed, &h->d_nonacgt, &h->d_counters, &h->d_seg_in, &h->d_seg_out,
		&h->d_seg_len, &h->d_seg_c1, &h->d_seg_c2, &h->d_seg_nsel, &h->d_sel_begin, &h->d_sym, &h->d_codes, &h->d_selmask, &h->d_pos1, &h->d_pos2,
		&h->d_inst_pos, &h->d_inst_seg, &h->d_inst_flag, &h->d_inst_hash, &h->d_inst_slot, &h->d_pal_list, &h->d_nodes, &h->d_edges, &h->d_node_out,
		&h->d_edge_out, &h->d_records, &h->d_recv, &h->d_bucket_count, &h->d_bucket_begin, &h->d_scan_tot, &h->d_scan_off, &h->d_node_hash_sorted, &h->d_node_cov,
		&h->d_node_tip, &h->d_edges_sorted, &h->d_inst_packpos, &h->d_ids_out, &h->d_id_begin, &h->d_edge_records, &h->d_xbuf[0], &h->d_xbuf[1], &h->d_xbuf[2],
		&h->d_xbuf[3], &h->d_var_hash, &h->d_var_offset, &h->d_var_len_begin, &h->d_var_len, &h->d_g_deg, &h->d_g_nb, &h->d_g_link, &h->d_g_st[0], &h->d_g_st[1],
		&h->d_g_kept, &h->d_g_newid, &h->d_g_flag, &h->d_g_at, &h->d_g_flag2, &h->d_g_at2, &h->d_g_flag_byte, &h->d_f_hash, &h->d_f_cov, &h->d_f_tip, &h->d_f_edges, &h->d_f_orig, &h->d_t_hash, &h->d_t_cov, &h->d_t_tip, &h->d_t_edges, &h->d_t_orig, &h->d_g_kept8, &h->d_u_pos, &h->d_th_final, &h->d_th_node, &h->d_th_unitig, &h->d_th_upos, &h->d_th_urest, &h->d_th_step, &h->d_thin, &h->d_thin_pos,
		&h->d_u_begin, &h->d_u_nodes, &h->d_u_of_node, &h->d_u_fw, &h->d_u_edges, &h->d_append_base, &h->d_range, &h->d_xrow, &h->d_xmatrix, &h->d_rec_slot, &h->d_thin_scan, &h->d_tpos, &h->d_tu_nodes, &h->d_tu_off, &h->d_tu_unitig, &h->d_tu_begin, &h->d_t_newov }) release(*b);
	for (HostBuf* b : { &h->p_node_hashes, &h->p_node_cov, &h->p_node_tip, &h->p_edges, &h->p_stage }) release_host(*b);
	for (auto& e : h->ev) if (e) cudaEventDestroy(e);
	for (auto& e : h->piece_events) cudaEventDestroy(e);
	if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
	for (auto& fs : h->front_stream) if (fs) cudaStreamDestroy(fs);
	for (auto& e : h->front_events) cudaEventDestroy(e);
	for (auto& e : h->pack_events) cudaEventDestroy(e);
	if (h->stream) cudaStreamDestroy(h->stream);
	for (auto* v : h->owned_bases) delete v;
	for (auto* v : h->owned_offsets) delete v;
	close_files(h->files); close_files(h->image_files);
	destroy_comm(h);
	delete h;
}

void* mbgfe_alloc_pinned(uint64_t bytes)
{
	void* p = nullptr;
	if (cudaMallocHost(&p, bytes) != cudaSuccess) return nullptr;
	return p;
}

void mbgfe_free_pinned(void* p) { if (p) cudaFreeHost(p); }

int mbgfe_add_reads(mbgfe* h, const char* bases, const uint64_t* offsets, uint64_t n_reads)
{
	if (!h) return MBGFE_E_PARAM;
	if (n_reads == 0) return 0;
	if (!bases || !offsets) return fail(h, MBGFE_E_PARAM, "null read buffer");
	for (uint64_t r = 0; r < n_reads; r++)
	{
		if (offsets[r + 1] < offsets[r]) return fail(h, MBGFE_E_PARAM, "offsets must be non-decreasing (read %llu)", (unsigned long long)r);
		if (offsets[r + 1] - offsets[r] >= (1ull << 29)) return fail(h, MBGFE_E_LIMIT, "read %llu is longer than 2^29-1 bases", (unsigned long long)r);
	}
	if (!h->files.empty()) if (int rc = materialize_files(h)) return rc;      // keep the order the caller added things in
	HostChunk c = { bases, offsets, n_reads };
	h->pending.push_back(c);
	return 0;
}

int mbgfe_add_file(mbgfe* h, const char* path)
{
	if (!h || !path) return MBGFE_E_PARAM;
	// a plain (uncompressed) FASTA file is not read here: mbgfe_run streams it -- read, scan and 2-bit packing in one pass on
	// worker threads, overlapped with the copies and the kernels (run_files).  Everything else (.gz: zlib is the bound; FASTQ)
	// is parsed now.
	{
		std::string base = path;
		auto ends = [&](const char* suf) { size_t n = strlen(suf); return base.size() >= n && base.compare(base.size() - n, n, suf) == 0; };
		struct stat sb;
		if ((ends(".fasta") || ends(".fa")) && h->pending.empty() && !getenv("MBGFE_FILE_LEGACY") && stat(path, &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size > 0)
		{
			const int fd = open(path, O_RDONLY);
			if (fd < 0) return fail(h, MBGFE_E_IO, "cannot read %s", path);
			h->files.push_back({ path, fd, (uint64_t)sb.st_size, 0 });
			return 0;
		}
	}
	if (!h->files.empty()) if (int rc = materialize_files(h)) return rc;
	// names are indexed by read; reads added without names before this file get empty names
	uint64_t have = 0;
	for (auto& c : h->pending) have += c.n_reads;
	if (h->names.size() < have) h->names.resize(have);
	return load_file(h, path);
}

const char* mbgfe_read_name(const mbgfe* h_, uint64_t read_idx)
{
	mbgfe* h = const_cast<mbgfe*>(h_);
	if (!h) return "";
	if (!h->files.empty()) materialize_files(h);      // names asked for before the run: parse now
	if (read_idx < h->name_off.size() && h->names.size() <= read_idx)
	{
		// reads of a streamed file: the header text is fetched from the file the first time a name is wanted
		h->names.resize(h->name_off.size());
		for (uint64_t r = 0; r < h->name_off.size(); r++)
		{
			std::string& n = h->names[r];
			n.resize(h->name_len[r]);
			const mbgfe::DeferredFile& f = h->image_files[h->name_file[r]];
			if (h->name_len[r] && pread(f.fd, &n[0], h->name_len[r], (off_t)(h->name_off[r] - f.base)) != (ssize_t)h->name_len[r]) n.clear();
			if (!n.empty() && n.back() == '\r') n.pop_back();
		}
	}
	if (read_idx >= h->names.size()) return "";
	return h->names[read_idx].c_str();
}

int mbgfe_reset(mbgfe* h)
{
	if (!h) return MBGFE_E_PARAM;
	h->pending.clear();
	h->names.clear();
	close_files(h->files); close_files(h->image_files);
	h->name_off.clear(); h->name_len.clear(); h->name_file.clear(); h->read_end.clear();
	for (auto* v : h->owned_bases) delete v;
	for (auto* v : h->owned_offsets) delete v;
	h->owned_bases.clear(); h->owned_offsets.clear();
	h->resident = false;
	h->n_reads = 0; h->n_bases = 0;
	h->append_base = 0; h->appending = false;
	clear_results(h);
	return 0;
}

int mbgfe_upload(mbgfe* h)
{
	if (!h) return MBGFE_E_PARAM;
	if (!h->files.empty()) if (int rc = materialize_files(h)) return rc;
	clear_results(h);
	h->resident = false;
	h->stats.h2d_bytes = 0; h->stats.d2h_bytes = 0;
	uint64_t n_reads = 0, n_bases = 0;
	for (auto& c : h->pending) { n_reads += c.n_reads; n_bases += c.offsets[c.n_reads] - c.offsets[0]; }
	if (n_reads >= (1ull << 32)) return fail(h, MBGFE_E_LIMIT, "more than 2^32-1 reads in one batch");
	h->n_reads = n_reads; h->n_bases = n_bases;
	h->read_off.resize(n_reads + 1);
	h->read_end.clear();
	h->name_off.clear(); h->name_len.clear(); h->name_file.clear(); close_files(h->image_files);
	std::vector<const char*> read_ptr(n_reads);
	if (int rc = ensure(h, h->d_ascii, n_bases + 64)) return rc;
	if (int rc = ensure(h, h->d_packed, n_bases / 4 + 64)) return rc;
	const uint64_t bad_cap = std::max<uint64_t>(1 << 20, n_bases / 16);
	if (int rc = ensure(h, h->d_nonacgt, bad_cap * 8)) return rc;
	tick(h, 0);
	CU(h, cudaMemsetAsync(h->d_counters.p, 0, h->d_counters.cap, h->stream));
	uint64_t pos = 0, r = 0;
	for (auto& c : h->pending)
	{
		const uint64_t first = c.offsets[0], bytes = c.offsets[c.n_reads] - first;
		if (bytes) CU(h, cudaMemcpyAsync(h->d_ascii.as<char>() + pos, c.bases + first, bytes, cudaMemcpyHostToDevice, h->stream));
		for (uint64_t i = 0; i < c.n_reads; i++, r++) { h->read_off[r] = pos + (c.offsets[i] - first); read_ptr[r] = c.bases + c.offsets[i]; }
		pos += bytes;
	}
	h->read_off[n_reads] = pos;
	h->stats.h2d_bytes += n_bases;
	tick(h, 1);
	if (n_bases)
	{
		const uint64_t chunks = (n_bases + 31) / 32;
		const unsigned grid = (unsigned)std::min<uint64_t>((chunks + 255) / 256, (uint64_t)h->sm_count * 16);
		MBGFE_LAUNCH(k_pack_ascii, grid, 256, 0, h->stream, (const uint4*)h->d_ascii.p, n_bases, (uint64_t)0, chunks, h->d_packed.as<uint64_t>(), h->d_nonacgt.as<uint64_t>(),
			h->d_counters.as<unsigned long long>() + CTR_NONACGT, bad_cap);
	}
	tick(h, 2);
	uint64_t n_bad = 0;
	if (int rc = read_counter(h, CTR_NONACGT, &n_bad)) return rc;
	CU(h, cudaGetLastError());
	if (n_bad > bad_cap) return fail(h, MBGFE_E_LIMIT, "%llu characters outside ACGT in %llu bases: more than 1 in 16", (unsigned long long)n_bad, (unsigned long long)n_bases);
	std::vector<uint64_t> bad;
	if (n_bad) { if (int rc = download(h, bad, h->d_nonacgt, n_bad)) return rc; std::sort(bad.begin(), bad.end()); }
	h->nonacgt_count = n_bad;
	build_segments(h, bad, [&](uint64_t r_) { return read_ptr[r_]; });
	if (int rc = upload_segments(h)) return rc;
	CU(h, cudaStreamSynchronize(h->stream));
	h->stats.ms_upload = elapsed(h, 0, 1);
	h->stats.ms_pack = elapsed(h, 1, 2);
	h->stats.reads = n_reads; h->stats.raw_bases = n_bases; h->stats.segments = h->seg_len.size(); h->stats.nonacgt = n_bad;
	h->pending.clear();
	h->resident = true;
	return 0;
}

int mbgfe_process(mbgfe* h)
{
	if (!h) return MBGFE_E_PARAM;
	if (!h->resident) return fail(h, MBGFE_E_STATE, "mbgfe_process before mbgfe_upload");
	clear_results(h);
	if (int rc = prepare_process(h)) return rc;
	tick(h, 3);
	const uint32_t launches = launch_front(h, 0, (uint32_t)h->seg_len.size(), true);
	return process_tail(h, launches);
}

}

namespace {

// device buffers of one pass over the resident batch + counter reset
int prepare_process(mbgfe* h)
{
	const bool want_poses = want_poses_of(h);
	const bool need_pos1 = want_poses && !h->collapse;
	const uint64_t tot = h->total_out;
	if (int rc = ensure(h, h->d_sym, tot / 4 + 64)) return rc;
	if (int rc = ensure(h, h->d_codes, tot * 2 + 64)) return rc;
	if (int rc = ensure(h, h->d_selmask, tot / 8 + 64)) return rc;
	if (need_pos1) if (int rc = ensure(h, h->d_pos1, tot * 4 + 64)) return rc;
	if (want_poses) if (int rc = ensure(h, h->d_pos2, tot * 4 + 64)) return rc;
	unsigned long long* ctr = h->d_counters.as<unsigned long long>();
	CU(h, cudaMemsetAsync(ctr + CTR_PALINDROME, 0, (CTR_OWNER + h->prm.world - CTR_PALINDROME) * 8, h->stream));
	return 0;
}

// K1 -> K2 -> K3a on segments [g0, g0 + cnt).  Segments are independent, so the pipelined upload can run each piece's
// segments while later pieces are still crossing PCIe.  timed: record the per-stage events (whole-batch launches only).
uint32_t launch_front(mbgfe* h, uint32_t g0, uint32_t cnt, bool timed, cudaStream_t on)
{
	if (!on) on = h->stream;
	if (cnt == 0) { if (timed) { tick(h, 4); tick(h, 5); } return 0; }
	const uint32_t k = h->prm.k, w = h->prm.w;
	const bool want_poses = want_poses_of(h);
	const bool need_pos1 = want_poses && !h->collapse;
	SegTable st;
	st.in_start = h->d_seg_in.as<uint64_t>() + g0;
	st.len = h->d_seg_len.as<uint32_t>() + g0;
	st.out_base = h->d_seg_out.as<uint64_t>() + g0;
	st.n = cnt;
	uint32_t* pos1 = need_pos1 ? h->d_pos1.as<uint32_t>() : nullptr;
	uint32_t* pos2 = want_poses ? h->d_pos2.as<uint32_t>() : nullptr;
	uint32_t* c1 = h->d_seg_c1.as<uint32_t>() + g0;
	uint32_t* c2 = h->d_seg_c2.as<uint32_t>() + g0;
	if (!pos1 && !getenv("MBGFE_COLLAPSE_BITWISE"))
	{
#ifdef MBGFE_EMU_BUILD
		const unsigned grid1 = cnt;
#else
		const unsigned grid1 = std::min<unsigned>((cnt + COLLAPSE_LUT_WARPS - 1) / COLLAPSE_LUT_WARPS, (unsigned)h->sm_count * 3);      // persistent: 3 CTAs of 72 KB per SM
#endif
		MBGFE_LAUNCH(k_collapse_lut, grid1, COLLAPSE_LUT_WARPS * 32, 65536 + COLLAPSE_LUT_WARPS * 2 * COLLAPSE_STAGE * 4, on, st, h->d_packed.as<uint64_t>(), h->d_sym.as<uint64_t>(), c1,
			h->tab_pext8.as<uint8_t>(), h->prm.error_masking == MBGFE_MASK_NO ? 1 : 0);
	}
	else
		MBGFE_LAUNCH(k_collapse, (cnt + COLLAPSE_WARPS - 1) / COLLAPSE_WARPS, COLLAPSE_WARPS * 32, 0, on, st, h->d_packed.as<uint64_t>(), h->d_sym.as<uint64_t>(), c1, pos1,
			h->prm.error_masking == MBGFE_MASK_NO ? 1 : 0);
	if (timed) tick(h, 4);
	if (h->msat)
		MBGFE_LAUNCH(k_msat, (cnt + MSAT_WARPS - 1) / MSAT_WARPS, MSAT_WARPS * 32, 0, on, st, h->d_sym.as<uint64_t>(), c1, h->d_codes.as<uint16_t>(),
			c2, (const uint32_t*)pos1, pos2, h->max_motif);
	else
		MBGFE_LAUNCH(k_unpack_symbols<NT_MSAT>, cnt, NT_MSAT, 0, on, st, h->d_sym.as<uint64_t>(), c1, h->d_codes.as<uint16_t>(),
			c2, (const uint32_t*)pos1, pos2);
	if (timed) tick(h, 5);
#define MBGFE_SELECT_CASE(SR) case SR: MBGFE_LAUNCH((k_select<NT_SELECT, SR>), cnt, NT_SELECT, h->plan.smem_bytes, on, st, h->d_codes.as<uint16_t>(), c2, h->tab_pair.as<uint4>(), \
		h->tab_in.as<uint4>(), h->tab_out.as<uint4>(), h->d_selmask.as<uint8_t>(), h->d_seg_nsel.as<uint32_t>() + g0, k, w, h->plan); break;
	switch ((k - w + 1) & 7)      // s mod 8
	{
		MBGFE_SELECT_CASE(0) MBGFE_SELECT_CASE(1) MBGFE_SELECT_CASE(2) MBGFE_SELECT_CASE(3) MBGFE_SELECT_CASE(4) MBGFE_SELECT_CASE(5) MBGFE_SELECT_CASE(6) MBGFE_SELECT_CASE(7)
	}
#undef MBGFE_SELECT_CASE
	return 3;
}

// ---- everything after selection: instance list, k-mer hashes, palindrome rule, node and edge tables.
// All of it is driven by instance ranges read from device memory (entries of sel_begin), so the same launches serve the
// whole batch (mbgfe_process: exact sizes known) and one piece of it (mbgfe_run: sizes estimated, checked afterwards).

// grows a buffer and keeps its first `keep` bytes (the instances of earlier batches)
int ensure_keep(mbgfe* h, DevBuf& b, size_t bytes, size_t keep)
{
	bytes = (bytes + 255) & ~(size_t)255;
	if (b.cap >= bytes) return 0;
	if (!keep || !b.p) return ensure(h, b, bytes);
	void* np = nullptr;
	CU(h, cudaMalloc(&np, bytes));
	CU(h, cudaMemcpyAsync(np, b.p, std::min(keep, b.cap), cudaMemcpyDeviceToDevice, h->stream));
	CU(h, cudaStreamSynchronize(h->stream));
	CU(h, cudaFree(b.p));
	b.p = np; b.cap = bytes;
	return 0;
}

int ensure_instances(mbgfe* h, uint64_t cap)
{
	const uint64_t keep = h->appending ? h->append_base : 0;
	cap += keep > cap ? keep : 0;
	if (int rc = ensure_keep(h, h->d_inst_pos, cap * 4 + 64, keep * 4)) return rc;
	if (int rc = ensure_keep(h, h->d_inst_seg, cap * 4 + 64, keep * 4)) return rc;
	if (int rc = ensure_keep(h, h->d_inst_flag, cap + 64, keep)) return rc;
	if (int rc = ensure_keep(h, h->d_inst_hash, cap * 16 + 64, keep * 16)) return rc;
	if (int rc = ensure(h, h->d_inst_slot, cap * 4 + 64)) return rc;
	if (int rc = ensure(h, h->d_pal_list, PAL_CAP * 8)) return rc;
	h->inst_cap = cap;
	return 0;
}

// empty node and edge tables of that many slots (powers of two)
int clear_tables(mbgfe* h, uint64_t node_slots, uint64_t edge_slots, uint32_t& launches)
{
	if (node_slots >= (1ull << 31) || edge_slots >= (1ull << 31))
		return fail(h, MBGFE_E_LIMIT, "the tables would need %llu / %llu slots (limit 2^31 per GPU)", (unsigned long long)node_slots, (unsigned long long)edge_slots);
	if (int rc = ensure(h, h->d_nodes, node_slots * sizeof(NodeSlot))) return rc;
	h->node_slots = node_slots;
	MBGFE_LAUNCH(k_table_clear, (unsigned)std::min<uint64_t>((node_slots + 255) / 256, (uint64_t)h->sm_count * 8), 256, 0, h->stream, h->d_nodes.as<NodeSlot>(), node_slots);
	launches++;
	if (!(h->prm.flags & MBGFE_NO_EDGES))
	{
		if (int rc = ensure(h, h->d_edges, edge_slots * sizeof(EdgeSlot))) return rc;
		h->edge_slots = edge_slots;
		MBGFE_LAUNCH(k_edge_clear, (unsigned)std::min<uint64_t>((edge_slots + 255) / 256, (uint64_t)h->sm_count * 8), 256, 0, h->stream, h->d_edges.as<EdgeSlot>(), edge_slots);
		launches++;
	}
	else h->edge_slots = 0;
	return 0;
}

// distinct keys behind `ones` set bits of an M-bit linear-counting bitmap, with a little head room
uint64_t sketch_estimate(uint64_t ones, uint64_t M)
{
	if (ones >= M) return 8 * M;
	return (uint64_t)(-(double)M * std::log(1.0 - (double)ones / (double)M) * 1.03) + 256;
}

// runs a sketch launch `fill` (or two: nodes and edges) over a fresh bitmap pair of M bits each and returns the two estimates
template <typename Fill>
int run_sketch(mbgfe* h, uint64_t n_keys, bool second, uint64_t est[2], uint32_t& launches, Fill fill)
{
	const uint64_t M = std::max<uint64_t>(next_pow2(n_keys), 1 << 16), words = M / 32;
	if (int rc = ensure(h, h->d_sketch, 256 + 2 * words * 4)) return rc;
	CU(h, cudaMemsetAsync(h->d_sketch.p, 0, 256 + (second ? 2 : 1) * words * 4, h->stream));
	uint32_t* bm0 = (uint32_t*)(h->d_sketch.as<char>() + 256);
	uint32_t* bm1 = second ? bm0 + words : (uint32_t*)nullptr;
	fill(bm0, bm1, M - 1);
	unsigned long long* ones = h->d_sketch.as<unsigned long long>();
	const unsigned grid = (unsigned)std::min<uint64_t>((words + 255) / 256, (uint64_t)h->sm_count * 8);
	MBGFE_LAUNCH(k_sketch_count, grid, 256, 0, h->stream, bm0, words, ones);
	if (second) MBGFE_LAUNCH(k_sketch_count, grid, 256, 0, h->stream, bm1, words, ones + 1);
	launches += second ? 3 : 2;
	unsigned long long got[2] = { 0, 0 };
	CU(h, cudaMemcpyAsync(got, ones, 16, cudaMemcpyDeviceToHost, h->stream));
	CU(h, cudaStreamSynchronize(h->stream));
	h->stats.d2h_bytes += 16;
	est[0] = sketch_estimate(got[0], M); est[1] = second ? sketch_estimate(got[1], M) : 0;
	return 0;
}

// (experiments and tests: MBGFE_NO_SKETCH = size from the instances as round 1 did; MBGFE_SKETCH_MIN = smallest input that is sketched)
bool sketch_wanted(uint64_t n_keys)
{
	// Off unless MBGFE_SKETCH_MIN says from how many keys on.  Measured: chr20 bench (15.6 M instances, 7.6 M distinct k-mers) -- the sketch
	// halves both tables but costs 0.5 ms itself and the fuller node table probes longer, 17.2 -> 17.8 ms; HG002-shaped quarter genome on 2
	// GPUs (189 M records per GPU, 92 M distinct: at k = 1001 most k-mers of reads with 0.1 % errors hold an error and are unique) --
	// exchange 84.1 -> 85.7 ms.  It pays in memory only (the two tables of a rank of the full HG002 run: 34 GB -> 17 GB), and in time only
	// where k-mers repeat 4x or more.
	const char* min_env = getenv("MBGFE_SKETCH_MIN");
	return n_keys > 0 && min_env && !getenv("MBGFE_NO_SKETCH") && n_keys >= (uint64_t)atoll(min_env);
}

// Table sizes for the resident instances: two slots per DISTINCT k-mer / edge expected (a sketch of the hashes, k_sketch_instances),
// never more than two per instance, never less than the caller asked for.  full = what the number of instances alone would give.
int size_tables(mbgfe* h, uint64_t n_inst, uint64_t* node_slots, uint64_t* edge_slots, uint64_t* full_out, uint32_t& launches)
{
	const uint64_t floor_slots = h->prm.table_slots ? next_pow2(h->prm.table_slots) : 0;
	const uint64_t full = std::max<uint64_t>(next_pow2(std::max<uint64_t>(2 * n_inst, 1024)), floor_slots);
	*node_slots = full; *edge_slots = full; *full_out = full;
	if (!sketch_wanted(n_inst)) return 0;
	const bool edges = !(h->prm.flags & MBGFE_NO_EDGES);
	uint64_t est[2];
	if (int rc = run_sketch(h, n_inst, edges, est, launches, [&](uint32_t* bm0, uint32_t* bm1, uint64_t mask)
	{
		MBGFE_LAUNCH(k_sketch_instances, (unsigned)std::min<uint64_t>((n_inst + 255) / 256, (uint64_t)h->sm_count * 16), 256, 0, h->stream, h->d_inst_hash.as<uint64_t>(),
			h->d_inst_flag.as<uint8_t>(), n_inst, bm0, bm1, mask);
	})) return rc;
	if (const char* d = getenv("MBGFE_SKETCH_DIV")) { const uint64_t by = std::max(1, atoi(d)); est[0] /= by; est[1] /= by; }      // tests: an estimate that is too low
	*node_slots = std::min(full, std::max<uint64_t>(next_pow2(std::max<uint64_t>(2 * est[0], 1024)), floor_slots));
	if (edges) *edge_slots = std::min(full, std::max<uint64_t>(next_pow2(std::max<uint64_t>(2 * est[1], 1024)), floor_slots));
	return 0;
}

// selected-count scan of segments [g0, g1): sel_begin[g0] must hold the instances before g0 (0 for g0 == 0)
void launch_scan(mbgfe* h, uint32_t g0, uint32_t g1, uint32_t& launches)
{
	uint64_t* sb = h->d_sel_begin.as<uint64_t>();
	// the scan of the first segments starts from the instances earlier batches left (mbgfe_run_append), else from 0
	const uint64_t* first_carry = (h->appending && h->append_base) ? h->d_append_base.as<uint64_t>() : (const uint64_t*)nullptr;
	MBGFE_LAUNCH(k_scan_u32<1024>, 1, 1024, 0, h->stream, h->d_seg_nsel.as<uint32_t>() + g0, sb + g0, g1 - g0, g0 ? (const uint64_t*)(sb + g0) : first_carry);
	launches++;
}

// K3b2, one CTA per segment (also behind the two known-answer entry points)
void launch_kmer_hash(cudaStream_t stream, uint32_t n_seg, const uint64_t* seg_out_base, const uint16_t* codes, const uint16_t* tab_comp, const uint32_t* inst_pos,
	const uint64_t* sel_begin, uint64_t inst_cap, uint64_t* inst_hash, uint32_t k)
{
	MBGFE_LAUNCH(k_kmer_hash, n_seg, HASH_NT, 0, stream, seg_out_base, codes, tab_comp, inst_pos, sel_begin, inst_cap, inst_hash, k);
}

// K3b for the instances of segments [g0, g1).  n_hint: instance count when the host knows it (exact grids), else 0.
void launch_list_hash(mbgfe* h, uint32_t g0, uint32_t g1, uint64_t n_hint, uint32_t& launches)
{
	if (g1 == g0) return;
	const uint32_t k = h->prm.k;
	unsigned long long* ctr = h->d_counters.as<unsigned long long>();
	SegTable st = seg_table(h);
	st.in_start += g0; st.len += g0; st.out_base += g0; st.n = g1 - g0;
	const uint64_t* sb = h->d_sel_begin.as<uint64_t>();
	MBGFE_LAUNCH(k_list_selected<NT_LIST>, g1 - g0, NT_LIST, 0, h->stream, st, g0, h->d_selmask.as<uint32_t>(), h->d_seg_c2.as<uint32_t>() + g0, h->d_seg_nsel.as<uint32_t>() + g0,
		sb + g0, h->d_inst_pos.as<uint32_t>(), h->d_inst_seg.as<uint32_t>(), h->d_inst_flag.as<uint8_t>(), k, h->inst_cap, ctr + CTR_OVERFLOW);
	launch_kmer_hash(h->stream, g1 - g0, h->d_seg_out.as<uint64_t>() + g0, h->d_codes.as<uint16_t>(), h->tab_comp.as<uint16_t>(), h->d_inst_pos.as<uint32_t>(),
		sb + g0, h->inst_cap, h->d_inst_hash.as<uint64_t>(), k);
	const unsigned pal_grid = n_hint ? (unsigned)((n_hint + 255) / 256) : (unsigned)h->sm_count * 4;
	MBGFE_LAUNCH(k_find_palindromes, pal_grid, 256, 0, h->stream, h->d_inst_hash.as<uint64_t>(), sb + g0, sb + g1, h->inst_cap, ctr + CTR_PALINDROME, h->d_pal_list.as<uint64_t>(), PAL_CAP);
	launches += 3;
}

// Pass 3 of runMBG: hashes of k-mers with registered hpc variants are perturbed after the palindrome rule
// (MBG/ReadHelper.h:672-737) and before anything is counted.  Runs per piece between K3b and K4, so pass 3 keeps the
// pipelined upload; a palindrome in the batch sends the run through process_tail, which re-hashes and re-applies.
void launch_variants(mbgfe* h, uint32_t g0, uint32_t g1, uint64_t n_hint, uint32_t& launches)
{
	if (!h->n_var || g1 == g0) return;
	VariantTable vt;
	vt.hash = h->d_var_hash.as<uint64_t>(); vt.offset = h->d_var_offset.as<uint32_t>();
	vt.len_begin = h->d_var_len_begin.as<uint64_t>(); vt.len = h->d_var_len.as<uint32_t>(); vt.n = h->n_var;
	const uint64_t* sb = h->d_sel_begin.as<uint64_t>();
	const unsigned grid = n_hint ? (unsigned)std::min<uint64_t>((n_hint + 255) / 256, (uint64_t)h->sm_count * 16) : (unsigned)h->sm_count * 8;
	MBGFE_LAUNCH(k_apply_variants, grid, 256, 0, h->stream, h->d_inst_hash.as<uint64_t>(), h->d_inst_pos.as<uint32_t>(), h->d_inst_seg.as<uint32_t>(), sb + g0, sb + g1, h->inst_cap,
		h->d_seg_out.as<uint64_t>(), h->d_pos2.as<uint32_t>(), vt, h->prm.k);
	launches++;
}

// K4a + K4b for the instances of segments [g0, g1)
void launch_tables(mbgfe* h, uint32_t g0, uint32_t g1, uint64_t n_hint, uint32_t& launches, bool timed)
{
	if ((h->prm.flags & MBGFE_NO_TABLE) || h->appending) { if (timed) { tick(h, 8); tick(h, 9); } return; }
	unsigned long long* ctr = h->d_counters.as<unsigned long long>();
	const uint64_t* sb = h->d_sel_begin.as<uint64_t>();
	const unsigned grid = n_hint ? (unsigned)((n_hint + 255) / 256) : (unsigned)h->sm_count * 8;
	if (g1 > g0)
	{
		MBGFE_LAUNCH(k_insert_nodes, grid, 256, 0, h->stream, h->d_inst_hash.as<uint64_t>(), h->d_inst_flag.as<uint8_t>(), sb + g0, sb + g1, h->inst_cap, h->d_nodes.as<NodeSlot>(),
			h->node_slots - 1, h->d_inst_slot.as<uint32_t>(), ctr + CTR_DISTINCT, ctr + CTR_OVERFLOW);
		launches++;
	}
	if (timed) tick(h, 8);
	if (g1 > g0 && !(h->prm.flags & MBGFE_NO_EDGES))
	{
		MBGFE_LAUNCH(k_insert_edges, grid, 256, 0, h->stream, h->d_inst_hash.as<uint64_t>(), h->d_inst_pos.as<uint32_t>(), h->d_inst_flag.as<uint8_t>(), h->d_inst_slot.as<uint32_t>(),
			sb + g0, sb + g1, h->inst_cap, h->d_edges.as<EdgeSlot>(), h->edge_slots - 1, ctr + CTR_EDGES, ctr + CTR_OVERFLOW, h->prm.k);
		launches++;
	}
	if (timed) tick(h, 9);
}

void finish_stats(mbgfe* h, uint64_t n_inst, uint32_t launches, bool staged)
{
	h->stats.ms_collapse = staged ? elapsed(h, 3, 4) : 0;
	h->stats.ms_msat = staged ? elapsed(h, 4, 5) : 0;
	h->stats.ms_select = staged ? elapsed(h, 5, 6) : 0;
	h->stats.ms_hash = staged ? elapsed(h, 6, 7) : 0;
	h->stats.ms_nodes = staged ? elapsed(h, 7, 8) : 0;
	h->stats.ms_edges = staged ? elapsed(h, 8, 9) : 0;
	h->stats.ms_process = elapsed(h, 3, 9);
	h->stats.selected = n_inst;
	h->stats.launches = launches;
	h->processed = true;
}

// The tail with exact sizes: scan, read the instance count, size the lists, hash, apply the palindrome rule, build the
// tables.  Used by mbgfe_process, and by mbgfe_run when its estimates did not hold or palindromes were found.
int process_tail(mbgfe* h, uint32_t launches)
{
	const uint32_t k = h->prm.k;
	const uint32_t n_seg = (uint32_t)h->seg_len.size();
	unsigned long long* ctr = h->d_counters.as<unsigned long long>();
	const SegTable st = seg_table(h);
	CU(h, cudaMemsetAsync(ctr + CTR_PALINDROME, 0, (CTR_OWNER - CTR_PALINDROME) * 8, h->stream));
	launch_scan(h, 0, n_seg, launches);
	tick(h, 6);
	uint64_t n_inst = 0;
	CU(h, cudaMemcpyAsync(&n_inst, h->d_sel_begin.as<uint64_t>() + n_seg, 8, cudaMemcpyDeviceToHost, h->stream));
	CU(h, cudaStreamSynchronize(h->stream));
	CU(h, cudaGetLastError());
	h->stats.d2h_bytes += 8;
	h->n_inst = n_inst;
	if (int rc = ensure_instances(h, std::max<uint64_t>(n_inst, h->inst_cap))) return rc;
	for (int pass = 0; pass < 2; pass++)
	{
		if (n_inst) launch_list_hash(h, 0, n_seg, n_inst, launches);
		if (pass == 0) tick(h, 7);
		uint64_t n_pal = 0;
		if (int rc = read_counter(h, CTR_PALINDROME, &n_pal)) return rc;
		if (n_pal == 0) break;
		// rare path (MBG/ReadHelper.h:755-785): drop what the reference drops, fail where it aborts, rebuild the list
		if (pass == 1) return fail(h, MBGFE_E_PALINDROME, "palindromic k-mers remain after dropping");
		if (n_pal > PAL_CAP) return fail(h, MBGFE_E_LIMIT, "%llu palindromic k-mers (limit %llu)", (unsigned long long)n_pal, (unsigned long long)PAL_CAP);
		MBGFE_LAUNCH(k_resolve_palindromes, (unsigned)((n_pal + 127) / 128), 128, 0, h->stream, h->d_pal_list.as<uint64_t>(), n_pal, h->d_inst_pos.as<uint32_t>(),
			h->d_inst_seg.as<uint32_t>(), h->d_inst_flag.as<uint8_t>(), h->d_seg_out.as<uint64_t>(), h->d_codes.as<uint16_t>(), h->tab_comp.as<uint16_t>(),
			h->d_selmask.as<uint32_t>(), k, ctr + CTR_CURSOR);
		unsigned long long errs[2] = { 0, 0 };
		CU(h, cudaMemcpyAsync(errs, ctr + CTR_CURSOR, 16, cudaMemcpyDeviceToHost, h->stream));
		CU(h, cudaStreamSynchronize(h->stream));
		if (errs[1]) return fail(h, MBGFE_E_UNHASHABLE, "unhashable k-mer: forward and reverse hashes agree on a non-palindrome (%llu)", errs[1]);
		if (errs[0]) return fail(h, MBGFE_E_PALINDROME, "the genome has a palindromic k-mer (%llu); cannot build a graph, try a different -w", errs[0]);
		MBGFE_LAUNCH(k_recount_selected<NT_LIST>, n_seg, NT_LIST, 0, h->stream, st, h->d_selmask.as<uint32_t>(), h->d_seg_c2.as<uint32_t>(), h->d_seg_nsel.as<uint32_t>(), k);
		launches++;
		launch_scan(h, 0, n_seg, launches);
		CU(h, cudaMemsetAsync(ctr + CTR_PALINDROME, 0, 8, h->stream));
		CU(h, cudaMemcpyAsync(&n_inst, h->d_sel_begin.as<uint64_t>() + n_seg, 8, cudaMemcpyDeviceToHost, h->stream));
		CU(h, cudaStreamSynchronize(h->stream));
		h->n_inst = n_inst;
		launches++;
	}
	h->stats.distinct = 0; h->stats.distinct_edges = 0;
	if (n_inst) launch_variants(h, 0, n_seg, n_inst, launches);
	if (!(h->prm.flags & MBGFE_NO_TABLE) && !h->appending)
	{
		uint64_t node_slots, edge_slots, full;
		if (int rc = size_tables(h, n_inst, &node_slots, &edge_slots, &full, launches)) return rc;
		for (int attempt = 0; ; attempt++)
		{
			if (int rc = clear_tables(h, node_slots, edge_slots, launches)) return rc;
			launch_tables(h, 0, n_seg, n_inst, launches, true);
			unsigned long long five[5] = { 0, 0, 0, 0, 0 };      // CTR_DISTINCT, CTR_EDGES, cursor x2, CTR_OVERFLOW
			CU(h, cudaMemcpyAsync(five, ctr + CTR_DISTINCT, sizeof(five), cudaMemcpyDeviceToHost, h->stream));
			CU(h, cudaStreamSynchronize(h->stream));
			h->stats.d2h_bytes += sizeof(five);
			h->stats.distinct = five[0]; h->stats.distinct_edges = five[1];
			if (!five[CTR_OVERFLOW - CTR_DISTINCT]) break;
			// the probe limit (MAX_PROBES) gave up on some k-mer or edge: the result would silently miss it.  A sketch that came out low is
			// mended by doubling; tables already at their full size cannot be.
			if ((node_slots >= full && edge_slots >= full) || attempt == 4) return fail(h, MBGFE_E_LIMIT, "k-mer table of %llu slots overflowed (%llu inserts gave up); pass a larger mbgfe_params.table_slots",
				(unsigned long long)node_slots, five[CTR_OVERFLOW - CTR_DISTINCT]);
			node_slots = std::min(full, 2 * node_slots); edge_slots = std::min(full, 2 * edge_slots);
			h->stats.table_grows++;
			CU(h, cudaMemsetAsync(ctr + CTR_DISTINCT, 0, (CTR_OVERFLOW - CTR_DISTINCT + 1) * 8, h->stream));
		}
		h->stats.table_slots = node_slots;
		h->have_table = true;
	}
	else { tick(h, 8); tick(h, 9); CU(h, cudaStreamSynchronize(h->stream)); }
	CU(h, cudaGetLastError());
	finish_stats(h, n_inst, launches, true);
	return 0;
}

}

extern "C" {

// mbgfe_run: upload and front end pipelined.  The ASCII reads cross PCIe in pieces on a copy stream; as soon as a piece
// has landed it is packed and the segments that are complete run K1 -> K2 -> K3a on the compute stream, so the front
// end hides behind the transfer.  The segment table is built before anything is copied, from the read lengths alone,
// assuming no characters outside ACGT (true for corrected HiFi reads); K0 counts such characters, and if it finds any
// the batch -- resident by then -- is cut at them and reprocessed the ordinary way.
int mbgfe_run(mbgfe* h)
{
	if (!h) return MBGFE_E_PARAM;
	if (!h->files.empty())
	{
		const int rc = run_files(h);
		if (rc != 1) return rc;
		// 1: a file the streaming loader does not take (sequence split over several lines, too many short reads): parse it the plain way
		if (int rc2 = materialize_files(h)) return rc2;
	}
	clear_results(h);
	h->resident = false;
	h->stats.h2d_bytes = 0; h->stats.d2h_bytes = 0;
	uint64_t n_reads = 0, n_bases = 0;
	for (auto& c : h->pending) { n_reads += c.n_reads; n_bases += c.offsets[c.n_reads] - c.offsets[0]; }
	if (n_reads >= (1ull << 32)) return fail(h, MBGFE_E_LIMIT, "more than 2^32-1 reads in one batch");
	h->n_reads = n_reads; h->n_bases = n_bases;
	h->read_off.resize(n_reads + 1);
	h->read_end.clear();
	h->name_off.clear(); h->name_len.clear(); h->name_file.clear(); close_files(h->image_files);
	std::vector<const char*> read_ptr(n_reads);
	{
		uint64_t pos = 0, r = 0;
		for (auto& c : h->pending)
		{
			const uint64_t first = c.offsets[0];
			for (uint64_t i = 0; i < c.n_reads; i++, r++) { h->read_off[r] = pos + (c.offsets[i] - first); read_ptr[r] = c.bases + c.offsets[i]; }
			pos += c.offsets[c.n_reads] - first;
		}
		h->read_off[n_reads] = pos;
	}
	if (int rc = ensure(h, h->d_ascii, n_bases + 64)) return rc;
	if (int rc = ensure(h, h->d_packed, n_bases / 4 + 64)) return rc;
	const uint64_t bad_cap = std::max<uint64_t>(1 << 20, n_bases / 16);
	if (int rc = ensure(h, h->d_nonacgt, bad_cap * 8)) return rc;
	const std::vector<uint64_t> none;
	build_segments(h, none, [&](uint64_t r_) { return read_ptr[r_]; });
	const uint32_t n_seg = (uint32_t)h->seg_len.size();
	CU(h, cudaMemsetAsync(h->d_counters.p, 0, h->d_counters.cap, h->stream));
	if (int rc = upload_segments(h)) return rc;
	if (int rc = prepare_process(h)) return rc;
	tick(h, 0);
	tick(h, 3);
	CU(h, cudaStreamWaitEvent(h->copy_stream, h->ev[0], 0));
	// ---- phase A: the pieces of the batch and their routes.  A piece either crosses PCIe as ASCII (packed by K0 on the
	// device) or is packed to 2 bits by host worker threads first (hostpack.cpp) and crosses as a quarter of the bytes.
	// Both resources work at once: on the bench box 16 cores pack 83 GB/s of ASCII and PCIe carries 55 GB/s of it, so about
	// 7 pieces in 10 go through the host.  H2D copies are queued in piece order, a few pieces ahead of the kernels.
	uint64_t PIECE = h->prm.piece_bytes ? std::max<uint64_t>(h->prm.piece_bytes, 64) : (48ull << 20);
	if (const char* e = getenv("MBGFE_PIECE_MB")) if (!h->prm.piece_bytes && atoi(e) > 0) PIECE = (uint64_t)atoi(e) << 20;
	struct Piece { const char* src; uint64_t pos, len; bool host; uint64_t stage_word; };
	std::vector<Piece> pieces;
	std::vector<uint64_t> piece_end;
	const unsigned hw = std::thread::hardware_concurrency();
	// one process per GPU: the ranks of a box share its cores
	const unsigned share = std::max(1u, h->prm.world);
	const unsigned n_workers = hw / share > 2 ? std::min(hw / share - 1, 31u) : 0;
	// share H of the bytes through the host so that both routes finish together: H / (r T) = (1 - 3/4 H) / P with r = 5.5 GB/s
	// of ASCII packed per worker, T workers, P = 55 GB/s of PCIe (measured on the bench box; 15 workers -> 65 %, 7 -> 46 %, 3 -> 24 %)
	uint32_t host_pct = (uint32_t)std::min(65.0, 100.0 * 5.5 * n_workers / (55.0 + 0.75 * 5.5 * n_workers));
	if (const char* e = getenv("MBGFE_HOST_PACK")) host_pct = (uint32_t)std::min(100, std::max(0, atoi(e)));
	if (n_workers == 0) host_pct = 0;
	uint64_t stage_words = 0;
	{
		uint64_t pos = 0, credit = 0;
		for (const HostChunk& c : h->pending)
		{
			const uint64_t first = c.offsets[0], bytes = c.offsets[c.n_reads] - first;
			for (uint64_t off = 0; off < bytes; off += PIECE)
			{
				const uint64_t len = std::min(PIECE, bytes - off);
				Piece pc { c.bases + first + off, pos, len, false, 0 };
				// the host route needs word-aligned pieces (only the very last one may end inside a word); the first piece
				// always goes as ASCII so that the device starts at once
				const bool eligible = pos % 32 == 0 && (len % 32 == 0 || pos + len == n_bases) && !pieces.empty();
				credit += host_pct;
				if (eligible && credit >= 100) { credit -= 100; pc.host = true; pc.stage_word = stage_words; stage_words += (len + 31) / 32; }
				else if (credit >= 100) credit = 100;
				pieces.push_back(pc);
				pos += len;
				piece_end.push_back(pos);
			}
		}
	}
	mbgfe_host::PackPool pool;
	if (stage_words)
	{
		if (int rc = ensure_host(h, h->p_stage, stage_words * 8)) return rc;
		std::vector<mbgfe_host::PackTask> tasks;
		const uint64_t SUB = 1ull << 20;                                   // characters per task (a multiple of 32)
		for (size_t i = 0; i < pieces.size(); i++) if (pieces[i].host)
			for (uint64_t o = 0; o < pieces[i].len; o += SUB)
				tasks.push_back({ i, pieces[i].src + o, std::min(SUB, pieces[i].len - o), (uint64_t*)h->p_stage.p + pieces[i].stage_word + o / 32, pieces[i].pos + o });
		pool.start(std::move(tasks), pieces.size(), n_workers);
	}
	while (h->piece_events.size() < pieces.size())
	{
		cudaEvent_t e;
		CU(h, cudaEventCreate(&e)); h->piece_events.push_back(e);
		CU(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->front_events.push_back(e);
		CU(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->pack_events.push_back(e);
	}
	size_t queued = 0;
	uint64_t h2d_bytes = 0;
	auto queue_upto = [&](size_t upto) -> int
	{
		for (; queued < std::min(upto, pieces.size()); queued++)
		{
			const Piece& pc = pieces[queued];
			if (pc.host)
			{
				pool.wait_piece(queued);
				const uint64_t bytes = (pc.len + 31) / 32 * 8;
				CU(h, cudaMemcpyAsync(h->d_packed.as<char>() + pc.pos / 4, (const char*)h->p_stage.p + pc.stage_word * 8, bytes, cudaMemcpyHostToDevice, h->copy_stream));
				h2d_bytes += bytes;
			}
			else
			{
				CU(h, cudaMemcpyAsync(h->d_ascii.as<char>() + pc.pos, pc.src, pc.len, cudaMemcpyHostToDevice, h->copy_stream));
				h2d_bytes += pc.len;
			}
			CU(h, cudaEventRecord(h->piece_events[queued], h->copy_stream));
			if (queued + 1 == pieces.size()) CU(h, cudaEventRecord(h->ev[1], h->copy_stream));
		}
		return 0;
	};
	constexpr size_t LOOKAHEAD = 6;
	if (pieces.empty()) CU(h, cudaEventRecord(h->ev[1], h->copy_stream));
	// ---- phase B: behind each piece, K0 on its bytes and K1 .. K4 on the segments it completes.  The instance lists and
	// the tables are sized from the first piece that completes segments (selected k-mers per base, extrapolated); the
	// kernels refuse to write past those sizes and count what they refused, and the exact path reruns the tail then.
	const bool want_table = !(h->prm.flags & MBGFE_NO_TABLE) && !h->appending;
	unsigned long long* ctr = h->d_counters.as<unsigned long long>();
	uint64_t packed_upto = 0;
	uint32_t g_done = 0, g_tail = 0, launches = 0;
	bool sized = false;
	for (size_t piece = 0; piece < piece_end.size(); piece++)
	{
		// the copies run a few pieces ahead; at the start only as far as the host packers have got, so that the device
		// is not kept waiting for them
		if (int rc = queue_upto(piece + (piece < 2 ? 1 : LOOKAHEAD))) return rc;
		// K0 .. K3a of consecutive pieces alternate between two streams: the tail of one piece's kernels overlaps the
		// head of the next piece's (every kernel here is a few waves of CTAs).  A piece's segments may begin in the piece
		// before it, so its stream also waits for that piece's bases (landed packed, or packed by K0 on the other stream).
		const cudaStream_t fs = h->front_stream[piece & 1];
		CU(h, cudaStreamWaitEvent(fs, h->piece_events[piece], 0));
		if (piece) { CU(h, cudaStreamWaitEvent(fs, h->piece_events[piece - 1], 0)); CU(h, cudaStreamWaitEvent(fs, h->pack_events[piece - 1], 0)); }
		const uint64_t pos = piece_end[piece];
		const bool last = pos == n_bases;
		const uint64_t pack_to = last ? n_bases : (pos & ~(uint64_t)31);
		if (pieces[piece].host) packed_upto = pack_to;            // arrived packed: nothing for K0 to do
		else if (pack_to > packed_upto)
		{
			const uint64_t c0 = packed_upto / 32, c1 = (pack_to + 31) / 32;
			const unsigned grid = (unsigned)std::min<uint64_t>((c1 - c0 + 255) / 256, (uint64_t)h->sm_count * 16);
			MBGFE_LAUNCH(k_pack_ascii, grid, 256, 0, fs, (const uint4*)h->d_ascii.p, n_bases, c0, c1, h->d_packed.as<uint64_t>(), h->d_nonacgt.as<uint64_t>(),
				h->d_counters.as<unsigned long long>() + CTR_NONACGT, bad_cap);
			launches++;
			packed_upto = pack_to;
		}
		CU(h, cudaEventRecord(h->pack_events[piece], fs));
		uint32_t g_end = g_done;
		while (g_end < n_seg && read_end_of(h, h->seg_read[g_end]) <= packed_upto) g_end++;
		if (g_end == g_done) { CU(h, cudaStreamWaitEvent(h->stream, h->pack_events[piece], 0)); continue; }
		launches += launch_front(h, g_done, g_end - g_done, false, fs);
		CU(h, cudaEventRecord(h->front_events[piece], fs));
		CU(h, cudaStreamWaitEvent(h->stream, h->front_events[piece], 0));
		launch_scan(h, g_done, g_end, launches);
		g_done = g_end;
		if (!sized)
		{
			uint64_t n0 = 0;
			CU(h, cudaMemcpyAsync(&n0, h->d_sel_begin.as<uint64_t>() + g_end, 8, cudaMemcpyDeviceToHost, h->stream));
			CU(h, cudaStreamSynchronize(h->stream));
			h->stats.d2h_bytes += 8;
			const uint64_t bases0 = read_end_of(h, h->seg_read[g_end - 1]);
			const double per_base = (double)n0 / (double)std::max<uint64_t>(bases0, 1);
			n0 -= std::min<uint64_t>(n0, h->appending ? h->append_base : 0);
			const double per_base2 = (double)n0 / (double)std::max<uint64_t>(bases0, 1);
			const uint64_t est = (uint64_t)(per_base2 * (double)n_bases) + 1;
			(void)per_base;
			if (int rc = ensure_instances(h, std::max<uint64_t>(est + est / 4 + 4096 + (h->appending ? h->append_base : 0), h->inst_cap))) return rc;
			if (want_table)
			{
				const uint64_t slots = std::max<uint64_t>(next_pow2(std::max<uint64_t>(2 * est, 1024)), h->prm.table_slots ? next_pow2(h->prm.table_slots) : 0);
				if (int rc = clear_tables(h, slots, slots, launches)) return rc;
				h->stats.table_slots = slots;
			}
			sized = true;
		}
		launch_list_hash(h, g_tail, g_end, 0, launches);
		launch_variants(h, g_tail, g_end, 0, launches);      // pass 3 keeps the pipelined upload
		launch_tables(h, g_tail, g_end, 0, launches, false);
		g_tail = g_end;
	}
	tick(h, 9);
	// ---- one read-back decides whether the pipelined result stands
	unsigned long long c8[8] = { 0 };
	uint64_t n_inst = 0;
	CU(h, cudaMemcpyAsync(c8, ctr, sizeof(c8), cudaMemcpyDeviceToHost, h->stream));
	if (n_seg) CU(h, cudaMemcpyAsync(&n_inst, h->d_sel_begin.as<uint64_t>() + n_seg, 8, cudaMemcpyDeviceToHost, h->stream));
	CU(h, cudaStreamSynchronize(h->stream));
	CU(h, cudaGetLastError());
	h->stats.d2h_bytes += sizeof(c8) + 8;
	pool.finish();
	h->stats.h2d_bytes += h2d_bytes;
	const uint64_t n_bad_dev = c8[CTR_NONACGT];
	const uint64_t n_bad = n_bad_dev + pool.bad.size();
	h->stats.ms_upload = elapsed(h, 0, 1);
	h->stats.ms_pack = 0;
	h->stats.reads = n_reads; h->stats.raw_bases = n_bases; h->stats.nonacgt = n_bad;
	h->nonacgt_count = n_bad;
	h->pending.clear();
	if (n_bad)
	{
		// rare: cut the reads at the characters outside ACGT and run the resident batch again
		if (n_bad > bad_cap) return fail(h, MBGFE_E_LIMIT, "%llu characters outside ACGT in %llu bases: more than 1 in 16", (unsigned long long)n_bad, (unsigned long long)n_bases);
		std::vector<uint64_t> bad;
		if (int rc = download(h, bad, h->d_nonacgt, n_bad_dev)) return rc;
		bad.insert(bad.end(), pool.bad.begin(), pool.bad.end());      // found by the host packers
		std::sort(bad.begin(), bad.end());
		build_segments(h, bad, [&](uint64_t r_) { return read_ptr[r_]; });
		if (int rc = upload_segments(h)) return rc;
		CU(h, cudaStreamSynchronize(h->stream));
		h->stats.segments = h->seg_len.size();
		h->resident = true;
		return mbgfe_process(h);
	}
	h->stats.segments = n_seg;
	h->resident = true;
	if (!sized || g_tail != n_seg || c8[CTR_PALINDROME] || c8[CTR_OVERFLOW] || n_inst > h->inst_cap)
	{
		// estimates exceeded, or palindromic k-mers to resolve (MBG/ReadHelper.h:755-785): the exact tail on the resident results
		tick(h, 4); tick(h, 5);
		h->stats.table_grows += (c8[CTR_OVERFLOW] || n_inst > h->inst_cap) ? 1 : 0;
		return process_tail(h, launches);
	}
	h->n_inst = n_inst;
	h->stats.distinct = want_table ? c8[CTR_DISTINCT] : 0;
	h->stats.distinct_edges = (want_table && h->edge_slots) ? c8[CTR_EDGES] : 0;
	h->have_table = want_table;
	finish_stats(h, n_inst, launches, false);
	return 0;
}

}

namespace {

// deferred files -> parsed into memory the plain way (mbgfe_upload, variants, mixed input, files the streaming loader refuses)
int materialize_files(mbgfe* h)
{
	std::vector<mbgfe::DeferredFile> todo;
	todo.swap(h->files);
	for (auto& f : todo)
	{
		if (f.fd >= 0) close(f.fd);
		uint64_t have = 0;
		for (auto& c : h->pending) have += c.n_reads;
		if (h->names.size() < have) h->names.resize(have);
		if (int rc = load_file(h, f.path.c_str())) return rc;
	}
	return 0;
}

// Record structure of a FASTA file from the scan events of its tasks, in file order (MBG/fastqloader.h:38-80): a record is a
// '>' line and the sequence lines up to the next one; lines before the first header are skipped; a trailing '\r' is dropped;
// empty lines are skipped.  The streaming loader only takes records whose sequence sits on ONE line (a contiguous range of the
// file image); `multiline` reports the others.
struct FastaStitch
{
	uint64_t line_start = 0, seq_start = 0, seq_end = 0, name_start = 0, name_end = 0, pending_cr = ~0ull;
	bool in_header = false, have_record = false, have_seq = false, multiline = false;
	size_t bad_at_line = 0;
	std::vector<uint64_t>* bad = nullptr;                                  // positions of the characters outside ACGT, ascending ...
	std::vector<unsigned char>* bad_ch = nullptr;                          // ... and the characters themselves
	std::function<void(uint64_t, uint64_t, uint64_t, uint64_t)> emit;      // seq_start, seq_end, name_start, name_end

	void flush_cr() { if (pending_cr != ~0ull) { bad->push_back(pending_cr); bad_ch->push_back('\r'); pending_cr = ~0ull; } }
	void finish_record() { if (have_record && !have_seq) emit(name_end, name_end, name_start, name_end); have_record = false; have_seq = false; }
	void end_line(uint64_t pos)          // the current line is [line_start, pos)
	{
		if (in_header) { name_end = pos; in_header = false; have_record = true; have_seq = false; }
		else
		{
			uint64_t end = pos;
			if (pending_cr != ~0ull && pending_cr + 1 == pos) { end = pos - 1; pending_cr = ~0ull; }      // "\r\n"
			flush_cr();
			if (end > line_start)
			{
				if (!have_record) { bad->resize(bad_at_line); bad_ch->resize(bad_at_line); }      // text before the first header: not a read
				else if (have_seq) multiline = true;
				else { seq_start = line_start; seq_end = end; have_seq = true; emit(seq_start, seq_end, name_start, name_end); }
			}
		}
		line_start = pos + 1;
		bad_at_line = bad->size();
	}
	void event(uint64_t pos, unsigned char ch)
	{
		if (ch == '\n') { end_line(pos); return; }
		if (in_header) return;
		if (ch == '>' && pos == line_start) { flush_cr(); finish_record(); in_header = true; name_start = pos + 1; return; }
		flush_cr();
		if (ch == '\r') pending_cr = pos; else { bad->push_back(pos); bad_ch->push_back(ch); }
	}
	void end_of_file(uint64_t pos) { if (pos > line_start || in_header) end_line(pos); flush_cr(); finish_record(); }
};

// mbgfe_run on plain FASTA files: the file never exists as a parsed copy in host memory.  Worker threads pread() it in tasks
// of 1 MiB, pack EVERY byte to 2 bits in file order (hostpack.cpp: pack_scan_fasta) straight into pinned staging and report the
// handful of bytes per read that are not ACGT; this thread stitches those into reads (FastaStitch), appends the segment table,
// queues the H2D copy of the piece and launches K1 .. K4 on the segments the piece completed -- the same per-piece pipeline as
// the in-memory path.  The device image keeps file coordinates (header bytes are junk entries nobody reads), so parse == pack
// and the bus carries a quarter byte per file byte.  Returns 1 when a file needs the plain parser instead.
int run_files(mbgfe* h)
{
	clear_results(h);
	h->resident = false;
	h->stats.h2d_bytes = 0; h->stats.d2h_bytes = 0;
	close_files(h->image_files);
	h->names.clear(); h->name_off.clear(); h->name_len.clear(); h->name_file.clear();
	uint64_t total = 0;
	for (auto& f : h->files) { f.base = total; total += (f.size + 31) & ~(uint64_t)31; }
	const uint64_t n_image = total;
	uint64_t PIECE = h->prm.piece_bytes ? std::max<uint64_t>((h->prm.piece_bytes + 31) & ~31u, 64) : (32ull << 20);
	if (const char* e = getenv("MBGFE_PIECE_MB")) if (!h->prm.piece_bytes && atoi(e) > 0) PIECE = (uint64_t)atoi(e) << 20;
	const uint64_t TASK = std::min<uint64_t>(PIECE, 1ull << 20);
	// capacity of the segment table and of the per-symbol arrays: nobody knows the reads yet.  One segment per 256 file bytes
	// is far more than HiFi reads need; a file of shorter reads goes to the plain parser.
	const uint64_t seg_cap = std::max<uint64_t>(1 << 16, n_image / 256);
	const uint64_t out_cap = n_image + 64 * seg_cap;
	if (seg_cap >= (1ull << 32) - 2) return 1;
	{
		const size_t n = seg_cap;
		if (int rc = ensure(h, h->d_seg_in, (n + 1) * 8)) return rc;
		if (int rc = ensure(h, h->d_seg_out, (n + 1) * 8)) return rc;
		if (int rc = ensure(h, h->d_seg_len, (n + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_seg_c1, (n + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_seg_c2, (n + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_seg_nsel, (n + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_sel_begin, (n + 2) * 8)) return rc;
	}
	if (int rc = ensure(h, h->d_packed, n_image / 4 + 64)) return rc;
	if (int rc = ensure(h, h->d_nonacgt, 8 << 20)) return rc;
	if (int rc = ensure_host(h, h->p_stage, n_image / 4 + 64)) return rc;
	h->read_off.clear(); h->read_end.clear();
	h->seg_in.clear(); h->seg_out.clear(); h->seg_len.clear(); h->seg_read.clear(); h->seg_rawoff.clear(); h->seg_nonacgt_before.clear();
	h->total_out = out_cap;
	CU(h, cudaMemsetAsync(h->d_counters.p, 0, h->d_counters.cap, h->stream));
	if (int rc = prepare_process(h)) return rc;
	h->total_out = 0;
	tick(h, 0);
	tick(h, 3);
	CU(h, cudaStreamWaitEvent(h->copy_stream, h->ev[0], 0));
	// ---- pieces and tasks
	struct Piece { size_t file; uint64_t pos, len; size_t task0, task1; };
	std::vector<Piece> pieces;
	std::vector<mbgfe_host::FileTask> tasks;
	for (size_t fi = 0; fi < h->files.size(); fi++)
	{
		const auto& f = h->files[fi];
		for (uint64_t off = 0; off < f.size; off += PIECE)
		{
			Piece pc { fi, f.base + off, std::min(PIECE, f.size - off), tasks.size(), 0 };
			for (uint64_t o = 0; o < pc.len; o += TASK)
			{
				mbgfe_host::FileTask t;
				t.piece = pieces.size(); t.fd = f.fd; t.file_off = off + o; t.n = std::min(TASK, pc.len - o);
				t.dst = (uint64_t*)h->p_stage.p + (pc.pos + o) / 32; t.pos = pc.pos + o;
				tasks.push_back(std::move(t));
			}
			pc.task1 = tasks.size();
			pieces.push_back(pc);
		}
	}
	const unsigned hw = std::thread::hardware_concurrency();
	const unsigned share = std::max(1u, h->prm.world);
	unsigned n_workers = hw / share > 2 ? std::min(hw / share - 1, 63u) : 1;
	if (const char* e = getenv("MBGFE_FILE_THREADS")) if (atoi(e) > 0) n_workers = (unsigned)atoi(e);
	mbgfe_host::FilePool pool;
	pool.start(std::move(tasks), pieces.size(), n_workers);
	while (h->piece_events.size() < pieces.size())
	{
		cudaEvent_t e;
		CU(h, cudaEventCreate(&e)); h->piece_events.push_back(e);
		CU(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->front_events.push_back(e);
		CU(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->pack_events.push_back(e);
	}
	// ---- the stitcher: reads and segments as the pieces come in
	std::vector<uint64_t> bad;
	std::vector<unsigned char> bad_ch;
	uint64_t out = 0, raw_bases = 0;
	bool refuse = false;
	size_t cur_file = 0;
	FastaStitch st;
	st.bad = &bad; st.bad_ch = &bad_ch;
	const std::function<void(uint64_t, uint64_t, uint64_t, uint64_t)> emit_read = [&](uint64_t s0, uint64_t s1, uint64_t n0, uint64_t n1)
	{
		const uint64_t r = h->read_off.size();
		h->read_off.push_back(s0); h->read_end.push_back(s1);
		h->name_off.push_back(n0); h->name_len.push_back((uint32_t)(n1 - n0)); h->name_file.push_back((uint32_t)cur_file);
		raw_bases += s1 - s0;
		if (s1 - s0 >= (1ull << 29)) { refuse = true; return; }
		if (s1 - s0 < 32) return;                                     // MBG/ReadHelper.h:838
		auto add = [&](uint64_t start, uint64_t len, uint32_t nonacgt_before)
		{
			if (h->seg_len.size() >= seg_cap) { refuse = true; return; }
			h->seg_in.push_back(start); h->seg_len.push_back((uint32_t)len); h->seg_read.push_back((uint32_t)r); h->seg_rawoff.push_back((uint32_t)(start - s0));
			h->seg_nonacgt_before.push_back(nonacgt_before); h->seg_out.push_back(out);
			out += (len + 1 + 63) & ~(uint64_t)63;
		};
		// the characters outside ACGT of this line are the last entries of `bad` (the scan reported them before the line's end)
		size_t bi = st.bad_at_line;
		while (bi < bad.size() && bad[bi] < s0) bi++;
		if (bi >= bad.size() || bad[bi] >= s1) { add(s0, s1 - s0, 0); return; }
		// a read with such characters is cut at them right here, as build_segments does for a resident batch (MBG/ReadHelper.h:864-897):
		// the run never has to go back over the image.  A character next to an ACGT base always differs from it; inside a run of
		// other characters the scan's own bytes decide what the collapse step keeps.
		uint64_t cur = s0;
		uint32_t collapsed_other = 0;
		while (cur < s1)
		{
			const uint64_t stop = (bi < bad.size() && bad[bi] < s1) ? bad[bi] : s1;
			if (stop > cur) add(cur, stop - cur, collapsed_other);
			if (stop >= s1) break;
			uint64_t e = stop;
			while (bi < bad.size() && bad[bi] == e && e < s1)
			{
				unsigned char c = bad_ch[bi], pch = (e > stop) ? bad_ch[bi - 1] : 0;
				if (c >= 'a' && c <= 'z') c = (unsigned char)(c - 32);
				if (pch >= 'a' && pch <= 'z') pch = (unsigned char)(pch - 32);
				if (e == stop || c != pch) collapsed_other++;
				e++; bi++;
			}
			cur = e;
		}
	};
	st.emit = emit_read;
	const bool want_table = !(h->prm.flags & MBGFE_NO_TABLE) && !h->appending;
	unsigned long long* ctr = h->d_counters.as<unsigned long long>();
	uint32_t g_done = 0, g_tail = 0, launches = 0;
	uint64_t h2d_bytes = 0;
	bool sized = false;
	auto drain = [&]() { pool.finish(); cudaStreamSynchronize(h->copy_stream); cudaStreamSynchronize(h->front_stream[0]); cudaStreamSynchronize(h->front_stream[1]); cudaStreamSynchronize(h->stream); };
	for (size_t piece = 0; piece < pieces.size(); piece++)
	{
		const Piece& pc = pieces[piece];
		pool.wait_piece(piece);
		if (pc.file != cur_file) { cur_file = pc.file; st = FastaStitch(); st.bad = &bad; st.bad_ch = &bad_ch; st.bad_at_line = bad.size(); st.emit = emit_read; st.line_start = pc.pos; }
		if (piece == 0) st.line_start = pc.pos;
		const size_t seg_before = h->seg_len.size();
		for (size_t t = pc.task0; t < pc.task1; t++)
		{
			mbgfe_host::FileTask& k = pool.tasks[t];
			if (k.error) { drain(); return fail(h, MBGFE_E_IO, "read error in %s", h->files[pc.file].path.c_str()); }
			for (uint64_t ev : k.events) st.event(ev >> 8, (unsigned char)(ev & 0xff));
			std::vector<uint64_t>().swap(k.events);
		}
		const bool file_ends = piece + 1 == pieces.size() || pieces[piece + 1].file != pc.file;
		if (file_ends) st.end_of_file(h->files[pc.file].base + h->files[pc.file].size);
		if (st.multiline || refuse || out > out_cap) { drain(); return 1; }
		// ---- new rows of the segment table, then the piece itself, in order on the copy stream
		const size_t seg_now = h->seg_len.size();
		if (seg_now > seg_before)
		{
			const size_t n = seg_now - seg_before;
			CU(h, cudaMemcpyAsync(h->d_seg_in.as<uint64_t>() + seg_before, h->seg_in.data() + seg_before, n * 8, cudaMemcpyHostToDevice, h->copy_stream));
			CU(h, cudaMemcpyAsync(h->d_seg_out.as<uint64_t>() + seg_before, h->seg_out.data() + seg_before, n * 8, cudaMemcpyHostToDevice, h->copy_stream));
			CU(h, cudaMemcpyAsync(h->d_seg_out.as<uint64_t>() + seg_now, &out, 8, cudaMemcpyHostToDevice, h->copy_stream));
			CU(h, cudaMemcpyAsync(h->d_seg_len.as<uint32_t>() + seg_before, h->seg_len.data() + seg_before, n * 4, cudaMemcpyHostToDevice, h->copy_stream));
			h2d_bytes += n * 20 + 8;
		}
		{
			const uint64_t bytes = (pc.len + 31) / 32 * 8;
			CU(h, cudaMemcpyAsync(h->d_packed.as<char>() + pc.pos / 4, (const char*)h->p_stage.p + pc.pos / 4, bytes, cudaMemcpyHostToDevice, h->copy_stream));
			h2d_bytes += bytes;
		}
		CU(h, cudaEventRecord(h->piece_events[piece], h->copy_stream));
		if (piece + 1 == pieces.size()) CU(h, cudaEventRecord(h->ev[1], h->copy_stream));
		// ---- K1 .. K3a on the segments this piece completed (a read is only known once its line has ended: all of them)
		const cudaStream_t fs = h->front_stream[piece & 1];
		CU(h, cudaStreamWaitEvent(fs, h->piece_events[piece], 0));
		const uint32_t g_end = (uint32_t)seg_now;
		if (g_end == g_done) continue;
		launches += launch_front(h, g_done, g_end - g_done, false, fs);
		CU(h, cudaEventRecord(h->front_events[piece], fs));
		CU(h, cudaStreamWaitEvent(h->stream, h->front_events[piece], 0));
		launch_scan(h, g_done, g_end, launches);
		g_done = g_end;
		if (!sized)
		{
			uint64_t n0 = 0;
			CU(h, cudaMemcpyAsync(&n0, h->d_sel_begin.as<uint64_t>() + g_end, 8, cudaMemcpyDeviceToHost, h->stream));
			CU(h, cudaStreamSynchronize(h->stream));
			h->stats.d2h_bytes += 8;
			n0 -= std::min<uint64_t>(n0, h->appending ? h->append_base : 0);
			const double per_byte = (double)n0 / (double)std::max<uint64_t>(pc.pos + pc.len, 1);
			const uint64_t est = (uint64_t)(per_byte * (double)n_image) + 1;
			if (int rc = ensure_instances(h, std::max<uint64_t>(est + est / 4 + 4096 + (h->appending ? h->append_base : 0), h->inst_cap))) return rc;
			if (want_table)
			{
				const uint64_t slots = std::max<uint64_t>(next_pow2(std::max<uint64_t>(2 * est, 1024)), h->prm.table_slots ? next_pow2(h->prm.table_slots) : 0);
				if (int rc = clear_tables(h, slots, slots, launches)) return rc;
				h->stats.table_slots = slots;
			}
			sized = true;
		}
		launch_list_hash(h, g_tail, g_end, 0, launches);
		launch_variants(h, g_tail, g_end, 0, launches);      // pass 3 keeps the pipelined upload
		launch_tables(h, g_tail, g_end, 0, launches, false);
		g_tail = g_end;
	}
	if (pieces.empty()) CU(h, cudaEventRecord(h->ev[1], h->copy_stream));
	tick(h, 9);
	pool.finish();
	const uint32_t n_seg = (uint32_t)h->seg_len.size();
	h->total_out = out;
	h->n_reads = h->read_off.size(); h->n_bases = n_image;
	h->image_files.swap(h->files);
	h->files.clear();
	// ---- one read-back decides whether the pipelined result stands
	unsigned long long c8[8] = { 0 };
	uint64_t n_inst = 0;
	CU(h, cudaMemcpyAsync(c8, ctr, sizeof(c8), cudaMemcpyDeviceToHost, h->stream));
	if (n_seg) CU(h, cudaMemcpyAsync(&n_inst, h->d_sel_begin.as<uint64_t>() + n_seg, 8, cudaMemcpyDeviceToHost, h->stream));
	CU(h, cudaStreamSynchronize(h->stream));
	CU(h, cudaStreamSynchronize(h->copy_stream));
	CU(h, cudaGetLastError());
	h->stats.d2h_bytes += sizeof(c8) + 8;
	h->stats.h2d_bytes += h2d_bytes;
	h->stats.ms_upload = elapsed(h, 0, 1);
	h->stats.ms_pack = 0;
	h->stats.reads = h->n_reads; h->stats.raw_bases = raw_bases; h->stats.nonacgt = bad.size();
	h->nonacgt_count = bad.size();
	h->pending.clear();
	h->stats.segments = n_seg;
	h->resident = true;
	if ((n_seg && !sized) || g_tail != n_seg || c8[CTR_PALINDROME] || c8[CTR_OVERFLOW] || n_inst > h->inst_cap)
	{
		tick(h, 4); tick(h, 5);
		h->stats.table_grows += (c8[CTR_OVERFLOW] || n_inst > h->inst_cap) ? 1 : 0;
		return process_tail(h, launches);
	}
	if (!n_seg) { if (int rc = ensure_instances(h, std::max<uint64_t>(4096, h->inst_cap))) return rc; if (want_table) { if (int rc = clear_tables(h, 1024, 1024, launches)) return rc; h->stats.table_slots = 1024; CU(h, cudaStreamSynchronize(h->stream)); } }
	h->n_inst = n_inst;
	h->stats.distinct = want_table ? c8[CTR_DISTINCT] : 0;
	h->stats.distinct_edges = (want_table && h->edge_slots) ? c8[CTR_EDGES] : 0;
	h->have_table = want_table;
	finish_stats(h, n_inst, launches, false);
	return 0;
}

// part table: per-segment results -> reference-shaped parts (MBG/ReadHelper.h:81-88, 838-862, 976-977, 1058-1059)
int export_parts(mbgfe* h)
{
	if (h->exported_parts) return 0;
	const size_t n_seg = h->seg_len.size();
	if (int rc = download(h, h->h_seg_c1, h->d_seg_c1, n_seg)) return rc;
	if (int rc = download(h, h->h_seg_c2, h->d_seg_c2, n_seg)) return rc;
	if (int rc = download(h, h->h_seg_nsel, h->d_seg_nsel, n_seg)) return rc;
	if (int rc = download(h, h->h_sel_begin, h->d_sel_begin, n_seg + 1)) return rc;
	h->parts.clear();
	uint64_t code_cursor = 0, pose_cursor = 0, c1_total = 0, hashed = 0;
	uint32_t cur_read = 0xffffffffu, collapsed_before = 0;
	const uint32_t em = h->prm.error_masking;
	for (size_t g = 0; g < n_seg; g++)
	{
		if (h->seg_read[g] != cur_read) { cur_read = h->seg_read[g]; collapsed_before = 0; }
		const uint32_t c1 = h->h_seg_c1[g], c2 = h->h_seg_c2[g];
		c1_total += c1;
		uint32_t offset;
		if (h->collapse) offset = collapsed_before + h->seg_nonacgt_before[g];
		else if (em == MBGFE_MASK_NO) offset = h->seg_rawoff[g] == 0 ? 0 : h->seg_rawoff[g] - 1; // index of the last non-ACGT character in front (MBG/ReadHelper.h:1064)
		else offset = h->seg_rawoff[g];
		collapsed_before += c1;
		if (c2 == 0) continue; // microsatellite modes drop stretches with fewer than 32 symbols
		mbgfe_part p;
		p.read_idx = h->seg_read[g];
		p.part_offset = offset;
		p.raw_len = (uint32_t)(read_end_of(h, p.read_idx) - h->read_off[p.read_idx]);
		p.n_codes = c2;
		p.code_begin = code_cursor;
		p.pose_begin = pose_cursor;
		p.sel_begin = h->h_sel_begin[g];
		p.n_sel = h->h_seg_nsel[g];
		p.hashed = c2 >= h->prm.k ? 1 : 0;
		hashed += p.hashed;
		code_cursor += c2;
		pose_cursor += c2 + 1;
		h->parts.push_back(p);
	}
	h->stats.collapsed_symbols = c1_total;
	h->stats.codes = code_cursor;
	h->stats.parts = h->parts.size();
	h->stats.hashed_parts = hashed;
	h->exported_parts = true;
	return 0;
}

int export_codes(mbgfe* h)
{
	if (h->exported_codes) return 0;
	if (int rc = export_parts(h)) return rc;
	const bool want_poses = (h->prm.flags & MBGFE_KEEP_POSES) != 0;
	std::vector<uint16_t> all;
	std::vector<uint32_t> allp;
	if (int rc = download(h, all, h->d_codes, h->total_out)) return rc;
	if (want_poses) if (int rc = download(h, allp, h->d_pos2, h->total_out)) return rc;
	h->h_codes.resize(h->stats.codes);
	h->h_poses.resize(want_poses ? h->stats.codes + h->parts.size() : 0);
	size_t pi = 0;
	for (size_t g = 0; g < h->seg_len.size(); g++)
	{
		const uint32_t c2 = h->h_seg_c2[g];
		if (c2 == 0) continue;
		const mbgfe_part& p = h->parts[pi++];
		memcpy(h->h_codes.data() + p.code_begin, all.data() + h->seg_out[g], (size_t)c2 * 2);
		if (want_poses)
		{
			// device poses are relative to the stretch; the reference reports positions in the (collapsed) read
			const uint32_t base = h->collapse ? p.part_offset : h->seg_rawoff[g];
			for (uint32_t i = 0; i <= c2; i++) h->h_poses[p.pose_begin + i] = allp[h->seg_out[g] + i] + base;
		}
	}
	h->exported_codes = true;
	return 0;
}

int export_instances(mbgfe* h)
{
	if (h->h_positions.size() == h->n_inst && h->h_hashes.size() == h->n_inst) return 0;
	if (int rc = download(h, h->h_positions, h->d_inst_pos, h->n_inst)) return rc;
	std::vector<uint64_t> raw;
	if (int rc = download(h, raw, h->d_inst_hash, 2 * h->n_inst)) return rc;
	h->h_hashes.resize(h->n_inst);
	if (h->n_inst) memcpy(h->h_hashes.data(), raw.data(), h->n_inst * 16);
	return 0;
}

// exclusive prefix sum of n uint32 -> n+1 uint64 on the device (two-level)
int device_scan(mbgfe* h, const uint32_t* in, uint64_t* out, uint32_t n)
{
	const uint32_t tile = SCAN_NT * SCAN_IPT;
	const uint32_t n_tiles = std::max<uint32_t>(1, (n + tile - 1) / tile);
	if (int rc = ensure(h, h->d_scan_tot, (size_t)(n_tiles + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_scan_off, (size_t)(n_tiles + 2) * 8)) return rc;
	MBGFE_LAUNCH(k_scan_tile_totals, n_tiles, SCAN_NT, 0, h->stream, in, h->d_scan_tot.as<uint32_t>(), n);
	MBGFE_LAUNCH(k_scan_u32<1024>, 1, 1024, 0, h->stream, h->d_scan_tot.as<uint32_t>(), h->d_scan_off.as<uint64_t>(), n_tiles, (const uint64_t*)nullptr);
	MBGFE_LAUNCH(k_scan_tiles, n_tiles, SCAN_NT, 0, h->stream, in, h->d_scan_off.as<uint64_t>(), out, n);
	return 0;
}

// HashList::sortByHash (MBG/HashList.cpp:245-353): ids = ranks of the canonical hashes; edges renumbered with those
// ids (already canonical: the device compared the hashes themselves), listed per node forward side first.
// Everything is ordered on the device (kernels.cuh "export"); the host only receives the finished arrays by DMA.
// node ids: device-side bucket + rank sort of the table; leaves the sorted node arrays in HBM and every entry's rank
// in its table slot.  Idempotent until the table changes.
int rank_nodes(mbgfe* h)
{
	if (h->nodes_ranked) return 0;
	const uint64_t n_nodes = h->stats.distinct;
	const unsigned grid = (unsigned)std::min<uint64_t>((h->node_slots + 255) / 256, (uint64_t)h->sm_count * 16);
	// buckets of ~32 entries on the top bits of the canonical hash
	uint32_t bits = 0;
	while (bits < 26 && (n_nodes >> bits) > 32) bits++;
	const uint32_t n_buckets = 1u << bits;
	const uint32_t shift = 64 - bits;
	if (int rc = ensure(h, h->d_bucket_count, (size_t)n_buckets * 4 + 64)) return rc;
	if (int rc = ensure(h, h->d_bucket_begin, (size_t)(n_buckets + 2) * 8)) return rc;
	if (int rc = ensure(h, h->d_node_out, (n_nodes + 1) * sizeof(NodeOut))) return rc;
	if (int rc = ensure(h, h->d_node_hash_sorted, (n_nodes + 1) * 16)) return rc;
	if (int rc = ensure(h, h->d_node_cov, (n_nodes + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_node_tip, n_nodes + 64)) return rc;
	const unsigned rank_grid = (unsigned)std::min<uint64_t>(((uint64_t)n_buckets * 32 + 255) / 256, (uint64_t)h->sm_count * 32);
	CU(h, cudaMemsetAsync(h->d_bucket_count.p, 0, (size_t)n_buckets * 4, h->stream));
	MBGFE_LAUNCH(k_node_bucket_count, grid, 256, 0, h->stream, h->d_nodes.as<NodeSlot>(), h->node_slots, shift, h->table_sharded ? h->prm.world : 1u, h->d_bucket_count.as<uint32_t>());
	if (int rc = device_scan(h, h->d_bucket_count.as<uint32_t>(), h->d_bucket_begin.as<uint64_t>(), n_buckets)) return rc;
	CU(h, cudaMemsetAsync(h->d_bucket_count.p, 0, (size_t)n_buckets * 4, h->stream));
	MBGFE_LAUNCH(k_node_bucket_scatter, grid, 256, 0, h->stream, h->d_nodes.as<NodeSlot>(), h->node_slots, shift, h->table_sharded ? h->prm.world : 1u, h->d_bucket_begin.as<uint64_t>(),
		h->d_bucket_count.as<uint32_t>(), h->d_node_out.as<NodeOut>());
	MBGFE_LAUNCH(k_node_bucket_rank, rank_grid, 256, 0, h->stream, h->d_node_out.as<NodeOut>(), h->d_bucket_begin.as<uint64_t>(), n_buckets,
		h->d_node_hash_sorted.as<uint64_t>(), h->d_node_cov.as<uint32_t>(), h->d_node_tip.as<uint8_t>(), h->d_nodes.as<NodeSlot>());
	h->nodes_ranked = true;
	return 0;
}

int export_table(mbgfe* h)
{
	if (h->exported_table) return 0;
	if (!h->have_table) return fail(h, MBGFE_E_STATE, "table requested but none was built (MBGFE_NO_TABLE without mbgfe_exchange_insert)");
	tick(h, 10);
	const uint64_t n_nodes = h->stats.distinct, n_edges = h->edge_slots ? h->stats.distinct_edges : 0;
	if (int rc = rank_nodes(h)) return rc;
	if (int rc = ensure_host(h, h->p_node_hashes, (n_nodes + 1) * 16)) return rc;
	if (int rc = ensure_host(h, h->p_node_cov, (n_nodes + 1) * 4)) return rc;
	if (int rc = ensure_host(h, h->p_node_tip, n_nodes + 64)) return rc;
	if (n_nodes)
	{
		// the node arrays leave on the copy stream while the edges are still being ordered on the compute stream
		CU(h, cudaEventRecord(h->ev[2], h->stream));
		CU(h, cudaStreamWaitEvent(h->copy_stream, h->ev[2], 0));
		CU(h, cudaMemcpyAsync(h->p_node_hashes.p, h->d_node_hash_sorted.p, n_nodes * 16, cudaMemcpyDeviceToHost, h->copy_stream));
		CU(h, cudaMemcpyAsync(h->p_node_cov.p, h->d_node_cov.p, n_nodes * 4, cudaMemcpyDeviceToHost, h->copy_stream));
		CU(h, cudaMemcpyAsync(h->p_node_tip.p, h->d_node_tip.p, n_nodes, cudaMemcpyDeviceToHost, h->copy_stream));
		CU(h, cudaEventRecord(h->ev[2], h->copy_stream));
		h->stats.d2h_bytes += n_nodes * 21;
	}
	// ---- edges: node slots -> node ids, buckets of ~32 edges on `from`
	if (n_edges)
	{
		// (a sharded partition holds edges whose `from` may belong to another rank: its buckets span the global id range)
		const uint64_t id_span = h->edges_global && !h->id_begin.empty() ? h->id_begin.back() : n_nodes;
		const uint32_t id_first = h->edges_global ? 0u : h->id_base;
		uint32_t eshift = 0;
		while (eshift < 31 && (id_span >> eshift) > std::max<uint64_t>(1, n_edges / 32)) eshift++;
		const uint32_t n_eb = (uint32_t)(id_span >> eshift) + 1;
		if (int rc = ensure(h, h->d_bucket_count, (size_t)n_eb * 4 + 64)) return rc;
		if (int rc = ensure(h, h->d_bucket_begin, (size_t)(n_eb + 2) * 8)) return rc;
		if (int rc = ensure(h, h->d_edge_out, (n_edges + 1) * sizeof(EdgeRec))) return rc;
		if (int rc = ensure(h, h->d_edges_sorted, (n_edges + 1) * sizeof(EdgeRec))) return rc;
		if (int rc = ensure_host(h, h->p_edges, (n_edges + 1) * sizeof(mbgfe_edge))) return rc;
		const unsigned egrid = (unsigned)std::min<uint64_t>((h->edge_slots + 255) / 256, (uint64_t)h->sm_count * 16);
		const unsigned erank_grid = (unsigned)std::min<uint64_t>(((uint64_t)n_eb * 32 + 255) / 256, (uint64_t)h->sm_count * 32);
		CU(h, cudaMemsetAsync(h->d_bucket_count.p, 0, (size_t)n_eb * 4, h->stream));
		const NodeSlot* slot_names = h->edges_global ? nullptr : h->d_nodes.as<NodeSlot>();     // sharded table: edge keys already hold global ids
		MBGFE_LAUNCH(k_edge_bucket_count, egrid, 256, 0, h->stream, h->d_edges.as<EdgeSlot>(), h->edge_slots, slot_names, eshift, id_first, h->d_bucket_count.as<uint32_t>());
		if (int rc = device_scan(h, h->d_bucket_count.as<uint32_t>(), h->d_bucket_begin.as<uint64_t>(), n_eb)) return rc;
		CU(h, cudaMemsetAsync(h->d_bucket_count.p, 0, (size_t)n_eb * 4, h->stream));
		MBGFE_LAUNCH(k_edge_bucket_scatter, egrid, 256, 0, h->stream, h->d_edges.as<EdgeSlot>(), h->edge_slots, slot_names, eshift, id_first,
			h->d_bucket_begin.as<uint64_t>(), h->d_bucket_count.as<uint32_t>(), h->d_edge_out.as<EdgeRec>());
		MBGFE_LAUNCH(k_edge_bucket_rank, erank_grid, 256, 0, h->stream, h->d_edge_out.as<EdgeRec>(), h->d_bucket_begin.as<uint64_t>(), n_eb, h->d_edges_sorted.as<EdgeRec>());
		CU(h, cudaMemcpyAsync(h->p_edges.p, h->d_edges_sorted.p, n_edges * sizeof(mbgfe_edge), cudaMemcpyDeviceToHost, h->stream));
		h->stats.d2h_bytes += n_edges * sizeof(mbgfe_edge);
	}
	if (n_nodes) CU(h, cudaStreamWaitEvent(h->stream, h->ev[2], 0));
	tick(h, 11);
	CU(h, cudaStreamSynchronize(h->stream));
	CU(h, cudaGetLastError());
	h->n_nodes_out = n_nodes; h->n_edges_out = n_edges;
	h->stats.ms_export = elapsed(h, 10, 11);
	h->exported_table = true;
	return 0;
}


// ---- the step after the table: abundance filter + unitig chains (graph.cuh)

unsigned grid_for(const mbgfe* h, uint64_t n) { return (unsigned)std::max<uint64_t>(1, std::min<uint64_t>((n + 255) / 256, (uint64_t)h->sm_count * 16)); }

// degrees -> links -> pointer doubling; leaves the final states in d_g_st[*which].  cov/edges describe the graph to rank.
int rank_chains(mbgfe* h, const EdgeRec* edges, uint64_t n_edges, const uint32_t* cov, uint64_t n_nodes, uint32_t min_cov, int* which, uint32_t& launches)
{
	const uint64_t n2 = 2 * n_nodes;
	if (int rc = ensure(h, h->d_g_deg, n2 * 4 + 64)) return rc;
	if (int rc = ensure(h, h->d_g_nb, n2 * 4 + 64)) return rc;
	if (int rc = ensure(h, h->d_g_link, n2 * 4 + 64)) return rc;
	for (int i = 0; i < 2; i++) if (int rc = ensure(h, h->d_g_st[i], (n2 + 1) * sizeof(ChainState))) return rc;
	if (int rc = ensure(h, h->d_g_flag_byte, 64)) return rc;
	CU(h, cudaMemsetAsync(h->d_g_deg.p, 0, n2 * 4 + 64, h->stream));
	if (n_edges) { MBGFE_LAUNCH(k_graph_degrees, grid_for(h, n_edges), 256, 0, h->stream, edges, n_edges, min_cov, h->d_g_deg.as<uint32_t>(), h->d_g_nb.as<uint32_t>()); launches++; }
	MBGFE_LAUNCH(k_graph_links, grid_for(h, n2), 256, 0, h->stream, h->d_g_deg.as<uint32_t>(), h->d_g_nb.as<uint32_t>(), n2, h->d_g_link.as<uint32_t>());
	launches++;
	uint32_t* flag = h->d_g_flag_byte.as<uint32_t>();
	for (int attempt = 0; attempt < 2; attempt++)
	{
		int cur = 0;
		MBGFE_LAUNCH(k_chain_init, grid_for(h, n2), 256, 0, h->stream, h->d_g_link.as<uint32_t>(), cov, n2, h->d_g_st[0].as<ChainState>());
		launches++;
		uint32_t open = 1;
		// a chain has at most n2 states, so ceil(log2 n2) + 1 doublings finish every open one; what is left then is circular.
		// The flag comes back every third round only (a host synchronisation costs more than a round on a short list).
		int limit = 2;
		while ((1ull << limit) < n2) limit++;
		limit += 1;
		for (int round = 0; round < limit && open; round++)
		{
			CU(h, cudaMemsetAsync(flag, 0, 4, h->stream));
			MBGFE_LAUNCH(k_chain_jump, grid_for(h, n2), 256, 0, h->stream, h->d_g_st[cur].as<ChainState>(), h->d_g_st[cur ^ 1].as<ChainState>(), n2, flag);
			launches++;
			cur ^= 1;
			if (round % 3 == 2 || round + 1 == limit)
			{
				CU(h, cudaMemcpyAsync(&open, flag, 4, cudaMemcpyDeviceToHost, h->stream));
				CU(h, cudaStreamSynchronize(h->stream));
			}
		}
		*which = cur;
		if (!open) return 0;
		// what is still open after that many doublings is circular: cut every cycle in front of its smallest node, rank again
		if (attempt == 1) return fail(h, MBGFE_E_STATE, "chain ranking did not terminate");
		MBGFE_LAUNCH(k_chain_cut, grid_for(h, n2), 256, 0, h->stream, h->d_g_st[cur].as<ChainState>(), n2, h->d_g_link.as<uint32_t>());
		launches++;
	}
	return 0;
}

template <typename T>
int fetch(mbgfe* h, std::vector<T>& dst, const DevBuf& src, size_t count)
{
	dst.resize(count);
	if (count) CU(h, cudaMemcpyAsync(dst.data(), src.p, count * sizeof(T), cudaMemcpyDeviceToHost, h->stream));
	h->stats.d2h_bytes += count * sizeof(T);
	return 0;
}


// a HashList on the device while the graph stage filters it: the exported table, or one of the two scratch sets
struct GList
{
	const uint4* hash; const uint32_t* cov; const uint8_t* tip; const EdgeRec* edges; const uint32_t* orig;     // orig == nullptr: the exported table itself
	uint64_t n, m;
	int set;      // -1 the exported table, 0 the d_f_* buffers, 1 the d_t_* buffers
};

// HashList::filter (MBG/HashList.cpp:159-232): d_g_kept holds the 0/1 flags and d_g_newid their exclusive scan (total at [n]);
// writes the survivors into the scratch set `in` does not live in
int filter_list(mbgfe* h, GList& l, uint64_t n_kept, uint32_t& launches)
{
	const int to = l.set == 0 ? 1 : 0;
	DevBuf& o_hash = to ? h->d_t_hash : h->d_f_hash; DevBuf& o_cov = to ? h->d_t_cov : h->d_f_cov; DevBuf& o_tip = to ? h->d_t_tip : h->d_f_tip;
	DevBuf& o_edges = to ? h->d_t_edges : h->d_f_edges; DevBuf& o_orig = to ? h->d_t_orig : h->d_f_orig;
	if (int rc = ensure(h, o_hash, (n_kept + 1) * 16)) return rc;
	if (int rc = ensure(h, o_cov, (n_kept + 1) * 4)) return rc;
	if (int rc = ensure(h, o_tip, n_kept + 64)) return rc;
	if (int rc = ensure(h, o_orig, (n_kept + 1) * 4)) return rc;
	if (int rc = ensure(h, o_edges, (l.m + 1) * sizeof(EdgeRec))) return rc;
	MBGFE_LAUNCH(k_filter_nodes, grid_for(h, l.n), 256, 0, h->stream, h->d_g_kept.as<uint32_t>(), h->d_g_newid.as<uint64_t>(), l.n, l.hash, l.cov, l.tip, l.orig,
		o_hash.as<uint4>(), o_cov.as<uint32_t>(), o_tip.as<uint8_t>(), o_orig.as<uint32_t>());
	launches++;
	uint64_t m_kept = 0;
	if (l.m)
	{
		MBGFE_LAUNCH(k_filter_edge_flags, grid_for(h, l.m), 256, 0, h->stream, l.edges, l.m, h->d_g_kept.as<uint32_t>(), h->d_g_flag.as<uint32_t>());
		if (int rc = device_scan(h, h->d_g_flag.as<uint32_t>(), h->d_g_at.as<uint64_t>(), (uint32_t)l.m)) return rc;
		MBGFE_LAUNCH(k_filter_edges, grid_for(h, l.m), 256, 0, h->stream, l.edges, l.m, h->d_g_flag.as<uint32_t>(), h->d_g_at.as<uint64_t>(), h->d_g_newid.as<uint64_t>(), o_edges.as<EdgeRec>());
		launches += 5;
		CU(h, cudaMemcpyAsync(&m_kept, h->d_g_at.as<uint64_t>() + l.m, 8, cudaMemcpyDeviceToHost, h->stream));
		CU(h, cudaStreamSynchronize(h->stream));
	}
	l.hash = o_hash.as<uint4>(); l.cov = o_cov.as<uint32_t>(); l.tip = o_tip.as<uint8_t>(); l.edges = o_edges.as<EdgeRec>(); l.orig = o_orig.as<uint32_t>();
	l.n = n_kept; l.m = m_kept; l.set = to;
	return 0;
}

// scan of the kept flags in d_g_kept; returns how many survive
int scan_kept(mbgfe* h, uint64_t n, uint64_t& n_kept, uint32_t& launches)
{
	n_kept = 0;
	if (!n) return 0;
	if (int rc = device_scan(h, h->d_g_kept.as<uint32_t>(), h->d_g_newid.as<uint64_t>(), (uint32_t)n)) return rc;
	launches += 3;
	CU(h, cudaMemcpyAsync(&n_kept, h->d_g_newid.as<uint64_t>() + n, 8, cudaMemcpyDeviceToHost, h->stream));
	CU(h, cudaStreamSynchronize(h->stream));
	return 0;
}

// removeOnecovNodes (MBG/MBG.cpp:548-627), one pass: kept flags -> filter when anything goes
int onecov_pass(mbgfe* h, GList& l, bool only_tips, uint32_t& launches)
{
	if (!l.n) return 0;
	const uint64_t n2 = 2 * l.n;
	if (n2 >= 0x80000000ull) return fail(h, MBGFE_E_PARAM, "oneCovHeuristic: more than 2^30 k-mers in one table");
	if (int rc = ensure(h, h->d_g_deg, n2 * 4 + 64)) return rc;
	if (int rc = ensure(h, h->d_g_nb, n2 * 4 + 64)) return rc;
	if (int rc = ensure(h, h->d_g_link, n2 * 4 + 64)) return rc;
	for (int i = 0; i < 2; i++) if (int rc = ensure(h, h->d_g_st[i], (n2 + 1) * sizeof(ChainState))) return rc;
	if (int rc = ensure(h, h->d_g_flag_byte, 64)) return rc;
	uint32_t* run_end = h->d_g_st[0].as<uint32_t>();
	uint32_t* strong = h->d_g_st[1].as<uint32_t>();
	uint32_t* bad = strong + n2;
	CU(h, cudaMemsetAsync(h->d_g_deg.p, 0, n2 * 4 + 64, h->stream));
	CU(h, cudaMemsetAsync(strong, 0, 2 * n2 * 4, h->stream));
	if (l.m)
	{
		MBGFE_LAUNCH(k_graph_degrees, grid_for(h, l.m), 256, 0, h->stream, l.edges, l.m, 1u, h->d_g_deg.as<uint32_t>(), h->d_g_nb.as<uint32_t>());
		MBGFE_LAUNCH(k_onecov_strong, grid_for(h, l.m), 256, 0, h->stream, l.edges, l.m, l.cov, strong);
		MBGFE_LAUNCH(k_onecov_bad, grid_for(h, l.m), 256, 0, h->stream, l.edges, l.m, strong, bad);
		launches += 3;
	}
	MBGFE_LAUNCH(k_graph_links, grid_for(h, n2), 256, 0, h->stream, h->d_g_deg.as<uint32_t>(), h->d_g_nb.as<uint32_t>(), n2, h->d_g_link.as<uint32_t>());
	MBGFE_LAUNCH(k_onecov_init, grid_for(h, n2), 256, 0, h->stream, h->d_g_link.as<uint32_t>(), l.cov, n2, run_end);
	launches += 2;
	uint32_t* flag = h->d_g_flag_byte.as<uint32_t>();
	uint32_t open = 1;
	int limit = 2;
	while ((1ull << limit) < n2) limit++;
	for (int round = 0; round <= limit && open; round++)      // what is still open after that is a circular run
	{
		CU(h, cudaMemsetAsync(flag, 0, 4, h->stream));
		MBGFE_LAUNCH(k_onecov_jump, grid_for(h, n2), 256, 0, h->stream, run_end, n2, flag);
		launches++;
		CU(h, cudaMemcpyAsync(&open, flag, 4, cudaMemcpyDeviceToHost, h->stream));
		CU(h, cudaStreamSynchronize(h->stream));
	}
	if (int rc = ensure(h, h->d_g_kept, (l.n + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_g_newid, (l.n + 2) * 8)) return rc;
	MBGFE_LAUNCH(k_onecov_verdict, grid_for(h, l.n), 256, 0, h->stream, run_end, h->d_g_deg.as<uint32_t>(), bad, l.cov, l.n, only_tips ? 1u : 0u, h->d_g_kept.as<uint32_t>());
	launches++;
	uint64_t n_kept = 0;
	if (int rc = scan_kept(h, l.n, n_kept, launches)) return rc;
	if (n_kept == l.n) return 0;                                // removedAny == false: the list stays as it is
	return filter_list(h, l, n_kept, launches);
}

}

extern "C" {

int mbgfe_build_unitigs(mbgfe* h, uint32_t kmer_abundance, double unitig_abundance) { return mbgfe_build_unitigs_ex(h, kmer_abundance, unitig_abundance, 0); }

// getUnitigGraph (MBG/MBG.cpp:629-760) without keepGaps, on the ordered table
int mbgfe_build_unitigs_ex(mbgfe* h, uint32_t kmer_abundance, double unitig_abundance, uint32_t flags)
{
	if (!h) return MBGFE_E_PARAM;
	if (!h->processed || !h->have_table) return fail(h, MBGFE_E_STATE, "mbgfe_build_unitigs needs the k-mer table of a finished run");
	if (h->prm.world != 1) return fail(h, MBGFE_E_STATE, "mbgfe_build_unitigs runs on a whole table (world == 1); gather the partitions first");
	if (h->prm.flags & MBGFE_NO_EDGES) return fail(h, MBGFE_E_STATE, "mbgfe_build_unitigs needs the edge table");
	if (kmer_abundance == 0) kmer_abundance = std::max<uint32_t>(1, h->prm.kmer_abundance);
	if (int rc = export_table(h)) return rc;
	const uint64_t n = h->n_nodes_out, m = h->n_edges_out;
	h->have_unitigs = false;
	h->have_threads = false;
	h->have_thin = false;
	h->unitig_min_cov = kmer_abundance;
	uint32_t launches = 0;
	tick(h, 12);
	if (int rc = ensure(h, h->d_g_flag, (std::max(n, m) + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_g_at, (std::max(n, m) + 2) * 8)) return rc;
	if (int rc = ensure(h, h->d_g_kept, (n + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_g_newid, (n + 2) * 8)) return rc;
	GList l { h->d_node_hash_sorted.as<uint4>(), h->d_node_cov.as<uint32_t>(), h->d_node_tip.as<uint8_t>(), h->d_edges_sorted.as<EdgeRec>(), nullptr, n, m, -1 };
	// ---- oneCovHeuristic (MBG/MBG.cpp:631-635): tips first, then every run
	if (flags & MBGFE_UNITIG_ONECOV_HEURISTIC)
	{
		if (int rc = onecov_pass(h, l, true, launches)) return rc;
		if (int rc = onecov_pass(h, l, false, launches)) return rc;
	}
	// ---- which nodes survive: chains over edges with coverage >= -a, mean coverage >= -u
	int which = 0;
	uint64_t n_kept = 0, n_fedges = 0, n_unitigs = 0, n_unitig_nodes = 0, n_uedges = 0;
	if (l.n)
	{
		if (int rc = rank_chains(h, l.edges, l.m, l.cov, l.n, kmer_abundance, &which, launches)) return rc;
		MBGFE_LAUNCH(k_chain_kept, grid_for(h, l.n), 256, 0, h->stream, h->d_g_st[which].as<ChainState>(), l.cov, l.n, unitig_abundance, h->d_g_kept.as<uint32_t>());
		launches++;
		if (int rc = scan_kept(h, l.n, n_kept, launches)) return rc;
		// ---- HashList::filter
		if (int rc = filter_list(h, l, n_kept, launches)) return rc;
		n_fedges = l.m;
		if (l.set == 1)      // the result is read from the d_f_* set
		{
			std::swap(h->d_f_hash, h->d_t_hash); std::swap(h->d_f_cov, h->d_t_cov); std::swap(h->d_f_tip, h->d_t_tip); std::swap(h->d_f_edges, h->d_t_edges); std::swap(h->d_f_orig, h->d_t_orig);
		}
	}
	// kept flags of the exported table: the survivors of every filter
	if (int rc = ensure(h, h->d_g_kept8, n + 64)) return rc;
	if (n)
	{
		CU(h, cudaMemsetAsync(h->d_g_kept8.p, 0, n, h->stream));
		if (n_kept) { MBGFE_LAUNCH(k_mark_kept, grid_for(h, n_kept), 256, 0, h->stream, h->d_f_orig.as<uint32_t>(), n_kept, h->d_g_kept8.as<uint8_t>()); launches++; }
	}
	fetch(h, h->h_kept, h->d_g_kept8, n);
	if (int rc = ensure(h, h->d_g_flag2, (n_kept + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_g_at2, (n_kept + 2) * 8)) return rc;
	if (int rc = ensure(h, h->d_u_begin, (n_kept + 2) * 8)) return rc;
	if (int rc = ensure(h, h->d_u_nodes, (n_kept + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_u_of_node, (n_kept + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_u_fw, (n_kept + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_u_pos, (n_kept + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_u_edges, (n_fedges + 1) * sizeof(EdgeRec))) return rc;
	// ---- unitigs of the filtered list
	if (n_kept)
	{
		const uint32_t* fcov = h->d_f_cov.as<uint32_t>();
		if (int rc = rank_chains(h, h->d_f_edges.as<EdgeRec>(), n_fedges, fcov, n_kept, kmer_abundance, &which, launches)) return rc;
		const ChainState* st = h->d_g_st[which].as<ChainState>();
		MBGFE_LAUNCH(k_unitig_starts, grid_for(h, n_kept), 256, 0, h->stream, st, fcov, n_kept, kmer_abundance, h->d_g_flag.as<uint32_t>(), h->d_g_flag2.as<uint32_t>());
		if (int rc = device_scan(h, h->d_g_flag.as<uint32_t>(), h->d_g_at.as<uint64_t>(), (uint32_t)n_kept)) return rc;
		if (int rc = device_scan(h, h->d_g_flag2.as<uint32_t>(), h->d_g_at2.as<uint64_t>(), (uint32_t)n_kept)) return rc;
		MBGFE_LAUNCH(k_unitig_write, grid_for(h, n_kept), 256, 0, h->stream, st, fcov, n_kept, kmer_abundance, h->d_g_flag.as<uint32_t>(), h->d_g_at.as<uint64_t>(),
			h->d_g_at2.as<uint64_t>(), h->d_u_begin.as<uint64_t>(), h->d_u_nodes.as<uint32_t>(), h->d_u_of_node.as<uint32_t>(), h->d_u_fw.as<uint32_t>(), h->d_u_pos.as<uint32_t>());
		launches += 8;
		CU(h, cudaMemcpyAsync(&n_unitigs, h->d_g_at.as<uint64_t>() + n_kept, 8, cudaMemcpyDeviceToHost, h->stream));       // before the scans are reused
		CU(h, cudaMemcpyAsync(&n_unitig_nodes, h->d_g_at2.as<uint64_t>() + n_kept, 8, cudaMemcpyDeviceToHost, h->stream));
		if (n_fedges)
		{
			// unitig edges: the covered edges that are not links (d_g_link still holds the links of this ranking, cycles cut)
			MBGFE_LAUNCH(k_unitig_edge_flags, grid_for(h, n_fedges), 256, 0, h->stream, h->d_f_edges.as<EdgeRec>(), n_fedges, kmer_abundance, h->d_g_link.as<uint32_t>(),
				h->d_u_of_node.as<uint32_t>(), h->d_u_fw.as<uint32_t>(), h->d_g_flag.as<uint32_t>());
			if (int rc = device_scan(h, h->d_g_flag.as<uint32_t>(), h->d_g_at.as<uint64_t>(), (uint32_t)n_fedges)) return rc;
			MBGFE_LAUNCH(k_unitig_edge_write, grid_for(h, n_fedges), 256, 0, h->stream, h->d_f_edges.as<EdgeRec>(), n_fedges, kmer_abundance, h->d_g_link.as<uint32_t>(),
				h->d_u_of_node.as<uint32_t>(), h->d_u_fw.as<uint32_t>(), h->d_g_at.as<uint64_t>(), h->d_u_edges.as<EdgeRec>());
			launches += 5;
			CU(h, cudaMemcpyAsync(&n_uedges, h->d_g_at.as<uint64_t>() + n_fedges, 8, cudaMemcpyDeviceToHost, h->stream));
		}
		CU(h, cudaStreamSynchronize(h->stream));
		CU(h, cudaMemcpyAsync(h->d_u_begin.as<uint64_t>() + n_unitigs, &n_unitig_nodes, 8, cudaMemcpyHostToDevice, h->stream));
	}
	tick(h, 13);
	fetch(h, h->h_f_hash, h->d_f_hash, n_kept);
	fetch(h, h->h_f_cov, h->d_f_cov, n_kept);
	fetch(h, h->h_f_tip, h->d_f_tip, n_kept);
	fetch(h, h->h_f_edges, h->d_f_edges, n_fedges);
	fetch(h, h->h_u_nodes, h->d_u_nodes, n_unitig_nodes);
	fetch(h, h->h_u_of_node, h->d_u_of_node, n_kept ? n_kept : 0);
	if (n_kept) fetch(h, h->h_u_begin, h->d_u_begin, n_unitigs + 1); else h->h_u_begin.assign(1, 0);
	{
		std::vector<EdgeRec> ue;
		fetch(h, ue, h->d_u_edges, n_uedges);
		CU(h, cudaStreamSynchronize(h->stream));
		// listed like MBGFE_EDGES: by (from, forward side first, to, to_fw); unitig edges are few, the host orders them
		std::sort(ue.begin(), ue.end(), [](const EdgeRec& a, const EdgeRec& b)
		{
			if (a.from != b.from) return a.from < b.from;
			if (a.from_fw != b.from_fw) return a.from_fw > b.from_fw;
			if (a.to != b.to) return a.to < b.to;
			return a.to_fw < b.to_fw;
		});
		h->h_u_edges.resize(ue.size());
		if (!ue.empty()) memcpy(h->h_u_edges.data(), ue.data(), ue.size() * sizeof(mbgfe_edge));
	}
	CU(h, cudaGetLastError());
	h->n_kept = n_kept; h->n_fedges = n_fedges; h->n_unitigs = n_unitigs; h->n_unitig_nodes = n_unitig_nodes; h->n_uedges = n_uedges;
	h->stats.ms_unitigs = elapsed(h, 12, 13);
	h->stats.launches = launches;
	h->have_unitigs = true;
	return 0;
}

// filterKmersToUnitigKmers (MBG/MBG.cpp:1065-1185) with filterWithinUnitig = true: which k-mers of the filtered list stay, and
// where every k-mer sits inside its unitig (graph.cuh)
int mbgfe_thin_unitigs(mbgfe* h)
{
	if (!h) return MBGFE_E_PARAM;
	if (!h->have_unitigs) return fail(h, MBGFE_E_STATE, "mbgfe_thin_unitigs needs mbgfe_build_unitigs first");
	const uint64_t n_kept = h->n_kept, n_pos = h->n_unitig_nodes, n_fedges = h->n_fedges;
	uint32_t launches = 0;
	h->have_thin = false; h->have_threads = false;
	tick(h, 12);
	if (int rc = ensure(h, h->d_thin, n_kept + 64)) return rc;
	if (int rc = ensure(h, h->d_thin_pos, (n_kept + 1) * 4)) return rc;
	if (n_kept)
	{
		CU(h, cudaMemsetAsync(h->d_thin.p, 0, n_kept, h->stream));
		CU(h, cudaMemsetAsync(h->d_thin_pos.p, 0, n_kept * 4, h->stream));
	}
	if (n_pos)
	{
		if (int rc = ensure(h, h->d_g_flag, (n_pos + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_g_at, (n_pos + 2) * 8)) return rc;
		if (int rc = ensure(h, h->d_g_flag2, (n_pos + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_g_kept, (n_pos + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_g_deg, (n_pos + 1) * 4)) return rc;
		uint32_t* delta = h->d_g_flag.as<uint32_t>();
		uint32_t* jump[2] = { h->d_g_flag2.as<uint32_t>(), h->d_g_kept.as<uint32_t>() };
		uint32_t* mark = h->d_g_deg.as<uint32_t>();
		CU(h, cudaMemsetAsync(delta + n_pos, 0, 4, h->stream));
		MBGFE_LAUNCH(k_thin_delta, grid_for(h, n_pos), 256, 0, h->stream, h->d_u_nodes.as<uint32_t>(), n_pos, h->d_u_of_node.as<uint32_t>(), h->d_u_begin.as<uint64_t>(),
			h->d_f_edges.as<EdgeRec>(), n_fedges, h->prm.k, delta);
		if (int rc = device_scan(h, delta, h->d_g_at.as<uint64_t>(), (uint32_t)n_pos)) return rc;
		MBGFE_LAUNCH(k_thin_next, grid_for(h, n_pos), 256, 0, h->stream, h->d_u_nodes.as<uint32_t>(), n_pos, h->d_u_of_node.as<uint32_t>(), h->d_u_begin.as<uint64_t>(),
			h->d_f_tip.as<uint8_t>(), delta, h->d_g_at.as<uint64_t>(), h->prm.k, jump[0], mark, h->d_thin_pos.as<uint32_t>());
		launches += 5;
		int cur = 0;
		for (uint64_t reach = 1; reach < n_pos; reach <<= 1)      // after the round every k-mer within 2 * reach - 1 steps of its head is marked
		{
			MBGFE_LAUNCH(k_thin_mark, grid_for(h, n_pos), 256, 0, h->stream, jump[cur], n_pos, mark);
			MBGFE_LAUNCH(k_thin_double, grid_for(h, n_pos), 256, 0, h->stream, jump[cur], n_pos, jump[cur ^ 1]);
			cur ^= 1;
			launches += 2;
		}
		MBGFE_LAUNCH(k_thin_mark, grid_for(h, n_pos), 256, 0, h->stream, jump[cur], n_pos, mark);
		MBGFE_LAUNCH(k_thin_flags, grid_for(h, n_pos), 256, 0, h->stream, h->d_u_nodes.as<uint32_t>(), n_pos, mark, h->d_thin.as<uint8_t>());
		launches += 2;
	}
	// ---- apply it (MBG/MBG.cpp:1166-1184): reads.filter(kept) into the other scratch set, the unitigs renumbered, the new overlaps
	uint64_t n_thin = 0, n_tedges = 0, n_tpos = 0;
	if (n_pos)
	{
		uint32_t* mark = h->d_g_deg.as<uint32_t>();
		if (int rc = ensure(h, h->d_thin_scan, (n_pos + 2) * 8)) return rc;
		if (int rc = ensure(h, h->d_tpos, (n_kept + 1) * 4)) return rc;
		if (int rc = device_scan(h, mark, h->d_thin_scan.as<uint64_t>(), (uint32_t)n_pos)) return rc;
		MBGFE_LAUNCH(k_thin_kept32, grid_for(h, n_kept), 256, 0, h->stream, h->d_thin.as<uint8_t>(), n_kept, h->d_g_kept.as<uint32_t>());
		launches += 4;
		if (int rc = ensure(h, h->d_g_newid, (n_kept + 2) * 8)) return rc;
		if (int rc = scan_kept(h, n_kept, n_thin, launches)) return rc;
		GList l { h->d_f_hash.as<uint4>(), h->d_f_cov.as<uint32_t>(), h->d_f_tip.as<uint8_t>(), h->d_f_edges.as<EdgeRec>(), h->d_f_orig.as<uint32_t>(), n_kept, n_fedges, 0 };
		if (int rc = ensure(h, h->d_g_flag, (std::max(n_fedges, n_pos) + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_g_at, (std::max(n_fedges, n_pos) + 2) * 8)) return rc;
		if (int rc = filter_list(h, l, n_thin, launches)) return rc;        // -> the d_t_* set
		n_tedges = l.m;
		n_tpos = n_thin;               // every k-mer that stays sits in exactly one unitig position
		if (int rc = ensure(h, h->d_tu_nodes, (n_tpos + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_tu_off, (n_tpos + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_tu_unitig, (n_tpos + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_tu_begin, (h->n_unitigs + 2) * 8)) return rc;
		if (int rc = ensure(h, h->d_t_newov, (n_tpos + 1) * sizeof(EdgeRec))) return rc;
		MBGFE_LAUNCH(k_thin_compact, grid_for(h, n_pos), 256, 0, h->stream, h->d_u_nodes.as<uint32_t>(), n_pos, mark, h->d_thin_scan.as<uint64_t>(), h->d_u_of_node.as<uint32_t>(),
			h->d_u_begin.as<uint64_t>(), h->d_g_newid.as<uint64_t>(), h->d_thin_pos.as<uint32_t>(), h->d_tpos.as<uint32_t>(), h->d_tu_nodes.as<uint32_t>(), h->d_tu_off.as<uint32_t>(),
			h->d_tu_unitig.as<uint32_t>());
		MBGFE_LAUNCH(k_thin_begins, grid_for(h, h->n_unitigs + 1), 256, 0, h->stream, h->d_u_begin.as<uint64_t>(), h->n_unitigs, h->d_thin_scan.as<uint64_t>(), h->d_tu_begin.as<uint64_t>());
		MBGFE_LAUNCH(k_thin_overlaps, grid_for(h, n_tpos), 256, 0, h->stream, h->d_tu_nodes.as<uint32_t>(), h->d_tu_off.as<uint32_t>(), h->d_tu_unitig.as<uint32_t>(),
			h->d_tu_begin.as<uint64_t>(), n_tpos, h->prm.k, h->d_t_newov.as<EdgeRec>());
		launches += 3;
	}
	tick(h, 13);
	fetch(h, h->h_thin, h->d_thin, n_kept);
	fetch(h, h->h_thin_pos, h->d_thin_pos, n_kept);
	fetch(h, h->h_t_hash, h->d_t_hash, n_thin);
	fetch(h, h->h_t_cov, h->d_t_cov, n_thin);
	fetch(h, h->h_t_tip, h->d_t_tip, n_thin);
	fetch(h, h->h_t_edges, h->d_t_edges, n_tedges);
	fetch(h, h->h_tu_nodes, h->d_tu_nodes, n_tpos);
	if (n_pos) fetch(h, h->h_tu_begin, h->d_tu_begin, h->n_unitigs + 1); else h->h_tu_begin.assign(h->n_unitigs + 1, 0);
	fetch(h, h->h_t_newov, h->d_t_newov, n_tpos - std::min<uint64_t>(n_tpos, h->n_unitigs));
	h->n_thin = n_thin; h->n_tedges = n_tedges; h->n_tpos = n_tpos;
	CU(h, cudaStreamSynchronize(h->stream));
	CU(h, cudaGetLastError());
	h->stats.ms_unitigs = elapsed(h, 12, 13);
	h->stats.launches = launches;
	h->have_thin = true;
	return 0;
}

// getReadPaths (MBG/MBG.cpp:1235-1251) as a batched lookup: every selected k-mer instance of the last run located in the
// unitigs of the last mbgfe_build_unitigs[_ex] and classified the way iterateReadPaths (MBG/KmerMatcher.h:12-144) would
int mbgfe_thread_reads(mbgfe* h)
{
	if (!h) return MBGFE_E_PARAM;
	if (!h->have_unitigs) return fail(h, MBGFE_E_STATE, "mbgfe_thread_reads needs mbgfe_build_unitigs first");
	if (h->prm.world != 1 || !h->edge_slots) return fail(h, MBGFE_E_STATE, "mbgfe_thread_reads runs on a whole table with edges (world == 1)");
	const uint64_t n_inst = h->n_inst, n = h->n_nodes_out, n_kept = h->n_kept;
	uint32_t launches = 0;
	tick(h, 12);
	if (int rc = ensure(h, h->d_th_final, (n + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_th_node, (n_inst + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_th_unitig, (n_inst + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_th_upos, (n_inst + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_th_urest, (n_inst + 1) * 4)) return rc;
	if (int rc = ensure(h, h->d_th_step, n_inst + 64)) return rc;
	if (n_inst)
	{
		CU(h, cudaMemsetAsync(h->d_th_final.p, 0xff, (n + 1) * 4, h->stream));
		if (n_kept) { MBGFE_LAUNCH(k_invert_orig, grid_for(h, n_kept), 256, 0, h->stream, h->d_f_orig.as<uint32_t>(), n_kept, h->d_th_final.as<uint32_t>()); launches++; }
		MBGFE_LAUNCH(k_thread_locate, grid_for(h, n_inst), 256, 0, h->stream, h->d_inst_slot.as<uint32_t>(), h->d_inst_hash.as<uint64_t>(), n_inst, h->d_nodes.as<NodeSlot>(),
			h->d_th_final.as<uint32_t>(), h->d_u_of_node.as<uint32_t>(), h->d_u_fw.as<uint32_t>(), h->d_u_pos.as<uint32_t>(), h->d_u_begin.as<uint64_t>(),
			h->have_thin ? h->d_thin.as<uint8_t>() : (const uint8_t*)nullptr, h->have_thin ? h->d_tpos.as<uint32_t>() : (const uint32_t*)nullptr,
			h->have_thin ? h->d_thin_scan.as<uint64_t>() : (const uint64_t*)nullptr, h->d_th_node.as<uint32_t>(), h->d_th_unitig.as<uint32_t>(), h->d_th_upos.as<uint32_t>(), h->d_th_urest.as<uint32_t>());
		MBGFE_LAUNCH(k_thread_steps, grid_for(h, n_inst), 256, 0, h->stream, h->d_th_node.as<uint32_t>(), h->d_th_unitig.as<uint32_t>(), h->d_th_upos.as<uint32_t>(),
			h->d_th_urest.as<uint32_t>(), h->d_inst_flag.as<uint8_t>(), h->d_inst_pos.as<uint32_t>(), h->d_inst_slot.as<uint32_t>(), h->d_inst_hash.as<uint64_t>(), n_inst,
			h->d_edges.as<EdgeSlot>(), h->edge_slots - 1, h->prm.k, h->unitig_min_cov, h->have_thin ? h->d_thin_pos.as<uint32_t>() : (const uint32_t*)nullptr, h->d_th_step.as<uint8_t>());
		launches += 2;
	}
	tick(h, 13);
	fetch(h, h->h_th_unitig, h->d_th_unitig, n_inst);
	fetch(h, h->h_th_upos, h->d_th_upos, n_inst);
	fetch(h, h->h_th_urest, h->d_th_urest, n_inst);
	fetch(h, h->h_th_step, h->d_th_step, n_inst);
	CU(h, cudaStreamSynchronize(h->stream));
	CU(h, cudaGetLastError());
	h->stats.ms_unitigs = elapsed(h, 12, 13);
	h->stats.launches = launches;
	h->have_threads = true;
	return 0;
}

}

namespace {

}

extern "C" {

uint64_t mbgfe_count(mbgfe* h, int what)
{
	if (!h || !h->processed) return 0;
	switch (what)
	{
		case MBGFE_PARTS: if (export_parts(h)) return 0; return h->parts.size();
		case MBGFE_CODES: if (export_parts(h)) return 0; return (h->prm.flags & MBGFE_KEEP_CODES) ? h->stats.codes : 0;
		case MBGFE_POSES: if (export_parts(h)) return 0; return (h->prm.flags & MBGFE_KEEP_POSES) ? h->stats.codes + h->parts.size() : 0;
		case MBGFE_POSITIONS: case MBGFE_HASHES: return h->n_inst;
		case MBGFE_NODE_HASHES: case MBGFE_NODE_COVERAGE: case MBGFE_NODE_TIP: return h->have_table ? h->stats.distinct : 0;
		case MBGFE_EDGES: return (h->have_table && h->edge_slots) ? h->stats.distinct_edges : 0;
		case MBGFE_NODE_KEPT: return h->have_unitigs ? h->h_kept.size() : 0;
		case MBGFE_F_NODE_HASHES: case MBGFE_F_NODE_COVERAGE: case MBGFE_F_NODE_TIP: case MBGFE_F_NODE_UNITIG: return h->have_unitigs ? h->n_kept : 0;
		case MBGFE_F_EDGES: return h->have_unitigs ? h->n_fedges : 0;
		case MBGFE_UNITIG_BEGIN: return h->have_unitigs ? h->n_unitigs + 1 : 0;
		case MBGFE_UNITIG_NODES: return h->have_unitigs ? h->n_unitig_nodes : 0;
		case MBGFE_UNITIG_EDGES: return h->have_unitigs ? h->n_uedges : 0;
		case MBGFE_INST_UNITIG: case MBGFE_INST_UNITIG_POS: case MBGFE_INST_UNITIG_REST: case MBGFE_INST_STEP: return h->have_threads ? h->h_th_step.size() : 0;
		case MBGFE_F_NODE_THIN: case MBGFE_F_NODE_UNITIG_OFFSET: return h->have_thin ? h->h_thin.size() : 0;
		case MBGFE_T_NODE_HASHES: case MBGFE_T_NODE_COVERAGE: case MBGFE_T_NODE_TIP: return h->have_thin ? h->n_thin : 0;
		case MBGFE_T_EDGES: return h->have_thin ? h->n_tedges : 0;
		case MBGFE_T_UNITIG_BEGIN: return h->have_thin ? h->h_tu_begin.size() : 0;
		case MBGFE_T_UNITIG_NODES: return h->have_thin ? h->n_tpos : 0;
		case MBGFE_T_NEW_OVERLAPS: return h->have_thin ? h->h_t_newov.size() : 0;
	}
	return 0;
}

const void* mbgfe_data(mbgfe* h, int what)
{
	if (!h) return nullptr;
	if (!h->processed) { fail(h, MBGFE_E_STATE, "results requested before mbgfe_process"); return nullptr; }
	switch (what)
	{
		case MBGFE_PARTS: if (export_parts(h)) return nullptr; return h->parts.data();
		case MBGFE_CODES:
			if (!(h->prm.flags & MBGFE_KEEP_CODES)) { fail(h, MBGFE_E_STATE, "codes requested without MBGFE_KEEP_CODES"); return nullptr; }
			if (export_codes(h)) return nullptr; return h->h_codes.data();
		case MBGFE_POSES:
			if (!(h->prm.flags & MBGFE_KEEP_POSES)) { fail(h, MBGFE_E_STATE, "poses requested without MBGFE_KEEP_POSES"); return nullptr; }
			if (export_codes(h)) return nullptr; return h->h_poses.data();
		case MBGFE_POSITIONS: if (export_instances(h)) return nullptr; return h->h_positions.data();
		case MBGFE_HASHES: if (export_instances(h)) return nullptr; return h->h_hashes.data();
		case MBGFE_NODE_HASHES: if (export_table(h)) return nullptr; return h->p_node_hashes.p;
		case MBGFE_NODE_COVERAGE: if (export_table(h)) return nullptr; return h->p_node_cov.p;
		case MBGFE_NODE_TIP: if (export_table(h)) return nullptr; return h->p_node_tip.p;
		case MBGFE_EDGES: if (export_table(h)) return nullptr; return h->p_edges.p;
		case MBGFE_NODE_KEPT: case MBGFE_F_NODE_HASHES: case MBGFE_F_NODE_COVERAGE: case MBGFE_F_NODE_TIP: case MBGFE_F_NODE_UNITIG: case MBGFE_F_EDGES:
		case MBGFE_UNITIG_BEGIN: case MBGFE_UNITIG_NODES: case MBGFE_UNITIG_EDGES:
			if (!h->have_unitigs) { fail(h, MBGFE_E_STATE, "filtered table / unitigs requested before mbgfe_build_unitigs"); return nullptr; }
			switch (what)
			{
				case MBGFE_NODE_KEPT: return h->h_kept.data();
				case MBGFE_F_NODE_HASHES: return h->h_f_hash.data();
				case MBGFE_F_NODE_COVERAGE: return h->h_f_cov.data();
				case MBGFE_F_NODE_TIP: return h->h_f_tip.data();
				case MBGFE_F_NODE_UNITIG: return h->h_u_of_node.data();
				case MBGFE_F_EDGES: return h->h_f_edges.data();
				case MBGFE_UNITIG_BEGIN: return h->h_u_begin.data();
				case MBGFE_UNITIG_EDGES: return h->h_u_edges.data();
				default: return h->h_u_nodes.data();
			}
		case MBGFE_INST_UNITIG: case MBGFE_INST_UNITIG_POS: case MBGFE_INST_UNITIG_REST: case MBGFE_INST_STEP:
			if (!h->have_threads) { fail(h, MBGFE_E_STATE, "read threading requested before mbgfe_thread_reads"); return nullptr; }
			switch (what)
			{
				case MBGFE_INST_UNITIG: return h->h_th_unitig.data();
				case MBGFE_INST_UNITIG_POS: return h->h_th_upos.data();
				case MBGFE_INST_UNITIG_REST: return h->h_th_urest.data();
				default: return h->h_th_step.data();
			}
	}
	if (what == MBGFE_F_NODE_THIN || what == MBGFE_F_NODE_UNITIG_OFFSET || (what >= MBGFE_T_NODE_HASHES && what <= MBGFE_T_NEW_OVERLAPS))
	{
		if (!h->have_thin) { fail(h, MBGFE_E_STATE, "thinning requested before mbgfe_thin_unitigs"); return nullptr; }
		switch (what)
		{
			case MBGFE_F_NODE_THIN: return h->h_thin.data();
			case MBGFE_F_NODE_UNITIG_OFFSET: return h->h_thin_pos.data();
			case MBGFE_T_NODE_HASHES: return h->h_t_hash.data();
			case MBGFE_T_NODE_COVERAGE: return h->h_t_cov.data();
			case MBGFE_T_NODE_TIP: return h->h_t_tip.data();
			case MBGFE_T_EDGES: return h->h_t_edges.data();
			case MBGFE_T_UNITIG_BEGIN: return h->h_tu_begin.data();
			case MBGFE_T_UNITIG_NODES: return h->h_tu_nodes.data();
			default: return h->h_t_newov.data();
		}
	}
	fail(h, MBGFE_E_PARAM, "unknown result array %d", what);
	return nullptr;
}

int mbgfe_get_stats(const mbgfe* h, mbgfe_stats* out)
{
	if (!h || !out) return MBGFE_E_PARAM;
	*out = h->stats;
	return 0;
}

int mbgfe_timer_start(mbgfe* h)
{
	if (!h) return MBGFE_E_PARAM;
	CU(h, cudaEventRecord(h->ev[14], h->stream));
	return 0;
}

int mbgfe_timer_stop(mbgfe* h, float* ms)
{
	if (!h || !ms) return MBGFE_E_PARAM;
	CU(h, cudaEventRecord(h->ev[15], h->stream));
	CU(h, cudaEventSynchronize(h->ev[15]));
	CU(h, cudaEventElapsedTime(ms, h->ev[14], h->ev[15]));
	return 0;
}

// ====================================================================================== multi-GPU exchange
// The whole exchange runs inside the library: counts of both collectives from one kernel and one all-gather, node records to
// their owners, ids answered from the slot each record landed in, edge records to the owner of their `from` node.  Three
// transports move the bytes: NCCL (one process per GPU; grouped ncclSend / ncclRecv over NVLink), peer copies between the
// handles of ONE process (mbgfe_exchange_group), or the caller's own collectives (callbacks; the CPU tests use gloo).
// Host synchronisations per exchange: the count matrix, the partition sizes, the final counters.

}

namespace {

#ifndef MBGFE_EMU_BUILD
struct NcclApi
{
	void* lib = nullptr;
	ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
	ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
	ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
	ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*GroupStart)() = nullptr;
	ncclResult_t (*GroupEnd)() = nullptr;
	const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi* nccl_api()
{
	static NcclApi api;
	static bool tried = false;
	if (!tried)
	{
		tried = true;
		// the process may already hold a libnccl.so.2 (torch's); dlopen then hands back that one
		void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
		if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
		if (lib)
		{
			api.lib = lib;
			*(void**)&api.GetUniqueId = dlsym(lib, "ncclGetUniqueId");
			*(void**)&api.CommInitRank = dlsym(lib, "ncclCommInitRank");
			*(void**)&api.CommDestroy = dlsym(lib, "ncclCommDestroy");
			*(void**)&api.AllGather = dlsym(lib, "ncclAllGather");
			*(void**)&api.Send = dlsym(lib, "ncclSend");
			*(void**)&api.Recv = dlsym(lib, "ncclRecv");
			*(void**)&api.GroupStart = dlsym(lib, "ncclGroupStart");
			*(void**)&api.GroupEnd = dlsym(lib, "ncclGroupEnd");
			*(void**)&api.GetErrorString = dlsym(lib, "ncclGetErrorString");
			if (!api.GetUniqueId || !api.CommInitRank || !api.AllGather || !api.Send || !api.Recv || !api.GroupStart || !api.GroupEnd) api.lib = nullptr;
		}
	}
	return api.lib ? &api : nullptr;
}
#define NCCLCHK(h, call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) return fail(h, MBGFE_E_CUDA, "%s failed: %s", #call, nccl_api()->GetErrorString ? nccl_api()->GetErrorString(r_) : "nccl error"); } while (0)
#endif

void destroy_comm(mbgfe* h)
{
#ifndef MBGFE_EMU_BUILD
	if (h->nccl_comm && nccl_api() && nccl_api()->CommDestroy) nccl_api()->CommDestroy((ncclComm_t)h->nccl_comm);
#endif
	h->nccl_comm = nullptr; h->comm_kind = 0;
}

// ---- transport: all-gather of `bytes` per rank and all-to-all-v between device buffers.  hs = the handles taking part IN THIS
// PROCESS: one (NCCL, callbacks: the other ranks are other processes) or all `world` of them (group).
int x_allgather(std::vector<mbgfe*>& hs, std::vector<const void*>& send, std::vector<void*>& recv, size_t bytes)
{
	mbgfe* h0 = hs[0];
	if (h0->comm_kind == 2)
	{
		for (mbgfe* h : hs) { CU(h, cudaSetDevice(h->prm.device)); CU(h, cudaStreamSynchronize(h->stream)); }
		for (size_t j = 0; j < hs.size(); j++)
			for (size_t i = 0; i < hs.size(); i++)
				CU(hs[j], cudaMemcpyPeerAsync((char*)recv[j] + i * bytes, hs[j]->prm.device, send[i], hs[i]->prm.device, bytes, hs[j]->stream));
		return 0;
	}
	if (h0->comm_kind == 3)
	{
		CU(h0, cudaStreamSynchronize(h0->stream));
		if (h0->cb_allgather(h0->cb_ctx, send[0], recv[0], bytes)) return fail(h0, MBGFE_E_CUDA, "all-gather callback failed");
		return 0;
	}
#ifndef MBGFE_EMU_BUILD
	if (h0->comm_kind == 1) { NCCLCHK(h0, nccl_api()->AllGather(send[0], recv[0], bytes, ncclUint8, (ncclComm_t)h0->nccl_comm, h0->stream)); return 0; }
#endif
	return fail(h0, MBGFE_E_STATE, "no communicator: call mbgfe_comm_init_nccl / mbgfe_comm_init_callbacks, or use mbgfe_exchange_group");
}

// counts in elements of elem bytes; per handle: send[p] elements for rank p from send_buf (grouped by rank), recv likewise
int x_alltoallv(std::vector<mbgfe*>& hs, std::vector<const void*>& send_buf, std::vector<std::vector<uint64_t>>& send_cnt, std::vector<void*>& recv_buf,
	std::vector<std::vector<uint64_t>>& recv_cnt, size_t elem)
{
	mbgfe* h0 = hs[0];
	const uint32_t world = h0->prm.world;
	if (h0->comm_kind == 2)
	{
		for (mbgfe* h : hs) { CU(h, cudaSetDevice(h->prm.device)); CU(h, cudaStreamSynchronize(h->stream)); }
		for (size_t j = 0; j < hs.size(); j++)      // receiver j pulls what every sender i packed for it
		{
			uint64_t roff = 0;
			for (size_t i = 0; i < hs.size(); i++)
			{
				uint64_t soff = 0;
				for (size_t q = 0; q < j; q++) soff += send_cnt[i][q];
				const uint64_t n = send_cnt[i][j];
				if (n) CU(hs[j], cudaMemcpyPeerAsync((char*)recv_buf[j] + roff * elem, hs[j]->prm.device, (const char*)send_buf[i] + soff * elem, hs[i]->prm.device, n * elem, hs[j]->stream));
				roff += n;
			}
		}
		return 0;
	}
	if (h0->comm_kind == 3)
	{
		CU(h0, cudaStreamSynchronize(h0->stream));
		std::vector<uint64_t> sb(world), rb(world);
		for (uint32_t p = 0; p < world; p++) { sb[p] = send_cnt[0][p] * elem; rb[p] = recv_cnt[0][p] * elem; }
		if (h0->cb_alltoallv(h0->cb_ctx, send_buf[0], sb.data(), recv_buf[0], rb.data())) return fail(h0, MBGFE_E_CUDA, "all-to-all callback failed");
		return 0;
	}
#ifndef MBGFE_EMU_BUILD
	if (h0->comm_kind == 1)
	{
		NcclApi* n = nccl_api();
		NCCLCHK(h0, n->GroupStart());
		uint64_t soff = 0, roff = 0;
		for (uint32_t p = 0; p < world; p++)
		{
			if (send_cnt[0][p]) NCCLCHK(h0, n->Send((const char*)send_buf[0] + soff * elem, send_cnt[0][p] * elem, ncclUint8, (int)p, (ncclComm_t)h0->nccl_comm, h0->stream));
			if (recv_cnt[0][p]) NCCLCHK(h0, n->Recv((char*)recv_buf[0] + roff * elem, recv_cnt[0][p] * elem, ncclUint8, (int)p, (ncclComm_t)h0->nccl_comm, h0->stream));
			soff += send_cnt[0][p]; roff += recv_cnt[0][p];
		}
		NCCLCHK(h0, n->GroupEnd());
		return 0;
	}
#endif
	return fail(h0, MBGFE_E_STATE, "no communicator");
}

int exchange_run(std::vector<mbgfe*>& hs)
{
	const size_t nh = hs.size();
	const uint32_t world = hs[0]->prm.world;
	const size_t row = 2 * world + 1;                       // node counts, edge counts, instances
	std::vector<const void*> sp(nh); std::vector<void*> rp(nh);
	auto on = [&](mbgfe* h) { return cudaSetDevice(h->prm.device); };
	// MBGFE_EXCHANGE_TRACE=1: device time of every phase of rank hs[0] on stderr (synchronises after each phase: for diagnosis only)
	const bool trace = getenv("MBGFE_EXCHANGE_TRACE") != nullptr;
	cudaEvent_t tr0 = nullptr, tr1 = nullptr;
	if (trace) { cudaEventCreate(&tr0); cudaEventCreate(&tr1); cudaEventRecord(tr0, hs[0]->stream); }
	auto mark = [&](const char* name)
	{
		if (!trace) return;
		cudaEventRecord(tr1, hs[0]->stream); cudaEventSynchronize(tr1);
		float ms = 0; cudaEventElapsedTime(&ms, tr0, tr1);
		fprintf(stderr, "[exchange rank %u] %-22s %8.3f ms\n", hs[0]->prm.rank, name, ms);
		cudaEventRecord(tr0, hs[0]->stream);
	};
	// ---- S1: counts of both collectives
	for (size_t x = 0; x < nh; x++)
	{
		mbgfe* h = hs[x];
		CU(h, on(h));
		if (!h->processed) return fail(h, MBGFE_E_STATE, "exchange before mbgfe_process / mbgfe_run");
		if (int rc = ensure(h, h->d_xrow, row * 8)) return rc;
		if (int rc = ensure(h, h->d_xmatrix, row * world * 8)) return rc;
		tick(h, 12);
		CU(h, cudaMemsetAsync(h->d_xrow.p, 0, row * 8, h->stream));
		if (h->n_inst) MBGFE_LAUNCH(k_owner_count2, (unsigned)((h->n_inst + 255) / 256), 256, 0, h->stream, h->d_inst_hash.as<uint64_t>(), h->d_inst_flag.as<uint8_t>(), h->n_inst, world,
			h->d_xrow.as<unsigned long long>());
		CU(h, cudaMemcpyAsync(h->d_xrow.as<uint64_t>() + 2 * world, &h->n_inst, 8, cudaMemcpyHostToDevice, h->stream));
		sp[x] = h->d_xrow.p; rp[x] = h->d_xmatrix.p;
	}
	mark("count kernel");
	if (int rc = x_allgather(hs, sp, rp, row * 8)) return rc;
	mark("count all-gather");
	std::vector<std::vector<uint64_t>> M(nh, std::vector<uint64_t>(row * world));
	for (size_t x = 0; x < nh; x++) { mbgfe* h = hs[x]; CU(h, on(h)); CU(h, cudaMemcpyAsync(M[x].data(), h->d_xmatrix.p, row * world * 8, cudaMemcpyDeviceToHost, h->stream)); }
	for (size_t x = 0; x < nh; x++) { mbgfe* h = hs[x]; CU(h, on(h)); CU(h, cudaStreamSynchronize(h->stream)); }
	// ---- S2: node records, grouped by owner
	std::vector<std::vector<uint64_t>> nsend(nh, std::vector<uint64_t>(world)), nrecv(nh, std::vector<uint64_t>(world)), esend(nh, std::vector<uint64_t>(world)), erecv(nh, std::vector<uint64_t>(world));
	std::vector<uint64_t> n_send(nh), n_recv(nh), e_send(nh), e_recv(nh), inst_base(nh);
	for (size_t x = 0; x < nh; x++)
	{
		mbgfe* h = hs[x];
		CU(h, on(h));
		const uint32_t me = h->prm.rank;
		const uint64_t* m = M[x].data();
		n_send[x] = n_recv[x] = e_send[x] = e_recv[x] = inst_base[x] = 0;
		for (uint32_t p = 0; p < world; p++)
		{
			nsend[x][p] = m[me * row + p]; nrecv[x][p] = m[p * row + me]; esend[x][p] = m[me * row + world + p]; erecv[x][p] = m[p * row + world + me];
			n_send[x] += nsend[x][p]; n_recv[x] += nrecv[x][p]; e_send[x] += esend[x][p]; e_recv[x] += erecv[x][p];
			if (p < me) inst_base[x] += m[p * row + 2 * world];
		}
		if (int rc = ensure(h, h->d_records, (n_send[x] + 1) * 32)) return rc;
		if (int rc = ensure(h, h->d_recv, (n_recv[x] + 1) * 32)) return rc;
		if (int rc = ensure(h, h->d_inst_packpos, (h->n_inst + 1) * 4)) return rc;
		std::vector<unsigned long long> cursors(world);
		uint64_t at = 0;
		for (uint32_t p = 0; p < world; p++) { cursors[p] = at; at += nsend[x][p]; }
		unsigned long long* ctr = h->d_counters.as<unsigned long long>();
		CU(h, cudaMemcpyAsync(ctr + CTR_OWNER, cursors.data(), world * 8, cudaMemcpyHostToDevice, h->stream));
		if (h->n_inst) MBGFE_LAUNCH(k_owner_pack, (unsigned)((h->n_inst + 255) / 256), 256, 0, h->stream, h->d_inst_hash.as<uint64_t>(), h->d_inst_flag.as<uint8_t>(), h->n_inst, world,
			ctr + CTR_OWNER, h->d_records.as<uint64_t>(), h->d_inst_packpos.as<uint32_t>());
		CU(h, cudaStreamSynchronize(h->stream));      // `cursors` is a host temporary of this iteration
		sp[x] = h->d_records.p; rp[x] = h->d_recv.p;
	}
	mark("node pack");
	if (int rc = x_alltoallv(hs, sp, nsend, rp, nrecv, 32)) return rc;
	mark("node all-to-all");
	// ---- S3: this rank's partition of the node table
	for (size_t x = 0; x < nh; x++)
	{
		mbgfe* h = hs[x];
		CU(h, on(h));
		unsigned long long* ctr = h->d_counters.as<unsigned long long>();
		const uint64_t floor_slots = h->prm.table_slots ? next_pow2(h->prm.table_slots) : 0;
		uint64_t slots = std::max<uint64_t>(next_pow2(std::max<uint64_t>(2 * n_recv[x], 1024)), floor_slots);
		if (sketch_wanted(n_recv[x]))
		{
			// two slots per DISTINCT k-mer this rank received, not two per record (size_tables)
			uint64_t est[2]; uint32_t dummy = 0;
			if (int rc = run_sketch(h, n_recv[x], false, est, dummy, [&](uint32_t* bm0, uint32_t*, uint64_t mask)
			{
				MBGFE_LAUNCH(k_sketch_records, (unsigned)std::min<uint64_t>((n_recv[x] + 255) / 256, (uint64_t)h->sm_count * 16), 256, 0, h->stream, h->d_recv.as<uint64_t>(), n_recv[x], 4u, 1u, bm0, mask);
			})) return rc;
			slots = std::min(slots, std::max<uint64_t>(next_pow2(std::max<uint64_t>(2 * est[0], 1024)), floor_slots));
		}
		if (slots >= (1ull << 31)) return fail(h, MBGFE_E_LIMIT, "node table would need %llu slots (limit 2^31 per GPU)", (unsigned long long)slots);
		if (int rc = ensure(h, h->d_nodes, slots * sizeof(NodeSlot))) return rc;
		if (int rc = ensure(h, h->d_rec_slot, (n_recv[x] + 1) * 4)) return rc;
		h->node_slots = slots; h->edge_slots = 0;
		CU(h, cudaMemsetAsync(ctr + CTR_DISTINCT, 0, 16, h->stream));
		CU(h, cudaMemsetAsync(ctr + CTR_OVERFLOW, 0, 8, h->stream));
		const unsigned clear_grid = (unsigned)std::min<uint64_t>((slots + 255) / 256, (uint64_t)h->sm_count * 8);
		MBGFE_LAUNCH(k_table_clear, clear_grid, 256, 0, h->stream, h->d_nodes.as<NodeSlot>(), slots);
		if (n_recv[x]) MBGFE_LAUNCH(k_insert_records2, (unsigned)((n_recv[x] + 255) / 256), 256, 0, h->stream, h->d_recv.as<uint64_t>(), n_recv[x], h->d_nodes.as<NodeSlot>(), slots - 1,
			h->d_rec_slot.as<uint32_t>(), ctr + CTR_DISTINCT, ctr + CTR_OVERFLOW);
		h->stats.table_slots = slots;
		sp[x] = ctr + CTR_DISTINCT; rp[x] = h->d_xmatrix.p;
	}
	mark("table clear + insert");
	if (int rc = x_allgather(hs, sp, rp, 8)) return rc;
	mark("sizes all-gather");
	std::vector<std::vector<uint64_t>> sizes(nh, std::vector<uint64_t>(world));
	for (size_t x = 0; x < nh; x++) { mbgfe* h = hs[x]; CU(h, on(h)); CU(h, cudaMemcpyAsync(sizes[x].data(), h->d_xmatrix.p, world * 8, cudaMemcpyDeviceToHost, h->stream)); }
	for (size_t x = 0; x < nh; x++) { mbgfe* h = hs[x]; CU(h, on(h)); CU(h, cudaStreamSynchronize(h->stream)); }
	// ---- S4: global node ids of the received records, back to their senders
	for (size_t x = 0; x < nh; x++)
	{
		mbgfe* h = hs[x];
		CU(h, on(h));
		const uint32_t me = h->prm.rank;
		uint64_t base = 0, all = 0;
		h->id_begin.assign(world + 1, 0);
		for (uint32_t p = 0; p < world; p++) { if (p < me) base += sizes[x][p]; all += sizes[x][p]; h->id_begin[p + 1] = all; }
		if (all >= (1ull << 31)) return fail(h, MBGFE_E_LIMIT, "more than 2^31-1 nodes in the sharded table");
		h->stats.distinct = sizes[x][me]; h->stats.distinct_edges = 0;
		h->id_base = (uint32_t)base;
		h->exported_table = false; h->nodes_ranked = false; h->edges_global = false; h->table_sharded = true; h->have_table = true;
		h->stats.selected = h->n_inst;
		if (h->prm.flags & MBGFE_NO_EDGES) { sp[x] = nullptr; continue; }
		if (int rc = rank_nodes(h)) return rc;
		if (int rc = ensure(h, h->d_ids_out, (n_recv[x] + 1) * 4)) return rc;
		if (int rc = ensure(h, h->d_xbuf[0], (n_send[x] + 1) * 4)) return rc;
		if (n_recv[x]) MBGFE_LAUNCH(k_slot_ids, (unsigned)((n_recv[x] + 255) / 256), 256, 0, h->stream, h->d_rec_slot.as<uint32_t>(), n_recv[x], h->d_nodes.as<NodeSlot>(), h->id_base,
			h->d_ids_out.as<uint32_t>());
		sp[x] = h->d_ids_out.p; rp[x] = h->d_xbuf[0].p;
	}
	if (!(hs[0]->prm.flags & MBGFE_NO_EDGES))
	{
		mark("rank nodes + ids");
		if (int rc = x_alltoallv(hs, sp, nrecv, rp, nsend, 4)) return rc;
		mark("id all-to-all");
		// ---- S5: edge observations to the owner of their `from` node
		for (size_t x = 0; x < nh; x++)
		{
			mbgfe* h = hs[x];
			CU(h, on(h));
			if (int rc = ensure(h, h->d_edge_records, (e_send[x] + 1) * 16)) return rc;
			if (int rc = ensure(h, h->d_xbuf[1], (e_recv[x] + 1) * 16)) return rc;
			std::vector<unsigned long long> cursors(world);
			uint64_t at = 0;
			for (uint32_t p = 0; p < world; p++) { cursors[p] = at; at += esend[x][p]; }
			unsigned long long* ctr = h->d_counters.as<unsigned long long>();
			CU(h, cudaMemcpyAsync(ctr + CTR_OWNER, cursors.data(), world * 8, cudaMemcpyHostToDevice, h->stream));
			if (h->n_inst) MBGFE_LAUNCH(k_edge_owner_pack2, (unsigned)((h->n_inst + 255) / 256), 256, 0, h->stream, h->d_inst_hash.as<uint64_t>(), h->d_inst_pos.as<uint32_t>(),
				h->d_inst_flag.as<uint8_t>(), h->d_inst_packpos.as<uint32_t>(), h->d_xbuf[0].as<uint32_t>(), h->n_inst, inst_base[x], h->prm.k, world, ctr + CTR_OWNER,
				h->d_edge_records.as<unsigned long long>());
			CU(h, cudaStreamSynchronize(h->stream));
			sp[x] = h->d_edge_records.p; rp[x] = h->d_xbuf[1].p;
		}
		mark("edge pack");
		if (int rc = x_alltoallv(hs, sp, esend, rp, erecv, 16)) return rc;
		mark("edge all-to-all");
		for (size_t x = 0; x < nh; x++)
		{
			mbgfe* h = hs[x];
			CU(h, on(h));
			unsigned long long* ctr = h->d_counters.as<unsigned long long>();
			uint64_t slots = next_pow2(std::max<uint64_t>(2 * e_recv[x], 1024));
			if (sketch_wanted(e_recv[x]))
			{
				uint64_t est[2]; uint32_t dummy = 0;
				if (int rc = run_sketch(h, e_recv[x], false, est, dummy, [&](uint32_t* bm0, uint32_t*, uint64_t mask)
				{
					MBGFE_LAUNCH(k_sketch_records, (unsigned)std::min<uint64_t>((e_recv[x] + 255) / 256, (uint64_t)h->sm_count * 16), 256, 0, h->stream, h->d_xbuf[1].as<uint64_t>(), e_recv[x], 2u, 0u, bm0, mask);
				})) return rc;
				slots = std::min(slots, next_pow2(std::max<uint64_t>(2 * est[0], 1024)));
			}
			if (slots >= (1ull << 31)) return fail(h, MBGFE_E_LIMIT, "edge table would need %llu slots (limit 2^31 per GPU)", (unsigned long long)slots);
			if (int rc = ensure(h, h->d_edges, slots * sizeof(EdgeSlot))) return rc;
			h->edge_slots = slots;
			const unsigned clear_grid = (unsigned)std::min<uint64_t>((slots + 255) / 256, (uint64_t)h->sm_count * 8);
			MBGFE_LAUNCH(k_edge_clear, clear_grid, 256, 0, h->stream, h->d_edges.as<EdgeSlot>(), slots);
			if (e_recv[x]) MBGFE_LAUNCH(k_insert_edge_records, (unsigned)((e_recv[x] + 255) / 256), 256, 0, h->stream, h->d_xbuf[1].as<unsigned long long>(), e_recv[x],
				h->d_edges.as<EdgeSlot>(), slots - 1, ctr + CTR_EDGES);
			h->edges_global = true;
		}
	}
	mark("edge clear + insert");
	if (trace) { cudaEventDestroy(tr0); cudaEventDestroy(tr1); }
	// ---- the final counters
	for (size_t x = 0; x < nh; x++)
	{
		mbgfe* h = hs[x];
		CU(h, on(h));
		unsigned long long* ctr = h->d_counters.as<unsigned long long>();
		tick(h, 13);
		unsigned long long five[5] = { 0, 0, 0, 0, 0 };
		CU(h, cudaMemcpyAsync(five, ctr + CTR_DISTINCT, sizeof(five), cudaMemcpyDeviceToHost, h->stream));
		CU(h, cudaStreamSynchronize(h->stream));
		CU(h, cudaGetLastError());
		if (five[CTR_OVERFLOW - CTR_DISTINCT]) return fail(h, MBGFE_E_LIMIT, "the sharded k-mer table overflowed (%llu inserts gave up)", five[CTR_OVERFLOW - CTR_DISTINCT]);
		h->stats.distinct_edges = five[1];
		h->stats.ms_exchange = elapsed(h, 12, 13);
		h->stats.exchange_bytes = n_send[x] * 32 + n_recv[x] * 4 + e_send[x] * 16;
		h->stats.exchange_launches = 3 + (h->edge_slots ? 11 : 0);
	}
	return 0;
}

}

extern "C" {

int mbgfe_comm_unique_id(void* id128)
{
#ifndef MBGFE_EMU_BUILD
	NcclApi* n = nccl_api();
	if (!n || !id128) return MBGFE_E_CUDA;
	static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
	return n->GetUniqueId((ncclUniqueId*)id128) == ncclSuccess ? 0 : MBGFE_E_CUDA;
#else
	(void)id128; return MBGFE_E_CUDA;
#endif
}

int mbgfe_comm_init_nccl(mbgfe* h, const void* id128)
{
	if (!h || !id128) return MBGFE_E_PARAM;
#ifndef MBGFE_EMU_BUILD
	NcclApi* n = nccl_api();
	if (!n) return fail(h, MBGFE_E_CUDA, "libnccl.so.2 cannot be loaded");
	CU(h, cudaSetDevice(h->prm.device));
	ncclUniqueId id;
	memcpy(&id, id128, sizeof(id));
	ncclComm_t comm = nullptr;
	NCCLCHK(h, n->CommInitRank(&comm, (int)h->prm.world, id, (int)h->prm.rank));
	h->nccl_comm = comm;
	h->comm_kind = 1;
	return 0;
#else
	return fail(h, MBGFE_E_CUDA, "the emulator build has no NCCL");
#endif
}

int mbgfe_comm_init_callbacks(mbgfe* h, mbgfe_alltoallv_fn alltoallv, mbgfe_allgather_fn allgather, void* ctx)
{
	if (!h || !alltoallv || !allgather) return MBGFE_E_PARAM;
	h->cb_alltoallv = alltoallv; h->cb_allgather = allgather; h->cb_ctx = ctx;
	h->comm_kind = 3;
	return 0;
}

int mbgfe_exchange(mbgfe* h)
{
	if (!h) return MBGFE_E_PARAM;
	if (h->comm_kind != 1 && h->comm_kind != 3) return fail(h, MBGFE_E_STATE, "mbgfe_exchange needs mbgfe_comm_init_nccl or mbgfe_comm_init_callbacks first");
	std::vector<mbgfe*> hs { h };
	return exchange_run(hs);
}

int mbgfe_exchange_group(mbgfe** handles, uint32_t n)
{
	if (!handles || n == 0) return MBGFE_E_PARAM;
	std::vector<mbgfe*> hs(handles, handles + n);
	for (uint32_t i = 0; i < n; i++)
	{
		if (!hs[i]) return MBGFE_E_PARAM;
		if (hs[i]->prm.world != n || hs[i]->prm.rank != i) return fail(hs[i], MBGFE_E_PARAM, "mbgfe_exchange_group: handle %u must be rank %u of %u", i, i, n);
		hs[i]->comm_kind = 2;
	}
	return exchange_run(hs);
}

// Input larger than one resident batch (the reference streams any volume through its queue into one HashList,
// MBG/ReadHelper.h:102-160): the front end runs batch by batch, the selected k-mers (position, flags, 128-bit hash: 25 bytes each,
// a few per thousand bases) of every batch stay on the device, and the tables are built once over all of them.
static int append_call(mbgfe* h, int (*what)(mbgfe*), bool fresh_batch)
{
	if (!h) return MBGFE_E_PARAM;
	if (h->n_var) return fail(h, MBGFE_E_STATE, "append runs with hpc variants registered: not supported");
	// mbgfe_process_append after mbgfe_upload: the upload cleared the results of the batch before but not its instances
	if (fresh_batch) h->append_base = h->processed ? h->n_inst : 0;
	h->appending = true;
	if (int rc = ensure(h, h->d_append_base, 64)) { h->appending = false; return rc; }
	cudaMemcpyAsync(h->d_append_base.p, &h->append_base, 8, cudaMemcpyHostToDevice, h->stream);
	cudaStreamSynchronize(h->stream);
	const int rc = what(h);
	h->appending = false;
	return rc;
}
int mbgfe_run_append(mbgfe* h) { return append_call(h, mbgfe_run, true); }
// resident form: mbgfe_upload_append() brings the next batch in (the instances of the batches before it stay), mbgfe_process_append()
// runs the front end on it and may be repeated (each repetition replaces this batch's instances, not the earlier ones)
int mbgfe_upload_append(mbgfe* h)
{
	if (!h) return MBGFE_E_PARAM;
	h->append_base = h->processed ? h->n_inst : 0;
	const uint64_t keep = h->append_base;
	const int rc = mbgfe_upload(h);
	h->append_base = keep;
	return rc;
}
int mbgfe_process_append(mbgfe* h) { return append_call(h, mbgfe_process, false); }

int mbgfe_build_table(mbgfe* h)
{
	if (!h) return MBGFE_E_PARAM;
	if (!h->processed) return fail(h, MBGFE_E_STATE, "mbgfe_build_table before a run");
	if (h->prm.flags & MBGFE_NO_TABLE) return fail(h, MBGFE_E_STATE, "mbgfe_build_table on a handle created with MBGFE_NO_TABLE (sharded tables are built by mbgfe_exchange)");
	uint32_t launches = 0;
	unsigned long long* ctr = h->d_counters.as<unsigned long long>();
	const uint64_t n_inst = h->n_inst;
	if (int rc = ensure(h, h->d_range, 64)) return rc;
	const uint64_t range[2] = { 0, n_inst };
	CU(h, cudaMemcpyAsync(h->d_range.p, range, 16, cudaMemcpyHostToDevice, h->stream));
	if (int rc = ensure(h, h->d_inst_slot, (n_inst + 1) * 4 + 64)) return rc;
	tick(h, 7);
	uint64_t node_slots, edge_slots, full;
	if (int rc = size_tables(h, n_inst, &node_slots, &edge_slots, &full, launches)) return rc;
	const unsigned grid = (unsigned)std::max<uint64_t>(1, (n_inst + 255) / 256);
	const uint64_t* rg = h->d_range.as<uint64_t>();
	for (int attempt = 0; ; attempt++)
	{
		CU(h, cudaMemsetAsync(ctr + CTR_DISTINCT, 0, (CTR_OVERFLOW - CTR_DISTINCT + 1) * 8, h->stream));
		if (int rc = clear_tables(h, node_slots, edge_slots, launches)) return rc;
		if (n_inst)
		{
			MBGFE_LAUNCH(k_insert_nodes, grid, 256, 0, h->stream, h->d_inst_hash.as<uint64_t>(), h->d_inst_flag.as<uint8_t>(), rg, rg + 1, h->inst_cap, h->d_nodes.as<NodeSlot>(),
				h->node_slots - 1, h->d_inst_slot.as<uint32_t>(), ctr + CTR_DISTINCT, ctr + CTR_OVERFLOW);
			launches++;
		}
		tick(h, 8);
		if (n_inst && !(h->prm.flags & MBGFE_NO_EDGES))
		{
			MBGFE_LAUNCH(k_insert_edges, grid, 256, 0, h->stream, h->d_inst_hash.as<uint64_t>(), h->d_inst_pos.as<uint32_t>(), h->d_inst_flag.as<uint8_t>(), h->d_inst_slot.as<uint32_t>(),
				rg, rg + 1, h->inst_cap, h->d_edges.as<EdgeSlot>(), h->edge_slots - 1, ctr + CTR_EDGES, ctr + CTR_OVERFLOW, h->prm.k);
			launches++;
		}
		tick(h, 9);
		unsigned long long five[5] = { 0, 0, 0, 0, 0 };
		CU(h, cudaMemcpyAsync(five, ctr + CTR_DISTINCT, sizeof(five), cudaMemcpyDeviceToHost, h->stream));
		CU(h, cudaStreamSynchronize(h->stream));
		CU(h, cudaGetLastError());
		h->stats.distinct = five[0]; h->stats.distinct_edges = five[1];
		if (!five[CTR_OVERFLOW - CTR_DISTINCT]) break;
		if ((node_slots >= full && edge_slots >= full) || attempt == 4) return fail(h, MBGFE_E_LIMIT, "k-mer table of %llu slots overflowed", (unsigned long long)node_slots);
		node_slots = std::min(full, 2 * node_slots); edge_slots = std::min(full, 2 * edge_slots);
		h->stats.table_grows++;
	}
	h->stats.table_slots = node_slots;
	h->stats.ms_nodes = elapsed(h, 7, 8); h->stats.ms_edges = elapsed(h, 8, 9);
	h->stats.launches = launches;
	h->have_table = true; h->exported_table = false; h->nodes_ranked = false; h->edges_global = false; h->table_sharded = false; h->id_base = 0;
	h->have_unitigs = false; h->have_thin = false; h->have_threads = false;
	return 0;
}

int mbgfe_set_variants(mbgfe* h, const mbgfe_hash128* hashes, const uint32_t* offsets, const uint64_t* len_begin, const uint32_t* lengths, uint64_t n_entries)
{
	if (!h) return MBGFE_E_PARAM;
	if (n_entries && (!hashes || !offsets || !len_begin || !lengths)) return fail(h, MBGFE_E_PARAM, "mbgfe_set_variants: null array");
	h->n_var = 0;
	if (n_entries == 0) return 0;
	for (uint64_t e = 0; e < n_entries; e++)
	{
		if (offsets[e] >= h->prm.k) return fail(h, MBGFE_E_PARAM, "variant offset %u outside the k-mer", offsets[e]);
		if (len_begin[e + 1] <= len_begin[e]) return fail(h, MBGFE_E_PARAM, "variant entry %llu has no lengths", (unsigned long long)e);
	}
	// what addHpcVariants leaves (MBG/ReadHelper.cpp:63-70): per hash the entries ordered by offset, lengths ascending
	std::vector<uint64_t> order(n_entries);
	for (uint64_t e = 0; e < n_entries; e++) order[e] = e;
	std::sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b)
	{
		if (hashes[a].hi != hashes[b].hi) return hashes[a].hi < hashes[b].hi;
		if (hashes[a].lo != hashes[b].lo) return hashes[a].lo < hashes[b].lo;
		if (offsets[a] != offsets[b]) return offsets[a] < offsets[b];
		return a < b;
	});
	std::vector<uint64_t> vh(2 * n_entries), vb(n_entries + 1);
	std::vector<uint32_t> vo(n_entries), vl;
	vl.reserve(len_begin[n_entries]);
	for (uint64_t e = 0; e < n_entries; e++)
	{
		const uint64_t src = order[e];
		vh[2 * e] = hashes[src].lo; vh[2 * e + 1] = hashes[src].hi; vo[e] = offsets[src];
		vb[e] = vl.size();
		vl.insert(vl.end(), lengths + len_begin[src], lengths + len_begin[src + 1]);
		std::sort(vl.begin() + vb[e], vl.end());
	}
	vb[n_entries] = vl.size();
	if (int rc = ensure(h, h->d_var_hash, vh.size() * 8)) return rc;
	if (int rc = ensure(h, h->d_var_offset, vo.size() * 4)) return rc;
	if (int rc = ensure(h, h->d_var_len_begin, vb.size() * 8)) return rc;
	if (int rc = ensure(h, h->d_var_len, vl.size() * 4)) return rc;
	CU(h, cudaMemcpyAsync(h->d_var_hash.p, vh.data(), vh.size() * 8, cudaMemcpyHostToDevice, h->stream));
	CU(h, cudaMemcpyAsync(h->d_var_offset.p, vo.data(), vo.size() * 4, cudaMemcpyHostToDevice, h->stream));
	CU(h, cudaMemcpyAsync(h->d_var_len_begin.p, vb.data(), vb.size() * 8, cudaMemcpyHostToDevice, h->stream));
	CU(h, cudaMemcpyAsync(h->d_var_len.p, vl.data(), vl.size() * 4, cudaMemcpyHostToDevice, h->stream));
	CU(h, cudaStreamSynchronize(h->stream));
	h->n_var = n_entries;
	return 0;
}

int mbgfe_kat_tables(mbgfe* h, uint64_t* char_hash_30948, uint16_t* complement_30948)
{
	if (!h) return MBGFE_E_PARAM;
	if (char_hash_30948) CU(h, cudaMemcpy(char_hash_30948, h->tab_hash.p, MAX_CODE * 8, cudaMemcpyDeviceToHost));
	if (complement_30948) CU(h, cudaMemcpy(complement_30948, h->tab_comp.p, MAX_CODE * 2, cudaMemcpyDeviceToHost));
	return 0;
}

int mbgfe_kat_kmer_hash(mbgfe* h, const uint16_t* codes, uint32_t k, mbgfe_hash128* out)
{
	if (!h || !codes || !out) return MBGFE_E_PARAM;
	if (k == 0) return fail(h, MBGFE_E_PARAM, "k must be positive");
	DevBuf d_codes, d_misc;
	int rc = ensure(h, d_codes, (size_t)k * 2 + 256);
	if (!rc) rc = ensure(h, d_misc, 128);
	if (rc) { release(d_codes); release(d_misc); return rc; }
	// one synthetic instance: a single segment whose slot holds the k codes (slots are multiples of 64 codes), position 0
	const uint64_t bases[2] = { 0, ((uint64_t)k + 1 + 63) & ~63ull };
	const uint32_t zero32 = 0;
	cudaMemsetAsync(d_codes.p, 0, d_codes.cap, h->stream);
	cudaMemcpyAsync(d_codes.p, codes, (size_t)k * 2, cudaMemcpyHostToDevice, h->stream);
	cudaMemcpyAsync(d_misc.as<char>() + 0, bases, 16, cudaMemcpyHostToDevice, h->stream);   // seg_out_base[0..1]
	cudaMemcpyAsync(d_misc.as<char>() + 16, &zero32, 4, cudaMemcpyHostToDevice, h->stream);  // inst_pos[0]
	const uint64_t range[2] = { 0, 1 };
	cudaMemcpyAsync(d_misc.as<char>() + 64, range, 16, cudaMemcpyHostToDevice, h->stream);   // instance range [0, 1)
	launch_kmer_hash(h->stream, 1, (const uint64_t*)d_misc.p, d_codes.as<uint16_t>(), h->tab_comp.as<uint16_t>(), (const uint32_t*)(d_misc.as<char>() + 16),
		(const uint64_t*)(d_misc.as<char>() + 64), (uint64_t)1, (uint64_t*)(d_misc.as<char>() + 32), k);
	cudaStreamSynchronize(h->stream);
	uint64_t res[2] = { 0, 0 };
	cudaError_t e = cudaMemcpy(res, d_misc.as<char>() + 32, 16, cudaMemcpyDeviceToHost);
	release(d_codes); release(d_misc);
	if (e != cudaSuccess) return fail(h, MBGFE_E_CUDA, "kat kernel failed: %s", cudaGetErrorString(e));
	out->lo = res[0]; out->hi = res[1];
	return 0;
}

// ---- mbgfe_kat_kmer_hash_noalloc: the same k-mer hash with no allocation of any kind (no counterpart in the reference;
// added for callers that cannot call cudaMalloc: CUDA graph capture, tight serving
// loops, harnesses that forbid it).  The staging that mbgfe_kat_kmer_hash allocates per call -- the code slot, the
// one-segment table, the instance and the result -- lives in this file-scope array instead: NOALLOC_SCRATCH_BYTES on every
// GPU that loads the library, until the process exits.  In host code `noalloc_scratch` is only a stand-in for the symbol:
// the device address comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr uint32_t NOALLOC_MISC_BYTES = 256;              // seg_out_base[2] @0, inst_pos @16, result @32, instance range @64
constexpr size_t NOALLOC_SCRATCH_BYTES = NOALLOC_MISC_BYTES + (((size_t)MBGFE_NOALLOC_MAX_K + 1 + 63) & ~(size_t)63) * 2;
static __device__ __align__(256) unsigned char noalloc_scratch[NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex noalloc_lock;
static unsigned char* noalloc_address[NOALLOC_MAX_DEVICES];      // per device, filled on first use

int mbgfe_kat_kmer_hash_noalloc(mbgfe* h, const uint16_t* codes, uint32_t k, mbgfe_hash128* out)
{
	if (!h || !codes || !out) return MBGFE_E_PARAM;
	if (k == 0) return fail(h, MBGFE_E_PARAM, "k must be positive");
	if (k > MBGFE_NOALLOC_MAX_K) return fail(h, MBGFE_E_LIMIT, "mbgfe_kat_kmer_hash_noalloc: k = %u exceeds MBGFE_NOALLOC_MAX_K = %u (the static scratch holds no more); use mbgfe_kat_kmer_hash",
		k, (unsigned)MBGFE_NOALLOC_MAX_K);
	const int dev = h->prm.device;
	if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail(h, MBGFE_E_LIMIT, "mbgfe_kat_kmer_hash_noalloc: device %d outside 0..%d", dev, NOALLOC_MAX_DEVICES - 1);
	CU(h, cudaSetDevice(dev));
	unsigned char* scratch = nullptr;
	{
		std::lock_guard<std::mutex> guard(noalloc_lock);
		if (!noalloc_address[dev])
		{
			void* p_ = nullptr;
			const cudaError_t e = cudaGetSymbolAddress((void**)&p_, noalloc_scratch);
			if (e != cudaSuccess || !p_) return fail(h, MBGFE_E_CUDA, "cudaGetSymbolAddress(noalloc_scratch) on device %d: %s", dev, cudaGetErrorString(e));
			noalloc_address[dev] = (unsigned char*)p_;
		}
		scratch = noalloc_address[dev];
	}
	// the same synthetic instance as above: one segment whose slot holds the k codes, position 0.  The array keeps what the
	// last call left, and K3b reads whole aligned 16-byte words of the slot: clear the slot (and the result) first.
	const size_t slot_codes = ((size_t)k + 1 + 63) & ~(size_t)63;
	uint16_t* d_codes = (uint16_t*)(scratch + NOALLOC_MISC_BYTES);
	const uint64_t bases[2] = { 0, slot_codes };
	const uint32_t zero32 = 0;
	const uint64_t range[2] = { 0, 1 };
	CU(h, cudaMemsetAsync(scratch, 0, NOALLOC_MISC_BYTES + slot_codes * 2, h->stream));
	CU(h, cudaMemcpyAsync(d_codes, codes, (size_t)k * 2, cudaMemcpyHostToDevice, h->stream));
	CU(h, cudaMemcpyAsync(scratch + 0, bases, 16, cudaMemcpyHostToDevice, h->stream));    // seg_out_base[0..1]
	CU(h, cudaMemcpyAsync(scratch + 16, &zero32, 4, cudaMemcpyHostToDevice, h->stream));  // inst_pos[0]
	CU(h, cudaMemcpyAsync(scratch + 64, range, 16, cudaMemcpyHostToDevice, h->stream));   // instance range [0, 1)
	launch_kmer_hash(h->stream, 1, (const uint64_t*)scratch, (const uint16_t*)d_codes, h->tab_comp.as<uint16_t>(), (const uint32_t*)(scratch + 16),
		(const uint64_t*)(scratch + 64), (uint64_t)1, (uint64_t*)(scratch + 32), k);
	CU(h, cudaGetLastError());
	uint64_t res[2] = { 0, 0 };
	CU(h, cudaMemcpyAsync(res, scratch + 32, 16, cudaMemcpyDeviceToHost, h->stream));
	CU(h, cudaStreamSynchronize(h->stream));
	out->lo = res[0]; out->hi = res[1];
	return 0;
}

}
