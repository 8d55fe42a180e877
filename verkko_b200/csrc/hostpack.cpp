// hostpack.cpp -- the host half of the read loader: ASCII bases -> planar 2-bit words, the layout K0 (k_pack_ascii in
// kernels.cuh) writes on the device.  mbgfe_run routes part of the pieces of a batch through here on worker threads while
// the rest crosses PCIe as ASCII: a packed piece costs a quarter of the PCIe time, so host cores and the copy engine share
// the upload (north_star subsystem 1: "2-bit-packed read loader with pinned-host-to-HBM streaming").
// Same mapping as the reference's read parser feeds its masks with (MBG/ReadHelper.h:899-1018: A0 C1 G2 T3, either case);
// characters outside ACGTacgt pack as A and are reported, exactly like K0.
// Plain host C++ (compiled by g++, linked into libmbgfe.so); AVX2 when the CPU has it, portable code otherwise.
#include "hostpack.h"
#include <algorithm>
#include <unistd.h>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace mbgfe_host {

static inline bool is_acgt(unsigned char c)
{
	const unsigned char u = c & 0xdf;
	return u == 'A' || u == 'C' || u == 'G' || u == 'T';
}

// one word (up to 32 characters), portable
static inline uint64_t pack_word_scalar(const char* src, uint32_t n, uint64_t pos, std::vector<uint64_t>* bad)
{
	uint32_t lo = 0, hi = 0;
	for (uint32_t i = 0; i < n; i++)
	{
		unsigned char c = (unsigned char)src[i];
		if (!is_acgt(c)) { if (bad) bad->push_back(pos + i); c = 'A'; }
		const uint32_t t = ((c >> 1) ^ (c >> 2)) & 3;       // A0 C1 G2 T3
		lo |= (t & 1) << i; hi |= (t >> 1) << i;
	}
	return ((uint64_t)hi << 32) | lo;
}

#if defined(__x86_64__)
__attribute__((target("avx2"))) static void pack_avx2(const char* src, uint64_t n_words, uint64_t* dst, uint64_t pos, std::vector<uint64_t>* bad)
{
	const __m256i df = _mm256_set1_epi8((char)0xdf);
	const __m256i A = _mm256_set1_epi8('A'), C = _mm256_set1_epi8('C'), G = _mm256_set1_epi8('G'), T = _mm256_set1_epi8('T');
	for (uint64_t i = 0; i < n_words; i++)
	{
		const __m256i x = _mm256_loadu_si256((const __m256i*)(src + 32 * i));
		const __m256i u = _mm256_and_si256(x, df);
		const __m256i ok = _mm256_or_si256(_mm256_or_si256(_mm256_cmpeq_epi8(u, A), _mm256_cmpeq_epi8(u, C)), _mm256_or_si256(_mm256_cmpeq_epi8(u, G), _mm256_cmpeq_epi8(u, T)));
		if ((uint32_t)_mm256_movemask_epi8(ok) != 0xffffffffu) { dst[i] = pack_word_scalar(src + 32 * i, 32, pos + 32 * i, bad); continue; }
		const __m256i t = _mm256_xor_si256(_mm256_srli_epi16(x, 1), _mm256_srli_epi16(x, 2));     // bits 0,1 of every byte = the 2-bit base
		const uint32_t lo = (uint32_t)_mm256_movemask_epi8(_mm256_slli_epi16(t, 7));
		const uint32_t hi = (uint32_t)_mm256_movemask_epi8(_mm256_slli_epi16(t, 6));
		dst[i] = ((uint64_t)hi << 32) | lo;
	}
}
#endif

void pack_planar(const char* src, uint64_t n, uint64_t* dst, uint64_t pos, std::vector<uint64_t>* bad)
{
	const uint64_t whole = n / 32;
#if defined(__x86_64__)
	static const bool have_avx2 = __builtin_cpu_supports("avx2");
	if (have_avx2) pack_avx2(src, whole, dst, pos, bad);
	else
#endif
	for (uint64_t i = 0; i < whole; i++) dst[i] = pack_word_scalar(src + 32 * i, 32, pos + 32 * i, bad);
	if (n % 32) dst[whole] = pack_word_scalar(src + 32 * whole, (uint32_t)(n % 32), pos + 32 * whole, bad);
}

PackPool::~PackPool() { finish(); }

void PackPool::start(std::vector<PackTask>&& t, size_t n_pieces, unsigned n_threads)
{
	tasks = std::move(t);
	left.reset(new std::atomic<uint32_t>[n_pieces ? n_pieces : 1]);
	for (size_t i = 0; i < n_pieces; i++) left[i].store(0, std::memory_order_relaxed);
	for (const PackTask& k : tasks) left[k.piece].fetch_add(1, std::memory_order_relaxed);
	next.store(0);
	for (unsigned i = 0; i < n_threads && !tasks.empty(); i++)
		workers.emplace_back([this]()
		{
			std::vector<uint64_t> mine;
			for (;;)
			{
				const size_t j = next.fetch_add(1);
				if (j >= tasks.size()) break;
				const PackTask& k = tasks[j];
				pack_planar(k.src, k.n, k.dst, k.pos, &mine);
				left[k.piece].fetch_sub(1, std::memory_order_release);
			}
			if (!mine.empty()) { std::lock_guard<std::mutex> g(lock); bad.insert(bad.end(), mine.begin(), mine.end()); }
		});
}

void PackPool::wait_piece(size_t piece)
{
	while (left[piece].load(std::memory_order_acquire) != 0)
	{
#if defined(__x86_64__)
		_mm_pause();
#endif
	}
}

void PackPool::finish()
{
	for (auto& w : workers) w.join();
	workers.clear();
}


// ------------------------------------------------------------------------------------------- file tasks

namespace {

struct ScanState { uint64_t line_start; bool in_header; };      // line_start: index in the task of the current line's first byte, ~0 = unknown

// one word (up to 32 bytes) that holds something other than ACGTacgt
inline uint64_t scan_word(const char* src, uint32_t n, uint64_t idx0, uint64_t pos0, ScanState& st, std::vector<uint64_t>* events)
{
	uint32_t lo = 0, hi = 0;
	for (uint32_t i = 0; i < n; i++)
	{
		unsigned char c = (unsigned char)src[i];
		if (!is_acgt(c))
		{
			const uint64_t idx = idx0 + i;
			if (c == '\n') { events->push_back(((pos0 + i) << 8) | c); st.in_header = false; st.line_start = idx + 1; }
			else if (st.in_header) { }
			else if (c == '>' && st.line_start == idx) { events->push_back(((pos0 + i) << 8) | c); st.in_header = true; }
			else events->push_back(((pos0 + i) << 8) | c);
			c = 'A';
		}
		const uint32_t t = ((c >> 1) ^ (c >> 2)) & 3;
		lo |= (t & 1) << i; hi |= (t >> 1) << i;
	}
	return ((uint64_t)hi << 32) | lo;
}

#if defined(__x86_64__)
__attribute__((target("avx2"))) void scan_avx2(const char* src, uint64_t n_words, uint64_t* dst, uint64_t pos, ScanState& st, std::vector<uint64_t>* events)
{
	const __m256i df = _mm256_set1_epi8((char)0xdf);
	const __m256i A = _mm256_set1_epi8('A'), C = _mm256_set1_epi8('C'), G = _mm256_set1_epi8('G'), T = _mm256_set1_epi8('T');
	for (uint64_t i = 0; i < n_words; i++)
	{
		const __m256i x = _mm256_loadu_si256((const __m256i*)(src + 32 * i));
		const __m256i u = _mm256_and_si256(x, df);
		const __m256i ok = _mm256_or_si256(_mm256_or_si256(_mm256_cmpeq_epi8(u, A), _mm256_cmpeq_epi8(u, C)), _mm256_or_si256(_mm256_cmpeq_epi8(u, G), _mm256_cmpeq_epi8(u, T)));
		if ((uint32_t)_mm256_movemask_epi8(ok) != 0xffffffffu) { dst[i] = scan_word(src + 32 * i, 32, 32 * i, pos + 32 * i, st, events); continue; }
		const __m256i t = _mm256_xor_si256(_mm256_srli_epi16(x, 1), _mm256_srli_epi16(x, 2));
		const uint32_t lo = (uint32_t)_mm256_movemask_epi8(_mm256_slli_epi16(t, 7));
		const uint32_t hi = (uint32_t)_mm256_movemask_epi8(_mm256_slli_epi16(t, 6));
		dst[i] = ((uint64_t)hi << 32) | lo;
	}
}
#endif

}

void pack_scan_fasta(const char* src, uint64_t n, uint64_t pos, bool starts_line, uint64_t* dst, std::vector<uint64_t>* events)
{
	ScanState st { starts_line ? 0 : ~0ull, false };
	const uint64_t whole = n / 32;
#if defined(__x86_64__)
	static const bool have_avx2 = __builtin_cpu_supports("avx2");
	if (have_avx2) scan_avx2(src, whole, dst, pos, st, events);
	else
#endif
	for (uint64_t i = 0; i < whole; i++)
	{
		bool plain = true;
		for (uint32_t j = 0; j < 32 && plain; j++) plain = is_acgt((unsigned char)src[32 * i + j]);
		dst[i] = plain ? pack_word_scalar(src + 32 * i, 32, 0, nullptr) : scan_word(src + 32 * i, 32, 32 * i, pos + 32 * i, st, events);
	}
	if (n % 32) dst[whole] = scan_word(src + 32 * whole, (uint32_t)(n % 32), 32 * whole, pos + 32 * whole, st, events);
}

FilePool::~FilePool() { finish(); }

void FilePool::start(std::vector<FileTask>&& t, size_t n_pieces, unsigned n_threads)
{
	tasks = std::move(t);
	left.reset(new std::atomic<uint32_t>[n_pieces ? n_pieces : 1]);
	for (size_t i = 0; i < n_pieces; i++) left[i].store(0, std::memory_order_relaxed);
	uint64_t largest = 0;
	for (const FileTask& k : tasks) { left[k.piece].fetch_add(1, std::memory_order_relaxed); largest = std::max(largest, k.n); }
	next.store(0);
	for (unsigned i = 0; i < std::max(1u, n_threads) && !tasks.empty(); i++)
		workers.emplace_back([this, largest]()
		{
			// the task's bytes plus the one before them (is the first byte a line start?) land in a buffer the thread keeps
			std::vector<char> buf(largest + 96);
			for (;;)
			{
				const size_t j = next.fetch_add(1);
				if (j >= tasks.size()) break;
				FileTask& k = tasks[j];
				const uint64_t lead = k.file_off ? 1 : 0;
				char* at = buf.data() + 32;
				uint64_t got = 0;
				while (got < k.n + lead)
				{
					const ssize_t r = pread(k.fd, at - lead + got, k.n + lead - got, (off_t)(k.file_off - lead + got));
					if (r <= 0) { k.error = 1; break; }
					got += (uint64_t)r;
				}
				if (!k.error) pack_scan_fasta(at, k.n, k.pos, lead == 0 || at[-1] == '\n', k.dst, &k.events);
				left[k.piece].fetch_sub(1, std::memory_order_release);
			}
		});
}

void FilePool::wait_piece(size_t piece)
{
	while (left[piece].load(std::memory_order_acquire) != 0)
	{
#if defined(__x86_64__)
		_mm_pause();
#endif
	}
}

void FilePool::finish()
{
	for (auto& w : workers) w.join();
	workers.clear();
}

}
