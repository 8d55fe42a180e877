// synth.cpp -- deterministic synthetic genome / HiFi-shaped read generator (host only, multi-threaded).
//
// There is no network for datasets, so every workload BASELINE.json names is synthesised:
//   genome = uniform ACGT, then loci overwritten with homopolymer runs and microsatellites (motif 1..6);
//   reads  = uniform start, random strand, span length uniform in [len_min, len_max], sequencing errors at
//            `error_ppm` per base: 60 % homopolymer-length changes, 20 % other indels, 20 % substitutions.
// Every read is a pure function of (seed, read index), so the CPU baseline, the oracle and the CUDA path can each
// regenerate exactly the same reads without sharing files.  Used by bench.py and the tests; not on the hot path.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {

inline uint64_t mix64(uint64_t x)
{
	x += 0x9e3779b97f4a7c15ull;
	x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
	x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
	return x ^ (x >> 31);
}

struct Rng
{
	uint64_t state;
	explicit Rng(uint64_t seed) : state(seed) {}
	uint64_t next() { state += 0x9e3779b97f4a7c15ull; uint64_t x = state; x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull; x = (x ^ (x >> 27)) * 0x94d049bb133111ebull; return x ^ (x >> 31); }
	uint64_t below(uint64_t n) { return n ? (uint64_t)(((unsigned __int128)next() * n) >> 64) : 0; }
};

const char BASES[5] = "ACGT";

inline char complement_base(char c)
{
	switch (c) { case 'A': return 'T'; case 'C': return 'G'; case 'G': return 'C'; case 'T': return 'A'; }
	return c;
}

template <typename F>
void parallel_for(uint64_t n, unsigned threads, F body)
{
	if (threads <= 1 || n < 2) { body(0, n); return; }
	std::vector<std::thread> pool;
	uint64_t chunk = (n + threads - 1) / threads;
	for (unsigned t = 0; t < threads; t++)
	{
		uint64_t lo = std::min(n, t * chunk), hi = std::min(n, lo + chunk);
		if (lo < hi) pool.emplace_back(body, lo, hi);
	}
	for (auto& th : pool) th.join();
}

struct ReadSpec { uint64_t start; uint32_t span; bool reverse; uint64_t rng_seed; };

ReadSpec read_spec(uint64_t seed, uint64_t r, uint64_t genome_len, uint32_t len_min, uint32_t len_max)
{
	Rng rng(mix64(seed ^ mix64(r * 0x632be59bd9b4e019ull + 0x1234)));
	ReadSpec s;
	uint32_t span = len_min + (uint32_t)rng.below((uint64_t)len_max - len_min + 1);
	if (span > genome_len) span = (uint32_t)genome_len;
	s.span = span;
	s.start = rng.below(genome_len - span + 1);
	s.reverse = rng.next() & 1;
	s.rng_seed = rng.next();
	return s;
}

// writes the read (or only measures it when out == nullptr); returns its length
uint64_t make_read(const char* genome, const ReadSpec& spec, uint32_t error_ppm, char* out)
{
	Rng rng(spec.rng_seed);
	uint64_t n = 0;
	auto put = [&](char c) { if (out) out[n] = c; n++; };
	auto base_at = [&](uint32_t i) -> char
	{
		return spec.reverse ? complement_base(genome[spec.start + spec.span - 1 - i]) : genome[spec.start + i];
	};
	for (uint32_t i = 0; i < spec.span; i++)
	{
		char c = base_at(i);
		if (error_ppm == 0 || rng.below(1000000) >= error_ppm) { put(c); continue; }
		uint64_t kind = rng.below(10);
		if (kind < 6)
		{
			// homopolymer-length error: lengthen the run by one, or shorten it when it is a run
			bool in_run = i + 1 < spec.span && base_at(i + 1) == c;
			if (in_run && (rng.next() & 1)) continue;
			put(c); put(c);
		}
		else if (kind < 8)
		{
			if (rng.next() & 1) { put(c); put(BASES[rng.below(4)]); }
			// else: deletion
		}
		else
		{
			char sub = BASES[rng.below(4)];
			if (sub == c) sub = BASES[((uint64_t)(strchr(BASES, c) - BASES) + 1 + rng.below(3)) % 4];
			put(sub);
		}
	}
	return n;
}

}

extern "C" {

// Fills genome[0..genome_len) with ASCII ACGT.  `locus_spacing` = average distance between repeat loci (0 = none).
void mbgsynth_genome(uint64_t seed, uint64_t genome_len, uint32_t locus_spacing, char* genome, uint32_t threads)
{
	parallel_for(genome_len, threads, [&](uint64_t lo, uint64_t hi)
	{
		for (uint64_t i = lo; i < hi; i += 32)
		{
			uint64_t bits = mix64(seed ^ mix64(i / 32));
			for (uint64_t j = i; j < std::min(hi, i + 32); j++) { genome[j] = BASES[bits & 3]; bits >>= 2; }
		}
	});
	if (locus_spacing == 0) return;
	uint64_t n_loci = genome_len / locus_spacing;
	parallel_for(n_loci, threads, [&](uint64_t lo, uint64_t hi)
	{
		for (uint64_t j = lo; j < hi; j++)
		{
			Rng rng(mix64(seed ^ 0xabcdef ^ mix64(j)));
			uint64_t pos = j * locus_spacing + rng.below(locus_spacing / 2 + 1);
			uint32_t motif_len, copies;
			if (rng.below(100) < 70)
			{
				motif_len = 1; // homopolymer run: geometric, mean ~3, tail to 20
				copies = 2;
				while (copies < 20 && rng.below(3) != 0) copies++;
			}
			else
			{
				motif_len = 2 + (uint32_t)rng.below(5);
				copies = 3 + (uint32_t)rng.below(18);
			}
			char motif[6];
			for (uint32_t m = 0; m < motif_len; m++) motif[m] = BASES[rng.below(4)];
			uint32_t overhang = motif_len > 1 ? (uint32_t)rng.below(motif_len) : 0;
			uint64_t total = (uint64_t)motif_len * copies + overhang;
			// loci never cross the next locus' window start, so threads write disjoint ranges
			uint64_t limit = std::min(genome_len, (j + 1) * locus_spacing);
			for (uint64_t t = 0; t < total && pos + t < limit; t++) genome[pos + t] = motif[t % motif_len];
		}
	});
}

// lengths[r] = length of read r after errors (call first, then prefix-sum into offsets and call mbgsynth_reads).
void mbgsynth_read_lengths(uint64_t seed, const char* genome, uint64_t genome_len, uint64_t first_read, uint64_t n_reads,
	uint32_t len_min, uint32_t len_max, uint32_t error_ppm, uint64_t* lengths, uint32_t threads)
{
	parallel_for(n_reads, threads, [&](uint64_t lo, uint64_t hi)
	{
		for (uint64_t r = lo; r < hi; r++)
		{
			ReadSpec spec = read_spec(seed, first_read + r, genome_len, len_min, len_max);
			lengths[r] = error_ppm ? make_read(genome, spec, error_ppm, nullptr) : spec.span;
		}
	});
}

void mbgsynth_reads(uint64_t seed, const char* genome, uint64_t genome_len, uint64_t first_read, uint64_t n_reads,
	uint32_t len_min, uint32_t len_max, uint32_t error_ppm, const uint64_t* offsets, char* bases, uint32_t threads)
{
	parallel_for(n_reads, threads, [&](uint64_t lo, uint64_t hi)
	{
		for (uint64_t r = lo; r < hi; r++)
		{
			ReadSpec spec = read_spec(seed, first_read + r, genome_len, len_min, len_max);
			make_read(genome, spec, error_ppm, bases + offsets[r]);
		}
	});
}

}
