// kernels.cuh -- sm_100a device code of libmbgfe (hand-written CUDA; no library kernels).
//
// Data layout in HBM (all indices are in bases / symbols / codes, never bytes, unless stated):
//   packed   2-bit bases of the whole batch in PLANAR 64-bit words: word i/32 holds bases 32(i/32)..+31, the low bit of
//            base i at bit i%32 and its high bit at bit 32 + i%32 (A0 C1 G2 T3).  One bit per position in each 32-bit
//            plane keeps every later mask operation (neighbour compares, period tests, compaction) 32 bits wide.
//            Reads are NOT padded: read r occupies [offsets[r], offsets[r+1]) exactly as in the ASCII buffer.
//   segment  maximal ACGT stretch of a read (the whole read when it has no other characters).  Segment g owns the
//            output range [out_base[g], out_base[g] + len[g] + 1) in every per-symbol array; out_base is a multiple
//            of 64, so each segment starts on a 128-bit boundary in the 2-bit symbol array and on a 128-byte boundary
//            in the 16-bit code array.
//   sym      2-bit run-length / collapsed symbols (output of K1), same planar packing as `packed`, indexed by out_base.
//   codes    uint16 final codes (output of K2), indexed by out_base.
//   selmask  1 bit per code position: k-mer starting there is selected (output of K3a), indexed by out_base.
//   inst_*   one entry per selected k-mer instance, in (segment, position) order (output of K3b).
//
// Reference citations: MBG/ = /root/reference/src/hifioverlapper/MBG/src (commit 7bef84dd).
#pragma once
#include <cstdint>
#ifdef MBGFE_EMU_BUILD
#include "cuda_emu.h"   // tests/emu: GPU-less debugging build, never shipped (see that header)
#else
#include <cuda_runtime.h>
#define MBGFE_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#define MBGFE_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#endif

namespace mbgfe_dev {

constexpr uint32_t FULL = 0xffffffffu;

__host__ __device__ __forceinline__ uint64_t rotl64(uint64_t v, uint32_t r) { r &= 63; return (v << r) | (v >> ((64 - r) & 63)); }
__host__ __device__ __forceinline__ uint64_t rotr64(uint64_t v, uint32_t r) { r &= 63; return (v >> r) | (v << ((64 - r) & 63)); }

struct Planes { uint32_t lo, hi; };    // 32 symbols: bit i of lo / hi = low / high bit of symbol i
__device__ __forceinline__ Planes planes_of(uint64_t w) { Planes p; p.lo = (uint32_t)w; p.hi = (uint32_t)(w >> 32); return p; }
__device__ __forceinline__ uint64_t word_of(uint32_t lo, uint32_t hi) { return ((uint64_t)hi << 32) | lo; }
// low n bits set (n in 0..32)
// (one clamped funnel shift on the device: the high word of 0x00000000ffffffff << min(n, 32))
__device__ __forceinline__ uint32_t low_bits(uint32_t n)
{
#ifdef MBGFE_EMU_BUILD
	return n >= 32 ? 0xffffffffu : ((1u << n) - 1);
#else
	return __funnelshift_lc(0xffffffffu, 0u, n);
#endif
}
// low max(a - b, 0) bits set
__device__ __forceinline__ uint32_t low_bits_sat(uint32_t a, uint32_t b) { return low_bits(max(a, b) - b); }
// (hi:lo) >> sh, low 32 bits (sh in 0..31)
__device__ __forceinline__ uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh)
{
#ifdef MBGFE_EMU_BUILD
	return sh ? ((lo >> sh) | (hi << (32 - sh))) : lo;
#else
	return __funnelshift_r(lo, hi, sh);
#endif
}
// planes of 32 symbols a[sh..] ++ b[..sh)
__device__ __forceinline__ Planes shift_in(Planes a, Planes b, uint32_t sh) { Planes r; r.lo = funnel_r(a.lo, b.lo, sh); r.hi = funnel_r(a.hi, b.hi, sh); return r; }
// 32 symbols of a planar array starting at symbol index `sym` (the array is padded by two words)
__device__ __forceinline__ Planes window32(const uint64_t* __restrict__ p, uint64_t sym)
{
	const uint64_t q = sym >> 5;
	const uint32_t sh = (uint32_t)(sym & 31);
	const Planes a = planes_of(__ldg(p + q));
	if (sh == 0) return a;
	return shift_in(a, planes_of(__ldg(p + q + 1)), sh);
}
__device__ __forceinline__ uint32_t symbol_at(const uint64_t* __restrict__ p, uint64_t sym)
{
	const uint64_t w = __ldg(p + (sym >> 5));
	const uint32_t b = (uint32_t)(sym & 31);
	return ((uint32_t)(w >> b) & 1) | (((uint32_t)(w >> (32 + b)) & 1) << 1);
}

// ---------------------------------------------------------------------------------------------- block scans

template <int NT>
__device__ __forceinline__ uint32_t block_exclusive_sum(uint32_t v, uint32_t* warp_tmp, uint32_t& total)
{
	const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	uint32_t inc = v;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(FULL, inc, d); if (lane >= d) inc += t; }
	if (lane == 31) warp_tmp[wid] = inc;
	__syncthreads();
	uint32_t base = 0, tot = 0;
#pragma unroll
	for (int i = 0; i < NT / 32; i++) { uint32_t t = warp_tmp[i]; if (i < (int)wid) base += t; tot += t; }
	__syncthreads();
	total = tot;
	return base + inc - v;
}

template <int NT>
__device__ __forceinline__ uint32_t block_exclusive_max(uint32_t v, uint32_t* warp_tmp, uint32_t& total)
{
	const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	uint32_t inc = v;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(FULL, inc, d); if (lane >= d) inc = max(inc, t); }
	uint32_t prev = __shfl_up_sync(FULL, inc, 1);
	if (lane == 0) prev = 0;
	if (lane == 31) warp_tmp[wid] = inc;
	__syncthreads();
	uint32_t base = 0, tot = 0;
#pragma unroll
	for (int i = 0; i < NT / 32; i++) { uint32_t t = warp_tmp[i]; if (i < (int)wid) base = max(base, t); tot = max(tot, t); }
	__syncthreads();
	total = tot;
	return max(base, prev);
}

// exclusive "last valid value" scan: result = the value of the closest preceding thread with valid != 0
struct LastRun { uint32_t start, prev_end, m; }; // m == 0 -> empty
template <int NT>
__device__ __forceinline__ LastRun block_exclusive_last(LastRun v, LastRun* warp_tmp, LastRun& total)
{
	const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	LastRun inc = v;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1)
	{
		LastRun t;
		t.start = __shfl_up_sync(FULL, inc.start, d);
		t.prev_end = __shfl_up_sync(FULL, inc.prev_end, d);
		t.m = __shfl_up_sync(FULL, inc.m, d);
		if (lane >= d && inc.m == 0) inc = t;
	}
	LastRun prev;
	prev.start = __shfl_up_sync(FULL, inc.start, 1);
	prev.prev_end = __shfl_up_sync(FULL, inc.prev_end, 1);
	prev.m = __shfl_up_sync(FULL, inc.m, 1);
	if (lane == 0) prev.m = 0;
	if (lane == 31) warp_tmp[wid] = inc;
	__syncthreads();
	LastRun base; base.m = 0; base.start = 0; base.prev_end = 0;
	LastRun tot = base;
#pragma unroll
	for (int i = 0; i < NT / 32; i++) { LastRun t = warp_tmp[i]; if (t.m) { if (i < (int)wid) base = t; tot = t; } }
	__syncthreads();
	total = tot;
	return prev.m ? prev : base;
}

struct Xor2 { uint64_t a, b; };
template <int NT>
__device__ __forceinline__ Xor2 block_exclusive_xor(Xor2 v, Xor2* warp_tmp, Xor2& total)
{
	const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	Xor2 inc = v;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1)
	{
		uint64_t ta = __shfl_up_sync(FULL, inc.a, d), tb = __shfl_up_sync(FULL, inc.b, d);
		if (lane >= d) { inc.a ^= ta; inc.b ^= tb; }
	}
	if (lane == 31) warp_tmp[wid] = inc;
	__syncthreads();
	Xor2 base = { 0, 0 }, tot = { 0, 0 };
#pragma unroll
	for (int i = 0; i < NT / 32; i++) { Xor2 t = warp_tmp[i]; if (i < (int)wid) { base.a ^= t.a; base.b ^= t.b; } tot.a ^= t.a; tot.b ^= t.b; }
	__syncthreads();
	total = tot;
	Xor2 r = { base.a ^ inc.a ^ v.a, base.b ^ inc.b ^ v.b };
	return r;
}

// ------------------------------------------------------------------------------------------------ tables

// Alphabet of MBG's microsatellite coding (MBG/ErrorMaskHelper.cpp:9-25): codes below CODES_BEFORE[m+1] have motif
// length <= m.
__constant__ uint32_t CODES_BEFORE[8] = { 0, 0, 4, 36, 228, 1252, 6372, 30948 };
constexpr uint32_t MAX_CODE = 30948;

// per-code mix of the rolling hasher (MBG/FastHasher.cpp:33-43)
__host__ __device__ __forceinline__ uint64_t char_hash(uint64_t key)
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

// reverse complement of one code (MBG/ErrorMaskHelper.cpp:43-73): rotate the motif left by `overhang` symbols,
// complement every symbol, reverse the symbol order; motif length and overhang are kept
__device__ __forceinline__ uint32_t complement_code(uint32_t code)
{
	uint32_t m = 1;
	while (m < 6 && CODES_BEFORE[m + 1] <= code) m++;
	uint32_t v = code - CODES_BEFORE[m];
	uint32_t motif = v / m, overhang = v - motif * m;
	uint32_t out = 0;
	for (uint32_t i = 0; i < m; i++)
	{
		uint32_t src_digit = (overhang + (m - 1 - i)) % m;                 // digit index, 0 = most significant
		uint32_t d = (motif >> (2 * (m - 1 - src_digit))) & 3;
		out = (out << 2) | (3 - d);
	}
	return CODES_BEFORE[m] + out * m + overhang;
}

// tab_hash[c] = char_hash(c), tab_comp[c] = complement(c), tab_pair[c] = { char_hash(c), char_hash(complement(c)) }
__global__ void k_build_tables(uint64_t* __restrict__ tab_hash, uint4* __restrict__ tab_pair, uint16_t* __restrict__ tab_comp)
{
	uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= MAX_CODE) return;
	const uint32_t rc = complement_code(c);
	const uint64_t hf = char_hash(c), hb = char_hash(rc);
	tab_hash[c] = hf;
	tab_pair[c] = make_uint4((uint32_t)hf, (uint32_t)(hf >> 32), (uint32_t)hb, (uint32_t)(hb >> 32));
	tab_comp[c] = (uint16_t)rc;
}

// ---------------------------------------------------------------------------------------------- K0: pack

// ASCII -> planar 2-bit.  One thread converts 32 characters (two 128-bit loads) into one 64-bit planar word; characters
// outside ACGTacgt are packed as A and their batch positions appended to `nonacgt` (rare; the host splits segments
// there).  Per 4 characters: byte-SIMD test, (x>>1 ^ x>>2) & 3 per byte, and one multiplication per plane that gathers
// the four bits (bit 8i -> bit 28+i).   Algorithmic bytes: 1 read + 1/4 written per base.
__global__ void __launch_bounds__(256) k_pack_ascii(const uint4* __restrict__ ascii, uint64_t n_bases, uint64_t first_chunk, uint64_t end_chunk,
	uint64_t* __restrict__ packed, uint64_t* __restrict__ nonacgt, unsigned long long* __restrict__ nonacgt_count, uint64_t nonacgt_cap)
{
	// chunks of 32 characters [first_chunk, end_chunk): the pipelined upload packs each piece as it arrives
	for (uint64_t i = first_chunk + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < end_chunk; i += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint4 v0 = __ldcs(ascii + 2 * i), v1 = __ldcs(ascii + 2 * i + 1);
		const uint32_t in[8] = { v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w };
		uint32_t lo = 0, hi = 0;
#pragma unroll
		for (int j = 0; j < 8; j++)
		{
			uint32_t x = in[j];
			const uint32_t u = x & 0xdfdfdfdfu;
			const uint32_t ok = __vcmpeq4(u, 0x41414141u) | __vcmpeq4(u, 0x43434343u) | __vcmpeq4(u, 0x47474747u) | __vcmpeq4(u, 0x54545454u);
			if (ok != 0xffffffffu)
			{
				for (int b = 0; b < 4; b++)
				{
					const uint64_t pos = i * 32 + j * 4 + b;
					if (((ok >> (8 * b)) & 0xff) == 0 && pos < n_bases)
					{
						const unsigned long long slot = atomicAdd(nonacgt_count, 1ull);
						if (slot < nonacgt_cap) nonacgt[slot] = pos;
					}
				}
				x &= ok; // non-ACGT -> 0 -> packs as A
			}
			const uint32_t t = ((x >> 1) ^ (x >> 2)) & 0x03030303u; // A0 C1 G2 T3
			lo |= (((t & 0x01010101u) * 0x10204080u) >> 28) << (4 * j);
			hi |= ((((t >> 1) & 0x01010101u) * 0x10204080u) >> 28) << (4 * j);
		}
		packed[i] = word_of(lo, hi);
	}
}

// ------------------------------------------------------------------------------------------ K1: collapse

// bit-compress of both planes: gathers the bits selected by `m` into the low bits, preserving order.
// Parallel-suffix method (Hacker's Delight 7-4); the move masks are computed once and applied to both planes.
__device__ __forceinline__ void compress_planes(uint32_t& lo, uint32_t& hi, uint32_t m)
{
	lo &= m; hi &= m;
	uint32_t mk = ~m << 1;
#pragma unroll
	for (int i = 0; i < 5; i++)
	{
		uint32_t mp = mk ^ (mk << 1);
		mp ^= mp << 2; mp ^= mp << 4; mp ^= mp << 8; mp ^= mp << 16;
		const uint32_t mv = mp & m;
		m = (m ^ mv) | (mv >> (1 << i));
		uint32_t t = lo & mv;
		lo = (lo ^ t) | (t >> (1 << i));
		t = hi & mv;
		hi = (hi ^ t) | (t >> (1 << i));
		mk &= ~mp;
	}
}

struct SegTable
{
	const uint64_t* in_start;  // first base of the segment in `packed`
	const uint32_t* len;       // bases
	const uint64_t* out_base;  // first symbol / code slot owned by the segment (n + 1 entries)
	uint32_t n;
};

// K1: homopolymer collapse (MBG/ReadHelper.h:864-876 collapse, :899-1018 run-length symbols).  One WARP per segment,
// each lane owns 32 bases of a 1024-base tile: neighbour compare on both planes -> keep mask -> popcount -> warp prefix
// sum -> bit-compress -> merged into a shared-memory staging tile -> coalesced 64-bit stores.  No CTA barrier anywhere:
// the base before a lane's 32 comes from the lane below (shuffle; lane 0: the last base of the previous tile, kept in a
// register), the partial output word between tiles lives in a register, the next tile's bases are requested one tile
// ahead.  (The earlier CTA-per-segment form with a block scan and three barriers per tile took 0.73 ms on the bench, this
// one 0.63 ms; ncu: ALU pipe, the 5-round compress is ~110 of the ~290 instructions per 32 bases.)
// keep_all != 0 (--error-masking=no): every base is a symbol (the kernel only re-aligns the segment).
// pos1 != nullptr: also writes the segment-relative raw position of every kept base (the run starts, i.e. poses).
// Algorithmic bytes: 1/4 read per base + 1/4 written per symbol (+ 4 per symbol when pos1 is requested).
#ifdef MBGFE_EMU_BUILD
constexpr int COLLAPSE_WARPS = 1;      // the emulator's warp primitives are CTA-wide barriers
#else
constexpr int COLLAPSE_WARPS = 4;
#endif

__global__ void __launch_bounds__(COLLAPSE_WARPS * 32) k_collapse(SegTable seg, const uint64_t* __restrict__ packed, uint64_t* __restrict__ sym,
	uint32_t* __restrict__ seg_c1, uint32_t* __restrict__ pos1, int keep_all)
{
	__shared__ uint32_t s_lo[COLLAPSE_WARPS][34], s_hi[COLLAPSE_WARPS][34];
	const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	const uint32_t g = blockIdx.x * COLLAPSE_WARPS + wid;
	if (g >= seg.n) return;
	const uint64_t in0 = seg.in_start[g];
	const uint32_t len = seg.len[g];
	const uint64_t out0 = seg.out_base[g];
	uint64_t* __restrict__ out = sym + (out0 >> 5);
	uint32_t* st_lo = s_lo[wid];
	uint32_t* st_hi = s_hi[wid];
	uint32_t carry = 0;                     // symbols written so far
	uint32_t part_lo = 0, part_hi = 0;      // the unfinished output word
	uint32_t prev_lo = 0, prev_hi = 0;      // last base of the previous tile
	Planes next;                            // the lane's 32 bases of the next tile, requested one tile ahead
	next.lo = 0; next.hi = 0;
	if (lane * 32 < len) next = window32(packed, in0 + lane * 32);
	for (uint32_t tile = 0; tile < len; tile += 1024)
	{
		const uint32_t base0 = tile + lane * 32;
		uint32_t lo = next.lo, hi = next.hi;
		const uint32_t nvalid = base0 < len ? min(32u, len - base0) : 0u;
		next.lo = 0; next.hi = 0;
		if (base0 + 1024 < len) next = window32(packed, in0 + base0 + 1024);
		uint32_t below_lo = __shfl_up_sync(FULL, lo >> 31, 1), below_hi = __shfl_up_sync(FULL, hi >> 31, 1);
		if (lane == 0) { below_lo = prev_lo; below_hi = prev_hi; }
		uint32_t keep = (lo ^ ((lo << 1) | below_lo)) | (hi ^ ((hi << 1) | below_hi));
		if (base0 == 0) keep |= 1u;
		if (keep_all) keep = 0xffffffffu;
		keep &= low_bits(nvalid);
		prev_lo = __shfl_sync(FULL, lo >> 31, 31); prev_hi = __shfl_sync(FULL, hi >> 31, 31);
		const uint32_t cnt = __popc(keep);
		uint32_t inc = cnt;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, d); if (lane >= (uint32_t)d) inc += t; }
		const uint32_t tile_total = __shfl_sync(FULL, inc, 31);
		const uint32_t excl = inc - cnt;
		// the staging tile holds symbols [carry & ~31, ...): word 0 is the partial word of the previous tile
		st_lo[1 + lane] = 0; st_hi[1 + lane] = 0;
		if (lane == 0) { st_lo[0] = part_lo; st_hi[0] = part_hi; st_lo[33] = 0; st_hi[33] = 0; }
		__syncwarp();
		if (cnt)
		{
			if (pos1)
			{
				uint32_t idx = carry + excl;
				uint32_t kk = keep;
				while (kk) { const uint32_t b = __ffs(kk) - 1; kk &= kk - 1; pos1[out0 + idx++] = base0 + b; }
			}
			compress_planes(lo, hi, keep);
			const uint32_t o = (carry & 31) + excl;
			const uint32_t sh = o & 31;
			atomicOr(&st_lo[o >> 5], lo << sh);
			atomicOr(&st_hi[o >> 5], hi << sh);
			if (sh && cnt + sh > 32) { atomicOr(&st_lo[(o >> 5) + 1], lo >> (32 - sh)); atomicOr(&st_hi[(o >> 5) + 1], hi >> (32 - sh)); }
		}
		__syncwarp();
		const uint32_t filled = (carry & 31) + tile_total;       // symbols now in the staging tile
		const bool last = tile + 1024 >= len;
		const uint32_t n_words = last ? (filled + 31) / 32 : filled / 32;
		const uint32_t w0 = carry >> 5;
		if (lane < n_words) out[w0 + lane] = word_of(st_lo[lane], st_hi[lane]);             // at most 33 words
		if (lane == 0 && n_words == 33) out[w0 + 32] = word_of(st_lo[32], st_hi[32]);
		part_lo = (filled & 31) ? st_lo[filled >> 5] : 0u;
		part_hi = (filled & 31) ? st_hi[filled >> 5] : 0u;
		carry += tile_total;
		__syncwarp();      // the stage is rewritten at the top of the next tile
	}
	if (lane == 0)
	{
		seg_c1[g] = carry;
		if (pos1) pos1[out0 + carry] = len;
	}
}

// K1, byte-table form (the one that runs unless per-base positions are wanted; k_collapse above serves those runs).
// ncu on k_collapse (r01): ALU pipe, ~290 instructions per 32 bases, of which the 5-round bit-compress of the two planes is ~110 and
// the per-tile overheads (unaligned window, warp scan, staging, barriers) most of the rest -- 20 % of the HBM copy peak for a kernel
// that moves 0.44 bytes per base.  Two changes: (1) compaction by table: PEXT8[mask][value] = the bits of `value` selected by `mask`,
// gathered at the low end -- 64 KB, copied into shared memory once per CTA (CTAs are persistent: warps stride over the segments), two
// byte lookups per 8 bases and plane; (2) a lane takes FOUR words (128 bases) per tile of 4 096, so the warp scan, the staging and
// the synchronisation are paid once per 128 bases.  ~90 instructions per 32 bases.
// dynamic shared memory = 65 536 + warps * 2 * COLLAPSE_STAGE * 4 bytes.
#ifdef MBGFE_EMU_BUILD
constexpr int COLLAPSE_LUT_WARPS = 1;      // the emulator's warp primitives are CTA-wide barriers
#else
constexpr int COLLAPSE_LUT_WARPS = 8;
#endif
constexpr uint32_t COLLAPSE_STAGE = 132;   // staging words per plane and warp: a tile's 128 output words + the carried word + spill

__global__ void __launch_bounds__(256) k_build_pext8(uint8_t* __restrict__ lut)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= 65536) return;
	const uint32_t m = i >> 8, v = i & 0xff;
	uint32_t out = 0, n = 0;
	for (uint32_t b = 0; b < 8; b++) if ((m >> b) & 1) { out |= ((v >> b) & 1) << n; n++; }
	lut[i] = (uint8_t)out;
}

// both planes of one word through the table: the kept bits of lo / hi gathered at the low end (popc(keep) of them)
__device__ __forceinline__ void compress_planes_lut(const uint8_t* __restrict__ lut, uint32_t& lo, uint32_t& hi, uint32_t keep)
{
	uint32_t out_lo = 0, out_hi = 0, off = 0;
#pragma unroll
	for (int b = 0; b < 4; b++)
	{
		const uint32_t m = (keep >> (8 * b)) & 0xff;
		const uint8_t* __restrict__ row = lut + (m << 8);
		out_lo |= (uint32_t)row[(lo >> (8 * b)) & 0xff] << off;
		out_hi |= (uint32_t)row[(hi >> (8 * b)) & 0xff] << off;
		off += __popc(m);
	}
	lo = out_lo; hi = out_hi;
}

__global__ void __launch_bounds__(COLLAPSE_LUT_WARPS * 32) k_collapse_lut(SegTable seg, const uint64_t* __restrict__ packed, uint64_t* __restrict__ sym,
	uint32_t* __restrict__ seg_c1, const uint8_t* __restrict__ pext8, int keep_all)
{
	MBGFE_DYN_SMEM(smem_raw);
	uint8_t* lut = smem_raw;
	const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	uint32_t* st_lo = (uint32_t*)(smem_raw + 65536) + (size_t)wid * 2 * COLLAPSE_STAGE;
	uint32_t* st_hi = st_lo + COLLAPSE_STAGE;
	for (uint32_t i = threadIdx.x; i < 4096; i += blockDim.x) ((uint4*)lut)[i] = __ldg((const uint4*)pext8 + i);
	__syncthreads();
	for (uint32_t g = blockIdx.x * COLLAPSE_LUT_WARPS + wid; g < seg.n; g += gridDim.x * COLLAPSE_LUT_WARPS)
	{
		const uint64_t in0 = seg.in_start[g];
		const uint32_t len = seg.len[g];
		uint64_t* __restrict__ out = sym + (seg.out_base[g] >> 5);
		const uint64_t* __restrict__ src = packed + (in0 >> 5);
		const uint32_t sh = (uint32_t)(in0 & 31);
		uint32_t carry = 0, part_lo = 0, part_hi = 0, prev_lo = 0, prev_hi = 0;
		// the lane's 128 bases of a tile: five source words, realigned to four
		uint64_t raw[5];
		auto fetch = [&](uint32_t base0)
		{
#pragma unroll
			for (int j = 0; j < 5; j++) raw[j] = 0;
			if (base0 < len)
			{
				const uint64_t* q = src + (base0 >> 5);
#pragma unroll
				for (int j = 0; j < 5; j++) raw[j] = __ldg(q + j);      // the packed array is padded past the batch
			}
		};
		fetch(lane * 128);
		for (uint32_t tile = 0; tile < len; tile += 4096)
		{
			const uint32_t base0 = tile + lane * 128;
			Planes wd[4];
#pragma unroll
			for (int j = 0; j < 4; j++) wd[j] = sh ? shift_in(planes_of(raw[j]), planes_of(raw[j + 1]), sh) : planes_of(raw[j]);
			fetch(base0 + 4096);
			const uint32_t nvalid = base0 < len ? min(128u, len - base0) : 0u;
			uint32_t below_lo = __shfl_up_sync(FULL, wd[3].lo >> 31, 1), below_hi = __shfl_up_sync(FULL, wd[3].hi >> 31, 1);
			if (lane == 0) { below_lo = prev_lo; below_hi = prev_hi; }
			prev_lo = __shfl_sync(FULL, wd[3].lo >> 31, 31); prev_hi = __shfl_sync(FULL, wd[3].hi >> 31, 31);
			uint32_t keep[4], cnt[4], lane_total = 0;
#pragma unroll
			for (int j = 0; j < 4; j++)
			{
				uint32_t kp = (wd[j].lo ^ ((wd[j].lo << 1) | below_lo)) | (wd[j].hi ^ ((wd[j].hi << 1) | below_hi));
				if (base0 + 32 * j == 0) kp |= 1u;
				if (keep_all) kp = 0xffffffffu;
				kp &= low_bits(nvalid > 32u * j ? min(32u, nvalid - 32u * j) : 0u);
				keep[j] = kp; cnt[j] = __popc(kp); lane_total += cnt[j];
				below_lo = wd[j].lo >> 31; below_hi = wd[j].hi >> 31;
			}
			uint32_t inc = lane_total;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, d); if (lane >= (uint32_t)d) inc += t; }
			const uint32_t tile_total = __shfl_sync(FULL, inc, 31);
			// the staging strip holds symbols [carry & ~31, ...): word 0 is the unfinished word of the previous tile
			for (uint32_t i = 1 + lane; i < COLLAPSE_STAGE; i += 32) { st_lo[i] = 0; st_hi[i] = 0; }
			if (lane == 0) { st_lo[0] = part_lo; st_hi[0] = part_hi; }
			__syncwarp();
			uint32_t o = (carry & 31) + inc - lane_total;
#pragma unroll
			for (int j = 0; j < 4; j++)
			{
				if (cnt[j])
				{
					uint32_t lo = wd[j].lo, hi = wd[j].hi;
					compress_planes_lut(lut, lo, hi, keep[j]);
					const uint32_t bs = o & 31;
					atomicOr(&st_lo[o >> 5], lo << bs);
					atomicOr(&st_hi[o >> 5], hi << bs);
					if (bs && cnt[j] + bs > 32) { atomicOr(&st_lo[(o >> 5) + 1], lo >> (32 - bs)); atomicOr(&st_hi[(o >> 5) + 1], hi >> (32 - bs)); }
					o += cnt[j];
				}
			}
			__syncwarp();
			const uint32_t filled = (carry & 31) + tile_total;
			const bool last = tile + 4096 >= len;
			const uint32_t n_words = last ? (filled + 31) / 32 : filled / 32;
			const uint32_t w0 = carry >> 5;
			for (uint32_t i = lane; i < n_words; i += 32) out[w0 + i] = word_of(st_lo[i], st_hi[i]);
			part_lo = (filled & 31) ? st_lo[filled >> 5] : 0u;
			part_hi = (filled & 31) ? st_hi[filled >> 5] : 0u;
			carry += tile_total;
			__syncwarp();
		}
		if (lane == 0) seg_c1[g] = carry;
	}
}

// ------------------------------------------------------------------------------- K2: microsatellite coding

// 1 bit per position p of word a where sym[p] == sym[p+m]; b = the 32 symbols that follow a
__device__ __forceinline__ uint32_t eq_shift(Planes a, Planes b, uint32_t m)
{
	return ~((a.lo ^ funnel_r(a.lo, b.lo, m)) | (a.hi ^ funnel_r(a.hi, b.hi, m)));
}

// bits 0..5 of x moved to the even positions 0, 2, .., 10
__device__ __forceinline__ uint32_t spread6(uint32_t x)
{
	x = (x | (x << 4)) & 0x0f0fu;
	x = (x | (x << 2)) & 0x3333u;
	x = (x | (x << 1)) & 0x5555u;
	return x;
}

// floor(v / m) for m in 2..6 and v < 2^29 as one multiply-high by ceil(2^32 / m)
__constant__ uint32_t RECIP32[8] = { 0, 0, 0x80000000u, 0x55555556u, 0x40000000u, 0x33333334u, 0x2aaaaaabu, 0 };

// code of a run-length stretch (MBG/ErrorMaskHelper.cpp:107-127) from the first m symbols of the stretch (planes, first symbol
// at bit 0) and its length: code = numBefore[m] + motif * m + (len mod m), motif = the m symbols as base-4 digits, first symbol
// most significant.
__device__ __forceinline__ uint32_t stretch_code(Planes first, uint32_t m, uint32_t len)
{
	const uint32_t le = (spread6(first.lo & low_bits(m)) | (spread6(first.hi & low_bits(m)) << 1));   // digit j = symbol j
#ifdef MBGFE_EMU_BUILD
	uint32_t be = 0;
	for (uint32_t j = 0; j < m; j++) be = (be << 2) | ((le >> (2 * j)) & 3);
	const uint32_t oh = m > 1 ? len % m : 0;
#else
	const uint32_t y = __brev(le) >> (32 - 2 * m);                 // digit order reversed, bits inside each digit swapped
	const uint32_t be = ((y & 0xaaau) >> 1) | ((y & 0x555u) << 1);
	const uint32_t oh = m > 1 ? len - __umulhi(len, RECIP32[m]) * m : 0;
#endif
	return CODES_BEFORE[m] + be * m + oh;
}

// K2: multiRLECompress (MBG/ErrorMaskHelper.cpp:169-342) in scan-parallel form (DESIGN.md "K2"; the sequential
// statement of exactly this algorithm is oracle/mbg_oracle.c:orc_msat_scanform, fuzzed against the reference).
//   candidates : starts a of maximal stretches of eq_m (sym[p]==sym[p+m]) of length L >= m; run = [a, a+L+m)
//   emitted    : candidate end > exclusive prefix maximum of candidate ends           (max-scan + carry)
//   cover      : 1 = inside exactly one emitted run, else single-symbol codes        (from the two carried ends)
//   code start : p == 0 | cover(p) != 1 | cover(p-1) != 1 | a run is emitted at p     (popcount + sum-scan)
// One WARP per segment, one 32-symbol planar word per lane and tile (1024 symbols): every mask is one bit per
// position, the three scans are warp shuffles (no CTA barrier anywhere), carries between tiles live in registers.
// A lane walks its candidates once in position order (ends by count-trailing-zeros on the 64 positions it sees; longer
// stretches walk on through the staged words), then once more to mark the emitted ones and build the cover masks.
// Codes are not stored to HBM one by one: the tile's codes and reverse-complement codes are staged in shared memory
// (the staging index is kept congruent to the output index mod 8) and leave as coalesced 128-bit stores; an unfinished
// group of 8 is carried into the next tile.  A stretch is coded by the lane that owns the symbol after it; when its
// slot was flushed by an earlier tile (stretches longer than a tile) the code goes to HBM directly.
// Segments with fewer than 32 symbols produce no part (MBG/ReadHelper.h:847,858,882,893): seg_c2 = 0.
// Algorithmic bytes: 1/4 read per symbol + 2 written per code (+ 4 per code when poses are requested).
#ifdef MBGFE_EMU_BUILD
constexpr int MSAT_WARPS = 1;      // the emulator's warp primitives are CTA-wide barriers: one warp per CTA there
#else
constexpr int MSAT_WARPS = 4;
#endif
constexpr uint32_t MSAT_TILE = 1024, MSAT_STAGE = MSAT_TILE + 8;

__global__ void __launch_bounds__(MSAT_WARPS * 32) k_msat(SegTable seg, const uint64_t* __restrict__ sym, const uint32_t* __restrict__ seg_c1,
	uint16_t* __restrict__ codes, uint32_t* __restrict__ seg_c2, const uint32_t* __restrict__ pos1,
	uint32_t* __restrict__ pos2, uint32_t max_motif)
{
	__shared__ uint64_t s_words[MSAT_WARPS][36];                       // planar words tile_word0 - 1 .. tile_word0 + 34
	__shared__ uint32_t s_cand[MSAT_WARPS][32 * 33];                   // per lane: (run end << 3) | (7 - motif) of the candidate at each of its 32 positions, then of its emitted runs
	__align__(16) __shared__ uint16_t s_codes[MSAT_WARPS][MSAT_STAGE];
	const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	const uint32_t g = blockIdx.x * MSAT_WARPS + wid;
	if (g >= seg.n) return;
	const uint32_t n = seg_c1[g];
	if (n < 32) { if (lane == 0) seg_c2[g] = 0; return; }
	const uint64_t out0 = seg.out_base[g];
	const uint64_t* __restrict__ s = sym + (out0 >> 5);   // segment-local words
	uint16_t* __restrict__ codes_g = codes + out0;
	const uint32_t n_words = (n + 31) / 32;
	uint64_t* tile_w = s_words[wid];
	uint32_t* my_cand = s_cand[wid] + lane * 33;
	uint16_t* st_c = s_codes[wid];
	const uint32_t lt_mask = low_bits(lane);               // lanes below this one
	uint32_t carry_max = 0;            // maximum run end over all candidates of earlier tiles
	uint32_t carry_last_sm = 0, carry_last_pe = 0;   // latest emitted run so far: start | m << 29 (m == 0: none), end of the run emitted before it
	uint32_t carry_codes = 0;
	for (uint32_t tile = 0; tile < n; tile += MSAT_TILE)
	{
		const uint32_t jb = tile >> 5;     // first word of the tile
		for (uint32_t i = lane; i < 36; i += 32)
		{
			const uint32_t jj = jb + i;    // word jj - 1
			tile_w[i] = (jj >= 1 && jj - 1 < n_words) ? __ldg(s + (jj - 1)) : 0ull;
		}
		__syncwarp();
		// 32 symbols starting at segment position q, from the staged words when they cover it
		auto win = [&](uint32_t q) -> Planes
		{
			const uint32_t jq = q >> 5, sh = q & 31;
			if (jq + 1 >= jb && jq + 1 - jb + 1 < 36)
			{
				const Planes a = planes_of(tile_w[jq + 1 - jb]);
				return sh ? shift_in(a, planes_of(tile_w[jq + 2 - jb]), sh) : a;
			}
			Planes a = planes_of(jq < n_words ? __ldg(s + jq) : 0ull);
			if (sh) a = shift_in(a, planes_of(jq + 1 < n_words ? __ldg(s + jq + 1) : 0ull), sh);
			return a;
		};
		const uint32_t p0 = tile + lane * 32;
		const bool active = p0 < n;
		const Planes am = planes_of(tile_w[lane]), a0 = planes_of(tile_w[lane + 1]), a1 = planes_of(tile_w[lane + 2]), a2 = planes_of(tile_w[lane + 3]);
		// ---- candidates: starts of stretches of eq_m of length >= m (one mask per motif length)
		uint32_t e0[7], e1[7], cm[7];
		uint32_t cand_any = 0;
#pragma unroll
		for (uint32_t m = 2; m <= 6; m++)
		{
			e0[m] = 0; e1[m] = 0; cm[m] = 0;
			if (m <= max_motif && n >= 2 * m && active)
			{
				const uint32_t lim = n - m;            // eq_m(p) is defined for p < lim
				e0[m] = eq_shift(a0, a1, m) & low_bits_sat(lim, p0);
				e1[m] = eq_shift(a1, a2, m) & low_bits_sat(lim, p0 + 32);
				uint32_t prev = 0;
				if (p0 > 0 && p0 - 1 < lim) prev = eq_shift(am, a0, m) >> 31;
				uint32_t run = e0[m] & ~((e0[m] << 1) | prev);
#pragma unroll
				for (uint32_t t = 1; t < m; t++) run &= funnel_r(e0[m], e1[m], t);
				cm[m] = run;
				cand_any |= run;
			}
		}
		// ---- pass 1: per candidate position the largest end (then smallest motif) among the runs starting there, kept in my_cand[position].
		// One loop per motif length over the bits of its own candidate mask: a lane's e0 / e1 / cm of that length are registers, and the
		// body runs once per round (the loop over positions with five predicated motif branches inside it ran every branch for a handful
		// of lanes: 17 % of the kernel's instructions at 6 of 32 lanes)
		uint32_t local_max = 0;
		{
			uint32_t seen = 0;
#pragma unroll
			for (uint32_t m = 2; m <= 6; m++)
			{
				uint32_t cbm = cm[m];
				while (cbm)
				{
					const uint32_t b = __ffs(cbm) - 1;
					cbm &= cbm - 1;
					// eq_m holds from b to the first zero above it; beyond the 64 positions this lane sees, walk on
					const uint32_t stop_lo = ~funnel_r(e0[m], e1[m], b), stop_hi = ~(e1[m] >> b);
					uint32_t q = p0 + 64;
					if (stop_lo) q = p0 + b + (__ffs(stop_lo) - 1);
					else if (stop_hi) q = p0 + b + 32 + (__ffs(stop_hi) - 1);      // (the shifted-in zeros stop it at p0 + 64 at the latest)
					if (q >= p0 + 64)
					{
						const uint32_t lim = n - m;
						q = p0 + 64;
						while (q < lim)
						{
							const Planes x = win(q), y = win(q + m);
							uint32_t ne = (x.lo ^ y.lo) | (x.hi ^ y.hi);
							const uint32_t room = lim - q;
							if (room < 32) ne |= ~low_bits(room);
							if (ne) { q += __ffs(ne) - 1; break; }
							q += 32;
						}
						if (q > lim) q = lim;
					}
					uint32_t val = ((q + m) << 3) | (7 - m);   // larger end wins, then smaller motif
					if ((seen >> b) & 1) val = max(val, my_cand[b]);
					my_cand[b] = val;
					local_max = max(local_max, val >> 3);
				}
				seen |= cm[m];
			}
		}
		// ---- scan 1: exclusive prefix maximum of candidate ends
		uint32_t pm_in, tile_max;
		{
			uint32_t inc = local_max;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, d); if (lane >= (uint32_t)d) inc = max(inc, t); }
			uint32_t excl = __shfl_up_sync(FULL, inc, 1);
			if (lane == 0) excl = 0;
			tile_max = __shfl_sync(FULL, inc, 31);
			pm_in = max(excl, carry_max);
		}
		// ---- pass 2: emitted runs of this word (end beyond everything seen before), kept as a compacted list at the front of
		// my_cand, and their share of the cover masks:
		//   ge1(p): some run covers p      <=> (maximum run end over candidates at or before p) > p
		//   ge2(p): two runs cover p       <=> (prev_end of the latest emitted run at or before p) > p
		uint32_t emit_bits = 0, n_emit = 0, ge1 = 0, ge2_own = 0;
		uint32_t my_sm = 0, my_pe = 0;     // latest emitted run of this word: start | m << 29, end of the run emitted before it
		{
			uint32_t running = pm_in, cb = cand_any;
			while (cb)
			{
				const uint32_t b = __ffs(cb) - 1;
				cb &= cb - 1;
				const uint32_t val = my_cand[b];      // (the emitted ones are packed to the front below: never past the entry just read)
				const uint32_t end = val >> 3;
				if (end > running)
				{
					const uint32_t from = 0xffffffffu << b;
					emit_bits |= 1u << b;
					ge1 |= from & low_bits(end - p0);
					ge2_own = (ge2_own & ~from) | (from & low_bits_sat(running, p0));
					my_sm = (p0 + b) | ((7 - (val & 7)) << 29);
					my_pe = running;
					my_cand[n_emit++] = val;
					running = end;
				}
			}
		}
		// ---- scan 2: latest emitted run that starts before this word (the nearest lower lane that has one, else the carry)
		uint32_t in_sm, in_pe;
		const uint32_t has = __ballot_sync(FULL, my_sm != 0);
		{
			const uint32_t below = has & lt_mask;
			const uint32_t src = below ? 31 - __clz(below) : 0;
			in_sm = __shfl_sync(FULL, my_sm, src);
			in_pe = __shfl_sync(FULL, my_pe, src);
			if (!below) { in_sm = carry_last_sm; in_pe = carry_last_pe; }
		}
		// ---- coverage of the 32 symbols of this word
		uint32_t flags = 0, cover1 = 0;
		bool prev_cover1 = false;
		if (active)
		{
			const uint32_t nvalid = min(32u, n - p0);
			const uint32_t valid = low_bits(nvalid);
			const uint32_t pe_in = in_sm ? in_pe : 0;
			ge1 |= low_bits_sat(pm_in, p0);
			const uint32_t first_from = emit_bits ? (0xffffffffu << (__ffs(emit_bits) - 1)) : 0u;
			const uint32_t ge2 = (low_bits_sat(pe_in, p0) & ~first_from) | ge2_own;
			cover1 = ge1 & ~ge2 & valid;
			if (p0 > 0) prev_cover1 = (pm_in > p0 - 1) && !(pe_in > p0 - 1);
			const uint32_t not1 = ~cover1;
			const uint32_t prev_not1 = (not1 << 1) | (prev_cover1 ? 0u : 1u);
			flags = (not1 | prev_not1 | emit_bits) & valid;
		}
		// ---- scan 3: index of the first code that starts in this word
		uint32_t idx, tile_codes;
		{
			const uint32_t v = __popc(flags);
			uint32_t inc = v;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, d); if (lane >= (uint32_t)d) inc += t; }
			tile_codes = __shfl_sync(FULL, inc, 31);
			idx = carry_codes + inc - v;
		}
		const uint32_t base8 = carry_codes & ~7u;      // stage entry i holds code base8 + i
		if (active)
		{
			// ---- single-symbol codes: every code start outside cover == 1
			uint32_t sb = flags & ~cover1;
			if (!pos2)
			{
				// without positions: one predicated store per symbol, the output slot counted along -- no find-first-set / popcount per
				// code and no divergence between lanes (the loop below ran as long as the fullest word of the warp: 24 % of K2's
				// instructions at 18 of 32 lanes)
				uint32_t o = idx - base8;
#pragma unroll
				for (uint32_t b = 0; b < 32; b++)
				{
					if ((sb >> b) & 1) st_c[o] = (uint16_t)(((a0.lo >> b) & 1) | (((a0.hi >> b) & 1) << 1));
					o += (flags >> b) & 1;
				}
				sb = 0;
			}
			while (sb)
			{
				const uint32_t b = __ffs(sb) - 1; sb &= sb - 1;
				const uint32_t oi = idx + __popc(flags & low_bits(b));
				const uint32_t sc = ((a0.lo >> b) & 1) | (((a0.hi >> b) & 1) << 1);
				st_c[oi - base8] = (uint16_t)sc;
				if (pos2) pos2[out0 + oi] = pos1 ? pos1[out0 + p0 + b] : p0 + b;
			}
			// ---- stretches of cover == 1 that end in this word: at a code start whose predecessor is covered once, or at n
			uint32_t ends = flags & ((cover1 << 1) | (prev_cover1 ? 1u : 0u));
			const uint32_t nvalid = min(32u, n - p0);
			bool tail = p0 + 32 >= n && ((cover1 >> (nvalid - 1)) & 1);    // the segment ends inside a stretch
			while (ends || tail)
			{
				uint32_t b;
				if (ends) { b = __ffs(ends) - 1; ends &= ends - 1; } else { b = nvalid; tail = false; }
				const uint32_t q = p0 + b;
				const uint32_t oi = idx + __popc(flags & low_bits(b)) - 1;
				// the latest emitted run that starts before q, and the end of the run emitted before it
				uint32_t cur_start = in_sm & 0x1fffffffu, cur_m = in_sm >> 29, cur_pe = in_pe;
				const uint32_t below = emit_bits & low_bits(b);
				if (below)
				{
					const uint32_t ie = __popc(below) - 1;
					const uint32_t val = my_cand[ie];
					cur_start = p0 + (31 - __clz(below));
					cur_m = 7 - (val & 7);
					cur_pe = ie ? (my_cand[ie - 1] >> 3) : pm_in;
				}
				// symbols [ps, q) are covered by that run alone: one code with its motif length (clamped to the stretch)
				const uint32_t ps = max(cur_start, cur_pe);
				const uint32_t slen = q - ps;
				const uint32_t m = min(cur_m, slen);
				const uint32_t code = stretch_code(win(ps), m, slen);
				if (oi >= base8) st_c[oi - base8] = (uint16_t)code;
				else codes_g[oi] = (uint16_t)code;     // its slot left with an earlier tile
				if (pos2) pos2[out0 + oi] = pos1 ? pos1[out0 + ps] : ps;
			}
		}
		__syncwarp();
		// ---- the finished groups of 8 codes leave as 128-bit stores; the unfinished group moves to the front of the stage
		{
			const uint32_t filled = (carry_codes - base8) + tile_codes;
			const bool last = tile + MSAT_TILE >= n;
			const uint32_t n_groups = last ? (filled + 7) / 8 : filled / 8;
			for (uint32_t gq = lane; gq < n_groups; gq += 32)
			{
				*(uint4*)(codes_g + base8 + 8 * gq) = *(const uint4*)(st_c + 8 * gq);
			}
			const uint32_t keep = last ? 0 : filled - 8 * n_groups;
			uint16_t kc = 0;
			if (lane < keep) kc = st_c[8 * n_groups + lane];
			__syncwarp();
			if (lane < keep) st_c[lane] = kc;
		}
		carry_max = max(carry_max, tile_max);
		if (has)
		{
			const uint32_t top = 31 - __clz(has);
			carry_last_sm = __shfl_sync(FULL, my_sm, top);
			carry_last_pe = __shfl_sync(FULL, my_pe, top);
		}
		carry_codes += tile_codes;
		__syncwarp();    // the staged words are replaced at the top of the next tile
	}
	if (lane == 0)
	{
		seg_c2[g] = carry_codes;
		if (pos2) pos2[out0 + carry_codes] = pos1 ? pos1[out0 + n] : n;
	}
}

// modes without microsatellite coding: codes are the symbols themselves (2-bit -> uint16), poses are pos1
template <int NT>
__global__ void __launch_bounds__(NT) k_unpack_symbols(SegTable seg, const uint64_t* __restrict__ sym, const uint32_t* __restrict__ seg_c1,
	uint16_t* __restrict__ codes, uint32_t* __restrict__ seg_c2, const uint32_t* __restrict__ pos1, uint32_t* __restrict__ pos2)
{
	const uint32_t g = blockIdx.x;
	if (g >= seg.n) return;
	const uint32_t n = seg_c1[g];
	const uint64_t out0 = seg.out_base[g];
	const uint64_t* __restrict__ s = sym + (out0 >> 5);
	for (uint32_t p = threadIdx.x; p < n; p += NT)
	{
		const uint32_t sc = symbol_at(s, p);
		codes[out0 + p] = (uint16_t)sc;
		if (pos2) pos2[out0 + p] = pos1 ? pos1[out0 + p] : p;
	}
	if (threadIdx.x == 0)
	{
		seg_c2[g] = n;
		if (pos2) pos2[out0 + n] = pos1 ? pos1[out0 + n] : n;
	}
}

// ----------------------------------------------------------------- K3a: rolling s-mer hash + syncmer selection

// 64-bit rotations by a value known to be in 0..63
__device__ __forceinline__ uint64_t rotl_c(uint64_t v, uint32_t r) { return r ? ((v << r) | (v >> (64 - r))) : v; }
__device__ __forceinline__ uint64_t rotr_c(uint64_t v, uint32_t r) { return r ? ((v >> r) | (v << (64 - r))) : v; }
__device__ __forceinline__ uint64_t u64_of(uint32_t lo, uint32_t hi) { return ((uint64_t)hi << 32) | lo; }
// rotation by a run-time amount r in 0..63 as two funnel shifts (+ a word swap when r >= 32): 4 ALU instructions
// instead of the ~10 a generic 64-bit shift pair costs
__device__ __forceinline__ uint64_t rotl_v(uint64_t v, uint32_t r)
{
#ifdef MBGFE_EMU_BUILD
	return rotl64(v, r);
#else
	uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
	if (r & 32) { const uint32_t t = lo; lo = hi; hi = t; }
	return u64_of(__funnelshift_l(hi, lo, r), __funnelshift_l(lo, hi, r));
#endif
}
__device__ __forceinline__ uint64_t rotr_v(uint64_t v, uint32_t r) { return rotl_v(v, (64 - r) & 63); }

// K3a: FastHasher + findSyncmerPositions (MBG/FastHasher.h:16-32, MBG/ReadHelper.h:31-79) without the sequential
// dependence.  The reference rolls fw = rotl(fw,1)^H[c] with the outgoing character removed; with the "moving frame"
// prefixes  P'[n+1] = rotl(P'[n] ^ H[c_n], 1)  and  Q'[n+1] = rotr(Q'[n] ^ H[comp c_n], 1)  (P'[0] = Q'[0] = 0) the hash
// of the s-mer starting at i is
//     fw(i) = rotr(P'[i+s], 1) ^ rotl(P'[i], s-1),   bw(i) = rotl(Q'[i+s], s) ^ Q'[i],   hash = min(fw, bw)
// (SURVEY.md H3 rewritten so that no rotation depends on the position).  Prefixes are a scan with the operator
// (a, b) -> rotl(a, len b) ^ b; a thread's 8 codes, a warp's 256 and a tile's 8 NT codes span multiples of 8 / 64
// positions, so every rotation in the scan is by a compile-time constant and warp/tile carries are plain XORs.
// k-mer p is selected iff hash[p] or hash[p+w-1] equals the minimum of hash[p..p+w-1] (ties kept).  With
// w-1 = 8a + r the window of p = 8b+j is: the suffix of block b from j, blocks b+1..b+a-1 (block minima), and the
// first r+j+1 hashes counted from the start of block b+a (a running minimum over the 8 "last" hashes, seeded with the
// minimum of the first r hashes of that block).
// One CTA per segment, a tile is T = 8 NT codes: every thread owns 8 consecutive codes / s-mer hashes / k-mer decisions
// per tile, kept in registers.  Three stages run per tile with fixed lags (multiples of 8):
//   A  codes [t0, t0+T) -> code ring, prefixes at the block starts -> pq ring
//   B  s-mer hashes for [t0-LH, t0-LH+T)           -> h ring (8 B per s-mer) + per-block minimum + minimum of the
//                                                                first r hashes of each block of 8
//   C  k-mer decisions for [t0-LH-LW, t0-LH-LW+T)  -> 1 bit per position
// Output: 1 bit per code position + selected count per segment.  Algorithmic bytes: 2 read per code + 1/8 written per code.
//
// Rolling form.  Round 1's k_select kept P' and Q' of EVERY position in shared memory (16 bytes per position, two passes over
// the thread's 8 codes to produce them, two 128-bit loads and three run-time rotations per s-mer to consume them).  Here the
// ring holds the prefixes of one position in 8 -- those a thread needs for its FIRST s-mer, P'/Q' at 8j and at 8j + (s mod 8) --
// and the other seven hashes of the thread roll from it as the reference's FastHasher does, with the outgoing and incoming
// characters taken from a 2-byte-per-position code ring and their hashes from tables rotated in advance:
//     fw(i+1) = rotl(fw(i), 1) ^ H[c(i+s)] ^ rotl(H[c(i)], s)               tab_in  = { H[c],          rotl(H[comp c], s-1) }
//     bw(i+1) = rotr(bw(i), 1) ^ rotl(Hc[c(i+s)], s-1) ^ rotr(Hc[c(i)], 1)   tab_out = { rotl(H[c], s), rotr(H[comp c], 1)   }
// (both follow from fw(i) = rotr(P'[i+s],1) ^ rotl(P'[i],s-1), bw(i) = rotl(Q'[i+s],s) ^ Q'[i] and the prefix recurrences).
// Stage A loses its second pass and seven of its eight stores; shared memory per CTA falls from 50 KB to 25 KB (k=1001, w=100,
// 128 threads); with 64 threads per CTA 6.71 -> 5.75 ms on the bench (profiles/r02_select_experiments.md).
struct SelectPlan { uint32_t LH, LW, NBP, NBH, smem_bytes; };
constexpr uint32_t BM_MIRROR = 16;      // windows up to w = 128 read their block minima without wrapping

__global__ void k_build_roll_tables(const uint4* __restrict__ tab_pair, uint4* __restrict__ tab_in, uint4* __restrict__ tab_out, uint32_t s)
{
	const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= MAX_CODE) return;
	const uint4 v = tab_pair[c];
	const uint64_t hf = ((uint64_t)v.y << 32) | v.x, hb = ((uint64_t)v.w << 32) | v.z;
	const uint64_t in_b = rotl64(hb, s - 1), out_f = rotl64(hf, s), out_b = rotr64(hb, 1);
	tab_in[c] = make_uint4(v.x, v.y, (uint32_t)in_b, (uint32_t)(in_b >> 32));
	tab_out[c] = make_uint4((uint32_t)out_f, (uint32_t)(out_f >> 32), (uint32_t)out_b, (uint32_t)(out_b >> 32));
}

// SR = s mod 8 is a template parameter: where the second prefix of a block sits, which words of the 16-code window hold the incoming
// characters and whether they need the 16-bit funnel are then decided at compile time (as run-time values they cost 32 predicated
// moves per thread and tile for the capture alone).
template <int NT, int SR>
__global__ void __launch_bounds__(NT) k_select(SegTable seg, const uint16_t* __restrict__ codes, const uint32_t* __restrict__ seg_c2,
	const uint4* __restrict__ tab_pair, const uint4* __restrict__ tab_in, const uint4* __restrict__ tab_out, uint8_t* __restrict__ selmask,
	uint32_t* __restrict__ seg_nsel, uint32_t k, uint32_t w, SelectPlan plan)
{
	constexpr uint32_t T = NT * 8;
	constexpr uint64_t INF = ~0ull;
	MBGFE_DYN_SMEM(smem_raw);
	__shared__ uint4 small_tab[4], small_in[4], small_out[4];
	__shared__ Xor2 warp_tot[NT / 32];
	__shared__ Xor2 s_carry[2];                                      // P', Q' at the start of the tile (by tile parity)
	__shared__ uint32_t warp_tmp[NT / 32];
	const uint32_t g = blockIdx.x;
	if (g >= seg.n) return;
	const uint32_t n = seg_c2[g];
	if (n < k) { if (threadIdx.x == 0) seg_nsel[g] = 0; return; }
	const uint32_t NBP = plan.NBP, NBH = plan.NBH, LH = plan.LH, LW = plan.LW;
	uint4* pq = (uint4*)smem_raw;                                  // NBP blocks x { (P',Q')[8j], (P',Q')[8j + sr] }
	uint4* cr = pq + (size_t)NBP * 2;                              // NBP + 1 blocks of 8 codes; block NBP mirrors block 0
	uint64_t* hr = (uint64_t*)(cr + NBP + 1);                      // NBH blocks x 10 entries
	uint64_t* bm = hr + (size_t)NBH * 10;                          // NBH block minima + BM_MIRROR entries that repeat the first ones
	uint64_t* pre = bm + NBH + BM_MIRROR;                          // NBH minima of the first r hashes of a block
	const uint64_t out0 = seg.out_base[g];
	const uint16_t* __restrict__ c = codes + out0;
	uint8_t* __restrict__ mask_out = selmask + (out0 >> 3);
	const uint32_t s = k - w + 1;
	const uint32_t n_kmers = n - k + 1, n_smers = n - s + 1;
	const uint32_t a = (w - 1) >> 3, r = (w - 1) & 7;
	const uint32_t sa = s >> 3;
	constexpr uint32_t sr = SR;      // == s & 7 (the host picks the instantiation)
	const uint32_t rot_s1 = (s - 1) & 63, rot_s = s & 63;
	const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	if (threadIdx.x < 4) { small_tab[threadIdx.x] = __ldg(tab_pair + threadIdx.x); small_in[threadIdx.x] = __ldg(tab_in + threadIdx.x); small_out[threadIdx.x] = __ldg(tab_out + threadIdx.x); }
	if (threadIdx.x == 0) { Xor2 z; z.a = 0; z.b = 0; s_carry[0] = z; s_carry[1] = z; }      // P'[0] = Q'[0] = 0
	uint32_t curA = 0;                                             // pq / code block of position t0
	uint32_t curBp = (NBP - (LH >> 3) % NBP) % NBP;                // pq / code block of position t0 - LH
	uint32_t curBh = (NBH - (LH >> 3) % NBH) % NBH;                // h block of s-mer t0 - LH
	uint32_t curC = (NBH - ((LH + LW) >> 3) % NBH) % NBH;          // h block of k-mer t0 - LH - LW
	uint32_t my_sel = 0, parity = 0;
	uint4 cw_next = make_uint4(0, 0, 0, 0);
	if (8 * threadIdx.x < n) cw_next = __ldg((const uint4*)(c + 8 * threadIdx.x));
	__syncthreads();
	for (uint32_t t0 = 0; t0 < n + LH + LW; t0 += T)
	{
		// ------------------------------------------------------------------ stage A: codes and block prefixes of the new tile
		{
			uint32_t blk = curA + threadIdx.x; if (blk >= NBP) blk -= NBP;
			if (t0 <= n)
			{
				const uint32_t j0 = t0 + 8 * threadIdx.x;
				uint4 cw = cw_next;
				cw_next = make_uint4(0, 0, 0, 0);
				if (j0 + T < n) cw_next = __ldg((const uint4*)(c + j0 + T));
				// codes at and beyond n read as 0 (the slot holds whatever an earlier batch left there)
				const uint32_t nv = j0 < n ? min(8u, n - j0) : 0u;
				uint32_t cws[4] = { cw.x, cw.y, cw.z, cw.w };
				if (nv < 8)
				{
#pragma unroll
					for (int q = 0; q < 4; q++) { if (nv <= 2u * q) cws[q] = 0; else if (nv == 2u * q + 1) cws[q] &= 0xffffu; }
				}
				const uint4 cwm = make_uint4(cws[0], cws[1], cws[2], cws[3]);
				cr[blk] = cwm;
				if (blk == 0) cr[NBP] = cwm;
				uint64_t ip = 0, iq = 0, cap_p = 0, cap_q = 0;
#pragma unroll
				for (int e = 0; e < 8; e++)
				{
					const uint32_t code = (cws[e >> 1] >> ((e & 1) * 16)) & 0xffff;
					uint4 v = small_tab[code & 3];
					if (code >= 4) v = __ldg(tab_pair + code);
					ip = rotl_c(ip ^ u64_of(v.x, v.y), 1); iq = rotr_c(iq ^ u64_of(v.z, v.w), 1);
					if ((uint32_t)e + 1 == sr) { cap_p = ip; cap_q = iq; }          // local prefix after sr codes
				}
				const uint64_t lp = ip, lq = iq;                                        // this block alone
				// inclusive warp scan; the value coming from d lanes below is advanced over the 8d codes gathered so far
#pragma unroll
				for (int d = 1; d < 32; d <<= 1)
				{
					const uint64_t tp = __shfl_up_sync(FULL, ip, d), tq = __shfl_up_sync(FULL, iq, d);
					if (lane >= (uint32_t)d) { ip ^= rotl_c(tp, (8 * d) & 63); iq ^= rotr_c(tq, (8 * d) & 63); }
				}
				if (lane == 31) { Xor2 t; t.a = ip; t.b = iq; warp_tot[wid] = t; }
				uint64_t ep = __shfl_up_sync(FULL, ip, 1), eq = __shfl_up_sync(FULL, iq, 1);
				if (lane == 0) { ep = 0; eq = 0; }
				__syncthreads();
				// a warp spans 256 positions and a tile a multiple of that: the tile's and the lower warps' totals enter as they are
				Xor2 wbase = s_carry[parity];
#pragma unroll
				for (int i = 0; i < NT / 32 - 1; i++) { const Xor2 t = warp_tot[i]; if (i < (int)wid) { wbase.a ^= t.a; wbase.b ^= t.b; } }
				const uint64_t pv = rotl_c(wbase.a, (8 * lane) & 63) ^ ep, qv = rotr_c(wbase.b, (8 * lane) & 63) ^ eq;      // P'[j0], Q'[j0]
				const uint64_t ps = rotl_c(pv, sr) ^ cap_p, qs = rotr_c(qv, sr) ^ cap_q;                                  // P'[j0 + sr], Q'[j0 + sr]
				pq[2 * blk] = make_uint4((uint32_t)pv, (uint32_t)(pv >> 32), (uint32_t)qv, (uint32_t)(qv >> 32));
				pq[2 * blk + 1] = make_uint4((uint32_t)ps, (uint32_t)(ps >> 32), (uint32_t)qs, (uint32_t)(qs >> 32));
				if (threadIdx.x == NT - 1) { Xor2 t; t.a = rotl_c(pv, 8) ^ lp; t.b = rotr_c(qv, 8) ^ lq; s_carry[parity ^ 1] = t; }      // P'[t0 + T]
			}
			else
			{
				const uint4 z = make_uint4(0, 0, 0, 0);
				cr[blk] = z;
				if (blk == 0) cr[NBP] = z;
			}
		}
		__syncthreads();
		// ------------------------------------------------------------------ stage B: s-mer hashes
		{
			const int64_t i0 = (int64_t)t0 - (int64_t)LH + 8 * (int64_t)threadIdx.x;
			if (i0 >= 0)
			{
				uint32_t b0 = curBp + threadIdx.x; if (b0 >= NBP) b0 -= NBP;
				uint32_t b1 = b0 + sa; if (b1 >= NBP) b1 -= NBP;
				const uint32_t i0u = (uint32_t)i0;
				const uint32_t n_ok = i0u < n_smers ? min(8u, n_smers - i0u) : 0u;    // valid s-mers of this block
				uint64_t fw, bw;
				{
					const uint4 x = pq[2 * b0], y = pq[2 * b1 + 1];
					fw = rotr_c(u64_of(y.x, y.y), 1) ^ rotl_v(u64_of(x.x, x.y), rot_s1);
					bw = rotl_v(u64_of(y.z, y.w), rot_s) ^ u64_of(x.z, x.w);
				}
				// outgoing codes c[i0 .. i0+7]: this block; incoming codes c[i0+s .. i0+s+7]: from offset sr of block b1 on
				const uint4 oc = cr[b0];
				uint32_t iw[4];
				{
					const uint4 A = cr[b1], B = cr[b1 + 1];
					uint32_t w5[5];
					switch (sr >> 1)
					{
						case 0: w5[0] = A.x; w5[1] = A.y; w5[2] = A.z; w5[3] = A.w; w5[4] = B.x; break;
						case 1: w5[0] = A.y; w5[1] = A.z; w5[2] = A.w; w5[3] = B.x; w5[4] = B.y; break;
						case 2: w5[0] = A.z; w5[1] = A.w; w5[2] = B.x; w5[3] = B.y; w5[4] = B.z; break;
						default: w5[0] = A.w; w5[1] = B.x; w5[2] = B.y; w5[3] = B.z; w5[4] = B.w; break;
					}
					if (sr & 1)
					{
#pragma unroll
						for (int q = 0; q < 4; q++) iw[q] = (w5[q] >> 16) | (w5[q + 1] << 16);
					}
					else
					{
#pragma unroll
						for (int q = 0; q < 4; q++) iw[q] = w5[q];
					}
				}
				const uint32_t ow[4] = { oc.x, oc.y, oc.z, oc.w };
				uint64_t hv[8];
				uint64_t m = INF, mpre = INF;
#pragma unroll
				for (int e = 0; e < 8; e++)
				{
					hv[e] = fw < bw ? fw : bw;
					if (e < 7)
					{
						const uint32_t co = (ow[e >> 1] >> ((e & 1) * 16)) & 0xffff, ci = (iw[e >> 1] >> ((e & 1) * 16)) & 0xffff;
						uint4 vo = small_out[co & 3], vi = small_in[ci & 3];
						if (co >= 4) vo = __ldg(tab_out + co);
						if (ci >= 4) vi = __ldg(tab_in + ci);
						fw = rotl_c(fw, 1) ^ u64_of(vi.x, vi.y) ^ u64_of(vo.x, vo.y);
						bw = rotr_c(bw, 1) ^ u64_of(vi.z, vi.w) ^ u64_of(vo.z, vo.w);
					}
				}
				if (n_ok < 8)      // only where the segment ends
				{
#pragma unroll
					for (int e = 0; e < 8; e++) if ((uint32_t)e >= n_ok) hv[e] = INF;
				}
#pragma unroll
				for (int e = 0; e < 8; e++)
				{
					if ((uint32_t)e == r) mpre = m;      // minimum of the first r hashes of the block
					m = hv[e] < m ? hv[e] : m;
				}
				uint32_t hbk = curBh + threadIdx.x; if (hbk >= NBH) hbk -= NBH;
				uint4* dst = (uint4*)(hr + (size_t)hbk * 10);
#pragma unroll
				for (int e = 0; e < 4; e++)
					dst[e] = make_uint4((uint32_t)hv[2 * e], (uint32_t)(hv[2 * e] >> 32), (uint32_t)hv[2 * e + 1], (uint32_t)(hv[2 * e + 1] >> 32));
				bm[hbk] = m;
				if (hbk < BM_MIRROR) bm[NBH + hbk] = m;
				pre[hbk] = mpre;
			}
		}
		__syncthreads();
		// ------------------------------------------------------------------ stage C: decisions
		{
			const int64_t p0 = (int64_t)t0 - (int64_t)LH - (int64_t)LW + 8 * (int64_t)threadIdx.x;
			if (p0 >= 0 && p0 < (int64_t)n_kmers)
			{
				uint32_t d0 = curC + threadIdx.x; if (d0 >= NBH) d0 -= NBH;
				uint64_t own[8];
				{
					const uint4* src = (const uint4*)(hr + (size_t)d0 * 10);
#pragma unroll
					for (int e = 0; e < 4; e++) { const uint4 v = src[e]; own[2 * e] = u64_of(v.x, v.y); own[2 * e + 1] = u64_of(v.z, v.w); }
				}
				uint32_t bits = 0;
				if (a == 0)
				{
					// window inside this block and the next one
					uint32_t d1 = d0 + 1; if (d1 >= NBH) d1 -= NBH;
					uint64_t ext[16];
#pragma unroll
					for (int e = 0; e < 8; e++) { ext[e] = own[e]; ext[8 + e] = hr[(size_t)d1 * 10 + e]; }
#pragma unroll
					for (int j = 0; j < 8; j++)
					{
						uint64_t wmin = ext[j], last = ext[j];
#pragma unroll
						for (int t = 1; t < 8; t++) if ((uint32_t)t <= r) { last = ext[j + t]; wmin = last < wmin ? last : wmin; }
						if (own[j] == wmin || last == wmin) bits |= 1u << j;
					}
				}
				else
				{
					uint64_t core = INF;                                 // blocks b+1 .. b+a-1
					{
						// blocks d0+1 .. d0+a-1 of the ring, read through its mirrored head when a <= BM_MIRROR (no wrap), else in two runs
						const uint32_t first = d0 + 1, cnt = a - 1;
						if (a <= BM_MIRROR)
						{
							const uint64_t* q1 = bm + first;
#pragma unroll 4
							for (uint32_t t = 0; t < cnt; t++) { const uint64_t v = q1[t]; core = v < core ? v : core; }
						}
						else
						{
							const uint32_t run1 = first >= NBH ? 0 : min(cnt, NBH - first);
							const uint64_t* q1 = bm + first;
							for (uint32_t t = 0; t < run1; t++) { const uint64_t v = q1[t]; core = v < core ? v : core; }
							const uint64_t* q2 = bm + (first >= NBH ? first - NBH : 0);
							for (uint32_t t = 0; t < cnt - run1; t++) { const uint64_t v = q2[t]; core = v < core ? v : core; }
						}
					}
					uint32_t e0 = d0 + a; if (e0 >= NBH) e0 -= NBH;      // block that holds the window ends
					uint32_t e1 = e0 + 1; if (e1 >= NBH) e1 -= NBH;
					uint64_t suf[8];                                     // suf[j] = min(own[j..7], inner blocks)
					suf[7] = own[7] < core ? own[7] : core;
#pragma unroll
					for (int t = 6; t >= 0; t--) suf[t] = own[t] < suf[t + 1] ? own[t] : suf[t + 1];
					uint64_t run = pre[e0];
					const uint64_t* l0 = hr + (size_t)e0 * 10;
					const uint64_t* l1 = hr + (size_t)e1 * 10;
#pragma unroll
					for (int j = 0; j < 8; j++)
					{
						const uint64_t last = (j + r < 8) ? l0[j + r] : l1[j + r - 8];
						run = last < run ? last : run;
						const uint64_t wmin = run < suf[j] ? run : suf[j];
						if (own[j] == wmin || last == wmin) bits |= 1u << j;
					}
				}
				const uint32_t left = n_kmers - (uint32_t)p0;
				if (left < 8) bits &= (1u << left) - 1;
				mask_out[p0 >> 3] = (uint8_t)bits;
				my_sel += __popc(bits);
			}
		}
		__syncthreads();
		parity ^= 1;
		curA += NT; if (curA >= NBP) curA -= NBP;
		curBp += NT; if (curBp >= NBP) curBp -= NBP;
		curBh += NT; if (curBh >= NBH) curBh -= NBH;
		curC += NT; if (curC >= NBH) curC -= NBH;
	}
	uint32_t total_sel;
	block_exclusive_sum<NT>(my_sel, warp_tmp, total_sel);
	if (threadIdx.x == 0) seg_nsel[g] = total_sel;
}

// --------------------------------------------------------------------------- K3b: positions and k-mer hashes

// exclusive prefix sum of a small uint32 array into uint64 (one CTA; segment tables are small).  carry_in: the sum
// starts from *carry_in (it may alias out[0]: the pipelined run scans one piece of the segment table at a time and
// each piece continues from the total the previous one left in its out[n]); nullptr = 0.
template <int NT>
__global__ void __launch_bounds__(NT) k_scan_u32(const uint32_t* __restrict__ in, uint64_t* out, uint32_t n, const uint64_t* carry_in)
{
	__shared__ uint32_t warp_tmp[NT / 32];
	uint64_t carry = carry_in ? *carry_in : 0;
	__syncthreads();
	for (uint32_t base = 0; base < n; base += NT)
	{
		uint32_t i = base + threadIdx.x;
		uint32_t v = i < n ? in[i] : 0;
		uint32_t tot;
		uint32_t ex = block_exclusive_sum<NT>(v, warp_tmp, tot);
		if (i < n) out[i] = carry + ex;
		carry += tot;
	}
	if (threadIdx.x == 0) out[n] = carry;
}

// two-level exclusive prefix sum for large arrays: tile totals, (k_scan_u32 over the totals), then tile rescans
constexpr int SCAN_NT = 256, SCAN_IPT = 8;
__global__ void __launch_bounds__(SCAN_NT) k_scan_tile_totals(const uint32_t* __restrict__ in, uint32_t* __restrict__ tile_total, uint32_t n)
{
	__shared__ uint32_t warp_tmp[SCAN_NT / 32];
	const uint32_t base = blockIdx.x * (SCAN_NT * SCAN_IPT) + threadIdx.x * SCAN_IPT;
	uint32_t v = 0;
#pragma unroll
	for (int j = 0; j < SCAN_IPT; j++) if (base + j < n) v += in[base + j];
	uint32_t tot;
	block_exclusive_sum<SCAN_NT>(v, warp_tmp, tot);
	if (threadIdx.x == 0) tile_total[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(SCAN_NT) k_scan_tiles(const uint32_t* __restrict__ in, const uint64_t* __restrict__ tile_offset, uint64_t* __restrict__ out, uint32_t n)
{
	__shared__ uint32_t warp_tmp[SCAN_NT / 32];
	const uint32_t base = blockIdx.x * (SCAN_NT * SCAN_IPT) + threadIdx.x * SCAN_IPT;
	uint32_t vals[SCAN_IPT];
	uint32_t v = 0;
#pragma unroll
	for (int j = 0; j < SCAN_IPT; j++) { vals[j] = base + j < n ? in[base + j] : 0; v += vals[j]; }
	uint32_t tot;
	uint64_t run = tile_offset[blockIdx.x] + block_exclusive_sum<SCAN_NT>(v, warp_tmp, tot);
#pragma unroll
	for (int j = 0; j < SCAN_IPT; j++) { if (base + j < n) out[base + j] = run; run += vals[j]; }
	if (blockIdx.x == gridDim.x - 1 && threadIdx.x == SCAN_NT - 1) out[n] = run;
}

// K3b1: selection bitmask -> instance list.  One CTA per segment; inst_pos = part-local k-mer position,
// inst_seg = segment, inst_flag bit0 = first selected k-mer of its part, bit1 = last (the tips, MBG/MBG.cpp:58).
template <int NT>
__global__ void __launch_bounds__(NT) k_list_selected(SegTable seg, uint32_t g_first, const uint32_t* __restrict__ selmask32, const uint32_t* __restrict__ seg_c2,
	const uint32_t* __restrict__ seg_nsel, const uint64_t* __restrict__ sel_begin, uint32_t* __restrict__ inst_pos, uint32_t* __restrict__ inst_seg,
	uint8_t* __restrict__ inst_flag, uint32_t k, uint64_t inst_cap, unsigned long long* __restrict__ overflow)
{
	// seg, seg_c2, seg_nsel and sel_begin are already offset to segment g_first (a piece of the batch); inst_seg holds
	// batch-wide segment numbers.  Instances beyond inst_cap are not written (the caller sized the lists from an
	// estimate; it sees the overflow count and reruns with the exact size).
	__shared__ uint32_t warp_tmp[NT / 32];
	const uint32_t g = blockIdx.x;
	if (g >= seg.n) return;
	const uint32_t nsel = seg_nsel[g];
	if (nsel == 0) return;
	const uint64_t begin = sel_begin[g];
	if (begin + nsel > inst_cap) { if (threadIdx.x == 0) atomicAdd(overflow, 1ull); return; }
	const uint32_t n_kmers = seg_c2[g] - k + 1;
	const uint64_t out0 = seg.out_base[g];
	const uint32_t* __restrict__ mw = selmask32 + (out0 >> 5);
	uint32_t carry = 0;
	for (uint32_t base = 0; base < n_kmers; base += NT * 32)
	{
		uint32_t p0 = base + threadIdx.x * 32;
		uint32_t bits = 0;
		if (p0 < n_kmers)
		{
			bits = mw[p0 >> 5];
			uint32_t nv = n_kmers - p0;
			if (nv < 32) bits &= (1u << nv) - 1;
		}
		uint32_t tot;
		uint32_t idx = block_exclusive_sum<NT>(__popc(bits), warp_tmp, tot) + carry;
		while (bits)
		{
			uint32_t b = __ffs(bits) - 1; bits &= bits - 1;
			inst_pos[begin + idx] = p0 + b;
			inst_seg[begin + idx] = g_first + g;
			inst_flag[begin + idx] = (uint8_t)((idx == 0 ? 1 : 0) | (idx == nsel - 1 ? 2 : 0));
			idx++;
		}
		carry += tot;
	}
}

// libstdc++ std::hash<std::string_view> == _Hash_bytes (MurmurHash64A variant, seed 0xc70f6907) as called by
// MBG/MBGCommon.cpp:25-35 on the raw little-endian bytes of `half` uint16 codes.
// x * Murmur::MUL (mod 2^64) in three multiply-adds: lo*Mlo as a wide product, then lo*Mhi and hi*Mlo added into its high half.
// nvcc's own expansion is IMAD + IMAD + IMAD.WIDE + IMAD.IADD, four slots of the same pipe (K3b: 5.45 -> 5.17 ms).
__device__ __forceinline__ uint64_t mul_murmur(uint64_t x)
{
#if defined(MBGFE_EMU_BUILD)
	return x * 0xc6a4a7935bd1e995ull;
#else
	uint64_t r;
	asm("{\n\t.reg .u32 xl, xh, pl, ph;\n\tmov.b64 {xl, xh}, %1;\n\tmul.wide.u32 %0, xl, 0x5bd1e995;\n\tmov.b64 {pl, ph}, %0;\n\t"
		"mad.lo.u32 ph, xl, 0xc6a4a793, ph;\n\tmad.lo.u32 ph, xh, 0x5bd1e995, ph;\n\tmov.b64 %0, {pl, ph};\n\t}" : "=l"(r) : "l"(x));
	return r;
#endif
}

struct Murmur
{
	static constexpr uint64_t MUL = 0xc6a4a7935bd1e995ull;
	uint64_t h;
	__device__ __forceinline__ void init(uint64_t len_bytes) { h = 0xc70f6907ull ^ (len_bytes * MUL); }
	__device__ __forceinline__ void block(uint64_t wd) { wd *= MUL; wd ^= wd >> 47; wd *= MUL; h ^= wd; h *= MUL; }
	__device__ __forceinline__ void tail(uint64_t wd) { h ^= wd; h *= MUL; }
	__device__ __forceinline__ uint64_t finish() { h ^= h >> 47; h *= MUL; h ^= h >> 47; return h; }
};

// K3b2: 128-bit k-mer hash (MBG/MBGCommon.cpp:25-35): low 64 bits = Murmur over the first (k+1)/2 codes of the k-mer,
// high 64 bits = Murmur over the first (k+1)/2 codes of its reverse complement.  The reverse strand is the same code array
// read backwards and complemented while it is staged, so both halves are the same forward walk over a staged stream.
//
// Murmur64A is h = (h ^ f(block)) * M with f (two of the three multiplications) independent of h, and the selected
// k-mers of a read overlap heavily (one every ~w/2 codes, each (k+1)/2 codes long), so f is shared: one CTA per segment
// walks the code array (and, in step, the reverse-complement array) in tiles of HASH_TILE offsets and alternates
//   premix  f of the 8-byte block at EVERY code offset of the tile -> shared memory (all four alignments at once;
//           coalesced, no divergence, each f serves every k-mer of that alignment that covers it)
//   chain   one thread per k-mer (and direction) whose blocks fall into the tile folds them into its running hash:
//           one shared-memory load, one xor and one multiplication per block
// A k-mer that continues into the next tile parks its running hash in its (not yet final) result slot.  The k-mers of a
// tile are found by binary search in the segment's sorted position list (kept in shared memory when it fits).
// Algorithmic bytes: 2 read per code and direction + 16 written per instance.
constexpr int HASH_NT = 256;
constexpr uint32_t HASH_TILE = 2048, HASH_POS_CAP = 1024;

__device__ __forceinline__ uint32_t lower_bound_u32(const uint32_t* p, uint32_t n, int64_t key)
{
	// first index with p[i] >= key (p ascending)
	uint32_t lo = 0, hi = n;
	while (lo < hi)
	{
		const uint32_t mid = (lo + hi) >> 1;
		if ((int64_t)p[mid] < key) lo = mid + 1; else hi = mid;
	}
	return lo;
}

__device__ __forceinline__ uint32_t complement_pair_swapped(uint32_t wd, const uint16_t* __restrict__ tab_comp)
{
	// two codes in one word -> their complements, halves swapped (the reverse strand reads them in the other order).  Branch-free:
	// the table holds plain bases too (3 - c), so every code is one clamped lookup (entries past the end of a part are whatever the
	// slot held: never used by a chain, but they must not index past the table)
	const uint32_t a = min(wd & 0xffffu, MAX_CODE - 1), b = min(wd >> 16, MAX_CODE - 1);
	return ((uint32_t)__ldg(tab_comp + a) << 16) | (uint32_t)__ldg(tab_comp + b);
}

__global__ void __launch_bounds__(HASH_NT) k_kmer_hash(const uint64_t* __restrict__ seg_out_base, const uint16_t* __restrict__ codes, const uint16_t* __restrict__ tab_comp,
	const uint32_t* __restrict__ inst_pos, const uint64_t* __restrict__ sel_begin, uint64_t inst_cap, uint64_t* __restrict__ inst_hash, uint32_t k)
{
	// seg_out_base and sel_begin are offset to the first segment of the launch; blockIdx.x = segment within it
	__shared__ uint64_t s_f[2][HASH_TILE];
	__align__(16) __shared__ uint16_t s_c[2][HASH_TILE + 8];
	__shared__ uint32_t s_pos[HASH_POS_CAP];
	__shared__ uint32_t s_range[4];
	const uint32_t g = blockIdx.x;
	const uint64_t begin = sel_begin[g], end = sel_begin[g + 1];
	if (end <= begin || end > inst_cap) return;
	const uint32_t n_sel = (uint32_t)(end - begin);
	const uint64_t ob = seg_out_base[g];
	const uint32_t cap = (uint32_t)(seg_out_base[g + 1] - ob);
	// Two streams over the segment's slot of `cap` codes: forward R0[x] = code[x]; reverse R1[x] = complement(code[cap - 1 - x]), in which
	// the reverse complement of the k-mer at pos starts at cap - k - pos and reads forwards.  R1 is made on the fly while staging (cap is
	// a multiple of 64, tiles start on multiples of 8: eight entries of R1 are one aligned 128-bit load of the code array, read backwards
	// and complemented through tab_comp) -- round 1 kept a second code array for it (2 bytes written + 2 read per code).
	const uint16_t* __restrict__ fwd = codes + ob;
	const uint32_t* __restrict__ gpos = inst_pos + begin;
	uint64_t* __restrict__ out = inst_hash + 2 * begin;
	const uint32_t half = (k + 1) / 2;
	const uint32_t nb = half / 4, rem = half & 3;
	const uint32_t reach = nb ? 4 * (nb - 1) : 0;          // offset of a k-mer's last whole block, from its start
	const bool pos_in_smem = n_sel <= HASH_POS_CAP;
	if (pos_in_smem) for (uint32_t i = threadIdx.x; i < n_sel; i += HASH_NT) s_pos[i] = gpos[i];
	const uint32_t* pos = pos_in_smem ? s_pos : gpos;
	const uint32_t pos_first = gpos[0], pos_last = gpos[n_sel - 1];
	// both directions cover the same number of offsets; their first tiles start at the (8-aligned) first block of either
	const uint32_t t0f = pos_first & ~7u, t0r = (cap - k - pos_last) & ~7u;
	const uint32_t span = (pos_last - pos_first) + reach + 4 + 8;
	uint64_t h_init;
	{ Murmur mm; mm.init((uint64_t)half * 2); h_init = mm.h; }
	__syncthreads();
	for (uint32_t toff = 0; toff < span; toff += HASH_TILE)
	{
		const uint32_t T0[2] = { t0f + toff, t0r + toff };
		// ---- stage the codes of both tiles (8 extra for the blocks that start in the last offsets)
		for (uint32_t i = threadIdx.x; i < 2 * (HASH_TILE / 8 + 1); i += HASH_NT)
		{
			const uint32_t d = i >= HASH_TILE / 8 + 1 ? 1 : 0, q = i - d * (HASH_TILE / 8 + 1);
			const uint32_t o = T0[d] + 8 * q;
			uint4 v = make_uint4(0, 0, 0, 0);
			if (o < cap)
			{
				if (d == 0) v = __ldg((const uint4*)(fwd + o));
				else
				{
					const uint4 x = __ldg((const uint4*)(fwd + (cap - 8 - o)));
					v = make_uint4(complement_pair_swapped(x.w, tab_comp), complement_pair_swapped(x.z, tab_comp), complement_pair_swapped(x.y, tab_comp),
						complement_pair_swapped(x.x, tab_comp));
				}
			}
			*(uint4*)(s_c[d] + 8 * q) = v;
		}
		// ---- warp 0 meanwhile finds the k-mers of the tile: forward starts in [T0f - reach, T0f + TILE), reverse starts
		// cap - k - pos in [T0r - reach, T0r + TILE)
		if (threadIdx.x < 4)
		{
			int64_t key;
			switch (threadIdx.x)
			{
				case 0: key = (int64_t)T0[0] - reach; break;
				case 1: key = (int64_t)T0[0] + HASH_TILE; break;
				case 2: key = (int64_t)cap - k - (int64_t)T0[1] - HASH_TILE + 1; break;       // pos > cap - k - T0r - TILE
				default: key = (int64_t)cap - k - (int64_t)T0[1] + reach + 1; break;          // pos <= cap - k - T0r + reach
			}
			s_range[threadIdx.x] = lower_bound_u32(pos, n_sel, key);
		}
		__syncthreads();
		// ---- premix: f(block) for every offset of both tiles that a k-mer can still need.  A thread takes two
		// neighbouring offsets: the same three words give the even block as they are and the odd one shifted by 16 bits
		const uint32_t need = min(HASH_TILE, span - toff);
#pragma unroll
		for (uint32_t d = 0; d < 2; d++)
		{
			const uint32_t* cw = (const uint32_t*)s_c[d];
			uint64_t* fo = s_f[d];
#pragma unroll 2
			for (uint32_t pr = threadIdx.x; 2 * pr < need; pr += HASH_NT)
			{
				const uint32_t w0 = cw[pr], w1 = cw[pr + 1], w2 = cw[pr + 2];
#ifdef MBGFE_EMU_BUILD
				const uint32_t lo1 = (w0 >> 16) | (w1 << 16), hi1 = (w1 >> 16) | (w2 << 16);
#else
				const uint32_t lo1 = __funnelshift_r(w0, w1, 16), hi1 = __funnelshift_r(w1, w2, 16);
#endif
				uint64_t wa = u64_of(w0, w1), wb = u64_of(lo1, hi1);
				wa = mul_murmur(wa); wb = mul_murmur(wb);
				wa ^= wa >> 47; wb ^= wb >> 47;
				wa = mul_murmur(wa); wb = mul_murmur(wb);
				ulonglong2 v; v.x = wa; v.y = wb;
				*(ulonglong2*)(fo + 2 * pr) = v;
			}
		}
		__syncthreads();
		// ---- chain: the tile's share of every k-mer that touches it
		const uint32_t fa = s_range[0], fb = s_range[1], ra = s_range[2], rb = s_range[3];
		const uint32_t nf = fb > fa ? fb - fa : 0, nr = rb > ra ? rb - ra : 0;
		for (uint32_t it = threadIdx.x; it < nf + nr; it += HASH_NT)
		{
			const uint32_t d = it >= nf ? 1 : 0;
			const uint32_t i = d ? ra + (it - nf) : fa + it;
			const uint32_t st = d ? cap - k - pos[i] : pos[i];
			const int64_t rel = (int64_t)st - (int64_t)T0[d];                   // start relative to the tile
			const uint32_t t_lo = rel >= 0 ? 0u : (uint32_t)((-rel + 3) >> 2);
			const int64_t room = (int64_t)HASH_TILE - rel;                       // offsets of the tile from the start on
			const uint32_t t_hi = (uint32_t)min((int64_t)nb, (room + 3) >> 2);
			if (t_lo >= t_hi) continue;
			uint64_t hcur = t_lo == 0 ? h_init : out[2 * i + d];
			const uint64_t* f = s_f[d] + (rel + 4 * (int64_t)t_lo);
			const uint32_t cnt = t_hi - t_lo;
#pragma unroll 4
			for (uint32_t t = 0; t < cnt; t++) { hcur ^= f[4 * t]; hcur = mul_murmur(hcur); }
			if (t_hi == nb)
			{
				Murmur mm; mm.h = hcur;
				if (rem)
				{
					uint64_t wd = 0;
					for (uint32_t j = 0; j < rem; j++)
					{
						const uint32_t x = st + 4 * nb + j;
						const uint32_t cd = d == 0 ? fwd[x] : (uint32_t)__ldg(tab_comp + min((uint32_t)fwd[cap - 1 - x], MAX_CODE - 1));
						wd |= (uint64_t)cd << (16 * j);
					}
					mm.tail(wd);
				}
				hcur = mm.finish();
			}
			out[2 * i + d] = hcur;
		}
		__syncthreads();
	}
}

// ------------------------------------------------------------------ a11: hpc-variant re-hash (pass 3 of runMBG)

// The variant map of ReadpartIterator (MBG/ReadHelper.h:250, filled by addHpcVariants, MBG/ReadHelper.cpp:63-70) as
// sorted arrays: entries ordered by (hash.hi, hash.lo, offset), each with a sorted list of run-length clusters.
struct VariantTable
{
	const uint64_t* hash;        // lo, hi per entry
	const uint32_t* offset;      // code offset inside the k-mer
	const uint64_t* len_begin;   // n + 1
	const uint32_t* len;         // cluster lengths, ascending per entry
	uint64_t n;
};

__device__ __forceinline__ uint64_t variant_lower_bound(const VariantTable& vt, uint64_t lo, uint64_t hi)
{
	uint64_t a = 0, b = vt.n;
	while (a < b)
	{
		const uint64_t m = (a + b) >> 1;
		const uint64_t mlo = __ldg(vt.hash + 2 * m), mhi = __ldg(vt.hash + 2 * m + 1);
		if (mhi < hi || (mhi == hi && mlo < lo)) a = m + 1; else b = m;
	}
	return a;
}

// *std::lower_bound(clusters, lengthHere) (MBG/ReadHelper.h:699, 718)
__device__ __forceinline__ uint64_t variant_cluster(const VariantTable& vt, uint64_t e, uint32_t here)
{
	const uint64_t b = __ldg(vt.len_begin + e), n = __ldg(vt.len_begin + e + 1) - b;
	uint64_t j = 0;
	while (j + 1 < n && __ldg(vt.len + b + j) < here) j++;
	return __ldg(vt.len + b + j);
}

// std::hash<const std::vector<size_t>&> step (MBG/MBGCommon.h:33-45)
__device__ __forceinline__ uint64_t variant_fold(uint64_t h, uint64_t x) { h ^= x; return (h << 3) + (h >> 61); }

// One thread per selected k-mer: if its forward hash (or the swapped, reverse-complement hash) is in the variant map,
// XOR the hash of the clustered run lengths at the recorded offsets into the two halves (MBG/ReadHelper.h:684-730).
// pos2 = per-code positions (poses) of every segment slot; only differences of neighbours are used.
__global__ void __launch_bounds__(256) k_apply_variants(uint64_t* __restrict__ inst_hash, const uint32_t* __restrict__ inst_pos, const uint32_t* __restrict__ inst_seg,
	const uint64_t* __restrict__ range_begin, const uint64_t* __restrict__ range_end, uint64_t inst_cap, const uint64_t* __restrict__ seg_out_base,
	const uint32_t* __restrict__ pos2, VariantTable vt, uint32_t k)
{
	// the instance range is read from device memory (entries of sel_begin), like the other kernels after selection
	const uint64_t i0 = *range_begin, i1 = min(*range_end, inst_cap);
	for (uint64_t i = i0 + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < i1; i += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint64_t lo = inst_hash[2 * i], hi = inst_hash[2 * i + 1];
		const uint32_t* __restrict__ poses = pos2 + seg_out_base[inst_seg[i]] + inst_pos[i];
		uint64_t first = 0, second = 0;
		uint64_t e = variant_lower_bound(vt, lo, hi);
		if (e < vt.n && __ldg(vt.hash + 2 * e) == lo && __ldg(vt.hash + 2 * e + 1) == hi)
		{
			uint64_t end = e;
			while (end < vt.n && __ldg(vt.hash + 2 * end) == lo && __ldg(vt.hash + 2 * end + 1) == hi) end++;
			for (uint64_t j = e; j < end; j++)
			{
				const uint32_t off = __ldg(vt.offset + j);
				if (off <= k / 2) first = variant_fold(first, variant_cluster(vt, j, poses[off + 1] - poses[off]));
			}
			for (uint64_t j = end; j-- > e; )
			{
				const uint32_t off = __ldg(vt.offset + j);
				if (off >= k / 2) second = variant_fold(second, variant_cluster(vt, j, poses[off + 1] - poses[off]));
			}
		}
		else
		{
			e = variant_lower_bound(vt, hi, lo);
			if (!(e < vt.n && __ldg(vt.hash + 2 * e) == hi && __ldg(vt.hash + 2 * e + 1) == lo)) continue;
			uint64_t end = e;
			while (end < vt.n && __ldg(vt.hash + 2 * end) == hi && __ldg(vt.hash + 2 * end + 1) == lo) end++;
			for (uint64_t j = end; j-- > e; )
			{
				const uint32_t off = __ldg(vt.offset + j);
				if (off >= k / 2) first = variant_fold(first, variant_cluster(vt, j, poses[k - off] - poses[k - off - 1]));
			}
			for (uint64_t j = e; j < end; j++)
			{
				const uint32_t off = __ldg(vt.offset + j);
				if (off <= k / 2) second = variant_fold(second, variant_cluster(vt, j, poses[k - off] - poses[k - off - 1]));
			}
		}
		inst_hash[2 * i] = lo ^ first;
		inst_hash[2 * i + 1] = hi ^ second;
	}
}


// instances whose two halves hash alike: palindromes or (practically impossible) collisions
__global__ void __launch_bounds__(256) k_find_palindromes(const uint64_t* __restrict__ inst_hash, const uint64_t* __restrict__ range_begin, const uint64_t* __restrict__ range_end,
	uint64_t inst_cap, unsigned long long* __restrict__ count, uint64_t* __restrict__ list, uint64_t cap)
{
	const uint64_t first = *range_begin, last = min(*range_end, inst_cap);
	for (uint64_t i = first + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < last; i += (uint64_t)gridDim.x * blockDim.x)
	{
		if (inst_hash[2 * i] == inst_hash[2 * i + 1])
		{
			unsigned long long slot = atomicAdd(count, 1ull);
			if (slot < cap) list[slot] = i;
		}
	}
}

// The reference's palindrome rule (MBG/ReadHelper.h:755-785) for the instances found above: a k-mer whose two hashes
// agree must be a true reverse-complement palindrome (else: "unhashable", errors[1]); it is dropped silently when it
// is the first or last selected k-mer of its part or when its neighbours are less than k apart, otherwise the
// reference aborts (errors[0]).  Dropping = clearing its bit in the selection mask; the instance list is then rebuilt.
__global__ void __launch_bounds__(128) k_resolve_palindromes(const uint64_t* __restrict__ list, uint64_t n_list, const uint32_t* __restrict__ inst_pos,
	const uint32_t* __restrict__ inst_seg, const uint8_t* __restrict__ inst_flag, const uint64_t* __restrict__ seg_out_base, const uint16_t* __restrict__ codes,
	const uint16_t* __restrict__ tab_comp, uint32_t* __restrict__ selmask32, uint32_t k, unsigned long long* __restrict__ errors)
{
	const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= n_list) return;
	const uint64_t i = list[t];
	const uint64_t out0 = seg_out_base[inst_seg[i]];
	const uint32_t pos = inst_pos[i];
	const uint16_t* __restrict__ c = codes + out0 + pos;
	bool palindrome = true;
	for (uint32_t j = 0; j < k / 2; j++)
	{
		if (c[j] != tab_comp[c[k - 1 - j]]) { palindrome = false; break; }
	}
	if (!palindrome) { atomicAdd(errors + 1, 1ull); return; }
	const uint32_t fl = inst_flag[i];
	const bool droppable = (fl & 3) != 0 || inst_pos[i + 1] - inst_pos[i - 1] < k;
	if (!droppable) { atomicAdd(errors + 0, 1ull); return; }
	atomicAnd(selmask32 + ((out0 + pos) >> 5), ~(1u << (pos & 31)));
}

// selected count per segment straight from the selection mask (after palindromes were dropped)
template <int NT>
__global__ void __launch_bounds__(NT) k_recount_selected(SegTable seg, const uint32_t* __restrict__ selmask32, const uint32_t* __restrict__ seg_c2, uint32_t* __restrict__ seg_nsel, uint32_t k)
{
	__shared__ uint32_t warp_tmp[NT / 32];
	const uint32_t g = blockIdx.x;
	if (g >= seg.n) return;
	const uint32_t n = seg_c2[g];
	if (n < k) { if (threadIdx.x == 0) seg_nsel[g] = 0; return; }
	const uint32_t n_kmers = n - k + 1;
	const uint32_t* __restrict__ mw = selmask32 + (seg.out_base[g] >> 5);
	uint32_t cnt = 0;
	for (uint32_t p0 = threadIdx.x * 32; p0 < n_kmers; p0 += NT * 32)
	{
		uint32_t bits = mw[p0 >> 5];
		if (n_kmers - p0 < 32) bits &= (1u << (n_kmers - p0)) - 1;
		cnt += __popc(bits);
	}
	uint32_t total;
	block_exclusive_sum<NT>(cnt, warp_tmp, total);
	if (threadIdx.x == 0) seg_nsel[g] = total;
}

// ---------------------------------------------------------------------------------------- K4: k-mer table

// Node table: open addressing, linear probing, 32-byte slots (one DRAM sector):
//   key_lo, key_hi : canonical hash (min of forward and swapped hash); all-ones = empty (cannot occur: it would need
//                    fw == bw, which is a palindrome and never reaches the table)
//   count          : coverage (HashList::coverage, MBG/HashList.cpp:120-149)
//   flags          : bit 0 tip (HashList::setTipKmer)
//   rank           : node id after the export sort (HashList::sortByHash)
struct __align__(32) NodeSlot { uint64_t key_lo, key_hi; uint32_t count, flags; uint64_t rank; };

__device__ __forceinline__ uint64_t slot_hash(uint64_t lo, uint64_t hi)
{
	uint64_t x = lo ^ (hi * 0x9e3779b97f4a7c15ull);
	x ^= x >> 32;
	return x;
}

// 128-bit compare-and-swap (sm_90+): returns true when the slot held {exp_lo, exp_hi} and now holds {lo, hi};
// old_* always receive what the slot held.  load128 is one 16-byte access (never a torn pair of halves).
#ifdef MBGFE_EMU_BUILD
__device__ __forceinline__ bool cas128(NodeSlot* slot, uint64_t exp_lo, uint64_t exp_hi, uint64_t lo, uint64_t hi, uint64_t& old_lo, uint64_t& old_hi)
{
	std::lock_guard<std::mutex> guard(emu::big_lock());
	old_lo = slot->key_lo; old_hi = slot->key_hi;
	if (old_lo == exp_lo && old_hi == exp_hi) { slot->key_lo = lo; slot->key_hi = hi; return true; }
	return false;
}
__device__ __forceinline__ void load128(const NodeSlot* slot, uint64_t& lo, uint64_t& hi)
{
	std::lock_guard<std::mutex> guard(emu::big_lock());
	lo = slot->key_lo; hi = slot->key_hi;
}
#else
__device__ __forceinline__ bool cas128(NodeSlot* slot, uint64_t exp_lo, uint64_t exp_hi, uint64_t lo, uint64_t hi, uint64_t& old_lo, uint64_t& old_hi)
{
	asm volatile(
		"{\n\t"
		".reg .b128 cmp, val, old;\n\t"
		"mov.b128 cmp, {%2, %3};\n\t"
		"mov.b128 val, {%4, %5};\n\t"
		"atom.global.relaxed.gpu.cas.b128 old, [%6], cmp, val;\n\t"
		"mov.b128 {%0, %1}, old;\n\t"
		"}"
		: "=l"(old_lo), "=l"(old_hi)
		: "l"(exp_lo), "l"(exp_hi), "l"(lo), "l"(hi), "l"(slot)
		: "memory");
	return old_lo == exp_lo && old_hi == exp_hi;
}
__device__ __forceinline__ void load128(const NodeSlot* slot, uint64_t& lo, uint64_t& hi)
{
	asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(lo), "=l"(hi) : "l"(slot) : "memory");
}
#endif

__device__ __forceinline__ void canonical_key(uint64_t lo, uint64_t hi, uint64_t& key_lo, uint64_t& key_hi, uint32_t& fw)
{
	// forward hash = hi:lo, backward hash = lo:hi (halves swapped, MBG/HashList.cpp:122-127); fw == forward is smaller
	fw = hi < lo ? 1u : 0u;
	key_hi = fw ? hi : lo;
	key_lo = fw ? lo : hi;
}

// Lock-free insert-or-update.  Slots only ever go from empty to one key, so a loaded value with no all-ones half is a
// complete key; a value with an all-ones half may be empty or half-visible and is settled by the CAS, whose return
// value is authoritative.  Returns the slot; is_new reports whether this call created the entry.
// Probe limit: with the tables at most half full a run of this many occupied slots does not happen; it does when the
// pipelined run under-estimated the table, and then the insert gives up (reported, and the exact path reruns) instead of
// crawling through a full table.
constexpr uint32_t MAX_PROBES = 2048;
__device__ __forceinline__ uint64_t node_insert(NodeSlot* __restrict__ table, uint64_t mask, uint64_t key_lo, uint64_t key_hi, uint32_t add, uint32_t flags, bool& is_new)
{
	uint64_t slot = slot_hash(key_lo, key_hi) & mask;
	is_new = false;
	for (uint32_t probes = 0; probes < MAX_PROBES; probes++)
	{
		NodeSlot* sp = table + slot;
		uint64_t cur_lo, cur_hi;
		load128(sp, cur_lo, cur_hi);
		if (cur_lo == ~0ull || cur_hi == ~0ull)
		{
			uint64_t old_lo, old_hi;
			if (cas128(sp, ~0ull, ~0ull, key_lo, key_hi, old_lo, old_hi)) { is_new = true; cur_lo = key_lo; cur_hi = key_hi; }
			else { cur_lo = old_lo; cur_hi = old_hi; }
		}
		if (cur_lo == key_lo && cur_hi == key_hi)
		{
			if (add) atomicAdd(&sp->count, add);
			if (flags) atomicOr(&sp->flags, flags);
			return slot;
		}
		slot = (slot + 1) & mask;
	}
	return ~0ull;     // MAX_PROBES occupied slots in a row: the table (sized from an estimate by the pipelined run) is as good as full
}

// one atomicAdd per warp for a per-thread 0/1 contribution; every thread of the warp must call it
__device__ __forceinline__ void warp_count(unsigned long long* counter, bool pred)
{
	const uint32_t m = __ballot_sync(FULL, pred);
	if ((threadIdx.x & 31) == 0 && m) atomicAdd(counter, (unsigned long long)__popc(m));
}

// warp-aggregated slot allocation: returns the output index for threads with pred; every thread must call it
__device__ __forceinline__ unsigned long long warp_alloc(unsigned long long* cursor, bool pred)
{
	const uint32_t m = __ballot_sync(FULL, pred);
	unsigned long long base = 0;
	if ((threadIdx.x & 31) == 0 && m) base = atomicAdd(cursor, (unsigned long long)__popc(m));
	base = __shfl_sync(FULL, base, 0);
	return base + __popc(m & ((1u << (threadIdx.x & 31)) - 1));
}

__global__ void __launch_bounds__(256) k_table_clear(NodeSlot* __restrict__ table, uint64_t n_slots)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
	{
		NodeSlot s; s.key_lo = ~0ull; s.key_hi = ~0ull; s.count = 0; s.flags = 0; s.rank = 0;
		table[i] = s;
	}
}

// How many DISTINCT k-mers and edges are about to be inserted, before the tables are sized (linear counting: every key sets one bit of
// a bitmap with at least as many bits as there are keys; distinct ~ -M ln(1 - ones / M)).  Sizing the tables from the number of
// INSTANCES gives a table 2 x coverage too large, and every slot of it is filled, scanned twice by the ranking and twice by the
// export: at 30x coverage that is most of the memory traffic of the step after K3b.
__device__ __forceinline__ void sketch_set(uint32_t* __restrict__ bitmap, uint64_t mask, uint64_t h)
{
	const uint64_t bit = (h ^ (h >> 29)) & mask;
	atomicOr(bitmap + (bit >> 5), 1u << (bit & 31));
}

// instances [0, n): nodes into bitmap, edge observations (unordered pair of the two canonical k-mers) into edge_bitmap
__global__ void __launch_bounds__(256) k_sketch_instances(const uint64_t* __restrict__ inst_hash, const uint8_t* __restrict__ inst_flag, uint64_t n_inst,
	uint32_t* __restrict__ bitmap, uint32_t* __restrict__ edge_bitmap, uint64_t mask)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_inst; i += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint64_t lo = inst_hash[2 * i], hi = inst_hash[2 * i + 1];
		uint64_t k_lo, k_hi; uint32_t fw;
		canonical_key(lo, hi, k_lo, k_hi, fw);
		const uint64_t a = slot_hash(k_lo, k_hi);
		sketch_set(bitmap, mask, a);
		if (edge_bitmap && i > 0 && !(inst_flag[i] & 1))
		{
			uint64_t p_lo, p_hi;
			canonical_key(inst_hash[2 * (i - 1)], inst_hash[2 * (i - 1) + 1], p_lo, p_hi, fw);
			const uint64_t b = slot_hash(p_lo, p_hi);
			sketch_set(edge_bitmap, mask, (a + b) * 0x9e3779b97f4a7c15ull ^ (a ^ b));      // symmetric in the two k-mers
		}
	}
}

// records of `stride` 64-bit words whose first one (narrow) or two (wide) words are the key
__global__ void __launch_bounds__(256) k_sketch_records(const uint64_t* __restrict__ rec, uint64_t n_rec, uint32_t stride, uint32_t wide, uint32_t* __restrict__ bitmap, uint64_t mask)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_rec; i += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint64_t a = rec[(uint64_t)stride * i], b = wide ? rec[(uint64_t)stride * i + 1] : (a >> 17);
		sketch_set(bitmap, mask, slot_hash(a, b));
	}
}

__global__ void __launch_bounds__(256) k_sketch_count(const uint32_t* __restrict__ bitmap, uint64_t n_words, unsigned long long* __restrict__ ones)
{
	uint32_t c = 0;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (uint64_t)gridDim.x * blockDim.x) c += __popc(bitmap[i]);
	if (c) atomicAdd(ones, (unsigned long long)c);
}

// K4a: count k-mers.  One thread per instance of [*range_begin, *range_end): canonical key, probe, 128-bit CAS claim,
// atomic add / or (HashList::addNode + setTipKmer, MBG/HashList.cpp:120-149,23-26).
// Algorithmic bytes: 16 read per instance + one 32-byte slot read and written (random access: sector-bound).
__global__ void __launch_bounds__(256) k_insert_nodes(const uint64_t* __restrict__ inst_hash, const uint8_t* __restrict__ inst_flag, const uint64_t* __restrict__ range_begin,
	const uint64_t* __restrict__ range_end, uint64_t inst_cap, NodeSlot* __restrict__ table, uint64_t mask, uint32_t* __restrict__ inst_slot,
	unsigned long long* __restrict__ distinct, unsigned long long* __restrict__ overflow)
{
	const uint64_t first = *range_begin, last = min(*range_end, inst_cap);
	for (uint64_t base = first + (uint64_t)blockIdx.x * blockDim.x; base < last; base += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint64_t i = base + threadIdx.x;
		bool is_new = false;
		if (i < last)
		{
			const uint64_t lo = inst_hash[2 * i], hi = inst_hash[2 * i + 1];
			if (lo == hi) inst_slot[i] = 0xffffffffu; // palindrome (rare; resolved before the table pass that counts)
			else
			{
				uint64_t key_lo, key_hi; uint32_t fw;
				canonical_key(lo, hi, key_lo, key_hi, fw);
				const uint64_t slot = node_insert(table, mask, key_lo, key_hi, 1, inst_flag[i] ? 1u : 0u, is_new);
				if (slot == ~0ull) atomicAdd(overflow, 1ull);
				inst_slot[i] = (uint32_t)slot;
			}
		}
		warp_count(distinct, is_new);
	}
}

// Edge table: 32-byte slots keyed by the canonical (from, to) pair of (node slot << 1 | fw).
//   key      : from << 32 | to ; all-ones = empty
//   ordered  : (global instance index of the later k-mer << 24) | overlap ; atomicMax keeps the LAST observation
//              in input order == what a single reference worker thread leaves (MBG/HashList.cpp:67-72)
//   count    : HashList::addEdgeCoverage (MBG/HashList.cpp:103-112)
struct __align__(32) EdgeSlot { unsigned long long key; unsigned long long ordered; uint32_t count, pad0; uint64_t pad1; };

__global__ void __launch_bounds__(256) k_edge_clear(EdgeSlot* __restrict__ table, uint64_t n_slots)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
	{
		EdgeSlot s; s.key = ~0ull; s.ordered = 0; s.count = 0; s.pad0 = 0; s.pad1 = 0;
		table[i] = s;
	}
}

// Canonical edge between two consecutive selected k-mers of a part (MBG/MBGCommon.cpp:78-89): the pair is reversed
// when the target's canonical hash is smaller (== smaller id after sortByHash), or when both are the same node in
// reverse orientation.  a_id / b_id are whatever names the nodes in the table being built (table slots on one GPU,
// global node ids when the table is sharded); the decision itself only looks at the hashes.
__device__ __forceinline__ unsigned long long canonical_edge_key(uint64_t a_lo, uint64_t a_hi, uint64_t b_lo, uint64_t b_hi, uint32_t a_id, uint32_t b_id)
{
	uint64_t ak_lo, ak_hi, bk_lo, bk_hi; uint32_t afw, bfw;
	canonical_key(a_lo, a_hi, ak_lo, ak_hi, afw);
	canonical_key(b_lo, b_hi, bk_lo, bk_hi, bfw);
	uint32_t from = a_id, to = b_id;
	uint32_t from_fw = afw, to_fw = bfw;
	const bool to_smaller = bk_hi < ak_hi || (bk_hi == ak_hi && bk_lo < ak_lo);
	const bool same = bk_hi == ak_hi && bk_lo == ak_lo;
	if (to_smaller || (same && !to_fw && !from_fw))
	{
		const uint32_t t = from; from = to; to = t;
		const uint32_t f = from_fw; from_fw = !to_fw; to_fw = !f;
	}
	return ((unsigned long long)(((uint64_t)from << 1) | from_fw) << 32) | (((uint64_t)to << 1) | to_fw);
}

// insert-or-update of one edge observation: coverage += 1, overlap of the latest observation in input order.
// Returns 1 when the edge is new, 0 when it was known, 2 when the table is full.
__device__ __forceinline__ uint32_t edge_insert(EdgeSlot* __restrict__ table, uint64_t mask, unsigned long long key, unsigned long long ordered)
{
	uint64_t slot = slot_hash(key, key >> 17) & mask;
	uint32_t is_new = 0;
	for (uint32_t probes = 0; probes < MAX_PROBES; probes++)
	{
		EdgeSlot* sp = table + slot;
		unsigned long long cur = *(volatile unsigned long long*)&sp->key;
		if (cur == ~0ull)
		{
			const unsigned long long old = atomicCAS(&sp->key, ~0ull, key);
			if (old == ~0ull) { is_new = 1; cur = key; } else cur = old;
		}
		if (cur == key)
		{
			atomicAdd(&sp->count, 1u);
			atomicMax(&sp->ordered, ordered);
			return is_new;
		}
		slot = (slot + 1) & mask;
	}
	return 2;
}

// K4b: adjacency.  One thread per instance of the range that has a predecessor in its part: canonical edge,
// coverage += 1, overlap = prev_pos + k - pos (MBG/MBG.cpp:59-67).
__global__ void __launch_bounds__(256) k_insert_edges(const uint64_t* __restrict__ inst_hash, const uint32_t* __restrict__ inst_pos, const uint8_t* __restrict__ inst_flag,
	const uint32_t* __restrict__ inst_slot, const uint64_t* __restrict__ range_begin, const uint64_t* __restrict__ range_end, uint64_t inst_cap,
	EdgeSlot* __restrict__ table, uint64_t mask, unsigned long long* __restrict__ distinct, unsigned long long* __restrict__ overflow, uint32_t k)
{
	const uint64_t first = *range_begin, last = min(*range_end, inst_cap);
	for (uint64_t base = first + (uint64_t)blockIdx.x * blockDim.x; base < last; base += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint64_t i = base + threadIdx.x;
		uint32_t res = 0;
		if (i < last && i > 0 && !(inst_flag[i] & 1) && inst_slot[i] != 0xffffffffu && inst_slot[i - 1] != 0xffffffffu)
		{
			const unsigned long long key = canonical_edge_key(inst_hash[2 * (i - 1)], inst_hash[2 * (i - 1) + 1], inst_hash[2 * i], inst_hash[2 * i + 1], inst_slot[i - 1], inst_slot[i]);
			const uint32_t overlap = inst_pos[i - 1] + k - inst_pos[i];
			res = edge_insert(table, mask, key, ((unsigned long long)i << 24) | overlap);
			if (res == 2) atomicAdd(overflow, 1ull);
		}
		warp_count(distinct, res == 1);
	}
}

// ------------------------------------------------------------------------------------------------ export

// HashList::sortByHash (MBG/HashList.cpp:245-353) on the device: node ids are the ranks of the canonical hashes.
// Canonical hashes are Murmur output, i.e. uniform, so one MSD bucket pass on the top bits of key_hi leaves buckets
// of a few dozen entries; each bucket is then ordered by counting, for every entry, the entries of its bucket with a
// smaller key (keys are distinct).  count -> scan -> scatter -> rank; the rank is also written back into the table
// slot so that edges (keyed by slot) can be renumbered.
struct NodeOut { uint64_t key_lo, key_hi; uint32_t count, flags; uint64_t slot; };

// Bucket coordinate of a canonical hash inside its owner's range: the canonical high word is the smaller of two uniform
// words (see owner_of below), F(x) = x + x (1 - x) makes it uniform again, F * world has the owner in its high word and the
// position inside the owner's range in its low word.  Non-decreasing in key_hi within one owner, so buckets stay ordered.
__device__ __forceinline__ uint64_t bucket_coordinate(uint64_t key_hi, uint32_t world)
{
	const uint64_t f = key_hi + __umul64hi(key_hi, ~key_hi);
	return f * (uint64_t)world;
}

__global__ void __launch_bounds__(256) k_node_bucket_count(const NodeSlot* __restrict__ table, uint64_t n_slots, uint32_t shift, uint32_t world, uint32_t* __restrict__ counts)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint64_t hi = table[i].key_hi, lo = table[i].key_lo;
		if (hi == ~0ull && lo == ~0ull) continue;
		atomicAdd(counts + (shift >= 64 ? 0 : (bucket_coordinate(hi, world) >> shift)), 1u);
	}
}

__global__ void __launch_bounds__(256) k_node_bucket_scatter(const NodeSlot* __restrict__ table, uint64_t n_slots, uint32_t shift, uint32_t world, const uint64_t* __restrict__ bucket_begin,
	uint32_t* __restrict__ fill, NodeOut* __restrict__ tmp)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
	{
		const NodeSlot s = table[i];
		if (s.key_hi == ~0ull && s.key_lo == ~0ull) continue;
		const uint64_t b = shift >= 64 ? 0 : (bucket_coordinate(s.key_hi, world) >> shift);
		const uint64_t o = bucket_begin[b] + atomicAdd(fill + b, 1u);
		NodeOut r; r.key_lo = s.key_lo; r.key_hi = s.key_hi; r.count = s.count; r.flags = s.flags; r.slot = i;
		tmp[o] = r;
	}
}

// one warp per bucket; lanes take entries round-robin and count smaller keys (uniform-address loads: L1 broadcasts)
__global__ void __launch_bounds__(256) k_node_bucket_rank(const NodeOut* __restrict__ tmp, const uint64_t* __restrict__ bucket_begin, uint32_t n_buckets,
	uint64_t* __restrict__ out_hash, uint32_t* __restrict__ out_cov, uint8_t* __restrict__ out_tip, NodeSlot* __restrict__ table)
{
	const uint32_t lane = threadIdx.x & 31;
	const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
	for (uint32_t b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < n_buckets; b += warps)
	{
		const uint64_t begin = bucket_begin[b];
		const uint32_t n = (uint32_t)(bucket_begin[b + 1] - begin);
		const NodeOut* __restrict__ e = tmp + begin;
		for (uint32_t i = lane; i < n; i += 32)
		{
			const NodeOut mine = e[i];
			uint32_t rank = 0;
			for (uint32_t j = 0; j < n; j++)
			{
				const uint64_t hi = e[j].key_hi, lo = e[j].key_lo;
				rank += (hi < mine.key_hi || (hi == mine.key_hi && lo < mine.key_lo)) ? 1u : 0u;
			}
			const uint64_t o = begin + rank;
			out_hash[2 * o] = mine.key_lo; out_hash[2 * o + 1] = mine.key_hi;
			out_cov[o] = mine.count;
			out_tip[o] = (uint8_t)(mine.flags & 1);
			table[mine.slot].rank = o;
		}
	}
}

// Edges: translated from slot pairs to node ids, then ordered by (from, from_fw descending, to, to_fw) with the same
// bucket scheme on `from` (MBG writes a node's forward-side edges first, MBG/MBG.cpp:967-989).
struct EdgeRec { uint32_t from, from_fw, to, to_fw, coverage, overlap; };   // == mbgfe_edge

__device__ __forceinline__ EdgeRec edge_record(const EdgeSlot& s, const NodeSlot* __restrict__ nodes)
{
	const uint32_t a = (uint32_t)(s.key >> 32), b = (uint32_t)s.key;
	EdgeRec r;
	// single GPU: the key names table slots, translated to ids through the ranks left by the node export;
	// sharded table (nodes == nullptr): the key already holds global node ids
	r.from = nodes ? (uint32_t)nodes[a >> 1].rank : (a >> 1); r.from_fw = a & 1;
	r.to = nodes ? (uint32_t)nodes[b >> 1].rank : (b >> 1); r.to_fw = b & 1;
	r.coverage = s.count;
	r.overlap = (uint32_t)(s.ordered & 0xffffff);
	return r;
}

__device__ __forceinline__ uint64_t edge_order_key(const EdgeRec& r)
{
	return ((uint64_t)((r.from << 1) | (1u - r.from_fw)) << 32) | (uint64_t)((r.to << 1) | r.to_fw);
}

__global__ void __launch_bounds__(256) k_edge_bucket_count(const EdgeSlot* __restrict__ table, uint64_t n_slots, const NodeSlot* __restrict__ nodes, uint32_t shift,
	uint32_t id_base, uint32_t* __restrict__ counts)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
	{
		const unsigned long long key = table[i].key;
		if (key == ~0ull) continue;
		const uint32_t from = nodes ? (uint32_t)nodes[key >> 33].rank : (uint32_t)(key >> 33);
		atomicAdd(counts + ((from - id_base) >> shift), 1u);
	}
}

__global__ void __launch_bounds__(256) k_edge_bucket_scatter(const EdgeSlot* __restrict__ table, uint64_t n_slots, const NodeSlot* __restrict__ nodes, uint32_t shift,
	uint32_t id_base, const uint64_t* __restrict__ bucket_begin, uint32_t* __restrict__ fill, EdgeRec* __restrict__ tmp)
{
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
	{
		const EdgeSlot s = table[i];
		if (s.key == ~0ull) continue;
		const EdgeRec r = edge_record(s, nodes);
		const uint32_t b = (r.from - id_base) >> shift;
		tmp[bucket_begin[b] + atomicAdd(fill + b, 1u)] = r;
	}
}

__global__ void __launch_bounds__(256) k_edge_bucket_rank(const EdgeRec* __restrict__ tmp, const uint64_t* __restrict__ bucket_begin, uint32_t n_buckets, EdgeRec* __restrict__ out)
{
	const uint32_t lane = threadIdx.x & 31;
	const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
	for (uint32_t b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < n_buckets; b += warps)
	{
		const uint64_t begin = bucket_begin[b];
		const uint32_t n = (uint32_t)(bucket_begin[b + 1] - begin);
		const EdgeRec* __restrict__ e = tmp + begin;
		for (uint32_t i = lane; i < n; i += 32)
		{
			const EdgeRec mine = e[i];
			const uint64_t mk = edge_order_key(mine);
			uint32_t rank = 0;
			for (uint32_t j = 0; j < n; j++) rank += edge_order_key(e[j]) < mk ? 1u : 0u;
			out[begin + rank] = mine;
		}
	}
}

// ------------------------------------------------------------------------------------- multi-GPU exchange

// Range partition of the canonical hash space with EQUAL SHARES.  The canonical hash is the smaller of (hi:lo) and (lo:hi),
// so its high word is the minimum of two uniform words: P(key_hi < x) = 1 - (1 - x)^2 = x + x (1 - x), far from uniform
// (cutting the range itself into equal parts gives rank 0 of 8 a share of 23 %).  Ranks own equal parts of that
// distribution instead: owner = floor(world * F(key_hi)), F in 64-bit fixed point.  F is non-decreasing in key_hi (its
// increment is 1 + a floor term that drops by at most 1), so ranks still own consecutive hash ranges and concatenating
// their sorted partitions is the global sortByHash order.
__device__ __forceinline__ uint32_t owner_of(uint64_t key_hi, uint32_t world)
{
	const uint64_t f = key_hi + __umul64hi(key_hi, ~key_hi);
	return (uint32_t)__umul64hi(f, (uint64_t)world);
}

// Per-owner allocation without hammering `world` global counters from millions of threads: a shared-memory histogram
// per CTA, one global atomicAdd per owner and CTA, then base + local rank.  Every thread of the CTA must call it
// (owner == NO_OWNER: nothing to place).  Used with a counter array (counting pass: result ignored) and with a
// cursor array (packing pass: result = output index).
constexpr uint32_t MAX_WORLD = 64, NO_OWNER = 0xffffffffu;
__device__ __forceinline__ unsigned long long block_owner_alloc(uint32_t owner, unsigned long long* __restrict__ cursors, uint32_t world)
{
	__shared__ uint32_t hist[MAX_WORLD];
	__shared__ unsigned long long base[MAX_WORLD];
	if (threadIdx.x < MAX_WORLD) hist[threadIdx.x] = 0;
	__syncthreads();
	uint32_t local = 0;
	if (owner != NO_OWNER) local = atomicAdd(&hist[owner], 1u);
	__syncthreads();
	if (threadIdx.x < world && hist[threadIdx.x]) base[threadIdx.x] = atomicAdd(cursors + threadIdx.x, (unsigned long long)hist[threadIdx.x]);
	__syncthreads();
	return owner != NO_OWNER ? base[owner] + local : 0ull;
}

__global__ void __launch_bounds__(256) k_owner_pack(const uint64_t* __restrict__ inst_hash, const uint8_t* __restrict__ inst_flag, uint64_t n_inst, uint32_t world,
	unsigned long long* __restrict__ cursors, uint64_t* __restrict__ records, uint32_t* __restrict__ inst_packpos)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t owner = NO_OWNER;
	uint64_t key_lo = 0, key_hi = 0;
	if (i < n_inst)
	{
		const uint64_t lo = inst_hash[2 * i], hi = inst_hash[2 * i + 1];
		if (lo != hi) { uint32_t fw; canonical_key(lo, hi, key_lo, key_hi, fw); owner = owner_of(key_hi, world); }
	}
	const unsigned long long o = block_owner_alloc(owner, cursors, world);
	if (i >= n_inst) return;
	if (owner == NO_OWNER) { inst_packpos[i] = 0xffffffffu; return; }
	ulonglong4 r; r.x = key_lo; r.y = key_hi; r.z = inst_flag[i] ? 1 : 0; r.w = 0;
	*(ulonglong4*)(records + 4 * o) = r;
	inst_packpos[i] = (uint32_t)o;     // where the owner's answer (the global node id) will come back
}

// owner side: edge records {canonical (from id, to id) key, (global instance index << 24) | overlap} into this rank's edge partition
__global__ void __launch_bounds__(256) k_insert_edge_records(const unsigned long long* __restrict__ rec, uint64_t n_rec, EdgeSlot* __restrict__ table, uint64_t mask,
	unsigned long long* __restrict__ distinct)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	bool is_new = false;
	if (i < n_rec) is_new = edge_insert(table, mask, rec[2 * i], rec[2 * i + 1]) == 1;
	warp_count(distinct, is_new);
}

// ---- the exchange as libmbgfe runs it itself (mbgfe_exchange): counts of BOTH collectives from the hashes alone, one pass

// The rank that will own the edge observation (i - 1, i): the owner of one of its two nodes, `from` (canonical_edge_key's choice, the
// smaller canonical hash) when the low bits of the two canonical hashes agree, else `to`.  Always `from` was round 1's rule, and
// with k-mers owned by hash RANGE it sent the edges to the low ranks (the smaller of two hashes is in the lowest 1/world of the range
// with probability 1 - (1 - 1/world)^2: 75 % of the edges on rank 0 of 2, 23 % instead of 12.5 % on rank 0 of 8 -- the "inserts of
// received records run 1.5-1.9x slower" of profiles/r02_exchange_trace_n2.txt).  The bit is symmetric in the two nodes and the same for
// both strands of an observation, and min and max together are uniform over the ranks.
__device__ __forceinline__ uint32_t edge_owner_of(uint64_t a_lo, uint64_t a_hi, uint64_t b_lo, uint64_t b_hi, uint32_t world)
{
	uint64_t ak_lo, ak_hi, bk_lo, bk_hi; uint32_t afw, bfw;
	canonical_key(a_lo, a_hi, ak_lo, ak_hi, afw);
	canonical_key(b_lo, b_hi, bk_lo, bk_hi, bfw);
	const bool to_smaller = bk_hi < ak_hi || (bk_hi == ak_hi && bk_lo < ak_lo);
	const bool same = bk_hi == ak_hi && bk_lo == ak_lo;
	const bool from_is_b = to_smaller || (same && !bfw && !afw);
	const bool take_from = ((ak_lo ^ bk_lo) & 1) == 0;
	return owner_of((from_is_b == take_from) ? bk_hi : ak_hi, world);
}

__device__ __forceinline__ bool has_edge_observation(uint64_t i, uint64_t n_inst, const uint64_t* __restrict__ inst_hash, const uint8_t* __restrict__ inst_flag)
{
	if (i >= n_inst || i == 0 || (inst_flag[i] & 1)) return false;
	return inst_hash[2 * i] != inst_hash[2 * i + 1] && inst_hash[2 * (i - 1)] != inst_hash[2 * (i - 1) + 1];      // neither is a palindrome
}

// counts[0 .. world) = node records per owner, counts[world .. 2 world) = edge records per owner
__global__ void __launch_bounds__(256) k_owner_count2(const uint64_t* __restrict__ inst_hash, const uint8_t* __restrict__ inst_flag, uint64_t n_inst, uint32_t world,
	unsigned long long* __restrict__ counts)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t owner = NO_OWNER, eowner = NO_OWNER;
	if (i < n_inst)
	{
		const uint64_t lo = inst_hash[2 * i], hi = inst_hash[2 * i + 1];
		if (lo != hi) { uint64_t key_lo, key_hi; uint32_t fw; canonical_key(lo, hi, key_lo, key_hi, fw); owner = owner_of(key_hi, world); }
		if (has_edge_observation(i, n_inst, inst_hash, inst_flag)) eowner = edge_owner_of(inst_hash[2 * (i - 1)], inst_hash[2 * (i - 1) + 1], lo, hi, world);
	}
	block_owner_alloc(owner, counts, world);
	block_owner_alloc(eowner, counts + world, world);
}

// received node records into this rank's partition; the slot each one landed in is kept (ids are answered from it, no second probe)
__global__ void __launch_bounds__(256) k_insert_records2(const uint64_t* __restrict__ rec, uint64_t n_rec, NodeSlot* __restrict__ table, uint64_t mask,
	uint32_t* __restrict__ rec_slot, unsigned long long* __restrict__ distinct, unsigned long long* __restrict__ overflow)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	bool is_new = false;
	if (i < n_rec)
	{
		const uint64_t slot = node_insert(table, mask, rec[4 * i], rec[4 * i + 1], 1, (uint32_t)rec[4 * i + 2] & 1u, is_new);
		if (slot == ~0ull) atomicAdd(overflow, 1ull);
		rec_slot[i] = (uint32_t)slot;
	}
	warp_count(distinct, is_new);
}

__global__ void __launch_bounds__(256) k_slot_ids(const uint32_t* __restrict__ rec_slot, uint64_t n_rec, const NodeSlot* __restrict__ table, uint32_t id_base, uint32_t* __restrict__ ids)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n_rec) ids[i] = id_base + (uint32_t)table[rec_slot[i]].rank;
}

// edge records {canonical (from id, to id) key, (global instance index << 24) | overlap} grouped by edge_owner_of (the rule
// k_owner_count2 counted with)
__global__ void __launch_bounds__(256) k_edge_owner_pack2(const uint64_t* __restrict__ inst_hash, const uint32_t* __restrict__ inst_pos, const uint8_t* __restrict__ inst_flag,
	const uint32_t* __restrict__ inst_packpos, const uint32_t* __restrict__ packed_ids, uint64_t n_inst, uint64_t inst_base, uint32_t k, uint32_t world,
	unsigned long long* __restrict__ cursors, unsigned long long* __restrict__ records)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const bool have = has_edge_observation(i, n_inst, inst_hash, inst_flag);
	uint32_t owner = NO_OWNER;
	if (have) owner = edge_owner_of(inst_hash[2 * (i - 1)], inst_hash[2 * (i - 1) + 1], inst_hash[2 * i], inst_hash[2 * i + 1], world);
	const unsigned long long o = block_owner_alloc(owner, cursors, world);
	if (have)
	{
		ulonglong2 r;
		r.x = canonical_edge_key(inst_hash[2 * (i - 1)], inst_hash[2 * (i - 1) + 1], inst_hash[2 * i], inst_hash[2 * i + 1], packed_ids[inst_packpos[i - 1]], packed_ids[inst_packpos[i]]);
		r.y = ((unsigned long long)(inst_base + i) << 24) | (inst_pos[i - 1] + k - inst_pos[i]);
		*(ulonglong2*)(records + 2 * o) = r;
	}
}

// KAT helper: complement table and per-code hash are read back directly; the k-mer hash runs k_kmer_hash on one
// synthetic instance.

} // namespace mbgfe_dev
