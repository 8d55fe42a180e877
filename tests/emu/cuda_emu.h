// tests/emu/cuda_emu.h -- TEST INFRASTRUCTURE ONLY.
//
// A minimal emulation of the CUDA execution model, so that the kernels in verkko_b200/csrc/kernels.cuh and the
// host orchestration in mbgfe.cu can be compiled with g++ and DEBUGGED in this GPU-less container (thread-block
// barriers, warp shuffles, shared memory, atomics).  It exists only to find logic errors before spending GPU time;
// it is built into tests/emu/libmbgfe_emu.so by tests/emu/Makefile, is never installed next to the product
// library, and nothing in verkko_b200/ loads it.  The product library has no CPU path.
//
// Model: one CTA at a time; every CUDA thread of the CTA is a FIBER (ucontext) of the launching OS thread, run round-robin and
// switched only at barriers: __syncthreads() waits for the CTA, warp collectives (__shfl_*_sync / __ballot_sync / __syncwarp
// with the full mask) for the 32 threads of the warp.  (OS threads with futex barriers, the first form, spent most of the
// CPU suite in the kernel: a barrier over 256 threads on 8 cores is hundreds of context switches.)  Nothing here may spin on
// another thread's progress without reaching a barrier -- the kernels do not.  `__shared__` variables become function-level
// statics (valid because CTAs run one after another).
#pragma once
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <atomic>
#include <memory>
#include <ucontext.h>
#include <thread>
#include <vector>

#define MBGFE_EMU 1
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __shared__ static
#define __constant__ static const
#define __align__(n) __attribute__((aligned(n)))

struct uint3 { unsigned x, y, z; };
struct uint4 { unsigned x, y, z, w; };
struct uint2 { unsigned x, y; };
struct alignas(16) ulonglong2 { unsigned long long x, y; };
struct alignas(32) ulonglong4 { unsigned long long x, y, z, w; };
inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { uint4 v = { x, y, z, w }; return v; }
struct dim3 { unsigned x, y, z; dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };

namespace emu {

void yield();      // to the scheduler: the next fiber of the CTA runs

// generation barrier between fibers of one OS thread: the last arriver opens it, the others yield until it has
struct Barrier
{
	unsigned count = 0, gen = 0, n = 0;
	void reset(unsigned c) { n = c; count = 0; }
	void wait()
	{
		const unsigned g = gen;
		if (++count == n) { count = 0; gen++; return; }
		while (gen == g) yield();
	}
};

inline Barrier& barrier() { static Barrier b; return b; }
// one barrier per warp of the CTA: warp collectives synchronise the 32 threads of a warp only (as the hardware does), which is also
// what keeps the emulator usable -- a barrier over all the OS threads of a 256-thread CTA costs a few hundred context switches
inline Barrier* warp_barriers() { static Barrier b[32]; return b; }
inline uint64_t* xchg() { static uint64_t buf[2048]; return buf; }
inline unsigned char*& dyn_smem() { static unsigned char* p = nullptr; return p; }
inline std::mutex& big_lock() { static std::mutex m; return m; }

}

extern thread_local uint3 threadIdx;
extern thread_local uint3 blockIdx;
extern thread_local dim3 blockDim;
extern thread_local dim3 gridDim;

inline void __syncthreads() { emu::barrier().wait(); }

template <typename T>
inline T __shfl_up_sync(unsigned, T v, unsigned d)
{
	static_assert(sizeof(T) <= 8, "shuffle of <= 8 bytes");
	uint64_t raw = 0; memcpy(&raw, &v, sizeof(T));
	emu::xchg()[threadIdx.x] = raw;
	emu::warp_barriers()[threadIdx.x >> 5].wait();
	unsigned lane = threadIdx.x & 31;
	uint64_t got = lane >= d ? emu::xchg()[threadIdx.x - d] : raw;
	emu::warp_barriers()[threadIdx.x >> 5].wait();
	T out; memcpy(&out, &got, sizeof(T));
	return out;
}

template <typename T>
inline T __shfl_sync(unsigned, T v, unsigned src)
{
	static_assert(sizeof(T) <= 8, "shuffle of <= 8 bytes");
	uint64_t raw = 0; memcpy(&raw, &v, sizeof(T));
	emu::xchg()[threadIdx.x] = raw;
	emu::warp_barriers()[threadIdx.x >> 5].wait();
	uint64_t got = emu::xchg()[(threadIdx.x & ~31u) + (src & 31)];
	emu::warp_barriers()[threadIdx.x >> 5].wait();
	T out; memcpy(&out, &got, sizeof(T));
	return out;
}

inline void __syncwarp(unsigned = 0xffffffffu) { emu::warp_barriers()[threadIdx.x >> 5].wait(); }
inline unsigned long long __umul64hi(unsigned long long a, unsigned long long b) { return (unsigned long long)(((unsigned __int128)a * b) >> 64); }

inline unsigned __ballot_sync(unsigned, int pred)
{
	emu::xchg()[threadIdx.x] = pred ? 1 : 0;
	emu::warp_barriers()[threadIdx.x >> 5].wait();
	unsigned base = threadIdx.x & ~31u, r = 0;
	for (unsigned i = 0; i < 32 && base + i < blockDim.x; i++) if (emu::xchg()[base + i]) r |= 1u << i;
	emu::warp_barriers()[threadIdx.x >> 5].wait();
	return r;
}

inline int __clz(unsigned v) { return v ? __builtin_clz(v) : 32; }
inline int __ffs(unsigned v) { return __builtin_ffs((int)v); }
inline int __ffsll(long long v) { return __builtin_ffsll(v); }
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __popcll(unsigned long long v) { return __builtin_popcountll(v); }
template <typename T> inline T __ldg(const T* p) { return *p; }
template <typename T> inline T __ldcs(const T* p) { return *p; }
inline unsigned __vcmpeq4(unsigned a, unsigned b)
{
	unsigned r = 0;
	for (int i = 0; i < 4; i++) if (((a >> (8 * i)) & 0xff) == ((b >> (8 * i)) & 0xff)) r |= 0xffu << (8 * i);
	return r;
}
using std::max;
using std::min;

template <typename T> inline T atomicAdd(T* p, T v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
template <typename T> inline T atomicOr(T* p, T v) { return __atomic_fetch_or(p, v, __ATOMIC_RELAXED); }
template <typename T> inline T atomicAnd(T* p, T v) { return __atomic_fetch_and(p, v, __ATOMIC_RELAXED); }
template <typename T> inline T atomicMax(T* p, T v)
{
	T old = __atomic_load_n(p, __ATOMIC_RELAXED);
	while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
	return old;
}
template <typename T> inline T atomicCAS(T* p, T expected, T desired)
{
	__atomic_compare_exchange_n(p, &expected, desired, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED);
	return expected;
}

// ------------------------------------------------------------------------------------------- runtime API
typedef int cudaError_t;
typedef void* cudaStream_t;
typedef struct emuEvent { double t; }* cudaEvent_t;
enum { cudaSuccess = 0 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
struct cudaDeviceProp { char name[64]; int major, minor, multiProcessorCount; size_t sharedMemPerBlockOptin; };
inline const char* cudaGetErrorString(cudaError_t) { return "emulator"; }
inline cudaError_t cudaGetLastError() { return 0; }
inline cudaError_t cudaGetDeviceCount(int* n) { *n = 1; return 0; }
inline cudaError_t cudaSetDevice(int) { return 0; }
template <class T> inline cudaError_t cudaGetSymbolAddress(void** p, const T& symbol) { *p = (void*)&symbol; return 0; }
inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp* p, int) { memset(p, 0, sizeof(*p)); strcpy(p->name, "emulated"); p->major = 10; p->minor = 0; p->multiProcessorCount = 4; p->sharedMemPerBlockOptin = 227 * 1024; return 0; }
inline cudaError_t cudaStreamCreate(cudaStream_t* s) { *s = nullptr; return 0; }
inline cudaError_t cudaStreamDestroy(cudaStream_t) { return 0; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
inline cudaError_t cudaDeviceSynchronize() { return 0; }
inline cudaError_t cudaMalloc(void** p, size_t n) { n = n ? n : 1; if (posix_memalign(p, 256, n)) { *p = nullptr; return 2; } memset(*p, 0, n); return 0; }
inline cudaError_t cudaFree(void* p) { free(p); return 0; }
inline cudaError_t cudaMallocHost(void** p, size_t n) { *p = malloc(n ? n : 1); return *p ? 0 : 2; }
inline cudaError_t cudaFreeHost(void* p) { free(p); return 0; }
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t) { memcpy(d, s, n); return 0; }
inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { memcpy(d, s, n); return 0; }
inline cudaError_t cudaMemcpyPeerAsync(void* d, int, const void* s, int, size_t n, cudaStream_t) { memcpy(d, s, n); return 0; }
inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t) { memset(d, v, n); return 0; }
inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = new emuEvent{0}; return 0; }
constexpr unsigned cudaEventDisableTiming = 2;
inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { return cudaEventCreate(e); }
inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return 0; }
inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t) { return 0; }
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return 0; }
inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return 0; }
inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0; return 0; }
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize };
template <typename F> inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return 0; }

namespace emu {

struct Fiber { ucontext_t ctx; bool done = false; };
struct Sched
{
	ucontext_t main_ctx;
	std::vector<Fiber> fibers;
	std::function<void()> body;
	unsigned cur = 0, alive = 0;
};
inline Sched*& sched() { static thread_local Sched* s = nullptr; return s; }
inline std::vector<std::unique_ptr<char[]>>& stacks() { static thread_local std::vector<std::unique_ptr<char[]>> v; return v; }
constexpr size_t FIBER_STACK = 256 * 1024;
inline void yield() { Sched* s = sched(); swapcontext(&s->fibers[s->cur].ctx, &s->main_ctx); }
inline void fiber_entry() { Sched* s = sched(); s->body(); s->fibers[s->cur].done = true; s->alive--; }      // returns to main_ctx (uc_link)

template <typename K, typename... Args>
void launch(K kernel, dim3 grid, dim3 block, size_t smem_bytes, Args... args)
{
	std::vector<unsigned char> smem(smem_bytes + 64);
	dyn_smem() = smem.data();
	Sched s;
	Sched* outer = sched();
	sched() = &s;
	s.body = [&]() { kernel(args...); };
	s.fibers.resize(block.x);
	while (stacks().size() < block.x) stacks().emplace_back(new char[FIBER_STACK]);
	ucontext_t proto;
	getcontext(&proto);
	for (unsigned b = 0; b < grid.x; b++)
	{
		barrier().reset(block.x);
		for (unsigned w = 0; w * 32 < block.x; w++) warp_barriers()[w].reset(std::min(32u, block.x - 32 * w));
		for (unsigned t = 0; t < block.x; t++)
		{
			Fiber& f = s.fibers[t];
			f.ctx = proto;
			f.ctx.uc_stack.ss_sp = stacks()[t].get();
			f.ctx.uc_stack.ss_size = FIBER_STACK;
			f.ctx.uc_link = &s.main_ctx;
			makecontext(&f.ctx, (void (*)())fiber_entry, 0);
			f.done = false;
		}
		s.alive = block.x;
		while (s.alive)
		{
			for (unsigned t = 0; t < block.x; t++)
			{
				if (s.fibers[t].done) continue;
				s.cur = t;
				blockDim = block; gridDim = grid; threadIdx = { t, 0, 0 }; blockIdx = { b, 0, 0 };
				swapcontext(&s.main_ctx, &s.fibers[t].ctx);
			}
		}
	}
	sched() = outer;
}

}

#define MBGFE_LAUNCH(kernel, grid, block, smem, stream, ...) emu::launch(kernel, dim3(grid), dim3(block), (size_t)(smem), __VA_ARGS__)
#define MBGFE_DYN_SMEM(name) unsigned char* name = emu::dyn_smem()
