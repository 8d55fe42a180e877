// hostpack.h -- host-side 2-bit packing of read pieces on worker threads (see hostpack.cpp)
#pragma once
#include <atomic>
#include <cstdint>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

namespace mbgfe_host {

// n characters at src -> ceil(n / 32) planar words at dst (kernels.cuh layout: bit i of the low / high half = low / high
// bit of base i; A0 C1 G2 T3).  pos = batch position of src[0]; positions of characters outside ACGTacgt (packed as A)
// are appended to *bad.
void pack_planar(const char* src, uint64_t n, uint64_t* dst, uint64_t pos, std::vector<uint64_t>* bad);

struct PackTask { size_t piece; const char* src; uint64_t n; uint64_t* dst; uint64_t pos; };

// Workers take tasks in order; a piece is ready when all of its tasks are done.
class PackPool
{
public:
	~PackPool();
	void start(std::vector<PackTask>&& tasks, size_t n_pieces, unsigned n_threads);
	void wait_piece(size_t piece);
	void finish();                      // joins the workers; `bad` is complete afterwards
	std::vector<uint64_t> bad;
private:
	std::vector<PackTask> tasks;
	std::unique_ptr<std::atomic<uint32_t>[]> left;
	std::atomic<size_t> next { 0 };
	std::vector<std::thread> workers;
	std::mutex lock;
};

}
