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


// ---- the file half of the loader (row a1): a plain FASTA file is read, scanned and packed in one pass by worker threads.
// The packed image keeps FILE coordinates: every byte of the file (header lines, newlines and all) becomes one 2-bit
// entry, so a task needs nothing from the tasks before it and a read whose sequence sits on one line is a contiguous range of
// the image -- parse == pack.  What the scan reports are the few characters that are not ACGTacgt: '>' at a line start,
// newlines, and anything else outside a header line; mbgfe_run stitches the record structure together from them in file
// order (MBG/fastqloader.h:38-80 semantics) while later tasks are still being read.
// event = file position << 8 | character
void pack_scan_fasta(const char* src, uint64_t n, uint64_t pos, bool starts_line, uint64_t* dst, std::vector<uint64_t>* events);

struct FileTask { size_t piece; int fd; uint64_t file_off, n; uint64_t* dst; uint64_t pos; std::vector<uint64_t> events; int error = 0; };

class FilePool
{
public:
	~FilePool();
	void start(std::vector<FileTask>&& tasks, size_t n_pieces, unsigned n_threads);
	void wait_piece(size_t piece);
	void finish();
	std::vector<FileTask> tasks;
private:
	std::unique_ptr<std::atomic<uint32_t>[]> left;
	std::atomic<size_t> next { 0 };
	std::vector<std::thread> workers;
};

}
