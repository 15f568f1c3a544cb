// Device bytes -> files through a ring of pinned chunks and a writer thread (PrintHaplotypes, haplotype2reads).
#pragma once
#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "common.cuh"

namespace sp {

class FileSink {
public:
    FileSink(int device, int n_slots, int64_t chunk_bytes);
    ~FileSink();
    // Appends bytes [d_src, d_src+n) as the whole content of `path`; copies are issued on `st` in call order. Returns once
    // every copy has been ISSUED into a free pinned chunk (not necessarily written).
    void write_file(const std::string& path, const uint8_t* d_src, int64_t n, cudaStream_t st);
    void drain();    // every issued chunk copied and written; rethrows a writer error
    void finish();

private:
    struct Slot { uint8_t* buf = nullptr; cudaEvent_t ev = nullptr; bool busy = false; };
    struct Task { int slot; int64_t bytes; bool open, close; std::string path; };
    void run();
    int device_;
    int64_t chunk_;
    std::vector<Slot> slots_;
    std::deque<Task> queue_;
    std::mutex m_;
    std::condition_variable cv_;
    std::thread writer_;
    FILE* file_ = nullptr;
    int next_ = 0, pending_ = 0;
    bool stop_ = false, failed_ = false;
    std::string error_;
};

// The context's sink (4 chunks of 8 MiB), created on first use.
FileSink& file_sink(sp_ctx* c);
void free_file_sink(sp_ctx* c);

}  // namespace sp
