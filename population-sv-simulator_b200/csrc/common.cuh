// Shared internals of libsynthpop_cuda.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <stdexcept>
#include "../../include/synthpop_cuda.h"

namespace sp {

struct PopulationState;
struct ReadState;
struct LengthBinState;
struct ReadMixState;
struct AnnotateState;
class FileSink;
struct LengthBinBuilderState;

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

#define SP_CUDA(call)                                                                                      \
    do {                                                                                                   \
        cudaError_t e__ = (call);                                                                          \
        if (e__ != cudaSuccess)                                                                            \
            throw sp::Error(SP_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__) + " (" +     \
                                             __FILE__ + ":" + std::to_string(__LINE__) + ")");            \
    } while (0)

// Device buffer with capacity-only growth (no shrink); contents are not preserved across grow().
template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;  // elements
    // headroom: fraction of extra capacity allocated when the buffer has to grow (avoids cudaFree/cudaMalloc
    // ping-pong when successive batches differ slightly in size)
    void reserve(size_t n, double headroom = 0.25, size_t slack_bytes = 64) {
        if (n <= cap && p) return;
        release();
        const size_t want = n + (size_t)((double)n * headroom);
        if (cudaMalloc((void**)&p, want * sizeof(T) + slack_bytes) != cudaSuccess) {
            cudaGetLastError();
            p = nullptr;
            SP_CUDA(cudaMalloc((void**)&p, n * sizeof(T) + slack_bytes));
            cap = n;
            return;
        }
        cap = want;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
    }
    void swap(DevBuf& o) { T* tp = p; p = o.p; o.p = tp; size_t tc = cap; cap = o.cap; o.cap = tc; }
    ~DevBuf() { release(); }
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
};

// Segment source encoding (seg_src, int64): bits 62..60 = kind, bits 47..0 = byte offset in that space.
enum SegKind : int { SEG_REF_FWD = 0, SEG_REF_RC = 1, SEG_PAYLOAD = 2, SEG_AUX = 3, SEG_PAD = 4 };
constexpr int SEG_KIND_SHIFT = 60;
constexpr int64_t SEG_OFF_MASK = (int64_t(1) << 48) - 1;

constexpr int TILE_BYTES = 16384;  // output bytes per CTA tile of the materialisation kernel

// The copy kernel works on RUNS: maximal pieces of one segment inside one 16 KiB output tile. A run is a plain copy
// (one source pointer, one length), so a tile needs no segment table, no search and no barrier: its descriptor holds the
// first run (offset 0 of the tile) and says where the tile's other runs, if any, sit in the run array. Built per launch by
// tile_count_kernel + a scan + tile_fill_kernel.
struct alignas(16) TileDesc {
    int64_t srcw0;        // first run: kind (bits 62..60) + source byte offset of the tile's first byte (SEG_PAD: nothing to copy)
    uint32_t len0_more;   // bits 0..14: length of the first run (0..16384), bits 15..31: number of further runs
    int32_t item_first;   // index of the tile's second run in the run array
};
struct alignas(16) RunItem {
    int64_t srcw;         // kind + source offset of the run's first output byte (reverse-complemented runs walk downwards from it)
    int32_t off, len;     // output bytes [off, off+len) of the tile
};

}  // namespace sp

struct sp_ctx {
    int device = 0;
    int64_t seed = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    // optional second stream for rows G/C/N (sp_set_population_stream): the genotype kernels and the serial collision
    // fixer of batch s+1 then run under the materialisation kernel of batch s
    // The plan (segment scan, tile table) of a batch is issued on the same second stream, into the plan set the copy kernel
    // of the batch before is NOT reading, so the host round trip that sizes the slots never waits behind a copy kernel.
    cudaStream_t pop_stream = nullptr;
    cudaEvent_t ev_planned = nullptr;   // recorded on the planning stream after tile_runs_kernel of the current plan
    std::string err;
    int64_t launches = 0;
    int sm_count = 148;

    // row A1
    std::vector<uint8_t> h_ref;
    sp::DevBuf<uint8_t> d_ref;
    int64_t ref_len = 0;

    // row V0
    int32_t n_variants = 0;
    std::vector<int8_t> h_type;
    std::vector<int32_t> h_pos, h_len;
    std::vector<int64_t> h_payload_off;
    std::vector<uint8_t> h_payload;
    sp::DevBuf<int8_t> d_type;
    sp::DevBuf<int32_t> d_pos, d_len;
    sp::DevBuf<int64_t> d_payload_off;
    sp::DevBuf<uint8_t> d_payload;
    int32_t n_haplotypes = 0;
    std::vector<int64_t> h_list_off;
    std::vector<int32_t> h_list_ids;
    sp::DevBuf<int64_t> d_list_off;
    sp::DevBuf<int32_t> d_list_ids;
    // the lists sp_plan_haplotypes consumes: a view over the buffers above (rows [0,n_haplotypes)) or over a fresh
    // genotype shard still resident on the device (rows [v_h0, v_h0+v_n))
    const int64_t* v_off_dev = nullptr;
    const int32_t* v_ids_dev = nullptr;
    const int64_t* v_off_host = nullptr;
    int32_t v_h0 = 0, v_n = 0;
    void view_own_lists() {
        v_off_dev = d_list_off.p; v_ids_dev = d_list_ids.p; v_off_host = h_list_off.data(); v_h0 = 0; v_n = n_haplotypes;
    }

    // row M plan (current batch of haplotypes [plan_first..plan_last])
    int32_t plan_first = 0, plan_last = -1;
    int32_t plan_framing = 0;
    int64_t slot_stride = 0;
    int64_t tiles_per_slot = 0;
    std::vector<int64_t> h_hap_len;      // framed lengths
    std::vector<int32_t> h_hdr_len;      // header bytes preceding the sequence in each slot
    struct PlanSet {   // everything sp_materialise reads besides the reference, the payloads and the slots
        sp::DevBuf<int64_t> seg_dst, seg_src, seg_base, aux_off;
        sp::DevBuf<sp::TileDesc> tile_desc;
        sp::DevBuf<sp::RunItem> items;
        sp::DevBuf<uint8_t> aux;
        cudaEvent_t ev_used = nullptr;   // recorded on the materialisation stream after the last copy kernel that read this set
    } plan_set[2];
    int plan_cur = 0;
    sp::DevBuf<int64_t> d_hap_len;
    // run tables of the record-gather uses of the copy kernel (launch_tiled_copy) and scratch of the table builders
    sp::DevBuf<sp::TileDesc> d_copy_desc;
    sp::DevBuf<sp::RunItem> d_copy_items;
    sp::DevBuf<int32_t> d_tile_more, d_tile_s0, d_tile_scan;
    sp::DevBuf<uint8_t> d_tile_tmp;
    int64_t* h_pinned = nullptr;         // small pinned staging area for the size read-backs (plan, rows G/C/N)
    size_t h_pinned_bytes = 0;
    int64_t* pinned(size_t bytes) {
        if (bytes > h_pinned_bytes) {
            if (h_pinned) cudaFreeHost(h_pinned);
            h_pinned = nullptr; h_pinned_bytes = 0;
            const size_t want = bytes < (1u << 16) ? (1u << 16) : bytes * 2;
            if (cudaHostAlloc((void**)&h_pinned, want, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); throw sp::Error(SP_ERR_CUDA, "cudaHostAlloc of the read-back staging area failed"); }
            h_pinned_bytes = want;
        }
        return h_pinned;
    }
    sp::DevBuf<int32_t> d_err;
    sp::DevBuf<uint8_t> d_out;
    bool materialised = false;

    // rows S1/G/C/N/I/R1/W/Z: one SimulateHaplotypes run
    sp::PopulationState* pop = nullptr;

    // row P: read simulation (pbsim2 model mode)
    sp::ReadState* reads = nullptr;
    sp::ReadState* reads_b = nullptr;   // the second batch in flight of sp_reads_mix_add_pair (same configuration and tables)

    // row D: SampleReadsFromLengthBins
    sp::LengthBinState* bins = nullptr;

    // §8(f) rank 1: the tail of docker/haplotype2reads.sh (cat, shuf, tr) and the coverage chunks of reads2svs.sh
    sp::ReadMixState* mix = nullptr;

    // §8(f) rank 2: AnnotateVCF (row A3's interval sets resident in HBM)
    sp::AnnotateState* annot = nullptr;

    // §8(f) rank 3: BuildReadLengthBins — device milliseconds of the last call (line index + binning, draws + swap forest,
    // gather copy), records and bytes written
    double bb_ms[3] = {0.0, 0.0, 0.0};
    int64_t bb_records = 0, bb_bytes = 0;

    // device -> pinned ring -> files, created on first use and kept (pinning memory costs milliseconds per 10 MB)
    sp::FileSink* sink = nullptr;
};

// Entry/exit of every extern "C" function: bind the context's device (failing loudly), turn exceptions into codes.
#define SP_API_BEGIN if (!ctx) return SP_ERR_ARG; try { SP_CUDA(cudaSetDevice(ctx->device));
#define SP_API_END                                                                         \
    return SP_OK; }                                                                        \
    catch (const sp::Error& e) { ctx->err = e.what(); return e.code; }                     \
    catch (const std::exception& e) { ctx->err = e.what(); return SP_ERR; }
