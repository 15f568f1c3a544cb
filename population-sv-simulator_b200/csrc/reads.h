// Row P of SURVEY.md §8a — state of the read simulator (the reference's patched PBSIM2 in model mode,
// docker/pbsim2.cpp) held inside sp_ctx: run parameters, the integer lookup tables simulate_by_model builds with the
// host libm (:1455-1637), their device copies, and the buffers of the current batch of reads.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "common.cuh"

namespace sp {

constexpr int PB_ACCURACY_MAX = 100;  // docker/pbsim2.cpp:28
constexpr int PB_STATE_MAX = 50;      // :29
constexpr int PB_NQ = 94;             // quality codes 0..93

// a % d for a < 2^31 by one widening multiply: q = (a*M) >> sh, M = ceil(2^(31+l)/d), l = ceil(log2 d)
struct FastMod {
    uint32_t M, sh, d, pad;
};
inline FastMod make_fastmod(uint32_t d) {
    FastMod f{0, 0, d, 0};
    if (d == 0) return f;
    uint32_t l = 0;
    while ((1ull << l) < d) l++;
    f.sh = 31 + l;
    f.M = (uint32_t)(((1ull << f.sh) + d - 1) / d);  // l == 0 (d == 1): M = 2^31
    return f;
}

struct ReadParams {  // sim_t after set_sim_param (docker/pbsim2.cpp:1047-1226)
    double depth = 20.0, len_mean = 9000, len_sd = 7000;
    long len_min = 100, len_max = 1000000;
    double accuracy_mean = 0.85;
    long sub_ratio = 6, ins_ratio = 54, del_ratio = 40;
    double sub_rate = 0, ins_rate = 0, del_rate = 0;
    unsigned seed = 0;
    std::string hmm_model_file, prefix = "sd", id_prefix = "S";
    bool no_error = false;
    int stream_mode = 0;  // SP_READS_PER_READ_STREAMS | SP_READS_REPLAY
};

struct ReadTablesHost {
    int32_t thre[3 * PB_NQ];                 // err_thre, sub_thre, ins_thre (:1455-1459)
    std::vector<int32_t> prob2len;           // [100001] (:1461-1490)
    std::vector<uint8_t> prob2accuracy;      // [100001] (:1497-1527)
    long len_rand_value = 0, accuracy_rand_value = 0, accuracy_min = 0, accuracy_max = 0;
    uint8_t exist_hmm[PB_ACCURACY_MAX + 1];
    uint8_t nstates[PB_ACCURACY_MAX + 1];    // highest state index the chain of that accuracy can visit
    std::vector<uint8_t> freq2qc;            // [101][1001] (:1613-1637)
    std::vector<uint8_t> init2state;         // [101][101]  (:1534-1552)
    std::vector<uint8_t> emis2qc, tran2state;  // [101][51][101] (:1553-1612)
    std::vector<FastMod> mod_freq, mod_init; // [101]
    std::vector<FastMod> mod_emis, mod_tran; // [101][51]
};

struct ReadTablesDev {  // what the kernels see
    const int32_t* thre;
    const uint8_t *exist_hmm, *nstates, *freq2qc, *init2state, *emis2qc, *tran2state;
    const FastMod *mod_freq, *mod_init, *mod_emis, *mod_tran;
};

struct ReadState {
    bool configured = false;
    ReadParams par;
    ReadTablesHost tab;
    DevBuf<uint8_t> d_tab;  // one blob, ReadTablesDev points into it
    ReadTablesDev dev{};

    // replay input for the next sp_reads_simulate_* call
    std::vector<int32_t> replay_draws;
    std::vector<int64_t> replay_off;
    bool have_replay = false;

    // current batch
    int64_t n_reads = 0, n_dropped = 0, out_bytes = 0, emitted_bases = 0;
    long ref_num = 1;
    std::vector<int32_t> h_len, h_ndel;
    std::vector<uint8_t> h_acc, h_dropped;
    std::vector<int64_t> h_rec_off;
    std::vector<uint32_t> h_used;
    std::vector<int32_t> h_order;
    bool results_on_host = false;
    DevBuf<int32_t> d_len, d_ndel, d_replay, d_first, d_order;
    DevBuf<uint8_t> d_acc, d_dropped, d_out, d_scratch, d_seq, d_compact;
    DevBuf<int64_t> d_rec_off, d_replay_off, d_koff;
    DevBuf<uint32_t> d_used;
    // per-read plan made on the device (per-read-stream mode): tables of simulate_by_model's loop header and work arrays
    DevBuf<int32_t> d_prob2len, d_delthre, d_key, d_key2, d_idx;
    DevBuf<uint8_t> d_prob2acc, d_cubtmp;
    DevBuf<int64_t> d_len64, d_scan, d_reclen, d_planres;
    bool plan_tables_stale = true;   // set by sp_reads_configure
    bool plan_on_device = false;     // h_len / h_rec_off have not been fetched for the current batch
    DevBuf<int32_t> d_ctr;  // [0] work counter, [1] error code, [2] error read
    // device time of the last read kernel (CUDA events on the launching stream) and host time of the per-read plan
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float kernel_ms = 0.f;
    double plan_ms = 0.0;
    // what launch_batch / collect_batch need of the batch plan_batch made (reads.cu)
    const uint8_t* cur_seq = nullptr;
    int64_t cur_L = 0;
    int32_t cur_max_len = 2;
    bool cur_replay = false;
    unsigned long long* h_ctr = nullptr;    // pinned copy of d_ctr: the read-back does not wait for the kernel
    uint64_t config_epoch = 0;              // bumped by sp_reads_configure: the second state follows the first
    cudaStream_t kernel_stream = nullptr;   // second state only (sp_reads_mix_add_pair): the stream its kernel runs on
    cudaEvent_t ev_ready = nullptr;         // second state only: the context's stream has finished both plans
    float pair_span_ms = 0.f;               // first kernel's start to second kernel's end of the last pair
};

void reads_configure(sp_ctx* c, const sp_read_args* a);
void reads_set_params(ReadState& R, const sp_read_args* a);  // host half of reads_configure
void reads_build_tables(ReadState& R);
void free_reads(sp_ctx* c);
// read_mix.cu: the records of R's last batch become the next source of the mix (the buffers change owner)
void mix_add_from(sp_ctx* c, ReadState& R);

}  // namespace sp
