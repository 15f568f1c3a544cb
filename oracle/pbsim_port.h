// ORACLE — TEST INFRASTRUCTURE ONLY. CPU restatement of the reference's patched PBSIM2 in model mode
// (docker/pbsim2.cpp: set_sim_param :1047-1226, main :391-433,:566, set_hmm_model :2583-2647,
// simulate_by_model :1421-1788, mutate :2260-2413), with glibc's rand() restated (TYPE_3 additive feedback).
// Pinned against the reference itself: oracle/_ref/pbsim2 and pbsim2noError are the UNMODIFIED reference sources
// compiled here, and tests/test_pbsim_port.py requires byte-identical read files for the same arguments and seed.
// Beyond the reads it logs, per read, the index of its first draw in the rand() stream: the input of the GPU
// read kernel's replay mode.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace pbsim {

constexpr int ACCURACY_MAX = 100;
constexpr int STATE_MAX = 50;

struct Params {  // the options docker/haplotype2reads.sh:32 passes, with pbsim's defaults (:1052-1107)
    double depth = 20.0;
    double len_mean = 9000, len_sd = 7000;
    long len_min = 100, len_max = 1000000;
    double accuracy_mean = 0.85;  // already through int(x*100)*0.01 when given on the command line (:1078)
    long sub_ratio = 6, ins_ratio = 54, del_ratio = 40;
    unsigned seed = 0;
    std::string hmm_model_file;
    std::string id_prefix = "S";
    bool no_error = false;  // pbsim2noError.cpp: print mut.seq instead of mut.new_seq
    // Production mode of the GPU read kernel (NOT a reference behaviour; the reference has one global rand() stream
    // whose per-read start is data-dependent): every read draws from its own glibc stream, srand(read_seed(seed,
    // ref_num, read_index)). Everything else is the reference's algorithm, so this mode pins the kernel's
    // per-read-stream output the same way the default mode pins its replay output.
    bool per_read_streams = false;
};

// Seed of read `read_index` (0-based) of sequence `ref_num` (1-based) in per-read-stream mode: splitmix64 finaliser
// of the three values, mapped to [1, 2^31-2].
inline unsigned read_seed(unsigned seed, long ref_num, long read_index) {
    uint64_t z = (uint64_t)seed * 0x9E3779B97F4A7C15ULL + (uint64_t)ref_num * 0xBF58476D1CE4E5B9ULL +
                 (uint64_t)read_index * 0x94D049BB133111EBULL;
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ULL;
    z ^= z >> 27; z *= 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return (unsigned)(1 + z % 2147483646ULL);
}

struct HmmModel {
    double ip[STATE_MAX + 1];
    double ep[STATE_MAX + 1][94];
    double tp[STATE_MAX + 1][STATE_MAX + 1];
    int exist_hmm;
};

// Integer lookup tables built once per run with the host libm (simulate_by_model :1455-1637).
struct Tables {
    long err_thre[94], sub_thre[94], ins_thre[94];
    double sub_rate, ins_rate, del_rate;
    std::vector<long> prob2len;       // [100001]
    std::vector<long> prob2accuracy;  // [100001]
    long len_rand_value = 0, accuracy_rand_value = 0;
    long accuracy_min = 0, accuracy_max = 0;
    int exist_hmm[ACCURACY_MAX + 1];
    std::vector<long> freq2qc;        // [101][1001]
    long qc_rand_value_freq[ACCURACY_MAX + 1];
    std::vector<long> init2state;     // [101][1001]
    std::vector<long> emis2qc;        // [101][51][101]
    std::vector<long> tran2state;     // [101][51][101]
    long qc_rand_value_init[ACCURACY_MAX + 1];
    long qc_rand_value_emis[ACCURACY_MAX + 1][STATE_MAX + 1];
    long qc_rand_value_tran[ACCURACY_MAX + 1][STATE_MAX + 1];
};

struct GlibcRand {  // srand(seed); rand()
    uint32_t r[34];
    int f, b;
    uint64_t count = 0;  // draws taken so far
    explicit GlibcRand(unsigned seed);
    int next();
};

struct ReadLog {
    std::vector<uint64_t> draw_offset;  // index of the read's first rand() call
    std::vector<int64_t> length;        // emitted length
    std::vector<uint8_t> dropped;       // 1 = contained "NNNNN", not printed
    std::vector<uint64_t> draws_used;   // rand() calls the read consumed
    uint64_t total_draws = 0;
};

void set_hmm_model(const std::string& path, std::vector<HmmModel>& hmm);
void build_tables(const Params& p, const std::vector<HmmModel>& hmm, Tables& t);
// Simulates one reference sequence (ref_num is 1-based, as in the read ids); appends the printed lines to `out`.
void simulate_by_model(const Params& p, const Tables& t, const std::string& ref_seq, long ref_num, GlibcRand& rng,
                       std::string& out, ReadLog* log);
// Whole program: every sequence of the FASTA, one output string per sequence (= <prefix>_%04ld.fastq).
void run(const Params& p, const std::string& fasta_path, std::vector<std::string>& outputs, std::vector<ReadLog>* logs);

}  // namespace pbsim
