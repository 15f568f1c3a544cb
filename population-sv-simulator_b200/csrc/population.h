// State of one SimulateHaplotypes run (reference src/SimulateHaplotypes.java:75-116) held inside sp_ctx.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "common.cuh"
#include "model.h"

namespace sp {

struct VariantRec {  // Variant, src/SimulateHaplotypes.java:1181-1198
    int32_t svType = 0, length = 0;
    double alleleFrequency = 0.0;
    int32_t position = 0, intervalType = 0;
    bool isActive = false, hasSequence = false;
    int64_t seq_off = 0, seq_len = 0;  // payload in PopulationState::payload
    int32_t intervalType_start = 0, intervalType_end = 0, intervalSurface = 0;
};

struct PopulationState {
    // arguments (src/SimulateHaplotypes.java:76-86)
    std::string model_dir, population_name, reference_file, repeats_file, segdups_file, insertions_file, output_dir;
    int32_t n_insertion_strings = 0, n_frequent = 0, n_individuals = 0;
    double af_threshold = 0.0;

    Model model;
    PositionSets positions;
    std::vector<VariantRec> variants;
    std::vector<uint8_t> payload;
    int32_t n_variants = 0;
    int32_t nVariants_frequent = 0, nVariants_rare = 0;

    // draw-stream bookkeeping
    uint64_t steps_after_sampling = 0;   // S1
    uint64_t state_after_sampling = 0;   // LCG state after S1 steps
    int32_t candidates = 0;

    // sharding: this context owns haplotypes [h0,h1) of n_haplotypes_total
    int32_t n_haplotypes_total = 0, h0 = 0, h1 = 0;

    // row G/C/N device buffers
    std::vector<int64_t> h_thr;
    DevBuf<int64_t> d_thr, d_off, d_eoff, d_off2;
    DevBuf<uint32_t> d_bits;
    DevBuf<int32_t> d_cnt, d_ids, d_fdeg, d_nedges, d_adj, d_adj_off, d_q, d_where, d_degree, d_cnt2, d_ids2, d_counts;
    DevBuf<int32_t> d_scalars;            // [0] longest pre-collision list, [1] haplotypes the fixer changed
    bool thr_uploaded = false;
    DevBuf<int8_t> d_removed;
    int64_t pre_total = 0;
    bool keep_pre_collision = false;
    std::vector<int64_t> h_pre_off;
    std::vector<int32_t> h_pre_ids;

    // results of rows G/C/N for the local shard
    std::vector<int64_t> h_off;           // local CSR after collision fixing (h1-h0+1)
    std::vector<int32_t> h_ids;
    std::vector<int32_t> h_counts_local;  // carriers per variant among local haplotypes
    bool host_lists_valid = false;        // h_ids / h_counts_local mirror the device (fetch_host_lists)
    int32_t n_haplotypes_with_collision = 0;

    // global view used by the writers (filled by finish(): from the local shard, or from all ranks)
    std::vector<int32_t> counts_global;
    std::vector<int64_t> all_off;         // CSR over ALL haplotypes (only where the joint outputs are written)
    std::vector<int32_t> all_ids;
    bool have_all_lists = false;
    bool begun = false, finished = false;

    // wall-clock seconds per row of SURVEY.md §8a for the last begin/finish (sp_population_timing)
    enum { T_A1 = 0, T_A2, T_A3, T_S1, T_GCN, T_I, T_R1, T_W, T_Z, T_N_ROWS };
    double row_seconds[T_N_ROWS] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
};

void run_genotype_and_collisions(sp_ctx* c, PopulationState& P);
void fetch_host_lists(sp_ctx* c, PopulationState& P);
void sample_variants(PopulationState& P, int64_t seed, int32_t reference_length);
void assign_insertion_strings(PopulationState& P, int64_t seed);
void annotate_variants_with_repeats(PopulationState& P, int maxDistance);
void write_vcfs_device(sp_ctx* c, const PopulationState& P, bool write_joint, bool individuals, int32_t h_first, int32_t h_last,
                       const int64_t* off, const int32_t* ids, int32_t base);
void write_population_files(const PopulationState& P);
void write_histograms(const PopulationState& P, int32_t L);
void load_reference_fasta(sp_ctx* c, const char* path);
void free_population(sp_ctx* c);

}  // namespace sp
