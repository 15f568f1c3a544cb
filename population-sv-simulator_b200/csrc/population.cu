// Orchestration and C ABI of one SimulateHaplotypes run (reference src/SimulateHaplotypes.java:75-116).
#include "population.h"
#include "java_random.cuh"
#include "textio.h"
#include <algorithm>
#include <chrono>
#include <cmath>

namespace sp {

namespace {
struct Lap {   // seconds since the previous lap, added to a row's total
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    void operator()(double& into) {
        const auto u = std::chrono::steady_clock::now();
        into += std::chrono::duration<double>(u - t).count();
        t = u;
    }
};
}  // namespace

void free_population(sp_ctx* c) {
    if (!c->pop) return;
    // the plan view may point at this state's device lists (sp_population_genotypes): fall back to the context's own
    const bool viewing = c->v_off_dev == c->pop->d_off2.p || c->v_off_host == c->pop->h_off.data();
    if (viewing) { c->view_own_lists(); c->plan_last = -1; c->plan_first = 0; }
    if (c->pop_stream) cudaStreamSynchronize(c->pop_stream);
    cudaStreamSynchronize(c->stream);
    delete c->pop;
    c->pop = nullptr;
}

static std::string need(const char* s, const char* what) {
    if (!s) throw Error(SP_ERR_ARG, std::string("missing argument: ") + what);
    return s;
}

// Ctx variant table (row V0) from the population's variants; payloads are set after assign_insertion_strings.
static void publish_variants(sp_ctx* c, PopulationState& P) {
    const int32_t n = P.n_variants;
    c->n_variants = n;
    c->h_type.resize(n); c->h_pos.resize(n); c->h_len.resize(n);
    c->h_payload_off.assign((size_t)n + 1, 0);
    c->h_payload.clear();
    for (int32_t i = 0; i < n; i++) {
        const VariantRec& v = P.variants[i];
        c->h_type[i] = (int8_t)v.svType; c->h_pos[i] = v.position; c->h_len[i] = v.length;
        if (v.hasSequence) c->h_payload.insert(c->h_payload.end(), P.payload.begin() + v.seq_off, P.payload.begin() + v.seq_off + v.seq_len);
        else if (v.svType == SP_INS) { const char* s = "null"; c->h_payload.insert(c->h_payload.end(), s, s + 4); }
        c->h_payload_off[i + 1] = (int64_t)c->h_payload.size();
    }
    c->d_type.reserve(std::max(n, 1)); c->d_pos.reserve(std::max(n, 1)); c->d_len.reserve(std::max(n, 1));
    c->d_payload_off.reserve(n + 1);
    c->d_payload.reserve(std::max<size_t>(c->h_payload.size(), 1));
    cudaStream_t st = c->stream;
    if (n > 0) {
        SP_CUDA(cudaMemcpyAsync(c->d_type.p, c->h_type.data(), n, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(c->d_pos.p, c->h_pos.data(), (size_t)n * 4, cudaMemcpyHostToDevice, st));
        SP_CUDA(cudaMemcpyAsync(c->d_len.p, c->h_len.data(), (size_t)n * 4, cudaMemcpyHostToDevice, st));
    }
    SP_CUDA(cudaMemcpyAsync(c->d_payload_off.p, c->h_payload_off.data(), ((size_t)n + 1) * 8, cudaMemcpyHostToDevice, st));
    if (!c->h_payload.empty())
        SP_CUDA(cudaMemcpyAsync(c->d_payload.p, c->h_payload.data(), c->h_payload.size(), cudaMemcpyHostToDevice, st));
    SP_CUDA(cudaStreamSynchronize(st));
    c->plan_last = -1; c->plan_first = 0;
}

void population_begin(sp_ctx* c, const sp_population_args* a, int32_t first_h, int32_t last_h) {
    if (!a) throw Error(SP_ERR_ARG, "null sp_population_args");
    free_population(c);
    c->pop = new PopulationState();
    PopulationState& P = *c->pop;
    P.model_dir = need(a->model_dir, "MODEL_DIR");
    P.population_name = need(a->population_name, "POPULATION_NAME");
    P.reference_file = need(a->reference_file, "REFERENCE_FILE");
    P.repeats_file = need(a->repeat_intervals_file, "REPEAT_INTERVALS_FILE");
    P.segdups_file = need(a->segdups_file, "SEGDUPS_FILE");
    P.insertions_file = need(a->insertion_strings_file, "INSERTION_STRINGS_FILE");
    P.output_dir = need(a->output_dir, "OUTPUT_DIR");
    P.n_insertion_strings = a->n_insertion_strings;
    P.af_threshold = a->allele_frequency_threshold;
    P.n_frequent = a->n_variants_frequent;
    P.n_individuals = a->n_individuals;
    if (P.n_insertion_strings < 0) throw Error(SP_ERR_JAVA_EXCEPTION, "NegativeArraySizeException: N_INSERTION_STRINGS");
    if (P.n_individuals < 0 || P.n_individuals > (INT32_MAX >> 1)) throw Error(SP_ERR_JAVA_EXCEPTION, "NegativeArraySizeException: N_INDIVIDUALS");
    P.n_haplotypes_total = P.n_individuals << 1;
    if (last_h < 0) last_h = P.n_haplotypes_total - 1;
    if (first_h < 0 || first_h > last_h + 1 || last_h >= P.n_haplotypes_total) throw Error(SP_ERR_ARG, "bad haplotype shard");
    P.h0 = first_h; P.h1 = last_h + 1;
    P.keep_pre_collision = getenv("SYNTHPOP_KEEP_PRE_COLLISION") != nullptr;

    Lap lap;
    load_reference_fasta(c, P.reference_file.c_str());                  // :93
    lap(P.row_seconds[PopulationState::T_A1]);
    P.model.load_statistics(P.model_dir);                               // :95
    P.model.load_repeat_intervals(P.repeats_file);                      // :96
    P.model.load_segdups(P.segdups_file);                               // :97
    lap(P.row_seconds[PopulationState::T_A2]);
    const int32_t L = (int32_t)c->ref_len;
    P.positions.build(P.model, c->h_ref.data(), L, MAX_DISTANCE_SR);    // :99
    lap(P.row_seconds[PopulationState::T_A3]);
    sample_variants(P, c->seed, L);                                     // :101
    publish_variants(c, P);
    lap(P.row_seconds[PopulationState::T_S1]);
    // nextDouble() <= AF  <=>  draw53 <= floor(AF * 2^53): both sides are exact in binary64
    P.h_thr.resize(P.n_variants);
    for (int32_t v = 0; v < P.n_variants; v++) {
        const double af = P.variants[v].alleleFrequency;
        int64_t t;
        if (!(af >= 0.0)) t = -1;                       // negative (or NaN): never
        else if (af >= 1.0) t = (int64_t)1 << 53;       // always
        else t = (int64_t)std::floor(std::ldexp(af, 53));
        P.h_thr[v] = t;
    }
    run_genotype_and_collisions(c, P);                                  // :104 (loops A and the collision fixer)
    lap(P.row_seconds[PopulationState::T_GCN]);
    P.begun = true; P.finished = false;
}

static void upload_context_lists(sp_ctx* c) {
    const int32_t H = c->n_haplotypes;
    c->d_list_off.reserve(H + 1);
    c->d_list_ids.reserve(std::max<size_t>(c->h_list_ids.size(), 1));
    SP_CUDA(cudaMemcpyAsync(c->d_list_off.p, c->h_list_off.data(), ((size_t)H + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    if (!c->h_list_ids.empty())
        SP_CUDA(cudaMemcpyAsync(c->d_list_ids.p, c->h_list_ids.data(), c->h_list_ids.size() * 4, cudaMemcpyHostToDevice, c->stream));
    SP_CUDA(cudaStreamSynchronize(c->stream));
    c->view_own_lists();
    c->plan_last = -1; c->plan_first = 0;
}

void population_finish(sp_ctx* c, const int32_t* global_counts, const int64_t* all_off, const int32_t* all_ids, int32_t write_joint) {
    if (!c->pop || !c->pop->begun) throw Error(SP_ERR_ARG, "sp_population_begin has not been called");
    PopulationState& P = *c->pop;
    const int32_t V = P.n_variants, H = P.n_haplotypes_total, L = (int32_t)c->ref_len;
    const bool local_is_global = P.h0 == 0 && P.h1 == H;
    fetch_host_lists(c, P);
    if (!global_counts && !local_is_global) throw Error(SP_ERR_ARG, "a sharded run needs the all-gathered carrier counts");
    P.counts_global.assign(V, 0);
    for (int32_t v = 0; v < V; v++) P.counts_global[v] = global_counts ? global_counts[v] : P.h_counts_local[v];
    if (all_off && all_ids) {
        P.all_off.assign(all_off, all_off + H + 1);
        P.all_ids.assign(all_ids, all_ids + all_off[H]);
        P.have_all_lists = true;
    } else if (local_is_global) {
        P.all_off = P.h_off;
        P.all_ids = P.h_ids;
        P.have_all_lists = true;
    }
    // isActive (:693-697) and the empirical allele frequency (:986)
    P.nVariants_frequent = 0; P.nVariants_rare = 0;
    for (int32_t v = 0; v < V; v++) {
        VariantRec& r = P.variants[v];
        r.isActive = P.counts_global[v] > 0;
        if (r.isActive) { if (r.alleleFrequency >= P.af_threshold) P.nVariants_frequent++; else P.nVariants_rare++; }
    }
    for (int32_t v = 0; v < V; v++) P.variants[v].alleleFrequency = ((double)(P.counts_global[v] - 1) + 1.0) / (double)H;
    Lap lap;
    assign_insertion_strings(P, c->seed);                               // :107-109
    lap(P.row_seconds[PopulationState::T_I]);
    annotate_variants_with_repeats(P, 100);                             // :111
    lap(P.row_seconds[PopulationState::T_R1]);
    const bool skip_individual = (write_joint & 2) != 0;
    write_joint &= 1;
    if ((write_joint || !skip_individual) && (int64_t)c->h_ref.size() != c->ref_len)
        throw Error(SP_ERR_ARG, "the VCF writers need the host copy of the reference this population was begun on");
    if (write_joint && !P.have_all_lists) throw Error(SP_ERR_ARG, "the joint outputs need every haplotype's list");
    // the variant table with its payloads goes to the device now: the VCF gather reads the INS payloads from there, and the
    // context is left ready for sp_plan_haplotypes
    publish_variants(c, P);
    if (write_joint) write_histograms(P, L);
    lap(P.row_seconds[PopulationState::T_Z]);
    if (write_joint || !skip_individual)                               // :113 buildVCF, bytes gathered on the device
        write_vcfs_device(c, P, write_joint != 0, !skip_individual, P.h0, P.h1 - 1, P.h_off.data(), P.h_ids.data(), P.h0);
    lap(P.row_seconds[PopulationState::T_W]);
    if (write_joint) write_population_files(P);                        // :114
    lap(P.row_seconds[PopulationState::T_Z]);
    c->n_haplotypes = H;
    if (P.have_all_lists) { c->h_list_off = P.all_off; c->h_list_ids = P.all_ids; }
    else {
        c->h_list_off.assign((size_t)H + 1, 0);
        for (int32_t h = 0; h < H; h++) {
            int64_t n = 0;
            if (h >= P.h0 && h < P.h1) n = P.h_off[h - P.h0 + 1] - P.h_off[h - P.h0];
            c->h_list_off[h + 1] = c->h_list_off[h] + n;
        }
        c->h_list_ids = P.h_ids;
    }
    upload_context_lists(c);
    P.finished = true;
}

// Row T of SURVEY.md §8a (an EXTENSION: the reference's trios are real families, src/GetTrioChildren.java:5-17; no code
// there derives a child). The child of individuals `father` and `mother` inherits one whole haplotype from each parent
// (no recombination), chosen by one nextBoolean() per parent from java.util.Random(seed) — father first. Its two
// haplotypes are appended to the context's lists as haplotypes H and H+1 (so sp_plan_haplotypes / sp_materialise /
// sp_haplotype2reads work on them unchanged, and a child can be a parent in turn) and its truth VCF is the same
// buildVCF merge (:1027-1066) the reference applies to every individual.
void trio_add_child(sp_ctx* c, int32_t father, int32_t mother, int64_t seed, int32_t write_vcf, int32_t* child_first,
                    int32_t* from_father, int32_t* from_mother) {
    if (!c->pop || !c->pop->finished) throw Error(SP_ERR_ARG, "sp_population_finish has not been called");
    PopulationState& P = *c->pop;
    const int32_t H = c->n_haplotypes;
    if (father < 0 || mother < 0 || 2 * father + 1 >= H || 2 * mother + 1 >= H)
        throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: parent individual outside [0," + std::to_string(H / 2) + ")");
    auto local = [&](int32_t ind) { return P.have_all_lists || ind >= P.n_haplotypes_total / 2 || (2 * ind >= P.h0 && 2 * ind + 1 < P.h1); };
    if (!local(father) || !local(mother)) throw Error(SP_ERR_ARG, "trio: the parents' haplotype lists are not on this rank");
    JavaRandom rng(seed);
    const int32_t hf = 2 * father + (rng.nextBoolean() ? 1 : 0);
    const int32_t hm = 2 * mother + (rng.nextBoolean() ? 1 : 0);
    for (int32_t h : {hf, hm}) {
        const int64_t a = c->h_list_off[h], b = c->h_list_off[h + 1];
        c->h_list_ids.reserve(c->h_list_ids.size() + (size_t)(b - a));
        for (int64_t k = a; k < b; k++) c->h_list_ids.push_back(c->h_list_ids[(size_t)k]);
        c->h_list_off.push_back((int64_t)c->h_list_ids.size());
    }
    c->n_haplotypes = H + 2;
    upload_context_lists(c);
    if (write_vcf && (int64_t)c->h_ref.size() != c->ref_len)
        throw Error(SP_ERR_ARG, "the VCF writers need the host copy of the reference this population was begun on");
    if (write_vcf) write_vcfs_device(c, P, false, true, H, H + 1, c->h_list_off.data(), c->h_list_ids.data(), 0);
    if (child_first) *child_first = H;
    if (from_father) *from_father = hf;
    if (from_mother) *from_mother = hm;
}

// Rows G/C/N again for haplotypes [first..last] with the model, variants and thresholds already resident; the
// fresh lists stay on the device and become the view sp_plan_haplotypes consumes (no host round trip of the ids).
void population_genotypes(sp_ctx* c, int32_t first, int32_t last) {
    if (!c->pop || !c->pop->begun) throw Error(SP_ERR_ARG, "sp_population_begin has not been called");
    PopulationState& P = *c->pop;
    if (first < 0 || last >= P.n_haplotypes_total || first > last) throw Error(SP_ERR_ARG, "bad haplotype shard");
    P.h0 = first; P.h1 = last + 1;
    run_genotype_and_collisions(c, P);
    c->v_off_dev = P.d_off2.p; c->v_ids_dev = P.d_ids2.p; c->v_off_host = P.h_off.data();
    c->v_h0 = first; c->v_n = last - first + 1;
    // the plan, segment tables and output slots of the previous batch stay valid until the next sp_plan_haplotypes
}

}  // namespace sp

using sp::Error;


static sp::PopulationState& pop_of(sp_ctx* ctx) {
    if (!ctx->pop || !ctx->pop->begun) throw Error(SP_ERR_ARG, "sp_population_begin has not been called");
    return *ctx->pop;
}

extern "C" {

int sp_simulate_population(sp_ctx* ctx, const sp_population_args* args) {
    SP_API_BEGIN
    sp::population_begin(ctx, args, 0, -1);
    sp::population_finish(ctx, nullptr, nullptr, nullptr, 1);
    SP_API_END
}

int sp_population_begin(sp_ctx* ctx, const sp_population_args* args, int32_t first_haplotype, int32_t last_haplotype) {
    SP_API_BEGIN
    sp::population_begin(ctx, args, first_haplotype, last_haplotype);
    SP_API_END
}

int sp_population_genotypes(sp_ctx* ctx, int32_t first_haplotype, int32_t last_haplotype) {
    SP_API_BEGIN
    sp::population_genotypes(ctx, first_haplotype, last_haplotype);
    SP_API_END
}

int sp_population_info(sp_ctx* ctx, int32_t* n_variants, int32_t* n_haplotypes, int64_t* local_entries, int32_t* n_fixed) {
    SP_API_BEGIN
    sp::PopulationState& P = pop_of(ctx);
    if (n_variants) *n_variants = P.n_variants;
    if (n_haplotypes) *n_haplotypes = P.n_haplotypes_total;
    if (local_entries) *local_entries = P.h_off.empty() ? 0 : P.h_off.back();
    if (n_fixed) *n_fixed = P.n_haplotypes_with_collision;
    SP_API_END
}

int sp_population_local_counts(sp_ctx* ctx, int32_t* counts) {
    SP_API_BEGIN
    sp::PopulationState& P = pop_of(ctx);
    sp::fetch_host_lists(ctx, P);
    if (P.n_variants > 0) memcpy(counts, P.h_counts_local.data(), (size_t)P.n_variants * 4);
    SP_API_END
}

int sp_population_counts_device(sp_ctx* ctx, void** device_counts) {
    SP_API_BEGIN
    sp::PopulationState& P = pop_of(ctx);
    if (device_counts) *device_counts = P.d_counts.p;
    SP_API_END
}

int sp_population_local_lists(sp_ctx* ctx, int64_t* off, int32_t* ids) {
    SP_API_BEGIN
    sp::PopulationState& P = pop_of(ctx);
    sp::fetch_host_lists(ctx, P);
    if (off) memcpy(off, P.h_off.data(), P.h_off.size() * 8);
    if (ids && !P.h_ids.empty()) memcpy(ids, P.h_ids.data(), P.h_ids.size() * 4);
    SP_API_END
}

int sp_population_pre_collision_lists(sp_ctx* ctx, int64_t* off, int32_t* ids, int64_t ids_capacity, int64_t* n_ids) {
    SP_API_BEGIN
    sp::PopulationState& P = pop_of(ctx);
    if (!P.keep_pre_collision) throw Error(SP_ERR_ARG, "set SYNTHPOP_KEEP_PRE_COLLISION=1 before sp_population_begin");
    if (n_ids) *n_ids = (int64_t)P.h_pre_ids.size();
    if (off) memcpy(off, P.h_pre_off.data(), P.h_pre_off.size() * 8);
    if (ids) {
        if ((int64_t)P.h_pre_ids.size() > ids_capacity) throw Error(SP_ERR_ARG, "ids buffer too small");
        if (!P.h_pre_ids.empty()) memcpy(ids, P.h_pre_ids.data(), P.h_pre_ids.size() * 4);
    }
    SP_API_END
}

int sp_population_draw_log(sp_ctx* ctx, uint64_t* steps_after_sample_variants, int32_t* candidates) {
    SP_API_BEGIN
    sp::PopulationState& P = pop_of(ctx);
    if (steps_after_sample_variants) *steps_after_sample_variants = P.steps_after_sampling;
    if (candidates) *candidates = P.candidates;
    SP_API_END
}

int sp_population_timing(sp_ctx* ctx, double* row_seconds, int32_t capacity, int32_t* n_rows) {
    SP_API_BEGIN
    sp::PopulationState& P = pop_of(ctx);
    if (n_rows) *n_rows = sp::PopulationState::T_N_ROWS;
    for (int i = 0; row_seconds && i < capacity && i < sp::PopulationState::T_N_ROWS; i++) row_seconds[i] = P.row_seconds[i];
    SP_API_END
}

int sp_trio_add_child(sp_ctx* ctx, int32_t father_individual, int32_t mother_individual, int64_t seed, int32_t write_vcf,
                      int32_t* child_first_haplotype, int32_t* from_father_haplotype, int32_t* from_mother_haplotype) {
    SP_API_BEGIN
    sp::trio_add_child(ctx, father_individual, mother_individual, seed, write_vcf, child_first_haplotype,
                       from_father_haplotype, from_mother_haplotype);
    SP_API_END
}

int sp_population_finish(sp_ctx* ctx, const int32_t* global_counts, const int64_t* all_off, const int32_t* all_ids,
                         int32_t write_joint_outputs) {
    SP_API_BEGIN
    sp::population_finish(ctx, global_counts, all_off, all_ids, write_joint_outputs);
    SP_API_END
}

}  // extern "C"
