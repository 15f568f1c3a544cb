// ORACLE — TEST INFRASTRUCTURE ONLY. extern "C" doors into the CPU restatement for ctypes-based tests
// (known-answer vectors of the JDK/glibc primitives, and in-memory calls of the per-row algorithms).
#include "simulate_haplotypes.h"
#include "pbsim_port.h"

extern "C" {

// ---- java.util.Random --------------------------------------------------------------------------
// kind: 0 nextInt(), 1 nextDouble() (bits returned as the 53-bit integer), 2 nextInt(bound), 3 nextBoolean
void oracle_java_random_draws(int64_t seed, int kind, int bound, int n, int64_t* out, uint64_t* state_out) {
    jc::JavaRandom r(seed);
    for (int i = 0; i < n; i++) {
        if (kind == 0) out[i] = r.nextInt();
        else if (kind == 1) { double d = r.nextDouble(); out[i] = (int64_t)(d * 0x1.0p53); }
        else if (kind == 2) out[i] = r.nextInt(bound);
        else out[i] = r.nextBoolean() ? 1 : 0;
    }
    if (state_out) *state_out = r.seed;
}
uint64_t oracle_java_random_state_after(int64_t seed, uint64_t steps) {
    jc::JavaRandom r(seed);
    for (uint64_t i = 0; i < steps; i++) r.next(32);
    return r.seed;
}
int oracle_double_to_string(double d, char* out, int cap) {
    std::string s = jc::doubleToString(d);
    snprintf(out, cap, "%s", s.c_str());
    return (int)s.size();
}
int oracle_number_format4(double d, char* out, int cap) {
    std::string s = jc::numberFormat4(d);
    snprintf(out, cap, "%s", s.c_str());
    return (int)s.size();
}
int oracle_binary_search(const double* a, int from, int to, double key) { return jc::binarySearch(a, from, to, key); }

// ---- java.util.PriorityQueue script runner: ops[i] = (code<<24)|id ; code 0 add, 1 poll, 2 remove,
// 3 = decrement degree[id]. Writes the heap array after every op into trace (n_ops rows of cap ints, -1 padded)
// and poll/remove results into results.
void oracle_pq_script(int n_ids, const int* degree_in, int n_ops, const int* ops, int cap, int* trace, int* results) {
    std::vector<int> degree(degree_in, degree_in + n_ids);
    jc::JavaPriorityQueue q(cap, degree.data());
    for (int i = 0; i < n_ops; i++) {
        int code = (int)((unsigned)ops[i] >> 24), id = ops[i] & 0xFFFFFF;
        int res = -1;
        if (code == 0) q.add(id);
        else if (code == 1) res = q.isEmpty() ? -1 : q.poll();
        else if (code == 2) res = q.remove(id) ? 1 : 0;
        else degree[id]--;
        results[i] = res;
        for (int k = 0; k < cap; k++) trace[(size_t)i * cap + k] = k < q.size ? q.q[k] : -1;
    }
}

// ---- glibc rand() (TYPE_3 additive feedback), restated; pbsim2 uses it at docker/pbsim2.cpp:391,1642-1678,2284-2387
void oracle_glibc_rand(unsigned seed, int n, int* out) {
    int32_t r[34];
    if (seed == 0) seed = 1;
    r[0] = (int32_t)seed;
    for (int i = 1; i < 31; i++) {
        int64_t hi = r[i - 1] / 127773, lo = r[i - 1] % 127773;
        int64_t w = 16807 * lo - 2836 * hi;
        if (w < 0) w += 2147483647;
        r[i] = (int32_t)w;
    }
    std::vector<uint32_t> s(344 + n);
    for (int i = 0; i < 31; i++) s[i] = (uint32_t)r[i];
    for (int i = 31; i < 34; i++) s[i] = s[i - 31];
    for (int i = 34; i < 344 + n; i++) s[i] = s[i - 31] + s[i - 3];
    for (int i = 0; i < n; i++) out[i] = (int)(s[344 + i] >> 1);
}

// ---- Row M in memory: buildHaplotypes loop body (src/SimulateHaplotypes.java:926-953) ---------------
// payload_off has n_variants+1 entries (offsets into payload_blob; equal consecutive offsets with
// has_payload[i]==0 mean sequence==null). Returns the haplotype length, or -1 on a Java exception
// (message in err), or -2 if out_cap is too small.
int64_t oracle_build_haplotype(const char* reference, int reference_length, int n_variants, const int8_t* sv_type,
                               const int32_t* position, const int32_t* length, const int64_t* payload_off,
                               const char* payload_blob, const int8_t* has_payload, const int32_t* ids, int n_ids,
                               char* out, int64_t out_cap, char* err, int err_cap) {
    static thread_local oracle::SimulateHaplotypes* sh = nullptr;
    static thread_local const char* cached_ref = nullptr;
    static thread_local const int32_t* cached_pos = nullptr;
    try {
        if (!sh || cached_ref != reference || cached_pos != position || sh->referenceLength != reference_length ||
            sh->lastVariant != n_variants - 1) {
            delete sh;
            sh = new oracle::SimulateHaplotypes(0);
            sh->verbose = false;
            sh->reference.assign(reference, (size_t)reference_length);
            sh->referenceLength = reference_length;
            sh->variants.assign(n_variants, oracle::Variant());
            sh->lastVariant = n_variants - 1;
            for (int i = 0; i < n_variants; i++) {
                auto& v = sh->variants[i];
                v.svType = sv_type[i]; v.position = position[i]; v.length = length[i];
                v.hasSequence = has_payload ? has_payload[i] != 0 : payload_off[i + 1] > payload_off[i];
                if (v.hasSequence) v.sequence.assign(payload_blob + payload_off[i], (size_t)(payload_off[i + 1] - payload_off[i]));
            }
            cached_ref = reference; cached_pos = position;
        }
        sh->haplotype2variants.assign(1, std::vector<int>(ids, ids + n_ids));
        sh->haplotype2variants_last.assign(1, n_ids - 1);
        std::string s;
        sh->buildHaplotypeString(0, s);
        if ((int64_t)s.size() > out_cap) return -2;
        memcpy(out, s.data(), s.size());
        return (int64_t)s.size();
    } catch (const std::exception& e) {
        if (err && err_cap > 0) snprintf(err, err_cap, "%s", e.what());
        return -1;
    }
}


// ---- Row P in memory: the pbsim2 restatement on one sequence (a single-record FASTA), srand(seed) at the start.
// Returns the bytes of <prefix>_0001.fastq written to `out`, -1 on error (message in err), -2 if a capacity is too
// small. Per-read log arrays (capacity log_cap): first-draw index, emitted length, dropped flag, draws consumed.
int64_t oracle_pbsim_simulate(const char* hmm_model_file, double depth, double len_mean, double len_sd, long len_min,
                              long len_max, double accuracy_mean, long sub_ratio, long ins_ratio, long del_ratio,
                              unsigned seed, const char* id_prefix, int no_error, int per_read_streams, const char* seq,
                              int64_t seq_len, long ref_num, char* out, int64_t out_cap, int64_t* n_reads,
                              uint64_t* draw_offset, int64_t* length, uint8_t* dropped, uint64_t* draws_used,
                              int64_t log_cap, uint64_t* total_draws, char* err, int err_cap) {
    try {
        pbsim::Params p;
        p.depth = depth; p.len_mean = len_mean; p.len_sd = len_sd; p.len_min = len_min; p.len_max = len_max;
        p.accuracy_mean = accuracy_mean; p.sub_ratio = sub_ratio; p.ins_ratio = ins_ratio; p.del_ratio = del_ratio;
        p.seed = seed; p.hmm_model_file = hmm_model_file; p.id_prefix = id_prefix ? id_prefix : "S";
        p.no_error = no_error != 0; p.per_read_streams = per_read_streams != 0;
        std::vector<pbsim::HmmModel> hmm;
        pbsim::set_hmm_model(p.hmm_model_file, hmm);
        pbsim::Tables t;
        pbsim::build_tables(p, hmm, t);
        pbsim::GlibcRand rng(p.seed);
        std::string o;
        pbsim::ReadLog lg;
        pbsim::simulate_by_model(p, t, std::string(seq, (size_t)seq_len), ref_num, rng, o, &lg);
        const int64_t R = (int64_t)lg.length.size();
        if (n_reads) *n_reads = R;
        if (total_draws) *total_draws = lg.total_draws;
        if ((int64_t)o.size() > out_cap || R > log_cap) return -2;
        memcpy(out, o.data(), o.size());
        for (int64_t r = 0; r < R; r++) {
            if (draw_offset) draw_offset[r] = lg.draw_offset[r];
            if (length) length[r] = lg.length[r];
            if (dropped) dropped[r] = lg.dropped[r];
            if (draws_used) draws_used[r] = lg.draws_used[r];
        }
        return (int64_t)o.size();
    } catch (const std::exception& e) {
        if (err && err_cap > 0) snprintf(err, err_cap, "%s", e.what());
        return -1;
    }
}

}  // extern "C"
