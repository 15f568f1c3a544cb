// Row P, host half: run parameters (set_sim_param, docker/pbsim2.cpp:1047-1226), the quality-code HMM file
// (set_hmm_model :2583-2647) and every lookup table simulate_by_model derives from them with libm
// (:1455-1637; qc/uni_ep from main :407-433). The tables are integers by the time the kernels see them, so the
// device never evaluates pow/exp/tgamma: they are computed here, by the same glibc libm the reference links.
#include "reads.h"
#include <cmath>
#include <cstdlib>

namespace sp {

namespace {

struct Hmm {  // hmm_model[accuracy]
    double ip[PB_STATE_MAX + 1];
    double ep[PB_STATE_MAX + 1][PB_NQ];
    double tp[PB_STATE_MAX + 1][PB_STATE_MAX + 1];
    int exist;
};

// set_hmm_model: "<accuracy> IP <state> <p>" | "<accuracy> EP <state> <p0..p93>" | "<accuracy> TP <state> <p1..pN>"
void load_hmm(const std::string& path, std::vector<Hmm>& hmm) {
    hmm.resize(PB_ACCURACY_MAX + 1);
    for (auto& m : hmm) memset(&m, 0, sizeof m);
    FILE* fp = fopen(path.c_str(), "r");
    if (!fp) throw Error(SP_ERR, "ERROR: Cannot open file: " + path);
    std::vector<char> line(10240);  // BUF_SIZE
    auto bad = [&](const char* what) {
        fclose(fp);
        throw Error(SP_ERR_ARG, std::string("hmm model: ") + what + " (the reference reads out of bounds or dereferences NULL here)");
    };
    while (fgets(line.data(), (int)line.size(), fp)) {
        size_t n = strlen(line.data());
        if (n && line[n - 1] == '\n') line[--n] = 0;  // trim() :604-612 strips the line feed only
        char* save = nullptr;
        char* tok = strtok_r(line.data(), " ", &save);
        if (!tok) bad("empty line");
        const int acc = atoi(tok);
        if (acc < 0 || acc > PB_ACCURACY_MAX) bad("accuracy outside 0..100");
        hmm[acc].exist = 1;
        tok = strtok_r(nullptr, " ", &save);
        if (!tok) bad("missing record type");
        const bool ip = !strcmp(tok, "IP"), ep = !strcmp(tok, "EP"), tp = !strcmp(tok, "TP");
        if (!ip && !ep && !tp) continue;
        tok = strtok_r(nullptr, " ", &save);
        if (!tok) bad("missing state");
        const int state = atoi(tok);
        if (state < 0 || state > PB_STATE_MAX) bad("state outside 0..50");
        tok = strtok_r(nullptr, " ", &save);
        if (ip) {
            if (!tok) bad("IP without a value");
            hmm[acc].ip[state] = atof(tok);
        } else if (ep) {
            for (int num = 0; tok; num++, tok = strtok_r(nullptr, " ", &save)) {
                if (num >= PB_NQ) bad("more than 94 EP columns");
                hmm[acc].ep[state][num] = atof(tok);
            }
        } else {
            for (int num = 1; tok; num++, tok = strtok_r(nullptr, " ", &save)) {
                if (num > PB_STATE_MAX) bad("more than 50 TP columns");
                hmm[acc].tp[state][num] = atof(tok);
            }
        }
    }
    fclose(fp);
}

// The reference's recurring idiom: walk a discrete distribution, give outcome `v` the table slots
// (previous cumulative, int(cumulative*scale+0.5)], stop at `scale`. `end` persists across calls exactly where the
// reference does not reset end_wk.
struct SlotFiller {
    long start = 1, end = 0;
    double total = 0.0;
    long scale;
    explicit SlotFiller(long s) : scale(s) {}
    template <class Put>
    bool add(double p, Put&& put) {  // returns true when the table is full
        total += p;
        end = int(total * scale + 0.5);
        if (end > scale) end = scale;
        for (long k = start; k <= end; k++) put(k);
        if (end >= scale) return true;
        start = end + 1;
        return false;
    }
};

}  // namespace

void reads_build_tables(ReadState& R) {
    ReadParams& p = R.par;
    ReadTablesHost& t = R.tab;
    std::vector<Hmm> hmm;
    load_hmm(p.hmm_model_file, hmm);

    double qc_prob[PB_NQ];
    for (int i = 0; i < PB_NQ; i++) qc_prob[i] = pow(10, (double)i / -10);  // main :407-410
    std::vector<double> uni_ep((size_t)(PB_ACCURACY_MAX + 1) * PB_NQ, 0.0);  // main :413-433
    for (int i = 0; i <= PB_ACCURACY_MAX; i++) {
        double* u = &uni_ep[(size_t)i * PB_NQ];
        if (i == PB_ACCURACY_MAX) { u[93] = 1.0; continue; }
        const double prob = 1.0 - i / 100.0;
        for (int j = 0; j < PB_NQ; j++) {
            if (prob == qc_prob[j]) { u[j] = 1.0; break; }
            if (prob > qc_prob[j]) {
                if (j == 0) throw Error(SP_ERR_ARG, "uniform error table: probability above qc[0] (the reference indexes qc[-1])");
                const double rate = (prob - qc_prob[j]) / (qc_prob[j - 1] - qc_prob[j]);
                u[j - 1] = rate;
                u[j] = 1 - rate;
                break;
            }
        }
    }
    for (int i = 0; i < PB_NQ; i++) {  // :1455-1459
        t.thre[i] = int(qc_prob[i] * 1000000 + 0.5);
        t.thre[PB_NQ + i] = int((qc_prob[i] * p.sub_rate) * 1000000 + 0.5);
        t.thre[2 * PB_NQ + i] = int((qc_prob[i] * (p.sub_rate + p.ins_rate)) * 1000000 + 0.5);
    }

    // read length: gamma density over [len_min, len_max] -> 100 000 quantile slots (:1461-1490)
    t.prob2len.assign(100001, 0);
    if (p.len_sd == 0.0) {
        t.prob2len[1] = int(p.len_mean + 0.5);
        t.len_rand_value = 1;
    } else {
        const double variance = pow(p.len_sd, 2);
        const double kappa = pow(p.len_mean, 2) / variance, theta = variance / p.len_mean, gamma = tgamma(kappa);
        SlotFiller f(100000);
        for (long i = p.len_min; i <= p.len_max; i++)
            if (f.add(pow(i, kappa - 1) * exp(-1 * i / theta) / pow(theta, kappa) / gamma, [&](long k) { t.prob2len[k] = (int32_t)i; })) break;
        t.len_rand_value = f.end;
    }
    if (t.len_rand_value < 1) throw Error(SP_ERR_ARG, "ERROR: length parameters are not appropriate.");

    // accuracy class: weights e^(0.22 a) over [floor(.75 m), min(100, floor(1.05 m))] (:1497-1527)
    t.prob2accuracy.assign(100001, 0);
    const double mean = p.accuracy_mean * 100;
    long amax = floor(mean * 1.05), amin = floor(mean * 0.75);
    if (amax > 100) amax = 100;
    if (amin < 0) throw Error(SP_ERR_ARG, "accuracy range below 0");
    t.accuracy_min = amin; t.accuracy_max = amax;
    {
        double freq_total = 0.0;
        for (long i = amin; i <= amax; i++) freq_total += exp(0.22 * i);
        SlotFiller f(100000);
        for (long i = amin; i <= amax; i++)
            if (f.add(exp(0.22 * i) / freq_total, [&](long k) { t.prob2accuracy[k] = (uint8_t)i; })) break;
        t.accuracy_rand_value = f.end;
    }
    if (t.accuracy_rand_value < 1) throw Error(SP_ERR_ARG, "ERROR: accuracy parameters are not appropriate.");

    // quality codes (:1534-1637)
    const size_t NA = PB_ACCURACY_MAX + 1, NS = PB_STATE_MAX + 1;
    t.freq2qc.assign(NA * 1001, 0);
    t.init2state.assign(NA * 101, 0);
    t.emis2qc.assign(NA * NS * 101, 0);
    t.tran2state.assign(NA * NS * 101, 0);
    t.mod_freq.assign(NA, make_fastmod(0));
    t.mod_init.assign(NA, make_fastmod(0));
    t.mod_emis.assign(NA * NS, make_fastmod(0));
    t.mod_tran.assign(NA * NS, make_fastmod(0));
    memset(t.exist_hmm, 0, sizeof t.exist_hmm);
    memset(t.nstates, 0, sizeof t.nstates);
    long carried_end = t.accuracy_rand_value;  // end_wk is one variable in the reference and is not reset between loops
    for (long a = amin; a <= amax; a++) {
        t.exist_hmm[a] = (uint8_t)hmm[a].exist;
        if (hmm[a].exist == 1) {
            {
                SlotFiller f(100);
                f.end = carried_end;
                for (int j = 1; j <= PB_STATE_MAX; j++) {
                    if (hmm[a].ip[j] == 0) continue;
                    if (f.add(hmm[a].ip[j], [&](long k) { t.init2state[a * 101 + k] = (uint8_t)j; })) break;
                }
                carried_end = f.end;
                t.mod_init[a] = make_fastmod((uint32_t)carried_end);
            }
            for (int j = 1; j <= PB_STATE_MAX; j++) {
                SlotFiller f(100);
                f.end = carried_end;
                for (int k = 0; k < PB_NQ; k++) {
                    if (hmm[a].ep[j][k] == 0) continue;
                    if (f.add(hmm[a].ep[j][k], [&](long l) { t.emis2qc[(a * NS + j) * 101 + l] = (uint8_t)k; })) break;
                }
                carried_end = f.end;
                t.mod_emis[a * NS + j] = make_fastmod((uint32_t)carried_end);
            }
            for (int j = 1; j <= PB_STATE_MAX; j++) {
                SlotFiller f(100);
                f.end = carried_end;
                for (int k = 1; k <= PB_STATE_MAX; k++) {
                    if (hmm[a].tp[j][k] == 0) continue;
                    if (f.add(hmm[a].tp[j][k], [&](long l) { t.tran2state[(a * NS + j) * 101 + l] = (uint8_t)k; })) break;
                }
                carried_end = f.end;
                t.mod_tran[a * NS + j] = make_fastmod((uint32_t)carried_end);
            }
            // states the chain can visit: the closure of the initial states under tran2state (slot 0 is never drawn)
            bool seen[PB_STATE_MAX + 1] = {false};
            std::vector<int> todo;
            auto visit = [&](int s) { if (!seen[s]) { seen[s] = true; todo.push_back(s); } };
            if (t.mod_init[a].d == 0 || t.mod_init[a].d > 100) throw Error(SP_ERR_ARG, "hmm model: no initial state for accuracy " + std::to_string(a) + " (the reference divides by zero)");
            for (uint32_t k = 1; k <= t.mod_init[a].d; k++) visit(t.init2state[a * 101 + k]);
            int hi = 0;
            while (!todo.empty()) {
                const int s = todo.back(); todo.pop_back();
                hi = std::max(hi, s);
                const FastMod& mt = t.mod_tran[a * NS + s];
                const FastMod& me = t.mod_emis[a * NS + s];
                if (mt.d == 0 || me.d == 0 || mt.d > 100 || me.d > 100)
                    throw Error(SP_ERR_ARG, "hmm model: state " + std::to_string(s) + " of accuracy " + std::to_string(a) +
                                                " has no transition or emission slots (the reference divides by zero)");
                for (uint32_t k = 1; k <= mt.d; k++) visit(t.tran2state[(a * NS + s) * 101 + k]);
            }
            (void)hi;
            // The state index never leaves the chain (only quality codes are emitted), so the device tables use the
            // visited states renumbered 1..K in increasing order: the nibble maps of the kernel then cover exactly the
            // states that can occur.
            int remap[PB_STATE_MAX + 1] = {0}, K = 0;
            for (int s = 1; s <= PB_STATE_MAX; s++) if (seen[s]) remap[s] = ++K;
            if (seen[0]) throw Error(SP_ERR_ARG, "hmm model: state 0 is reachable for accuracy " + std::to_string(a));
            std::vector<uint8_t> tr(NS * 101, 0), em(NS * 101, 0);
            std::vector<FastMod> mt(NS, make_fastmod(0)), me(NS, make_fastmod(0));
            for (int s = 1; s <= PB_STATE_MAX; s++) {
                if (!seen[s]) continue;
                const int cs = remap[s];
                mt[cs] = t.mod_tran[a * NS + s]; me[cs] = t.mod_emis[a * NS + s];
                for (uint32_t k = 1; k <= mt[cs].d; k++) tr[cs * 101 + k] = (uint8_t)remap[t.tran2state[(a * NS + s) * 101 + k]];
                for (uint32_t k = 1; k <= me[cs].d; k++) em[cs * 101 + k] = t.emis2qc[(a * NS + s) * 101 + k];
            }
            for (size_t s = 0; s < NS; s++) {
                t.mod_tran[a * NS + s] = mt[s]; t.mod_emis[a * NS + s] = me[s];
                for (int k = 0; k < 101; k++) {
                    t.tran2state[(a * NS + s) * 101 + k] = tr[s * 101 + k];
                    t.emis2qc[(a * NS + s) * 101 + k] = em[s * 101 + k];
                }
            }
            for (uint32_t k = 1; k <= t.mod_init[a].d; k++) t.init2state[a * 101 + k] = (uint8_t)remap[t.init2state[a * 101 + k]];
            t.nstates[a] = (uint8_t)K;
        } else {
            SlotFiller f(1000);
            f.end = carried_end;
            const double* u = &uni_ep[(size_t)a * PB_NQ];
            for (int j = 0; j < PB_NQ; j++) {
                if (u[j] == 0) continue;
                if (f.add(u[j], [&](long k) { t.freq2qc[a * 1001 + k] = (uint8_t)j; })) break;
            }
            carried_end = f.end;
            if (carried_end < 1 || carried_end > 1000) throw Error(SP_ERR_ARG, "uniform quality table is empty (the reference divides by zero)");
            t.mod_freq[a] = make_fastmod((uint32_t)carried_end);
        }
    }
    // entries no read can reach still get a divisor, so that a speculative lookup stays inside its table row
    for (auto* v : {&t.mod_freq, &t.mod_init, &t.mod_emis, &t.mod_tran})
        for (auto& m : *v) if (m.d == 0) m = make_fastmod(1);
}

// set_sim_param + the option checks of main (docker/pbsim2.cpp:217-389, :1047-1107). Errors carry pbsim's messages.
// Host half only (parameters + tables, no device): what tests/emu links to run the kernel's per-read code on the CPU.
void reads_set_params(ReadState& R, const sp_read_args* a) {
    if (!a) throw Error(SP_ERR_ARG, "null sp_read_args");
    R.configured = false;
    ReadParams p;
    const int FASTQ_LEN_MAX = 1000000;
    auto range_err = [](const char* opt, const std::string& v, const std::string& range) {
        return Error(SP_ERR_ARG, std::string("ERROR (") + opt + ": " + v + "): Acceptable range is " + range + ".");
    };
    if (a->depth >= 0) { p.depth = a->depth; if (p.depth <= 0.0) throw Error(SP_ERR_ARG, "ERROR (depth: " + std::to_string(a->depth) + "): Acceptable range is more than 0."); }
    if (a->length_mean >= 0) { p.len_mean = a->length_mean; if (p.len_mean < 1 || p.len_mean > FASTQ_LEN_MAX) throw range_err("length-mean", std::to_string(a->length_mean), "1-1000000"); }
    if (a->length_sd >= 0) { p.len_sd = a->length_sd; if (p.len_sd > FASTQ_LEN_MAX) throw range_err("length-sd", std::to_string(a->length_sd), "0-1000000"); }
    if (a->length_min >= 0) { p.len_min = (long)a->length_min; if (p.len_min < 1 || p.len_min > FASTQ_LEN_MAX) throw range_err("length-min", std::to_string(a->length_min), "1-1000000"); }
    if (a->length_max >= 0) { p.len_max = (long)a->length_max; if (p.len_max < 1 || p.len_max > FASTQ_LEN_MAX) throw range_err("length-max", std::to_string(a->length_max), "1-1000000"); }
    if (a->accuracy_mean >= 0) {
        if (a->accuracy_mean > 1.0) throw range_err("accuracy-mean", std::to_string(a->accuracy_mean), "0.0-1.0");
        p.accuracy_mean = int(a->accuracy_mean * 100) * 0.01;  // :1078
    }
    if (a->accuracy_min > 1.0) throw range_err("accuracy-min", std::to_string(a->accuracy_min), "0.0-1.0");
    if (a->accuracy_max > 1.0) throw range_err("accuracy-max", std::to_string(a->accuracy_max), "0.0-1.0");
    if (a->sub_ratio >= 0) {
        const int RATIO_MAX = 1000;
        if (a->sub_ratio > RATIO_MAX || a->ins_ratio < 0 || a->ins_ratio > RATIO_MAX || a->del_ratio < 0 || a->del_ratio > RATIO_MAX)
            throw Error(SP_ERR_ARG, "ERROR (difference-ratio): Acceptable range is 0-1000.");
        p.sub_ratio = a->sub_ratio; p.ins_ratio = a->ins_ratio; p.del_ratio = a->del_ratio;
    }
    const long sum = p.sub_ratio + p.ins_ratio + p.del_ratio;
    if (sum == 0) throw Error(SP_ERR_ARG, "ERROR (difference-ratio): all ratios are zero.");
    p.sub_rate = (double)p.sub_ratio / sum; p.ins_rate = (double)p.ins_ratio / sum; p.del_rate = (double)p.del_ratio / sum;
    if (!a->hmm_model || !*a->hmm_model)
        throw Error(SP_ERR_ARG, "ERROR: either --sample-fastq(and/or --sample-profile-id)(sampling-based) or --hmm_model(model-based) must be set.");
    p.hmm_model_file = a->hmm_model;
    if (a->prefix) p.prefix = a->prefix;
    if (a->id_prefix) p.id_prefix = a->id_prefix;
    if (p.id_prefix.size() > 40) throw Error(SP_ERR_ARG, "id prefix longer than 40 characters");
    if (p.len_min > p.len_max)
        throw Error(SP_ERR_ARG, "ERROR: length min(" + std::to_string(p.len_min) + ") is greater than max(" + std::to_string(p.len_max) + ").");
    p.seed = a->seed;
    p.no_error = a->no_error != 0;
    if (a->stream_mode != SP_READS_PER_READ_STREAMS && a->stream_mode != SP_READS_REPLAY) throw Error(SP_ERR_ARG, "bad stream_mode");
    p.stream_mode = a->stream_mode;
    R.par = p;
    reads_build_tables(R);
}

void reads_configure(sp_ctx* c, const sp_read_args* a) {
    if (!a) throw Error(SP_ERR_ARG, "null sp_read_args");
    if (!c->reads) c->reads = new ReadState();
    ReadState& R = *c->reads;
    reads_set_params(R, a);

    // one device blob
    const ReadTablesHost& t = R.tab;
    auto al = [](size_t x) { return (x + 255) / 256 * 256; };
    size_t off = 0;
    auto place = [&](size_t bytes) { size_t o = off; off = al(off + bytes); return o; };
    const size_t o_thre = place(sizeof t.thre), o_ex = place(sizeof t.exist_hmm), o_ns = place(sizeof t.nstates),
                 o_freq = place(t.freq2qc.size()), o_init = place(t.init2state.size()), o_emis = place(t.emis2qc.size()),
                 o_tran = place(t.tran2state.size()), o_mf = place(t.mod_freq.size() * sizeof(FastMod)),
                 o_mi = place(t.mod_init.size() * sizeof(FastMod)), o_me = place(t.mod_emis.size() * sizeof(FastMod)),
                 o_mt = place(t.mod_tran.size() * sizeof(FastMod));
    std::vector<uint8_t> blob(off, 0);
    memcpy(&blob[o_thre], t.thre, sizeof t.thre);
    memcpy(&blob[o_ex], t.exist_hmm, sizeof t.exist_hmm);
    memcpy(&blob[o_ns], t.nstates, sizeof t.nstates);
    memcpy(&blob[o_freq], t.freq2qc.data(), t.freq2qc.size());
    memcpy(&blob[o_init], t.init2state.data(), t.init2state.size());
    memcpy(&blob[o_emis], t.emis2qc.data(), t.emis2qc.size());
    memcpy(&blob[o_tran], t.tran2state.data(), t.tran2state.size());
    memcpy(&blob[o_mf], t.mod_freq.data(), t.mod_freq.size() * sizeof(FastMod));
    memcpy(&blob[o_mi], t.mod_init.data(), t.mod_init.size() * sizeof(FastMod));
    memcpy(&blob[o_me], t.mod_emis.data(), t.mod_emis.size() * sizeof(FastMod));
    memcpy(&blob[o_mt], t.mod_tran.data(), t.mod_tran.size() * sizeof(FastMod));
    R.d_tab.reserve(blob.size(), 0.0);
    SP_CUDA(cudaMemcpyAsync(R.d_tab.p, blob.data(), blob.size(), cudaMemcpyHostToDevice, c->stream));
    SP_CUDA(cudaStreamSynchronize(c->stream));
    const uint8_t* b = R.d_tab.p;
    R.dev.thre = (const int32_t*)(b + o_thre);
    R.dev.exist_hmm = b + o_ex; R.dev.nstates = b + o_ns; R.dev.freq2qc = b + o_freq; R.dev.init2state = b + o_init;
    R.dev.emis2qc = b + o_emis; R.dev.tran2state = b + o_tran;
    R.dev.mod_freq = (const FastMod*)(b + o_mf); R.dev.mod_init = (const FastMod*)(b + o_mi);
    R.dev.mod_emis = (const FastMod*)(b + o_me); R.dev.mod_tran = (const FastMod*)(b + o_mt);
    R.have_replay = false;
    R.n_reads = 0;
    R.configured = true;
    R.plan_tables_stale = true;
    static uint64_t epoch = 0;
    R.config_epoch = ++epoch;
}

void free_reads(sp_ctx* c) {
    for (ReadState** pr : {&c->reads, &c->reads_b}) {
        ReadState* r = *pr;
        if (!r) continue;
        if (r->kernel_stream) { cudaStreamSynchronize(r->kernel_stream); cudaStreamDestroy(r->kernel_stream); }
        if (r->ev0) { cudaEventDestroy(r->ev0); cudaEventDestroy(r->ev1); }
        if (r->ev_ready) cudaEventDestroy(r->ev_ready);
        if (r->h_ctr) cudaFreeHost(r->h_ctr);
        delete r;
        *pr = nullptr;
    }
}

}  // namespace sp
