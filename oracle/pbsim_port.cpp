// ORACLE — TEST INFRASTRUCTURE ONLY. See pbsim_port.h.
#include "pbsim_port.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>

namespace pbsim {

// glibc random_r.c, TYPE_3 (degree 31, separation 3): r[i] = r[i-31] + r[i-3] mod 2^32, result = r[i] >> 1,
// after seeding by the Lehmer generator 16807 (Schrage) and discarding 310 outputs.
GlibcRand::GlibcRand(unsigned seed) {
    if (seed == 0) seed = 1;
    int32_t w = (int32_t)seed;
    r[0] = (uint32_t)w;
    for (int i = 1; i < 31; i++) {
        const long hi = w / 127773, lo = w % 127773;
        long x = 16807 * lo - 2836 * hi;
        if (x < 0) x += 2147483647;
        w = (int32_t)x;
        r[i] = (uint32_t)w;
    }
    f = 3; b = 0;
    for (int i = 0; i < 310; i++) next();
    count = 0;
}
int GlibcRand::next() {
    r[f] += r[b];
    const int result = (int)(r[f] >> 1);
    if (++f >= 31) f = 0;
    if (++b >= 31) b = 0;
    count++;
    return result;
}

// docker/pbsim2.cpp:2583-2647
void set_hmm_model(const std::string& path, std::vector<HmmModel>& hmm) {
    hmm.assign(ACCURACY_MAX + 1, HmmModel());
    for (auto& m : hmm) memset(&m, 0, sizeof m);
    FILE* fp = fopen(path.c_str(), "r");
    if (!fp) throw std::runtime_error("ERROR: Cannot open file: " + path);
    std::vector<char> line(1 << 20);
    while (fgets(line.data(), 10240, fp) != nullptr) {
        size_t n = strlen(line.data());
        if (n && line[n - 1] == '\n') line[n - 1] = 0;
        char* tp = strtok(line.data(), " ");
        if (!tp) throw std::runtime_error("hmm model: empty line (the reference dereferences NULL here)");
        const int accuracy = atoi(tp);
        if (accuracy < 0 || accuracy > ACCURACY_MAX) throw std::runtime_error("hmm model: accuracy out of range");
        hmm[accuracy].exist_hmm = 1;
        tp = strtok(nullptr, " ");
        if (!tp) throw std::runtime_error("hmm model: truncated line");
        if (strcmp(tp, "IP") == 0) {
            tp = strtok(nullptr, " ");
            const int state = atoi(tp);
            tp = strtok(nullptr, " ");
            if (state < 0 || state > STATE_MAX || !tp) throw std::runtime_error("hmm model: bad IP line");
            hmm[accuracy].ip[state] = atof(tp);
        } else if (strcmp(tp, "EP") == 0) {
            tp = strtok(nullptr, " ");
            const int state = atoi(tp);
            if (state < 0 || state > STATE_MAX) throw std::runtime_error("hmm model: bad EP state");
            int num = 0;
            tp = strtok(nullptr, " ");
            while (tp != nullptr) {
                if (num > 93) throw std::runtime_error("hmm model: too many EP columns");
                hmm[accuracy].ep[state][num] = atof(tp);
                num++;
                tp = strtok(nullptr, " ");
            }
        } else if (strcmp(tp, "TP") == 0) {
            tp = strtok(nullptr, " ");
            const int state = atoi(tp);
            if (state < 0 || state > STATE_MAX) throw std::runtime_error("hmm model: bad TP state");
            int num = 0;
            tp = strtok(nullptr, " ");
            while (tp != nullptr) {
                num++;
                if (num > STATE_MAX) throw std::runtime_error("hmm model: too many TP columns");
                hmm[accuracy].tp[state][num] = atof(tp);
                tp = strtok(nullptr, " ");
            }
        }
    }
    fclose(fp);
}

// main :407-433 (qc, uni_ep) + simulate_by_model :1455-1637
void build_tables(const Params& p, const std::vector<HmmModel>& hmm, Tables& t) {
    double qc_prob[94];
    for (int i = 0; i <= 93; i++) qc_prob[i] = pow(10, (double)i / -10);
    static double uni_ep[ACCURACY_MAX + 1][94];
    for (int i = 0; i <= ACCURACY_MAX; i++) {
        for (int j = 0; j <= 93; j++) uni_ep[i][j] = 0;
        if (i == ACCURACY_MAX) { uni_ep[i][93] = 1.0; continue; }
        const double prob = 1.0 - i / 100.0;
        for (int j = 0; j <= 93; j++) {
            if (prob == qc_prob[j]) { uni_ep[i][j] = 1.0; break; }
            else if (prob > qc_prob[j]) {
                const double rate = (prob - qc_prob[j]) / (qc_prob[j - 1] - qc_prob[j]);
                uni_ep[i][j - 1] = rate;
                uni_ep[i][j] = 1 - rate;
                break;
            }
        }
    }
    const long sum = p.sub_ratio + p.ins_ratio + p.del_ratio;
    t.sub_rate = (double)p.sub_ratio / sum;
    t.ins_rate = (double)p.ins_ratio / sum;
    t.del_rate = (double)p.del_ratio / sum;
    for (int i = 0; i <= 93; i++) {
        t.err_thre[i] = int(qc_prob[i] * 1000000 + 0.5);
        t.sub_thre[i] = int((qc_prob[i] * t.sub_rate) * 1000000 + 0.5);
        t.ins_thre[i] = int((qc_prob[i] * (t.sub_rate + t.ins_rate)) * 1000000 + 0.5);
    }
    // length distribution
    t.prob2len.assign(100001, 0);
    t.prob2accuracy.assign(100001, 0);
    double variance = pow(p.len_sd, 2);
    const double kappa = pow(p.len_mean, 2) / variance;
    const double theta = variance / p.len_mean;
    const double gamma = tgamma(kappa);
    long start_wk, end_wk = 0;
    if (p.len_sd == 0.0) {
        t.prob2len[1] = int(p.len_mean + 0.5);
        t.len_rand_value = 1;
    } else {
        start_wk = 1;
        double len_prob_total = 0.0;
        for (long i = p.len_min; i <= p.len_max; i++) {
            len_prob_total += pow(i, kappa - 1) * exp(-1 * i / theta) / pow(theta, kappa) / gamma;
            end_wk = int(len_prob_total * 100000 + 0.5);
            if (end_wk > 100000) end_wk = 100000;
            for (long j = start_wk; j <= end_wk; j++) t.prob2len[j] = i;
            if (end_wk >= 100000) break;
            start_wk = end_wk + 1;
        }
        t.len_rand_value = end_wk;
    }
    if (t.len_rand_value < 1) throw std::runtime_error("ERROR: length parameters are not appropriate.");
    // accuracy distribution
    const double mean = p.accuracy_mean * 100;
    long accuracy_max = floor(mean * 1.05);
    long accuracy_min = floor(mean * 0.75);
    if (accuracy_max > 100) accuracy_max = 100;
    t.accuracy_min = accuracy_min; t.accuracy_max = accuracy_max;
    double freq_total = 0.0;
    for (long i = accuracy_min; i <= accuracy_max; i++) freq_total += exp(0.22 * i);
    start_wk = 1;
    end_wk = 0;
    double accuracy_prob_total = 0.0;
    for (long i = accuracy_min; i <= accuracy_max; i++) {
        accuracy_prob_total += exp(0.22 * i) / freq_total;
        end_wk = int(accuracy_prob_total * 100000 + 0.5);
        if (end_wk > 100000) end_wk = 100000;
        for (long j = start_wk; j <= end_wk; j++) t.prob2accuracy[j] = i;
        if (end_wk >= 100000) break;
        start_wk = end_wk + 1;
    }
    t.accuracy_rand_value = end_wk;
    if (t.accuracy_rand_value < 1) throw std::runtime_error("ERROR: accuracy parameters are not appropriate.");
    // quality code distribution
    t.freq2qc.assign((size_t)(ACCURACY_MAX + 1) * 1001, 0);
    t.init2state.assign((size_t)(ACCURACY_MAX + 1) * 1001, 0);
    t.emis2qc.assign((size_t)(ACCURACY_MAX + 1) * (STATE_MAX + 1) * 101, 0);
    t.tran2state.assign((size_t)(ACCURACY_MAX + 1) * (STATE_MAX + 1) * 101, 0);
    memset(t.qc_rand_value_freq, 0, sizeof t.qc_rand_value_freq);
    memset(t.qc_rand_value_init, 0, sizeof t.qc_rand_value_init);
    memset(t.qc_rand_value_emis, 0, sizeof t.qc_rand_value_emis);
    memset(t.qc_rand_value_tran, 0, sizeof t.qc_rand_value_tran);
    for (int i = 0; i <= ACCURACY_MAX; i++) t.exist_hmm[i] = hmm[i].exist_hmm;
    if (accuracy_min < 0) throw std::runtime_error("accuracy_min < 0");
    for (long i = accuracy_min; i <= accuracy_max; i++) {
        if (hmm[i].exist_hmm == 1) {
            start_wk = 1;
            double qc_prob_total = 0.0;  // end_wk is deliberately not reset (nor is it in the reference)
            for (int j = 1; j <= STATE_MAX; j++) {
                if (hmm[i].ip[j] == 0) continue;
                qc_prob_total += hmm[i].ip[j];
                end_wk = int(qc_prob_total * 100 + 0.5);
                if (end_wk > 100) end_wk = 100;
                for (long k = start_wk; k <= end_wk; k++) t.init2state[i * 1001 + k] = j;
                if (end_wk >= 100) break;
                start_wk = end_wk + 1;
            }
            t.qc_rand_value_init[i] = end_wk;
            for (int j = 1; j <= STATE_MAX; j++) {
                start_wk = 1;
                qc_prob_total = 0.0;
                // NB: end_wk is NOT reset here in the reference: a state with no emission keeps the previous value
                for (int k = 0; k <= 93; k++) {
                    if (hmm[i].ep[j][k] == 0) continue;
                    qc_prob_total += hmm[i].ep[j][k];
                    end_wk = int(qc_prob_total * 100 + 0.5);
                    if (end_wk > 100) end_wk = 100;
                    for (long l = start_wk; l <= end_wk; l++) t.emis2qc[((size_t)i * (STATE_MAX + 1) + j) * 101 + l] = k;
                    if (end_wk >= 100) break;
                    start_wk = end_wk + 1;
                }
                t.qc_rand_value_emis[i][j] = end_wk;
            }
            for (int j = 1; j <= STATE_MAX; j++) {
                start_wk = 1;
                qc_prob_total = 0.0;
                for (int k = 1; k <= STATE_MAX; k++) {
                    if (hmm[i].tp[j][k] == 0) continue;
                    qc_prob_total += hmm[i].tp[j][k];
                    end_wk = int(qc_prob_total * 100 + 0.5);
                    if (end_wk > 100) end_wk = 100;
                    for (long l = start_wk; l <= end_wk; l++) t.tran2state[((size_t)i * (STATE_MAX + 1) + j) * 101 + l] = k;
                    if (end_wk >= 100) break;
                    start_wk = end_wk + 1;
                }
                t.qc_rand_value_tran[i][j] = end_wk;
            }
        } else {
            start_wk = 1;
            double qc_prob_total = 0.0;
            for (int j = 0; j <= 93; j++) {
                if (uni_ep[i][j] == 0) continue;
                qc_prob_total += uni_ep[i][j];
                end_wk = int(qc_prob_total * 1000 + 0.5);
                if (end_wk > 1000) end_wk = 1000;
                for (long k = start_wk; k <= end_wk; k++) t.freq2qc[i * 1001 + k] = j;
                if (end_wk >= 1000) break;
                start_wk = end_wk + 1;
            }
            t.qc_rand_value_freq[i] = end_wk;
        }
    }
}


// mutate :2260-2413 — returns new_seq (or seq when no_error); seq_len_out = strlen(mut.new_seq)
namespace {
struct Mut {
    std::vector<char> qc, tmp_qc, seq, new_seq;
    std::vector<long> del_num;
};
}

static void mutate(const Params& p, const Tables& t, const std::string& ref, long res_num, long del_thre, long len_in,
                   Mut& m, GlibcRand& rng, long& out_len, long& seq_len) {
    long len = len_in;  // strlen(mut.qc) == tmp_len_max here
    m.del_num.assign(len, 0);
    const long del_num_total = int((float)len * del_thre / 1000000 + 0.5);
    long num = 0;
    while (num < del_num_total) {
        const long pos = rng.next() % (len - 1);
        const long rand_value = rng.next() % 1000000;
        const long qc_value = (int)m.qc[pos] - 33;
        if (rand_value < t.err_thre[qc_value]) { m.del_num[pos]++; num++; }
    }
    long offset = 0;
    m.tmp_qc.clear();
    for (long i = 0; i < len - 1; i++) {
        m.tmp_qc.push_back(m.qc[i]);
        for (long j = 0; j < m.del_num[i]; j++) m.tmp_qc.push_back(' ');
    }
    m.tmp_qc.push_back(m.qc[len - 1]);
    len = (long)m.tmp_qc.size();
    const long ref_len = (long)ref.size();
    if (len >= ref_len) { offset = 0; len = ref_len; }
    else offset = rng.next() % (ref_len - len + 1);
    m.seq.assign(len + 1, 0);
    char strand;
    if (res_num % 2 == 0) {
        strand = '+';
        for (long i = 0; i < len; i++) m.seq[i] = (char)toupper((unsigned char)ref[offset + i]);
    } else {
        strand = '-';
        for (long i = 0; i < len; i++) {
            const char nt = (char)toupper((unsigned char)ref[offset + i]);
            char c;
            if (nt == 'A') c = 'T'; else if (nt == 'T') c = 'A'; else if (nt == 'G') c = 'C'; else if (nt == 'C') c = 'G'; else c = nt;
            m.seq[len - 1 - i] = c;
        }
    }
    (void)strand;
    seq_len = len;
    static const char* sub_nt_a = "TGC"; static const char* sub_nt_t = "AGC"; static const char* sub_nt_g = "ATC";
    static const char* sub_nt_c = "ATG"; static const char* sub_nt_n = "ATGC"; static const char* ins_nt = "ATGC";
    m.new_seq.assign(len + 1, 0);
    offset = 0;
    long seq_offset = 0;
    for (long i = 0; i < len; i++) {
        const char nt = m.seq[seq_offset++];
        if (m.tmp_qc[i] == ' ') continue;
        const long rand_value = rng.next() % 1000000;
        const long qc_value = (int)m.tmp_qc[i] - 33;
        if (rand_value < t.sub_thre[qc_value]) {
            long index = rng.next() % 3;
            if (nt == 'A') m.new_seq[offset] = sub_nt_a[index];
            else if (nt == 'T') m.new_seq[offset] = sub_nt_t[index];
            else if (nt == 'G') m.new_seq[offset] = sub_nt_g[index];
            else if (nt == 'C') m.new_seq[offset] = sub_nt_c[index];
            else { index = rng.next() % 4; m.new_seq[offset] = sub_nt_n[index]; }
        } else if (rand_value < t.ins_thre[qc_value]) {
            const long index = rng.next() % 8;
            m.new_seq[offset] = index >= 4 ? nt : ins_nt[index];
            seq_offset--;
        } else m.new_seq[offset] = nt;
        offset++;
    }
    out_len = offset;
    (void)p;
}

// simulate_by_model :1641-1711 (the simulation loop)
void simulate_by_model(const Params& p, const Tables& t, const std::string& ref_seq, long ref_num, GlibcRand& rng,
                       std::string& out, ReadLog* log) {
    const long long len_quota = (long long)(p.depth * (long)ref_seq.size());
    long long len_total = 0;
    long res_num = 0;
    Mut m;
    char idbuf[256];
    while (len_total < len_quota) {
        if (p.per_read_streams) rng = GlibcRand(read_seed(p.seed, ref_num, res_num));
        const uint64_t draws_at_start = rng.count;
        if (log) log->draw_offset.push_back(rng.count);
        long index = rng.next() % t.len_rand_value + 1;
        long len = t.prob2len[index];
        if (len_total + len > len_quota) {
            len = len_quota - len_total;
            if (len < p.len_min) len = p.len_min;
        }
        index = rng.next() % t.accuracy_rand_value + 1;
        const int accuracy = (int)t.prob2accuracy[index];
        m.qc.assign(len, 0);
        long num = 0;
        if (t.exist_hmm[accuracy] == 1) {
            index = rng.next() % t.qc_rand_value_init[accuracy] + 1;
            long state = t.init2state[accuracy * 1001 + index];
            index = rng.next() % t.qc_rand_value_emis[accuracy][state] + 1;
            index = t.emis2qc[((size_t)accuracy * (STATE_MAX + 1) + state) * 101 + index];
            m.qc[num++] = (char)(index + 33);
            while (num < len) {
                index = rng.next() % t.qc_rand_value_tran[accuracy][state] + 1;
                state = t.tran2state[((size_t)accuracy * (STATE_MAX + 1) + state) * 101 + index];
                index = rng.next() % t.qc_rand_value_emis[accuracy][state] + 1;
                index = t.emis2qc[((size_t)accuracy * (STATE_MAX + 1) + state) * 101 + index];
                m.qc[num++] = (char)(index + 33);
            }
        } else {
            while (num < len) {
                index = rng.next() % t.qc_rand_value_freq[accuracy] + 1;
                index = t.freq2qc[accuracy * 1001 + index];
                m.qc[num++] = (char)(index + 33);
            }
        }
        const long del_thre = int(((100 - accuracy) * t.del_rate) * 10000 + 0.5);
        long out_len = 0, seq_len = 0;
        mutate(p, t, ref_seq, res_num, del_thre, len, m, rng, out_len, seq_len);
        len = out_len;
        len_total += len;
        res_num++;
        const char* s = p.no_error ? m.seq.data() : m.new_seq.data();
        const long slen = p.no_error ? seq_len : out_len;
        bool drop = false;
        for (long i = 0; i + 5 <= slen; i++)
            if (s[i] == 'N' && s[i + 1] == 'N' && s[i + 2] == 'N' && s[i + 3] == 'N' && s[i + 4] == 'N') { drop = true; break; }
        if (!drop) {
            snprintf(idbuf, sizeof idbuf, ">%s%ld_%ldZ", p.id_prefix.c_str(), ref_num, res_num);
            out += idbuf;
            out.append(s, (size_t)slen);
            out.push_back('\n');
        }
        if (log) { log->length.push_back(len); log->dropped.push_back(drop ? 1 : 0); log->draws_used.push_back(rng.count - draws_at_start); }
    }
    if (log) log->total_draws = rng.count;
}

// get_ref_inf / get_ref_seq :618-750: every '>' line starts a sequence; other lines are concatenated verbatim
// (fgets chunks of 10239 characters re-join to the same bytes).
void run(const Params& p, const std::string& fasta_path, std::vector<std::string>& outputs, std::vector<ReadLog>* logs) {
    std::vector<HmmModel> hmm;
    set_hmm_model(p.hmm_model_file, hmm);
    Tables t;
    build_tables(p, hmm, t);
    FILE* fp = fopen(fasta_path.c_str(), "rb");
    if (!fp) throw std::runtime_error("ERROR: Cannot open file: " + fasta_path);
    std::vector<std::string> seqs;
    std::string cur;
    bool have = false, in_header = false, at_line_start = true;
    int c;
    while ((c = getc_unlocked(fp)) != EOF) {
        if (at_line_start && c == '>') {
            if (have) seqs.push_back(cur);
            cur.clear(); have = true; in_header = true; at_line_start = false;
            continue;
        }
        if (c == '\n') { in_header = false; at_line_start = true; continue; }
        at_line_start = false;
        if (!in_header) {
            if (!have) have = true;  // sequence bytes before any header are counted by the reference as well
            cur.push_back((char)c);
        }
    }
    fclose(fp);
    if (have) seqs.push_back(cur);
    GlibcRand rng(p.seed);
    outputs.clear();
    if (logs) logs->clear();
    for (size_t k = 0; k < seqs.size(); k++) {
        if ((long)seqs[k].size() < 100) throw std::runtime_error("ERROR: Reference is too short. Acceptable length >= 100.");
        outputs.emplace_back();
        ReadLog lg;
        simulate_by_model(p, t, seqs[k], (long)k + 1, rng, outputs.back(), logs ? &lg : nullptr);
        if (logs) logs->push_back(std::move(lg));
    }
}

}  // namespace pbsim
