// TEST INFRASTRUCTURE, not product code: runs the read kernel's per-read functions (csrc/reads_core.cuh, the same source the
// device kernel compiles) on the CPU, one lane at a time, so that their logic — ring rotation, one draw per step, the
// record's output words — can be compared with the pbsim2 restatement without a GPU (tests/test_reads_emu.py). What only
// a GPU can show (warp votes, shared-memory layout, the launch) stays with tests/test_reads_gpu.py. Built by
// tests/emu/Makefile into libreads_emu.so; nothing in population-sv-simulator_b200/ links or loads it.
#include "../../population-sv-simulator_b200/csrc/reads_core.cuh"
#include <cstring>
#include <string>
#include <vector>

using namespace sp;

namespace {

template <bool REPLAY, bool NOERR>
void run_reads(const ReadJob& J, const ReadTablesDev& T, const int32_t* tables, int64_t words) {
    std::vector<uint32_t> state((size_t)RING * RD_THREADS, 0u);
    std::vector<uint2> scratch((size_t)words * 32);
    CtaTables C;
    C.bind(tables);
    Draws<REPLAY> S;
    S.st = state.data() + 5;   // any column
    S.replay = nullptr; S.used = 0; S.pos = 31;
    for (int64_t r = 0; r < J.n_reads; r++) {
        for (auto& w : scratch) w = make_uint2(0xdeadbeefu, 0xdeadbeefu);   // nothing may rely on a clean scratch
        simulate_read_sync<REPLAY, NOERR>(J, T, C, S, scratch.data() + 7, (int32_t)r, 1u);
    }
}

}  // namespace

extern "C" int emu_reads(const sp_read_args* args, const uint8_t* seq, int64_t L, int64_t ref_num, const int32_t* replay_draws,
                         int64_t n_draws, const int64_t* replay_off, int64_t n_replay, uint8_t* out, int64_t cap,
                         int64_t* out_bytes, int64_t* n_reads_out, int32_t* len_out, uint8_t* dropped_out, uint32_t* used_out,
                         int64_t log_cap, char* err, int errcap) {
    try {
        ReadState R;
        reads_set_params(R, args);
        const ReadParams& p = R.par;
        const ReadTablesHost& t = R.tab;
        ReadTablesDev T{};
        T.thre = t.thre; T.exist_hmm = t.exist_hmm; T.nstates = t.nstates; T.freq2qc = t.freq2qc.data();
        T.init2state = t.init2state.data(); T.emis2qc = t.emis2qc.data(); T.tran2state = t.tran2state.data();
        T.mod_freq = t.mod_freq.data(); T.mod_init = t.mod_init.data(); T.mod_emis = t.mod_emis.data(); T.mod_tran = t.mod_tran.data();
        const bool replay = p.stream_mode == SP_READS_REPLAY;
        if (L < 100) throw Error(SP_ERR_ARG, "ERROR: Reference is too short. Acceptable length >= 100.");

        // the loop header of simulate_by_model (:1641-1655) and the per-read plan (:1682, :2282), as csrc/reads.cu makes it
        const long long quota = (long long)(p.depth * (long)L);
        std::vector<int32_t> len, ndel;
        std::vector<uint8_t> acc;
        long long len_total = 0;
        for (int64_t r = 0; len_total < quota; r++) {
            long d0, d1;
            if (replay) {
                if (r >= n_replay) throw Error(SP_ERR_ARG, "replay: the log has fewer reads than the quota needs");
                const int64_t o = replay_off[r];
                if (o < 0 || o + 1 >= n_draws) throw Error(SP_ERR_ARG, "replay: read offset outside the draw log");
                d0 = replay_draws[o]; d1 = replay_draws[o + 1];
            } else {
                uint32_t g[31];
                glibc_seed(read_seed(p.seed, ref_num, r), g);
                g[3] += g[0];
                d0 = (long)(g[3] >> 1); d1 = (long)((g[4] + g[1]) >> 1);
            }
            long l = t.prob2len[d0 % t.len_rand_value + 1];
            if (len_total + l > quota) {
                l = quota - len_total;
                if (l < p.len_min) l = p.len_min;
            }
            const int a = t.prob2accuracy[d1 % t.accuracy_rand_value + 1];
            const long del_thre = int(((100 - a) * p.del_rate) * 10000 + 0.5);
            const long nd = int((float)l * del_thre / 1000000 + 0.5);
            len.push_back((int32_t)l); acc.push_back((uint8_t)a); ndel.push_back((int32_t)nd);
            len_total += l;
        }
        const int64_t n = (int64_t)len.size();
        if (replay && n != n_replay) throw Error(SP_ERR_ARG, "replay: the log has more reads than the quota needs");
        const std::string id = ">" + p.id_prefix + std::to_string(ref_num) + "_";
        std::vector<int64_t> rec_off((size_t)n + 1, 0);
        int32_t max_len = 2;
        for (int64_t i = 0; i < n; i++) {
            const int64_t seqlen = p.no_error ? (int64_t)len[i] + ndel[i] : len[i];
            rec_off[i + 1] = rec_off[i] + (int64_t)id.size() + (int64_t)std::to_string(i + 1).size() + 1 + seqlen + 1;
            max_len = std::max(max_len, len[i]);
        }
        std::vector<uint8_t> rec((size_t)rec_off[n] + 64, 0xEE);   // 0xEE: a byte the kernel must overwrite
        std::vector<uint8_t> padded((size_t)L + 32, 0);
        memcpy(padded.data() + 16, seq, (size_t)L);
        std::vector<int32_t> order((size_t)n);
        for (int64_t i = 0; i < n; i++) order[(size_t)i] = (int32_t)i;
        std::vector<uint8_t> dropped((size_t)n + 1, 0);
        std::vector<uint32_t> used((size_t)n + 1, 0);
        unsigned long long ctr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        int32_t tables[RD_TABLE_WORDS];
        fill_cta_tables(tables, t.thre, 0, 1, 0);
        fill_cta_tables(tables, t.thre, 0, 1, 1);

        ReadJob J{};
        J.seq = padded.data() + 16; J.L = L; J.n_reads = n; J.order = order.data();
        J.len = len.data(); J.acc = acc.data(); J.ndel = ndel.data(); J.rec_off = rec_off.data();
        J.seed = p.seed; J.ref_num = ref_num; J.no_error = p.no_error ? 1 : 0;
        J.replay = replay ? replay_draws : nullptr; J.replay_off = replay ? replay_off : nullptr;
        J.out = rec.data(); J.dropped = dropped.data(); J.used = used.data();
        J.scratch = nullptr; J.scratch_words = 0; J.ctr = ctr;
        memcpy(J.id, id.data(), id.size()); J.id_len = (int32_t)id.size();
        const int64_t words = ((int64_t)max_len + 3) / 4 + 1;
        if (replay) { if (p.no_error) run_reads<true, true>(J, T, tables, words); else run_reads<true, false>(J, T, tables, words); }
        else { if (p.no_error) run_reads<false, true>(J, T, tables, words); else run_reads<false, false>(J, T, tables, words); }
        if (ctr[1] != 0) throw Error(SP_ERR_ARG, "read kernel error code " + std::to_string(ctr[1]) + ", read " + std::to_string(ctr[2] + 1));

        int64_t w = 0;
        for (int64_t i = 0; i < n; i++) {
            if (dropped[(size_t)i]) continue;
            const int64_t b = rec_off[i + 1] - rec_off[i];
            if (w + b > cap) throw Error(SP_ERR_ARG, "output buffer too small");
            memcpy(out + w, rec.data() + rec_off[i], (size_t)b);
            w += b;
        }
        if (n > log_cap) throw Error(SP_ERR_ARG, "log buffers too small");
        for (int64_t i = 0; i < n; i++) { len_out[i] = len[(size_t)i]; dropped_out[i] = dropped[(size_t)i]; used_out[i] = used[(size_t)i]; }
        *out_bytes = w; *n_reads_out = n;
        return 0;
    } catch (const std::exception& e) {
        if (err && errcap > 0) { strncpy(err, e.what(), (size_t)errcap - 1); err[errcap - 1] = 0; }
        return 1;
    }
}
