// ORACLE — TEST INFRASTRUCTURE ONLY. Command line of the pbsim2 restatement, same long options as
// docker/pbsim2.cpp:195-213 for model mode, plus --no-error (pbsim2noError.cpp) and --log FILE.
#include "pbsim_port.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>

int main(int argc, char** argv) {
    pbsim::Params p;
    std::string prefix = "sd", logfile, fasta;
    bool seed_set = false;
    for (int a = 1; a < argc; a++) {
        std::string s = argv[a];
        auto val = [&]() -> const char* { if (a + 1 >= argc) { fprintf(stderr, "missing value for %s\n", s.c_str()); exit(2); } return argv[++a]; };
        if (s == "--depth") p.depth = atof(val());
        else if (s == "--length-mean") p.len_mean = atof(val());
        else if (s == "--length-sd") p.len_sd = atof(val());
        else if (s == "--length-min") p.len_min = atoi(val());
        else if (s == "--length-max") p.len_max = atoi(val());
        else if (s == "--accuracy-mean") p.accuracy_mean = int(atof(val()) * 100) * 0.01;
        else if (s == "--accuracy-min" || s == "--accuracy-max") val();  // parsed and unused in model mode
        else if (s == "--difference-ratio") {
            long r[3] = {6, 54, 40};
            sscanf(val(), "%ld:%ld:%ld", &r[0], &r[1], &r[2]);
            p.sub_ratio = r[0]; p.ins_ratio = r[1]; p.del_ratio = r[2];
        } else if (s == "--hmm_model") p.hmm_model_file = val();
        else if (s == "--prefix") prefix = val();
        else if (s == "--seed") { p.seed = (unsigned)atoi(val()); seed_set = true; }
        else if (s == "--id-prefix") p.id_prefix = val();
        else if (s == "--no-error") p.no_error = true;
        else if (s == "--per-read-streams") p.per_read_streams = true;
        else if (s == "--log") logfile = val();
        else fasta = s;
    }
    if (!seed_set || fasta.empty() || p.hmm_model_file.empty()) {
        fprintf(stderr, "usage: pbsim_port --seed S --hmm_model FILE [pbsim2 model-mode options] [--no-error] [--log FILE] ref.fa\n");
        return 2;
    }
    try {
        std::vector<std::string> outs;
        std::vector<pbsim::ReadLog> logs;
        pbsim::run(p, fasta, outs, &logs);
        for (size_t k = 0; k < outs.size(); k++) {
            char name[4096];
            snprintf(name, sizeof name, "%s_%04zu.fastq", prefix.c_str(), k + 1);
            FILE* f = fopen(name, "wb");
            if (!f) { fprintf(stderr, "ERROR: Cannot open output file: %s\n", name); return 255; }
            fwrite(outs[k].data(), 1, outs[k].size(), f);
            fclose(f);
        }
        if (!logfile.empty()) {
            FILE* f = fopen(logfile.c_str(), "w");
            for (size_t k = 0; k < logs.size(); k++)
                for (size_t r = 0; r < logs[k].draw_offset.size(); r++)
                    fprintf(f, "%zu\t%zu\t%llu\t%lld\t%d\t%llu\n", k + 1, r + 1, (unsigned long long)logs[k].draw_offset[r],
                            (long long)logs[k].length[r], (int)logs[k].dropped[r], (unsigned long long)logs[k].draws_used[r]);
            fclose(f);
        }
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 255;
    }
    return 0;
}
