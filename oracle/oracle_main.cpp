// ORACLE — TEST INFRASTRUCTURE ONLY. Command-line front end of the CPU restatement:
//   synthpop_oracle [--seed S] [--quiet] [--dump-pre-collision FILE] [--draw-log FILE] [--cdf FILE]
//                   <SimulateHaplotypes|PrintHaplotypes|SampleReadsFromLengthBins|AnnotateVCF|BuildReadLengthBins|Fastq2LengthHistogram|BuildModel> <positional argv of the Java class>
// Exit codes follow the Java: 0 ok, 1 uncaught exception, 3 = SampleReadsFromLengthBins coverage unreachable.
// The seed replaces the reference's unseeded `new Random()` (src/SimulateHaplotypes.java:71) — the one
// documented deviation from the reference argv (SURVEY.md §8b "Seed").
#include "simulate_haplotypes.h"

namespace oracle {
int sampleReadsFromLengthBinsMain(const std::vector<std::string>& args, int64_t seed, const std::string& cdfFile);
int annotateVcfMain(const std::vector<std::string>& args);
int buildReadLengthBinsMain(const std::vector<std::string>& args, int64_t seed);
int fastq2LengthHistogramMain(const std::vector<std::string>& args);
int buildModelMain(const std::vector<std::string>& args);
}

int main(int argc, char** argv) {
    int64_t seed = 42;
    if (const char* e = getenv("SYNTHPOP_SEED")) seed = atoll(e);
    bool quiet = false;
    std::string dumpPre, drawLogFile, cdfFile;
    int a = 1;
    for (; a < argc; a++) {
        std::string s = argv[a];
        if (s == "--seed" && a + 1 < argc) seed = atoll(argv[++a]);
        else if (s == "--quiet") quiet = true;
        else if (s == "--dump-pre-collision" && a + 1 < argc) dumpPre = argv[++a];
        else if (s == "--draw-log" && a + 1 < argc) drawLogFile = argv[++a];
        else if (s == "--cdf" && a + 1 < argc) cdfFile = argv[++a];
        else break;
    }
    if (a >= argc) {
        fprintf(stderr, "usage: synthpop_oracle [--seed S] [--quiet] <SimulateHaplotypes|PrintHaplotypes|SampleReadsFromLengthBins> args...\n");
        return 2;
    }
    std::string cls = argv[a++];
    std::vector<std::string> args;
    for (; a < argc; a++) args.push_back(argv[a]);
    try {
        if (cls == "SimulateHaplotypes") {
            oracle::SimulateHaplotypes sh(seed);
            sh.verbose = !quiet;
            sh.dumpPreCollisionFile = dumpPre;
            int rc = sh.mainSimulate(args);
            if (!drawLogFile.empty()) {
                FILE* f = fopen(drawLogFile.c_str(), "w");
                if (f) {
                    fprintf(f, "candidates=%d\nsteps_after_sampleVariants=%llu\nsteps_after_buildPopulation=%llu\nsteps_after_assignInsertionStrings=%llu\nnVariants=%d\nnHaplotypes=%d\n",
                            sh.drawLog.candidates, (unsigned long long)sh.drawLog.steps_after_sampleVariants,
                            (unsigned long long)sh.drawLog.steps_after_buildPopulation,
                            (unsigned long long)sh.drawLog.steps_after_assignInsertionStrings, sh.lastVariant + 1, sh.N_HAPLOTYPES);
                    for (size_t k = 0; k < sh.phaseNames.size(); k++) fprintf(f, "seconds_%s=%.6f\n", sh.phaseNames[k].c_str(), sh.phaseSeconds[k]);
                    fclose(f);
                }
            }
            return rc;
        } else if (cls == "PrintHaplotypes") {
            oracle::SimulateHaplotypes sh(seed);
            sh.verbose = !quiet;
            return sh.mainPrint(args);
        } else if (cls == "SampleReadsFromLengthBins") {
            return oracle::sampleReadsFromLengthBinsMain(args, seed, cdfFile);
        } else if (cls == "AnnotateVCF") {
            return oracle::annotateVcfMain(args);
        } else if (cls == "BuildReadLengthBins") {
            return oracle::buildReadLengthBinsMain(args, seed);
        } else if (cls == "Fastq2LengthHistogram") {
            return oracle::fastq2LengthHistogramMain(args);
        } else if (cls == "BuildModel") {
            return oracle::buildModelMain(args);
        }
        fprintf(stderr, "unknown class %s\n", cls.c_str());
        return 2;
    } catch (const std::exception& e) {
        fprintf(stderr, "Exception in thread \"main\" %s\n", e.what());
        return 1;
    }
}
