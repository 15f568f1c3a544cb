// C++ command-line host above the C ABI (this image has no JVM): same class names and positional argv as the
// reference's Java CLIs. Seed: optional `--seed S` before the class name, else $SYNTHPOP_SEED, else 42.
//   synthpop PrintHaplotypes FIRST LAST REFERENCE_FILE HAPLOTYPE2VARIANTS_FILE VARIANTS_FILE OUTPUT_DIR
//     (src/PrintHaplotypes.java:7-12)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <string>
#include <vector>
#include "synthpop_cuda.h"

static int fail(sp_ctx* ctx, int rc) {
    // Java: uncaught exception => stack trace on stderr, exit status 1
    fprintf(stderr, "Exception in thread \"main\" %s\n", sp_last_error(ctx));
    return rc == SP_ERR_COVERAGE ? 3 : 1;
}

int main(int argc, char** argv) {
    long long seed = 42;
    if (const char* e = getenv("SYNTHPOP_SEED")) seed = atoll(e);
    int device = 0;
    if (const char* e = getenv("SYNTHPOP_DEVICE")) device = atoi(e);
    int a = 1;
    for (; a < argc; a++) {
        if (!strcmp(argv[a], "--seed") && a + 1 < argc) seed = atoll(argv[++a]);
        else if (!strcmp(argv[a], "--device") && a + 1 < argc) device = atoi(argv[++a]);
        else break;
    }
    if (a >= argc) {
        fprintf(stderr, "usage: synthpop [--seed S] [--device D] <SimulateHaplotypes|PrintHaplotypes|SampleReadsFromLengthBins|AnnotateVCF|BuildReadLengthBins|Fastq2LengthHistogram|BuildModel|pbsim|haplotype2reads> <argv of the reference's command>\n");
        return 2;
    }
    const std::string cls = argv[a++];
    std::vector<const char*> args(argv + a, argv + argc);
    sp_ctx* ctx = nullptr;
    if (sp_create(&ctx, device, seed) != SP_OK) {
        fprintf(stderr, "synthpop: no usable CUDA device (there is no CPU fallback)\n");
        return 1;
    }
    int rc = SP_OK;
    if (cls == "PrintHaplotypes") {
        if (args.size() < 6) { fprintf(stderr, "Exception in thread \"main\" java.lang.ArrayIndexOutOfBoundsException\n"); return 1; }
        rc = sp_print_haplotypes(ctx, atoi(args[0]), atoi(args[1]), args[2], args[3], args[4], args[5]);
    } else if (cls == "SimulateHaplotypes") {
        // MODEL_DIR POPULATION_NAME REFERENCE_FILE REPEAT_INTERVALS_FILE SEGDUPS_FILE INSERTION_STRINGS_FILE
        // N_INSERTION_STRINGS ALLELE_FREQUENCY_THRESHOLD N_VARIANTS_FREQUENT N_INDIVIDUALS OUTPUT_DIR
        // (src/SimulateHaplotypes.java:76-86)
        if (args.size() < 11) { fprintf(stderr, "Exception in thread \"main\" java.lang.ArrayIndexOutOfBoundsException\n"); return 1; }
        sp_population_args pa;
        pa.model_dir = args[0]; pa.population_name = args[1]; pa.reference_file = args[2];
        pa.repeat_intervals_file = args[3]; pa.segdups_file = args[4]; pa.insertion_strings_file = args[5];
        pa.n_insertion_strings = atoi(args[6]); pa.allele_frequency_threshold = atof(args[7]);
        pa.n_variants_frequent = atoi(args[8]); pa.n_individuals = atoi(args[9]); pa.output_dir = args[10];
        rc = sp_simulate_population(ctx, &pa);
    } else if (cls == "SampleReadsFromLengthBins") {
        // MEAN_LEFT STD_LEFT MEAN_RIGHT STD_RIGHT WEIGHT_LEFT BIN_LENGTH MAX_READ_LENGTH BINS_PREFIX
        // GENOME_LENGTH_EACH_HAPLOTYPE TARGET_COVERAGE_EACH_HAPLOTYPE OUTPUT_FASTQ_FILE (src/SampleReadsFromLengthBins.java:20-30)
        if (args.size() < 11) { fprintf(stderr, "Exception in thread \"main\" java.lang.ArrayIndexOutOfBoundsException\n"); return 1; }
        const double ml = atof(args[0]), sl = atof(args[1]), mr = atof(args[2]), sr = atof(args[3]), wl = atof(args[4]);
        sp_length_bin_args la;
        la.bin_length = atoi(args[5]); la.max_read_length = atoi(args[6]); la.bins_prefix = args[7];
        la.genome_length_each_haplotype = atoll(args[8]); la.target_coverage_each_haplotype = atof(args[9]);
        la.output_fastq_file = args[10];
        if (la.bin_length <= 0) { fprintf(stderr, "Exception in thread \"main\" java.lang.ArithmeticException: / by zero\n"); return 1; }
        // buildCumulativeBimodal (:87-102). The Java host uses commons-math3's NormalDistribution.probability; this C++
        // host has only libm erf, which can differ from it in the last ulp (documented in INTEGRATION.md).
        const int nb = (la.max_read_length + la.bin_length - 1) / la.bin_length;
        std::vector<double> c(nb > 0 ? nb : 0);
        auto prob = [](double m, double sd, double x0, double x1) {
            const double d = sd * std::sqrt(2.0);
            return 0.5 * (std::erf((x1 - m) / d) - std::erf((x0 - m) / d));
        };
        double sum = 0.0;
        for (int i = 0; i < nb; i++) {
            c[i] = prob(ml, sl, (double)i * la.bin_length, (double)(i + 1) * la.bin_length) * wl +
                   prob(mr, sr, (double)i * la.bin_length, (double)(i + 1) * la.bin_length) * (1.0 - wl);
            sum += c[i];
        }
        for (int i = 0; i < nb; i++) c[i] /= sum;
        for (int i = 1; i < nb; i++) c[i] += c[i - 1];
        la.cumulative = c.data();
        rc = sp_sample_reads_from_bins(ctx, &la);
    } else if (cls == "pbsim") {
        // the long options of docker/pbsim2.cpp:195-213 that model mode uses, then the FASTA (docker/haplotype2reads.sh:53)
        sp_read_args ra;
        memset(&ra, 0, sizeof ra);
        ra.depth = ra.length_mean = ra.length_sd = ra.accuracy_mean = ra.accuracy_min = ra.accuracy_max = -1.0;
        ra.length_min = ra.length_max = -1; ra.sub_ratio = ra.ins_ratio = ra.del_ratio = -1;
        ra.seed = (uint32_t)seed;
        const char* fasta = nullptr;
        for (size_t i = 0; i < args.size(); i++) {
            const std::string o = args[i];
            if (o.rfind("--", 0) != 0) { fasta = args[i]; continue; }
            if (i + 1 >= args.size()) { fprintf(stderr, "ERROR: option %s needs a value\n", o.c_str()); return 255; }
            const char* v = args[++i];
            if (o == "--depth") ra.depth = atof(v);
            else if (o == "--length-mean") ra.length_mean = atof(v);
            else if (o == "--length-sd") ra.length_sd = atof(v);
            else if (o == "--length-min") ra.length_min = atoll(v);
            else if (o == "--length-max") ra.length_max = atoll(v);
            else if (o == "--accuracy-mean") ra.accuracy_mean = atof(v);
            else if (o == "--accuracy-min") ra.accuracy_min = atof(v);
            else if (o == "--accuracy-max") ra.accuracy_max = atof(v);
            else if (o == "--difference-ratio") { if (sscanf(v, "%d:%d:%d", &ra.sub_ratio, &ra.ins_ratio, &ra.del_ratio) != 3) { fprintf(stderr, "ERROR: (difference-ratio: %s) is not valid.\n", v); return 255; } }
            else if (o == "--hmm_model") ra.hmm_model = v;
            else if (o == "--prefix") ra.prefix = v;
            else if (o == "--id-prefix") ra.id_prefix = v;
            else if (o == "--seed") ra.seed = (uint32_t)strtoul(v, nullptr, 10);
            else if (o == "--no-error") ra.no_error = atoi(v);  // docker/pbsim2noError.cpp
            else { fprintf(stderr, "ERROR: option %s is not part of model mode\n", o.c_str()); return 255; }
        }
        if (!fasta) { fprintf(stderr, "ERROR: no reference FASTA\n"); return 255; }
        rc = sp_simulate_reads(ctx, &ra, fasta);
        if (rc != SP_OK) {  // pbsim: message on stderr + exit(-1) (docker/pbsim2.cpp:233-236)
            fprintf(stderr, "%s\n", sp_last_error(ctx));
            sp_destroy(ctx);
            return 255;
        }
    } else if (cls == "AnnotateVCF") {
        // INPUT_FILES_LIST MIN_DISTANCE MAX_DISTANCE REFERENCE_FILE REPEAT_INTERVALS_FILE SEGDUPS_FILE (src/AnnotateVCF.java:96-102)
        if (args.size() < 6) { fprintf(stderr, "Exception in thread \"main\" java.lang.ArrayIndexOutOfBoundsException\n"); return 1; }
        sp_annotate_args aa;
        aa.input_files_list = args[0]; aa.min_distance = atoi(args[1]); aa.max_distance = atoi(args[2]);
        aa.reference_file = args[3]; aa.repeat_intervals_file = args[4]; aa.segdups_file = args[5];
        rc = sp_annotate_vcf(ctx, &aa);
    } else if (cls == "BuildReadLengthBins") {
        // INPUT_FILES_LIST BIN_LENGTH MAX_READ_LENGTH GENOME_LENGTH_HAPLOID OUTPUT_PREFIX (src/BuildReadLengthBins.java:21-27)
        if (args.size() < 5) { fprintf(stderr, "Exception in thread \"main\" java.lang.ArrayIndexOutOfBoundsException\n"); return 1; }
        sp_build_bins_args ba;
        ba.input_files_list = args[0]; ba.bin_length = atoi(args[1]); ba.max_read_length = atoi(args[2]);
        ba.genome_length_haploid = atoll(args[3]); ba.output_prefix = args[4];
        rc = sp_build_read_length_bins(ctx, &ba);
    } else if (cls == "Fastq2LengthHistogram") {
        // FASTQ_FILE BIN_LENGTH MAX_READ_LENGTH OUTPUT_FILE_HISTOGRAM OUTPUT_FILE_MAX (src/Fastq2LengthHistogram.java:11-16)
        if (args.size() < 5) { fprintf(stderr, "Exception in thread \"main\" java.lang.ArrayIndexOutOfBoundsException\n"); return 1; }
        rc = sp_fastq2length_histogram(ctx, args[0], atoi(args[1]), atoi(args[2]), args[3], args[4]);
    } else if (cls == "BuildModel") {
        // REPEATMASKER_FILE NROWS TRF_FILE NROWS SEGDUPS_FILE NROWS REFERENCE_FAI REFERENCE_CYTOBANDS GNOMADSV_FILE ONLY_PASS
        // OUTPUT_DIR VERBOSE (src/BuildModel.java:22-33)
        if (args.size() < 12) { fprintf(stderr, "Exception in thread \"main\" java.lang.ArrayIndexOutOfBoundsException\n"); return 1; }
        sp_build_model_args ma;
        ma.repeatmasker_file = args[0]; ma.repeatmasker_rows = atoi(args[1]); ma.trf_file = args[2]; ma.trf_rows = atoi(args[3]);
        ma.segdups_file = args[4]; ma.segdups_rows = atoi(args[5]); ma.reference_fai = args[6]; ma.reference_cytobands = args[7];
        ma.gnomad_vcf = args[8]; ma.only_pass = atoi(args[9]) == 1; ma.output_dir = args[10]; ma.verbose = atoi(args[11]) == 1;
        rc = sp_build_model(ctx, &ma);
    } else if (cls == "haplotype2reads") {
        // docker/haplotype2reads.sh:16-24,31-32,53-64 for one individual, without the haplotype_<i>.fa files and the bucket:
        // ID1 LENGTH_MIN LENGTH_MAX LENGTH_MEAN LENGTH_STDEV MAX_COVERAGE REFERENCE_FILE HAPLOTYPE2VARIANTS_FILE VARIANTS_FILE
        // HMM_MODEL OUTPUT_FILE   (ID2 = ID1 + 1; the pbsim options are the script's PBSIM2_COMMAND)
        if (args.size() < 11) { fprintf(stderr, "usage: haplotype2reads ID1 LENGTH_MIN LENGTH_MAX LENGTH_MEAN LENGTH_STDEV MAX_COVERAGE REFERENCE_FILE HAPLOTYPE2VARIANTS_FILE VARIANTS_FILE HMM_MODEL OUTPUT_FILE\n"); return 2; }
        sp_read_args ra;
        memset(&ra, 0, sizeof ra);
        ra.length_min = atoll(args[1]); ra.length_max = atoll(args[2]); ra.length_mean = atof(args[3]); ra.length_sd = atof(args[4]);
        ra.depth = atof(args[5]);
        ra.accuracy_min = 0.99; ra.accuracy_max = 1.0; ra.accuracy_mean = 0.995;
        ra.sub_ratio = 6; ra.ins_ratio = 50; ra.del_ratio = 54;
        ra.hmm_model = args[9];
        ra.seed = (uint32_t)seed;
        rc = sp_load_reference_fasta(ctx, args[6]);
        if (rc == SP_OK) rc = sp_load_population(ctx, args[7], args[8]);
        if (rc == SP_OK) rc = sp_reads_configure(ctx, &ra);
        int64_t n_reads = 0, n_bytes = 0;
        if (rc == SP_OK) rc = sp_haplotype2reads(ctx, atoi(args[0]), seed, args[10], &n_reads, &n_bytes);
        if (rc == SP_OK) fprintf(stderr, "%lld reads, %lld bytes\n", (long long)n_reads, (long long)n_bytes);
    } else {
        fprintf(stderr, "synthpop: unknown class %s\n", cls.c_str());
        sp_destroy(ctx);
        return 2;
    }
    int code = rc == SP_OK ? 0 : fail(ctx, rc);
    sp_destroy(ctx);
    return code;
}
