// C++ command-line host above the C ABI (this image has no JVM): same class names and positional argv as the
// reference's Java CLIs. Seed: optional `--seed S` before the class name, else $SYNTHPOP_SEED, else 42.
//   synthpop PrintHaplotypes FIRST LAST REFERENCE_FILE HAPLOTYPE2VARIANTS_FILE VARIANTS_FILE OUTPUT_DIR
//     (src/PrintHaplotypes.java:7-12)
#include <cstdio>
#include <cstdlib>
#include <cstring>
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
        fprintf(stderr, "usage: synthpop [--seed S] [--device D] <SimulateHaplotypes|PrintHaplotypes> <positional argv of the Java class>\n");
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
    } else {
        fprintf(stderr, "synthpop: unknown class %s\n", cls.c_str());
        sp_destroy(ctx);
        return 2;
    }
    int code = rc == SP_OK ? 0 : fail(ctx, rc);
    sp_destroy(ctx);
    return code;
}
