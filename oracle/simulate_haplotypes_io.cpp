// ORACLE — TEST INFRASTRUCTURE ONLY. See simulate_haplotypes.h.
// Truth VCFs, histograms, (de)serialisation and the two main() entry points.
#include "simulate_haplotypes.h"
#include <chrono>

namespace oracle {

static const char* VCF_SEPARATOR = "\t";
static const char* CHROMOSOME = "chr1";
static const char* ID_PREFIX = "simulated_";
static const char* QUALITY = "999";
static const char* FILTER = "PASS";
static const char* INFO_SEPARATOR = ";";
static const char* GENOTYPE_STR = "GT";

// Variant.writeVCF, src/SimulateHaplotypes.java:1243-1279
void SimulateHaplotypes::writeVCF(const Variant& v, int id, FILE* bw) {
    auto charAt = [&](int p) -> char {
        if (p < 0 || p >= referenceLength) throw std::runtime_error("StringIndexOutOfBoundsException: index " + std::to_string(p));
        return reference[p];
    };
    fputs(CHROMOSOME, bw); fputs(VCF_SEPARATOR, bw);
    if (v.svType == 0) {
        fprintf(bw, "%d%s", v.position, VCF_SEPARATOR);
        fprintf(bw, "%s%d%s", ID_PREFIX, id, VCF_SEPARATOR);
        int b = v.position - 1, e = v.position + v.length;
        if (b < 0 || e > referenceLength || b > e) throw std::runtime_error("StringIndexOutOfBoundsException: begin " + std::to_string(b) + ", end " + std::to_string(e));
        fwrite(reference.data() + b, 1, (size_t)(e - b), bw); fputs(VCF_SEPARATOR, bw);
        fputc(charAt(v.position - 1), bw); fputs(VCF_SEPARATOR, bw);
        fputs(QUALITY, bw); fputs(VCF_SEPARATOR, bw);
        fputs(FILTER, bw); fputs(VCF_SEPARATOR, bw);
        fprintf(bw, "END=%d%s", v.position + v.length, INFO_SEPARATOR);
        fprintf(bw, "SVLEN=%d%s", -v.length, INFO_SEPARATOR);
    } else if (v.svType == 2) {
        fprintf(bw, "%d%s", v.position + 1, VCF_SEPARATOR);
        fprintf(bw, "%s%d%s", ID_PREFIX, id, VCF_SEPARATOR);
        fputc(charAt(v.position), bw); fputs(VCF_SEPARATOR, bw);
        fputc(charAt(v.position), bw);
        if (!v.hasSequence) throw std::runtime_error("NullPointerException: sequence");
        fwrite(v.sequence.data(), 1, v.sequence.size(), bw); fputs(VCF_SEPARATOR, bw);
        fputs(QUALITY, bw); fputs(VCF_SEPARATOR, bw);
        fputs(FILTER, bw); fputs(VCF_SEPARATOR, bw);
        fprintf(bw, "END=%d%s", v.position + 1, INFO_SEPARATOR);
        fprintf(bw, "SVLEN=%d%s", v.length, INFO_SEPARATOR);
    } else {
        if (v.svType < 0 || v.svType > 3) throw std::runtime_error("ArrayIndexOutOfBoundsException: SV_TYPES");
        fprintf(bw, "%d%s", v.position + 1, VCF_SEPARATOR);
        fprintf(bw, "%s%d%s", ID_PREFIX, id, VCF_SEPARATOR);
        fputc(charAt(v.position), bw); fputs(VCF_SEPARATOR, bw);
        fprintf(bw, "<%s>%s", BuildModel::SV_TYPES[v.svType], VCF_SEPARATOR);
        fputs(QUALITY, bw); fputs(VCF_SEPARATOR, bw);
        fputs(FILTER, bw); fputs(VCF_SEPARATOR, bw);
        fprintf(bw, "END=%d%s", v.position + v.length, INFO_SEPARATOR);
    }
    fprintf(bw, "SVTYPE=%s%s", BuildModel::SV_TYPES[v.svType], INFO_SEPARATOR);
    fprintf(bw, "AF=%s%s", jc::doubleToString(v.alleleFrequency).c_str(), INFO_SEPARATOR);
    fprintf(bw, "REPEATS_START=%d%s", v.intervalType_start, INFO_SEPARATOR);
    fprintf(bw, "REPEATS_END=%d%s", v.intervalType_end, INFO_SEPARATOR);
    fprintf(bw, "REPEATS_FRACTION=%s", jc::numberFormat4(((double)v.intervalSurface) / v.length).c_str());
}

// src/SimulateHaplotypes.java:1077-1096
void SimulateHaplotypes::printVcfHeader(bool printGenotype, const std::string& sampleName, FILE* bw) {
    fputs("##fileformat=VCFv4.2\n", bw);
    fprintf(bw, "##contig=<ID=%s,length=%d>\n", CHROMOSOME, referenceLength);
    fputs("##ALT=<ID=INS,Description=\"Insertion\">\n", bw);
    fputs("##ALT=<ID=DEL,Description=\"Deletion\">\n", bw);
    fputs("##ALT=<ID=DUP,Description=\"Duplication\">\n", bw);
    fputs("##ALT=<ID=INV,Description=\"Inversion\">\n", bw);
    fputs("##FILTER=<ID=PASS,Description=\"All filters passed\">\n", bw);
    fputs("##INFO=<ID=END,Number=1,Type=Integer,Description=\"End position of structural variation\">\n", bw);
    fputs("##INFO=<ID=SVTYPE,Number=1,Type=String,Description=\"Type of structural variant\">\n", bw);
    fputs("##INFO=<ID=SVLEN,Number=1,Type=Integer,Description=\"Length of structural variation\">\n", bw);
    fputs("##INFO=<ID=AF,Number=1,Type=Float,Description=\"Allele Frequency\">\n", bw);
    fputs("##INFO=<ID=REPEATS_START,Number=1,Type=Integer,Description=\"Repetitive context around the first position\">\n", bw);
    fputs("##INFO=<ID=REPEATS_END,Number=1,Type=Integer,Description=\"Repetitive context around the last position\">\n", bw);
    fputs("##INFO=<ID=REPEATS_FRACTION,Number=1,Type=Float,Description=\"Fraction of the SV covered by repeats of any type\">\n", bw);
    fputs("##FORMAT=<ID=GT,Number=1,Type=String,Description=\"Genotype\">\n", bw);
    fputs("#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO", bw);
    if (printGenotype) fprintf(bw, "\tFORMAT\t%s", sampleName.c_str());
    fputc('\n', bw);
}

// src/SimulateHaplotypes.java:1005-1067
void SimulateHaplotypes::buildVCF(const std::string& outDir) {
    FILE* bw = fopen((outDir + "/groundTruth_joint.vcf").c_str(), "wb");
    if (!bw) throw std::runtime_error("IOException: groundTruth_joint.vcf");
    printVcfHeader(false, "", bw);
    for (int i = 0; i <= lastVariant; i++) {
        if (!variants[i].isActive) continue;
        writeVCF(variants[i], i, bw);
        fputs(";HAPLOTYPES=", bw);
        for (int j = 0; j < variant2haplotypes_last[i]; j++) fprintf(bw, "%d,", variant2haplotypes[i][j]);
        fprintf(bw, "%d", variant2haplotypes[i][variant2haplotypes_last[i]]);
        fputc('\n', bw);
    }
    fclose(bw);
    for (int i = 0; i < N_HAPLOTYPES; i += 2) {
        bw = fopen((outDir + "/groundTruth_individual_" + std::to_string(i) + ".vcf").c_str(), "wb");
        if (!bw) throw std::runtime_error("IOException: groundTruth_individual");
        printVcfHeader(true, "individual_" + std::to_string(i), bw);
        int last1 = haplotype2variants_last[i], last2 = haplotype2variants_last[i + 1];
        int j = 0, k = 0, variant;
        auto gt = [&](const char* g) { fprintf(bw, "%s%s%s%s\n", VCF_SEPARATOR, GENOTYPE_STR, VCF_SEPARATOR, g); };
        while (j <= last1 && k <= last2) {
            if (haplotype2variants[i][j] < haplotype2variants[i + 1][k]) {
                variant = haplotype2variants[i][j]; writeVCF(variants[variant], variant, bw); gt("1|0"); j++;
            } else if (haplotype2variants[i + 1][k] < haplotype2variants[i][j]) {
                variant = haplotype2variants[i + 1][k]; writeVCF(variants[variant], variant, bw); gt("0|1"); k++;
            } else {
                variant = haplotype2variants[i][j]; writeVCF(variants[variant], variant, bw); gt("1|1"); j++; k++;
            }
        }
        while (j <= last1) { variant = haplotype2variants[i][j]; writeVCF(variants[variant], variant, bw); gt("1|0"); j++; }
        while (k <= last2) { variant = haplotype2variants[i + 1][k]; writeVCF(variants[variant], variant, bw); gt("0|1"); k++; }
        fclose(bw);
    }
}

// src/SimulateHaplotypes.java:1104-1120
void SimulateHaplotypes::printPositionHistogram(const std::string& outputFile, int binSize) {
    std::vector<std::vector<int>> histogram(4, std::vector<int>(referenceLength / binSize + 1, 0));
    for (int i = 0; i <= lastVariant; i++)
        if (variants[i].isActive) histogram[variants[i].svType][variants[i].position / binSize]++;
    FILE* bw = fopen(outputFile.c_str(), "w");
    if (!bw) throw std::runtime_error("IOException: " + outputFile);
    for (size_t i = 0; i < histogram.size(); i++) {
        for (size_t j = 0; j < histogram[i].size(); j++) fprintf(bw, "%d,", histogram[i][j]);
        fputc('\n', bw);
    }
    fclose(bw);
}

// src/SimulateHaplotypes.java:1123-1144, Variant.serialize :1281-1292
void SimulateHaplotypes::serialize(const std::string& h2vFile, const std::string& variantsFile) {
    FILE* bw = fopen(h2vFile.c_str(), "wb");
    if (!bw) throw std::runtime_error("IOException: " + h2vFile);
    fprintf(bw, "%d\n", N_HAPLOTYPES);
    for (int i = 0; i < N_HAPLOTYPES; i++) {
        int last = haplotype2variants_last[i];
        fprintf(bw, "%d", last);
        for (int j = 0; j <= last; j++) fprintf(bw, ",%d", haplotype2variants[i][j]);
        fputc('\n', bw);
    }
    fclose(bw);
    bw = fopen(variantsFile.c_str(), "wb");
    if (!bw) throw std::runtime_error("IOException: " + variantsFile);
    fprintf(bw, "%d\n", lastVariant);
    for (int i = 0; i <= lastVariant; i++) {
        const Variant& v = variants[i];
        fprintf(bw, "%d\n%d\n%d\n%s\n%d\n%s\n", v.svType, v.intervalType, v.length,
                jc::doubleToString(v.alleleFrequency).c_str(), v.position, v.isActive ? "true" : "false");
        if (v.hasSequence) fwrite(v.sequence.data(), 1, v.sequence.size(), bw); else fputs("null", bw);
        fputc('\n', bw);
        fprintf(bw, "%d\n%d\n%d\n", v.intervalType_start, v.intervalType_end, v.intervalSurface);
    }
    fclose(bw);
}

// src/SimulateHaplotypes.java:1147-1178, Variant.deserialize :1294-1306
void SimulateHaplotypes::deserialize(const std::string& h2vFile, const std::string& variantsFile) {
    auto must = [](jc::LineReader& br) { std::string s; if (!br.readLine(s)) throw std::runtime_error("NullPointerException: unexpected end of file"); return s; };
    {
        jc::LineReader br(h2vFile);
        N_HAPLOTYPES = jc::parseInt(must(br));
        haplotype2variants.assign(N_HAPLOTYPES, std::vector<int>());
        haplotype2variants_last.assign(N_HAPLOTYPES, -1);
        for (int i = 0; i < N_HAPLOTYPES; i++) {
            auto tokens = jc::splitComma(must(br));
            if (tokens.empty()) throw std::runtime_error("ArrayIndexOutOfBoundsException: 0");
            int last = jc::parseInt(tokens[0]);
            haplotype2variants_last[i] = last;
            if (last + 1 < 0) throw std::runtime_error("NegativeArraySizeException");
            haplotype2variants[i].assign(last + 1, 0);
            for (int j = 0; j <= last; j++) {
                if (1 + j >= (int)tokens.size()) throw std::runtime_error("ArrayIndexOutOfBoundsException: token");
                haplotype2variants[i][j] = jc::parseInt(tokens[1 + j]);
            }
        }
    }
    {
        jc::LineReader br(variantsFile);
        lastVariant = jc::parseInt(must(br));
        variants.assign(lastVariant + 1, Variant());
        for (int i = 0; i <= lastVariant; i++) {
            Variant& v = variants[i];
            v.svType = jc::parseInt(must(br));
            v.intervalType = jc::parseInt(must(br));
            v.length = jc::parseInt(must(br));
            v.alleleFrequency = jc::parseDouble(must(br));
            v.position = jc::parseInt(must(br));
            std::string b = must(br);
            v.isActive = strcasecmp(b.c_str(), "true") == 0;  // Boolean.parseBoolean
            v.sequence = must(br);
            v.hasSequence = strcasecmp(v.sequence.c_str(), "null") != 0;
            if (!v.hasSequence) v.sequence.clear();
            v.intervalType_start = jc::parseInt(must(br));
            v.intervalType_end = jc::parseInt(must(br));
            v.intervalSurface = jc::parseInt(must(br));
        }
    }
}

// src/SimulateHaplotypes.java:75-116
int SimulateHaplotypes::mainSimulate(const std::vector<std::string>& args) {
    if (args.size() < 11) throw std::runtime_error("ArrayIndexOutOfBoundsException: args");
    const std::string MODEL_DIR = args[0], POPULATION_NAME = args[1], REFERENCE_FILE = args[2];
    const std::string REPEAT_INTERVALS_FILE = args[3], SEGDUPS_FILE = args[4], INSERTION_STRINGS_FILE = args[5];
    const int N_INSERTION_STRINGS = jc::parseInt(args[6]);
    const double ALLELE_FREQUENCY_THRESHOLD = jc::parseDouble(args[7]);
    const int N_VARIANTS_FREQUENT = jc::parseInt(args[8]);
    const int N_INDIVIDUALS = jc::parseInt(args[9]);
    const std::string OUTPUT_DIR = args[10];
    const int MAX_DISTANCE_TO_REPEAT_INTERVAL = 100;
    const int HISTOGRAM_BIN_SIZE = 100000;
    // wall-clock seconds per phase (bench.py's prologue / reference-arm report): phaseSeconds[k], names in phaseNames
    auto now = []() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t = now();
    auto lap = [&](const char* name) { const double u = now(); phaseNames.push_back(name); phaseSeconds.push_back(u - t); t = u; };
    loadReference(REFERENCE_FILE); lap("loadReference");
    model.deserializeStatistics(MODEL_DIR);
    model.deserializeRepeatIntervals(REPEAT_INTERVALS_FILE);
    model.deserializeSegdups(SEGDUPS_FILE); lap("deserializeModel");
    loadPositions(referenceLength, BuildModel::MAX_DISTANCE_SR); lap("loadPositions");
    sampleVariants(ALLELE_FREQUENCY_THRESHOLD, N_VARIANTS_FREQUENT, jc::binarySearchStr(BuildModel::AF_NAMES, 6, POPULATION_NAME));
    lap("sampleVariants");
    drawLog.steps_after_sampleVariants = random.steps;
    if (verbose) fprintf(stderr, "DONE  Total variants=%d. Frequent=%d Rare=%d\n", lastVariant + 1, nVariants_frequent, nVariants_rare);
    buildPopulation(N_INDIVIDUALS, ALLELE_FREQUENCY_THRESHOLD, OUTPUT_DIR);
    buildVariant2haplotypes(); lap("buildPopulation");
    loadInsertions(INSERTION_STRINGS_FILE, N_INSERTION_STRINGS, BuildModel::SV_LENGTH_QUANTUM);
    assignInsertionStrings(); lap("assignInsertionStrings");
    drawLog.steps_after_assignInsertionStrings = random.steps;
    annotateVariantsWithRepeats(MAX_DISTANCE_TO_REPEAT_INTERVAL); lap("annotateVariantsWithRepeats");
    buildVCF(OUTPUT_DIR); lap("buildVCF");
    serialize(OUTPUT_DIR + "/haplotype2variants.txt", OUTPUT_DIR + "/variants.txt");
    printPositionHistogram(OUTPUT_DIR + "/histogram_startPositions.txt", HISTOGRAM_BIN_SIZE); lap("serialize");
    return 0;
}

// src/PrintHaplotypes.java:6-18
int SimulateHaplotypes::mainPrint(const std::vector<std::string>& args) {
    if (args.size() < 6) throw std::runtime_error("ArrayIndexOutOfBoundsException: args");
    const int FIRST_HAPLOTYPE = jc::parseInt(args[0]);
    const int LAST_HAPLOTYPE = jc::parseInt(args[1]);
    loadReference(args[2]);
    deserialize(args[3], args[4]);
    buildHaplotypes(FIRST_HAPLOTYPE, LAST_HAPLOTYPE, args[5]);
    return 0;
}

}  // namespace oracle
