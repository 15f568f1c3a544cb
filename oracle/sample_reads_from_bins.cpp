// ORACLE — TEST INFRASTRUCTURE ONLY. CPU restatement of src/SampleReadsFromLengthBins.java:19-102.
// The cumulative bimodal table (buildCumulativeBimodal :87-102) depends on Apache commons-math3 3.6.1
// NormalDistribution.probability (not vendored in the reference, docker/Dockerfile:228-232); the drop-in
// boundary therefore takes the table as an input. When no table is supplied this restatement computes it
// with libm erf() following NormalDistribution.probability's formula 0.5*erf(v0,v1) — values can differ
// from commons-math in the last ulp; parity for this row is on the sampling given the same table.
#include "java_compat.h"
#include <memory>

namespace oracle {

std::vector<double> buildCumulativeBimodal(int nBins, int binLength, double meanLeft, double stdLeft,
                                           double meanRight, double stdRight, double weightLeft) {
    const double weightRight = 1.0 - weightLeft;
    auto prob = [](double mean, double sd, double x0, double x1) {
        const double denom = sd * std::sqrt(2.0);
        const double v0 = (x0 - mean) / denom, v1 = (x1 - mean) / denom;
        // Erf.erf(x1,x2) of commons-math: picks the numerically best of erf/erfc differences.
        double e;
        if (v0 > 0.476936 && v1 > 0.476936) e = std::erfc(v0) - std::erfc(v1);
        else if (v0 < -0.476936 && v1 < -0.476936) e = std::erfc(-v1) - std::erfc(-v0);
        else e = std::erf(v1) - std::erf(v0);
        return 0.5 * e;
    };
    std::vector<double> c(nBins);
    for (int i = 0; i < nBins; i++)
        c[i] = prob(meanLeft, stdLeft, (double)(i * binLength), (double)((i + 1) * binLength)) * weightLeft +
               prob(meanRight, stdRight, (double)(i * binLength), (double)((i + 1) * binLength)) * weightRight;
    double sum = 0.0;
    for (int i = 0; i < nBins; i++) sum += c[i];
    for (int i = 0; i < nBins; i++) c[i] /= sum;
    for (int i = 1; i < nBins; i++) c[i] += c[i - 1];
    return c;
}

// Returns the process exit code (3 = bins ran dry, src/SampleReadsFromLengthBins.java:64-67).
// cdfFile: optional text file with N_BINS lines holding the cumulative table (Double.toString format).
int sampleReadsFromLengthBinsMain(const std::vector<std::string>& args, int64_t seed, const std::string& cdfFile) {
    if (args.size() < 11) throw std::runtime_error("ArrayIndexOutOfBoundsException: args");
    const double MEAN_LEFT = jc::parseDouble(args[0]), STD_LEFT = jc::parseDouble(args[1]);
    const double MEAN_RIGHT = jc::parseDouble(args[2]), STD_RIGHT = jc::parseDouble(args[3]);
    const double WEIGHT_LEFT = jc::parseDouble(args[4]);
    const int BIN_LENGTH = jc::parseInt(args[5]);
    const int MAX_READ_LENGTH = jc::parseInt(args[6]);
    const std::string BINS_PREFIX = args[7];
    const int64_t GENOME_LENGTH_EACH_HAPLOTYPE = std::stoll(args[8]);
    const double TARGET_COVERAGE_EACH_HAPLOTYPE = jc::parseDouble(args[9]);
    const std::string OUTPUT_FASTQ_FILE = args[10];

    const int N_BINS = (MAX_READ_LENGTH + BIN_LENGTH - 1) / BIN_LENGTH;
    const int64_t TARGET_COVERAGE_BP = (int64_t)((double)GENOME_LENGTH_EACH_HAPLOTYPE * (TARGET_COVERAGE_EACH_HAPLOTYPE * 2));
    const int MAX_SAMPLING_ATTEMPTS = N_BINS * 10;

    jc::JavaRandom random(seed);
    std::vector<double> cumulativeBimodal;
    if (cdfFile.empty())
        cumulativeBimodal = buildCumulativeBimodal(N_BINS, BIN_LENGTH, MEAN_LEFT, STD_LEFT, MEAN_RIGHT, STD_RIGHT, WEIGHT_LEFT);
    else {
        jc::LineReader br(cdfFile);
        std::string s;
        while (br.readLine(s)) if (!jc::trim(s).empty()) cumulativeBimodal.push_back(jc::parseDouble(s));
        if ((int)cumulativeBimodal.size() != N_BINS) throw std::runtime_error("cdf table must have N_BINS entries");
    }
    std::vector<std::unique_ptr<jc::LineReader>> bins(N_BINS);
    for (int i = 0; i < N_BINS; i++) bins[i].reset(new jc::LineReader(BINS_PREFIX + std::to_string(i) + ".bin"));
    std::vector<int> histogram(N_BINS, 0);
    int64_t coverage = 0;
    int read = -1;
    FILE* bw = fopen(OUTPUT_FASTQ_FILE.c_str(), "wb");
    if (!bw) throw std::runtime_error("IOException: " + OUTPUT_FASTQ_FILE);
    std::string str;
    while (coverage < TARGET_COVERAGE_BP) {
        int j = 0, bin = -1;
        while (j < MAX_SAMPLING_ATTEMPTS) {
            bin = jc::binarySearch(cumulativeBimodal.data(), 0, N_BINS, random.nextDouble());
            if (bin < 0) bin = -1 - bin;
            else if (bin == N_BINS) bin = N_BINS - 1;
            if (bin >= N_BINS) { fclose(bw); throw std::runtime_error("ArrayIndexOutOfBoundsException: bins[" + std::to_string(bin) + "]"); }
            // bins[bin].mark(10); c=bins[bin].read(); bins[bin].reset();
            int c = getc_unlocked(bins[bin]->f);
            if (c != EOF) { ungetc(c, bins[bin]->f); break; }
            j++;
        }
        if (j == MAX_SAMPLING_ATTEMPTS) {
            fprintf(stderr, "Not enough reads to sample %sx coverage for each haplotype from the given bins and distribution.\n",
                    jc::doubleToString(TARGET_COVERAGE_EACH_HAPLOTYPE).c_str());
            fflush(bw);  // System.exit(3) does not flush the BufferedWriter; the partial file is unspecified
            fclose(bw);
            return 3;
        }
        histogram[bin]++;
        bins[bin]->readLine(str);
        fprintf(bw, "@read%d\n", ++read);
        if (!bins[bin]->readLine(str)) { fclose(bw); throw std::runtime_error("NullPointerException: truncated bin file"); }
        coverage += (int64_t)str.size();
        fwrite(str.data(), 1, str.size(), bw); fputc('\n', bw);
        if (!bins[bin]->readLine(str)) { fclose(bw); throw std::runtime_error("NullPointerException: truncated bin file"); }
        fwrite(str.data(), 1, str.size(), bw); fputc('\n', bw);
        if (!bins[bin]->readLine(str)) { fclose(bw); throw std::runtime_error("NullPointerException: truncated bin file"); }
        fwrite(str.data(), 1, str.size(), bw); fputc('\n', bw);
    }
    fclose(bw);
    bw = fopen((OUTPUT_FASTQ_FILE + ".histogram").c_str(), "wb");
    if (!bw) throw std::runtime_error("IOException: histogram");
    for (int i = 0; i < N_BINS; i++) fprintf(bw, "%d\n", histogram[i]);
    fclose(bw);
    return 0;
}

}  // namespace oracle
