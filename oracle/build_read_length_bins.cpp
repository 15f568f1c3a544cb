// ORACLE — TEST INFRASTRUCTURE ONLY (see java_compat.h). CPU restatement of the reference's src/BuildReadLengthBins.java
// (main :21-105, shuffle :108-116, printStats :139-208, std :219-239), src/Fastq.java and src/Fastq2LengthHistogram.java
// (SURVEY.md §8f rank 3). Same control flow: bins are written in input order, re-read one by one, and permuted by the
// sequential Fisher-Yates with ONE java.util.Random shared by all bins (bins with fewer than three records are left as
// they are and draw nothing, :95). The seed replaces the reference's unseeded `new Random()` (:34). Parity unpinned by the
// reference (no tests, no JVM here); java.util.Random and Double.toString are pinned by tests/test_oracle_kat.py.
#include "java_compat.h"
#include <cmath>

namespace oracle {

namespace {

struct LineReader {   // BufferedReader.readLine
    std::string buf;
    size_t pos = 0;
    explicit LineReader(const std::string& path) {
        FILE* f = fopen(path.c_str(), "rb");
        if (!f) throw std::runtime_error("java.io.FileNotFoundException: " + path + " (No such file or directory)");
        char tmp[1 << 16];
        size_t n;
        while ((n = fread(tmp, 1, sizeof tmp, f)) > 0) buf.append(tmp, n);
        fclose(f);
    }
    bool next(std::string& line) {
        if (pos >= buf.size()) return false;
        size_t e = pos;
        while (e < buf.size() && buf[e] != '\n' && buf[e] != '\r') e++;
        line.assign(buf, pos, e - pos);
        if (e < buf.size()) { if (buf[e] == '\r' && e + 1 < buf.size() && buf[e + 1] == '\n') e++; e++; }
        pos = e;
        return true;
    }
};

struct Fastq { std::string header, sequence, separator, quality; };

struct Writer {
    FILE* f;
    explicit Writer(const std::string& path) : f(fopen(path.c_str(), "wb")) {
        if (!f) throw std::runtime_error("java.io.FileNotFoundException: " + path);
    }
    ~Writer() { if (f) fclose(f); }
    void line(const std::string& s) { fwrite(s.data(), 1, s.size(), f); fputc('\n', f); }
    void str(const std::string& s) { fwrite(s.data(), 1, s.size(), f); }
    void close() { fclose(f); f = nullptr; }
};

const int MIN_READS_IN_LOCAL_MAXIMUM = 500, MIN_LOCAL_MAXIMUM = 1000, MAX_LOCAL_MAXIMUM = 25000;  // :14-16

[[noreturn]] void npe() { throw std::runtime_error("java.lang.NullPointerException"); }

// the four readLine calls of one record; false at a clean end of file
bool readRecord(LineReader& br, Fastq& r) {
    if (!br.next(r.header)) return false;
    if (!br.next(r.sequence)) npe();      // sequence.length() on null (:57)
    if (!br.next(r.separator)) npe();     // BufferedWriter.write((String)null) (:62)
    if (!br.next(r.quality)) npe();
    return true;
}

double stdOneSide(int from, int to, int bin, const std::vector<int>& buffers_last, int BIN_LENGTH, int HALF_BIN_LENGTH) {  // :219-239
    double k, n, std_;
    const double mean = bin * BIN_LENGTH + HALF_BIN_LENGTH;
    std_ = 0.0; n = 0.0;
    for (int i = from; i <= to; i++) {
        n += buffers_last[i] + 1;
        k = i * BIN_LENGTH + HALF_BIN_LENGTH - mean;
        std_ += (buffers_last[i] + 1) * k * k;
    }
    n = n + (buffers_last[bin] + 1) + n;
    std_ = std_ + 0 + std_;
    return std::sqrt(std_ / n);
}

void printStats(const std::vector<int>& buffers_last, const std::vector<long long>& buffers_stringLength, const std::string& outputFile,
                int N_BINS, int BIN_LENGTH, int HALF_BIN_LENGTH, long long GENOME_LENGTH_HAPLOID) {  // :139-208
    const int WINDOW_RADIUS = 1;
    int i, j;
    int nMaxima, lastMaximum, previousMaximum, leftMinBin, leftRightMinBin;
    double sum, mass;
    std::vector<int> smoothed(N_BINS);
    for (i = 0; i < WINDOW_RADIUS; i++) smoothed[i] = buffers_last[i];
    for (i = WINDOW_RADIUS; i < N_BINS - WINDOW_RADIUS; i++) {
        smoothed[i] = 0;
        for (j = i - WINDOW_RADIUS; j <= i + WINDOW_RADIUS; j++) smoothed[i] += buffers_last[j];
        smoothed[i] /= WINDOW_RADIUS + 1 + WINDOW_RADIUS;   // int division truncates towards zero, also for the -1 entries
    }
    for (i = N_BINS - WINDOW_RADIUS; i < N_BINS; i++) smoothed[i] = buffers_last[i];
    Writer bw(outputFile);
    nMaxima = 0; lastMaximum = -1; previousMaximum = -1;
    for (i = 1; i < N_BINS - 1; i++) {
        if ((i + 1) * BIN_LENGTH - 1 >= MIN_LOCAL_MAXIMUM && i * (BIN_LENGTH) <= MAX_LOCAL_MAXIMUM &&
            buffers_last[i] + 1 >= MIN_READS_IN_LOCAL_MAXIMUM && smoothed[i] > smoothed[i - 1] && smoothed[i] > smoothed[i + 1]) {
            nMaxima++;
            previousMaximum = lastMaximum; lastMaximum = i;
        }
    }
    bw.str(std::to_string(nMaxima) + "\n");
    if (nMaxima < 2) { bw.close(); return; }
    leftRightMinBin = -1;
    for (i = previousMaximum + 1; i < lastMaximum; i++) {
        if (smoothed[i] < smoothed[i - 1] && smoothed[i] < smoothed[i + 1]) { leftRightMinBin = i; break; }
    }
    leftMinBin = 0;
    for (i = 1; i < previousMaximum; i++) {
        if (smoothed[i] < smoothed[i - 1] && smoothed[i] < smoothed[i + 1]) { leftMinBin = i; break; }
    }
    bw.str(std::to_string(previousMaximum * BIN_LENGTH + HALF_BIN_LENGTH) + "\n");
    bw.str(jc::doubleToString(stdOneSide(leftMinBin, previousMaximum - 1, previousMaximum, buffers_last, BIN_LENGTH, HALF_BIN_LENGTH)) + "\n");
    bw.str(std::to_string(lastMaximum * BIN_LENGTH + HALF_BIN_LENGTH) + "\n");
    bw.str(jc::doubleToString(stdOneSide(lastMaximum + 1, N_BINS - 1, lastMaximum, buffers_last, BIN_LENGTH, HALF_BIN_LENGTH)) + "\n");
    bw.str(std::to_string(leftMinBin) + "\n");
    bw.str(std::to_string(leftRightMinBin) + "\n");
    sum = 0.0;
    for (i = 0; i < N_BINS; i++) sum += buffers_last[i] + 1;
    mass = 0.0;
    for (i = leftMinBin; i <= leftRightMinBin; i++) mass += buffers_last[i] + 1;
    mass /= sum;
    bw.str(jc::doubleToString(mass) + "\n");
    mass = 0.0;
    for (i = leftRightMinBin + 1; i < N_BINS; i++) mass += buffers_last[i] + 1;
    mass /= sum;
    bw.str(jc::doubleToString(mass) + "\n");
    sum = 0.0;
    for (i = leftMinBin; i <= leftRightMinBin; i++) sum += (double)buffers_stringLength[i];
    bw.str(jc::doubleToString(sum / (double)(GENOME_LENGTH_HAPLOID << 1)) + "\n");
    sum = 0.0;
    for (i = leftRightMinBin + 1; i < N_BINS; i++) sum += (double)buffers_stringLength[i];
    bw.str(jc::doubleToString(sum / (double)(GENOME_LENGTH_HAPLOID << 1)) + "\n");
    bw.close();
}

}  // namespace

// BuildReadLengthBins.main: INPUT_FILES_LIST BIN_LENGTH MAX_READ_LENGTH GENOME_LENGTH_HAPLOID OUTPUT_PREFIX
int buildReadLengthBinsMain(const std::vector<std::string>& args, int64_t seed) {
    if (args.size() < 5) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: args");
    const std::string INPUT_FILES_LIST = args[0];
    const int BIN_LENGTH = jc::parseInt(args[1]);
    const int MAX_READ_LENGTH = jc::parseInt(args[2]);
    const long long GENOME_LENGTH_HAPLOID = atoll(args[3].c_str());
    const std::string OUTPUT_PREFIX = args[4];
    if (BIN_LENGTH == 0) throw std::runtime_error("java.lang.ArithmeticException: / by zero");
    const int N_BINS = (MAX_READ_LENGTH + BIN_LENGTH - 1) / BIN_LENGTH;
    const int HALF_BIN_LENGTH = BIN_LENGTH >> 1;
    if (N_BINS < 0) throw std::runtime_error("java.lang.NegativeArraySizeException: " + std::to_string(N_BINS));
    jc::JavaRandom random(seed);

    // Building bins (:43-75)
    std::vector<Writer*> buffers(N_BINS, nullptr);
    struct Closer { std::vector<Writer*>& b; ~Closer() { for (Writer* w : b) delete w; } } closer{buffers};
    for (int i = 0; i < N_BINS; i++) buffers[i] = new Writer(OUTPUT_PREFIX + std::to_string(i) + ".bin");
    std::vector<int> buffers_last(N_BINS, -1);
    std::vector<long long> buffers_stringLength(N_BINS, 0);
    {
        LineReader br1(INPUT_FILES_LIST);
        std::string file;
        while (br1.next(file)) {
            LineReader br2(file);
            Fastq r;
            while (readRecord(br2, r)) {
                int bin = (int)r.sequence.size() / BIN_LENGTH;
                if (bin >= N_BINS) bin = N_BINS - 1;
                if (bin < 0) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: Index " + std::to_string(bin));
                buffers_last[bin]++;
                buffers[bin]->line(r.header);
                buffers[bin]->line(r.sequence);
                buffers_stringLength[bin] += (long long)r.sequence.size();
                buffers[bin]->line(r.separator);
                buffers[bin]->line(r.quality);
            }
        }
    }
    for (int i = 0; i < N_BINS; i++) buffers[i]->close();
    int maxLength = 0;
    {
        Writer bw(OUTPUT_PREFIX + ".histogram");
        for (int i = 0; i < N_BINS; i++) {
            const int length = buffers_last[i] + 1;
            if (length > maxLength) maxLength = length;
            bw.str(std::to_string(i * BIN_LENGTH + HALF_BIN_LENGTH) + "," + std::to_string(length) + "\n");
        }
        bw.close();
    }
    printStats(buffers_last, buffers_stringLength, OUTPUT_PREFIX + ".stats", N_BINS, BIN_LENGTH, HALF_BIN_LENGTH, GENOME_LENGTH_HAPLOID);

    // Permuting reads in each bin (:86-103)
    std::vector<Fastq> tmpArray((size_t)maxLength);
    for (int i = 0; i < N_BINS; i++) {
        int last = -1;
        {
            LineReader br1(OUTPUT_PREFIX + std::to_string(i) + ".bin");
            Fastq r;
            while (readRecord(br1, r)) tmpArray[(size_t)++last] = r;
        }
        if (last <= 1) continue;
        for (int k = last; k > 0; k--) {   // shuffle :108-116
            const int j = random.nextInt(k + 1);
            std::swap(tmpArray[(size_t)j], tmpArray[(size_t)k]);
        }
        Writer bw(OUTPUT_PREFIX + std::to_string(i) + ".bin");
        for (int j = 0; j <= last; j++) {   // Fastq.writeTo
            bw.line(tmpArray[(size_t)j].header); bw.line(tmpArray[(size_t)j].sequence);
            bw.line(tmpArray[(size_t)j].separator); bw.line(tmpArray[(size_t)j].quality);
        }
        bw.close();
    }
    return 0;
}

// Fastq2LengthHistogram.main (src/Fastq2LengthHistogram.java:11-51): FASTQ_FILE BIN_LENGTH MAX_READ_LENGTH
// OUTPUT_FILE_HISTOGRAM OUTPUT_FILE_MAX
int fastq2LengthHistogramMain(const std::vector<std::string>& args) {
    if (args.size() < 5) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: args");
    const int BIN_LENGTH = jc::parseInt(args[1]), MAX_READ_LENGTH = jc::parseInt(args[2]);
    if (BIN_LENGTH == 0) throw std::runtime_error("java.lang.ArithmeticException: / by zero");
    const int N_BINS = (MAX_READ_LENGTH + BIN_LENGTH - 1) / BIN_LENGTH;
    if (N_BINS < 0) throw std::runtime_error("java.lang.NegativeArraySizeException: " + std::to_string(N_BINS));
    std::vector<int> histogram(N_BINS, 0);
    LineReader br(args[0]);
    std::string str;
    bool have = br.next(str);
    while (have) {
        if (!br.next(str)) npe();                 // str.length() on null (:33)
        const int length = (int)str.size();
        br.next(str); br.next(str);               // separator and quality: a missing one is only a null assignment (:34)
        int bin = length / BIN_LENGTH;
        if (bin >= N_BINS) bin = N_BINS - 1;
        if (bin < 0) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: Index " + std::to_string(bin));
        histogram[bin]++;
        have = br.next(str);
    }
    {
        Writer bw(args[4]);
        for (int i = 1; i < N_BINS - 1; i++) {
            if ((i + 1) * BIN_LENGTH - 1 >= MIN_LOCAL_MAXIMUM && i * (BIN_LENGTH) <= MAX_LOCAL_MAXIMUM &&
                histogram[i] + 1 >= MIN_READS_IN_LOCAL_MAXIMUM && histogram[i] > histogram[i - 1] && histogram[i] > histogram[i + 1])
                bw.str(std::to_string(i * BIN_LENGTH) + "," + std::to_string((i + 1) * BIN_LENGTH - 1) + "\n");
        }
        bw.close();
    }
    Writer bw(args[3]);
    for (int i = 0; i < N_BINS; i++) bw.str(std::to_string(i * BIN_LENGTH) + "," + std::to_string(histogram[i]) + "\n");
    bw.close();
    return 0;
}

}  // namespace oracle
