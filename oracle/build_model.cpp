// ORACLE — TEST INFRASTRUCTURE ONLY (see java_compat.h). CPU restatement of the reference's src/BuildModel.java main path
// (SURVEY.md §8f rank 4): RepeatMasker + TRF tracks -> repeat intervals, segdups, gnomAD-SV records -> the statistics files
// SimulateHaplotypes reads. Same arrays, same loop order, same floating-point operation order as the Java (doubles are
// summed in index order; -ffp-contract=off). Parity unpinned by the reference (no tests, no JVM here, gnomAD-SV inputs not
// in the image): pinned only through the JDK primitives (tests/test_oracle_kat.py) and tests/test_build_model.py's
// hand-computed cases.
// Restrictions (stated): ASCII case folding; Double.parseDouble over what strtod accepts; a record that would raise a Java
// exception (missing ';' after a key, an SV type without autosomal records) raises here too.
#include "java_compat.h"
#include <climits>
#include <fstream>

namespace oracle {

namespace {

const int MAX_DISTANCE_REPEATMASKER = 25, MIN_OVERLAP_TRF = 10, MIN_OVERLAP_INTERVALS = 25;            // :12-14
const int MAX_DISTANCE_SR = 25, MAX_DISTANCE_PE = 250, MAX_DISTANCE_OTHER = 1000, MAX_DISTANCE_SEGDUPS = 25;  // :15-18
const int ADD_TYPE[3][3] = {{0, 2, 2}, {2, 1, 2}, {2, 2, 2}};                                          // :83
const int N_CHROMOSOMES = 24, N_AUTOSOMES = 22, SV_POSITION_QUANTUM = 100000;                          // :628-634
const int SMOOTHING_RADIUS_POSITIONS = 5, SMOOTHING_WINDOW_POSITIONS = 11;                             // :637-638
const int N_SV_TYPES = 4;
const char* const SV_TYPES[N_SV_TYPES] = {"DEL", "DUP", "INS", "INV"};                                 // :837
const int MIN_SV_LENGTH = 50, MAX_SV_LENGTH = 1000000, SV_LENGTH_QUANTUM = 50, ALLELE_FREQUENCY_NBINS = 10000;  // :838-841
const int N_AF = 6;
const char* const AF_FILE_NAMES[N_AF] = {"statistics_alleleFrequency_ALL", "statistics_alleleFrequency_AFR", "statistics_alleleFrequency_AMR",
                                         "statistics_alleleFrequency_EAS", "statistics_alleleFrequency_EUR", "statistics_alleleFrequency_OTH"};  // :843
const char* const ALLELE_FREQUENCY_STR[N_AF] = {";AF=", ";AFR_AF=", ";AMR_AF=", ";EAS_AF=", ";EUR_AF=", ";OTH_AF="};  // :844
const int ALLELE_FREQUENCY_STR_LENGTH[N_AF] = {4, 8, 8, 8, 8, 8};                                      // :870
const int N_INTERVAL_TYPES = 5, SMOOTHING_RADIUS = 4, SMOOTHING_WINDOW = 9;                            // :854,871-872
const double MIN_FREQUENCY = 0.01;                                                                     // :873
const char* const ARTIFICIAL_SHORT_PERIOD = "ARTIFICIAL_SHORT_PERIOD";                                 // :332

std::string toLower(std::string s) { for (auto& ch : s) if (ch >= 'A' && ch <= 'Z') ch = (char)(ch - 'A' + 'a'); return s; }
bool eqIgnoreCase(const std::string& a, const std::string& b) { return a.size() == b.size() && strcasecmp(a.c_str(), b.c_str()) == 0; }
int indexOf(const std::string& s, const std::string& what, int from = 0) {
    if (from < 0) from = 0;
    if ((size_t)from > s.size()) return -1;
    size_t p = s.find(what, (size_t)from);
    return p == std::string::npos ? -1 : (int)p;
}
std::string substring(const std::string& s, int b, int e) {
    if (b < 0 || e > (int)s.size() || b > e) throw std::runtime_error("java.lang.StringIndexOutOfBoundsException: begin " + std::to_string(b) + ", end " + std::to_string(e) + ", length " + std::to_string(s.size()));
    return s.substr((size_t)b, (size_t)(e - b));
}
// String.split(<one char>): trailing empty strings removed
std::vector<std::string> split(const std::string& s, char sep) {
    std::vector<std::string> out;
    size_t a = 0;
    for (;;) {
        size_t b = s.find(sep, a);
        if (b == std::string::npos) { out.push_back(s.substr(a)); break; }
        out.push_back(s.substr(a, b - a));
        a = b + 1;
    }
    while (out.size() > 1 && out.back().empty()) out.pop_back();
    return out;
}
const std::string& tok(const std::vector<std::string>& t, size_t i) {
    if (i >= t.size()) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: Index " + std::to_string(i) + " out of bounds for length " + std::to_string(t.size()));
    return t[i];
}
int parseInt(const std::string& s) {
    auto bad = [&]() { return std::runtime_error("java.lang.NumberFormatException: For input string: \"" + s + "\""); };
    if (s.empty()) throw bad();
    size_t i = 0;
    bool neg = false;
    if (s[0] == '-' || s[0] == '+') { neg = s[0] == '-'; i = 1; if (s.size() == 1) throw bad(); }
    int64_t v = 0;
    for (; i < s.size(); i++) {
        if (s[i] < '0' || s[i] > '9') throw bad();
        v = v * 10 + (s[i] - '0');
        if (v > (int64_t)INT_MAX + 1) throw bad();
    }
    if (neg) v = -v;
    if (v > INT_MAX || v < INT_MIN) throw bad();
    return (int)v;
}
double parseDouble(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && (unsigned char)s[a] <= ' ') a++;
    while (b > a && (unsigned char)s[b - 1] <= ' ') b--;
    std::string t = s.substr(a, b - a);
    if (t.empty()) throw std::runtime_error("java.lang.NumberFormatException: empty String");
    char* end = nullptr;
    double d = strtod(t.c_str(), &end);
    if (*end != 0) throw std::runtime_error("java.lang.NumberFormatException: For input string: \"" + s + "\"");
    return d;
}
bool readLine(std::ifstream& in, std::string& str) {
    if (!std::getline(in, str)) return false;
    if (!str.empty() && str.back() == '\r') str.pop_back();
    return true;
}

struct Tuple {  // :331-396
    std::string chromosome, repeat, repeatFamily;
    int referenceStart, referenceEnd, repeatStart, repeatEnd, orientation;
    bool isSatellite;
    void merge(const Tuple& o) {  // :352-374
        bool changed = false;
        if (o.referenceStart < referenceStart) { referenceStart = o.referenceStart; changed = true; }
        if (o.referenceEnd > referenceEnd) { referenceEnd = o.referenceEnd; changed = true; }
        isSatellite |= o.isSatellite;
        if (changed) {
            repeatStart = -1; repeatEnd = -1;
            if (repeat != ARTIFICIAL_SHORT_PERIOD && o.repeat == ARTIFICIAL_SHORT_PERIOD) { repeat = ARTIFICIAL_SHORT_PERIOD; repeatFamily = ""; orientation = 1; }
        }
    }
    bool operator<(const Tuple& o) const {  // compareTo :380-390
        int i = chromosome.compare(o.chromosome);
        if (i != 0) return i < 0;
        if (referenceStart != o.referenceStart) return referenceStart < o.referenceStart;
        return referenceEnd < o.referenceEnd;
    }
};

struct Model {
    // repeat track (:58-70)
    std::vector<int> repeatTrack[5];
    int lastRepeatTrack = -1;
    std::vector<std::string> chromosomeIDs;
    std::vector<char> isSatellite;
    std::vector<std::array<std::string, 2>> repeatTrackIDs;
    // repeat intervals (:77-81)
    std::vector<std::array<int, 3>> repeatIntervals;
    int lastRepeatInterval = -1;
    std::vector<std::string> repeatIntervals_chromosomes;
    // segdups (:491-495)
    std::vector<std::array<int, 2>> segdups;
    int lastSegdup = -1;
    std::vector<std::string> segdups_chromosomeIDs;
    // positions (:643-654)
    std::vector<int> chromosomeLengths, chromosomePLengths;
    std::vector<std::vector<std::vector<double>>> svs_per_position;
    std::vector<std::vector<double>> statistics_position;
    // statistics (:810-833)
    std::vector<std::vector<std::vector<double>>> statistics_length;
    std::vector<std::vector<double>> statistics_intervalType;
    std::vector<double> statistics_svType;
    std::vector<std::vector<std::vector<double>>> statistics_alleleFrequency;
    std::vector<std::vector<int>> statistics_lastAlleleFrequency;
    std::vector<std::vector<std::array<double, 2>>> statistics_alleleFrequency_minmax;
    int n_chr1 = 0, nFrequent_chr1 = 0;

    // src/BuildModel.java:101-192
    void loadRepeatMaskerAnnotations(const std::string& path, int nRecords, int maxDistance) {
        chromosomeIDs.assign(nRecords, "");
        for (auto& r : repeatTrack) r.assign(nRecords, 0);
        isSatellite.assign(nRecords, 0);
        repeatTrackIDs.assign(nRecords, {});
        std::ifstream br(path);
        if (!br) throw std::runtime_error("java.io.FileNotFoundException: " + path);
        std::string str;
        lastRepeatTrack = -1;
        while (readLine(br, str)) {
            auto tokens = split(str, ',');
            if (tok(tokens, 4).size() > 5 || eqIgnoreCase(tokens[4], "chrM")) break;
            lastRepeatTrack++;
            if (lastRepeatTrack >= nRecords) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: more RepeatMasker rows than declared");
            chromosomeIDs[lastRepeatTrack] = tokens[4];
            int referenceStart = parseInt(tok(tokens, 5)) - 1, referenceEnd = parseInt(tok(tokens, 6)) - 1;
            repeatTrack[0][lastRepeatTrack] = referenceStart;
            repeatTrack[1][lastRepeatTrack] = referenceEnd;
            if (tok(tokens, 8).empty()) throw std::runtime_error("java.lang.StringIndexOutOfBoundsException: strand");
            bool orientation = tokens[8][0] == '+';
            int repeatStart, repeatEnd;
            if (orientation) { repeatStart = parseInt(tok(tokens, 11)) - 1; repeatEnd = parseInt(tok(tokens, 12)) - 1; }
            else { repeatEnd = parseInt(tok(tokens, 12)) - 1; repeatStart = parseInt(tok(tokens, 13)) - 1; }
            repeatTrack[2][lastRepeatTrack] = repeatStart;
            repeatTrack[3][lastRepeatTrack] = repeatEnd;
            repeatTrack[4][lastRepeatTrack] = orientation ? 1 : 0;
            repeatTrackIDs[lastRepeatTrack][0] = toLower(tok(tokens, 9));
            repeatTrackIDs[lastRepeatTrack][1] = toLower(tok(tokens, 10));
            const std::string& fam = repeatTrackIDs[lastRepeatTrack][1];
            isSatellite[lastRepeatTrack] = indexOf(fam, "satellite") >= 0 || indexOf(fam, "sat") >= 0 || indexOf(fam, "simple") >= 0 || indexOf(fam, "low_complexity") >= 0;
        }
        // compaction :159-190
        int j = 0;
        for (int i = 1; i <= lastRepeatTrack; i++) {
            if (repeatTrackIDs[i][0] == repeatTrackIDs[j][0] && repeatTrackIDs[i][1] == repeatTrackIDs[j][1] && repeatTrack[4][i] == repeatTrack[4][j] &&
                eqIgnoreCase(chromosomeIDs[i], chromosomeIDs[j]) && std::abs(repeatTrack[0][i] - repeatTrack[1][j]) <= maxDistance &&
                (isSatellite[i] || (repeatTrack[4][i] == 1 && std::abs(repeatTrack[2][i] - repeatTrack[3][j]) <= maxDistance) ||
                 (repeatTrack[4][i] == 0 && std::abs(repeatTrack[3][i] - repeatTrack[2][j]) <= maxDistance))) {
                repeatTrack[1][j] = repeatTrack[1][i];
                repeatTrack[2][j] = std::min(repeatTrack[2][j], repeatTrack[2][i]);
                repeatTrack[3][j] = std::max(repeatTrack[3][j], repeatTrack[3][i]);
            } else {
                j++;
                for (int k = 0; k < 5; k++) repeatTrack[k][j] = repeatTrack[k][i];
                isSatellite[j] = isSatellite[i];
                repeatTrackIDs[j] = repeatTrackIDs[i];
                chromosomeIDs[j] = chromosomeIDs[i];
            }
        }
        lastRepeatTrack = j;
    }

    // src/BuildModel.java:220-323
    void loadTrfAnnotations(const std::string& path, int nRecords, int minOverlap) {
        std::vector<std::array<int, 2>> trfTrack(nRecords);
        std::vector<std::string> chrIDs(nRecords);
        std::ifstream br(path);
        if (!br) throw std::runtime_error("java.io.FileNotFoundException: " + path);
        std::string str;
        int lastTrfTrack = -1;
        while (readLine(br, str)) {
            auto tokens = split(str, ',');
            lastTrfTrack++;
            if (lastTrfTrack >= nRecords) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: more TRF rows than declared");
            chrIDs[lastTrfTrack] = tok(tokens, 0);
            trfTrack[lastTrfTrack][0] = parseInt(tok(tokens, 1)) - 1;
            trfTrack[lastTrfTrack][1] = parseInt(tok(tokens, 2)) - 1;
        }
        int j = 0;
        for (int i = 1; i <= lastTrfTrack; i++) {  // :251-266
            if (eqIgnoreCase(chrIDs[i], chrIDs[j]) && trfTrack[i][0] <= trfTrack[j][1] - minOverlap) {
                if (trfTrack[i][1] > trfTrack[j][1]) trfTrack[j][1] = trfTrack[i][1];
            } else {
                j++;
                trfTrack[j] = trfTrack[i];
                chrIDs[j] = chrIDs[i];
            }
        }
        lastTrfTrack = j;
        // union :274-294
        std::vector<Tuple> tuples;
        tuples.reserve(lastRepeatTrack + lastTrfTrack + 2);
        for (int i = 0; i <= lastRepeatTrack; i++)
            tuples.push_back(Tuple{chromosomeIDs[i], repeatTrackIDs[i][0], repeatTrackIDs[i][1], repeatTrack[0][i], repeatTrack[1][i], repeatTrack[2][i], repeatTrack[3][i], repeatTrack[4][i], (bool)isSatellite[i]});
        for (int i = 0; i <= lastTrfTrack; i++)
            tuples.push_back(Tuple{chrIDs[i], ARTIFICIAL_SHORT_PERIOD, "", trfTrack[i][0], trfTrack[i][1], -1, -1, 1, true});
        int last = (int)tuples.size() - 1;
        std::stable_sort(tuples.begin(), tuples.end());  // Arrays.sort(Object[]) is stable
        j = 0;
        for (int i = 1; i <= last; i++) {
            if ((tuples[i].repeat == ARTIFICIAL_SHORT_PERIOD || tuples[j].repeat == ARTIFICIAL_SHORT_PERIOD) && tuples[i].chromosome == tuples[j].chromosome &&
                tuples[i].referenceStart <= tuples[j].referenceEnd - minOverlap) tuples[j].merge(tuples[i]);
            else {
                j++;
                std::swap(tuples[j], tuples[i]);
            }
        }
        last = j;
        // :297-311
        for (auto& r : repeatTrack) if ((int)r.size() < last + 1) r.assign(last + 1, 0);
        if ((int)isSatellite.size() < last + 1) isSatellite.assign(last + 1, 0);
        if ((int)repeatTrackIDs.size() < last + 1) repeatTrackIDs.assign(last + 1, {});
        if ((int)chromosomeIDs.size() < last + 1) chromosomeIDs.assign(last + 1, "");
        for (int i = 0; i <= last; i++) {
            repeatTrack[0][i] = tuples[i].referenceStart; repeatTrack[1][i] = tuples[i].referenceEnd;
            repeatTrack[2][i] = tuples[i].repeatStart; repeatTrack[3][i] = tuples[i].repeatEnd;
            repeatTrack[4][i] = tuples[i].orientation;
            isSatellite[i] = tuples[i].isSatellite;
            repeatTrackIDs[i][0] = tuples[i].repeat; repeatTrackIDs[i][1] = tuples[i].repeatFamily;
            chromosomeIDs[i] = tuples[i].chromosome;
        }
        lastRepeatTrack = last;
    }

    // src/BuildModel.java:401-440 (chromosome2firstInterval is only read by the commented-out getIntervalType path)
    void buildRepeatIntervals(int minOverlap) {
        if (lastRepeatTrack < 1) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: Index 0 out of bounds for length 0");
        repeatIntervals.assign(lastRepeatTrack, {});
        repeatIntervals_chromosomes.assign(lastRepeatTrack, "");
        repeatIntervals_chromosomes[0] = chromosomeIDs[0];
        repeatIntervals[0] = {isSatellite[0] ? 1 : 0, repeatTrack[0][0], repeatTrack[1][0]};
        lastRepeatInterval = 0;
        for (int i = 1; i <= lastRepeatTrack; i++) {
            if (chromosomeIDs[i] == chromosomeIDs[lastRepeatInterval] && repeatTrack[0][i] <= repeatIntervals[lastRepeatInterval][2] - minOverlap) {
                if (repeatTrack[1][i] > repeatIntervals[lastRepeatInterval][2]) repeatIntervals[lastRepeatInterval][2] = repeatTrack[1][i];
                repeatIntervals[lastRepeatInterval][0] = ADD_TYPE[repeatIntervals[lastRepeatInterval][0]][isSatellite[i] ? 1 : 0];
            } else {
                lastRepeatInterval++;
                if (lastRepeatInterval >= lastRepeatTrack) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: repeatIntervals");
                repeatIntervals[lastRepeatInterval] = {isSatellite[i] ? 1 : 0, repeatTrack[0][i], repeatTrack[1][i]};
                repeatIntervals_chromosomes[lastRepeatInterval] = chromosomeIDs[i];
            }
        }
    }

    // src/BuildModel.java:506-557
    void loadSegdups(const std::string& path, int nSegdups, int maxDistance) {
        struct Segdup { std::string chromosome; int first, last; };
        std::vector<Segdup> array;
        std::ifstream br(path);
        if (!br) throw std::runtime_error("java.io.FileNotFoundException: " + path);
        std::string str;
        readLine(br, str);  // header
        while (readLine(br, str)) {
            auto tokens = split(str, '\t');
            if ((int)array.size() >= nSegdups) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: more segdup rows than declared");
            array.push_back({tok(tokens, 1), parseInt(tok(tokens, 2)) - 1, parseInt(tok(tokens, 3)) - 2});
        }
        if ((int)array.size() != nSegdups) throw std::runtime_error("java.lang.NullPointerException: fewer segdup rows than declared");
        std::stable_sort(array.begin(), array.end(), [](const Segdup& a, const Segdup& b) {
            int c = a.chromosome.compare(b.chromosome);
            if (c != 0) return c < 0;
            if (a.first != b.first) return a.first < b.first;
            return a.last < b.last;
        });
        int j = 0;
        lastSegdup = (int)array.size() - 1;
        for (int i = 1; i <= lastSegdup; i++) {
            if (array[i].first > array[j].last + maxDistance) array[++j] = array[i];
            else if (array[i].last > array[j].last) array[j].last = array[i].last;
        }
        lastSegdup = j;
        segdups.assign(lastSegdup + 1, {});
        segdups_chromosomeIDs.assign(lastSegdup + 1, "");
        for (int i = 0; i <= lastSegdup; i++) { segdups[i] = {array[i].first, array[i].last}; segdups_chromosomeIDs[i] = array[i].chromosome; }
    }

    // src/BuildModel.java:715-729
    static int str2chromosomeID(const std::string& str) {
        char c;
        if (str.size() > 3) {
            if (!eqIgnoreCase(str.substr(0, 3), "chr")) return -1;
            if (str.size() > 5) return -1;
            c = str[3];
        } else {
            if (str.empty()) throw std::runtime_error("java.lang.StringIndexOutOfBoundsException: empty chromosome");
            c = str[0];
        }
        if (c == 'M' || c == 'm') return -1;
        if (c == 'X' || c == 'x') return N_CHROMOSOMES - 2;
        if (c == 'Y' || c == 'y') return N_CHROMOSOMES - 1;
        int id = parseInt(str.size() > 3 ? str.substr(3) : str) - 1;
        if (id < 0 || id >= N_CHROMOSOMES) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: chromosome " + str);
        return id;
    }

    // src/BuildModel.java:666-708
    void loadChromosomeLengths(const std::string& faiFile, const std::string& cytobandFile) {
        chromosomeLengths.assign(N_CHROMOSOMES, 0);
        std::ifstream br(faiFile);
        if (!br) throw std::runtime_error("java.io.FileNotFoundException: " + faiFile);
        std::string str;
        while (readLine(br, str)) {
            auto tokens = split(str, '\t');
            int i = str2chromosomeID(tok(tokens, 0));
            if (i == -1) continue;
            chromosomeLengths[i] = parseInt(tok(tokens, 1));
        }
        svs_per_position.assign(N_SV_TYPES, std::vector<std::vector<double>>(N_CHROMOSOMES));
        for (int i = 0; i < N_SV_TYPES; i++)
            for (int j = 0; j < N_CHROMOSOMES; j++) svs_per_position[i][j].assign((chromosomeLengths[j] + SV_POSITION_QUANTUM) / SV_POSITION_QUANTUM, 0.0);
        chromosomePLengths.assign(N_CHROMOSOMES, 0);
        std::ifstream cr(cytobandFile);
        if (!cr) throw std::runtime_error("java.io.FileNotFoundException: " + cytobandFile);
        while (readLine(cr, str)) {
            auto tokens = split(str, '\t');
            int i = str2chromosomeID(tok(tokens, 0));
            if (i == -1) continue;
            int length = parseInt(tok(tokens, 2));
            if (tok(tokens, 3).empty()) throw std::runtime_error("java.lang.StringIndexOutOfBoundsException: band name");
            if (tokens[3][0] == 'p' && length > chromosomePLengths[i]) chromosomePLengths[i] = length;
        }
    }

    // one moving-average pass, :1028-1039 / :1042-1055 (sum is carried, not recomputed)
    static void smooth(std::vector<double>& a, int radius, int window, std::vector<double>& smoothedArray) {
        const int n = (int)a.size();
        if (n < window + 1) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: array shorter than the smoothing window");
        std::fill(smoothedArray.begin(), smoothedArray.begin() + n, 0.0);
        double sum = 0.0;
        for (int k = 0; k < window; k++) sum += a[k];
        int k;
        for (k = radius; k <= n - radius - 2; k++) {
            smoothedArray[k] = sum / window;
            sum += a[k + radius + 1] - a[k - radius];
        }
        smoothedArray[n - radius - 1] = sum / window;
        for (k = radius; k < n - radius; k++) a[k] = smoothedArray[k];
    }

    // src/BuildModel.java:909-1097
    void collectStatistics(const std::string& gnomadVCFfile, int maxDistance_SR, int maxDistance_PE, int maxDistance_other, int maxDistance_segdups, bool onlyPass) {
        const int nLengthBins = 1 + (MAX_SV_LENGTH - MIN_SV_LENGTH + 1) / SV_LENGTH_QUANTUM;
        statistics_length.assign(N_SV_TYPES, std::vector<std::vector<double>>(N_INTERVAL_TYPES, std::vector<double>(nLengthBins, 0.0)));
        statistics_intervalType.assign(N_SV_TYPES, std::vector<double>(N_INTERVAL_TYPES, 0.0));
        statistics_svType.assign(N_SV_TYPES, 0.0);
        statistics_alleleFrequency.assign(N_AF, std::vector<std::vector<double>>(N_SV_TYPES, std::vector<double>(ALLELE_FREQUENCY_NBINS, 0.0)));
        statistics_lastAlleleFrequency.assign(N_AF, std::vector<int>(N_SV_TYPES, -1));
        statistics_alleleFrequency_minmax.assign(N_AF, std::vector<std::array<double, 2>>(N_SV_TYPES));

        std::ifstream br(gnomadVCFfile);
        if (!br) throw std::runtime_error("java.io.FileNotFoundException: " + gnomadVCFfile);
        std::string str;
        int k = 0, x = 0;
        n_chr1 = 0; nFrequent_chr1 = 0;
        while (readLine(br, str)) {
            if (str.empty()) throw std::runtime_error("java.lang.StringIndexOutOfBoundsException: empty line");
            if (str[0] == '#') continue;
            auto tokens = split(str, '\t');
            if (tokens[0].size() > 5 || eqIgnoreCase(tokens[0], "chrM")) break;
            if (onlyPass && !eqIgnoreCase(tok(tokens, 6), "PASS") && !eqIgnoreCase(tokens[6], "MULTIALLELIC")) continue;
            const std::string& info = tok(tokens, 7);
            int p = indexOf(info, "SVLEN=");
            if (p < 0) continue;
            int q = indexOf(info, ";", p + 6);
            int length = parseInt(substring(info, p + 6, q));
            if (length < 0) length = -length;
            if (length < MIN_SV_LENGTH) continue;
            p = indexOf(info, "SVTYPE=");
            if (p < 0) continue;
            q = indexOf(info, ";", p + 7);
            int svType = jc::binarySearchStr(SV_TYPES, N_SV_TYPES, substring(info, p + 7, q));
            if (svType < 0) continue;
            std::string chromosome = "chr" + tokens[0];
            if (eqIgnoreCase(tokens[0], "1")) n_chr1++;
            int position = parseInt(tok(tokens, 1));
            p = indexOf(info, "EVIDENCE");
            q = indexOf(info, ";", p + 8);
            std::string evidence = substring(info, p + 8, q);
            bool hasSR = indexOf(evidence, "SR") >= 0, hasPE = indexOf(evidence, "PE") >= 0;
            int maxDistance = hasSR ? maxDistance_SR : (hasPE ? maxDistance_PE : maxDistance_other);
            while (k <= lastRepeatInterval && (repeatIntervals_chromosomes[k].compare(chromosome) < 0 || repeatIntervals[k][2] + maxDistance < position)) k++;
            int intervalType;
            if (k > lastRepeatInterval || repeatIntervals_chromosomes[k].compare(chromosome) > 0 || repeatIntervals[k][1] > position + maxDistance) intervalType = 4;
            else {
                intervalType = repeatIntervals[k][0];
                for (int h = k - 1; h >= 0; h--) {
                    if (repeatIntervals[h][2] + maxDistance < position) break;
                    intervalType = ADD_TYPE[intervalType][repeatIntervals[h][0]];
                }
                for (int h = k + 1; h <= lastRepeatInterval; h++) {
                    if (repeatIntervals[h][1] > position + maxDistance) break;
                    intervalType = ADD_TYPE[intervalType][repeatIntervals[h][0]];
                }
            }
            if (intervalType == 4) {
                while (x <= lastSegdup && (segdups_chromosomeIDs[x].compare(chromosome) < 0 || segdups[x][1] + maxDistance_segdups < position)) x++;
                if (x > lastSegdup || segdups_chromosomeIDs[x].compare(chromosome) > 0 || segdups[x][0] > position + maxDistance_segdups) intervalType = 4;
                else intervalType = 3;
            }
            length = (length - MIN_SV_LENGTH) / SV_LENGTH_QUANTUM;
            statistics_length[svType][intervalType][length >= nLengthBins ? nLengthBins - 1 : length]++;
            if (!eqIgnoreCase(tokens[0], "X") && !eqIgnoreCase(tokens[0], "Y")) {
                bool isFrequent = false;
                for (int i = 0; i < N_AF; i++) {
                    p = indexOf(info, ALLELE_FREQUENCY_STR[i]);
                    q = indexOf(info, ";", p + ALLELE_FREQUENCY_STR_LENGTH[i]);
                    double alleleFrequency = parseDouble(substring(info, p + ALLELE_FREQUENCY_STR_LENGTH[i], q));
                    if (eqIgnoreCase(tokens[0], "1") && alleleFrequency >= MIN_FREQUENCY) isFrequent = true;
                    int& lastAF = statistics_lastAlleleFrequency[i][svType];
                    std::vector<double>& arr = statistics_alleleFrequency[i][svType];
                    lastAF++;
                    if (lastAF == (int)arr.size()) arr.resize(arr.size() << 1, 0.0);
                    arr[lastAF] = alleleFrequency;
                }
                if (isFrequent) nFrequent_chr1++;
            }
            int chrID = str2chromosomeID(tokens[0]);
            if (chrID < 0) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: Index -1 out of bounds for length 24");
            std::vector<double>& bins = svs_per_position[svType][chrID];
            if (position < 0 || position / SV_POSITION_QUANTUM >= (int)bins.size()) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: position bin");
            bins[position / SV_POSITION_QUANTUM]++;
        }
        fprintf(stderr, "Total SVs in chr1: %d. With allele frequency >=%s: %d\n", n_chr1, jc::doubleToString(MIN_FREQUENCY).c_str(), nFrequent_chr1);

        // smoothing :1024-1056
        std::vector<double> smoothedArray(nLengthBins);
        for (int i = 0; i < N_SV_TYPES; i++)
            for (int j = 0; j < N_INTERVAL_TYPES; j++) smooth(statistics_length[i][j], SMOOTHING_RADIUS, SMOOTHING_WINDOW, smoothedArray);
        smoothedArray.assign(svs_per_position[0][0].size(), 0.0);
        for (int i = 0; i < N_SV_TYPES; i++)
            for (int j = 0; j < N_CHROMOSOMES; j++) {
                if (svs_per_position[i][j].size() > smoothedArray.size()) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: a chromosome longer than chr1");
                smooth(svs_per_position[i][j], SMOOTHING_RADIUS_POSITIONS, SMOOTHING_WINDOW_POSITIONS, smoothedArray);
            }

        // probabilities :1058-1075
        for (int i = 0; i < N_SV_TYPES; i++)
            for (int j = 0; j < N_INTERVAL_TYPES; j++) {
                double sum = 0.0;
                for (int kk = 0; kk < nLengthBins; kk++) sum += statistics_length[i][j][kk];
                for (int kk = 0; kk < nLengthBins; kk++) statistics_length[i][j][kk] /= sum;
                statistics_intervalType[i][j] = sum;
            }
        for (int i = 0; i < N_SV_TYPES; i++) {
            double sum = 0.0;
            for (int j = 0; j < N_INTERVAL_TYPES; j++) sum += statistics_intervalType[i][j];
            for (int j = 0; j < N_INTERVAL_TYPES; j++) statistics_intervalType[i][j] /= sum;
            statistics_svType[i] = sum;
        }
        {
            double sum = 0.0;
            for (int i = 0; i < N_SV_TYPES; i++) sum += statistics_svType[i];
            for (int i = 0; i < N_SV_TYPES; i++) statistics_svType[i] /= sum;
        }
        // allele frequencies :1076-1094 (the histogram is written over the sorted values, in place, as the Java does)
        for (int h = 0; h < N_AF; h++)
            for (int i = 0; i < N_SV_TYPES; i++) {
                std::vector<double>& arr = statistics_alleleFrequency[h][i];
                const int lastAF = statistics_lastAlleleFrequency[h][i];
                if (lastAF < 0) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: Index -1 (no autosomal record of type " + std::string(SV_TYPES[i]) + ")");
                if (lastAF > 0) std::sort(arr.begin(), arr.begin() + lastAF + 1);
                double min = arr[0], max = arr[lastAF];
                double quantum = (max - min) / (ALLELE_FREQUENCY_NBINS - 1);
                statistics_alleleFrequency_minmax[h][i] = {min, max};
                int kk = 0;
                for (int j = 0; j < ALLELE_FREQUENCY_NBINS; j++) {
                    double boundary = min + j * quantum, sum = 0.0;
                    while (kk <= lastAF && arr[kk] < boundary) { sum++; kk++; }
                    arr[j] = sum / (lastAF + 1);
                }
            }
        buildStatisticsPosition();
    }

    // src/BuildModel.java:740-802
    void buildStatisticsPosition() {
        const int Q = SV_POSITION_QUANTUM;
        statistics_position.assign(N_SV_TYPES, std::vector<double>(svs_per_position[0][0].size(), 0.0));
        for (int i = 0; i < N_SV_TYPES; i++) {
            std::vector<double>& out = statistics_position[i];
            auto add = [&](int bin, double v) {
                if (bin < 0 || bin >= (int)out.size()) throw std::runtime_error("java.lang.ArrayIndexOutOfBoundsException: projected bin");
                out[bin] += v;
            };
            std::copy(svs_per_position[i][0].begin(), svs_per_position[i][0].end(), out.begin());
            for (int j = 1; j < N_AUTOSOMES; j++) {
                double ratioP = ((double)chromosomePLengths[0]) / chromosomePLengths[j];
                double ratioQ = ((double)(chromosomeLengths[0] - chromosomePLengths[0])) / (chromosomeLengths[j] - chromosomePLengths[j]);
                int length = (int)svs_per_position[i][j].size();
                for (int k = 0; k < length; k++) {
                    int first = k * Q, last = std::min((k + 1) * Q, chromosomeLengths[j]) - 1;
                    int firstP, lastP, firstQ, lastQ;
                    double massP, massQ;
                    const double c = svs_per_position[i][j][k];
                    if (last < chromosomePLengths[j]) { firstP = first; lastP = last; firstQ = -1; lastQ = -1; massP = c; massQ = 0.0; }
                    else if (first < chromosomePLengths[j] && last >= chromosomePLengths[j]) {
                        firstP = first; lastP = chromosomePLengths[j] - 1; firstQ = chromosomePLengths[j]; lastQ = last;
                        massP = ((double)((lastP - firstP + 1) * c)) / Q;
                        massQ = ((double)((lastQ - firstQ + 1) * c)) / Q;
                    } else { firstP = -1; lastP = -1; firstQ = first; lastQ = last; massP = 0.0; massQ = c; }
                    int projectedFirst, projectedLast, projectedFirstBin, projectedLastBin;
                    double projectedMassPerBp;
                    if (firstP != -1) {
                        projectedFirst = (int)(firstP * ratioP); projectedLast = (int)(lastP * ratioP);
                        if (projectedFirst < chromosomePLengths[0]) {
                            projectedMassPerBp = massP / (projectedLast - projectedFirst + 1);
                            projectedLast = std::min(projectedLast, chromosomePLengths[0] - 1);
                            projectedFirstBin = projectedFirst / Q; projectedLastBin = projectedLast / Q;
                            if (projectedLastBin == projectedFirstBin) add(projectedFirstBin, projectedMassPerBp * (projectedLast - projectedFirst + 1));
                            else {
                                add(projectedFirstBin, projectedMassPerBp * ((projectedFirstBin + 1) * Q - projectedFirst));
                                for (int h = projectedFirstBin + 1; h < projectedLastBin; h++) add(h, projectedMassPerBp * Q);
                                add(projectedLastBin, projectedMassPerBp * (projectedLast + 1 - projectedLastBin * Q));
                            }
                        }
                    }
                    if (firstQ != -1) {
                        projectedFirst = chromosomePLengths[0] + (int)((firstQ - chromosomePLengths[j]) * ratioQ);
                        projectedLast = chromosomePLengths[0] + (int)((lastQ - chromosomePLengths[j]) * ratioQ);
                        projectedMassPerBp = massQ / (projectedLast - projectedFirst + 1);
                        projectedLast = std::min(projectedLast, chromosomeLengths[0] - 1);
                        projectedFirstBin = projectedFirst / Q; projectedLastBin = projectedLast / Q;
                        if (projectedLastBin == projectedFirstBin) add(projectedFirstBin, projectedMassPerBp * (projectedLast - projectedFirst + 1));
                        else {
                            add(projectedFirstBin, projectedMassPerBp * ((projectedFirstBin + 1) * Q - projectedFirst));
                            for (int h = projectedFirstBin + 1; h < projectedLastBin; h++) add(h, projectedMassPerBp * Q);
                            add(projectedLastBin, projectedMassPerBp * (projectedLast + 1 - projectedLastBin * Q));
                        }
                    }
                }
            }
        }
        for (int i = 0; i < N_SV_TYPES; i++) {
            double sum = 0.0;
            for (double v : statistics_position[i]) sum += v;
            for (double& v : statistics_position[i]) v /= sum;
        }
    }

    // src/BuildModel.java:1100-1132
    void makeStatisticsCumulative() {
        for (int i = 1; i < N_SV_TYPES; i++) statistics_svType[i] += statistics_svType[i - 1];
        for (auto& row : statistics_intervalType) for (size_t j = 1; j < row.size(); j++) row[j] += row[j - 1];
        for (auto& a : statistics_length) for (auto& row : a) for (size_t k = 1; k < row.size(); k++) row[k] += row[k - 1];
        for (auto& a : statistics_alleleFrequency) for (auto& row : a) for (size_t k = 1; k < row.size(); k++) row[k] += row[k - 1];
        for (auto& row : statistics_position) for (size_t j = 1; j < row.size(); j++) row[j] += row[j - 1];
    }

    static void put(FILE* f, double d) { std::string s = jc::doubleToString(d); fwrite(s.data(), 1, s.size(), f); fputc(',', f); }
    static FILE* open(const std::string& path) {
        FILE* f = fopen(path.c_str(), "wb");
        if (!f) throw std::runtime_error("java.io.FileNotFoundException: " + path);
        return f;
    }

    // src/BuildModel.java:1135-1197
    void serializeStatistics(const std::string& outputDir) {
        FILE* bw = open(outputDir + "/statistics_svType.txt");
        fprintf(bw, "%d\n", N_SV_TYPES);
        for (double d : statistics_svType) put(bw, d);
        fputc('\n', bw); fclose(bw);
        bw = open(outputDir + "/statistics_intervalType.txt");
        fprintf(bw, "%d,%d\n", N_SV_TYPES, N_INTERVAL_TYPES);
        for (auto& row : statistics_intervalType) { for (double d : row) put(bw, d); fputc('\n', bw); }
        fclose(bw);
        bw = open(outputDir + "/statistics_length.txt");
        fprintf(bw, "%d,%d,%d\n", N_SV_TYPES, N_INTERVAL_TYPES, (int)statistics_length[0][0].size());
        for (auto& a : statistics_length) for (auto& row : a) { for (double d : row) put(bw, d); fputc('\n', bw); }
        fclose(bw);
        for (int k = 0; k < N_AF; k++) {
            bw = open(outputDir + "/" + AF_FILE_NAMES[k] + ".txt");
            fprintf(bw, "%d,%d\n", N_SV_TYPES, ALLELE_FREQUENCY_NBINS);
            for (int i = 0; i < N_SV_TYPES; i++) {
                put(bw, statistics_alleleFrequency_minmax[k][i][0]); put(bw, statistics_alleleFrequency_minmax[k][i][1]);
                for (int j = 0; j < ALLELE_FREQUENCY_NBINS; j++) put(bw, statistics_alleleFrequency[k][i][j]);
                fputc('\n', bw);
            }
            fclose(bw);
        }
        bw = open(outputDir + "/statistics_position.txt");
        fprintf(bw, "%d,%d\n", N_SV_TYPES, (int)statistics_position[0].size());
        for (auto& row : statistics_position) { for (double d : row) put(bw, d); fputc('\n', bw); }
        fclose(bw);
    }

    // src/BuildModel.java:446-458
    void serializeRepeatIntervals(const std::string& outFile) {
        FILE* bw = open(outFile);
        int i;
        for (i = 0; i <= lastRepeatInterval; i++) if (!eqIgnoreCase(repeatIntervals_chromosomes[i], "chr1")) break;
        fprintf(bw, "%d\n", i);
        for (i = 0; i <= lastRepeatInterval; i++) {
            if (!eqIgnoreCase(repeatIntervals_chromosomes[i], "chr1")) break;
            fprintf(bw, "%d,%d,%d\n", repeatIntervals[i][0], repeatIntervals[i][1], repeatIntervals[i][2]);
        }
        fclose(bw);
    }

    // src/BuildModel.java:563-577
    void serializeSegdups(const std::string& outFile) {
        FILE* bw = open(outFile);
        int i;
        for (i = 0; i <= lastSegdup; i++) if (!eqIgnoreCase(segdups_chromosomeIDs[i], "chr1")) break;
        fprintf(bw, "%d\n", i);
        for (i = 0; i <= lastSegdup; i++) {
            if (!eqIgnoreCase(segdups_chromosomeIDs[i], "chr1")) break;
            fprintf(bw, "%d,%d\n", segdups[i][0], segdups[i][1]);
        }
        fclose(bw);
    }
};

}  // namespace

// src/BuildModel.java:21-47. args as the Java main's argv.
int buildModelMain(const std::vector<std::string>& args) {
    if (args.size() < 12) throw std::runtime_error("usage: BuildModel repeatmasker nrows trf nrows segdups nrows reference.fai cytobands gnomad.vcf onlyPass outputDir verbose");
    Model m;
    m.loadRepeatMaskerAnnotations(args[0], parseInt(args[1]), MAX_DISTANCE_REPEATMASKER);
    m.loadTrfAnnotations(args[2], parseInt(args[3]), MIN_OVERLAP_TRF);
    m.buildRepeatIntervals(MIN_OVERLAP_INTERVALS);
    m.loadSegdups(args[4], parseInt(args[5]), MAX_DISTANCE_SEGDUPS);
    m.loadChromosomeLengths(args[6], args[7]);
    m.collectStatistics(args[8], MAX_DISTANCE_SR, MAX_DISTANCE_PE, MAX_DISTANCE_OTHER, MAX_DISTANCE_SEGDUPS, parseInt(args[9]) == 1);
    m.makeStatisticsCumulative();
    m.serializeStatistics(args[10]);
    m.serializeRepeatIntervals(args[10] + "/repeatIntervals.txt");
    m.serializeSegdups(args[10] + "/segdups.txt");
    return 0;
}

}  // namespace oracle
