// ORACLE — TEST INFRASTRUCTURE ONLY (see java_compat.h). CPU restatement of the reference's
// src/SimulateHaplotypes.java, src/PrintHaplotypes.java and the loader half of src/BuildModel.java.
// Same control flow, same quirks (SURVEY.md "Bug compatibility"); parity unpinned by the reference
// (it has no tests); JDK primitives pinned by KATs.
#pragma once
#include "java_compat.h"
#include <memory>

namespace oracle {

// --- BuildModel.java loader half -------------------------------------------------------------
struct BuildModel {
    // constants: src/BuildModel.java:12-18, :83, :837-854
    static constexpr int MAX_DISTANCE_SR = 25;
    static constexpr int MIN_SV_LENGTH = 50;
    static constexpr int SV_LENGTH_QUANTUM = 50;
    static constexpr int ALLELE_FREQUENCY_NBINS = 10000;
    static constexpr int N_INTERVAL_TYPES = 5;
    static const char* const SV_TYPES[4];
    static const char* const AF_NAMES[6];
    static const char* const AF_FILE_NAMES[6];
    static const int ADD_TYPE[3][3];

    std::vector<double> statistics_svType;
    std::vector<std::vector<double>> statistics_intervalType;
    std::vector<std::vector<std::vector<double>>> statistics_length;
    std::vector<std::vector<std::vector<double>>> statistics_alleleFrequency;         // [pop][svType][bin]
    std::vector<std::vector<std::vector<double>>> statistics_alleleFrequency_minmax;  // [pop][svType][2]
    std::vector<std::vector<double>> statistics_position;
    std::vector<std::array<int, 3>> repeatIntervals;  // 0=type, 1=start, 2=end
    int lastRepeatInterval = -1;
    std::vector<std::array<int, 2>> segdups;
    int lastSegdup = -1;

    void deserializeStatistics(const std::string& inputDir);       // src/BuildModel.java:1200-1282
    void deserializeRepeatIntervals(const std::string& inFile);    // src/BuildModel.java:464-480
    void deserializeSegdups(const std::string& inFile);            // src/BuildModel.java:581-598
};

struct Variant {  // src/SimulateHaplotypes.java:1181-1198
    int svType = 0, length = 0;
    double alleleFrequency = 0;
    int position = 0;
    int intervalType = 0;
    bool hasSequence = false;
    std::string sequence;
    bool isActive = false;
    int intervalType_start = 0, intervalType_end = 0;
    int intervalSurface = 0;
};

struct Pair { int first = 0, last = 0; };

// Every draw the reference's single java.util.Random makes, logged for replay-mode tests.
struct DrawLog {
    uint64_t steps_after_sampleVariants = 0;
    uint64_t steps_after_buildPopulation = 0;
    uint64_t steps_after_assignInsertionStrings = 0;
    int candidates = 0;  // sampleVariant() calls, including those discarded by the do/while
};

struct SimulateHaplotypes {
    BuildModel model;
    std::string reference;
    int referenceLength = 0;
    std::vector<std::vector<int>> positions;
    std::vector<int> positions_last;
    bool positionsAllocated = false;
    std::vector<std::string> insertions;
    int lastInsertion = -1;
    std::vector<int> firstInsertionWithLength;
    std::vector<Variant> variants;
    int lastVariant = -1, nVariants_frequent = 0, nVariants_rare = 0;
    int N_HAPLOTYPES = 0;
    std::vector<std::vector<int>> haplotype2variants;
    std::vector<int> haplotype2variants_last;
    std::vector<std::vector<int>> variant2haplotypes;
    std::vector<int> variant2haplotypes_last;
    jc::JavaRandom random;
    std::vector<Pair> tmpPairs;  // null until first use (size()==0)
    DrawLog drawLog;
    bool verbose = true;
    // Optional dump of the genotype matrix before collision fixing (row G of SURVEY.md §8a), one
    // line per haplotype, for kernel-level parity tests.
    std::string dumpPreCollisionFile;
    std::vector<std::string> phaseNames;   // mainSimulate's wall-clock seconds per phase (reporting only)
    std::vector<double> phaseSeconds;

    explicit SimulateHaplotypes(int64_t seed) : random(seed) {}

    int mainSimulate(const std::vector<std::string>& args);  // :75-116
    int mainPrint(const std::vector<std::string>& args);     // src/PrintHaplotypes.java:6-18

    void loadReference(const std::string& path);                         // :119-134
    void loadPositions(int chromosomeLength, int maxDistance);           // :143-218
    void loadInsertions(const std::string& path, int nInsertions, int quantum);  // :313-337
    void sampleVariants(double alleleFrequencyThreshold, int nFrequent, int populationID);  // :353-374
    Variant sampleVariant(int populationID);                             // :384-400
    void annotateVariantsWithRepeats(int maxDistance);                   // :412-432
    int getIntervalType(int position, int maxDistance, int* cursors, bool updateCursors);  // :442-473
    int getIntervalSurface(int first, int last, int* cursors, int* tmpArray);              // :484-515
    template <class M> void getIntersectingIntervals(int first, int last, const M& matrix, int lastRow,
                                                     int column_first, int column_last, int* cursors);  // :527-571
    void buildPopulation(int nIndividuals, double alleleFrequencyThreshold, const std::string& outputDir);  // :603-718
    void assignInsertionStrings();                                       // :786-814
    int getInsertionString(int length);                                  // :827-875
    static void reverseComplement(std::string& buffer);                  // :878-899
    void buildHaplotypes(int firstHaplotype, int lastHaplotype, const std::string& outDir);  // :917-960
    void buildHaplotypeString(int i, std::string& out);                  // body of the loop at :926-953
    void buildVariant2haplotypes();                                      // :967-987
    void buildVCF(const std::string& outDir);                            // :1005-1067
    void printVcfHeader(bool printGenotype, const std::string& sampleName, FILE* bw);  // :1077-1096
    void writeVCF(const Variant& v, int id, FILE* bw);                   // :1243-1279
    void printPositionHistogram(const std::string& outputFile, int binSize);  // :1104-1120
    void serialize(const std::string& h2vFile, const std::string& variantsFile);    // :1123-1144
    void deserialize(const std::string& h2vFile, const std::string& variantsFile);  // :1147-1178
};

}  // namespace oracle
