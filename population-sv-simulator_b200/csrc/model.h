// Host-side model of the population sampler: BuildModel statistics, repeat/segdup tracks, and the five
// repeat-context position sets of SimulateHaplotypes.loadPositions held as sorted disjoint intervals
// (never materialised as int[5][L] like the reference, src/SimulateHaplotypes.java:143-218).
#pragma once
#include <array>
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

namespace sp {

constexpr int N_SV_TYPES = 4;
constexpr int N_INTERVAL_TYPES = 5;        // src/BuildModel.java:854
constexpr int MAX_DISTANCE_SR = 25;        // src/BuildModel.java:15
constexpr int MIN_SV_LENGTH = 50;          // src/BuildModel.java:838
constexpr int SV_LENGTH_QUANTUM = 50;      // src/BuildModel.java:840
constexpr int ALLELE_FREQUENCY_NBINS = 10000;  // src/BuildModel.java:841
extern const char* const SV_TYPE_NAMES[4];
extern const char* const AF_NAMES[6];
extern const int ADD_TYPE[3][3];           // src/BuildModel.java:83

struct Model {
    std::vector<double> svType;                               // cumulative
    std::vector<std::vector<double>> intervalType;            // [svType][5] cumulative
    std::vector<std::vector<std::vector<double>>> length;     // [svType][intervalType][bins] cumulative
    std::vector<std::vector<std::vector<double>>> af;         // [population][svType][bins] cumulative
    std::vector<std::vector<std::array<double, 2>>> af_minmax;  // [population][svType]
    std::vector<std::array<int32_t, 3>> repeats;              // type, start, end
    std::vector<std::array<int32_t, 2>> segdups;              // start, end

    void load_statistics(const std::string& dir);             // BuildModel.deserializeStatistics :1200-1282
    void load_repeat_intervals(const std::string& path);      // BuildModel.deserializeRepeatIntervals :464-480
    void load_segdups(const std::string& path);               // BuildModel.deserializeSegdups :581-598
};

// Sorted, disjoint, non-adjacent closed intervals [first,last] over int32 positions.
struct IntervalSet {
    std::vector<std::pair<int32_t, int32_t>> iv;
    std::vector<int64_t> before;  // before[i] = number of positions in iv[0..i)
    int64_t count() const { return before.empty() ? 0 : before.back(); }
    void normalise();             // sort + merge (also adjacent), rebuild prefix counts
    void finalise();              // prefix counts only
    int32_t kth(int64_t k) const; // k-th smallest position (0-based)
    static IntervalSet unite(const IntervalSet& a, const IntervalSet& b);
    static IntervalSet intersect(const IntervalSet& a, const IntervalSet& b);
    static IntervalSet minus(const IntervalSet& a, const IntervalSet& b);
};

// positions[0..4] of the reference, after the REP/SAT/BOTH/SEGDUP/NONE algebra and the ACGT filter.
struct PositionSets {
    IntervalSet set[N_INTERVAL_TYPES];
    void build(const Model& m, const uint8_t* reference, int32_t reference_length, int max_distance);
};

}  // namespace sp
