// ORACLE — TEST INFRASTRUCTURE ONLY. See simulate_haplotypes.h.
// Variant sampling, population building and collision fixing, insertion strings, repeat annotation.
#include "simulate_haplotypes.h"
#include <array>

namespace oracle {

// src/SimulateHaplotypes.java:313-337
void SimulateHaplotypes::loadInsertions(const std::string& path, int nInsertions, int quantum) {
    insertions.assign((size_t)nInsertions, std::string());
    jc::LineReader br(path);
    std::string str;
    lastInsertion = -1;
    while (br.readLine(str)) {
        if (++lastInsertion >= nInsertions) throw std::runtime_error("ArrayIndexOutOfBoundsException: insertions");
        insertions[lastInsertion] = jc::trim(str);
    }
    if (lastInsertion < 0) throw std::runtime_error("ArrayIndexOutOfBoundsException: insertions[-1]");
    int nBins = 1 + (int)insertions[lastInsertion].size() / quantum;
    firstInsertionWithLength.assign(nBins, -1);
    for (int i = 0; i <= lastInsertion; i++) {
        int j = (int)insertions[i].size() / quantum;
        if (j >= nBins) throw std::runtime_error("ArrayIndexOutOfBoundsException: firstInsertionWithLength");
        if (firstInsertionWithLength[j] == -1) firstInsertionWithLength[j] = i;
    }
}

// src/SimulateHaplotypes.java:353-374
void SimulateHaplotypes::sampleVariants(double alleleFrequencyThreshold, int nFrequent, int populationID) {
    variants.clear();
    nVariants_frequent = 0; nVariants_rare = 0; lastVariant = -1;
    while (nVariants_frequent < nFrequent) {
        lastVariant++;
        Variant variant;
        do { variant = sampleVariant(populationID); drawLog.candidates++; }
        while (variant.svType != 2 && variant.position + variant.length > referenceLength);
        variants.push_back(variant);
        if (variants[lastVariant].alleleFrequency >= alleleFrequencyThreshold) nVariants_frequent++;
        else nVariants_rare++;
    }
    // Arrays.sort(Object[]) is a stable merge sort (TimSort); Variant.compareTo orders by position only.
    std::stable_sort(variants.begin(), variants.end(),
                     [](const Variant& a, const Variant& b) { return a.position < b.position; });
    positions.clear(); positions.shrink_to_fit(); positions_last.clear();
}

// src/SimulateHaplotypes.java:384-400
Variant SimulateHaplotypes::sampleVariant(int populationID) {
    Variant out;
    auto idx = [](int r) { return r < 0 ? -1 - r : r; };
    if (populationID < 0 || populationID >= 6) throw std::runtime_error("ArrayIndexOutOfBoundsException: populationID");
    out.svType = idx(jc::binarySearch(model.statistics_svType.data(), 0, (int)model.statistics_svType.size(), random.nextDouble()));
    if (out.svType >= (int)model.statistics_intervalType.size()) throw std::runtime_error("ArrayIndexOutOfBoundsException: svType");
    const auto& it = model.statistics_intervalType[out.svType];
    out.intervalType = idx(jc::binarySearch(it.data(), 0, (int)it.size(), random.nextDouble()));
    if (out.intervalType >= (int)model.statistics_length[out.svType].size()) throw std::runtime_error("ArrayIndexOutOfBoundsException: intervalType");
    const auto& sl = model.statistics_length[out.svType][out.intervalType];
    int p = idx(jc::binarySearch(sl.data(), 0, (int)sl.size(), random.nextDouble()));
    out.length = BuildModel::MIN_SV_LENGTH + BuildModel::SV_LENGTH_QUANTUM * p + random.nextInt(BuildModel::SV_LENGTH_QUANTUM);
    const double min = model.statistics_alleleFrequency_minmax[populationID][out.svType][0];
    const double quantum = (model.statistics_alleleFrequency_minmax[populationID][out.svType][1] - min) / (BuildModel::ALLELE_FREQUENCY_NBINS - 1);
    const auto& af = model.statistics_alleleFrequency[populationID][out.svType];
    if ((int)af.size() < BuildModel::ALLELE_FREQUENCY_NBINS) throw std::runtime_error("ArrayIndexOutOfBoundsException: alleleFrequency");
    p = idx(jc::binarySearch(af.data(), 0, BuildModel::ALLELE_FREQUENCY_NBINS, random.nextDouble()));
    // min+quantum*p+random.nextDouble()*quantum : left-to-right, strict IEEE double (-ffp-contract=off)
    volatile double t1 = quantum * p;
    volatile double t2 = min + t1;
    volatile double t3 = random.nextDouble() * quantum;
    out.alleleFrequency = t2 + t3;
    if (out.intervalType >= (int)positions.size()) throw std::runtime_error("ArrayIndexOutOfBoundsException: positions");
    out.position = positions[out.intervalType][random.nextInt(positions_last[out.intervalType] + 1)];
    return out;
}

// src/SimulateHaplotypes.java:412-432
void SimulateHaplotypes::annotateVariantsWithRepeats(int maxDistance) {
    int tmpArray1[2] = {0, 0}, tmpArray2[2] = {0, 0}, tmpArray3[2] = {0, 0};
    for (int i = 0; i <= lastVariant; i++) {
        variants[i].intervalType_start = getIntervalType(variants[i].position, maxDistance, tmpArray1, true);
        if (variants[i].svType == 2) {
            variants[i].intervalType_end = variants[i].intervalType_start;
            variants[i].intervalSurface = 0;
        } else {
            int last = variants[i].position + variants[i].length - 1;
            variants[i].intervalType_end = getIntervalType(last, maxDistance, tmpArray1, false);
            variants[i].intervalSurface = getIntervalSurface(variants[i].position, last, tmpArray2, tmpArray3);
        }
    }
}

// src/SimulateHaplotypes.java:442-473
int SimulateHaplotypes::getIntervalType(int position, int maxDistance, int* cursors, bool updateCursors) {
    const auto& R = model.repeatIntervals; const int lastR = model.lastRepeatInterval;
    const auto& S = model.segdups; const int lastS = model.lastSegdup;
    int i1 = cursors[0], i2 = cursors[1], intervalType;
    while (i1 <= lastR && R[i1][2] + maxDistance < position) i1++;
    if (i1 > lastR || R[i1][1] > position + maxDistance) intervalType = 4;
    else {
        intervalType = R[i1][0];
        for (int j = i1 - 1; j >= 0; j--) {
            if (R[j][2] + maxDistance < position) break;
            intervalType = BuildModel::ADD_TYPE[intervalType][R[j][0]];
        }
        for (int j = i1 + 1; j <= lastR; j++) {
            if (R[j][1] > position + maxDistance) break;
            intervalType = BuildModel::ADD_TYPE[intervalType][R[j][0]];
        }
    }
    if (intervalType == 4) {
        while (i2 <= lastS && S[i2][1] + maxDistance < position) i2++;
        if (i2 > lastS || S[i2][0] > position + maxDistance) intervalType = 4;
        else intervalType = 3;
    }
    if (updateCursors) { cursors[0] = i1; cursors[1] = i2; }
    return intervalType;
}

// src/SimulateHaplotypes.java:484-515 (stale tmpPairs[0] when nothing intersects: :513)
int SimulateHaplotypes::getIntervalSurface(int first, int last, int* cursors, int* tmpArray) {
    if (tmpPairs.empty()) tmpPairs.assign(100, Pair());
    tmpArray[0] = cursors[0]; tmpArray[1] = -1;
    getIntersectingIntervals(first, last, model.repeatIntervals, model.lastRepeatInterval, 1, 2, tmpArray);
    cursors[0] = tmpArray[0];
    tmpArray[0] = cursors[1];
    getIntersectingIntervals(first, last, model.segdups, model.lastSegdup, 0, 1, tmpArray);
    cursors[1] = tmpArray[0];
    int lastPair = tmpArray[1];
    if (lastPair != 0) {
        // Arrays.sort(tmpPairs,0,lastPair+1): fromIndex>toIndex throws only when lastPair+1<0, never here.
        if (lastPair + 1 > 0)
            std::stable_sort(tmpPairs.begin(), tmpPairs.begin() + (lastPair + 1), [](const Pair& a, const Pair& b) {
                return a.first < b.first || (a.first == b.first && a.last < b.last);
            });
    }
    int j = 0, sum = 0;
    for (int i = 1; i <= lastPair; i++) {
        if (tmpPairs[i].first <= tmpPairs[j].last) tmpPairs[j].last = std::max(tmpPairs[j].last, tmpPairs[i].last);
        else {
            sum += tmpPairs[j].last - tmpPairs[j].first + 1;
            j++;
            tmpPairs[j].first = tmpPairs[i].first;
            tmpPairs[j].last = tmpPairs[i].last;
        }
    }
    sum += tmpPairs[j].last - tmpPairs[j].first + 1;
    return sum;
}

// src/SimulateHaplotypes.java:527-571, including the grow path that clobbers loop variable k
// (:537-542, :548-553, :560-565).
template <class M>
void SimulateHaplotypes::getIntersectingIntervals(int first, int last, const M& matrix, int lastRow,
                                                  int column_first, int column_last, int* cursors) {
    int i = cursors[0], j = cursors[1], k;
    auto row = [&](int r) -> const auto& {
        if (r < 0 || r >= (int)matrix.size()) throw std::runtime_error("ArrayIndexOutOfBoundsException: interval row " + std::to_string(r));
        return matrix[r];
    };
    auto grow = [&]() -> int {
        size_t old = tmpPairs.size();
        tmpPairs.resize(old << 1);
        return (int)tmpPairs.size();  // value of k after "for (k=old; k<newLength; k++)"
    };
    while (i <= lastRow && row(i)[column_last] < first) i++;
    if (i <= lastRow && row(i)[column_first] <= last) {
        j++;
        if (j == (int)tmpPairs.size()) k = grow();
        tmpPairs[j].first = std::max(first, row(i)[column_first]);
        tmpPairs[j].last = std::min(last, row(i)[column_last]);
        for (k = i - 1; k >= 0; k--) {
            if (row(k)[column_last] < first) break;
            j++;
            if (j == (int)tmpPairs.size()) k = grow();
            tmpPairs[j].first = std::max(first, row(k)[column_first]);
            tmpPairs[j].last = std::min(last, row(k)[column_last]);
        }
        for (k = i + 1; k <= lastRow; k++) {
            if (row(k)[column_first] > last) break;
            j++;
            if (j == (int)tmpPairs.size()) k = grow();
            tmpPairs[j].first = std::max(first, row(k)[column_first]);
            tmpPairs[j].last = std::min(last, row(k)[column_last]);
        }
    }
    cursors[0] = i; cursors[1] = j;
}
template void SimulateHaplotypes::getIntersectingIntervals(int, int, const std::vector<std::array<int, 3>>&, int, int, int, int*);
template void SimulateHaplotypes::getIntersectingIntervals(int, int, const std::vector<std::array<int, 2>>&, int, int, int, int*);

}  // namespace oracle
