// ORACLE — TEST INFRASTRUCTURE ONLY. See simulate_haplotypes.h.
#include "simulate_haplotypes.h"
#include <array>

namespace oracle {

const char* const BuildModel::SV_TYPES[4] = {"DEL", "DUP", "INS", "INV"};
const char* const BuildModel::AF_NAMES[6] = {"ALL", "AFR", "AMR", "EAS", "EUR", "OTH"};
const char* const BuildModel::AF_FILE_NAMES[6] = {
    "statistics_alleleFrequency_ALL", "statistics_alleleFrequency_AFR", "statistics_alleleFrequency_AMR",
    "statistics_alleleFrequency_EAS", "statistics_alleleFrequency_EUR", "statistics_alleleFrequency_OTH"};
const int BuildModel::ADD_TYPE[3][3] = {{0, 2, 2}, {2, 1, 2}, {2, 2, 2}};

static std::string mustReadLine(jc::LineReader& br) {
    std::string s;
    if (!br.readLine(s)) throw std::runtime_error("NullPointerException: unexpected end of file");
    return s;
}
static const std::string& tok(const std::vector<std::string>& t, size_t i) {
    if (i >= t.size()) throw std::runtime_error("ArrayIndexOutOfBoundsException: token " + std::to_string(i));
    return t[i];
}

// src/BuildModel.java:1200-1282
void BuildModel::deserializeStatistics(const std::string& inputDir) {
    int lengthI, lengthJ, lengthK;
    std::string str;
    std::vector<std::string> tokens;
    {
        jc::LineReader br(inputDir + "/statistics_svType.txt");
        lengthI = jc::parseInt(mustReadLine(br));
        statistics_svType.assign(lengthI, 0.0);
        tokens = jc::splitComma(mustReadLine(br));
        for (int i = 0; i < lengthI; i++) statistics_svType[i] = jc::parseDouble(tok(tokens, i));
    }
    {
        jc::LineReader br(inputDir + "/statistics_intervalType.txt");
        tokens = jc::splitComma(mustReadLine(br));
        lengthI = jc::parseInt(tok(tokens, 0)); lengthJ = jc::parseInt(tok(tokens, 1));
        statistics_intervalType.assign(lengthI, std::vector<double>(lengthJ, 0.0));
        for (int i = 0; i < lengthI; i++) {
            tokens = jc::splitComma(mustReadLine(br));
            for (int j = 0; j < lengthJ; j++) statistics_intervalType[i][j] = jc::parseDouble(tok(tokens, j));
        }
    }
    {
        jc::LineReader br(inputDir + "/statistics_length.txt");
        tokens = jc::splitComma(mustReadLine(br));
        lengthI = jc::parseInt(tok(tokens, 0)); lengthJ = jc::parseInt(tok(tokens, 1)); lengthK = jc::parseInt(tok(tokens, 2));
        statistics_length.assign(lengthI, std::vector<std::vector<double>>(lengthJ, std::vector<double>(lengthK, 0.0)));
        for (int i = 0; i < lengthI; i++)
            for (int j = 0; j < lengthJ; j++) {
                tokens = jc::splitComma(mustReadLine(br));
                for (int k = 0; k < lengthK; k++) statistics_length[i][j][k] = jc::parseDouble(tok(tokens, k));
            }
    }
    statistics_alleleFrequency.assign(6, {});
    statistics_alleleFrequency_minmax.assign(6, {});
    for (int k = 0; k < 6; k++) {
        jc::LineReader br(inputDir + "/" + AF_FILE_NAMES[k] + ".txt");
        tokens = jc::splitComma(mustReadLine(br));
        lengthI = jc::parseInt(tok(tokens, 0)); lengthJ = jc::parseInt(tok(tokens, 1));
        statistics_alleleFrequency[k].assign(lengthI, std::vector<double>(lengthJ, 0.0));
        statistics_alleleFrequency_minmax[k].assign(lengthI, std::vector<double>(2, 0.0));
        for (int i = 0; i < lengthI; i++) {
            tokens = jc::splitComma(mustReadLine(br));
            statistics_alleleFrequency_minmax[k][i][0] = jc::parseDouble(tok(tokens, 0));
            statistics_alleleFrequency_minmax[k][i][1] = jc::parseDouble(tok(tokens, 1));
            for (int j = 0; j < lengthJ; j++) statistics_alleleFrequency[k][i][j] = jc::parseDouble(tok(tokens, 2 + j));
        }
    }
    {
        jc::LineReader br(inputDir + "/statistics_position.txt");
        tokens = jc::splitComma(mustReadLine(br));
        lengthI = jc::parseInt(tok(tokens, 0)); lengthJ = jc::parseInt(tok(tokens, 1));
        statistics_position.assign(lengthI, std::vector<double>(lengthJ, 0.0));
        for (int i = 0; i < lengthI; i++) {
            tokens = jc::splitComma(mustReadLine(br));
            for (int j = 0; j < lengthJ; j++) statistics_position[i][j] = jc::parseDouble(tok(tokens, j));
        }
    }
}

// src/BuildModel.java:464-480
void BuildModel::deserializeRepeatIntervals(const std::string& inFile) {
    jc::LineReader br(inFile);
    lastRepeatInterval = jc::parseInt(mustReadLine(br)) - 1;
    repeatIntervals.assign(lastRepeatInterval + 1, {0, 0, 0});
    for (int i = 0; i <= lastRepeatInterval; i++) {
        auto tokens = jc::splitComma(mustReadLine(br));
        for (int j = 0; j < 3; j++) repeatIntervals[i][j] = jc::parseInt(tok(tokens, j));
    }
}

// src/BuildModel.java:581-598
void BuildModel::deserializeSegdups(const std::string& inFile) {
    jc::LineReader br(inFile);
    lastSegdup = jc::parseInt(mustReadLine(br)) - 1;
    segdups.assign(lastSegdup + 1, {0, 0});
    for (int i = 0; i <= lastSegdup; i++) {
        auto tokens = jc::splitComma(mustReadLine(br));
        segdups[i][0] = jc::parseInt(tok(tokens, 0));
        segdups[i][1] = jc::parseInt(tok(tokens, 1));
    }
}

// src/SimulateHaplotypes.java:119-134
void SimulateHaplotypes::loadReference(const std::string& path) {
    jc::LineReader br(path);
    std::string str;
    reference.clear();
    reference.reserve(250000000 / 64);  // (the Java pre-sizes 250 M chars; capacity is not observable)
    br.readLine(str);  // Skipping header
    while (br.readLine(str)) reference += jc::trim(str);
    referenceLength = (int)reference.size();
    if (verbose) fprintf(stderr, "Length of the reference: %d\n", referenceLength);
}

// src/SimulateHaplotypes.java:227-304
static int setUnion(const int* from1, int last1, const int* from2, int last2, int* to) {
    int i1 = 0, i2 = 0, j = 0;
    while (i1 <= last1 && i2 <= last2) {
        if (from1[i1] < from2[i2]) to[j++] = from1[i1++];
        else if (from1[i1] > from2[i2]) to[j++] = from2[i2++];
        else { to[j++] = from1[i1]; i1++; i2++; }
    }
    while (i1 <= last1) to[j++] = from1[i1++];
    while (i2 <= last2) to[j++] = from2[i2++];
    return j - 1;
}
static int setMinus(const int* from1, int last1, const int* from2, int last2, int* to) {
    int i1 = 0, i2 = 0, j = 0;
    while (i1 <= last1 && i2 <= last2) {
        if (from1[i1] < from2[i2]) to[j++] = from1[i1++];
        else if (from1[i1] > from2[i2]) i2++;
        else { i1++; i2++; }
    }
    while (i1 <= last1) to[j++] = from1[i1++];
    return j - 1;
}
static int setMinusAll(int chromosomeLength, const int* from2, int last2, int* to) {
    int i1 = 0, i2 = 0, j = 0;
    while (i1 < chromosomeLength && i2 <= last2) {
        if (i1 < from2[i2]) to[j++] = i1++;
        else if (i1 > from2[i2]) i2++;
        else { i1++; i2++; }
    }
    while (i1 < chromosomeLength) to[j++] = i1++;
    return j - 1;
}
static int setIntersection(const int* x1, int from1, int last1, const int* x2, int from2, int last2, int* y, int fromY) {
    int i1 = from1, i2 = from2, j = fromY;
    while (i1 <= last1 && i2 <= last2) {
        if (x1[i1] < x2[i2]) i1++;
        else if (x1[i1] > x2[i2]) i2++;
        else { y[j++] = x1[i1]; i1++; i2++; }
    }
    return j - 1;
}

// src/SimulateHaplotypes.java:143-218. Arrays of capacity chromosomeLength, as in the Java: an append
// past that capacity is Java's ArrayIndexOutOfBoundsException.
void SimulateHaplotypes::loadPositions(int chromosomeLength, int maxDistance) {
    const int NT = BuildModel::N_INTERVAL_TYPES;
    positions.assign(NT, std::vector<int>((size_t)chromosomeLength, 0));
    positions_last.assign(NT, -1);
    auto push = [&](int type, int v) {
        if (++positions_last[type] >= chromosomeLength)
            throw std::runtime_error("ArrayIndexOutOfBoundsException: positions[" + std::to_string(type) + "]");
        positions[type][positions_last[type]] = v;
    };
    for (int i = 0; i <= model.lastRepeatInterval; i++) {
        int type = model.repeatIntervals[i][0];
        if (type < 0 || type >= NT) throw std::runtime_error("ArrayIndexOutOfBoundsException: interval type");
        int from = model.repeatIntervals[i][1] - maxDistance;
        if (from < 0) from = 0;
        int to = model.repeatIntervals[i][2] + maxDistance;
        if (to >= referenceLength) to = referenceLength - 1;
        for (int j = from; j <= to; j++) push(type, j);
    }
    for (int i = 0; i <= model.lastSegdup; i++) {
        int from = model.segdups[i][0] - maxDistance;
        if (from < 0) from = 0;
        int to = model.segdups[i][1] + maxDistance;
        if (to >= referenceLength) to = referenceLength - 1;
        for (int j = from; j <= to; j++) push(3, j);
    }
    for (int i = 0; i < NT; i++) {
        std::sort(positions[i].begin(), positions[i].begin() + (positions_last[i] + 1));
        int j = 0;
        for (int k = 1; k <= positions_last[i]; k++)
            if (positions[i][k] != positions[i][k - 1]) positions[i][++j] = positions[i][k];
        positions_last[i] = j;  // NB: 0 even when the set was empty (Java quirk, :170-174)
    }
    std::vector<int> tmpArray1((size_t)chromosomeLength), tmpArray2((size_t)chromosomeLength);
    int last1, last2;
    // Adding to BOTH the intersection of REP and SAT
    last1 = setIntersection(positions[0].data(), 0, positions_last[0], positions[1].data(), 0, positions_last[1], tmpArray1.data(), 0);
    last2 = setUnion(tmpArray1.data(), last1, positions[2].data(), positions_last[2], tmpArray2.data());
    std::copy(tmpArray2.begin(), tmpArray2.begin() + (last2 + 1), positions[2].begin());
    positions_last[2] = last2;
    // Subtracting BOTH from REP and SAT
    last1 = setMinus(positions[0].data(), positions_last[0], positions[2].data(), positions_last[2], tmpArray1.data());
    std::copy(tmpArray1.begin(), tmpArray1.begin() + (last1 + 1), positions[0].begin());
    positions_last[0] = last1;
    last1 = setMinus(positions[1].data(), positions_last[1], positions[2].data(), positions_last[2], tmpArray1.data());
    std::copy(tmpArray1.begin(), tmpArray1.begin() + (last1 + 1), positions[1].begin());
    positions_last[1] = last1;
    // Subtracting REP,SAT,BOTH from SEGDUP.
    last1 = setMinus(positions[3].data(), positions_last[3], positions[0].data(), positions_last[0], tmpArray1.data());
    last2 = setMinus(tmpArray1.data(), last1, positions[1].data(), positions_last[1], tmpArray2.data());
    last1 = setMinus(tmpArray2.data(), last2, positions[2].data(), positions_last[2], tmpArray1.data());
    std::copy(tmpArray1.begin(), tmpArray1.begin() + (last1 + 1), positions[3].begin());
    positions_last[3] = last1;
    // Building nonrepetitive positions
    last1 = setUnion(positions[0].data(), positions_last[0], positions[1].data(), positions_last[1], tmpArray1.data());
    last2 = setUnion(tmpArray1.data(), last1, positions[2].data(), positions_last[2], tmpArray2.data());
    last1 = setUnion(tmpArray2.data(), last2, positions[3].data(), positions_last[3], tmpArray1.data());
    positions_last[4] = setMinusAll(chromosomeLength, tmpArray1.data(), last1, positions[4].data());
    // Filtering positions by character
    for (int i = 0; i < NT; i++) {
        int k = -1;
        last1 = positions_last[i];
        for (int j = 0; j <= last1; j++) {
            int p = positions[i][j];
            if (p >= referenceLength) continue;
            char c = reference[p];
            if (c != 'A' && c != 'C' && c != 'G' && c != 'T' && c != 'a' && c != 'c' && c != 'g' && c != 't') continue;
            positions[i][++k] = p;
        }
        positions_last[i] = k;
    }
    positionsAllocated = true;
}

}  // namespace oracle
