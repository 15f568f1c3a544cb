// Sequential, data-dependent parts of SimulateHaplotypes that run once per population on the host, replicated
// on every rank from the same seed: variant sampling, insertion-string assignment,
// repeat annotation. Their draw counts depend on the draws themselves (nextInt rejections, resampling), so they
// are not index-addressable; they are O(V) and feed the kernels, they are not the hot path.
#include "population.h"
#include "java_random.cuh"
#include "textio.h"
#include <algorithm>

namespace sp {

// Arrays.binarySearch(double[],from,to,key) followed by the callers' `if (p<0) p=-1-p` (:386-396).
static int search_cdf(const double* a, int from, int to, double key) {
    int lo = from, hi = to - 1;
    while (lo <= hi) {
        const int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
        const double v = a[mid];
        if (v < key) lo = mid + 1;
        else if (v > key) hi = mid - 1;
        else return mid;  // keys are finite, non-negative: equal values are bit-equal
    }
    return lo;
}

// sampleVariants + sampleVariant, src/SimulateHaplotypes.java:353-400
void sample_variants(PopulationState& P, int64_t seed, int32_t reference_length) {
    const Model& M = P.model;
    int population_id = -1;
    {   // Arrays.binarySearch(BuildModel.AF_NAMES,POPULATION_NAME) on the (unsorted!) name table, :101
        int lo = 0, hi = 5;
        population_id = -1;
        while (lo <= hi) {
            const int mid = (lo + hi) >> 1;
            const int cmp = strcmp(AF_NAMES[mid], P.population_name.c_str());
            if (cmp < 0) lo = mid + 1; else if (cmp > 0) hi = mid - 1; else { population_id = mid; break; }
        }
        if (population_id < 0) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: population " + P.population_name);
    }
    if ((int)M.af.size() <= population_id) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: population table");
    JavaRandom random(seed);
    P.variants.clear();
    P.candidates = 0;
    int32_t n_frequent = 0, n_rare = 0;
    auto oob = [](const char* what) { return Error(SP_ERR_JAVA_EXCEPTION, std::string("ArrayIndexOutOfBoundsException: ") + what); };
    while (n_frequent < P.n_frequent) {
        VariantRec v;
        do {
            P.candidates++;
            v = VariantRec();
            v.svType = search_cdf(M.svType.data(), 0, (int)M.svType.size(), random.nextDouble());
            if (v.svType >= (int)M.intervalType.size() || v.svType >= (int)M.length.size()) throw oob("svType");
            const auto& it = M.intervalType[v.svType];
            v.intervalType = search_cdf(it.data(), 0, (int)it.size(), random.nextDouble());
            if (v.intervalType >= (int)M.length[v.svType].size()) throw oob("intervalType");
            const auto& sl = M.length[v.svType][v.intervalType];
            int p = search_cdf(sl.data(), 0, (int)sl.size(), random.nextDouble());
            v.length = MIN_SV_LENGTH + SV_LENGTH_QUANTUM * p + random.nextInt(SV_LENGTH_QUANTUM);
            if (v.svType >= (int)M.af[population_id].size()) throw oob("alleleFrequency svType");
            const double mn = M.af_minmax[population_id][v.svType][0];
            const double quantum = (M.af_minmax[population_id][v.svType][1] - mn) / (ALLELE_FREQUENCY_NBINS - 1);
            const auto& af = M.af[population_id][v.svType];
            if ((int)af.size() < ALLELE_FREQUENCY_NBINS) throw oob("alleleFrequency bins");
            p = search_cdf(af.data(), 0, ALLELE_FREQUENCY_NBINS, random.nextDouble());
            // min+quantum*p+random.nextDouble()*quantum, left to right, IEEE double without contraction
            volatile double a = quantum * (double)p;
            volatile double b = mn + a;
            volatile double c = random.nextDouble() * quantum;
            v.alleleFrequency = b + c;
            if (v.intervalType >= N_INTERVAL_TYPES) throw oob("positions");
            const IntervalSet& S = P.positions.set[v.intervalType];
            const int64_t cnt = S.count();
            if (cnt <= 0) throw Error(SP_ERR_JAVA_EXCEPTION, "IllegalArgumentException: bound must be positive (no position of interval type " + std::to_string(v.intervalType) + ")");
            v.position = S.kth(random.nextInt((int32_t)cnt));
        } while (v.svType != SP_INS && (int64_t)v.position + v.length > reference_length);
        P.variants.push_back(v);
        if (v.alleleFrequency >= P.af_threshold) n_frequent++; else n_rare++;
    }
    std::stable_sort(P.variants.begin(), P.variants.end(),
                     [](const VariantRec& x, const VariantRec& y) { return x.position < y.position; });
    P.n_variants = (int32_t)P.variants.size();
    P.nVariants_frequent = n_frequent; P.nVariants_rare = n_rare;
    P.steps_after_sampling = random.steps;
    P.state_after_sampling = random.state;
}

// loadInsertions :313-337
struct Insertions {
    std::vector<char> buf;
    std::vector<std::pair<int64_t, int32_t>> str;  // offset, length
    std::vector<int32_t> first_with_length;
    int32_t last = -1;
    const char* data(int i) const { return buf.data() + str[i].first; }
    int32_t len(int i) const { return str[i].second; }
};

static void load_insertions(Insertions& I, const std::string& path, int32_t n_insertions) {
    I.buf = slurp(path);
    Lines lines(I.buf.data(), I.buf.data() + I.buf.size());
    const char *b, *e;
    I.str.clear();
    I.last = -1;
    while (lines.next(b, e)) {
        if (++I.last >= n_insertions) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: insertions[" + std::to_string(I.last) + "]");
        trim_span(b, e);
        I.str.emplace_back((int64_t)(b - I.buf.data()), (int32_t)(e - b));
    }
    if (I.last < 0) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: insertions[-1]");
    const int nbins = 1 + I.len(I.last) / SV_LENGTH_QUANTUM;
    I.first_with_length.assign(nbins, -1);
    for (int i = 0; i <= I.last; i++) {
        const int j = I.len(i) / SV_LENGTH_QUANTUM;
        if (j >= nbins) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: firstInsertionWithLength[" + std::to_string(j) + "]");
        if (I.first_with_length[j] == -1) I.first_with_length[j] = i;
    }
}

// getInsertionString :827-875, with its clobbered scan index (:856 vs :860) and unchecked [bin] (:833)
static int get_insertion_string(const Insertions& I, JavaRandom& random, int length) {
    const std::vector<int32_t>& F = I.first_with_length;
    const int FL = (int)F.size();
    auto at = [&](int b) -> int {
        if (b < 0 || b >= FL) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: firstInsertionWithLength[" + std::to_string(b) + "]");
        return F[b];
    };
    auto slen = [&](int s) -> int {
        if (s < 0 || s > I.last) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: insertions[" + std::to_string(s) + "]");
        return I.len(s);
    };
    auto next_int = [&](int bound) -> int {
        if (bound <= 0) throw Error(SP_ERR_JAVA_EXCEPTION, "IllegalArgumentException: bound must be positive");
        return random.nextInt(bound);
    };
    int i, j, lastString;
    int bin = length / SV_LENGTH_QUANTUM;
    const int firstString = at(bin);
    if (firstString == -1) {
        while (bin < FL && F[bin] == -1) bin++;
        if (bin == FL) {
            bin = length / SV_LENGTH_QUANTUM;
            while (bin >= 0 && F[bin] == -1) bin--;
            return at(bin) + next_int(I.last - at(bin) + 1);
        }
        if (bin == FL - 1) lastString = I.last;
        else {
            lastString = -1;
            for (i = bin + 1; i < FL; i++)
                if (F[i] != -1) { lastString = F[i] - 1; break; }
        }
        return F[bin] + next_int(lastString - F[bin] + 1);
    }
    i = firstString;
    if (bin == FL - 1) lastString = I.last;
    else {
        lastString = -1;
        for (i = bin + 1; i < FL; i++)
            if (F[i] != -1) { lastString = F[i] - 1; break; }
    }
    while (i <= lastString && slen(i) < length) i++;
    j = lastString;
    while (j >= firstString && slen(j) > length) j--;
    if (i == j) return i;
    if (i < j) return i + next_int(j - i + 1);
    if (slen(lastString) >= length) return j + 1 + next_int(lastString - (j + 1) + 1);
    return firstString + next_int(lastString - firstString + 1);
}

static inline char complement_char(char c) {
    switch (c) {
        case 'a': return 't'; case 'c': return 'g'; case 'g': return 'c'; case 't': return 'a';
        case 'A': return 'T'; case 'C': return 'G'; case 'G': return 'C'; case 'T': return 'A';
        default: return 'N';
    }
}

// loadInsertions + assignInsertionStrings :786-814; the generator continues the single stream after the
// 2*V*H genotype steps.
void assign_insertion_strings(PopulationState& P, int64_t seed) {
    Insertions I;
    load_insertions(I, P.insertions_file, P.n_insertion_strings);
    JavaRandom random(seed);
    random.skip(P.steps_after_sampling + 2ull * (uint64_t)P.n_variants * (uint64_t)P.n_haplotypes_total);
    P.payload.clear();
    for (int32_t v = 0; v < P.n_variants; v++) {
        VariantRec& r = P.variants[v];
        if (!r.isActive || r.svType != SP_INS) continue;
        const int j = get_insertion_string(I, random, r.length);
        if (j < 0 || j > I.last) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: insertions[" + std::to_string(j) + "]");
        const int insertionLength = I.len(j), variantLength = r.length;
        const char* s = I.data(j);
        int from = 0, n = insertionLength;
        if (insertionLength > variantLength) {
            from = random.nextInt(insertionLength - variantLength);
            n = variantLength;
        } else if (insertionLength < variantLength) r.length = insertionLength;
        r.seq_off = (int64_t)P.payload.size();
        r.seq_len = n;
        r.hasSequence = true;
        if (random.nextBoolean()) {
            for (int k = n - 1; k >= 0; k--) P.payload.push_back((uint8_t)complement_char(s[from + k]));
        } else P.payload.insert(P.payload.end(), (const uint8_t*)s + from, (const uint8_t*)s + from + n);
    }
}

// annotateVariantsWithRepeats :412-432 with getIntervalType :442-473, getIntervalSurface :484-515 and
// getIntersectingIntervals :527-571 (stale tmpPairs[0] at :513, loop variable clobbered by the grow path).
namespace {
struct PairT { int32_t first = 0, last = 0; };
struct Annotator {
    const Model& M;
    std::vector<PairT> tmpPairs;
    explicit Annotator(const Model& m) : M(m) {}

    int interval_type(int position, int maxDistance, int* cursors, bool update) {
        const auto& R = M.repeats; const int lastR = (int)R.size() - 1;
        const auto& S = M.segdups; const int lastS = (int)S.size() - 1;
        int i1 = cursors[0], i2 = cursors[1], t;
        while (i1 <= lastR && R[i1][2] + maxDistance < position) i1++;
        if (i1 > lastR || R[i1][1] > position + maxDistance) t = 4;
        else {
            t = R[i1][0];
            auto add = [&](int a, int b) {
                if (a < 0 || a > 2 || b < 0 || b > 2) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: ADD_TYPE");
                return ADD_TYPE[a][b];
            };
            for (int j = i1 - 1; j >= 0; j--) {
                if (R[j][2] + maxDistance < position) break;
                t = add(t, R[j][0]);
            }
            for (int j = i1 + 1; j <= lastR; j++) {
                if (R[j][1] > position + maxDistance) break;
                t = add(t, R[j][0]);
            }
        }
        if (t == 4) {
            while (i2 <= lastS && S[i2][1] + maxDistance < position) i2++;
            t = (i2 > lastS || S[i2][0] > position + maxDistance) ? 4 : 3;
        }
        if (update) { cursors[0] = i1; cursors[1] = i2; }
        return t;
    }

    template <class Mx>
    void intersecting(int first, int last, const Mx& matrix, int cf, int cl, int* cursors) {
        const int lastRow = (int)matrix.size() - 1;
        int i = cursors[0], j = cursors[1], k = 0;
        auto row = [&](int r) -> const typename Mx::value_type& {
            if (r < 0 || r > lastRow) throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: interval row " + std::to_string(r));
            return matrix[r];
        };
        auto grow = [&]() { tmpPairs.resize(tmpPairs.size() << 1); return (int)tmpPairs.size(); };
        while (i <= lastRow && row(i)[cl] < first) i++;
        if (i <= lastRow && row(i)[cf] <= last) {
            j++;
            if (j == (int)tmpPairs.size()) k = grow();
            tmpPairs[j].first = std::max(first, row(i)[cf]);
            tmpPairs[j].last = std::min(last, row(i)[cl]);
            for (k = i - 1; k >= 0; k--) {
                if (row(k)[cl] < first) break;
                j++;
                if (j == (int)tmpPairs.size()) k = grow();
                tmpPairs[j].first = std::max(first, row(k)[cf]);
                tmpPairs[j].last = std::min(last, row(k)[cl]);
            }
            for (k = i + 1; k <= lastRow; k++) {
                if (row(k)[cf] > last) break;
                j++;
                if (j == (int)tmpPairs.size()) k = grow();
                tmpPairs[j].first = std::max(first, row(k)[cf]);
                tmpPairs[j].last = std::min(last, row(k)[cl]);
            }
        }
        (void)k;
        cursors[0] = i; cursors[1] = j;
    }

    int interval_surface(int first, int last, int* cursors, int* tmp) {
        if (tmpPairs.empty()) tmpPairs.assign(100, PairT());
        tmp[0] = cursors[0]; tmp[1] = -1;
        intersecting(first, last, M.repeats, 1, 2, tmp);
        cursors[0] = tmp[0];
        tmp[0] = cursors[1];
        intersecting(first, last, M.segdups, 0, 1, tmp);
        cursors[1] = tmp[0];
        const int lastPair = tmp[1];
        if (lastPair > 0)
            std::sort(tmpPairs.begin(), tmpPairs.begin() + (lastPair + 1), [](const PairT& a, const PairT& b) {
                return a.first < b.first || (a.first == b.first && a.last < b.last);
            });
        int j = 0, sum = 0;
        for (int i = 1; i <= lastPair; i++) {
            if (tmpPairs[i].first <= tmpPairs[j].last) tmpPairs[j].last = std::max(tmpPairs[j].last, tmpPairs[i].last);
            else {
                sum += tmpPairs[j].last - tmpPairs[j].first + 1;
                j++;
                tmpPairs[j] = tmpPairs[i];
            }
        }
        sum += tmpPairs[j].last - tmpPairs[j].first + 1;
        return sum;
    }
};
}  // namespace

void annotate_variants_with_repeats(PopulationState& P, int maxDistance) {
    Annotator A(P.model);
    int c1[2] = {0, 0}, c2[2] = {0, 0}, c3[2] = {0, 0};
    for (int32_t i = 0; i < P.n_variants; i++) {
        VariantRec& v = P.variants[i];
        v.intervalType_start = A.interval_type(v.position, maxDistance, c1, true);
        if (v.svType == SP_INS) {
            v.intervalType_end = v.intervalType_start;
            v.intervalSurface = 0;
        } else {
            const int last = v.position + v.length - 1;
            v.intervalType_end = A.interval_type(last, maxDistance, c1, false);
            v.intervalSurface = A.interval_surface(v.position, last, c2, c3);
        }
    }
}

}  // namespace sp
