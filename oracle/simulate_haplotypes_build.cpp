// ORACLE — TEST INFRASTRUCTURE ONLY. See simulate_haplotypes.h.
// buildPopulation (genotype matrix + collision fixer), insertion strings, haplotype materialisation.
#include "simulate_haplotypes.h"

namespace oracle {

// Variant.collidesWith (src/SimulateHaplotypes.java:1214-1234); "self" is `this`.
static bool collidesWith(const Variant& self, const Variant& other, int maxDistance) {
    if (self.svType == 2) {
        if (other.svType == 2) return std::abs(self.position - other.position) <= maxDistance;
        else return self.position >= other.position - maxDistance && self.position <= other.position + other.length - 1 + maxDistance;
    } else {
        if (other.svType == 2) {
            return other.position >= self.position - maxDistance && other.position <= self.position + self.length - 1 + maxDistance;
        } else {
            int first1 = std::min(self.position, other.position);
            int last1 = std::max(self.position + self.length - 1, other.position + other.length - 1);
            if (first1 <= last1) return true;
            first1 = self.position; last1 = self.position + self.length - 1;
            int first2 = other.position, last2 = other.position + other.length - 1;
            return (first1 > last2 && first1 <= last2 + maxDistance) || (last1 < first2 && last1 >= first2 - maxDistance);
        }
    }
}

// src/SimulateHaplotypes.java:603-718
void SimulateHaplotypes::buildPopulation(int nIndividuals, double alleleFrequencyThreshold, const std::string& outputDir) {
    N_HAPLOTYPES = nIndividuals << 1;
    const int COLLISION_DISTANCE = 5;
    const int HISTOGRAM_MAX_LENGTH = 4000;
    haplotype2variants.assign(N_HAPLOTYPES, std::vector<int>());
    haplotype2variants_last.assign(N_HAPLOTYPES, -1);
    for (int j = 0; j <= lastVariant; j++) {
        for (int i = 0; i < N_HAPLOTYPES; i++) {
            if (random.nextDouble() <= variants[j].alleleFrequency) {
                haplotype2variants_last[i]++;
                haplotype2variants[i].push_back(j);
            }
        }
    }
    drawLog.steps_after_buildPopulation = random.steps;
    if (!dumpPreCollisionFile.empty()) {
        FILE* f = fopen(dumpPreCollisionFile.c_str(), "w");
        if (!f) throw std::runtime_error("IOException: " + dumpPreCollisionFile);
        fprintf(f, "%d\n", N_HAPLOTYPES);
        for (int i = 0; i < N_HAPLOTYPES; i++) {
            fprintf(f, "%d", haplotype2variants_last[i]);
            for (int j = 0; j <= haplotype2variants_last[i]; j++) fprintf(f, ",%d", haplotype2variants[i][j]);
            fputc('\n', f);
        }
        fclose(f);
    }

    // Fixing collisions inside each haplotype
    const int nV = lastVariant + 1;
    std::vector<std::vector<int>> neighbors(nV);
    std::vector<int> lastNeighbor(nV, -1);
    std::vector<int> degree(nV, 0);
    std::vector<char> removed(nV, 0);
    jc::JavaPriorityQueue queue(nV, degree.data());
    int nHaplotypesWithCollision = 0;
    auto addEdge = [&](int node1, int node2) {  // :721-736
        lastNeighbor[node1]++;
        if ((int)neighbors[node1].size() <= lastNeighbor[node1]) neighbors[node1].resize(lastNeighbor[node1] + 1);
        neighbors[node1][lastNeighbor[node1]] = node2;
        lastNeighbor[node2]++;
        if ((int)neighbors[node2].size() <= lastNeighbor[node2]) neighbors[node2].resize(lastNeighbor[node2] + 1);
        neighbors[node2][lastNeighbor[node2]] = node1;
    };
    for (int i = 0; i < N_HAPLOTYPES; i++) {
        std::fill(lastNeighbor.begin(), lastNeighbor.end(), -1);
        int nEdges = 0;
        const int h2vLast = haplotype2variants_last[i];
        for (int j = 0; j <= h2vLast; j++) {
            int variantJ = haplotype2variants[i][j];
            int lastPosition = variants[variantJ].position + (variants[variantJ].svType == 2 ? 0 : variants[variantJ].length - 1);
            for (int k = j + 1; k <= h2vLast; k++) {
                int variantK = haplotype2variants[i][k];
                if (variants[variantK].position > lastPosition + COLLISION_DISTANCE) break;
                if (collidesWith(variants[variantK], variants[variantJ], COLLISION_DISTANCE)) {
                    addEdge(variantJ, variantK);
                    nEdges++;
                }
            }
        }
        if (nEdges == 0) continue;
        nHaplotypesWithCollision++;
        queue.clear();
        for (int j = 0; j <= h2vLast; j++) {
            int variantJ = haplotype2variants[i][j];
            degree[variantJ] = lastNeighbor[variantJ] + 1;
            removed[variantJ] = 0;
            queue.add(variantJ);
        }
        while (!queue.isEmpty()) {
            int element = queue.poll();
            if (removed[element]) continue;
            int variantJ = element;
            for (int k = 0; k <= lastNeighbor[variantJ]; k++) {
                int neighbor = neighbors[variantJ][k];
                if (removed[neighbor]) continue;
                removed[neighbor] = 1;
                for (int h = 0; h <= lastNeighbor[neighbor]; h++) {
                    int e = neighbors[neighbor][h];
                    if (!queue.remove(e)) continue;
                    degree[e]--;
                    queue.add(e);
                }
            }
        }
        int k = -1;
        for (int j = 0; j <= h2vLast; j++) {
            int variantJ = haplotype2variants[i][j];
            if (!removed[variantJ]) haplotype2variants[i][++k] = variantJ;
        }
        haplotype2variants_last[i] = k;
        haplotype2variants[i].resize(k + 1);
    }
    if (verbose) fprintf(stderr, "Fixed %d haplotypes with collisions\n", nHaplotypesWithCollision);

    // Marking unused variants
    for (int i = 0; i <= lastVariant; i++) variants[i].isActive = false;
    for (int i = 0; i < N_HAPLOTYPES; i++)
        for (int j = 0; j <= haplotype2variants_last[i]; j++) variants[haplotype2variants[i][j]].isActive = true;
    nVariants_frequent = 0; nVariants_rare = 0;
    for (int i = 0; i <= lastVariant; i++) {
        if (!variants[i].isActive) continue;
        if (variants[i].alleleFrequency >= alleleFrequencyThreshold) nVariants_frequent++;
        else nVariants_rare++;
    }
    // Histogram of nVariants per haplotype
    std::vector<int> histogram(HISTOGRAM_MAX_LENGTH + 1, 0);
    for (int i = 0; i < N_HAPLOTYPES; i++) {
        int j = haplotype2variants_last[i] + 1;
        if (j > HISTOGRAM_MAX_LENGTH) j = HISTOGRAM_MAX_LENGTH;
        histogram[j]++;
    }
    FILE* bw = fopen((outputDir + "/histogram_nVariantsPerHaplotype.txt").c_str(), "w");
    if (!bw) throw std::runtime_error("IOException: histogram_nVariantsPerHaplotype.txt");
    for (int i = 0; i <= HISTOGRAM_MAX_LENGTH; i++) fprintf(bw, "%d,%d\n", i, histogram[i]);
    fclose(bw);
}

// src/SimulateHaplotypes.java:786-814
void SimulateHaplotypes::assignInsertionStrings() {
    std::string buffer;
    for (int i = 0; i <= lastVariant; i++) {
        if (!variants[i].isActive || variants[i].svType != 2) continue;
        int j = getInsertionString(variants[i].length);
        if (j < 0 || j > (int)insertions.size() - 1) throw std::runtime_error("ArrayIndexOutOfBoundsException: insertions[" + std::to_string(j) + "]");
        buffer.clear();
        int insertionLength = (int)insertions[j].size();
        int variantLength = variants[i].length;
        if (insertionLength == variantLength) buffer += insertions[j];
        else if (insertionLength > variantLength) {
            int p = random.nextInt(insertionLength - variantLength);
            buffer.append(insertions[j], p, variantLength);
        } else {
            buffer += insertions[j];
            variants[i].length = insertionLength;
        }
        if (random.nextBoolean()) reverseComplement(buffer);
        variants[i].sequence = buffer;
        variants[i].hasSequence = true;
    }
    insertions.clear(); insertions.shrink_to_fit(); firstInsertionWithLength.clear();
}

// src/SimulateHaplotypes.java:827-875 (index i clobbered at :856 vs :860; unchecked [bin] at :833)
int SimulateHaplotypes::getInsertionString(int length) {
    const std::vector<int>& F = firstInsertionWithLength;
    const int FL = (int)F.size();
    auto at = [&](int b) -> int {
        if (b < 0 || b >= FL) throw std::runtime_error("ArrayIndexOutOfBoundsException: firstInsertionWithLength[" + std::to_string(b) + "]");
        return F[b];
    };
    auto insLen = [&](int s) -> int {
        if (s < 0 || s >= (int)insertions.size()) throw std::runtime_error("ArrayIndexOutOfBoundsException: insertions[" + std::to_string(s) + "]");
        return (int)insertions[s].size();
    };
    int i, j, bin, firstString, lastString, nBins;
    bin = length / BuildModel::SV_LENGTH_QUANTUM;
    nBins = FL;
    firstString = at(bin);
    if (firstString == -1) {
        while (bin < FL && F[bin] == -1) bin++;
        if (bin == FL) {
            bin = length / BuildModel::SV_LENGTH_QUANTUM;
            while (bin >= 0 && F[bin] == -1) bin--;
            return at(bin) + random.nextInt(lastInsertion - at(bin) + 1);
        } else {
            if (bin == FL - 1) lastString = lastInsertion;
            else {
                lastString = -1;
                for (i = bin + 1; i < nBins; i++) {
                    if (F[i] != -1) { lastString = F[i] - 1; break; }
                }
            }
            return F[bin] + random.nextInt(lastString - F[bin] + 1);
        }
    } else {
        i = firstString;
        if (bin == FL - 1) lastString = lastInsertion;
        else {
            lastString = -1;
            for (i = bin + 1; i < nBins; i++) {
                if (F[i] != -1) { lastString = F[i] - 1; break; }
            }
        }
        while (i <= lastString && insLen(i) < length) i++;
        j = lastString;
        while (j >= firstString && insLen(j) > length) j--;
        if (i == j) return i;
        else if (i < j) return i + random.nextInt(j - i + 1);
        else if (insLen(lastString) >= length) return j + 1 + random.nextInt(lastString - (j + 1) + 1);
        else return firstString + random.nextInt(lastString - firstString + 1);
    }
}

// src/SimulateHaplotypes.java:878-899
void SimulateHaplotypes::reverseComplement(std::string& buffer) {
    std::reverse(buffer.begin(), buffer.end());
    for (size_t i = 0; i < buffer.size(); i++) {
        char c = buffer[i], d;
        switch (c) {
            case 'a': d = 't'; break; case 'c': d = 'g'; break; case 'g': d = 'c'; break; case 't': d = 'a'; break;
            case 'A': d = 'T'; break; case 'C': d = 'G'; break; case 'G': d = 'C'; break; case 'T': d = 'A'; break;
            default: d = 'N';
        }
        buffer[i] = d;
    }
}

// Loop body of src/SimulateHaplotypes.java:926-953
void SimulateHaplotypes::buildHaplotypeString(int i, std::string& out) {
    out.clear();
    std::string buffer;
    int firstInReference = 0;
    auto appendRef = [&](int start, int end) {  // StringBuilder.append(CharSequence,int,int)
        if (start < 0 || start > end || end > referenceLength)
            throw std::runtime_error("IndexOutOfBoundsException: start " + std::to_string(start) + ", end " + std::to_string(end));
        out.append(reference, (size_t)start, (size_t)(end - start));
    };
    if (i < 0 || i >= (int)haplotype2variants.size()) throw std::runtime_error("ArrayIndexOutOfBoundsException: haplotype " + std::to_string(i));
    for (int j = 0; j <= haplotype2variants_last[i]; j++) {
        int id = haplotype2variants[i][j];
        if (id < 0 || id > lastVariant) throw std::runtime_error("ArrayIndexOutOfBoundsException: variant " + std::to_string(id));
        const Variant& variant = variants[id];
        appendRef(firstInReference, variant.position);
        if (variant.svType == 0) {
            firstInReference = variant.position + variant.length;
        } else if (variant.svType == 1) {
            int k = variant.position + variant.length;
            appendRef(variant.position, k);
            appendRef(variant.position, k);
            firstInReference = k;
        } else if (variant.svType == 2) {
            if (!variant.hasSequence) out += "null";  // StringBuilder.append((String)null)
            else out += variant.sequence;
            firstInReference = variant.position;
        } else if (variant.svType == 3) {
            buffer.clear();
            int e = variant.position + variant.length;
            if (variant.position < 0 || variant.position > e || e > referenceLength)
                throw std::runtime_error("IndexOutOfBoundsException: INV");
            buffer.append(reference, (size_t)variant.position, (size_t)variant.length);
            reverseComplement(buffer);
            out += buffer;
            firstInReference = variant.position + variant.length;
        }
    }
    if (firstInReference < referenceLength) appendRef(firstInReference, referenceLength);
}

// src/SimulateHaplotypes.java:917-960
void SimulateHaplotypes::buildHaplotypes(int firstHaplotype, int lastHaplotype, const std::string& outDir) {
    std::string out;
    out.reserve((size_t)referenceLength);
    for (int i = firstHaplotype; i <= lastHaplotype; i++) {
        buildHaplotypeString(i, out);
        FILE* bw = fopen((outDir + "/haplotype_" + std::to_string(i) + ".fa").c_str(), "wb");
        if (!bw) throw std::runtime_error("IOException: haplotype file");
        fprintf(bw, ">chr1_%d\n", i);
        fwrite(out.data(), 1, out.size(), bw);
        fputc('\n', bw);
        fclose(bw);
    }
}

// src/SimulateHaplotypes.java:967-987
void SimulateHaplotypes::buildVariant2haplotypes() {
    variant2haplotypes.assign(lastVariant + 1, std::vector<int>());
    variant2haplotypes_last.assign(lastVariant + 1, -1);
    for (int i = 0; i < N_HAPLOTYPES; i++) {
        for (int j = 0; j <= haplotype2variants_last[i]; j++) {
            int variantID = haplotype2variants[i][j];
            variant2haplotypes_last[variantID]++;
            variant2haplotypes[variantID].push_back(i);
        }
    }
    for (int i = 0; i <= lastVariant; i++) variants[i].alleleFrequency = (variant2haplotypes_last[i] + 1.0) / N_HAPLOTYPES;
}

}  // namespace oracle
