// Byte-exact writers of SimulateHaplotypes' small outputs (row Z of SURVEY.md §8a): haplotype2variants.txt, variants.txt
// and the two histograms (reference src/SimulateHaplotypes.java:707-717, :1104-1144, :1281-1292). The VCFs (row W) are
// assembled on the device: csrc/population_vcf.cu.
#include "population.h"
#include "textio.h"
#include <algorithm>

namespace sp {

namespace {
struct Out {
    FILE* f;
    std::string path;
    explicit Out(const std::string& p) : f(fopen(p.c_str(), "wb")), path(p) {
        if (!f) throw Error(SP_ERR_JAVA_EXCEPTION, "FileNotFoundException: " + p);
        setvbuf(f, nullptr, _IOFBF, 4 << 20);
    }
    ~Out() { if (f) fclose(f); }
    void s(const char* x) { fputs(x, f); }
    void s(const std::string& x) { fwrite(x.data(), 1, x.size(), f); }
    void b(const void* p, size_t n) { fwrite(p, 1, n, f); }
    void c(char x) { fputc(x, f); }
    void i(long long v) { fprintf(f, "%lld", v); }
    void close() {
        const bool bad = ferror(f) != 0;
        const int rc = fclose(f);
        f = nullptr;
        if (bad || rc != 0) throw Error(SP_ERR, "IOException: write failed: " + path);
    }
};
}  // namespace

// serialize :1123-1144, Variant.serialize :1281-1292
void write_population_files(const PopulationState& P) {
    {
        Out o(P.output_dir + "/haplotype2variants.txt");
        o.i(P.n_haplotypes_total); o.c('\n');
        for (int32_t h = 0; h < P.n_haplotypes_total; h++) {
            const int64_t a = P.all_off[h], b = P.all_off[h + 1];
            o.i(b - a - 1);
            for (int64_t k = a; k < b; k++) { o.c(','); o.i(P.all_ids[k]); }
            o.c('\n');
        }
        o.close();
    }
    {
        Out o(P.output_dir + "/variants.txt");
        o.i((int64_t)P.n_variants - 1); o.c('\n');
        for (int32_t i = 0; i < P.n_variants; i++) {
            const VariantRec& v = P.variants[i];
            o.i(v.svType); o.c('\n');
            o.i(v.intervalType); o.c('\n');
            o.i(v.length); o.c('\n');
            o.s(java_double_to_string(v.alleleFrequency)); o.c('\n');
            o.i(v.position); o.c('\n');
            o.s(v.isActive ? "true\n" : "false\n");
            if (v.hasSequence) o.b(P.payload.data() + v.seq_off, (size_t)v.seq_len); else o.s("null");
            o.c('\n');
            o.i(v.intervalType_start); o.c('\n');
            o.i(v.intervalType_end); o.c('\n');
            o.i(v.intervalSurface); o.c('\n');
        }
        o.close();
    }
}

// histogram_nVariantsPerHaplotype.txt :707-717 and histogram_startPositions.txt :1104-1120
void write_histograms(const PopulationState& P, int32_t L) {
    {
        const int MAXL = 4000;
        std::vector<int32_t> hist(MAXL + 1, 0);
        for (int32_t h = 0; h < P.n_haplotypes_total; h++) {
            int64_t j = P.all_off[h + 1] - P.all_off[h];
            if (j > MAXL) j = MAXL;
            hist[j]++;
        }
        Out o(P.output_dir + "/histogram_nVariantsPerHaplotype.txt");
        for (int i = 0; i <= MAXL; i++) { o.i(i); o.c(','); o.i(hist[i]); o.c('\n'); }
        o.close();
    }
    {
        const int bin = 100000;
        const int nb = L / bin + 1;
        std::vector<int32_t> hist((size_t)4 * nb, 0);
        for (int32_t i = 0; i < P.n_variants; i++) {
            const VariantRec& v = P.variants[i];
            if (!v.isActive) continue;
            if (v.svType < 0 || v.svType > 3 || v.position / bin >= nb || v.position < 0)
                throw Error(SP_ERR_JAVA_EXCEPTION, "ArrayIndexOutOfBoundsException: position histogram");
            hist[(size_t)v.svType * nb + v.position / bin]++;
        }
        Out o(P.output_dir + "/histogram_startPositions.txt");
        for (int t = 0; t < 4; t++) {
            for (int j = 0; j < nb; j++) { o.i(hist[(size_t)t * nb + j]); o.c(','); }
            o.c('\n');
        }
        o.close();
    }
}

}  // namespace sp
