// Byte-exact writers of SimulateHaplotypes' outputs (rows W and Z of SURVEY.md §8a): groundTruth_joint.vcf,
// groundTruth_individual_<2k>.vcf, haplotype2variants.txt, variants.txt and the two histograms
// (reference src/SimulateHaplotypes.java:707-717, :1005-1144, :1243-1292).
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

// Variant.writeVCF :1243-1279
static void write_vcf_record(Out& o, const PopulationState& P, const uint8_t* ref, int32_t L, int32_t id) {
    const VariantRec& v = P.variants[id];
    auto char_at = [&](int64_t p) -> char {
        if (p < 0 || p >= L) throw Error(SP_ERR_JAVA_EXCEPTION, "StringIndexOutOfBoundsException: index " + std::to_string(p));
        return (char)ref[p];
    };
    o.s("chr1\t");
    if (v.svType == SP_DEL) {
        o.i(v.position); o.c('\t');
        o.s("simulated_"); o.i(id); o.c('\t');
        const int64_t b = (int64_t)v.position - 1, e = (int64_t)v.position + v.length;
        if (b < 0 || e > L || b > e)
            throw Error(SP_ERR_JAVA_EXCEPTION, "StringIndexOutOfBoundsException: begin " + std::to_string(b) + ", end " + std::to_string(e) + ", length " + std::to_string(L));
        o.b(ref + b, (size_t)(e - b)); o.c('\t');
        o.c(char_at(b)); o.c('\t');
        o.s("999\tPASS\t");
        o.s("END="); o.i((int64_t)v.position + v.length); o.c(';');
        o.s("SVLEN="); o.i(-(int64_t)v.length); o.c(';');
    } else if (v.svType == SP_INS) {
        o.i((int64_t)v.position + 1); o.c('\t');
        o.s("simulated_"); o.i(id); o.c('\t');
        o.c(char_at(v.position)); o.c('\t');
        o.c(char_at(v.position));
        if (!v.hasSequence) throw Error(SP_ERR_JAVA_EXCEPTION, "NullPointerException: INS without sequence");
        o.b(P.payload.data() + v.seq_off, (size_t)v.seq_len); o.c('\t');
        o.s("999\tPASS\t");
        o.s("END="); o.i((int64_t)v.position + 1); o.c(';');
        o.s("SVLEN="); o.i(v.length); o.c(';');
    } else {
        o.i((int64_t)v.position + 1); o.c('\t');
        o.s("simulated_"); o.i(id); o.c('\t');
        o.c(char_at(v.position)); o.c('\t');
        o.c('<'); o.s(SV_TYPE_NAMES[v.svType]); o.s(">\t");
        o.s("999\tPASS\t");
        o.s("END="); o.i((int64_t)v.position + v.length); o.c(';');
    }
    o.s("SVTYPE="); o.s(SV_TYPE_NAMES[v.svType]); o.c(';');
    o.s("AF="); o.s(java_double_to_string(v.alleleFrequency)); o.c(';');
    o.s("REPEATS_START="); o.i(v.intervalType_start); o.c(';');
    o.s("REPEATS_END="); o.i(v.intervalType_end); o.c(';');
    o.s("REPEATS_FRACTION="); o.s(java_number_format4(((double)v.intervalSurface) / v.length));
}

// printVcfHeader :1077-1096
static void write_vcf_header(Out& o, int32_t L, bool genotype, const std::string& sample) {
    o.s("##fileformat=VCFv4.2\n");
    o.s("##contig=<ID=chr1,length="); o.i(L); o.s(">\n");
    o.s("##ALT=<ID=INS,Description=\"Insertion\">\n"
        "##ALT=<ID=DEL,Description=\"Deletion\">\n"
        "##ALT=<ID=DUP,Description=\"Duplication\">\n"
        "##ALT=<ID=INV,Description=\"Inversion\">\n"
        "##FILTER=<ID=PASS,Description=\"All filters passed\">\n"
        "##INFO=<ID=END,Number=1,Type=Integer,Description=\"End position of structural variation\">\n"
        "##INFO=<ID=SVTYPE,Number=1,Type=String,Description=\"Type of structural variant\">\n"
        "##INFO=<ID=SVLEN,Number=1,Type=Integer,Description=\"Length of structural variation\">\n"
        "##INFO=<ID=AF,Number=1,Type=Float,Description=\"Allele Frequency\">\n"
        "##INFO=<ID=REPEATS_START,Number=1,Type=Integer,Description=\"Repetitive context around the first position\">\n"
        "##INFO=<ID=REPEATS_END,Number=1,Type=Integer,Description=\"Repetitive context around the last position\">\n"
        "##INFO=<ID=REPEATS_FRACTION,Number=1,Type=Float,Description=\"Fraction of the SV covered by repeats of any type\">\n"
        "##FORMAT=<ID=GT,Number=1,Type=String,Description=\"Genotype\">\n"
        "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO");
    if (genotype) { o.s("\tFORMAT\t"); o.s(sample); }
    o.c('\n');
}

// Joint VCF (:1014-1024). all_off/all_ids: CSR over ALL haplotypes (needed for ";HAPLOTYPES=").
void write_joint_vcf(const PopulationState& P, const uint8_t* ref, int32_t L) {
    const int32_t H = P.n_haplotypes_total, V = P.n_variants;
    // variant2haplotypes (:967-985): transpose, haplotypes ascending
    std::vector<int64_t> voff((size_t)V + 1, 0);
    for (int64_t k = 0; k < (int64_t)P.all_ids.size(); k++) voff[P.all_ids[k] + 1]++;
    for (int32_t v = 0; v < V; v++) voff[v + 1] += voff[v];
    std::vector<int32_t> v2h(P.all_ids.size());
    std::vector<int64_t> cur(voff.begin(), voff.end() - 1);
    for (int32_t h = 0; h < H; h++)
        for (int64_t k = P.all_off[h]; k < P.all_off[h + 1]; k++) v2h[cur[P.all_ids[k]]++] = h;
    Out o(P.output_dir + "/groundTruth_joint.vcf");
    write_vcf_header(o, L, false, "");
    for (int32_t v = 0; v < V; v++) {
        if (!P.variants[v].isActive) continue;
        write_vcf_record(o, P, ref, L, v);
        o.s(";HAPLOTYPES=");
        for (int64_t k = voff[v]; k < voff[v + 1]; k++) {
            if (k > voff[v]) o.c(',');
            o.i(v2h[k]);
        }
        o.c('\n');
    }
    o.close();
}

// Individual VCFs (:1027-1066) for haplotype pairs (i,i+1), i even, inside [h_first, h_last]; lists from a CSR
// whose row 0 is haplotype `base`.
void write_individual_vcfs(const PopulationState& P, const uint8_t* ref, int32_t L, int32_t h_first, int32_t h_last,
                           const int64_t* off, const int32_t* ids, int32_t base) {
    for (int32_t i = h_first + (h_first & 1); i + 1 <= h_last; i += 2) {
        Out o(P.output_dir + "/groundTruth_individual_" + std::to_string(i) + ".vcf");
        write_vcf_header(o, L, true, "individual_" + std::to_string(i));
        const int32_t* a = ids + off[i - base];
        const int32_t* b = ids + off[i + 1 - base];
        const int64_t na = off[i + 1 - base] - off[i - base], nb = off[i + 2 - base] - off[i + 1 - base];
        int64_t j = 0, k = 0;
        while (j < na && k < nb) {
            if (a[j] < b[k]) { write_vcf_record(o, P, ref, L, a[j]); o.s("\tGT\t1|0\n"); j++; }
            else if (b[k] < a[j]) { write_vcf_record(o, P, ref, L, b[k]); o.s("\tGT\t0|1\n"); k++; }
            else { write_vcf_record(o, P, ref, L, a[j]); o.s("\tGT\t1|1\n"); j++; k++; }
        }
        while (j < na) { write_vcf_record(o, P, ref, L, a[j]); o.s("\tGT\t1|0\n"); j++; }
        while (k < nb) { write_vcf_record(o, P, ref, L, b[k]); o.s("\tGT\t0|1\n"); k++; }
        o.close();
    }
}

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
