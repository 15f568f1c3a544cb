// Row W of SURVEY.md §8a on the device — the truth VCFs of SimulateHaplotypes.buildVCF (reference
// src/SimulateHaplotypes.java:1005-1067, Variant.writeVCF :1243-1279).
//
// A VCF is almost entirely bytes that already sit in HBM: the REF column of a deletion is reference[position-1 ..
// position+length) (:1248, tens of kilobases per record) and the ALT column of an insertion is its payload. A variant's
// record text is the same in every file it appears in (the joint VCF and the VCF of each carrier), so the files are built
// by the tiled copy kernel of row M in two gathers:
//   1. record blob: per active variant [prefix text | reference bytes or payload | suffix text], the texts (numbers,
//      Double.toString of the allele frequency, NumberFormat of the repeat fraction) formatted by the host into a small
//      auxiliary blob;
//   2. files: header + per record [its bytes in the record blob | "\tGT\t1|0\n" or ";HAPLOTYPES=...\n"], several
//      individuals per launch (one slot each), then device -> pinned ring -> file (csrc/file_sink.h).
// The merge of the two haplotype lists of an individual (:1031-1062) and the variant2haplotypes transpose (:967-985) are
// index work on the host: a few entries per kilobase of output.
#include "population.h"
#include "textio.h"
#include "file_sink.h"
#include <algorithm>

namespace sp {

void launch_tiled_copy(sp_ctx* c, const uint8_t* src, const uint8_t* payload, const uint8_t* aux, const int64_t* seg_dst,
                       const int64_t* seg_src, const int64_t* seg_base, int64_t n_segments, uint8_t* out, int64_t stride, int n_slots);

namespace {

// Segment tables in the layout plan_kernel produces: per slot ascending destinations, a PAD segment at the slot's length,
// a sentinel.
struct SegTables {
    std::vector<int64_t> dst, src, base{0};
    int64_t cur = 0;
    void add(int kind, int64_t off, int64_t len) {
        if (len <= 0) return;
        dst.push_back(cur);
        src.push_back((int64_t(kind) << SEG_KIND_SHIFT) | (off & SEG_OFF_MASK));
        cur += len;
    }
    int64_t end_slot() {   // returns the slot's length
        const int64_t n = cur;
        dst.push_back(cur); src.push_back(int64_t(SEG_PAD) << SEG_KIND_SHIFT);
        dst.push_back(INT64_MAX); src.push_back(int64_t(SEG_PAD) << SEG_KIND_SHIFT);
        base.push_back((int64_t)dst.size());
        cur = 0;
        return n;
    }
};

struct Text {   // the auxiliary blob; every piece is referenced by (offset, length)
    std::string s;
    int64_t mark() const { return (int64_t)s.size(); }
    void i(long long v) { s += std::to_string(v); }
};

void header_text(Text& t, int32_t L, bool genotype, const std::string& sample) {   // printVcfHeader :1077-1096
    t.s += "##fileformat=VCFv4.2\n##contig=<ID=chr1,length=";
    t.i(L);
    t.s += ">\n"
           "##ALT=<ID=INS,Description=\"Insertion\">\n"
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
           "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO";
    if (genotype) { t.s += "\tFORMAT\t"; t.s += sample; }
    t.s += '\n';
}

void upload(sp_ctx* c, DevBuf<int64_t>& d, const std::vector<int64_t>& h) {
    d.reserve(std::max<size_t>(h.size(), 1));
    if (!h.empty()) SP_CUDA(cudaMemcpyAsync(d.p, h.data(), h.size() * 8, cudaMemcpyHostToDevice, c->stream));
}

}  // namespace

struct VcfDeviceState {
    DevBuf<uint8_t> d_text, d_records, d_files;
    DevBuf<int64_t> d_seg_dst, d_seg_src, d_seg_base;
    std::vector<int64_t> rec_off;   // V+1 offsets into d_records (inactive variants have empty records)
    double gather_ms = 0.0;
    int64_t bytes_out = 0;
};

static void run_copy(sp_ctx* c, VcfDeviceState& S, SegTables& T, const uint8_t* src, const uint8_t* payload, const std::string& text,
                     DevBuf<uint8_t>& out, int64_t stride, int n_slots) {
    S.d_text.reserve(std::max<size_t>(text.size(), 1));
    if (!text.empty()) SP_CUDA(cudaMemcpyAsync(S.d_text.p, text.data(), text.size(), cudaMemcpyHostToDevice, c->stream));
    upload(c, S.d_seg_dst, T.dst); upload(c, S.d_seg_src, T.src); upload(c, S.d_seg_base, T.base);
    out.reserve((size_t)stride * n_slots, 0.05);
    launch_tiled_copy(c, src, payload, S.d_text.p, S.d_seg_dst.p, S.d_seg_src.p, S.d_seg_base.p, (int64_t)T.dst.size(), out.p, stride, n_slots);
    SP_CUDA(cudaStreamSynchronize(c->stream));   // the host vectors above go out of scope with the caller
}

// Gather 1: the record of every active variant, once. Requires the variant table with payloads on the device
// (publish_variants after assign_insertion_strings).
static void build_record_blob(sp_ctx* c, const PopulationState& P, VcfDeviceState& S, int32_t L) {
    const uint8_t* ref = c->h_ref.data();
    auto char_at = [&](int64_t p) -> char {
        if (p < 0 || p >= L) throw Error(SP_ERR_JAVA_EXCEPTION, "StringIndexOutOfBoundsException: index " + std::to_string(p));
        return (char)ref[p];
    };
    Text t;
    SegTables T;
    const int32_t V = P.n_variants;
    S.rec_off.assign((size_t)V + 1, 0);
    for (int32_t id = 0; id < V; id++) {
        S.rec_off[id] = T.cur;
        const VariantRec& v = P.variants[id];
        if (!v.isActive) continue;
        int64_t a = t.mark();
        t.s += "chr1\t";
        if (v.svType == SP_DEL) {
            t.i(v.position); t.s += "\tsimulated_"; t.i(id); t.s += '\t';
            const int64_t b = (int64_t)v.position - 1, e = (int64_t)v.position + v.length;
            if (b < 0 || e > L || b > e)
                throw Error(SP_ERR_JAVA_EXCEPTION, "StringIndexOutOfBoundsException: begin " + std::to_string(b) + ", end " + std::to_string(e) + ", length " + std::to_string(L));
            T.add(SEG_AUX, a, t.mark() - a);
            T.add(SEG_REF_FWD, b, e - b);                       // REF = reference.substring(position-1, position+length)
            a = t.mark();
            t.s += '\t'; t.s += char_at(b); t.s += "\t999\tPASS\tEND="; t.i((int64_t)v.position + v.length);
            t.s += ";SVLEN="; t.i(-(int64_t)v.length); t.s += ';';
        } else if (v.svType == SP_INS) {
            t.i((int64_t)v.position + 1); t.s += "\tsimulated_"; t.i(id); t.s += '\t';
            t.s += char_at(v.position); t.s += '\t'; t.s += char_at(v.position);
            if (!v.hasSequence) throw Error(SP_ERR_JAVA_EXCEPTION, "NullPointerException: INS without sequence");
            T.add(SEG_AUX, a, t.mark() - a);
            T.add(SEG_PAYLOAD, c->h_payload_off[id], v.seq_len);   // ALT = REF + sequence
            a = t.mark();
            t.s += "\t999\tPASS\tEND="; t.i((int64_t)v.position + 1); t.s += ";SVLEN="; t.i(v.length); t.s += ';';
        } else {
            t.i((int64_t)v.position + 1); t.s += "\tsimulated_"; t.i(id); t.s += '\t';
            t.s += char_at(v.position); t.s += "\t<"; t.s += SV_TYPE_NAMES[v.svType]; t.s += ">\t999\tPASS\tEND=";
            t.i((int64_t)v.position + v.length); t.s += ';';
        }
        t.s += "SVTYPE="; t.s += SV_TYPE_NAMES[v.svType];
        t.s += ";AF="; t.s += java_double_to_string(v.alleleFrequency);
        t.s += ";REPEATS_START="; t.i(v.intervalType_start);
        t.s += ";REPEATS_END="; t.i(v.intervalType_end);
        t.s += ";REPEATS_FRACTION="; t.s += java_number_format4(((double)v.intervalSurface) / v.length);
        T.add(SEG_AUX, a, t.mark() - a);
    }
    S.rec_off[V] = T.cur;
    const int64_t total = T.end_slot();
    const int64_t stride = std::max<int64_t>(TILE_BYTES, (total + TILE_BYTES - 1) / TILE_BYTES * TILE_BYTES);
    run_copy(c, S, T, c->d_ref.p, c->d_payload.p, t.s, S.d_records, stride, 1);
}

// Gather 2 + emission for a set of files; `fill(k, text, T)` appends file k's segments and returns nothing, the slot is closed
// here. Files are assembled in groups that fit `budget` bytes of HBM.
template <class Fill>
static void emit_files(sp_ctx* c, VcfDeviceState& S, const std::vector<std::string>& paths, Fill fill) {
    const int64_t budget = 2ll << 30;
    FileSink& sink = file_sink(c);
    size_t k = 0;
    while (k < paths.size()) {
        Text t;
        SegTables T;
        std::vector<int64_t> lens;
        int64_t mx = 0, sum = 0;
        const size_t k0 = k;
        while (k < paths.size()) {
            fill(k, t, T);
            const int64_t n = T.end_slot();
            lens.push_back(n); mx = std::max(mx, n); sum += n;
            k++;
            const int64_t stride_now = std::max<int64_t>(TILE_BYTES, (mx + TILE_BYTES - 1) / TILE_BYTES * TILE_BYTES);
            if (stride_now * (int64_t)(k - k0 + 1) > budget) break;
        }
        const int64_t stride = std::max<int64_t>(TILE_BYTES, (mx + TILE_BYTES - 1) / TILE_BYTES * TILE_BYTES);
        sink.drain();   // the previous group's slots are about to be overwritten
        run_copy(c, S, T, S.d_records.p, S.d_records.p, t.s, S.d_files, stride, (int)(k - k0));
        for (size_t q = k0; q < k; q++) sink.write_file(paths[q], S.d_files.p + (int64_t)(q - k0) * stride, lens[q - k0], c->stream);
        S.bytes_out += sum;
    }
    sink.finish();
}

// buildVCF (:1005-1067) with the bytes gathered on the device. write_joint: groundTruth_joint.vcf (needs every haplotype's
// list); individuals: the pairs (i, i+1), i even, inside [h_first, h_last], lists from a CSR whose row 0 is haplotype `base`.
void write_vcfs_device(sp_ctx* c, const PopulationState& P, bool write_joint, bool individuals, int32_t h_first, int32_t h_last,
                       const int64_t* off, const int32_t* ids, int32_t base) {
    const int32_t L = (int32_t)c->ref_len, V = P.n_variants;
    VcfDeviceState S;
    build_record_blob(c, P, S, L);
    auto record = [&](SegTables& T, int32_t v) { T.add(SEG_REF_FWD, S.rec_off[v], S.rec_off[v + 1] - S.rec_off[v]); };
    if (write_joint) {
        const int32_t H = P.n_haplotypes_total;
        // variant2haplotypes (:967-985): transpose, haplotypes ascending
        std::vector<int64_t> voff((size_t)V + 1, 0);
        for (int64_t k = 0; k < (int64_t)P.all_ids.size(); k++) voff[P.all_ids[k] + 1]++;
        for (int32_t v = 0; v < V; v++) voff[v + 1] += voff[v];
        std::vector<int32_t> v2h(P.all_ids.size());
        std::vector<int64_t> cur(voff.begin(), voff.end() - 1);
        for (int32_t h = 0; h < H; h++)
            for (int64_t k = P.all_off[h]; k < P.all_off[h + 1]; k++) v2h[cur[P.all_ids[k]]++] = h;
        emit_files(c, S, {P.output_dir + "/groundTruth_joint.vcf"}, [&](size_t, Text& t, SegTables& T) {
            int64_t a = t.mark();
            header_text(t, L, false, "");
            T.add(SEG_AUX, a, t.mark() - a);
            for (int32_t v = 0; v < V; v++) {
                if (!P.variants[v].isActive) continue;
                record(T, v);
                a = t.mark();
                t.s += ";HAPLOTYPES=";
                for (int64_t k = voff[v]; k < voff[v + 1]; k++) {
                    if (k > voff[v]) t.s += ',';
                    t.i(v2h[k]);
                }
                t.s += '\n';
                T.add(SEG_AUX, a, t.mark() - a);
            }
        });
    }
    if (individuals) {
        std::vector<std::string> paths;
        std::vector<int32_t> first_hap;
        for (int32_t i = h_first + (h_first & 1); i + 1 <= h_last; i += 2) {
            paths.push_back(P.output_dir + "/groundTruth_individual_" + std::to_string(i) + ".vcf");
            first_hap.push_back(i);
        }
        emit_files(c, S, paths, [&](size_t k, Text& t, SegTables& T) {
            const int32_t i = first_hap[k];
            int64_t a = t.mark();
            header_text(t, L, true, "individual_" + std::to_string(i));
            T.add(SEG_AUX, a, t.mark() - a);
            const int64_t gt = t.mark();   // the three genotype endings of this file, 8 bytes each
            t.s += "\tGT\t1|0\n\tGT\t0|1\n\tGT\t1|1\n";
            auto emit = [&](int32_t v, int which) { record(T, v); T.add(SEG_AUX, gt + 8 * which, 8); };
            const int32_t* x = ids + off[i - base];
            const int32_t* y = ids + off[i + 1 - base];
            const int64_t nx = off[i + 1 - base] - off[i - base], ny = off[i + 2 - base] - off[i + 1 - base];
            int64_t j = 0, q = 0;
            while (j < nx && q < ny) {   // the merge of :1031-1062
                if (x[j] < y[q]) emit(x[j++], 0);
                else if (y[q] < x[j]) emit(y[q++], 1);
                else { emit(x[j], 2); j++; q++; }
            }
            while (j < nx) emit(x[j++], 0);
            while (q < ny) emit(y[q++], 1);
        });
    }
}

}  // namespace sp
