/*
 * libsynthpop_cuda.so — C ABI of the B200-native population-simulation hot path.
 *
 * The reference (fabio-cunial/population-sv-simulator) has no FFI: its boundary is three Java CLIs with
 * positional argv, the static methods other classes call, and on-disk formats. Every entry
 * point below names the reference interface it replaces (file:line relative to the reference root). A Java
 * host (JNI / Panama FFM) or the C++ CLI host in population-sv-simulator_b200/host binds exactly these.
 *
 * Conventions: plain pointers and sizes only; every function returns int (0 = ok, 3 = SampleReadsFromLengthBins
 * "coverage unreachable" [src/SampleReadsFromLengthBins.java:64-67], any other non-zero = error, message via
 * sp_last_error). The library owns all device memory inside sp_ctx; the caller owns host buffers it passes.
 * One sp_ctx per host thread; calls are blocking unless stated. There is no CPU fallback: without a CUDA
 * device sp_create fails.
 */
#ifndef SYNTHPOP_CUDA_H
#define SYNTHPOP_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sp_ctx sp_ctx;

enum { SP_OK = 0, SP_ERR = 1, SP_ERR_COVERAGE = 3, SP_ERR_CUDA = 4, SP_ERR_ARG = 5, SP_ERR_JAVA_EXCEPTION = 6 };
enum { SP_DEL = 0, SP_DUP = 1, SP_INS = 2, SP_INV = 3 }; /* BuildModel.SV_TYPES, src/BuildModel.java:837 */

/* ---- context --------------------------------------------------------------------------------------- */
/* Replaces `private static Random random = new Random()` (src/SimulateHaplotypes.java:71): the seed is the one
 * documented deviation from the reference argv. device = CUDA ordinal. */
int sp_create(sp_ctx** out, int device, int64_t seed);
void sp_destroy(sp_ctx* ctx);
const char* sp_last_error(const sp_ctx* ctx);
const char* sp_version(void);
/* Launch on the caller's CUDA stream (cudaStream_t as void*); NULL = the context's own stream. */
int sp_set_stream(sp_ctx* ctx, void* cuda_stream);
/* Second stream for sp_population_genotypes (rows G/C/N): the genotype kernels and the serial collision fixer
 * (src/SimulateHaplotypes.java:617-690) of the next batch then run under sp_materialise of the current one, which stays
 * on the stream of sp_set_stream. Give this stream a HIGHER priority (cudaStreamCreateWithPriority): the hardware
 * dispatches every tile of the copy kernel before a later grid of equal priority. The call still returns with the
 * batch's lists complete. NULL = same stream. */
int sp_set_population_stream(sp_ctx* ctx, void* cuda_stream);
/* Number of kernels this library launched on ctx since creation (bench.py's gpu_launches). */
int64_t sp_kernel_launches(const sp_ctx* ctx);

/* Test hooks (no device work): the library's Double.toString (JDK >= 19 rules: "AF=", variants.txt) and
 * NumberFormat.getInstance() with 4 fraction digits ("REPEATS_FRACTION=", src/SimulateHaplotypes.java:1010-1011, :1278).
 * Return the string length. */
int sp_test_double_to_string(double x, char* out, int32_t capacity);
int sp_test_number_format4(double x, char* out, int32_t capacity);
/* Bandwidth probes behind the design of the tiled copy kernel (csrc/probe.cu, tools/probe_copy_variants.py): n_slots slots of
 * slot_bytes filled region-major from one shared source shifted by `misalign` bytes. kind 0 = stores only, 1 = the register
 * path of materialise_kernel, 2 = cp.async.bulk into shared memory under an mbarrier. Device milliseconds per launch. */
int sp_probe_bandwidth(sp_ctx* ctx, int32_t kind, int64_t slot_bytes, int32_t n_slots, int32_t misalign, int32_t reps, double* ms_per_launch);

/* ---- row A1: reference ------------------------------------------------------------------------------ */
/* SimulateHaplotypes.loadReference(String) (src/SimulateHaplotypes.java:119-134): skip one header line, trim and
 * concatenate the rest; bytes go to HBM once. */
int sp_load_reference_fasta(sp_ctx* ctx, const char* path);
/* Same, from bytes already in host memory (one byte per base, case preserved). */
int sp_set_reference(sp_ctx* ctx, const uint8_t* bases, int64_t length);
int64_t sp_reference_length(const sp_ctx* ctx);
/* The resident reference as a device pointer (read-only for the caller; valid until the reference is replaced), e.g. the
 * source of an NCCL broadcast to the other GPUs of a box. */
int sp_reference_device(sp_ctx* ctx, void** device_ptr, int64_t* length);
/* Makes room for a reference of `length` bytes that the caller writes ON THE DEVICE (the destination of that broadcast) and
 * returns where; the caller must have finished writing before the next call on this context. Replaces any reference and
 * drops any population, like sp_set_reference; no host copy exists afterwards (the VCF writers need one). */
int sp_reference_device_buffer(sp_ctx* ctx, int64_t length, void** device_ptr);

/* ---- row V0: variant table + per-haplotype lists ------------------------------------------------------ */
/* The `variants` array and `haplotype2variants[_last]` (src/SimulateHaplotypes.java:52-60) as SoA + CSR.
 * payload_off has n+1 entries into payload_blob (INS sequence, already oriented, :809-810); has_payload[i]==0
 * means sequence==null (an INS without payload splices the literal "null", as StringBuilder.append(null) does). */
int sp_set_variants(sp_ctx* ctx, int32_t n, const int8_t* sv_type, const int32_t* position, const int32_t* length,
                    const int64_t* payload_off, const uint8_t* payload_blob, const int8_t* has_payload);
int sp_set_haplotype_lists(sp_ctx* ctx, int32_t n_haplotypes, const int64_t* list_off, const int32_t* ids);
/* Read back what the context holds (host copies). Call once with NULL arrays to get the sizes. */
int sp_get_variants(sp_ctx* ctx, int32_t* n, int64_t* payload_bytes, int8_t* sv_type, int32_t* position, int32_t* length,
                    int64_t* payload_off, uint8_t* payload_blob);
int sp_get_haplotype_lists(sp_ctx* ctx, int32_t* n_haplotypes, int64_t* n_ids, int64_t* list_off, int32_t* ids);
/* SimulateHaplotypes.deserialize(String,String) (src/SimulateHaplotypes.java:1147-1178). */
int sp_load_population(sp_ctx* ctx, const char* haplotype2variants_file, const char* variants_file);

/* ---- row M: haplotype materialisation ----------------------------------------------------------------- */
/* SimulateHaplotypes.buildHaplotypes(first,last,outDir,buffer) (src/SimulateHaplotypes.java:917-960), split in
 * (1) plan = edit lists -> segment tables by a segmented exclusive scan, (2) materialise = tiled copy kernel.
 * Haplotypes [first..last] (inclusive, as the Java) become device-resident slots of `*slot_stride` bytes each.
 * fasta_framing!=0 lays out the exact bytes of haplotype_<i>.fa (">chr1_<i>\n" + sequence + "\n") in the slot.
 * out_lengths (host, last-first+1 entries): bytes of each slot that are meaningful (framed length). */
int sp_plan_haplotypes(sp_ctx* ctx, int32_t first, int32_t last, int32_t fasta_framing, int64_t* out_lengths,
                       int64_t* slot_stride);
int sp_materialise(sp_ctx* ctx); /* asynchronous on the context's stream */
int sp_synchronize(sp_ctx* ctx);
/* Device pointer/length of the sequence (framing excluded) of planned haplotype h: input of the read kernel. */
int sp_haplotype_device(sp_ctx* ctx, int32_t h, void** device_ptr, int64_t* sequence_length);
/* D2H of the framed bytes of haplotype h. */
int sp_copy_haplotype(sp_ctx* ctx, int32_t h, uint8_t* host_out, int64_t capacity, int64_t* written);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it). ONE haplotype, through the same plan and
 * copy kernels as sp_plan_haplotypes + sp_materialise, with every plan table in a file-scope device array of
 * about 3.1 MB (csrc/materialise.cu, `static __device__ __align__(256)`; every GPU that loads the library holds it until
 * exit). The call allocates nothing: no cudaMalloc, cudaFree or allocating library routine. Inputs: the reference and the
 * variant table resident in the context (sp_set_reference / sp_load_reference_fasta, sp_set_variants) and the haplotype's
 * edit list as n_ids ascending variant ids in DEVICE memory. Output: the haplotype's bases (no FASTA framing) into the
 * caller's 16-byte aligned DEVICE buffer of `capacity` bytes; *length = bytes written (set even when capacity is too small).
 * Limits: n_ids <= SP_STATIC_MAX_EDITS and a haplotype of at most SP_STATIC_MAX_OUTPUT_BYTES bytes, beyond which the call
 * fails with SP_ERR_ARG and a message naming the limit. Runs on the context's stream and returns after it has drained.
 * The array is shared by every context of the process on that GPU and keeps the last call's contents (what must start at
 * zero is cleared per call): calls that overlap on the GPU race — serialise them. */
#define SP_STATIC_MAX_EDITS 16384
#define SP_STATIC_MAX_OUTPUT_BYTES (512ll << 20)
int sp_materialise_haplotype_static(sp_ctx* ctx, const int32_t* device_ids, int32_t n_ids, void* device_out, int64_t capacity,
                                    int64_t* length);
/* plan + materialise + D2H of all of [first..last] into one host buffer, back to back (offsets: n+1 entries). */
int sp_build_haplotypes_host(sp_ctx* ctx, int32_t first, int32_t last, int32_t fasta_framing, uint8_t* host_out,
                             int64_t capacity, int64_t* offsets);
/* PrintHaplotypes.main (src/PrintHaplotypes.java:6-18): writes OUTPUT_DIR/haplotype_<i>.fa. */
int sp_print_haplotypes(sp_ctx* ctx, int32_t first, int32_t last, const char* reference_file,
                        const char* haplotype2variants_file, const char* variants_file, const char* output_dir);

/* ---- rows A2/A3/S1/G/C/N/I/R1/W/Z: SimulateHaplotypes ----------------------------------------------------- */
/* Positional argv of SimulateHaplotypes.main (src/SimulateHaplotypes.java:76-86), same order and meaning. */
typedef struct sp_population_args {
    const char* model_dir;              /* args[0]  BuildModel.deserializeStatistics directory */
    const char* population_name;        /* args[1]  ALL|AFR|AMR|EAS|EUR|OTH (resolved like Arrays.binarySearch, :101) */
    const char* reference_file;         /* args[2] */
    const char* repeat_intervals_file;  /* args[3] */
    const char* segdups_file;           /* args[4] */
    const char* insertion_strings_file; /* args[5] */
    int32_t n_insertion_strings;        /* args[6] */
    double allele_frequency_threshold;  /* args[7] */
    int32_t n_variants_frequent;        /* args[8] */
    int32_t n_individuals;              /* args[9] */
    const char* output_dir;             /* args[10] */
} sp_population_args;

/* SimulateHaplotypes.main, whole run on one GPU: writes groundTruth_joint.vcf, groundTruth_individual_<2k>.vcf,
 * haplotype2variants.txt, variants.txt, histogram_nVariantsPerHaplotype.txt, histogram_startPositions.txt into
 * output_dir (:114-115, :715, :1014, :1028). Afterwards the context holds the reference, the variant table with
 * payloads and every haplotype list, so sp_plan_haplotypes can follow without re-reading the files. */
int sp_simulate_population(sp_ctx* ctx, const sp_population_args* args);

/* The same run split at its one exchange point so that individuals can be sharded across GPUs
 * (SURVEY.md §8e). begin: loadReference .. sampleVariants (replicated, same seed => same variants on every rank),
 * then the genotype kernel + collision fixer + carrier counts for haplotypes [first_haplotype..last_haplotype] only
 * (loop A :621-634 is index-addressable: cell (j,i) uses LCG steps S1+2(j*H+i)+1,+2). */
int sp_population_begin(sp_ctx* ctx, const sp_population_args* args, int32_t first_haplotype, int32_t last_haplotype);
/* Rows G/C/N once more for haplotypes [first..last] of an already begun population (no file I/O): the fresh lists
 * stay in HBM and are what the next sp_plan_haplotypes consumes ("genotype sampling + materialisation" per batch). */
int sp_population_genotypes(sp_ctx* ctx, int32_t first_haplotype, int32_t last_haplotype);
/* n_variants = lastVariant+1, n_haplotypes = 2*N_INDIVIDUALS, local_entries = list entries of the local shard. */
int sp_population_info(sp_ctx* ctx, int32_t* n_variants, int32_t* n_haplotypes, int64_t* local_entries,
                       int32_t* n_local_haplotypes_with_collision);
/* Per-variant carrier counts of the local shard: host copy, and the device buffer (int32[n_variants]) that the
 * host's NCCL all-gather reads (the single inter-GPU exchange). */
int sp_population_local_counts(sp_ctx* ctx, int32_t* counts);
int sp_population_counts_device(sp_ctx* ctx, void** device_counts);
/* Local haplotype lists after collision fixing (CSR: off has last-first+2 entries). */
int sp_population_local_lists(sp_ctx* ctx, int64_t* off, int32_t* ids);
/* Lists before collision fixing = the raw genotype matrix (row G), for kernel-level parity tests. */
int sp_population_pre_collision_lists(sp_ctx* ctx, int64_t* off, int32_t* ids, int64_t ids_capacity, int64_t* n_ids);
/* Draw-stream bookkeeping of the replicated prologue. */
int sp_population_draw_log(sp_ctx* ctx, uint64_t* steps_after_sample_variants, int32_t* candidates);
/* Wall-clock seconds of the last begin/finish by row of the hot-path table: [0] A1 loadReference (:93), [1] A2 model parsers
 * (:95-97), [2] A3 loadPositions (:99), [3] S1 sampleVariants (:101), [4] G/C/N buildPopulation on the device (:104),
 * [5] I assignInsertionStrings (:107-109), [6] R1 annotateVariantsWithRepeats (:111), [7] W buildVCF (:113), [8] Z serialize
 * + histograms (:114-115). n_rows (may be NULL) = 9. */
int sp_population_timing(sp_ctx* ctx, double* row_seconds, int32_t capacity, int32_t* n_rows);
/* finish: AF := carriers/H and isActive from global_counts (NULL = the local counts are global), then
 * assignInsertionStrings, annotateVariantsWithRepeats, the individual VCFs of the local shard and, when
 * all_off/all_ids (CSR over ALL haplotypes) are given and write_joint_outputs!=0, the joint VCF, haplotype2variants.txt,
 * variants.txt and the histograms. write_joint_outputs: 0 = individual VCFs only, 1 = joint outputs too,
 * 2 = write no file at all (benchmarking rows G/C/N/I/R1 without file emission). */
int sp_population_finish(sp_ctx* ctx, const int32_t* global_counts, const int64_t* all_off, const int32_t* all_ids,
                         int32_t write_joint_outputs);

/* ---- row P: CCS-like reads (the reference's patched PBSIM2 in model mode, docker/pbsim2.cpp) --------------------- */
/* How a read finds its rand() draws. The reference has ONE glibc rand() stream per process and the index of a
 * read's first draw depends on every earlier read (docker/pbsim2.cpp:1641-1711).
 *   SP_READS_REPLAY ........... the caller supplies the reference's draw stream and the index of each read's first
 *                               draw (sp_reads_set_replay): output is byte-identical to the reference's file.
 *   SP_READS_PER_READ_STREAMS . production mode: read r of sequence s draws from its own glibc stream,
 *                               srand(read_seed(seed, s, r)); same algorithm, same distributions, reads independent
 *                               of one another (and of how they are sharded over warps or GPUs). */
enum { SP_READS_PER_READ_STREAMS = 0, SP_READS_REPLAY = 1 };

/* The long options of `pbsim` (docker/pbsim2.cpp:195-213) that model mode uses, as docker/haplotype2reads.sh:32 passes
 * them. A negative number (NULL string) means "option not given": set_sim_param's default applies (:1047-1107). */
typedef struct sp_read_args {
    double depth;           /* --depth            default 20.0 */
    double length_mean;     /* --length-mean      default 9000 */
    double length_sd;       /* --length-sd        default 7000 */
    int64_t length_min;     /* --length-min       default 100 */
    int64_t length_max;     /* --length-max       default 1000000 */
    double accuracy_mean;   /* --accuracy-mean    default 0.85; int(x*100)*0.01 is applied as at :1078 */
    double accuracy_min;    /* --accuracy-min     range-checked, unused in model mode */
    double accuracy_max;    /* --accuracy-max     range-checked, unused in model mode */
    int32_t sub_ratio, ins_ratio, del_ratio; /* --difference-ratio sub:ins:del, default 6:54:40 (sub_ratio<0) */
    const char* hmm_model;  /* --hmm_model        required (selects PROCESS_MODEL, :1175-1180) */
    const char* prefix;     /* --prefix           default "sd" */
    const char* id_prefix;  /* --id-prefix        default "S" */
    uint32_t seed;          /* --seed             (the reference defaults to time(NULL), :191) */
    int32_t no_error;       /* !=0: docker/pbsim2noError.cpp (prints the error-free oriented substring) */
    int32_t stream_mode;    /* SP_READS_* */
} sp_read_args;

/* set_sim_param + set_hmm_model + the lookup tables of simulate_by_model (:1047-1226, :2583-2647, :1455-1637),
 * built on the host with libm and uploaded as integers. pbsim's own error messages come back via sp_last_error. */
int sp_reads_configure(sp_ctx* ctx, const sp_read_args* args);
/* Replay input for the NEXT sp_reads_simulate_* call: draws[k] = k-th value rand() returned after srand(seed) in the
 * reference run, read_offsets[r] = index of read r's first draw (n_reads entries). */
int sp_reads_set_replay(sp_ctx* ctx, const int32_t* draws, int64_t n_draws, const int64_t* read_offsets, int64_t n_reads);
/* simulate_by_model (:1421-1788) + mutate (:2260-2413) for one sequence; reads stay in HBM as the exact records of
 * <prefix>_%04ld.fastq (">S<ref_num>_<n>Z<seq>\n"). Sources: a planned + materialised haplotype (never leaves the
 * device), a device buffer of the caller, or host bytes. Outputs (any may be NULL): number of reads simulated,
 * bytes of the file (dropped reads excluded), reads dropped for containing "NNNNN" (:1711), bases emitted. */
int sp_reads_simulate_haplotype(sp_ctx* ctx, int32_t h, int64_t ref_num, int64_t* n_reads, int64_t* file_bytes,
                                int64_t* n_dropped, int64_t* bases);
int sp_reads_simulate_device(sp_ctx* ctx, const void* device_seq, int64_t seq_len, int64_t ref_num, int64_t* n_reads,
                             int64_t* file_bytes, int64_t* n_dropped, int64_t* bases);
int sp_reads_simulate_sequence(sp_ctx* ctx, const uint8_t* seq, int64_t seq_len, int64_t ref_num, int64_t* n_reads,
                               int64_t* file_bytes, int64_t* n_dropped, int64_t* bases);
/* D2H of the file bytes of the last simulate call (dropped reads removed on the device first). */
int sp_reads_copy(sp_ctx* ctx, uint8_t* host_out, int64_t capacity, int64_t* written);
/* Per-read log of the last simulate call (n_reads entries each, any may be NULL): emitted length, dropped flag,
 * rand() calls consumed (in replay mode this must equal read_offsets[r+1]-read_offsets[r]). */
int sp_reads_log(sp_ctx* ctx, int32_t* length, uint8_t* dropped, uint32_t* draws_used);
/* Timing of the last simulate call: host milliseconds of the per-read plan (:1641-1655 for every read) and device
 * milliseconds of the read kernel (CUDA events on the launching stream). */
int sp_reads_timing(sp_ctx* ctx, double* plan_ms, double* kernel_ms);
/* Device view of the last simulate call: record buffer and the n_reads+1 record offsets (int64) into it. */
int sp_reads_device(sp_ctx* ctx, void** records, void** record_offsets, void** dropped_flags);
/* The whole `pbsim` command of docker/haplotype2reads.sh:53: every sequence of the FASTA, output
 * <prefix>_%04ld.fastq per sequence. (The <prefix>_%04ld.ref copy the reference also leaves is deleted by the
 * calling script at :54 and is not written.) */
int sp_simulate_reads(sp_ctx* ctx, const sp_read_args* args, const char* fasta_path);

/* ---- row T (extension): a Mendelian child ------------------------------------------------------------------------ */
/* Nothing in the reference derives a child (its trios are real families, src/GetTrioChildren.java:5-17), so this row
 * has no reference arithmetic to match; it is defined here. After sp_population_finish: the child
 * of individuals `father` and `mother` inherits one whole haplotype from each (no recombination), chosen by one
 * nextBoolean() per parent from java.util.Random(seed), father first: child haplotype 0 = haplotype 2*father + b_f,
 * child haplotype 1 = haplotype 2*mother + b_m. The two lists are appended to the context as haplotypes H and H+1
 * (returned in child_first_haplotype), so sp_plan_haplotypes / sp_materialise / sp_haplotype2reads apply unchanged and a
 * child can be a parent in turn. write_vcf != 0 writes OUTPUT_DIR/groundTruth_individual_<H>.vcf with the reference's
 * own buildVCF merge (src/SimulateHaplotypes.java:1027-1066). Outputs may be NULL. */
int sp_trio_add_child(sp_ctx* ctx, int32_t father_individual, int32_t mother_individual, int64_t seed, int32_t write_vcf,
                      int32_t* child_first_haplotype, int32_t* from_father_haplotype, int32_t* from_mother_haplotype);

/* ---- the tail of docker/haplotype2reads.sh and the coverage chunks of reads2svs.sh ------- */
/* `cat reads_ID2.fa >> reads_ID1.fa; shuf; tr 'Z' '\n'` (docker/haplotype2reads.sh:59-63) on the device.
 * begin: start a new individual. add: take the records of the last sp_reads_simulate_* call (ownership of the device
 * buffers moves, nothing is copied); call it once per haplotype, ID1 first. finish: permute the lines and write the
 * two-line FASTA records (">S1_<n>\n<seq>\n") into one device buffer. `shuf` is unseeded in the reference; here line k
 * of the concatenated file gets the key "state of java.util.Random(shuffle_seed) after k+1 steps" and the lines are
 * written in ascending key order. */
int sp_reads_mix_begin(sp_ctx* ctx);
int sp_reads_mix_add(sp_ctx* ctx);
/* docker/haplotype2reads.sh:53-58 for both haplotypes of an individual in one call: the pbsim runs of planned haplotypes h0
 * (ID1) and h1 (ID2) with BOTH read kernels in flight (the second one's CTAs fill the thinning tail of the first), then both
 * batches added to the mix in that order; equivalent to sp_reads_simulate_haplotype(h0) + sp_reads_mix_add +
 * sp_reads_simulate_haplotype(h1) + sp_reads_mix_add, record for record. Per-read streams only. seed_per_haplotype != 0:
 * haplotype h is sequenced with sp_haplotype_read_seed(configured seed, h), as sp_haplotype2reads does. The four outputs are arrays
 * of 2 (may be NULL); kernels_span_ms = first kernel's start to second kernel's end on the device. */
int sp_reads_mix_add_pair(sp_ctx* ctx, int32_t h0, int32_t h1, int64_t ref_num0, int64_t ref_num1, int32_t seed_per_haplotype,
                          int64_t* n_reads, int64_t* file_bytes, int64_t* n_dropped, int64_t* bases, double* kernels_span_ms);
int sp_reads_mix_finish(sp_ctx* ctx, int64_t shuffle_seed, int64_t* n_reads, int64_t* file_bytes);
/* The 1x chunks of docker/reads2svs.sh:65-74 (N_ROWS_1X = N_ROWS / MAX_COVERAGE, rounded up to a multiple of 4;
 * `split -d -l`) and the coverage sweep of :98-128 (coverage c = the first c chunks): rows per chunk, rows and bytes
 * of the prefix of the mixed file that makes up `coverage`. Outputs may be NULL. */
int sp_reads_mix_coverage_prefix(sp_ctx* ctx, int32_t max_coverage, int32_t coverage, int64_t* n_rows_1x, int64_t* n_rows,
                                 int64_t* n_bytes);
/* Bytes [offset, offset+n_bytes) of the mixed file -> host. */
int sp_reads_mix_copy(sp_ctx* ctx, int64_t offset, int64_t n_bytes, uint8_t* host_out);
/* Device view: the file bytes and the n_reads+1 byte offsets (int64) of its records. Outputs may be NULL. */
int sp_reads_mix_device(sp_ctx* ctx, void** file_bytes, void** read_offsets, int64_t* n_reads, int64_t* n_bytes);
/* The n_reads+1 record offsets -> host. */
int sp_reads_mix_offsets(sp_ctx* ctx, int64_t* read_offsets);
/* Device time of the last finish: ordering (keys, sort, scans, segment table) and the tiled copy. */
int sp_reads_mix_timing(sp_ctx* ctx, double* order_ms, double* copy_ms);
/* The pbsim seed sp_haplotype2reads uses for haplotype `haplotype` when sp_reads_configure was given `seed`. The reference
 * runs one pbsim PROCESS per haplotype, each seeded with its own time(NULL) (docker/pbsim2.cpp:191,
 * docker/haplotype2reads.sh:53-58): the two haplotypes of an individual, the individuals of a cohort and a child and its
 * parents all get independent sequencing errors and coverage fluctuations. Here that independence comes from this pure
 * function (splitmix64 finaliser of seed and haplotype id, mapped to [1, 2^31-2]); ref_num stays 1 so the record names
 * match the reference's. */
uint32_t sp_haplotype_read_seed(uint32_t seed, int32_t haplotype);
/* docker/haplotype2reads.sh:53-64 for the individual with haplotypes id1 and id1+1 (sp_reads_configure holds the pbsim
 * options): materialise both, pbsim on each (ref_num 1, as each is its own process there; seed
 * sp_haplotype_read_seed(seed, h) for haplotype h), mix, write output_file (NULL = keep it on the device only). */
int sp_haplotype2reads(sp_ctx* ctx, int32_t id1, int64_t shuffle_seed, const char* output_file, int64_t* n_reads,
                       int64_t* file_bytes);

/* ---- row D: SampleReadsFromLengthBins ------------------------------------------------------------------------------ */
/* Positional argv of SampleReadsFromLengthBins.main (src/SampleReadsFromLengthBins.java:20-30). args[0..4] (the two
 * normal distributions and the left weight) only feed buildCumulativeBimodal (:87-102), which needs commons-math3's
 * NormalDistribution; the host computes that table (N_BINS doubles, cumulative) and passes it, so that the native side
 * never re-implements erf. N_BINS = (MAX_READ_LENGTH+BIN_LENGTH-1)/BIN_LENGTH (:36). */
typedef struct sp_length_bin_args {
    const double* cumulative;               /* cumulativeBimodal[N_BINS] (:47) */
    int32_t bin_length;                     /* args[5] */
    int32_t max_read_length;                /* args[6] */
    const char* bins_prefix;                /* args[7]: files <prefix><i>.bin, 4-line FASTQ records */
    int64_t genome_length_each_haplotype;   /* args[8] */
    double target_coverage_each_haplotype;  /* args[9] */
    const char* output_fastq_file;          /* args[10]; <file>.histogram is written next to it (:78-80) */
} sp_length_bin_args;
/* The whole main (:19-81) with `new Random(seed)` of the context: returns 0, or SP_ERR_COVERAGE (= 3, the reference's
 * System.exit(3), :64-67) when N_BINS*10 consecutive attempts hit bins that ran dry. */
int sp_sample_reads_from_bins(sp_ctx* ctx, const sp_length_bin_args* args);
/* The same in steps, with the bins resident in HBM. load: reads <prefix><i>.bin, i in [0,n_bins). set: the bin files
 * back to back in host memory (bin_off: n_bins+1 byte offsets). */
int sp_length_bins_load(sp_ctx* ctx, const char* bins_prefix, int32_t n_bins);
int sp_length_bins_set(sp_ctx* ctx, int32_t n_bins, const uint8_t* blob, const int64_t* bin_off);
/* The sampling loop (:53-76) until `coverage >= target_coverage_bp`; the FASTQ bytes stay on the device. Outputs (any may
 * be NULL): reads written, bytes of the FASTQ, nextDouble() draws consumed, sequence bases, histogram[n_bins]. */
int sp_length_bins_sample(sp_ctx* ctx, const double* cumulative, int32_t n_bins, int64_t target_coverage_bp, int64_t* n_reads,
                          int64_t* fastq_bytes, int64_t* attempts, int64_t* bases, int32_t* histogram);
int sp_length_bins_copy(sp_ctx* ctx, uint8_t* host_out, int64_t capacity, int64_t* written);
/* Device milliseconds of the last sample call: draw/rank/scan/emit kernels, and the tiled gather-copy. */
int sp_length_bins_timing(sp_ctx* ctx, double* sample_ms, double* copy_ms);

/* ---- AnnotateVCF ------------------------------------------------------------------------------ */
/* Positional argv of AnnotateVCF.main (src/AnnotateVCF.java:96-102). Every DEL/DUP/INS/INV record of every listed VCF gets
 * ";REPEATS_START=t;REPEATS_END=t;REPEATS_FRACTION=x.xxxx" appended to INFO (:195); other lines are copied. Output file =
 * the input path up to its first ".vcf" + "_annotated.vcf" (:119). */
typedef struct sp_annotate_args {
    const char* input_files_list;       /* args[0]: one VCF path per line */
    int32_t min_distance;               /* args[1] */
    int32_t max_distance;               /* args[2] */
    const char* reference_file;         /* args[3] */
    const char* repeat_intervals_file;  /* args[4] */
    const char* segdups_file;           /* args[5] */
} sp_annotate_args;
int sp_annotate_vcf(sp_ctx* ctx, const sp_annotate_args* args);
/* The same in steps. load = :112-115 (loadReference, deserializeRepeatIntervals, deserializeSegdups, loadPositions): the five
 * repeat-context position sets go to HBM as sorted disjoint intervals with prefix counts. file = annotateVCF (:127-205) for
 * one VCF: the host parses the text, one GPU thread per record answers getRepeatType (:437-457) twice and getRepeatSurface
 * (:480-495) by rank queries. n_annotated (may be NULL) = records that received the three fields. */
int sp_annotate_load(sp_ctx* ctx, const char* reference_file, const char* repeat_intervals_file, const char* segdups_file);
int sp_annotate_file(sp_ctx* ctx, const char* input_vcf, const char* output_vcf, int32_t min_distance, int32_t max_distance,
                     int64_t* n_annotated);
/* Seconds of the last load, device milliseconds of the last file's kernel, records annotated since the load. */
int sp_annotate_timing(sp_ctx* ctx, double* load_seconds, double* last_kernel_ms, int64_t* records_total);

/* ---- BuildReadLengthBins, Fastq2LengthHistogram ------------------------------------------------ */
/* Positional argv of BuildReadLengthBins.main (src/BuildReadLengthBins.java:21-27). Writes <prefix><i>.bin for i in
 * [0, N_BINS), N_BINS = (MAX_READ_LENGTH+BIN_LENGTH-1)/BIN_LENGTH (:28): the records of all listed FASTQ files whose
 * sequence length falls in bin i (the last bin takes everything longer, :57-58), permuted by the reference's Fisher-Yates
 * with `new Random(seed)` of the context shared by all bins (:95-99, :108-116; bins with fewer than three records keep their
 * input order and draw nothing); <prefix>.histogram (:77-84) and <prefix>.stats (printStats :139-208). These are the
 * files sp_sample_reads_from_bins / sp_length_bins_load consume. */
typedef struct sp_build_bins_args {
    const char* input_files_list;    /* args[0]: one FASTQ path per line */
    int32_t bin_length;              /* args[1] */
    int32_t max_read_length;         /* args[2] */
    int64_t genome_length_haploid;   /* args[3] */
    const char* output_prefix;       /* args[4] */
} sp_build_bins_args;
int sp_build_read_length_bins(sp_ctx* ctx, const sp_build_bins_args* args);
/* Fastq2LengthHistogram.main (src/Fastq2LengthHistogram.java:11-51): FASTQ_FILE BIN_LENGTH MAX_READ_LENGTH
 * OUTPUT_FILE_HISTOGRAM OUTPUT_FILE_MAX. */
int sp_fastq2length_histogram(sp_ctx* ctx, const char* fastq_file, int32_t bin_length, int32_t max_read_length,
                              const char* output_file_histogram, const char* output_file_max);
/* Device milliseconds of the last sp_build_read_length_bins: line index + binning, draws + swap forest, gather copy; records
 * and bytes written. Outputs may be NULL. */
int sp_build_bins_timing(sp_ctx* ctx, double* index_ms, double* shuffle_ms, double* copy_ms, int64_t* n_records, int64_t* n_bytes);

/* BuildModel.main (src/BuildModel.java:21-47; buildModel/buildModel.sh:16): REPEATMASKER_FILE NROWS TRF_FILE NROWS SEGDUPS_FILE
 * NROWS REFERENCE_FAI REFERENCE_CYTOBANDS GNOMADSV_FILE ONLY_PASS OUTPUT_DIR VERBOSE. Writes into output_dir the files
 * SimulateHaplotypes loads (serializeStatistics :1135-1197, serializeRepeatIntervals :446-458, serializeSegdups :563-577):
 * statistics_svType.txt, statistics_intervalType.txt, statistics_length.txt, statistics_alleleFrequency_{ALL,AFR,AMR,EAS,EUR,OTH}.txt,
 * statistics_position.txt, repeatIntervals.txt, segdups.txt. The record histograms (collectStatistics :933-1020) and the
 * allele-frequency sorts (:1079) run on the device; the carried-state scans and every floating-point sum keep the Java's order. */
typedef struct sp_build_model_args {
    const char* repeatmasker_file;   /* args[0]: cleaned RepeatMasker table, comma-separated */
    int32_t repeatmasker_rows;       /* args[1]: capacity the Java allocates; more rows than this is an exception */
    const char* trf_file;            /* args[2]: cleaned TRF table, comma-separated */
    int32_t trf_rows;                /* args[3] */
    const char* segdups_file;        /* args[4]: UCSC segmental duplications, tab-separated with a header line */
    int32_t segdups_rows;            /* args[5]: must equal the number of data rows */
    const char* reference_fai;       /* args[6] */
    const char* reference_cytobands; /* args[7] */
    const char* gnomad_vcf;          /* args[8] */
    int32_t only_pass;               /* args[9] == 1 */
    const char* output_dir;          /* args[10] */
    int32_t verbose;                 /* args[11] == 1 (the Java passes false to the loaders; only the chr1 summary line is printed) */
} sp_build_model_args;
int sp_build_model(sp_ctx* ctx, const sp_build_model_args* args);
/* Of the last sp_build_model of this process: host seconds of the loaders and the record scan, device milliseconds of the
 * histogram + sort launches, host seconds of the floating-point passes and the file writes; records counted and allele
 * frequencies sorted. Outputs may be NULL. */
int sp_build_model_timing(sp_ctx* ctx, double* parse_seconds, double* device_ms, double* finish_seconds, int64_t* n_records, int64_t* n_allele_frequencies);

#ifdef __cplusplus
}
#endif
#endif
