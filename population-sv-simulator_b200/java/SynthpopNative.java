import java.lang.foreign.*;
import java.lang.invoke.MethodHandle;
import static java.lang.foreign.ValueLayout.*;

/**
 * Panama FFM (JDK 22+) binding of libsynthpop_cuda.so: one downcall handle per entry point of include/synthpop_cuda.h
 * that the command-line classes of this directory need. Not compiled in the build image (no JDK there); compile where a JDK
 * exists with `javac -cp commons-math3-3.6.1.jar *.java` and run with `--enable-native-access=ALL-UNNAMED
 * -Djava.library.path=<dir of libsynthpop_cuda.so>`.
 */
final class SynthpopNative {
    static final Linker LINKER = Linker.nativeLinker();
    static final SymbolLookup LIB = SymbolLookup.libraryLookup(System.mapLibraryName("synthpop_cuda"), Arena.global());

    private static MethodHandle h(String name, FunctionDescriptor d) {
        return LINKER.downcallHandle(LIB.find(name).orElseThrow(() -> new UnsatisfiedLinkError(name)), d);
    }

    static final MethodHandle sp_create = h("sp_create", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_LONG));
    static final MethodHandle sp_destroy = h("sp_destroy", FunctionDescriptor.ofVoid(ADDRESS));
    static final MethodHandle sp_last_error = h("sp_last_error", FunctionDescriptor.of(ADDRESS, ADDRESS));
    static final MethodHandle sp_simulate_population = h("sp_simulate_population", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    static final MethodHandle sp_print_haplotypes = h("sp_print_haplotypes",
            FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    static final MethodHandle sp_sample_reads_from_bins = h("sp_sample_reads_from_bins", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    // the read path without intermediate files: docker/haplotype2reads.sh:53-64 and the coverage prefixes of reads2svs.sh:65-128
    static final MethodHandle sp_load_reference_fasta = h("sp_load_reference_fasta", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    static final MethodHandle sp_load_population = h("sp_load_population", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    static final MethodHandle sp_reads_configure = h("sp_reads_configure", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    static final MethodHandle sp_haplotype2reads = h("sp_haplotype2reads",
            FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS));
    static final MethodHandle sp_reads_mix_coverage_prefix = h("sp_reads_mix_coverage_prefix",
            FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    static final MethodHandle sp_reads_mix_copy = h("sp_reads_mix_copy", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, JAVA_LONG, ADDRESS));
    // row T (extension): a Mendelian child of two simulated individuals
    static final MethodHandle sp_trio_add_child = h("sp_trio_add_child",
            FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, JAVA_LONG, JAVA_INT, ADDRESS, ADDRESS, ADDRESS));

    /** sp_population_args (include/synthpop_cuda.h): argv of SimulateHaplotypes.main, src/SimulateHaplotypes.java:76-86. */
    static final StructLayout POPULATION_ARGS = MemoryLayout.structLayout(
            ADDRESS.withName("model_dir"), ADDRESS.withName("population_name"), ADDRESS.withName("reference_file"),
            ADDRESS.withName("repeat_intervals_file"), ADDRESS.withName("segdups_file"), ADDRESS.withName("insertion_strings_file"),
            JAVA_INT.withName("n_insertion_strings"), MemoryLayout.paddingLayout(4),
            JAVA_DOUBLE.withName("allele_frequency_threshold"), JAVA_INT.withName("n_variants_frequent"),
            JAVA_INT.withName("n_individuals"), ADDRESS.withName("output_dir"));

    /** sp_length_bin_args: argv[5..10] of SampleReadsFromLengthBins.main + the cumulative table. */
    static final StructLayout LENGTH_BIN_ARGS = MemoryLayout.structLayout(
            ADDRESS.withName("cumulative"), JAVA_INT.withName("bin_length"), JAVA_INT.withName("max_read_length"),
            ADDRESS.withName("bins_prefix"), JAVA_LONG.withName("genome_length_each_haplotype"),
            JAVA_DOUBLE.withName("target_coverage_each_haplotype"), ADDRESS.withName("output_fastq_file"));

    // the rows either side of the path (SURVEY.md section 8f ranks 2-4)
    static final MethodHandle sp_annotate_vcf = h("sp_annotate_vcf", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    static final MethodHandle sp_build_read_length_bins = h("sp_build_read_length_bins", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    static final MethodHandle sp_fastq2length_histogram = h("sp_fastq2length_histogram",
            FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS));
    static final MethodHandle sp_build_model = h("sp_build_model", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));

    /** sp_annotate_args: argv of AnnotateVCF.main, src/AnnotateVCF.java:96-110. */
    static final StructLayout ANNOTATE_ARGS = MemoryLayout.structLayout(
            ADDRESS.withName("input_files_list"), JAVA_INT.withName("min_distance"), JAVA_INT.withName("max_distance"),
            ADDRESS.withName("reference_file"), ADDRESS.withName("repeat_intervals_file"), ADDRESS.withName("segdups_file"));

    /** sp_build_bins_args: argv of BuildReadLengthBins.main, src/BuildReadLengthBins.java:21-27. */
    static final StructLayout BUILD_BINS_ARGS = MemoryLayout.structLayout(
            ADDRESS.withName("input_files_list"), JAVA_INT.withName("bin_length"), JAVA_INT.withName("max_read_length"),
            JAVA_LONG.withName("genome_length_haploid"), ADDRESS.withName("output_prefix"));

    /** sp_build_model_args: argv of BuildModel.main, src/BuildModel.java:22-33 (an int after a pointer is followed by 4 bytes of padding). */
    static final StructLayout BUILD_MODEL_ARGS = MemoryLayout.structLayout(
            ADDRESS.withName("repeatmasker_file"), JAVA_INT.withName("repeatmasker_rows"), MemoryLayout.paddingLayout(4),
            ADDRESS.withName("trf_file"), JAVA_INT.withName("trf_rows"), MemoryLayout.paddingLayout(4),
            ADDRESS.withName("segdups_file"), JAVA_INT.withName("segdups_rows"), MemoryLayout.paddingLayout(4),
            ADDRESS.withName("reference_fai"), ADDRESS.withName("reference_cytobands"), ADDRESS.withName("gnomad_vcf"),
            JAVA_INT.withName("only_pass"), MemoryLayout.paddingLayout(4), ADDRESS.withName("output_dir"),
            JAVA_INT.withName("verbose"), MemoryLayout.paddingLayout(4));

    /** The one deviation from the reference argv: the RNG seed (the reference uses `new Random()`). */
    static long seed() {
        String s = System.getenv("SYNTHPOP_SEED");
        return s == null ? System.nanoTime() : Long.parseLong(s);
    }

    static MemorySegment create(Arena arena) throws Throwable {
        MemorySegment out = arena.allocate(ADDRESS);
        String dev = System.getenv("SYNTHPOP_DEVICE");
        int rc = (int) sp_create.invokeExact(out, dev == null ? 0 : Integer.parseInt(dev), seed());
        if (rc != 0) throw new IllegalStateException("sp_create failed: no usable CUDA device (there is no CPU fallback)");
        return out.get(ADDRESS, 0);
    }

    /** 0 = ok; 3 = System.exit(3) of SampleReadsFromLengthBins; anything else = the Java exception named in the message. */
    static void check(MemorySegment ctx, int rc) throws Throwable {
        if (rc == 0) return;
        MemorySegment msg = ((MemorySegment) sp_last_error.invokeExact(ctx)).reinterpret(Long.MAX_VALUE);
        String text = msg.getString(0);
        if (rc == 3) { System.err.println(text); System.exit(3); }
        throw new RuntimeException(text);
    }
}
