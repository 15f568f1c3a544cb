import java.lang.foreign.*;

/** Drop-in for src/SimulateHaplotypes.java `main`: same class name, same 11 positional arguments (:76-86), same output
 *  files; the work happens in libsynthpop_cuda.so. */
public class SimulateHaplotypes {
    public static void main(String[] args) throws Throwable {
        try (Arena arena = Arena.ofConfined()) {
            MemorySegment ctx = SynthpopNative.create(arena);
            MemorySegment a = arena.allocate(SynthpopNative.POPULATION_ARGS);
            long[] off = new long[11];
            String[] names = {"model_dir", "population_name", "reference_file", "repeat_intervals_file", "segdups_file",
                              "insertion_strings_file", "n_insertion_strings", "allele_frequency_threshold",
                              "n_variants_frequent", "n_individuals", "output_dir"};
            for (int i = 0; i < 11; i++) off[i] = SynthpopNative.POPULATION_ARGS.byteOffset(MemoryLayout.PathElement.groupElement(names[i]));
            for (int i : new int[] {0, 1, 2, 3, 4, 5, 10}) a.set(ValueLayout.ADDRESS, off[i], arena.allocateFrom(args[i]));
            a.set(ValueLayout.JAVA_INT, off[6], Integer.parseInt(args[6]));
            a.set(ValueLayout.JAVA_DOUBLE, off[7], Double.parseDouble(args[7]));
            a.set(ValueLayout.JAVA_INT, off[8], Integer.parseInt(args[8]));
            a.set(ValueLayout.JAVA_INT, off[9], Integer.parseInt(args[9]));
            SynthpopNative.check(ctx, (int) SynthpopNative.sp_simulate_population.invokeExact(ctx, a));
            SynthpopNative.sp_destroy.invokeExact(ctx);
        }
    }
}
