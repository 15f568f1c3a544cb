import java.lang.foreign.*;
import static java.lang.foreign.ValueLayout.*;

/** Drop-in for src/BuildReadLengthBins.java: INPUT_FILES_LIST BIN_LENGTH MAX_READ_LENGTH GENOME_LENGTH_HAPLOID OUTPUT_PREFIX (:21-27). */
public class BuildReadLengthBins {
    public static void main(String[] args) throws Throwable {
        try (Arena arena = Arena.ofConfined()) {
            MemorySegment ctx = SynthpopNative.create(arena);
            MemorySegment a = arena.allocate(SynthpopNative.BUILD_BINS_ARGS);
            a.set(ADDRESS, 0, arena.allocateFrom(args[0]));
            a.set(JAVA_INT, 8, Integer.parseInt(args[1]));
            a.set(JAVA_INT, 12, Integer.parseInt(args[2]));
            a.set(JAVA_LONG, 16, Long.parseLong(args[3]));
            a.set(ADDRESS, 24, arena.allocateFrom(args[4]));
            SynthpopNative.check(ctx, (int) SynthpopNative.sp_build_read_length_bins.invokeExact(ctx, a));
            SynthpopNative.sp_destroy.invokeExact(ctx);
        }
    }
}
