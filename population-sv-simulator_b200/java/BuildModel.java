import java.lang.foreign.*;
import static java.lang.foreign.ValueLayout.*;

/**
 * Drop-in for src/BuildModel.java: REPEATMASKER_FILE NROWS TRF_FILE NROWS SEGDUPS_FILE NROWS REFERENCE_FAI REFERENCE_CYTOBANDS
 * GNOMADSV_FILE ONLY_PASS OUTPUT_DIR VERBOSE (:22-33).
 */
public class BuildModel {
    public static void main(String[] args) throws Throwable {
        try (Arena arena = Arena.ofConfined()) {
            MemorySegment ctx = SynthpopNative.create(arena);
            MemorySegment a = arena.allocate(SynthpopNative.BUILD_MODEL_ARGS);
            a.set(ADDRESS, 0, arena.allocateFrom(args[0]));
            a.set(JAVA_INT, 8, Integer.parseInt(args[1]));
            a.set(ADDRESS, 16, arena.allocateFrom(args[2]));
            a.set(JAVA_INT, 24, Integer.parseInt(args[3]));
            a.set(ADDRESS, 32, arena.allocateFrom(args[4]));
            a.set(JAVA_INT, 40, Integer.parseInt(args[5]));
            a.set(ADDRESS, 48, arena.allocateFrom(args[6]));
            a.set(ADDRESS, 56, arena.allocateFrom(args[7]));
            a.set(ADDRESS, 64, arena.allocateFrom(args[8]));
            a.set(JAVA_INT, 72, Integer.parseInt(args[9]) == 1 ? 1 : 0);
            a.set(ADDRESS, 80, arena.allocateFrom(args[10]));
            a.set(JAVA_INT, 88, Integer.parseInt(args[11]) == 1 ? 1 : 0);
            SynthpopNative.check(ctx, (int) SynthpopNative.sp_build_model.invokeExact(ctx, a));
            SynthpopNative.sp_destroy.invokeExact(ctx);
        }
    }
}
