import java.lang.foreign.*;
import static java.lang.foreign.ValueLayout.*;

/** Drop-in for src/AnnotateVCF.java: INPUT_FILES_LIST MIN_DISTANCE MAX_DISTANCE REFERENCE_FILE REPEAT_INTERVALS_FILE SEGDUPS_FILE (:96-110). */
public class AnnotateVCF {
    public static void main(String[] args) throws Throwable {
        try (Arena arena = Arena.ofConfined()) {
            MemorySegment ctx = SynthpopNative.create(arena);
            MemorySegment a = arena.allocate(SynthpopNative.ANNOTATE_ARGS);
            a.set(ADDRESS, 0, arena.allocateFrom(args[0]));
            a.set(JAVA_INT, 8, Integer.parseInt(args[1]));
            a.set(JAVA_INT, 12, Integer.parseInt(args[2]));
            a.set(ADDRESS, 16, arena.allocateFrom(args[3]));
            a.set(ADDRESS, 24, arena.allocateFrom(args[4]));
            a.set(ADDRESS, 32, arena.allocateFrom(args[5]));
            SynthpopNative.check(ctx, (int) SynthpopNative.sp_annotate_vcf.invokeExact(ctx, a));
            SynthpopNative.sp_destroy.invokeExact(ctx);
        }
    }
}
