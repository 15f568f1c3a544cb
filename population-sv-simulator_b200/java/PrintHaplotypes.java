import java.lang.foreign.*;

/** Drop-in for src/PrintHaplotypes.java: FIRST LAST REFERENCE_FILE HAPLOTYPE2VARIANTS_FILE VARIANTS_FILE OUTPUT_DIR (:7-12). */
public class PrintHaplotypes {
    public static void main(String[] args) throws Throwable {
        try (Arena arena = Arena.ofConfined()) {
            MemorySegment ctx = SynthpopNative.create(arena);
            int rc = (int) SynthpopNative.sp_print_haplotypes.invokeExact(ctx, Integer.parseInt(args[0]), Integer.parseInt(args[1]),
                    arena.allocateFrom(args[2]), arena.allocateFrom(args[3]), arena.allocateFrom(args[4]), arena.allocateFrom(args[5]));
            SynthpopNative.check(ctx, rc);
            SynthpopNative.sp_destroy.invokeExact(ctx);
        }
    }
}
