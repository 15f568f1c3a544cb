import java.lang.foreign.*;

/** Drop-in for src/Fastq2LengthHistogram.java: FASTQ_FILE BIN_LENGTH MAX_READ_LENGTH OUTPUT_FILE_HISTOGRAM OUTPUT_FILE_MAX (:11-16). */
public class Fastq2LengthHistogram {
    public static void main(String[] args) throws Throwable {
        try (Arena arena = Arena.ofConfined()) {
            MemorySegment ctx = SynthpopNative.create(arena);
            int rc = (int) SynthpopNative.sp_fastq2length_histogram.invokeExact(ctx, arena.allocateFrom(args[0]), Integer.parseInt(args[1]),
                    Integer.parseInt(args[2]), arena.allocateFrom(args[3]), arena.allocateFrom(args[4]));
            SynthpopNative.check(ctx, rc);
            SynthpopNative.sp_destroy.invokeExact(ctx);
        }
    }
}
