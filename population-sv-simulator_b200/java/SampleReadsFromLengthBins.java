import java.lang.foreign.*;
import org.apache.commons.math3.distribution.NormalDistribution;

/** Drop-in for src/SampleReadsFromLengthBins.java: same 11 positional arguments (:20-30). The cumulative bimodal table
 *  is computed HERE with commons-math3 exactly as buildCumulativeBimodal does (:87-102) and handed to the native side. */
public class SampleReadsFromLengthBins {
    public static void main(String[] args) throws Throwable {
        final double meanLeft = Double.parseDouble(args[0]), stdLeft = Double.parseDouble(args[1]);
        final double meanRight = Double.parseDouble(args[2]), stdRight = Double.parseDouble(args[3]);
        final double weightLeft = Double.parseDouble(args[4]), weightRight = 1.0 - weightLeft;
        final int binLength = Integer.parseInt(args[5]), maxReadLength = Integer.parseInt(args[6]);
        final int nBins = (maxReadLength + binLength - 1) / binLength;
        final NormalDistribution left = new NormalDistribution(meanLeft, stdLeft), right = new NormalDistribution(meanRight, stdRight);
        double[] c = new double[nBins];
        for (int i = 0; i < nBins; i++)
            c[i] = left.probability(i * binLength, (i + 1) * binLength) * weightLeft + right.probability(i * binLength, (i + 1) * binLength) * weightRight;
        double sum = 0.0;
        for (int i = 0; i < nBins; i++) sum += c[i];
        for (int i = 0; i < nBins; i++) c[i] /= sum;
        System.err.println("Sampling from the following distribution:");
        for (int i = 0; i < nBins; i++) System.err.println((i * binLength) + "," + c[i]);
        for (int i = 1; i < nBins; i++) c[i] += c[i - 1];
        try (Arena arena = Arena.ofConfined()) {
            MemorySegment ctx = SynthpopNative.create(arena);
            MemorySegment a = arena.allocate(SynthpopNative.LENGTH_BIN_ARGS);
            StructLayout L = SynthpopNative.LENGTH_BIN_ARGS;
            a.set(ValueLayout.ADDRESS, L.byteOffset(MemoryLayout.PathElement.groupElement("cumulative")), arena.allocateFrom(ValueLayout.JAVA_DOUBLE, c));
            a.set(ValueLayout.JAVA_INT, L.byteOffset(MemoryLayout.PathElement.groupElement("bin_length")), binLength);
            a.set(ValueLayout.JAVA_INT, L.byteOffset(MemoryLayout.PathElement.groupElement("max_read_length")), maxReadLength);
            a.set(ValueLayout.ADDRESS, L.byteOffset(MemoryLayout.PathElement.groupElement("bins_prefix")), arena.allocateFrom(args[7]));
            a.set(ValueLayout.JAVA_LONG, L.byteOffset(MemoryLayout.PathElement.groupElement("genome_length_each_haplotype")), Long.parseLong(args[8]));
            a.set(ValueLayout.JAVA_DOUBLE, L.byteOffset(MemoryLayout.PathElement.groupElement("target_coverage_each_haplotype")), Double.parseDouble(args[9]));
            a.set(ValueLayout.ADDRESS, L.byteOffset(MemoryLayout.PathElement.groupElement("output_fastq_file")), arena.allocateFrom(args[10]));
            SynthpopNative.check(ctx, (int) SynthpopNative.sp_sample_reads_from_bins.invokeExact(ctx, a));
            SynthpopNative.sp_destroy.invokeExact(ctx);
        }
    }
}
