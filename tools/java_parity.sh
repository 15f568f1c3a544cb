#!/usr/bin/env bash
# First thing to run on a box that has a JDK (none exists in the build image or on the GPU box, so the Java half of the
# oracle is pinned only by known answers of the JDK primitives and by invariants: DESIGN.md section 5).
#
# Compiles the REFERENCE's own src/*.java (a scratch copy; the one edit is the seed of `new Random()`), runs the path on the
# seeded synthetic inputs, and byte-compares every output with the oracle restatement; with --cli also with synthpop_cli.
#
#   tools/java_parity.sh /path/to/population-sv-simulator /path/to/commons-math3-3.6.1.jar [--cli]
set -euo pipefail
REF=${1:?reference checkout}; CM=${2:?commons-math3 jar}; WITH_CLI=${3:-}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
W=$(mktemp -d); trap 'rm -rf "$W"' EXIT
SEED=42
command -v javac >/dev/null || { echo "no javac on this box"; exit 2; }
mkdir -p "$W/src" "$W/classes" "$W/in" "$W/java" "$W/oracle" "$W/cli"
cp "$REF"/src/*.java "$W/src/"
# the documented deviation: a fixed seed instead of `new Random()` (src/SimulateHaplotypes.java:71, BuildReadLengthBins.java:34,
# SampleReadsFromLengthBins.java:41)
sed -i "s/new Random()/new Random(${SEED}L)/" "$W/src/SimulateHaplotypes.java" "$W/src/BuildReadLengthBins.java" "$W/src/SampleReadsFromLengthBins.java"
javac -nowarn -cp "$CM" -d "$W/classes" "$W"/src/*.java
make -C "$ROOT/oracle" -s all
ORACLE="$ROOT/oracle/_build/synthpop_oracle"
CLI="$ROOT/population-sv-simulator_b200/synthpop_cli"
python - "$ROOT" "$W" <<'PY'
import sys, os
root, w = sys.argv[1], sys.argv[2]
sys.path.insert(0, os.path.join(root, "population-sv-simulator_b200"))
from synthpop import synth
inp = synth.generate_population_inputs(os.path.join(w, "in"), 5_000_000, n_insertions=5000)
open(os.path.join(w, "argv.txt"), "w").write("\n".join(inp.simulate_argv("OUT", 6, n_frequent=800)))
synth.write_length_bins  # (bins come from BuildReadLengthBins below)
PY
mapfile -t ARGV < "$W/argv.txt"
run() {  # name, output dir, command...
  local name=$1 out=$2; shift 2
  local a=("${ARGV[@]}"); a[10]="$out"
  "$@" "${a[@]}"
}
JAVA=(java -Xmx24g -cp "$W/classes:$CM")
run sim "$W/java"   "${JAVA[@]}" SimulateHaplotypes
run sim "$W/oracle" "$ORACLE" --quiet --seed $SEED SimulateHaplotypes
[ -n "$WITH_CLI" ] && run sim "$W/cli" "$CLI" --seed $SEED SimulateHaplotypes
for d in java oracle ${WITH_CLI:+cli}; do
  mkdir -p "$W/$d/haps"
  case $d in
    java)   "${JAVA[@]}" PrintHaplotypes 0 11 "${ARGV[2]}" "$W/$d/haplotype2variants.txt" "$W/$d/variants.txt" "$W/$d/haps";;
    oracle) "$ORACLE" --quiet PrintHaplotypes 0 11 "${ARGV[2]}" "$W/$d/haplotype2variants.txt" "$W/$d/variants.txt" "$W/$d/haps";;
    cli)    "$CLI" PrintHaplotypes 0 11 "${ARGV[2]}" "$W/$d/haplotype2variants.txt" "$W/$d/variants.txt" "$W/$d/haps";;
  esac
done
# AnnotateVCF on the truth VCFs
for d in java oracle ${WITH_CLI:+cli}; do
  ls "$W/$d"/groundTruth_*.vcf > "$W/$d/list.txt"
  case $d in
    java)   "${JAVA[@]}" AnnotateVCF "$W/$d/list.txt" 10 100 "${ARGV[2]}" "${ARGV[3]}" "${ARGV[4]}";;
    oracle) "$ORACLE" AnnotateVCF "$W/$d/list.txt" 10 100 "${ARGV[2]}" "${ARGV[3]}" "${ARGV[4]}";;
    cli)    "$CLI" AnnotateVCF "$W/$d/list.txt" 10 100 "${ARGV[2]}" "${ARGV[3]}" "${ARGV[4]}";;
  esac
done
# BuildModel on the synthetic tracks and gnomAD-SV-shaped records
python - "$ROOT" "$W" <<'PY'
import sys, os
root, w = sys.argv[1], sys.argv[2]
sys.path.insert(0, os.path.join(root, "population-sv-simulator_b200"))
from synthpop import synth
r = synth.write_build_model_inputs(os.path.join(w, "bm_in"))
open(os.path.join(w, "bm_argv.txt"), "w").write("\n".join(str(x) for x in r["argv"]))
PY
mapfile -t BM < "$W/bm_argv.txt"
for d in java oracle ${WITH_CLI:+cli}; do
  mkdir -p "$W/$d/model"
  BM[10]="$W/$d/model"
  case $d in
    java)   "${JAVA[@]}" BuildModel "${BM[@]}";;
    oracle) "$ORACLE" BuildModel "${BM[@]}";;
    cli)    "$CLI" BuildModel "${BM[@]}";;
  esac
done
# BuildReadLengthBins on a synthetic FASTQ, then SampleReadsFromLengthBins from each implementation's own bins
python - "$W" <<'PY'
import sys, os
import numpy as np
w = sys.argv[1]
rng = np.random.default_rng(3)
acgt = np.frombuffer(b"ACGT", np.uint8)
with open(os.path.join(w, "reads.fastq"), "wb") as f:
    for k in range(20000):
        ln = max(60, int(rng.normal(1750, 350) if rng.random() < 0.45 else rng.normal(3750, 350)))
        f.write(b"@r%d\n" % k + acgt[rng.integers(0, 4, ln)].tobytes() + b"\n+\n" + rng.integers(53, 93, ln).astype(np.uint8).tobytes() + b"\n")
open(os.path.join(w, "fastq_list.txt"), "w").write(os.path.join(w, "reads.fastq") + "\n")
PY
for d in java oracle ${WITH_CLI:+cli}; do
  mkdir -p "$W/$d/bins"
  B=("$W/fastq_list.txt" 500 6000 3100000 "$W/$d/bins/b_")
  S=(1750 350 3750 350 0.45 500 6000 "$W/$d/bins/b_" 200000 3.0 "$W/$d/bins/sampled.fastq")
  case $d in
    java)   "${JAVA[@]}" BuildReadLengthBins "${B[@]}"; "${JAVA[@]}" SampleReadsFromLengthBins "${S[@]}";;
    oracle) "$ORACLE" --seed $SEED BuildReadLengthBins "${B[@]}"; "$ORACLE" --seed $SEED SampleReadsFromLengthBins "${S[@]}";;
    cli)    "$CLI" --seed $SEED BuildReadLengthBins "${B[@]}"; "$CLI" --seed $SEED SampleReadsFromLengthBins "${S[@]}";;
  esac
done
fail=0
for d in oracle ${WITH_CLI:+cli}; do
  while IFS= read -r f; do
    rel=${f#"$W/java/"}
    if ! cmp -s "$f" "$W/$d/$rel"; then echo "DIFFERS: $rel (java vs $d)"; fail=1; fi
  done < <(find "$W/java" -type f ! -name list.txt)
done
[ $fail -eq 0 ] && echo "java parity: every output of the reference matches (oracle${WITH_CLI:+, synthpop_cli})"
exit $fail
