#!/usr/bin/env python
"""Generates the golden vectors of row P by running the REFERENCE ITSELF: oracle/_ref/pbsim2 and pbsim2noError are the
reference's docker/pbsim2.cpp / docker/pbsim2noError.cpp compiled unmodified by oracle/Makefile. Run in the build
container (where /root/reference exists):  python tests/golden/make_golden.py
Each case stores the inputs (sequence, HMM model text, command line) and the reference's output file, gzip-compressed.
"""
import gzip
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "population-sv-simulator_b200"))
from synthpop import synth  # noqa: E402

CASES = {
    "ccs_hmm_seed42": dict(L=150_000, ref_seed=3, model="hmm", exe="pbsim2", seed=42,
                           opts=["--depth", "3", "--length-mean", "15000", "--length-sd", "2000", "--length-min", "5000",
                                 "--length-max", "30000", "--accuracy-mean", "0.995", "--accuracy-min", "0.99",
                                 "--accuracy-max", "1.0", "--difference-ratio", "6:50:54"]),
    "clr_empty_model_seed7": dict(L=120_000, ref_seed=5, model="empty", exe="pbsim2", seed=7,
                                  opts=["--depth", "2", "--length-mean", "9000", "--length-sd", "7000", "--length-min", "100",
                                        "--length-max", "1000000", "--accuracy-mean", "0.85", "--difference-ratio", "6:54:40"]),
    "noerror_hmm_seed1": dict(L=100_000, ref_seed=9, model="hmm", exe="pbsim2noError", seed=1,
                              opts=["--depth", "2", "--length-mean", "15000", "--length-sd", "2000", "--length-min", "5000",
                                    "--length-max", "30000", "--accuracy-mean", "0.995", "--accuracy-min", "0.99",
                                    "--accuracy-max", "1.0", "--difference-ratio", "6:50:54"]),
}


def main():
    for name, c in CASES.items():
        with tempfile.TemporaryDirectory() as tmp:
            bases, _, _ = synth.random_reference(c["L"], c["ref_seed"])
            fa = os.path.join(tmp, "hap.fa")
            synth.write_fasta(fa, ">chr1_0", bases)
            model = os.path.join(tmp, "m.model")
            if c["model"] == "hmm":
                synth.write_hmm_model(model, quality="phred")
            else:
                open(model, "w").close()
            exe = os.path.join(ROOT, "oracle", "_ref", c["exe"])
            subprocess.run([exe] + c["opts"] + ["--hmm_model", model, "--seed", str(c["seed"]), "--prefix",
                                                os.path.join(tmp, "out"), fa], check=True, stderr=subprocess.DEVNULL)
            d = os.path.join(HERE, name)
            os.makedirs(d, exist_ok=True)
            for src, dst in [(fa, "hap.fa.gz"), (model, "model.txt.gz"), (os.path.join(tmp, "out_0001.fastq"), "out_0001.fastq.gz")]:
                with open(src, "rb") as f, gzip.GzipFile(os.path.join(d, dst), "wb", mtime=0) as g:
                    g.write(f.read())
            json.dump(dict(exe=c["exe"], seed=c["seed"], opts=c["opts"], no_error=c["exe"] == "pbsim2noError"),
                      open(os.path.join(d, "case.json"), "w"), indent=1)
            print(name, os.path.getsize(os.path.join(d, "out_0001.fastq.gz")), "bytes")


if __name__ == "__main__":
    main()
