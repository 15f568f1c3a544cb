"""Row D on the GPU: SampleReadsFromLengthBins through the C ABI against the CPU restatement
(oracle/sample_reads_from_bins.cpp, src/SampleReadsFromLengthBins.java:19-81) on the same seed, bin files and
cumulative table: the FASTQ and the .histogram must be byte-identical, and the exit status 3 must be reproduced."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _argv(prefix, out, genome, cov, bin_length=500, max_len=35000):
    return ["9000", "1500", "15000", "2000", "0.3", str(bin_length), str(max_len), prefix, str(genome), str(cov), out]


def _cdf_file(path, cum):
    with open(path, "w") as f:
        for v in cum:
            f.write(repr(float(v)) + "\n")


def _run_oracle(oracle_bin, seed, cdf_path, argv):
    return subprocess.run([oracle_bin, "--seed", str(seed), "--cdf", cdf_path, "SampleReadsFromLengthBins"] + argv,
                          stderr=subprocess.DEVNULL).returncode


@pytest.mark.parametrize("seed,per_bin,genome,cov", [(42, 40, 2_000_000, 2.0), (7, 300, 20_000_000, 1.5), (123, 3, 100_000, 1.0)])
def test_fastq_and_histogram_byte_identical(sp, oracle_bin, tmp_path, seed, per_bin, genome, cov):
    from synthpop import synth
    n_bins, bl = 70, 500
    prefix = str(tmp_path / "bins_")
    synth.write_length_bins(prefix, n_bins, bl, per_bin, seed=seed)
    cum = synth.cumulative_bimodal(n_bins, bl, 9000, 1500, 15000, 2000, 0.3)
    cdf = str(tmp_path / "cdf.txt"); _cdf_file(cdf, cum)
    want_out, got_out = str(tmp_path / "want.fastq"), str(tmp_path / "got.fastq")
    rc = _run_oracle(oracle_bin, seed, cdf, _argv(prefix, want_out, genome, cov))
    with sp.Context(0, seed) as ctx:
        if rc == 3:
            with pytest.raises(sp.SynthpopError) as e:
                ctx.sample_reads_from_length_bins(_argv(prefix, got_out, genome, cov), cum)
            assert e.value.code == 3
            return
        assert rc == 0
        ctx.sample_reads_from_length_bins(_argv(prefix, got_out, genome, cov), cum)
    assert open(got_out, "rb").read() == open(want_out, "rb").read()
    assert open(got_out + ".histogram", "rb").read() == open(want_out + ".histogram", "rb").read()


def test_bins_run_dry_exit_status_3(sp, oracle_bin, tmp_path):
    from synthpop import synth
    n_bins, bl = 20, 500
    prefix = str(tmp_path / "b_")
    synth.write_length_bins(prefix, n_bins, bl, 2, seed=5)
    cum = synth.cumulative_bimodal(n_bins, bl, 4000, 800, 7000, 900, 0.5)
    cdf = str(tmp_path / "cdf.txt"); _cdf_file(cdf, cum)
    argv = _argv(prefix, str(tmp_path / "o.fastq"), 10_000_000, 5.0, bl, n_bins * bl)
    assert _run_oracle(oracle_bin, 9, cdf, argv) == 3
    with sp.Context(0, 9) as ctx:
        with pytest.raises(sp.SynthpopError) as e:
            ctx.sample_reads_from_length_bins(argv, cum)
    assert e.value.code == sp.SP_ERR_COVERAGE


def test_ragged_bins_and_resident_api(sp, oracle_bin, tmp_path):
    """Empty bins, a one-record bin, zero target; the device-resident API returns the same bytes as the file API."""
    from synthpop import synth
    n_bins, bl = 30, 400
    per = [0 if b % 3 == 0 else (1 if b % 3 == 1 else 25) for b in range(n_bins)]
    prefix = str(tmp_path / "r_")
    synth.write_length_bins(prefix, n_bins, bl, per, seed=77, with_names=False)
    cum = synth.cumulative_bimodal(n_bins, bl, 5000, 2500, 9000, 1500, 0.6)
    cdf = str(tmp_path / "cdf.txt"); _cdf_file(cdf, cum)
    want_out = str(tmp_path / "w.fastq")
    argv = _argv(prefix, want_out, 150_000, 0.5, bl, n_bins * bl)
    rc = _run_oracle(oracle_bin, 31, cdf, argv)
    assert rc == 0
    want = np.fromfile(want_out, dtype=np.uint8)
    with sp.Context(0, 31) as ctx:
        ctx.length_bins_load(prefix, n_bins)
        res = ctx.length_bins_sample(cum, int(150_000 * (0.5 * 2)))
        got = ctx.length_bins_copy(res["fastq_bytes"])
        assert np.array_equal(got, want)
        assert res["n_reads"] == int(res["histogram"].sum()) and res["bases"] >= 150_000
        zero = ctx.length_bins_sample(cum, 0)
        assert zero["n_reads"] == 0 and zero["fastq_bytes"] == 0
