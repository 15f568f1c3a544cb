"""ctypes binding of libsynthpop_cuda.so (the C ABI declared in include/synthpop_cuda.h).

Host-side mirror of the reference's static-method interface for the hot path
(SimulateHaplotypes.loadReference / deserialize / buildHaplotypes, PrintHaplotypes.main ...).
There is no CPU fallback: if the CUDA extension is missing, importing `load()` raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libsynthpop_cuda.so")

SP_OK, SP_ERR, SP_ERR_COVERAGE, SP_ERR_CUDA, SP_ERR_ARG, SP_ERR_JAVA_EXCEPTION = 0, 1, 3, 4, 5, 6
DEL, DUP, INS, INV = 0, 1, 2, 3

_lib = None


class SynthpopError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"[{code}] {message}")
        self.code = code
        self.message = message


class PopulationArgs(C.Structure):
    """sp_population_args: the positional argv of SimulateHaplotypes.main (src/SimulateHaplotypes.java:76-86)."""
    _fields_ = [("model_dir", C.c_char_p), ("population_name", C.c_char_p), ("reference_file", C.c_char_p),
                ("repeat_intervals_file", C.c_char_p), ("segdups_file", C.c_char_p),
                ("insertion_strings_file", C.c_char_p), ("n_insertion_strings", C.c_int32),
                ("allele_frequency_threshold", C.c_double), ("n_variants_frequent", C.c_int32),
                ("n_individuals", C.c_int32), ("output_dir", C.c_char_p)]

    @classmethod
    def from_argv(cls, argv):
        a = [str(x) for x in argv]
        if len(a) < 11:
            raise IndexError("SimulateHaplotypes takes 11 positional arguments")
        return cls(a[0].encode(), a[1].encode(), a[2].encode(), a[3].encode(), a[4].encode(), a[5].encode(),
                   int(a[6]), float(a[7]), int(a[8]), int(a[9]), a[10].encode())


READS_PER_READ_STREAMS, READS_REPLAY = 0, 1


class LengthBinArgs(C.Structure):
    """sp_length_bin_args: argv[5..10] of SampleReadsFromLengthBins.main + the cumulative table (src/SampleReadsFromLengthBins.java:20-47)."""
    _fields_ = [("cumulative", C.POINTER(C.c_double)), ("bin_length", C.c_int32), ("max_read_length", C.c_int32),
                ("bins_prefix", C.c_char_p), ("genome_length_each_haplotype", C.c_int64),
                ("target_coverage_each_haplotype", C.c_double), ("output_fastq_file", C.c_char_p)]


class ReadArgs(C.Structure):
    """sp_read_args: the long options of `pbsim` in model mode (docker/pbsim2.cpp:195-213); negative / None = not given."""
    _fields_ = [("depth", C.c_double), ("length_mean", C.c_double), ("length_sd", C.c_double),
                ("length_min", C.c_int64), ("length_max", C.c_int64), ("accuracy_mean", C.c_double),
                ("accuracy_min", C.c_double), ("accuracy_max", C.c_double), ("sub_ratio", C.c_int32),
                ("ins_ratio", C.c_int32), ("del_ratio", C.c_int32), ("hmm_model", C.c_char_p), ("prefix", C.c_char_p),
                ("id_prefix", C.c_char_p), ("seed", C.c_uint32), ("no_error", C.c_int32), ("stream_mode", C.c_int32)]

    @classmethod
    def from_options(cls, argv, no_error=False, stream_mode=READS_PER_READ_STREAMS):
        """Parse a pbsim command line (list of strings, the trailing FASTA is returned separately)."""
        a = cls(-1.0, -1.0, -1.0, -1, -1, -1.0, -1.0, -1.0, -1, -1, -1, None, None, None, 0, int(no_error), stream_mode)
        fasta = None
        it = iter([str(x) for x in argv])
        for opt in it:
            if not opt.startswith("--"):
                fasta = opt
                continue
            val = next(it)
            if opt == "--depth": a.depth = float(val)
            elif opt == "--length-mean": a.length_mean = float(val)
            elif opt == "--length-sd": a.length_sd = float(val)
            elif opt == "--length-min": a.length_min = int(val)
            elif opt == "--length-max": a.length_max = int(val)
            elif opt == "--accuracy-mean": a.accuracy_mean = float(val)
            elif opt == "--accuracy-min": a.accuracy_min = float(val)
            elif opt == "--accuracy-max": a.accuracy_max = float(val)
            elif opt == "--difference-ratio":
                a.sub_ratio, a.ins_ratio, a.del_ratio = (int(x) for x in val.split(":"))
            elif opt == "--hmm_model": a.hmm_model = val.encode()
            elif opt == "--prefix": a.prefix = val.encode()
            elif opt == "--id-prefix": a.id_prefix = val.encode()
            elif opt == "--seed": a.seed = int(val) & 0xFFFFFFFF
            else: raise ValueError(f"pbsim option {opt} is not part of model mode")
        return a, fasta


def _p(dtype):
    return np.ctypeslib.ndpointer(dtype=dtype, flags="C_CONTIGUOUS")


class AnnotateArgs(C.Structure):
    """sp_annotate_args: positional argv of AnnotateVCF.main (src/AnnotateVCF.java:96-102)."""
    _fields_ = [("input_files_list", C.c_char_p), ("min_distance", C.c_int32), ("max_distance", C.c_int32),
                ("reference_file", C.c_char_p), ("repeat_intervals_file", C.c_char_p), ("segdups_file", C.c_char_p)]


class BuildBinsArgs(C.Structure):
    """sp_build_bins_args: positional argv of BuildReadLengthBins.main (src/BuildReadLengthBins.java:21-27)."""
    _fields_ = [("input_files_list", C.c_char_p), ("bin_length", C.c_int32), ("max_read_length", C.c_int32),
                ("genome_length_haploid", C.c_int64), ("output_prefix", C.c_char_p)]


class BuildModelArgs(C.Structure):
    """sp_build_model_args: positional argv of BuildModel.main (src/BuildModel.java:22-33)."""
    _fields_ = [("repeatmasker_file", C.c_char_p), ("repeatmasker_rows", C.c_int32), ("trf_file", C.c_char_p), ("trf_rows", C.c_int32),
                ("segdups_file", C.c_char_p), ("segdups_rows", C.c_int32), ("reference_fai", C.c_char_p), ("reference_cytobands", C.c_char_p),
                ("gnomad_vcf", C.c_char_p), ("only_pass", C.c_int32), ("output_dir", C.c_char_p), ("verbose", C.c_int32)]


def load() -> C.CDLL:
    """dlopen the in-tree libsynthpop_cuda.so and declare every prototype of include/synthpop_cuda.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, cp = C.c_void_p, C.c_int32, C.c_int64, C.c_char_p
    P = C.POINTER
    sig = {
        "sp_create": (C.c_int, [P(vp), C.c_int, i64]),
        "sp_destroy": (None, [vp]),
        "sp_last_error": (cp, [vp]),
        "sp_version": (cp, []),
        "sp_set_stream": (C.c_int, [vp, vp]),
        "sp_set_population_stream": (C.c_int, [vp, vp]),
        "sp_kernel_launches": (i64, [vp]),
        "sp_trio_add_child": (C.c_int, [vp, i32, i32, i64, i32, vp, vp, vp]),
        "sp_reads_mix_begin": (C.c_int, [vp]),
        "sp_reads_mix_add": (C.c_int, [vp]),
        "sp_reads_mix_add_pair": (C.c_int, [vp, i32, i32, i64, i64, i32, P(i64), P(i64), P(i64), P(i64), P(C.c_double)]),
        "sp_reads_mix_finish": (C.c_int, [vp, i64, vp, vp]),
        "sp_reads_mix_coverage_prefix": (C.c_int, [vp, i32, i32, vp, vp, vp]),
        "sp_reads_mix_copy": (C.c_int, [vp, i64, i64, vp]),
        "sp_reads_mix_device": (C.c_int, [vp, vp, vp, vp, vp]),
        "sp_reads_mix_offsets": (C.c_int, [vp, vp]),
        "sp_reads_mix_timing": (C.c_int, [vp, vp, vp]),
        "sp_haplotype2reads": (C.c_int, [vp, i32, i64, cp, vp, vp]),
        "sp_haplotype_read_seed": (C.c_uint32, [C.c_uint32, i32]),
        "sp_test_double_to_string": (C.c_int, [C.c_double, cp, i32]),
        "sp_test_number_format4": (C.c_int, [C.c_double, cp, i32]),
        "sp_probe_bandwidth": (C.c_int, [vp, i32, i64, i32, i32, i32, P(C.c_double)]),
        "sp_load_reference_fasta": (C.c_int, [vp, cp]),
        "sp_set_reference": (C.c_int, [vp, vp, i64]),
        "sp_materialise_haplotype_static": (C.c_int, [vp, vp, i32, vp, i64, P(i64)]),
        "sp_reference_device": (C.c_int, [vp, P(C.c_void_p), P(i64)]),
        "sp_reference_device_buffer": (C.c_int, [vp, i64, P(C.c_void_p)]),
        "sp_reference_length": (i64, [vp]),
        "sp_set_variants": (C.c_int, [vp, i32, vp, vp, vp, vp, vp, vp]),
        "sp_set_haplotype_lists": (C.c_int, [vp, i32, vp, vp]),
        "sp_get_variants": (C.c_int, [vp, P(i32), P(i64), vp, vp, vp, vp, vp]),
        "sp_get_haplotype_lists": (C.c_int, [vp, P(i32), P(i64), vp, vp]),
        "sp_load_population": (C.c_int, [vp, cp, cp]),
        "sp_plan_haplotypes": (C.c_int, [vp, i32, i32, i32, vp, P(i64)]),
        "sp_materialise": (C.c_int, [vp]),
        "sp_synchronize": (C.c_int, [vp]),
        "sp_haplotype_device": (C.c_int, [vp, i32, P(vp), P(i64)]),
        "sp_copy_haplotype": (C.c_int, [vp, i32, vp, i64, P(i64)]),
        "sp_build_haplotypes_host": (C.c_int, [vp, i32, i32, i32, vp, i64, vp]),
        "sp_print_haplotypes": (C.c_int, [vp, i32, i32, cp, cp, cp, cp]),
        "sp_simulate_population": (C.c_int, [vp, P(PopulationArgs)]),
        "sp_population_begin": (C.c_int, [vp, P(PopulationArgs), i32, i32]),
        "sp_population_genotypes": (C.c_int, [vp, i32, i32]),
        "sp_population_info": (C.c_int, [vp, P(i32), P(i32), P(i64), P(i32)]),
        "sp_population_local_counts": (C.c_int, [vp, vp]),
        "sp_population_counts_device": (C.c_int, [vp, P(vp)]),
        "sp_population_local_lists": (C.c_int, [vp, vp, vp]),
        "sp_population_pre_collision_lists": (C.c_int, [vp, vp, vp, i64, P(i64)]),
        "sp_population_draw_log": (C.c_int, [vp, P(C.c_uint64), P(i32)]),
        "sp_population_finish": (C.c_int, [vp, vp, vp, vp, i32]),
        "sp_population_timing": (C.c_int, [vp, vp, i32, P(i32)]),
        "sp_reads_configure": (C.c_int, [vp, P(ReadArgs)]),
        "sp_reads_set_replay": (C.c_int, [vp, vp, i64, vp, i64]),
        "sp_reads_simulate_haplotype": (C.c_int, [vp, i32, i64, P(i64), P(i64), P(i64), P(i64)]),
        "sp_reads_simulate_device": (C.c_int, [vp, vp, i64, i64, P(i64), P(i64), P(i64), P(i64)]),
        "sp_reads_simulate_sequence": (C.c_int, [vp, vp, i64, i64, P(i64), P(i64), P(i64), P(i64)]),
        "sp_reads_copy": (C.c_int, [vp, vp, i64, P(i64)]),
        "sp_reads_log": (C.c_int, [vp, vp, vp, vp]),
        "sp_reads_device": (C.c_int, [vp, P(vp), P(vp), P(vp)]),
        "sp_reads_timing": (C.c_int, [vp, P(C.c_double), P(C.c_double)]),
        "sp_simulate_reads": (C.c_int, [vp, P(ReadArgs), cp]),
        "sp_sample_reads_from_bins": (C.c_int, [vp, P(LengthBinArgs)]),
        "sp_length_bins_load": (C.c_int, [vp, cp, i32]),
        "sp_length_bins_set": (C.c_int, [vp, i32, vp, vp]),
        "sp_length_bins_sample": (C.c_int, [vp, vp, i32, i64, P(i64), P(i64), P(i64), P(i64), vp]),
        "sp_length_bins_copy": (C.c_int, [vp, vp, i64, P(i64)]),
        "sp_length_bins_timing": (C.c_int, [vp, P(C.c_double), P(C.c_double)]),
        "sp_annotate_vcf": (C.c_int, [vp, P(AnnotateArgs)]),
        "sp_annotate_load": (C.c_int, [vp, cp, cp, cp]),
        "sp_annotate_file": (C.c_int, [vp, cp, cp, i32, i32, P(i64)]),
        "sp_annotate_timing": (C.c_int, [vp, P(C.c_double), P(C.c_double), P(i64)]),
        "sp_build_read_length_bins": (C.c_int, [vp, P(BuildBinsArgs)]),
        "sp_fastq2length_histogram": (C.c_int, [vp, cp, i32, i32, cp, cp]),
        "sp_build_bins_timing": (C.c_int, [vp, P(C.c_double), P(C.c_double), P(C.c_double), P(i64), P(i64)]),
        "sp_build_model": (C.c_int, [vp, P(BuildModelArgs)]),
        "sp_build_model_timing": (C.c_int, [vp, P(C.c_double), P(C.c_double), P(C.c_double), P(i64), P(i64)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    # optional entry points added by later rows; declared when present
    lib._declared = set(sig)
    _lib = lib
    return lib


def haplotype_read_seed(seed: int, haplotype: int) -> int:
    """The pbsim seed sp_haplotype2reads derives for one haplotype (the reference runs one time(NULL)-seeded process each)."""
    return int(load().sp_haplotype_read_seed(seed, haplotype))


def declared_symbols() -> list[str]:
    """Every `sp_*` function declared in include/synthpop_cuda.h (parsed from the header)."""
    import re
    hdr = os.path.join(os.path.dirname(os.path.dirname(_HERE)), "include", "synthpop_cuda.h")
    text = open(hdr).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sp_[a-z0-9_]+)\s*\(", text)))


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Context:
    """One sp_ctx. Method names follow the reference's static methods."""

    def __init__(self, device: int = 0, seed: int = 42):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.sp_create(C.byref(h), device, seed)
        if rc != SP_OK:
            raise SynthpopError(rc, "sp_create failed: no usable CUDA device (there is no CPU fallback)")
        self.h = h
        self._keep = []

    def close(self):
        if getattr(self, "h", None):
            self.lib.sp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc: int):
        if rc != SP_OK:
            raise SynthpopError(rc, (self.lib.sp_last_error(self.h) or b"").decode("utf-8", "replace"))

    # -- plumbing
    def set_stream(self, cuda_stream: int | None):
        self._check(self.lib.sp_set_stream(self.h, C.c_void_p(cuda_stream or 0)))

    def set_population_stream(self, cuda_stream: int | None):
        self._check(self.lib.sp_set_population_stream(self.h, C.c_void_p(cuda_stream or 0)))

    def probe_bandwidth(self, kind: int, slot_bytes: int, n_slots: int, misalign: int = 9, reps: int = 5) -> float:
        """Device ms per launch of one of the copy-kernel design probes (csrc/probe.cu)."""
        ms = C.c_double()
        self._check(self.lib.sp_probe_bandwidth(self.h, kind, slot_bytes, n_slots, misalign, reps, C.byref(ms)))
        return ms.value

    def synchronize(self):
        self._check(self.lib.sp_synchronize(self.h))

    @property
    def kernel_launches(self) -> int:
        return int(self.lib.sp_kernel_launches(self.h))

    # -- row A1
    def load_reference(self, path: str):
        """SimulateHaplotypes.loadReference (src/SimulateHaplotypes.java:119-134)."""
        self._check(self.lib.sp_load_reference_fasta(self.h, path.encode()))

    def set_reference(self, bases: np.ndarray):
        bases = np.ascontiguousarray(bases, dtype=np.uint8)
        self._check(self.lib.sp_set_reference(self.h, _ptr(bases), bases.size))

    def reference_device(self):
        """(device pointer, length) of the resident reference."""
        p, n = C.c_void_p(), C.c_int64()
        self._check(self.lib.sp_reference_device(self.h, C.byref(p), C.byref(n)))
        return int(p.value or 0), int(n.value)

    def reference_device_buffer(self, length: int) -> int:
        """Device pointer of a fresh reference buffer of `length` bytes for the caller to fill on the device."""
        p = C.c_void_p()
        self._check(self.lib.sp_reference_device_buffer(self.h, int(length), C.byref(p)))
        return int(p.value or 0)

    @property
    def reference_length(self) -> int:
        return int(self.lib.sp_reference_length(self.h))

    # -- row V0
    def set_variants(self, sv_type, position, length, payload_off=None, payload_blob=None, has_payload=None):
        sv_type = np.ascontiguousarray(sv_type, dtype=np.int8)
        position = np.ascontiguousarray(position, dtype=np.int32)
        length = np.ascontiguousarray(length, dtype=np.int32)
        n = sv_type.size
        if payload_off is None:
            payload_off = np.zeros(n + 1, dtype=np.int64)
            payload_blob = np.zeros(1, dtype=np.uint8)
        payload_off = np.ascontiguousarray(payload_off, dtype=np.int64)
        payload_blob = np.ascontiguousarray(payload_blob, dtype=np.uint8)
        hp = None if has_payload is None else np.ascontiguousarray(has_payload, dtype=np.int8)
        self._check(self.lib.sp_set_variants(self.h, n, _ptr(sv_type), _ptr(position), _ptr(length),
                                             _ptr(payload_off), _ptr(payload_blob), _ptr(hp)))

    def set_haplotype_lists(self, list_off, ids):
        list_off = np.ascontiguousarray(list_off, dtype=np.int64)
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        if ids.size == 0:
            ids = np.zeros(1, dtype=np.int32)
        self._check(self.lib.sp_set_haplotype_lists(self.h, list_off.size - 1, _ptr(list_off), _ptr(ids)))

    def get_variants(self):
        n, nb = C.c_int32(), C.c_int64()
        self._check(self.lib.sp_get_variants(self.h, C.byref(n), C.byref(nb), None, None, None, None, None))
        t = np.zeros(max(n.value, 1), np.int8); p = np.zeros(max(n.value, 1), np.int32); l = np.zeros(max(n.value, 1), np.int32)
        po = np.zeros(n.value + 1, np.int64); pb = np.zeros(max(nb.value, 1), np.uint8)
        self._check(self.lib.sp_get_variants(self.h, None, None, _ptr(t), _ptr(p), _ptr(l), _ptr(po), _ptr(pb)))
        return dict(sv_type=t[: n.value], position=p[: n.value], length=l[: n.value], payload_off=po, payload=pb[: max(nb.value, 1)])

    def get_haplotype_lists(self):
        h, n = C.c_int32(), C.c_int64()
        self._check(self.lib.sp_get_haplotype_lists(self.h, C.byref(h), C.byref(n), None, None))
        off = np.zeros(h.value + 1, np.int64); ids = np.zeros(max(n.value, 1), np.int32)
        self._check(self.lib.sp_get_haplotype_lists(self.h, None, None, _ptr(off), _ptr(ids)))
        return off, ids[: n.value]

    def deserialize(self, haplotype2variants_file: str, variants_file: str):
        """SimulateHaplotypes.deserialize (src/SimulateHaplotypes.java:1147-1178)."""
        self._check(self.lib.sp_load_population(self.h, haplotype2variants_file.encode(), variants_file.encode()))

    # -- row M
    def plan_haplotypes(self, first: int, last: int, fasta_framing: bool = False):
        n = max(0, last - first + 1)
        lens = np.zeros(max(n, 1), dtype=np.int64)
        stride = C.c_int64(0)
        self._check(self.lib.sp_plan_haplotypes(self.h, first, last, int(fasta_framing), _ptr(lens), C.byref(stride)))
        return lens[:n], int(stride.value)

    def materialise(self):
        self._check(self.lib.sp_materialise(self.h))

    def haplotype_device(self, h: int):
        p = C.c_void_p()
        n = C.c_int64()
        self._check(self.lib.sp_haplotype_device(self.h, h, C.byref(p), C.byref(n)))
        return int(p.value or 0), int(n.value)

    def materialise_haplotype_static(self, device_ids: int, n_ids: int, device_out: int, capacity: int) -> int:
        """sp_materialise_haplotype_static: one haplotype from a DEVICE edit list into a DEVICE buffer, no allocation (plan
        tables in the library's static device array). Pointers are integers (e.g. tensor.data_ptr()). Returns the length."""
        n = C.c_int64(0)
        self._check(self.lib.sp_materialise_haplotype_static(self.h, C.c_void_p(device_ids or 0), int(n_ids), C.c_void_p(device_out or 0),
                                                             int(capacity), C.byref(n)))
        return int(n.value)

    def copy_haplotype(self, h: int, capacity: int) -> np.ndarray:
        out = np.empty(capacity, dtype=np.uint8)
        w = C.c_int64()
        self._check(self.lib.sp_copy_haplotype(self.h, h, _ptr(out), capacity, C.byref(w)))
        return out[: w.value]

    def build_haplotypes_host(self, first: int, last: int, out: np.ndarray, fasta_framing: bool = False):
        """buildHaplotypes into one host buffer; returns offsets (n+1)."""
        n = last - first + 1
        offs = np.zeros(n + 1, dtype=np.int64)
        self._check(self.lib.sp_build_haplotypes_host(self.h, first, last, int(fasta_framing), _ptr(out), out.size, _ptr(offs)))
        return offs

    def print_haplotypes(self, first: int, last: int, reference_file: str, haplotype2variants_file: str,
                         variants_file: str, output_dir: str):
        """PrintHaplotypes.main (src/PrintHaplotypes.java:6-18)."""
        self._check(self.lib.sp_print_haplotypes(self.h, first, last, reference_file.encode(),
                                                 haplotype2variants_file.encode(), variants_file.encode(),
                                                 output_dir.encode()))

    # -- rows S1/G/C/N/I/R1/W/Z
    def simulate_haplotypes(self, argv):
        """SimulateHaplotypes.main(args) (src/SimulateHaplotypes.java:75-116): same positional argv."""
        args = PopulationArgs.from_argv(argv)
        self._check(self.lib.sp_simulate_population(self.h, C.byref(args)))

    def population_begin(self, argv, first_haplotype: int = 0, last_haplotype: int = -1):
        args = PopulationArgs.from_argv(argv)
        self._check(self.lib.sp_population_begin(self.h, C.byref(args), first_haplotype, last_haplotype))

    def population_genotypes(self, first_haplotype: int, last_haplotype: int):
        self._check(self.lib.sp_population_genotypes(self.h, first_haplotype, last_haplotype))

    def population_info(self):
        nv, nh, ne, nf = C.c_int32(), C.c_int32(), C.c_int64(), C.c_int32()
        self._check(self.lib.sp_population_info(self.h, C.byref(nv), C.byref(nh), C.byref(ne), C.byref(nf)))
        return dict(n_variants=nv.value, n_haplotypes=nh.value, local_entries=ne.value, n_fixed=nf.value)

    POPULATION_ROWS = ("A1_loadReference", "A2_model", "A3_loadPositions", "S1_sampleVariants", "GCN_buildPopulation",
                       "I_assignInsertionStrings", "R1_annotateVariantsWithRepeats", "W_buildVCF", "Z_serialize")

    def population_timing(self) -> dict:
        """Wall-clock seconds of the last begin/finish per row of SURVEY.md §8a."""
        out = np.zeros(len(self.POPULATION_ROWS), np.float64)
        self._check(self.lib.sp_population_timing(self.h, _ptr(out), out.size, None))
        return {k: float(v) for k, v in zip(self.POPULATION_ROWS, out)}

    def population_local_counts(self) -> np.ndarray:
        info = self.population_info()
        out = np.zeros(max(info["n_variants"], 1), dtype=np.int32)
        self._check(self.lib.sp_population_local_counts(self.h, _ptr(out)))
        return out[: info["n_variants"]]

    def population_counts_device(self) -> int:
        p = C.c_void_p()
        self._check(self.lib.sp_population_counts_device(self.h, C.byref(p)))
        return int(p.value or 0)

    def population_local_lists(self, n_local_haplotypes: int):
        info = self.population_info()
        off = np.zeros(n_local_haplotypes + 1, dtype=np.int64)
        ids = np.zeros(max(info["local_entries"], 1), dtype=np.int32)
        self._check(self.lib.sp_population_local_lists(self.h, _ptr(off), _ptr(ids)))
        return off, ids[: info["local_entries"]]

    def population_pre_collision_lists(self, n_local_haplotypes: int):
        n = C.c_int64()
        self._check(self.lib.sp_population_pre_collision_lists(self.h, None, None, 0, C.byref(n)))
        off = np.zeros(n_local_haplotypes + 1, dtype=np.int64)
        ids = np.zeros(max(n.value, 1), dtype=np.int32)
        self._check(self.lib.sp_population_pre_collision_lists(self.h, _ptr(off), _ptr(ids), ids.size, C.byref(n)))
        return off, ids[: n.value]

    def population_draw_log(self):
        s, c = C.c_uint64(), C.c_int32()
        self._check(self.lib.sp_population_draw_log(self.h, C.byref(s), C.byref(c)))
        return dict(steps_after_sample_variants=s.value, candidates=c.value)

    def population_finish(self, global_counts=None, all_off=None, all_ids=None, write_joint_outputs: bool = True):
        gc = None if global_counts is None else np.ascontiguousarray(global_counts, dtype=np.int32)
        ao = None if all_off is None else np.ascontiguousarray(all_off, dtype=np.int64)
        ai = None if all_ids is None else np.ascontiguousarray(all_ids, dtype=np.int32)
        if ai is not None and ai.size == 0:
            ai = np.zeros(1, dtype=np.int32)
        self._check(self.lib.sp_population_finish(self.h, _ptr(gc), _ptr(ao), _ptr(ai), int(write_joint_outputs)))

    # ---- row T (extension): Mendelian child -----------------------------------------------------------------------
    def trio_add_child(self, father_individual: int, mother_individual: int, seed: int, write_vcf: bool = True):
        h, f, m = C.c_int32(), C.c_int32(), C.c_int32()
        self._check(self.lib.sp_trio_add_child(self.h, father_individual, mother_individual, seed, int(write_vcf),
                                               C.byref(h), C.byref(f), C.byref(m)))
        return dict(child_first_haplotype=h.value, from_father_haplotype=f.value, from_mother_haplotype=m.value)

    # ---- §8(f) rank 1: haplotype2reads.sh tail (cat, shuf, tr) + reads2svs.sh coverage chunks -------------------------
    def reads_mix_begin(self):
        self._check(self.lib.sp_reads_mix_begin(self.h))

    def reads_mix_add(self):
        self._check(self.lib.sp_reads_mix_add(self.h))

    def reads_mix_add_pair(self, h0: int, h1: int, ref_num0: int = 1, ref_num1: int = 1, seed_per_haplotype: bool = False):
        """Both haplotypes of an individual sequenced with both read kernels in flight, then added to the mix (h0 first).
        Returns a list of two result dicts (as reads_simulate_haplotype) and the device span of the two kernels in ms."""
        n = (C.c_int64 * 2)(); fb = (C.c_int64 * 2)(); nd = (C.c_int64 * 2)(); nb = (C.c_int64 * 2)()
        span = C.c_double()
        self._check(self.lib.sp_reads_mix_add_pair(self.h, h0, h1, ref_num0, ref_num1, int(seed_per_haplotype), n, fb, nd, nb, C.byref(span)))
        return [dict(n_reads=n[k], file_bytes=fb[k], n_dropped=nd[k], bases=nb[k]) for k in range(2)], span.value

    def reads_mix_finish(self, shuffle_seed: int):
        n, b = C.c_int64(), C.c_int64()
        self._check(self.lib.sp_reads_mix_finish(self.h, shuffle_seed, C.byref(n), C.byref(b)))
        return n.value, b.value

    def reads_mix_coverage_prefix(self, max_coverage: int, coverage: int):
        a, r, b = C.c_int64(), C.c_int64(), C.c_int64()
        self._check(self.lib.sp_reads_mix_coverage_prefix(self.h, max_coverage, coverage, C.byref(a), C.byref(r), C.byref(b)))
        return dict(rows_1x=a.value, rows=r.value, bytes=b.value)

    def reads_mix_copy(self, offset: int, n_bytes: int) -> np.ndarray:
        out = np.empty(max(n_bytes, 1), dtype=np.uint8)
        self._check(self.lib.sp_reads_mix_copy(self.h, offset, n_bytes, _ptr(out)))
        return out[:n_bytes]

    def reads_mix_offsets(self, n_reads: int) -> np.ndarray:
        out = np.zeros(n_reads + 1, dtype=np.int64)
        self._check(self.lib.sp_reads_mix_offsets(self.h, _ptr(out)))
        return out

    def reads_mix_timing(self):
        a, b = C.c_double(), C.c_double()
        self._check(self.lib.sp_reads_mix_timing(self.h, C.byref(a), C.byref(b)))
        return dict(order_ms=a.value, copy_ms=b.value)

    def haplotype2reads(self, id1: int, shuffle_seed: int, output_file: str | None = None):
        n, b = C.c_int64(), C.c_int64()
        self._check(self.lib.sp_haplotype2reads(self.h, id1, shuffle_seed, output_file.encode() if output_file else None,
                                                C.byref(n), C.byref(b)))
        return n.value, b.value

    # ---- row P: reads (pbsim2 model mode) --------------------------------------------------------------------
    def reads_configure(self, args: "ReadArgs"):
        self._read_args = args  # keeps the byte strings alive
        self._check(self.lib.sp_reads_configure(self.h, C.byref(args)))

    def reads_set_replay(self, draws: np.ndarray, read_offsets: np.ndarray):
        draws = np.ascontiguousarray(draws, dtype=np.int32)
        off = np.ascontiguousarray(read_offsets, dtype=np.int64)
        self._check(self.lib.sp_reads_set_replay(self.h, _ptr(draws), draws.size, _ptr(off), off.size))

    def _reads_result(self, fn, *args):
        n, b, d, e = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        self._check(fn(self.h, *args, C.byref(n), C.byref(b), C.byref(d), C.byref(e)))
        return dict(n_reads=n.value, file_bytes=b.value, n_dropped=d.value, bases=e.value)

    def reads_simulate_haplotype(self, h: int, ref_num: int = 1):
        return self._reads_result(self.lib.sp_reads_simulate_haplotype, h, ref_num)

    def reads_simulate_device(self, device_ptr: int, length: int, ref_num: int = 1):
        return self._reads_result(self.lib.sp_reads_simulate_device, C.c_void_p(device_ptr), length, ref_num)

    def reads_simulate_sequence(self, seq: np.ndarray, ref_num: int = 1):
        seq = np.ascontiguousarray(seq, dtype=np.uint8)
        return self._reads_result(self.lib.sp_reads_simulate_sequence, _ptr(seq), seq.size, ref_num)

    def reads_copy(self, capacity: int, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.empty(max(capacity, 1), dtype=np.uint8)
        w = C.c_int64()
        self._check(self.lib.sp_reads_copy(self.h, _ptr(out), out.size, C.byref(w)))
        return out[:w.value]

    def reads_log(self, n_reads: int):
        ln = np.zeros(max(n_reads, 1), np.int32); dr = np.zeros(max(n_reads, 1), np.uint8); us = np.zeros(max(n_reads, 1), np.uint32)
        self._check(self.lib.sp_reads_log(self.h, _ptr(ln), _ptr(dr), _ptr(us)))
        return ln[:n_reads], dr[:n_reads], us[:n_reads]

    def reads_timing(self):
        a, b = C.c_double(), C.c_double()
        self._check(self.lib.sp_reads_timing(self.h, C.byref(a), C.byref(b)))
        return dict(plan_ms=a.value, kernel_ms=b.value)

    def simulate_reads(self, args: "ReadArgs", fasta_path: str):
        """The whole `pbsim` command (docker/haplotype2reads.sh:53): writes <prefix>_%04d.fastq per sequence."""
        self._read_args = args
        self._check(self.lib.sp_simulate_reads(self.h, C.byref(args), fasta_path.encode()))

    # ---- row D: SampleReadsFromLengthBins -------------------------------------------------------------------------
    def sample_reads_from_length_bins(self, argv, cumulative: np.ndarray):
        """SampleReadsFromLengthBins.main(args) (src/SampleReadsFromLengthBins.java:19-81): same positional argv; the
        cumulative bimodal table (buildCumulativeBimodal :87-102) is computed by the caller. Raises SynthpopError with
        code 3 where the reference calls System.exit(3)."""
        a = [str(x) for x in argv]
        if len(a) < 11:
            raise IndexError("SampleReadsFromLengthBins takes 11 positional arguments")
        cum = np.ascontiguousarray(cumulative, dtype=np.float64)
        args = LengthBinArgs(cum.ctypes.data_as(C.POINTER(C.c_double)), int(a[5]), int(a[6]), a[7].encode(), int(a[8]),
                             float(a[9]), a[10].encode())
        n_bins = (int(a[6]) + int(a[5]) - 1) // int(a[5])
        if cum.size != n_bins:
            raise ValueError(f"cumulative table must have N_BINS={n_bins} entries")
        self._check(self.lib.sp_sample_reads_from_bins(self.h, C.byref(args)))

    def length_bins_load(self, bins_prefix: str, n_bins: int):
        self._check(self.lib.sp_length_bins_load(self.h, bins_prefix.encode(), n_bins))

    def length_bins_set(self, blob: np.ndarray, bin_off: np.ndarray):
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        off = np.ascontiguousarray(bin_off, dtype=np.int64)
        if blob.size == 0:
            blob = np.zeros(1, np.uint8)
        self._check(self.lib.sp_length_bins_set(self.h, off.size - 1, _ptr(blob), _ptr(off)))

    def length_bins_sample(self, cumulative: np.ndarray, target_coverage_bp: int):
        cum = np.ascontiguousarray(cumulative, dtype=np.float64)
        n, b, a, e = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        hist = np.zeros(cum.size, np.int32)
        self._check(self.lib.sp_length_bins_sample(self.h, _ptr(cum), cum.size, int(target_coverage_bp), C.byref(n), C.byref(b),
                                                   C.byref(a), C.byref(e), _ptr(hist)))
        return dict(n_reads=n.value, fastq_bytes=b.value, attempts=a.value, bases=e.value, histogram=hist)

    def length_bins_copy(self, capacity: int, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.empty(max(capacity, 1), dtype=np.uint8)
        w = C.c_int64()
        self._check(self.lib.sp_length_bins_copy(self.h, _ptr(out), out.size, C.byref(w)))
        return out[:w.value]

    # ---- §8(f) rank 2: AnnotateVCF ---------------------------------------------------------------------------------------
    def annotate_vcf(self, argv):
        """AnnotateVCF.main(args) (src/AnnotateVCF.java:96-124): INPUT_FILES_LIST MIN_DISTANCE MAX_DISTANCE REFERENCE_FILE
        REPEAT_INTERVALS_FILE SEGDUPS_FILE; writes <input up to .vcf>_annotated.vcf beside every listed file."""
        a = [str(x) for x in argv]
        if len(a) < 6:
            raise IndexError("AnnotateVCF takes 6 positional arguments")
        args = AnnotateArgs(a[0].encode(), int(a[1]), int(a[2]), a[3].encode(), a[4].encode(), a[5].encode())
        self._check(self.lib.sp_annotate_vcf(self.h, C.byref(args)))

    def annotate_load(self, reference_file: str, repeat_intervals_file: str, segdups_file: str):
        self._check(self.lib.sp_annotate_load(self.h, reference_file.encode(), repeat_intervals_file.encode(), segdups_file.encode()))

    def annotate_file(self, input_vcf: str, output_vcf: str, min_distance: int, max_distance: int) -> int:
        n = C.c_int64()
        self._check(self.lib.sp_annotate_file(self.h, input_vcf.encode(), output_vcf.encode(), min_distance, max_distance, C.byref(n)))
        return n.value

    def annotate_timing(self):
        a, b, n = C.c_double(), C.c_double(), C.c_int64()
        self._check(self.lib.sp_annotate_timing(self.h, C.byref(a), C.byref(b), C.byref(n)))
        return dict(load_s=a.value, kernel_ms=b.value, records=n.value)

    # ---- §8(f) rank 3: BuildReadLengthBins / Fastq2LengthHistogram -----------------------------------------------------------
    def build_read_length_bins(self, argv):
        """BuildReadLengthBins.main(args) (src/BuildReadLengthBins.java:21-105): INPUT_FILES_LIST BIN_LENGTH MAX_READ_LENGTH
        GENOME_LENGTH_HAPLOID OUTPUT_PREFIX."""
        a = [str(x) for x in argv]
        if len(a) < 5:
            raise IndexError("BuildReadLengthBins takes 5 positional arguments")
        args = BuildBinsArgs(a[0].encode(), int(a[1]), int(a[2]), int(a[3]), a[4].encode())
        self._check(self.lib.sp_build_read_length_bins(self.h, C.byref(args)))

    def fastq2length_histogram(self, argv):
        """Fastq2LengthHistogram.main(args): FASTQ_FILE BIN_LENGTH MAX_READ_LENGTH OUTPUT_FILE_HISTOGRAM OUTPUT_FILE_MAX."""
        a = [str(x) for x in argv]
        self._check(self.lib.sp_fastq2length_histogram(self.h, a[0].encode(), int(a[1]), int(a[2]), a[3].encode(), a[4].encode()))

    def build_model(self, argv):
        """BuildModel.main(args) (src/BuildModel.java:21-47): REPEATMASKER_FILE NROWS TRF_FILE NROWS SEGDUPS_FILE NROWS REFERENCE_FAI
        REFERENCE_CYTOBANDS GNOMADSV_FILE ONLY_PASS OUTPUT_DIR VERBOSE."""
        a = [str(x) for x in argv]
        if len(a) < 12:
            raise IndexError("BuildModel takes 12 positional arguments")
        args = BuildModelArgs(a[0].encode(), int(a[1]), a[2].encode(), int(a[3]), a[4].encode(), int(a[5]), a[6].encode(), a[7].encode(),
                              a[8].encode(), 1 if int(a[9]) == 1 else 0, a[10].encode(), 1 if int(a[11]) == 1 else 0)
        self._check(self.lib.sp_build_model(self.h, C.byref(args)))

    def build_model_timing(self):
        a, b, d = C.c_double(), C.c_double(), C.c_double()
        n, m = C.c_int64(), C.c_int64()
        self._check(self.lib.sp_build_model_timing(self.h, C.byref(a), C.byref(b), C.byref(d), C.byref(n), C.byref(m)))
        return dict(parse_s=a.value, device_ms=b.value, finish_s=d.value, records=n.value, allele_frequencies=m.value)

    def build_bins_timing(self):
        a, b, d = C.c_double(), C.c_double(), C.c_double()
        n, nb = C.c_int64(), C.c_int64()
        self._check(self.lib.sp_build_bins_timing(self.h, C.byref(a), C.byref(b), C.byref(d), C.byref(n), C.byref(nb)))
        return dict(index_ms=a.value, shuffle_ms=b.value, copy_ms=d.value, records=n.value, bytes=nb.value)

    def length_bins_timing(self):
        a, b = C.c_double(), C.c_double()
        self._check(self.lib.sp_length_bins_timing(self.h, C.byref(a), C.byref(b)))
        return dict(sample_ms=a.value, copy_ms=b.value)
