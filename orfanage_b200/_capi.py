"""ctypes bindings for include/orfx.h (device library) and include/orfx_pack.h (host batcher)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

MODES = {"START_MATCH": 0, "ALL": 1, "LONGEST": 2, "LONGEST_MATCH": 3, "BEST": 4, "FIRST": 5}
STATUS_NOTES = ["-", "no-overlap", "segment_len<3", "cds_len<3", "not_rescued", "missing_stop",
                "missing_start", "minlen", "dup"]

# aa_table of the reference (common.hpp:268-277): order ARNDCQEGHILKMFPSTWYVBZX*, unknown -> 23
_AA_ORDER = "ARNDCQEGHILKMFPSTWYVBZX*"


def encode_protein(s):
    """ASCII protein -> 0..23 codes through the reference's aa_table."""
    lut = np.full(256, 23, dtype=np.uint8)
    for i, ch in enumerate(_AA_ORDER):
        lut[ord(ch)] = i
        lut[ord(ch.lower())] = i
    return lut[np.frombuffer(s.encode("ascii") if isinstance(s, str) else bytes(s), dtype=np.uint8)]


class OrfxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("orfx error %d: %s" % (code, msg))
        self.code = code


def device_library_path():
    # ORFX_LIB: a differently-tuned build of the same library (kernel experiments); never a CPU stand-in
    return os.environ.get("ORFX_LIB") or os.path.join(_HERE, "liborfx.so")


def host_library_path():
    return os.path.join(_HERE, "liborfx_host.so")


def cli_path():
    return os.path.join(_HERE, "bin", "orfanage")


class Params(C.Structure):
    """orfx_params (include/orfx.h); replaces GlobalParameters, orfanage.cpp:99-142."""
    _fields_ = [("n_modes", C.c_int32), ("mode_array", C.c_int32 * 8), ("lpi", C.c_int32), ("ilpi", C.c_int32),
                ("mlpi", C.c_int32), ("cds_minlen", C.c_int32), ("non_aug", C.c_int32), ("overhang", C.c_int32),
                ("spliced_overhang", C.c_int32), ("has_reference", C.c_int32), ("want_rows", C.c_int32),
                ("align_enabled", C.c_int32), ("align_gapo", C.c_int32), ("align_gape", C.c_int32), ("host_results_pageable", C.c_int32)]

    @classmethod
    def make(cls, mode="LONGEST_MATCH,BEST,START_MATCH,LONGEST,FIRST,ALL", lpi=-1, ilpi=-1, mlpi=-1, minlen=-1,
             non_aug=False, has_reference=True, want_rows=True, overhang=0, spliced_overhang=False, align=False, gapo=0, gape=0):
        p = cls()
        names = [m.strip().upper() for m in mode.split(",")] if isinstance(mode, str) else list(mode)
        p.n_modes = len(names)
        for i, m in enumerate(names):
            p.mode_array[i] = MODES[m] if isinstance(m, str) else int(m)
        p.lpi, p.ilpi, p.mlpi, p.cds_minlen = lpi, ilpi, mlpi, minlen
        p.non_aug = int(non_aug)
        p.overhang = overhang
        p.spliced_overhang = int(spliced_overhang)
        p.has_reference = int(has_reference)
        p.want_rows = int(want_rows)
        p.align_enabled, p.align_gapo, p.align_gape = int(align), int(gapo), int(gape)
        return p


class _Seg(C.Structure):
    _fields_ = [("start", C.c_int32), ("end", C.c_int32)]


class _Batch(C.Structure):
    _fields_ = [("n_bundles", C.c_int32), ("n_tx", C.c_int32), ("n_segs", C.c_int64), ("seq_bytes", C.c_int64),
                ("seq", C.c_void_p), ("seq_off", C.c_void_p), ("bundle_start", C.c_void_p), ("bundle_tx_off", C.c_void_p),
                ("tx_kind", C.c_void_p), ("tx_strand", C.c_void_p), ("tx_seg_off", C.c_void_p), ("tx_cds_start", C.c_void_p),
                ("tx_cds_end", C.c_void_p), ("tx_tid_class", C.c_void_p), ("segs", C.c_void_p), ("seg_phase", C.c_void_p),
                ("seg_seq_off", C.c_void_p), ("seq_format", C.c_int32), ("reserved0", C.c_int32),
                ("tx_ext_lo", C.c_void_p), ("tx_ext_hi", C.c_void_p), ("tx_ext_flags", C.c_void_p),
                ("seq_exc_pos", C.c_void_p), ("seq_exc_code", C.c_void_p), ("n_seq_exc", C.c_int64),
                ("seg_ref", C.c_void_p), ("bundle_dict_off", C.c_void_p), ("dict_segs", C.c_void_p), ("dict_phase", C.c_void_p),
                ("dict_seq_off", C.c_void_p), ("n_dict", C.c_int64)]


class _Best(C.Structure):
    _fields_ = [("q_tx", C.c_int32), ("t_tx", C.c_int32), ("cds_start", C.c_int32), ("cds_end", C.c_int32),
                ("duplicity", C.c_int32), ("n_rows", C.c_int32), ("row_off", C.c_int64), ("best_row", C.c_int64),
                ("lpi", C.c_float), ("ilpi", C.c_float), ("mlpi", C.c_float), ("start_match", C.c_int32),
                ("ext_lo", C.c_int32), ("ext_hi", C.c_int32)]


BEST_DTYPE = np.dtype([("q_tx", "<i4"), ("t_tx", "<i4"), ("cds_start", "<i4"), ("cds_end", "<i4"), ("duplicity", "<i4"),
                       ("n_rows", "<i4"), ("row_off", "<i8"), ("best_row", "<i8"), ("lpi", "<f4"), ("ilpi", "<f4"),
                       ("mlpi", "<f4"), ("start_match", "<i4"), ("ext_lo", "<i4"), ("ext_hi", "<i4")])
assert BEST_DTYPE.itemsize == C.sizeof(_Best)

_ROW_I32 = ["row_q_tx", "row_t_tx", "row_seg_start", "row_seg_end", "row_cds_start", "row_cds_end"]
_ROW_U8 = ["row_status", "row_pass", "row_start_match"]
_ROW_I32B = ["row_n_dup", "row_qlen", "row_tlen", "row_ulen", "row_query_start", "row_extra", "row_missing", "row_match",
             "row_outframe", "row_inframe"]
_ROW_F32 = ["row_lpi", "row_ilpi", "row_mlpi"]
_ROW_ALN = ["row_aln_max", "row_aln_score", "row_aln_mqe", "row_aln_mte"]


class _Results(C.Structure):
    _fields_ = ([("n_queries", C.c_int32), ("n_rows", C.c_int64), ("best", C.c_void_p)]
                + [(n, C.c_void_p) for n in _ROW_I32] + [(n, C.c_void_p) for n in _ROW_U8]
                + [(n, C.c_void_p) for n in _ROW_I32B] + [(n, C.c_void_p) for n in _ROW_F32]
                + [("row_ext_lo", C.c_void_p), ("row_ext_hi", C.c_void_p)]
                + [("print_order", C.c_void_p), ("ms_h2d", C.c_float), ("ms_kernels", C.c_float), ("ms_d2h", C.c_float)]
                + [(n, C.c_void_p) for n in _ROW_ALN] + [("ms_align", C.c_float), ("reserved1", C.c_float)])


class _AlnOut(C.Structure):
    _fields_ = [("max", C.c_void_p), ("score", C.c_void_p), ("mqe", C.c_void_p), ("mte", C.c_void_p)]


_BATCH_ARRAYS = [("seq", np.uint8), ("seq_off", np.int64), ("bundle_start", np.int32), ("bundle_tx_off", np.int32),
                 ("tx_kind", np.uint8), ("tx_strand", np.uint8), ("tx_seg_off", np.int32), ("tx_cds_start", np.int32),
                 ("tx_cds_end", np.int32), ("tx_tid_class", np.int32), ("segs", np.int32), ("seg_phase", np.int8),
                 ("seg_seq_off", np.int32), ("tx_ext_lo", np.int32), ("tx_ext_hi", np.int32), ("tx_ext_flags", np.uint8),
                 ("seq_exc_pos", np.int64), ("seq_exc_code", np.uint8),
                 ("seg_ref", np.uint16), ("bundle_dict_off", np.int32), ("dict_segs", np.int32), ("dict_phase", np.int8),
                 ("dict_seq_off", np.int32)]
# with a segment dictionary (seg_ref) the device library takes the chains from it and never reads these
_DICT_REPLACES = ("segs", "seg_phase", "seg_seq_off")


def _view(ptr, n, dtype):
    """numpy view (no copy) of n elements at a raw address."""
    if not ptr or n <= 0:
        return np.zeros(0, dtype=dtype)
    nbytes = int(n) * np.dtype(dtype).itemsize
    buf = (C.c_char * nbytes).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=int(n))


class _CleanOut(C.Structure):
    _fields_ = [("keep", C.c_void_p), ("cds_start", C.c_void_p), ("cds_end", C.c_void_p)]


class Batch:
    """A packed set of locus bundles (orfx_batch).  Holds numpy arrays (or views of library-owned
    memory) and the ctypes struct that points at them."""

    def __init__(self, **arrays):
        self.arrays = {}
        for name, dt in _BATCH_ARRAYS:
            a = arrays.get(name)
            if a is None:
                a = np.zeros(0, dtype=dt)
            a = np.ascontiguousarray(a, dtype=dt)
            self.arrays[name] = a
        self.owner = arrays.get("owner")
        # ORFX_SEQ_NT4 / ORFX_SEQ_NT2 batches: `seq` holds two / four bases per byte and `seq_bases` counts the bases
        self.seq_format = int(np.asarray(arrays.get("seq_format", 0)).reshape(-1)[0]) if arrays.get("seq_format") is not None else 0
        a = self.arrays
        if a["segs"].ndim == 1:
            a["segs"] = a["segs"].reshape(-1, 2)
        s = _Batch()
        s.n_bundles = max(len(a["bundle_tx_off"]) - 1, 0)
        s.n_tx = len(a["tx_kind"])
        s.n_segs = a["segs"].shape[0] or a["seg_ref"].size
        s.seq_bytes = int(np.asarray(arrays["seq_bases"]).reshape(-1)[0]) if self.seq_format in (1, 2) else a["seq"].size
        s.seq_format = self.seq_format
        s.n_seq_exc = a["seq_exc_pos"].size
        if a["dict_segs"].ndim == 1:
            a["dict_segs"] = a["dict_segs"].reshape(-1, 2)
        s.n_dict = a["dict_segs"].shape[0]
        for name, _ in _BATCH_ARRAYS:
            arr = a[name]
            setattr(s, name, arr.ctypes.data if arr.size else None)
        # offset arrays must exist even for an empty batch
        if a["seq_off"].size == 0:
            self.arrays["seq_off"] = np.zeros(1, dtype=np.int64)
            s.seq_off = self.arrays["seq_off"].ctypes.data
        if a["bundle_tx_off"].size == 0:
            self.arrays["bundle_tx_off"] = np.zeros(1, dtype=np.int32)
            s.bundle_tx_off = self.arrays["bundle_tx_off"].ctypes.data
        if a["tx_seg_off"].size == 0:
            self.arrays["tx_seg_off"] = np.zeros(1, dtype=np.int32)
            s.tx_seg_off = self.arrays["tx_seg_off"].ctypes.data
        self.struct = s

    @classmethod
    def from_struct(cls, s, owner=None):
        nb, nt, ns = s.n_bundles, s.n_tx, s.n_segs
        arrays = dict(
            seq=_view(s.seq, {1: (s.seq_bytes + 1) // 2, 2: (s.seq_bytes + 3) // 4}.get(s.seq_format, s.seq_bytes), np.uint8), seq_off=_view(s.seq_off, nb + 1, np.int64),
            bundle_start=_view(s.bundle_start, nb, np.int32), bundle_tx_off=_view(s.bundle_tx_off, nb + 1, np.int32),
            tx_kind=_view(s.tx_kind, nt, np.uint8), tx_strand=_view(s.tx_strand, nt, np.uint8),
            tx_seg_off=_view(s.tx_seg_off, nt + 1, np.int32), tx_cds_start=_view(s.tx_cds_start, nt, np.int32),
            tx_cds_end=_view(s.tx_cds_end, nt, np.int32), tx_tid_class=_view(s.tx_tid_class, nt, np.int32),
            segs=_view(s.segs, ns * 2, np.int32), seg_phase=_view(s.seg_phase, ns, np.int8),
            seg_seq_off=_view(s.seg_seq_off, ns if s.seg_seq_off else 0, np.int32),
            tx_ext_lo=_view(s.tx_ext_lo, nt if s.tx_ext_lo else 0, np.int32), tx_ext_hi=_view(s.tx_ext_hi, nt if s.tx_ext_hi else 0, np.int32),
            tx_ext_flags=_view(s.tx_ext_flags, nt if s.tx_ext_flags else 0, np.uint8),
            seq_exc_pos=_view(s.seq_exc_pos, s.n_seq_exc if s.seq_exc_pos else 0, np.int64),
            seq_exc_code=_view(s.seq_exc_code, s.n_seq_exc if s.seq_exc_code else 0, np.uint8),
            seg_ref=_view(s.seg_ref, ns if s.seg_ref else 0, np.uint16), bundle_dict_off=_view(s.bundle_dict_off, nb + 1 if s.seg_ref else 0, np.int32),
            dict_segs=_view(s.dict_segs, s.n_dict * 2 if s.seg_ref else 0, np.int32), dict_phase=_view(s.dict_phase, s.n_dict if s.seg_ref else 0, np.int8),
            dict_seq_off=_view(s.dict_seq_off, s.n_dict if (s.seg_ref and s.dict_seq_off) else 0, np.int32))
        b = cls.__new__(cls)
        b.arrays = arrays
        b.arrays["segs"] = b.arrays["segs"].reshape(-1, 2)
        b.arrays["dict_segs"] = b.arrays["dict_segs"].reshape(-1, 2)
        b.owner = owner
        b.seq_format = int(s.seq_format)
        cp = _Batch()
        C.memmove(C.byref(cp), C.byref(s), C.sizeof(_Batch))
        b.struct = cp
        return b

    @property
    def n_queries(self):
        return int((self.arrays["tx_kind"] == 0).sum())

    @property
    def h2d_bytes(self):
        """Bytes orfx_submit copies to the device (Context.last_upload_bytes() is the library's own count)."""
        skip = _DICT_REPLACES if self.arrays["seg_ref"].size else ()
        return int(sum(a.nbytes for k, a in self.arrays.items() if k not in skip))

    def without_segment_arrays(self):
        """The same batch with segs / seg_phase / seg_seq_off dropped: what a caller that only has the dictionary passes."""
        assert self.arrays["seg_ref"].size, "no segment dictionary in this batch"
        a = {k: v for k, v in self.arrays.items() if k not in _DICT_REPLACES}
        extra = dict(seq_format=np.array([self.seq_format], np.int32), seq_bases=np.array([self.struct.seq_bytes], np.int64)) if self.seq_format else {}
        return Batch(**a, **extra)

    def query_nt(self):
        a = self.arrays
        seg_len = (a["segs"][:, 1] - a["segs"][:, 0] + 1).astype(np.int64)
        cum = np.concatenate([[0], np.cumsum(seg_len)])
        tx_len = cum[a["tx_seg_off"][1:]] - cum[a["tx_seg_off"][:-1]]
        return int(tx_len[a["tx_kind"] == 0].sum())

    def save(self, path):
        extra = dict(seq_format=np.array([self.seq_format], np.int32), seq_bases=np.array([self.struct.seq_bytes], np.int64)) if self.seq_format else {}
        np.savez_compressed(path, **self.arrays, **extra)

    @classmethod
    def load(cls, path):
        with np.load(path) as z:
            return cls(**{k: z[k] for k in z.files})


class Results:
    """Copies of one orfx_results (the library's buffers are only valid until the next submit)."""

    def __init__(self, r, want_rows=True):
        self.n_queries = int(r.n_queries)
        self.n_rows = int(r.n_rows)
        self.best = _view(r.best, self.n_queries, BEST_DTYPE).copy()
        self.rows = {}
        if want_rows and r.row_q_tx:
            n = self.n_rows
            for name in _ROW_I32 + _ROW_I32B:
                self.rows[name[4:]] = _view(getattr(r, name), n, np.int32).copy()
            for name in _ROW_U8:
                self.rows[name[4:]] = _view(getattr(r, name), n, np.uint8).copy()
            for name in _ROW_F32:
                self.rows[name[4:]] = _view(getattr(r, name), n, np.float32).copy()
            self.rows["print_order"] = _view(r.print_order, n, np.int64).copy()
            if r.row_ext_lo:   # --overhang only
                self.rows["ext_lo"] = _view(r.row_ext_lo, n, np.int32).copy()
                self.rows["ext_hi"] = _view(r.row_ext_hi, n, np.int32).copy()
            if r.row_aln_max:   # params.align_enabled only
                for name in _ROW_ALN:
                    self.rows[name[4:]] = _view(getattr(r, name), n, np.int32).copy()
        self.ms_h2d, self.ms_kernels, self.ms_d2h, self.ms_align = float(r.ms_h2d), float(r.ms_kernels), float(r.ms_d2h), float(r.ms_align)

    def d2h_bytes(self):
        return int(self.best.nbytes + sum(a.nbytes for a in self.rows.values()))


# ---- library loading ---------------------------------------------------------------------------
_DEVICE_LIB = None
_HOST_LIB = None

_SUFFIXES = ["last_error", "create", "submit", "wait", "release", "destroy"]


def _bind_backend(lib, prefix):
    """argtypes/restypes of the five backend entry points + last_error under a symbol prefix."""
    f = lambda n: getattr(lib, prefix + n)  # noqa: E731
    f("last_error").restype = C.c_char_p
    f("last_error").argtypes = []
    f("create").restype = C.c_int
    f("create").argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(Params)]
    f("submit").restype = C.c_int
    f("submit").argtypes = [C.c_void_p, C.POINTER(_Batch)]
    f("wait").restype = C.c_int
    f("wait").argtypes = [C.c_void_p, C.POINTER(_Results)]
    f("release").restype = C.c_int
    f("release").argtypes = [C.c_void_p]
    f("destroy").restype = None
    f("destroy").argtypes = [C.c_void_p]


def load_device_library(path=None):
    """Loads liborfx.so.  Raises when it is missing: there is no CPU path to fall back to."""
    global _DEVICE_LIB
    if _DEVICE_LIB is not None and path is None:
        return _DEVICE_LIB
    p = path or device_library_path()
    if not os.path.exists(p):
        raise OrfxError(-3, "CUDA library %s is not built (run `make cuda` / __graft_entry__.build()); "
                            "this package has no CPU fallback" % p)
    lib = C.CDLL(p)
    _bind_backend(lib, "orfx_")
    lib.orfx_abi_version.restype = C.c_int
    lib.orfx_set_params.restype = C.c_int
    lib.orfx_set_params.argtypes = [C.c_void_p, C.POINTER(Params)]
    lib.orfx_host_alloc.restype = C.c_void_p
    lib.orfx_host_alloc.argtypes = [C.c_size_t]
    lib.orfx_host_free.restype = None
    lib.orfx_host_free.argtypes = [C.c_void_p]
    lib.orfx_upload.restype = C.c_int
    lib.orfx_upload.argtypes = [C.c_void_p, C.POINTER(_Batch)]
    lib.orfx_run_resident.restype = C.c_int
    lib.orfx_run_resident.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
    lib.orfx_fetch.restype = C.c_int
    lib.orfx_fetch.argtypes = [C.c_void_p, C.POINTER(_Results)]
    lib.orfx_kernel_times.restype = C.c_int
    lib.orfx_kernel_times.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_float), C.POINTER(C.c_int)]
    lib.orfx_last_run_ms.restype = C.c_float
    lib.orfx_last_run_ms.argtypes = [C.c_void_p]
    if hasattr(lib, "orfx_last_upload_bytes"):
        lib.orfx_last_upload_bytes.restype = C.c_int64
        lib.orfx_last_upload_bytes.argtypes = [C.c_void_p]
    lib.orfx_last_run_launches.restype = C.c_int
    lib.orfx_last_run_launches.argtypes = [C.c_void_p]
    lib.orfx_resident_query_nt.restype = C.c_int64
    lib.orfx_resident_query_nt.argtypes = [C.c_void_p]
    lib.orfx_get_masks.restype = C.c_int64
    lib.orfx_get_masks.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int64]
    lib.orfx_align_batch.restype = C.c_int
    lib.orfx_align_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64,
                                     C.c_int, C.c_int, C.POINTER(_AlnOut), C.POINTER(C.c_double), C.POINTER(C.c_float)]
    if hasattr(lib, "orfx_align_pairs_static"):   # the CUDA library only
        lib.orfx_align_pairs_static.restype = C.c_int
        lib.orfx_align_pairs_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64,
                                                C.c_int, C.c_int, C.POINTER(_AlnOut)]
    lib.orfx_debug_set.restype = C.c_int
    lib.orfx_debug_set.argtypes = [C.c_void_p, C.c_char_p, C.c_int64]
    if hasattr(lib, "orfx_int_alu_peak"):
        lib.orfx_int_alu_peak.restype = C.c_int
        lib.orfx_int_alu_peak.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
    if path is None:
        _DEVICE_LIB = lib
    return lib


def load_host_library(path=None):
    global _HOST_LIB
    if _HOST_LIB is not None and path is None:
        return _HOST_LIB
    p = path or host_library_path()
    if not os.path.exists(p):
        raise OrfxError(-1, "host library %s is not built (run `make host`)" % p)
    lib = C.CDLL(p)
    lib.orfxh_last_error.restype = C.c_char_p
    lib.orfxh_load.restype = C.c_int
    lib.orfxh_load.argtypes = [C.POINTER(C.c_void_p), C.c_char_p, C.POINTER(C.c_char_p), C.c_int, C.c_char_p, C.c_int, C.c_int]
    lib.orfxh_load_overhang.restype = C.c_int
    lib.orfxh_load_overhang.argtypes = lib.orfxh_load.argtypes + [C.c_int, C.c_int]
    lib.orfxh_count.restype = C.c_int64
    lib.orfxh_count.argtypes = [C.c_void_p, C.c_int]
    lib.orfxh_pack.restype = C.c_int
    lib.orfxh_pack.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(_Batch)]
    lib.orfxh_truncate.restype = C.c_int
    lib.orfxh_truncate.argtypes = [C.c_void_p, C.c_int64]
    lib.orfxh_tx_id.restype = C.c_char_p
    lib.orfxh_tx_id.argtypes = [C.c_void_p, C.c_int, C.c_int32]
    lib.orfxh_tx_global.restype = C.c_int64
    lib.orfxh_tx_global.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64]
    lib.orfxh_free.restype = None
    lib.orfxh_free.argtypes = [C.c_void_p]
    if path is None:
        _HOST_LIB = lib
    return lib


class Context:
    """One GPU, one stream (orfx_ctx).  `lib`/`prefix` exist so that the test-suite can drive a
    second implementation of the same five entry points (the CPU oracle) through this wrapper."""

    def __init__(self, params=None, device=0, lib=None, prefix="orfx_"):
        self.lib = lib if lib is not None else load_device_library()
        self.prefix = prefix
        if lib is not None:
            _bind_backend(self.lib, prefix)
        self.params = params if params is not None else Params.make()
        self.handle = C.c_void_p()
        self._keep = None
        self._check(self._f("create")(C.byref(self.handle), device, C.byref(self.params)))

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    def _check(self, rc):
        if rc != 0:
            msg = self._f("last_error")()
            raise OrfxError(rc, msg.decode() if msg else "")
        return rc

    def set_params(self, params):
        self.params = params
        self._check(self.lib.orfx_set_params(self.handle, C.byref(params)))

    def debug_set(self, key, value):
        """orfx_debug_set: test hook that lowers an internal capacity (see include/orfx.h)."""
        self._check(self.lib.orfx_debug_set(self.handle, key.encode(), int(value)))

    def last_upload_bytes(self):
        return int(self.lib.orfx_last_upload_bytes(self.handle))

    def submit(self, batch):
        self._keep = batch
        self._check(self._f("submit")(self.handle, C.byref(batch.struct)))

    def wait(self, copy=True):
        r = _Results()
        self._check(self._f("wait")(self.handle, C.byref(r)))
        return Results(r, bool(self.params.want_rows)) if copy else r

    def run(self, batch):
        """orfx_submit + orfx_wait: the whole path on host buffers."""
        self.submit(batch)
        return self.wait()

    # device-resident variant
    def upload(self, batch):
        self._keep = batch
        self._check(self.lib.orfx_upload(self.handle, C.byref(batch.struct)))

    def run_resident(self):
        n = C.c_int(0)
        self._check(self.lib.orfx_run_resident(self.handle, C.byref(n)))
        return n.value

    def fetch(self):
        r = _Results()
        self._check(self.lib.orfx_fetch(self.handle, C.byref(r)))
        return Results(r, bool(self.params.want_rows))

    def kernel_times(self):
        names = (C.c_char_p * 32)()
        ms = (C.c_float * 32)()
        n = C.c_int(0)
        self._check(self.lib.orfx_kernel_times(self.handle, 32, names, ms, C.byref(n)))
        return {names[i].decode(): float(ms[i]) for i in range(n.value)}

    def last_run_ms(self):
        return float(self.lib.orfx_last_run_ms(self.handle))

    def last_run_launches(self):
        return int(self.lib.orfx_last_run_launches(self.handle))

    def resident_query_nt(self):
        v = self.lib.orfx_resident_query_nt(self.handle)
        if v < 0:
            self._check(int(v))
        return int(v)

    def masks(self, tx, max_len):
        words = (max_len + 31) // 32 + 1
        stop = np.zeros(words, dtype=np.uint32)
        start = np.zeros(words, dtype=np.uint32)
        L = self.lib.orfx_get_masks(self.handle, tx, stop.ctypes.data, start.ctypes.data, words)
        if L < 0:
            self._check(int(L))
        nw = (L + 31) // 32
        return int(L), stop[:nw], start[:nw]

    def align(self, aa, off, pair_q, pair_t, gapo=12, gape=2):
        """orfx_align_batch: returns dict(max, score, mqe, mte), cells, device ms."""
        aa = np.ascontiguousarray(aa, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.int64)
        pq = np.ascontiguousarray(pair_q, dtype=np.int32)
        pt = np.ascontiguousarray(pair_t, dtype=np.int32)
        n = len(pq)
        out = {k: np.zeros(max(n, 1), dtype=np.int32) for k in ("max", "score", "mqe", "mte")}
        o = _AlnOut(out["max"].ctypes.data, out["score"].ctypes.data, out["mqe"].ctypes.data, out["mte"].ctypes.data)
        cells = C.c_double(0)
        ms = C.c_float(0)
        self._check(self.lib.orfx_align_batch(self.handle, aa.ctypes.data, off.ctypes.data, len(off) - 1, pq.ctypes.data,
                                              pt.ctypes.data, n, gapo, gape, C.byref(o), C.byref(cells), C.byref(ms)))
        return {k: v[:n] for k, v in out.items()}, cells.value, ms.value

    def align_static(self, d_aa, d_off, n_seq, d_pair_q, d_pair_t, n_pairs, d_out, gapo=12, gape=2):
        """orfx_align_pairs_static: every pointer is a DEVICE address (int) of the context's device; d_out = four addresses
        (max, score, mqe, mte).  Nothing is allocated inside the call (scratch = a static __device__ array of the library);
        limits ORFX_STATIC_MAX_PAIRS / ORFX_STATIC_MAX_LEN of include/orfx.h."""
        o = _AlnOut(*[int(x) for x in d_out])
        self._check(self.lib.orfx_align_pairs_static(self.handle, int(d_aa), int(d_off), int(n_seq), int(d_pair_q), int(d_pair_t), int(n_pairs),
                                                     gapo, gape, C.byref(o)))

    def clean(self, batch, rescue=False):
        """orfx_clean_templates on a cleaning batch: returns (keep u8, cds_start i32, cds_end i32), one entry per
        ORFX_TX_QUERY transcript in batch order."""
        n = batch.n_queries
        keep = np.zeros(max(n, 1), dtype=np.uint8)
        cs = np.zeros(max(n, 1), dtype=np.int32)
        ce = np.zeros(max(n, 1), dtype=np.int32)
        out = _CleanOut(keep.ctypes.data, cs.ctypes.data, ce.ctypes.data)
        n_out = C.c_int64(0)
        f = self._f("clean_templates")
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.POINTER(_Batch), C.c_int, C.POINTER(_CleanOut), C.POINTER(C.c_int64)]
        self._keep = batch
        self._check(f(self.handle, C.byref(batch.struct), 1 if rescue else 0, C.byref(out), C.byref(n_out)))
        assert n_out.value == n
        return keep[:n], cs[:n], ce[:n]

    def score_pairs(self, batch):
        """orfx_score_pairs on a compare batch: dict of per-pair arrays (q_tx, t_tx, qlen, tlen, ulen, match, inframe,
        outframe, extra, missing as int32; lpi, ilpi, mlpi as float32), query-major."""
        r = _Results()
        f = self._f("score_pairs")
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.POINTER(_Batch), C.POINTER(_Results)]
        self._keep = batch
        self._check(f(self.handle, C.byref(batch.struct), C.byref(r)))
        n = int(r.n_rows)
        out = {}
        for name in ("q_tx", "t_tx", "qlen", "tlen", "ulen", "match", "inframe", "outframe", "extra", "missing"):
            out[name] = _view(getattr(r, "row_" + name), n, np.int32).copy()
        for name in ("lpi", "ilpi", "mlpi"):
            out[name] = _view(getattr(r, "row_" + name), n, np.float32).copy()
        return out

    def int_alu_peak(self):
        """orfx_int_alu_peak: measured integer issue rates of this GPU, operations/s."""
        ops = (C.c_double * 4)()
        self._check(self.lib.orfx_int_alu_peak(self.handle, ops))
        return {"iadd3": ops[0], "iadd_vimnmx3": ops[1], "viaddmnmx": ops[2], "idp4a": ops[3]}

    def close(self):
        if self.handle:
            self._f("release")(self.handle)
            self._f("destroy")(self.handle)
            self.handle = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def shard_pairs(off, pair_q, pair_t, n_shards):
    """Deals protein pairs to n_shards GPUs ("length-bucketed pairs dealt round-robin by cells"): pairs in
    descending order of DP cells (qlen x tlen), dealt in snake order so that every shard gets the same mix of lengths.
    Pairs are independent, so there is no exchange step.  Returns one index array (into pair_q / pair_t) per shard."""
    off = np.asarray(off, dtype=np.int64)
    pq = np.asarray(pair_q, dtype=np.int64)
    pt = np.asarray(pair_t, dtype=np.int64)
    cells = (off[pq + 1] - off[pq]) * (off[pt + 1] - off[pt])
    order = np.argsort(-cells, kind="stable")
    pos = np.arange(len(order))
    rnd, col = pos // n_shards, pos % n_shards
    owner = np.where(rnd % 2 == 0, col, n_shards - 1 - col)
    return [order[owner == k] for k in range(n_shards)]


class Workload:
    """Loaded + prepared + bundled transcriptome (orfxh_workload): orfanage.cpp:313-367."""

    def __init__(self, query, templates, reference=None, use_id=False, keep_all_cds=False, overhang=0, spliced_overhang=False):
        self.lib = load_host_library()
        self.handle = C.c_void_p()
        arr = (C.c_char_p * len(templates))(*[t.encode() for t in templates])
        rc = self.lib.orfxh_load_overhang(C.byref(self.handle), query.encode(), arr, len(templates),
                                          reference.encode() if reference else None, int(use_id), int(keep_all_cds),
                                          int(overhang), int(spliced_overhang))
        if rc != 0:
            raise OrfxError(rc, self.lib.orfxh_last_error().decode())

    def count(self, what):
        return int(self.lib.orfxh_count(self.handle, {"bundles": 0, "active_bundles": 1, "queries": 2, "templates": 3,
                                                      "all_queries": 4, "query_nt": 5}[what]))

    def truncate(self, n_active):
        self.lib.orfxh_truncate(self.handle, int(n_active))

    def pack(self, shard=0, n_shards=1, pinned=False, layout="islands"):
        """Returns a Batch viewing arrays owned by this workload (valid until the next pack of the
        same shard or close())."""
        s = _Batch()
        alloc = dealloc = None
        if pinned:
            dev = load_device_library()
            alloc = C.cast(dev.orfx_host_alloc, C.c_void_p)
            dealloc = C.cast(dev.orfx_host_free, C.c_void_p)
        rc = self.lib.orfxh_pack(self.handle, shard, n_shards, {"windows": 0, "islands": 1, "islands4": 2, "islands_t": 3, "islands2": 4, "islands2d": 5}[layout], alloc, dealloc, C.byref(s))
        if rc != 0:
            raise OrfxError(rc, self.lib.orfxh_last_error().decode())
        return Batch.from_struct(s, owner=self)

    def tx_global(self, n_tx, shard=0):
        """Workload-wide transcript index of every transcript of the batch last packed for `shard`."""
        out = np.zeros(max(int(n_tx), 1), dtype=np.int32)
        n = self.lib.orfxh_tx_global(self.handle, shard, out.ctypes.data, len(out))
        if n < 0:
            raise OrfxError(int(n), self.lib.orfxh_last_error().decode())
        return out[:n]

    def tx_id(self, tx, shard=0):
        v = self.lib.orfxh_tx_id(self.handle, shard, tx)
        return v.decode() if v is not None else None

    def close(self):
        if self.handle:
            self.lib.orfxh_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
