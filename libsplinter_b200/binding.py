"""ctypes binding of libsplinter_gpu.so (C ABI: include/splinter_gpu.h).

This is the harness-side view used by tests/, bench.py and __graft_entry__.py; the product is
the shared library itself.  There is no CPU fallback: if the library is missing the import of
this module raises, and every entry point fails loudly without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "libsplinter_gpu.so"

# symbols include/splinter_gpu.h declares; tests check every one is exported
ABI_SYMBOLS = [
    "spl_gpu_attach", "spl_gpu_attach_image", "spl_gpu_open_dense", "spl_gpu_detach", "spl_gpu_get_info",
    "spl_gpu_last_error", "spl_gpu_sync", "spl_gpu_wait", "spl_gpu_dense_upload", "spl_gpu_dense_fill_synthetic",
    "spl_gpu_read_mirror", "spl_gpu_search", "spl_gpu_search_device", "spl_gpu_merge_topk_device", "spl_gpu_merge_topk_batch_device",
    "spl_gpu_launch_count", "spl_gpu_set_timing", "spl_gpu_last_scan_ms", "spl_gpu_scan_noalloc", "spl_gpu_slot_meta", "spl_gpu_last_totals",
    "spl_gpu_get_counters", "spl_gpu_find_key", "spl_gpu_merge_topk_packed_device",
    "spl_gpu_group_attach", "spl_gpu_group_open_dense", "spl_gpu_group_detach", "spl_gpu_group_size", "spl_gpu_group_member",
    "spl_gpu_group_sync", "spl_gpu_group_fill_synthetic", "spl_gpu_group_search", "spl_gpu_group_reissued",
    "spl_gpu_ingest_embeddings",
]

ORDER_COSINE, ORDER_L2 = 0, 1
MODE_AUTO, MODE_FAST, MODE_EXACT, MODE_BATCH_BF16 = 0, 1, 2, 3
OPT_PERSISTENT, OPT_BF16_MIRROR = 0x1, 0x2
ROW_LIST, ROW_ENUM, ROW_EMB = 0x1, 0x2, 0x4
NO_SLOT = 0xFFFFFFFF
KMAX_FUSED = 256
COUNT_UNANSWERED = 0xFFFFFFFF
SEARCH_PIPELINED = 0x1


class Opts(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("flags", C.c_uint32),
                ("slot_begin", C.c_uint32), ("slot_count", C.c_uint32), ("sync_threads", C.c_uint32)]


class SearchParams(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("nq", C.c_uint32), ("dim", C.c_uint32), ("k", C.c_uint32),
                ("order", C.c_uint32), ("mode", C.c_uint32), ("bloom_mask", C.c_uint64),
                ("min_similarity", C.c_float), ("max_distance", C.c_float), ("exclude_slot", C.c_uint32),
                ("flags", C.c_uint32)]


class SyncStats(C.Structure):
    _fields_ = [("generation", C.c_uint64), ("slots_swept", C.c_uint64), ("rows_uploaded", C.c_uint64),
                ("meta_updated", C.c_uint64), ("skipped_busy", C.c_uint64), ("aba_rechecked", C.c_uint64),
                ("bytes_h2d", C.c_uint64), ("sweep_ms", C.c_double), ("upload_ms", C.c_double),
                ("header_epoch", C.c_uint64), ("rows_revalidated", C.c_uint64), ("stale_rows_found", C.c_uint64)]


class IngestStats(C.Structure):
    _fields_ = [("candidates", C.c_uint64), ("embedded", C.c_uint64), ("skipped_busy", C.c_uint64), ("raced", C.c_uint64),
                ("write_failed", C.c_uint64), ("batches", C.c_uint64), ("gather_ms", C.c_double), ("embed_ms", C.c_double),
                ("write_ms", C.c_double), ("sync_ms", C.c_double)]


class Counters(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("reserved", C.c_uint32), ("batch_queries", C.c_uint64),
                ("batch_unproven", C.c_uint64), ("exact_unproven", C.c_uint64), ("stale_rows_found", C.c_uint64),
                ("batch_rescored_extra", C.c_uint64)]


class Info(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("dim", C.c_uint32), ("total_slots", C.c_uint64),
                ("slot_begin", C.c_uint64), ("slot_count", C.c_uint64), ("generation", C.c_uint64),
                ("device", C.c_int32), ("sm_count", C.c_int32), ("has_bf16_mirror", C.c_uint32),
                ("is_dense", C.c_uint32), ("hbm_bytes", C.c_uint64)]


RESULT_DTYPE = np.dtype([("slot_idx", "<u4"), ("similarity", "<f4"), ("distance", "<f4"),
                         ("has_embedding", "<u4"), ("epoch", "<u8"), ("bloom", "<u8")])
assert RESULT_DTYPE.itemsize == 32

_lib = None


def load() -> C.CDLL:
    """dlopen the in-tree library (built by __graft_entry__.build()).  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    lib = C.CDLL(str(LIB_PATH), mode=os.RTLD_GLOBAL if hasattr(os, "RTLD_GLOBAL") else 0)
    vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
    lib.spl_gpu_attach.restype = vp
    lib.spl_gpu_attach.argtypes = [C.c_char_p, C.POINTER(Opts)]
    lib.spl_gpu_attach_image.restype = vp
    lib.spl_gpu_attach_image.argtypes = [vp, C.c_size_t, C.POINTER(Opts)]
    lib.spl_gpu_open_dense.restype = vp
    lib.spl_gpu_open_dense.argtypes = [u64, u32, u64, C.POINTER(Opts)]
    lib.spl_gpu_detach.restype = None
    lib.spl_gpu_detach.argtypes = [vp]
    lib.spl_gpu_get_info.restype = C.c_int
    lib.spl_gpu_get_info.argtypes = [vp, C.POINTER(Info)]
    lib.spl_gpu_get_counters.restype = C.c_int
    lib.spl_gpu_get_counters.argtypes = [vp, C.POINTER(Counters)]
    lib.spl_gpu_last_error.restype = C.c_char_p
    lib.spl_gpu_last_error.argtypes = []
    lib.spl_gpu_sync.restype = C.c_int
    lib.spl_gpu_sync.argtypes = [vp, C.POINTER(SyncStats)]
    lib.spl_gpu_wait.restype = C.c_int
    lib.spl_gpu_wait.argtypes = [vp, u32]
    lib.spl_gpu_dense_upload.restype = C.c_int
    lib.spl_gpu_dense_upload.argtypes = [vp, u64, u64, vp, vp, vp, vp]
    lib.spl_gpu_dense_fill_synthetic.restype = C.c_int
    lib.spl_gpu_dense_fill_synthetic.argtypes = [vp, u64, u64]
    lib.spl_gpu_read_mirror.restype = C.c_int
    lib.spl_gpu_read_mirror.argtypes = [vp, u64, u64, vp, vp, vp, vp]
    lib.spl_gpu_search.restype = C.c_int
    lib.spl_gpu_search.argtypes = [vp, C.POINTER(SearchParams), vp, vp, u32, vp]
    lib.spl_gpu_search_device.restype = C.c_int
    lib.spl_gpu_search_device.argtypes = [vp, C.POINTER(SearchParams), vp, vp, u32, vp, vp]
    lib.spl_gpu_merge_topk_device.restype = C.c_int
    lib.spl_gpu_merge_topk_device.argtypes = [vp, vp, vp, u32, u32, u32, u32, vp, vp, vp]
    lib.spl_gpu_merge_topk_batch_device.restype = C.c_int
    lib.spl_gpu_merge_topk_batch_device.argtypes = [vp, vp, vp, u32, u32, u32, u32, u32, vp, u32, vp, vp]
    lib.spl_gpu_merge_topk_packed_device.restype = C.c_int
    lib.spl_gpu_merge_topk_packed_device.argtypes = [vp, vp, u32, u32, u32, u32, u32, vp, u32, vp, vp]
    lib.spl_gpu_group_attach.restype = vp
    lib.spl_gpu_group_attach.argtypes = [C.c_char_p, u32, C.POINTER(C.c_int32), C.POINTER(Opts)]
    lib.spl_gpu_group_open_dense.restype = vp
    lib.spl_gpu_group_open_dense.argtypes = [u64, u32, u32, C.POINTER(C.c_int32), C.POINTER(Opts)]
    lib.spl_gpu_group_detach.restype = None
    lib.spl_gpu_group_detach.argtypes = [vp]
    lib.spl_gpu_group_size.restype = u32
    lib.spl_gpu_group_size.argtypes = [vp]
    lib.spl_gpu_group_member.restype = vp
    lib.spl_gpu_group_member.argtypes = [vp, u32]
    lib.spl_gpu_group_sync.restype = C.c_int
    lib.spl_gpu_group_sync.argtypes = [vp, C.POINTER(SyncStats)]
    lib.spl_gpu_group_fill_synthetic.restype = C.c_int
    lib.spl_gpu_group_fill_synthetic.argtypes = [vp, u64, u64]
    lib.spl_gpu_group_search.restype = C.c_int
    lib.spl_gpu_group_search.argtypes = [vp, C.POINTER(SearchParams), vp, vp, u32, vp]
    lib.spl_gpu_group_reissued.restype = u64
    lib.spl_gpu_group_reissued.argtypes = [vp]
    lib.spl_gpu_ingest_embeddings.restype = C.c_int
    lib.spl_gpu_ingest_embeddings.argtypes = [vp, u64, u32, vp, vp, vp, C.POINTER(IngestStats)]
    lib.spl_gpu_scan_noalloc.restype = C.c_int
    lib.spl_gpu_scan_noalloc.argtypes = [vp, vp, vp, vp, u64, u64, C.POINTER(SearchParams), vp, vp, u32, vp, vp]
    lib.spl_gpu_slot_meta.restype = C.c_int
    lib.spl_gpu_slot_meta.argtypes = [vp, u32, C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint8)]
    lib.spl_gpu_find_key.restype = C.c_int
    lib.spl_gpu_find_key.argtypes = [vp, C.c_char_p, C.POINTER(C.c_uint32)]
    lib.spl_gpu_last_totals.restype = C.c_int
    lib.spl_gpu_last_totals.argtypes = [vp, vp, u32]
    lib.spl_gpu_launch_count.restype = u64
    lib.spl_gpu_launch_count.argtypes = [vp]
    lib.spl_gpu_set_timing.restype = C.c_int
    lib.spl_gpu_set_timing.argtypes = [vp, C.c_int]
    lib.spl_gpu_last_scan_ms.restype = C.c_double
    lib.spl_gpu_last_scan_ms.argtypes = [vp]
    _lib = lib
    return lib


def packed_counts_offset(nq: int, stride: int) -> int:
    """SPL_GPU_PACKED_COUNTS_OFFSET"""
    return nq * stride * 32


def packed_block_bytes(nq: int, stride: int) -> int:
    """SPL_GPU_PACKED_BLOCK_BYTES: [nq][stride] records, then [nq] counts padded to a whole record"""
    return packed_counts_offset(nq, stride) + ((nq * 4 + 31) & ~31)


class SplinterGpuError(RuntimeError):
    pass


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _opts(device=-1, flags=0, slot_begin=0, slot_count=0, sync_threads=0):
    return Opts(C.sizeof(Opts), device, flags, slot_begin, slot_count, sync_threads)


class Mirror:
    """One HBM mirror context (a store shard or a dense device-only shard)."""

    def __init__(self, handle):
        self._lib = load()
        if not handle:
            raise SplinterGpuError(self._lib.spl_gpu_last_error().decode() or "attach failed")
        self._h = C.c_void_p(handle)
        self._keep = None
        i = self.info()                 # geometry is fixed for the life of a context
        self._dim, self._slot_count = int(i.dim), int(i.slot_count)

    # ---- constructors ----
    @classmethod
    def attach(cls, name_or_path: str, **kw):
        lib = load()
        o = _opts(**kw)
        return cls(lib.spl_gpu_attach(name_or_path.encode(), C.byref(o)))

    @classmethod
    def attach_image(cls, image: np.ndarray, **kw):
        lib = load()
        o = _opts(**kw)
        assert image.dtype == np.uint8 and image.flags.c_contiguous
        m = cls(lib.spl_gpu_attach_image(_ptr(image), image.nbytes, C.byref(o)))
        m._keep = image
        return m

    @classmethod
    def open_dense(cls, n_rows: int, dim: int, global_row0: int = 0, **kw):
        lib = load()
        o = _opts(**kw)
        return cls(lib.spl_gpu_open_dense(n_rows, dim, global_row0, C.byref(o)))

    # ---- plumbing ----
    def _check(self, rc):
        if rc != 0:
            raise SplinterGpuError(f"rc={rc} errno={C.get_errno()}: {self._lib.spl_gpu_last_error().decode()}")

    def close(self):
        if self._h:
            self._lib.spl_gpu_detach(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def handle(self):
        return self._h

    def info(self) -> Info:
        i = Info()
        self._check(self._lib.spl_gpu_get_info(self._h, C.byref(i)))
        return i

    def counters(self) -> Counters:
        cnt = Counters(C.sizeof(Counters))
        self._check(self._lib.spl_gpu_get_counters(self._h, C.byref(cnt)))
        return cnt

    def wait(self, timeout_ms: int) -> bool:
        """True when the store moved since the last sync (event bus / H->epoch), False on timeout."""
        rc = self._lib.spl_gpu_wait(self._h, int(timeout_ms))
        if rc < 0:
            self._check(rc)
        return rc == 1

    def sync(self) -> SyncStats:
        s = SyncStats()
        self._check(self._lib.spl_gpu_sync(self._h, C.byref(s)))
        return s

    def dense_upload(self, first_row, E, bloom=None, flags=None, epoch=None):
        E = np.ascontiguousarray(E, dtype=np.float32)
        n = E.shape[0]
        bloom = None if bloom is None else np.ascontiguousarray(bloom, dtype=np.uint64)
        flags = None if flags is None else np.ascontiguousarray(flags, dtype=np.uint8)
        epoch = None if epoch is None else np.ascontiguousarray(epoch, dtype=np.uint64)
        self._check(self._lib.spl_gpu_dense_upload(self._h, first_row, n, _ptr(E), _ptr(bloom), _ptr(flags), _ptr(epoch)))

    def fill_synthetic(self, seed_corpus=1234, seed_bloom=99):
        self._check(self._lib.spl_gpu_dense_fill_synthetic(self._h, seed_corpus, seed_bloom))

    def read_mirror(self, first_slot, n, want_E=True):
        dim = self.info().dim
        E = np.empty((n, dim), np.float32) if want_E else None
        bloom = np.empty(n, np.uint64)
        flags = np.empty(n, np.uint8)
        epoch = np.empty(n, np.uint64)
        self._check(self._lib.spl_gpu_read_mirror(self._h, first_slot, n, _ptr(E), _ptr(bloom), _ptr(flags), _ptr(epoch)))
        return E, bloom, flags, epoch

    def params(self, nq, k, order=ORDER_COSINE, mode=MODE_AUTO, bloom_mask=0, min_similarity=0.0,
               max_distance=0.0, exclude_slot=NO_SLOT, flags=0) -> SearchParams:
        return SearchParams(C.sizeof(SearchParams), nq, self._dim, k, order, mode, bloom_mask,
                            min_similarity, max_distance, exclude_slot, flags)

    def search(self, q, k, out_stride=None, **kw):
        """Host-buffer search.  Returns (results[nq, stride] structured array, counts[nq])."""
        q = np.ascontiguousarray(q, dtype=np.float32)
        if q.ndim == 1:
            q = q[None, :]
        nq = q.shape[0]
        p = self.params(nq, k, **kw)
        p.dim = q.shape[1]  # a wrong dimension must be rejected by the library, not papered over here
        stride = out_stride if out_stride is not None else (k if k > 0 else self._slot_count)
        out = np.zeros((nq, stride), RESULT_DTYPE)
        counts = np.zeros(nq, np.uint32)
        self._check(self._lib.spl_gpu_search(self._h, C.byref(p), q.ctypes.data, out.ctypes.data, stride,
                                             counts.ctypes.data))
        return out, counts

    def search_into(self, p: SearchParams, q: np.ndarray, out: np.ndarray, out_stride: int, counts: np.ndarray):
        """spl_gpu_search with caller-owned host buffers and a prepared parameter block: the C call and nothing else
        (what a C host pays per query; `search` above adds array set-up on every call)."""
        rc = self._lib.spl_gpu_search(self._h, C.byref(p), q.ctypes.data, out.ctypes.data, out_stride, counts.ctypes.data)
        if rc != 0:
            self._check(rc)

    def search_ptrs(self, p_ref, q_addr: int, out_addr: int, out_stride: int, counts_addr: int):
        """spl_gpu_search on raw host addresses (p_ref = ctypes.byref(params), made once): nothing but the C call, for
        loops that would otherwise time numpy's `.ctypes` accessors (about a microsecond each) along with it"""
        rc = self._lib.spl_gpu_search(self._h, p_ref, q_addr, out_addr, out_stride, counts_addr)
        if rc != 0:
            self._check(rc)

    def search_device(self, p: SearchParams, d_q: int, d_out: int, out_stride: int, d_counts: int, stream: int = 0):
        self._check(self._lib.spl_gpu_search_device(self._h, C.byref(p), C.c_void_p(d_q), C.c_void_p(d_out),
                                                    out_stride, C.c_void_p(d_counts), C.c_void_p(stream)))

    def merge_topk_device(self, d_lists, d_counts, nlists, stride, order, k, d_out, d_out_count, stream=0):
        self._check(self._lib.spl_gpu_merge_topk_device(self._h, C.c_void_p(d_lists), C.c_void_p(d_counts), nlists,
                                                        stride, order, k, C.c_void_p(d_out), C.c_void_p(d_out_count),
                                                        C.c_void_p(stream)))

    def merge_topk_batch_device(self, d_lists, d_counts, nlists, nq, stride, order, k, d_out, out_stride, d_out_counts,
                                stream=0):
        self._check(self._lib.spl_gpu_merge_topk_batch_device(self._h, C.c_void_p(d_lists), C.c_void_p(d_counts), nlists, nq,
                                                              stride, order, k, C.c_void_p(d_out), out_stride,
                                                              C.c_void_p(d_out_counts), C.c_void_p(stream)))

    def merge_topk_packed_device(self, d_packed, nlists, nq, stride, order, k, d_out, out_stride, d_out_counts, stream=0):
        self._check(self._lib.spl_gpu_merge_topk_packed_device(self._h, C.c_void_p(d_packed), nlists, nq, stride, order, k,
                                                               C.c_void_p(d_out), out_stride, C.c_void_p(d_out_counts),
                                                               C.c_void_p(stream)))

    def ingest_embeddings(self, label_mask, batch, embed_fn_ptr, user_ptr, set_embedding_fn_ptr) -> IngestStats:
        """embed_fn_ptr / set_embedding_fn_ptr: C function pointers (ctypes functions or addresses)"""
        st = IngestStats()
        as_vp = lambda f: C.cast(f, C.c_void_p)  # noqa: E731
        self._check(self._lib.spl_gpu_ingest_embeddings(self._h, label_mask, batch, as_vp(embed_fn_ptr), C.c_void_p(user_ptr),
                                                        as_vp(set_embedding_fn_ptr), C.byref(st)))
        return st

    def find_key(self, key: str):
        """slot index of `key` in the live store, or None"""
        idx = C.c_uint32(0)
        rc = self._lib.spl_gpu_find_key(self._h, key.encode(), C.byref(idx))
        return int(idx.value) if rc == 0 else None

    def last_totals(self, nq=1):
        t = np.zeros(nq, np.uint32)
        self._check(self._lib.spl_gpu_last_totals(self._h, _ptr(t), nq))
        return t

    def launch_count(self) -> int:
        return int(self._lib.spl_gpu_launch_count(self._h))

    def set_timing(self, on=True):
        self._check(self._lib.spl_gpu_set_timing(self._h, int(on)))

    def last_scan_ms(self) -> float:
        return float(self._lib.spl_gpu_last_scan_ms(self._h))


class Group:
    """Several GPUs behind one call (spl_gpu_group_*): one context per device over contiguous slot ranges."""

    def __init__(self, handle, dim):
        self._lib = load()
        if not handle:
            raise SplinterGpuError(self._lib.spl_gpu_last_error().decode() or "group attach failed")
        self._h = C.c_void_p(handle)
        self._dim = dim

    @staticmethod
    def _devs(devices):
        return (C.c_int32 * len(devices))(*devices)

    @classmethod
    def attach(cls, name_or_path: str, devices, dim=768, **kw):
        o = _opts(**kw)
        return cls(load().spl_gpu_group_attach(name_or_path.encode(), len(devices), cls._devs(devices), C.byref(o)), dim)

    @classmethod
    def open_dense(cls, n_rows: int, dim: int, devices, **kw):
        o = _opts(**kw)
        return cls(load().spl_gpu_group_open_dense(n_rows, dim, len(devices), cls._devs(devices), C.byref(o)), dim)

    def _check(self, rc):
        if rc != 0:
            raise SplinterGpuError(f"rc={rc} errno={C.get_errno()}: {self._lib.spl_gpu_last_error().decode()}")

    def close(self):
        if self._h:
            self._lib.spl_gpu_group_detach(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def size(self) -> int:
        return int(self._lib.spl_gpu_group_size(self._h))

    def member(self, i) -> "Mirror":
        """a NON-owning view of shard i's context (do not close it)"""
        m = Mirror.__new__(Mirror)
        m._lib, m._h, m._keep = self._lib, C.c_void_p(self._lib.spl_gpu_group_member(self._h, i)), None
        inf = m.info()
        m._dim, m._slot_count = int(inf.dim), int(inf.slot_count)
        return m

    def sync(self) -> SyncStats:
        s = SyncStats()
        self._check(self._lib.spl_gpu_group_sync(self._h, C.byref(s)))
        return s

    def fill_synthetic(self, seed_corpus=1234, seed_bloom=99):
        self._check(self._lib.spl_gpu_group_fill_synthetic(self._h, seed_corpus, seed_bloom))

    def reissued(self) -> int:
        return int(self._lib.spl_gpu_group_reissued(self._h))

    def search(self, q, k, **kw):
        q = np.ascontiguousarray(q, dtype=np.float32)
        if q.ndim == 1:
            q = q[None, :]
        nq = q.shape[0]
        p = SearchParams(C.sizeof(SearchParams), nq, q.shape[1], k, kw.get("order", ORDER_COSINE), kw.get("mode", MODE_AUTO),
                         kw.get("bloom_mask", 0), kw.get("min_similarity", 0.0), kw.get("max_distance", 0.0),
                         kw.get("exclude_slot", NO_SLOT), 0)
        out = np.zeros((nq, k), RESULT_DTYPE)
        counts = np.zeros(nq, np.uint32)
        self._check(self._lib.spl_gpu_group_search(self._h, C.byref(p), q.ctypes.data, out.ctypes.data, k, counts.ctypes.data))
        return out, counts
