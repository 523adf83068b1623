"""ctypes binding of include/mrgpu.h (libmrgpu.so).

There is no CPU path: if the CUDA library cannot be loaded, or no sm_100 device is
present, every compute entry fails loudly (NativeError).
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

MR_OK, MR_ERR_INVALID, MR_ERR_CUDA, MR_ERR_NOMEM, MR_ERR_OVERFLOW, MR_ERR_UNSUPPORTED, MR_ERR_END = range(7)
MR_MAP_TOKENIZE_WS, MR_MAP_RECORDS_U32, MR_MAP_BIGRAM_WS = 0, 1, 2
MR_REDUCE_SUM_U64, MR_REDUCE_MIN_U64, MR_REDUCE_MAX_U64, MR_REDUCE_COUNT = 0, 1, 2, 3
MR_PARTITION_FNV1_LUA52, MR_PARTITION_MOD_U32 = 0, 1

# every symbol include/mrgpu.h declares
EXPORTS = (
    "mr_abi_version", "mr_ctx_create", "mr_ctx_destroy", "mr_last_error", "mr_job_begin",
    "mr_map_split", "mr_map_split_path", "mr_map_split_device", "mr_map_records", "mr_map_records_device", "mr_map_pairs", "mr_map_finish", "mr_reduce_all", "mr_job_sync",
    "mr_device_rows_get", "mr_rows_fetch", "mr_reduce_partition", "mr_result_next",
    "mr_result_release", "mr_shuffle_count", "mr_shuffle_pack", "mr_shuffle_recv_buffer", "mr_shuffle_import",
    "mr_shuffle_pack_fixed", "mr_shuffle_import_fixed",
    "mr_nccl_unique_id", "mr_nccl_init", "mr_shuffle_nccl",
    "mr_p2p_setup", "mr_p2p_attach", "mr_shuffle_p2p",
    "mr_shuffle_set_owners", "mr_row_sort_word", "mr_owner_of",
    "mr_stats_get", "mr_partition_of", "mr_count_tokens_static",
)


class NativeError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"mrgpu error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("num_reducers", C.c_uint32),
                ("map_kind", C.c_uint32), ("reduce_kind", C.c_uint32), ("partition_kind", C.c_uint32),
                ("arena_bytes", C.c_uint64), ("max_keys", C.c_uint64), ("table_slots", C.c_uint64),
                ("stream", C.c_void_p), ("world_size", C.c_uint32), ("rank", C.c_uint32)]


class JobInfo(C.Structure):
    _fields_ = [("n_keys", C.c_uint64), ("n_tokens", C.c_uint64), ("input_bytes", C.c_uint64),
                ("key_bytes", C.c_uint64), ("map_keys", C.c_uint64), ("tie_rounds", C.c_uint32),
                ("flags", C.c_uint32)]


class DeviceRows(C.Structure):
    _fields_ = [("n", C.c_uint64), ("part", C.c_void_p), ("count", C.c_void_p),
                ("key_off", C.c_void_p), ("key_bytes", C.c_void_p), ("part_begin", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [("map_ms", C.c_float), ("reduce_ms", C.c_float), ("h2d_bytes", C.c_uint64),
                ("d2h_bytes", C.c_uint64), ("launches", C.c_uint64)]


class ShufflePlan(C.Structure):
    _fields_ = [("send_base", C.c_void_p), ("send_off", C.POINTER(C.c_uint64)),
                ("send_nrec", C.POINTER(C.c_uint64))]


_lib = None


def library_path():
    return _build.LIB_GPU


def load(build_if_missing=True):
    """Loads libmrgpu.so (building it in-tree with nvcc when the sources are newer)."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_GPU
    if os.environ.get("MRGPU_LIB"):          # a variant build of the same ABI (scripts/map_variants.py)
        path, build_if_missing = os.environ["MRGPU_LIB"], False
    if build_if_missing:
        try:
            path = _build.build_gpu()
        except Exception as e:  # no nvcc on this box: use the prebuilt library if it is there
            if not os.path.exists(path):
                raise NativeError(-1, f"libmrgpu.so is missing and cannot be built: {e}")
    if not os.path.exists(path):
        raise NativeError(-1, "libmrgpu.so is missing; run `python __graft_entry__.py build`")
    L = C.CDLL(path)
    vp, u64, u32 = C.c_void_p, C.c_uint64, C.c_uint32
    L.mr_abi_version.restype = C.c_int
    L.mr_ctx_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    L.mr_ctx_destroy.argtypes = [vp]
    L.mr_ctx_destroy.restype = None
    L.mr_last_error.argtypes = [vp]
    L.mr_last_error.restype = C.c_char_p
    L.mr_job_begin.argtypes = [vp, C.c_int]
    L.mr_map_split.argtypes = [vp, u64, vp, u64]
    L.mr_map_split_device.argtypes = [vp, u64, vp, u64]
    L.mr_map_split_path.argtypes = [vp, u64, C.c_char_p]
    L.mr_map_records.argtypes = [vp, u64, vp, vp, u64]
    L.mr_map_records_device.argtypes = [vp, u64, vp, vp, u64]
    L.mr_map_pairs.argtypes = [vp, u64, vp, vp, vp, u64]
    L.mr_map_finish.argtypes = [vp]
    L.mr_reduce_all.argtypes = [vp]
    L.mr_job_sync.argtypes = [vp, C.POINTER(JobInfo)]
    L.mr_device_rows_get.argtypes = [vp, C.POINTER(DeviceRows)]
    L.mr_rows_fetch.argtypes = [vp, vp, vp, vp, vp, vp]
    L.mr_reduce_partition.argtypes = [vp, u32, C.POINTER(vp)]
    L.mr_result_next.argtypes = [vp, C.POINTER(vp), C.POINTER(u64), C.POINTER(u64)]
    L.mr_result_release.argtypes = [vp]
    L.mr_result_release.restype = None
    L.mr_shuffle_count.argtypes = [vp, C.POINTER(vp)]
    L.mr_shuffle_pack.argtypes = [vp, C.POINTER(u64), C.POINTER(ShufflePlan)]
    L.mr_shuffle_pack_fixed.argtypes = [vp, u64, C.POINTER(vp)]
    L.mr_shuffle_import_fixed.argtypes = [vp, u64]
    L.mr_nccl_unique_id.argtypes = [C.c_char_p]
    L.mr_nccl_init.argtypes = [vp, C.c_char_p]
    L.mr_shuffle_nccl.argtypes = [vp, u64, C.POINTER(u64)]
    L.mr_p2p_setup.argtypes = [vp, u64, C.c_char_p]
    L.mr_p2p_attach.argtypes = [vp, C.c_char_p]
    L.mr_shuffle_p2p.argtypes = [vp]
    L.mr_shuffle_set_owners.argtypes = [vp, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(u64)]
    L.mr_row_sort_word.argtypes = [vp, u64, C.POINTER(C.c_uint32), C.POINTER(u64)]
    L.mr_owner_of.argtypes = [vp, C.c_char_p, u64, C.POINTER(C.c_uint32)]
    L.mr_shuffle_recv_buffer.argtypes = [vp, u64, C.POINTER(vp)]
    L.mr_shuffle_import.argtypes = [vp, C.POINTER(u64), C.POINTER(u64)]
    L.mr_stats_get.argtypes = [vp, C.POINTER(Stats)]
    L.mr_partition_of.argtypes = [C.c_char_p, u64, u32]
    L.mr_partition_of.restype = u32
    L.mr_count_tokens_static.argtypes = [vp, u64, vp, u64, vp, C.POINTER(u64), vp, C.c_char_p, u64]
    _lib = L
    return L


class Rows:
    """Host copy of the reduced rows, in the reference's final order."""

    def __init__(self, part, count, key_off, key_bytes, part_begin):
        self.part, self.count, self.key_off = part, count, key_off
        self.key_bytes, self.part_begin = key_bytes, part_begin

    @property
    def n(self):
        return len(self.part)

    def keys(self):
        kb = self.key_bytes.tobytes()
        off = self.key_off
        return [kb[int(off[i]):int(off[i + 1])] for i in range(self.n)]

    def keys_u32(self):
        """Number keys of a MR_MAP_RECORDS_U32 job (4 big-endian bytes per row)."""
        return np.frombuffer(self.key_bytes.tobytes(), dtype=">u4").astype(np.uint32)


class Context:
    """One per worker process per GPU (include/mrgpu.h: mr_ctx)."""

    def __init__(self, num_reducers, arena_bytes, max_keys=0, table_slots=0, device=0, stream=None,
                 world_size=1, rank=0, map_kind=MR_MAP_TOKENIZE_WS, partition_kind=MR_PARTITION_FNV1_LUA52,
                 reduce_kind=MR_REDUCE_SUM_U64):
        self._L = load()
        cfg = Config(C.sizeof(Config), device, num_reducers, map_kind, reduce_kind,
                     partition_kind, arena_bytes, max_keys, table_slots, stream, world_size, rank)
        h = C.c_void_p()
        rc = self._L.mr_ctx_create(C.byref(cfg), C.byref(h))
        if rc != MR_OK:
            raise NativeError(rc, (self._L.mr_last_error(None) or b"").decode())
        self._h = h
        self.num_reducers, self.world_size, self.rank = num_reducers, world_size, rank
        self.stream = stream    # cudaStream_t handle the context runs on; None = a private stream of its own

    def _ck(self, rc):
        if rc != MR_OK:
            raise NativeError(rc, (self._L.mr_last_error(self._h) or b"").decode())

    def close(self):
        if getattr(self, "_h", None):
            self._L.mr_ctx_destroy(self._h)
            self._h = None

    __del__ = close

    def job_begin(self, keep_input=False):
        self._ck(self._L.mr_job_begin(self._h, int(keep_input)))

    def map_split(self, split_id, data):
        """data: bytes-like or a uint8 numpy array (host memory)."""
        arr = data if isinstance(data, np.ndarray) else np.frombuffer(data, dtype=np.uint8)
        self._ck(self._L.mr_map_split(self._h, split_id, arr.ctypes.data if arr.size else None, arr.size))

    def map_split_path(self, split_id, path):
        """The library reads the file itself (what the reference's mapfn does with its value)."""
        self._ck(self._L.mr_map_split_path(self._h, split_id, os.fsencode(path)))

    def map_split_ptr(self, split_id, ptr, n, device=False):
        fn = self._L.mr_map_split_device if device else self._L.mr_map_split
        self._ck(fn(self._h, split_id, ptr, n))

    def map_records(self, split_id, keys, values):
        """keys, values: uint32 numpy arrays of equal length (host memory)."""
        k = np.ascontiguousarray(keys, dtype=np.uint32)
        v = np.ascontiguousarray(values, dtype=np.uint32)
        assert k.shape == v.shape and k.ndim == 1
        self._ck(self._L.mr_map_records(self._h, split_id, k.ctypes.data, v.ctypes.data, k.size))

    def map_records_ptr(self, split_id, keys_ptr, values_ptr, n, device=False):
        fn = self._L.mr_map_records_device if device else self._L.mr_map_records
        self._ck(fn(self._h, split_id, keys_ptr, values_ptr, n))

    def map_pairs(self, split_id, keys, values):
        """keys: list of bytes (1 byte or more each); values: one integer per key (a run line's value
        list already reduced by the caller).  mr_map_pairs: the reduce job's read of other mappers' runs."""
        n = len(keys)
        assert n == len(values)
        if not n:
            return
        blob = np.frombuffer(b"".join(keys), dtype=np.uint8)
        off = np.zeros(n + 1, np.uint64)
        np.cumsum([len(k) for k in keys], out=off[1:])
        val = np.ascontiguousarray(values, dtype=np.uint64)
        self._ck(self._L.mr_map_pairs(self._h, split_id, blob.ctypes.data if blob.size else None, off.ctypes.data,
                                      val.ctypes.data, n))

    def map_finish(self):
        self._ck(self._L.mr_map_finish(self._h))

    def reduce_all(self):
        self._ck(self._L.mr_reduce_all(self._h))

    def job_sync(self):
        info = JobInfo()
        self._ck(self._L.mr_job_sync(self._h, C.byref(info)))
        return info

    def device_rows(self):
        r = DeviceRows()
        self._ck(self._L.mr_device_rows_get(self._h, C.byref(r)))
        return r

    def rows(self, out=None):
        """Bulk copy of all rows to host numpy arrays (optionally into preallocated `out`)."""
        info = self.job_sync()
        n, kb = int(info.n_keys), int(info.key_bytes)
        if out is None:
            out = Rows(np.empty(n, np.uint32), np.empty(n, np.uint64), np.empty(n + 1, np.uint64),
                       np.empty(kb, np.uint8), np.empty(self.num_reducers + 1, np.uint64))
        self._ck(self._L.mr_rows_fetch(self._h, out.part.ctypes.data, out.count.ctypes.data,
                                       out.key_off.ctypes.data, out.key_bytes.ctypes.data,
                                       out.part_begin.ctypes.data))
        return out

    def rows_part_begin(self):
        """part_begin of the finished job: rows of partition p are [part_begin[p], part_begin[p+1])."""
        self.job_sync()
        pb = np.empty(self.num_reducers + 1, np.uint64)
        self._ck(self._L.mr_rows_fetch(self._h, None, None, None, None, pb.ctypes.data))
        return pb

    def partition_cursor(self, part):
        """Iterates (key bytes, count) of one partition, keys ascending (result.P<part>)."""
        h = C.c_void_p()
        self._ck(self._L.mr_reduce_partition(self._h, part, C.byref(h)))
        try:
            key, klen, cnt = C.c_void_p(), C.c_uint64(), C.c_uint64()
            while True:
                rc = self._L.mr_result_next(h, C.byref(key), C.byref(klen), C.byref(cnt))
                if rc == MR_ERR_END:
                    break
                self._ck(rc)
                yield C.string_at(key.value, klen.value), int(cnt.value)
        finally:
            self._L.mr_result_release(h)

    def stats(self):
        s = Stats()
        self._ck(self._L.mr_stats_get(self._h, C.byref(s)))
        return s

    # ---- shuffle seam
    def shuffle_count(self):
        """Device address of this rank's [2 * world_size] counts (records, then long-key bytes, per destination)."""
        p = C.c_void_p()
        self._ck(self._L.mr_shuffle_count(self._h, C.byref(p)))
        return p.value

    def shuffle_pack(self, counts):
        """counts: host copy of the array shuffle_count points at.  Returns (device base, byte offsets, records)."""
        c = np.ascontiguousarray(counts, dtype=np.uint64)
        plan = ShufflePlan()
        self._ck(self._L.mr_shuffle_pack(self._h, c.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(plan)))
        W = self.world_size
        off = np.array([plan.send_off[i] for i in range(W + 1)], dtype=np.uint64)
        nrec = np.array([plan.send_nrec[i] for i in range(W)], dtype=np.uint64)
        return plan.send_base, off, nrec

    def shuffle_pack_fixed(self, slot_bytes):
        p = C.c_void_p()
        self._ck(self._L.mr_shuffle_pack_fixed(self._h, slot_bytes, C.byref(p)))
        return p.value

    def shuffle_import_fixed(self, slot_bytes):
        self._ck(self._L.mr_shuffle_import_fixed(self._h, slot_bytes))

    def nccl_init(self, unique_id):
        """Collective: joins the library's own NCCL communicator (mr_nccl_init)."""
        if len(unique_id) != NCCL_ID_BYTES:
            raise ValueError("unique_id must be %d bytes" % NCCL_ID_BYTES)
        self._ck(self._L.mr_nccl_init(self._h, bytes(unique_id)))

    def shuffle_nccl(self, slot_bytes=0):
        """pack -> all-to-all -> import inside the library; returns the slot size used."""
        used = C.c_uint64(0)
        self._ck(self._L.mr_shuffle_nccl(self._h, int(slot_bytes), C.byref(used)))
        return used.value

    def set_owners(self, splitters):
        """splitters: ascending list of (partition, sort word), world_size - 1 of them, or empty for
        the default (owner = partition mod world_size).  Every rank installs the same list."""
        n = len(splitters)
        parts = (C.c_uint32 * max(n, 1))(*[int(p) for p, _ in splitters])
        words = (C.c_uint64 * max(n, 1))(*[int(w) for _, w in splitters])
        self._ck(self._L.mr_shuffle_set_owners(self._h, n, parts, words))

    def row_sort_word(self, row):
        """(partition, sort word) of one row of the finished job."""
        p, w = C.c_uint32(), C.c_uint64()
        self._ck(self._L.mr_row_sort_word(self._h, int(row), C.byref(p), C.byref(w)))
        return int(p.value), int(w.value)

    def owner_of(self, key):
        o = C.c_uint32()
        self._ck(self._L.mr_owner_of(self._h, bytes(key), len(key), C.byref(o)))
        return int(o.value)

    def p2p_setup(self, slot_bytes):
        """Allocates this rank's receive window; returns the 128-byte mr_peer_info blob for the peers."""
        blob = C.create_string_buffer(PEER_INFO_BYTES)
        self._ck(self._L.mr_p2p_setup(self._h, int(slot_bytes), blob))
        return blob.raw

    def p2p_attach(self, blobs):
        """blobs: the mr_peer_info of every rank, in rank order."""
        if len(blobs) != self.world_size or any(len(b) != PEER_INFO_BYTES for b in blobs):
            raise ValueError("need one %d-byte blob per rank" % PEER_INFO_BYTES)
        self._ck(self._L.mr_p2p_attach(self._h, b"".join(blobs)))

    def shuffle_p2p(self):
        """pack straight into the owners' windows -> flags -> index in place; asynchronous."""
        self._ck(self._L.mr_shuffle_p2p(self._h))

    def shuffle_recv_buffer(self, nbytes):
        p = C.c_void_p()
        self._ck(self._L.mr_shuffle_recv_buffer(self._h, nbytes, C.byref(p)))
        return p.value

    def shuffle_import(self, recv_off, recv_nrec):
        ro = np.ascontiguousarray(recv_off, dtype=np.uint64)
        rn = np.ascontiguousarray(recv_nrec, dtype=np.uint64)
        self._ck(self._L.mr_shuffle_import(self._h, ro.ctypes.data_as(C.POINTER(C.c_uint64)),
                                           rn.ctypes.data_as(C.POINTER(C.c_uint64))))


def partition_of(key: bytes, num_reducers: int) -> int:
    return load().mr_partition_of(key, len(key), num_reducers)


STATIC_MAX_BYTES = 128 * 1024
NCCL_ID_BYTES = 128
PEER_INFO_BYTES = 128


def nccl_unique_id():
    """mr_nccl_unique_id: the rendezvous blob rank 0 hands to the other ranks."""
    buf = C.create_string_buffer(NCCL_ID_BYTES)
    L = load()
    rc = L.mr_nccl_unique_id(buf)
    if rc != MR_OK:
        raise NativeError(rc, (L.mr_last_error(None) or b"").decode())
    return buf.raw


TOKEN_COUNT_DTYPE = np.dtype([("key", "S15"), ("len", np.uint8), ("count", np.uint64)])


def count_tokens_static(dev_text, n_bytes, dev_out, out_capacity, dev_n_out, stream=None):
    """mr_count_tokens_static: every pointer is a raw device address owned by the caller; nothing is
    allocated on the device.  Returns the number of tokens; raises NativeError with the library's text."""
    err = C.create_string_buffer(256)
    ntok = C.c_uint64(0)
    rc = load().mr_count_tokens_static(dev_text, n_bytes, dev_out, out_capacity, dev_n_out, C.byref(ntok),
                                       stream, err, len(err))
    if rc != MR_OK:
        raise NativeError(rc, err.value.decode())
    return int(ntok.value)
